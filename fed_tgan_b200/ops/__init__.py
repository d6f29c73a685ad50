"""Op surface of the framework.

Every op exists twice with one signature:

* ``fed_tgan_b200.ops.native``     hand-written sm_100a kernels (``csrc/``, C ABI, loaded
                                    from the in-tree ``libfedtgan_b200.so``)
* ``fed_tgan_b200.ops.reference``  plain fp32 PyTorch (oracle + CPU path)

``ops.<name>(...)`` dispatches on the device of the first tensor argument: CUDA tensors go
to the native library and **fail loudly** if it cannot be loaded (no silent PyTorch
fallback on a GPU box); CPU tensors go to the reference.  Set ``FEDTGAN_FORCE_REFERENCE=1``
to run the oracle on CUDA tensors (used by numerics tests only).
"""

from __future__ import annotations

import os

import torch

from . import reference
from .reference import RngState  # noqa: F401

_NAMES = [
    "linear", "linear_bn_relu", "rowdot", "d_head", "d_mid", "colsum", "scale_outer_mask", "bn_relu_fwd", "bn_relu_bwd", "activate_fwd",
    "activate_bwd", "gen_output_bwd", "cond_loss_fwd_bwd", "slerp", "gp_penalty", "adam_step", "fill_normal",
    "sample_condvec", "write_condvec", "sample_rows", "expand_rows", "argsort_perm",
    "vgm_decode", "vgm_encode",
]


def _first_tensor(args, kwargs):
    for a in list(args) + list(kwargs.values()):
        if torch.is_tensor(a):
            return a
    return None


def _force_reference() -> bool:
    return os.environ.get("FEDTGAN_FORCE_REFERENCE", "0") == "1"


def _make(name):
    ref_fn = getattr(reference, name)

    def op(*args, **kwargs):
        t = _first_tensor(args, kwargs)
        if t is not None and t.is_cuda and not _force_reference():
            from . import native  # raises if the .so is missing

            return getattr(native, name)(*args, **kwargs)
        return ref_fn(*args, **kwargs)

    op.__name__ = name
    op.__doc__ = ref_fn.__doc__
    return op


for _n in _NAMES:
    globals()[_n] = _make(_n)


def native_available() -> bool:
    try:
        from . import native  # noqa: F401

        return True
    except Exception:
        return False
