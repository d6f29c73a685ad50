"""ctypes bindings of ``libfedtgan_b200.so`` with the same signatures as ``ops.reference``.

Importing this module loads the in-tree shared object (building it first if the sources
are newer) and raises if that is impossible -- on a GPU box there is no silent PyTorch
fallback.  All launches go to ``torch.cuda.current_stream()``, so they order with the
caller's PyTorch work and are CUDA-graph capturable.
"""

from __future__ import annotations

import ctypes as C
import os

import torch

from . import build as _build

_LIB_PATH = _build.LIB
if os.environ.get("FEDTGAN_NO_BUILD", "0") != "1":
    try:
        _LIB_PATH = _build.build()
    except Exception:  # no nvcc on this box: use the shipped .so
        if not os.path.exists(_LIB_PATH):
            raise
lib = C.CDLL(_LIB_PATH)

UNSUPPORTED = 1000
_vp, _i, _f, _ll = C.c_void_p, C.c_int, C.c_float, C.c_longlong


class GemmArgs(C.Structure):
    _fields_ = [
        ("a", _vp), ("b", _vp), ("d", _vp),
        ("M", _i), ("N", _i), ("K", _i),
        ("lda", _i), ("ldb", _i), ("ldd", _i),
        ("a_t", _i), ("b_t", _i),
        ("bias", _vp),
        ("act", _i),
        ("mask_in", _vp), ("ld_mask_in", _i),
        ("mask_out", _vp), ("ld_mask_out", _i),
        ("drop_u", _vp), ("ld_drop_u", _i),
        ("rng_state", _vp), ("rng_stream", _i),
        ("alpha", _f), ("slope", _f), ("drop_p", _f),
        ("accumulate", _i),
        ("x3", _i),
    ]


class BnArgs(C.Structure):
    _fields_ = [
        ("gamma", _vp), ("beta", _vp), ("running_mean", _vp), ("running_var", _vp),
        ("save_mean", _vp), ("save_invstd", _vp), ("out", _vp), ("ldo", _i),
        ("eps", _f), ("momentum", _f), ("nbt", _vp),
    ]


def _sig(name, argtypes, restype=_i):
    fn = getattr(lib, name)
    fn.argtypes = argtypes
    fn.restype = restype
    return fn


_gemm_simt = _sig("ftg_gemm_simt", [C.POINTER(GemmArgs), _vp])
_gemm_tc = _sig("ftg_gemm_tc", [C.POINTER(GemmArgs), _i, _vp, _vp, _vp])
_gemm_tc_bn = _sig("ftg_gemm_tc_bn", [C.POINTER(GemmArgs), C.POINTER(BnArgs), _vp])
_gemm_tc_bn_plan = _sig("ftg_gemm_tc_bn_plan", [_i, _i])
_choose_split = _sig("ftg_gemm_tc_choose_split", [_i, _i, _i])
_choose_cluster_split = _sig("ftg_gemm_tc_choose_cluster_split", [_i, _i, _i])
_ws_bytes = _sig("ftg_gemm_tc_workspace_bytes", [_i, _i, _i], _ll)
_num_tiles = _sig("ftg_gemm_tc_num_tiles", [_i, _i])
_rowdot = _sig("ftg_rowdot", [_vp, _i, _i, _i, _vp, _vp, _vp, _vp])
_rowdot_loss = _sig("ftg_rowdot_loss", [_vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp])
_prep_select = _sig("ftg_prep_select", [_vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _i, _i, _i, _i, _vp])
_act_fwd_slerp = _sig("ftg_activate_fwd_slerp", [_vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _i, _f, _vp, _i, _vp, _i, _vp,
                                                 _i, _i, _i, _vp, _i, _vp, _i, _vp, _i, _vp])
_gen_out_bwd = _sig("ftg_gen_output_bwd", [_vp, _i, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _i, _f, _vp, _i, _vp, _i, _vp, _i,
                                           _vp, _vp, _vp, _vp])
_prep_rows = _sig("ftg_prep_rows", [_i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp, _i, _i, _vp, _i, _i, _vp, _i, _vp,
                                    _vp, _vp, _vp, _i, _i, _vp, _i, _vp, _i, _vp, _i, _i, _vp])
_d_head = _sig("ftg_d_head", [_vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _i, _vp])
_d_mid = _sig("ftg_d_mid", [_vp, _i, _i, _i, _i, _vp, _i, _vp, _vp, _i, _vp, _i, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i,
                            _vp, _i, _vp, _i, _vp, _i, _f, _f, _vp, _vp])
_d_mid_supported = _sig("ftg_d_mid_supported", [_i, _i])
_colsum = _sig("ftg_colsum", [_vp, _i, _i, _i, _vp, _vp, _i, _vp])
_som = _sig("ftg_scale_outer_mask", [_vp, _vp, _vp, _i, _vp, _i, _i, _i, _vp])
_bn_fwd = _sig("ftg_bn_relu_fwd", [_vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _f, _f, _vp, _vp])
_bn_bwd = _sig("ftg_bn_relu_bwd", [_vp, _i, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp])
_bn_bwd_bias = _sig("ftg_bn_relu_bwd_bias", [_vp, _i, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp])
_act_fwd = _sig("ftg_activate_fwd", [_vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _i, _f, _vp, _i, _vp, _i, _vp, _i, _i, _i, _vp])
_act_bwd = _sig("ftg_activate_bwd", [_vp, _i, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _i, _f, _i, _vp])
_cond_loss = _sig("ftg_cond_loss", [_vp, _i, _vp, _i, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _i, _vp, _i, _vp])
_slerp = _sig("ftg_slerp", [_vp, _i, _vp, _i, _vp, _i, _i, _i, _vp, _vp, _i, _vp])
_gp = _sig("ftg_gp_penalty", [_vp, _i, _i, _i, _vp, _f, _vp])
_adam = _sig("ftg_adam", [_vp, _vp, _vp, _vp, _ll, _vp, _f, _f, _f, _f, _f, _vp])
_tick = _sig("ftg_tick", [_vp, _vp])
_fill_normal = _sig("ftg_fill_normal", [_vp, _i, _i, _i, _vp, _i, _vp])
_sample_condvec = _sig("ftg_sample_condvec", [_vp, _i, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _vp])
_write_condvec = _sig("ftg_write_condvec", [_vp, _vp, _vp, _vp, _i, _i, _vp, _i, _i, _vp, _i, _vp])
_argsort = _sig("ftg_argsort_perm", [_vp, _i, _vp, _vp, _i, _vp])
_sample_rows = _sig("ftg_sample_rows", [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _vp])
_expand_rows = _sig("ftg_expand_rows", [_vp, _i, _vp, _i, _vp, _vp, _vp, _vp, _i, _i, _i, _vp])
_vgm_decode = _sig("ftg_vgm_decode", [_vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _vp])
_vgm_encode = _sig("ftg_vgm_encode", [_vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _i, _vp, _i, _vp])
_fedavg_local = _sig("ftg_fedavg_local", [_vp, _i, _vp, _ll, _vp, _i, _vp])
_fedavg_twoshot = _sig("ftg_fedavg_twoshot", [_vp, _vp, _i, _i, _vp, _ll, _vp, _i, _i, _vp, _i, _vp])
_fedavg_oneshot = _sig("ftg_fedavg_oneshot", [_vp, _vp, _vp, _i, _i, _vp, _ll, _vp, _i, _i, _vp, _i, _vp])
_fedavg_nvls = _sig("ftg_fedavg_nvls", [_vp, _vp, _vp, _i, _i, _vp, _ll, _vp, _i, _i, _vp, _i, _vp])
_symm_alloc = _sig("ftg_symm_alloc", [_ll, C.POINTER(_vp)])
_symm_free = _sig("ftg_symm_free", [_vp])
_symm_export = _sig("ftg_symm_export", [_vp, C.c_char_p])
_symm_open = _sig("ftg_symm_open", [C.c_char_p, C.POINTER(_vp)])
_symm_close = _sig("ftg_symm_close", [_vp])
_sort_pad = _sig("ftg_sort_pad", [_i])
_gmm_sample_sort = _sig("ftg_gmm_sample_sort", [_vp, _i, _i, _i, C.c_ulonglong, _i, _vp, _vp])
_sort_rows = _sig("ftg_sort_rows", [_vp, _i, _i, _vp, _vp])
_fed_stats = _sig("ftg_fed_stats_to_weights", [_vp, _vp, _vp, _vp, _i, _i, _vp, _i, _vp, _i, _vp, _i, _vp, _vp, _vp, _vp, _vp])
_stats = _sig("ftg_stats_to_weights", [_vp, _i, _i, _vp, _vp, _i, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp])
_vbgmm = _sig("ftg_vbgmm_fit", [_vp, _i, _i, _vp, _i, C.c_double, _i, C.c_double, C.c_double, _i, _vp, _vp, _vp])
_launch_count = _sig("ftg_last_launch_count", [], _ll)
lib.ftg_version.restype = C.c_char_p


def version() -> str:
    return lib.ftg_version().decode()


def launch_count() -> int:
    return int(_launch_count())


class NativeError(RuntimeError):
    pass


def _ck(rc, what):
    if rc != 0:
        raise NativeError(f"{what} failed with code {rc}")


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _p(t):
    return None if t is None else t.data_ptr()


def _m(t):
    """(ptr, ld) of a 2-D fp32 view with unit column stride."""
    if t is None:
        return None, 0
    assert t.dtype == torch.float32 and t.dim() == 2 and (t.shape[1] == 1 or t.stride(1) == 1), (t.dtype, t.shape, t.stride())
    return t.data_ptr(), (t.stride(0) if t.shape[0] > 1 else max(t.shape[1], 1))


def _ld(t):
    return t.stride(0) if t.shape[0] > 1 else max(t.shape[1], 1)


# ----------------------------------------------------------------- precision
# "fp32x3": tcgen05 with the 3xTF32 hi/lo split (fp32-equivalent products; the default -- the reference
#           trains with fp32 matmuls);  "tf32": single-pass tcgen05 kind::tf32 (opt-in, 10-bit mantissa
#           products);  "fp32": exact fp32 on the CUDA cores (k_gemm_simt, the oracle / fallback).
PRECISIONS = ("fp32x3", "tf32", "fp32")
_precision = os.environ.get("FEDTGAN_PRECISION", "fp32x3")
# fp32x3 splitter: 2 (default) = the stage as TMA wrote it is the hi operand (the tensor core drops the 13 low
# mantissa bits of an fp32 word) and only lo = rna_tf32(x - trunc_tf32(x)) is written; 1 = rewrite hi = rna_tf32(x)
# in place as well (one third more shared-memory traffic; measured 0.272 vs 0.260 ms per Adult iteration)
_x3_mode = int(os.environ.get("FEDTGAN_X3_MODE", "2"))
_simt_fallbacks = 0          # tensor-core launches that fell back to k_gemm_simt (unaligned operands)
_splitk_mode = os.environ.get("FEDTGAN_SPLITK", "cluster")     # "cluster" (DSMEM) | "workspace"
_scratch = {}
MAX_WS_BYTES = 64 << 20


def set_precision(p: str):
    global _precision
    assert p in PRECISIONS, p
    _precision = p


def get_precision() -> str:
    return _precision


def set_x3_mode(mode: int):
    """fp32x3 splitter variant: 1 = rewrite the stage as rna_tf32 hi + lo copy, 2 = keep the stage as the
    hi operand (the tensor core truncates fp32 words to tf32) and write only the lo copy."""
    global _x3_mode
    assert mode in (1, 2)
    _x3_mode = mode


def simt_fallback_count() -> int:
    """How many GEMMs asked for the tensor cores but ran on the CUDA-core kernel because an operand was
    not 16 B aligned / its leading dimension not a multiple of 4 floats.  Should stay 0."""
    return _simt_fallbacks


def _scratch_for(device, stream):
    key = (device.index, stream)
    if key not in _scratch:
        _scratch[key] = (torch.empty(MAX_WS_BYTES // 4, dtype=torch.float32, device=device),
                         torch.zeros(8192, dtype=torch.int32, device=device))
    return _scratch[key]


def preallocate_scratch(device):
    """Allocate the split-K scratch for the current stream (call before graph capture)."""
    return _scratch_for(torch.device(device), _stream())


# ---------------------------------------------------------------------- GEMM
def _gemm_args(a, b, out, a_t, b_t, bias, act, mask_in, mask_out, u, rng, stream, accumulate, alpha, slope, p):
    M, N = out.shape
    K = a.shape[0] if a_t else a.shape[1]
    if out.device.index != torch.cuda.current_device():
        # launches go to the *current* device's stream: a tensor of another GPU would be dereferenced there
        raise NativeError(f"tensor on cuda:{out.device.index} but current device is cuda:{torch.cuda.current_device()}")
    assert (a.shape[1] if a_t else a.shape[0]) == M, (a.shape, out.shape, a_t)
    assert (b.shape[1] if b_t else b.shape[0]) == N and (b.shape[0] if b_t else b.shape[1]) == K, (b.shape, N, K, b_t)
    g = GemmArgs()
    g.a, g.lda = _m(a)
    g.b, g.ldb = _m(b)
    g.d, g.ldd = _m(out)
    g.M, g.N, g.K = M, N, K
    g.a_t, g.b_t = int(a_t), int(b_t)
    g.bias = _p(bias)
    g.act = {None: 0, "leaky": 1, "leaky_dropout": 2}[act]
    g.mask_in, g.ld_mask_in = _m(mask_in)
    g.mask_out, g.ld_mask_out = _m(mask_out)
    g.drop_u, g.ld_drop_u = _m(u)
    g.rng_state = _p(rng.state) if rng is not None else None
    g.rng_stream = int(stream)
    g.alpha, g.slope, g.drop_p = float(alpha), float(slope), float(p)
    g.accumulate = int(accumulate)
    if g.act == 2 and u is None and rng is None:
        raise ValueError("dropout needs uniforms or an rng")
    return g


def linear(a, b, out, *, a_t=False, b_t=False, bias=None, act=None, mask_in=None, mask_out=None, u=None, rng=None,
           stream=0, accumulate=False, alpha=1.0, slope=0.2, p=0.5, precision=None, split_k=None, splitk_mode=None):
    splitk_mode = splitk_mode or _splitk_mode
    g = _gemm_args(a, b, out, a_t, b_t, bias, act, mask_in, mask_out, u, rng, stream, accumulate, alpha, slope, p)
    prec = precision or _precision
    st = _stream()
    if prec in ("tf32", "fp32x3"):
        g.x3 = _x3_mode if prec == "fp32x3" else 0
        ws = ctr = None
        if split_k is None and splitk_mode == "cluster":
            sk = _choose_cluster_split(g.M, g.N, g.K)        # DSMEM reduction inside a cluster: no scratch
        else:
            sk = split_k if split_k is not None else _choose_split(g.M, g.N, g.K)
        if sk > 1 and not (split_k is None and splitk_mode == "cluster"):
            wsb, ctrb = _scratch_for(out.device, st)
            while sk > 1 and _ws_bytes(g.M, g.N, sk) > MAX_WS_BYTES:
                sk -= 1
            if _num_tiles(g.M, g.N) > ctrb.numel():
                sk = 1
            ws, ctr = wsb.data_ptr(), ctrb.data_ptr()
        rc = _gemm_tc(C.byref(g), sk, ws, ctr, st)
        if rc == 0:
            return out
        if rc != UNSUPPORTED:
            raise NativeError(f"ftg_gemm_tc failed with code {rc}")
        global _simt_fallbacks
        _simt_fallbacks += 1
        if _simt_fallbacks == 1:
            import warnings

            warnings.warn(f"tcgen05 GEMM not applicable to M={g.M} N={g.N} K={g.K} lda={g.lda} ldb={g.ldb} "
                          "(pointers must be 16 B aligned, leading dimensions multiples of 4 floats): running the "
                          "CUDA-core fp32 kernel instead; see native.simt_fallback_count()")
    _ck(_gemm_simt(C.byref(g), st), "ftg_gemm_simt")
    return out


_fused_bn = os.environ.get("FEDTGAN_FUSED_BN", "1") != "0"


def linear_bn_relu(x, w, a_out, out, *, bias, gamma, beta, running_mean, running_var, save_mean=None, save_invstd=None,
                   eps=1e-5, momentum=0.1, num_batches_tracked=None, precision=None):
    """Residual block forward: ``a_out = x @ w.T + bias`` and ``out = relu(batchnorm_train(a_out))`` (``out`` may be a
    column window of a wider buffer).  One tcgen05 launch with the batch statistics reduced inside a thread-block
    cluster (``ftg_gemm_tc_bn``) when the shape allows it (batch <= 512, tensor-core precision); otherwise the GEMM
    followed by ``bn_relu_fwd``.  Reference: ``Residual.forward``, ``ctgan.py:40-44``."""
    prec = precision or _precision
    if _fused_bn and prec in ("tf32", "fp32x3") and _gemm_tc_bn_plan(a_out.shape[0], x.shape[1]) > 0:
        g = _gemm_args(x, w, a_out, False, False, bias, None, None, None, None, None, 0, False, 1.0, 0.2, 0.5)
        g.x3 = _x3_mode if prec == "fp32x3" else 0
        b = BnArgs()
        b.gamma, b.beta = _p(gamma), _p(beta)
        b.running_mean, b.running_var = _p(running_mean), _p(running_var)
        b.save_mean, b.save_invstd = _p(save_mean), _p(save_invstd)
        b.out, b.ldo = _m(out)
        b.eps, b.momentum = float(eps), float(momentum)
        b.nbt = _p(num_batches_tracked)
        rc = _gemm_tc_bn(C.byref(g), C.byref(b), _stream())
        if rc == 0:
            return out
        if rc != UNSUPPORTED:
            raise NativeError(f"ftg_gemm_tc_bn failed with code {rc}")
    linear(x, w, a_out, bias=bias, precision=precision)
    bn_relu_fwd(a_out, gamma, beta, running_mean, running_var, out, save_mean, save_invstd, training=True, eps=eps,
                momentum=momentum, num_batches_tracked=num_batches_tracked)
    return out


def set_fused_bn(on: bool):
    global _fused_bn
    _fused_bn = bool(on)


def set_gemm_stages(stages: int):
    """Force the tcgen05 GEMM's TMA ring depth (2..6); 0 restores the automatic policy."""
    lib.ftg_gemm_tc_set_stages(int(stages))


def set_cluster_push(on: bool):
    """In-cluster split-K reduction protocol of the tcgen05 GEMM: push (st.async, default) or pull."""
    lib.ftg_gemm_tc_set_cluster_push(int(bool(on)))


def set_gemm_persistent(on: bool):
    """Experimental: multi-wave unsplit GEMMs through the persistent double-buffered-TMEM kernel."""
    lib.ftg_gemm_tc_set_persistent(int(bool(on)))


def rowdot(a, w, bias, out, rowscale=None, loss_out=None):
    pa, lda = _m(a)
    if loss_out is not None:
        _ck(_rowdot_loss(pa, lda, a.shape[0], a.shape[1], _p(w), _p(bias), _p(out), _p(rowscale), _p(loss_out),
                         _stream()), "rowdot_loss")
    else:
        _ck(_rowdot(pa, lda, a.shape[0], a.shape[1], _p(w), _p(bias), _p(out), _stream()), "rowdot")
    return out


def d_head(a, w, bias, y, rs_loss, loss_out, rs_sig, mask, z):
    pa, lda = _m(a)
    pm, ldm = _m(mask)
    pz, ldz = _m(z)
    _ck(_d_head(pa, lda, a.shape[0], a.shape[1], _p(w), _p(bias), _p(y), _p(rs_loss), _p(loss_out), _p(rs_sig), pm,
                ldm, pz, ldz, _stream()), "d_head")
    return y


_dmid_counters = {}


def d_mid_supported(h0: int, h1: int) -> bool:
    """Shape check of the fused discriminator-middle kernel (hidden widths <= 512, H0 % 4 == 0)."""
    return bool(_d_mid_supported(int(h0), int(h1)))


def d_mid_enabled() -> bool:
    """Opt-in (``FEDTGAN_FUSED_DMID=1``): measured 26-30 us per launch inside the iteration's graph against ~24 us for
    the three launches it replaces (tcgen05 GEMM + k_d_head + tcgen05 GEMM), see profiles/gemm_cluster_protocols.md."""
    return os.environ.get("FEDTGAN_FUSED_DMID", "0") == "1"


def d_mid(a0, w1, b1, a1, mask1, mask0, w_head, b_head, y, rs_loss, loss_out, rs_sig, z1, z0, *, u=None, rng=None,
          stream=0, slope=0.2, p=0.5, counter=None):
    """Discriminator layer 1 + head + backward signals of both in one launch (exact fp32), see ``ftg_d_mid``."""
    pa0, lda0 = _m(a0)
    pw, ldw = _m(w1)
    pa1, lda1 = _m(a1)
    pm1, ldm1 = _m(mask1)
    pm0, ldm0 = _m(mask0)
    pz1, ldz1 = _m(z1)
    pz0, ldz0 = _m(z0)
    pu, ldu = _m(u)
    if counter is None:                      # callers that may overlap on the GPU pass their own counter
        dev = a0.device.index
        if dev not in _dmid_counters:
            _dmid_counters[dev] = torch.zeros(1, dtype=torch.int32, device=a0.device)
        counter = _dmid_counters[dev]
    M, H0 = a0.shape
    H1 = w1.shape[0]
    _ck(_d_mid(pa0, lda0, M, H0, H1, pw, ldw, _p(b1), pa1, lda1, pm1, ldm1, pm0, ldm0, _p(w_head), _p(b_head), _p(y),
               _p(rs_loss), _p(loss_out), _p(rs_sig), pz1, ldz1, pz0, ldz0, pu, ldu,
               _p(rng.state) if rng is not None else None, int(stream), float(slope), float(p),
               _p(counter), _stream()), "d_mid")
    return y


def colsum(a, out, rowscale=None, accumulate=False):
    pa, lda = _m(a)
    _ck(_colsum(pa, lda, a.shape[0], a.shape[1], _p(rowscale), _p(out), int(accumulate), _stream()), "colsum")
    return out


def scale_outer_mask(rowscale, v, mask, out):
    pm, ldm = _m(mask)
    po, ldo = _m(out)
    _ck(_som(_p(rowscale), _p(v), pm, ldm, po, ldo, out.shape[0], out.shape[1], _stream()), "scale_outer_mask")
    return out


def bn_relu_fwd(a, gamma, beta, running_mean, running_var, out, save_mean, save_invstd, *, training=True, eps=1e-5,
                momentum=0.1, num_batches_tracked=None):
    pa, lda = _m(a)
    po, ldo = _m(out)
    _ck(_bn_fwd(pa, lda, a.shape[0], a.shape[1], _p(gamma), _p(beta), _p(running_mean), _p(running_var), po, ldo,
                _p(save_mean), _p(save_invstd), int(training), float(eps), float(momentum),
                _p(num_batches_tracked), _stream()), "bn_relu_fwd")
    return out


def bn_relu_bwd(dout, a, save_mean, save_invstd, gamma, beta, da, dgamma, dbeta, dbias_fc=None):
    pd, ldd = _m(dout)
    pa, lda = _m(a)
    pda, ldda = _m(da)
    if dbias_fc is not None:
        _ck(_bn_bwd_bias(pd, ldd, pa, lda, a.shape[0], a.shape[1], _p(save_mean), _p(save_invstd), _p(gamma),
                         _p(beta), pda, ldda, _p(dgamma), _p(dbeta), _p(dbias_fc), _stream()), "bn_relu_bwd_bias")
    else:
        _ck(_bn_bwd(pd, ldd, pa, lda, a.shape[0], a.shape[1], _p(save_mean), _p(save_invstd), _p(gamma), _p(beta),
                    pda, ldda, _p(dgamma), _p(dbeta), _stream()), "bn_relu_bwd")
    return da


def activate_fwd(fake, tables, out, *, tau=0.2, u=None, rng=None, stream=0, cond=None, hard_noise=True,
                 slerp_real=None, slerp_out=None, alpha=None, slerp_stream=0):
    dd = tables["col_span"].numel()
    pf, ldf = _m(fake)
    po, ldo = _m(out)
    pu, ldu = _m(u)
    pc, ldc = _m(cond)
    n_opt = cond.shape[1] if cond is not None else 0
    if slerp_real is not None:
        pr, ldr = _m(slerp_real)
        pi, ldi = _m(slerp_out)
        _ck(_act_fwd_slerp(pf, ldf, fake.shape[0], dd, tables["span_start"].numel(), _p(tables["span_start"]),
                           _p(tables["span_width"]), _p(tables["span_kind"]), _p(tables["col_span"]), po, ldo,
                           float(tau), pu, ldu, _p(rng.state) if rng is not None else None, int(stream), pc, ldc,
                           n_opt, int(hard_noise), pr, ldr, pi, ldi, _p(alpha), int(slerp_stream), _stream()),
            "activate_fwd_slerp")
        return out
    _ck(_act_fwd(pf, ldf, fake.shape[0], dd, tables["span_start"].numel(), _p(tables["span_start"]),
                 _p(tables["span_width"]), _p(tables["span_kind"]), _p(tables["col_span"]), po, ldo, float(tau), pu,
                 ldu, _p(rng.state) if rng is not None else None, int(stream), pc, ldc, n_opt, int(hard_noise),
                 _stream()), "activate_fwd")
    return out


def activate_bwd(dout, act, tables, dfake, *, tau=0.2, accumulate=False):
    pd, ldd = _m(dout)
    pa, lda = _m(act)
    pf, ldf = _m(dfake)
    _ck(_act_bwd(pd, ldd, pa, lda, dout.shape[0], tables["span_start"].numel(), _p(tables["span_start"]),
                 _p(tables["span_width"]), _p(tables["span_kind"]), pf, ldf, float(tau), int(accumulate), _stream()),
        "activate_bwd")
    return dfake


def gen_output_bwd(dout, act, fake, cond, mask, tables, dfake, loss_out, *, tau=0.2):
    pd, ldd = _m(dout)
    pa, lda = _m(act)
    pdf, ldf = _m(dfake)
    pfk, ldfk = _m(fake)
    pc, ldc = _m(cond)
    pm, ldm = _m(mask)
    _ck(_gen_out_bwd(pd, ldd, pa, lda, dout.shape[0], tables["span_start"].numel(), _p(tables["span_start"]),
                     _p(tables["span_width"]), _p(tables["span_kind"]), pdf, ldf, float(tau), pfk, ldfk, pc, ldc, pm,
                     ldm, _p(tables["col_soft"]), _p(tables["soft_cstart"]), _p(loss_out), _stream()), "gen_output_bwd")
    return dfake


def cond_loss_fwd_bwd(fake, cond, mask, tables, dfake, loss_out, *, accumulate=True):
    pf, ldf = _m(fake)
    pc, ldc = _m(cond)
    pm, ldm = _m(mask)
    pdf, lddf = _m(dfake)
    dd = tables["col_span"].numel()
    _ck(_cond_loss(pf, ldf, pc, ldc, pm, ldm, fake.shape[0], dd, tables["soft_start"].numel(),
                   _p(tables["soft_start"]), _p(tables["soft_width"]), _p(tables["soft_cstart"]), pdf, lddf,
                   _p(loss_out), int(accumulate), _stream()), "cond_loss")
    return loss_out


def slerp(low, high, out, *, alpha=None, rng=None, stream=0):
    pl, ldl = _m(low)
    ph, ldh = _m(high)
    po, ldo = _m(out)
    _ck(_slerp(pl, ldl, ph, ldh, po, ldo, low.shape[0], low.shape[1], _p(alpha),
               _p(rng.state) if rng is not None else None, int(stream), _stream()), "slerp")
    return out


def gp_penalty(g, pen_out, *, lambda_=10.0):
    pg, ldg = _m(g)
    _ck(_gp(pg, ldg, g.shape[0], g.shape[1], _p(pen_out), float(lambda_), _stream()), "gp_penalty")
    return pen_out


def adam_step(param, grad, exp_avg, exp_avg_sq, step, *, lr, beta1, beta2, eps, weight_decay=0.0):
    """``step`` must be a view of the device rng/step state whose element [0] is the step
    count; the kernel reads ``state[1]`` so we pass ``step`` minus one int64."""
    assert torch.is_tensor(step) and step.dtype == torch.int64
    state_ptr = step.data_ptr() - 8
    _ck(_adam(_p(param), _p(grad), _p(exp_avg), _p(exp_avg_sq), param.numel(), state_ptr, float(lr), float(beta1),
              float(beta2), float(eps), float(weight_decay), _stream()), "adam")
    return param


def tick(state):
    _ck(_tick(_p(state), _stream()), "tick")


def fill_normal(out, *, z=None, rng=None, stream=0):
    if z is not None:
        out.copy_(z)
        return out
    po, ldo = _m(out)
    _ck(_fill_normal(po, out.shape[0], out.shape[1], ldo, _p(rng.state), int(stream), _stream()), "fill_normal")
    return out


def sample_condvec(cum_p, soft_width, col_out, opt_out, *, u_col=None, u_opt=None, rng=None, stream=0):
    _ck(_sample_condvec(_p(cum_p), cum_p.shape[0], cum_p.shape[1], _p(soft_width), _p(col_out), _p(opt_out),
                        col_out.shape[0], _p(u_col), _p(u_opt), _p(rng.state) if rng is not None else None,
                        int(stream), _stream()), "sample_condvec")
    return col_out, opt_out


def write_condvec(col, opt, soft_cstart, c_out, m_out=None, perm=None):
    pc, ldc = _m(c_out)
    pm, ldm = _m(m_out)
    n_col = m_out.shape[1] if m_out is not None else 0
    _ck(_write_condvec(_p(col), _p(opt), _p(soft_cstart), pc, ldc, c_out.shape[1], pm, ldm, n_col, _p(perm),
                       c_out.shape[0], _stream()), "write_condvec")
    return c_out


def argsort_perm(perm_out, *, u=None, rng=None, stream=0):
    _ck(_argsort(_p(perm_out), perm_out.shape[0], _p(u), _p(rng.state) if rng is not None else None, int(stream),
                 _stream()), "argsort_perm")
    return perm_out


def sample_rows(row_ptr, row_ids, soft_cstart, col, opt, rid_out, *, perm=None, u=None, rng=None, stream=0):
    _ck(_sample_rows(_p(row_ptr), _p(row_ids), _p(soft_cstart), _p(col), _p(opt), _p(perm), _p(rid_out),
                     rid_out.shape[0], _p(u), _p(rng.state) if rng is not None else None, int(stream), _stream()),
        "sample_rows")
    return rid_out


def expand_rows(vals, codes, rid, tables, out):
    po, ldo = _m(out)
    dd = tables["col_span"].numel()
    _ck(_expand_rows(_p(vals), vals.shape[1], _p(codes), codes.shape[1], _p(rid), _p(tables["tanh_start"]),
                     _p(tables["soft_start"]), po, ldo, dd, out.shape[0], _stream()), "expand_rows")
    return out


def prep_g_batch(table, tables, z_out, c_out, m_out, col, opt, rng, s_z, s_col):
    """Fused G-step preparation: (col, opt) draw, z ~ N(0,1), one-hot c1, column mask m1."""
    pz, ldz = _m(z_out)
    pc, ldc = _m(c_out)
    pm, ldm = _m(m_out)
    _ck(_prep_rows(0, _p(col), _p(opt), None, None, _p(tables["soft_cstart"]), pz, ldz, z_out.shape[1], pc, ldc,
                   c_out.shape[1], pm, ldm, m_out.shape[1], None, 0, None, None, None, None, 0, 0,
                   _p(table.cum_p), table.cum_p.shape[1], _p(tables["soft_width"]), z_out.shape[0], _p(rng.state),
                   int(s_z), int(s_col), _stream()), "prep_rows(G)")


def prep_d_batch(table, tables, z_out, c_out, real_out, col, opt, perm, rid, rng, s_z, s_col, s_perm, s_row,
                 tick=False):
    """Fused D-step preparation: select kernel + per-row writer (z, c1, real row | c2)."""
    B = z_out.shape[0]
    dd = tables["col_span"].numel()
    _ck(_prep_select(_p(table.cum_p), table.cum_p.shape[0], table.cum_p.shape[1], _p(tables["soft_width"]),
                     _p(tables["soft_cstart"]), _p(table.row_ptr), _p(table.row_ids), _p(col), _p(opt), _p(perm),
                     _p(rid), B, _p(rng.state), int(s_col), int(s_perm), int(s_row), int(tick), _stream()),
        "prep_select")
    pz, ldz = _m(z_out)
    pc, ldc = _m(c_out)
    pr, ldr = _m(real_out)
    _ck(_prep_rows(1, _p(col), _p(opt), _p(perm), _p(rid), _p(tables["soft_cstart"]), pz, ldz, z_out.shape[1], pc,
                   ldc, c_out.shape[1], None, 0, table.cum_p.shape[0], _p(table.vals), table.vals.shape[1],
                   _p(table.codes), _p(tables["tanh_start"]), _p(tables["soft_start"]), pr, ldr, dd,
                   _p(table.cum_p), table.cum_p.shape[1], _p(tables["soft_width"]), B, _p(rng.state), int(s_z),
                   int(s_col), _stream()), "prep_rows(D)")


_decode_cache = {}


def _decode_tables(tables, col_is_cont, cat_off, device):
    key = (id(tables), col_is_cont.data_ptr())
    if key not in _decode_cache:
        cic = col_is_cont.cpu().tolist()
        starts = tables["span_start"].cpu().tolist()
        widths = tables["span_width"].cpu().tolist()
        coff = cat_off.cpu().tolist()
        col_start, col_width, col_aux = [], [], []
        s = ci = ki = 0
        for c in cic:
            if c:
                col_start.append(starts[s])
                col_width.append(widths[s + 1])
                col_aux.append(ci)
                ci += 1
                s += 2
            else:
                col_start.append(starts[s])
                col_width.append(widths[s])
                col_aux.append(coff[ki])
                ki += 1
                s += 1
        mk = lambda v: torch.tensor(v, dtype=torch.int32, device=device)  # noqa: E731
        _decode_cache[key] = (mk(cic), mk(col_start), mk(col_width), mk(col_aux))
    return _decode_cache[key]


def vgm_decode(act, tables, mu, sd, cat_ids, cat_off, col_is_cont, out):
    cic, cs, cw, ca = _decode_tables(tables, col_is_cont, cat_off, act.device)
    pa, lda = _m(act)
    po, ldo = _m(out)
    k = mu.shape[1] if mu.dim() == 2 and mu.numel() else 1
    _ck(_vgm_decode(pa, lda, act.shape[0], cic.numel(), _p(cic), _p(cs), _p(cw), _p(ca), _p(mu), _p(sd), k,
                    _p(cat_ids), po, ldo, _stream()), "vgm_decode")
    return out


def vgm_encode(x, logit_const, mu, var, active, u, val_out, code_out, rng=None, stream=0):
    lc, m, v = (t[active].contiguous().float() for t in (logit_const, mu, var))
    _ck(_vgm_encode(_p(x), x.shape[0], _p(lc), _p(m), _p(v), lc.numel(), _p(u),
                    _p(rng.state) if rng is not None else None, int(stream), _p(val_out), val_out.stride(0),
                    _p(code_out), code_out.stride(0), _stream()), "vgm_encode")
    return val_out, code_out


# ------------------------------------------------------------ federated average
def _ptr_array(ptrs):
    arr = (C.c_void_p * len(ptrs))(*ptrs)
    return arr


def _weights_arg(weights, world):
    """Host list -> by-value float array; CUDA tensor -> its device pointer (the kernel reads the weights when it
    runs, e.g. straight from ``fed_stats_to_weights``)."""
    if torch.is_tensor(weights) and weights.is_cuda:
        assert weights.dtype == torch.float32 and weights.numel() >= world and weights.is_contiguous()
        return C.c_void_p(weights.data_ptr())
    return C.cast((C.c_float * world)(*[float(x) for x in weights]), C.c_void_p)


def fedavg_local(arenas, weights, int_offsets=()):
    """All replicas on this GPU: one launch over all ranks' data."""
    n = arenas[0].numel()
    pa = _ptr_array([a.data_ptr() for a in arenas])
    w = _weights_arg(weights, len(arenas))
    io = (C.c_int * max(len(int_offsets), 1))(*[int(x) for x in int_offsets])
    _ck(_fedavg_local(pa, len(arenas), w, n, io, len(int_offsets), _stream()), "fedavg_local")
    return arenas


def fedavg_twoshot(peer_ptrs, flag_ptrs, rank, weights, n, int_offsets, round_id, status, timeout_ms=20000):
    world = len(peer_ptrs)
    pa, fa = _ptr_array(peer_ptrs), _ptr_array(flag_ptrs)
    w = _weights_arg(weights, world)
    io = (C.c_int * max(len(int_offsets), 1))(*[int(x) for x in int_offsets])
    _ck(_fedavg_twoshot(pa, fa, rank, world, w, n, io, len(int_offsets), int(round_id), _p(status), int(timeout_ms),
                        _stream()), "fedavg_twoshot")


def fedavg_oneshot(peer_ptrs, flag_ptrs, staging, rank, weights, n, int_offsets, round_id, status, timeout_ms=20000):
    world = len(peer_ptrs)
    pa, fa = _ptr_array(peer_ptrs), _ptr_array(flag_ptrs)
    w = _weights_arg(weights, world)
    io = (C.c_int * max(len(int_offsets), 1))(*[int(x) for x in int_offsets])
    _ck(_fedavg_oneshot(pa, fa, _p(staging), rank, world, w, n, io, len(int_offsets), int(round_id), _p(status),
                        int(timeout_ms), _stream()), "fedavg_oneshot")


def fedavg_nvls(peer_ptrs, flag_ptrs, multicast_ptr, rank, weights, n, int_offsets, round_id, status, timeout_ms=20000):
    world = len(peer_ptrs)
    pa, fa = _ptr_array(peer_ptrs), _ptr_array(flag_ptrs)
    w = _weights_arg(weights, world)
    io = (C.c_int * max(len(int_offsets), 1))(*[int(x) for x in int_offsets])
    _ck(_fedavg_nvls(pa, fa, multicast_ptr, rank, world, w, n, io, len(int_offsets), int(round_id), _p(status),
                     int(timeout_ms), _stream()), "fedavg_nvls")


def symm_alloc(nbytes):
    ptr = _vp()
    _ck(_symm_alloc(int(nbytes), C.byref(ptr)), "symm_alloc")
    return ptr.value


def symm_free(ptr):
    _ck(_symm_free(ptr), "symm_free")


def symm_export(ptr) -> bytes:
    buf = C.create_string_buffer(64)
    _ck(_symm_export(ptr, buf), "symm_export")
    return buf.raw


def symm_open(handle: bytes):
    ptr = _vp()
    _ck(_symm_open(C.c_char_p(handle), C.byref(ptr)), "symm_open")
    return ptr.value


def symm_close(ptr):
    _ck(_symm_close(ptr), "symm_close")


def stats_to_weights(hist, cat_off, pool_sorted, client_sorted, smp_off, rows):
    """Device computation of (JSD_norm [W,n_cat], WD_norm [W,n_cont], weights [W])."""
    world = hist.shape[0]
    n_cat = cat_off.numel() - 1
    n_cont = pool_sorted.shape[0] if pool_sorted is not None and pool_sorted.numel() else 0
    n_total = pool_sorted.shape[1] if n_cont else 0
    dev = hist.device
    jsd = torch.zeros(world, max(n_cat, 1), dtype=torch.float32, device=dev)
    wd = torch.zeros(world, max(n_cont, 1), dtype=torch.float32, device=dev)
    w = torch.zeros(world, dtype=torch.float32, device=dev)
    _ck(_stats(_p(hist), world, n_cat, _p(cat_off), _p(pool_sorted) if n_cont else None, n_total,
               _p(client_sorted) if n_cont else None, _p(smp_off) if n_cont else None, n_cont, _p(rows), _p(jsd),
               _p(wd), _p(w), _stream()), "stats_to_weights")
    return jsd[:, :n_cat], wd[:, :n_cont], w


def sort_pad(n: int) -> int:
    """Row pitch (power of two >= 4096) of the sorted-sample buffers of ``gmm_sample_sort``."""
    return int(_sort_pad(int(n)))


def sort_rows(x):
    """Every row of ``x`` ([C, n] fp32 CUDA) sorted ascending by the in-kernel bitonic network -> ``[C, sort_pad(n)]``
    (padding +inf).  No library sort."""
    assert x.is_cuda and x.dtype == torch.float32 and x.is_contiguous() and x.dim() == 2
    Cn, n = x.shape
    out = torch.empty(Cn, sort_pad(n), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _ck(_sort_rows(_p(x), Cn, n, _p(out), _stream()), "sort_rows")
    return out


def quantile_centers(x, K):
    """``[C, K]`` quantiles (k+0.5)/K of every row of ``x``: the k-means seeds of ``vbgmm_fit``."""
    n = x.shape[1]
    srt = sort_rows(x)
    q = ((torch.arange(K, device=x.device, dtype=torch.float64) + 0.5) / K * (n - 1)).round().long()
    return srt[:, q].contiguous()


def gmm_sample_sort(gmm, n, seed, client_id, out=None):
    """Draw ``n`` samples from each of the C mixtures ``gmm`` ([C, 7, K] fp64 CUDA, ``vbgmm_fit`` layout) and sort
    them inside the kernel.  Returns ``[C, sort_pad(n)]`` fp32 (first ``n`` of a row ascending, padding +inf)."""
    assert gmm.is_cuda and gmm.dtype == torch.float64 and gmm.is_contiguous() and gmm.dim() == 3 and gmm.shape[1] == 7
    Cn, _, K = gmm.shape
    pad = sort_pad(n)
    if out is None:
        out = torch.empty(Cn, pad, dtype=torch.float32, device=gmm.device)
    assert out.shape == (Cn, pad) and out.is_contiguous()
    with torch.cuda.device(gmm.device):
        _ck(_gmm_sample_sort(_p(gmm), Cn, K, int(n), int(seed) & ((1 << 64) - 1), int(client_id), _p(out), _stream()),
            "gmm_sample_sort")
    return out


def fed_stats_to_weights(hist_ptrs, seg_ptrs, n_i, rows, cat_off, n_cont, device, K=10, want_pool=False, chunks=None):
    """Fused federated statistics -> aggregation weights.  ``hist_ptrs`` / ``seg_ptrs``: per-client
    device pointers (ints; local or peer-mapped) to the category histogram and to the sorted samples of
    ``gmm_sample_sort``.  Returns ``(jsd [W, n_cat], wd [W, n_cont], weights [W], centers [n_cont, K] | None,
    pool [n_cont, sum n_i] | None)`` -- all device tensors; ``weights`` can be handed to ``fedavg_*`` as is."""
    world = len(rows)
    n_cat = int(cat_off.numel()) - 1 if cat_off is not None else 0
    dev = torch.device(device)
    n_total = int(sum(int(x) for x in n_i)) if n_cont else 0
    if chunks is None:
        chunks = max(1, min(256, (n_total + 1023) // 1024))
    f32 = dict(dtype=torch.float32, device=dev)
    jsd = torch.zeros(world, max(n_cat, 1), **f32)
    wd = torch.zeros(world, max(n_cont, 1), **f32)
    w = torch.zeros(world, **f32)
    partial = torch.zeros(max(n_cont, 1) * chunks * world, **f32)
    centers = torch.zeros(n_cont, K, **f32) if n_cont else None
    pool = torch.empty(n_cont, n_total, **f32) if (n_cont and want_pool) else None
    hp = _ptr_array([int(x) for x in hist_ptrs]) if n_cat else None
    sp = _ptr_array([int(x) for x in seg_ptrs]) if n_cont else None
    ni = (C.c_int * world)(*[int(x) for x in (n_i if n_cont else [0] * world)])
    rw = (C.c_float * world)(*[float(x) for x in rows])
    with torch.cuda.device(dev):
        _ck(_fed_stats(hp, sp, ni, rw, world, n_cat, _p(cat_off) if n_cat else None, int(n_cont), _p(partial), chunks,
                       _p(centers), K, _p(pool), _p(jsd), _p(wd), _p(w), _stream()), "fed_stats_to_weights")
    return jsd[:, :n_cat], wd[:, :n_cont], w, centers, pool


def vbgmm_fit(x, init_centers, dp_prior=1e-3, max_iter=100, tol=1e-3, reg_covar=1e-6, lloyd_iters=50):
    """Fit a K-component DP-GMM to every row of ``x`` ([C, n] fp32 CUDA); returns ([C, 7, K] fp64, iterations)."""
    assert x.is_cuda and x.dtype == torch.float32 and x.is_contiguous()
    Cn, n = x.shape
    K = init_centers.shape[1]
    ic = init_centers.to(device=x.device, dtype=torch.float32).contiguous()
    out = torch.zeros(Cn, 7, K, dtype=torch.float64, device=x.device)
    iters = torch.zeros(Cn, dtype=torch.int32, device=x.device)
    with torch.cuda.device(x.device):
        _ck(_vbgmm(_p(x), Cn, n, _p(ic), K, float(dp_prior), int(max_iter), float(tol), float(reg_covar),
                   int(lloyd_iters), _p(out), _p(iters), _stream()), "vbgmm_fit")
    return out, iters
