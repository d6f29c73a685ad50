"""Process entry for the UNMODIFIED reference (baseline/_ref/dtds): applies import-time
compatibility shims that the image needs, then calls the reference's own ``run()``
(``dtds/distributed.py:838``) -- its public entry point -- with the same arguments its
``__main__`` would pass.  Nothing of fed_tgan_b200 is imported here.

Shims (none touches the measured work):
  * ``pickle5``                          -> stdlib ``pickle`` (module alias)
  * ``np.float_`` / ``np.complex_``      -> ``np.float64`` / ``np.complex128`` (removed in NumPy 2)
  * ``rpc.BackendType.PROCESS_GROUP`` /
    ``rpc.ProcessGroupRpcBackendOptions`` -> TensorPipe equivalents (removed from torch)
"""
import json
import os
import pickle
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")


def apply_shims():
    sys.modules.setdefault("pickle5", pickle)
    import numpy as np

    if not hasattr(np, "float_"):
        np.float_ = np.float64
    if not hasattr(np, "complex_"):
        np.complex_ = np.complex128
    import torch.distributed.rpc as rpc

    if not hasattr(rpc.BackendType, "PROCESS_GROUP"):
        rpc.BackendType.PROCESS_GROUP = rpc.BackendType.TENSORPIPE
    if not hasattr(rpc, "ProcessGroupRpcBackendOptions"):
        def _pg_options(num_send_recv_threads=8, rpc_timeout=600, init_method=None):
            return rpc.TensorPipeRpcBackendOptions(num_worker_threads=num_send_recv_threads,
                                                   rpc_timeout=rpc_timeout, init_method=init_method)
        rpc.ProcessGroupRpcBackendOptions = _pg_options


def main():
    spec = json.loads(sys.argv[1])
    # repo root must NOT shadow the reference's `dtds` package
    sys.path = [p for p in sys.path if os.path.abspath(p or ".") != os.path.dirname(HERE)]
    sys.path.insert(0, REF)
    os.chdir(spec["cwd"])
    apply_shims()
    import dtds.distributed as ref

    assert os.path.abspath(ref.__file__).startswith(REF), ref.__file__
    ref.run(rank=spec["rank"], einterval=1, world_size=spec["world_size"], ip="127.0.0.1", port=spec["port"],
            name="Intrusion_train", datapath="data/raw/Intrusion_train.csv",
            selected_variables=spec["columns"], categorical_list=spec["categorical"],
            nonnegative_list=spec["nonnegative"], date_dic={}, target_column=spec["target"], problem_type="",
            epochs=spec["epochs"], report=False)


if __name__ == "__main__":
    main()
