"""2+-GPU check of the peer-reading statistics kernels (run under torchrun, one rank per GPU).

Every rank publishes ONLY its own category histogram and sorted GMM sample in a CUDA-IPC block; the fused
kernels of every rank read all peers' blocks over NVLink (devstats.spmd_statistics).  Checks: (i) the weights
are bit-identical on every rank, (ii) they equal what one process computes when it builds all clients'
samples locally (devstats.local_statistics: same Philox streams), (iii) they match the numpy/scipy
restatement of the reference (Server/dtds/distributed.py:592-783) on the device-drawn samples.
"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fed_tgan_b200.features.gmm import GMMParams  # noqa: E402
from fed_tgan_b200.parallel import weights as W  # noqa: E402
from fed_tgan_b200.parallel.devstats import local_statistics, spmd_statistics  # noqa: E402

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
K = 10
one = np.ones(K)
rng = np.random.default_rng(7)


def gm(mus, sds, ws):
    pad = K - len(mus)
    return GMMParams(np.array(list(mus) + [0.0] * pad), np.array([s * s for s in sds] + [1.0] * pad),
                     np.array(list(ws) + [0.0] * pad), one, one, one, one)


rows = [int(3000 + 1700 * i) for i in range(world)]
models = [[None, gm([0.3 * i], [1.0 + 0.1 * i], [1.0]), gm([1.0, 5.0 + 0.2 * i], [0.5, 1.0], [0.6, 0.4]), None] for i in range(world)]
count_vectors = [[rng.integers(1, 500, 6).astype(np.float64) for _ in range(world)],
                 [rng.integers(0, 90, 3).astype(np.float64) + 1 for _ in range(world)]]
st = spmd_statistics(rank, models[rank], [1, 2], count_vectors, rows, dev, seed=5, K=K)
jsd, wd, w = st.host()
mine = st.weights.clone()
allw = [torch.empty_like(mine) for _ in range(world)]
dist.all_gather(allw, mine)
identical = all(torch.equal(allw[0], x) for x in allw)
ok = identical
detail = {"world": world, "weights": [float(x) for x in w], "identical_on_all_ranks": identical}
if rank == 0:
    loc = local_statistics(models, [1, 2], count_vectors, rows, dev, seed=5, K=K)
    same_as_local = torch.equal(loc.weights, st.weights) and torch.equal(loc.wd, st.wd) and torch.equal(loc.jsd, st.jsd)
    host = W.final_weights(jsd, wd, rows)
    detail.update({"equals_single_process_build": bool(same_as_local), "max_abs_diff_vs_numpy_final_weights": float(np.abs(host - w).max())})
    ok = ok and same_as_local and float(np.abs(host - w).max()) < 1e-5
    print(json.dumps(detail), flush=True)
    print("ALL OK" if ok else "FAILED", flush=True)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
