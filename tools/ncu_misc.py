"""One launch of every kernel family that the training-iteration capture (tools/ncu_iter.py) does not reach,
between cudaProfilerStart/Stop, for `ncu --profile-from-start off`:

  k_vbgmm_fit, k_gmm_sample_sort (sampler + in-kernel bitonic sort, and the plain row sort), k_fed_jsd / k_fed_wd /
  k_fed_finish (fused statistics -> weights), k_distances / k_finish (older host-gathered variant),
  k_fedavg_local, k_fedavg_twoshot / k_fedavg_oneshot (world = 1: same code path incl. flag barriers),
  k_gemm_tc_persist (persistent double-buffered-TMEM GEMM), k_gemm_tc<..., EPI_BN> (fused BatchNorm epilogue, tf32 and
  fp32x3), k_gemm_simt, k_d_mid.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fed_tgan_b200.features.gmm import GMMParams  # noqa: E402
from fed_tgan_b200.ops import native as N  # noqa: E402
from fed_tgan_b200.parallel.devstats import local_statistics  # noqa: E402
from fed_tgan_b200.parallel.symm import SymmetricBlock  # noqa: E402

dev = torch.device("cuda:0")
rng = np.random.default_rng(0)
K = 10
one = np.ones(K)


def gm(mus, sds, ws):
    pad = K - len(mus)
    return GMMParams(np.array(list(mus) + [0.0] * pad), np.array([s * s for s in sds] + [1.0] * pad),
                     np.array(list(ws) + [0.0] * pad), one, one, one, one)


x = torch.as_tensor(np.stack([np.concatenate([rng.normal(0, 1, 12000), rng.normal(8, 0.5, 6000), rng.normal(-6, 2, 6421)])
                              for _ in range(6)]), dtype=torch.float32).to(dev).contiguous()
rows = [24421, 24421]
models = [[gm([0.0, 4.0], [1.0, 0.5], [0.7, 0.3]) for _ in range(6)] for _ in range(2)]
cv = [[rng.integers(1, 500, 9).astype(np.float64) for _ in range(2)] for _ in range(8)]
reps = [torch.randn(1_005_472, device=dev) for _ in range(2)]
blk = SymmetricBlock(1_005_472, dev)
blk.connect(0, [blk.export_handle()])
a, b = torch.randn(8192, 508, device=dev), torch.randn(256, 508, device=dev)
d, bias = torch.empty(8192, 256, device=dev), torch.randn(256, device=dev)
xb, wb = torch.randn(500, 252, device=dev), torch.randn(256, 252, device=dev) * 0.1
ab, hb = torch.empty(500, 256, device=dev), torch.empty(500, 256, device=dev)
g_, b_ = torch.ones(256, device=dev), torch.zeros(256, device=dev)
rm, rv = torch.zeros(256, device=dev), torch.ones(256, device=dev)


from fed_tgan_b200.ops import reference as Rf  # noqa: E402

dm = {"a0": torch.randn(150, 256, device=dev), "w1": torch.randn(256, 256, device=dev) * 0.1,
      "b1": torch.randn(256, device=dev), "wh": torch.randn(256, device=dev), "bh": torch.randn(1, device=dev),
      "m0": torch.ones(150, 256, device=dev), "rs": torch.randn(150, device=dev), "y": torch.zeros(150, device=dev),
      "loss": torch.zeros(1, device=dev), "rng": Rf.RngState(1, dev)}
for k_ in ("a1", "m1", "z1", "z0"):
    dm[k_] = torch.empty(150, 256, device=dev)


def run(rid):
    cen = N.quantile_centers(x, K)                                   # k_gmm_sample_sort (row-sort mode)
    N.vbgmm_fit(x, cen)                                              # k_vbgmm_fit
    local_statistics(models, list(range(6)), cv, rows, dev, seed=1)  # k_gmm_sample_sort, k_fed_jsd, k_fed_wd, k_fed_finish
    N.fedavg_local(reps, [0.4, 0.6], [7])
    blk.fedavg([1.0], 1_005_472, [7], rid, mode="twoshot")
    blk.fedavg([1.0], 1_005_472, [7], rid + 1, mode="oneshot")
    N.set_gemm_persistent(True)
    N.linear(a, b, d, bias=bias, precision="tf32")                   # k_gemm_tc_persist
    N.set_gemm_persistent(False)
    for prec in ("tf32", "fp32x3"):                                  # fused BatchNorm epilogue
        N.linear_bn_relu(xb, wb, ab, hb, bias=bias, gamma=g_, beta=b_, running_mean=rm, running_var=rv, precision=prec)
    N.linear(xb, wb, ab, bias=bias, precision="fp32")                # k_gemm_simt
    N.d_mid(dm["a0"], dm["w1"], dm["b1"], dm["a1"], dm["m1"], dm["m0"], dm["wh"], dm["bh"], dm["y"], dm["rs"], dm["loss"],
            dm["rs"], dm["z1"], dm["z0"], rng=dm["rng"], stream=3)   # k_d_mid (fused discriminator middle)


run(1)
torch.cuda.synchronize()
torch.cuda.profiler.start()
run(3)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
blk.check()
print("done")
