"""Single-GPU launches of the aggregation kernels for ncu (the multi-rank path must not run under
ncu): two-shot with world=1 (barriers pass trivially; loads/stores hit this GPU's own arena) and the
all-replicas-on-one-GPU variant with 4 replicas."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fed_tgan_b200.ops import native  # noqa: E402
from fed_tgan_b200.parallel.symm import SymmetricBlock  # noqa: E402

n = 42_000_000
blk = SymmetricBlock(n, "cuda:0")
blk.connect(0, [blk.export_handle()])
blk.tensor.normal_()
arenas = [torch.randn(8_000_000, device="cuda:0") for _ in range(4)]
for r in range(1, 3):
    blk.fedavg([1.0], n, [7], r)
    native.fedavg_local(arenas, [0.1, 0.2, 0.3, 0.4], [7])
torch.cuda.synchronize()
torch.cuda.profiler.start()
blk.fedavg([1.0], n, [7], 3)
native.fedavg_local(arenas, [0.1, 0.2, 0.3, 0.4], [7])
torch.cuda.synchronize()
torch.cuda.profiler.stop()
blk.check()
print("done")
