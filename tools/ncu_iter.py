"""Two eager training iterations (+ one 8192-row generation chunk) between
cudaProfilerStart/Stop, for `ncu --profile-from-start off`."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fed_tgan_b200.config import CTGANConfig  # noqa: E402
from fed_tgan_b200.engine import DeviceTable, TrainEngine  # noqa: E402
from fed_tgan_b200.engine.generate import RowGenerator  # noqa: E402
from fed_tgan_b200.features.layout import TableLayout  # noqa: E402
from fed_tgan_b200.ops import native  # noqa: E402

prec = sys.argv[1] if len(sys.argv) > 1 else "tf32"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 2
native.set_precision(prec)
dev = torch.device("cuda:0")
rng = np.random.default_rng(0)
# Adult-like layout: 6 continuous (1+~5 modes) + 8 categorical -> data_dim 128, n_opt 122
info = []
for k in (5, 6, 5, 7, 6, 5):
    info += [(1, "tanh"), (k, "softmax")]
for k in (9, 16, 7, 15, 6, 5, 2, 2):
    info += [(k, "softmax")]
lay = TableLayout(info)
rows = 20000
vals = rng.uniform(-0.9, 0.9, (rows, lay.n_tanh)).astype(np.float32)
codes = np.stack([rng.integers(0, w, rows) for w in lay.soft_width], 1).astype(np.int32)
cfg = CTGANConfig()
table = DeviceTable(vals, codes, lay, dev)
eng = TrainEngine(cfg, lay, table, dev)
dec = {"mu": np.zeros((6, 10), np.float32), "sd": np.ones((6, 10), np.float32),
       "cat_ids": np.arange(62, dtype=np.float32), "cat_off": np.cumsum([0, 9, 16, 7, 15, 6, 5, 2, 2]).astype(np.int32),
       "col_is_cont": np.array([1] * 6 + [0] * 8, np.int32)}
gen = RowGenerator(cfg, lay, dec, table.cum_emp, dev, chunk=8192)
for _ in range(3):
    eng.iteration()
gen.generate(eng.G, 8192)
torch.cuda.synchronize()
torch.cuda.profiler.start()
for _ in range(iters):
    eng.iteration()
gen.generate(eng.G, 8192)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("done", eng.losses())
