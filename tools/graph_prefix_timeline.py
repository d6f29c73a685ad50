"""Per-kernel latency INSIDE the CUDA graph of one training iteration (no profiler).

The iteration is a chain of ~40 dependent launches; eager timings are dominated by Python launch overhead and ncu
serialises with cold caches.  Here the iteration is captured on ONE stream (no parallel branches) truncated after its
first k ops, for k = 1..n, each prefix graph is replayed 300 times between CUDA events, and the marginal time
T(k) - T(k-1) is what op k costs in the chain with programmatic dependent launch on.  Also reports the full iteration
with and without the parallel branches.  Writes gpurun_out/graph_prefix_<precision>.json.

    python tools/graph_prefix_timeline.py [fp32x3|tf32]
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fed_tgan_b200 import ops  # noqa: E402
from fed_tgan_b200.config import CTGANConfig  # noqa: E402
from fed_tgan_b200.engine import DeviceTable, TrainEngine  # noqa: E402
from fed_tgan_b200.features.layout import TableLayout  # noqa: E402
from fed_tgan_b200.ops import native  # noqa: E402

prec = sys.argv[1] if len(sys.argv) > 1 else "fp32x3"
native.set_precision(prec)
dev = torch.device("cuda:0")
rng = np.random.default_rng(0)
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

budget = {"left": 10 ** 9, "names": []}


def gate(name, fn):
    def inner(*a, **k):
        if budget["left"] <= 0:
            return None
        budget["left"] -= 1
        tag = name
        if name == "linear":
            o = a[2]
            K = a[0].shape[0] if k.get("a_t") else a[0].shape[1]
            tag = f"linear M{o.shape[0]} N{o.shape[1]} K{K} {'T' if k.get('a_t') else 'N'}{'T' if k.get('b_t') else 'N'}"
        elif name == "linear_bn_relu":
            tag = f"linear_bn_relu M{a[2].shape[0]} N{a[2].shape[1]} K{a[0].shape[1]}"
        budget["names"].append(tag)
        return fn(*a, **k)
    return inner


for n in ops._NAMES:
    setattr(ops, n, gate(n, getattr(ops, n)))
native.prep_d_batch = gate("prep_d_batch (select + rows)", native.prep_d_batch)
native.prep_g_batch = gate("prep_g_batch", native.prep_g_batch)


def graph_time(k, concurrent):
    eng.concurrent = concurrent
    s = torch.cuda.Stream(device=dev)
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        native.preallocate_scratch(dev)
        budget["left"], budget["names"] = k, []
        eng.iteration()
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        budget["left"], budget["names"] = k, []
        with torch.cuda.graph(g, stream=s):
            eng.iteration()
        names = list(budget["names"])
        for _ in range(20):
            g.replay()
        s.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(300):
            g.replay()
        e1.record(s)
        s.synchronize()
    return e0.elapsed_time(e1) / 300 * 1e3, names


full_us, names = graph_time(10 ** 9, False)
n_ops = len(names)
rows_out, prev = [], 0.0
for k in range(1, n_ops + 1):
    t, nm = graph_time(k, False)
    rows_out.append({"k": k, "op": nm[-1], "prefix_us": round(t, 2), "marginal_us": round(t - prev, 2)})
    prev = t
par_us, _ = graph_time(10 ** 9, True)
res = {"precision": prec, "ops": n_ops, "single_stream_iteration_us": round(full_us, 2),
       "with_parallel_branches_us": round(par_us, 2), "per_op": rows_out}
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", f"graph_prefix_{prec}.json"), "w"), indent=1)
print(json.dumps({k: v for k, v in res.items() if k != "per_op"}))
for r in rows_out:
    print(f"{r['k']:3d} {r['marginal_us']:7.2f} us  {r['op']}")
