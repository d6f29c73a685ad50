"""Where does a federated round spend its time?  (CUDA-event timing, no profiler.)

Phases: per-op breakdown of one eager iteration, graph-replayed iteration, 40 000-row
generation, D2H, decode + CSV.  Writes gpurun_out/profile_step.json.
"""
import collections
import json
import os
import sys
import tempfile
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from fed_tgan_b200 import ops  # noqa: E402
from fed_tgan_b200.config import CTGANConfig, wide_config  # noqa: E402
from fed_tgan_b200.data.schemas import get_schema, make_table  # noqa: E402
from fed_tgan_b200.ops import native  # noqa: E402
from fed_tgan_b200.parallel.client import MDGANClient  # noqa: E402
from fed_tgan_b200.parallel.federator import MDGANServer  # noqa: E402

prec = sys.argv[1] if len(sys.argv) > 1 else "tf32"
cfgname = sys.argv[2] if len(sys.argv) > 2 else "adult"
native.set_precision(prec)
dev = torch.device("cuda:0")
schema = get_schema(cfgname, rows=48842 if cfgname == "adult" else 20000)
df = make_table(schema, seed=0)
cfg = wide_config() if cfgname == "wide" else CTGANConfig()
client = MDGANClient(datapath=f"{schema.name}_train.csv", selected_variables=schema.columns,
                     categorical_list=schema.categorical, nonnegative_list=schema.nonnegative, date_dic={},
                     target_column=schema.target, problem_type="", epochs=1, device=dev, config=cfg, dataframe=df,
                     write_side_files=False, client_id=0, vgm_backend="vb")
out_dir = tempfile.mkdtemp()
server = MDGANServer([client], epochs=1, config=cfg, name=schema.name, out_dir=out_dir, vgm_backend="vb", device=dev)
server.uniform_meta_category(); server.uniform_continuous_gmm(); server.refit_local_transformer()
server.calculate_final_weights_for_aggregation(); server.server_local_synthesizer_initialization()
eng = client.engine
res = {"precision": prec, "config": cfgname, "data_dim": eng.layout.data_dim, "n_opt": eng.layout.n_opt,
       "Din": eng.Din, "G_params": eng.G.n_trainable, "D_params": eng.D.n_trainable}

# ---- per-op breakdown (eager)
records = collections.defaultdict(list)
orig = {}


def wrap(name, fn):
    def inner(*a, **k):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        out = fn(*a, **k)
        e.record()
        tag = name
        if name == "linear":
            o = a[2]
            K = a[0].shape[0] if k.get("a_t") else a[0].shape[1]
            tag = f"linear M{o.shape[0]} N{o.shape[1]} K{K} {'T' if k.get('a_t') else 'N'}{'T' if k.get('b_t') else 'N'}"
        records[tag].append((s, e))
        return out
    return inner


for n in ops._NAMES:
    orig[n] = getattr(ops, n)
for _ in range(3):
    eng.iteration()
torch.cuda.synchronize()
for n in ops._NAMES:
    setattr(ops, n, wrap(n, orig[n]))
for _ in range(5):
    eng.iteration()
torch.cuda.synchronize()
for n in ops._NAMES:
    setattr(ops, n, orig[n])
rows = []
for tag, evs in records.items():
    ms = [s.elapsed_time(e) for s, e in evs]
    rows.append((tag, len(evs) / 5, float(np.median(ms)) * 1e3))
rows.sort(key=lambda r: -r[1] * r[2])
res["per_op_us"] = [{"op": t, "calls_per_iter": c, "median_us": round(u, 1), "total_us": round(c * u, 1)} for t, c, u in rows]
res["eager_sum_us"] = round(sum(c * u for _, c, u in rows), 1)

# ---- eager vs graph iteration
def timeit(fn, n):
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / n

res["eager_iter_ms"] = timeit(eng.iteration, 20)
from fed_tgan_b200.engine.graph import GraphedIteration  # noqa: E402
g = GraphedIteration(eng)
res["graph_launches_per_iter"] = g.launches_per_replay
g.run(10)
res["graph_iter_ms"] = timeit(lambda: g.graph.replay(), 200)

# ---- generation / output
gen_ms = timeit(lambda: server.sample(cfg.sample_rows), 5)
res["generate_40000_ms"] = gen_ms
tab = server.sample(cfg.sample_rows)
torch.cuda.synchronize()
t0 = time.time(); host = tab.cpu().numpy(); res["d2h_ms"] = (time.time() - t0) * 1e3
t0 = time.time(); server._decode_and_write(host, "prof"); res["decode_csv_ms"] = (time.time() - t0) * 1e3
t0 = time.time(); server.sample_data("prof2", wait=True); res["sample_data_sync_ms"] = (time.time() - t0) * 1e3
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
with open(os.path.join(ROOT, "gpurun_out", f"profile_step_{prec}_{cfgname}.json"), "w") as fh:
    json.dump(res, fh, indent=1)
print(json.dumps({k: v for k, v in res.items() if k != "per_op_us"}))
for r in res["per_op_us"][:40]:
    print(f"{r['total_us']:9.1f} us  {r['calls_per_iter']:5.1f} x {r['median_us']:8.1f}  {r['op']}")
