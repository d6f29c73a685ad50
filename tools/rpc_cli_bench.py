"""Federated epochs/s through the north-star entry point: ``python -m dtds.distributed`` (torch-RPC star: rank 0 =
federator, ranks 1..N = clients on GPUs 0..N-1), the same command line as the reference's README.

    python tools/rpc_cli_bench.py <clients> [epochs] [config] [extra CLI flags ...]

Writes the synthetic table of the named shape to a temp CSV, launches the whole world on this box
(``-rank`` omitted), every client keeps its own IID shard (``--split iid``), and reports the per-round time from
``timestamp_experiment.csv`` (time between consecutive epoch CSVs on disk; the first 2 rounds are dropped) next
to metrics.jsonl's host-side phase times.  Prints one JSON line; also usable with ``--cpu`` for config 1.
"""
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fed_tgan_b200.data.schemas import get_schema, make_table  # noqa: E402

clients = int(sys.argv[1]) if len(sys.argv) > 1 else 2
epochs = int(sys.argv[2]) if len(sys.argv) > 2 else 12
config = sys.argv[3] if len(sys.argv) > 3 else "adult"
extra = sys.argv[4:]
schema = get_schema(config)
base = tempfile.mkdtemp(prefix="rpc_cli_")
csv = os.path.join(base, f"{schema.name}_train.csv")
make_table(schema, seed=0).to_csv(csv, index=False)
port = 7788 + (os.getpid() % 500)
cmd = [sys.executable, "-m", "dtds.distributed", "-ip", "127.0.0.1", "-port", str(port), "-world_size", str(clients + 1),
       "-epoch", str(epochs), "-datapath", csv, "-selected_variables", ",".join(schema.columns),
       "-categorical_list", ",".join(schema.categorical), "-nonnegative_list", ",".join(schema.nonnegative) or ",",
       "-target_column", schema.target, "--split", "iid", "--out_dir", base, "--vgm_backend", "vb", *extra]
t0 = time.time()
res = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=3000)
wall = time.time() - t0
out = {"entry": "python -m dtds.distributed (RPC star)", "clients": clients, "epochs": epochs, "config": config,
       "rc": res.returncode, "wall_s": round(wall, 1), "extra": extra}
ts = os.path.join(base, "timestamp_experiment.csv")
if res.returncode == 0 and os.path.exists(ts):
    times = [float(x) for x in open(ts).read().split()]
    steady = times[2:] if len(times) > 4 else times
    out.update({"per_round_s": [round(t, 4) for t in times], "federated_epochs_per_s": len(steady) / sum(steady),
                "ms_per_round": 1e3 * sum(steady) / len(steady)})
    mj = os.path.join(base, "metrics.jsonl")
    if os.path.exists(mj):
        rows = [json.loads(l) for l in open(mj) if l.strip()]
        rows = [r for r in rows if "train_s" in r][2:]
        if rows:
            out["host_phase_ms"] = {k: round(1e3 * sum(r[k] for r in rows) / len(rows), 3)
                                    for k in ("train_s", "aggregate_s", "distribute_s", "total_s")}
else:
    out["tail"] = res.stdout[-1500:]
print(json.dumps(out))
