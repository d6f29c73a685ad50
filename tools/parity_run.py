"""Statistical parity: Avg_JSD / Avg_WD per epoch, ours vs the unmodified reference, same
synthetic table and IID split, 2 clients, both arms on one GPU.  Writes
gpurun_out/parity_<config>.json.

    python tools/parity_run.py [epochs] [config] [seeds]

Environment: PARITY_CORRELATED=<float> makes the target column depend on four feature columns and
adds the ML-utility gap (difference in weighted F1, real-trained minus synthetic-trained, mean over
LR/DT/RF/MLP, tested on a held-out table) of the last epoch's CSV; PARITY_TAG names the output file;
PARITY_ARMS=ours,reference selects arms.  Without a GPU both arms run on the CPU (ours through the
torch restatement of the kernels, fp32).
"""
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "baseline"))

import numpy as np  # noqa: E402
import torch  # noqa: E402

from fed_tgan_b200.config import CTGANConfig  # noqa: E402
from fed_tgan_b200.data.schemas import get_schema, make_table, split_table  # noqa: E402
from fed_tgan_b200.eval import stat_sim_normalize  # noqa: E402
from fed_tgan_b200.parallel.client import MDGANClient  # noqa: E402
from fed_tgan_b200.parallel.federator import MDGANServer  # noqa: E402
from fed_tgan_b200.parallel.launcher import federate  # noqa: E402
import ref_harness  # noqa: E402

epochs = int(sys.argv[1]) if len(sys.argv) > 1 else 10
config = sys.argv[2] if len(sys.argv) > 2 else "adult"
seeds = int(sys.argv[3]) if len(sys.argv) > 3 else 2
n_clients = 2
schema = get_schema(config)
correlated = float(os.environ.get("PARITY_CORRELATED", "0"))
tag = os.environ.get("PARITY_TAG", config)
arms = os.environ.get("PARITY_ARMS", "ours,reference").split(",")
df = make_table(schema, seed=0, correlated=correlated)
df_test = make_table(schema, seed=1, rows=10000, correlated=correlated) if correlated > 0 else None
shards = split_table(df, n_clients, "iid", seed=0)
dev = "cuda:0" if torch.cuda.is_available() else "cpu"
precision = None
if dev != "cpu":
    from fed_tgan_b200.ops import native  # noqa: E402

    native.set_precision(os.environ.get("PARITY_PRECISION", native.get_precision()))
    precision = native.get_precision()
out = {"config": config, "epochs": epochs, "clients": n_clients, "rows": len(df), "device": dev, "precision_ours": precision,
       "correlated_target": correlated, "ours": [], "reference": []}


def utility_gap(result_dir, name, arm="ours", seed=0):
    if df_test is None:
        return None
    import shutil

    import pandas as pd

    last = os.path.join(result_dir, f"{name}_synthesis_epoch_{epochs - 1}.csv")
    save = os.environ.get("PARITY_SAVE_LAST")
    if save:                       # keep the CSV so that the (CPU-only) utility scoring can run off the GPU box
        os.makedirs(save, exist_ok=True)
        shutil.copy(last, os.path.join(save, f"{tag}_{arm}_seed{seed}_epoch{epochs - 1}.csv"))
    if os.environ.get("PARITY_SKIP_UTILITY") == "1":
        return None

    from fed_tgan_b200.eval.utility import utility_difference

    fake = pd.read_csv(os.path.join(result_dir, f"{name}_synthesis_epoch_{epochs - 1}.csv"))
    fake.columns = list(df.columns)
    diff, f1 = utility_difference(df, df_test, fake, schema.categorical, schema.target, verbose=False)
    return {"delta_f1": f1, "delta_acc_f1_per_model": np.asarray(diff).tolist()}


def curves(result_dir, name):
    res = []
    for i in range(epochs):
        p = os.path.join(result_dir, f"{name}_synthesis_epoch_{i}.csv")
        res.append(stat_sim_normalize(df, p, schema.categorical))
    return res


for seed in range(seeds if "ours" in arms else 0):
    # ---------------- ours
    base = tempfile.mkdtemp(prefix="parity_ours_")
    cfg = CTGANConfig(seed=seed)
    t0 = time.time()
    clients = [MDGANClient(datapath=f"{schema.name}_train.csv", selected_variables=schema.columns,
                           categorical_list=schema.categorical, nonnegative_list=schema.nonnegative, date_dic={},
                           target_column=schema.target, problem_type="", epochs=epochs, device=dev, config=cfg,
                           dataframe=sh, write_side_files=False, client_id=i, vgm_backend="sklearn")
               for i, sh in enumerate(shards)]
    server = MDGANServer(clients, epochs=epochs, config=cfg, name=schema.name, out_dir=base, device=dev)
    federate(server, verbose=False)
    c = curves(os.path.join(base, f"{schema.name}_result"), schema.name)
    out["ours"].append({"seed": seed, "wall_s": round(time.time() - t0, 1), "avg_jsd": [x[0] for x in c],
                        "avg_wd": [x[1] for x in c], "round_s": server.round_times,
                        "weights": [float(w) for w in server.weights_con_cat_combination],
                        "utility": utility_gap(os.path.join(base, f"{schema.name}_result"), schema.name, "ours", seed)})
    print("ours seed", seed, "final JSD/WD", c[-1], out["ours"][-1]["utility"] and out["ours"][-1]["utility"]["delta_f1"], flush=True)

for seed in range(seeds if "reference" in arms else 0):
    # ---------------- reference (its own run(); torch/numpy seeds are whatever it draws)
    base = tempfile.mkdtemp(prefix="parity_ref_")
    ref_harness._prepare_dirs(base, schema, df, shards)
    common = dict(world_size=n_clients + 1, port=24500 + seed, columns=schema.columns, categorical=schema.categorical,
                  nonnegative=schema.nonnegative, target=schema.target, epochs=epochs)
    t0 = time.time()
    procs = [ref_harness._spawn(dict(common, rank=0, cwd=os.path.join(base, "server")), None, os.path.join(base, "server.log"))]
    for i in range(n_clients):
        procs.append(ref_harness._spawn(dict(common, rank=i + 1, cwd=os.path.join(base, f"client{i + 1}")),
                                        0 if dev != "cpu" else None, os.path.join(base, f"client{i + 1}.log")))
    rc = [p.wait() for p in procs]
    if any(rc):
        print("reference failed", rc, open(os.path.join(base, "server.log")).read()[-2000:])
        continue
    c = curves(os.path.join(base, "server", "Intrusion_result"), "Intrusion")
    times = [float(x) for x in open(os.path.join(base, "server", "timestamp_experiment.csv")).read().split()]
    out["reference"].append({"seed": seed, "wall_s": round(time.time() - t0, 1), "avg_jsd": [x[0] for x in c],
                             "avg_wd": [x[1] for x in c], "round_s": times,
                             "utility": utility_gap(os.path.join(base, "server", "Intrusion_result"), "Intrusion", "reference", seed)})
    print("reference seed", seed, "final JSD/WD", c[-1], out["reference"][-1]["utility"] and out["reference"][-1]["utility"]["delta_f1"], flush=True)


def summary(arm):
    if not out[arm]:
        return None
    j = np.array([r["avg_jsd"] for r in out[arm]])
    w = np.array([r["avg_wd"] for r in out[arm]])
    return {"final_avg_jsd_mean": float(j[:, -1].mean()), "final_avg_jsd_std": float(j[:, -1].std()),
            "final_avg_wd_mean": float(w[:, -1].mean()), "final_avg_wd_std": float(w[:, -1].std()),
            "last5_avg_jsd_mean": float(j[:, -5:].mean(1).mean()), "last5_avg_jsd_std": float(j[:, -5:].mean(1).std()),
            "last5_avg_wd_mean": float(w[:, -5:].mean(1).mean()), "last5_avg_wd_std": float(w[:, -5:].mean(1).std()),
            "epoch0_avg_jsd": float(j[:, 0].mean()), "epoch0_avg_wd": float(w[:, 0].mean()),
            "delta_f1_mean": (float(np.mean([r["utility"]["delta_f1"] for r in out[arm]])) if out[arm][0].get("utility") else None),
            "delta_f1_std": (float(np.std([r["utility"]["delta_f1"] for r in out[arm]])) if out[arm][0].get("utility") else None),
            "mean_round_s": float(np.mean([np.mean(r["round_s"][1:]) for r in out[arm]])) if epochs > 1 else None}


out["summary"] = {"ours": summary("ours"), "reference": summary("reference")}
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"parity_{tag}.json"), "w"), indent=1)
print(json.dumps(out["summary"], indent=1))
