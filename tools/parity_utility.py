"""Score saved last-epoch CSVs of tools/parity_run.py (PARITY_SAVE_LAST=<dir>, PARITY_SKIP_UTILITY=1) for ML utility
on the CPU: difference in weighted F1 (real-trained minus synthetic-trained, mean of LR / DT / RF / MLP, tested on a
held-out real table) -- the reference's utility_analysis.py:15-91 metric.  Fills `utility` into the parity JSON.

    python tools/parity_utility.py gpurun_out/parity_<tag>.json <csv_dir> [correlated=1.0]
"""
import glob
import json
import os
import re
import sys

import numpy as np
import pandas as pd

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fed_tgan_b200.data.schemas import get_schema, make_table  # noqa: E402
from fed_tgan_b200.eval.utility import utility_difference  # noqa: E402

path, csv_dir = sys.argv[1], sys.argv[2]
correlated = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
out = json.load(open(path))
schema = get_schema(out["config"])
df = make_table(schema, seed=0, correlated=correlated)
df_test = make_table(schema, seed=1, rows=10000, correlated=correlated)
tag = os.path.basename(path)[len("parity_"):-len(".json")]
for arm in ("ours", "reference"):
    for rec in out.get(arm, []):
        hits = glob.glob(os.path.join(csv_dir, f"{tag}_{arm}_seed{rec['seed']}_epoch*.csv"))
        if not hits or rec.get("utility"):
            continue
        fake = pd.read_csv(hits[0])
        fake.columns = list(df.columns)
        diff, f1 = utility_difference(df, df_test, fake, schema.categorical, schema.target, verbose=False)
        rec["utility"] = {"delta_f1": f1, "delta_acc_f1_per_model": np.asarray(diff).tolist()}
        print(arm, rec["seed"], "delta_f1", f1, flush=True)
    vals = [r["utility"]["delta_f1"] for r in out.get(arm, []) if r.get("utility")]
    if vals and out.get("summary", {}).get(arm):
        out["summary"][arm]["delta_f1_mean"] = float(np.mean(vals))
        out["summary"][arm]["delta_f1_std"] = float(np.std(vals))
json.dump(out, open(path, "w"), indent=1)
print(json.dumps(out.get("summary"), indent=1))
