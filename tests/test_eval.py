import sys

import numpy as np
import pandas as pd

from fed_tgan_b200.eval import stat_sim_normalize, utility_difference
from fed_tgan_b200.data.schemas import get_schema, make_table


def test_similarity_matches_reference_script(tmp_path):
    schema = get_schema("adult", rows=600)
    real = make_table(schema, seed=0)
    fake = make_table(schema, seed=1, skew=0.4)
    real.to_csv(tmp_path / "r.csv", index=False)
    fake.to_csv(tmp_path / "f.csv", index=False)
    sys.path.insert(0, "/root/reference/Server")
    try:
        import similarity_analysis as ref  # the reference's own script
        ref_out = ref.stat_sim_normalize(str(tmp_path / "r.csv"), str(tmp_path / "f.csv"), list(schema.categorical))
    finally:
        sys.path.remove("/root/reference/Server")
        sys.modules.pop("similarity_analysis", None)
    mine = stat_sim_normalize(str(tmp_path / "r.csv"), str(tmp_path / "f.csv"), schema.categorical)
    assert np.allclose(mine, ref_out, atol=1e-9)
    same = stat_sim_normalize(real, real, schema.categorical)
    assert same[0] < 1e-9 and same[1] < 1e-12


def test_utility_runs():
    schema = get_schema("adult", rows=500)
    df = make_table(schema, seed=0)
    tr, te = df.iloc[:350], df.iloc[350:]
    diff, f1 = utility_difference(tr, te, tr.copy(), schema.categorical, schema.target, verbose=False)
    assert abs(f1) < 1e-9 and diff.shape == (4, 2)
