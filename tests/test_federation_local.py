"""End-to-end on CPU, in one process: federated init -> rounds -> CSVs -> analysis."""

import os

import numpy as np
import pandas as pd

from fed_tgan_b200.config import CTGANConfig
from fed_tgan_b200.data.schemas import get_schema, make_table, split_table
from fed_tgan_b200.parallel.client import MDGANClient
from fed_tgan_b200.parallel.federator import MDGANServer
from fed_tgan_b200.parallel.launcher import federate


def _clients(schema, shards, cfg, **kw):
    return [MDGANClient(datapath=f"{schema.name}_train.csv", selected_variables=schema.columns,
                        categorical_list=schema.categorical, nonnegative_list=schema.nonnegative, date_dic={},
                        target_column=schema.target, problem_type="", epochs=2, device="cpu", config=cfg,
                        dataframe=sh, write_side_files=False, client_id=i, **kw)
            for i, sh in enumerate(shards)]


def test_two_clients_adult_small(tmp_path):
    schema = get_schema("adult", rows=1500)
    df = make_table(schema, seed=1)
    shards = split_table(df, 2, "skewed")
    cfg = CTGANConfig(batch_size=100, gen_dim=(32, 32), dis_dim=(32, 32), embedding_dim=16, sample_rows=300)
    clients = _clients(schema, shards, cfg)
    server = MDGANServer(clients, epochs=2, config=cfg, name="Adult", out_dir=str(tmp_path), async_csv=True)
    times = federate(server, verbose=False)
    assert len(times) == 2
    w = server.weights_con_cat_combination
    assert abs(w.sum() - 1) < 1e-9 and (w > 0).all()
    # replicas identical after aggregation
    a, b = clients[0].engine, clients[1].engine
    assert np.array_equal(a.G.data.numpy(), b.G.data.numpy())
    assert np.array_equal(a.D.data.numpy(), b.D.data.numpy())
    for i in range(2):
        out = pd.read_csv(tmp_path / "Adult_result" / f"Adult_synthesis_epoch_{i}.csv")
        assert list(out.columns) == schema.columns and len(out) == 300
        for c, k in zip(schema.categorical, schema.cat_sizes):
            assert set(out[c]).issubset(set(df[c]))
    assert os.path.exists(tmp_path / "timestamp_experiment.csv")
    assert os.path.exists(tmp_path / "models" / "Adult.json")
    assert os.path.exists(tmp_path / "models" / "label_encoders_Adult.pickle")
    # same output_info on every client
    assert clients[0].out_info == clients[1].out_info


def test_single_client_shortcut(tmp_path):
    schema = get_schema("adult", rows=400)
    df = make_table(schema, seed=2)
    cfg = CTGANConfig(batch_size=50, gen_dim=(16,), dis_dim=(16,), embedding_dim=8, sample_rows=60, pack=5)
    clients = _clients(schema, [df], cfg)
    server = MDGANServer(clients, epochs=1, config=cfg, name="Adult", out_dir=str(tmp_path), async_csv=False)
    federate(server, verbose=False)
    assert np.allclose(server.weights_con_cat_combination, [1.0])
    assert os.path.exists(tmp_path / "Adult_result" / "Adult_synthesis_epoch_0.csv")


def test_rebinding_arena_storage_keeps_training_state(tmp_path):
    """Moving the arenas into another block (what happens when the symmetric allocation is set up)
    must not reset weights, Adam moments, the RNG counter or the iteration count: replaying an
    epoch from the same state gives bit-identical results with and without a rebind in between."""
    import torch

    schema = get_schema("adult", rows=600)
    df = make_table(schema, seed=3)
    cfg = CTGANConfig(batch_size=50, gen_dim=(16, 16), dis_dim=(16,), embedding_dim=8, sample_rows=60, pack=5)
    c = _clients(schema, [df], cfg)[0]
    server = MDGANServer([c], epochs=1, config=cfg, name="Adult", out_dir=str(tmp_path), async_csv=False)
    server.uniform_meta_category(); server.uniform_continuous_gmm(); server.refit_local_transformer()
    c.run_local_epochs(1)
    start = c.checkpoint_state()
    c.run_local_epochs(1)
    want = c.checkpoint_state()
    c.load_checkpoint_state(start)
    block = torch.full((c.engine.G.numel + c.engine.D.numel + 64,), float("nan"))
    c._rebind_storage(block)
    assert c.engine.G.data.data_ptr() == block.data_ptr()
    c.run_local_epochs(1)
    got = c.checkpoint_state()
    for net in ("G", "D"):
        for k in want[net]:
            assert torch.equal(want[net][k], got[net][k]), (net, k)
        assert torch.equal(want["opt" + net]["exp_avg"], got["opt" + net]["exp_avg"])
        assert want["opt" + net]["step"] == got["opt" + net]["step"]
    assert torch.equal(want["rng"], got["rng"]) and want["iterations"] == got["iterations"]


def test_csv_fast_path_matches_general_decode(tmp_path):
    """The native strided writer (numeric table -> text in one call) must produce byte-identical files to the
    general path (per-column decode + column writer), including out-of-range / negative category codes."""
    import torch

    from fed_tgan_b200.data.transform import Transform
    from fed_tgan_b200.utils.csvio import write_table_csv

    schema = get_schema("adult", rows=500)
    df = make_table(schema, seed=4)
    cfg = CTGANConfig(batch_size=50, gen_dim=(16,), dis_dim=(16,), embedding_dim=8, sample_rows=700, pack=5)
    c = _clients(schema, [df], cfg)[0]
    server = MDGANServer([c], epochs=1, config=cfg, name="Adult", out_dir=str(tmp_path), async_csv=False)
    server.uniform_meta_category(); server.uniform_continuous_gmm(); server.refit_local_transformer()
    server.calculate_final_weights_for_aggregation(); server.server_local_synthesizer_initialization()
    table = server.sample(cfg.sample_rows).to(torch.float32).cpu().numpy()
    cat_cols = [j for j, n in enumerate(schema.columns) if n in schema.categorical]
    table[3, cat_cols[0]] = 1e6           # beyond the last class -> clamped
    table[5, cat_cols[1]] = -4.0          # negative -> class 0
    plan = server._csv_plan()
    assert plan is not None
    fast = plan.write(str(tmp_path / "fast.csv"), table, n_threads=3)
    names, cols = Transform.decode_arrays(table, server.meta_base, server.label_encoder, as_codes=True)
    slow = write_table_csv(str(tmp_path / "slow.csv"), names, cols, n_threads=2)
    assert open(fast, "rb").read() == open(slow, "rb").read()
    back = pd.read_csv(fast)
    assert list(back.columns) == schema.columns and len(back) == cfg.sample_rows
    # values no integer can hold (numpy's astype is undefined there): the native writer clamps in float space --
    # huge / +inf -> last class, NaN / -inf -> class 0 -- and never indexes outside the class table
    wild = table.copy()
    j = cat_cols[2]
    wild[6, j], wild[7, j], wild[8, j], wild[9, j] = 3.0e38, float("inf"), float("nan"), float("-inf")
    got = pd.read_csv(plan.write(str(tmp_path / "wild.csv"), wild, n_threads=2), keep_default_na=False)
    classes = [str(c) for c in server.label_encoder[2]["label_encoder"].classes_]
    assert list(got.iloc[6:10, j]) == [classes[-1], classes[-1], classes[0], classes[0]]
    # header-less part files concatenate to the same bytes
    a = plan.write(str(tmp_path / "p0.csv"), table[:300], n_threads=2, header=True)
    b = plan.write(str(tmp_path / "p1.csv"), table[300:], n_threads=3, header=False)
    assert open(a, "rb").read() + open(b, "rb").read() == open(fast, "rb").read()
