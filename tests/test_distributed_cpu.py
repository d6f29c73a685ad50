"""Multi-process CPU runs: the RPC star (CLI entry point, world_size=3) and the SPMD/gloo path."""

import os
import socket
import subprocess
import sys

import numpy as np
import pandas as pd
import pytest

from fed_tgan_b200.data.schemas import get_schema, make_table

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _env():
    env = dict(os.environ)
    env["PYTHONPATH"] = ROOT + os.pathsep + env.get("PYTHONPATH", "")
    env["CUDA_VISIBLE_DEVICES"] = ""
    env["OMP_NUM_THREADS"] = "2"
    return env


def test_rpc_star_intrusion_world3(tmp_path):
    """BASELINE config 1 in miniature: Intrusion schema, server + 2 clients over torch RPC."""
    schema = get_schema("intrusion")
    os.makedirs(tmp_path / "data" / "raw")
    make_table(schema, seed=0, rows=700).to_csv(tmp_path / "data" / "raw" / "Intrusion_train.csv", index=False)
    port = _free_port()
    cmd = [sys.executable, "-m", "dtds.distributed", "-ip", "127.0.0.1", "-port", str(port), "-world_size", "3",
           "-epoch", "2", "-datapath", "data/raw/Intrusion_train.csv", "--cpu", "--vgm_backend", "vb",
           "--batch_size", "100", "--sample_rows", "200", "--gen_dim", "32,32", "--dis_dim", "32,32"]
    procs = [subprocess.Popen(cmd + ["-rank", str(r)], cwd=tmp_path, env=_env(), stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT) for r in range(3)]
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=600)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append(out.decode(errors="replace"))
    assert all(p.returncode == 0 for p in procs), "\n".join(o[-3000:] for o in outs)
    for i in range(2):
        df = pd.read_csv(tmp_path / "Intrusion_result" / f"Intrusion_synthesis_epoch_{i}.csv")
        assert list(df.columns) == schema.columns and len(df) == 200
    ts = pd.read_csv(tmp_path / "timestamp_experiment.csv", header=None)
    assert len(ts) == 2
    assert "final aggregation weights" in outs[0]


def test_spmd_gloo_two_ranks(tmp_path):
    schema = get_schema("adult", rows=900)
    make_table(schema, seed=3).to_csv(tmp_path / "Adult_train.csv", index=False)
    port = _free_port()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port), "-m", "dtds.distributed", "--spmd", "--cpu",
           "-epochs", "2", "-datapath", str(tmp_path / "Adult_train.csv"), "--out_dir", str(tmp_path),
           "-selected_variables", ",".join(schema.columns), "-categorical_list", ",".join(schema.categorical),
           "-nonnegative_list", "", "-target_column", schema.target, "--vgm_backend", "vb",
           "--batch_size", "100", "--sample_rows", "150", "--gen_dim", "32,32", "--dis_dim", "32,32"]
    res = subprocess.run(cmd, cwd=tmp_path, env=_env(), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=600)
    assert res.returncode == 0, res.stdout.decode(errors="replace")[-4000:]
    for i in range(2):
        df = pd.read_csv(tmp_path / "Adult_result" / f"Adult_synthesis_epoch_{i}.csv")
        assert list(df.columns) == schema.columns and len(df) == 150


def test_spmd_gloo_checkpoint_resume(tmp_path):
    """SPMD mode: a run stopped after epoch 0 resumes from the checkpoint (states all-gathered to rank 0 for
    saving, broadcast back on load) and only trains the missing epoch."""
    schema = get_schema("adult", rows=700)
    make_table(schema, seed=4).to_csv(tmp_path / "Adult_train.csv", index=False)

    def launch(epochs, extra):
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
               "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), "-m", "dtds.distributed", "--spmd",
               "--cpu", "-epochs", str(epochs), "-datapath", str(tmp_path / "Adult_train.csv"), "--out_dir", str(tmp_path),
               "-selected_variables", ",".join(schema.columns), "-categorical_list", ",".join(schema.categorical),
               "-nonnegative_list", "", "-target_column", schema.target, "--vgm_backend", "vb", "--batch_size", "100",
               "--sample_rows", "120", "--gen_dim", "16,16", "--dis_dim", "16,16", "--checkpoint_every", "1"] + extra
        res = subprocess.run(cmd, cwd=tmp_path, env=_env(), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, timeout=600)
        assert res.returncode == 0, res.stdout.decode(errors="replace")[-4000:]

    launch(1, [])
    ck = tmp_path / "models" / "checkpoint_Adult.pt"
    assert ck.exists()
    first = (tmp_path / "Adult_result" / "Adult_synthesis_epoch_0.csv").stat().st_mtime_ns
    launch(2, ["--resume"])
    assert (tmp_path / "Adult_result" / "Adult_synthesis_epoch_1.csv").exists()
    assert (tmp_path / "Adult_result" / "Adult_synthesis_epoch_0.csv").stat().st_mtime_ns == first   # epoch 0 not redone
    ts = pd.read_csv(tmp_path / "timestamp_experiment.csv", header=None)
    assert len(ts) == 2
