"""``bench.py --impl reference``: time the unmodified reference on the same table / split.

Topology is the reference's own: one federator process (rank 0 of the RPC world, CPU) plus
one client process per GPU (``cuda:0`` of its own ``CUDA_VISIBLE_DEVICES``), all driven by
the reference's ``run()``.  Under ``torchrun`` every rank starts its own client; rank 0 also
starts the federator.  The per-round time is the reference's own measurement
(``timestamp_experiment.csv``, ``Server/dtds/distributed.py:795-829``: wall clock of the whole
round incl. sampling 40 000 rows and writing the CSV); the first ``warmup`` rounds are
dropped.  Files the reference expects but does not create are prepared here
(``models/``, ``Intrusion_result/``, ``models/Intrusion_train.npz`` built with the
reference's own ``FileGenerator``).
"""
from __future__ import annotations

import json
import os
import subprocess
import sys
import tempfile
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.path.join(HERE, "_ref")
PUBLISHED_EPOCHS_PER_S = 1.0 / 24.2


def _prepare_dirs(base, schema, df, shards):
    import numpy as np

    os.makedirs(os.path.join(base, "server", "models"), exist_ok=True)
    os.makedirs(os.path.join(base, "server", "Intrusion_result"), exist_ok=True)
    for i, sh in enumerate(shards):
        d = os.path.join(base, f"client{i + 1}", "data", "raw")
        os.makedirs(d, exist_ok=True)
        sh.to_csv(os.path.join(d, "Intrusion_train.csv"), index=False)
    # models/Intrusion_train.npz = the full table label-encoded exactly as the reference's
    # FileGenerator would (sorted-string LabelEncoder ids, log(x+1) on non-negative columns)
    code = f"""
import sys, pickle, json
sys.path.insert(0, {REF!r}); sys.path.insert(0, {HERE!r})
import ref_worker; ref_worker.apply_shims()
import numpy as np, pandas as pd
from sklearn import preprocessing
from dtds.data.utils.file_generator import FileGenerator
df = pd.read_csv({os.path.join(base, 'full.csv')!r})
cols, cats, nonneg = json.loads({json.dumps(json.dumps([schema.columns, schema.categorical, schema.nonnegative]))})
fg = FileGenerator(dataframe=df[cols], file_name='Intrusion_train', problem_type='', target_col={schema.target!r},
                   categorical_list=list(cats), non_negative_columns=list(nonneg), date_columns={{}})
les = []
for c in fg.df.columns:
    if c in cats:
        les.append(preprocessing.LabelEncoder().fit(fg.df[c].astype(str)))
fg.generate_data(fg.generate_meta_data(), les, 'prep')
import glob, shutil
shutil.copy(glob.glob('dtds/saved_models/*prep/*.npz')[0], {os.path.join(base, 'server', 'models', 'Intrusion_train.npz')!r})
"""
    df.to_csv(os.path.join(base, "full.csv"), index=False)
    prep_dir = os.path.join(base, "prep")
    os.makedirs(prep_dir, exist_ok=True)
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", PYTHONPATH="")
    subprocess.run([sys.executable, "-c", code], cwd=prep_dir, env=env, check=True,
                   stdout=subprocess.PIPE, stderr=subprocess.STDOUT)


def _spawn(spec, gpu, log):
    env = dict(os.environ)
    env["CUDA_VISIBLE_DEVICES"] = "" if gpu is None else str(gpu)
    env["PYTHONPATH"] = ""
    # Children must not inherit torchrun's environment: TORCHELASTIC_USE_AGENT_STORE makes the
    # reference's own tcp:// rendezvous wait for a store nobody hosts (hang), and OMP_NUM_THREADS=1
    # would throttle the CPU-heavy reference.
    drop = ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT", "GROUP_RANK", "LOCAL_WORLD_SIZE",
            "ROLE_RANK", "ROLE_WORLD_SIZE", "ROLE_NAME", "GROUP_WORLD_SIZE", "OMP_NUM_THREADS")
    for k in list(env):
        if k in drop or k.startswith("TORCHELASTIC") or k.startswith("TORCH_NCCL") or k.startswith("NCCL_ASYNC"):
            env.pop(k, None)
    fh = open(log, "w")
    return subprocess.Popen([sys.executable, os.path.join(HERE, "ref_worker.py"), json.dumps(spec)],
                            env=env, stdout=fh, stderr=subprocess.STDOUT, start_new_session=True)


def _kill(p):
    try:
        os.killpg(p.pid, 9)          # the worker is its own session leader (start_new_session)
    except Exception:
        try:
            p.kill()
        except Exception:
            pass


def _round_csvs(base):
    d = os.path.join(base, "server", "Intrusion_result")
    try:
        fs = [os.path.join(d, f) for f in os.listdir(d) if f.endswith(".csv")]
    except OSError:
        return []
    out = []
    for f in fs:
        try:
            idx = int(os.path.splitext(f)[0].rsplit("_", 1)[1])
            out.append((idx, os.path.getmtime(f)))
        except Exception:
            pass
    return sorted(out)


def run_reference_bench(args):
    """Budgeted run.  The reference's round is ~30 s at N=1 on the Adult shape, so warm-up + steps can
    exceed the driver's per-run limit.  The harness therefore watches the per-epoch CSVs the reference
    writes (``Intrusion_result/Intrusion_synthesis_epoch_<i>.csv``, one per round, written at the same
    point of every round): if the whole run will not fit ``FEDTGAN_REF_BUDGET_S`` (default 720 s) it
    stops the reference at the budget and reports the rounds that did complete (period between
    consecutive CSVs = one full round), saying so in the JSON (``truncated``, actual ``steps``).  A run
    that fits is reported from the reference's own clock (``timestamp_experiment.csv``)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    try:
        sys.path.insert(0, HERE)
        import install_ref

        ok = install_ref.ensure_installed() if rank == 0 else True
    except Exception as exc:
        ok = False
        why = f"installing baseline/_ref failed: {exc}"[:300]
    else:
        why = "baseline/_ref is not installed and /root/reference is not on this machine (run __graft_entry__.build() first)"
    sys.path.insert(0, ROOT)
    from fed_tgan_b200.data.schemas import get_schema, make_table, split_table   # data generation only

    schema = get_schema(args.config, rows=args.rows)
    n_clients = max(world, 1)
    warm = max(args.warmup, 1)
    epochs = warm + args.steps
    base_port = 23000 + (os.getpid() % 2000 if world == 1 else int(os.environ.get("MASTER_PORT", "29500")) % 2000)
    budget = float(os.environ.get("FEDTGAN_REF_BUDGET_S", "720"))
    t_begin = time.time()
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("gloo", rank=rank, world_size=world)
        box = [tempfile.mkdtemp(prefix="fedtgan_ref_") if rank == 0 else None, ok]
        dist.broadcast_object_list(box, src=0)
        base, ok = box
    else:
        base = tempfile.mkdtemp(prefix="fedtgan_ref_")
    if not ok or not os.path.isdir(os.path.join(REF, "dtds")):
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": why}), flush=True)
        if world > 1:
            dist.destroy_process_group()
        return
    mode = "skewed" if args.config == "covertype" else "iid"
    n_rows = None
    if rank == 0:
        df = make_table(schema, seed=0)
        n_rows = len(df)
        shards = split_table(df, n_clients, mode, seed=0)
        shard_rows = [len(s) for s in shards]
        _prepare_dirs(base, schema, df, shards)
    if world > 1:
        dist.barrier()
    common = dict(world_size=n_clients + 1, port=base_port, columns=schema.columns, categorical=schema.categorical,
                  nonnegative=schema.nonnegative, target=schema.target, epochs=epochs)
    procs = []
    t0 = time.time()
    if rank == 0:
        procs.append(_spawn(dict(common, rank=0, cwd=os.path.join(base, "server")), None,
                            os.path.join(base, "server.log")))
    procs.append(_spawn(dict(common, rank=rank + 1, cwd=os.path.join(base, f"client{rank + 1}")), local,
                        os.path.join(base, f"client{rank + 1}.log")))
    stop_file = os.path.join(base, "STOP")
    truncated = False
    while any(p.poll() is None for p in procs):
        time.sleep(0.5)
        if rank == 0 and time.time() - t_begin > budget and not os.path.exists(stop_file):
            open(stop_file, "w").close()
        if os.path.exists(stop_file):
            truncated = True
            for p in procs:
                if p.poll() is None:
                    _kill(p)
    rc = [p.wait() for p in procs]
    wall = time.time() - t0
    if world > 1:
        dist.barrier()
    if rank == 0:
        out = {"impl": "reference", "metric": "federated_epochs_per_sec", "unit": "federated_epochs/s",
               "n_gpus": world, "steps": args.steps, "warmup": warm, "higher_is_better": True,
               "scaling": "strong", "dtype": "fp32", "data": "synthetic", "gpu_launches": None}
        ts_path = os.path.join(base, "server", "timestamp_experiment.csv")
        times, how = None, None
        if not truncated and not any(rc) and os.path.exists(ts_path):
            times = [float(x) for x in open(ts_path).read().split()]
            timed = times[-args.steps:]
            how = "reference's own per-round wall clock (timestamp_experiment.csv), warm-up rounds dropped"
        elif truncated:
            marks = _round_csvs(base)
            periods = [b[1] - a[1] for a, b in zip(marks, marks[1:]) if b[0] == a[0] + 1]
            # period k covers round k+1 (CSV k -> CSV k+1); round 0 is never inside a period
            drop = max(0, min(warm - 1, len(periods) - 1))
            timed = periods[drop:]
            times = periods
            if timed:
                out.update({"truncated": True, "requested_steps": args.steps, "requested_warmup": warm,
                            "steps": len(timed), "warmup": drop + 1})
                how = (f"stopped at the {budget:.0f} s budget after {len(marks)} of {epochs} rounds; per-round time = "
                       "period between consecutive per-epoch CSVs written by the reference")
        if not times or not timed:
            tail = ""
            try:
                tail = open(os.path.join(base, "server.log")).read()[-400:].replace("\n", " | ")
            except Exception:
                pass
            out["unavailable"] = f"reference run failed rc={rc} truncated={truncated}: {tail}"
        else:
            sec = sum(timed)
            k = len(timed)
            out.update({
                "value": k / sec, "ms_per_step": 1e3 * sec / k, "vs_baseline": (k / sec) / PUBLISHED_EPOCHS_PER_S,
                "e2e": {"value": k / sec, "unit": "federated_epochs/s",
                        "note": "the reference's round is end to end by construction (numpy sampling + H2D every step, D2H + CPU sampling + CSV every round)"},
                "timing": how, "total_wall_s": round(wall, 1), "per_round_s": [round(t, 3) for t in times],
                "config": {"model": "CTGAN G(128+n_opt,(256,256)) / PacGAN-10 D((256,256))",
                           "table": f"{schema.name}-shape synthetic {n_rows}x{len(schema.columns)}", "clients": n_clients,
                           "split": mode, "global_batch": 500 * n_clients, "batch_per_client": 500, "seq_len": None,
                           "local_steps_per_round": max(shard_rows) // 500, "rows_per_client": shard_rows[0],
                           "sample_rows_per_round": 40000, "parallelism": f"federated-dp{n_clients}",
                           "aggregation": "torch RPC (TensorPipe) gather + CPU average_model + RPC broadcast",
                           "code": "baseline/_ref/dtds (unmodified Server/dtds), own run() entry point"},
            })
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()
