"""Process bootstrap and CLI.

Reference: ``run()`` and ``__main__`` of ``Server/dtds/distributed.py:838-971`` (client
trees: ``Client/.../dtds/distributed.py:395-491``).  The command line is kept:

    python -m dtds.distributed -ip <server ip> -port 7788 -rank 0 -world_size 3 -epoch 500 \\
           -datapath data/raw/Intrusion_train.csv            # federator
    python -m dtds.distributed -ip <server ip> -rank 1 -world_size 3        # client 1 ...

(both ``-epoch`` and ``-epochs`` are accepted; defaults are the KDD-Intrusion schema).
rank 0 = federator driving clients ``client1..clientN`` through ``torch.distributed.rpc``
(TensorPipe; ``PROCESS_GROUP`` no longer exists in torch); rank k>0 maps to GPU k-1.
torch RPC is the control plane only: on GPUs the per-round model exchange is the in-kernel
NVLink all-reduce.

Extras: ``-rank`` omitted spawns the whole world locally (the reference's ``mp.spawn``
branch is broken, ``:956-971``); ``--spmd`` runs one process per GPU under ``torchrun``.
List/dict flags can really be overridden from the shell (comma separated / JSON).
"""

from __future__ import annotations

import argparse
import json
import os

import torch
import torch.distributed as dist
import torch.distributed.rpc as rpc
import torch.multiprocessing as mp

from ..config import CTGANConfig
from ..data.schemas import INTRUSION_CATEGORICAL, INTRUSION_COLUMNS, INTRUSION_NONNEG
from .client import MDGANClient
from .federator import MDGANServer, dataset_name
from .groups import RpcGroup, SpmdGroup


def _csv_list(text):
    if isinstance(text, (list, tuple)):
        return list(text)
    text = text.strip()
    if text.startswith("["):
        return json.loads(text)
    return [t.strip() for t in text.split(",") if t.strip()]


def _json_dict(text):
    return text if isinstance(text, dict) else json.loads(text or "{}")


def build_parser():
    p = argparse.ArgumentParser(prog="dtds.distributed", description="Fed-TGAN on B200")
    p.add_argument("-rank", type=int, default=None)
    p.add_argument("-ip", type=str, default="127.0.0.1")
    p.add_argument("-port", type=int, default=7788)
    p.add_argument("-name", type=str, default="Intrusion_train")
    p.add_argument("-datapath", type=str, default="data/raw/Intrusion_train.csv")
    p.add_argument("-epochs", "-epoch", dest="epochs", type=int, default=10)
    p.add_argument("-E_interval", type=int, default=1)
    p.add_argument("-world_size", type=int, default=2)
    p.add_argument("-report", action="store_true")
    p.add_argument("-problem_type", type=str, default="")
    p.add_argument("-target_column", type=str, default="class")
    p.add_argument("-selected_variables", type=_csv_list, default=list(INTRUSION_COLUMNS))
    p.add_argument("-categorical_list", type=_csv_list, default=list(INTRUSION_CATEGORICAL))
    p.add_argument("-nonnegative_list", type=_csv_list, default=list(INTRUSION_NONNEG))
    p.add_argument("-date_dic", type=_json_dict, default={})
    # extensions
    p.add_argument("--spmd", action="store_true", help="one process per GPU under torchrun")
    p.add_argument("--out_dir", type=str, default=".")
    p.add_argument("--vgm_backend", choices=["sklearn", "vb"], default="sklearn")
    p.add_argument("--sample_rows", type=int, default=40000)
    p.add_argument("--batch_size", type=int, default=500)
    p.add_argument("--gen_dim", type=_csv_list, default=[256, 256])
    p.add_argument("--dis_dim", type=_csv_list, default=[256, 256])
    p.add_argument("--precision", choices=["fp32x3", "tf32", "fp32"], default="fp32x3",
                   help="GEMM arithmetic: fp32x3 = 3xTF32 split on the tensor cores (fp32-equivalent, matches the "
                        "reference's fp32 matmuls; default), tf32 = single-pass tensor cores (opt-in), fp32 = CUDA cores")
    p.add_argument("--split", choices=["none", "iid", "skewed"], default="none",
                   help="all clients read the same -datapath and each keeps its own horizontal shard of it "
                        "(iid: equal random shards; skewed: unequal, sorted by the first column)")
    p.add_argument("--split_seed", type=int, default=0)
    p.add_argument("--seed", type=int, default=0)
    p.add_argument("--rpc_timeout", type=float, default=600.0)
    p.add_argument("--checkpoint_every", type=int, default=0)
    p.add_argument("--aggregation", choices=["twoshot", "nvls"], default=None,
                   help="SPMD aggregation kernel: P2P two-shot (default) or NVSwitch multimem reduce")
    p.add_argument("--resume", action="store_true")
    p.add_argument("--cpu", action="store_true", help="force CPU clients (gloo/RPC test config)")
    return p


def config_from_args(args) -> CTGANConfig:
    return CTGANConfig(epochs=args.epochs, batch_size=args.batch_size, gen_dim=tuple(map(int, args.gen_dim)),
                       dis_dim=tuple(map(int, args.dis_dim)), sample_rows=args.sample_rows,
                       precision=args.precision, seed=args.seed)


def resolve_datapath(template: str, client_id: int) -> str:
    """``-datapath`` may carry ``{rank}`` (1-based RPC rank of the client, as in the reference's
    ``Client<k>`` folders) and/or ``{client}`` (0-based client id) so that every client of a one-box
    launch reads its OWN file; the reference gets the same effect from one working directory per client
    (``README.md:12``)."""
    return template.replace("{rank}", str(client_id + 1)).replace("{client}", str(client_id))


def client_kwargs(args, cfg, client_id, device, n_clients=None):
    split = None
    if getattr(args, "split", "none") != "none" and n_clients and n_clients > 1:
        split = (args.split, int(n_clients), int(client_id), int(getattr(args, "split_seed", 0)))
    return dict(datapath=resolve_datapath(args.datapath, client_id), split=split,
                selected_variables=args.selected_variables,
                categorical_list=args.categorical_list, nonnegative_list=args.nonnegative_list,
                date_dic=args.date_dic, target_column=args.target_column, problem_type=args.problem_type,
                epochs=args.epochs, config=cfg.to_dict(), device=device, vgm_backend=args.vgm_backend,
                client_id=client_id)


def federate(server: MDGANServer, verbose=True):
    """The 7-step federator script (reference ``distributed.py:865-875``)."""
    say = print if verbose and server.group.is_root else (lambda *a, **k: None)
    say("Begin initialization for categorical columns")
    server.uniform_meta_category()
    say("Begin initialization for continuous columns")
    server.uniform_continuous_gmm()
    server.refit_local_transformer()
    say("Calculate aggregation weights for participants")
    w = server.calculate_final_weights_for_aggregation()
    say("final aggregation weights", w)
    server.server_local_synthesizer_initialization()
    say("Finishing initialization")
    return server.fit()


def _warn_shared_datapath(args, n_clients):
    if n_clients > 1 and getattr(args, "split", "none") == "none" and "{" not in args.datapath:
        import warnings

        warnings.warn(f"all {n_clients} clients resolve -datapath to the same file '{args.datapath}': unless every "
                      "client process runs in its own working directory they train on identical data and the "
                      "similarity weights degenerate to 1/N.  Use '{rank}'/'{client}' in -datapath or --split iid|skewed.")


def _client_device(rank, force_cpu):
    if force_cpu or not torch.cuda.is_available():
        return "cpu"
    return f"cuda:{(rank - 1) % torch.cuda.device_count()}"


def _apply_precision(cfg):
    """The CLI's --precision reaches the kernels through the process-wide switch of ops.native."""
    if torch.cuda.is_available():
        from ..ops import native

        native.set_precision(cfg["precision"] if isinstance(cfg, dict) else cfg.precision)


def _make_remote_client(kwargs):
    if str(kwargs.get("device", "cpu")).startswith("cuda"):
        torch.cuda.set_device(torch.device(kwargs["device"]))
        _apply_precision(kwargs["config"])
    return MDGANClient(**kwargs)


def run(rank, einterval, world_size, ip, port, name, datapath, selected_variables, categorical_list,
        nonnegative_list, date_dic, target_column, problem_type, epochs, report, args=None):
    """RPC star bootstrap (same signature as the reference's ``run``)."""
    if args is None:
        args = build_parser().parse_args([])
        args.datapath, args.selected_variables, args.categorical_list = datapath, selected_variables, categorical_list
        args.nonnegative_list, args.date_dic, args.target_column = nonnegative_list, date_dic, target_column
        args.problem_type, args.epochs = problem_type, epochs
    os.environ["MASTER_ADDR"] = ip
    os.environ["MASTER_PORT"] = str(port)
    os.environ["WORLD_SIZE"] = str(world_size)
    os.environ["RANK"] = str(rank)
    opts = rpc.TensorPipeRpcBackendOptions(num_worker_threads=8, rpc_timeout=args.rpc_timeout,
                                           init_method=f"tcp://{ip}:{port}")
    cfg = config_from_args(args)
    if rank == 0:
        if torch.cuda.is_available() and not args.cpu:
            _apply_precision(cfg)
        print("number of training epochs: ", epochs)
        print("world size: ", world_size, f"tcp://{ip}:{port}")
        rpc.init_rpc("server", rank=0, world_size=world_size, rpc_backend_options=opts)
        rrefs = []
        for worker in range(1, world_size):
            kw = client_kwargs(args, cfg, worker - 1, _client_device(worker, args.cpu), world_size - 1)
            rrefs.append(rpc.remote(f"client{worker}", _make_remote_client, args=(kw,)))
            print(f"register remote client{worker}")
        on_gpu = torch.cuda.is_available() and not args.cpu
        _warn_shared_datapath(args, world_size - 1)
        # the in-kernel barrier spins: it needs every client on its own GPU (two spinning kernels
        # time-slicing one GPU can only time out), otherwise fall back to the state-dict exchange
        one_gpu_each = on_gpu and (world_size - 1) <= torch.cuda.device_count()
        group = RpcGroup(rrefs, use_symmetric=one_gpu_each and world_size > 2,
                         server_device="cuda:0" if on_gpu else "cpu")
        server = MDGANServer(group, epochs=epochs, config=cfg, name=dataset_name(datapath), out_dir=args.out_dir,
                             vgm_backend=args.vgm_backend, checkpoint_every=args.checkpoint_every,
                             resume=args.resume, device="cuda:0" if on_gpu else "cpu")
        federate(server)
    else:
        rpc.init_rpc(f"client{rank}", rank=rank, world_size=world_size, rpc_backend_options=opts)
        print(f"client{rank} is joining")
    rpc.shutdown()


def run_spmd(args):
    """One process per GPU under ``torchrun``: rank r is client r; rank 0 also writes files."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    on_gpu = torch.cuda.is_available() and not args.cpu
    if on_gpu:
        torch.cuda.set_device(local)
    if not dist.is_initialized():
        dist.init_process_group("nccl" if on_gpu else "gloo", rank=rank, world_size=world)
    cfg = config_from_args(args)
    if on_gpu:
        _apply_precision(cfg)
    device = f"cuda:{local}" if on_gpu else "cpu"
    if getattr(args, "aggregation", None):
        os.environ["FEDTGAN_AGGREGATION"] = args.aggregation
    if rank == 0:
        _warn_shared_datapath(args, world)
    client = MDGANClient(**client_kwargs(args, cfg, rank, device, world))
    server = MDGANServer(SpmdGroup(client), epochs=args.epochs, config=cfg, name=dataset_name(args.datapath),
                         out_dir=args.out_dir, vgm_backend=args.vgm_backend, device=device,
                         checkpoint_every=args.checkpoint_every, resume=args.resume)
    out = federate(server)
    dist.barrier()
    return out


def _spawn_entry(rank, args):
    run(rank, args.E_interval, args.world_size, args.ip, args.port, args.name, args.datapath,
        args.selected_variables, args.categorical_list, args.nonnegative_list, args.date_dic,
        args.target_column, args.problem_type, args.epochs, args.report, args=args)


def main(argv=None):
    args = build_parser().parse_args(argv)
    if args.spmd:
        return run_spmd(args)
    if args.rank is not None:
        return _spawn_entry(args.rank, args)
    # whole world on this machine (server + world_size-1 clients)
    mp.spawn(_spawn_entry, args=(args,), nprocs=args.world_size, join=True)


if __name__ == "__main__":
    main()
