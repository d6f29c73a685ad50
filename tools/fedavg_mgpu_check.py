"""Multi-GPU check of the fused aggregation kernel (run under torchrun, one rank per GPU).

Verifies two-shot and one-shot against `sum_i w_i theta_i` computed with torch on gathered
replicas, bitwise replica equality after write-back, and times the kernel (CUDA events, max
over ranks) across arena sizes, next to an NCCL all-reduce of pre-scaled arenas (the
"NCCL equivalent" baseline).
"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fed_tgan_b200.parallel.symm import NvlsBlock, SymmetricBlock  # noqa: E402

MODES = os.environ.get("FEDAVG_MODES", "twoshot,oneshot").split(",")

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)

sizes = [1_005_472, 2_900_000, 8_000_000, 42_000_000]      # floats: Adult G+D, Intrusion, ..., wide 1024 nets
if os.environ.get("FEDAVG_SIZES"):
    sizes = [int(x) for x in os.environ["FEDAVG_SIZES"].split(",")]
n_max = max(sizes)
ipc_blk = SymmetricBlock(n_max, dev)
handles = [None] * world
dist.all_gather_object(handles, ipc_blk.export_handle())
ipc_blk.connect(rank, handles)
dist.barrier()
nvls_blk = None
if "nvls" in MODES:
    try:
        nvls_blk = NvlsBlock(n_max, dev)
    except Exception as exc:                                   # noqa: BLE001
        if rank == 0:
            print(json.dumps({"mode": "nvls", "unavailable": repr(exc)[:300]}), flush=True)
        MODES = [m for m in MODES if m != "nvls"]
w = torch.softmax(torch.arange(world, dtype=torch.float64) * 0.3, 0).tolist()
if os.environ.get("FEDAVG_UNIFORM") == "1":        # IID shards: equal weights (NVLS then skips its pre-scale pass)
    w = [1.0 / world] * world
round_id = 0
results = []
ok_all = True
for mode in MODES:
    blk = nvls_blk if mode == "nvls" else ipc_blk
    for n in sizes:
        n4 = n // 4 * 4
        g = torch.Generator(device="cpu").manual_seed(100 + rank)
        mine = torch.randn(n4, generator=g).to(dev)
        mine[7] = 40.0 + rank          # integer slot
        blk.tensor[:n4].copy_(mine)
        gathered = [torch.empty(n4, device=dev) for _ in range(world)]
        dist.all_gather(gathered, mine)
        ref = gathered[0] * w[0]
        for i in range(1, world):
            ref += gathered[i] * w[i]
        ref[7] = torch.trunc(ref[7])
        torch.cuda.synchronize(); dist.barrier()
        round_id += 1
        blk.fedavg(w, n4, [7], round_id, mode=mode)
        torch.cuda.synchronize()
        blk.check()
        got = blk.tensor[:n4]
        err = (got - ref).abs().max().item()
        same = [torch.empty(n4, device=dev) for _ in range(world)]
        dist.all_gather(same, got.clone())
        equal = all(torch.equal(same[0], s) for s in same)
        ok = err < 1e-5 and equal
        ok_all &= ok
        # timing
        times = []
        for it in range(12):
            torch.cuda.synchronize(); dist.barrier()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            round_id += 1
            s.record(); blk.fedavg(w, n4, [7], round_id, mode=mode); e.record()
            torch.cuda.synchronize()
            times.append(s.elapsed_time(e))
        t = torch.tensor(sorted(times)[len(times) // 2], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        # NCCL equivalent: pre-scale + all_reduce
        buf = mine.clone()
        nt = []
        for it in range(12):
            torch.cuda.synchronize(); dist.barrier()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); buf.mul_(w[rank]); dist.all_reduce(buf); e.record()
            torch.cuda.synchronize()
            nt.append(s.elapsed_time(e))
        tn = torch.tensor(sorted(nt)[len(nt) // 2], device=dev, dtype=torch.float64)
        dist.all_reduce(tn, op=dist.ReduceOp.MAX)
        # bytes crossing THIS GPU's links per direction: two-shot = my loads of shard r from N-1 peers + the peers'
        # stores of their shards into me (and the mirror image outbound): 2 (N-1)/N P; one-shot: (N-1) P inbound;
        # NVLS: the reduced shard + the other ranks' multicasts inbound, my contributions + my multicast outbound: P
        P_bytes = n4 * 4
        per_dir = {"oneshot": (world - 1) * P_bytes, "nvls": P_bytes}.get(mode, 2 * (world - 1) / world * P_bytes)
        gbps = per_dir / (float(t) * 1e-3) / 1e9
        row = {"mode": mode, "world": world, "floats": n4, "uniform_weights": os.environ.get("FEDAVG_UNIFORM") == "1",
               "max_err": err, "replicas_equal": equal, "ok": ok, "ms": float(t), "nccl_ms": float(tn),
               "link_bytes_per_direction": per_dir, "nvlink_GBps_per_direction": gbps,
               "fraction_of_770GBps_peer_copy": gbps / 770.0}
        results.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
if rank == 0:
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    tagu = "_uniform" if os.environ.get("FEDAVG_UNIFORM") == "1" else ""
    with open(os.path.join(ROOT, "gpurun_out", f"fedavg_mgpu_{world}{'_nvls' if 'nvls' in MODES else ''}{tagu}.json"), "w") as fh:
        json.dump(results, fh, indent=1)
    print("ALL OK" if ok_all else "FAILED")
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok_all else 1)
