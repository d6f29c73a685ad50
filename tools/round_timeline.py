"""Where does a federated ROUND spend its time at N GPUs?  (Run under torchrun; CUDA events, no profiler.)

Per rank and per round, events on the main stream bracket: L2 flush | local epoch (graph replays) |
aggregation kernel | generation hand-off (snapshot + side-stream enqueue).  Host wall time of each phase is
taken next to it, so a phase whose device time is small but whose host time is large is launch/GIL bound.
Also reports the ideal round (steps x graph iteration time measured back to back).  Writes
gpurun_out/round_timeline_<world>.json (one row per rank).

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/round_timeline.py [rounds]
"""
import json
import os
import sys
import tempfile
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fed_tgan_b200.config import CTGANConfig  # noqa: E402
from fed_tgan_b200.data.schemas import get_schema, make_table, split_table  # noqa: E402
from fed_tgan_b200.parallel.client import MDGANClient  # noqa: E402
from fed_tgan_b200.parallel.federator import MDGANServer  # noqa: E402
from fed_tgan_b200.parallel.groups import LocalGroup, SpmdGroup  # noqa: E402

rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 20
rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", str(rank)))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)

schema = get_schema("adult")
df = make_table(schema, seed=0)
shard = split_table(df, max(world, 1), "iid", seed=0)[rank]
cfg = CTGANConfig()
box = [tempfile.mkdtemp(prefix="fedtgan_tl_") if rank == 0 else None]
if world > 1:
    dist.broadcast_object_list(box, src=0)
client = MDGANClient(datapath=f"{schema.name}_train.csv", selected_variables=schema.columns,
                     categorical_list=schema.categorical, nonnegative_list=schema.nonnegative, date_dic={},
                     target_column=schema.target, problem_type="", epochs=rounds, device=dev, config=cfg,
                     dataframe=shard, write_side_files=False, client_id=rank, vgm_backend="vb")
group = SpmdGroup(client) if world > 1 else LocalGroup([client])
server = MDGANServer(group, epochs=rounds, config=cfg, name=schema.name, out_dir=box[0], vgm_backend="vb", device=dev,
                     async_csv=True)
server.uniform_meta_category()
server.uniform_continuous_gmm()
server.refit_local_transformer()
weights = server.calculate_final_weights_for_aggregation()
server.server_local_synthesizer_initialization()
group.prepare()
server.warm_generation()
flush = torch.empty(192 << 20, dtype=torch.uint8, device=dev)

PHASES = ["flush", "train", "aggregate", "generate"]


def one_round(i, record):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(len(PHASES) + 1)]
    host = [time.perf_counter()]
    ev[0].record()
    flush.zero_()
    ev[1].record(); host.append(time.perf_counter())
    client.run_local_epochs(1)
    ev[2].record(); host.append(time.perf_counter())
    group.aggregate(weights, server._round_id())
    ev[3].record(); host.append(time.perf_counter())
    server.sample_data(i, wait=False)
    ev[4].record(); host.append(time.perf_counter())
    if record is not None:
        record.append((ev, host))


for i in range(4):
    one_round(f"w{i}", None)
server.drain()
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
rec = []
t0 = time.perf_counter()
for i in range(rounds):
    one_round(f"t{i}", rec)
torch.cuda.synchronize()
wall = time.perf_counter() - t0
server.drain()
dev_ms = np.array([[e[k].elapsed_time(e[k + 1]) for k in range(len(PHASES))] for e, _ in rec])
host_ms = np.array([[(h[k + 1] - h[k]) * 1e3 for k in range(len(PHASES))] for _, h in rec])

# ideal: the same number of graph replays back to back, nothing else on the GPU
steps = client.steps_per_epoch
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
client.run_local_epochs(3)
e.record()
torch.cuda.synchronize()
ideal_ms = s.elapsed_time(e) / 3

row = {"rank": rank, "world": world, "rounds": rounds, "steps_per_round": steps,
       "wall_ms_per_round": wall * 1e3 / rounds, "ideal_train_ms_per_round": ideal_ms,
       "device_ms": {p: float(np.median(dev_ms[:, k])) for k, p in enumerate(PHASES)},
       "device_ms_sum": float(np.median(dev_ms.sum(1))),
       "host_ms": {p: float(np.median(host_ms[:, k])) for k, p in enumerate(PHASES)},
       "host_ms_sum": float(np.median(host_ms.sum(1)))}
rows = [None] * world
if world > 1:
    dist.all_gather_object(rows, row)
else:
    rows = [row]
if rank == 0:
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", f"round_timeline_{world}.json"), "w") as fh:
        json.dump(rows, fh, indent=1)
    for r in rows:
        print(json.dumps(r))
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
