"""Does the multi-GPU (SPMD, in-kernel aggregation) path actually train?  Run under torchrun.

Adult-shape synthetic table, IID split over the ranks, E federated epochs; rank 0 then scores every
epoch's CSV against the real table (Avg_JSD / Avg_WD) and checks that all replicas are bit-identical.
The same table through the single-process path (profiles/parity_adult.json, 2 clients) reaches
Avg_JSD ~0.03 / Avg_WD ~0.04 around epochs 3-5; a broken discriminator stays at the epoch-0 level.
Writes gpurun_out/spmd_quality_<world>.json.
"""
import json
import os
import sys
import tempfile

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fed_tgan_b200.config import CTGANConfig  # noqa: E402
from fed_tgan_b200.data.schemas import get_schema, make_table, split_table  # noqa: E402
from fed_tgan_b200.eval import stat_sim_normalize  # noqa: E402
from fed_tgan_b200.parallel.client import MDGANClient  # noqa: E402
from fed_tgan_b200.parallel.federator import MDGANServer  # noqa: E402
from fed_tgan_b200.parallel.groups import SpmdGroup  # noqa: E402
from fed_tgan_b200.parallel.launcher import federate  # noqa: E402

epochs = int(sys.argv[1]) if len(sys.argv) > 1 else 6
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
local = int(os.environ.get("LOCAL_RANK", rank))
on_gpu = torch.cuda.is_available()
if on_gpu:
    torch.cuda.set_device(local)
dist.init_process_group("nccl" if on_gpu else "gloo", rank=rank, world_size=world)
dev = f"cuda:{local}" if on_gpu else "cpu"
schema = get_schema("adult")
df = make_table(schema, seed=0)
shard = split_table(df, world, "iid", seed=0)[rank]
cfg = CTGANConfig(seed=0)
box = [tempfile.mkdtemp(prefix="spmdq_") if rank == 0 else None]
dist.broadcast_object_list(box, src=0)
client = MDGANClient(datapath=f"{schema.name}_train.csv", selected_variables=schema.columns,
                     categorical_list=schema.categorical, nonnegative_list=schema.nonnegative, date_dic={},
                     target_column=schema.target, problem_type="", epochs=epochs, device=dev, config=cfg,
                     dataframe=shard, write_side_files=False, client_id=rank, vgm_backend="sklearn")
server = MDGANServer(SpmdGroup(client), epochs=epochs, config=cfg, name=schema.name, out_dir=box[0], device=dev)
times = federate(server, verbose=False)
e = client.engine
sig = torch.stack([e.G.data.double().sum(), e.G.data.double().abs().sum(), e.D.data.double().sum(),
                   e.D.data.double().abs().sum()]).to(dev)
sigs = [torch.empty_like(sig) for _ in range(world)]
dist.all_gather(sigs, sig)
same = all(torch.equal(sigs[0], s) for s in sigs)
losses = client.get_losses()
if rank == 0:
    res = {"world": world, "epochs": epochs, "device": dev, "replicas_identical": bool(same), "losses_rank0": losses,
           "aggregation": os.environ.get("FEDTGAN_AGGREGATION", "twoshot") if on_gpu and world > 1 else "host",
           "round_s": times, "avg_jsd": [], "avg_wd": []}
    for i in range(epochs):
        j, w = stat_sim_normalize(df, os.path.join(box[0], f"{schema.name}_result", f"{schema.name}_synthesis_epoch_{i}.csv"),
                                  schema.categorical)
        res["avg_jsd"].append(float(j)); res["avg_wd"].append(float(w))
    # distributed generation: every rank contributed a contiguous share of the rows -- all shares must be there and
    # look alike (score each share of the last CSV on its own)
    import pandas as pd

    last = pd.read_csv(os.path.join(box[0], f"{schema.name}_result", f"{schema.name}_synthesis_epoch_{epochs - 1}.csv"))
    res["csv_rows"] = int(len(last))
    per = (len(last) + world - 1) // world
    res["share_avg_jsd"] = []
    for r in range(world):
        part = last.iloc[r * per:(r + 1) * per]
        pth = os.path.join(box[0], f"share_{r}.csv")
        part.to_csv(pth, index=False)
        res["share_avg_jsd"].append(float(stat_sim_normalize(df, pth, schema.categorical)[0]))
    res["distributed_generation"] = getattr(server, "_dg", None) is not None
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", f"spmd_quality_{world}.json"), "w") as fh:
        json.dump(res, fh, indent=1)
    print(json.dumps(res))
    ok = same and min(res["avg_jsd"]) < 0.06 and all(abs(v) < 1e4 for v in losses.values())
    ok = ok and res["csv_rows"] == cfg.sample_rows and max(res["share_avg_jsd"]) < 2.0 * min(res["share_avg_jsd"]) + 0.02
    print("QUALITY OK" if ok else "QUALITY FAILED")
dist.barrier()
dist.destroy_process_group()
