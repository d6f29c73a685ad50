"""How the federator reaches its clients.

The reference has exactly one topology: a star of blocking/async RPCs from rank 0
(``Server/dtds/distributed.py:595-597,661-664,696-701,763-765,799-806,821-823,832-835``;
SURVEY R1-R10).  The federator logic here is written once against ``ClientGroup`` and runs
over three transports:

``LocalGroup``   clients are objects in this process (tests, 1-GPU runs, emulated ranks).
``RpcGroup``     ``torch.distributed.rpc`` RRefs -- the CLI-compatible star.  Control plane
                 only: on GPUs the per-round model exchange is the in-kernel NVLink
                 all-reduce, triggered by one async RPC per client.
``SpmdGroup``    one process per GPU under ``torchrun``; every rank runs the *same*
                 federator code on its own client and results are all-gathered, so the
                 federator state is replicated and rank 0 only adds file output.  No
                 per-round host communication at all (the barrier is inside the kernel).
"""

from __future__ import annotations

import os
import warnings

import torch
import torch.distributed as dist

from ..ops import reference as ref_ops


class ClientGroup:
    n: int
    is_root: bool = True
    # True when generator_arena() is the live training arena on this process' GPU (the
    # federator may then snapshot it and generate asynchronously); False when it is a copy
    # fetched over the control plane.
    local_generator: bool = True

    def call(self, method, *args, per_client=None):
        raise NotImplementedError

    def call_one(self, idx, method, *args):
        raise NotImplementedError

    def run(self, method, *args):
        """Invoke on every client, results not needed (no gather)."""
        self.call(method, *args)

    def root_compute(self, fn):
        return fn()

    def aggregate(self, weights, round_id):
        raise NotImplementedError

    def generator_arena(self):
        """Arena (on this process' device) holding the current global generator."""
        raise NotImplementedError

    def prepare(self):
        """Called once after the clients built their networks and before the first round."""

    def is_writer(self, round_no: int) -> bool:
        """Does this process generate + write the synthetic CSV of round ``round_no``?"""
        return self.is_root

    def barrier(self):
        pass

    def shutdown(self):
        pass

    # True when ``aggregate`` can take the weights as a CUDA tensor (read by the kernel in place)
    accepts_device_weights = False

    def device_statistics(self, models, cont_ids, count_vectors, rows, device, seed=0, K=10):
        """Initialisation statistics on the device.  Default: this process holds every
        client's mixtures, so it draws + sorts all samples on its own GPU (``devstats.local_statistics``)."""
        if not self.is_root:
            return None
        from .devstats import local_statistics

        return local_statistics(models, cont_ids, count_vectors, rows, device, seed=seed, K=K)


def _average_state_dicts(dicts, weights):
    """Reference ``average_model`` (``distributed.py:86-106``): every key, fixed client order,
    fp32; 1-client short circuit."""
    if len(dicts) == 1:
        return dicts[0]
    out = {}
    for k in dicts[0]:
        acc = dicts[0][k].to(torch.float32) * float(weights[0])
        for d, w in zip(dicts[1:], weights[1:]):
            acc = acc + d[k].to(torch.float32) * float(w)
        out[k] = acc.to(dicts[0][k].dtype) if dicts[0][k].dtype != torch.float32 else acc
    return out


class LocalGroup(ClientGroup):
    accepts_device_weights = True

    def __init__(self, clients):
        self.clients = list(clients)
        self.n = len(self.clients)

    def call(self, method, *args, per_client=None):
        out = []
        for i, c in enumerate(self.clients):
            a = per_client[i] if per_client is not None else args
            out.append(getattr(c, method)(*a))
        return out

    def call_one(self, idx, method, *args):
        return getattr(self.clients[idx], method)(*args)

    def aggregate(self, weights, round_id):
        if self.n == 1:
            return
        from .. import ops

        engines = [c.engine for c in self.clients]
        dev = engines[0].device
        if dev.type == "cuda" and ops.native_available():
            from ..ops import native

            for which in ("G", "D"):
                arenas = [getattr(e, which) for e in engines]
                native.fedavg_local([a.data for a in arenas], weights, arenas[0].int_slot_offsets())
        else:
            if torch.is_tensor(weights):
                weights = weights.detach().cpu().tolist()
            for which in ("G", "D"):
                arenas = [getattr(e, which) for e in engines]
                ref_ops.fedavg([a.data for a in arenas], weights, arenas[0].int_slot_offsets())

    def generator_arena(self):
        return self.clients[0].engine.G


class RpcGroup(ClientGroup):
    local_generator = False

    def __init__(self, rrefs, use_symmetric=False, server_device="cpu"):
        self.rrefs = list(rrefs)
        self.n = len(self.rrefs)
        self.use_symmetric = use_symmetric
        self.server_device = server_device
        self._arena = None

    def call(self, method, *args, per_client=None):
        futs = []
        for i, r in enumerate(self.rrefs):
            a = per_client[i] if per_client is not None else args
            futs.append(getattr(r.rpc_async(), method)(*a))
        return [f.wait() for f in futs]

    def call_one(self, idx, method, *args):
        return getattr(self.rrefs[idx].rpc_sync(), method)(*args)

    def prepare(self):
        if self.use_symmetric and self.n > 1:
            self.setup_symmetric()

    def setup_symmetric(self):
        handles = self.call("symm_export")
        self.call("symm_connect", per_client=[(i, handles) for i in range(self.n)])
        self.use_symmetric = True
        self._map_generator(handles[0])

    def _map_generator(self, handle):
        """Map client 1's arena block into the federator process (CUDA IPC): after every aggregation it holds the
        averaged generator, so the per-epoch generation reads it in place instead of pulling a pickled state_dict
        over RPC, re-uploading it and generating synchronously (the reference's data flow, R8/R9)."""
        if not str(self.server_device).startswith("cuda"):
            return
        try:
            from ..features.layout import TableLayout
            from ..models.arena import GeneratorArena
            from ..ops import native
            from .symm import _CudaBlob

            dev = torch.device(self.server_device)
            info = self.call_one(0, "get_layout_info")
            probe = GeneratorArena(self._cfg, TableLayout(info["output_info"]), "cpu", init=False)
            with torch.cuda.device(dev):
                ptr = native.symm_open(handle)
                view = torch.as_tensor(_CudaBlob(ptr, probe.numel), device=dev)
            self._arena = GeneratorArena(self._cfg, TableLayout(info["output_info"]), dev, storage=view, init=False)
            self._mapped_ptr = ptr
            self.local_generator = True          # federator may snapshot it and generate asynchronously
            self.snapshot_sync = True            # ... but the clients are other processes: finish the snapshot
                                                 # before the next round is started over RPC
        except Exception as exc:                 # same-GPU IPC mapping unavailable: keep the state_dict pull
            warnings.warn(f"could not map the generator arena into the federator ({exc}); using the RPC pull")
            self._arena, self.local_generator = None, False

    def aggregate(self, weights, round_id):
        if self.n == 1:
            return
        if self.use_symmetric:
            self.call("symm_aggregate", list(map(float, weights)), int(round_id), True)
            return
        # CPU / no-peer-access path == the reference's data flow (gather, average, push)
        pairs = self.call("get_state_dicts")
        g = _average_state_dicts([p[0] for p in pairs], weights)
        d = _average_state_dicts([p[1] for p in pairs], weights)
        self.call("set_discriminator_weights", d)
        self.call("set_generator_weights", g)
        self._last_g = g

    def generator_arena(self):
        from ..features.layout import TableLayout
        from ..models.arena import GeneratorArena

        if self.local_generator and self._arena is not None:
            return self._arena                   # live view of client 1's (averaged) generator
        sd = getattr(self, "_last_g", None)
        if sd is None or self.use_symmetric:
            sd = self.call_one(0, "get_generator_weights")
        if self._arena is None:
            info = self.call_one(0, "get_layout_info")
            cfg = self._cfg
            self._arena = GeneratorArena(cfg, TableLayout(info["output_info"]), self.server_device, init=False)
        self._arena.load_state_dict(sd)
        self._last_g = None
        return self._arena


class SpmdGroup(ClientGroup):
    """Every rank owns one client; collective calls all-gather their results."""

    accepts_device_weights = True

    def device_statistics(self, models, cont_ids, count_vectors, rows, device, seed=0, K=10):
        """Collective: each rank publishes only its own histogram + sorted GMM sample in a CUDA-IPC block and
        every rank's fused kernels read the peers' statistics in place over NVLink."""
        from .devstats import spmd_statistics

        return spmd_statistics(self.rank, models[self.rank], cont_ids, count_vectors, rows, device, pg=self.pg,
                               seed=seed, K=K)

    def __init__(self, client, pg=None, use_symmetric=None, rotate_writers=False):
        self.rotate_writers = rotate_writers
        self.client = client
        self.pg = pg
        self.rank = dist.get_rank(pg)
        self.n = dist.get_world_size(pg)
        self.is_root = self.rank == 0
        self.use_symmetric = use_symmetric
        self._symm_ready = False

    def call(self, method, *args, per_client=None):
        a = per_client[self.rank] if per_client is not None else args
        mine = getattr(self.client, method)(*a)
        out = [None] * self.n
        dist.all_gather_object(out, mine, group=self.pg)
        return out

    def call_one(self, idx, method, *args):
        box = [getattr(self.client, method)(*args) if self.rank == idx else None]
        dist.broadcast_object_list(box, src=idx, group=self.pg)
        return box[0]

    def run(self, method, *args):
        getattr(self.client, method)(*args)

    def is_writer(self, round_no: int) -> bool:
        # After aggregation every rank holds the same generator and the federator state is
        # replicated, so the per-epoch generation + CSV duty CAN rotate over the ranks (all ranks
        # share the box's file system).  Measured at 8 GPUs it is no faster than rank 0 doing it
        # asynchronously (244 vs 255 epochs/s), so it is off by default.
        if self.rotate_writers:
            return round_no % self.n == self.rank
        return self.is_root

    def root_compute(self, fn):
        box = [fn() if self.is_root else None]
        dist.broadcast_object_list(box, src=0, group=self.pg)
        return box[0]

    def barrier(self):
        if self._symm_ready:
            self.client.symm.check()
        dist.barrier(group=self.pg)

    def host_barrier(self):
        dist.barrier(group=self.pg)

    def gather_block(self, n_floats):
        """Collective: a second CUDA-IPC block per rank (used as rank 0's gather buffer for the distributed
        generation of the epoch CSV); returns the connected block (``peer_ptrs[r]`` = rank r's copy)."""
        from .symm import SymmetricBlock

        blk = SymmetricBlock(n_floats, self.client.device)
        handles = [None] * self.n
        dist.all_gather_object(handles, blk.export_handle(), group=self.pg)
        blk.connect(self.rank, handles)
        dist.barrier(group=self.pg)
        self._gather_blocks = getattr(self, "_gather_blocks", []) + [blk]
        return blk

    def prepare(self):
        e = self.client.engine
        symmetric = self.use_symmetric if self.use_symmetric is not None else e.device.type == "cuda"
        if symmetric and self.n > 1 and not self._symm_ready:
            self._setup_symmetric()

    def _setup_symmetric(self):
        # FEDTGAN_AGGREGATION=nvls: in-switch reduction (multimem) instead of the P2P two-shot
        # kernel; measured only marginally faster at 8 GPUs / 168 MB and slower below that
        # (profiles/fedavg_mgpu_8_nvls.json), and not run-to-run bitwise reproducible -> opt-in.
        if os.environ.get("FEDTGAN_AGGREGATION", "twoshot") == "nvls":
            try:
                self.client.symm_setup_nvls(self.pg)
                self._symm_ready = True
                return
            except (RuntimeError, ImportError) as exc:
                warnings.warn(f"NVLS aggregation unavailable ({exc}); using the P2P two-shot kernel")
        handle = self.client.symm_export()
        handles = [None] * self.n
        dist.all_gather_object(handles, handle, group=self.pg)
        self.client.symm_connect(self.rank, handles)
        dist.barrier(group=self.pg)
        self._symm_ready = True

    def aggregate(self, weights, round_id):
        if self.n == 1:
            return
        e = self.client.engine
        symmetric = self.use_symmetric if self.use_symmetric is not None else e.device.type == "cuda"
        if symmetric:
            if not self._symm_ready:
                self._setup_symmetric()
            if getattr(self, "pre_aggregate", None) is not None:
                self.pre_aggregate()
            self.client.symm_aggregate(weights if torch.is_tensor(weights) else list(map(float, weights)), int(round_id))
            return
        # gloo / CPU: all-gather the flat arenas, reduce in fixed client order (bit-identical
        # in every rank)
        if torch.is_tensor(weights):
            weights = weights.detach().cpu().tolist()
        for arena in (e.G, e.D):
            parts = [torch.empty_like(arena.data) for _ in range(self.n)]
            dist.all_gather(parts, arena.data, group=self.pg)
            ref_ops.fedavg(parts, weights, arena.int_slot_offsets())
            arena.data.copy_(parts[0])

    def generator_arena(self):
        return self.client.engine.G
