"""Peer-visible ("symmetric") device memory for the fused aggregation kernel.

Each rank allocates one block ``[arena floats | flag page]`` with ``cudaMalloc`` through the
native library, exports its CUDA-IPC handle, and maps every peer's block.  The control
plane (torch RPC or ``torch.distributed``) only ships the 64-byte handles once; after that
the per-round model exchange is a single kernel per rank that loads/stores peer memory over
NVLink and synchronises with flags in these same blocks.
"""

from __future__ import annotations

import torch

import os

FLAG_INTS = 256
# how long a rank spins at the in-kernel barrier before it gives up and sets the status word.  Shards of very different
# size (non-IID splits) legitimately arrive far apart; FEDTGAN_BARRIER_TIMEOUT_MS overrides.
BARRIER_TIMEOUT_MS = int(os.environ.get("FEDTGAN_BARRIER_TIMEOUT_MS", "60000"))


class _CudaBlob:
    """Minimal ``__cuda_array_interface__`` carrier so torch can wrap a raw device pointer."""

    def __init__(self, ptr, n_floats):
        self.__cuda_array_interface__ = {
            "shape": (n_floats,), "typestr": "<f4", "data": (int(ptr), False), "version": 3, "strides": None,
        }


class _StatusWord:
    """Host side of the kernels' status word (0 = ok, 1/2 = barrier phase that timed out)."""

    status: torch.Tensor
    device: torch.device
    _host_status = None
    _status_evt = None

    def check(self):
        """Blocking: synchronise and raise if any launch so far gave up at a barrier."""
        code = int(self.status.item())
        if code:
            raise RuntimeError(f"fedavg peer barrier timed out (phase {code})")

    def poll(self):
        """Non-blocking: raise if a status copy enqueued by an earlier ``poll`` has landed with an
        error, then enqueue the next copy.  Called once per round, so a rank that missed a barrier
        surfaces one round later without ever stalling the launch thread."""
        if self._host_status is None:
            self._host_status = torch.zeros(1, dtype=torch.int32).pin_memory()
        if self._status_evt is not None:
            if not self._status_evt.query():
                return
            self._status_evt = None
            code = int(self._host_status[0])
            if code:
                raise RuntimeError(f"fedavg peer barrier timed out (phase {code})")
        with torch.cuda.device(self.device):
            self._host_status.copy_(self.status, non_blocking=True)
            self._status_evt = torch.cuda.Event()
            self._status_evt.record()


class SymmetricBlock(_StatusWord):
    def __init__(self, n_floats: int, device):
        from ..ops import native

        self.native = native
        self.device = torch.device(device)
        self.n_floats = (int(n_floats) + 31) // 32 * 32
        self.nbytes = self.n_floats * 4 + FLAG_INTS * 4
        with torch.cuda.device(self.device):
            self.ptr = native.symm_alloc(self.nbytes)
            self._blob = _CudaBlob(self.ptr, self.n_floats)
            self.tensor = torch.as_tensor(self._blob, device=self.device)
        self.flag_ptr = self.ptr + self.n_floats * 4
        self.status = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.staging = None
        self.peer_ptrs = None
        self.rank = None
        self._opened = []

    def export_handle(self) -> bytes:
        return self.native.symm_export(self.ptr)

    def connect(self, rank: int, handles):
        self.rank = int(rank)
        self.world = len(handles)
        ptrs = []
        with torch.cuda.device(self.device):
            for r, h in enumerate(handles):
                if r == self.rank:
                    ptrs.append(self.ptr)
                else:
                    p = self.native.symm_open(h)
                    self._opened.append(p)
                    ptrs.append(p)
        self.peer_ptrs = ptrs
        self.peer_flag_ptrs = [p + self.n_floats * 4 for p in ptrs]

    def fedavg(self, weights, n, int_offsets, round_id, mode="twoshot", timeout_ms=None):
        timeout_ms = BARRIER_TIMEOUT_MS if timeout_ms is None else timeout_ms
        n = (int(n) + 3) // 4 * 4
        assert n <= self.n_floats
        with torch.cuda.device(self.device):
            if mode == "twoshot":
                self.native.fedavg_twoshot(self.peer_ptrs, self.peer_flag_ptrs, self.rank, weights, n, int_offsets,
                                           round_id, self.status, timeout_ms)
            else:
                if self.staging is None:
                    self.staging = torch.empty(self.n_floats, dtype=torch.float32, device=self.device)
                self.native.fedavg_oneshot(self.peer_ptrs, self.peer_flag_ptrs, self.staging, self.rank, weights, n,
                                           int_offsets, round_id, self.status, timeout_ms)

    def close(self):
        for p in self._opened:
            try:
                self.native.symm_close(p)
            except Exception:
                pass
        self._opened = []


class NvlsBlock(_StatusWord):
    """Multicast-bound symmetric block (``torch.distributed._symmetric_memory`` does the
    cuMulticast* / fabric-handle plumbing); the aggregation itself is our ``ftg_fedavg_nvls``
    kernel.  Collective constructor: every rank of ``group`` must call it.  Raises
    ``RuntimeError`` if the platform exposes no multicast pointer (NVLS unavailable)."""

    def __init__(self, n_floats: int, device, group=None):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem

        from ..ops import native

        self.native = native
        self.device = torch.device(device)
        self.n_floats = (int(n_floats) + 31) // 32 * 32
        group = group if group is not None else dist.group.WORLD
        with torch.cuda.device(self.device):
            self._all = symm_mem.empty(self.n_floats + FLAG_INTS, dtype=torch.float32, device=self.device)
            self._all.zero_()
            self.hdl = symm_mem.rendezvous(self._all, group)
        self.rank, self.world = self.hdl.rank, self.hdl.world_size
        mc = int(getattr(self.hdl, "multicast_ptr", 0) or 0)
        if mc == 0:
            raise RuntimeError("NVLS multicast is not available on this platform")
        self.multicast_ptr = mc
        self.tensor = self._all[: self.n_floats]
        self.peer_ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        self.peer_flag_ptrs = [p + self.n_floats * 4 for p in self.peer_ptrs]
        self.status = torch.zeros(1, dtype=torch.int32, device=self.device)
        dist.barrier(group=group)

    def fedavg(self, weights, n, int_offsets, round_id, mode="nvls", timeout_ms=None):
        timeout_ms = BARRIER_TIMEOUT_MS if timeout_ms is None else timeout_ms
        n = (int(n) + 3) // 4 * 4
        with torch.cuda.device(self.device):
            if mode == "nvls":
                self.native.fedavg_nvls(self.peer_ptrs, self.peer_flag_ptrs, self.multicast_ptr, self.rank, weights, n,
                                        int_offsets, round_id, self.status, timeout_ms)
            else:
                self.native.fedavg_twoshot(self.peer_ptrs, self.peer_flag_ptrs, self.rank, weights, n, int_offsets,
                                           round_id, self.status, timeout_ms)
