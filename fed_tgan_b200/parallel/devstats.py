"""Device path of the federated initialisation statistics (SURVEY K14 sampler + K15).

Reference: ``MDGANServer.uniform_meta_category`` (JSD part), ``uniform_continuous_gmm`` (GMM sampling + WD
part) and ``calculate_final_weights_for_aggregation`` (``Server/dtds/distributed.py:592-783``): the server
samples every client's pickled sklearn mixtures on the CPU, sorts inside scipy's ``wasserstein_distance``
and computes the weights with numpy.

Here each client's sample is drawn AND sorted by one kernel (``ftg_gmm_sample_sort``), and a fused pair of
kernels (``ftg_fed_stats_to_weights``) reads every client's histogram and sorted sample through device
pointers and leaves the aggregation weights in device memory, where ``ftg_fedavg_*`` read them.

Two ways to obtain the pointers:

``local_statistics``  the calling process holds every client's mixture parameters (LocalGroup, RpcGroup, or
                      a single client) and builds all samples on its own GPU;
``spmd_statistics``   one process per GPU: every rank builds only ITS OWN histogram + sorted sample inside a
                      CUDA-IPC block, the ranks exchange the 64-byte handles, and every rank's kernels read
                      the peers' statistics in place over NVLink.  All ranks end up with bit-identical
                      weights without any host gather of samples.
"""

from __future__ import annotations

import numpy as np
import torch


class DeviceStats:
    """Result of the fused kernel: device tensors + lazily fetched host copies."""

    def __init__(self, jsd, wd, weights, centers, pool, keep=()):
        self.jsd, self.wd, self.weights, self.centers, self.pool = jsd, wd, weights, centers, pool
        self._keep = keep            # buffers the pointers referred to (must outlive the kernels)

    def host(self):
        return (self.jsd.double().cpu().numpy(), self.wd.double().cpu().numpy(), self.weights.double().cpu().numpy())


def _gmm_block(models, cont_ids, device):
    """[C, 7, K] fp64 device array of one client's continuous-column mixtures."""
    arr = np.stack([models[ci].to_array() for ci in cont_ids]) if cont_ids else np.zeros((0, 7, 1))
    return torch.as_tensor(arr, dtype=torch.float64).to(device).contiguous()


def _hist_vector(count_vectors, client):
    """Client's category counts, columns concatenated in global label order, and the column offsets."""
    if not count_vectors:
        return np.zeros(4, np.float32), np.zeros(1, np.int32)
    hist = np.concatenate([count_vectors[c][client] for c in range(len(count_vectors))]).astype(np.float32)
    off = np.concatenate([[0], np.cumsum([len(count_vectors[c][0]) for c in range(len(count_vectors))])]).astype(np.int32)
    return hist, off


def sample_counts(rows):
    """``int(sum_rows * frac_i)`` per client, exactly as the reference computes it (``distributed.py:732-734``)."""
    n_total = int(np.sum(rows))
    return [int(n_total * (r / n_total)) for r in rows]


def local_statistics(models_by_client, cont_ids, count_vectors, rows, device, seed=0, K=10, want_pool=True):
    from ..ops import native

    dev = torch.device(device)
    world = len(rows)
    n_i = sample_counts(rows)
    keep, hist_ptrs, seg_ptrs = [], [], []
    cat_off = None
    with torch.cuda.device(dev):
        for i in range(world):
            h, off = _hist_vector(count_vectors, i)
            ht = torch.as_tensor(h).to(dev)
            keep.append(ht)
            hist_ptrs.append(ht.data_ptr())
            cat_off = torch.as_tensor(off).to(dev)
            if cont_ids:
                seg = native.gmm_sample_sort(_gmm_block(models_by_client[i], cont_ids, dev), n_i[i], seed, i)
                keep.append(seg)
                seg_ptrs.append(seg.data_ptr())
        keep.append(cat_off)
        jsd, wd, w, centers, pool = native.fed_stats_to_weights(
            hist_ptrs, seg_ptrs, n_i, rows, cat_off if count_vectors else None, len(cont_ids), dev, K=K,
            want_pool=want_pool)
    return DeviceStats(jsd, wd, w, centers, pool, keep)


def spmd_statistics(rank, my_models, cont_ids, count_vectors, rows, device, pg=None, seed=0, K=10, want_pool=True):
    """Collective over ``pg``: every rank publishes its own statistics block and reads the peers' in place."""
    import torch.distributed as dist

    from ..ops import native
    from .symm import SymmetricBlock

    dev = torch.device(device)
    world = len(rows)
    n_i = sample_counts(rows)
    h, off = _hist_vector(count_vectors, rank)
    hist_pad = (len(h) + 3) // 4 * 4
    pad = native.sort_pad(n_i[rank]) if cont_ids else 0
    with torch.cuda.device(dev):
        block = SymmetricBlock(hist_pad + len(cont_ids) * pad, dev)
        block.tensor[: len(h)].copy_(torch.as_tensor(h).to(dev))
        if cont_ids:
            seg = block.tensor[hist_pad: hist_pad + len(cont_ids) * pad].view(len(cont_ids), pad)
            native.gmm_sample_sort(_gmm_block(my_models, cont_ids, dev), n_i[rank], seed, rank, out=seg)
        torch.cuda.synchronize()
        handles = [None] * world
        dist.all_gather_object(handles, (block.export_handle(), hist_pad), group=pg)
        block.connect(rank, [hd for hd, _ in handles])
        dist.barrier(group=pg)                       # every rank's block is complete and mapped
        hist_ptrs = list(block.peer_ptrs)
        seg_ptrs = [p + hp * 4 for p, (_, hp) in zip(block.peer_ptrs, handles)]
        cat_off = torch.as_tensor(off).to(dev)
        jsd, wd, w, centers, pool = native.fed_stats_to_weights(
            hist_ptrs, seg_ptrs, n_i, rows, cat_off if count_vectors else None, len(cont_ids), dev, K=K,
            want_pool=want_pool)
        torch.cuda.synchronize()
        dist.barrier(group=pg)                       # nobody unmaps while a peer's kernels still read
        block.close()
    return DeviceStats(jsd, wd, w, centers, pool, (block, cat_off))
