"""The federator (reference: ``MDGANServer``, ``Server/dtds/distributed.py:543-835``).

Seven-step script, same public method names as the reference so a launcher reads like
``distributed.py:865-875``:

    uniform_meta_category()                     federated categorical encoding + JSD similarity
    uniform_continuous_gmm()                    federated VGM encoding + WD similarity
    refit_local_transformer()                   clients adopt the global encoders, build nets
    calculate_final_weights_for_aggregation()   similarity x row-fraction -> softmax
    server_local_synthesizer_initialization()   sampling distribution for the CSVs
    fit()                                       rounds: local epoch -> weighted average -> CSV
    sample_data(i) / save_model(name)

Deliberate changes: dataset name is derived from the data path instead of the
hard-coded ``'Intrusion'``; output directories are created; the federator's ``Cond`` is
built from summed per-client option counts, not from a server-side copy of the training
table; CSV decoding/writing overlaps the next round; the round loop has checkpoint /
resume and per-phase metrics.
"""

from __future__ import annotations

import copy
import csv
import datetime
import json
import os
import pickle
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch
from sklearn import preprocessing

from ..config import CTGANConfig
from ..data.constants import CATEGORICAL, CONTINUOUS
from ..data.file_generator import NumpyEncoder
from ..data.transform import Transform
from ..engine.generate import RowGenerator
from ..engine.table import cumulative_tables
from ..features.gmm import GMMParams, fit_sklearn_1d, fit_vbgmm_1d
from ..features.layout import TableLayout
from ..utils.csvio import write_table_csv
from ..utils.metrics import MetricsLog
from . import weights as W
from .groups import ClientGroup, LocalGroup


def dataset_name(datapath: str) -> str:
    stem = os.path.basename(datapath).split(".")[0]
    return stem[:-6] if stem.endswith("_train") else stem


class MDGANServer:
    def __init__(self, clients, epochs, config: CTGANConfig | None = None, name="Intrusion", out_dir=".",
                 device=None, vgm_backend="sklearn", async_csv=True, write_csv=True, checkpoint_every=0,
                 resume=False, seed=None, **kwargs):
        self.cfg = config if config is not None else CTGANConfig(**kwargs)
        self.epochs = int(epochs)
        self.embedding_dim, self.gen_dim, self.dis_dim = self.cfg.embedding_dim, self.cfg.gen_dim, self.cfg.dis_dim
        self.l2scale, self.batch_size = self.cfg.l2scale, self.cfg.batch_size
        self.group: ClientGroup = clients if isinstance(clients, ClientGroup) else LocalGroup(clients)
        if hasattr(self.group, "_cfg") is False:
            self.group._cfg = self.cfg
        self.client_rrefs = getattr(self.group, "rrefs", getattr(self.group, "clients", []))
        self.name = name
        self.out_dir = out_dir
        self.models_dir = os.path.join(out_dir, "models")
        self.result_dir = os.path.join(out_dir, f"{name}_result")
        self.device = torch.device(device) if device is not None else None
        self.vgm_backend = vgm_backend
        self.async_csv = async_csv
        self.write_csv = write_csv
        self.checkpoint_every = checkpoint_every
        self.resume = resume
        self.seed = self.cfg.seed if seed is None else seed
        self.np_rng = np.random.default_rng(self.seed + 99)
        self.metrics = MetricsLog(os.path.join(out_dir, "metrics.jsonl") if self.group.is_root else None)
        self._writer = ThreadPoolExecutor(max_workers=6) if async_csv else None
        self._slots = None
        self._pending = []
        self.start_epoch = 0
        self.round_times = []
        self._rid = 0
        self._csv_done = {}          # epoch index -> wall-clock time at which its CSV was on disk
        os.makedirs(self.models_dir, exist_ok=True)
        os.makedirs(self.result_dir, exist_ok=True)

    # ------------------------------------------------------------ categorical
    def uniform_meta_category(self):
        meta_list = self.group.call("get_meta")
        self.meta_list = meta_list
        meta_base = copy.deepcopy(meta_list[0])
        encoders, count_vectors = [], []
        for ci, column in enumerate(meta_base["columns"]):
            if column["type"] != CATEGORICAL:
                continue
            merged = {}
            for meta in meta_list:
                for key, cnt in meta["columns"][ci]["i2s"].items():
                    merged[key] = merged.get(key, 0) + cnt
            order = [k for k, _ in sorted(merged.items(), key=lambda kv: kv[1], reverse=True)]
            column["i2s"] = order
            column["size"] = len(order)
            le = preprocessing.LabelEncoder().fit(order)
            encoders.append(le)
            slot = {k: i for i, k in enumerate(le.classes_)}
            per_client = []
            for meta in meta_list:
                vec = np.zeros(len(order))
                for key, cnt in meta["columns"][ci]["i2s"].items():
                    vec[slot[key]] = cnt
                per_client.append(vec)
            count_vectors.append(per_client)
        self.category_count_vectors = count_vectors
        self.distribution_similarity_vector = W.categorical_similarity(count_vectors) if count_vectors else \
            np.zeros((self.group.n, 0))
        self.meta_base = meta_base
        self.label_encoders_data = encoders
        stamps = [str(datetime.datetime.now().timestamp()).replace(".", "") + f"{i}" for i in range(self.group.n)]
        stamps = self.group.root_compute(lambda: stamps)
        self.group.call("register_meta", per_client=[(meta_base, stamps[i], encoders) for i in range(self.group.n)])
        cursor = 0
        self.label_encoder = []
        for column in meta_base["columns"]:
            if column["type"] == CATEGORICAL:
                self.label_encoder.append({"column_name": column["column_name"], "label_encoder": encoders[cursor]})
                cursor += 1
        if self.group.is_root:
            with open(os.path.join(self.models_dir, f"label_encoders_{self.name}.pickle"), "wb") as fh:
                pickle.dump(self.label_encoder, fh, protocol=pickle.HIGHEST_PROTOCOL)
            self.meta_path = os.path.join(self.models_dir, f"{self.name}.json")
            with open(self.meta_path, "w") as fh:
                json.dump(meta_base, fh, sort_keys=True, indent=4, separators=(",", ": "), cls=NumpyEncoder)

    # ------------------------------------------------------------- continuous
    def uniform_continuous_gmm(self):
        info = self.group.call("get_transformer_information")
        models = [[None if m is None else GMMParams.from_array(m) for m in item[0]] for item in info]
        self.meta_transformer = info[0][2]
        rows = [int(item[3]) for item in info]
        self.rows_number_list = rows
        n_total = int(np.sum(rows))
        self.model_weights_by_number = [r / n_total for r in rows]
        cfg = self.cfg

        cont_ids_all = [ci for ci, col in enumerate(self.meta_transformer) if col["type"] == CONTINUOUS]
        dev_stats = None
        if self._device_stats_ok():
            # K14/K15 on the device: every client's sample is drawn + sorted by a kernel and the fused
            # statistics kernels read all clients' histograms / samples through device (peer) pointers
            dev_stats = self.group.device_statistics(models, cont_ids_all, self.category_count_vectors, rows,
                                                     self.device, seed=self.seed + 7, K=cfg.n_clusters)
        self.device_stats = dev_stats

        def build():
            if dev_stats is not None:
                jsd_d, wd_d, w_d = dev_stats.host()
                if not cont_ids_all:
                    return wd_d, cont_ids_all, [], (jsd_d, wd_d, w_d)
                if self.vgm_backend == "vb":
                    fitted = fit_vbgmm_1d(dev_stats.pool, cfg.n_clusters, cfg.dp_prior, device=self.device,
                                          init_centers=dev_stats.centers)
                else:
                    pool_h = dev_stats.pool.cpu().numpy()
                    fitted = [fit_sklearn_1d(p, cfg.n_clusters, cfg.dp_prior, random_state=self.seed) for p in pool_h]
                return wd_d, cont_ids_all, [f.to_array() for f in fitted], (jsd_d, wd_d, w_d)
            samples, pooled, cont_ids = [], [], []
            for ci, col in enumerate(self.meta_transformer):
                if col["type"] != CONTINUOUS:
                    continue
                cont_ids.append(ci)
                per_client = [models[i][ci].sample(int(n_total * self.model_weights_by_number[i]), self.np_rng)
                              for i in range(self.group.n)]
                samples.append(per_client)
                pooled.append(np.concatenate(per_client))
            wd = W.continuous_similarity(samples) if samples else np.zeros((self.group.n, 0))
            if self.vgm_backend == "vb" and pooled:
                fitted = fit_vbgmm_1d(np.stack(pooled), cfg.n_clusters, cfg.dp_prior, device=self.device)
            else:
                fitted = [fit_sklearn_1d(p, cfg.n_clusters, cfg.dp_prior, random_state=self.seed) for p in pooled]
            return wd, cont_ids, [f.to_array() for f in fitted], None

        wd, cont_ids, fitted, dev_w = self.group.root_compute(build)
        self.distribution_similarity_vector_continuous = wd
        self.device_similarity = dev_w          # (JSD_norm, WD_norm, weights) from the fused kernel, or None
        agg_models, agg_comps = [], []
        it = iter(fitted)
        for ci, col in enumerate(self.meta_transformer):
            if col["type"] == CONTINUOUS:
                gm = GMMParams.from_array(next(it))
                agg_models.append(gm)
                agg_comps.append(np.asarray(gm.weights > cfg.mode_eps))
            else:
                agg_models.append(None)
                agg_comps.append(None)
        self.model_aggregate, self.components_aggregate = agg_models, agg_comps
        wire = [None if m is None else m.to_array() for m in agg_models]
        self.group.call("register_transformer_information", wire, agg_comps)

    def _device_stats_ok(self):
        """Device path of the initialisation statistics: needs a CUDA federator device and the native library
        (``FEDTGAN_DEVICE_STATS=0`` forces the numpy / scipy restatement of the reference)."""
        dev = self.device
        if dev is None or dev.type != "cuda" or os.environ.get("FEDTGAN_DEVICE_STATS", "1") == "0":
            return False
        if self.group.n > 16:
            return False
        try:
            from ..ops import native  # noqa: F401
        except Exception:
            return False
        return True

    def refit_local_transformer(self):
        # the reference walks the clients sequentially (:832-835); they are independent, so
        # the group runs them concurrently
        self.group.call("refit_transformer")

    def calculate_final_weights_for_aggregation(self):
        host = W.final_weights(self.distribution_similarity_vector, self.distribution_similarity_vector_continuous,
                               self.rows_number_list)
        dev_w = getattr(self, "device_similarity", None)
        self.weights_device = None
        if dev_w is not None:
            # the fused kernel's result is the one used (and it stays on the device: the aggregation kernels
            # read it from there); the numpy restatement is the cross-check
            if np.allclose(dev_w[2], host, atol=2e-3):
                self.weights_con_cat_combination = dev_w[2] / dev_w[2].sum()
                self.weights_device = self.device_stats.weights
            else:                                   # never train on weights the two paths disagree about
                import warnings

                warnings.warn(f"device weights {dev_w[2]} disagree with host weights {host}; using host weights")
                self.weights_con_cat_combination = host
        else:
            self.weights_con_cat_combination = host
        return self.weights_con_cat_combination

    def aggregation_weights(self):
        """What ``group.aggregate`` should be given: the device tensor written by the fused statistics kernel
        when there is one (the fedavg kernels read it in place), else the host array."""
        wdev = getattr(self, "weights_device", None)
        if wdev is not None and getattr(self.group, "accepts_device_weights", False):
            return wdev
        return np.asarray(self.weights_con_cat_combination, dtype=np.float64)

    # ----------------------------------------------------------- sampling side
    def server_local_synthesizer_initialization(self):
        infos = self.group.call("get_layout_info")
        counts = self.group.call("get_option_counts")
        self.layout = TableLayout(infos[0]["output_info"])
        self.decode_tables = infos[0]["decode_tables"]
        self.steps_per_epoch = [i["steps_per_epoch"] for i in infos]
        self.option_counts = np.sum(counts, axis=0)
        _, cum_emp = cumulative_tables(self.option_counts, self.layout)
        self.cum_emp = cum_emp
        self.cond_generator = {"cum_emp": cum_emp, "counts": self.option_counts}
        self._rowgen = None

    def _row_generator(self, device):
        if self._rowgen is None or self._rowgen.device != torch.device(device):
            self._rowgen = RowGenerator(self.cfg, self.layout, self.decode_tables, self.cum_emp, device,
                                        seed=self.seed + 4242)
        return self._rowgen

    def sample(self, n):
        """Numeric synthetic table ``[n, n_columns]`` (device tensor) from the global G."""
        arena = self.group.generator_arena()
        return self._row_generator(arena.device).generate(arena, n)

    def _decode_and_write(self, table_np, index):
        path = os.path.join(self.result_dir, f"{self.name}_synthesis_epoch_{index}.csv")
        if self.meta_base.get("date_info"):
            df, _, _ = Transform.inverse(table_np, self.meta_base, self.label_encoder)
            df.to_csv(path, index=False)
            self._csv_done[index] = time.time()
            return path
        busy = sum(1 for f in self._pending if not f.done())
        plan = self._csv_plan()
        if plan is not None and table_np.dtype == np.float32:
            # fast path: the numeric table goes to text in ONE native call (no per-column numpy passes)
            plan.write(path, table_np, n_threads=(min(16, os.cpu_count() or 1) if busy <= 1 else 4))
            self._csv_done[index] = time.time()
            return path
        names, cols = Transform.decode_arrays(table_np, self.meta_base, self.label_encoder, as_codes=True)
        if any(isinstance(c, np.ndarray) and c.dtype == object for c in cols):
            names, cols = Transform.decode_arrays(table_np, self.meta_base, self.label_encoder)
        # a file that nothing else competes with (the last one of a run, or a slow-round configuration) is
        # formatted by many threads: it is the serial tail of the run; in steady state several files are in
        # flight at once and share the cores
        write_table_csv(path, names, cols, n_threads=(min(16, os.cpu_count() or 1) if busy <= 1 else 4))
        self._csv_done[index] = time.time()
        return path

    def _csv_plan(self):
        """Per-run constants of the epoch CSV for the native strided writer, or None when a column needs a host-side
        transformation (log-space non-negative columns, dates) -- then the general decode path runs."""
        if hasattr(self, "_csv_plan_cache"):
            return self._csv_plan_cache
        plan = None
        try:
            from ..data.transform import EMPTY_TOKEN
            from ..ops import hostlib

            meta = self.meta_base
            le_by_col = {item["column_name"]: item["label_encoder"] for item in self.label_encoder}
            if hostlib.available() and not meta.get("date_info") and not meta.get("non_negative_cols"):
                names = [c["column_name"] for c in meta["columns"]]
                specs = []
                for name in names:
                    if name in le_by_col:
                        specs.append(("cat", [" " if str(c) == EMPTY_TOKEN else str(c) for c in le_by_col[name].classes_]))
                    else:
                        specs.append(("real",))
                plan = hostlib.TablePlan(names, specs)
        except Exception:
            plan = None
        self._csv_plan_cache = plan
        return plan

    # ---- asynchronous generation pipeline (device path) --------------------------------
    def _async_setup(self, arena):
        from ..models.arena import GeneratorArena

        dev = arena.device
        self._gen_stream = torch.cuda.Stream(device=dev)
        self._snap = GeneratorArena(self.cfg, self.layout, dev, init=False)
        n, ncol = self.cfg.sample_rows, int(np.asarray(self.decode_tables["col_is_cont"]).size)
        self._slots = [{"dev": torch.empty(n, ncol, dtype=torch.float32, device=dev),
                        "host": torch.empty(n, ncol, dtype=torch.float32).pin_memory(),
                        "done": torch.cuda.Event(), "future": None} for _ in range(8)]
        self._slot_i = 0
        self._gen_graph = None

    def warm_generation(self):
        """One-time set-up of the asynchronous generation pipeline (pinned slots, CUDA-graph capture,
        kernel module load) on every process that will ever be a writer, so that no round pays it."""
        arena = self.group.generator_arena()
        if self._writer is None or arena.device.type != "cuda" or not self.group.local_generator:
            return
        if self._dist_gen_ok():
            self._sample_data_dist("warm", write=False)
            self.drain()
            return
        if not any(self.group.is_writer(r) for r in range(max(self.group.n, 1))):
            return
        self._sample_data_async("warm", write=False)
        self.drain()

    def _sample_data_async(self, index, write=True):
        """Snapshot G, then generate + copy to pinned host memory on a side stream (one CUDA
        graph) while the caller goes on with the next round; a worker thread waits for the
        copy and writes the CSV."""
        arena = self.group.generator_arena()
        if getattr(self, "_slots", None) is None:
            self._async_setup(arena)
        slot = self._slots[self._slot_i]
        self._slot_i = (self._slot_i + 1) % len(self._slots)
        if slot["future"] is not None:
            slot["future"].result()                 # the slot's previous CSV is on disk
        cur = torch.cuda.current_stream(arena.device)
        gen = self._row_generator(arena.device)
        self._gen_stream.wait_stream(cur)
        with torch.cuda.stream(self._gen_stream):
            self._snap.data.copy_(arena.data, non_blocking=True)
            snap_done = torch.cuda.Event()
            snap_done.record(self._gen_stream)
            cur.wait_event(snap_done)         # the next round may overwrite the arena only after the snapshot
            if getattr(self.group, "snapshot_sync", False):
                snap_done.synchronize()       # the trainer is another process (RPC star): no stream orders it
            if self._gen_graph is None:
                from ..ops import native

                native.preallocate_scratch(arena.device)
                gen.generate(self._snap, self.cfg.sample_rows, out=self._slots[0]["dev"])      # warm-up, uncaptured
                self._gen_stream.synchronize()
                self._gen_graph = torch.cuda.CUDAGraph()
                n0 = native.launch_count()
                with torch.cuda.graph(self._gen_graph, stream=self._gen_stream):
                    gen.generate(self._snap, self.cfg.sample_rows, out=self._gen_out())
                self.gen_graph_launches = native.launch_count() - n0
            self._gen_graph.replay()
            self.gen_graph_replays = getattr(self, "gen_graph_replays", 0) + 1
            slot["dev"].copy_(self._gen_out(), non_blocking=True)
            slot["host"].copy_(slot["dev"], non_blocking=True)
            slot["done"].record(self._gen_stream)
        self.d2h_bytes_last = slot["host"].numel() * 4

        def finish(slot=slot, index=index):
            slot["done"].synchronize()
            if self.write_csv and write:
                return self._decode_and_write(slot["host"].numpy(), index)
            return None

        slot["future"] = self._writer.submit(finish)
        self._pending = [f for f in self._pending if not f.done()]
        self._pending.append(slot["future"])
        return None

    def _gen_out(self):
        if getattr(self, "_gen_out_buf", None) is None:
            s0 = self._slots[0]["dev"]
            self._gen_out_buf = torch.empty_like(s0)
        return self._gen_out_buf

    def sample_data(self, index, wait=True):
        """Generate ``cfg.sample_rows`` rows with the current global G and write
        ``<Name>_result/<Name>_synthesis_epoch_<index>.csv`` (reference ``:582-590``)."""
        self._sample_calls = getattr(self, "_sample_calls", -1) + 1
        if not wait and self._dist_gen_ok():
            return self._sample_data_dist(index)        # every rank generates its share of the rows
        if not self.group.is_writer(self._sample_calls):
            return None
        if (not wait and self._writer is not None and self.group.generator_arena().device.type == "cuda"
                and self.group.local_generator):
            return self._sample_data_async(index)
        t0 = time.time()
        dev_table = self.sample(self.cfg.sample_rows)
        host = torch.empty(dev_table.shape, dtype=dev_table.dtype, pin_memory=dev_table.is_cuda)
        host.copy_(dev_table, non_blocking=False)
        self.d2h_bytes_last = host.numel() * host.element_size()
        table_np = host.numpy()
        self.metrics.add("generate_s", time.time() - t0)
        if not self.write_csv:
            return None
        if self._writer is not None and not wait:
            self._pending.append(self._writer.submit(self._decode_and_write, table_np, index))
            return None
        return self._decode_and_write(table_np, index)

    def _round_id(self):
        """Monotonic id of the next aggregation (the in-kernel barrier flags count rounds)."""
        self._rid += 1
        return self._rid

    def drain(self):
        if getattr(self, "_dg", None) is not None:
            self._dist_flush()
        for f in self._pending:
            f.result()
        self._pending = []

    # ---- distributed generation (one process per GPU) ----------------------------------------
    # After the aggregation every rank holds the same generator, so the 40 000 rows of the epoch CSV are
    # split over the ranks: rank r generates rows [r*per, (r+1)*per) with its own Philox stream on a side
    # stream and stores them straight into rank 0's gather buffer (peer-mapped CUDA-IPC memory, NVLink).
    # The hand-over needs no extra synchronisation: every rank's main stream waits for its own side stream
    # right before the NEXT round's aggregation kernel, whose in-kernel barrier then tells rank 0 that all
    # shares of the previous round have landed.  Rank 0 copies the assembled table to pinned memory and a
    # writer thread formats the CSV.  (With a single writer the 2.3 ms generation of every round sat on
    # rank 0 next to its 3 ms of training, and every rank waited for it at the barrier: SCALE_r01 0.82.)
    def _dist_gen_ok(self):
        g = self.group
        return (self._writer is not None and g.n > 1 and hasattr(g, "gather_block") and g.local_generator
                and os.environ.get("FEDTGAN_DIST_GEN", "1") != "0"
                and g.generator_arena().device.type == "cuda" and getattr(g, "_symm_ready", False))

    def _dist_setup(self, arena):
        from ..models.arena import GeneratorArena
        from .symm import _CudaBlob

        dev, g = arena.device, self.group
        n, N = self.cfg.sample_rows, g.n
        per = (n + N - 1) // N
        ncol = int(np.asarray(self.decode_tables["col_is_cont"]).size)
        slot_floats = N * per * ncol
        block = g.gather_block(2 * slot_floats)                      # collective; [2 slots][N*per rows][ncol]
        root = block.peer_ptrs[0]
        dst = [torch.as_tensor(_CudaBlob(root + (s_ * slot_floats + g.rank * per * ncol) * 4, per * ncol),
                               device=dev).view(per, ncol) for s_ in range(2)]
        self._dg = {"per": per, "ncol": ncol, "block": block, "dst": dst, "pending": None, "d2h_done": None,
                    "out": torch.empty(per, ncol, dtype=torch.float32, device=dev), "round": 0, "graph": None,
                    "gen": RowGenerator(self.cfg, self.layout, self.decode_tables, self.cum_emp, dev,
                                        seed=self.seed + 4242 + 7919 * g.rank),
                    "tail_host": torch.zeros(per, ncol, dtype=torch.float32).pin_memory()}
        if g.is_root:
            self._dg["tables"] = [block.tensor[s_ * slot_floats:(s_ + 1) * slot_floats].view(N * per, ncol)[:n]
                                  for s_ in range(2)]
            self._dg["host"] = [{"host": torch.empty(n, ncol, dtype=torch.float32).pin_memory(),
                                 "done": torch.cuda.Event(), "future": None} for _ in range(8)]
            self._dg["host_i"] = 0
        self._gen_stream = torch.cuda.Stream(device=dev)
        self._snap = GeneratorArena(self.cfg, self.layout, dev, init=False)
        g.pre_aggregate = self._dist_pre_aggregate
        plan = self._csv_plan() if self.write_csv else None
        if plan is not None:
            # the native writer keeps a per-thread formatting buffer: fault it in now on this (the launching) thread,
            # which formats this rank's share of the LAST table of a run (_dist_flush)
            warm = os.path.join(self.result_dir, f".warm_{g.rank}.csv")
            plan.write(warm, self._dg["tail_host"].numpy(), n_threads=self._tail_threads(), header=False)
            os.remove(warm)

    def _tail_threads(self):
        return max(1, min(8, (os.cpu_count() or 8) // max(self.group.n, 1)))

    def _dist_pre_aggregate(self):
        """Called by the group right before it launches the aggregation kernel: order this rank's share of the
        previous table (and, on rank 0, the previous device->host copy) before the in-kernel barrier."""
        dg = self._dg
        cur = torch.cuda.current_stream(self._snap.device)
        if dg["pending"] is not None:
            cur.wait_event(dg["pending"][2])
        if dg["d2h_done"] is not None:
            cur.wait_event(dg["d2h_done"])

    def _dist_collect(self, cur, write=True):
        """Rank 0, after an aggregation kernel has been enqueued on ``cur``: the table announced by that barrier
        is complete -> device->host copy on the side stream, CSV in a writer thread."""
        dg = self._dg
        pend, dg["pending"] = dg["pending"], None
        if pend is None or not self.group.is_root:
            return
        slot, index, _ = pend
        agg_done = torch.cuda.Event()
        agg_done.record(cur)
        h = dg["host"][dg["host_i"]]
        dg["host_i"] = (dg["host_i"] + 1) % len(dg["host"])
        if h["future"] is not None:
            h["future"].result()
        with torch.cuda.stream(self._gen_stream):
            self._gen_stream.wait_event(agg_done)
            h["host"].copy_(dg["tables"][slot], non_blocking=True)
            h["done"].record(self._gen_stream)
            dg["d2h_done"] = h["done"]
        self.d2h_bytes_last = h["host"].numel() * 4

        def finish(h=h, index=index):
            h["done"].synchronize()
            if self.write_csv and write and index is not None:
                return self._decode_and_write(h["host"].numpy(), index)
            return None

        h["future"] = self._writer.submit(finish)
        self._pending = [f for f in self._pending if not f.done()]
        self._pending.append(h["future"])

    def _sample_data_dist(self, index, write=True):
        arena = self.group.generator_arena()
        if getattr(self, "_dg", None) is None:
            self._dist_setup(arena)
        dg = self._dg
        cur = torch.cuda.current_stream(arena.device)
        self._dist_collect(cur)                          # previous round's table (announced by the aggregation just enqueued)
        slot = dg["round"] % 2
        dg["round"] += 1
        self._gen_stream.wait_stream(cur)
        with torch.cuda.stream(self._gen_stream):
            self._snap.data.copy_(arena.data, non_blocking=True)
            snap_done = torch.cuda.Event()
            snap_done.record(self._gen_stream)
            cur.wait_event(snap_done)                    # the next round may overwrite the arena only after the snapshot
            if dg["graph"] is None:
                from ..ops import native

                native.preallocate_scratch(arena.device)
                dg["gen"].generate(self._snap, dg["per"], out=dg["out"])           # warm-up, uncaptured
                self._gen_stream.synchronize()
                dg["graph"] = torch.cuda.CUDAGraph()
                n0 = native.launch_count()
                with torch.cuda.graph(dg["graph"], stream=self._gen_stream):
                    dg["gen"].generate(self._snap, dg["per"], out=dg["out"])
                self.gen_graph_launches = native.launch_count() - n0
            dg["graph"].replay()
            self.gen_graph_replays = getattr(self, "gen_graph_replays", 0) + 1
            dg["dst"][slot].copy_(dg["out"], non_blocking=True)                    # NVLink store into rank 0's buffer
            ev = torch.cuda.Event()
            ev.record(self._gen_stream)
        dg["pending"] = (slot, index if write else None, ev)
        return None

    def _dist_flush(self):
        """End of a run / of a timed region: there is no next aggregation to announce the last table."""
        dg = self._dg
        if dg["pending"] is None:
            return
        dev = self._snap.device
        slot, index, ev = dg["pending"]
        ev.synchronize()
        torch.cuda.synchronize(dev)
        plan = self._csv_plan() if (self.write_csv and index is not None) else None
        if plan is None or os.environ.get("FEDTGAN_PARALLEL_TAIL", "1") == "0":
            self.group.host_barrier()
            self._dist_collect(torch.cuda.current_stream(dev))
            return
        # The last table of a run is the serial tail of the whole job: instead of rank 0 formatting all rows, every
        # rank turns ITS share (still in its own generation buffer) into text and rank 0 concatenates the parts.
        g, n, per = self.group, self.cfg.sample_rows, dg["per"]
        rows = max(0, min(n, (g.rank + 1) * per) - g.rank * per)
        tm = [time.perf_counter()]
        dg["tail_host"].copy_(dg["out"])
        tm.append(time.perf_counter())
        final = os.path.join(self.result_dir, f"{self.name}_synthesis_epoch_{index}.csv")
        nt = self._tail_threads()
        if rows > 0:
            plan.write(f"{final}.part{g.rank}", dg["tail_host"].numpy()[:rows], n_threads=nt, header=g.is_root)
        tm.append(time.perf_counter())
        g.host_barrier()
        tm.append(time.perf_counter())
        dg["pending"] = None
        if g.is_root:
            # part 0 (with the header) becomes the file; the other parts are appended inside the kernel (sendfile)
            os.replace(f"{final}.part0", final)
            with open(final, "r+b") as out:          # (not "ab": sendfile refuses O_APPEND descriptors)
                out.seek(0, os.SEEK_END)
                for r in range(1, g.n):
                    part = f"{final}.part{r}"
                    if os.path.exists(part):
                        with open(part, "rb") as fh:
                            size, off = os.fstat(fh.fileno()).st_size, 0
                            while off < size:
                                off += os.sendfile(out.fileno(), fh.fileno(), off, size - off)
                        os.remove(part)
            self._csv_done[index] = time.time()
            self.d2h_bytes_last = n * dg["ncol"] * 4
        tm.append(time.perf_counter())
        self.tail_breakdown_ms = dict(zip(("d2h_share", "format_share", "host_barrier", "concatenate"),
                                          [round((b - a) * 1e3, 3) for a, b in zip(tm, tm[1:])]))

    # ------------------------------------------------------------------ rounds
    def fit(self):
        self.group.prepare()
        self.warm_generation()
        if self.resume:
            self._try_resume()
        w = self.aggregation_weights()
        t_fit = time.time()
        for i in range(self.start_epoch, self.epochs):
            t_in = time.time()
            self.group.run("run_local_epochs", 1)
            t_train = time.time()
            self.group.aggregate(w, self._round_id())
            t_agg = time.time()
            self.sample_data(i, wait=not self.async_csv)
            t_out = time.time()
            self.round_times.append(t_out - t_in)
            self.metrics.log(epoch=i, train_s=t_train - t_in, aggregate_s=t_agg - t_train,
                             distribute_s=t_out - t_agg, total_s=t_out - t_in)
            if self.checkpoint_every and (i + 1) % self.checkpoint_every == 0:
                self.save_checkpoint(i + 1)
        self.drain()
        self.group.barrier()
        # timestamp_experiment.csv keeps the reference's meaning (``distributed.py:795-829``: the round's clock stops
        # after its CSV is written, and similarity_analysis.py cumulative-sums the rows into the time at which
        # each epoch's output existed).  Rounds overlap here (the CSV of round i is written while round i+1
        # trains), so row i is the time between CSV i-1 and CSV i being on disk; the host-side enqueue times
        # stay in metrics.jsonl.
        done = [self._csv_done.get(i) for i in range(self.start_epoch, self.epochs)]
        if self.write_csv and self.group.is_root and done and all(d is not None for d in done):
            stamps = np.maximum.accumulate(np.asarray(done)) - t_fit
            restored = list(self.round_times[: self.start_epoch])          # rounds of a resumed run's first leg
            self.round_times = restored + [float(x) for x in np.diff(np.concatenate([[0.0], stamps]))]
        if self.group.is_root:
            with open(os.path.join(self.out_dir, "timestamp_experiment.csv"), "w", newline="") as fh:
                csv.writer(fh, dialect="excel").writerows(np.asarray(self.round_times).reshape(-1, 1))
        return self.round_times

    # -------------------------------------------------------------- persistence
    def save_model(self, filename):
        """``torch.save([generator, cond_generator, transformer-info, batch_size, embedding_dim])``
        under ``models/model_list_rpc_<filename>.pt`` (reference ``:560-563``)."""
        arena = self.group.generator_arena()
        obj = [arena.to_module(), self.cond_generator,
               {"output_info": self.layout.output_info, "decode_tables": self.decode_tables,
                "meta": self.meta_base}, self.batch_size, self.embedding_dim]
        path = os.path.join(self.models_dir, f"model_list_rpc_{filename}.pt")
        if self.group.is_root:
            torch.save(obj, path)
        return path

    def checkpoint_path(self):
        return os.path.join(self.models_dir, f"checkpoint_{self.name}.pt")

    def save_checkpoint(self, next_epoch):
        states = self.group.call("checkpoint_state")
        if self.group.is_root:
            torch.save({"epoch": int(next_epoch), "clients": states,
                        "weights": np.asarray(self.weights_con_cat_combination),
                        "round_times": list(self.round_times), "config": self.cfg.to_dict()},
                       self.checkpoint_path())

    def _try_resume(self):
        path = self.checkpoint_path()
        ck = self.group.root_compute(lambda: torch.load(path, weights_only=False) if os.path.exists(path) else None)
        if ck is None:
            return False
        self.group.call("load_checkpoint_state", per_client=[(s,) for s in ck["clients"]])
        self.start_epoch = int(ck["epoch"])
        self.round_times = list(ck.get("round_times", []))
        return True
