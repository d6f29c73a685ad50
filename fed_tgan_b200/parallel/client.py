"""Client worker: owns one data shard, one GPU, one replica of G and D.

RPC-visible surface kept from the reference's ``MDGANClient``
(``Server/dtds/distributed.py:210-542`` == ``Client/.../dtds/distributed.py:62-393``):
``get_meta``, ``register_meta``, ``get_transformer_information``,
``register_transformer_information``, ``refit_transformer``, ``get_transformer``,
``get_models``, ``train_model``, ``set_generator_weights``, ``set_discriminator_weights``
and the legacy MD-GAN / GDTS split-training API (``send_client_refs``, ``register_G``,
``train_D``, ``loss_G``, ``get_discriminator_weights``, ``reset_on_cuda``,
``save_time_stamp``, ``encode_data``, ``get_cond_generator``).

What differs underneath:

* the encoded shard lives on the device in compact form (``DeviceTable``); conditional
  vectors, permutations and real rows are sampled by kernels -- no numpy in the loop;
* parameters live in flat arenas (``GeneratorArena`` / ``DiscriminatorArena``), optionally
  inside a peer-mapped symmetric allocation, and the local epoch is the closed-form op
  program of ``engine.TrainEngine`` (CUDA-graph replayed on the device);
* ``train_model`` does not move the models to the CPU; ``state_dict`` conversion happens
  only when a caller actually asks for one (reference: ``.cpu()`` every round, ``:414-417``);
* mixtures travel as ``GMMParams`` arrays instead of pickled sklearn objects;
* ``get_option_counts`` exports per-option counts so the federator can build its sampling
  distribution without the raw rows (reference loads the full table on the server,
  ``:565-580``).
"""

from __future__ import annotations

import csv
import os
import time

import numpy as np
import torch

from ..config import CTGANConfig
from ..data.load import encode_data_with_meta_labelencoder, prepare_data
from ..engine import DeviceTable, TrainEngine
from ..features.gmm import GMMParams
from ..features.transformers import BGM_CTGAN_Transformer
from ..models import ctgan


def _pick_device(device=None):
    if device is not None:
        return torch.device(device)
    if torch.cuda.is_available():
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def _on_device(fn):
    """RPC handlers run on arbitrary worker threads whose current CUDA device is 0: make the
    client's own GPU current for the duration of the call (streams, graphs and the native
    library all follow the current device)."""
    import functools

    @functools.wraps(fn)
    def wrapper(self, *a, **k):
        if self.device.type == "cuda":
            with torch.cuda.device(self.device):
                return fn(self, *a, **k)
        return fn(self, *a, **k)

    return wrapper


class MDGANClient:
    def __init__(self, datapath, selected_variables, categorical_list, nonnegative_list, date_dic, target_column,
                 problem_type, epochs, device=None, config: CTGANConfig | dict | None = None, dataframe=None,
                 vgm_backend="sklearn", write_side_files=True, client_id=0, split=None, **kwargs):
        if isinstance(config, dict):
            config = CTGANConfig.from_dict(config)
        self.cfg = config if config is not None else CTGANConfig(**kwargs)
        # hyper-parameter attributes under the reference's names (CTGANSynthesizer.__init__)
        self.embedding_dim, self.gen_dim, self.dis_dim = self.cfg.embedding_dim, self.cfg.gen_dim, self.cfg.dis_dim
        self.l2scale, self.batch_size = self.cfg.l2scale, self.cfg.batch_size
        self.device = _pick_device(device)
        self.client_id = client_id
        self.datapath = datapath
        self.selected_variables = list(selected_variables)
        self.categorical_list = list(categorical_list)
        self.nonnegative_list = list(nonnegative_list)
        self.date_dic = dict(date_dic or {})
        self.epochs = epochs
        self.target_column = target_column
        self.problem_type = problem_type
        self.vgm_backend = vgm_backend
        self.write_side_files = write_side_files
        if split is not None and dataframe is None:
            # (mode, n_clients, index, seed): every client reads the same file and keeps its own shard
            import pandas as pd

            from ..data.schemas import split_table

            mode, n_shards, idx, sseed = split
            dataframe = split_table(pd.read_csv(datapath), int(n_shards), mode, seed=int(sseed))[int(idx)]
        self._dataframe = dataframe
        self.meta_file = prepare_data(self.datapath, self.selected_variables, self.categorical_list,
                                      self.nonnegative_list, self.date_dic, self.target_column, self.problem_type,
                                      dataframe=dataframe)
        self.filename = os.path.basename(self.datapath).split(".")[0]
        self.time_train_d, self.time_loss_g = [], []
        self.engine: TrainEngine | None = None
        self.G_rref = None
        self.symm = None
        self._graph = None

    # ------------------------------------------------------------ meta / encoders
    def get_meta(self):
        return self.meta_file

    def register_meta(self, meta, timestamp, label_encoder):
        self.meta_file = meta
        self.timestamp = timestamp
        self.label_encoder = label_encoder
        return True

    def encode_data(self):
        self.train, self.cat_cols, self.ord_cols = encode_data_with_meta_labelencoder(
            self.datapath, self.meta_file, self.label_encoder, self.selected_variables, self.categorical_list,
            self.nonnegative_list, self.date_dic, self.target_column, self.problem_type, self.timestamp,
            dataframe=self._dataframe, write_side_files=self.write_side_files)
        self.rows = self.train.shape[0]

    @_on_device
    def get_transformer_information(self):
        """Encode with the global label encoders, fit local per-column mixtures and return
        ``(models, component masks, transformer meta, n_rows)``."""
        self.encode_data()
        self.transformer = BGM_CTGAN_Transformer(
            n_clusters=self.cfg.n_clusters, eps=self.cfg.mode_eps, backend=self.vgm_backend,
            dp_prior=self.cfg.dp_prior, device=self.device if self.vgm_backend == "vb" else None,
            seed=self.cfg.seed + 17 * (self.client_id + 1))
        self.transformer.fit(self.train, self.cat_cols, self.ord_cols)
        self.model, self.components, self.meta_transformer = self.transformer.get_information()
        wire = [None if m is None else m.to_array() for m in self.model]
        return wire, self.components, self.meta_transformer, self.rows

    def register_transformer_information(self, model, components):
        self.model = [None if m is None else (GMMParams.from_array(m) if isinstance(m, np.ndarray) else GMMParams.coerce(m))
                      for m in model]
        self.components = components
        return True

    @_on_device
    def refit_transformer(self):
        """Adopt the global encoders, encode the shard, build sampler tables, networks and
        optimiser state (reference ``:295-317``)."""
        self.transformer.refit(self.train, self.meta_file, self.label_encoder, self.cat_cols, self.ord_cols,
                               self.model, self.components)
        if self.device.type == "cuda" and self.vgm_backend == "vb":
            dv, dc = self.transformer.transform_compact_device(self.train, self.device)     # K12 on the device
            vals, codes = dv.cpu().numpy(), dc.cpu().numpy()
        else:
            vals, codes = self.transformer.transform_compact(self.train)
        self.out_info = self.transformer.output_info
        self.data_dim = self.transformer.output_dim
        self.layout = self.transformer.layout
        self.table = DeviceTable(vals, codes, self.layout, self.device, pinned_host=self.device.type == "cuda")
        self.steps_per_epoch = len(self.train) // self.batch_size
        self._build_engine()
        return True

    def _build_engine(self, g_storage=None, d_storage=None):
        from ..models.arena import DiscriminatorArena, GeneratorArena

        seed = self.cfg.seed + 1000 * (self.client_id + 1)   # every client initialises its own replica
        G = GeneratorArena(self.cfg, self.layout, self.device, storage=g_storage, seed=seed)
        D = DiscriminatorArena(self.cfg, self.layout, self.device, storage=d_storage, seed=seed + 1)
        self.engine = TrainEngine(self.cfg, self.layout, self.table, self.device, G, D, seed=seed)
        self._graph = None

    def get_option_counts(self):
        """Per-option counts of the encoded shard (== ``Cond.model`` statistics)."""
        return self.table.counts.copy()

    def get_layout_info(self):
        return {"output_info": self.out_info, "decode_tables": self.transformer.decode_tables(),
                "rows": self.rows, "steps_per_epoch": self.steps_per_epoch}

    def get_transformer(self):
        return self.transformer

    def get_cond_generator(self):
        return {"cum_p": self.table.cum_p.cpu().numpy(), "cum_emp": self.table.cum_emp.cpu().numpy(),
                "n_col": self.layout.n_col, "n_opt": self.layout.n_opt}

    # ----------------------------------------------------------------- weights
    @_on_device
    def get_models(self):
        """Reference-layout CPU modules (``ctgan.Generator``, ``ctgan.Discriminator``)."""
        return self.engine.G.to_module(), self.engine.D.to_module()

    @_on_device
    def get_state_dicts(self):
        return ({k: v.cpu() for k, v in self.engine.G.state_dict().items()},
                {k: v.cpu() for k, v in self.engine.D.state_dict().items()})

    @_on_device
    def get_discriminator_weights(self):
        return {k: v.cpu() for k, v in self.engine.D.state_dict().items()}

    @_on_device
    def get_generator_weights(self):
        return {k: v.cpu() for k, v in self.engine.G.state_dict().items()}

    @_on_device
    def set_discriminator_weights(self, state_dict):
        self.engine.D.load_state_dict(state_dict)
        return True

    @_on_device
    def set_generator_weights(self, state_dict):
        self.engine.G.load_state_dict(state_dict)
        return True

    def reset_on_cuda(self):
        return True   # arenas never leave the device

    # ------------------------------------------------------------------ training
    @_on_device
    def run_local_epochs(self, number_of_epoch=1, use_graph=None):
        """Local training without any host read-back (the hot path)."""
        steps = int(number_of_epoch) * self.steps_per_epoch
        if steps <= 0:
            return 0
        if use_graph is None:
            use_graph = self.device.type == "cuda"
        if use_graph:
            from ..engine.graph import GraphedIteration

            if self._graph is None:
                self._graph = GraphedIteration(self.engine)
            self._graph.run(steps)
        else:
            self.engine.train(steps)
        return steps

    @_on_device
    def train_model(self, number_of_epoch=1):
        """Reference API: train ``number_of_epoch`` local epochs and return CPU state_dicts."""
        self.run_local_epochs(number_of_epoch)
        return self.get_state_dicts()

    @_on_device
    def get_losses(self):
        return self.engine.losses() if self.engine is not None else {}

    def read_losses(self):
        """Device->host read of the round's losses (16 bytes)."""
        self.last_losses = self.get_losses()
        return self.last_losses

    @_on_device
    def restage_shard(self):
        """Host(pinned)->device copy of this client's compact shard (out-of-core / e2e path)."""
        return self.table.restage()

    def graph_replays(self):
        return self._graph.replays if self._graph is not None else 0

    def graph_launches(self):
        return self._graph.launches_per_replay if self._graph is not None else 0

    # --------------------------------------------------------------- checkpoint
    @_on_device
    def checkpoint_state(self):
        e = self.engine
        if self.symm is not None:
            self.symm.check()        # never checkpoint an arena a timed-out aggregation left half-averaged
        return {
            "G": e.G.state_dict(), "D": e.D.state_dict(),
            "optG": e.G.optimizer_state(int(e.rng.state[1].item())),
            "optD": e.D.optimizer_state(int(e.rng.state[1].item())),
            "rng": e.rng.state.cpu().clone(), "iterations": e.iterations,
        }

    @_on_device
    def load_checkpoint_state(self, st):
        e = self.engine
        e.G.load_state_dict(st["G"])
        e.D.load_state_dict(st["D"])
        e.G.load_optimizer_state(st["optG"])
        e.D.load_optimizer_state(st["optD"])
        e.rng.state.copy_(st["rng"].to(e.rng.state.device))
        e.rng.seed = int(st["rng"][0])
        e.rng.resync()
        e.iterations = int(st.get("iterations", 0))
        return True

    # ------------------------------------------------- symmetric memory (GPU only)
    @_on_device
    def symm_export(self):
        """Allocate the peer-visible arena block and return its IPC handle bytes."""
        from .symm import SymmetricBlock

        e = self.engine
        need = e.G.numel + e.D.numel
        self.symm = SymmetricBlock(need, self.device)
        self._rebind_storage(self.symm.tensor)
        return self.symm.export_handle()

    def _rebind_storage(self, block):
        """Move both arenas into ``block`` (G first, then D), keeping weights, Adam moments, the
        Philox counter and the iteration count, and rebuild the engine over the new storage."""
        e = self.engine
        st = self.checkpoint_state()
        g_raw, d_raw = e.G.data.clone(), e.D.data.clone()
        self._build_engine(block[: e.G.numel], block[e.G.numel: e.G.numel + e.D.numel])
        self.load_checkpoint_state(st)
        self.engine.G.data.copy_(g_raw)          # bit-exact weights, padding included
        self.engine.D.data.copy_(d_raw)

    @_on_device
    def symm_setup_nvls(self, group=None):
        """Collective (SPMD only): put the arenas into a multicast-bound symmetric allocation so the
        aggregation can run through the NVSwitch reduction engine (``ftg_fedavg_nvls``)."""
        from .symm import NvlsBlock

        e = self.engine
        self.symm = NvlsBlock(e.G.numel + e.D.numel, self.device, group)
        self._rebind_storage(self.symm.tensor)
        return True

    @_on_device
    def symm_connect(self, rank, handles):
        self.symm.connect(rank, handles)
        return True

    @_on_device
    def symm_aggregate(self, weights, round_id, sync=False):
        """Launch the fused weighted all-reduce on this client's GPU (all clients must be
        launched concurrently; they synchronise inside the kernel over NVLink)."""
        e = self.engine
        ints = [o for o in e.G.int_slot_offsets()]
        self.symm.fedavg(weights, e.G.numel + e.D.numel, ints, round_id)
        if sync:
            self.symm.check()        # raises if a peer never arrived at the in-kernel barrier
        else:
            self.symm.poll()         # same, one round late, without a host-device sync
        return True

    # ------------------------------------- legacy MD-GAN / GDTS split-training API
    def send_client_refs(self):
        import torch.distributed.rpc as rpc

        self.discriminator = self.engine.D.to_module().to(self.device)
        return rpc.RRef(self.discriminator), self.get_cond_generator(), self.transformer, self.steps_per_epoch

    def register_G(self, G_rref):
        self.G_rref = G_rref
        return True

    def _legacy_setup(self):
        if not hasattr(self, "discriminator"):
            self.discriminator = self.engine.D.to_module().to(self.device)
        if not hasattr(self, "optD"):
            self.optD = torch.optim.Adam(self.discriminator.parameters(), lr=self.cfg.lr, betas=self.cfg.betas)

    def _remote_generate(self, fakez, timings):
        t0 = time.time()
        if self.G_rref is None:
            raise RuntimeError("register_G() must be called before the MD-GAN split API is used")
        out = self.G_rref.rpc_sync().forward(fakez.cpu()).to(self.device)
        fakeact = ctgan.apply_activate(out, self.out_info, tau=self.cfg.tau)
        timings.append(time.time() - t0)
        return out, fakeact

    @_on_device
    def train_D(self):
        """MD-GAN mode: the generator lives on the server, D is trained here (``:421-480``)."""
        self._legacy_setup()
        e, B = self.engine, self.batch_size
        loss_d = pen = torch.zeros(())
        for _ in range(self.steps_per_epoch):
            e.rng.tick()
            c1 = e._prep(None, "d", with_real=True)
            h0 = e.off[0]
            fakez = e.h[:, h0:h0 + e.d0].clone()
            _, fakeact = self._remote_generate(fakez, self.time_train_d)
            fake_cat = torch.cat([fakeact, c1], dim=1)
            real_cat = e.X[B:2 * B, : e.din].clone()
            y_fake, y_real = self.discriminator(fake_cat), self.discriminator(real_cat)
            loss_d = -(y_real.mean() - y_fake.mean())
            pen = ctgan.calc_gradient_penalty(self.discriminator, real_cat, fake_cat, self.device,
                                              pac=self.cfg.pack, lambda_=self.cfg.gp_lambda)
            self.optD.zero_grad()
            pen.backward(retain_graph=True)
            loss_d.backward()
            self.optD.step()
        return loss_d.detach().cpu(), pen.detach().cpu()

    @_on_device
    def loss_G(self):
        """MD-GAN mode: generator loss evaluated against the local D (``:483-525``)."""
        self._legacy_setup()
        e = self.engine
        total = 0.0
        for _ in range(self.steps_per_epoch):
            e.rng.tick()
            c1 = e._prep(None, "g", with_real=False)
            h0 = e.off[0]
            fakez = e.h[:, h0:h0 + e.d0].clone()
            fake, fakeact = self._remote_generate(fakez, self.time_loss_g)
            y_fake = self.discriminator(torch.cat([fakeact, c1], dim=1))
            ce = ctgan.cond_loss(fake, self.out_info, c1, e.m1)
            total = total + (-y_fake.mean() + ce).detach().cpu()
        return total

    def save_time_stamp(self):
        for name, series in (("time_train_d.csv", self.time_train_d), ("time_loss_g.csv", self.time_loss_g)):
            with open(name, "w", newline="") as fh:
                csv.writer(fh, dialect="excel").writerows(np.asarray(series).reshape(-1, 1))
        return True
