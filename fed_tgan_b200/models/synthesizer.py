"""Stand-alone (non-federated) CTGAN trainer with ``fit`` / ``sample``.

Parity: ``CTGANSynthesizer`` (``Server/dtds/synthesizers/ctgan.py:309-487``) and
``BaseSynthesizer`` (``synthesizers/base.py:1-30``; its methods recurse into themselves in
the reference -- here they are a real abstract interface).  ``fit`` uses ``BGMTransformer``
and the same closed-form engine as the federated clients, ``sample`` the device row
generator.  ``local_main`` is the stand-alone CLI (``Server/dtds/local.py`` is dead code in
the reference because its imports do not exist).
"""

from __future__ import annotations

import time

import numpy as np
import torch

from ..config import CTGANConfig


class BaseSynthesizer:
    """Interface of every synthesizer (SDGym heritage)."""

    def fit(self, data, categorical_columns=tuple(), ordinal_columns=tuple(), bimodal_columns=None):
        raise NotImplementedError

    def sample(self, samples):
        raise NotImplementedError

    def sample_indicating_column(self, samples, column):
        raise NotImplementedError

    def sample_indicating_pair(self, samples, pair):
        raise NotImplementedError

    def fit_sample(self, data, categorical_columns=tuple(), ordinal_columns=tuple(), bimodal_columns=tuple()):
        self.fit(data, categorical_columns, ordinal_columns)
        return self.sample(np.asarray(data).shape[0])


class CTGANSynthesizer(BaseSynthesizer):
    def __init__(self, embedding_dim=128, gen_dim=(256, 256), dis_dim=(256, 256), l2scale=1e-6, batch_size=500,
                 epochs=3, device=None, config: CTGANConfig | None = None, vgm_backend="sklearn", verbose=True):
        self.__name__ = "CTGANSynth"
        self.cfg = config or CTGANConfig(embedding_dim=embedding_dim, gen_dim=gen_dim, dis_dim=dis_dim,
                                         l2scale=l2scale, batch_size=batch_size, epochs=epochs)
        self.embedding_dim, self.gen_dim, self.dis_dim = self.cfg.embedding_dim, self.cfg.gen_dim, self.cfg.dis_dim
        self.l2scale, self.batch_size, self.epochs = self.cfg.l2scale, self.cfg.batch_size, self.cfg.epochs
        self.device = torch.device(device) if device is not None else torch.device(
            "cuda:0" if torch.cuda.is_available() else "cpu")
        self.vgm_backend = vgm_backend
        self.verbose = verbose

    def fit(self, train_data, categorical_columns=tuple(), ordinal_columns=tuple(), bimodal_columns=None):
        from ..engine import DeviceTable, TrainEngine
        from ..features.transformers import BGMTransformer

        train_data = np.asarray(train_data, dtype=np.float64)
        self.transformer = BGMTransformer(n_clusters=self.cfg.n_clusters, eps=self.cfg.mode_eps,
                                          backend=self.vgm_backend, dp_prior=self.cfg.dp_prior, seed=self.cfg.seed)
        self.transformer.fit(train_data, categorical_columns, ordinal_columns)
        self.out_info = self.transformer.output_info
        vals, codes = self.transformer.transform_compact(train_data)
        self.table = DeviceTable(vals, codes, self.transformer.layout, self.device)
        self.engine = TrainEngine(self.cfg, self.transformer.layout, self.table, self.device)
        self.generator = self.engine.G
        steps = len(train_data) // self.batch_size
        for i in range(self.epochs):
            t0 = time.time()
            self.engine.train(steps)
            if self.verbose and steps:
                ls = self.engine.losses()
                print(f"EPOCH {i}:   loss_d:{ls['loss_d']:>6.2f}   loss_g:{ls['loss_g']:>6.2f}   "
                      f"time taken: {time.time() - t0:.2f} sec")
        return self

    def sample(self, n):
        from ..engine.generate import RowGenerator

        gen = RowGenerator(self.cfg, self.transformer.layout, self.transformer.decode_tables(),
                           self.table.cum_emp, self.device, seed=self.cfg.seed + 7)
        return gen.generate(self.engine.G, int(n)).cpu().numpy().astype(np.float64)


def local_main(argv=None):
    """``python -m fed_tgan_b200.models.synthesizer -dataset <csv> -epochs N [-report]``."""
    import argparse
    import pandas as pd
    from sklearn.preprocessing import LabelEncoder

    p = argparse.ArgumentParser()
    p.add_argument("-dataset", required=True, help="CSV path")
    p.add_argument("-categorical_list", default="")
    p.add_argument("-epochs", default=3, type=int)
    p.add_argument("-report", action="store_true")
    p.add_argument("-out", default=None)
    a = p.parse_args(argv)
    df = pd.read_csv(a.dataset)
    cats = [c for c in a.categorical_list.split(",") if c]
    encoders = {c: LabelEncoder().fit(df[c].astype(str)) for c in cats}
    enc = df.copy()
    for c, le in encoders.items():
        enc[c] = le.transform(df[c].astype(str))
    synth = CTGANSynthesizer(epochs=a.epochs)
    synth.fit(enc.to_numpy(dtype=np.float64), [df.columns.get_loc(c) for c in cats])
    if a.report:
        out = pd.DataFrame(synth.sample(1000), columns=df.columns)
        for c, le in encoders.items():
            out[c] = le.inverse_transform(out[c].astype(int))
        print(out.head())
        if a.out:
            out.to_csv(a.out, index=False)
    return synth


if __name__ == "__main__":
    local_main()
