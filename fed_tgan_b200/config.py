"""Hyper-parameters.

The reference keeps these as constructor defaults of ``CTGANSynthesizer``
(``Server/dtds/synthesizers/ctgan.py:310-334``) plus literals scattered through the code
(``pack=10`` ctgan.py:16, ``tau=0.2`` ctgan.py:77, ``lambda_=10`` ctgan.py:240, Adam
``lr=2e-4, betas=(0.5,0.9)`` distributed.py:315-317, ``n_clusters=10, eps=0.005``
transformers.py:316, sample size 40000 distributed.py:583).  Here they are one dataclass
that every layer takes, so the wide-table config (1024-wide nets) is just another instance.
"""

from __future__ import annotations

from dataclasses import dataclass, asdict
from typing import Tuple


@dataclass
class CTGANConfig:
    embedding_dim: int = 128
    gen_dim: Tuple[int, ...] = (256, 256)
    dis_dim: Tuple[int, ...] = (256, 256)
    l2scale: float = 1e-6          # Adam weight_decay of the generator (L2-in-gradient)
    batch_size: int = 500
    epochs: int = 3
    pack: int = 10                 # PacGAN packing of the discriminator
    tau: float = 0.2               # gumbel-softmax temperature
    gp_lambda: float = 10.0        # gradient-penalty weight
    lr: float = 2e-4
    betas: Tuple[float, float] = (0.5, 0.9)
    adam_eps: float = 1e-8
    leaky_slope: float = 0.2
    dropout_p: float = 0.5
    bn_eps: float = 1e-5
    bn_momentum: float = 0.1
    # encoder
    n_clusters: int = 10
    mode_eps: float = 0.005
    dp_prior: float = 0.001
    # generation
    sample_rows: int = 40000
    # GEMM arithmetic on the device path (applied by TrainEngine): "fp32x3" (default; tcgen05 with the
    # 3xTF32 hi/lo split, fp32-equivalent like the reference's fp32 matmuls), "tf32" (single-pass
    # tcgen05 kind::tf32, opt-in) or "fp32" (CUDA-core FMA kernels)
    precision: str = "fp32x3"
    seed: int = 0

    def __post_init__(self):
        self.gen_dim = tuple(int(x) for x in self.gen_dim)
        self.dis_dim = tuple(int(x) for x in self.dis_dim)
        self.betas = tuple(float(b) for b in self.betas)
        if self.batch_size % self.pack != 0:
            raise ValueError("batch_size must be a multiple of pack")

    def to_dict(self):
        return asdict(self)

    @classmethod
    def from_dict(cls, d):
        known = {k: v for k, v in d.items() if k in cls.__dataclass_fields__}
        return cls(**known)


def wide_config() -> CTGANConfig:
    """BASELINE.json config 5: enlarged generator/discriminator (1024-wide Residual)."""
    return CTGANConfig(gen_dim=(1024, 1024), dis_dim=(1024, 1024))
