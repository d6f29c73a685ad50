"""fed_tgan_b200 -- Blackwell-native federated tabular-GAN framework.

Capabilities mirror zhao-zilong/Fed-TGAN (federated CTGAN with similarity-weighted
aggregation and federated table encoding); the architecture is B200-first:

* ``data/``, ``features/``  table ingestion, federated VGM / one-hot encoders
* ``models/``               Generator / Discriminator (reference ``state_dict`` layout),
                            flat parameter arenas
* ``ops/``                  op surface: hand-written sm_100a kernels (``csrc/``) with a
                            plain fp32 PyTorch restatement of every op as oracle / CPU path
* ``engine/``               closed-form CTGAN train step (no autograd on the hot path),
                            device-resident sampling, CUDA-graph capture
* ``parallel/``             federator / client, symmetric peer memory, fused weighted
                            all-reduce over NVLink, RPC control plane, launcher
* ``eval/``                 Avg_JSD / Avg_WD / ML-utility analysis
"""

from .config import CTGANConfig  # noqa: F401

__version__ = "0.1.0"
