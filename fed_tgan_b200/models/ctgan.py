"""CTGAN networks and losses as plain PyTorch modules (the oracle / CPU path).

Parity targets (all in ``Server/dtds/synthesizers/ctgan.py``): ``Discriminator`` (:15-30),
``Residual`` (:33-44), ``Generator`` (:47-64), ``apply_activate`` (:67-82), ``cond_loss``
(:174-194), ``slerp`` (:231-237), ``calc_gradient_penalty`` (:240-258).  The modules keep
the reference's ``state_dict`` keys/shapes exactly (``seq.{i}.fc.weight`` ...,
``seq.{0,3,6}.weight`` ...), because that layout is the checkpoint / interchange format.

On the device path these modules are *not* executed: ``fed_tgan_b200.engine`` runs the same
maths as hand-written kernels over a flat parameter arena
(``fed_tgan_b200.models.arena``), and converts to/from these ``state_dict``s at the edges.
"""

from __future__ import annotations

import torch
from torch import nn
from torch.nn import functional as F


class Discriminator(nn.Module):
    """PacGAN discriminator: ``pack`` rows are concatenated into one sample."""

    def __init__(self, input_dim, dis_dims, pack=10, leaky_slope=0.2, dropout_p=0.5):
        super().__init__()
        self.pack = pack
        self.packdim = input_dim * pack
        layers = []
        width = self.packdim
        for h in dis_dims:
            layers += [nn.Linear(width, h), nn.LeakyReLU(leaky_slope), nn.Dropout(dropout_p)]
            width = h
        layers.append(nn.Linear(width, 1))
        self.seq = nn.Sequential(*layers)

    def forward(self, x):
        if x.size(0) % self.pack != 0:
            raise ValueError("batch must be a multiple of pack")
        return self.seq(x.view(-1, self.packdim))


class Residual(nn.Module):
    def __init__(self, i, o):
        super().__init__()
        self.fc = nn.Linear(i, o)
        self.bn = nn.BatchNorm1d(o)
        self.relu = nn.ReLU()

    def forward(self, x):
        return torch.cat([self.relu(self.bn(self.fc(x))), x], dim=1)


class Generator(nn.Module):
    def __init__(self, embedding_dim, gen_dims, data_dim):
        super().__init__()
        width = embedding_dim
        blocks = []
        for h in gen_dims:
            blocks.append(Residual(width, h))
            width += h
        blocks.append(nn.Linear(width, data_dim))
        self.seq = nn.Sequential(*blocks)

    def forward(self, x):
        return self.seq(x)


def apply_activate(data, output_info, tau=0.2):
    """tanh on scalar spans, soft gumbel-softmax(tau) on one-hot spans."""
    parts, st = [], 0
    for width, kind in output_info:
        ed = st + width
        if kind == "tanh":
            parts.append(torch.tanh(data[:, st:ed]))
        elif kind == "softmax":
            parts.append(F.gumbel_softmax(data[:, st:ed], tau=tau))
        else:
            raise ValueError(kind)
        st = ed
    return torch.cat(parts, dim=1)


def cond_loss(data, output_info, c, m):
    """Masked cross-entropy between generated logits and the conditional vector."""
    losses, st, st_c = [], 0, 0
    for width, kind in output_info:
        if kind == "tanh":
            st += width
            continue
        ed, ed_c = st + width, st_c + width
        target = torch.argmax(c[:, st_c:ed_c], dim=1)
        losses.append(F.cross_entropy(data[:, st:ed], target, reduction="none"))
        st, st_c = ed, ed_c
    stacked = torch.stack(losses, dim=1)
    return (stacked * m).sum() / data.size(0)


def slerp(val, low, high):
    """Row-wise spherical interpolation (the reference's replacement for the linear
    WGAN-GP mix)."""
    low_n = low / torch.norm(low, dim=1, keepdim=True)
    high_n = high / torch.norm(high, dim=1, keepdim=True)
    omega = torch.acos((low_n * high_n).sum(1)).view(val.size(0), 1)
    so = torch.sin(omega)
    return (torch.sin((1.0 - val) * omega) / so) * low + (torch.sin(val * omega) / so) * high


def calc_gradient_penalty(netD, real_data, fake_data, device="cpu", pac=10, lambda_=10, alpha=None):
    if alpha is None:
        alpha = torch.rand(real_data.size(0), 1, device=device)
    interpolates = slerp(alpha, real_data, fake_data)
    if not interpolates.requires_grad:           # inputs without history (e.g. fake rows received over RPC)
        interpolates.requires_grad_(True)
    disc = netD(interpolates)
    grads = torch.autograd.grad(
        outputs=disc, inputs=interpolates, grad_outputs=torch.ones_like(disc),
        create_graph=True, retain_graph=True, only_inputs=True,
    )[0]
    return ((grads.view(-1, pac * real_data.size(1)).norm(2, dim=1) - 1) ** 2).mean() * lambda_
