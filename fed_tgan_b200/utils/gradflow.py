"""Gradient-flow diagnostics (reference: ``update_grad_flow`` / ``plot_grad_flow``,
``Server/dtds/synthesizers/ctgan.py:261-306``, dead code there behind a TkAgg import).
Here the statistics come from the arena gradients and the plot is rendered headless (Agg)
to a PNG when matplotlib is installed; otherwise the series are dumped as JSON."""

from __future__ import annotations

import json
import os

ave_grads = {"G": [], "D": []}
layers = {"G": [], "D": []}


def update_grad_flow(arena, net: str):
    grads = arena.grad_state_dict()
    layers[net] = list(grads.keys())
    ave_grads[net].append([float(g.abs().mean()) for g in grads.values()])


def plot_grad_flow(save_path: str):
    os.makedirs(save_path, exist_ok=True)
    try:
        import matplotlib

        matplotlib.use("Agg")
        from matplotlib import pyplot as plt
    except Exception:
        with open(os.path.join(save_path, "grad_flow.json"), "w") as fh:
            json.dump({"layers": layers, "ave_grads": ave_grads}, fh)
        return os.path.join(save_path, "grad_flow.json")
    _, ax = plt.subplots(1, 2, figsize=(12, 6))
    for p, net in enumerate(["D", "G"]):
        for series in ave_grads[net]:
            ax[p].plot(layers[net], series, alpha=0.5)
        ax[p].tick_params(axis="x", rotation=30)
        ax[p].set_xlabel(f"Layers {net}")
    ax[0].set_ylabel("Average gradient")
    plt.suptitle("Gradient flow")
    out = os.path.join(save_path, "grad_flow.png")
    plt.savefig(out)
    return out
