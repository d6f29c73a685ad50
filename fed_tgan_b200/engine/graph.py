"""CUDA-graph capture of one full training iteration.

The reference's step is dominated by Python/launch latency.  The engine's
iteration is ~100 launches of the native library on one stream with no host decisions and
device-resident RNG/step counters, so it captures into a single ``cudaGraph`` that is
replayed ``steps_per_epoch`` times with one host call each.
"""

from __future__ import annotations

import torch


class GraphedIteration:
    def __init__(self, engine, warmup: int = 2):
        self.engine = engine
        self.graph = None
        dev = engine.device
        assert dev.type == "cuda"
        from ..ops import native

        # High-priority stream: the iteration's tiny latency-bound kernels must not queue behind the
        # federator's big generation GEMMs that run concurrently on a default-priority stream.
        self.stream = torch.cuda.Stream(device=dev, priority=-1)
        self.stream.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(self.stream):
            native.preallocate_scratch(dev)
            # The warm-up iterations (module load, scratch sizing) must not count as training: the reference's
            # first local epoch is exactly rows // batch steps (`distributed.py:334`).  Snapshot everything an
            # iteration changes -- weights + BN buffers, Adam moments, the Philox (seed, step) state -- and put it
            # back after the warm-up.
            saved = [t.clone() for t in (engine.G.data, engine.D.data, engine.G.exp_avg, engine.G.exp_avg_sq,
                                         engine.D.exp_avg, engine.D.exp_avg_sq, engine.rng.state)]
            it0 = engine.iterations
            for _ in range(warmup):            # un-captured warm-up: loads modules, sizes scratch
                engine.iteration()
            self.stream.synchronize()
            for t, sv in zip((engine.G.data, engine.D.data, engine.G.exp_avg, engine.G.exp_avg_sq,
                              engine.D.exp_avg, engine.D.exp_avg_sq, engine.rng.state), saved):
                t.copy_(sv)
            engine.iterations = it0
            self.stream.synchronize()
            g = torch.cuda.CUDAGraph()
            n0 = native.launch_count()
            with torch.cuda.graph(g, stream=self.stream):
                engine.iteration()
            self.launches_per_replay = native.launch_count() - n0
            engine.iterations = it0            # the capture pass executed nothing
            self.graph = g
        torch.cuda.current_stream(dev).wait_stream(self.stream)
        self.warmup_iterations = 0             # (restored: the warm-up leaves no trace in the training state)
        self.replays = 0

    def run(self, steps: int):
        cur = torch.cuda.current_stream(self.engine.device)
        self.stream.wait_stream(cur)
        with torch.cuda.stream(self.stream):
            for _ in range(int(steps)):
                self.graph.replay()
        cur.wait_stream(self.stream)
        self.replays += int(steps)
        self.engine.iterations += int(steps)
