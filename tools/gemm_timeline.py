"""In-kernel timeline of the tcgen05 GEMM (clock64 at phase boundaries of CTA 0) + CUDA-event
latency per shape, to see where a latency-bound launch spends its microseconds."""
import ctypes as C
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fed_tgan_b200.ops import native  # noqa: E402

lib = native.lib
lib.ftg_gemm_tc_timeline.argtypes = [C.c_int, C.POINTER(C.c_longlong)]
dev = torch.device("cuda:0")
names = ["start", "prod_begin", "prod_end", "mma_first_full", "mma_commit", "epi_tmem_full", "epi_staged", "epi_begin",
         "epi_end", "exit", "cl_bar1", "cl_reduced", "cl_epi_done"]
shapes = [(500, 256, 252, 0, 0, None), (500, 256, 508, 0, 0, None), (150, 256, 2520, 0, 0, None),
          (150, 256, 2520, 0, 0, 1), (50, 2520, 256, 0, 1, None), (256, 2520, 150, 1, 1, None), (500, 128, 764, 0, 0, None),
          (8192, 256, 508, 0, 0, None), (500, 256, 252, 0, 0, 1), (500, 256, 508, 0, 0, 1)]
out = []
variants = [("tf32", True), ("tf32", False), ("fp32x3", True), ("fp32x3", False)] if "--all" in sys.argv else [("tf32", True)]
for prec, push in variants:
  native.set_precision(prec)
  native.set_cluster_push(push)
  for M, N, K, at, bt, sk in shapes:
      a = torch.randn((K, M) if at else (M, K), device=dev)
      b = torch.randn((K, N) if bt else (N, K), device=dev)
      d = torch.empty(M, N, device=dev)
      bias = torch.randn(N, device=dev)
      run = lambda: native.linear(a, b, d, a_t=bool(at), b_t=bool(bt), bias=bias, split_k=sk)  # noqa: E731
      for _ in range(5):
          run()
      torch.cuda.synchronize()
      lib.ftg_gemm_tc_timeline(1, None)
      run()
      torch.cuda.synchronize()
      buf = (C.c_longlong * 16)()
      lib.ftg_gemm_tc_timeline(0, buf)
      t = list(buf)[:13]
      rel = {n: t[i] - t[0] for i, n in enumerate(names)}
      # graph-replayed latency (no host overhead)
      g = torch.cuda.CUDAGraph()
      s = torch.cuda.Stream()
      s.wait_stream(torch.cuda.current_stream())
      with torch.cuda.stream(s):
          native.preallocate_scratch(dev)
          run()
          s.synchronize()
          with torch.cuda.graph(g, stream=s):
              for _ in range(20):
                  run()
      torch.cuda.synchronize()
      e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
      e0.record()
      for _ in range(10):
          g.replay()
      e1.record()
      torch.cuda.synchronize()
      us = e0.elapsed_time(e1) * 1e3 / 200
      split = sk if sk is not None else native._choose_cluster_split(M, N, K)
      row = {"precision": prec, "cluster_push": push, "M": M, "N": N, "K": K, "a_t": at, "b_t": bt, "split": split, "us_per_launch_in_graph": round(us, 2), "cycles": rel}
      out.append(row)
      print(json.dumps(row))
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "gemm_timeline.json"), "w"), indent=1)
