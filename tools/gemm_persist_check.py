"""Experimental persistent tcgen05 GEMM: numerics vs torch and time vs the default launch policy."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fed_tgan_b200.ops import native  # noqa: E402

dev = torch.device("cuda:0")
SHAPES = [(1024, 31480, 152, True, True), (52, 31480, 1024, False, True), (1624, 3700, 500, True, True),
          (8192, 256, 252, False, False), (1024, 2676, 500, True, True)]
rows = []
for (M, N, K, a_t, b_t) in SHAPES:
    A = torch.randn((K, M) if a_t else (M, K), device=dev)
    B = torch.randn((K, N) if b_t else (N, K), device=dev)
    bias = torch.randn(N, device=dev)
    ref = (A.t() if a_t else A) @ (B if b_t else B.t()) + bias
    rec = {"M": M, "N": N, "K": K, "a_t": a_t, "b_t": b_t}
    for mode in (0, 1):
        native.set_gemm_persistent(mode)
        out = torch.full((M, N), float("nan"), device=dev)
        native.linear(A, B, out, a_t=a_t, b_t=b_t, bias=bias)
        torch.cuda.synchronize()
        err = ((out - ref).abs().max() / ref.abs().max()).item()
        ts = []
        for r in range(12):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); native.linear(A, B, out, a_t=a_t, b_t=b_t, bias=bias); e.record(); e.synchronize()
            ts.append(s.elapsed_time(e) * 1e3)
        ts.sort()
        rec["persistent" if mode else "default"] = {"rel_err": err, "median_us": round(ts[6], 1), "min_us": round(ts[0], 1)}
    native.set_gemm_persistent(0)
    rows.append(rec)
    print(json.dumps(rec), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "gemm_persist_check.json"), "w"), indent=1)
