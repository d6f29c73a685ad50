"""Ring-depth sweep of the tcgen05 GEMM over the shapes of the Adult and 1024-wide CTGAN iterations.

For every (shape, stages) pair: median / min device time over REPS launches (CUDA events, operands
rotated through a pool larger than L2 so every launch streams from HBM).  Writes
gpurun_out/gemm_stage_sweep.json; the launch policy in csrc/gemm_tc.cu::choose_stages is read off it.
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fed_tgan_b200.ops import native  # noqa: E402

dev = torch.device("cuda:0")
REPS = 24
# (M, N, K, a_t, b_t): logical out[M,N] = A[M,K] . B[N,K]^T; dimensions padded to multiples of 4 floats as in
# the arenas (an unaligned leading dimension takes the SIMT fallback, which is not what this tool measures)
SHAPES = [
    (1024, 31480, 150, True, True), (50, 31480, 1024, False, True), (1624, 3700, 500, True, True),
    (1024, 1652, 500, True, True), (1024, 2676, 500, True, True), (500, 2048, 1624, False, True),
    (50, 1024, 31480, False, False), (150, 1024, 31480, False, False), (500, 1624, 3700, False, False),
    (500, 1024, 2676, False, False), (500, 1024, 1024, False, True),
    # Adult iteration / generation shapes
    (500, 256, 252, False, False), (150, 256, 2500, False, False), (256, 2500, 150, True, True),
    (8192, 256, 252, False, False), (8192, 128, 764, False, False),
]


def operands(M, N, K, a_t, b_t, copies):
    A = [torch.randn((K, M) if a_t else (M, K), device=dev) for _ in range(copies)]
    B = [torch.randn((K, N) if b_t else (N, K), device=dev) for _ in range(copies)]
    return A, B


rows = []
for (M, N, K, a_t, b_t) in SHAPES:
    bytes_once = 4 * (M * K + N * K + M * N)
    copies = max(2, min(8, int(300e6 // bytes_once) + 1))
    A, B = operands(M, N, K, a_t, b_t, copies)
    out = torch.empty(M, N, device=dev)
    ref = (A[0].t() if a_t else A[0]) @ (B[0] if b_t else B[0].t())
    rec = {"M": M, "N": N, "K": K, "a_t": a_t, "b_t": b_t, "split": native._choose_cluster_split(M, N, K),
           "tiles": ((M + 127) // 128) * ((N + 63) // 64), "hbm_bytes": bytes_once, "us": {}}
    for st in (0, 2, 3, 4, 6):
        native.set_gemm_stages(st)
        native.linear(A[0], B[0], out, a_t=a_t, b_t=b_t)
        torch.cuda.synchronize()
        err = ((out - ref).abs().max() / ref.abs().max()).item()
        assert err < 2e-2, (M, N, K, st, err)
        ts = []
        for r in range(REPS):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            native.linear(A[r % copies], B[r % copies], out, a_t=a_t, b_t=b_t)
            e.record()
            e.synchronize()
            ts.append(s.elapsed_time(e) * 1e3)
        ts.sort()
        rec["us"]["auto" if st == 0 else str(st)] = {"median": round(ts[len(ts) // 2], 2), "min": round(ts[0], 2)}
    native.set_gemm_stages(0)
    best = min((k for k in rec["us"] if k != "auto"), key=lambda k: rec["us"][k]["median"])
    rec["best"] = best
    rec["GBps_best"] = round(bytes_once / rec["us"][best]["median"] / 1e3, 1)
    rec["TFLOPs_best"] = round(2.0 * M * N * K / rec["us"][best]["median"] / 1e6, 1)
    rows.append(rec)
    print(f"M{M:5d} N{N:6d} K{K:6d} {'T' if a_t else 'N'}{'T' if b_t else 'N'} tiles {rec['tiles']:5d} split {rec['split']} | "
          + "  ".join(f"{k}:{v['median']:7.1f}" for k, v in rec["us"].items())
          + f" | best {best}  {rec['GBps_best']} GB/s  {rec['TFLOPs_best']} TFLOP/s", flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "gemm_stage_sweep.json"), "w"), indent=1)

# ---- split-K sweep for the skinny weight-streaming shapes: cluster (DSMEM) splits vs workspace splits that
# put two CTAs on every SM.  Leading dimensions are multiples of 4 floats (tcgen05 path, never the SIMT fallback).
SKINNY = [(50, 1024, 31480), (150, 1024, 31480), (50, 256, 2500), (150, 256, 2500)]
srows = []
for (M, N, K) in SKINNY:
    A = [torch.randn(M, K, device=dev) for _ in range(3)]
    B = [torch.randn(N, K, device=dev) for _ in range(3)]
    out = torch.empty(M, N, device=dev)
    ref = A[0] @ B[0].t()
    rec = {"M": M, "N": N, "K": K, "us": {}}
    for label, kw in [("cluster", {})] + [(f"ws{sk}", {"split_k": sk}) for sk in (4, 8, 16, 24, 32)]:
        native.linear(A[0], B[0], out, **kw)
        torch.cuda.synchronize()
        err = ((out - ref).abs().max() / ref.abs().max()).item()
        assert err < 2e-2, (M, N, K, label, err)
        ts = []
        for r in range(REPS):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            native.linear(A[r % 3], B[r % 3], out, **kw)
            e.record()
            e.synchronize()
            ts.append(s.elapsed_time(e) * 1e3)
        ts.sort()
        rec["us"][label] = {"median": round(ts[len(ts) // 2], 2), "min": round(ts[0], 2)}
    srows.append(rec)
    print(f"M{M:5d} N{N:6d} K{K:6d} NN | " + "  ".join(f"{k}:{v['median']:7.1f}" for k, v in rec["us"].items()), flush=True)
json.dump(srows, open(os.path.join(ROOT, "gpurun_out", "gemm_splitk_sweep.json"), "w"), indent=1)
