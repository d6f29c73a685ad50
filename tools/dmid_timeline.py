"""clock64 timeline of CTA 0 of the fused discriminator-middle kernel (k_d_mid) + its latency inside a CUDA graph."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fed_tgan_b200.ops import native as N  # noqa: E402
from fed_tgan_b200.ops import reference as R  # noqa: E402

lib = N.lib
lib.ftg_d_mid_timeline.argtypes = [C.c_int, C.POINTER(C.c_longlong)]
dev = torch.device("cuda:0")
for M in (150, 50):
    H0 = H1 = 256
    a0, w1 = torch.randn(M, H0, device=dev), torch.randn(H1, H0, device=dev) * 0.1
    b1, wh, bh = torch.randn(H1, device=dev), torch.randn(H1, device=dev), torch.randn(1, device=dev)
    mask0 = torch.ones(M, H0, device=dev)
    rs = torch.randn(M, device=dev)
    a1, m1, z1, z0 = (torch.empty(M, H1, device=dev) for _ in range(4))
    y, loss = torch.zeros(M, device=dev), torch.zeros(1, device=dev)
    rng = R.RngState(1, dev)
    run = lambda: N.d_mid(a0, w1, b1, a1, m1, mask0, wh, bh, y, rs, loss, rs, z1, z0, rng=rng, stream=3)  # noqa: E731
    for _ in range(5):
        run()
    torch.cuda.synchronize()
    lib.ftg_d_mid_timeline(1, None)
    run()
    torch.cuda.synchronize()
    buf = (C.c_longlong * 8)()
    lib.ftg_d_mid_timeline(0, buf)
    t = list(buf)
    names = ["entry", "after_pdl_wait", "chunk0_ready", "chunk1_ready", "chunks_done", "stores_done", "chunk0_forward_done_thread0", "chunk0_forward_all_warps"]
    print(M, {n: t[i] - t[0] for i, n in enumerate(names)})
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
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
    print(M, "us per launch in graph", round(e0.elapsed_time(e1) * 1e3 / 200, 2))
