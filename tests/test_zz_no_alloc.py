"""ftg_gemm_tc_noalloc -- the GEMM entry point for callers that cannot allocate.

Some users of libfedtgan_b200 (a plugin inside a host that owns the whole device heap, an
out-of-memory fail-safe path, a loop that cannot take cudaMalloc's synchronisation) cannot
provide the split-K workspace that ``ftg_gemm_tc`` asks for.  ``ftg_gemm_tc_noalloc`` takes it
from a fixed device array inside the library.  This test drives both entry points through the
plain C ABI (ctypes, raw device pointers) and checks that the no-allocation variant returns what
the caller-workspace variant returns, on two calls in a row (the second call re-uses the
scratch the first one left behind), and that an over-sized request is refused with a message.

The file name sorts last on purpose: if this test fails under ``-x`` every other test has
already run.
"""

import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from fed_tgan_b200.ops import native

    lib = native.lib
    lib.ftg_gemm_tc_noalloc.argtypes = [C.POINTER(native.GemmArgs), C.c_int, C.c_void_p]
    lib.ftg_gemm_tc_noalloc.restype = C.c_int
    lib.ftg_gemm_tc.argtypes = [C.POINTER(native.GemmArgs), C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.ftg_gemm_tc.restype = C.c_int
    lib.ftg_gemm_tc_workspace_bytes.argtypes = [C.c_int, C.c_int, C.c_int]
    lib.ftg_gemm_tc_workspace_bytes.restype = C.c_longlong
    lib.ftg_gemm_tc_noalloc_max_workspace_bytes.restype = C.c_longlong
    lib.ftg_last_error_message.restype = C.c_char_p
    return lib, native.GemmArgs


def _args(GemmArgs, a, b, d, bias):
    g = GemmArgs()
    g.a, g.b, g.d = a.data_ptr(), b.data_ptr(), d.data_ptr()
    g.M, g.N, g.K = d.shape[0], d.shape[1], a.shape[1]
    g.lda, g.ldb, g.ldd = a.stride(0), b.stride(0), d.stride(0)
    g.a_t = g.b_t = 0
    g.bias = bias.data_ptr()
    g.act = 0
    g.alpha, g.slope, g.drop_p = 1.0, 0.2, 0.5
    g.accumulate = 0
    return g


def test_noalloc_entry_matches_workspace_entry():
    lib, GemmArgs = _lib()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream().cuda_stream
    M, N, K, split = 150, 256, 2520, 8
    gen = torch.Generator().manual_seed(0)
    for call in range(2):                          # two calls in a row share the built-in scratch
        a = torch.randn(M, K, generator=gen).to(dev)
        b = torch.randn(N, K, generator=gen).to(dev)
        bias = torch.randn(N, generator=gen).to(dev)
        want = torch.empty(M, N, device=dev)
        got = torch.full((M, N), float("nan"), device=dev)
        ws = torch.empty(lib.ftg_gemm_tc_workspace_bytes(M, N, split) // 4, device=dev)
        ctr = torch.zeros(64, dtype=torch.int32, device=dev)
        rc = lib.ftg_gemm_tc(C.byref(_args(GemmArgs, a, b, want, bias)), split, ws.data_ptr(), ctr.data_ptr(), stream)
        assert rc == 0
        rc = lib.ftg_gemm_tc_noalloc(C.byref(_args(GemmArgs, a, b, got, bias)), split, stream)
        assert rc == 0, lib.ftg_last_error_message()
        torch.cuda.synchronize()
        # same kernel, same fixed-order split-K reduction: the two entry points agree exactly;
        # the tolerance below is the tf32 tolerance the other GEMM tests use against fp32
        assert torch.equal(got, want), (call, (got - want).abs().max().item())
        ref = a @ b.t() + bias
        err = (got - ref).abs().max().item()
        assert err <= 4e-3 * (K ** 0.5) / 8 + 4e-3 * ref.abs().max().item(), (call, err)


def test_noalloc_rejects_oversized_request():
    lib, GemmArgs = _lib()
    dev = torch.device("cuda:0")
    stream = torch.cuda.current_stream().cuda_stream
    M, N, K, split = 2048, 2048, 4096, 32           # 512 tiles * 32 splits * 32 KiB = 512 MiB of partials
    assert lib.ftg_gemm_tc_workspace_bytes(M, N, split) > lib.ftg_gemm_tc_noalloc_max_workspace_bytes()
    a = torch.zeros(M, K, device=dev)
    b = torch.zeros(N, K, device=dev)
    d = torch.zeros(M, N, device=dev)
    bias = torch.zeros(N, device=dev)
    rc = lib.ftg_gemm_tc_noalloc(C.byref(_args(GemmArgs, a, b, d, bias)), split, stream)
    assert rc == 1001
    msg = lib.ftg_last_error_message().decode()
    assert "split_k" in msg and "scratch" in msg
    # the same problem without split-K needs no scratch and is accepted
    assert lib.ftg_gemm_tc_noalloc(C.byref(_args(GemmArgs, a, b, d, bias)), 1, stream) == 0
    torch.cuda.synchronize()
