"""The allocation-free split-K GEMM entry point, driven through the C ABI exactly as a non-Python caller would.

Some users of this library cannot allocate device memory at the moment they need a GEMM: a training or serving
loop that must never stall the GPU (cudaMalloc / cudaFree synchronise the device), or a C/C++ plug-in running
inside a process whose memory pool has already claimed the GPU and cannot be reached without libtorch.  For them
`dfb_gemm_bf16_splitk_noalloc` keeps the split-K workspace in a 32 MiB array inside the library.  This test calls
it twice in a row with plain ctypes (no torch allocation between the two calls), a large problem first and a
smaller one second so the second call finds the first one's leftovers in the shared array, and compares each
result with the ordinary entry point `dfb_gemm_bf16` fed a caller-allocated workspace.  It also checks that an
oversized request is refused with a readable message.  The file name sorts last on purpose: if this test fails
under `pytest -x`, every other test has already run.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_BYTES = 32 << 20


def _lib():
    from distriflow_b200.ops import _lib as L
    from distriflow_b200.ops.dense import _init_sigs
    _init_sigs()
    lib = L.lib()
    c_int, c_void_p = ctypes.c_int, ctypes.c_void_p
    lib.dfb_gemm_bf16_splitk_noalloc.restype = c_int
    lib.dfb_gemm_bf16_splitk_noalloc.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p] + [c_int] * 14 + [c_void_p]
    return lib


def _ref_call(lib, A, B, D, M, N, K, splits, out_dtype):
    nbytes = int(lib.dfb_gemm_workspace_bytes(M, N, splits))
    ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=A.device)
    rc = lib.dfb_gemm_bf16(A.data_ptr(), B.data_ptr(), D.data_ptr(), None, 0, M, N, K, A.stride(0), B.stride(0), D.stride(0),
                           1, 1, out_dtype, 0, 128, splits, ws.data_ptr(), nbytes, 0,
                           ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, lib.dfb_last_error()
    return ws


def test_noalloc_splitk_matches_regular_entry_point_twice():
    lib = _lib()
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    # wgrad-shaped problems (both operands MN-major, long K): first a bigger one, then a smaller one
    cases = [(768, 768, 8192, 8, 1), (256, 384, 4096, 4, 0)]         # M, N, K, splits, out dtype (1 = bf16, 0 = fp32)
    inputs, want, got = [], [], []
    for M, N, K, splits, odt in cases:                                 # everything is allocated up front ...
        A = torch.randn(K, M, device=dev).bfloat16()
        B = torch.randn(K, N, device=dev).bfloat16()
        tdt = torch.bfloat16 if odt == 1 else torch.float32
        inputs.append((A, B))
        want.append(torch.empty(M, N, device=dev, dtype=tdt))
        got.append(torch.full((M, N), float("nan"), device=dev, dtype=tdt))
    keep = [_ref_call(lib, A, B, w, M, N, K, s, odt) for (A, B), w, (M, N, K, s, odt) in zip(inputs, want, cases)]
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    for (A, B), g, (M, N, K, splits, odt) in zip(inputs, got, cases):  # ... so these two calls run back to back
        rc = lib.dfb_gemm_bf16_splitk_noalloc(A.data_ptr(), B.data_ptr(), g.data_ptr(), None, 0, M, N, K, A.stride(0),
                                              B.stride(0), g.stride(0), 1, 1, odt, 0, 128, splits, 0, stream)
        assert rc == 0, lib.dfb_last_error()
    assert torch.cuda.memory_allocated() == before                     # nothing was taken from torch's pool either
    torch.cuda.synchronize()
    for (A, B), w, g, (M, N, K, splits, odt) in zip(inputs, want, got, cases):
        assert not torch.isnan(g.float()).any()
        ref32 = A.float().t() @ B.float()
        scale = float(ref32.abs().max())
        assert float((g.float() - w.float()).abs().max()) <= 1e-2 * scale       # same tolerance as test_gpu_kernels.py
        assert float((g.float() - ref32).abs().max()) <= 1e-2 * scale
    del keep


def test_noalloc_rejects_oversized_request():
    lib = _lib()
    dev = torch.device("cuda", 0)
    M, N, K, splits = 4096, 4096, 4096, 8
    assert int(lib.dfb_gemm_workspace_bytes(M, N, splits)) > MAX_BYTES
    A = torch.zeros(K, M, device=dev, dtype=torch.bfloat16)
    B = torch.zeros(K, N, device=dev, dtype=torch.bfloat16)
    D = torch.zeros(M, N, device=dev, dtype=torch.bfloat16)
    rc = lib.dfb_gemm_bf16_splitk_noalloc(A.data_ptr(), B.data_ptr(), D.data_ptr(), None, 0, M, N, K, M, N, N, 1, 1, 1, 0, 128,
                                          splits, 0, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    msg = lib.dfb_last_error().decode()
    assert rc != 0 and "scratch" in msg and str(MAX_BYTES) in msg, msg
