"""tfd_gemm_bf16_splitk_noalloc: the split-K GEMM entry point that allocates nothing.

Some users of this library cannot call cudaMalloc at the moment they need a GEMM: a host application whose own
memory pool already owns all of the GPU's memory (or that must stay inside a fixed budget) has nothing left to
allocate from, and a latency-critical loop cannot afford the device-wide synchronisation cudaMalloc/cudaFree
imply.  For them the partial-sum planes of the split-K GEMM live in a fixed 8 MiB array inside the library.

This test drives the entry point through the plain C ABI (ctypes, raw device pointers) exactly as such a caller
would, twice in a row on the same shared scratch, and checks each result against the ordinary entry point
(tfd_gemm_bf16, one split, caller-owned output) within the tolerance the other GEMM tests use.  The file name
keeps it last in collection order, so a run that stops at the first failure has already run every other test.
"""
import ctypes
from ctypes import c_float, c_int, c_longlong, c_void_p

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 2e-3  # same bound as tests/test_gpu_kernels.py uses for the fp32-output GEMMs


def _rel(a, b):
    return (a.float() - b.float()).abs().max().item() / (b.float().abs().max().item() + 1e-20)


def _lib():
    from tensorflow_distributed_b200.ops import native
    lib = native.lib()
    native.set_device(torch.cuda.current_device())
    lib.tfd_gemm_bf16_splitk_noalloc.restype = c_int
    lib.tfd_gemm_bf16.restype = c_int
    return lib


def _stream():
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def _noalloc(lib, A, W, M, N, K, splits, out, bias=None, relu=0, bn=64):
    return lib.tfd_gemm_bf16_splitk_noalloc(
        c_void_p(A.data_ptr()), c_longlong(A.stride(0)), c_int(0), c_void_p(W.data_ptr()), c_longlong(W.stride(0)),
        c_int(1), c_int(M), c_int(N), c_int(K), c_int(bn), c_int(splits), c_void_p(out.data_ptr()),
        c_longlong(out.stride(0)), c_void_p(bias.data_ptr() if bias is not None else 0), c_int(relu), _stream())


def _existing(lib, A, W, M, N, K, out, bn=64):
    # the existing entry point: one split, TFD_EPI_STORE_F32 (mode 0) into caller-owned memory
    return lib.tfd_gemm_bf16(
        c_void_p(A.data_ptr()), c_longlong(A.stride(0)), c_int(0), c_void_p(W.data_ptr()), c_longlong(W.stride(0)),
        c_int(1), c_int(M), c_int(N), c_int(K), c_int(bn), c_int(1), c_int(0), c_float(1.0), c_void_p(out.data_ptr()),
        c_longlong(out.stride(0)), c_void_p(0), c_void_p(0), c_longlong(0), c_longlong(0), c_void_p(0),
        ctypes.c_uint(0), _stream())


def test_noalloc_splitk_matches_existing_entry_point_twice_in_a_row():
    assert torch.cuda.is_available()
    lib = _lib()
    dev = "cuda"
    torch.manual_seed(21)
    # call 1: the fc1 forward shape of the CNN (8 splits -> 1 Mi floats of scratch)
    M, N, K = 128, 1024, 3136
    A = torch.randn(M, K, device=dev).to(torch.bfloat16)
    W = torch.randn(K, N, device=dev).to(torch.bfloat16)
    # every buffer the calls touch exists before the first call: the entry point itself must not allocate
    out1, ref1 = torch.zeros(M, N, device=dev), torch.zeros(M, N, device=dev)
    M2, N2, K2 = 100, 72, 520
    A2 = torch.randn(M2, K2, device=dev).to(torch.bfloat16)
    W2 = torch.randn(K2, N2, device=dev).to(torch.bfloat16)
    bias2 = torch.randn(N2, device=dev)
    out2, ref2 = torch.zeros(M2, N2, device=dev), torch.zeros(M2, N2, device=dev)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()

    rc = _noalloc(lib, A, W, M, N, K, 8, out1)
    assert rc == 0, lib.tfd_last_error().decode()
    # call 2 straight after, on the same shared scratch (which still holds call 1's partial sums): smaller, ragged
    # shape, fewer splits, bias + ReLU
    rc = _noalloc(lib, A2, W2, M2, N2, K2, 3, out2, bias=bias2, relu=1)
    assert rc == 0, lib.tfd_last_error().decode()
    assert _existing(lib, A, W, M, N, K, ref1) == 0, lib.tfd_last_error().decode()
    assert _existing(lib, A2, W2, M2, N2, K2, ref2) == 0, lib.tfd_last_error().decode()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before  # nothing went through torch's allocator either

    assert _rel(out1, ref1) < TOL
    assert _rel(out2, torch.relu(ref2 + bias2)) < TOL
    # and both agree with a plain fp32 PyTorch matmul of the same bf16 inputs
    assert _rel(out1, A.float() @ W.float()) < TOL


def test_noalloc_rejects_requests_beyond_the_builtin_scratch():
    lib = _lib()
    dev = "cuda"
    M, N, K = 1024, 1024, 1024  # 3 splits * 1 Mi floats > 2 Mi floats of scratch
    A = torch.zeros(M, K, device=dev, dtype=torch.bfloat16)
    W = torch.zeros(K, N, device=dev, dtype=torch.bfloat16)
    out = torch.full((M, N), 5.0, device=dev)
    rc = _noalloc(lib, A, W, M, N, K, 3, out, bn=128)
    assert rc != 0
    msg = lib.tfd_last_error().decode()
    assert "exceeds the built-in scratch" in msg and "8 MiB" in msg
    torch.cuda.synchronize()
    assert torch.all(out == 5.0)  # rejected before anything ran


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs 2 GPUs in one process")
def test_noalloc_first_call_on_a_second_device_of_the_same_process():
    """The header promises per-device set-up on the first call on EACH device (scratch address and the kernel's
    shared-memory opt-in are both per device): drive device 0, then device 1, from one process."""
    lib = _lib()
    from tensorflow_distributed_b200.ops import native
    M, N, K = 128, 256, 640
    for d in (0, 1):
        torch.cuda.set_device(d)
        native.set_device(d)
        dev = "cuda:%d" % d
        torch.manual_seed(3 + d)
        A = torch.randn(M, K, device=dev).to(torch.bfloat16)
        W = torch.randn(K, N, device=dev).to(torch.bfloat16)
        out = torch.zeros(M, N, device=dev)
        assert _noalloc(lib, A, W, M, N, K, 4, out) == 0, lib.tfd_last_error().decode()
        torch.cuda.synchronize(d)
        assert _rel(out, A.float() @ W.float()) < TOL
    torch.cuda.set_device(0)
    native.set_device(0)
