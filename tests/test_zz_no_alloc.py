"""The allocation-free Krum Gram entry point, called through the C ABI.

`b2_krum_gram_static` exists for hosts that cannot allocate device memory at the point where they need the Krum
scores (the GPU's memory belongs to another runtime's allocator, or a device-synchronising cudaMalloc next to live
training streams is not acceptable): its split-K scratch is a fixed array inside the library.  This test drives it
the way such a host would - raw pointers through ctypes, no workspace - and checks that it returns what the
workspace-taking `b2_krum_gram` returns, twice in a row (the second call meets the scratch as the first one left it).
The file sorts last on purpose: a run that stops at the first failure has then already run every other test.
"""

import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from myfyp_b200.ops import native

    lib = native.load()
    lib.b2_krum_gram_static_max_n.restype = ctypes.c_int64
    lib.b2_krum_gram_workspace_bytes.restype = ctypes.c_int64
    lib.b2_last_error.restype = ctypes.c_char_p
    return lib


def _ptrs(tensors):
    arr = (ctypes.c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = t.data_ptr()
    return arr


@pytest.mark.parametrize("k,n", [(5, 235264), (16, 1 << 20)])
def test_static_scratch_matches_workspace_entry_point_twice(k, n):
    lib = _lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    g = torch.Generator().manual_seed(k)
    base = torch.randn(n, generator=g)
    slabs = [(base + 0.01 * torch.randn(n, generator=g)).to(dev) for _ in range(k)]
    ws = torch.zeros(int(lib.b2_krum_gram_workspace_bytes()) // 4, dtype=torch.float32, device=dev)
    ref = torch.empty(64 * 64, dtype=torch.float32, device=dev)
    rc = lib.b2_krum_gram(ctypes.c_void_p(ref.data_ptr()), _ptrs(slabs), k, ctypes.c_int64(n), ctypes.c_void_p(ws.data_ptr()), 3, stream)
    assert rc == 0, lib.b2_last_error()
    for call in range(2):
        out = torch.full((64 * 64,), float("nan"), dtype=torch.float32, device=dev)
        rc = lib.b2_krum_gram_static(ctypes.c_void_p(out.data_ptr()), _ptrs(slabs), k, ctypes.c_int64(n), 3, stream)
        assert rc == 0, lib.b2_last_error()
        torch.cuda.synchronize(dev)
        # same kernel, fewer CTAs => the partial sums are grouped differently: equal up to fp32 summation order
        scale = ref.abs().max().item()
        torch.testing.assert_close(out, ref, rtol=1e-4, atol=1e-4 * scale, msg=f"call {call}")
    # and against an fp64 oracle (looser: with at most 64 CTAs each tensor-core accumulation chain is longer, and
    # fp32 accumulation in the MMA pipeline drifts by a few 1e-4 relative over thousands of steps)
    m = torch.stack([s.double().cpu() for s in slabs])
    m = m - m[0:1]
    oracle = m @ m.T
    got = out.view(64, 64)[:k, :k].double().cpu()
    torch.testing.assert_close(got, oracle, rtol=2e-3, atol=2e-3 * oracle.abs().max().item())


def test_static_scratch_rejects_oversize_requests():
    lib = _lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    slab = torch.zeros(64, device=dev)
    out = torch.zeros(64 * 64, device=dev)
    too_big = int(lib.b2_krum_gram_static_max_n()) + 4
    rc = lib.b2_krum_gram_static(ctypes.c_void_p(out.data_ptr()), _ptrs([slab, slab]), 2, ctypes.c_int64(too_big), 3, stream)
    assert rc != 0 and b"exceeds" in lib.b2_last_error()
    rc = lib.b2_krum_gram_static(ctypes.c_void_p(out.data_ptr()), _ptrs([slab] * 65), 65, ctypes.c_int64(64), 3, stream)
    assert rc != 0 and b"outside" in lib.b2_last_error()
