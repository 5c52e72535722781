"""GPU test of the no-allocation entry point fsb_rows_gemm_static (include/fish_b200.h): the same Linear as fsb_rows_gemm
with the stream-K partial tiles in a file-scope device array of the library instead of caller scratch.

Checked through the C ABI against the existing entry point on two calls in a row (the second on different inputs and a
different shape, so stale contents of the shared array from the first call would show), plus the size limit.
Collected last (file name) so that a failure here cannot hide other tests under `pytest -x`; passes alone in a fresh
process and after the other tests (it keeps no state besides the library's array, which every call overwrites before
reading)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    from fish_speech_b200 import _lib
    _lib.require_cuda()
    return _lib.load()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _existing(lib, W, x, res, epi):
    M, K = x.shape
    out = torch.empty((M, W.shape[0]), dtype=torch.bfloat16, device="cuda")
    nbytes = lib.fsb_rows_gemm_scratch_bytes()
    scratch = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
    rc = lib.fsb_rows_gemm(W.data_ptr(), None, x.data_ptr(), None if res is None else res.data_ptr(), out.data_ptr(), M, W.shape[0],
                           K, epi, scratch.data_ptr(), nbytes, _stream())
    assert rc == 0, rc
    torch.cuda.synchronize()
    return out


def _static(lib, W, x, res, epi, out):
    """`out` is allocated by the caller BEFORE the call: the entry point itself allocates nothing."""
    M, K = x.shape
    err = C.create_string_buffer(512)
    rc = lib.fsb_rows_gemm_static(W.data_ptr(), x.data_ptr(), None if res is None else res.data_ptr(), out.data_ptr(), M,
                                  W.shape[0], K, epi, _stream(), err, len(err))
    return rc, err.value.decode()


def _case(seed, M, N, K, epi):
    g = torch.Generator().manual_seed(seed)
    W = (torch.randn(N, K, generator=g) * 0.02).bfloat16().cuda()
    x = torch.randn(M, K, generator=g).bfloat16().cuda()
    res = torch.randn(M, N, generator=g).bfloat16().cuda() if epi == 1 else None
    return W, x, res


def test_static_scratch_matches_existing_entry_point_twice_in_a_row(lib):
    import gc
    gc.collect()  # models of earlier tests go now, not in the middle of a call
    # (M, N, K, epi): wo with residual at the S2 width (every tile split 7 ways -> the finish pass reads the static array),
    # then wqkv with a plain store and a ragged row count (3-way split), then the first shape again with new data
    calls = [(1, 12, 2560, 4096, 1), (2, 31, 6144, 2560, 0), (3, 12, 2560, 4096, 1), (4, 32, 1000, 512, 0)]
    for seed, M, N, K, epi in calls:
        W, x, res = _case(seed, M, N, K, epi)
        ref = _existing(lib, W, x, res, epi)
        out = torch.full((M, N), 7.0, dtype=torch.bfloat16, device="cuda")
        # number of allocation REQUESTS torch's caching allocator has served (monotonic: frees by the garbage collector of
        # objects left over from earlier tests do not move it, unlike memory_allocated())
        before = torch.cuda.memory_stats().get("allocation.all.allocated", 0)
        rc, msg = _static(lib, W, x, res, epi, out)
        torch.cuda.synchronize()
        assert rc == 0, msg
        assert torch.cuda.memory_stats().get("allocation.all.allocated", 0) == before  # nothing allocated through torch
        # same kernels, same split, same summation order: bit-identical; the stated tolerance of the rows_gemm tests
        # (one bf16 ulp of the value + 2e-3) is the fallback bound that is asserted
        err = (out.float() - ref.float()).abs()
        tol = ref.float().abs() * 2.0 ** -7 + 2e-3
        print(f"[parity] rows_gemm_static M={M} N={N} K={K} epi={epi}: max-abs {float(err.max()):.5f}, "
              f"bit-identical {bool(torch.equal(out, ref))}")
        assert int((err > tol).sum()) == 0, float(err.max())
        # and against an fp32 reference of the op, like tests/test_gpu_kernels.py
        acc = x.float() @ W.float().t()
        want = acc.bfloat16() if epi == 0 else (res.float() + acc.bfloat16().float()).bfloat16()
        mag = want.float().abs() + (acc.abs() + res.float().abs() if epi == 1 else 0)
        assert int(((out.float() - want.float()).abs() > mag * 2.0 ** -7 + 2e-3).sum()) == 0


def test_static_scratch_rejects_sizes_beyond_its_limit(lib):
    W, x, res = _case(9, 33, 512, 512, 0)
    out = torch.empty((33, 512), dtype=torch.bfloat16, device="cuda")
    rc, msg = _static(lib, W, x, None, 0, out)
    assert rc != 0 and "32 rows" in msg and "fsb_rows_gemm" in msg, (rc, msg)
    W, x, _ = _case(10, 8, 512, 512, 0)
    rc, msg = _static(lib, W, x, None, 2, torch.empty((8, 512), dtype=torch.bfloat16, device="cuda"))
    assert rc != 0 and "SwiGLU" in msg, (rc, msg)
    rc, msg = _static(lib, W, x[:, :500].contiguous(), None, 0, torch.empty((8, 512), dtype=torch.bfloat16, device="cuda"))
    assert rc != 0 and "multiple of 64" in msg, (rc, msg)
