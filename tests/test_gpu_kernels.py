"""GPU parity tests for the stage entry points of the C ABI (fsb_gemv, fsb_sample) against the oracle."""
import pytest
import torch
import torch.nn.functional as F

from oracle import dual_ar_ref as R
from tests.util import load_golden

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    from fish_speech_b200 import _lib
    _lib.require_cuda()
    return _lib.load()


def _gemv(lib, W, x, W2=None, norm_w=None, res=None, pro=0, epi=0, eps=1e-6):
    M, K = x.shape
    N = W.shape[0]
    out = torch.empty((M, N), dtype=torch.bfloat16, device="cuda")
    p = lambda t: None if t is None else t.data_ptr()
    rc = lib.fsb_gemv(p(W), p(W2), p(x), p(norm_w), p(res), p(out), M, N, K, pro, epi, eps,
                      torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    return out


# shapes: (N, K) of every Linear on the path at tiny and S2-Pro sizes (+ an odd N for the row partition)
SHAPES = [(768, 512), (512, 512), (1536, 512), (512, 1536), (6144, 2560), (2560, 4096), (9728, 2560), (2560, 9728),
          (4096, 2560), (1000, 512)]


@pytest.mark.parametrize("N,K", SHAPES)
@pytest.mark.parametrize("M", [1, 2, 3, 8, 11])
def test_gemv_store_and_residual(lib, N, K, M):
    g = torch.Generator().manual_seed(N * 7 + K + M)
    W = (torch.randn(N, K, generator=g) * 0.02).bfloat16()
    x = torch.randn(M, K, generator=g).bfloat16()
    res = torch.randn(M, N, generator=g).bfloat16()
    ref = F.linear(x.float(), W.float())                      # fp32 reference of the same op
    out = _gemv(lib, W.cuda(), x.cuda()).cpu()
    # bf16 output: tolerance = 1 bf16 ulp of the value (2^-8 relative) + accumulate-order noise
    tol = ref.abs() * 2 ** -7 + 1e-3
    assert bool(((out.float() - ref).abs() <= tol).all())
    out2 = _gemv(lib, W.cuda(), x.cuda(), res=res.cuda(), epi=1).cpu()
    ref2 = (ref.bfloat16() + res).float()
    assert bool(((out2.float() - ref2).abs() <= ref2.abs() * 2 ** -6 + 2e-2).all())
    # rows are independent of the batch they travel in (M buckets share one accumulation order)
    out1 = _gemv(lib, W.cuda(), x[:1].cuda()).cpu()
    if K <= 4096:  # same kernel family for every M: bit-identical; K=9728 with M=8 uses the direct-load kernel
        assert torch.equal(out1[0], out[0])
    else:
        assert bool(((out1[0].float() - out[0].float()).abs() <= out[0].float().abs() * 2 ** -7 + 1e-3).all())


@pytest.mark.parametrize("N,K", [(1536, 512), (9728, 2560)])
@pytest.mark.parametrize("M", [1, 4, 8])
def test_gemv_rmsnorm_swiglu_vs_oracle(lib, N, K, M):
    g = torch.Generator().manual_seed(N + K + M)
    w1 = (torch.randn(N, K, generator=g) * 0.02).bfloat16()
    w3 = (torch.randn(N, K, generator=g) * 0.02).bfloat16()
    nw = (1 + 0.1 * torch.randn(K, generator=g)).bfloat16()
    x = torch.randn(M, K, generator=g).bfloat16()
    n = R.rms_norm_double_round(x, nw, 1e-6)                   # llama.py:995-1000
    ref = F.silu(F.linear(n, w1)) * F.linear(n, w3)            # llama.py:986, bf16 op by op on CPU
    out = _gemv(lib, w1.cuda(), x.cuda(), W2=w3.cuda(), norm_w=nw.cuda(), pro=1, epi=2).cpu()
    d = (out.float() - ref.float()).abs()
    assert float(d.max()) <= float(ref.float().abs().max()) * 2 ** -6 + 1e-3
    assert float((d > 0).float().mean()) < 0.2  # most outputs are bit-identical to the CPU reference
    # norm + store (wqkv / LM-head pattern)
    ref2 = F.linear(n, w1)
    out2 = _gemv(lib, w1.cuda(), x.cuda(), norm_w=nw.cuda(), pro=1, epi=0).cpu()
    assert float((out2.float() - ref2.float()).abs().max()) <= float(ref2.float().abs().max()) * 2 ** -6 + 1e-3


def _sample(lib, logits, T, P, K, u, bias=None, n_variants=1, T2=1.0, P2=1.0, segs=(0, 0, 0, 0)):
    B, V = logits.shape
    out = torch.full((B, 2), -7, dtype=torch.int32, device="cuda")
    bf = lambda v: float(torch.tensor(v, dtype=torch.bfloat16).float())
    rc = lib.fsb_sample(logits.data_ptr(), None if bias is None else bias.data_ptr(), B, V, K, bf(T), bf(P), n_variants,
                        bf(T2), bf(P2), None if u is None else u.data_ptr(), 0, *segs, out.data_ptr(),
                        torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    return out.cpu()


def _oracle_sample_stable(logits, T, P, K, u):
    """oracle sample() with ties in ascending index order (stable sort = what torch.sort does on CUDA)."""
    import oracle.dual_ar_ref as M
    real = torch.sort
    torch.sort = lambda x, descending=False, **k: real(x, descending=descending, stable=True)
    try:
        idx, _ = M.sample(logits[None, None], torch.tensor(T, dtype=torch.bfloat16), torch.tensor(P, dtype=torch.bfloat16),
                          K, rand_like=lambda t: u)
    finally:
        torch.sort = real
    return int(idx)


def test_sampling_golden_cases(lib):
    """Reference sample() outputs (inference.py:80-93, recorded with injected bf16 uniforms) vs fsb_sample."""
    cases = load_golden("sampling_cases")
    mism_ref, mism_stable = [], []
    for i, c in enumerate(cases):
        got = int(_sample(lib, c["logits"][None].cuda(), c["temperature"], c["top_p"], c["top_k"], c["u"][None, None].cuda())[0, 0])
        if got != c["idx"]:
            mism_ref.append((i, c["tag"]))
        if got != _oracle_sample_stable(c["logits"], c["temperature"], c["top_p"], c["top_k"], c["u"]):
            mism_stable.append((i, c["tag"], got))
    # bit-exact against the oracle with a defined (stable) tie order
    assert not mism_stable, mism_stable
    # against the recorded CPU reference only tie-order-dependent cases may differ (torch's CPU sort is unstable)
    assert all(t in ("all_equal", "quantised_ties", "random", "biased") for _, t in mism_ref), mism_ref
    assert len(mism_ref) <= max(3, len(cases) // 20), mism_ref


# ------------------------------------------------------------------------------------------- many-row Linear (tcgen05 stream-K)
def _rows_gemm(lib, W, x, W2=None, res=None, epi=0, scratch=None):
    M, K = x.shape
    N = W.shape[0]
    out = torch.empty((M, N), dtype=torch.bfloat16, device="cuda")
    nbytes = lib.fsb_rows_gemm_scratch_bytes()
    if scratch is None:
        scratch = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
    p = lambda t: None if t is None else t.data_ptr()
    rc = lib.fsb_rows_gemm(p(W), p(W2), p(x), p(res), p(out), M, N, K, epi, scratch.data_ptr(), nbytes,
                           torch.cuda.current_stream().cuda_stream)
    assert rc == 0, rc
    torch.cuda.synchronize()
    return out


def _rows_ref(W, x, W2=None, res=None, epi=0):
    """fp32 reference of the same op with the reference's rounding points (one rounding per Linear, llama.py:838-843,985-986)."""
    acc = x.float() @ W.float().t()
    if epi == 0:
        return acc.bfloat16()
    if epi == 1:
        return (res.float() + acc.bfloat16().float()).bfloat16()
    g = acc.bfloat16().float()
    up = (x.float() @ W2.float().t()).bfloat16().float()
    return ((g / (1 + torch.exp(-g))).bfloat16().float() * up).bfloat16()


ROWS_SHAPES = [  # (M, N, K): every Linear of the path at tiny and S2-Pro sizes, ragged N and M, tile-count edge cases
    (12, 768, 512), (12, 512, 512), (12, 1536, 512), (12, 512, 1536), (33, 1024, 1536), (9, 130, 64), (64, 128, 9728),
    (64, 6144, 2560), (64, 2560, 4096), (64, 9728, 2560), (64, 2560, 9728), (64, 4096, 2560), (17, 2560, 4096),
    (128, 6144, 2560), (100, 2560, 9728),
]


@pytest.mark.parametrize("M,N,K", ROWS_SHAPES)
@pytest.mark.parametrize("epi", [0, 1, 2])
def test_rows_gemm_matches_fp32_reference(lib, M, N, K, epi):
    if epi == 2 and N % 128:
        pytest.skip("gate/up runs as one launch over both matrices: rows per matrix must be a multiple of the 128-row tile")
    g = torch.Generator().manual_seed(M * 131 + N * 7 + K + epi)
    W = (torch.randn(N, K, generator=g) * 0.02).bfloat16().cuda()
    W2 = (torch.randn(N, K, generator=g) * 0.02).bfloat16().cuda() if epi == 2 else None
    x = torch.randn(M, K, generator=g).bfloat16().cuda()
    res = torch.randn(M, N, generator=g).bfloat16().cuda() if epi == 1 else None
    scratch = torch.zeros(lib.fsb_rows_gemm_scratch_bytes(), dtype=torch.uint8, device="cuda")
    got = _rows_gemm(lib, W, x, W2, res, epi, scratch)
    ref = _rows_ref(W, x, W2, res, epi)
    # fp32 accumulation order differs from torch's, so the Linear output may round to the neighbouring bf16 value:
    # allow one bf16 ulp of the LARGEST operand of the epilogue (2^-7 relative) + a small absolute term
    err = (got.float() - ref.float()).abs()
    lin = (x.float() @ W.float().t()).abs()
    mag = ref.float().abs()
    if epi == 1:
        mag = mag + lin + res.float().abs()
    if epi == 2:
        mag = mag + lin * (x.float() @ W2.float().t()).abs()
    tol = mag * 2.0 ** -7 + 2e-3
    bad = int((err > tol).sum())
    print(f"[parity] rows_gemm M={M} N={N} K={K} epi={epi}: max-abs {float(err.max()):.5f}, exact {float((got == ref).float().mean()):.4f}")
    assert bad == 0, (bad, float(err.max()))
    # stream-K partial sums are added in CTA order: a second launch on the same scratch is bit-identical
    again = _rows_gemm(lib, W, x, W2, res, epi, scratch)
    assert torch.equal(got, again)


def test_rows_gemm_lm_head_shape(lib):
    """The S2-Pro LM head (N = 155776 rows, 1217 weight tiles) over 64 rows."""
    g = torch.Generator(device="cuda").manual_seed(5)
    N, K, M = 155776, 2560, 64
    W = torch.empty(N, K, dtype=torch.bfloat16, device="cuda").normal_(0, 0.02, generator=g)
    x = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    got = _rows_gemm(lib, W, x)
    ref = _rows_ref(W, x)
    err = (got.float() - ref.float()).abs()
    assert int((err > ref.float().abs() * 2.0 ** -7 + 2e-3).sum()) == 0, float(err.max())


@pytest.mark.parametrize("M,N,K,nw", [(64, 6144, 2560, 48), (12, 4096, 2560, 32), (100, 768, 512, 16), (64, 1024, 64, 64),
                                      (1, 96, 128, 48), (128, 6144, 2560, 48), (33, 4096, 9728, 32)])
def test_rows_gemm_narrow_matches_fp32_reference(lib, M, N, K, nw):
    """No-split variant (activations as the A operand, nw weight rows per CTA as the B operand)."""
    g = torch.Generator().manual_seed(M * 17 + N + K + nw)
    W = (torch.randn(N, K, generator=g) * 0.02).bfloat16().cuda()
    x = torch.randn(M, K, generator=g).bfloat16().cuda()
    out = torch.full((M, N), 7.0, dtype=torch.bfloat16, device="cuda")
    rc = lib.fsb_rows_gemm_narrow(W.data_ptr(), x.data_ptr(), out.data_ptr(), M, N, K, nw, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, rc
    torch.cuda.synchronize()
    ref = _rows_ref(W, x)
    err = (out.float() - ref.float()).abs()
    tol = ref.float().abs() * 2.0 ** -7 + 2e-3
    print(f"[parity] rows_gemm_narrow M={M} N={N} K={K} nw={nw}: max-abs {float(err.max()):.5f}, exact {float((out == ref).float().mean()):.4f}")
    assert int((err > tol).sum()) == 0, float(err.max())


# ------------------------------------------------------------------------------------------- sampling at scale (>= 1e5 cases)
def _make_sampling_batch(kind, B, V, gen):
    """bf16 logits [B][V] + bf16 uniforms [B][V] of one family of cases"""
    lg = torch.randn(B, V, generator=gen) * 2.0
    u = torch.randint(0, 256, (B, V), generator=gen).float() / 256
    if kind == "ties":            # heavy quantisation: many exact ties at every rank, incl. across the top-k boundary
        lg = (lg * 2).round() / 2
    elif kind == "u0":            # a quarter of the draws are u = 0 (score 0), some rows entirely
        u = torch.where(torch.rand(B, V, generator=gen) < 0.25, torch.zeros(()), u)
        u[: B // 16] = 0.0
    elif kind == "top_p_edge":    # a few dominant candidates whose cumulative mass sits near top_p
        lg[:, :4] += torch.tensor([6.0, 5.5, 5.0, 4.5])
        lg = lg[:, torch.randperm(V, generator=gen)]
    elif kind == "masked":        # -inf-biased vocabulary: only a short range is finite (constrained decode)
        lo = int(torch.randint(0, V - 40, (1,), generator=gen))
        keep = torch.zeros(V, dtype=torch.bool)
        keep[lo: lo + 33] = True
        keep[int(torch.randint(0, V, (1,), generator=gen))] = True
        lg = torch.where(keep, lg, torch.full((), float("-inf")))
    elif kind == "flat":          # all equal
        lg = torch.zeros(B, V) + 0.5
    return lg.bfloat16(), u.bfloat16()


def test_sampling_100k_on_the_fly_cases_vs_oracle(lib):
    """>= 1e5 random cases of sample() (inference.py:80-93): fsb_sample on the GPU against the oracle's restatement with a
    stable sort (ties by ascending index, what torch.sort yields on CUDA), identical bf16 logits and uniforms -> the drawn
    index must be IDENTICAL in every case.  Families: random, quantised ties, u = 0, top-p boundary, -inf-masked vocabulary,
    all-equal; top_k in {1, 30}; T in {0.8, 1.0, 0.3}; top_p in {0.8, 0.9, 0.5}."""
    import oracle.dual_ar_ref as M
    V, B = 512, 2048
    gen = torch.Generator().manual_seed(2024)
    combos = [(0.8, 0.8, 30), (1.0, 0.9, 30), (0.8, 0.8, 1), (0.3, 0.5, 30), (1.0, 0.9, 7)]
    kinds = ["random", "ties", "u0", "top_p_edge", "masked", "flat", "random", "ties", "u0", "masked"]
    real = torch.sort
    torch.sort = lambda x, descending=False, **k: real(x, descending=descending, stable=True)
    total = bad = 0
    worst = []
    try:
        for kind in kinds:
            for (T, P, K) in combos:
                lg, u = _make_sampling_batch(kind, B, V, gen)
                got = _sample(lib, lg.cuda(), T, P, K, u.cuda().view(B, 1, V))[:, 0]
                t_, p_ = torch.tensor(T, dtype=torch.bfloat16), torch.tensor(P, dtype=torch.bfloat16)
                for b in range(B):
                    row_u = u[b]
                    idx, _ = M.sample(lg[b][None, None], t_, p_, K, rand_like=lambda t: row_u)
                    total += 1
                    if int(idx) != int(got[b]):
                        bad += 1
                        if len(worst) < 5:
                            worst.append((kind, T, P, K, b, int(idx), int(got[b])))
    finally:
        torch.sort = real
    print(f"[parity] sampling: {total} on-the-fly cases, {bad} mismatches")
    assert total >= 100000
    assert bad == 0, worst


def test_philox_mode_draws_follow_the_reference_distribution(lib):
    """Production mode (no injected uniforms: Philox4x32-10 keyed by seed, counters = index / draw / frame / request).
    Two-sample chi-square of 2^17 GPU draws against 2^17 draws of the oracle's multinomial_sample_one_no_sync with torch's
    bf16 rand_like (the reference's RNG, 8 random bits per draw) on the same 24-way distribution; and two seeds must differ."""
    import oracle.dual_ar_ref as M
    V, N = 24, 1 << 17
    g = torch.Generator().manual_seed(77)
    lg = (torch.randn(V, generator=g) * 1.5).bfloat16()
    T, P, K = 1.0, 0.9, 30
    rows = lg[None].expand(N, V).contiguous().cuda()
    out = torch.full((N, 2), -7, dtype=torch.int32, device="cuda")
    bf = lambda v: float(torch.tensor(v, dtype=torch.bfloat16).float())

    def draw(seed):
        rc = lib.fsb_sample(rows.data_ptr(), None, N, V, K, bf(T), bf(P), 1, 1.0, 1.0, None, seed, 0, 0, 0, 0, out.data_ptr(),
                            torch.cuda.current_stream().cuda_stream)
        assert rc == 0
        torch.cuda.synchronize()
        return out[:, 0].cpu().long().clone()

    a, a2 = draw(12345), draw(999)
    assert not torch.equal(a, a2)                                   # the seed matters
    assert torch.equal(a, draw(12345))                              # and the stream is reproducible
    probs = M.logits_to_probs(lg, torch.tensor(T, dtype=torch.bfloat16), torch.tensor(P, dtype=torch.bfloat16), K)
    torch.manual_seed(5)
    q = -torch.log(torch.rand((N, V), dtype=torch.bfloat16))
    b = torch.argmax(probs[None] / q, dim=-1)
    ca, cb_ = torch.bincount(a, minlength=V).double(), torch.bincount(b, minlength=V).double()
    live = (ca + cb_) > 0
    chi2 = float((((ca - cb_) ** 2) / (ca + cb_).clamp_min(1))[live].sum())
    dof = int(live.sum()) - 1
    print(f"[parity] Philox draws vs torch bf16 rand_like draws: chi2 {chi2:.1f} on {dof} dof; support {int(live.sum())} of {V}")
    # 99.99 % quantile of chi-square is below dof + 4 * sqrt(2 * dof) + 12 for dof <= 30
    assert chi2 < dof + 4 * (2 * dof) ** 0.5 + 12, (chi2, dof)
    assert int((ca > 0).sum()) == int((cb_ > 0).sum())              # same support (the top-p mask removes the same tail)
