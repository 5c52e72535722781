"""GPU parity at the S2-Pro WIDTH (dim 2560, ffn 9728, 32 heads / 8 kv heads, head_dim 128) on a shallow model
(2 slow + 1 fast layers, vocab 8192) so that the CPU oracle finishes in seconds.  This is the geometry bench.py runs:
  * frame_mega_kernel<128,4> with all three row layouts of its weight ring (K = 2560 -> 4 rows per stage, K = 4096 (wo) ->
    2 rows per stage / two warps per row, K = 9728 (w2) -> 1 row per stage / four warps per row with the cross-warp
    reduction) and attn_split_body<128,4>;
  * a 64-row step through rows_gemm_kernel + attn_tc_kernel with contexts of 1.5k-2k tokens.
Tolerance: the oracle's own prompt-route vs decode-route logits differ by 0.031 max-abs here (one bf16 ulp at |logit| ~ 4.4),
so logits are accepted within max(0.05, 4 ulp of max |logit|) AND 2.5 % of max |logit| (the tiny-model criterion, 0.05 at
|logit| ~ 2, is 6.4 ulp); both are printed."""
import pytest
import torch

from oracle import dual_ar_ref as R
from tests.util import err_stats

pytestmark = pytest.mark.gpu

REL_TOL = 0.025


def s2w_config(max_seq_len=512):
    return R.DualARConfig(vocab_size=8192, n_layer=2, n_head=32, dim=2560, intermediate_size=9728, n_local_heads=8,
                          head_dim=128, max_seq_len=max_seq_len, codebook_size=1024, num_codebooks=10,
                          semantic_begin_id=4096, semantic_end_id=5119, n_fast_layer=1, im_end_id=7)


def _abs_tol(ref):
    mx = float(ref.float().abs().max())
    ulp = 2.0 ** (torch.tensor(mx).log2().floor().item() - 7)
    return max(0.05, 4 * ulp)


def _check(name, got, ref):
    mabs, mrel, frac = err_stats(got, ref)
    tol = _abs_tol(ref)
    print(f"[parity] {name}: max-abs {mabs:.4f} (tol {tol:.4f})  rel-to-max {mrel:.4f}  frac-differing {frac:.3f}")
    assert mabs <= tol and mrel <= REL_TOL, (name, mabs, mrel)


@pytest.fixture(scope="module")
def s2w():
    from fish_speech_b200 import DualARTransformerB200
    cfg = s2w_config()
    sd = R.make_state_dict(cfg, seed=11)
    m = DualARTransformerB200(cfg.__dict__, sd, device="cuda:0", max_rows=1)
    m.setup_caches(1, cfg.max_seq_len)
    om = R.DualARRef(cfg, sd)
    om.setup_caches(1, cfg.max_seq_len)
    return cfg, sd, m, om


def _frame(m, seq, uni, **opts):
    """prompt pass over seq[:, :-1] + ONE fused frame on the last column; returns slow logits, fast logits, tokens"""
    from fish_speech_b200 import dual_ar
    cfg = m.config
    for k, v in opts.items():
        m.set_option(k, v)
    m._session_frames = 8  # decode_one_token_ar picks its session by this size: same object as below
    sess = m.session(1, 8)
    sess.uniforms = uni.cuda()
    bias = dual_ar.make_semantic_logit_bias(m)
    tok = dual_ar.decode_one_token_ar(m, seq[None].cuda(), torch.arange(seq.shape[1]), 0.8, 0.8, 30, bias)
    torch.cuda.synchronize()
    sess.check_device()
    sess.uniforms = None
    V, cb, C = cfg.vocab_size, cfg.codebook_size, cfg.num_codebooks
    slow = m._buffer("slow_logits")[:V].clone().cpu()
    fl = m._buffer("fast_logits")
    fast = fl.view(C, fl.numel() // (C * cb), cb)[1:, 0].clone().cpu()
    return slow, fast, tok.cpu()


def test_frame_kernel_128_at_s2_width_vs_oracle_and_chain(s2w):
    cfg, sd, m, om = s2w
    V, cb, C = cfg.vocab_size, cfg.codebook_size, cfg.num_codebooks
    assert m.launches_per_frame(1) == 1  # the persistent frame kernel is what runs at this width
    gen = torch.Generator().manual_seed(21)
    lo = min(cfg.im_end_id, cfg.semantic_begin_id) & ~7
    hi = (max(cfg.im_end_id, cfg.semantic_end_id) + 1 + 7) & ~7
    for T, span in ((300, 40), (97, 0), (511, 200)):  # 511: the last position of the cache, every attention split in use
        seq = R.make_prompt(cfg, T, seed=T, voice_span=span)
        uni = (torch.randint(0, 256, (1, 1, 2 * V + (C - 1) * cb), generator=gen).float() / 256).to(torch.bfloat16)
        slow_m, fast_m, tok_m = _frame(m, seq, uni, mega=1, masked_head=0)
        slow_c, fast_c, tok_c = _frame(m, seq, uni, mega=0, masked_head=0)
        slow_r, fast_r, tok_r = _frame(m, seq, uni, mega=1, masked_head=1)
        m.set_option("mega", 1)
        # oracle: prompt-route slow logits; fast logits teacher-forced with the codes the GPU frame sampled
        ref, hid = om.forward_generate(seq[None], torch.arange(T))
        _check(f"frame kernel<128,4> slow logits T={T}", slow_m, ref[0, -1])
        _check(f"kernel chain slow logits T={T}", slow_c, ref[0, -1])
        mabs, _, _ = err_stats(slow_m, slow_c)
        print(f"[parity] frame kernel vs chain slow logits T={T}: max-abs {mabs:.4f}")
        assert mabs <= _abs_tol(ref[0, -1])
        assert torch.equal(slow_r[lo:hi], slow_m[lo:hi]) and torch.equal(tok_r, tok_m)  # restricted head: same rows bit for bit
        assert cfg.semantic_begin_id <= int(tok_m[0]) <= cfg.semantic_end_id or int(tok_m[0]) == cfg.im_end_id
        assert int(tok_m[1]) == max(0, min(int(tok_m[0]) - cfg.semantic_begin_id, cb - 1))  # inference.py:151-155
        om.forward_generate_fast(hid, torch.tensor([0]))
        h = om.fast_embeddings(tok_m[1].long().view(1))
        for i in range(1, C):
            fl = om.forward_generate_fast(h, torch.tensor([i]))
            _check(f"frame kernel<128,4> fast logits T={T} cb {i}", fast_m[i - 1], fl[0, -1])
            h = om.fast_embeddings(tok_m[i + 1].long().view(1))
        if torch.equal(tok_m, tok_c):
            mabs, _, _ = err_stats(fast_m, fast_c)
            print(f"[parity] frame kernel vs chain fast logits T={T}: max-abs {mabs:.4f}")
            assert mabs <= 0.1


def test_tiled_frame_kernel_vs_one_tile_kernel_chain_and_oracle(s2w):
    """frame_mega_kernel<128,4,true>: a split walks its keys in tiles of 128 with a running softmax (caches longer than 16 x 128
    tokens).  Forced onto the 512-token cache (mega_long = 2 keeps the requested split count): 1 split x 4 tiles and 3 splits
    x 2 tiles (ragged last tile, the new key in the last tile of the last split) against the one-tile kernel, the kernel
    chain and the oracle."""
    cfg, sd, m, om = s2w
    V, cb, C = cfg.vocab_size, cfg.codebook_size, cfg.num_codebooks
    gen = torch.Generator().manual_seed(5)
    try:
        for T, span in ((511, 200), (300, 40), (130, 0)):
            seq = R.make_prompt(cfg, T, seed=T, voice_span=span)
            uni = (torch.randint(0, 256, (1, 1, 2 * V + (C - 1) * cb), generator=gen).float() / 256).to(torch.bfloat16)
            slow_1, fast_1, tok_1 = _frame(m, seq, uni, mega=1, masked_head=0, mega_long=0, attn_splits=0)
            ref, _ = om.forward_generate(seq[None], torch.arange(T))
            for splits in (1, 3):
                slow_t, fast_t, tok_t = _frame(m, seq, uni, mega=1, masked_head=0, mega_long=2, attn_splits=splits)
                assert m.launches_per_frame(1) == 1
                _check(f"tiled frame kernel ({splits} split(s)) slow logits T={T}", slow_t, ref[0, -1])
                mabs, _, _ = err_stats(slow_t, slow_1)
                print(f"[parity] tiled ({splits} split(s)) vs one-tile frame kernel slow logits T={T}: max-abs {mabs:.4f}")
                assert mabs <= _abs_tol(ref[0, -1])
                if torch.equal(tok_t, tok_1):
                    assert float((fast_t - fast_1).abs().max()) <= 0.1
    finally:
        m.set_option("mega_long", 0)
        m.set_option("attn_splits", 0)


def test_frame_kernel_past_2048_tokens_runs_persistent_and_matches_chain_and_oracle():
    """A 2560-token cache (more than 16 splits x 128 keys): with option mega_long = 1 the frame is still ONE launch (tiled
    instantiation), and its slow / fast logits agree with the kernel chain (the default at this length) and the oracle at
    context 2300."""
    from fish_speech_b200 import DualARTransformerB200
    cfg = s2w_config(max_seq_len=2560)
    sd = R.make_state_dict(cfg, seed=12)
    m = DualARTransformerB200(cfg.__dict__, sd, device="cuda:0", max_rows=1)
    m.setup_caches(1, cfg.max_seq_len)
    om = R.DualARRef(cfg, sd)
    om.setup_caches(1, cfg.max_seq_len)
    V, cb, C = cfg.vocab_size, cfg.codebook_size, cfg.num_codebooks
    T = 2300
    seq = R.make_prompt(cfg, T, seed=8, voice_span=900)
    uni = (torch.randint(0, 256, (1, 1, 2 * V + (C - 1) * cb), generator=torch.Generator().manual_seed(9)).float() / 256).to(torch.bfloat16)
    assert m.launches_per_frame(1) > 1
    slow_c, fast_c, tok_c = _frame(m, seq, uni, mega=1, masked_head=0)
    slow_m, fast_m, tok_m = _frame(m, seq, uni, mega=1, masked_head=0, mega_long=1)
    assert m.launches_per_frame(1) == 1
    m.set_option("mega_long", 0)
    ref, _ = om.forward_generate(seq[None], torch.arange(T))
    _check("tiled frame kernel slow logits T=2300", slow_m, ref[0, -1])
    _check("kernel chain slow logits T=2300", slow_c, ref[0, -1])
    mabs, _, _ = err_stats(slow_m, slow_c)
    print(f"[parity] tiled frame kernel vs chain slow logits T=2300: max-abs {mabs:.4f}")
    assert mabs <= _abs_tol(ref[0, -1])
    if torch.equal(tok_m, tok_c):
        assert float((fast_m - fast_c).abs().max()) <= 0.1


def test_frame_kernel_128_free_run_is_deterministic_and_in_range(s2w):
    """48 frames back to back from one graph at the S2 width: replay-deterministic, graph == eager, codes in range."""
    from fish_speech_b200 import generate
    cfg, sd, m, om = s2w
    prompt = R.make_prompt(cfg, 64, seed=3, voice_span=10)
    runs = []
    for graph in (1, 0, 1):
        m.set_option("graph", graph)
        m.set_option("mega", 1)
        torch.manual_seed(0)  # generate() keys its Philox stream off torch's CPU generator (like utils.set_seed + rand_like)
        seq = generate(model=m, prompt=prompt.cuda(), max_new_tokens=48, temperature=0.8, top_p=0.8, top_k=30)
        m._session.check_device()
        runs.append(seq.cpu())
    m.set_option("graph", 1)
    assert torch.equal(runs[0], runs[1]) and torch.equal(runs[0], runs[2])
    new = runs[0][:, 64:]
    assert bool((new[1:] >= 0).all()) and bool((new[1:] < cfg.codebook_size).all())


def test_64_row_step_long_context_at_s2_width_vs_oracle():
    """BASELINE configs[2] geometry: 64 requests in flight with 1.5k-2k-token contexts, one decode step through
    rows_gemm_kernel (tcgen05) + attn_tc_kernel.  8 of the rows against the oracle's prompt-route logits (the oracle needs
    ~2 s per row), all 64 rows against the 8-row GEMV / CUDA-core attention chain (option batch_gemm = 0) on the same state."""
    from fish_speech_b200 import DualARTransformerB200
    from fish_speech_b200.dual_ar import _stream
    cfg = s2w_config(max_seq_len=2048)
    sd = R.make_state_dict(cfg, seed=11)
    B, V = 64, cfg.vocab_size
    lens = [1536 + ((i * 37) % 512) for i in range(B)]
    lens[0], lens[1], lens[2] = 2048, 1536, 1537        # the cache's last position; split boundaries
    prompts = [R.make_prompt(cfg, T, seed=700 + i, voice_span=(i % 3) * 50) for i, T in enumerate(lens)]
    m = DualARTransformerB200(cfg.__dict__, sd, device="cuda:0", max_rows=B)
    m.setup_caches(B, cfg.max_seq_len)
    for k, p in enumerate(prompts):
        m.prefill(p[:, :-1].cuda(), 0, k)
    tok = torch.stack([p[:, -1] for p in prompts]).to(torch.int32).cuda().contiguous()
    pos = torch.tensor([T - 1 for T in lens], dtype=torch.int32).cuda()

    def step(mode):
        m.set_option("batch_gemm", mode)
        m._ck(m.lib.fsb_dual_ar_slow_step(m._h, B, tok.data_ptr(), pos.data_ptr(), None, _stream()))
        torch.cuda.synchronize()
        return m._buffer("slow_logits")[: B * V].view(B, V).clone().cpu()

    lg = step(1)
    assert m.launches_per_frame(B) > 100       # the many-row path (not the frame kernel, not 8-row passes)
    base = step(0)                             # re-appends the same K/V at the same positions: same state
    m.set_option("batch_gemm", 1)
    mabs, mrel, _ = err_stats(lg, base)
    print(f"[parity] 64-row step vs 8-row chain, slow logits: max-abs {mabs:.4f} rel {mrel:.4f}")
    assert mabs <= 0.1 and mrel <= REL_TOL
    om = R.DualARRef(cfg, sd)
    om.setup_caches(1, cfg.max_seq_len)
    for k in (0, 1, 2, 9, 23, 40, 57, 63):
        ref, _ = om.forward_generate(prompts[k][None], torch.arange(lens[k]))
        _check(f"64-row step row {k} ctx {lens[k]}", lg[k], ref[0, -1])


def test_prompt_pass_on_own_tcgen05_kernel_vs_oracle_and_library(s2w):
    """Prompt pass (forward_generate with Tq > 1, llama.py:390-466) with every Linear on tcv_kernel (option prefill_gemm = 1,
    the default: bf16 store, residual add and SwiGLU in the GEMM epilogue with the reference's rounding points) against the
    oracle, against the same pass with cuBLAS GEMMs (prefill_gemm = 2) and against 8-row passes through the decode kernels
    (prefill_gemm = 0).  Ragged lengths: one tile, tile edge, several chunks."""
    cfg, sd, m, om = s2w
    for T in (2, 128, 129, 400):
        seq = R.make_prompt(cfg, T, seed=900 + T, voice_span=T // 3)
        ref, _ = om.forward_generate(seq[None], torch.arange(T))
        outs = {}
        for mode in (1, 2, 0):
            m.set_option("prefill_gemm", mode)
            outs[mode] = m.forward_generate(seq[None].cuda(), torch.arange(T)).logits[0, 0].float().cpu()
        m.set_option("prefill_gemm", 1)
        _check(f"prompt pass (own kernel) T={T}", outs[1], ref[0, -1])
        for other in (2, 0):
            mabs, mrel, _ = err_stats(outs[1], outs[other])
            print(f"[parity] prompt pass own kernel vs mode {other} T={T}: max-abs {mabs:.4f}")
            assert mabs <= _abs_tol(ref[0, -1])


def test_short_prompt_pieces_through_the_many_row_kernels_vs_prompt_pass_and_oracle():
    """A handle with 16 rows (a scheduler's) takes prompt pieces of <= 32 columns through the many-row decode kernels in passes of
    16 rows of ONE KV slot (option prefill_rows): the first frame's logits must agree with the oracle and with the prompt-pass
    kernels (prefill_rows = 0), for one pass, two passes, a last pass of one row, into slot 0 and slot 5, from position 0 and
    behind resident columns."""
    from fish_speech_b200 import DualARTransformerB200
    import ctypes as C
    from fish_speech_b200.dual_ar import _stream
    cfg = s2w_config()
    sd = R.make_state_dict(cfg, seed=13)
    m = DualARTransformerB200(cfg.__dict__, sd, device="cuda:0", max_rows=16)
    m.setup_caches(16, cfg.max_seq_len)
    om = R.DualARRef(cfg, sd)
    om.setup_caches(1, cfg.max_seq_len)
    V = cfg.vocab_size

    def first_logits(seq, slot, rows, split_at=0):
        m.set_option("prefill_rows", rows)
        T = seq.shape[1]
        if split_at:
            m.prefill(seq[:, :split_at].cuda(), 0, slot)                 # resident columns (prompt-pass kernels: 40 > 2 x 16)
        m.prefill(seq[:, split_at: T - 1].cuda(), split_at, slot)
        tok = torch.zeros((16, cfg.num_codebooks + 1), dtype=torch.int32, device="cuda")
        pos = torch.zeros((16,), dtype=torch.int32, device="cuda")
        tok[slot] = seq[:, -1].to(torch.int32).cuda()
        pos[slot] = T - 1
        m._ck(m.lib.fsb_dual_ar_slow_step(m._h, 16, tok.data_ptr(), pos.data_ptr(), None, _stream()))
        torch.cuda.synchronize()
        return m._buffer("slow_logits")[: 16 * V].view(16, V)[slot].float().cpu().clone()

    for T, slot, split_at in ((12, 0, 0), (17, 5, 0), (33, 5, 0), (34, 3, 0), (70, 2, 40)):
        seq = R.make_prompt(cfg, T, seed=40 + T, voice_span=T // 3)
        ref, _ = om.forward_generate(seq[None], torch.arange(T))
        a = first_logits(seq, slot, 1, split_at)
        b = first_logits(seq, slot, 0, split_at)
        _check(f"short prompt through the many-row kernels T={T} slot {slot}", a, ref[0, -1])
        mabs, _, _ = err_stats(a, b)
        print(f"[parity] many-row route vs prompt-pass kernels T={T}: max-abs {mabs:.4f}")
        assert mabs <= _abs_tol(ref[0, -1])
    m.set_option("prefill_rows", 1)


def test_short_prompt_in_one_many_row_pass_on_a_batch_1_handle(s2w):
    """A prompt piece of <= 128 columns goes through the many-row decode kernels in one pass on ANY handle (rows in the
    prompt-pass scratch, one attention item per row and head): first-frame logits against the oracle and against the
    prompt-pass kernels (prefill_rows = 0) for 2 .. 128 columns from position 0, and for 50 new columns behind 300 resident
    ones (the prefix-cache case)."""
    cfg, sd, m, om = s2w
    try:
        for T, resident in ((3, 0), (13, 0), (100, 0), (129, 0), (351, 300)):
            seq = R.make_prompt(cfg, T, seed=60 + T, voice_span=T // 3)
            ref, _ = om.forward_generate(seq[None], torch.arange(T))
            outs = {}
            for rows in (1, 0):
                m.set_option("prefill_rows", rows)
                m.prefix_cache = False
                if resident:
                    m.prefill(seq[:, :resident].cuda(), 0, 0)
                    m.prefill(seq[:, resident: T - 1].cuda(), resident, 0)
                    out = m.forward_generate(seq[None, :, -1:].cuda(), torch.tensor([T - 1]))
                else:
                    out = m.forward_generate(seq[None].cuda(), torch.arange(T))
                outs[rows] = out.logits[0, 0].float().cpu()
            _check(f"short prompt in one many-row pass T={T} (resident {resident})", outs[1], ref[0, -1])
            mabs, _, _ = err_stats(outs[1], outs[0])
            print(f"[parity] one many-row pass vs prompt-pass kernels T={T}: max-abs {mabs:.4f}")
            assert mabs <= _abs_tol(ref[0, -1])
    finally:
        m.set_option("prefill_rows", 1)
        m.prefix_cache = True


def test_flash_attention_prompt_pass_vs_per_row_attention_and_oracle(s2w):
    """fa_prefill_kernel (tcgen05 causal flash attention, option prefill_fa = 1, the default at head_dim 128) against the
    per-row decode attention (prefill_fa = 0) and the oracle: lengths around the 128-row / 128-key block edges, and a prompt
    continued at a non-zero offset (second call attends to the keys of the first through the cache)."""
    cfg, sd, m, om = s2w
    for T in (1, 127, 128, 129, 255, 300, 511):
        seq = R.make_prompt(cfg, T + 1, seed=1200 + T, voice_span=T // 4)
        ref, _ = om.forward_generate(seq[None], torch.arange(T + 1))
        out = {}
        for fa in (1, 0):
            m.set_option("prefill_fa", fa)
            out[fa] = m.forward_generate(seq[None].cuda(), torch.arange(T + 1)).logits[0, 0].float().cpu()
        m.set_option("prefill_fa", 1)
        _check(f"flash prompt pass T={T}", out[1], ref[0, -1])
        mabs, _, _ = err_stats(out[1], out[0])
        print(f"[parity] flash vs per-row attention T={T}: max-abs {mabs:.4f}")
        assert mabs <= _abs_tol(ref[0, -1])
    # continuation: 200 tokens, then 150 more at offset 200, then the decode step
    seq = R.make_prompt(cfg, 351, seed=77, voice_span=60)
    ref, _ = om.forward_generate(seq[None], torch.arange(351))
    m.prefill(seq[:, :200].cuda(), 0, 0)
    m.prefill(seq[:, 200:350].cuda(), 200, 0)
    got = m.forward_generate(seq[None, :, 350:].cuda(), torch.tensor([350])).logits[0, 0].float().cpu()
    _check("flash prompt pass in two calls (offset 200)", got, ref[0, -1])
