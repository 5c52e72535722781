"""CPU restatement of the reference Dual-AR decode path (TEST INFRASTRUCTURE).

Functional, state-dict driven (no nn.Module), written to execute the SAME
sequence of torch CPU ops — and therefore the same bf16 rounding points — as

  /root/reference/fish_speech/models/text2semantic/llama.py
  /root/reference/fish_speech/models/text2semantic/inference.py

`oracle/gen_golden.py` asserts bit-identity against the imported reference.
Each function cites the reference lines it follows.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional

import torch
import torch.nn.functional as F
from torch import Tensor
from torch.nn.attention import SDPBackend, sdpa_kernel

RAS_WIN_SIZE = 10  # inference.py:49
RAS_HIGH_TEMP = 1.0  # inference.py:50
RAS_HIGH_TOP_P = 0.9  # inference.py:51


def find_multiple(n: int, k: int) -> int:  # llama.py:21-24
    return n if n % k == 0 else n + k - (n % k)


@dataclass
class DualARConfig:
    """Subset of DualARModelArgs (llama.py:28-123) the decode path reads."""

    vocab_size: int = 8192
    n_layer: int = 4
    n_head: int = 8
    dim: int = 512
    intermediate_size: int = 1536
    n_local_heads: int = 2
    head_dim: int = 64
    rope_base: float = 1e6
    norm_eps: float = 1e-6
    max_seq_len: int = 512
    codebook_size: int = 1024
    num_codebooks: int = 10
    semantic_begin_id: int = 4096
    semantic_end_id: int = 5119
    n_fast_layer: int = 2
    attention_qk_norm: bool = True
    tie_word_embeddings: bool = True
    scale_codebook_embeddings: bool = True  # forced by _from_fish_qwen3_omni, llama.py:101-143
    norm_fastlayer_input: bool = True
    im_end_id: int = 7  # tokenizer.get_token_id("<|im_end|>") — 3rd int the hot path needs
    initializer_range: float = 0.02

    # fast_* default to the slow values (llama.py:100-123); kept equal here.
    @property
    def fast_dim(self) -> int:
        return self.dim

    def to_ref_kwargs(self) -> dict:
        return dict(
            vocab_size=self.vocab_size, n_layer=self.n_layer, n_head=self.n_head, dim=self.dim,
            intermediate_size=self.intermediate_size, n_local_heads=self.n_local_heads,
            head_dim=self.head_dim, rope_base=self.rope_base, norm_eps=self.norm_eps,
            max_seq_len=self.max_seq_len, codebook_size=self.codebook_size,
            num_codebooks=self.num_codebooks, semantic_begin_id=self.semantic_begin_id,
            semantic_end_id=self.semantic_end_id, n_fast_layer=self.n_fast_layer,
            attention_qk_norm=self.attention_qk_norm, tie_word_embeddings=self.tie_word_embeddings,
            scale_codebook_embeddings=self.scale_codebook_embeddings,
            norm_fastlayer_input=self.norm_fastlayer_input, attention_qkv_bias=False,
            attention_o_bias=False, initializer_range=self.initializer_range,
        )


def tiny_config(**kw) -> DualARConfig:
    """BASELINE.json configs[0]: 4 slow layers d=512, 2 fast layers, 10 codebooks."""
    return DualARConfig(**kw)


def s2_config(max_seq_len: int = 2048, vocab_size: int = 155776, **kw) -> DualARConfig:
    """S2-Pro shape (assumed): Qwen3-4B backbone + 4 fast layers."""
    return DualARConfig(
        vocab_size=vocab_size, n_layer=36, n_head=32, dim=2560, intermediate_size=9728,
        n_local_heads=8, head_dim=128, rope_base=1e6, norm_eps=1e-6, max_seq_len=max_seq_len,
        codebook_size=4096, num_codebooks=10, semantic_begin_id=vocab_size - 4096,
        semantic_end_id=vocab_size - 1, n_fast_layer=4, im_end_id=151645, **kw)


def state_dict_keys(cfg: DualARConfig) -> Dict[str, tuple]:
    """name -> shape, names exactly as the reference nn.Module registers them."""
    d, hd = cfg.dim, cfg.head_dim
    qkv = (cfg.n_head + 2 * cfg.n_local_heads) * hd
    out: Dict[str, tuple] = {
        "embeddings.weight": (cfg.vocab_size, d),
        "codebook_embeddings.weight": (cfg.codebook_size * cfg.num_codebooks, d),
        "norm.weight": (d,),
        "fast_embeddings.weight": (cfg.codebook_size, d),
        "fast_norm.weight": (d,),
        "fast_output.weight": (cfg.codebook_size, d),
    }
    if not cfg.tie_word_embeddings:
        out["output.weight"] = (cfg.vocab_size, d)
    for pre, n in (("layers", cfg.n_layer), ("fast_layers", cfg.n_fast_layer)):
        for i in range(n):
            p = f"{pre}.{i}."
            out[p + "attention.wqkv.weight"] = (qkv, d)
            out[p + "attention.wo.weight"] = (d, cfg.n_head * hd)
            if cfg.attention_qk_norm:
                out[p + "attention.q_norm.weight"] = (hd,)
                out[p + "attention.k_norm.weight"] = (hd,)
            out[p + "feed_forward.w1.weight"] = (cfg.intermediate_size, d)
            out[p + "feed_forward.w3.weight"] = (cfg.intermediate_size, d)
            out[p + "feed_forward.w2.weight"] = (d, cfg.intermediate_size)
            out[p + "ffn_norm.weight"] = (d,)
            out[p + "attention_norm.weight"] = (d,)
    return out


def make_state_dict(cfg: DualARConfig, seed: int = 0, dtype=torch.bfloat16,
                    norm_jitter: float = 0.1, scale: float = 1.0) -> Dict[str, Tensor]:
    """Seeded synthetic weights: N(0, initializer_range) for Linear/Embedding as
    llama.py:468-477; norm weights 1 + U(-j, j) (the reference inits them to 1; a
    non-trivial value exercises the weight multiply)."""
    g = torch.Generator().manual_seed(seed)
    sd: Dict[str, Tensor] = {}
    for name, shape in state_dict_keys(cfg).items():
        if len(shape) == 1:
            w = 1.0 + norm_jitter * (2 * torch.rand(shape, generator=g) - 1)
        else:
            w = torch.randn(shape, generator=g) * (cfg.initializer_range * scale)
        sd[name] = w.to(dtype)
    return sd


# ----------------------------------------------------------------------------- numerics
def rms_norm_double_round(x: Tensor, w: Tensor, eps: float) -> Tensor:
    """llama.py:989-1000: fp32 normalise -> cast to x.dtype -> * weight (2 roundings)."""
    xf = x.float()
    o = (xf * torch.rsqrt(torch.mean(xf * xf, dim=-1, keepdim=True) + eps)).type_as(x)
    return o * w


def precompute_freqs_cis(seq_len: int, n_elem: int, base: float) -> Tensor:
    """llama.py:1003-1022 (table stored bf16)."""
    freqs = 1.0 / (base ** (torch.arange(0, n_elem, 2)[: (n_elem // 2)].float() / n_elem))
    t = torch.arange(seq_len, device=freqs.device)
    freqs = torch.outer(t, freqs)
    freqs_cis = torch.polar(torch.ones_like(freqs), freqs)
    cache = torch.stack([freqs_cis.real, freqs_cis.imag], dim=-1)
    return cache.to(dtype=torch.bfloat16)


def apply_rotary_emb(x: Tensor, freqs_cis: Tensor) -> Tensor:
    """llama.py:1025-1037: interleaved pairs, fp32 math on the bf16 table."""
    xshaped = x.float().reshape(*x.shape[:-1], -1, 2)
    freqs_cis = freqs_cis.view(1, xshaped.size(1), 1, xshaped.size(3), 2)
    x_out2 = torch.stack(
        [
            xshaped[..., 0] * freqs_cis[..., 0] - xshaped[..., 1] * freqs_cis[..., 1],
            xshaped[..., 1] * freqs_cis[..., 0] + xshaped[..., 0] * freqs_cis[..., 1],
        ],
        -1,
    )
    return x_out2.flatten(3).type_as(x)


def eq_sdpa(q: Tensor, k: Tensor, v: Tensor, mask: Tensor) -> Tensor:
    """llama.py:947-975: fast-AR attention, every op in the model dtype."""
    L, S = q.size(-2), k.size(-2)
    scale_factor = 1 / math.sqrt(q.size(-1))
    attn_bias = torch.zeros(1, 1, L, S, dtype=q.dtype, device=q.device)
    attn_bias = torch.where(mask.logical_not(), float("-inf"), attn_bias)
    attn_weight = q @ k.transpose(-2, -1) * scale_factor
    attn_weight += attn_bias
    attn_weight = torch.softmax(attn_weight, dim=-1)
    return attn_weight @ v


# ----------------------------------------------------------------------------- model
class DualARRef:
    """Stateful (KV caches) functional mirror of DualARTransformer's inference
    methods (llama.py:659-827).  Weights: a state dict with the reference's names."""

    def __init__(self, cfg: DualARConfig, sd: Dict[str, Tensor]):
        self.config = cfg
        self.sd = sd
        self.dtype = sd["embeddings.weight"].dtype
        # llama.py:273-298 buffers
        self.freqs_cis = precompute_freqs_cis(cfg.max_seq_len, cfg.head_dim, cfg.rope_base)
        self.fast_freqs_cis = precompute_freqs_cis(cfg.num_codebooks, cfg.head_dim, cfg.rope_base)
        self.causal_mask = torch.tril(torch.ones(cfg.max_seq_len, cfg.max_seq_len, dtype=torch.bool))
        self.max_seq_len = -1
        self.max_batch_size = -1
        self.kv: List[List[Tensor]] = []
        self.fast_kv: List[List[Tensor]] = []

    # llama.py:307-324, 707-721
    def setup_caches(self, max_batch_size: int, max_seq_len: int, dtype=torch.bfloat16):
        if self.max_seq_len >= max_seq_len and self.max_batch_size >= max_batch_size:
            return
        c = self.config
        max_seq_len = find_multiple(max_seq_len, 8)
        self.max_seq_len, self.max_batch_size = max_seq_len, max_batch_size
        shp = (max_batch_size, c.n_local_heads, max_seq_len, c.head_dim)
        self.kv = [[torch.zeros(shp, dtype=dtype), torch.zeros(shp, dtype=dtype)] for _ in range(c.n_layer)]
        fshp = (max_batch_size, c.n_local_heads, c.num_codebooks, c.head_dim)
        self.fast_kv = [[torch.zeros(fshp, dtype=dtype), torch.zeros(fshp, dtype=dtype)]
                        for _ in range(c.n_fast_layer)]

    # llama.py:883-945
    def _attention(self, p: str, x: Tensor, freqs_cis: Tensor, mask: Tensor, input_pos: Tensor,
                   cache: List[Tensor], use_sdpa: bool) -> Tensor:
        c, sd = self.config, self.sd
        bsz, seqlen, _ = x.shape
        q_size, kv_size = c.n_head * c.head_dim, c.n_local_heads * c.head_dim
        q, k, v = F.linear(x, sd[p + "wqkv.weight"]).split([q_size, kv_size, kv_size], dim=-1)
        q = q.view(bsz, seqlen, c.n_head, c.head_dim)
        k = k.view(bsz, seqlen, c.n_local_heads, c.head_dim)
        v = v.view(bsz, seqlen, c.n_local_heads, c.head_dim)
        if c.attention_qk_norm:  # nn.RMSNorm: single rounding (llama.py:862-863)
            q = F.rms_norm(q, (c.head_dim,), sd[p + "q_norm.weight"], c.norm_eps)
            k = F.rms_norm(k, (c.head_dim,), sd[p + "k_norm.weight"], c.norm_eps)
        q = apply_rotary_emb(q, freqs_cis)
        k = apply_rotary_emb(k, freqs_cis)
        q, k, v = map(lambda t: t.transpose(1, 2), (q, k, v))
        cache[0][:, :, input_pos] = k  # KVCache.update, llama.py:205-214
        cache[1][:, :, input_pos] = v
        k, v = cache[0], cache[1]
        rep = c.n_head // c.n_local_heads
        k = k.repeat_interleave(rep, dim=1)
        v = v.repeat_interleave(rep, dim=1)
        if use_sdpa:
            y = F.scaled_dot_product_attention(q, k, v, attn_mask=mask, dropout_p=0.0)
        else:
            y = eq_sdpa(q, k, v, mask)
        y = y.transpose(1, 2).contiguous().view(bsz, seqlen, q_size)
        return F.linear(y, sd[p + "wo.weight"])

    # llama.py:838-843, 985-986
    def _block(self, p: str, x, freqs_cis, mask, input_pos, cache, use_sdpa) -> Tensor:
        c, sd = self.config, self.sd
        h = x + self._attention(
            p + "attention.", rms_norm_double_round(x, sd[p + "attention_norm.weight"], c.norm_eps),
            freqs_cis, mask, input_pos, cache, use_sdpa)
        n = rms_norm_double_round(h, sd[p + "ffn_norm.weight"], c.norm_eps)
        ff = F.linear(
            F.silu(F.linear(n, sd[p + "feed_forward.w1.weight"])) * F.linear(n, sd[p + "feed_forward.w3.weight"]),
            sd[p + "feed_forward.w2.weight"])
        return h + ff

    # llama.py:390-466 (+ Identity fast_project_in, llama.py:818-827)
    def forward_generate(self, inp: Tensor, input_pos: Tensor, return_layers: bool = False):
        c, sd = self.config, self.sd
        embeds = []
        for i in range(c.num_codebooks):
            embeds.append(F.embedding(inp[:, i + 1] + i * c.codebook_size, sd["codebook_embeddings.weight"]))
        vq_embeds_sum = torch.stack(embeds, dim=1).sum(dim=1)
        vq_masks = (inp[:, 0] >= c.semantic_begin_id) & (inp[:, 0] <= c.semantic_end_id)
        vq_embeds_sum[~vq_masks] = 0
        x = F.embedding(inp[:, 0], sd["embeddings.weight"]) + vq_embeds_sum
        if c.scale_codebook_embeddings:
            x = torch.where(vq_masks.unsqueeze(-1).expand_as(x), x / math.sqrt(c.num_codebooks + 1), x)
        mask = self.causal_mask[None, None, input_pos, : self.max_seq_len]
        freqs_cis = self.freqs_cis[input_pos]
        layers_out = [x]
        for i in range(c.n_layer):
            x = self._block(f"layers.{i}.", x, freqs_cis, mask, input_pos, self.kv[i], True)
            if return_layers:
                layers_out.append(x)
        if x.size(1) > 1:
            x = x[:, -1:]
        slow_out = rms_norm_double_round(x, sd["norm.weight"], c.norm_eps)
        w = sd["embeddings.weight"] if c.tie_word_embeddings else sd["output.weight"]
        logits = F.linear(slow_out, w)
        hidden = slow_out if c.norm_fastlayer_input else x
        if return_layers:
            return logits, hidden, layers_out
        return logits, hidden

    # llama.py:798-816
    def forward_generate_fast(self, x: Tensor, input_pos: Tensor) -> Tensor:
        c, sd = self.config, self.sd
        x = x.view(x.shape[0], 1, -1)
        fast_mask = self.causal_mask[None, None, input_pos, : c.num_codebooks]
        fast_freqs_cis = self.fast_freqs_cis[input_pos]
        for i in range(c.n_fast_layer):
            x = self._block(f"fast_layers.{i}.", x, fast_freqs_cis, fast_mask, input_pos, self.fast_kv[i], False)
        fast_out = rms_norm_double_round(x, sd["fast_norm.weight"], c.norm_eps)
        return F.linear(fast_out, sd["fast_output.weight"])

    def fast_embeddings(self, a: Tensor) -> Tensor:
        return F.embedding(a, self.sd["fast_embeddings.weight"])


# ----------------------------------------------------------------------------- sampling
RandLike = Callable[[Tensor], Tensor]


def multinomial_sample_one_no_sync(probs_sort: Tensor, rand_like: RandLike = torch.rand_like) -> Tensor:
    """inference.py:43-46: exponential-race argmax; `rand_like` is injectable."""
    q = rand_like(probs_sort)
    q = -torch.log(q)
    return torch.argmax(probs_sort / q, dim=-1, keepdim=True).to(dtype=torch.int)


def logits_to_probs(logits: Tensor, temperature: Tensor, top_p: Tensor, top_k: int) -> Tensor:
    """inference.py:54-77.  Temperature is applied AFTER the top-p/top-k mask."""
    sorted_logits, sorted_indices = torch.sort(logits, descending=True)
    cum_probs = torch.cumsum(torch.nn.functional.softmax(sorted_logits, dim=-1), dim=-1)
    indices = torch.arange(sorted_logits.shape[-1], device=sorted_logits.device)
    top_k_mask = indices >= top_k
    sorted_indices_to_remove = (cum_probs > top_p) | top_k_mask
    sorted_indices_to_remove[0] = False
    indices_to_remove = sorted_indices_to_remove.scatter(
        dim=-1, index=sorted_indices, src=sorted_indices_to_remove)
    logits = torch.where(indices_to_remove, float("-Inf"), logits)
    logits = logits / torch.clip(temperature, min=1e-5)
    return torch.nn.functional.softmax(logits, dim=-1)


def sample(logits: Tensor, temperature: Tensor, top_p: Tensor, top_k: int,
           rand_like: RandLike = torch.rand_like):
    """inference.py:80-93 (batch 0, last position only)."""
    probs = logits_to_probs(logits[0, -1], temperature, top_p, top_k)
    return multinomial_sample_one_no_sync(probs, rand_like), probs


# ----------------------------------------------------------------------------- one frame
def decode_one_token_ar(model: DualARRef, x: Tensor, input_pos: Tensor, temperature: Tensor,
                        top_p: Tensor, top_k: int, semantic_logit_bias: Tensor,
                        audio_masks=None, audio_parts=None, previous_tokens: Optional[Tensor] = None,
                        rand_like: RandLike = torch.rand_like, trace: Optional[dict] = None) -> Tensor:
    """inference.py:96-181.  `trace`, when given, receives slow/fast logits for parity."""
    c = model.config
    logits, hidden_states = model.forward_generate(x, input_pos)
    biased_logits = logits + semantic_logit_bias
    main_token_normal = sample(biased_logits, temperature, top_p, top_k, rand_like)[0]
    high_temp = torch.tensor(RAS_HIGH_TEMP, dtype=temperature.dtype)
    high_top_p = torch.tensor(RAS_HIGH_TOP_P, dtype=top_p.dtype)
    main_token_high = sample(biased_logits, high_temp, high_top_p, top_k, rand_like)[0]
    if previous_tokens is not None:
        in_window = (previous_tokens[0] == main_token_normal).any()
        is_semantic = (main_token_normal >= c.semantic_begin_id) & (main_token_normal <= c.semantic_end_id)
        main_token_normal = torch.where(in_window & is_semantic, main_token_high, main_token_normal)
    codebooks = [main_token_normal]
    if trace is not None:
        trace["slow_logits"] = logits[0, -1].clone()
        trace["hidden"] = hidden_states[0, -1].clone()
        trace["fast_logits"] = []

    model.forward_generate_fast(hidden_states, torch.tensor([0], dtype=torch.long))  # output discarded
    a = codebooks[0] - c.semantic_begin_id
    a = torch.clamp(a, min=0, max=c.codebook_size - 1)
    hidden_states = model.fast_embeddings(a)
    codebooks.append(a)
    for codebook_idx in range(1, c.num_codebooks):
        fl = model.forward_generate_fast(hidden_states, torch.tensor([codebook_idx], dtype=torch.long))
        if trace is not None:
            trace["fast_logits"].append(fl[0, -1].clone())
        a = sample(fl, temperature, top_p, top_k, rand_like)[0]
        hidden_states = model.fast_embeddings(a)
        codebooks.append(a)
    return torch.stack(codebooks, dim=1).T


def decode_n_tokens(model: DualARRef, cur_token: Tensor, input_pos: Tensor, num_new_tokens: int,
                    temperature: Tensor, top_p: Tensor, top_k: int, semantic_logit_bias: Tensor,
                    rand_like: RandLike = torch.rand_like, traces: Optional[list] = None,
                    decode_one_token=decode_one_token_ar) -> Tensor:
    """inference.py:184-238 (RAS window roll, EOS break)."""
    c = model.config
    previous_tokens = torch.zeros((c.num_codebooks + 1, RAS_WIN_SIZE), dtype=torch.int)
    new_tokens = []
    for _ in range(num_new_tokens):
        tr = {} if traces is not None else None
        with sdpa_kernel(SDPBackend.MATH):
            next_token = decode_one_token(
                model, cur_token, input_pos, temperature, top_p, top_k, semantic_logit_bias,
                previous_tokens=previous_tokens, rand_like=rand_like, trace=tr).clone()
        if traces is not None:
            traces.append(tr)
        input_pos += 1
        cur_token = next_token.view(1, c.num_codebooks + 1, -1)
        previous_tokens = previous_tokens.roll(-1, dims=1)
        previous_tokens[:, -1] = next_token.view(c.num_codebooks + 1, -1)[:, 0]
        new_tokens.append(next_token)
        if cur_token[0, 0, -1] == c.im_end_id:
            break
    if not new_tokens:
        return torch.zeros((c.num_codebooks + 1, 0), dtype=torch.int)
    return torch.cat(new_tokens, dim=1)


def make_semantic_logit_bias(cfg: DualARConfig, dtype=torch.bfloat16) -> Tensor:
    """inference.py:308-320."""
    b = torch.full((1, 1, cfg.vocab_size), float("-inf"), dtype=dtype)
    b[0, 0, cfg.semantic_begin_id: cfg.semantic_end_id + 1] = 0.0
    b[0, 0, cfg.im_end_id] = 0.0
    return b


@torch.inference_mode()
def generate(*, model: DualARRef, prompt: Tensor, max_new_tokens: int, temperature: float = 1.0,
             top_p: float = 0.9, top_k: int = 30, rand_like: RandLike = torch.rand_like,
             traces: Optional[list] = None) -> Tensor:
    """inference.py:243-359."""
    c = model.config
    T = prompt.size(1)
    if T >= c.max_seq_len:
        raise ValueError(f"Input sequence length {T} exceeds max_seq_len {c.max_seq_len}")
    if max_new_tokens:
        if T + max_new_tokens > c.max_seq_len:
            max_new_tokens = c.max_seq_len - T
    else:
        max_new_tokens = c.max_seq_len - T
    dtype = model.dtype
    model.setup_caches(1, c.max_seq_len, dtype)
    codebook_dim = 1 + c.num_codebooks
    input_pos = torch.arange(0, T, dtype=torch.long)
    seq = torch.empty((codebook_dim, c.max_seq_len), dtype=prompt.dtype)
    seq[:, :T] = prompt
    temperature_t = torch.tensor(temperature, dtype=dtype)
    top_p_t = torch.tensor(top_p, dtype=dtype)
    bias = make_semantic_logit_bias(c, dtype)
    tr0 = {} if traces is not None else None
    first_token = decode_one_token_ar(model, prompt.view(1, codebook_dim, -1), input_pos, temperature_t,
                                      top_p_t, top_k, bias, rand_like=rand_like, trace=tr0)
    if traces is not None:
        traces.append(tr0)
    seq[:, T: T + 1] = first_token
    input_pos = torch.tensor([T], dtype=torch.int)
    x = decode_n_tokens(model, first_token.view(1, codebook_dim, -1), input_pos, max_new_tokens - 1,
                        temperature_t, top_p_t, top_k, bias, rand_like=rand_like, traces=traces)
    seq = seq[:, : T + 1 + x.size(1)]
    seq[:, T + 1:] = x
    return seq


def make_prompt(cfg: DualARConfig, T: int, seed: int = 1, voice_span: int = 0) -> Tensor:
    """Synthetic prompt: row 0 random text ids below the semantic
    range, rows 1..C zero; optional trailing `voice_span` VQ frames whose row 0 is
    semantic_begin_id + code0 (content_sequence.py:206-209,318)."""
    g = torch.Generator().manual_seed(seed)
    C = cfg.num_codebooks
    p = torch.zeros((C + 1, T), dtype=torch.long)
    p[0] = torch.randint(0, cfg.semantic_begin_id, (T,), generator=g)
    if voice_span:
        s = T - voice_span
        codes = torch.randint(0, cfg.codebook_size, (C, voice_span), generator=g)
        p[1:, s:] = codes
        p[0, s:] = cfg.semantic_begin_id + codes[0]
    return p
