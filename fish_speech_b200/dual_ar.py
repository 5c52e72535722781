"""Host-side mirror of the reference's Dual-AR inference interface over the CUDA engine.

Drop-in surface, same names / argument meaning / error behaviour as
  fish_speech/models/text2semantic/llama.py      DualARTransformer (:659) inference methods
  fish_speech/models/text2semantic/inference.py  decode_one_token_ar (:96), decode_n_tokens (:184), generate (:243)

PyTorch is used for device memory and streams only; every arithmetic op on the path runs in
libfish_b200.so (csrc/*.cu) through the C ABI in include/fish_b200.h.  No CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass, fields
from typing import Dict, Iterator, List, Optional

import torch
from torch import Tensor

from . import _lib

RAS_WIN_SIZE = 10  # inference.py:49


def find_multiple(n: int, k: int) -> int:  # llama.py:21-24
    return n if n % k == 0 else n + k - (n % k)


@dataclass
class DualARConfig:
    """The DualARModelArgs fields (llama.py:28-123,155-193) the decode path reads."""

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
    scale_codebook_embeddings: bool = True
    norm_fastlayer_input: bool = True
    im_end_id: int = 7
    initializer_range: float = 0.02

    @classmethod
    def from_any(cls, cfg) -> "DualARConfig":
        """Accepts this class, a dict, or any object with the reference's attribute names."""
        if isinstance(cfg, cls):
            return cfg
        names = [f.name for f in fields(cls)]
        get = (lambda k: cfg.get(k)) if isinstance(cfg, dict) else (lambda k: getattr(cfg, k, None))
        kw = {k: get(k) for k in names if get(k) is not None}
        return cls(**kw)


def s2_pro_config(max_seq_len: int = 2048, vocab_size: int = 155776) -> DualARConfig:
    """S2-Pro shape (assumed): Qwen3-4B backbone + 4 fast layers, 10 codebooks x 4096."""
    return DualARConfig(vocab_size=vocab_size, n_layer=36, n_head=32, dim=2560, intermediate_size=9728,
                        n_local_heads=8, head_dim=128, rope_base=1e6, norm_eps=1e-6, max_seq_len=max_seq_len,
                        codebook_size=4096, num_codebooks=10, semantic_begin_id=vocab_size - 4096,
                        semantic_end_id=vocab_size - 1, n_fast_layer=4, im_end_id=151645)


def state_dict_shapes(c: "DualARConfig") -> Dict[str, tuple]:
    """name -> shape of every tensor the engine binds, named as the reference nn.Module registers them
    (llama.py:254-271 BaseTransformer, :659-705 DualARTransformer, :846-881 Attention, :978-986 FeedForward)."""
    d, hd = c.dim, c.head_dim
    qkv = (c.n_head + 2 * c.n_local_heads) * hd
    out: Dict[str, tuple] = {
        "embeddings.weight": (c.vocab_size, d),
        "codebook_embeddings.weight": (c.codebook_size * c.num_codebooks, d),
        "norm.weight": (d,),
        "fast_embeddings.weight": (c.codebook_size, d),
        "fast_norm.weight": (d,),
        "fast_output.weight": (c.codebook_size, d),
    }
    if not c.tie_word_embeddings:
        out["output.weight"] = (c.vocab_size, d)
    for pre, n in (("layers", c.n_layer), ("fast_layers", c.n_fast_layer)):
        for i in range(n):
            p = f"{pre}.{i}."
            out[p + "attention.wqkv.weight"] = (qkv, d)
            out[p + "attention.wo.weight"] = (d, c.n_head * hd)
            if c.attention_qk_norm:
                out[p + "attention.q_norm.weight"] = (hd,)
                out[p + "attention.k_norm.weight"] = (hd,)
            out[p + "feed_forward.w1.weight"] = (c.intermediate_size, d)
            out[p + "feed_forward.w3.weight"] = (c.intermediate_size, d)
            out[p + "feed_forward.w2.weight"] = (d, c.intermediate_size)
            out[p + "ffn_norm.weight"] = (d,)
            out[p + "attention_norm.weight"] = (d,)
    return out


def precompute_freqs_cis(seq_len: int, n_elem: int, base: float) -> Tensor:
    """RoPE table exactly as the reference stores it (llama.py:1003-1022): bf16 [seq, n_elem/2, 2]."""
    freqs = 1.0 / (base ** (torch.arange(0, n_elem, 2)[: (n_elem // 2)].float() / n_elem))
    t = torch.arange(seq_len)
    freqs = torch.outer(t, freqs)
    cis = torch.polar(torch.ones_like(freqs), freqs)
    return torch.stack([cis.real, cis.imag], dim=-1).to(torch.bfloat16)


@dataclass
class ForwardResult:  # llama.py BaseTransformerForwardResult
    logits: Tensor
    hidden_states: Tensor


class _Tokenizer:
    """The three ints the hot path reads from FishTokenizer (inference.py:207,308-320)."""

    def __init__(self, cfg: DualARConfig):
        self.semantic_begin_id, self.semantic_end_id, self._im_end = cfg.semantic_begin_id, cfg.semantic_end_id, cfg.im_end_id

    def get_token_id(self, _name: str) -> int:
        return self._im_end


def _ptr(t: Optional[Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


class DualARTransformerB200:
    """Engine-backed stand-in for DualARTransformer's inference API (llama.py:659-827)."""

    def __init__(self, config, state_dict: Dict[str, Tensor], device="cuda:0", max_rows: int = 1,
                 max_seq_len: Optional[int] = None):
        _lib.require_cuda()
        self.lib = _lib.load()
        self.config = DualARConfig.from_any(config)
        c = self.config
        self.device = torch.device(device)
        self.tokenizer = _Tokenizer(c)
        self.max_rows = int(max_rows)
        self._engine_max_seq = find_multiple(int(max_seq_len or c.max_seq_len), 8)
        self.max_seq_len = -1
        self.max_batch_size = -1
        self._cache_setup_done = False
        # prefix caching: tokens [C+1, n] whose KV the PROMPT PASS left in slot 0 at positions [0, n).  A
        # later prompt that starts with the same columns skips them (bit-identical to recomputing: a prompt row's result
        # does not depend on which rows share its chunk).  "all" also reuses positions written by decode frames (the growing
        # conversation of generate_long, inference.py:620-688): their KV came through the decode kernels, equal within the
        # logit tolerance rather than bit for bit.  False switches it off.
        self.prefix_cache = True
        self._kv_tokens: Optional[Tensor] = None
        self.prefix_hits = 0           # prompt columns skipped so far
        self._h = C.c_void_p()
        self._keep: List[Tensor] = []
        with torch.cuda.device(self.device):
            self._create(state_dict)

    # ------------------------------------------------------------------ construction
    def _err(self) -> str:
        return (self.lib.fsb_dual_ar_last_error(self._h) or b"").decode()

    def _ck(self, rc: int, exc=RuntimeError):
        if rc != 0:
            raise exc(f"fish_b200: {self._err()}")

    def _create(self, sd: Dict[str, Tensor]):
        c = self.config
        d = _lib.DualARDesc(
            dim=c.dim, n_layer=c.n_layer, n_fast_layer=c.n_fast_layer, n_head=c.n_head, n_kv_head=c.n_local_heads,
            head_dim=c.head_dim, ffn=c.intermediate_size, vocab=c.vocab_size, num_codebooks=c.num_codebooks,
            codebook_size=c.codebook_size, semantic_begin_id=c.semantic_begin_id, semantic_end_id=c.semantic_end_id,
            im_end_id=c.im_end_id, qk_norm=int(c.attention_qk_norm),
            scale_codebook_embeddings=int(c.scale_codebook_embeddings), norm_fastlayer_input=int(c.norm_fastlayer_input),
            tie_word_embeddings=int(c.tie_word_embeddings), max_rows=self.max_rows, norm_eps=c.norm_eps)
        self._desc = d
        nbytes = self.lib.fsb_dual_ar_workspace_bytes(C.byref(d), self._engine_max_seq)
        self._workspace = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        rc = self.lib.fsb_dual_ar_create(C.byref(d), self._engine_max_seq, self._workspace.data_ptr(), nbytes,
                                         C.byref(self._h))
        self._ck(rc)
        # weights: bf16, contiguous, device-resident; names are the reference's state-dict keys
        self.weights: Dict[str, Tensor] = {}
        for name, w in sd.items():
            if name.endswith(("freqs_cis", "causal_mask")):
                continue
            t = w.detach().to(device=self.device, dtype=torch.bfloat16).contiguous()
            self.weights[name] = t
            rows, cols = (t.shape[0], t.shape[1]) if t.dim() == 2 else (t.shape[0], 1)
            self._ck(self.lib.fsb_dual_ar_bind_weight(self._h, name.encode(), t.data_ptr(), rows, cols))
        self._rope = precompute_freqs_cis(self._engine_max_seq, c.head_dim, c.rope_base).to(self.device).contiguous()
        self._fast_rope = precompute_freqs_cis(c.num_codebooks, c.head_dim, c.rope_base).to(self.device).contiguous()
        self._ck(self.lib.fsb_dual_ar_bind_rope(self._h, 0, self._rope.data_ptr(), self._rope.shape[0]))
        self._ck(self.lib.fsb_dual_ar_bind_rope(self._h, 1, self._fast_rope.data_ptr(), self._fast_rope.shape[0]))
        self._session: Optional[DecodeSession] = None

    def __del__(self):
        try:
            if self._h:
                self.lib.fsb_dual_ar_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def set_option(self, name: str, value: int):
        self._ck(self.lib.fsb_dual_ar_set_option(self._h, name.encode(), int(value)))

    # ------------------------------------------------------------------ reference API
    def parameters(self) -> Iterator[Tensor]:
        return iter(self.weights.values())

    def setup_caches(self, max_batch_size: int, max_seq_len: int, dtype: torch.dtype = torch.bfloat16):
        """llama.py:707-721 / :307-324: allocate [B, KV, S, hd] K/V per layer (torch-owned), bind to the engine."""
        if self.max_seq_len >= max_seq_len and self.max_batch_size >= max_batch_size:
            return
        if dtype != torch.bfloat16:
            raise ValueError("fish_b200 engine is bf16 only")
        c = self.config
        max_seq_len = find_multiple(max_seq_len, 8)
        if max_seq_len > self._engine_max_seq:
            raise ValueError(f"max_seq_len {max_seq_len} exceeds the engine's {self._engine_max_seq}")
        if max_batch_size > self.max_rows:
            raise ValueError(f"max_batch_size {max_batch_size} exceeds the engine's max_rows {self.max_rows}")
        self.max_seq_len, self.max_batch_size = max_seq_len, max_batch_size
        shp = (max_batch_size, c.n_local_heads, max_seq_len, c.head_dim)
        fshp = (max_batch_size, c.n_local_heads, c.num_codebooks, c.head_dim)
        self.kv_cache = [(torch.zeros(shp, dtype=dtype, device=self.device), torch.zeros(shp, dtype=dtype, device=self.device))
                         for _ in range(c.n_layer)]
        self.fast_kv_cache = [(torch.zeros(fshp, dtype=dtype, device=self.device),
                               torch.zeros(fshp, dtype=dtype, device=self.device)) for _ in range(c.n_fast_layer)]
        for i, (k, v) in enumerate(self.kv_cache):
            self._ck(self.lib.fsb_dual_ar_bind_kv(self._h, 0, i, k.data_ptr(), v.data_ptr(), max_batch_size, max_seq_len))
        for i, (k, v) in enumerate(self.fast_kv_cache):
            self._ck(self.lib.fsb_dual_ar_bind_kv(self._h, 1, i, k.data_ptr(), v.data_ptr(), max_batch_size,
                                                  c.num_codebooks))
        self._cache_setup_done = True
        self._session = None
        self._kv_tokens = None

    def _buffer(self, name: str, dtype=torch.bfloat16) -> Tensor:
        p, n = C.c_void_p(), C.c_size_t()
        self._ck(self.lib.fsb_dual_ar_buffer(self._h, name.encode(), C.byref(p), C.byref(n)))
        return _view(self._workspace, p.value, n.value, dtype)

    def prefill(self, prompt_cols: Tensor, pos0: int = 0, slot: int = 0):
        """Slow-stack pass over prompt columns [C+1, T] (any int dtype): fills KV, produces no logits."""
        T = prompt_cols.shape[-1]
        if slot == 0:
            self._kv_tokens = None   # slot 0 is rewritten by a caller that does not go through prefill_cached
        if T == 0:
            return
        tok = prompt_cols.reshape(self.config.num_codebooks + 1, T).t().to(device=self.device, dtype=torch.int32).contiguous()
        self._keep = [tok]
        self._ck(self.lib.fsb_dual_ar_prefill(self._h, tok.data_ptr(), T, slot, pos0, _stream()), ValueError)

    def prefill_cached(self, prompt_cols: Tensor) -> int:
        """Prompt pass into slot 0 from position 0 that skips the leading columns whose KV is already there (see
        `prefix_cache`); returns how many were skipped."""
        C1 = self.config.num_codebooks + 1
        cols = prompt_cols.reshape(C1, -1).to(device=self.device, dtype=torch.int32)
        T = cols.shape[1]
        have = self._kv_tokens
        L = 0
        if self.prefix_cache and have is not None and T > 0:
            m = min(T, have.shape[1])
            same = (cols[:, :m] == have[:, :m]).all(0)
            L = m if bool(same.all()) else int((~same).nonzero()[0])
        self.prefill(cols[:, L:], L, 0)
        self._kv_tokens = cols.clone() if self.prefix_cache else None
        self.prefix_hits += L
        return L

    def copy_kv_prefix(self, src_slot: int, dst_slot: int, n: int):
        """KV of positions [0, n) of one slot copied to another (all slow layers): a request that starts with the columns of
        a resident one (the same reference voice, the same system prompt) takes its keys and values instead of recomputing
        them; 147 KB per token at the S2-Pro shape against a prompt pass over n tokens."""
        if n <= 0 or src_slot == dst_slot:
            return
        if dst_slot == 0:
            self._kv_tokens = None
        for k, v in self.kv_cache:
            k[dst_slot, :, :n].copy_(k[src_slot, :, :n])
            v[dst_slot, :, :n].copy_(v[src_slot, :, :n])

    def _kv_written_from(self, pos: int):
        """A decode frame writes slot 0 at `pos` and after: the prompt-pass record ends there (mode "all" keeps what generate()
        recorded for the frames it fed)."""
        if self._kv_tokens is not None and self._kv_tokens.shape[1] > pos:
            self._kv_tokens = self._kv_tokens[:, :pos]

    def forward_generate(self, inp: Tensor, input_pos: Optional[Tensor] = None, audio_masks=None, audio_parts=None,
                         return_all: bool = False) -> ForwardResult:
        """llama.py:390-466 + :818-827. inp [B, C+1, Tq]; Tq > 1 (prompt) supports B = 1 like the reference."""
        if audio_parts is not None:
            raise NotImplementedError("audio_parts needs an audio_projector the reference does not define (llama.py:423-433)")
        if return_all:
            raise NotImplementedError("return_all is a training/reward path, not on the decode path")
        B, C1, Tq = inp.shape
        if input_pos is None:
            input_pos = torch.arange(Tq, device=self.device)
        if int(input_pos[-1]) >= self.max_seq_len or int(input_pos[0]) < 0:
            raise ValueError(f"input_pos {int(input_pos[-1])} is outside the KV cache (max_seq_len {self.max_seq_len})")
        if Tq > 1:
            if B != 1:
                raise ValueError("prompt pass supports batch 1 (reference: inference.py:284-288)")
            self.prefill(inp[0, :, :-1], int(input_pos[0]), 0)
        self._kv_written_from(int(input_pos[-1]))
        tok = inp[:, :, -1].to(device=self.device, dtype=torch.int32).contiguous()
        pos = input_pos[-1:].to(device=self.device, dtype=torch.int32).expand(B).contiguous()
        self._ck(self.lib.fsb_dual_ar_slow_step(self._h, B, tok.data_ptr(), pos.data_ptr(), None, _stream()))
        V, D = self.config.vocab_size, self.config.dim
        logits = self._buffer("slow_logits")[: B * V].view(B, 1, V).clone()
        hidden = self._buffer("hidden")[: B * D].view(B, 1, D).clone()
        return ForwardResult(logits=logits, hidden_states=hidden)

    def forward_generate_fast(self, x: Tensor, input_pos: Optional[Tensor] = None) -> Tensor:
        """llama.py:798-816. x [B, 1, dim] bf16, input_pos [1] = codebook index."""
        B = x.shape[0]
        xi = x.reshape(B, -1).to(device=self.device, dtype=torch.bfloat16).contiguous()
        idx = int(input_pos[0])
        self._ck(self.lib.fsb_dual_ar_fast_step(self._h, B, xi.data_ptr(), idx, 1, _stream()))
        cb = self.config.codebook_size
        fl = self._buffer("fast_logits")
        R = fl.numel() // (self.config.num_codebooks * cb)
        return fl.view(self.config.num_codebooks, R, cb)[idx, :B].reshape(B, 1, cb).clone()

    def fast_embeddings(self, a: Tensor) -> Tensor:
        """nn.Embedding lookup (llama.py:693): pure gather, done by torch indexing (host-side glue; the fused
        frame gathers it inside the sampling kernel)."""
        return self.weights["fast_embeddings.weight"][a.to(self.device).long()]

    # ------------------------------------------------------------------ engine extras
    def frame_bytes(self, batch: int, ctx_sum: int, masked_head: bool = False) -> float:
        return float(self.lib.fsb_dual_ar_frame_bytes(self._h, batch, int(ctx_sum), int(masked_head)))

    def launches_per_frame(self, batch: int) -> int:
        return int(self.lib.fsb_dual_ar_launches_per_frame(self._h, batch))

    def session(self, batch: int, max_frames: int) -> "DecodeSession":
        s = self._session
        if s is None or s.batch != batch or s.max_frames < max_frames:
            s = DecodeSession(self, batch, max_frames)
            self._session = s
        return s


def _view(ws: Tensor, ptr: int, nbytes: int, dtype: torch.dtype) -> Tensor:
    """Zero-copy typed view of a sub-range of the (torch-owned) workspace the engine carved up."""
    off = ptr - ws.data_ptr()
    assert 0 <= off and off + nbytes <= ws.numel()
    return ws[off: off + nbytes].view(dtype)


class DecodeSession:
    """Device-resident state of the fused frame loop (decode_n_tokens bookkeeping, inference.py:184-238)."""

    def __init__(self, model: DualARTransformerB200, batch: int, max_frames: int):
        c, dev = model.config, model.device
        self.model, self.batch, self.max_frames = model, batch, max_frames
        C1 = c.num_codebooks + 1
        self.cur_tokens = torch.zeros((batch, C1), dtype=torch.int32, device=dev)
        self.pos = torch.zeros((batch,), dtype=torch.int32, device=dev)
        self.ras_window = torch.zeros((batch, RAS_WIN_SIZE), dtype=torch.int32, device=dev)
        self.out_tokens = torch.zeros((max_frames, batch, C1), dtype=torch.int32, device=dev)
        self.frame_counter = torch.zeros((1,), dtype=torch.int32, device=dev)
        self.done = torch.zeros((batch,), dtype=torch.int32, device=dev)
        self.uniforms: Optional[Tensor] = None
        self.bias: Optional[Tensor] = None
        self.seed = 0
        self.restrict_scan = True
        # per-request sampling parameters (fsb_frame_args.row_*): float32 / float32 / int32 device tensors [batch] that replace
        # the scalars of run() for each request's normal draw, or None.  Values as generate() would round them (bf16).
        self.row_temperature: Optional[Tensor] = None
        self.row_top_p: Optional[Tensor] = None
        self.row_top_k: Optional[Tensor] = None

    def use_row_parameters(self):
        """Allocates the per-request parameter arrays (filled with generate()'s defaults); the caller writes rows."""
        dev = self.model.device
        self.row_temperature = torch.full((self.batch,), _as_bf16_float(1.0), dtype=torch.float32, device=dev)
        self.row_top_p = torch.full((self.batch,), _as_bf16_float(0.9), dtype=torch.float32, device=dev)
        self.row_top_k = torch.full((self.batch,), 30, dtype=torch.int32, device=dev)

    def args(self, temperature: float, top_p: float, top_k: int, use_ras: bool) -> _lib.FrameArgs:
        return _lib.FrameArgs(
            batch=self.batch, cur_tokens=self.cur_tokens.data_ptr(), pos=self.pos.data_ptr(), slots=None,
            ras_window=self.ras_window.data_ptr() if use_ras else None, out_tokens=self.out_tokens.data_ptr(),
            max_frames=self.max_frames, frame_counter=self.frame_counter.data_ptr(), done=self.done.data_ptr(),
            logit_bias=_ptr(self.bias), restrict_scan=int(self.restrict_scan and self.bias is not None),
            temperature=temperature, top_p=top_p, top_k=top_k, uniforms=_ptr(self.uniforms), seed=self.seed,
            row_temperature=_ptr(self.row_temperature), row_top_p=_ptr(self.row_top_p), row_top_k=_ptr(self.row_top_k))

    def run(self, n_frames: int, temperature: float, top_p: float, top_k: int, use_ras: bool,
            pos_bound: Optional[int] = None):
        """n_frames fused frames.  The reference's KVCache.update raises an index error past the end of the cache
        (llama.py:205-214); here the positions are checked on the host BEFORE the launch: `pos_bound` = the largest
        current position if the caller knows it, else it is read from the device (one small synchronising copy)."""
        m = self.model
        hi = int(self.pos.max().item()) if pos_bound is None else int(pos_bound)
        if hi + n_frames > m.max_seq_len:
            raise ValueError(f"decode would write position {hi + n_frames - 1} past the KV cache (max_seq_len {m.max_seq_len})")
        a = self.args(temperature, top_p, top_k, use_ras)
        m._ck(m.lib.fsb_dual_ar_decode_frames(m._h, C.byref(a), n_frames, _stream()))

    def check_device(self):
        """Raises if a device-side watchdog expired since the last check (synchronises the stream)."""
        m = self.model
        code = C.c_int32(0)
        m._ck(m.lib.fsb_dual_ar_device_error(m._h, C.byref(code), _stream()))


def _as_bf16_float(v) -> float:
    """generate() builds temperature / top_p as tensors in the MODEL dtype (inference.py:301-306)."""
    if isinstance(v, Tensor):
        return float(v.detach().to(torch.bfloat16).float())
    return float(torch.tensor(float(v), dtype=torch.bfloat16).float())


def make_semantic_logit_bias(model: DualARTransformerB200) -> Tensor:
    """inference.py:308-320.  Extension: `model.ignore_eos = True` leaves <|im_end|> masked (-inf) so that a run decodes
    exactly max_new_tokens frames (fixed-length benchmarking); default off = the reference's bias."""
    c = model.config
    b = torch.full((1, 1, c.vocab_size), float("-inf"), dtype=torch.bfloat16, device=model.device)
    b[0, 0, c.semantic_begin_id: c.semantic_end_id + 1] = 0.0
    if not getattr(model, "ignore_eos", False):
        b[0, 0, model.tokenizer.get_token_id("<|im_end|>")] = 0.0
    return b


def _bias_is_standard(model: DualARTransformerB200, bias: Tensor) -> bool:
    """True when `bias` has the support of the constrained-decode bias of inference.py:308-320: every entry outside
    [semantic_begin, semantic_end] and <|im_end|> is -inf.  Then no token outside can ever be drawn, the sampler may scan
    only that support and the fused frame may compute only those LM-head rows (the values INSIDE the support are free:
    the sampler adds them itself).  Cached per tensor content (data_ptr, in-place version counter); tensors built under
    inference_mode carry no version counter and are checked every time."""
    def check() -> bool:
        outside = getattr(model, "_bias_outside", None)
        if outside is None:
            c = model.config
            outside = torch.ones(c.vocab_size, dtype=torch.bool, device=model.device)
            outside[c.semantic_begin_id: c.semantic_end_id + 1] = False
            outside[model.tokenizer.get_token_id("<|im_end|>")] = False
            model._bias_outside = outside
        flat = bias.reshape(-1).to(model.device)
        return flat.numel() == outside.numel() and bool(torch.isneginf(flat[outside]).all())

    if bias.is_inference():
        return check()
    key = "_fsb_standard"
    stamp = (bias.data_ptr(), bias._version, tuple(bias.shape))
    cached = getattr(bias, key, None)
    if cached is None or cached[0] != stamp:  # re-checked when the tensor was mutated in place or re-allocated
        cached = (stamp, check())
        setattr(bias, key, cached)
    return cached[1]


def decode_one_token_ar(model: DualARTransformerB200, x: Tensor, input_pos: Tensor, temperature, top_p, top_k: int,
                        semantic_logit_bias: Tensor, audio_masks=None, audio_parts=None,
                        previous_tokens: Optional[Tensor] = None) -> Tensor:
    """inference.py:96-181 on the engine: prompt pass (if Tq > 1) + ONE fused frame.  Returns (C+1, 1) int32."""
    if audio_parts is not None:
        raise NotImplementedError("audio_parts is not on the decode path (llama.py:423-433)")
    c = model.config
    C1 = c.num_codebooks + 1
    x = x.reshape(1, C1, -1)
    Tq = x.shape[-1]
    p_last = int(input_pos.reshape(-1)[-1])
    if p_last >= model.max_seq_len or p_last - (Tq - 1) < 0:
        raise ValueError(f"input_pos {p_last} is outside the KV cache (max_seq_len {model.max_seq_len})")
    if Tq > 1:
        if int(input_pos[0]) == 0:
            model.prefill_cached(x[0, :, :-1])
        else:
            model.prefill(x[0, :, :-1], int(input_pos[0]), 0)
    model._kv_written_from(p_last)
    s = model.session(1, getattr(model, "_session_frames", 4096))
    if Tq > 1:  # a new generation starts: frame records (injected uniforms / Philox counters) restart at 0
        s.frame_counter.zero_()
    s.cur_tokens.copy_(x[0, :, -1].to(torch.int32).view(1, C1))
    s.pos.fill_(p_last)
    use_ras = previous_tokens is not None
    if use_ras:
        s.ras_window.copy_(previous_tokens[0].to(torch.int32).view(1, RAS_WIN_SIZE))
    s.bias = semantic_logit_bias.reshape(-1).to(device=model.device, dtype=torch.bfloat16).contiguous()
    s.restrict_scan = _bias_is_standard(model, semantic_logit_bias)
    s.run(1, _as_bf16_float(temperature), _as_bf16_float(top_p), int(top_k), use_ras, pos_bound=p_last)
    return s.cur_tokens.view(C1, 1).clone()


def decode_n_tokens(model: DualARTransformerB200, cur_token: Tensor, input_pos: Tensor, num_new_tokens: int,
                    temperature, top_p, top_k: int, semantic_logit_bias: Tensor, audio_masks=None, audio_parts=None,
                    decode_one_token=None, eos_poll: int = 32) -> Tensor:
    """inference.py:184-238 with the frame loop on the device: frames run back to back from one CUDA graph;
    the <|im_end|> test (inference.py:233, a device->host sync per frame in the reference) becomes a device flag
    polled every `eos_poll` frames.  Output is trimmed at the first <|im_end|> frame, so it equals the reference's."""
    c = model.config
    C1 = c.num_codebooks + 1
    if num_new_tokens <= 0:
        return torch.zeros((C1, 0), dtype=torch.int32, device=model.device)
    p0 = int(input_pos.reshape(-1)[-1])
    if p0 < 0 or p0 + num_new_tokens > model.max_seq_len:  # the reference fails in KVCache.update (llama.py:205-214)
        raise ValueError(f"decoding {num_new_tokens} frames from position {p0} exceeds the KV cache "
                         f"(max_seq_len {model.max_seq_len})")
    model._kv_written_from(p0)
    s = model.session(1, max(num_new_tokens, getattr(model, "_session_frames", 0)))
    s.cur_tokens.copy_(cur_token.reshape(C1).to(torch.int32).view(1, C1))
    s.pos.fill_(p0)
    s.ras_window.zero_()  # previous_tokens = zeros((C+1, RAS_WIN_SIZE)) (inference.py:199-203)
    s.done.zero_()
    f0 = int(s.frame_counter.item())
    if f0 + num_new_tokens > s.max_frames:
        s.frame_counter.zero_()
        f0 = 0
    s.bias = semantic_logit_bias.reshape(-1).to(device=model.device, dtype=torch.bfloat16).contiguous()
    s.restrict_scan = _bias_is_standard(model, semantic_logit_bias)
    T, P, K = _as_bf16_float(temperature), _as_bf16_float(top_p), int(top_k)
    n = 0
    while n < num_new_tokens:
        step = min(eos_poll, num_new_tokens - n)
        s.run(step, T, P, K, True, pos_bound=p0 + n)
        n += step
        if bool(s.done.any().item()):
            break
    s.check_device()
    out = s.out_tokens[f0: f0 + n, 0, :]  # [n, C+1]
    hit = (out[:, 0] == c.im_end_id).nonzero()
    if hit.numel() > 0:
        out = out[: int(hit[0]) + 1]
    return out.t().contiguous().clone()


@torch.inference_mode()
def generate(*, model: DualARTransformerB200, prompt: Tensor, max_new_tokens: int, audio_masks=None, audio_parts=None,
             decode_one_token=decode_one_token_ar, num_samples: int = 1, **sampling_kwargs) -> Tensor:
    """inference.py:243-359: same checks, same output layout (C+1, T + N) in prompt.dtype."""
    c = model.config
    T = prompt.size(1)
    prompt = prompt[None].repeat(num_samples, 1, 1)
    if T >= c.max_seq_len:
        raise ValueError(f"Input sequence length {T} exceeds max_seq_len {c.max_seq_len}")
    if max_new_tokens:
        if T + max_new_tokens > c.max_seq_len:
            max_new_tokens = c.max_seq_len - T
        T_new = T + max_new_tokens
    else:
        T_new = c.max_seq_len
        max_new_tokens = T_new - T
    device = model.device
    if not model._cache_setup_done:
        model.setup_caches(max_batch_size=1, max_seq_len=c.max_seq_len, dtype=torch.bfloat16)
    codebook_dim = 1 + c.num_codebooks
    input_pos = torch.arange(0, T, device=device, dtype=torch.long)
    seq = torch.empty((codebook_dim, c.max_seq_len), dtype=prompt.dtype, device=device)
    seq[:, :T] = prompt[0].to(device)
    temperature = sampling_kwargs.get("temperature", 1.0)
    top_p = sampling_kwargs.get("top_p", 0.9)
    top_k = sampling_kwargs.get("top_k", 30)
    bias = make_semantic_logit_bias(model)
    model._session_frames = max(max_new_tokens + 1, getattr(model, "_session_frames", 0))
    sess = model.session(1, model._session_frames)
    if sess.uniforms is None:  # Philox stream keyed by torch's CPU generator (reference: utils.set_seed + rand_like)
        sess.seed = int(torch.randint(0, 2 ** 62, (1,)).item())
    first = decode_one_token(model, prompt[:1].to(device).view(1, codebook_dim, -1), input_pos, temperature, top_p,
                             top_k, bias, audio_masks, audio_parts)
    seq[:, T: T + 1] = first.to(seq.dtype)
    input_pos = torch.tensor([T], device=device, dtype=torch.int)
    x = decode_n_tokens(model, first.view(1, codebook_dim, -1), input_pos, max_new_tokens - 1, temperature, top_p,
                        top_k, bias, audio_masks, audio_parts, decode_one_token=decode_one_token)
    seq = seq[:, : T + 1 + x.size(1)]
    seq[:, T + 1:] = x.to(seq.dtype)
    if model.prefix_cache == "all" and decode_one_token is decode_one_token_ar:
        # every column but the last one produced was fed as an input frame: its KV sits at its position in slot 0
        model._kv_tokens = seq[:, : seq.shape[1] - 1].to(torch.int32).clone()
    return seq


def init_model_from_state_dict(config, state_dict: Dict[str, Tensor], device="cuda:0", max_rows: int = 1,
                               max_seq_len: Optional[int] = None):
    """Counterpart of init_model (inference.py:362-392) for an in-memory state dict: returns
    (model, decode_one_token) so callers keep passing the callable through generate()/generate_long()."""
    model = DualARTransformerB200(config, state_dict, device=device, max_rows=max_rows, max_seq_len=max_seq_len)
    return model, decode_one_token_ar
