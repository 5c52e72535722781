"""Host-side mirror of the reference's DAC codec decode interface over the CUDA engine.

Drop-in surface: an object with `from_indices(LongTensor[B, 10, T]) -> Tensor[B, 1, 2048*T]`,
`sample_rate`, `device`, `quantizer.decode`-free (the whole of DAC.from_indices, modded_dac.py:925-927, runs in
libfish_b200.so).  Host work done here, once, at construction: fold weight norm (g * v / ||v||, recomputed on every
forward by the reference, modded_dac.py:553-556) and repack each conv / transposed conv / linear weight into the
engine's implicit-GEMM layout [taps][Cin][N] (csrc/dac_kernels.cuh).  No CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass, fields
from typing import Dict, Optional, Tuple

import torch
from torch import Tensor

from . import _lib


@dataclass
class DACConfig:
    """fish_speech/configs/modded_dac_vq.yaml values the decode path uses."""

    sample_rate: int = 44100
    latent_dim: int = 1024
    decoder_dim: int = 1536
    decoder_rates: Tuple[int, ...] = (8, 8, 4, 2)
    n_codebooks: int = 9
    codebook_size: int = 1024
    semantic_codebook_size: int = 4096
    codebook_dim: int = 8
    upsample_factors: Tuple[int, ...] = (2, 2)
    tf_layers: int = 8
    tf_heads: int = 16
    tf_dim: int = 1024
    tf_ffn: int = 3072
    tf_head_dim: int = 64
    tf_rope_base: float = 10000.0
    tf_norm_eps: float = 1e-5
    tf_window: int = 128
    convnext_ratio: int = 4
    # encode side (DAC.encode): Encoder(encoder_dim, encoder_rates), window transformer after the blocks named in
    # encoder_tf_layers (heads = dim // 64, ffn = 3 * dim, window 512: modded_dac.py:636-650)
    encoder_dim: int = 64
    encoder_rates: Tuple[int, ...] = (2, 4, 8, 8)
    encoder_tf_layers: Tuple[int, ...] = (0, 0, 0, 4)
    encoder_tf_window: int = 512
    encoder_tf_head_dim: int = 64

    @property
    def hop(self) -> int:
        return int(math.prod(self.upsample_factors) * math.prod(self.decoder_rates))

    @classmethod
    def from_any(cls, cfg) -> "DACConfig":
        if isinstance(cfg, cls):
            return cfg
        names = [f.name for f in fields(cls)]
        get = (lambda k: cfg.get(k)) if isinstance(cfg, dict) else (lambda k: getattr(cfg, k, None))
        kw = {k: get(k) for k in names if get(k) is not None}
        for k in ("decoder_rates", "upsample_factors", "encoder_rates", "encoder_tf_layers"):
            if k in kw:
                kw[k] = tuple(kw[k])
        return cls(**kw)


def _weight_norm(v: Tensor, g: Tensor) -> Tensor:
    """torch weight_norm(dim=0): g * v / ||v|| with the norm over every dim but 0."""
    n = v.flatten(1).norm(dim=1).view(-1, *([1] * (v.dim() - 1)))
    return v * (g / n)


def _rope_table(seq_len: int, n_elem: int, base: float) -> Tensor:
    """modded_dac.py:442-452: table rounded to bf16 even when the codec runs fp32; stored as fp32 for the engine."""
    freqs = 1.0 / (base ** (torch.arange(0, n_elem, 2)[: (n_elem // 2)].float() / n_elem))
    t = torch.arange(seq_len)
    freqs = torch.outer(t, freqs)
    cis = torch.polar(torch.ones_like(freqs), freqs)
    return torch.stack([cis.real, cis.imag], dim=-1).to(torch.bfloat16).float()


def pack_weights(c: DACConfig, sd: Dict[str, Tensor], max_frames: int) -> Dict[str, Tensor]:
    """Reference state dict (names of modded_dac.DAC / rvq.DownsampleResidualVectorQuantize) -> engine tensors."""
    f32 = lambda t: t.detach().float()
    out: Dict[str, Tensor] = {}

    def wn_new(p):  # parametrizations.weight_norm (modded convs)
        return _weight_norm(f32(sd[p + "conv.parametrizations.weight.original1"]),
                            f32(sd[p + "conv.parametrizations.weight.original0"]))

    conv_pack = lambda w: w.permute(2, 1, 0).contiguous()           # [Cout,Cin,k] -> [k][Cin][Cout]

    def convt_pack(w, s):                                           # [Cin,Cout,k] -> [k/s][Cin][s*Cout]
        cin, cout, k = w.shape
        # kernel tap j reads x[q - (taps-1-j)]: x[q] pairs with w[..., r], x[q-1] with w[..., r+s] -> flip the tap axis
        return w.view(cin, cout, k // s, s).permute(2, 0, 3, 1).reshape(k // s, cin, s * cout).flip(0).contiguous()

    # ---- RVQ (descript from_codes): codebooks + weight-normed 1x1 out_proj
    prefixes = ["quantizer.semantic_quantizer.quantizers.0."] + [f"quantizer.quantizer.quantizers.{i}." for i in range(c.n_codebooks)]
    out["rvq.codebooks"] = torch.cat([f32(sd[p + "codebook.weight"]) for p in prefixes], 0).contiguous()
    out["rvq.proj_w"] = torch.stack([_weight_norm(f32(sd[p + "out_proj.weight_v"]), f32(sd[p + "out_proj.weight_g"]))[:, :, 0]
                                     for p in prefixes]).contiguous()
    out["rvq.proj_b"] = torch.stack([f32(sd[p + "out_proj.bias"]) for p in prefixes]).contiguous()
    # ---- post_module transformer
    out["tf.rope"] = _rope_table(max_frames, c.tf_head_dim, c.tf_rope_base).contiguous()
    for i in range(c.tf_layers):
        p, q = f"quantizer.post_module.layers.{i}.", f"tf.{i}."
        out[q + "attn_norm"] = f32(sd[p + "attention_norm.weight"])
        out[q + "ffn_norm"] = f32(sd[p + "ffn_norm.weight"])
        out[q + "gamma1"] = f32(sd[p + "attention_layer_scale.gamma"])
        out[q + "gamma2"] = f32(sd[p + "ffn_layer_scale.gamma"])
        out[q + "wqkv"] = f32(sd[p + "attention.wqkv.weight"]).t().contiguous()
        out[q + "wo"] = f32(sd[p + "attention.wo.weight"]).t().contiguous()
        out[q + "w13"] = torch.cat([f32(sd[p + "feed_forward.w1.weight"]), f32(sd[p + "feed_forward.w3.weight"])], 0).t().contiguous()
        out[q + "w2"] = f32(sd[p + "feed_forward.w2.weight"]).t().contiguous()
    out["tf.norm"] = f32(sd["quantizer.post_module.norm.weight"])
    # ---- upsample stages
    for i, f in enumerate(c.upsample_factors):
        p, q = f"quantizer.upsample.{i}.", f"up.{i}."
        out[q + "ct_w"] = convt_pack(f32(sd[p + "0.conv.weight"]), f)
        out[q + "ct_b"] = f32(sd[p + "0.conv.bias"])
        out[q + "dw_w"] = f32(sd[p + "1.dwconv.conv.weight"])[:, 0, :].contiguous()
        out[q + "dw_b"] = f32(sd[p + "1.dwconv.conv.bias"])
        out[q + "ln_w"], out[q + "ln_b"] = f32(sd[p + "1.norm.weight"]), f32(sd[p + "1.norm.bias"])
        out[q + "pw1_w"], out[q + "pw1_b"] = f32(sd[p + "1.pwconv1.weight"]).t().contiguous(), f32(sd[p + "1.pwconv1.bias"])
        out[q + "pw2_w"], out[q + "pw2_b"] = f32(sd[p + "1.pwconv2.weight"]).t().contiguous(), f32(sd[p + "1.pwconv2.bias"])
        out[q + "gamma"] = f32(sd[p + "1.gamma"])
    # ---- conv decoder
    out["dec.in_w"], out["dec.in_b"] = conv_pack(wn_new("decoder.model.0.")), f32(sd["decoder.model.0.conv.bias"])
    for i, s in enumerate(c.decoder_rates):
        p, q = f"decoder.model.{i + 1}.block.", f"dec.{i}."
        out[q + "alpha"] = f32(sd[p + "0.alpha"]).flatten()
        out[q + "ct_w"], out[q + "ct_b"] = convt_pack(wn_new(p + "1."), s), f32(sd[p + "1.conv.bias"])
        for j in range(3):
            u, v = p + f"{2 + j}.block.", q + f"r{j}."
            out[v + "a1"], out[v + "a2"] = f32(sd[u + "0.alpha"]).flatten(), f32(sd[u + "2.alpha"]).flatten()
            out[v + "w1"], out[v + "b1"] = conv_pack(wn_new(u + "1.")), f32(sd[u + "1.conv.bias"])
            out[v + "w2"], out[v + "b2"] = conv_pack(wn_new(u + "3.")), f32(sd[u + "3.conv.bias"])
    n = len(c.decoder_rates)
    out["dec.out_alpha"] = f32(sd[f"decoder.model.{n + 1}.alpha"]).flatten()
    out["dec.out_w"], out["dec.out_b"] = conv_pack(wn_new(f"decoder.model.{n + 2}.")), f32(sd[f"decoder.model.{n + 2}.conv.bias"])
    return out


def has_encode_side(sd: Dict[str, Tensor]) -> bool:
    return "encoder.block.0.conv.bias" in sd and "quantizer.semantic_quantizer.quantizers.0.in_proj.bias" in sd


def pack_encode_weights(c: DACConfig, sd: Dict[str, Tensor], max_positions: int) -> Dict[str, Tensor]:
    """Encode-side tensors of the reference state dict (Encoder, quantizer.downsample, quantizer.pre_module, the codebooks'
    in_proj) -> engine tensors, fp32.  Convs as [taps][Cin][N]; a strided conv (kernel k = m * s, stride s) is regrouped for
    the [len / s][s * Cin] view of its channels-last input: tap j of m, row r * Cin + c  <-  w[n][c][j * s + r]."""
    f32 = lambda t: t.detach().float()
    out: Dict[str, Tensor] = {}

    def wn_new(p):
        return _weight_norm(f32(sd[p + "conv.parametrizations.weight.original1"]),
                            f32(sd[p + "conv.parametrizations.weight.original0"]))

    conv_pack = lambda w: w.permute(2, 1, 0).contiguous()           # [Cout,Cin,k] -> [k][Cin][Cout]

    def strided_pack(w, s):                                         # [Cout,Cin,m*s] -> [m][s*Cin][Cout]
        cout, cin, k = w.shape
        return w.view(cout, cin, k // s, s).permute(2, 3, 1, 0).reshape(k // s, s * cin, cout).contiguous()

    def transformer(p, q, layers):
        for i in range(layers):
            u, v = f"{p}layers.{i}.", f"{q}{i}."
            out[v + "attn_norm"] = f32(sd[u + "attention_norm.weight"])
            out[v + "ffn_norm"] = f32(sd[u + "ffn_norm.weight"])
            out[v + "gamma1"] = f32(sd[u + "attention_layer_scale.gamma"])
            out[v + "gamma2"] = f32(sd[u + "ffn_layer_scale.gamma"])
            out[v + "wqkv"] = f32(sd[u + "attention.wqkv.weight"]).t().contiguous()
            out[v + "wo"] = f32(sd[u + "attention.wo.weight"]).t().contiguous()
            out[v + "w13"] = torch.cat([f32(sd[u + "feed_forward.w1.weight"]), f32(sd[u + "feed_forward.w3.weight"])], 0).t().contiguous()
            out[v + "w2"] = f32(sd[u + "feed_forward.w2.weight"]).t().contiguous()
        out[q + "norm"] = f32(sd[p + "norm.weight"])

    out["enc.in_w"], out["enc.in_b"] = conv_pack(wn_new("encoder.block.0.")), f32(sd["encoder.block.0.conv.bias"])
    for i, (s, nl) in enumerate(zip(c.encoder_rates, c.encoder_tf_layers)):
        p, q = f"encoder.block.{i + 1}.block.", f"enc.{i}."
        for j in range(3):
            u, v = p + f"{j}.block.", q + f"r{j}."
            out[v + "a1"], out[v + "a2"] = f32(sd[u + "0.alpha"]).flatten(), f32(sd[u + "2.alpha"]).flatten()
            out[v + "w1"], out[v + "b1"] = conv_pack(wn_new(u + "1.")), f32(sd[u + "1.conv.bias"])
            out[v + "w2"], out[v + "b2"] = conv_pack(wn_new(u + "3.")), f32(sd[u + "3.conv.bias"])
        out[q + "down_alpha"] = f32(sd[p + "3.alpha"]).flatten().repeat(s)
        out[q + "down_w"], out[q + "down_b"] = strided_pack(wn_new(p + "4."), s), f32(sd[p + "4.conv.bias"])
        if nl:
            transformer(p + "5.", q + "tf.", nl)
    n = len(c.encoder_rates)
    out["enc.out_alpha"] = f32(sd[f"encoder.block.{n + 1}.alpha"]).flatten()
    out["enc.out_w"], out["enc.out_b"] = conv_pack(wn_new(f"encoder.block.{n + 2}.")), f32(sd[f"encoder.block.{n + 2}.conv.bias"])
    out["enc.rope"] = _rope_table(max_positions, c.encoder_tf_head_dim, 10000.0).contiguous()
    for i, f in enumerate(c.upsample_factors):
        p, q = f"quantizer.downsample.{i}.", f"down.{i}."
        out[q + "w"], out[q + "b"] = strided_pack(f32(sd[p + "0.conv.weight"]), f), f32(sd[p + "0.conv.bias"])
        out[q + "dw_w"] = f32(sd[p + "1.dwconv.conv.weight"])[:, 0, :].contiguous()
        out[q + "dw_b"] = f32(sd[p + "1.dwconv.conv.bias"])
        out[q + "ln_w"], out[q + "ln_b"] = f32(sd[p + "1.norm.weight"]), f32(sd[p + "1.norm.bias"])
        out[q + "pw1_w"], out[q + "pw1_b"] = f32(sd[p + "1.pwconv1.weight"]).t().contiguous(), f32(sd[p + "1.pwconv1.bias"])
        out[q + "pw2_w"], out[q + "pw2_b"] = f32(sd[p + "1.pwconv2.weight"]).t().contiguous(), f32(sd[p + "1.pwconv2.bias"])
        out[q + "gamma"] = f32(sd[p + "1.gamma"])
    transformer("quantizer.pre_module.", "pre.", c.tf_layers)
    prefixes = ["quantizer.semantic_quantizer.quantizers.0."] + [f"quantizer.quantizer.quantizers.{i}." for i in range(c.n_codebooks)]
    out["vq.in_w"] = torch.stack([_weight_norm(f32(sd[p + "in_proj.weight_v"]), f32(sd[p + "in_proj.weight_g"]))[:, :, 0]
                                  for p in prefixes]).contiguous()                      # [NQ][CD][D]
    out["vq.in_b"] = torch.stack([f32(sd[p + "in_proj.bias"]) for p in prefixes]).contiguous()
    cbn = torch.cat([torch.nn.functional.normalize(f32(sd[p + "codebook.weight"])) for p in prefixes], 0).contiguous()
    out["vq.cbn"] = cbn
    out["vq.cbsq"] = cbn.pow(2).sum(1).contiguous()
    return out


def state_dict_shapes(c: "DACConfig") -> Dict[str, tuple]:
    """Reference state-dict keys the decode path consumes and their shapes (modded_dac.py / rvq.py module registration
    order; weight norm in the `parametrizations.weight.original0/1` style for the decoder and `weight_g/weight_v` for
    the descript quantizers).  Used for synthetic-weight benchmarks and as the list of tensors `pack_weights` needs."""
    D = c.latent_dim
    out: Dict[str, tuple] = {}

    def vq(p, size):
        out[p + "codebook.weight"] = (size, c.codebook_dim)
        out[p + "out_proj.weight_g"] = (D, 1, 1)
        out[p + "out_proj.weight_v"] = (D, c.codebook_dim, 1)
        out[p + "out_proj.bias"] = (D,)

    vq("quantizer.semantic_quantizer.quantizers.0.", c.semantic_codebook_size)
    for i in range(c.n_codebooks):
        vq(f"quantizer.quantizer.quantizers.{i}.", c.codebook_size)
    for i in range(c.tf_layers):
        p = f"quantizer.post_module.layers.{i}."
        out[p + "attention.wqkv.weight"] = (3 * c.tf_heads * c.tf_head_dim, c.tf_dim)
        out[p + "attention.wo.weight"] = (c.tf_dim, c.tf_heads * c.tf_head_dim)
        out[p + "feed_forward.w1.weight"] = (c.tf_ffn, c.tf_dim)
        out[p + "feed_forward.w3.weight"] = (c.tf_ffn, c.tf_dim)
        out[p + "feed_forward.w2.weight"] = (c.tf_dim, c.tf_ffn)
        for k in ("ffn_norm.weight", "attention_norm.weight", "attention_layer_scale.gamma", "ffn_layer_scale.gamma"):
            out[p + k] = (c.tf_dim,)
    out["quantizer.post_module.norm.weight"] = (c.tf_dim,)
    for i, f in enumerate(c.upsample_factors):
        p = f"quantizer.upsample.{i}."
        out[p + "0.conv.weight"] = (D, D, f)
        out[p + "0.conv.bias"] = (D,)
        out[p + "1.dwconv.conv.weight"] = (D, 1, 7)
        out[p + "1.dwconv.conv.bias"] = (D,)
        out[p + "1.norm.weight"] = (D,)
        out[p + "1.norm.bias"] = (D,)
        out[p + "1.pwconv1.weight"] = (c.convnext_ratio * D, D)
        out[p + "1.pwconv1.bias"] = (c.convnext_ratio * D,)
        out[p + "1.pwconv2.weight"] = (D, c.convnext_ratio * D)
        out[p + "1.pwconv2.bias"] = (D,)
        out[p + "1.gamma"] = (D,)

    def wnconv(p, cout, cin, k, transposed=False):
        shape = (cin, cout, k) if transposed else (cout, cin, k)
        out[p + "conv.parametrizations.weight.original0"] = (shape[0], 1, 1)
        out[p + "conv.parametrizations.weight.original1"] = shape
        out[p + "conv.bias"] = (cout,)

    wnconv("decoder.model.0.", c.decoder_dim, D, 7)
    ch = c.decoder_dim
    for i, s_ in enumerate(c.decoder_rates):
        cin, cout = ch // 2 ** i, ch // 2 ** (i + 1)
        p = f"decoder.model.{i + 1}.block."
        out[p + "0.alpha"] = (1, cin, 1)
        wnconv(p + "1.", cout, cin, 2 * s_, transposed=True)
        for j in range(3):
            q = p + f"{2 + j}.block."
            out[q + "0.alpha"] = (1, cout, 1)
            wnconv(q + "1.", cout, cout, 7)
            out[q + "2.alpha"] = (1, cout, 1)
            wnconv(q + "3.", cout, cout, 1)
    n = len(c.decoder_rates)
    out[f"decoder.model.{n + 1}.alpha"] = (1, ch // 2 ** n, 1)
    wnconv(f"decoder.model.{n + 2}.", 1, ch // 2 ** n, 7)
    return out


def encode_state_dict_shapes(c: "DACConfig") -> Dict[str, tuple]:
    """Reference state-dict keys of the encode side (Encoder modded_dac.py:670-709, quantizer.downsample / pre_module
    rvq.py:249-283, the codebooks' in_proj) and their shapes: what `pack_encode_weights` needs."""
    D = c.latent_dim
    out: Dict[str, tuple] = {}

    def wnconv(p, cout, cin, k):
        out[p + "conv.parametrizations.weight.original0"] = (cout, 1, 1)
        out[p + "conv.parametrizations.weight.original1"] = (cout, cin, k)
        out[p + "conv.bias"] = (cout,)

    def transformer(p, layers, dim, heads, hd, ffn):
        for i in range(layers):
            q = f"{p}layers.{i}."
            out[q + "attention.wqkv.weight"] = (3 * heads * hd, dim)
            out[q + "attention.wo.weight"] = (dim, heads * hd)
            for w_, shp in (("w1", (ffn, dim)), ("w3", (ffn, dim)), ("w2", (dim, ffn))):
                out[q + f"feed_forward.{w_}.weight"] = shp
            for n_ in ("ffn_norm.weight", "attention_norm.weight", "attention_layer_scale.gamma", "ffn_layer_scale.gamma"):
                out[q + n_] = (dim,)
        out[p + "norm.weight"] = (dim,)

    wnconv("encoder.block.0.", c.encoder_dim, 1, 7)
    ch = c.encoder_dim
    for i, (s, nl) in enumerate(zip(c.encoder_rates, c.encoder_tf_layers)):
        p = f"encoder.block.{i + 1}.block."
        for j in range(3):
            out[p + f"{j}.block.0.alpha"] = (1, ch, 1)
            wnconv(p + f"{j}.block.1.", ch, ch, 7)
            out[p + f"{j}.block.2.alpha"] = (1, ch, 1)
            wnconv(p + f"{j}.block.3.", ch, ch, 1)
        out[p + "3.alpha"] = (1, ch, 1)
        wnconv(p + "4.", 2 * ch, ch, 2 * s)
        ch *= 2
        if nl:
            transformer(p + "5.", nl, ch, ch // c.encoder_tf_head_dim, c.encoder_tf_head_dim, 3 * ch)
    n = len(c.encoder_rates)
    out[f"encoder.block.{n + 1}.alpha"] = (1, ch, 1)
    wnconv(f"encoder.block.{n + 2}.", D, ch, 3)
    for i, f in enumerate(c.upsample_factors):
        p = f"quantizer.downsample.{i}."
        out[p + "0.conv.weight"], out[p + "0.conv.bias"] = (D, D, f), (D,)
        out[p + "1.dwconv.conv.weight"], out[p + "1.dwconv.conv.bias"] = (D, 1, 7), (D,)
        out[p + "1.norm.weight"], out[p + "1.norm.bias"] = (D,), (D,)
        out[p + "1.pwconv1.weight"], out[p + "1.pwconv1.bias"] = (c.convnext_ratio * D, D), (c.convnext_ratio * D,)
        out[p + "1.pwconv2.weight"], out[p + "1.pwconv2.bias"] = (D, c.convnext_ratio * D), (D,)
        out[p + "1.gamma"] = (D,)
    transformer("quantizer.pre_module.", c.tf_layers, c.tf_dim, c.tf_heads, c.tf_head_dim, c.tf_ffn)
    for p in ["quantizer.semantic_quantizer.quantizers.0."] + [f"quantizer.quantizer.quantizers.{i}." for i in range(c.n_codebooks)]:
        out[p + "in_proj.weight_g"], out[p + "in_proj.weight_v"], out[p + "in_proj.bias"] = (c.codebook_dim, 1, 1), (c.codebook_dim, D, 1), (c.codebook_dim,)
    return out


def synthetic_state_dict(c: "DACConfig", seed: int = 0, device="cpu", encode_side: bool = False) -> Dict[str, Tensor]:
    """Random weights of the right shapes (benchmarks only: bounded activations, no claim of audio); `encode_side` adds the
    encoder, the quantizer's downsample / pre_module and the codebooks' in_proj."""
    g = torch.Generator().manual_seed(seed)
    sd: Dict[str, Tensor] = {}
    shapes = dict(state_dict_shapes(c))
    if encode_side:
        shapes.update(encode_state_dict_shapes(c))
    for name, shape in shapes.items():
        if name.endswith("alpha"):
            w = 0.5 + torch.rand(shape, generator=g)
        elif name.endswith(("gamma", "original0", "weight_g")):
            w = 0.3 * (0.5 + torch.rand(shape, generator=g))
        elif name.endswith("norm.weight"):
            w = torch.ones(shape)
        elif name.endswith("bias"):
            w = torch.zeros(shape)
        else:
            fan = shape[1] * (shape[2] if len(shape) == 3 else 1) if len(shape) > 1 else 1
            w = torch.randn(shape, generator=g) / math.sqrt(max(1, fan))
        sd[name] = w.to(device)
    return sd


class DACB200:
    """Engine-backed stand-in for modded_dac.DAC: from_indices (decode side, the hot path) and, when the state dict has the
    encode side's weights, encode (reference voice -> codes)."""

    def __init__(self, config, state_dict: Dict[str, Tensor], device="cuda:0", max_batch: int = 4, max_frames: int = 1024,
                 operand: str = "f16", encode_side: Optional[bool] = None):
        """`encode_side`: None = bind the encode side when the state dict has it (~0.8 GB of fp32 weights at the full config),
        False = decode only, True = required (KeyError when it is missing)."""
        _lib.require_cuda()
        self.lib = _lib.load()
        self.config = c = DACConfig.from_any(config)
        self.device = torch.device(device)
        self.sample_rate = c.sample_rate
        self.hop_length = c.hop
        self.frame_length = c.hop
        self.max_batch, self.max_frames = int(max_batch), int(max_frames)
        rates = list(c.decoder_rates) + [1] * (4 - len(c.decoder_rates))
        ups = list(c.upsample_factors) + [1] * (2 - len(c.upsample_factors))
        self._desc = d = _lib.DACDesc(
            latent_dim=c.latent_dim, decoder_dim=c.decoder_dim, n_rates=len(c.decoder_rates), rate0=rates[0], rate1=rates[1],
            rate2=rates[2], rate3=rates[3], n_codebooks=c.n_codebooks, codebook_size=c.codebook_size,
            semantic_codebook_size=c.semantic_codebook_size, codebook_dim=c.codebook_dim, n_up=len(c.upsample_factors),
            up0=ups[0], up1=ups[1], tf_layers=c.tf_layers, tf_heads=c.tf_heads, tf_dim=c.tf_dim, tf_ffn=c.tf_ffn,
            tf_head_dim=c.tf_head_dim, tf_window=c.tf_window, convnext_ratio=c.convnext_ratio, max_batch=self.max_batch,
            max_frames=self.max_frames, tf_norm_eps=c.tf_norm_eps)
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            nbytes = self.lib.fsb_dac_workspace_bytes(C.byref(d))
            self._workspace = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
            rc = self.lib.fsb_dac_create(C.byref(d), self._workspace.data_ptr(), nbytes, C.byref(self._h))
            self._ck(rc)
            dev_sd = {k: v.to(self.device) for k, v in state_dict.items()}
            self.weights = {k: v.to(self.device, torch.float32).contiguous()
                            for k, v in pack_weights(c, dev_sd, self.max_frames).items()}
            # tcgen05 pipeline: every GEMM weight again as [taps][N][Cin] in the 16-bit operand type (K-major rows, what a TMA
            # box reads); the transformer's gate/up matrix with its columns interleaved (SwiGLU in the GEMM epilogue)
            self.operand = operand
            hdt = torch.float16 if operand == "f16" else torch.bfloat16
            self.can_encode = has_encode_side(dev_sd) if encode_side is None else bool(encode_side)
            for k in [k for k, v in self.weights.items() if v.dim() == 3 and k != "dec.out_w"] + \
                     [k for k in self.weights if k.startswith("tf.") and k.split(".")[-1] in ("wqkv", "wo", "w13", "w2")] + \
                     [k for k in self.weights if k.endswith(("pw1_w", "pw2_w"))]:
                w = self.weights[k]
                w = w[None] if w.dim() == 2 else w                      # [taps][Cin][N]
                w = w.permute(0, 2, 1)                                  # [taps][N][Cin]
                if k.endswith("w13"):
                    F_ = w.shape[1] // 2
                    w = torch.stack([w[:, :F_], w[:, F_:]], dim=2).reshape(w.shape[0], 2 * F_, w.shape[2])
                self.weights[k + ".h"] = w.to(hdt).contiguous()
            if self.can_encode:        # fp32 only: the encode side's result is a set of arg-max decisions
                down = int(math.prod(c.upsample_factors))
                for k, v in pack_encode_weights(c, dev_sd, self.max_frames * down).items():
                    self.weights[k] = v.to(self.device, torch.float32).contiguous()
            for k, v in self.weights.items():
                self._ck(self.lib.fsb_dac_bind_weight(self._h, k.encode(), v.data_ptr(), v.numel()))
            self.set_option("half", 1 if operand == "f16" else 0)
            if self.can_encode:
                if len(c.encoder_rates) != len(c.encoder_tf_layers) or not 1 <= len(c.encoder_rates) <= 4:
                    raise ValueError("encoder_rates / encoder_tf_layers: 1..4 entries each, same length")
                self.set_option("enc_dim", c.encoder_dim)
                self.set_option("enc_n_rates", len(c.encoder_rates))
                for i, (r_, n_) in enumerate(zip(c.encoder_rates, c.encoder_tf_layers)):
                    self.set_option(f"enc_rate{i}", r_)
                    self.set_option(f"enc_tf{i}", n_)
                self.set_option("enc_window", c.encoder_tf_window)
                self.set_option("enc_head_dim", c.encoder_tf_head_dim)

    def _ck(self, rc: int):
        if rc != 0:
            raise RuntimeError("fish_b200 dac: " + (self.lib.fsb_dac_last_error(self._h) or b"").decode())

    def __del__(self):
        try:
            if self._h:
                self.lib.fsb_dac_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def set_option(self, name: str, value: int):
        """"tc": 1 = tcgen05 pipeline (default: 16-bit operands of the type chosen at construction, fp32 accumulation and
        residual stream), 0 = the round-1 kernels, for which "tf32": 1 = mma.sync TF32 GEMMs, 0 = fp32 CUDA-core path (parity
        anchor); "force_simt": 1 = the tcgen05 pipeline's data flow with its CUDA-core twin kernel (cross-check)."""
        self._ck(self.lib.fsb_dac_set_option(self._h, name.encode(), int(value)))

    def flops(self, B: int, T: int) -> float:
        return float(self.lib.fsb_dac_flops(self._h, B, T))

    @torch.inference_mode()
    def from_indices(self, indices: Tensor) -> Tensor:
        """DAC.from_indices (modded_dac.py:925-927): [B, 1+n_codebooks, T] int64 -> [B, 1, hop*T] fp32.
        Like the reference (rvq.py:354-359) the caller's `indices` tensor is clamped in place when it is already a
        contiguous int64 CUDA tensor."""
        if indices.dim() != 3 or indices.shape[1] != self.config.n_codebooks + 1:
            raise ValueError(f"expected [B, {self.config.n_codebooks + 1}, T] indices, got {tuple(indices.shape)}")
        B, _, T = indices.shape
        idx = indices
        if idx.device != self.device or idx.dtype != torch.int64 or not idx.is_contiguous():
            idx = indices.to(device=self.device, dtype=torch.int64).contiguous()
        wav = torch.empty((B, 1, self.hop_length * T), dtype=torch.float32, device=self.device)
        st = torch.cuda.current_stream().cuda_stream
        if T <= self.max_frames:
            for b0 in range(0, B, self.max_batch):
                nb = min(self.max_batch, B - b0)
                self._ck(self.lib.fsb_dac_decode(self._h, idx[b0: b0 + nb].data_ptr(), nb, T, wav[b0: b0 + nb].data_ptr(), st))
            return wav
        # A clip longer than the workspace: overlapping pieces.  Every layer of the decode side is causal with a finite
        # reach (context_frames), so a piece that starts `halo` frames early reproduces the frames after the halo exactly,
        # given their absolute RoPE positions (option "pos_offset").
        halo = self.context_frames()
        stride = self.max_frames - halo
        if stride < 1:
            raise ValueError(f"T={T} exceeds max_frames={self.max_frames}, and pieces need max_frames > {halo} (the decode "
                             f"side's context in frames)")
        self._ensure_rope(T)
        hop = self.hop_length
        try:
            for b0 in range(0, B, self.max_batch):
                nb = min(self.max_batch, B - b0)
                t0 = 0
                while t0 < T:
                    a = max(0, t0 - halo)                      # piece = frames [a, t1), kept = [t0, t1)
                    t1 = min(T, a + self.max_frames)
                    piece = idx[b0: b0 + nb, :, a:t1].contiguous()
                    out = torch.empty((nb, 1, hop * (t1 - a)), dtype=torch.float32, device=self.device)
                    self.set_option("pos_offset", a)
                    self._ck(self.lib.fsb_dac_decode(self._h, piece.data_ptr(), nb, t1 - a, out.data_ptr(), st))
                    wav[b0: b0 + nb, :, hop * t0: hop * t1] = out[:, :, hop * (t0 - a):]
                    if idx is indices:                         # keep the reference's in-place clamp visible to the caller
                        idx[b0: b0 + nb, :, t0:t1] = piece[:, :, t0 - a:]
                    t0 = t1
        finally:
            self.set_option("pos_offset", 0)
        return wav

    @torch.inference_mode()
    def encode(self, audio_data: Tensor, audio_lengths: Optional[Tensor] = None, n_quantizers: Optional[int] = None, **kwargs):
        """DAC.encode (modded_dac.py:874-923): [B, 1, n] or [B, n] waveform -> (indices [B, 1+n_codebooks, ceil(n / frame)] int64,
        indices_lens).  fp32 throughout.  Needs the encode side's weights in the state dict the engine was built from."""
        if not self.can_encode:
            raise RuntimeError("fish_b200 dac: built from a state dict without the encode side (encoder.*, quantizer.downsample.*, "
                               "quantizer.pre_module.*, in_proj.*)")
        if n_quantizers is not None and n_quantizers != self.config.n_codebooks:
            # rvq.py:309-311 passes it to the residual codebooks only; fewer codebooks is a training-time feature
            raise NotImplementedError("n_quantizers other than all of them")
        if audio_data.dim() == 2:
            audio_data = audio_data.unsqueeze(1)
        if audio_data.dim() != 3 or audio_data.shape[1] != 1:
            raise ValueError(f"expected [B, 1, n] or [B, n] audio, got {tuple(audio_data.shape)}")
        B, _, length = audio_data.shape
        frame = self.frame_length
        right_pad = -length % frame
        T = (length + right_pad) // frame
        if T < 1 or T > self.max_frames:
            raise ValueError(f"{length} samples = {T} code frames, outside 1..max_frames={self.max_frames}")
        x = torch.zeros((B, length + right_pad), dtype=torch.float32, device=self.device)
        x[:, :length] = audio_data[:, 0].to(device=self.device, dtype=torch.float32)
        if audio_lengths is None:
            audio_lengths = torch.LongTensor([length + right_pad]).to(self.device)
        codes = torch.empty((B, self.config.n_codebooks + 1, T), dtype=torch.int64, device=self.device)
        st = torch.cuda.current_stream().cuda_stream
        for b0 in range(0, B, self.max_batch):
            nb = min(self.max_batch, B - b0)
            self._ck(self.lib.fsb_dac_encode(self._h, x[b0: b0 + nb].data_ptr(), nb, length + right_pad,
                                             codes[b0: b0 + nb].data_ptr(), st))
        return codes, torch.ceil(audio_lengths / frame).long()

    def context_frames(self) -> int:
        """How many earlier code frames an output sample can depend on: the causal convs' taps walked back from the waveform
        to the code frames (modded_dac.py:599-620, 712-801; rvq.py:263-276) plus the window transformer's layers x (window - 1)
        (modded_dac.py:380-398)."""
        c = self.config
        ctx = 6                                                # output conv, 7 taps
        for s_ in reversed(list(c.decoder_rates)):
            ctx += 6 * (1 + 3 + 9)                             # three residual units: 7 taps at dilation 1, 3, 9 (+ 1-tap convs)
            ctx = -(-ctx // s_) + 1                            # transposed conv, kernel 2s stride s: one earlier input row
        ctx += 6                                               # input conv, 7 taps
        for f in reversed(list(c.upsample_factors)):
            ctx += 6                                           # ConvNeXt depthwise conv, 7 taps
            ctx = -(-ctx // f)                                 # transposed conv, kernel = stride
        return ctx + c.tf_layers * (c.tf_window - 1)

    def stream(self, batch: int = 1, halo: Optional[int] = None) -> "DACStream":
        """A DACStream over this engine (audio out while code frames still arrive)."""
        return DACStream(self, batch, halo)

    def _ensure_rope(self, positions: int):
        c = self.config
        if self.weights["tf.rope"].shape[0] >= positions:
            return
        n = max(positions, 2 * self.weights["tf.rope"].shape[0])
        torch.cuda.current_stream().synchronize()              # earlier launches may still read the old table
        t = _rope_table(n, c.tf_head_dim, c.tf_rope_base).to(self.device, torch.float32).contiguous()
        self.weights["tf.rope"] = t
        self._ck(self.lib.fsb_dac_bind_weight(self._h, b"tf.rope", t.data_ptr(), t.numel()))



class DACStream:
    """Waveform out as code frames arrive: `push(codes[B, 1+n_codebooks, n])` returns the `hop * n` new samples of every clip.

    The reference's engine decodes a whole segment when its last token is there (inference_engine/__init__.py:104-126 feeding
    vq_manager.py:20-33); its decode side is causal with a finite reach (DACB200.context_frames), so the samples of new frames
    only need the last `halo` code frames again.  With halo = context_frames() (the default) the stream equals the one-pass
    decode of the whole clip bit for bit; a shorter halo trades exactness of the first samples of each push for less
    recomputation (the window transformer's reach, layers x (window - 1) frames, is the long part)."""

    def __init__(self, dac: "DACB200", batch: int = 1, halo: Optional[int] = None):
        self.dac = dac
        self.batch = int(batch)
        if self.batch < 1 or self.batch > dac.max_batch:
            raise ValueError(f"batch={batch} outside 1..max_batch={dac.max_batch}")
        self.halo = dac.context_frames() if halo is None else int(halo)
        if self.halo < 0 or self.halo >= dac.max_frames:
            raise ValueError(f"halo={self.halo} needs 0 <= halo < max_frames={dac.max_frames}")
        self.frames = 0                                        # code frames taken so far
        self._tail = torch.empty((self.batch, dac.config.n_codebooks + 1, 0), dtype=torch.int64, device=dac.device)

    @torch.inference_mode()
    def push(self, codes: Tensor) -> Tensor:
        d = self.dac
        if codes.dim() != 3 or codes.shape[0] != self.batch or codes.shape[1] != d.config.n_codebooks + 1:
            raise ValueError(f"expected [{self.batch}, {d.config.n_codebooks + 1}, n] codes, got {tuple(codes.shape)}")
        codes = codes.to(device=d.device, dtype=torch.int64)
        n, hop = codes.shape[2], d.hop_length
        wav = torch.empty((self.batch, 1, hop * n), dtype=torch.float32, device=d.device)
        st = torch.cuda.current_stream().cuda_stream
        done = 0
        while done < n:
            keep = self._tail.shape[2]
            take = min(n - done, d.max_frames - keep)
            piece = torch.cat([self._tail, codes[:, :, done: done + take]], dim=2).contiguous()
            a = self.frames - keep                             # absolute index of the piece's first frame
            d._ensure_rope(self.frames + take)
            out = torch.empty((self.batch, 1, hop * (keep + take)), dtype=torch.float32, device=d.device)
            try:
                d.set_option("pos_offset", a)
                d._ck(d.lib.fsb_dac_decode(d._h, piece.data_ptr(), self.batch, keep + take, out.data_ptr(), st))
            finally:
                d.set_option("pos_offset", 0)
            wav[:, :, hop * done: hop * (done + take)] = out[:, :, hop * keep:]
            self._tail = piece[:, :, max(0, keep + take - self.halo):]
            self.frames += take
            done += take
        return wav

    def reset(self):
        self.frames = 0
        self._tail = self._tail[:, :, :0]
