"""CPU restatement of the reference DAC *decode* path (TEST INFRASTRUCTURE).

Follows, op for op,
  /root/reference/fish_speech/models/dac/modded_dac.py   (DAC.from_indices:925, Decoder:760,
      DecoderBlock:712, ResidualUnit:599, CausalConvNet:521, CausalTransConvNet:563,
      WindowLimitedTransformer:349, Transformer:97, Attention:198, RMSNorm:320)
  /root/reference/fish_speech/models/dac/rvq.py          (DownsampleResidualVectorQuantize.decode:352,
      upsample:263-276, ConvNeXtBlock:129)
  descript-audio-codec==1.0.0 (absent; restated): Snake1d, ResidualVectorQuantize.from_codes
and the architecture in fish_speech/configs/modded_dac_vq.yaml:1-50.

State-dict driven; key names are the reference nn.Module's.  **Parity unpinned** w.r.t.
the real descript wheel (see oracle/__init__.py); pinned against the reference classes
(+shim) by oracle/gen_golden_dac.py.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Tuple

import torch
import torch.nn.functional as F
from torch import Tensor


@dataclass
class DACConfig:
    """modded_dac_vq.yaml values the decode path uses."""

    sample_rate: int = 44100
    latent_dim: int = 1024  # encoder_dim 64 * 2**4 (modded_dac.py:832-835)
    decoder_dim: int = 1536
    decoder_rates: Tuple[int, ...] = (8, 8, 4, 2)
    n_codebooks: int = 9  # residual codebooks; +1 semantic
    codebook_size: int = 1024
    semantic_codebook_size: int = 4096
    codebook_dim: int = 8
    upsample_factors: Tuple[int, ...] = (2, 2)  # downsample_factor reversed (both 2)
    # post_module: WindowLimitedTransformer(causal, window 128)
    tf_layers: int = 8
    tf_heads: int = 16
    tf_dim: int = 1024
    tf_ffn: int = 3072
    tf_head_dim: int = 64
    tf_rope_base: float = 10000.0
    tf_norm_eps: float = 1e-5
    tf_window: int = 128
    convnext_ratio: int = 4
    # encode side (DAC.encode, modded_dac.py:874-923): Encoder(encoder_dim, encoder_rates) with a window transformer after the
    # blocks named in encoder_tf_layers (heads = dim // 64, ffn = 3 * dim, window 512: modded_dac.py:636-650), the quantizer's
    # downsample stages (= upsample_factors) and its pre_module (the post_module's twin, own weights)
    encoder_dim: int = 64
    encoder_rates: Tuple[int, ...] = (2, 4, 8, 8)
    encoder_tf_layers: Tuple[int, ...] = (0, 0, 0, 4)
    encoder_tf_window: int = 512
    encoder_tf_head_dim: int = 64

    @property
    def hop(self) -> int:  # samples per code frame: 2*2 * 8*8*4*2 = 2048
        return int(math.prod(self.upsample_factors) * math.prod(self.decoder_rates))


def small_config() -> DACConfig:
    """Reduced widths for fast CPU-side tests (same topology, every layer type present)."""
    return DACConfig(latent_dim=128, decoder_dim=192, tf_layers=2, tf_heads=2, tf_dim=128, tf_ffn=384,
                     tf_window=8, n_codebooks=9, codebook_size=64, semantic_codebook_size=128,
                     encoder_dim=8, encoder_tf_layers=(0, 0, 0, 2), encoder_tf_window=6)


def state_dict_keys(c: DACConfig) -> Dict[str, tuple]:
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
        out[p + "ffn_norm.weight"] = (c.tf_dim,)
        out[p + "attention_norm.weight"] = (c.tf_dim,)
        out[p + "attention_layer_scale.gamma"] = (c.tf_dim,)
        out[p + "ffn_layer_scale.gamma"] = (c.tf_dim,)
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
    for i, s in enumerate(c.decoder_rates):
        cin, cout = ch // 2 ** i, ch // 2 ** (i + 1)
        p = f"decoder.model.{i + 1}.block."
        out[p + "0.alpha"] = (1, cin, 1)
        wnconv(p + "1.", cout, cin, 2 * s, transposed=True)
        for j in range(3):
            q = p + f"{2 + j}.block."
            out[q + "0.alpha"] = (1, cout, 1)
            wnconv(q + "1.", cout, cout, 7)
            out[q + "2.alpha"] = (1, cout, 1)
            wnconv(q + "3.", cout, cout, 1)
    n = len(c.decoder_rates)
    cl = ch // 2 ** n
    out[f"decoder.model.{n + 1}.alpha"] = (1, cl, 1)
    wnconv(f"decoder.model.{n + 2}.", 1, cl, 7)
    return out


def encode_state_dict_keys(c: DACConfig) -> Dict[str, tuple]:
    """The encode side's parameters (reference names), kept apart from state_dict_keys: the seeded decode-side weights and the
    fixtures made from them do not change when these are drawn after them."""
    D = c.latent_dim
    out: Dict[str, tuple] = {}

    def wnconv(p, cout, cin, k):
        out[p + "conv.parametrizations.weight.original0"] = (cout, 1, 1)
        out[p + "conv.parametrizations.weight.original1"] = (cout, cin, k)
        out[p + "conv.bias"] = (cout,)

    def transformer(p, layers, dim, heads, hd, ffn):
        for i in range(layers):
            q = p + f"layers.{i}."
            out[q + "attention.wqkv.weight"] = (3 * heads * hd, dim)
            out[q + "attention.wo.weight"] = (dim, heads * hd)
            out[q + "feed_forward.w1.weight"] = (ffn, dim)
            out[q + "feed_forward.w3.weight"] = (ffn, dim)
            out[q + "feed_forward.w2.weight"] = (dim, ffn)
            out[q + "ffn_norm.weight"] = (dim,)
            out[q + "attention_norm.weight"] = (dim,)
            out[q + "attention_layer_scale.gamma"] = (dim,)
            out[q + "ffn_layer_scale.gamma"] = (dim,)
        out[p + "norm.weight"] = (dim,)

    wnconv("encoder.block.0.", c.encoder_dim, 1, 7)
    ch = c.encoder_dim
    for i, (s, nl) in enumerate(zip(c.encoder_rates, c.encoder_tf_layers)):
        p = f"encoder.block.{i + 1}.block."
        for j in range(3):
            q = p + f"{j}.block."
            out[q + "0.alpha"] = (1, ch, 1)
            wnconv(q + "1.", ch, ch, 7)
            out[q + "2.alpha"] = (1, ch, 1)
            wnconv(q + "3.", ch, ch, 1)
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
    transformer("quantizer.pre_module.", c.tf_layers, c.tf_dim, c.tf_heads, c.tf_head_dim, c.tf_ffn)
    for p in ["quantizer.semantic_quantizer.quantizers.0."] + [f"quantizer.quantizer.quantizers.{i}." for i in range(c.n_codebooks)]:
        out[p + "in_proj.weight_g"] = (c.codebook_dim, 1, 1)
        out[p + "in_proj.weight_v"] = (c.codebook_dim, D, 1)
        out[p + "in_proj.bias"] = (c.codebook_dim,)
    return out


def make_state_dict(c: DACConfig, seed: int = 0, dtype=torch.float32, gain: float = 1.0) -> Dict[str, Tensor]:
    """Seeded synthetic decode-path weights.  The reference's default init (trunc_normal
    std .02 convs, zero biases, alpha=1, layer-scale 1e-2, ConvNeXt gamma 1e-6) makes whole
    branches numerically invisible, so parity fixtures use livelier values: fan-in scaled
    normal weights, small random biases, alpha in [0.5,1.5], layer-scale ~0.3, gamma ~0.3."""
    g = torch.Generator().manual_seed(seed)
    sd: Dict[str, Tensor] = {}
    for name, shape in state_dict_keys(c).items():
        if name.endswith("alpha"):
            w = 0.5 + torch.rand(shape, generator=g)
        elif name.endswith(("layer_scale.gamma", "1.gamma")):
            w = 0.3 * (0.5 + torch.rand(shape, generator=g))
        elif name.endswith(("norm.weight",)) and len(shape) == 1:
            w = 1.0 + 0.1 * (2 * torch.rand(shape, generator=g) - 1)
        elif name.endswith("bias"):
            w = 0.02 * torch.randn(shape, generator=g)
        elif name.endswith("weight_g"):
            w = gain * (0.8 + 0.4 * torch.rand(shape, generator=g))
        elif name.endswith("original0"):
            # decoder weight-norm gains chosen so activations stay O(1) into the final tanh:
            # transposed convs spread one input over k/2 phases (gain sqrt(k/4)), residual
            # branches are damped (0.45), first conv 1.0, last conv 0.5
            v_shape = state_dict_keys(c)[name.replace("original0", "original1")]
            if ".block.1.conv." in name:
                base = 0.55 * math.sqrt(v_shape[2] / 4.0)
            elif ".block." in name:
                base = 0.45
            elif name.startswith("decoder.model.0."):
                base = 1.0
            else:
                base = 0.35
            w = gain * base * (0.8 + 0.4 * torch.rand(shape, generator=g))
        elif name.endswith("codebook.weight"):
            w = torch.randn(shape, generator=g)
        elif name.endswith(("weight_v", "original1")):
            w = torch.randn(shape, generator=g)
        else:  # plain linear / conv weights: fan-in scaled
            fan_in = shape[1] * (shape[2] if len(shape) == 3 else 1)
            if name.endswith("0.conv.weight") and len(shape) == 3:  # ConvTranspose1d [Cin, Cout, k]
                fan_in = shape[0]
            w = torch.randn(shape, generator=g) * (gain / math.sqrt(fan_in))
        sd[name] = w.to(dtype)
    return sd


def make_encode_state_dict(c: DACConfig, seed: int = 0, dtype=torch.float32) -> Dict[str, Tensor]:
    """Seeded synthetic encode-side weights (own generator: the decode-side draw of make_state_dict stays what it was).  Same
    idea as there: lively values, weight-norm gains chosen so that activations stay O(1) from the waveform to the latents."""
    g = torch.Generator().manual_seed(seed + 7919)
    keys = encode_state_dict_keys(c)
    sd: Dict[str, Tensor] = {}
    for name, shape in keys.items():
        if name.endswith("alpha"):
            w = 0.5 + torch.rand(shape, generator=g)
        elif name.endswith(("layer_scale.gamma", "1.gamma")):
            w = 0.3 * (0.5 + torch.rand(shape, generator=g))
        elif name.endswith("norm.weight") and len(shape) == 1:
            w = 1.0 + 0.1 * (2 * torch.rand(shape, generator=g) - 1)
        elif name.endswith("bias"):
            w = 0.02 * torch.randn(shape, generator=g)
        elif name.endswith("weight_g"):                       # VQ in_proj
            w = 0.8 + 0.4 * torch.rand(shape, generator=g)
        elif name.endswith("original0"):
            if name.startswith("encoder.block.0."):
                base = 3.0                                    # waveform (std ~0.3) -> O(1)
            elif ".block.4.conv." in name:
                base = 0.8                                    # strided convs
            elif ".block." in name:
                base = 0.4                                    # residual branches
            else:
                base = 1.0                                    # last conv
            w = base * (0.8 + 0.4 * torch.rand(shape, generator=g))
        elif name.endswith(("weight_v", "original1")):
            w = torch.randn(shape, generator=g)
        else:                                                 # plain linear / conv weights: fan-in scaled
            fan_in = shape[1] * (shape[2] if len(shape) == 3 else 1)
            w = torch.randn(shape, generator=g) / math.sqrt(fan_in)
        sd[name] = w.to(dtype)
    return sd


def make_audio(c: DACConfig, B: int, n_samples: int, seed: int = 4) -> Tensor:
    """A synthetic waveform [B, 1, n]: a few sines plus noise, std ~0.3."""
    g = torch.Generator().manual_seed(seed)
    t = torch.arange(n_samples, dtype=torch.float32) / c.sample_rate
    f = 100.0 + 3000.0 * torch.rand((B, 4, 1), generator=g)
    ph = 6.2831853 * torch.rand((B, 4, 1), generator=g)
    x = torch.sin(6.2831853 * f * t + ph).sum(1, keepdim=True) * 0.15
    return (x + 0.1 * torch.randn((B, 1, n_samples), generator=g)).contiguous()


def make_codes(c: DACConfig, B: int, T: int, seed: int = 2) -> Tensor:
    """row 0 ~ U[0,4096), rows 1-9 ~ U[0,1024), int64."""
    g = torch.Generator().manual_seed(seed)
    codes = torch.randint(0, c.codebook_size, (B, c.n_codebooks + 1, T), generator=g)
    codes[:, 0] = torch.randint(0, c.semantic_codebook_size, (B, T), generator=g)
    return codes


# ----------------------------------------------------------------------------- building blocks
def snake(x: Tensor, alpha: Tensor) -> Tensor:
    """descript-audio-codec 1.0.0 dac/nn/layers.py: x + (alpha+1e-9)^-1 * sin(alpha x)^2."""
    shape = x.shape
    x = x.reshape(shape[0], shape[1], -1)
    x = x + (alpha + 1e-9).reciprocal() * torch.sin(alpha * x).pow(2)
    return x.reshape(shape)


def _wn(sd, p) -> Tensor:
    """parametrizations.weight_norm(dim=0): g * v / ||v|| (modded_dac.py:553-556)."""
    return torch._weight_norm(sd[p + "conv.parametrizations.weight.original1"],
                              sd[p + "conv.parametrizations.weight.original0"], 0)


def causal_conv(x: Tensor, w: Tensor, b: Tensor, dilation: int = 1, groups: int = 1) -> Tensor:
    """CausalConvNet.forward, stride 1 (modded_dac.py:546-552 / rvq.py:83-89): left-pad (k-1)*d."""
    k = (w.shape[-1] - 1) * dilation + 1
    x = F.pad(x, (k - 1, 0), mode="constant", value=0)
    return F.conv1d(x, w, b, dilation=dilation, groups=groups).contiguous()


def causal_trans_conv(x: Tensor, w: Tensor, b: Tensor, stride: int) -> Tensor:
    """CausalTransConvNet.forward (modded_dac.py:574-580 / rvq.py:110-117): trim k-s on the right."""
    x = F.conv_transpose1d(x, w, b, stride=stride)
    pad = w.shape[-1] - stride
    if pad > 0:
        x = x[..., : x.shape[-1] - pad]
    return x.contiguous()


def rms_norm_double_round(x: Tensor, w: Tensor, eps: float) -> Tensor:
    """modded_dac.py:320-331."""
    xf = x.float()
    o = (xf * torch.rsqrt(torch.mean(xf * xf, dim=-1, keepdim=True) + eps)).type_as(x)
    return o * w


def precompute_freqs_cis(seq_len: int, n_elem: int, base: float) -> Tensor:
    """modded_dac.py:442-452 (bf16 table even when the codec runs fp32)."""
    freqs = 1.0 / (base ** (torch.arange(0, n_elem, 2)[: (n_elem // 2)].float() / n_elem))
    t = torch.arange(seq_len)
    freqs = torch.outer(t, freqs)
    freqs_cis = torch.polar(torch.ones_like(freqs), freqs)
    return torch.stack([freqs_cis.real, freqs_cis.imag], dim=-1).to(torch.bfloat16)


def apply_rotary_emb(x: Tensor, freqs_cis: Tensor) -> Tensor:
    """modded_dac.py:455-467."""
    xshaped = x.float().reshape(*x.shape[:-1], -1, 2)
    freqs_cis = freqs_cis.view(1, xshaped.size(1), 1, xshaped.size(3), 2)
    x_out2 = torch.stack(
        [xshaped[..., 0] * freqs_cis[..., 0] - xshaped[..., 1] * freqs_cis[..., 1],
         xshaped[..., 1] * freqs_cis[..., 0] + xshaped[..., 0] * freqs_cis[..., 1]], -1)
    return x_out2.flatten(3).type_as(x)


def window_mask(T: int, window: int) -> Tensor:
    """WindowLimitedTransformer.make_window_limited_mask (modded_dac.py:380-398)."""
    mask = torch.tril(torch.ones(T, T))
    row = torch.arange(T).view(-1, 1)
    valid = (row - window + 1).clamp(min=0)
    col = torch.arange(T)
    return ((col >= valid) & mask.bool())[None, None]


# ----------------------------------------------------------------------------- stages
def rvq_dequant(c: DACConfig, sd: Dict[str, Tensor], indices: Tensor) -> Tensor:
    """rvq.py:352-363 + descript from_codes.  NB clamps `indices` IN PLACE like the reference."""
    indices[:, 0] = torch.clamp(indices[:, 0], max=c.semantic_codebook_size - 1)
    indices[:, 1:] = torch.clamp(indices[:, 1:], max=c.codebook_size - 1)

    def from_codes(prefix, codes):
        z_q = 0.0
        for i in range(codes.shape[1]):
            p = f"{prefix}.quantizers.{i}."
            z_p_i = F.embedding(codes[:, i, :], sd[p + "codebook.weight"]).transpose(1, 2)
            w = torch._weight_norm(sd[p + "out_proj.weight_v"], sd[p + "out_proj.weight_g"], 0)
            z_q = z_q + F.conv1d(z_p_i, w, sd[p + "out_proj.bias"])
        return z_q

    return from_codes("quantizer.semantic_quantizer", indices[:, :1]) + \
        from_codes("quantizer.quantizer", indices[:, 1:])


def post_module(c: DACConfig, sd: Dict[str, Tensor], z: Tensor) -> Tensor:
    return window_transformer(sd, "quantizer.post_module.", z, c.tf_layers, c.tf_heads, c.tf_head_dim, c.tf_window,
                              c.tf_norm_eps, c.tf_rope_base)


def window_transformer(sd: Dict[str, Tensor], prefix: str, z: Tensor, layers: int, H: int, hd: int, window: int,
                       eps: float, rope_base: float) -> Tensor:
    """WindowLimitedTransformer.forward (modded_dac.py:418-439) over Transformer.forward (:145-171), channels first."""
    x = z.transpose(1, 2)  # channels_first
    B, T, _ = x.shape
    mask = window_mask(T, window)
    freqs = precompute_freqs_cis(T, hd, rope_base)
    for i in range(layers):
        p = f"{prefix}layers.{i}."
        n = rms_norm_double_round(x, sd[p + "attention_norm.weight"], eps)
        q, k, v = F.linear(n, sd[p + "attention.wqkv.weight"]).split([H * hd] * 3, dim=-1)
        q, k, v = (t.view(B, T, H, hd) for t in (q, k, v))
        q, k = apply_rotary_emb(q, freqs), apply_rotary_emb(k, freqs)
        q, k, v = (t.transpose(1, 2) for t in (q, k, v))
        y = F.scaled_dot_product_attention(q, k, v, dropout_p=0.0, attn_mask=mask)
        y = y.transpose(1, 2).contiguous().view(B, T, H * hd)
        y = F.linear(y, sd[p + "attention.wo.weight"])
        h = x + y.mul_(sd[p + "attention_layer_scale.gamma"])  # LayerScale(inplace=True)
        n = rms_norm_double_round(h, sd[p + "ffn_norm.weight"], eps)
        ff = F.linear(F.silu(F.linear(n, sd[p + "feed_forward.w1.weight"])) *
                      F.linear(n, sd[p + "feed_forward.w3.weight"]), sd[p + "feed_forward.w2.weight"])
        x = h + ff.mul_(sd[p + "ffn_layer_scale.gamma"])
    x = rms_norm_double_round(x, sd[prefix + "norm.weight"], eps)
    return x.transpose(1, 2)


def upsample(c: DACConfig, sd: Dict[str, Tensor], z: Tensor) -> Tensor:
    """rvq.py:263-276: [CausalTransConvNet(k=s=f), ConvNeXtBlock] per factor."""
    for i, f in enumerate(c.upsample_factors):
        p = f"quantizer.upsample.{i}."
        z = causal_trans_conv(z, sd[p + "0.conv.weight"], sd[p + "0.conv.bias"], f)
        inp = z  # ConvNeXtBlock.forward, rvq.py:173-191
        x = causal_conv(z, sd[p + "1.dwconv.conv.weight"], sd[p + "1.dwconv.conv.bias"], groups=z.shape[1])
        x = x.permute(0, 2, 1)
        x = F.layer_norm(x, (x.shape[-1],), sd[p + "1.norm.weight"], sd[p + "1.norm.bias"], 1e-6)
        x = F.linear(x, sd[p + "1.pwconv1.weight"], sd[p + "1.pwconv1.bias"])
        x = F.gelu(x)
        x = F.linear(x, sd[p + "1.pwconv2.weight"], sd[p + "1.pwconv2.bias"])
        x = sd[p + "1.gamma"] * x
        z = inp + x.permute(0, 2, 1)
    return z


def quantizer_decode(c: DACConfig, sd, indices: Tensor) -> Tensor:
    """DownsampleResidualVectorQuantize.decode (rvq.py:352-366)."""
    return upsample(c, sd, post_module(c, sd, rvq_dequant(c, sd, indices)))


def residual_unit(sd, p: str, x: Tensor, dilation: int) -> Tensor:
    """ResidualUnit (modded_dac.py:599-620), causal: snake, conv k7 dil d, snake, conv k1, +x."""
    y = snake(x, sd[p + "0.alpha"])
    y = causal_conv(y, _wn(sd, p + "1."), sd[p + "1.conv.bias"], dilation=dilation)
    y = snake(y, sd[p + "2.alpha"])
    y = causal_conv(y, _wn(sd, p + "3."), sd[p + "3.conv.bias"])
    return x + y


def decoder(c: DACConfig, sd, z: Tensor, stages: List[Tensor] | None = None) -> Tensor:
    """Decoder.forward (modded_dac.py:760-801); DecoderBlock (:712-757; transformer unused, :741-742)."""
    x = causal_conv(z, _wn(sd, "decoder.model.0."), sd["decoder.model.0.conv.bias"])
    if stages is not None:
        stages.append(x)
    for i, s in enumerate(c.decoder_rates):
        p = f"decoder.model.{i + 1}.block."
        x = snake(x, sd[p + "0.alpha"])
        x = causal_trans_conv(x, _wn(sd, p + "1."), sd[p + "1.conv.bias"], s)
        for j, d in enumerate((1, 3, 9)):
            x = residual_unit(sd, p + f"{2 + j}.block.", x, d)
        if stages is not None:
            stages.append(x)
    n = len(c.decoder_rates)
    x = snake(x, sd[f"decoder.model.{n + 1}.alpha"])
    x = causal_conv(x, _wn(sd, f"decoder.model.{n + 2}."), sd[f"decoder.model.{n + 2}.conv.bias"])
    return torch.tanh(x)


@torch.inference_mode()
def from_indices(c: DACConfig, sd, indices: Tensor, stages: List[Tensor] | None = None) -> Tensor:
    """DAC.from_indices (modded_dac.py:925-927): [B, 1+n_codebooks, T] int64 -> [B, 1, hop*T]."""
    z = quantizer_decode(c, sd, indices)
    if stages is not None:
        stages.append(z)
    return decoder(c, sd, z, stages)


# ----------------------------------------------------------------------------- encode side (DAC.encode, modded_dac.py:874-923)
def strided_causal_conv(x: Tensor, w: Tensor, b: Tensor, stride: int) -> Tensor:
    """CausalConvNet.forward with a stride (modded_dac.py:546-552 / rvq.py:83-89): k - s zeros on the left, and on the right
    what get_extra_padding_for_conv1d (modded_dac.py:485-492) asks for to end on a whole output frame."""
    k = w.shape[-1]
    pad = k - stride
    length = x.shape[-1]
    n_frames = (length - k + pad) / stride + 1
    ideal = (math.ceil(n_frames) - 1) * stride + (k - pad)
    x = F.pad(x, (pad, ideal - length), mode="constant", value=0)
    return F.conv1d(x, w, b, stride=stride).contiguous()


def encoder(c: DACConfig, sd, audio: Tensor, stages: List[Tensor] | None = None) -> Tensor:
    """Encoder.forward (modded_dac.py:670-709) over EncoderBlock (:623-667): three residual units, Snake, strided conv, and
    the window transformer where encoder_tf_layers says so; then Snake and a 3-tap conv to the latent width."""
    x = causal_conv(audio, _wn(sd, "encoder.block.0."), sd["encoder.block.0.conv.bias"])
    ch = c.encoder_dim
    for i, (s, nl) in enumerate(zip(c.encoder_rates, c.encoder_tf_layers)):
        p = f"encoder.block.{i + 1}.block."
        for j, d in enumerate((1, 3, 9)):
            x = residual_unit(sd, p + f"{j}.block.", x, d)
        x = snake(x, sd[p + "3.alpha"])
        x = strided_causal_conv(x, _wn(sd, p + "4."), sd[p + "4.conv.bias"], s)
        ch *= 2
        if nl:
            x = window_transformer(sd, p + "5.", x, nl, ch // c.encoder_tf_head_dim, c.encoder_tf_head_dim,
                                   c.encoder_tf_window, 1e-5, 10000.0)
        if stages is not None:
            stages.append(x)
    n = len(c.encoder_rates)
    x = snake(x, sd[f"encoder.block.{n + 1}.alpha"])
    return causal_conv(x, _wn(sd, f"encoder.block.{n + 2}."), sd[f"encoder.block.{n + 2}.conv.bias"])


def downsample(c: DACConfig, sd, z: Tensor) -> Tensor:
    """rvq.py:249-261: [CausalConvNet(k = s = f), ConvNeXtBlock] per factor."""
    for i, f in enumerate(c.upsample_factors):
        p = f"quantizer.downsample.{i}."
        z = strided_causal_conv(z, sd[p + "0.conv.weight"], sd[p + "0.conv.bias"], f)
        inp = z  # ConvNeXtBlock.forward, rvq.py:173-191
        x = causal_conv(z, sd[p + "1.dwconv.conv.weight"], sd[p + "1.dwconv.conv.bias"], groups=z.shape[1])
        x = x.permute(0, 2, 1)
        x = F.layer_norm(x, (x.shape[-1],), sd[p + "1.norm.weight"], sd[p + "1.norm.bias"], 1e-6)
        x = F.linear(x, sd[p + "1.pwconv1.weight"], sd[p + "1.pwconv1.bias"])
        x = F.gelu(x)
        x = F.linear(x, sd[p + "1.pwconv2.weight"], sd[p + "1.pwconv2.bias"])
        x = sd[p + "1.gamma"] * x
        z = inp + x.permute(0, 2, 1)
    return z


def vq_step(sd, p: str, residual: Tensor) -> Tuple[Tensor, Tensor]:
    """descript-audio-codec 1.0.0 VectorQuantize.forward in eval mode: project to codebook_dim, nearest entry by distance
    between L2-normalised vectors, entry (not normalised) through the straight-through sum and out_proj.
    Returns (z_q_i [B, D, T], indices [B, T])."""
    z_e = F.conv1d(residual, torch._weight_norm(sd[p + "in_proj.weight_v"], sd[p + "in_proj.weight_g"], 0), sd[p + "in_proj.bias"])
    B, Dc, T = z_e.shape
    enc = F.normalize(z_e.permute(0, 2, 1).reshape(B * T, Dc))
    cb = F.normalize(sd[p + "codebook.weight"])
    dist = enc.pow(2).sum(1, keepdim=True) - 2 * enc @ cb.t() + cb.pow(2).sum(1, keepdim=True).t()
    idx = (-dist).max(1)[1].reshape(B, T)
    z_q = F.embedding(idx, sd[p + "codebook.weight"]).transpose(1, 2)
    z_q = z_e + (z_q - z_e)                                  # the straight-through sum: it rounds, so it is kept
    z_q = F.conv1d(z_q, torch._weight_norm(sd[p + "out_proj.weight_v"], sd[p + "out_proj.weight_g"], 0), sd[p + "out_proj.bias"])
    return z_q, idx


def quantizer_encode(c: DACConfig, sd, z: Tensor, stages: List[Tensor] | None = None) -> Tensor:
    """DownsampleResidualVectorQuantize.forward up to the codes (rvq.py:293-318): downsample, pre_module, the semantic
    codebook on z, the residual codebooks on what it leaves.  -> [B, 1 + n_codebooks, T] int64."""
    z = downsample(c, sd, z)
    z = window_transformer(sd, "quantizer.pre_module.", z, c.tf_layers, c.tf_heads, c.tf_head_dim, c.tf_window,
                           c.tf_norm_eps, c.tf_rope_base)
    if stages is not None:
        stages.append(z)
    codes = []
    z_q, idx = vq_step(sd, "quantizer.semantic_quantizer.quantizers.0.", z)
    codes.append(idx)
    residual = z - z_q
    for i in range(c.n_codebooks):
        z_q, idx = vq_step(sd, f"quantizer.quantizer.quantizers.{i}.", residual)
        residual = residual - z_q
        codes.append(idx)
    return torch.stack(codes, dim=1)


@torch.inference_mode()
def encode(c: DACConfig, sd, audio: Tensor, stages: List[Tensor] | None = None) -> Tuple[Tensor, Tensor]:
    """DAC.encode (modded_dac.py:874-923): [B, 1, n] (or [B, n]) waveform -> ([B, 1 + n_codebooks, ceil(n / hop)] int64 codes,
    their lengths): zeros on the right up to a whole code frame, encoder, quantizer."""
    if audio.dim() == 2:
        audio = audio.unsqueeze(1)
    length = audio.shape[-1]
    right = math.ceil(length / c.hop) * c.hop - length
    audio = F.pad(audio, (0, right))
    z = encoder(c, sd, audio, stages)
    if stages is not None:
        stages.append(z)
    codes = quantizer_encode(c, sd, z, stages)
    return codes, torch.ceil(torch.tensor([length + right]) / c.hop).long()
