"""Checkpoint ingestion: the reference's `init_model(checkpoint_path, device, precision, compile)` and codec `load_model` over
the CUDA engine, so that the reference-side binding is a one-line import.

Follows, host-side only (no arithmetic):
  fish_speech/models/text2semantic/llama.py:76-143    BaseModelArgs.from_pretrained / _from_fish_qwen3_omni (config.json)
  fish_speech/models/text2semantic/llama.py:64-73     __post_init__ defaults (n_local_heads, intermediate_size, head_dim)
  fish_speech/models/text2semantic/llama.py:229-246   _remap_fish_qwen3_omni_keys
  fish_speech/models/text2semantic/llama.py:545-593   sharded / single safetensors, model.pth (state_dict, "model." prefix, audio_*)
  fish_speech/models/text2semantic/llama.py:876-881   Attention.load_hook: wq / wk / wv -> wqkv
  fish_speech/tokenizer.py:56-80                      semantic token ids and <|im_end|> from the tokenizer's vocabulary
  fish_speech/models/text2semantic/inference.py:362-392  init_model
  fish_speech/models/dac/inference.py:23-48           codec load_model ("state_dict", "generator." prefix)
"""
from __future__ import annotations

import json
from collections import OrderedDict
from pathlib import Path
from typing import Dict, Optional, Tuple

import torch
from torch import Tensor

from .dual_ar import DualARConfig, DualARTransformerB200, decode_one_token_ar, find_multiple

IM_END_TOKEN = "<|im_end|>"                    # tokenizer.py:18
SEMANTIC_TOKEN_TEMPLATE = "<|semantic:{i}|>"   # tokenizer.py:36


# ----------------------------------------------------------------------------------------------- config.json
def _flatten_fish_qwen3_omni(data: dict) -> dict:
    """llama.py:101-143: text_config / audio_decoder_config -> the flat DualARModelArgs fields."""
    tc, adc = data["text_config"], data["audio_decoder_config"]
    flat = dict(
        model_type="dual_ar", vocab_size=tc["vocab_size"], n_layer=tc["n_layer"], n_head=tc["n_head"],
        n_local_heads=tc.get("n_local_heads", -1), head_dim=tc.get("head_dim"), dim=tc["dim"],
        intermediate_size=tc.get("intermediate_size"), rope_base=tc.get("rope_base", 10000),
        norm_eps=tc.get("norm_eps", 1e-5), max_seq_len=tc.get("max_seq_len", 2048),
        tie_word_embeddings=tc.get("tie_word_embeddings", True),
        attention_qkv_bias=tc.get("attention_qkv_bias", False), attention_o_bias=tc.get("attention_o_bias", False),
        attention_qk_norm=tc.get("attention_qk_norm", False), initializer_range=tc.get("initializer_range", 0.02),
        semantic_begin_id=data.get("semantic_start_token_id", 0), semantic_end_id=data.get("semantic_end_token_id", 0),
        scale_codebook_embeddings=True, norm_fastlayer_input=True,
        codebook_size=adc["vocab_size"], num_codebooks=adc["num_codebooks"], n_fast_layer=adc["n_layer"],
        fast_dim=adc.get("dim"), fast_n_head=adc.get("n_head"), fast_n_local_heads=adc.get("n_local_heads"),
        fast_head_dim=adc.get("head_dim"), fast_intermediate_size=adc.get("intermediate_size"),
        fast_attention_qkv_bias=adc.get("attention_qkv_bias"), fast_attention_qk_norm=adc.get("attention_qk_norm"),
        fast_attention_o_bias=adc.get("attention_o_bias"),
    )
    return {k: v for k, v in flat.items() if v is not None}


def load_config(path) -> DualARConfig:
    """BaseModelArgs.from_pretrained (llama.py:76-99) for `dual_ar` and `fish_qwen3_omni` configs, reduced to the fields the
    decode path reads.  Raises for what the engine does not implement instead of loading it wrongly."""
    path = Path(path)
    if path.is_dir():
        path = path / "config.json"
    data = json.loads(path.read_text(encoding="utf-8"))
    mt = data.get("model_type")
    if mt == "fish_qwen3_omni":
        data = _flatten_fish_qwen3_omni(data)
    elif mt != "dual_ar":
        raise ValueError(f"Unknown model type: {mt}")  # llama.py:96 (naive models have no fast AR: not this engine)
    d = dict(data)
    # __post_init__ (llama.py:64-73)
    if d.get("n_local_heads", -1) == -1:
        d["n_local_heads"] = d["n_head"]
    if d.get("intermediate_size") is None:
        d["intermediate_size"] = find_multiple(int(2 * (4 * d["dim"]) / 3), 256)
    if d.get("head_dim") is None:
        d["head_dim"] = d["dim"] // d["n_head"]
    # DualARModelArgs fast_* default to the slow values (llama.py:100-123); the engine runs one geometry for both stacks
    for fk, sk in (("fast_dim", "dim"), ("fast_n_head", "n_head"), ("fast_n_local_heads", "n_local_heads"), ("fast_head_dim", "head_dim"),
                   ("fast_intermediate_size", "intermediate_size"), ("fast_attention_qk_norm", "attention_qk_norm")):
        if d.get(fk) is not None and d[fk] != d[sk]:
            raise ValueError(f"{fk}={d[fk]} differs from {sk}={d[sk]}: the engine runs the fast stack at the slow stack's geometry")
    for k in ("attention_qkv_bias", "attention_o_bias", "fast_attention_qkv_bias", "fast_attention_o_bias"):
        if d.get(k):
            raise ValueError(f"{k}=True is not implemented by the engine (S2-Pro has no attention biases)")
    cfg = DualARConfig.from_any({k: v for k, v in d.items()})
    if "norm_eps" not in d:
        cfg.norm_eps = 1e-5     # the reference's defaults (llama.py:41-42), not this mirror's
    if "rope_base" not in d:
        cfg.rope_base = 10000
    return cfg


def read_tokenizer_ids(path) -> Optional[Dict[str, int]]:
    """semantic_begin_id / semantic_end_id / im_end_id as FishTokenizer derives them from the vocabulary (tokenizer.py:56-80,
    inference.py:207), read from the checkpoint's tokenizer.json without instantiating a tokenizer."""
    tj = Path(path) / "tokenizer.json"
    if not tj.exists():
        return None
    data = json.loads(tj.read_text(encoding="utf-8"))
    vocab: Dict[str, int] = dict(data.get("model", {}).get("vocab", {}) or {})
    for t in data.get("added_tokens", []) or []:
        vocab[t["content"]] = t["id"]
    sem = [vocab[SEMANTIC_TOKEN_TEMPLATE.format(i=i)] for i in range(4096) if SEMANTIC_TOKEN_TEMPLATE.format(i=i) in vocab]
    out: Dict[str, int] = {}
    if sem:
        out["semantic_begin_id"], out["semantic_end_id"] = min(sem), max(sem)
    if IM_END_TOKEN in vocab:
        out["im_end_id"] = vocab[IM_END_TOKEN]
    return out or None


# ----------------------------------------------------------------------------------------------- weights
def remap_fish_qwen3_omni_keys(weights: "OrderedDict[str, Tensor]") -> "OrderedDict[str, Tensor]":
    """llama.py:229-246."""
    if not any(k.startswith(("text_model.", "audio_decoder.")) for k in weights):
        return weights
    out: "OrderedDict[str, Tensor]" = OrderedDict()
    for k, v in weights.items():
        if k.startswith("text_model.model."):
            nk = k[len("text_model.model."):]
        elif k.startswith("audio_decoder."):
            suffix = k[len("audio_decoder."):]
            nk = suffix if suffix.startswith("codebook_embeddings.") else "fast_" + suffix
        else:
            nk = k
        out[nk] = v
    return out


def fuse_qkv(weights: "OrderedDict[str, Tensor]") -> "OrderedDict[str, Tensor]":
    """Attention.load_hook (llama.py:876-881): separate wq / wk / wv -> one wqkv = cat([wq, wk, wv])."""
    for k in [k for k in weights if k.endswith("attention.wq.weight")]:
        p = k[: -len("wq.weight")]
        weights[p + "wqkv.weight"] = torch.cat([weights.pop(p + "wq.weight"), weights.pop(p + "wk.weight"), weights.pop(p + "wv.weight")])
    return weights


def load_weights(path) -> "OrderedDict[str, Tensor]":
    """The weight files of BaseTransformer.from_pretrained (llama.py:545-593), in its order of preference."""
    p = Path(path)
    index_json, single_st, pth = p / "model.safetensors.index.json", p / "model.safetensors", p / "model.pth"
    if index_json.exists():
        from safetensors.torch import load_file
        shards = sorted(set(json.loads(index_json.read_text())["weight_map"].values()))
        weights: "OrderedDict[str, Tensor]" = OrderedDict()
        for shard in shards:
            weights.update(load_file(str(p / shard), device="cpu"))
        weights = remap_fish_qwen3_omni_keys(weights)
    elif single_st.exists():
        from safetensors.torch import load_file
        weights = remap_fish_qwen3_omni_keys(OrderedDict(load_file(str(single_st), device="cpu")))
    elif pth.exists():
        weights = torch.load(pth, map_location="cpu", mmap=True, weights_only=True)
        if "state_dict" in weights:
            weights = weights["state_dict"]
        if weights and next(iter(weights.keys())).startswith("model."):
            weights = OrderedDict((k.replace("model.", ""), v) for k, v in weights.items())
        for k in list(weights.keys()):
            if "audio_" in k:
                weights.pop(k)
        weights = OrderedDict(weights)
    else:
        raise FileNotFoundError(f"No model weights found in {p}")
    return fuse_qkv(weights)


def init_model(checkpoint_path, device, precision=torch.bfloat16, compile: bool = False, max_rows: int = 1,
               max_length: Optional[int] = None) -> Tuple[DualARTransformerB200, object]:
    """inference.py:362-392 on the engine: returns (model, decode_one_token) with the reference's signature.  `compile` is
    accepted and ignored (there is no tracing compiler on this path: the frame is already one CUDA-graph node)."""
    if precision not in (torch.bfloat16, "bf16", "bfloat16"):
        raise ValueError("the engine computes in bf16 (the reference CLI's default, inference.py:860)")
    cfg = load_config(checkpoint_path)
    if max_length is not None:
        cfg.max_seq_len = int(max_length)  # llama.py:493-495
    ids = read_tokenizer_ids(checkpoint_path)
    if ids:  # llama.py:501-507: the tokenizer's ids win over config.json
        for k, v in ids.items():
            setattr(cfg, k, v)
    weights = load_weights(checkpoint_path)
    sd = {k: v for k, v in weights.items() if not k.endswith(("freqs_cis", "causal_mask"))}
    model = DualARTransformerB200(cfg, sd, device=device, max_rows=max_rows)
    model.fixed_temperature = torch.tensor(0.7, device=device, dtype=torch.float)   # inference.py:376-378
    model.fixed_top_p = torch.tensor(0.7, device=device, dtype=torch.float)
    model.fixed_repetition_penalty = torch.tensor(1.5, device=device, dtype=torch.float)
    model._cache_setup_done = False
    return model, decode_one_token_ar


# ----------------------------------------------------------------------------------------------- codec
def normalise_codec_state_dict(sd: Dict[str, Tensor]) -> Dict[str, Tensor]:
    """dac/inference.py:29-42 ("state_dict" wrapper, "generator." prefix) plus both weight-norm spellings: the decoder's and the encoder's convs
    are read as `...conv.parametrizations.weight.original0/1`, the descript quantizers' projections as `...weight_g/weight_v`
    (fish_speech_b200/dac.py: pack_weights); a checkpoint saved in the other spelling is renamed, never re-computed."""
    if "state_dict" in sd:
        sd = sd["state_dict"]
    if any("generator" in k for k in sd):
        sd = {k.replace("generator.", ""): v for k, v in sd.items() if "generator." in k}
    out: Dict[str, Tensor] = {}
    for k, v in sd.items():
        if k.startswith(("decoder.", "encoder.")) and k.endswith(".weight_g"):
            k = k[: -len("weight_g")] + "parametrizations.weight.original0"
        elif k.startswith(("decoder.", "encoder.")) and k.endswith(".weight_v"):
            k = k[: -len("weight_v")] + "parametrizations.weight.original1"
        elif "_proj.parametrizations.weight.original0" in k:       # the quantizers' in_proj / out_proj
            k = k.replace("parametrizations.weight.original0", "weight_g")
        elif "_proj.parametrizations.weight.original1" in k:
            k = k.replace("parametrizations.weight.original1", "weight_v")
        out[k] = v
    return out


def load_codec(checkpoint_path, device="cuda", config=None, **kw):
    """dac/inference.py:23-48 on the engine: codec.pth -> DACB200 with `from_indices`, `sample_rate`, `device`, and `encode`
    when the file carries the encode side (encoder.*, quantizer.downsample.*, quantizer.pre_module.*, in_proj.*: the released
    codec.pth does; `can_encode` says).  `config`: the modded_dac_vq.yaml values (default: fish_speech_b200.dac.DACConfig,
    which restates that file)."""
    from .dac import DACB200, DACConfig
    sd = torch.load(checkpoint_path, map_location="cpu", mmap=True, weights_only=True)
    return DACB200(config or DACConfig(), normalise_codec_state_dict(sd), device=device, **kw)
