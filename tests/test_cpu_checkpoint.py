"""Host-side checkpoint ingestion (fish_speech_b200/checkpoint.py) on synthetic checkpoints written by the test: config.json of
model_type fish_qwen3_omni / dual_ar, sharded safetensors with the published key prefixes and separate wq / wk / wv, model.pth,
tokenizer.json; cross-checked against the reference's own functions when /root/reference is importable (this container)."""
import json
import sys
import types
from collections import OrderedDict
from pathlib import Path

import pytest
import torch

from fish_speech_b200 import checkpoint as CK
from oracle import dual_ar_ref as R

REF = Path("/root/reference")


def _tiny():
    return R.tiny_config(n_layer=2, n_fast_layer=1, vocab_size=2048, dim=256, n_head=4, n_local_heads=2, head_dim=64,
                         intermediate_size=512, codebook_size=64, num_codebooks=4, semantic_begin_id=1024, semantic_end_id=1087,
                         im_end_id=5, max_seq_len=128)


def _omni_config(c):
    return {"model_type": "fish_qwen3_omni", "semantic_start_token_id": 3, "semantic_end_token_id": 4,
            "text_config": {"vocab_size": c.vocab_size, "n_layer": c.n_layer, "n_head": c.n_head, "n_local_heads": c.n_local_heads,
                            "head_dim": c.head_dim, "dim": c.dim, "intermediate_size": c.intermediate_size, "rope_base": 1000000,
                            "norm_eps": 1e-6, "max_seq_len": c.max_seq_len, "attention_qk_norm": True, "tie_word_embeddings": True},
            "audio_decoder_config": {"vocab_size": c.codebook_size, "num_codebooks": c.num_codebooks, "n_layer": c.n_fast_layer,
                                     "dim": c.dim, "n_head": c.n_head, "n_local_heads": c.n_local_heads, "head_dim": c.head_dim,
                                     "intermediate_size": c.intermediate_size, "attention_qk_norm": True}}


def _to_omni_keys(c, sd):
    """inverse of the reference's remap + load hook: published names, separate q / k / v"""
    out = OrderedDict()
    q, kv = c.n_head * c.head_dim, c.n_local_heads * c.head_dim
    for k, v in sd.items():
        parts = {k: v}
        if k.endswith("attention.wqkv.weight"):
            p = k[: -len("wqkv.weight")]
            parts = {p + "wq.weight": v[:q].clone(), p + "wk.weight": v[q: q + kv].clone(), p + "wv.weight": v[q + kv:].clone()}
        for kk, vv in parts.items():
            if kk.startswith("fast_"):
                out["audio_decoder." + kk[len("fast_"):]] = vv
            elif kk.startswith("codebook_embeddings."):
                out["audio_decoder." + kk] = vv
            else:
                out["text_model.model." + kk] = vv
    return out


def _write_tokenizer(path, c):
    added = [{"id": c.im_end_id, "content": "<|im_end|>"}] + \
            [{"id": c.semantic_begin_id + i, "content": f"<|semantic:{i}|>"} for i in range(c.semantic_end_id - c.semantic_begin_id + 1)]
    (path / "tokenizer.json").write_text(json.dumps({"model": {"vocab": {"a": 0, "b": 1}}, "added_tokens": added}))


def write_omni_checkpoint(path: Path, c, sd, shards=2):
    from safetensors.torch import save_file
    path.mkdir(parents=True, exist_ok=True)
    (path / "config.json").write_text(json.dumps(_omni_config(c)))
    _write_tokenizer(path, c)
    w = _to_omni_keys(c, sd)
    keys = list(w)
    wm = {}
    for s in range(shards):
        name = f"model-{s + 1:05d}-of-{shards:05d}.safetensors"
        part = {k: w[k].contiguous() for k in keys[s::shards]}
        save_file(part, str(path / name))
        wm.update({k: name for k in part})
    (path / "model.safetensors.index.json").write_text(json.dumps({"weight_map": wm}))


def test_fish_qwen3_omni_sharded_checkpoint_round_trip(tmp_path):
    c = _tiny()
    sd = R.make_state_dict(c, seed=4)
    write_omni_checkpoint(tmp_path, c, sd)
    cfg = CK.load_config(tmp_path)
    for k in ("vocab_size", "n_layer", "n_fast_layer", "n_head", "n_local_heads", "head_dim", "dim", "intermediate_size",
              "codebook_size", "num_codebooks", "max_seq_len", "attention_qk_norm", "tie_word_embeddings"):
        assert getattr(cfg, k) == getattr(c, k), k
    assert cfg.scale_codebook_embeddings and cfg.norm_fastlayer_input          # forced by _from_fish_qwen3_omni
    assert (cfg.semantic_begin_id, cfg.semantic_end_id) == (3, 4)              # config.json values ...
    ids = CK.read_tokenizer_ids(tmp_path)
    assert ids == {"semantic_begin_id": c.semantic_begin_id, "semantic_end_id": c.semantic_end_id, "im_end_id": c.im_end_id}  # ... the tokenizer's win
    w = CK.load_weights(tmp_path)
    assert set(w) == set(sd)
    for k in sd:
        assert torch.equal(w[k], sd[k]), k


def test_defaults_pth_and_rejections(tmp_path):
    c = _tiny()
    sd = R.make_state_dict(c, seed=5)
    # dual_ar config with the fields __post_init__ derives left out; model.pth with "state_dict" + "model." prefix + audio_* keys
    (tmp_path / "config.json").write_text(json.dumps({"model_type": "dual_ar", "vocab_size": 2048, "n_layer": 2, "n_head": 4, "dim": 256,
                                                      "codebook_size": 64, "num_codebooks": 4, "n_fast_layer": 1, "unknown_key": 1}))
    cfg = CK.load_config(tmp_path)
    assert cfg.n_local_heads == 4 and cfg.head_dim == 64 and cfg.intermediate_size == 768     # find_multiple(int(2 * 1024 / 3), 256)
    assert cfg.norm_eps == 1e-5 and cfg.rope_base == 10000                                   # the reference's defaults
    pth = {"state_dict": OrderedDict([("model." + k, v) for k, v in sd.items()] + [("model.audio_projector.weight", torch.zeros(1))])}
    torch.save(pth, tmp_path / "model.pth")
    w = CK.load_weights(tmp_path)
    assert set(w) == set(sd) and all(torch.equal(w[k], sd[k]) for k in sd)
    with pytest.raises(ValueError):
        (tmp_path / "config.json").write_text(json.dumps({"model_type": "naive"}))
        CK.load_config(tmp_path)
    bad = _omni_config(c)
    bad["audio_decoder_config"]["dim"] = 128
    (tmp_path / "config.json").write_text(json.dumps(bad))
    with pytest.raises(ValueError):
        CK.load_config(tmp_path)
    with pytest.raises(FileNotFoundError):
        CK.load_weights(tmp_path / "nowhere")


def test_codec_state_dict_normalisation():
    g, v = torch.rand(4, 1, 1), torch.rand(4, 3, 7)
    sd = {"state_dict": {"generator.decoder.model.0.conv.weight_g": g, "generator.decoder.model.0.conv.weight_v": v,
                         "generator.quantizer.quantizer.quantizers.0.out_proj.parametrizations.weight.original0": g,
                         "discriminator.x": g}}
    out = CK.normalise_codec_state_dict(sd)
    assert set(out) == {"decoder.model.0.conv.parametrizations.weight.original0", "decoder.model.0.conv.parametrizations.weight.original1",
                        "quantizer.quantizer.quantizers.0.out_proj.weight_g"}
    # the encode side's keys take the same two spellings, and a whole encode-side state dict is recognised as one
    enc = {"generator.encoder.block.0.conv.weight_g": g, "generator.encoder.block.0.conv.weight_v": v,
           "generator.quantizer.semantic_quantizer.quantizers.0.in_proj.parametrizations.weight.original1": v}
    out = CK.normalise_codec_state_dict(enc)
    assert set(out) == {"encoder.block.0.conv.parametrizations.weight.original0", "encoder.block.0.conv.parametrizations.weight.original1",
                        "quantizer.semantic_quantizer.quantizers.0.in_proj.weight_v"}
    from fish_speech_b200.dac import DACConfig, encode_state_dict_shapes, has_encode_side, state_dict_shapes
    c = DACConfig()
    full = {**state_dict_shapes(c), **encode_state_dict_shapes(c)}
    assert has_encode_side(full) and not has_encode_side(state_dict_shapes(c))
    assert full["encoder.block.4.block.4.conv.parametrizations.weight.original1"] == (1024, 512, 16)       # modded_dac.py:655-663
    assert full["encoder.block.4.block.5.layers.3.attention.wqkv.weight"] == (3072, 1024)                    # :636-650
    assert full["quantizer.downsample.1.0.conv.weight"] == (1024, 1024, 2)                                    # rvq.py:249-261


@pytest.mark.skipif(not REF.exists(), reason="the reference tree is only in the build container")
def test_against_the_reference_functions(tmp_path):
    """Same config.json and keys through the reference's BaseModelArgs.from_pretrained / _remap_fish_qwen3_omni_keys."""
    sys.path.insert(0, str(REF))
    sys.modules.setdefault("loralib", types.ModuleType("loralib"))
    try:
        from fish_speech.models.text2semantic import llama as L
    finally:
        sys.path.remove(str(REF))
    c = _tiny()
    sd = R.make_state_dict(c, seed=6)
    write_omni_checkpoint(tmp_path, c, sd)
    ref = L.BaseModelArgs.from_pretrained(str(tmp_path))
    mine = CK.load_config(tmp_path)
    for k in ("vocab_size", "n_layer", "n_fast_layer", "n_head", "n_local_heads", "head_dim", "dim", "intermediate_size", "rope_base",
              "norm_eps", "max_seq_len", "codebook_size", "num_codebooks", "semantic_begin_id", "semantic_end_id",
              "attention_qk_norm", "tie_word_embeddings", "scale_codebook_embeddings", "norm_fastlayer_input"):
        assert getattr(mine, k) == getattr(ref, k), k
    omni = _to_omni_keys(c, sd)
    assert list(CK.remap_fish_qwen3_omni_keys(OrderedDict(omni))) == list(L._remap_fish_qwen3_omni_keys(OrderedDict(omni)))
