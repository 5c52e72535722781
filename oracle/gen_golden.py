"""Generate tests/golden/*.pt by running THE REFERENCE ITSELF (imported unmodified
from /root/reference, CPU) and assert the oracle restatement is bit-identical.

Run in the build container only (the GPU box has no /root/reference):

    python -m oracle.gen_golden            # writes tests/golden/*.pt

TEST INFRASTRUCTURE — see oracle/__init__.py.
"""
from __future__ import annotations

import os
import sys
import types
from pathlib import Path

import torch

from oracle import dual_ar_ref as R

REF = os.environ.get("FISH_REFERENCE", "/root/reference")
GOLD = Path(__file__).resolve().parent.parent / "tests" / "golden"


def import_reference():
    """stub `loralib` (imported at lora.py:3, never used at inference)."""
    if REF not in sys.path:
        sys.path.insert(0, REF)
    sys.modules.setdefault("loralib", types.ModuleType("loralib"))
    shim = str(Path(__file__).resolve().parent / "refshim")
    if shim not in sys.path:
        sys.path.insert(0, shim)
    import fish_speech.models.text2semantic.inference as inf
    import fish_speech.models.text2semantic.llama as llama
    inf.tqdm = lambda it, *a, **k: it
    return llama, inf


class _Tok:
    def __init__(self, cfg):
        self.cfg = cfg
        self.semantic_begin_id, self.semantic_end_id = cfg.semantic_begin_id, cfg.semantic_end_id

    def get_token_id(self, _name):
        return self.cfg.im_end_id


def build_reference_model(cfg: R.DualARConfig, sd):
    llama, _ = import_reference()
    args = llama.DualARModelArgs(**cfg.to_ref_kwargs())
    m = llama.DualARTransformer(args).eval().to(torch.bfloat16)
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected and all(k.endswith(("freqs_cis", "causal_mask")) for k in missing), (missing, unexpected)
    m.tokenizer = _Tok(cfg)
    return m


class UniformFeeder:
    """Deterministic replacement for torch.rand_like: bf16 uniforms from a seeded
    generator, recorded so that the CUDA path can be fed the very same draws."""

    def __init__(self, seed: int):
        self.g = torch.Generator().manual_seed(seed)
        self.draws = []

    def __call__(self, t):
        if t.dtype == torch.bfloat16:
            # torch's bf16 uniform has 8 random bits: multiples of 1/256 in [0, 1)
            u = (torch.randint(0, 256, t.shape, generator=self.g).float() / 256.0).to(t.dtype)
        else:
            u = torch.rand(t.shape, generator=self.g, dtype=torch.float32).to(t.dtype)
        self.draws.append(u)
        return u


def golden_dual_ar(name: str, cfg: R.DualARConfig, T: int, frames: int, voice_span: int,
                   top_k: int, temperature: float, top_p: float, wseed: int = 0, useed: int = 1234,
                   keep_logits_every: int = 16):
    _, inf = import_reference()
    sd = R.make_state_dict(cfg, seed=wseed)
    prompt = R.make_prompt(cfg, T, seed=1, voice_span=voice_span)

    # --- reference run, hooks capture logits; uniforms injected via torch.rand_like patch
    ref = build_reference_model(cfg, sd)
    ref_tr = []
    orig_fg, orig_ff = ref.forward_generate, ref.forward_generate_fast

    def fg(*a, **k):
        r = orig_fg(*a, **k)
        ref_tr.append({"slow_logits": r.logits[0, -1].clone(), "hidden": r.hidden_states[0, -1].clone(),
                       "fast_logits": []})
        return r

    def ff(x, pos):
        r = orig_ff(x, pos)
        if int(pos[0]) > 0:
            ref_tr[-1]["fast_logits"].append(r[0, -1].clone())
        return r

    ref.forward_generate, ref.forward_generate_fast = fg, ff
    feeder = UniformFeeder(useed)
    real = torch.rand_like
    torch.rand_like = lambda t, **k: feeder(t)
    try:
        with torch.inference_mode():
            ref_seq = inf.generate(model=ref, prompt=prompt.clone(), max_new_tokens=frames, audio_masks=None,
                                   audio_parts=None, temperature=temperature, top_p=top_p, top_k=top_k)
    finally:
        torch.rand_like = real

    # --- oracle run with the same uniforms
    feeder2 = UniformFeeder(useed)
    om = R.DualARRef(cfg, sd)
    tr = []
    seq = R.generate(model=om, prompt=prompt.clone(), max_new_tokens=frames, temperature=temperature,
                     top_p=top_p, top_k=top_k, rand_like=feeder2, traces=tr)
    assert torch.equal(ref_seq, seq), f"{name}: oracle sequence differs from the reference"
    assert len(tr) == len(ref_tr)
    for a, b in zip(tr, ref_tr):
        assert torch.equal(a["slow_logits"], b["slow_logits"]), name
        assert torch.equal(a["hidden"], b["hidden"]), name
        for x, y in zip(a["fast_logits"], b["fast_logits"]):
            assert torch.equal(x, y), name
    print(f"[golden] {name}: oracle == reference bit-exact over {len(tr)} frames; seq {tuple(seq.shape)}")

    keep = sorted(set(list(range(0, len(tr), keep_logits_every)) + [len(tr) - 1]))
    out = {
        "cfg": cfg.__dict__, "wseed": wseed, "useed": useed, "prompt": prompt, "seq": ref_seq.clone(),
        "sampling": {"temperature": temperature, "top_p": top_p, "top_k": top_k},
        "frames_kept": keep,
        "slow_logits": {i: ref_tr[i]["slow_logits"] for i in keep},
        "fast_logits": {i: torch.stack(ref_tr[i]["fast_logits"]) for i in keep},
        # top-2 margin of every greedy decision (slow biased + 9 fast), for near-tie aware comparison
        "weights_checksum": float(sum(v.float().abs().sum() for v in sd.values())),
    }
    bias = R.make_semantic_logit_bias(cfg)[0, 0].float()
    margins = []
    for t in ref_tr:
        top2 = torch.topk(t["slow_logits"].float() + bias, 2).values
        row = [float(top2[0] - top2[1])]
        for fl in t["fast_logits"]:
            t2 = torch.topk(fl.float(), 2).values
            row.append(float(t2[0] - t2[1]))
        margins.append(row)
    out["margins"] = torch.tensor(margins)
    GOLD.mkdir(parents=True, exist_ok=True)
    torch.save(out, GOLD / f"{name}.pt")
    return out


def main():
    torch.manual_seed(0)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    cfg = R.tiny_config()
    # config 1 (BASELINE.json configs[0]): tiny, prompt 16, 128-frame greedy decode
    golden_dual_ar("dual_ar_tiny_greedy", cfg, T=16, frames=128, voice_span=6, top_k=1,
                   temperature=1.0, top_p=0.9)
    # stochastic path with injected uniforms (server defaults T=0.8 top_p=0.8 top_k=30; RAS live)
    golden_dual_ar("dual_ar_tiny_sampled", cfg, T=24, frames=48, voice_span=0, top_k=30,
                   temperature=0.8, top_p=0.8, useed=99)
    from oracle import gen_golden_sampling, gen_golden_dac  # noqa
    gen_golden_sampling.main()
    gen_golden_dac.main()


if __name__ == "__main__":
    main()
