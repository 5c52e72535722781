"""Golden vectors for the sampling contract, produced by the
REFERENCE's own `sample()` (inference.py:80-93) with `torch.rand_like` patched to
return recorded bf16 uniforms.  TEST INFRASTRUCTURE.
"""
from __future__ import annotations

import torch

from oracle import dual_ar_ref as R
from oracle.gen_golden import GOLD, import_reference


def bf16_uniform(g, shape):
    """Same support as torch.rand_like on bf16: multiples of 1/256 in [0, 1)."""
    return (torch.randint(0, 256, shape, generator=g).float() / 256.0).to(torch.bfloat16)


def make_cases(seed: int = 7, n_random: int = 160):
    g = torch.Generator().manual_seed(seed)
    cases = []

    def add(logits, T, P, K, u=None, tag=""):
        logits = logits.to(torch.bfloat16)
        if u is None:
            u = bf16_uniform(g, logits.shape)
        cases.append({"logits": logits, "temperature": T, "top_p": P, "top_k": K, "u": u, "tag": tag})

    for i in range(n_random):
        n = [1024, 4096, 8192][i % 3]
        sigma = [0.5, 1.0, 3.0, 8.0][(i // 3) % 4]
        T, P = [(0.8, 0.8), (1.0, 0.9), (0.7, 0.7), (1.5, 0.95)][(i // 12) % 4]
        K = [30, 30, 1, 5][(i // 48) % 4]
        add(torch.randn(n, generator=g) * sigma, T, P, K, tag="random")
    # constrained-decode bias: -inf outside a 1024+1 window
    for i in range(16):
        l = torch.randn(8192, generator=g) * 2
        b = torch.full((8192,), float("-inf"))
        b[4096:5120] = 0
        b[7] = 0
        add((l.to(torch.bfloat16) + b.to(torch.bfloat16)), 0.8, 0.8, 30, tag="biased")
    # u = 0 on the winner (score -> 0), all-zero uniforms, heavy ties, one dominant logit
    l = torch.randn(4096, generator=g)
    add(l, 0.8, 0.8, 30, u=torch.zeros(4096, dtype=torch.bfloat16), tag="u_all_zero")
    l = torch.randn(4096, generator=g)
    l[1234] = 30.0
    add(l, 1.0, 0.9, 30, tag="dominant")
    u = bf16_uniform(g, (4096,))
    u[1234] = 0
    add(l, 1.0, 0.9, 30, u=u, tag="dominant_u0")
    add(torch.zeros(1024), 1.0, 0.9, 30, tag="all_equal")
    add(torch.round(torch.randn(4096, generator=g) * 2) / 2, 0.8, 0.8, 30, tag="quantised_ties")
    add(torch.randn(4096, generator=g) * 1e-3, 1e-6, 0.8, 30, tag="tiny_temperature")  # clip(T, 1e-5)
    add(torch.randn(4096, generator=g), 0.8, 1.0, 30, tag="top_p_one")
    add(torch.randn(4096, generator=g), 0.8, 0.0, 30, tag="top_p_zero")
    return cases


def main():
    _, inf = import_reference()
    cases = make_cases()
    real = torch.rand_like
    out = []
    for c in cases:
        T = torch.tensor(c["temperature"], dtype=torch.bfloat16)
        P = torch.tensor(c["top_p"], dtype=torch.bfloat16)
        torch.rand_like = lambda t, **k: c["u"]
        try:
            idx, probs = inf.sample(c["logits"][None, None], temperature=T, top_p=P, top_k=c["top_k"])
        finally:
            torch.rand_like = real
        idx2, probs2 = R.sample(c["logits"][None, None], T, P, c["top_k"], rand_like=lambda t: c["u"])
        assert torch.equal(idx, idx2) and torch.equal(probs, probs2), c["tag"]
        nz = probs.nonzero().flatten()
        out.append({**c, "idx": int(idx), "kept": nz.to(torch.int32), "kept_probs": probs[nz].clone()})
    torch.save(out, GOLD / "sampling_cases.pt")
    print(f"[golden] sampling: {len(out)} cases, oracle == reference bit-exact")


if __name__ == "__main__":
    main()
