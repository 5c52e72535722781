"""Time the weight-streaming GEMV entry point (fsb_gemv) on the S2-Pro decode shapes with rotating weights
(so nothing is served from L2) across kernel configurations.  Usage: python tools/gemv_sweep.py"""
import itertools
import os
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from fish_speech_b200 import _lib  # noqa: E402

lib = _lib.load()
dev = "cuda"
SHAPES = [("wqkv", 6144, 2560, 1, 0), ("wo", 2560, 4096, 0, 1), ("w13", 9728, 2560, 1, 2), ("w2", 2560, 9728, 0, 1),
          ("head", 155776, 2560, 1, 0)]
PEAK = 6549.8


def run(name, N, K, pro, epi, reps=3):
    nW = max(2, int(600e6 // (N * K * 2)))  # >= 600 MB of distinct weights per sweep
    if name == "head":
        nW = 2
    Ws = [torch.randn(N, K, device=dev).mul_(0.02).bfloat16() for _ in range(nW)]
    W2s = [torch.randn(N, K, device=dev).mul_(0.02).bfloat16() for _ in range(nW)] if epi == 2 else [None] * nW
    x = torch.randn(1, K, device=dev).bfloat16()
    nw = torch.ones(K, device=dev).bfloat16()
    res = torch.randn(1, N, device=dev).bfloat16()
    out = torch.empty(1, N, device=dev, dtype=torch.bfloat16)
    st = torch.cuda.current_stream().cuda_stream
    p = lambda t: None if t is None else t.data_ptr()

    def sweep():
        for W, W2 in zip(Ws, W2s):
            assert lib.fsb_gemv(p(W), p(W2), p(x), p(nw), p(res), p(out), 1, N, K, pro, epi, 1e-6, st) == 0
    sweep()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        sweep()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (reps * nW)
    gb = N * K * 2 * (2 if epi == 2 else 1) / 1e9
    return us, gb / (us * 1e-6)


configs = [dict(FSB_GEMV_TMA="0")]
for nst, cps in [(4, 1), (8, 1), (10, 1), (4, 2), (5, 2)]:
    configs.append(dict(FSB_GEMV_TMA="1", FSB_GEMV_NST=str(nst), FSB_GEMV_CPS=str(cps), FSB_GEMV_STAGE_KB="20"))
if len(sys.argv) > 1:
    SHAPES = [s for s in SHAPES if s[0] in sys.argv[1].split(",")]
if os.environ.get("SWEEP_ONE"):
    configs = [dict()]
for cfg in configs:
    os.environ.update(cfg)
    row = []
    for s in SHAPES:
        us, gbs = run(*s)
        row.append(f"{s[0]} {us:7.1f}us {gbs:6.0f}GB/s {gbs / PEAK:4.2f}")
    print(cfg, " | ".join(row), flush=True)
