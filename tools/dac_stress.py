"""Determinism stress of the DAC decode: the same clips decoded N times must give the same bytes every time (a race in a
kernel shows up as a rare difference).  usage: python tools/dac_stress.py [iters] [B] [T]"""
import sys

import torch

sys.path.insert(0, ".")
from fish_speech_b200.dac import DACB200, DACConfig, synthetic_state_dict  # noqa: E402

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 30
B = int(sys.argv[2]) if len(sys.argv) > 2 else 8
T = int(sys.argv[3]) if len(sys.argv) > 3 else 646
cfg = DACConfig()
m = DACB200(cfg, synthetic_state_dict(cfg, 0), device="cuda:0", max_batch=B, max_frames=T)
g = torch.Generator().manual_seed(2)
codes = torch.randint(0, cfg.codebook_size, (B, cfg.n_codebooks + 1, T), generator=g).cuda()
first = m.from_indices(codes.clone()).clone()
bad = 0
for i in range(iters):
    w = m.from_indices(codes.clone())
    if not torch.equal(w, first):
        d = (w - first).abs()
        nz = d.nonzero()
        bad += 1
        print(f"iter {i}: {int((d > 0).sum())} samples differ, max {float(d.max()):.3e}, first at {nz[0].tolist()}, last at {nz[-1].tolist()}")
print(f"{iters} decodes of {B} x {T} frames: {bad} differed from the first")
