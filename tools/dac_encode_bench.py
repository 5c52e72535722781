"""Encode side of the codec at the full config (synthetic weights): time of DACB200.encode for clips of `seconds` seconds.
usage: python tools/dac_encode_bench.py [seconds] [batch] [repeats]"""
import json
import sys

import torch

sys.path.insert(0, ".")
from fish_speech_b200 import DACB200, DACConfig  # noqa: E402
from fish_speech_b200.dac import synthetic_state_dict  # noqa: E402

seconds = float(sys.argv[1]) if len(sys.argv) > 1 else 10.0
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 1
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
c = DACConfig()
n = int(seconds * c.sample_rate)
frames = -(-n // c.hop)
m = DACB200(c, synthetic_state_dict(c, seed=0, encode_side=True), max_batch=batch, max_frames=frames)
g = torch.Generator().manual_seed(0)
audio = (0.3 * torch.randn((batch, 1, n), generator=g)).cuda()
codes, _ = m.encode(audio)                       # warm-up
torch.cuda.synchronize()
ts = []
for _ in range(reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    again, _ = m.encode(audio)
    e1.record()
    e1.synchronize()
    ts.append(e0.elapsed_time(e1))
assert torch.equal(again, codes)
ms = sorted(ts)[len(ts) // 2]
m.set_option("enc_exact", 0)                     # GEMM-shaped layers on the TF32 mma.sync kernels: how many codes move, how fast
m.set_option("tf32", 1)
tf, _ = m.encode(audio)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
tf, _ = m.encode(audio)
e1.record()
e1.synchronize()
tf32 = {"ms": round(e0.elapsed_time(e1), 2), "codes_equal": round(float((tf == codes).float().mean()), 4),
        "semantic_codes_equal": round(float((tf[:, 0] == codes[:, 0]).float().mean()), 4),
        "frames_fully_equal": round(float((tf == codes).all(dim=1).float().mean()), 4)}
m.set_option("enc_exact", 1)
print(json.dumps({"tf32_variant": tf32}))
print(json.dumps({"seconds_per_clip": seconds, "clips": batch, "frames": frames, "ms": round(ms, 2),
                  "audio_s_per_s": round(batch * seconds / (ms * 1e-3), 1),
                  "codes_used": [int(torch.unique(codes[:, i]).numel()) for i in range(codes.shape[1])]}))
