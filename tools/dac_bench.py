"""DAC codec decode throughput (BASELINE configs[3]: B=32, T=646 code frames = 30 s clips), synthetic weights.
Prints audio-s/s, achieved TFLOP/s (algorithmic GEMM-shaped FLOPs, fsb_dac_flops) and the fraction of the measured
bf16 tensor peak.  usage: python tools/dac_bench.py [B] [T] [iters] [operand f16|bf16] [max_batch] [name=value options ...]"""
import json
import sys

import torch

sys.path.insert(0, ".")
from fish_speech_b200.dac import DACB200, DACConfig, synthetic_state_dict  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
T = int(sys.argv[2]) if len(sys.argv) > 2 else 646
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 3
operand = sys.argv[4] if len(sys.argv) > 4 else "f16"
max_batch = int(sys.argv[5]) if len(sys.argv) > 5 else 4
opts = dict(kv.split("=") for kv in sys.argv[6:])
cfg = DACConfig()
dev = "cuda:0"
m = DACB200(cfg, synthetic_state_dict(cfg, 0), device=dev, max_batch=max_batch, max_frames=max(T, 8), operand=operand)
for k, v in opts.items():
    m.set_option(k, int(v))
g = torch.Generator().manual_seed(2)
codes = torch.randint(0, cfg.codebook_size, (B, cfg.n_codebooks + 1, T), generator=g)
codes[:, 0] = torch.randint(0, cfg.semantic_codebook_size, (B, T), generator=g)
codes = codes.to(dev)
for _ in range(2):
    wav = m.from_indices(codes.clone())
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(iters):
    wav = m.from_indices(codes.clone())
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
audio_s = B * T * cfg.hop / cfg.sample_rate
flops = m.flops(B, T)
peaks = json.load(open("MEASURED_PEAKS.json")) if __import__("os").path.exists("MEASURED_PEAKS.json") else {}
peak = float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1375.8)))
print(json.dumps({"workload": f"DAC decode B={B} T={T}", "ms": ms, "audio_s_per_s": audio_s / (ms / 1e3),
                  "tflops": flops / (ms * 1e-3) / 1e12, "frac_of_bf16_peak": flops / (ms * 1e-3) / 1e12 / peak,
                  "finite": bool(torch.isfinite(wav).all()), "operand": operand, "max_batch": max_batch, "options": opts}))
