"""Per-phase cycle breakdown of the persistent frame kernel (CTA 0), S2-Pro shape, batch 1.
usage: python tools/mega_prof.py [frames]"""
import sys

import torch

sys.path.insert(0, ".")
from bench import PROMPT_T, synth_prompt, synth_state_dict  # noqa: E402
from fish_speech_b200 import DualARTransformerB200, s2_pro_config  # noqa: E402

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dev = torch.device("cuda", 0)
cfg = s2_pro_config(max_seq_len=2048)
model = DualARTransformerB200(cfg, synth_state_dict(cfg, dev, seed=0), device=dev, max_rows=1)
model.setup_caches(1, cfg.max_seq_len)
prompt = synth_prompt(cfg, PROMPT_T, seed=1)
model.prefill(prompt[:, :-1].to(dev), 0, 0)
sess = model.session(1, frames * 2 + 2)
bias = torch.full((cfg.vocab_size,), float("-inf"), dtype=torch.bfloat16, device=dev)
bias[cfg.semantic_begin_id: cfg.semantic_end_id + 1] = 0.0
sess.bias, sess.restrict_scan, sess.seed = bias, True, 1
sess.cur_tokens.copy_(prompt[:, -1].to(torch.int32).view(1, -1).to(dev))
sess.pos.fill_(PROMPT_T - 1)
sess.run(frames, 0.8008, 0.8008, 30, True)   # warm
torch.cuda.synchronize()
model.set_option("mega_prof", 1)
prof = model._buffer("mega_prof", torch.int64)
prof.zero_()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
sess.run(frames, 0.8008, 0.8008, 30, True)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / frames
p = prof.cpu().tolist()
names = ["x-staging", "stage loop", "grid barrier", "attention", "fast attn", "sampling", "embed/end", "total"]
n = max(1, p[8])
print(f"frames {n}  event ms/frame {ms:.3f}")
for i, nm in enumerate(names):
    print(f"  {nm:14s} {p[i] / n / 1.965e6:8.3f} ms/frame  ({100.0 * p[i] / max(1, p[7]):5.1f} %)")
