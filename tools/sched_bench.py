"""Continuous batching at the S2-Pro shape: N requests (prompts of 64-512 tokens, 48-192 frames each, <|im_end|> masked so the
lengths are the budgets) through ContinuousBatcher with `rows` rows; delivered audio-s/s against the static many-row step.
usage: python tools/sched_bench.py [rows] [requests] [chunk] [voice] [share_min] [prefill_budget]     voice > 0: every prompt starts with the same `voice`
columns (a reference voice), followed by its own 32-128 columns: the prefix is computed once and copied afterwards"""
import json
import queue
import sys
import time

import torch

sys.path.insert(0, ".")
from bench import FRAME_SEC, synth_prompt, synth_state_dict  # noqa: E402
from fish_speech_b200 import DualARTransformerB200, s2_pro_config  # noqa: E402
from fish_speech_b200.scheduler import ContinuousBatcher, GenerateRequest  # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 64
n_req = int(sys.argv[2]) if len(sys.argv) > 2 else 192
chunk = int(sys.argv[3]) if len(sys.argv) > 3 else 8
voice = int(sys.argv[4]) if len(sys.argv) > 4 else 0
dev = torch.device("cuda", 0)
cfg = s2_pro_config(max_seq_len=1024 if voice <= 384 else (voice + 128 + 192 + 8 + 7) // 8 * 8)
model = DualARTransformerB200(cfg, synth_state_dict(cfg, dev, seed=0), device=dev, max_rows=rows)
model.setup_caches(rows, cfg.max_seq_len)
model.ignore_eos = True
g = torch.Generator().manual_seed(0)
reqs = []
for i in range(n_req):
    T = int(torch.randint(64, 513, (1,), generator=g)) if not voice else voice + int(torch.randint(32, 129, (1,), generator=g))
    F = int(torch.randint(48, 193, (1,), generator=g))
    p = synth_prompt(cfg, T, seed=10 + i)
    if voice:
        p[:, :voice] = synth_prompt(cfg, voice, seed=3)
    reqs.append(GenerateRequest(request=dict(prompt=p.to(dev), max_new_tokens=F, temperature=0.8,
                                             top_p=0.8, top_k=30), response_queue=queue.Queue()))
share = int(sys.argv[5]) if len(sys.argv) > 5 else 32
budget = int(sys.argv[6]) if len(sys.argv) > 6 and int(sys.argv[6]) > 0 else None     # prompt columns per chunk boundary
b = ContinuousBatcher(model, chunk_frames=chunk, share_min=share, prefill_budget=budget)
for it in reqs[:rows]:       # warm-up: graph capture, tensor maps
    b.submit(it)
b.step()
b.active.clear(); b.waiting.clear(); b.filling.clear(); b.resident.clear(); b.park.clear(); b.key = None
b.steps = b.frames_run = b.frames_kept = b.prompt_cols = b.shared_cols = 0
torch.cuda.synchronize()
t0 = time.perf_counter()
for it in reqs:
    b.submit(it)
done = b.run_until_idle()
torch.cuda.synchronize()
dt = time.perf_counter() - t0
frames = sum(int(c.shape[1]) + 1 for _, c in done)
print(json.dumps({"rows": rows, "requests": n_req, "chunk_frames": chunk, "prefill_budget": budget, "wall_s": round(dt, 3), "frames_delivered": frames,
                  "audio_s_per_s": round(frames * FRAME_SEC / dt, 1), "steps": b.steps,
                  "row_occupancy": round(b.frames_kept / max(1, b.frames_run), 3),
                  "ms_per_chunk_step": round(dt / max(1, b.steps) * 1e3, 2),
                  "prompt_tokens": sum(int(r.request["prompt"].shape[1]) for r in reqs), "voice_prefix": voice,
                  "prompt_columns_computed": b.prompt_cols - b.shared_cols, "prompt_columns_shared": b.shared_cols}))
