"""Time to first audio at the S2-Pro shape: one request (a 384-column reference voice + 64 own columns) through the scheduler
with request["stream"], its code portions through DACStream; prints the time from submit to the first waveform samples on the
host, and the delivered rate after that.  usage: python tools/first_audio_bench.py [rows] [chunk_frames] [frames] [prefill_budget]
(rows > 1: the other rows run requests of their own at the same time)"""
import json
import queue
import sys
import time

import torch

sys.path.insert(0, ".")
from bench import FRAME_SEC, synth_prompt, synth_state_dict  # noqa: E402
from fish_speech_b200 import DACB200, DACConfig, DualARTransformerB200, s2_pro_config  # noqa: E402
from fish_speech_b200.dac import synthetic_state_dict  # noqa: E402
from fish_speech_b200.scheduler import ContinuousBatcher, GenerateRequest  # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1
chunk = int(sys.argv[2]) if len(sys.argv) > 2 else 8
frames = int(sys.argv[3]) if len(sys.argv) > 3 else 128
budget = int(sys.argv[4]) if len(sys.argv) > 4 and int(sys.argv[4]) > 0 else None      # prompt columns per chunk boundary
dev = torch.device("cuda", 0)
cfg = s2_pro_config(max_seq_len=1024)
model = DualARTransformerB200(cfg, synth_state_dict(cfg, dev, seed=0), device=dev, max_rows=rows)
model.setup_caches(rows, cfg.max_seq_len)
model.ignore_eos = True
dc = DACConfig()
dac = DACB200(dc, synthetic_state_dict(dc, seed=0), device=dev, max_batch=1, max_frames=1536)
assert dc.n_codebooks + 1 == cfg.num_codebooks, (dc.n_codebooks, cfg.num_codebooks)


def make(i, stream=False):
    p = synth_prompt(cfg, 384 + 64, seed=10 + i)
    p[:, :384] = synth_prompt(cfg, 384, seed=3)
    return GenerateRequest(request=dict(prompt=p.to(dev), max_new_tokens=frames, temperature=0.8, top_p=0.8, top_k=30,
                                        stream=stream), response_queue=queue.Queue())


def run(reqs):
    b = ContinuousBatcher(model, chunk_frames=chunk, prefill_budget=budget)
    stream = dac.stream(batch=1)
    marks = []                                     # (host time, frames of audio on the host so far)
    got = [0]

    def on_partial(item, codes):
        if item is not reqs[0]:
            return
        idx = codes.to(torch.int64)[None]           # [1, C, n], semantic row first (inference.py:708); the engine clamps
        wav = stream.push(idx).cpu()                # the samples are on the host when push + copy return
        got[0] += idx.shape[2]
        marks.append((time.perf_counter(), got[0], wav.shape[-1]))

    b.on_partial = on_partial
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for it in reqs:
        b.submit(it)
    b.run_until_idle()
    torch.cuda.synchronize()
    return t0, marks, time.perf_counter()


run([make(i, i == 0) for i in range(rows)])                # warm-up: graph capture, tensor maps, rope tables
t0, marks, t1 = run([make(100 + i, i == 0) for i in range(rows)])
first = marks[0]
print(json.dumps({"rows": rows, "chunk_frames": chunk, "prefill_budget": budget, "prompt_columns": 448, "frames": frames,
                  "first_audio_ms": round((first[0] - t0) * 1e3, 2), "first_audio_frames": first[1],
                  "first_audio_seconds_of_audio": round(first[1] * FRAME_SEC, 3),
                  "portions": len(marks), "wall_ms": round((t1 - t0) * 1e3, 1),
                  "audio_s_per_s_streamed": round(marks[-1][1] * FRAME_SEC / (marks[-1][0] - t0), 2)}))
