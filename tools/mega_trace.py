"""Per-op timeline of ONE frame of the persistent frame kernel, all CTAs (globaltimer stamps), S2-Pro shape, batch 1.
Prints, per op kind: mean over ops of [slowest CTA's op time, fastest, spread of finish times, barrier latency]."""
import sys
from collections import defaultdict

import torch

sys.path.insert(0, ".")
from bench import PROMPT_T, synth_prompt, synth_state_dict  # noqa: E402
from fish_speech_b200 import DualARTransformerB200, s2_pro_config  # noqa: E402

dev = torch.device("cuda", 0)
cfg = s2_pro_config(max_seq_len=2048)
model = DualARTransformerB200(cfg, synth_state_dict(cfg, dev, seed=0), device=dev, max_rows=1)
model.setup_caches(1, cfg.max_seq_len)
prompt = synth_prompt(cfg, PROMPT_T, seed=1)
model.prefill(prompt[:, :-1].to(dev), 0, 0)
sess = model.session(1, 64)
bias = torch.full((cfg.vocab_size,), float("-inf"), dtype=torch.bfloat16, device=dev)
bias[cfg.semantic_begin_id: cfg.semantic_end_id + 1] = 0.0
sess.bias, sess.restrict_scan, sess.seed = bias, True, 1
sess.cur_tokens.copy_(prompt[:, -1].to(torch.int32).view(1, -1).to(dev))
sess.pos.fill_(PROMPT_T - 1)
sess.run(8, 0.8008, 0.8008, 30, True)
torch.cuda.synchronize()
model.set_option("mega_trace", 1)
sess.run(1, 0.8008, 0.8008, 30, True)
torch.cuda.synchronize()
n_sm = torch.cuda.get_device_properties(0).multi_processor_count
# op list: rebuild the kinds on the host (same order as build_mega_plan)
L, FL, C = cfg.n_layer, cfg.n_fast_layer, cfg.num_codebooks
layer = ["wqkv", "attn", "wo", "w13", "w2"]
flayer = ["f.wqkv", "f.attn", "f.wo", "f.w13", "f.w2"]
kinds = ["embed"] + layer * L + ["head", "sample"] + flayer * FL
for i in range(1, C):
    kinds += flayer * FL + ["f.head", "f.sample"]
kinds += ["end"]
n_ops = len(kinds)
tr = model._buffer("mega_trace", torch.int64)[: n_sm * n_ops * 4].view(n_sm, n_ops, 4).cpu().double()
t0 = tr[:, 0, 0].min()
tr = (tr - t0) / 1e3  # us
print(f"ops {n_ops}  frame {float(tr[:, -1, 2].max()):.1f} us")
agg = defaultdict(list)
for o, k in enumerate(kinds):
    st, xs_, en, ba = tr[:, o, 0], tr[:, o, 1], tr[:, o, 2], tr[:, o, 3]
    dur = en - st
    agg[k].append((float(st.min()), float(dur.max()), float(dur.min()), float(dur.mean()), float(en.max() - en.min()),
                   float(ba.max() - en.max()), float(ba.max() - st.min()), float((xs_ - st).mean()) if k.endswith(("wqkv", "wo", "w13", "w2", "head")) else 0.0))
print(f"{'op':9s} {'n':>4s} {'op max':>8s} {'op min':>8s} {'op mean':>8s} {'end spread':>10s} {'barrier':>8s} {'wall':>8s} {'xstage':>7s}   [us, mean over ops]   total wall ms")
tot = 0.0
for k, v in agg.items():
    t = torch.tensor(v)
    m = t.mean(0)
    tot += float(t[:, 6].sum())
    print(f"{k:9s} {len(v):4d} {m[1]:8.2f} {m[2]:8.2f} {m[3]:8.2f} {m[4]:10.2f} {m[5]:8.2f} {m[6]:8.2f} {m[7]:7.2f}   {float(t[:, 6].sum()) / 1e3:8.3f}")
print(f"sum of walls {tot / 1e3:.3f} ms")
