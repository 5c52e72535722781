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
import os
for kv in os.environ.get("TRACE_OPTS", "").split(","):
    if kv:
        model.set_option(kv.split("=")[0], int(kv.split("=")[1]))
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

# detail of a few consecutive ops in the middle of the slow stack: per-CTA percentiles (us, relative to the first start)
import os
if os.environ.get("TRACE_DETAIL", "1") == "1":
    base = 1 + 5 * 10
    for o in range(base, base + 6):
        st, x1, en, ba = tr[:, o, 0], tr[:, o, 1], tr[:, o, 2], tr[:, o, 3]
        ref = float(tr[:, base, 0].min())
        q = lambda v: "%.2f/%.2f/%.2f" % (float(v.min()) - ref, float(v.median()) - ref, float(v.max()) - ref)
        print(f"op {o} {kinds[o]:6s} start {q(st)}  xstaged {q(x1)}  end {q(en)}")

# per-CTA consistency: is the same CTA slow in every op of a kind?
for kind in ("w13", "w2", "f.w13"):
    idx = [o for o, k in enumerate(kinds) if k == kind]
    loop = (tr[:, idx, 2] - tr[:, idx, 1])          # [cta, ops] loop duration
    per_cta = loop.mean(1)
    rank = loop.argsort(0).argsort(0).double()       # rank of each CTA within each op
    rank_mean = rank.mean(1)
    order = per_cta.argsort()
    print(f"{kind}: loop us per CTA min {float(per_cta.min()):.2f} median {float(per_cta.median()):.2f} max {float(per_cta.max()):.2f};"
          f" rank stability: slowest CTAs {order[-8:].tolist()} mean rank {[round(float(rank_mean[i])) for i in order[-8:]]}"
          f" fastest {order[:8].tolist()} mean rank {[round(float(rank_mean[i])) for i in order[:8]]}")

# attention ops: time from op start to "q ready" (tr[1]) and from there to the end, CTAs with an item only
idx = [o for o, k in enumerate(kinds) if k == "attn"]
a0, a1, a2 = tr[:, idx, 0], tr[:, idx, 1], tr[:, idx, 2]
has = (a2 - a0) > 1.0
has = has & (a1 > 0)
wait_q = (a1 - a0)[has]
rest = (a2 - a1)[has]
print(f"attn items: start->q ready mean {float(wait_q.mean()):.2f} max {float(wait_q.max()):.2f} us; q ready->end mean {float(rest.mean()):.2f} median {float(rest.median()):.2f} max {float(rest.max()):.2f} us")
comb = has & ((a2 - a0) > (a2 - a0)[has].median() + 2.0)
if comb.any():
    print(f"  combiner items only: start->point mean {float((a1 - a0)[comb].mean()):.2f} max {float((a1 - a0)[comb].max()):.2f}; point->end mean {float((a2 - a1)[comb].mean()):.2f} max {float((a2 - a1)[comb].max()):.2f}  (n={int(comb.sum())})")

# per-iteration detail of one op (TRACE_OPTS=mega_detail_op=<index>): group 0 / lane 0 of every CTA
dop = [int(kv.split("=")[1]) for kv in os.environ.get("TRACE_OPTS", "").split(",") if kv.startswith("mega_detail_op=")]
if dop:
    o = dop[0]
    raw = model._buffer("mega_trace", torch.int64)[n_sm * n_ops * 4: n_sm * n_ops * 4 + n_sm * 64].view(n_sm, 64).cpu().double()
    det = (raw - t0) / 1e3
    ref = float(tr[:, o, 1].median())
    print(f"detail of op {o} ({kinds[o]}): times relative to the median x-staged stamp; per iteration of group 0: stage seen / dot done / epilogue done (median over CTAs, [min..max])")
    for itn in range(6):
        cols = det[:, itn * 3: itn * 3 + 3]
        ok = raw[:, itn * 3] > 0
        if not bool(ok.any()):
            break
        f = lambda v: "%.2f [%.2f..%.2f]" % (float(v[ok].median()) - ref, float(v[ok].min()) - ref, float(v[ok].max()) - ref)
        print(f"  it {itn}: seen {f(cols[:, 0])}  dot {f(cols[:, 1])}  epi {f(cols[:, 2])}   (CTAs {int(ok.sum())})")
    print(f"  op end (tid 0): {float(tr[:, o, 2].median()) - ref:.2f}")
    print("  producer: request time of the op's stages (median over CTAs [min..max]), same reference; previous op ended at "
          f"{float(tr[:, o - 1, 2].median()) - ref:.2f}, the op before it at {float(tr[:, o - 2, 2].median()) - ref:.2f}")
    for k in range(16):
        col, ok = det[:, 32 + k], raw[:, 32 + k] > 0
        if bool(ok.any()):
            print(f"    stage {k:2d}: {float(col[ok].median()) - ref:7.2f} [{float(col[ok].min()) - ref:.2f}..{float(col[ok].max()) - ref:.2f}]")
