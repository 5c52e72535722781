#!/usr/bin/env python
"""Phase timing of attn_tc_kernel CTAs (clock64 stamps, engine option attn_trace) in one many-row decode step:
entry -> q prep done -> key loop done -> partial written / arrival counted.  python tools/attn_trace.py [batch] [ctx] [splits]"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from bench import synth_state_dict
from fish_speech_b200 import DualARTransformerB200, s2_pro_config
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
ctx = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
splits = int(sys.argv[3]) if len(sys.argv) > 3 else 0
dev = torch.device("cuda", 0)
cfg = s2_pro_config(max_seq_len=ctx + 64)
m = DualARTransformerB200(cfg, synth_state_dict(cfg, dev, 0), device=dev, max_rows=B)
m.setup_caches(B, ctx + 64)
m.set_option("attn_trace", 1)
m.set_option("graph", 0)
if splits:
    m.set_option("attn_splits", splits)
tok = torch.zeros((B, cfg.num_codebooks + 1), dtype=torch.int32, device=dev)
pos = torch.full((B,), ctx, dtype=torch.int32, device=dev)
st = torch.cuda.current_stream().cuda_stream
for _ in range(2):
    m._ck(m.lib.fsb_dual_ar_slow_step(m._h, B, tok.data_ptr(), pos.data_ptr(), None, st))
torch.cuda.synchronize()
tr = m._buffer("mega_trace", torch.int64).view(-1, 4).cpu().double()
tr = tr[(tr[:, 0] > 0) & (tr[:, 3] > 0)]
rel = (tr - tr[:, :1]) / 1.965e3
print(f"{tr.shape[0]} CTAs of the last layer's launch (B={B}, ctx={ctx})")
for i, nm in enumerate(["entry", "q prep done", "key loop done", "partial written + counted"]):
    c = rel[:, i]
    print(f"  {nm:28s} median {c.median():7.2f} us  p10 {c.quantile(0.1):7.2f}  p90 {c.quantile(0.9):7.2f}  max {c.max():7.2f}")
span = (tr[:, 3].max() - tr[:, 0].min()) / 1.965e3
print(f"  first entry -> last arrival (SM clocks are not synchronised: indicative) {span:.1f} us")
