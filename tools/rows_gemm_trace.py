#!/usr/bin/env python
"""Per-CTA phase timing of one fsb_rows_gemm launch (clock64 stamps, FSB_RG_TRACE=1): where the fixed cost of a short
weight-streaming launch goes.  python tools/rows_gemm_trace.py"""
import os
import sys
from pathlib import Path
os.environ["FSB_RG_TRACE"] = "1"
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from fish_speech_b200 import _lib
_lib.require_cuda()
lib = _lib.load()
nb = lib.fsb_rows_gemm_scratch_bytes()
names = ["entry", "setup done", "first stage landed", "last MMA issued", "first seg accum ready", "last seg accum ready",
         "epilogue done"]
for name, N, K, epi in [("wqkv", 6144, 2560, 0), ("wo", 2560, 4096, 1), ("w1w3", 9728, 2560, 2), ("w2", 2560, 9728, 1), ("lm_head", 155776, 2560, 0)]:
    M = 64
    nw = 2 if epi == 2 else 1
    Ws = [torch.empty(nw * N, K, dtype=torch.bfloat16, device="cuda").normal_(0, 0.02) for _ in range(3)]
    x = torch.randn(M, K, device="cuda").bfloat16()
    res = torch.randn(M, N, device="cuda").bfloat16()
    out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    scratch = torch.zeros(nb, dtype=torch.uint8, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for W in Ws:
        scratch.zero_()
        flush.fill_(1)
        torch.cuda.synchronize()
        rc = lib.fsb_rows_gemm(W.data_ptr(), W[N:].data_ptr() if nw == 2 else None, x.data_ptr(), res.data_ptr(), out.data_ptr(),
                               M, N, K, epi, scratch.data_ptr(), nb, torch.cuda.current_stream().cuda_stream)
        assert rc == 0
        torch.cuda.synchronize()
    off = 320 * 2 * 128 * 128 * 4  # rg_partial_bytes(2 x 160 CTAs, 128 rows)
    tr = scratch[off: off + 320 * 8 * 8].view(torch.int64).view(320, 8).cpu().double()
    tr = tr[tr[:, 0] > 0]  # CTAs of this launch
    rel = (tr - tr[:, :1]) / 1.965e3  # us at 1965 MHz
    print(f"== {name}: N={N*nw} K={K} ({nw*N*K*2/1e6:.1f} MB, ideal {nw*N*K*2/6554.6e3:.1f} us), {tr.shape[0]} CTAs")
    for i, nm in enumerate(names):
        col = rel[:, i]
        print(f"   {nm:24s} median {col.median():7.2f} us   min {col.min():7.2f}   max {col.max():7.2f}")
