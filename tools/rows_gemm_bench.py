#!/usr/bin/env python
"""Times fsb_rows_gemm (tcgen05 stream-K many-row Linear) alone on the S2-Pro Linear shapes, cycling over enough distinct
weight matrices that every launch streams from HBM (not L2), next to torch.matmul (cuBLAS) on the same data.

    python tools/rows_gemm_bench.py [--rows 64] [--json out.json]
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=64)
    ap.add_argument("--json", default="")
    args = ap.parse_args()
    import torch
    from bench import measured_peaks
    from fish_speech_b200 import _lib
    _lib.require_cuda()
    lib = _lib.load()
    peak, _ = measured_peaks()
    M = args.rows
    st = torch.cuda.current_stream().cuda_stream
    nb = lib.fsb_rows_gemm_scratch_bytes()
    scratch = torch.zeros(nb, dtype=torch.uint8, device="cuda")
    shapes = [("wqkv", 6144, 2560, 0), ("wo", 2560, 4096, 1), ("w1w3", 9728, 2560, 2), ("w2", 2560, 9728, 1),
              ("fast_head", 4096, 2560, 0), ("lm_head", 155776, 2560, 0)]
    rows = []
    for name, N, K, epi in shapes:
        nw = 2 if epi == 2 else 1
        wbytes = nw * N * K * 2
        copies = max(2, min(24, int(1.5e9 // wbytes)))
        Ws = [torch.empty(nw * N, K, dtype=torch.bfloat16, device="cuda").normal_(0, 0.02) for _ in range(copies)]
        x = torch.randn(M, K, device="cuda").bfloat16()
        res = torch.randn(M, N, device="cuda").bfloat16()
        out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")

        def ours():
            for W in Ws:
                rc = lib.fsb_rows_gemm(W.data_ptr(), W[N:].data_ptr() if nw == 2 else None, x.data_ptr(), res.data_ptr(),
                                       out.data_ptr(), M, N, K, epi, scratch.data_ptr(), nb, st)
                assert rc == 0

        def blas():
            for W in Ws:
                torch.matmul(x, W.t())

        def timeit(fn, it=5):
            fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(it):
                fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) * 1e3 / (it * len(Ws))
        t_o, t_b = timeit(ours), timeit(blas)
        r = {"op": name, "M": M, "N": N * nw, "K": K, "weight_MB": wbytes / 1e6, "ours_us": t_o, "cublas_us": t_b,
             "ours_gbs": wbytes / t_o / 1e3, "cublas_gbs": wbytes / t_b / 1e3, "ours_frac": wbytes / t_o / 1e3 / peak,
             "ideal_us": wbytes / peak / 1e3}
        rows.append(r)
        print(json.dumps(r))
        del Ws
    if args.json:
        Path(args.json).write_text("\n".join(json.dumps(r) for r in rows) + "\n")


if __name__ == "__main__":
    main()
