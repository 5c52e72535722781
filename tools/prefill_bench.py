"""Prompt pass of the S2-Pro-shape model: ms for T prompt tokens per mode (1 = own tcgen05 GEMM, 2 = cuBLAS GEMMs).
usage: python tools/prefill_bench.py [T ...]"""
import sys

import torch

sys.path.insert(0, ".")
from bench import synth_prompt, synth_state_dict  # noqa: E402
from fish_speech_b200 import DualARTransformerB200, s2_pro_config  # noqa: E402

Ts = [int(x) for x in sys.argv[1:]] or [511, 2047]
S = max(Ts) + 9
dev = torch.device("cuda", 0)
cfg = s2_pro_config(max_seq_len=(S + 7) // 8 * 8)
model = DualARTransformerB200(cfg, synth_state_dict(cfg, dev, seed=0), device=dev, max_rows=1)
model.setup_caches(1, cfg.max_seq_len)
c = cfg
p_layer = c.dim * (c.n_head + 2 * c.n_local_heads) * c.head_dim + c.n_head * c.head_dim * c.dim + 3 * c.dim * c.intermediate_size
for T in Ts:
    prompt = synth_prompt(cfg, T + 1, seed=1)[:, :T].to(dev)
    for mode in (1, 2):
        model.set_option("prefill_gemm", mode)
        model.prefill(prompt, 0, 0)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            model.prefill(prompt, 0, 0)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        flops = 2.0 * c.n_layer * p_layer * T + 4.0 * c.n_layer * c.n_head * c.head_dim * T * T / 2
        print(f"T={T} prefill_gemm={mode}: {ms:.2f} ms  {flops / ms / 1e9:.0f} TFLOP/s (GEMM + causal attention FLOPs)", flush=True)
