#!/usr/bin/env python
"""BASELINE configs[2]: S2-Pro-shape Dual-AR batched decode — B concurrent requests, each with a CTX-token KV context
already resident (synthetic prefill: the caches are filled with N(0, 0.5) bf16), decoded together for F frames.

    python tools/batch_bench.py [--batch 64] [--ctx 2048] [--frames 256] [--steps 2] [--mode 1] [--json out.json]

Prints one JSON object: audio-s/s over all requests, ms/step, algorithmic bytes per step (fsb_dual_ar_frame_bytes) and
the fraction of the measured HBM bandwidth.  Under torchrun every rank decodes its own B requests (weak scaling,
no collective on the data path); rank 0 reports the sum over ranks over the slowest rank's time.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
FRAME_SEC = 2048.0 / 44100.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--ctx", type=int, default=2048)
    ap.add_argument("--frames", type=int, default=256)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--mode", type=int, default=1, help="batch_gemm option: 1 = every Linear once over all rows, 0 = 8-row GEMV passes")
    ap.add_argument("--json", default="")
    ap.add_argument("--max-rows", type=int, default=0, help="row capacity of the handle (default: the batch)")
    ap.add_argument("--opt", action="append", default=[], help="engine option name=value (e.g. attn_splits=8), repeatable")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from bench import ClockSampler, measured_peaks, synth_state_dict
    from fish_speech_b200 import DualARTransformerB200, s2_pro_config
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, F = args.batch, args.frames
    S = (args.ctx + F + 8 + 7) // 8 * 8
    cfg = s2_pro_config(max_seq_len=S)
    sd = synth_state_dict(cfg, dev, seed=0)
    model = DualARTransformerB200(cfg, sd, device=dev, max_rows=max(B, args.max_rows))
    del sd
    model.setup_caches(max(B, args.max_rows), S)
    model.set_option("batch_gemm", args.mode)
    for kv in args.opt:
        k, v = kv.split("=")
        model.set_option(k, int(v))
    g = torch.Generator(device=dev).manual_seed(3 + rank)
    for k, v in model.kv_cache:
        k[:, :, : args.ctx].normal_(0.0, 0.5, generator=g)
        v[:, :, : args.ctx].normal_(0.0, 0.5, generator=g)
    sess = model.session(B, F + 1)
    bias = torch.full((cfg.vocab_size,), float("-inf"), dtype=torch.bfloat16, device=dev)
    bias[cfg.semantic_begin_id: cfg.semantic_end_id + 1] = 0.0  # <|im_end|> masked: fixed frame count
    sess.bias, sess.restrict_scan, sess.seed = bias, True, 99 + rank
    C1 = cfg.num_codebooks + 1
    start = torch.zeros((B, C1), dtype=torch.int32, device=dev)
    start[:, 0] = cfg.semantic_begin_id + torch.randint(0, cfg.codebook_size, (B,), device=dev, generator=g).int()
    start[:, 1:] = torch.randint(0, cfg.codebook_size, (B, C1 - 1), device=dev, generator=g).int()

    def reset():
        sess.cur_tokens.copy_(start)
        sess.pos.fill_(args.ctx)
        sess.frame_counter.zero_()
        sess.ras_window.zero_()
        sess.done.zero_()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(1, args.warmup)):
        reset()
        sess.run(min(F, 8), 0.8008, 0.8008, 30, True)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        reset()
        sess.run(F, 0.8008, 0.8008, 30, True)
    e1.record()
    barrier()
    clocks = sampler.stop()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms[0])
    ok = bool((sess.out_tokens[:F, :, 1:] >= 0).all()) and bool((sess.out_tokens[:F, :, 1:] < cfg.codebook_size).all())
    if rank == 0:
        peak, src = measured_peaks()
        ms_step = ms / (args.steps * F)
        ctx_avg = args.ctx + (F - 1) / 2.0
        fbytes = model.frame_bytes(B, int(ctx_avg * B), masked_head=True)  # counts the LM-head rows the step actually streams
        out = {
            "workload": f"S2-Pro-shape Dual-AR batched decode, {B} requests/GPU x {world} GPU(s), {args.ctx}-token KV context, {F} frames",
            "batch_gemm": args.mode, "n_gpus": world, "audio_s_per_s": world * B * args.steps * F * FRAME_SEC / (ms / 1e3),
            "ms_per_step": ms_step, "frames_per_s": world * B / (ms_step / 1e3), "bytes_per_step": fbytes,
            "achieved_gbs": fbytes / (ms_step * 1e-3) / 1e9, "peak_gbs": peak, "frac": fbytes / (ms_step * 1e-3) / 1e9 / peak,
            "peak_source": src, "codes_in_range": ok, "launches_per_step": model.launches_per_frame(B), "clocks": clocks,
        }
        print(json.dumps(out))
        if args.json:
            Path(args.json).write_text(json.dumps(out) + "\n")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
