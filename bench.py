#!/usr/bin/env python
"""bench.py — audio-seconds per second of the S2-Pro-shape Dual-AR decode on B200 (BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--frames F]

A "step" = one F-frame (default 1024) batch-1 decode job after a 512-token prompt, S2-Pro shape, random-init bf16,
synthetic prompt.  N > 1 (torchrun): one independent utterance per GPU, no collective on the data path (weak scaling).
Prints ONE JSON line (contract in the task statement; extra keys: roofline, roofline_frame, cpu_baseline, p50_ms_per_frame).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

FRAME_SEC = 2048.0 / 44100.0  # one Dual-AR frame = 2048 samples @ 44.1 kHz
PROMPT_T = 512


def ncu_traffic(kernel: str):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed `ncu --set full` summary
    (profiles/r2_traffic.json, written by tools/ncu_summary.py traffic); None when there is no capture of it."""
    p = ROOT / "profiles" / "r2_traffic.json"
    if not p.exists():
        return None, None
    d = json.loads(p.read_text()).get(kernel)
    return (d["dram_bytes_per_launch"], d.get("source")) if d else (None, None)


METRIC = "audio_sec_per_sec_dual_ar_decode_bs1"


def measured_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu: int):
        super().__init__(daemon=True)
        self.gpu, self.rows, self._stop_evt = gpu, [], threading.Event()

    def run(self):
        while not self._stop_evt.is_set():
            try:
                o = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.gpu)],
                                   capture_output=True, text=True, timeout=5).stdout.strip()
                if o:
                    self.rows.append([x.strip() for x in o.split(",")])
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=6)
        sm = [float(r[1]) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 8 for n, v in zip(names, r[4:8]) if v.lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def synth_state_dict(cfg, device, seed=0):
    """Random-init S2-Pro-shape weights straight on the device: N(0, 0.02) for Linear/Embedding (llama.py:468-477),
    norm weights 1."""
    import torch
    from fish_speech_b200.dual_ar import state_dict_shapes
    g = torch.Generator(device=device).manual_seed(seed)
    sd = {}
    for name, shape in state_dict_shapes(cfg).items():
        if len(shape) == 1:
            sd[name] = torch.ones(shape, dtype=torch.bfloat16, device=device)
        else:
            w = torch.empty(shape, dtype=torch.bfloat16, device=device)
            w.normal_(0.0, cfg.initializer_range, generator=g)
            sd[name] = w
    return sd


def synth_prompt(cfg, T, seed=1):
    import torch
    g = torch.Generator().manual_seed(seed)
    p = torch.zeros((cfg.num_codebooks + 1, T), dtype=torch.long)
    p[0] = torch.randint(0, cfg.semantic_begin_id, (T,), generator=g)
    return p


# ----------------------------------------------------------------------------------------------- CPU reference leg
def cpu_reference_leg(frames: int, steps: int = 1, warmup: int = 0):
    """The reference's CPU path for this workload: oracle/dual_ar_ref.py (a torch-CPU restatement pinned bit-exact
    to the reference's own generate(); the Python reference itself cannot travel to the GPU box), S2-Pro shape,
    bf16, all host threads, the same 512-token prompt pass as the GPU arm (timed separately) and `frames` decode frames
    after it with max_seq_len 2048.  Weights are a tiled random block (values do not affect timing; building 4.5 B fresh normals would
    dominate the run).  Returns (audio-sec/sec, cores, sample description, seconds per frame)."""
    import torch
    from oracle import dual_ar_ref as R
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = R.s2_config(max_seq_len=2048)
    g = torch.Generator().manual_seed(0)
    block = (torch.randn(1 << 20, generator=g) * cfg.initializer_range).to(torch.bfloat16)
    sd = {}
    for name, shape in R.state_dict_keys(cfg).items():
        n = 1
        for s in shape:
            n *= s
        if len(shape) == 1:
            sd[name] = torch.ones(shape, dtype=torch.bfloat16)
        else:
            sd[name] = block.repeat((n + block.numel() - 1) // block.numel())[:n].view(shape)
    m = R.DualARRef(cfg, sd)
    m.setup_caches(1, cfg.max_seq_len)
    prompt = R.make_prompt(cfg, PROMPT_T, seed=1)
    T = torch.tensor(0.8, dtype=torch.bfloat16)
    P = torch.tensor(0.8, dtype=torch.bfloat16)
    bias = R.make_semantic_logit_bias(cfg)
    bias[0, 0, cfg.im_end_id] = float("-inf")  # fixed frame count
    with torch.inference_mode():
        t0 = time.perf_counter()
        first = R.decode_one_token_ar(m, prompt[None], torch.arange(PROMPT_T), T, P, 30, bias)
        prefill_s = time.perf_counter() - t0
        times = []
        pos = PROMPT_T
        cur = first.view(1, cfg.num_codebooks + 1, -1)
        for it in range(warmup + steps):
            t0 = time.perf_counter()
            out = R.decode_n_tokens(m, cur, torch.tensor([pos], dtype=torch.int), frames, T, P, 30, bias)
            dt = time.perf_counter() - t0
            pos += out.shape[1]
            cur = out[:, -1:].reshape(1, cfg.num_codebooks + 1, 1)
            if it >= warmup:
                times.append(dt / max(1, out.shape[1]))
    spf = statistics.mean(times)
    sample = (f"{frames} decode frames/step x {steps} step(s) after the {PROMPT_T}-token prompt pass ({prefill_s:.1f} s, not counted), "
              f"S2-Pro shape bf16, bs1, max_seq_len 2048, oracle port of generate()")
    return FRAME_SEC / spf, cores, sample, spf


def cpu_reference_leg_ref(frames: int, steps: int = 1, warmup: int = 0):
    """The same workload through THE REFERENCE'S OWN code: `generate()` of fish_speech/models/text2semantic/inference.py
    (vendored unmodified into oracle/_ref by oracle/make_ref.py) on a `DualARTransformer` of the S2-Pro shape, bf16, all host
    threads, 512-token prompt, `frames` new tokens per call.  The frames after the prompt pass are what is timed
    (`decode_n_tokens`, inference.py:165-238), like the GPU arm's `value`; the prompt pass is reported in the sample text.
    Returns (audio-sec/sec, cores, sample description, seconds per frame)."""
    import torch
    from oracle import dual_ar_ref as R
    from oracle import make_ref
    llama, inf = make_ref.import_vendored()
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = R.s2_config(max_seq_len=2048)
    g = torch.Generator().manual_seed(0)
    block = (torch.randn(1 << 20, generator=g) * cfg.initializer_range).to(torch.bfloat16)
    with torch.device("meta"):   # 4.5 B parameters: no default initialisation pass, the tiled block below is assigned instead
        m = llama.DualARTransformer(llama.DualARModelArgs(**cfg.to_ref_kwargs())).eval().to(torch.bfloat16)
    sd = {}
    for name, prm in m.state_dict().items():
        n = prm.numel()
        sd[name] = (torch.ones(prm.shape, dtype=torch.bfloat16) if prm.dim() == 1
                    else block.repeat((n + block.numel() - 1) // block.numel())[:n].view(prm.shape))
    m.load_state_dict(sd, assign=True, strict=True)
    # the non-persistent buffers were built on the meta device: rebuild them as the constructors do (llama.py:279-297, 696-704)
    m.freqs_cis = llama.precompute_freqs_cis(cfg.max_seq_len, cfg.head_dim, cfg.rope_base)
    m.causal_mask = torch.tril(torch.ones(cfg.max_seq_len, cfg.max_seq_len, dtype=torch.bool))
    m.fast_freqs_cis = llama.precompute_freqs_cis(cfg.num_codebooks, m.config.fast_head_dim, cfg.rope_base)

    class Tok:   # generate() asks the tokenizer for <|im_end|> only (inference.py:207, 320)
        def get_token_id(self, _name):
            return cfg.im_end_id
    m.tokenizer = Tok()
    prompt = R.make_prompt(cfg, PROMPT_T, seed=1)
    real_n = inf.decode_n_tokens
    spans = []

    def timed_n(*a, **k):
        t0 = time.perf_counter()
        r = real_n(*a, **k)
        spans.append((time.perf_counter() - t0, r.shape[1]))
        return r
    inf.decode_n_tokens = timed_n
    times, prefill_s = [], 0.0
    try:
        with torch.inference_mode():
            for it in range(warmup + steps):
                t0 = time.perf_counter()
                inf.generate(model=m, prompt=prompt, max_new_tokens=frames + 1, audio_masks=None, audio_parts=None,
                             temperature=0.8, top_p=0.8, top_k=30)
                total = time.perf_counter() - t0
                dt, n = spans[-1]
                prefill_s = total - dt
                if it >= warmup and n > 0:
                    times.append(dt / n)
    finally:
        inf.decode_n_tokens = real_n
    spf = statistics.mean(times)
    sample = (f"reference generate() (oracle/_ref, unmodified): {frames} decode frames/step x {steps} step(s) after its {PROMPT_T}-token "
              f"prompt pass ({prefill_s:.1f} s, not counted), S2-Pro shape bf16, bs1, max_seq_len 2048")
    return FRAME_SEC / spf, cores, sample, spf


def cpu_leg(frames: int, steps: int = 1, warmup: int = 0):
    """(value, cores, sample, s/frame, kind): the vendored reference when oracle/_ref exists, else the oracle port."""
    from oracle import make_ref
    if make_ref.available():
        try:
            return (*cpu_reference_leg_ref(frames, steps, warmup), "reference")
        except Exception as e:  # a broken vendored copy must not cost the baseline
            sys.stderr.write(f"[bench] vendored reference failed ({e}); timing the oracle port instead\n")
    return (*cpu_reference_leg(frames, steps, warmup), "port")


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    frames = max(1, args.ref_frames)
    v, cores, sample, spf, kind = cpu_leg(frames, steps=max(1, args.steps), warmup=min(args.warmup, 1))
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "audio-s/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": spf * 1e3 * frames, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "S2-Pro-shape Dual-AR decode, batch 1, 512-token prompt, %d-frame decode (BASELINE configs[1])" % args.frames,
                   "prompt_tokens": PROMPT_T, "frames_per_step": frames,
                   "sample": "host cores, %d frames per step after the prompt pass (a bounded sample of the %d-frame workload)" % (frames, args.frames)},
        "cpu_baseline": {"value": v, "unit": "audio-s/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ----------------------------------------------------------------------------------------------- GPU arm
def time_dominant_kernel(model, iters=4):
    """Dominant kernel = the fused RMSNorm + gate/up GEMV + SiLU*mul (gemv_kernel<1,4,RMSNORM,SWIGLU>): 2 x ffn x dim
    bf16 weights per launch.  Timed ALONE with CUDA events on the launch stream, cycling over all slow layers so
    every launch streams weights that are not in L2 (36 x 99.6 MB >> 126 MB)."""
    import torch
    c = model.config
    lib = model.lib
    x = torch.randn(1, c.dim, device=model.device).bfloat16()
    out = torch.empty(1, c.intermediate_size, dtype=torch.bfloat16, device=model.device)
    st = torch.cuda.current_stream().cuda_stream
    layers = [(model.weights[f"layers.{i}.feed_forward.w1.weight"], model.weights[f"layers.{i}.feed_forward.w3.weight"],
               model.weights[f"layers.{i}.ffn_norm.weight"]) for i in range(c.n_layer)]

    def sweep():
        for w1, w3, nw in layers:
            rc = lib.fsb_gemv(w1.data_ptr(), w3.data_ptr(), x.data_ptr(), nw.data_ptr(), None, out.data_ptr(), 1,
                              c.intermediate_size, c.dim, 1, 2, c.norm_eps, st)
            assert rc == 0
    sweep()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        sweep()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / (iters * len(layers))
    nbytes = 2.0 * c.intermediate_size * c.dim * 2 + 2.0 * c.dim * 2 + c.intermediate_size * 2 + c.dim * 2
    return nbytes, ms


def measure_batched(dev, rank, world, batch=64, ctx=2048, frames=48, steps=2):
    """BASELINE configs[2] beside the headline: `batch` requests per GPU, each with a `ctx`-token KV context already
    resident (synthetic prefill: caches filled with N(0, 0.5)), decoded together for `frames` frames per step; every
    Linear runs once over all rows on the tcgen05 stream-K kernel (csrc/rows_gemm.cuh), attention on attn_tc_kernel.
    Returns (ms per step measured with CUDA events, algorithmic bytes per step, launches per step)."""
    import torch
    from fish_speech_b200 import DualARTransformerB200, s2_pro_config
    S = (ctx + frames + 8 + 7) // 8 * 8
    cfg = s2_pro_config(max_seq_len=S)
    sd = synth_state_dict(cfg, dev, seed=0)
    model = DualARTransformerB200(cfg, sd, device=dev, max_rows=batch)
    del sd
    model.setup_caches(batch, S)
    g = torch.Generator(device=dev).manual_seed(3 + rank)
    for k, v in model.kv_cache:
        k[:, :, :ctx].normal_(0.0, 0.5, generator=g)
        v[:, :, :ctx].normal_(0.0, 0.5, generator=g)
    sess = model.session(batch, frames + 1)
    bias = torch.full((cfg.vocab_size,), float("-inf"), dtype=torch.bfloat16, device=dev)
    bias[cfg.semantic_begin_id: cfg.semantic_end_id + 1] = 0.0
    sess.bias, sess.restrict_scan, sess.seed = bias, True, 99 + rank
    C1 = cfg.num_codebooks + 1
    start = torch.zeros((batch, C1), dtype=torch.int32, device=dev)
    start[:, 0] = cfg.semantic_begin_id + torch.randint(0, cfg.codebook_size, (batch,), device=dev, generator=g).int()
    start[:, 1:] = torch.randint(0, cfg.codebook_size, (batch, C1 - 1), device=dev, generator=g).int()

    def reset():
        sess.cur_tokens.copy_(start)
        sess.pos.fill_(ctx)
        sess.frame_counter.zero_()
        sess.ras_window.zero_()
        sess.done.zero_()
    for _ in range(3):
        reset()
        sess.run(min(frames, 8), 0.8008, 0.8008, 30, True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        reset()
        sess.run(frames, 0.8008, 0.8008, 30, True)
    e1.record()
    torch.cuda.synchronize()
    ms_step = e0.elapsed_time(e1) / (steps * frames)
    ok = bool((sess.out_tokens[:frames, :, 1:] >= 0).all()) and bool((sess.out_tokens[:frames, :, 1:] < cfg.codebook_size).all())
    fbytes = model.frame_bytes(batch, int((ctx + (frames - 1) / 2.0) * batch), masked_head=True)  # restricted LM head (V_eff rows)
    launches = model.launches_per_frame(batch)
    del sess, model
    torch.cuda.empty_cache()
    return ms_step, fbytes, launches, ok


def measure_dac(dev, B=32, T=646, iters=2):
    """BASELINE configs[3]: DAC codec decode of B clips x T code frames (30 s each) -> 44.1 kHz waveform, synthetic weights
    and codes; the step includes the H2D of the codes (int64 [B][10][T]) - the waveform stays on the device.
    Returns (ms per decode, audio seconds per decode, GEMM-shaped FLOPs per decode, finite)."""
    import torch
    from fish_speech_b200.dac import DACB200, DACConfig, synthetic_state_dict
    cfg = DACConfig()
    m = DACB200(cfg, synthetic_state_dict(cfg, 0), device=dev, max_batch=8, max_frames=max(T, 8))
    g = torch.Generator().manual_seed(2)
    codes = torch.randint(0, cfg.codebook_size, (B, cfg.n_codebooks + 1, T), generator=g)
    codes[:, 0] = torch.randint(0, cfg.semantic_codebook_size, (B, T), generator=g)
    codes = codes.pin_memory()
    for _ in range(2):
        wav = m.from_indices(codes.to(dev, non_blocking=True))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        wav = m.from_indices(codes.to(dev, non_blocking=True))
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    out = (ms, B * T * cfg.hop / cfg.sample_rate, m.flops(B, T), bool(torch.isfinite(wav).all()))
    del m, wav
    torch.cuda.empty_cache()
    return out


def measure_dac_encode(dev, seconds=30.0, iters=3):
    """The codec's encode side (the reference voice of a cloning request -> codes): one clip of `seconds` seconds
    from pinned host memory -> codes on the device; fp32 CUDA-core path.  -> {"ms", "audio_s_per_s", ...}"""
    import torch
    from fish_speech_b200.dac import DACB200, DACConfig, synthetic_state_dict
    cfg = DACConfig()
    n = int(seconds * cfg.sample_rate)
    m = DACB200(cfg, synthetic_state_dict(cfg, 0, encode_side=True), device=dev, max_batch=1, max_frames=-(-n // cfg.hop))
    g = torch.Generator().manual_seed(3)
    audio = (0.3 * torch.randn((1, 1, n), generator=g)).pin_memory()
    codes, _ = m.encode(audio.to(dev, non_blocking=True))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        again, _ = m.encode(audio.to(dev, non_blocking=True))
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    out = {"workload": f"DAC.encode, 1 clip x {seconds:.0f} s at 44.1 kHz -> {codes.shape[2]} code frames x {codes.shape[1]} codebooks",
           "ms": ms, "value": seconds / (ms * 1e-3), "unit": "audio-s/s", "dtype": "f32",
           "deterministic": bool(torch.equal(again, codes))}
    del m
    torch.cuda.empty_cache()
    return out


def cpu_dac_encode_leg(seconds=2.0):
    """The reference's CPU encode path on a bounded sample: `seconds` of audio, fp32, all host threads: `DAC.encode`
    (modded_dac.py:874-923) of the vendored classes when oracle/_ref is present, else the oracle port pinned to them."""
    import torch
    from oracle import dac_ref as D
    from oracle import make_ref
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    c = D.DACConfig()
    sd = {**D.make_state_dict(c, seed=0), **D.make_encode_state_dict(c, seed=0)}
    audio = D.make_audio(c, 1, int(seconds * c.sample_rate))
    kind, fn = "port", (lambda: D.encode(c, sd, audio))
    if make_ref.available() and (make_ref.DEST / "fish_speech/models/dac/modded_dac.py").exists():
        try:
            from oracle.gen_golden_dac import build_reference_dac
            ref = build_reference_dac(c, importer=make_ref.import_vendored, encode_side=True)
            ref.load_state_dict(sd, strict=False)
            kind, fn = "reference", (lambda: ref.encode(audio))
        except Exception as e:
            sys.stderr.write(f"[bench] vendored reference codec (encode) failed ({e}); timing the oracle port instead\n")
    with torch.inference_mode():
        t0 = time.perf_counter()
        fn()
        dt = time.perf_counter() - t0
    what = "reference DAC.encode (oracle/_ref, unmodified)" if kind == "reference" else "oracle port of DAC.encode"
    return {"value": seconds / dt, "unit": "audio-s/s", "cores": cores, "kind": kind,
            "sample": f"1 clip x {seconds:.0f} s of audio, fp32, {what}"}


def cpu_dac_leg(frames=215):
    """The reference's CPU codec path on a bounded sample: ONE clip of `frames` code frames (10 s of audio), fp32, all host
    threads; audio-s/s.  With oracle/_ref present it is the reference's own `DAC.from_indices` (modded_dac.py:925-927, classes
    instantiated as configs/modded_dac_vq.yaml describes), else oracle/dac_ref.py, the restatement pinned to those classes."""
    import torch
    from oracle import dac_ref as D
    from oracle import make_ref
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    c = D.DACConfig()
    sd = D.make_state_dict(c, seed=0)
    codes = D.make_codes(c, 1, frames)
    kind, fn = "port", (lambda: D.from_indices(c, sd, codes))
    if make_ref.available() and (make_ref.DEST / "fish_speech/models/dac/modded_dac.py").exists():
        try:
            from oracle.gen_golden_dac import build_reference_dac
            ref = build_reference_dac(c, importer=make_ref.import_vendored)
            ref.load_state_dict(sd, strict=False)
            kind, fn = "reference", (lambda: ref.from_indices(codes))
        except Exception as e:
            sys.stderr.write(f"[bench] vendored reference codec failed ({e}); timing the oracle port instead\n")
    with torch.inference_mode():
        t0 = time.perf_counter()
        fn()
        dt = time.perf_counter() - t0
    what = "reference DAC.from_indices (oracle/_ref, unmodified)" if kind == "reference" else "oracle port of DAC.from_indices"
    return {"value": frames * c.hop / c.sample_rate / dt, "unit": "audio-s/s", "cores": cores, "kind": kind,
            "sample": f"1 clip x {frames} code frames ({frames * c.hop / c.sample_rate:.1f} s of audio), fp32, {what}; "
                      "32 x 30 s clips would take 96x as long"}


def measure_scheduler(dev, rows=64, n_req=192, voice=384):
    """Continuous batching (fish_speech_b200/scheduler.py) on one GPU: `n_req` requests through `rows` rows, each a
    prompt of a shared `voice`-column prefix (a reference voice) + 32-128 own columns and 48-192 frames (<|im_end|> masked so
    the lengths are the budgets); wall clock from the first submit to the last result, prompts resident on the device."""
    import queue as _q
    import torch
    from fish_speech_b200 import DualARTransformerB200, s2_pro_config
    from fish_speech_b200.scheduler import ContinuousBatcher, GenerateRequest
    cfg = s2_pro_config(max_seq_len=1024)
    model = DualARTransformerB200(cfg, synth_state_dict(cfg, dev, seed=0), device=dev, max_rows=rows)
    model.setup_caches(rows, cfg.max_seq_len)
    model.ignore_eos = True
    g = torch.Generator().manual_seed(0)
    reqs, vp = [], synth_prompt(cfg, voice, seed=3)
    for i in range(n_req):
        T = voice + int(torch.randint(32, 129, (1,), generator=g))
        p = synth_prompt(cfg, T, seed=10 + i)
        p[:, :voice] = vp
        reqs.append(GenerateRequest(request=dict(prompt=p.to(dev), max_new_tokens=int(torch.randint(48, 193, (1,), generator=g)),
                                                 temperature=0.8, top_p=0.8, top_k=30), response_queue=_q.Queue()))
    b = ContinuousBatcher(model, chunk_frames=8)
    for it in reqs[:rows]:      # warm-up: graph capture, tensor maps
        b.submit(it)
    b.step()
    b.active.clear(); b.waiting.clear(); b.resident.clear(); b.park.clear(); b.key = None
    b.steps = b.frames_run = b.frames_kept = b.prompt_cols = b.shared_cols = 0
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for it in reqs:
        b.submit(it)
    done = b.run_until_idle()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    frames = sum(int(c.shape[1]) + 1 for _, c in done)
    out = {"workload": f"{n_req} requests through {rows} rows (continuous batching behind the reference's queue protocol): a shared {voice}-column "
                       "voice prefix + 32-128 own prompt columns, 48-192 frames each, S2-Pro shape, one GPU",
           "value": frames * FRAME_SEC / dt, "unit": "audio-s/s", "wall_s": dt, "frames_delivered": frames, "chunk_steps": b.steps,
           "row_occupancy": b.frames_kept / max(1, b.frames_run), "prompt_columns": b.prompt_cols, "prompt_columns_shared": b.shared_cols}
    del b, model
    torch.cuda.empty_cache()
    return out


def measure_config5(dev, rank, T=8192, F=4096):
    """BASELINE configs[4]: long-form voice clone on one GPU: T-token prompt (pinned host), prompt pass (own tcgen05 GEMM + flash
    attention), F decode frames at max_seq_len 16384 (context 8k -> 12k: kernel chain; the persistent frame kernel's tiled
    instantiation, option mega_long = 1, measured slower at this length), then the DAC on the F frames; wall clock around everything incl. H2D of the prompt and D2H of the
    waveform, plus CUDA-event times of the three phases."""
    import torch
    from fish_speech_b200 import DualARTransformerB200, generate, s2_pro_config
    from fish_speech_b200.dac import DACB200, DACConfig, synthetic_state_dict
    import fish_speech_b200.dual_ar as DA
    cfg = s2_pro_config(max_seq_len=16384)
    model = DualARTransformerB200(cfg, synth_state_dict(cfg, dev, seed=0), device=dev, max_rows=1)
    model.setup_caches(1, cfg.max_seq_len)
    model.ignore_eos = True
    dcfg = DACConfig()
    span = T - T // 8
    dac = DACB200(dcfg, synthetic_state_dict(dcfg, 0, encode_side=True), device=dev, max_batch=1, max_frames=max(F, span))
    prompt = synth_prompt(cfg, T, seed=11 + rank)
    # voice-clone prompt: the last 7/8 are reference-audio frames (semantic tokens + codes, content_sequence.py:206-209), and the
    # codes are what the codec's encode side makes of a (synthetic) reference waveform of that length (vq_manager.py:24-52);
    # timed on its own: the reference voice is encoded once, not per request
    g = torch.Generator().manual_seed(5 + rank)
    ref_wav = (0.3 * torch.randn((1, 1, span * dcfg.hop), generator=g)).pin_memory()
    dac.encode(ref_wav[:, :, : 8 * dcfg.hop].to(dev))
    torch.cuda.synchronize()
    ee = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ee[0].record()
    codes = dac.encode(ref_wav.to(dev, non_blocking=True))[0][0].cpu()          # [10, span]
    ee[1].record()
    torch.cuda.synchronize()
    enc_ms = ee[0].elapsed_time(ee[1])
    prompt[1:, T - span:] = codes
    prompt[0, T - span:] = cfg.semantic_begin_id + codes[0]
    hp = prompt.pin_memory()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    marks = {}
    real_n = DA.decode_n_tokens

    def timed_n(*a, **k):
        ev[1].record()
        r = real_n(*a, **k)
        ev[2].record()
        return r
    # warm-up on a short prompt: graph capture, tensor maps, lazy module loading
    generate(model=model, prompt=hp[:, :256].to(dev), max_new_tokens=8, temperature=0.8, top_p=0.8, top_k=30)
    dac.from_indices(torch.zeros((1, dcfg.n_codebooks + 1, 64), dtype=torch.long, device=dev))
    torch.cuda.synchronize()
    DA.decode_n_tokens = timed_n
    try:
        t0 = time.perf_counter()
        ev[0].record()
        seq = generate(model=model, prompt=hp.to(dev, non_blocking=True), max_new_tokens=F, temperature=0.8, top_p=0.8, top_k=30)
        new = seq[1:, T:][None].contiguous()     # [1, 10, F] codes of the generated frames (the codec clamps them like rvq.py:354-359)
        wav = dac.from_indices(new)
        ev[3].record()
        host = wav.cpu()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
    finally:
        DA.decode_n_tokens = real_n
    pre_ms, dec_ms, dac_ms = ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]), ev[2].elapsed_time(ev[3])
    n_new = seq.shape[1] - T
    c = cfg
    p_layer = c.dim * (c.n_head + 2 * c.n_local_heads) * c.head_dim + c.n_head * c.head_dim * c.dim + 3 * c.dim * c.intermediate_size
    pre_flops = 2.0 * c.n_layer * p_layer * T + 4.0 * c.n_layer * c.n_head * c.head_dim * T * T / 2
    peak, _ = measured_peaks()
    pk = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
    tpeak = float(pk.get("bf16_tflops_sustained", pk.get("bf16_tflops", 1395.3)))
    fbytes = model.frame_bytes(1, int(T + (n_new - 1) / 2.0), masked_head=True)
    ms_frame = dec_ms / max(1, n_new - 1)
    out = {
        "workload": f"long-form voice clone (BASELINE configs[4]) per GPU: {T}-token prompt, {n_new} frames decoded at max_seq_len 16384, DAC on the frames",
        "value": n_new * FRAME_SEC / wall, "unit": "audio-s/s", "wall_s": wall, "audio_s": n_new * FRAME_SEC,
        "h2d_bytes": int(prompt.numel() * 8), "d2h_bytes": int(host.numel() * 4), "finite": bool(torch.isfinite(host).all()),
        "prefill": {"ms": pre_ms, "tflops": pre_flops / pre_ms / 1e9, "frac_of_tensor_peak": pre_flops / pre_ms / 1e9 / tpeak,
                    "note": "includes the first frame; GEMM + causal attention FLOPs (SURVEY 8d); tcv_kernel + fa_prefill_kernel (tcgen05)"},
        "decode": {"ms_per_frame": ms_frame, "gbs": fbytes / ms_frame / 1e6, "frac_of_hbm": fbytes / ms_frame / 1e6 / peak,
                   "launches_per_frame": model.launches_per_frame(1),
                   "note": ("context 8k -> 12k: frame_mega_kernel<128,4,true> (persistent, attention splits walk their keys in tiles)"
                            if model.launches_per_frame(1) == 1 else
                            "context 8k -> 12k: kernel chain (PDL-chained launches per frame in a CUDA graph)") + ", restricted LM head"},
        "dac": {"ms": dac_ms, "tflops": dac.flops(1, n_new) / dac_ms / 1e9, "frac_of_tensor_peak": dac.flops(1, n_new) / dac_ms / 1e9 / tpeak},
        "reference_encode": {"ms": enc_ms, "frames": span, "audio_s": span * FRAME_SEC,
                             "note": "DAC.encode of the reference waveform whose codes fill the prompt's audio span (fp32; once per voice, "
                                     "outside wall_s; H2D of the waveform and D2H of the codes inside ms)"},
    }
    del model, dac
    torch.cuda.empty_cache()
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist
    from fish_speech_b200 import DualARTransformerB200, generate, partition, s2_pro_config
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    F = args.frames
    cfg = s2_pro_config(max_seq_len=2048)
    sd = synth_state_dict(cfg, dev, seed=0)
    model = DualARTransformerB200(cfg, sd, device=dev, max_rows=1)
    del sd
    model.setup_caches(1, cfg.max_seq_len)
    # one utterance per GPU, handed out by the request partitioner (fish_speech_b200/partition.py)
    utterance = partition.assign(world, world)[rank][0]
    prompt = synth_prompt(cfg, PROMPT_T, seed=1 + utterance)
    C1 = cfg.num_codebooks + 1

    # ---- device-resident leg (`value`): prompt pass once, then K timed steps of F frames from position 512
    model.prefill(prompt[:, :-1].to(dev), 0, 0)
    sess = model.session(1, F + 1)
    bias = torch.full((cfg.vocab_size,), float("-inf"), dtype=torch.bfloat16, device=dev)
    bias[cfg.semantic_begin_id: cfg.semantic_end_id + 1] = 0.0  # <|im_end|> masked: fixed frame count
    sess.bias, sess.restrict_scan, sess.seed = bias, True, 1234 + rank
    last_col = prompt[:, -1].to(torch.int32).view(1, C1).to(dev)

    def reset():
        sess.cur_tokens.copy_(last_col)
        sess.pos.fill_(PROMPT_T - 1)
        sess.frame_counter.zero_()
        sess.ras_window.zero_()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        reset()
        sess.run(min(F, 64), 0.8008, 0.8008, 30, True)
    torch.cuda.synchronize()
    # p50 ms/frame: per-frame CUDA events over one step
    reset()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(min(F, 256) + 1)]
    evs[0].record()
    for i in range(1, len(evs)):
        sess.run(1, 0.8008, 0.8008, 30, True)
        evs[i].record()
    torch.cuda.synchronize()
    p50 = statistics.median(evs[i].elapsed_time(evs[i + 1]) for i in range(len(evs) - 1))

    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        reset()
        sess.run(F, 0.8008, 0.8008, 30, True)
    e1.record()
    barrier()
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1)
    codes_ok = bool((sess.out_tokens[:F, 0, 1:] >= 0).all()) and bool((sess.out_tokens[:F, 0, 1:] < cfg.codebook_size).all())

    # ---- end-to-end leg (`e2e`): public generate() with a HOST prompt (pinned), result read back to the host
    hp = prompt.pin_memory()
    e2e_frames, e2e_t = 0, 0.0
    # same fixed-length workload as the device-resident leg: generate() builds its own constrained-decode bias
    # (inference.py:308-320); the <|im_end|> entry is masked here so every run decodes exactly F frames
    model.ignore_eos = True  # supported switch (dual_ar.make_semantic_logit_bias): same restricted LM head / scan as the device-resident leg
    def one_utterance(u):  # the public call a user makes; result back on the host
        assert u == utterance
        return generate(model=model, prompt=hp.to(dev, non_blocking=True), max_new_tokens=F, temperature=0.8, top_p=0.8, top_k=30).cpu()

    for it in range(1 + args.e2e_steps):
        barrier()
        t0 = time.perf_counter()
        (_, host), = partition.run_partitioned(list(range(world)), one_utterance, gather=False)  # this rank's share: one utterance
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if it > 0:
            e2e_frames += host.shape[1] - PROMPT_T
            e2e_t += dt
    model.ignore_eos = False
    b64 = b64_strong = None
    if not args.no_batched:
        fb1 = model.frame_bytes(1, int(PROMPT_T + (F - 1) / 2.0), masked_head=True)
        launches_bs1 = model.launches_per_frame(1)
        kb_kms = time_dominant_kernel(model)
        del sess, model
        torch.cuda.empty_cache()
        b64 = measure_batched(dev, rank, world, batch=args.batch, ctx=args.batch_ctx, frames=args.batch_frames)
        if world > 1:
            # strong-scaling pair (SURVEY §8e "report both"): the SAME pool of `batch` requests spread over the GPUs
            pool = partition.assign(args.batch, world)[rank]
            b64_strong = measure_batched(dev, rank, world, batch=len(pool), ctx=args.batch_ctx, frames=args.batch_frames)
    dac = None
    if not args.no_dac:
        try:
            dac = measure_dac(dev)
        except Exception as e:  # the Dual-AR numbers stand on their own
            dac = e
    dac_cpu = None
    if not args.no_dac and not args.no_cpu and rank == 0:
        try:
            dac_cpu = cpu_dac_leg()
        except Exception as e:
            dac_cpu = {"value": None, "sample": f"failed: {e}"}
    dac_enc = None
    if not args.no_dac and world == 1:
        try:
            dac_enc = measure_dac_encode(dev)
            if not args.no_cpu:
                dac_enc["cpu_baseline"] = cpu_dac_encode_leg()
        except Exception as e:
            dac_enc = {"error": str(e)}
    sched = None
    if args.sched and world == 1:
        try:
            sched = measure_scheduler(dev)
        except Exception as e:
            sched = {"error": str(e)}
    c5 = None
    if args.config5:
        try:
            c5 = measure_config5(dev, rank)
        except Exception as e:
            c5 = {"error": str(e)}
    tms = torch.tensor([ms, e2e_t, b64[0] if b64 else 0.0, dac[0] if isinstance(dac, tuple) else 0.0,
                        b64_strong[0] if b64_strong else 0.0], dtype=torch.float64, device=dev)
    cnt = torch.tensor([float(args.steps * F), float(e2e_frames)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    ms_max, e2e_t_max, b64_ms_max, dac_ms_max, strong_ms_max = (float(x) for x in tms)
    total_frames, total_e2e_frames = float(cnt[0]), float(cnt[1])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peak, peak_src = measured_peaks()
    value = total_frames * FRAME_SEC / (ms_max / 1e3)
    ms_per_frame = ms_max / (args.steps * F)
    ctx_avg = PROMPT_T + (F - 1) / 2.0
    if b64 is not None:
        fbytes, launches, (kb, kms) = fb1, launches_bs1, kb_kms
    else:
        fbytes = model.frame_bytes(1, int(ctx_avg), masked_head=True)
        kb, kms = time_dominant_kernel(model)
        launches = model.launches_per_frame(1)
    persistent = launches == 1
    line = {
        "metric": METRIC, "value": value, "unit": "audio-s/s", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
        "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "S2-Pro-shape Dual-AR decode, batch 1, 512-token prompt, %d-frame decode (BASELINE configs[1])" % F,
                   "frames_per_step": F, "prompt_tokens": PROMPT_T, "utterances": world, "parallelism": f"replica x{world}, request-partitioned, no collective",
                   "l2_policy": "inputs larger than L2: 16.3 GB of weights streamed per frame vs 126 MB L2",
                   "sampling": "T=0.8 top_p=0.8 top_k=30, RAS on, <|im_end|> masked for a fixed frame count",
                   "lm_head": "fused frames compute only the LM-head rows the constrained-decode sampler can read (semantic range + <|im_end|>: "
                              "4136 of 155776 rows, SURVEY §8d V_eff); the roofline bytes count those rows only"},
        "p50_ms_per_frame": p50, "mean_ms_per_frame": ms_per_frame, "codes_in_range": codes_ok,
        "clocks": clocks,
        "e2e": {"value": (total_e2e_frames * FRAME_SEC / e2e_t_max) if e2e_t_max > 0 else None, "unit": "audio-s/s",
                "h2d_bytes_per_step": int(prompt.numel() * 8), "d2h_bytes_per_step": int(C1 * (PROMPT_T + F) * 8),
                "frames": total_e2e_frames, "includes": "pinned-host prompt H2D, 512-token prompt pass, decode loop, D2H of codes"},
        "gpu_launches": int(launches * F * args.steps),
        # dominant kernel: with the persistent frame kernel ONE launch is the whole frame (every weight byte of the frame
        # streams through it), so its launch duration is the frame time measured above with CUDA events
        "roofline": ({"bound": "hbm", "kernel": "frame_mega_kernel<128,4> (persistent cooperative kernel, one launch = one frame)",
                      "achieved": fbytes / (ms_per_frame * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                      "frac": fbytes / (ms_per_frame * 1e-3) / 1e9 / peak, "traffic": ncu_traffic("frame_mega_kernel")[0],
                      "traffic_source": ncu_traffic("frame_mega_kernel")[1],
                      "bytes_per_launch": fbytes, "ms_per_launch": ms_per_frame, "peak_source": peak_src}
                     if persistent else
                     {"bound": "hbm", "kernel": "gemv_tma_kernel<M=1,RMSNORM,SWIGLU> (w1/w3 gate-up, timed alone over 36 layers)",
                      "achieved": kb / (kms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s", "frac": kb / (kms * 1e-3) / 1e9 / peak,
                      "traffic": None, "bytes_per_launch": kb, "ms_per_launch": kms, "peak_source": peak_src}),
        "roofline_gemv_alone": {"kernel": "gemv_tma_kernel<M=1,RMSNORM,SWIGLU> (w1/w3, standalone launches over 36 layers)",
                                "achieved": kb / (kms * 1e-3) / 1e9, "frac": kb / (kms * 1e-3) / 1e9 / peak, "unit": "GB/s",
                                "bytes_per_launch": kb, "ms_per_launch": kms},
        "roofline_frame": {"bound": "hbm", "achieved": fbytes / (ms_per_frame * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                           "frac": fbytes / (ms_per_frame * 1e-3) / 1e9 / peak, "bytes_per_frame": fbytes,
                           "note": "algorithmic weight+KV bytes per frame (SURVEY §8d, restricted LM head: V_eff rows) / mean frame time"},
    }
    if b64 is not None:
        # BASELINE configs[2] (metric: "bs1 & bs64"): weak scaling, every GPU decodes its own `batch` requests
        line["bs64"] = {
            "workload": "S2-Pro-shape Dual-AR batched decode, %d requests per GPU x %d GPU(s), %d-token KV context resident, %d frames/step (BASELINE configs[2])"
                        % (args.batch, world, args.batch_ctx, args.batch_frames),
            "value": world * args.batch * FRAME_SEC / (b64_ms_max * 1e-3), "unit": "audio-s/s", "ms_per_step": b64_ms_max,
            "frames_per_s": world * args.batch / (b64_ms_max * 1e-3), "codes_in_range": b64[3], "launches_per_step": b64[2],
            "roofline": {"bound": "hbm", "achieved": b64[1] / (b64_ms_max * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": b64[1] / (b64_ms_max * 1e-3) / 1e9 / peak, "bytes_per_step": b64[1],
                         "note": "algorithmic weight + KV bytes of one step for all requests (SURVEY §8d) / CUDA-event step time; "
                                 "kernels: rows_gemm_kernel (tcgen05 stream-K), attn_tc_kernel, finish kernels"},
        }
        if b64_strong is not None:
            per_gpu = -(-args.batch // world)
            line["bs64"]["strong"] = {
                "workload": "the same %d requests spread over %d GPUs by partition.assign (%d per GPU), same context and frames" % (args.batch, world, per_gpu),
                "value": args.batch * FRAME_SEC / (strong_ms_max * 1e-3), "unit": "audio-s/s", "ms_per_step": strong_ms_max,
                "speedup_vs_one_gpu": b64_ms_max / strong_ms_max,
                "roofline_limit": b64[1] / b64_strong[1],
                "note": "one GPU's step over all %d requests is this run's weak-scaling step (every GPU runs it); the limit is the ratio of the "
                        "algorithmic bytes of the two steps: the weights stream once per step for any number of rows" % args.batch}
    if isinstance(dac, tuple):
        pk = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
        tpeak = float(pk.get("bf16_tflops_sustained", pk.get("bf16_tflops", 1375.8)))
        tf = dac[2] / (dac_ms_max * 1e-3) / 1e12
        line["dac"] = {"workload": "DAC codec decode, 32 clips x 646 code frames (30 s) per GPU x %d GPU(s) -> 44.1 kHz (BASELINE configs[3])" % world,
                       "value": world * dac[1] / (dac_ms_max * 1e-3), "unit": "audio-s/s", "ms_per_step": dac_ms_max, "finite": dac[3],
                       "roofline": {"bound": "tensor", "achieved": tf, "peak": tpeak, "unit": "TFLOP/s", "frac": tf / tpeak,
                                    "kernel": "tcv_kernel (csrc/tc_conv.cuh): tcgen05 implicit-GEMM conv, fp16 operands, fp32 TMEM accumulation",
                                    "note": "algorithmic GEMM-shaped FLOPs of the decode (fsb_dac_flops, SURVEY 8d: 6.76 GFLOP per code frame) / CUDA-event time, "
                                            "against the measured sustained bf16/fp16 tensor peak"}}
        if dac_cpu is not None:
            line["dac"]["cpu_baseline"] = dac_cpu
        if dac_enc is not None:
            line["dac"]["encode"] = dac_enc
    elif dac is not None:
        line["dac"] = {"error": str(dac)}
    if c5 is not None:
        line["config5"] = c5
    if sched is not None:
        line["scheduler"] = sched
    if not args.no_cpu and world >= 1:
        try:
            v, cores, sample, _, kind = cpu_leg(args.ref_frames)
            line["cpu_baseline"] = {"value": v, "unit": "audio-s/s", "cores": cores, "kind": kind, "sample": sample}
        except Exception as e:  # the GPU numbers stand on their own
            line["cpu_baseline"] = {"value": None, "unit": "audio-s/s", "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {e}"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--frames", type=int, default=1024)
    ap.add_argument("--ref-frames", type=int, default=2)
    ap.add_argument("--e2e-steps", type=int, default=1)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-batched", action="store_true", help="skip the config-3 (bs64) section")
    ap.add_argument("--no-dac", action="store_true", help="skip the config-4 (DAC decode) section")
    ap.add_argument("--no-config5", dest="config5", action="store_false", help="skip the long-form section (8k-token prompt + 4k frames + DAC, BASELINE configs[4], ~30 s)")
    ap.add_argument("--no-sched", dest="sched", action="store_false", help="skip the continuous-batching section (N = 1 only, ~15 s)")
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--batch-ctx", type=int, default=2048)
    ap.add_argument("--batch-frames", type=int, default=48)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
