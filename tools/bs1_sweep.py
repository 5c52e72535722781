"""ms/frame of the batch-1 fused frame (S2-Pro shape, 512-token prompt) for a list of option settings, one process.
usage: python tools/bs1_sweep.py "name=v,name=v" "name=v" ...     (each argument = one setting; "" = defaults)"""
import sys

import torch

sys.path.insert(0, ".")
from bench import PROMPT_T, synth_prompt, synth_state_dict  # noqa: E402
from fish_speech_b200 import DualARTransformerB200, s2_pro_config  # noqa: E402

frames = 256
dev = torch.device("cuda", 0)
cfg = s2_pro_config(max_seq_len=2048)
model = DualARTransformerB200(cfg, synth_state_dict(cfg, dev, seed=0), device=dev, max_rows=1)
model.setup_caches(1, cfg.max_seq_len)
prompt = synth_prompt(cfg, PROMPT_T, seed=1)
model.prefill(prompt[:, :-1].to(dev), 0, 0)
sess = model.session(1, frames + 2)
bias = torch.full((cfg.vocab_size,), float("-inf"), dtype=torch.bfloat16, device=dev)
bias[cfg.semantic_begin_id: cfg.semantic_end_id + 1] = 0.0
sess.bias, sess.restrict_scan, sess.seed = bias, True, 1
last = prompt[:, -1].to(torch.int32).view(1, -1).to(dev)


def reset():
    sess.cur_tokens.copy_(last)
    sess.pos.fill_(PROMPT_T - 1)
    sess.frame_counter.zero_()
    sess.ras_window.zero_()


def measure():
    reset()
    sess.run(32, 0.8008, 0.8008, 30, True)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(2):
        reset()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        sess.run(frames, 0.8008, 0.8008, 30, True)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / frames)
    toks = sess.out_tokens[:frames, 0].clone()
    return best, toks


base_toks = None
defaults = {}
for setting in (sys.argv[1:] or [""]):
    kv = dict(x.split("=") for x in setting.split(",") if x)
    for k in list(defaults):
        if k not in kv:
            model.set_option(k, defaults[k])
    for k, v in kv.items():
        defaults.setdefault(k, 0 if k != "masked_head" else 1)
        model.set_option(k, int(v))
    ms, toks = measure()
    same = "-" if base_toks is None else str(bool(torch.equal(toks, base_toks)))
    if base_toks is None:
        base_toks = toks
    fb = model.frame_bytes(1, int(PROMPT_T + (frames - 1) / 2.0), masked_head=True)
    print(f"{setting or 'defaults':40s} {ms:.4f} ms/frame  {fb / ms / 1e6:7.1f} GB/s  frac {fb / ms / 1e6 / 6544.3:.3f}  same tokens as first: {same}", flush=True)
model._session.check_device()
