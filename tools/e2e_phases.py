"""Where the end-to-end time of generate() goes (S2-Pro shape, 512-token prompt, 256 frames)."""
import sys, time
import torch
sys.path.insert(0, ".")
from bench import PROMPT_T, synth_prompt, synth_state_dict  # noqa: E402
from fish_speech_b200 import DualARTransformerB200, generate, s2_pro_config  # noqa: E402
import fish_speech_b200.dual_ar as D  # noqa: E402

dev = torch.device("cuda", 0)
cfg = s2_pro_config(max_seq_len=2048)
model = DualARTransformerB200(cfg, synth_state_dict(cfg, dev, seed=0), device=dev, max_rows=1)
model.setup_caches(1, cfg.max_seq_len)
prompt = synth_prompt(cfg, PROMPT_T, seed=1).to(dev)

def t(fn, n=1):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): r = fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n, r

for it in range(2):
    dt, _ = t(lambda: model.prefill(prompt[:, :-1], 0, 0))
    print(f"prefill 511 tokens: {dt*1e3:.1f} ms")
F = 256
for it in range(2):
    dt, seq = t(lambda: generate(model=model, prompt=prompt, max_new_tokens=F, temperature=0.8, top_p=0.8, top_k=30))
    print(f"generate {F} frames: {dt*1e3:.1f} ms  ({seq.shape})")
real = D.decode_n_tokens
def timed(*a, **k):
    dt, r = t(lambda: real(*a, **k)); print(f"  decode_n_tokens: {dt*1e3:.1f} ms"); return r
D.decode_n_tokens = timed
real1 = D.decode_one_token_ar
def timed1(*a, **k):
    dt, r = t(lambda: real1(*a, **k)); print(f"  decode_one_token_ar: {dt*1e3:.1f} ms"); return r
dt, seq = t(lambda: generate(model=model, prompt=prompt, max_new_tokens=F, temperature=0.8, top_p=0.8, top_k=30, decode_one_token=timed1))
print(f"generate {F} frames (instrumented): {dt*1e3:.1f} ms")
