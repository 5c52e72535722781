"""One layer of the DAC decoder on the tcgen05 conv kernel (fsb_tc_conv), timed alone with CUDA events; for ncu captures.
usage: python tools/tc_conv_bench.py <name> [iters]     names: see LAYERS (shapes at B = 8 clips x 646 frames)"""
import ctypes as C
import json
import sys

import torch

sys.path.insert(0, ".")
from fish_speech_b200 import _lib  # noqa: E402

# name -> (rows per clip-frame, Cin, N, taps, dil, res, alpha, bias_mod)
LAYERS = {
    "b1_conv7": (32, 768, 768, 7, 1, 0, 1, 0), "b1_conv1": (32, 768, 768, 1, 1, 1, 1, 0),
    "b2_conv7": (256, 384, 384, 7, 3, 0, 1, 0), "b2_conv1": (256, 384, 384, 1, 1, 1, 1, 0),
    "b3_conv7": (1024, 192, 192, 7, 9, 0, 1, 0), "b3_conv1": (1024, 192, 192, 1, 1, 1, 1, 0),
    "b4_conv7": (2048, 96, 96, 7, 9, 0, 1, 0), "b4_conv1": (2048, 96, 96, 1, 1, 1, 1, 0),
    "b1_ct": (4, 1536, 8 * 768, 2, 1, 0, 1, 768), "b3_ct": (256, 384, 4 * 192, 2, 1, 0, 1, 192),
    # what the 1-tap layer's time is made of: no residual / no Snake / neither
    "b4_conv1_nores": (2048, 96, 96, 1, 1, 0, 1, 0), "b4_conv1_nosnake": (2048, 96, 96, 1, 1, 1, 0, 0),
    "b4_conv1_plain": (2048, 96, 96, 1, 1, 0, 0, 0), "b3_conv1_nores": (1024, 192, 192, 1, 1, 0, 1, 0),
    "b3_conv1_plain": (1024, 192, 192, 1, 1, 0, 0, 0),
}
name = sys.argv[1] if len(sys.argv) > 1 else "b3_conv7"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 5
direct = len(sys.argv) > 3 and sys.argv[3] == "direct"      # the unstaged epilogue, for A/B timing
rpf, Cin, N, taps, dil, res, alpha, bias_mod = LAYERS[name]
B, T = 8, 646 * rpf
lib = _lib.load()
g = torch.Generator(device="cuda").manual_seed(0)
A = torch.randn(B, T, Cin, device="cuda", generator=g).half()
W = (torch.randn(taps, N, Cin, device="cuda", generator=g) / (taps * Cin) ** 0.5).half()
bias = torch.zeros(bias_mod or N, device="cuda")
al = torch.ones(bias_mod or N, device="cuda")
out16 = torch.empty(B, T, N, device="cuda", dtype=torch.float16)
x16 = torch.zeros(B, T, N, device="cuda", dtype=torch.float16) if res else None
p = lambda t: None if t is None else t.data_ptr()
a = _lib.TcConvArgs(A16=p(A), W16=p(W), B=B, T=T, Cin=Cin, N=N, taps=taps, dil=dil, f16=1, use_simt=-1 if direct else 0, bias=p(bias),
                    bias_mod=bias_mod, gamma=None, res32=None, res16=p(x16), ldres=N, alpha=p(al) if alpha else None,
                    alpha_mod=bias_mod, act=0, round_lin=0, out16=p(out16), ld16=N, out32=None, ld32=N, out16r=p(x16), ld16r=N)
st = torch.cuda.current_stream().cuda_stream
for _ in range(2):
    assert lib.fsb_tc_conv(C.byref(a), st) == 0
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(iters):
    lib.fsb_tc_conv(C.byref(a), st)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
flops = 2.0 * B * T * N * Cin * taps
byts = B * T * (Cin * 2 + N * 2 + (N * 4 if res else 0))
print(json.dumps({"layer": name, "epilogue": "direct" if direct else "staged", "rows": B * T, "Cin": Cin, "N": N, "taps": taps, "ms": round(ms, 4), "tflops": round(flops / ms / 1e9, 1),
                  "hbm_gbs_min": round(byts / ms / 1e6, 1)}))
