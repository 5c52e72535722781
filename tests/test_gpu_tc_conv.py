"""GPU parity of the tcgen05 implicit-GEMM conv (csrc/tc_conv.cuh) through its C-ABI stage entry point fsb_tc_conv, against a
plain PyTorch fp32 reference of the same op (F.conv1d with causal left padding on the same 16-bit-rounded operands) and
against its CUDA-core twin kernel.  Shapes: every layer class of the DAC decoder (modded_dac.py:521-801, rvq.py:58-191) at
the real widths, and the Dual-AR prompt-pass Linears (llama.py:856-860, 981-983) with the bf16 rounding points.
Tolerance: the operands are identical 16-bit values on both sides and accumulation is fp32, so the only differences are the
summation order and the 16-bit rounding of the output: |err| <= 2^-8 (bf16) / 2^-10 (fp16) of |value| + 1e-3 of the
reduction's scale."""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    from fish_speech_b200 import _lib
    _lib.require_cuda()
    return _lib.load()


def _dt(f16):
    return torch.float16 if f16 else torch.bfloat16


def _run(lib, A, W, f16, dil=1, simt=0, bias=None, bias_mod=0, gamma=None, res32=None, res16=None, alpha=None, alpha_mod=0,
         act=0, round_lin=0, want32=True, want16=True, want16r=False):
    from fish_speech_b200._lib import TcConvArgs
    B, T, Cin = A.shape
    taps, N, _ = W.shape
    n16 = N // 2 if act == 3 else N
    out16 = torch.full((B, T, n16), 7.0, dtype=_dt(f16), device="cuda") if want16 else None
    out32 = torch.full((B, T, N), 7.0, dtype=torch.float32, device="cuda") if (want32 and act != 3) else None
    out16r = torch.full((B, T, N), 7.0, dtype=_dt(f16), device="cuda") if want16r else None
    p = lambda t: None if t is None else t.data_ptr()
    a = TcConvArgs(A16=p(A), W16=p(W), B=B, T=T, Cin=Cin, N=N, taps=taps, dil=dil, f16=int(f16), use_simt=simt,
                   bias=p(bias), bias_mod=bias_mod, gamma=p(gamma), res32=p(res32), res16=p(res16), ldres=N,
                   alpha=p(alpha), alpha_mod=alpha_mod, act=act, round_lin=round_lin, out16=p(out16), ld16=n16,
                   out32=p(out32), ld32=N, out16r=p(out16r), ld16r=N)
    rc = lib.fsb_tc_conv(C.byref(a), torch.cuda.current_stream().cuda_stream)
    assert rc == 0, rc
    torch.cuda.synchronize()
    if want16r:
        return out32, out16, out16r
    return out32, out16


def _ref(A, W, dil=1, bias=None, bias_mod=0, gamma=None, res=None, alpha=None, alpha_mod=0, act=0):
    """fp32 reference: causal conv on the 16-bit-rounded operands; returns (raw fp32, fp32 value of the 16-bit copy)"""
    cpu = lambda t: None if t is None else t.detach().cpu()   # fp32 on the CPU: cuDNN would use TF32 for an fp32 conv
    A, W, bias, gamma, res, alpha = cpu(A), cpu(W), cpu(bias), cpu(gamma), cpu(res), cpu(alpha)
    B, T, Cin = A.shape
    taps, N, _ = W.shape
    x = F.pad(A.float().transpose(1, 2), ((taps - 1) * dil, 0))
    w = W.float().permute(1, 2, 0).contiguous()                      # [N][Cin][taps]
    y = F.conv1d(x, w, dilation=dil).transpose(1, 2)                 # [B][T][N]
    if bias is not None:
        y = y + bias.float().repeat(N // (bias_mod or N))
    if act == 1:
        y = F.gelu(y)
    if act == 3:
        g, u = y[..., 0::2], y[..., 1::2]
        return None, F.silu(g) * u
    if gamma is not None:
        y = y * gamma.float()
    if res is not None:
        y = y + res.float()
    if act == 2:
        y = torch.tanh(y)
    z = y
    if alpha is not None:
        al = alpha.float().repeat(N // (alpha_mod or N))
        z = y + torch.sin(al * y) ** 2 / (al + 1e-9)
    return y, z


def _close(got, ref, f16, scale, name):
    got, ref = got.float().cpu(), ref.float().cpu()
    err = (got - ref).abs()
    tol = ref.abs() * (2.0 ** -10 if f16 else 2.0 ** -8) + 1e-3 * scale
    bad = int((err > tol).sum())
    print(f"[parity] tc_conv {name}: max-abs {float(err.max()):.5f} (scale {scale:.3f}), outside tolerance {bad}")
    assert bad == 0, (name, float(err.max()))


# (B, T, Cin, N, taps, dil, extras) -- DAC layer classes at the real widths (modded_dac_vq.yaml), ragged T
CASES = [
    ("dec.in 7-tap 1024->1536 + snake", 2, 300, 1024, 1536, 7, 1, dict(bias=True, alpha=True)),
    ("block1 ct 1536->8x768 (transposed, 2 taps)", 1, 130, 1536, 8 * 768, 2, 1, dict(bias=True, bias_mod=768, alpha=True, alpha_mod=768)),
    ("block3 unit conv7 dil9 192->192 + snake", 3, 1000, 192, 192, 7, 9, dict(bias=True, alpha=True)),
    ("block3 unit conv1 192->192 + res", 3, 1000, 192, 192, 1, 1, dict(bias=True, res32=True, alpha=True)),
    ("block4 unit conv7 dil3 96->96", 2, 777, 96, 96, 7, 3, dict(bias=True, alpha=True)),
    ("block4 ct 192->2x96", 2, 515, 192, 192, 2, 1, dict(bias=True, bias_mod=96, alpha=True, alpha_mod=96)),
    ("block2 unit conv7 dil1 384->384", 1, 260, 384, 384, 7, 1, dict(bias=True, alpha=True)),
    ("tf wqkv 1024->3072", 2, 129, 1024, 3072, 1, 1, dict()),
    ("tf wo + LayerScale + res", 2, 129, 1024, 1024, 1, 1, dict(gamma=True, res32=True)),
    ("tf w13 SwiGLU 1024->2x3072", 2, 129, 1024, 6144, 1, 1, dict(act=3)),
    ("convnext pw1 GELU 1024->4096", 1, 200, 1024, 4096, 1, 1, dict(bias=True, act=1)),
    ("convnext pw2 4096->1024 gamma res", 1, 200, 4096, 1024, 1, 1, dict(bias=True, gamma=True, res32=True)),
    ("ragged N=40, Cin=24", 2, 70, 24, 40, 3, 2, dict(bias=True, alpha=True)),
    ("single row, single tile", 1, 1, 64, 16, 7, 1, dict(bias=True)),
]


@pytest.mark.parametrize("f16", [1, 0])
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_tc_conv_matches_fp32_reference(lib, case, f16):
    name, B, T, Cin, N, taps, dil, ex = case
    g = torch.Generator().manual_seed(sum(map(ord, name)) + f16)
    dt = _dt(f16)
    A = torch.randn(B, T, Cin, generator=g).to(dt).cuda()
    W = (torch.randn(taps, N, Cin, generator=g) / (taps * Cin) ** 0.5).to(dt).cuda()
    bias = torch.randn(ex.get("bias_mod") or N, generator=g).cuda() * 0.1 if ex.get("bias") else None
    gamma = (0.5 + torch.rand(N, generator=g)).cuda() if ex.get("gamma") else None
    res32 = torch.randn(B, T, N, generator=g).cuda() if ex.get("res32") else None
    alpha = (0.5 + torch.rand(ex.get("alpha_mod") or N, generator=g)).cuda() if ex.get("alpha") else None
    act = ex.get("act", 0)
    kw = dict(dil=dil, bias=bias, bias_mod=ex.get("bias_mod", 0), gamma=gamma, alpha=alpha, alpha_mod=ex.get("alpha_mod", 0), act=act)
    y_ref, z_ref = _ref(A, W, res=res32, **kw)
    o32, o16 = _run(lib, A, W, f16, res32=res32, **kw)
    scale = 1.0
    if act != 3:
        _close(o32, y_ref, True, scale, name + " fp32 out")      # the fp32 copy carries no output rounding
    _close(o16, z_ref, f16, scale, name + " 16-bit out")
    # CUDA-core twin: same operands, same epilogue code -> equal up to the summation order
    s32, s16 = _run(lib, A, W, f16, simt=1, res32=res32, **kw)
    d = (o16.float() - s16.float()).abs()
    assert float(d.max()) <= float(z_ref.abs().max()) * (2.0 ** -9 if f16 else 2.0 ** -7) + 2e-3, float(d.max())
    del s32
    # re-launch is bit-identical
    o32b, o16b = _run(lib, A, W, f16, res32=res32, **kw)
    assert torch.equal(o16, o16b)


@pytest.mark.parametrize("T,N,K", [(512, 6144, 2560), (511, 2560, 4096), (300, 9728, 2560), (512, 2560, 9728), (37, 768, 512)])
def test_tc_conv_prompt_pass_linear_bf16_rounding(lib, T, N, K):
    """Dual-AR prompt-pass Linear: bf16 in / fp32 accumulate / ONE rounding to bf16, then bf16(residual + linear)
    (llama.py:838-843) - against the same two roundings done by torch on an fp32 product."""
    g = torch.Generator().manual_seed(T + N + K)
    x = torch.randn(1, T, K, generator=g).bfloat16().cuda()
    W = (torch.randn(1, N, K, generator=g) * 0.02).bfloat16().cuda()
    res = torch.randn(1, T, N, generator=g).bfloat16().cuda()
    _, o = _run(lib, x, W, 0, want32=False)
    ref = (x.float().cpu() @ W[0].float().cpu().t())
    _close(o, ref.bfloat16().float(), False, 4 * float(ref.abs().max()), f"linear T={T} N={N} K={K}")
    _, o2 = _run(lib, x, W, 0, res16=res, round_lin=1, want32=False)
    ref2 = (res.float().cpu() + ref.bfloat16().float()).bfloat16().float()
    o2 = o2.cpu()
    err = (o2.float() - ref2).abs()
    # a flip of the Linear's rounding (summation order) moves the sum by one ulp of the Linear's value
    tol = (ref.abs() + ref2.abs()) * 2.0 ** -7 + 2e-3
    assert int((err > tol).sum()) == 0, float(err.max())
    exact = float((o2.float() == ref2).float().mean())
    print(f"[parity] tc_conv linear+residual T={T} N={N} K={K}: bit-identical to the fp32 reference in {exact:.4f} of the outputs")
    assert exact > 0.9


@pytest.mark.parametrize("f16", [1, 0])
def test_tc_conv_16bit_residual_stream_in_place(lib, f16):
    """ResidualUnit's second conv as the decoder blocks run it (modded_dac.py:599-620): 16-bit residual in, raw 16-bit result
    written IN PLACE over the residual, and the consumer's snaked copy; ragged T, N = 96 (16-column blocks) and 192."""
    for N in (96, 192):
        g = torch.Generator().manual_seed(N + f16)
        dt = _dt(f16)
        B, T = 2, 700
        A = torch.randn(B, T, N, generator=g).to(dt).cuda()
        W = (torch.randn(1, N, N, generator=g) / N ** 0.5).to(dt).cuda()
        bias = (torch.randn(N, generator=g) * 0.1).cuda()
        alpha = (0.5 + torch.rand(N, generator=g)).cuda()
        res = torch.randn(B, T, N, generator=g).to(dt).cuda()
        y_ref, z_ref = _ref(A, W, bias=bias, res=res, alpha=alpha)
        from fish_speech_b200._lib import TcConvArgs
        x = res.clone()
        out16 = torch.empty(B, T, N, dtype=dt, device="cuda")
        a = TcConvArgs(A16=A.data_ptr(), W16=W.data_ptr(), B=B, T=T, Cin=N, N=N, taps=1, dil=1, f16=int(f16), use_simt=0,
                       bias=bias.data_ptr(), bias_mod=0, gamma=None, res32=None, res16=x.data_ptr(), ldres=N, alpha=alpha.data_ptr(),
                       alpha_mod=0, act=0, round_lin=0, out16=out16.data_ptr(), ld16=N, out32=None, ld32=N, out16r=x.data_ptr(), ld16r=N)
        assert lib.fsb_tc_conv(C.byref(a), torch.cuda.current_stream().cuda_stream) == 0
        torch.cuda.synchronize()
        _close(x, y_ref, f16, 1.0, f"in-place raw stream N={N}")
        _close(out16, z_ref, f16, 1.0, f"snaked copy N={N}")
