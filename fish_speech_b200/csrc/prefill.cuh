// Prompt pass (Tq > 1 of BaseTransformer.forward_generate, llama.py:390-466) over whole chunks of prompt rows.
//
// The prompt pass is GEMM-shaped (T rows x every Linear of the slow stack); its plain GEMMs go to cuBLAS (bf16 operands,
// fp32 accumulation, bf16 output = one rounding per Linear like the reference), and everything around them — RMSNorm with
// the reference's double rounding, residual adds, SwiGLU, q/k-norm + RoPE + KV scatter, causal attention against the
// cache — stays in this library's own kernels.  (The decode path never calls a library; a tcgen05 GEMM for the prompt
// pass is a "next" row in DESIGN.md.)
#pragma once
#include "common.cuh"

#define PREFILL_KS_ROWS 4096  // rows of fp32 split-K partial sums (pieces x chunk rows) the prompt pass may hold
#define PREFILL_ROWS 2048  // prompt rows per chunk: bounds the activation scratch (about 100 KB per row at the S2 shape)

// out[t][:] = bf16(bf16(x * rsqrt(mean(x^2) + eps)) * w)      (llama.py:995-1000)
__global__ void __launch_bounds__(256) rmsnorm_rows_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w,
                                                           bf16* __restrict__ out, int K, float eps) {
    __shared__ float red[32];
    const int t = blockIdx.x, K8 = K >> 3;
    const uint4* xr = reinterpret_cast<const uint4*>(x + (size_t)t * K);
    float ss = 0.f;
    for (int c = threadIdx.x; c < K8; c += blockDim.x) {
        float f[8];
        unpack8(xr[c], f);
#pragma unroll
        for (int i = 0; i < 8; ++i) ss += f[i] * f[i];
    }
    ss = block_sum(ss, red);
    const float rs = __frsqrt_rn(__fadd_rn(__fdiv_rn(ss, (float)K), eps));
    const uint4* wr = reinterpret_cast<const uint4*>(w);
    uint4* o = reinterpret_cast<uint4*>(out + (size_t)t * K);
    for (int c = threadIdx.x; c < K8; c += blockDim.x) {
        float f[8], g[8];
        unpack8(xr[c], f);
        unpack8(wr[c], g);
#pragma unroll
        for (int i = 0; i < 8; ++i) f[i] = rbf(__fmul_rn(f[i], rs)) * g[i];
        uint4 v;
        v.x = pack2(f[0], f[1]); v.y = pack2(f[2], f[3]); v.z = pack2(f[4], f[5]); v.w = pack2(f[6], f[7]);
        o[c] = v;
    }
}

// out = bf16(res + lin)   (h = x + attention(...), out = h + feed_forward(...): llama.py:841-842); lin is a bf16 Linear output
__global__ void __launch_bounds__(256) add_rows_kernel(const bf16* __restrict__ res, const bf16* __restrict__ lin,
                                                       bf16* __restrict__ out, size_t n8) {
    for (size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x; c < n8; c += (size_t)gridDim.x * blockDim.x) {
        float a[8], b[8];
        unpack8(reinterpret_cast<const uint4*>(res)[c], a);
        unpack8(reinterpret_cast<const uint4*>(lin)[c], b);
        uint4 v;
        v.x = pack2(a[0] + b[0], a[1] + b[1]); v.y = pack2(a[2] + b[2], a[3] + b[3]);
        v.z = pack2(a[4] + b[4], a[5] + b[5]); v.w = pack2(a[6] + b[6], a[7] + b[7]);
        reinterpret_cast<uint4*>(out)[c] = v;
    }
}

// out = bf16(bf16(silu(a)) * b), silu(a) = a / (1 + exp(-a))   (llama.py:986)
__global__ void __launch_bounds__(256) swiglu_rows_kernel(const bf16* a, const bf16* b, bf16* out, size_t n8) {  // out may alias a
    for (size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x; c < n8; c += (size_t)gridDim.x * blockDim.x) {
        float x[8], y[8];
        unpack8(reinterpret_cast<const uint4*>(a)[c], x);
        unpack8(reinterpret_cast<const uint4*>(b)[c], y);
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = rbf(__fdiv_rn(x[i], 1.f + expf(-x[i]))) * y[i];
        uint4 v;
        v.x = pack2(x[0], x[1]); v.y = pack2(x[2], x[3]); v.z = pack2(x[4], x[5]); v.w = pack2(x[6], x[7]);
        reinterpret_cast<uint4*>(out)[c] = v;
    }
}
