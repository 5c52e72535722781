// DAC codec decode kernels (round-1 correctness path: fp32 activations, channels-last [B][T][C]).
//
// Every conv / transposed conv / linear of the decode path is ONE implicit-GEMM shape
//     out[b][t][n] = epi( bias[n % bias_mod] + sum_j sum_k pro(A[b][t - shift_j][k]) * W[j][k][n] )
// with shift_j = (taps-1-j)*dil (causal left padding = rows t < 0 read as zero):
//   * CausalConvNet            (modded_dac.py:521-552, rvq.py:58-89)   taps = k, N = Cout
//   * CausalTransConvNet k=2s  (modded_dac.py:563-580)                 taps = 2, K = Cin, N = s*Cout: output row q holds
//                               the s phases of samples q*s..q*s+s-1, which IS channels-last [T*s][Cout]
//   * CausalTransConvNet k=s   (rvq.py:100-117)                        taps = 1, N = s*Cout
//   * nn.Linear                                                          taps = 1
// Prologue: Snake1d (descript-audio-codec 1.0.0: x + sin^2(alpha x)/(alpha + 1e-9)) on the A operand.
// Epilogue: bias, exact GELU (rvq.py:180), LayerScale / ConvNeXt gamma, residual add, tanh (modded_dac.py:795).
// This file is the SIMT fp32 tile version used as the parity anchor; the tcgen05 version keeps the same tap/shift plan.
#pragma once
#include "common.cuh"

enum { DEPI_NONE = 0, DEPI_TANH = 1, DEPI_GELU = 2 };

struct ConvArgs {
    const float* A;      // [B][T][Cin]
    const float* W;      // [taps][Cin][N]
    const float* bias;   // [bias_mod] or nullptr
    const float* alpha;  // [Cin] snake prologue or nullptr
    const float* res;    // [B][T][N] or nullptr
    const float* gamma;  // [N] or nullptr
    float* out;          // [B][T][N]
    int B, T, Cin, N, taps, dil, bias_mod, epi;
};

__device__ __forceinline__ float snake_f(float x, float a, float inv) {
    const float s = sinf(a * x);
    return x + inv * s * s;
}

#define CG_BM 64
#define CG_BN 64
#define CG_BK 16

__global__ void __launch_bounds__(256) conv_gemm_kernel(ConvArgs a) {
    __shared__ float As[CG_BK][CG_BM + 4];
    __shared__ float Bs[CG_BK][CG_BN];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int t0 = blockIdx.x * CG_BM, n0 = blockIdx.y * CG_BN, b = blockIdx.z;
    const float* Ab = a.A + (size_t)b * a.T * a.Cin;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    const int am = tid >> 2, ak = (tid & 3) * 4;   // A tile: row am, 4 consecutive k
    const int bk = tid >> 4, bn = (tid & 15) * 4;  // B tile: row bk, 4 consecutive n
    for (int j = 0; j < a.taps; ++j) {
        const int shift = (a.taps - 1 - j) * a.dil;
        const float* Wj = a.W + (size_t)j * a.Cin * a.N;
        for (int k0 = 0; k0 < a.Cin; k0 += CG_BK) {
            float4 av = make_float4(0.f, 0.f, 0.f, 0.f);
            const int t = t0 + am - shift, k = k0 + ak;
            if (t >= 0 && t < a.T && (t0 + am) < a.T && k < a.Cin) {
                av = *reinterpret_cast<const float4*>(Ab + (size_t)t * a.Cin + k);
                if (a.alpha) {
                    const float4 al = *reinterpret_cast<const float4*>(a.alpha + k);
                    av.x = snake_f(av.x, al.x, 1.f / (al.x + 1e-9f));
                    av.y = snake_f(av.y, al.y, 1.f / (al.y + 1e-9f));
                    av.z = snake_f(av.z, al.z, 1.f / (al.z + 1e-9f));
                    av.w = snake_f(av.w, al.w, 1.f / (al.w + 1e-9f));
                }
            }
            float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
            if (k0 + bk < a.Cin && n0 + bn < a.N)
                bv = *reinterpret_cast<const float4*>(Wj + (size_t)(k0 + bk) * a.N + n0 + bn);
            __syncthreads();
            As[ak + 0][am] = av.x; As[ak + 1][am] = av.y; As[ak + 2][am] = av.z; As[ak + 3][am] = av.w;
            *reinterpret_cast<float4*>(&Bs[bk][bn]) = bv;
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < CG_BK; ++kk) {
                const float4 x = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
                const float4 w = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
                const float xs[4] = {x.x, x.y, x.z, x.w}, ws[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj) acc[i][jj] = fmaf(xs[i], ws[jj], acc[i][jj]);
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int t = t0 + ty * 4 + i;
        if (t >= a.T) continue;
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
            const int n = n0 + tx * 4 + jj;
            if (n >= a.N) continue;
            float v = acc[i][jj];
            if (a.bias) v += a.bias[n % a.bias_mod];
            if (a.epi == DEPI_GELU) v = 0.5f * v * (1.f + erff(v * 0.70710678118654752f));
            if (a.gamma) v *= a.gamma[n];
            const size_t o = ((size_t)b * a.T + t) * a.N + n;
            if (a.res) v += a.res[o];
            if (a.epi == DEPI_TANH) v = tanhf(v);
            a.out[o] = v;
        }
    }
}

// ----------------------------------------------------------------------------- tensor-core implicit GEMM (TF32)
// Same op as conv_gemm_kernel on the tensor cores: mma.sync m16n8k8 TF32 (10-bit mantissa operands, fp32 accumulate),
// activations stay fp32 in memory.  CTA tile 128 (time) x 128 (N) x 32 (channels); the A tile is staged ONCE per channel
// chunk with its causal halo ((taps-1)*dil rows), Snake applied once per element, and every tap reads the same
// shared-memory tile at a row offset — the SIMT version re-read and re-Snaked A for every tap.
#define CM_BM 128
#define CM_BN 128                 // widest N tile (NT = 4 n8-tiles per warp); NT = 3 gives 96-wide tiles
#define CM_BK 32
#define CM_HALO 56                // >= (taps - 1) * dil of every layer (7-tap convs with dilation <= 9)
#define CM_AS (CM_BK + 4)         // A row stride in floats: conflict-free fragment loads
#define CM_BS (CM_BN + 8)         // B row stride
#define CM_SMEM (((CM_BM + CM_HALO) * CM_AS + CM_BK * CM_BS) * 4)

__device__ __forceinline__ uint32_t f2tf32(float f) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(f));
    return r;
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

// Snake for the tensor-core path: MUFU sine (abs error ~1e-7 on the O(1) arguments here, far below TF32 rounding)
__device__ __forceinline__ float snake_fast(float x, float a, float inv) {
    const float s = __sinf(a * x);
    return fmaf(inv * s, s, x);
}

// NT: n8-tiles per warp -> CTA N tile = 32 * NT (128 or 96: every decoder width 1536/768/384/192/96 is a multiple of 96)
template <int NT>
__global__ void __launch_bounds__(256, 2) conv_mma_kernel(ConvArgs a) {
    constexpr int BN = 32 * NT;
    extern __shared__ __align__(16) float cm_smem[];
    uint32_t* As = reinterpret_cast<uint32_t*>(cm_smem);                   // [(BM + HALO)][AS] tf32 bits
    uint32_t* Bs = As + (CM_BM + CM_HALO) * CM_AS;                         // [BK][BS]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t4 = lane & 3;
    const int wm = warp & 1, wn = warp >> 1;                               // 2 x 4 warps: 64 rows x 32 columns each
    const int t0 = blockIdx.x * CM_BM, n0 = blockIdx.y * BN, b = blockIdx.z;
    const float* Ab = a.A + (size_t)b * a.T * a.Cin;
    const int halo = (a.taps - 1) * a.dil;
    const int rows = CM_BM + halo;

    float acc[4][NT][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < NT; ++j)
#pragma unroll
            for (int k = 0; k < 4; ++k) acc[i][j][k] = 0.f;

    for (int k0 = 0; k0 < a.Cin; k0 += CM_BK) {
        __syncthreads();  // the previous chunk's MMAs are done with As / Bs
        for (int e = tid; e < rows * (CM_BK / 4); e += 256) {
            const int r = e >> 3, c4 = (e & 7) * 4;
            const int t = t0 - halo + r, k = k0 + c4;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (t >= 0 && t < a.T && k < a.Cin) {
                v = *reinterpret_cast<const float4*>(Ab + (size_t)t * a.Cin + k);
                if (a.alpha) {
                    const float4 al = *reinterpret_cast<const float4*>(a.alpha + k);
                    v.x = snake_fast(v.x, al.x, 1.f / (al.x + 1e-9f));
                    v.y = snake_fast(v.y, al.y, 1.f / (al.y + 1e-9f));
                    v.z = snake_fast(v.z, al.z, 1.f / (al.z + 1e-9f));
                    v.w = snake_fast(v.w, al.w, 1.f / (al.w + 1e-9f));
                }
            }
            *reinterpret_cast<uint4*>(&As[r * CM_AS + c4]) = make_uint4(f2tf32(v.x), f2tf32(v.y), f2tf32(v.z), f2tf32(v.w));
        }
        for (int j = 0; j < a.taps; ++j) {
            const float* Wj = a.W + ((size_t)j * a.Cin + k0) * a.N;
            if (j > 0) __syncthreads();  // MMAs of the previous tap are done with Bs
            for (int e = tid; e < CM_BK * (BN / 4); e += 256) {
                const int kk = e / (BN / 4), c4 = (e % (BN / 4)) * 4;
                float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
                if (k0 + kk < a.Cin && n0 + c4 < a.N) w = *reinterpret_cast<const float4*>(Wj + (size_t)kk * a.N + n0 + c4);
                *reinterpret_cast<uint4*>(&Bs[kk * CM_BS + c4]) = make_uint4(f2tf32(w.x), f2tf32(w.y), f2tf32(w.z), f2tf32(w.w));
            }
            __syncthreads();
            const int roff = halo - (a.taps - 1 - j) * a.dil;  // output row r reads staged row r + roff
#pragma unroll
            for (int ks = 0; ks < CM_BK; ks += 8) {
                uint32_t af[4][4], bf[NT][2];
#pragma unroll
                for (int mt = 0; mt < 4; ++mt) {
                    const uint32_t* p = As + (wm * 64 + mt * 16 + g + roff) * CM_AS + ks + t4;
                    af[mt][0] = p[0]; af[mt][1] = p[8 * CM_AS]; af[mt][2] = p[4]; af[mt][3] = p[8 * CM_AS + 4];
                }
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    const uint32_t* p = Bs + (ks + t4) * CM_BS + wn * (8 * NT) + nt * 8 + g;
                    bf[nt][0] = p[0]; bf[nt][1] = p[4 * CM_BS];
                }
#pragma unroll
                for (int mt = 0; mt < 4; ++mt)
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt) mma_tf32(acc[mt][nt], af[mt], bf[nt]);
            }
        }
    }
#pragma unroll
    for (int mt = 0; mt < 4; ++mt)
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            const int t = t0 + wm * 64 + mt * 16 + g + half * 8;
            if (t >= a.T) continue;
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    const int n = n0 + wn * (8 * NT) + nt * 8 + 2 * t4 + c;
                    if (n >= a.N) continue;
                    float v = acc[mt][nt][half * 2 + c];
                    if (a.bias) v += a.bias[n % a.bias_mod];
                    if (a.epi == DEPI_GELU) v = 0.5f * v * (1.f + erff(v * 0.70710678118654752f));
                    if (a.gamma) v *= a.gamma[n];
                    const size_t o = ((size_t)b * a.T + t) * a.N + n;
                    if (a.res) v += a.res[o];
                    if (a.epi == DEPI_TANH) v = tanhf(v);
                    a.out[o] = v;
                }
        }
}

// Final conv (Cout = 1) + tanh: one warp per output sample (modded_dac.py:793-795)
__global__ void __launch_bounds__(256) conv_out1_kernel(ConvArgs a) {
    const int lane = threadIdx.x & 31;
    const long long gw = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (gw >= (long long)a.B * a.T) return;
    const int b = (int)(gw / a.T), t = (int)(gw % a.T);
    float s = 0.f;
    for (int j = 0; j < a.taps; ++j) {
        const int tt = t - (a.taps - 1 - j) * a.dil;
        if (tt < 0) continue;
        const float* row = a.A + ((size_t)b * a.T + tt) * a.Cin;
        for (int k = lane; k < a.Cin; k += 32) {
            float x = row[k];
            if (a.alpha) { const float al = a.alpha[k]; x = snake_f(x, al, 1.f / (al + 1e-9f)); }
            s = fmaf(x, a.W[(size_t)j * a.Cin + k], s);
        }
    }
    s = warp_sum(s);
    if (lane == 0) {
        s += a.bias ? a.bias[0] : 0.f;
        a.out[(size_t)b * a.T + t] = (a.epi == DEPI_TANH) ? tanhf(s) : s;
    }
}

// ----------------------------------------------------------------------------- RVQ dequantisation (K11)
// rvq.py:352-363 + descript ResidualVectorQuantize.from_codes: clamp (in place, like the reference), per codebook
// gather the 8-d code vector, weight-normed 1x1 out_proj (+bias), sum over the codebooks.
struct RvqArgs {
    long long* codes;        // [B][NQ][T] int64, clamped in place
    const float* codebooks;  // concatenated [sum sizes][CD]
    const float* proj_w;     // [NQ][D][CD] (weight norm folded)
    const float* proj_b;     // [NQ][D]
    float* z;                // [B][T][D]
    int B, T, NQ, D, CD, size0, size_rest;
};
__global__ void __launch_bounds__(256) rvq_dequant_kernel(RvqArgs a) {
    __shared__ float e[16][16];  // [NQ<=16][CD<=16]
    const int t = blockIdx.x, b = blockIdx.y;
    if (threadIdx.x < a.NQ) {
        const int q = threadIdx.x;
        long long* cp = a.codes + ((size_t)b * a.NQ + q) * a.T + t;
        long long c = *cp;
        const long long mx = (q == 0 ? a.size0 : a.size_rest) - 1;
        if (c > mx) { c = mx; *cp = c; }  // torch.clamp(max=...) only (rvq.py:354-359)
        const size_t row = (q == 0) ? (size_t)c : (size_t)a.size0 + (size_t)(q - 1) * a.size_rest + (size_t)c;
        for (int i = 0; i < a.CD; ++i) e[q][i] = a.codebooks[row * a.CD + i];
    }
    __syncthreads();
    for (int d = threadIdx.x; d < a.D; d += blockDim.x) {
        float s = 0.f;
        for (int q = 0; q < a.NQ; ++q) {
            const float* w = a.proj_w + ((size_t)q * a.D + d) * a.CD;
            float p = 0.f;
            for (int i = 0; i < a.CD; ++i) p = fmaf(w[i], e[q][i], p);
            s += p + a.proj_b[(size_t)q * a.D + d];
        }
        a.z[((size_t)b * a.T + t) * a.D + d] = s;
    }
}

// ----------------------------------------------------------------------------- row-wise norms
// RMSNorm (modded_dac.py:320-331), one warp per row
__global__ void __launch_bounds__(256) rmsnorm_rows_kernel(const float* x, const float* w, float* y, long long rows, int C,
                                                          float eps) {
    const int lane = threadIdx.x & 31;
    const long long r = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (r >= rows) return;
    const float* xr = x + r * C;
    float ss = 0.f;
    for (int c = lane; c < C; c += 32) ss += xr[c] * xr[c];
    ss = warp_sum(ss);
    const float rs = rsqrtf(ss / C + eps);
    for (int c = lane; c < C; c += 32) y[r * C + c] = xr[c] * rs * w[c];
}

// ConvNeXt front half (rvq.py:173-179): depthwise causal conv k=7 (+bias) then LayerNorm(eps 1e-6) over channels.
// One CTA per (b, t) row.
__global__ void __launch_bounds__(256) dwconv_ln_kernel(const float* x, const float* dw_w, const float* dw_b,
                                                       const float* ln_w, const float* ln_b, float* y, int T, int C) {
    extern __shared__ float v[];  // [C]
    __shared__ float red[32];
    const int t = blockIdx.x, b = blockIdx.y;
    float s = 0.f;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float acc = dw_b[c];
#pragma unroll
        for (int j = 0; j < 7; ++j) {
            const int tt = t - 6 + j;
            if (tt >= 0) acc = fmaf(x[((size_t)b * T + tt) * C + c], dw_w[c * 7 + j], acc);
        }
        v[c] = acc;
        s += acc;
    }
    const float mean = block_sum(s, red) / C;
    float q = 0.f;
    for (int c = threadIdx.x; c < C; c += blockDim.x) { const float d = v[c] - mean; q += d * d; }
    const float var = block_sum(q, red) / C;
    const float rs = rsqrtf(var + 1e-6f);
    for (int c = threadIdx.x; c < C; c += blockDim.x)
        y[((size_t)b * T + t) * C + c] = (v[c] - mean) * rs * ln_w[c] + ln_b[c];
}

// SwiGLU on a fused [rows][2F] gate|up GEMM output (modded_dac.py FeedForward: w2(silu(w1 x) * w3 x))
__global__ void swiglu_kernel(const float* gu, float* out, long long rows, int F) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * F) return;
    const long long r = i / F;
    const int f = (int)(i % F);
    const float g = gu[r * 2 * F + f], u = gu[r * 2 * F + F + f];
    out[i] = g / (1.f + expf(-g)) * u;
}

// ----------------------------------------------------------------------------- window attention (K12)
// RoPE in place on the q and k thirds of qkv [B][T][3][H][HD] (modded_dac.py:455-467, table bf16-rounded)
// `pos0` = position of row 0 (a clip decoded in overlapping pieces keeps its absolute positions)
__global__ void rope_qk_kernel(float* qkv, const float* table, int B, int T, int H, int HD, int pos0) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int half = HD / 2;
    const long long total = (long long)B * T * 2 * H * half;
    if (i >= total) return;
    const int p = (int)(i % half);
    const int h = (int)((i / half) % H);
    const int which = (int)((i / ((long long)half * H)) % 2);
    const long long bt = i / ((long long)half * H * 2);
    const int t = (int)(bt % T) + pos0;
    float* v = qkv + (bt * 3 + which) * H * HD + (size_t)h * HD + 2 * p;
    const float c = table[((size_t)t * half + p) * 2], s = table[((size_t)t * half + p) * 2 + 1];
    const float x0 = v[0], x1 = v[1];
    v[0] = x0 * c - x1 * s;
    v[1] = x1 * c + x0 * s;
}

// causal attention limited to the last `window` keys (WindowLimitedTransformer mask, modded_dac.py:380-398);
// one warp per (b, h, t); lane scores keys start+lane+32i, then owns output dims lane*EPL..
// MAXI: rounds of 32 keys a lane scores, window <= 32 * MAXI (4 for the quantizer's transformers, 16 for the encoder's 512)
template <int HD, int MAXI = 4>
__global__ void __launch_bounds__(128) window_attn_kernel(const float* qkv, float* y, int B, int T, int H, int window,
                                                         float scale) {
    constexpr int EPL = HD / 32;
    __shared__ float qs[4][HD];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int t = blockIdx.x * 4 + w, h = blockIdx.y, b = blockIdx.z;
    if (t >= T) return;
    const size_t rs = (size_t)3 * H * HD;
    const float* base = qkv + (size_t)b * T * rs;
#pragma unroll
    for (int i = 0; i < EPL; ++i) qs[w][lane * EPL + i] = base[(size_t)t * rs + (size_t)h * HD + lane * EPL + i];
    __syncwarp();
    const int start = max(0, t - window + 1), n = t - start + 1;
    float sc[MAXI], mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < MAXI; ++i) {
        const int j = lane + 32 * i;
        sc[i] = -INFINITY;
        if (j < n) {
            const float* k = base + (size_t)(start + j) * rs + (size_t)(H + h) * HD;
            float d = 0.f;
#pragma unroll 8
            for (int e = 0; e < HD; ++e) d = fmaf(qs[w][e], k[e], d);
            sc[i] = d * scale;
            mx = fmaxf(mx, sc[i]);
        }
    }
    mx = warp_max(mx);
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < MAXI; ++i) { sc[i] = (sc[i] == -INFINITY) ? 0.f : expf(sc[i] - mx); sum += sc[i]; }
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    float o[EPL];
#pragma unroll
    for (int i = 0; i < EPL; ++i) o[i] = 0.f;
#pragma unroll
    for (int i = 0; i < MAXI; ++i) {
        for (int src = 0; src < 32; ++src) {
            const int j = src + 32 * i;
            if (j >= n) break;  // uniform across the warp
            const float p = __shfl_sync(0xffffffffu, sc[i], src);
            const float* v = base + (size_t)(start + j) * rs + (size_t)(2 * H + h) * HD + lane * EPL;
#pragma unroll
            for (int e = 0; e < EPL; ++e) o[e] = fmaf(p, v[e], o[e]);
        }
    }
#pragma unroll
    for (int e = 0; e < EPL; ++e) y[((size_t)b * T + t) * H * HD + (size_t)h * HD + lane * EPL + e] = o[e] * inv;
}

// ============================================================================= encode side (DAC.encode, modded_dac.py:874-923)
// The encoder's first conv (modded_dac.py:683): 1 -> N channels, `taps` taps, causal.  One thread per output element.
__global__ void __launch_bounds__(256) conv_in1_kernel(const float* x, const float* W, const float* bias, float* out, int B,
                                                      long long T, int N, int taps) {
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    if (i >= (long long)B * T * N) return;
    const int n = (int)(i % N);
    const long long bt = i / N, t = bt % T, b = bt / T;
    float acc = 0.f;
    for (int j = 0; j < taps; ++j) {
        const long long tt = t - (taps - 1 - j);
        if (tt >= 0) acc = fmaf(x[b * T + tt], W[j * N + n], acc);
    }
    out[i] = acc + bias[n];
}

// The codebooks (rvq.py:300-312 over descript VectorQuantize.forward / decode_latents): for one latent row, codebook after
// codebook on what the previous ones left: e = in_proj(r); nearest entry between L2-normalised vectors (distance written as
// the reference writes it: |e|^2 - 2 e.c + |c|^2, first index on ties); u = e + (entry - e) (the straight-through sum, it
// rounds); r -= out_proj(u).  One CTA per (frame, clip); the row never leaves shared memory.
struct VqEncArgs {
    const float* z;       // [B][T][D]
    const float* in_w;    // [NQ][CD][D]   weight-normed
    const float* in_b;    // [NQ][CD]
    const float* cb;      // [size0 + (NQ-1)*size_rest][CD]   entries as stored
    const float* cbn;     // same shape, rows L2-normalised (F.normalize)
    const float* cbsq;    // [..]  |normalised row|^2 (fp32: not exactly 1)
    const float* out_w;   // [NQ][D][CD]   weight-normed
    const float* out_b;   // [NQ][D]
    long long* codes;     // [B][NQ][T]
    int B, T, NQ, D, CD, size0, size_rest;
};

__global__ void __launch_bounds__(256) vq_encode_kernel(VqEncArgs a) {
    extern __shared__ float vq_r[];            // [D] residual
    __shared__ float e_s[16], en_s[16], u_s[16];
    __shared__ float best_v[8];
    __shared__ int best_i[8];
    const int t = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float* zr = a.z + ((size_t)b * a.T + t) * a.D;
    for (int c = tid; c < a.D; c += 256) vq_r[c] = zr[c];
    __syncthreads();
    int off = 0;
    for (int q = 0; q < a.NQ; ++q) {
        const int size = q == 0 ? a.size0 : a.size_rest;
        // ---- e = in_proj(r): one warp per codebook dimension
        for (int k = warp; k < a.CD; k += 8) {
            const float* w = a.in_w + ((size_t)q * a.CD + k) * a.D;
            float acc = 0.f;
            for (int c = lane; c < a.D; c += 32) acc = fmaf(w[c], vq_r[c], acc);
            acc = warp_sum(acc);
            if (lane == 0) e_s[k] = acc + a.in_b[q * a.CD + k];
        }
        __syncthreads();
        if (tid == 0) {                        // F.normalize: e / max(|e|, 1e-12)
            float ss = 0.f;
            for (int k = 0; k < a.CD; ++k) ss += e_s[k] * e_s[k];
            const float nrm = fmaxf(sqrtf(ss), 1e-12f);
            for (int k = 0; k < a.CD; ++k) en_s[k] = e_s[k] / nrm;
        }
        __syncthreads();
        float en[16], esq = 0.f;
#pragma unroll
        for (int k = 0; k < 16; ++k) { en[k] = k < a.CD ? en_s[k] : 0.f; if (k < a.CD) esq += en[k] * en[k]; }
        // ---- nearest entry: the largest -dist, the smallest index among equals
        float bv = -INFINITY;
        int bi = 0x7fffffff;
        for (int j = tid; j < size; j += 256) {
            const float* c = a.cbn + (size_t)(off + j) * a.CD;
            float m = 0.f;
            for (int k = 0; k < a.CD; ++k) m = fmaf(en[k], c[k], m);
            const float v = -((esq - 2.f * m) + a.cbsq[off + j]);
            if (v > bv) { bv = v; bi = j; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (lane == 0) { best_v[warp] = bv; best_i[warp] = bi; }
        __syncthreads();
        if (tid == 0) {
            float v = best_v[0];
            int i = best_i[0];
            for (int w = 1; w < 8; ++w)
                if (best_v[w] > v || (best_v[w] == v && best_i[w] < i)) { v = best_v[w]; i = best_i[w]; }
            a.codes[((size_t)b * a.NQ + q) * a.T + t] = i;
            const float* c = a.cb + (size_t)(off + i) * a.CD;
            for (int k = 0; k < a.CD; ++k) u_s[k] = e_s[k] + (c[k] - e_s[k]);
        }
        __syncthreads();
        // ---- r -= out_proj(u)
        for (int c = tid; c < a.D; c += 256) {
            const float* w = a.out_w + ((size_t)q * a.D + c) * a.CD;
            float acc = 0.f;
            for (int k = 0; k < a.CD; ++k) acc = fmaf(w[k], u_s[k], acc);
            vq_r[c] -= acc + a.out_b[q * a.D + c];
        }
        __syncthreads();
        off += size;
    }
}

// ============================================================================= tcgen05 pipeline: the non-GEMM pieces
// Data flow of the tensor-core decode path (csrc/tc_conv.cuh): GEMM operands are 16-bit (fp16 or bf16) channels-last
// tensors written by their producer; the residual stream stays fp32.  These kernels are the fp32 -> 16-bit producers
// that are not GEMM epilogues.  H = Tcv16<F16> (tc_conv.cuh) converts.
#include "tc_conv.cuh"

// RMSNorm (modded_dac.py:320-331) of fp32 rows -> 16-bit rows, one warp per row, float4 loads
template <bool F16>
__global__ void __launch_bounds__(256) rmsnorm_rows16_kernel(const float* x, const float* w, void* y16, long long rows, int C, float eps) {
    typedef Tcv16<F16> H;
    const int lane = threadIdx.x & 31;
    const long long r = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (r >= rows) return;
    const float* xr = x + r * C;
    uint16_t* yr = reinterpret_cast<uint16_t*>(y16) + r * C;
    float ss = 0.f;
    if ((C & 3) == 0) {
        for (int c = lane * 4; c < C; c += 128) { const float4 v = *reinterpret_cast<const float4*>(xr + c); ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w; }
    } else {
        for (int c = lane; c < C; c += 32) ss += xr[c] * xr[c];
    }
    ss = warp_sum(ss);
    const float rs = rsqrtf(ss / C + eps);
    if ((C & 3) == 0) {
        for (int c = lane * 4; c < C; c += 128) {
            const float4 v = *reinterpret_cast<const float4*>(xr + c), g = __ldg(reinterpret_cast<const float4*>(w + c));
            const uint32_t lo = (uint32_t)H::pack(v.x * rs * g.x) | ((uint32_t)H::pack(v.y * rs * g.y) << 16);
            const uint32_t hi = (uint32_t)H::pack(v.z * rs * g.z) | ((uint32_t)H::pack(v.w * rs * g.w) << 16);
            *reinterpret_cast<uint2*>(yr + c) = make_uint2(lo, hi);
        }
    } else {
        for (int c = lane; c < C; c += 32) yr[c] = H::pack(xr[c] * rs * w[c]);
    }
}

// ConvNeXt front half (rvq.py:173-179) fp32 in -> 16-bit out; one CTA per (b, t) row
template <bool F16>
__global__ void __launch_bounds__(256) dwconv_ln16_kernel(const float* x, const float* dw_w, const float* dw_b, const float* ln_w,
                                                          const float* ln_b, void* y16, int T, int C) {
    typedef Tcv16<F16> H;
    extern __shared__ float v[];  // [C]
    __shared__ float red[32];
    const int t = blockIdx.x, b = blockIdx.y;
    float s = 0.f;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float acc = dw_b[c];
#pragma unroll
        for (int j = 0; j < 7; ++j) {
            const int tt = t - 6 + j;
            if (tt >= 0) acc = fmaf(x[((size_t)b * T + tt) * C + c], dw_w[c * 7 + j], acc);
        }
        v[c] = acc;
        s += acc;
    }
    const float mean = block_sum(s, red) / C;
    float q = 0.f;
    for (int c = threadIdx.x; c < C; c += blockDim.x) { const float d = v[c] - mean; q += d * d; }
    const float var = block_sum(q, red) / C;
    const float rs = rsqrtf(var + 1e-6f);
    uint16_t* yr = reinterpret_cast<uint16_t*>(y16) + ((size_t)b * T + t) * C;
    for (int c = threadIdx.x; c < C; c += blockDim.x) yr[c] = H::pack((v[c] - mean) * rs * ln_w[c] + ln_b[c]);
}

// window attention as window_attn_kernel, output rows in the 16-bit operand type (the wo GEMM reads them)
template <int HD, bool F16>
__global__ void __launch_bounds__(128) window_attn16_kernel(const float* qkv, void* y16, int B, int T, int H, int window, float scale) {
    typedef Tcv16<F16> Hh;
    constexpr int EPL = HD / 32;
    constexpr int MAXI = 4;  // window <= 128
    __shared__ float qs[4][HD];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int t = blockIdx.x * 4 + w, h = blockIdx.y, b = blockIdx.z;
    if (t >= T) return;
    const size_t rs = (size_t)3 * H * HD;
    const float* base = qkv + (size_t)b * T * rs;
#pragma unroll
    for (int i = 0; i < EPL; ++i) qs[w][lane * EPL + i] = base[(size_t)t * rs + (size_t)h * HD + lane * EPL + i];
    __syncwarp();
    const int start = max(0, t - window + 1), n = t - start + 1;
    float sc[MAXI], mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < MAXI; ++i) {
        const int j = lane + 32 * i;
        sc[i] = -INFINITY;
        if (j < n) {
            const float4* k = reinterpret_cast<const float4*>(base + (size_t)(start + j) * rs + (size_t)(H + h) * HD);
            float d = 0.f;
#pragma unroll
            for (int e = 0; e < HD / 4; ++e) {
                const float4 kv = k[e];
                d = fmaf(qs[w][4 * e], kv.x, d); d = fmaf(qs[w][4 * e + 1], kv.y, d);
                d = fmaf(qs[w][4 * e + 2], kv.z, d); d = fmaf(qs[w][4 * e + 3], kv.w, d);
            }
            sc[i] = d * scale;
            mx = fmaxf(mx, sc[i]);
        }
    }
    mx = warp_max(mx);
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < MAXI; ++i) { sc[i] = (sc[i] == -INFINITY) ? 0.f : expf(sc[i] - mx); sum += sc[i]; }
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    float o[EPL];
#pragma unroll
    for (int i = 0; i < EPL; ++i) o[i] = 0.f;
#pragma unroll
    for (int i = 0; i < MAXI; ++i) {
        for (int src = 0; src < 32; ++src) {
            const int j = src + 32 * i;
            if (j >= n) break;  // uniform across the warp
            const float p = __shfl_sync(0xffffffffu, sc[i], src);
            const float* v = base + (size_t)(start + j) * rs + (size_t)(2 * H + h) * HD + lane * EPL;
#pragma unroll
            for (int e = 0; e < EPL; ++e) o[e] = fmaf(p, v[e], o[e]);
        }
    }
    uint16_t* yr = reinterpret_cast<uint16_t*>(y16) + ((size_t)b * T + t) * H * HD + (size_t)h * HD + lane * EPL;
#pragma unroll
    for (int e = 0; e < EPL; ++e) yr[e] = Hh::pack(o[e] * inv);
}

// Final conv (Cout = 1) + tanh over the already-snaked 16-bit input (modded_dac.py:793-795): one warp per output sample
template <bool F16>
__global__ void __launch_bounds__(256) conv_out1_16_kernel(const void* A16, const float* W, const float* bias, float* out, int B, int T,
                                                           int Cin, int taps) {
    typedef Tcv16<F16> H;
    const int lane = threadIdx.x & 31;
    const long long gw = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (gw >= (long long)B * T) return;
    const int b = (int)(gw / T), t = (int)(gw % T);
    const uint16_t* A = reinterpret_cast<const uint16_t*>(A16);
    float s = 0.f;
    for (int j = 0; j < taps; ++j) {
        const int tt = t - (taps - 1 - j);
        if (tt < 0) continue;
        const uint16_t* row = A + ((size_t)b * T + tt) * Cin;
        for (int k = lane; k < Cin; k += 32) s = fmaf(H::unpack(row[k]), __ldg(W + (size_t)j * Cin + k), s);
    }
    s = warp_sum(s);
    if (lane == 0) out[(size_t)b * T + t] = tanhf(s + (bias ? bias[0] : 0.f));
}

// ----------------------------------------------------------------------------- tiled forms of the two kernels above
// window_attn16_kernel reads every key and value row of a query's window from L2 once per query: 128 x 2 x 256 B per (b, h, t),
// 5.4 GB per layer at 8 clips x 646 frames - 0.8 ms per layer for 63 MB of q/k/v.  Here a CTA owns WA_QT consecutive queries
// of one (clip, head): the keys and values of their windows (WA_QT + window - 1 rows) are staged in shared memory once, and
// every warp walks its queries with the SAME arithmetic, in the same order, as window_attn16_kernel (bit-identical output).
#define WA_QT 64
#define WA_THREADS 256
template <int HD>
__host__ __device__ constexpr size_t wa_smem_bytes(int window) {
    return (size_t)(WA_QT + window - 1) * ((HD + 4) + HD) * sizeof(float) + (WA_THREADS / 32) * HD * sizeof(float);
}
template <int HD, bool F16>
__global__ void __launch_bounds__(WA_THREADS) window_attn16_tiled_kernel(const float* qkv, void* y16, int B, int T, int H, int window,
                                                                         float scale) {
    typedef Tcv16<F16> Hh;
    constexpr int EPL = HD / 32;
    constexpr int MAXI = 4;      // window <= 128
    constexpr int KS = HD + 4;   // key row stride in floats: lanes read different rows with LDS.128, 4 j mod 32 spreads them
    extern __shared__ __align__(16) float wa_smem[];
    const int nrows = WA_QT + window - 1;
    float* ks = wa_smem;
    float* vs = ks + (size_t)nrows * KS;
    float* qs = vs + (size_t)nrows * HD;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int t0 = blockIdx.x * WA_QT, h = blockIdx.y, b = blockIdx.z;
    const size_t rs = (size_t)3 * H * HD;
    const float* base = qkv + (size_t)b * T * rs;
    const int first = t0 - window + 1;                       // key position of shared-memory row 0 (may be negative)
    const int t_end = min(T, t0 + WA_QT);
    for (int i = threadIdx.x; i < nrows * (HD / 4); i += WA_THREADS) {
        const int r = i / (HD / 4), c = i % (HD / 4);
        const int kp = first + r;
        if (kp >= 0 && kp < t_end) {
            const float* row = base + (size_t)kp * rs + (size_t)h * HD;
            *reinterpret_cast<float4*>(ks + (size_t)r * KS + 4 * c) = *reinterpret_cast<const float4*>(row + (size_t)H * HD + 4 * c);
            *reinterpret_cast<float4*>(vs + (size_t)r * HD + 4 * c) = *reinterpret_cast<const float4*>(row + (size_t)2 * H * HD + 4 * c);
        }
    }
    __syncthreads();
    float* q = qs + w * HD;
    for (int t = t0 + w; t < t_end; t += WA_THREADS / 32) {
#pragma unroll
        for (int i = 0; i < EPL; ++i) q[lane * EPL + i] = base[(size_t)t * rs + (size_t)h * HD + lane * EPL + i];
        __syncwarp();
        const int start = max(0, t - window + 1), n = t - start + 1;
        const int r0 = start - first;                        // shared-memory row of the window's first key
        float sc[MAXI], mx = -INFINITY;
#pragma unroll
        for (int i = 0; i < MAXI; ++i) {
            const int j = lane + 32 * i;
            sc[i] = -INFINITY;
            if (j < n) {
                const float4* k = reinterpret_cast<const float4*>(ks + (size_t)(r0 + j) * KS);
                float d = 0.f;
#pragma unroll
                for (int e = 0; e < HD / 4; ++e) {
                    const float4 kv = k[e];
                    d = fmaf(q[4 * e], kv.x, d); d = fmaf(q[4 * e + 1], kv.y, d);
                    d = fmaf(q[4 * e + 2], kv.z, d); d = fmaf(q[4 * e + 3], kv.w, d);
                }
                sc[i] = d * scale;
                mx = fmaxf(mx, sc[i]);
            }
        }
        mx = warp_max(mx);
        float sum = 0.f;
#pragma unroll
        for (int i = 0; i < MAXI; ++i) { sc[i] = (sc[i] == -INFINITY) ? 0.f : expf(sc[i] - mx); sum += sc[i]; }
        sum = warp_sum(sum);
        const float inv = 1.f / sum;
        float o[EPL];
#pragma unroll
        for (int i = 0; i < EPL; ++i) o[i] = 0.f;
#pragma unroll
        for (int i = 0; i < MAXI; ++i) {
            for (int src = 0; src < 32; ++src) {
                const int j = src + 32 * i;
                if (j >= n) break;  // uniform across the warp
                const float p = __shfl_sync(0xffffffffu, sc[i], src);
                const float* v = vs + (size_t)(r0 + j) * HD + lane * EPL;
#pragma unroll
                for (int e = 0; e < EPL; ++e) o[e] = fmaf(p, v[e], o[e]);
            }
        }
        uint16_t* yr = reinterpret_cast<uint16_t*>(y16) + ((size_t)b * T + t) * H * HD + (size_t)h * HD + lane * EPL;
#pragma unroll
        for (int e = 0; e < EPL; ++e) yr[e] = Hh::pack(o[e] * inv);
        __syncwarp();
    }
}

// Final conv (Cout = 1) + tanh, tiled: conv_out1_16_kernel runs one WARP per output sample and re-reads each input row `taps`
// times from L2 (6.4 ms per 8 clips for 2 GB of input).  Here a CTA stages CO_TILE + taps - 1 rows once and one THREAD computes
// one sample; rows are padded to an odd number of 16-byte units so that neighbouring threads' LDS.128 do not collide.
#define CO_TILE 256
__host__ __device__ inline int co_row_units(int Cin) { const int u = Cin / 8; return (u & 1) ? u : u + 1; }
__host__ __device__ inline size_t co_smem_bytes(int Cin, int taps) {
    return (size_t)(CO_TILE + taps - 1) * co_row_units(Cin) * 16 + (size_t)taps * Cin * sizeof(float);
}
template <bool F16>
__global__ void __launch_bounds__(CO_TILE) conv_out1_16_tiled_kernel(const void* A16, const float* W, const float* bias, float* out,
                                                                    int B, int T, int Cin, int taps) {
    typedef Tcv16<F16> H;
    extern __shared__ __align__(16) unsigned char co_smem[];
    const int units = Cin / 8, ru = co_row_units(Cin);
    const int nrows = CO_TILE + taps - 1;
    uint4* rows = reinterpret_cast<uint4*>(co_smem);
    float* ws = reinterpret_cast<float*>(co_smem + (size_t)nrows * ru * 16);
    const int tiles = (T + CO_TILE - 1) / CO_TILE;
    const int b = blockIdx.x / tiles, t0 = (blockIdx.x % tiles) * CO_TILE;
    const uint4* A = reinterpret_cast<const uint4*>(A16) + (size_t)b * T * units;
    for (int i = threadIdx.x; i < nrows * units; i += CO_TILE) {
        const int r = i / units, c = i % units;
        const int tt = t0 - (taps - 1) + r;
        rows[(size_t)r * ru + c] = (tt >= 0 && tt < T) ? A[(size_t)tt * units + c] : make_uint4(0, 0, 0, 0);  // zeros = causal padding
    }
    for (int i = threadIdx.x; i < taps * Cin; i += CO_TILE) ws[i] = W[i];
    __syncthreads();
    const int t = t0 + threadIdx.x;
    if (t >= T) return;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int j = 0; j < taps; ++j) {
        const uint4* row = rows + (size_t)(threadIdx.x + j) * ru;
        const float* wj = ws + (size_t)j * Cin;
        for (int c = 0; c < units; ++c) {
            const uint4 v = row[c];
            const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                acc[e] = fmaf(H::unpack((uint16_t)(u[e] & 0xffffu)), wj[8 * c + 2 * e], acc[e]);
                acc[e] = fmaf(H::unpack((uint16_t)(u[e] >> 16)), wj[8 * c + 2 * e + 1], acc[e]);
            }
        }
    }
    out[(size_t)b * T + t] = tanhf((acc[0] + acc[1]) + (acc[2] + acc[3]) + (bias ? bias[0] : 0.f));
}
