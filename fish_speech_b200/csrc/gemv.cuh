// Weight-streaming GEMV for the batch-1 / few-row decode regime (HBM-bound).
//
//   out[m][n] = epilogue( sum_k W[n][k] * prologue(x)[m][k] )          m < M (<= 8 rows), n < N
//
// Replaces the reference's F.linear call sites on the decode path
// (llama.py:894 wqkv, :945 wo, :986 w1/w3/w2, :450-457 LM head, :816 fast_output) together with the
// RMSNorm in front of them (llama.py:995-1000) and the SiLU*mul / residual add behind them
// (llama.py:986, :841-842).  Rounding points follow the reference op by op: every Linear output is
// rounded to bf16 before the next elementwise op, each elementwise op rounds to bf16 again.
//
// Design (bandwidth-bound, no tensor cores on purpose):
//   * W is [N][K] row-major bf16, so a CTA's share of rows is ONE contiguous byte range.  Rows are split
//     evenly over the grid (grid = 2 CTAs x 148 SMs) so every CTA streams the same number of bytes.
//   * 128-bit ld.global.nc.L1::no_allocate loads, G rows in flight per thread per step.
//   * The slab is prefetched to L2 with cp.async.bulk.prefetch.L2 BEFORE griddepcontrol.wait: weights
//     do not depend on the previous kernel, so HBM stays busy across the kernel boundary (PDL).
//   * x (M rows) is staged once per CTA in shared memory as bf16 (after the fused RMSNorm prologue).
//   * fp32 accumulation, warp-shuffle + shared-memory reduction, fused epilogue.
#pragma once
#include "common.cuh"

enum { PRO_NONE = 0, PRO_RMSNORM = 1 };
enum { EPI_STORE = 0, EPI_RESIDUAL = 1, EPI_SWIGLU = 2 };

struct GemvArgs {
    const bf16* W;        // [N][K]
    const bf16* W2;       // [N][K] second matrix (EPI_SWIGLU: W = w1 gate, W2 = w3 up)
    const bf16* x;        // [M][K]
    const bf16* norm_w;   // [K]   (PRO_RMSNORM)
    const bf16* res;      // [M][N] residual (EPI_RESIDUAL)
    bf16* out;            // [M][N]
    bf16* xnorm_out;      // optional [M][K]: prologue output written by CTA 0 (hidden states)
    int N, K, M_valid;
    float eps;
    const void* pf_ptr;   // optional prefetch-ahead region (weights of a LATER kernel), shared over CTAs
    size_t pf_bytes;
    size_t self_pf_bytes; // cap on the per-CTA prefetch of its own slab
};

#define GEMV_THREADS 256

template <int M, int G, int PRO, int EPI>
__global__ void __launch_bounds__(GEMV_THREADS, (M <= 2 ? 3 : 1))
gemv_kernel(GemvArgs a) {
    extern __shared__ uint4 xs[];  // [M][K/8] bf16x8
    constexpr int NW = GEMV_THREADS / 32;
    constexpr int NACC = (EPI == EPI_SWIGLU) ? 2 : 1;
    __shared__ float red[32];
    __shared__ float part[NW][NACC * G * M];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nb = gridDim.x, b = blockIdx.x;
    const int K = a.K, KC = K >> 3;
    const int r0 = (int)(((long long)a.N * b) / nb);
    const int r1 = (int)(((long long)a.N * (b + 1)) / nb);
    const int nrows = r1 - r0;

    // ---- pre-dependency: pull this CTA's weight slab towards L2 (immutable data only)
    if (tid == 0 && nrows > 0) {
        size_t slab = (size_t)nrows * K * sizeof(bf16);
        size_t n = slab < a.self_pf_bytes ? slab : a.self_pf_bytes;
        n &= ~(size_t)15;
        const char* p = (const char*)(a.W + (size_t)r0 * K);
        const char* p2 = (EPI == EPI_SWIGLU) ? (const char*)(a.W2 + (size_t)r0 * K) : nullptr;
        for (size_t o = 0; o < n; o += 32768) {
            uint32_t c = (uint32_t)((n - o) < 32768 ? (n - o) : 32768);
            l2_prefetch_bulk(p + o, c);
            if (EPI == EPI_SWIGLU) l2_prefetch_bulk(p2 + o, c);
        }
        l2_prefetch_share(a.pf_ptr, a.pf_bytes, b, nb);
    }
    pdl_wait();
    pdl_launch_dependents();

    // ---- stage x (optionally RMSNorm'd) into shared memory as bf16
    for (int m = 0; m < M; ++m) {
        const uint4* xg = reinterpret_cast<const uint4*>(a.x + (size_t)m * K);
        const bool valid = m < a.M_valid;
        if (PRO == PRO_RMSNORM) {
            float ss = 0.f;
            if (valid)
                for (int c = tid; c < KC; c += GEMV_THREADS) {
                    float f[8];
                    unpack8(xg[c], f);
#pragma unroll
                    for (int i = 0; i < 8; ++i) ss += f[i] * f[i];
                }
            ss = block_sum(ss, red);
            // llama.py:995-1000: fp32 normalise -> round to bf16 -> * weight (bf16 mul, second rounding)
            const float rs = __frsqrt_rn(__fadd_rn(__fdiv_rn(ss, (float)K), a.eps));
            const uint4* wg = reinterpret_cast<const uint4*>(a.norm_w);
            for (int c = tid; c < KC; c += GEMV_THREADS) {
                uint4 o = make_uint4(0, 0, 0, 0);
                if (valid) {
                    float f[8], w[8];
                    unpack8(xg[c], f);
                    unpack8(wg[c], w);
#pragma unroll
                    for (int i = 0; i < 8; ++i) f[i] = rbf(__fmul_rn(f[i], rs)) * w[i];
                    o.x = pack2(f[0], f[1]); o.y = pack2(f[2], f[3]);
                    o.z = pack2(f[4], f[5]); o.w = pack2(f[6], f[7]);
                    if (a.xnorm_out != nullptr && b == 0)
                        reinterpret_cast<uint4*>(a.xnorm_out + (size_t)m * K)[c] = o;
                }
                xs[m * KC + c] = o;
            }
        } else {
            for (int c = tid; c < KC; c += GEMV_THREADS)
                xs[m * KC + c] = valid ? xg[c] : make_uint4(0, 0, 0, 0);
        }
    }
    __syncthreads();
    if (nrows <= 0) return;

    // ---- stream the slab: groups of <= G rows, all 256 threads split K
    const int ngroups = (nrows + G - 1) / G;
    for (int gi = 0; gi < ngroups; ++gi) {
        const int g0 = r0 + (int)(((long long)nrows * gi) / ngroups);
        const int g1 = r0 + (int)(((long long)nrows * (gi + 1)) / ngroups);
        const int ng = g1 - g0;  // <= G
        float acc[NACC][G][M];
#pragma unroll
        for (int q = 0; q < NACC; ++q)
#pragma unroll
            for (int g = 0; g < G; ++g)
#pragma unroll
                for (int m = 0; m < M; ++m) acc[q][g][m] = 0.f;

        const bf16* wbase = a.W + (size_t)g0 * K;
        const bf16* wbase2 = (EPI == EPI_SWIGLU) ? a.W2 + (size_t)g0 * K : nullptr;
        for (int c = tid; c < KC; c += GEMV_THREADS) {
            uint4 wv[NACC][G];
#pragma unroll
            for (int g = 0; g < G; ++g) {
                if (g < ng) {
                    wv[0][g] = ld_stream(wbase + (size_t)g * K + (size_t)c * 8);
                    if (EPI == EPI_SWIGLU) wv[NACC - 1][g] = ld_stream(wbase2 + (size_t)g * K + (size_t)c * 8);
                } else {
                    wv[0][g] = make_uint4(0, 0, 0, 0);
                    if (EPI == EPI_SWIGLU) wv[NACC - 1][g] = make_uint4(0, 0, 0, 0);
                }
            }
#pragma unroll
            for (int m = 0; m < M; ++m) {
                float xf[8];
                unpack8(xs[m * KC + c], xf);
#pragma unroll
                for (int q = 0; q < NACC; ++q)
#pragma unroll
                    for (int g = 0; g < G; ++g) {
                        float wf[8];
                        unpack8(wv[q][g], wf);
#pragma unroll
                        for (int i = 0; i < 8; ++i) acc[q][g][m] = fmaf(wf[i], xf[i], acc[q][g][m]);
                    }
            }
        }
        // ---- reduce over the CTA
#pragma unroll
        for (int q = 0; q < NACC; ++q)
#pragma unroll
            for (int g = 0; g < G; ++g)
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    float v = warp_sum(acc[q][g][m]);
                    if (lane == 0) part[warp][(q * G + g) * M + m] = v;
                }
        __syncthreads();
        if (tid < G * M) {
            const int g = tid / M, m = tid % M;
            if (g < ng && m < a.M_valid) {
                float s = 0.f, s2 = 0.f;
#pragma unroll
                for (int w = 0; w < NW; ++w) {
                    s += part[w][g * M + m];
                    if (EPI == EPI_SWIGLU) s2 += part[w][(G + g) * M + m];
                }
                const size_t oi = (size_t)m * a.N + (g0 + g);
                float o = rbf(s);  // Linear output, bf16
                if (EPI == EPI_RESIDUAL) {
                    o = __bfloat162float(a.res[oi]) + o;  // llama.py:841-842 residual add in bf16
                } else if (EPI == EPI_SWIGLU) {
                    const float u = rbf(s2);
                    const float sl = rbf(__fdiv_rn(o, 1.f + expf(-o)));  // F.silu in bf16
                    o = sl * u;
                }
                a.out[oi] = __float2bfloat16_rn(o);
            }
        }
        __syncthreads();
    }
}
