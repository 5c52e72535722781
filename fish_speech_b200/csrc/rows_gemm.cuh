// Weight-streaming skinny GEMM for decode steps with many rows (config 3: up to 128 requests in flight), on the
// 5th-generation tensor cores:  out[m][n] = epi( sum_k X[m][k] * W[n][k] ),  M <= 128 rows, W row-major [N][K] bf16
// exactly as the reference state dict stores a Linear weight (llama.py:856-860, 981-983).
//
// The op is HBM-bound (every weight byte is used for M <= 128 rows, ~2*M FLOP per byte against a ridge of ~220), so the
// design goal is the same as for the batch-1 GEMV: keep ~190 KB of weight bytes in flight per SM and never re-read a
// byte.  Tensor cores are used because CUDA-core FMAs cannot keep up with the stream at M = 64.
//
//  * Orientation: the WEIGHT tile is the MMA's A operand (128 weight rows = the 128 TMEM lanes), the activation rows are
//    the B operand (N = BN = 32/64/128 batch rows), both K-major, so neither needs a transpose:
//    D[n_local][m] (fp32, TMEM, 128 lanes x BN columns) += W[n0 + n_local][k] * X[m][k].
//  * tcgen05.mma.cta_group::1.kind::f16 (bf16 x bf16 -> fp32), issued by one thread; operands in shared memory in the
//    canonical 128-byte-swizzled K-major layout, which is what a TMA tensor-map box {64 k, rows} with SWIZZLE_128B writes.
//  * Pipeline: TMA producer warp -> ring of NST stages {W tile 16 KB (x2 for gate/up), X tile BN x 128 B} -> MMA warp ->
//    two TMEM accumulator buffers -> 4 epilogue warps (tcgen05.ld).  Weight tiles are requested BEFORE
//    griddepcontrol.wait (they do not depend on the previous kernel), activations after it.
//  * Work split = stream-K: the (weight tile, k block) units are cut into one contiguous range per CTA, so every SM
//    streams the same number of bytes whatever N is (2560 rows = 20 tiles would otherwise use 20 of 148 SMs).  A tile
//    whose k range is shared by several CTAs is finished by the LAST CTA to arrive: every contributor writes its fp32
//    partial tile to a per-CTA slab, and the finisher adds the partials in CTA order (its own from TMEM), so the result
//    does not depend on arrival order.  No second launch, no atomics on data.
//  * Epilogues (one rounding per Linear like the reference, llama.py:838-843, 985-986): store bf16; bf16(res + bf16(acc));
//    SwiGLU over two accumulators bf16(bf16(silu(bf16 a)) * bf16 b).
#pragma once
#include <cuda.h>

#include "common.cuh"
#include "gemv_tma.cuh"  // smem_u32, mbar_*

#define RG_BM 128
#define RG_BK 64
#define RG_THREADS 192
#define RG_EPI_THREADS 128
#define RG_MAX_STAGES 10
#define RG_MAX_TILES 2048
#define RG_SMEM_BUDGET (222 * 1024)

enum { RG_EPI_STORE = 0, RG_EPI_RESIDUAL = 1, RG_EPI_SWIGLU = 2 };

struct RowsGemmArgs {
    bf16* out;         // [M][N]
    const bf16* res;   // [M][N] (RG_EPI_RESIDUAL), may alias out
    float* partial;    // [grid][2][NW][BN][128] fp32 partial tiles of the stream-K split
    int* counters;     // [n_tiles] arrivals per tile; zero between launches
    int M, N, K, epi;
    int n_tiles, kb;   // tiles of 128 weight rows, k blocks of 64
    int nst;           // ring depth
    unsigned long long* trace;  // profiling aid: 8 clock64 stamps per CTA (nullptr = off)
    int defer;         // 1: every tile segment only writes its fp32 partial; rows_finish_norm_kernel sums them (few-tile ops)
};

__host__ __device__ constexpr uint32_t rg_stage_bytes(int bn, int nw) { return (uint32_t)(nw * RG_BM * RG_BK * 2 + bn * RG_BK * 2); }
__host__ __device__ constexpr int rg_tmem_cols(int bn, int nw) {
    const int c = 2 * nw * bn;
    return c <= 32 ? 32 : c <= 64 ? 64 : c <= 128 ? 128 : c <= 256 ? 256 : 512;
}
inline int rg_stages(int bn, int nw) {
    int n = (int)((RG_SMEM_BUDGET - 1024) / rg_stage_bytes(bn, nw));
    return n > RG_MAX_STAGES ? RG_MAX_STAGES : n;
}
inline size_t rg_smem_bytes(int bn, int nw) { return (size_t)rg_stages(bn, nw) * rg_stage_bytes(bn, nw) + 1024; }
inline size_t rg_partial_bytes(int grid, int bn) { return (size_t)grid * 2 * 2 * bn * RG_BM * sizeof(float); }

// ----------------------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ void rg_tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "l"(policy) : "memory");
}
__device__ __forceinline__ void rg_prefetch_map(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[smem] . B[smem]^T, bf16 inputs, fp32 accumulate; `acc` = 0 overwrites D
__device__ __forceinline__ void tc_mma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
// shared-memory matrix descriptor: K-major tile of rows x 64 bf16 (128 B per row), SWIZZLE_128B, 8-row groups 1024 B apart
__device__ __forceinline__ uint64_t tc_smem_desc(uint32_t saddr) {
    return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)(1024u >> 4) << 32) | ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
// instruction descriptor: D fp32, A/B bf16, both K-major, M = 128, N = bn
__host__ __device__ constexpr uint32_t tc_idesc_bf16(int bn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(bn >> 3) << 17) | ((uint32_t)(RG_BM >> 4) << 24);
}
// 16 consecutive fp32 columns of this thread's TMEM lane
__device__ __forceinline__ void tc_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void rg_epi_bar() { asm volatile("bar.sync 1, %0;" ::"n"(RG_EPI_THREADS) : "memory"); }

// CTA that owns unit u when U units are cut into G contiguous ranges [U*c/G, U*(c+1)/G)
__device__ __forceinline__ int rg_owner(long long u, long long U, int G) {
    int c = (int)((u * G) / U);
    while (c + 1 < G && (U * (c + 1)) / G <= u) ++c;
    while (c > 0 && (U * c) / G > u) --c;
    return c;
}

#define RG_STAMP(i) do { if (a.trace) a.trace[(size_t)blockIdx.x * 8 + (i)] = clock64(); } while (0)
#define RG_POLICY_EVICT_FIRST 0x12F0000000000000ull
#define RG_POLICY_EVICT_LAST 0x14F0000000000000ull

template <int BN, int NW>
__global__ void __launch_bounds__(RG_THREADS, 1)
rows_gemm_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmW2,
                 const __grid_constant__ CUtensorMap tmX, RowsGemmArgs a) {
    constexpr uint32_t W_BYTES = RG_BM * RG_BK * 2;
    constexpr uint32_t STAGE = rg_stage_bytes(BN, NW);
    constexpr int TCOLS = rg_tmem_cols(BN, NW);
    extern __shared__ uint8_t rg_smem_raw[];
    __shared__ uint64_t full_bar[RG_MAX_STAGES], empty_bar[RG_MAX_STAGES], tfull_bar[2], tempty_bar[2];
    __shared__ uint32_t tmem_base_s;
    __shared__ int s_last, s_cf, s_cl;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) RG_STAMP(0);
    const uint32_t sbase = (smem_u32(rg_smem_raw) + 1023u) & ~1023u;
    const int nst = a.nst;
    const long long U = (long long)a.n_tiles * a.kb;
    const int G = gridDim.x, c = blockIdx.x;
    const int u0 = (int)((U * c) / G), u1 = (int)((U * (c + 1)) / G);

    if (tid == 0) {
        for (int s = 0; s < nst; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&tfull_bar[b], 1); mbar_init(&tempty_bar[b], 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(TCOLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;
    pdl_launch_dependents();
    if (tid == 0) RG_STAMP(1);

    if (warp == 0) {
        // ===== TMA producer
        if (lane == 0) {
            rg_prefetch_map(&tmW);
            if (NW == 2) rg_prefetch_map(&tmW2);
            rg_prefetch_map(&tmX);
            const int n_units = u1 - u0;
            const int npre = n_units < nst ? n_units : nst;
            // weights of the first ring fill do not depend on the previous kernel: request them before the dependency wait
            for (int i = 0; i < npre; ++i) {
                const int u = u0 + i, t = u / a.kb, kbi = u - t * a.kb;
                const uint32_t st = sbase + (uint32_t)i * STAGE, bar = smem_u32(&full_bar[i]);
                mbar_expect_tx(&full_bar[i], STAGE);
                rg_tma_load_2d(st, &tmW, bar, kbi * RG_BK, t * RG_BM, RG_POLICY_EVICT_FIRST);
                if (NW == 2) rg_tma_load_2d(st + W_BYTES, &tmW2, bar, kbi * RG_BK, t * RG_BM, RG_POLICY_EVICT_FIRST);
            }
            pdl_wait();
            for (int i = 0; i < npre; ++i) {
                const int u = u0 + i, t = u / a.kb, kbi = u - t * a.kb;
                rg_tma_load_2d(sbase + (uint32_t)i * STAGE + NW * W_BYTES, &tmX, smem_u32(&full_bar[i]), kbi * RG_BK, 0,
                               RG_POLICY_EVICT_LAST);
            }
            for (int i = npre; i < n_units; ++i) {
                const int u = u0 + i, t = u / a.kb, kbi = u - t * a.kb;
                const int s = i % nst;
                const uint32_t use = (uint32_t)(i / nst);
                mbar_wait(&empty_bar[s], (use & 1u) ^ 1u);
                const uint32_t st = sbase + (uint32_t)s * STAGE, bar = smem_u32(&full_bar[s]);
                mbar_expect_tx(&full_bar[s], STAGE);
                rg_tma_load_2d(st, &tmW, bar, kbi * RG_BK, t * RG_BM, RG_POLICY_EVICT_FIRST);
                if (NW == 2) rg_tma_load_2d(st + W_BYTES, &tmW2, bar, kbi * RG_BK, t * RG_BM, RG_POLICY_EVICT_FIRST);
                rg_tma_load_2d(st + NW * W_BYTES, &tmX, bar, kbi * RG_BK, 0, RG_POLICY_EVICT_LAST);
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer
        constexpr uint32_t IDESC = tc_idesc_bf16(BN);
        int iseg = 0;
        for (int u = u0; u < u1; ++iseg) {
            const int t = u / a.kb, k0 = u - t * a.kb;
            const int nk = min(a.kb - k0, u1 - u);
            const int buf = iseg & 1;
            const uint32_t use = (uint32_t)(iseg >> 1);
            mbar_wait(&tempty_bar[buf], (use & 1u) ^ 1u);  // the epilogue has drained this accumulator buffer
            tc_fence_after();
            for (int i = 0; i < nk; ++i) {
                const int idx = (u - u0) + i, s = idx % nst;
                mbar_wait(&full_bar[s], (uint32_t)(idx / nst) & 1u);
                tc_fence_after();
                if (lane == 0 && idx == 0) RG_STAMP(2);
                if (lane == 0) {
                    const uint32_t st = sbase + (uint32_t)s * STAGE;
                    const uint64_t bdesc = tc_smem_desc(st + NW * W_BYTES);
#pragma unroll
                    for (int w = 0; w < NW; ++w) {
                        const uint64_t adesc = tc_smem_desc(st + w * W_BYTES);
                        const uint32_t d = tmem + (uint32_t)((buf * NW + w) * BN);
#pragma unroll
                        for (int k = 0; k < RG_BK / 16; ++k)  // K = 16 per instruction: +32 B inside the swizzle atom
                            tc_mma_bf16(d, adesc + 2 * k, bdesc + 2 * k, IDESC, (i > 0 || k > 0) ? 1u : 0u);
                    }
                    tc_commit(&empty_bar[s]);  // arrives when the MMAs above have read the stage
                }
                __syncwarp();
            }
            if (lane == 0) tc_commit(&tfull_bar[buf]);
            __syncwarp();
            u += nk;
        }
        if (lane == 0) RG_STAMP(3);
    } else {
        // ===== epilogue: warp w may touch TMEM lanes 32*(w%4) .. +31; thread = one weight row of the tile
        const int et = tid - 64;
        const int lg = warp & 3;
        const int r = lg * 32 + lane;
        const uint32_t tlane = (uint32_t)(lg * 32) << 16;
        pdl_wait();
        int iseg = 0;
        for (int u = u0; u < u1; ++iseg) {
            const int t = u / a.kb, k0 = u - t * a.kb;
            const int nk = min(a.kb - k0, u1 - u);
            const int buf = iseg & 1;
            const bool whole = (nk == a.kb);
            mbar_wait(&tfull_bar[buf], (uint32_t)(iseg >> 1) & 1u);
            tc_fence_after();
            if (et == 0) { if (iseg == 0) RG_STAMP(4); if (u + nk >= u1) RG_STAMP(5); }
            const uint32_t tacc = tmem + tlane + (uint32_t)(buf * NW * BN);
            bool finish = whole && !a.defer;
            int c_first = c, c_last = c;
            if (!finish) {
                const int slot = (t == u0 / a.kb) ? 0 : 1;
                float* P = a.partial + (size_t)(c * 2 + slot) * NW * BN * RG_BM;
#pragma unroll 1
                for (int w = 0; w < NW; ++w)
#pragma unroll 1
                    for (int ch = 0; ch < BN / 16; ++ch) {
                        if (ch * 16 >= a.M) break;
                        float v[16];
                        tc_ld16(tacc + (uint32_t)(w * BN + ch * 16), v);
#pragma unroll
                        for (int j = 0; j < 16; ++j)
                            if (ch * 16 + j < a.M) P[(size_t)(w * BN + ch * 16 + j) * RG_BM + r] = v[j];
                    }
            }
            if (!finish && !a.defer) {
                __threadfence();
                rg_epi_bar();
                if (et == 0) {
                    const int cf = rg_owner((long long)t * a.kb, U, G), cl = rg_owner((long long)t * a.kb + a.kb - 1, U, G);
                    const int prev = atomicAdd(&a.counters[t], 1);
                    const int last = (prev == cl - cf);
                    if (last) a.counters[t] = 0;  // re-arm for the next launch
                    s_last = last;
                    s_cf = cf;
                    s_cl = cl;
                }
                rg_epi_bar();
                finish = s_last != 0;
                if (finish) {
                    __threadfence();
                    c_first = s_cf;
                    c_last = s_cl;
                }
            }
            if (finish) {
                const int n = t * RG_BM + r;
                const bool nv = n < a.N;
                constexpr int GRP = (NW == 2) ? 2 : 4;  // contributors whose partial tiles are in flight together
#pragma unroll 1
                for (int ch = 0; ch < BN / 16; ++ch) {
                    if (ch * 16 >= a.M) break;
                    float acc[NW][16];
#pragma unroll
                    for (int w = 0; w < NW; ++w) tc_ld16(tacc + (uint32_t)(w * BN + ch * 16), acc[w]);
                    if (!whole) {
                        float own[NW][16];
#pragma unroll
                        for (int w = 0; w < NW; ++w)
#pragma unroll
                            for (int j = 0; j < 16; ++j) { own[w][j] = acc[w][j]; acc[w][j] = 0.f; }
                        for (int cc0 = c_first; cc0 <= c_last; cc0 += GRP) {
                            float pv[GRP][NW][16];
#pragma unroll
                            for (int g = 0; g < GRP; ++g) {
                                const int cc = cc0 + g;
                                if (cc <= c_last && cc != c) {
                                    const int slot = ((int)((U * cc) / G) / a.kb == t) ? 0 : 1;
                                    const float* P = a.partial + (size_t)(cc * 2 + slot) * NW * BN * RG_BM;
#pragma unroll
                                    for (int w = 0; w < NW; ++w)
#pragma unroll
                                        for (int j = 0; j < 16; ++j)
                                            pv[g][w][j] = (ch * 16 + j < a.M) ? __ldcg(&P[(size_t)(w * BN + ch * 16 + j) * RG_BM + r]) : 0.f;
                                }
                            }
#pragma unroll
                            for (int g = 0; g < GRP; ++g) {  // fixed CTA order: the sum does not depend on who finishes
                                const int cc = cc0 + g;
                                if (cc > c_last) break;
#pragma unroll
                                for (int w = 0; w < NW; ++w)
#pragma unroll
                                    for (int j = 0; j < 16; ++j) acc[w][j] += (cc == c) ? own[w][j] : pv[g][w][j];
                            }
                        }
                    }
                    if (nv) {
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            const int m = ch * 16 + j;
                            if (m >= a.M) break;
                            const size_t oi = (size_t)m * a.N + n;
                            float o;
                            if (NW == 2) {  // SwiGLU (llama.py:986): silu(w1 x) * w3 x, each op rounded to bf16
                                const float g = rbf(acc[0][j]), up = rbf(acc[NW - 1][j]);
                                o = rbf(__fdiv_rn(g, 1.f + expf(-g))) * up;
                            } else if (a.epi == RG_EPI_RESIDUAL) {
                                o = lda(&a.res[oi]) + rbf(acc[0][j]);
                            } else {
                                o = acc[0][j];
                            }
                            a.out[oi] = __float2bfloat16_rn(o);
                        }
                    }
                }
            }
            tc_fence_before();
            rg_epi_bar();
            if (et == 0) mbar_arrive(&tempty_bar[buf]);
            u += nk;
        }
        if (et == 0) RG_STAMP(6);
    }
    tc_fence_before();
    __syncthreads();
    if (tid == 0) RG_STAMP(7);
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TCOLS) : "memory");
    }
}

// ----------------------------------------------------------------------------- deferred finish + RMSNorm
// Row-parallel second half of a few-tile Linear (wo, w2: 20 weight tiles shared by ~8 CTAs each): sums the fp32 partial
// tiles rows_gemm_kernel left behind (defer = 1) in CTA order, applies the residual (llama.py:841-842) and, since an
// RMSNorm follows every such Linear on the path (ffn_norm, the next layer's attention_norm, the final norms), produces
// the normalised rows for the next Linear in the same pass (llama.py:995-1000, double rounding).  One CTA per row: the
// partial reads are spread over M SMs instead of landing on the 20 SMs that would finish the tiles in-kernel.
#define RF_THREADS 512
#define RF_MAX_TILES 128
struct RowsFinishArgs {
    const float* partial;  // partial tiles of the pending Linear, or nullptr: x_in already holds finished rows
    const bf16* res;       // [M][N] residual of the pending Linear
    const bf16* x_in;      // [M][N] finished rows (partial == nullptr)
    bf16* x_out;           // [M][N] finished rows (written when partial != nullptr)
    const bf16* norm_w;    // [N] RMSNorm weight, or nullptr: no norm
    bf16* xn;              // [M][N] normalised rows
    int N, bn, G, kb, n_tiles;
    float eps;
};
__global__ void __launch_bounds__(RF_THREADS) rows_finish_norm_kernel(RowsFinishArgs a) {
    __shared__ float red[32];
    __shared__ int s_cf[RF_MAX_TILES], s_cl[RF_MAX_TILES], s_sf[RF_MAX_TILES];
    const int m = blockIdx.x, tid = threadIdx.x;
    const long long U = (long long)a.n_tiles * a.kb;
    if (a.partial != nullptr)
        for (int t = tid; t < a.n_tiles; t += RF_THREADS) {
            const int cf = rg_owner((long long)t * a.kb, U, a.G);
            s_cf[t] = cf;
            s_cl[t] = rg_owner((long long)t * a.kb + a.kb - 1, U, a.G);
            s_sf[t] = ((int)((U * cf) / a.G) / a.kb == t) ? 0 : 1;
        }
    pdl_wait();
    pdl_launch_dependents();
    __syncthreads();
    float ss = 0.f;
    const size_t row = (size_t)m * a.N;
    if (a.partial != nullptr) {
        // all of a thread's partial values are requested before the first one is used (8 contributors per pass): the loop
        // is a chain of L2 round trips otherwise
        constexpr int SEG = 8;
        for (int n = tid; n < a.N; n += RF_THREADS) {
            const int t = n >> 7, r = n & 127;
            const int cf = s_cf[t], cl = s_cl[t], slot_f = s_sf[t];
            const float resv = lda(&a.res[row + n]);
            float v = 0.f;
            for (int c0 = cf; c0 <= cl; c0 += SEG) {
                float pv[SEG];
#pragma unroll
                for (int g = 0; g < SEG; ++g) {
                    const int cc = c0 + g;
                    const int slot = (cc == cf) ? slot_f : 0;  // a CTA that starts inside this tile holds it in slot 0
                    pv[g] = (cc <= cl) ? __ldcg(&a.partial[((size_t)(cc * 2 + slot) * a.bn + m) * RG_BM + r]) : 0.f;
                }
#pragma unroll
                for (int g = 0; g < SEG; ++g)
                    if (c0 + g <= cl) v += pv[g];  // CTA order
            }
            const float o = rbf(resv + rbf(v));
            a.x_out[row + n] = __float2bfloat16_rn(o);
            ss += o * o;
        }
    } else {
        for (int n = tid; n < a.N; n += RF_THREADS) {
            const float o = lda(&a.x_in[row + n]);
            ss += o * o;
        }
    }
    if (a.norm_w == nullptr) return;
    ss = block_sum(ss, red);
    const float rs = __frsqrt_rn(__fadd_rn(__fdiv_rn(ss, (float)a.N), a.eps));
    const bf16* src = a.partial != nullptr ? a.x_out : a.x_in;
    for (int n = tid; n < a.N; n += RF_THREADS) {
        const float o = __bfloat162float(src[row + n]);  // this thread's own store above, or an input row
        a.xn[row + n] = __float2bfloat16_rn(rbf(__fmul_rn(o, rs)) * __bfloat162float(a.norm_w[n]));
    }
}

// ----------------------------------------------------------------------------- host side
typedef CUresult (*rg_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline rg_encode_fn rg_get_encode() {
    static rg_encode_fn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (rg_encode_fn)p;
    }
    return fn;
}
// 2-D map over a row-major bf16 matrix [rows][cols]: box = 64 columns (one 128-byte swizzle atom) x box_rows rows;
// rows beyond `rows` read as zero (weight tiles past N, activation rows past M)
inline bool rg_make_map(CUtensorMap* m, const void* ptr, uint64_t rows, uint64_t cols, uint32_t box_rows) {
    rg_encode_fn enc = rg_get_encode();
    if (!enc) return false;
    const cuuint64_t gdim[2] = {cols, rows};
    const cuuint64_t gstr[1] = {cols * 2};
    const cuuint32_t box[2] = {RG_BK, box_rows};
    const cuuint32_t est[2] = {1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstr, box, est, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

typedef void (*rows_gemm_fn)(const CUtensorMap, const CUtensorMap, const CUtensorMap, RowsGemmArgs);
inline rows_gemm_fn rg_pick(int bn, int nw) {
    if (nw == 2) return bn == 32 ? rows_gemm_kernel<32, 2> : bn == 64 ? rows_gemm_kernel<64, 2> : rows_gemm_kernel<128, 2>;
    return bn == 32 ? rows_gemm_kernel<32, 1> : bn == 64 ? rows_gemm_kernel<64, 1> : rows_gemm_kernel<128, 1>;
}
inline int rg_bn_for(int M) { return M <= 32 ? 32 : M <= 64 ? 64 : 128; }
inline cudaError_t rg_set_attrs() {
    static bool done = false;
    if (done) return cudaSuccess;
    const int bns[3] = {32, 64, 128};
    for (int nw = 1; nw <= 2; ++nw)
        for (int i = 0; i < 3; ++i) {
            cudaError_t e = cudaFuncSetAttribute((const void*)rg_pick(bns[i], nw), cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 (int)rg_smem_bytes(bns[i], nw));
            if (e != cudaSuccess) return e;
        }
    done = true;
    return cudaSuccess;
}

// One launch.  Requirements (checked by the callers): K % 64 == 0, N <= 128 * RG_MAX_TILES, M <= 128, 16-byte aligned
// pointers; `partial` holds rg_partial_bytes(grid, BN) bytes and `counters` RG_MAX_TILES zeroed ints.
inline cudaError_t rg_launch(cudaStream_t s, bool pdl, int n_sm, const CUtensorMap& mW, const CUtensorMap& mW2, const CUtensorMap& mX,
                             int M, int N, int K, int epi, const bf16* res, bf16* out, float* partial, int* counters, int defer = 0,
                             unsigned long long* trace = nullptr) {
    const int bn = rg_bn_for(M), nw = (epi == RG_EPI_SWIGLU) ? 2 : 1;
    RowsGemmArgs a{};
    a.out = out; a.res = res; a.partial = partial; a.counters = counters;
    a.M = M; a.N = N; a.K = K; a.epi = epi;
    a.n_tiles = (N + RG_BM - 1) / RG_BM; a.kb = K / RG_BK; a.nst = rg_stages(bn, nw); a.defer = defer; a.trace = trace;
    const long long U = (long long)a.n_tiles * a.kb;
    const int grid = (int)(U < n_sm ? U : n_sm);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(RG_THREADS); cfg.dynamicSmemBytes = rg_smem_bytes(bn, nw); cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, rg_pick(bn, nw), mW, mW2, mX, a);
}
