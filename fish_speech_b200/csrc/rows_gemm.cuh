// Weight-streaming skinny GEMM for decode steps with many rows (config 3: up to 128 requests in flight), on the
// 5th-generation tensor cores:  out[m][n] = epi( sum_k X[m][k] * W[n][k] ),  M <= 128 rows, W row-major [N][K] bf16
// exactly as the reference state dict stores a Linear weight (llama.py:856-860, 981-983).
//
// The op is HBM-bound (every weight byte is used for M <= 128 rows, ~2*M FLOP per byte against a ridge of ~220), so the
// design goal is the same as for the batch-1 GEMV: keep the weight stream running and never re-read a byte.  Tensor cores
// are used because CUDA-core FMAs cannot keep up with the stream at M = 64.
//
//  * Orientation: the WEIGHT tile is the MMA's A operand (128 weight rows = the 128 TMEM lanes), the activation rows are
//    the B operand (N = BN = 32/64/128 batch rows), both K-major, so neither needs a transpose:
//    D[n_local][m] (fp32, TMEM, 128 lanes x BN columns) += W[n0 + n_local][k] * X[m][k].
//  * tcgen05.mma.cta_group::1.kind::f16 (bf16 x bf16 -> fp32), issued by one thread; operands in shared memory in the
//    canonical 128-byte-swizzled K-major layout, which is what a TMA tensor-map box {64 k, rows} with SWIZZLE_128B writes.
//  * Pipeline: TMA producer warp -> ring of 3 stages {W tile 16 KB, X tile BN x 128 B} -> MMA warp -> two TMEM
//    accumulator buffers -> 4 epilogue warps (tcgen05.ld).  Weight tiles are requested BEFORE griddepcontrol.wait (they
//    do not depend on the previous kernel), activations after it.  A CTA needs 73 KB of shared memory and 128 TMEM
//    columns at 64 rows, so three fit one SM: a launch uses two CTAs per SM and the third slot lets the NEXT kernel of
//    the PDL chain start its weight stream while this one drains (a short op is otherwise dominated by the ~5 us it takes
//    the first ring fill of all SMs to come back from HBM).
//  * Work split = stream-K: the (weight tile, k block) units are cut into one contiguous range per CTA, so every SM
//    streams the same number of bytes whatever N is (2560 rows = 20 tiles would otherwise use 20 of 148 SMs).  A tile
//    whose k range belongs to one CTA is finished in place (bf16, optional residual); a tile shared by several CTAs is
//    left as fp32 partial tiles in per-CTA slabs and finished by the small kernel that follows (rows_finish_tiles_kernel,
//    or rows_finish_norm_kernel fused with the RMSNorm after wo / w2), which adds the pieces in CTA order: deterministic,
//    no atomics, nothing serial on the GEMM's critical path.
//  * gate/up (w1, w3) run as ONE launch over the 2F rows of both matrices; the finish kernel applies SwiGLU
//    bf16(bf16(silu(bf16 a)) * bf16 b) (llama.py:986).  One rounding per Linear like the reference (llama.py:838-843).
#pragma once
#include <cuda.h>

#include "common.cuh"
#include "gemv_tma.cuh"
#include "tc_common.cuh"  // smem_u32, mbar_*, tcgen05 / TMA wrappers

#define RG_BM 128
#define RG_BK 64
#define RG_THREADS 192
#define RG_STAGES_MAX 9      // barrier arrays; the ring depth of a launch comes from rg_stages_for()
#define RG_STAGES_MIN 3      //   of the NEXT kernel fits beside them and starts its weight stream early)
#define RG_CTAS_PER_SM 2     // most CTAs of one launch per SM
#define RG_MAX_TILES 4096

enum { RG_EPI_STORE = 0, RG_EPI_RESIDUAL = 1, RG_EPI_SWIGLU = 2 };

struct RowsGemmArgs {
    bf16* out;         // [M][N] bf16 outputs of tiles whose whole k range belongs to one CTA (N = rows of W + rows of W2)
    const bf16* res;   // [M][N] residual (RG_EPI_RESIDUAL), may alias out
    float* partial;    // [grid][2][BN][128] fp32 partial tiles of the stream-K split
    int M, N, K, epi;
    int ldo;           // row stride of out / res in elements (N, or the full width when the launch covers a column range)
    int n_tiles, kb;   // tiles of 128 weight rows over W then W2, k blocks of 64
    int tiles_w;       // tiles that belong to W (the rest index W2)
    int nst;           // ring depth of this launch
    int all_partial;   // 1: every tile segment is left as an fp32 partial (rows_finish_norm_kernel finishes the op)
    unsigned long long* trace;  // profiling aid: 8 clock64 stamps per CTA (nullptr = off)
};

__host__ __device__ constexpr uint32_t rg_stage_bytes(int bn) { return (uint32_t)(RG_BM * RG_BK * 2 + bn * RG_BK * 2); }
__host__ __device__ constexpr int rg_tmem_cols(int bn) { return 2 * bn <= 32 ? 32 : 2 * bn <= 64 ? 64 : 2 * bn <= 128 ? 128 : 256; }
inline size_t rg_smem_bytes(int bn, int nst) { return (size_t)nst * rg_stage_bytes(bn) + 1024; }
inline size_t rg_partial_bytes(int grid, int bn) { return (size_t)grid * 2 * bn * RG_BM * sizeof(float); }

// ----------------------------------------------------------------------------- stream-K bookkeeping
// U = n_tiles * kb units are cut into G contiguous ranges [U*c/G, U*(c+1)/G), one per CTA.
// (32-bit arithmetic: U * G < 2^31 is checked on the host; 64-bit divisions cost ~100 instructions each on the device)
struct RgSplit {
    unsigned U;
    int G, kb;
    __host__ __device__ int start(int c) const { return (int)((U * (unsigned)c) / (unsigned)G); }
    __host__ __device__ int owner(int u) const {
        int c = (int)(((unsigned)u * (unsigned)G) / U);
        while (c + 1 < G && start(c + 1) <= u) ++c;
        while (c > 0 && start(c) > u) --c;
        return c;
    }
    // contributors of tile t: CTAs cf..cl; the first one holds its piece in slot sf, every later one in slot 0 (a CTA's
    // slot 0 is the tile its range starts in, slot 1 the tile it ends in)
    __host__ __device__ void tile(int t, int& cf, int& cl, int& sf) const {
        cf = owner(t * kb);
        cl = owner(t * kb + kb - 1);
        sf = (start(cf) / kb == t) ? 0 : 1;
    }
};
// CTAs of a launch.  Few tiles (<= 2 per SM): G = n_tiles * s, which makes the ranges tile-aligned (start(c) = kb*c/s), so
// every CTA holds exactly one partial tile and a tile has exactly s contributors - s = 3 (wqkv, 144 CTAs), 7 (wo / w2, 140),
// 2 (gate/up, 304).  Many tiles (LM head): plain stream-K over 2 CTAs per SM, ~300 split tiles out of 1217.
inline int rg_grid(int n_tiles, int kb, int n_sm) {
    if (n_tiles > RG_CTAS_PER_SM * n_sm) return RG_CTAS_PER_SM * n_sm;
    int s = n_tiles <= n_sm ? n_sm / n_tiles : (2 * n_tiles <= 3 * n_sm ? 2 : 1);
    if (s > kb) s = kb;
    return n_tiles * s;
}
inline int rg_stages_for(int grid, int n_sm, int bn) {
    if (grid > n_sm) return RG_STAGES_MIN;
    // one CTA per SM: 6 stages (144 KB in flight per SM at 64 rows).  Deeper is slower: every SM requests its whole ring at
    // launch, and the first stage of a CTA only lands when most of that burst has (9 stages: wqkv 11.1 us instead of 9.4)
    const int want = 6;
    const int fit = (int)((225 * 1024 - 1024) / rg_stage_bytes(bn));
    const int n = want < RG_STAGES_MAX ? want : RG_STAGES_MAX;
    return fit < n ? fit : n;
}

#define RG_STAMP(i) do { if (a.trace) a.trace[(size_t)blockIdx.x * 8 + (i)] = clock64(); } while (0)

template <int BN>
__global__ void __launch_bounds__(RG_THREADS, 3)
rows_gemm_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmW2,
                 const __grid_constant__ CUtensorMap tmX, RowsGemmArgs a) {
    constexpr uint32_t W_BYTES = RG_BM * RG_BK * 2;
    constexpr uint32_t STAGE = rg_stage_bytes(BN);
    constexpr int TCOLS = rg_tmem_cols(BN);
    extern __shared__ uint8_t rg_smem_raw[];
    __shared__ uint64_t full_bar[RG_STAGES_MAX], empty_bar[RG_STAGES_MAX], tfull_bar[2], tempty_bar[2];
    __shared__ uint32_t tmem_base_s;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) RG_STAMP(0);
    const uint32_t sbase = (smem_u32(rg_smem_raw) + 1023u) & ~1023u;
    const unsigned U = (unsigned)(a.n_tiles * a.kb);
    const int G = gridDim.x, c = blockIdx.x;
    const int u0 = (int)((U * (unsigned)c) / (unsigned)G), u1 = (int)((U * (unsigned)(c + 1)) / (unsigned)G);

    if (tid == 0) {
        for (int s = 0; s < a.nst; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
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
            rg_prefetch_map(&tmW2);
            rg_prefetch_map(&tmX);
            const int n_units = u1 - u0;
            const int npre = n_units < a.nst ? n_units : a.nst;
            // weights of the first ring fill do not depend on the previous kernel: request them before the dependency wait
            for (int i = 0; i < npre; ++i) {
                const int u = u0 + i, t = u / a.kb, kbi = u - t * a.kb;
                mbar_expect_tx(&full_bar[i], STAGE);
                if (t < a.tiles_w) rg_tma_load_2d(sbase + (uint32_t)i * STAGE, &tmW, smem_u32(&full_bar[i]), kbi * RG_BK, t * RG_BM, RG_POLICY_EVICT_FIRST);
                else rg_tma_load_2d(sbase + (uint32_t)i * STAGE, &tmW2, smem_u32(&full_bar[i]), kbi * RG_BK, (t - a.tiles_w) * RG_BM, RG_POLICY_EVICT_FIRST);
            }
            pdl_wait();
            for (int i = 0; i < npre; ++i) {
                const int u = u0 + i, t = u / a.kb, kbi = u - t * a.kb;
                rg_tma_load_2d(sbase + (uint32_t)i * STAGE + W_BYTES, &tmX, smem_u32(&full_bar[i]), kbi * RG_BK, 0, RG_POLICY_EVICT_LAST);
            }
            for (int i = npre; i < n_units; ++i) {
                const int u = u0 + i, t = u / a.kb, kbi = u - t * a.kb;
                const int s = i % a.nst;
                const uint32_t use = (uint32_t)(i / a.nst);
                mbar_wait(&empty_bar[s], (use & 1u) ^ 1u);
                const uint32_t st = sbase + (uint32_t)s * STAGE, bar = smem_u32(&full_bar[s]);
                mbar_expect_tx(&full_bar[s], STAGE);
                if (t < a.tiles_w) rg_tma_load_2d(st, &tmW, bar, kbi * RG_BK, t * RG_BM, RG_POLICY_EVICT_FIRST);
                else rg_tma_load_2d(st, &tmW2, bar, kbi * RG_BK, (t - a.tiles_w) * RG_BM, RG_POLICY_EVICT_FIRST);
                rg_tma_load_2d(st + W_BYTES, &tmX, bar, kbi * RG_BK, 0, RG_POLICY_EVICT_LAST);
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
                const int idx = (u - u0) + i, s = idx % a.nst;
                mbar_wait(&full_bar[s], (uint32_t)(idx / a.nst) & 1u);
                tc_fence_after();
                if (lane == 0) {
                    if (idx == 0) RG_STAMP(2);
                    const uint32_t st = sbase + (uint32_t)s * STAGE;
                    const uint64_t adesc = tc_smem_desc(st), bdesc = tc_smem_desc(st + W_BYTES);
                    const uint32_t d = tmem + (uint32_t)(buf * BN);
#pragma unroll
                    for (int k = 0; k < RG_BK / 16; ++k)  // K = 16 per instruction: +32 B inside the swizzle atom
                        tc_mma_bf16(d, adesc + 2 * k, bdesc + 2 * k, IDESC, (i > 0 || k > 0) ? 1u : 0u);
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
            const bool whole = (nk == a.kb) && !a.all_partial;
            mbar_wait(&tfull_bar[buf], (uint32_t)(iseg >> 1) & 1u);
            tc_fence_after();
            if (et == 0) { if (iseg == 0) RG_STAMP(4); if (u + nk >= u1) RG_STAMP(5); }
            const uint32_t tacc = tmem + tlane + (uint32_t)(buf * BN);
            const int n = t * RG_BM + r;
            float* P = a.partial + ((size_t)(c * 2 + ((t == u0 / a.kb) ? 0 : 1)) * BN) * RG_BM + r;
#pragma unroll 1
            for (int ch = 0; ch < BN / 16; ++ch) {
                if (ch * 16 >= a.M) break;
                float v[16];
                tc_ld16(tacc + (uint32_t)(ch * 16), v);
                if (!whole) {  // a piece of the tile's k range: fp32, finished by the kernel that follows
#pragma unroll
                    for (int j = 0; j < 16; ++j)
                        if (ch * 16 + j < a.M) P[(size_t)(ch * 16 + j) * RG_BM] = v[j];
                } else if (n < a.N) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const int m = ch * 16 + j;
                        if (m >= a.M) break;
                        const size_t oi = (size_t)m * a.ldo + n;
                        const float o = (a.epi == RG_EPI_RESIDUAL) ? lda(&a.res[oi]) + rbf(v[j]) : v[j];
                        a.out[oi] = __float2bfloat16_rn(o);
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
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TCOLS) : "memory");
    }
}

// ----------------------------------------------------------------------------- narrow weight tiles, no k split
// Mid-size Linears (wqkv: 6144 rows, fast head: 4096) have too few 128-row weight tiles for 148 SMs, and a k split costs
// partial-tile traffic plus a finish.  Here the operands are swapped: the ACTIVATION rows are the MMA's A operand (M = 128
// TMEM lanes = batch rows; rows past the batch are TMA out-of-bounds zero fill and cost no traffic) and the weight tile is
// the B operand, whose height NW only has to be a multiple of 16 - 48 rows give 128 CTAs for wqkv, each streaming its rows
// over the whole K with no split, no partials and a direct bf16 store (what cuBLAS picks for this shape, 9.4 us vs 12.8 us
// + finish for the 3-way split).  Price: an 8 KB activation block from L2 per 6 KB weight block - and that price decided
// it: measured in the step, this kernel for wqkv and the fast head is 3 % SLOWER than the 3-way split whose partial tiles
// are summed by the readers of q / k / v (10.54 vs 10.23 ms per step), so it is off by default (option rows_narrow).
//   D[m][n_local] (fp32, TMEM, 128 lanes x NW columns) += X[m][k] * W[n0 + n_local][k]
#define RGN_STAGES 8
#define RGN_STAGE_BYTES (16384 + 8192)  // X block 128 x 128 B | W block (<= 64 rows x 128 B)
struct RowsGemmNArgs {
    bf16* out;   // [M][N]
    int M, N, K, nw, kb;
};
__global__ void __launch_bounds__(RG_THREADS, 1)
rows_gemm_narrow_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmX, RowsGemmNArgs a) {
    extern __shared__ uint8_t rgn_smem_raw[];
    __shared__ uint64_t full_bar[RGN_STAGES], empty_bar[RGN_STAGES], tfull_bar;
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t sbase = (smem_u32(rgn_smem_raw) + 1023u) & ~1023u;
    const int n0 = blockIdx.x * a.nw;
    const uint32_t stage_tx = 16384u + (uint32_t)a.nw * 128u;
    if (tid == 0) {
        for (int s = 0; s < RGN_STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(&tfull_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(64) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;
    pdl_launch_dependents();
    if (warp == 0) {
        if (lane == 0) {
            rg_prefetch_map(&tmW);
            rg_prefetch_map(&tmX);
            const int npre = a.kb < RGN_STAGES ? a.kb : RGN_STAGES;
            for (int i = 0; i < npre; ++i) {  // weights first: they do not depend on the previous kernel
                mbar_expect_tx(&full_bar[i], stage_tx);
                rg_tma_load_2d(sbase + (uint32_t)i * RGN_STAGE_BYTES + 16384, &tmW, smem_u32(&full_bar[i]), i * RG_BK, n0, RG_POLICY_EVICT_FIRST);
            }
            pdl_wait();
            for (int i = 0; i < npre; ++i)
                rg_tma_load_2d(sbase + (uint32_t)i * RGN_STAGE_BYTES, &tmX, smem_u32(&full_bar[i]), i * RG_BK, 0, RG_POLICY_EVICT_LAST);
            for (int i = npre; i < a.kb; ++i) {
                const int s = i % RGN_STAGES;
                mbar_wait(&empty_bar[s], ((uint32_t)(i / RGN_STAGES) & 1u) ^ 1u);
                const uint32_t st = sbase + (uint32_t)s * RGN_STAGE_BYTES, bar = smem_u32(&full_bar[s]);
                mbar_expect_tx(&full_bar[s], stage_tx);
                rg_tma_load_2d(st + 16384, &tmW, bar, i * RG_BK, n0, RG_POLICY_EVICT_FIRST);
                rg_tma_load_2d(st, &tmX, bar, i * RG_BK, 0, RG_POLICY_EVICT_LAST);
            }
        }
    } else if (warp == 1) {
        const uint32_t idesc = tc_idesc_bf16(a.nw);  // M = 128 (batch rows), N = nw (weight rows)
        for (int i = 0; i < a.kb; ++i) {
            const int s = i % RGN_STAGES;
            mbar_wait(&full_bar[s], (uint32_t)(i / RGN_STAGES) & 1u);
            tc_fence_after();
            if (lane == 0) {
                const uint32_t st = sbase + (uint32_t)s * RGN_STAGE_BYTES;
                const uint64_t adesc = tc_smem_desc(st), bdesc = tc_smem_desc(st + 16384);
#pragma unroll
                for (int k = 0; k < RG_BK / 16; ++k) tc_mma_bf16(tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (i > 0 || k > 0) ? 1u : 0u);
                tc_commit(&empty_bar[s]);
            }
            __syncwarp();
        }
        if (lane == 0) tc_commit(&tfull_bar);
        __syncwarp();
    } else {
        // epilogue: TMEM lane = batch row; warps 4, 5 own lanes 0-63, warps 2, 3 lanes 64-127 (rows 64.. only when M > 64)
        const int lg = warp & 3, m = lg * 32 + lane;
        pdl_wait();
        mbar_wait(&tfull_bar, 0);
        tc_fence_after();
        if (lg * 32 < a.M) {
            const uint32_t tacc = tmem + ((uint32_t)(lg * 32) << 16);
            for (int ch = 0; ch < a.nw / 16; ++ch) {
                float v[16];
                tc_ld16(tacc + (uint32_t)(ch * 16), v);
                if (m < a.M) {
                    uint4 o0, o1;
                    o0.x = pack2(v[0], v[1]); o0.y = pack2(v[2], v[3]); o0.z = pack2(v[4], v[5]); o0.w = pack2(v[6], v[7]);
                    o1.x = pack2(v[8], v[9]); o1.y = pack2(v[10], v[11]); o1.z = pack2(v[12], v[13]); o1.w = pack2(v[14], v[15]);
                    uint4* dst = reinterpret_cast<uint4*>(a.out + (size_t)m * a.N + n0 + ch * 16);
                    dst[0] = o0; dst[1] = o1;
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(64) : "memory");
    }
}
inline size_t rgn_smem_bytes() { return (size_t)RGN_STAGES * RGN_STAGE_BYTES + 1024; }
// weight-tile height for a no-split launch: a multiple of 16 that divides N, N >= 16-byte aligned rows, and gives between
// 0.8 and 1.0 CTAs per SM; 0 if there is none (then the 128-row stream-K kernel is used)
inline int rgn_pick_nw(int N, int n_sm) {
    for (int nw = 64; nw >= 16; nw -= 16) {
        if (N % nw) continue;
        const int ctas = N / nw;
        if (ctas <= n_sm && ctas * 5 >= n_sm * 4) return nw;
    }
    return 0;
}
inline cudaError_t rgn_launch(cudaStream_t s, bool pdl, const CUtensorMap& mW, const CUtensorMap& mX, int M, int N, int K, int nw, bf16* out) {
    {
        FsbOnce once(3);
        if (once.first) {
            cudaError_t e = cudaFuncSetAttribute((const void*)rows_gemm_narrow_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rgn_smem_bytes());
            if (e != cudaSuccess) return e;
        }
    }
    RowsGemmNArgs a{};
    a.out = out; a.M = M; a.N = N; a.K = K; a.nw = nw; a.kb = K / RG_BK;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(N / nw); cfg.blockDim = dim3(RG_THREADS); cfg.dynamicSmemBytes = rgn_smem_bytes(); cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, rows_gemm_narrow_kernel, mW, mX, a);
}

// ----------------------------------------------------------------------------- finish of split tiles
// rows_gemm_kernel leaves every tile whose k range is shared by several CTAs as fp32 partial tiles.  The finish is a
// separate, embarrassingly parallel pass (a last-arriver fix-up inside the GEMM put 10-25 us of serial L2 round trips on
// the critical path of a 15 us kernel): one CTA per (tile, 16 rows), one thread per weight row, the contributors' values
// are added in CTA order (deterministic).  PDL-chained: it starts while the GEMM drains, and the next GEMM's CTAs start
// their weight stream while it runs.
//   mode RG_EPI_STORE / RG_EPI_RESIDUAL : out[m][n] = epi(bf16(sum))            for split tiles only
//   mode RG_EPI_SWIGLU                  : ff[m][n] = bf16(bf16(silu(g)) * up), g / up = the w1 / w3 halves of the GEMM's
//                                         N = 2F columns, each read from `pre` (whole tiles) or summed (split tiles)
#define RFT_ROWS 4    // rows per CTA, unrolled: 4 x <= 16 partial loads in flight per thread, one L2 round trip per CTA
struct RowsFinishTilesArgs {
    const float* partial;
    const bf16* pre;   // [M][N] bf16 outputs of whole tiles (SwiGLU mode input; aliases out in the other modes)
    const bf16* res;
    bf16* out;         // [M][N] (store / residual) or [M][N/2] (SwiGLU)
    int M, N, bn, G, kb, n_tiles, mode, ldo;  // ldo: row stride of out / res (store, residual modes)
};
__device__ __forceinline__ float rft_ld(const float* p) {  // volatile: the loads of one sum are issued back to back
    float v;
    asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}
template <int SEG>  // contributors whose pieces are in flight together
__device__ __forceinline__ float rft_sum(const float* partial, int bn, int cf, int cl, int sf, int m, int r) {
    float v = 0.f;
    for (int c0 = cf; c0 <= cl; c0 += SEG) {
        float pv[SEG];
#pragma unroll
        for (int g = 0; g < SEG; ++g) {
            const int cc = min(c0 + g, cl);  // past the end: re-read the last piece (dropped below) instead of branching
            pv[g] = rft_ld(&partial[((size_t)(cc * 2 + ((cc == cf) ? sf : 0)) * bn + m) * RG_BM + r]);
        }
#pragma unroll
        for (int g = 0; g < SEG; ++g) v += (c0 + g <= cl) ? pv[g] : 0.f;
    }
    return v;
}
// sums of RFT_ROWS consecutive rows at once: load phase (RFT_ROWS x 4 volatile loads) then sum phase; contributors past
// the 4th (not on the S2 shapes of the ops that use this kernel) go through rft_sum
__device__ __forceinline__ void rft_sum_rows(const float* partial, int bn, int cf, int cl, int sf, int m0, int M, int r, float (&v)[RFT_ROWS]) {
    float pv[RFT_ROWS][4];
#pragma unroll
    for (int i = 0; i < RFT_ROWS; ++i) {
        const int m = min(m0 + i, M - 1);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            const int cc = min(cf + g, cl);
            pv[i][g] = rft_ld(&partial[((size_t)(cc * 2 + ((cc == cf) ? sf : 0)) * bn + m) * RG_BM + r]);
        }
    }
    const int cnt = cl - cf + 1;
#pragma unroll
    for (int i = 0; i < RFT_ROWS; ++i) {
        float x = 0.f;
#pragma unroll
        for (int g = 0; g < 4; ++g) x += (g < cnt) ? pv[i][g] : 0.f;  // CTA order
        if (cnt > 4) x += rft_sum<4>(partial, bn, cf + 4, cl, 0, min(m0 + i, M - 1), r);
        v[i] = x;
    }
}
// grid.x: SwiGLU mode = the F/128 output tiles; otherwise the G-1 boundaries between CTA ranges (only a boundary that
// falls inside a tile splits it, and the first boundary inside a tile owns its finish), so a 1217-tile LM head launches
// ~300 small CTAs per row group, not one per tile.
__global__ void __launch_bounds__(RG_BM) rows_finish_tiles_kernel(RowsFinishTilesArgs a) {
    __shared__ int s_seg[2][3];
    pdl_launch_dependents();
    const int r = threadIdx.x;
    const int m0 = blockIdx.y * RFT_ROWS;
    const RgSplit sp{(unsigned)(a.n_tiles * a.kb), a.G, a.kb};
    const bool swiglu = a.mode == RG_EPI_SWIGLU;
    const int half_tiles = a.n_tiles / 2;
    int t = blockIdx.x;
    if (!swiglu) {
        const int u = sp.start(blockIdx.x + 1);
        if (u % a.kb == 0) return;  // the boundary coincides with a tile boundary
        t = u / a.kb;
    }
    if (r < (swiglu ? 2 : 1)) sp.tile(t + r * half_tiles, s_seg[r][0], s_seg[r][1], s_seg[r][2]);
    __syncthreads();
    if (!swiglu && s_seg[0][0] != (int)blockIdx.x) return;  // an earlier boundary inside the same tile finishes it
    pdl_wait();
    const int n = t * RG_BM + r;
    if (!swiglu) {
        if (n >= a.N) return;
        float v[RFT_ROWS];
        rft_sum_rows(a.partial, a.bn, s_seg[0][0], s_seg[0][1], s_seg[0][2], m0, a.M, r, v);
#pragma unroll
        for (int i = 0; i < RFT_ROWS; ++i) {
            if (m0 + i >= a.M) break;
            const size_t oi = (size_t)(m0 + i) * a.ldo + n;
            a.out[oi] = __float2bfloat16_rn(a.mode == RG_EPI_RESIDUAL ? lda(&a.res[oi]) + rbf(v[i]) : v[i]);
        }
        return;
    }
    const int F = a.N / 2;
    const bool split_g = s_seg[0][0] != s_seg[0][1], split_u = s_seg[1][0] != s_seg[1][1];
    float g[RFT_ROWS], up[RFT_ROWS];
    if (split_g) rft_sum_rows(a.partial, a.bn, s_seg[0][0], s_seg[0][1], s_seg[0][2], m0, a.M, r, g);
    if (split_u) rft_sum_rows(a.partial, a.bn, s_seg[1][0], s_seg[1][1], s_seg[1][2], m0, a.M, r, up);
#pragma unroll
    for (int i = 0; i < RFT_ROWS; ++i) {
        const int m = min(m0 + i, a.M - 1);
        g[i] = split_g ? rbf(g[i]) : lda(&a.pre[(size_t)m * a.N + n]);
        up[i] = split_u ? rbf(up[i]) : lda(&a.pre[(size_t)m * a.N + F + n]);
    }
#pragma unroll
    for (int i = 0; i < RFT_ROWS; ++i)
        if (m0 + i < a.M)
            a.out[(size_t)(m0 + i) * F + n] = __float2bfloat16_rn(rbf(__fdiv_rn(g[i], 1.f + expf(-g[i]))) * up[i]);  // llama.py:986
}

// ----------------------------------------------------------------------------- row-parallel finish + RMSNorm
// For the Linears with few weight tiles (wo, w2: N = dim = 20 tiles, every tile shared by ~15 CTAs) the GEMM leaves ALL
// tiles as partials and this kernel finishes whole rows: sum in CTA order, residual (llama.py:841-842) and, since an
// RMSNorm follows every such Linear on the path (ffn_norm, the next layer's attention_norm, the final norms), the
// normalised rows for the next Linear in the same pass (llama.py:995-1000, double rounding).  One CTA per row.
#define RF_THREADS 512
#define RF_EPT 5
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
    pdl_launch_dependents();
    const int m = blockIdx.x, tid = threadIdx.x;
    if (a.partial != nullptr) {
        const RgSplit sp{(unsigned)(a.n_tiles * a.kb), a.G, a.kb};
        for (int t = tid; t < a.n_tiles; t += RF_THREADS) sp.tile(t, s_cf[t], s_cl[t], s_sf[t]);
    }
    pdl_wait();
    __syncthreads();
    float ss = 0.f;
    const size_t row = (size_t)m * a.N;
    if (a.partial != nullptr) {
        // up to RF_EPT elements per thread; ALL of their partial values (8 contributors each) are requested before the first
        // one is used: written as a load phase of volatile loads followed by a sum phase, otherwise the compiler keeps the
        // round trips in a chain (one per element and group)
        for (int nb = 0; nb < a.N; nb += RF_THREADS * RF_EPT) {
            float pv[RF_EPT][8], resv[RF_EPT];
            int cnt[RF_EPT];
#pragma unroll
            for (int e = 0; e < RF_EPT; ++e) {
                const int n = min(nb + e * RF_THREADS + tid, a.N - 1);
                const int t = n >> 7, r = n & 127, cf = s_cf[t], cl = s_cl[t], sf = s_sf[t];
                cnt[e] = cl - cf + 1;
                resv[e] = lda(&a.res[row + n]);
#pragma unroll
                for (int g2 = 0; g2 < 8; ++g2) {
                    const int cc = min(cf + g2, cl);
                    pv[e][g2] = rft_ld(&a.partial[((size_t)(cc * 2 + ((cc == cf) ? sf : 0)) * a.bn + m) * RG_BM + r]);
                }
            }
#pragma unroll
            for (int e = 0; e < RF_EPT; ++e) {
                const int n = nb + e * RF_THREADS + tid;
                if (n < a.N) {
                    float v = 0.f;
#pragma unroll
                    for (int g2 = 0; g2 < 8; ++g2) v += (g2 < cnt[e]) ? pv[e][g2] : 0.f;  // CTA order
                    if (cnt[e] > 8) {  // more than 8 contributors (not on the S2 shapes): the rest in a second pass
                        const int t = n >> 7;
                        v += rft_sum<8>(a.partial, a.bn, s_cf[t] + 8, s_cl[t], 0, m, n & 127);
                    }
                    const float o = rbf(resv[e] + rbf(v));
                    a.x_out[row + n] = __float2bfloat16_rn(o);
                    ss += o * o;
                }
            }
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
inline rows_gemm_fn rg_pick(int bn) { return bn == 32 ? rows_gemm_kernel<32> : bn == 64 ? rows_gemm_kernel<64> : rows_gemm_kernel<128>; }
inline int rg_bn_for(int M) { return M <= 32 ? 32 : M <= 64 ? 64 : 128; }
inline cudaError_t rg_set_attrs() {
    FsbOnce once(2);
    if (!once.first) return cudaSuccess;
    const int bns[3] = {32, 64, 128};
    for (int i = 0; i < 3; ++i) {
        cudaError_t e = cudaFuncSetAttribute((const void*)rg_pick(bns[i]), cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)rg_smem_bytes(bns[i], rg_stages_for(1, 1, bns[i])));
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}
template <typename Args>
inline cudaError_t rg_launch_pdl(void (*k)(Args), dim3 grid, dim3 block, cudaStream_t s, bool pdl, const Args& a) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = 0; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, k, a);
}

// The GEMM launch.  N = rows of W (+ rows of W2, which must equal rows of W and be a multiple of 128 when W2 is given).
// Requirements (checked by the callers): K % 64 == 0, tiles <= RG_MAX_TILES, M <= 128, 16-byte aligned pointers; `partial`
// holds rg_partial_bytes(RG_CTAS_PER_SM * n_sm, BN) bytes.  Split tiles are left for rg_finish_tiles / rows_finish_norm_kernel.
inline cudaError_t rg_launch(cudaStream_t s, bool pdl, int n_sm, const CUtensorMap& mW, const CUtensorMap& mW2, const CUtensorMap& mX,
                             int M, int N, int tiles_w, int K, int epi, const bf16* res, bf16* out, float* partial, int all_partial,
                             unsigned long long* trace = nullptr, int ldo = 0) {
    const int bn = rg_bn_for(M);
    RowsGemmArgs a{};
    a.out = out; a.res = res; a.partial = partial;
    a.M = M; a.N = N; a.K = K; a.epi = epi; a.ldo = ldo > 0 ? ldo : N;
    a.n_tiles = (N + RG_BM - 1) / RG_BM; a.kb = K / RG_BK; a.tiles_w = tiles_w; a.all_partial = all_partial; a.trace = trace;
    const int grid = rg_grid(a.n_tiles, a.kb, n_sm);
    a.nst = rg_stages_for(grid, n_sm, bn);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(RG_THREADS); cfg.dynamicSmemBytes = rg_smem_bytes(bn, a.nst); cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, rg_pick(bn), mW, mW2, mX, a);
}
// finish of the split tiles of the launch above (same M, N, K, n_sm); mode SWIGLU: pre = the GEMM's out, out = [M][N/2]
inline cudaError_t rg_finish_tiles(cudaStream_t s, bool pdl, int n_sm, int M, int N, int K, int mode, const float* partial,
                                   const bf16* pre, const bf16* res, bf16* out, int ldo = 0) {
    RowsFinishTilesArgs f{};
    f.partial = partial; f.pre = pre; f.res = res; f.out = out;
    f.M = M; f.N = N; f.bn = rg_bn_for(M); f.kb = K / RG_BK; f.n_tiles = (N + RG_BM - 1) / RG_BM; f.mode = mode; f.ldo = ldo > 0 ? ldo : N;
    f.G = rg_grid(f.n_tiles, f.kb, n_sm);
    const int gx = mode == RG_EPI_SWIGLU ? f.n_tiles / 2 : f.G - 1;
    if (gx < 1) return cudaSuccess;
    return rg_launch_pdl(rows_finish_tiles_kernel, dim3(gx, (M + RFT_ROWS - 1) / RFT_ROWS), dim3(RG_BM), s, pdl, f);
}
