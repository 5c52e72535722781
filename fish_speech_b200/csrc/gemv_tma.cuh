// Weight-streaming GEMV, bulk-async-copy (TMA engine, cp.async.bulk -> SASS UBLKCP) version.
//
// Same contract and rounding points as gemv.cuh (see there for the reference call sites).  Difference: the
// weight slab of the CTA is staged through a ring of shared-memory stages by ONE producer warp with
// cp.async.bulk + mbarrier complete_tx, so the bytes in flight per SM (NST stages x ~20 KB x 2 CTAs) no longer
// depend on what the math warps are doing, and the producer starts streaming BEFORE griddepcontrol.wait
// (weights are immutable), which keeps HBM busy across the PDL kernel boundary.
//
//   CTA    = 1 producer warp + 2 teams of 8 consumer warps; grid = one CTA per SM, <= ~100 KB of shared memory and
//            <= 60 registers per thread, so the NEXT kernel's CTA fits on the SM beside it and fills its ring while this
//            one drains (the weights never depend on the previous kernel)
//   stage  = 8 rows x KC columns of ONE matrix: one bulk copy per row slice (KC*2 bytes, contiguous, >= 2 KB)
//   team t = row groups 2p+t; warp w of the team = row w of the group; accumulates over the K/KC stages
//            (x2 matrices for SwiGLU), then finalises.  Stage order interleaves the two teams so both stay busy.
//   full[s]/empty[s] mbarriers: producer arrives with expect_tx; ALL 16 consumer warps wait on and release every stage
//            (only the owning team computes on it), so nobody can run a phase ahead of or behind a barrier
#pragma once
#include "common.cuh"
#include "gemv.cuh"
#include "tc_common.cuh"  // smem_u32, mbar_*, bulk_g2s

#define GT_ROWS 8
#define GT_TEAM_WARPS 8
#define GT_TEAMS 2
#define GT_CONSUMERS (32 * GT_TEAM_WARPS * GT_TEAMS)
#define GT_THREADS (GT_CONSUMERS + 32)
#define GT_MAX_STAGES 10
#define GT_CPL 10  // 128-bit chunks per lane per row slice: KC <= 1280 columns

__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(GT_CONSUMERS) : "memory"); }

// sum over the 256 consumer threads (named barrier 1), result broadcast
__device__ __forceinline__ float consumer_sum(float v, float* red) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    v = warp_sum(v);
    consumer_sync();
    if (lane == 0) red[w] = v;
    consumer_sync();
    float t = (lane < GT_CONSUMERS / 32) ? red[lane] : 0.f;
    return warp_sum(t);
}

struct GemvTmaGeom {
    int KC;   // columns per stage (divides K, multiple of 8)
    int NST;  // ring depth
};

// stage iteration space shared by producer and consumers: it -> (team, matrix q, k-chunk, row group)
struct GtStage { int team, q, kc, gi; };
template <int NACC>
__device__ __forceinline__ GtStage gt_decode(int it, int NK) {
    GtStage s;
    s.team = it % GT_TEAMS;
    int r = it / GT_TEAMS;
    s.q = r % NACC; r /= NACC;
    s.kc = r % NK;
    s.gi = (r / NK) * GT_TEAMS + s.team;
    return s;
}

template <int M, int PRO, int EPI>
__global__ void __launch_bounds__(GT_THREADS, 1) gemv_tma_kernel(GemvArgs a, GemvTmaGeom g) {
    constexpr int NACC = (EPI == EPI_SWIGLU) ? 2 : 1;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[GT_MAX_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[GT_MAX_STAGES];
    __shared__ float red[32];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nb = gridDim.x, b = blockIdx.x;
    const int K = a.K, KC8 = K >> 3;
    const int r0 = (int)(((long long)a.N * b) / nb);
    const int r1 = (int)(((long long)a.N * (b + 1)) / nb);
    const int nrows = r1 - r0;
    const int ngroups = (nrows + GT_ROWS - 1) / GT_ROWS;
    const int NK = K / g.KC;
    const int total = ((ngroups + GT_TEAMS - 1) / GT_TEAMS) * NK * NACC * GT_TEAMS;
    const uint32_t slice_bytes = (uint32_t)g.KC * 2u;
    const uint32_t stage_bytes = (uint32_t)GT_ROWS * slice_bytes;
    unsigned char* stages = smem_raw;                                              // [NST][8][KC] bf16
    uint4* xs = reinterpret_cast<uint4*>(smem_raw + (size_t)g.NST * stage_bytes);  // [M][K/8]

    if (tid == 0) {
        for (int s = 0; s < g.NST; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], GT_CONSUMERS / 32); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (warp == GT_CONSUMERS / 32) {
        // ================= producer warp: weights only, never waits on the previous kernel =================
        int n = 0;
        for (int it = 0; it < total; ++it) {
            const GtStage st = gt_decode<NACC>(it, NK);
            if (st.gi >= ngroups) continue;
            const int slot = n % g.NST, ph = (n / g.NST) & 1;
            ++n;
            const int g0 = r0 + st.gi * GT_ROWS;
            const int ng = min(GT_ROWS, r1 - g0);
            mbar_wait(&empty_bar[slot], ph ^ 1);
            if (lane == 0) mbar_expect_tx(&full_bar[slot], (uint32_t)ng * slice_bytes);
            __syncwarp();
            if (lane < ng) {
                const bf16* src = (st.q == 0 ? a.W : a.W2) + (size_t)(g0 + lane) * K + (size_t)st.kc * g.KC;
                bulk_g2s(stages + (size_t)slot * stage_bytes + (size_t)lane * slice_bytes, src, slice_bytes, &full_bar[slot]);
            }
        }
        if (lane == 0) l2_prefetch_share(a.pf_ptr, a.pf_bytes, b, nb);
        return;
    }

    // ================= consumer warps =================
    pdl_wait();
    pdl_launch_dependents();
    for (int m = 0; m < M; ++m) {
        const uint4* xg = reinterpret_cast<const uint4*>(a.x + (size_t)m * K);
        const bool valid = m < a.M_valid;
        if (PRO == PRO_RMSNORM) {
            float ss = 0.f;
            if (valid)
                for (int c = tid; c < KC8; c += GT_CONSUMERS) {
                    float f[8];
                    unpack8(xg[c], f);
#pragma unroll
                    for (int i = 0; i < 8; ++i) ss += f[i] * f[i];
                }
            ss = consumer_sum(ss, red);
            const float rs = __frsqrt_rn(__fadd_rn(__fdiv_rn(ss, (float)K), a.eps));
            const uint4* wg = reinterpret_cast<const uint4*>(a.norm_w);
            for (int c = tid; c < KC8; c += GT_CONSUMERS) {
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
                xs[m * KC8 + c] = o;
            }
        } else {
            for (int c = tid; c < KC8; c += GT_CONSUMERS) xs[m * KC8 + c] = valid ? xg[c] : make_uint4(0, 0, 0, 0);
        }
    }
    consumer_sync();

    const int my_team = warp / GT_TEAM_WARPS, row = warp % GT_TEAM_WARPS;
    const int SC8 = g.KC >> 3;  // uint4 per row slice
    float acc[NACC][M];
    int n = 0;
    for (int it = 0; it < total; ++it) {
        const GtStage st = gt_decode<NACC>(it, NK);
        if (st.gi >= ngroups) continue;
        const int slot = n % g.NST, ph = (n / g.NST) & 1;
        ++n;
        if (st.team != my_team) {
            // every consumer warp observes and releases every stage, in order: an mbarrier waiter may be neither more
            // than one phase ahead (parity aliasing) nor behind (slot recycled under it) of the barrier it waits on
            mbar_wait(&full_bar[slot], ph);
            if (lane == 0) mbar_arrive(&empty_bar[slot]);
            continue;
        }
        const int g0 = r0 + st.gi * GT_ROWS;
        const int ng = min(GT_ROWS, r1 - g0);
        if (st.kc == 0 && st.q == 0) {
#pragma unroll
            for (int q = 0; q < NACC; ++q)
#pragma unroll
                for (int m = 0; m < M; ++m) acc[q][m] = 0.f;
        }
        mbar_wait(&full_bar[slot], ph);
        if (row < ng) {
            const uint4* w0 = reinterpret_cast<const uint4*>(stages + (size_t)slot * stage_bytes + (size_t)row * slice_bytes);
            const uint4* xk = xs + st.kc * SC8;
            // whole row slice of this lane in registers first (<= GT_CPL 128-bit LDS in flight), then the FMAs on two
            // independent partial sums per row of x: a single serial chain per warp was latency-bound
            float part[M];
            uint4 wv[GT_CPL];
#pragma unroll
            for (int i = 0; i < GT_CPL; ++i) {
                const int c = lane + 32 * i;
                wv[i] = (c < SC8) ? w0[c] : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int m = 0; m < M; ++m) {
                uint4 xv[GT_CPL];
#pragma unroll
                for (int i = 0; i < GT_CPL; ++i) {
                    const int c = lane + 32 * i;
                    xv[i] = (c < SC8) ? xk[m * KC8 + c] : make_uint4(0, 0, 0, 0);
                }
                float pa = 0.f, pb = 0.f;
#pragma unroll
                for (int i = 0; i < GT_CPL; ++i) {
                    float wf[8], xf[8];
                    unpack8(wv[i], wf);
                    unpack8(xv[i], xf);
#pragma unroll
                    for (int e = 0; e < 8; e += 2) {
                        pa = fmaf(wf[e], xf[e], pa);
                        pb = fmaf(wf[e + 1], xf[e + 1], pb);
                    }
                }
                part[m] = pa + pb;
            }
#pragma unroll
            for (int m = 0; m < M; ++m) {
                if (NACC == 2 && st.q == 1) acc[NACC - 1][m] += part[m];
                else acc[0][m] += part[m];
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[slot]);
        if (st.kc == NK - 1 && st.q == NACC - 1 && row < ng) {
#pragma unroll
            for (int m = 0; m < M; ++m) {
                const float s = warp_sum(acc[0][m]);
                const float s2 = (NACC == 2) ? warp_sum(acc[NACC - 1][m]) : 0.f;
                if (lane == 0 && m < a.M_valid) {
                    const size_t oi = (size_t)m * a.N + (g0 + row);
                    float o = rbf(s);
                    if (EPI == EPI_RESIDUAL) {
                        o = __bfloat162float(a.res[oi]) + o;
                    } else if (EPI == EPI_SWIGLU) {
                        const float u = rbf(s2);
                        const float sl = rbf(__fdiv_rn(o, 1.f + expf(-o)));
                        o = sl * u;
                    }
                    a.out[oi] = __float2bfloat16_rn(o);
                }
            }
        }
    }
}
