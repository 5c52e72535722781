// Weight-streaming GEMV, bulk-async-copy (TMA engine, cp.async.bulk -> SASS UBLKCP) version.
//
// Same contract and rounding points as gemv.cuh (see there for the reference call sites).  Difference: the
// weight slab of the CTA is staged through a ring of shared-memory stages by ONE producer warp with
// cp.async.bulk + mbarrier complete_tx, so the bytes in flight per SM (NST stages x ~20 KB x 2 CTAs) no longer
// depend on what the math warps are doing, and the producer starts streaming BEFORE griddepcontrol.wait
// (weights are immutable), which keeps HBM busy across the PDL kernel boundary.
//
//   stage  = 8 rows x KC columns (x2 matrices for SwiGLU): one bulk copy per row slice (KC*2 bytes, contiguous)
//   warp w = row w of the current 8-row group; accumulates over the K/KC stages of the group, then finalises
//   full[s]/empty[s] mbarriers: producer arrives with expect_tx, 8 consumer warps arrive to release a stage
#pragma once
#include "common.cuh"
#include "gemv.cuh"

#define GT_ROWS 8
#define GT_CONSUMERS 256
#define GT_THREADS (GT_CONSUMERS + 32)
#define GT_MAX_STAGES 6

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(smem_u32(b)), "r"(parity)
            : "memory");
    } while (!ok);
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
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
    int KC;       // columns per stage (divides K, multiple of 8)
    int NST;      // ring depth
};

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
    const int total = ngroups * NK;
    const uint32_t slice_bytes = (uint32_t)g.KC * 2u;
    const uint32_t stage_bytes = (uint32_t)GT_ROWS * slice_bytes * NACC;
    unsigned char* stages = smem_raw;                                         // [NST][NACC][8][KC] bf16
    uint4* xs = reinterpret_cast<uint4*>(smem_raw + (size_t)g.NST * stage_bytes);  // [M][K/8]

    if (tid == 0) {
        for (int s = 0; s < g.NST; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], GT_CONSUMERS / 32); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (warp == GT_CONSUMERS / 32) {
        // ================= producer warp: weights only, never waits on the previous kernel =================
        for (int it = 0; it < total; ++it) {
            const int slot = it % g.NST, ph = (it / g.NST) & 1;
            const int gi = it / NK, kc = it % NK;
            const int g0 = r0 + gi * GT_ROWS;
            const int ng = min(GT_ROWS, r1 - g0);
            mbar_wait(&empty_bar[slot], ph ^ 1);
            if (lane == 0) mbar_expect_tx(&full_bar[slot], (uint32_t)ng * slice_bytes * NACC);
            __syncwarp();
            unsigned char* sb = stages + (size_t)slot * stage_bytes;
            for (int c = lane; c < ng * NACC; c += 32) {
                const int q = c / ng, i = c % ng;
                const bf16* src = (q == 0 ? a.W : a.W2) + (size_t)(g0 + i) * K + (size_t)kc * g.KC;
                bulk_g2s(sb + ((size_t)q * GT_ROWS + i) * slice_bytes, src, slice_bytes, &full_bar[slot]);
            }
        }
        if (lane == 0) l2_prefetch_share(a.pf_ptr, a.pf_bytes, b, nb);
        pdl_launch_dependents();
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

    const int SC8 = g.KC >> 3;  // uint4 per row slice
    float acc[NACC][M];
    for (int it = 0; it < total; ++it) {
        const int slot = it % g.NST, ph = (it / g.NST) & 1;
        const int gi = it / NK, kc = it % NK;
        const int g0 = r0 + gi * GT_ROWS;
        const int ng = min(GT_ROWS, r1 - g0);
        if (kc == 0) {
#pragma unroll
            for (int q = 0; q < NACC; ++q)
#pragma unroll
                for (int m = 0; m < M; ++m) acc[q][m] = 0.f;
        }
        mbar_wait(&full_bar[slot], ph);
        if (warp < ng) {
            const uint4* w0 = reinterpret_cast<const uint4*>(stages + (size_t)slot * stage_bytes + (size_t)warp * slice_bytes);
            const uint4* w1 = reinterpret_cast<const uint4*>(stages + (size_t)slot * stage_bytes +
                                                            ((size_t)GT_ROWS + warp) * slice_bytes);
            const uint4* xk = xs + kc * SC8;
#pragma unroll 2
            for (int c = lane; c < SC8; c += 32) {
                float wf[8], wf2[8];
                unpack8(w0[c], wf);
                if (NACC == 2) unpack8(w1[c], wf2);
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    float xf[8];
                    unpack8(xk[m * KC8 + c], xf);
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        acc[0][m] = fmaf(wf[i], xf[i], acc[0][m]);
                        if (NACC == 2) acc[NACC - 1][m] = fmaf(wf2[i], xf[i], acc[NACC - 1][m]);
                    }
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[slot]);
        if (kc == NK - 1 && warp < ng) {
#pragma unroll
            for (int m = 0; m < M; ++m) {
                const float s = warp_sum(acc[0][m]);
                const float s2 = (NACC == 2) ? warp_sum(acc[NACC - 1][m]) : 0.f;
                if (lane == 0 && m < a.M_valid) {
                    const size_t oi = (size_t)m * a.N + (g0 + warp);
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
