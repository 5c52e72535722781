// Causal attention of the prompt pass (forward_generate with Tq > 1: Attention.forward, llama.py:883-945, SDPA over the cache
// with the causal mask, :909-933) as a flash-attention forward on the 5th-generation tensor cores, head_dim 128.
//
//   qk_prep_kernel     q-norm / k-norm (single rounding, llama.py:862-863,907-909) + RoPE on the bf16 table (:1025-1037) for
//                      all rows of a chunk: q -> scratch [H][T][128], k -> KV cache, v -> KV cache (KVCache.update, :205-214)
//   vt_transpose_kernel V of the cache rows [0, n_keys) -> scratch [KV][128][S] (keys contiguous), so that BOTH operands of
//                      P.V are K-major like every other tile of this library
//   fa_prefill_kernel  one CTA = one query head x 128 query rows; per block of 128 keys (up to the diagonal):
//                        S = Q . K^T         tcgen05.mma  M 128 (queries) x N 128 (keys) x K 128 (head_dim), fp32 in TMEM
//                        softmax             one thread per query row reads its S row from TMEM, online max / sum in fp32
//                                            (base-2 exponent with the 1/sqrt(d) scale folded in), causal mask by position
//                        P (bf16) -> smem    written in the swizzled K-major operand layout, fence to the async proxy
//                        O_j = P . V         tcgen05.mma  M 128 x N 128 (head_dim) x K 128 (keys), fresh accumulator
//                        O = O * corr + O_j  in registers (128 fp32 per thread): no TMEM read-modify-write, no correction pass
//                      warp roles: TMA producer (Q once, K / V^T double-buffered) / MMA issuer / 4 softmax warps; S is double
//                      buffered in TMEM so Q.K^T of block j+1 runs under the softmax of block j.
#pragma once
#include "tc_common.cuh"
#include "attn.cuh"

#define FA_BQ 128
#define FA_BK 128
#define FA_HD 128
#define FA_SOFT_WARPS 8                               // two per TMEM lane quadrant: the quadrant's query rows, one half of the columns each
#define FA_THREADS (64 + 32 * FA_SOFT_WARPS)
#define FA_TILE_BYTES (128 * 128 * 2)                 // one 128 x 128 bf16 tile = two 16 KB swizzled boxes
#define FA_SMEM (4 * FA_TILE_BYTES + 2 * FA_TILE_BYTES + 1024)  // K, V^T double-buffered + Q + P

struct FaArgs {
    int T;            // query rows of this chunk
    int pos0;         // absolute position of query row 0
    int H, KV, S;     // heads, kv heads, cache length (row pitch of the V^T scratch)
    int slot;         // kv slot of the request
    int nqb;          // query blocks
    float scale_log2; // log2(e) / sqrt(head_dim)
    bf16* y;          // [T][H * 128]
};

// ----------------------------------------------------------------------------- q / k prep for whole chunks
struct QkPrepArgs {
    const bf16* qkv;   // [T][(H + 2 KV) * 128]
    bf16* q_out;       // [H][T][128]
    bf16 *kcache, *vcache;
    const bf16 *q_norm, *k_norm, *rope;
    const int32_t* pos;
    int T, H, KV, S, slot, qk_norm;
    float eps;
};
// grid (H + KV, T), 64 threads: warp 0 = the head's norm + RoPE; for kv heads warp 1 copies V
__global__ void __launch_bounds__(64) qk_prep_kernel(QkPrepArgs a) {
    constexpr int HD = FA_HD, EPL = HD / 32;
    const int hh = blockIdx.x, t = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int pos = max(0, min(a.pos[t], a.S - 1));
    const int qkv_w = (a.H + 2 * a.KV) * HD;
    const bf16* row = a.qkv + (size_t)t * qkv_w;
    if (hh < a.H) {
        if (warp != 0) return;
        float in[EPL], out[EPL];
        load_head<HD, false>(row + hh * HD, nullptr, 0, in);
        head_norm_rope<HD>(in, a.q_norm, a.rope + (size_t)pos * HD, a.qk_norm, a.eps, out);
        bf16* dst = a.q_out + ((size_t)hh * a.T + t) * HD + lane * EPL;
#pragma unroll
        for (int i = 0; i < EPL; ++i) dst[i] = __float2bfloat16_rn(out[i]);
        return;
    }
    const int h = hh - a.H;
    const size_t co = (((size_t)a.slot * a.KV + h) * a.S + pos) * HD;
    if (warp == 0) {
        float in[EPL], out[EPL];
        load_head<HD, false>(row + (a.H + h) * HD, nullptr, 0, in);
        head_norm_rope<HD>(in, a.k_norm, a.rope + (size_t)pos * HD, a.qk_norm, a.eps, out);
#pragma unroll
        for (int i = 0; i < EPL; ++i) a.kcache[co + lane * EPL + i] = __float2bfloat16_rn(out[i]);
    } else {
        const bf16* v = row + (a.H + a.KV + h) * HD;
#pragma unroll
        for (int i = 0; i < EPL; ++i) a.vcache[co + lane * EPL + i] = v[lane * EPL + i];
    }
}

// V^T scratch: vt[kvh][d][key] = vcache[slot][kvh][key][d] for key < n_keys.  grid (ceil(n_keys / 64), 2, KV), 256 threads
__global__ void __launch_bounds__(256) vt_transpose_kernel(const bf16* vcache, bf16* vt, int KV, int S, int slot, int n_keys) {
    __shared__ bf16 tile[64][64 + 2];
    const int k0 = blockIdx.x * 64, d0 = blockIdx.y * 64, h = blockIdx.z;
    const bf16* src = vcache + (((size_t)slot * KV + h) * S) * FA_HD;
    for (int e = threadIdx.x; e < 64 * 64; e += 256) {
        const int kk = e >> 6, dd = e & 63;
        tile[kk][dd] = (k0 + kk < n_keys) ? src[(size_t)(k0 + kk) * FA_HD + d0 + dd] : __float2bfloat16_rn(0.f);
    }
    __syncthreads();
    bf16* dst = vt + ((size_t)h * FA_HD + d0) * S;
    for (int e = threadIdx.x; e < 64 * 64; e += 256) {
        const int dd = e >> 6, kk = e & 63;
        if (k0 + kk < S) dst[(size_t)dd * S + k0 + kk] = tile[kk][dd];
    }
}

// ----------------------------------------------------------------------------- the attention kernel
__device__ __forceinline__ void fa_tma_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
// 2^x on the special-function unit (arguments here are <= 0: scores minus the running maximum, or -inf)
__device__ __forceinline__ float fa_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ void fa_soft_bar() { asm volatile("bar.sync 1, %0;" ::"n"(32 * FA_SOFT_WARPS) : "memory"); }

__global__ void __launch_bounds__(FA_THREADS, 1)
fa_prefill_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                  const __grid_constant__ CUtensorMap tmVT, FaArgs a) {
    extern __shared__ uint8_t fa_smem_raw[];
    __shared__ uint64_t q_full, kv_full[2], kv_empty[2], s_full[2], s_empty[2], p_full, p_empty, o_full, o_empty;
    __shared__ uint32_t tmem_base_s;
    __shared__ float ex_max[2][2][FA_BQ];   // [block parity][column half][query row]: the halves' row maxima of a key block
    __shared__ float ex_sum[FA_BQ];         // the upper half's row sum, handed to the lower half at the end
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t sbase = (smem_u32(fa_smem_raw) + 1023u) & ~1023u;
    const uint32_t sQ = sbase, sP = sbase + FA_TILE_BYTES, sKV = sbase + 2 * FA_TILE_BYTES;  // stage s: K at sKV + s*64K, V^T 32 KB after it
    const int h = blockIdx.x % a.H, qb = a.nqb - 1 - (int)(blockIdx.x / a.H);  // long (late) query blocks first
    const int q0 = qb * FA_BQ, kvh = h / (a.H / a.KV);
    const int last_q = a.pos0 + min(q0 + FA_BQ, a.T) - 1;   // largest absolute query position of the block
    const int nblk = last_q / FA_BK + 1;                      // key blocks up to the diagonal

    if (tid == 0) {
        mbar_init(&q_full, 1);
        for (int s = 0; s < 2; ++s) { mbar_init(&kv_full[s], 1); mbar_init(&kv_empty[s], 1); mbar_init(&s_full[s], 1); mbar_init(&s_empty[s], 32 * FA_SOFT_WARPS); }
        mbar_init(&p_full, 32 * FA_SOFT_WARPS); mbar_init(&p_empty, 1); mbar_init(&o_full, 1); mbar_init(&o_empty, 32 * FA_SOFT_WARPS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;  // columns: S buffers at 0 and 128, O_j at 256

    if (warp == 0) {
        // ===== TMA producer
        if (lane == 0) {
            rg_prefetch_map(&tmQ); rg_prefetch_map(&tmK); rg_prefetch_map(&tmVT);
            mbar_expect_tx(&q_full, FA_TILE_BYTES);
            fa_tma_3d(sQ, &tmQ, smem_u32(&q_full), 0, q0, h);
            fa_tma_3d(sQ + 16384, &tmQ, smem_u32(&q_full), 64, q0, h);
            for (int j = 0; j < nblk; ++j) {
                const int s = j & 1;
                mbar_wait(&kv_empty[s], (((uint32_t)j >> 1) & 1u) ^ 1u);
                const uint32_t st = sKV + (uint32_t)s * 2 * FA_TILE_BYTES, bar = smem_u32(&kv_full[s]);
                mbar_expect_tx(&kv_full[s], 2 * FA_TILE_BYTES);
                fa_tma_3d(st, &tmK, bar, 0, j * FA_BK, a.slot * a.KV + kvh);                 // keys x head_dim 0-63
                fa_tma_3d(st + 16384, &tmK, bar, 64, j * FA_BK, a.slot * a.KV + kvh);        //      x head_dim 64-127
                fa_tma_3d(st + FA_TILE_BYTES, &tmVT, bar, j * FA_BK, 0, kvh);                // head_dim rows x keys 0-63 of the block
                fa_tma_3d(st + FA_TILE_BYTES + 16384, &tmVT, bar, j * FA_BK + 64, 0, kvh);   //               x keys 64-127
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer
        constexpr uint32_t IDESC = tc_idesc_bf16(128);
        auto issue_qk = [&](int j) {  // S[j & 1] = Q . K_j^T
            const uint32_t st = sKV + (uint32_t)(j & 1) * 2 * FA_TILE_BYTES, d = tmem + (uint32_t)(j & 1) * 128;
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {
                const uint32_t off = (uint32_t)(kk >> 2) * 16384u;
                tc_mma_bf16(d, tc_smem_desc(sQ + off) + 2 * (kk & 3), tc_smem_desc(st + off) + 2 * (kk & 3), IDESC, kk > 0 ? 1u : 0u);
            }
            tc_commit(&s_full[j & 1]);
        };
        mbar_wait(&q_full, 0);
        mbar_wait(&kv_full[0], 0);
        tc_fence_after();
        if (lane == 0) issue_qk(0);
        __syncwarp();
        for (int j = 0; j < nblk; ++j) {
            if (j + 1 < nblk) {  // the next block's scores run under this block's softmax
                const int s1 = (j + 1) & 1;
                mbar_wait(&kv_full[s1], ((uint32_t)(j + 1) >> 1) & 1u);
                mbar_wait(&s_empty[s1], ((((uint32_t)(j + 1)) >> 1) & 1u) ^ 1u);
                tc_fence_after();
                if (lane == 0) issue_qk(j + 1);
                __syncwarp();
            }
            mbar_wait(&p_full, (uint32_t)j & 1u);
            mbar_wait(&o_empty, ((uint32_t)j & 1u) ^ 1u);
            tc_fence_after();
            if (lane == 0) {  // O_j = P . V_j (fresh accumulator)
                const uint32_t st = sKV + (uint32_t)(j & 1) * 2 * FA_TILE_BYTES + FA_TILE_BYTES, d = tmem + 256;
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {
                    const uint32_t off = (uint32_t)(kk >> 2) * 16384u;
                    tc_mma_bf16(d, tc_smem_desc(sP + off) + 2 * (kk & 3), tc_smem_desc(st + off) + 2 * (kk & 3), IDESC, kk > 0 ? 1u : 0u);
                }
                tc_commit(&o_full);
                tc_commit(&kv_empty[j & 1]);
                tc_commit(&p_empty);
            }
            __syncwarp();
        }
    } else {
        // ===== softmax + output: thread = query row r of the block (TMEM lane r) x one half of the columns.  Four warps (one
        // per scheduler) could not keep up with the tensor pipe: ~3000 dependent instructions per key block and thread left it
        // 11 % busy; eight warps halve the chain, the two halves of a row meet once per block (row maximum, through shared
        // memory) and once at the end (row sum).
        const int r = (warp & 3) * 32 + lane, hf = (warp - 2) >> 2;
        const uint32_t tlane = (uint32_t)((warp & 3) * 32) << 16;
        const int q_abs = a.pos0 + q0 + r;
        const int c0 = hf * 64;                      // first column (key inside the block / head dim) of this thread's half
        float m = -INFINITY, l = 0.f, corr_prev = 0.f;
        float O[64];
#pragma unroll
        for (int i = 0; i < 64; ++i) O[i] = 0.f;
        for (int j = 0; j < nblk; ++j) {
            const uint32_t sb = (uint32_t)j & 1u;
            mbar_wait(&s_full[sb], ((uint32_t)j >> 1) & 1u);
            tc_fence_after();
            const uint32_t ts = tmem + tlane + sb * 128 + (uint32_t)c0;
            const int kbase = j * FA_BK + c0;
            const bool need_mask = j * FA_BK + FA_BK - 1 > a.pos0 + q0;  // uniform: the block reaches past the first query of the tile
            uint32_t u0[32], u1[32];
            tc_ld32_nowait(ts, u0);
            tc_ld32_nowait(ts + 32, u1);
            tc_ld_wait();
            float mx = -INFINITY;
            if (need_mask) {   // only the blocks on the diagonal pay for the per-element comparisons
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const float s0 = (kbase + i > q_abs) ? -INFINITY : __uint_as_float(u0[i]);
                    const float s1 = (kbase + 32 + i > q_abs) ? -INFINITY : __uint_as_float(u1[i]);
                    mx = fmaxf(mx, fmaxf(s0, s1));
                }
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) mx = fmaxf(mx, fmaxf(__uint_as_float(u0[i]), __uint_as_float(u1[i])));
            }
            ex_max[sb][hf][r] = mx;
            fa_soft_bar();
            mx = fmaxf(mx, ex_max[sb][hf ^ 1][r]);
            const float m_new = fmaxf(m, mx * a.scale_log2);
            const float corr = exp2f(m - m_new);  // 0 for the first block (m = -inf)
            if (j > 0) {
                // the previous block's P.V ran under this block's maximum pass: fold its result in now (its own factor), which
                // also means that P.V has read P and the P tile may be overwritten
                mbar_wait(&o_full, (uint32_t)(j - 1) & 1u);
                tc_fence_after();
                uint32_t o0[32], o1[32];
                tc_ld32_nowait(tmem + tlane + 256 + (uint32_t)c0, o0);
                tc_ld32_nowait(tmem + tlane + 256 + (uint32_t)c0 + 32, o1);
                tc_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    O[i] = fmaf(O[i], corr_prev, __uint_as_float(o0[i]));
                    O[32 + i] = fmaf(O[32 + i], corr_prev, __uint_as_float(o1[i]));
                }
                tc_fence_before();
                mbar_arrive(&o_empty);
                mbar_wait(&p_empty, ((uint32_t)(j - 1)) & 1u);
            }
            tc_ld32_nowait(ts, u0);      // the scores again (keeping them in registers across the update above would spill)
            tc_ld32_nowait(ts + 32, u1);
            tc_ld_wait();
            float lsum = 0.f;
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                const uint32_t (&v)[32] = c ? u1 : u0;
                uint32_t w[16];
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    float p0 = fa_ex2(fmaf(__uint_as_float(v[i]), a.scale_log2, -m_new));
                    float p1 = fa_ex2(fmaf(__uint_as_float(v[i + 1]), a.scale_log2, -m_new));
                    if (need_mask) {
                        p0 = (kbase + c * 32 + i > q_abs) ? 0.f : p0;
                        p1 = (kbase + c * 32 + i + 1 > q_abs) ? 0.f : p1;
                    }
                    const __nv_bfloat162 pb = __floats2bfloat162_rn(p0, p1);     // one two-way convert; lo = p0
                    const float2 pf = __bfloat1622float2(pb);
                    lsum += pf.x + pf.y;                                         // the sum of what P.V will use
                    w[i >> 1] = *reinterpret_cast<const uint32_t*>(&pb);
                }
                // swizzled K-major operand layout: 16-byte unit u of row r of a 64-key box sits at unit u ^ (r & 7);
                // this half's 64 keys are box `hf`
                const uint32_t box = sP + (uint32_t)hf * 16384u + (uint32_t)r * 128u;
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const uint32_t unit = (uint32_t)(c * 4 + u) ^ (uint32_t)(r & 7);
                    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(box + (unit << 4)), "r"(w[4 * u]), "r"(w[4 * u + 1]), "r"(w[4 * u + 2]), "r"(w[4 * u + 3]) : "memory");
                }
            }
            tc_fence_before();
            mbar_arrive(&s_empty[sb]);                                   // S[sb] may be overwritten
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // P visible to the tensor core's reads
            mbar_arrive(&p_full);
            l = l * corr + lsum;
            m = m_new;
            corr_prev = corr;
        }
        {   // the last block's P.V
            mbar_wait(&o_full, (uint32_t)(nblk - 1) & 1u);
            tc_fence_after();
            uint32_t o0[32], o1[32];
            tc_ld32_nowait(tmem + tlane + 256 + (uint32_t)c0, o0);
            tc_ld32_nowait(tmem + tlane + 256 + (uint32_t)c0 + 32, o1);
            tc_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                O[i] = fmaf(O[i], corr_prev, __uint_as_float(o0[i]));
                O[32 + i] = fmaf(O[32 + i], corr_prev, __uint_as_float(o1[i]));
            }
            tc_fence_before();
            mbar_arrive(&o_empty);
        }
        if (hf == 1) ex_sum[r] = l;
        fa_soft_bar();
        if (hf == 0) l += ex_sum[r];
        else l += 0.f;
        if (hf == 0) ex_sum[r] = l;          // the lower half publishes the full sum for its partner
        fa_soft_bar();
        l = ex_sum[r];
        if (q0 + r < a.T) {
            const float inv = 1.f / l;
            uint4* dst = reinterpret_cast<uint4*>(a.y + ((size_t)(q0 + r) * a.H + h) * FA_HD + c0);
#pragma unroll
            for (int u = 0; u < 8; ++u)
                dst[u] = make_uint4(pack2(O[8 * u] * inv, O[8 * u + 1] * inv), pack2(O[8 * u + 2] * inv, O[8 * u + 3] * inv),
                                    pack2(O[8 * u + 4] * inv, O[8 * u + 5] * inv), pack2(O[8 * u + 6] * inv, O[8 * u + 7] * inv));
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
    }
}
