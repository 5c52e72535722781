// Attention kernels for the Dual-AR decode path.
//
//  * embed_kernel          : llama.py:400-420  (token + masked codebook-embedding sum, 1/sqrt(C+1) scale)
//  * rope_append_kernel    : llama.py:900-905,862-863,205-214 (k-norm, RoPE, KV-cache scatter) for prompt rows
//  * attn_decode_kernel    : llama.py:883-945 slow-AR attention for one query row per (request|prompt token):
//                            q-norm + RoPE fused, optional fused K/V append, GQA-aware split-KV flash-decode over
//                            keys [0,pos] ONLY (the reference attends over the whole allocated cache with a mask),
//                            fp32 softmax like the MATH SDPA backend the reference forces (inference.py:210),
//                            partials merged by the last CTA to arrive (no second launch).
//  * fast_attn_kernel      : llama.py:947-975 fast-AR attention, context <= num_codebooks, every op rounded to bf16
//                            exactly where eq_scaled_dot_product_attention rounds.
#pragma once
#include "common.cuh"

// ----------------------------------------------------------------------------- embed (K1)
struct EmbedArgs {
    const int32_t* tokens;   // [M][C+1]
    const bf16* emb;         // [V][D]
    const bf16* cb_emb;      // [C*cb][D]
    bf16* x;                 // [M][D]
    uint32_t* x_t;           // flag-in-data copy of x for the persistent frame kernel (written instead of x), or nullptr
    int D, C, cb, sem_begin, sem_end, scale;
    float inv_div;           // sqrt(C+1) as float (the reference divides)
    const void* pf_ptr; size_t pf_bytes;
};

__device__ __forceinline__ void embed_body(const EmbedArgs& a, int m, int tid, int nt, unsigned tag = 0) {
    const int32_t* t = a.tokens + (size_t)m * (a.C + 1);
    const int tok = lda(&t[0]);
    const bool sem = tok >= a.sem_begin && tok <= a.sem_end;
    const int DC = a.D >> 3;
    for (int c = tid; c < DC; c += nt) {
        float e[8], s[8];
        unpack8(reinterpret_cast<const uint4*>(a.emb + (size_t)tok * a.D)[c], e);
#pragma unroll
        for (int i = 0; i < 8; ++i) s[i] = 0.f;
        if (sem) {
            for (int k = 0; k < a.C; ++k) {
                float f[8];
                const size_t row = (size_t)lda(&t[k + 1]) + (size_t)k * a.cb;
                unpack8(reinterpret_cast<const uint4*>(a.cb_emb + row * a.D)[c], f);
#pragma unroll
                for (int i = 0; i < 8; ++i) s[i] += f[i];
            }
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            float v = rbf(e[i] + rbf(s[i]));                      // stack().sum() rounds once; + rounds again
            if (sem && a.scale) v = rbf(__fdiv_rn(v, a.inv_div)); // x / sqrt(C+1)
            e[i] = v;
        }
        uint4 o;
        o.x = pack2(e[0], e[1]); o.y = pack2(e[2], e[3]); o.z = pack2(e[4], e[5]); o.w = pack2(e[6], e[7]);
        if (a.x_t != nullptr) {
            uint32_t* d = a.x_t + (size_t)m * a.D + (size_t)c * 8;
            st_tagged4(d, o.x, o.y, tag);
            st_tagged4(d + 4, o.z, o.w, tag);
        } else {
            reinterpret_cast<uint4*>(a.x + (size_t)m * a.D)[c] = o;
        }
    }
}
__global__ void __launch_bounds__(256) embed_kernel(EmbedArgs a) {
    if (threadIdx.x == 0) l2_prefetch_share(a.pf_ptr, a.pf_bytes, blockIdx.x, gridDim.x);
    pdl_wait();
    pdl_launch_dependents();
    embed_body(a, blockIdx.x, threadIdx.x, blockDim.x);
}

// ----------------------------------------------------------------------------- head prep: per-head RMSNorm (single rounding) + RoPE
// One warp per head vector; lane owns EPL = HD/32 contiguous elements (whole interleaved pairs).
// F.rms_norm on bf16 computes x*rsqrt(mean(x^2)+eps)*w in fp32 and rounds once [probe]; RoPE multiplies
// fp32 values with the bf16 table and rounds once (llama.py:1025-1037); products are not fused (the
// reference issues separate mul/sub ops).
// this lane's EPL elements of one head vector: plain bf16 (through L2) or flag-in-data words (spins until current)
template <int HD, bool TAG>
__device__ __forceinline__ void load_head(const bf16* src, const uint32_t* src_t, unsigned tag, float (&v)[HD / 32]) {
    constexpr int EPL = HD / 32;
    const int lane = threadIdx.x & 31;
    if (TAG) {
        if (EPL == 4) {
            const uint2 pk = ld_tagged4(src_t + lane * 4, tag);
            v[0] = bf16_lo(pk.x); v[1] = bf16_hi(pk.x); v[EPL - 2] = bf16_lo(pk.y); v[EPL - 1] = bf16_hi(pk.y);
        } else {
#pragma unroll
            for (int i = 0; i < EPL; i += 2) {
                const uint32_t pk = ld_tagged2(src_t + lane * EPL + i, tag);
                v[i] = bf16_lo(pk); v[i + 1] = bf16_hi(pk);
            }
        }
    } else {
#pragma unroll
        for (int i = 0; i < EPL; ++i) v[i] = lda(&src[lane * EPL + i]);
    }
}

// this lane's EPL elements of the head that starts at column col0 of row m of qkv, from the partial tiles of the wqkv GEMM
template <int HD>
__device__ __forceinline__ void load_head_part(const float* part, int s, int bn, int m, int col0, float (&v)[HD / 32]) {
    constexpr int EPL = HD / 32;
    static_assert(EPL == 4 || EPL == 2, "a lane's elements are one 16- or 8-byte piece of a partial-tile row");
    const int n = col0 + (threadIdx.x & 31) * EPL, t = n >> 7, r = n & 127;
#pragma unroll
    for (int i = 0; i < EPL; ++i) v[i] = 0.f;
    for (int g = 0; g < s; ++g) {  // CTA order
        const float* p = part + ((size_t)((t * s + g) * 2) * bn + m) * 128 + r;
        if (EPL == 4) {
            const float4 x = __ldcg(reinterpret_cast<const float4*>(p));
            v[0] += x.x; v[1] += x.y; v[EPL - 2] += x.z; v[EPL - 1] += x.w;
        } else {
            const float2 x = __ldcg(reinterpret_cast<const float2*>(p));
            v[0] += x.x; v[1] += x.y;
        }
    }
#pragma unroll
    for (int i = 0; i < EPL; ++i) v[i] = rbf(v[i]);
}

// norm weights and RoPE (cos, sin) of this lane's elements, fetched ahead of the data they will be applied to
template <int HD>
struct HeadCoef {
    float w[HD / 32], c[HD / 64], s[HD / 64];
    __device__ __forceinline__ void load(const bf16* norm_w, const bf16* rope_row, bool qk_norm) {
        constexpr int EPL = HD / 32;
        const int lane = threadIdx.x & 31;
#pragma unroll
        for (int i = 0; i < EPL; ++i) w[i] = qk_norm ? __bfloat162float(norm_w[lane * EPL + i]) : 1.f;
#pragma unroll
        for (int i = 0; i < EPL; i += 2) {
            const int pair = (lane * EPL + i) >> 1;
            c[i >> 1] = __bfloat162float(rope_row[pair * 2 + 0]);
            s[i >> 1] = __bfloat162float(rope_row[pair * 2 + 1]);
        }
    }
};
template <int HD>
__device__ __forceinline__ void head_norm_rope(const float (&in)[HD / 32], const HeadCoef<HD>& k, bool qk_norm, float eps,
                                               float (&out)[HD / 32]) {
    constexpr int EPL = HD / 32;
    float v[EPL];
#pragma unroll
    for (int i = 0; i < EPL; ++i) v[i] = in[i];
    if (qk_norm) {
        float ss = 0.f;
#pragma unroll
        for (int i = 0; i < EPL; ++i) ss += v[i] * v[i];
        ss = warp_sum(ss);
        const float rs = __frsqrt_rn(__fadd_rn(__fdiv_rn(ss, (float)HD), eps));
#pragma unroll
        for (int i = 0; i < EPL; ++i) v[i] = rbf(__fmul_rn(__fmul_rn(v[i], rs), k.w[i]));
    }
#pragma unroll
    for (int i = 0; i < EPL; i += 2) {
        const float c = k.c[i >> 1], s = k.s[i >> 1];
        const float x0 = v[i], x1 = v[i + 1];
        out[i] = rbf(__fsub_rn(__fmul_rn(x0, c), __fmul_rn(x1, s)));
        out[i + 1] = rbf(__fadd_rn(__fmul_rn(x1, c), __fmul_rn(x0, s)));
    }
}

template <int HD>
__device__ __forceinline__ void head_norm_rope(const float (&in)[HD / 32], const bf16* norm_w, const bf16* rope_row,
                                               bool qk_norm, float eps, float (&out)[HD / 32]) {
    constexpr int EPL = HD / 32;
    const int lane = threadIdx.x & 31;
    float v[EPL];
#pragma unroll
    for (int i = 0; i < EPL; ++i) v[i] = in[i];
    if (qk_norm) {
        float ss = 0.f;
#pragma unroll
        for (int i = 0; i < EPL; ++i) ss += v[i] * v[i];
        ss = warp_sum(ss);
        const float rs = __frsqrt_rn(__fadd_rn(__fdiv_rn(ss, (float)HD), eps));
#pragma unroll
        for (int i = 0; i < EPL; ++i)
            v[i] = rbf(__fmul_rn(__fmul_rn(v[i], rs), __bfloat162float(norm_w[lane * EPL + i])));
    }
#pragma unroll
    for (int i = 0; i < EPL; i += 2) {
        const int pair = (lane * EPL + i) >> 1;
        const float c = __bfloat162float(rope_row[pair * 2 + 0]);
        const float s = __bfloat162float(rope_row[pair * 2 + 1]);
        const float x0 = v[i], x1 = v[i + 1];
        out[i] = rbf(__fsub_rn(__fmul_rn(x0, c), __fmul_rn(x1, s)));
        out[i + 1] = rbf(__fadd_rn(__fmul_rn(x1, c), __fmul_rn(x0, s)));
    }
}

struct AttnArgs {
    const bf16* qkv;        // [M][(H+2KV)*HD]
    bf16* kcache;           // [maxB][KV][S][HD]
    bf16* vcache;
    const bf16* q_norm;     // [HD]
    const bf16* k_norm;     // [HD]
    const bf16* rope;       // [S][HD/2][2]
    const int32_t* pos;     // [M] position of each row (device)
    const int32_t* slots;   // [M] kv slot per row, or nullptr -> slot = row
    bf16* y;                // [M][H*HD]
    const uint32_t* qkv_t;  // flag-in-data versions of qkv / y used by the persistent frame kernel (else nullptr)
    uint32_t* y_t;
    int tag_in_op, tag_out_op;  // op indices whose tags qkv_t carries / y_t gets
    uint2* ll_acc;          // [M][KV][splits][GQ*HD] {fp32, tag} partial numerators of the frame kernel's split attention
    uint2* ll_ml;           // [M][KV][splits][GQ][2]  {max, sum}
    const bf16 *next_k, *next_v;  // KV cache of the next slow layer (L2 prefetch one layer ahead), or nullptr
    float* ws_acc;          // [M][KV][splits][GQ][HD]
    float* ws_ml;           // [M][KV][splits][GQ][2]
    int32_t* counters;      // [M][KV]
    int H, KV, S, splits, qk_norm, fused_append, pos_fixed;  // pos_fixed >= 0 overrides pos[] (fast AR)
    int tc_u1, tc_s2;       // attn_tc_kernel with a 1-D grid: (row, head) units [0, tc_u1) are cut into `splits` items, the rest into tc_s2
    float eps, scale;
    const void* pf_ptr; size_t pf_bytes;
    unsigned long long* trace;  // profiling aid (attn_tc_kernel): 4 clock64 stamps per CTA, or nullptr
    // many-row steps: qkv left by rows_gemm_kernel as fp32 partial tiles (tile-aligned split: tile t = CTAs t*s .. t*s+s-1, slot 0),
    // summed in CTA order and rounded to bf16 (one rounding per Linear) by the kernels that read q / k / v; nullptr: read `qkv`
    const float* qkv_part;
    int qkv_s, qkv_bn;
};

// position of row m, kept inside the bound cache: a position past the end (the reference raises an index error there,
// llama.py:205-214; the host mirror raises ValueError before the launch) must never scatter K/V outside the buffer
__device__ __forceinline__ int attn_pos(const AttnArgs& a, int m) { return max(0, min(a.pos[m], a.S - 1)); }

// ----------------------------------------------------------------------------- prompt rows: k-norm + RoPE + scatter
template <int HD>
__global__ void __launch_bounds__(64) rope_append_kernel(AttnArgs a) {
    if (threadIdx.x == 0)
        l2_prefetch_share(a.pf_ptr, a.pf_bytes, blockIdx.y * gridDim.x + blockIdx.x, gridDim.x * gridDim.y);
    pdl_wait();
    pdl_launch_dependents();
    constexpr int EPL = HD / 32;
    const int h = blockIdx.x, m = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int pos = attn_pos(a, m);
    const int slot = a.slots ? a.slots[m] : m;
    const int qkv_w = (a.H + 2 * a.KV) * HD;
    const size_t co = (((size_t)slot * a.KV + h) * a.S + pos) * HD;
    if (warp == 0) {
        float kin[EPL], k[EPL];
        if (a.qkv_part) load_head_part<HD>(a.qkv_part, a.qkv_s, a.qkv_bn, m, (a.H + h) * HD, kin);
        else load_head<HD, false>(a.qkv + (size_t)m * qkv_w + (a.H + h) * HD, nullptr, 0, kin);
        head_norm_rope<HD>(kin, a.k_norm, a.rope + (size_t)pos * HD, a.qk_norm, a.eps, k);
#pragma unroll
        for (int i = 0; i < EPL; ++i) a.kcache[co + lane * EPL + i] = __float2bfloat16_rn(k[i]);
    } else if (a.qkv_part) {
        float vin[EPL];
        load_head_part<HD>(a.qkv_part, a.qkv_s, a.qkv_bn, m, (a.H + a.KV + h) * HD, vin);
#pragma unroll
        for (int i = 0; i < EPL; ++i) a.vcache[co + lane * EPL + i] = __float2bfloat16_rn(vin[i]);
    } else {
        const bf16* v = a.qkv + (size_t)m * qkv_w + (a.H + a.KV + h) * HD;
#pragma unroll
        for (int i = 0; i < EPL; ++i) a.vcache[co + lane * EPL + i] = lda_raw(&v[lane * EPL + i]);
    }
}

// ----------------------------------------------------------------------------- slow-AR decode attention
#define ATTN_THREADS 128
#define ATTN_TILE 128
#define ATTN_UNROLL 4  // key rows in flight per thread (16 B each): the loop is latency-bound without them

// one (split, kv head, row) work item on ATTN_THREADS threads synchronised through SY; tid = index inside the group
// QREG: keep q in registers (standalone kernel) or read it from shared memory in the FMA loop (persistent frame kernel,
// which is capped at 96 registers per thread by its 17 warps per SM)
template <int HD, int GQ, class SY, bool QREG = true, bool TAG = false>
__device__ __forceinline__ void attn_decode_body(const AttnArgs& a, int split, int h, int m, int tid, unsigned tag_in = 0,
                                                 unsigned tag_out = 0) {
    constexpr int EPL = HD / 32;
    constexpr int LPK = HD / 8;              // lanes per key row (16 B each)
    constexpr int KPI = ATTN_THREADS / LPK;  // keys per CTA step
    __shared__ float qs[GQ][HD];
    __shared__ float knew[HD], vnew[HD];
    __shared__ float sc[GQ][ATTN_TILE];
    __shared__ float accs[KPI][GQ][HD + 4];
    __shared__ float mrun[GQ], lrun[GQ], corr[GQ];
    __shared__ int is_last;

    const int lane = tid & 31, warp = tid >> 5;

    const int pos = attn_pos(a, m);
    const int slot = a.slots ? a.slots[m] : m;
    const int qkv_w = (a.H + 2 * a.KV) * HD;
    const bf16* row = a.qkv + (size_t)m * qkv_w;
    const uint32_t* row_t = TAG ? a.qkv_t + (size_t)m * qkv_w : nullptr;
    const int n_keys = pos + 1;
    const int per = (n_keys + a.splits - 1) / a.splits;
    const int k0 = split * per;
    const int k1 = min(n_keys, k0 + per);
    const bool owns_new = a.fused_append && (pos >= k0) && (pos < k1);
    const bf16* rope_row = a.rope + (size_t)pos * HD;

    // ---- q prep (one warp per q head of this GQA group), optional fused K/V append
    for (int g = warp; g < GQ; g += ATTN_THREADS / 32) {
        float qin[EPL], q[EPL];
        load_head<HD, TAG>(row + (h * GQ + g) * HD, row_t + (h * GQ + g) * HD, tag_in, qin);
        head_norm_rope<HD>(qin, a.q_norm, rope_row, a.qk_norm, a.eps, q);
#pragma unroll
        for (int i = 0; i < EPL; ++i) qs[g][lane * EPL + i] = q[i];
    }
    if (owns_new) {
        const size_t co = (((size_t)slot * a.KV + h) * a.S + pos) * HD;
        if (warp == (GQ % (ATTN_THREADS / 32))) {  // spread the extra work to a different warp when possible
            float kin[EPL], k[EPL];
            load_head<HD, TAG>(row + (a.H + h) * HD, row_t + (a.H + h) * HD, tag_in, kin);
            head_norm_rope<HD>(kin, a.k_norm, rope_row, a.qk_norm, a.eps, k);
#pragma unroll
            for (int i = 0; i < EPL; ++i) {
                knew[lane * EPL + i] = k[i];
                a.kcache[co + lane * EPL + i] = __float2bfloat16_rn(k[i]);
            }
        }
        if (warp == ((GQ + 1) % (ATTN_THREADS / 32))) {
            float vin[EPL];
            load_head<HD, TAG>(row + (a.H + a.KV + h) * HD, row_t + (a.H + a.KV + h) * HD, tag_in, vin);
#pragma unroll
            for (int i = 0; i < EPL; ++i) {
                vnew[lane * EPL + i] = vin[i];
                a.vcache[co + lane * EPL + i] = __float2bfloat16_rn(vin[i]);  // exact: vin is a bf16 value
            }
        }
    }
    if (tid < GQ) { mrun[tid] = -INFINITY; lrun[tid] = 0.f; }
    SY::sync();

    const int kslot = tid / LPK, chunk = tid % LPK;
    float qreg[QREG ? GQ : 1][8];
    if (QREG) {
#pragma unroll
        for (int g = 0; g < GQ; ++g)
#pragma unroll
            for (int i = 0; i < 8; ++i) qreg[g][i] = qs[g][chunk * 8 + i];
    }
    float acc[GQ][8];
#pragma unroll
    for (int g = 0; g < GQ; ++g)
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[g][i] = 0.f;

    const bf16* kbase = a.kcache + ((size_t)slot * a.KV + h) * a.S * HD;
    const bf16* vbase = a.vcache + ((size_t)slot * a.KV + h) * a.S * HD;
    // keys read from the cache: [k0, kc1); the new key (if owned) comes from shared memory
    const int kc1 = owns_new ? min(k1, pos) : k1;

    for (int t0 = k0; t0 < k1; t0 += ATTN_TILE) {
        const int t1 = min(k1, t0 + ATTN_TILE);
        // -- scores for this tile: ATTN_UNROLL key rows are requested per thread before the first one is used
        for (int jb = t0; jb < t1; jb += KPI * ATTN_UNROLL) {  // warp-uniform trip count (shuffles below)
            uint4 kraw[ATTN_UNROLL];
#pragma unroll
            for (int u = 0; u < ATTN_UNROLL; ++u) {
                const int j = jb + u * KPI + kslot;
                kraw[u] = (j < t1 && j < kc1) ? ld_stream(kbase + (size_t)j * HD + chunk * 8) : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int u = 0; u < ATTN_UNROLL; ++u) {
                const int j = jb + u * KPI + kslot;
                const bool jv = j < t1;
                float kf[8];
                if (jv && j < kc1) {
                    unpack8(kraw[u], kf);
                } else {
#pragma unroll
                    for (int i = 0; i < 8; ++i) kf[i] = jv ? knew[chunk * 8 + i] : 0.f;
                }
                float d[GQ];
#pragma unroll
                for (int g = 0; g < GQ; ++g) {
                    float s = 0.f;
#pragma unroll
                    for (int i = 0; i < 8; ++i) s = fmaf(QREG ? qreg[QREG ? g : 0][i] : qs[g][chunk * 8 + i], kf[i], s);
                    d[g] = s;
                }
#pragma unroll
                for (int o = LPK / 2; o > 0; o >>= 1)
#pragma unroll
                    for (int g = 0; g < GQ; ++g) d[g] += __shfl_xor_sync(0xffffffffu, d[g], o);
                if (chunk == 0 && jv)
#pragma unroll
                    for (int g = 0; g < GQ; ++g) sc[g][j - t0] = d[g] * a.scale;
            }
        }
        SY::sync();
        // -- online softmax bookkeeping: warp g owns head g
        for (int g = warp; g < GQ; g += ATTN_THREADS / 32) {
            float mx = -INFINITY;
            for (int j = lane; j < t1 - t0; j += 32) mx = fmaxf(mx, sc[g][j]);
            mx = warp_max(mx);
            const float mold = mrun[g];
            const float mnew = fmaxf(mold, mx);
            float sum = 0.f;
            for (int j = lane; j < t1 - t0; j += 32) {
                const float p = expf(sc[g][j] - mnew);
                sc[g][j] = p;
                sum += p;
            }
            sum = warp_sum(sum);
            if (lane == 0) {
                const float c = (mold == -INFINITY) ? 0.f : expf(mold - mnew);
                corr[g] = c;
                mrun[g] = mnew;
                lrun[g] = lrun[g] * c + sum;
            }
        }
        SY::sync();
        // -- rescale and accumulate P.V
#pragma unroll
        for (int g = 0; g < GQ; ++g) {
            const float c = corr[g];
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[g][i] *= c;
        }
        for (int jb = t0; jb < t1; jb += KPI * ATTN_UNROLL) {
            uint4 vraw[ATTN_UNROLL];
#pragma unroll
            for (int u = 0; u < ATTN_UNROLL; ++u) {
                const int j = jb + u * KPI + kslot;
                vraw[u] = (j < t1 && j < kc1) ? ld_stream(vbase + (size_t)j * HD + chunk * 8) : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int u = 0; u < ATTN_UNROLL; ++u) {
                const int j = jb + u * KPI + kslot;
                if (j >= t1) continue;
                float vf[8];
                if (j < kc1) {
                    unpack8(vraw[u], vf);
                } else {
#pragma unroll
                    for (int i = 0; i < 8; ++i) vf[i] = vnew[chunk * 8 + i];
                }
#pragma unroll
                for (int g = 0; g < GQ; ++g) {
                    const float p = sc[g][j - t0];
#pragma unroll
                    for (int i = 0; i < 8; ++i) acc[g][i] = fmaf(p, vf[i], acc[g][i]);
                }
            }
        }
        SY::sync();
    }

    // ---- reduce acc over the KPI key slots
#pragma unroll
    for (int g = 0; g < GQ; ++g)
#pragma unroll
        for (int i = 0; i < 8; ++i) accs[kslot][g][chunk * 8 + i] = acc[g][i];
    SY::sync();

    const int H = a.H;
    if (a.splits == 1) {
        for (int e = tid; e < GQ * HD; e += ATTN_THREADS) {
            const int g = e / HD, d = e % HD;
            float s = 0.f;
#pragma unroll
            for (int k = 0; k < KPI; ++k) s += accs[k][g][d];
            const size_t yi = (size_t)m * H * HD + (h * GQ + g) * HD + d;
            if (TAG) st_tagged(a.y_t + yi, s / lrun[g], tag_out);
            else a.y[yi] = __float2bfloat16_rn(s / lrun[g]);
        }
        return;
    }
    // ---- write partial, last CTA of this (row, kv head) merges
    const size_t pbase = (((size_t)m * a.KV + h) * a.splits + split);
    for (int e = tid; e < GQ * HD; e += ATTN_THREADS) {
        const int g = e / HD, d = e % HD;
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < KPI; ++k) s += accs[k][g][d];
        a.ws_acc[pbase * GQ * HD + e] = s;
    }
    if (tid < GQ) {
        a.ws_ml[(pbase * GQ + tid) * 2 + 0] = mrun[tid];
        a.ws_ml[(pbase * GQ + tid) * 2 + 1] = lrun[tid];
    }
    __threadfence();
    SY::sync();
    if (tid == 0) {
        const int prev = atomicAdd(&a.counters[m * a.KV + h], 1);
        is_last = (prev == a.splits - 1);
    }
    SY::sync();
    if (!is_last) return;
    __threadfence();
    const size_t b0 = ((size_t)m * a.KV + h) * a.splits;
    for (int e = tid; e < GQ * HD; e += ATTN_THREADS) {
        const int g = e / HD;
        float M = -INFINITY;
        for (int s = 0; s < a.splits; ++s) M = fmaxf(M, __ldcg(&a.ws_ml[((b0 + s) * GQ + g) * 2]));
        float num = 0.f, den = 0.f;
        for (int s = 0; s < a.splits; ++s) {
            const float ms = __ldcg(&a.ws_ml[((b0 + s) * GQ + g) * 2]);
            const float ls = __ldcg(&a.ws_ml[((b0 + s) * GQ + g) * 2 + 1]);
            const float w = (ms == -INFINITY) ? 0.f : expf(ms - M);
            num = fmaf(w, __ldcg(&a.ws_acc[(b0 + s) * GQ * HD + e]), num);
            den = fmaf(w, ls, den);
        }
        const size_t yi = (size_t)m * H * HD + (h * GQ) * HD + e;
        if (TAG) st_tagged(a.y_t + yi, num / den, tag_out);
        else a.y[yi] = __float2bfloat16_rn(num / den);
    }
    if (tid == 0) a.counters[m * a.KV + h] = 0;  // re-arm for the next launch
}
template <int HD, int GQ>
__global__ void __launch_bounds__(ATTN_THREADS) attn_decode_kernel(AttnArgs a) {
    if (threadIdx.x == 0)
        l2_prefetch_share(a.pf_ptr, a.pf_bytes, (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x,
                          gridDim.x * gridDim.y * gridDim.z);
    pdl_wait();
    pdl_launch_dependents();
    attn_decode_body<HD, GQ, SyncBlock<ATTN_THREADS>>(a, blockIdx.x, blockIdx.y, blockIdx.z, threadIdx.x);
}

// ----------------------------------------------------------------------------- slow-AR decode attention, many rows
// Same math and the same (split, kv head, row) work item as attn_decode_body, for steps with many requests in flight
// (config 3: 64 rows x 2k keys = 537 MB of cached K/V per layer).  attn_decode_body spends ~61 warp instructions per key
// (fp32 FMAs for 4 query heads plus a 16-lane shuffle reduction per key), which caps it at ~40 % of the HBM rate; here
// Q.K^T and P.V run on the tensor cores (mma.sync.m16n8k16, bf16 x bf16 -> fp32): ~12 instructions per key.
//  * The 4 query heads of a GQA group are rows 0-3 of the 16-row A operand (rows 4-15 are zero).
//  * No shared-memory staging and no ldmatrix: the reduction index of an MMA may be permuted freely as long as both
//    operands agree, so every lane reads whole 16-byte pieces of the K / V rows straight from global memory
//    (ld.global.nc, each byte read once, 64 B or 128 B contiguous per row and instruction):
//      S = Q.K^T : lane (n = lane/4, q = lane%4) holds K[key n][8(4j+q) .. +8), j = 0..3; MMA k-step 2j uses its first four
//                  elements as k = {2q, 2q+1, 2q+8, 2q+9}, k-step 2j+1 the last four; the Q fragments are built with the
//                  same assignment.
//      O += P.V  : lane (g = lane/4, q) holds V[keys 2q, 2q+1, 2q+8, 2q+9][8(g+8j) .. +8), j = 0..1; MMA (j, e) takes element
//                  e of each piece (two PRMTs make its B registers), so its output column g is d = 8(g+8j)+e.
//  * P is split into two bf16 terms (p = hi + lo, 16 mantissa bits) so that P.V keeps fp32-level accuracy like the
//    reference's MATH SDPA (inference.py:210); Q and K are bf16 values, so the products of Q.K^T are exact.
//  * A warp owns every 4th block of 16 keys of the item with its own running (max, sum, O); the four warps are merged
//    through shared memory, then the split partials by the last CTA to arrive (as in attn_decode_body).
//  * The new key / value of the step must already be in the cache (rope_append_kernel runs first).
#define ATC_THREADS 128
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// D rows 0-7 (+)= A rows 0-7 . B; rows 8-15 of A are zero here (only <= 8 query heads), so their accumulators are not kept
__device__ __forceinline__ void mma_bf16_top(float& c0, float& c1, uint32_t a0, uint32_t a2, uint32_t b0, uint32_t b1) {
    asm volatile(
        "{\n\t.reg .f32 z2, z3;\n\t"
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,z2,z3}, {%2,%6,%3,%6}, {%4,%5}, {%0,%1,%7,%7};\n\t}"
        : "+f"(c0), "+f"(c1) : "r"(a0), "r"(a2), "r"(b0), "r"(b1), "r"(0u), "r"(0u), "r"(0u), "f"(0.f));
}
template <int HD, int GQ>
__global__ void __launch_bounds__(ATC_THREADS, 3) attn_tc_kernel(AttnArgs a) {
    static_assert(HD == 128 && GQ <= 8, "tensor-core decode attention: head_dim 128, up to 8 query heads per kv head");
    constexpr int EPL = HD / 32;
    constexpr int NWARP = ATC_THREADS / 32;
    __shared__ float qs[GQ][HD];
    __shared__ float so[NWARP][GQ][HD];   // per-warp numerators
    __shared__ float sm[NWARP][GQ], sl[NWARP][GQ];
    __shared__ int is_last;
    if (threadIdx.x == 0) l2_prefetch_share(a.pf_ptr, a.pf_bytes, blockIdx.x, gridDim.x);
    pdl_wait();
    pdl_launch_dependents();
    // 1-D grid, big items first: units [0, tc_u1) have `splits` items each and fill whole rounds of resident CTAs, the
    // remaining units are cut finer (tc_s2 items) so that the last, partial round is short (64 rows x 8 heads x 2 splits =
    // 1024 equal items on 444 slots would cost three rounds for 2.3 rounds of work)
    int unit, split, nsp;
    {
        const int i = blockIdx.x, n1 = a.tc_u1 * a.splits;
        if (i < n1) { unit = i / a.splits; split = i - unit * a.splits; nsp = a.splits; }
        else { unit = a.tc_u1 + (i - n1) / a.tc_s2; split = (i - n1) % a.tc_s2; nsp = a.tc_s2; }
    }
    const int m = unit / a.KV, h = unit - m * a.KV;
    const int stride = max(a.splits, a.tc_s2);  // partial slots per unit in the workspace
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, q = lane & 3;
    unsigned long long* tr = a.trace ? a.trace + 4 * (size_t)blockIdx.x : nullptr;
    if (tr && tid == 0) tr[0] = clock64();
    const int pos = attn_pos(a, m);
    const int slot = a.slots ? a.slots[m] : m;
    const int qkv_w = (a.H + 2 * a.KV) * HD;
    const bf16* row = a.qkv + (size_t)m * qkv_w;
    const int n_keys = pos + 1;
    const int per = (n_keys + nsp - 1) / nsp;
    const int k0 = split * per;
    const int k1 = min(n_keys, k0 + per);
    const bf16* rope_row = a.rope + (size_t)pos * HD;
    for (int gg = warp; gg < GQ; gg += NWARP) {
        float qin[EPL], qv[EPL];
        if (a.qkv_part) load_head_part<HD>(a.qkv_part, a.qkv_s, a.qkv_bn, m, (h * GQ + gg) * HD, qin);
        else load_head<HD, false>(row + (h * GQ + gg) * HD, nullptr, 0, qin);
        head_norm_rope<HD>(qin, a.q_norm, rope_row, a.qk_norm, a.eps, qv);
#pragma unroll
        for (int i = 0; i < EPL; ++i) qs[gg][lane * EPL + i] = qv[i];
    }
    __syncthreads();
    // Q fragments: k-step 2j -> (a0 | a2) = Q[g][8(4j+q) + {0,1 | 2,3}], k-step 2j+1 -> + {4,5 | 6,7}; rows >= GQ are zero
    uint32_t qa[8][2];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int d0 = 8 * (4 * j + q);
#pragma unroll
        for (int e = 0; e < 4; ++e)
            qa[2 * j + (e >> 1)][e & 1] = (g < GQ) ? pack2(qs[g < GQ ? g : 0][d0 + 2 * e], qs[g < GQ ? g : 0][d0 + 2 * e + 1]) : 0u;
    }
    const bf16* kbase = a.kcache + ((size_t)slot * a.KV + h) * a.S * HD;
    const bf16* vbase = a.vcache + ((size_t)slot * a.KV + h) * a.S * HD;
    // O is accumulated TRANSPOSED: O^T[d][head] = V^T[d][key] . P^T[key][head], V^T as the A operand (all 16 rows are d
    // values), P^T as the B operand - which is exactly the C layout of S = Q.K^T (lane (g, q) holds P[head g][keys 2q, 2q+1]).
    // MMA (j, ee): rows g / g+8 <-> d = 8(g+8j) + ee / + ee + 4, columns 2q, 2q+1 <-> heads.  8 MMAs per 16 keys (x2 for hi/lo).
    float o[2][4][4];
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int e = 0; e < 4; ++e)
#pragma unroll
            for (int i = 0; i < 4; ++i) o[j][e][i] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;  // of head g, replicated in the 4 lanes of the row (l_run: this lane's share)
    if (tr && tid == 0) tr[1] = clock64();

    for (int kb = k0 + warp * 16; kb < k1; kb += NWARP * 16) {
        // ---- loads: K rows kb + g and kb + 8 + g (4 pieces each), V rows kb + {2q, 2q+1, 2q+8, 2q+9} (2 pieces each); all 16
        //      requests of the block are in flight together (requesting block i+1 early was measured slower: spills)
        uint4 kr[2][4], vr[4][2];
#pragma unroll
        for (int nb = 0; nb < 2; ++nb) {
            const int key = kb + nb * 8 + g;
#pragma unroll
            for (int j = 0; j < 4; ++j)
                kr[nb][j] = key < k1 ? ld_stream(kbase + (size_t)key * HD + 8 * (4 * j + q)) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int key = kb + 2 * q + (r & 1) + (r >> 1) * 8;
#pragma unroll
            for (int j = 0; j < 2; ++j)
                vr[r][j] = key < k1 ? ld_stream(vbase + (size_t)key * HD + 8 * (g + 8 * j)) : make_uint4(0, 0, 0, 0);
        }
        // ---- S = Q.K^T for the two 8-key blocks
        float sc[2][4];  // [2], [3]: the unused rows 8-15 (A rows are zero there, they stay 0)
#pragma unroll
        for (int nb = 0; nb < 2; ++nb) {
            sc[nb][0] = sc[nb][1] = sc[nb][2] = sc[nb][3] = 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                mma_bf16_16816(sc[nb], qa[2 * j][0], 0u, qa[2 * j][1], 0u, kr[nb][j].x, kr[nb][j].y);
                mma_bf16_16816(sc[nb], qa[2 * j + 1][0], 0u, qa[2 * j + 1][1], 0u, kr[nb][j].z, kr[nb][j].w);
            }
        }
        // ---- online softmax of row g over the 16 keys (this lane: keys kb + nb*8 + 2q + {0,1})
        float p[2][2];
        float mx = -INFINITY;
#pragma unroll
        for (int nb = 0; nb < 2; ++nb)
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const int key = kb + nb * 8 + 2 * q + i;
                p[nb][i] = key < k1 ? sc[nb][i] * a.scale : -INFINITY;
                mx = fmaxf(mx, p[nb][i]);
            }
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
        const float m_new = fmaxf(m_run, mx);  // finite: key kb of this block is valid for every lane's row
        const float corr = (m_run == -INFINITY) ? 0.f : expf(m_run - m_new);
        float psum = 0.f;
#pragma unroll
        for (int nb = 0; nb < 2; ++nb)
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                p[nb][i] = expf(p[nb][i] - m_new);
                psum += p[nb][i];
            }
        l_run = l_run * corr + psum;
        if (__any_sync(0xffffffffu, m_new != m_run)) {  // this lane's accumulator columns are heads 2q and 2q+1
            const float c0 = __shfl_sync(0xffffffffu, corr, (2 * q) * 4), c1 = __shfl_sync(0xffffffffu, corr, (2 * q + 1) * 4);
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int e = 0; e < 4; ++e) { o[j][e][0] *= c0; o[j][e][1] *= c1; o[j][e][2] *= c0; o[j][e][3] *= c1; }
        }
        m_run = m_new;
        // ---- P = hi + lo (bf16 each); B fragment of P^T: b0 = keys 2q, 2q+1 (block 0), b1 = keys 8+2q, 9+2q (block 1)
        uint32_t ph[2], pl[2];
#pragma unroll
        for (int nb = 0; nb < 2; ++nb) {
            const float h0 = rbf(p[nb][0]), h1 = rbf(p[nb][1]);
            ph[nb] = pack2(h0, h1);
            pl[nb] = pack2(p[nb][0] - h0, p[nb][1] - h1);
        }
        // ---- O^T += V^T.P^T
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const uint32_t va[4] = {vr[0][j].x, vr[0][j].y, vr[0][j].z, vr[0][j].w};  // key 2q
            const uint32_t vb[4] = {vr[1][j].x, vr[1][j].y, vr[1][j].z, vr[1][j].w};  // key 2q+1
            const uint32_t vc[4] = {vr[2][j].x, vr[2][j].y, vr[2][j].z, vr[2][j].w};  // key 2q+8
            const uint32_t vd[4] = {vr[3][j].x, vr[3][j].y, vr[3][j].z, vr[3][j].w};  // key 2q+9
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const uint32_t sel = (e & 1) ? 0x7632u : 0x5410u;  // element e (row g) and e + 4 (row g + 8) of the 8-d piece
                const uint32_t a0 = __byte_perm(va[e >> 1], vb[e >> 1], sel), a1 = __byte_perm(va[2 + (e >> 1)], vb[2 + (e >> 1)], sel);
                const uint32_t a2 = __byte_perm(vc[e >> 1], vd[e >> 1], sel), a3 = __byte_perm(vc[2 + (e >> 1)], vd[2 + (e >> 1)], sel);
                mma_bf16_16816(o[j][e], a0, a1, a2, a3, ph[0], ph[1]);
                mma_bf16_16816(o[j][e], a0, a1, a2, a3, pl[0], pl[1]);
            }
        }
    }
    // ---- merge the warps: row sums over the 4 lanes of a row, then (max, sum, O) through shared memory
    if (tr && tid == 0) tr[2] = clock64();
    l_run += __shfl_xor_sync(0xffffffffu, l_run, 1);
    l_run += __shfl_xor_sync(0xffffffffu, l_run, 2);
    if (g < GQ && q == 0) { sm[warp][g] = m_run; sl[warp][g] = l_run; }
#pragma unroll
    for (int c = 0; c < 2; ++c) {
        const int head = 2 * q + c;
        if (head < GQ) {
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    so[warp][head][8 * (g + 8 * j) + e] = o[j][e][c];
                    so[warp][head][8 * (g + 8 * j) + e + 4] = o[j][e][2 + c];
                }
        }
    }
    __syncthreads();
    const int H = a.H;
    const size_t pbase = (((size_t)m * a.KV + h) * stride + split);
    for (int e = tid; e < GQ * HD; e += ATC_THREADS) {
        const int gg = e / HD, d = e % HD;
        float M = -INFINITY;
#pragma unroll
        for (int w = 0; w < NWARP; ++w) M = fmaxf(M, sm[w][gg]);
        float num = 0.f, den = 0.f;
#pragma unroll
        for (int w = 0; w < NWARP; ++w) {
            const float wgt = (sm[w][gg] == -INFINITY) ? 0.f : expf(sm[w][gg] - M);
            num = fmaf(wgt, so[w][gg][d], num);
            den = fmaf(wgt, sl[w][gg], den);
        }
        if (nsp == 1) {
            a.y[(size_t)m * H * HD + (h * GQ + gg) * HD + d] = __float2bfloat16_rn(num / den);
        } else {
            a.ws_acc[pbase * GQ * HD + e] = num;
            if (d == 0) {
                a.ws_ml[(pbase * GQ + gg) * 2 + 0] = M;
                a.ws_ml[(pbase * GQ + gg) * 2 + 1] = den;
            }
        }
    }
    if (nsp == 1) return;
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const int prev = atomicAdd(&a.counters[m * a.KV + h], 1);
        is_last = (prev == nsp - 1);
        if (tr) tr[3] = clock64();
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    const size_t b0 = ((size_t)m * a.KV + h) * stride;
    for (int e = tid; e < GQ * HD; e += ATC_THREADS) {
        const int gg = e / HD;
        float M = -INFINITY;
        for (int s2 = 0; s2 < nsp; ++s2) M = fmaxf(M, __ldcg(&a.ws_ml[((b0 + s2) * GQ + gg) * 2]));
        float num = 0.f, den = 0.f;
        for (int s2 = 0; s2 < nsp; ++s2) {
            const float ms = __ldcg(&a.ws_ml[((b0 + s2) * GQ + gg) * 2]);
            const float ls = __ldcg(&a.ws_ml[((b0 + s2) * GQ + gg) * 2 + 1]);
            const float w = (ms == -INFINITY) ? 0.f : expf(ms - M);
            num = fmaf(w, __ldcg(&a.ws_acc[(b0 + s2) * GQ * HD + e]), num);
            den = fmaf(w, ls, den);
        }
        a.y[(size_t)m * H * HD + (h * GQ) * HD + e] = __float2bfloat16_rn(num / den);
    }
    if (tid == 0) a.counters[m * a.KV + h] = 0;  // re-arm for the next launch
}

// ----------------------------------------------------------------------------- slow-AR decode attention, latency form
// Same math as attn_decode_body for ONE (split, kv head, row) item, shaped for the persistent frame kernel where this
// op sits on a ~400-step dependent chain: the split's whole key range (<= ASP_MAXKEYS) is one tile, and every K and V
// row of the tile is requested from memory BEFORE the item waits for this frame's q/k/v (the cache does not depend on
// them), so the DRAM latency is paid once and overlaps the wait for the producer op.  Activations are flag-in-data.
#define ASP_THREADS 512
#define ASP_KPT 4  // keys per thread: 4 K + 4 V 128-bit loads in flight per thread
#define ASP_MAXS 16  // splits per (row, kv head) the merge step is unrolled for
__host__ __device__ constexpr int asp_max_keys(int hd) { return (ASP_THREADS / (hd / 8)) * ASP_KPT; }  // 128 at hd 128

// shared-memory index of element d of a head vector, arranged so that a lane's 8 elements (chunk = d / 8) are two
// float4 whose neighbours in the warp are contiguous: [half][chunk][4]  (conflict-free LDS.128 / STS.128)
template <int HD>
__device__ __forceinline__ int asp_idx(int d) {
    return (((d & 7) >> 2) * (HD / 8) + (d >> 3)) * 4 + (d & 3);
}

template <int HD, int GQ>
struct AttnSplitSmem {
    alignas(16) float qs[GQ][HD];
    alignas(16) float knew[HD], vnew[HD];
    alignas(16) float sc[GQ][asp_max_keys(HD)];
    alignas(16) float red[ASP_THREADS / 64][GQ][HD];
    float mrun[GQ], lrun[GQ];
    float corr[GQ];  // TILED: rescale factor of the running numerators for the current key tile
    int is_last;
};

// L2 prefetch of the cached keys/values one (split, kv head, row) item will read: a demand load that misses L2 queues
// behind the weight stream's outstanding bulk copies (~4 us under load), so the item's KV range is pulled in a layer ahead
template <int HD>
__device__ __forceinline__ void attn_split_prefetch(const AttnArgs& a, const bf16* kc, const bf16* vc, int split, int h, int m) {
    const int pos = attn_pos(a, m);
    const int slot = a.slots ? a.slots[m] : m;
    const int per = (pos + 1 + a.splits - 1) / a.splits;
    const int k0 = split * per, k1 = min(pos, k0 + per);
    l2_prefetch_bulk(a.rope + (size_t)pos * HD, HD * 2);
    if (k1 <= k0) return;
    const size_t off = (((size_t)slot * a.KV + h) * a.S + k0) * HD;
    const uint32_t bytes = (uint32_t)(k1 - k0) * HD * 2;
    l2_prefetch_bulk(kc + off, bytes);
    l2_prefetch_bulk(vc + off, bytes);
}

// TILED = true: a split may hold more than asp_max_keys keys (long caches): its range is walked in tiles of asp_max_keys with
// a running (max, sum) per head and rescaled numerators (flash-decoding); the next tile's keys are requested as soon as this
// tile's are consumed.  With one tile it computes what the untiled form computes.
template <int HD, int GQ, class SY, bool TILED = false>
__device__ __forceinline__ void attn_split_body(const AttnArgs& a, AttnSplitSmem<HD, GQ>& sm, int split, int h, int m,
                                                int tid, unsigned tag_in, unsigned tag_out,
                                                unsigned long long* trace = nullptr) {
    constexpr int EPL = HD / 32;
    constexpr int LPK = HD / 8;               // lanes per key row (16 B each)
    constexpr int KPI = ASP_THREADS / LPK;    // key slots
    constexpr int NW = ASP_THREADS / 32;
    const int lane = tid & 31, warp = tid >> 5;
    const int kslot = tid / LPK, chunk = tid % LPK;

    const int pos = attn_pos(a, m);
    const int slot = a.slots ? a.slots[m] : m;
    const int qkv_w = (a.H + 2 * a.KV) * HD;
    const uint32_t* row_t = a.qkv_t + (size_t)m * qkv_w;
    const int n_keys = pos + 1;
    const int per = (n_keys + a.splits - 1) / a.splits;
    const int k0 = split * per;
    const int k1 = min(n_keys, k0 + per);
    const int nk = max(0, k1 - k0);                       // keys of this split (<= asp_max_keys, checked by the host)
    const bool owns_new = (pos >= k0) && (pos < k1);      // the new key/value come from this frame's qkv
    const int kc1 = owns_new ? pos : k1;                  // keys [k0, kc1) are read from the cache
    const bf16* kbase = a.kcache + ((size_t)slot * a.KV + h) * a.S * HD;
    const bf16* vbase = a.vcache + ((size_t)slot * a.KV + h) * a.S * HD;

    // ---- 1. request the whole K and V tile now (independent of this frame's activations)
    uint4 kq[ASP_KPT], vq[ASP_KPT];
#pragma unroll
    for (int i = 0; i < ASP_KPT; ++i) {
        const int j = k0 + kslot + KPI * i;
        kq[i] = (j < kc1) ? ld_stream(kbase + (size_t)j * HD + chunk * 8) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int i = 0; i < ASP_KPT; ++i) {
        const int j = k0 + kslot + KPI * i;
        vq[i] = (j < kc1) ? ld_stream(vbase + (size_t)j * HD + chunk * 8) : make_uint4(0, 0, 0, 0);
    }

    // ---- 2. q heads (warps 0..GQ-1), new k (warp GQ), new v (warp GQ+1): spin on the producer's tags; the norm
    //         weights and the RoPE row are fetched first (they do not depend on the frame)
    const bf16* rope_row = a.rope + (size_t)pos * HD;
    HeadCoef<HD> coef;
    if (warp <= GQ) coef.load(warp < GQ ? a.q_norm : a.k_norm, rope_row, a.qk_norm);
    if (warp < GQ) {
        float qin[EPL], q[EPL];
        load_head<HD, true>(nullptr, row_t + (h * GQ + warp) * HD, tag_in, qin);
        head_norm_rope<HD>(qin, coef, a.qk_norm, a.eps, q);
#pragma unroll
        for (int i = 0; i < EPL; ++i) sm.qs[warp][asp_idx<HD>(lane * EPL + i)] = q[i];
    } else if (owns_new && warp == GQ) {
        const size_t co = (((size_t)slot * a.KV + h) * a.S + pos) * HD;
        float kin[EPL], k[EPL];
        load_head<HD, true>(nullptr, row_t + (a.H + h) * HD, tag_in, kin);
        head_norm_rope<HD>(kin, coef, a.qk_norm, a.eps, k);
#pragma unroll
        for (int i = 0; i < EPL; ++i) {
            sm.knew[asp_idx<HD>(lane * EPL + i)] = k[i];
            a.kcache[co + lane * EPL + i] = __float2bfloat16_rn(k[i]);
        }
    } else if (owns_new && warp == GQ + 1) {
        const size_t co = (((size_t)slot * a.KV + h) * a.S + pos) * HD;
        float vin[EPL];
        load_head<HD, true>(nullptr, row_t + (a.H + a.KV + h) * HD, tag_in, vin);
#pragma unroll
        for (int i = 0; i < EPL; ++i) {
            sm.vnew[asp_idx<HD>(lane * EPL + i)] = vin[i];
            a.vcache[co + lane * EPL + i] = __float2bfloat16_rn(vin[i]);
        }
    }
    SY::sync();
#define ASP_TRACE(point)                                              \
    if (trace && tid == 0 && a.pf_bytes == (size_t)(point)) {         \
        unsigned long long t_;                                        \
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_));         \
        trace[1] = t_;                                                \
    }
    ASP_TRACE(0)

    float acc[GQ][8];
#pragma unroll
    for (int g = 0; g < GQ; ++g)
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[g][e] = 0.f;
    if constexpr (!TILED) {
        // ---- 3. scores
    #pragma unroll
        for (int i = 0; i < ASP_KPT; ++i) {
            const int jl = kslot + KPI * i;       // key index inside the split (warp-uniform validity: nk is uniform)
            if (KPI * i >= nk) break;
            float kf[8];
            if (k0 + jl < kc1) unpack8(kq[i], kf);
            else {
    #pragma unroll
                for (int e = 0; e < 8; ++e) kf[e] = (jl < nk) ? sm.knew[asp_idx<HD>(chunk * 8 + e)] : 0.f;
            }
            float d[GQ];
    #pragma unroll
            for (int g = 0; g < GQ; ++g) {
                const float4 qa = *reinterpret_cast<const float4*>(&sm.qs[g][chunk * 4]);
                const float4 qb = *reinterpret_cast<const float4*>(&sm.qs[g][(LPK + chunk) * 4]);
                float s = qa.x * kf[0];
                s = fmaf(qa.y, kf[1], s); s = fmaf(qa.z, kf[2], s); s = fmaf(qa.w, kf[3], s);
                s = fmaf(qb.x, kf[4], s); s = fmaf(qb.y, kf[5], s); s = fmaf(qb.z, kf[6], s); s = fmaf(qb.w, kf[7], s);
                d[g] = s;
            }
    #pragma unroll
            for (int o = LPK / 2; o > 0; o >>= 1)
    #pragma unroll
                for (int g = 0; g < GQ; ++g) d[g] += __shfl_xor_sync(0xffffffffu, d[g], o);
            if (chunk == 0 && jl < nk)
    #pragma unroll
                for (int g = 0; g < GQ; ++g) sm.sc[g][jl] = d[g] * a.scale;
        }
        SY::sync();
        ASP_TRACE(1)

        // ---- 4. softmax over the split (fp32, like the MATH SDPA backend): warp g owns head g
        if (warp < GQ) {
            float mx = -INFINITY;
            for (int j = lane; j < nk; j += 32) mx = fmaxf(mx, sm.sc[warp][j]);
            mx = warp_max(mx);
            float sum = 0.f;
            for (int j = lane; j < nk; j += 32) {
                const float pj = expf(sm.sc[warp][j] - mx);
                sm.sc[warp][j] = pj;
                sum += pj;
            }
            sum = warp_sum(sum);
            if (lane == 0) { sm.mrun[warp] = mx; sm.lrun[warp] = sum; }
        }
        SY::sync();
        ASP_TRACE(2)

        // ---- 5. P.V
    #pragma unroll
        for (int i = 0; i < ASP_KPT; ++i) {
            const int jl = kslot + KPI * i;
            if (jl < nk) {
                float vf[8];
                if (k0 + jl < kc1) unpack8(vq[i], vf);
                else {
    #pragma unroll
                    for (int e = 0; e < 8; ++e) vf[e] = sm.vnew[asp_idx<HD>(chunk * 8 + e)];
                }
    #pragma unroll
                for (int g = 0; g < GQ; ++g) {
                    const float pj = sm.sc[g][jl];
    #pragma unroll
                    for (int e = 0; e < 8; ++e) acc[g][e] = fmaf(pj, vf[e], acc[g][e]);
                }
            }
        }

    } else {
        constexpr int MAXK = asp_max_keys(HD);
        const int ntiles = max(1, (nk + MAXK - 1) / MAXK);
        for (int t = 0; t < ntiles; ++t) {
            const int tb = t * MAXK;                 // first key of the tile, counted inside the split
            const int nkt = min(MAXK, nk - tb);      // keys of the tile (0 only for an empty split)
            // ---- 3. scores of the tile
#pragma unroll
            for (int i = 0; i < ASP_KPT; ++i) {
                const int jl = kslot + KPI * i;
                if (KPI * i >= nkt) break;
                float kf[8];
                if (k0 + tb + jl < kc1) unpack8(kq[i], kf);
                else {
#pragma unroll
                    for (int e = 0; e < 8; ++e) kf[e] = (jl < nkt) ? sm.knew[asp_idx<HD>(chunk * 8 + e)] : 0.f;
                }
                float d[GQ];
#pragma unroll
                for (int g = 0; g < GQ; ++g) {
                    const float4 qa = *reinterpret_cast<const float4*>(&sm.qs[g][chunk * 4]);
                    const float4 qb = *reinterpret_cast<const float4*>(&sm.qs[g][(LPK + chunk) * 4]);
                    float s = qa.x * kf[0];
                    s = fmaf(qa.y, kf[1], s); s = fmaf(qa.z, kf[2], s); s = fmaf(qa.w, kf[3], s);
                    s = fmaf(qb.x, kf[4], s); s = fmaf(qb.y, kf[5], s); s = fmaf(qb.z, kf[6], s); s = fmaf(qb.w, kf[7], s);
                    d[g] = s;
                }
#pragma unroll
                for (int o = LPK / 2; o > 0; o >>= 1)
#pragma unroll
                    for (int g = 0; g < GQ; ++g) d[g] += __shfl_xor_sync(0xffffffffu, d[g], o);
                if (chunk == 0 && jl < nkt)
#pragma unroll
                    for (int g = 0; g < GQ; ++g) sm.sc[g][jl] = d[g] * a.scale;
            }
            uint4 vnext[ASP_KPT];
            if (t + 1 < ntiles) {  // the key registers are free: request the next tile's keys AND values now, so that both
#pragma unroll                     // travel under this tile's softmax and P.V (a load that misses L2 waits ~2-4 us behind the weight stream)
                for (int i = 0; i < ASP_KPT; ++i) {
                    const int j = k0 + tb + MAXK + kslot + KPI * i;
                    kq[i] = (j < kc1) ? ld_stream(kbase + (size_t)j * HD + chunk * 8) : make_uint4(0, 0, 0, 0);
                }
#pragma unroll
                for (int i = 0; i < ASP_KPT; ++i) {
                    const int j = k0 + tb + MAXK + kslot + KPI * i;
                    vnext[i] = (j < kc1) ? ld_stream(vbase + (size_t)j * HD + chunk * 8) : make_uint4(0, 0, 0, 0);
                }
            }
            SY::sync();
            // ---- 4. softmax of the tile against the running (max, sum): warp g owns head g
            if (warp < GQ) {
                const float m_old = t ? sm.mrun[warp] : -INFINITY, l_old = t ? sm.lrun[warp] : 0.f;
                float mx = -INFINITY;
                for (int j = lane; j < nkt; j += 32) mx = fmaxf(mx, sm.sc[warp][j]);
                mx = fmaxf(warp_max(mx), m_old);
                float sum = 0.f;
                for (int j = lane; j < nkt; j += 32) {
                    const float pj = expf(sm.sc[warp][j] - mx);
                    sm.sc[warp][j] = pj;
                    sum += pj;
                }
                sum = warp_sum(sum);
                const float corr = t ? expf(m_old - mx) : 0.f;
                if (lane == 0) { sm.mrun[warp] = mx; sm.lrun[warp] = fmaf(l_old, corr, sum); sm.corr[warp] = corr; }
            }
            SY::sync();
            // ---- 5. P.V of the tile on top of the rescaled numerators
            if (t) {
#pragma unroll
                for (int g = 0; g < GQ; ++g) {
                    const float c = sm.corr[g];
#pragma unroll
                    for (int e = 0; e < 8; ++e) acc[g][e] *= c;
                }
            }
#pragma unroll
            for (int i = 0; i < ASP_KPT; ++i) {
                const int jl = kslot + KPI * i;
                if (jl < nkt) {
                    float vf[8];
                    if (k0 + tb + jl < kc1) unpack8(vq[i], vf);
                    else {
#pragma unroll
                        for (int e = 0; e < 8; ++e) vf[e] = sm.vnew[asp_idx<HD>(chunk * 8 + e)];
                    }
#pragma unroll
                    for (int g = 0; g < GQ; ++g) {
                        const float pj = sm.sc[g][jl];
#pragma unroll
                        for (int e = 0; e < 8; ++e) acc[g][e] = fmaf(pj, vf[e], acc[g][e]);
                    }
                }
            }
            if (t + 1 < ntiles) {
#pragma unroll
                for (int i = 0; i < ASP_KPT; ++i) vq[i] = vnext[i];
                SY::sync();  // the scores and factors of this tile are overwritten by the next one
            }
        }
    }
    // key slots of one warp (32 / LPK of them), then the warps through shared memory, in a fixed order
#pragma unroll
    for (int o = LPK; o < 32; o <<= 1)
#pragma unroll
        for (int g = 0; g < GQ; ++g)
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[g][e] += __shfl_xor_sync(0xffffffffu, acc[g][e], o);
    if (warp >= NW / 2 && lane < LPK) {  // upper half of the warps hands its partials to the lower half
#pragma unroll
        for (int g = 0; g < GQ; ++g) {
            float* r = sm.red[warp - NW / 2][g];
            *reinterpret_cast<float4*>(&r[chunk * 4]) = make_float4(acc[g][0], acc[g][1], acc[g][2], acc[g][3]);
            *reinterpret_cast<float4*>(&r[(LPK + chunk) * 4]) = make_float4(acc[g][4], acc[g][5], acc[g][6], acc[g][7]);
        }
    }
    SY::sync();
    if (warp < NW / 2 && lane < LPK) {
#pragma unroll
        for (int g = 0; g < GQ; ++g) {
            float* r = sm.red[warp][g];
            float4 lo = *reinterpret_cast<float4*>(&r[chunk * 4]), hi = *reinterpret_cast<float4*>(&r[(LPK + chunk) * 4]);
            lo.x += acc[g][0]; lo.y += acc[g][1]; lo.z += acc[g][2]; lo.w += acc[g][3];
            hi.x += acc[g][4]; hi.y += acc[g][5]; hi.z += acc[g][6]; hi.w += acc[g][7];
            *reinterpret_cast<float4*>(&r[chunk * 4]) = lo;
            *reinterpret_cast<float4*>(&r[(LPK + chunk) * 4]) = hi;
        }
    }
    SY::sync();

    ASP_TRACE(3)
    const int H = a.H;
    const size_t pbase = (((size_t)m * a.KV + h) * a.splits + split);
    const uint32_t ptag = tag_out | 0x10000u;  // 32-bit tag of this op's partials (never 0)
    for (int e = tid; e < GQ * HD; e += ASP_THREADS) {
        const int g = e / HD, d = e % HD;
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < NW / 2; ++w) s += sm.red[w][g][asp_idx<HD>(d)];
        if (a.splits == 1) st_tagged(a.y_t + (size_t)m * H * HD + (h * GQ + g) * HD + d, s / sm.lrun[g], tag_out);
        else st_ll64(a.ll_acc + pbase * GQ * HD + e, s, ptag);
    }
    if (a.splits == 1) return;
    // the (max, sum) records are published after the numerators: barrier + fence by the publishing threads make
    // "record visible" imply "numerators visible", so the merging item fetches them once instead of re-polling
    SY::sync();
    if (tid < GQ) {
        __threadfence();
        st_ll64(a.ll_ml + (pbase * GQ + tid) * 2 + 0, (nk > 0) ? sm.mrun[tid] : -INFINITY, ptag);
        st_ll64(a.ll_ml + (pbase * GQ + tid) * 2 + 1, (nk > 0) ? sm.lrun[tid] : 0.f, ptag);
    }
    // ---- 6. the split-0 item of each (row, kv head) merges the partials: it spins on their tags (no fence, no atomic,
    //         no "last arriver" election), all loads of a thread in flight together
    ASP_TRACE(4)
    if (split != 0) return;
    static_assert(GQ * HD <= ASP_THREADS, "one output element per thread in the merge");
    const size_t b0 = ((size_t)m * a.KV + h) * a.splits;
    float* mlbuf = &sm.red[0][0][0];                 // [splits][GQ][2]; the reduction buffer is free after the sync
    float* wbuf = mlbuf + ASP_MAXS * GQ * 2;         // [splits][GQ] merge weights, then [GQ] 1 / denominators
    const int e = tid;
    const bool has = e < GQ * HD;
    SY::sync();
    {
        // wait on the small (max, sum) records only — a few threads, with back-off — then fetch the numerators, which the
        // same CTAs stored just before (their tags are still checked: stores are not ordered)
        unsigned spins = 0;
        long long t0 = 0;
        for (int i = tid; i < a.splits * GQ * 2; i += ASP_THREADS) {
            float v;
            while (!ld_ll64_try(a.ll_ml + b0 * GQ * 2 + i, ptag, v))
                if (spin_watchdog(spins, t0, FSB_ERR_PARTIAL_WAIT)) break;
            mlbuf[i] = v;
        }
    }
    SY::sync();
    ASP_TRACE(5)
    float pv[ASP_MAXS];
    {
        unsigned spins = 0;
        long long t0 = 0;
        bool ok;
        do {
            ok = true;
#pragma unroll
            for (int s = 0; s < ASP_MAXS; ++s) {
                pv[s] = 0.f;
                if (has && s < a.splits) ok &= ld_ll64_try(a.ll_acc + (b0 + s) * GQ * HD + e, ptag, pv[s]);
            }
            if (!ok && spin_watchdog(spins, t0, FSB_ERR_PARTIAL_WAIT)) break;
        } while (!ok);
    }
    if (warp < GQ) {  // warp g, lane s: global max, weights exp(m_s - M), sum(w_s l_s) — a serial loop on one thread
        static_assert(ASP_MAXS <= 32, "one lane per split");      // costs microseconds on this dependent chain
        const bool live = lane < a.splits;
        const float ms = live ? mlbuf[(lane * GQ + warp) * 2] : -INFINITY;
        const float ls = live ? mlbuf[(lane * GQ + warp) * 2 + 1] : 0.f;
        const float M = warp_max(ms);
        const float w = (ms == -INFINITY) ? 0.f : expf(ms - M);
        const float den = warp_sum(w * ls);
        if (lane < ASP_MAXS) wbuf[lane * GQ + warp] = w;
        if (lane == 0) wbuf[ASP_MAXS * GQ + warp] = den;
    }
    SY::sync();
    ASP_TRACE(6)
    if (has) {
        const int g = e / HD;
        float num = 0.f;
#pragma unroll
        for (int s = 0; s < ASP_MAXS; ++s) num = fmaf(wbuf[s * GQ + g], pv[s], num);
        st_tagged(a.y_t + (size_t)m * H * HD + (h * GQ) * HD + e, num / wbuf[ASP_MAXS * GQ + g], tag_out);
    }
}

// ----------------------------------------------------------------------------- fast-AR attention (context <= 16)
#define FAST_MAX_CTX 16

template <int HD, int GQ, class SY, bool TAG = false>
__device__ __forceinline__ void fast_attn_body(const AttnArgs& a, int h, int m, int tid, unsigned tag_in = 0,
                                               unsigned tag_out = 0) {
    constexpr int EPL = HD / 32;
    __shared__ float knew[HD], vnew[HD];
    const int lane = tid & 31, g = tid >> 5;

    const int pos = a.pos_fixed;  // codebook index of this pass
    const int slot = a.slots ? a.slots[m] : m;
    const int qkv_w = (a.H + 2 * a.KV) * HD;
    const bf16* row = a.qkv + (size_t)m * qkv_w;
    const uint32_t* row_t = TAG ? a.qkv_t + (size_t)m * qkv_w : nullptr;
    const bf16* rope_row = a.rope + (size_t)pos * HD;
    const size_t cbase = ((size_t)slot * a.KV + h) * a.S * HD;

    // everything that does not depend on this pass's activations is requested first: norm weights, the RoPE row and
    // the cached keys/values of the earlier codebook positions (latency hides behind the wait for the producer op)
    HeadCoef<HD> cq, ck;
    cq.load(a.q_norm, rope_row, a.qk_norm);
    if (g == 0) ck.load(a.k_norm, rope_row, a.qk_norm);
    uint32_t kraw[FAST_MAX_CTX][EPL / 2], vraw[FAST_MAX_CTX][EPL / 2];
#pragma unroll
    for (int j = 0; j < FAST_MAX_CTX; ++j)
#pragma unroll
        for (int i = 0; i < EPL / 2; ++i) {
            kraw[j][i] = 0; vraw[j][i] = 0;
            if (j < pos) {
                kraw[j][i] = __ldcg(reinterpret_cast<const uint32_t*>(a.kcache + cbase + (size_t)j * HD + lane * EPL) + i);
                vraw[j][i] = __ldcg(reinterpret_cast<const uint32_t*>(a.vcache + cbase + (size_t)j * HD + lane * EPL) + i);
            }
        }

    float qin[EPL], q[EPL];
    const bool part = !TAG && a.qkv_part != nullptr;
    if (part) load_head_part<HD>(a.qkv_part, a.qkv_s, a.qkv_bn, m, (h * GQ + g) * HD, qin);
    else load_head<HD, TAG>(row + (h * GQ + g) * HD, row_t + (h * GQ + g) * HD, tag_in, qin);
    head_norm_rope<HD>(qin, cq, a.qk_norm, a.eps, q);
    if (g == 0) {
        float kin[EPL], k[EPL];
        if (part) load_head_part<HD>(a.qkv_part, a.qkv_s, a.qkv_bn, m, (a.H + h) * HD, kin);
        else load_head<HD, TAG>(row + (a.H + h) * HD, row_t + (a.H + h) * HD, tag_in, kin);
        head_norm_rope<HD>(kin, ck, a.qk_norm, a.eps, k);
#pragma unroll
        for (int i = 0; i < EPL; ++i) {
            knew[lane * EPL + i] = k[i];
            a.kcache[cbase + (size_t)pos * HD + lane * EPL + i] = __float2bfloat16_rn(k[i]);
        }
    }
    if (g == GQ - 1) {
        float vin[EPL];
        if (part) load_head_part<HD>(a.qkv_part, a.qkv_s, a.qkv_bn, m, (a.H + a.KV + h) * HD, vin);
        else load_head<HD, TAG>(row + (a.H + a.KV + h) * HD, row_t + (a.H + a.KV + h) * HD, tag_in, vin);
#pragma unroll
        for (int i = 0; i < EPL; ++i) {
            vnew[lane * EPL + i] = vin[i];
            a.vcache[cbase + (size_t)pos * HD + lane * EPL + i] = __float2bfloat16_rn(vin[i]);
        }
    }
    SY::sync();

    // scores: bf16(q.k) then bf16(* scale)   (llama.py:964-966; masked entries never enter the softmax)
    float s[FAST_MAX_CTX];
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < FAST_MAX_CTX; ++j) {
        s[j] = -INFINITY;
        if (j <= pos) {
            float d = 0.f;
#pragma unroll
            for (int i = 0; i < EPL; ++i) {
                const uint32_t kw = kraw[j][i >> 1];
                const float kv = (j == pos) ? knew[lane * EPL + i] : ((i & 1) ? bf16_hi(kw) : bf16_lo(kw));
                d = fmaf(q[i], kv, d);
            }
            d = warp_sum(d);
            s[j] = rbf(rbf(d) * a.scale);
            mx = fmaxf(mx, s[j]);
        }
    }
    float den = 0.f;
#pragma unroll
    for (int j = 0; j < FAST_MAX_CTX; ++j)
        if (j <= pos) { s[j] = expf(s[j] - mx); den += s[j]; }
    float o[EPL];
#pragma unroll
    for (int i = 0; i < EPL; ++i) o[i] = 0.f;
#pragma unroll
    for (int j = 0; j < FAST_MAX_CTX; ++j)
        if (j <= pos) {
            const float p = rbf(__fdiv_rn(s[j], den));  // softmax output in bf16
#pragma unroll
            for (int i = 0; i < EPL; ++i) {
                const uint32_t vw = vraw[j][i >> 1];
                const float vv = (j == pos) ? vnew[lane * EPL + i] : ((i & 1) ? bf16_hi(vw) : bf16_lo(vw));
                o[i] = fmaf(p, vv, o[i]);
            }
        }
#pragma unroll
    for (int i = 0; i < EPL; ++i) {
        const size_t yi = (size_t)m * a.H * HD + (h * GQ + g) * HD + lane * EPL + i;
        if (TAG) st_tagged(a.y_t + yi, o[i], tag_out);
        else a.y[yi] = __float2bfloat16_rn(o[i]);
    }
}
template <int HD, int GQ>
__global__ void __launch_bounds__(32 * GQ) fast_attn_kernel(AttnArgs a) {
    if (threadIdx.x == 0)
        l2_prefetch_share(a.pf_ptr, a.pf_bytes, blockIdx.y * gridDim.x + blockIdx.x, gridDim.x * gridDim.y);
    pdl_wait();
    pdl_launch_dependents();
    fast_attn_body<HD, GQ, SyncBlock<32 * GQ>>(a, blockIdx.x, blockIdx.y, threadIdx.x);
}
