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
    int D, C, cb, sem_begin, sem_end, scale;
    float inv_div;           // sqrt(C+1) as float (the reference divides)
    const void* pf_ptr; size_t pf_bytes;
};

__device__ __forceinline__ void embed_body(const EmbedArgs& a, int m, int tid, int nt) {
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
        reinterpret_cast<uint4*>(a.x + (size_t)m * a.D)[c] = o;
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
template <int HD>
__device__ __forceinline__ void head_norm_rope(const bf16* src, const bf16* norm_w, const bf16* rope_row,
                                               bool qk_norm, float eps, float (&out)[HD / 32]) {
    constexpr int EPL = HD / 32;
    const int lane = threadIdx.x & 31;
    float v[EPL];
#pragma unroll
    for (int i = 0; i < EPL; ++i) v[i] = lda(&src[lane * EPL + i]);
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
    float* ws_acc;          // [M][KV][splits][GQ][HD]
    float* ws_ml;           // [M][KV][splits][GQ][2]
    int32_t* counters;      // [M][KV]
    int H, KV, S, splits, qk_norm, fused_append, pos_fixed;  // pos_fixed >= 0 overrides pos[] (fast AR)
    float eps, scale;
    const void* pf_ptr; size_t pf_bytes;
};

// ----------------------------------------------------------------------------- prompt rows: k-norm + RoPE + scatter
template <int HD>
__global__ void __launch_bounds__(64) rope_append_kernel(AttnArgs a) {
    if (threadIdx.x == 0)
        l2_prefetch_share(a.pf_ptr, a.pf_bytes, blockIdx.y * gridDim.x + blockIdx.x, gridDim.x * gridDim.y);
    pdl_wait();
    pdl_launch_dependents();
    constexpr int EPL = HD / 32;
    const int h = blockIdx.x, m = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int pos = a.pos[m];
    const int slot = a.slots ? a.slots[m] : m;
    const int qkv_w = (a.H + 2 * a.KV) * HD;
    const size_t co = (((size_t)slot * a.KV + h) * a.S + pos) * HD;
    if (warp == 0) {
        float k[EPL];
        head_norm_rope<HD>(a.qkv + (size_t)m * qkv_w + (a.H + h) * HD, a.k_norm, a.rope + (size_t)pos * HD,
                           a.qk_norm, a.eps, k);
#pragma unroll
        for (int i = 0; i < EPL; ++i) a.kcache[co + lane * EPL + i] = __float2bfloat16_rn(k[i]);
    } else {
        const bf16* v = a.qkv + (size_t)m * qkv_w + (a.H + a.KV + h) * HD;
#pragma unroll
        for (int i = 0; i < EPL; ++i) a.vcache[co + lane * EPL + i] = lda_raw(&v[lane * EPL + i]);
    }
}

// ----------------------------------------------------------------------------- slow-AR decode attention
#define ATTN_THREADS 128
#define ATTN_TILE 128

// one (split, kv head, row) work item on ATTN_THREADS threads synchronised through SY; tid = index inside the group
// QREG: keep q in registers (standalone kernel) or read it from shared memory in the FMA loop (persistent frame kernel,
// which is capped at 96 registers per thread by its 17 warps per SM)
template <int HD, int GQ, class SY, bool QREG = true>
__device__ __forceinline__ void attn_decode_body(const AttnArgs& a, int split, int h, int m, int tid) {
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

    const int pos = a.pos[m];
    const int slot = a.slots ? a.slots[m] : m;
    const int qkv_w = (a.H + 2 * a.KV) * HD;
    const bf16* row = a.qkv + (size_t)m * qkv_w;
    const int n_keys = pos + 1;
    const int per = (n_keys + a.splits - 1) / a.splits;
    const int k0 = split * per;
    const int k1 = min(n_keys, k0 + per);
    const bool owns_new = a.fused_append && (pos >= k0) && (pos < k1);
    const bf16* rope_row = a.rope + (size_t)pos * HD;

    // ---- q prep (one warp per q head of this GQA group), optional fused K/V append
    for (int g = warp; g < GQ; g += ATTN_THREADS / 32) {
        float q[EPL];
        head_norm_rope<HD>(row + (h * GQ + g) * HD, a.q_norm, rope_row, a.qk_norm, a.eps, q);
#pragma unroll
        for (int i = 0; i < EPL; ++i) qs[g][lane * EPL + i] = q[i];
    }
    if (owns_new) {
        const size_t co = (((size_t)slot * a.KV + h) * a.S + pos) * HD;
        if (warp == (GQ % (ATTN_THREADS / 32))) {  // spread the extra work to a different warp when possible
            float k[EPL];
            head_norm_rope<HD>(row + (a.H + h) * HD, a.k_norm, rope_row, a.qk_norm, a.eps, k);
#pragma unroll
            for (int i = 0; i < EPL; ++i) {
                knew[lane * EPL + i] = k[i];
                a.kcache[co + lane * EPL + i] = __float2bfloat16_rn(k[i]);
            }
        }
        if (warp == ((GQ + 1) % (ATTN_THREADS / 32))) {
            const bf16* v = row + (a.H + a.KV + h) * HD;
#pragma unroll
            for (int i = 0; i < EPL; ++i) {
                const bf16 t = lda_raw(&v[lane * EPL + i]);
                vnew[lane * EPL + i] = __bfloat162float(t);
                a.vcache[co + lane * EPL + i] = t;
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
        // -- scores for this tile
        for (int jb = t0; jb < t1; jb += KPI) {  // warp-uniform trip count (shuffles below)
            const int j = jb + kslot;
            const bool jv = j < t1;
            float kf[8];
            if (jv && j < kc1) {
                unpack8(*reinterpret_cast<const uint4*>(kbase + (size_t)j * HD + chunk * 8), kf);
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
        for (int j = t0 + kslot; j < t1; j += KPI) {
            float vf[8];
            if (j < kc1) {
                unpack8(*reinterpret_cast<const uint4*>(vbase + (size_t)j * HD + chunk * 8), vf);
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
            a.y[(size_t)m * H * HD + (h * GQ + g) * HD + d] = __float2bfloat16_rn(s / lrun[g]);
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
        a.y[(size_t)m * H * HD + (h * GQ) * HD + e] = __float2bfloat16_rn(num / den);
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

// ----------------------------------------------------------------------------- fast-AR attention (context <= 16)
#define FAST_MAX_CTX 16

template <int HD, int GQ, class SY>
__device__ __forceinline__ void fast_attn_body(const AttnArgs& a, int h, int m, int tid) {
    constexpr int EPL = HD / 32;
    __shared__ float knew[HD], vnew[HD];
    const int lane = tid & 31, g = tid >> 5;

    const int pos = a.pos_fixed;  // codebook index of this pass
    const int slot = a.slots ? a.slots[m] : m;
    const int qkv_w = (a.H + 2 * a.KV) * HD;
    const bf16* row = a.qkv + (size_t)m * qkv_w;
    const bf16* rope_row = a.rope + (size_t)pos * HD;
    const size_t cbase = ((size_t)slot * a.KV + h) * a.S * HD;

    float q[EPL];
    head_norm_rope<HD>(row + (h * GQ + g) * HD, a.q_norm, rope_row, a.qk_norm, a.eps, q);
    if (g == 0) {
        float k[EPL];
        head_norm_rope<HD>(row + (a.H + h) * HD, a.k_norm, rope_row, a.qk_norm, a.eps, k);
#pragma unroll
        for (int i = 0; i < EPL; ++i) {
            knew[lane * EPL + i] = k[i];
            a.kcache[cbase + (size_t)pos * HD + lane * EPL + i] = __float2bfloat16_rn(k[i]);
        }
    }
    if (g == GQ - 1) {
        const bf16* v = row + (a.H + a.KV + h) * HD;
#pragma unroll
        for (int i = 0; i < EPL; ++i) {
            const bf16 t = lda_raw(&v[lane * EPL + i]);
            vnew[lane * EPL + i] = __bfloat162float(t);
            a.vcache[cbase + (size_t)pos * HD + lane * EPL + i] = t;
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
                const float kv = (j == pos) ? knew[lane * EPL + i]
                                            : lda(&a.kcache[cbase + (size_t)j * HD + lane * EPL + i]);
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
                const float vv = (j == pos) ? vnew[lane * EPL + i]
                                            : lda(&a.vcache[cbase + (size_t)j * HD + lane * EPL + i]);
                o[i] = fmaf(p, vv, o[i]);
            }
        }
#pragma unroll
    for (int i = 0; i < EPL; ++i)
        a.y[(size_t)m * a.H * HD + (h * GQ + g) * HD + lane * EPL + i] = __float2bfloat16_rn(o[i]);
}
template <int HD, int GQ>
__global__ void __launch_bounds__(32 * GQ) fast_attn_kernel(AttnArgs a) {
    if (threadIdx.x == 0)
        l2_prefetch_share(a.pf_ptr, a.pf_bytes, blockIdx.y * gridDim.x + blockIdx.x, gridDim.x * gridDim.y);
    pdl_wait();
    pdl_launch_dependents();
    fast_attn_body<HD, GQ, SyncBlock<32 * GQ>>(a, blockIdx.x, blockIdx.y, threadIdx.x);
}
