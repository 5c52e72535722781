// On-device logit processing + sampling: one CTA per request, no host round trip.
//
// Restates inference.py:43-93 (multinomial_sample_one_no_sync, logits_to_probs, sample) and the
// RAS selection in decode_one_token_ar (inference.py:118-155) with the reference's bf16 rounding
// points:
//   sort desc -> softmax (bf16) -> cumsum (fp32 running sum, bf16 outputs) -> (cum > top_p) | (rank >= top_k),
//   rank 0 always kept -> -inf fill -> / clip(T, 1e-5) (bf16) -> softmax (bf16) -> argmax(p / -log(u)) (bf16).
// Only the first top_k ranks can survive the mask, so the full sort is replaced by an exact top-k
// selection (two-level radix select on the 16-bit bf16 key), the softmax denominator over all V
// entries, and a <= 64-term prefix sum.  Ties are ordered by ascending index (what a stable sort,
// e.g. torch.sort on CUDA, yields; torch's CPU sort is not stable, so the CPU reference itself is
// implementation-defined there).  Entries that are -inf contribute exp()=0 and can never be
// sampled, so they are skipped; callers may therefore restrict the scan to the finite ranges of
// the constrained-decode bias (inference.py:308-320) without changing the result.
#pragma once
#include "common.cuh"

#define SAMPLE_THREADS 1024
#define SAMPLE_MAXK 64
#define SAMPLE_CAND_CAP 1024
#define RAS_WIN 10

struct SampleArgs {
    const bf16* logits;        // [B][ld_logits]
    const bf16* bias;          // [V] added in bf16 (semantic_logit_bias) or nullptr
    int V, ld_logits;
    int seg_lo[2], seg_hi[2];  // ascending, disjoint index ranges to scan; others are known to be -inf
    int top_k;
    int n_variants;            // 1, or 2 = (normal, RAS high-temperature) for the slow token
    float temp[2], top_p[2];   // bf16-representable values (generate() builds them in the model dtype)
    const float* row_temp;     // [B] per-request values for variant 0 (batches that mix requests), or nullptr
    const float* row_top_p;
    const int32_t* row_top_k;
    const bf16* uniforms;      // injected draws [frames][B][u_stride]; nullptr -> Philox
    long long u_stride;        // elements per (frame, request)
    long long u_offset;        // offset of this call's first draw inside a (frame, request) record
    unsigned long long seed;
    const int32_t* frame_counter;  // device scalar: frame index (selects the uniform record / Philox counter)
    int draw_id;               // distinguishes the draws of one frame in Philox mode
    // ---- outputs
    int mode;                  // 0: fast codebook sample -> tokens[b][out_col]; 1: slow main token (+RAS, +code 0); 2: raw draws
    int32_t* tokens;           // [B][C+1] current frame column, written in place
    int out_col;
    int n_cols;                // C+1
    const int32_t* ras_window; // [B][RAS_WIN] previous main tokens or nullptr (prefill frame: no RAS)
    int sem_begin, sem_end, im_end, codebook_size;
    int32_t* done;             // [B] EOS flag (mode 1) or nullptr
    const bf16* fast_emb;      // [cb][D]: next fast-AR input row gathered here (llama.py fast_embeddings), or nullptr
    bf16* xfast;               // [B][D]
    uint32_t* xfast_t;         // flag-in-data copy for the persistent frame kernel (written instead of xfast), or nullptr
    int D;
};

__device__ __forceinline__ uint32_t order_key(uint16_t b) {
    if (b == 0x8000u) b = 0;  // -0 == +0 for comparisons
    return (b & 0x8000u) ? (uint32_t)(uint16_t)~b : (uint32_t)(b | 0x8000u);
}
__device__ __forceinline__ uint16_t key_to_bits(uint32_t k) {
    return (k & 0x8000u) ? (uint16_t)(k & 0x7fffu) : (uint16_t)~k;
}

// Philox4x32-10 (Salmon et al.), counter = (idx, draw, frame, request), key = seed
__device__ __forceinline__ uint32_t philox_u32(unsigned long long seed, uint32_t c0, uint32_t c1, uint32_t c2,
                                               uint32_t c3) {
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        c0 = hi1 ^ c1 ^ k0; c1 = lo1; c2 = hi0 ^ c3 ^ k1; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return c0;
}

// visit every scanned element: f(index, bf16 bits of logit(+bias))
template <int NT, typename F>
__device__ __forceinline__ void sample_scan(const SampleArgs& a, const bf16* lg, int tid, F f) {
    for (int s = 0; s < 2; ++s) {
        const int lo = a.seg_lo[s], hi = a.seg_hi[s];
        if (hi <= lo) continue;
        const int alo = min(hi, (lo + 7) & ~7), ahi = max(alo, hi & ~7);
        for (int i = lo + tid; i < alo; i += NT) {  // unaligned head
            float v = lda(&lg[i]);
            if (a.bias) v = rbf(v + __bfloat162float(a.bias[i]));
            f(i, f_to_bf16_bits(v));
        }
        for (int c = (alo >> 3) + tid; c < (ahi >> 3); c += NT) {  // 128-bit body
            float v[8];
            unpack8(lda(&reinterpret_cast<const uint4*>(lg)[c]), v);
            if (a.bias) {
                float bb[8];
                unpack8(reinterpret_cast<const uint4*>(a.bias)[c], bb);
#pragma unroll
                for (int i = 0; i < 8; ++i) v[i] = rbf(v[i] + bb[i]);
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) f(c * 8 + i, f_to_bf16_bits(v[i]));
        }
        for (int i = ahi + tid; i < hi; i += NT) {  // tail
            if (i < alo) continue;
            float v = lda(&lg[i]);
            if (a.bias) v = rbf(v + __bfloat162float(a.bias[i]));
            f(i, f_to_bf16_bits(v));
        }
    }
}

// group-wide exclusive prefix sum of one int per thread (SY::NT threads); returns exclusive prefix, total in *total
template <class SY>
__device__ __forceinline__ int block_excl_scan(int v, int* wsum, int* total, int tid) {
    const int lane = tid & 31, w = tid >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    SY::sync();
    if (lane == 31) wsum[w] = inc;
    SY::sync();
    if (w == 0) {
        int s = (lane < SY::NT / 32) ? wsum[lane] : 0;
        int si = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, si, o);
            if (lane >= o) si += t;
        }
        wsum[lane] = si - s;  // exclusive warp offsets
        if (lane == 31) wsum[32] = si;
    }
    SY::sync();
    *total = wsum[32];
    return wsum[w] + inc - v;
}

// one request (row b of nB) on SY::NT threads; tid = index inside the group
template <class SY>
__device__ __forceinline__ void sample_body(const SampleArgs& a, int b, int nB, int tid, unsigned xtag = 0) {
    constexpr int NT = SY::NT;
    __shared__ uint32_t hist[2048];
    __shared__ unsigned long long cand[SAMPLE_CAND_CAP];
    __shared__ unsigned long long top[SAMPLE_MAXK];
    __shared__ float red[32];
    __shared__ int wsum[33];
    __shared__ int sh_i[8];
    __shared__ float probs[SAMPLE_MAXK], cums[SAMPLE_MAXK];
    __shared__ int sh_tok[2];

    const int lane = tid & 31, warp = tid >> 5;
    const bf16* lg = a.logits + (size_t)b * a.ld_logits;

    // ---- pass A: 11-bit histogram of the order key over finite entries, running max
    for (int i = tid; i < 2048; i += NT) hist[i] = 0;
    if (tid < 8) sh_i[tid] = 0;
    SY::sync();
    uint32_t kmax = 0;
    int nfin = 0;
    sample_scan<NT>(a, lg, tid, [&](int, uint16_t bits) {
        if (bits != 0xFF80u) {
            const uint32_t k = order_key(bits);
            atomicAdd(&hist[k >> 5], 1u);
            kmax = max(kmax, k);
            ++nfin;
        }
    });
    kmax = __reduce_max_sync(0xffffffffu, kmax);
    if (lane == 0) wsum[warp] = (int)kmax;
    SY::sync();
    kmax = (lane < NT / 32) ? (uint32_t)wsum[lane] : 0u;
    kmax = __reduce_max_sync(0xffffffffu, kmax);
    int nfin_total;
    block_excl_scan<SY>(nfin, wsum, &nfin_total, tid);
    const int k_eff = min(min(a.row_top_k ? a.row_top_k[b] : a.top_k, SAMPLE_MAXK), nfin_total);
    // no finite logit at all: the reference's softmax is NaN everywhere and its argmax returns index 0; the token is
    // still written through the common output section (code, EOS flag, the next fast-AR input row and its tag — a
    // consumer of the persistent frame kernel waits for that row)
    const bool degenerate = k_eff <= 0;
    if (degenerate) {
        if (warp != 0) return;
        if (lane == 0) { sh_tok[0] = 0; sh_tok[1] = 0; }
        __syncwarp();
    } else {
    const float maxv = bf16_bits_to_f(key_to_bits(kmax));

    // ---- locate the 11-bit bin holding the k-th largest key (suffix scan from the top bin)
    {
        constexpr int BPT = 2048 / NT;  // bins per thread, highest bins first
        const int bh = 2047 - BPT * tid;
        int cnt[BPT], sum = 0;
#pragma unroll
        for (int j = 0; j < BPT; ++j) { cnt[j] = (int)hist[bh - j]; sum += cnt[j]; }
        int tot;
        int cum = block_excl_scan<SY>(sum, wsum, &tot, tid);
#pragma unroll
        for (int j = 0; j < BPT; ++j) {
            if (cum < k_eff && k_eff <= cum + cnt[j]) { sh_i[0] = bh - j; sh_i[1] = cum; }
            cum += cnt[j];
        }
    }
    SY::sync();
    const int bin1 = sh_i[0], above1 = sh_i[1];
    SY::sync();

    // ---- pass B: 5-bit histogram inside that bin + softmax denominator over everything
    if (tid < 32) hist[tid] = 0;
    SY::sync();
    float den = 0.f;
    sample_scan<NT>(a, lg, tid, [&](int, uint16_t bits) {
        if (bits != 0xFF80u) {
            const uint32_t k = order_key(bits);
            if ((int)(k >> 5) == bin1) atomicAdd(&hist[k & 31], 1u);
            den += expf(bf16_bits_to_f(bits) - maxv);
        }
    });
    den = group_sum<SY>(den, red, tid);
    SY::sync();
    if (tid == 0) {
        int cum = above1, b2 = 31;
        for (; b2 >= 0; --b2) {
            if (cum + (int)hist[b2] >= k_eff) break;
            cum += (int)hist[b2];
        }
        sh_i[2] = (bin1 << 5) | b2;  // threshold key t
        sh_i[3] = cum;               // #keys > t   (< k_eff)
        sh_i[4] = (int)hist[b2];     // #keys == t
        sh_i[5] = 0;                 // candidate counter
        sh_i[6] = 0;                 // ordered-tie counter
    }
    SY::sync();
    const uint32_t tkey = (uint32_t)sh_i[2];
    const int n_gt = sh_i[3], n_eq = sh_i[4];
    const int need_eq = k_eff - n_gt;
    const bool small = (n_gt + n_eq) <= SAMPLE_CAND_CAP;

    // ---- pass C: collect candidates (composite = key << 32 | ~index: larger is earlier in sort order)
    sample_scan<NT>(a, lg, tid, [&](int i, uint16_t bits) {
        if (bits != 0xFF80u) {
            const uint32_t k = order_key(bits);
            if (k > tkey || (small && k == tkey)) {
                const int p = atomicAdd(&sh_i[5], 1);
                cand[p] = ((unsigned long long)k << 32) | (uint32_t)(~(uint32_t)i);
            }
        }
    });
    SY::sync();
    if (!small) {
        // massive tie at the threshold: take the `need_eq` lowest indices in index order
        int taken = 0;
        for (int s = 0; s < 2 && taken < need_eq; ++s) {
            for (int base = a.seg_lo[s]; base < a.seg_hi[s] && taken < need_eq; base += NT) {
                const int i = base + tid;
                int eq = 0;
                if (i < a.seg_hi[s]) {
                    float v = lda(&lg[i]);
                    if (a.bias) v = rbf(v + __bfloat162float(a.bias[i]));
                    const uint16_t bits = f_to_bf16_bits(v);
                    eq = (bits != 0xFF80u && order_key(bits) == tkey) ? 1 : 0;
                }
                int tot;
                const int ex = block_excl_scan<SY>(eq, wsum, &tot, tid);
                if (eq && taken + ex < need_eq)
                    cand[n_gt + taken + ex] = ((unsigned long long)tkey << 32) | (uint32_t)(~(uint32_t)i);
                taken += tot;
                SY::sync();
            }
        }
    }
    SY::sync();
    const int n_cand = small ? sh_i[5] : k_eff;

    // ---- rank by counting (composites are unique), keep the first k_eff
    for (int c = tid; c < n_cand; c += NT) {
        const unsigned long long me = cand[c];
        int rank = 0;
        for (int j = 0; j < n_cand; ++j) rank += (cand[j] > me) ? 1 : 0;
        if (rank < k_eff) top[rank] = me;
    }
    SY::sync();
    if (warp != 0) return;

    // ---- warp 0: the <= 64 survivors
    float val[2], pr[2];
    int idx[2];
#pragma unroll
    for (int s = 0; s < 2; ++s) {
        const int j = lane + 32 * s;
        val[s] = 0.f; idx[s] = 0x7fffffff; pr[s] = 0.f;
        if (j < k_eff) {
            const unsigned long long c = top[j];
            val[s] = bf16_bits_to_f(key_to_bits((uint32_t)(c >> 32)));
            idx[s] = (int)(~(uint32_t)c);
            pr[s] = rbf(__fdiv_rn(expf(val[s] - maxv), den));  // first softmax, bf16
            probs[j] = pr[s];
        }
    }
    __syncwarp();
    if (lane == 0) {  // torch.cumsum: fp32 running sum, each output rounded to bf16
        float c = 0.f;
        for (int j = 0; j < k_eff; ++j) { c += probs[j]; cums[j] = rbf(c); }
    }
    __syncwarp();

    const float tmin = rbf(1e-5f);
    for (int v = 0; v < a.n_variants; ++v) {
        const float T = fmaxf((v == 0 && a.row_temp) ? a.row_temp[b] : a.temp[v], tmin);
        const float P = (v == 0 && a.row_top_p) ? a.row_top_p[b] : a.top_p[v];
        bool keep[2];
        float l2[2];
        float m2 = -INFINITY;
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            const int j = lane + 32 * s;
            keep[s] = (j < k_eff) && (j == 0 || !(cums[j] > P));
            l2[s] = keep[s] ? rbf(__fdiv_rn(val[s], T)) : -INFINITY;
            m2 = fmaxf(m2, l2[s]);
        }
        m2 = warp_max(m2);
        float e2[2], s2 = 0.f;
#pragma unroll
        for (int s = 0; s < 2; ++s) { e2[s] = keep[s] ? expf(l2[s] - m2) : 0.f; s2 += e2[s]; }
        s2 = warp_sum(s2);
        float best = -1.f;
        int besti = 0x7fffffff;
        const long long fr = a.frame_counter ? (long long)(*a.frame_counter) : 0;
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            if (!keep[s]) continue;
            const float p = rbf(__fdiv_rn(e2[s], s2));
            float u;
            if (a.uniforms) {
                u = __bfloat162float(a.uniforms[((size_t)fr * nB + b) * a.u_stride + a.u_offset +
                                                (size_t)v * a.V + idx[s]]);
            } else {
                const uint32_t r = philox_u32(a.seed, (uint32_t)idx[s], (uint32_t)(a.draw_id * 2 + v), (uint32_t)fr,
                                              (uint32_t)b);
                u = (float)(r & 0xffu) * (1.f / 256.f);  // bf16 rand_like support: multiples of 1/256 in [0,1)
            }
            const float q = -rbf(logf(u));               // u == 0 -> +inf -> score 0
            const float sc = rbf(__fdiv_rn(p, q));
            if (sc > best || (sc == best && idx[s] < besti)) { best = sc; besti = idx[s]; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
            if (ob > best || (ob == best && oi < besti)) { best = ob; besti = oi; }
        }
        if (!(best > 0.f)) besti = 0;  // every score is 0: argmax returns the first element of the whole vector
        if (lane == 0) sh_tok[v] = besti;
    }
    __syncwarp();
    }  // !degenerate

    // ---- outputs
    int tok = sh_tok[0];
    if (a.mode == 2) {  // stage entry point (fsb_sample): both draws, nothing else
        if (lane == 0) {
            a.tokens[(size_t)b * 2 + 0] = sh_tok[0];
            a.tokens[(size_t)b * 2 + 1] = (a.n_variants == 2) ? sh_tok[1] : -1;
        }
        return;
    }
    if (a.mode == 1) {
        if (a.n_variants == 2 && a.ras_window != nullptr) {  // inference.py:125-144
            bool in_win = false;
            if (lane < RAS_WIN) in_win = a.ras_window[(size_t)b * RAS_WIN + lane] == tok;
            in_win = __any_sync(0xffffffffu, in_win);
            const bool is_sem = tok >= a.sem_begin && tok <= a.sem_end;
            if (in_win && is_sem) tok = sh_tok[1];
        }
        int code = tok - a.sem_begin;  // inference.py:151-155 (also when tok is <|im_end|>)
        code = max(0, min(code, a.codebook_size - 1));
        if (lane == 0) {
            a.tokens[(size_t)b * a.n_cols + 0] = tok;
            a.tokens[(size_t)b * a.n_cols + 1] = code;
            if (a.done && tok == a.im_end) a.done[b] = 1;
        }
        tok = code;
    } else if (lane == 0) {
        a.tokens[(size_t)b * a.n_cols + a.out_col] = tok;
    }
    if (a.fast_emb != nullptr) {  // next fast-AR input = fast_embeddings(code)
        const uint4* src = reinterpret_cast<const uint4*>(a.fast_emb + (size_t)tok * a.D);
        if (a.xfast_t != nullptr) {
            for (int c = lane; c < (a.D >> 3); c += 32) {
                const uint4 v = src[c];
                uint32_t* d = a.xfast_t + (size_t)b * a.D + (size_t)c * 8;
                st_tagged4(d, v.x, v.y, xtag);
                st_tagged4(d + 4, v.z, v.w, xtag);
            }
        } else {
            uint4* dst = reinterpret_cast<uint4*>(a.xfast + (size_t)b * a.D);
            for (int c = lane; c < (a.D >> 3); c += 32) dst[c] = src[c];
        }
    }
}
__global__ void __launch_bounds__(SAMPLE_THREADS, 1) sample_kernel(SampleArgs a) {
    pdl_wait();
    pdl_launch_dependents();
    sample_body<SyncBlock<SAMPLE_THREADS>>(a, blockIdx.x, gridDim.x, threadIdx.x);
}

// ----------------------------------------------------------------------------- frame epilogue (decode_n_tokens bookkeeping)
struct FrameEndArgs {
    int32_t* tokens;       // [B][C+1]
    int32_t* pos;          // [B]
    int32_t* ras_window;   // [B][RAS_WIN] or nullptr
    int32_t* out_tokens;   // [max_frames][B][C+1] or nullptr
    int32_t* frame_counter;
    int B, n_cols, max_frames;
};
// inference.py:224-230: input_pos += 1; previous_tokens.roll(-1); previous_tokens[:, -1] = next_token
template <class SY>
__device__ __forceinline__ void frame_end_body(const FrameEndArgs& a, int tid) {
    const int fr = *a.frame_counter;
    for (int b = tid; b < a.B; b += SY::NT) {
        a.pos[b] += 1;
        if (a.ras_window) {
            int32_t* w = a.ras_window + (size_t)b * RAS_WIN;
            for (int i = 0; i + 1 < RAS_WIN; ++i) w[i] = w[i + 1];
            w[RAS_WIN - 1] = lda(&a.tokens[(size_t)b * a.n_cols]);
        }
        if (a.out_tokens && fr < a.max_frames)
            for (int c = 0; c < a.n_cols; ++c)
                a.out_tokens[((size_t)fr * a.B + b) * a.n_cols + c] = lda(&a.tokens[(size_t)b * a.n_cols + c]);
    }
    SY::sync();
    if (tid == 0) *a.frame_counter = fr + 1;
}
__global__ void __launch_bounds__(64) frame_end_kernel(FrameEndArgs a) {
    pdl_wait();
    pdl_launch_dependents();
    frame_end_body<SyncBlock<64>>(a, threadIdx.x);
}
