// Dual-AR decode engine: handle, launch plumbing (PDL-chained kernels, CUDA-graph replay) and the C ABI.
// See include/fish_b200.h for the boundary and DESIGN.md for the kernel inventory.
#include <cublas_v2.h>
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/fish_b200.h"
#include "attn.cuh"
#include "gemv.cuh"
#include "frame_mega.cuh"
#include "gemv_tma.cuh"
#include "prefill.cuh"
#include "rows_gemm.cuh"
#include "sample.cuh"
#include "tc_conv.cuh"
#include "fa_prefill.cuh"

namespace {

struct LayerW {
    const bf16 *attn_norm = nullptr, *wqkv = nullptr, *q_norm = nullptr, *k_norm = nullptr, *wo = nullptr,
               *ffn_norm = nullptr, *w1 = nullptr, *w3 = nullptr, *w2 = nullptr;
    bf16 *kc = nullptr, *vc = nullptr;
};

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
constexpr size_t MEGA_BUF_BYTES = 512 * 1024;
constexpr size_t CUBLAS_WS_BYTES = 32u << 20;
constexpr int RG_SM_MAX = 160;                           // SMs the stream-K partial slabs are sized for
constexpr int RG_GRID_MAX = RG_CTAS_PER_SM * RG_SM_MAX;   // (the constant is defined in rows_gemm.cuh, included below)
constexpr size_t MEGA_TRACE_WORDS = 160 * 512 * 4;  // [CTAs][ops][4] time stamps of one frame (profiling aid)

struct GemvCfg {
    int n_sm = 148;
    size_t self_pf = 512 * 1024;
    int tma = 1;          // 1: bulk-async-copy staged kernel (gemv_tma.cuh), 0: direct-load kernel (gemv.cuh)
    int nst = 4;          // ring depth of the staged kernel (4 x 40 KB = 160 KB in flight per SM)
    int ctas_per_sm = 1;  // grid = ctas_per_sm * n_sm
    int stage_kb = 40;    // target stage size: 8 row slices of 5 KB (bulk copies below ~4 KB are issue-bound)
};

}  // namespace

struct fsb_dual_ar {
    fsb_dual_ar_desc d;
    int max_seq_len = 0;
    std::vector<LayerW> slow, fast;
    const bf16 *emb = nullptr, *cb_emb = nullptr, *norm = nullptr, *fast_emb = nullptr, *fast_norm = nullptr,
               *fast_out = nullptr, *out_w = nullptr;
    const bf16 *rope = nullptr, *fast_rope = nullptr;
    int rope_n = 0, fast_rope_n = 0;
    int kv_max_batch = 0, kv_max_seq = 0, fkv_max_batch = 0, fkv_max_seq = 0;
    // workspace carve-up
    bf16 *x = nullptr, *x2 = nullptr, *qkv = nullptr, *y = nullptr, *ff = nullptr, *hidden = nullptr,
         *xfast = nullptr, *xf = nullptr, *xf2 = nullptr, *slow_logits = nullptr, *fast_logits = nullptr;
    float *ws_acc = nullptr, *ws_ml = nullptr;
    int32_t *counters = nullptr, *pf_pos = nullptr, *pf_slot = nullptr;
    char* mega_buf = nullptr;        // device copy of the persistent frame kernel's op list (MEGA_BUF_BYTES)
    unsigned* mega_sync = nullptr;   // grid barrier counters
    uint2 *ll_acc = nullptr, *ll_ml = nullptr;  // {fp32, tag} split-attention partials of the frame kernel
    uint32_t* tg_base = nullptr;     // flag-in-data activations: x, x2, qkv, y, ff, hidden, xfast, xf, xf2
    size_t tg_words = 0;
    unsigned long long* mega_prof = nullptr;  // cycle counters of CTA 0 when option "mega_prof" is on
    int mega_prof_on = 0;
    unsigned long long* mega_trace = nullptr;
    int mega_trace_on = 0;
    // prompt pass scratch (PREFILL_ROWS rows each) and the cuBLAS handle used for its plain GEMMs
    bf16 *pxa = nullptr, *pxb = nullptr, *pxn = nullptr, *pqkv = nullptr, *py = nullptr, *ptmp = nullptr, *pf1 = nullptr,
         *pf3 = nullptr;
    bf16 *pq = nullptr, *pvt = nullptr;  // flash attention of the prompt pass: normed + roped q [H][rows][hd], V^T [KV][hd][max_seq]
    int prefill_fa = 1;                  // prompt-pass attention on fa_prefill_kernel (tcgen05) when head_dim = 128; 0: one decode-attention CTA per row
    std::map<std::tuple<const void*, uint64_t, uint64_t, uint64_t>, CUtensorMap> fa_maps;
    char* cublas_ws = nullptr;
    cublasHandle_t blas = nullptr;
    int prefill_gemm = 1;            // prompt pass: 1 = whole chunks, every Linear on this library's tcgen05 kernel (tc_conv.cuh: tcv_kernel),
                                     // 2 = whole chunks with cuBLAS GEMMs (comparison only), 0 = 8-row passes through the decode kernels
    int batch_gemm = 1;              // decode steps with more than 8 rows: 1 = every Linear once over all rows (weights streamed once
                                     // per step) on this library's tcgen05 stream-K kernel (rows_gemm.cuh), 2 = the same through
                                     // cuBLAS, 0 = 8-row passes through the GEMV kernels (weights re-streamed per pass)
    float* rg_partial = nullptr;     // stream-K partial tiles of rows_gemm_kernel
    float* pks = nullptr;            // split-K partial sums of the prompt pass (PREFILL_KS_ROWS x dim)
    int masked_head = 1;             // fused frames with the constrained-decode bias: compute only the LM-head rows the sampler can
                                     // read (semantic range + <|im_end|>) instead of all `vocab` rows; batch 1 paths
    int rows_narrow = 0;             // many-row Linears with a plain store on narrow weight tiles without a k split
                                     // (rows_gemm_narrow_kernel); measured slower than the 3-way split + folded finish: off
    int qkv_fold = 1;                // many-row steps: q / k / v partial tiles summed by their readers instead of a finish launch
    int attn_tail = 1;               // attn_tc_kernel: cut the units of the last, partial round of CTAs finer
    int attn_trace_point = 0;        // profiling: which point of attn_split_body the frame kernel's trace records
    int attn_trace = 0;              // profiling: attn_tc_kernel writes 4 clock stamps per CTA into the mega_trace buffer
    int attn_tc = 1;                 // many-row steps: tensor-core decode attention (attn_tc_kernel) instead of attn_decode_kernel
    bool rg_ok = false;              // tensor-map encoder available, attributes set
    std::map<std::tuple<const void*, int, int, int>, CUtensorMap> rg_maps;  // (ptr, rows, cols, box rows) -> TMA map
    std::map<std::tuple<const void*, int, int, int>, CUtensorMap> tcv_maps; // the same for tcv_kernel's 3-D maps (prompt pass)
    int max_splits = 32;
    // options
    int use_pdl = 1, use_graph = 1, attn_splits = 0;
    int use_mega = 1;                // batch-1 frames as ONE persistent cooperative kernel (frame_mega.cuh)
    int rows_min = 2;                // fewest rows that take the many-row path (every Linear once over all rows on the tcgen05 kernel).
                                     // Measured at 2k context: 2 rows 5.64 ms per step against 7.59 through the GEMV chain's 8-row
                                     // pass, 4 rows 5.71 / 10.6, 8 rows 6.27 / 22.9 - the chain (rows_min = 9) is kept for comparison
    int chain_attn_tc = 2048;        // kernel chain: caches longer than this use the tensor-core decode attention (0 = never)
    int prefill_rows = 1;            // prompt pieces of <= 2 x max_rows rows on handles with >= 16 rows go through the many-row decode kernels
    int prefill_splitk = 1;          // prompt pass: wo / w2 of a short chunk as split-K pieces + a finish kernel
    int mega_long = 0;               // caches longer than ASP_MAXS x asp_max_keys tokens: 0 = kernel chain (faster as measured: 6.04
                                     // vs 6.29 ms per frame at 8k context), 1 = tiled frame kernel, 2 = tiled kernel at any length (tests)
    int mega_skip_dot = 0;           // experiment (results are garbage): no dot products in the frame kernel
    int mega_detail_op = -1;         // profiling: op index whose per-iteration stamps are recorded (frame kernel)
    int mega_l2_ahead = 0;           // frame kernel: stages per producer lane prefetched into L2 while the ring is full
    bool mega_supported = false;     // shape fits the frame kernel's stage geometry and it is resident at 1 CTA/SM
    MegaPlan mega_plan{};
    std::vector<char> mega_host;
    GemvCfg gc;
    int n_sm = 148;
    // graph cache
    cudaGraphExec_t exec = nullptr;
    cudaStream_t cap_stream = nullptr;  // capture happens here: the caller's stream may be the legacy default stream
    size_t R = 8;                       // rows the workspace is sized for
    fsb_frame_args cached{};
    bool cached_valid = false;
    std::string err;
};

namespace {

int fail(fsb_dual_ar* h, const std::string& m) {
    if (h) h->err = m;
    return 1;
}
#define CK(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess)                                                                           \
            return fail(h, std::string(#call) + ": " + cudaGetErrorString(e_) + " @" + std::to_string(__LINE__)); \
    } while (0)

template <typename Args>
cudaError_t launch(void (*k)(Args), dim3 grid, dim3 block, size_t smem, cudaStream_t s, bool pdl, const Args& a) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, k, a);
}

// ------------------------------------------------------------------------------------------- GEMV dispatch
typedef void (*gemv_fn)(GemvArgs);
template <int PRO, int EPI>
gemv_fn pick_gemv(int M) {
    constexpr bool SW = (EPI == EPI_SWIGLU);
    if (M <= 1) return gemv_kernel<1, SW ? 4 : 8, PRO, EPI>;
    if (M <= 2) return gemv_kernel<2, SW ? 2 : 4, PRO, EPI>;
    if (M <= 4) return gemv_kernel<4, SW ? 2 : 4, PRO, EPI>;
    return gemv_kernel<8, SW ? 1 : 2, PRO, EPI>;
}
gemv_fn pick_gemv(int M, int pro, int epi) {
    if (pro == PRO_RMSNORM) {
        if (epi == EPI_STORE) return pick_gemv<PRO_RMSNORM, EPI_STORE>(M);
        if (epi == EPI_RESIDUAL) return pick_gemv<PRO_RMSNORM, EPI_RESIDUAL>(M);
        return pick_gemv<PRO_RMSNORM, EPI_SWIGLU>(M);
    }
    if (epi == EPI_STORE) return pick_gemv<PRO_NONE, EPI_STORE>(M);
    if (epi == EPI_RESIDUAL) return pick_gemv<PRO_NONE, EPI_RESIDUAL>(M);
    return pick_gemv<PRO_NONE, EPI_SWIGLU>(M);
}
int m_bucket(int M) { return M <= 1 ? 1 : M <= 2 ? 2 : M <= 4 ? 4 : 8; }

cudaError_t set_gemv_attrs() {
    FsbOnce once(0);
    if (!once.first) return cudaSuccess;
    const int ms[4] = {1, 2, 4, 8};
    for (int pro = 0; pro < 2; ++pro)
        for (int epi = 0; epi < 3; ++epi)
            for (int i = 0; i < 4; ++i) {
                cudaError_t e = cudaFuncSetAttribute((const void*)pick_gemv(ms[i], pro, epi),
                                                     cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
                if (e != cudaSuccess) return e;
            }
    return cudaSuccess;
}

typedef void (*gemv_tma_fn)(GemvArgs, GemvTmaGeom);
template <int PRO, int EPI>
gemv_tma_fn pick_gemv_tma(int M) {
    if (M <= 1) return gemv_tma_kernel<1, PRO, EPI>;
    if (M <= 2) return gemv_tma_kernel<2, PRO, EPI>;
    if (M <= 4) return gemv_tma_kernel<4, PRO, EPI>;
    return gemv_tma_kernel<8, PRO, EPI>;
}
gemv_tma_fn pick_gemv_tma(int M, int pro, int epi) {
    if (pro == PRO_RMSNORM) {
        if (epi == EPI_STORE) return pick_gemv_tma<PRO_RMSNORM, EPI_STORE>(M);
        if (epi == EPI_RESIDUAL) return pick_gemv_tma<PRO_RMSNORM, EPI_RESIDUAL>(M);
        return pick_gemv_tma<PRO_RMSNORM, EPI_SWIGLU>(M);
    }
    if (epi == EPI_STORE) return pick_gemv_tma<PRO_NONE, EPI_STORE>(M);
    if (epi == EPI_RESIDUAL) return pick_gemv_tma<PRO_NONE, EPI_RESIDUAL>(M);
    return pick_gemv_tma<PRO_NONE, EPI_SWIGLU>(M);
}
cudaError_t set_gemv_tma_attrs() {
    FsbOnce once(1);
    if (!once.first) return cudaSuccess;
    const int ms[4] = {1, 2, 4, 8};
    for (int pro = 0; pro < 2; ++pro)
        for (int epi = 0; epi < 3; ++epi)
            for (int i = 0; i < 4; ++i) {
                cudaError_t e = cudaFuncSetAttribute((const void*)pick_gemv_tma(ms[i], pro, epi),
                                                     cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
                if (e != cudaSuccess) return e;
            }
    return cudaSuccess;
}
// largest KC <= limit that divides K and is a multiple of 8 (0 if none)
int pick_kc(int K, int limit) {
    if (K <= limit) return K;
    for (int kc = limit - (limit % 8); kc >= 64; kc -= 8)
        if (K % kc == 0) return kc;
    return 0;
}

// rows M are processed in passes of <= 8; x rows have stride K, out/res rows stride N
cudaError_t run_gemv(const GemvCfg& gc, bool pdl, cudaStream_t s, int M, int N, int K, int pro, int epi,
                     const bf16* W, const bf16* W2, const bf16* x, const bf16* norm_w, const bf16* res, bf16* out,
                     bf16* xnorm_out, float eps, const void* pf_ptr, size_t pf_bytes) {
    for (int m0 = 0; m0 < M; m0 += 8) {
        const int mv = std::min(8, M - m0);
        const int mb = m_bucket(mv);
        GemvArgs a{};
        a.W = W; a.W2 = W2; a.x = x + (size_t)m0 * K; a.norm_w = norm_w;
        a.res = res ? res + (size_t)m0 * N : nullptr;
        a.out = out + (size_t)m0 * N;
        a.xnorm_out = xnorm_out ? xnorm_out + (size_t)m0 * K : nullptr;
        a.N = N; a.K = K; a.M_valid = mv; a.eps = eps;
        a.pf_ptr = (m0 == 0) ? pf_ptr : nullptr; a.pf_bytes = pf_bytes; a.self_pf_bytes = gc.self_pf;
        const size_t xsmem = (size_t)mb * K * sizeof(bf16);
        if (gc.tma) {
            const int nacc = (epi == EPI_SWIGLU) ? 2 : 1;
            const int kc = pick_kc(K, std::min(gc.stage_kb * 1024 / (GT_ROWS * 2), 32 * 8 * GT_CPL));
            (void)nacc;
            int nst = std::max(2, std::min(gc.nst, GT_MAX_STAGES));
            if (kc > 0 && xsmem < 200 * 1024)
                nst = std::min<int>(nst, (int)((200 * 1024 - xsmem) / ((size_t)GT_ROWS * kc * 2)));
            const size_t smem = (size_t)nst * GT_ROWS * kc * 2 + xsmem;
            if (kc > 0 && nst >= 2 && smem <= 200 * 1024) {
                GemvTmaGeom g{kc, nst};
                const int nb = std::min(gc.ctas_per_sm * gc.n_sm, std::max(1, N / GT_ROWS));
                cudaLaunchConfig_t cfg{};
                cfg.gridDim = dim3(nb); cfg.blockDim = dim3(GT_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = s;
                cudaLaunchAttribute at[1];
                at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
                at[0].val.programmaticStreamSerializationAllowed = 1;
                cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
                cudaError_t e = cudaLaunchKernelEx(&cfg, pick_gemv_tma(mv, pro, epi), a, g);
                if (e != cudaSuccess) return e;
                continue;
            }
        }
        int nb = std::min(2 * gc.n_sm, std::max(1, N / 4));
        cudaError_t e = launch(pick_gemv(mv, pro, epi), dim3(nb), dim3(GEMV_THREADS), xsmem, s, pdl, a);
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

// ------------------------------------------------------------------------------------------- attention dispatch
typedef void (*attn_fn)(AttnArgs);
template <int HD, int GQ>
attn_fn attn_tc_pick() {
    if constexpr (HD == 128) return attn_tc_kernel<HD, GQ>;
    else return nullptr;
}
struct AttnKernels { attn_fn decode, append, fast, tc; int threads_fast; };  // tc: tensor-core many-row decode (head_dim 128) or nullptr
bool pick_attn(int hd, int gq, AttnKernels* k) {
#define CASE(HD, GQ)                                                               \
    if (hd == HD && gq == GQ) {                                                    \
        k->decode = attn_decode_kernel<HD, GQ>;                                    \
        k->append = rope_append_kernel<HD>;                                        \
        k->fast = fast_attn_kernel<HD, GQ>;                                        \
        k->tc = attn_tc_pick<HD, GQ>();                                            \
        k->threads_fast = 32 * GQ;                                                 \
        return true;                                                               \
    }
    CASE(128, 4) CASE(64, 4) CASE(128, 1) CASE(64, 1) CASE(128, 2) CASE(64, 2) CASE(128, 8) CASE(64, 8)
#undef CASE
    return false;
}

int splits_for(const fsb_dual_ar* h, int M) {
    if (h->attn_splits > 0) return std::min(h->attn_splits, h->max_splits);
    // enough CTAs to cover the chip (about 12 resident CTAs per SM keep ~80 KB of key/value rows in flight per SM),
    // tiles of >= 128 keys at the longest context
    int by_len = std::max(1, (h->max_seq_len + 255) / 256);
    int by_sm = std::max(1, (12 * h->n_sm + M * h->d.n_kv_head - 1) / std::max(1, M * h->d.n_kv_head));
    return std::max(1, std::min(std::min(by_len, by_sm), h->max_splits));
}

// ------------------------------------------------------------------------------------------- op argument builders
AttnArgs make_attn_args(const fsb_dual_ar* h, const LayerW& w, bool fast, const int32_t* pos, const int32_t* slots,
                        int pos_fixed) {
    const fsb_dual_ar_desc& d = h->d;
    AttnArgs a{};
    a.qkv = h->qkv; a.kcache = w.kc; a.vcache = w.vc; a.q_norm = w.q_norm; a.k_norm = w.k_norm;
    a.rope = fast ? h->fast_rope : h->rope;
    a.pos = pos; a.slots = slots; a.y = h->y; a.ws_acc = h->ws_acc; a.ws_ml = h->ws_ml;
    a.counters = h->counters; a.H = d.n_head; a.KV = d.n_kv_head;
    a.S = fast ? h->fkv_max_seq : h->kv_max_seq;
    a.qk_norm = d.qk_norm; a.eps = d.norm_eps; a.scale = (float)(1.0 / std::sqrt((double)d.head_dim));
    a.pos_fixed = pos_fixed;
    a.splits = 1; a.fused_append = 1;
    return a;
}

EmbedArgs make_embed_args(const fsb_dual_ar* h, const int32_t* tokens) {
    const fsb_dual_ar_desc& d = h->d;
    EmbedArgs e{};
    e.tokens = tokens; e.emb = h->emb; e.cb_emb = h->cb_emb; e.x = h->x2;  // x2: layer 0 reads it as x_in
    e.D = d.dim; e.C = d.num_codebooks; e.cb = d.codebook_size;
    e.sem_begin = d.semantic_begin_id; e.sem_end = d.semantic_end_id; e.scale = d.scale_codebook_embeddings;
    e.inv_div = (float)std::sqrt((double)(d.num_codebooks + 1));
    e.pf_ptr = nullptr; e.pf_bytes = 0;
    return e;
}

// LM-head rows a fused frame has to compute: with the constrained-decode bias (inference.py:308-320) every logit outside
// [semantic_begin, semantic_end] + {im_end} is -inf after the bias add, so the sampler (restrict_scan) never reads it.
// One contiguous 8-aligned row range covers both (im_end sits 35 rows below the semantic range at the S2 ids).
void head_rows(const fsb_dual_ar_desc& d, bool restricted, int* lo, int* n) {
    if (!restricted) { *lo = 0; *n = d.vocab; return; }
    const int a = std::min(d.im_end_id, d.semantic_begin_id) & ~7;
    const int b = std::min(d.vocab, (std::max(d.im_end_id, d.semantic_end_id) + 1 + 7) & ~7);
    *lo = a; *n = b - a;
}
bool rows_tc_layers(const fsb_dual_ar* h, int M);
bool head_restricted(const fsb_dual_ar* h, const fsb_frame_args& fa) {
    return h->masked_head && fa.restrict_scan && fa.logit_bias && (fa.batch == 1 || rows_tc_layers(h, fa.batch));
}

// sampling of the main (slow) token: two draws + RAS selection + code 0 (inference.py:118-155)
SampleArgs make_slow_sample(const fsb_dual_ar* h, const fsb_frame_args& fa) {
    const fsb_dual_ar_desc& d = h->d;
    const int C = d.num_codebooks, V = d.vocab, cb = d.codebook_size;
    SampleArgs sa{};
    sa.logits = h->slow_logits; sa.ld_logits = V; sa.V = V;
    sa.bias = (const bf16*)fa.logit_bias;
    if (fa.restrict_scan && fa.logit_bias) {
        // finite support of semantic_logit_bias (inference.py:308-320), ascending
        int lo0 = d.im_end_id, hi0 = d.im_end_id + 1, lo1 = d.semantic_begin_id, hi1 = d.semantic_end_id + 1;
        if (lo0 >= lo1 && lo0 < hi1) { lo0 = hi0 = 0; }
        else if (lo0 > lo1) { std::swap(lo0, lo1); std::swap(hi0, hi1); }
        sa.seg_lo[0] = lo0; sa.seg_hi[0] = hi0; sa.seg_lo[1] = lo1; sa.seg_hi[1] = hi1;
    } else {
        sa.seg_lo[0] = 0; sa.seg_hi[0] = V; sa.seg_lo[1] = sa.seg_hi[1] = 0;
    }
    sa.top_k = fa.top_k;
    sa.n_variants = 2;
    sa.temp[0] = fa.temperature; sa.top_p[0] = fa.top_p;
    sa.row_temp = fa.row_temperature; sa.row_top_p = fa.row_top_p; sa.row_top_k = fa.row_top_k;
    sa.temp[1] = 1.0f;         // RAS_HIGH_TEMP  (inference.py:50)
    sa.top_p[1] = 0.8984375f;  // RAS_HIGH_TOP_P 0.9 as bf16 (inference.py:51, tensor built in the model dtype)
    sa.uniforms = (const bf16*)fa.uniforms;
    sa.u_stride = 2LL * V + (long long)(C - 1) * cb;
    sa.u_offset = 0;
    sa.seed = fa.seed; sa.frame_counter = fa.frame_counter; sa.draw_id = 0;
    sa.mode = 1; sa.tokens = fa.cur_tokens; sa.out_col = 0; sa.n_cols = C + 1;
    sa.ras_window = fa.ras_window;
    sa.sem_begin = d.semantic_begin_id; sa.sem_end = d.semantic_end_id; sa.im_end = d.im_end_id;
    sa.codebook_size = cb; sa.done = fa.done;
    sa.fast_emb = h->fast_emb; sa.xfast = h->xfast; sa.D = d.dim;
    return sa;
}

// sampling of codebook i (1 <= i < C) from the fast logits (inference.py:157-176)
SampleArgs make_fast_sample(const fsb_dual_ar* h, const fsb_frame_args& fa, int i) {
    const fsb_dual_ar_desc& d = h->d;
    const int C = d.num_codebooks, V = d.vocab, cb = d.codebook_size;
    SampleArgs fs{};
    fs.logits = h->fast_logits + (size_t)i * h->R * cb; fs.ld_logits = cb; fs.V = cb;
    fs.bias = nullptr;
    fs.seg_lo[0] = 0; fs.seg_hi[0] = cb; fs.seg_lo[1] = fs.seg_hi[1] = 0;
    fs.top_k = fa.top_k; fs.n_variants = 1;
    fs.temp[0] = fa.temperature; fs.top_p[0] = fa.top_p; fs.temp[1] = 1.f; fs.top_p[1] = 1.f;
    fs.row_temp = fa.row_temperature; fs.row_top_p = fa.row_top_p; fs.row_top_k = fa.row_top_k;
    fs.uniforms = (const bf16*)fa.uniforms;
    fs.u_stride = 2LL * V + (long long)(C - 1) * cb; fs.u_offset = 2LL * V + (long long)(i - 1) * cb;
    fs.seed = fa.seed; fs.frame_counter = fa.frame_counter; fs.draw_id = i;
    fs.mode = 0; fs.tokens = fa.cur_tokens; fs.out_col = i + 1; fs.n_cols = C + 1;
    fs.ras_window = nullptr;
    fs.sem_begin = d.semantic_begin_id; fs.sem_end = d.semantic_end_id; fs.im_end = d.im_end_id;
    fs.codebook_size = cb; fs.done = nullptr;
    fs.fast_emb = (i + 1 < C) ? h->fast_emb : nullptr; fs.xfast = h->xfast; fs.D = d.dim;
    return fs;
}

FrameEndArgs make_frame_end(const fsb_dual_ar* h, const fsb_frame_args& fa) {
    FrameEndArgs fe{};
    fe.tokens = fa.cur_tokens; fe.pos = fa.pos; fe.ras_window = fa.ras_window; fe.out_tokens = fa.out_tokens;
    fe.frame_counter = fa.frame_counter; fe.B = fa.batch; fe.n_cols = h->d.num_codebooks + 1; fe.max_frames = fa.max_frames;
    return fe;
}

// ------------------------------------------------------------------------------------------- the passes
int ensure_blas(fsb_dual_ar* h, cudaStream_t s) {
    if (!h->blas) {
        if (cublasCreate(&h->blas) != CUBLAS_STATUS_SUCCESS) { h->blas = nullptr; return fail(h, "cublasCreate failed"); }
        cublasSetWorkspace(h->blas, h->cublas_ws, CUBLAS_WS_BYTES);  // caller-owned scratch: no hidden allocations
    }
    if (cublasSetStream(h->blas, s) != CUBLAS_STATUS_SUCCESS) return fail(h, "cublasSetStream failed");
    return 0;
}

// C[T][N] = X[T][K] . W[N][K]^T, bf16 in / fp32 accumulate / bf16 out (one rounding per Linear)
int rows_linear(fsb_dual_ar* h, const bf16* W, const bf16* X, bf16* C, int T, int N, int K) {
    const float one = 1.f, zero = 0.f;
    cublasStatus_t st = cublasGemmEx(h->blas, CUBLAS_OP_T, CUBLAS_OP_N, N, T, K, &one, W, CUDA_R_16BF, K, X, CUDA_R_16BF, K,
                                     &zero, C, CUDA_R_16BF, N, CUBLAS_COMPUTE_32F, CUBLAS_GEMM_DEFAULT);
    if (st != CUBLAS_STATUS_SUCCESS) return fail(h, "cublasGemmEx failed with status " + std::to_string((int)st));
    return 0;
}

// the many-row path
bool batch_gemm_active(const fsb_dual_ar* h, int M) { return h->batch_gemm && M >= h->rows_min && M <= PREFILL_ROWS; }
bool rows_tc_active(const fsb_dual_ar* h, int M) { return h->batch_gemm == 1 && h->rg_ok && M <= 128; }

// the whole layer stack runs on the tcgen05 kernel (every Linear's K is a multiple of the 64-column k block)
bool rows_tc_layers(const fsb_dual_ar* h, int M) {
    const fsb_dual_ar_desc& d = h->d;
    return batch_gemm_active(h, M) && rows_tc_active(h, M) && d.dim % RG_BK == 0 && d.ffn % RG_BM == 0 &&
           (d.n_head * d.head_dim) % RG_BK == 0;
}

const CUtensorMap* rows_map(fsb_dual_ar* h, const void* ptr, int rows, int cols, int box_rows) {
    auto key = std::make_tuple(ptr, rows, cols, box_rows);
    auto it = h->rg_maps.find(key);
    if (it != h->rg_maps.end()) return &it->second;
    CUtensorMap m;
    if (!rg_make_map(&m, ptr, (uint64_t)rows, (uint64_t)cols, (uint32_t)box_rows)) return nullptr;
    return &h->rg_maps.emplace(key, m).first->second;
}

// out = epi(X[M][K] . W[N][K]^T) on the tcgen05 stream-K kernel (+ the finish of its split tiles); returns -1 if the
// shape is not covered.  epi SWIGLU: W2 = up projection, `pre` = [M][2N] scratch for the pre-activation halves, out = [M][N].
// all_partial = 1 leaves every tile as fp32 partials for rows_finish_norm (then `out` is not written here).
int rows_linear_tc(fsb_dual_ar* h, cudaStream_t s, const bf16* W, const bf16* W2, const bf16* X, int M, int N, int K, int epi,
                   const bf16* res, bf16* out, bf16* pre = nullptr, int all_partial = 0, int ldo = 0) {
    const bool sw = epi == RG_EPI_SWIGLU;
    const int NV = sw ? 2 * N : N;  // rows of the virtual weight matrix [W; W2]
    if (K % RG_BK || (NV + RG_BM - 1) / RG_BM > RG_MAX_TILES || (sw && (N % RG_BM || !pre || !W2))) return -1;
    if ((long long)((NV + RG_BM - 1) / RG_BM) * (K / RG_BK) * RG_GRID_MAX >= (1LL << 31)) return -1;  // 32-bit range arithmetic
    const bool pdl0 = h->use_pdl != 0;
    if (epi == RG_EPI_STORE && !all_partial && h->rows_narrow && ldo == 0) {  // enough narrow weight tiles for one CTA per SM: no k split
        const int nw = rgn_pick_nw(N, h->n_sm);
        if (nw) {
            const CUtensorMap* nW = rows_map(h, W, N, K, nw);
            const CUtensorMap* nX = rows_map(h, X, M, K, 128);
            if (!nW || !nX) return fail(h, "cuTensorMapEncodeTiled failed");
            CK(rgn_launch(s, pdl0, *nW, *nX, M, N, K, nw, out));
            return 0;
        }
    }
    const int bn = rg_bn_for(M);
    const CUtensorMap* mW = rows_map(h, W, N, K, RG_BM);
    const CUtensorMap* mW2 = W2 ? rows_map(h, W2, N, K, RG_BM) : mW;
    const CUtensorMap* mX = rows_map(h, X, M, K, bn);
    if (!mW || !mW2 || !mX) return fail(h, "cuTensorMapEncodeTiled failed");
    const bool pdl = h->use_pdl != 0;
    const int tiles_w = (N + RG_BM - 1) / RG_BM;
    CK(rg_launch(s, pdl, h->n_sm, *mW, *mW2, *mX, M, NV, tiles_w, K, sw ? RG_EPI_STORE : epi, res, sw ? pre : out, h->rg_partial,
                 all_partial, nullptr, sw ? 0 : ldo));
    if (!all_partial) CK(rg_finish_tiles(s, pdl, h->n_sm, M, NV, K, epi, h->rg_partial, sw ? pre : out, res, out, sw ? 0 : ldo));
    return 0;
}

// every tile of out = X . W^T can stay an fp32 partial for rows_finish_norm: a CTA's range then spans at most two tiles
bool rows_all_partial_ok(const fsb_dual_ar* h, int N, int K) {
    const int n_tiles = (N + RG_BM - 1) / RG_BM;
    return n_tiles <= RF_MAX_TILES && n_tiles <= RG_CTAS_PER_SM * h->n_sm;  // tile-aligned ranges: one partial tile per CTA
}

// finish of a deferred Linear (sum of its partial tiles + residual -> x_out) and/or the RMSNorm that follows it
int rows_finish_norm(fsb_dual_ar* h, cudaStream_t s, int M, int N, int K_pending, const bf16* res, const bf16* x_in, bf16* x_out,
                     const bf16* norm_w, bf16* xn) {
    RowsFinishArgs f{};
    f.partial = K_pending > 0 ? h->rg_partial : nullptr;
    f.res = res; f.x_in = x_in; f.x_out = x_out; f.norm_w = norm_w; f.xn = xn;
    f.N = N; f.bn = rg_bn_for(M); f.kb = K_pending > 0 ? K_pending / RG_BK : 1; f.n_tiles = (N + RG_BM - 1) / RG_BM;
    f.G = rg_grid(f.n_tiles, f.kb, h->n_sm);
    f.eps = h->d.norm_eps;
    if (f.partial == nullptr && norm_w == nullptr) return 0;
    CK(launch(rows_finish_norm_kernel, dim3(M), dim3(RF_THREADS), 0, s, h->use_pdl != 0, f));
    return 0;
}

// slow-AR attention of a many-row step: K/V append + tensor-core decode attention (head_dim 128), else the fused kernel
// force_splits > 0: exactly that many splits per (row, head) unit (1 = no split workspace is touched: rows beyond the handle's
// row capacity, as in a prompt piece on a batch-1 handle)
int rows_attention(fsb_dual_ar* h, cudaStream_t s, const AttnKernels& ak, AttnArgs a, int splits, int KV, int M, bool pdl,
                   int force_splits = 0) {
    if (force_splits > 0) splits = force_splits;
    a.splits = splits;
    if (ak.tc && h->attn_tc) {
        if (h->attn_splits <= 0 && force_splits <= 0) {
            // a (split, kv head, row) item costs ~2 us of q prep / merge besides its stream, and 3 CTAs fit an SM: about two
            // rounds of CTAs measured best (64 rows x 8 kv heads: 2 splits 11.7 ms per step, 4: 12.1, 8: 13.1)
            const int by_len = std::max(1, (h->max_seq_len + 255) / 256);
            const int want = (6 * h->n_sm + M * KV - 1) / (M * KV);
            splits = std::max(1, std::min(std::min(want, by_len), h->max_splits));
            a.splits = splits;
        }
        AttnArgs ap = a;
        ap.pf_ptr = nullptr; ap.pf_bytes = 0; ap.fused_append = 0;
        CK(launch(ak.append, dim3(KV, M), dim3(64), 0, s, pdl, ap));
        a.fused_append = 0;
        // items: `splits` per (row, head) unit while they fill whole rounds of the 3 resident CTAs per SM, the remaining
        // units cut finer so that the last round is short
        const int units = M * KV, slots = 3 * h->n_sm;
        int u1 = units, s2 = splits;
        if (h->attn_splits <= 0 && force_splits <= 0 && h->attn_tail && units * splits > slots) {
            u1 = std::min(units, (units * splits / slots) * slots / splits);
            if (u1 < units) s2 = std::max(splits, std::min(h->max_splits, std::min((slots + units - u1 - 1) / (units - u1), std::max(1, (h->max_seq_len + 255) / 256))));
        }
        a.tc_u1 = u1; a.tc_s2 = s2;
        const int G = u1 * splits + (units - u1) * s2;
        a.trace = (h->attn_trace && (size_t)G * 4 <= MEGA_TRACE_WORDS) ? h->mega_trace : nullptr;
        CK(launch(ak.tc, dim3(G), dim3(ATC_THREADS), 0, s, pdl, a));
        return 0;
    }
    a.fused_append = 1;
    CK(launch(ak.decode, dim3(splits, KV, M), dim3(ATTN_THREADS), 0, s, pdl, a));
    return 0;
}

// transformer layers over 8 < M <= PREFILL_ROWS decode rows (one row per request, each with its own position and KV
// slot): every Linear runs ONCE over all rows, so a step streams the weights once however many requests are in flight.
// Same contract as run_layers (result in h->x / h->xf); scratch = the prompt-pass buffers.
// `rb`: activation buffers for more rows than the handle's row capacity (a prompt piece: the prompt-pass scratch), with
// `force_splits` = 1 so that the attention needs no per-row split workspace; tensor-core path only.
struct RowsBufs { bf16 *xa, *xb, *qkv, *y, *ff; };
int run_layers_rows(fsb_dual_ar* h, cudaStream_t s, bool fast, int M, const bf16* x_in, const int32_t* pos,
                    const int32_t* slots, int pos_fixed, const bf16* tail_norm_w, bf16* tail_xn, const RowsBufs* rb = nullptr,
                    int force_splits = 0) {
    const fsb_dual_ar_desc& d = h->d;
    const int D = d.dim, H = d.n_head, KV = d.n_kv_head, HD = d.head_dim, F = d.ffn, QKV = (H + 2 * KV) * HD;
    AttnKernels ak;
    if (!pick_attn(HD, H / KV, &ak)) return fail(h, "unsupported head_dim / GQA ratio");
    const bool tc = rows_tc_layers(h, M);
    if (!tc)
        if (int r = ensure_blas(h, s)) return r;
    const std::vector<LayerW>& L = fast ? h->fast : h->slow;
    if (rb && !tc) return fail(h, "run_layers_rows: external buffers need the tensor-core path");
    bf16* xa = rb ? rb->xa : (fast ? h->xf : h->x);
    bf16* xb = rb ? rb->xb : (fast ? h->xf2 : h->x2);
    bf16* qkv = rb ? rb->qkv : h->qkv;
    bf16* yb = rb ? rb->y : h->y;
    bf16* ffb = rb ? rb->ff : h->ff;
    const int ew_blocks = 2 * h->n_sm;
    const int splits = force_splits > 0 ? force_splits : splits_for(h, M);
    const bf16* cur = x_in;
    for (size_t l = 0; l < L.size(); ++l) {
        const LayerW& w = L[l];
        if (!w.wqkv || !w.wo || !w.w1 || !w.w2 || !w.w3 || !w.attn_norm || !w.ffn_norm || !w.kc || !w.vc)
            return fail(h, "unbound weights / kv cache in layer " + std::to_string(l));
        if (!tc) rmsnorm_rows_kernel<<<M, 256, 0, s>>>(cur, w.attn_norm, h->pxn, D, d.norm_eps);
        AttnArgs a = make_attn_args(h, w, fast, pos, slots, pos_fixed);
        a.qkv = qkv; a.y = yb;
        a.splits = splits;
        a.fused_append = 1;
        if (tc) {
            // Linears with few weight tiles (wo, w2: N = dim) leave fp32 partial tiles behind; the RMSNorm pass that follows
            // them anyway finishes them (sum in CTA order + residual) row-parallel.  wqkv / gate-up finish in-kernel.
            const bool pdl = h->use_pdl != 0;
            const bool last = l + 1 == L.size();
            if (l == 0)
                if (int r = rows_finish_norm(h, s, M, D, 0, nullptr, cur, nullptr, w.attn_norm, h->pxn)) return r;
            // wqkv: its split tiles are finished by the kernels that read q / k / v (K/V append, attention), not by a
            // finish launch, whenever those kernels can (tile-aligned split; the CUDA-core slow attention cannot)
            const int qkv_tiles = (QKV + RG_BM - 1) / RG_BM;
            const bool qkv_part = (fast || (ak.tc && h->attn_tc)) && h->qkv_fold && qkv_tiles <= RG_CTAS_PER_SM * h->n_sm &&
                                  QKV % RG_BM == 0 && !(h->rows_narrow && rgn_pick_nw(QKV, h->n_sm));
            a.qkv_part = nullptr;
            if (int r = rows_linear_tc(h, s, w.wqkv, nullptr, h->pxn, M, QKV, D, RG_EPI_STORE, nullptr, qkv, nullptr, qkv_part)) return r;
            if (qkv_part) {
                a.qkv_part = h->rg_partial;
                a.qkv_s = rg_grid(qkv_tiles, D / RG_BK, h->n_sm) / qkv_tiles;
                a.qkv_bn = rg_bn_for(M);
            }
            if (fast) CK(launch(ak.fast, dim3(KV, M), dim3(ak.threads_fast), 0, s, pdl, a));
            else if (int r = rows_attention(h, s, ak, a, splits, KV, M, pdl, force_splits)) return r;
            const bool dfr_o = rows_all_partial_ok(h, D, H * HD), dfr_2 = rows_all_partial_ok(h, D, F);
            if (int r = rows_linear_tc(h, s, w.wo, nullptr, yb, M, D, H * HD, RG_EPI_RESIDUAL, cur, xb, nullptr, dfr_o)) return r;
            if (int r = rows_finish_norm(h, s, M, D, dfr_o ? H * HD : 0, cur, xb, xb, w.ffn_norm, h->pxn)) return r;
            if (int r = rows_linear_tc(h, s, w.w1, w.w3, h->pxn, M, F, D, RG_EPI_SWIGLU, nullptr, ffb, h->pf1)) return r;
            if (int r = rows_linear_tc(h, s, w.w2, nullptr, ffb, M, D, F, RG_EPI_RESIDUAL, xb, xa, nullptr, dfr_2)) return r;
            const bf16* nw = last ? tail_norm_w : L[l + 1].attn_norm;
            bf16* nx = last ? tail_xn : h->pxn;
            if (int r = rows_finish_norm(h, s, M, D, dfr_2 ? F : 0, xb, xa, xa, nw, nx)) return r;
            cur = xa;
            continue;
        }
        if (int r = rows_linear(h, w.wqkv, h->pxn, h->qkv, M, QKV, D)) return r;
        if (fast) CK(launch(ak.fast, dim3(KV, M), dim3(ak.threads_fast), 0, s, false, a));
        else if (int r = rows_attention(h, s, ak, a, splits, KV, M, false)) return r;
        if (int r = rows_linear(h, w.wo, h->y, h->ptmp, M, D, H * HD)) return r;
        add_rows_kernel<<<ew_blocks, 256, 0, s>>>(cur, h->ptmp, xb, (size_t)M * D / 8);
        rmsnorm_rows_kernel<<<M, 256, 0, s>>>(xb, w.ffn_norm, h->pxn, D, d.norm_eps);
        if (int r = rows_linear(h, w.w1, h->pxn, h->pf1, M, F, D)) return r;
        if (int r = rows_linear(h, w.w3, h->pxn, h->pf3, M, F, D)) return r;
        swiglu_rows_kernel<<<ew_blocks, 256, 0, s>>>(h->pf1, h->pf3, h->ff, (size_t)M * F / 8);
        if (int r = rows_linear(h, w.w2, h->ff, h->ptmp, M, D, F)) return r;
        add_rows_kernel<<<ew_blocks, 256, 0, s>>>(xb, h->ptmp, xa, (size_t)M * D / 8);
        cur = xa;
    }
    CK(cudaGetLastError());
    return 0;
}

// logits[M][N] = head . (RMSNorm(x) * norm_w); the normalised rows are kept in xnorm (the slow stack's `hidden`)
int run_head_rows(fsb_dual_ar* h, cudaStream_t s, int M, const bf16* head, int N, const bf16* x, const bf16* norm_w,
                  bf16* xnorm, bf16* logits, int lo = 0, int n_rows = 0) {
    const fsb_dual_ar_desc& d = h->d;
    if (n_rows > 0 && n_rows < N && rows_tc_layers(h, M)) {  // restricted LM head: rows [lo, lo + n_rows) of the weight, same columns of the logits
        const int r = rows_linear_tc(h, s, head + (size_t)lo * d.dim, nullptr, xnorm, M, n_rows, d.dim, RG_EPI_STORE, nullptr, logits + lo,
                                     nullptr, 0, N);
        if (r >= 0) return r;
    }
    // on the tcgen05 path the layer stack's last finish pass has already produced the normalised rows (tail norm)
    if (!rows_tc_layers(h, M)) rmsnorm_rows_kernel<<<M, 256, 0, s>>>(x, norm_w, xnorm, d.dim, d.norm_eps);
    if (rows_tc_active(h, M)) {
        const int r = rows_linear_tc(h, s, head, nullptr, xnorm, M, N, d.dim, RG_EPI_STORE, nullptr, logits);
        if (r >= 0) return r;
    }
    if (int r = ensure_blas(h, s)) return r;
    if (int r = rows_linear(h, head, xnorm, logits, M, N, d.dim)) return r;
    CK(cudaGetLastError());
    return 0;
}

// transformer layers over M rows. x_in: [M][D] input (not modified); result in h->x (slow) / h->xf (fast)
int run_layers(fsb_dual_ar* h, cudaStream_t s, bool fast, int M, const bf16* x_in, const int32_t* pos,
               const int32_t* slots, bool fused_append, int pos_fixed, const bf16* tail_norm_w = nullptr,
               bf16* tail_xn = nullptr) {
    if (fused_append && batch_gemm_active(h, M))
        return run_layers_rows(h, s, fast, M, x_in, pos, slots, pos_fixed, tail_norm_w, tail_xn);
    const fsb_dual_ar_desc& d = h->d;
    const int D = d.dim, H = d.n_head, KV = d.n_kv_head, HD = d.head_dim, F = d.ffn;
    const int QKV = (H + 2 * KV) * HD;
    AttnKernels ak;
    if (!pick_attn(HD, H / KV, &ak)) return fail(h, "unsupported head_dim / GQA ratio");
    const std::vector<LayerW>& L = fast ? h->fast : h->slow;
    bf16* xa = fast ? h->xf : h->x;
    bf16* xb = fast ? h->xf2 : h->x2;
    const bool pdl = h->use_pdl;
    const bf16* cur = x_in;
    const int splits = splits_for(h, M);
    for (size_t l = 0; l < L.size(); ++l) {
        const LayerW& w = L[l];
        if (!w.wqkv || !w.wo || !w.w1 || !w.w2 || !w.w3 || !w.attn_norm || !w.ffn_norm || !w.kc || !w.vc)
            return fail(h, "unbound weights / kv cache in layer " + std::to_string(l));
        CK(run_gemv(h->gc, pdl, s, M, QKV, D, PRO_RMSNORM, EPI_STORE, w.wqkv, nullptr, cur,
                    w.attn_norm, nullptr, h->qkv, nullptr, d.norm_eps, nullptr, 0));
        AttnArgs a = make_attn_args(h, w, fast, pos, slots, pos_fixed);
        a.pf_ptr = w.wo; a.pf_bytes = (size_t)D * H * HD * sizeof(bf16);
        if (fast) {
            CK(launch(ak.fast, dim3(KV, M), dim3(ak.threads_fast), 0, s, pdl, a));
        } else if (h->chain_attn_tc && ak.tc && h->attn_tc && h->kv_max_seq > h->chain_attn_tc) {
            // long caches (a single request past the frame kernel's 2048 tokens): the tensor-core decode attention of the
            // many-row path (K / V append as its own launch, then attn_tc_kernel) instead of the CUDA-core kernel
            a.pf_ptr = nullptr; a.pf_bytes = 0;
            if (int r = rows_attention(h, s, ak, a, splits, KV, M, pdl)) return r;
        } else {
            a.splits = splits;
            a.fused_append = fused_append ? 1 : 0;
            if (!fused_append) {
                AttnArgs ap = a;
                ap.pf_ptr = nullptr; ap.pf_bytes = 0;
                CK(launch(ak.append, dim3(KV, M), dim3(64), 0, s, pdl, ap));
            }
            CK(launch(ak.decode, dim3(splits, KV, M), dim3(ATTN_THREADS), 0, s, pdl, a));
        }
        CK(run_gemv(h->gc, pdl, s, M, D, H * HD, PRO_NONE, EPI_RESIDUAL, w.wo, nullptr, h->y,
                    nullptr, cur, xb, nullptr, d.norm_eps, nullptr, 0));
        CK(run_gemv(h->gc, pdl, s, M, F, D, PRO_RMSNORM, EPI_SWIGLU, w.w1, w.w3, xb, w.ffn_norm,
                    nullptr, h->ff, nullptr, d.norm_eps, nullptr, 0));
        CK(run_gemv(h->gc, pdl, s, M, D, F, PRO_NONE, EPI_RESIDUAL, w.w2, nullptr, h->ff, nullptr,
                    xb, xa, nullptr, d.norm_eps, nullptr, 0));
        cur = xa;
    }
    return 0;
}

int run_embed(fsb_dual_ar* h, cudaStream_t s, int M, const int32_t* tokens) {
    const fsb_dual_ar_desc& d = h->d;
    if (!h->emb || !h->cb_emb) return fail(h, "embeddings not bound");
    (void)d;
    EmbedArgs e = make_embed_args(h, tokens);
    CK(launch(embed_kernel, dim3(M), dim3(256), 0, s, h->use_pdl != 0, e));
    return 0;
}

// slow stack + final norm + LM head for M decode rows
int run_slow_step(fsb_dual_ar* h, cudaStream_t s, int M, const int32_t* tokens, const int32_t* pos,
                  const int32_t* slots, bool restricted_head = false) {
    const fsb_dual_ar_desc& d = h->d;
    if (int r = run_embed(h, s, M, tokens)) return r;
    // embed wrote x2; layer 0 uses it as input and as the wo residual, writes xb = x2?  -> use a distinct buffer
    // (run_layers writes xb = h->x2 in layer 0 while `cur` == h->x2 is still the residual source of the same
    //  kernel: same element read-then-written by the same thread only, which is safe for EPI_RESIDUAL.)
    const bf16* head = d.tie_word_embeddings ? h->emb : h->out_w;
    if (!head || !h->norm) return fail(h, "LM head / final norm not bound");
    if (int r = run_layers(h, s, false, M, h->x2, pos, slots, true, -1, h->norm, h->hidden)) return r;
    if (batch_gemm_active(h, M)) {
        int lo = 0, n = d.vocab;
        head_rows(d, restricted_head, &lo, &n);
        return run_head_rows(h, s, M, head, d.vocab, h->x, h->norm, h->hidden, h->slow_logits, lo, n);
    }
    int lo = 0, n = d.vocab;
    head_rows(d, restricted_head && M == 1, &lo, &n);
    CK(run_gemv(h->gc, h->use_pdl != 0, s, M, n, d.dim, PRO_RMSNORM, EPI_STORE, head + (size_t)lo * d.dim,
                nullptr, h->x, h->norm, nullptr, h->slow_logits + lo, h->hidden, d.norm_eps, nullptr, 0));
    return 0;
}

int run_fast_step(fsb_dual_ar* h, cudaStream_t s, int M, const bf16* x_in, int cb_idx, bool want_logits) {
    const fsb_dual_ar_desc& d = h->d;
    if (cb_idx < 0 || cb_idx >= d.num_codebooks) return fail(h, "codebook_idx out of range");
    if (want_logits && (!h->fast_out || !h->fast_norm)) return fail(h, "fast head not bound");
    if (int r = run_layers(h, s, true, M, x_in, nullptr, nullptr, true, cb_idx, want_logits ? h->fast_norm : nullptr, h->pxb)) return r;
    if (want_logits) {
        bf16* lg = h->fast_logits + (size_t)cb_idx * h->R * d.codebook_size;
        if (batch_gemm_active(h, M)) return run_head_rows(h, s, M, h->fast_out, d.codebook_size, h->xf, h->fast_norm, h->pxb, lg);
        CK(run_gemv(h->gc, h->use_pdl != 0, s, M, d.codebook_size, d.dim, PRO_RMSNORM, EPI_STORE,
                    h->fast_out, nullptr, h->xf, h->fast_norm, nullptr, lg, nullptr, d.norm_eps, nullptr, 0));
    }
    return 0;
}

int enqueue_frame(fsb_dual_ar* h, cudaStream_t s, const fsb_frame_args& fa) {
    const fsb_dual_ar_desc& d = h->d;
    const int B = fa.batch, C = d.num_codebooks, V = d.vocab, cb = d.codebook_size;
    const bool pdl = h->use_pdl != 0;
    if (int r = run_slow_step(h, s, B, fa.cur_tokens, fa.pos, fa.slots, head_restricted(h, fa))) return r;

    SampleArgs sa = make_slow_sample(h, fa);
    CK(launch(sample_kernel, dim3(B), dim3(SAMPLE_THREADS), 0, s, pdl, sa));

    // fast AR: pass 0 on the slow hidden state (output discarded, inference.py:148-149), then C-1 sampled passes
    const bf16* fin = d.norm_fastlayer_input ? h->hidden : h->x;
    if (int r = run_fast_step(h, s, B, fin, 0, false)) return r;
    for (int i = 1; i < C; ++i) {
        if (int r = run_fast_step(h, s, B, h->xfast, i, true)) return r;
        SampleArgs fs = make_fast_sample(h, fa, i);
        CK(launch(sample_kernel, dim3(B), dim3(SAMPLE_THREADS), 0, s, pdl, fs));
    }
    FrameEndArgs fe = make_frame_end(h, fa);
    CK(launch(frame_end_kernel, dim3(1), dim3(64), 0, s, pdl, fe));
    return 0;
}

// ------------------------------------------------------------------------------------------- persistent frame kernel
typedef void (*mega_fn)(MegaPlan);
// `tiled`: the attention walks a split's keys in tiles (caches longer than ASP_MAXS x asp_max_keys tokens)
mega_fn pick_mega(int hd, int gq, bool tiled = false) {
    if (hd == 128 && gq == 4) return tiled ? frame_mega_kernel<128, 4, true> : frame_mega_kernel<128, 4, false>;
    if (hd == 64 && gq == 4) return tiled ? frame_mega_kernel<64, 4, true> : frame_mega_kernel<64, 4, false>;
    return nullptr;
}
// attention splits per kv head inside the frame kernel: one CTA per item while they last, more splits when a split's key range
// would not fit one tile (mega_long = 2 keeps the requested count: the tiled kernel on short caches, for tests)
int mega_splits(const fsb_dual_ar* h) {
    const int msplit = std::min(h->max_splits, ASP_MAXS), mk = asp_max_keys(h->d.head_dim);
    int splits = h->attn_splits > 0 ? h->attn_splits : std::max(1, h->n_sm / h->d.n_kv_head);
    splits = std::max(1, std::min(splits, msplit));
    if (h->mega_long != 2)
        while (splits < msplit && (h->kv_max_seq + splits - 1) / splits > mk) ++splits;
    return splits;
}
bool mega_tiled(const fsb_dual_ar* h) {
    const int splits = mega_splits(h);
    return (h->kv_max_seq + splits - 1) / splits > asp_max_keys(h->d.head_dim);
}

// stage geometry of one GEMV inside the frame kernel: rows per stage (log2) or -1 if the shape does not fit
int mega_rshift(int K) {
    if (K % 8 || (size_t)K * 2 > MG_X_BYTES || (size_t)K * 2 > MG_STAGE_BYTES) return -1;
    int rs = 3;
    while (rs > 0 && ((size_t)K * 2 << rs) > MG_STAGE_BYTES) --rs;
    const int segw = rs >= 2 ? 1 : (4 >> rs);  // warps of a group sharing one row
    const int cpw = (K / 8 + segw - 1) / segw;
    return cpw <= 32 * MG_CPL ? rs : -1;
}

bool mega_shape_ok(const fsb_dual_ar_desc& d) {
    return pick_mega(d.head_dim, d.n_head / d.n_kv_head) != nullptr && mega_rshift(d.dim) >= 0 &&
           mega_rshift(d.n_head * d.head_dim) >= 0 && mega_rshift(d.ffn) >= 0;
}

struct MegaTagged {  // flag-in-data activation buffers carved from tg_base
    uint32_t *x, *x2, *qkv, *y, *ff, *hidden, *xfast, *xf, *xf2;
    explicit MegaTagged(const fsb_dual_ar* h) {
        const fsb_dual_ar_desc& d = h->d;
        uint32_t* p = h->tg_base;
        auto take = [&](size_t n) { uint32_t* r = p; p += n; return r; };
        x = take(d.dim); x2 = take(d.dim); qkv = take((size_t)(d.n_head + 2 * d.n_kv_head) * d.head_dim);
        y = take((size_t)d.n_head * d.head_dim); ff = take(d.ffn); hidden = take(d.dim); xfast = take(d.dim);
        xf = take(d.dim); xf2 = take(d.dim);
    }
};

struct MegaBuilder {
    int n_cta = 148;
    std::vector<MegaOp> ops;
    std::vector<MegaGemv> gemv;
    std::vector<AttnArgs> attn;
    std::vector<SampleArgs> samp;
    std::vector<std::pair<const uint32_t*, int>> writer;  // last op that wrote each tagged buffer
    std::vector<uint4> hot;                               // {ptr lo, ptr hi, bytes, 0}
    void add_hot(const void* p, size_t bytes) {
        if (!p || bytes < 16) return;
        for (const uint4& q : hot)
            if ((((unsigned long long)q.y << 32) | q.x) == (unsigned long long)(uintptr_t)p) return;
        for (size_t off = 0; off < bytes; off += 32768) {  // one bulk prefetch per <= 32 KB piece
            const unsigned long long v = (unsigned long long)(uintptr_t)p + off;
            const size_t n = std::min<size_t>(32768, bytes - off) & ~(size_t)15;
            if (n) hot.push_back(make_uint4((unsigned)(v & 0xffffffffu), (unsigned)(v >> 32), (unsigned)n, 0u));
        }
    }
    int writer_of(const uint32_t* p) const {
        for (auto it = writer.rbegin(); it != writer.rend(); ++it)
            if (it->first == p) return it->second;
        return -1;
    }
    void wrote(const uint32_t* p) { writer.emplace_back(p, (int)ops.size() - 1); }
    void op(int type, int idx, int sync) { ops.push_back(MegaOp{type, idx, sync, 0}); }
    // out_t != nullptr: tagged output vector; else plain bf16 logits at `out` followed by a grid barrier
    void add_gemv(int N, int K, int pro, int epi, const bf16* W, const bf16* W2, const uint32_t* x_t, const bf16* norm_w,
                  const uint32_t* res_t, uint32_t* out_t, bf16* out, bf16* xnorm_out, uint32_t* xnorm_out_t, float eps) {
        MegaGemv g{};
        g.W = W; g.W2 = W2; g.x_t = x_t; g.norm_w = norm_w; g.res_t = res_t; g.out_t = out_t; g.out = out;
        g.xnorm_out = xnorm_out; g.xnorm_out_t = xnorm_out_t;
        g.N = N; g.K = K; g.pro = pro; g.epi = epi; g.eps = eps; g.rshift = mega_rshift(K);
        g.x_op = writer_of(x_t);
        g.rows_q = N / n_cta; g.rows_rem = N % n_cta;
        if (pro == PRO_RMSNORM) add_hot(norm_w, (size_t)K * 2);
        gemv.push_back(g);
        op(MOP_GEMV, (int)gemv.size() - 1, out_t ? MSYNC_CTA : MSYNC_GRID);
        if (out_t) wrote(out_t);
        if (xnorm_out_t) wrote(xnorm_out_t);
    }
    void add_layers(const fsb_dual_ar* h, const MegaTagged& t, bool fast, const uint32_t* x_in, const int32_t* pos,
                    const int32_t* slots, int pos_fixed, int splits) {
        const fsb_dual_ar_desc& d = h->d;
        const int D = d.dim, H = d.n_head, KV = d.n_kv_head, HD = d.head_dim, F = d.ffn, QKV = (H + 2 * KV) * HD;
        const std::vector<LayerW>& L = fast ? h->fast : h->slow;
        uint32_t* xa = fast ? t.xf : t.x;
        uint32_t* xb = fast ? t.xf2 : t.x2;
        const uint32_t* cur = x_in;
        for (const LayerW& w : L) {
            add_gemv(QKV, D, PRO_RMSNORM, EPI_STORE, w.wqkv, nullptr, cur, w.attn_norm, nullptr, t.qkv, nullptr, nullptr,
                     nullptr, d.norm_eps);
            AttnArgs a = make_attn_args(h, w, fast, pos, slots, pos_fixed);
            a.splits = fast ? 1 : splits;
            a.qkv_t = t.qkv; a.y_t = t.y; a.ll_acc = h->ll_acc; a.ll_ml = h->ll_ml;
            a.pf_ptr = nullptr; a.pf_bytes = (size_t)h->attn_trace_point;
            if (!fast && (&w - &L[0]) + 1 < (long)L.size()) { a.next_k = (&w)[1].kc; a.next_v = (&w)[1].vc; }
            if (d.qk_norm) { add_hot(w.q_norm, (size_t)HD * 2); add_hot(w.k_norm, (size_t)HD * 2); }
            if (fast) add_hot(h->fast_rope + (size_t)pos_fixed * HD, (size_t)HD * 2);
            a.tag_in_op = writer_of(t.qkv); a.tag_out_op = (int)ops.size();
            attn.push_back(a);
            op(fast ? MOP_FATTN : MOP_ATTN, (int)attn.size() - 1, MSYNC_CTA);
            wrote(t.y);
            add_gemv(D, H * HD, PRO_NONE, EPI_RESIDUAL, w.wo, nullptr, t.y, nullptr, cur, xb, nullptr, nullptr, nullptr,
                     d.norm_eps);
            add_gemv(F, D, PRO_RMSNORM, EPI_SWIGLU, w.w1, w.w3, xb, w.ffn_norm, nullptr, t.ff, nullptr, nullptr, nullptr,
                     d.norm_eps);
            add_gemv(D, F, PRO_NONE, EPI_RESIDUAL, w.w2, nullptr, t.ff, nullptr, xb, xa, nullptr, nullptr, nullptr,
                     d.norm_eps);
            cur = xa;
        }
    }
};

int check_bound(fsb_dual_ar* h) {
    const fsb_dual_ar_desc& d = h->d;
    for (int f = 0; f < 2; ++f) {
        const std::vector<LayerW>& L = f ? h->fast : h->slow;
        for (size_t l = 0; l < L.size(); ++l) {
            const LayerW& w = L[l];
            if (!w.wqkv || !w.wo || !w.w1 || !w.w2 || !w.w3 || !w.attn_norm || !w.ffn_norm || !w.kc || !w.vc)
                return fail(h, std::string("unbound weights / kv cache in ") + (f ? "fast " : "") + "layer " + std::to_string(l));
        }
    }
    if (!h->emb || !h->cb_emb) return fail(h, "embeddings not bound");
    if (!(d.tie_word_embeddings ? h->emb : h->out_w) || !h->norm) return fail(h, "LM head / final norm not bound");
    if (!h->fast_out || !h->fast_norm) return fail(h, "fast head not bound");
    return 0;
}

// op list of one batch-1 frame, in the order enqueue_frame launches the same work; uploaded on `s` (stream-ordered
// after earlier launches that still read the previous plan)
int build_mega_plan(fsb_dual_ar* h, cudaStream_t s, const fsb_frame_args& fa) {
    const fsb_dual_ar_desc& d = h->d;
    if (int r = check_bound(h)) return r;
    const int C = d.num_codebooks;
    MegaBuilder mb;
    mb.n_cta = h->n_sm;
    const MegaTagged t(h);
    mb.op(MOP_EMBED, 0, MSYNC_CTA);
    mb.wrote(t.x2);
    // short caches: a split's whole key range is one tile; longer ones take the tiled instantiation (mega_tiled)
    const int splits = mega_splits(h);
    mb.add_layers(h, t, false, t.x2, fa.pos, fa.slots, -1, splits);
    const bf16* head = d.tie_word_embeddings ? h->emb : h->out_w;
    int head_lo = 0, head_n = d.vocab;
    head_rows(d, head_restricted(h, fa), &head_lo, &head_n);
    mb.add_gemv(head_n, d.dim, PRO_RMSNORM, EPI_STORE, head + (size_t)head_lo * d.dim, nullptr, t.x, h->norm, nullptr, nullptr,
                h->slow_logits + head_lo, h->hidden, t.hidden, d.norm_eps);
    {
        SampleArgs sa = make_slow_sample(h, fa);
        sa.xfast_t = t.xfast;
        mb.samp.push_back(sa);
        mb.op(MOP_SAMPLE, 0, MSYNC_CTA);
        mb.wrote(t.xfast);
    }
    const uint32_t* fin = d.norm_fastlayer_input ? t.hidden : t.x;
    mb.add_layers(h, t, true, fin, nullptr, nullptr, 0, 1);
    for (int i = 1; i < C; ++i) {
        mb.add_layers(h, t, true, t.xfast, nullptr, nullptr, i, 1);
        mb.add_gemv(d.codebook_size, d.dim, PRO_RMSNORM, EPI_STORE, h->fast_out, nullptr, t.xf, h->fast_norm, nullptr,
                    nullptr, h->fast_logits + (size_t)i * h->R * d.codebook_size, nullptr, nullptr, d.norm_eps);
        SampleArgs fs = make_fast_sample(h, fa, i);
        fs.xfast_t = (i + 1 < C) ? t.xfast : nullptr;
        mb.samp.push_back(fs);
        mb.op(MOP_SAMPLE, (int)mb.samp.size() - 1, MSYNC_CTA);
        if (i + 1 < C) mb.wrote(t.xfast);
    }
    mb.op(MOP_END, 0, MSYNC_NONE);
    if (fa.logit_bias) mb.add_hot(fa.logit_bias, (size_t)d.vocab * 2);
    if ((int)mb.ops.size() >= 512) return fail(h, "frame has too many ops for the 9-bit op field of the activation tags");
    EmbedArgs emb = make_embed_args(h, fa.cur_tokens);
    emb.x_t = t.x2;
    const FrameEndArgs fend = make_frame_end(h, fa);

    // pack: [ops][gemv][attn][samp][embed][fend], each 16-byte aligned
    size_t off = 0;
    auto place = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 16); return o; };
    const size_t o_ops = place(mb.ops.size() * sizeof(MegaOp)), o_g = place(mb.gemv.size() * sizeof(MegaGemv)),
                 o_a = place(mb.attn.size() * sizeof(AttnArgs)), o_s = place(mb.samp.size() * sizeof(SampleArgs)),
                 o_e = place(sizeof(EmbedArgs)), o_f = place(sizeof(FrameEndArgs)),
                 o_h = place(mb.hot.size() * sizeof(uint4));
    if (off > MEGA_BUF_BYTES) return fail(h, "frame op list exceeds its device buffer");
    h->mega_host.assign(off, 0);
    char* hp = h->mega_host.data();
    memcpy(hp + o_ops, mb.ops.data(), mb.ops.size() * sizeof(MegaOp));
    memcpy(hp + o_g, mb.gemv.data(), mb.gemv.size() * sizeof(MegaGemv));
    memcpy(hp + o_a, mb.attn.data(), mb.attn.size() * sizeof(AttnArgs));
    memcpy(hp + o_s, mb.samp.data(), mb.samp.size() * sizeof(SampleArgs));
    memcpy(hp + o_e, &emb, sizeof(EmbedArgs));
    memcpy(hp + o_f, &fend, sizeof(FrameEndArgs));
    memcpy(hp + o_h, mb.hot.data(), mb.hot.size() * sizeof(uint4));
    CK(cudaMemcpyAsync(h->mega_buf, hp, off, cudaMemcpyHostToDevice, s));
    MegaPlan& p = h->mega_plan;
    p.ops = (const MegaOp*)(h->mega_buf + o_ops); p.n_ops = (int)mb.ops.size(); p.B = 1;
    p.gemv = (const MegaGemv*)(h->mega_buf + o_g); p.n_gemv = (int)mb.gemv.size(); p.attn = (const AttnArgs*)(h->mega_buf + o_a);
    p.samp = (const SampleArgs*)(h->mega_buf + o_s); p.embed = (const EmbedArgs*)(h->mega_buf + o_e);
    p.fend = (const FrameEndArgs*)(h->mega_buf + o_f); p.sync = h->mega_sync;
    p.prof = h->mega_prof_on ? h->mega_prof : nullptr;
    p.hot = (const uint4*)(h->mega_buf + o_h); p.n_hot = (int)mb.hot.size(); p.n_attn_slow = d.n_layer;
    p.l2_ahead = h->mega_l2_ahead;
    p.trace = (h->mega_trace_on && (size_t)h->n_sm * mb.ops.size() * 4 <= MEGA_TRACE_WORDS) ? h->mega_trace : nullptr;
    p.detail_op = h->mega_detail_op;
    p.skip_dot = h->mega_skip_dot;
    p.detail_gemv = (h->mega_detail_op >= 0 && h->mega_detail_op < (int)mb.ops.size() && mb.ops[h->mega_detail_op].type == MOP_GEMV) ? mb.ops[h->mega_detail_op].idx : -1;
    p.detail = (p.trace && h->mega_detail_op >= 0 && (size_t)h->n_sm * (mb.ops.size() * 4 + 64) <= MEGA_TRACE_WORDS)
                   ? h->mega_trace + (size_t)h->n_sm * mb.ops.size() * 4 : nullptr;  // behind the op stamps
    return 0;
}

int launch_mega(fsb_dual_ar* h, cudaStream_t s) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(h->n_sm); cfg.blockDim = dim3(MG_THREADS); cfg.dynamicSmemBytes = MG_SMEM; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeCooperative;  // all CTAs co-resident: the in-kernel grid barrier depends on it
    at[0].val.cooperative = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    CK(cudaLaunchKernelEx(&cfg, pick_mega(h->d.head_dim, h->d.n_head / h->d.n_kv_head, mega_tiled(h)), h->mega_plan));
    return 0;
}

// batch 1 runs the persistent frame kernel while a split's keys fit one tile (caches up to ASP_MAXS x asp_max_keys tokens).
// Longer caches (config 5: 8k-token prompts) take the kernel chain, whose attention kernel streams long key ranges better than
// the frame kernel's tiled instantiation does today (option mega_long = 1 selects the latter)
bool mega_active(const fsb_dual_ar* h, int batch) {
    return h->use_mega && h->mega_supported && batch == 1 && (h->mega_long || !mega_tiled(h));
}

__global__ void fill_prefill_kernel(int32_t* pos, int32_t* slot, int n, int pos0, int slot_v) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { pos[i] = pos0 + i; slot[i] = slot_v; }
}

}  // namespace

// =============================================================================================== C ABI
extern "C" {

int fsb_abi_version(void) { return FSB_ABI_VERSION; }

static void carve(const fsb_dual_ar_desc& d, int max_seq_len, int max_splits, char* base, fsb_dual_ar* h, size_t* total) {
    size_t off = 0;
    const size_t R = (size_t)std::max(d.max_rows, 8);
    const int QKV = (d.n_head + 2 * d.n_kv_head) * d.head_dim;
    auto take = [&](size_t bytes) {
        char* p = base ? base + off : nullptr;
        off = align_up(off + bytes, 256);
        return p;
    };
#define T(field, type, count) { char* p_ = take((size_t)(count) * sizeof(type)); if (h) h->field = (type*)p_; }
    T(x, bf16, R * d.dim) T(x2, bf16, R * d.dim) T(qkv, bf16, R * QKV) T(y, bf16, R * d.n_head * d.head_dim)
    T(ff, bf16, R * d.ffn) T(hidden, bf16, R * d.dim) T(xfast, bf16, R * d.dim) T(xf, bf16, R * d.dim)
    T(xf2, bf16, R * d.dim) T(slow_logits, bf16, R * d.vocab)
    T(fast_logits, bf16, (size_t)d.num_codebooks * R * d.codebook_size)
    const int gq = d.n_head / std::max(1, d.n_kv_head);
    T(ws_acc, float, R * d.n_kv_head * max_splits * gq * d.head_dim)
    T(ws_ml, float, R * d.n_kv_head * max_splits * gq * 2)
    T(counters, int32_t, R * d.n_kv_head)
    {
        const size_t PR = PREFILL_ROWS;
        T(pxa, bf16, PR * d.dim) T(pxb, bf16, PR * d.dim) T(pxn, bf16, PR * d.dim) T(pqkv, bf16, PR * QKV)
        T(py, bf16, PR * d.n_head * d.head_dim) T(ptmp, bf16, PR * d.dim) T(pf1, bf16, PR * d.ffn) T(pf3, bf16, PR * d.ffn)
        T(pq, bf16, PR * d.n_head * d.head_dim) T(pvt, bf16, (size_t)d.n_kv_head * d.head_dim * ((size_t)max_seq_len + 8))
        T(cublas_ws, char, CUBLAS_WS_BYTES)
        T(pks, float, (size_t)PREFILL_KS_ROWS * d.dim)   // split-K partial sums of the prompt pass's wo / w2 (short chunks)
    }
    {
        T(rg_partial, float, rg_partial_bytes(RG_GRID_MAX, 128) / sizeof(float))
    }
    T(mega_buf, char, MEGA_BUF_BYTES) T(mega_sync, unsigned, 64) T(mega_prof, unsigned long long, 16)
    // flag-in-data activations of the persistent frame kernel (batch 1): one u32 per element
    const size_t tg = (size_t)2 * d.dim + QKV + (size_t)d.n_head * d.head_dim + d.ffn + (size_t)4 * d.dim;
    T(tg_base, uint32_t, tg)
    T(ll_acc, uint2, (size_t)d.n_kv_head * max_splits * gq * d.head_dim) T(ll_ml, uint2, (size_t)d.n_kv_head * max_splits * gq * 2)
    T(mega_trace, unsigned long long, MEGA_TRACE_WORDS)
    T(pf_pos, int32_t, (size_t)max_seq_len + 8) T(pf_slot, int32_t, (size_t)max_seq_len + 8)
#undef T
    *total = off;
}

size_t fsb_dual_ar_workspace_bytes(const fsb_dual_ar_desc* desc, int32_t max_seq_len) {
    size_t total = 0;
    carve(*desc, max_seq_len, 32, nullptr, nullptr, &total);
    return total;
}

int fsb_dual_ar_create(const fsb_dual_ar_desc* desc, int32_t max_seq_len, void* workspace, size_t workspace_bytes,
                       fsb_dual_ar** out) {
    if (!desc || !out || !workspace) return 1;
    fsb_dual_ar* h = new fsb_dual_ar();
    h->d = *desc;
    h->max_seq_len = max_seq_len;
    h->R = (size_t)std::max(desc->max_rows, 8);
    *out = h;
    const fsb_dual_ar_desc& d = h->d;
    if (d.dim % 8 || d.ffn % 8 || d.head_dim % 32 || d.n_head % d.n_kv_head || d.num_codebooks > FAST_MAX_CTX ||
        d.num_codebooks < 1 || d.max_rows < 1)
        return fail(h, "unsupported model shape (dim/ffn % 8, head_dim % 32, num_codebooks <= 16 required)");
    AttnKernels ak;
    if (!pick_attn(d.head_dim, d.n_head / d.n_kv_head, &ak)) return fail(h, "unsupported head_dim / GQA ratio");
    size_t need = 0;
    carve(d, max_seq_len, h->max_splits, (char*)workspace, h, &need);
    if (need > workspace_bytes) return fail(h, "workspace too small");
    int dev = 0;
    CK(cudaGetDevice(&dev));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, dev));
    if (prop.major < 10) return fail(h, "fish_speech_b200 requires an sm_100a device (B200)");
    h->n_sm = prop.multiProcessorCount;
    h->gc.n_sm = h->n_sm;
    CK(set_gemv_attrs());
    CK(set_gemv_tma_attrs());
    CK(cudaMemset(h->counters, 0, (size_t)std::max(d.max_rows, 8) * d.n_kv_head * sizeof(int32_t)));
    CK(cudaMemset(h->mega_sync, 0, 64 * sizeof(unsigned)));
    {
        const unsigned one = 1;  // launch sequence starts at 1: tag 0 (the cleared buffers) is never expected
        CK(cudaMemcpy(h->mega_sync + 2, &one, sizeof(one), cudaMemcpyHostToDevice));
        h->tg_words = (size_t)2 * d.dim + (size_t)(d.n_head + 2 * d.n_kv_head) * d.head_dim +
                      (size_t)d.n_head * d.head_dim + d.ffn + (size_t)4 * d.dim;
        CK(cudaMemset(h->tg_base, 0, h->tg_words * sizeof(uint32_t)));
        const size_t gq_ = d.n_head / d.n_kv_head;
        CK(cudaMemset(h->ll_acc, 0, (size_t)d.n_kv_head * h->max_splits * gq_ * d.head_dim * sizeof(uint2)));
        CK(cudaMemset(h->ll_ml, 0, (size_t)d.n_kv_head * h->max_splits * gq_ * 2 * sizeof(uint2)));
    }
    CK(cudaMemset(h->mega_prof, 0, 16 * sizeof(unsigned long long)));
    if (h->rg_partial) h->rg_ok = h->n_sm <= RG_SM_MAX && rg_get_encode() != nullptr && rg_set_attrs() == cudaSuccess;
    if (mega_shape_ok(d)) {
        int coop = 0;
        CK(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
        h->mega_supported = coop != 0;
        for (int tiled = 0; tiled < 2; ++tiled) {
            mega_fn mf = pick_mega(d.head_dim, d.n_head / d.n_kv_head, tiled != 0);
            int occ = 0;
            CK(cudaFuncSetAttribute((const void*)mf, cudaFuncAttributeMaxDynamicSharedMemorySize, MG_SMEM));
            CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, (const void*)mf, MG_THREADS, MG_SMEM));
            h->mega_supported = h->mega_supported && occ >= 1;
        }
    }
    h->slow.resize(d.n_layer);
    h->fast.resize(d.n_fast_layer);
    // the one environment hook: FSB_OPTIONS="name=value,name=value" is applied through fsb_dual_ar_set_option
    if (const char* e = getenv("FSB_OPTIONS")) {
        std::string all(e);
        size_t i = 0;
        while (i < all.size()) {
            size_t j = all.find(',', i);
            if (j == std::string::npos) j = all.size();
            const std::string kv = all.substr(i, j - i);
            const size_t eq = kv.find('=');
            if (eq != std::string::npos && eq > 0)
                if (fsb_dual_ar_set_option(h, kv.substr(0, eq).c_str(), atoll(kv.c_str() + eq + 1))) return 1;
            i = j + 1;
        }
    }
    return 0;
}

int fsb_dual_ar_destroy(fsb_dual_ar* h) {
    if (!h) return 0;
    if (h->exec) cudaGraphExecDestroy(h->exec);
    if (h->cap_stream) cudaStreamDestroy(h->cap_stream);
    if (h->blas) cublasDestroy(h->blas);
    delete h;
    return 0;
}

const char* fsb_dual_ar_last_error(const fsb_dual_ar* h) { return h ? h->err.c_str() : "null handle"; }

int fsb_dual_ar_set_option(fsb_dual_ar* h, const char* name, int64_t value) {
    if (!h || !name) return 1;
    std::string n(name);
    if (n == "pdl") h->use_pdl = (int)value;
    else if (n == "graph") h->use_graph = (int)value;
    else if (n == "self_prefetch_kb") h->gc.self_pf = (size_t)value * 1024;
    else if (n == "gemv_tma") h->gc.tma = (int)value;
    else if (n == "gemv_stages") h->gc.nst = (int)value;
    else if (n == "gemv_ctas_per_sm") h->gc.ctas_per_sm = (int)value;
    else if (n == "gemv_stage_kb") h->gc.stage_kb = (int)value;
    else if (n == "attn_splits") h->attn_splits = (int)value;
    else if (n == "mega") h->use_mega = (int)value;
    else if (n == "prefill_gemm") h->prefill_gemm = (int)value;
    else if (n == "prefill_fa") h->prefill_fa = (int)value;
    else if (n == "batch_gemm") h->batch_gemm = (int)value;
    else if (n == "attn_tc") h->attn_tc = (int)value;
    else if (n == "masked_head") h->masked_head = (int)value;
    else if (n == "attn_trace") h->attn_trace = (int)value;
    else if (n == "attn_tail") h->attn_tail = (int)value;
    else if (n == "qkv_fold") h->qkv_fold = (int)value;
    else if (n == "rows_narrow") h->rows_narrow = (int)value;
    else if (n == "mega_prof") h->mega_prof_on = (int)value;
    else if (n == "mega_trace") h->mega_trace_on = (int)value;
    else if (n == "mega_l2_ahead") h->mega_l2_ahead = (int)value;
    else if (n == "mega_detail_op") h->mega_detail_op = (int)value;
    else if (n == "mega_skip_dot") h->mega_skip_dot = (int)value;
    else if (n == "mega_long") h->mega_long = (int)value;
    else if (n == "prefill_splitk") h->prefill_splitk = (int)value;
    else if (n == "prefill_rows") h->prefill_rows = (int)value;
    else if (n == "chain_attn_tc") h->chain_attn_tc = (int)value;
    else if (n == "rows_min") h->rows_min = (int)std::max<int64_t>(1, value);
    else if (n == "attn_trace_point") h->attn_trace_point = (int)value;
    else if (n == "spin_ns") {  // poll back-off of tagged-load spins on the current device
        const unsigned ns = (unsigned)value;
        if (cudaMemcpyToSymbol(fsb_spin_ns, &ns, sizeof(ns)) != cudaSuccess) return fail(h, "cudaMemcpyToSymbol(fsb_spin_ns) failed");
    }
    else return fail(h, "unknown option " + n);
    h->cached_valid = false;
    return 0;
}

int fsb_dual_ar_bind_weight(fsb_dual_ar* h, const char* name, const void* dptr, int64_t rows, int64_t cols) {
    if (!h || !name || !dptr) return 1;
    const fsb_dual_ar_desc& d = h->d;
    const bf16* p = (const bf16*)dptr;
    std::string n(name);
    if (((uintptr_t)dptr) & 15) return fail(h, n + ": pointer must be 16-byte aligned");
    auto expect = [&](int64_t r, int64_t c) {
        if (rows != r || cols != c) {
            fail(h, n + ": expected [" + std::to_string(r) + "," + std::to_string(c) + "] got [" +
                        std::to_string(rows) + "," + std::to_string(cols) + "]");
            return false;
        }
        return true;
    };
    const int64_t D = d.dim, HD = d.head_dim, QKV = (int64_t)(d.n_head + 2 * d.n_kv_head) * HD;
    h->cached_valid = false;
    if (n == "embeddings.weight") { if (!expect(d.vocab, D)) return 1; h->emb = p; return 0; }
    if (n == "output.weight") { if (!expect(d.vocab, D)) return 1; h->out_w = p; return 0; }
    if (n == "codebook_embeddings.weight") { if (!expect((int64_t)d.codebook_size * d.num_codebooks, D)) return 1; h->cb_emb = p; return 0; }
    if (n == "norm.weight") { if (!expect(D, 1)) return 1; h->norm = p; return 0; }
    if (n == "fast_embeddings.weight") { if (!expect(d.codebook_size, D)) return 1; h->fast_emb = p; return 0; }
    if (n == "fast_norm.weight") { if (!expect(D, 1)) return 1; h->fast_norm = p; return 0; }
    if (n == "fast_output.weight") { if (!expect(d.codebook_size, D)) return 1; h->fast_out = p; return 0; }
    bool fast = false;
    size_t pos = 0;
    if (n.rfind("fast_layers.", 0) == 0) { fast = true; pos = 12; }
    else if (n.rfind("layers.", 0) == 0) { pos = 7; }
    else return fail(h, "unknown weight name " + n);
    size_t dot = n.find('.', pos);
    if (dot == std::string::npos) return fail(h, "bad weight name " + n);
    int li = atoi(n.substr(pos, dot - pos).c_str());
    std::vector<LayerW>& L = fast ? h->fast : h->slow;
    if (li < 0 || li >= (int)L.size()) return fail(h, "layer index out of range in " + n);
    LayerW& w = L[li];
    std::string k = n.substr(dot + 1);
    if (k == "attention.wqkv.weight") { if (!expect(QKV, D)) return 1; w.wqkv = p; }
    else if (k == "attention.wo.weight") { if (!expect(D, (int64_t)d.n_head * HD)) return 1; w.wo = p; }
    else if (k == "attention.q_norm.weight") { if (!expect(HD, 1)) return 1; w.q_norm = p; }
    else if (k == "attention.k_norm.weight") { if (!expect(HD, 1)) return 1; w.k_norm = p; }
    else if (k == "feed_forward.w1.weight") { if (!expect(d.ffn, D)) return 1; w.w1 = p; }
    else if (k == "feed_forward.w3.weight") { if (!expect(d.ffn, D)) return 1; w.w3 = p; }
    else if (k == "feed_forward.w2.weight") { if (!expect(D, d.ffn)) return 1; w.w2 = p; }
    else if (k == "ffn_norm.weight") { if (!expect(D, 1)) return 1; w.ffn_norm = p; }
    else if (k == "attention_norm.weight") { if (!expect(D, 1)) return 1; w.attn_norm = p; }
    else return fail(h, "unknown weight name " + n);
    return 0;
}

int fsb_dual_ar_bind_rope(fsb_dual_ar* h, int32_t is_fast, const void* table, int32_t n_pos) {
    if (!h || !table) return 1;
    if (is_fast) { h->fast_rope = (const bf16*)table; h->fast_rope_n = n_pos; }
    else { h->rope = (const bf16*)table; h->rope_n = n_pos; }
    h->cached_valid = false;
    return 0;
}

int fsb_dual_ar_bind_kv(fsb_dual_ar* h, int32_t is_fast, int32_t layer, void* k, void* v, int32_t max_batch,
                        int32_t max_seq) {
    if (!h || !k || !v) return 1;
    std::vector<LayerW>& L = is_fast ? h->fast : h->slow;
    if (layer < 0 || layer >= (int)L.size()) return fail(h, "kv layer out of range");
    if ((((uintptr_t)k) | ((uintptr_t)v)) & 15) return fail(h, "kv pointers must be 16-byte aligned");
    L[layer].kc = (bf16*)k; L[layer].vc = (bf16*)v;
    if (is_fast) {
        if (max_seq < h->d.num_codebooks) return fail(h, "fast kv cache shorter than num_codebooks");
        h->fkv_max_batch = max_batch; h->fkv_max_seq = max_seq;
    } else {
        if (max_seq > h->max_seq_len) return fail(h, "kv max_seq exceeds the handle's max_seq_len");
        h->kv_max_batch = max_batch; h->kv_max_seq = max_seq;
    }
    h->cached_valid = false;
    return 0;
}

static int check_ready(fsb_dual_ar* h, int B) {
    if (!h) return 1;
    if (B < 1 || B > h->d.max_rows) return fail(h, "batch out of range (max_rows)");
    if (!h->rope || !h->fast_rope) return fail(h, "RoPE tables not bound");
    if (B > h->kv_max_batch || B > h->fkv_max_batch) return fail(h, "batch exceeds the bound KV caches");
    return 0;
}

// C[T][N] = X[T][K] . W[N][K]^T, bf16 in / fp32 accumulate / bf16 out (one rounding per Linear)
static int prefill_linear(fsb_dual_ar* h, const bf16* W, const bf16* X, bf16* C, int T, int N, int K) {
    const float one = 1.f, zero = 0.f;
    cublasStatus_t st = cublasGemmEx(h->blas, CUBLAS_OP_T, CUBLAS_OP_N, N, T, K, &one, W, CUDA_R_16BF, K, X, CUDA_R_16BF, K,
                                     &zero, C, CUDA_R_16BF, N, CUBLAS_COMPUTE_32F, CUBLAS_GEMM_DEFAULT);
    if (st != CUBLAS_STATUS_SUCCESS) return fail(h, "cublasGemmEx failed with status " + std::to_string((int)st));
    return 0;
}

// C = epi(X[T][K] . W[N][K]^T) on tcv_kernel (bf16 in, fp32 accumulate): epi NONE = bf16 store (one rounding per Linear),
// residual: C = bf16(res + bf16(lin)) (llama.py:841-842), GATE: C = bf16(bf16(silu(gate)) * bf16(lin)) (llama.py:986)
static const CUtensorMap* tcv_map(fsb_dual_ar* h, const void* ptr, int rows, int cols, int box_rows) {
    auto key = std::make_tuple(ptr, rows, cols, box_rows);
    auto it = h->tcv_maps.find(key);
    if (it != h->tcv_maps.end()) return &it->second;
    CUtensorMap m;
    if (!tcv_make_map(&m, ptr, false, (uint64_t)cols, (uint64_t)rows, 1, (uint32_t)box_rows)) return nullptr;
    return &h->tcv_maps.emplace(key, m).first->second;
}
// out = bf16(bf16(sum of the split-K partial sums) + res): the rounding points of a Linear followed by the residual add (llama.py:838-843)
__global__ void __launch_bounds__(256) splitk_finish_kernel(const float* __restrict__ part, int ks, size_t n, const bf16* __restrict__ res,
                                                            bf16* __restrict__ out) {
    const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= n) return;
    float4 acc = *reinterpret_cast<const float4*>(part + i);
    for (int k = 1; k < ks; ++k) {   // fixed order: deterministic
        const float4 p = *reinterpret_cast<const float4*>(part + (size_t)k * n + i);
        acc.x += p.x; acc.y += p.y; acc.z += p.z; acc.w += p.w;
    }
    const uint2 r = *reinterpret_cast<const uint2*>(res + i);
    const float o0 = rbf(acc.x) + bf16_lo(r.x), o1 = rbf(acc.y) + bf16_hi(r.x), o2 = rbf(acc.z) + bf16_lo(r.y), o3 = rbf(acc.w) + bf16_hi(r.y);
    *reinterpret_cast<uint2*>(out + i) = make_uint2(pack2(o0, o1), pack2(o2, o3));
}

static int prefill_linear_tc(fsb_dual_ar* h, cudaStream_t s, const bf16* W, const bf16* X, bf16* C, int T, int N, int K,
                             const bf16* res = nullptr, const bf16* gate = nullptr) {
    if (K % 8) return fail(h, "prompt-pass Linear needs K % 8 == 0");
    // A Linear with few output columns on a short chunk (wo, w2: 2560 columns) has too few tiles for the chip, and a tile's time
    // is its number of 64-channel steps (~900 cycles each whatever the tile width): the widest tile, its channel range cut into
    // pieces that run side by side, fp32 partial sums added by a finish kernel (511 tokens: w2 94 -> ~35 us).
    if (h->prefill_splitk && res && !gate && N % 4 == 0 && N <= h->d.dim) {
        const int bnw = tcv_pick_bn(N);
        const int tiles = ((T + TCV_BM - 1) / TCV_BM) * ((N + bnw - 1) / bnw), kchunks = (K + TCV_BK - 1) / TCV_BK;
        int ks = std::min(std::min(h->n_sm / std::max(1, tiles), kchunks / 8), 8);
        ks = std::min(ks, PREFILL_KS_ROWS / std::max(1, T));
        if (ks >= 2) {
            const CUtensorMap* mA = tcv_map(h, X, T, K, TCV_BM);
            const CUtensorMap* mW = tcv_map(h, W, N, K, bnw);
            if (!mA || !mW) return fail(h, "cuTensorMapEncodeTiled failed");
            TcConvArgs a{};
            a.B = 1; a.T = T; a.N = N; a.Cin = K; a.taps = 1; a.dil = 1;
            a.out32 = h->pks; a.ld32 = N; a.ld16 = N; a.ldres = N; a.bias_mod = N; a.alpha_mod = N;
            a.ksplit = ks;
            CK(tcv_launch(s, h->n_sm, false, bnw, *mA, *mW, a));
            const size_t n = (size_t)T * N;
            const int ks_used = (kchunks + ((kchunks + ks - 1) / ks) - 1) / ((kchunks + ks - 1) / ks);
            splitk_finish_kernel<<<(unsigned)((n / 4 + 255) / 256), 256, 0, s>>>(h->pks, ks_used, n, res, C);
            CK(cudaGetLastError());
            return 0;
        }
    }
    // column-tile width: enough tiles for the 148 SMs when the chunk is short
    int bn = tcv_pick_bn(N);
    const int tiles_t = (T + TCV_BM - 1) / TCV_BM;
    while (bn > 64 && tiles_t * ((N + bn - 1) / bn) < h->n_sm && N % (bn / 2) == 0) bn /= 2;
    const CUtensorMap* mA = tcv_map(h, X, T, K, TCV_BM);
    const CUtensorMap* mW = tcv_map(h, W, N, K, bn);
    if (!mA || !mW) return fail(h, "cuTensorMapEncodeTiled failed");
    TcConvArgs a{};
    a.B = 1; a.T = T; a.N = N; a.Cin = K; a.taps = 1; a.dil = 1;
    a.out16 = C; a.ld16 = N; a.ldres = N; a.bias_mod = N; a.alpha_mod = N;
    if (gate) { a.res16 = gate; a.act = TCV_ACT_GATE; }
    else if (res) { a.res16 = res; a.round_lin = 1; }
    CK(tcv_launch(s, h->n_sm, false, bn, *mA, *mW, a));
    return 0;
}

// attention of T prompt rows at absolute positions pos0 .. pos0 + T - 1 of `slot` on the tcgen05 flash kernel (fa_prefill.cuh):
// q/k prep + KV append, V^T scratch of all keys so far, then one CTA per (head, 128-query block).  qkv in h->pqkv, result in h->py
static const CUtensorMap* fa_map(fsb_dual_ar* h, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2) {
    auto key = std::make_tuple(ptr, d0, d1, d2);
    auto it = h->fa_maps.find(key);
    if (it != h->fa_maps.end()) return &it->second;
    CUtensorMap m;
    if (!tcv_make_map(&m, ptr, false, d0, d1, d2, 128)) return nullptr;
    return &h->fa_maps.emplace(key, m).first->second;
}
static int prefill_attention_fa(fsb_dual_ar* h, cudaStream_t s, const LayerW& w, int T, const int32_t* pos, int pos0, int slot) {
    const fsb_dual_ar_desc& d = h->d;
    const int H = d.n_head, KV = d.n_kv_head, S = h->kv_max_seq;
    {
        FsbOnce once(5);
        if (once.first) CK(cudaFuncSetAttribute((const void*)fa_prefill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FA_SMEM));
    }
    QkPrepArgs qa{};
    qa.qkv = h->pqkv; qa.q_out = h->pq; qa.kcache = w.kc; qa.vcache = w.vc; qa.q_norm = w.q_norm; qa.k_norm = w.k_norm; qa.rope = h->rope;
    qa.pos = pos; qa.T = T; qa.H = H; qa.KV = KV; qa.S = S; qa.slot = slot; qa.qk_norm = d.qk_norm; qa.eps = d.norm_eps;
    qk_prep_kernel<<<dim3(H + KV, T), 64, 0, s>>>(qa);
    const int n_keys = pos0 + T;
    vt_transpose_kernel<<<dim3(((n_keys + FA_BK - 1) / FA_BK) * 2, 2, KV), 256, 0, s>>>(w.vc, h->pvt, KV, S, slot, n_keys);
    const CUtensorMap* mQ = fa_map(h, h->pq, FA_HD, (uint64_t)T, (uint64_t)H);
    const CUtensorMap* mK = fa_map(h, w.kc, FA_HD, (uint64_t)S, (uint64_t)h->kv_max_batch * KV);
    const CUtensorMap* mV = fa_map(h, h->pvt, (uint64_t)S, FA_HD, (uint64_t)KV);
    if (!mQ || !mK || !mV) return fail(h, "cuTensorMapEncodeTiled failed (flash attention maps)");
    FaArgs fa{};
    fa.T = T; fa.pos0 = pos0; fa.H = H; fa.KV = KV; fa.S = S; fa.slot = slot; fa.nqb = (T + FA_BQ - 1) / FA_BQ;
    fa.scale_log2 = (float)(1.4426950408889634 / std::sqrt((double)FA_HD)); fa.y = h->py;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(H * fa.nqb); cfg.blockDim = dim3(FA_THREADS); cfg.dynamicSmemBytes = FA_SMEM; cfg.stream = s;
    CK(cudaLaunchKernelEx(&cfg, fa_prefill_kernel, *mQ, *mK, *mV, fa));
    return 0;
}

// slow stack over a chunk of T <= PREFILL_ROWS prompt rows already embedded in h->pxa; fills the KV cache
static int prefill_chunk(fsb_dual_ar* h, cudaStream_t s, int T, const int32_t* pos, const int32_t* slots, int pos0, int slot) {
    const fsb_dual_ar_desc& d = h->d;
    const int D = d.dim, H = d.n_head, KV = d.n_kv_head, HD = d.head_dim, F = d.ffn, QKV = (H + 2 * KV) * HD;
    AttnKernels ak;
    if (!pick_attn(HD, H / KV, &ak)) return fail(h, "unsupported head_dim / GQA ratio");
    const int ew_blocks = 4 * h->n_sm;
    bf16* cur = h->pxa;
    for (size_t l = 0; l < h->slow.size(); ++l) {
        const LayerW& w = h->slow[l];
        const bool own = h->prefill_gemm == 1;  // this library's tcgen05 kernel with the adds / SwiGLU in its epilogue
        rmsnorm_rows_kernel<<<T, 256, 0, s>>>(cur, w.attn_norm, h->pxn, D, d.norm_eps);
        if (int r = own ? prefill_linear_tc(h, s, w.wqkv, h->pxn, h->pqkv, T, QKV, D) : prefill_linear(h, w.wqkv, h->pxn, h->pqkv, T, QKV, D)) return r;
        if (own && h->prefill_fa && HD == FA_HD) {
            if (int r = prefill_attention_fa(h, s, w, T, pos, pos0, slot)) return r;
        } else {
            AttnArgs a = make_attn_args(h, w, false, pos, slots, -1);
            a.qkv = h->pqkv; a.y = h->py; a.splits = 1; a.fused_append = 0;
            CK(launch(ak.append, dim3(KV, T), dim3(64), 0, s, false, a));
            CK(launch(ak.decode, dim3(1, KV, T), dim3(ATTN_THREADS), 0, s, false, a));
        }
        if (own) {
            if (int r = prefill_linear_tc(h, s, w.wo, h->py, h->pxb, T, D, H * HD, cur)) return r;           // pxb = cur + wo(y)
            rmsnorm_rows_kernel<<<T, 256, 0, s>>>(h->pxb, w.ffn_norm, h->pxn, D, d.norm_eps);
            if (int r = prefill_linear_tc(h, s, w.w1, h->pxn, h->pf1, T, F, D)) return r;
            if (int r = prefill_linear_tc(h, s, w.w3, h->pxn, h->pf1, T, F, D, nullptr, h->pf1)) return r;   // pf1 = silu(pf1) * w3(x), in place
            if (int r = prefill_linear_tc(h, s, w.w2, h->pf1, h->pxa, T, D, F, h->pxb)) return r;            // pxa = pxb + w2(ff)
            cur = h->pxa;
            continue;
        }
        if (int r = prefill_linear(h, w.wo, h->py, h->ptmp, T, D, H * HD)) return r;
        add_rows_kernel<<<ew_blocks, 256, 0, s>>>(cur, h->ptmp, h->pxb, (size_t)T * D / 8);
        rmsnorm_rows_kernel<<<T, 256, 0, s>>>(h->pxb, w.ffn_norm, h->pxn, D, d.norm_eps);
        if (int r = prefill_linear(h, w.w1, h->pxn, h->pf1, T, F, D)) return r;
        if (int r = prefill_linear(h, w.w3, h->pxn, h->pf3, T, F, D)) return r;
        swiglu_rows_kernel<<<ew_blocks, 256, 0, s>>>(h->pf1, h->pf3, h->pf1, (size_t)T * F / 8);
        if (int r = prefill_linear(h, w.w2, h->pf1, h->ptmp, T, D, F)) return r;
        add_rows_kernel<<<ew_blocks, 256, 0, s>>>(h->pxb, h->ptmp, h->pxa, (size_t)T * D / 8);
        cur = h->pxa;
    }
    CK(cudaGetLastError());
    return 0;
}

int fsb_dual_ar_prefill(fsb_dual_ar* h, const int32_t* tokens, int32_t T, int32_t slot, int32_t pos0, void* stream) {
    if (int r = check_ready(h, 1)) return r;
    if (T <= 0) return 0;
    if (pos0 < 0 || pos0 + T > h->kv_max_seq || pos0 + T > h->rope_n) return fail(h, "prefill exceeds max_seq_len");
    if (slot < 0 || slot >= h->kv_max_batch) return fail(h, "slot out of range");
    cudaStream_t s = (cudaStream_t)stream;
    fill_prefill_kernel<<<(T + 255) / 256, 256, 0, s>>>(h->pf_pos, h->pf_slot, T, pos0, slot);
    CK(cudaGetLastError());
    const int C1 = h->d.num_codebooks + 1;
    if (h->prefill_gemm == 1 && (rg_get_encode() == nullptr || h->d.dim % 8 || h->d.ffn % 8)) h->prefill_gemm = 0;  // no tensor-map encoder: decode kernels
    // A short prompt piece (<= 128 columns: the own columns behind a resident prefix, a short prompt) goes through the many-row
    // decode kernels in ONE pass: rows of one KV slot at consecutive positions; the K / V append is its own launch before the
    // tensor-core attention, so a row sees the keys of the rows before it.  A pass streams the weights once over all SMs (~75 us
    // per layer), the 128-row tiles of the prompt-pass GEMM need ~190 us per layer however few rows there are.  The rows live in
    // the prompt-pass scratch (the handle's own activation and split-attention buffers hold max_rows rows only: one attention
    // item per (row, head), no split workspace).
    if (h->prefill_gemm == 1 && h->prefill_rows && T >= 2 && T <= 128) {
        AttnKernels ak;
        if (pick_attn(h->d.head_dim, h->d.n_head / h->d.n_kv_head, &ak) && ak.tc && h->attn_tc && rows_tc_layers(h, T)) {
            if (int r = check_bound(h)) return r;
            EmbedArgs e = make_embed_args(h, tokens);
            e.x = h->pxa;
            CK(launch(embed_kernel, dim3(T), dim3(256), 0, s, false, e));
            const RowsBufs rb{h->ptmp, h->pxb, h->pqkv, h->py, h->pf3};
            return run_layers_rows(h, s, false, T, h->pxa, h->pf_pos, h->pf_slot, -1, nullptr, nullptr, &rb, 1);
        }
    }
    // ... and a handle with many rows (a scheduler's) takes up to 2 x max_rows columns in passes of max_rows rows on its own buffers
    if (h->prefill_gemm == 1 && h->prefill_rows && h->d.max_rows >= 16 && T <= 2 * h->d.max_rows) {
        const int Mr = std::min<int>(T, h->d.max_rows);
        AttnKernels ak;
        if (pick_attn(h->d.head_dim, h->d.n_head / h->d.n_kv_head, &ak) && ak.tc && h->attn_tc && rows_tc_layers(h, Mr)) {
            if (int r = check_bound(h)) return r;
            for (int t0 = 0; t0 < T; t0 += h->d.max_rows) {
                const int M = std::min<int>(h->d.max_rows, T - t0);
                if (int r = run_embed(h, s, M, tokens + (size_t)t0 * C1)) return r;
                if (rows_tc_layers(h, M)) {
                    if (int r = run_layers_rows(h, s, false, M, h->x2, h->pf_pos + t0, h->pf_slot + t0, -1, nullptr, nullptr)) return r;
                } else if (int r = run_layers(h, s, false, M, h->x2, h->pf_pos + t0, h->pf_slot + t0, false, -1)) {  // a last piece of one row
                    return r;
                }
            }
            return 0;
        }
    }
    if (h->prefill_gemm) {
        if (int r = check_bound(h)) return r;
        if (h->prefill_gemm == 1) CK(tcv_set_attrs());
        else if (int r = ensure_blas(h, s)) return r;
        for (int t0 = 0; t0 < T; t0 += PREFILL_ROWS) {
            const int M = std::min<int>(PREFILL_ROWS, T - t0);
            EmbedArgs e = make_embed_args(h, tokens + (size_t)t0 * C1);
            e.x = h->pxa;
            CK(launch(embed_kernel, dim3(M), dim3(256), 0, s, false, e));
            if (int r = prefill_chunk(h, s, M, h->pf_pos + t0, h->pf_slot + t0, pos0 + t0, slot)) return r;
        }
        return 0;
    }
    const int chunk = 8;
    for (int t0 = 0; t0 < T; t0 += chunk) {
        const int M = std::min(chunk, T - t0);
        if (int r = run_embed(h, s, M, tokens + (size_t)t0 * C1)) return r;
        if (int r = run_layers(h, s, false, M, h->x2, h->pf_pos + t0, h->pf_slot + t0, false, -1)) return r;
    }
    return 0;
}

int fsb_dual_ar_slow_step(fsb_dual_ar* h, int32_t B, const int32_t* tokens, const int32_t* pos, const int32_t* slots,
                          void* stream) {
    if (int r = check_ready(h, B)) return r;
    return run_slow_step(h, (cudaStream_t)stream, B, tokens, pos, slots);
}

int fsb_dual_ar_fast_step(fsb_dual_ar* h, int32_t B, const void* x, int32_t codebook_idx, int32_t want_logits,
                          void* stream) {
    if (int r = check_ready(h, B)) return r;
    const bf16* xin = x ? (const bf16*)x : (h->d.norm_fastlayer_input ? h->hidden : h->x);
    return run_fast_step(h, (cudaStream_t)stream, B, xin, codebook_idx, want_logits != 0);
}

int fsb_dual_ar_buffer(fsb_dual_ar* h, const char* name, void** dptr, size_t* bytes) {
    if (!h || !name || !dptr || !bytes) return 1;
    const fsb_dual_ar_desc& d = h->d;
    const size_t R = (size_t)std::max(d.max_rows, 8);
    std::string n(name);
    if (n == "slow_logits") { *dptr = h->slow_logits; *bytes = R * d.vocab * 2; }
    else if (n == "hidden") { *dptr = h->hidden; *bytes = R * d.dim * 2; }
    else if (n == "fast_logits") { *dptr = h->fast_logits; *bytes = (size_t)d.num_codebooks * R * d.codebook_size * 2; }
    else if (n == "x") { *dptr = h->x; *bytes = R * d.dim * 2; }
    else if (n == "xfast") { *dptr = h->xfast; *bytes = R * d.dim * 2; }
    else if (n == "mega_prof") { *dptr = h->mega_prof; *bytes = 16 * sizeof(unsigned long long); }
    else if (n == "mega_trace") { *dptr = h->mega_trace; *bytes = MEGA_TRACE_WORDS * sizeof(unsigned long long); }
    else if (n == "mega_plan") { *dptr = h->mega_buf; *bytes = MEGA_BUF_BYTES; }
    else return fail(h, "unknown buffer " + n);
    return 0;
}

int fsb_dual_ar_decode_frames(fsb_dual_ar* h, const fsb_frame_args* args, int32_t n_frames, void* stream) {
    if (!args) return 1;
    if (int r = check_ready(h, args->batch)) return r;
    if (!args->cur_tokens || !args->pos || !args->frame_counter) return fail(h, "cur_tokens/pos/frame_counter required");
    if (args->top_k < 1 || args->top_k > SAMPLE_MAXK) return fail(h, "top_k must be in [1, 64]");
    if (!h->fast_emb) return fail(h, "fast_embeddings not bound");
    cudaStream_t s = (cudaStream_t)stream;
    const bool mega = mega_active(h, args->batch);
    if (batch_gemm_active(h, args->batch) && !rows_tc_active(h, args->batch))  // the handle must exist before a stream capture begins
        if (int r = ensure_blas(h, s)) return r;
    const bool fresh = !h->cached_valid || memcmp(&h->cached, args, sizeof(fsb_frame_args)) != 0;
    if (mega && fresh) {
        if (int r = build_mega_plan(h, s, *args)) return r;
        if (!h->use_graph) {  // the plan is the cache in this mode
            memcpy(&h->cached, args, sizeof(fsb_frame_args));
            h->cached_valid = true;
        }
    }
    if (!h->use_graph) {
        for (int i = 0; i < n_frames; ++i)
            if (int r = mega ? launch_mega(h, s) : enqueue_frame(h, s, *args)) return r;
        return 0;
    }
    if (fresh || !h->exec) {
        if (h->exec) { cudaGraphExecDestroy(h->exec); h->exec = nullptr; }
        cudaGraph_t g = nullptr;
        if (!h->cap_stream) CK(cudaStreamCreateWithFlags(&h->cap_stream, cudaStreamNonBlocking));
        CK(cudaStreamBeginCapture(h->cap_stream, cudaStreamCaptureModeThreadLocal));
        int r = mega ? launch_mega(h, h->cap_stream) : enqueue_frame(h, h->cap_stream, *args);
        cudaError_t e = cudaStreamEndCapture(h->cap_stream, &g);
        if (r) { if (g) cudaGraphDestroy(g); return r; }
        if (e != cudaSuccess) return fail(h, std::string("cudaStreamEndCapture: ") + cudaGetErrorString(e));
        e = cudaGraphInstantiate(&h->exec, g, 0);
        cudaGraphDestroy(g);
        if (e != cudaSuccess) { h->exec = nullptr; return fail(h, std::string("cudaGraphInstantiate: ") + cudaGetErrorString(e)); }
        memcpy(&h->cached, args, sizeof(fsb_frame_args));
        h->cached_valid = true;
    }
    for (int i = 0; i < n_frames; ++i) CK(cudaGraphLaunch(h->exec, s));
    return 0;
}

int fsb_dual_ar_device_error(fsb_dual_ar* h, int32_t* code, void* stream) {
    if (!h) return 1;
    CK(cudaStreamSynchronize((cudaStream_t)stream));
    unsigned v = 0;
    CK(cudaMemcpyFromSymbol(&v, fsb_dev_error, sizeof(v)));
    if (code) *code = (int32_t)v;
    if (v == 0) return 0;
    const unsigned zero = 0;
    CK(cudaMemcpyToSymbol(fsb_dev_error, &zero, sizeof(zero)));
    static const char* what[] = {"", "activation tag", "shared-memory stage (mbarrier)", "grid barrier", "weight ring slot",
                                 "split-attention partial"};
    return fail(h, std::string("device watchdog: a wait for a ") + (v < 6 ? what[v] : "dependency") +
                       " timed out; the results of that launch are invalid");
}

double fsb_dual_ar_frame_bytes(const fsb_dual_ar* h, int32_t batch, int64_t ctx_sum, int32_t masked_head) {
    if (!h) return 0;
    const fsb_dual_ar_desc& d = h->d;
    const double D = d.dim, HD = d.head_dim, H = d.n_head, KV = d.n_kv_head, F = d.ffn, C = d.num_codebooks;
    const double p_layer = D * (H + 2 * KV) * HD + H * HD * D + 3 * D * F;
    int hl = 0, hn = d.vocab;
    head_rows(d, masked_head != 0 && h->masked_head && (batch == 1 || rows_tc_layers(h, batch)), &hl, &hn);
    const double v_eff = hn;  // rows the fused frame's LM head streams (4136 of 155776 at the S2 ids when restricted)
    double bytes = 2.0 * (d.n_layer * p_layer + v_eff * D) + 2.0 * (C * d.n_fast_layer * p_layer + (C - 1) * d.codebook_size * D);
    bytes += (double)ctx_sum * 2.0 * d.n_layer * 2.0 * KV * HD;
    (void)batch;
    return bytes;
}

int32_t fsb_dual_ar_launches_per_frame(const fsb_dual_ar* h, int32_t batch) {
    if (!h) return 0;
    const fsb_dual_ar_desc& d = h->d;
    if (mega_active(h, batch)) return 1;  // the whole frame is one persistent kernel
    const int C_ = d.num_codebooks;
    if (rows_tc_layers(h, batch)) {
        // many-row step on this library's kernels: per slow layer wqkv GEMM + tile finish, K/V append, attention, wo GEMM,
        // finish+norm, gate/up GEMM + SwiGLU finish, w2 GEMM, finish+norm (10; the fused attention kernel saves the append);
        // per fast layer the same with one attention kernel (9); +1 leading norm per stack, heads = GEMM + tile finish
        AttnKernels ak{};
        const bool tc_attn = pick_attn(d.head_dim, d.n_head / d.n_kv_head, &ak) && ak.tc && h->attn_tc;
        const int fold = h->qkv_fold ? 1 : 0;  // wqkv tile finish folded into the readers of q / k / v
        const int slow = 1 + d.n_layer * ((tc_attn ? 10 : 9) - (tc_attn ? fold : 0)), fast = 1 + d.n_fast_layer * (9 - fold);
        return 1 /*embed*/ + slow + 2 /*LM head*/ + 1 /*sample*/ + C_ * fast + (C_ - 1) * 3 /*fast head + sample*/ + 1 /*frame end*/;
    }
    const int passes = (batch + 7) / 8;  // GEMV passes of <= 8 rows
    const int per_layer = 4 * passes + 1;
    const int C = d.num_codebooks;
    AttnKernels ak2{};
    const bool long_tc = h->chain_attn_tc && pick_attn(d.head_dim, d.n_head / d.n_kv_head, &ak2) && ak2.tc && h->attn_tc &&
                         h->kv_max_seq > h->chain_attn_tc;  // K/V append as its own launch before the tensor-core attention
    return 1 + d.n_layer * (per_layer + (long_tc ? 1 : 0)) + passes /*head*/ + 1 /*slow sample*/ + C * d.n_fast_layer * per_layer +
           (C - 1) * (passes + 1) + 1 /*frame end*/;
}

int fsb_gemv(const void* W, const void* W2, const void* x, const void* norm_w, const void* res, void* out, int32_t M,
             int32_t N, int32_t K, int32_t pro, int32_t epi, float eps, void* stream) {
    fsb_dual_ar* h = nullptr;
    if (!W || !x || !out || M < 1 || K % 8 || pro < 0 || pro > 1 || epi < 0 || epi > 2) return 1;
    if (epi == EPI_SWIGLU && !W2) return 1;
    if (epi == EPI_RESIDUAL && !res) return 1;
    if (pro == PRO_RMSNORM && !norm_w) return 1;
    if (set_gemv_attrs() != cudaSuccess || set_gemv_tma_attrs() != cudaSuccess) return 2;
    int dev = 0, n_sm = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
    GemvCfg gc;
    gc.n_sm = n_sm;
    cudaError_t e = run_gemv(gc, false, (cudaStream_t)stream, M, N, K, pro, epi, (const bf16*)W,
                             (const bf16*)W2, (const bf16*)x, (const bf16*)norm_w, (const bf16*)res, (bf16*)out, nullptr,
                             eps, nullptr, 0);
    (void)h;
    return e == cudaSuccess ? 0 : 3;
}

size_t fsb_rows_gemm_scratch_bytes(void) {  // partial tiles | 8 clock stamps per CTA (written when FSB_RG_TRACE is set) | SwiGLU pre-activations
    return rg_partial_bytes(RG_GRID_MAX, 128) + RG_GRID_MAX * 8 * sizeof(unsigned long long) + (size_t)128 * 2 * 32768 * sizeof(bf16);
}

int fsb_rows_gemm(const void* W, const void* W2, const void* x, const void* res, void* out, int32_t M, int32_t N, int32_t K,
                  int32_t epi, void* scratch, size_t scratch_bytes, void* stream) {
    if (!W || !x || !out || !scratch || M < 1 || M > 128 || N < 1 || K < RG_BK || K % RG_BK) return 1;
    const bool sw = epi == RG_EPI_SWIGLU;
    const int NV = sw ? 2 * N : N;
    if ((NV + RG_BM - 1) / RG_BM > RG_MAX_TILES || scratch_bytes < fsb_rows_gemm_scratch_bytes()) return 1;
    if ((long long)((NV + RG_BM - 1) / RG_BM) * (K / RG_BK) * RG_GRID_MAX >= (1LL << 31)) return 1;
    if (epi < RG_EPI_STORE || epi > RG_EPI_SWIGLU || (sw && (!W2 || N % RG_BM || N > 32768)) || (epi == RG_EPI_RESIDUAL && !res)) return 1;
    int dev = 0, n_sm = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 2;
    if (n_sm > RG_SM_MAX || rg_set_attrs() != cudaSuccess) return 2;
    CUtensorMap mW, mW2, mX;
    if (!rg_make_map(&mW, W, (uint64_t)N, (uint64_t)K, RG_BM)) return 3;
    if (!rg_make_map(&mW2, W2 ? W2 : W, (uint64_t)N, (uint64_t)K, RG_BM)) return 3;
    if (!rg_make_map(&mX, x, (uint64_t)M, (uint64_t)K, (uint32_t)rg_bn_for(M))) return 3;
    float* partial = (float*)scratch;
    static const bool want_trace = getenv("FSB_RG_TRACE") != nullptr;
    unsigned long long* trace = (unsigned long long*)((char*)scratch + rg_partial_bytes(RG_GRID_MAX, 128));
    bf16* pre = (bf16*)(trace + RG_GRID_MAX * 8);
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e = rg_launch(s, false, n_sm, mW, mW2, mX, M, NV, (N + RG_BM - 1) / RG_BM, K, sw ? RG_EPI_STORE : epi, (const bf16*)res,
                              sw ? pre : (bf16*)out, partial, 0, want_trace ? trace : nullptr);
    if (e != cudaSuccess) return 4;
    e = rg_finish_tiles(s, false, n_sm, M, NV, K, epi, partial, sw ? pre : (const bf16*)out, (const bf16*)res, (bf16*)out);
    return e == cudaSuccess ? 0 : 4;
}

// ------------------------------------------------------------------------------------------- no-allocation variant
// fsb_rows_gemm with its stream-K partial tiles in a file-scope device array instead of caller scratch, for callers that
// cannot allocate (code under graph capture, serving loops, harnesses that forbid it).  Limit: M <= 32 rows, which bounds
// the partial tiles at RG_GRID_MAX CTAs x 2 slots x 32 rows x 128 columns x 4 B = 10 MiB; every GPU that loads the library
// holds those 10 MiB until exit.  The array's name is only a host-side stand-in in host code: its device address is looked
// up once per device with cudaGetSymbolAddress (that device current) and that pointer is what the kernels get.
#define FSB_STATIC_ROWS_MAX 32
#define FSB_STATIC_SCRATCH_BYTES ((size_t)RG_GRID_MAX * 2 * FSB_STATIC_ROWS_MAX * RG_BM * sizeof(float))
static __device__ __align__(256) unsigned char fsb_static_scratch[FSB_STATIC_SCRATCH_BYTES];

static int static_msg(char* err, size_t err_len, int rc, const std::string& m) {
    if (err && err_len) snprintf(err, err_len, "%s", m.c_str());
    return rc;
}
// device address of fsb_static_scratch on the CURRENT device, resolved once per device
static cudaError_t static_scratch_ptr(void** out) {
    static std::mutex mu;
    static std::map<int, void*> ptrs;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lock(mu);
    auto it = ptrs.find(dev);
    if (it == ptrs.end()) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, fsb_static_scratch);
        if (e != cudaSuccess) return e;
        if (p == nullptr) return cudaErrorInvalidSymbol;
        it = ptrs.emplace(dev, p).first;
    }
    *out = it->second;
    return cudaSuccess;
}

int fsb_rows_gemm_static(const void* W, const void* x, const void* res, void* out, int32_t M, int32_t N, int32_t K, int32_t epi,
                         void* stream, char* err, size_t err_len) {
    if (err && err_len) err[0] = 0;
    if (!W || !x || !out) return static_msg(err, err_len, 1, "fsb_rows_gemm_static: W, x and out are required");
    if (M < 1 || N < 1 || K < RG_BK || K % RG_BK)
        return static_msg(err, err_len, 1, "fsb_rows_gemm_static: need M >= 1, N >= 1, K a positive multiple of 64");
    if (M > FSB_STATIC_ROWS_MAX)
        return static_msg(err, err_len, 1, "fsb_rows_gemm_static: M = " + std::to_string(M) + " rows exceeds the static scratch limit of " +
                          std::to_string(FSB_STATIC_ROWS_MAX) + " rows (" + std::to_string(FSB_STATIC_SCRATCH_BYTES >> 20) +
                          " MiB of partial tiles); use fsb_rows_gemm with caller scratch");
    if (epi != RG_EPI_STORE && epi != RG_EPI_RESIDUAL)
        return static_msg(err, err_len, 1, "fsb_rows_gemm_static: epi must be 0 (store) or 1 (residual); SwiGLU needs fsb_rows_gemm");
    if (epi == RG_EPI_RESIDUAL && !res) return static_msg(err, err_len, 1, "fsb_rows_gemm_static: residual epilogue without res");
    const int n_tiles = (N + RG_BM - 1) / RG_BM, kb = K / RG_BK;
    if (n_tiles > RG_MAX_TILES || (long long)n_tiles * kb * RG_GRID_MAX >= (1LL << 31))
        return static_msg(err, err_len, 1, "fsb_rows_gemm_static: N x K too large for the 32-bit stream-K range arithmetic");
    int dev = 0, n_sm = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
        return static_msg(err, err_len, 2, "fsb_rows_gemm_static: no current CUDA device");
    if (n_sm > RG_SM_MAX) return static_msg(err, err_len, 2, "fsb_rows_gemm_static: device has more SMs than the scratch is sized for");
    const size_t need = rg_partial_bytes(rg_grid(n_tiles, kb, n_sm), rg_bn_for(M));
    if (need > FSB_STATIC_SCRATCH_BYTES)
        return static_msg(err, err_len, 1, "fsb_rows_gemm_static: this shape needs " + std::to_string(need) + " bytes of partial tiles, the static scratch holds " +
                          std::to_string(FSB_STATIC_SCRATCH_BYTES));
    cudaError_t e = rg_set_attrs();
    if (e != cudaSuccess) return static_msg(err, err_len, 2, std::string("fsb_rows_gemm_static: cudaFuncSetAttribute: ") + cudaGetErrorString(e));
    void* scratch = nullptr;
    e = static_scratch_ptr(&scratch);
    if (e != cudaSuccess) return static_msg(err, err_len, 2, std::string("fsb_rows_gemm_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
    CUtensorMap mW, mX;
    if (!rg_make_map(&mW, W, (uint64_t)N, (uint64_t)K, RG_BM) || !rg_make_map(&mX, x, (uint64_t)M, (uint64_t)K, (uint32_t)rg_bn_for(M)))
        return static_msg(err, err_len, 3, "fsb_rows_gemm_static: cuTensorMapEncodeTiled failed (16-byte aligned pointers required)");
    // nothing in the scratch has to start at zero: the GEMM writes every partial-tile element the finish pass reads (rows < M
    // of the split tiles) before that pass runs, so the previous call's contents are never observed
    float* partial = (float*)scratch;
    cudaStream_t s = (cudaStream_t)stream;
    e = rg_launch(s, false, n_sm, mW, mW, mX, M, N, n_tiles, K, epi, (const bf16*)res, (bf16*)out, partial, 0, nullptr);
    if (e == cudaSuccess) e = rg_finish_tiles(s, false, n_sm, M, N, K, epi, partial, (const bf16*)out, (const bf16*)res, (bf16*)out);
    if (e != cudaSuccess) return static_msg(err, err_len, 4, std::string("fsb_rows_gemm_static: launch failed: ") + cudaGetErrorString(e));
    return 0;
}

int fsb_rows_gemm_narrow(const void* W, const void* x, void* out, int32_t M, int32_t N, int32_t K, int32_t nw, void* stream) {
    if (!W || !x || !out || M < 1 || M > 128 || K < RG_BK || K % RG_BK || nw < 16 || nw > 64 || nw % 16 || N < nw || N % nw) return 1;
    CUtensorMap mW, mX;
    if (!rg_make_map(&mW, W, (uint64_t)N, (uint64_t)K, (uint32_t)nw)) return 3;
    if (!rg_make_map(&mX, x, (uint64_t)M, (uint64_t)K, 128)) return 3;
    return rgn_launch((cudaStream_t)stream, false, mW, mX, M, N, K, nw, (bf16*)out) == cudaSuccess ? 0 : 4;
}

int fsb_sample(const void* logits, const void* bias, int32_t B, int32_t V, int32_t top_k, float temperature, float top_p,
               int32_t n_variants, float temp2, float top_p2, const void* uniforms, uint64_t seed, int32_t seg0_lo,
               int32_t seg0_hi, int32_t seg1_lo, int32_t seg1_hi, int32_t* out, void* stream) {
    if (!logits || !out || B < 1 || V < 1 || top_k < 1 || top_k > SAMPLE_MAXK || n_variants < 1 || n_variants > 2) return 1;
    SampleArgs sa{};
    sa.logits = (const bf16*)logits; sa.bias = (const bf16*)bias; sa.V = V; sa.ld_logits = V;
    if (seg0_hi <= seg0_lo && seg1_hi <= seg1_lo) { seg0_lo = 0; seg0_hi = V; }
    sa.seg_lo[0] = seg0_lo; sa.seg_hi[0] = seg0_hi; sa.seg_lo[1] = seg1_lo; sa.seg_hi[1] = seg1_hi;
    sa.top_k = top_k; sa.n_variants = n_variants;
    sa.temp[0] = temperature; sa.top_p[0] = top_p; sa.temp[1] = temp2; sa.top_p[1] = top_p2;
    sa.uniforms = (const bf16*)uniforms; sa.u_stride = (long long)n_variants * V; sa.u_offset = 0;
    sa.seed = seed; sa.frame_counter = nullptr; sa.draw_id = 0;
    sa.mode = 2; sa.tokens = out; sa.out_col = 0; sa.n_cols = 2;
    cudaError_t e = launch(sample_kernel, dim3(B), dim3(SAMPLE_THREADS), 0, (cudaStream_t)stream, false, sa);
    return e == cudaSuccess ? 0 : 3;
}

}  // extern "C"
