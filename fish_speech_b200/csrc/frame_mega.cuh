// Persistent frame kernel: one cooperative launch executes a whole Dual-AR frame at batch 1
// (decode_one_token_ar, inference.py:96-181 + the bookkeeping of decode_n_tokens, :224-233).
//
// Why: at batch 1 a frame is ~400 dependent ops over 16.3 GB of weights.  As separate kernels every boundary drains
// the SM's bytes-in-flight (HBM3e needs ~160 KB in flight per SM) and costs 3-5 us; measured 0.44 of the HBM roofline.
// Here the weight stream never stops:
//   * grid = one CTA per SM (cooperative launch), 16 consumer warps + 1 producer warp;
//   * the producer walks the frame's op list and streams EVERY GEMV's weight rows of this CTA through one ring of
//     MG_NST x 20 KB shared-memory stages, one contiguous cp.async.bulk per stage (whole rows: R rows x K columns,
//     16-20 KB per copy), throttled only by the ring — it runs ahead of the dependent chain across op boundaries;
//   * consumers execute the ops in order.  Activations travel between ops as flag-in-data words (common.cuh): the
//     consumer of a vector spins on the tags of the words it has to load anyway, so dependent ops need no fence, no
//     grid barrier and no reload; a grid barrier (release/acquire counter) remains only in front of the sampling ops,
//     which read plain bf16 logits;
//   * a stage holds R whole rows and belongs to ONE of four consumer groups (4 warps each): stage n -> group n % 4, ring
//     slot n % 8, so a slot's full/empty barriers are only ever used by its own group (no cross-group coupling) and
//     four stages are in different phases of wait / FMA / reduce / epilogue at any time;
//     R >= 4: a warp owns whole rows (no cross-warp reduction); R < 4 (long rows): the group's warps split a row into
//     column segments and meet in shared memory in a fixed order (deterministic);
//   * attention / sampling / embedding / frame bookkeeping run as in-kernel ops on a subset of warps (the same device
//     bodies the standalone kernels use).
#pragma once
#include "attn.cuh"
#include "gemv_tma.cuh"
#include "sample.cuh"

#define MG_CONSUMERS 512
#define MG_THREADS (MG_CONSUMERS + 32)
#define MG_STAGE_BYTES 20480
#define MG_NST 8  // power of two
#define MG_CPL 10  // 128-bit chunks per lane per row segment (segment <= 2560 columns)
#define MG_GROUPS 4
#define MG_GROUP_WARPS 4
#define MG_X_BYTES 20480
#define MG_SMEM (MG_NST * MG_STAGE_BYTES + MG_X_BYTES)
#define MG_RES_MAX 64  // rows of one CTA whose residual values are prefetched per op (dim <= 64 x grid)

enum { MOP_EMBED = 0, MOP_GEMV = 1, MOP_ATTN = 2, MOP_FATTN = 3, MOP_SAMPLE = 4, MOP_END = 5 };
enum { MSYNC_NONE = 0, MSYNC_GRID = 1, MSYNC_CTA = 2 };

struct MegaGemv {
    const bf16 *W, *W2, *norm_w;
    const uint32_t *x_t, *res_t;  // flag-in-data input vector [K] / residual [N]
    uint32_t* out_t;              // flag-in-data output [N], or nullptr when the output is plain bf16 logits
    bf16* out;                    // plain output (LM heads), or nullptr
    bf16* xnorm_out;              // plain copy of the normalised input ("hidden"), CTA 0 only, or nullptr
    uint32_t* xnorm_out_t;        // flag-in-data copy of the same (tag = this op)
    int N, K, pro, epi;
    float eps;
    int rshift;  // log2(rows per stage); rows * K * 2 <= MG_STAGE_BYTES
    int x_op;    // index of the op that produced x_t (its tag)
    int rows_q, rows_rem;  // N / grid, N % grid: CTA b owns rows [b*q + min(b, rem), +q + (b < rem)) — no division on device
    int pad_[5];
};
static_assert(sizeof(MegaGemv) % 16 == 0, "MegaGemv is copied in 16-byte words");
struct MegaOp { int type, idx, sync, pad; };
struct MegaPlan {
    const MegaOp* ops;
    int n_ops, B;
    const MegaGemv* gemv;
    int n_gemv;  // GEMV ops are stored in frame order: the producer walks this array alone
    int n_attn_slow;  // attn[0 .. n_attn_slow) are the slow layers in order
    const AttnArgs* attn;
    const SampleArgs* samp;
    const EmbedArgs* embed;
    const FrameEndArgs* fend;
    unsigned* sync;  // [0] barrier arrivals, [1] exits (both zero between launches), [2] launch sequence number
    unsigned long long* prof;  // optional [16] cycle counters of CTA 0 (profiling aid), or nullptr
    const uint4* hot;          // [n_hot] {ptr lo, ptr hi, bytes, 0}: small operands every CTA reads on the critical path
    int n_hot;                 // (norm weights, RoPE rows, sampling bias): prefetched into L2 with evict_last at launch
    unsigned long long* trace; // optional [grid][n_ops][4] globaltimer stamps (op start, x staged, op end, barrier passed)
    int l2_ahead;              // stages per producer lane prefetched into L2 while the ring is full (0 = off)
    int skip_dot;              // experiment: consumers wait for and release the stages without reading them (results are garbage)
    int detail_op;             // profiling: op whose per-iteration stamps go to `detail` (-1 = none)
    int detail_gemv;           //            its index in the GEMV list (producer side)
    unsigned long long* detail; // [grid][64]: group 0's lane 0: per iteration {stage full seen, dot done, epilogue done}; producer lane 0: [32..] issue times
};
enum { MP_XSTAGE = 0, MP_LOOP, MP_GRID, MP_ATTN, MP_FATTN, MP_SAMPLE, MP_MISC, MP_TOTAL, MP_FRAMES };

typedef SyncNamed<1, MG_CONSUMERS> MgAll;

// tag of op `oi` in the launch with sequence number `seq`: unique among everything written in this and the previous
// launch (every activation word is rewritten at least once per launch), never 0 in the first launches
__device__ __forceinline__ unsigned mg_tag(unsigned seq, int oi) { return ((seq << 9) + (unsigned)oi) & 0xffffu; }

__device__ __forceinline__ unsigned long long mg_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// all consumer threads of all CTAs; `epoch` counts the barriers this launch has passed
__device__ __forceinline__ void mg_grid_sync(unsigned* ctr, unsigned& epoch, int nb, int tid) {
    ++epoch;
    MgAll::sync();
    if (tid == 0) {
        __threadfence();
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
        const unsigned target = epoch * (unsigned)nb;
        unsigned v, spins = 0;
        long long t0 = 0;
        do {
            asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
            if (v < target && fsb_watchdog(spins, t0, 0x3ffu, FSB_ERR_GRID_WAIT)) break;  // a lost CTA must end, not hang the GPU
        } while (v < target);
        __threadfence();
    }
    MgAll::sync();
}

// stage iteration space of one GEMV op for one CTA, shared by producer and consumers.
// it -> (block of 4 row groups, matrix q, j): row group gi = 4*blk + j.  `total` is a multiple of 4, so with the global
// stage counter n the owner group of a stage is n % 4 == j for every op; for SwiGLU the w1 and w3 stages of a row group
// are 4 apart, i.e. they reach the same warps.  Row groups past the end are EMPTY stages (0 bytes) so that nothing is
// ever skipped and slot ownership stays fixed.
struct MgGeom {
    int r0, r1, ngroups, total, ns, rshift;
    __device__ __forceinline__ MgGeom(int rows_q, int rows_rem, int epi, int rshift_, int b) {
        rshift = rshift_;
        r0 = b * rows_q + min(b, rows_rem);
        r1 = r0 + rows_q + (b < rows_rem ? 1 : 0);
        ngroups = (r1 - r0 + (1 << rshift) - 1) >> rshift;
        ns = (epi == EPI_SWIGLU) ? 1 : 0;
        total = ((ngroups + 3) >> 2) << (2 + ns);
    }
    __device__ __forceinline__ void decode(int it, int& q, int& gi) const {
        q = (it >> 2) & ns;
        gi = ((it >> (2 + ns)) << 2) + (it & 3);
    }
};

__device__ __forceinline__ void bulk_g2s_evict_first(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar,
                                                     uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
            smem_u32(dst_smem)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
        : "memory");
}
__device__ __forceinline__ bool mbar_test(uint64_t* b, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(b)), "r"(parity)
        : "memory");
    return ok != 0;
}

// Producer warp: lane l (< MG_NST) owns ring slot l, i.e. the stages n = l, l + 8, l + 16, ... of this CTA's stream over
// ALL GEMV ops of the frame.  Eight lanes poll their slots' empty barriers and issue their bulk copies independently
// (a single in-order producer thread was the bottleneck: ~450 cycles of dependent integer code per 20 KB stage).
// Each lane walks the GEMV list on its own; the list is immutable, so it is read through the read-only path (L1).
struct MgLaneCursor {
    const MegaGemv* gl;
    int n_gemv, b, j, it;
    int K, rshift, r0, r1, ngroups, total, ns;
    const bf16 *W, *W2;
    bool live;
    __device__ __forceinline__ void load(int jj) {
        const MegaGemv* g = gl + jj;
        const MgGeom geo(__ldg(&g->rows_q), __ldg(&g->rows_rem), __ldg(&g->epi), __ldg(&g->rshift), b);
        K = __ldg(&g->K);
        W = (const bf16*)__ldg((const unsigned long long*)&g->W);
        W2 = (const bf16*)__ldg((const unsigned long long*)&g->W2);
        rshift = geo.rshift; r0 = geo.r0; r1 = geo.r1; ngroups = geo.ngroups; total = geo.total; ns = geo.ns;
    }
    __device__ __forceinline__ void normalise() {
        while (live && it >= total) {
            it -= total;
            if (++j >= n_gemv) { live = false; break; }
            load(j);
        }
    }
    __device__ __forceinline__ MgLaneCursor(const MegaPlan& p, int b_, int first)
        : gl(p.gemv), n_gemv(p.n_gemv), b(b_), j(0), it(first) {
        live = n_gemv > 0;
        if (live) { load(0); normalise(); }
    }
    __device__ __forceinline__ void stage(const bf16*& src, uint32_t& bytes) const {
        const int q = (it >> 2) & ns, gi = ((it >> (2 + ns)) << 2) + (it & 3);
        if (gi >= ngroups) { src = W; bytes = 0; return; }
        const int g0 = r0 + (gi << rshift);
        const int ng = min(1 << rshift, r1 - g0);
        src = (q ? W2 : W) + (size_t)g0 * K;
        bytes = (uint32_t)ng * (uint32_t)K * 2u;
    }
    __device__ __forceinline__ void advance(int k) { it += k; normalise(); }
};

template <int HD>
__device__ __forceinline__ void mg_producer(const MegaPlan& p, unsigned char* stages, uint64_t* full_bar,
                                            uint64_t* empty_bar, int b, int lane) {
    if (lane >= MG_NST) return;
    uint64_t policy;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
    MgLaneCursor cur(p, b, lane);
    // L2 look-ahead: while this lane's slot is still full (the consumers sit in an attention / sampling op or wait for an
    // activation), it prefetches its OWN future stages into L2, at most `l2_ahead` stages beyond the one it will load next.
    // HBM keeps streaming through the bubbles of the dependent chain and the ring refills from L2 afterwards; nothing is
    // prefetched while the lane can load (no competition with the demand stream).
    MgLaneCursor pf = cur;
    int ahead = 0;  // stages `pf` is ahead of `cur`
    const int max_ahead = p.l2_ahead;
    unsigned char* dst = stages + (size_t)lane * MG_STAGE_BYTES;
    unsigned round = 0, spins = 0;
    long long t0 = 0;
    while (cur.live) {
        if (!mbar_test(&empty_bar[lane], (round & 1u) ^ 1u)) {
            if (ahead < max_ahead) {
                if (ahead == 0) { pf = cur; pf.advance(MG_NST); ahead = 1; }
                else if (pf.live) { pf.advance(MG_NST); ++ahead; }
                if (pf.live) {
                    const bf16* psrc;
                    uint32_t pbytes;
                    pf.stage(psrc, pbytes);
                    if (pbytes) l2_prefetch_bulk(psrc, pbytes);
                } else {
                    ahead = max_ahead;  // nothing left to prefetch
                }
                continue;
            }
            if (spin_watchdog(spins, t0, FSB_ERR_RING_WAIT)) break;
            continue;
        }
        const bf16* src;
        uint32_t bytes;
        cur.stage(src, bytes);
        if (p.detail && cur.j == p.detail_gemv && cur.it < 16)  // profiling: when each stage of the chosen op was requested
            p.detail[(size_t)b * 64 + 32 + cur.it] = mg_now();
        if (bytes == 0) {
            mbar_arrive(&full_bar[lane]);  // empty stage: the phase completes at once
        } else {
            mbar_expect_tx(&full_bar[lane], bytes);
            bulk_g2s_evict_first(dst, src, bytes, &full_bar[lane], policy);
        }
        ++round;
        cur.advance(MG_NST);
        if (ahead > 0) --ahead;  // `pf` keeps its position: it is now one stage less ahead of `cur`
    }
}

__device__ __forceinline__ void mg_group_sync(int group) {
    asm volatile("bar.sync %0, %1;" ::"r"(2 + group), "n"(32 * MG_GROUP_WARPS) : "memory");
}

// epilogue of one output row (lane 0 of the finalising warp); rounding points as in gemv.cuh
__device__ __forceinline__ void mg_finish_row(const MegaGemv& g, size_t oi, float s, float s2, float resv,
                                              unsigned tag_out) {
    float o = rbf(s);
    if (g.epi == EPI_RESIDUAL) {
        o = resv + o;
    } else if (g.epi == EPI_SWIGLU) {
        const float u = rbf(s2);
        const float sl = rbf(__fdiv_rn(o, 1.f + expf(-o)));
        o = sl * u;
    }
    if (g.out_t != nullptr) st_tagged(g.out_t + oi, o, tag_out);
    else g.out[oi] = __float2bfloat16_rn(o);
}

// dot product of one row segment in shared memory with the activation chunks of the same columns: chunk c of the segment
// goes to lane (c - c_lo) % 32; the math is the native bf16 x bf16 + fp32 FMA (dot8_bf16), bit-identical to unpack + FFMA.
__device__ __forceinline__ float mg_dot(const uint4* w0, const uint4* xs, int c_lo, int c_hi, int lane) {
    // MG_CPL chunks per lane in two batches of loads-then-math (x is re-read from shared memory: holding it in registers
    // does not fit the 96-register cap); a fixed shape with predicated loads measured faster than a warp-uniform trip count
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int hb = 0; hb < 2; ++hb) {
        uint4 wv[MG_CPL / 2], xv[MG_CPL / 2];
#pragma unroll
        for (int i = 0; i < MG_CPL / 2; ++i) {
            const int c = c_lo + lane + 32 * (hb * (MG_CPL / 2) + i);
            const bool ok = c < c_hi;
            wv[i] = ok ? w0[c] : make_uint4(0, 0, 0, 0);
            xv[i] = ok ? xs[c] : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int i = 0; i < MG_CPL / 2; ++i) dot8_bf16(wv[i], xv[i], acc);
    }
    return (acc[0] + acc[1]) + (acc[2] + acc[3]);
}
__device__ __forceinline__ void mg_gemv(const MegaGemv* gp, unsigned char* stages, uint4* xs, uint64_t* full_bar,
                                        uint64_t* empty_bar, float* red, float* res_s, float (*pbuf)[2][2][4], unsigned& n, int b,
                                        int nb, int tid, unsigned seq, int oi, unsigned long long* prof,
                                        unsigned long long* trace, unsigned long long* detail, bool skip_dot) {
    const long long tp0 = prof ? clock64() : 0;
    const int lane = tid & 31, warp = tid >> 5;
    const int grp = warp >> 2, wg = warp & 3;
    MegaGemv g;
    {   // immutable plan data: read-only path (L1); the next op's entry was prefetched while the previous op ran
        const uint4* src = reinterpret_cast<const uint4*>(gp);
        uint4* dst = reinterpret_cast<uint4*>(&g);
#pragma unroll
        for (int i = 0; i < (int)(sizeof(MegaGemv) / 16); ++i) dst[i] = __ldg(src + i);
    }
    const int K = g.K, K8 = K >> 3;
    const unsigned tag_in = mg_tag(seq, g.x_op), tag_out = mg_tag(seq, oi);
    // ---- stage the activation row: spin on the producer's tags while loading (this IS the inter-op dependency), keep
    //      the chunks in registers for the RMSNorm second pass (llama.py:995-1000, double rounding)
    constexpr int XC = (MG_X_BYTES / 16 + MG_CONSUMERS - 1) / MG_CONSUMERS;  // chunks of 8 elements per thread
    uint4 nw[XC];  // RMSNorm weights of this thread's chunks: requested before the spin, they do not depend on x
#pragma unroll
    for (int i = 0; i < XC; ++i) {
        const int c = tid + i * MG_CONSUMERS;
        nw[i] = (g.pro == PRO_RMSNORM && c < K8) ? __ldg(reinterpret_cast<const uint4*>(g.norm_w) + c) : make_uint4(0, 0, 0, 0);
    }
    // residual values of this CTA's rows (EPI_RESIDUAL: wo, w2): ONE load per row, requested before the spin below so that
    // its L2 round trip hides behind the tag wait.  A load per stage inside the loop put ~0.7 us on every round of the
    // multi-warp-per-row layouts (ncu/trace: w2 9.2 us for 7.6 us of streaming, wo 6 us on stages already resident).
    const MgGeom geo0(g.rows_q, g.rows_rem, g.epi, g.rshift, b);
    const bool res_pre = (g.epi == EPI_RESIDUAL) && (geo0.r1 - geo0.r0 <= MG_RES_MAX);
    float res_mine = 0.f;
    if (res_pre && tid < geo0.r1 - geo0.r0) res_mine = ld_tagged_value(g.res_t + (size_t)(geo0.r0 + tid));
    uint4 xv[XC];
    {
        // all of this thread's words are requested together and re-requested together until every tag is current: one
        // L2 round trip once the producer is done (a word-by-word spin would serialise up to 2*XC round trips)
        unsigned spins = 0;
        long long t0 = 0;
        bool ok;
        do {
            ok = true;
#pragma unroll
            for (int i = 0; i < XC; ++i) {
                const int c = tid + i * MG_CONSUMERS;
                uint2 lo = make_uint2(0, 0), hi = make_uint2(0, 0);
                if (c < K8) {
                    ok &= ld_tagged4_try(g.x_t + (size_t)c * 8, tag_in, lo);
                    ok &= ld_tagged4_try(g.x_t + (size_t)c * 8 + 4, tag_in, hi);
                }
                xv[i] = make_uint4(lo.x, lo.y, hi.x, hi.y);
            }
            if (!ok && spin_watchdog(spins, t0)) break;
        } while (!ok);
    }
    if (g.pro == PRO_RMSNORM) {
        float ss = 0.f;
#pragma unroll
        for (int i = 0; i < XC; ++i) {
            float f[8];
            unpack8(xv[i], f);
#pragma unroll
            for (int e = 0; e < 8; ++e) ss += f[e] * f[e];
        }
        ss = group_sum<MgAll>(ss, red, tid);
        const float rs = __frsqrt_rn(__fadd_rn(__fdiv_rn(ss, (float)K), g.eps));
#pragma unroll
        for (int i = 0; i < XC; ++i) {
            const int c = tid + i * MG_CONSUMERS;
            if (c < K8) {
                float f[8], w[8];
                unpack8(xv[i], f);
                unpack8(nw[i], w);
#pragma unroll
                for (int e = 0; e < 8; ++e) f[e] = rbf(__fmul_rn(f[e], rs)) * w[e];
                uint4 o;
                o.x = pack2(f[0], f[1]); o.y = pack2(f[2], f[3]); o.z = pack2(f[4], f[5]); o.w = pack2(f[6], f[7]);
                if (b == 0 && g.xnorm_out != nullptr) reinterpret_cast<uint4*>(g.xnorm_out)[c] = o;
                if (b == 0 && g.xnorm_out_t != nullptr) {
                    st_tagged4(g.xnorm_out_t + (size_t)c * 8, o.x, o.y, tag_out);
                    st_tagged4(g.xnorm_out_t + (size_t)c * 8 + 4, o.z, o.w, tag_out);
                }
                xs[c] = o;
            }
        }
    } else {
#pragma unroll
        for (int i = 0; i < XC; ++i) {
            const int c = tid + i * MG_CONSUMERS;
            if (c < K8) xs[c] = xv[i];
        }
    }
    if (res_pre && tid < geo0.r1 - geo0.r0) res_s[tid] = res_mine;
    MgAll::sync();
    const long long tp1 = prof ? clock64() : 0;
    if (trace && tid == 0) trace[1] = mg_now();

    const MgGeom geo(g.rows_q, g.rows_rem, g.epi, g.rshift, b);
    const int R = 1 << geo.rshift;
    // rows per warp / warps per row inside a stage: R = 8: 2 rows per warp; 4: 1; 2: two warps per row; 1: four
    const int rpw = (R >= 8) ? 2 : 1;
    const int seg_shift = (geo.rshift >= 2) ? 0 : 2 - geo.rshift, SEGW = 1 << seg_shift;
    const int row0 = (geo.rshift >= 2) ? wg : (wg >> seg_shift);
    const int seg = wg & (SEGW - 1);
    const int cpw = (K8 + SEGW - 1) >> seg_shift;  // chunks per segment (<= 32 * MG_CPL)
    const int c_lo = seg * cpw, c_hi = min(K8, c_lo + cpw);
    const size_t row_bytes = (size_t)K * 2;
    float acc0[2] = {0.f, 0.f};  // w1 row sums waiting for the w3 stage (SwiGLU, whole-row warps)
    int par = 0;
    const unsigned n0 = n;
    for (int it = grp; it < geo.total; it += MG_GROUPS) {
        int q, gi;
        geo.decode(it, q, gi);
        const unsigned ns_ = n0 + (unsigned)it;
        const unsigned slot = ns_ & (MG_NST - 1), ph = (ns_ / MG_NST) & 1u;
        const int g0 = geo.r0 + (gi << geo.rshift);
        const int ng = (gi < geo.ngroups) ? min(R, geo.r1 - g0) : 0;
        const bool last_q = (q == geo.ns);
        // the residual is fetched before the stage is waited for: its L2 latency hides behind the wait and the FMAs
        float resv[2] = {0.f, 0.f};
        if (g.epi == EPI_RESIDUAL && lane == 0 && seg == 0) {
#pragma unroll
            for (int r = 0; r < 2; ++r)
                if (r < rpw && row0 + 4 * r < ng)
                    resv[r] = res_pre ? res_s[g0 + row0 + 4 * r - geo.r0] : ld_tagged_value(g.res_t + (size_t)(g0 + row0 + 4 * r));
        }
        mbar_wait(&full_bar[slot], ph);
        const int dit = (it - grp) / MG_GROUPS;
        if (detail && tid == 0 && dit < 10) detail[dit * 3 + 0] = mg_now();
        const unsigned char* st = stages + (size_t)slot * MG_STAGE_BYTES;
        float part[2] = {0.f, 0.f};
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            const int row = row0 + 4 * r;
            if (r < rpw && row < ng && !skip_dot)
                part[r] = warp_sum(mg_dot(reinterpret_cast<const uint4*>(st + row * row_bytes), xs, c_lo, c_hi, lane));
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[slot]);
        if (detail && tid == 0 && dit < 10) detail[dit * 3 + 1] = mg_now();
        if (SEGW == 1) {
            if (lane == 0) {
#pragma unroll
                for (int r = 0; r < 2; ++r) {
                    const int row = row0 + 4 * r;
                    if (r < rpw && row < ng) {
                        if (!last_q) acc0[r] = part[r];
                        else mg_finish_row(g, (size_t)(g0 + row), geo.ns ? acc0[r] : part[r], part[r], resv[r], tag_out);
                    }
                }
            }
        } else {
            if (lane == 0) pbuf[grp][par][q][wg] = part[0];
            if (last_q) {
                mg_group_sync(grp);
                if (seg == 0 && row0 < ng && lane == 0) {
                    float s = 0.f, s2 = 0.f;
                    for (int j = 0; j < SEGW; ++j) s += pbuf[grp][par][0][wg + j];
                    if (geo.ns)
                        for (int j = 0; j < SEGW; ++j) s2 += pbuf[grp][par][1][wg + j];
                    mg_finish_row(g, (size_t)(g0 + row0), s, s2, resv[0], tag_out);
                }
                par ^= 1;
            }
        }
        if (detail && tid == 0 && dit < 10) detail[dit * 3 + 2] = mg_now();
    }
    n = n0 + (unsigned)geo.total;
    if (prof && tid == 0) {
        prof[MP_XSTAGE] += (unsigned long long)(tp1 - tp0);
        prof[MP_LOOP] += (unsigned long long)(clock64() - tp1);
    }
}

// LONG = true: the slow-AR attention walks a split's keys in tiles (caches beyond ASP_MAXS x asp_max_keys tokens); a separate
// instantiation so that the short-context kernel keeps its code and register allocation
template <int HD, int GQ, bool LONG = false>
__global__ void __launch_bounds__(MG_THREADS, 1) frame_mega_kernel(MegaPlan p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[MG_NST];
    __shared__ __align__(8) uint64_t empty_bar[MG_NST];
    __shared__ float red[32];
    __shared__ float res_s[MG_RES_MAX];
    __shared__ float pbuf[MG_GROUPS][2][2][4];  // [group][parity][matrix][warp of group]
    __shared__ AttnSplitSmem<HD, GQ> asm_;

    const int tid = threadIdx.x, warp = tid >> 5;
    const int nb = gridDim.x, b = blockIdx.x;
    unsigned char* stages = smem_raw;
    uint4* xs = reinterpret_cast<uint4*>(smem_raw + (size_t)MG_NST * MG_STAGE_BYTES);

    if (tid == 0) {
        for (int s = 0; s < MG_NST; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], MG_GROUP_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (warp == MG_CONSUMERS / 32) {  // producer warp: weights only, never waits for an activation
        mg_producer<HD>(p, stages, full_bar, empty_bar, b, tid & 31);
        return;
    }

    typedef SyncNamed<6, 32 * GQ> FastSync;
    unsigned n = 0, epoch = 0;
    const unsigned seq = p.sync[2];  // written by the previous launch's last CTA
    if (tid < 32) {  // launch-time L2 prefetches (warp 0; nothing waits for them)
        uint64_t keep;
        asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(keep));
        for (int i = b * 32 + tid; i < p.n_hot; i += nb * 32) {
            const uint4 hr = __ldg(p.hot + i);
            const void* ptr = (const void*)(((unsigned long long)hr.y << 32) | hr.x);
            asm volatile("cp.async.bulk.prefetch.L2.global.L2::cache_hint [%0], %1, %2;" ::"l"(ptr), "r"(hr.z), "l"(keep) : "memory");
        }
        if (tid == 0 && p.n_attn_slow > 0) {
            const AttnArgs* a0 = p.attn;
            const int splits = a0->splits;
            if (b < a0->KV * splits * p.B) {
                const int hm = b / splits;
                attn_split_prefetch<HD>(*a0, a0->kcache, a0->vcache, b % splits, hm % a0->KV, hm / a0->KV);
            }
        }
    }
    unsigned long long* prof = (b == 0) ? p.prof : nullptr;
    const long long tk0 = prof ? clock64() : 0;
    for (int oi = 0; oi < p.n_ops; ++oi) {
        MegaOp op;
        *reinterpret_cast<uint4*>(&op) = __ldg(reinterpret_cast<const uint4*>(p.ops + oi));
        if (tid == 0 && oi + 1 < p.n_ops) {  // pull the next op's descriptors into L1 while this op runs
            const MegaOp nx = p.ops[oi + 1];
            asm volatile("prefetch.global.L1 [%0];" ::"l"(p.ops + oi + 2));
            if (op.type == MOP_GEMV && nx.type == MOP_ATTN) {  // pull this CTA's KV split into L2 one op ahead
                const AttnArgs* an = p.attn + nx.idx;
                const int splits = __ldg(&an->splits);
                if (b < __ldg(&an->KV) * splits * p.B) {
                    const AttnArgs a2 = *an;
                    const int hm = b / splits;
                    attn_split_prefetch<HD>(a2, a2.kcache, a2.vcache, b % splits, hm % a2.KV, hm / a2.KV);
                }
            }
            if (nx.type == MOP_GEMV) {
                asm volatile("prefetch.global.L1 [%0];" ::"l"(p.gemv + nx.idx));
                asm volatile("prefetch.global.L1 [%0];" ::"l"((const char*)(p.gemv + nx.idx) + 64));
            }
        }
        const long long to0 = prof ? clock64() : 0;
        unsigned long long* tr = p.trace ? p.trace + ((size_t)b * p.n_ops + oi) * 4 : nullptr;
        if (tr && tid == 0) tr[0] = mg_now();
        switch (op.type) {
        case MOP_GEMV:
            mg_gemv(p.gemv + op.idx, stages, xs, full_bar, empty_bar, red, res_s, pbuf, n, b, nb, tid, seq, oi, prof, tr, (p.detail && oi == p.detail_op) ? p.detail + (size_t)b * 64 : nullptr, p.skip_dot != 0);
            break;
        case MOP_EMBED:
            if (b < p.B) embed_body(*p.embed, b, tid, MG_CONSUMERS, mg_tag(seq, oi));
            break;
        case MOP_ATTN: {
            const AttnArgs a = p.attn[op.idx];
            const int items = a.KV * a.splits * p.B;
            for (int item = b; item < items; item += nb) {
                const int split = item % a.splits, hm = item / a.splits;
                if (tid == 0 && a.next_k != nullptr) attn_split_prefetch<HD>(a, a.next_k, a.next_v, split, hm % a.KV, hm / a.KV);
                attn_split_body<HD, GQ, MgAll, LONG>(a, asm_, split, hm % a.KV, hm / a.KV, tid, mg_tag(seq, a.tag_in_op),
                                               mg_tag(seq, a.tag_out_op), tr);
                MgAll::sync();
            }
            break;
        }
        case MOP_FATTN:
            if (tid < 32 * GQ) {
                const AttnArgs a = p.attn[op.idx];
                const int items = a.KV * p.B;
                for (int item = b; item < items; item += nb) {
                    fast_attn_body<HD, GQ, FastSync, true>(a, item % a.KV, item / a.KV, tid, mg_tag(seq, a.tag_in_op),
                                                           mg_tag(seq, a.tag_out_op));
                    FastSync::sync();
                }
            }
            break;
        case MOP_SAMPLE:
            for (int r = b; r < p.B; r += nb) {
                sample_body<MgAll>(p.samp[op.idx], r, p.B, tid, mg_tag(seq, oi));
                MgAll::sync();
            }
            break;
        case MOP_END:
            if (b == 0) frame_end_body<MgAll>(*p.fend, tid);
            break;
        default: break;
        }
        const long long to1 = prof ? clock64() : 0;
        if (tr && tid == 0) tr[2] = mg_now();
        if (op.sync == MSYNC_GRID) mg_grid_sync(p.sync, epoch, nb, tid);
        else if (op.sync == MSYNC_CTA) { __threadfence_block(); MgAll::sync(); }
        if (tr && tid == 0) tr[3] = mg_now();
        if (prof && tid == 0) {
            const int slot = op.type == MOP_ATTN ? MP_ATTN : op.type == MOP_FATTN ? MP_FATTN
                           : op.type == MOP_SAMPLE ? MP_SAMPLE : MP_MISC;
            if (op.type != MOP_GEMV) prof[slot] += (unsigned long long)(to1 - to0);
            prof[MP_GRID] += (unsigned long long)(clock64() - to1);
        }
    }
    if (prof && tid == 0) { prof[MP_TOTAL] += (unsigned long long)(clock64() - tk0); prof[MP_FRAMES] += 1; }
    // leave the counters zeroed for the next launch: the last CTA out resets them (everyone has passed every barrier)
    MgAll::sync();
    if (tid == 0) {
        const unsigned old = atomicAdd(p.sync + 1, 1u);
        if (old == (unsigned)nb - 1u) {
            p.sync[0] = 0u;
            p.sync[1] = 0u;
            p.sync[2] = seq + 1u;
            __threadfence();
        }
    }
}
