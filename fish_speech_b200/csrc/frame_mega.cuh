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
//   * consumers execute the ops in order; dependent ops are separated by a grid barrier (release/acquire counter);
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

enum { MOP_EMBED = 0, MOP_GEMV = 1, MOP_ATTN = 2, MOP_FATTN = 3, MOP_SAMPLE = 4, MOP_END = 5 };
enum { MSYNC_NONE = 0, MSYNC_GRID = 1, MSYNC_CTA = 2 };

struct MegaGemv {
    const bf16 *W, *W2, *x, *norm_w, *res;
    bf16 *out, *xnorm_out;
    int N, K, pro, epi;
    float eps;
    int rshift;  // log2(rows per stage); rows * K * 2 <= MG_STAGE_BYTES
};
struct MegaOp { int type, idx, sync, pad; };
struct MegaPlan {
    const MegaOp* ops;
    int n_ops, B;
    const MegaGemv* gemv;
    const AttnArgs* attn;
    const SampleArgs* samp;
    const EmbedArgs* embed;
    const FrameEndArgs* fend;
    unsigned* sync;  // [0] barrier arrivals, [1] exits; both zero between launches
    unsigned long long* prof;  // optional [16] cycle counters of CTA 0 (profiling aid), or nullptr
    int pf_stages;             // L2 prefetch look-ahead of the weight stream, in ring stages (0 = off)
    int copy_bytes;            // bulk-copy granularity inside a stage (multiple of 16)
    unsigned long long* trace; // optional [grid][n_ops][4] globaltimer stamps (op start, x staged, op end, barrier passed)
};
enum { MP_XSTAGE = 0, MP_LOOP, MP_GRID, MP_ATTN, MP_FATTN, MP_SAMPLE, MP_MISC, MP_TOTAL, MP_FRAMES };

typedef SyncNamed<1, MG_CONSUMERS> MgAll;

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
            if (v < target && (++spins & 0x3ffu) == 0) {  // watchdog: a lost CTA must fault, not hang the GPU
                if (t0 == 0) t0 = clock64();
                else if (clock64() - t0 > 4000000000LL) __trap();
            }
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
    __device__ __forceinline__ MgGeom(int N, int epi, int rshift_, int b, int nb) {
        rshift = rshift_;
        r0 = (int)(((long long)N * b) / nb);
        r1 = (int)(((long long)N * (b + 1)) / nb);
        ngroups = (r1 - r0 + (1 << rshift) - 1) >> rshift;
        ns = (epi == EPI_SWIGLU) ? 1 : 0;
        total = ((ngroups + 3) >> 2) << (2 + ns);
    }
    __device__ __forceinline__ void decode(int it, int& q, int& gi) const {
        q = (it >> 2) & ns;
        gi = ((it >> (2 + ns)) << 2) + (it & 3);
    }
};

// walks the stage sequence of this CTA over all GEMV ops of the frame (the order consumers use); non-GEMV ops are
// reported through `other` so the look-ahead cursor can prefetch their operands
struct MgCursor {
    const MegaPlan& p;
    int b, nb, oi, it;
    bool have;
    int K, R, rshift, r0, r1, ngroups, total, ns;
    const bf16 *W, *W2;
    __device__ __forceinline__ MgCursor(const MegaPlan& p_, int b_, int nb_)
        : p(p_), b(b_), nb(nb_), oi(0), it(0), have(false) {}
    // false at the end of the frame; bytes == 0 for an empty stage
    template <class OtherOp>
    __device__ __forceinline__ bool next(const bf16*& src, uint32_t& bytes, OtherOp other) {
        for (;;) {
            if (!have) {
                while (oi < p.n_ops) {
                    const MegaOp op = p.ops[oi];
                    if (op.type == MOP_GEMV) break;
                    other(op);
                    ++oi;
                }
                if (oi >= p.n_ops) return false;
                const MegaGemv* g = p.gemv + p.ops[oi].idx;
                const MgGeom geo(g->N, g->epi, g->rshift, b, nb);
                K = g->K; W = g->W; W2 = g->W2; rshift = geo.rshift; R = 1 << rshift;
                r0 = geo.r0; r1 = geo.r1; ngroups = geo.ngroups; total = geo.total; ns = geo.ns;
                it = 0;
                have = true;
            }
            if (it < total) {
                const int i = it++;
                const int q = (i >> 2) & ns, gi = ((i >> (2 + ns)) << 2) + (i & 3);
                if (gi >= ngroups) { src = W; bytes = 0; return true; }
                const int g0 = r0 + (gi << rshift);
                const int ng = min(R, r1 - g0);
                src = (q ? W2 : W) + (size_t)g0 * K;
                bytes = (uint32_t)ng * (uint32_t)K * 2u;
                return true;
            }
            have = false;
            ++oi;
        }
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
__device__ __forceinline__ void l2_prefetch_range(const void* p, size_t bytes) {
    const char* c = (const char*)((uintptr_t)p & ~(uintptr_t)15);
    size_t n = (bytes + ((uintptr_t)p & 15) + 15) & ~(size_t)15;
    while (n > 0) {
        const uint32_t k = n > 32768 ? 32768u : (uint32_t)n;
        l2_prefetch_bulk(c, k);
        c += k; n -= k;
    }
}

// operands of the attention ops that do not depend on this frame's activations: this CTA's KV split, RoPE row, norms
template <int HD>
__device__ __forceinline__ void mg_prefetch_attn(const MegaPlan& p, const MegaOp& op, int b) {
    if (op.type != MOP_ATTN && op.type != MOP_FATTN) return;
    const AttnArgs* a = p.attn + op.idx;
    const int KV = a->KV;
    if (op.type == MOP_ATTN) {
        const int splits = a->splits;
        if (b >= KV * splits * p.B) return;
        const int split = b % splits, hm = b / splits, h = hm % KV, m = hm / KV;
        const int pos = a->pos[m];
        const int slot = a->slots ? a->slots[m] : m;
        const int per = (pos + 1 + splits - 1) / splits;
        const int k0 = split * per, k1 = min(pos, k0 + per);  // key `pos` itself is produced in this frame
        const size_t base = ((size_t)slot * KV + h) * a->S * HD;
        if (k1 > k0) {
            l2_prefetch_range(a->kcache + base + (size_t)k0 * HD, (size_t)(k1 - k0) * HD * 2);
            l2_prefetch_range(a->vcache + base + (size_t)k0 * HD, (size_t)(k1 - k0) * HD * 2);
        }
        l2_prefetch_range(a->rope + (size_t)pos * HD, HD * 2);
    } else {
        if (b >= KV * p.B) return;
        l2_prefetch_range(a->rope + (size_t)a->pos_fixed * HD, HD * 2);
    }
    if (a->qk_norm) { l2_prefetch_range(a->q_norm, HD * 2); l2_prefetch_range(a->k_norm, HD * 2); }
}

// producer thread: the smem ring is filled in order; an L2 prefetch cursor runs `pf_stages` stages ahead of it, so the
// HBM stream keeps going while the consumers sit in a barrier / attention / sampling op and the ring is full
template <int HD>
__device__ __forceinline__ void mg_producer(const MegaPlan& p, unsigned char* stages, uint64_t* full_bar,
                                            uint64_t* empty_bar, int b, int nb) {
    uint64_t policy;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
    MgCursor cp(p, b, nb), pf(p, b, nb);
    const bf16* src;
    uint32_t bytes;
    bool pf_live = p.pf_stages > 0;
    auto pf_other = [&](const MegaOp& op) { mg_prefetch_attn<HD>(p, op, b); };
    auto no_other = [](const MegaOp&) {};
    for (int i = 0; i < p.pf_stages && pf_live; ++i) {
        pf_live = pf.next(src, bytes, pf_other);
        if (pf_live && bytes) l2_prefetch_bulk(src, bytes);
    }
    unsigned n = 0;
    const uint32_t chunk = (uint32_t)p.copy_bytes;
    while (cp.next(src, bytes, no_other)) {
        if (pf_live) {
            const bf16* s2;
            uint32_t b2;
            pf_live = pf.next(s2, b2, pf_other);
            if (pf_live && b2) l2_prefetch_bulk(s2, b2);
        }
        const unsigned slot = n & (MG_NST - 1), ph = (n / MG_NST) & 1u;
        ++n;
        mbar_wait(&empty_bar[slot], ph ^ 1u);
        if (bytes == 0) { mbar_arrive(&full_bar[slot]); continue; }  // empty stage: the phase completes at once
        mbar_expect_tx(&full_bar[slot], bytes);
        unsigned char* dst = stages + (size_t)slot * MG_STAGE_BYTES;
        const char* sp = (const char*)src;
        for (uint32_t o = 0; o < bytes; o += chunk) {
            const uint32_t c = min(chunk, bytes - o);
            bulk_g2s_evict_first(dst + o, sp + o, c, &full_bar[slot], policy);
        }
    }
}

__device__ __forceinline__ void mg_group_sync(int group) {
    asm volatile("bar.sync %0, %1;" ::"r"(2 + group), "n"(32 * MG_GROUP_WARPS) : "memory");
}

// epilogue of one output row (lane 0 of the finalising warp); rounding points as in gemv.cuh
__device__ __forceinline__ void mg_finish_row(const MegaGemv& g, size_t oi, float s, float s2, float resv) {
    float o = rbf(s);
    if (g.epi == EPI_RESIDUAL) {
        o = resv + o;
    } else if (g.epi == EPI_SWIGLU) {
        const float u = rbf(s2);
        const float sl = rbf(__fdiv_rn(o, 1.f + expf(-o)));
        o = sl * u;
    }
    g.out[oi] = __float2bfloat16_rn(o);
}

// dot product of one row segment in shared memory with the activation chunks of the same columns, MG_CPL chunks per lane
// in two batches (x is re-read from shared memory: holding it in registers does not fit the 96-register cap)
__device__ __forceinline__ float mg_dot(const uint4* w0, const uint4* xs, int c_lo, int c_hi, int lane) {
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
        for (int i = 0; i < MG_CPL / 2; ++i) {
            float wf[8], xf[8];
            unpack8(wv[i], wf);
            unpack8(xv[i], xf);
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e & 3] = fmaf(wf[e], xf[e], acc[e & 3]);
        }
    }
    return (acc[0] + acc[1]) + (acc[2] + acc[3]);
}

__device__ __forceinline__ void mg_gemv(const MegaGemv* gp, unsigned char* stages, uint4* xs, uint64_t* full_bar,
                                        uint64_t* empty_bar, float* red, float (*pbuf)[2][2][4], unsigned& n, int b,
                                        int nb, int tid, unsigned long long* prof, unsigned long long* trace) {
    const long long tp0 = prof ? clock64() : 0;
    const int lane = tid & 31, warp = tid >> 5;
    const int grp = warp >> 2, wg = warp & 3;
    const MegaGemv g = *gp;
    const int K = g.K, K8 = K >> 3;
    // ---- stage the activation row (one rounding chain: llama.py:995-1000 RMSNorm, double rounding)
    const uint4* xg = reinterpret_cast<const uint4*>(g.x);
    if (g.pro == PRO_RMSNORM) {
        float ss = 0.f;
        for (int c = tid; c < K8; c += MG_CONSUMERS) {
            float f[8];
            unpack8(lda(&xg[c]), f);
#pragma unroll
            for (int i = 0; i < 8; ++i) ss += f[i] * f[i];
        }
        ss = group_sum<MgAll>(ss, red, tid);
        const float rs = __frsqrt_rn(__fadd_rn(__fdiv_rn(ss, (float)K), g.eps));
        const uint4* wgt = reinterpret_cast<const uint4*>(g.norm_w);
        for (int c = tid; c < K8; c += MG_CONSUMERS) {
            float f[8], w[8];
            unpack8(lda(&xg[c]), f);
            unpack8(wgt[c], w);
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] = rbf(__fmul_rn(f[i], rs)) * w[i];
            uint4 o;
            o.x = pack2(f[0], f[1]); o.y = pack2(f[2], f[3]); o.z = pack2(f[4], f[5]); o.w = pack2(f[6], f[7]);
            if (g.xnorm_out != nullptr && b == 0) reinterpret_cast<uint4*>(g.xnorm_out)[c] = o;
            xs[c] = o;
        }
    } else {
        for (int c = tid; c < K8; c += MG_CONSUMERS) xs[c] = lda(&xg[c]);
    }
    MgAll::sync();
    const long long tp1 = prof ? clock64() : 0;
    if (trace && tid == 0) trace[1] = mg_now();

    const MgGeom geo(g.N, g.epi, g.rshift, b, nb);
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
                if (r < rpw && row0 + 4 * r < ng) resv[r] = lda(&g.res[(size_t)(g0 + row0 + 4 * r)]);
        }
        mbar_wait(&full_bar[slot], ph);
        const unsigned char* st = stages + (size_t)slot * MG_STAGE_BYTES;
        float part[2] = {0.f, 0.f};
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            const int row = row0 + 4 * r;
            if (r < rpw && row < ng)
                part[r] = warp_sum(mg_dot(reinterpret_cast<const uint4*>(st + row * row_bytes), xs, c_lo, c_hi, lane));
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[slot]);
        if (SEGW == 1) {
            if (lane == 0) {
#pragma unroll
                for (int r = 0; r < 2; ++r) {
                    const int row = row0 + 4 * r;
                    if (r < rpw && row < ng) {
                        if (!last_q) acc0[r] = part[r];
                        else mg_finish_row(g, (size_t)(g0 + row), geo.ns ? acc0[r] : part[r], part[r], resv[r]);
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
                    mg_finish_row(g, (size_t)(g0 + row0), s, s2, resv[0]);
                }
                par ^= 1;
            }
        }
    }
    n = n0 + (unsigned)geo.total;
    if (prof && tid == 0) {
        prof[MP_XSTAGE] += (unsigned long long)(tp1 - tp0);
        prof[MP_LOOP] += (unsigned long long)(clock64() - tp1);
    }
}

template <int HD, int GQ>
__global__ void __launch_bounds__(MG_THREADS, 1) frame_mega_kernel(MegaPlan p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[MG_NST];
    __shared__ __align__(8) uint64_t empty_bar[MG_NST];
    __shared__ float red[32];
    __shared__ float pbuf[MG_GROUPS][2][2][4];  // [group][parity][matrix][warp of group]

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
        if ((tid & 31) == 0) mg_producer<HD>(p, stages, full_bar, empty_bar, b, nb);
        return;
    }

    typedef SyncNamed<6, ATTN_THREADS> AttnSync;
    typedef SyncNamed<6, 32 * GQ> FastSync;
    unsigned n = 0, epoch = 0;
    unsigned long long* prof = (b == 0) ? p.prof : nullptr;
    const long long tk0 = prof ? clock64() : 0;
    for (int oi = 0; oi < p.n_ops; ++oi) {
        const MegaOp op = p.ops[oi];
        const long long to0 = prof ? clock64() : 0;
        unsigned long long* tr = p.trace ? p.trace + ((size_t)b * p.n_ops + oi) * 4 : nullptr;
        if (tr && tid == 0) tr[0] = mg_now();
        switch (op.type) {
        case MOP_GEMV:
            mg_gemv(p.gemv + op.idx, stages, xs, full_bar, empty_bar, red, pbuf, n, b, nb, tid, prof, tr);
            break;
        case MOP_EMBED:
            if (b < p.B) embed_body(*p.embed, b, tid, MG_CONSUMERS);
            break;
        case MOP_ATTN:
            if (tid < ATTN_THREADS) {
                const AttnArgs a = p.attn[op.idx];
                const int items = a.KV * a.splits * p.B;
                for (int item = b; item < items; item += nb) {
                    const int split = item % a.splits, hm = item / a.splits;
                    attn_decode_body<HD, GQ, AttnSync, false>(a, split, hm % a.KV, hm / a.KV, tid);
                    AttnSync::sync();
                }
            }
            break;
        case MOP_FATTN:
            if (tid < 32 * GQ) {
                const AttnArgs a = p.attn[op.idx];
                const int items = a.KV * p.B;
                for (int item = b; item < items; item += nb) {
                    fast_attn_body<HD, GQ, FastSync>(a, item % a.KV, item / a.KV, tid);
                    FastSync::sync();
                }
            }
            break;
        case MOP_SAMPLE:
            for (int r = b; r < p.B; r += nb) {
                sample_body<MgAll>(p.samp[op.idx], r, p.B, tid);
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
            __threadfence();
        }
    }
}
