// Causal conv / transposed conv / Linear over MANY rows as one implicit GEMM on the 5th-generation tensor cores:
//
//     out[b][t][n] = epi( sum_j sum_k A[b][t - (taps-1-j)*dil][k] * W[j][n][k] ),   rows t < 0 read as zero
//
//   * the DAC decoder's dense layers (modded_dac.py:521-580 causal conv / transposed conv, :599-620 ResidualUnit,
//     :712-801 DecoderBlock / Decoder; rvq.py:58-191 ConvNeXt / upsample convs; modded_dac.py:97-346 transformer Linears);
//   * the Linears of the Dual-AR prompt pass (llama.py:390-466 with Tq > 1), where W is the state dict's [N][K] as it is
//     (taps = 1).
//
// Operands are 16-bit (bf16 for the LLM, fp16 or bf16 for the codec), K-major, accumulation is fp32 in TMEM:
//   * A tile = 128 time rows x 64 channels, fetched by TMA from a 3-D tensor map {C, T, B} at row coordinate
//     t0 - shift_j: rows before the start of the clip are OUT OF BOUNDS and arrive as zeros, which IS the causal left
//     padding (no halo logic, no masks); rows past T and channels past C are zeros too;
//   * W tile = BN output channels x 64 input channels of tap j from a 3-D map {Cin, N, taps};
//   * every (tap, channel chunk) step is 4 tcgen05.mma (M = 128, N = BN, K = 16) into the same TMEM accumulator, so a
//     7-tap conv is just a 7x longer reduction;
//   * warp roles: TMA producer / MMA issuer (one thread) / 4 epilogue warps; a ring of shared-memory stages; TWO TMEM
//     accumulators so the epilogue of tile i overlaps the MMAs of tile i+1; persistent CTAs walk the tile list;
//   * epilogue (fp32): bias, exact GELU, LayerScale / ConvNeXt gamma, residual (fp32 stream or 16-bit), tanh, SwiGLU over
//     interleaved gate/up columns, and Snake1d of the CONSUMING layer (x + sin^2(alpha x) / (alpha + 1e-9)) applied to the
//     16-bit copy that the next GEMM reads, while the raw value goes to the fp32 residual stream: activations are
//     transformed once, by their producer, and the next layer's operand needs no prologue.
//
// tcv_simt_kernel is the same op with the same epilogue on CUDA cores for shapes TMA cannot address (channel counts that
// are not multiples of 8, as in the reduced test configuration) and the cross-check of the tensor-core kernel in tests.
#pragma once
#include <cuda_fp16.h>

#include "tc_common.cuh"  // tcgen05 / TMA / mbarrier wrappers

#define TCV_BM 128
#define TCV_BK 64
#define TCV_EPI_WARPS 8          // two per TMEM lane quadrant: the quadrant's rows, alternate column blocks
#define TCV_THREADS (64 + 32 * TCV_EPI_WARPS)
#define TCV_STAGES_MAX 8
#define TCV_A_BYTES (TCV_BM * TCV_BK * 2)
#define TCV_POLICY_EVICT_NORMAL 0x1000000000000000ull

enum { TCV_ACT_NONE = 0, TCV_ACT_GELU = 1, TCV_ACT_TANH = 2, TCV_ACT_SWIGLU = 3, TCV_ACT_GATE = 4 };

struct TcConvArgs {
    int B, T, N, Cin, taps, dil;
    int tiles_t, tiles_n, n_tiles, kchunks, nst;
    int ksplit, kper;    // split-K: a tile's channel chunks are cut into `ksplit` ranges of `kper` chunks, one work item each (n_tiles
                         // counts work items); item i = (tile i / ksplit, range i % ksplit) writes its fp32 partial sums to
                         // out32 + range * B*T rows.  For few-tile Linears of a short prompt chunk (set by the caller, 0/1 = off)
    // ---- operands of the SIMT variant (the tensor-core kernel reads them through its tensor maps)
    const void* A16;     // [B][T][Cin] 16-bit
    const void* W16;     // [taps][N][Cin] 16-bit
    // ---- epilogue
    void* out16;         // [B*T][ld16] 16-bit copy (snaked when alpha != nullptr), or nullptr
    void* out16r;        // [B*T][ld16r] raw 16-bit copy (never snaked: the residual stream of the decoder blocks), or nullptr
    float* out32;        // [B*T][ld32] fp32 copy (never snaked), or nullptr
    int ld16, ld16r, ld32;
    const float* bias;   // [bias_mod] or nullptr (transposed conv: bias_mod = Cout, N = stride * Cout)
    const float* gamma;  // [N] or nullptr
    const float* res32;  // [B*T][ldres] fp32 residual, may alias out32, or nullptr
    const void* res16;   // [B*T][ldres] 16-bit residual, may alias out16, or nullptr
    const float* alpha;  // [alpha_mod] Snake of the consumer applied to out16, or nullptr
    int bias_mod, alpha_mod, ldres;
    int act;             // TCV_ACT_*; SWIGLU: columns (2i, 2i+1) = (gate, up) -> out column i, out width N / 2;
                         // GATE: out16 = rnd(rnd(silu(g)) * rnd(v)) with g read from res16 (the w1 GEMM's output): FeedForward of the
                         // bf16 LLM with a rounding after every op like the reference (llama.py:985-986)
    int vec_ok;          // 1: every row pitch / pointer allows the 16-byte vector paths of the epilogue
    int staged_ok;       // caller's switch for the staged epilogue (option "staged_epi")
    int staged;          // 1: the 16-bit-only epilogue through shared memory (set by tcv_launch: see tcv_kernel)
    int kgroup;          // ring slots filled / consumed per barrier round (set by tcv_plan; nst is a multiple of it)
    int round_lin;       // 1: round the reduction to the 16-bit type BEFORE the residual add (the bf16 LLM's rounding points:
                         //    one rounding per Linear, one per add, llama.py:838-843)
};

// ----------------------------------------------------------------------------- 16-bit type policy
template <bool F16> struct Tcv16;
template <> struct Tcv16<true> {
    static __device__ __forceinline__ uint16_t pack(float v) {
        v = fminf(fmaxf(v, -65504.f), 65504.f);  // fp16 operands saturate instead of turning into inf
        return __half_as_ushort(__float2half_rn(v));
    }
    static __device__ __forceinline__ float unpack(uint16_t b) { return __half2float(__ushort_as_half(b)); }
};
template <> struct Tcv16<false> {
    static __device__ __forceinline__ uint16_t pack(float v) { return f_to_bf16_bits(v); }
    static __device__ __forceinline__ float unpack(uint16_t b) { return bf16_bits_to_f(b); }
};
// two values -> one 32-bit word (lo = a), same rounding / saturation as pack(): one two-way convert instead of two + shifts
template <bool F16>
__device__ __forceinline__ uint32_t tcv_pack2(float a, float b) {
    if (F16) {   // one instruction: round to nearest even, +-65504 past the range (F2FP.SATFINITE.F16.F32.PACK_AB)
        uint32_t r;
        asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
        return r;
    }
    const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&h);
}

__device__ __forceinline__ float tcv_snake(float x, float a) {
    const float s = __sinf(a * x);  // MUFU sine: abs error ~1e-7 on the O(1) arguments here, far below the 16-bit rounding
    return fmaf(__fdividef(1.f, a + 1e-9f) * s, s, x);
}
__device__ __forceinline__ float tcv_gelu(float v) { return 0.5f * v * (1.f + erff(v * 0.70710678118654752f)); }

// ----------------------------------------------------------------------------- epilogue of NC consecutive columns of one row
// Per-column parameters of the block (index 0 = column n0), neutral where absent (bias 0, gamma 1): shared memory in the
// tensor-core kernel (staged once per column tile, read as broadcasts), a few registers in the CUDA-core twin.
struct TcvParams { const float *bias, *gamma, *alpha, *ainv; };

// v[i] = reduction of column n0 + i of row `row` (global row index b*T + t); nv = valid columns (NC except at a ragged edge).
// Everything is unrolled with compile-time indices (a run-time loop over v[] would put it in local memory).
template <bool F16, int NC>
__device__ __forceinline__ void tcv_epilogue(const TcConvArgs& a, const TcvParams& p, size_t row, int n0, float (&v)[NC], int nv) {
    typedef Tcv16<F16> H;
    const bool vec = (nv == NC) && a.vec_ok;
    if (a.round_lin) {
#pragma unroll
        for (int i = 0; i < NC; ++i) v[i] = H::unpack(H::pack(v[i]));
    }
#pragma unroll
    for (int i = 0; i < NC; ++i) v[i] += p.bias[i];
    if (a.act == TCV_ACT_GELU) {
#pragma unroll
        for (int i = 0; i < NC; ++i) v[i] = tcv_gelu(v[i]);
    }
    if (a.act == TCV_ACT_SWIGLU) {  // (gate, up) pairs -> NC / 2 outputs at columns n0 / 2 ..
        uint16_t* dst = reinterpret_cast<uint16_t*>(a.out16) + row * a.ld16 + (n0 >> 1);
        uint32_t w[NC / 4];
#pragma unroll
        for (int q = 0; q < NC / 4; ++q) {
            const float g0 = v[4 * q], u0 = v[4 * q + 1], g1 = v[4 * q + 2], u1 = v[4 * q + 3];
            w[q] = (uint32_t)H::pack(__fdividef(g0, 1.f + __expf(-g0)) * u0) | ((uint32_t)H::pack(__fdividef(g1, 1.f + __expf(-g1)) * u1) << 16);
        }
        if (vec) {
#pragma unroll
            for (int q = 0; q < NC / 16; ++q) *reinterpret_cast<uint4*>(dst + 8 * q) = make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
        } else {
#pragma unroll
            for (int i = 0; i < NC / 2; ++i)
                if (2 * i + 1 < nv) dst[i] = (uint16_t)((i & 1) ? (w[i >> 1] >> 16) : (w[i >> 1] & 0xffffu));
        }
        return;
    }
#pragma unroll
    for (int i = 0; i < NC; ++i) v[i] *= p.gamma[i];
    if (a.res32) {
        const float* rp = a.res32 + row * a.ldres + n0;
        if (vec) {
#pragma unroll
            for (int q = 0; q < NC / 4; ++q) { const float4 r4 = *reinterpret_cast<const float4*>(rp + 4 * q); v[4 * q] += r4.x; v[4 * q + 1] += r4.y; v[4 * q + 2] += r4.z; v[4 * q + 3] += r4.w; }
        } else {
#pragma unroll
            for (int i = 0; i < NC; ++i)
                if (i < nv) v[i] += rp[i];
        }
    } else if (a.res16 && a.act == TCV_ACT_GATE) {
        const uint16_t* rp = reinterpret_cast<const uint16_t*>(a.res16) + row * a.ldres + n0;
        float g[NC];
        if (vec) {
#pragma unroll
            for (int q = 0; q < NC / 8; ++q) {
                const uint4 r = *reinterpret_cast<const uint4*>(rp + 8 * q);
                g[8 * q] = H::unpack((uint16_t)(r.x & 0xffffu)); g[8 * q + 1] = H::unpack((uint16_t)(r.x >> 16));
                g[8 * q + 2] = H::unpack((uint16_t)(r.y & 0xffffu)); g[8 * q + 3] = H::unpack((uint16_t)(r.y >> 16));
                g[8 * q + 4] = H::unpack((uint16_t)(r.z & 0xffffu)); g[8 * q + 5] = H::unpack((uint16_t)(r.z >> 16));
                g[8 * q + 6] = H::unpack((uint16_t)(r.w & 0xffffu)); g[8 * q + 7] = H::unpack((uint16_t)(r.w >> 16));
            }
        } else {
#pragma unroll
            for (int i = 0; i < NC; ++i) g[i] = (i < nv) ? H::unpack(rp[i]) : 0.f;
        }
#pragma unroll
        for (int i = 0; i < NC; ++i) {
            const float sl = H::unpack(H::pack(__fdiv_rn(g[i], 1.f + expf(-g[i]))));  // silu rounded to the 16-bit type
            v[i] = sl * H::unpack(H::pack(v[i]));                                      // times the rounded up projection
        }
    } else if (a.res16) {
        const uint16_t* rp = reinterpret_cast<const uint16_t*>(a.res16) + row * a.ldres + n0;
        if (vec) {
#pragma unroll
            for (int q = 0; q < NC / 8; ++q) {
                const uint4 r = *reinterpret_cast<const uint4*>(rp + 8 * q);
                v[8 * q] += H::unpack((uint16_t)(r.x & 0xffffu)); v[8 * q + 1] += H::unpack((uint16_t)(r.x >> 16));
                v[8 * q + 2] += H::unpack((uint16_t)(r.y & 0xffffu)); v[8 * q + 3] += H::unpack((uint16_t)(r.y >> 16));
                v[8 * q + 4] += H::unpack((uint16_t)(r.z & 0xffffu)); v[8 * q + 5] += H::unpack((uint16_t)(r.z >> 16));
                v[8 * q + 6] += H::unpack((uint16_t)(r.w & 0xffffu)); v[8 * q + 7] += H::unpack((uint16_t)(r.w >> 16));
            }
        } else {
#pragma unroll
            for (int i = 0; i < NC; ++i)
                if (i < nv) v[i] += H::unpack(rp[i]);
        }
    }
    if (a.act == TCV_ACT_TANH) {
#pragma unroll
        for (int i = 0; i < NC; ++i) v[i] = tanhf(v[i]);
    }
    if (a.out32) {
        float* op = a.out32 + row * a.ld32 + n0;
        if (vec) {
#pragma unroll
            for (int q = 0; q < NC / 4; ++q) *reinterpret_cast<float4*>(op + 4 * q) = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
        } else {
#pragma unroll
            for (int i = 0; i < NC; ++i)
                if (i < nv) op[i] = v[i];
        }
    }
#pragma unroll
    for (int pass = 0; pass < 2; ++pass) {  // 0: raw 16-bit copy (residual stream), 1: the consumer's operand (Snake applied)
        uint16_t* base = reinterpret_cast<uint16_t*>(pass == 0 ? a.out16r : a.out16);
        if (base == nullptr) continue;
        if (pass == 1 && a.alpha) {
#pragma unroll
            for (int i = 0; i < NC; ++i) { const float sn = __sinf(p.alpha[i] * v[i]); v[i] = fmaf(p.ainv[i] * sn, sn, v[i]); }
        }
        uint16_t* op = base + row * (pass == 0 ? a.ld16r : a.ld16) + n0;
        if (vec) {
#pragma unroll
            for (int q = 0; q < NC / 8; ++q)
                *reinterpret_cast<uint4*>(op + 8 * q) =
                    make_uint4((uint32_t)H::pack(v[8 * q]) | ((uint32_t)H::pack(v[8 * q + 1]) << 16), (uint32_t)H::pack(v[8 * q + 2]) | ((uint32_t)H::pack(v[8 * q + 3]) << 16),
                               (uint32_t)H::pack(v[8 * q + 4]) | ((uint32_t)H::pack(v[8 * q + 5]) << 16), (uint32_t)H::pack(v[8 * q + 6]) | ((uint32_t)H::pack(v[8 * q + 7]) << 16));
        } else {
#pragma unroll
            for (int i = 0; i < NC; ++i)
                if (i < nv) op[i] = H::pack(v[i]);
        }
    }
}
// per-column parameter values of column n (neutral when the array is absent or n is past the end)
__device__ __forceinline__ void tcv_param(const TcConvArgs& a, int n, float& bias, float& gamma, float& alpha, float& ainv) {
    const bool in = n < a.N;
    bias = (a.bias && in) ? __ldg(a.bias + (n % a.bias_mod)) : 0.f;
    gamma = (a.gamma && in) ? __ldg(a.gamma + n) : 1.f;
    alpha = (a.alpha && in) ? __ldg(a.alpha + (n % a.alpha_mod)) : 0.f;
    ainv = __fdividef(1.f, alpha + 1e-9f);
}

// ----------------------------------------------------------------------------- the tensor-core kernel
__host__ __device__ constexpr uint32_t tcv_stage_bytes(int bn) { return (uint32_t)(TCV_A_BYTES + bn * TCV_BK * 2); }
__host__ __device__ constexpr int tcv_tmem_cols(int bn) { return 2 * bn <= 32 ? 32 : 2 * bn <= 64 ? 64 : 2 * bn <= 128 ? 128 : 2 * bn <= 256 ? 256 : 512; }
// Staged epilogue (decoder convs: 16-bit residual in, 16-bit outputs only): a panel of PW columns of the tile at a time.  The
// residual panel is copied global -> shared with cp.async one panel ahead (coalesced 16-byte pieces, no registers, its
// latency off the chain), every thread adds its row from there, and the packed outputs go through a staging panel so that the
// global stores are whole contiguous runs instead of 32 rows x 16 bytes per warp instruction.  (One bulk async store per row
// from two staging panels was measured too: the same on the 1-tap layers, 24 % slower on the transposed convs - 192-byte bulk
// copies are too small for the copy engine.)  Rows carry 16 bytes of padding
// (an odd number of 16-byte units: a thread per row and a thread per piece are both conflict-free).
__host__ __device__ constexpr int tcv_panel_cols(int bn) { return bn % 96 == 0 ? 96 : (bn % 64 == 0 ? 64 : 32); }
__host__ __device__ constexpr uint32_t tcv_panel_row_bytes(int bn) { return (uint32_t)tcv_panel_cols(bn) * 2 + 16; }
// one staging panel, plus two residual panels when the layer has a 16-bit residual (a layer without one keeps that shared memory
// for its stage ring: at N = 96 a stage is only ~190 tensor-pipe cycles of work, the ring is what covers the TMA latency)
__host__ __device__ constexpr uint32_t tcv_epi_bytes(int bn, bool has_res) { return (has_res ? 3u : 1u) * TCV_BM * tcv_panel_row_bytes(bn); }
inline size_t tcv_smem_bytes(int bn, int nst, bool staged, bool has_res) { return (size_t)nst * tcv_stage_bytes(bn) + (staged ? tcv_epi_bytes(bn, has_res) : 0) + 1024; }

__device__ __forceinline__ void tcv_epi_bar() { asm volatile("bar.sync 1, %0;" ::"n"(32 * TCV_EPI_WARPS) : "memory"); }
__device__ __forceinline__ void tcv_tma_load_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4, %5}], [%2], %6;"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "l"(policy) : "memory");
}
// instruction descriptor: D fp32, A/B both bf16 (1) or both fp16 (0), K-major, M = 128, N = bn
__host__ __device__ constexpr uint32_t tcv_idesc(int bn, bool f16) {
    return (1u << 4) | ((f16 ? 0u : 1u) << 7) | ((f16 ? 0u : 1u) << 10) | ((uint32_t)(bn >> 3) << 17) | ((uint32_t)(TCV_BM >> 4) << 24);
}

__device__ __forceinline__ void tcv_cp_async16(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void tcv_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tcv_cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

template <int BN, bool F16>
__device__ __forceinline__ void tcv_staged_epilogue(const TcConvArgs& a, uint32_t tmem, uint8_t* epi, uint64_t* tfull_bar, uint64_t* tempty_bar,
                                                    float* ep_bias, float* ep_gamma, float* ep_alpha, float* ep_ainv, int et, int lg,
                                                    int half, int lane) {
    typedef Tcv16<F16> H;
    constexpr int PW = tcv_panel_cols(BN), NP = BN / PW;     // panel width, panels per tile
    constexpr int HW = PW / 2;                               // columns of one warp of a lane quadrant
    constexpr int NCB = (HW % 32 == 0) ? 32 : 16, NB = HW / NCB;
    constexpr int CPR = PW / 8;                              // 16-byte pieces per panel row
    constexpr uint32_t RB = tcv_panel_row_bytes(BN);
    constexpr int NT = 32 * TCV_EPI_WARPS;
    uint8_t* const sbuf = epi + (a.res16 != nullptr ? 2 : 0) * (size_t)TCV_BM * RB;   // staging panel behind the residual panels, if any
    const uint32_t tlane = (uint32_t)(lg * 32) << 16;
    const int myrow = lg * 32 + lane;                        // row of the tile this thread reads from TMEM
    const uint16_t* res = reinterpret_cast<const uint16_t*>(a.res16);
    const int row0 = et / CPR, c0 = et % CPR;                // this thread's first 16-byte piece of a panel (then + NT pieces)

    // The cooperative copies (residual panel in, staging panel out) move 16-byte pieces, NT at a time; which pieces a thread
    // takes is the same for every panel, so their (row, offset) pairs are worked out once: the loops below were a third of
    // the kernel's instructions when they derived them per piece (ncu source page, 1-tap layers: issue-bound epilogue).
    constexpr int ITER = TCV_BM * CPR / NT;
    static_assert(ITER * NT == TCV_BM * CPR, "whole rounds of pieces");
    int prow[ITER], pcol8[ITER], soff[ITER];
    {
        int row = row0, c = c0;
#pragma unroll
        for (int k = 0; k < ITER; ++k) {
            prow[k] = row; pcol8[k] = c * 8; soff[k] = row * (int)RB + c * 16;
            row += NT / CPR; c += NT % CPR;
            if (c >= CPR) { c -= CPR; ++row; }
        }
    }
    // residual panel `p` of tile `tile` -> rbuf[slot]: coalesced 16-byte pieces, rows past the end of the clip skipped
    auto request = [&](int tile, int p, int slot) {
        if (res == nullptr || tile >= a.n_tiles) return;
        const int tn = tile % a.tiles_n, r = tile / a.tiles_n, tt = r % a.tiles_t, b = r / a.tiles_t;
        const uint16_t* gb = res + ((size_t)b * a.T + (size_t)tt * TCV_BM) * a.ldres + tn * BN + p * PW;
        uint8_t* sb = epi + (size_t)slot * TCV_BM * RB;
        const int left = a.T - tt * TCV_BM;
#pragma unroll
        for (int k = 0; k < ITER; ++k)
            if (prow[k] < left) tcv_cp_async16(sb + soff[k], gb + (size_t)prow[k] * a.ldres + pcol8[k]);
    };
    // staging panel -> global, whole rows of the panel as contiguous runs
    auto copy_out = [&](uint16_t* base, int ld, int b, int tt, int col0) {
        uint16_t* gb = base + ((size_t)b * a.T + (size_t)tt * TCV_BM) * ld + col0;
        const int left = a.T - tt * TCV_BM;
#pragma unroll
        for (int k = 0; k < ITER; ++k)
            if (prow[k] < left)
                *reinterpret_cast<uint4*>(gb + (size_t)prow[k] * ld + pcol8[k]) = *reinterpret_cast<const uint4*>(sbuf + soff[k]);
    };
    uint32_t iseg = 0, g = 0;   // tiles / panels done by this CTA
    int tn_loaded = -1;
    request(blockIdx.x, 0, 0);
    tcv_cp_async_commit();
    for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x, ++iseg) {
        const int tn = tile % a.tiles_n, r = tile / a.tiles_n, tt = r % a.tiles_t, b = r / a.tiles_t;
        const int n0 = tn * BN;
        if (tn != tn_loaded) {  // uniform; everyone is past the previous tile (barriers below)
            for (int i = et; i < BN; i += NT) tcv_param(a, n0 + i, ep_bias[i], ep_gamma[i], ep_alpha[i], ep_ainv[i]);
            tn_loaded = tn;
        }
        const uint32_t buf = iseg & 1u;
        mbar_wait(&tfull_bar[buf], (iseg >> 1) & 1u);
        tc_fence_after();
        const uint32_t tacc = tmem + tlane + buf * BN;
#pragma unroll 1
        for (int p = 0; p < NP; ++p, ++g) {
            const int slot = (int)(g & 1u);
            tcv_cp_async_wait_all();         // my pieces of this panel's residual have landed
            tcv_epi_bar();                   // ... everybody's have; the staging panel and the other residual slot are free
            if (p + 1 < NP) request(tile, p + 1, slot ^ 1);
            else request(tile + (int)gridDim.x, 0, slot ^ 1);
            tcv_cp_async_commit();
            const int pc = p * PW + half * HW;             // first column (inside the tile) of this thread's share
            float v[NB][NCB];
            {   // all of this thread's accumulator columns requested before ONE wait
                uint32_t (&u)[NB][NCB] = reinterpret_cast<uint32_t(&)[NB][NCB]>(v);
#pragma unroll
                for (int k = 0; k < NB; ++k) {
                    if (NCB == 32) tc_ld32_nowait(tacc + (uint32_t)(pc + k * NCB), reinterpret_cast<uint32_t(&)[32]>(u[k]));
                    else tc_ld16_nowait(tacc + (uint32_t)(pc + k * NCB), reinterpret_cast<uint32_t(&)[16]>(u[k]));
                }
                tc_ld_wait();
            }
            if (p == NP - 1) {               // the accumulator is in registers: hand it back before the stores
                tc_fence_before();
                tcv_epi_bar();
                if (et == 0) mbar_arrive(&tempty_bar[buf]);
            }
            const uint8_t* rrow = epi + (size_t)slot * TCV_BM * RB + (size_t)myrow * RB + (size_t)(half * HW) * 2;
#pragma unroll
            for (int k = 0; k < NB; ++k) {
#pragma unroll
                for (int q = 0; q < NCB / 4; ++q) {   // per-column parameters as 16-byte broadcast reads
                    const float4 b4 = *reinterpret_cast<const float4*>(ep_bias + pc + k * NCB + 4 * q);
                    v[k][4 * q] += b4.x; v[k][4 * q + 1] += b4.y; v[k][4 * q + 2] += b4.z; v[k][4 * q + 3] += b4.w;
                }
                if (res) {
#pragma unroll
                    for (int q = 0; q < NCB / 8; ++q) {
                        const uint4 rr = *reinterpret_cast<const uint4*>(rrow + (k * NCB + 8 * q) * 2);
                        const uint32_t u[4] = {rr.x, rr.y, rr.z, rr.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            v[k][8 * q + 2 * e] += H::unpack((uint16_t)(u[e] & 0xffffu));
                            v[k][8 * q + 2 * e + 1] += H::unpack((uint16_t)(u[e] >> 16));
                        }
                    }
                }
            }
            uint8_t* srow = sbuf + (size_t)myrow * RB + (size_t)(half * HW) * 2;
#pragma unroll
            for (int pass = 0; pass < 2; ++pass) {   // 0: raw copy (the residual stream), 1: the consumer's operand (Snake applied)
                uint16_t* base = reinterpret_cast<uint16_t*>(pass == 0 ? a.out16r : a.out16);
                if (base == nullptr) continue;       // uniform
                if (pass == 1 && a.alpha) {
#pragma unroll
                    for (int k = 0; k < NB; ++k)
#pragma unroll
                        for (int q = 0; q < NCB / 4; ++q) {
                            const float4 al = *reinterpret_cast<const float4*>(ep_alpha + pc + k * NCB + 4 * q);
                            const float4 ai = *reinterpret_cast<const float4*>(ep_ainv + pc + k * NCB + 4 * q);
                            const float av[4] = {al.x, al.y, al.z, al.w}, iv[4] = {ai.x, ai.y, ai.z, ai.w};
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                const float sn = __sinf(av[e] * v[k][4 * q + e]);
                                v[k][4 * q + e] = fmaf(iv[e] * sn, sn, v[k][4 * q + e]);
                            }
                        }
                }
#pragma unroll
                for (int k = 0; k < NB; ++k)
#pragma unroll
                    for (int q = 0; q < NCB / 8; ++q)
                        *reinterpret_cast<uint4*>(srow + (k * NCB + 8 * q) * 2) =
                            make_uint4(tcv_pack2<F16>(v[k][8 * q], v[k][8 * q + 1]), tcv_pack2<F16>(v[k][8 * q + 2], v[k][8 * q + 3]),
                                       tcv_pack2<F16>(v[k][8 * q + 4], v[k][8 * q + 5]), tcv_pack2<F16>(v[k][8 * q + 6], v[k][8 * q + 7]));
                tcv_epi_bar();
                copy_out(base, pass == 0 ? a.ld16r : a.ld16, b, tt, n0 + p * PW);
                tcv_epi_bar();               // the staging panel is free again
            }
        }
    }
    tcv_cp_async_wait_all();
}

template <int BN, bool F16>
__global__ void __launch_bounds__(TCV_THREADS, 1)
tcv_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, TcConvArgs a) {
    constexpr uint32_t STAGE = tcv_stage_bytes(BN);
    constexpr int TCOLS = tcv_tmem_cols(BN);
    extern __shared__ uint8_t tcv_smem_raw[];
    __shared__ uint64_t full_bar[TCV_STAGES_MAX], empty_bar[TCV_STAGES_MAX], tfull_bar[2], tempty_bar[2];
    __shared__ uint32_t tmem_base_s;
    __shared__ __align__(16) float ep_bias[BN], ep_gamma[BN], ep_alpha[BN], ep_ainv[BN];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t sbase = (smem_u32(tcv_smem_raw) + 1023u) & ~1023u;
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
    const int ks = a.ksplit > 1 ? a.ksplit : 1;

    if (warp == 0) {
        // ===== TMA producer
        if (lane == 0) {
            rg_prefetch_map(&tmA);
            rg_prefetch_map(&tmW);
            uint32_t s = 0, ph = 1;   // ring slot and the parity its empty barrier is waited with (no divisions on this path)
            for (int item = blockIdx.x; item < a.n_tiles; item += gridDim.x) {
                const int sp = item % ks, tile = item / ks;
                const int tn = tile % a.tiles_n, r = tile / a.tiles_n, tt = r % a.tiles_t, b = r / a.tiles_t;
                const int t0 = tt * TCV_BM, n0 = tn * BN;
                const int kc_lo = sp * a.kper, kc_hi = min(a.kchunks, kc_lo + a.kper);
                if (a.kgroup == 1) {
                for (int j = 0; j < a.taps; ++j) {
                    const int trow = t0 - (a.taps - 1 - j) * a.dil;  // negative rows: out of bounds = zeros = causal padding
                    for (int kc = kc_lo; kc < kc_hi; ++kc) {
                        mbar_wait(&empty_bar[s], ph);
                        const uint32_t st = sbase + s * STAGE, bar = smem_u32(&full_bar[s]);
                        mbar_expect_tx(&full_bar[s], STAGE);
                        tcv_tma_load_3d(st, &tmA, bar, kc * TCV_BK, trow, b, TCV_POLICY_EVICT_NORMAL);  // re-read by the other taps / column tiles: keep in L2
                        tcv_tma_load_3d(st + TCV_A_BYTES, &tmW, bar, kc * TCV_BK, n0, j, RG_POLICY_EVICT_LAST);
                        if (++s == (uint32_t)a.nst) { s = 0; ph ^= 1u; }
                    }
                }
                } else {
                // reduction steps (tap j, channel chunk kc) in groups of a.kgroup ring slots behind ONE barrier pair: what a
                // narrow 7-tap layer costs is this thread's (and the issuer's) latency per barrier round, not bytes or MMAs
                // (measured by switching loads / MMAs / the epilogue off: DESIGN.md 5.4).  The per-slot loop above is kept
                // as it was for kgroup = 1: anything added to it shows 1:1 in the plain GEMMs' time.
                const int nk = a.taps * (kc_hi - kc_lo);
                int j = 0, kc = kc_lo;
                for (int i = 0; i < nk; i += a.kgroup) {
                    const int cnt = min(a.kgroup, nk - i);
                    mbar_wait(&empty_bar[s], ph);
                    const uint32_t bar = smem_u32(&full_bar[s]);
                    mbar_expect_tx(&full_bar[s], (uint32_t)cnt * STAGE);
                    for (int c = 0; c < cnt; ++c) {
                        const uint32_t st = sbase + (s + c) * STAGE;
                        const int trow = t0 - (a.taps - 1 - j) * a.dil;  // negative rows: out of bounds = zeros = causal padding
                        tcv_tma_load_3d(st, &tmA, bar, kc * TCV_BK, trow, b, TCV_POLICY_EVICT_NORMAL);  // re-read by the other taps / column tiles: keep in L2
                        tcv_tma_load_3d(st + TCV_A_BYTES, &tmW, bar, kc * TCV_BK, n0, j, RG_POLICY_EVICT_LAST);
                        if (++kc == kc_hi) { kc = kc_lo; ++j; }
                    }
                    s += (uint32_t)a.kgroup;
                    if (s >= (uint32_t)a.nst) { s = 0; ph ^= 1u; }
                }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer
        constexpr uint32_t IDESC = tcv_idesc(BN, F16);
        // One thread issues every MMA of the CTA: at N = 96 a 16-channel step is ~48 tensor-pipe cycles, so anything this
        // loop spends per stage beyond that (integer divisions, descriptor rebuilds) shows up 1:1 in the layer time.
        uint32_t iseg = 0, s = 0, ph = 0;
        uint64_t adesc = tc_smem_desc(sbase), bdesc = tc_smem_desc(sbase + TCV_A_BYTES);
        const uint64_t dstep = (uint64_t)(STAGE >> 4);     // descriptor address field counts 16-byte units
        for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x, ++iseg) {
            const int kc_lo = (tile % ks) * a.kper;
            const int nk = a.taps * (min(a.kchunks, kc_lo + a.kper) - kc_lo);   // reduction steps of this work item
            const uint32_t buf = iseg & 1u;
            mbar_wait(&tempty_bar[buf], ((iseg >> 1) & 1u) ^ 1u);  // the epilogue has drained this accumulator
            tc_fence_after();
            const uint32_t d = tmem + buf * BN;
            if (a.kgroup == 1) {
            for (int i = 0; i < nk; ++i) {
                mbar_wait(&full_bar[s], ph);
                tc_fence_after();
                if (lane == 0) {
                    // (skipping the zero-filled 16-channel steps of a last chunk past Cin - the 96-channel block - was measured:
                    // no gain, the layer is not bound by the number of MMAs)
#pragma unroll
                    for (int k = 0; k < TCV_BK / 16; ++k)
                        tc_mma_bf16(d, adesc + 2 * k, bdesc + 2 * k, IDESC, (i > 0 || k > 0) ? 1u : 0u);
                    tc_commit(&empty_bar[s]);
                }
                __syncwarp();
                if (++s == (uint32_t)a.nst) { s = 0; ph ^= 1u; adesc = tc_smem_desc(sbase); bdesc = tc_smem_desc(sbase + TCV_A_BYTES); }
                else { adesc += dstep; bdesc += dstep; }
            }
            } else {
            for (int i = 0; i < nk; i += a.kgroup) {
                mbar_wait(&full_bar[s], ph);
                tc_fence_after();
                if (lane == 0) {
                    // (skipping the zero-filled 16-channel steps of a last chunk past Cin - the 96-channel block - was measured:
                    // no gain, the layer is not bound by the number of MMAs)
#pragma unroll
                    for (int k = 0; k < TCV_BK / 16; ++k)
                        tc_mma_bf16(d, adesc + 2 * k, bdesc + 2 * k, IDESC, (i > 0 || k > 0) ? 1u : 0u);
                    const int cnt = min(a.kgroup, nk - i);
                    uint64_t ad = adesc, bd = bdesc;
                    for (int c = 1; c < cnt; ++c) {      // the group's other ring slots
                        ad += dstep; bd += dstep;
#pragma unroll
                        for (int k = 0; k < TCV_BK / 16; ++k)
                            tc_mma_bf16(d, ad + 2 * k, bd + 2 * k, IDESC, 1u);
                    }
                    tc_commit(&empty_bar[s]);
                }
                __syncwarp();
                s += (uint32_t)a.kgroup;
                if (s >= (uint32_t)a.nst) { s = 0; ph ^= 1u; adesc = tc_smem_desc(sbase); bdesc = tc_smem_desc(sbase + TCV_A_BYTES); }
                else { adesc += dstep * (uint64_t)a.kgroup; bdesc += dstep * (uint64_t)a.kgroup; }
            }
            }
            if (lane == 0) tc_commit(&tfull_bar[buf]);
            __syncwarp();
        }
    } else {
        // ===== epilogue: warp w may touch TMEM lanes 32*(w%4) .. +31 (one thread = one time row of the tile); the two warps
        // of a lane quadrant take alternate column blocks.  Per-column parameters come from shared memory (broadcast reads):
        // a global load per block on this dependent chain cost more than the tile's MMAs (ncu: long-scoreboard stalls).
        constexpr int NCB = (BN % 64 == 0) ? 32 : 16, NBLK = BN / NCB;
        const int et = tid - 64, lg = warp & 3, half = (warp - 2) >> 2;
        const uint32_t tlane = (uint32_t)(lg * 32) << 16;
        uint32_t iseg = 0;
        int tn_loaded = -1;
        if (a.staged) {
            tcv_staged_epilogue<BN, F16>(a, tmem, tcv_smem_raw + ((sbase - smem_u32(tcv_smem_raw)) + (size_t)a.nst * STAGE), tfull_bar, tempty_bar,
                                         ep_bias, ep_gamma, ep_alpha, ep_ainv, et, lg, half, lane);
        } else
        for (int item = blockIdx.x; item < a.n_tiles; item += gridDim.x, ++iseg) {
            const int sp = item % ks, tile = item / ks;
            const int tn = tile % a.tiles_n, r = tile / a.tiles_n, tt = r % a.tiles_t, b = r / a.tiles_t;
            const int t = tt * TCV_BM + lg * 32 + lane, n0 = tn * BN;
            if (tn != tn_loaded) {  // uniform over the epilogue threads; everyone is past the previous tile (barrier below)
                for (int i = et; i < BN; i += 32 * TCV_EPI_WARPS) tcv_param(a, n0 + i, ep_bias[i], ep_gamma[i], ep_alpha[i], ep_ainv[i]);
                tn_loaded = tn;
                tcv_epi_bar();
            }
            const uint32_t buf = iseg & 1u;
            mbar_wait(&tfull_bar[buf], (iseg >> 1) & 1u);
            tc_fence_after();
            const uint32_t tacc = tmem + tlane + buf * BN;
            const size_t row = (size_t)b * a.T + t + (size_t)sp * a.B * a.T;   // split-K: partial sums of range sp go to their own rows
#pragma unroll 1
            for (int k = half; k < NBLK; k += 2) {
                const int nc = n0 + k * NCB;
                if (nc >= a.N) break;  // uniform
                float v[NCB];
                if (NCB == 32) tc_ld32(tacc + (uint32_t)(k * NCB), reinterpret_cast<float(&)[32]>(v));
                else tc_ld16(tacc + (uint32_t)(k * NCB), reinterpret_cast<float(&)[16]>(v));
                const TcvParams ps{ep_bias + k * NCB, ep_gamma + k * NCB, ep_alpha + k * NCB, ep_ainv + k * NCB};
                if (t < a.T) tcv_epilogue<F16, NCB>(a, ps, row, nc, v, min(NCB, a.N - nc));
            }
            tc_fence_before();
            tcv_epi_bar();
            if (et == 0) mbar_arrive(&tempty_bar[buf]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TCOLS) : "memory");
    }
}

// ----------------------------------------------------------------------------- the same op on CUDA cores
// one thread = 16 consecutive output columns of one row; for channel counts TMA cannot address and as the cross-check
template <bool F16>
__global__ void __launch_bounds__(128) tcv_simt_kernel(TcConvArgs a) {
    typedef Tcv16<F16> H;
    const int chunks = (a.N + 15) / 16;
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)a.B * a.T * chunks) return;
    const int ch = (int)(idx % chunks);
    const long long bt = idx / chunks;
    const int t = (int)(bt % a.T), b = (int)(bt / a.T);
    const int n0 = ch * 16, nv = min(16, a.N - n0);
    const uint16_t* A = reinterpret_cast<const uint16_t*>(a.A16);
    const uint16_t* W = reinterpret_cast<const uint16_t*>(a.W16);
    float v[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = 0.f;
    for (int j = 0; j < a.taps; ++j) {
        const int tr = t - (a.taps - 1 - j) * a.dil;
        if (tr < 0) continue;
        const uint16_t* ar = A + ((size_t)b * a.T + tr) * a.Cin;
        const uint16_t* wj = W + ((size_t)j * a.N + n0) * a.Cin;
        for (int k = 0; k < a.Cin; ++k) {
            const float x = H::unpack(ar[k]);
#pragma unroll
            for (int i = 0; i < 16; ++i)
                if (i < nv) v[i] = fmaf(x, H::unpack(wj[(size_t)i * a.Cin + k]), v[i]);
        }
    }
    float pb[16], pg[16], pa[16], pi[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) tcv_param(a, n0 + i, pb[i], pg[i], pa[i], pi[i]);
    const TcvParams ps{pb, pg, pa, pi};
    tcv_epilogue<F16, 16>(a, ps, (size_t)b * a.T + t, n0, v, nv);
}

// ----------------------------------------------------------------------------- host side
// 3-D map over a 16-bit tensor [d2][d1][d0] (d0 contiguous): box = 64 x box_rows x 1, 128-byte swizzle, zero fill out of bounds
inline bool tcv_make_map(CUtensorMap* m, const void* ptr, bool f16, uint64_t d0, uint64_t d1, uint64_t d2, uint32_t box_rows) {
    rg_encode_fn enc = rg_get_encode();
    if (!enc) return false;
    const cuuint64_t gdim[3] = {d0, d1, d2};
    const cuuint64_t gstr[2] = {d0 * 2, d0 * d1 * 2};
    const cuuint32_t box[3] = {TCV_BK, box_rows, 1};
    const cuuint32_t est[3] = {1, 1, 1};
    return enc(m, f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), gdim, gstr, box, est,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
// column-tile width: the widest of the UMMA widths that divides N (no wasted columns), else 256 / the next multiple of 16
inline int tcv_pick_bn(int N) {
    const int cands[6] = {256, 192, 128, 96, 64, 32};
    if (N >= 96)
        for (int i = 0; i < 4; ++i)
            if (N % cands[i] == 0) return cands[i];
    if (N > 256) return 256;
    for (int i = 5; i >= 0; --i)
        if (N <= cands[i]) return cands[i];
    return 256;
}
// shapes the tensor-core kernel can address: 16-byte row pitches for TMA; anything else goes to tcv_simt_kernel
inline bool tcv_tc_ok(int Cin, int N) { return Cin % 8 == 0 && Cin >= 8 && N >= 1; }

typedef void (*tcv_fn)(const CUtensorMap, const CUtensorMap, TcConvArgs);
template <bool F16>
inline tcv_fn tcv_pick(int bn) {
    switch (bn) {
    case 256: return tcv_kernel<256, F16>;
    case 192: return tcv_kernel<192, F16>;
    case 128: return tcv_kernel<128, F16>;
    case 96: return tcv_kernel<96, F16>;
    case 64: return tcv_kernel<64, F16>;
    default: return tcv_kernel<32, F16>;
    }
}
inline int tcv_stages_for(int bn, bool staged, bool has_res) {
    const int fit = (int)((225 * 1024 - 1024 - 16 * bn - 256 - (staged ? (int)tcv_epi_bytes(bn, has_res) : 0)) / tcv_stage_bytes(bn));  // static: 4 parameter rows + barriers
    return fit < TCV_STAGES_MAX ? fit : TCV_STAGES_MAX;
}
inline cudaError_t tcv_set_attrs() {
    FsbOnce once(4);
    if (!once.first) return cudaSuccess;
    const int bns[6] = {256, 192, 128, 96, 64, 32};
    for (int f = 0; f < 2; ++f)
        for (int i = 0; i < 6; ++i) {
            const void* k = f ? (const void*)tcv_pick<true>(bns[i]) : (const void*)tcv_pick<false>(bns[i]);
            size_t need = tcv_smem_bytes(bns[i], tcv_stages_for(bns[i], false, false), false, false);
            for (int r = 0; r < 2; ++r) {
                const size_t st = tcv_smem_bytes(bns[i], tcv_stages_for(bns[i], true, r != 0), true, r != 0);
                need = st > need ? st : need;
            }
            cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)need);
            if (e != cudaSuccess) return e;
        }
    return cudaSuccess;
}
// fills the tiling fields of `a` for column-tile width bn (a.staged decides how much shared memory the stage ring gets)
inline void tcv_plan(TcConvArgs& a, int bn) {
    a.tiles_t = (a.T + TCV_BM - 1) / TCV_BM;
    a.tiles_n = (a.N + bn - 1) / bn;
    a.kchunks = (a.Cin + TCV_BK - 1) / TCV_BK;
    if (a.ksplit > 1) {
        a.kper = (a.kchunks + a.ksplit - 1) / a.ksplit;
        a.ksplit = (a.kchunks + a.kper - 1) / a.kper;   // no empty ranges
    } else {
        a.ksplit = 1;
        a.kper = a.kchunks;
    }
    a.n_tiles = a.B * a.tiles_t * a.tiles_n * a.ksplit;
    a.nst = tcv_stages_for(bn, a.staged != 0, a.res16 != nullptr);
    // a.kgroup on entry: the caller's wish (0 = automatic: 2 for convs with more than one tap, 1 for plain GEMMs)
    const int want = a.kgroup > 0 ? a.kgroup : (a.taps > 1 ? 2 : 1);
    a.kgroup = 1;
    if (want >= 2 && a.nst >= 2 * want) { a.kgroup = want; a.nst -= a.nst % want; }
}
inline bool tcv_vec_ok(const TcConvArgs& a) {
    auto al = [](const void* p) { return (((uintptr_t)p) & 15) == 0; };
    return al(a.out16) && al(a.out16r) && a.ld16r % 8 == 0 && al(a.out32) && al(a.res32) && al(a.res16) &&
           a.ld16 % 8 == 0 && a.ld32 % 4 == 0 && a.ldres % 8 == 0;
}
// tensor-core launch (maps built by the caller with box rows 128 for A and `bn` for W); persistent grid of <= n_sm CTAs
inline cudaError_t tcv_launch(cudaStream_t s, int n_sm, bool f16, int bn, const CUtensorMap& mA, const CUtensorMap& mW, TcConvArgs a) {
    a.vec_ok = tcv_vec_ok(a) ? 1 : 0;
    // 16-bit-only epilogues (the decoder's convs) go through shared memory: see tcv_staged_epilogue
    a.staged = (a.staged_ok && a.vec_ok && a.out32 == nullptr && a.res32 == nullptr && a.gamma == nullptr && a.act == TCV_ACT_NONE &&
                !a.round_lin && a.ksplit <= 1 && a.N % tcv_panel_cols(bn) == 0 && (a.out16 != nullptr || a.out16r != nullptr)) ? 1 : 0;
    tcv_plan(a, bn);
    if (a.n_tiles <= 0) return cudaSuccess;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(a.n_tiles < n_sm ? a.n_tiles : n_sm); cfg.blockDim = dim3(TCV_THREADS);
    cfg.dynamicSmemBytes = tcv_smem_bytes(bn, a.nst, a.staged != 0, a.res16 != nullptr); cfg.stream = s;
    tcv_fn k = f16 ? tcv_pick<true>(bn) : tcv_pick<false>(bn);
    return cudaLaunchKernelEx(&cfg, k, mA, mW, a);
}
inline cudaError_t tcv_launch_simt(cudaStream_t s, bool f16, TcConvArgs a) {
    a.vec_ok = tcv_vec_ok(a) ? 1 : 0;
    const long long n = (long long)a.B * a.T * ((a.N + 15) / 16);
    if (n <= 0) return cudaSuccess;
    const int blocks = (int)((n + 127) / 128);
    if (f16) tcv_simt_kernel<true><<<blocks, 128, 0, s>>>(a);
    else tcv_simt_kernel<false><<<blocks, 128, 0, s>>>(a);
    return cudaGetLastError();
}
