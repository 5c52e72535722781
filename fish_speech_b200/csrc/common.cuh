// Shared device helpers for the fish-speech hot-path kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>
#include <condition_variable>
#include <map>
#include <set>
#include <utility>
// Function attributes (dynamic shared memory opt-in) are per device: one-time setup is keyed by (current device, key) so
// that a thread-per-GPU host sets them on every device it uses.  `first` is true for the first thread that asks for this
// key on the current device; it does the setup while the object lives.  Any other thread that asks meanwhile WAITS until
// that setup is finished (two host threads starting on a cold device: the second must not launch before the first has set
// the attribute).
struct FsbOnce {
    bool first = true;
    FsbOnce(int key) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) return;          // no device: the caller's own CUDA call reports it
        id_ = std::make_pair(dev, key);
        tracked_ = true;
        std::unique_lock<std::mutex> lock(mu());
        auto it = state().find(id_);
        if (it == state().end()) { state()[id_] = 1; return; }  // 1: being set up by this thread
        first = false;
        cv().wait(lock, [&] { return state()[id_] == 2; });
    }
    ~FsbOnce() {
        if (!tracked_ || !first) return;
        { std::lock_guard<std::mutex> lock(mu()); state()[id_] = 2; }
        cv().notify_all();
    }
    FsbOnce(const FsbOnce&) = delete;
    FsbOnce& operator=(const FsbOnce&) = delete;

private:
    std::pair<int, int> id_{0, 0};
    bool tracked_ = false;
    static std::mutex& mu() { static std::mutex m; return m; }
    static std::condition_variable& cv() { static std::condition_variable c; return c; }
    static std::map<std::pair<int, int>, int>& state() { static std::map<std::pair<int, int>, int> s; return s; }
};

#ifndef __CUDA_ARCH__
#define FSB_ARCH_OK 1
#elif __CUDA_ARCH__ >= 1000
#define FSB_ARCH_OK 1
#else
#error "fish_speech_b200 kernels are written for sm_100a only"
#endif

typedef __nv_bfloat16 bf16;

#define FSB_WARP 32

// ----------------------------------------------------------------------------- bf16 <-> fp32
__device__ __forceinline__ float bf16_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
__device__ __forceinline__ float bf16_bits_to_f(uint16_t b) { return __uint_as_float(((uint32_t)b) << 16); }
__device__ __forceinline__ uint16_t f_to_bf16_bits(float f) {
    return __bfloat16_as_ushort(__float2bfloat16_rn(f));
}
// round-trip through bf16 (one RNE rounding), the reference's per-op rounding point
__device__ __forceinline__ float rbf(float f) { return __bfloat162float(__float2bfloat16_rn(f)); }

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
    f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x);
    f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
    f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z);
    f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint32_t pack2(float a, float b) {
    return (uint32_t)f_to_bf16_bits(a) | ((uint32_t)f_to_bf16_bits(b) << 16);
}

// acc += w[e] * x[e] for the 8 bf16 pairs of two 128-bit words, e -> acc[e & 3].  sm_100 has a native mixed-precision FMA
// (PTX fma.rn.f32.bf16 -> SASS FHFMA.BF16 Rd, Ra.H0/.H1, Rb.H0/.H1, Rc): fp32 += bf16 * bf16 straight from the packed
// register halves.  The product of two bf16 values is exact in fp32, so this equals fmaf(float(w), float(x), acc) bit for bit
// at 8 instructions per 8 weights instead of 24 (16 unpacks + 8 FMAs).
__device__ __forceinline__ void dot8_bf16(const uint4& w, const uint4& x, float (&acc)[4]) {
    const uint32_t ww[4] = {w.x, w.y, w.z, w.w}, xw[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        asm("{\n\t.reg .b16 wl, wh, xl, xh;\n\t"
            "mov.b32 {wl, wh}, %2;\n\tmov.b32 {xl, xh}, %3;\n\t"
            "fma.rn.f32.bf16 %0, wl, xl, %0;\n\t"
            "fma.rn.f32.bf16 %1, wh, xh, %1;\n\t}"
            : "+f"(acc[(2 * i) & 3]), "+f"(acc[(2 * i + 1) & 3])
            : "r"(ww[i]), "r"(xw[i]));
    }
}

// ----------------------------------------------------------------------------- memory
// streaming 128-bit load: read-only path, do not allocate in L1 (weights are touched once per step)
__device__ __forceinline__ uint4 ld_stream(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
// bulk L2 prefetch of a contiguous byte range (16 B aligned, size multiple of 16)
__device__ __forceinline__ void l2_prefetch_bulk(const void* p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(p), "r"(bytes) : "memory");
}
// spread a prefetch of [base, base+bytes) over the CTAs of the grid; called by one thread per CTA
__device__ __forceinline__ void l2_prefetch_share(const void* base, size_t bytes, int cta, int n_cta) {
    if (base == nullptr || bytes == 0) return;
    size_t per = ((bytes / n_cta) + 15) & ~(size_t)15;
    size_t off = (size_t)cta * per;
    if (off >= bytes) return;
    size_t n = bytes - off < per ? bytes - off : per;
    n &= ~(size_t)15;
    const char* p = (const char*)base + off;
    while (n > 0) {
        uint32_t c = n > 65536 ? 65536u : (uint32_t)n;
        l2_prefetch_bulk(p, c);
        p += c; n -= c;
    }
}

// ----------------------------------------------------------------------------- programmatic dependent launch
// Everything before pdl_wait() may only touch immutable data (weights); all writes come after it.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ----------------------------------------------------------------------------- reductions
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
// block-wide sum; `red` is >= 32 floats of shared memory; result broadcast to every thread
__device__ __forceinline__ float block_sum(float v, float* red) {
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    float t = (lane < nw) ? red[lane] : 0.f;
    t = warp_sum(t);
    return t;
}
__device__ __forceinline__ float block_max(float v, float* red) {
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    float t = (lane < nw) ? red[lane] : -INFINITY;
    t = warp_max(t);
    return t;
}

// ----------------------------------------------------------------------------- thread-group sync policies
// Device "bodies" (attention, sampling, ...) are written against a policy so that the same code runs as a whole
// CTA of a standalone kernel (SyncBlock) or on a subset of the warps of the persistent frame kernel (SyncNamed).
template <int NT_>
struct SyncBlock {
    static constexpr int NT = NT_;
    static __device__ __forceinline__ void sync() { __syncthreads(); }
};
template <int ID, int NT_>
struct SyncNamed {
    static constexpr int NT = NT_;
    static __device__ __forceinline__ void sync() { asm volatile("bar.sync %0, %1;" ::"n"(ID), "n"(NT_) : "memory"); }
};
// group-wide sum over SY::NT threads (tid = index inside the group); `red` >= 32 floats of shared memory
template <class SY>
__device__ __forceinline__ float group_sum(float v, float* red, int tid) {
    const int lane = tid & 31, w = tid >> 5;
    v = warp_sum(v);
    SY::sync();
    if (lane == 0) red[w] = v;
    SY::sync();
    float t = (lane < SY::NT / 32) ? red[lane] : 0.f;
    return warp_sum(t);
}

// ----------------------------------------------------------------------------- activation loads
// Activations are produced by other CTAs of the same launch in the persistent frame kernel, so they are always read
// through L2 (ld.global.cg): L1 is not coherent across SMs within a launch.
__device__ __forceinline__ float lda(const bf16* p) {
    return bf16_bits_to_f(__ldcg(reinterpret_cast<const unsigned short*>(p)));
}
__device__ __forceinline__ bf16 lda_raw(const bf16* p) {
    return __ushort_as_bfloat16(__ldcg(reinterpret_cast<const unsigned short*>(p)));
}
__device__ __forceinline__ uint4 lda(const uint4* p) { return __ldcg(p); }
__device__ __forceinline__ int lda(const int* p) { return __ldcg(p); }

// ----------------------------------------------------------------------------- flag-in-data activations
// Inside the persistent frame kernel an activation element travels as ONE 32-bit word (tag << 16) | bf16 bits, where
// the tag names the op instance that produced it.  A 32-bit store is single-copy atomic, so a reader that sees the
// expected tag also sees the value: producers need no fence and no flag, consumers spin on the data they have to
// load anyway.  This replaces "fence + grid barrier + reload" between dependent ops (the same idea as the
// low-latency protocol of collective libraries).
// poll back-off in ns (tunable per device: option "spin_ns")
static __device__ unsigned fsb_spin_ns = 100;
// Watchdog word: a wait that outlives FSB_WATCHDOG_CLOCKS (a lost producer, a protocol bug) records a code here and
// gives up; once it is non-zero every other wait of the launch gives up at its next check, so the kernel drains and
// ends normally with garbage results instead of hanging the GPU or killing the context with a trap.  The host reads and
// clears it (fsb_dual_ar_device_error) and reports it through *_last_error.
static __device__ unsigned fsb_dev_error = 0;
#define FSB_WATCHDOG_CLOCKS 4000000000LL
enum { FSB_ERR_TAG_WAIT = 1, FSB_ERR_MBAR_WAIT = 2, FSB_ERR_GRID_WAIT = 3, FSB_ERR_RING_WAIT = 4, FSB_ERR_PARTIAL_WAIT = 5 };
// called every `mask + 1` failed polls; true = stop waiting
__device__ __forceinline__ bool fsb_watchdog(unsigned& spins, long long& t0, unsigned mask, unsigned code) {
    if ((++spins & mask) != 0) return false;
    if (*(volatile unsigned*)&fsb_dev_error != 0) return true;
    if (t0 == 0) { t0 = clock64(); return false; }
    if (clock64() - t0 <= FSB_WATCHDOG_CLOCKS) return false;
    atomicCAS(&fsb_dev_error, 0u, code);
    return true;
}
__device__ __forceinline__ void st_tagged(uint32_t* p, float v, unsigned tag) {
    const uint32_t w = (tag << 16) | (uint32_t)f_to_bf16_bits(v);
    asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(w) : "memory");
}
__device__ __forceinline__ void st_tagged4(uint32_t* p, uint32_t lo_pair, uint32_t hi_pair, unsigned tag) {
    const uint32_t t = tag << 16;
    asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(t | (lo_pair & 0xffffu)),
                 "r"(t | (lo_pair >> 16)), "r"(t | (hi_pair & 0xffffu)), "r"(t | (hi_pair >> 16))
                 : "memory");
}
// back off between polls (tens of thousands of threads polling the same L2 slices delay the very stores they wait
// for); true = the watchdog expired, stop waiting
__device__ __forceinline__ bool spin_watchdog(unsigned& spins, long long& t0, unsigned code = FSB_ERR_TAG_WAIT) {
    __nanosleep(fsb_spin_ns);
    return fsb_watchdog(spins, t0, 0x3ffu, code);
}
// 4 consecutive elements (16-byte aligned); returns them as two packed bf16 pairs once all four carry `tag`
__device__ __forceinline__ uint2 ld_tagged4(const uint32_t* p, unsigned tag) {
    uint32_t a, b, c, d;
    unsigned spins = 0;
    long long t0 = 0;
    for (;;) {
        asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "l"(p) : "memory");
        if ((a >> 16) == tag && (b >> 16) == tag && (c >> 16) == tag && (d >> 16) == tag) break;
        if (spin_watchdog(spins, t0)) break;
    }
    return make_uint2((a & 0xffffu) | (b << 16), (c & 0xffffu) | (d << 16));
}
// one attempt: false if any of the 4 words does not carry `tag` yet
__device__ __forceinline__ bool ld_tagged4_try(const uint32_t* p, unsigned tag, uint2& out) {
    uint32_t a, b, c, d;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "l"(p) : "memory");
    out = make_uint2((a & 0xffffu) | (b << 16), (c & 0xffffu) | (d << 16));
    return (a >> 16) == tag && (b >> 16) == tag && (c >> 16) == tag && (d >> 16) == tag;
}
__device__ __forceinline__ uint32_t ld_tagged2(const uint32_t* p, unsigned tag) {
    uint32_t a, b;
    unsigned spins = 0;
    long long t0 = 0;
    for (;;) {
        asm volatile("ld.volatile.global.v2.u32 {%0,%1}, [%2];" : "=r"(a), "=r"(b) : "l"(p) : "memory");
        if ((a >> 16) == tag && (b >> 16) == tag) break;
        if (spin_watchdog(spins, t0)) break;
    }
    return (a & 0xffffu) | (b << 16);
}
// value of an element whose tag is already known to be current (validated earlier by this CTA)
__device__ __forceinline__ float ld_tagged_value(const uint32_t* p) {
    uint32_t a;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(a) : "l"(p) : "memory");
    return bf16_bits_to_f((uint16_t)(a & 0xffffu));
}

// 64-bit flag-in-data words for fp32 payloads (split-attention partials): {value, tag}, one 8-byte store
__device__ __forceinline__ void st_ll64(uint2* p, float v, uint32_t tag) {
    asm volatile("st.volatile.global.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(__float_as_uint(v)), "r"(tag) : "memory");
}
__device__ __forceinline__ bool ld_ll64_try(const uint2* p, uint32_t tag, float& v) {
    uint32_t a, b;
    asm volatile("ld.volatile.global.v2.u32 {%0,%1}, [%2];" : "=r"(a), "=r"(b) : "l"(p) : "memory");
    v = __uint_as_float(a);
    return b == tag;
}
