// DAC codec decode engine: DAC.from_indices (modded_dac.py:925-927) = quantizer.decode (rvq.py:352-366) + Decoder
// (modded_dac.py:760-801) behind the C ABI of include/fish_b200.h.  Weights arrive folded (weight norm applied) and
// packed [taps][Cin][N] by the host mirror (fish_speech_b200/dac.py); activations are fp32 channels-last.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <map>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/fish_b200.h"
#include "dac_kernels.cuh"
#include "tc_conv.cuh"

struct fsb_dac {
    fsb_dac_desc d;
    std::map<std::string, const float*> w;
    std::map<std::string, long long> wn;
    float* buf[4] = {nullptr, nullptr, nullptr, nullptr};
    size_t buf_elems = 0;
    int tf32 = 1;  // 1: GEMM-shaped layers on the tensor cores (TF32 operands, fp32 accumulate); 0: fp32 SIMT (parity anchor)
    int tc = 1;    // 1: tcgen05 pipeline (csrc/tc_conv.cuh: 16-bit operands, fp32 TMEM accumulation, fp32 residual stream) when the
                   //    16-bit weights are bound and the device has the tensor-map encoder; 0: the round-1 kernels above
    int half = 1;  // operand type of the tcgen05 pipeline: 1 fp16 (11-bit significand, as TF32), 0 bf16; set by the host mirror to
                   //    match the type it packed the ".h" weights in
    int pos_offset = 0;  // position of frame 0 of the next decode calls in the window transformer's RoPE table (pieces of a long clip)
    int staged_epi = 1;   // tcv_kernel: 16-bit-only epilogues (the decoder's convs) through shared memory (coalesced residual loads / stores)
    int tiled_small = 1;  // window attention and the output conv of the tcgen05 pipeline as shared-memory-tiled kernels (0: per-query / per-sample)
    int force_simt = 0;  // tcgen05 pipeline with every GEMM on the CUDA-core twin kernel (cross-check)
    int conv_group = 0;  // tcv_kernel: ring slots per barrier round (0: automatic = 2 for convs with taps, 1: one slot per round)
    // encode side (DAC.encode): set through fsb_dac_set_option ("enc_dim", "enc_n_rates", "enc_rate0..3", "enc_tf0..3",
    // "enc_window", "enc_head_dim"); 0 = not configured
    int enc_dim = 0, enc_n = 0, enc_rate[4] = {0, 0, 0, 0}, enc_tf[4] = {0, 0, 0, 0}, enc_window = 512, enc_hd = 64;
    int enc_exact = 1;   // 1: fp32 CUDA-core GEMMs on the encode side; 0: they follow "tf32" (measurement of how many codes move)
    int n_sm = 148;
    std::map<std::tuple<const void*, uint64_t, uint64_t, uint64_t, uint32_t, int>, CUtensorMap> maps;
    std::string err;
};

namespace {

int dfail(fsb_dac* h, const std::string& m) {
    if (h) h->err = m;
    return 1;
}
#define DCK(call)                                                                                           \
    do {                                                                                                    \
        cudaError_t e_ = (call);                                                                            \
        if (e_ != cudaSuccess) return dfail(h, std::string(#call) + ": " + cudaGetErrorString(e_) + " @" + std::to_string(__LINE__)); \
    } while (0)

int rates(const fsb_dac_desc& d, int i) { const int r[4] = {d.rate0, d.rate1, d.rate2, d.rate3}; return r[i]; }
int ups(const fsb_dac_desc& d, int i) { return i == 0 ? d.up0 : d.up1; }

// largest activation (floats) per batch item and frame over the whole plan
size_t per_frame_elems(const fsb_dac_desc& d) {
    size_t m = std::max<size_t>({(size_t)3 * d.tf_dim, (size_t)2 * d.tf_ffn, (size_t)d.latent_dim});
    size_t len = 1;
    for (int i = 0; i < d.n_up; ++i) {
        len *= ups(d, i);
        m = std::max(m, len * (size_t)d.convnext_ratio * d.latent_dim);
    }
    m = std::max(m, len * (size_t)d.decoder_dim);
    size_t ch = d.decoder_dim;
    for (int i = 0; i < d.n_rates; ++i) {
        len *= rates(d, i);
        ch /= 2;
        m = std::max(m, len * ch);
    }
    return m;
}

struct Runner {
    fsb_dac* h;
    cudaStream_t s;
    int B;
    int rc = 0;
    bool exact = false;   // fp32 CUDA-core GEMMs whatever "tf32" says (the encode side: codes are arg-max results)
    const float* W(const std::string& n) {
        auto it = h->w.find(n);
        if (it == h->w.end()) { if (!rc) rc = dfail(h, "weight not bound: " + n); return nullptr; }
        return it->second;
    }
    void conv(const float* A, float* out, int T, int Cin, int N, int taps, int dil, const std::string& wn,
              const std::string& bn, int bias_mod, const float* alpha, const float* res, const float* gamma, int epi) {
        ConvArgs a{};
        a.A = A; a.out = out; a.W = W(wn); a.bias = bn.empty() ? nullptr : W(bn);
        a.alpha = alpha; a.res = res; a.gamma = gamma;
        a.B = B; a.T = T; a.Cin = Cin; a.N = N; a.taps = taps; a.dil = dil; a.bias_mod = bias_mod > 0 ? bias_mod : N; a.epi = epi;
        if (rc) return;
        if (h->tf32 && !exact && (taps - 1) * dil <= CM_HALO && Cin % 4 == 0 && N % 4 == 0) {
            if (N % 96 == 0 && N % 128 != 0) {
                dim3 grid((T + CM_BM - 1) / CM_BM, N / 96, B);
                conv_mma_kernel<3><<<grid, 256, CM_SMEM, s>>>(a);
            } else {
                dim3 grid((T + CM_BM - 1) / CM_BM, (N + CM_BN - 1) / CM_BN, B);
                conv_mma_kernel<4><<<grid, 256, CM_SMEM, s>>>(a);
            }
            return;
        }
        dim3 grid((T + CG_BM - 1) / CG_BM, (N + CG_BN - 1) / CG_BN, B);
        conv_gemm_kernel<<<grid, 256, 0, s>>>(a);
    }
    // ---- tcgen05 pipeline: one GEMM-shaped layer.  A16 [B][T][Cin] 16-bit; weight `wn` + ".h" = [taps][N][Cin] 16-bit.
    const CUtensorMap* map(const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2, uint32_t box_rows) {
        auto key = std::make_tuple(ptr, d0, d1, d2, box_rows, h->half);
        auto it = h->maps.find(key);
        if (it != h->maps.end()) return &it->second;
        CUtensorMap m;
        if (!tcv_make_map(&m, ptr, h->half != 0, d0, d1, d2, box_rows)) { if (!rc) rc = dfail(h, "cuTensorMapEncodeTiled failed"); return nullptr; }
        return &h->maps.emplace(key, m).first->second;
    }
    void tc(const void* A16, int T, int Cin, int N, int taps, int dil, const std::string& wn, const std::string& bn, int bias_mod,
            const float* gamma, const float* res32, const float* alpha, int alpha_mod, int act, void* out16, float* out32,
            const void* res16 = nullptr, void* out16r = nullptr) {
        TcConvArgs a{};
        a.B = B; a.T = T; a.N = N; a.Cin = Cin; a.taps = taps; a.dil = dil;
        a.A16 = A16; a.W16 = W(wn + ".h");
        a.bias = bn.empty() ? nullptr : W(bn);
        a.bias_mod = bias_mod > 0 ? bias_mod : N; a.alpha_mod = alpha_mod > 0 ? alpha_mod : N;
        a.gamma = gamma; a.res32 = res32; a.alpha = alpha; a.act = act;
        a.out16 = out16; a.out32 = out32; a.ld16 = act == TCV_ACT_SWIGLU ? N / 2 : N; a.ld32 = N; a.ldres = N;
        a.res16 = res16; a.out16r = out16r; a.ld16r = N;
        a.staged_ok = h->staged_epi;
        a.kgroup = h->conv_group;
        if (rc) return;
        cudaError_t e;
        if (h->force_simt || !tcv_tc_ok(Cin, N)) {
            e = tcv_launch_simt(s, h->half != 0, a);
        } else {
            const int bnw = tcv_pick_bn(N);
            const CUtensorMap* mA = map(A16, (uint64_t)Cin, (uint64_t)T, (uint64_t)B, TCV_BM);
            const CUtensorMap* mW = map(a.W16, (uint64_t)Cin, (uint64_t)N, (uint64_t)taps, (uint32_t)bnw);
            if (!mA || !mW) return;
            e = tcv_launch(s, h->n_sm, h->half != 0, bnw, *mA, *mW, a);
        }
        if (e != cudaSuccess && !rc) rc = dfail(h, std::string("tc_conv launch (") + wn + "): " + cudaGetErrorString(e));
    }
};

int decode_tc(fsb_dac* h, void* codes, int B, int T, float* wav, cudaStream_t s);

}  // namespace

extern "C" {

size_t fsb_dac_workspace_bytes(const fsb_dac_desc* d) {
    return 4 * ((size_t)d->max_batch * d->max_frames * per_frame_elems(*d) * sizeof(float) + 256);
}

int fsb_dac_create(const fsb_dac_desc* desc, void* workspace, size_t workspace_bytes, fsb_dac** out) {
    if (!desc || !workspace || !out) return 1;
    fsb_dac* h = new fsb_dac();
    h->d = *desc;
    *out = h;
    const fsb_dac_desc& d = h->d;
    if (d.n_rates < 1 || d.n_rates > 4 || d.n_up < 0 || d.n_up > 2 || d.n_codebooks + 1 > 16 || d.codebook_dim > 16 ||
        d.tf_window > 128 || (d.tf_head_dim != 64 && d.tf_head_dim != 32) || d.latent_dim % 4 || d.decoder_dim % 64)
        return dfail(h, "unsupported DAC shape");
    if (workspace_bytes < fsb_dac_workspace_bytes(desc)) return dfail(h, "workspace too small");
    {
        int dev = 0, n_sm = 0;
        if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n_sm > 0)
            h->n_sm = n_sm;
    }
    h->buf_elems = (size_t)d.max_batch * d.max_frames * per_frame_elems(d);
    char* p = (char*)workspace;
    for (int i = 0; i < 4; ++i) {
        h->buf[i] = (float*)p;
        p += (h->buf_elems * sizeof(float) + 255) / 256 * 256;
    }
    return 0;
}

int fsb_dac_destroy(fsb_dac* h) { delete h; return 0; }
const char* fsb_dac_last_error(const fsb_dac* h) { return h ? h->err.c_str() : "null handle"; }

int fsb_dac_bind_weight(fsb_dac* h, const char* name, const void* dptr, int64_t numel) {
    if (!h || !name || !dptr) return 1;
    if (((uintptr_t)dptr) & 15) return dfail(h, std::string(name) + ": pointer must be 16-byte aligned");
    h->w[name] = (const float*)dptr;
    h->wn[name] = numel;
    return 0;
}

double fsb_dac_flops(const fsb_dac* h, int32_t B, int32_t T) {
    if (!h) return 0;
    const fsb_dac_desc& d = h->d;
    double f = 0, len = T;
    const double D = d.latent_dim;
    f += len * d.tf_layers * 2.0 * (D * 3 * D + D * D + 3.0 * D * d.tf_ffn);  // transformer GEMMs
    for (int i = 0; i < d.n_up; ++i) {
        f += len * 2.0 * D * ups(d, i) * D;  // transposed conv
        len *= ups(d, i);
        f += len * 2.0 * (2.0 * D * d.convnext_ratio * D + 7 * D);
    }
    double ch = d.decoder_dim;
    f += len * 2.0 * 7 * D * ch;
    for (int i = 0; i < d.n_rates; ++i) {
        const int s = rates(d, i);
        f += len * 2.0 * 2 * ch * s * (ch / 2);
        len *= s; ch /= 2;
        f += len * 3 * 2.0 * (7 * ch * ch + ch * ch);
    }
    f += len * 2.0 * 7 * ch;
    return f * B;
}

int fsb_dac_set_option(fsb_dac* h, const char* name, int64_t value) {
    if (!h || !name) return 1;
    const std::string n(name);
    if (n == "tf32") h->tf32 = (int)value;
    else if (n == "tc") h->tc = (int)value;
    else if (n == "half") { h->half = (int)value; h->maps.clear(); }
    else if (n == "force_simt") h->force_simt = (int)value;
    else if (n == "tiled_small") h->tiled_small = (int)value;
    else if (n == "staged_epi") h->staged_epi = (int)value;
    else if (n == "conv_group") { if (value < 0 || value > 3) return dfail(h, "conv_group must be 0..3"); h->conv_group = (int)value; }
    else if (n == "pos_offset") { if (value < 0) return dfail(h, "pos_offset < 0"); h->pos_offset = (int)value; }
    else if (n == "enc_dim") h->enc_dim = (int)value;
    else if (n == "enc_n_rates") { if (value < 1 || value > 4) return dfail(h, "enc_n_rates must be 1..4"); h->enc_n = (int)value; }
    else if (n.size() == 9 && n.compare(0, 8, "enc_rate") == 0 && n[8] >= '0' && n[8] <= '3') h->enc_rate[n[8] - '0'] = (int)value;
    else if (n.size() == 7 && n.compare(0, 6, "enc_tf") == 0 && n[6] >= '0' && n[6] <= '3') h->enc_tf[n[6] - '0'] = (int)value;
    else if (n == "enc_window") h->enc_window = (int)value;
    else if (n == "enc_head_dim") h->enc_hd = (int)value;
    else if (n == "enc_exact") h->enc_exact = (int)value;
    else return dfail(h, "unknown option " + n);
    return 0;
}

int fsb_dac_stage(fsb_dac* h, const char* name, void** dptr, size_t* bytes) {
    if (!h || !name || !dptr || !bytes) return 1;
    std::string n(name);
    if (n.size() == 4 && n.compare(0, 3, "buf") == 0 && n[3] >= '0' && n[3] <= '3') {
        *dptr = h->buf[n[3] - '0'];
        *bytes = h->buf_elems * sizeof(float);
        return 0;
    }
    return dfail(h, "unknown stage buffer " + n);
}

static TcConvArgs tcv_from_abi(const fsb_tc_conv_args& x) {
    TcConvArgs a{};
    a.B = x.B; a.T = x.T; a.N = x.N; a.Cin = x.Cin; a.taps = x.taps; a.dil = x.dil;
    a.A16 = x.A16; a.W16 = x.W16; a.out16 = x.out16; a.out32 = x.out32; a.ld16 = x.ld16; a.ld32 = x.ld32;
    a.out16r = x.out16r; a.ld16r = x.ld16r;
    a.bias = x.bias; a.gamma = x.gamma; a.res32 = x.res32; a.res16 = x.res16; a.alpha = x.alpha;
    a.bias_mod = x.bias_mod > 0 ? x.bias_mod : x.N; a.alpha_mod = x.alpha_mod > 0 ? x.alpha_mod : x.N; a.ldres = x.ldres;
    a.act = x.act; a.round_lin = x.round_lin;
    a.staged_ok = x.use_simt >= 0 ? 1 : 0;   // use_simt = -1: the tensor-core kernel with the direct (unstaged) epilogue, for A/B timing
    return a;
}

int fsb_tc_conv(const fsb_tc_conv_args* x, void* stream) {
    if (!x || !x->A16 || !x->W16 || (!x->out16 && !x->out32 && !x->out16r)) return 1;
    if (x->B < 1 || x->T < 1 || x->Cin < 1 || x->N < 1 || x->taps < 1 || x->dil < 1) return 1;
    if (x->act == TCV_ACT_SWIGLU && (!x->out16 || (x->N & 1))) return 1;
    TcConvArgs a = tcv_from_abi(*x);
    cudaStream_t s = (cudaStream_t)stream;
    if (x->use_simt > 0 || !tcv_tc_ok(x->Cin, x->N)) return tcv_launch_simt(s, x->f16 != 0, a) == cudaSuccess ? 0 : 4;
    int dev = 0, n_sm = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 2;
    if (tcv_set_attrs() != cudaSuccess) return 2;
    const int bn = tcv_pick_bn(x->N);
    CUtensorMap mA, mW;
    if (!tcv_make_map(&mA, x->A16, x->f16 != 0, (uint64_t)x->Cin, (uint64_t)x->T, (uint64_t)x->B, TCV_BM)) return 3;
    if (!tcv_make_map(&mW, x->W16, x->f16 != 0, (uint64_t)x->Cin, (uint64_t)x->N, (uint64_t)x->taps, (uint32_t)bn)) return 3;
    return tcv_launch(s, n_sm, x->f16 != 0, bn, mA, mW, a) == cudaSuccess ? 0 : 4;
}

// codes: device int64 [B][1+n_codebooks][T] (clamped in place like rvq.py:354-359); wav: device fp32 [B][hop*T]
int fsb_dac_decode(fsb_dac* h, void* codes, int32_t B, int32_t T, void* wav, void* stream) {
    if (!h || !codes || !wav) return 1;
    const fsb_dac_desc& d = h->d;
    if (B < 1 || B > d.max_batch || T < 1 || T > d.max_frames) return dfail(h, "B/T exceed the workspace (max_batch/max_frames)");
    {
        auto it = h->wn.find("tf.rope");
        if (it != h->wn.end() && (int64_t)(h->pos_offset + T) * d.tf_head_dim > it->second)
            return dfail(h, "pos_offset + T exceeds the bound tf.rope table");
    }
    cudaStream_t s = (cudaStream_t)stream;
    if (h->tc && h->w.count("dec.in_w.h")) return decode_tc(h, codes, B, T, (float*)wav, s);
    Runner r{h, s, B};
    const int D = d.latent_dim, NQ = d.n_codebooks + 1;
    float *b0 = h->buf[0], *b1 = h->buf[1], *b2 = h->buf[2], *b3 = h->buf[3];

    // ---- K11: RVQ dequant -> b0 [B][T][D]
    {
        RvqArgs a{};
        a.codes = (long long*)codes; a.codebooks = r.W("rvq.codebooks"); a.proj_w = r.W("rvq.proj_w");
        a.proj_b = r.W("rvq.proj_b"); a.z = b0;
        a.B = B; a.T = T; a.NQ = NQ; a.D = D; a.CD = d.codebook_dim; a.size0 = d.semantic_codebook_size;
        a.size_rest = d.codebook_size;
        if (r.rc) return r.rc;
        rvq_dequant_kernel<<<dim3(T, B), 256, 0, s>>>(a);
    }
    // ---- K12: post_module window transformer (modded_dac.py:418-439,145-171)
    const int H = d.tf_heads, HD = d.tf_head_dim, F = d.tf_ffn;
    const long long rows = (long long)B * T;
    const int nwb = (int)((rows * 32 + 255) / 256);
    const float* rope = r.W("tf.rope");
    for (int l = 0; l < d.tf_layers; ++l) {
        const std::string p = "tf." + std::to_string(l) + ".";
        const float *an = r.W(p + "attn_norm"), *fn = r.W(p + "ffn_norm"), *g1 = r.W(p + "gamma1"), *g2 = r.W(p + "gamma2");
        if (r.rc) return r.rc;
        rmsnorm_rows_kernel<<<nwb, 256, 0, s>>>(b0, an, b1, rows, D, d.tf_norm_eps);
        r.conv(b1, b2, T, D, 3 * H * HD, 1, 1, p + "wqkv", "", 0, nullptr, nullptr, nullptr, DEPI_NONE);
        {
            const long long n = rows * 2 * H * (HD / 2);
            rope_qk_kernel<<<(int)((n + 255) / 256), 256, 0, s>>>(b2, rope, B, T, H, HD, h->pos_offset);
            dim3 grid((T + 3) / 4, H, B);
            const float scale = 1.f / sqrtf((float)HD);
            if (HD == 64) window_attn_kernel<64><<<grid, 128, 0, s>>>(b2, b1, B, T, H, d.tf_window, scale);
            else window_attn_kernel<32><<<grid, 128, 0, s>>>(b2, b1, B, T, H, d.tf_window, scale);
        }
        r.conv(b1, b3, T, H * HD, D, 1, 1, p + "wo", "", 0, nullptr, b0, g1, DEPI_NONE);      // h = x + gamma1 * wo(y)
        rmsnorm_rows_kernel<<<nwb, 256, 0, s>>>(b3, fn, b1, rows, D, d.tf_norm_eps);
        r.conv(b1, b2, T, D, 2 * F, 1, 1, p + "w13", "", 0, nullptr, nullptr, nullptr, DEPI_NONE);
        swiglu_kernel<<<(int)((rows * F + 255) / 256), 256, 0, s>>>(b2, b1, rows, F);
        r.conv(b1, b0, T, F, D, 1, 1, p + "w2", "", 0, nullptr, b3, g2, DEPI_NONE);           // x = h + gamma2 * w2(..)
    }
    {
        const float* nw = r.W("tf.norm");
        if (r.rc) return r.rc;
        rmsnorm_rows_kernel<<<nwb, 256, 0, s>>>(b0, nw, b1, rows, D, d.tf_norm_eps);
    }
    // ---- K13: upsample stages (rvq.py:263-276): transposed conv k=s=f, ConvNeXt block.  in b1 -> out b1
    int len = T;
    for (int i = 0; i < d.n_up; ++i) {
        const int f = ups(d, i);
        const std::string p = "up." + std::to_string(i) + ".";
        r.conv(b1, b0, len, D, f * D, 1, 1, p + "ct_w", p + "ct_b", D, nullptr, nullptr, nullptr, DEPI_NONE);
        len *= f;
        const float *dww = r.W(p + "dw_w"), *dwb = r.W(p + "dw_b"), *lw = r.W(p + "ln_w"), *lb = r.W(p + "ln_b");
        const float* gm = r.W(p + "gamma");
        if (r.rc) return r.rc;
        dwconv_ln_kernel<<<dim3(len, B), 256, D * sizeof(float), s>>>(b0, dww, dwb, lw, lb, b2, len, D);
        r.conv(b2, b3, len, D, d.convnext_ratio * D, 1, 1, p + "pw1_w", p + "pw1_b", 0, nullptr, nullptr, nullptr, DEPI_GELU);
        r.conv(b3, b1, len, d.convnext_ratio * D, D, 1, 1, p + "pw2_w", p + "pw2_b", 0, nullptr, b0, gm, DEPI_NONE);
    }
    // ---- K14: conv decoder (modded_dac.py:760-801)
    int ch = d.decoder_dim;
    r.conv(b1, b0, len, D, ch, 7, 1, "dec.in_w", "dec.in_b", 0, nullptr, nullptr, nullptr, DEPI_NONE);
    float *cur = b0, *t1 = b1, *t2 = b2, *t3 = b3;
    for (int i = 0; i < d.n_rates; ++i) {
        const int st = rates(d, i), co = ch / 2;
        const std::string p = "dec." + std::to_string(i) + ".";
        r.conv(cur, t1, len, ch, st * co, 2, 1, p + "ct_w", p + "ct_b", co, r.W(p + "alpha"), nullptr, nullptr, DEPI_NONE);
        len *= st; ch = co;
        float *x = t1, *y = t2, *o = t3, *spare = cur;
        const int dil[3] = {1, 3, 9};
        for (int j = 0; j < 3; ++j) {
            const std::string q = p + "r" + std::to_string(j) + ".";
            r.conv(x, y, len, ch, ch, 7, dil[j], q + "w1", q + "b1", 0, r.W(q + "a1"), nullptr, nullptr, DEPI_NONE);
            r.conv(y, o, len, ch, ch, 1, 1, q + "w2", q + "b2", 0, r.W(q + "a2"), x, nullptr, DEPI_NONE);
            float* tmp = x; x = o; o = spare; spare = tmp;
        }
        cur = x; t1 = y; t2 = o; t3 = spare;
        if (r.rc) return r.rc;
    }
    {
        ConvArgs a{};
        a.A = cur; a.W = r.W("dec.out_w"); a.bias = r.W("dec.out_b"); a.alpha = r.W("dec.out_alpha"); a.out = (float*)wav;
        a.B = B; a.T = len; a.Cin = ch; a.N = 1; a.taps = 7; a.dil = 1; a.bias_mod = 1; a.epi = DEPI_TANH;
        if (r.rc) return r.rc;
        const long long warps = (long long)B * len;
        conv_out1_kernel<<<(int)((warps * 32 + 255) / 256), 256, 0, s>>>(a);
    }
    DCK(cudaGetLastError());
    return r.rc;
}

}  // extern "C"

// =============================================================================================== tcgen05 decode pipeline
// Same layer plan as fsb_dac_decode above (modded_dac.py:925-927) with every GEMM-shaped layer on tcv_kernel: 16-bit operands
// written by their producers (GEMM epilogues with the consumer's Snake folded in, or the small fp32 -> 16-bit kernels), fp32
// residual stream, fp32 accumulation.  Buffers are the handle's four scratch slabs, used as raw bytes.
namespace {

template <bool F16>
int decode_tc_t(fsb_dac* h, void* codes, int B, int T, float* wav, cudaStream_t s) {
    const fsb_dac_desc& d = h->d;
    if (!h->force_simt) {
        if (rg_get_encode() == nullptr) return dfail(h, "tensor-map encoder not available (set option tc = 0)");
        DCK(tcv_set_attrs());
    }
    Runner r{h, s, B};
    const int D = d.latent_dim, NQ = d.n_codebooks + 1;
    float *b0 = h->buf[0], *b1 = h->buf[1], *b2 = h->buf[2], *b3 = h->buf[3];
    {   // K11: RVQ dequant -> b0 fp32 [B][T][D]
        RvqArgs a{};
        a.codes = (long long*)codes; a.codebooks = r.W("rvq.codebooks"); a.proj_w = r.W("rvq.proj_w");
        a.proj_b = r.W("rvq.proj_b"); a.z = b0;
        a.B = B; a.T = T; a.NQ = NQ; a.D = D; a.CD = d.codebook_dim; a.size0 = d.semantic_codebook_size;
        a.size_rest = d.codebook_size;
        if (r.rc) return r.rc;
        rvq_dequant_kernel<<<dim3(T, B), 256, 0, s>>>(a);
    }
    // K12: window transformer.  residual stream x32 in b0 / h32 in b3; b1, b2 hold 16-bit operands and the fp32 qkv
    const int H = d.tf_heads, HD = d.tf_head_dim, F = d.tf_ffn;
    const long long rows = (long long)B * T;
    const int nwb = (int)((rows * 32 + 255) / 256);
    const float* rope = r.W("tf.rope");
    for (int l = 0; l < d.tf_layers; ++l) {
        const std::string p = "tf." + std::to_string(l) + ".";
        const float *an = r.W(p + "attn_norm"), *fn = r.W(p + "ffn_norm"), *g1 = r.W(p + "gamma1"), *g2 = r.W(p + "gamma2");
        if (r.rc) return r.rc;
        rmsnorm_rows16_kernel<F16><<<nwb, 256, 0, s>>>(b0, an, b1, rows, D, d.tf_norm_eps);
        r.tc(b1, T, D, 3 * H * HD, 1, 1, p + "wqkv", "", 0, nullptr, nullptr, nullptr, 0, TCV_ACT_NONE, nullptr, b2);
        {
            const long long n = rows * 2 * H * (HD / 2);
            rope_qk_kernel<<<(int)((n + 255) / 256), 256, 0, s>>>(b2, rope, B, T, H, HD, h->pos_offset);
            dim3 grid((T + 3) / 4, H, B);
            const float scale = 1.f / sqrtf((float)HD);
            if (h->tiled_small) {  // keys / values of a query tile's windows staged once per CTA (bit-identical to the per-query kernel)
                dim3 tg((T + WA_QT - 1) / WA_QT, H, B);
                if (HD == 64) {
                    DCK(cudaFuncSetAttribute((const void*)window_attn16_tiled_kernel<64, F16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wa_smem_bytes<64>(128)));
                    window_attn16_tiled_kernel<64, F16><<<tg, WA_THREADS, wa_smem_bytes<64>(d.tf_window), s>>>(b2, b1, B, T, H, d.tf_window, scale);
                } else {
                    DCK(cudaFuncSetAttribute((const void*)window_attn16_tiled_kernel<32, F16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wa_smem_bytes<32>(128)));
                    window_attn16_tiled_kernel<32, F16><<<tg, WA_THREADS, wa_smem_bytes<32>(d.tf_window), s>>>(b2, b1, B, T, H, d.tf_window, scale);
                }
            } else if (HD == 64) window_attn16_kernel<64, F16><<<grid, 128, 0, s>>>(b2, b1, B, T, H, d.tf_window, scale);
            else window_attn16_kernel<32, F16><<<grid, 128, 0, s>>>(b2, b1, B, T, H, d.tf_window, scale);
        }
        r.tc(b1, T, H * HD, D, 1, 1, p + "wo", "", 0, g1, b0, nullptr, 0, TCV_ACT_NONE, nullptr, b3);      // h = x + gamma1 * wo(y)
        rmsnorm_rows16_kernel<F16><<<nwb, 256, 0, s>>>(b3, fn, b1, rows, D, d.tf_norm_eps);
        r.tc(b1, T, D, 2 * F, 1, 1, p + "w13", "", 0, nullptr, nullptr, nullptr, 0, TCV_ACT_SWIGLU, b2, nullptr);  // interleaved gate/up
        r.tc(b2, T, F, D, 1, 1, p + "w2", "", 0, g2, b3, nullptr, 0, TCV_ACT_NONE, nullptr, b0);            // x = h + gamma2 * w2(..)
    }
    {
        const float* nw = r.W("tf.norm");
        if (r.rc) return r.rc;
        rmsnorm_rows16_kernel<F16><<<nwb, 256, 0, s>>>(b0, nw, b1, rows, D, d.tf_norm_eps);
    }
    // K13: upsample stages.  16-bit input in b1 -> 16-bit output in b1
    int len = T;
    for (int i = 0; i < d.n_up; ++i) {
        const int f = ups(d, i);
        const std::string p = "up." + std::to_string(i) + ".";
        r.tc(b1, len, D, f * D, 1, 1, p + "ct_w", p + "ct_b", D, nullptr, nullptr, nullptr, 0, TCV_ACT_NONE, nullptr, b0);
        len *= f;
        const float *dww = r.W(p + "dw_w"), *dwb = r.W(p + "dw_b"), *lw = r.W(p + "ln_w"), *lb = r.W(p + "ln_b");
        const float* gm = r.W(p + "gamma");
        if (r.rc) return r.rc;
        dwconv_ln16_kernel<F16><<<dim3(len, B), 256, D * sizeof(float), s>>>(b0, dww, dwb, lw, lb, b2, len, D);
        r.tc(b2, len, D, d.convnext_ratio * D, 1, 1, p + "pw1_w", p + "pw1_b", 0, nullptr, nullptr, nullptr, 0, TCV_ACT_GELU, b3, nullptr);
        r.tc(b3, len, d.convnext_ratio * D, D, 1, 1, p + "pw2_w", p + "pw2_b", 0, gm, b0, nullptr, 0, TCV_ACT_NONE, b1, nullptr);
    }
    // K14: conv decoder.  Every producer applies the Snake of its consumer to the 16-bit copy (modded_dac.py:599-620, 712-757)
    int ch = d.decoder_dim;
    const int n_blk = d.n_rates;
    r.tc(b1, len, D, ch, 7, 1, "dec.in_w", "dec.in_b", 0, nullptr, nullptr, r.W("dec.0.alpha"), 0, TCV_ACT_NONE, b0, nullptr);
    void* cur16 = b0;
    float* fr[3] = {b1, b2, b3};
    for (int i = 0; i < n_blk; ++i) {
        const int st = rates(d, i), co = ch / 2;
        const std::string p = "dec." + std::to_string(i) + ".";
        void *X16 = fr[0], *A16 = fr[1], *Y16 = fr[2];  // raw residual stream (16-bit), snaked operand, unit-internal operand
        r.tc(cur16, len, ch, st * co, 2, 1, p + "ct_w", p + "ct_b", co, nullptr, nullptr, r.W(p + "r0.a1"), co, TCV_ACT_NONE, A16, nullptr,
             nullptr, X16);
        len *= st; ch = co;
        const int dil[3] = {1, 3, 9};
        for (int j = 0; j < 3; ++j) {
            const std::string q = p + "r" + std::to_string(j) + ".";
            r.tc(A16, len, ch, ch, 7, dil[j], q + "w1", q + "b1", 0, nullptr, nullptr, r.W(q + "a2"), 0, TCV_ACT_NONE, Y16, nullptr);
            const bool last = j == 2;
            const float* an = !last ? r.W(p + "r" + std::to_string(j + 1) + ".a1")
                                    : (i + 1 < n_blk ? r.W("dec." + std::to_string(i + 1) + ".alpha") : r.W("dec.out_alpha"));
            r.tc(Y16, len, ch, ch, 1, 1, q + "w2", q + "b2", 0, nullptr, nullptr, an, 0, TCV_ACT_NONE, A16, nullptr, X16, last ? nullptr : X16);
        }
        // the block's output (16-bit, snaked for its consumer) is in A16; the old input buffer becomes free
        float* old_in = (float*)cur16;
        cur16 = A16;
        fr[0] = old_in; fr[1] = (float*)X16; fr[2] = (float*)Y16;
        if (r.rc) return r.rc;
    }
    {
        const float *ow = r.W("dec.out_w"), *ob = r.W("dec.out_b");
        if (r.rc) return r.rc;
        const long long warps = (long long)B * len;
        const size_t co_bytes = co_smem_bytes(ch, 7);
        if (h->tiled_small && ch % 8 == 0 && co_bytes <= 200 * 1024) {
            DCK(cudaFuncSetAttribute((const void*)conv_out1_16_tiled_kernel<F16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
            conv_out1_16_tiled_kernel<F16><<<B * ((len + CO_TILE - 1) / CO_TILE), CO_TILE, co_bytes, s>>>(cur16, ow, ob, wav, B, len, ch, 7);
        } else {
            conv_out1_16_kernel<F16><<<(int)((warps * 32 + 255) / 256), 256, 0, s>>>(cur16, ow, ob, wav, B, len, ch, 7);
        }
    }
    DCK(cudaGetLastError());
    return r.rc;
}

int decode_tc(fsb_dac* h, void* codes, int B, int T, float* wav, cudaStream_t s) {
    return h->half ? decode_tc_t<true>(h, codes, B, T, wav, s) : decode_tc_t<false>(h, codes, B, T, wav, s);
}

}  // namespace

// =============================================================================================== encode side
// DAC.encode (modded_dac.py:874-923): waveform -> codes, for the reference voice of a cloning request (runs once per voice, in
// front of the hot path).  Everything in fp32 on the CUDA-core kernels of the decode side's parity anchor: the result is a set of
// arg-max decisions, and 16-bit or TF32 operands would move the near-ties.  Strided convs (EncoderBlock's kernel 2s / stride s,
// the quantizer's kernel = stride) need no kernel of their own in a channels-last layout: s consecutive rows of [len][C] ARE one
// row of [len / s][s * C], so they are 2-tap / 1-tap causal convs over that view with the weights regrouped at pack time.
namespace {

// x in bx, temporaries ba / bb / bc; the normed output lands in ba (Transformer.forward, modded_dac.py:145-171)
int window_transformer32(Runner& r, const std::string& pre, int layers, int D, int H, int HD, int F, int window, const float* rope,
                         int T, float eps, float* bx, float* ba, float* bb, float* bc) {
    fsb_dac* h = r.h;
    cudaStream_t s = r.s;
    const int B = r.B;
    const long long rows = (long long)B * T;
    const int nwb = (int)((rows * 32 + 255) / 256);
    for (int l = 0; l < layers; ++l) {
        const std::string p = pre + std::to_string(l) + ".";
        const float *an = r.W(p + "attn_norm"), *fn = r.W(p + "ffn_norm"), *g1 = r.W(p + "gamma1"), *g2 = r.W(p + "gamma2");
        if (r.rc) return r.rc;
        rmsnorm_rows_kernel<<<nwb, 256, 0, s>>>(bx, an, ba, rows, D, eps);
        r.conv(ba, bb, T, D, 3 * H * HD, 1, 1, p + "wqkv", "", 0, nullptr, nullptr, nullptr, DEPI_NONE);
        {
            const long long n = rows * 2 * H * (HD / 2);
            rope_qk_kernel<<<(int)((n + 255) / 256), 256, 0, s>>>(bb, rope, B, T, H, HD, 0);
            dim3 grid((T + 3) / 4, H, B);
            const float scale = 1.f / sqrtf((float)HD);
            if (HD == 64 && window <= 128) window_attn_kernel<64, 4><<<grid, 128, 0, s>>>(bb, ba, B, T, H, window, scale);
            else if (HD == 64) window_attn_kernel<64, 16><<<grid, 128, 0, s>>>(bb, ba, B, T, H, window, scale);
            else if (window <= 128) window_attn_kernel<32, 4><<<grid, 128, 0, s>>>(bb, ba, B, T, H, window, scale);
            else window_attn_kernel<32, 16><<<grid, 128, 0, s>>>(bb, ba, B, T, H, window, scale);
        }
        r.conv(ba, bc, T, H * HD, D, 1, 1, p + "wo", "", 0, nullptr, bx, g1, DEPI_NONE);      // h = x + gamma1 * wo(y)
        rmsnorm_rows_kernel<<<nwb, 256, 0, s>>>(bc, fn, ba, rows, D, eps);
        r.conv(ba, bb, T, D, 2 * F, 1, 1, p + "w13", "", 0, nullptr, nullptr, nullptr, DEPI_NONE);
        swiglu_kernel<<<(int)((rows * F + 255) / 256), 256, 0, s>>>(bb, ba, rows, F);
        r.conv(ba, bx, T, F, D, 1, 1, p + "w2", "", 0, nullptr, bc, g2, DEPI_NONE);           // x = h + gamma2 * w2(..)
    }
    const float* nw = r.W(pre + "norm");
    if (r.rc) return r.rc;
    rmsnorm_rows_kernel<<<nwb, 256, 0, s>>>(bx, nw, ba, rows, D, eps);
    DCK(cudaGetLastError());
    return 0;
}

}  // namespace

extern "C" int fsb_dac_encode(fsb_dac* h, const void* audio, int32_t B, int64_t n_samples, void* codes, void* stream) {
    if (!h || !audio || !codes) return 1;
    const fsb_dac_desc& d = h->d;
    if (h->enc_dim < 4 || h->enc_dim % 4 || h->enc_n < 1)
        return dfail(h, "encode side not configured (options enc_dim, enc_n_rates, enc_rate0.., enc_tf0.., enc_window, enc_head_dim)");
    long long hop = 1;
    for (int i = 0; i < h->enc_n; ++i) {
        if (h->enc_rate[i] < 1) return dfail(h, "enc_rate not set");
        hop *= h->enc_rate[i];
    }
    long long down = 1;
    for (int i = 0; i < d.n_up; ++i) down *= ups(d, i);
    if (B < 1 || B > d.max_batch || n_samples < hop * down || n_samples % (hop * down))
        return dfail(h, "encode: B in 1..max_batch and n_samples a positive multiple of the code frame (pad on the host)");
    const int T = (int)(n_samples / (hop * down));
    if (T > d.max_frames) return dfail(h, "encode: more code frames than max_frames");
    if ((h->enc_hd != 64 && h->enc_hd != 32) || h->enc_window > 512 || h->enc_window < 1) return dfail(h, "encode: unsupported encoder transformer shape");
    const int D = d.latent_dim, NQ = d.n_codebooks + 1;
    {   // the four scratch slabs were sized for the decode side: check what this plan needs
        size_t need = 0, len = (size_t)n_samples, ch = h->enc_dim;
        for (int i = 0; i < h->enc_n; ++i) {
            need = std::max(need, len * ch);
            len /= h->enc_rate[i];
            ch *= 2;
            if (h->enc_tf[i]) need = std::max(need, len * 6 * ch);
        }
        need = std::max(need, len * std::max<size_t>(ch, D));
        for (int i = 0; i < d.n_up; ++i) {
            len /= ups(d, i);
            need = std::max(need, len * (size_t)d.convnext_ratio * D);
        }
        need = std::max(need, len * std::max<size_t>(3 * (size_t)d.tf_heads * d.tf_head_dim, 2 * (size_t)d.tf_ffn));
        if (need * B > h->buf_elems) return dfail(h, "encode: the workspace (sized by max_batch x max_frames for the decode side) is too small for this clip");
    }
    cudaStream_t s = (cudaStream_t)stream;
    Runner r{h, s, B};
    r.exact = h->enc_exact != 0;
    float *cur = h->buf[0], *t1 = h->buf[1], *t2 = h->buf[2], *t3 = h->buf[3];
    int len = (int)n_samples, ch = h->enc_dim;
    {
        const float *w = r.W("enc.in_w"), *b = r.W("enc.in_b");
        if (r.rc) return r.rc;
        const long long n = (long long)B * len * ch;
        conv_in1_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>((const float*)audio, w, b, cur, B, len, ch, 7);
    }
    const int dil[3] = {1, 3, 9};
    for (int i = 0; i < h->enc_n; ++i) {
        const int st = h->enc_rate[i];
        const std::string p = "enc." + std::to_string(i) + ".";
        float *x = cur, *y = t1, *o = t2, *spare = t3;
        for (int j = 0; j < 3; ++j) {   // ResidualUnit (modded_dac.py:599-620)
            const std::string q = p + "r" + std::to_string(j) + ".";
            r.conv(x, y, len, ch, ch, 7, dil[j], q + "w1", q + "b1", 0, r.W(q + "a1"), nullptr, nullptr, DEPI_NONE);
            r.conv(y, o, len, ch, ch, 1, 1, q + "w2", q + "b2", 0, r.W(q + "a2"), x, nullptr, DEPI_NONE);
            float* tmp = x; x = o; o = spare; spare = tmp;
        }
        if (len % st) return dfail(h, "encode: length not a multiple of the encoder stride");
        // Snake + strided conv (kernel 2s, stride s, s zeros on the left) = 2-tap causal conv over the [len/s][s*ch] view
        r.conv(x, y, len / st, st * ch, 2 * ch, 2, 1, p + "down_w", p + "down_b", 0, r.W(p + "down_alpha"), nullptr, nullptr, DEPI_NONE);
        len /= st;
        ch *= 2;
        if (r.rc) return r.rc;
        if (h->enc_tf[i]) {
            const float* rope = r.W("enc.rope");
            if (r.rc) return r.rc;
            if ((long long)len * h->enc_hd > h->wn["enc.rope"]) return dfail(h, "encode: enc.rope table shorter than the clip");
            if (ch % h->enc_hd) return dfail(h, "encode: encoder transformer width not a multiple of the head size");
            if (window_transformer32(r, p + "tf.", h->enc_tf[i], ch, ch / h->enc_hd, h->enc_hd, 3 * ch, h->enc_window, rope, len, 1e-5f,
                                     y, x, o, spare)) return 1;
            cur = x; t1 = y; t2 = o; t3 = spare;
        } else {
            cur = y; t1 = x; t2 = o; t3 = spare;
        }
    }
    // Snake + 3-tap conv to the latent width (modded_dac.py:699-702)
    r.conv(cur, t1, len, ch, D, 3, 1, "enc.out_w", "enc.out_b", 0, r.W("enc.out_alpha"), nullptr, nullptr, DEPI_NONE);
    // quantizer.downsample (rvq.py:249-261): conv kernel = stride = f over the [len/f][f*D] view, ConvNeXt block
    for (int i = 0; i < d.n_up; ++i) {
        const int f = ups(d, i);
        const std::string p = "down." + std::to_string(i) + ".";
        if (len % f) return dfail(h, "encode: length not a multiple of the downsample factor");
        r.conv(t1, t2, len / f, f * D, D, 1, 1, p + "w", p + "b", 0, nullptr, nullptr, nullptr, DEPI_NONE);
        len /= f;
        const float *dww = r.W(p + "dw_w"), *dwb = r.W(p + "dw_b"), *lw = r.W(p + "ln_w"), *lb = r.W(p + "ln_b");
        const float* gm = r.W(p + "gamma");
        if (r.rc) return r.rc;
        dwconv_ln_kernel<<<dim3(len, B), 256, D * sizeof(float), s>>>(t2, dww, dwb, lw, lb, t3, len, D);
        r.conv(t3, cur, len, D, d.convnext_ratio * D, 1, 1, p + "pw1_w", p + "pw1_b", 0, nullptr, nullptr, nullptr, DEPI_GELU);
        r.conv(cur, t1, len, d.convnext_ratio * D, D, 1, 1, p + "pw2_w", p + "pw2_b", 0, nullptr, t2, gm, DEPI_NONE);
    }
    if (len != T) return dfail(h, "encode: internal length mismatch");
    // pre_module (the post_module's twin) then the codebooks
    {
        const float* rope = r.W("tf.rope");
        if (r.rc) return r.rc;
        if ((long long)T * d.tf_head_dim > h->wn["tf.rope"]) return dfail(h, "encode: tf.rope table shorter than the clip");
        if (window_transformer32(r, "pre.", d.tf_layers, D, d.tf_heads, d.tf_head_dim, d.tf_ffn, d.tf_window, rope, T, d.tf_norm_eps,
                                 t1, t2, t3, cur)) return 1;
    }
    {
        VqEncArgs a{};
        a.z = t2; a.in_w = r.W("vq.in_w"); a.in_b = r.W("vq.in_b"); a.cb = r.W("rvq.codebooks"); a.cbn = r.W("vq.cbn");
        a.cbsq = r.W("vq.cbsq"); a.out_w = r.W("rvq.proj_w"); a.out_b = r.W("rvq.proj_b"); a.codes = (long long*)codes;
        a.B = B; a.T = T; a.NQ = NQ; a.D = D; a.CD = d.codebook_dim; a.size0 = d.semantic_codebook_size; a.size_rest = d.codebook_size;
        if (r.rc) return r.rc;
        vq_encode_kernel<<<dim3(T, B), 256, D * sizeof(float), s>>>(a);
    }
    DCK(cudaGetLastError());
    return r.rc;
}
