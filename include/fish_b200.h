/* fish_b200.h — C ABI of the B200-native fish-speech hot path (Dual-AR decode + DAC codec decode).
 *
 * Boundary rules:
 *   - extern "C", plain pointers and sizes; no torch / C++ types cross the ABI.
 *   - every device buffer (weights, KV caches, I/O, workspace) is allocated by the caller (PyTorch) and
 *     passed as a raw device pointer; the library borrows it while the handle lives and never frees it.
 *   - every call takes an explicit cudaStream_t (passed as void*), is asynchronous, never synchronises.
 *   - return value: 0 = ok, non-zero = error; fsb_*_last_error(h) gives the message.  No exceptions,
 *     no abort().  CUDA errors are reported, not swallowed.
 *   - no global mutable state: two threads may drive two handles on one device (the reference runs the
 *     LLM on a worker thread and the codec on the request thread, inference_engine/__init__.py:86-119).
 *
 * Each entry point names the reference interface it replaces (paths relative to the reference repo).
 */
#ifndef FISH_B200_H
#define FISH_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FSB_ABI_VERSION 1
int fsb_abi_version(void);

/* ------------------------------------------------------------------------------------------------
 * Dual-AR transformer (fish_speech/models/text2semantic/llama.py: DualARTransformer:659)
 * ---------------------------------------------------------------------------------------------- */
typedef struct fsb_dual_ar fsb_dual_ar;

/* Subset of DualARModelArgs (llama.py:28-123,155-193) the decode path reads. fast_* == slow values. */
typedef struct {
    int32_t dim, n_layer, n_fast_layer, n_head, n_kv_head, head_dim, ffn, vocab;
    int32_t num_codebooks, codebook_size;
    int32_t semantic_begin_id, semantic_end_id, im_end_id; /* tokenizer ids used at inference.py:207,308-320 */
    int32_t qk_norm, scale_codebook_embeddings, norm_fastlayer_input, tie_word_embeddings;
    int32_t max_rows; /* max concurrent requests (decode) / rows per internal pass */
    float norm_eps;
} fsb_dual_ar_desc;

size_t fsb_dual_ar_workspace_bytes(const fsb_dual_ar_desc* desc, int32_t max_seq_len);
int fsb_dual_ar_create(const fsb_dual_ar_desc* desc, int32_t max_seq_len, void* workspace, size_t workspace_bytes,
                       fsb_dual_ar** out);
int fsb_dual_ar_destroy(fsb_dual_ar* h);
const char* fsb_dual_ar_last_error(const fsb_dual_ar* h);

/* Weights by the reference's state-dict names (llama.py nn.Module registration order), bf16 row-major
 * [rows][cols], e.g. "layers.3.attention.wqkv.weight", "fast_layers.0.feed_forward.w2.weight",
 * "embeddings.weight", "codebook_embeddings.weight", "norm.weight", "fast_embeddings.weight",
 * "fast_norm.weight", "fast_output.weight", "output.weight" (untied only).
 * Replaces: load_state_dict(assign=True) in BaseTransformer.from_pretrained (llama.py:480-593). */
int fsb_dual_ar_bind_weight(fsb_dual_ar* h, const char* name, const void* dptr, int64_t rows, int64_t cols);
/* RoPE tables as the reference stores them: bf16 [n_pos][head_dim/2][2] (llama.py:1003-1022). */
int fsb_dual_ar_bind_rope(fsb_dual_ar* h, int32_t is_fast, const void* table, int32_t n_pos);
/* KVCache buffers, bf16 [max_batch][n_kv_head][max_seq][head_dim] (llama.py:196-214, setup_caches:707-721). */
int fsb_dual_ar_bind_kv(fsb_dual_ar* h, int32_t is_fast, int32_t layer, void* k, void* v, int32_t max_batch,
                        int32_t max_seq);
/* Options: "pdl" (0/1), "graph" (0/1), "self_prefetch_kb", "attn_splits" (0 = auto), "mega", "masked_head", "batch_gemm",
 * "prefill_gemm", "spin_ns", ... — the full list with defaults is in DESIGN.md §2; every FSB_* environment override of
 * round 1 is an option here (the environment is no longer read, except FSB_OPTIONS="name=value,..." applied at create
 * and FSB_RG_TRACE, which makes the stage entry point fsb_rows_gemm record clock stamps for tools/rows_gemm_trace.py). */
int fsb_dual_ar_set_option(fsb_dual_ar* h, const char* name, int64_t value);

/* Prompt pass over the slow stack: fills KV for positions [pos0, pos0+T) of `slot`; no logits.
 * tokens: device int32 [T][C+1] (the transpose of the reference's (C+1, T) prompt).
 * Replaces: the Tq>1 call of BaseTransformer.forward_generate (llama.py:390-466) inside
 * decode_one_token_ar (inference.py:96-118) — the last prompt column is then run as a decode frame. */
int fsb_dual_ar_prefill(fsb_dual_ar* h, const int32_t* tokens, int32_t T, int32_t slot, int32_t pos0, void* stream);

/* One slow-AR step for B rows (DualARTransformer.forward_generate with Tq=1, llama.py:818-827).
 * tokens int32 [B][C+1], pos int32 [B], slots int32 [B] or NULL (slot = row). Results stay in the
 * handle's buffers "slow_logits" bf16 [B][vocab] and "hidden" bf16 [B][dim]. */
int fsb_dual_ar_slow_step(fsb_dual_ar* h, int32_t B, const int32_t* tokens, const int32_t* pos,
                          const int32_t* slots, void* stream);
/* One fast-AR position (DualARTransformer.forward_generate_fast, llama.py:798-816). x: bf16 [B][dim] or
 * NULL (= "hidden"); logits land in "fast_logits" [num_codebooks][max_rows][codebook_size] at [codebook_idx].
 * want_logits = 0 reproduces the discarded pass at inference.py:148-149 without the head GEMV. */
int fsb_dual_ar_fast_step(fsb_dual_ar* h, int32_t B, const void* x, int32_t codebook_idx, int32_t want_logits,
                          void* stream);
/* Named internal device buffers for parity tests: "slow_logits", "hidden", "fast_logits", "x". */
int fsb_dual_ar_buffer(fsb_dual_ar* h, const char* name, void** dptr, size_t* bytes);

/* Arguments of the fused frame (decode_one_token_ar, inference.py:96-181, + the per-frame bookkeeping of
 * decode_n_tokens, inference.py:224-233). All pointers are device pointers owned by the caller. */
typedef struct {
    int32_t batch;           /* requests in flight (rows) */
    int32_t* cur_tokens;     /* [B][C+1] in: the input column; out: the sampled frame (row 0 = main token) */
    int32_t* pos;            /* [B] input position, incremented per frame */
    const int32_t* slots;    /* [B] KV slot per request or NULL */
    int32_t* ras_window;     /* [B][10] previous main tokens (RAS, inference.py:49-51,125-144) or NULL = off */
    int32_t* out_tokens;     /* [max_frames][B][C+1] or NULL */
    int32_t max_frames;
    int32_t* frame_counter;  /* [1] frame index, incremented per frame */
    int32_t* done;           /* [B] set to 1 when the main token is <|im_end|> (inference.py:233), or NULL */
    const void* logit_bias;  /* bf16 [vocab] semantic_logit_bias (inference.py:308-320) */
    int32_t restrict_scan;   /* 1: sample only over [semantic_begin, semantic_end] U {im_end} (result-equivalent) */
    float temperature, top_p; /* already rounded to bf16 like generate() does (inference.py:301-306) */
    int32_t top_k;
    const void* uniforms;    /* optional injected bf16 draws [max_frames][B][2*vocab + (C-1)*codebook_size] */
    uint64_t seed;           /* Philox seed when uniforms == NULL */
    /* per-request sampling parameters (a batch that mixes requests, scheduler.py): device arrays [B] that take the place
     * of temperature / top_p / top_k for the request's normal draw (the RAS variant keeps its constants), or NULL */
    const float* row_temperature;
    const float* row_top_p;
    const int32_t* row_top_k;
} fsb_frame_args;

/* n_frames back-to-back frames, no host synchronisation in between (CUDA graph replay when enabled). */
int fsb_dual_ar_decode_frames(fsb_dual_ar* h, const fsb_frame_args* args, int32_t n_frames, void* stream);
/* Device watchdog: a wait inside a kernel that outlived its limit (a lost producer; never seen on a healthy device)
 * records a code and lets the launch drain instead of hanging or trapping.  This call SYNCHRONISES `stream`, reads and
 * clears the word of the current device; returns 0 and *code = 0 when nothing happened, else 1 with the message in
 * fsb_dual_ar_last_error (results of that launch are invalid).  The host mirror calls it where the reference's loop
 * synchronises anyway (the <|im_end|> test, inference.py:233). */
int fsb_dual_ar_device_error(fsb_dual_ar* h, int32_t* code, void* stream);

/* Algorithmic HBM bytes of one frame (SURVEY.md §8d shared formula): weights + KV for `ctx_sum` = sum of
 * context lengths over the batch. masked_head = 1 counts the LM-head rows a fused batch-1 frame streams when the
 * constrained-decode bias is in use (option masked_head, default on): the 8-aligned row range that covers
 * [semantic_begin, semantic_end] and <|im_end|> (4136 of 155776 rows at the S2-Pro ids), else all rows. */
double fsb_dual_ar_frame_bytes(const fsb_dual_ar* h, int32_t batch, int64_t ctx_sum, int32_t masked_head);
int32_t fsb_dual_ar_launches_per_frame(const fsb_dual_ar* h, int32_t batch);

/* ---- stage entry points (tests / profiling) */
/* out[m][n] = epi(sum_k W[n][k] * pro(x)[m][k]); pro: 0 none, 1 RMSNorm(norm_w, eps) (llama.py:995-1000);
 * epi: 0 store, 1 + res (llama.py:841-842), 2 SwiGLU with W2 (llama.py:986). All bf16. */
int fsb_gemv(const void* W, const void* W2, const void* x, const void* norm_w, const void* res, void* out,
             int32_t M, int32_t N, int32_t K, int32_t pro, int32_t epi, float eps, void* stream);
/* Many-row Linear of a batched decode step on the tcgen05 stream-K kernel (csrc/rows_gemm.cuh):
 * out[m][n] = epi(sum_k W[n][k] * x[m][k]), 1 <= M <= 128, K % 64 == 0, all bf16, W row-major [N][K] (llama.py:856-860,
 * 981-983).  epi: 0 store, 1 bf16(res + bf16(acc)) (llama.py:841-842), 2 SwiGLU with W2 (llama.py:986).
 * `scratch`: fsb_rows_gemm_scratch_bytes() bytes of device memory, zeroed once by the caller. */
size_t fsb_rows_gemm_scratch_bytes(void);
int fsb_rows_gemm(const void* W, const void* W2, const void* x, const void* res, void* out, int32_t M, int32_t N,
                  int32_t K, int32_t epi, void* scratch, size_t scratch_bytes, void* stream);
/* The same Linear with a plain store on narrow weight tiles (nw = 16/32/48/64 rows per CTA, N % nw == 0) and no k split:
 * activations as the MMA's A operand, weights as the B operand (csrc/rows_gemm.cuh, rows_gemm_narrow_kernel). */
int fsb_rows_gemm_narrow(const void* W, const void* x, void* out, int32_t M, int32_t N, int32_t K, int32_t nw, void* stream);
/* No counterpart in the reference: added for callers that cannot allocate device
 * memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The same Linear as
 * fsb_rows_gemm (epi 0 store / 1 residual only) with the stream-K partial tiles in a file-scope device array of the library
 * instead of caller scratch: it allocates nothing (no cudaMalloc / cudaFree / allocating library call).  Limit: M <= 32 rows
 * (10 MiB of partial tiles, held by every GPU that loads the library until exit); larger M, SwiGLU or a shape whose partial
 * tiles exceed 10 MiB are rejected with rc != 0 and a message in `err` (caller buffer of err_len bytes, may be NULL).
 * All calls on a device share the one array: calls that overlap on the GPU (different streams) race; calls on one stream
 * are ordered and need no clearing in between (every element read is written first by the same call). */
int fsb_rows_gemm_static(const void* W, const void* x, const void* res, void* out, int32_t M, int32_t N, int32_t K,
                         int32_t epi, void* stream, char* err, size_t err_len);
/* sample() of inference.py:80-93 on device. logits bf16 [B][V]; uniforms bf16 [B][n_variants][V] or NULL;
 * out int32 [B][2]: the index drawn with (temperature, top_p) and, if n_variants == 2, with (temp2, top_p2). */
int fsb_sample(const void* logits, const void* bias, int32_t B, int32_t V, int32_t top_k, float temperature,
               float top_p, int32_t n_variants, float temp2, float top_p2, const void* uniforms, uint64_t seed,
               int32_t seg0_lo, int32_t seg0_hi, int32_t seg1_lo, int32_t seg1_hi, int32_t* out, void* stream);

/* ------------------------------------------------------------------------------------------------
 * DAC codec decode (fish_speech/models/dac/modded_dac.py: DAC.from_indices:925 ; rvq.py decode:352)
 * ---------------------------------------------------------------------------------------------- */
typedef struct fsb_dac fsb_dac;

/* fish_speech/configs/modded_dac_vq.yaml values the decode path uses. */
typedef struct {
    int32_t latent_dim, decoder_dim, n_rates, rate0, rate1, rate2, rate3;
    int32_t n_codebooks /* residual */, codebook_size, semantic_codebook_size, codebook_dim;
    int32_t n_up, up0, up1;
    int32_t tf_layers, tf_heads, tf_dim, tf_ffn, tf_head_dim, tf_window, convnext_ratio;
    int32_t max_batch, max_frames; /* workspace sizing: clips per call, code frames per clip */
    float tf_norm_eps;
} fsb_dac_desc;

size_t fsb_dac_workspace_bytes(const fsb_dac_desc* desc);
int fsb_dac_create(const fsb_dac_desc* desc, void* workspace, size_t workspace_bytes, fsb_dac** out);
int fsb_dac_destroy(fsb_dac* h);
const char* fsb_dac_last_error(const fsb_dac* h);
/* fp32 device tensors, weight norm folded, packed [taps][Cin][N] by the host mirror (fish_speech_b200/dac.py lists
 * the name of every tensor and the reference state-dict keys it is built from). */
int fsb_dac_bind_weight(fsb_dac* h, const char* name, const void* dptr, int64_t numel);
/* DAC.from_indices: codes int64 [B][1+n_codebooks][T] (clamped IN PLACE like rvq.py:354-359) -> wav fp32 [B][hop*T]. */
int fsb_dac_decode(fsb_dac* h, void* codes, int32_t B, int32_t T, void* wav, void* stream);
/* DAC.encode (modded_dac.py:874-923: Encoder :670-709, DownsampleResidualVectorQuantize.forward rvq.py:293-318 up to the codes):
 * audio fp32 [B][n_samples], n_samples a positive multiple of the code frame (the caller pads with zeros on the right like
 * :909-911) -> codes int64 [B][1+n_codebooks][n_samples / hop].  fp32 throughout.  The encode side's shape is set with
 * fsb_dac_set_option first: "enc_dim", "enc_n_rates", "enc_rate0".."enc_rate3", "enc_tf0".."enc_tf3" (transformer layers after
 * each block), "enc_window", "enc_head_dim"; its weights are bound under the names fish_speech_b200/dac.py lists. */
int fsb_dac_encode(fsb_dac* h, const void* audio, int32_t B, int64_t n_samples, void* codes, void* stream);
/* Options: "tf32" (1 = GEMM-shaped layers on the tensor cores with TF32 operands and fp32 accumulation, the default;
 * 0 = fp32 CUDA-core path, the parity anchor). */
int fsb_dac_set_option(fsb_dac* h, const char* name, int64_t value);
/* internal activation buffers "buf0".."buf3" (tests / profiling) */
int fsb_dac_stage(fsb_dac* h, const char* name, void** dptr, size_t* bytes);
/* algorithmic FLOPs of one decode call (GEMM-shaped work only) */
double fsb_dac_flops(const fsb_dac* h, int32_t B, int32_t T);

/* ---- stage entry point of the tcgen05 implicit-GEMM conv (csrc/tc_conv.cuh): tests / profiling.
 *   out[b][t][n] = epi( sum_j sum_k A[b][t - (taps-1-j)*dil][k] * W[j][n][k] ),  rows t < 0 read as zero
 * = CausalConvNet / CausalTransConvNet (modded_dac.py:521-580) in the channels-last packing of fish_speech_b200/dac.py, and
 * nn.Linear (taps = 1, W = the state dict's [N][K]).  epi (fp32): [round to the 16-bit type], + bias[n % bias_mod], GELU,
 * * gamma[n], + residual, tanh | SwiGLU over interleaved (gate, up) columns; out32 = raw fp32, out16 = 16-bit copy with the
 * consumer's Snake1d (alpha[n % alpha_mod]) applied when alpha != NULL. */
typedef struct {
    const void* A16;   /* [B][T][Cin] fp16 (f16 = 1) or bf16 (f16 = 0) */
    const void* W16;   /* [taps][N][Cin] same type */
    int32_t B, T, Cin, N, taps, dil;
    int32_t f16, use_simt; /* use_simt = 1: the CUDA-core twin kernel (also chosen automatically when Cin % 8 != 0) */
    const float* bias; int32_t bias_mod;
    const float* gamma;
    const float* res32; const void* res16; int32_t ldres;
    const float* alpha; int32_t alpha_mod;
    int32_t act;       /* 0 none, 1 GELU, 2 tanh, 3 SwiGLU (out16 width N / 2) */
    int32_t round_lin;
    void* out16; int32_t ld16;
    float* out32; int32_t ld32;
    void* out16r; int32_t ld16r; /* optional raw 16-bit copy (no Snake): the residual stream of the decoder blocks */
} fsb_tc_conv_args;
int fsb_tc_conv(const fsb_tc_conv_args* a, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FISH_B200_H */
