/*
 * pi0b200.h - C ABI of the B200-native Pi0 `infer_action` hot path (libpi0b200.so).
 *
 * The reference (chenbadream/INT-ACT) is pure Python and has no FFI for this path; the entry points
 * below are what a ctypes binding on the reference side binds in place of the PyTorch modules it
 * calls today. Each entry cites the reference code it replaces (paths relative to the reference root;
 * "[mirror]" = the in-tree derivative of the missing lerobot Pi0 source).
 *
 * Conventions
 *   - extern "C", plain pointers and sizes only. Every device buffer (weights, workspace, inputs,
 *     outputs) is owned by the caller (PyTorch); the library never allocates caller-visible memory.
 *   - every function returns 0 on success and a negative code on failure; the message is available
 *     from pi0_last_error() (thread-local).
 *   - every launch goes to the cudaStream_t passed as `stream` (void* here to keep the header free of
 *     CUDA includes). No entry point synchronises the device or allocates device memory, so
 *     pi0_infer / pi0_denoise_loop can be captured in a CUDA graph.
 *   - dtype: 0 = bf16 activations + weights (tcgen05 product path), 1 = fp32 check mode (SIMT).
 *   - there is no CPU fallback anywhere behind this ABI.
 */
#ifndef PI0B200_H_
#define PI0B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PI0_ABI_VERSION 4

#define PI0_SIGLIP_LAYERS 27
#define PI0_GEMMA_LAYERS 18

/* ---- model geometry (config keys of config/models/pi0_finetune_bridge.json:34-48) -------------- */
typedef struct pi0_model_cfg {
  int32_t n_cams;         /* len(image_features) + empty cameras actually fed                      */
  int32_t lang_len;       /* tokenizer_max_length                                                  */
  int32_t horizon;        /* n_action_steps                                                        */
  int32_t max_state_dim;  /* 32                                                                    */
  int32_t max_action_dim; /* 32                                                                    */
  int32_t num_steps;      /* flow-matching Euler steps (10)                                        */
  int32_t dtype;          /* 0 = bf16, 1 = fp32 check mode                                         */
  int32_t gemm_impl;      /* 0 = tcgen05 (bf16) / SIMT (fp32); 1 = force SIMT (debug cross-check)  */
  int32_t max_batch;      /* workspace is carved for this many observations                        */
  int32_t num_sms;        /* SM count of the device (persistent-grid size)                         */
  int32_t skip_dead_work; /* 1 = skip the last prefill layer's attention/MLP (outputs discarded)  */
  int32_t n_siglip_layers;/* <= PI0_SIGLIP_LAYERS (27 in the shipped model; fewer in depth-reduced tests) */
  int32_t n_gemma_layers; /* <= PI0_GEMMA_LAYERS, shared by Gemma-2B and the action expert (18)     */
  int32_t images_pre_normalized; /* 1 = skip prepare_images' img*2-1                 */
  int32_t n_euler_iters;  /* iterations of `while time >= -dt/2` with fp32-accumulated time (0 = num_steps) */
  int32_t reserved;
} pi0_model_cfg;

/* ---- weights: packed by the host loader (int-act_b200/weights.py) into device tensors ----------
 * W* matrices are [out, in] row-major in the activation dtype; biases and norm gains are fp32.    */
typedef struct pi0_siglip_layer_w {
  const float* ln1_w; const float* ln1_b;
  const void* wqkv;   const float* bqkv;   /* [3456, 1152] q|k|v rows                               */
  const void* wo;     const float* bo;     /* [1152, 1152]                                          */
  const float* ln2_w; const float* ln2_b;
  const void* wfc1;   const float* bfc1;   /* [4304, 1152]                                          */
  const void* wfc2;   const float* bfc2;   /* [1152, 4304]                                          */
} pi0_siglip_layer_w;

typedef struct pi0_gemma_layer_w {
  const float* g1;     /* 1 + input_layernorm.weight, fp32 [hidden]                                 */
  const void* wqkv;    /* [8*256 + 256 + 256, hidden] q|k|v rows                                    */
  const void* wo;      /* [hidden, 2048]                                                            */
  const float* g2;     /* 1 + post_attention_layernorm.weight                                       */
  const void* wgu;     /* [2*inter, hidden]: per 256 rows, 128 gate rows then the 128 matching up   */
  const void* wdown;   /* [hidden, inter]                                                           */
} pi0_gemma_layer_w;

typedef struct pi0_weights {
  /* SigLIP-So400m/14 + projector ([mirror] src/model/mvla/mllm.py:98-110,182-187) */
  const void* patch_w;  const float* patch_b;   /* [1152, 592] (3*14*14 = 588 zero-padded to 592)  */
  const void* pos_emb;                          /* [256, 1152] activation dtype                     */
  pi0_siglip_layer_w sig[PI0_SIGLIP_LAYERS];
  const float* post_ln_w; const float* post_ln_b;
  const void* proj_w;   const float* proj_b;    /* [2048, 1152]                                     */
  /* Gemma-2B ([mirror] mllm.py:86-97,193-249) */
  const void* embed;                            /* [257152, 2048]                                   */
  int64_t vocab;
  pi0_gemma_layer_w gem[PI0_GEMMA_LAYERS];
  /* action expert, Gemma-300M ([mirror] src/model/mvla/expert.py:70-93) */
  pi0_gemma_layer_w exp[PI0_GEMMA_LAYERS];
  const float* exp_norm_g;                      /* 1 + model.norm.weight                            */
  /* projections ([mirror] src/model/mvla/modeling_mvla.py:538-543) */
  const void* state_w;      const float* state_b;       /* [1024, 32]                               */
  const void* action_in_w;  const float* action_in_b;   /* [1024, 32]                               */
  const void* time_in_wa;                                /* [1024, 1024] action half of mlp_in      */
  const float* time_bias;                                /* [num_steps, 1024] b + W_t . emb(t_k)    */
  const void* time_out_w;   const float* time_out_b;    /* [1024, 1024]                             */
  const void* action_out_w; const float* action_out_b;  /* [32, 1024]                               */
  /* RoPE table ([mirror] mllm.py:30-53): cos/sin of pos / 10000^(2i/256), fp32 [max_pos, 128] */
  const float* rope_cos; const float* rope_sin;
  int32_t max_pos; int32_t reserved;
} pi0_weights;

typedef struct pi0_ctx pi0_ctx;

int pi0_abi_version(void);
const char* pi0_last_error(void);

/* bytes of device workspace (activations + KV cache) needed for cfg->max_batch observations */
size_t pi0_workspace_bytes(const pi0_model_cfg* cfg);

/* Host-side context: records geometry, weight pointers and the carved workspace. No device work. */
int pi0_ctx_create(const pi0_model_cfg* cfg, const pi0_weights* w, void* workspace, size_t workspace_bytes,
                   pi0_ctx** out);
void pi0_ctx_destroy(pi0_ctx* ctx);

/* SigLIP + projector for one camera: images f32 [B,3,224,224] -> prefix embedding rows
 * [b, cam*256 .. cam*256+255]. `scale`/`shift` fuse prepare_images' img*2-1.
 * Replaces PaliGemmaModel.embed_image + the two sqrt(2048) scalings
 * ([mirror] mllm.py:182-187; modeling_mvla.py:366-367,596-601). */
int pi0_siglip_forward(pi0_ctx* ctx, int B, int cam, const float* images, float scale, float shift, void* stream);

/* Token embedding gather * sqrt(2048) into prefix rows [b, ncam*256 ..]
 * ([mirror] mllm.py:189-191; modeling_mvla.py:612-616). tokens: int64 [B, lang_len] on device. */
int pi0_embed_language(pi0_ctx* ctx, int B, const int64_t* tokens, void* stream);

/* Gemma-2B prefill over the prefix: builds pad mask / position ids, runs the 18 layers and writes the
 * KV cache in place. img_masks u8 [n_cams, B], lang_masks u8 [B, lang_len] on device.
 * Replaces make_att_2d_masks + cumsum position ids + PaliGemmaWithExpertModel.forward(fill_kv_cache=True)
 * ([mirror] modeling_mvla.py:98-128,765-766; mllm.py:193-249,265-317). */
int pi0_prefill(pi0_ctx* ctx, int B, const uint8_t* img_masks, const uint8_t* lang_masks, void* stream);

/* The flow-matching Euler loop: x_t f32 [B, horizon, max_action_dim] in/out (noise in, actions out),
 * state f32 [B, max_state_dim]. Replaces embed_suffix + denoise_step + the while loop
 * ([mirror] modeling_mvla.py:643-694,787-839). */
int pi0_denoise_loop(pi0_ctx* ctx, int B, const float* state, float* x_t, void* stream);

/* sample_actions end to end = siglip (per camera) + embed_language + prefill + denoise_loop
 * ([mirror] modeling_mvla.py:753-804). images: n_cams device pointers (host array), each f32 [B,3,224,224]. */
int pi0_infer(pi0_ctx* ctx, int B, const float* const* images, float img_scale, float img_shift,
              const uint8_t* img_masks, const int64_t* tokens, const uint8_t* lang_masks, const float* state,
              float* x_t, void* stream);

/* Introspection for the parity tests (device pointers into the workspace; element type = cfg dtype) */
const void* pi0_debug_buffer(pi0_ctx* ctx, const char* name, int64_t* n_elems);

/* Per-launch CUDA-event timing for bench.py's roofline: enable, run pi0_infer once (NOT under graph capture), then
 * pi0_profile_report writes {"label": [total_ms, launches], ...} (JSON) into buf. Reporting synchronises on the
 * recorded events; the product path leaves profiling off. */
int pi0_profile(pi0_ctx* ctx, int enable);
int pi0_profile_report(pi0_ctx* ctx, char* buf, size_t buf_bytes);

/* Which kernel variant each phase of pi0_infer takes at batch B with this ctx, as one JSON object: {"gemm", "siglip_attention",
 * "prefill_attention", "denoise", "denoise_attention", ...}. The reference has one eager path
 * (config/models/pi0_finetune_bridge.json:50 "attention_implementation": "eager"); here shape limits select between
 * fused and per-op kernels (P <= 336 for the fused prefix attention, M = B*S <= 16 for the persistent denoise kernel, ...):
 * this call makes every such fallback visible. Host only, no launch. */
int pi0_path_report(pi0_ctx* ctx, int B, char* buf, size_t buf_bytes);

/* Bounds check of our own (compute-sanitizer is not available on the GPU pool): with PI0_GUARD=1 in the environment when
 * pi0_workspace_bytes / pi0_ctx_create run, every workspace buffer is followed by a guard zone filled with 0xA5;
 * this call synchronises the device and verifies the zones. 0 = intact, 1 = overwritten (description in buf), < 0 = error. */
int pi0_debug_check_guards(pi0_ctx* ctx, char* buf, size_t buf_bytes);

/* kernels launched by this library since the last call with reset != 0 (bench "gpu_launches") */
long long pi0_launch_count(int reset);

/* ---- single-kernel entry points for unit tests ------------------------------------------------- */
typedef struct pi0_gemm_desc {
  const void* A; int64_t lda, a_z0, a_z1;
  const void* W; int64_t ldw, w_z0, w_z1;
  int32_t M, N, K, nz0, nz1;
  int32_t epi_mode;  /* 0 store, 1 geglu, 2 rope(qkv), 3 euler */
  int32_t bn;        /* 32 / 64 / 128 / 256; 0 = pick from the shape */
  int32_t act;       /* 0 none, 1 gelu_tanh, 2 silu */
  void* out; int64_t ldc, out_z0, out_z1;
  const float* bias;
  const void* res; int64_t ldr; int32_t res_row_mod;
  float alpha, post_div, post_mul;
  int32_t rpb_in, rpb_out, row_off;
  void* vt; int32_t vt_col0, vt_hd, vt_heads, vt_rpb, vt_toff; int64_t vt_pitch;
  const float* rope_cos; const float* rope_sin; const int32_t* pos_ids;
  void* kcache; int32_t n_q_heads; int64_t kv_pitch;
  float dt;
  int32_t splits;    /* split-K factor = thread-block cluster size, 1..8; 0 = pick from the shape */
  /* ABI 4. Producer side (epi_mode 0): rowsq_out[row * rowsq_ld + col / 32] = sum of squares of the stored values of every
   * 32-column chunk. Consumer side (epi_mode 1 / 2, bn 256, bf16): GemmaRMSNorm fused into the A operand - A is the RAW
   * residual stream, a <- rnd((a * rstd[row]) * norm_gain[k]) in shared memory before the MMA, rstd[row] = rsqrt(sum of
   * norm_rowsq[row * norm_slots + 0 .. norm_slots) / K + norm_eps). Replaces GemmaRMSNorm + Linear
   * ([mirror] src/model/mvla/mllm.py:30-53,206-215). */
  float* rowsq_out; int32_t rowsq_ld;
  const float* norm_gain; const float* norm_rowsq; int32_t norm_slots; float norm_eps;
} pi0_gemm_desc;

/* C[z] = A[z] . W[z]^T with a fused epilogue. impl: 0 = default for dtype, 1 = SIMT, 2 = tcgen05 single-CTA kernels
   only, 3 = force the tcgen05 CTA-pair (cta_group::2) kernel. */
int pi0_gemm(const pi0_gemm_desc* d, int dtype, int impl, int num_sms, void* stream);
/* The GEMM of the fp32 regime on the tensor cores (csrc/gemm_x3.cu): fp32 A, W, bias / residual / out; every operand is
 * split into three bf16 pieces (hi + mid + lo = the fp32 value exactly), the six piece products >= 2^-16 are one bf16 tcgen05
 * GEMM over K' = 6 K with fp32 accumulation and fp32 epilogues. Replaces torch.matmul / nn.Linear in fp32 (the reference with
 * `--use_bf16 False`, slurms/eval_scripts/simpler/ev_pi0_bridge_simpler.sh:99). scratch_a / scratch_w: device buffers of
 * pi0_gemm_f32x3_scratch_bytes(nz0 * nz1, M | N, K) bytes; scratch_p: optional fp32 workspace (any size, may be null) that
 * lets skinny problems split K over the idle SMs (partials summed in split order by a second launch: deterministic). */
int pi0_gemm_f32x3(const pi0_gemm_desc* d, void* scratch_a, size_t bytes_a, void* scratch_w, size_t bytes_w, void* scratch_p,
                   size_t bytes_p, int num_sms, void* stream);
size_t pi0_gemm_f32x3_scratch_bytes(int64_t nz, int64_t rows, int K);

int pi0_rmsnorm(const void* x, void* y, const float* g, int M, int D, int dtype, void* stream);
int pi0_layernorm(const void* x, void* y, const float* w, const float* b, int M, int D, int dtype, void* stream);
int pi0_softmax(void* scores, int64_t rows_total, int rows_per_b, int64_t pitch, int n_keys, int n_prefix,
                const uint8_t* pad, int pad_pitch, const int32_t* cum, int heads, int dtype, void* stream);
/* Fused suffix attention of one denoise step (bf16): q [B*S, 8*256], K cache [B, kv_pitch, 256], V^T cache
 * [B, 256, vt_pitch], pad u8 [B, P], cum i32 [S] (cumsum of the suffix att mask) -> att [B*S, 8*256].
 * Replaces eager_attention_forward for the suffix ([mirror] src/model/mvla/mllm.py:265-317). */
int pi0_decode_attention(const void* q, const void* kcache, const void* vtcache, void* att, const uint8_t* pad,
                         const int32_t* cum, int B, int S, int P, int kv_pitch, int vt_pitch, int num_sms, void* stream);
/* The same attention as two wide launches for rollout batches (attn_split.cu): scores of all heads per 16-key block,
 * then masked softmax + P.V per (sample, head, 64 dims). scratch: bf16 [B*8*S, roundup(P+S, 32)]. S <= 6, P+S <= 384.
 * Bit-identical to pi0_decode_attention. */
int pi0_decode_attention_split(const void* q, const void* kcache, const void* vtcache, void* att, void* scratch,
                               const uint8_t* pad, const int32_t* cum, int B, int S, int P, int kv_pitch, int vt_pitch,
                               void* stream);
/* pi0_decode_attention_split without a scratch argument and without any allocation: the score scratch lives in a 4 MiB
 * static __device__ array inside the library (pi0_decode_attention_split_static_max_bytes()), for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it). No counterpart in the reference.
 * Needs B*8*S*roundup(P+S, 32)*2 bytes <= the maximum, else -2 with a message
 * (bridge shapes: up to batch 145). Every call on a GPU shares the one array and nothing is cleared (the scores kernel
 * writes every element the second kernel reads): calls that overlap on the same GPU - two streams, two threads - RACE;
 * serialise them or use pi0_decode_attention_split with caller-owned scratch. Output bit-identical to that entry. */
int pi0_decode_attention_split_static(const void* q, const void* kcache, const void* vtcache, void* att, const uint8_t* pad,
                                      const int32_t* cum, int B, int S, int P, int kv_pitch, int vt_pitch, void* stream);
long long pi0_decode_attention_split_static_max_bytes(void);
/* The same attention with one CTA per sample and the sample's K / V^T tiles staged in shared memory (attn_batch.cu):
 * the rollout-batch path (B >= 4). S <= 6, P+S <= 352. */
int pi0_decode_attention_batch(const void* q, const void* kcache, const void* vtcache, void* att, const uint8_t* pad,
                               const int32_t* cum, int B, int S, int P, int kv_pitch, int vt_pitch, void* stream);
/* Fused SigLIP self-attention of one encoder layer (bf16): qkv [B*256, 3456] (q | k | v per token, 16 heads x 72 each,
 * exactly what the fused q|k|v Linear writes) -> att [B*256, 1152]; scores = rnd(rnd(q.k) * 72^-0.5), fp32 softmax, bf16
 * probabilities.
 * Replaces HF SiglipAttention's matmul / softmax / matmul ([mirror] src/model/mvla/mllm.py:182-187). */
int pi0_siglip_attention(const void* qkv, void* att, int B, void* stream);
/* Fused prefix-prefill attention on tcgen05 (bf16, attn_prefill.cu): q / att [B*T, 8*256] (8 query heads per token against
 * one K / V, MQA), kcache [B, kv_pitch, 256], vtcache [B, 256, vt_pitch], pad [B, P] (key j visible iff pad[b][j]);
 * scores = rnd(rnd(q.k) * 256^-0.5), finite mask value, fp32 softmax, bf16 probabilities, P <= 336, P % 8 == 0.
 * Replaces eager_attention_forward of the prefix pass ([mirror] src/model/mvla/mllm.py:265-317). */
int pi0_prefill_attention(const void* q, const void* kcache, const void* vtcache, void* att, const uint8_t* pad, int B,
                          int T, int P, int kv_pitch, int vt_pitch, int num_sms, void* stream);
/* The same attention for LONG key ranges (csrc/attn_long.cu, tcgen05, keys in blocks of 256, two passes: row max / sum, then
 * probabilities and P.V): up to 1024 keys. n_suffix == 0: prefix pass, T query tokens against the P prefix keys (pad mask).
 * n_suffix > 0: denoise-step suffix attention, T == n_suffix suffix tokens against P + n_suffix keys (pad mask + cumsum rule,
 * cum i32 [n_suffix]). Replaces eager_attention_forward ([mirror] src/model/mvla/mllm.py:265-317) for the LIBERO (P = 584,
 * 17 suffix tokens) and 3-camera (P = 840) shapes that exceed pi0_prefill_attention / pi0_decode_attention_batch. P % 8 == 0. */
int pi0_attention_long(const void* q, const void* kcache, const void* vtcache, void* att, const uint8_t* pad, const int32_t* cum,
                       int B, int T, int P, int n_suffix, int kv_pitch, int vt_pitch, int num_sms, void* stream);
int pi0_position_ids(const uint8_t* pad, int B, int P, int S, int32_t* prefix_pos, int32_t* suffix_pos, void* stream);


/* ---- environment adapters on the GPU ------------------------------------------------------------
 * Replace, for a whole rollout batch, the per-sample host code of BridgeSimplerBatchAdapter
 * (/root/reference/src/experiments/env_adapters/simplerMS3.py:40-132,153-172) and BridgeSimplerAdapter.postprocess
 * (src/experiments/env_adapters/simpler.py:96-139). */
/* images u8 [B,H,W,3] -> out f32 [B,3,out_h,out_w] = ((resize_u8(img) * rescale) - mean) / std. The resize is torch's
 * uint8 bilinear (align_corners=False): separable, int16 fixed-point weights, uint8 rounding after each pass; the
 * caller passes the per-axis tables idx i32 [out,2], w i16 [out,2] and their precision (simplerMS3.py:50-63). */
int pi0_adapter_resize_normalize(const uint8_t* images, float* out, int B, int H, int W, int out_h, int out_w,
                                 const int32_t* x_idx, const int16_t* x_w, int x_prec, const int32_t* y_idx,
                                 const int16_t* y_w, int y_prec, float rescale, float mean, float stdv, void* stream);
/* The single-observation adapter's resize: cv2.resize(img, size, interpolation=cv2.INTER_LANCZOS4) on uint8
 * (/root/reference/src/experiments/env_adapters/simpler.py:47-51) + the same normalisation. 8-tap separable Lanczos,
 * 11-bit fixed-point coefficients; the caller passes per-axis tables ofs i32 [out] (floor of the source coordinate) and
 * coef i16 [out,8]. Bit-exact with OpenCV's generic 8U path. */
int pi0_adapter_lanczos4_normalize(const uint8_t* images, float* out, int B, int H, int W, int out_h, int out_w,
                                   const int32_t* x_ofs, const int16_t* x_coef, const int32_t* y_ofs,
                                   const int16_t* y_coef, float rescale, float mean, float stdv, void* stream);
/* eef [B,8] (f32 or f64: pos | quat wxyz | gripper) -> state f32 [B,7]: quat -> R -> R.default_rot^T -> euler sxyz, then
 * norm_type 0: bound with (p01, p99), clipped to [-1,1]; 1: gaussian with (mean, std). round_bf16 rounds f64 -> bf16
 * like torch.as_tensor(proprio, dtype=bfloat16) (simplerMS3.py:66-88,153-165). */
int pi0_adapter_proprio(const void* eef, int eef_is_f64, float* state, int B, const double* stat_a, const double* stat_b,
                        int norm_type, int round_bf16, void* stream);
/* actions f32 [N,7] -> env actions f64 [N,7]: dims 0-5 denormalised, rpy -> axis*angle, gripper 2*(g>0.5)-1
 * (simplerMS3.py:95-132,166-172). stat_a / stat_b hold 6 entries used. */
int pi0_adapter_postprocess(const float* actions, double* out, int N, const double* stat_a, const double* stat_b,
                            int norm_type, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PI0B200_H_ */
