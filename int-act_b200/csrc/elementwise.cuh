// Launchers for the HBM-bound kernels (see elementwise.cu). T = bf16 (product) or float (check mode).
#pragma once
#include "gemm.cuh"

namespace pi0 {

template <typename T>
int im2col_patch(const float* img, T* out, int B, int kpad, float scale, float shift, cudaStream_t s);
template <typename T>
int layernorm(const T* x, T* y, const float* w, const float* b, int M, int D, float eps, cudaStream_t s);
template <typename T>
int rmsnorm(const T* x, T* y, const float* g, int M, int D, float eps, int rpb_in, int rpb_out, int off,
            cudaStream_t s);
template <typename T>
int softmax_rows(T* sc, int64_t rows_total, int rows_per_b, int64_t pitch, int n_keys, int n_prefix,
                 const uint8_t* pad, int pad_pitch, const int* cum, int heads, cudaStream_t s);
template <typename T>
int embed_tokens(const int64_t* tok, const T* table, T* out, int B, int L, int P, int row_off, int D, int64_t vocab,
                 float scale, cudaStream_t s);
template <typename T>
int cast_f32(const float* in, T* out, int64_t n, cudaStream_t s);

// fused suffix attention of the denoise step (attn_decode.cu), bf16 only
int decode_attention(const void* q, const void* kcache, const void* vtcache, void* att, const uint8_t* pad,
                     const int* cum, int B, int S, int P, int kv_pitch, int vt_pitch, int num_sms, cudaStream_t s);

// the same attention as two wide launches (attn_split.cu); scratch = bf16 [B*8*S, roundup(P+S, 32)]
bool decode_attention_split_supported(int S, int P);
int decode_attention_split(const void* q, const void* kcache, const void* vtcache, void* att, void* scratch,
                           const uint8_t* pad, const int* cum, int B, int S, int P, int kv_pitch, int vt_pitch,
                           cudaStream_t s);

// one CTA per sample with K / V^T staged in shared memory (attn_batch.cu): the rollout-batch path
bool decode_attention_batch_supported(int S, int P);
int decode_attention_batch(const void* q, const void* kcache, const void* vtcache, void* att, const uint8_t* pad,
                           const int* cum, int B, int S, int P, int kv_pitch, int vt_pitch, cudaStream_t s);

// fused SigLIP self-attention (attn_siglip.cu), bf16 only: fused q|k|v [B*256, 3456] -> att [B*256, 1152]
int siglip_attention(const void* qkv, void* att, int B, int num_sms, cudaStream_t s);   // dispatcher (attn_siglip.cu)
int siglip_attention_mma(const void* qkv, void* att, int B, cudaStream_t s);
int siglip_attention_tc(const void* qkv, void* att, int B, int num_sms, cudaStream_t s);
bool siglip_attention_use_tc();
// fused prefix-prefill attention on tcgen05 (attn_prefill.cu); q / att [B*T][8*256], P <= 336 prefix keys
// long key ranges (<= 1024 keys), prefix pass (n_suffix == 0) or denoise-step suffix attention (attn_long.cu)
bool attention_long_supported(int T, int P, int n_suffix);
int attention_long(const void* q, const void* kcache, const void* vtcache, void* att, const uint8_t* pad, const int* cum, int B,
                   int T, int P, int n_suffix, int kv_pitch, int vt_pitch, int num_sms, cudaStream_t s);
bool prefill_attention_supported(int T, int P);
int prefill_attention(const void* q, const void* kcache, const void* vtcache, void* att, const uint8_t* pad, int B, int T,
                      int P, int kv_pitch, int vt_pitch, int num_sms, cudaStream_t s);

// weight-streaming mma.sync GEMMs for M <= 16 suffix rows (gemm_skinny.cu), bf16 only
bool skinny_supported(int M);
int skinny_qkv_rope(const void* x, int64_t ldx, const float* gain, const void* wqkv, int M, int K, void* q,
                    const float* rope_cos, const float* rope_sin, const int* pos_ids, void* kcache, void* vtcache, int S,
                    int P, int64_t kv_pitch, int64_t vt_pitch, cudaStream_t s);
int skinny_linear_res(const void* A, int64_t lda, const void* W, int M, int N, int K, const void* res, int64_t ldr,
                      void* out, int64_t ldc, cudaStream_t s);
int skinny_gateup_geglu(const void* y, int64_t ldy, const float* gain, const void* wgu, int M, int K, int inter, void* out,
                        int64_t ldc, cudaStream_t s);

int position_ids(const uint8_t* pad, int B, int P, int S, int* prefix_pos, int* suffix_pos, cudaStream_t s);
int build_pad(const uint8_t* img_masks, const uint8_t* lang_masks, uint8_t* pad, int B, int ncam, int L,
              cudaStream_t s);

}  // namespace pi0
