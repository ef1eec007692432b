// Fused attention of the denoise step: the S suffix tokens of one sample (x 8 query heads, MQA) against the cached
// prefix K/V plus the suffix K/V written this step. One launch replaces the scores GEMM + masked softmax + P.V GEMM
// of the per-op path (reference: eager_attention_forward, [mirror] src/model/mvla/mllm.py:265-317, with the
// block-causal suffix mask of modeling_mvla.py:686-692,817-821).
//
// Rounding points are the reference's under autocast(bf16): scores = rnd(rnd(q.k) * 256^-0.5) in bf16, softmax in
// fp32 over the masked row (finite -2.3819763e38), probabilities rounded to bf16, P.V accumulated in fp32 and rounded.
//
// Work split: grid = (head groups, samples). A CTA owns R = S * (8 / groups) query rows (<= 48 = three m16 tiles) and
// streams K once for all of them (MQA: every head reads the same K/V). Tensor work is mma.sync m16n8k16 (the rows are
// far too few for a 128-row tcgen05 tile); K and V^T are read straight from HBM/L2 as B fragments with 16-byte loads
// (the k index is permuted identically on both operands so a lane's 8 consecutive k values feed two MMAs).
#include "attn_decode.cuh"

namespace pi0 {

namespace {

constexpr int kRMax = 48;       // query rows per CTA (3 m16 tiles)

__global__ void __launch_bounds__(256) decode_attn_kernel(const bf16* q, const bf16* kc, const bf16* vt, bf16* att,
                                                          const uint8_t* __restrict__ pad, const int* __restrict__ cum,
                                                          int S, int P, int kv_pitch, int vt_pitch, int hpg, int kp,
                                                          int sc_pitch) {
  pdl_trigger();
  extern __shared__ __align__(16) uint8_t smem_raw[];
  pdl_wait();
  decode_attn_body<3>(smem_raw, q, kc, vt, att, pad, cum, blockIdx.y, blockIdx.x * hpg, S, P, kv_pitch, vt_pitch, hpg,
                      kp, sc_pitch);
}

}  // namespace

// heads per CTA: as many as fit 48 rows, fewer when the batch alone cannot fill the SMs
int decode_attention(const void* q, const void* kcache, const void* vtcache, void* att, const uint8_t* pad,
                     const int* cum, int B, int S, int P, int kv_pitch, int vt_pitch, int num_sms, cudaStream_t s) {
  PI0_REQUIRE(S >= 1 && S <= kRMax, "decode_attention: S=%d outside [1,%d]", S, kRMax);
  const int n_keys = P + S;
  const int kp = (n_keys + 31) / 32 * 32;
  PI0_REQUIRE(kp <= 640, "decode_attention: %d keys exceed the 640 the kernel holds per row", n_keys);
  PI0_REQUIRE(vt_pitch % 8 == 0 && vt_pitch >= n_keys && kv_pitch >= n_keys, "decode_attention: bad cache pitches");
  int hpg = 8;
  while (S * hpg > kRMax) hpg >>= 1;
  while (hpg > 1 && B * (8 / hpg) < num_sms / 2) hpg >>= 1;
  const int sc_pitch = kp + 16;  // fp32 elements; (kp + 16) * 4 B = 64 mod 128
  const size_t smem = attn_smem_bytes(3, sc_pitch);
  static PerDeviceOnce smem_once;
  if (smem_once.need_more(smem)) {
    PI0_CHECK_CUDA(cudaFuncSetAttribute(decode_attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    smem_once.raise(smem);
  }
  PI0_CHECK_CUDA(launch_pdl(decode_attn_kernel, dim3(8 / hpg, B), dim3(256), smem, s, 1, (const bf16*)q,
                            (const bf16*)kcache, (const bf16*)vtcache, (bf16*)att, pad, cum, S, P, kv_pitch, vt_pitch,
                            hpg, kp, sc_pitch));
  ++g_launch_count;
  return 0;
}

}  // namespace pi0
