// Denoise-step attention for rollout batches: ONE CTA per sample, K and V^T of the sample staged in shared memory.
//
// At batch 64 the K / V^T caches of a layer are 21.8 MB, far beyond what the registers-only kernels (attn_decode.cu,
// attn_split.cu) keep in flight: they chain several dependent rounds of 16-byte loads per warp and ran at 34 - 45 us
// per layer against 3.4 us of HBM time. Here the whole K tile of a sample (P+S rows x 512 B, <= 176 KB) is brought in
// with one burst of cp.async (all of it in flight at once), the 8 heads x S suffix rows (<= 48 rows, MQA: one K for
// all heads) are scored out of shared memory with mma.sync, and while the masked softmax runs the V^T tile of the
// sample streams into the space K and Q occupied. Rounding points are the reference's under autocast(bf16)
// (eager_attention_forward, [mirror] src/model/mvla/mllm.py:265-317; suffix mask modeling_mvla.py:686-692,817-821):
// scores = rnd(rnd(q.k) * 256^-0.5), softmax in fp32 over the masked row (finite -2.3819763e38), probabilities rounded
// to bf16, P.V accumulated in fp32 and rounded.
#include "attn_batch.cuh"

namespace pi0 {

namespace {

using namespace battn;

__global__ void __launch_bounds__(256, 1) batch_attn_kernel(const BatchAttnParams p) {
  pdl_trigger();
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[2];
  if (threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_barrier_init();
  }
  __syncthreads();
  pdl_wait();
  uint32_t par = 0;
  batch_attn_unit(smem, bars, par, p, (int)blockIdx.y, (int)blockIdx.x * p.hpg, [] { __syncthreads(); });
}

}  // namespace


bool decode_attention_batch_supported(int S, int P) { return S >= 1 && S <= 6 && P + S <= kKeysMax; }

int decode_attention_batch(const void* q, const void* kcache, const void* vtcache, void* att, const uint8_t* pad,
                           const int* cum, int B, int S, int P, int kv_pitch, int vt_pitch, cudaStream_t s) {
  PI0_REQUIRE(decode_attention_batch_supported(S, P), "decode_attention_batch: S=%d P=%d unsupported", S, P);
  PI0_REQUIRE(vt_pitch % 8 == 0 && vt_pitch >= P + S && kv_pitch >= P + S, "decode_attention_batch: bad cache pitches");
  BatchAttnParams p;
  p.q = (const bf16*)q; p.kc = (const bf16*)kcache; p.vt = (const bf16*)vtcache; p.att = (bf16*)att;
  p.pad = pad; p.cum = cum; p.S = S; p.P = P; p.kv_pitch = kv_pitch; p.vt_pitch = vt_pitch;
  p.kp = (P + S + 31) / 32 * 32;
  const int vpitch = vt_pitch * 2;
  const size_t kq = (size_t)((P + S + 15) / 16 * 16) * kRowB + (size_t)kRows * kRowB;
  const size_t vbytes = (size_t)256 * vpitch + 64;  // + the tail the last row's final chunk may touch
  const size_t smem = (kq > vbytes ? kq : vbytes) + (size_t)kRows * kSPitch;
  PI0_REQUIRE(vbytes <= kq, "decode_attention_batch: V^T tile (%zu B) exceeds the K|Q region (%zu B)", vbytes, kq);
  static PerDeviceOnce smem_once;
  if (smem_once.need_more(smem)) {
    PI0_CHECK_CUDA(cudaFuncSetAttribute(batch_attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    smem_once.raise(smem);
  }
  p.hpg = B * 2 <= 160 ? 4 : 8;  // two CTAs per sample while they still fit one wave
  PI0_CHECK_CUDA(launch_pdl(batch_attn_kernel, dim3(8 / p.hpg, B), dim3(256), smem, s, 1, p));
  ++g_launch_count;
  return 0;
}

}  // namespace pi0
