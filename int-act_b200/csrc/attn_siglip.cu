// Fused SigLIP self-attention (bidirectional, 256 tokens, 16 heads x 72 dims): scores + softmax + P.V in ONE launch per
// encoder layer, replacing the scores GEMM -> softmax -> P.V GEMM chain (3 launches and ~0.5 GB of score traffic per
// layer at batch 64). Reference: HF SiglipAttention as called by get_image_features ([mirror]
// /root/reference/src/model/mvla/mllm.py:182-187): scale 72^-0.5, no mask.
//
// One CTA = (128 query rows, head, image); 8 warps x 16 rows. K, V (256 x 72 each, row-major exactly as the fused q|k|v
// GEMM wrote them - no transposed copy of V exists any more) and the CTA's Q rows are staged in shared memory with
// cp.async; the B fragments of P.V come out of the row-major V tile through ldmatrix.trans; the 16 x 256 score tile of a warp lives in registers (mma.sync m16n8k16, k padded
// 72 -> 80 with zeros), the softmax runs on the accumulator fragments (row statistics over the 4 lanes of a quad), and the
// bf16 probabilities are re-used directly as the A fragments of P.V (accumulator layout = A-operand layout).
// Rounding points of the reference under autocast(bf16): scores = rnd(rnd(q.k) * 72^-0.5), softmax in fp32,
// probabilities rounded to bf16, P.V accumulated in fp32 and rounded.
#include "attn_decode.cuh"

namespace pi0 {

namespace {

constexpr int kT = 256;          // tokens per image
constexpr int kHdS = 72;         // head dim
constexpr int kHeadsS = 16;
constexpr int kHid = kHeadsS * kHdS;   // 1152
constexpr int kQKPitch = 88;     // bf16 elements per staged Q / K row (176 B: conflict-free 4-byte fragment loads)
constexpr int kQRows = 128;      // query rows per CTA
constexpr int kLd = 3 * kHid;    // row pitch of the fused q|k|v activation (3456)
constexpr size_t kSmemBytes = (size_t)(2 * kT + kQRows) * kQKPitch * 2;

__device__ __forceinline__ void cp16(void* dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ uint32_t lds32(const bf16* p) { return *reinterpret_cast<const uint32_t*>(p); }
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

__device__ __forceinline__ void ldmatrix_x2_trans(uint32_t& r0, uint32_t& r1, const void* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0, %1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(smem_u32(p)));
}

// qkv [B*256, 3456]: q(b, t, h, d) at (b*256 + t) * 3456 + h*72 + d, k at + 1152, v at + 2304
// att [B*256, 1152]
__global__ void __launch_bounds__(256, 1) siglip_attn_kernel(const bf16* __restrict__ qk, bf16* __restrict__ att,
                                                             float scale) {
  pdl_trigger();
  extern __shared__ __align__(16) uint8_t smem_raw[];
  bf16* Ks = reinterpret_cast<bf16*>(smem_raw);            // [256][88]
  bf16* Vs = Ks + kT * kQKPitch;                           // [256][88] row-major V
  bf16* Qs = Vs + kT * kQKPitch;                           // [128][88]
  const int qblk = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t4 = lane & 3;
  pdl_wait();

  // ---- stage K, V, Q (9 chunks of 16 B per row; dims 72..87 zero) -----------------------------------------------------
  const bf16* kbase = qk + (int64_t)b * kT * kLd + kHid + h * kHdS;            // V rows are kHid further
  const bf16* qbase = qk + ((int64_t)b * kT + qblk * kQRows) * kLd + h * kHdS;
  for (int i = threadIdx.x; i < (2 * kT + kQRows) * 11; i += 256) {
    const int row = i / 11, ci = i - row * 11;
    bf16* dst = Ks + row * kQKPitch + ci * 8;               // Ks | Vs | Qs are contiguous
    if (ci < 9) {
      const bf16* src = row < kT ? kbase + (int64_t)row * kLd
                                 : row < 2 * kT ? kbase + kHid + (int64_t)(row - kT) * kLd
                                                : qbase + (int64_t)(row - 2 * kT) * kLd;
      cp16(dst, src + ci * 8);
    } else {
      *reinterpret_cast<uint4*>(dst) = make_uint4(0, 0, 0, 0);
    }
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();

  // ---- S = Q K^T for this warp's 16 rows: 32 n-tiles of 8 keys, 5 k-steps of 16 dims ------------------------------------
  const int r0 = warp * 16;
  uint32_t aq[5][4];
#pragma unroll
  for (int ks = 0; ks < 5; ++ks) {
    aq[ks][0] = lds32(Qs + (r0 + g) * kQKPitch + ks * 16 + 2 * t4);
    aq[ks][1] = lds32(Qs + (r0 + g + 8) * kQKPitch + ks * 16 + 2 * t4);
    aq[ks][2] = lds32(Qs + (r0 + g) * kQKPitch + ks * 16 + 8 + 2 * t4);
    aq[ks][3] = lds32(Qs + (r0 + g + 8) * kQKPitch + ks * 16 + 8 + 2 * t4);
  }
  float acc[32][4];
#pragma unroll
  for (int nt = 0; nt < 32; ++nt) {
#pragma unroll
    for (int e = 0; e < 4; ++e) acc[nt][e] = 0.0f;
    const bf16* kr = Ks + (nt * 8 + g) * kQKPitch + 2 * t4;
#pragma unroll
    for (int ks = 0; ks < 5; ++ks) {
      const uint32_t b0 = lds32(kr + ks * 16), b1 = lds32(kr + ks * 16 + 8);
      mma_bf16_16816(acc[nt], aq[ks][0], aq[ks][1], aq[ks][2], aq[ks][3], b0, b1);
    }
  }

  // ---- softmax over the 256 keys of rows g (elements 0, 1) and g + 8 (elements 2, 3) ---------------------------------------
  float m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
  for (int nt = 0; nt < 32; ++nt) {
#pragma unroll
    for (int e = 0; e < 4; ++e) acc[nt][e] = bf16_round(bf16_round(acc[nt][e]) * scale);
    m0 = fmaxf(m0, fmaxf(acc[nt][0], acc[nt][1]));
    m1 = fmaxf(m1, fmaxf(acc[nt][2], acc[nt][3]));
  }
  m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1));
  m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
  m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1));
  m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
  float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
  for (int nt = 0; nt < 32; ++nt) {
    acc[nt][0] = expf(acc[nt][0] - m0);
    acc[nt][1] = expf(acc[nt][1] - m0);
    acc[nt][2] = expf(acc[nt][2] - m1);
    acc[nt][3] = expf(acc[nt][3] - m1);
    s0 += acc[nt][0] + acc[nt][1];
    s1 += acc[nt][2] + acc[nt][3];
  }
  s0 += __shfl_xor_sync(0xffffffffu, s0, 1);
  s0 += __shfl_xor_sync(0xffffffffu, s0, 2);
  s1 += __shfl_xor_sync(0xffffffffu, s1, 1);
  s1 += __shfl_xor_sync(0xffffffffu, s1, 2);
  const float i0 = 1.0f / s0, i1 = 1.0f / s1;
  // probabilities as A fragments of P.V: k-step kk covers keys [16 kk, 16 kk + 16) = n-tiles 2 kk, 2 kk + 1
  uint32_t pa[16][4];
#pragma unroll
  for (int kk = 0; kk < 16; ++kk) {
    pa[kk][0] = pack_bf16(acc[2 * kk][0] * i0, acc[2 * kk][1] * i0);
    pa[kk][1] = pack_bf16(acc[2 * kk][2] * i1, acc[2 * kk][3] * i1);
    pa[kk][2] = pack_bf16(acc[2 * kk + 1][0] * i0, acc[2 * kk + 1][1] * i0);
    pa[kk][3] = pack_bf16(acc[2 * kk + 1][2] * i1, acc[2 * kk + 1][3] * i1);
  }

  // ---- O = P V: 9 n-tiles of 8 dims, 16 k-steps of 16 keys; B fragments straight from V^T rows ----------------------------
  float o[9][4];
#pragma unroll
  for (int dt = 0; dt < 9; ++dt)
#pragma unroll
    for (int e = 0; e < 4; ++e) o[dt][e] = 0.0f;
#pragma unroll
  for (int kk = 0; kk < 16; ++kk) {
    // lanes 0-15 address the 16 key rows of this k-step; ldmatrix.trans hands every lane (keys 2 t4, 2 t4 + 1; dim g)
    const bf16* vrow = Vs + (kk * 16 + (lane & 15)) * kQKPitch;
#pragma unroll
    for (int dt = 0; dt < 9; ++dt) {
      uint32_t b0, b1;
      ldmatrix_x2_trans(b0, b1, vrow + dt * 8);
      mma_bf16_16816(o[dt], pa[kk][0], pa[kk][1], pa[kk][2], pa[kk][3], b0, b1);
    }
  }
  bf16* out0 = att + ((int64_t)b * kT + qblk * kQRows + r0 + g) * kHid + h * kHdS + 2 * t4;
  bf16* out1 = out0 + (int64_t)8 * kHid;
#pragma unroll
  for (int dt = 0; dt < 9; ++dt) {
    *reinterpret_cast<uint32_t*>(out0 + dt * 8) = pack_bf16(o[dt][0], o[dt][1]);
    *reinterpret_cast<uint32_t*>(out1 + dt * 8) = pack_bf16(o[dt][2], o[dt][3]);
  }
}

}  // namespace

int siglip_attention_mma(const void* qkv, void* att, int B, cudaStream_t s) {
  PI0_REQUIRE(qkv && att && B >= 1, "siglip_attention: bad argument");
  static PerDeviceOnce attr_once;
  if (attr_once.need()) {
    PI0_CHECK_CUDA(cudaFuncSetAttribute(siglip_attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    attr_once.mark();
  }
  PI0_CHECK_CUDA(launch_pdl(siglip_attn_kernel, dim3(kT / kQRows, kHeadsS, B), dim3(256), kSmemBytes, s, 1,
                            (const bf16*)qkv, (bf16*)att, 0.11785113019775793f));
  ++g_launch_count;
  return 0;
}

// The tcgen05 kernel (attn_siglip_tc.cu) is the product path; PI0_SIGLIP_ATTN=mma keeps this one selectable for comparison.
bool siglip_attention_use_tc() {
  const char* e = getenv("PI0_SIGLIP_ATTN");
  return !(e && e[0] == 'm');
}
int siglip_attention(const void* qkv, void* att, int B, int num_sms, cudaStream_t s) {
  if (siglip_attention_use_tc()) return siglip_attention_tc(qkv, att, B, num_sms, s);
  return siglip_attention_mma(qkv, att, B, s);
}

}  // namespace pi0
