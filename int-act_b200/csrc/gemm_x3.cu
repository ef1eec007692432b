// The GEMM of the fp32 regime (`--use_bf16 False`, what the reference's shipped eval scripts pass:
// /root/reference/slurms/eval_scripts/simpler/ev_pi0_bridge_simpler.sh:99, doc/evaluation.md:152) on the tensor cores.
//
// A B200 has no fp32 tensor path, and its CUDA cores give ~70 TFLOP/s of fp32 FMA against 2250 of bf16 MMA. An fp32 number
// is the exact sum of three bf16 numbers (8 + 8 + 8 mantissa bits): x = hi + mid + lo with hi = rn_bf16(x),
// mid = rn_bf16(x - hi), lo = rn_bf16(x - hi - mid). A product a.w then is the sum of nine piece products, of which the six
// with weight >= 2^-16 relative are kept (hi.hi, hi.mid, mid.hi, hi.lo, lo.hi, mid.mid; the dropped ones are <= 2^-24
// relative, i.e. below fp32 rounding). Every piece product is exact in fp32 (8 x 8 bits), the tensor core accumulates in fp32.
// Laid out along K the six products are ONE bf16 GEMM with K' = 6 K:
//     A' = [ lo | hi | mid | mid | hi  | hi ]        W' = [ hi | lo | mid | hi  | mid | hi ]
// (small terms first, so they meet a small accumulator), which the unchanged tcgen05 pipeline of gemm_tc.cu runs with the
// epilogues instantiated for fp32 storage: 6 bf16 MMAs per fp32 FMA is still >5x the CUDA-core peak.
// The split operands go to caller-owned scratch (the ctx workspace). Weights are split on their first use into an optional
// cache (X3WeightCache, 12 bytes per parameter) and found by pointer afterwards; without it they are split on every call.
#include "gemm.cuh"

namespace pi0 {

namespace {

constexpr int kPieces = 6;

// piece p of the A-side / W-side layout: 0 = hi, 1 = mid, 2 = lo
__device__ __constant__ int kPieceA[kPieces] = {2, 0, 1, 1, 0, 0};
__device__ __constant__ int kPieceW[kPieces] = {0, 2, 1, 0, 1, 0};

__device__ __forceinline__ void split3(float x, float& hi, float& mid, float& lo) {
  hi = bf16_round(x);
  const float r1 = x - hi;        // exact: hi shares x's leading bits
  mid = bf16_round(r1);
  lo = bf16_round(r1 - mid);      // r1 - mid is exact; lo holds its top 8 bits (all of them when x has 24)
}

// src: [nz1][nz0] matrices of rows x K fp32 (row pitch ld, batch strides z0 / z1, elements)
// dst: [nz1 * nz0][rows][6][Kp] bf16, zero for k in [K, Kp)
template <bool W_SIDE, bool VEC>
__global__ void __launch_bounds__(256) split3_kernel(const float* __restrict__ src, int64_t ld, int64_t sz0, int64_t sz1,
                                                     int nz0, int64_t rows, int K, int Kp, int64_t total_groups,
                                                     bf16* __restrict__ dst) {
  pdl_trigger();
  pdl_wait();
  const int gpr = Kp >> 3;  // groups of 8 k per row
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total_groups; i += (int64_t)gridDim.x * 256) {
    const int g = (int)(i % gpr);
    const int64_t rz = i / gpr;
    const int64_t row = rz % rows;
    const int64_t z = rz / rows;
    const int z0 = (int)(z % nz0), z1 = (int)(z / nz0);
    const float* s = src + (int64_t)z0 * sz0 + (int64_t)z1 * sz1 + row * ld + g * 8;
    float x[8];
    if (VEC) {
      const float4 a = __ldg(reinterpret_cast<const float4*>(s));
      const float4 b = __ldg(reinterpret_cast<const float4*>(s + 4));
      x[0] = a.x; x[1] = a.y; x[2] = a.z; x[3] = a.w; x[4] = b.x; x[5] = b.y; x[6] = b.z; x[7] = b.w;
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) x[j] = g * 8 + j < K ? __ldg(s + j) : 0.0f;
    }
    float p[3][8];
#pragma unroll
    for (int j = 0; j < 8; ++j) split3(x[j], p[0][j], p[1][j], p[2][j]);
    uint4 pk[3];
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      pk[q].x = bf16x2_bits(p[q][1], p[q][0]);
      pk[q].y = bf16x2_bits(p[q][3], p[q][2]);
      pk[q].z = bf16x2_bits(p[q][5], p[q][4]);
      pk[q].w = bf16x2_bits(p[q][7], p[q][6]);
    }
    bf16* d = dst + (rz * kPieces) * Kp + g * 8;
#pragma unroll
    for (int q = 0; q < kPieces; ++q) {
      const int which = W_SIDE ? kPieceW[q] : kPieceA[q];
      *reinterpret_cast<uint4*>(d + (int64_t)q * Kp) = which == 0 ? pk[0] : which == 1 ? pk[1] : pk[2];
    }
  }
}

template <bool W_SIDE>
int launch_split(const void* src, int64_t ld, int64_t sz0, int64_t sz1, int nz0, int nz1, int64_t rows, int K, int Kp,
                 void* dst, int num_sms, cudaStream_t s) {
  const int64_t groups = (int64_t)nz0 * nz1 * rows * (Kp >> 3);
  const bool vec = K % 8 == 0 && ld % 4 == 0 && sz0 % 4 == 0 && sz1 % 4 == 0 && (reinterpret_cast<uintptr_t>(src) & 15) == 0;
  int64_t blocks = (groups + 255) / 256;
  const int64_t cap = (int64_t)num_sms * 16;
  if (blocks > cap) blocks = cap;
  if (vec)
    PI0_CHECK_CUDA(launch_pdl(split3_kernel<W_SIDE, true>, dim3((unsigned)blocks), dim3(256), 0, s, 1, (const float*)src, ld,
                              sz0, sz1, nz0, rows, K, Kp, groups, (bf16*)dst));
  else
    PI0_CHECK_CUDA(launch_pdl(split3_kernel<W_SIDE, false>, dim3((unsigned)blocks), dim3(256), 0, s, 1, (const float*)src, ld,
                              sz0, sz1, nz0, rows, K, Kp, groups, (bf16*)dst));
  ++g_launch_count;
  return 0;
}


// ---------------------------------------------------------------------------------------------
// Skinny problems of the fp32 regime (M <= 64 rows: the denoise steps up to batch 12): the weight matrix is read ONCE, in
// fp32, by CUDA-core FMAs - at M = 40 the tensor-core form above spends 4 launches and a 16-byte-per-parameter split pass
// on a problem whose floor is one pass over 4 bytes per parameter. Work item = (64 output columns, 256-wide k slice): the
// CTA stages the A slice [M x 256] and the W slice [64 x 256] in shared memory with cp.async, thread (col, q) accumulates
// rows q, q + 4, ... with 16-byte shared loads (A broadcast across the warp, W conflict-free through a 260-float pitch),
// and the k slices' partial sums go to the fp32 workspace that x3_reduce_kernel (gemm_tc.cu) sums in slice order before
// the fused epilogue. Deterministic, plain fp32 arithmetic.
// ---------------------------------------------------------------------------------------------
constexpr int kSkCols = 64, kSkK = 256, kSkWPitch = kSkK + 4, kSkMaxM = 64;

// RPT = rows per thread = ceil(M / 4), a template parameter: predicated-off FMAs still take issue slots (the 16-row form took
// 23 us at M = 5)
template <int RPT>
__global__ void __launch_bounds__(256) skinny_f32_kernel(const float* __restrict__ A, int64_t lda,
                                                         const float* __restrict__ W, int64_t ldw, int M, int N,
                                                         float* __restrict__ partial, int ldp) {
  extern __shared__ __align__(16) float sk_smem[];
  float* Ws = sk_smem;                         // [64][260]
  float* As = sk_smem + kSkCols * kSkWPitch;   // [M][256]
  pdl_trigger();
  const int cb = blockIdx.x, ks = blockIdx.y;
  const int col = threadIdx.x & 63, q = threadIdx.x >> 6;
  // the weights are never written by a preceding kernel: their slice is requested before the dependency wait
  for (int i = threadIdx.x; i < kSkCols * (kSkK / 4); i += 256) {
    const int c = i / (kSkK / 4), k4 = i % (kSkK / 4);
    const float* src = W + (int64_t)(cb * kSkCols + c) * ldw + ks * kSkK + k4 * 4;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(Ws + c * kSkWPitch + k4 * 4)), "l"(src) : "memory");
  }
  pdl_wait();
  for (int i = threadIdx.x; i < M * (kSkK / 4); i += 256) {
    const int r = i / (kSkK / 4), k4 = i % (kSkK / 4);
    const float* src = A + (int64_t)r * lda + ks * kSkK + k4 * 4;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(As + r * kSkK + k4 * 4)), "l"(src) : "memory");
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  const int nr = (M - q + 3) / 4;   // rows q, q + 4, ... < M
  float acc[RPT];
#pragma unroll
  for (int i = 0; i < RPT; ++i) acc[i] = 0.0f;
  const float4* wrow = reinterpret_cast<const float4*>(Ws + col * kSkWPitch);
#pragma unroll 2
  for (int k4 = 0; k4 < kSkK / 4; ++k4) {
    const float4 w = wrow[k4];
#pragma unroll
    for (int i = 0; i < RPT; ++i) {
      if (i < nr) {
        const float4 a = *reinterpret_cast<const float4*>(As + (q + 4 * i) * kSkK + k4 * 4);
        acc[i] = fmaf(a.x, w.x, acc[i]);
        acc[i] = fmaf(a.y, w.y, acc[i]);
        acc[i] = fmaf(a.z, w.z, acc[i]);
        acc[i] = fmaf(a.w, w.w, acc[i]);
      }
    }
  }
  float* dst = partial + ((int64_t)ks * M) * ldp + cb * kSkCols + col;
#pragma unroll
  for (int i = 0; i < RPT; ++i)
    if (i < nr) dst[(int64_t)(q + 4 * i) * ldp] = acc[i];
}

// 0 = launched, 1 = not a skinny problem / does not fit
int skinny_f32_launch(const GemmProblem& g, void* partial, size_t bytes_p, cudaStream_t stream) {
  static const bool off = getenv("PI0_FP32_SKINNY") && getenv("PI0_FP32_SKINNY")[0] == '0';
  if (off || partial == nullptr || g.nz0 * g.nz1 != 1 || g.M > kSkMaxM || g.N % kSkCols != 0 || g.K % kSkK != 0) return 1;
  if (g.lda % 4 != 0 || g.ldw % 4 != 0 || ((reinterpret_cast<uintptr_t>(g.A) | reinterpret_cast<uintptr_t>(g.W)) & 15) != 0)
    return 1;
  const int S = g.K / kSkK, ldp = (g.N + 31) / 32 * 32;
  if ((size_t)S * g.M * ldp * sizeof(float) > bytes_p) return 1;
  const size_t smem = (size_t)(kSkCols * kSkWPitch + g.M * kSkK) * sizeof(float);
  const int max_smem = (int)((kSkCols * kSkWPitch + kSkMaxM * kSkK) * sizeof(float));
  const dim3 grid((unsigned)(g.N / kSkCols), (unsigned)S);
  const int rpt = (g.M + 3) / 4;
#define PI0_SKINNY_CASE(R)                                                                                              \
  {                                                                                                                     \
    static PerDeviceOnce once;                                                                                          \
    if (once.need()) {                                                                                                  \
      PI0_CHECK_CUDA(cudaFuncSetAttribute(skinny_f32_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem)); \
      once.mark();                                                                                                      \
    }                                                                                                                   \
    PI0_CHECK_CUDA(launch_pdl(skinny_f32_kernel<R>, grid, dim3(256), smem, stream, 1, (const float*)g.A, g.lda,         \
                              (const float*)g.W, g.ldw, g.M, g.N, (float*)partial, ldp));                               \
  }
  if (rpt <= 1) PI0_SKINNY_CASE(1)
  else if (rpt <= 2) PI0_SKINNY_CASE(2)
  else if (rpt <= 4) PI0_SKINNY_CASE(4)
  else if (rpt <= 6) PI0_SKINNY_CASE(6)
  else if (rpt <= 8) PI0_SKINNY_CASE(8)
  else if (rpt <= 10) PI0_SKINNY_CASE(10)
  else if (rpt <= 12) PI0_SKINNY_CASE(12)
  else PI0_SKINNY_CASE(16)
#undef PI0_SKINNY_CASE
  ++g_launch_count;
  return gemm_f32_reduce_launch(g, (const float*)partial, S, ldp, stream);
}

}  // namespace

size_t gemm_x3_operand_bytes(int64_t nz, int64_t rows, int K) {
  const int64_t Kp = ((int64_t)K + 7) / 8 * 8;
  return (size_t)(nz * rows * kPieces * Kp * 2);
}

int gemm_x3_launch(const GemmProblem& g, void* scratch_a, size_t bytes_a, void* scratch_w, size_t bytes_w, void* partial,
                   size_t bytes_p, int num_sms, cudaStream_t stream, X3WeightCache* cache, bool w_is_weight) {
  PI0_REQUIRE(g.M > 0 && g.N > 0 && g.K > 0, "gemm_x3: empty problem M=%d N=%d K=%d", g.M, g.N, g.K);
  const int64_t nz = (int64_t)g.nz0 * g.nz1;
  {
    const int rc = skinny_f32_launch(g, partial, bytes_p, stream);
    if (rc != 1) return rc;
  }
  // what the tcgen05 launcher needs from the epilogue side (the SIMT kernel has no such constraints)
  const bool ragged_ok = g.epi_mode == EPI_STORE && g.epi.bias == nullptr && g.epi.res == nullptr &&
                         g.epi.ldc >= (int64_t)((g.N + 7) / 8) * 8;
  if (!(g.N % 8 == 0 || g.epi_mode == EPI_EULER || g.epi.vt != nullptr || ragged_ok)) return 1;
  if (scratch_a == nullptr || scratch_w == nullptr) return 1;
  const size_t w_bytes = gemm_x3_operand_bytes(nz, g.N, g.K);
  if (gemm_x3_operand_bytes(nz, g.M, g.K) > bytes_a) return 1;
  const int Kp = (g.K + 7) / 8 * 8;
  // W side: a cached split of a weight, a fresh cache entry, or the per-call scratch
  void* w_split = nullptr;
  bool w_ready = false;
  if (cache != nullptr && w_is_weight && nz == 1 && cache->base != nullptr) {
    for (int i = 0; i < cache->n; ++i)
      if (cache->key[i] == g.W) { w_split = cache->base + cache->off[i]; w_ready = true; break; }
    if (!w_ready && cache->n < X3WeightCache::kMaxEntries && cache->used + w_bytes <= cache->bytes) {
      w_split = cache->base + cache->used;
      cache->key[cache->n] = g.W;
      cache->off[cache->n] = cache->used;
      ++cache->n;
      cache->used += (w_bytes + 1023) & ~(size_t)1023;
    }
  }
  if (w_split == nullptr) {
    if (w_bytes > bytes_w) return 1;
    w_split = scratch_w;
  }
  PI0_TRY(launch_split<false>(g.A, g.lda, g.a_z0, g.a_z1, g.nz0, g.nz1, g.M, g.K, Kp, scratch_a, num_sms, stream));
  if (!w_ready) PI0_TRY(launch_split<true>(g.W, g.ldw, g.w_z0, g.w_z1, g.nz0, g.nz1, g.N, g.K, Kp, w_split, num_sms, stream));
  GemmProblem t = g;
  const int64_t K6 = (int64_t)kPieces * Kp;
  t.A = scratch_a; t.lda = K6; t.a_z0 = (int64_t)g.M * K6; t.a_z1 = (int64_t)g.nz0 * g.M * K6;
  t.W = w_split; t.ldw = K6; t.w_z0 = (int64_t)g.N * K6; t.w_z1 = (int64_t)g.nz0 * g.N * K6;
  PI0_REQUIRE(K6 <= 0x7fffffff, "gemm_x3: K=%d too long", g.K);
  t.K = (int)K6;
  return gemm_tc_launch_f32out(t, num_sms, partial, bytes_p, stream);
}

}  // namespace pi0
