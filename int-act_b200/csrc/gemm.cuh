// Host-side description of one (optionally batched) GEMM  C[z][M,N] = A[z][M,K] . W[z][N,K]^T
// with a fused epilogue. Both operands are K-major (nn.Linear weight layout), which is the layout
// TMA + tcgen05 consume directly.
#pragma once
#include "epilogue.cuh"

namespace pi0 {

struct GemmProblem {
  const void* A;
  int64_t lda, a_z0, a_z1;  // elements
  const void* W;
  int64_t ldw, w_z0, w_z1;  // elements
  int M, N, K;
  int nz0, nz1;             // batch grid (1,1 for a plain GEMM)
  int epi_mode;             // EpiMode
  int bn;                   // N tile of the tcgen05 kernel (32/64/128/256; 0 = auto); EPI_ROPE requires 256
  int splits;               // split-K factor = cluster size (1..8; 0 = auto)
  int pair;                 // CTA-pair (cta_group::2) kernel: 0 = auto, 1 = force, -1 = never
  int w_const;              // 1: W is a constant parameter matrix (never written by a preceding kernel): its first tiles may be
                            // fetched before the programmatic-dependency wait
  // Gemma RMSNorm fused into the A operand (tcgen05 path, EPI_ROPE / EPI_GEGLU, bn 256): A holds the RAW residual stream, the
  // kernel forms rnd((a * rstd[row]) * norm_gain[k]) in shared memory before the MMA; rstd[row] = rsqrt(sum of the row's
  // norm_slots entries of norm_rowsq / K + eps), the per-chunk sums of squares left by the producing GEMM (EpiParams.rowsq_out)
  const float* norm_gain;   // [K] fp32 (1 + w), nullptr = no fused norm
  const float* norm_rowsq;  // [M][norm_slots]
  int norm_slots;
  float norm_eps;
  EpiParams epi;
};

inline GemmProblem gemm_problem(const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K) {
  GemmProblem g;
  g.A = A; g.lda = lda; g.a_z0 = 0; g.a_z1 = 0;
  g.W = W; g.ldw = ldw; g.w_z0 = 0; g.w_z1 = 0;
  g.M = M; g.N = N; g.K = K; g.nz0 = 1; g.nz1 = 1;
  g.epi_mode = EPI_STORE; g.bn = 256; g.splits = 0; g.pair = 0;
  g.w_const = 0;
  g.norm_gain = nullptr; g.norm_rowsq = nullptr; g.norm_slots = 0; g.norm_eps = 1e-6f;
  g.epi = epi_default();
  return g;
}

// dtype: 0 = bf16 activations/weights (tcgen05 path), 1 = fp32 check mode (SIMT path).
// impl:  0 = default for the dtype, 1 = force the SIMT kernel (debug cross-check of the tcgen05 path),
//        2 = tcgen05 single-CTA kernels only, 3 = force the tcgen05 CTA-pair kernel.
int gemm_launch(const GemmProblem& g, int dtype, int impl, int num_sms, cudaStream_t stream);

int gemm_tc_launch(const GemmProblem& g, int num_sms, cudaStream_t stream);
// bf16 operands, epilogue and output in fp32 storage (EpiParams pointers are float buffers)
// `partial`: optional fp32 workspace for the split-K form (skinny problems), may be null
int gemm_tc_launch_f32out(const GemmProblem& g, int num_sms, void* partial, size_t partial_bytes, cudaStream_t stream);
// The fp32 regime's GEMM (gemm_x3.cu): fp32 operands split into three bf16 pieces each, six piece products accumulated in
// fp32 on the tensor cores. scratch_a / scratch_w: caller-owned device buffers for the split operands. Returns 1 when the
// problem does not fit the scratch or the tcgen05 constraints (caller falls back to the SIMT kernel), 0 on success, < 0 on error.
// Optional cache of split WEIGHT operands (12 bytes per parameter, caller-owned memory): a weight is split the first time a GEMM
// uses it and found by its pointer afterwards, which removes the W-side split pass from every later call.
struct X3WeightCache {
  uint8_t* base = nullptr;
  size_t bytes = 0, used = 0;
  static constexpr int kMaxEntries = 1024;
  const void* key[kMaxEntries];
  size_t off[kMaxEntries];
  int n = 0;
};
// w_is_weight: g.W is a constant parameter matrix (not an activation), i.e. eligible for `cache`
int gemm_x3_launch(const GemmProblem& g, void* scratch_a, size_t bytes_a, void* scratch_w, size_t bytes_w, void* partial,
                   size_t bytes_p, int num_sms, cudaStream_t stream, X3WeightCache* cache = nullptr, bool w_is_weight = false);
size_t gemm_x3_operand_bytes(int64_t nz, int64_t rows, int K);
// sums S fp32 partial slices [S][nz][M][ldp] in slice order and runs g's fused epilogue in fp32 storage (gemm_tc.cu)
int gemm_f32_reduce_launch(const GemmProblem& g, const float* partial, int S, int ldp, cudaStream_t stream);
// rank-4 bf16 tensor map {k, rows, z0, z1} with box {box_k, box_rows, 1, 1}, 64- or 128-byte swizzle, zero OOB fill
int tc_make_map(CUtensorMap* map, const void* base, int K, int rows, int nz0, int nz1, int64_t ld, int64_t sz0, int64_t sz1,
                int box_rows, int box_k, int swizzle_bytes);
template <typename T>
int gemm_simt_launch(const GemmProblem& g, cudaStream_t stream);

// number of tcgen05 GEMM / SIMT GEMM / other kernels launched since the last reset (bench "gpu_launches")
extern long long g_launch_count;

}  // namespace pi0
