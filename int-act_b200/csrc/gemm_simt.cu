// CUDA-core GEMM with the same fused epilogues as the tcgen05 kernel. It is the fp32 "check mode"
// of the path (north_star: <= 1e-4 vs the oracle) and, instantiated for bf16, a debugging
// cross-check for the tensor-core kernel. Never the product path in the bf16 regime.
#include "gemm.cuh"

namespace pi0 {

long long g_launch_count = 0;

constexpr int kSR = 64;    // rows per block
constexpr int kSC = 128;   // primary columns per block (4 chunks of 32)
constexpr int kSK = 32;    // K slab

template <typename T, int EPI>
__global__ void __launch_bounds__(256) gemm_simt_kernel(const T* __restrict__ A, const T* __restrict__ W, int64_t lda,
                                                        int64_t ldw, int64_t a_z0, int64_t a_z1, int64_t w_z0,
                                                        int64_t w_z1, int M, int N, int K, int nz0, int bn,
                                                        const EpiParams ep) {
  constexpr bool kPaired = (EPI == EPI_GEGLU || EPI == EPI_ROPE);
  __shared__ float As[kSR][kSK + 1];
  __shared__ float Ws[kPaired ? 2 : 1][kSC][kSK + 1];

  const int z = blockIdx.z;
  const int z0 = z % nz0, z1 = z / nz0;
  A += (int64_t)z0 * a_z0 + (int64_t)z1 * a_z1;
  W += (int64_t)z0 * w_z0 + (int64_t)z1 * w_z1;

  const int r_local = threadIdx.x & 63;
  const int chunk = threadIdx.x >> 6;  // 0..3
  const int row = blockIdx.y * kSR + r_local;

  // column bookkeeping: primary chunk start, partner chunk start (paired modes)
  // STORE/EULER : blockIdx.x covers 128 consecutive columns
  // GEGLU       : blockIdx.x covers 128 output columns; gate col = tile*bn + j, up col = gate + bn/2
  // ROPE        : blockIdx.x = head (256 cols); primary = dims [chunk*32,+32), partner = +128
  int col_p0, col_q0 = 0;  // first column of the primary / partner 32-chunk of this thread
  int blk_p0, blk_q0 = 0;  // first column of the block's primary / partner 128-panel
  if constexpr (EPI == EPI_GEGLU) {
    const int half = bn / 2;
    // 128 output columns per block; half is a multiple of 128
    int out0 = blockIdx.x * kSC;
    int tile = out0 / half, within = out0 % half;
    blk_p0 = tile * bn + within;
    blk_q0 = blk_p0 + half;
  } else if constexpr (EPI == EPI_ROPE) {
    blk_p0 = blockIdx.x * 256;
    blk_q0 = blk_p0 + 128;
  } else {
    blk_p0 = blockIdx.x * kSC;
  }
  col_p0 = blk_p0 + chunk * 32;
  col_q0 = blk_q0 + chunk * 32;

  float accp[32], accq[kPaired ? 32 : 1];
#pragma unroll
  for (int j = 0; j < 32; ++j) accp[j] = 0.0f;
  if constexpr (kPaired) {
#pragma unroll
    for (int j = 0; j < 32; ++j) accq[j] = 0.0f;
  }

  for (int k0 = 0; k0 < K; k0 += kSK) {
    // A slab: 64 x 32
    for (int i = threadIdx.x; i < kSR * kSK; i += 256) {
      int r = i / kSK, k = i % kSK;
      int gr = blockIdx.y * kSR + r, gk = k0 + k;
      As[r][k] = (gr < M && gk < K) ? Num<T>::ld(A + (int64_t)gr * lda + gk) : 0.0f;
    }
    for (int i = threadIdx.x; i < kSC * kSK; i += 256) {
      int c = i / kSK, k = i % kSK;
      int gk = k0 + k;
      int gc = blk_p0 + c;
      Ws[0][c][k] = (gc < N && gk < K) ? Num<T>::ld(W + (int64_t)gc * ldw + gk) : 0.0f;
      if constexpr (kPaired) {
        int gq = blk_q0 + c;
        Ws[1][c][k] = (gq < N && gk < K) ? Num<T>::ld(W + (int64_t)gq * ldw + gk) : 0.0f;
      }
    }
    __syncthreads();
#pragma unroll 4
    for (int k = 0; k < kSK; ++k) {
      float a = As[r_local][k];
#pragma unroll
      for (int j = 0; j < 32; ++j) accp[j] = fmaf(a, Ws[0][chunk * 32 + j][k], accp[j]);
      if constexpr (kPaired) {
#pragma unroll
        for (int j = 0; j < 32; ++j) accq[j] = fmaf(a, Ws[1][chunk * 32 + j][k], accq[j]);
      }
    }
    __syncthreads();
  }

  if (row >= M) return;
  if constexpr (EPI == EPI_STORE) {
    if (col_p0 < N) epi_store_chunk<T>(ep, row, col_p0, N, (int64_t)z0 * ep.out_z0 + (int64_t)z1 * ep.out_z1, accp);
  } else if constexpr (EPI == EPI_EULER) {
    if (col_p0 < N) epi_euler_chunk<T>(ep, row, col_p0, N, accp);
  } else if constexpr (EPI == EPI_GEGLU) {
    int out_col0 = blockIdx.x * kSC + chunk * 32;
    if (out_col0 < N / 2) epi_geglu_chunk<T>(ep, row, out_col0, N / 2, accp, accq);
  } else if constexpr (EPI == EPI_ROPE) {
    int head = blockIdx.x;
    int which = head < ep.n_q_heads ? 0 : (head == ep.n_q_heads ? 1 : 2);
    epi_rope_chunk<T>(ep, row, which, head, chunk, accp, accq);
  }
}

template <typename T>
int gemm_simt_launch(const GemmProblem& g, cudaStream_t stream) {
  dim3 block(256);
  dim3 grid;
  grid.y = (g.M + kSR - 1) / kSR;
  grid.z = g.nz0 * g.nz1;
  const T* A = reinterpret_cast<const T*>(g.A);
  const T* W = reinterpret_cast<const T*>(g.W);
#define PI0_SIMT_ARGS A, W, g.lda, g.ldw, g.a_z0, g.a_z1, g.w_z0, g.w_z1, g.M, g.N, g.K, g.nz0, g.bn, g.epi
  switch (g.epi_mode) {
    case EPI_STORE:
      grid.x = (g.N + kSC - 1) / kSC;
      gemm_simt_kernel<T, EPI_STORE><<<grid, block, 0, stream>>>(PI0_SIMT_ARGS);
      break;
    case EPI_EULER:
      grid.x = (g.N + kSC - 1) / kSC;
      gemm_simt_kernel<T, EPI_EULER><<<grid, block, 0, stream>>>(PI0_SIMT_ARGS);
      break;
    case EPI_GEGLU:
      PI0_REQUIRE(g.bn == 256 && g.N % 256 == 0, "gemm_simt: GeGLU expects bn=256 packing");
      grid.x = (g.N / 2 + kSC - 1) / kSC;
      gemm_simt_kernel<T, EPI_GEGLU><<<grid, block, 0, stream>>>(PI0_SIMT_ARGS);
      break;
    case EPI_ROPE:
      PI0_REQUIRE(g.N == (g.epi.n_q_heads + 2) * 256, "gemm_simt: RoPE epilogue expects N = (q_heads + 2) * 256");
      grid.x = g.N / 256;
      gemm_simt_kernel<T, EPI_ROPE><<<grid, block, 0, stream>>>(PI0_SIMT_ARGS);
      break;
    default:
      set_error("gemm_simt: unknown epilogue %d", g.epi_mode);
      return -2;
  }
#undef PI0_SIMT_ARGS
  PI0_CHECK_CUDA(cudaGetLastError());
  ++g_launch_count;
  return 0;
}

template int gemm_simt_launch<float>(const GemmProblem&, cudaStream_t);
template int gemm_simt_launch<bf16>(const GemmProblem&, cudaStream_t);

int gemm_launch(const GemmProblem& g, int dtype, int impl, int num_sms, cudaStream_t stream) {
  if (dtype == 1) return gemm_simt_launch<float>(g, stream);
  if (impl == 1) return gemm_simt_launch<bf16>(g, stream);
  if (impl == 2 || impl == 3) {
    GemmProblem f = g;
    f.pair = impl == 3 ? 1 : -1;
    return gemm_tc_launch(f, num_sms, stream);
  }
  return gemm_tc_launch(g, num_sms, stream);
}

}  // namespace pi0
