// Host-visible interface of the fused denoise-step kernel for rollout batches (fused_denoise.cu).
#pragma once
#include <vector>

#include "elementwise.cuh"

namespace pi0 {

// One GEMM of an expert layer: A = activations [rows, K] (bf16, row-major), W = packed weights [N, K]; the tensor maps are
// built once per ctx for the largest batch (rows beyond M are never stored).
struct alignas(128) FusedGemmDesc {
  CUtensorMap tmA;      // box 128 rows x 64 k, 128-byte swizzle
  CUtensorMap tmB;      // box bn rows x 64 k
  int N, K, bn, n_tiles, num_k;
  unsigned stage_tx;    // bytes per pipeline stage (A box + B box)
  int reserved[2];
};

struct FusedParams {
  const FusedGemmDesc* gd;     // device array [n_layers][4]: qkv, o_proj, gate/up, down
  const float* g1[18];         // 1 + w of the input RMSNorm of every layer, fp32
  const float* g2[18];         // ... of the post-attention RMSNorm
  int n_layers;
  int M, S, P, B;              // M = B * S suffix rows
  const bf16* x0;              // layer-0 input [M, 1024] (suffix embeddings)
  bf16 *x, *y, *h, *q, *att, *act;   // residual stream, post-attention stream, normed operand, q / attention out [M, 2048], act [M, 4096]
  float* partial;              // split-K partial sums [split][M][N] fp32
  long long partial_elems;
  bf16* kcache;                // [layer][B_max][kv_pitch][256]
  bf16* vtcache;               // [layer][B_max][256][vt_pitch]
  long long kc_layer_stride, vt_layer_stride;
  int kv_pitch, vt_pitch;
  const float* rope_cos;
  const float* rope_sin;
  const int* pos_ids;          // [M]
  const uint8_t* pad;          // [B, P]
  const int* cum;              // [S]
  unsigned* sync;              // 64-bit arrival counter of the grid barrier, zero-initialised once
  int kp, hpg;
};

bool fused_supported(int M, int S, int P, int B, int num_ctas);
size_t fused_desc_bytes();
size_t fused_partial_bytes(int max_rows, int num_ctas);
// fills `host` (18 * 4 descriptors) for the given buffers; `max_rows` = rows of the activation buffers
int fused_build_descs(std::vector<FusedGemmDesc>& host, int n_layers, int max_rows, const void* h, const void* att, const void* act,
                      const void* const* wqkv, const void* const* wo, const void* const* wgu, const void* const* wdown);
int fused_denoise_layers(const FusedParams& p, int num_ctas, cudaStream_t s);

}  // namespace pi0
