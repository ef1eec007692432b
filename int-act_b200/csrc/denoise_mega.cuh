// Host-visible interface of the persistent denoise kernel (denoise_mega.cu).
#pragma once
#include <vector>

#include "elementwise.cuh"

namespace pi0 {

enum MegaPhase : int { MEGA_QKV = 0, MEGA_O = 1, MEGA_GU = 2, MEGA_DOWN = 3 };
constexpr int kMegaMaxUnits = 15;  // work units one CTA may own per phase (table row = count + 15 ids = 32 bytes)

struct MegaLayerW {
  const bf16 *wqkv, *wo, *wgu, *wdown;  // packed as for the tcgen05 path (fused q|k|v rows, gate/up interleaved per 256)
  const float *g1, *g2;                 // 1 + w of the two RMSNorms, fp32
};

struct MegaParams {
  MegaLayerW L[18];
  int n_layers;
  int M, S, P, B;       // M = B * S suffix rows
  const bf16* x0;       // layer-0 input [M, 1024] (suffix embeddings)
  bf16* x;              // residual stream after every layer [M, 1024]
  bf16* y;              // after attention [M, 1024]
  bf16* q;              // [M, 2048]
  bf16* att;            // [M, 2048]
  bf16* act;            // [M, 4096]
  bf16* kcache;         // [layer][B_max][kv_pitch][256]
  bf16* vtcache;        // [layer][B_max][256][vt_pitch]
  int64_t kc_layer_stride, vt_layer_stride;  // elements
  int kv_pitch, vt_pitch;
  const float* rope_cos;
  const float* rope_sin;
  const int* pos_ids;   // [M]
  const uint8_t* pad;   // [B, P]
  const int* cum;       // [S]
  const int16_t* units; // device copy of the table built by mega_build_units
  unsigned* sync;       // 64-bit arrival counter of the grid barrier, zero-initialised once, 8-byte aligned
  bf16* sc;             // score scratch [B*8*S][kp]
  int kp;               // key count padded to a multiple of 32
};

bool mega_supported(int M, int S, int P);
int mega_build_units(int G, std::vector<int16_t>& tab);
int mega_denoise_layers(const MegaParams& p, int num_ctas, cudaStream_t s);

}  // namespace pi0
