// Host-visible interface of the whole-loop persistent denoise kernel (denoise_loop.cu).
#pragma once
#include <vector>

#include "denoise_mega.cuh"

namespace pi0 {

// GEMM phases that consume streamed weights, in table order. The first four are the per-layer phases of
// denoise_mega.cuh (same unit numbering); the last three are the per-step projections of embed_suffix / the action head.
enum LoopPhase : int {
  LOOP_QKV = 0, LOOP_O = 1, LOOP_GU = 2, LOOP_DOWN = 3, LOOP_TIN = 4, LOOP_TOUT = 5, LOOP_OUT = 6, LOOP_PHASES = 7
};

struct LoopParams {
  MegaLayerW L[18];
  int n_layers, n_steps;
  int M, S, P, B, H;        // M = B * S suffix rows, H = S - 1 action rows per sample
  int layers_only;          // 1: one pass over the layers, x0 -> x (drop-in for mega_denoise_layers; unit tests)
  // embed_suffix / action head ([mirror] modeling_mvla.py:643-694, 806-839)
  const bf16 *state_w, *action_in_w, *time_in_wa, *time_out_w, *action_out_w;
  const float *state_b, *action_in_b, *time_bias, *time_out_b, *action_out_b, *exp_norm_g;
  const float* state;       // [B, 32] fp32
  float* x_t;               // [B, H, 32] fp32, in (noise) / out (actions)
  float dt;
  const bf16* x0;           // layers_only: layer-0 input [M, 1024]
  bf16* x;                  // residual stream after the last layer [M, 1024] (also written in loop mode: debug buffer)
  bf16* kcache;             // [layer][B_max][kv_pitch][256]
  bf16* vtcache;            // [layer][B_max][256][vt_pitch]
  int64_t kc_layer_stride, vt_layer_stride;
  int kv_pitch, vt_pitch;
  const float* rope_cos;
  const float* rope_sin;
  const int* pos_ids;       // [M]
  const uint8_t* pad;       // [B, P]
  const int* cum;           // [S]
  const int16_t* units;     // device copy of the table built by loop_build_units
  uint4* ll;                // flag-in-data exchange arena (loop_ll_bytes), zero-initialised once
  unsigned* epoch;          // running epoch of the arena (one u32, zero-initialised once)
  int kp;                   // key count padded to a multiple of 32
  int pace_ns;              // producer: pause after every streamed chunk (spreads the ring-refill bursts), 0 = none
};

bool loop_supported(int M, int S, int P, int num_ctas);
int loop_build_units(int G, std::vector<int16_t>& tab);
size_t loop_ll_bytes();
size_t loop_units_bytes();
int loop_denoise(const LoopParams& p, int num_ctas, cudaStream_t s);

}  // namespace pi0
