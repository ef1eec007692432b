// Host orchestration of the Pi0 `sample_actions` path behind the C ABI: SigLIP -> prefix embedding ->
// Gemma-2B prefill (KV cache written in place by the QKV GEMM epilogue) -> 10-step flow-matching Euler
// loop of the action expert over the cached prefix. No device allocation, no synchronisation: every
// buffer is carved from the caller's workspace and every launch goes to the caller's stream, so
// pi0_infer / pi0_denoise_loop are CUDA-graph capturable.
//
// Reference arithmetic followed ([mirror] = /root/reference/src/model/mvla/*):
//   embed_prefix        modeling_mvla.py:582-624,636-641      embed_image      mllm.py:182-187
//   Gemma layer body    mllm.py:206-245                        attention        mllm.py:265-317
//   embed_suffix        modeling_mvla.py:643-694               denoise_step     modeling_mvla.py:806-839
//   Euler loop          modeling_mvla.py:787-803
#include <cstdlib>
#include <cstring>
#include <map>
#include <new>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "../../include/pi0b200.h"
#include "denoise_loop.cuh"
#include "fused_denoise.cuh"
#include "denoise_mega.cuh"
#include "elementwise.cuh"

using namespace pi0;

namespace {

constexpr int kPatches = 256;
constexpr int kVisHidden = 1152, kVisInter = 4304, kVisHeads = 16, kVisHd = 72, kPatchK = 592;
constexpr int kHidden = 2048, kInter = 16384, kExpHidden = 1024, kExpInter = 4096;
constexpr int kHeads = 8, kHd = 256, kProjW = 1024;
constexpr float kEps = 1e-6f;

inline int64_t round_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

struct Buf {
  const char* name;
  size_t offset;
  size_t bytes;
};

}  // namespace

struct pi0_ctx {
  pi0_model_cfg cfg;
  pi0_weights w;
  uint8_t* base;
  size_t ws_bytes;
  int es;  // element size of the activation type
  int P, S, H, L, ncam, maxB;
  int kv_pitch;   // rows per sample in the K cache (P + S)
  int sc_pitch;   // score row pitch (>= P + S, multiple of 8)
  int vt_pitch;   // V^T row pitch (>= P + S, multiple of 8)
  // SigLIP
  void *col, *xs, *hs, *qk_s, *vt_s, *sc_s, *att_s, *mlp_s;
  // Gemma prefill
  void *xp, *hp, *qp, *sc_p, *att_p, *act_p;
  void *kcache, *vtcache;  // [layers][B][kv_pitch][256], [layers][B][256][vt_pitch]
  // expert
  void *suf, *xe, *ye, *he, *qe, *sc_e, *att_e, *act_e, *xt_act, *st_act, *a1, *a2, *hfin;
  uint8_t* pad;
  int *prefix_pos, *suffix_pos, *suffix_cum;
  // persistent denoise kernel (denoise_mega.cu): per-CTA work-unit table and the grid-barrier words
  int16_t* mega_units;
  unsigned* mega_sync;
  void* mega_sc;
  int mega_ctas;  // 0 = unavailable / disabled (PI0_MEGA=0)
  int mega_max_rows;  // largest M = B * S the persistent per-step kernel takes (PI0_MEGA_MAX_ROWS, default 32)
  // whole-loop persistent denoise kernel (denoise_loop.cu): unit table, flag-in-data exchange arena and its epoch word
  int16_t* loop_units;
  uint4* loop_ll;
  unsigned* loop_epoch;
  // fused denoise-step kernel for rollout batches (fused_denoise.cu): GEMM descriptors, split-K scratch, barrier counter
  FusedGemmDesc* fused_descs;
  float* fused_partial;
  unsigned* fused_sync;
  bool fused_on;  // PI0_FUSED=1 enables (opt-in: measured slower than the per-op launches)
  int fused_min_batch;  // PI0_FUSED_MIN_BATCH (default 7)
  int loop_pace_ns;  // PI0_LOOP_PACE (development knob)
  int loop_mode;  // 0 = off, 1 = per-step layers-only launches (PI0_MEGA=2), 2 = the whole Euler loop in one launch (default)
  bool siglip_fused;  // fused SigLIP attention kernel (PI0_SIGLIP_FUSED=0 disables)
  void* x3_a;  // fp32 regime: split-bf16 operand scratch of the tensor-core GEMM (gemm_x3.cu), null when it is off
  void* x3_w;
  void* x3_p;  // fp32 partial sums of the split-K form (skinny problems)
  size_t x3_a_bytes, x3_w_bytes, x3_p_bytes;
  float* rowsq_x;  // fused RMSNorm of the rollout-batch denoise path: per-chunk sums of squares of the residual stream rows
  float* rowsq_y;  // ... and of the post-attention stream (written by the o_proj / down epilogues)
  bool norm_fuse;  // PI0_NORM_FUSE=0 disables
  int norm_fuse_max_rows;  // largest M = B x S that takes the fused norms (PI0_NORM_FUSE_MAX_ROWS, default 64)
  X3WeightCache x3_cache;  // split copies of the GEMM weights (PI0_FP32_WCACHE=0 disables)
  bool attn_split;  // two-launch denoise attention for B >= 4 (PI0_ATTN_SPLIT=0 disables)
  bool prefill_fused;  // fused tcgen05 prefix attention (PI0_PREFILL_FUSED=0 disables)
  int prefill_fused_min_tiles;  // ... for at least this many 128-row tiles (PI0_PREFILL_FUSED_MIN_TILES, default 2 x SMs)
  Buf bufs[56];
  int nbufs;
  bool guarded;  // PI0_GUARD: 1 KiB 0xA5 zones behind every buffer
  // per-launch CUDA-event profiling (pi0_profile / pi0_profile_report); off on the product path
  bool prof_on;
  bool nvtx_on;  // PI0_NVTX=1
  struct ProfRec { const char* label; cudaEvent_t e0, e1; };
  std::vector<ProfRec> prof;
};

namespace {

// Records a CUDA-event pair around the launches issued while it is alive (only when profiling is enabled).
// With PI0_NVTX=1 every phase is also an NVTX range (same labels), so that `ncu --nvtx --nvtx-include "prefill.attention/"`
// or a timeline tool can select a phase by name (tracing).
struct ProfScope {
  pi0_ctx* c; cudaStream_t s; cudaEvent_t e1; bool nvtx;
  ProfScope(pi0_ctx* ctx, const char* label, cudaStream_t st) : c(ctx), s(st), e1(nullptr), nvtx(ctx->nvtx_on) {
    if (nvtx) nvtxRangePushA(label);
    if (!c->prof_on) return;
    cudaEvent_t e0;
    if (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess) { e1 = nullptr; return; }
    cudaEventRecord(e0, s);
    c->prof.push_back(pi0_ctx::ProfRec{label, e0, e1});
  }
  ~ProfScope() {
    if (e1) cudaEventRecord(e1, s);
    if (nvtx) nvtxRangePop();
  }
};
#define PROF(label) ProfScope _prof_scope(c, label, s)

// ---- workspace carving (shared by pi0_workspace_bytes and pi0_ctx_create) ------------------------------
// PI0_GUARD=1 (development / tests): every workspace buffer is followed by a 1 KiB guard zone filled with 0xA5 at ctx
// creation; pi0_debug_check_guards verifies them after a run. compute-sanitizer is closed on this GPU pool, so this is the
// bounds check of our own: any kernel writing past the end of its buffer (or before the start of the next) trips it.
constexpr size_t kGuardBytes = 1024;
inline bool guard_enabled() {
  const char* e = getenv("PI0_GUARD");
  return e && strcmp(e, "0") != 0;
}

struct Carver {
  size_t off = 0;
  uint8_t* base = nullptr;
  pi0_ctx* ctx = nullptr;
  bool guard = guard_enabled();
  void* take(const char* name, size_t bytes) {
    size_t o = off;
    off = round_up((int64_t)(off + bytes), 1024);
    if (guard) off += kGuardBytes;
    if (ctx && ctx->nbufs < 56) ctx->bufs[ctx->nbufs++] = Buf{name, o, bytes};
    return base ? base + o : nullptr;
  }
};

int validate_cfg(const pi0_model_cfg* c) {
  PI0_REQUIRE(c != nullptr, "null cfg");
  PI0_REQUIRE(c->n_cams >= 1 && c->n_cams <= 4, "n_cams=%d out of range [1,4]", c->n_cams);
  PI0_REQUIRE(c->lang_len >= 1 && c->lang_len <= 1024 && c->lang_len % 8 == 0,
              "lang_len=%d must be a positive multiple of 8", c->lang_len);
  PI0_REQUIRE(c->horizon >= 1 && c->horizon <= 64, "horizon=%d out of range [1,64]", c->horizon);
  PI0_REQUIRE(c->max_state_dim == 32 && c->max_action_dim == 32, "max_state_dim/max_action_dim must be 32");
  PI0_REQUIRE(c->num_steps >= 1 && c->num_steps <= 100, "num_steps=%d out of range", c->num_steps);
  PI0_REQUIRE(c->dtype == 0 || c->dtype == 1, "dtype=%d must be 0 (bf16) or 1 (fp32)", c->dtype);
  PI0_REQUIRE(c->max_batch >= 1, "max_batch=%d must be >= 1", c->max_batch);
  PI0_REQUIRE(c->n_siglip_layers >= 0 && c->n_siglip_layers <= PI0_SIGLIP_LAYERS, "n_siglip_layers=%d",
              c->n_siglip_layers);
  PI0_REQUIRE(c->n_gemma_layers >= 1 && c->n_gemma_layers <= PI0_GEMMA_LAYERS, "n_gemma_layers=%d",
              c->n_gemma_layers);
  return 0;
}

// fp32 regime (dtype 1): the GEMMs run on the tensor cores as split-bf16 products (gemm_x3.cu) unless PI0_FP32_GEMM=simt asks
// for the CUDA-core check kernel
bool fp32_x3_enabled() {
  const char* e = getenv("PI0_FP32_GEMM");
  return !(e && e[0] == 's');
}

void carve(Carver& cv, pi0_ctx* c, const pi0_model_cfg& cfg) {
  const size_t es = cfg.dtype == 1 ? 4 : 2;
  const int64_t B = cfg.max_batch;
  const int P = cfg.n_cams * kPatches + cfg.lang_len, S = 1 + cfg.horizon, H = cfg.horizon;
  const int kvp = P + S;
  const int scp = (int)round_up(kvp, 8), vtp = (int)round_up(kvp, 8);
  const int NL = cfg.n_gemma_layers;
  const int64_t Ms = B * kPatches, Mg = B * P, Me = B * S;
  pi0_ctx dummy;
  pi0_ctx& x = c ? *c : dummy;
  x.col = cv.take("siglip.col", Ms * kPatchK * es);
  x.xs = cv.take("siglip.x", Ms * kVisHidden * es);
  x.hs = cv.take("siglip.h", Ms * kVisHidden * es);
  x.qk_s = cv.take("siglip.qk", Ms * 3 * kVisHidden * es);  // q|k (+v on the fused-attention path)
  x.vt_s = cv.take("siglip.vt", B * kVisHeads * kVisHd * kPatches * es);
  x.sc_s = cv.take("siglip.scores", B * kVisHeads * kPatches * kPatches * es);
  x.att_s = cv.take("siglip.att", Ms * kVisHidden * es);
  x.mlp_s = cv.take("siglip.mlp", Ms * kVisInter * es);
  x.xp = cv.take("prefix.x", Mg * kHidden * es);
  x.hp = cv.take("prefix.h", Mg * kHidden * es);
  x.qp = cv.take("prefix.q", Mg * kHidden * es);
  x.sc_p = cv.take("prefix.scores", B * P * kHeads * scp * es);
  x.att_p = cv.take("prefix.att", Mg * kHidden * es);
  x.act_p = cv.take("prefix.act", Mg * kInter * es);
  x.kcache = cv.take("kcache", (int64_t)NL * B * kvp * kHd * es);
  x.vtcache = cv.take("vtcache", (int64_t)NL * B * kHd * vtp * es);
  x.suf = cv.take("suffix.embs", Me * kExpHidden * es);
  x.xe = cv.take("suffix.x", Me * kExpHidden * es);
  x.ye = cv.take("suffix.y", Me * kExpHidden * es);
  x.he = cv.take("suffix.h", Me * kExpHidden * es);
  x.qe = cv.take("suffix.q", Me * kHidden * es);
  x.sc_e = cv.take("suffix.scores", B * S * kHeads * round_up(kvp, 32) * es);
  x.att_e = cv.take("suffix.att", Me * kHidden * es);
  x.act_e = cv.take("suffix.act", Me * kExpInter * es);
  x.xt_act = cv.take("suffix.xt", B * H * 32 * es);
  x.st_act = cv.take("suffix.state", B * 32 * es);
  x.a1 = cv.take("suffix.a1", B * H * kProjW * es);
  x.a2 = cv.take("suffix.a2", B * H * kProjW * es);
  x.hfin = cv.take("suffix.hfin", B * H * kExpHidden * es);
  x.pad = (uint8_t*)cv.take("pad", B * P);
  x.prefix_pos = (int*)cv.take("prefix_pos", B * P * 4);
  x.suffix_pos = (int*)cv.take("suffix_pos", B * S * 4);
  x.suffix_cum = (int*)cv.take("suffix_cum", S * 4);
  x.mega_units = (int16_t*)cv.take("mega.units", (size_t)1024 * 4 * (kMegaMaxUnits + 1) * 2);
  x.mega_sync = (unsigned*)cv.take("mega.sync", 64);
  x.mega_sc = cv.take("mega.scores", (size_t)48 * 8 * 640 * 2);
  x.fused_descs = (FusedGemmDesc*)cv.take("fused.descs", fused_desc_bytes());
  x.fused_partial = (float*)cv.take("fused.partial", fused_partial_bytes((int)Me, cfg.num_sms > 0 ? cfg.num_sms : 148));
  x.fused_sync = (unsigned*)cv.take("fused.sync", 64);
  x.loop_units = (int16_t*)cv.take("loop.units", loop_units_bytes());
  x.loop_ll = (uint4*)cv.take("loop.ll", loop_ll_bytes());
  x.loop_epoch = (unsigned*)cv.take("loop.epoch", 64);
  x.rowsq_x = (float*)cv.take("suffix.rowsq_x", Me * (kExpHidden / 32) * 4);
  x.rowsq_y = (float*)cv.take("suffix.rowsq_y", Me * (kExpHidden / 32) * 4);
  x.x3_a = x.x3_w = x.x3_p = nullptr;
  x.x3_a_bytes = x.x3_w_bytes = x.x3_p_bytes = 0;
  if (cfg.dtype == 1 && fp32_x3_enabled()) {
    // largest A-side operand: the prefix down projection [B * P, 16384] (the attention products, batched over heads, and
    // the SigLIP MLP are smaller); largest W-side operand: the prefix gate/up weight [32768, 2048], or an activation used as
    // W (K cache, V^T, SigLIP k / v^T of all heads) at large batches. A problem that does not fit falls back to the SIMT kernel.
    const int64_t a_elems = std::max<int64_t>(std::max<int64_t>(Mg * kInter, Ms * (int64_t)(kVisInter + 8)),
                                              std::max<int64_t>(B * P * kHeads * round_up(kvp, 8), B * kVisHeads * kPatches * (int64_t)kPatches));
    const int64_t w_elems = std::max<int64_t>(std::max<int64_t>(2ll * kInter * kHidden, B * kVisHeads * kPatches * (int64_t)kPatches),
                                              B * kHd * (int64_t)round_up(kvp, 8) * 2);
    x.x3_a_bytes = (size_t)a_elems * 12;
    x.x3_w_bytes = (size_t)w_elems * 12;
    x.x3_a = cv.take("x3.a", x.x3_a_bytes);
    x.x3_w = cv.take("x3.w", x.x3_w_bytes);
    x.x3_p_bytes = (size_t)64 << 20;
    x.x3_p = cv.take("x3.partial", x.x3_p_bytes);
    const char* wc = getenv("PI0_FP32_WCACHE");
    if (!(wc && wc[0] == '0')) {
      // 12 bytes per GEMM weight element: SigLIP layers + patch embedding + projector, Gemma layers, expert layers + heads
      const int64_t sig = (int64_t)cfg.n_siglip_layers * ((int64_t)3 * kVisHidden * kVisHidden + (int64_t)kVisHidden * kVisHidden +
                                                          2ll * kVisInter * kVisHidden) +
                          (int64_t)kVisHidden * kPatchK + (int64_t)kHidden * kVisHidden;
      const int64_t gem = (int64_t)NL * ((int64_t)(kHeads + 2) * kHd * kHidden + (int64_t)kHidden * kHidden + 3ll * kInter * kHidden);
      const int64_t exp_ = (int64_t)NL * ((int64_t)(kHeads + 2) * kHd * kExpHidden + (int64_t)kExpHidden * kHidden +
                                          3ll * kExpInter * kExpHidden) + (8ll << 20);
      const size_t cache_bytes = (size_t)(sig + gem + exp_) * 12 + ((size_t)1024 << 10);
      uint8_t* cb = (uint8_t*)cv.take("x3.wcache", cache_bytes);
      if (c) { c->x3_cache.base = cb; c->x3_cache.bytes = cache_bytes; c->x3_cache.used = 0; c->x3_cache.n = 0; }
    }
  }
  if (c) {
    c->P = P; c->S = S; c->H = H; c->L = cfg.lang_len; c->ncam = cfg.n_cams; c->maxB = (int)B;
    c->kv_pitch = kvp; c->sc_pitch = scp; c->vt_pitch = vtp; c->es = (int)es;
  }
}

__global__ void suffix_cum_kernel(int* cum, int S) {
  pdl_trigger();
  pdl_wait();
  // att_masks of the suffix = [1, 1, 0, ...] (modeling_mvla.py:686-692) => cumsum = [1, 2, 2, ...]
  for (int i = threadIdx.x; i < S; i += blockDim.x) cum[i] = i == 0 ? 1 : 2;
}

inline void* at(void* p, int64_t elems, int es) { return (uint8_t*)p + elems * es; }
inline const void* at(const void* p, int64_t elems, int es) { return (const uint8_t*)p + elems * es; }

// The long-key tcgen05 attention (attn_long.cu) takes over where the one-tile kernels stop (P > 336 in the prefix pass; more
// than 6 suffix tokens or more than 352 keys in the denoise step) once a launch has enough (sample, 128-row tile) items to
// occupy the GPU; PI0_ATTN_LONG=0 disables, PI0_ATTN_LONG_MIN_ITEMS overrides the threshold (default 16).
bool attn_long_enabled() {
  const char* e = getenv("PI0_ATTN_LONG");
  return !(e && e[0] == '0');
}
int attn_long_min_items() {
  const char* e = getenv("PI0_ATTN_LONG_MIN_ITEMS");
  return e ? atoi(e) : 16;
}
bool use_long_suffix(const pi0_ctx* c, int B) {
  return attn_long_enabled() && attention_long_supported(c->S, c->P, c->S) &&
         (int64_t)B * ((c->S * 8 + 127) / 128) >= attn_long_min_items();
}
// prefix pass: the blocked two-pass kernel also beats the one-tile kernel at the shapes that one covers (bridge P = 328:
// 156 vs 174 us per launch at batch 64, tools/prefill_attn_compare.py), so it takes every prefix pass with at least
// PI0_ATTN_LONG_PREFIX_MIN_ITEMS (sample, tile) items (default 128, about one item per SM: at batch 1 - 21 items of ~20 us - the
// three-launch scores -> softmax -> P.V chain, which spreads the work over every SM, stays faster: 9.32 vs 9.49 ms end to end;
// at batch 4 the two are equal)
bool use_long_prefix(const pi0_ctx* c, int B, int T) {
  const int min_items = getenv("PI0_ATTN_LONG_PREFIX_MIN_ITEMS") ? atoi(getenv("PI0_ATTN_LONG_PREFIX_MIN_ITEMS")) : 128;
  return attn_long_enabled() && c->prefill_fused && attention_long_supported(T, c->P, 0) &&
         (int64_t)B * ((T * 8 + 127) / 128) >= min_items;
}

int run_gemm(pi0_ctx* c, const char* label, const GemmProblem& g, cudaStream_t s) {
  PROF(label);
  const int sms = c->cfg.num_sms > 0 ? c->cfg.num_sms : 148;
  if (c->cfg.dtype == 1 && c->x3_a != nullptr && c->cfg.gemm_impl == 0) {
    // W is a weight iff it lies outside the workspace (activations used as W - K cache, V^T, SigLIP k / v^T - live inside)
    const uint8_t* wp = (const uint8_t*)g.W;
    const bool is_weight = wp < c->base || wp >= c->base + c->ws_bytes;
    const int rc = gemm_x3_launch(g, c->x3_a, c->x3_a_bytes, c->x3_w, c->x3_w_bytes, c->x3_p, c->x3_p_bytes, sms, s,
                                  &c->x3_cache, is_weight);
    if (rc != 1) return rc;   // 1: not eligible (ragged N with a bias, operand larger than the scratch) -> SIMT kernel
  }
  static const bool w_pre = !(getenv("PI0_W_PREFETCH") && getenv("PI0_W_PREFETCH")[0] == '0');
  if (w_pre && c->cfg.dtype == 0) {
    // W outside the workspace = a weight (activations used as W - K cache, V^T - live inside): eligible for the early fetch
    const uint8_t* wp = (const uint8_t*)g.W;
    if (wp < c->base || wp >= c->base + c->ws_bytes) {
      GemmProblem gw = g;
      gw.w_const = 1;
      return gemm_launch(gw, c->cfg.dtype, c->cfg.gemm_impl, sms, s);
    }
  }
  return gemm_launch(g, c->cfg.dtype, c->cfg.gemm_impl, sms, s);
}

int run_rmsnorm(pi0_ctx* c, const char* label, const void* x, void* y, const float* g, int M, int D, int rpb_in, int rpb_out, int off,
                cudaStream_t s) {
  PROF(label);
  if (c->cfg.dtype == 1) return rmsnorm<float>((const float*)x, (float*)y, g, M, D, kEps, rpb_in, rpb_out, off, s);
  return rmsnorm<bf16>((const bf16*)x, (bf16*)y, g, M, D, kEps, rpb_in, rpb_out, off, s);
}
int run_layernorm(pi0_ctx* c, const char* label, const void* x, void* y, const float* w, const float* b, int M, int D,
                  cudaStream_t s) {
  PROF(label);
  if (c->cfg.dtype == 1) return layernorm<float>((const float*)x, (float*)y, w, b, M, D, kEps, s);
  return layernorm<bf16>((const bf16*)x, (bf16*)y, w, b, M, D, kEps, s);
}
int run_softmax(pi0_ctx* c, const char* label, void* sc, int64_t rows_total, int rows_per_b, int64_t pitch, int n_keys, int n_prefix,
                const uint8_t* pad, int pad_pitch, const int* cum, int heads, cudaStream_t s) {
  PROF(label);
  if (c->cfg.dtype == 1)
    return softmax_rows<float>((float*)sc, rows_total, rows_per_b, pitch, n_keys, n_prefix, pad, pad_pitch, cum, heads,
                               s);
  return softmax_rows<bf16>((bf16*)sc, rows_total, rows_per_b, pitch, n_keys, n_prefix, pad, pad_pitch, cum, heads, s);
}
int run_cast(pi0_ctx* c, const char* label, const float* in, void* out, int64_t n, cudaStream_t s) {
  PROF(label);
  if (c->cfg.dtype == 1) return cast_f32<float>(in, (float*)out, n, s);
  return cast_f32<bf16>(in, (bf16*)out, n, s);
}

// ---- one Gemma-style layer (prefill when suffix == false, expert over the cached prefix otherwise) ----
struct LayerIO {
  const void* x_in;  // residual stream in  [M, hidden]
  void* y;           // after attention     [M, hidden]
  void* x_out;       // after MLP           [M, hidden]
};

// x_stats: the rows of io.x_in come with per-chunk sums of squares in c->rowsq_x (left by the previous layer's down projection),
// so the input RMSNorm can be fused into the q|k|v GEMM
int gemma_layer(pi0_ctx* c, int B, int layer, bool suffix, const LayerIO& io, bool kv_only, cudaStream_t s,
                bool x_stats = false, bool* out_stats = nullptr) {
  const int es = c->es;
  const int T = suffix ? c->S : c->P;            // tokens per sample in this pass
  const int M = B * T;
  const int hidden = suffix ? kExpHidden : kHidden;
  const int inter = suffix ? kExpInter : kInter;
  const pi0_gemma_layer_w& w = suffix ? c->w.exp[layer] : c->w.gem[layer];
  void* h = suffix ? c->he : c->hp;
  void* q = suffix ? c->qe : c->qp;
  void* sc = suffix ? c->sc_e : c->sc_p;
  void* att = suffix ? c->att_e : c->att_p;
  void* act = suffix ? c->act_e : c->act_p;
  const int toff = suffix ? c->P : 0;
  static const char* kLbl[2][9] = {
      {"prefill.rmsnorm", "prefill.qkv_rope", "prefill.scores", "prefill.softmax", "prefill.pv", "prefill.o_proj",
       "prefill.rmsnorm", "prefill.gateup_geglu", "prefill.down"},
      {"denoise.rmsnorm", "denoise.qkv_rope", "denoise.scores", "denoise.softmax", "denoise.pv", "denoise.o_proj",
       "denoise.rmsnorm", "denoise.gateup_geglu", "denoise.down"}};
  const char* const* L = kLbl[suffix ? 1 : 0];
  const int n_keys = suffix ? c->P + c->S : c->P;
  void* kc = at(c->kcache, (int64_t)layer * c->maxB * c->kv_pitch * kHd, es);
  void* vt = at(c->vtcache, (int64_t)layer * c->maxB * kHd * c->vt_pitch, es);

  // latency regime (batch 1-3): weight-streaming mma.sync kernels with the norms fused in, see gemm_skinny.cu
  if (suffix && c->cfg.dtype == 0 && c->cfg.gemm_impl == 0 && skinny_supported(M) && c->S <= 48 &&
      c->P + c->S <= 640) {
    {
      PROF(L[1]);
      PI0_TRY(skinny_qkv_rope(io.x_in, hidden, w.g1, w.wqkv, M, hidden, q, c->w.rope_cos, c->w.rope_sin, c->suffix_pos,
                              kc, vt, c->S, c->P, c->kv_pitch, c->vt_pitch, s));
    }
    {
      PROF("denoise.attention");
      PI0_TRY(decode_attention(q, kc, vt, att, c->pad, c->suffix_cum, B, c->S, c->P, c->kv_pitch, c->vt_pitch,
                               c->cfg.num_sms > 0 ? c->cfg.num_sms : 148, s));
    }
    {
      PROF(L[5]);
      PI0_TRY(skinny_linear_res(att, kHidden, w.wo, M, hidden, kHidden, io.x_in, hidden, io.y, hidden, s));
    }
    {
      PROF(L[7]);
      PI0_TRY(skinny_gateup_geglu(io.y, hidden, w.g2, w.wgu, M, hidden, inter, act, inter, s));
    }
    {
      PROF(L[8]);
      PI0_TRY(skinny_linear_res(act, inter, w.wdown, M, hidden, inter, io.y, hidden, io.x_out, hidden, s));
    }
    return 0;
  }

  // Denoise step of a rollout batch on the tcgen05 per-op path: both RMSNorms of the layer are fused into the A operand of the
  // GEMM that consumes them (norm_a_operand in gemm_tc.cu); the row statistics come from the epilogues of the o_proj / down
  // projections that produce the streams. Saves two launches per layer-step (360 per denoise loop) but puts ~0.3 us per
  // k-block on the critical path of two latency-bound GEMMs: measured -2.6 % per infer at batch 8, -1 % at 16, 0 at 24 / 32,
  // +2 % at batch 64 (the stand-alone norm kernels hide behind PDL there). Hence the row limit (default 64 = batch 12).
  const bool fuse = suffix && c->norm_fuse && M <= c->norm_fuse_max_rows && c->cfg.dtype == 0 && c->cfg.gemm_impl == 0;
  const int slots = hidden / 32;
  if (out_stats) *out_stats = fuse;

  // input RMSNorm
  if (!(fuse && x_stats)) PI0_TRY(run_rmsnorm(c, L[0], io.x_in, h, w.g1, M, hidden, 0, 0, 0, s));

  // fused QKV + RoPE; K and V^T land in the cache (mllm.py:216-221 + fill_kv_cache)
  {
    GemmProblem g = fuse && x_stats ? gemm_problem(io.x_in, hidden, w.wqkv, hidden, M, (kHeads + 2) * kHd, hidden)
                                    : gemm_problem(h, hidden, w.wqkv, hidden, M, (kHeads + 2) * kHd, hidden);
    if (fuse && x_stats) { g.norm_gain = w.g1; g.norm_rowsq = c->rowsq_x; g.norm_slots = slots; g.norm_eps = kEps; }
    g.epi_mode = EPI_ROPE; g.bn = 256;
    g.epi.out = q; g.epi.ldc = kHidden;
    g.epi.rope_cos = c->w.rope_cos; g.epi.rope_sin = c->w.rope_sin;
    g.epi.pos_ids = suffix ? c->suffix_pos : c->prefix_pos;
    g.epi.kcache = kc; g.epi.kv_pitch = c->kv_pitch; g.epi.n_q_heads = kHeads;
    g.epi.vt = vt; g.epi.vt_pitch = c->vt_pitch; g.epi.vt_rpb = T; g.epi.vt_toff = toff;
    PI0_TRY(run_gemm(c, L[1], g, s));
  }
  if (kv_only) return 0;

  if (suffix && c->cfg.dtype == 0 && c->cfg.gemm_impl == 0 && c->S <= 48 && c->P + c->S <= 640) {
    // denoise step, bf16: scores + masked softmax + P.V in one launch over the cached prefix
    PROF("denoise.attention");
    // rollout batches: two wide launches up to B = 12 (latency-bound regime), one CTA pair per sample with K / V^T staged
    // in shared memory above (bandwidth-bound regime); the one-CTA-per-(sample, head group) kernel otherwise
    if (B > 96 && use_long_suffix(c, B))
      // more samples than ~2/3 of the SMs: the blocked tcgen05 kernel (one item per sample here) passes the batch kernel
      // (two CTAs per sample): 21.3 vs 24.3 us at B = 128, 38.4 vs 48.1 us at B = 256 (tools/suffix_attn_compare.py)
      PI0_TRY(attention_long(q, kc, vt, att, c->pad, c->suffix_cum, B, c->S, c->P, c->S, c->kv_pitch, c->vt_pitch,
                             c->cfg.num_sms > 0 ? c->cfg.num_sms : 148, s));
    else if (B > 12 && c->attn_split && decode_attention_batch_supported(c->S, c->P))
      PI0_TRY(decode_attention_batch(q, kc, vt, att, c->pad, c->suffix_cum, B, c->S, c->P, c->kv_pitch, c->vt_pitch, s));
    else if (B >= 4 && c->attn_split && decode_attention_split_supported(c->S, c->P))
      PI0_TRY(decode_attention_split(q, kc, vt, att, c->sc_e, c->pad, c->suffix_cum, B, c->S, c->P, c->kv_pitch,
                                     c->vt_pitch, s));
    else if (use_long_suffix(c, B))
      // long horizons / multi-camera prefixes (LIBERO: 17 suffix tokens over 601 keys): tcgen05, keys in blocks of 256
      PI0_TRY(attention_long(q, kc, vt, att, c->pad, c->suffix_cum, B, c->S, c->P, c->S, c->kv_pitch, c->vt_pitch,
                             c->cfg.num_sms > 0 ? c->cfg.num_sms : 148, s));
    else
      PI0_TRY(decode_attention(q, kc, vt, att, c->pad, c->suffix_cum, B, c->S, c->P, c->kv_pitch, c->vt_pitch,
                               c->cfg.num_sms > 0 ? c->cfg.num_sms : 148, s));
  } else if (suffix && c->cfg.dtype == 0 && c->cfg.gemm_impl == 0 && use_long_suffix(c, B)) {
    PROF("denoise.attention");
    PI0_TRY(attention_long(q, kc, vt, att, c->pad, c->suffix_cum, B, c->S, c->P, c->S, c->kv_pitch, c->vt_pitch,
                           c->cfg.num_sms > 0 ? c->cfg.num_sms : 148, s));
  } else if (!suffix && c->cfg.dtype == 0 && c->cfg.gemm_impl == 0 && use_long_prefix(c, B, T)) {
    // prefix pass, rollout batches: the blocked two-pass tcgen05 kernel of attn_long.cu (any P up to 1024)
    PROF("prefill.attention");
    PI0_TRY(attention_long(q, kc, vt, att, c->pad, nullptr, B, T, c->P, 0, c->kv_pitch, c->vt_pitch,
                           c->cfg.num_sms > 0 ? c->cfg.num_sms : 148, s));
  } else if (!suffix && c->prefill_fused && c->cfg.dtype == 0 && c->cfg.gemm_impl == 0 &&
             prefill_attention_supported(T, c->P) &&
             (int64_t)B * ((T * kHeads + 127) / 128) >= c->prefill_fused_min_tiles) {
    // prefix pass, bf16, rollout batches (>= 2 waves of 128-row tiles): scores + masked softmax + P.V of a layer in one
    // tcgen05 launch. A tile is a serial S -> softmax -> P.V chain of ~19 us, so with fewer tiles than SMs (batch 1: 21
    // tiles) the three-launch GEMM chain, which spreads the same work over every SM, is faster (2.30 vs 2.55 ms prefill)
    PROF("prefill.attention");
    PI0_TRY(prefill_attention(q, kc, vt, att, c->pad, B, T, c->P, c->kv_pitch, c->vt_pitch,
                              c->cfg.num_sms > 0 ? c->cfg.num_sms : 148, s));
  } else {
  // scores[b, t*8+h, j] = rnd(rnd(q . k) * 256^-0.5)   (MQA: the 8 heads of a token are 8 rows against one K)
  {
    GemmProblem g = gemm_problem(q, kHd, kc, kHd, T * kHeads, n_keys, kHd);
    g.nz0 = B; g.a_z0 = (int64_t)T * kHeads * kHd; g.w_z0 = (int64_t)c->kv_pitch * kHd;
    g.bn = 0;
    g.epi.out = sc; g.epi.ldc = c->sc_pitch; g.epi.out_z0 = (int64_t)T * kHeads * c->sc_pitch;
    g.epi.alpha = 0.0625f;
    PI0_TRY(run_gemm(c, L[2], g, s));
  }
  // masked softmax (mllm.py:299-307); prefix keys gated by pad, suffix keys by the block-causal cumsum rule
  PI0_TRY(run_softmax(c, L[3], sc, (int64_t)B * T * kHeads, T * kHeads, c->sc_pitch, n_keys, c->P, c->pad, c->P,
                      suffix ? c->suffix_cum : nullptr, kHeads, s));
  // att[b, t*8+h, :] = probs . V   (V^T is the K-major operand)
  {
    GemmProblem g = gemm_problem(sc, c->sc_pitch, vt, c->vt_pitch, T * kHeads, kHd, n_keys);
    g.nz0 = B; g.a_z0 = (int64_t)T * kHeads * c->sc_pitch; g.w_z0 = (int64_t)kHd * c->vt_pitch;
    g.bn = 256;
    g.epi.out = att; g.epi.ldc = kHd; g.epi.out_z0 = (int64_t)T * kHeads * kHd;
    PI0_TRY(run_gemm(c, L[4], g, s));
  }
  }
  // o_proj + residual
  {
    GemmProblem g = gemm_problem(att, kHidden, w.wo, kHidden, M, hidden, kHidden);
    g.bn = 0;
    g.epi.out = io.y; g.epi.ldc = hidden; g.epi.res = io.x_in; g.epi.ldr = hidden;
    if (fuse) { g.epi.rowsq_out = c->rowsq_y; g.epi.rowsq_ld = slots; }
    PI0_TRY(run_gemm(c, L[5], g, s));
  }
  if (!fuse) PI0_TRY(run_rmsnorm(c, L[6], io.y, h, w.g2, M, hidden, 0, 0, 0, s));
  // GeGLU: gate/up interleaved per 256-row tile at pack time
  {
    GemmProblem g = fuse ? gemm_problem(io.y, hidden, w.wgu, hidden, M, 2 * inter, hidden)
                         : gemm_problem(h, hidden, w.wgu, hidden, M, 2 * inter, hidden);
    if (fuse) { g.norm_gain = w.g2; g.norm_rowsq = c->rowsq_y; g.norm_slots = slots; g.norm_eps = kEps; }
    g.epi_mode = EPI_GEGLU; g.bn = 256;
    g.epi.out = act; g.epi.ldc = inter;
    PI0_TRY(run_gemm(c, L[7], g, s));
  }
  {
    GemmProblem g = gemm_problem(act, inter, w.wdown, inter, M, hidden, inter);
    g.bn = 0;
    g.epi.out = io.x_out; g.epi.ldc = hidden; g.epi.res = io.y; g.epi.ldr = hidden;
    if (fuse) { g.epi.rowsq_out = c->rowsq_x; g.epi.rowsq_ld = slots; }
    PI0_TRY(run_gemm(c, L[8], g, s));
  }
  return 0;
}

int check_batch(pi0_ctx* c, int B) {
  PI0_REQUIRE(c != nullptr, "null ctx");
  PI0_REQUIRE(B >= 1 && B <= c->maxB, "batch %d outside [1, max_batch=%d]", B, c->maxB);
  return 0;
}

}  // namespace

extern "C" {

size_t pi0_workspace_bytes(const pi0_model_cfg* cfg) {
  if (validate_cfg(cfg) != 0) return 0;
  Carver cv;
  carve(cv, nullptr, *cfg);
  return cv.off + 1024;
}

int pi0_ctx_create(const pi0_model_cfg* cfg, const pi0_weights* w, void* workspace, size_t workspace_bytes,
                   pi0_ctx** out) {
  PI0_TRY(validate_cfg(cfg));
  PI0_REQUIRE(w != nullptr && workspace != nullptr && out != nullptr, "pi0_ctx_create: null argument");
  const size_t need = pi0_workspace_bytes(cfg);
  PI0_REQUIRE(workspace_bytes >= need, "workspace too small: %zu < %zu bytes", workspace_bytes, need);
  PI0_REQUIRE(w->max_pos >= cfg->n_cams * kPatches + cfg->lang_len + 1 + cfg->horizon,
              "rope table covers %d positions, need %d", w->max_pos,
              cfg->n_cams * kPatches + cfg->lang_len + 1 + cfg->horizon);
  pi0_ctx* c = new (std::nothrow) pi0_ctx();
  PI0_REQUIRE(c != nullptr, "out of host memory");
  c->cfg = *cfg;
  c->w = *w;
  c->nbufs = 0;
  c->prof_on = false;
  c->nvtx_on = getenv("PI0_NVTX") && strcmp(getenv("PI0_NVTX"), "0") != 0;
  uintptr_t b = ((uintptr_t)workspace + 1023) & ~(uintptr_t)1023;
  c->base = (uint8_t*)b;
  c->ws_bytes = workspace_bytes;
  Carver cv;
  cv.base = c->base;
  cv.ctx = c;
  carve(cv, c, *cfg);
  c->guarded = cv.guard;
  if (c->guarded) {
    for (int i = 0; i < c->nbufs; ++i) {
      const size_t g0 = (size_t)round_up((int64_t)(c->bufs[i].offset + c->bufs[i].bytes), 1024);
      // the slack between the end of the buffer and the 1 KiB boundary belongs to the guard as well
      PI0_CHECK_CUDA(cudaMemset(c->base + c->bufs[i].offset + c->bufs[i].bytes, 0xA5,
                                g0 + kGuardBytes - (c->bufs[i].offset + c->bufs[i].bytes)));
    }
  }
  // persistent denoise kernel: deal the work units to one CTA per SM (bf16 product path only)
  c->mega_ctas = 0;
  {
    const char* e = getenv("PI0_ATTN_SPLIT");
    c->attn_split = !(e && strcmp(e, "0") == 0);
    const char* f = getenv("PI0_SIGLIP_FUSED");
    c->siglip_fused = !(f && strcmp(f, "0") == 0);
    const char* nf = getenv("PI0_NORM_FUSE");
    c->norm_fuse = !(nf && strcmp(nf, "0") == 0);
    c->norm_fuse_max_rows = getenv("PI0_NORM_FUSE_MAX_ROWS") ? atoi(getenv("PI0_NORM_FUSE_MAX_ROWS")) : 64;
    const char* pf = getenv("PI0_PREFILL_FUSED");
    c->prefill_fused = !(pf && strcmp(pf, "0") == 0);
    const char* pm = getenv("PI0_PREFILL_FUSED_MIN_TILES");
    c->prefill_fused_min_tiles = pm ? atoi(pm) : 2 * (cfg->num_sms > 0 ? cfg->num_sms : 148);
  }
  // PI0_MEGA: 0 = per-op launches, unset / 1 = per-step persistent kernel (grid barriers, denoise_mega.cu), 2 = flag-in-data
  // kernel launched per step (layers only), 3 = the whole Euler loop in one flag-in-data launch (denoise_loop.cu; measured
  // slower than 1 so far - 41 vs 28 us per layer - see DESIGN.md 4.2b, so it is opt-in)
  const char* mega_env = getenv("PI0_MEGA");
  const int mega_sel = mega_env ? atoi(mega_env) : 1;
  const int sms = cfg->num_sms > 0 ? cfg->num_sms : 148;
  c->loop_mode = 0;
  // measured in graph replay (full depth): batch 4 / 5 / 6 = 16.2 / 18.9 / 19.8 ms with the persistent kernel against 18.7 /
  // 19.8 / 20.7 ms per op, but 22.9 against 21.2 ms at batch 7 (three m16 tiles, five ring slots): it stops at 32 rows
  c->mega_max_rows = getenv("PI0_MEGA_MAX_ROWS") ? atoi(getenv("PI0_MEGA_MAX_ROWS")) : 32;
  c->loop_pace_ns = getenv("PI0_LOOP_PACE") ? atoi(getenv("PI0_LOOP_PACE")) : 0;
  if (cfg->dtype == 0 && cfg->gemm_impl == 0 && mega_sel != 0 && sms >= 64 && sms <= 1024) {
    std::vector<int16_t> tab;
    if (mega_build_units(sms, tab) == 0 &&
        cudaMemcpy(c->mega_units, tab.data(), tab.size() * sizeof(int16_t), cudaMemcpyHostToDevice) == cudaSuccess &&
        cudaMemset(c->mega_sync, 0, 64) == cudaSuccess) {
      c->mega_ctas = sms;
    } else {
      cudaGetLastError();
    }
    if (mega_sel >= 2 && sms >= 128 && sms <= 512 && loop_build_units(sms, tab) == 0 &&
        tab.size() * sizeof(int16_t) <= loop_units_bytes() &&
        cudaMemcpy(c->loop_units, tab.data(), tab.size() * sizeof(int16_t), cudaMemcpyHostToDevice) == cudaSuccess &&
        cudaMemset(c->loop_ll, 0, loop_ll_bytes()) == cudaSuccess && cudaMemset(c->loop_epoch, 0, 64) == cudaSuccess) {
      c->loop_mode = mega_sel == 2 ? 1 : 2;
    } else {
      cudaGetLastError();
    }
  }
  c->fused_on = false;
  c->fused_min_batch = getenv("PI0_FUSED_MIN_BATCH") ? atoi(getenv("PI0_FUSED_MIN_BATCH")) : 7;
  {
    const char* fe = getenv("PI0_FUSED");
    if (cfg->dtype == 0 && cfg->gemm_impl == 0 && fe && strcmp(fe, "1") == 0) {
      std::vector<FusedGemmDesc> descs;
      const void* wq[18]; const void* wo[18]; const void* wg[18]; const void* wd[18];
      for (int l = 0; l < cfg->n_gemma_layers; ++l) {
        wq[l] = w->exp[l].wqkv; wo[l] = w->exp[l].wo; wg[l] = w->exp[l].wgu; wd[l] = w->exp[l].wdown;
      }
      if (fused_build_descs(descs, cfg->n_gemma_layers, c->maxB * c->S, c->he, c->att_e, c->act_e, wq, wo, wg, wd) == 0 &&
          cudaMemcpy(c->fused_descs, descs.data(), descs.size() * sizeof(FusedGemmDesc), cudaMemcpyHostToDevice) == cudaSuccess &&
          cudaMemset(c->fused_sync, 0, 64) == cudaSuccess) {
        c->fused_on = true;
      } else {
        cudaGetLastError();
      }
    }
  }
  *out = c;
  return 0;
}

static void prof_clear(pi0_ctx* c) {
  for (auto& r : c->prof) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); }
  c->prof.clear();
}

void pi0_ctx_destroy(pi0_ctx* ctx) {
  if (ctx) prof_clear(ctx);
  delete ctx;
}

int pi0_profile(pi0_ctx* c, int enable) {
  PI0_REQUIRE(c != nullptr, "null ctx");
  prof_clear(c);
  c->prof_on = enable != 0;
  return 0;
}

int pi0_profile_report(pi0_ctx* c, char* buf, size_t buf_bytes) {
  PI0_REQUIRE(c != nullptr && buf != nullptr && buf_bytes > 2, "pi0_profile_report: bad arguments");
  std::map<std::string, std::pair<double, int>> agg;
  std::vector<std::string> order;
  for (auto& r : c->prof) {
    PI0_CHECK_CUDA(cudaEventSynchronize(r.e1));
    float ms = 0.f;
    PI0_CHECK_CUDA(cudaEventElapsedTime(&ms, r.e0, r.e1));
    auto it = agg.find(r.label);
    if (it == agg.end()) { agg[r.label] = {ms, 1}; order.push_back(r.label); }
    else { it->second.first += ms; it->second.second += 1; }
  }
  std::string out = "{";
  for (size_t i = 0; i < order.size(); ++i) {
    char tmp[160];
    snprintf(tmp, sizeof(tmp), "%s\"%s\": [%.6f, %d]", i ? ", " : "", order[i].c_str(), agg[order[i]].first,
             agg[order[i]].second);
    out += tmp;
  }
  out += "}";
  PI0_REQUIRE(out.size() + 1 <= buf_bytes, "pi0_profile_report: buffer too small (%zu needed)", out.size() + 1);
  memcpy(buf, out.c_str(), out.size() + 1);
  return 0;
}

int pi0_debug_check_guards(pi0_ctx* c, char* buf, size_t buf_bytes) {
  // Returns 0 when every guard zone still holds its pattern, 1 when one was overwritten (its buffer's name and the first bad
  // byte offset go to `buf`), < 0 on errors (guards not enabled: PI0_GUARD=1 before pi0_workspace_bytes / pi0_ctx_create).
  PI0_REQUIRE(c != nullptr && buf != nullptr && buf_bytes > 0, "pi0_debug_check_guards: bad arguments");
  PI0_REQUIRE(c->guarded, "pi0_debug_check_guards: the ctx was created without PI0_GUARD=1");
  buf[0] = 0;
  PI0_CHECK_CUDA(cudaDeviceSynchronize());
  std::vector<uint8_t> host(2 * kGuardBytes);
  for (int i = 0; i < c->nbufs; ++i) {
    const size_t end = c->bufs[i].offset + c->bufs[i].bytes;
    const size_t g1 = (size_t)round_up((int64_t)end, 1024) + kGuardBytes;
    const size_t n = g1 - end;
    PI0_CHECK_CUDA(cudaMemcpy(host.data(), c->base + end, n, cudaMemcpyDeviceToHost));
    for (size_t j = 0; j < n; ++j) {
      if (host[j] != 0xA5) {
        snprintf(buf, buf_bytes, "guard behind '%s' overwritten at +%zu bytes past its end (value 0x%02x)", c->bufs[i].name,
                 j, (unsigned)host[j]);
        return 1;
      }
    }
  }
  return 0;
}

int pi0_path_report(pi0_ctx* c, int B, char* buf, size_t buf_bytes) {
  // Which kernel variant every phase of pi0_infer takes at batch B with this ctx: the same predicates as gemma_layer /
  // pi0_siglip_forward / pi0_denoise_loop, so that silent fallbacks (P > 336, P + S > 352, S > 16, fp32 ...) are visible.
  PI0_TRY(check_batch(c, B));
  PI0_REQUIRE(buf != nullptr && buf_bytes > 2, "pi0_path_report: bad arguments");
  const bool bf = c->cfg.dtype == 0 && c->cfg.gemm_impl == 0;
  const int S = c->S, P = c->P, M = B * S;
  const int sms = c->cfg.num_sms > 0 ? c->cfg.num_sms : 148;
  const char* gemm = c->cfg.dtype == 1 && c->x3_a != nullptr && c->cfg.gemm_impl == 0
                         ? "tcgen05_split_bf16x3_fp32"
                         : (c->cfg.dtype == 1 || c->cfg.gemm_impl == 1 ? "simt_fp32_check" : "tcgen05");
  const char* siglip_attn = bf && c->siglip_fused ? (siglip_attention_use_tc() ? "fused_tcgen05" : "fused_mma_sync")
                                                  : "gemm_softmax_gemm";
  const bool pf = bf && c->prefill_fused && prefill_attention_supported(P, P) &&
                  (int64_t)B * ((P * kHeads + 127) / 128) >= c->prefill_fused_min_tiles;
  const char* prefill_attn = bf && use_long_prefix(c, B, P) ? "fused_tcgen05_long" : (pf ? "fused_tcgen05" : "gemm_softmax_gemm");
  const char* denoise;
  const char* denoise_attn;
  if (c->loop_mode > 0 && c->mega_ctas > 0 && loop_supported(M, S, P, c->mega_ctas)) {
    denoise = c->loop_mode == 2 ? "loop_kernel_whole_euler_loop" : "loop_kernel_per_step";
    denoise_attn = "in_kernel";
  } else if (c->mega_ctas > 0 && M <= c->mega_max_rows && mega_supported(M, S, P)) {
    denoise = "persistent_layers_kernel_per_step";
    denoise_attn = "in_kernel";
  } else if (c->fused_on && M > c->mega_max_rows && B >= c->fused_min_batch && fused_supported(M, S, P, B, sms)) {
    denoise = "fused_tcgen05_layers_kernel_per_step";
    denoise_attn = "in_kernel";
  } else if (bf && skinny_supported(M) && S <= 48 && P + S <= 640) {
    denoise = "skinny_mma_sync_per_op";
    denoise_attn = "fused_single_cta";
  } else {
    denoise = bf ? "tcgen05_per_op" : (c->cfg.dtype == 1 && c->x3_a != nullptr && c->cfg.gemm_impl == 0 ? "fp32_per_op" : "simt_fp32_check");
    if (bf && S <= 48 && P + S <= 640) {
      if (B > 96 && use_long_suffix(c, B)) denoise_attn = "fused_tcgen05_long";
      else if (B > 12 && c->attn_split && decode_attention_batch_supported(S, P)) denoise_attn = "batch_cta_pair_per_sample";
      else if (B >= 4 && c->attn_split && decode_attention_split_supported(S, P)) denoise_attn = "split_two_launches";
      else if (use_long_suffix(c, B)) denoise_attn = "fused_tcgen05_long";
      else denoise_attn = "fused_single_cta";
    } else {
      denoise_attn = bf && use_long_suffix(c, B) ? "fused_tcgen05_long" : "gemm_softmax_gemm";
    }
  }
  char tmp[640];
  const int n = snprintf(tmp, sizeof(tmp),
                         "{\"batch\": %d, \"prefix_tokens\": %d, \"suffix_tokens\": %d, \"sms\": %d, \"gemm\": \"%s\", "
                         "\"siglip_attention\": \"%s\", \"prefill_attention\": \"%s\", \"denoise\": \"%s\", "
                         "\"denoise_attention\": \"%s\"}",
                         B, P, S, sms, gemm, siglip_attn, prefill_attn, denoise, denoise_attn);
  PI0_REQUIRE(n > 0 && (size_t)n + 1 <= buf_bytes, "pi0_path_report: buffer too small (%d needed)", n + 1);
  memcpy(buf, tmp, (size_t)n + 1);
  return 0;
}

const void* pi0_debug_buffer(pi0_ctx* ctx, const char* name, int64_t* n_elems) {
  if (!ctx || !name) return nullptr;
  for (int i = 0; i < ctx->nbufs; ++i) {
    if (strcmp(ctx->bufs[i].name, name) == 0) {
      const bool raw = strcmp(name, "pad") == 0 || strcmp(name, "prefix_pos") == 0 ||
                       strcmp(name, "suffix_pos") == 0 || strcmp(name, "suffix_cum") == 0;
      const size_t unit = strcmp(name, "pad") == 0 ? 1 : (raw ? 4 : (size_t)ctx->es);
      if (n_elems) *n_elems = (int64_t)(ctx->bufs[i].bytes / unit);
      return ctx->base + ctx->bufs[i].offset;
    }
  }
  return nullptr;
}

int pi0_siglip_forward(pi0_ctx* c, int B, int cam, const float* images, float scale, float shift, void* stream) {
  PI0_TRY(check_batch(c, B));
  PI0_REQUIRE(cam >= 0 && cam < c->ncam, "camera index %d outside [0,%d)", cam, c->ncam);
  PI0_REQUIRE(images != nullptr, "null images");
  cudaStream_t s = (cudaStream_t)stream;
  const int es = c->es;
  const int M = B * kPatches;
  const pi0_weights& w = c->w;

  // patch embedding as a GEMM over the fused im2col / normalise (modeling_mvla.py:366-367 + Conv2d k14 s14),
  // bias and position embedding in the epilogue
  {
    PROF("siglip.im2col");
    if (c->cfg.dtype == 1) PI0_TRY(im2col_patch<float>(images, (float*)c->col, B, kPatchK, scale, shift, s));
    else PI0_TRY(im2col_patch<bf16>(images, (bf16*)c->col, B, kPatchK, scale, shift, s));
  }
  {
    GemmProblem g = gemm_problem(c->col, kPatchK, w.patch_w, kPatchK, M, kVisHidden, kPatchK);
    g.bn = 0;
    g.epi.out = c->xs; g.epi.ldc = kVisHidden; g.epi.bias = w.patch_b;
    g.epi.res = w.pos_emb; g.epi.ldr = kVisHidden; g.epi.res_row_mod = kPatches;
    PI0_TRY(run_gemm(c, "siglip.patch_embed", g, s));
  }
  const int64_t qk_ld = 2 * kVisHidden;
  for (int l = 0; l < c->cfg.n_siglip_layers; ++l) {
    const pi0_siglip_layer_w& lw = w.sig[l];
    PI0_TRY(run_layernorm(c, "siglip.layernorm", c->xs, c->hs, lw.ln1_w, lw.ln1_b, M, kVisHidden, s));
    const bool fused_attn = c->cfg.dtype == 0 && c->cfg.gemm_impl == 0 && c->siglip_fused;
    {  // fused q|k|v projection. Fused-attention path: plain [M, 3456] output. Per-op path: the V columns are written
       // transposed per head for the P.V GEMM
      GemmProblem g = gemm_problem(c->hs, kVisHidden, lw.wqkv, kVisHidden, M, 3 * kVisHidden, kVisHidden);
      g.bn = 0;
      g.epi.out = c->qk_s; g.epi.bias = lw.bqkv;
      if (fused_attn) {
        g.epi.ldc = 3 * kVisHidden;
      } else {
        g.epi.ldc = qk_ld;
        g.epi.vt = c->vt_s; g.epi.vt_col0 = 2 * kVisHidden; g.epi.vt_hd = kVisHd; g.epi.vt_heads = kVisHeads;
        g.epi.vt_rpb = kPatches; g.epi.vt_toff = 0; g.epi.vt_pitch = kPatches;
      }
      PI0_TRY(run_gemm(c, "siglip.qkv", g, s));
    }
    if (fused_attn) {
      // scores + softmax + P.V of all heads in one launch (attn_siglip.cu)
      PROF("siglip.attention");
      PI0_TRY(siglip_attention(c->qk_s, c->att_s, B, c->cfg.num_sms > 0 ? c->cfg.num_sms : 148, s));
    } else {
    {  // scores[b,h] = rnd(rnd(q k^T) * 72^-0.5), bidirectional
      GemmProblem g = gemm_problem(c->qk_s, qk_ld, at(c->qk_s, kVisHidden, es), qk_ld, kPatches, kPatches, kVisHd);
      g.nz0 = kVisHeads; g.nz1 = B;
      g.a_z0 = kVisHd; g.a_z1 = (int64_t)kPatches * qk_ld; g.w_z0 = kVisHd; g.w_z1 = (int64_t)kPatches * qk_ld;
      g.bn = 256;
      g.epi.out = c->sc_s; g.epi.ldc = kPatches; g.epi.out_z0 = (int64_t)kPatches * kPatches;
      g.epi.out_z1 = (int64_t)kVisHeads * kPatches * kPatches;
      g.epi.alpha = 0.11785113019775793f;  // 72 ** -0.5
      PI0_TRY(run_gemm(c, "siglip.scores", g, s));
    }
    PI0_TRY(run_softmax(c, "siglip.softmax", c->sc_s, (int64_t)B * kVisHeads * kPatches, kVisHeads * kPatches, kPatches, kPatches,
                        kPatches, nullptr, 0, nullptr, 1, s));
    {  // att[b, t, h*72 + d] = probs[b,h] . V[b,h]
      GemmProblem g = gemm_problem(c->sc_s, kPatches, c->vt_s, kPatches, kPatches, kVisHd, kPatches);
      g.nz0 = kVisHeads; g.nz1 = B;
      g.a_z0 = (int64_t)kPatches * kPatches; g.a_z1 = (int64_t)kVisHeads * kPatches * kPatches;
      g.w_z0 = (int64_t)kVisHd * kPatches; g.w_z1 = (int64_t)kVisHeads * kVisHd * kPatches;
      g.bn = 128;
      g.epi.out = c->att_s; g.epi.ldc = kVisHidden; g.epi.out_z0 = kVisHd;
      g.epi.out_z1 = (int64_t)kPatches * kVisHidden;
      PI0_TRY(run_gemm(c, "siglip.pv", g, s));
    }
    }
    {  // out_proj + residual (in place on the residual stream: each element is read then written by one thread)
      GemmProblem g = gemm_problem(c->att_s, kVisHidden, lw.wo, kVisHidden, M, kVisHidden, kVisHidden);
      g.bn = 0;
      g.epi.out = c->xs; g.epi.ldc = kVisHidden; g.epi.bias = lw.bo; g.epi.res = c->xs; g.epi.ldr = kVisHidden;
      PI0_TRY(run_gemm(c, "siglip.out_proj", g, s));
    }
    PI0_TRY(run_layernorm(c, "siglip.layernorm", c->xs, c->hs, lw.ln2_w, lw.ln2_b, M, kVisHidden, s));
    {
      GemmProblem g = gemm_problem(c->hs, kVisHidden, lw.wfc1, kVisHidden, M, kVisInter, kVisHidden);
      g.bn = 0;
      g.epi.out = c->mlp_s; g.epi.ldc = kVisInter; g.epi.bias = lw.bfc1; g.epi.act = ACT_GELU_TANH;
      PI0_TRY(run_gemm(c, "siglip.fc1_gelu", g, s));
    }
    {
      GemmProblem g = gemm_problem(c->mlp_s, kVisInter, lw.wfc2, kVisInter, M, kVisHidden, kVisInter);
      g.bn = 0;
      g.epi.out = c->xs; g.epi.ldc = kVisHidden; g.epi.bias = lw.bfc2; g.epi.res = c->xs; g.epi.ldr = kVisHidden;
      PI0_TRY(run_gemm(c, "siglip.fc2", g, s));
    }
  }
  PI0_TRY(run_layernorm(c, "siglip.layernorm", c->xs, c->hs, w.post_ln_w, w.post_ln_b, M, kVisHidden, s));
  {  // projector, "/ sqrt(2048)" of HF get_image_features and "* sqrt(2048)" of embed_prefix, written into the
     // prefix rows [b*P + cam*256, +256)
    GemmProblem g = gemm_problem(c->hs, kVisHidden, w.proj_w, kVisHidden, M, kHidden, kVisHidden);
    g.bn = 0;
    g.epi.out = c->xp; g.epi.ldc = kHidden; g.epi.bias = w.proj_b;
    g.epi.post_div = 45.254833995939045f;                    // sqrt(2048) in fp32
    g.epi.post_mul = c->cfg.dtype == 1 ? 45.254833995939045f : 45.25f;  // sqrt(2048) as a 0-dim tensor of the act dtype
    g.epi.rpb_in = kPatches; g.epi.rpb_out = c->P; g.epi.row_off = cam * kPatches;
    PI0_TRY(run_gemm(c, "siglip.projector", g, s));
  }
  return 0;
}

int pi0_embed_language(pi0_ctx* c, int B, const int64_t* tokens, void* stream) {
  PI0_TRY(check_batch(c, B));
  PI0_REQUIRE(tokens != nullptr, "null tokens");
  cudaStream_t s = (cudaStream_t)stream;
  const float scale = 45.254833995939045f;
  PROF("embed.language");
  if (c->cfg.dtype == 1)
    return embed_tokens<float>(tokens, (const float*)c->w.embed, (float*)c->xp, B, c->L, c->P, c->ncam * kPatches,
                               kHidden, c->w.vocab, scale, s);
  return embed_tokens<bf16>(tokens, (const bf16*)c->w.embed, (bf16*)c->xp, B, c->L, c->P, c->ncam * kPatches, kHidden,
                            c->w.vocab, scale, s);
}

int pi0_prefill(pi0_ctx* c, int B, const uint8_t* img_masks, const uint8_t* lang_masks, void* stream) {
  PI0_TRY(check_batch(c, B));
  PI0_REQUIRE(img_masks != nullptr && lang_masks != nullptr, "null masks");
  cudaStream_t s = (cudaStream_t)stream;
  {
  PROF("prefill.masks_posids");
  PI0_TRY(build_pad(img_masks, lang_masks, c->pad, B, c->ncam, c->L, s));
  PI0_TRY(position_ids(c->pad, B, c->P, c->S, c->prefix_pos, c->suffix_pos, s));
  PI0_CHECK_CUDA(launch_pdl(suffix_cum_kernel, dim3(1), dim3(64), 0, s, 1, c->suffix_cum, c->S));
  ++g_launch_count;
  }
  const int NL = c->cfg.n_gemma_layers;
  for (int l = 0; l < NL; ++l) {
    // the prefill output is discarded by sample_actions (modeling_mvla.py:776): the last layer only contributes K/V
    const bool kv_only = c->cfg.skip_dead_work != 0 && l == NL - 1;
    // the residual stream is updated in place (each epilogue thread reads then writes its own elements)
    LayerIO io{c->xp, c->xp, c->xp};
    PI0_TRY(gemma_layer(c, B, l, false, io, kv_only, s));
  }
  return 0;
}

int pi0_denoise_loop(pi0_ctx* c, int B, const float* state, float* x_t, void* stream) {
  PI0_TRY(check_batch(c, B));
  PI0_REQUIRE(state != nullptr && x_t != nullptr, "null state / x_t");
  cudaStream_t s = (cudaStream_t)stream;
  const int H = c->H, S = c->S;
  const pi0_weights& w = c->w;
  const float dt = -1.0f / (float)c->cfg.num_steps;

  const bool use_loop = c->loop_mode > 0 && c->mega_ctas > 0 && loop_supported(B * S, S, c->P, c->mega_ctas);
  const int sms_n = c->cfg.num_sms > 0 ? c->cfg.num_sms : 148;
  const bool use_fused = c->fused_on && !use_loop && B * S > c->mega_max_rows && B >= c->fused_min_batch &&
                         fused_supported(B * S, S, c->P, B, sms_n);
  auto fill_loop = [&](LoopParams& lp) {
    for (int l = 0; l < c->cfg.n_gemma_layers; ++l) {
      const pi0_gemma_layer_w& lw = w.exp[l];
      lp.L[l] = MegaLayerW{(const bf16*)lw.wqkv, (const bf16*)lw.wo, (const bf16*)lw.wgu, (const bf16*)lw.wdown, lw.g1,
                           lw.g2};
    }
    lp.n_layers = c->cfg.n_gemma_layers;
    lp.M = B * S; lp.S = S; lp.P = c->P; lp.B = B; lp.H = H;
    lp.state_w = (const bf16*)w.state_w; lp.action_in_w = (const bf16*)w.action_in_w;
    lp.time_in_wa = (const bf16*)w.time_in_wa; lp.time_out_w = (const bf16*)w.time_out_w;
    lp.action_out_w = (const bf16*)w.action_out_w;
    lp.state_b = w.state_b; lp.action_in_b = w.action_in_b; lp.time_bias = w.time_bias; lp.time_out_b = w.time_out_b;
    lp.action_out_b = w.action_out_b; lp.exp_norm_g = w.exp_norm_g;
    lp.state = state; lp.x_t = x_t; lp.dt = dt;
    lp.x0 = (const bf16*)c->suf; lp.x = (bf16*)c->xe;
    lp.kcache = (bf16*)c->kcache; lp.vtcache = (bf16*)c->vtcache;
    lp.kc_layer_stride = (int64_t)c->maxB * c->kv_pitch * kHd;
    lp.vt_layer_stride = (int64_t)c->maxB * kHd * c->vt_pitch;
    lp.kv_pitch = c->kv_pitch; lp.vt_pitch = c->vt_pitch;
    lp.rope_cos = w.rope_cos; lp.rope_sin = w.rope_sin; lp.pos_ids = c->suffix_pos; lp.pad = c->pad;
    lp.cum = c->suffix_cum; lp.units = c->loop_units; lp.ll = c->loop_ll; lp.epoch = c->loop_epoch;
    lp.kp = (c->P + S + 31) / 32 * 32;
    lp.pace_ns = c->loop_pace_ns;
  };
  const int n_iters = c->cfg.n_euler_iters > 0 ? c->cfg.n_euler_iters : c->cfg.num_steps;
  if (use_loop && c->loop_mode == 2) {
    // latency regime: embed_suffix, the 18 expert layers, the action head and the Euler update of ALL steps in one
    // persistent cooperative launch (denoise_loop.cu)
    PROF("denoise.loop_mega");
    LoopParams lp = {};
    fill_loop(lp);
    lp.n_steps = n_iters;
    lp.layers_only = 0;
    return loop_denoise(lp, c->mega_ctas, s);
  }

  // state_proj -> suffix row 0 of every sample (input-independent across steps)
  PI0_TRY(run_cast(c, "denoise.cast", state, c->st_act, (int64_t)B * 32, s));
  {
    GemmProblem g = gemm_problem(c->st_act, 32, w.state_w, 32, B, kExpHidden, 32);
    g.bn = 0;
    g.epi.out = c->suf; g.epi.ldc = kExpHidden; g.epi.bias = w.state_b;
    g.epi.rpb_in = 1; g.epi.rpb_out = S; g.epi.row_off = 0;
    PI0_TRY(run_gemm(c, "denoise.state_proj", g, s));
  }
  for (int step = 0; step < n_iters; ++step) {
    // embed_suffix: action_in_proj -> (time half folded into the bias) mlp_in -> SiLU -> mlp_out -> suffix rows 1..H
    PI0_TRY(run_cast(c, "denoise.cast", x_t, c->xt_act, (int64_t)B * H * 32, s));
    {
      GemmProblem g = gemm_problem(c->xt_act, 32, w.action_in_w, 32, B * H, kProjW, 32);
      g.bn = 0;
      g.epi.out = c->a1; g.epi.ldc = kProjW; g.epi.bias = w.action_in_b;
      PI0_TRY(run_gemm(c, "denoise.action_in", g, s));
    }
    {
      GemmProblem g = gemm_problem(c->a1, kProjW, w.time_in_wa, kProjW, B * H, kProjW, kProjW);
      g.bn = 0;
      g.epi.out = c->a2; g.epi.ldc = kProjW; g.epi.bias = w.time_bias + (int64_t)step * kProjW;
      g.epi.act = ACT_SILU;
      PI0_TRY(run_gemm(c, "denoise.time_mlp_in_silu", g, s));
    }
    {
      GemmProblem g = gemm_problem(c->a2, kProjW, w.time_out_w, kProjW, B * H, kExpHidden, kProjW);
      g.bn = 0;
      g.epi.out = c->suf; g.epi.ldc = kExpHidden; g.epi.bias = w.time_out_b;
      g.epi.rpb_in = H; g.epi.rpb_out = S; g.epi.row_off = 1;
      PI0_TRY(run_gemm(c, "denoise.time_mlp_out", g, s));
    }
    if (use_loop) {
      // PI0_MEGA=2: the flag-in-data kernel as a per-step drop-in for the round-1 kernel (A/B tests)
      PROF("denoise.layers_mega");
      LoopParams lp = {};
      fill_loop(lp);
      lp.n_steps = 1;
      lp.layers_only = 1;
      PI0_TRY(loop_denoise(lp, c->mega_ctas, s));
    } else if (use_fused) {
      // rollout batches: all expert layers of this step in one persistent tcgen05 launch (fused_denoise.cu)
      PROF("denoise.layers_fused");
      FusedParams fp = {};
      fp.gd = c->fused_descs;
      for (int l = 0; l < c->cfg.n_gemma_layers; ++l) { fp.g1[l] = w.exp[l].g1; fp.g2[l] = w.exp[l].g2; }
      fp.n_layers = c->cfg.n_gemma_layers;
      fp.M = B * S; fp.S = S; fp.P = c->P; fp.B = B;
      fp.x0 = (const bf16*)c->suf; fp.x = (bf16*)c->xe; fp.y = (bf16*)c->ye; fp.h = (bf16*)c->he; fp.q = (bf16*)c->qe;
      fp.att = (bf16*)c->att_e; fp.act = (bf16*)c->act_e;
      fp.partial = c->fused_partial;
      fp.partial_elems = (long long)(fused_partial_bytes(c->maxB * S, sms_n) / sizeof(float));
      fp.kcache = (bf16*)c->kcache; fp.vtcache = (bf16*)c->vtcache;
      fp.kc_layer_stride = (long long)c->maxB * c->kv_pitch * kHd;
      fp.vt_layer_stride = (long long)c->maxB * kHd * c->vt_pitch;
      fp.kv_pitch = c->kv_pitch; fp.vt_pitch = c->vt_pitch;
      fp.rope_cos = w.rope_cos; fp.rope_sin = w.rope_sin; fp.pos_ids = c->suffix_pos; fp.pad = c->pad; fp.cum = c->suffix_cum;
      fp.sync = c->fused_sync;
      fp.kp = (c->P + S + 31) / 32 * 32;
      fp.hpg = B * 2 <= sms_n ? 4 : 8;
      PI0_TRY(fused_denoise_layers(fp, sms_n, s));
    } else if (c->mega_ctas > 0 && B * S <= c->mega_max_rows && mega_supported(B * S, S, c->P)) {
      // latency regime: all expert layers of this step in one persistent cooperative launch
      PROF("denoise.layers_mega");
      MegaParams mp = {};
      for (int l = 0; l < c->cfg.n_gemma_layers; ++l) {
        const pi0_gemma_layer_w& lw = w.exp[l];
        mp.L[l] = MegaLayerW{(const bf16*)lw.wqkv, (const bf16*)lw.wo, (const bf16*)lw.wgu, (const bf16*)lw.wdown,
                             lw.g1, lw.g2};
      }
      mp.n_layers = c->cfg.n_gemma_layers;
      mp.M = B * S; mp.S = S; mp.P = c->P; mp.B = B;
      mp.x0 = (const bf16*)c->suf; mp.x = (bf16*)c->xe; mp.y = (bf16*)c->ye; mp.q = (bf16*)c->qe;
      mp.att = (bf16*)c->att_e; mp.act = (bf16*)c->act_e;
      mp.kcache = (bf16*)c->kcache; mp.vtcache = (bf16*)c->vtcache;
      mp.kc_layer_stride = (int64_t)c->maxB * c->kv_pitch * kHd;
      mp.vt_layer_stride = (int64_t)c->maxB * kHd * c->vt_pitch;
      mp.kv_pitch = c->kv_pitch; mp.vt_pitch = c->vt_pitch;
      mp.rope_cos = w.rope_cos; mp.rope_sin = w.rope_sin; mp.pos_ids = c->suffix_pos; mp.pad = c->pad;
      mp.cum = c->suffix_cum; mp.units = c->mega_units; mp.sync = c->mega_sync;
      mp.kp = (c->P + S + 31) / 32 * 32; mp.sc = (bf16*)c->mega_sc;
      PI0_TRY(mega_denoise_layers(mp, c->mega_ctas, s));
    } else {
      bool stats = false;   // layer 0 reads the suffix embeddings, which come without row statistics
      for (int l = 0; l < c->cfg.n_gemma_layers; ++l) {
        LayerIO io{l == 0 ? c->suf : c->xe, c->ye, c->xe};
        bool produced = false;
        PI0_TRY(gemma_layer(c, B, l, true, io, false, s, stats, &produced));
        stats = produced;
      }
    }
    // final norm on the last H rows of every sample, action_out_proj and the Euler update x_t += dt * v_t
    PI0_TRY(run_rmsnorm(c, "denoise.final_norm", c->xe, c->hfin, w.exp_norm_g, B * H, kExpHidden, H, S, 1, s));
    {
      GemmProblem g = gemm_problem(c->hfin, kExpHidden, w.action_out_w, kExpHidden, B * H, 32, kExpHidden);
      g.epi_mode = EPI_EULER; g.bn = 32;
      g.epi.out = x_t; g.epi.ldc = 32; g.epi.bias = w.action_out_b; g.epi.dt = dt;
      PI0_TRY(run_gemm(c, "denoise.out_proj_euler", g, s));
    }
  }
  return 0;
}

int pi0_infer(pi0_ctx* c, int B, const float* const* images, float img_scale, float img_shift,
              const uint8_t* img_masks, const int64_t* tokens, const uint8_t* lang_masks, const float* state,
              float* x_t, void* stream) {
  PI0_TRY(check_batch(c, B));
  PI0_REQUIRE(images != nullptr, "null images");
  for (int cam = 0; cam < c->ncam; ++cam)
    PI0_TRY(pi0_siglip_forward(c, B, cam, images[cam], img_scale, img_shift, stream));
  PI0_TRY(pi0_embed_language(c, B, tokens, stream));
  PI0_TRY(pi0_prefill(c, B, img_masks, lang_masks, stream));
  return pi0_denoise_loop(c, B, state, x_t, stream);
}

}  // extern "C"
