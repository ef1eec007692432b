// Whole-loop persistent kernel for the latency regime of the denoise loop (M = B * S <= 16 suffix rows, i.e. batch 1-3
// at the bridge shapes): ALL Euler steps x (embed_suffix + 18 action-expert layers + action head + Euler update) in ONE
// cooperative launch, one CTA per SM.
//
// Round 1 (denoise_mega.cu) kept one CTA per SM resident for the 18 layers of ONE step and separated the six phases of
// a layer with grid-wide barriers: 28 us per layer against 5.3 us of weight streaming (19 % of HBM bandwidth), because
// each exchange cost a release fence + an arrival atomic + an acquire poll + a dependent re-read of the activations.
// Here every cross-CTA activation travels as FLAG-IN-DATA words (the "LL" protocol of collective libraries): a 16-byte
// word is two 8-byte halves {4 bytes of payload, 4 bytes of epoch}; 8-byte accesses are single-copy atomic, so a reader
// that sees the expected epoch in a half has its payload - no fence, no barrier, no second round trip. A consumer polls
// exactly the words it needs (the whole activation vector for a GEMM phase, one head for an attention unit), so the
// exchange costs one L2 write + one L2 read and late producers only delay their own consumers. Epochs are taken from
// a device counter that the kernel advances at exit, so CUDA-graph replays need no memset and stale words never match.
//   * warp 8 of every CTA is the weight PRODUCER (unchanged from round 1): 1-D bulk TMA copies of 8 weight rows x 1024 k
//     into a shared-memory ring, L2 evict-first, throttled only by the ring - HBM keeps streaming through every exchange
//     and across step boundaries;
//   * warps 0-7 are CONSUMERS: poll + stage the activation operand (GemmaRMSNorm fused into the staging), mma.sync over
//     the chunks as they land, fixed-order cross-warp reduction, epilogue (RoPE + cache write, residual, GeGLU, SiLU,
//     Euler update) that publishes its 8 output columns as LL words;
//   * the per-step projections (action_in_proj, action_time_mlp_in + SiLU with the time half folded into a per-step
//     bias, action_time_mlp_out, final norm + action_out_proj + x_t += dt * v_t) are phases of the same kernel, so a
//     batch-1 denoise loop is 1 launch instead of 70.
// Arithmetic and rounding points are those of denoise_mega.cu / gemm_skinny.cu / attn_decode.cu; with
// layers_only = 1 the kernel is a drop-in for mega_denoise_layers and matches it bit for bit (tests/test_denoise_loop_gpu.py).
// Reference arithmetic: [mirror] src/model/mvla/modeling_mvla.py:643-694,787-839; expert.py:128-179,387-409;
// mllm.py:30-53,265-317.
#include <algorithm>
#include <vector>

#include "attn_split.cuh"
#include "denoise_loop.cuh"

namespace pi0 {

namespace {

constexpr int kThreads = 384;                   // 3 warpgroups: 8 consumer warps + the producer warp (and 3 idle warps)
constexpr int kChunkK = 1024;                   // k elements per streamed chunk
constexpr int kRowPitch = kChunkK + 32;         // bf16 elements; 2112 B = 64 mod 128 (conflict-free 16-byte loads)
constexpr int kSlotBytes = 8 * kRowPitch * 2;   // 16896
constexpr int kAK = 4096;                       // largest K (down projection)
constexpr int kAPitch = kAK + 32;               // 8256 B = 64 mod 128
constexpr int kHidden = 1024, kInter = 4096, kQDim = 2048;
constexpr int kMaxRows = 16;
constexpr int kTabW = kMegaMaxUnits + 1;

#ifdef PI0_TIMELINE
__device__ unsigned long long g_loop_tl[4][32];
#define LOOP_STAMP(i) do { if (gl == 1 && threadIdx.x == 0 && (cta == 0 || cta == 7 || cta == 64 || cta == 140)) \
  g_loop_tl[cta == 0 ? 0 : cta == 7 ? 1 : cta == 64 ? 2 : 3][i] = global_timer_ns(); } while (0)
#else
#define LOOP_STAMP(i) do { } while (0)
#endif

// ---- flag-in-data exchange arena -------------------------------------------------------------------------------------
// Two sets (alternating per producing layer / step) of one buffer per activation kind; sizes in 16-byte words = 4 bf16
// (or 2 fp32 for x_t) each, dimensioned for 16 rows and 384 keys.
enum LLBuf : int { LL_X = 0, LL_Q, LL_KN, LL_VN, LL_SC, LL_ATT, LL_Y, LL_ACT, LL_A1, LL_A2, LL_XT, LL_NBUF };
__host__ __device__ constexpr int ll_words(int b) {
  return b == LL_X ? kMaxRows * 256 : b == LL_Q ? kMaxRows * 512 : b == LL_KN ? kMaxRows * 64 : b == LL_VN ? kMaxRows * 64
       : b == LL_SC ? kMaxRows * 8 * (32 * kMegaKeyChunks) / 4 : b == LL_ATT ? kMaxRows * 512 : b == LL_Y ? kMaxRows * 256
       : b == LL_ACT ? kMaxRows * 1024 : b == LL_A1 ? kMaxRows * 256 : b == LL_A2 ? kMaxRows * 256 : kMaxRows * 16;
}
__host__ __device__ constexpr int ll_offset(int b) {
  int o = 0;
  for (int i = 0; i < b; ++i) o += ll_words(i);
  return o;
}
constexpr int kLLSetWords = ll_offset(LL_NBUF);

struct LLW { uint64_t a, b; };  // two {payload : 32, epoch : 32} halves

__device__ __forceinline__ void ll_st(uint4* p, uint32_t d0, uint32_t d1, uint32_t f) {
  const uint64_t a = (uint64_t)d0 | ((uint64_t)f << 32), b = (uint64_t)d1 | ((uint64_t)f << 32);
  asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(a), "l"(b) : "memory");
}
__device__ __forceinline__ LLW ll_ld(const uint4* p) {
  LLW w;
  asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w.a), "=l"(w.b) : "l"(p) : "memory");
  return w;
}
__device__ __forceinline__ bool ll_ok(const LLW& w, uint32_t f) {
  return (uint32_t)(w.a >> 32) == f && (uint32_t)(w.b >> 32) == f;
}
#ifdef PI0_TIMELINE
__device__ unsigned long long g_loop_stats[4];  // polls issued, polls that missed, retry rounds, gathers
#define LOOP_STAT(i, v) atomicAdd(&g_loop_stats[i], (unsigned long long)(v))
#else
#define LOOP_STAT(i, v) do { } while (0)
#endif

// Bounded spins: a protocol bug must surface as a launch failure, never as a hung GPU. No printf here: a call (vprintf
// follows the ABI) anywhere inside a polling loop makes ptxas spill every value that is live across it AT ITS DEFINITION,
// i.e. a store to the stack right behind each LL load, which serialised the batched loads (8.7 us instead of 1.5 us to stage
// 80 KB). The failing word's address / epoch go to a debug slot next to the epoch word instead.
__device__ unsigned long long g_loop_fail[4];
__device__ __forceinline__ void ll_timeout(const uint4* p, uint32_t f, uint32_t saw_a, uint32_t saw_b) {
  g_loop_fail[0] = (unsigned long long)p;
  g_loop_fail[1] = ((unsigned long long)f << 32) | saw_a;
  g_loop_fail[2] = ((unsigned long long)blockIdx.x << 32) | threadIdx.x;
  g_loop_fail[3] = saw_b;
  __threadfence_system();
  asm volatile("trap;");
}
__device__ __forceinline__ void mbar_wait_quiet(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  uint64_t t0 = 0;
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (((++spins) & 0xfff) == 0) {
      if (t0 == 0) t0 = global_timer_ns();
      else if (global_timer_ns() - t0 > 4000000000ull) asm volatile("trap;");
    }
  }
}
// Polls N words together: every round re-issues ALL still-missing loads back to back (one L2 round trip per round
// whatever N is), `addr(j)` gives the address of word j, bit j of `need` says whether word j is wanted.
template <int N, typename AddrFn>
__device__ __forceinline__ void ll_gather(LLW (&w)[N], uint32_t need, uint32_t f, AddrFn&& addr) {
#pragma unroll
  for (int j = 0; j < N; ++j)
    if ((need >> j) & 1u) w[j] = ll_ld(addr(j));
#pragma unroll
  for (int j = 0; j < N; ++j)
    if (((need >> j) & 1u) && ll_ok(w[j], f)) need &= ~(1u << j);
  LOOP_STAT(3, 1);
  if (need == 0) return;
  LOOP_STAT(1, __popc(need));
  uint32_t rounds = 0;
  uint64_t t0 = 0;
  do {
#pragma unroll
    for (int j = 0; j < N; ++j)
      if ((need >> j) & 1u) w[j] = ll_ld(addr(j));
#pragma unroll
    for (int j = 0; j < N; ++j)
      if (((need >> j) & 1u) && ll_ok(w[j], f)) need &= ~(1u << j);
    if (((++rounds) & 0x3ff) == 0) {
      if (t0 == 0) t0 = global_timer_ns();
      else if (global_timer_ns() - t0 > 2000000000ull) {
#pragma unroll
        for (int j = 0; j < N; ++j)
          if ((need >> j) & 1u) ll_timeout(addr(j), f, (uint32_t)(w[j].a >> 32), (uint32_t)(w[j].b >> 32));
      }
    }
  } while (need);
  LOOP_STAT(2, rounds);
}
__device__ __forceinline__ LLW ll_wait(const uint4* p, uint32_t f) {
  LLW w[1];
  ll_gather<1>(w, 1u, f, [&](int) { return p; });
  return w[0];
}
// one bf16 element `e` (index in elements) of a bf16 buffer
__device__ __forceinline__ float ll_elem_bf16(const uint4* buf, int64_t e, uint32_t f) {
  const LLW w = ll_wait(buf + (e >> 2), f);
  const uint32_t d = (e & 2) ? (uint32_t)w.b : (uint32_t)w.a;
  return __uint_as_float((e & 1) ? (d & 0xffff0000u) : (d << 16));
}
__device__ __forceinline__ uint32_t bf16_bits(float x) { return bf16x2_bits(x, 0.0f) >> 16; }

// The 8 lanes of a row hold 8 consecutive bf16 columns (lane `col` = column n0 + col, n0 % 8 == 0): lanes 0 and 4 publish
// four columns each. Must be executed by all 32 lanes of the warp.
__device__ __forceinline__ void ll_publish4(uint4* buf, int64_t e, int col, uint32_t bits, bool valid, uint32_t f) {
  const uint32_t b1 = __shfl_down_sync(0xffffffffu, bits, 1), b2 = __shfl_down_sync(0xffffffffu, bits, 2),
                 b3 = __shfl_down_sync(0xffffffffu, bits, 3);
  if (valid && (col & 3) == 0) ll_st(buf + (e >> 2), bits | (b1 << 16), b2 | (b3 << 16), f);
}

template <bool M16>
struct LoopCfg {
  static constexpr int kRows = M16 ? 16 : 8;
  static constexpr int kSlots = M16 ? 4 : 8;
  static constexpr int kRingBytes = kSlots * kSlotBytes;
  static constexpr int kABytes = kRows * kAPitch * 2;
  static constexpr int kRedBytes = 2 * 8 * 2 * 16 * 8 * 4;  // [parity][warp][acc][row][col] fp32
  static constexpr int kSmem = kRingBytes + kABytes + kRedBytes + 256 + 256 + 128;
};

__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
      ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// weight row `j` (0..7) of chunk `ch` of unit `u` in phase `ph`
__device__ __forceinline__ int unit_row(int ph, int u, int ch, int j) {
  if (ph == LOOP_QKV) return (u >> 4) * 256 + (u & 15) * 8 + ch * 128 + j;  // u = head*16 + dc; ch = RoPE half
  if (ph == LOOP_GU) {
    const int i = u * 8 + j;  // inter column; packed per 256 rows as 128 gate rows then the 128 matching up rows
    return (i >> 7) * 256 + (i & 127) + ch * 128;
  }
  return u * 8 + j;  // 8 consecutive output columns, chunks run along K
}

template <bool M16>
__device__ __forceinline__ void mma_chunk(float* acc, const bf16* As, int kc, const bf16* slot, int warp, int g,
                                          int t4) {
  const int k0 = warp * 128 + t4 * 8;
#pragma unroll
  for (int cc = 0; cc < 4; ++cc) {
    const uint4 w = *reinterpret_cast<const uint4*>(slot + g * kRowPitch + k0 + cc * 32);
    const uint4 lo = *reinterpret_cast<const uint4*>(As + g * kAPitch + kc * kChunkK + k0 + cc * 32);
    uint4 hi = make_uint4(0, 0, 0, 0);
    if constexpr (M16) hi = *reinterpret_cast<const uint4*>(As + (g + 8) * kAPitch + kc * kChunkK + k0 + cc * 32);
    mma_bf16_16816(acc, lo.x, hi.x, lo.y, hi.y, w.x, w.y);
    mma_bf16_16816(acc, lo.z, hi.z, lo.w, hi.w, w.z, w.w);
  }
}

// A = rnd((x * rsqrt(mean(x^2) + eps)) * (1 + w)) for M rows of 1024, x polled from an LL buffer -> shared memory.
// Warp w normalises rows w, w + 8; same summation order as denoise_mega.cu's stage_norm.
template <bool M16>
__device__ __noinline__ void stage_norm_ll(bf16* As, const uint4* X, uint32_t f, const float* gain, int M) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int rr = 0; rr < (M16 ? 2 : 1); ++rr) {
    const int row = warp + rr * 8;
    if (row >= M) continue;
    const uint4* base = X + row * 256 + lane * 2;
    LLW w[8];
    ll_gather<8>(w, 0xffu, f, [&](int j) { return base + (j >> 1) * 64 + (j & 1); });
    float fv[4][8];
    float ss = 0.0f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint32_t u[4] = {(uint32_t)w[2 * i].a, (uint32_t)w[2 * i].b, (uint32_t)w[2 * i + 1].a,
                             (uint32_t)w[2 * i + 1].b};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u[j]));
        fv[i][2 * j] = t.x;
        fv[i][2 * j + 1] = t.y;
        ss += t.x * t.x + t.y * t.y;
      }
    }
    ss = warp_sum_f(ss);
    const float r = rsqrtf(ss / (float)kHidden + 1e-6f);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int k = (i * 32 + lane) * 8;
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gain + k));
      const float4 g1 = __ldg(reinterpret_cast<const float4*>(gain + k + 4));
      const float gv[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
      uint4 o;
      __nv_bfloat162* oh = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
      for (int j = 0; j < 4; ++j)
        oh[j] = __floats2bfloat162_rn((fv[i][2 * j] * r) * gv[2 * j], (fv[i][2 * j + 1] * r) * gv[2 * j + 1]);
      *reinterpret_cast<uint4*>(As + row * kAPitch + k) = o;
    }
  }
}

// plain copy of M rows x K bf16 (K multiple of 4) from an LL buffer into the staged operand; 8 polls in flight per thread
__device__ __noinline__ void stage_copy_ll(bf16* As, const uint4* buf, uint32_t f, int M, int K) {
  const int per_row = K >> 2;
  const int total = M * per_row;
  for (int i0 = threadIdx.x; i0 < total; i0 += 256 * 8) {
    LLW w[8];
    uint32_t need = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (i0 + j * 256 < total) need |= 1u << j;
    ll_gather<8>(w, need, f, [&](int j) { return buf + i0 + j * 256; });
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int i = i0 + j * 256;
      if (i < total) {
        const int row = i / per_row, c4 = i - row * per_row;
        *reinterpret_cast<uint2*>(As + row * kAPitch + c4 * 4) = make_uint2((uint32_t)w[j].a, (uint32_t)w[j].b);
      }
    }
  }
}

__device__ __forceinline__ uint4 ll_pack8(const LLW& a, const LLW& b) {
  return make_uint4((uint32_t)a.a, (uint32_t)a.b, (uint32_t)b.a, (uint32_t)b.b);
}

// Everything the attention units need, passed BY VALUE to the (deliberately not inlined) unit functions: each gets its
// own register allocation instead of competing with the ~180 long-lived registers of the kernel body.
struct AttnCtx {
  const uint4 *Q, *KN, *VN;
  uint4* SC;
  uint4* ATT;
  uint32_t fq, fsc, fatt;
  int S, P, kp, kv_pitch, vt_pitch;
  const uint8_t* pad;
  const int* cum;
};

// scores of one (sample, head, 16-key block), one warp: rnd(rnd(q.k) * 256^-0.5) -> LL score buffer [bh][S][kp].
// NC head-dim chunks of 32 per round of loads (8 / NC dependent round trips); keys >= P come from this step's K rows
// (LL), keys < P from the prefix cache. Same MMA order as attn_scores_unit (attn_split.cuh).
template <int NC, bool KNEW, bool HI>
__device__ __noinline__ void ll_scores_unit(const AttnCtx a, const bf16* kc, int u, int lane) {
  const AttnCtx& p = a;
  const int g = lane >> 2, t4 = lane & 3;
  const int n_kb = p.kp >> 4;
  const int kb = u % n_kb, bh = u / n_kb, h = bh & 7, b = bh >> 3;
  const int n_keys = p.P + p.S;
  const int j0 = kb * 16;
  int ja = j0 + g, jb = j0 + 8 + g;
  ja = ja < n_keys ? ja : n_keys - 1;
  jb = jb < n_keys ? jb : n_keys - 1;
  const bool a_new = KNEW && ja >= p.P, b_new = KNEW && jb >= p.P;
  const bf16* kbase = kc + (int64_t)b * p.kv_pitch * 256;
  const bf16* ka_p = kbase + (int64_t)ja * 256 + t4 * 8;
  const bf16* kb_p = kbase + (int64_t)jb * 256 + t4 * 8;
  const int64_t ka_e = ((int64_t)(b * p.S + (a_new ? ja - p.P : 0))) * 256 + t4 * 8;
  const int64_t kb_e = ((int64_t)(b * p.S + (b_new ? jb - p.P : 0))) * 256 + t4 * 8;
  const bool lo_ok = g < p.S, hi_ok = g + 8 < p.S;
  constexpr bool any_hi = HI;
  const int64_t qlo_e = ((int64_t)(b * p.S + (lo_ok ? g : 0)) * 8 + h) * 256 + t4 * 8;
  const int64_t qhi_e = ((int64_t)(b * p.S + (hi_ok ? g + 8 : 0)) * 8 + h) * 256 + t4 * 8;
  float acc[2][4];
#pragma unroll
  for (int nt = 0; nt < 2; ++nt)
#pragma unroll
    for (int e = 0; e < 4; ++e) acc[nt][e] = 0.0f;
#pragma unroll 1
  for (int c0 = 0; c0 < 8; c0 += NC) {
    uint4 ka[NC], kbv[NC], lo[NC], hi[NC];
    // one gather for everything this round needs from LL buffers: q rows g (and g + 8), this step's K rows
    constexpr int kQW = (HI ? 4 : 2) * NC;             // q words
    constexpr int kW = kQW + (KNEW ? 4 * NC : 0);      // + K words (keys ja, jb)
    LLW w[kW];
    uint32_t need = (1u << kQW) - 1u;
    if constexpr (KNEW) {
      if (a_new) need |= ((1u << (2 * NC)) - 1u) << kQW;
      if (b_new) need |= ((1u << (2 * NC)) - 1u) << (kQW + 2 * NC);
    }
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      if (!a_new) ka[c] = ldcg_u4(ka_p + (c0 + c) * 32);
      if (!b_new) kbv[c] = ldcg_u4(kb_p + (c0 + c) * 32);
    }
    ll_gather<kW>(w, need, a.fq, [&](int j) -> const uint4* {
      if (j < 2 * NC) return a.Q + ((qlo_e + (c0 + (j >> 1)) * 32) >> 2) + (j & 1);
      if (HI && j < 4 * NC) return a.Q + ((qhi_e + (c0 + ((j - 2 * NC) >> 1)) * 32) >> 2) + (j & 1);
      const int k = j - kQW;
      if (k < 2 * NC) return a.KN + ((ka_e + (c0 + (k >> 1)) * 32) >> 2) + (k & 1);
      return a.KN + ((kb_e + (c0 + ((k - 2 * NC) >> 1)) * 32) >> 2) + (k & 1);
    });
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      lo[c] = lo_ok ? ll_pack8(w[2 * c], w[2 * c + 1]) : make_uint4(0, 0, 0, 0);
      hi[c] = make_uint4(0, 0, 0, 0);
      if constexpr (HI) {
        if (hi_ok) hi[c] = ll_pack8(w[2 * NC + 2 * c], w[2 * NC + 2 * c + 1]);
      }
      if constexpr (KNEW) {
        if (a_new) ka[c] = ll_pack8(w[kQW + 2 * c], w[kQW + 2 * c + 1]);
        if (b_new) kbv[c] = ll_pack8(w[kQW + 2 * NC + 2 * c], w[kQW + 2 * NC + 2 * c + 1]);
      }
    }
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      mma_bf16_16816(acc[0], lo[c].x, hi[c].x, lo[c].y, hi[c].y, ka[c].x, ka[c].y);
      mma_bf16_16816(acc[0], lo[c].z, hi[c].z, lo[c].w, hi[c].w, ka[c].z, ka[c].w);
      mma_bf16_16816(acc[1], lo[c].x, hi[c].x, lo[c].y, hi[c].y, kbv[c].x, kbv[c].y);
      mma_bf16_16816(acc[1], lo[c].z, hi[c].z, lo[c].w, hi[c].w, kbv[c].z, kbv[c].w);
    }
  }
  const int64_t srow = (int64_t)bh * p.S * p.kp;
#pragma unroll
  for (int nt = 0; nt < 2; ++nt) {
    const int col = j0 + nt * 8 + t4 * 2;
    // lanes with even t4 publish 4 consecutive keys (their pair and the next lane's)
    const __nv_bfloat162 plo = __floats2bfloat162_rn(bf16_round(acc[nt][0]) * 0.0625f, bf16_round(acc[nt][1]) * 0.0625f);
    const __nv_bfloat162 phi = __floats2bfloat162_rn(bf16_round(acc[nt][2]) * 0.0625f, bf16_round(acc[nt][3]) * 0.0625f);
    const uint32_t ulo = *reinterpret_cast<const uint32_t*>(&plo), uhi = *reinterpret_cast<const uint32_t*>(&phi);
    const uint32_t nlo = __shfl_down_sync(0xffffffffu, ulo, 1), nhi = __shfl_down_sync(0xffffffffu, uhi, 1);
    if ((t4 & 1) == 0) {
      if (lo_ok) ll_st(a.SC + ((srow + (int64_t)g * p.kp + col) >> 2), ulo, nlo, a.fsc);
      if (hi_ok) ll_st(a.SC + ((srow + (int64_t)(g + 8) * p.kp + col) >> 2), uhi, nhi, a.fsc);
    }
  }
}

// unit = (sample, head, quarter of the 256 output dims), executed by the 8 consumer warps of one CTA: warp w does the masked
// softmax of rows w and w + 8 into the shared probability tile, then (after a CTA barrier) P.V for 8 of the unit's 64 dims.
// Same arithmetic, lane-to-key assignment and MMA order as attn_pv_unit (attn_split.cuh); the V^T columns of this step's
// suffix rows are patched in from the LL buffer.
__device__ __noinline__ void ll_pv_unit(const AttnCtx a, const bf16* vt, uint8_t* ptile, int u, int warp, int lane) {
  const AttnCtx& p = a;
  const int g = lane >> 2, t4 = lane & 3;
  const int dq = u & 3, bh = u >> 2, h = bh & 7, b = bh >> 3;
  const int n_keys = p.P + p.S, kp = p.kp;
  const int ppb = attn_p_pitch_bytes(kp);
  const float kNeg = -2.3819763e38f;
  const int64_t srow = (int64_t)bh * p.S * kp;
  const int d0 = dq * 64 + warp * 8;
  const bf16* va = vt + ((int64_t)b * 256 + d0 + g) * p.vt_pitch;
  const int n_chunks = kp >> 5;
  uint4 fa[kMegaKeyChunks];
#pragma unroll
  for (int cc = 0; cc < kMegaKeyChunks; ++cc) {
    const int j = cc * 32 + t4 * 8;
    fa[cc] = make_uint4(0, 0, 0, 0);
    if (cc < n_chunks && j + 8 <= p.vt_pitch) fa[cc] = ldcg_u4(va + j);
  }
  // patch keys [P, P + S): element (key - P, d0 + g) of this step's V rows
#pragma unroll
  for (int cc = 0; cc < kMegaKeyChunks; ++cc) {
    const int j = cc * 32 + t4 * 8;
    if (cc < n_chunks && j + 8 > p.P && j < n_keys) {
      uint32_t wv[4] = {fa[cc].x, fa[cc].y, fa[cc].z, fa[cc].w};
      LLW r[8];
      uint32_t need = 0;
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (j + i >= p.P && j + i < n_keys) need |= 1u << i;
      const int64_t e0 = ((int64_t)(b * p.S + j - p.P)) * 256 + d0 + g;  // element of key j (may be "negative key": masked)
      ll_gather<8>(r, need, a.fq, [&](int i) { return a.VN + ((e0 + (int64_t)i * 256) >> 2); });
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if ((need >> i) & 1u) {
          const int64_t e = e0 + (int64_t)i * 256;
          const uint32_t d = (e & 2) ? (uint32_t)r[i].b : (uint32_t)r[i].a;
          const uint32_t bits = (e & 1) ? (d >> 16) : (d & 0xffffu);
          wv[i >> 1] = (i & 1) ? ((wv[i >> 1] & 0x0000ffffu) | (bits << 16)) : ((wv[i >> 1] & 0xffff0000u) | bits);
        }
      }
      fa[cc] = make_uint4(wv[0], wv[1], wv[2], wv[3]);
    }
  }
#pragma unroll
  for (int rr = 0; rr < 2; ++rr) {
    const int r = warp + rr * 8;
    bf16* prow = reinterpret_cast<bf16*>(ptile + r * ppb);
    if (r >= p.S) {
      for (int j = lane; j < kp; j += 32) prow[j] = to_bf16(0.0f);
      continue;
    }
    const int cq = p.cum[r];
    float v[kMegaKeyChunks];
    unsigned okbits = 0;
    constexpr int kHalf = kMegaKeyChunks / 2;
#pragma unroll
    for (int hb = 0; hb < 2; ++hb) {  // two batches of polls: 6 words in flight per lane
      LLW sw[kHalf];
#pragma unroll
      for (int ii = 0; ii < kHalf; ++ii) {
        const int i = hb * kHalf + ii;
        const int j = i * 32 + lane;
        const int jp = j < p.P ? j : p.P - 1;
        const bool okp = __ldg(p.pad + (int64_t)b * p.P + jp) != 0;
        const int js = min(max(j - p.P, 0), p.S - 1);
        const bool oks = __ldg(p.cum + js) <= cq;
        if (j < n_keys && (j < p.P ? okp : oks)) okbits |= 1u << i;
      }
      auto sc_elem = [&](int ii) -> int64_t {
        const int j = (hb * kHalf + ii) * 32 + lane;
        return srow + (int64_t)r * kp + (j < kp ? j : kp - 1);
      };
      ll_gather<kHalf>(sw, (1u << kHalf) - 1u, a.fsc, [&](int ii) { return a.SC + (sc_elem(ii) >> 2); });
#pragma unroll
      for (int ii = 0; ii < kHalf; ++ii) {
        const int64_t e = sc_elem(ii);
        const uint32_t d = (e & 2) ? (uint32_t)sw[ii].b : (uint32_t)sw[ii].a;
        v[hb * kHalf + ii] = __uint_as_float((e & 1) ? (d & 0xffff0000u) : (d << 16));
      }
    }
    float mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < kMegaKeyChunks; ++i) {
      const int j = i * 32 + lane;
      v[i] = j < n_keys ? (((okbits >> i) & 1u) ? v[i] : kNeg) : -INFINITY;
      mx = fmaxf(mx, v[i]);
    }
    mx = warp_max_f(mx);
    float sum = 0.0f;
#pragma unroll
    for (int i = 0; i < kMegaKeyChunks; ++i) {
      v[i] = (i * 32 + lane < n_keys) ? expf(v[i] - mx) : 0.0f;
      sum += v[i];
    }
    sum = warp_sum_f(sum);
    const float inv = 1.0f / sum;
#pragma unroll
    for (int i = 0; i < kMegaKeyChunks; ++i) {
      const int j = i * 32 + lane;
      if (j < kp) prow[j] = to_bf16(v[i] * inv);
    }
  }
  attn_sync256();
  float acc[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
  for (int cc = 0; cc < kMegaKeyChunks; ++cc) {
    if (cc < n_chunks) {
      const int j = cc * 32 + t4 * 8;
      const uint4 lo = *reinterpret_cast<const uint4*>(ptile + g * ppb + j * 2);
      const uint4 hi = *reinterpret_cast<const uint4*>(ptile + (g + 8) * ppb + j * 2);
      mma_bf16_16816(acc, lo.x, hi.x, lo.y, hi.y, fa[cc].x, fa[cc].y);
      mma_bf16_16816(acc, lo.z, hi.z, lo.w, hi.w, fa[cc].z, fa[cc].w);
    }
  }
  {
    const int d = d0 + t4 * 2;
    const __nv_bfloat162 plo = __floats2bfloat162_rn(acc[0], acc[1]), phi = __floats2bfloat162_rn(acc[2], acc[3]);
    const uint32_t ulo = *reinterpret_cast<const uint32_t*>(&plo), uhi = *reinterpret_cast<const uint32_t*>(&phi);
    const uint32_t nlo = __shfl_down_sync(0xffffffffu, ulo, 1), nhi = __shfl_down_sync(0xffffffffu, uhi, 1);
    if ((t4 & 1) == 0) {
      if (g < p.S) ll_st(a.ATT + ((((int64_t)(b * p.S + g) * 8 + h) * 256 + d) >> 2), ulo, nlo, a.fatt);
      if (g + 8 < p.S) ll_st(a.ATT + ((((int64_t)(b * p.S + g + 8) * 8 + h) * 256 + d) >> 2), uhi, nhi, a.fatt);
    }
  }
  attn_sync256();
}

template <bool M16>
__global__ void __launch_bounds__(kThreads, 1) denoise_loop_kernel(const LoopParams p) {
  using Cfg = LoopCfg<M16>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  uint8_t* ring = smem;
  bf16* As = reinterpret_cast<bf16*>(smem + Cfg::kRingBytes);
  float* red = reinterpret_cast<float*>(smem + Cfg::kRingBytes + Cfg::kABytes);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kRingBytes + Cfg::kABytes + Cfg::kRedBytes);
  uint64_t* empty_bar = full_bar + Cfg::kSlots;
  float* semb = reinterpret_cast<float*>(smem + Cfg::kRingBytes + Cfg::kABytes + Cfg::kRedBytes + 256);  // [8][8]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cta = blockIdx.x, G = gridDim.x;
  if (threadIdx.x == 0) {
    for (int s = 0; s < Cfg::kSlots; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 8);
    }
    fence_barrier_init();
  }
  __syncthreads();

  const int16_t* tab = p.units + (int64_t)cta * LOOP_PHASES * kTabW;
  const bool lo_mode = p.layers_only != 0;
  const int n_steps = lo_mode ? 1 : p.n_steps;

  // Register budget: 9 warps would cap every thread at 168 registers (3 warps per scheduler), which made ptxas serialise
  // the consumers' batched LL loads. With warpgroup granularity the producer group hands its registers over instead:
  // 2 x 128 consumer threads at 232 + 128 producer-group threads at 40 = the 64512 registers the CTA was launched with.
  if (warp >= 8) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
    if (warp > 8) return;
    // ------------------------------ producer: stream this CTA's weight chunks in consumption order ---------------
    // (weights are never written by a preceding kernel: no griddepcontrol.wait on this path)
    const uint64_t policy = policy_evict_first();
    int slot = 0;
    uint32_t phase = 0;
    auto stream = [&](int ph, const bf16* W, int K, int nch, bool along_k) {
      const int16_t* t = tab + ph * kTabW;
      const int cnt = t[0];
      for (int i = 0; i < cnt; ++i) {
        const int u = t[1 + i];
        for (int ch = 0; ch < nch; ++ch) {
          mbar_wait_quiet(&empty_bar[slot], phase ^ 1);
          if (lane == 0) mbar_arrive_expect_tx(&full_bar[slot], 8 * kChunkK * 2);
          __syncwarp();
          if (lane < 8) {
            const int row = unit_row(ph, u, along_k ? 0 : ch, lane);
            const bf16* src = W + (int64_t)row * K + (along_k ? ch * kChunkK : 0);
            bulk_g2s(ring + slot * kSlotBytes + lane * kRowPitch * 2, src, kChunkK * 2, &full_bar[slot], policy);
          }
          if (++slot == Cfg::kSlots) { slot = 0; phase ^= 1; }
          if (p.pace_ns > 0) __nanosleep((unsigned)p.pace_ns);
        }
      }
    };
    for (int step = 0; step < n_steps; ++step) {
      if (!lo_mode) {
        stream(LOOP_TIN, p.time_in_wa, kHidden, 1, true);
        stream(LOOP_TOUT, p.time_out_w, kHidden, 1, true);
      }
#pragma unroll 1
      for (int layer = 0; layer < p.n_layers; ++layer) {
        const MegaLayerW& lw = p.L[layer];
        stream(LOOP_QKV, lw.wqkv, kHidden, 2, false);
        stream(LOOP_O, lw.wo, kQDim, 2, true);
        stream(LOOP_GU, lw.wgu, kHidden, 2, false);
        stream(LOOP_DOWN, lw.wdown, kInter, 4, true);
      }
      if (!lo_mode) stream(LOOP_OUT, p.action_out_w, kHidden, 1, true);
    }
    return;
  }

  // ------------------------------ consumers ------------------------------------------------------------------------
  asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
  pdl_wait();
  const int tid = threadIdx.x;
  const int g = lane >> 2, t4 = lane & 3;
  int slot = 0;
  uint32_t phase = 0;
  int parity = 0;  // double-buffered cross-warp reduction scratch
  uint32_t epoch_base;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(epoch_base) : "l"(p.epoch) : "memory");
  uint32_t ep = 0;  // productions so far (identical sequence in every CTA); flag of a production = epoch_base + ep
  // Arrival counter = a HINT that a production is (probably) complete, so that only ONE thread per CTA spins while the
  // data is in flight (256 threads x 148 CTAs polling their LL words saturated the L2 request path and starved the weight
  // stream). Producers bump it with a relaxed, unfenced red after issuing their LL stores; the LL epochs remain the
  // only correctness mechanism, so an early / late / miscounted hint costs a few polls, never a wrong read. Its base
  // value is the running total the previous launch stored next to the epoch at exit.
  unsigned* const ctr = p.epoch + 8;
  uint32_t ctr_t;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(ctr_t) : "l"(p.epoch + 1) : "memory");
  auto arrive = [&]() { asm volatile("red.relaxed.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory"); };
  auto hint_wait = [&](uint32_t target) {
    if (threadIdx.x == 0) {
      for (int i = 0; i < 8192; ++i) {
        uint32_t v;
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
        if ((int32_t)(v - target) >= 0) break;
      }
    }
    consumer_sync();
  };

  auto llbuf = [&](int buf, int set) -> uint4* { return p.ll + (int64_t)(set & 1) * kLLSetWords + ll_offset(buf); };

  // runs this CTA's units of one GEMM phase; `epi(u, row, col, valid, t0, t1)` is called by ALL threads < 128 with
  // row = tid >> 3, col = tid & 7 (valid = row < rows)
  auto run_units = [&](int ph, int nch, bool along_k, int rows, auto&& epi) {
    const int16_t* t = tab + ph * kTabW;
    const int cnt = t[0];
    for (int i = 0; i < cnt; ++i) {
      const int u = t[1 + i];
      float acc[2][4];
#pragma unroll
      for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[a][e] = 0.0f;
      for (int ch = 0; ch < nch; ++ch) {
        mbar_wait_quiet(&full_bar[slot], phase);
        const bf16* sl = reinterpret_cast<const bf16*>(ring + slot * kSlotBytes);
        if (along_k) mma_chunk<M16>(acc[0], As, ch, sl, warp, g, t4);
        else if (ch == 0) mma_chunk<M16>(acc[0], As, 0, sl, warp, g, t4);
        else mma_chunk<M16>(acc[1], As, 0, sl, warp, g, t4);
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[slot]);
        if (++slot == Cfg::kSlots) { slot = 0; phase ^= 1; }
      }
      float* rp = red + parity * (8 * 2 * 16 * 8);
#pragma unroll
      for (int a = 0; a < 2; ++a) {
        *reinterpret_cast<float2*>(rp + ((warp * 2 + a) * 16 + g) * 8 + t4 * 2) = make_float2(acc[a][0], acc[a][1]);
        *reinterpret_cast<float2*>(rp + ((warp * 2 + a) * 16 + g + 8) * 8 + t4 * 2) = make_float2(acc[a][2], acc[a][3]);
      }
      consumer_sync();
      if (tid < 128) {
        const int row = tid >> 3, col = tid & 7;
        float t0 = 0.0f, t1 = 0.0f;
#pragma unroll
        for (int w = 0; w < 8; ++w) {
          t0 += rp[((w * 2 + 0) * 16 + row) * 8 + col];
          t1 += rp[((w * 2 + 1) * 16 + row) * 8 + col];
        }
        epi(u, row, col, row < rows, t0, t1);
      }
      parity ^= 1;
    }
  };
  auto owns = [&](int ph) -> int { return tab[ph * kTabW] > 0 ? tab[ph * kTabW + 1] : -1; };

  const int M = p.M, S = p.S, H = p.H, B = p.B, Ma = p.B * p.H;
  // producers (arrivals) per production kind
  const uint32_t n128 = G < 128 ? G : 128, nQ = G < 160 ? G : 160, nSC = (uint32_t)(B * 8 * (p.kp >> 4)),
                 nATT = (uint32_t)(B * 32 < G ? B * 32 : G), nACT = G < 512 ? G : 512;
  uint32_t tX = 0;        // hint target of the latest X
  int nx = 0;             // X productions so far: X number n lives in set n & 1
  uint32_t fX = 0;        // flag of the latest X
  float xt_reg = 0.0f;    // OUT owners: x_t[(row -> b, t - 1)][u * 8 + col], carried across the steps
  const int u_tin = owns(LOOP_TIN), u_tout = owns(LOOP_TOUT), u_out = owns(LOOP_OUT);

  if (lo_mode) {
    // publish the layer-0 input (written by the preceding kernels) as X number 0
    ++ep;
    fX = epoch_base + ep;
    ctr_t += 1;
    tX = ctr_t;
    if (cta == 0) {
      uint4* X = llbuf(LL_X, nx);
      for (int w = tid; w < M * 256; w += 256) {
        const uint2 d = __ldcg(reinterpret_cast<const uint2*>(p.x0) + w);
        ll_st(X + w, d.x, d.y, fX);
      }
      if (tid == 0) arrive();
    }
  } else {
    // state_proj for the columns of this CTA's time_mlp_out unit (input-independent across the steps)
    if (u_tout >= 0 && tid < 128) {
      const int b = tid >> 3, j = tid & 7;
      if (b < B && b < 8) {
        const int n = u_tout * 8 + j;
        float acc = 0.0f;
        const bf16* wr = p.state_w + (int64_t)n * 32;
#pragma unroll 8
        for (int k = 0; k < 32; ++k) acc += bf16_round(__ldg(p.state + b * 32 + k)) * __bfloat162float(wr[k]);
        semb[b * 8 + j] = bf16_round(acc + __ldg(p.state_b + n));
      }
    }
    if (u_out >= 0 && tid < 128) {
      const int row = tid >> 3, col = tid & 7;
      const int b = row / S, t = row % S;
      if (row < M && t >= 1) xt_reg = p.x_t[((int64_t)b * H + t - 1) * 32 + u_out * 8 + col];
    }
    consumer_sync();
  }

  for (int step = 0; step < n_steps; ++step) {
    if (!lo_mode) {
      // ---- action_in_proj: a1 = rnd(rnd_bf16(x_t) . W^T + b), K = 32 (weights straight from L2, no ring) ------------
      float* xs = reinterpret_cast<float*>(As);
      if (step == 0) {
        for (int i = tid; i < Ma * 32; i += 256) xs[i] = p.x_t[i];
      } else {
        const uint4* XT = llbuf(LL_XT, step - 1);
        const uint32_t fxt = epoch_base + ep;  // the OUT phase was the previous production
        hint_wait(ctr_t);
        for (int w = tid; w < Ma * 16; w += 256) {
          const LLW v = ll_wait(XT + w, fxt);
          xs[2 * w] = __uint_as_float((uint32_t)v.a);
          xs[2 * w + 1] = __uint_as_float((uint32_t)v.b);
        }
      }
      consumer_sync();
      ++ep;
      const uint32_t fA1 = epoch_base + ep;
      ctr_t += n128;
      const uint32_t tA1 = ctr_t;
      if (u_tin >= 0 && tid < 128) {
        const int r = tid >> 3, j = tid & 7;
        const bool valid = r < Ma;
        const int rr = valid ? r : 0;
        const int n = u_tin * 8 + j;
        const uint4* wr = reinterpret_cast<const uint4*>(p.action_in_w + (int64_t)n * 32);
        float acc = 0.0f;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint4 wq = __ldg(wr + q);
          const __nv_bfloat162* wh = reinterpret_cast<const __nv_bfloat162*>(&wq);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 wf = __bfloat1622float2(wh[e]);
            acc += bf16_round(xs[rr * 32 + q * 8 + 2 * e]) * wf.x;
            acc += bf16_round(xs[rr * 32 + q * 8 + 2 * e + 1]) * wf.y;
          }
        }
        const float v = bf16_round(acc + __ldg(p.action_in_b + n));
        ll_publish4(llbuf(LL_A1, step), (int64_t)r * kHidden + n, j, bf16_bits(v), valid, fA1);
        if (tid == 0) arrive();
      }
      consumer_sync();

      // ---- action_time_mlp_in + SiLU (time half folded into the per-step bias) ---------------------------------------
      ++ep;
      const uint32_t fA2 = epoch_base + ep;
      ctr_t += n128;
      const uint32_t tA2 = ctr_t;
      if (u_tin >= 0) {
        hint_wait(tA1);
        stage_copy_ll(As, llbuf(LL_A1, step), fA1, Ma, kHidden);
        consumer_sync();
      }
      {
        uint4* A2 = llbuf(LL_A2, step);
        const float* tb = p.time_bias + (int64_t)step * kHidden;
        run_units(LOOP_TIN, 1, true, Ma, [&](int u, int row, int col, bool valid, float t0, float) {
          const int n = u * 8 + col;
          float v = bf16_round(t0 + __ldg(tb + n));
          v = bf16_round(silu(v));
          ll_publish4(A2, (int64_t)row * kHidden + n, col, bf16_bits(v), valid, fA2);
        });
        if (u_tin >= 0 && tid == 0) arrive();
      }
      // ---- action_time_mlp_out -> suffix rows 1..H; state embedding -> suffix row 0 ------------------------------------
      ++ep;
      ++nx;
      fX = epoch_base + ep;
      ctr_t += n128;
      tX = ctr_t;
      if (u_tout >= 0) {
        hint_wait(tA2);
        stage_copy_ll(As, llbuf(LL_A2, step), fA2, Ma, kHidden);
        consumer_sync();
      }
      {
        uint4* X = llbuf(LL_X, nx);
        run_units(LOOP_TOUT, 1, true, Ma, [&](int u, int row, int col, bool valid, float t0, float) {
          const int n = u * 8 + col;
          const float v = bf16_round(t0 + __ldg(p.time_out_b + n));
          const int orow = valid ? (row / H) * S + 1 + row % H : 0;
          ll_publish4(X, (int64_t)orow * kHidden + n, col, bf16_bits(v), valid, fX);
        });
        if (u_tout >= 0 && tid < 128) {
          const int b = tid >> 3, j = tid & 7;
          const bool valid = b < B && b < 8;
          const float v = semb[(valid ? b : 0) * 8 + j];
          ll_publish4(X, (int64_t)(valid ? b * S : 0) * kHidden + u_tout * 8 + j, j, bf16_bits(v), valid, fX);
          if (tid == 0) arrive();
        }
      }
    }

    for (int layer = 0; layer < p.n_layers; ++layer) {
      const int gl = step * p.n_layers + layer;
      const MegaLayerW& lw = p.L[layer];
      bf16* kc = p.kcache + (int64_t)layer * p.kc_layer_stride;
      bf16* vt = p.vtcache + (int64_t)layer * p.vt_layer_stride;
      const uint4* Xin = llbuf(LL_X, nx);
      const uint32_t fXin = fX;
      AttnCtx ac;
      ac.Q = llbuf(LL_Q, gl); ac.KN = llbuf(LL_KN, gl); ac.VN = llbuf(LL_VN, gl);
      ac.SC = llbuf(LL_SC, gl); ac.ATT = llbuf(LL_ATT, gl);
      ac.S = S; ac.P = p.P; ac.kp = p.kp; ac.kv_pitch = p.kv_pitch; ac.vt_pitch = p.vt_pitch; ac.pad = p.pad; ac.cum = p.cum;

      // ---- phase 1: q|k|v = RMSNorm(x) . Wqkv^T, RoPE, caches ------------------------------------------------------
      LOOP_STAMP(0);
      hint_wait(tX);
      LOOP_STAMP(11);
      stage_norm_ll<M16>(As, Xin, fXin, lw.g1, M);
      consumer_sync();
      LOOP_STAMP(1);
      ++ep;
      ac.fq = epoch_base + ep;
      ctr_t += nQ;
      const uint32_t tQ = ctr_t;
      {
        uint4* Qw = llbuf(LL_Q, gl);
        uint4* KNw = llbuf(LL_KN, gl);
        uint4* VNw = llbuf(LL_VN, gl);
        const uint32_t fq = ac.fq;
        run_units(LOOP_QKV, 2, false, M, [&](int u, int row, int col, bool valid, float t0, float t1) {
          const int head = u >> 4, d = (u & 15) * 8 + col;
          const float a = bf16_round(t0), b = bf16_round(t1);
          float o1 = a, o2 = b;
          if (head < 9) {
            int pos = valid ? p.pos_ids[row] : 0;
            if (pos < 0) pos = 0;
            const float cs = __ldg(p.rope_cos + (int64_t)pos * 128 + d), sn = __ldg(p.rope_sin + (int64_t)pos * 128 + d);
            o1 = a * cs - b * sn;
            o2 = b * cs + a * sn;
          }
          const uint32_t b1 = bf16_bits(o1), b2 = bf16_bits(o2);
          if (head < 8) {
            const int64_t e = (int64_t)row * kQDim + head * 256 + d;
            ll_publish4(Qw, e, col, b1, valid, fq);
            ll_publish4(Qw, e + 128, col, b2, valid, fq);
          } else {
            uint4* dstb = head == 8 ? KNw : VNw;
            const int64_t e = (int64_t)row * 256 + d;
            ll_publish4(dstb, e, col, b1, valid, fq);
            ll_publish4(dstb, e + 128, col, b2, valid, fq);
            if (valid) {  // the caches also receive the suffix rows (the reference's cat[cache, k/v])
              const int bb = row / S, t = row % S;
              if (head == 8) {
                bf16* dst = kc + ((int64_t)bb * p.kv_pitch + p.P + t) * 256 + d;
                dst[0] = __ushort_as_bfloat16((unsigned short)b1);
                dst[128] = __ushort_as_bfloat16((unsigned short)b2);
              } else {
                bf16* dst = vt + (int64_t)bb * 256 * p.vt_pitch + p.P + t;
                dst[(int64_t)d * p.vt_pitch] = __ushort_as_bfloat16((unsigned short)b1);
                dst[(int64_t)(d + 128) * p.vt_pitch] = __ushort_as_bfloat16((unsigned short)b2);
              }
            }
          }
        });
      }

      if (owns(LOOP_QKV) >= 0 && tid == 0) arrive();
      // ---- phase 2a: scores, one warp per (sample, head, 16 keys) ---------------------------------------------------
      LOOP_STAMP(2);
      ++ep;
      ac.fsc = epoch_base + ep;
      ctr_t += nSC;
      const uint32_t tSC = ctr_t;
      {
        const int n_kb = p.kp >> 4;
        const int n_units = B * 8 * n_kb;
        if (cta < n_units) hint_wait(tQ);
        LOOP_STAMP(12);
        bool did = false;
        for (int u = warp * G + cta; u < n_units; u += 8 * G) {
          did = true;
          const bool knew = ((u % n_kb) + 1) * 16 > p.P;  // block reaches into this step's suffix keys
          if constexpr (M16) {
            if (S > 8) {
              if (knew) ll_scores_unit<2, true, true>(ac, kc, u, lane);
              else ll_scores_unit<2, false, true>(ac, kc, u, lane);
              continue;
            }
          }
          if (knew) ll_scores_unit<2, true, false>(ac, kc, u, lane);
          else ll_scores_unit<4, false, false>(ac, kc, u, lane);
        }
        if (did && lane == 0) arrive();
      }
      // ---- phase 2b: masked softmax + P.V, one CTA per (sample, head, 64 dims) ----------------------------------------
      LOOP_STAMP(3);
      ++ep;
      ac.fatt = epoch_base + ep;
      ctr_t += nATT;
      const uint32_t tATT = ctr_t;
      if (cta < B * 32) {
        hint_wait(tSC);
        LOOP_STAMP(13);
        for (int u = cta; u < B * 32; u += G) ll_pv_unit(ac, vt, reinterpret_cast<uint8_t*>(As), u, warp, lane);
        if (tid == 0) arrive();
      }

      // ---- phase 3: y = x + att . Wo^T -------------------------------------------------------------------------------
      LOOP_STAMP(4);
      ++ep;
      const uint32_t fY = epoch_base + ep;
      ctr_t += n128;
      const uint32_t tY = ctr_t;
      uint4* Yb = llbuf(LL_Y, gl);
      if (owns(LOOP_O) >= 0) {
        hint_wait(tATT);
        LOOP_STAMP(14);
        stage_copy_ll(As, ac.ATT, ac.fatt, M, kQDim);
        consumer_sync();
      }
      LOOP_STAMP(5);
      run_units(LOOP_O, 2, true, M, [&](int u, int row, int col, bool valid, float t0, float) {
        const int n = u * 8 + col;
        const int64_t e = (int64_t)(valid ? row : 0) * kHidden + n;
        const float v = bf16_round(t0) + ll_elem_bf16(Xin, e, fXin);
        ll_publish4(Yb, e, col, bf16_bits(v), valid, fY);
      });

      if (owns(LOOP_O) >= 0 && tid == 0) arrive();
      // ---- phase 4: act = gelu_tanh(h . Wg^T) * (h . Wu^T), h = RMSNorm(y) -------------------------------------------
      LOOP_STAMP(6);
      hint_wait(tY);
      LOOP_STAMP(15);
      stage_norm_ll<M16>(As, Yb, fY, lw.g2, M);
      consumer_sync();
      LOOP_STAMP(7);
      ++ep;
      const uint32_t fAct = epoch_base + ep;
      ctr_t += nACT;
      const uint32_t tACT = ctr_t;
      uint4* ACTb = llbuf(LL_ACT, gl);
      run_units(LOOP_GU, 2, false, M, [&](int u, int row, int col, bool valid, float t0, float t1) {
        const float gt = bf16_round(t0), up = bf16_round(t1);
        const float act = bf16_round(gelu_tanh_fast(gt));
        ll_publish4(ACTb, (int64_t)row * kInter + u * 8 + col, col, bf16_bits(act * up), valid, fAct);
      });

      if (owns(LOOP_GU) >= 0 && tid == 0) arrive();
      // ---- phase 5: x = y + act . Wdown^T ----------------------------------------------------------------------------
      LOOP_STAMP(8);
      ++ep;
      ++nx;
      fX = epoch_base + ep;
      ctr_t += n128;
      tX = ctr_t;
      uint4* Xout = llbuf(LL_X, nx);
      if (owns(LOOP_DOWN) >= 0) {
        hint_wait(tACT);
        LOOP_STAMP(16);
        stage_copy_ll(As, ACTb, fAct, M, kInter);
        consumer_sync();
      }
      LOOP_STAMP(9);
      {
        const uint32_t fXo = fX;
        run_units(LOOP_DOWN, 4, true, M, [&](int u, int row, int col, bool valid, float t0, float) {
          const int n = u * 8 + col;
          const int64_t e = (int64_t)(valid ? row : 0) * kHidden + n;
          const float v = bf16_round(t0) + ll_elem_bf16(Yb, e, fY);
          const uint32_t bits = bf16_bits(v);
          ll_publish4(Xout, e, col, bits, valid, fXo);
          if (valid) p.x[e] = __ushort_as_bfloat16((unsigned short)bits);
        });
      }
      if (owns(LOOP_DOWN) >= 0 && tid == 0) arrive();
      LOOP_STAMP(10);
    }

    if (!lo_mode) {
      // ---- final RMSNorm, action_out_proj and the Euler update x_t += rnd(dt * rnd(v_t)) (modeling_mvla.py:802) ------------
      ++ep;
      const uint32_t fXT = epoch_base + ep;
      ctr_t += 4;
      if (u_out >= 0) {
        hint_wait(tX);
        stage_norm_ll<M16>(As, llbuf(LL_X, nx), fX, p.exp_norm_g, M);
        consumer_sync();
      }
      uint4* XT = llbuf(LL_XT, step);
      run_units(LOOP_OUT, 1, true, M, [&](int u, int row, int col, bool valid, float t0, float) {
        const int n = u * 8 + col;
        const int b = row / S, t = row % S;
        const bool act_row = valid && t >= 1;
        const float vt_ = bf16_round(t0 + __ldg(p.action_out_b + n));
        if (act_row) xt_reg += bf16_round(p.dt * vt_);
        const int64_t e = act_row ? ((int64_t)b * H + t - 1) * 32 + n : 0;
        const uint32_t mine = __float_as_uint(xt_reg);
        const uint32_t next = __shfl_down_sync(0xffffffffu, mine, 1);
        if (act_row) {
          p.x_t[e] = xt_reg;
          if ((col & 1) == 0) ll_st(XT + (e >> 1), mine, next, fXT);
        }
      });
      if (u_out >= 0 && tid == 0) arrive();
    }
  }
  // advance the arena's epoch for the next launch: CTA 0 has consumed data that depends on every CTA's epoch read
  if (cta == 0 && tid == 0) {
    asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p.epoch), "r"(epoch_base + ep) : "memory");
    asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p.epoch + 1), "r"(ctr_t) : "memory");
  }
}

}  // namespace

// Layout: [cta][phase][1 + kMegaMaxUnits] = count, unit ids. The four per-layer phases are dealt as in
// mega_build_units (balanced per phase); time_mlp_in / time_mlp_out unit u (8 of the 1024 columns) goes to CTA u, the
// four action_out_proj units to the last CTAs (idle in o_proj / down).
int loop_build_units(int G, std::vector<int16_t>& tab) {
  PI0_REQUIRE(G >= 128 && G <= 1024, "denoise loop kernel: %d CTAs outside [128, 1024]", G);
  tab.assign((size_t)G * LOOP_PHASES * kTabW, 0);
  auto push = [&](int cta, int ph, int u) -> bool {
    int16_t* t = tab.data() + ((size_t)cta * LOOP_PHASES + ph) * kTabW;
    if (t[0] >= kMegaMaxUnits) return false;
    t[1 + t[0]] = (int16_t)u;
    ++t[0];
    return true;
  };
  bool ok = true;
  for (int u = 0; u < kHidden / 8; ++u)
    ok = ok && push(u % G, LOOP_O, u) && push(u % G, LOOP_DOWN, u) && push(u, LOOP_TIN, u) && push(u, LOOP_TOUT, u);
  for (int u = 0; u < kInter / 8; ++u) ok = ok && push(G - 1 - (u % G), LOOP_GU, u);
  for (int u = 0; u < 10 * 16; ++u) ok = ok && push(G - 1 - (u % G), LOOP_QKV, u);
  for (int u = 0; u < 4; ++u) ok = ok && push(G - 1 - u, LOOP_OUT, u);
  PI0_REQUIRE(ok, "denoise loop kernel: unit table overflow");
  return 0;
}

size_t loop_ll_bytes() { return (size_t)2 * kLLSetWords * 16; }
size_t loop_units_bytes() { return (size_t)1024 * LOOP_PHASES * kTabW * sizeof(int16_t); }

bool loop_supported(int M, int S, int P, int num_ctas) {
  return M >= 1 && M <= 16 && S >= 2 && S <= 16 && P + S <= 32 * kMegaKeyChunks && P % 8 == 0 && num_ctas >= 128 &&
         num_ctas <= 512 && M / S <= 8;
}

int loop_denoise(const LoopParams& p, int num_ctas, cudaStream_t s) {
  PI0_REQUIRE(loop_supported(p.M, p.S, p.P, num_ctas), "denoise loop kernel: M=%d S=%d P=%d CTAs=%d unsupported", p.M,
              p.S, p.P, num_ctas);
  PI0_REQUIRE(p.n_layers >= 1 && p.n_layers <= 18 && p.n_steps >= 1, "denoise loop kernel: n_layers=%d n_steps=%d",
              p.n_layers, p.n_steps);
  PI0_REQUIRE(p.H == p.S - 1 && p.M == p.B * p.S, "denoise loop kernel: inconsistent shapes");
  const bool m16 = p.M > 8;
  const size_t smem = m16 ? LoopCfg<true>::kSmem : LoopCfg<false>::kSmem;
  PI0_REQUIRE(16 * attn_p_pitch_bytes(p.kp) <= (m16 ? LoopCfg<true>::kABytes : LoopCfg<false>::kABytes) &&
                  p.kp <= 32 * kMegaKeyChunks && p.kp % 32 == 0 && p.ll != nullptr && p.epoch != nullptr,
              "denoise loop kernel: attention scratch does not fit");
  static PerDeviceOnce attr_once;
  if (attr_once.need()) {
    PI0_CHECK_CUDA(cudaFuncSetAttribute(denoise_loop_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)LoopCfg<true>::kSmem));
    PI0_CHECK_CUDA(cudaFuncSetAttribute(denoise_loop_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)LoopCfg<false>::kSmem));
    attr_once.mark();
  }
  cudaLaunchConfig_t lc = {};
  lc.gridDim = dim3((unsigned)num_ctas);
  lc.blockDim = dim3(kThreads);
  lc.dynamicSmemBytes = smem;
  lc.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeCooperative;  // all CTAs co-resident: the polling exchanges cannot deadlock
  at[0].val.cooperative = 1;
  lc.attrs = at;
  lc.numAttrs = 1;
  if (m16) PI0_CHECK_CUDA(cudaLaunchKernelEx(&lc, denoise_loop_kernel<true>, p));
  else PI0_CHECK_CUDA(cudaLaunchKernelEx(&lc, denoise_loop_kernel<false>, p));
  ++g_launch_count;
  return 0;
}

}  // namespace pi0

#ifdef PI0_TIMELINE
extern "C" int pi0_debug_loop_timeline(unsigned long long* out) {
  return cudaMemcpyFromSymbol(out, pi0::g_loop_tl, sizeof(pi0::g_loop_tl)) == cudaSuccess ? 0 : -1;
}
extern "C" int pi0_debug_loop_stats(unsigned long long* out, int reset) {
  if (cudaMemcpyFromSymbol(out, pi0::g_loop_stats, sizeof(pi0::g_loop_stats)) != cudaSuccess) return -1;
  if (reset) {
    unsigned long long z[4] = {0, 0, 0, 0};
    if (cudaMemcpyToSymbol(pi0::g_loop_stats, z, sizeof(z)) != cudaSuccess) return -1;
  }
  return 0;
}
#endif
