// Persistent kernel for the latency regime of the denoise loop: ALL 18 action-expert layers of one Euler step in ONE
// cooperative launch (M = B * S <= 16 suffix rows, i.e. batch 1-3 at the bridge shapes).
//
// Why: at batch 1 a denoise step is 623 MB of expert weights against ~10 KB of activations. As ~90 separate kernels
// (5 per layer) every launch pays drain + launch + ramp-up and the HBM stream stops at every kernel boundary; measured
// 40 us per layer against 5.3 us of pure weight streaming. Here one CTA per SM stays resident for the whole step:
//   * warp 8 of every CTA is a PRODUCER that streams the CTA's share of the weights through a shared-memory ring with
//     1-D bulk TMA copies (cp.async.bulk, 8 weight rows x 1024 k per 16 KB chunk, L2 evict-first). It runs ahead of the
//     consumers by the depth of the ring and is never blocked by a grid barrier, so HBM keeps streaming while the
//     consumers synchronise;
//   * warps 0-7 are CONSUMERS: per phase they stage the (tiny) activation operand in shared memory (GemmaRMSNorm fused
//     into the staging), run mma.sync m16n8k16 over the chunks as they land (the 8 warps split K; the k index is
//     permuted identically on both operands so a lane's 8 consecutive k feed two MMAs), reduce across warps in a
//     fixed order and apply the epilogue (RoPE + in-place K / V^T cache write, residual add, GeGLU);
//   * phases of a layer: qkv+rope | attention | o_proj+res | gate/up GeGLU | down+res, separated by a grid barrier
//     (sense-reversing counter in global memory, self-resetting => CUDA-graph replay needs no memset).
// Work units (8 output columns over the whole K) are dealt to the CTAs by a host-built table that balances the bytes
// every CTA streams per layer.
// Reference arithmetic: [mirror] src/model/mvla/expert.py:128-179,387-409; mllm.py:30-53,265-317 (rounding points of
// SURVEY.md 9.7, identical to gemm_skinny.cu / attn_decode.cu which this kernel replaces for M <= 16).
#include <algorithm>
#include <vector>

#include "attn_split.cuh"
#include "denoise_mega.cuh"

namespace pi0 {

namespace {

constexpr int kThreads = 288;                   // 8 consumer warps + 1 producer warp
constexpr int kChunkK = 1024;                   // k elements per streamed chunk
constexpr int kRowPitch = kChunkK + 32;         // bf16 elements; 2112 B = 64 mod 128 (conflict-free 16-byte loads)
constexpr int kSlotBytes = 8 * kRowPitch * 2;   // 16896
constexpr int kAK = 4096;                       // largest K (down projection)
constexpr int kAPitch = kAK + 32;               // 8256 B = 64 mod 128
constexpr int kHidden = 1024, kInter = 4096, kQDim = 2048;

// Optional in-kernel timeline (development aid, -DPI0_TIMELINE): one CTA stamps %globaltimer at the phase boundaries
// of layer 1; pi0_debug_mega_timeline() copies the stamps back.
#ifdef PI0_TIMELINE
__device__ unsigned long long g_mega_tl[4][32];
#define MEGA_STAMP(i) do { if (layer == 1 && threadIdx.x == 0 && (cta == 0 || cta == 7 || cta == 64 || cta == 140)) \
  g_mega_tl[cta == 0 ? 0 : cta == 7 ? 1 : cta == 64 ? 2 : 3][i] = global_timer_ns(); } while (0)
#else
#define MEGA_STAMP(i) do { } while (0)
#endif

// R = rows of the staged activation operand: 8 (M <= 8: half an MMA tile), 16, or 32 / 48 (two / three m16 tiles, batch 4-9
// at the bridge shapes). Up to 16 rows the operand is staged for the whole K (up to 4096) at once; above that it would not
// fit next to the weight ring, so it is staged one 1024-k chunk at a time and the along-K phases (o_proj, down) re-stage
// between the chunks of their unit.
template <int R>
struct MegaCfg {
  static constexpr int kRows = R;
  static constexpr bool kChunkedA = R > 16;
  static constexpr int kPitch = kChunkedA ? kChunkK + 32 : kAPitch;  // bf16 elements per staged row (64 mod 128 bytes)
  static constexpr int kMT = R <= 16 ? 1 : R / 16;                   // m16 tiles
  static constexpr int kSlots = R == 8 ? 8 : R == 16 ? 4 : R == 32 ? 8 : 6;
  static constexpr int kRingBytes = kSlots * kSlotBytes;
  static constexpr int kABytes = kRows * kPitch * 2;
  static constexpr int kRedBytes = R <= 16 ? 2 * 8 * 2 * 16 * 8 * 4 : 8 * 2 * R * 8 * 4;  // [unit][warp][acc][row][col] fp32
  static constexpr int kSmem = kRingBytes + kABytes + kRedBytes + 256 + 128;
  static_assert(kSmem <= 232448, "shared memory budget of one SM");
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
// The RMSNorm gains of a phase (4 KB, first touched right after a grid barrier, i.e. a cold HBM read on the critical
// path) are pulled into L2 by 32 threads BEFORE the barrier; costs no registers.
__device__ __forceinline__ void prefetch_gain_l2(const float* gain) {
  if (threadIdx.x < 32) asm volatile("prefetch.global.L2 [%0];" ::"l"(gain + threadIdx.x * 32));
}

__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_add_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Grid barrier over the consumer warps of all CTAs: one monotonic 64-bit arrival counter, never reset (so a CUDA-graph
// replay needs no memset and the last arriver has no reset-then-release chain): barrier number k of a launch is passed
// once count >= base + k * G, where base = the counter at launch start rounded down to a multiple of G (a CTA that
// starts late sees at most G - 1 arrivals of barrier 1 on top of base). One fire-and-forget red.release per CTA and
// one ld.acquire poll loop; the CTA-wide bar.sync on both sides extends the release / acquire to the other threads.
// Bounded spin: a bug must surface as a launch failure, never as a hung GPU.
struct GridBarrier {
  unsigned long long* count;
  unsigned long long target;  // threshold of the NEXT barrier (thread 0 only)
  unsigned long long step;    // G
};
__device__ __forceinline__ void grid_barrier_init(GridBarrier& gb, unsigned long long* count, unsigned n_ctas) {
  gb.count = count;
  gb.step = n_ctas;
  gb.target = 0;
  if (threadIdx.x == 0) gb.target = (ld_acquire_u64(count) / n_ctas) * n_ctas + n_ctas;
}
__device__ __forceinline__ void grid_barrier(GridBarrier& gb) {
  consumer_sync();
  if (threadIdx.x == 0) {
    red_release_add_u64(gb.count, 1ull);
    if (ld_acquire_u64(gb.count) < gb.target) {
      const uint64_t t0 = global_timer_ns();
      uint32_t spins = 0;
      while (ld_acquire_u64(gb.count) < gb.target) {
        if (((++spins) & 0x3ff) == 0 && global_timer_ns() - t0 > 2000000000ull) {
          printf("pi0: denoise grid barrier timed out (block %d)\n", blockIdx.x);
          __trap();
        }
      }
    }
    gb.target += gb.step;
  }
  consumer_sync();
}

// weight row `j` (0..7) of chunk `ch` of unit `u` in phase `ph`, and the k offset of the chunk
__device__ __forceinline__ int unit_row(int ph, int u, int ch, int j) {
  if (ph == MEGA_QKV) return (u >> 4) * 256 + (u & 15) * 8 + ch * 128 + j;  // u = head*16 + dc; ch = RoPE half
  if (ph == MEGA_GU) {
    const int i = u * 8 + j;  // inter column; packed per 256 rows as 128 gate rows then the 128 matching up rows
    return (i >> 7) * 256 + (i & 127) + ch * 128;
  }
  return u * 8 + j;  // o_proj / down: 8 consecutive output columns, chunks run along K
}

struct RingState {
  int slot;
  uint32_t phase;
};

template <int R>
__device__ __forceinline__ void mma_chunk(float* acc, const bf16* As, int kc, const bf16* slot, int warp, int g,
                                          int t4) {
  using Cfg = MegaCfg<R>;
  const int k0 = warp * 128 + t4 * 8 + (Cfg::kChunkedA ? 0 : kc * kChunkK);
  const int w0 = warp * 128 + t4 * 8;
#pragma unroll
  for (int cc = 0; cc < 4; ++cc) {
    const uint4 w = *reinterpret_cast<const uint4*>(slot + g * kRowPitch + w0 + cc * 32);
#pragma unroll
    for (int mt = 0; mt < Cfg::kMT; ++mt) {
      const uint4 lo = *reinterpret_cast<const uint4*>(As + (g + 16 * mt) * Cfg::kPitch + k0 + cc * 32);
      uint4 hi = make_uint4(0, 0, 0, 0);
      if constexpr (R >= 16) hi = *reinterpret_cast<const uint4*>(As + (g + 8 + 16 * mt) * Cfg::kPitch + k0 + cc * 32);
      mma_bf16_16816(acc + 4 * mt, lo.x, hi.x, lo.y, hi.y, w.x, w.y);
      mma_bf16_16816(acc + 4 * mt, lo.z, hi.z, lo.w, hi.w, w.z, w.w);
    }
  }
}

// A = rnd((x * rsqrt(mean(x^2) + eps)) * (1 + w)) for M rows of 1024 -> shared memory. Warp w normalises rows w, w + 8.
template <int R>
__device__ __forceinline__ void stage_norm(bf16* As, const bf16* x, const float* gain, int M) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll 1
  for (int rr = 0; rr < R / 8; ++rr) {
    const int row = warp + rr * 8;
    if (row >= M) continue;
    uint4 v[4];
    float f[4][8];
    float ss = 0.0f;
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = ldcg_u4(x + (int64_t)row * kHidden + (i * 32 + lane) * 8);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v[i]);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 t = __bfloat1622float2(h[j]);
        f[i][2 * j] = t.x;
        f[i][2 * j + 1] = t.y;
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
        oh[j] = __floats2bfloat162_rn((f[i][2 * j] * r) * gv[2 * j], (f[i][2 * j + 1] * r) * gv[2 * j + 1]);
      *reinterpret_cast<uint4*>(As + row * MegaCfg<R>::kPitch + k) = o;
    }
  }
}

// plain copy of M rows x K (K multiple of 8) starting at column `kofs` of `src` into the staged operand (row pitch `pitch`
// elements); 8 independent 16-byte loads in flight per thread
__device__ __forceinline__ void stage_copy(bf16* As, int pitch, const bf16* src, int ld, int M, int K, int kofs) {
  const int per_row = K >> 3;
  const int total = M * per_row;
  for (int i0 = threadIdx.x; i0 < total; i0 += 256 * 8) {
    uint4 v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int i = i0 + j * 256;
      if (i < total) {
        const int row = i / per_row, c8 = i - row * per_row;
        v[j] = ldcg_u4(src + (int64_t)row * ld + kofs + c8 * 8);
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int i = i0 + j * 256;
      if (i < total) {
        const int row = i / per_row, c8 = i - row * per_row;
        *reinterpret_cast<uint4*>(As + row * pitch + c8 * 8) = v[j];
      }
    }
  }
}

template <int R>
__global__ void __launch_bounds__(kThreads, 1) denoise_layers_kernel(const MegaParams p) {
  using Cfg = MegaCfg<R>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
  uint8_t* ring = smem;
  bf16* As = reinterpret_cast<bf16*>(smem + Cfg::kRingBytes);
  float* red = reinterpret_cast<float*>(smem + Cfg::kRingBytes + Cfg::kABytes);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kRingBytes + Cfg::kABytes + Cfg::kRedBytes);
  uint64_t* empty_bar = full_bar + Cfg::kSlots;

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
  pdl_wait();

  const int16_t* tab = p.units + (int64_t)cta * 4 * (kMegaMaxUnits + 1);
  GridBarrier gb;
  if (warp < 8) grid_barrier_init(gb, reinterpret_cast<unsigned long long*>(p.sync), (unsigned)G);

  if (warp == 8) {
    // ------------------------------ producer: stream this CTA's weight chunks in consumption order ---------------
    const uint64_t policy = policy_evict_first();
    int slot = 0;
    uint32_t phase = 0;
    for (int layer = 0; layer < p.n_layers; ++layer) {
      const MegaLayerW& lw = p.L[layer];
#pragma unroll 1
      for (int ph = 0; ph < 4; ++ph) {
        const bf16* W = ph == MEGA_QKV ? lw.wqkv : ph == MEGA_O ? lw.wo : ph == MEGA_GU ? lw.wgu : lw.wdown;
        const int K = ph == MEGA_O ? kQDim : ph == MEGA_DOWN ? kInter : kHidden;
        const int nch = ph == MEGA_DOWN ? 4 : 2;
        const bool along_k = ph == MEGA_O || ph == MEGA_DOWN;
        const int16_t* t = tab + ph * (kMegaMaxUnits + 1);
        const int cnt = t[0];
        for (int i = 0; i < cnt; ++i) {
          const int u = t[1 + i];
          for (int ch = 0; ch < nch; ++ch) {
            mbar_wait(&empty_bar[slot], phase ^ 1);
            if (lane == 0) mbar_arrive_expect_tx(&full_bar[slot], 8 * kChunkK * 2);
            __syncwarp();
            if (lane < 8) {
              const int row = unit_row(ph, u, along_k ? 0 : ch, lane);
              const bf16* src = W + (int64_t)row * K + (along_k ? ch * kChunkK : 0);
              bulk_g2s(ring + slot * kSlotBytes + lane * kRowPitch * 2, src, kChunkK * 2, &full_bar[slot], policy);
            }
            if (++slot == Cfg::kSlots) { slot = 0; phase ^= 1; }
          }
        }
      }
    }
    return;
  }

  // ------------------------------ consumers ------------------------------------------------------------------------
  const int g = lane >> 2, t4 = lane & 3;
  int slot = 0;
  uint32_t phase = 0;

  // Runs this CTA's units of one GEMM phase. Units are processed in batches (4 per batch for M <= 8 rows, 2 for M <= 16):
  // every warp accumulates its k-slice of all the batch's chunks, then ONE cross-warp reduction + ONE CTA barrier + ONE
  // epilogue pass serve the whole batch, the epilogues of the batch's units running side by side on disjoint thread groups
  // (round 1 paid a reduction, a barrier and a 128-thread epilogue per unit: ~1.2 us each on the critical path, 4 units in
  // the gate/up phase). `epi(u, row, col, t0, t1)` is called with row < M; the fp32 sum over the 8 warps keeps its fixed
  // order, so results are bit-identical to the per-unit form.
  constexpr int kUB = R == 8 ? 4 : R == 16 ? 2 : 1;   // units per batch
  constexpr int kEpiT = R * 8;                        // epilogue (row, col) pairs per unit
  constexpr int kAccN = 4 * Cfg::kMT;
  // `a_src` / `a_ld`: the activation source of an along-K phase, needed only when the operand is staged chunk by chunk
  auto run_units = [&](int ph, const bf16* a_src, int a_ld, auto&& epi) {
    const int nch = ph == MEGA_DOWN ? 4 : 2;
    const bool along_k = ph == MEGA_O || ph == MEGA_DOWN;
    const int16_t* t = tab + ph * (kMegaMaxUnits + 1);
    const int cnt = t[0];
    for (int i0 = 0; i0 < cnt; i0 += kUB) {
      const int nb = min(kUB, cnt - i0);
      float acc[kUB][2][kAccN];
#pragma unroll
      for (int b = 0; b < kUB; ++b)
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
          for (int e = 0; e < kAccN; ++e) acc[b][a][e] = 0.0f;
#pragma unroll
      for (int b = 0; b < kUB; ++b) {
        if (b < nb) {
          for (int ch = 0; ch < nch; ++ch) {
            if constexpr (Cfg::kChunkedA) {
              if (along_k && (ch > 0 || i0 > 0)) {  // next 1024-k slice of the activations (slice 0 was staged by the caller)
                consumer_sync();
                stage_copy(As, Cfg::kPitch, a_src, a_ld, p.M, kChunkK, ch * kChunkK);
                consumer_sync();
              }
            }
            mbar_wait(&full_bar[slot], phase);
            const bf16* sl = reinterpret_cast<const bf16*>(ring + slot * kSlotBytes);
            if (along_k) mma_chunk<R>(acc[b][0], As, ch, sl, warp, g, t4);
            else if (ch == 0) mma_chunk<R>(acc[b][0], As, 0, sl, warp, g, t4);
            else mma_chunk<R>(acc[b][1], As, 0, sl, warp, g, t4);
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_bar[slot]);
            if (++slot == Cfg::kSlots) { slot = 0; phase ^= 1; }
          }
        }
      }
      // scratch layout [unit][warp][acc][row][col]
#pragma unroll
      for (int b = 0; b < kUB; ++b) {
        if (b < nb) {
          float* rp = red + b * (8 * 2 * R * 8);
#pragma unroll
          for (int a = 0; a < 2; ++a) {
#pragma unroll
            for (int mt = 0; mt < Cfg::kMT; ++mt) {
              *reinterpret_cast<float2*>(rp + ((warp * 2 + a) * R + g + 16 * mt) * 8 + t4 * 2) =
                  make_float2(acc[b][a][4 * mt + 0], acc[b][a][4 * mt + 1]);
              if constexpr (R >= 16)
                *reinterpret_cast<float2*>(rp + ((warp * 2 + a) * R + g + 8 + 16 * mt) * 8 + t4 * 2) =
                    make_float2(acc[b][a][4 * mt + 2], acc[b][a][4 * mt + 3]);
            }
          }
        }
      }
      consumer_sync();
      for (int e = threadIdx.x; e < kUB * kEpiT; e += 256) {
        const int b = e / kEpiT, rem = e % kEpiT;
        const int row = rem >> 3, col = rem & 7;
        if (b < nb && row < p.M) {
          const float* rp = red + b * (8 * 2 * R * 8);
          float t0 = 0.0f, t1 = 0.0f;
#pragma unroll
          for (int w = 0; w < 8; ++w) {
            t0 += rp[((w * 2 + 0) * R + row) * 8 + col];
            t1 += rp[((w * 2 + 1) * R + row) * 8 + col];
          }
          epi((int)t[1 + i0 + b], row, col, t0, t1);
        }
      }
      if (i0 + kUB < cnt) consumer_sync();  // the scratch is rewritten by the next batch
    }
  };
  auto has_units = [&](int ph) -> bool { return tab[ph * (kMegaMaxUnits + 1)] > 0; };

  for (int layer = 0; layer < p.n_layers; ++layer) {
    const MegaLayerW& lw = p.L[layer];
    const bf16* x_in = layer == 0 ? p.x0 : p.x;
    bf16* kc = p.kcache + (int64_t)layer * p.kc_layer_stride;
    bf16* vt = p.vtcache + (int64_t)layer * p.vt_layer_stride;

    // ---- phase 1: q|k|v = RMSNorm(x) . Wqkv^T, RoPE, caches ------------------------------------------------------
    MEGA_STAMP(0);
    prefetch_gain_l2(lw.g2);
    if (layer > 0) grid_barrier(gb);
    MEGA_STAMP(1);
    stage_norm<R>(As, x_in, lw.g1, p.M);
    consumer_sync();
    MEGA_STAMP(2);
    run_units(MEGA_QKV, nullptr, 0, [&](int u, int row, int col, float t0, float t1) {
      const int head = u >> 4, d = (u & 15) * 8 + col;
      const float a = bf16_round(t0), b = bf16_round(t1);
      float o1 = a, o2 = b;
      if (head < 9) {
        int pos = p.pos_ids[row];
        if (pos < 0) pos = 0;
        const float cs = __ldg(p.rope_cos + (int64_t)pos * 128 + d), sn = __ldg(p.rope_sin + (int64_t)pos * 128 + d);
        o1 = a * cs - b * sn;
        o2 = b * cs + a * sn;
      }
      const int bb = row / p.S, t = row % p.S;
      if (head < 8) {
        bf16* dst = p.q + (int64_t)row * kQDim + head * 256 + d;
        dst[0] = to_bf16(o1);
        dst[128] = to_bf16(o2);
      } else if (head == 8) {
        bf16* dst = kc + ((int64_t)bb * p.kv_pitch + p.P + t) * 256 + d;
        dst[0] = to_bf16(o1);
        dst[128] = to_bf16(o2);
      } else {
        bf16* dst = vt + (int64_t)bb * 256 * p.vt_pitch + p.P + t;
        dst[(int64_t)d * p.vt_pitch] = to_bf16(o1);
        dst[(int64_t)(d + 128) * p.vt_pitch] = to_bf16(o2);
      }
    });

    // ---- phase 2a: scores, one warp per (sample, head, 16 keys) ---------------------------------------------------
    MEGA_STAMP(3);
    grid_barrier(gb);
    MEGA_STAMP(4);
    {
      const int n_units = p.B * 8 * (p.kp >> 4);
      for (int u = warp * G + cta; u < n_units; u += 8 * G) attn_scores_unit(p, kc, u, lane);
    }
    // ---- phase 2b: masked softmax + P.V, one warp per (sample, head, 16 dims) ----------------------------------------
    MEGA_STAMP(15);
    grid_barrier(gb);
    MEGA_STAMP(16);
    for (int u = cta; u < p.B * 32; u += G) attn_pv_unit(p, vt, reinterpret_cast<uint8_t*>(As), u, warp, lane);

    // ---- phase 3: y = x + att . Wo^T -------------------------------------------------------------------------------
    MEGA_STAMP(5);
    grid_barrier(gb);
    MEGA_STAMP(6);
    if (has_units(MEGA_O)) stage_copy(As, Cfg::kPitch, p.att, kQDim, p.M, Cfg::kChunkedA ? kChunkK : kQDim, 0);
    consumer_sync();
    MEGA_STAMP(7);
    run_units(MEGA_O, p.att, kQDim, [&](int u, int row, int col, float t0, float) {
      const int n = u * 8 + col;
      const unsigned short xb = __ldcg(reinterpret_cast<const unsigned short*>(x_in) + (int64_t)row * kHidden + n);
      const float v = bf16_round(t0) + __bfloat162float(__ushort_as_bfloat16(xb));
      p.y[(int64_t)row * kHidden + n] = to_bf16(v);
    });

    // ---- phase 4: act = gelu_tanh(h . Wg^T) * (h . Wu^T), h = RMSNorm(y) -------------------------------------------
    MEGA_STAMP(8);
    grid_barrier(gb);
    MEGA_STAMP(9);
    stage_norm<R>(As, p.y, lw.g2, p.M);
    consumer_sync();
    MEGA_STAMP(10);
    run_units(MEGA_GU, nullptr, 0, [&](int u, int row, int col, float t0, float t1) {
      const float gt = bf16_round(t0), up = bf16_round(t1);
      const float act = bf16_round(gelu_tanh_fast(gt));
      p.act[(int64_t)row * kInter + u * 8 + col] = to_bf16(act * up);
    });

    // ---- phase 5: x = y + act . Wdown^T ----------------------------------------------------------------------------
    MEGA_STAMP(11);
    grid_barrier(gb);
    MEGA_STAMP(12);
    if (has_units(MEGA_DOWN)) stage_copy(As, Cfg::kPitch, p.act, kInter, p.M, Cfg::kChunkedA ? kChunkK : kInter, 0);
    consumer_sync();
    MEGA_STAMP(13);
    run_units(MEGA_DOWN, p.act, kInter, [&](int u, int row, int col, float t0, float) {
      const int n = u * 8 + col;
      const unsigned short yb = __ldcg(reinterpret_cast<const unsigned short*>(p.y) + (int64_t)row * kHidden + n);
      const float v = bf16_round(t0) + __bfloat162float(__ushort_as_bfloat16(yb));
      p.x[(int64_t)row * kHidden + n] = to_bf16(v);
    });
    MEGA_STAMP(14);
    if (layer + 1 < p.n_layers) prefetch_gain_l2(p.L[layer + 1].g1);
  }
}

}  // namespace

// Deals the work units of the four GEMM phases to `G` CTAs, balanced PER PHASE (the kernel is latency- not bandwidth-
// bound: a phase lasts as long as its busiest CTA): o_proj / down units (128 each) go to CTAs 0.. round-robin, the
// qkv (160) and gate/up (512) units round-robin from the LAST CTA downwards, so that the CTAs with a spare unit there
// are the ones idle in o_proj / down. Layout: [cta][phase][1 + kMegaMaxUnits] = count, unit ids.
int mega_build_units(int G, std::vector<int16_t>& tab) {
  PI0_REQUIRE(G >= 64 && G <= 1024, "denoise mega kernel: %d CTAs outside [64, 1024]", G);
  const int W = kMegaMaxUnits + 1;
  tab.assign((size_t)G * 4 * W, 0);
  auto push = [&](int cta, int ph, int u) -> bool {
    int16_t* t = tab.data() + ((size_t)cta * 4 + ph) * W;
    if (t[0] >= kMegaMaxUnits) return false;
    t[1 + t[0]] = (int16_t)u;
    ++t[0];
    return true;
  };
  bool ok = true;
  for (int u = 0; u < kHidden / 8; ++u) ok = ok && push(u % G, MEGA_O, u) && push(u % G, MEGA_DOWN, u);
  for (int u = 0; u < kInter / 8; ++u) ok = ok && push(G - 1 - (u % G), MEGA_GU, u);
  for (int u = 0; u < 10 * 16; ++u) ok = ok && push(G - 1 - (u % G), MEGA_QKV, u);
  PI0_REQUIRE(ok, "denoise mega kernel: unit table overflow");
  return 0;
}

bool mega_supported(int M, int S, int P) { return M >= 1 && M <= 48 && S <= 16 && P + S <= 32 * kMegaKeyChunks; }

int mega_denoise_layers(const MegaParams& p, int num_ctas, cudaStream_t s) {
  PI0_REQUIRE(mega_supported(p.M, p.S, p.P), "denoise mega kernel: M=%d S=%d P=%d unsupported", p.M, p.S, p.P);
  PI0_REQUIRE(p.n_layers >= 1 && p.n_layers <= 18, "denoise mega kernel: n_layers=%d", p.n_layers);
  const int R = p.M <= 8 ? 8 : p.M <= 16 ? 16 : p.M <= 32 ? 32 : 48;
  const size_t smem = R == 8 ? MegaCfg<8>::kSmem : R == 16 ? MegaCfg<16>::kSmem : R == 32 ? MegaCfg<32>::kSmem : MegaCfg<48>::kSmem;
  const size_t abytes = R == 8 ? MegaCfg<8>::kABytes : R == 16 ? MegaCfg<16>::kABytes : R == 32 ? MegaCfg<32>::kABytes
                                                                                                : MegaCfg<48>::kABytes;
  PI0_REQUIRE((size_t)16 * attn_p_pitch_bytes(p.kp) <= abytes && p.kp <= 32 * kMegaKeyChunks && p.kp % 32 == 0 && p.sc != nullptr,
              "denoise mega kernel: attention scratch does not fit");
  static PerDeviceOnce attr_once;
  if (attr_once.need()) {
    PI0_CHECK_CUDA(cudaFuncSetAttribute(denoise_layers_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MegaCfg<8>::kSmem));
    PI0_CHECK_CUDA(cudaFuncSetAttribute(denoise_layers_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MegaCfg<16>::kSmem));
    PI0_CHECK_CUDA(cudaFuncSetAttribute(denoise_layers_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MegaCfg<32>::kSmem));
    PI0_CHECK_CUDA(cudaFuncSetAttribute(denoise_layers_kernel<48>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MegaCfg<48>::kSmem));
    attr_once.mark();
  }
  cudaLaunchConfig_t lc = {};
  lc.gridDim = dim3((unsigned)num_ctas);
  lc.blockDim = dim3(kThreads);
  lc.dynamicSmemBytes = smem;
  lc.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeCooperative;  // all CTAs co-resident: the grid barrier cannot deadlock
  at[0].val.cooperative = 1;
  lc.attrs = at;
  lc.numAttrs = 1;
  if (R == 8) PI0_CHECK_CUDA(cudaLaunchKernelEx(&lc, denoise_layers_kernel<8>, p));
  else if (R == 16) PI0_CHECK_CUDA(cudaLaunchKernelEx(&lc, denoise_layers_kernel<16>, p));
  else if (R == 32) PI0_CHECK_CUDA(cudaLaunchKernelEx(&lc, denoise_layers_kernel<32>, p));
  else PI0_CHECK_CUDA(cudaLaunchKernelEx(&lc, denoise_layers_kernel<48>, p));
  ++g_launch_count;
  return 0;
}

}  // namespace pi0

#ifdef PI0_TIMELINE
extern "C" int pi0_debug_mega_timeline(unsigned long long* out) {
  return cudaMemcpyFromSymbol(out, pi0::g_mega_tl, sizeof(pi0::g_mega_tl)) == cudaSuccess ? 0 : -1;
}
#endif
