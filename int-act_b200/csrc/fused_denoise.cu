// Fused denoise step for rollout batches (16 < M = B * S <= ~640 suffix rows): ALL 18 action-expert layers of one Euler step
// in ONE persistent cooperative launch.
//
// Why: at batch 64 the denoise loop is 2 % of the FLOPs and 27 % of the step (round 1: 26.8 of 100.7 ms): 1260 launches of
// 12 - 22 us each at M = 320 rows (7 per layer: rmsnorm, q|k|v + RoPE, attention, o_proj, rmsnorm, gate/up, down) against a
// roofline of ~1.6 ms. Every one of those kernels is latency-, not throughput-bound (few tiles, a short K loop, a cluster
// reduction, launch + drain), so the fix is not a faster GEMM but fewer kernel boundaries. Here one CTA per SM stays resident
// and walks the layer as a sequence of grid-wide phases separated by grid barriers (~2 us each instead of ~8 us of launch
// gap + prologue + drain):
//   * GEMM phases run the tcgen05 pipeline of gemm_tc.cu (warp 0 = TMA producer, warp 1 = MMA issuer with a double-buffered
//     TMEM accumulator, warps 4-7 = epilogue) over (m tile, n tile, k split) work items; every GEMM is split along K so that
//     ~one item per SM exists whatever M is, and its epilogue is uniform: the fp32 partial tile goes to an L2-resident scratch;
//   * REDUCE phases (all threads, CUDA cores) sum the partials in fixed split order (deterministic) and apply what used to be
//     the GEMM epilogues AND the stand-alone norm kernels: RoPE + in-place K / V^T cache write for q|k|v, residual + the NEXT
//     RMSNorm for o_proj / down (rows are complete here, so the norm is free), GeGLU for gate/up;
//   * the ATTENTION phase runs the rollout-batch attention unit of attn_batch.cuh (K / V^T of a sample staged in shared
//     memory by bulk TMA copies, mma.sync) over (sample, head group) units.
// Rounding points are unchanged: Linear outputs are rounded once (here: after the split sum), norms / RoPE /
// softmax in fp32, rounded where the reference rounds. Reference arithmetic: [mirror] src/model/mvla/expert.py:128-179,
// 387-409; mllm.py:30-53,265-317.
#include <algorithm>
#include <vector>

#include "attn_batch.cuh"
#include "fused_denoise.cuh"

namespace pi0 {

namespace {

constexpr int kFThreads = 256;   // warp 0 = TMA producer, warp 1 = MMA issuer, warps 4-7 = epilogue (2-3 idle in GEMM phases);
                                 // 8 warps keep the attention / reduce phases at up to 255 registers per thread
constexpr int kBM = 128, kBK = 64;
constexpr int kStages = 4;
constexpr int kABytes = kBM * kBK * 2;          // 16 KB
constexpr int kBBytesMax = 256 * kBK * 2;       // 32 KB
constexpr int kStageBytes = kABytes + kBBytesMax;
constexpr int kRingBytes = kStages * kStageBytes;                                        // 196608
constexpr int kAttnBytes = (352 + battn::kRows) * battn::kRowB + battn::kRows * battn::kSPitch;  // 238592 upper bound
constexpr int kHidden = 1024, kInter = 4096, kQDim = 2048, kQkvN = 2560;
constexpr float kEps = 1e-6f;

#ifdef PI0_TIMELINE
__device__ unsigned long long g_fused_tl[4][32];
#define F_STAMP(i) do { if (layer == 1 && threadIdx.x == 0 && (blockIdx.x == 0 || blockIdx.x == 7 || blockIdx.x == 64 || blockIdx.x == 140)) \
  g_fused_tl[blockIdx.x == 0 ? 0 : blockIdx.x == 7 ? 1 : blockIdx.x == 64 ? 2 : 3][i] = global_timer_ns(); } while (0)
#else
#define F_STAMP(i) do { } while (0)
#endif

__host__ __device__ inline int attn_smem_need(int P, int S) {
  return ((P + S + 15) / 16 * 16 + battn::kRows) * battn::kRowB + battn::kRows * battn::kSPitch;
}

// split-K plan of one GEMM for M rows on G CTAs: about one work item per CTA, at least 2 k-blocks per item
struct SplitPlan { int m_tiles, splits, kb_per_split, items; };
__host__ __device__ inline SplitPlan plan_splits(int M, int n_tiles, int num_k, int G) {
  SplitPlan sp;
  sp.m_tiles = (M + kBM - 1) / kBM;
  const int base = sp.m_tiles * n_tiles;
  int want = G / base;
  if (want < 1) want = 1;
  int kbps = (num_k + want - 1) / want;
  if (kbps < 2) kbps = 2;
  if (kbps > num_k) kbps = num_k;
  sp.kb_per_split = kbps;
  sp.splits = (num_k + kbps - 1) / kbps;
  sp.items = base * sp.splits;
  return sp;
}

__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_add_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void named_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

// Grid barrier over all CTAs (monotonic 64-bit counter, never reset: CUDA-graph replays need no memset; see denoise_mega.cu).
// The generic-proxy stores of a phase are read through TMA (async proxy) in the next one: proxy fences on both sides.
struct GridBarrier {
  unsigned long long* count;
  unsigned long long target, step;
};
__device__ __forceinline__ void grid_sync(GridBarrier& gb) {
  fence_proxy_async_all();
  __syncthreads();
  if (threadIdx.x == 0) {
    red_release_add_u64(gb.count, 1ull);
    if (ld_acquire_u64(gb.count) < gb.target) {
      const uint64_t t0 = global_timer_ns();
      uint32_t spins = 0;
      while (ld_acquire_u64(gb.count) < gb.target) {
        if (((++spins) & 0x3ff) == 0 && global_timer_ns() - t0 > 2000000000ull) asm volatile("trap;");
      }
    }
    gb.target += gb.step;
  }
  __syncthreads();
  fence_proxy_async_all();
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

__device__ __forceinline__ void ld8_bf16(const bf16* p, float* v) {
  const uint4 u = __ldcg(reinterpret_cast<const uint4*>(p));
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __bfloat1622float2(h[i]);
    v[2 * i] = f.x;
    v[2 * i + 1] = f.y;
  }
}
__device__ __forceinline__ void st8_bf16(bf16* p, const float* v) {
  uint4 o;
  o.x = bf16x2_bits(v[1], v[0]);
  o.y = bf16x2_bits(v[3], v[2]);
  o.z = bf16x2_bits(v[5], v[4]);
  o.w = bf16x2_bits(v[7], v[6]);
  *reinterpret_cast<uint4*>(p) = o;
}
// sum of the split partials of 8 consecutive columns, in split order; the loads of 4 splits are in flight together
// (a rolled loop pays one L2 round trip per split: the reduce phases took ~10 us instead of ~2)
__device__ __forceinline__ void sum_partials8(const float* partial, int splits, int64_t split_stride, int64_t off, float* v) {
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = 0.0f;
  for (int s0 = 0; s0 < splits; s0 += 4) {
    float4 a[4], b[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (s0 + k < splits) {
        a[k] = __ldcg(reinterpret_cast<const float4*>(partial + (s0 + k) * split_stride + off));
        b[k] = __ldcg(reinterpret_cast<const float4*>(partial + (s0 + k) * split_stride + off + 4));
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (s0 + k < splits) {
        v[0] += a[k].x; v[1] += a[k].y; v[2] += a[k].z; v[3] += a[k].w;
        v[4] += b[k].x; v[5] += b[k].y; v[6] += b[k].z; v[7] += b[k].w;
      }
    }
  }
}
// two column groups at once (16 loads in flight)
__device__ __forceinline__ void sum_partials8x2(const float* partial, int splits, int64_t split_stride, int64_t off0, int64_t off1,
                                                float* v0, float* v1) {
#pragma unroll
  for (int i = 0; i < 8; ++i) { v0[i] = 0.0f; v1[i] = 0.0f; }
  for (int s0 = 0; s0 < splits; s0 += 4) {
    float4 a[4], b[4], c[4], d[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (s0 + k < splits) {
        const float* base = partial + (s0 + k) * split_stride;
        a[k] = __ldcg(reinterpret_cast<const float4*>(base + off0));
        b[k] = __ldcg(reinterpret_cast<const float4*>(base + off0 + 4));
        c[k] = __ldcg(reinterpret_cast<const float4*>(base + off1));
        d[k] = __ldcg(reinterpret_cast<const float4*>(base + off1 + 4));
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (s0 + k < splits) {
        v0[0] += a[k].x; v0[1] += a[k].y; v0[2] += a[k].z; v0[3] += a[k].w;
        v0[4] += b[k].x; v0[5] += b[k].y; v0[6] += b[k].z; v0[7] += b[k].w;
        v1[0] += c[k].x; v1[1] += c[k].y; v1[2] += c[k].z; v1[3] += c[k].w;
        v1[4] += d[k].x; v1[5] += d[k].y; v1[6] += d[k].z; v1[7] += d[k].w;
      }
    }
  }
}

struct PipeState {
  int stage = 0;
  uint32_t phase = 0;
  int acc = 0;
  uint32_t acc_phase = 0;
};

struct Smem {
  uint8_t* ring;
  uint64_t *full_bar, *empty_bar, *tfull_bar, *tempty_bar, *attn_bar;
  uint32_t tmem_base;
};

// ---- GEMM phase: fp32 partial tiles of A[M, K] . W[N, K]^T ----------------------------------------------------------------
__device__ __noinline__ void gemm_phase(const FusedGemmDesc* d, const FusedParams& p, const Smem& sm, PipeState& ps, int G) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int N = d->N, bn = d->bn, n_tiles = d->n_tiles, num_k = d->num_k;
  const SplitPlan sp = plan_splits(p.M, n_tiles, num_k, G);
  const int tiles_mn = sp.m_tiles * n_tiles;
  if (warp == 0) {
    if (lane == 0) {
      tma_prefetch_desc(&d->tmA);
      tma_prefetch_desc(&d->tmB);
      const uint32_t tx = d->stage_tx;
      for (int item = blockIdx.x; item < sp.items; item += G) {
        const int s = item / tiles_mn, t = item - s * tiles_mn;
        const int tm = t / n_tiles, tn = t - tm * n_tiles;
        const int kb0 = s * sp.kb_per_split, kb1 = min(num_k, kb0 + sp.kb_per_split);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait_quiet(&sm.empty_bar[ps.stage], ps.phase ^ 1);
          uint8_t* sa = sm.ring + ps.stage * kStageBytes;
          mbar_arrive_expect_tx(&sm.full_bar[ps.stage], tx);
          tma_load_4d(sa, &d->tmA, &sm.full_bar[ps.stage], kb * kBK, tm * kBM, 0, 0);
          tma_load_4d(sa + kABytes, &d->tmB, &sm.full_bar[ps.stage], kb * kBK, tn * bn, 0, 0);
          if (++ps.stage == kStages) { ps.stage = 0; ps.phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc = umma_idesc_bf16(kBM, bn);
    for (int item = blockIdx.x; item < sp.items; item += G) {
      const int s = item / tiles_mn;
      const int kb0 = s * sp.kb_per_split, kb1 = min(num_k, kb0 + sp.kb_per_split);
      mbar_wait_quiet(&sm.tempty_bar[ps.acc], ps.acc_phase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = sm.tmem_base + ps.acc * 256;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait_quiet(&sm.full_bar[ps.stage], ps.phase);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t sa = smem_u32(sm.ring + ps.stage * kStageBytes);
          const uint64_t da = umma_desc_k_sw128(sa);
          const uint64_t db = umma_desc_k_sw128(sa + kABytes);
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k)
            umma_bf16(tmem_d, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb > kb0 || k != 0) ? 1u : 0u);
          umma_commit(&sm.empty_bar[ps.stage]);
          if (kb == kb1 - 1) umma_commit(&sm.tfull_bar[ps.acc]);
        }
        __syncwarp();
        if (++ps.stage == kStages) { ps.stage = 0; ps.phase ^= 1; }
      }
      ps.acc ^= 1;
      if (ps.acc == 0) ps.acc_phase ^= 1;
    }
  } else if (warp >= 4) {
    const int quad = warp & 3;          // TMEM lanes [32 * quad, +32) are the ones this warp may read
    const int64_t split_stride = (int64_t)p.M * N;
    // Stores go through a per-warp transpose scratch: with the 32x32b TMEM layout a lane owns a ROW, so a direct
    // st.global.v4 touches 32 different 128-byte lines per instruction (LSU-bound); after the transpose
    // 8 consecutive lanes write one row's 128 bytes, 4 rows per instruction.
    float* scr = reinterpret_cast<float*>(sm.ring + kRingBytes) + quad * (32 * 36);
    for (int item = blockIdx.x; item < sp.items; item += G) {
      const int s = item / tiles_mn, t = item - s * tiles_mn;
      const int tm = t / n_tiles, tn = t - tm * n_tiles;
      const int row0 = tm * kBM + quad * 32;
      mbar_wait_quiet(&sm.tfull_bar[ps.acc], ps.acc_phase);
      tc_fence_after();
      const uint32_t taddr = sm.tmem_base + ((uint32_t)(quad * 32) << 16) + ps.acc * 256;
      float* pbase = p.partial + s * split_stride + (int64_t)row0 * N + tn * bn;
      for (int c = 0; c < bn / 32; ++c) {
        float v[32];
        tmem_ld_32x32(taddr + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<float4*>(scr + lane * 36 + j * 4) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        __syncwarp();
        if (tn * bn + c * 32 < N) {
#pragma unroll
          for (int pass = 0; pass < 8; ++pass) {
            const int r = pass * 4 + (lane >> 3), c4 = (lane & 7) * 4;
            const float4 o = *reinterpret_cast<const float4*>(scr + r * 36 + c4);
            if (row0 + r < p.M) *reinterpret_cast<float4*>(pbase + (int64_t)r * N + c * 32 + c4) = o;
          }
        }
        __syncwarp();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.tempty_bar[ps.acc]);
      ps.acc ^= 1;
      if (ps.acc == 0) ps.acc_phase ^= 1;
    }
  }
}

// ---- h = rnd((x * rsqrt(mean(x^2) + eps)) * g), one warp per row --------------------------------------------------------
__device__ __noinline__ void norm_phase(const bf16* x, bf16* h, const float* gain, int M, int G) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int row = blockIdx.x * (kFThreads / 32) + warp; row < M; row += G * (kFThreads / 32)) {
    float v[4][8];
    float ss = 0.0f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      ld8_bf16(x + (int64_t)row * kHidden + (i * 32 + lane) * 8, v[i]);
#pragma unroll
      for (int j = 0; j < 8; ++j) ss += v[i][j] * v[i][j];
    }
    ss = warp_sum_f(ss);
    const float r = rsqrtf(ss / (float)kHidden + kEps);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int k = (i * 32 + lane) * 8;
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gain + k)), g1 = __ldg(reinterpret_cast<const float4*>(gain + k + 4));
      const float gv[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = (v[i][j] * r) * gv[j];
      st8_bf16(h + (int64_t)row * kHidden + k, o);
    }
  }
}

// ---- out = rnd(rnd(sum partials) + res); h = RMSNorm(out) * gain (skipped when gain == nullptr); one warp per row --------
__device__ __noinline__ void reduce_res_norm_phase(const FusedParams& p, int splits, const bf16* res, bf16* out, bf16* h,
                                                      const float* gain, int G) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t split_stride = (int64_t)p.M * kHidden;
  for (int row = blockIdx.x * (kFThreads / 32) + warp; row < p.M; row += G * (kFThreads / 32)) {
    float v[4][8];
    float ss = 0.0f;
#pragma unroll
    for (int i = 0; i < 4; i += 2) {
      const int k0 = (i * 32 + lane) * 8, k1 = ((i + 1) * 32 + lane) * 8;
      float acc0[8], acc1[8], r0[8], r1[8];
      ld8_bf16(res + (int64_t)row * kHidden + k0, r0);
      ld8_bf16(res + (int64_t)row * kHidden + k1, r1);
      sum_partials8x2(p.partial, splits, split_stride, (int64_t)row * kHidden + k0, (int64_t)row * kHidden + k1, acc0, acc1);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        v[i][j] = bf16_round(bf16_round(acc0[j]) + r0[j]);
        v[i + 1][j] = bf16_round(bf16_round(acc1[j]) + r1[j]);
        ss += v[i][j] * v[i][j];
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) ss += v[i + 1][j] * v[i + 1][j];
      st8_bf16(out + (int64_t)row * kHidden + k0, v[i]);
      st8_bf16(out + (int64_t)row * kHidden + k1, v[i + 1]);
    }
    if (gain != nullptr) {
      ss = warp_sum_f(ss);
      const float r = rsqrtf(ss / (float)kHidden + kEps);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int k = (i * 32 + lane) * 8;
        const float4 g0 = __ldg(reinterpret_cast<const float4*>(gain + k)), g1 = __ldg(reinterpret_cast<const float4*>(gain + k + 4));
        const float gv[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (v[i][j] * r) * gv[j];
        st8_bf16(h + (int64_t)row * kHidden + k, o);
      }
    }
  }
}

// ---- q|k|v: sum partials, round, RoPE (q and k), q -> [M, 2048], k -> K cache row P + t, v -> V^T cache column P + t --------
__device__ __noinline__ void reduce_rope_phase(const FusedParams& p, int splits, bf16* kc, bf16* vt, int G) {
  const int64_t split_stride = (int64_t)p.M * kQkvN;
  const int total = p.M * 10 * 16;  // (row, head, 8-wide d group of the first 128 dims)
  for (int it = blockIdx.x * kFThreads + threadIdx.x; it < total; it += G * kFThreads) {
    const int dg = it & 15, head = (it >> 4) % 10, row = it / 160;
    const int d0 = dg * 8;
    float x1[8], x2[8], o1[8], o2[8];
    sum_partials8x2(p.partial, splits, split_stride, (int64_t)row * kQkvN + head * 256 + d0,
                    (int64_t)row * kQkvN + head * 256 + 128 + d0, x1, x2);
    if (head < 9) {
      int pos = p.pos_ids[row];
      if (pos < 0) pos = 0;
      const float4 c0 = __ldg(reinterpret_cast<const float4*>(p.rope_cos + (int64_t)pos * 128 + d0));
      const float4 c1 = __ldg(reinterpret_cast<const float4*>(p.rope_cos + (int64_t)pos * 128 + d0 + 4));
      const float4 s0 = __ldg(reinterpret_cast<const float4*>(p.rope_sin + (int64_t)pos * 128 + d0));
      const float4 s1 = __ldg(reinterpret_cast<const float4*>(p.rope_sin + (int64_t)pos * 128 + d0 + 4));
      const float cs[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
      const float sn[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float a = bf16_round(x1[j]), b = bf16_round(x2[j]);
        o1[j] = a * cs[j] - b * sn[j];
        o2[j] = b * cs[j] + a * sn[j];
      }
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) { o1[j] = x1[j]; o2[j] = x2[j]; }
    }
    const int bb = row / p.S, t = row % p.S;
    if (head < 8) {
      bf16* dst = p.q + (int64_t)row * kQDim + head * 256 + d0;
      st8_bf16(dst, o1);
      st8_bf16(dst + 128, o2);
    } else if (head == 8) {
      bf16* dst = kc + ((int64_t)bb * p.kv_pitch + p.P + t) * 256 + d0;
      st8_bf16(dst, o1);
      st8_bf16(dst + 128, o2);
    } else {
      bf16* dst = vt + (int64_t)bb * 256 * p.vt_pitch + p.P + t;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        dst[(int64_t)(d0 + j) * p.vt_pitch] = to_bf16(o1[j]);
        dst[(int64_t)(128 + d0 + j) * p.vt_pitch] = to_bf16(o2[j]);
      }
    }
  }
}

// ---- act = rnd(rnd(gelu_tanh(rnd(gate))) * rnd(up)); gate / up rows interleaved per 256-row weight tile (128 + 128) ------
__device__ __noinline__ void reduce_geglu_phase(const FusedParams& p, int splits, int G) {
  const int64_t split_stride = (int64_t)p.M * (2 * kInter);
  const int total = p.M * (kInter / 8);
  for (int it = blockIdx.x * kFThreads + threadIdx.x; it < total; it += G * kFThreads) {
    const int jg = it % (kInter / 8), row = it / (kInter / 8);
    const int j0 = jg * 8;
    const int gcol = (j0 >> 7) * 256 + (j0 & 127);
    float gt[8], up[8], o[8];
    sum_partials8x2(p.partial, splits, split_stride, (int64_t)row * (2 * kInter) + gcol,
                    (int64_t)row * (2 * kInter) + gcol + 128, gt, up);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = bf16_round(gelu_tanh_fast(bf16_round(gt[j]))) * bf16_round(up[j]);
    st8_bf16(p.act + (int64_t)row * kInter + j0, o);
  }
}

// own register allocation for the attention unit (185 registers stand-alone): not inlined into the kernel body
__device__ __noinline__ void attention_phase(uint8_t* smem, uint64_t* bars, uint32_t* apar, const battn::BatchAttnParams ap, int B,
                                             int G) {
  const int groups = 8 / ap.hpg;
  uint32_t par = *apar;
  for (int u = blockIdx.x; u < B * groups; u += G)
    battn::batch_attn_unit(smem, bars, par, ap, u / groups, (u % groups) * ap.hpg, [] { __syncthreads(); });
  *apar = par;
}

__global__ void __launch_bounds__(kFThreads, 1) fused_denoise_kernel(const FusedParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int main_bytes = max(kRingBytes, attn_smem_need(p.P, p.S));
  Smem sm;
  sm.ring = smem;
  sm.full_bar = reinterpret_cast<uint64_t*>(smem + ((main_bytes + 127) & ~127));
  sm.empty_bar = sm.full_bar + kStages;
  sm.tfull_bar = sm.empty_bar + kStages;
  sm.tempty_bar = sm.tfull_bar + 2;
  sm.attn_bar = sm.tempty_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm.attn_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int G = gridDim.x;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&sm.full_bar[s], 1);
      mbar_init(&sm.empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&sm.tfull_bar[a], 1);
      mbar_init(&sm.tempty_bar[a], 4);
      mbar_init(&sm.attn_bar[a], 1);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  sm.tmem_base = *tmem_slot;
  pdl_wait();

  GridBarrier gb;
  gb.count = reinterpret_cast<unsigned long long*>(p.sync);
  gb.step = (unsigned long long)G;
  gb.target = 0;
  if (threadIdx.x == 0) gb.target = (ld_acquire_u64(gb.count) / G) * G + G;

  PipeState ps;
  uint32_t apar = 0;
  battn::BatchAttnParams ap;
  ap.q = p.q; ap.att = p.att; ap.pad = p.pad; ap.cum = p.cum; ap.S = p.S; ap.P = p.P; ap.kv_pitch = p.kv_pitch;
  ap.vt_pitch = p.vt_pitch; ap.kp = p.kp; ap.hpg = p.hpg;
  norm_phase(p.x0, p.h, p.g1[0], p.M, G);
  grid_sync(gb);
  for (int layer = 0; layer < p.n_layers; ++layer) {
    const FusedGemmDesc* gd = p.gd + layer * 4;
    bf16* kc = p.kcache + (int64_t)layer * p.kc_layer_stride;
    bf16* vt = p.vtcache + (int64_t)layer * p.vt_layer_stride;
    const bf16* x_in = layer == 0 ? p.x0 : p.x;
    // q|k|v
    F_STAMP(0);
    gemm_phase(gd + 0, p, sm, ps, G);
    F_STAMP(1);
    grid_sync(gb);
    F_STAMP(2);
    reduce_rope_phase(p, plan_splits(p.M, gd[0].n_tiles, gd[0].num_k, G).splits, kc, vt, G);
    F_STAMP(3);
    grid_sync(gb);
    F_STAMP(4);
    // attention over the cached prefix + this step's suffix rows
    ap.kc = kc;
    ap.vt = vt;
    attention_phase(smem, sm.attn_bar, &apar, ap, p.B, G);
    F_STAMP(5);
    grid_sync(gb);
    F_STAMP(6);
    // o_proj + residual + post-attention norm
    gemm_phase(gd + 1, p, sm, ps, G);
    F_STAMP(7);
    grid_sync(gb);
    F_STAMP(8);
    reduce_res_norm_phase(p, plan_splits(p.M, gd[1].n_tiles, gd[1].num_k, G).splits, x_in, p.y, p.h, p.g2[layer], G);
    F_STAMP(9);
    grid_sync(gb);
    F_STAMP(10);
    // gate/up + GeGLU
    gemm_phase(gd + 2, p, sm, ps, G);
    F_STAMP(11);
    grid_sync(gb);
    F_STAMP(12);
    reduce_geglu_phase(p, plan_splits(p.M, gd[2].n_tiles, gd[2].num_k, G).splits, G);
    F_STAMP(13);
    grid_sync(gb);
    F_STAMP(14);
    // down + residual + the next layer's input norm
    gemm_phase(gd + 3, p, sm, ps, G);
    F_STAMP(15);
    grid_sync(gb);
    F_STAMP(16);
    reduce_res_norm_phase(p, plan_splits(p.M, gd[3].n_tiles, gd[3].num_k, G).splits, p.y, p.x, p.h,
                          layer + 1 < p.n_layers ? p.g1[layer + 1] : nullptr, G);
    F_STAMP(17);
    if (layer + 1 < p.n_layers) grid_sync(gb);
    F_STAMP(18);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(sm.tmem_base);
}

constexpr int kBnQkv = 256, kBnO = 128, kBnGu = 256, kBnDown = 128;

size_t smem_bytes(int P, int S) { return (size_t)std::max(kRingBytes, attn_smem_need(P, S)) + 128 + 256 + 1024; }

}  // namespace

bool fused_supported(int M, int S, int P, int B, int num_ctas) {
  return M > 16 && M <= 1024 && S >= 2 && S <= 6 && P + S <= battn::kKeysMax && P % 8 == 0 && B >= 1 && num_ctas >= 32 &&
         smem_bytes(P, S) <= 232448;
}

size_t fused_desc_bytes() { return (size_t)18 * 4 * sizeof(FusedGemmDesc); }

size_t fused_partial_bytes(int max_rows, int num_ctas) {
  // splits * M * N <= (G / base + 1) * M * N <= G * 128 * bn + M * N  (base >= M / 128 * N / bn items without splitting)
  return ((size_t)num_ctas * kBM * 256 + (size_t)max_rows * 2 * kInter) * sizeof(float) + 1024;
}

int fused_build_descs(std::vector<FusedGemmDesc>& host, int n_layers, int max_rows, const void* h, const void* att, const void* act,
                      const void* const* wqkv, const void* const* wo, const void* const* wgu, const void* const* wdown) {
  host.assign((size_t)18 * 4, FusedGemmDesc{});
  struct Op { const void* A; const void* const* W; int N, K, bn; };
  const Op ops[4] = {{h, wqkv, kQkvN, kHidden, kBnQkv}, {att, wo, kHidden, kQDim, kBnO}, {h, wgu, 2 * kInter, kHidden, kBnGu},
                     {act, wdown, kHidden, kInter, kBnDown}};
  for (int l = 0; l < n_layers; ++l) {
    for (int o = 0; o < 4; ++o) {
      FusedGemmDesc& d = host[(size_t)l * 4 + o];
      d.N = ops[o].N; d.K = ops[o].K; d.bn = ops[o].bn;
      d.n_tiles = (d.N + d.bn - 1) / d.bn;
      d.num_k = d.K / kBK;
      d.stage_tx = (unsigned)(kABytes + d.bn * kBK * 2);
      PI0_TRY(tc_make_map(&d.tmA, ops[o].A, d.K, max_rows, 1, 1, d.K, 0, 0, kBM, kBK, 128));
      PI0_TRY(tc_make_map(&d.tmB, ops[o].W[l], d.K, d.N, 1, 1, d.K, 0, 0, d.bn, kBK, 128));
    }
  }
  return 0;
}

int fused_denoise_layers(const FusedParams& p, int num_ctas, cudaStream_t s) {
  PI0_REQUIRE(fused_supported(p.M, p.S, p.P, p.B, num_ctas), "fused denoise kernel: M=%d S=%d P=%d unsupported", p.M, p.S, p.P);
  PI0_REQUIRE(p.n_layers >= 1 && p.n_layers <= 18 && p.gd != nullptr && p.partial != nullptr && p.sync != nullptr,
              "fused denoise kernel: bad arguments");
  // the split plan must fit the partial scratch
  const int nt[4] = {kQkvN / kBnQkv, kHidden / kBnO, 2 * kInter / kBnGu, kHidden / kBnDown};
  const int nk[4] = {kHidden / kBK, kQDim / kBK, kHidden / kBK, kInter / kBK};
  const int nn[4] = {kQkvN, kHidden, 2 * kInter, kHidden};
  for (int o = 0; o < 4; ++o) {
    const SplitPlan sp = plan_splits(p.M, nt[o], nk[o], num_ctas);
    PI0_REQUIRE((long long)sp.splits * p.M * nn[o] <= p.partial_elems, "fused denoise kernel: partial scratch too small");
  }
  PI0_REQUIRE((size_t)256 * p.vt_pitch * 2 + 64 <= (size_t)((p.P + p.S + 15) / 16 * 16 + battn::kRows) * battn::kRowB &&
                  p.vt_pitch % 8 == 0 && p.vt_pitch >= p.P + p.S && p.kv_pitch >= p.P + p.S && (p.hpg == 4 || p.hpg == 8),
              "fused denoise kernel: cache pitches / head grouping unsupported");
  const size_t smem = smem_bytes(p.P, p.S);
  static PerDeviceOnce attr_once;
  if (attr_once.need()) {
    PI0_CHECK_CUDA(cudaFuncSetAttribute(fused_denoise_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
    attr_once.mark();
  }
  cudaLaunchConfig_t lc = {};
  lc.gridDim = dim3((unsigned)num_ctas);
  lc.blockDim = dim3(kFThreads);
  lc.dynamicSmemBytes = smem;
  lc.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeCooperative;  // all CTAs co-resident: the grid barrier cannot deadlock
  at[0].val.cooperative = 1;
  lc.attrs = at;
  lc.numAttrs = 1;
  PI0_CHECK_CUDA(cudaLaunchKernelEx(&lc, fused_denoise_kernel, p));
  ++g_launch_count;
  return 0;
}

}  // namespace pi0

#ifdef PI0_TIMELINE
extern "C" int pi0_debug_fused_timeline(unsigned long long* out) {
  return cudaMemcpyFromSymbol(out, pi0::g_fused_tl, sizeof(pi0::g_fused_tl)) == cudaSuccess ? 0 : -1;
}
#endif
