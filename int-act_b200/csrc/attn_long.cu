// Fused attention on tcgen05 for LONG key ranges (up to 1024 keys): the prefix pass and the denoise-step suffix attention of
// the multi-camera / long-horizon configurations (LIBERO: P = 584 prefix keys, 17 suffix tokens; 3-camera hf_pi0: P = 840),
// which exceed the one-TMEM-tile kernels (attn_prefill.cu: 336 keys; attn_batch.cu: 352 keys, 6 suffix tokens) and fell back to
// the scores GEMM -> softmax -> P.V GEMM chain / the one-CTA-per-sample mma.sync kernel (27 ms of a 193 ms LIBERO step).
// Reference: eager_attention_forward ([mirror] /root/reference/src/model/mvla/mllm.py:265-317) under the masks of
// make_att_2d_masks (modeling_mvla.py:98-128): prefix rows see the valid prefix keys; suffix rows (modeling_mvla.py:686-692,
// 817-821) see the valid prefix keys and the suffix keys whose mask cumsum is <= their own.
//
// MQA: the 8 query heads of a token are 8 rows against one K / V. One CTA per SM, persistent over (sample, 128-query-row tile)
// items. Keys are processed in blocks of 256 (fp32 S block = 256 TMEM columns, O = the other 256). The reference rounds the
// NORMALISED probabilities to bf16 before P.V, so the row maximum and sum must be known before the first probability is formed:
//   pass 1   S_kb = Q K_kb^T per key block (tcgen05.mma M = 128, N = 256, head dim in 4 k-blocks of 64; Q stays resident in
//            shared memory, K streams through a 3 x 32 KiB TMA ring); 16 softmax warps (4 per TMEM lane quadrant, 64 keys each per
//            block) form s = rnd(rnd(acc) * 256^-0.5) under the mask and keep running (max, sum) per row
//   exchange one shared-memory exchange of the four key parts' (max, sum) -> row max M, row sum L
//   pass 2   S_kb again (S_kb+1 is issued BEFORE P.V of block kb, so it runs under the softmax of block kb);
//            p = rnd(exp(s - M) / L) as bf16 into the 128-byte-swizzled K-major P tile; O += P V_kb (V^T cache
//            rows staged by TMA through the same ring, M = 128, N = 256 dims, 4 k-blocks of 64 keys)
//   epilogue O -> rnd -> bf16 -> the idle P tile (four 128-byte-swizzled [128 rows x 64 dims] blocks) -> one TMA store per block
// Rounding points are those of the chain it replaces: scores and probabilities materialised in bf16, softmax in fp32.
#include "elementwise.cuh"
#include "gemm.cuh"

namespace pi0 {

namespace {

constexpr int kAlParts = 4;                       // softmax warps per TMEM lane quadrant
constexpr int kAlSmThreads = 128 * kAlParts;
constexpr int kAlThreads = 64 + kAlSmThreads;     // TMA warp, MMA warp, 16 softmax / epilogue warps
constexpr int kAlKeyBlock = 256;
constexpr int kAlMaxKeys = 1024;
constexpr uint32_t kAlQBytes = 4 * 128 * 128;     // Q tile: 4 head-dim blocks of [128 rows x 128 B]
constexpr uint32_t kAlStageBytes = 256 * 128;     // K block [256 keys x 64 dims] or V^T block [256 dims x 64 keys]
constexpr int kAlStages = 3;
constexpr uint32_t kAlPBytes = 4 * 128 * 128;     // P tile: 4 key blocks of [128 rows x 64 keys]
constexpr uint32_t kAlSmem = kAlQBytes + kAlStages * kAlStageBytes + kAlPBytes + 256 + 1024;   // 225.25 KiB
static_assert(kAlSmem <= 232448, "shared memory budget of one SM");

struct AlParams {
  const uint8_t* pad;   // [B][P]
  const int* cum;       // [n_suffix] cumsum of the suffix attention mask, or nullptr in prefix mode
  bf16* att;            // [B][R][256]
  int B, P, R;          // samples, prefix keys, query rows per sample (tokens * 8)
  int n_keys;           // P (prefix mode) or P + suffix tokens
  int n_suffix;         // 0 = prefix mode
  int tiles_per_b, total_items, n_kb;
};

__device__ __forceinline__ float al_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void al_bar() { asm volatile("bar.sync 1, %0;" ::"n"(kAlSmThreads) : "memory"); }

// visibility bits of the PREFIX keys in [key0, key0 + 64) (bit = pad byte != 0); P % 8 == 0, key0 % 64 == 0
__device__ __forceinline__ unsigned long long al_pad_bits64(const uint8_t* padb, int P, int key0) {
  unsigned long long m = 0;
#pragma unroll
  for (int q8 = 0; q8 < 8; ++q8) {
    const int k = key0 + q8 * 8;
    if (k < P) {   // an 8-byte group never straddles the prefix end
      const uint2 w = __ldg(reinterpret_cast<const uint2*>(padb + k));
      // byte != 0 -> bit: per-byte compare, then gather the four 0/1 bytes into a nibble with one multiply
      const uint32_t lo = ((__vcmpne4(w.x, 0u) & 0x01010101u) * 0x10204080u) >> 28;
      const uint32_t hi = ((__vcmpne4(w.y, 0u) & 0x01010101u) * 0x10204080u) >> 28;
      m |= (unsigned long long)(lo | (hi << 4)) << (q8 * 8);
    }
  }
  return m;
}
// suffix keys of [key0, key0 + 64) visible to query token tq: cum[key] <= cum[tq] (modeling_mvla.py:98-128)
__device__ __forceinline__ unsigned long long al_suffix_bits64(const AlParams& p, int key0, int tq) {
  unsigned long long m = 0;
  if (p.n_suffix > 0 && key0 + 64 > p.P && key0 < p.n_keys) {
    const int cq = __ldg(p.cum + tq);
    const int j0 = max(key0, p.P), j1 = min(key0 + 64, p.n_keys);
    for (int j = j0; j < j1; ++j)
      if (__ldg(p.cum + (j - p.P)) <= cq) m |= 1ull << (j - key0);
  }
  return m;
}

// the thread's 64 scores of one key block: TMEM -> s = rnd(rnd(acc) * 2^-4) under the mask (finite -2.3819763e38 like the
// reference, rounded to bf16 because the scores are materialised in bf16; keys beyond n_keys are -inf so that they drop out
// of the softmax). FULL: the whole 64-key range lies inside n_keys.
template <bool FULL>
__device__ __forceinline__ void al_scores(uint32_t taddr, unsigned long long vis, int key0, int n_keys, float* s) {
  const float kNegR = bf16_round(-2.3819763e38f);
  const uint32_t vlo = (uint32_t)vis, vhi = (uint32_t)(vis >> 32);
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    float v[32];
    tmem_ld_32x32(taddr + c * 32, v);
    tmem_ld_wait();
    const uint32_t vw = c == 0 ? vlo : vhi;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const float x = bf16_round(v[j]) * 0.0625f;   // exact in bf16: the reference's second rounding changes nothing
      const float y = (vw & (1u << j)) ? x : kNegR;
      if constexpr (FULL) s[c * 32 + j] = y;
      else s[c * 32 + j] = key0 + c * 32 + j < n_keys ? y : -INFINITY;
    }
  }
}

#ifdef PI0_TIMELINE
__device__ unsigned long long g_al_tl[64];
#define AL_STAMP(i) do { if (blockIdx.x == 0 && threadIdx.x == 64 && idx == (int)(blockIdx.x + gridDim.x)) g_al_tl[i] = global_timer_ns(); } while (0)
#else
#define AL_STAMP(i) do { } while (0)
#endif

__global__ void __launch_bounds__(kAlThreads, 1)
attn_long_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO, const AlParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* qs = smem;                                     // [4 dim blocks][128 rows][128 B] SW128
  uint8_t* ring = qs + kAlQBytes;                         // kAlStages x 32 KiB
  uint8_t* ptile = ring + kAlStages * kAlStageBytes;      // [4 key blocks of 64][128 rows][128 B] SW128
  // [max | sum][part][row] of the exchange between the passes, 4 KiB: aliases the head of the P tile, which is idle then (the
  // previous item's P.V products precede this item's S products in the tensor pipe, and these have all been consumed)
  float* red = reinterpret_cast<float*>(ptile);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(ptile + kAlPBytes);
  uint64_t* empty_bar = full_bar + kAlStages;
  uint64_t* q_full = empty_bar + kAlStages;
  uint64_t* q_empty = q_full + 1;    // every S product of the item complete -> Q may be overwritten
  uint64_t* s_full = q_empty + 1;    // S block accumulated                 (MMA -> softmax)
  uint64_t* s_free = s_full + 1;     // S block read out of TMEM            (softmax -> MMA)
  uint64_t* p_full = s_free + 1;     // P block in shared memory            (softmax -> MMA)
  uint64_t* p_free = p_full + 1;     // P.V of the block complete           (MMA -> softmax)
  uint64_t* o_full = p_free + 1;     // O accumulated                       (MMA -> epilogue)
  uint64_t* o_free = o_full + 1;     // O read out of TMEM                  (epilogue -> MMA)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_free + 1);
  unsigned long long* maskw = reinterpret_cast<unsigned long long*>(o_free + 2);   // [n_kb * 4] pad bits per 64 keys

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_trigger();
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmO);
    for (int s = 0; s < kAlStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(q_full, 1);
    mbar_init(q_empty, 1);
    mbar_init(s_full, 1);
    mbar_init(s_free, 4 * kAlParts);
    mbar_init(p_full, 4 * kAlParts);
    mbar_init(p_free, 1);
    mbar_init(o_full, 1);
    mbar_init(o_free, 4 * kAlParts);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int n_kb = p.n_kb;
  pdl_wait();

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0, ipar = 0;
      for (int idx = blockIdx.x; idx < p.total_items; idx += gridDim.x) {
        const int b = idx / p.tiles_per_b, mt = idx - b * p.tiles_per_b;
        mbar_wait(q_empty, ipar ^ 1);
        mbar_arrive_expect_tx(q_full, kAlQBytes);
#pragma unroll
        for (int dk = 0; dk < 4; ++dk) tma_load_4d(qs + dk * (128 * 128), &tmQ, q_full, dk * 64, mt * 128, b, 0);
        auto load_k = [&](int kb) {
          for (int dk = 0; dk < 4; ++dk) {          // K block kb, head-dim block dk
            mbar_wait(&empty_bar[stage], phase ^ 1);
            mbar_arrive_expect_tx(&full_bar[stage], kAlStageBytes);
            tma_load_4d(ring + stage * kAlStageBytes, &tmK, &full_bar[stage], dk * 64, kb * kAlKeyBlock, b, 0);
            if (++stage == kAlStages) { stage = 0; phase ^= 1; }
          }
        };
        auto load_v = [&](int kb) {
          for (int kk = 0; kk < 4; ++kk) {          // V^T, 64 keys of block kb
            mbar_wait(&empty_bar[stage], phase ^ 1);
            mbar_arrive_expect_tx(&full_bar[stage], kAlStageBytes);
            tma_load_4d(ring + stage * kAlStageBytes, &tmV, &full_bar[stage], kb * kAlKeyBlock + kk * 64, 0, b, 0);
            if (++stage == kAlStages) { stage = 0; phase ^= 1; }
          }
        };
        for (int kb = 0; kb < n_kb; ++kb) load_k(kb);   // pass 1
        load_k(0);                                      // pass 2, in the MMA warp's order: S(kb + 1) is issued before P.V(kb)
        for (int kb = 0; kb < n_kb; ++kb) {
          if (kb + 1 < n_kb) load_k(kb + 1);
          load_v(kb);
        }
        ipar ^= 1;
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer --------------------------------
    constexpr uint32_t idesc = umma_idesc_bf16(128, 256);
    const uint32_t tmem_s = tmem_base, tmem_o = tmem_base + 256;
    int stage = 0;
    uint32_t phase = 0, ipar = 0, s_par = 0, p_par = 0;
    for (int idx = blockIdx.x; idx < p.total_items; idx += gridDim.x) {
      mbar_wait(q_full, ipar);
      tc_fence_after();
      auto issue_s = [&](bool last_of_item) {
        mbar_wait(s_free, s_par ^ 1);                // the previous S block has been read out
        tc_fence_after();
        for (int dk = 0; dk < 4; ++dk) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint64_t da = umma_desc_k_sw128(smem_u32(qs + dk * (128 * 128)));
            const uint64_t db = umma_desc_k_sw128(smem_u32(ring + stage * kAlStageBytes));
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma_bf16(tmem_s, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (dk | k) != 0 ? 1u : 0u);
            umma_commit(&empty_bar[stage]);
            if (dk == 3) {
              umma_commit(s_full);
              if (last_of_item) umma_commit(q_empty);
            }
          }
          __syncwarp();
          if (++stage == kAlStages) { stage = 0; phase ^= 1; }
        }
        s_par ^= 1;
      };
      for (int kb = 0; kb < n_kb; ++kb) issue_s(false);       // pass 1
      issue_s(n_kb == 1);                                       // pass 2: S(kb + 1) runs under the softmax of block kb
      for (int kb = 0; kb < n_kb; ++kb) {
        if (kb + 1 < n_kb) issue_s(kb + 2 == n_kb);
        mbar_wait(p_full, p_par);                    // P block written (async-proxy visible)
        tc_fence_after();
        if (kb == 0) {
          mbar_wait(o_free, ipar ^ 1);               // the previous item's O has been read out
          tc_fence_after();
        }
        for (int kk = 0; kk < 4; ++kk) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint64_t da = umma_desc_k_sw128(smem_u32(ptile + kk * (128 * 128)));
            const uint64_t db = umma_desc_k_sw128(smem_u32(ring + stage * kAlStageBytes));
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma_bf16(tmem_o, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb | kk | k) != 0 ? 1u : 0u);
            umma_commit(&empty_bar[stage]);
            if (kk == 3) {
              umma_commit(p_free);
              if (kb == n_kb - 1) umma_commit(o_full);
            }
          }
          __syncwarp();
          if (++stage == kAlStages) { stage = 0; phase ^= 1; }
        }
        p_par ^= 1;
      }
      ipar ^= 1;
    }
  } else {
    // ------------------------------ softmax + epilogue ------------------------
    const int quad = warp & 3, part = (warp - 2) >> 2;
    const int r = quad * 32 + lane;                       // row of the tile = TMEM lane
    const uint32_t trow = tmem_base + ((uint32_t)(quad * 32) << 16);
    const float kLog2e = 1.4426950408889634f;
    float* rm = red;
    float* rs = red + kAlParts * 128;
    uint32_t ipar = 0, s_par = 0, p_par = 0;
    for (int idx = blockIdx.x; idx < p.total_items; idx += gridDim.x) {
      const int b = idx / p.tiles_per_b, mt = idx - b * p.tiles_per_b;
      const int row = mt * 128 + r;                       // query row of the sample: token row / 8, head row % 8
      const uint8_t* padb = p.pad + (int64_t)b * p.P;
      const int tq = min(row >> 3, max(p.n_suffix - 1, 0));
      AL_STAMP(0);
      // pad bits of the sample, one word per 64 keys, computed once per item (every row of a part shares them); the readers
      // of the previous item's table are past the exchange barriers
      if ((int)threadIdx.x - 64 < n_kb * 4) maskw[threadIdx.x - 64] = al_pad_bits64(padb, p.P, ((int)threadIdx.x - 64) * 64);
      if (threadIdx.x == 64) bulk_wait_read<0>();   // the previous item's O stores have left the P tile
      al_bar();
      // ---- pass 1: running (max, sum) over this thread's 64 keys of every block ----
      float m = -INFINITY, l = 0.0f;
      for (int kb = 0; kb < n_kb; ++kb) {
        const int key0 = kb * kAlKeyBlock + part * 64;
        const unsigned long long vis = maskw[kb * 4 + part] | al_suffix_bits64(p, key0, tq);
        AL_STAMP(1 + kb * 4);
        mbar_wait(s_full, s_par);
        tc_fence_after();
        AL_STAMP(2 + kb * 4);
        float s[64];
        if (key0 + 64 <= p.n_keys) al_scores<true>(trow + part * 64, vis, key0, p.n_keys, s);
        else al_scores<false>(trow + part * 64, vis, key0, p.n_keys, s);
        AL_STAMP(3 + kb * 4);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(s_free);
        s_par ^= 1;
        float mb = -INFINITY;
#pragma unroll
        for (int j = 0; j < 64; ++j) mb = fmaxf(mb, s[j]);
        if (mb > -INFINITY) {
          const float mn = fmaxf(m, mb);
          float sum = 0.0f;
#pragma unroll
          for (int j = 0; j < 64; ++j) sum += al_ex2((s[j] - mn) * kLog2e);   // subtract first: the mask value times log2e overflows
          l = l * al_ex2((m - mn) * kLog2e) + sum;        // m = -inf: l is 0 and ex2(-inf) = 0
          m = mn;
        }
        AL_STAMP(4 + kb * 4);
      }
      // ---- exchange between the four key parts ----
      rm[part * 128 + r] = m;
      rs[part * 128 + r] = l;
      al_bar();
      float M = rm[r];
#pragma unroll
      for (int q = 1; q < kAlParts; ++q) M = fmaxf(M, rm[q * 128 + r]);
      float L = 0.0f;
#pragma unroll
      for (int q = 0; q < kAlParts; ++q) {
        const float mq = rm[q * 128 + r];
        if (mq > -INFINITY) L += rs[q * 128 + r] * al_ex2((mq - M) * kLog2e);
      }
      al_bar();                                            // `red` lies in the P tile: all reads before the first P write
      AL_STAMP(20);
      const float l2 = __log2f(L);                         // p = 2^((s - M) log2e - log2 L)
      // ---- pass 2: probabilities -> P tile ----
      for (int kb = 0; kb < n_kb; ++kb) {
        const int key0 = kb * kAlKeyBlock + part * 64;
        const unsigned long long vis = maskw[kb * 4 + part] | al_suffix_bits64(p, key0, tq);
        mbar_wait(s_full, s_par);
        tc_fence_after();
        float s[64];
        if (key0 + 64 <= p.n_keys) al_scores<true>(trow + part * 64, vis, key0, p.n_keys, s);
        else al_scores<false>(trow + part * 64, vis, key0, p.n_keys, s);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(s_free);
        s_par ^= 1;
        AL_STAMP(21 + kb * 4);
        // the probabilities are formed in registers first: P.V of the previous block may still be reading the P tile
#pragma unroll
        for (int j = 0; j < 64; ++j) s[j] = al_ex2((s[j] - M) * kLog2e - l2);
        mbar_wait(p_free, p_par ^ 1);                      // P.V of the previous block has finished reading the P tile
        AL_STAMP(22 + kb * 4);
        uint8_t* prow = ptile + part * (128 * 128) + r * 128;
#pragma unroll
        for (int g = 0; g < 8; ++g) {
          uint4 o;
          o.x = bf16x2_bits(s[g * 8 + 1], s[g * 8 + 0]);
          o.y = bf16x2_bits(s[g * 8 + 3], s[g * 8 + 2]);
          o.z = bf16x2_bits(s[g * 8 + 5], s[g * 8 + 4]);
          o.w = bf16x2_bits(s[g * 8 + 7], s[g * 8 + 6]);
          *reinterpret_cast<uint4*>(prow + ((g ^ (r & 7)) << 4)) = o;
        }
        fence_proxy_async_smem();   // generic-proxy writes of P -> visible to the tensor core's async-proxy reads
        __syncwarp();
        if (lane == 0) mbar_arrive(p_full);
        p_par ^= 1;
        AL_STAMP(23 + kb * 4);
      }
      // ---- O epilogue: this part's 64 dims ----
      mbar_wait(o_full, ipar);
      tc_fence_after();
      AL_STAMP(40);
      // O -> bf16 -> the (idle) P tile as four [128 rows x 64 dims] 128-byte-swizzled blocks -> one TMA store per part (a
      // lane owns a row: direct 16-byte global stores would touch 32 lines per warp instruction); rows beyond the sample
      // are clipped by the tensor map
      uint8_t* obase = ptile + part * (128 * 128) + r * 128;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float v[32];
        tmem_ld_32x32(trow + 256 + part * 64 + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 o;
          o.x = bf16x2_bits(v[g * 8 + 1], v[g * 8 + 0]);
          o.y = bf16x2_bits(v[g * 8 + 3], v[g * 8 + 2]);
          o.z = bf16x2_bits(v[g * 8 + 5], v[g * 8 + 4]);
          o.w = bf16x2_bits(v[g * 8 + 7], v[g * 8 + 6]);
          *reinterpret_cast<uint4*>(obase + (((c * 4 + g) ^ (r & 7)) << 4)) = o;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(o_free);
      fence_proxy_async_smem();
      al_bar();
      if (threadIdx.x == 64) {
#pragma unroll
        for (int q = 0; q < kAlParts; ++q) tma_store_4d(&tmO, ptile + q * (128 * 128), q * 64, mt * 128, b, 0);
        bulk_commit();
      }
      AL_STAMP(41);
      ipar ^= 1;
    }
    if (threadIdx.x == 64) bulk_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

}  // namespace

bool attention_long_supported(int T, int P, int n_suffix) {
  return T >= 1 && P >= 8 && P % 8 == 0 && n_suffix >= 0 && P + n_suffix <= kAlMaxKeys;
}

// q / att: [B][T * 8 rows][256]; kcache: [B][kv_pitch][256]; vtcache: [B][256][vt_pitch]; pad: [B][P].
// n_suffix == 0: prefix pass (T query tokens against the P prefix keys). n_suffix > 0: denoise step (T = n_suffix suffix tokens
// against P + n_suffix keys, cum = cumsum of the suffix attention mask).
int attention_long(const void* q, const void* kcache, const void* vtcache, void* att, const uint8_t* pad, const int* cum,
                   int B, int T, int P, int n_suffix, int kv_pitch, int vt_pitch, int num_sms, cudaStream_t s) {
  PI0_REQUIRE(q && kcache && vtcache && att && pad && B >= 1, "attention_long: bad argument");
  PI0_REQUIRE(attention_long_supported(T, P, n_suffix), "attention_long: unsupported shape T=%d P=%d suffix=%d", T, P, n_suffix);
  PI0_REQUIRE(n_suffix == 0 || (cum != nullptr && T == n_suffix), "attention_long: suffix mode needs cum and T == n_suffix");
  const int n_keys = P + n_suffix;
  PI0_REQUIRE(vt_pitch % 8 == 0 && vt_pitch >= n_keys && kv_pitch >= n_keys, "attention_long: bad cache pitches");
  static PerDeviceOnce attr_once;
  if (attr_once.need()) {
    PI0_CHECK_CUDA(cudaFuncSetAttribute(attn_long_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kAlSmem));
    attr_once.mark();
  }
  const int R = T * 8;
  CUtensorMap tmQ, tmK, tmV, tmO;
  PI0_TRY(tc_make_map(&tmQ, q, 256, R, B, 1, 256, (int64_t)R * 256, 0, 128, 64, 128));
  PI0_TRY(tc_make_map(&tmK, kcache, 256, n_keys, B, 1, 256, (int64_t)kv_pitch * 256, 0, 256, 64, 128));
  PI0_TRY(tc_make_map(&tmV, vtcache, n_keys, 256, B, 1, vt_pitch, (int64_t)256 * vt_pitch, 0, 256, 64, 128));
  PI0_TRY(tc_make_map(&tmO, att, 256, R, B, 1, 256, (int64_t)R * 256, 0, 128, 64, 128));
  AlParams p;
  p.pad = pad; p.cum = cum; p.att = (bf16*)att;
  p.B = B; p.P = P; p.R = R; p.n_keys = n_keys; p.n_suffix = n_suffix;
  p.tiles_per_b = (R + 127) / 128;
  p.total_items = p.tiles_per_b * B;
  p.n_kb = (n_keys + kAlKeyBlock - 1) / kAlKeyBlock;
  const int grid = p.total_items < num_sms ? p.total_items : num_sms;
  PI0_CHECK_CUDA(launch_pdl(attn_long_kernel, dim3((unsigned)grid), dim3(kAlThreads), kAlSmem, s, 1, tmQ, tmK, tmV, tmO, p));
  ++g_launch_count;
  return 0;
}

}  // namespace pi0

#ifdef PI0_TIMELINE
extern "C" int pi0_debug_attn_long_timeline(unsigned long long* out) {
  return cudaMemcpyFromSymbol(out, pi0::g_al_tl, sizeof(pi0::g_al_tl)) == cudaSuccess ? 0 : -1;
}
#endif
