// Fused prefix-prefill attention on tcgen05 (bidirectional over the valid prefix keys, MQA: the 8 query heads of a token
// are 8 rows against one K / V): scores + masked softmax + P.V in ONE launch per Gemma layer, replacing the
// scores GEMM -> softmax -> P.V GEMM chain and its ~0.35 GB of score traffic per layer at batch 64.
// Reference: eager_attention_forward of the prefix pass ([mirror] /root/reference/src/model/mvla/mllm.py:265-317) with the
// prefix rows of make_att_2d_masks (modeling_mvla.py:98-128): key j is visible iff pad[b][j].
//
// One CTA per SM, persistent over (sample, 128-query-row tile) work items; rows of a sample are r = token * 8 + head of the
// [tokens * 8, 256] view of q. Per tile:
//   S = Q K^T          tcgen05.mma, M = 128, keys padded to 336 = 3 x N=112, head dim in 4 k-blocks of 64 staged by TMA
//                      (Q block 16 KB + K block 42 KB per stage, 2 stages); fp32 S in TMEM columns [0, 336)
//   softmax            12 warps (three per TMEM lane quadrant, a third of the key columns each), thread = row:
//                      s = rnd(rnd(acc) * 256^-0.5), mask (finite -2.3819763e38 like the reference), row max -> exp kept in
//                      TMEM (tcgen05.st) -> row sum -> p = rnd(e / sum) written as bf16 into the 128-byte-swizzled K-major
//                      P tile in shared memory (the A operand of the second MMA)
//   O = P V            tcgen05.mma, M = 128, N = 256 dims, keys in 6 k-blocks of 64 (V^T cache rows staged by TMA into the
//                      same ring; keys >= P are zero-filled by the tensor map and P is zero there); O overwrites S in TMEM
//   epilogue           O -> rnd -> bf16 -> swizzled staging (the dead P tile) -> TMA store, rows beyond the sample clipped
// Rounding points are those of the chain it replaces (scores and probabilities materialised in bf16).
#include "elementwise.cuh"
#include "gemm.cuh"

namespace pi0 {

namespace {

constexpr int kPfParts = 3;                // column parts of the softmax (warps per TMEM lane quadrant)
constexpr int kPfCpp = 12 / kPfParts;      // 32-key chunks per part (12 x 32 = 384 >= 336 padded keys)
constexpr int kPfThreads = 64 + 128 * kPfParts;  // TMA warp, MMA warp, 4 x kPfParts softmax / epilogue warps
constexpr int kPfKeys = 336;               // padded key count (3 MMAs of N = 112)
constexpr int kPfKeyBlocks = 6;            // 64-key blocks of the P.V product (384 >= 336)
constexpr uint32_t kPfQBytes = 128 * 64 * 2;
constexpr uint32_t kPfKBytes = kPfKeys * 64 * 2;
constexpr uint32_t kPfStageBytes = kPfQBytes + kPfKBytes;   // 59392 = 58 KiB (V^T block: 32 KiB, fits)
constexpr uint32_t kPfVBytes = 256 * 64 * 2;
constexpr int kPfStages = 2;
constexpr uint32_t kPfPBytes = kPfKeyBlocks * 128 * 64 * 2; // 96 KiB
constexpr uint32_t kPfSmem = kPfStages * kPfStageBytes + kPfPBytes + 1024 + 2048 + 1024;

struct PfParams {
  const uint8_t* pad;   // [B][P]
  int B, P, R;          // samples, prefix keys, query rows per sample (tokens * 8)
  int tiles_per_b, total_tiles;
  float scale;
};

__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const float* v) {
  const uint32_t* r = reinterpret_cast<const uint32_t*>(v);
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
        "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
        "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
// 32 lanes x 16 consecutive fp32 columns
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// bit j = pad[key0 + j] != 0 for the 32 keys of chunk c (keys >= P read as 0); P % 8 == 0 and 8-byte aligned rows
__device__ __forceinline__ uint32_t chunk_mask(const uint8_t* padb, int c, int P) {
  uint32_t m = 0;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int key0 = c * 32 + q * 8;
    if (key0 < P) {
      const uint2 w = __ldg(reinterpret_cast<const uint2*>(padb + key0));
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint32_t wd = j < 4 ? w.x : w.y;
        if ((wd >> (8 * (j & 3))) & 0xffu) m |= 1u << (q * 8 + j);
      }
    }
  }
  return m;
}
__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, %0;" ::"n"(128 * kPfParts) : "memory"); }

#ifdef PI0_TIMELINE
__device__ unsigned long long g_pf_tl[16];
#define PF_STAMP(i) do { if (blockIdx.x == 0 && threadIdx.x == 64 && idx == (int)blockIdx.x) g_pf_tl[i] = global_timer_ns(); } while (0)
#else
#define PF_STAMP(i) do { } while (0)
#endif

__global__ void __launch_bounds__(kPfThreads, 1)
prefill_attn_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                    const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmO, const PfParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* ring = smem;                                          // kPfStages x (Q block | K block)  /  V^T blocks
  uint8_t* ptile = smem + kPfStages * kPfStageBytes;             // 6 x [128 rows x 64 keys], SW128 K-major
  float* red = reinterpret_cast<float*>(ptile + kPfPBytes);      // [kPfParts][128 rows] row max, then row sum
  uint32_t* maskw = reinterpret_cast<uint32_t*>(ptile + kPfPBytes + 1536);   // [12] visibility bits per 32-key chunk
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(ptile + kPfPBytes + 2048);
  uint64_t* empty_bar = full_bar + kPfStages;
  uint64_t* s_full = empty_bar + kPfStages;    // S accumulated            (MMA -> softmax warps)
  uint64_t* p_full = s_full + 1;               // P tile written, S dead   (softmax warps -> MMA)
  uint64_t* o_full = p_full + 1;               // O accumulated            (MMA -> epilogue)
  uint64_t* o_empty = o_full + 1;              // O drained from TMEM      (epilogue -> MMA of the next tile)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_empty + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_trigger();
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
    tma_prefetch_desc(&tmO);
    for (int s = 0; s < kPfStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(p_full, 4 * kPfParts);
    mbar_init(o_full, 1);
    mbar_init(o_empty, 4 * kPfParts);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int idx = blockIdx.x; idx < p.total_tiles; idx += gridDim.x) {
        const int b = idx / p.tiles_per_b, mt = idx - b * p.tiles_per_b;
        for (int kb = 0; kb < 4; ++kb) {          // Q and K, head-dim block kb
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sq = ring + stage * kPfStageBytes;
          mbar_arrive_expect_tx(&full_bar[stage], kPfStageBytes);
          tma_load_4d(sq, &tmQ, &full_bar[stage], kb * 64, mt * 128, b, 0);
#pragma unroll
          for (int j = 0; j < 3; ++j)
            tma_load_4d(sq + kPfQBytes + j * (112 * 128), &tmK, &full_bar[stage], kb * 64, j * 112, b, 0);
          if (++stage == kPfStages) { stage = 0; phase ^= 1; }
        }
        for (int vb = 0; vb < kPfKeyBlocks; ++vb) {  // V^T, key block vb
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full_bar[stage], kPfVBytes);
          tma_load_4d(ring + stage * kPfStageBytes, &tmV, &full_bar[stage], vb * 64, 0, b, 0);
          if (++stage == kPfStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer --------------------------------
    constexpr uint32_t idesc_s = umma_idesc_bf16(128, 112);
    constexpr uint32_t idesc_o = umma_idesc_bf16(128, 256);
    int stage = 0;
    uint32_t phase = 0, tile_par = 0;
    for (int idx = blockIdx.x; idx < p.total_tiles; idx += gridDim.x) {
      mbar_wait(o_empty, tile_par ^ 1);          // previous tile's O (aliasing S) has been read out
      tc_fence_after();
      for (int kb = 0; kb < 4; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sq = smem_u32(ring + stage * kPfStageBytes);
          const uint64_t da = umma_desc_k_sw128(sq);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
#pragma unroll
            for (int j = 0; j < 3; ++j) {
              const uint64_t db = umma_desc_k_sw128(sq + kPfQBytes + j * (112 * 128));
              umma_bf16(tmem_base + j * 112, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc_s,
                        (kb | k) != 0 ? 1u : 0u);
            }
          }
          umma_commit(&empty_bar[stage]);
          if (kb == 3) umma_commit(s_full);
        }
        __syncwarp();
        if (++stage == kPfStages) { stage = 0; phase ^= 1; }
      }
      mbar_wait(p_full, tile_par);               // P in shared memory (async-proxy visible), S consumed
      tc_fence_after();
      for (int vb = 0; vb < kPfKeyBlocks; ++vb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t da = umma_desc_k_sw128(smem_u32(ptile + vb * (128 * 128)));
          const uint64_t db = umma_desc_k_sw128(smem_u32(ring + stage * kPfStageBytes));
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_bf16(tmem_base, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc_o, (vb | k) != 0 ? 1u : 0u);
          umma_commit(&empty_bar[stage]);
          if (vb == kPfKeyBlocks - 1) umma_commit(o_full);
        }
        __syncwarp();
        if (++stage == kPfStages) { stage = 0; phase ^= 1; }
      }
      tile_par ^= 1;
    }
  } else {
    // ------------------------------ softmax + epilogue ------------------------
    const int quad = warp & 3, part = (warp - 2) >> 2;
    const int r = quad * 32 + lane;                       // row of the tile = TMEM lane
    const uint32_t trow = tmem_base + ((uint32_t)(quad * 32) << 16);
    const float kNeg = -2.3819763e38f;
    uint8_t* stage2 = ptile + (warp - 2) * 2 * 2048;      // output staging (P tile is dead by then): 4 KiB per warp
    uint32_t tile_par = 0;
    for (int idx = blockIdx.x; idx < p.total_tiles; idx += gridDim.x) {
      const int b = idx / p.tiles_per_b, mt = idx - b * p.tiles_per_b;
      const uint8_t* padb = p.pad + (int64_t)b * p.P;
      PF_STAMP(0);
      // key-visibility bits of the sample, one word per 32-key chunk, fetched while the S MMAs are still running
      // (the previous tile's readers are past their last use: they all arrived on o_empty's predecessor barriers)
      if (threadIdx.x - 64 < 12) maskw[threadIdx.x - 64] = chunk_mask(padb, (int)threadIdx.x - 64, p.P);
      epi_bar();
      mbar_wait(s_full, tile_par);
      tc_fence_after();
      PF_STAMP(1);
      // ---- pass 1 (the only TMEM pass): masked scaled scores of this part's 4 chunks, kept as packed bf16 (they ARE
      //      bf16 values: the reference materialises the scores in bf16), and the row max ----
      uint32_t sp[kPfCpp][16];
      float mx = -INFINITY;
#pragma unroll
      for (int ci = 0; ci < kPfCpp; ++ci) {
        const int c = part * kPfCpp + ci;                // chunk 11 lies beyond the padded keys: all -inf
        const uint32_t msk = maskw[c];
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          float v[16];
          if (c < 11) {
            tmem_ld_32x16(trow + c * 32 + hh * 16, v);
            tmem_ld_wait();
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = 0.0f;
          }
#pragma unroll
          for (int j = 0; j < 16; j += 2) {
            const int jj = hh * 16 + j, key = c * 32 + jj;
            // scale = 2^-4: rnd(acc) * scale is exact in bf16, the reference's second rounding changes nothing
            float s0 = bf16_round(v[j]) * p.scale, s1 = bf16_round(v[j + 1]) * p.scale;
            s0 = key < p.P ? (((msk >> jj) & 1u) ? s0 : kNeg) : -INFINITY;
            s1 = key + 1 < p.P ? (((msk >> (jj + 1)) & 1u) ? s1 : kNeg) : -INFINITY;
            const uint32_t w = bf16x2_bits(s1, s0);
            sp[ci][jj >> 1] = w;
            // the max is taken over what later passes read back (kNeg is not a bf16 value: it rounds here, for all keys alike)
            mx = fmaxf(mx, fmaxf(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u)));
          }
        }
      }
      PF_STAMP(2);
      red[part * 128 + r] = mx;
      epi_bar();
      mx = red[r];
#pragma unroll
      for (int q = 1; q < kPfParts; ++q) mx = fmaxf(mx, red[q * 128 + r]);
      epi_bar();
      // ---- pass 2: row sum of exp(s - max) (ex2.approx: relative error 2^-22, far below the bf16 rounding of p) ----
      const float kLog2e = 1.4426950408889634f;
      float sum = 0.0f;
#pragma unroll
      for (int ci = 0; ci < kPfCpp; ++ci) {
#pragma unroll
        for (int jj = 0; jj < 16; ++jj) {
          const uint32_t w = sp[ci][jj];
          sum += exp2f((__uint_as_float(w << 16) - mx) * kLog2e) + exp2f((__uint_as_float(w & 0xffff0000u) - mx) * kLog2e);
        }
      }
      PF_STAMP(3);
      red[part * 128 + r] = sum;
      epi_bar();
      sum = red[r];
#pragma unroll
      for (int q = 1; q < kPfParts; ++q) sum += red[q * 128 + r];
      const float lsum = __log2f(sum);
      // ---- pass 3: p = rnd(exp(s - max) / sum) = rnd(2^((s - max) log2e - log2 sum)) -> swizzled K-major P tile ----
#pragma unroll
      for (int ci = 0; ci < kPfCpp; ++ci) {
        const int c = part * kPfCpp + ci;
        uint8_t* blk = ptile + (c >> 1) * (128 * 128) + r * 128;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint32_t o[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const uint32_t w = sp[ci][g * 4 + q];
            const float p0 = exp2f((__uint_as_float(w << 16) - mx) * kLog2e - lsum);
            const float p1 = exp2f((__uint_as_float(w & 0xffff0000u) - mx) * kLog2e - lsum);
            o[q] = bf16x2_bits(p1, p0);
          }
          const int piece = (c & 1) * 4 + g;
          *reinterpret_cast<uint4*>(blk + ((piece ^ (r & 7)) << 4)) = make_uint4(o[0], o[1], o[2], o[3]);
        }
      }
      fence_proxy_async_smem();   // generic-proxy writes of P -> visible to the tensor core's async-proxy reads
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
      PF_STAMP(4);
      // ---- O epilogue: 4 chunks of 32 dims per warp ----
      mbar_wait(o_full, tile_par);
      tc_fence_after();
      PF_STAMP(5);
      int sbuf = 0;
#pragma unroll 1
      constexpr int kOPer = (8 + kPfParts - 1) / kPfParts;   // 8 chunks of 32 dims: 3 + 3 + 2
      for (int c = part * kOPer; c < min(8, (part + 1) * kOPer); ++c) {
        float v[32];
        tmem_ld_32x32(trow + c * 32, v);
        tmem_ld_wait();
        uint8_t* buf = stage2 + sbuf * 2048;
        sbuf ^= 1;
        if (lane == 0) bulk_wait_read<1>();
        __syncwarp();
        uint8_t* mine = buf + lane * 64;
        const int sw = (lane >> 1) & 3;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 pk;
          pk.x = bf16x2_bits(v[g * 8 + 1], v[g * 8 + 0]);
          pk.y = bf16x2_bits(v[g * 8 + 3], v[g * 8 + 2]);
          pk.z = bf16x2_bits(v[g * 8 + 5], v[g * 8 + 4]);
          pk.w = bf16x2_bits(v[g * 8 + 7], v[g * 8 + 6]);
          *reinterpret_cast<uint4*>(mine + ((g ^ sw) << 4)) = pk;
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          tma_store_4d(&tmO, buf, c * 32, mt * 128 + quad * 32, b, 0);
          bulk_commit();
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(o_empty);
        bulk_wait_read<0>();   // the staging buffers live in the P tile: drained before the next tile's P is written
      }
      __syncwarp();
      PF_STAMP(6);
      tile_par ^= 1;
    }
    if (lane == 0) bulk_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

}  // namespace

bool prefill_attention_supported(int T, int P) { return P >= 16 && P <= kPfKeys && P % 8 == 0 && T >= 1; }

// q / att: [B * T rows][8 * 256]; kcache: [B][kv_pitch][256]; vtcache: [B][256][vt_pitch]; pad: [B][P]
int prefill_attention(const void* q, const void* kcache, const void* vtcache, void* att, const uint8_t* pad, int B, int T,
                      int P, int kv_pitch, int vt_pitch, int num_sms, cudaStream_t s) {
  PI0_REQUIRE(q && kcache && vtcache && att && pad && B >= 1, "prefill_attention: bad argument");
  PI0_REQUIRE(prefill_attention_supported(T, P), "prefill_attention: unsupported shape T=%d P=%d", T, P);
  static PerDeviceOnce attr_once;
  if (attr_once.need()) {
    PI0_CHECK_CUDA(cudaFuncSetAttribute(prefill_attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPfSmem));
    attr_once.mark();
  }
  const int R = T * 8;
  CUtensorMap tmQ, tmK, tmV, tmO;
  PI0_TRY(tc_make_map(&tmQ, q, 256, R, B, 1, 256, (int64_t)R * 256, 0, 128, 64, 128));
  PI0_TRY(tc_make_map(&tmK, kcache, 256, P, B, 1, 256, (int64_t)kv_pitch * 256, 0, 112, 64, 128));
  PI0_TRY(tc_make_map(&tmV, vtcache, P, 256, B, 1, vt_pitch, (int64_t)256 * vt_pitch, 0, 256, 64, 128));
  PI0_TRY(tc_make_map(&tmO, att, 256, R, B, 1, 256, (int64_t)R * 256, 0, 32, 32, 64));
  PfParams p;
  p.pad = pad; p.B = B; p.P = P; p.R = R;
  p.tiles_per_b = (R + 127) / 128;
  p.total_tiles = p.tiles_per_b * B;
  p.scale = 0.0625f;
  const int grid = p.total_tiles < num_sms ? p.total_tiles : num_sms;
  PI0_CHECK_CUDA(launch_pdl(prefill_attn_kernel, dim3((unsigned)grid), dim3(kPfThreads), kPfSmem, s, 1, tmQ, tmK, tmV, tmO,
                            p));
  ++g_launch_count;
  return 0;
}

}  // namespace pi0

#ifdef PI0_TIMELINE
extern "C" int pi0_debug_prefill_timeline(unsigned long long* out) {
  return cudaMemcpyFromSymbol(out, pi0::g_pf_tl, sizeof(pi0::g_pf_tl)) == cudaSuccess ? 0 : -1;
}
#endif
