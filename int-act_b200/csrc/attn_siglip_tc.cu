// Fused SigLIP self-attention on tcgen05 (bidirectional, 256 tokens, 16 heads x 72 dims): scores + softmax + P.V of one
// encoder layer in ONE persistent launch, the tensor-core form of attn_siglip.cu (mma.sync, 88 TFLOP/s at batch 64).
// Reference: HF SiglipAttention as called by get_image_features ([mirror] /root/reference/src/model/mvla/mllm.py:182-187;
// SURVEY.md 9.2): scale 72^-0.5, no mask.
//
// One CTA per SM, persistent over (image, head) items; an item is two 128-query-row halves that share K and V:
//   operands   Q, K, V tiles of the head come straight out of the fused q|k|v activation [tokens, 3456] by TMA: dims 0..63 as
//              a [256 tokens x 128 B] 128-byte-swizzled tile, dims 64..71 as a [256 tokens x 16 B] unswizzled tile followed
//              by a zero block (dims 72..79) that is written once per CTA. The head dim 72 is therefore covered by four
//              k-steps on the swizzled tile and a fifth on the {tail, zero} pair (K-major, no swizzle: k-blocks 4 KiB apart).
//   S = Q K^T  tcgen05.mma M = 128, N = 256 keys, 5 k-steps; fp32 S of half h in TMEM columns [256 h, 256 h + 256)
//   softmax    16 warps (four per TMEM lane quadrant, 64 keys each), thread = row: s = rnd(rnd(acc) * 72^-0.5) (the
//              reference materialises the scores in bf16), e = exp(s - local max), ONE exchange of (local max, local sum)
//              between the four key parts through shared memory, p = rnd(e * f / sum) as bf16 into the 128-byte-swizzled
//              K-major P tile (A operand of the second product)
//   O = P V    V is consumed as it lies in the activation (row-major [key][dim] = MN-major B operand): N = 64 dims from the
//              swizzled tile + N = 16 from {tail, zero}, 16 k-steps of 16 keys; O of half h overwrites S_h in TMEM
//   epilogue   O -> rnd -> bf16 -> [128 rows x 144 B] staging tile (conflict-free 16-byte stores) -> one TMA store per half
//              into att [tokens, 1152]
// The S product of half 1 runs under the softmax of half 0, P.V of half 0 under the softmax of half 1, the next item's Q | K
// load as soon as both S products have been issued.
#include "elementwise.cuh"
#include "gemm.cuh"

namespace pi0 {

namespace {

constexpr int kStT = 256;                 // tokens per image
constexpr int kStHd = 72;                 // head dim
constexpr int kStHeads = 16;
constexpr int kStHid = kStHeads * kStHd;  // 1152
constexpr int kStLd = 3 * kStHid;         // row pitch of the fused q|k|v activation
constexpr int kStParts = 4;               // softmax warps per TMEM lane quadrant (64 keys each)
constexpr int kStSmThreads = 128 * kStParts;
constexpr int kStThreads = 64 + kStSmThreads;       // TMA warp, MMA warp, 16 softmax / epilogue warps
constexpr uint32_t kStMainBytes = kStT * 128;       // [256 tokens][64 dims], SW128
constexpr uint32_t kStTailBytes = kStT * 16;        // [256 tokens][8 dims] (dims 64..71); the zero block has the same size
constexpr uint32_t kStOperandBytes = kStMainBytes + 2 * kStTailBytes;   // 40 KiB
constexpr uint32_t kStPBytes = 4 * 128 * 128;       // four 64-key blocks of [128 rows x 128 B]
constexpr uint32_t kStRedBytes = 2 * 2 * kStParts * 128 * 4;   // [half][max | sum][part][row]
constexpr uint32_t kStOutBytes = 128 * kStHd * 2;     // one half's O as bf16 [128 rows][144 B], source of a TMA store
constexpr uint32_t kStSmem = 3 * kStOperandBytes + kStPBytes + kStOutBytes + kStRedBytes + 256 + 1024;
static_assert(kStSmem <= 232448, "shared memory budget of one SM");

// shared-memory matrix descriptor without swizzle: 8 x 16-byte core matrices; `lbo` / `sbo` in bytes.
// K-major operand:  core matrix = 8 rows x 8 k, rows 16 B apart; sbo = stride between 8-row groups, lbo = between the two
//                   8-element k-blocks of one MMA.
// MN-major operand: core matrix = 8 k x 8 n (8 n contiguous), k 16 B apart; sbo = stride between 8-element n-blocks,
//                   lbo = between 8-element k-blocks.
__device__ __forceinline__ uint64_t umma_desc_noswz(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)(lbo >> 4) << 16;
  d |= (uint64_t)(sbo >> 4) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
// MN-major operand in a 128-byte-swizzled tile ([k rows][64 n] = 128 B per k): 8-k groups 1024 B apart (sbo); a single
// 64-element n group, so lbo is unused
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__device__ __forceinline__ void tmem_ld_32x8(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
}
__device__ __forceinline__ void tmem_ld_32x16s(uint32_t taddr, float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
// tcgen05.wait::ld that also "produces" the 16 loaded registers, so that no use of them can be scheduled above the wait when
// other work sits between the load and the wait
__device__ __forceinline__ void tmem_ld_wait16(float* v) {
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]),
                 "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
               :
               : "memory");
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void sm_bar() { asm volatile("bar.sync 1, %0;" ::"n"(kStSmThreads) : "memory"); }

#ifdef PI0_TIMELINE
__device__ unsigned long long g_st_tl[32];
#define ST_STAMP(i) do { if (blockIdx.x == 0 && threadIdx.x == 64 && idx == (int)(blockIdx.x + 2 * gridDim.x)) g_st_tl[i] = global_timer_ns(); } while (0)
#else
#define ST_STAMP(i) do { } while (0)
#endif

__global__ void __launch_bounds__(kStThreads, 1)
siglip_attn_tc_kernel(const __grid_constant__ CUtensorMap tmMain, const __grid_constant__ CUtensorMap tmTail,
                      const __grid_constant__ CUtensorMap tmOut, int total_items, float scale) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* qs = smem;                                  // main | tail | zero
  uint8_t* ks = qs + kStOperandBytes;
  uint8_t* vs = ks + kStOperandBytes;
  uint8_t* ptile = vs + kStOperandBytes;
  // O staging [128 rows][144 B]: half 0 has its own buffer (P.V of half 1 may still be reading the P tile), half 1 uses the
  // P tile, which is dead once its O is complete and is not rewritten before the next item's exchange barrier
  uint8_t* ostage0 = ptile + kStPBytes;
  float* red = reinterpret_cast<float*>(ostage0 + kStOutBytes);
  uint64_t* qk_full = reinterpret_cast<uint64_t*>(ostage0 + kStOutBytes + kStRedBytes);
  uint64_t* v_full = qk_full + 1;
  uint64_t* qk_empty = v_full + 1;      // both S products issued and complete -> Q | K may be overwritten
  uint64_t* v_empty = qk_empty + 1;     // both P.V products complete          -> V may be overwritten
  uint64_t* s_full = v_empty + 1;       // [2] S of half h accumulated
  uint64_t* p_full = s_full + 2;        // [2] P of half h in shared memory, S_h consumed
  uint64_t* o_full = p_full + 2;        // [2] O of half h accumulated (and the P tile is free again)
  uint64_t* s_free = o_full + 2;        // [2] O_h drained from TMEM -> S_h of the next item may be accumulated
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(s_free + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_trigger();
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmMain);
    tma_prefetch_desc(&tmTail);
    tma_prefetch_desc(&tmOut);
    mbar_init(qk_full, 1);
    mbar_init(v_full, 1);
    mbar_init(qk_empty, 1);
    mbar_init(v_empty, 1);
    for (int h = 0; h < 2; ++h) {
      mbar_init(&s_full[h], 1);
      mbar_init(&p_full[h], 4 * kStParts);
      mbar_init(&o_full[h], 1);
      mbar_init(&s_free[h], 4 * kStParts);
    }
    fence_barrier_init();
  }
  // the zero blocks (dims 72..79 of Q, K and V) are never touched by the TMA loads
  for (int i = threadIdx.x; i < 3 * (int)(kStTailBytes / 16); i += kStThreads) {
    const int op = i / (int)(kStTailBytes / 16), j = i - op * (int)(kStTailBytes / 16);
    *reinterpret_cast<uint4*>(smem + op * kStOperandBytes + kStMainBytes + kStTailBytes + j * 16) = make_uint4(0, 0, 0, 0);
  }
  fence_proxy_async_smem();
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (elect_one()) {
      uint32_t par = 0;
      for (int idx = blockIdx.x; idx < total_items; idx += gridDim.x) {
        const int b = idx / kStHeads, h = idx - b * kStHeads;
        const int cq = h * kStHd;
        mbar_wait(qk_empty, par ^ 1);
        mbar_arrive_expect_tx(qk_full, 2 * (kStMainBytes + kStTailBytes));
        tma_load_4d(qs, &tmMain, qk_full, cq, 0, b, 0);
        tma_load_4d(qs + kStMainBytes, &tmTail, qk_full, cq + 64, 0, b, 0);
        tma_load_4d(ks, &tmMain, qk_full, kStHid + cq, 0, b, 0);
        tma_load_4d(ks + kStMainBytes, &tmTail, qk_full, kStHid + cq + 64, 0, b, 0);
        mbar_wait(v_empty, par ^ 1);
        mbar_arrive_expect_tx(v_full, kStMainBytes + kStTailBytes);
        tma_load_4d(vs, &tmMain, v_full, 2 * kStHid + cq, 0, b, 0);
        tma_load_4d(vs + kStMainBytes, &tmTail, v_full, 2 * kStHid + cq + 64, 0, b, 0);
        par ^= 1;
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer --------------------------------
    constexpr uint32_t idesc_s = umma_idesc_bf16(128, 256);
    constexpr uint32_t idesc_o64 = umma_idesc_bf16(128, 64) | (1u << 16);   // B operand MN-major
    constexpr uint32_t idesc_o16 = umma_idesc_bf16(128, 16) | (1u << 16);
    uint32_t par = 0;
    for (int idx = blockIdx.x; idx < total_items; idx += gridDim.x) {
      mbar_wait(qk_full, par);
      tc_fence_after();
      for (int half = 0; half < 2; ++half) {
        mbar_wait(&s_free[half], par ^ 1);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t d = tmem_base + half * 256;
          const uint64_t da = umma_desc_k_sw128(smem_u32(qs) + half * (128 * 128));
          const uint64_t db = umma_desc_k_sw128(smem_u32(ks));
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16(d, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc_s, k != 0 ? 1u : 0u);
          const uint64_t ta = umma_desc_noswz(smem_u32(qs) + kStMainBytes + half * (128 * 16), kStTailBytes, 128);
          const uint64_t tb = umma_desc_noswz(smem_u32(ks) + kStMainBytes, kStTailBytes, 128);
          umma_bf16(d, ta, tb, idesc_s, 1u);
          umma_commit(&s_full[half]);
          if (half == 1) umma_commit(qk_empty);
        }
        __syncwarp();
      }
      mbar_wait(v_full, par);
      tc_fence_after();
      for (int half = 0; half < 2; ++half) {
        mbar_wait(&p_full[half], par);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t d = tmem_base + half * 256;
#pragma unroll
          for (int kk = 0; kk < 16; ++kk) {   // 16 keys per step
            const uint64_t da = umma_desc_k_sw128(smem_u32(ptile) + (kk >> 2) * (128 * 128)) + (uint64_t)((kk & 3) * 2);
            const uint64_t dm = umma_desc_mn_sw128(smem_u32(vs) + kk * 2048);
            const uint64_t dt = umma_desc_noswz(smem_u32(vs) + kStMainBytes + kk * 256, 128, kStTailBytes);
            umma_bf16(d, da, dm, idesc_o64, kk != 0 ? 1u : 0u);
            umma_bf16(d + 64, da, dt, idesc_o16, kk != 0 ? 1u : 0u);
          }
          umma_commit(&o_full[half]);
          if (half == 1) umma_commit(v_empty);
        }
        __syncwarp();
      }
      par ^= 1;
    }
  } else {
    // ------------------------------ softmax + epilogue ------------------------
    const int quad = warp & 3, part = (warp - 2) >> 2;
    const int r = quad * 32 + lane;                       // row of the half = TMEM lane
    const uint32_t trow = tmem_base + ((uint32_t)(quad * 32) << 16);
    const float kLog2e = 1.4426950408889634f;
    uint32_t par = 0;
    for (int idx = blockIdx.x; idx < total_items; idx += gridDim.x) {
      const int b = idx / kStHeads, h = idx - b * kStHeads;
#pragma unroll 1
      for (int half = 0; half < 2; ++half) {
        ST_STAMP(half * 6 + 0);
        mbar_wait(&s_full[half], par);
        tc_fence_after();
        ST_STAMP(half * 6 + 1);
        // ---- scores of this thread's 64 keys, s = rnd(rnd(acc) * scale) (bf16 values, as the reference materialises them),
        //      kept as fp32 bit patterns of the rounded values, and their max ----
        float e[64];
        float mx = -INFINITY;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          float v[32];
          tmem_ld_32x16s(trow + half * 256 + part * 64 + c * 32, v);
          tmem_ld_32x16s(trow + half * 256 + part * 64 + c * 32 + 16, v + 16);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            e[c * 32 + j] = bf16_round(bf16_round(v[j]) * scale);
            mx = fmaxf(mx, e[c * 32 + j]);
          }
        }
        ST_STAMP(half * 6 + 2);
        // ---- e = exp(s - local max), local sum; one exchange of (max, sum) between the four key parts ----
        float sum = 0.0f;
        const float mxl = mx * kLog2e;
#pragma unroll
        for (int j = 0; j < 64; ++j) {
          e[j] = ex2_approx(fmaf(e[j], kLog2e, -mxl));
          sum += e[j];
        }
        ST_STAMP(half * 6 + 3);
        float* rm = red + half * (2 * kStParts * 128);
        float* rs = rm + kStParts * 128;
        rm[part * 128 + r] = mx;
        rs[part * 128 + r] = sum;
        if (half == 0 && threadIdx.x == 64) bulk_wait_read<0>();   // the previous item's O stores have left the staging
        sm_bar();
        float gmx = rm[r];
#pragma unroll
        for (int q = 1; q < kStParts; ++q) gmx = fmaxf(gmx, rm[q * 128 + r]);
        float tot = 0.0f;
#pragma unroll
        for (int q = 0; q < kStParts; ++q) tot += rs[q * 128 + r] * ex2_approx((rm[q * 128 + r] - gmx) * kLog2e);
        const float f = ex2_approx((mx - gmx) * kLog2e) / tot;
        // the P tile is shared by the two halves: P.V of half 0 must have finished reading it
        if (half == 1) mbar_wait(&o_full[0], par);
        ST_STAMP(half * 6 + 4);
        uint8_t* prow = ptile + part * (128 * 128) + r * 128;
#pragma unroll
        for (int g = 0; g < 8; ++g) {
          uint4 o;
          o.x = bf16x2_bits(e[g * 8 + 1] * f, e[g * 8 + 0] * f);
          o.y = bf16x2_bits(e[g * 8 + 3] * f, e[g * 8 + 2] * f);
          o.z = bf16x2_bits(e[g * 8 + 5] * f, e[g * 8 + 4] * f);
          o.w = bf16x2_bits(e[g * 8 + 7] * f, e[g * 8 + 6] * f);
          *reinterpret_cast<uint4*>(prow + ((g ^ (r & 7)) << 4)) = o;
        }
        fence_proxy_async_smem();   // generic-proxy writes of P -> visible to the tensor core's async-proxy reads
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_full[half]);
        ST_STAMP(half * 6 + 5);
      }
      // ---- O epilogue: nine 8-dim blocks per row, dealt 3 / 2 / 2 / 2 to the four parts ----
#pragma unroll 1
      for (int half = 0; half < 2; ++half) {
        mbar_wait(&o_full[half], par);
        tc_fence_after();
        ST_STAMP(12 + half * 2);
        const int blk0 = part == 0 ? 0 : 1 + 2 * part, blk1 = 3 + 2 * part;
        uint8_t* ostage = half == 0 ? ostage0 : ptile;
        uint8_t* orow = ostage + r * (kStHd * 2);   // 144-byte rows: conflict-free 16-byte stores
        for (int blk = blk0; blk < blk1; ++blk) {
          float v[8];
          tmem_ld_32x8(trow + half * 256 + blk * 8, v);
          tmem_ld_wait();
          uint4 o;
          o.x = bf16x2_bits(v[1], v[0]);
          o.y = bf16x2_bits(v[3], v[2]);
          o.z = bf16x2_bits(v[5], v[4]);
          o.w = bf16x2_bits(v[7], v[6]);
          *reinterpret_cast<uint4*>(orow + blk * 16) = o;
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&s_free[half]);
        fence_proxy_async_smem();
        sm_bar();
        if (threadIdx.x == 64) {
          tma_store_4d(&tmOut, ostage, h * kStHd, half * 128, b, 0);
          bulk_commit();
        }
        ST_STAMP(13 + half * 2);
      }
      par ^= 1;
    }
    if (threadIdx.x == 64) bulk_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

}  // namespace

// qkv [B*256, 3456]: q(b, t, h, d) at (b*256 + t) * 3456 + h*72 + d, k at + 1152, v at + 2304; att [B*256, 1152]
int siglip_attention_tc(const void* qkv, void* att, int B, int num_sms, cudaStream_t s) {
  PI0_REQUIRE(qkv && att && B >= 1 && num_sms >= 1, "siglip_attention_tc: bad argument");
  static PerDeviceOnce attr_once;
  if (attr_once.need()) {
    PI0_CHECK_CUDA(cudaFuncSetAttribute(siglip_attn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kStSmem));
    attr_once.mark();
  }
  CUtensorMap tmMain, tmTail, tmOut;
  PI0_TRY(tc_make_map(&tmMain, qkv, kStLd, kStT, B, 1, kStLd, (int64_t)kStT * kStLd, 0, kStT, 64, 128));
  PI0_TRY(tc_make_map(&tmTail, qkv, kStLd, kStT, B, 1, kStLd, (int64_t)kStT * kStLd, 0, kStT, 8, 0));
  PI0_TRY(tc_make_map(&tmOut, att, kStHid, kStT, B, 1, kStHid, (int64_t)kStT * kStHid, 0, 128, kStHd, 0));
  const int total = B * kStHeads;
  const int grid = total < num_sms ? total : num_sms;
  PI0_CHECK_CUDA(launch_pdl(siglip_attn_tc_kernel, dim3((unsigned)grid), dim3(kStThreads), kStSmem, s, 1, tmMain, tmTail,
                            tmOut, total, 0.11785113019775793f));
  ++g_launch_count;
  return 0;
}

}  // namespace pi0

#ifdef PI0_TIMELINE
extern "C" int pi0_debug_siglip_tc_timeline(unsigned long long* out) {
  return cudaMemcpyFromSymbol(out, pi0::g_st_tl, sizeof(pi0::g_st_tl)) == cudaSuccess ? 0 : -1;
}
#endif
