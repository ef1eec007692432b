// Device body of the rollout-batch denoise attention (one (sample, head group) unit: K and V^T of the sample staged in
// shared memory, see attn_batch.cu for the design), shared by the stand-alone batch_attn_kernel and the fused denoise-step
// kernel (fused_denoise.cu). Executed by 256 threads that synchronise through `SYNC()`; the two mbarriers are reused from
// unit to unit, `par` carries their phase parity.
#pragma once
#include "attn_decode.cuh"

namespace pi0 {
namespace battn {

constexpr int kRows = 48;        // 8 heads x S <= 6 suffix rows, three m16 tiles
constexpr int kKeysMax = 352;    // P + S padded to 32
constexpr int kRowB = 512;       // bytes of one K / Q row (256 bf16), stored unpadded with a parity swizzle
constexpr int kSPitch = kKeysMax * 2;  // bytes of one score / probability row (704 = 64 mod 128)

struct BatchAttnParams {
  const bf16* q;        // [B*S, 8*256]
  const bf16* kc;       // [B, kv_pitch, 256]
  const bf16* vt;       // [B, 256, vt_pitch]
  bf16* att;            // [B*S, 8*256]
  const uint8_t* pad;   // [B, P]
  const int* cum;       // [S]
  int S, P, kv_pitch, vt_pitch, kp;
  int hpg;              // heads per CTA (8 or 4): grid = (8 / hpg, B); row r of a CTA = token r / hpg, head h0 + r % hpg
};

__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// byte offset of 16-byte chunk `ci` (0..31) of row `row` in a 512-byte-row tile. The tiles are filled by 1-D bulk TMA
// copies, which cannot swizzle: fragment loads of a quarter warp (2 rows x 4 chunks) are 2-way bank-conflicted, which
// costs far less than issuing the tile as ~11k per-thread cp.async (measured 40-100 ns per warp instruction).
__device__ __forceinline__ int swz(int row, int ci) { return row * kRowB + (ci << 4); }
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}


// `bars`: two mbarriers initialised (count 1) by the caller; `par`: their current phase parity (flipped here)
template <typename SyncFn>
__device__ __forceinline__ void batch_attn_unit(uint8_t* smem, uint64_t* bars, uint32_t& par, const BatchAttnParams& p, int b,
                                                int h0, SyncFn&& SYNC) {
  const int hpg = p.hpg;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t4 = lane & 3;
  const int n_keys = p.P + p.S, kp = p.kp, R = p.S * hpg;
  const int kr = (n_keys + 15) & ~15;                   // K rows staged: the keys padded to the 16-key score blocks
  uint8_t* ks = smem;                                   // [kr][512] swizzled (rows >= n_keys zero)
  uint8_t* qs = smem + (size_t)kr * kRowB;              // [48][512] swizzled (rows >= R zero)
  uint8_t* ss = qs + kRows * kRowB;                     // [48][704] scores, then probabilities (bf16)
  uint8_t* vs = smem;                                   // [256][vpitch] V^T, overlays K and Q after the scores
  const int vpitch = p.vt_pitch * 2;                    // bytes: V^T is staged exactly as it lies in the cache

  // ---- K and Q of this sample: bulk TMA copies (K rows are contiguous in the cache; Q rows of a token are) ---------
  if (threadIdx.x == 0) {
    const bf16* kb = p.kc + (int64_t)b * p.kv_pitch * 256;
    const bf16* qb = p.q + (int64_t)b * p.S * 2048;     // (token t, head h) = 256 consecutive elements at (t*8 + h) * 256
    mbar_arrive_expect_tx(&bars[0], (uint32_t)(n_keys * kRowB + p.S * hpg * kRowB));
    bulk_g2s(ks, kb, (uint32_t)(n_keys * kRowB), &bars[0]);
    for (int t = 0; t < p.S; ++t)
      bulk_g2s(qs + t * hpg * kRowB, qb + (int64_t)(t * 8 + h0) * 256, (uint32_t)(hpg * kRowB), &bars[0]);
  }
  // zero the padding rows (keys n_keys..kr-1, query rows R..47) with ordinary stores
  for (int i = threadIdx.x; i < (kr - n_keys) * 32; i += 256)
    *reinterpret_cast<uint4*>(ks + swz(n_keys + (i >> 5), i & 31)) = make_uint4(0, 0, 0, 0);
  for (int i = threadIdx.x; i < (kRows - R) * 32; i += 256)
    *reinterpret_cast<uint4*>(qs + swz(R + (i >> 5), i & 31)) = make_uint4(0, 0, 0, 0);
  // which of this lane's keys are attendable for a query of cumsum 1 / 2 (suffix att mask = [1, 1, 0, ...]:
  // modeling_mvla.py:686-692): the prefix part comes from the pad bytes, the suffix part from the cumsum rule
  const float kNeg = -2.3819763e38f;
  unsigned ok_pref = 0;
  int cum_key[kKeysMax / 32];
#pragma unroll
  for (int i = 0; i < kKeysMax / 32; ++i) {
    const int j = i * 32 + lane;
    cum_key[i] = 0x7fffffff;
    if (j < p.P) {
      if (__ldg(p.pad + (int64_t)b * p.P + j) != 0) ok_pref |= 1u << i;
    } else if (j < n_keys) {
      cum_key[i] = __ldg(p.cum + (j - p.P));
    }
  }
  SYNC();                    // padding rows visible
  mbar_wait(&bars[0], par);

  // ---- scores: warp w takes the 16-key blocks w, w + 8, w + 16; the Q fragments of a k chunk are loaded once and
  // reused for all of the warp's key blocks (shared-memory bandwidth, not the tensor pipe, bounds this phase) ---------
  const int MT = (R + 15) >> 4;
  const int n_blk = kr >> 4;
  {
    float acc[3][3][2][4];  // [key block][m tile][n tile]
#pragma unroll
    for (int kb3 = 0; kb3 < 3; ++kb3)
#pragma unroll
      for (int mi = 0; mi < 3; ++mi)
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
#pragma unroll
          for (int e = 0; e < 4; ++e) acc[kb3][mi][nt][e] = 0.0f;
#pragma unroll 2
    for (int c = 0; c < 8; ++c) {
      const int ci = c * 4 + t4;
      uint4 lo[3], hi[3];
#pragma unroll
      for (int mi = 0; mi < 3; ++mi) {
        lo[mi] = *reinterpret_cast<const uint4*>(qs + swz(mi * 16 + g, ci));
        hi[mi] = *reinterpret_cast<const uint4*>(qs + swz(mi * 16 + g + 8, ci));
      }
#pragma unroll
      for (int kb3 = 0; kb3 < 3; ++kb3) {
        const int kblk = warp + kb3 * 8;
        if (kblk < n_blk) {
          const uint4 ka = *reinterpret_cast<const uint4*>(ks + swz(kblk * 16 + g, ci));
          const uint4 kbv = *reinterpret_cast<const uint4*>(ks + swz(kblk * 16 + 8 + g, ci));
#pragma unroll
          for (int mi = 0; mi < 3; ++mi) {
            if (mi < MT) {
              mma_bf16_16816(acc[kb3][mi][0], lo[mi].x, hi[mi].x, lo[mi].y, hi[mi].y, ka.x, ka.y);
              mma_bf16_16816(acc[kb3][mi][0], lo[mi].z, hi[mi].z, lo[mi].w, hi[mi].w, ka.z, ka.w);
              mma_bf16_16816(acc[kb3][mi][1], lo[mi].x, hi[mi].x, lo[mi].y, hi[mi].y, kbv.x, kbv.y);
              mma_bf16_16816(acc[kb3][mi][1], lo[mi].z, hi[mi].z, lo[mi].w, hi[mi].w, kbv.z, kbv.w);
            }
          }
        }
      }
    }
#pragma unroll
    for (int kb3 = 0; kb3 < 3; ++kb3) {
      const int kblk = warp + kb3 * 8;
      if (kblk < n_blk) {
#pragma unroll
        for (int mi = 0; mi < 3; ++mi) {
          if (mi < MT) {
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
              const int col = kblk * 16 + nt * 8 + t4 * 2;
              *reinterpret_cast<__nv_bfloat162*>(ss + (mi * 16 + g) * kSPitch + col * 2) = __floats2bfloat162_rn(
                  bf16_round(acc[kb3][mi][nt][0]) * 0.0625f, bf16_round(acc[kb3][mi][nt][1]) * 0.0625f);
              *reinterpret_cast<__nv_bfloat162*>(ss + (mi * 16 + g + 8) * kSPitch + col * 2) = __floats2bfloat162_rn(
                  bf16_round(acc[kb3][mi][nt][2]) * 0.0625f, bf16_round(acc[kb3][mi][nt][3]) * 0.0625f);
            }
          }
        }
      }
    }
  }
  fence_proxy_async_smem();  // this thread's generic-proxy accesses to K / Q are ordered before the async-proxy overwrite
  SYNC();                    // K and Q are dead: V^T may overwrite them

  // ---- V^T of this sample streams in while the softmax runs --------------------------------------------------------------
  // V^T rows are contiguous in the cache ([256][vt_pitch]): one bulk copy; the P.V loop reads keys up to kp, i.e. up to
  // 32 B past a row's end = the head of the next row (finite values) times probabilities that are exactly 0
  if (threadIdx.x == 0) {
    const bf16* vb = p.vt + (int64_t)b * 256 * p.vt_pitch;
    mbar_arrive_expect_tx(&bars[1], (uint32_t)(256 * p.vt_pitch * 2));
    bulk_g2s(vs, vb, (uint32_t)(256 * p.vt_pitch * 2), &bars[1]);
  }

  // ---- masked softmax, one warp per row; probabilities (bf16) replace the scores in place --------------------------------
  // warp w owns rows w, w + 8, ...; they are processed together (6 independent instruction streams) because one row
  // at a time is a chain of dependent shuffles / exps that leaves the SM idle
  {
    constexpr int NI = kKeysMax / 32, NR = kRows / 8;
    float v[NR][NI], mx[NR], sum[NR];
#pragma unroll
    for (int k = 0; k < NR; ++k) {
      const int r = warp + k * 8;
      mx[k] = -INFINITY;
      if (k * 8 >= R) continue;  // uniform: none of the CTA's rows uses this slot
      const bf16* row = reinterpret_cast<const bf16*>(ss + r * kSPitch);
      const int cq = r < R ? __ldg(p.cum + r / hpg) : 0;
#pragma unroll
      for (int i = 0; i < NI; ++i) {
        const int j = i * 32 + lane;
        float x = -INFINITY;
        if (r < R && j < n_keys) {
          const bool ok = j < p.P ? ((ok_pref >> i) & 1u) != 0 : cum_key[i] <= cq;
          x = ok ? __bfloat162float(row[j]) : kNeg;
        }
        v[k][i] = x;
        mx[k] = fmaxf(mx[k], x);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
#pragma unroll
      for (int k = 0; k < NR; ++k) mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
#pragma unroll
    for (int k = 0; k < NR; ++k) {
      const int r = warp + k * 8;
      sum[k] = 0.0f;
      if (k * 8 >= R) continue;
#pragma unroll
      for (int i = 0; i < NI; ++i) {
        v[k][i] = (r < R && i * 32 + lane < n_keys) ? __expf(v[k][i] - mx[k]) : 0.0f;
        sum[k] += v[k][i];
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
#pragma unroll
      for (int k = 0; k < NR; ++k) sum[k] += __shfl_xor_sync(0xffffffffu, sum[k], o);
#pragma unroll
    for (int k = 0; k < NR; ++k) {
      const int r = warp + k * 8;
      if (r < MT * 16) {
        bf16* row = reinterpret_cast<bf16*>(ss + r * kSPitch);
        const float inv = r < R ? 1.0f / sum[k] : 0.0f;
#pragma unroll
        for (int i = 0; i < NI; ++i) {
          const int j = i * 32 + lane;
          if (j < kp) row[j] = to_bf16(v[k][i] * inv);
        }
      }
    }
  }
  SYNC();                    // probabilities of all rows written
  mbar_wait(&bars[1], par);

  // ---- P.V: warp w owns dims [32 w, 32 w + 32) = 4 n-tiles; k runs over the keys ----------------------------------------
  float acc[3][4][4];
#pragma unroll
  for (int mi = 0; mi < 3; ++mi)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[mi][nt][e] = 0.0f;
  for (int cc = 0; cc < (kp >> 5); ++cc) {
    const int jb = (cc * 32 + t4 * 8) * 2;  // byte offset of this lane's 8 keys
    uint4 vf[4];
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) vf[nt] = *reinterpret_cast<const uint4*>(vs + (warp * 32 + nt * 8 + g) * vpitch + jb);
#pragma unroll
    for (int mi = 0; mi < 3; ++mi) {
      if (mi < MT) {
        const uint4 lo = *reinterpret_cast<const uint4*>(ss + (mi * 16 + g) * kSPitch + jb);
        const uint4 hi = *reinterpret_cast<const uint4*>(ss + (mi * 16 + g + 8) * kSPitch + jb);
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
          mma_bf16_16816(acc[mi][nt], lo.x, hi.x, lo.y, hi.y, vf[nt].x, vf[nt].y);
          mma_bf16_16816(acc[mi][nt], lo.z, hi.z, lo.w, hi.w, vf[nt].z, vf[nt].w);
        }
      }
    }
  }
#pragma unroll
  for (int mi = 0; mi < 3; ++mi) {
    if (mi < MT) {
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int r = mi * 16 + g + half * 8;
        if (r < R) {
          bf16* orow = p.att + (int64_t)b * p.S * 2048 + (int64_t)((r / hpg) * 8 + h0 + r % hpg) * 256 + warp * 32 + t4 * 2;
#pragma unroll
          for (int nt = 0; nt < 4; ++nt)
            *reinterpret_cast<__nv_bfloat162*>(orow + nt * 8) =
                __floats2bfloat162_rn(acc[mi][nt][half * 2], acc[mi][nt][half * 2 + 1]);
        }
      }
    }
  }
  par ^= 1;
  fence_proxy_async_smem();  // generic-proxy reads of V^T / scores precede the next unit's bulk copies into the same bytes
  SYNC();
}

}  // namespace battn
}  // namespace pi0
