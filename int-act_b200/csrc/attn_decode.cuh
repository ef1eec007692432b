// Device body of the fused suffix attention (scores + masked softmax + P.V) shared by the stand-alone
// decode_attn_kernel (attn_decode.cu) and the persistent denoise kernel (denoise_mega.cu). Executed by the 256 threads
// (8 warps) with threadIdx.x < 256 of a CTA; they synchronise on named barrier 1 so that extra warps of the CTA
// (the weight-streaming producer of the persistent kernel) are not involved. All global reads of q / K / V^T go
// through ld.global.cg: in the persistent kernel those rows were written by OTHER CTAs of the same launch.
// Reference: eager_attention_forward, [mirror] src/model/mvla/mllm.py:265-317; suffix mask modeling_mvla.py:686-692.
#pragma once
#include "elementwise.cuh"

namespace pi0 {

constexpr int kAttnHd = 256;
constexpr int kAttnQPitch = kAttnHd + 32;  // bf16 elements; 576 B = 64 mod 128 -> conflict-free 16-byte fragment loads

__device__ __forceinline__ void attn_sync256() { asm volatile("bar.sync 1, 256;" ::: "memory"); }
__device__ __forceinline__ uint4 ldcg_u4(const void* p) { return __ldcg(reinterpret_cast<const uint4*>(p)); }

__device__ __forceinline__ void mma_bf16_16816(float* c, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                               uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max_f(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// shared memory the body needs for MTM m16 row tiles and kp (padded) keys
__host__ __device__ constexpr size_t attn_smem_bytes(int mtm, int sc_pitch) {
  return (size_t)mtm * 16 * kAttnQPitch * 2 + (size_t)mtm * 16 * sc_pitch * 4;
}

// q    [B*S, 8*256]   (row = token, col = head*256 + d)
// kc   [B, kv_pitch, 256]       K cache of this layer (prefix rows 0..P-1, suffix rows P..P+S-1)
// vt   [B, 256, vt_pitch]       V^T cache of this layer
// att  [B*S, 8*256]
template <int MTM>
__device__ __forceinline__ void decode_attn_body(uint8_t* smem_raw, const bf16* q, const bf16* kc, const bf16* vt,
                                                 bf16* att, const uint8_t* pad, const int* cum, int b, int h0, int S,
                                                 int P, int kv_pitch, int vt_pitch, int hpg, int kp, int sc_pitch) {
  constexpr int kRMax = MTM * 16;
  constexpr int kHd = kAttnHd;
  constexpr int kQPitch = kAttnQPitch;
  bf16* qs = reinterpret_cast<bf16*>(smem_raw);                            // [kRMax][kQPitch]
  float* sc = reinterpret_cast<float*>(smem_raw + kRMax * kQPitch * 2);    // [kRMax][sc_pitch] fp32 (P aliases it)
  const int R = S * hpg;               // row r = t * hpg + hh  (token t, head h0 + hh)
  const int n_keys = P + S;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t4 = lane & 3;
  const int MT = (R + 15) >> 4;

  // ---- stage Q (zero rows beyond R) -------------------------------------------------------------------------
  for (int i = threadIdx.x; i < kRMax * (kHd / 8); i += 256) {
    const int r = i / (kHd / 8), c8 = i % (kHd / 8);
    uint4 v = make_uint4(0, 0, 0, 0);
    if (r < R) {
      const int t = r / hpg, hh = r % hpg;
      v = ldcg_u4(q + ((int64_t)(b * S + t) * 8 + h0 + hh) * kHd + c8 * 8);
    }
    *reinterpret_cast<uint4*>(qs + r * kQPitch + c8 * 8) = v;
  }
  attn_sync256();

  // ---- scores: each warp takes pairs of 8-key tiles ------------------------------------------------------------
  const bf16* kb = kc + (int64_t)b * kv_pitch * kHd;
  const int n_pairs = kp / 16;
  for (int pr = warp; pr < n_pairs; pr += 8) {
    const int j0 = pr * 16;
    int ja = j0 + g, jb = j0 + 8 + g;
    ja = ja < n_keys ? ja : n_keys - 1;
    jb = jb < n_keys ? jb : n_keys - 1;
    uint4 ka[8], kbv[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      ka[c] = ldcg_u4(kb + (int64_t)ja * kHd + c * 32 + t4 * 8);
      kbv[c] = ldcg_u4(kb + (int64_t)jb * kHd + c * 32 + t4 * 8);
    }
    float acc[MTM][2][4];
#pragma unroll
    for (int mi = 0; mi < MTM; ++mi)
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[mi][nt][e] = 0.0f;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
#pragma unroll
      for (int mi = 0; mi < MTM; ++mi) {
        if (mi < MT) {
          const uint4 lo = *reinterpret_cast<const uint4*>(qs + (mi * 16 + g) * kQPitch + c * 32 + t4 * 8);
          const uint4 hi = *reinterpret_cast<const uint4*>(qs + (mi * 16 + g + 8) * kQPitch + c * 32 + t4 * 8);
          mma_bf16_16816(acc[mi][0], lo.x, hi.x, lo.y, hi.y, ka[c].x, ka[c].y);
          mma_bf16_16816(acc[mi][0], lo.z, hi.z, lo.w, hi.w, ka[c].z, ka[c].w);
          mma_bf16_16816(acc[mi][1], lo.x, hi.x, lo.y, hi.y, kbv[c].x, kbv[c].y);
          mma_bf16_16816(acc[mi][1], lo.z, hi.z, lo.w, hi.w, kbv[c].z, kbv[c].w);
        }
      }
    }
#pragma unroll
    for (int mi = 0; mi < MTM; ++mi) {
      if (mi < MT) {
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
          const int col = j0 + nt * 8 + t4 * 2;
          float2 v0, v1;  // rnd(rnd(q.k) * 256^-0.5)
          v0.x = bf16_round(bf16_round(acc[mi][nt][0]) * 0.0625f);
          v0.y = bf16_round(bf16_round(acc[mi][nt][1]) * 0.0625f);
          v1.x = bf16_round(bf16_round(acc[mi][nt][2]) * 0.0625f);
          v1.y = bf16_round(bf16_round(acc[mi][nt][3]) * 0.0625f);
          *reinterpret_cast<float2*>(sc + (mi * 16 + g) * sc_pitch + col) = v0;
          *reinterpret_cast<float2*>(sc + (mi * 16 + g + 8) * sc_pitch + col) = v1;
        }
      }
    }
  }
  attn_sync256();

  // ---- masked softmax, one warp per row; probabilities (bf16) overwrite the head of the row's fp32 storage ------
  const float kNeg = -2.3819763e38f;
  for (int r = warp; r < MT * 16; r += 8) {
    float* row = sc + r * sc_pitch;
    bf16* prow = reinterpret_cast<bf16*>(row);
    if (r >= R) {  // padding rows of the last m16 tile: zero probabilities
      for (int j = lane; j < kp; j += 32) prow[j] = to_bf16(0.0f);
      continue;
    }
    const int cq = cum[r / hpg];
    float v[20];  // kp <= 640
    float mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < 20; ++i) {
      const int j = i * 32 + lane;
      float x = -INFINITY;
      if (j < n_keys) {
        const bool ok = j < P ? pad[(int64_t)b * P + j] != 0 : cum[j - P] <= cq;
        x = ok ? row[j] : kNeg;
      }
      v[i] = x;
      mx = fmaxf(mx, x);
    }
    mx = warp_max_f(mx);
    float sum = 0.0f;
#pragma unroll
    for (int i = 0; i < 20; ++i) {
      v[i] = (i * 32 + lane < n_keys) ? expf(v[i] - mx) : 0.0f;
      sum += v[i];
    }
    sum = warp_sum_f(sum);
    const float inv = 1.0f / sum;
    __syncwarp();  // every lane has read its fp32 scores before the row is overwritten
#pragma unroll
    for (int i = 0; i < 20; ++i) {
      const int j = i * 32 + lane;
      if (j < kp) prow[j] = to_bf16(v[i] * inv);
    }
  }
  attn_sync256();

  // ---- P.V: each warp takes pairs of 8-wide d tiles; k runs over the keys -----------------------------------------
  const bf16* vb = vt + (int64_t)b * kHd * vt_pitch;
  const int pp = sc_pitch * 2;  // pitch of the probability rows in bf16 elements
  const bf16* ps = reinterpret_cast<const bf16*>(sc);
  const int n_chunks = kp / 32;
  for (int pr = warp; pr < kHd / 16; pr += 8) {
    const int d0 = pr * 16;
    float acc[MTM][2][4];
#pragma unroll
    for (int mi = 0; mi < MTM; ++mi)
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[mi][nt][e] = 0.0f;
    const bf16* va = vb + (int64_t)(d0 + g) * vt_pitch;
    const bf16* vbp = vb + (int64_t)(d0 + 8 + g) * vt_pitch;
    for (int c0 = 0; c0 < n_chunks; c0 += 4) {
      uint4 fa[4], fb[4];
#pragma unroll
      for (int cc = 0; cc < 4; ++cc) {
        const int j = (c0 + cc) * 32 + t4 * 8;
        fa[cc] = make_uint4(0, 0, 0, 0);
        fb[cc] = make_uint4(0, 0, 0, 0);
        if (c0 + cc < n_chunks && j + 8 <= vt_pitch) {
          fa[cc] = ldcg_u4(va + j);
          fb[cc] = ldcg_u4(vbp + j);
        }
      }
#pragma unroll
      for (int cc = 0; cc < 4; ++cc) {
        if (c0 + cc < n_chunks) {
          const int j = (c0 + cc) * 32 + t4 * 8;
#pragma unroll
          for (int mi = 0; mi < MTM; ++mi) {
            if (mi < MT) {
              const uint4 lo = *reinterpret_cast<const uint4*>(ps + (mi * 16 + g) * pp + j);
              const uint4 hi = *reinterpret_cast<const uint4*>(ps + (mi * 16 + g + 8) * pp + j);
              mma_bf16_16816(acc[mi][0], lo.x, hi.x, lo.y, hi.y, fa[cc].x, fa[cc].y);
              mma_bf16_16816(acc[mi][0], lo.z, hi.z, lo.w, hi.w, fa[cc].z, fa[cc].w);
              mma_bf16_16816(acc[mi][1], lo.x, hi.x, lo.y, hi.y, fb[cc].x, fb[cc].y);
              mma_bf16_16816(acc[mi][1], lo.z, hi.z, lo.w, hi.w, fb[cc].z, fb[cc].w);
            }
          }
        }
      }
    }
#pragma unroll
    for (int mi = 0; mi < MTM; ++mi) {
      if (mi < MT) {
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
          const int d = d0 + nt * 8 + t4 * 2;
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            const int r = mi * 16 + g + half * 8;
            if (r < R) {
              const int t = r / hpg, hh = r % hpg;
              __nv_bfloat162 o = __floats2bfloat162_rn(acc[mi][nt][half * 2], acc[mi][nt][half * 2 + 1]);
              *reinterpret_cast<__nv_bfloat162*>(att + ((int64_t)(b * S + t) * 8 + h0 + hh) * kHd + d) = o;
            }
          }
        }
      }
    }
  }
}


}  // namespace pi0
