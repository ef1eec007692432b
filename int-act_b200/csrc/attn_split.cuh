// Split form of the denoise-step attention (scores | masked softmax + P.V), shared by the persistent denoise kernel
// (denoise_mega.cu: two grid-wide sub-phases) and the two stand-alone kernels of attn_split.cu (large rollout batches).
// The parameter type PT only needs the fields q, att, sc, pad, cum, S, P, kv_pitch, vt_pitch, kp.
#pragma once
#include "attn_decode.cuh"

namespace pi0 {

constexpr int kMegaKeyChunks = 12;  // keys (padded to 32) / 32 the split attention holds in registers: P + S <= 384

// ---- attention of the denoise step, spread over the whole grid in two warp-granular sub-phases -------------------
// (a) scores: unit = (sample, head, block of 16 keys); one warp computes S x 16 scores = rnd(rnd(q.k) * 256^-0.5) and
//     stores them (bf16) to the global scratch sc[((b*8 + h)*S + t)*kp + j];
// (b) after a grid barrier: unit = (sample, head, 128 output dims) per CTA; its warps do the masked softmax of the head's S
//     rows (fp32, finite -2.3819763e38 mask, probabilities rounded to bf16 into a per-warp shared-memory tile) and
//     multiplies by its 16 rows of V^T.
// Same arithmetic and rounding points as decode_attn_body (attn_decode.cuh); only the work split differs.
template <typename PT>
__device__ __forceinline__ void attn_scores_unit(const PT& p, const bf16* kc, int u, int lane) {
  const int g = lane >> 2, t4 = lane & 3;
  const int n_kb = p.kp >> 4;
  const int kb = u % n_kb, bh = u / n_kb, h = bh & 7, b = bh >> 3;
  const int n_keys = p.P + p.S;
  const int j0 = kb * 16;
  int ja = j0 + g, jb = j0 + 8 + g;
  ja = ja < n_keys ? ja : n_keys - 1;
  jb = jb < n_keys ? jb : n_keys - 1;
  const bf16* kbase = kc + (int64_t)b * p.kv_pitch * 256;
  const bf16* qlo = p.q + ((int64_t)(b * p.S + g) * 8 + h) * 256 + t4 * 8;
  const bf16* qhi = p.q + ((int64_t)(b * p.S + g + 8) * 8 + h) * 256 + t4 * 8;
  const bool lo_ok = g < p.S, hi_ok = g + 8 < p.S;
  float acc[2][4];
#pragma unroll
  for (int nt = 0; nt < 2; ++nt)
#pragma unroll
    for (int e = 0; e < 4; ++e) acc[nt][e] = 0.0f;
#pragma unroll
  for (int c0 = 0; c0 < 8; c0 += 4) {  // two rounds of 16 independent 16-byte loads per lane
    uint4 ka[4], kbv[4], lo[4], hi[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      ka[c] = ldcg_u4(kbase + (int64_t)ja * 256 + (c0 + c) * 32 + t4 * 8);
      kbv[c] = ldcg_u4(kbase + (int64_t)jb * 256 + (c0 + c) * 32 + t4 * 8);
      lo[c] = lo_ok ? ldcg_u4(qlo + (c0 + c) * 32) : make_uint4(0, 0, 0, 0);
      hi[c] = hi_ok ? ldcg_u4(qhi + (c0 + c) * 32) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      mma_bf16_16816(acc[0], lo[c].x, hi[c].x, lo[c].y, hi[c].y, ka[c].x, ka[c].y);
      mma_bf16_16816(acc[0], lo[c].z, hi[c].z, lo[c].w, hi[c].w, ka[c].z, ka[c].w);
      mma_bf16_16816(acc[1], lo[c].x, hi[c].x, lo[c].y, hi[c].y, kbv[c].x, kbv[c].y);
      mma_bf16_16816(acc[1], lo[c].z, hi[c].z, lo[c].w, hi[c].w, kbv[c].z, kbv[c].w);
    }
  }
  bf16* srow = p.sc + (int64_t)bh * p.S * p.kp;
#pragma unroll
  for (int nt = 0; nt < 2; ++nt) {
    const int col = j0 + nt * 8 + t4 * 2;
    if (lo_ok)
      *reinterpret_cast<__nv_bfloat162*>(srow + (int64_t)g * p.kp + col) = __floats2bfloat162_rn(
          bf16_round(acc[nt][0]) * 0.0625f, bf16_round(acc[nt][1]) * 0.0625f);
    if (hi_ok)
      *reinterpret_cast<__nv_bfloat162*>(srow + (int64_t)(g + 8) * p.kp + col) = __floats2bfloat162_rn(
          bf16_round(acc[nt][2]) * 0.0625f, bf16_round(acc[nt][3]) * 0.0625f);
  }
}

// bytes of the per-warp probability tile: 16 rows, pitch = 64 mod 128 bytes
__host__ __device__ constexpr int attn_p_pitch_bytes(int kp) { return kp * 2 + (((kp * 2) & 127) ? 128 : 64); }

// unit = (sample, head, half of the 256 output dims), executed by the 8 consumer warps of one CTA: warp w does the
// masked softmax of rows w and w + 8 into the shared probability tile, then (after a CTA barrier) the P.V product for
// 16 of the unit's 128 dims.
template <typename PT>
__device__ __forceinline__ void attn_pv_unit(const PT& p, const bf16* vt, uint8_t* ptile, int u, int warp,
                                             int lane) {
  const int g = lane >> 2, t4 = lane & 3;
  const int dq = u & 3, bh = u >> 2, h = bh & 7, b = bh >> 3;
  const int n_keys = p.P + p.S, kp = p.kp;
  const int ppb = attn_p_pitch_bytes(kp);
  const float kNeg = -2.3819763e38f;
  const bf16* srow = p.sc + (int64_t)bh * p.S * kp;
  // V^T fragments of the first round are independent of the softmax: issue them first
  const int d0 = dq * 64 + warp * 8;
  const bf16* va = vt + ((int64_t)b * 256 + d0 + g) * p.vt_pitch;
  const int n_chunks = kp >> 5;
  uint4 fa[kMegaKeyChunks];  // kp <= 384: the whole key range of this warp's 8 dims in one round of loads
#pragma unroll
  for (int cc = 0; cc < kMegaKeyChunks; ++cc) {
    const int j = cc * 32 + t4 * 8;
    fa[cc] = make_uint4(0, 0, 0, 0);
    if (cc < n_chunks && j + 8 <= p.vt_pitch) fa[cc] = ldcg_u4(va + j);
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
#pragma unroll
    for (int i = 0; i < kMegaKeyChunks; ++i) {  // all loads unconditional and independent: one L2 round trip
      const int j = i * 32 + lane;
      const int jc = j < kp ? j : kp - 1;
      const unsigned short raw = __ldcg(reinterpret_cast<const unsigned short*>(srow) + (int64_t)r * kp + jc);
      v[i] = __bfloat162float(__ushort_as_bfloat16(raw));
      const int jp = j < p.P ? j : p.P - 1;
      const bool okp = __ldg(p.pad + (int64_t)b * p.P + jp) != 0;
      const int js = min(max(j - p.P, 0), p.S - 1);
      const bool oks = __ldg(p.cum + js) <= cq;
      if (j < n_keys && (j < p.P ? okp : oks)) okbits |= 1u << i;
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
    if (g < p.S)
      *reinterpret_cast<__nv_bfloat162*>(p.att + ((int64_t)(b * p.S + g) * 8 + h) * 256 + d) =
          __floats2bfloat162_rn(acc[0], acc[1]);
    if (g + 8 < p.S)
      *reinterpret_cast<__nv_bfloat162*>(p.att + ((int64_t)(b * p.S + g + 8) * 8 + h) * 256 + d) =
          __floats2bfloat162_rn(acc[2], acc[3]);
  }
  attn_sync256();
}


}  // namespace pi0
