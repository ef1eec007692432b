// GEMM epilogues shared by the tcgen05 kernel (bf16 regime) and the SIMT fp32 check-mode kernel.
// Every epilogue consumes one 32-column fp32 accumulator chunk of one output row (the layout a
// tcgen05.ld 32x32b gives each epilogue thread) and reproduces the reference's rounding points
// (SURVEY.md §9.7): Linear output rounded to the activation type, then bias/residual/activation.
#pragma once
#include "common.cuh"

namespace pi0 {

enum EpiMode : int { EPI_STORE = 0, EPI_GEGLU = 1, EPI_ROPE = 2, EPI_EULER = 3 };
enum ActKind : int { ACT_NONE = 0, ACT_GELU_TANH = 1, ACT_SILU = 2 };

struct EpiParams {
  void* out;            // activation type T, row-major
  int64_t ldc;          // leading dimension of out (elements)
  int64_t out_z0;       // batch strides of out (elements)
  int64_t out_z1;
  const float* bias;    // [N] fp32 or nullptr
  const void* res;      // residual (type T) or nullptr
  int64_t ldr;
  int res_row_mod;      // residual row = row % res_row_mod when > 0 (SigLIP position embedding)
  float alpha;          // t = rnd(t * alpha) when != 1 (attention scale)
  float post_div;       // t = rnd(t / post_div) when != 0   (PaliGemma "/ sqrt(hidden)")
  float post_mul;       // t = rnd(t * post_mul) when != 0   (Pi0 "* sqrt(hidden)")
  int act;              // ActKind applied to the rounded Linear output
  // output row remap: out_row = (row / rpb_in) * rpb_out + row_off + row % rpb_in (identity if rpb_in == 0)
  int rpb_in, rpb_out, row_off;
  // transposed region (V^T for the P.V GEMM): columns >= vt_col0 go to
  //   vt[((b * vt_heads + h) * vt_hd + d) * vt_pitch + vt_toff + t],  b = row / vt_rpb, t = row % vt_rpb
  void* vt;
  int vt_col0, vt_hd, vt_heads, vt_rpb, vt_toff;
  int64_t vt_pitch;
  // EPI_ROPE (Gemma / expert fused QKV, one 256-wide head per N tile)
  const float* rope_cos;  // [max_pos, 128]
  const float* rope_sin;
  const int* pos_ids;     // [M]
  void* kcache;           // [B, kv_pitch, 256] type T
  int n_q_heads;          // tiles [0, n_q_heads) are Q, then one K tile, then one V tile
  int64_t kv_pitch;       // rows per sample in kcache
  // EPI_EULER
  float dt;
  // EPI_STORE, optional: sum of squares of the stored (rounded) values of every 32-column chunk -> rowsq_out[row * rowsq_ld +
  // col0 / 32]; the next GEMM's fused RMSNorm sums the slots of a row in slot order (deterministic, no atomics)
  float* rowsq_out;
  int rowsq_ld;
};

__host__ inline EpiParams epi_default() {
  EpiParams e;
  e.out = nullptr; e.ldc = 0; e.out_z0 = 0; e.out_z1 = 0; e.bias = nullptr; e.res = nullptr; e.ldr = 0;
  e.res_row_mod = 0; e.alpha = 1.0f; e.post_div = 0.0f; e.post_mul = 0.0f; e.act = ACT_NONE;
  e.rpb_in = 0; e.rpb_out = 0; e.row_off = 0;
  e.vt = nullptr; e.vt_col0 = 0x7fffffff; e.vt_hd = 1; e.vt_heads = 1; e.vt_rpb = 1; e.vt_toff = 0; e.vt_pitch = 0;
  e.rope_cos = nullptr; e.rope_sin = nullptr; e.pos_ids = nullptr; e.kcache = nullptr; e.n_q_heads = 0; e.kv_pitch = 0;
  e.dt = 0.0f;
  e.rowsq_out = nullptr; e.rowsq_ld = 0;
  return e;
}

template <typename T>
__device__ __forceinline__ void store8(T* dst, const float* v);
template <>
__device__ __forceinline__ void store8<bf16>(bf16* dst, const float* v) {
  uint4 pk;
  __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]);
  __nv_bfloat162 b = __floats2bfloat162_rn(v[2], v[3]);
  __nv_bfloat162 c = __floats2bfloat162_rn(v[4], v[5]);
  __nv_bfloat162 d = __floats2bfloat162_rn(v[6], v[7]);
  pk.x = *reinterpret_cast<uint32_t*>(&a);
  pk.y = *reinterpret_cast<uint32_t*>(&b);
  pk.z = *reinterpret_cast<uint32_t*>(&c);
  pk.w = *reinterpret_cast<uint32_t*>(&d);
  *reinterpret_cast<uint4*>(dst) = pk;
}
template <>
__device__ __forceinline__ void store8<float>(float* dst, const float* v) {
  *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(dst + 4) = make_float4(v[4], v[5], v[6], v[7]);
}
template <typename T>
__device__ __forceinline__ void load8(const T* src, float* v);
template <>
__device__ __forceinline__ void load8<bf16>(const bf16* src, float* v) {
  uint4 pk = *reinterpret_cast<const uint4*>(src);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&pk);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 f = __bfloat1622float2(h[i]);
    v[2 * i] = f.x;
    v[2 * i + 1] = f.y;
  }
}
template <>
__device__ __forceinline__ void load8<float>(const float* src, float* v) {
  float4 a = *reinterpret_cast<const float4*>(src);
  float4 b = *reinterpret_cast<const float4*>(src + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}

__device__ __forceinline__ int64_t epi_out_row(const EpiParams& ep, int row) {
  if (ep.rpb_in <= 0) return row;
  return (int64_t)(row / ep.rpb_in) * ep.rpb_out + ep.row_off + (row % ep.rpb_in);
}

// ---- EPI_STORE: bias, alpha, post scale, activation, residual; optional transposed V region -----
// Written as whole-chunk passes: every feature test is one uniform branch per 32-column chunk and every pass is a
// straight-line loop over 32 independent elements, so a single epilogue warp per scheduler keeps its ALU pipe
// busy (the element-at-a-time form was latency-bound at ~300 cycles per element).
// direct_store == false: stop after the arithmetic (v[] then holds the final values, to be rounded by the caller's pack);
// used by the tcgen05 kernels' TMA-store path, which never sees a transposed-V region.
template <typename T>
__device__ __forceinline__ void epi_store_chunk(const EpiParams& ep, int row, int col0, int N, int64_t zoff,
                                                float* v, bool direct_store = true) {
  if (col0 >= ep.vt_col0) {  // V^T region: lanes of a warp hold consecutive t => coalesced per d
    const int b = row / ep.vt_rpb, t = row % ep.vt_rpb;
    T* vt = reinterpret_cast<T*>(ep.vt);
    const int dall0 = col0 - ep.vt_col0;
    int h = dall0 / ep.vt_hd, d = dall0 - h * ep.vt_hd;
    T* dst = vt + ((int64_t)(b * ep.vt_heads + h) * ep.vt_hd + d) * ep.vt_pitch + ep.vt_toff + t;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      if (col0 + j < N) {
        float x = v[j] + (ep.bias ? ep.bias[col0 + j] : 0.0f);
        Num<T>::st(dst, x);
        dst += ep.vt_pitch;  // (h, d) -> (h, d + 1) and (h, hd - 1) -> (h + 1, 0) are both one row further
      }
    }
    return;
  }
  const int ngroups = min(4, (N - col0 + 7) >> 3);  // 8-column groups of this chunk that hold real columns
  if (ep.bias) {
    if (ngroups == 4) {
      const float4* bp = reinterpret_cast<const float4*>(ep.bias + col0);
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        float4 bq = __ldg(bp + q);
        v[4 * q] += bq.x; v[4 * q + 1] += bq.y; v[4 * q + 2] += bq.z; v[4 * q + 3] += bq.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (j < ngroups * 8) v[j] += ep.bias[col0 + j];
    }
  }
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = Num<T>::rnd(v[j]);
  if (ep.alpha != 1.0f) {
    const float a = ep.alpha;
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = Num<T>::rnd(v[j] * a);
  }
  if (ep.post_div != 0.0f) {
    const float dv = ep.post_div;
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = Num<T>::rnd(v[j] / dv);
  }
  if (ep.post_mul != 0.0f) {
    const float m = ep.post_mul;
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = Num<T>::rnd(v[j] * m);
  }
  if (ep.act == ACT_GELU_TANH) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = Num<T>::rnd(Num<T>::gelu(v[j]));
  } else if (ep.act == ACT_SILU) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = Num<T>::rnd(silu(v[j]));
  }
  if (ep.res) {
    const int rrow = ep.res_row_mod > 0 ? row % ep.res_row_mod : row;
    const T* res = reinterpret_cast<const T*>(ep.res) + (int64_t)rrow * ep.ldr + col0;  // residuals are never batched
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      if (g < ngroups) {
        float r[8];
        load8<T>(res + g * 8, r);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[g * 8 + j] += r[j];
      }
    }
  }
  if (ep.rowsq_out != nullptr) {
    float q = 0.0f;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const float r = Num<T>::rnd(v[j]);
      if (j < ngroups * 8) q = fmaf(r, r, q);
    }
    ep.rowsq_out[(int64_t)row * ep.rowsq_ld + (col0 >> 5)] = q;
  }
  if (!direct_store) return;
  T* out = reinterpret_cast<T*>(ep.out) + zoff + epi_out_row(ep, row) * ep.ldc + col0;
#pragma unroll
  for (int g = 0; g < 4; ++g)
    if (g < ngroups) store8<T>(out + g * 8, v + g * 8);
}

// ---- EPI_GEGLU: out = gelu_tanh(gate) * up, gate/up interleaved per N tile at pack time ------------
template <typename T>
__device__ __forceinline__ void epi_geglu_chunk(const EpiParams& ep, int row, int out_col0, int n_out, float* gate,
                                                const float* up) {
  T* out = reinterpret_cast<T*>(ep.out) + epi_out_row(ep, row) * ep.ldc + out_col0;
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    if (out_col0 + g * 8 >= n_out) break;
    float t[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float a = Num<T>::rnd(Num<T>::gelu(Num<T>::rnd(gate[g * 8 + j])));
      float u = Num<T>::rnd(up[g * 8 + j]);
      t[j] = a * u;
    }
    store8<T>(out + g * 8, t);
  }
}

// arithmetic only (TMA-store path): gate[j] <- rnd(gelu(rnd(gate))) * rnd(up), rounded by the caller's pack
template <typename T>
__device__ __forceinline__ void epi_geglu_math(float* gate, const float* up) {
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    const float a = Num<T>::rnd(Num<T>::gelu(Num<T>::rnd(gate[j])));
    gate[j] = a * Num<T>::rnd(up[j]);
  }
}

// ---- EPI_ROPE: x1 = dims [c*32, c*32+32), x2 = the same + 128, of one 256-wide head ---------------
// which: 0 = Q head `head`, 1 = K, 2 = V (V is not rotated; x1/x2 are stored as they are)
template <typename T>
__device__ __forceinline__ void epi_rope_chunk(const EpiParams& ep, int row, int which, int head, int c, float* x1,
                                               float* x2) {
  float o1[32], o2[32];
  if (which < 2) {
    int pos = ep.pos_ids[row];
    if (pos < 0) pos = 0;  // leading pad tokens have cumsum(pad)-1 == -1; their K/V are never attended
    const float* cs = ep.rope_cos + (int64_t)pos * 128 + c * 32;
    const float* sn = ep.rope_sin + (int64_t)pos * 128 + c * 32;
    const float4* c4 = reinterpret_cast<const float4*>(cs);
    const float4* s4 = reinterpret_cast<const float4*>(sn);
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const float4 cq = __ldg(c4 + q), sq = __ldg(s4 + q);
      const float cc[4] = {cq.x, cq.y, cq.z, cq.w}, ss[4] = {sq.x, sq.y, sq.z, sq.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int j = 4 * q + e;
        float a = Num<T>::rnd(x1[j]), b = Num<T>::rnd(x2[j]);
        o1[j] = a * cc[e] - b * ss[e];
        o2[j] = b * cc[e] + a * ss[e];
      }
    }
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j) { o1[j] = x1[j]; o2[j] = x2[j]; }
  }
  if (which == 0) {
    T* out = reinterpret_cast<T*>(ep.out) + (int64_t)row * ep.ldc + head * 256 + c * 32;
#pragma unroll
    for (int g = 0; g < 4; ++g) { store8<T>(out + g * 8, o1 + g * 8); store8<T>(out + 128 + g * 8, o2 + g * 8); }
  } else {
    int b = row / ep.vt_rpb, t = row % ep.vt_rpb;
    if (which == 1) {
      T* out = reinterpret_cast<T*>(ep.kcache) + ((int64_t)b * ep.kv_pitch + ep.vt_toff + t) * 256 + c * 32;
#pragma unroll
      for (int g = 0; g < 4; ++g) { store8<T>(out + g * 8, o1 + g * 8); store8<T>(out + 128 + g * 8, o2 + g * 8); }
    } else {
      T* vt = reinterpret_cast<T*>(ep.vt) + (int64_t)b * 256 * ep.vt_pitch + ep.vt_toff + t;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        Num<T>::st(vt + (int64_t)(c * 32 + j) * ep.vt_pitch, o1[j]);
        Num<T>::st(vt + (int64_t)(128 + c * 32 + j) * ep.vt_pitch, o2[j]);
      }
    }
  }
}

// ---- EPI_EULER: x_t += rnd(dt * rnd(v_t)), x_t kept in fp32 (modeling_mvla.py:802) -----------------
template <typename T>
__device__ __forceinline__ void epi_euler_chunk(const EpiParams& ep, int row, int col0, int N, const float* v) {
  float* x = reinterpret_cast<float*>(ep.out) + (int64_t)row * ep.ldc + col0;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    if (col0 + j < N) {
      float vt = Num<T>::rnd(v[j] + (ep.bias ? ep.bias[col0 + j] : 0.0f));
      x[j] += Num<T>::rnd(ep.dt * vt);
    }
  }
}

}  // namespace pi0
