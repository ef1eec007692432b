// Weight-streaming GEMMs for the latency regime of the denoise loop: M = B * S <= 16 suffix rows (batch 1-3) against
// the action expert's [N, K] weights. A 128-row tcgen05 tile would waste > 87 % of the tensor pipe and pay the TMEM /
// barrier prologue on every one of the ~900 launches of a denoise loop, so this path is a bandwidth kernel:
//   * every CTA owns a few 8-column tiles of the output over the whole K; its 8 warps split K, so one round of
//     16-byte loads per lane (8-16 in flight) brings the CTA's whole weight slice into registers as mma.sync B fragments
//     (k is permuted identically on both operands, so 8 consecutive k per lane feed two m16n8k16 MMAs);
//   * the weight (and RMSNorm gain) loads are issued BEFORE griddepcontrol.wait: weights do not depend on the previous
//     kernel, so under programmatic dependent launch the HBM stream of kernel N+1 overlaps the tail of kernel N;
//   * GemmaRMSNorm is fused as a prologue (sum of squares over the K split, cross-warp reduction in shared memory),
//     RoPE + in-place K / V^T cache write, GeGLU and the residual add are fused as epilogues, with the reference's
//     rounding points.
// Replaces, for M <= 16: GemmaRMSNorm + q/k/v_proj + apply_rope, o_proj, GemmaMLP of the expert
// ([mirror] src/model/mvla/expert.py:128-179,387-409; mllm.py:30-53).
#include "elementwise.cuh"

namespace pi0 {

namespace {

enum SkinnyEpi : int { SK_RES = 0, SK_ROPE = 1, SK_GEGLU = 2 };

struct SkinnyParams {
  const bf16* A;      // [M, K]
  int64_t lda;
  const bf16* W;      // [N, K] row-major (K contiguous)
  int M, K;
  const float* gain;  // NORM: 1 + w, fp32 [K]
  float eps;
  // SK_RES
  bf16* out; int64_t ldc;
  const bf16* res; int64_t ldr;
  // SK_ROPE
  bf16* q; int64_t ldq;
  const float* rope_cos; const float* rope_sin; const int* pos_ids;
  bf16* kcache; bf16* vtcache;
  int S, P; int64_t kv_pitch, vt_pitch;
};

__device__ __forceinline__ void mma16816(float* c, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                         uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

__device__ __forceinline__ uint4 ldg_nc_v4(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

__device__ __forceinline__ void unpack8(const uint4& u, float* f) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(h[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 u;
  __nv_bfloat162 a = __floats2bfloat162_rn(f[0], f[1]), b = __floats2bfloat162_rn(f[2], f[3]);
  __nv_bfloat162 c = __floats2bfloat162_rn(f[4], f[5]), d = __floats2bfloat162_rn(f[6], f[7]);
  u.x = *reinterpret_cast<uint32_t*>(&a); u.y = *reinterpret_cast<uint32_t*>(&b);
  u.z = *reinterpret_cast<uint32_t*>(&c); u.w = *reinterpret_cast<uint32_t*>(&d);
  return u;
}

// weight row of tile `nt` column `g` for CTA `c`
template <int EPI, int NT>
__device__ __forceinline__ int weight_row(int c, int nt, int g) {
  if constexpr (EPI == SK_RES) {
    return (c * NT + nt) * 8 + g;
  } else if constexpr (EPI == SK_ROPE) {
    // c = head * 16 + dc; tile 0 = dims [dc*8, +8), tile 1 = the same + 128 (the RoPE partner half)
    return (c >> 4) * 256 + (c & 15) * 8 + nt * 128 + g;
  } else {
    // inter columns [c*16, +16): tiles 0,1 = gate, 2,3 = up; packed per 256 rows as 128 gate rows then 128 up rows
    const int i = c * 16 + (nt & 1) * 8 + g;
    return (i >> 7) * 256 + (i & 127) + (nt >> 1) * 128;
  }
}

template <int NT, int CH, int EPI, bool NORM, bool M16>
__global__ void __launch_bounds__(256) skinny_gemm_kernel(const SkinnyParams p) {
  pdl_trigger();
  __shared__ float red[8][NT][16][8];
  __shared__ float ssq[8][16];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t4 = lane & 3;
  const int c = blockIdx.x;
  const int k0 = warp * (32 * CH) + t4 * 8;

  // ---- weights (and gains): independent of the previous kernel -> in flight before the dependency wait ---------
  uint4 wf[NT][CH];
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) {
    const bf16* wr = p.W + (int64_t)weight_row<EPI, NT>(c, nt, g) * p.K + k0;
#pragma unroll
    for (int cc = 0; cc < CH; ++cc) wf[nt][cc] = ldg_nc_v4(wr + cc * 32);
  }
  float4 gn[NORM ? CH : 1][2];
  if constexpr (NORM) {
#pragma unroll
    for (int cc = 0; cc < CH; ++cc) {
      gn[cc][0] = __ldg(reinterpret_cast<const float4*>(p.gain + k0 + cc * 32));
      gn[cc][1] = __ldg(reinterpret_cast<const float4*>(p.gain + k0 + cc * 32 + 4));
    }
  }
  pdl_wait();

  // ---- activations: rows g (and g + 8) of A, this warp's K slice ---------------------------------------------------
  uint4 alo[CH], ahi[M16 ? CH : 1];
  const bool lo_ok = g < p.M, hi_ok = M16 && (g + 8 < p.M);
#pragma unroll
  for (int cc = 0; cc < CH; ++cc) {
    alo[cc] = make_uint4(0, 0, 0, 0);
    if (lo_ok) alo[cc] = *reinterpret_cast<const uint4*>(p.A + (int64_t)g * p.lda + k0 + cc * 32);
    if constexpr (M16) {
      ahi[cc] = make_uint4(0, 0, 0, 0);
      if (hi_ok) ahi[cc] = *reinterpret_cast<const uint4*>(p.A + (int64_t)(g + 8) * p.lda + k0 + cc * 32);
    }
  }

  if constexpr (NORM) {
    // GemmaRMSNorm prologue: h = rnd((x * rsqrt(mean(x^2) + eps)) * (1 + w))
    float slo = 0.0f, shi = 0.0f;
#pragma unroll
    for (int cc = 0; cc < CH; ++cc) {
      float f[8];
      unpack8(alo[cc], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) slo += f[j] * f[j];
      if constexpr (M16) {
        unpack8(ahi[cc], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) shi += f[j] * f[j];
      }
    }
    slo += __shfl_xor_sync(0xffffffffu, slo, 1);
    slo += __shfl_xor_sync(0xffffffffu, slo, 2);
    shi += __shfl_xor_sync(0xffffffffu, shi, 1);
    shi += __shfl_xor_sync(0xffffffffu, shi, 2);
    if (t4 == 0) { ssq[warp][g] = slo; ssq[warp][g + 8] = shi; }
    __syncthreads();
    float tlo = 0.0f, thi = 0.0f;
#pragma unroll
    for (int w = 0; w < 8; ++w) { tlo += ssq[w][g]; thi += ssq[w][g + 8]; }
    const float rlo = rsqrtf(tlo / (float)p.K + p.eps), rhi = rsqrtf(thi / (float)p.K + p.eps);
#pragma unroll
    for (int cc = 0; cc < CH; ++cc) {
      const float gv[8] = {gn[cc][0].x, gn[cc][0].y, gn[cc][0].z, gn[cc][0].w,
                           gn[cc][1].x, gn[cc][1].y, gn[cc][1].z, gn[cc][1].w};
      float f[8];
      unpack8(alo[cc], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = (f[j] * rlo) * gv[j];
      alo[cc] = pack8(f);
      if constexpr (M16) {
        unpack8(ahi[cc], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = (f[j] * rhi) * gv[j];
        ahi[cc] = pack8(f);
      }
    }
  }

  // ---- MMA over the K slice -----------------------------------------------------------------------------------
  float acc[NT][4];
#pragma unroll
  for (int nt = 0; nt < NT; ++nt)
#pragma unroll
    for (int e = 0; e < 4; ++e) acc[nt][e] = 0.0f;
#pragma unroll
  for (int cc = 0; cc < CH; ++cc) {
    const uint4 lo = alo[cc];
    uint4 hi = make_uint4(0, 0, 0, 0);
    if constexpr (M16) hi = ahi[cc];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      mma16816(acc[nt], lo.x, hi.x, lo.y, hi.y, wf[nt][cc].x, wf[nt][cc].y);
      mma16816(acc[nt], lo.z, hi.z, lo.w, hi.w, wf[nt][cc].z, wf[nt][cc].w);
    }
  }
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) {
    *reinterpret_cast<float2*>(&red[warp][nt][g][t4 * 2]) = make_float2(acc[nt][0], acc[nt][1]);
    *reinterpret_cast<float2*>(&red[warp][nt][g + 8][t4 * 2]) = make_float2(acc[nt][2], acc[nt][3]);
  }
  __syncthreads();

  // ---- cross-warp reduction (fixed order) + epilogue ---------------------------------------------------------------
  auto total = [&](int nt, int row, int col) {
    float s = 0.0f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[w][nt][row][col];
    return s;
  };
  if constexpr (EPI == SK_RES) {
    for (int i = threadIdx.x; i < NT * 128; i += 256) {
      const int nt = i >> 7, row = (i >> 3) & 15, col = i & 7;
      if (row < p.M) {
        const int n = (c * NT + nt) * 8 + col;
        float v = bf16_round(total(nt, row, col));
        v += __bfloat162float(p.res[(int64_t)row * p.ldr + n]);
        p.out[(int64_t)row * p.ldc + n] = to_bf16(v);
      }
    }
  } else if constexpr (EPI == SK_ROPE) {
    static_assert(EPI != SK_ROPE || NT == 2, "RoPE pairs dims d and d + 128");
    if (threadIdx.x < 128) {
      const int row = threadIdx.x >> 3, col = threadIdx.x & 7;
      if (row < p.M) {
        const int head = c >> 4, d = (c & 15) * 8 + col;
        const float a = bf16_round(total(0, row, col)), b = bf16_round(total(1, row, col));
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
          bf16* dst = p.q + (int64_t)row * p.ldq + head * 256 + d;
          dst[0] = to_bf16(o1);
          dst[128] = to_bf16(o2);
        } else if (head == 8) {
          bf16* dst = p.kcache + ((int64_t)bb * p.kv_pitch + p.P + t) * 256 + d;
          dst[0] = to_bf16(o1);
          dst[128] = to_bf16(o2);
        } else {
          bf16* dst = p.vtcache + (int64_t)bb * 256 * p.vt_pitch + p.P + t;
          dst[(int64_t)d * p.vt_pitch] = to_bf16(o1);
          dst[(int64_t)(d + 128) * p.vt_pitch] = to_bf16(o2);
        }
      }
    }
  } else {
    static_assert(EPI != SK_GEGLU || NT == 4, "GeGLU: two gate tiles + the two matching up tiles");
    {
      const int row = threadIdx.x >> 4, col = threadIdx.x & 15;
      if (row < p.M) {
        const float gt = bf16_round(total(col >> 3, row, col & 7));
        const float up = bf16_round(total(2 + (col >> 3), row, col & 7));
        const float act = bf16_round(gelu_tanh_fast(gt));
        p.out[(int64_t)row * p.ldc + c * 16 + col] = to_bf16(act * up);
      }
    }
  }
}

template <int NT, int CH, int EPI, bool NORM>
int launch_skinny(const SkinnyParams& p, int grid, cudaStream_t s) {
  if (p.M > 8) PI0_CHECK_CUDA(launch_pdl(skinny_gemm_kernel<NT, CH, EPI, NORM, true>, dim3(grid), dim3(256), 0, s, 1, p));
  else PI0_CHECK_CUDA(launch_pdl(skinny_gemm_kernel<NT, CH, EPI, NORM, false>, dim3(grid), dim3(256), 0, s, 1, p));
  ++g_launch_count;
  return 0;
}

}  // namespace

bool skinny_supported(int M) { return M >= 1 && M <= 16; }

// q|k|v = RMSNorm(x) . Wqkv^T with RoPE on q, k; q -> [M, 2048], k / v -> the suffix rows of the caches
int skinny_qkv_rope(const void* x, int64_t ldx, const float* gain, const void* wqkv, int M, int K, void* q,
                    const float* rope_cos, const float* rope_sin, const int* pos_ids, void* kcache, void* vtcache, int S,
                    int P, int64_t kv_pitch, int64_t vt_pitch, cudaStream_t s) {
  PI0_REQUIRE(skinny_supported(M) && K == 1024, "skinny_qkv_rope: M=%d K=%d unsupported", M, K);
  SkinnyParams p = {};
  p.A = (const bf16*)x; p.lda = ldx; p.W = (const bf16*)wqkv; p.M = M; p.K = K; p.gain = gain; p.eps = 1e-6f;
  p.q = (bf16*)q; p.ldq = 2048; p.rope_cos = rope_cos; p.rope_sin = rope_sin; p.pos_ids = pos_ids;
  p.kcache = (bf16*)kcache; p.vtcache = (bf16*)vtcache; p.S = S; p.P = P; p.kv_pitch = kv_pitch; p.vt_pitch = vt_pitch;
  return launch_skinny<2, 4, SK_ROPE, true>(p, 10 * 16, s);
}

// out = res + A . W^T   (o_proj: K = 2048, down: K = 4096), N = 1024-multiple of 8
int skinny_linear_res(const void* A, int64_t lda, const void* W, int M, int N, int K, const void* res, int64_t ldr,
                      void* out, int64_t ldc, cudaStream_t s) {
  PI0_REQUIRE(skinny_supported(M) && N % 8 == 0 && (K == 2048 || K == 4096), "skinny_linear_res: M=%d N=%d K=%d", M, N, K);
  SkinnyParams p = {};
  p.A = (const bf16*)A; p.lda = lda; p.W = (const bf16*)W; p.M = M; p.K = K;
  p.out = (bf16*)out; p.ldc = ldc; p.res = (const bf16*)res; p.ldr = ldr;
  if (K == 2048) return launch_skinny<1, 8, SK_RES, false>(p, N / 8, s);
  return launch_skinny<1, 16, SK_RES, false>(p, N / 8, s);
}

// out = gelu_tanh(h . Wg^T) * (h . Wu^T), h = RMSNorm(y); wgu packed per 256 rows as 128 gate + 128 up rows
int skinny_gateup_geglu(const void* y, int64_t ldy, const float* gain, const void* wgu, int M, int K, int inter, void* out,
                        int64_t ldc, cudaStream_t s) {
  PI0_REQUIRE(skinny_supported(M) && K == 1024 && inter % 128 == 0, "skinny_gateup_geglu: M=%d K=%d inter=%d", M, K, inter);
  SkinnyParams p = {};
  p.A = (const bf16*)y; p.lda = ldy; p.W = (const bf16*)wgu; p.M = M; p.K = K; p.gain = gain; p.eps = 1e-6f;
  p.out = (bf16*)out; p.ldc = ldc;
  return launch_skinny<4, 4, SK_GEGLU, true>(p, inter / 16, s);
}

}  // namespace pi0
