// HBM-bound kernels around the GEMMs: patch im2col, LayerNorm / GemmaRMSNorm, masked softmax,
// token-embedding gather, position ids, suffix staging. All statistics in fp32; outputs rounded to
// the activation type exactly where the reference rounds.
#include "elementwise.cuh"

namespace pi0 {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
template <int kThreads>
__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  float t = (threadIdx.x < kThreads / 32) ? red[threadIdx.x] : 0.0f;
  if (w == 0) {
    t = warp_sum(t);
    if (l == 0) red[0] = t;
  }
  __syncthreads();
  return red[0];
}
template <int kThreads>
__device__ __forceinline__ float block_max(float v, float* red) {
  v = warp_max(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  float t = (threadIdx.x < kThreads / 32) ? red[threadIdx.x] : -INFINITY;
  if (w == 0) {
    t = warp_max(t);
    if (l == 0) red[0] = t;
  }
  __syncthreads();
  return red[0];
}

// ---------------------------------------------------------------------------------------------
// SigLIP patch im2col: pixels f32 [B,3,224,224] -> A [B*256, kpad], column = c*196 + ky*14 + kx
// (the flattening of the Conv2d(3,1152,14,14) weight), value = rnd(img * scale + shift)
// (modeling_mvla.py:366-367 "img * 2.0 - 1.0" fused here).
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void im2col_patch_kernel(const float* __restrict__ img, T* __restrict__ out, int B, int kpad, float scale,
                                    float shift) {
  pdl_trigger();
  pdl_wait();
  const int patch = blockIdx.x;            // b*256 + py*16 + px
  const int b = patch >> 8, pp = patch & 255;
  const int py = pp >> 4, px = pp & 15;
  T* o = out + (int64_t)patch * kpad;
  for (int i = threadIdx.x; i < kpad; i += blockDim.x) {
    float v = 0.0f;
    if (i < 588) {
      int c = i / 196, r = i % 196;
      int ky = r / 14, kx = r % 14;
      float x = img[(((int64_t)b * 3 + c) * 224 + (py * 14 + ky)) * 224 + (px * 14 + kx)];
      v = x * scale + shift;
    }
    Num<T>::st(o + i, v);
  }
}
template <typename T>
int im2col_patch(const float* img, T* out, int B, int kpad, float scale, float shift, cudaStream_t s) {
  PI0_CHECK_CUDA(launch_pdl(im2col_patch_kernel<T>, dim3(B * 256), dim3(128), 0, s, 1, img, out, B, kpad, scale, shift));
  PI0_CHECK_CUDA(cudaGetLastError());
  ++g_launch_count;
  return 0;
}

// ---------------------------------------------------------------------------------------------
// LayerNorm (SigLIP; eps 1e-6, affine, fp32 statistics, torch formulation: biased variance)
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) layernorm_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                        const float* __restrict__ w, const float* __restrict__ bias,
                                                        int D, float eps) {
  pdl_trigger();
  pdl_wait();
  __shared__ float red[8];
  const int64_t row = blockIdx.x;
  const T* xr = x + row * D;
  float s = 0.0f;
  for (int i = threadIdx.x; i < D; i += 256) s += Num<T>::ld(xr + i);
  const float mean = block_sum<256>(s, red) / D;
  float q = 0.0f;
  for (int i = threadIdx.x; i < D; i += 256) {
    float d = Num<T>::ld(xr + i) - mean;
    q += d * d;
  }
  const float var = block_sum<256>(q, red) / D;
  const float rstd = rsqrtf(var + eps);
  T* yr = y + row * D;
  for (int i = threadIdx.x; i < D; i += 256) {
    float v = (Num<T>::ld(xr + i) - mean) * rstd * w[i] + bias[i];
    Num<T>::st(yr + i, v);
  }
}
// Register-resident variant: one warp per row, the row is read from HBM exactly once (16-byte loads), statistics by
// warp shuffles. D <= 32 * 8 * NCH.
template <typename T, int NCH>
__global__ void __launch_bounds__(256) layernorm_warp_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                             const float* __restrict__ w,
                                                             const float* __restrict__ bias, int M, int D, float eps) {
  pdl_trigger();
  pdl_wait();
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= M) return;
  const int lane = threadIdx.x & 31;
  const T* xr = x + row * D;
  float v[NCH][8];
  float s = 0.0f;
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int i = (c * 32 + lane) * 8;
    if (i < D) {
      load8<T>(xr + i, v[c]);
#pragma unroll
      for (int j = 0; j < 8; ++j) s += v[c][j];
    }
  }
  const float mean = warp_sum(s) / D;
  float q = 0.0f;
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int i = (c * 32 + lane) * 8;
    if (i < D) {
#pragma unroll
      for (int j = 0; j < 8; ++j) { float d = v[c][j] - mean; q += d * d; }
    }
  }
  const float rstd = rsqrtf(warp_sum(q) / D + eps);
  T* yr = y + row * D;
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int i = (c * 32 + lane) * 8;
    if (i < D) {
      float wv[8], bv[8], o[8];
      load8<float>(w + i, wv);
      load8<float>(bias + i, bv);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = (v[c][j] - mean) * rstd * wv[j] + bv[j];
      store8<T>(yr + i, o);
    }
  }
}
template <typename T>
int layernorm(const T* x, T* y, const float* w, const float* b, int M, int D, float eps, cudaStream_t s) {
  if (D % 8 == 0 && D <= 32 * 8 * 5) {
    PI0_CHECK_CUDA(launch_pdl(layernorm_warp_kernel<T, 5>, dim3((M + 7) / 8), dim3(256), 0, s, 1, x, y, w, b, M, D, eps));
  } else {
    PI0_CHECK_CUDA(launch_pdl(layernorm_kernel<T>, dim3(M), dim3(256), 0, s, 1, x, y, w, b, D, eps));
  }
  PI0_CHECK_CUDA(cudaGetLastError());
  ++g_launch_count;
  return 0;
}

// ---------------------------------------------------------------------------------------------
// GemmaRMSNorm: y = rnd((x * rsqrt(mean(x^2) + eps)) * g), g = 1 + w precomputed in fp32.
// Input rows may be gathered: in_row = (r / rpb_in) * rpb_out + off + r % rpb_in.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) rmsnorm_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                      const float* __restrict__ g, int D, float eps, int rpb_in,
                                                      int rpb_out, int off) {
  pdl_trigger();
  pdl_wait();
  __shared__ float red[8];
  const int r = blockIdx.x;
  const int64_t in_row = rpb_in > 0 ? (int64_t)(r / rpb_in) * rpb_out + off + (r % rpb_in) : r;
  const T* xr = x + in_row * D;
  float q = 0.0f;
  for (int i = threadIdx.x * 8; i < D; i += 256 * 8) {
    float v[8];
    load8<T>(xr + i, v);
#pragma unroll
    for (int j = 0; j < 8; ++j) q += v[j] * v[j];
  }
  const float ms = block_sum<256>(q, red) / D;
  const float rs = rsqrtf(ms + eps);
  T* yr = y + (int64_t)r * D;
  for (int i = threadIdx.x * 8; i < D; i += 256 * 8) {
    float v[8];
    load8<T>(xr + i, v);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = (v[j] * rs) * g[i + j];
    store8<T>(yr + i, v);
  }
}
template <typename T, int NCH>
__global__ void __launch_bounds__(256) rmsnorm_warp_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                           const float* __restrict__ g, int M, int D, float eps,
                                                           int rpb_in, int rpb_out, int off) {
  pdl_trigger();
  pdl_wait();
  const int r = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (r >= M) return;
  const int lane = threadIdx.x & 31;
  const int64_t in_row = rpb_in > 0 ? (int64_t)(r / rpb_in) * rpb_out + off + (r % rpb_in) : r;
  const T* xr = x + in_row * D;
  float v[NCH][8];
  float q = 0.0f;
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int i = (c * 32 + lane) * 8;
    if (i < D) {
      load8<T>(xr + i, v[c]);
#pragma unroll
      for (int j = 0; j < 8; ++j) q += v[c][j] * v[c][j];
    }
  }
  const float rs = rsqrtf(warp_sum(q) / D + eps);
  T* yr = y + (int64_t)r * D;
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int i = (c * 32 + lane) * 8;
    if (i < D) {
      float gv[8], o[8];
      load8<float>(g + i, gv);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = (v[c][j] * rs) * gv[j];
      store8<T>(yr + i, o);
    }
  }
}
template <typename T>
int rmsnorm(const T* x, T* y, const float* g, int M, int D, float eps, int rpb_in, int rpb_out, int off,
            cudaStream_t s) {
  PI0_REQUIRE(D % 8 == 0, "rmsnorm: D=%d must be a multiple of 8", D);
  if (D <= 1024) PI0_CHECK_CUDA(launch_pdl(rmsnorm_warp_kernel<T, 4>, dim3((M + 7) / 8), dim3(256), 0, s, 1, x, y, g, M, D, eps, rpb_in, rpb_out, off));
  else if (D <= 2048) PI0_CHECK_CUDA(launch_pdl(rmsnorm_warp_kernel<T, 8>, dim3((M + 7) / 8), dim3(256), 0, s, 1, x, y, g, M, D, eps, rpb_in, rpb_out, off));
  else PI0_CHECK_CUDA(launch_pdl(rmsnorm_kernel<T>, dim3(M), dim3(256), 0, s, 1, x, y, g, D, eps, rpb_in, rpb_out, off));
  PI0_CHECK_CUDA(cudaGetLastError());
  ++g_launch_count;
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Masked softmax over attention scores, in place. One warp per row.
//   scores [nb, rows_per_b, pitch], keys n_keys.
//   key j < n_prefix  : allowed iff pad[b*pad_pitch + j]            (pad == nullptr: all allowed)
//   key j >= n_prefix : allowed iff cum[j - n_prefix] <= cum[q]     q = (row / heads) suffix token
// Masked logits take the reference's finite -2.3819763e38 (mllm.py:301-303), so a fully masked row
// yields a uniform distribution instead of NaN.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) softmax_kernel(T* __restrict__ sc, int64_t rows_total, int rows_per_b,
                                                      int64_t pitch, int n_keys, int n_prefix,
                                                      const uint8_t* __restrict__ pad, int pad_pitch,
                                                      const int* __restrict__ cum, int heads) {
  pdl_trigger();
  pdl_wait();
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows_total) return;
  const int lane = threadIdx.x & 31;
  const int b = (int)(row / rows_per_b);
  const int rq = (int)(row % rows_per_b);
  T* p = sc + row * pitch;
  const float kNeg = -2.3819763e38f;
  int cq = 0;
  if (cum != nullptr) cq = cum[rq / heads];
  float mx = -INFINITY;
  for (int j = lane; j < n_keys; j += 32) {
    bool ok;
    if (j < n_prefix) ok = (pad == nullptr) || pad[(int64_t)b * pad_pitch + j] != 0;
    else ok = cum[j - n_prefix] <= cq;
    float v = ok ? Num<T>::ld(p + j) : kNeg;
    mx = fmaxf(mx, v);
  }
  mx = warp_max(mx);
  float sum = 0.0f;
  for (int j = lane; j < n_keys; j += 32) {
    bool ok;
    if (j < n_prefix) ok = (pad == nullptr) || pad[(int64_t)b * pad_pitch + j] != 0;
    else ok = cum[j - n_prefix] <= cq;
    float v = ok ? Num<T>::ld(p + j) : kNeg;
    sum += expf(v - mx);
  }
  sum = warp_sum(sum);
  const float inv = 1.0f / sum;
  for (int j = lane; j < n_keys; j += 32) {
    bool ok;
    if (j < n_prefix) ok = (pad == nullptr) || pad[(int64_t)b * pad_pitch + j] != 0;
    else ok = cum[j - n_prefix] <= cq;
    float v = ok ? Num<T>::ld(p + j) : kNeg;
    Num<T>::st(p + j, expf(v - mx) * inv);
  }
}
// Register-resident variant: one warp per row, each lane owns NCH chunks of 8 consecutive keys, so the row crosses
// HBM once in each direction with 16-byte accesses. Needs n_keys <= 256 * NCH, n_prefix % 8 == 0, pitch % 8 == 0.
template <typename T, int NCH>
__global__ void __launch_bounds__(256) softmax_warp_kernel(T* __restrict__ sc, int64_t rows_total, int rows_per_b,
                                                           int64_t pitch, int n_keys, int n_prefix,
                                                           const uint8_t* __restrict__ pad, int pad_pitch,
                                                           const int* __restrict__ cum, int heads) {
  pdl_trigger();
  pdl_wait();
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= rows_total) return;
  const int lane = threadIdx.x & 31;
  const int b = (int)(row / rows_per_b);
  const int rq = (int)(row % rows_per_b);
  T* p = sc + row * pitch;
  const float kNeg = -2.3819763e38f;
  int cq = 0;
  if (cum != nullptr) cq = cum[rq / heads];
  float v[NCH][8];
  float mx = -INFINITY;
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int j0 = (c * 32 + lane) * 8;
    if (j0 < n_keys) {
      load8<T>(p + j0, v[c]);
      if (j0 < n_prefix) {
        if (pad != nullptr) {
          const uint2 pm = *reinterpret_cast<const uint2*>(pad + (int64_t)b * pad_pitch + j0);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t wd = j < 4 ? pm.x : pm.y;
            if (((wd >> (8 * (j & 3))) & 0xffu) == 0) v[c][j] = kNeg;
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if (j0 + j >= n_keys) v[c][j] = -INFINITY;          // beyond the row: not part of the softmax
          else if (!(cum[j0 + j - n_prefix] <= cq)) v[c][j] = kNeg;
        }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) mx = fmaxf(mx, v[c][j]);
    }
  }
  mx = warp_max(mx);
  float sum = 0.0f;
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int j0 = (c * 32 + lane) * 8;
    if (j0 < n_keys) {
#pragma unroll
      for (int j = 0; j < 8; ++j) { v[c][j] = expf(v[c][j] - mx); sum += v[c][j]; }
    }
  }
  sum = warp_sum(sum);
  const float inv = 1.0f / sum;
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int j0 = (c * 32 + lane) * 8;
    if (j0 < n_keys) {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[c][j] *= inv;
      store8<T>(p + j0, v[c]);
    }
  }
}
template <typename T>
int softmax_rows(T* sc, int64_t rows_total, int rows_per_b, int64_t pitch, int n_keys, int n_prefix,
                 const uint8_t* pad, int pad_pitch, const int* cum, int heads, cudaStream_t s) {
  int64_t blocks = (rows_total + 7) / 8;
  // keys in [n_prefix, n_keys) need the cumsum rule; a prefix-only row with n_keys % 8 != 0 takes the generic kernel
  const bool vec_ok = pitch % 8 == 0 && n_prefix % 8 == 0 && pitch >= (int64_t)((n_keys + 7) / 8) * 8 &&
                      (pad == nullptr || pad_pitch % 8 == 0) && (n_keys == n_prefix ? n_keys % 8 == 0 : cum != nullptr) &&
                      (reinterpret_cast<uintptr_t>(sc) % 32 == 0) &&
                      (pad == nullptr || reinterpret_cast<uintptr_t>(pad) % 8 == 0);
#define PI0_SM(NCH) PI0_CHECK_CUDA(launch_pdl(softmax_warp_kernel<T, NCH>, dim3((unsigned)blocks), dim3(256), 0, s, 1, \
                                              sc, rows_total, rows_per_b, pitch, n_keys, n_prefix, pad, pad_pitch, cum, heads))
  if (vec_ok && n_keys <= 256) PI0_SM(1);
  else if (vec_ok && n_keys <= 512) PI0_SM(2);
  else if (vec_ok && n_keys <= 768) PI0_SM(3);
  else if (vec_ok && n_keys <= 1024) PI0_SM(4);
  else
    PI0_CHECK_CUDA(launch_pdl(softmax_kernel<T>, dim3((unsigned)blocks), dim3(256), 0, s, 1, sc, rows_total, rows_per_b,
                              pitch, n_keys, n_prefix, pad, pad_pitch, cum, heads));
#undef PI0_SM
  PI0_CHECK_CUDA(cudaGetLastError());
  ++g_launch_count;
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Language embedding gather: out[(b*P + row_off + t), :] = rnd(embed[tok] * scale)
// (modeling_mvla.py:612-616; scale = sqrt(2048) as an fp32 scalar)
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void embed_tokens_kernel(const int64_t* __restrict__ tok, const T* __restrict__ table, T* __restrict__ out,
                                    int L, int P, int row_off, int D, int64_t vocab, float scale) {
  pdl_trigger();
  pdl_wait();
  const int bt = blockIdx.x;
  const int b = bt / L, t = bt % L;
  int64_t id = tok[bt];
  if (id < 0) id = 0;
  if (id >= vocab) id = vocab - 1;
  const T* src = table + id * D;
  T* dst = out + ((int64_t)b * P + row_off + t) * D;
  for (int i = threadIdx.x * 8; i < D; i += blockDim.x * 8) {
    float v[8];
    load8<T>(src + i, v);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] *= scale;
    store8<T>(dst + i, v);
  }
}
template <typename T>
int embed_tokens(const int64_t* tok, const T* table, T* out, int B, int L, int P, int row_off, int D, int64_t vocab,
                 float scale, cudaStream_t s) {
  PI0_CHECK_CUDA(launch_pdl(embed_tokens_kernel<T>, dim3(B * L), dim3(256), 0, s, 1, tok, table, out, L, P, row_off, D, vocab, scale));
  PI0_CHECK_CUDA(cudaGetLastError());
  ++g_launch_count;
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Position ids (bit-exact integer work): prefix pos = cumsum(pad) - 1 (modeling_mvla.py:766),
// suffix pos = sum(pad) + s. One block per sample.
// ---------------------------------------------------------------------------------------------
__global__ void position_ids_kernel(const uint8_t* __restrict__ pad, int P, int S, int* __restrict__ prefix_pos,
                                    int* __restrict__ suffix_pos) {
  pdl_trigger();
  pdl_wait();
  const int b = blockIdx.x;
  if (threadIdx.x == 0) {
    int c = 0;
    for (int j = 0; j < P; ++j) {
      c += pad[(int64_t)b * P + j] != 0;
      prefix_pos[(int64_t)b * P + j] = c - 1;
    }
    for (int s = 0; s < S; ++s) suffix_pos[(int64_t)b * S + s] = c + s;
  }
}
int position_ids(const uint8_t* pad, int B, int P, int S, int* prefix_pos, int* suffix_pos, cudaStream_t s) {
  PI0_CHECK_CUDA(launch_pdl(position_ids_kernel, dim3(B), dim3(32), 0, s, 1, pad, P, S, prefix_pos, suffix_pos));
  PI0_CHECK_CUDA(cudaGetLastError());
  ++g_launch_count;
  return 0;
}

// build the [B, P] pad mask from per-camera image masks and the language mask (modeling_mvla.py:603-619)
__global__ void build_pad_kernel(const uint8_t* __restrict__ img_masks, const uint8_t* __restrict__ lang_masks,
                                 uint8_t* __restrict__ pad, int B, int ncam, int L) {
  pdl_trigger();
  pdl_wait();
  const int P = ncam * 256 + L;
  const int b = blockIdx.x;
  for (int j = threadIdx.x; j < P; j += blockDim.x) {
    uint8_t v;
    if (j < ncam * 256) v = img_masks[(int64_t)(j >> 8) * B + b];
    else v = lang_masks[(int64_t)b * L + (j - ncam * 256)];
    pad[(int64_t)b * P + j] = v != 0;
  }
}
int build_pad(const uint8_t* img_masks, const uint8_t* lang_masks, uint8_t* pad, int B, int ncam, int L,
              cudaStream_t s) {
  PI0_CHECK_CUDA(launch_pdl(build_pad_kernel, dim3(B), dim3(128), 0, s, 1, img_masks, lang_masks, pad, B, ncam, L));
  PI0_CHECK_CUDA(cudaGetLastError());
  ++g_launch_count;
  return 0;
}

// fp32 -> activation type (x_t and state enter the Linear layers in the activation type under autocast)
template <typename T>
__global__ void cast_kernel(const float* __restrict__ in, T* __restrict__ out, int64_t n) {
  pdl_trigger();
  pdl_wait();
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) Num<T>::st(out + i, in[i]);
}
template <typename T>
int cast_f32(const float* in, T* out, int64_t n, cudaStream_t s) {
  PI0_CHECK_CUDA(launch_pdl(cast_kernel<T>, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, s, 1, in, out, n));
  PI0_CHECK_CUDA(cudaGetLastError());
  ++g_launch_count;
  return 0;
}

#define PI0_INST(T)                                                                                              \
  template int im2col_patch<T>(const float*, T*, int, int, float, float, cudaStream_t);                           \
  template int layernorm<T>(const T*, T*, const float*, const float*, int, int, float, cudaStream_t);             \
  template int rmsnorm<T>(const T*, T*, const float*, int, int, float, int, int, int, cudaStream_t);              \
  template int softmax_rows<T>(T*, int64_t, int, int64_t, int, int, const uint8_t*, int, const int*, int,         \
                               cudaStream_t);                                                                     \
  template int embed_tokens<T>(const int64_t*, const T*, T*, int, int, int, int, int, int64_t, float,             \
                               cudaStream_t);                                                                     \
  template int cast_f32<T>(const float*, T*, int64_t, cudaStream_t);
PI0_INST(float)
PI0_INST(bf16)

}  // namespace pi0
