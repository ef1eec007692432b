// GPU side of the environment adapters: the per-observation pre / post-processing the
// reference does on the host, one sample at a time, as three small kernels so that a rollout batch crosses PCIe once as
// raw uint8 frames and comes back as the final float64 environment actions.
//   * adapter_resize_norm_kernel   BridgeSimplerBatchAdapter.preprocess images
//                                  (/root/reference/src/experiments/env_adapters/simplerMS3.py:50-63,
//                                  src/utils/pipeline.py:34-69): torch's uint8 bilinear resize (separable, 16-bit
//                                  fixed-point weights, uint8 rounding after each pass - bit-exact) + x/255 + (x-0.5)/0.5
//   * adapter_proprio_kernel       preprocess_proprio + normalize_bound (simplerMS3.py:153-165, base.py:8-18,
//                                  src/utils/geometry.py:49-116,118-175): quat(wxyz) -> R -> R.default_rot^T -> euler sxyz
//   * adapter_postprocess_kernel   postprocess (simplerMS3.py:95-132; simpler.py:96-139,181-190; geometry.py:261-435):
//                                  denormalize_bound, euler -> axis*angle, gripper binarisation; float64 out
// All three are HBM/latency-trivial integer / fp64 work; the point is removing the host loop, not the FLOPs.
#include "../../include/pi0b200.h"
#include "gemm.cuh"

namespace pi0 {

namespace {

// one thread per output pixel (b, oy, ox), all three channels
__global__ void adapter_resize_norm_kernel(const uint8_t* __restrict__ img, float* __restrict__ out, int B, int H, int W,
                                           int oh, int ow, const int* __restrict__ xi, const short* __restrict__ xw,
                                           int xprec, const int* __restrict__ yi, const short* __restrict__ yw,
                                           int yprec, float rescale, float mean, float stdv) {
  pdl_trigger();
  pdl_wait();
  const int64_t n = (int64_t)B * oh * ow;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int ox = (int)(i % ow), oy = (int)((i / ow) % oh), b = (int)(i / ((int64_t)ow * oh));
    const int x0 = xi[2 * ox], x1 = xi[2 * ox + 1], y0 = yi[2 * oy], y1 = yi[2 * oy + 1];
    const int wx0 = xw[2 * ox], wx1 = xw[2 * ox + 1], wy0 = yw[2 * oy], wy1 = yw[2 * oy + 1];
    const uint8_t* base = img + (int64_t)b * H * W * 3;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      int h0, h1;
      if (ow != W) {  // horizontal pass, rounded to uint8 like the temporary image of the separable resize
        const int p00 = base[((int64_t)y0 * W + x0) * 3 + c], p01 = base[((int64_t)y0 * W + x1) * 3 + c];
        const int p10 = base[((int64_t)y1 * W + x0) * 3 + c], p11 = base[((int64_t)y1 * W + x1) * 3 + c];
        h0 = min(max(((1 << (xprec - 1)) + p00 * wx0 + p01 * wx1) >> xprec, 0), 255);
        h1 = min(max(((1 << (xprec - 1)) + p10 * wx0 + p11 * wx1) >> xprec, 0), 255);
      } else {
        h0 = base[((int64_t)y0 * W + ox) * 3 + c];
        h1 = base[((int64_t)y1 * W + ox) * 3 + c];
      }
      int v = h0;
      if (oh != H) v = min(max(((1 << (yprec - 1)) + h0 * wy0 + h1 * wy1) >> yprec, 0), 255);
      // process_images: image * (1/255) then (image - mean) / std, separate fp32 roundings (no FMA contraction)
      const float f = __fdiv_rn(__fsub_rn(__fmul_rn((float)v, rescale), mean), stdv);
      out[(((int64_t)b * 3 + c) * oh + oy) * ow + ox] = f;
    }
  }
}

// cv2.resize(..., INTER_LANCZOS4) on uint8 (the call of simpler.py:47-51; OpenCV imgproc resize.cpp, 8-tap separable
// Lanczos with 11-bit fixed-point coefficients): horizontal pass in int32 over the 8 clamped source rows, vertical pass,
// then (v + 2^21) >> 22 saturated to uint8. One thread per output pixel, all three channels. Bit-exact integer work.
__global__ void adapter_lanczos4_norm_kernel(const uint8_t* __restrict__ img, float* __restrict__ out, int B, int H, int W,
                                             int oh, int ow, const int* __restrict__ xofs, const short* __restrict__ xc,
                                             const int* __restrict__ yofs, const short* __restrict__ yc, float rescale,
                                             float mean, float stdv) {
  pdl_trigger();
  pdl_wait();
  const int64_t n = (int64_t)B * oh * ow;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int ox = (int)(i % ow), oy = (int)((i / ow) % oh), b = (int)(i / ((int64_t)ow * oh));
    const uint8_t* base = img + (int64_t)b * H * W * 3;
    int xi[8], cx[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      xi[k] = min(max(xofs[ox] - 3 + k, 0), W - 1);
      cx[k] = xc[ox * 8 + k];
    }
    long long acc[3] = {0, 0, 0};
#pragma unroll 1
    for (int ky = 0; ky < 8; ++ky) {
      const int y = min(max(yofs[oy] - 3 + ky, 0), H - 1);
      const uint8_t* row = base + (int64_t)y * W * 3;
      int h[3] = {0, 0, 0};
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint8_t* px = row + xi[k] * 3;
        h[0] += px[0] * cx[k];
        h[1] += px[1] * cx[k];
        h[2] += px[2] * cx[k];
      }
      const long long cy = yc[oy * 8 + ky];
      acc[0] += cy * h[0];
      acc[1] += cy * h[1];
      acc[2] += cy * h[2];
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const long long v = (acc[c] + (1ll << 21)) >> 22;
      const int u = (int)min(max(v, 0ll), 255ll);
      out[(((int64_t)b * 3 + c) * oh + oy) * ow + ox] = __fdiv_rn(__fsub_rn(__fmul_rn((float)u, rescale), mean), stdv);
    }
  }
}

struct Mat3 { double m[3][3]; };

__device__ __forceinline__ Mat3 quat2mat_d(double w, double x, double y, double z) {
  Mat3 r;
  const double Nq = w * w + x * x + y * y + z * z;
  if (Nq < 2.220446049250313e-16) {
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) r.m[i][j] = i == j ? 1.0 : 0.0;
    return r;
  }
  const double s = 2.0 / Nq;
  const double X = x * s, Y = y * s, Z = z * s;
  const double wX = w * X, wY = w * Y, wZ = w * Z, xX = x * X, xY = x * Y, xZ = x * Z, yY = y * Y, yZ = y * Z, zZ = z * Z;
  r.m[0][0] = 1.0 - (yY + zZ); r.m[0][1] = xY - wZ; r.m[0][2] = xZ + wY;
  r.m[1][0] = xY + wZ; r.m[1][1] = 1.0 - (xX + zZ); r.m[1][2] = yZ - wX;
  r.m[2][0] = xZ - wY; r.m[2][1] = yZ + wX; r.m[2][2] = 1.0 - (xX + yY);
  return r;
}

// mat2euler, axes "sxyz" (i, j, k = 0, 1, 2; no parity / repetition / frame swap)
__device__ __forceinline__ void mat2euler_sxyz(const Mat3& M, double* e) {
  const double kEps4 = 8.881784197001252e-16;
  const double cy = sqrt(M.m[0][0] * M.m[0][0] + M.m[1][0] * M.m[1][0]);
  if (cy > kEps4) {
    e[0] = atan2(M.m[2][1], M.m[2][2]);
    e[1] = atan2(-M.m[2][0], cy);
    e[2] = atan2(M.m[1][0], M.m[0][0]);
  } else {
    e[0] = atan2(-M.m[1][2], M.m[1][1]);
    e[1] = atan2(-M.m[2][0], cy);
    e[2] = 0.0;
  }
}

// eef [B, 8] = pos(3) | quat wxyz(4) | gripper -> state [B, 7], normalised and rounded to the adapter dtype
template <typename TIn>
__global__ void adapter_proprio_kernel(const TIn* __restrict__ eef, float* __restrict__ state, int B,
                                       const double* __restrict__ a, const double* __restrict__ b, int norm_type,
                                       int round_bf16) {
  pdl_trigger();
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  const TIn* p = eef + (int64_t)i * 8;
  const Mat3 R = quat2mat_d((double)p[3], (double)p[4], (double)p[5], (double)p[6]);
  // R @ default_rot^T with default_rot = [[0,0,1],[0,1,0],[-1,0,0]]: column j of the product = R . default_rot[j,:]
  Mat3 Q;
  for (int r = 0; r < 3; ++r) {
    Q.m[r][0] = R.m[r][0] * 0.0 + R.m[r][1] * 0.0 + R.m[r][2] * 1.0;
    Q.m[r][1] = R.m[r][0] * 0.0 + R.m[r][1] * 1.0 + R.m[r][2] * 0.0;
    Q.m[r][2] = R.m[r][0] * -1.0 + R.m[r][1] * 0.0 + R.m[r][2] * 0.0;
  }
  double raw[7];
  raw[0] = (double)p[0]; raw[1] = (double)p[1]; raw[2] = (double)p[2];
  mat2euler_sxyz(Q, raw + 3);
  raw[6] = (double)p[7];
  for (int d = 0; d < 7; ++d) {
    double v;
    if (norm_type == 0) {  // bound: 2 (x - p01) / (p99 - p01 + 1e-8) - 1, clipped to [-1, 1]
      v = 2.0 * (raw[d] - a[d]) / (b[d] - a[d] + 1e-8) - 1.0;
      v = fmin(fmax(v, -1.0), 1.0);
    } else {               // gaussian: (x - mean) / (std + 1e-8)
      v = (raw[d] - a[d]) / (b[d] + 1e-8);
    }
    state[(int64_t)i * 7 + d] = round_bf16 ? __bfloat162float(__double2bfloat16(v)) : (float)v;
  }
}

// actions f32 [N, 7] (N = B * action_step) -> env actions f64 [N, 7]
__global__ void adapter_postprocess_kernel(const float* __restrict__ act, double* __restrict__ out, int N,
                                           const double* __restrict__ a, const double* __restrict__ b, int norm_type) {
  pdl_trigger();
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const float* p = act + (int64_t)i * 7;
  double raw[6];
  for (int d = 0; d < 6; ++d) {
    const double x = (double)p[d];
    raw[d] = norm_type == 0 ? (x - (-1.0)) / 2.0 * (b[d] - a[d]) + a[d]   // denormalize_bound, clip range [-1, 1]
                            : x * (b[d] + 1e-8) + a[d];                   // denormalize_gaussian
  }
  // euler2quat "sxyz"
  const double ai = raw[3] / 2.0, aj = raw[4] / 2.0, ak = raw[5] / 2.0;
  const double ci = cos(ai), si = sin(ai), cj = cos(aj), sj = sin(aj), ck = cos(ak), sk = sin(ak);
  const double cc = ci * ck, cs = ci * sk, sc = si * ck, ss = si * sk;
  double q[4] = {cj * cc + sj * ss, cj * sc - sj * cs, cj * ss + sj * cc, cj * cs - sj * sc};
  // quat2axangle
  double ax[3] = {1.0, 0.0, 0.0}, theta = 0.0;
  const double kEps = 2.220446049250313e-16;
  const double Nq = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
  if (!isfinite(Nq)) {
    theta = nan("");
  } else if (!(Nq < kEps * kEps)) {
    if (Nq != 1.0) {
      const double s = sqrt(Nq);
      for (int k = 0; k < 4; ++k) q[k] = q[k] / s;
    }
    const double len2 = q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
    if (!(len2 < (kEps * 3) * (kEps * 3))) {
      theta = 2.0 * acos(fmax(fmin(q[0], 1.0), -1.0));
      const double l = sqrt(len2);
      ax[0] = q[1] / l; ax[1] = q[2] / l; ax[2] = q[3] / l;
    }
  }
  double* o = out + (int64_t)i * 7;
  o[0] = raw[0]; o[1] = raw[1]; o[2] = raw[2];
  o[3] = ax[0] * theta; o[4] = ax[1] * theta; o[5] = ax[2] * theta;
  o[6] = 2.0 * ((double)p[6] > 0.5 ? 1.0 : 0.0) - 1.0;  // postprocess_gripper: 2 * (g > 0.5) - 1
}

}  // namespace
}  // namespace pi0

using namespace pi0;

extern "C" {

int pi0_adapter_resize_normalize(const uint8_t* images, float* out, int B, int H, int W, int out_h, int out_w,
                                 const int32_t* x_idx, const int16_t* x_w, int x_prec, const int32_t* y_idx,
                                 const int16_t* y_w, int y_prec, float rescale, float mean, float stdv, void* stream) {
  PI0_REQUIRE(images && out && x_idx && x_w && y_idx && y_w, "pi0_adapter_resize_normalize: null argument");
  PI0_REQUIRE(B >= 1 && H >= 1 && W >= 1 && out_h >= 1 && out_w >= 1, "pi0_adapter_resize_normalize: bad sizes");
  PI0_REQUIRE(x_prec >= 1 && x_prec <= 22 && y_prec >= 1 && y_prec <= 22, "pi0_adapter_resize_normalize: bad precision");
  const int64_t n = (int64_t)B * out_h * out_w;
  const int grid = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
  PI0_CHECK_CUDA(launch_pdl(adapter_resize_norm_kernel, dim3(grid), dim3(256), 0, (cudaStream_t)stream, 1, images, out, B,
                            H, W, out_h, out_w, (const int*)x_idx, (const short*)x_w, x_prec, (const int*)y_idx,
                            (const short*)y_w, y_prec, rescale, mean, stdv));
  ++g_launch_count;
  return 0;
}

int pi0_adapter_lanczos4_normalize(const uint8_t* images, float* out, int B, int H, int W, int out_h, int out_w,
                                   const int32_t* x_ofs, const int16_t* x_coef, const int32_t* y_ofs,
                                   const int16_t* y_coef, float rescale, float mean, float stdv, void* stream) {
  PI0_REQUIRE(images && out && x_ofs && x_coef && y_ofs && y_coef, "pi0_adapter_lanczos4_normalize: null argument");
  PI0_REQUIRE(B >= 1 && H >= 1 && W >= 1 && out_h >= 1 && out_w >= 1, "pi0_adapter_lanczos4_normalize: bad sizes");
  const int64_t n = (int64_t)B * out_h * out_w;
  const int grid = (int)((n + 127) / 128 < 148 * 16 ? (n + 127) / 128 : 148 * 16);
  PI0_CHECK_CUDA(launch_pdl(adapter_lanczos4_norm_kernel, dim3(grid), dim3(128), 0, (cudaStream_t)stream, 1, images, out,
                            B, H, W, out_h, out_w, (const int*)x_ofs, (const short*)x_coef, (const int*)y_ofs,
                            (const short*)y_coef, rescale, mean, stdv));
  ++g_launch_count;
  return 0;
}

int pi0_adapter_proprio(const void* eef, int eef_is_f64, float* state, int B, const double* stat_a, const double* stat_b,
                        int norm_type, int round_bf16, void* stream) {
  PI0_REQUIRE(eef && state && stat_a && stat_b && B >= 1, "pi0_adapter_proprio: bad argument");
  PI0_REQUIRE(norm_type == 0 || norm_type == 1, "pi0_adapter_proprio: norm_type %d (0 = bound, 1 = gaussian)", norm_type);
  const int grid = (B + 127) / 128;
  if (eef_is_f64)
    PI0_CHECK_CUDA(launch_pdl(adapter_proprio_kernel<double>, dim3(grid), dim3(128), 0, (cudaStream_t)stream, 1,
                              (const double*)eef, state, B, stat_a, stat_b, norm_type, round_bf16));
  else
    PI0_CHECK_CUDA(launch_pdl(adapter_proprio_kernel<float>, dim3(grid), dim3(128), 0, (cudaStream_t)stream, 1,
                              (const float*)eef, state, B, stat_a, stat_b, norm_type, round_bf16));
  ++g_launch_count;
  return 0;
}

int pi0_adapter_postprocess(const float* actions, double* out, int N, const double* stat_a, const double* stat_b,
                            int norm_type, void* stream) {
  PI0_REQUIRE(actions && out && stat_a && stat_b && N >= 1, "pi0_adapter_postprocess: bad argument");
  PI0_REQUIRE(norm_type == 0 || norm_type == 1, "pi0_adapter_postprocess: norm_type %d", norm_type);
  PI0_CHECK_CUDA(launch_pdl(adapter_postprocess_kernel, dim3((N + 127) / 128), dim3(128), 0, (cudaStream_t)stream, 1,
                            actions, out, N, stat_a, stat_b, norm_type));
  ++g_launch_count;
  return 0;
}

}  // extern "C"
