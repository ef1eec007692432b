// Denoise-step attention for rollout batches (B >= 4), as two wide launches instead of one CTA per (sample, head group):
//   decode_scores_kernel   one WARP per (sample, block of 16 keys), all 8 query heads at once (MQA: every head reads the
//                          same K, so K streams from HBM exactly once): scores = rnd(rnd(q.k) * 256^-0.5) -> bf16 scratch;
//   decode_pv_kernel       one CTA per (sample, head, quarter of the 256 dims): masked softmax of the head's S rows
//                          (fp32, finite -2.3819763e38 mask, probabilities rounded to bf16) and P.V against 64 rows of V^T.
// The one-CTA-per-(sample, head group) kernel (attn_decode.cu) chains ~8 dependent rounds of L2/HBM loads per CTA with
// at most one CTA per SM; at batch 64 it took ~45 us of the ~98 us of a denoise layer. Here every unit is one or two
// rounds of independent loads and there are thousands of units, so the launch is bandwidth- not latency-bound.
// Same arithmetic and rounding points as decode_attn_body (reference: eager_attention_forward,
// [mirror] src/model/mvla/mllm.py:265-317; suffix mask modeling_mvla.py:686-692,817-821).
#include "attn_split.cuh"

// Scratch of the allocation-free entry point (pi0_decode_attention_split_static, bottom of this file): the bf16 score
// matrix [B*8*S, roundup(P+S, 32)] that pi0_decode_attention_split takes from the caller. File scope, 4 MiB: every GPU that
// loads the library holds it until the process exits. In HOST code this name is only a stand-in for the symbol - never a
// device address; the host side resolves it with cudaGetSymbolAddress (once per device, below) and hands that pointer on.
constexpr size_t kSplitStaticScratchBytes = (size_t)4 << 20;
static __device__ __align__(256) unsigned char g_split_static_scratch[kSplitStaticScratchBytes];

namespace pi0 {

namespace {

struct SplitParams {
  const bf16* q;        // [B*S, 8*256]
  const bf16* kc;       // [B, kv_pitch, 256]
  const bf16* vt;       // [B, 256, vt_pitch]
  bf16* att;            // [B*S, 8*256]
  bf16* sc;             // scratch [B*8*S, kp]
  const uint8_t* pad;   // [B, P]
  const int* cum;       // [S]
  int B, S, P, kv_pitch, vt_pitch, kp;
};

// rows r = t * 8 + h (token t, head h), up to 48 rows = 3 m16 tiles
__global__ void __launch_bounds__(256) decode_scores_kernel(const SplitParams p) {
  pdl_trigger();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t4 = lane & 3;
  const int n_kb = p.kp >> 4;
  const int u = blockIdx.x * 8 + warp;
  pdl_wait();
  if (u >= p.B * n_kb) return;
  const int kb = u % n_kb, b = u / n_kb;
  const int n_keys = p.P + p.S, R = p.S * 8;
  const int MT = (R + 15) >> 4;
  const int j0 = kb * 16;
  int ja = j0 + g, jb = j0 + 8 + g;
  ja = ja < n_keys ? ja : n_keys - 1;
  jb = jb < n_keys ? jb : n_keys - 1;
  const bf16* kbase = p.kc + (int64_t)b * p.kv_pitch * 256;
  const bf16* qb = p.q + (int64_t)b * p.S * 2048;  // row r of the sample = 256 consecutive elements at r * 256
  float acc[3][2][4];
#pragma unroll
  for (int mi = 0; mi < 3; ++mi)
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[mi][nt][e] = 0.0f;
#pragma unroll
  for (int c0 = 0; c0 < 8; c0 += 4) {
    uint4 ka[4], kbv[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      ka[c] = ldcg_u4(kbase + (int64_t)ja * 256 + (c0 + c) * 32 + t4 * 8);
      kbv[c] = ldcg_u4(kbase + (int64_t)jb * 256 + (c0 + c) * 32 + t4 * 8);
    }
#pragma unroll
    for (int mi = 0; mi < 3; ++mi) {
      if (mi < MT) {
        const int rlo = mi * 16 + g, rhi = rlo + 8;
        uint4 lo[4], hi[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          lo[c] = rlo < R ? ldcg_u4(qb + (int64_t)rlo * 256 + (c0 + c) * 32 + t4 * 8) : make_uint4(0, 0, 0, 0);
          hi[c] = rhi < R ? ldcg_u4(qb + (int64_t)rhi * 256 + (c0 + c) * 32 + t4 * 8) : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          mma_bf16_16816(acc[mi][0], lo[c].x, hi[c].x, lo[c].y, hi[c].y, ka[c].x, ka[c].y);
          mma_bf16_16816(acc[mi][0], lo[c].z, hi[c].z, lo[c].w, hi[c].w, ka[c].z, ka[c].w);
          mma_bf16_16816(acc[mi][1], lo[c].x, hi[c].x, lo[c].y, hi[c].y, kbv[c].x, kbv[c].y);
          mma_bf16_16816(acc[mi][1], lo[c].z, hi[c].z, lo[c].w, hi[c].w, kbv[c].z, kbv[c].w);
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
          const int t = r >> 3, h = r & 7;
          bf16* srow = p.sc + ((int64_t)(b * 8 + h) * p.S + t) * p.kp;
#pragma unroll
          for (int nt = 0; nt < 2; ++nt)
            *reinterpret_cast<__nv_bfloat162*>(srow + j0 + nt * 8 + t4 * 2) = __floats2bfloat162_rn(
                bf16_round(acc[mi][nt][half * 2]) * 0.0625f, bf16_round(acc[mi][nt][half * 2 + 1]) * 0.0625f);
        }
      }
    }
  }
}

__global__ void __launch_bounds__(256) decode_pv_kernel(const SplitParams p) {
  pdl_trigger();
  extern __shared__ __align__(128) uint8_t ptile[];
  pdl_wait();
  attn_pv_unit(p, p.vt, ptile, blockIdx.x, threadIdx.x >> 5, threadIdx.x & 31);
}

}  // namespace

bool decode_attention_split_supported(int S, int P) { return S >= 1 && S <= 6 && P + S <= 32 * kMegaKeyChunks; }

int decode_attention_split(const void* q, const void* kcache, const void* vtcache, void* att, void* scratch,
                           const uint8_t* pad, const int* cum, int B, int S, int P, int kv_pitch, int vt_pitch,
                           cudaStream_t s) {
  PI0_REQUIRE(decode_attention_split_supported(S, P), "decode_attention_split: S=%d P=%d unsupported", S, P);
  PI0_REQUIRE(vt_pitch % 8 == 0 && vt_pitch >= P + S && kv_pitch >= P + S, "decode_attention_split: bad cache pitches");
  SplitParams p;
  p.q = (const bf16*)q; p.kc = (const bf16*)kcache; p.vt = (const bf16*)vtcache; p.att = (bf16*)att;
  p.sc = (bf16*)scratch; p.pad = pad; p.cum = cum;
  p.B = B; p.S = S; p.P = P; p.kv_pitch = kv_pitch; p.vt_pitch = vt_pitch; p.kp = (P + S + 31) / 32 * 32;
  const int n_units = B * (p.kp >> 4);
  PI0_CHECK_CUDA(launch_pdl(decode_scores_kernel, dim3((n_units + 7) / 8), dim3(256), 0, s, 1, p));
  ++g_launch_count;
  const size_t smem = 16 * (size_t)attn_p_pitch_bytes(p.kp);
  PI0_CHECK_CUDA(launch_pdl(decode_pv_kernel, dim3(B * 32), dim3(256), smem, s, 1, p));
  ++g_launch_count;
  return 0;
}

// Device address of g_split_static_scratch on the CURRENT device, resolved once per device (a symbol has one instance per
// GPU that loaded the module, each with its own address).
static int split_static_scratch(void** out) {
  static void* ptr[64] = {};
  int dev = -1;
  PI0_CHECK_CUDA(cudaGetDevice(&dev));
  PI0_REQUIRE(dev >= 0 && dev < 64, "decode_attention_split_static: device ordinal %d outside [0, 64)", dev);
  if (ptr[dev] == nullptr) {
    void* p = nullptr;
    PI0_CHECK_CUDA(cudaGetSymbolAddress(&p, g_split_static_scratch));
    PI0_REQUIRE(p != nullptr, "decode_attention_split_static: null symbol address");
    ptr[dev] = p;
  }
  *out = ptr[dev];
  return 0;
}

int decode_attention_split_static(const void* q, const void* kcache, const void* vtcache, void* att, const uint8_t* pad,
                                  const int* cum, int B, int S, int P, int kv_pitch, int vt_pitch, cudaStream_t s) {
  PI0_REQUIRE(q && kcache && vtcache && att && pad && cum && B >= 1, "decode_attention_split_static: bad argument");
  PI0_REQUIRE(decode_attention_split_supported(S, P), "decode_attention_split_static: S=%d P=%d unsupported", S, P);
  const size_t need = (size_t)B * 8 * S * ((size_t)(P + S + 31) / 32 * 32) * sizeof(bf16);
  PI0_REQUIRE(need <= kSplitStaticScratchBytes,
              "decode_attention_split_static: B=%d S=%d P=%d needs %zu bytes of score scratch, the built-in buffer holds %zu "
              "(use pi0_decode_attention_split with caller-owned scratch)", B, S, P, need, kSplitStaticScratchBytes);
  void* scratch = nullptr;
  PI0_TRY(split_static_scratch(&scratch));
  // nothing to clear: decode_scores_kernel writes every score element that decode_pv_kernel reads
  return decode_attention_split(q, kcache, vtcache, att, scratch, pad, cum, B, S, P, kv_pitch, vt_pitch, s);
}

}  // namespace pi0

extern "C" int pi0_decode_attention_split_static(const void* q, const void* kcache, const void* vtcache, void* att,
                                                 const uint8_t* pad, const int32_t* cum, int B, int S, int P, int kv_pitch,
                                                 int vt_pitch, void* stream) {
  return pi0::decode_attention_split_static(q, kcache, vtcache, att, pad, cum, B, S, P, kv_pitch, vt_pitch,
                                            (cudaStream_t)stream);
}

extern "C" long long pi0_decode_attention_split_static_max_bytes(void) { return (long long)kSplitStaticScratchBytes; }
