// tcgen05 / TMEM / TMA GEMM for sm_100a:  C[z][M,N] = A[z][M,K] . W[z][N,K]^T, bf16 in, fp32 accumulate.
//
// Persistent, warp-specialised: warp 0 = TMA producer, warp 1 = MMA issuer (+ TMEM owner),
// warps 2..9 = epilogue (two warps per TMEM lane quadrant, each taking half of the tile's columns). Operand tiles are 128 x 64 (A) and BN x 64 (W),
// K-major with the 128-byte swizzle that TMA writes and the UMMA smem descriptor reads.
// The fp32 accumulator is double-buffered in TMEM (2 x BN columns) so the epilogue of tile i overlaps
// the main loop of tile i+1. Tiles are visited in groups of 8 M-tiles per N sweep to keep both operand
// panels L2-resident. Replaces every nn.Linear / torch.matmul on the reference's hot path
// (reference: src/model/mvla/mllm.py:216-242, 299-311).
#include "gemm.cuh"

namespace pi0 {

// Optional in-kernel timeline (development aid, compiled in with -DPI0_TIMELINE): CTA 0 stamps %globaltimer at the
// phases of one launch; pi0_debug_timeline() copies the stamps back.
#ifdef PI0_TIMELINE
__device__ unsigned long long g_timeline[16];
#define PI0_STAMP(i) do { if (blockIdx.x == 0) g_timeline[i] = global_timer_ns(); } while (0)
__device__ unsigned long long g_pair_timeline[2][16][4];  // [rank][tile][mma first issue, last commit, tfull seen, drained]
#define PAIR_STAMP(r, t, k) do { if (pair == 0 && (t) < 16) g_pair_timeline[r][t][k] = global_timer_ns(); } while (0)
#else
#define PI0_STAMP(i) do { } while (0)
#define PAIR_STAMP(r, t, k) do { } while (0)
#endif

constexpr int kBM = 128;
constexpr int kBK = 64;
constexpr int kGroupM = 8;
constexpr int kTcSegThreads = 384;  // gemm_tc_seg_kernel: warpgroup 0 (TMA, MMA, 2 idle warps) + 8 epilogue warps
constexpr int kTcThreads = 320;  // TMA warp + MMA warp + 8 epilogue warps (2 per TMEM lane quadrant)
// TMA-store staging: every epilogue warp owns two 32-row x 64-byte (one 32-column bf16 chunk) buffers, written in the
// 64-byte-swizzle pattern of the output tensor map (16-byte piece g of row r lands at piece g ^ ((r >> 1) & 3): the
// row-per-lane st.shared.v4 is then bank-conflict free) and drained by cp.async.bulk.tensor stores.
constexpr uint32_t kStoreBufBytes = 32 * 64;
constexpr uint32_t kStoreStageBytes = 8 * 2 * kStoreBufBytes;

template <int BN>
struct TcCfg {
  static constexpr uint32_t kABytes = kBM * kBK * 2;
  static constexpr uint32_t kBBytes = BN * kBK * 2;
  static constexpr uint32_t kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (BN == 256) ? 4 : (BN == 128) ? 6 : 8;
  static constexpr int kTmemCols = (2 * BN < 32) ? 32 : 2 * BN;
  static constexpr uint32_t kSmemBytes = kStages * kStageBytes + 1024 + 256;
  // persistent kernel: + the epilogue warps' TMA-store staging buffers behind the (1 KiB padded) barrier block
  static constexpr uint32_t kSmemBytesStaged = kStages * kStageBytes + 1024 + 1024 + kStoreStageBytes;
};

struct TcShape {
  int M, N, K;
  int nz0, nz1;
  int tiles_m, tiles_n;
  int total_tiles;
  int splits;        // split-K: CTAs (= cluster size) per output tile, 1 = none
  int kb_per_split;  // k-blocks (of kBK) per split
  uint32_t stage_tx; // bytes one pipeline stage receives: A box (rows clipped to M for a single short M tile) + W box
  int tma_store;     // 1: EPI_STORE / EPI_GEGLU output goes through shared memory and tensor-map stores (tmC valid)
  int super_m, super_n;  // second-level tile blocking for wide problems (tiles), see tile_coords
  int w_prefetch;    // 1: W is a constant weight matrix - the TMA warp requests the W boxes of the first ring stages BEFORE the
                     // programmatic-dependency wait, so the weight fetch (cold HBM) overlaps the previous kernel's tail
};

// Tile order of the persistent kernels. Small problems: groups of kGroupM M-tiles, M fastest inside a group, so that the
// CTAs in flight share a W tile column-wise and a handful of A tiles. Wide problems (prefill gate/up: 82 x 128 tiles of
// 256 x 256 at batch 64) add a second level: super-blocks of kSuperM x kSuperN tiles whose A block (32 x 256 rows x K) and W
// block (32 x 256 rows x K) are 33.5 MB each at K = 2048 and stay resident in the 126 MB L2 while the block's 1024 tiles
// run; the A block is reused by the next super-block of the same rows. With the one-level order the whole W panel (134 MB,
// larger than L2) was re-streamed from HBM for every group of 8 M-tiles: 3.0 GB of DRAM reads per launch against 0.22 GB of
// operands (ncu, profiles/r02_gateup_pair_b64_full_raw.csv).
// Defaults from a sweep on the batch-64 gate/up GEMM (tools/sweep_super_tiles.sh, DRAM read per launch / kernel time):
// one level 1.9-3.0 GB / 1.81 ms; 32 x 32 1.07 GB / 1.766 ms; 41 x 32 0.96 GB / 1.764 ms; 82 x 16 0.93 GB / 1.756 ms.
// => N panels of 16 tiles (W block 16.8 MB at K = 2048, resident) and an A block of ~84 MB worth of rows streamed per
// panel and largely kept by the L2 from one panel to the next. PI0_SUPER_M / PI0_SUPER_N override (development knobs).
constexpr int kSuperN = 16;
constexpr long long kSuperABytes = 84ll << 20;
static void set_super(TcShape& sh, int tile_rows) {
  static int sm = getenv("PI0_SUPER_M") ? atoi(getenv("PI0_SUPER_M")) : 0;
  static int sn = getenv("PI0_SUPER_N") ? atoi(getenv("PI0_SUPER_N")) : 0;
  long long auto_m = kSuperABytes / ((long long)tile_rows * sh.K * 2);
  if (auto_m < kGroupM) auto_m = kGroupM;
  if (auto_m > (1 << 20)) auto_m = 1 << 20;
  sh.super_m = sm > 0 ? sm : (int)auto_m;
  sh.super_n = sn > 0 ? sn : kSuperN;
}
__device__ __forceinline__ void tile_coords(const TcShape& sh, int idx, int& z, int& tm, int& tn) {
  int per_z = sh.tiles_m * sh.tiles_n;
  z = idx / per_z;
  int r = idx - z * per_z;
  int first_sm = 0, sm_sz = sh.tiles_m, first_n = 0, sn_sz = sh.tiles_n;
  if (sh.tiles_n > sh.super_n && sh.tiles_m > kGroupM) {
    const int sm_span = sh.super_m * sh.tiles_n;
    const int sg = r / sm_span;
    first_sm = sg * sh.super_m;
    sm_sz = min(sh.tiles_m - first_sm, sh.super_m);
    r -= sg * sm_span;
    const int sn_span = sm_sz * sh.super_n;
    const int ng = r / sn_span;
    first_n = ng * sh.super_n;
    sn_sz = min(sh.tiles_n - first_n, sh.super_n);
    r -= ng * sn_span;
  }
  int span = kGroupM * sn_sz;
  int g = r / span;
  int first_m = g * kGroupM;
  int gsz = min(sm_sz - first_m, kGroupM);
  int rr = r - g * span;
  tm = first_sm + first_m + rr % gsz;
  tn = first_n + rr / gsz;
}

// Epilogue of one 128-row accumulator tile: TMEM lanes of this warp's quadrant -> registers -> fused epilogue -> global.
// `part` selects which of the NPARTS slices of the tile's column chunks this warp drains (NPARTS warps share a lane
// quadrant).
// One 32-row x 32-column bf16 chunk of this warp (lane = row) -> swizzled staging buffer -> TMA store at (x, y, z0, z1).
// Rows >= M and columns >= N are clipped by the tensor map. `sbuf` alternates between the warp's two buffers: before a
// buffer is rewritten, the store issued from it two chunks ago must have finished READING it (wait_group.read 1).
__device__ __forceinline__ void stage_and_store(const CUtensorMap* tmC, uint8_t* stage2, int& sbuf, const float* v, int x,
                                                int y, int z0, int z1) {
  const int lane = threadIdx.x & 31;
  uint8_t* buf = stage2 + sbuf * kStoreBufBytes;
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
    tma_store_4d(tmC, buf, x, y, z0, z1);
    bulk_commit();
  }
}

template <int BN, int EPI, int NPARTS = 2>
__device__ __forceinline__ void drain_tile(const TcShape& sh, const EpiParams& ep, uint32_t taddr, int row, bool row_ok,
                                           int tn, int z0, int z1, int part, const CUtensorMap* tmC, uint8_t* stage2,
                                           int& sbuf) {
  const int n0 = tn * BN;
  const int row0 = row - (int)(threadIdx.x & 31);
  if constexpr (EPI == EPI_STORE || EPI == EPI_EULER) {
    const int64_t zoff = (int64_t)z0 * ep.out_z0 + (int64_t)z1 * ep.out_z1;
    constexpr int kTotal = BN / 32;
    constexpr int kPer = (kTotal + NPARTS - 1) / NPARTS;  // chunks per part
#pragma unroll 1
    for (int c = part * kPer; c < min(kTotal, (part + 1) * kPer); ++c) {
      const int col0 = n0 + c * 32;
      if (col0 >= sh.N) break;
      float v[32];
      tmem_ld_32x32(taddr + c * 32, v);
      tmem_ld_wait();
      if constexpr (EPI == EPI_STORE) {
        if (sh.tma_store) {
          if (row_ok) epi_store_chunk<bf16>(ep, row, col0, sh.N, zoff, v, false);
          stage_and_store(tmC, stage2, sbuf, v, col0, row0, z0, z1);
        } else if (row_ok) {
          epi_store_chunk<bf16>(ep, row, col0, sh.N, zoff, v);
        }
      } else {
        if (row_ok) epi_euler_chunk<bf16>(ep, row, col0, sh.N, v);
      }
    }
  } else if constexpr (EPI == EPI_GEGLU) {
    // tile columns [0, BN/2) are gate rows, [BN/2, BN) the matching up rows (interleaved at pack time)
    const int n_out = sh.N / 2;
    constexpr int kTotal = BN / 64;  // (gate, up) chunk pairs
    constexpr int kPer = (kTotal + NPARTS - 1) / NPARTS;
#pragma unroll 1
    for (int c = part * kPer; c < min(kTotal, (part + 1) * kPer); ++c) {
      const int out_col0 = tn * (BN / 2) + c * 32;
      if (out_col0 >= n_out) break;
      float g[32], u[32];
      tmem_ld_32x32(taddr + c * 32, g);
      tmem_ld_32x32(taddr + BN / 2 + c * 32, u);
      tmem_ld_wait();
      if (sh.tma_store) {
        epi_geglu_math<bf16>(g, u);
        stage_and_store(tmC, stage2, sbuf, g, out_col0, row0, z0, z1);
      } else if (row_ok) {
        epi_geglu_chunk<bf16>(ep, row, out_col0, n_out, g, u);
      }
    }
  } else if constexpr (EPI == EPI_ROPE) {
    static_assert(EPI != EPI_ROPE || BN == 256, "RoPE epilogue needs one 256-wide head per tile");
    const int which = tn < ep.n_q_heads ? 0 : (tn == ep.n_q_heads ? 1 : 2);
    constexpr int kPer = (4 + NPARTS - 1) / NPARTS;
#pragma unroll 1
    for (int c = part * kPer; c < min(4, (part + 1) * kPer); ++c) {
      float x1[32], x2[32];
      tmem_ld_32x32(taddr + c * 32, x1);
      tmem_ld_32x32(taddr + 128 + c * 32, x2);
      tmem_ld_wait();
      if (row_ok) epi_rope_chunk<bf16>(ep, row, which, tn, c, x1, x2);
    }
  }
}

// NormArgs: Gemma's RMSNorm fused into the A operand. The A tiles hold the RAW residual stream; the eight epilogue warps - idle
// during the main loop when a CTA owns a single tile - rewrite every landed A tile in shared memory as
// rnd((a * rstd[row]) * gain[k]) (the arithmetic of rmsnorm_warp_kernel) between the TMA's arrival and the MMA.
struct NormArgs {
  const float* gain;    // [K] fp32, 1 + w
  const float* rowsq;   // [M][slots] per-chunk sums of squares of the rows of A (EpiParams.rowsq_out of the producing GEMM)
  int slots;
  float eps;
};

// executed by the 256 threads of warps 2..9 (t = threadIdx.x - 64) for k-blocks [kb0, kb1) of M-tile tm
template <int BN>
__device__ __forceinline__ void norm_a_operand(uint8_t* smem, uint64_t* full_bar, uint64_t* ready_bar, float* rstd_s,
                                               const NormArgs& na, const TcShape& sh, int tm, int kb0, int kb1, int t) {
  using Cfg = TcCfg<BN>;
  const int lane = threadIdx.x & 31;
  {
    // rstd of the tile's rows: two threads per row, each sums half of the row's slots from independent 16-byte loads (a
    // scalar loop over the slots is a chain of dependent L2 round trips: measured +18 us per launch), fixed order
    const int r = t >> 1, hf = t & 1;
    const int row = tm * kBM + r;
    const int n4 = na.slots >> 3;          // float4 groups per half (slots % 8 == 0)
    float q = 0.0f;
    if (row < sh.M) {
      const float4* rq = reinterpret_cast<const float4*>(na.rowsq + (int64_t)row * na.slots) + hf * n4;
      float4 v[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = i < n4 ? __ldcg(rq + i) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll
      for (int i = 0; i < 8; ++i) q += (v[i].x + v[i].y) + (v[i].z + v[i].w);
      for (int i = 8; i < n4; ++i) { const float4 w = __ldcg(rq + i); q += (w.x + w.y) + (w.z + w.w); }
    }
    q += __shfl_xor_sync(0xffffffffu, q, 1);
    if (hf == 0) rstd_s[r] = row < sh.M ? rsqrtf(q / (float)sh.K + na.eps) : 0.0f;
  }
  asm volatile("bar.sync 2, 256;" ::: "memory");
  // piece i = t + 256 * qq of the [128 rows][8 pieces] tile: row (t >> 3) + 32 qq, physical piece t & 7. Under the 128-byte
  // swizzle its logical k offset (ph ^ (row & 7)) * 8 is the SAME for all four pieces of a thread, so a thread needs eight
  // gains per k-block: they are fetched one k-block ahead (a first touch of the gain vector costs an L2 round trip per
  // k-block otherwise: measured +10 us on a 16-k-block GEMM).
  const int ph = t & 7, r0 = t >> 3;
  const int kofs = (ph ^ (r0 & 7)) << 3;
  float4 g0 = __ldg(reinterpret_cast<const float4*>(na.gain + kb0 * kBK + kofs));
  float4 g1 = __ldg(reinterpret_cast<const float4*>(na.gain + kb0 * kBK + kofs + 4));
  int stage = 0;
  uint32_t phase = 0;
  for (int kb = kb0; kb < kb1; ++kb) {
    const int kn = (kb + 1 < kb1 ? kb + 1 : kb) * kBK + kofs;
    const float4 n0 = __ldg(reinterpret_cast<const float4*>(na.gain + kn));
    const float4 n1 = __ldg(reinterpret_cast<const float4*>(na.gain + kn + 4));
    mbar_wait(&full_bar[stage], phase);
    uint8_t* sa = smem + stage * Cfg::kStageBytes;
#pragma unroll
    for (int qq = 0; qq < (kBM * 8) / 256; ++qq) {
      const int r = r0 + 32 * qq;
      uint4* pp = reinterpret_cast<uint4*>(sa + r * 128 + ph * 16);
      const uint4 x = *pp;
      const float rs = rstd_s[r];
      const __nv_bfloat162* xh = reinterpret_cast<const __nv_bfloat162*>(&x);
      const float2 a0 = __bfloat1622float2(xh[0]), a1 = __bfloat1622float2(xh[1]);
      const float2 a2 = __bfloat1622float2(xh[2]), a3 = __bfloat1622float2(xh[3]);
      uint4 o;
      o.x = bf16x2_bits((a0.y * rs) * g0.y, (a0.x * rs) * g0.x);
      o.y = bf16x2_bits((a1.y * rs) * g0.w, (a1.x * rs) * g0.z);
      o.z = bf16x2_bits((a2.y * rs) * g1.y, (a2.x * rs) * g1.x);
      o.w = bf16x2_bits((a3.y * rs) * g1.w, (a3.x * rs) * g1.z);
      *pp = o;
    }
    fence_proxy_async_smem();   // generic-proxy writes -> visible to the tensor core's async-proxy reads
    __syncwarp();
    if (lane == 0) mbar_arrive(&ready_bar[stage]);
    g0 = n0;
    g1 = n1;
    if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
  }
}

// NORM (fused RMSNorm of the A operand, norm_a_operand): only launched with at most ONE tile per CTA, so that the epilogue warps
// are free during the tile's main loop.
template <int BN, int EPI, bool NORM = false>
__global__ void __launch_bounds__(kTcThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmC, const TcShape sh,
               const EpiParams ep, const NormArgs na) {
  using Cfg = TcCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  uint64_t* ready_bar = tempty_bar + 3;                                  // NORM: A tile of the stage normalised
  float* rstd_s = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(full_bar) + 256);   // NORM: [128]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  pdl_trigger();

  if (threadIdx.x == 0) {
#ifdef PI0_TIMELINE
    if (blockIdx.x == 0) g_timeline[8] = g_timeline[7];
#endif
    PI0_STAMP(0);
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], 8);
    }
    if constexpr (NORM)
      for (int s = 0; s < Cfg::kStages; ++s) mbar_init(&ready_bar[s], 8);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int num_k = (sh.K + kBK - 1) / kBK;
  // weights do not depend on the previous kernel: the W boxes of the CTA's first tile, first ring pass, are requested now
  const int pre = sh.w_prefetch && (int)blockIdx.x < sh.total_tiles ? min(num_k, Cfg::kStages) : 0;
  if (pre > 0 && warp == 0 && elect_one()) {
    int z, tm, tn;
    tile_coords(sh, blockIdx.x, z, tm, tn);
    for (int kb = 0; kb < pre; ++kb) {
      mbar_arrive_expect_tx(&full_bar[kb], sh.stage_tx);
      tma_load_4d(smem + kb * Cfg::kStageBytes + Cfg::kABytes, &tmB, &full_bar[kb], kb * kBK, tn * BN, z % sh.nz0, z / sh.nz0);
    }
  }
  pdl_wait();  // everything above overlapped the previous kernel's tail; activations are touched only below
  if (threadIdx.x == 0) PI0_STAMP(1);

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int idx = blockIdx.x; idx < sh.total_tiles; idx += gridDim.x) {
        int z, tm, tn;
        tile_coords(sh, idx, z, tm, tn);
        const int z0 = z % sh.nz0, z1 = z / sh.nz0;
        for (int kb = 0; kb < num_k; ++kb) {
#ifdef PI0_TIMELINE
          long long c0 = clock64();
#endif
          mbar_wait(&empty_bar[stage], phase ^ 1);
#ifdef PI0_TIMELINE
          long long c1 = clock64();
#endif
          uint8_t* sa = smem + stage * Cfg::kStageBytes;
          uint8_t* sb = sa + Cfg::kABytes;
          const bool prefetched = idx == (int)blockIdx.x && kb < pre;   // barrier armed and W requested before the wait
          if (!prefetched) mbar_arrive_expect_tx(&full_bar[stage], sh.stage_tx);
#ifdef PI0_TIMELINE
          long long c2 = clock64();
#endif
          tma_load_4d(sa, &tmA, &full_bar[stage], kb * kBK, tm * kBM, z0, z1);
#ifdef PI0_TIMELINE
          long long c3 = clock64();
#endif
          if (!prefetched) tma_load_4d(sb, &tmB, &full_bar[stage], kb * kBK, tn * BN, z0, z1);
#ifdef PI0_TIMELINE
          long long c4 = clock64();
          if (blockIdx.x == 0 && idx == 0 && kb == 2) {
            g_timeline[14] = (unsigned long long)(((c1 - c0) << 48) | ((c2 - c1) << 32) | ((c3 - c2) << 16) | (c4 - c3));
          }
#endif
          if (idx == (int)blockIdx.x && kb == 0) PI0_STAMP(2);
          if (idx == (int)blockIdx.x && kb == Cfg::kStages - 1) PI0_STAMP(13);
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer --------------------------------
    constexpr uint32_t idesc = umma_idesc_bf16(kBM, BN);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int idx = blockIdx.x; idx < sh.total_tiles; idx += gridDim.x) {
      mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * BN;
      for (int kb = 0; kb < num_k; ++kb) {
#ifdef PI0_TIMELINE
        long long m0 = clock64();
#endif
        mbar_wait(NORM ? &ready_bar[stage] : &full_bar[stage], phase);
#ifdef PI0_TIMELINE
        long long m1 = clock64();
#endif
        tc_fence_after();
#ifdef PI0_TIMELINE
        long long m2 = clock64();
#endif
        if (elect_one()) {
          if (idx == (int)blockIdx.x && kb == 0) PI0_STAMP(3);
          if (idx == (int)blockIdx.x && kb == 8) PI0_STAMP(9);
          if (idx == (int)blockIdx.x && kb == 16) PI0_STAMP(10);
          if (idx == (int)blockIdx.x && kb == 32) PI0_STAMP(11);
          const uint32_t sa = smem_u32(smem + stage * Cfg::kStageBytes);
          const uint64_t da = umma_desc_k_sw128(sa);
          const uint64_t db = umma_desc_k_sw128(sa + Cfg::kABytes);
#ifdef PI0_EXP_ONE_MMA
#pragma unroll
          for (int k = 0; k < 1; ++k) {
#else
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {
#endif
            // advancing 16 bf16 (32 B) along K inside the 128 B swizzle row: +2 in the (>>4) address field
            umma_bf16(tmem_d, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb | k) != 0 ? 1u : 0u);
          }
#ifdef PI0_TIMELINE
          long long m3 = clock64();
#endif
          umma_commit(&empty_bar[stage]);
#ifdef PI0_TIMELINE
          long long m4 = clock64();
          if (blockIdx.x == 0 && idx == 0 && kb == 20) {
            g_timeline[15] = (unsigned long long)(((m1 - m0) << 48) | ((m2 - m1) << 32) | ((m3 - m2) << 16) | (m4 - m3));
          }
#endif
          if (kb == num_k - 1) { umma_commit(&tfull_bar[acc]); PI0_STAMP(4); }
        }
        __syncwarp();
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  } else {
    // ------------------------------ epilogue ----------------------------------
    const int quad = warp & 3;          // TMEM lanes [32*quad, 32*quad+32) are the ones this warp may read
    const int half = (warp - 2) >> 2;   // which half of the tile's columns this warp drains
    uint8_t* stage2 = reinterpret_cast<uint8_t*>(full_bar) + 1024 + (warp - 2) * 2 * kStoreBufBytes;
    int sbuf = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int idx = blockIdx.x; idx < sh.total_tiles; idx += gridDim.x) {
      int z, tm, tn;
      tile_coords(sh, idx, z, tm, tn);
      const int z0 = z % sh.nz0, z1 = z / sh.nz0;
      const int row = tm * kBM + quad * 32 + lane;
      const bool row_ok = row < sh.M;
      if constexpr (NORM) norm_a_operand<BN>(smem, full_bar, ready_bar, rstd_s, na, sh, tm, 0, num_k, (int)threadIdx.x - 64);
#ifdef PI0_TIMELINE
      if (threadIdx.x == 64 && idx == (int)blockIdx.x && num_k >= Cfg::kStages) {
        mbar_wait(&full_bar[Cfg::kStages - 1], 0);  // when does the LAST stage of the first ring pass land?
        PI0_STAMP(12);
      }
      __syncwarp();
#endif
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      if (threadIdx.x == 64) PI0_STAMP(5);
      const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + acc * BN;
      drain_tile<BN, EPI>(sh, ep, taddr, row, row_ok, tn, z0, z1, half, &tmC, stage2, sbuf);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[acc]);
      if (threadIdx.x == 64) PI0_STAMP(6);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
    if (sh.tma_store && lane == 0) bulk_wait_all();  // this warp's output stores have landed before the CTA exits
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
  if (threadIdx.x == 32) PI0_STAMP(7);
}

// ---------------------------------------------------------------------------------------------
// fp32-output variant with SEGMENTED accumulation, for the split-bf16 GEMM of the fp32 regime (gemm_x3.cu).
// The tensor core adds into its fp32 accumulator with truncation, so one accumulator run over a long K drifts towards zero
// by ~2^-24 per MMA step (measured: mean relative error 1.8e-5 at K' = 6 x 16384 against 2e-6 for an fp32 FMA chain). Here
// the MMA warp closes the TMEM accumulator every `seg_kb` k-blocks (the two accumulator buffers alternate, so the next
// segment is issued while the previous one drains) and the epilogue warps add the segments into REGISTERS with round-to-
// nearest fp32 adds: 4 chunks x 32 columns per thread. The fused epilogue then runs from the registers, in fp32 storage.
// ---------------------------------------------------------------------------------------------
template <int BN, int EPI>
__global__ void __launch_bounds__(kTcSegThreads, 1)
gemm_tc_seg_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const TcShape sh,
                   const EpiParams ep, const int seg_kb, float* __restrict__ partial, const int ldp) {
  using Cfg = TcCfg<BN>;
  constexpr bool kPaired = EPI == EPI_GEGLU || EPI == EPI_ROPE;
  static_assert(!kPaired || BN == 256, "paired epilogues need one 256-wide tile");
  constexpr int kChunks = BN / 32;                              // 32-column chunks of the tile
  constexpr int kSlots = kPaired ? 4 : (kChunks + 1) / 2;       // chunks per epilogue warp (two warps share a lane quadrant)
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  pdl_trigger();
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], 8);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int num_k = (sh.K + kBK - 1) / kBK;
  const int total_items = sh.total_tiles * sh.splits;
  pdl_wait();

  // 12 warps: warpgroup 0 = TMA warp, MMA warp and two idle warps, warpgroups 1-2 = the eight epilogue warps. A thread of a
  // 384-thread CTA gets 168 registers, the epilogue needs 128 accumulators + a 32-column TMEM chunk: warpgroup 0 hands its
  // registers over (128 x 88 + 256 x 208 = the 64512 the CTA was launched with).
  // (the low-register group leaves through its own return so that the register bound of each side is unambiguous; the
  // CTA-wide rendezvous before the TMEM release is a named barrier both sides arrive on)
  if (warp < 4) {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 88;");
  if (warp == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int idx = blockIdx.x; idx < total_items; idx += gridDim.x) {
        int z, tm, tn;
        tile_coords(sh, idx / sh.splits, z, tm, tn);
        const int z0 = z % sh.nz0, z1 = z / sh.nz0;
        const int kb_lo = (idx % sh.splits) * sh.kb_per_split, kb_hi = min(num_k, kb_lo + sh.kb_per_split);
        for (int kb = kb_lo; kb < kb_hi; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * Cfg::kStageBytes;
          mbar_arrive_expect_tx(&full_bar[stage], sh.stage_tx);
          tma_load_4d(sa, &tmA, &full_bar[stage], kb * kBK, tm * kBM, z0, z1);
          tma_load_4d(sa + Cfg::kABytes, &tmB, &full_bar[stage], kb * kBK, tn * BN, z0, z1);
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = umma_idesc_bf16(kBM, BN);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int idx = blockIdx.x; idx < total_items; idx += gridDim.x) {
      const int kb_lo = (idx % sh.splits) * sh.kb_per_split, kb_hi = min(num_k, kb_lo + sh.kb_per_split);
      for (int kb0 = kb_lo; kb0 < kb_hi; kb0 += seg_kb) {
        const int kb1 = min(kb_hi, kb0 + seg_kb);
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t sa = smem_u32(smem + stage * Cfg::kStageBytes);
            const uint64_t da = umma_desc_k_sw128(sa);
            const uint64_t db = umma_desc_k_sw128(sa + Cfg::kABytes);
#pragma unroll
            for (int k = 0; k < kBK / 16; ++k)
              umma_bf16(tmem_d, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb != kb0 || k != 0) ? 1u : 0u);
            umma_commit(&empty_bar[stage]);
            if (kb == kb1 - 1) umma_commit(&tfull_bar[acc]);
          }
          __syncwarp();
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  }
  tc_fence_before();
  asm volatile("bar.sync 2, %0;" ::"n"(kTcSegThreads) : "memory");
  if (warp == 1) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
  return;
  }

  asm volatile("setmaxnreg.inc.sync.aligned.u32 208;");
  {
    const int quad = warp & 3;
    const int part = (warp - 4) >> 2;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int idx = blockIdx.x; idx < total_items; idx += gridDim.x) {
      int z, tm, tn;
      tile_coords(sh, idx / sh.splits, z, tm, tn);
      const int z0 = z % sh.nz0, z1 = z / sh.nz0;
      const int sp = idx % sh.splits;
      const int kb_lo = sp * sh.kb_per_split, kb_hi = min(num_k, kb_lo + sh.kb_per_split);
      const int row = tm * kBM + quad * 32 + lane;
      const bool row_ok = row < sh.M;
      const int n0 = tn * BN;
      float r[kSlots][32];
      // TMEM column offset of slot s inside the tile, -1 if this warp has no such chunk / it lies beyond N
      auto slot_off = [&](int s) -> int {
        if constexpr (kPaired) {
          const int c = part * 2 + (s >> 1);      // pair index 0..3; halves of a pair are 128 columns apart
          return (s & 1) * 128 + c * 32;
        } else {
          const int c = part * kSlots + s;
          return (c < kChunks && n0 + c * 32 < sh.N) ? c * 32 : -1;
        }
      };
      for (int kb0 = kb_lo; kb0 < kb_hi; kb0 += seg_kb) {
        mbar_wait(&tfull_bar[acc], acc_phase);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + acc * BN;
#pragma unroll
        for (int s = 0; s < kSlots; ++s) {
          const int off = slot_off(s);
          if (off >= 0) {
            float v[32];
            tmem_ld_32x32(taddr + off, v);
            tmem_ld_wait();
            if (kb0 == kb_lo) {
#pragma unroll
              for (int j = 0; j < 32; ++j) r[s][j] = v[j];
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) r[s][j] += v[j];
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty_bar[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
      if (partial != nullptr) {
        // split-K: this CTA's k-slice of the tile goes to the fp32 workspace [split][z][row][ldp]; x3_reduce_kernel sums the
        // slices in split order and runs the epilogue
        if (row_ok) {
          float* prow = partial + (((int64_t)sp * (sh.nz0 * sh.nz1) + z) * sh.M + row) * ldp + n0;
#pragma unroll
          for (int s = 0; s < kSlots; ++s) {
            const int off = slot_off(s);
            if (off >= 0 && n0 + off < ldp) {
#pragma unroll
              for (int j = 0; j < 32; j += 4)
                *reinterpret_cast<float4*>(prow + off + j) = make_float4(r[s][j], r[s][j + 1], r[s][j + 2], r[s][j + 3]);
            }
          }
        }
      } else if (row_ok) {
        if constexpr (EPI == EPI_STORE || EPI == EPI_EULER) {
          const int64_t zoff = (int64_t)z0 * ep.out_z0 + (int64_t)z1 * ep.out_z1;
#pragma unroll
          for (int s = 0; s < kSlots; ++s) {
            const int off = slot_off(s);
            if (off >= 0) {
              if constexpr (EPI == EPI_STORE) epi_store_chunk<float>(ep, row, n0 + off, sh.N, zoff, r[s]);
              else epi_euler_chunk<float>(ep, row, n0 + off, sh.N, r[s]);
            }
          }
        } else if constexpr (EPI == EPI_GEGLU) {
          const int n_out = sh.N / 2;
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const int out_col0 = tn * (BN / 2) + (part * 2 + i) * 32;
            if (out_col0 < n_out) epi_geglu_chunk<float>(ep, row, out_col0, n_out, r[2 * i], r[2 * i + 1]);
          }
        } else {
          const int which = tn < ep.n_q_heads ? 0 : (tn == ep.n_q_heads ? 1 : 2);
#pragma unroll
          for (int i = 0; i < 2; ++i) epi_rope_chunk<float>(ep, row, which, tn, part * 2 + i, r[2 * i], r[2 * i + 1]);
        }
      }
    }
  }

  tc_fence_before();
  asm volatile("bar.sync 2, %0;" ::"n"(kTcSegThreads) : "memory");
}


// Second pass of the split-K forms (gemm_tc_seg_kernel, skinny_f32_kernel): v = sum over the k slices of the fp32 partials of
// one (row, 32-column chunk [pair]), then the ordinary fused epilogue in fp32 storage. One WARP per unit: lane (g, f) loads
// float4 f of the chunk from slices g, g + 4, g + 8, ... (all loads independent: one memory round trip whatever S is - a
// thread-per-unit loop over S took 15 us at S = 16 with a few hundred threads in the grid), the four slice groups are combined
// with shuffles in a fixed order (deterministic), and lane 0 runs the epilogue on the gathered 32 values.
__device__ __forceinline__ void x3_fetch_chunk(const float* __restrict__ partial, int S, int64_t slice_stride, int64_t off,
                                               int lane, float* v) {
  const int g = lane >> 3, f = lane & 7;
  float4 acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
  const float* p = partial + off + f * 4;
  for (int s0 = g; s0 < S; s0 += 16) {   // up to four independent loads per pass
    float4 t[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int s = s0 + 4 * u;
      t[u] = s < S ? __ldcg(reinterpret_cast<const float4*>(p + (int64_t)s * slice_stride)) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) { acc.x += t[u].x; acc.y += t[u].y; acc.z += t[u].z; acc.w += t[u].w; }
  }
#pragma unroll
  for (int sh = 16; sh >= 8; sh >>= 1) {
    acc.x += __shfl_xor_sync(0xffffffffu, acc.x, sh);
    acc.y += __shfl_xor_sync(0xffffffffu, acc.y, sh);
    acc.z += __shfl_xor_sync(0xffffffffu, acc.z, sh);
    acc.w += __shfl_xor_sync(0xffffffffu, acc.w, sh);
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    v[4 * j] = __shfl_sync(0xffffffffu, acc.x, j);
    v[4 * j + 1] = __shfl_sync(0xffffffffu, acc.y, j);
    v[4 * j + 2] = __shfl_sync(0xffffffffu, acc.z, j);
    v[4 * j + 3] = __shfl_sync(0xffffffffu, acc.w, j);
  }
}

template <int EPI>
__global__ void __launch_bounds__(256) x3_reduce_kernel(const float* __restrict__ partial, int S, int nz0, int nz1, int M,
                                                        int N, int ldp, int units, const EpiParams ep) {
  pdl_trigger();
  pdl_wait();
  const int64_t total = (int64_t)nz0 * nz1 * M * units;
  const int64_t i = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (i >= total) return;
  const int lane = threadIdx.x & 31;
  const int row = (int)(i % M);
  const int unit = (int)((i / M) % units);
  const int z = (int)(i / ((int64_t)M * units));
  const int64_t slice_stride = (int64_t)nz0 * nz1 * M * ldp;
  const int64_t row_off = ((int64_t)z * M + row) * ldp;
  float a[32];
  if constexpr (EPI == EPI_STORE || EPI == EPI_EULER) {
    const int col0 = unit * 32;
    x3_fetch_chunk(partial, S, slice_stride, row_off + col0, lane, a);
    if (lane == 0) {
      if constexpr (EPI == EPI_STORE)
        epi_store_chunk<float>(ep, row, col0, N, (int64_t)(z % nz0) * ep.out_z0 + (int64_t)(z / nz0) * ep.out_z1, a);
      else epi_euler_chunk<float>(ep, row, col0, N, a);
    }
  } else if constexpr (EPI == EPI_GEGLU) {
    float u[32];
    const int out_col0 = unit * 32;                       // tile = 128 output columns: 128 gate rows then the 128 up rows
    const int gcol = (out_col0 / 128) * 256 + out_col0 % 128;
    x3_fetch_chunk(partial, S, slice_stride, row_off + gcol, lane, a);
    x3_fetch_chunk(partial, S, slice_stride, row_off + gcol + 128, lane, u);
    if (lane == 0) epi_geglu_chunk<float>(ep, row, out_col0, N / 2, a, u);
  } else {
    float x2[32];
    const int head = unit >> 2, c = unit & 3;
    x3_fetch_chunk(partial, S, slice_stride, row_off + head * 256 + c * 32, lane, a);
    x3_fetch_chunk(partial, S, slice_stride, row_off + head * 256 + 128 + c * 32, lane, x2);
    const int which = head < ep.n_q_heads ? 0 : (head == ep.n_q_heads ? 1 : 2);
    if (lane == 0) epi_rope_chunk<float>(ep, row, which, head, c, a, x2);
  }
}

// ---------------------------------------------------------------------------------------------
// Split-K variant for skinny problems (fewer output tiles than SMs): one thread-block CLUSTER per output tile, CTA
// `rank` of the cluster accumulates k-blocks [rank*kb_per_split, ...) of the tile in its own TMEM, spills the fp32
// partial into its shared memory ([col][row], reusing the drained operand ring), and after a cluster barrier every
// CTA finalises a share of the tile's column chunks by summing the partials of all ranks IN RANK ORDER over
// distributed shared memory (deterministic), then runs the ordinary epilogue. No global workspace, no atomics.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  __syncwarp();
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t dsmem_map(uint32_t local_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ float dsmem_ld_f32(uint32_t addr) {
  float v;
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
  return v;
}
// Partial tiles live column-major ([col][row], 128 rows) in shared memory: the lanes of a warp own consecutive rows,
// so both the spill and the remote reads are 128-byte coalesced (a row-major layout with 16-byte remote loads was
// measured 2x slower: lanes then stride a whole row). v[j] = sum over ranks (ascending) of partial[rank][chunk*32+j][r].
template <int BN>
__device__ __forceinline__ void splitk_fetch(uint32_t stage_u32, int splits, int chunk, int r, float* v) {
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = 0.0f;
  for (int s = 0; s < splits; ++s) {
    const uint32_t base = dsmem_map(stage_u32, (uint32_t)s) + (uint32_t)(((chunk * 32) * kBM + r) * 4);
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] += dsmem_ld_f32(base + (uint32_t)(j * kBM * 4));
  }
}

// NORM: fused RMSNorm of the A operand (norm_a_operand above)
template <int BN, int EPI, bool NORM = false>
__global__ void __launch_bounds__(kTcThreads, 1)
gemm_tc_splitk_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const TcShape sh,
                      const EpiParams ep, const NormArgs na) {
  using Cfg = TcCfg<BN>;
  static_assert((uint32_t)BN * kBM * 4 <= Cfg::kStages * Cfg::kStageBytes, "partial tile must fit the operand ring");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tfull_bar + 4);
  uint64_t* ready_bar = tfull_bar + 5;                                   // NORM: A tile of the stage normalised
  float* rstd_s = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(full_bar) + 256);   // NORM: [128] (launch adds 1 KiB)

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int splits = sh.splits;
  const int rank = (int)cluster_ctarank();
  const int tile = blockIdx.x / splits;
  pdl_trigger();

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&tfull_bar[0], 1);
    if constexpr (NORM)
      for (int s = 0; s < Cfg::kStages; ++s) mbar_init(&ready_bar[s], 8);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<(BN < 32 ? 32 : BN)>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int num_k = (sh.K + kBK - 1) / kBK;
  const int kb0 = rank * sh.kb_per_split;
  const int kb1 = min(num_k, kb0 + sh.kb_per_split);
  int z, tm, tn;
  tile_coords(sh, tile, z, tm, tn);
  const int z0 = z % sh.nz0, z1 = z / sh.nz0;
  // weights do not depend on the previous kernel: the W boxes of the first ring pass are requested before the wait
  const int pre = sh.w_prefetch ? min(kb1 - kb0, Cfg::kStages) : 0;
  if (pre > 0 && warp == 0 && elect_one()) {
    for (int i = 0; i < pre; ++i) {
      mbar_arrive_expect_tx(&full_bar[i], sh.stage_tx);
      tma_load_4d(smem + i * Cfg::kStageBytes + Cfg::kABytes, &tmB, &full_bar[i], (kb0 + i) * kBK, tn * BN, z0, z1);
    }
  }
  pdl_wait();

  if (warp == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sa = smem + stage * Cfg::kStageBytes;
        uint8_t* sb = sa + Cfg::kABytes;
        const bool prefetched = kb - kb0 < pre;
        if (!prefetched) mbar_arrive_expect_tx(&full_bar[stage], sh.stage_tx);
        tma_load_4d(sa, &tmA, &full_bar[stage], kb * kBK, tm * kBM, z0, z1);
        if (!prefetched) tma_load_4d(sb, &tmB, &full_bar[stage], kb * kBK, tn * BN, z0, z1);
        if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = umma_idesc_bf16(kBM, BN);
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = kb0; kb < kb1; ++kb) {
      mbar_wait(NORM ? &ready_bar[stage] : &full_bar[stage], phase);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t sa = smem_u32(smem + stage * Cfg::kStageBytes);
        const uint64_t da = umma_desc_k_sw128(sa);
        const uint64_t db = umma_desc_k_sw128(sa + Cfg::kABytes);
#pragma unroll
        for (int k = 0; k < kBK / 16; ++k)
          umma_bf16(tmem_base, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb > kb0 || k != 0) ? 1u : 0u);
        umma_commit(&empty_bar[stage]);
        if (kb == kb1 - 1) umma_commit(&tfull_bar[0]);
      }
      __syncwarp();
      if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
    }
  } else {
    if constexpr (NORM)
      norm_a_operand<BN>(smem, full_bar, ready_bar, rstd_s, na, sh, tm, kb0, kb1, (int)threadIdx.x - 64);
    // spill this CTA's partial accumulator: TMEM -> smem [col][row] (the operand ring is drained once tfull fires)
    const int quad = warp & 3;
    const int half = (warp - 2) >> 2;
    const int r = quad * 32 + lane;
    mbar_wait(&tfull_bar[0], 0);
    tc_fence_after();
    float* stage_f = reinterpret_cast<float*>(smem);
    const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16);
    constexpr int kChunks = BN / 32;
#pragma unroll 1
    for (int c = half; c < kChunks; c += 2) {
      if (tn * BN + c * 32 >= sh.N) break;
      float v[32];
      tmem_ld_32x32(taddr + c * 32, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) stage_f[(c * 32 + j) * kBM + r] = v[j];
    }
    tc_fence_before();
  }
  cluster_sync_all();  // every rank's partial is visible cluster-wide

  if (warp >= 2) {
    const int quad = warp & 3;
    const int half = (warp - 2) >> 2;
    const int r = quad * 32 + lane;
    const int row = tm * kBM + r;
    const int n0 = tn * BN;
    const uint32_t stage_u32 = smem_u32(smem);
    if (row < sh.M) {
      if constexpr (EPI == EPI_STORE || EPI == EPI_EULER) {
        const int64_t zoff = (int64_t)z0 * ep.out_z0 + (int64_t)z1 * ep.out_z1;
        int i = 0;
        for (int c = rank; c < BN / 32; c += splits, ++i) {
          if ((i & 1) != half) continue;
          const int col0 = n0 + c * 32;
          if (col0 >= sh.N) break;
          float v[32];
          splitk_fetch<BN>(stage_u32, splits, c, r, v);
          if constexpr (EPI == EPI_STORE) epi_store_chunk<bf16>(ep, row, col0, sh.N, zoff, v);
          else epi_euler_chunk<bf16>(ep, row, col0, sh.N, v);
        }
      } else if constexpr (EPI == EPI_GEGLU) {
        const int n_out = sh.N / 2;
        int i = 0;
        for (int c = rank; c < BN / 64; c += splits, ++i) {
          if ((i & 1) != half) continue;
          const int out_col0 = tn * (BN / 2) + c * 32;
          if (out_col0 >= n_out) break;
          float g[32], u[32];
          splitk_fetch<BN>(stage_u32, splits, c, r, g);
          splitk_fetch<BN>(stage_u32, splits, BN / 64 + c, r, u);
          epi_geglu_chunk<bf16>(ep, row, out_col0, n_out, g, u);
        }
      } else if constexpr (EPI == EPI_ROPE) {
        static_assert(EPI != EPI_ROPE || BN == 256, "RoPE epilogue needs one 256-wide head per tile");
        const int which = tn < ep.n_q_heads ? 0 : (tn == ep.n_q_heads ? 1 : 2);
        int i = 0;
        for (int c = rank; c < 4; c += splits, ++i) {
          if ((i & 1) != half) continue;
          float x1[32], x2[32];
          splitk_fetch<BN>(stage_u32, splits, c, r, x1);
          splitk_fetch<BN>(stage_u32, splits, 4 + c, r, x2);
          epi_rope_chunk<bf16>(ep, row, which, tn, c, x1, x2);
        }
      }
    }
  }
  cluster_sync_all();  // nobody leaves while a peer may still read its partial
  tc_fence_after();
  if (warp == 1) tmem_dealloc<(BN < 32 ? 32 : BN)>(tmem_base);
}

// ---------------------------------------------------------------------------------------------
// CTA-pair variant (tcgen05 cta_group::2) for problems with many tiles: a cluster of two CTAs (the two SMs of a TPC)
// owns a 256 x BN output tile. Each CTA stages ITS 128 rows of A and ITS half (BN/2 rows) of the W tile, so a k-block
// costs (128 + BN/2) x 64 operand elements per SM instead of (128 + BN) x 64: the L2 -> SM path (~64 B/clk per SM),
// which caps the single-CTA kernel at 128 x 256 tiles, is relieved by a third and the ring holds 6 stages instead of 4.
// The leader CTA (cluster rank 0) issues one M = 256 MMA per 16-wide k-step that reads both CTAs' shared memory and
// writes both CTAs' TMEM (128 accumulator rows each); tcgen05.commit multicasts the "stage free" and "accumulator
// ready" arrivals to both CTAs, both producers signal their TMA bytes on the LEADER's full barrier, and both CTAs'
// epilogue warps hand the accumulator buffer back on the leader's barrier.
// ---------------------------------------------------------------------------------------------
// EW = epilogue warps per TMEM lane quadrant (2 or 3). Short-K tiles (SigLIP, K = 1152: main loop 5.3 us per tile) are bound by
// the 8 epilogue warps (6.1 us for a bias + GELU tile): with 12 warps a warp drains 3 instead of 4 of the tile's 8 column
// chunks. The four extra staging buffers cost one ring stage.
template <int BN, int EW = 2>
struct Tc2Cfg {
  static constexpr uint32_t kABytes = kBM * kBK * 2;
  static constexpr uint32_t kBBytes = (BN / 2) * kBK * 2;
  static constexpr uint32_t kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (BN == 256) ? (EW == 2 ? 6 : 5) : 8;
  static constexpr int kTmemCols = 2 * BN;
  static constexpr int kThreads = 64 + 128 * EW;
  static constexpr uint32_t kSmemBytes = kStages * kStageBytes + 1024 + 1024 + 4 * EW * 2 * kStoreBufBytes;
  static_assert(kSmemBytes <= 232448, "shared memory budget of one SM");
};

template <int BN, int EPI, int EW = 2>
__global__ void __launch_bounds__(64 + 128 * EW, 1)
gemm_tc_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmC, const TcShape sh,
                    const EpiParams ep) {
  using Cfg = Tc2Cfg<BN, EW>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int rank = (int)cluster_ctarank();
  const int pair = blockIdx.x >> 1;
  const int n_pairs = gridDim.x >> 1;
  pdl_trigger();

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (sh.tma_store) tma_prefetch_desc(&tmC);
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);    // leader's producer (arrive + expected bytes of BOTH CTAs); unused in the peer
      mbar_init(&empty_bar[s], 1);   // one multicast tcgen05.commit per use
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], 8 * EW);  // 4 EW epilogue warps of each CTA; only the leader's copy is waited on
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc_2sm<Cfg::kTmemCols>(tmem_slot);
  tc_fence_before();
  cluster_sync_all();  // barriers of both CTAs initialised before any remote arrive / TMA completion can land
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int num_k = (sh.K + kBK - 1) / kBK;
  pdl_wait();

  if (warp == 0) {
    // ------------------------------ TMA producer (both CTAs) ------------------
    if (elect_one()) {
      const uint32_t full_leader = dsmem_map(smem_u32(&full_bar[0]), 0);
      int stage = 0;
      uint32_t phase = 0;
      for (int idx = pair; idx < sh.total_tiles; idx += n_pairs) {
        int z, tm, tn;
        tile_coords(sh, idx, z, tm, tn);
        const int z0 = z % sh.nz0, z1 = z / sh.nz0;
        const int a_row = (tm * 2 + rank) * kBM;
        const int w_row = tn * BN + rank * (BN / 2);
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * Cfg::kStageBytes;
          uint8_t* sb = sa + Cfg::kABytes;
          if (rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * Cfg::kStageBytes);
          tma_load_4d_2sm(sa, &tmA, full_leader + (uint32_t)(stage * 8), kb * kBK, a_row, z0, z1);
          tma_load_4d_2sm(sb, &tmB, full_leader + (uint32_t)(stage * 8), kb * kBK, w_row, z0, z1);
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer (leader CTA only) --------------
    if (rank == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(2 * kBM, BN);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int idx = pair; idx < sh.total_tiles; idx += n_pairs) {
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < num_k; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            if (kb == 0) PAIR_STAMP(0, (idx - pair) / n_pairs, 0);
            if (kb == num_k - 1) PAIR_STAMP(0, (idx - pair) / n_pairs, 1);
            const uint32_t sa = smem_u32(smem + stage * Cfg::kStageBytes);
            const uint64_t da = umma_desc_k_sw128(sa);
            const uint64_t db = umma_desc_k_sw128(sa + Cfg::kABytes);
#pragma unroll
            for (int k = 0; k < kBK / 16; ++k)
              umma_bf16_2sm(tmem_d, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb | k) != 0 ? 1u : 0u);
            umma_commit_2sm(&empty_bar[stage], 3);
            if (kb == num_k - 1) umma_commit_2sm(&tfull_bar[acc], 3);
          }
          __syncwarp();
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else {
    // ------------------------------ epilogue (both CTAs, own 128 rows) --------
    const int quad = warp & 3;
    const int half = (warp - 2) >> 2;
    uint8_t* stage2 = reinterpret_cast<uint8_t*>(full_bar) + 1024 + (warp - 2) * 2 * kStoreBufBytes;
    int sbuf = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int idx = pair; idx < sh.total_tiles; idx += n_pairs) {
      int z, tm, tn;
      tile_coords(sh, idx, z, tm, tn);
      const int z0 = z % sh.nz0, z1 = z / sh.nz0;
      const int row = (tm * 2 + rank) * kBM + quad * 32 + lane;
      const bool row_ok = row < sh.M;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      if (threadIdx.x == 64) PAIR_STAMP(rank, (idx - pair) / n_pairs, 2);
      const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + acc * BN;
      drain_tile<BN, EPI, EW>(sh, ep, taddr, row, row_ok, tn, z0, z1, half, &tmC, stage2, sbuf);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&tempty_bar[acc], 0);
      if (threadIdx.x == 64) PAIR_STAMP(rank, (idx - pair) / n_pairs, 3);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
    if (sh.tma_store && lane == 0) bulk_wait_all();
  }

  tc_fence_before();
  cluster_sync_all();  // the peer's shared memory and TMEM stay alive until the leader's last MMA has retired
  tc_fence_after();
  if (warp == 1) tmem_dealloc_2sm<Cfg::kTmemCols>(tmem_base);
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess || p == nullptr) {
    return nullptr;
  }
  fn = reinterpret_cast<EncodeTiledFn>(p);
  return fn;
}

// rank-4 bf16 tensor map {k, rows, z0, z1}, box {64, box_rows, 1, 1}, 128 B swizzle, zero OOB fill
static int make_map(CUtensorMap* map, const void* base, int K, int rows, int nz0, int nz1, int64_t ld, int64_t sz0,
                    int64_t sz1, int box_rows, int box_k = kBK, CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
  EncodeTiledFn fn = get_encode_fn();
  PI0_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled entry point not available");
  PI0_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA operand base %p not 16-byte aligned", base);
  PI0_REQUIRE((ld * 2) % 16 == 0, "TMA operand leading dimension %lld not a multiple of 8 elements", (long long)ld);
  cuuint64_t gdim[4] = {(cuuint64_t)K, (cuuint64_t)rows, (cuuint64_t)nz0, (cuuint64_t)nz1};
  cuuint64_t s1 = (cuuint64_t)ld * 2;
  cuuint64_t s2 = nz0 > 1 ? (cuuint64_t)sz0 * 2 : s1;
  cuuint64_t s3 = nz1 > 1 ? (cuuint64_t)sz1 * 2 : s1;
  PI0_REQUIRE(s2 % 16 == 0 && s3 % 16 == 0, "TMA batch strides must be multiples of 8 elements");
  cuuint64_t gstride[3] = {s1, s2, s3};
  cuuint32_t box[4] = {(cuuint32_t)box_k, (cuuint32_t)box_rows, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  PI0_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d): K=%d rows=%d nz=%dx%d ld=%lld", (int)r, K, rows,
              nz0, nz1, (long long)ld);
  return 0;
}

// the same encoder for the other tcgen05 kernels of the library (attn_prefill.cu)
int tc_make_map(CUtensorMap* map, const void* base, int K, int rows, int nz0, int nz1, int64_t ld, int64_t sz0, int64_t sz1,
                int box_rows, int box_k, int swizzle_bytes) {
  return make_map(map, base, K, rows, nz0, nz1, ld, sz0, sz1, box_rows, box_k,
                  swizzle_bytes == 0 ? CU_TENSOR_MAP_SWIZZLE_NONE
                                     : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B);
}

// Output tensor map of the TMA-store epilogue ({columns, rows, z0, z1}, 32 x 32 box, 64-byte swizzle) when the epilogue
// writes a plain row-major bf16 matrix: EPI_STORE without row remap / transposed-V region, or EPI_GEGLU. Returns 1 when
// `map` is valid, 0 when the kernel must keep its direct stores, < 0 on error.
static int make_out_map(CUtensorMap* map, const GemmProblem& g) {
  const bool store = g.epi_mode == EPI_STORE && g.epi.rpb_in <= 0 && (g.epi.vt == nullptr || g.epi.vt_col0 >= g.N);
  const bool geglu = g.epi_mode == EPI_GEGLU && g.epi.rpb_in <= 0;
  if (!store && !geglu) return 0;
  static int enabled = -1;
  if (enabled < 0) {
    const char* e = getenv("PI0_TMA_STORE");
    enabled = (e && e[0] == '0') ? 0 : 1;
  }
  if (!enabled) return 0;
  const int n_out = geglu ? g.N / 2 : g.N;
  const int64_t z0 = g.nz0 > 1 ? g.epi.out_z0 : 0, z1 = g.nz1 > 1 ? g.epi.out_z1 : 0;
  if ((reinterpret_cast<uintptr_t>(g.epi.out) & 15) != 0 || (g.epi.ldc * 2) % 16 != 0 || (z0 * 2) % 16 != 0 ||
      (z1 * 2) % 16 != 0 || g.epi.ldc < n_out)
    return 0;
  PI0_TRY(make_map(map, g.epi.out, n_out, g.M, g.nz0, g.nz1, g.epi.ldc, z0, z1, 32, 32, CU_TENSOR_MAP_SWIZZLE_64B));
  return 1;
}

template <int BN, int EPI>
static int launch_cfg(const GemmProblem& g, int num_sms, int splits, cudaStream_t stream) {
  using Cfg = TcCfg<BN>;
  static PerDeviceOnce attr_once;
  if (attr_once.need()) {
    PI0_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<BN, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)Cfg::kSmemBytesStaged));
    PI0_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_splitk_kernel<BN, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)Cfg::kSmemBytes));
    attr_once.mark();
  }
  CUtensorMap tmA, tmB;
  // a single short M tile only fetches the rows that exist (rounded to the 8-row swizzle atom): the MMA still reads
  // 128 smem rows, but rows >= M only feed accumulator rows that the epilogue never stores
  const int a_rows = g.M < kBM ? ((g.M + 7) / 8) * 8 : kBM;
  PI0_TRY(make_map(&tmA, g.A, g.K, g.M, g.nz0, g.nz1, g.lda, g.a_z0, g.a_z1, a_rows));
  PI0_TRY(make_map(&tmB, g.W, g.K, g.N, g.nz0, g.nz1, g.ldw, g.w_z0, g.w_z1, BN));
  TcShape sh;
  sh.M = g.M; sh.N = g.N; sh.K = g.K; sh.nz0 = g.nz0; sh.nz1 = g.nz1;
  sh.tiles_m = (g.M + kBM - 1) / kBM;
  sh.tiles_n = (g.N + BN - 1) / BN;
  sh.total_tiles = sh.tiles_m * sh.tiles_n * g.nz0 * g.nz1;
  set_super(sh, kBM);
  const int num_k = (g.K + kBK - 1) / kBK;
  sh.splits = 1;
  sh.kb_per_split = num_k;
  sh.stage_tx = (uint32_t)(a_rows * kBK * 2) + Cfg::kBBytes;
  sh.tma_store = 0;
  sh.w_prefetch = g.w_const ? 1 : 0;
  if (splits > 1) {
    // every rank of the cluster must own at least one k-block
    int per = (num_k + splits - 1) / splits;
    int eff = (num_k + per - 1) / per;
    sh.splits = eff;
    sh.kb_per_split = per;
  }
  if (sh.splits > 1) {
    PI0_CHECK_CUDA(launch_pdl(gemm_tc_splitk_kernel<BN, EPI>, dim3((unsigned)(sh.total_tiles * sh.splits)),
                              dim3(kTcThreads), Cfg::kSmemBytes, stream, sh.splits, tmA, tmB, sh, g.epi, NormArgs{nullptr, nullptr, 0, 0.0f}));
  } else {
    int grid = sh.total_tiles < num_sms ? sh.total_tiles : num_sms;
    CUtensorMap tmC = tmA;
    const int have_c = make_out_map(&tmC, g);
    if (have_c < 0) return have_c;
    sh.tma_store = have_c;
    PI0_CHECK_CUDA(launch_pdl(gemm_tc_kernel<BN, EPI>, dim3((unsigned)grid), dim3(kTcThreads), Cfg::kSmemBytesStaged,
                              stream, 1, tmA, tmB, tmC, sh, g.epi, NormArgs{nullptr, nullptr, 0, 0.0f}));
  }
  PI0_CHECK_CUDA(cudaGetLastError());
  ++g_launch_count;
  return 0;
}

// CTA pairs the device can hold at once (148 SMs = 74 TPCs on a full B200; fewer if a GPC has an odd SM out)
template <int BN, int EPI>
static int max_active_pairs(int num_sms) {
  static int cached = 0;
  if (cached > 0) return cached;
  using Cfg = Tc2Cfg<BN>;
  cudaLaunchConfig_t lc = {};
  lc.gridDim = dim3((unsigned)(2 * (num_sms / 2)), 1, 1);
  lc.blockDim = dim3(kTcThreads, 1, 1);
  lc.dynamicSmemBytes = Cfg::kSmemBytes;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  lc.attrs = at;
  lc.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, gemm_tc_pair_kernel<BN, EPI>, &lc) != cudaSuccess || n <= 0) {
    cudaGetLastError();
    n = num_sms / 2 - 2;
  }
  if (n > num_sms / 2) n = num_sms / 2;
  cached = n;
  return n;
}

template <int BN, int EPI, int EW>
static int launch_pair_ew(const GemmProblem& g, int num_sms, cudaStream_t stream) {
  using Cfg = Tc2Cfg<BN, EW>;
  static PerDeviceOnce attr_once;
  if (attr_once.need()) {
    PI0_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_pair_kernel<BN, EPI, EW>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)Cfg::kSmemBytes));
    attr_once.mark();
  }
  CUtensorMap tmA, tmB;
  PI0_TRY(make_map(&tmA, g.A, g.K, g.M, g.nz0, g.nz1, g.lda, g.a_z0, g.a_z1, kBM));
  PI0_TRY(make_map(&tmB, g.W, g.K, g.N, g.nz0, g.nz1, g.ldw, g.w_z0, g.w_z1, BN / 2));
  TcShape sh;
  sh.M = g.M; sh.N = g.N; sh.K = g.K; sh.nz0 = g.nz0; sh.nz1 = g.nz1;
  sh.tiles_m = (g.M + 2 * kBM - 1) / (2 * kBM);   // 256-row tiles, one per CTA pair
  sh.tiles_n = (g.N + BN - 1) / BN;
  sh.total_tiles = sh.tiles_m * sh.tiles_n * g.nz0 * g.nz1;
  set_super(sh, 2 * kBM);
  sh.splits = 1;
  sh.kb_per_split = (g.K + kBK - 1) / kBK;
  sh.stage_tx = 2 * Cfg::kStageBytes;
  sh.w_prefetch = 0;
  CUtensorMap tmC = tmA;
  const int have_c = make_out_map(&tmC, g);
  if (have_c < 0) return have_c;
  sh.tma_store = have_c;
  const int cap = max_active_pairs<BN, EPI>(num_sms);   // the EW = 2 kernel's figure: smem-bound (1 CTA per SM) for both
  const int pairs = sh.total_tiles < cap ? sh.total_tiles : cap;
  PI0_CHECK_CUDA(launch_pdl(gemm_tc_pair_kernel<BN, EPI, EW>, dim3((unsigned)(2 * pairs)), dim3(Cfg::kThreads),
                            Cfg::kSmemBytes, stream, 2, tmA, tmB, tmC, sh, g.epi));
  PI0_CHECK_CUDA(cudaGetLastError());
  ++g_launch_count;
  return 0;
}

template <int BN, int EPI>
static int launch_pair(const GemmProblem& g, int num_sms, cudaStream_t stream) {
  // short-K plain-store tiles (SigLIP q|k|v, out_proj, fc1: K = 1152) are epilogue-bound: 12 epilogue warps, 5 ring stages.
  // PI0_PAIR_EW3_MAX_K (default 2047; 0 = never) is the largest K that takes that variant.
  if constexpr (BN == 256 && EPI == EPI_STORE) {
    static const int max_k = getenv("PI0_PAIR_EW3_MAX_K") ? atoi(getenv("PI0_PAIR_EW3_MAX_K")) : 2047;
    if (g.K <= max_k) return launch_pair_ew<BN, EPI, 3>(g, num_sms, stream);
  }
  return launch_pair_ew<BN, EPI, 2>(g, num_sms, stream);
}

// The CTA-pair kernel pays off once the single-CTA kernel runs several waves of tiles (the regime where the L2 -> SM
// operand path, not latency, bounds it). PI0_GEMM_PAIR=0 in the environment turns it off (A/B measurements).
static bool pair_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("PI0_GEMM_PAIR");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}
static bool use_pair(const GemmProblem& g, int bn, int num_sms) {
  if (bn != 256 && !(bn == 128 && g.epi_mode == EPI_STORE)) return false;
  if (g.M <= kBM) return false;
  if (g.pair > 0) return true;
  if (g.pair < 0 || g.splits > 1 || !pair_enabled()) return false;  // an auto-picked split yields to the pair kernel
  // Very short K makes a tile epilogue-bound (a 128 x 256 bias + GELU epilogue takes ~6 us next to a running main loop, a
  // K = 1152 pair main loop 5.3 us) and the pair's coarser wave quantisation (72 slots) then costs more than its main loop
  // gains: measured break-even near K = 1024 with the TMA-store epilogue (SigLIP tower 18.2 -> 17.5 ms at batch 64 with
  // its K = 1152 GEMMs on the pair kernel). PI0_PAIR_MIN_K overrides.
  static int min_k = -1;
  if (min_k < 0) {
    const char* e = getenv("PI0_PAIR_MIN_K");
    min_k = e ? atoi(e) : 1024;
  }
  if (g.K < min_k) return false;
  const long tiles = (long)((g.M + kBM - 1) / kBM) * ((g.N + bn - 1) / bn) * g.nz0 * g.nz1;
  return tiles >= 4L * num_sms;
}

// How many clusters of `sp` GEMM CTAs (1 CTA per SM: ~195 KB of shared memory each) the device can hold at once.
// GPCs have 16-20 SMs, so e.g. only two clusters of 8 fit a GPC; a launch that needs more runs in waves.
int max_active_clusters(int sp, int num_sms) {
  static int cache[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  if (sp <= 1) return num_sms;
  if (cache[sp] > 0) return cache[sp];
  using Cfg = TcCfg<256>;
  cudaFuncSetAttribute(gemm_tc_splitk_kernel<256, EPI_STORE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                       (int)Cfg::kSmemBytes);
  cudaLaunchConfig_t lc = {};
  lc.gridDim = dim3((unsigned)(sp * 64), 1, 1);
  lc.blockDim = dim3(kTcThreads, 1, 1);
  lc.dynamicSmemBytes = Cfg::kSmemBytes;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = (unsigned)sp;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  lc.attrs = at;
  lc.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, gemm_tc_splitk_kernel<256, EPI_STORE>, &lc) != cudaSuccess || n <= 0) {
    cudaGetLastError();
    n = num_sms / sp > 2 ? num_sms / sp - 2 : 1;  // conservative guess
  }
  cache[sp] = n;
  return n;
}

// Tile / split choice (bn == 0 / splits == 0 => auto). Cost model in microseconds, calibrated with
// tools/small_bench.py and tools/timeline.py on a B200:
//   t = launch gap + waves * (CTA setup/exit + k-blocks per CTA * t_kb(bn) + epilogue + split reduction)
// One elected thread issues the TMA / tcgen05 instructions (~32 / ~65 cycles each), so a k-block costs >= ~0.21 us
// whatever the tile width; wide tiles are MMA-bound (0.47 us per 128x256x64 k-block under the power cap). A persistent launch
// (splits == 1) pays the CTA setup once. Split factors are thread-block cluster sizes (<= 8); every rank must own
// >= 2 k-blocks; the reduction moves rows x 32 x 4 B per chunk and rank over distributed shared memory.
void gemm_tc_pick(int M, int N, int K, int nz, int num_sms, bool fixed_bn, int* bn_io, int* splits_out,
                  bool rope_epilogue = false) {
  const int tm = (M + kBM - 1) / kBM;
  const int num_k = (K + kBK - 1) / kBK;
  const int rows = M < kBM ? M : kBM;
  int best_bn = fixed_bn ? *bn_io : 256, best_s = 1;
  double best_cost = 1e30;
  const int cands[4] = {256, 128, 64, 32};
  const double t_kb[4] = {0.47, 0.35, 0.30, 0.28};       // many waves: MMA / power bound
  const double t_kb_skinny[4] = {0.34, 0.25, 0.21, 0.21};  // one wave: issue bound
  for (int ci = 0; ci < 4; ++ci) {
    const int bn = cands[ci];
    if (fixed_bn && bn != *bn_io) continue;
    if (!fixed_bn && bn > 64 && N < bn) continue;
    const long tiles = (long)tm * ((N + bn - 1) / bn) * nz;
    // the RoPE / KV-cache epilogue costs ~4x a plain store per tile (measured: denoise q|k|v at M = 320, 30 tiles: 20 us
    // unsplit, 15 us as clusters of 4 that share the epilogue; clusters of 8 lose again to the reduction)
    const double t_epi = (rope_epilogue ? 1.2 : 0.3) * bn / 64.0;
    for (int sp = 1; sp <= (rope_epilogue ? 4 : 8); ++sp) {
      if (sp > 1 && num_k / sp < 2) break;
      const int per = (num_k + sp - 1) / sp;
      double cost;
      if (sp == 1) {
        const long waves = (tiles + num_sms - 1) / num_sms;
        // a single skinny wave is issue-bound (0.37 us per k-block), many waves are MMA/power-bound
        cost = 1.7 + 2.5 + (double)waves * per * (waves > 1 ? t_kb[ci] : t_kb_skinny[ci]) + t_epi;
      } else {
        const int cap = max_active_clusters(sp, num_sms);
        const long waves = (tiles + cap - 1) / cap;
        const int chunks = bn / 32;
        const double red_bytes = (double)rows * 32 * 4 * sp * ((chunks + sp - 1) / sp);
        // several waves of clusters keep every SM's tensor pipe busy: MMA / power bound like the persistent kernel
        cost = 1.7 + (double)waves * (2.5 + per * (waves > 1 ? t_kb[ci] : t_kb_skinny[ci]) + t_epi / sp + 0.5 +
                                      red_bytes / 21000.0);
      }
      cost += 1e-3 * sp;
      if (cost < best_cost) { best_cost = cost; best_bn = bn; best_s = sp; }
    }
  }
  *bn_io = best_bn;
  *splits_out = best_s;
}

// ---- Gemma RMSNorm fused into the A operand (GemmProblem.norm_gain): the one-tile-per-CTA cluster kernel with NORM = true,
// bn = 256, the q|k|v (RoPE) and gate/up (GeGLU) epilogues; splits = 1 is a cluster of one ----
template <int EPI>
static int launch_norm(const GemmProblem& g, int splits, int num_sms, cudaStream_t stream) {
  constexpr int BN = 256;
  using Cfg = TcCfg<BN>;
  constexpr uint32_t kSmem = Cfg::kSmemBytes + 1024;   // + the per-row rstd table
  static PerDeviceOnce attr_once;
  if (attr_once.need()) {
    PI0_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_splitk_kernel<BN, EPI, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)kSmem));
    attr_once.mark();
  }
  PI0_REQUIRE(g.K % kBK == 0 && g.nz0 * g.nz1 == 1, "gemm_tc (fused norm): K=%d must be a multiple of 64, no batching", g.K);
  PI0_REQUIRE(g.norm_rowsq != nullptr && g.norm_slots >= 8 && g.norm_slots % 8 == 0,
              "gemm_tc (fused norm): row statistics missing or slots=%d not a multiple of 8", g.norm_slots);
  CUtensorMap tmA, tmB;
  const int a_rows = g.M < kBM ? ((g.M + 7) / 8) * 8 : kBM;
  PI0_TRY(make_map(&tmA, g.A, g.K, g.M, 1, 1, g.lda, 0, 0, a_rows));
  PI0_TRY(make_map(&tmB, g.W, g.K, g.N, 1, 1, g.ldw, 0, 0, BN));
  TcShape sh;
  sh.M = g.M; sh.N = g.N; sh.K = g.K; sh.nz0 = 1; sh.nz1 = 1;
  sh.tiles_m = (g.M + kBM - 1) / kBM;
  sh.tiles_n = (g.N + BN - 1) / BN;
  sh.total_tiles = sh.tiles_m * sh.tiles_n;
  set_super(sh, kBM);
  const int num_k = g.K / kBK;
  const int per = (num_k + splits - 1) / splits;
  sh.splits = (num_k + per - 1) / per;
  sh.kb_per_split = per;
  sh.stage_tx = (uint32_t)(a_rows * kBK * 2) + Cfg::kBBytes;
  sh.tma_store = 0;
  sh.w_prefetch = g.w_const ? 1 : 0;
  const NormArgs na{g.norm_gain, g.norm_rowsq, g.norm_slots, g.norm_eps};
  if (sh.splits == 1 && sh.total_tiles <= num_sms) {
    // one tile per CTA without a k split: the persistent kernel, which keeps its TMA-store epilogue
    static PerDeviceOnce attr2;
    if (attr2.need()) {
      PI0_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<BN, EPI, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)Cfg::kSmemBytesStaged));
      attr2.mark();
    }
    CUtensorMap tmC = tmA;
    const int have_c = make_out_map(&tmC, g);
    if (have_c < 0) return have_c;
    sh.tma_store = have_c;
    PI0_CHECK_CUDA(launch_pdl(gemm_tc_kernel<BN, EPI, true>, dim3((unsigned)sh.total_tiles), dim3(kTcThreads),
                              Cfg::kSmemBytesStaged, stream, 1, tmA, tmB, tmC, sh, g.epi, na));
    PI0_CHECK_CUDA(cudaGetLastError());
    ++g_launch_count;
    return 0;
  }
  PI0_CHECK_CUDA(launch_pdl(gemm_tc_splitk_kernel<BN, EPI, true>, dim3((unsigned)(sh.total_tiles * sh.splits)),
                            dim3(kTcThreads), kSmem, stream, sh.splits, tmA, tmB, sh, g.epi, na));
  PI0_CHECK_CUDA(cudaGetLastError());
  ++g_launch_count;
  return 0;
}

int gemm_tc_launch(const GemmProblem& g, int num_sms, cudaStream_t stream) {
  PI0_REQUIRE(g.M > 0 && g.N > 0 && g.K > 0, "gemm_tc: empty problem M=%d N=%d K=%d", g.M, g.N, g.K);
  // a ragged N is fine for EPI_STORE when nothing is read per column and the row pitch covers the 8-wide store
  const bool ragged_ok = g.epi_mode == EPI_STORE && g.epi.bias == nullptr && g.epi.res == nullptr &&
                         g.epi.ldc >= (int64_t)((g.N + 7) / 8) * 8;
  PI0_REQUIRE(g.N % 8 == 0 || g.epi_mode == EPI_EULER || g.epi.vt != nullptr || ragged_ok,
              "gemm_tc: N=%d must be a multiple of 8", g.N);
  int bn = g.bn, sp = g.splits;
  if (sp == 0 || bn == 0) {
    int auto_s = 1;
    const bool fixed = bn != 0;
    gemm_tc_pick(g.M, g.N, g.K, g.nz0 * g.nz1, num_sms, fixed, &bn, &auto_s, g.epi_mode == EPI_ROPE);
    if (sp == 0) sp = auto_s;
  }
  PI0_REQUIRE(sp >= 1 && sp <= 8, "gemm_tc: splits=%d outside [1,8]", sp);
  if (g.norm_gain != nullptr) {
    PI0_REQUIRE(bn == 256 && (g.epi_mode == EPI_ROPE || g.epi_mode == EPI_GEGLU),
                "gemm_tc (fused norm): only the bn = 256 RoPE / GeGLU GEMMs take a fused RMSNorm");
    if (g.epi_mode == EPI_ROPE) return launch_norm<EPI_ROPE>(g, sp, num_sms, stream);
    return launch_norm<EPI_GEGLU>(g, sp, num_sms, stream);
  }
  if (use_pair(g, bn, num_sms)) {
    switch (g.epi_mode) {
      case EPI_STORE:
        if (bn == 256) return launch_pair<256, EPI_STORE>(g, num_sms, stream);
        return launch_pair<128, EPI_STORE>(g, num_sms, stream);
      case EPI_GEGLU:
        PI0_REQUIRE(g.N % 256 == 0, "gemm_tc: GeGLU N=%d must be a multiple of 256", g.N);
        return launch_pair<256, EPI_GEGLU>(g, num_sms, stream);
      case EPI_ROPE:
        PI0_REQUIRE(g.N == (g.epi.n_q_heads + 2) * 256, "gemm_tc: RoPE epilogue expects N = (q_heads + 2) * 256");
        return launch_pair<256, EPI_ROPE>(g, num_sms, stream);
      default: break;  // EPI_EULER is a 32-wide tile: single-CTA kernel below
    }
  }
  switch (g.epi_mode) {
    case EPI_STORE:
      switch (bn) {
        case 256: return launch_cfg<256, EPI_STORE>(g, num_sms, sp, stream);
        case 128: return launch_cfg<128, EPI_STORE>(g, num_sms, sp, stream);
        case 64: return launch_cfg<64, EPI_STORE>(g, num_sms, sp, stream);
        case 32: return launch_cfg<32, EPI_STORE>(g, num_sms, sp, stream);
      }
      break;
    case EPI_GEGLU:
      PI0_REQUIRE(g.N % 256 == 0, "gemm_tc: GeGLU N=%d must be a multiple of 256", g.N);
      if (bn == 256) return launch_cfg<256, EPI_GEGLU>(g, num_sms, sp, stream);
      break;
    case EPI_ROPE:
      PI0_REQUIRE(g.N == (g.epi.n_q_heads + 2) * 256, "gemm_tc: RoPE epilogue expects N = (q_heads + 2) * 256");
      if (bn == 256) return launch_cfg<256, EPI_ROPE>(g, num_sms, sp, stream);
      break;
    case EPI_EULER:
      if (bn == 32) return launch_cfg<32, EPI_EULER>(g, num_sms, sp, stream);
      break;
  }
  set_error("gemm_tc: unsupported epilogue/tile combination epi=%d bn=%d", g.epi_mode, bn);
  return -2;
}


// ---- bf16 operands, fp32 OUTPUT (the fp32 regime's split-bf16 GEMM, gemm_x3.cu): the single-CTA persistent kernel with the
// epilogues instantiated for float storage and segmented accumulation (gemm_tc_seg_kernel). No split-K, no CTA pairs, no
// TMA-store staging (fp32 rows are 128-byte lines per lane already). ----
template <int BN, int EPI>
static int launch_f32out(const GemmProblem& g, int num_sms, void* partial, size_t partial_bytes, cudaStream_t stream) {
  using Cfg = TcCfg<BN>;
  static PerDeviceOnce attr_once;
  if (attr_once.need()) {
    PI0_CHECK_CUDA(cudaFuncSetAttribute(gemm_tc_seg_kernel<BN, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)Cfg::kSmemBytes));
    attr_once.mark();
  }
  CUtensorMap tmA, tmB;
  const int a_rows = g.M < kBM ? ((g.M + 7) / 8) * 8 : kBM;
  PI0_TRY(make_map(&tmA, g.A, g.K, g.M, g.nz0, g.nz1, g.lda, g.a_z0, g.a_z1, a_rows));
  PI0_TRY(make_map(&tmB, g.W, g.K, g.N, g.nz0, g.nz1, g.ldw, g.w_z0, g.w_z1, BN));
  TcShape sh;
  sh.M = g.M; sh.N = g.N; sh.K = g.K; sh.nz0 = g.nz0; sh.nz1 = g.nz1;
  sh.tiles_m = (g.M + kBM - 1) / kBM;
  sh.tiles_n = (g.N + BN - 1) / BN;
  sh.total_tiles = sh.tiles_m * sh.tiles_n * g.nz0 * g.nz1;
  set_super(sh, kBM);
  const int num_k = (g.K + kBK - 1) / kBK;
  sh.splits = 1;
  sh.kb_per_split = num_k;
  sh.stage_tx = (uint32_t)(a_rows * kBK * 2) + Cfg::kBBytes;
  sh.tma_store = 0;
  sh.w_prefetch = 0;
  // split-K over CTAs when the tiles cover less than half of the SMs (denoise steps, batch-1 prefill): every CTA takes a
  // k-slice of >= 4 k-blocks and leaves its fp32 partial in the workspace; x3_reduce_kernel finishes
  const int ldp = (g.N + 31) / 32 * 32;
  if (BN == 256 && partial != nullptr && sh.total_tiles * 2 <= num_sms && num_k >= 8) {
    int want = num_sms / sh.total_tiles;
    if (want > num_k / 4) want = num_k / 4;
    const size_t per_split = (size_t)g.nz0 * g.nz1 * g.M * ldp * sizeof(float);
    if ((size_t)want * per_split > partial_bytes) want = (int)(partial_bytes / per_split);
    if (want > 1) {
      const int per = (num_k + want - 1) / want;
      sh.splits = (num_k + per - 1) / per;
      sh.kb_per_split = per;
    }
  }
  const int items = sh.total_tiles * sh.splits;
  const int grid = items < num_sms ? items : num_sms;
  static int seg = getenv("PI0_X3_SEG_KB") ? atoi(getenv("PI0_X3_SEG_KB")) : 16;   // k-blocks (of 64) per accumulator run
  float* part = sh.splits > 1 ? (float*)partial : nullptr;
  PI0_CHECK_CUDA(launch_pdl(gemm_tc_seg_kernel<BN, EPI>, dim3((unsigned)grid), dim3(kTcSegThreads), Cfg::kSmemBytes, stream, 1,
                            tmA, tmB, sh, g.epi, seg > 0 ? seg : 1 << 30, part, ldp));
  PI0_CHECK_CUDA(cudaGetLastError());
  ++g_launch_count;
  if (sh.splits > 1) {
    const int units = EPI == EPI_GEGLU ? g.N / 64 : EPI == EPI_ROPE ? (g.N / 256) * 4 : ldp / 32;
    const int64_t warps = (int64_t)g.nz0 * g.nz1 * g.M * units;
    PI0_CHECK_CUDA(launch_pdl(x3_reduce_kernel<EPI>, dim3((unsigned)((warps + 7) / 8)), dim3(256), 0, stream, 1,
                              (const float*)partial, sh.splits, g.nz0, g.nz1, g.M, g.N, ldp, units, g.epi));
    ++g_launch_count;
  }
  return 0;
}

// sums `S` fp32 partial slices [S][nz][M][ldp] in slice order and runs the fused epilogue of `g` in fp32 storage
int gemm_f32_reduce_launch(const GemmProblem& g, const float* partial, int S, int ldp, cudaStream_t stream) {
  const int units = g.epi_mode == EPI_GEGLU ? g.N / 64 : g.epi_mode == EPI_ROPE ? (g.N / 256) * 4 : ldp / 32;
  const int64_t warps = (int64_t)g.nz0 * g.nz1 * g.M * units;
  const dim3 grid((unsigned)((warps + 7) / 8));
  switch (g.epi_mode) {
    case EPI_STORE:
      PI0_CHECK_CUDA(launch_pdl(x3_reduce_kernel<EPI_STORE>, grid, dim3(256), 0, stream, 1, partial, S, g.nz0, g.nz1, g.M, g.N,
                                ldp, units, g.epi));
      break;
    case EPI_GEGLU:
      PI0_CHECK_CUDA(launch_pdl(x3_reduce_kernel<EPI_GEGLU>, grid, dim3(256), 0, stream, 1, partial, S, g.nz0, g.nz1, g.M, g.N,
                                ldp, units, g.epi));
      break;
    case EPI_ROPE:
      PI0_CHECK_CUDA(launch_pdl(x3_reduce_kernel<EPI_ROPE>, grid, dim3(256), 0, stream, 1, partial, S, g.nz0, g.nz1, g.M, g.N,
                                ldp, units, g.epi));
      break;
    case EPI_EULER:
      PI0_CHECK_CUDA(launch_pdl(x3_reduce_kernel<EPI_EULER>, grid, dim3(256), 0, stream, 1, partial, S, g.nz0, g.nz1, g.M, g.N,
                                ldp, units, g.epi));
      break;
    default:
      set_error("gemm_f32_reduce: unknown epilogue %d", g.epi_mode);
      return -2;
  }
  ++g_launch_count;
  return 0;
}

int gemm_tc_launch_f32out(const GemmProblem& g, int num_sms, void* partial, size_t partial_bytes, cudaStream_t stream) {
  PI0_REQUIRE(g.M > 0 && g.N > 0 && g.K > 0, "gemm_tc (fp32 out): empty problem M=%d N=%d K=%d", g.M, g.N, g.K);
  int bn = g.bn;
  if (g.epi_mode == EPI_STORE) {
    const long tm = (g.M + kBM - 1) / kBM, nz = (long)g.nz0 * g.nz1;
    const long t256 = tm * ((g.N + 255) / 256) * nz;
    if (partial != nullptr && g.N >= 256 && t256 * 2 <= num_sms && g.K >= 8 * kBK) {
      bn = 256;   // few wide tiles, the k range split over the idle SMs
    } else {      // tile width by wave count: the widest tile that still gives every SM work
      bn = 256;
      while (bn > 32 && (g.N < bn || tm * ((g.N + bn - 1) / bn) * nz < num_sms)) bn >>= 1;
    }
  }
  switch (g.epi_mode) {
    case EPI_STORE:
      switch (bn) {
        case 256: return launch_f32out<256, EPI_STORE>(g, num_sms, partial, partial_bytes, stream);
        case 128: return launch_f32out<128, EPI_STORE>(g, num_sms, partial, partial_bytes, stream);
        case 64: return launch_f32out<64, EPI_STORE>(g, num_sms, partial, partial_bytes, stream);
        default: return launch_f32out<32, EPI_STORE>(g, num_sms, partial, partial_bytes, stream);
      }
    case EPI_GEGLU:
      PI0_REQUIRE(g.bn == 256 && g.N % 256 == 0, "gemm_tc (fp32 out): GeGLU expects bn=256 packing, N=%d", g.N);
      return launch_f32out<256, EPI_GEGLU>(g, num_sms, partial, partial_bytes, stream);
    case EPI_ROPE:
      PI0_REQUIRE(g.N == (g.epi.n_q_heads + 2) * 256, "gemm_tc (fp32 out): RoPE epilogue expects N = (q_heads + 2) * 256");
      return launch_f32out<256, EPI_ROPE>(g, num_sms, partial, partial_bytes, stream);
    case EPI_EULER:
      return launch_f32out<32, EPI_EULER>(g, num_sms, partial, partial_bytes, stream);
  }
  set_error("gemm_tc (fp32 out): unknown epilogue %d", g.epi_mode);
  return -2;
}

}  // namespace pi0

// development aid (not part of the C ABI in include/pi0b200.h): cluster capacity the tile picker works with
extern "C" int pi0_debug_max_clusters(int sp) { return pi0::max_active_clusters(sp, 148); }
extern "C" int pi0_debug_max_pairs(void) { return pi0::max_active_pairs<256, pi0::EPI_STORE>(148); }

#ifdef PI0_TIMELINE
extern "C" int pi0_debug_pair_timeline(unsigned long long* out) {
  return cudaMemcpyFromSymbol(out, pi0::g_pair_timeline, sizeof(pi0::g_pair_timeline)) == cudaSuccess ? 0 : -1;
}
extern "C" int pi0_debug_timeline(unsigned long long* out) {
  return cudaMemcpyFromSymbol(out, pi0::g_timeline, sizeof(pi0::g_timeline)) == cudaSuccess ? 0 : -1;
}
#endif
