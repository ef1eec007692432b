// Micro-benchmark (development aid): cost of the all-gather-by-reading step of denoise_loop.cu.
// 148 CTAs x 256 threads. Each rep: every CTA writes its 1/G slice of a W-word buffer (16-byte words), grid barrier,
// then every CTA reads the WHOLE buffer, 8 loads in flight per thread; the read phase is timed with %globaltimer per CTA.
// Variants: load flavour (relaxed.gpu v2.u64 | ld.cg v4.u32 | plain ld v4) x store flavour (relaxed.gpu | plain st).
#include <cooperative_groups.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
namespace cg = cooperative_groups;

__device__ __forceinline__ uint64_t gtimer() { uint64_t t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }

template <int LD>
__device__ __forceinline__ uint4 ld16(const uint4* p) {
  uint4 v;
  if (LD == 0) { uint64_t a, b; asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
    v = make_uint4((uint32_t)a, (uint32_t)(a >> 32), (uint32_t)b, (uint32_t)(b >> 32)); }
  if (LD == 1) asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  if (LD == 2) asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  if (LD == 3) asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
template <int ST>
__device__ __forceinline__ void st16(uint4* p, uint4 v) {
  if (ST == 0) { uint64_t a = v.x | ((uint64_t)v.y << 32), b = v.z | ((uint64_t)v.w << 32);
    asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(a), "l"(b) : "memory"); }
  if (ST == 1) asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

template <int LD, int ST>
__global__ void __launch_bounds__(256, 1) k(uint4* buf, int W, int reps, unsigned long long* t_out, unsigned* sink, int readers) {
  cg::grid_group grid = cg::this_grid();
  const int G = gridDim.x, cta = blockIdx.x, tid = threadIdx.x;
  unsigned acc = 0;
  unsigned long long tsum = 0;
  for (int r = 0; r < reps; ++r) {
    uint4* b = buf + (size_t)(r & 1) * W;
    const int per = (W + G - 1) / G;
    for (int i = tid; i < per; i += 256) {
      const int w = cta * per + i;
      if (w < W) st16<ST>(b + w, make_uint4(w, r + 1, w * 3, r + 1));
    }
    __threadfence();
    grid.sync();
    const uint64_t t0 = gtimer();
    if (cta < readers) {
      for (int i0 = tid; i0 < W; i0 += 256 * 8) {
        uint4 v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) if (i0 + j * 256 < W) v[j] = ld16<LD>(b + i0 + j * 256);
#pragma unroll
        for (int j = 0; j < 8; ++j) if (i0 + j * 256 < W) acc += v[j].x ^ v[j].y ^ v[j].z ^ v[j].w;
      }
    }
    __syncthreads();
    const uint64_t t1 = gtimer();
    if (r > 0) tsum += t1 - t0;
    grid.sync();
  }
  if (tid == 0) t_out[cta] = tsum / (reps - 1);
  if (acc == 0x12345678u) sink[0] = acc;
}

template <int LD, int ST>
void run(const char* name, int W, int readers) {
  int G = 148;
  uint4* buf; unsigned long long* t; unsigned* sink;
  cudaMalloc(&buf, (size_t)2 * W * 16); cudaMalloc(&t, G * 8); cudaMalloc(&sink, 4);
  cudaMemset(buf, 0, (size_t)2 * W * 16);
  int reps = 21;
  void* args[] = {&buf, &W, &reps, &t, &sink, &readers};
  cudaError_t e = cudaLaunchCooperativeKernel((void*)k<LD, ST>, dim3(G), dim3(256), args, 0, 0);
  cudaDeviceSynchronize();
  e = cudaGetLastError();
  std::vector<unsigned long long> h(G);
  cudaMemcpy(h.data(), t, G * 8, cudaMemcpyDeviceToHost);
  unsigned long long mx = 0, mn = ~0ull, sum = 0;
  for (int i = 0; i < readers; ++i) { mx = h[i] > mx ? h[i] : mx; mn = h[i] < mn ? h[i] : mn; sum += h[i]; }
  printf("%-28s W=%6d (%4d KB) readers=%3d: read phase avg %6.2f us  min %6.2f  max %6.2f  (%s)\n", name, W, W * 16 / 1024,
         readers, sum / 1e3 / readers, mn / 1e3, mx / 1e3, cudaGetErrorString(e));
  cudaFree(buf); cudaFree(t); cudaFree(sink);
}

int main() {
  for (int W : {1280, 2560, 5120}) {
    for (int readers : {1, 16, 148}) {
      run<0, 0>("ld.relaxed.gpu / st.relaxed", W, readers);
      run<1, 0>("ld.cg / st.relaxed", W, readers);
      run<2, 1>("ld (L1) / st", W, readers);
      run<3, 0>("ld.volatile / st.relaxed", W, readers);
    }
  }
  return 0;
}
