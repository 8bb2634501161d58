// Compute ceiling of the lazy butterfly code as ptxas schedules it: radix-16 register tiles, no memory traffic.
// Reports butterflies per clock per SM.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../hectr_b200/csrc
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define HB_STREAM
#include "hb_dev.cuh"
using namespace hbdev;

template <int MODE, int MINB>
__global__ void __launch_bounds__(256, MINB) k(uint64_t *out, const ulonglong2 *tw, uint64_t p, int iters) {
  const Mod m = make_mod(p);
  uint64_t x[16];
#pragma unroll
  for (int t = 0; t < 16; t++) x[t] = (threadIdx.x * 16 + t) * 0x9E3779B97F4A7C15ull % p;
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < iters; it++) {
    if (MODE == 0) tile_fwd<4, false>(x, tw, 1u + (it & 3), m);
    if (MODE == 1) { tile_inv<4, true>(x, tw, 1u + (it & 3), m); }
    if (MODE == 2) {
#pragma unroll
      for (int t = 0; t < 16; t++) x[t] = fold(x[t], m);
      tile_fwd<4, false>(x, tw, 1u + (it & 3), m);
    }
    if (MODE == 0 && (it & 1)) {
#pragma unroll
      for (int t = 0; t < 16; t++) x[t] = fold(x[t], m);
    }
  }
  long long t1 = clock64();
  uint64_t r = 0;
#pragma unroll
  for (int t = 0; t < 16; t++) r += x[t];
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  if (threadIdx.x == 0) out[gridDim.x * blockDim.x + blockIdx.x] = t1 - t0;
}

// forward radix-16 tile with the 15 twiddles of tile K0 read from the lane-transposed table:
// slot s of tile K0 at ((K0 >> 5) * 15 + s) * 32 + (K0 & 31)
__device__ __forceinline__ void tile_fwd_tr(uint64_t (&x)[16], const ulonglong2 *__restrict__ tw2, unsigned K0, const Mod &m) {
  const ulonglong2 *base = tw2 + (size_t)(K0 >> 5) * 15 * 32 + (K0 & 31);
  int s = 0;
#pragma unroll
  for (int a = 0; a < 4; a++) {
    const int beta = 3 - a, half = 1 << beta;
#pragma unroll
    for (int u = 0; u < (8 >> beta); u++) {
      const ulonglong2 w = __ldg(&base[(s + u) * 32]);
#pragma unroll
      for (int t = 0; t < 16; t++)
        if ((t & half) == 0 && (t >> (beta + 1)) == u) ct_bfly<false>(x[t], x[t + half], w, m);
    }
    s += 8 >> beta;
  }
}

// MODE 3/4: the same tiles through padded shared memory with a barrier per pass (stride-256 and stride-1 passes)
template <int MODE, int MINB>
__global__ void __launch_bounds__(256, MINB) ks(uint64_t *out, const ulonglong2 *tw, uint64_t p, int iters) {
  extern __shared__ __align__(16) uint64_t sm[];
  const Mod m = make_mod(p);
  for (int i = threadIdx.x; i < 4096; i += 256) sm[pidx(i)] = (i * 0x9E3779B97F4A7C15ull) % p;
  __syncthreads();
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < iters; it++) {
    uint64_t x[16];
    if (MODE == 3) {
#pragma unroll
      for (int t = 0; t < 16; t++) x[t] = sm[pidx(threadIdx.x + t * 256)];
    } else {
#pragma unroll
      for (int t = 0; t < 16; t += 2) { ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(&sm[pidx(threadIdx.x * 16 + t)]); x[t] = v.x; x[t + 1] = v.y; }
    }
#pragma unroll
    for (int t = 0; t < 16; t++) x[t] = fold(x[t], m);
    if (MODE == 6) tile_fwd_tr(x, tw, 256u + threadIdx.x, m);
    else tile_fwd<4, false>(x, tw, MODE == 5 ? 256u + threadIdx.x : 1u + (it & 3), m);
    if (MODE == 3) {
#pragma unroll
      for (int t = 0; t < 16; t++) sm[pidx(threadIdx.x + t * 256)] = x[t];
    } else {
#pragma unroll
      for (int t = 0; t < 16; t += 2) *reinterpret_cast<ulonglong2 *>(&sm[pidx(threadIdx.x * 16 + t)]) = make_ulonglong2(x[t], x[t + 1]);
    }
    __syncthreads();
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = sm[pidx(threadIdx.x)];
  if (threadIdx.x == 0) out[gridDim.x * blockDim.x + blockIdx.x] = t1 - t0;
}

template <int MODE, int MINB>
void runs(const char *name, int ctas_per_sm) {
  const int sms = 148, grid = sms * ctas_per_sm, iters = 400;
  uint64_t *out; ulonglong2 *tw;
  const uint64_t p = 0x800000000004001ull;
  cudaMalloc(&out, (size_t)grid * 256 * 8 + grid * 8);
  ulonglong2 *h = new ulonglong2[4096];
  for (int i = 0; i < 4096; i++) { uint64_t w = (0x123456789abcdefull * (i + 1)) % p; h[i] = make_ulonglong2(w, (uint64_t)((((unsigned __int128)w) << 64) / p)); }
  cudaMalloc(&tw, 2 * 4096 * 16); cudaMemcpy(tw, h, 4096 * 16, cudaMemcpyHostToDevice); cudaMemcpy(tw + 4096, h, 4096 * 16, cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(ks<MODE, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, 36864);
  ks<MODE, MINB><<<grid, 256, 36864>>>(out, tw, p, iters);
  cudaDeviceSynchronize();
  ks<MODE, MINB><<<grid, 256, 36864>>>(out, tw, p, iters);
  cudaDeviceSynchronize();
  uint64_t *clk = new uint64_t[grid];
  cudaMemcpy(clk, out + (size_t)grid * 256, grid * 8, cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < grid; i++) avg += clk[i]; avg /= grid;
  double bf = (double)ctas_per_sm * 256 * 32.0 * iters;
  cudaFuncAttributes fa; cudaFuncGetAttributes(&fa, ks<MODE, MINB>);
  printf("%-28s regs=%3d warps/SM=%2d  %.2f bfly/clk/SM   (%s)\n", name, fa.numRegs, ctas_per_sm * 8, bf / avg, cudaGetErrorString(cudaGetLastError()));
  cudaFree(out); cudaFree(tw); delete[] clk; delete[] h;
}

template <int MODE, int MINB>
void run(const char *name, int ctas_per_sm) {
  const int sms = 148, grid = sms * ctas_per_sm, iters = 400;
  uint64_t *out; ulonglong2 *tw;
  const uint64_t p = 0x800000000004001ull;
  cudaMalloc(&out, (size_t)grid * 256 * 8 + grid * 8);
  ulonglong2 h[128];
  for (int i = 0; i < 128; i++) { uint64_t w = (0x123456789abcdefull * (i + 1)) % p; h[i] = make_ulonglong2(w, (uint64_t)((((unsigned __int128)w) << 64) / p)); }
  cudaMalloc(&tw, sizeof h); cudaMemcpy(tw, h, sizeof h, cudaMemcpyHostToDevice);
  k<MODE, MINB><<<grid, 256>>>(out, tw, p, iters);
  cudaDeviceSynchronize();
  k<MODE, MINB><<<grid, 256>>>(out, tw, p, iters);
  cudaDeviceSynchronize();
  uint64_t *clk = new uint64_t[grid];
  cudaMemcpy(clk, out + (size_t)grid * 256, grid * 8, cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < grid; i++) avg += clk[i]; avg /= grid;
  double bf = (double)ctas_per_sm * 256 * 32.0 * iters;      // butterflies per SM
  cudaFuncAttributes fa; cudaFuncGetAttributes(&fa, k<MODE, MINB>);
  printf("%-28s regs=%3d warps/SM=%2d  %.2f bfly/clk/SM\n", name, fa.numRegs, ctas_per_sm * 8, bf / avg);
  cudaFree(out); cudaFree(tw); delete[] clk;
}

int main() {
  run<0, 2>("fwd tile (fold every 2nd)", 1); run<0, 2>("fwd tile (fold every 2nd)", 2);
  run<0, 3>("fwd tile (fold every 2nd)", 3); run<0, 4>("fwd tile (fold every 2nd)", 4);
  run<1, 2>("inv tile", 1); run<1, 2>("inv tile", 2); run<1, 3>("inv tile", 3); run<1, 4>("inv tile", 4);
  runs<3, 3>("smem stride-256 + fold", 3); runs<3, 4>("smem stride-256 + fold", 4);
  runs<4, 3>("smem stride-1 + fold", 3); runs<4, 4>("smem stride-1 + fold", 4);
  runs<6, 3>("smem stride-1, transposed tw", 3); runs<6, 4>("smem stride-1, transposed tw", 4);
  runs<5, 3>("smem stride-1, own twiddles", 3); runs<5, 4>("smem stride-1, own twiddles", 4);
  return 0;
}
