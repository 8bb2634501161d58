// Register-only radix-16 forward tiles with variant lazy butterflies: how many butterflies per clock per SM does each
// formulation of  t = Y*w mod p (lazy)  reach?   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o modmul_variants modmul_variants.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t lo32(uint64_t x) { return (uint32_t)x; }
__device__ __forceinline__ uint32_t hi32(uint64_t x) { return (uint32_t)(x >> 32); }
__device__ __forceinline__ uint64_t pack64(uint32_t lo, uint32_t hi) { uint64_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(lo), "r"(hi)); return r; }
__device__ __forceinline__ uint32_t madlo(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
__device__ __forceinline__ uint64_t madwide(uint32_t a, uint32_t b, uint64_t c) { uint64_t r; asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(r) : "r"(a), "r"(b), "l"(c)); return r; }

struct Mod { uint64_t p, pn; uint32_t c; };

// V0: current product code: exact quotient (umulhi) + 6-IMAD remainder chain
__device__ __forceinline__ uint64_t mac_msub(uint64_t acc, uint64_t a, uint64_t w, uint64_t q, uint64_t pn) {
  uint32_t h = hi32(acc);
  h = madlo(hi32(q), lo32(pn), h);
  h = madlo(lo32(q), hi32(pn), h);
  h = madlo(hi32(a), lo32(w), h);
  h = madlo(lo32(a), hi32(w), h);
  uint64_t r = pack64(lo32(acc), h);
  r = madwide(lo32(q), lo32(pn), r);
  r = madwide(lo32(a), lo32(w), r);
  return r;
}
// quotient variants
__device__ __forceinline__ uint64_t q_exact(uint64_t y, uint64_t wq) { return __umul64hi(y, wq); }
// truncated: drops lo(y)*lo(wq) and the low halves of the cross products: q - 2 <= q~ <= q
__device__ __forceinline__ uint64_t q_trunc(uint64_t y, uint64_t wq) {
  const uint64_t t1 = (uint64_t)hi32(y) * lo32(wq);
  const uint64_t t2 = (uint64_t)lo32(y) * hi32(wq);
  uint64_t q = madwide(hi32(y), hi32(wq), (uint64_t)hi32(t1));
  return q + hi32(t2);
}
// remainder with the p = 2^59 + c structure: -q*p = -(q << 59) - q*c ; hi word: -(lo(q) << 27) - lo(q*c >> 32) ...
__device__ __forceinline__ uint64_t mac_msub_c(uint64_t acc, uint64_t a, uint64_t w, uint64_t q, const Mod &m) {
  // low 64 bits of acc + a*w - q*(2^59 + c)
  uint32_t h = hi32(acc) - (lo32(q) << 27);
  h = madlo(hi32(a), lo32(w), h);
  h = madlo(lo32(a), hi32(w), h);
  h = madlo(hi32(q), 0u - m.c, h);
  uint64_t r = pack64(lo32(acc), h);
  r = madwide(lo32(a), lo32(w), r);
  r -= (uint64_t)lo32(q) * m.c;
  return r;
}

template <int V>
__device__ __forceinline__ void bfly(uint64_t &X, uint64_t &Y, ulonglong2 w, const Mod &m) {
  const uint64_t x = X;
  uint64_t xn;
  if (V == 0) xn = mac_msub(x, Y, w.x, q_exact(Y, w.y), m.pn);
  if (V == 1) xn = mac_msub(x, Y, w.x, q_trunc(Y, w.y), m.pn);
  if (V == 2) xn = mac_msub_c(x, Y, w.x, q_exact(Y, w.y), m);
  if (V == 3) xn = mac_msub_c(x, Y, w.x, q_trunc(Y, w.y), m);
  Y = x + x + ((V & 1) ? 4 : 2) * m.p - xn;
  X = xn;
}

// 32-bit Harvey butterfly, p < 2^30, values kept in [0, 4p)   (two of these = one 59-bit limb's worth of modulus bits)
__device__ __forceinline__ void bfly32(uint32_t &X, uint32_t &Y, uint2 w, uint32_t p) {
  uint32_t x = X;
  x = min(x, x - 2 * p);
  const uint32_t q = __umulhi(Y, w.y);
  const uint32_t t = Y * w.x - q * p;
  X = x + t;
  Y = x - t + 2 * p;
}

template <int V, int MINB>
__global__ void __launch_bounds__(256, MINB) k64(uint64_t *out, const ulonglong2 *tw, uint64_t p, int iters) {
  Mod m; m.p = p; m.pn = 0 - p; m.c = (uint32_t)(p - (1ull << 59));
  uint64_t x[16];
#pragma unroll
  for (int t = 0; t < 16; t++) x[t] = (threadIdx.x * 16 + t) * 0x9E3779B97F4A7C15ull % p;
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < iters; it++) {
    const ulonglong2 *tb = tw + (it & 3) * 16;
#pragma unroll
    for (int a = 0; a < 4; a++) {
      const int half = 8 >> a;
#pragma unroll
      for (int t = 0; t < 16; t++)
        if ((t & half) == 0) bfly<V>(x[t], x[t + half], __ldg(&tb[(1 << a) + (t >> (4 - a))]), m);
    }
    if (it & 1) {
#pragma unroll
      for (int t = 0; t < 16; t++) {    // fold
        const uint32_t xh = hi32(x[t]);
        uint64_t r = pack64(lo32(x[t]), xh & 0x07ffffffu);
        r = madwide((xh >> 27) ^ 31u, m.c, r);
        x[t] = r + (m.p - 31ull * m.c);
      }
    }
  }
  long long t1 = clock64();
  uint64_t r = 0;
#pragma unroll
  for (int t = 0; t < 16; t++) r += x[t];
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  if (threadIdx.x == 0) out[gridDim.x * blockDim.x + blockIdx.x] = t1 - t0;
}

template <int MINB>
__global__ void __launch_bounds__(256, MINB) k32(uint64_t *out, const uint2 *tw, uint32_t p, int iters) {
  uint32_t x[32];
#pragma unroll
  for (int t = 0; t < 32; t++) x[t] = (threadIdx.x * 32 + t) * 0x9E3779B9u % p;
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < iters; it++) {
    const uint2 *tb = tw + (it & 3) * 32;
#pragma unroll
    for (int a = 0; a < 5; a++) {
      const int half = 16 >> a;
#pragma unroll
      for (int t = 0; t < 32; t++)
        if ((t & half) == 0) bfly32(x[t], x[t + half], __ldg(&tb[(1 << a) + (t >> (5 - a))]), p);
    }
  }
  long long t1 = clock64();
  uint32_t r = 0;
#pragma unroll
  for (int t = 0; t < 32; t++) r += x[t];
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
  if (threadIdx.x == 0) out[gridDim.x * blockDim.x + blockIdx.x] = t1 - t0;
}

template <class K, class TW, class P>
void run(const char *name, K kern, TW *tw, P p, int ctas_per_sm, double bfly_per_thread_iter) {
  const int sms = 148, grid = sms * ctas_per_sm, iters = 400;
  uint64_t *out;
  cudaMalloc(&out, (size_t)grid * 256 * 8 + grid * 8);
  kern<<<grid, 256>>>(out, tw, p, iters);
  cudaDeviceSynchronize();
  kern<<<grid, 256>>>(out, tw, p, iters);
  cudaDeviceSynchronize();
  uint64_t *clk = new uint64_t[grid];
  cudaMemcpy(clk, out + (size_t)grid * 256, grid * 8, cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < grid; i++) avg += clk[i]; avg /= grid;
  cudaFuncAttributes fa; cudaFuncGetAttributes(&fa, kern);
  printf("%-44s regs=%3d warps/SM=%2d  %.2f bfly/clk/SM  (%s)\n", name, fa.numRegs, ctas_per_sm * 8,
         ctas_per_sm * 256 * bfly_per_thread_iter * iters / avg, cudaGetErrorString(cudaGetLastError()));
  cudaFree(out); delete[] clk;
}

int main() {
  const uint64_t p = 0x800000000004001ull;
  ulonglong2 h[64];
  for (int i = 0; i < 64; i++) { uint64_t w = (0x123456789abcdefull * (i + 1)) % p; h[i] = make_ulonglong2(w, (uint64_t)((((unsigned __int128)w) << 64) / p)); }
  ulonglong2 *tw; cudaMalloc(&tw, sizeof h); cudaMemcpy(tw, h, sizeof h, cudaMemcpyHostToDevice);
  const uint32_t p32 = 1073479681u;   // < 2^30, == 1 mod 2^13
  uint2 h32[128];
  for (int i = 0; i < 128; i++) { uint32_t w = (uint32_t)((0x12345678ull * (i + 1)) % p32); h32[i] = make_uint2(w, (uint32_t)((((uint64_t)w) << 32) / p32)); }
  uint2 *tw32; cudaMalloc(&tw32, sizeof h32); cudaMemcpy(tw32, h32, sizeof h32, cudaMemcpyHostToDevice);
  for (int c = 2; c <= 4; c++) {
    if (c == 2) { run("V0 exact q, generic q*p", k64<0, 2>, tw, p, c, 32); run("V1 truncated q", k64<1, 2>, tw, p, c, 32);
                  run("V2 exact q, q*p via c", k64<2, 2>, tw, p, c, 32); run("V3 truncated q, q*p via c", k64<3, 2>, tw, p, c, 32);
                  run("32-bit Harvey (radix-32, per 30-bit limb)", k32<2>, tw32, p32, c, 80); }
    if (c == 3) { run("V0 exact q, generic q*p", k64<0, 3>, tw, p, c, 32); run("V1 truncated q", k64<1, 3>, tw, p, c, 32);
                  run("V2 exact q, q*p via c", k64<2, 3>, tw, p, c, 32); run("V3 truncated q, q*p via c", k64<3, 3>, tw, p, c, 32);
                  run("32-bit Harvey (radix-32, per 30-bit limb)", k32<3>, tw32, p32, c, 80); }
    if (c == 4) { run("V0 exact q, generic q*p", k64<0, 4>, tw, p, c, 32); run("V1 truncated q", k64<1, 4>, tw, p, c, 32);
                  run("V2 exact q, q*p via c", k64<2, 4>, tw, p, c, 32); run("V3 truncated q, q*p via c", k64<3, 4>, tw, p, c, 32);
                  run("32-bit Harvey (radix-32, per 30-bit limb)", k32<4>, tw32, p32, c, 80); }
  }
  return 0;
}
