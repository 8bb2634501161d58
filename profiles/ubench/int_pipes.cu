// Integer-pipe throughput microbenchmark for sm_100a: warp-instructions per clock per SM for the opcodes the
// modular butterflies are made of.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o int_pipes int_pipes.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 2048
#define CHAINS 8

template <int OP>
__global__ void k(uint32_t *out, uint32_t s0, uint32_t s1) {
  uint32_t a[CHAINS], b[CHAINS];
  uint64_t w[CHAINS];
#pragma unroll
  for (int i = 0; i < CHAINS; i++) { a[i] = threadIdx.x * 3 + i + s0; b[i] = threadIdx.x + 7 * i + s1; w[i] = ((uint64_t)a[i] << 32) | b[i]; }
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < CHAINS; i++) {
      if (OP == 0) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b[i]), "r"(s1));            // IMAD
      if (OP == 1) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(a[i]), "r"(b[i]));        // IMAD.WIDE.U32
      if (OP == 2) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b[i]));                             // IADD3
      if (OP == 3) { asm volatile("add.cc.u32 %0, %0, %2; addc.u32 %1, %1, %3;" : "+r"(a[i]), "+r"(b[i]) : "r"(s0), "r"(s1)); }   // 64-bit add = 2
      if (OP == 4) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b[i]), "r"(s1));            // IMAD.HI
      if (OP == 5) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b[i]), "r"(s1));        // LOP3
      if (OP == 6) { asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b[i]), "r"(s1)); asm volatile("add.u32 %0, %0, %1;" : "+r"(b[i]) : "r"(s0)); }  // IMAD + IADD3 pair
      if (OP == 7) { asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(a[i]), "r"(b[i])); asm volatile("add.u32 %0, %0, %1;" : "+r"(b[i]) : "r"(s0)); }  // WIDE + IADD3 pair
      if (OP == 8) asm volatile("shf.r.wrap.b32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b[i]), "r"(s1));        // SHF
      if (OP == 9) asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(w[i]) : "r"(a[i]), "r"((uint32_t)w[i]));   // IMAD.WIDE no acc (dependent via lo word)
      if (OP == 10) { double d = __longlong_as_double(w[i]); asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d) : "d"(1.0000001), "d"(0.5)); w[i] = __double_as_longlong(d); }   // DFMA
      if (OP == 11) { double d = __longlong_as_double(w[i]); asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d) : "d"(1.0000001), "d"(0.5)); w[i] = __double_as_longlong(d);
                      asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b[i]), "r"(s1)); }        // DFMA + IMAD pair
      if (OP == 12) { asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[i]) : "r"(a[i]), "r"(b[i])); asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b[i]), "r"(s1)); }  // WIDE + IMAD pair
      if (OP == 13) { asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b[i]), "r"(s1)); asm volatile("add.u32 %0, %0, %1;" : "+r"(b[i]) : "r"(s0)); asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(b[i]) : "r"(a[i]), "r"(s1)); }  // IMAD + 2 ALU
      if (OP == 14) { asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b[i]), "r"(s1)); float f = __uint_as_float(b[i]); asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(1.0001f), "f"(0.5f)); b[i] = __float_as_uint(f); }  // IMAD + FFMA pair (heavy + lite?)
    }
  }
  long long t1 = clock64();
  uint32_t r = 0;
#pragma unroll
  for (int i = 0; i < CHAINS; i++) r += a[i] + b[i] + (uint32_t)w[i] + (uint32_t)(w[i] >> 32);
  out[blockIdx.x * blockDim.x + threadIdx.x] = r + (uint32_t)(t1 - t0);
  if (threadIdx.x == 0) ((long long *)(out + gridDim.x * blockDim.x))[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char *name, int per_iter, int ctas_per_sm, int threads) {
  int sms = 148;
  int grid = sms * ctas_per_sm;
  uint32_t *out;
  cudaMalloc(&out, (size_t)grid * threads * 4 + grid * 8);
  k<OP><<<grid, threads>>>(out, 1, 2);
  cudaDeviceSynchronize();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k<OP><<<grid, threads>>>(out, 1, 2);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long *clk = new long long[grid];
  cudaMemcpy(clk, out + (size_t)grid * threads, grid * 8, cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < grid; i++) avg += clk[i]; avg /= grid;
  double warp_instr_per_sm = (double)ctas_per_sm * (threads / 32) * ITERS * CHAINS * per_iter;
  printf("%-22s warps/SM=%3d  %.3f warp-instr/clk/SM (by clock64 %.0f clk)  %.3f ms\n", name, ctas_per_sm * threads / 32,
         warp_instr_per_sm / avg, avg, ms);
  cudaFree(out); delete[] clk;
}

int main() {
  for (int c = 1; c <= 4; c *= 2) {
    run<0>("IMAD (mad.lo)", 1, c, 512);
    run<1>("IMAD.WIDE.U32 acc", 1, c, 512);
    run<9>("IMAD.WIDE.U32 noacc", 1, c, 512);
    run<4>("IMAD.HI", 1, c, 512);
    run<2>("IADD3", 1, c, 512);
    run<3>("IADD3+IADD3.X", 2, c, 512);
    run<5>("LOP3", 1, c, 512);
    run<8>("SHF", 1, c, 512);
    run<6>("IMAD+IADD3", 2, c, 512);
    run<7>("WIDE+IADD3", 2, c, 512);
    run<10>("DFMA", 1, c, 512);
    run<11>("DFMA+IMAD", 2, c, 512);
    run<12>("WIDE+IMAD", 2, c, 512);
    run<13>("IMAD+IADD3+LOP3", 3, c, 512);
    run<14>("IMAD+FFMA", 2, c, 512);
  }
  return 0;
}
