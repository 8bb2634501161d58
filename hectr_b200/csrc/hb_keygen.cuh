// Switching-key generation on the device: the arithmetic of he_genswk, GPQHE/src/he-kem.c:74-119.
//
//   swk.p1 = a                                   (the caller's sample_uniform draw, below 2^nbits(P q_L))
//   swk.p0 = -a*sk + e + P*sp   mod P q_L         (poly_mul on dim' primes, centred mod P_dim', then centred mod P q_L)
//   both decomposed onto the dimevk key primes and transformed (the reference's [0,p] representation)
//
// Unlike everything on the evaluation path the modulus P q_L is not a power of two and the operands are wider than a
// top-level coefficient, so none of the evaluation tables apply: the constants below are built per call on the host
// (hb_bigint.hpp) and the kernels are generic in the word count.  This runs once per key, not per control step:
// clarity over speed (128-bit remainders, per-thread word arrays in local memory).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace hbkg {

constexpr int MAXW = 48;           // words of the widest intermediate (P_dim' plus guard words); checked on the host

struct Consts {                    // device pointers
  const uint64_t *primes;          // [max(D, dimevk)]
  const uint64_t *inv;             // [D]      (P_D/p_d)^-1 mod p_d
  const uint64_t *phat;            // [D][WX]  P_D/p_d
  const uint64_t *PD, *PDh;        // [WX]     P_D, floor(P_D/2)
  const uint64_t *M, *Mh, *Mlo;    // [WX]     P q_L, floor(M/2), floor(M/2) - M (two's complement)
  const uint64_t *P;               // [WX]     P
  unsigned D, WX, WA, N;
};

__device__ __forceinline__ uint64_t mulmod(uint64_t a, uint64_t b, uint64_t p) { return (uint64_t)(((unsigned __int128)a * b) % p); }

__device__ inline int ucmp(const uint64_t *a, const uint64_t *b, int W) {
  for (int w = W - 1; w >= 0; w--)
    if (a[w] != b[w]) return a[w] < b[w] ? -1 : 1;
  return 0;
}
__device__ inline int scmp(const uint64_t *a, const uint64_t *b, int W) {      // two's complement
  const int64_t ta = (int64_t)a[W - 1], tb = (int64_t)b[W - 1];
  if (ta != tb) return ta < tb ? -1 : 1;
  return ucmp(a, b, W - 1);
}
__device__ inline void add(uint64_t *a, const uint64_t *b, int W) {
  uint64_t c = 0;
  for (int w = 0; w < W; w++) { const unsigned __int128 t = (unsigned __int128)a[w] + b[w] + c; a[w] = (uint64_t)t; c = (uint64_t)(t >> 64); }
}
__device__ inline void sub(uint64_t *a, const uint64_t *b, int W) {
  uint64_t br = 0;
  for (int w = 0; w < W; w++) {
    const uint64_t x = a[w], y = b[w];
    const uint64_t d = x - y - br;
    br = (x < y) || (x == y && br);
    a[w] = d;
  }
}
__device__ inline void neg(uint64_t *a, int W) {
  uint64_t c = 1;
  for (int w = 0; w < W; w++) { const uint64_t t = ~a[w] + c; c = (c && t == 0); a[w] = t; }
}
// acc += m * k   /   acc -= m * k     (mod 2^(64 W))
__device__ inline void addmul(uint64_t *acc, const uint64_t *m, uint64_t k, int W) {
  uint64_t c = 0;
  for (int w = 0; w < W; w++) { const unsigned __int128 t = (unsigned __int128)m[w] * k + acc[w] + c; acc[w] = (uint64_t)t; c = (uint64_t)(t >> 64); }
}
__device__ inline void submul(uint64_t *acc, const uint64_t *m, uint64_t k, int W) {
  uint64_t c = 0, br = 0;
  for (int w = 0; w < W; w++) {
    const unsigned __int128 t = (unsigned __int128)m[w] * k + c;
    const uint64_t y = (uint64_t)t, x = acc[w];
    c = (uint64_t)(t >> 64);
    acc[w] = x - y - br;
    br = (x < y) || (x == y && br);
  }
}
// the value divided by 2^(64 (W - TOPW)), as a double: only the TOPW most significant words count, so that quotients of two
// numbers of the same width stay inside the double range whatever W is (P q_L reaches 1100 bits on deep chains)
constexpr int TOPW = 6;
__device__ inline double to_double_top(const uint64_t *a, int W, bool is_signed) {
  uint64_t mag[MAXW];
  const bool negative = is_signed && ((int64_t)a[W - 1] < 0);
  for (int w = 0; w < W; w++) mag[w] = a[w];
  if (negative) neg(mag, W);
  double d = 0;
  const int lo = W > TOPW ? W - TOPW : 0;
  for (int w = W - 1; w >= lo; w--) d = d * 18446744073709551616.0 + (double)mag[w];
  return negative ? -d : d;
}

// mpi_smod (types.c:108-113) for an arbitrary modulus: u <- u mod M in [0, M), minus M when it reaches floor(M/2).
// The quotient estimate is within one of the truth (M keeps at least 128 significant bits inside the window: the host checks
// it); the correction loops are bounded so that a wrong constant can never hang the device.
__device__ inline void smod_big(uint64_t *u, const Consts &C) {
  const int W = (int)C.WX;
  const long long t = llrint(to_double_top(u, W, true) / to_double_top(C.M, W, false));
  if (t > 0) submul(u, C.M, (uint64_t)t, W);
  else if (t < 0) addmul(u, C.M, (uint64_t)(-t), W);
  for (int it = 0; it < 4 && scmp(u, C.Mh, W) >= 0; it++) sub(u, C.M, W);
  for (int it = 0; it < 4 && scmp(u, C.Mlo, W) < 0; it++) add(u, C.M, W);
}

// rns_decompose (rns.c:37-48) for coefficients of ANY width: src[npoly][W][N] two's complement -> out[npoly][D][N] in [0,p)
__global__ void k_decompose_wide(const uint64_t *__restrict__ src, int W, size_t npoly, const uint64_t *__restrict__ primes, unsigned D,
                                 uint64_t *__restrict__ out, unsigned N) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= npoly * D * N) return;
  const unsigned i = gid % N, d = (gid / N) % D;
  const size_t q = gid / ((size_t)N * D);
  const uint64_t p = primes[d];
  const uint64_t *a = src + q * (size_t)W * N + i;
  uint64_t r = 0, c = 1;
  for (int w = W - 1; w >= 0; w--) {
    r = (uint64_t)((((unsigned __int128)r << 64) | a[(size_t)w * N]) % p);
    c = (uint64_t)(((unsigned __int128)c << 64) % p);                  // ends as 2^(64 W) mod p
  }
  if ((int64_t)a[(size_t)(W - 1) * N] < 0) r = (r >= c) ? r - c : r + p - c;
  out[gid] = r;
}

// out[q][d][i] = a[q][d][i] * s[d][i] mod p_d
__global__ void k_pointmul(const uint64_t *__restrict__ a, const uint64_t *__restrict__ s, const uint64_t *__restrict__ primes, unsigned D,
                           uint64_t *__restrict__ out, unsigned N, size_t npoly) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= npoly * D * N) return;
  const unsigned i = gid % N, d = (gid / N) % D;
  out[gid] = mulmod(a[gid] % primes[d], s[(size_t)d * N + i] % primes[d], primes[d]);
}

// prod[K][D][N]: coefficient-domain residues of a*sk.  Per coefficient: exact CRT centred mod P_D (poly_rns2mpi,
// poly.c:109-120), then p0 = smod(-x + e + P sp, P q_L) and p1 = smod(a, P q_L) (he-kem.c:97-102), WA words each.
__global__ void k_tail(const uint64_t *__restrict__ prod, Consts C, const int64_t *__restrict__ e, const int64_t *__restrict__ sp,
                       const uint64_t *__restrict__ a_words, uint64_t *__restrict__ out_v, uint64_t *__restrict__ out_a, unsigned K) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (size_t)K * C.N) return;
  const unsigned i = gid % C.N, kk = gid / C.N;
  const int W = (int)C.WX;
  uint64_t u[MAXW];
  for (int w = 0; w < W; w++) u[w] = 0;
  for (unsigned d = 0; d < C.D; d++) {
    const uint64_t p = C.primes[d];
    const uint64_t y = mulmod(prod[((size_t)kk * C.D + d) * C.N + i] % p, C.inv[d], p);
    addmul(u, C.phat + (size_t)d * W, y, W);
  }
  for (unsigned it = 0; it <= C.D && ucmp(u, C.PD, W) >= 0; it++) sub(u, C.PD, W);
  if (ucmp(u, C.PDh, W) >= 0) sub(u, C.PD, W);          // centred: now two's complement
  neg(u, W);
  {
    const int64_t ev = e[gid];
    uint64_t ext[MAXW];
    ext[0] = (uint64_t)ev;
    for (int w = 1; w < W; w++) ext[w] = (uint64_t)(ev >> 63);
    add(u, ext, W);
  }
  {
    const int64_t s = sp[gid];
    if (s > 0) addmul(u, C.P, (uint64_t)s, W);
    else if (s < 0) submul(u, C.P, (uint64_t)(-s), W);
  }
  smod_big(u, C);
  for (unsigned w = 0; w < C.WA; w++) out_v[((size_t)kk * C.WA + w) * C.N + i] = u[w];
  for (int w = 0; w < W; w++) u[w] = ((unsigned)w < C.WA) ? a_words[((size_t)kk * C.WA + w) * C.N + i] : 0;
  smod_big(u, C);
  for (unsigned w = 0; w < C.WA; w++) out_a[((size_t)kk * C.WA + w) * C.N + i] = u[w];
}

// Bound of an evaluation key (hb_evk_upload).  coef[npoly][D][N]: the key's limbs back in the coefficient domain, canonical.
// Per coefficient the exact CRT over the D key primes, centred mod P_D -- the integer he_genswk decomposed (he-kem.c:97-110), as
// long as it was below P_D / 2, which smod(., P q_L) guarantees -- and the maximum bit length of its magnitude over the whole key.
// A key made by he_genswk stays below P q_L / 2; limbs that are merely uniform residues (synthetic keys) reach P_D.
// With res32 != nullptr the same pass also reduces the lift modulo the m32 primes q32[] of the engine's 30-bit basis
// (res32[q][i32][N], canonical): the key's limbs for the hybrid-basis key switch (hb_dev32.cuh).
__global__ void k_lift_bits(const uint64_t *__restrict__ coef, Consts C, unsigned npoly, unsigned *__restrict__ out_max,
                            const uint32_t *__restrict__ q32, unsigned m32, uint32_t *__restrict__ res32) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (size_t)npoly * C.N) return;
  const unsigned i = gid % C.N, q = gid / C.N;
  const int W = (int)C.WX;
  uint64_t u[MAXW];
  for (int w = 0; w < W; w++) u[w] = 0;
  for (unsigned d = 0; d < C.D; d++) {
    const uint64_t p = C.primes[d];
    const uint64_t y = mulmod(coef[((size_t)q * C.D + d) * C.N + i] % p, C.inv[d], p);
    addmul(u, C.phat + (size_t)d * W, y, W);
  }
  for (unsigned it = 0; it <= C.D && ucmp(u, C.PD, W) >= 0; it++) sub(u, C.PD, W);
  bool negative = false;
  if (ucmp(u, C.PDh, W) >= 0) { sub(u, C.PD, W); neg(u, W); negative = true; }         // magnitude of the centred value
  if (res32) {
    for (unsigned j = 0; j < m32; j++) {
      const uint64_t qq = q32[j];
      uint64_t r = 0;
      for (int w = W - 1; w >= 0; w--) r = (uint64_t)((((unsigned __int128)r << 64) | u[w]) % qq);
      if (negative && r) r = qq - r;
      res32[((size_t)q * m32 + j) * C.N + i] = (uint32_t)r;
    }
  }
  unsigned bits = 0;
  for (int w = W - 1; w >= 0; w--)
    if (u[w]) { bits = 64u * (unsigned)w + (64u - (unsigned)__clzll((long long)u[w])); break; }
  atomicMax(out_max, bits);
}

}  // namespace hbkg
