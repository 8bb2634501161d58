// Host tables of the engine's INTERNAL 30-bit RNS basis (plaintext-product half of he_gemv).
//
// The reference multiplies polynomials through 59-bit primes (GPQHE/src/precomp.c:357-376) and reconstructs the exact
// integer product, centred modulo P_D (poly.c:109-120).  When that integer provably does not wrap P_D (gemv_plan's
// bound), ANY basis whose product Q >= P_D reconstructs the same integer.  A 30-bit Harvey butterfly is one
// IMAD.HI + two IMAD against six IMAD.WIDE + four IMAD for the 59-bit one: measured 16.8 vs 4.1 butterflies/clk/SM
// (profiles/ubench/modmul_variants.txt), i.e. twice the throughput per modulus bit.
//
//   primes   q_j : the largest primes below 2^30 with q_j == 1 (mod 2N), descending
//   twiddles     : psi_j^brv(i) (psi_j = g^((q-1)/2N), g least generator) with 32-bit Shoup quotients, same indexing
//                  as the 59-bit tables (entry k of stage s at 2^s + k), so the transform is the same negacyclic map
//   pow32        : 2^(32 w) mod q_j with quotients: word weights of the decomposition (32-bit digits)
//   CRT for D limbs: (Q_D/q_j)^-1 mod q_j, low words of Q_D/q_j and of Q_D, floor(2^61/q_j)
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include "hb_bigint.hpp"
#include "hb_tables.hpp"

namespace hb {

struct Prime32 {
  uint32_t p, ninv, ninv_q, mu;     // N^-1 mod p (+ quotient), floor(2^61 / p)
};

struct HostTables32 {
  unsigned logn = 0, N = 0, nprimes = 0, WC = 0, ndig = 0;
  std::vector<Prime32> primes;                 // [nprimes]
  std::vector<uint32_t> tw, itw;               // [nprimes][N][2]
  std::vector<uint32_t> tw_tr, itw_tr;         // [nprimes][tr_stride][2]: lane-transposed twiddles of the stride-1 pass (hb_dev32.cuh Tw32)
  unsigned tr_R = 0, tr_stride = 0;            // stages of that pass; entries per prime
  std::vector<uint32_t> pow32;                 // [nprimes][ndig + 1][2]   ndig = 2 WC digits, + the sign digit weight
  std::vector<uint32_t> crt_inv;               // [nprimes + 1][nprimes][2]  (value, quotient)
  std::vector<uint64_t> crt_qhat;              // [nprimes + 1][nprimes][WC]
  std::vector<uint64_t> crt_Q;                 // [nprimes + 1][WC]
  std::vector<unsigned> qbits;                 // [nprimes + 1] bits of Q_D
};

inline uint32_t shoup_quotient32(uint32_t w, uint32_t p) { return (uint32_t)((((uint64_t)w) << 32) / p); }

// limbs needed so that Q_D32 >= 2^bits
inline unsigned limbs32_for_bits(const HostTables32 &t, unsigned bits) {
  for (unsigned D = 1; D <= t.nprimes; D++)
    if (t.qbits[D] > bits) return D;
  return 0;
}

inline std::string build_tables32(HostTables32 &t, unsigned logn, unsigned WC, unsigned max_bits) {
  t.logn = logn; t.N = 1u << logn; t.WC = WC; t.ndig = 2 * WC;
  const unsigned N = t.N;
  t.nprimes = max_bits / 29 + 2;
  t.primes.resize(t.nprimes);
  t.tw.assign((size_t)t.nprimes * N * 2, 0);
  t.itw.assign((size_t)t.nprimes * N * 2, 0);
  t.pow32.assign((size_t)t.nprimes * (t.ndig + 1) * 2, 0);
  auto brv = [&](uint32_t x) { uint32_t r = 0; for (unsigned b = 0; b < logn; b++) r |= ((x >> b) & 1u) << (logn - 1 - b); return r; };
  uint64_t p = ((1ULL << 30) / (2ULL * N)) * (2ULL * N) + 1;
  for (unsigned d = 0; d < t.nprimes; d++) {
    do { p -= 2ULL * N; } while (p > (1ULL << 29) && !is_prime_u64(p));
    if (p <= (1ULL << 29)) return "internal: ran out of 30-bit primes == 1 mod 2N";
    Prime32 &c = t.primes[d];
    c.p = (uint32_t)p;
    c.ninv = (uint32_t)powmod(N, p - 2, p);
    c.ninv_q = shoup_quotient32(c.ninv, c.p);
    c.mu = (uint32_t)((1ULL << 61) / p);
    const uint64_t g = least_generator(p);
    const uint64_t root = powmod(g, (p - 1) / (2ULL * N), p), rootinv = powmod(root, p - 2, p);
    uint64_t pw = 1, pwi = 1;
    uint32_t *tw = &t.tw[(size_t)d * N * 2], *itw = &t.itw[(size_t)d * N * 2];
    for (unsigned i = 0; i < N; i++) {
      const uint32_t j = brv(i);
      tw[2 * j] = (uint32_t)pw;   tw[2 * j + 1] = shoup_quotient32((uint32_t)pw, c.p);
      itw[2 * j] = (uint32_t)pwi; itw[2 * j + 1] = shoup_quotient32((uint32_t)pwi, c.p);
      pw = mulmod(pw, root, p);
      pwi = mulmod(pwi, rootinv, p);
    }
    {
      // the stride-1 pass of the device transform (hb_dev32.cuh fwd32 / inv32: passes of PassSize<remaining>::R stages)
      unsigned S = 0, R = 0;
      for (;;) {
        const unsigned rem = logn - S, np = (rem + 3) / 4;
        R = (rem + np - 1) / np;
        if (S + R == logn) break;
        S += R;
      }
      const unsigned tiles = N >> R, slots = (1u << R) - 1;
      t.tr_R = R;
      t.tr_stride = ((tiles + 31) / 32) * slots * 32;
      if (d == 0) {
        t.tw_tr.assign((size_t)t.nprimes * t.tr_stride * 2, 0);
        t.itw_tr.assign((size_t)t.nprimes * t.tr_stride * 2, 0);
      }
      uint32_t *ttr = &t.tw_tr[(size_t)d * t.tr_stride * 2], *itr = &t.itw_tr[(size_t)d * t.tr_stride * 2];
      for (unsigned K = 0; K < tiles; K++)
        for (unsigned A = 0; A < R; A++)
          for (unsigned u = 0; u < (1u << A); u++) {
            const size_t src = ((size_t)((1u << S) + K) << A) + u;
            const size_t dst = ((size_t)(K >> 5) * slots + ((1u << A) - 1) + u) * 32 + (K & 31);
            ttr[2 * dst] = tw[2 * src]; ttr[2 * dst + 1] = tw[2 * src + 1];
            itr[2 * dst] = itw[2 * src]; itr[2 * dst + 1] = itw[2 * src + 1];
          }
    }
    uint64_t r = 1;
    const uint64_t r32 = (1ULL << 32) % p;
    for (unsigned w = 0; w <= t.ndig; w++) {
      uint32_t *o = &t.pow32[((size_t)d * (t.ndig + 1) + w) * 2];
      o[0] = (uint32_t)r; o[1] = shoup_quotient32((uint32_t)r, c.p);
      r = mulmod(r, r32, p);
    }
  }
  const unsigned DU = t.nprimes;
  t.crt_inv.assign((size_t)(DU + 1) * DU * 2, 0);
  t.crt_qhat.assign((size_t)(DU + 1) * DU * WC, 0);
  t.crt_Q.assign((size_t)(DU + 1) * WC, 0);
  t.qbits.assign(DU + 1, 0);
  BigU Q(1);
  for (unsigned D = 1; D <= DU; D++) {
    Q = Q.mul_u64(t.primes[D - 1].p);
    t.qbits[D] = Q.bits();
    std::vector<uint64_t> lowQ = Q.low(WC);
    for (unsigned w = 0; w < WC; w++) t.crt_Q[(size_t)D * WC + w] = lowQ[w];
    for (unsigned d = 0; d < D; d++) {
      const uint64_t pd = t.primes[d].p;
      BigU qhat = Q.divmod_u64(pd, nullptr);
      const uint32_t inv = (uint32_t)powmod(qhat.mod_u64(pd), pd - 2, pd);
      t.crt_inv[((size_t)D * DU + d) * 2] = inv;
      t.crt_inv[((size_t)D * DU + d) * 2 + 1] = shoup_quotient32(inv, (uint32_t)pd);
      std::vector<uint64_t> lo = qhat.low(WC);
      for (unsigned w = 0; w < WC; w++) t.crt_qhat[((size_t)D * DU + d) * WC + w] = lo[w];
    }
  }
  return "";
}

}  // namespace hb
