// See hb_tables.hpp.  Pure host code, no CUDA, no libgcrypt.
#include <cmath>
#include <string>
#include "hb_tables.hpp"

namespace hb {

uint64_t mulmod(uint64_t a, uint64_t b, uint64_t m) { return (uint64_t)(((u128)a * b) % m); }

uint64_t powmod(uint64_t a, uint64_t e, uint64_t m) {
  uint64_t r = 1 % m;
  a %= m;
  while (e) {
    if (e & 1) r = mulmod(r, a, m);
    a = mulmod(a, a, m);
    e >>= 1;
  }
  return r;
}

// Deterministic Miller-Rabin for 64-bit inputs (first twelve prime bases).  The reference uses 50
// random rand() witnesses (precomp.c:153-191); both decide primality correctly for these inputs, and
// tests/test_host_tables.py checks the resulting chain against the compiled reference.
bool is_prime_u64(uint64_t n) {
  if (n < 2) return false;
  static const uint64_t small[] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
  for (uint64_t s : small) {
    if (n == s) return true;
    if (n % s == 0) return false;
  }
  uint64_t d = n - 1;
  unsigned r = 0;
  while (!(d & 1)) { d >>= 1; r++; }
  for (uint64_t a : small) {
    uint64_t x = powmod(a, d, n);
    if (x == 1 || x == n - 1) continue;
    bool comp = true;
    for (unsigned i = 1; i < r; i++) {
      x = mulmod(x, x, n);
      if (x == n - 1) { comp = false; break; }
    }
    if (comp) return false;
  }
  return true;
}

// Least primitive root of prime p: the distinct prime factors of p-1 by trial division, then the
// smallest g with g^((p-1)/f) != 1 for every factor f (same definition as precomp.c:194-226).
uint64_t least_generator(uint64_t p) {
  uint64_t phi = p - 1, n = phi;
  std::vector<uint64_t> f;
  for (uint64_t d = 2; d * d <= n; d += (d == 2 ? 1 : 2)) {
    if (n % d == 0) {
      f.push_back(d);
      while (n % d == 0) n /= d;
    }
  }
  if (n > 1) f.push_back(n);
  for (uint64_t g = 2; g <= phi; g++) {
    bool ok = true;
    for (uint64_t q : f)
      if (powmod(g, phi / q, p) == 1) { ok = false; break; }
    if (ok) return g;
  }
  return 0;
}

uint64_t shoup_quotient(uint64_t w, uint64_t p) { return (uint64_t)((((u128)w) << 64) / p); }

// The device splits the stages [first, logn) into passes of R = ceil(rem / ceil(rem / 4)) stages (hb_dev.cuh PassSize);
// this returns the R of the last one (the stride-1 pass).
unsigned last_pass_stages(unsigned logn, unsigned first) {
  unsigned s = first, r = 0;
  while (s < logn) {
    const unsigned rem = logn - s, np = (rem + 3) / 4;
    r = (rem + np - 1) / np;
    s += r;
  }
  return r;
}

unsigned security_logq_bound(unsigned logn) {  // 128-bit classical column, precomp.c:57-64
  switch (logn) {
    case 10: return 27;
    case 11: return 54;
    case 12: return 109;
    case 13: return 218;
    case 14: return 438;
    case 15: return 881;
  }
  return 0;
}

static uint32_t bitrev32(uint32_t a) {
  a = ((a >> 1) & 0x55555555u) | ((a & 0x55555555u) << 1);
  a = ((a >> 2) & 0x33333333u) | ((a & 0x33333333u) << 2);
  a = ((a >> 4) & 0x0f0f0f0fu) | ((a & 0x0f0f0f0fu) << 4);
  a = ((a >> 8) & 0x00ff00ffu) | ((a & 0x00ff00ffu) << 8);
  return (a >> 16) | (a << 16);
}

std::string build_tables(HostTables &t, unsigned logn, unsigned logq, unsigned slots, unsigned logDelta) {
  const unsigned LOGP = 59;
  if (logn < 2 || logn > 17) return "logn out of range [2,17]";
  if (logDelta == 0 || logDelta > 62) return "Delta must be 2^k with 1 <= k <= 62";
  if (logq == 0) return "q must be 2^k with k >= 1";
  t.logn = logn; t.N = 1u << logn; t.logq = logq; t.logDelta = logDelta; t.slots = slots;
  unsigned logqub = security_logq_bound(logn);
  if (logn < 10 || logn > 15) logqub = logq;                    // precomp.c:338-340
  if (logq > logqub) return "The input modulus q is too large. Must guarantee log(q)<=log(qub).";
  if (slots == 0 || (slots & (slots - 1))) return "The slots must be the power of 2.";
  if (slots > t.N / 2) return "Must guarantee slots<=(n/2).";
  {  // assert(Delta > n + 2*Bclean), precomp.c:411-416,452
    double sigma = 3.1915382432114616, h = 64, n = t.N;
    double Bclean = 8 * std::sqrt(2.0) * sigma * n + 6 * sigma * std::sqrt(n) + 16 * sigma * std::sqrt(h * n);
    if (!(std::ldexp(1.0, (int)logDelta) > n + 2 * Bclean)) return "Delta too small for the fresh-noise bound";
  }
  t.dimub = (1 + logn + 4 * logqub) / LOGP + 1;                 // precomp.c:357
  t.L = logq / logDelta;                                        // precomp.c:391
  t.logq_l.assign(t.L + 1, 0);
  for (unsigned l = 0; l <= t.L; l++) t.logq_l[l] = logq - (t.L - l) * logDelta;
  t.dim = (logq + 1 + logn) / LOGP + 1;                         // precomp.c:401
  t.WC = (logq + 1 + 63) / 64;

  // ---- primes and per-prime constants ------------------------------------------------------------
  const unsigned N = t.N;
  t.primes.resize(t.dimub);
  t.tw.assign((size_t)t.dimub * N * 2, 0);
  t.itw.assign((size_t)t.dimub * N * 2, 0);
  t.pow64.assign((size_t)t.dimub * (t.WC + 1), 0);
  // lane-transposed copy for the stride-1 pass: R stages, 2^S tiles (S = logn - R)
  t.last_R = last_pass_stages(logn, 0);
  if (logn > 4 && last_pass_stages(logn, 4) != t.last_R) return "internal: the two pass chains disagree on the last pass";
  const unsigned trR = t.last_R, trS = logn - trR, trSlots = (1u << trR) - 1;
  const unsigned trBlocks = ((1u << trS) + 31) / 32;
  t.tr_stride = trBlocks * trSlots * 32;
  t.tw_tr.assign((size_t)t.dimub * t.tr_stride * 2, 0);
  t.itw_tr.assign((size_t)t.dimub * t.tr_stride * 2, 0);
  uint64_t p = (1ULL << LOGP) + 1;
  for (unsigned d = 0; d < t.dimub; d++) {
    do { p += 2ULL * N; } while (!is_prime_u64(p));
    if (p - (1ULL << LOGP) >= (1ULL << 28)) return "internal: prime chain leaves 2^59 + 2^28 (the kernels fold with c = p - 2^59 < 2^28)";
    PrimeConsts &c = t.primes[d];
    c.p = p;
    c.mu = (uint64_t)((((u128)1) << 120) / p);
    c.one_q = (uint64_t)((((u128)1) << 64) / p);
    c.r64 = (uint64_t)((((u128)1) << 64) % p);
    c.r64_q = shoup_quotient(c.r64, p);
    c.ninv = powmod(N, p - 2, p);
    c.ninv_q = shoup_quotient(c.ninv, p);
    uint64_t g = least_generator(p);
    c.root = powmod(g, (p - 1) / (2ULL * N), p);
    uint64_t rootinv = powmod(c.root, p - 2, p);
    uint64_t pw = 1, pwi = 1;
    uint64_t *tw = &t.tw[(size_t)d * N * 2], *itw = &t.itw[(size_t)d * N * 2];
    for (unsigned i = 0; i < N; i++) {
      uint32_t j = bitrev32(i) >> (32 - logn);
      tw[2 * j] = pw;   tw[2 * j + 1] = shoup_quotient(pw, p);
      itw[2 * j] = pwi; itw[2 * j + 1] = shoup_quotient(pwi, p);
      pw = mulmod(pw, c.root, p);
      pwi = mulmod(pwi, rootinv, p);
    }
    uint64_t *ttr = &t.tw_tr[(size_t)d * t.tr_stride * 2], *itr = &t.itw_tr[(size_t)d * t.tr_stride * 2];
    for (unsigned K = 0; K < (1u << trS); K++)
      for (unsigned e = 0; e < trR; e++)
        for (unsigned u = 0; u < (1u << e); u++) {
          const size_t src = ((((size_t)1 << trS) + K) << e) + u;
          const size_t dst = ((size_t)(K >> 5) * trSlots + ((1u << e) - 1) + u) * 32 + (K & 31);
          ttr[2 * dst] = tw[2 * src];   ttr[2 * dst + 1] = tw[2 * src + 1];
          itr[2 * dst] = itw[2 * src];  itr[2 * dst + 1] = itw[2 * src + 1];
        }
    uint64_t r = 1;
    for (unsigned w = 0; w <= t.WC; w++) { t.pow64[(size_t)d * (t.WC + 1) + w] = r; r = mulmod(r, c.r64, p); }
  }

  // ---- CRT constants for every limb count D ---------------------------------------------------------
  const unsigned DU = t.dimub, WC = t.WC;
  t.crt_inv.assign((size_t)(DU + 1) * DU, 0);
  t.crt_inv_q.assign((size_t)(DU + 1) * DU, 0);
  t.crt_phat.assign((size_t)(DU + 1) * DU * WC, 0);
  t.crt_P.assign((size_t)(DU + 1) * WC, 0);
  t.Pfull.assign(DU + 1, BigU(1));
  for (unsigned D = 1; D <= DU; D++) {
    t.Pfull[D] = t.Pfull[D - 1].mul_u64(t.primes[D - 1].p);
    const BigU &P = t.Pfull[D];
    std::vector<uint64_t> lowP = P.low(WC);
    for (unsigned w = 0; w < WC; w++) t.crt_P[(size_t)D * WC + w] = lowP[w];
    for (unsigned d = 0; d < D; d++) {
      uint64_t pd = t.primes[d].p, rem;
      BigU phat = P.divmod_u64(pd, &rem);
      uint64_t inv = powmod(phat.mod_u64(pd), pd - 2, pd);
      t.crt_inv[(size_t)D * DU + d] = inv;
      t.crt_inv_q[(size_t)D * DU + d] = shoup_quotient(inv, pd);
      std::vector<uint64_t> lo = phat.low(WC);
      for (unsigned w = 0; w < WC; w++) t.crt_phat[((size_t)D * DU + d) * WC + w] = lo[w];
    }
  }

  // ---- key-switch constants --------------------------------------------------------------------------
  if (t.dim > DU) return "internal: dim exceeds the prime chain";
  const BigU &P = t.Pfull[t.dim];
  t.nbitsP = P.bits();
  t.nbitsPqL = t.nbitsP + logq;
  t.dimevk = (logq + 1 + t.nbitsPqL + logn) / LOGP + 1;         // precomp.c:407
  // NB: for small rings dimevk can exceed dimub (e.g. logn=11, logq=54); the reference's hectx_init accepts
  // that too and only its key generation walks off the prime list.  Key-switch entry points check D <= dimub.
  t.WP = (t.nbitsP + 63) / 64 + 1;
  t.ks_inv.assign(t.dim, 0); t.ks_inv_q.assign(t.dim, 0);
  t.ks_phat.assign((size_t)t.dim * t.WP, 0);
  for (unsigned d = 0; d < t.dim; d++) {
    t.ks_inv[d] = t.crt_inv[(size_t)t.dim * DU + d];
    t.ks_inv_q[d] = t.crt_inv_q[(size_t)t.dim * DU + d];
    BigU phat = P.divmod_u64(t.primes[d].p, nullptr);
    std::vector<uint64_t> f = phat.low(t.WP);
    for (unsigned w = 0; w < t.WP; w++) t.ks_phat[(size_t)d * t.WP + w] = f[w];
  }
  t.ks_P = P.low(t.WP);
  t.ks_Ph = P.shr1().low(t.WP);
  t.ks_Pinv = inv_mod_2k(P.low(WC), WC);

  // ---- pre-scaling constants ---------------------------------------------------------------------------
  t.dec_s.assign((size_t)(DU + 1) * DU * (WC + 1) * 2, 0);
  t.ks_z.assign((size_t)(DU + 1) * t.dim * 2, 0);
  for (unsigned D = 1; D <= DU; D++) {
    for (unsigned d = 0; d < D; d++) {
      const uint64_t pd = t.primes[d].p, inv = t.crt_inv[(size_t)D * DU + d];
      for (unsigned w = 0; w <= WC; w++) {
        const uint64_t c = mulmod(t.pow64[(size_t)d * (WC + 1) + w], inv, pd);
        uint64_t *o = &t.dec_s[(((size_t)D * DU + d) * (WC + 1) + w) * 2];
        o[0] = c; o[1] = shoup_quotient(c, pd);
      }
      if (d < t.dim) {
        const uint64_t z = mulmod(t.ks_inv[d], powmod(inv, pd - 2, pd), pd);
        uint64_t *o = &t.ks_z[((size_t)D * t.dim + d) * 2];
        o[0] = z; o[1] = shoup_quotient(z, pd);
      }
    }
  }
  return "";
}

unsigned dim_keyswitch(const HostTables &t, unsigned l) { return (t.logq_l[l] + 1 + t.nbitsPqL + t.logn) / 59 + 1; }
unsigned dim_mul(const HostTables &t, unsigned l) { return ((t.logq_l[l] + 1) * 2 + t.logn) / 59 + 1; }
unsigned dim_mulpt(const HostTables &t, unsigned l, double nu) {
  // unsigned + double + unsigned, divided by an unsigned, plus one, truncated (he-mult.c:168)
  return (unsigned)(((t.logq_l[l] + 1) + std::log2(nu) + t.logn) / 59u + 1);
}

}  // namespace hb
