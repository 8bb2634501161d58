// Minimal unsigned multiword integer for host-side table construction (CRT constants, P, P*q_L,
// P^-1 mod 2^k).  Little-endian vector of 64-bit words, always normalised (no leading zero words).
// The reference does this arithmetic with libgcrypt MPIs (GPQHE/src/precomp.c:266-293,386-409); the
// engine must not depend on libgcrypt, so it carries its own ~100 lines.
#pragma once
#include <cstdint>
#include <vector>
#include <algorithm>

namespace hb {

typedef unsigned __int128 u128;

struct BigU {
  std::vector<uint64_t> w;
  BigU() {}
  explicit BigU(uint64_t v) { if (v) w.push_back(v); }
  void trim() { while (!w.empty() && w.back() == 0) w.pop_back(); }
  bool zero() const { return w.empty(); }
  unsigned bits() const {
    if (w.empty()) return 0;
    return 64u * (unsigned)(w.size() - 1) + (64u - (unsigned)__builtin_clzll(w.back()));
  }
  uint64_t word(size_t i) const { return i < w.size() ? w[i] : 0; }
  static int cmp(const BigU &a, const BigU &b) {
    if (a.w.size() != b.w.size()) return a.w.size() < b.w.size() ? -1 : 1;
    for (size_t i = a.w.size(); i-- > 0;)
      if (a.w[i] != b.w[i]) return a.w[i] < b.w[i] ? -1 : 1;
    return 0;
  }
  BigU mul_u64(uint64_t m) const {
    BigU r; r.w.resize(w.size() + 1);
    uint64_t c = 0;
    for (size_t i = 0; i < w.size(); i++) { u128 t = (u128)w[i] * m + c; r.w[i] = (uint64_t)t; c = (uint64_t)(t >> 64); }
    r.w[w.size()] = c; r.trim(); return r;
  }
  // quotient, remainder by a single word
  BigU divmod_u64(uint64_t d, uint64_t *rem) const {
    BigU q; q.w.resize(w.size());
    u128 r = 0;
    for (size_t i = w.size(); i-- > 0;) { u128 cur = (r << 64) | w[i]; q.w[i] = (uint64_t)(cur / d); r = cur % d; }
    if (rem) *rem = (uint64_t)r;
    q.trim(); return q;
  }
  uint64_t mod_u64(uint64_t d) const { uint64_t r; divmod_u64(d, &r); return r; }
  BigU shl(unsigned s) const {
    if (w.empty()) return *this;
    BigU r; unsigned ws = s / 64, bs = s % 64;
    r.w.assign(w.size() + ws + 1, 0);
    for (size_t i = 0; i < w.size(); i++) {
      r.w[i + ws] |= w[i] << bs;
      if (bs) r.w[i + ws + 1] |= w[i] >> (64 - bs);
    }
    r.trim(); return r;
  }
  BigU shr1() const {
    BigU r; r.w.resize(w.size());
    for (size_t i = 0; i < w.size(); i++) r.w[i] = (w[i] >> 1) | (i + 1 < w.size() ? (w[i + 1] << 63) : 0);
    r.trim(); return r;
  }
  static BigU mul(const BigU &a, const BigU &b) {
    BigU r; r.w.assign(a.w.size() + b.w.size() + 1, 0);
    for (size_t i = 0; i < a.w.size(); i++) {
      uint64_t c = 0;
      for (size_t j = 0; j < b.w.size(); j++) {
        u128 t = (u128)a.w[i] * b.w[j] + r.w[i + j] + c; r.w[i + j] = (uint64_t)t; c = (uint64_t)(t >> 64);
      }
      size_t k = i + b.w.size();
      while (c) { u128 t = (u128)r.w[k] + c; r.w[k] = (uint64_t)t; c = (uint64_t)(t >> 64); k++; }
    }
    r.trim(); return r;
  }
  // low `nw` words, zero padded
  std::vector<uint64_t> low(size_t nw) const {
    std::vector<uint64_t> r(nw, 0);
    for (size_t i = 0; i < nw && i < w.size(); i++) r[i] = w[i];
    return r;
  }
};

// (a * b) mod 2^(64*nw), operands given as nw-word vectors
inline std::vector<uint64_t> mul_low(const std::vector<uint64_t> &a, const std::vector<uint64_t> &b, size_t nw) {
  std::vector<uint64_t> r(nw, 0);
  for (size_t i = 0; i < nw; i++) {
    uint64_t c = 0;
    for (size_t j = 0; i + j < nw; j++) {
      u128 t = (u128)a[i] * b[j] + r[i + j] + c; r[i + j] = (uint64_t)t; c = (uint64_t)(t >> 64);
    }
  }
  return r;
}

// inverse of an odd multiword integer modulo 2^(64*nw) by Newton iteration x <- x*(2 - a*x)
inline std::vector<uint64_t> inv_mod_2k(const std::vector<uint64_t> &a, size_t nw) {
  std::vector<uint64_t> x(nw, 0);
  uint64_t x0 = 1;                       // a*x0 == 1 mod 2; doubles precision per step
  for (int i = 0; i < 6; i++) x0 *= 2 - a[0] * x0;
  x[0] = x0;
  for (size_t prec = 1; prec < nw; prec *= 2) {
    std::vector<uint64_t> ax = mul_low(a, x, nw);
    // t = 2 - ax
    std::vector<uint64_t> t(nw);
    uint64_t borrow = 0;
    for (size_t i = 0; i < nw; i++) {
      uint64_t m = (i == 0) ? 2 : 0;
      u128 d = (u128)m - ax[i] - borrow;
      t[i] = (uint64_t)d; borrow = (uint64_t)(d >> 64) & 1;
    }
    x = mul_low(x, t, nw);
  }
  return x;
}

}  // namespace hb
