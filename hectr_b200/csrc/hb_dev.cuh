// Device-side arithmetic and the in-shared-memory negacyclic NTT core (sm_100a).
//
// Reference semantics being reproduced (results, not code):
//   forward  : GPQHE/src/ntt.c:37-52   Cooley-Tukey, natural in -> bit-reversed out, len = N/2..1
//   inverse  : GPQHE/src/ntt.c:54-73   Gentleman-Sande, len = 1..N/2, then * N^-1
//   modmul   : GPQHE/src/reduce.c:59-66 (Montgomery) / :88-106 (Barrett) -- here Shoup (fixed operand)
//              and a 2^120 Barrett (variable operands); a*b mod p is the same integer in [0,p).
//
// Arithmetic design (the path is bound by the integer pipes, not by HBM -- profiles/r01/):
//   * every prime is p = 2^59 + c with c < 2^28, so 2^64 = 31.99.. p: a 64-bit word holds lazily reduced
//     residues up to 31 p.  Forward butterflies never reduce (each stage adds 2p to the bound); inverse
//     butterflies double the bound of the sum leg and reduce only where the compile-time bound requires it.
//   * a * w - q * p is one chain of six IMADs accumulating onto the other butterfly leg (no carry fix-ups);
//     q = umulhi(a, w') with the precomputed Shoup quotient w' = floor(w 2^64 / p).
//
// Geometry.  One limb (N coefficients mod one prime) is processed by a cluster of C CTAs, each owning a
// contiguous chunk of M = N/C coefficients in (padded) shared memory; C = 1 up to N = 2^13, then 2/4/8 for
// 2^14/2^15/2^16 so that a 512 KiB limb still makes exactly one HBM round trip.  The first four forward
// stages (len N/2..N/16) are done in registers straight from global memory and scattered to the owning CTA's
// shared memory (distributed shared memory when C > 1); all later stages are chunk-local radix-2^R register
// tiles (R <= 4).  The inverse mirrors this: local tiles first, the last four stages gather across the cluster.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

// streaming accesses (evict-first) for data touched once, so the twiddle tables stay in L1
#if defined(HB_STREAM) && !defined(HB_NOSTREAM)
#define HB_LD(p) __ldcs(p)
#define HB_LDS2(p) __ldcs(p)
#define HB_ST(p, v) __stcs(p, v)
#else
#define HB_LD(p) (*(p))
#define HB_LDS2(p) __ldg(p)
#define HB_ST(p, v) (*(p) = (v))
#endif

namespace hbdev {

struct PrimeConsts {  // must match hb::PrimeConsts
  uint64_t p, mu, one_q, r64, r64_q, ninv, ninv_q, root;
};

struct Tables {  // device pointers, passed by value
  const PrimeConsts *primes;   // [dimub]
  const ulonglong2 *tw, *itw;  // [dimub][N] (w, floor(w 2^64 / p))
  const ulonglong2 *tw_tr, *itw_tr;   // [dimub][TR_STRIDE]: the stride-1 pass's twiddles, lane-transposed (hb_tables.hpp)
  unsigned tr_stride;
  const uint64_t *pow64;       // [dimub][WC+1]
  const uint64_t *crt_inv, *crt_inv_q;  // [dimub+1][dimub]
  const uint64_t *crt_phat;    // [dimub+1][dimub][WC]
  const uint64_t *crt_P;       // [dimub+1][WC]
  const uint64_t *ks_inv, *ks_inv_q;    // [dim]
  const uint64_t *ks_phat;     // [dim][WP]
  const uint64_t *ks_P, *ks_Ph;         // [WP]
  const uint64_t *ks_Pinv;     // [WC]
  const ulonglong2 *dec_s;     // [dimub+1][dimub][WC+1]  word weights of the scaled decomposition (hb_tables.hpp)
  const ulonglong2 *ks_z;      // [dimub+1][dim]
  unsigned N, logn, dimub, WC, WP, dim;
};

// ---------------------------------------------------------------------------------------------------------
// 32-bit building blocks
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t lo32(uint64_t x) { return (uint32_t)x; }
__device__ __forceinline__ uint32_t hi32(uint64_t x) { return (uint32_t)(x >> 32); }
__device__ __forceinline__ uint64_t pack64(uint32_t lo, uint32_t hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(lo), "r"(hi));
  return r;
}
__device__ __forceinline__ uint32_t madlo(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}
__device__ __forceinline__ uint64_t madwide(uint32_t a, uint32_t b, uint64_t c) {
  uint64_t r;
  asm("mad.wide.u32 %0, %1, %2, %3;" : "=l"(r) : "r"(a), "r"(b), "l"(c));
  return r;
}

// The modulus as the kernels carry it: p, -p mod 2^64, and c = p - 2^59.
struct Mod {
  uint64_t p, pn;
  uint32_t c;
};
__device__ __forceinline__ Mod make_mod(uint64_t p) {
  Mod m;
  m.p = p;
  m.pn = 0 - p;
  m.c = (uint32_t)(p - (1ULL << 59));
  return m;
}

// acc + a*w - q*p  (mod 2^64): six IMADs, the low product pair accumulates in 64 bits.
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

__device__ __forceinline__ uint64_t csub(uint64_t x, uint64_t m) { return x >= m ? x - m : x; }

// a*w mod p in [0, 2p) for ANY 64-bit a, given wq = floor(w 2^64 / p), w < p.
__device__ __forceinline__ uint64_t shoup_lazy(uint64_t a, uint64_t w, uint64_t wq, const Mod &m) {
  return mac_msub(0, a, w, __umul64hi(a, wq), m.pn);
}
__device__ __forceinline__ uint64_t shoup(uint64_t a, uint64_t w, uint64_t wq, const Mod &m) {
  return csub(shoup_lazy(a, w, wq, m), m.p);
}
// same, from the bare modulus (cold paths)
__device__ __forceinline__ uint64_t shoup(uint64_t a, uint64_t w, uint64_t wq, uint64_t p) {
  uint64_t q = __umul64hi(a, wq);
  return csub(a * w - q * p, p);
}

// Any 64-bit x -> the same residue in (0, 2p):  x = h 2^59 + r  ==  r - h c  (mod p); returned as
// r + (31 - h) c + (p - 31 c).
__device__ __forceinline__ uint64_t fold(uint64_t x, const Mod &m) {
  const uint32_t xh = hi32(x);
  const uint32_t e = (xh >> 27) ^ 31u;                       // 31 - h
  uint64_t r = pack64(lo32(x), xh & 0x07ffffffu);
  r = madwide(e, m.c, r);
  return r + (m.p - 31ull * m.c);
}
// any 64-bit x -> [0,p)
__device__ __forceinline__ uint64_t canon(uint64_t x, const Mod &m) { return csub(fold(x, m), m.p); }

// any 64-bit word -> [0,p)  (cold paths)
__device__ __forceinline__ uint64_t reduce_word(uint64_t a, const PrimeConsts &c) {
  uint64_t q = __umul64hi(a, c.one_q);
  return csub(a - q * c.p, c.p);
}

// a*b mod p in [0,p) for a,b <= p < 2^60 (operands may equal p: SURVEY F8).  mu = floor(2^120/p).
__device__ __forceinline__ uint64_t mulmod_barrett(uint64_t a, uint64_t b, uint64_t p, uint64_t mu) {
  uint64_t lo = a * b, hi = __umul64hi(a, b);              // z < 2^121
  uint64_t zh = (hi << 5) | (lo >> 59);                    // floor(z / 2^59) < 2^62
  uint64_t qhat = (__umul64hi(zh, mu) << 3) | ((zh * mu) >> 61);   // floor(zh*mu / 2^61), within 2 of floor(z/p)
  uint64_t r = lo - qhat * p;                              // exact mod 2^64; true value in [0, 3p)
  r = csub(r, 2 * p);
  return csub(r, p);
}

__device__ __forceinline__ uint64_t addmod(uint64_t a, uint64_t b, uint64_t p) { return csub(a + b, p); }

// Forward butterfly (ntt.c:44-48).  LAZY (default): X' = X + wY, Y' = X - wY + 2p with wY in [0,2p): the bound
// of both legs grows by 2p and nothing is reduced.  STRICT: the reference's exact data flow, values in
// [0,p] including its "sum == p stays p" quirk.
template <bool STRICT>
__device__ __forceinline__ void ct_bfly(uint64_t &X, uint64_t &Y, ulonglong2 w, const Mod &m) {
  if (STRICT) {
    uint64_t t = shoup(Y, w.x, w.y, m);
    uint64_t x = X;
    Y = (x >= t) ? x - t : x - t + m.p;
    X = (x <= m.p - t) ? x + t : x + t - m.p;
  } else {
    const uint64_t x = X;
    const uint64_t xn = mac_msub(x, Y, w.x, __umul64hi(Y, w.y), m.pn);   // x + t
    Y = x + x + 2 * m.p - xn;                                             // x - t + 2p
    X = xn;
  }
}

// Inverse butterfly (ntt.c:62-67): X' = X + Y, Y' = (X - Y) w.  BY = bound of Y in units of p.
template <int BY>
__device__ __forceinline__ void gs_bfly(uint64_t &X, uint64_t &Y, ulonglong2 w, const Mod &m) {
  const uint64_t s = X + Y;
  uint32_t h = madlo((uint32_t)BY, hi32(m.p), hi32(X));
  const uint64_t d = madwide((uint32_t)BY, lo32(m.p), pack64(lo32(X), h)) - Y;   // X + BY p - Y
  X = s;
  Y = shoup_lazy(d, w.x, w.y, m);
}

// ---------------------------------------------------------------------------------------------------------
// geometry
// ---------------------------------------------------------------------------------------------------------
template <int LOGN>
struct Cfg {
  static constexpr int N = 1 << LOGN;
  static constexpr int LOGC = (LOGN > 13) ? (LOGN - 13) : 0;
  static constexpr int C = 1 << LOGC;          // CTAs per limb (cluster size)
  static constexpr int LOGM = LOGN - LOGC;
  static constexpr int M = 1 << LOGM;          // coefficients per CTA
  static constexpr int TILES16 = (M >= 16) ? M / 16 : 1;
  static constexpr int THREADS = TILES16 > 256 ? 256 : TILES16;   // radix-16 tiles are looped over when M > 4096
  static constexpr int SMEM_WORDS = M + (M >> 4) * 2;   // 2 pad words per 16 (keeps 16-byte alignment)
  static constexpr int SMEM_BYTES = SMEM_WORDS * 8;
  #ifndef HB_MINB12
#define HB_MINB12 3
#endif
#ifndef HB_MINB_BIG
#define HB_MINB_BIG 3     /* 3 x 74 KB of shared memory and 80 registers: three resident CTAs for N >= 2^14 too */
#endif
  static constexpr int MINBLOCKS = (LOGN <= 12) ? HB_MINB12 : (LOGN == 13 ? 3 : HB_MINB_BIG);
};

__device__ __forceinline__ int pidx(int i) { return i + ((i >> 4) << 1); }

// Twiddle tables of one prime and one direction.  nat: index = bit-reversed exponent (entry k of stage s at 2^s + k).
// tr: the stride-1 pass (last forward pass, first inverse pass) gives every thread its own 2^R - 1 twiddles; read
// from `nat` that is up to 32 cache lines per warp load, so a lane-transposed copy keeps each load at 4 lines:
//   slot (2^e - 1) + u of global tile K  at  ((K >> 5) (2^R - 1) + slot) 32 + (K & 31),  holding nat[((2^S + K) << e) + u]
// (e = 0 is the coarsest stage of the tile, S = log2 N - R).  Built by hb::build_tables.
struct Tw {
  const ulonglong2 *nat, *tr;
};
__device__ __forceinline__ Tw make_tw(const Tables &T, unsigned d, bool inverse) {
  Tw t;
  t.nat = (inverse ? T.itw : T.tw) + (size_t)d * T.N;
  t.tr = (inverse ? T.itw_tr : T.tw_tr) + (size_t)d * T.tr_stride;
  return t;
}
// twiddle (e, u) of a tile: natural table addressed by the tile's coarsest block index K0, or the transposed copy
template <int R, bool TR>
struct TileTw {
  const ulonglong2 *base;
  unsigned K0;
  __device__ __forceinline__ TileTw(const Tw &tw, unsigned K0_, unsigned K) : K0(K0_) {
    base = TR ? tw.tr + (size_t)(K >> 5) * (((1 << R) - 1) * 32) + (K & 31) : tw.nat;
  }
  __device__ __forceinline__ ulonglong2 get(int e, int u) const {
    return TR ? __ldg(&base[(((1 << e) - 1) + u) * 32]) : __ldg(&base[(K0 << e) + u]);
  }
};

// Forward tile: 2^R elements, stages beta = R-1 .. 0.  Twiddle index of the butterfly at tile-bit `beta` for tile
// element t, tile whose coarsest block index is K0 (= 2^S + block):  k = (K0 << (R-1-beta)) + (t >> (beta+1)).
template <int R, bool STRICT, class TW>
__device__ __forceinline__ void tile_fwd(uint64_t (&x)[1 << R], const TW &tw, const Mod &m) {
#pragma unroll
  for (int a = 0; a < R; a++) {
    const int beta = R - 1 - a;
    const int half = 1 << beta;
#pragma unroll
    for (int t = 0; t < (1 << R); t++) {
      if ((t & half) == 0) {
        ulonglong2 w = tw.get(R - 1 - beta, t >> (beta + 1));
        ct_bfly<STRICT>(x[t], x[t + half], w, m);
      }
    }
  }
}

// Inverse tile: stages beta = 0 .. R-1.  Every element enters below 2p.  The sum leg doubles its bound each
// stage, the product leg returns to 2p; bounds are tracked at compile time (the loops are fully unrolled) and a
// leg is folded back below 2p only when the next sum could pass 31p.  With FOLD_OUT every element leaves below
// 2p again (needed between passes); the last pass multiplies by N^-1 instead, which accepts any input.
template <int R, bool FOLD_OUT, class TW>
__device__ __forceinline__ void tile_inv(uint64_t (&x)[1 << R], const TW &tw, const Mod &m) {
  int bnd[1 << R];
#pragma unroll
  for (int t = 0; t < (1 << R); t++) bnd[t] = 2;
#pragma unroll
  for (int a = 0; a < R; a++) {
    const int beta = a;
    const int half = 1 << beta;
#pragma unroll
    for (int t = 0; t < (1 << R); t++) {
      if ((t & half) == 0) {
        if (bnd[t] + bnd[t + half] > 31) {
          if (bnd[t] > 2) { x[t] = fold(x[t], m); bnd[t] = 2; }
          if (bnd[t + half] > 2) { x[t + half] = fold(x[t + half], m); bnd[t + half] = 2; }
        }
        ulonglong2 w = tw.get(R - 1 - beta, t >> (beta + 1));
        const int by = bnd[t + half];
        if (by == 2) gs_bfly<2>(x[t], x[t + half], w, m);
        else if (by == 4) gs_bfly<4>(x[t], x[t + half], w, m);
        else if (by == 8) gs_bfly<8>(x[t], x[t + half], w, m);
        else gs_bfly<16>(x[t], x[t + half], w, m);
        bnd[t] += by;
        bnd[t + half] = 2;
      }
    }
  }
  if (FOLD_OUT) {
#pragma unroll
    for (int t = 0; t < (1 << R); t++)
      if (bnd[t] > 2) x[t] = fold(x[t], m);
  }
}

// One chunk-local pass: stages [S, S+R) in forward numbering (len = N>>(S+1) ...), on this CTA's chunk.
// FOLD_IN (forward only): fold every element below 2p on entry (bound bookkeeping of ntt_forward).
template <int LOGN, int S, int R, bool INV, bool STRICT, bool FOLD_IN>
__device__ __forceinline__ void local_pass(uint64_t *sm, const Tw &tw, const Mod &m, unsigned chunk) {
  typedef Cfg<LOGN> G;
  constexpr int LOGST = LOGN - S - R;           // log2 of the element stride inside a tile
  constexpr int ST = 1 << LOGST;
  constexpr int TILES = G::M >> R;
  constexpr int LOGBLK = LOGN - S;              // log2 of the block a tile spans
  const unsigned blk0 = (1u << S) + (chunk << (G::LOGM - LOGBLK));
  for (int tile = threadIdx.x; tile < TILES; tile += G::THREADS) {
    const int b = tile >> LOGST, o = tile & (ST - 1);
    const int g = (b << LOGBLK) + o;
    uint64_t x[1 << R];
    // padded offsets as one base plus compile-time constants: pidx(g + t ST) = pidx(g) + t (ST + ST/8) when
    // 16 | ST, and pidx(g + t) = pidx(g) + t inside one group of 16 (g is a multiple of 2^R, R <= 4)
    const int pg = pidx(g);
    if (ST == 1) {
#pragma unroll
      for (int t = 0; t < (1 << R); t += 2) {
        ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(&sm[pg + t]);
        x[t] = v.x; x[t + 1] = v.y;
      }
    } else if (ST % 16 == 0) {
#pragma unroll
      for (int t = 0; t < (1 << R); t++) x[t] = sm[pg + t * (ST + ST / 8)];
    } else {
#pragma unroll
      for (int t = 0; t < (1 << R); t++) x[t] = sm[pidx(g + t * ST)];
    }
    if (FOLD_IN) {
#pragma unroll
      for (int t = 0; t < (1 << R); t++) x[t] = fold(x[t], m);
    }
    const TileTw<R, ST == 1> ttw(tw, blk0 + b, (chunk << (G::LOGM - R)) + tile);
    if (INV) tile_inv<R, true>(x, ttw, m);
    else tile_fwd<R, STRICT>(x, ttw, m);
    if (ST == 1) {
#pragma unroll
      for (int t = 0; t < (1 << R); t += 2)
        *reinterpret_cast<ulonglong2 *>(&sm[pg + t]) = make_ulonglong2(x[t], x[t + 1]);
    } else if (ST % 16 == 0) {
#pragma unroll
      for (int t = 0; t < (1 << R); t++) sm[pg + t * (ST + ST / 8)] = x[t];
    } else {
#pragma unroll
      for (int t = 0; t < (1 << R); t++) sm[pidx(g + t * ST)] = x[t];
    }
  }
}

template <int REM>
struct PassSize {  // stages per pass when REM stages remain: ceil(REM / ceil(REM/4))
  static constexpr int NP = (REM + 3) / 4;
  static constexpr int R = (REM + NP - 1) / NP;
};

// BND = compile-time bound (units of p) of the values entering stage S.
template <int LOGN, int S, bool STRICT, int BND>
__device__ __forceinline__ void fwd_local(uint64_t *sm, const Tw &tw, const Mod &m, unsigned chunk) {
  if constexpr (S < LOGN) {
    constexpr int R = PassSize<LOGN - S>::R;
    constexpr bool FOLD = !STRICT && (BND + 2 * R > 31);
    constexpr int OUT = (FOLD ? 2 : BND) + 2 * R;
    local_pass<LOGN, S, R, false, STRICT, FOLD>(sm, tw, m, chunk);
    __syncthreads();
    fwd_local<LOGN, S + R, STRICT, OUT>(sm, tw, m, chunk);
  }
}
// bound of the values ntt_forward leaves in shared memory, for inputs below BND0 p
template <int LOGN, int S, int BND>
struct FwdBound {
  static constexpr int R = PassSize<(LOGN - S > 0) ? LOGN - S : 1>::R;
  static constexpr int value = (S >= LOGN) ? BND : FwdBound<LOGN, (S >= LOGN) ? S : S + R, ((BND + 2 * R > 31) ? 2 : BND) + 2 * R>::value;
};
template <int LOGN, int BND>
struct FwdBound<LOGN, LOGN, BND> { static constexpr int value = BND; };

template <int LOGN, int S>
__device__ __forceinline__ void inv_local(uint64_t *sm, const Tw &itw, const Mod &m, unsigned chunk) {
  if constexpr (S < LOGN) {
    constexpr int R = PassSize<LOGN - S>::R;
    inv_local<LOGN, S + R>(sm, itw, m, chunk);
    local_pass<LOGN, S, R, true, false, false>(sm, itw, m, chunk);
    __syncthreads();
  }
}

template <int LOGN>
__device__ __forceinline__ unsigned chunk_rank() {
  if constexpr (Cfg<LOGN>::C > 1) return cg::this_cluster().block_rank();
  else return 0;
}

// Forward transform of one limb into the cluster's shared memory.  load(gi) returns coefficient gi of the
// natural-order input, below BND0 p (BND0 <= 7; STRICT: <= p).  On return (all CTAs synchronised) this CTA's chunk
// holds outputs [chunk*M, (chunk+1)*M) of the bit-reversed-order result as lazily reduced 64-bit words (any
// value; fold()/canon() them, or feed them to a Shoup multiplication) -- STRICT: the exact reference values.
template <int LOGN, bool STRICT, int BND0, class Load>
__device__ __forceinline__ void ntt_forward(uint64_t *sm, Load load, const Tw &tw, const Mod &m) {
  typedef Cfg<LOGN> G;
  static_assert(BND0 + 8 <= 31, "entry bound too large");
  const unsigned chunk = chunk_rank<LOGN>();
  if constexpr (LOGN >= 4) {
    constexpr int ST0 = G::N / 16;
    constexpr int PER = G::TILES16 / G::THREADS;     // radix-16 tiles per thread in the global pass
    if constexpr (G::C > 1) cg::this_cluster().sync();   // every CTA of the cluster is resident before remote writes
#pragma unroll 1
    for (int r = 0; r < PER; r++) {
      const int tau = chunk * G::TILES16 + r * G::THREADS + threadIdx.x;
      uint64_t x[16];
#pragma unroll
      for (int t = 0; t < 16; t++) x[t] = load(tau + t * ST0);
      tile_fwd<4, STRICT>(x, TileTw<4, false>(tw, 1u, 0u), m);
      if constexpr (G::C > 1) {
        cg::cluster_group cl = cg::this_cluster();
#pragma unroll
        for (int t = 0; t < 16; t++) {
          const int gi = tau + t * ST0;
          uint64_t *dst = cl.map_shared_rank(sm, gi >> G::LOGM);
          dst[pidx(gi & (G::M - 1))] = x[t];
        }
      } else {
#pragma unroll
        for (int t = 0; t < 16; t++)
          sm[(ST0 % 16 == 0) ? pidx(tau) + t * (ST0 + ST0 / 8) : pidx(tau + t * ST0)] = x[t];
      }
    }
    if constexpr (G::C > 1) cg::this_cluster().sync();
    else __syncthreads();
    fwd_local<LOGN, 4, STRICT, BND0 + 8>(sm, tw, m, chunk);
  }
}

// C == 1 only: forward transform of a limb that already sits in shared memory (natural order, padded with pidx,
// values below BND0 p, visible to the whole CTA after the barrier this function starts with).
template <int LOGN, bool STRICT, int BND0>
__device__ __forceinline__ void ntt_forward_resident(uint64_t *sm, const Tw &tw, const Mod &m) {
  static_assert(Cfg<LOGN>::C == 1, "resident forward needs the whole limb in one CTA");
  __syncthreads();
  fwd_local<LOGN, 0, STRICT, BND0>(sm, tw, m, 0);
}

// Inverse transform of one limb held in the cluster's shared memory (values below 2p, bit-reversed order);
// store(gi, v) receives natural-order coefficient gi:
//   INV_LAZY   scaled by N^-1, in [0,2p)        INV_CANON  scaled by N^-1, canonical [0,p)
//   INV_RAW    NOT scaled (the caller folded N^-1 into a fixed multiplicand beforehand), folded into (0,2p)
enum { INV_LAZY = 0, INV_CANON = 1, INV_RAW = 2 };
template <int LOGN, int MODE, class Store>
__device__ __forceinline__ void ntt_inverse(uint64_t *sm, Store store, const Tw &itw, const PrimeConsts &pc,
                                            const Mod &m) {
  typedef Cfg<LOGN> G;
  const unsigned chunk = chunk_rank<LOGN>();
  inv_local<LOGN, 4>(sm, itw, m, chunk);
  constexpr int ST0 = G::N / 16;
  constexpr int PER = G::TILES16 / G::THREADS;
  if constexpr (G::C > 1) cg::this_cluster().sync();
#pragma unroll 1
  for (int r = 0; r < PER; r++) {
    const int tau = chunk * G::TILES16 + r * G::THREADS + threadIdx.x;
    uint64_t x[16];
    if constexpr (G::C > 1) {
      cg::cluster_group cl = cg::this_cluster();
#pragma unroll
      for (int t = 0; t < 16; t++) {
        const int gi = tau + t * ST0;
        const uint64_t *src = cl.map_shared_rank(sm, gi >> G::LOGM);
        x[t] = src[pidx(gi & (G::M - 1))];
      }
    } else {
#pragma unroll
      for (int t = 0; t < 16; t++)
        x[t] = sm[(ST0 % 16 == 0) ? pidx(tau) + t * (ST0 + ST0 / 8) : pidx(tau + t * ST0)];
    }
    tile_inv<4, MODE == INV_RAW>(x, TileTw<4, false>(itw, 1u, 0u), m);
#pragma unroll
    for (int t = 0; t < 16; t++) {
      if (MODE == INV_RAW) {
        store(tau + t * ST0, x[t]);
      } else {
        uint64_t v = shoup_lazy(x[t], pc.ninv, pc.ninv_q, m);
        store(tau + t * ST0, MODE == INV_CANON ? csub(v, m.p) : v);
      }
    }
  }
  // nobody's shared memory is reused or released while peers still read it
  if constexpr (G::C > 1) cg::this_cluster().sync();
  else __syncthreads();
}

// ---------------------------------------------------------------------------------------------------------
// multiword helpers (little-endian, two's complement)
// ---------------------------------------------------------------------------------------------------------
// Residue (lazy, below 5p) of the signed W-word integer words[w*stride] modulo the prime -- the same residue class
// as gcry_mpi_mod in rns_decompose, GPQHE/src/rns.c:37-48.  Horner in 2^64 with Shoup steps; a negative top word
// subtracts 2^(64 W), i.e. adds p - (2^(64W) mod p) = negW.  WT > 0 fixes W at compile time (all loads issue first).
template <int WT>
__device__ __forceinline__ uint64_t decompose_words(const uint64_t *__restrict__ words, size_t stride, int W,
                                                    const PrimeConsts &pc, const Mod &m, uint64_t negW) {
  uint64_t acc, top;
  if constexpr (WT > 0) {
    uint64_t raw[WT];
#pragma unroll
    for (int w = 0; w < WT; w++) raw[w] = HB_LD(&words[(size_t)w * stride]);
    top = raw[WT - 1];
    acc = top;                                               // any 64-bit value is a valid lazy residue
#pragma unroll
    for (int w = WT - 2; w >= 0; w--) acc = shoup_lazy(acc, pc.r64, pc.r64_q, m) + fold(raw[w], m);   // < 2p + 2p
    if (WT == 1) acc = fold(acc, m);
  } else {
    top = HB_LD(&words[(size_t)(W - 1) * stride]);
    acc = top;
    for (int w = W - 2; w >= 0; w--)
      acc = shoup_lazy(acc, pc.r64, pc.r64_q, m) + fold(HB_LD(&words[(size_t)w * stride]), m);
    if (W == 1) acc = fold(acc, m);
  }
  const uint64_t mask = (uint64_t)((int64_t)top >> 63);
  return acc + (negW & mask);                                // < 5p
}

// The same residue times a fixed scalar s: sum_w word_w * (2^(64 w) s mod p), one Shoup product per word.  The word
// weights (pairs) are held in registers for WT > 0 (c[0..WT), negW = p - weight of 2^(64 WT)); WT == 0 reads them from dc.
// Returned below 5p for W <= 2, below 3p otherwise.
template <int WT>
struct DecScale {
  ulonglong2 c[WT > 0 ? WT : 1];
  uint64_t negW;
  const ulonglong2 *dc;
  __device__ __forceinline__ DecScale(const ulonglong2 *__restrict__ dc_, int W, const Mod &m) : dc(dc_) {
    if constexpr (WT > 0) {
#pragma unroll
      for (int w = 0; w < WT; w++) c[w] = __ldg(&dc_[w]);
    }
    negW = m.p - __ldg(&dc_[W]).x;
  }
};
template <int WT>
__device__ __forceinline__ uint64_t decompose_scaled(const uint64_t *__restrict__ words, size_t stride, int W,
                                                     const DecScale<WT> &ds, const Mod &m) {
  uint64_t acc = 0, top;
  if constexpr (WT > 0) {
    uint64_t raw[WT];
#pragma unroll
    for (int w = 0; w < WT; w++) raw[w] = HB_LD(&words[(size_t)w * stride]);
    top = raw[WT - 1];
#pragma unroll
    for (int w = 0; w < WT; w++) acc += shoup_lazy(raw[w], ds.c[w].x, ds.c[w].y, m);     // each below 2p
  } else {
    top = 0;
    for (int w = 0; w < W; w++) {
      const ulonglong2 c = __ldg(&ds.dc[w]);
      top = HB_LD(&words[(size_t)w * stride]);
      acc += shoup_lazy(top, c.x, c.y, m);
      if ((w & 7) == 7) acc = fold(acc, m);
    }
    acc = fold(acc, m);
  }
  return acc + (ds.negW & (uint64_t)((int64_t)top >> 63));
}

// C == 1: residues of sigma_g(poly) mod p into shared memory, natural order.  Coalesced reads of the source
// coefficients i, scattered to k = i g mod 2N (sign flipped when k >= N): poly_rot / poly_conj, poly.c:263-283.
template <int LOGN, int WT, bool SCALED>
__device__ __forceinline__ void stage_poly(uint64_t *sm, const uint64_t *__restrict__ poly, int W, unsigned g,
                                           const PrimeConsts &pc, const Mod &m, uint64_t negW,
                                           const ulonglong2 *__restrict__ dc) {
  typedef Cfg<LOGN> G;
  const uint64_t p5 = 5 * m.p;
  if constexpr (SCALED) {
    const DecScale<WT> ds(dc, W, m);
#pragma unroll 4
    for (int i = threadIdx.x; i < G::N; i += G::THREADS) {
      const uint64_t v = decompose_scaled<WT>(poly + i, G::N, W, ds, m);
      const unsigned k = ((unsigned)i * g) & (2 * G::N - 1);
      sm[pidx(k & (G::N - 1))] = (k >= (unsigned)G::N) ? p5 - v : v;
    }
  } else {
#pragma unroll 4
    for (int i = threadIdx.x; i < G::N; i += G::THREADS) {
      const uint64_t v = decompose_words<WT>(poly + i, G::N, W, pc, m, negW);
      const unsigned k = ((unsigned)i * g) & (2 * G::N - 1);
      sm[pidx(k & (G::N - 1))] = (k >= (unsigned)G::N) ? p5 - v : v;
    }
  }
}

// Forward NTT (lazy) of the residues mod p of sigma_g(poly) into the cluster's shared memory; ginv = g^-1 mod 2N.
// dc != nullptr: the residues are multiplied by the scalar whose word weights dc holds (decompose_scaled).
template <int LOGN, bool SCALED = false>
__device__ __forceinline__ void poly_forward(uint64_t *sm, const uint64_t *__restrict__ poly, int W, unsigned g,
                                             unsigned ginv, const PrimeConsts &pc, const Mod &m, uint64_t negW,
                                             const Tw &tw, const ulonglong2 *__restrict__ dc = nullptr) {
  typedef Cfg<LOGN> G;
  if constexpr (G::C == 1) {
    if (W == 1) stage_poly<LOGN, 1, SCALED>(sm, poly, W, g, pc, m, negW, dc);
    else if (W == 2) stage_poly<LOGN, 2, SCALED>(sm, poly, W, g, pc, m, negW, dc);
    else stage_poly<LOGN, 0, SCALED>(sm, poly, W, g, pc, m, negW, dc);
    ntt_forward_resident<LOGN, false, 5>(sm, tw, m);
  } else {
    const uint64_t p5 = 5 * m.p;
    if constexpr (SCALED) {
      const DecScale<0> ds(dc, W, m);
      ntt_forward<LOGN, false, 5>(
          sm,
          [&](int gi) {
            const unsigned ip = ((unsigned)gi * ginv) & (2 * G::N - 1);
            const uint64_t v = decompose_scaled<0>(poly + (ip & (G::N - 1)), G::N, W, ds, m);
            return (ip >= (unsigned)G::N) ? p5 - v : v;
          },
          tw, m);
    } else {
      ntt_forward<LOGN, false, 5>(
          sm,
          [&](int gi) {
            const unsigned ip = ((unsigned)gi * ginv) & (2 * G::N - 1);
            const uint64_t v = decompose_words<0>(poly + (ip & (G::N - 1)), G::N, W, pc, m, negW);
            return (ip >= (unsigned)G::N) ? p5 - v : v;
          },
          tw, m);
    }
  }
}

// sign-extend x (WK words) from bit k-1 in place; afterwards words >= ceil(k/64) hold the sign fill.
template <int WK>
__device__ __forceinline__ void sext_bits(uint64_t (&x)[WK], int k) {
  uint64_t fill = 0;
  bool done = false;
#pragma unroll
  for (int w = 0; w < WK; w++) {
    if (done) { x[w] = fill; continue; }
    int kk = k - 64 * w;       // bits of the value that live in word w and above
    if (kk <= 64) {
      if (kk < 64) x[w] = (uint64_t)(((int64_t)(x[w] << (64 - kk))) >> (64 - kk));
      fill = (uint64_t)(((int64_t)x[w]) >> 63);
      done = true;
    }
  }
}

}  // namespace hbdev
