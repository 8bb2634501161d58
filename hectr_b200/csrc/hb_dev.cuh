// Device-side arithmetic and the in-shared-memory negacyclic NTT core (sm_100a).
//
// Reference semantics being reproduced (results, not code):
//   forward  : GPQHE/src/ntt.c:37-52   Cooley-Tukey, natural in -> bit-reversed out, len = N/2..1
//   inverse  : GPQHE/src/ntt.c:54-73   Gentleman-Sande, len = 1..N/2, then * N^-1
//   modmul   : GPQHE/src/reduce.c:59-66 (Montgomery) / :88-106 (Barrett) -- here Shoup (fixed operand)
//              and a 2^120 Barrett (variable operands); a*b mod p is the same integer in [0,p).
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

namespace hbdev {

struct PrimeConsts {  // must match hb::PrimeConsts
  uint64_t p, mu, one_q, r64, r64_q, ninv, ninv_q, root;
};

struct Tables {  // device pointers, passed by value
  const PrimeConsts *primes;   // [dimub]
  const ulonglong2 *tw, *itw;  // [dimub][N] (w, floor(w 2^64 / p))
  const uint64_t *pow64;       // [dimub][WC+1]
  const uint64_t *crt_inv, *crt_inv_q;  // [dimub+1][dimub]
  const uint64_t *crt_phat;    // [dimub+1][dimub][WC]
  const uint64_t *crt_P;       // [dimub+1][WC]
  const uint64_t *ks_inv, *ks_inv_q;    // [dim]
  const uint64_t *ks_phat;     // [dim][WP]
  const uint64_t *ks_P, *ks_Ph;         // [WP]
  const uint64_t *ks_Pinv;     // [WC]
  unsigned N, logn, dimub, WC, WP, dim;
};

// ---------------------------------------------------------------------------------------------------------
// scalar modular arithmetic
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t csub(uint64_t x, uint64_t m) { return x >= m ? x - m : x; }

// a*w mod p in [0, 2p) for ANY 64-bit a, given wq = floor(w 2^64 / p), w < p.
__device__ __forceinline__ uint64_t shoup_lazy(uint64_t a, uint64_t w, uint64_t wq, uint64_t p) {
  uint64_t q = __umul64hi(a, wq);
  return a * w - q * p;
}
__device__ __forceinline__ uint64_t shoup(uint64_t a, uint64_t w, uint64_t wq, uint64_t p) {
  return csub(shoup_lazy(a, w, wq, p), p);
}

// any 64-bit word -> [0,p)
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

// Forward butterflies.  LAZY (default): Harvey, values in [0,4p).  STRICT: the reference's exact
// data flow, values in [0,p] including its "sum == p stays p" quirk (ntt.c:47-48).
template <bool STRICT>
__device__ __forceinline__ void ct_bfly(uint64_t &X, uint64_t &Y, ulonglong2 w, uint64_t p) {
  if (STRICT) {
    uint64_t t = shoup(Y, w.x, w.y, p);
    uint64_t x = X;
    Y = (x >= t) ? x - t : x - t + p;
    X = (x <= p - t) ? x + t : x + t - p;
  } else {
    uint64_t x = csub(X, 2 * p);
    uint64_t t = shoup_lazy(Y, w.x, w.y, p);
    X = x + t;
    Y = x + 2 * p - t;
  }
}

// Inverse butterfly (ntt.c:62-67): inputs in [0,2p), outputs in [0,2p).
__device__ __forceinline__ void gs_bfly(uint64_t &X, uint64_t &Y, ulonglong2 w, uint64_t p) {
  uint64_t s = csub(X + Y, 2 * p);
  uint64_t d = X + 2 * p - Y;
  X = s;
  Y = shoup_lazy(d, w.x, w.y, p);
}

__device__ __forceinline__ uint64_t canon4(uint64_t x, uint64_t p) { return csub(csub(x, 2 * p), p); }

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
  static constexpr int THREADS = M / 16;       // one radix-16 tile per thread in the global pass
  static constexpr int SMEM_WORDS = M + (M >> 4) * 2;   // 2 pad words per 16 (keeps 16-byte alignment)
  static constexpr int SMEM_BYTES = SMEM_WORDS * 8;
};

__device__ __forceinline__ int pidx(int i) { return i + ((i >> 4) << 1); }

// Twiddle index of the butterfly at tile-bit `beta` for tile element t, tile of 2^R elements whose coarsest
// block index is K0 (= 2^S + block):  k = (K0 << (R-1-beta)) + (t >> (beta+1)).
template <int R, bool INV, bool STRICT>
__device__ __forceinline__ void tile_compute(uint64_t (&x)[1 << R], const ulonglong2 *__restrict__ tw, unsigned K0,
                                             uint64_t p) {
#pragma unroll
  for (int a = 0; a < R; a++) {
    const int beta = INV ? a : (R - 1 - a);
    const int half = 1 << beta;
#pragma unroll
    for (int t = 0; t < (1 << R); t++) {
      if ((t & half) == 0) {
        ulonglong2 w = __ldg(&tw[(K0 << (R - 1 - beta)) + (t >> (beta + 1))]);
        if (INV) gs_bfly(x[t], x[t + half], w, p);
        else ct_bfly<STRICT>(x[t], x[t + half], w, p);
      }
    }
  }
}

// One chunk-local pass: stages [S, S+R) in forward numbering (len = N>>(S+1) ...), on this CTA's chunk.
template <int LOGN, int S, int R, bool INV, bool STRICT>
__device__ __forceinline__ void local_pass(uint64_t *sm, const ulonglong2 *__restrict__ tw, uint64_t p,
                                           unsigned chunk) {
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
    if (ST == 1) {
#pragma unroll
      for (int t = 0; t < (1 << R); t += 2) {
        ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(&sm[pidx(g + t)]);
        x[t] = v.x; x[t + 1] = v.y;
      }
    } else {
#pragma unroll
      for (int t = 0; t < (1 << R); t++) x[t] = sm[pidx(g + t * ST)];
    }
    tile_compute<R, INV, STRICT>(x, tw, blk0 + b, p);
    if (ST == 1) {
#pragma unroll
      for (int t = 0; t < (1 << R); t += 2)
        *reinterpret_cast<ulonglong2 *>(&sm[pidx(g + t)]) = make_ulonglong2(x[t], x[t + 1]);
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

template <int LOGN, int S, bool STRICT>
__device__ __forceinline__ void fwd_local(uint64_t *sm, const ulonglong2 *tw, uint64_t p, unsigned chunk) {
  if constexpr (S < LOGN) {
    constexpr int R = PassSize<LOGN - S>::R;
    local_pass<LOGN, S, R, false, STRICT>(sm, tw, p, chunk);
    __syncthreads();
    fwd_local<LOGN, S + R, STRICT>(sm, tw, p, chunk);
  }
}

template <int LOGN, int S>
__device__ __forceinline__ void inv_local(uint64_t *sm, const ulonglong2 *itw, uint64_t p, unsigned chunk) {
  if constexpr (S < LOGN) {
    constexpr int R = PassSize<LOGN - S>::R;
    inv_local<LOGN, S + R>(sm, itw, p, chunk);
    local_pass<LOGN, S, R, true, false>(sm, itw, p, chunk);
    __syncthreads();
  }
}

template <int LOGN>
__device__ __forceinline__ unsigned chunk_rank() {
  if constexpr (Cfg<LOGN>::C > 1) return cg::this_cluster().block_rank();
  else return 0;
}
template <int LOGN>
__device__ __forceinline__ void cluster_barrier() {
  if constexpr (Cfg<LOGN>::C > 1) cg::this_cluster().sync();
  else __syncthreads();
}

// Forward transform of one limb into the cluster's shared memory.  load(gi) returns coefficient gi of the
// natural-order input, < 4p (STRICT: <= p).  On return (all CTAs synchronised) this CTA's chunk holds outputs
// [chunk*M, (chunk+1)*M) of the bit-reversed-order result, lazily reduced (< 4p; STRICT: exact reference values).
template <int LOGN, bool STRICT, class Load>
__device__ __forceinline__ void ntt_forward(uint64_t *sm, Load load, const ulonglong2 *__restrict__ tw, uint64_t p) {
  typedef Cfg<LOGN> G;
  const unsigned chunk = chunk_rank<LOGN>();
  if constexpr (LOGN >= 4) {
    constexpr int ST0 = G::N / 16;
    const int tau = chunk * G::THREADS + threadIdx.x;
    uint64_t x[16];
#pragma unroll
    for (int t = 0; t < 16; t++) x[t] = load(tau + t * ST0);
    tile_compute<4, false, STRICT>(x, tw, 1u, p);
    if constexpr (G::C > 1) {
      cg::cluster_group cl = cg::this_cluster();
      cl.sync();   // every CTA of the cluster is resident before remote shared memory is written
#pragma unroll
      for (int t = 0; t < 16; t++) {
        const int gi = tau + t * ST0;
        uint64_t *dst = cl.map_shared_rank(sm, gi >> G::LOGM);
        dst[pidx(gi & (G::M - 1))] = x[t];
      }
      cl.sync();
    } else {
#pragma unroll
      for (int t = 0; t < 16; t++) sm[pidx(tau + t * ST0)] = x[t];
      __syncthreads();
    }
    fwd_local<LOGN, 4, STRICT>(sm, tw, p, chunk);
  }
}

// Inverse transform of one limb held in the cluster's shared memory (values < 2p, bit-reversed order);
// store(gi, v) receives natural-order coefficient gi, already scaled by N^-1 and canonical in [0,p).
template <int LOGN, class Store>
__device__ __forceinline__ void ntt_inverse(uint64_t *sm, Store store, const ulonglong2 *__restrict__ itw,
                                            const PrimeConsts &pc) {
  typedef Cfg<LOGN> G;
  const unsigned chunk = chunk_rank<LOGN>();
  const uint64_t p = pc.p;
  inv_local<LOGN, 4>(sm, itw, p, chunk);
  constexpr int ST0 = G::N / 16;
  const int tau = chunk * G::THREADS + threadIdx.x;
  uint64_t x[16];
  if constexpr (G::C > 1) {
    cg::cluster_group cl = cg::this_cluster();
    cl.sync();
#pragma unroll
    for (int t = 0; t < 16; t++) {
      const int gi = tau + t * ST0;
      const uint64_t *src = cl.map_shared_rank(sm, gi >> G::LOGM);
      x[t] = src[pidx(gi & (G::M - 1))];
    }
    cl.sync();   // nobody's shared memory is reused or released while peers still read it
  } else {
#pragma unroll
    for (int t = 0; t < 16; t++) x[t] = sm[pidx(tau + t * ST0)];
    __syncthreads();
  }
  tile_compute<4, true, false>(x, itw, 1u, p);
#pragma unroll
  for (int t = 0; t < 16; t++) store(tau + t * ST0, shoup(x[t], pc.ninv, pc.ninv_q, p));
}

// ---------------------------------------------------------------------------------------------------------
// multiword helpers (little-endian, two's complement)
// ---------------------------------------------------------------------------------------------------------
// residue of the signed W-word integer words[w*stride] modulo prime `pc` (non-negative, like gcry_mpi_mod in
// rns_decompose, GPQHE/src/rns.c:37-48).  pow64W = 2^(64 W) mod p.
__device__ __forceinline__ uint64_t decompose_words(const uint64_t *__restrict__ words, size_t stride, int W,
                                                    const PrimeConsts &pc, uint64_t pow64W) {
  const uint64_t p = pc.p;
  uint64_t top = words[(size_t)(W - 1) * stride];
  uint64_t acc = reduce_word(top, pc);
  for (int w = W - 2; w >= 0; w--) {
    uint64_t lo = reduce_word(words[(size_t)w * stride], pc);
    acc = csub(shoup(acc, pc.r64, pc.r64_q, p) + lo, p);
  }
  if ((int64_t)top < 0) acc = (acc >= pow64W) ? acc - pow64W : acc + p - pow64W;
  return acc;
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
