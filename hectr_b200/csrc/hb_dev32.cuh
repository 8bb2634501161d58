// 30-bit negacyclic NTT core and the plaintext-product kernels of he_gemv on the engine's internal 30-bit RNS basis
// (hb_tables32.hpp explains why the integers are the same as the reference's, GPQHE/src/he-algo.c:64-79 + he-mult.c:159-196).
//
// Harvey lazy butterflies on 32-bit words: every prime is below 2^30, values live in [0, 4p) (forward) / [0, 2p) (inverse),
// a Shoup product  a*w - umulhi(a, w')*p  is one IMAD.HI + two IMAD and accepts ANY 32-bit a.
// One limb (N <= 2^13 coefficients, 4 bytes each) per CTA in padded shared memory (one pad word per 32), radix-2^R
// register tiles with R <= 4 like the 59-bit core, same twiddle indexing (entry k of stage s at 2^s + k).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "hb_dev.cuh"
#include "hb_kernels.cuh"

namespace hb32 {

struct Prime32 { uint32_t p, ninv, ninv_q, mu; };   // must match hb::Prime32

struct Tables32 {                 // device pointers, by value
  const Prime32 *primes;          // [nprimes]
  const uint2 *tw, *itw;          // [nprimes][N] (w, floor(w 2^32 / p))
  const uint2 *tw_tr, *itw_tr;    // [nprimes][tr_stride] lane-transposed copy for the stride-1 pass (see Tw32)
  unsigned tr_stride;
  const uint2 *pow32;             // [nprimes][ndig + 1]
  const uint2 *crt_inv;           // [nprimes + 1][nprimes]
  const uint64_t *crt_qhat;       // [nprimes + 1][nprimes][WC]
  const uint64_t *crt_Q;          // [nprimes + 1][WC]
  unsigned N, logn, nprimes, WC, ndig;
};

__device__ __forceinline__ uint32_t shoup32(uint32_t a, uint32_t w, uint32_t wq, uint32_t p) {   // [0, 2p) for any a
  return a * w - __umulhi(a, wq) * p;
}
__device__ __forceinline__ uint32_t csub32(uint32_t x, uint32_t m) { return min(x, x - m); }    // x < 2m -> x mod-ish m
// a w mod p in (0, 2p) for any 32-bit a, from wR = w 2^32 mod p and pinv = p^-1 mod 2^32:  t = a wR, m = lo(t) pinv, so that
// m p = t (mod 2^32) and (t - m p) / 2^32 = hi(t) - hi(m p) exactly, in (-p, p)
__device__ __forceinline__ uint32_t mont32(uint32_t a, uint32_t wR, uint32_t p, uint32_t pinv) {
  const uint64_t t = (uint64_t)a * wR;
  const uint32_t m = (uint32_t)t * pinv;
  return (uint32_t)(t >> 32) - __umulhi(m, p) + p;
}
__device__ __forceinline__ uint32_t inv32_mod_2_32(uint32_t p) {      // p odd: Newton, 5 -> 10 -> 20 -> 40 correct bits
  uint32_t x = p * 3u ^ 2u;                                           // 5 correct bits
  x *= 2u - p * x; x *= 2u - p * x; x *= 2u - p * x;
  return x;
}

template <int LOGN>
struct Cfg32 {
  static constexpr int N = 1 << LOGN;
  static constexpr int TILES16 = N / 16;
  static constexpr int THREADS = TILES16 > 256 ? 256 : TILES16;
  static constexpr int SMEM_WORDS = N + (N >> 5);
  static constexpr int SMEM_BYTES = SMEM_WORDS * 4;
};
// one pad word per 32: a warp's 32 consecutive words, its 16-strided tiles and (with the lane swap in pass32) its two
// 256-blocks of the middle pass all fall on 32 different banks
__device__ __forceinline__ int pidx32(int i) { return i + (i >> 5); }
// Thread tid's coefficients c = tid + k THREADS (THREADS a multiple of 32) sit at padded offsets pidx32(tid) + k PSTEP32: the loops
// below index shared memory as base + k * constant instead of recomputing the padding per access.
#define HB32_PSTEP(THREADS) ((THREADS) + ((THREADS) >> 5))

// the 2^a consecutive twiddles of one stage of a tile, fetched 16 bytes at a time (the run starts at an even index)
template <int CNT>
__device__ __forceinline__ void load_tw32(uint2 (&w)[CNT], const uint2 *__restrict__ src) {
  if constexpr (CNT == 1) {
    w[0] = __ldg(src);
  } else {
    const uint4 *v = reinterpret_cast<const uint4 *>(src);
#pragma unroll
    for (int m = 0; m < CNT / 2; m++) {
      const uint4 x = __ldg(&v[m]);
      w[2 * m] = make_uint2(x.x, x.y);
      w[2 * m + 1] = make_uint2(x.z, x.w);
    }
  }
}

// Twiddles of one prime and one direction.  nat: index = bit-reversed exponent (entry k of stage s at 2^s + k).  In the
// stride-1 pass every thread has its own 2^R - 1 twiddles; read from `nat` a warp's 16-byte loads of stage A >= 2 are 32 / 64
// bytes apart (8 / 16 cache lines per instruction, 86 L1 wavefronts per tile), so a lane-transposed copy holds them as
//   slot (2^A - 1) + u of tile K  at  ((K >> 5) (2^R - 1) + slot) 32 + (K & 31)  =  nat[((2^S + K) << A) + u],  S = log2 N - R
// and every load is 256 contiguous bytes per warp (30 wavefronts per tile).  Built by hb::build_tables32.
struct Tw32 {
  const uint2 *nat, *tr;
};
__device__ __forceinline__ Tw32 tw32_of(const Tables32 &T, unsigned i, bool inverse) {
  Tw32 t;
  t.nat = (inverse ? T.itw : T.tw) + (size_t)i * T.N;
  t.tr = (inverse ? T.itw_tr : T.tw_tr) + (size_t)i * T.tr_stride;
  return t;
}
template <int R, int A, bool TR>
__device__ __forceinline__ void load_stage_tw32(uint2 (&w)[1 << A], const uint2 *__restrict__ tw, unsigned K0) {
  if constexpr (TR) {                         // tw: the tile's lane of the transposed copy
#pragma unroll
    for (int u = 0; u < (1 << A); u++) w[u] = __ldg(&tw[(((1 << A) - 1) + u) * 32]);
  } else {
    load_tw32<(1 << A)>(w, tw + ((size_t)K0 << A));
  }
}

template <int R, int A, bool TR = false>
__device__ __forceinline__ void stage_fwd32(uint32_t (&x)[1 << R], const uint2 *__restrict__ tw, unsigned K0, uint32_t p) {
  constexpr int beta = R - 1 - A, half = 1 << beta;
  const uint32_t p2 = 2 * p;
  uint2 w[1 << A];
  load_stage_tw32<R, A, TR>(w, tw, K0);
#pragma unroll
  for (int t = 0; t < (1 << R); t++) {
    if ((t & half) == 0) {
      const uint2 ww = w[t >> (beta + 1)];
      const uint32_t u = csub32(x[t], p2);
      const uint32_t v = shoup32(x[t + half], ww.x, ww.y, p);
      x[t] = u + v;
      x[t + half] = u - v + p2;
    }
  }
}
// forward tile: X' = X + wY, Y' = X - wY + 2p, X first brought below 2p; all values stay in [0, 4p)
template <int R, bool TR = false>
__device__ __forceinline__ void tile_fwd32(uint32_t (&x)[1 << R], const uint2 *__restrict__ tw, unsigned K0, uint32_t p) {
  stage_fwd32<R, 0, TR>(x, tw, K0, p);
  if constexpr (R > 1) stage_fwd32<R, 1, TR>(x, tw, K0, p);
  if constexpr (R > 2) stage_fwd32<R, 2, TR>(x, tw, K0, p);
  if constexpr (R > 3) stage_fwd32<R, 3, TR>(x, tw, K0, p);
}

template <int R, int BETA, bool TR = false>
__device__ __forceinline__ void stage_inv32(uint32_t (&x)[1 << R], const uint2 *__restrict__ tw, unsigned K0, uint32_t p) {
  constexpr int A = R - 1 - BETA, half = 1 << BETA;
  const uint32_t p2 = 2 * p;
  uint2 w[1 << A];
  load_stage_tw32<R, A, TR>(w, tw, K0);
#pragma unroll
  for (int t = 0; t < (1 << R); t++) {
    if ((t & half) == 0) {
      const uint2 ww = w[t >> (BETA + 1)];
      const uint32_t s = csub32(x[t] + x[t + half], p2);
      const uint32_t d = x[t] - x[t + half] + p2;
      x[t] = s;
      x[t + half] = shoup32(d, ww.x, ww.y, p);
    }
  }
}
// inverse tile (Gentleman-Sande): inputs and outputs in [0, 2p)
template <int R, bool TR = false>
__device__ __forceinline__ void tile_inv32(uint32_t (&x)[1 << R], const uint2 *__restrict__ tw, unsigned K0, uint32_t p) {
  stage_inv32<R, 0, TR>(x, tw, K0, p);
  if constexpr (R > 1) stage_inv32<R, 1, TR>(x, tw, K0, p);
  if constexpr (R > 2) stage_inv32<R, 2, TR>(x, tw, K0, p);
  if constexpr (R > 3) stage_inv32<R, 3, TR>(x, tw, K0, p);
}

// one pass over the limb in shared memory: stages [S, S+R) in forward numbering
#ifndef HB32_TW_TR
#define HB32_TW_TR 1
#endif
template <int LOGN, int S, int R, bool INV>
__device__ __forceinline__ void pass32(uint32_t *sm, const Tw32 &tw, uint32_t p) {
  typedef Cfg32<LOGN> G;
  constexpr int LOGST = LOGN - S - R, ST = 1 << LOGST, TILES = G::N >> R, LOGBLK = LOGN - S;
  for (int t0 = threadIdx.x; t0 < TILES; t0 += G::THREADS) {
    // 16-strided pass: the two half-warps take blocks two apart (bits 4 and 5 of the tile index swapped), whose padded
    // offsets differ by 16 banks
    const int tile = (LOGST == 4 && TILES >= 64) ? ((t0 & ~0x30) | ((t0 & 0x10) << 1) | ((t0 & 0x20) >> 1)) : t0;
    const int b = tile >> LOGST, o = tile & (ST - 1), g = (b << LOGBLK) + o;
    uint32_t x[1 << R];
    // Padded offsets of the tile's elements as base + compile-time constant, so that every access is one LDS / STS with an
    // immediate: pidx32(g + t ST) = pidx32(g) + t (ST + ST/32) when 32 | ST; two interleaved bases when ST = 16; pidx32(g) + t
    // for the 2^R <= 16 consecutive elements of the stride-1 pass (g is a multiple of 2^R: no carry into bit 5).
    int off[1 << R];
    if constexpr (LOGST >= 5) {
      const int base = pidx32(g);
#pragma unroll
      for (int t = 0; t < (1 << R); t++) off[t] = base + t * (ST + (ST >> 5));
    } else if constexpr (LOGST == 4) {
      const int b0 = pidx32(g), b1 = pidx32(g + 16);
#pragma unroll
      for (int t = 0; t < (1 << R); t++) off[t] = ((t & 1) ? b1 : b0) + (t >> 1) * 33;
    } else if constexpr (LOGST == 0 && R <= 4) {
      const int base = pidx32(g);
#pragma unroll
      for (int t = 0; t < (1 << R); t++) off[t] = base + t;
    } else {
#pragma unroll
      for (int t = 0; t < (1 << R); t++) off[t] = pidx32(g + t * ST);
    }
#pragma unroll
    for (int t = 0; t < (1 << R); t++) x[t] = sm[off[t]];
    if constexpr (LOGST == 0 && HB32_TW_TR && (TILES % 32 == 0)) {      // stride-1 pass: b == tile, one set of twiddles per thread
      const uint2 *lane_tw = tw.tr + (size_t)(tile >> 5) * (((1 << R) - 1) * 32) + (tile & 31);
      if (INV) tile_inv32<R, true>(x, lane_tw, 0u, p);
      else tile_fwd32<R, true>(x, lane_tw, 0u, p);
    } else {
      if (INV) tile_inv32<R>(x, tw.nat, (1u << S) + b, p);
      else tile_fwd32<R>(x, tw.nat, (1u << S) + b, p);
    }
#pragma unroll
    for (int t = 0; t < (1 << R); t++) sm[off[t]] = x[t];
  }
}

template <int LOGN, int S>
__device__ __forceinline__ void fwd32(uint32_t *sm, const Tw32 &tw, uint32_t p) {
  if constexpr (S < LOGN) {
    constexpr int R = hbdev::PassSize<LOGN - S>::R;
    pass32<LOGN, S, R, false>(sm, tw, p);
    __syncthreads();
    fwd32<LOGN, S + R>(sm, tw, p);
  }
}
template <int LOGN, int S>
__device__ __forceinline__ void inv32(uint32_t *sm, const Tw32 &itw, uint32_t p) {
  if constexpr (S < LOGN) {
    constexpr int R = hbdev::PassSize<LOGN - S>::R;
    inv32<LOGN, S + R>(sm, itw, p);
    pass32<LOGN, S, R, true>(sm, itw, p);
    __syncthreads();
  }
}
// natural order, values below 4p, visible after the barrier this starts with -> bit-reversed order, values below 4p
template <int LOGN>
__device__ __forceinline__ void ntt32_forward(uint32_t *sm, const Tw32 &tw, uint32_t p) {
  __syncthreads();
  fwd32<LOGN, 0>(sm, tw, p);
}
// bit-reversed order, values below 2p -> natural order, NOT scaled by N^-1 (folded into the fixed multiplicand), below 2p
template <int LOGN>
__device__ __forceinline__ void ntt32_inverse(uint32_t *sm, const Tw32 &itw, uint32_t p) {
  __syncthreads();
  inv32<LOGN, 0>(sm, itw, p);
}

}  // namespace hb32

namespace hb32 {

// ---------------------------------------------------------------------------------------------------------------
// rns_decompose (GPQHE/src/rns.c:37-48) onto a 30-bit prime + forward transform.
// src: npoly polynomials of W 64-bit words (planar, two's complement), polynomial q at src + q*poly_stride.
// out[q][D][N] u32, bit-reversed order, lazily reduced (below 4p) -- or canonical [0,p) when CANON (fixed multiplicands).
// ---------------------------------------------------------------------------------------------------------------
template <int LOGN, bool CANON>
__global__ void __launch_bounds__(Cfg32<LOGN>::THREADS)
k32_poly_ntt(const uint64_t *__restrict__ src, size_t poly_stride, int W, uint32_t *__restrict__ out, Tables32 T, unsigned D) {
  typedef Cfg32<LOGN> G;
  extern __shared__ __align__(16) uint32_t sm32[];
  const unsigned q = blockIdx.x / D, d = blockIdx.x % D;
  const uint32_t p = T.primes[d].p, p2 = 2 * p;
  const uint2 *pw = T.pow32 + (size_t)d * (T.ndig + 1);
  const uint32_t negW = p - __ldg(&pw[2 * W]).x;             // a negative coefficient subtracts 2^(64 W)
  const uint64_t *poly = src + (size_t)q * poly_stride;
  for (int i = threadIdx.x; i < G::N; i += G::THREADS) {
    uint32_t acc = 0;
    uint64_t top = 0;
    for (int w = 0; w < W; w++) {
      top = HB_LD(&poly[(size_t)w * G::N + i]);
      const uint2 c0 = __ldg(&pw[2 * w]), c1 = __ldg(&pw[2 * w + 1]);
      acc = csub32(acc + shoup32((uint32_t)top, c0.x, c0.y, p), p2);            // stays below 2p
      acc = csub32(acc + shoup32((uint32_t)(top >> 32), c1.x, c1.y, p), p2);
    }
    acc += (negW & (uint32_t)((int64_t)top >> 63));                            // below 3p
    sm32[pidx32(i)] = acc;
  }
  ntt32_forward<LOGN>(sm32, tw32_of(T, d, false), p);
  uint32_t *dst = out + ((size_t)q * D + d) * G::N;
  for (int i = threadIdx.x; i < G::N; i += G::THREADS) {
    uint32_t v = sm32[pidx32(i)];
    if (CANON) { v = csub32(v, p2); v = csub32(v, p); }
    dst[i] = v;
  }
}

// canonical residues -> Shoup pairs of  w * N^-1 * (Q_Dw/q_d)^-1  (the inverse transform then skips its scaling and the
// reconstruction its per-limb multiply).  limbs[j][N] uses prime j % D.
// Output is PLANAR per limb: out[limb][0][N] the multiplicands, out[limb][1][N] their quotients, so that a warp's 16-byte
// loads of either plane are contiguous.
// planes = 3 adds out[limb][2][N] = w 2^32 mod p: a fixed multiplicand in Montgomery form needs ONE word per product where a
// Shoup pair needs two (mont32 below); the diagonals of a gemv plan carry all three (HB32_PT_PLANES).
#define HB32_PT_PLANES 3
__global__ void k32_shoup_prep(const uint32_t *__restrict__ in, uint32_t *__restrict__ out, Tables32 T, unsigned D, size_t nlimbs, unsigned Dw,
                               unsigned planes) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= nlimbs * T.N) return;
  const unsigned d = (unsigned)((gid / T.N) % D);
  const Prime32 pc = T.primes[d];
  uint32_t w = in[gid];
  w = csub32(shoup32(w, pc.ninv, pc.ninv_q, pc.p), pc.p);
  if (Dw) {                                                 // Dw == 0: times N^-1 only (evaluation keys on the hybrid basis)
    const uint2 inv = T.crt_inv[(size_t)Dw * T.nprimes + d];
    w = csub32(shoup32(w, inv.x, inv.y, pc.p), pc.p);
  }
  const size_t limb = gid / T.N, i = gid % T.N;
  const uint64_t w32 = ((uint64_t)w) << 32;
  const uint32_t wq = (uint32_t)(w32 / pc.p);
  const uint32_t wR = (uint32_t)(w32 - (uint64_t)wq * pc.p);         // w 2^32 mod p (Montgomery form)
  if (planes == 1) { out[limb * T.N + i] = wR; return; }             // evaluation keys on the hybrid basis: mont32 only
  out[(planes * limb) * T.N + i] = w;
  out[(planes * limb + 1) * T.N + i] = wq;
  if (planes > 2) out[(planes * limb + 2) * T.N + i] = wR;
}

// ---------------------------------------------------------------------------------------------------------------
// all giant steps of he_gemv in one launch (GPQHE/src/he-algo.c:64-79):
//   chat[g][q][d] = y-weighted invntt( sum_j hat[j][q][d] o pthat[g*n1 + j][d] ),   g < ngroups, q < nq, j < n1
// ---------------------------------------------------------------------------------------------------------------
template <int LOGN>
__global__ void __launch_bounds__(Cfg32<LOGN>::THREADS)
k32_mulacc_intt(const uint32_t *__restrict__ hat, size_t hat_jstride, const uint32_t *__restrict__ pthat_a, const uint32_t *__restrict__ pthat_b,
                unsigned q_split, size_t pthat_gstride, const __grid_constant__ hbdev::GroupMap GM, unsigned n1, unsigned nq,
                uint32_t *__restrict__ chat, Tables32 T, unsigned D) {
  typedef Cfg32<LOGN> G;
  extern __shared__ __align__(16) uint32_t sm32[];
  const unsigned qq = blockIdx.x / D, d = blockIdx.x % D, q = qq % nq;
  const uint32_t *pthat_all = q < q_split ? pthat_a : pthat_b;
  const uint32_t p = T.primes[d].p, p2 = 2 * p;
  const uint32_t *a = hat + ((size_t)q * D + d) * G::N;
  const uint32_t *w = pthat_all + ((size_t)GM.idx[qq / nq] * pthat_gstride + d) * HB32_PT_PLANES * G::N;       // strides in limbs
#pragma unroll 2
  for (int i = threadIdx.x; i < G::N; i += G::THREADS) {
    uint32_t acc = 0;
    for (unsigned j = 0; j < n1; j++) {
      const uint32_t *wj = w + (size_t)j * D * HB32_PT_PLANES * G::N;
      acc = csub32(acc + shoup32(HB_LD(&a[(size_t)j * hat_jstride + i]), __ldg(&wj[i]), __ldg(&wj[G::N + i]), p), p2);
    }
    sm32[pidx32(i)] = acc;                                                       // below 2p
  }
  ntt32_inverse<LOGN>(sm32, tw32_of(T, d, true), p);
  uint32_t *dst = chat + ((size_t)qq * D + d) * G::N;
  for (int i = threadIdx.x; i < G::N; i += G::THREADS) dst[i] = sm32[pidx32(i)];
}

// ---------------------------------------------------------------------------------------------------------------
// exact CRT over D 30-bit limbs, centred mod Q_D, then centred mod 2^k (poly_rns2mpi, poly.c:109-120, on this basis):
//   x = sum_d y_d (Q_D/q_d) - v Q_D,  v = round(sum_d y_d / q_d);  |x| < Q_D/4 for every valid product, so a 56-bit
//   fixed-point fraction (error below 2 D units) decides v.  Limbs arrive pre-weighted (y_d, below 2p).
// ---------------------------------------------------------------------------------------------------------------
template <int WK>
__global__ void k32_recon_smod(const uint32_t *__restrict__ chat, unsigned D, uint64_t *__restrict__ out, int Wout, int k, Tables32 T,
                               unsigned npoly) {
  const unsigned N = T.N;
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (size_t)npoly * N) return;
  const unsigned q = gid / N, i = gid % N;
  const uint32_t *limbs = chat + (size_t)q * D * N + i;
  const uint64_t *qhat = T.crt_qhat + (size_t)D * T.nprimes * T.WC;
  uint64_t acc[WK];
#pragma unroll
  for (int w = 0; w < WK; w++) acc[w] = 0;
  uint64_t F = 0;
  for (unsigned d = 0; d < D; d++) {
    const Prime32 pc = T.primes[d];
    const uint32_t y = csub32(limbs[(size_t)d * N], pc.p);
    F += ((uint64_t)y * pc.mu) >> 5;                       // y / q_d in units of 2^-56
    uint64_t carry = 0;
#pragma unroll
    for (int w = 0; w < WK; w++) {
      const uint64_t pw = (w < (int)T.WC) ? qhat[(size_t)d * T.WC + w] : 0;
      const unsigned __int128 t = (unsigned __int128)y * pw + acc[w] + carry;
      acc[w] = (uint64_t)t;
      carry = (uint64_t)(t >> 64);
    }
  }
  const uint64_t v = (F + (1ULL << 55)) >> 56;
  const uint64_t *Q = T.crt_Q + (size_t)D * T.WC;
  uint64_t borrow = 0;
#pragma unroll
  for (int w = 0; w < WK; w++) {
    const uint64_t pw = (w < (int)T.WC) ? Q[w] : 0;
    const unsigned __int128 t = (unsigned __int128)v * pw + borrow;
    const uint64_t sub = (uint64_t)t;
    borrow = (uint64_t)(t >> 64) + (acc[w] < sub);
    acc[w] -= sub;
  }
  hbdev::sext_bits<WK>(acc, k);
  uint64_t *o = out + (size_t)q * Wout * N + i;
#pragma unroll
  for (int w = 0; w < WK; w++)
    if (w < Wout) o[(size_t)w * N] = acc[w];
}

}  // namespace hb32

namespace hb32 {

// ===============================================================================================================
// Memory-lean variants.  On the 30-bit basis the butterflies are cheap enough that global-load latency and DRAM
// traffic bind instead of the integer multiplier, so each CTA loads its inputs ONCE, 16 bytes per thread per load,
// and loops over everything that reuses them.
// ===============================================================================================================

// One CTA per polynomial: the W words of 16 coefficients per thread are read once and decomposed onto every limb.
// k32_poly_ntt_all keeps no digits in registers: the words are read again for every limb (from L2 after the first), four
// coefficients at a time, and five CTAs of 48 registers fit an SM where two of 128 did: 6.13 -> 5.45 ms per step
// [B200, 4096 instances; variants_misc.log, variants_reread.log].  HB32_PN_REREAD=0 restores the digits in registers.
#ifndef HB32_PN_MINB
#define HB32_PN_MINB 5
#endif
#ifndef HB32_PN_REREAD
#define HB32_PN_REREAD 4
#endif
// k32_mulacc_intt_all reads the n1 rotated transforms again for every giant step (from L2 after the first) instead of holding
// them in 64 registers: six CTAs of 40 registers per SM, 5.13 -> 4.65 ms per step [variants_reread_ma.log, variants_reread2.log].
// HB32_MA_REREAD=0 with HB32_MA_MINB=3 restores the register-resident version.
#ifndef HB32_MA_REREAD
#define HB32_MA_REREAD 1
#endif
#ifndef HB32_MA_MONT
#define HB32_MA_MONT 1
#endif
#ifndef HB32_MA_MINB
#define HB32_MA_MINB 6
#endif
template <int LOGN, int WT>
__global__ void __launch_bounds__(Cfg32<LOGN>::THREADS, HB32_PN_MINB)
k32_poly_ntt_all(const uint64_t *__restrict__ src, size_t poly_stride, uint32_t *__restrict__ out, Tables32 T, unsigned D) {
  typedef Cfg32<LOGN> G;
  constexpr int PER = G::N / G::THREADS;                      // coefficients per thread
  extern __shared__ __align__(16) uint32_t sm32[];
  const unsigned q = blockIdx.x;
  const uint64_t *poly = src + (size_t)q * poly_stride;
#if !HB32_PN_REREAD
  uint32_t dig[PER][2 * WT];                                  // the sign of a coefficient is the top bit of its last digit
#pragma unroll
  for (int k = 0; k < PER; k++) {
    const int i = threadIdx.x + k * G::THREADS;
#pragma unroll
    for (int w = 0; w < WT; w++) {
      const uint64_t word = HB_LD(&poly[(size_t)w * G::N + i]);
      dig[k][2 * w] = (uint32_t)word;
      dig[k][2 * w + 1] = (uint32_t)(word >> 32);
    }
  }
#endif
#pragma unroll 1
  for (unsigned d = 0; d < D; d++) {
    const uint32_t p = T.primes[d].p, p2 = 2 * p;
    const uint2 *pw = T.pow32 + (size_t)d * (T.ndig + 1);
    uint2 c[2 * WT];
#pragma unroll
    for (int w = 0; w < 2 * WT; w++) c[w] = __ldg(&pw[w]);
    const uint32_t negW = p - __ldg(&pw[2 * WT]).x;
#if HB32_PN_REREAD
    // the words are read again for every limb (from L2 after the first): 64 registers of digits less per thread
    constexpr int RR = HB32_PN_REREAD;
#pragma unroll RR
    for (int k = 0; k < PER; k++) {
      const int i = threadIdx.x + k * G::THREADS;
      uint32_t dg[2 * WT];
#pragma unroll
      for (int w = 0; w < WT; w++) {
        const uint64_t word = __ldg(&poly[(size_t)w * G::N + i]);
        dg[2 * w] = (uint32_t)word;
        dg[2 * w + 1] = (uint32_t)(word >> 32);
      }
      uint32_t acc = shoup32(dg[0], c[0].x, c[0].y, p);
#pragma unroll
      for (int w = 1; w < 2 * WT; w++) acc = csub32(acc + shoup32(dg[w], c[w].x, c[w].y, p), p2);
      sm32[pidx32(threadIdx.x) + k * HB32_PSTEP(G::THREADS)] = acc + (negW & (uint32_t)((int32_t)dg[2 * WT - 1] >> 31));      // below 3p
    }
#else
#pragma unroll
    for (int k = 0; k < PER; k++) {
      uint32_t acc = shoup32(dig[k][0], c[0].x, c[0].y, p);
#pragma unroll
      for (int w = 1; w < 2 * WT; w++) acc = csub32(acc + shoup32(dig[k][w], c[w].x, c[w].y, p), p2);
      sm32[pidx32(threadIdx.x) + k * HB32_PSTEP(G::THREADS)] = acc + (negW & (uint32_t)((int32_t)dig[k][2 * WT - 1] >> 31));      // below 3p
    }
#endif
    ntt32_forward<LOGN>(sm32, tw32_of(T, d, false), p);
    uint4 *dst = reinterpret_cast<uint4 *>(out + ((size_t)q * D + d) * G::N);
#pragma unroll
    for (int k = 0; k < PER / 4; k++) {
      const int i4 = threadIdx.x + k * G::THREADS;
      uint4 v;
      const int b4 = pidx32(4 * i4);                          // four consecutive words never straddle a pad word
      v.x = sm32[b4]; v.y = sm32[b4 + 1]; v.z = sm32[b4 + 2]; v.w = sm32[b4 + 3];
      __stcs(&dst[i4], v);
    }
    __syncthreads();                                          // the buffer is refilled for the next limb
  }
}

// One CTA per (polynomial q, limb d): the NJ rotated transforms hat[j][q][d] are read once into registers and serve every
// giant step g < ngroups:  chat[g][q][d] = y-weighted invntt( sum_j hat[j][q][d] o pthat[g*NJ + j][d] ).
template <int LOGN, int NJ>
__global__ void __launch_bounds__(Cfg32<LOGN>::THREADS, HB32_MA_MINB)
k32_mulacc_intt_all(const uint32_t *__restrict__ hat, size_t hat_jstride, const uint32_t *__restrict__ pthat_a, const uint32_t *__restrict__ pthat_b,
                    unsigned q_split, size_t pthat_gstride, const __grid_constant__ hbdev::GroupMap GM, unsigned ngroups, unsigned nq,
                    uint32_t *__restrict__ chat, Tables32 T, unsigned D) {
  // polynomials q < q_split multiply the diagonals pthat_a, the others pthat_b: the two matrix-vector products of ctr_hempc
  // (src/hempc.c:257-259) share every launch (q_split = nq: one matrix)
  typedef Cfg32<LOGN> G;
  constexpr int V = G::N / 4 / G::THREADS;                    // uint4 per thread
  extern __shared__ __align__(16) uint32_t sm32[];
  const unsigned q = blockIdx.x / D, d = blockIdx.x % D;
  const uint32_t *pthat_all = q < q_split ? pthat_a : pthat_b;
  const uint32_t p = T.primes[d].p, p2 = 2 * p;
#if HB32_MA_MONT
  const uint32_t pinv = inv32_mod_2_32(p);
#endif
#if !HB32_MA_REREAD
  uint4 h[NJ][V];
#pragma unroll
  for (int j = 0; j < NJ; j++) {
    const uint4 *a = reinterpret_cast<const uint4 *>(hat + (size_t)j * hat_jstride + ((size_t)q * D + d) * G::N);
#pragma unroll
    for (int k = 0; k < V; k++) h[j][k] = __ldcs(&a[threadIdx.x + k * G::THREADS]);
  }
#endif
  const Tw32 itw = tw32_of(T, d, true);
#pragma unroll 1
  for (unsigned g = 0; g < ngroups; g++) {
    const uint32_t *w = pthat_all + ((size_t)GM.idx[g] * pthat_gstride + d) * HB32_PT_PLANES * G::N;       // strides in limbs
    constexpr int KU = HB32_MA_REREAD ? HB32_MA_REREAD : V;
#pragma unroll KU
    for (int k = 0; k < V; k++) {
      const int i4 = threadIdx.x + k * G::THREADS;
      uint32_t acc[4] = {0, 0, 0, 0};
#pragma unroll
      for (int j = 0; j < NJ; j++) {
        const uint4 *wp = reinterpret_cast<const uint4 *>(w + (size_t)j * D * HB32_PT_PLANES * G::N);
#if HB32_MA_REREAD
        const uint4 hv = __ldg(reinterpret_cast<const uint4 *>(hat + (size_t)j * hat_jstride + ((size_t)q * D + d) * G::N) + i4);
#else
        const uint4 hv = h[j][k];
#endif
#if HB32_MA_MONT
        const uint4 kr = __ldg(&wp[2 * (G::N / 4) + i4]);                   // Montgomery plane: one 16-byte load per four products
        acc[0] = csub32(acc[0] + mont32(hv.x, kr.x, p, pinv), p2);
        acc[1] = csub32(acc[1] + mont32(hv.y, kr.y, p, pinv), p2);
        acc[2] = csub32(acc[2] + mont32(hv.z, kr.z, p, pinv), p2);
        acc[3] = csub32(acc[3] + mont32(hv.w, kr.w, p, pinv), p2);
#else
        const uint4 kw = __ldg(&wp[i4]), kq = __ldg(&wp[G::N / 4 + i4]);    // elements 4 i4 .. 4 i4 + 3 of both planes
        acc[0] = csub32(acc[0] + shoup32(hv.x, kw.x, kq.x, p), p2);
        acc[1] = csub32(acc[1] + shoup32(hv.y, kw.y, kq.y, p), p2);
        acc[2] = csub32(acc[2] + shoup32(hv.z, kw.z, kq.z, p), p2);
        acc[3] = csub32(acc[3] + shoup32(hv.w, kw.w, kq.w, p), p2);
#endif
      }
#pragma unroll
      for (int e = 0; e < 4; e++) sm32[pidx32(4 * i4) + e] = acc[e];
    }
    ntt32_inverse<LOGN>(sm32, itw, p);
    uint4 *dst = reinterpret_cast<uint4 *>(chat + (((size_t)g * nq + q) * D + d) * G::N);
#pragma unroll
    for (int k = 0; k < V; k++) {
      const int i4 = threadIdx.x + k * G::THREADS;
      uint4 v;
      const int b4 = pidx32(4 * i4);                          // four consecutive words never straddle a pad word
      v.x = sm32[b4]; v.y = sm32[b4 + 1]; v.z = sm32[b4 + 2]; v.w = sm32[b4 + 3];
      __stcs(&dst[i4], v);
    }
    __syncthreads();
  }
}

}  // namespace hb32

namespace hb32 {

// k32_recon_smod with the limb count fixed at compile time and four consecutive coefficients per thread: one 16-byte
// load per limb, 32-bit digit chains (one IMAD.WIDE per digit), 16-byte stores.  Same integers as k32_recon_smod.
#ifndef HB32_RS4_MINB
#define HB32_RS4_MINB 1
#endif
template <int WK, int DD>
__global__ void __launch_bounds__(128, HB32_RS4_MINB)
k32_recon_smod4(const uint32_t *__restrict__ chat, uint64_t *__restrict__ out, int Wout, int k, Tables32 T, unsigned npoly) {
  const unsigned N = T.N, N4 = N >> 2;
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (size_t)npoly * N4) return;
  const unsigned q = gid / N4, i4 = gid % N4;
  const uint4 *limbs = reinterpret_cast<const uint4 *>(chat + (size_t)q * DD * N) + i4;
  const uint64_t *qhat = T.crt_qhat + (size_t)DD * T.nprimes * T.WC;
  uint32_t acc[4][2 * WK];
  uint64_t F[4];
#pragma unroll
  for (int e = 0; e < 4; e++) {
    F[e] = 0;
#pragma unroll
    for (int w = 0; w < 2 * WK; w++) acc[e][w] = 0;
  }
  uint4 yv[DD];
#pragma unroll
  for (int d = 0; d < DD; d++) yv[d] = __ldcs(&limbs[(size_t)d * N4]);
#pragma unroll
  for (int d = 0; d < DD; d++) {
    const uint4 pcv = __ldg(reinterpret_cast<const uint4 *>(&T.primes[d]));     // p, ninv, ninv_q, mu
    const uint32_t p = pcv.x, mu = pcv.w;
    uint32_t qd[2 * WK];
#pragma unroll
    for (int w = 0; w < WK; w++) {
      const uint64_t v = (w < (int)T.WC) ? __ldg(&qhat[(size_t)d * T.WC + w]) : 0;
      qd[2 * w] = (uint32_t)v;
      qd[2 * w + 1] = (uint32_t)(v >> 32);
    }
    const uint32_t y[4] = {csub32(yv[d].x, p), csub32(yv[d].y, p), csub32(yv[d].z, p), csub32(yv[d].w, p)};
#pragma unroll
    for (int e = 0; e < 4; e++) {
      F[e] += ((uint64_t)y[e] * mu) >> 5;
      uint32_t carry = 0;
#pragma unroll
      for (int w = 0; w < 2 * WK; w++) {
        const uint64_t t = (uint64_t)y[e] * qd[w] + acc[e][w] + carry;
        acc[e][w] = (uint32_t)t;
        carry = (uint32_t)(t >> 32);
      }
    }
  }
  uint32_t Qd[2 * WK];
#pragma unroll
  for (int w = 0; w < WK; w++) {
    const uint64_t v = (w < (int)T.WC) ? __ldg(&T.crt_Q[(size_t)DD * T.WC + w]) : 0;
    Qd[2 * w] = (uint32_t)v;
    Qd[2 * w + 1] = (uint32_t)(v >> 32);
  }
  uint64_t x[4][WK];
#pragma unroll
  for (int e = 0; e < 4; e++) {
    const uint32_t v = (uint32_t)((F[e] + (1ULL << 55)) >> 56);
    uint32_t borrow = 0;
#pragma unroll
    for (int w = 0; w < 2 * WK; w++) {
      const uint64_t t = (uint64_t)v * Qd[w] + borrow;
      const uint32_t sub = (uint32_t)t;
      borrow = (uint32_t)(t >> 32) + (acc[e][w] < sub);
      acc[e][w] -= sub;
    }
#pragma unroll
    for (int w = 0; w < WK; w++) x[e][w] = (uint64_t)acc[e][2 * w] | ((uint64_t)acc[e][2 * w + 1] << 32);
    hbdev::sext_bits<WK>(x[e], k);
  }
  uint64_t *o = out + (size_t)q * Wout * N + 4 * (size_t)i4;
#pragma unroll
  for (int w = 0; w < WK; w++)
    if (w < Wout) {
      ulonglong2 *dst = reinterpret_cast<ulonglong2 *>(o + (size_t)w * N);
      dst[0] = make_ulonglong2(x[0][w], x[1][w]);
      dst[1] = make_ulonglong2(x[2][w], x[3][w]);
    }
}

}  // namespace hb32

namespace hb32 {

// ===============================================================================================================
// Key switching on a HYBRID basis (round 2).  A key switch needs the exact product x = sigma(c1) (*) key twice: modulo
// the primes of P, to rebuild R = x mod P exactly (he-automorphism.c:68-76: rdiv by P), and as an integer modulo 2^k.
// The reference takes D = 7 of its 59-bit primes so that x (below 2^406 at config A) does not wrap P_D.  Only the dim = 3
// primes of P are dictated by the result; for the rest ANY primes whose product, times P, exceeds 2 |x| reconstruct the
// same integer.  Here the rest are 30-bit primes of the engine's own basis (hb_tables32.hpp): a Harvey butterfly on them
// costs a quarter of a 59-bit one, and twice as many limbs are needed.  x fits because a switching key is smod(., P q_L)
// (he-kem.c:97-102) -- the bound is MEASURED per key at hb_evk_upload (hbkg::k_lift_convert), which also produces the key's
// residues on the 30-bit primes.  Basis M = P * Q_m; limbs arrive at the tail pre-weighted by (M/prime)^-1.
//
// The kernels below are the 30-bit halves; the limbs of P stay on k_ks_baby / k_ks_limb (hb_kernels.cuh) with a limb range.
// Keys: key32[h][i][2][N] u32 per key (multiplicand plane, quotient plane), times N^-1, transform domain.
// dec32[i][2 W + 1]: digit weights 2^(32 j) (M/q_i)^-1 mod q_i (the CRT weight rides on the decomposition).
// ===============================================================================================================
struct Baby32 {
  const uint32_t *key[4];        // rk[j] on the 30-bit primes
  const uint16_t *perm[4];       // perm_{g_j}[N] (hb_kernels.cuh BabySet)
  unsigned nj;
};
struct Groups32 {
  const uint32_t *key[HB_MAX_GROUPS];
  unsigned g[HB_MAX_GROUPS];     // automorphism exponent of each giant step
};

// residue (below 3q) of the signed W-word coefficient, times the weight folded into `pw`
// WT > 0: the word count as a compile-time constant (the loads of all words, and of the coefficients an unrolled caller
// handles together, are issued before the first product)
template <int WT = 0>
__device__ __forceinline__ uint32_t decompose32(const uint64_t *__restrict__ poly, int i, int Wdyn, unsigned N, const uint2 *__restrict__ pw,
                                                uint32_t p, uint32_t p2, uint32_t negW) {
  uint32_t acc = 0;
  uint64_t top = 0;
  const int W = WT > 0 ? WT : Wdyn;
#pragma unroll
  for (int w = 0; w < W; w++) {
    top = HB_LD(&poly[(size_t)w * N + i]);
    const uint2 c0 = __ldg(&pw[2 * w]), c1 = __ldg(&pw[2 * w + 1]);
    acc = csub32(acc + shoup32((uint32_t)top, c0.x, c0.y, p), p2);
    acc = csub32(acc + shoup32((uint32_t)(top >> 32), c1.x, c1.y, p), p2);
  }
  return acc + (negW & (uint32_t)((int64_t)top >> 63));
}

// Baby rotations of he_gemv (he-algo.c:66-68) on the 30-bit limbs: ONE decomposition and forward transform of ct.c1 per limb,
// then per rotation j and half h the permuted product with rk[j] and an inverse transform (same identity as k_ks_baby).
//   chat32[j][b][h][i][N] = y-weighted invntt( perm_j( ntt(c1 mod q_i) ) o rk[j].p_h[i] ),  lazily reduced (below 2q)
// resident CTAs per SM the two kernels are compiled for [B200, batch 4096]: k32_ks_baby 4.28 ms per launch at 5 (51 registers),
// 4.55 at 4, 4.42 at 6 (spills); k32_ks_limb_sum 4.33 ms at 4 (64 registers), 5.26 at 3  (profiles/r02/variants_hybrid.log).
// Measured and dropped: taking the first-pass tiles straight from global words / register digits / products and handing the
// last-pass tiles to global memory or the accumulators (two shared-memory round trips and two barriers fewer per transform)
// changed k32_ks_baby 8.55 -> 8.65, k32_ks_limb_sum 8.66 -> 8.48, k32_poly_ntt_all 6.76 -> 6.66 and k32_mulacc_intt_all
// 6.05 -> 8.01 ms per step (its 64-byte-per-thread runs halve the coalescing): the 30-bit kernels are not bound by their
// shared-memory traffic (profiles/r02/fused_passes.txt).
// planes of an evaluation key on the 30-bit primes: 1 = Montgomery form only (mont32: one load per product), 2 = Shoup pairs
#ifndef HB32_KEY_PLANES
#define HB32_KEY_PLANES 1
#endif
#ifndef HB32_KSB_MINB
#define HB32_KSB_MINB 6      /* after the transposed twiddles and the static word count: 7.11 ms at 5, 6.99 at 6, 7.34 at 4 (variants_occ.log) */
#endif
#ifndef HB32_KSS_MINB
#define HB32_KSS_MINB 4
#endif
#ifndef HB32_KSS_UNROLL
#define HB32_KSS_UNROLL 4
#endif
#ifndef HB32_KSS_PUNROLL
#define HB32_KSS_PUNROLL 8
#endif
#ifndef HB32_KSB_PUNROLL
#define HB32_KSB_PUNROLL 8
#endif
template <int LOGN, int WT>
__global__ void __launch_bounds__(Cfg32<LOGN>::THREADS, HB32_KSB_MINB)
k32_ks_baby(const uint64_t *__restrict__ src, size_t src_bstride, int W, const __grid_constant__ Baby32 S, unsigned m, unsigned mkey,
            const uint2 *__restrict__ dec32, unsigned decstride, uint32_t *__restrict__ chat, size_t chat_jstride, Tables32 T) {
  // m: limbs of this launch; mkey: limbs per half in the key arrays (hb_ctx::hyb_mmax)
  typedef Cfg32<LOGN> G;
  extern __shared__ __align__(16) uint32_t sm32[];
  uint32_t *smA = sm32, *smB = sm32 + G::SMEM_WORDS;
  // (limb-major CTA order -- CTAs resident together share a prime, its twiddles and keys -- measured slower: 4.36 vs 4.28 ms)
  const unsigned b = blockIdx.x / m, i = blockIdx.x % m;
  const uint32_t p = T.primes[i].p, p2 = 2 * p;
#if HB32_KEY_PLANES == 1
  const uint32_t pinv = inv32_mod_2_32(p);
#endif
  const uint2 *pw = dec32 + (size_t)i * decstride;
  const uint32_t negW = p - __ldg(&pw[2 * W]).x;
  const uint64_t *poly = src + (size_t)b * src_bstride;
  constexpr int PER = G::N / G::THREADS, PST = HB32_PSTEP(G::THREADS), KSB_PUNROLL = HB32_KSB_PUNROLL;
  const int pb = pidx32(threadIdx.x);
#pragma unroll 4
  for (int k = 0; k < PER; k++) smA[pb + k * PST] = decompose32<WT>(poly, threadIdx.x + k * G::THREADS, W, G::N, pw, p, p2, negW);
  ntt32_forward<LOGN>(smA, tw32_of(T, i, false), p);
  const Tw32 itw = tw32_of(T, i, true);
#pragma unroll 1
  for (unsigned j = 0; j < S.nj; j++) {
    const uint16_t *perm = S.perm[j];
#pragma unroll 1
    for (int h = 0; h < 2; h++) {
#if HB32_KEY_PLANES == 1
      const uint32_t *kr = S.key[j] + ((size_t)h * mkey + i) * G::N;
#pragma unroll KSB_PUNROLL
      for (int k = 0; k < PER; k++) {
        const int c = threadIdx.x + k * G::THREADS;
        const unsigned s = __ldg(&perm[c]);
        smB[pb + k * PST] = mont32(smA[pidx32(s)], __ldg(&kr[c]), p, pinv);                 // below 2q
      }
#else
      const uint32_t *kw = S.key[j] + ((size_t)h * mkey + i) * 2 * G::N, *kq = kw + G::N;
#pragma unroll KSB_PUNROLL
      for (int k = 0; k < PER; k++) {
        const int c = threadIdx.x + k * G::THREADS;
        const unsigned s = __ldg(&perm[c]);
        smB[pb + k * PST] = shoup32(smA[pidx32(s)], __ldg(&kw[c]), __ldg(&kq[c]), p);      // below 2q
      }
#endif
      ntt32_inverse<LOGN>(smB, itw, p);
      uint32_t *dst = chat + (size_t)j * chat_jstride + (((size_t)b * 2 + h) * m + i) * G::N;
#pragma unroll 4
      for (int k = 0; k < PER; k++) __stcs(&dst[threadIdx.x + k * G::THREADS], smB[pb + k * PST]);
      __syncthreads();                                       // smB is refilled
    }
  }
}

// Giant steps of he_gemv (he-algo.c:80-84) on the 30-bit limbs, summed over the giant steps in the transform domain (the
// 30-bit twin of k_ks_limb_sum):
//   chatX[b][h][i][N] = y-weighted invntt( sum_g ntt( sigma_g(inner_g[b].c1) mod q_i ) o rk[g n1].p_h[i] )
template <int LOGN, int WT>
__global__ void __launch_bounds__(Cfg32<LOGN>::THREADS, HB32_KSS_MINB)
k32_ks_limb_sum(const uint64_t *__restrict__ src, size_t src_bstride, size_t src_gstride, int W, const __grid_constant__ Groups32 S,
                unsigned ng, unsigned m, unsigned mkey, const uint2 *__restrict__ dec32, unsigned decstride, uint32_t *__restrict__ chatX, Tables32 T) {
  typedef Cfg32<LOGN> G;
  extern __shared__ __align__(16) uint32_t sm32[];
  uint32_t *smA = sm32;
  const unsigned b = blockIdx.x / m, i = blockIdx.x % m;      // (limb-major order measured slower: 4.92 vs 4.33 ms)
  const uint32_t p = T.primes[i].p, p2 = 2 * p;
#if HB32_KEY_PLANES == 1
  const uint32_t pinv = inv32_mod_2_32(p);
#endif
  const uint2 *pw = dec32 + (size_t)i * decstride;
  const uint32_t negW = p - __ldg(&pw[2 * W]).x;
  const Tw32 tw = tw32_of(T, i, false);
  constexpr int PER = G::N / G::THREADS, PST = HB32_PSTEP(G::THREADS), KSS_UNROLL = HB32_KSS_UNROLL, KSS_PUNROLL = HB32_KSS_PUNROLL;
  const int pb = pidx32(threadIdx.x);
#pragma unroll 1
  for (unsigned g = 0; g < ng; g++) {
    const uint64_t *poly = src + (size_t)g * src_gstride + (size_t)b * src_bstride;
    const unsigned ge = S.g[g];
    // sigma_g scatters coefficient c to c g mod 2N, negated past N (poly_rot, poly.c:263-283)
#pragma unroll KSS_UNROLL
    for (int c = threadIdx.x; c < G::N; c += G::THREADS) {
      const uint32_t v = decompose32<WT>(poly, c, W, G::N, pw, p, p2, negW);        // below 3q
      const unsigned k = ((unsigned)c * ge) & (2 * G::N - 1);
      smA[pidx32(k & (G::N - 1))] = (k >= (unsigned)G::N) ? 3 * p - v : v;
    }
    ntt32_forward<LOGN>(smA, tw, p);
#pragma unroll 1
    for (int h = 0; h < 2; h++) {
      uint32_t *acc = sm32 + (size_t)(1 + h) * G::SMEM_WORDS;
#if HB32_KEY_PLANES == 1
      const uint32_t *kr = S.key[g] + ((size_t)h * mkey + i) * G::N;
#else
      const uint32_t *kw = S.key[g] + ((size_t)h * mkey + i) * 2 * G::N, *kq = kw + G::N;
#endif
#pragma unroll KSS_PUNROLL
      for (int k = 0; k < PER; k++) {
        const int c = threadIdx.x + k * G::THREADS, o = pb + k * PST;
#if HB32_KEY_PLANES == 1
        uint32_t v = mont32(smA[o], __ldg(&kr[c]), p, pinv);                            // below 2q
#else
        uint32_t v = shoup32(smA[o], __ldg(&kw[c]), __ldg(&kq[c]), p);                  // below 2q
#endif
        if (g) v = csub32(v + acc[o], p2);
        acc[o] = v;
      }
    }
    __syncthreads();                                         // smA is refilled; the accumulators are thread-private
  }
  const Tw32 itw = tw32_of(T, i, true);
#pragma unroll 1
  for (int h = 0; h < 2; h++) {
    uint32_t *acc = sm32 + (size_t)(1 + h) * G::SMEM_WORDS;
    ntt32_inverse<LOGN>(acc, itw, p);
    uint32_t *dst = chatX + (((size_t)b * 2 + h) * m + i) * G::N;
#pragma unroll 4
    for (int k = 0; k < PER; k++) __stcs(&dst[threadIdx.x + k * G::THREADS], acc[pb + k * PST]);
  }
}

// plain forward transform of u32 limbs (limb j uses prime j % D), canonical output: the key conversion at hb_evk_upload
template <int LOGN>
__global__ void __launch_bounds__(Cfg32<LOGN>::THREADS)
k32_ntt_fwd_plain(const uint32_t *__restrict__ in, uint32_t *__restrict__ out, Tables32 T, unsigned D) {
  typedef Cfg32<LOGN> G;
  extern __shared__ __align__(16) uint32_t sm32[];
  const unsigned i = blockIdx.x % D;
  const uint32_t p = T.primes[i].p, p2 = 2 * p;
  const uint32_t *src = in + (size_t)blockIdx.x * G::N;
  for (int c = threadIdx.x; c < G::N; c += G::THREADS) sm32[pidx32(c)] = src[c];
  ntt32_forward<LOGN>(sm32, tw32_of(T, i, false), p);
  uint32_t *dst = out + (size_t)blockIdx.x * G::N;
  for (int c = threadIdx.x; c < G::N; c += G::THREADS) {
    uint32_t v = sm32[pidx32(c)];
    v = csub32(v, p2);
    dst[c] = csub32(v, p);
  }
}

}  // namespace hb32
