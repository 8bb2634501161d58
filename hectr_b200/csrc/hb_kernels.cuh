// sm_100a kernels of the CKKS evaluation path.  Every kernel cites the reference routine that defines its
// results.  Ciphertext polynomials are planar little-endian two's-complement words  poly[w*N + i]
// (w = 0 least significant), RNS limbs are u64[limb][N].
#pragma once
#include "hb_dev.cuh"

#ifndef HB_KS_MINB
#define HB_KS_MINB 3     /* resident CTAs per SM the two-buffer key-switch kernels are compiled for, N <= 2^12 */
#endif

namespace hbdev {

// Several key switches in one launch ("groups": the n1 baby or the n2 giant rotations of one he_gemv).  Instance b of
// the launch belongs to group b / gbatch; each group has its own automorphism and key.  Single key switch: one group.
#define HB_MAX_GROUPS 16
struct KsGroups {
  const ulonglong2 *evk[HB_MAX_GROUPS];
  unsigned g[HB_MAX_GROUPS], ginv[HB_MAX_GROUPS];
  unsigned gbatch;
};
// giant steps taken by a launch: slot s of the launch is giant step idx[s] of the matrix (identically zero giant steps of a
// sparse matrix are skipped: their inner sum is the zero ciphertext, whose key switch is zero again)
struct GroupMap {
  unsigned idx[HB_MAX_GROUPS];
};
struct TailGroups {
  unsigned ginv[HB_MAX_GROUPS];
  unsigned gbatch;
  unsigned add_shared;     // 1: every group adds sigma(add[b % gbatch]) (baby steps: the same source), 0: add[b]
};

// two shared-memory buffers per CTA: 3 x 74 KB fit one SM up to N = 2^12; from 2^13 (M = 8192, 147 KB) only one CTA does
template <int LOGN>
struct KsBlocks { static constexpr int value = (LOGN <= 12) ? HB_KS_MINB : 1; };

// =========================================================================================================
// K2 / K3: stand-alone batched transforms (config B microbench; building blocks of he_mul)
// =========================================================================================================
// Results = ntt() of GPQHE/src/ntt.c:37-52.  STRICT reproduces the reference's [0,p] representation bit for
// bit; the default returns the same residues canonicalised to [0,p) (p -> 0).  Inputs below 4p.
template <int LOGN, bool STRICT>
__global__ void __launch_bounds__(Cfg<LOGN>::THREADS, Cfg<LOGN>::MINBLOCKS)
k_ntt_fwd(const uint64_t *__restrict__ in, uint64_t *__restrict__ out, Tables T, unsigned d0, unsigned D) {
  typedef Cfg<LOGN> G;
  extern __shared__ __align__(16) uint64_t smem[];
  const unsigned limb = blockIdx.x / G::C, chunk = chunk_rank<LOGN>();
  const unsigned d = d0 + limb % D;
  const Mod m = make_mod(T.primes[d].p);
  const Tw tw = make_tw(T, d, false);
  const uint64_t *src = in + (size_t)limb * G::N;
  if constexpr (G::C == 1) {
#pragma unroll 8
    for (int i = threadIdx.x; i < G::N; i += G::THREADS) smem[pidx(i)] = HB_LD(&src[i]);
    ntt_forward_resident<LOGN, STRICT, 4>(smem, tw, m);
  } else {
    ntt_forward<LOGN, STRICT, 4>(smem, [&](int gi) { return HB_LD(&src[gi]); }, tw, m);
  }
  uint64_t *dst = out + (size_t)limb * G::N + (size_t)chunk * G::M;
  for (int i = threadIdx.x; i < G::M; i += G::THREADS) {
    uint64_t v = smem[pidx(i)];
    dst[i] = STRICT ? v : canon(v, m);
  }
}

// Results = invntt() of GPQHE/src/ntt.c:54-73 (canonical [0,p)).  Inputs may be any value below 2p.
template <int LOGN>
__global__ void __launch_bounds__(Cfg<LOGN>::THREADS, Cfg<LOGN>::MINBLOCKS)
k_ntt_inv(const uint64_t *__restrict__ in, uint64_t *__restrict__ out, Tables T, unsigned d0, unsigned D) {
  typedef Cfg<LOGN> G;
  extern __shared__ __align__(16) uint64_t smem[];
  const unsigned limb = blockIdx.x / G::C, chunk = chunk_rank<LOGN>();
  const unsigned d = d0 + limb % D;
  const PrimeConsts pc = T.primes[d];
  const Mod m = make_mod(pc.p);
  const uint64_t *src = in + (size_t)limb * G::N + (size_t)chunk * G::M;
#pragma unroll 8
  for (int i = threadIdx.x; i < G::M; i += G::THREADS) smem[pidx(i)] = HB_LD(&src[i]);
  __syncthreads();
  uint64_t *dst = out + (size_t)limb * G::N;
  ntt_inverse<LOGN, INV_CANON>(smem, [&](int gi, uint64_t v) { HB_ST(&dst[gi], v); }, make_tw(T, d, true), pc, m);
}

// =========================================================================================================
// K1 + K2: signed multiword polynomial -> NTT limbs  (rns_decompose rns.c:37-48, then ntt)
// =========================================================================================================
// src: npoly polynomials, polynomial q at src + q*poly_stride, W words each.  out[q][d][N], canonical.
template <int LOGN>
__global__ void __launch_bounds__(Cfg<LOGN>::THREADS, Cfg<LOGN>::MINBLOCKS)
k_poly_ntt(const uint64_t *__restrict__ src, size_t poly_stride, int W, uint64_t *__restrict__ out, Tables T,
           unsigned D) {
  typedef Cfg<LOGN> G;
  extern __shared__ __align__(16) uint64_t smem[];
  const unsigned idx = blockIdx.x / G::C, chunk = chunk_rank<LOGN>();
  const unsigned q = idx / D, d = idx % D;
  const PrimeConsts pc = T.primes[d];
  const Mod m = make_mod(pc.p);
  const uint64_t negW = pc.p - T.pow64[(size_t)d * (T.WC + 1) + W];
  const uint64_t *poly = src + (size_t)q * poly_stride;
  poly_forward<LOGN>(smem, poly, W, 1u, 1u, pc, m, negW, make_tw(T, d, false));
  uint64_t *dst = out + ((size_t)q * D + d) * G::N + (size_t)chunk * G::M;
  for (int i = threadIdx.x; i < G::M; i += G::THREADS) dst[i] = canon(smem[pidx(i)], m);
}

// canonical residues -> Shoup pairs (w, floor(w 2^64 / p)) for fixed multiplicands (keys, encoded diagonals).
// limbs[j][N] uses prime j % D.  A residue stored as p becomes 0: the same class.
// scale: 0 = as is; 1 = times N^-1 (the inverse transform after the product then skips its own scaling);
//        2 = times N^-1 (P_Dw/p_d)^-1, the CRT weight for Dw limbs as well (the reconstruction skips its multiply).
__global__ void k_shoup_prep(const uint64_t *__restrict__ in, ulonglong2 *__restrict__ out, Tables T, unsigned D,
                             size_t nlimbs, int scale, unsigned Dw) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= nlimbs * T.N) return;
  const unsigned d = (unsigned)((gid / T.N) % D);
  const PrimeConsts pc = T.primes[d];
  const uint64_t p = pc.p;
  uint64_t w = in[gid];
  if (w >= p) w -= p;
  if (scale >= 1) w = shoup(w, pc.ninv, pc.ninv_q, p);
  if (scale >= 2) w = shoup(w, T.crt_inv[(size_t)Dw * T.dimub + d], T.crt_inv_q[(size_t)Dw * T.dimub + d], p);
  const unsigned __int128 num = ((unsigned __int128)w) << 64;
  out[gid] = make_ulonglong2(w, (uint64_t)(num / p));
}

// =========================================================================================================
// key switch, per-limb half:  he_swk  GPQHE/src/he-automorphism.c:59-67  (== he_relin he-mult.c:58-66)
//   chat[b][h][d] = invntt( ntt( decompose( sigma(src_b) ) mod p_d ) o evk.p_h[d] ),  h = 0,1
// sigma is the automorphism X -> X^g fused into the load (poly_rot / poly_conj, poly.c:263-283), given as
// ginv = g^-1 mod 2N:  sigma(a)[k] = +a[k*ginv mod 2N]  if that index is < N, else  -a[(k*ginv mod 2N) - N].
// evk: Shoup pairs [2][dimevk][N] of key * N^-1 (hb_evk_upload), so the inverse transforms do not scale.  The source
// residues are decomposed with the CRT weight (P_D/p_d)^-1 folded into the word weights (decompose_scaled), so chat
// holds y_d = a_d (P_D/p_d)^-1 mod p_d directly, lazily reduced (below 2p): the tail starts at the CRT sum.
// =========================================================================================================
template <int LOGN>
__global__ void __launch_bounds__(Cfg<LOGN>::THREADS, KsBlocks<LOGN>::value)
k_ks_limb(const uint64_t *__restrict__ src, size_t src_bstride, int W, const __grid_constant__ KsGroups S,
          unsigned dimevk, uint64_t *__restrict__ chat, Tables T, unsigned D, unsigned d0, unsigned Dn,
          const ulonglong2 *__restrict__ dec_override) {
  // D: limbs of the CRT basis (selects the pre-weights); this launch covers limbs [d0, d0 + Dn) and writes them compactly:
  // chat[b][h][Dn][N].  A whole key switch has d0 = 0, Dn = D.  dec_override: pre-weights of another basis, [limb][WC + 1]
  // (the hybrid basis of hb_dev32.cuh), instead of the D-limb table.
  typedef Cfg<LOGN> G;
  extern __shared__ __align__(16) uint64_t smem[];
  uint64_t *smA = smem, *smB = smem + G::SMEM_WORDS;
  const unsigned idx = blockIdx.x / G::C, chunk = chunk_rank<LOGN>();
  const unsigned b = idx / Dn, dd = idx % Dn, d = d0 + dd;
  const PrimeConsts pc = T.primes[d];
  const Mod m = make_mod(pc.p);
  const uint64_t negW = pc.p - T.pow64[(size_t)d * (T.WC + 1) + W];
  const uint64_t *poly = src + (size_t)b * src_bstride;
  const ulonglong2 *dc = dec_override ? dec_override + (size_t)d * (T.WC + 1) : T.dec_s + ((size_t)D * T.dimub + d) * (T.WC + 1);
  const unsigned grp = b / S.gbatch;
  const ulonglong2 *evk = S.evk[grp];
  poly_forward<LOGN, true>(smA, poly, W, S.g[grp], S.ginv[grp], pc, m, negW, make_tw(T, d, false), dc);
#pragma unroll 1
  for (int h = 0; h < 2; h++) {
    const ulonglong2 *key = evk + ((size_t)h * dimevk + d) * G::N + (size_t)chunk * G::M;
#pragma unroll 4
    for (int i = threadIdx.x; i < G::M; i += G::THREADS) {
      const ulonglong2 k = HB_LDS2(&key[i]);
      smB[pidx(i)] = shoup_lazy(smA[pidx(i)], k.x, k.y, m);
    }
    __syncthreads();
    uint64_t *dst = chat + (((size_t)b * 2 + h) * Dn + dd) * G::N;
    ntt_inverse<LOGN, INV_RAW>(smB, [&](int gi, uint64_t v) { HB_ST(&dst[gi], v); }, make_tw(T, d, true), pc, m);
  }
}

// =========================================================================================================
// Giant steps of he_gemv (GPQHE/src/he-algo.c:80-84), limbs that do NOT divide P, summed over the giant steps:
//   chatX[b][h][d - d0] = y-weighted invntt( sum_g ntt( decompose( sigma_g(inner_g[b].c1) ) mod p_d ) o rk[g n1].p_h[d] )
// he_gemv adds the n2 key-switched inner sums.  Each key switch is  c_g = (x_g - R_g)/P + round_g (+ sigma_g(c0)), with x_g the
// exact product and R_g = x_g mod P, so  sum_g c_g = (X - sum_g R_g)/P + sum_g round_g,  X = sum_g x_g.  R_g needs x_g modulo
// the primes of P only (limbs d < dim: k_ks_limb with d0 = 0, Dn = dim); for every other limb the sum is taken in the transform
// domain and ONE inverse transform per half serves all giant steps.  X is the same integer as long as |X| < P_D / 2, which the
// host checks from the key bound recorded at hb_evk_upload (hb_api.cu gemv_fused).  The he_add chain only re-centres modulo
// q_l, so the words are the reference's.  One limb per CTA (N <= 2^13), three shared-memory buffers.
// =========================================================================================================
template <int LOGN>
struct KsSumBlocks { static constexpr int value = (LOGN <= 12) ? 2 : 1; };

template <int LOGN>
__global__ void __launch_bounds__(Cfg<LOGN>::THREADS, KsSumBlocks<LOGN>::value)
k_ks_limb_sum(const uint64_t *__restrict__ src, size_t src_bstride, size_t src_gstride, int W, const __grid_constant__ KsGroups S,
              unsigned ng, unsigned dimevk, uint64_t *__restrict__ chatX, Tables T, unsigned D, unsigned d0, unsigned Dn) {
  typedef Cfg<LOGN> G;
  static_assert(G::C == 1, "the summed giant-step kernel keeps one limb per CTA");
  extern __shared__ __align__(16) uint64_t smem[];
  uint64_t *smA = smem;
  const unsigned b = blockIdx.x / Dn, dd = blockIdx.x % Dn, d = d0 + dd;
  const PrimeConsts pc = T.primes[d];
  const Mod m = make_mod(pc.p);
  const uint64_t negW = pc.p - T.pow64[(size_t)d * (T.WC + 1) + W];
  const ulonglong2 *dc = T.dec_s + ((size_t)D * T.dimub + d) * (T.WC + 1);
  const Tw tw = make_tw(T, d, false);
#pragma unroll 1
  for (unsigned g = 0; g < ng; g++) {
    const uint64_t *poly = src + (size_t)g * src_gstride + (size_t)b * src_bstride;
    poly_forward<LOGN, true>(smA, poly, W, S.g[g], S.ginv[g], pc, m, negW, tw, dc);
    const bool last = g + 1 == ng;
#pragma unroll 1
    for (int h = 0; h < 2; h++) {
      uint64_t *acc = smem + (size_t)(1 + h) * G::SMEM_WORDS;
      const ulonglong2 *key = S.evk[g] + ((size_t)h * dimevk + d) * G::N;
#pragma unroll 4
      for (int i = threadIdx.x; i < G::N; i += G::THREADS) {
        const ulonglong2 k = HB_LDS2(&key[i]);
        uint64_t v = shoup_lazy(smA[pidx(i)], k.x, k.y, m);              // below 2p; at most 8 terms: below 16p
        if (g) v += acc[pidx(i)];
        acc[pidx(i)] = (last && ng > 1) ? fold(v, m) : v;                // the inverse transform takes values below 2p
      }
    }
    __syncthreads();                                                     // smA is refilled; the accumulators are thread-private
  }
  const Tw itw = make_tw(T, d, true);
#pragma unroll 1
  for (int h = 0; h < 2; h++) {
    uint64_t *acc = smem + (size_t)(1 + h) * G::SMEM_WORDS;
    uint64_t *dst = chatX + (((size_t)b * 2 + h) * Dn + dd) * G::N;
    ntt_inverse<LOGN, INV_RAW>(acc, [&](int gi, uint64_t v) { HB_ST(&dst[gi], v); }, itw, pc, m);
  }
}

// =========================================================================================================
// Baby rotations of he_gemv (GPQHE/src/he-algo.c:66-68), per-limb half of nj key switches of the SAME ciphertext:
//   chat[j][b][h][d] = y-weighted invntt( ntt( decompose( sigma_j(c1_b) ) mod p_d ) o rk[j].p_h[d] ),  j < nj
// sigma_j only permutes (and negates) coefficients, and the negacyclic transform of sigma_g(a) is a permutation of the
// transform of a:  ntt(sigma_g a)[i] = ntt(a)[perm_g[i]],  perm_g[i] = brv(((2 brv(i) + 1) g mod 2N - 1) / 2)
// (position i holds the evaluation at zeta^(2 brv(i) + 1)).  So ONE decomposition and ONE forward transform serve all
// nj rotations; every key product, inverse transform, reconstruction and rounding of the reference is still performed
// on the same residues (the identity is exact modulo p_d), hence the same ciphertexts.
// =========================================================================================================
struct BabySet {
  const ulonglong2 *evk[4];      // rk[j] (Shoup pairs [2][dimevk][N], times N^-1)
  const uint16_t *perm[4];       // perm_{g_j}[N]
  unsigned nj;
};

template <int LOGN>
__global__ void __launch_bounds__(Cfg<LOGN>::THREADS, KsBlocks<LOGN>::value)
k_ks_baby(const uint64_t *__restrict__ src, size_t src_bstride, int W, const __grid_constant__ BabySet S, unsigned dimevk,
          uint64_t *__restrict__ chat, size_t chat_jstride, Tables T, unsigned D, unsigned Dn, const ulonglong2 *__restrict__ dec_override) {
  // limbs [0, Dn) of the D-limb basis (Dn = D: all), written compactly as chat[j][b][h][Dn][N]; dec_override as in k_ks_limb
  typedef Cfg<LOGN> G;
  extern __shared__ __align__(16) uint64_t smem[];
  uint64_t *smA = smem, *smB = smem + G::SMEM_WORDS;
  const unsigned idx = blockIdx.x / G::C, chunk = chunk_rank<LOGN>();
  const unsigned b = idx / Dn, d = idx % Dn;
  const PrimeConsts pc = T.primes[d];
  const Mod m = make_mod(pc.p);
  const uint64_t negW = pc.p - T.pow64[(size_t)d * (T.WC + 1) + W];
  const uint64_t *poly = src + (size_t)b * src_bstride;
  const ulonglong2 *dc = dec_override ? dec_override + (size_t)d * (T.WC + 1) : T.dec_s + ((size_t)D * T.dimub + d) * (T.WC + 1);
  poly_forward<LOGN, true>(smA, poly, W, 1u, 1u, pc, m, negW, make_tw(T, d, false), dc);
  if constexpr (G::C > 1) cg::this_cluster().sync();        // peers' chunks of the transform are complete
  const Tw itw = make_tw(T, d, true);
#pragma unroll 1
  for (unsigned j = 0; j < S.nj; j++) {
    const uint16_t *perm = S.perm[j] + (size_t)chunk * G::M;
#pragma unroll 1
    for (int h = 0; h < 2; h++) {
      const ulonglong2 *key = S.evk[j] + ((size_t)h * dimevk + d) * G::N + (size_t)chunk * G::M;
#pragma unroll 4
      for (int i = threadIdx.x; i < G::M; i += G::THREADS) {
        const ulonglong2 k = HB_LDS2(&key[i]);
        const unsigned s = __ldg(&perm[i]);
        uint64_t v;
        if constexpr (G::C > 1) v = cg::this_cluster().map_shared_rank(smA, s >> G::LOGM)[pidx(s & (G::M - 1))];
        else v = smA[pidx(s)];
        smB[pidx(i)] = shoup_lazy(v, k.x, k.y, m);
      }
      __syncthreads();
      uint64_t *dst = chat + (size_t)j * chat_jstride + (((size_t)b * 2 + h) * Dn + d) * G::N;
      ntt_inverse<LOGN, INV_RAW>(smB, [&](int gi, uint64_t v) { HB_ST(&dst[gi], v); }, itw, pc, m);
    }
  }
}

// =========================================================================================================
// ct x pt, per-limb:  he_mulpt  GPQHE/src/he-mult.c:176-188
//   chat[b][h][d] = (P_D/p_d)^-1 invntt( ntt(decompose(c_h)) o pthat[d] )
// pthat: Shoup pairs [D][N] of ntt(pt) N^-1 (P_D/p_d)^-1 (k_shoup_prep scale 2): no scaling pass, no CRT multiply later.
// =========================================================================================================
template <int LOGN>
__global__ void __launch_bounds__(Cfg<LOGN>::THREADS, Cfg<LOGN>::MINBLOCKS)
k_mulpt_limb(const uint64_t *__restrict__ ct, size_t poly_stride, int W, const ulonglong2 *__restrict__ pthat,
             uint64_t *__restrict__ chat, Tables T, unsigned D) {
  typedef Cfg<LOGN> G;
  extern __shared__ __align__(16) uint64_t smem[];
  const unsigned idx = blockIdx.x / G::C, chunk = chunk_rank<LOGN>();
#ifdef HB_DMAJOR
  const unsigned nq = gridDim.x / (G::C * D);
  const unsigned d = idx / nq, q = idx % nq;        // q = b*2 + h
#else
  const unsigned q = idx / D, d = idx % D;          // q = b*2 + h
#endif
  const PrimeConsts pc = T.primes[d];
  const Mod m = make_mod(pc.p);
  const uint64_t negW = pc.p - T.pow64[(size_t)d * (T.WC + 1) + W];
  const uint64_t *poly = ct + (size_t)q * poly_stride;     // a ciphertext batch: W*N; one polynomial per instance (he_dec): 2*W*N
  poly_forward<LOGN>(smem, poly, W, 1u, 1u, pc, m, negW, make_tw(T, d, false));
  const ulonglong2 *pt = pthat + (size_t)d * G::N + (size_t)chunk * G::M;
#pragma unroll 4
  for (int i = threadIdx.x; i < G::M; i += G::THREADS) {
    const ulonglong2 k = HB_LDS2(&pt[i]);
    smem[pidx(i)] = shoup_lazy(smem[pidx(i)], k.x, k.y, m);
  }
  __syncthreads();
  uint64_t *dst = chat + ((size_t)q * D + d) * G::N;
  ntt_inverse<LOGN, INV_RAW>(smem, [&](int gi, uint64_t v) { HB_ST(&dst[gi], v); }, make_tw(T, d, true), pc, m);
}

// =========================================================================================================
// rotation-sum inner step of he_gemv (GPQHE/src/he-algo.c:64-79), accumulated in the NTT domain:
//   chat[q][d] = invntt( sum_j  hat[j][q][d] o pthat[j][d] ),   q = b*2 + h,  j < n1
// hat[j] = ntt(decompose(rot_j(ct))) (k_poly_ntt), pthat[j] = Shoup pairs of the j-th encoded diagonal of this giant
// step, pre-scaled like k_mulpt_limb's.  The reference multiplies, reconstructs and adds mod q_l one diagonal at a time; the transforms and the CRT
// are linear, so the sum of the exact products is the same integer as long as it stays inside (-P_D/2, P_D/2),
// which the host checks when it builds the plan (hb_api.cu gemv_plan).
// =========================================================================================================
template <int LOGN>
__global__ void __launch_bounds__(Cfg<LOGN>::THREADS, Cfg<LOGN>::MINBLOCKS)
k_mulacc_intt(const uint64_t *__restrict__ hat, size_t hat_jstride, const ulonglong2 *__restrict__ pthat_all, size_t pthat_gstride,
              const __grid_constant__ GroupMap GM, unsigned n1, unsigned nq, uint64_t *__restrict__ chat, Tables T, unsigned D) {
  typedef Cfg<LOGN> G;
  extern __shared__ __align__(16) uint64_t smem[];
  const unsigned idx = blockIdx.x / G::C, chunk = chunk_rank<LOGN>();
  // all giant steps in one launch: group i = qq / nq reads the SAME hat[j][q] with its own diagonals pthat_all + i * gstride
  const unsigned qq = idx / D, d = idx % D, q = qq % nq;
  const ulonglong2 *pthat = pthat_all + (size_t)GM.idx[qq / nq] * pthat_gstride;
  const PrimeConsts pc = T.primes[d];
  const Mod m = make_mod(pc.p);
  const size_t off = (size_t)chunk * G::M;
  const uint64_t *a = hat + ((size_t)q * D + d) * G::N + off;
  const ulonglong2 *w = pthat + (size_t)d * G::N + off;
#pragma unroll 2
  for (int i = threadIdx.x; i < G::M; i += G::THREADS) {
    uint64_t acc = 0;
    for (unsigned j = 0; j < n1; j++) {
      const ulonglong2 k = HB_LDS2(&w[(size_t)j * D * G::N + i]);
      acc += shoup_lazy(HB_LD(&a[(size_t)j * hat_jstride + i]), k.x, k.y, m);      // each term below 2p
      if ((j & 7) == 7) acc = fold(acc, m);
    }
    smem[pidx(i)] = fold(acc, m);
  }
  __syncthreads();
  uint64_t *dst = chat + ((size_t)qq * D + d) * G::N;
  ntt_inverse<LOGN, INV_RAW>(smem, [&](int gi, uint64_t v) { HB_ST(&dst[gi], v); }, make_tw(T, d, true), pc, m);
}

// =========================================================================================================
// ct x ct tensor, per-limb:  he_mul  GPQHE/src/he-mult.c:116-138
//   in  hatA[b][2][D][N], hatB[b][2][D][N] = ntt of (a.c0, a.c1), (b.c0, b.c1)      (k_poly_ntt, canonical)
//   out dhat[b][3][D][N] = invntt of (a0 b0, a0 b1 + a1 b0, a1 b1)                   (lazily reduced, below 2p)
// The reference inverts a0 b1 and a1 b0 separately and adds mod p afterwards; the inverse transform is linear
// mod p, so adding first gives the same residues.
// =========================================================================================================
template <int LOGN>
__global__ void __launch_bounds__(Cfg<LOGN>::THREADS, Cfg<LOGN>::MINBLOCKS)
k_tensor_intt(const uint64_t *__restrict__ hatA, const uint64_t *__restrict__ hatB, uint64_t *__restrict__ dhat, Tables T,
              unsigned D) {
  typedef Cfg<LOGN> G;
  extern __shared__ __align__(16) uint64_t smem[];
  const unsigned idx = blockIdx.x / G::C, chunk = chunk_rank<LOGN>();
  const unsigned d = idx % D, which = (idx / D) % 3, b = idx / (3 * D);
  const PrimeConsts pc = T.primes[d];
  const Mod m = make_mod(pc.p);
  const uint64_t p = pc.p;
  const size_t off = (size_t)d * G::N + (size_t)chunk * G::M;
  const uint64_t *a0 = hatA + ((size_t)b * 2 + 0) * D * G::N + off, *a1 = hatA + ((size_t)b * 2 + 1) * D * G::N + off;
  const uint64_t *b0 = hatB + ((size_t)b * 2 + 0) * D * G::N + off, *b1 = hatB + ((size_t)b * 2 + 1) * D * G::N + off;
  for (int i = threadIdx.x; i < G::M; i += G::THREADS) {
    uint64_t v;
    if (which == 0) v = mulmod_barrett(a0[i], b0[i], p, pc.mu);
    else if (which == 2) v = mulmod_barrett(a1[i], b1[i], p, pc.mu);
    else v = addmod(mulmod_barrett(a0[i], b1[i], p, pc.mu), mulmod_barrett(a1[i], b0[i], p, pc.mu), p);
    smem[pidx(i)] = v;
  }
  __syncthreads();
  uint64_t *dst = dhat + (((size_t)b * 3 + which) * D + d) * G::N;
  ntt_inverse<LOGN, INV_LAZY>(smem, [&](int gi, uint64_t v) { HB_ST(&dst[gi], v); }, make_tw(T, d, true), pc, m);
}

// =========================================================================================================
// K5: exact CRT reconstruction, low words only.
//   rns_reconstruct rns.c:60-75 + centring mod P_D (poly.c:116):  x = sum_d y_d (P_D/p_d) - v P_D,
//   y_d = a_d (P_D/p_d)^-1 mod p_d,  v = round(sum_d y_d / p_d)  (120-bit fixed point, error < D 2^-60).
//   Only x mod 2^(64 WK) is formed: every consumer reduces modulo a power of two afterwards.
// =========================================================================================================
// PRE: the limbs already carry the weight (a(d) = y_d, lazily reduced below 2p) -- k_ks_limb, k_mulpt_limb, k_mulacc_intt.
template <int WK, bool PRE, class Get>
__device__ __forceinline__ void crt_low(Get a, unsigned D, const Tables &T, uint64_t (&acc)[WK]) {
#pragma unroll
  for (int w = 0; w < WK; w++) acc[w] = 0;
  uint64_t Flo = 0, Fhi = 0;
  const uint64_t *inv = T.crt_inv + (size_t)D * T.dimub, *invq = T.crt_inv_q + (size_t)D * T.dimub;
  const uint64_t *phat = T.crt_phat + (size_t)D * T.dimub * T.WC;
  for (unsigned d = 0; d < D; d++) {
    const uint64_t p = T.primes[d].p, mu = T.primes[d].mu;
    const uint64_t y = PRE ? csub(a(d), p) : shoup(a(d), inv[d], invq[d], p);
    uint64_t lo = y * mu, hi = __umul64hi(y, mu);
    Flo += lo;
    Fhi += hi + (Flo < lo);
    uint64_t carry = 0;
#pragma unroll
    for (int w = 0; w < WK; w++) {
      const uint64_t pw = (w < (int)T.WC) ? phat[(size_t)d * T.WC + w] : 0;
      uint64_t l = y * pw, h = __umul64hi(y, pw);
      uint64_t s = acc[w] + l;
      h += (s < l);
      uint64_t s2 = s + carry;
      h += (s2 < carry);
      acc[w] = s2;
      carry = h;
    }
  }
  const uint64_t v = (Fhi + (1ULL << 55)) >> 56;
  const uint64_t *P = T.crt_P + (size_t)D * T.WC;
  uint64_t borrow = 0;
#pragma unroll
  for (int w = 0; w < WK; w++) {
    const uint64_t pw = (w < (int)T.WC) ? P[w] : 0;
    uint64_t l = v * pw, h = __umul64hi(v, pw);
    uint64_t sub = l + borrow;
    h += (sub < l);
    h += (acc[w] < sub);
    acc[w] -= sub;
    borrow = h;
  }
}

// out polynomial q (of npoly) = smod(reconstruct(chat[q][0..D)), 2^k)   -- poly_rns2mpi, poly.c:109-120
template <int WK, bool PRE>
__global__ void k_recon_smod(const uint64_t *__restrict__ chat, unsigned D, uint64_t *__restrict__ out, int Wout,
                             int k, Tables T, unsigned npoly) {
  const unsigned N = T.N;
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (size_t)npoly * N) return;
  const unsigned q = gid / N, i = gid % N;
  const uint64_t *limbs = chat + (size_t)q * D * N + i;
  uint64_t acc[WK];
  crt_low<WK, PRE>([&](unsigned d) { return limbs[(size_t)d * N]; }, D, T, acc);
  sext_bits<WK>(acc, k);
  uint64_t *o = out + (size_t)q * Wout * N + i;
#pragma unroll
  for (int w = 0; w < WK; w++)
    if (w < Wout) o[(size_t)w * N] = acc[w];
}

// Reconstruction plus addends, for the two client-side products:
//   he_enc_pk (he-encrypt.c:37-74):  c0 = smod(v pk.p0 + m + e0, q_L),  c1 = smod(v pk.p1 + e1, q_L)     H = 2
//   he_dec    (he-encrypt.c:107-123): m = smod(c1 sk + c0, q_l)                                           H = 1
// chat[inst][H][D][N] pre-weighted limbs of the product(s); addw_h: optional multiword addend of half h (instance stride
// given, Waddw words), adds_h: optional one-word signed addend [inst][N] (the sampled errors).  out[inst][H][Wout][N].
template <int WK>
__global__ void k_recon_add(const uint64_t *__restrict__ chat, unsigned D, uint64_t *__restrict__ out, int Wout, int k, Tables T,
                            unsigned npoly, unsigned H, const uint64_t *__restrict__ addw0, const uint64_t *__restrict__ addw1,
                            size_t addw_stride, int Waddw, const int64_t *__restrict__ adds0, const int64_t *__restrict__ adds1) {
  const unsigned N = T.N;
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (size_t)npoly * N) return;
  const unsigned q = gid / N, i = gid % N, inst = q / H, h = q % H;
  const uint64_t *limbs = chat + (size_t)q * D * N + i;
  uint64_t acc[WK];
  crt_low<WK, true>([&](unsigned d) { return limbs[(size_t)d * N]; }, D, T, acc);
  const uint64_t *aw = h ? addw1 : addw0;
  const int64_t *as = h ? adds1 : adds0;
  uint64_t carry = 0;
  if (aw) {
    const uint64_t *ap = aw + (size_t)inst * addw_stride + i;
    uint64_t fill = 0;
#pragma unroll
    for (int w = 0; w < WK; w++) {
      uint64_t a;
      if (w < Waddw) { a = ap[(size_t)w * N]; fill = (uint64_t)((int64_t)a >> 63); }
      else a = fill;
      const uint64_t s1 = acc[w] + a, c1 = s1 < a;
      const uint64_t s2 = s1 + carry, c2 = s2 < carry;
      acc[w] = s2;
      carry = c1 | c2;
    }
  }
  if (as) {
    const int64_t e = as[(size_t)inst * N + i];
    const uint64_t fill = (uint64_t)(e >> 63);
    carry = 0;
#pragma unroll
    for (int w = 0; w < WK; w++) {
      const uint64_t a = w == 0 ? (uint64_t)e : fill;
      const uint64_t s1 = acc[w] + a, c1 = s1 < a;
      const uint64_t s2 = s1 + carry, c2 = s2 < carry;
      acc[w] = s2;
      carry = c1 | c2;
    }
  }
  sext_bits<WK>(acc, k);
  uint64_t *o = out + (size_t)q * Wout * N + i;
#pragma unroll
  for (int w = 0; w < WK; w++)
    if (w < Wout) o[(size_t)w * N] = acc[w];
}

// =========================================================================================================
// K5 + K6: key-switch tail.  he_swk he-automorphism.c:68-76 / he_relin he-mult.c:68-77:
//   c_h = smod( rdiv( smod(reconstruct(chat_h), P q_l), P ) + add_h , q_l )
// With q_l = 2^k and P odd this equals  ((x - R) P^-1 + [R > floor(P/2)] + add_h) mod 2^k, sign-extended, where
// x is the centred reconstruction (only its low words are needed) and R = x mod P is rebuilt EXACTLY from the
// first `dim` limbs (P = p_0...p_{dim-1}); see DESIGN.md "key-switch tail" for the proof.
// add0/add1: optional polynomials (batch stride given) with the automorphism ginv applied on the fly.
// =========================================================================================================
// R = x mod P rebuilt exactly from the limbs of the primes of P (limbs[e * N], pre-weighted for a D-limb basis, any value
// below 2^64), and roundup = [R > floor(P/2)]  (mpi_rdiv, types.c:115-128)
template <int WPK>
__device__ __forceinline__ void ks_R_exact(const uint64_t *__restrict__ limbs, size_t lstride, unsigned D, const Tables &T,
                                           uint64_t (&R)[WPK], bool &roundup_out, const ulonglong2 *__restrict__ ztab = nullptr) {
#pragma unroll
  for (int w = 0; w < WPK; w++) R[w] = 0;
  uint64_t Flo = 0, Fhi = 0;
  const int WP = T.WP;
  for (unsigned d = 0; d < T.dim; d++) {
    const uint64_t p = T.primes[d].p, mu = T.primes[d].mu;
    const ulonglong2 kz = ztab ? ztab[d] : T.ks_z[(size_t)D * T.dim + d];     // limbs hold y_d: re-weight for P
    const uint64_t z = shoup(limbs[(size_t)d * lstride], kz.x, kz.y, p);
    uint64_t lo = z * mu, hi = __umul64hi(z, mu);
    Flo += lo;
    Fhi += hi + (Flo < lo);
    uint64_t carry = 0;
#pragma unroll
    for (int w = 0; w < WPK; w++) {
      if (w < WP) {
        const uint64_t pw = T.ks_phat[(size_t)d * WP + w];
        uint64_t l = z * pw, hh = __umul64hi(z, pw);
        uint64_t s = R[w] + l;
        hh += (s < l);
        uint64_t s2 = s + carry;
        hh += (s2 < carry);
        R[w] = s2;
        carry = hh;
      }
    }
  }
  {
    const uint64_t v = Fhi >> 56;     // floor estimate: true floor is v or v+1
    uint64_t borrow = 0;
#pragma unroll
    for (int w = 0; w < WPK; w++) {
      if (w < WP) {
        const uint64_t pw = T.ks_P[w];
        uint64_t l = v * pw, hh = __umul64hi(v, pw);
        uint64_t sub = l + borrow;
        hh += (sub < l);
        hh += (R[w] < sub);
        R[w] -= sub;
        borrow = hh;
      }
    }
    // if R >= P: R -= P
    bool ge = true;  // compare from the top word down
    {
      bool decided = false;
#pragma unroll
      for (int w = WPK - 1; w >= 0; w--) {
        if (w < WP && !decided) {
          const uint64_t pw = T.ks_P[w];
          if (R[w] != pw) { ge = R[w] > pw; decided = true; }
        }
      }
    }
    if (ge) {
      uint64_t bw = 0;
#pragma unroll
      for (int w = 0; w < WPK; w++) {
        if (w < WP) {
          const uint64_t pw = T.ks_P[w];
          uint64_t sub = pw + bw;
          uint64_t nb = (sub < pw) | (R[w] < sub);
          R[w] -= sub;
          bw = nb;
        }
      }
    }
  }
  bool roundup = false;  // R > floor(P/2)
  {
    bool decided = false;
#pragma unroll
    for (int w = WPK - 1; w >= 0; w--) {
      if (w < WP && !decided) {
        const uint64_t hw = T.ks_Ph[w];
        if (R[w] != hw) { roundup = R[w] > hw; decided = true; }
      }
    }
  }
  roundup_out = roundup;
}

// t = diff * P^-1 mod 2^(64 WK)
template <int WK>
__device__ __forceinline__ void mul_Pinv(const uint64_t (&diff)[WK], const Tables &T, uint64_t (&t)[WK]) {
#pragma unroll
  for (int w = 0; w < WK; w++) t[w] = 0;
#pragma unroll
  for (int a = 0; a < WK; a++) {
    uint64_t carry = 0;
#pragma unroll
    for (int c = 0; c + a < WK; c++) {
      const uint64_t pw = (c < (int)T.WC) ? T.ks_Pinv[c] : 0;
      uint64_t l = diff[a] * pw, hh = __umul64hi(diff[a], pw);
      uint64_t s = t[a + c] + l;
      hh += (s < l);
      uint64_t s2 = s + carry;
      hh += (s2 < carry);
      t[a + c] = s2;
      carry = hh;
    }
  }
}

// exact path for one coefficient: t = (x - R) P^-1 mod 2^(64 WK), roundup = [R > floor(P/2)]
template <int WK, int WPK>
__device__ __forceinline__ void ks_exact(const uint64_t *__restrict__ limbs, unsigned D, const Tables &T, uint64_t (&t)[WK],
                                      bool &roundup_out) {
  const unsigned N = T.N;
  uint64_t acc[WK];
  crt_low<WK, true>([&](unsigned d) { return limbs[(size_t)d * N]; }, D, T, acc);
  uint64_t R[WPK];
  ks_R_exact<WPK>(limbs, N, D, T, R, roundup_out);
  const int WP = T.WP;
  // ---- t = (x - R) * P^-1 mod 2^(64 WK) ----
  uint64_t diff[WK];
  {
    uint64_t bw = 0;
#pragma unroll
    for (int w = 0; w < WK; w++) {
      const uint64_t rw = (w < WPK && w < WP) ? R[w < WPK ? w : 0] : 0;
      uint64_t sub = rw + bw;
      uint64_t nb = (sub < rw) | (acc[w] < sub);
      diff[w] = acc[w] - sub;
      bw = nb;
    }
  }
  mul_Pinv<WK>(diff, T, t);
}

// sigma(add_h)[i] as WK sign-extended words (zero without an addend) and its sign under the automorphism.  Separate from
// ks_tail_apply so that a kernel can issue these gathered loads before its arithmetic instead of waiting for them at the end.
template <int WK>
__device__ __forceinline__ void ks_tail_load(const uint64_t *__restrict__ add, size_t add_bstride, int Wadd, unsigned ginv, unsigned N,
                                             unsigned badd, unsigned i, uint64_t (&addw)[WK], bool &negate) {
#pragma unroll
  for (int w = 0; w < WK; w++) addw[w] = 0;
  negate = false;
  if (add) {
    unsigned ip = (i * ginv) & (2 * N - 1);
    negate = ip >= N;
    const uint64_t *ap = add + (size_t)badd * add_bstride + (ip & (N - 1));
    uint64_t fill = 0;
#pragma unroll
    for (int w = 0; w < WK; w++) {
      if (w < Wadd) { addw[w] = ap[(size_t)w * N]; fill = (uint64_t)((int64_t)addw[w] >> 63); }
      else addw[w] = fill;
    }
  }
}

// t + roundup + (negate ? -addw : addw), centred mod 2^k, stored as Wout words
template <int WK>
__device__ __forceinline__ void ks_tail_apply(uint64_t (&t)[WK], bool roundup, const uint64_t (&addw)[WK], bool negate, unsigned N,
                                              unsigned b, unsigned h, unsigned i, uint64_t *__restrict__ out, int Wout, int k) {
  {
    // t += roundup + (negate ? -addw : addw)  ==  t + roundup + (addw ^ m) + (negate ? 1 : 0)
    const uint64_t m = negate ? ~0ULL : 0ULL;
    uint64_t carry = (roundup ? 1u : 0u);
    uint64_t carry2 = negate ? 1u : 0u;
#pragma unroll
    for (int w = 0; w < WK; w++) {
      uint64_t a = addw[w] ^ m;
      uint64_t s = t[w] + a;
      uint64_t c1 = (s < a);
      uint64_t s2 = s + carry;
      uint64_t c2 = (s2 < carry);
      uint64_t s3 = s2 + carry2;
      uint64_t c3 = (s3 < carry2);
      t[w] = s3;
      carry = c1 + c2;   // at most 1 in total with c3 handled separately
      carry2 = c3;
    }
  }
  sext_bits<WK>(t, k);
  uint64_t *o = out + ((size_t)b * 2 + h) * Wout * N + i;
#pragma unroll
  for (int w = 0; w < WK; w++)
    if (w < Wout) HB_ST(&o[(size_t)w * N], t[w]);
}

// t + roundup + sigma(add_h)[i], centred mod 2^k, stored as Wout words
template <int WK>
__device__ __forceinline__ void ks_tail(uint64_t (&t)[WK], bool roundup, const uint64_t *__restrict__ add, size_t add_bstride,
                                        int Wadd, unsigned ginv, unsigned N, unsigned b, unsigned badd, unsigned h, unsigned i,
                                        uint64_t *__restrict__ out, int Wout, int k) {
  uint64_t addw[WK];
  bool negate;
  ks_tail_load<WK>(add, add_bstride, Wadd, ginv, N, badd, i, addw, negate);
  ks_tail_apply<WK>(t, roundup, addw, negate, N, b, h, i, out, Wout, k);
}

// ---- fast path ------------------------------------------------------------------------------------------------
// Same integers as ks_exact for small rings (D <= 8 limbs, dim <= 4, q_l <= 2^128): all constants travel as a kernel
// parameter (constant bank, no table loads), the limb residues stay lazy (below 2p), and the two floors that ks_exact
// settles with multiword comparisons -- v = round(sum y_d / p_d) and R/P = frac(sum z_e / p_e) -- are read off 56-bit
// fixed-point sums whose error is below 2 D units.  When a sum lands within 64 units of a decision boundary (probability
// about 2^-48 per coefficient, or an exactly zero R) the thread takes ks_exact instead, so the result is always exact.
struct KsFast {
  uint64_t p[8], mu[8], phat[8][2], PD[2];                       // CRT over the D pre-weighted limbs (x mod 2^128)
  uint64_t kinv[4], kinvq[4], kphat[4][2], kP[2], Pinv[2];       // R = x mod P over the first dim limbs (kinv = ks_z: re-weight); P^-1 mod 2^128
  unsigned D, dim;
};

template <int WK>
__device__ __forceinline__ void mac_low(uint64_t (&acc)[WK], uint64_t y, const uint64_t (&w)[2]) {   // acc += y * w (low WK words)
  if (WK == 1) { acc[0] += y * w[0]; return; }
  const uint64_t lo = y * w[0];
  const uint64_t hi = __umul64hi(y, w[0]) + y * w[1];
  acc[0] += lo;
  acc[WK - 1] += hi + (acc[0] < lo);
}

// the whole exact tail for one coefficient, out of line so that its registers and arrays stay off the fast path
template <int WK, int WPK>
__device__ __noinline__ void ks_slow(const uint64_t *__restrict__ limbs, unsigned D, const Tables &T,
                                     const uint64_t *__restrict__ add, size_t add_bstride, int Wadd, unsigned ginv, unsigned b,
                                     unsigned badd, unsigned h, unsigned i, uint64_t *__restrict__ out, int Wout, int k) {
  uint64_t t[WK];
  bool roundup;
  ks_exact<WK, WPK>(limbs, D, T, t, roundup);
  ks_tail<WK>(t, roundup, add, add_bstride, Wadd, ginv, T.N, b, badd, h, i, out, Wout, k);
}

template <int WK, int WPK, int D, int DIM>
__global__ void __launch_bounds__(128) k_ks_finish_fast(const uint64_t *__restrict__ chat, const __grid_constant__ KsFast F,
                                                        const uint64_t *__restrict__ add0, const uint64_t *__restrict__ add1,
                                                        size_t add_bstride, int Wadd, const __grid_constant__ TailGroups TG,
                                                        uint64_t *__restrict__ out, int Wout, int k,
                                                        const __grid_constant__ Tables T, unsigned batch) {
  static_assert(WK <= 2 && D <= 8 && DIM <= 4 && DIM <= D, "fast key-switch tail: q_l <= 2^128, D <= 8, dim <= 4");
  const unsigned N = T.N;
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (size_t)batch * 2 * N) return;
  const unsigned i = gid % N, h = (gid / N) & 1, b = gid / (2 * (size_t)N);
  const unsigned grp = b / TG.gbatch, ginv = TG.ginv[grp], badd = TG.add_shared ? b - grp * TG.gbatch : b;
  const uint64_t *limbs = chat + ((size_t)b * 2 + h) * D * N + i;
  uint64_t a[D];
#pragma unroll
  for (int d = 0; d < D; d++) a[d] = HB_LD(&limbs[(size_t)d * N]);
  uint64_t acc[WK], Rl[WK];
#pragma unroll
  for (int w = 0; w < WK; w++) { acc[w] = 0; Rl[w] = 0; }
  uint64_t Fx = 0, Fr = 0;
#pragma unroll
  for (int d = 0; d < D; d++) {
    const uint64_t y = a[d];                                     // already a_d (P_D/p_d)^-1, below 2p
    Fx += (uint64_t)hi32(y) * hi32(F.mu[d]);
    mac_low<WK>(acc, y, F.phat[d]);
  }
#pragma unroll
  for (int e = 0; e < DIM; e++) {
    const Mod m = make_mod(F.p[e]);
    const uint64_t z = shoup_lazy(a[e], F.kinv[e], F.kinvq[e], m);
    Fr += (uint64_t)hi32(z) * hi32(F.mu[e]);
    mac_low<WK>(Rl, z, F.kphat[e]);
  }
  // The fractions come from the high halves only (one IMAD.WIDE per term): each term underestimates y mu / 2^64 by less
  // than 2^33 + 3 units, at most 12 terms -> below 2^37; a sum within MARGIN of a decision boundary takes the exact tail.
  const uint64_t ONE = 1ULL << 56, HALF = 1ULL << 55, MASK = ONE - 1, MARGIN = 1ULL << 38;
  const uint64_t fx = (Fx + HALF) & MASK, fr = Fr & MASK;
  const bool unsure = fx < MARGIN || fx > ONE - MARGIN || fr < MARGIN || fr > ONE - MARGIN ||
                      (fr > HALF - MARGIN && fr < HALF + MARGIN);
  const uint64_t *add = h ? add1 : add0;
  if (unsure) {
    ks_slow<WK, WPK>(limbs, D, T, add, add_bstride, Wadd, ginv, b, badd, h, i, out, Wout, k);
    return;
  }
  const uint64_t v = (Fx + HALF) >> 56, v2 = Fr >> 56;
  const bool roundup = fr > HALF;
  // t = (acc - v P_D) - (Rl - v2 P)  (mod 2^(64 WK)), every constant already times P^-1 (do_ks_finish)
  uint64_t t[WK];
  if (WK == 1) {
    t[0] = acc[0] - v * F.PD[0] - Rl[0] + v2 * F.kP[0];
  } else {
    uint64_t dlt[WK], sub[WK];
#pragma unroll
    for (int w = 0; w < WK; w++) { dlt[w] = 0; sub[w] = 0; }
    mac_low<WK>(dlt, v, F.PD);
    mac_low<WK>(sub, v2, F.kP);
    unsigned __int128 A = ((unsigned __int128)acc[WK - 1] << 64) | acc[0];
    A -= ((unsigned __int128)dlt[WK - 1] << 64) | dlt[0];
    A -= ((unsigned __int128)Rl[WK - 1] << 64) | Rl[0];
    A += ((unsigned __int128)sub[WK - 1] << 64) | sub[0];
    t[0] = (uint64_t)A;
    t[WK - 1] = (uint64_t)(A >> 64);
  }
  ks_tail<WK>(t, roundup, add, add_bstride, Wadd, ginv, N, b, badd, h, i, out, Wout, k);
}

// ---- summed giant steps -----------------------------------------------------------------------------------------
// Tail of the n2 giant-step key switches of he_gemv, their he_add chain and the closing he_rs (he-algo.c:80-90) in one pass:
//   out = smod( rdiv( smod( sum_g [ (x_g - R_g)/P + [R_g > P/2] + sigma_g(inner_g.c0) ], 2^k ), 2^shift ), 2^kout )
// X = sum_g x_g comes from ONE reconstruction: the limbs of the primes of P are the sums of the per-giant-step limbs
// (chatP, which also give every R_g exactly as in k_ks_finish_fast), the other limbs were summed in the transform domain by
// k_ks_limb_sum (chatX).  Fixed-point decisions with the same margins as the fast tail; a thread that lands within the margin
// of any boundary takes ks_sum_slow, which settles all of them with multiword arithmetic.  q_l <= 2^128.
struct SumGroups {
  unsigned ginv[HB_MAX_GROUPS];
  unsigned ng;
};

template <int WPK>
__device__ __noinline__ void ks_sum_slow(const uint64_t *__restrict__ limbsP, size_t gstride, const uint64_t *__restrict__ limbsX,
                                         unsigned D, unsigned ng, const Tables &T, uint64_t (&t)[2], unsigned &nround) {
  const unsigned N = T.N, dim = T.dim;
  uint64_t acc[2];
  crt_low<2, true>(
      [&](unsigned d) -> uint64_t {
        const Mod m = make_mod(T.primes[d].p);
        if (d >= dim) return canon(limbsX[(size_t)(d - dim) * N], m);
        uint64_t sum = 0;
        for (unsigned g = 0; g < ng; g++) sum = addmod(sum, canon(limbsP[(size_t)g * gstride + (size_t)d * N], m), m.p);
        return sum;
      },
      D, T, acc);
  nround = 0;
  const int WP = T.WP;
  for (unsigned g = 0; g < ng; g++) {
    uint64_t R[WPK];
    bool up;
    ks_R_exact<WPK>(limbsP + (size_t)g * gstride, N, D, T, R, up);
    nround += up ? 1u : 0u;
    uint64_t bw = 0;
#pragma unroll
    for (int w = 0; w < 2; w++) {
      const uint64_t rw = (w < WPK && w < WP) ? R[w] : 0;
      const uint64_t sub = rw + bw;
      const uint64_t nb = (sub < rw) | (acc[w] < sub);
      acc[w] -= sub;
      bw = nb;
    }
  }
  mul_Pinv<2>(acc, T, t);
}

template <int WPK, int D, int DIM>
__global__ void __launch_bounds__(128)
k_ks_finish_sum(const uint64_t *__restrict__ chatP, size_t chatP_gstride, const uint64_t *__restrict__ chatX, const __grid_constant__ KsFast F,
                const uint64_t *__restrict__ add0, size_t add_gstride, size_t add_bstride, int Wadd, const __grid_constant__ SumGroups SG,
                uint64_t *__restrict__ out, int Wout, int k, int shift, int kout, const __grid_constant__ Tables T, unsigned batch) {
  static_assert(D <= 8 && DIM <= 4 && DIM < D, "summed key-switch tail: D <= 8, dim <= 4");
  typedef unsigned __int128 u128;
  const unsigned N = T.N;
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (size_t)batch * 2 * N) return;
  const unsigned i = gid % N, h = (gid / N) & 1, b = gid / (2 * (size_t)N);
  const unsigned ng = SG.ng;
  const uint64_t *limbsP = chatP + ((size_t)b * 2 + h) * DIM * N + i;          // + g * chatP_gstride + e * N
  const uint64_t *limbsX = chatX + ((size_t)b * 2 + h) * (D - DIM) * N + i;    // + (d - DIM) * N
  const uint64_t ONE = 1ULL << 56, HALF = 1ULL << 55, MASK = ONE - 1, MARGIN = 1ULL << 38;
  uint64_t acc[2] = {0, 0}, Fx = 0;
#pragma unroll
  for (int d = DIM; d < D; d++) {
    const uint64_t y = HB_LD(&limbsX[(size_t)(d - DIM) * N]);                   // below 2p
    Fx += (uint64_t)hi32(y) * hi32(F.mu[d]);
    mac_low<2>(acc, y, F.phat[d]);
  }
  uint64_t ysum[DIM];
#pragma unroll
  for (int e = 0; e < DIM; e++) ysum[e] = 0;
  u128 Rsum = 0;
  unsigned nround = 0;
  bool unsure = false;
#pragma unroll 1
  for (unsigned g = 0; g < ng; g++) {
    uint64_t Rl[2] = {0, 0}, Fr = 0;
#pragma unroll
    for (int e = 0; e < DIM; e++) {
      const uint64_t a = HB_LD(&limbsP[(size_t)g * chatP_gstride + (size_t)e * N]);     // below 2p
      ysum[e] += a;                                                             // at most 8 giant steps: below 16p < 2^64
      const Mod m = make_mod(F.p[e]);
      const uint64_t z = shoup_lazy(a, F.kinv[e], F.kinvq[e], m);
      Fr += (uint64_t)hi32(z) * hi32(F.mu[e]);
      mac_low<2>(Rl, z, F.kphat[e]);
    }
    const uint64_t fr = Fr & MASK, v2 = Fr >> 56;
    unsure |= fr < MARGIN || fr > ONE - MARGIN || (fr > HALF - MARGIN && fr < HALF + MARGIN);
    nround += fr > HALF ? 1u : 0u;
    uint64_t sub[2] = {0, 0};
    mac_low<2>(sub, v2, F.kP);
    Rsum += (((u128)Rl[1] << 64) | Rl[0]) - (((u128)sub[1] << 64) | sub[0]);
  }
#pragma unroll
  for (int e = 0; e < DIM; e++) {
    Fx += (uint64_t)hi32(ysum[e]) * hi32(F.mu[e]);
    mac_low<2>(acc, ysum[e], F.phat[e]);
  }
  const uint64_t fx = (Fx + HALF) & MASK;
  unsure |= fx < MARGIN || fx > ONE - MARGIN;
  u128 tt;
  if (unsure) {
    uint64_t t[2];
    ks_sum_slow<WPK>(limbsP, chatP_gstride, limbsX, D, ng, T, t, nround);
    tt = ((u128)t[1] << 64) | t[0];
  } else {
    const uint64_t v = (Fx + HALF) >> 56;
    uint64_t dlt[2] = {0, 0};
    mac_low<2>(dlt, v, F.PD);
    tt = ((((u128)acc[1]) << 64) | acc[0]) - ((((u128)dlt[1]) << 64) | dlt[0]) - Rsum;      // every constant already times P^-1
  }
  tt += nround;
  if (h == 0 && add0) {                                                          // sum_g sigma_g(inner_g.c0)
#pragma unroll 1
    for (unsigned g = 0; g < ng; g++) {
      const unsigned ip = (i * SG.ginv[g]) & (2 * N - 1);
      const uint64_t *ap = add0 + (size_t)g * add_gstride + (size_t)b * add_bstride + (ip & (N - 1));
      const uint64_t lo = HB_LD(&ap[0]);
      const uint64_t hi = Wadd > 1 ? HB_LD(&ap[N]) : (uint64_t)((int64_t)lo >> 63);
      const u128 a = ((u128)hi << 64) | lo;
      tt = (ip >= N) ? tt - a : tt + a;
    }
  }
  uint64_t x[2] = {(uint64_t)tt, (uint64_t)(tt >> 64)};
  sext_bits<2>(x, k);                                                            // the centred sum of the he_add chain
  if (shift > 0 || kout != k) {                                                  // he_rs: rounded shift, ties toward -inf (types.c:115-128)
    const __int128 vs = (__int128)((((u128)x[1]) << 64) | x[0]);
    __int128 q = vs;
    if (shift > 0) {
      const u128 rem = (u128)vs & ((((u128)1) << shift) - 1), half = ((u128)1) << (shift - 1);
      q = (vs >> shift) + (rem > half ? 1 : 0);
    }
    x[0] = (uint64_t)(u128)q;
    x[1] = (uint64_t)((u128)q >> 64);
    sext_bits<2>(x, kout);
  }
  uint64_t *o = out + ((size_t)b * 2 + h) * Wout * N + i;
  HB_ST(&o[0], x[0]);
  if (Wout > 1) HB_ST(&o[N], x[1]);
}

// ---- hybrid basis (hb_dev32.cuh "Key switching on a HYBRID basis"): limbs of P on 59-bit primes, the rest on 30-bit primes ----
// Same tail as k_ks_finish_fast / k_ks_finish_sum with the CRT over M = P * Q_m.  The host guarantees |x| < M / 8 from the measured
// key bound, so the rounding v = round(sum y/prime) sits at least 3/8 away from a boundary and the fixed-point sum (error below
// 2^-18) always decides it; only the decisions about R = x mod P can land inside a margin, and those go to the exact path.
#define HB_HYB_MAX32 12
struct KsHyb {
  uint64_t p[4], mu[4], phat[4][2];                       // limbs of P: modulus, floor(2^120/p), M/p_e mod 2^128 (times P^-1)
  uint32_t q[HB_HYB_MAX32], mu32[HB_HYB_MAX32];           // 30-bit limbs: modulus, floor(2^61/q)
  float rq[HB_HYB_MAX32], rp[4];                          // 1/q_i and 2^32/p_e: the rounding v = round(sum y/prime) is decided in
                                                          // FP32 (|x| < M/8 keeps it 3/8 from a boundary; eleven terms err by < 1e-5),
                                                          // on the FP32 pipe instead of eleven wide integer multiplies
  uint64_t qhat[HB_HYB_MAX32][2];                         // M/q_i mod 2^128 (times P^-1)
  uint64_t M[2];                                          // M mod 2^128 (times P^-1)
  uint64_t kinv[4], kinvq[4], kphat[4][2], kP[2];         // y_e -> z_e re-weighting, P/p_e and P mod 2^128 (times P^-1)
  const ulonglong2 *ztab;                                 // DEVICE [dim]: kinv as Shoup pairs for the exact path
  unsigned dim, m;
};

__device__ __forceinline__ void mac_low32(uint64_t (&acc)[2], uint32_t y, const uint64_t (&w)[2]) {   // acc += y * w (mod 2^128)
  const uint64_t a = (uint64_t)y * lo32(w[0]);                       // bits 0..63
  const uint64_t b = (uint64_t)y * hi32(w[0]);                       // bits 32..95
  const uint64_t lo = a + (b << 32);
  const uint64_t hi = (b >> 32) + (lo < a) + (uint64_t)y * w[1];     // low 64 bits of y * w[1] suffice
  acc[0] += lo;
  acc[1] += hi + (acc[0] < lo);
}

// 128-bit accumulators (the compiler's carry chains: IADD3.X, no compare / select pairs)
typedef unsigned __int128 hb_u128;
__device__ __forceinline__ hb_u128 u128_of(const uint64_t (&w)[2]) { return (((hb_u128)w[1]) << 64) | w[0]; }
__device__ __forceinline__ void mac128(hb_u128 &acc, uint64_t y, const uint64_t (&w)[2]) {        // acc += y * w (mod 2^128)
  acc += (hb_u128)y * w[0] + (((hb_u128)(y * w[1])) << 64);
}
__device__ __forceinline__ void mac128_32(hb_u128 &acc, uint32_t y, const uint64_t (&w)[2]) {     // the same for a 32-bit y
  const uint64_t a = (uint64_t)y * lo32(w[0]), b = (uint64_t)y * hi32(w[0]);
  acc += (hb_u128)a + (((hb_u128)b) << 32) + (((hb_u128)((uint64_t)y * lo32(w[1]) + (((uint64_t)(y * hi32(w[1]))) << 32))) << 64);
}

// x part of the hybrid tail for one coefficient: acc += sum y_i (M/q_i) (mod 2^128), fx += sum y_i / q_i (FP32)
template <int DIM, int M32>
__device__ __forceinline__ void hyb_x32(const uint32_t *__restrict__ limbs32, unsigned N, const KsHyb &F, hb_u128 &acc, float &fx) {
  uint32_t y[M32];
#pragma unroll
  for (int i = 0; i < M32; i++) y[i] = __ldcs(&limbs32[(size_t)i * N]);
#pragma unroll
  for (int i = 0; i < M32; i++) {
    fx = fmaf(__uint2float_rn(y[i]), F.rq[i], fx);
    mac128_32(acc, y[i], F.qhat[i]);
  }
}

template <int WPK>
__device__ __noinline__ void hyb_R_slow(const uint64_t *__restrict__ limbs59, unsigned N, const Tables &T, const ulonglong2 *__restrict__ ztab,
                                        uint64_t (&RP)[2], bool &roundup) {
  uint64_t R[WPK];
  ks_R_exact<WPK>(limbs59, N, 0u, T, R, roundup, ztab);
  uint64_t r2[2] = {R[0], (WPK > 1 && T.WP > 1) ? R[1 < WPK ? 1 : 0] : 0};
  mul_Pinv<2>(r2, T, RP);                                   // R P^-1 mod 2^128
}

// nine CTAs of 128 threads per SM (56 registers, nothing spilled): 6.61 -> 6.49 ms per step [variants_tails.log]
#ifndef HB_FH_MINB
#define HB_FH_MINB 9
#endif
template <int WPK, int DIM, int M32>
__global__ void __launch_bounds__(128, HB_FH_MINB)
k_ks_finish_hyb(const uint64_t *__restrict__ chat59, const uint32_t *__restrict__ chat32, const __grid_constant__ KsHyb F,
                const uint64_t *__restrict__ add0, const uint64_t *__restrict__ add1, size_t add_bstride, int Wadd,
                const __grid_constant__ TailGroups TG, uint64_t *__restrict__ out, int Wout, int k, const __grid_constant__ Tables T, unsigned batch) {
  typedef unsigned __int128 u128;
  const unsigned N = T.N;
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (size_t)batch * 2 * N) return;
  const unsigned i = gid % N, h = (gid / N) & 1, b = gid / (2 * (size_t)N);
  const unsigned grp = b / TG.gbatch, ginv = TG.ginv[grp], badd = TG.add_shared ? b - grp * TG.gbatch : b;
  const uint64_t *limbs59 = chat59 + ((size_t)b * 2 + h) * DIM * N + i;
  const uint32_t *limbs32 = chat32 + ((size_t)b * 2 + h) * M32 * N + i;
  const uint64_t ONE = 1ULL << 56, HALF = 1ULL << 55, MASK = ONE - 1, MARGIN = 1ULL << 38;
  u128 acc = 0, Rl = 0;
  uint64_t Fr = 0;
  float fx = 0.5f;
  // every load of the coefficient is issued here: the gathered addend (needed last), the 59-bit limbs, then the 30-bit limbs
  uint64_t addw[2];
  bool negate;
  ks_tail_load<2>(h ? add1 : add0, add_bstride, Wadd, ginv, N, badd, i, addw, negate);
  uint64_t y59[DIM];
#pragma unroll
  for (int e = 0; e < DIM; e++) y59[e] = HB_LD(&limbs59[(size_t)e * N]);
  hyb_x32<DIM, M32>(limbs32, N, F, acc, fx);
#pragma unroll
  for (int e = 0; e < DIM; e++) {
    const uint64_t y = y59[e];                                           // below 2p
    fx = fmaf(__uint2float_rn(hi32(y)), F.rp[e], fx);
    mac128(acc, y, F.phat[e]);
    const Mod m = make_mod(F.p[e]);
    const uint64_t z = shoup_lazy(y, F.kinv[e], F.kinvq[e], m);
    Fr += (uint64_t)hi32(z) * hi32(F.mu[e]);
    mac128(Rl, z, F.kphat[e]);
  }
  const uint32_t v = (uint32_t)fx;                                       // round(sum y/prime): at most 2 (dim + m)
  u128 tt = acc - u128_of(F.M) * v;
  const uint64_t fr = Fr & MASK;
  bool roundup;
  if (fr < MARGIN || fr > ONE - MARGIN || (fr > HALF - MARGIN && fr < HALF + MARGIN)) {
    uint64_t RP[2];
    hyb_R_slow<WPK>(limbs59, N, T, F.ztab, RP, roundup);
    tt -= u128_of(RP);
  } else {
    const uint32_t v2 = (uint32_t)(Fr >> 56);
    roundup = fr > HALF;
    tt -= Rl - u128_of(F.kP) * v2;
  }
  uint64_t t[2] = {(uint64_t)tt, (uint64_t)(tt >> 64)};
  ks_tail_apply<2>(t, roundup, addw, negate, N, b, h, i, out, Wout, k);
}

// summed giant steps on the hybrid basis: chatP[g][b][h][DIM][N] (u64, limbs of P per giant step), chatX32[b][h][M32][N] (u32, summed)
// NGT > 0: the number of giant steps as a compile-time constant; the limbs of P of every giant step and the gathered addends
// are then read before the first product (with a run-time count each giant step waits for its own loads in turn).
// six CTAs of 128 threads per SM (80 registers, a few spilled): 3.10 -> 2.91 ms per step against the compiler's own choice
// of 102 [B200, 4096 instances, profiles/r02/variants_tails.log]
#ifndef HB_FSH_MINB
#define HB_FSH_MINB 6
#endif
template <int WPK, int DIM, int M32, int NGT>
__global__ void __launch_bounds__(128, HB_FSH_MINB)
k_ks_finish_sum_hyb(const uint64_t *__restrict__ chatP, size_t chatP_gstride, const uint32_t *__restrict__ chatX32, const __grid_constant__ KsHyb F,
                    const uint64_t *__restrict__ add0, size_t add_gstride, size_t add_bstride, int Wadd, const __grid_constant__ SumGroups SG,
                    uint64_t *__restrict__ out, int Wout, int k, int shift, int kout, const __grid_constant__ Tables T, unsigned batch) {
  typedef unsigned __int128 u128;
  const unsigned N = T.N;
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (size_t)batch * 2 * N) return;
  const unsigned i = gid % N, h = (gid / N) & 1, b = gid / (2 * (size_t)N);
  const unsigned ng = NGT > 0 ? (unsigned)NGT : SG.ng;
  constexpr int NB = NGT > 0 ? NGT : 1, UNR = NGT > 0 ? NGT : 1;
  const uint64_t *limbsP = chatP + ((size_t)b * 2 + h) * DIM * N + i;
  const uint32_t *limbs32 = chatX32 + ((size_t)b * 2 + h) * M32 * N + i;
  const uint64_t ONE = 1ULL << 56, HALF = 1ULL << 55, MASK = ONE - 1, MARGIN = 1ULL << 38;
  u128 acc = 0;
  float fx = 0.5f;
  const bool with_add = h == 0 && add0;
  uint64_t pre[NB][DIM], addlo[NB], addhi[NB];
  bool addneg[NB];
  if constexpr (NGT > 0) {
#pragma unroll
    for (int g = 0; g < NGT; g++) {
      addlo[g] = addhi[g] = 0;
      addneg[g] = false;
      if (with_add) {
        const unsigned ip = (i * SG.ginv[g]) & (2 * N - 1);
        const uint64_t *ap = add0 + (size_t)g * add_gstride + (size_t)b * add_bstride + (ip & (N - 1));
        addlo[g] = HB_LD(&ap[0]);
        addhi[g] = Wadd > 1 ? HB_LD(&ap[N]) : 0;
        addneg[g] = ip >= N;
      }
    }
#pragma unroll
    for (int g = 0; g < NGT; g++)
#pragma unroll
      for (int e = 0; e < DIM; e++) pre[g][e] = HB_LD(&limbsP[(size_t)g * chatP_gstride + (size_t)e * N]);
  }
  hyb_x32<DIM, M32>(limbs32, N, F, acc, fx);
  uint64_t ysum[DIM];
#pragma unroll
  for (int e = 0; e < DIM; e++) ysum[e] = 0;
  u128 Rsum = 0;
  unsigned nround = 0;
#pragma unroll UNR
  for (unsigned g = 0; g < ng; g++) {
    u128 Rl = 0;
    uint64_t Fr = 0;
    const uint64_t *lg = limbsP + (size_t)g * chatP_gstride;
#pragma unroll
    for (int e = 0; e < DIM; e++) {
      const uint64_t a = NGT > 0 ? pre[NGT > 0 ? g : 0][e] : HB_LD(&lg[(size_t)e * N]);       // below 2p
      ysum[e] += a;                                                      // at most 8 giant steps: below 16p < 2^64
      const Mod m = make_mod(F.p[e]);
      const uint64_t z = shoup_lazy(a, F.kinv[e], F.kinvq[e], m);
      Fr += (uint64_t)hi32(z) * hi32(F.mu[e]);
      mac128(Rl, z, F.kphat[e]);
    }
    const uint64_t fr = Fr & MASK;
    if (fr < MARGIN || fr > ONE - MARGIN || (fr > HALF - MARGIN && fr < HALF + MARGIN)) {
      uint64_t RP[2];
      bool up;
      hyb_R_slow<WPK>(lg, N, T, F.ztab, RP, up);
      Rsum += u128_of(RP);
      nround += up ? 1u : 0u;
    } else {
      const uint32_t v2 = (uint32_t)(Fr >> 56);
      nround += fr > HALF ? 1u : 0u;
      Rsum += Rl - u128_of(F.kP) * v2;
    }
  }
#pragma unroll
  for (int e = 0; e < DIM; e++) {
    fx = fmaf(__uint2float_rn(hi32(ysum[e])), F.rp[e], fx);
    mac128(acc, ysum[e], F.phat[e]);
  }
  const uint32_t v = (uint32_t)fx;
  u128 tt = acc - u128_of(F.M) * v - Rsum + nround;
  if (with_add) {                                                        // sum_g sigma_g(inner_g.c0)
    if constexpr (NGT > 0) {
#pragma unroll
      for (int g = 0; g < NGT; g++) {
        const uint64_t hi = Wadd > 1 ? addhi[g] : (uint64_t)((int64_t)addlo[g] >> 63);
        const u128 a = ((u128)hi << 64) | addlo[g];
        tt = addneg[g] ? tt - a : tt + a;
      }
    } else {
#pragma unroll 1
      for (unsigned g = 0; g < ng; g++) {
        const unsigned ip = (i * SG.ginv[g]) & (2 * N - 1);
        const uint64_t *ap = add0 + (size_t)g * add_gstride + (size_t)b * add_bstride + (ip & (N - 1));
        const uint64_t lo = HB_LD(&ap[0]);
        const uint64_t hi = Wadd > 1 ? HB_LD(&ap[N]) : (uint64_t)((int64_t)lo >> 63);
        const u128 a = ((u128)hi << 64) | lo;
        tt = (ip >= N) ? tt - a : tt + a;
      }
    }
  }
  uint64_t x[2] = {(uint64_t)tt, (uint64_t)(tt >> 64)};
  sext_bits<2>(x, k);
  if (shift > 0 || kout != k) {                                          // he_rs (types.c:115-128)
    const __int128 vs = (__int128)((((u128)x[1]) << 64) | x[0]);
    __int128 q = vs;
    if (shift > 0) {
      const u128 rem = (u128)vs & ((((u128)1) << shift) - 1), half = ((u128)1) << (shift - 1);
      q = (vs >> shift) + (rem > half ? 1 : 0);
    }
    x[0] = (uint64_t)(u128)q;
    x[1] = (uint64_t)((u128)q >> 64);
    sext_bits<2>(x, kout);
  }
  uint64_t *o = out + ((size_t)b * 2 + h) * Wout * N + i;
  HB_ST(&o[0], x[0]);
  if (Wout > 1) HB_ST(&o[N], x[1]);
}

template <int WK, int WPK>
__global__ void k_ks_finish(const uint64_t *__restrict__ chat, unsigned D, const uint64_t *__restrict__ add0,
                            const uint64_t *__restrict__ add1, size_t add_bstride, int Wadd, const __grid_constant__ TailGroups TG,
                            uint64_t *__restrict__ out, int Wout, int k, Tables T, unsigned batch) {
  const unsigned N = T.N;
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (size_t)batch * 2 * N) return;
  const unsigned i = gid % N, h = (gid / N) & 1, b = gid / (2 * (size_t)N);
  const unsigned grp = b / TG.gbatch, ginv = TG.ginv[grp], badd = TG.add_shared ? b - grp * TG.gbatch : b;
  const uint64_t *limbs = chat + ((size_t)b * 2 + h) * D * N + i;
  uint64_t t[WK];
  bool roundup;
  ks_exact<WK, WPK>(limbs, D, T, t, roundup);
  ks_tail<WK>(t, roundup, h ? add1 : add0, add_bstride, Wadd, ginv, N, b, badd, h, i, out, Wout, k);
}

// =========================================================================================================
// K7: coefficient-wise ciphertext arithmetic (he-add.c:32-142, he-rescale.c:33-70), q_l = 2^k.
//   op 0: a + b   op 1: a - b   op 2: -a      then smod(., 2^k) = sign-extension from bit k-1.
// One thread per (polynomial, coefficient); W words each.
// =========================================================================================================
// a, b and out may alias (he_neg works in place, he_add(s, s, x) he-algo.c:78): no __restrict__; every thread reads its own
// words before it writes them.
__global__ void k_ct_addsub(const uint64_t *a, const uint64_t *b, uint64_t *out,
                            int W, int k, int op, unsigned N, size_t npoly) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= npoly * N) return;
  const size_t q = gid / N, i = gid % N;
  const uint64_t *pa = a + q * W * N + i;
  const uint64_t *pb = b ? b + q * W * N + i : nullptr;
  uint64_t *po = out + q * W * N + i;
  uint64_t carry = (op == 0) ? 0 : 1;      // two's complement: x - y = x + ~y + 1 ; -x = ~x + 1
  uint64_t fill = 0;
  bool done = false;
  for (int w = 0; w < W; w++) {
    uint64_t r;
    if (done) r = fill;
    else {
      uint64_t x = pa[(size_t)w * N], y = pb ? pb[(size_t)w * N] : 0;
      uint64_t lhs = (op == 2) ? 0 : x;
      uint64_t rhs = (op == 0) ? y : (op == 1 ? ~y : ~x);
      uint64_t s = lhs + rhs;
      uint64_t c1 = s < rhs;
      uint64_t s2 = s + carry;
      uint64_t c2 = s2 < carry;
      carry = c1 | c2;
      r = s2;
      int kk = k - 64 * w;
      if (kk <= 64) {
        if (kk < 64) r = (uint64_t)(((int64_t)(r << (64 - kk))) >> (64 - kk));
        fill = (uint64_t)((int64_t)r >> 63);
        done = true;
      }
    }
    po[(size_t)w * N] = r;
  }
}

// k_ct_sum followed by he_rs (he-rescale.c:33-54) in one pass: the centred sum stays in registers, then the rounded shift by
// `shift` bits and the centring mod 2^kout -- the end of he_gemv (he-algo.c:81-90).  Win <= 16 words.
__global__ void k_ct_sum_rs(const uint64_t *__restrict__ a, size_t gstride, unsigned G, uint64_t *__restrict__ out, int Win, int k,
                            int Wout, int shift, int kout, unsigned N, size_t npoly) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= npoly * N) return;
  const size_t q = gid / N, i = gid % N;
  const uint64_t *pa = a + q * Win * N + i;
  uint64_t s[16];
  {
    uint64_t fill = 0;
    unsigned carry = 0;
    bool done = false;
    for (int w = 0; w < Win; w++) {
      uint64_t r;
      if (done) r = fill;
      else {
        unsigned __int128 t = carry;
        for (unsigned g = 0; g < G; g++) t += pa[(size_t)g * gstride + (size_t)w * N];
        r = (uint64_t)t;
        carry = (unsigned)(t >> 64);
        const int kk = k - 64 * w;
        if (kk <= 64) {
          if (kk < 64) r = (uint64_t)(((int64_t)(r << (64 - kk))) >> (64 - kk));
          fill = (uint64_t)((int64_t)r >> 63);
          done = true;
        }
      }
      s[w] = r;
    }
  }
  const uint64_t sign = (uint64_t)((int64_t)s[Win - 1] >> 63);
  auto word = [&](int w) -> uint64_t { return w < Win ? s[w] : sign; };
  uint64_t roundup = 0;                          // round up iff (x mod 2^shift) > 2^(shift-1): ties toward -infinity (types.c:123)
  if (shift > 0) {
    const int hb = shift - 1, hw = hb / 64, hbit = hb % 64;
    const uint64_t wv = word(hw);
    if ((wv >> hbit) & 1) {
      bool lower = (hbit > 0) && ((wv & ((1ULL << hbit) - 1)) != 0);
      for (int w = 0; w < hw && !lower; w++) lower = word(w) != 0;
      roundup = lower ? 1 : 0;
    }
  }
  const int ws = shift / 64, bs = shift % 64;
  uint64_t *po = out + q * Wout * N + i;
  uint64_t carry = roundup, fill = 0;
  bool done = false;
  for (int w = 0; w < Wout; w++) {
    uint64_t r;
    if (done) r = fill;
    else {
      const uint64_t lo = word(w + ws), hi = word(w + ws + 1);
      const uint64_t v = bs ? ((lo >> bs) | (hi << (64 - bs))) : lo;
      const uint64_t t = v + carry;
      carry = (t < carry);
      r = t;
      const int kk = kout - 64 * w;
      if (kk <= 64) {
        if (kk < 64) r = (uint64_t)(((int64_t)(r << (64 - kk))) >> (64 - kk));
        fill = (uint64_t)((int64_t)r >> 63);
        done = true;
      }
    }
    po[(size_t)w * N] = r;
  }
}

// The last four calls of ctr_hempc (src/hempc.c:261-266) in one pass: he_add, he_neg, he_moddown of a copy of u, he_add.
// Every step is arithmetic mod 2^kout followed by centring, so  out = smod(u - a - b, 2^kout)  on the low words of u.
__global__ void k_ct_combine(const uint64_t *__restrict__ u, int Wu, const uint64_t *__restrict__ a, const uint64_t *__restrict__ b,
                             uint64_t *__restrict__ out, int Wout, int kout, unsigned N, size_t npoly) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= npoly * N) return;
  const size_t q = gid / N, i = gid % N;
  const uint64_t *pu = u + q * Wu * N + i, *pa = a + q * Wout * N + i, *pb = b + q * Wout * N + i;
  uint64_t *po = out + q * Wout * N + i;
  uint64_t br = 0, fill = 0;
  bool done = false;
  for (int w = 0; w < Wout; w++) {
    uint64_t r;
    if (done) r = fill;
    else {
      const unsigned __int128 sub = (unsigned __int128)pa[(size_t)w * N] + pb[(size_t)w * N] + br;     // below 2^65 + small
      const uint64_t x = pu[(size_t)w * N], y = (uint64_t)sub;
      r = x - y;
      br = (uint64_t)(sub >> 64) + (x < y);
      const int kk = kout - 64 * w;
      if (kk <= 64) {
        if (kk < 64) r = (uint64_t)(((int64_t)(r << (64 - kk))) >> (64 - kk));
        fill = (uint64_t)((int64_t)r >> 63);
        done = true;
      }
    }
    po[(size_t)w * N] = r;
  }
}

// he_addpt / he_subpt (he-add.c:80-123): c0 +- pt, c1 only re-centred; pt is one polynomial [WC][N] shared by the batch.
__global__ void k_ct_addpt(const uint64_t *a, const uint64_t *__restrict__ pt, uint64_t *out, int W, int k,     // a, out may alias
                           int op, unsigned N, size_t npoly) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= npoly * N) return;
  const size_t q = gid / N, i = gid % N;
  const uint64_t *pa = a + q * W * N + i;
  const uint64_t *pb = (q & 1) ? nullptr : pt + i;
  uint64_t *po = out + q * W * N + i;
  uint64_t carry = op;                      // x - y = x + ~y + 1
  uint64_t fill = 0;
  bool done = false;
  for (int w = 0; w < W; w++) {
    uint64_t r;
    if (done) r = fill;
    else {
      const uint64_t x = pa[(size_t)w * N], y = pb ? pb[(size_t)w * N] : 0;
      const uint64_t rhs = op ? ~y : y;
      const uint64_t s = x + rhs;
      const uint64_t c1 = s < rhs;
      const uint64_t s2 = s + carry;
      const uint64_t c2 = s2 < carry;
      carry = c1 | c2;
      r = s2;
      const int kk = k - 64 * w;
      if (kk <= 64) {
        if (kk < 64) r = (uint64_t)(((int64_t)(r << (64 - kk))) >> (64 - kk));
        fill = (uint64_t)((int64_t)r >> 63);
        done = true;
      }
    }
    po[(size_t)w * N] = r;
  }
}

// out = smod(sum_{g < G} a[g], 2^k): the he_add chain over the giant steps of he_gemv (he-algo.c:81-84).  Each he_add
// centres its result; every representative is congruent mod 2^k, so one centring of the full sum gives the same words.
__global__ void k_ct_sum(const uint64_t *__restrict__ a, size_t gstride, unsigned G, uint64_t *__restrict__ out, int W, int k,
                         unsigned N, size_t npoly) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= npoly * N) return;
  const size_t q = gid / N, i = gid % N;
  const uint64_t *pa = a + q * W * N + i;
  uint64_t *po = out + q * W * N + i;
  uint64_t fill = 0;
  unsigned carry = 0;           // at most G - 1 + previous carry: fits easily
  bool done = false;
  for (int w = 0; w < W; w++) {
    uint64_t r;
    if (done) r = fill;
    else {
      unsigned __int128 s = carry;
      for (unsigned g = 0; g < G; g++) s += pa[(size_t)g * gstride + (size_t)w * N];
      r = (uint64_t)s;
      carry = (unsigned)(s >> 64);
      const int kk = k - 64 * w;
      if (kk <= 64) {
        if (kk < 64) r = (uint64_t)(((int64_t)(r << (64 - kk))) >> (64 - kk));
        fill = (uint64_t)((int64_t)r >> 63);
        done = true;
      }
    }
    po[(size_t)w * N] = r;
  }
}

// he_rs (shift = log2 Delta: rounded division, ties toward -inf on the remainder test r > Delta/2, types.c:115-128)
// and he_moddown (shift = 0), both followed by smod(., 2^kout).  Win words in, Wout words out (out of place).
__global__ void k_ct_rescale(const uint64_t *__restrict__ a, uint64_t *__restrict__ out, int Win, int Wout, int shift,
                             int kout, unsigned N, size_t npoly) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= npoly * N) return;
  const size_t q = gid / N, i = gid % N;
  const uint64_t *pa = a + q * Win * N + i;
  uint64_t *po = out + q * Wout * N + i;
  const uint64_t sign = (uint64_t)((int64_t)pa[(size_t)(Win - 1) * N] >> 63);
  auto word = [&](int w) -> uint64_t { return w < Win ? pa[(size_t)w * N] : sign; };
  // round up iff (a mod 2^shift) > 2^(shift-1)
  uint64_t roundup = 0;
  if (shift > 0) {
    const int hb = shift - 1;                   // bit index of the half
    const int hw = hb / 64, hbit = hb % 64;
    const uint64_t wv = word(hw);
    if ((wv >> hbit) & 1) {
      bool lower = (hbit > 0) && ((wv & ((1ULL << hbit) - 1)) != 0);
      for (int w = 0; w < hw && !lower; w++) lower = word(w) != 0;
      roundup = lower ? 1 : 0;
    }
  }
  const int ws = shift / 64, bs = shift % 64;
  uint64_t carry = roundup, fill = 0;
  bool done = false;
  for (int w = 0; w < Wout; w++) {
    uint64_t r;
    if (done) r = fill;
    else {
      uint64_t lo = word(w + ws), hi = word(w + ws + 1);
      uint64_t v = bs ? ((lo >> bs) | (hi << (64 - bs))) : lo;
      uint64_t s = v + carry;
      carry = (s < carry);
      r = s;
      int kk = kout - 64 * w;
      if (kk <= 64) {
        if (kk < 64) r = (uint64_t)(((int64_t)(r << (64 - kk))) >> (64 - kk));
        fill = (uint64_t)((int64_t)r >> 63);
        done = true;
      }
    }
    po[(size_t)w * N] = r;
  }
}

// poly_rot / poly_conj as a stand-alone permutation (poly.c:263-283) -- used by tests and he_rot's add path.
__global__ void k_poly_automorph(const uint64_t *__restrict__ a, uint64_t *__restrict__ out, int W, unsigned ginv,
                                 unsigned N, size_t npoly) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= npoly * N) return;
  const size_t q = gid / N;
  const unsigned i = gid % N;
  const unsigned ip = (i * ginv) & (2 * N - 1);
  const bool neg = ip >= N;
  const uint64_t *pa = a + q * W * N + (ip & (N - 1));
  uint64_t *po = out + q * W * N + i;
  uint64_t carry = neg ? 1 : 0;
  for (int w = 0; w < W; w++) {
    uint64_t x = pa[(size_t)w * N];
    if (neg) { x = ~x; uint64_t s = x + carry; carry = s < carry; x = s; }
    po[(size_t)w * N] = x;
  }
}

}  // namespace hbdev
