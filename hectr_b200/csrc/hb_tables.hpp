// Host-side parameter and table construction for the CKKS engine: the engine's own restatement of
// what hectx_init() derives (GPQHE/src/precomp.c:328-453), re-shaped for the device kernels:
//   * prime chain p_k == 1 (mod 2N), 2^59 < p_k < 2^60          (precomp.c:358,372-376)
//   * twiddles zeta^bitrev(i) in PLAIN form with Shoup quotients  (reference: Montgomery form,
//     precomp.c:244-264; a*zeta mod p is the same integer either way)
//   * CRT constants per limb count D: (P_D/p_d)^-1 mod p_d and the LOW words of P_D/p_d and P_D
//     (reference: full MPIs, precomp.c:266-293) -- the kernels only ever need x mod 2^k
//   * key-switch constants for P = P_dim: full words of P/p_d, P, floor(P/2), P^-1 mod 2^(64*WC)
//   * level table logq_l, limb-count formulas                     (precomp.c:386-409, he-mult.c:51,99,168)
#pragma once
#include <cstdint>
#include <vector>
#include "hb_bigint.hpp"

namespace hb {

struct PrimeConsts {
  uint64_t p;
  uint64_t mu;      // floor(2^120 / p)  (== the reference's pinv_barr, reduce.c:75-78)
  uint64_t one_q;   // floor(2^64 / p)   (Shoup quotient of 1: reduces any 64-bit word)
  uint64_t r64;     // 2^64 mod p
  uint64_t r64_q;   // Shoup quotient of r64
  uint64_t ninv;    // N^-1 mod p (plain)
  uint64_t ninv_q;  // Shoup quotient
  uint64_t root;    // primitive 2N-th root used (g^((p-1)/2N), g least generator)
};

struct HostTables {
  unsigned logn = 0, N = 0, logq = 0, logDelta = 0, slots = 0;
  unsigned L = 0, dim = 0, dimevk = 0, dimub = 0;
  unsigned WC = 0;      // words of a top-level coefficient: ceil((logq+1)/64)
  unsigned WP = 0;      // words of P (= P_dim) plus one guard word
  unsigned nbitsP = 0, nbitsPqL = 0;
  std::vector<unsigned> logq_l;          // [L+1]
  std::vector<PrimeConsts> primes;       // [dimub]
  std::vector<uint64_t> tw, itw;         // [dimub][N][2]: (w, floor(w*2^64/p)), index = bit-reversed exponent
  // the stride-1 NTT pass's twiddles, lane-transposed (hb_dev.cuh struct Tw): [dimub][tr_stride][2]
  std::vector<uint64_t> tw_tr, itw_tr;
  unsigned tr_stride = 0, last_R = 0;
  std::vector<uint64_t> pow64;           // [dimub][WC+1]: 2^(64*w) mod p
  std::vector<uint64_t> crt_inv, crt_inv_q;  // [dimub+1][dimub]
  std::vector<uint64_t> crt_phat;        // [dimub+1][dimub][WC]
  std::vector<uint64_t> crt_P;           // [dimub+1][WC]
  std::vector<uint64_t> ks_inv, ks_inv_q;    // [dim]
  std::vector<uint64_t> ks_phat;         // [dim][WP]
  std::vector<uint64_t> ks_P, ks_Ph;     // [WP]
  std::vector<uint64_t> ks_Pinv;         // [WC]
  // Pre-scaling constants (Shoup pairs).  The kernels push the CRT weight (P_D/p_d)^-1 mod p_d through the linear
  // transforms, so the limbs reach the reconstruction already multiplied by it:
  //   dec_s[D][d][w] = 2^(64 w) (P_D/p_d)^-1 mod p_d, w <= WC : word weights of the scaled rns_decompose
  //   ks_z[D][e]     = (P/p_e)^-1 ((P_D/p_e)^-1)^-1 mod p_e, e < dim : from a D-limb weight to the weight of R = x mod P
  std::vector<uint64_t> dec_s;           // [dimub+1][dimub][WC+1][2]
  std::vector<uint64_t> ks_z;            // [dimub+1][dim][2]
  std::vector<BigU> Pfull;               // [dimub+1] P_D (host only; tests / diagnostics)
};

uint64_t mulmod(uint64_t a, uint64_t b, uint64_t m);
uint64_t powmod(uint64_t a, uint64_t e, uint64_t m);
bool is_prime_u64(uint64_t n);
uint64_t least_generator(uint64_t p);
uint64_t shoup_quotient(uint64_t w, uint64_t p);
unsigned last_pass_stages(unsigned logn, unsigned first);   // stages of the stride-1 NTT pass (device PassSize chain)
unsigned security_logq_bound(unsigned logn);

// Returns empty string on success, otherwise the reason the parameters are rejected
// (same conditions that make hectx_init abort, precomp.c:344-350,434-445,452).
std::string build_tables(HostTables &t, unsigned logn, unsigned logq, unsigned slots, unsigned logDelta);

// limb-count formulas (host metadata)
unsigned dim_keyswitch(const HostTables &t, unsigned l);          // he-automorphism.c:52 / he-mult.c:51
unsigned dim_mul(const HostTables &t, unsigned l);                // he-mult.c:99
unsigned dim_mulpt(const HostTables &t, unsigned l, double nu);   // he-mult.c:168
inline unsigned words_of_level(const HostTables &t, unsigned l) { return (t.logq_l[l] + 1 + 63) / 64; }

}  // namespace hb
