// C ABI of the engine (include/hectr_b200.h): context, containers, launch orchestration.
#include <algorithm>
#include <cstdio>
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <map>
#include <string>
#include <vector>
#include <cuda_runtime.h>
#include "../../include/hectr_b200.h"
#include "hb_tables.hpp"
#include "hb_tables32.hpp"
#include "hb_kernels.cuh"
#include "hb_dev32.cuh"
#include "hb_keygen.cuh"

extern "C" {
struct hb_ring;
hb_ring *hb_ring_create(unsigned logn);
void hb_ring_destroy(hb_ring *r);
void hb_ring_encode(const hb_ring *r, unsigned slots, const double *m, double Delta, uint64_t *words, unsigned W);
void hb_ring_decode(const hb_ring *r, unsigned slots, const uint64_t *words, unsigned W, double Delta, double *m);
void hb_ring_const(const hb_ring *r, double re, double im, double Delta, uint64_t *words, unsigned W);
}

using hbdev::Tables;

// ---------------------------------------------------------------------------------------------------------
// error handling: the reference prints a red "error:" line and aborts (precomp.c:344-350)
// ---------------------------------------------------------------------------------------------------------
[[noreturn]] static void hb_die(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  fprintf(stderr, "\033[1m\033[31merror:\033[0m \033[1mhectr_b200\033[0m. ");
  vfprintf(stderr, fmt, ap);
  fprintf(stderr, "\n");
  va_end(ap);
  abort();
}
#define CUDA_OK(expr)                                                                          \
  do {                                                                                         \
    cudaError_t e_ = (expr);                                                                   \
    if (e_ != cudaSuccess) hb_die("CUDA failure %s at %s:%d (%s)", cudaGetErrorString(e_), __FILE__, __LINE__, #expr); \
  } while (0)

struct hb_ctx;
static void use_device(const hb_ctx *ctx);      // every entry point that touches the GPU selects the context's device first

struct GemvPlan {
  std::vector<double> A;     // the matrix the plan was built from
  unsigned l = 0, D = 0;
  bool fused = false;        // the n1 products of one giant step may be summed before the CRT (bound check below)
  ulonglong2 *pthat = nullptr; // device [slots][D][N] Shoup pairs: ntt of the encoded, pre-rotated diagonals
  // the same diagonals on the internal 30-bit basis (hb_tables32.hpp), used when the fused bound holds
  bool use32 = false;
  unsigned D32 = 0;
  uint32_t *pthat32 = nullptr; // device [slots][D32][2][N]: multiplicand plane, quotient plane
  // giant steps i (of n2) with at least one non-zero encoded diagonal.  A giant step whose n1 diagonals all encode to the zero
  // polynomial (the hempc gain matrices are 8x3 / 8x2 blocks in a 16x16 frame: src/hempc.c:232-238) has the zero ciphertext
  // as inner sum, and he_rot of the zero ciphertext is the zero ciphertext: skipping it changes no word of the result.
  std::vector<unsigned> giant_nz;
};

struct hb_ctx {
  hb::HostTables ht;
  Tables dt;
  std::vector<void *> table_allocs;
  int device = 0, sm_count = 0;
  cudaStream_t stream = nullptr, own_stream = nullptr;
  cudaStream_t up_stream = nullptr, down_stream = nullptr;        // host-buffer pipeline (hb_ctr_hempc_eval_host)
  cudaEvent_t ev_up[2] = {nullptr, nullptr}, ev_comp[2] = {nullptr, nullptr}, ev_down[2] = {nullptr, nullptr};
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  uint64_t launches = 0;
  // optional per-kernel CUDA-event timing (bench.py roofline): name -> (launches, event pairs)
  bool prof = false;
  struct ProfRec { uint64_t n = 0; double ms = 0; std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pending; };
  std::map<std::string, ProfRec> prof_recs;
  std::vector<cudaEvent_t> prof_free;
  // workspaces (grow only)
  uint64_t *wsA = nullptr, *wsB = nullptr, *wsC = nullptr;
  size_t wsA_bytes = 0, wsB_bytes = 0, wsC_bytes = 0;
  bool exact_tail = false;          // HB_KS_EXACT=1: always take the multiword key-switch tail (tests)
  bool stepwise = false;            // HB_GEMV_STEPWISE=1: evaluate he_gemv one reference call at a time (tests, A/B timing)
  // pool of ciphertext buffers keyed by capacity in bytes
  std::multimap<size_t, uint64_t *> pool;
  std::map<uint64_t *, size_t> live;
  hb_ring *ring = nullptr;
  bool have32 = false;              // the 30-bit plaintext-product path is built for this ring (2^11 <= N <= 2^13)
  hb::HostTables32 ht32;
  hb32::Tables32 dt32;
  std::map<unsigned, uint16_t *> perms;   // g -> device perm_g[N] (k_ks_baby)
  uint64_t *wsD = nullptr;
  size_t wsD_bytes = 0;
  bool no_hoist = false;            // HB_GEMV_NOHOIST=1: one key-switch launch per baby rotation (A/B timing, tests)
  std::vector<GemvPlan> plans;
  // bit length of the largest |coefficient| of each uploaded evaluation key, as an integer (k_lift_bits): a key made by
  // he_genswk stays below P q_L / 2, which is what lets the summed giant steps (and any other exact-product shortcut) be used
  std::map<const void *, unsigned> evk_bits;
  uint64_t *keycheck_consts = nullptr;
  hbkg::Consts keycheck;
  // hybrid-basis key switch (hb_dev32.cuh): every uploaded key also lives on the first hyb_mmax primes of the 30-bit basis
  unsigned hyb_mmax = 0;            // 0: the hybrid path is not available for this context
  uint32_t *hyb_q32 = nullptr;      // DEVICE [hyb_mmax] primes
  std::map<const void *, uint32_t *> evk32;     // evk.p -> DEVICE key32[2][hyb_mmax][HB32_KEY_PLANES][N], times N^-1 (Montgomery form, or multiplicand / quotient planes)
  struct HybBasis { unsigned m = 0, Mbits = 0; ulonglong2 *dec59 = nullptr; uint2 *dec32 = nullptr; hbdev::KsHyb F; };
  std::map<unsigned, HybBasis> hyb; // constants of M = P * Q_m, by m
  bool no_hyb = false;              // HB_GEMV_NOHYB=1: every key-switch limb on the reference's 59-bit primes
  bool no_pair = false;             // HB_HEMPC_NOPAIR=1: ctr_hempc's two matrix-vector products as two passes
  bool no_sum = false;              // HB_GEMV_NOSUM=1: one tail per giant step (A/B timing, tests)
  bool no_skip = false;             // HB_GEMV_NOSKIP=1: evaluate identically zero giant steps too
  std::vector<uint64_t> enc_host;   // [WC][N] scratch for hb_he_ecd / gemv plans
  // streaming host interface (hb_ctr_hempc_submit_host / hb_ctr_hempc_wait): two persistent device buffer sets
  hb_ct job_in[2][4], job_up[2];
  unsigned job_cap = 0;             // instances each set holds
  uint64_t job_next = 0;            // ticket of the next submission
  bool job_busy[2] = {false, false};
  cudaEvent_t job_done[2] = {nullptr, nullptr};
};

static void use_device(const hb_ctx *ctx) {
  if (ctx) CUDA_OK(cudaSetDevice(ctx->device));   // a no-op in the driver when the device is already current
}

static void need_ct(const hb_ct *ct, const char *what);
static void pipeline_streams(hb_ctx *ctx);
static size_t ct_words(const hb_ctx *ctx, const hb_ct *ct);
static void check_level(const hb_ctx *ctx, unsigned l);

// ---------------------------------------------------------------------------------------------------------
// launch helpers
// ---------------------------------------------------------------------------------------------------------
static cudaEvent_t prof_event(hb_ctx *ctx) {
  if (!ctx->prof_free.empty()) { cudaEvent_t e = ctx->prof_free.back(); ctx->prof_free.pop_back(); return e; }
  cudaEvent_t e;
  CUDA_OK(cudaEventCreate(&e));
  return e;
}

template <class... KArgs, class... Args>
static void launch(hb_ctx *ctx, const char *name, void (*kernel)(KArgs...), unsigned grid, unsigned threads, size_t smem,
                   int cluster, Args... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof cfg);
  if (grid == 0) return;
  cfg.gridDim = dim3(grid, 1, 1);
  cfg.blockDim = dim3(threads, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = ctx->stream;
  cudaLaunchAttribute at[1];
  if (cluster > 1) {
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cluster;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
  }
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (ctx->prof) { e0 = prof_event(ctx); e1 = prof_event(ctx); CUDA_OK(cudaEventRecord(e0, ctx->stream)); }
  CUDA_OK(cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...));
  if (ctx->prof) {
    CUDA_OK(cudaEventRecord(e1, ctx->stream));
    hb_ctx::ProfRec &r = ctx->prof_recs[name];
    r.n++;
    r.pending.push_back({e0, e1});
  }
  ctx->launches++;
}

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device attribute: the cache is keyed by (current device, kernel)
template <class K>
static void want_smem(K kernel, size_t bytes) {
  static std::map<std::pair<int, void *>, size_t> done;
  int dev = 0;
  CUDA_OK(cudaGetDevice(&dev));
  const std::pair<int, void *> key(dev, (void *)kernel);
  auto it = done.find(key);
  if (it != done.end() && it->second >= bytes) return;
  CUDA_OK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  done[key] = bytes;
}

#ifdef HB_DEV_FAST   /* development builds: fewer template instantiations */
#define HB_LOGN_CASES(X) X(12) X(16)
#else
#define HB_LOGN_CASES(X) X(7) X(10) X(11) X(12) X(13) X(14) X(15) X(16)
#endif

#ifdef HB_DEV_FAST
#define HB_LOGN32_CASES(X) X(12)
#else
// N = 2^10 admits q <= 2^27 only (precomp.c:338-350): no matvec fits, so the 30-bit kernels start at 2^11
#define HB_LOGN32_CASES(X) X(11) X(12) X(13)
#endif

// ---- 30-bit plaintext-product path (hb_dev32.cuh) -----------------------------------------------------------------
template <int LOGN>
struct Launch32 {
  typedef hb32::Cfg32<LOGN> G;
  static void poly_ntt(hb_ctx *c, const uint64_t *src, size_t stride, int W, uint32_t *out, unsigned D, size_t npoly, bool canon) {
    if (!canon && (W == 1 || W == 2) && G::N >= 1024) {      // one CTA per polynomial, every limb from one read
      if (W == 1) launch(c, "k32_poly_ntt_all", hb32::k32_poly_ntt_all<LOGN, 1>, (unsigned)npoly, G::THREADS, G::SMEM_BYTES, 1, src, stride, out, c->dt32, D);
      else launch(c, "k32_poly_ntt_all", hb32::k32_poly_ntt_all<LOGN, 2>, (unsigned)npoly, G::THREADS, G::SMEM_BYTES, 1, src, stride, out, c->dt32, D);
      return;
    }
    if (canon) launch(c, "k32_poly_ntt", hb32::k32_poly_ntt<LOGN, true>, (unsigned)(npoly * D), G::THREADS, G::SMEM_BYTES, 1, src, stride, W, out, c->dt32, D);
    else launch(c, "k32_poly_ntt", hb32::k32_poly_ntt<LOGN, false>, (unsigned)(npoly * D), G::THREADS, G::SMEM_BYTES, 1, src, stride, W, out, c->dt32, D);
  }
  static void mulacc_intt(hb_ctx *c, const uint32_t *hat, size_t jstride, const uint32_t *pthat, const uint32_t *pthat_b, unsigned q_split,
                          size_t gstride, const hbdev::GroupMap &GM, unsigned n1, unsigned ngroups, uint32_t *chat, unsigned D, unsigned batch) {
    if (gstride == (size_t)n1 * D && (n1 == 2 || n1 == 4) && G::N >= 1024) {      // rotated transforms read once, in registers
      const unsigned grid = (unsigned)((size_t)batch * 2 * D);
      if (n1 == 4) launch(c, "k32_mulacc_intt_all", hb32::k32_mulacc_intt_all<LOGN, 4>, grid, G::THREADS, G::SMEM_BYTES, 1, hat, jstride, pthat, pthat_b, q_split, gstride, GM, ngroups, batch * 2, chat, c->dt32, D);
      else launch(c, "k32_mulacc_intt_all", hb32::k32_mulacc_intt_all<LOGN, 2>, grid, G::THREADS, G::SMEM_BYTES, 1, hat, jstride, pthat, pthat_b, q_split, gstride, GM, ngroups, batch * 2, chat, c->dt32, D);
      return;
    }
    launch(c, "k32_mulacc_intt", hb32::k32_mulacc_intt<LOGN>, (unsigned)((size_t)ngroups * batch * 2 * D), G::THREADS, G::SMEM_BYTES, 1, hat, jstride,
           pthat, pthat_b, q_split, gstride, GM, n1, batch * 2, chat, c->dt32, D);
  }
};
template <int LOGN>
struct LaunchHyb {
  typedef hb32::Cfg32<LOGN> G;
  static void ks_baby(hb_ctx *c, const uint64_t *src, size_t bstride, int W, const hb32::Baby32 &S, unsigned m, const uint2 *dec32, uint32_t *chat,
                      size_t jstride, unsigned batch) {
#define X(WT)                                                                                                                      \
    want_smem(hb32::k32_ks_baby<LOGN, WT>, 2 * G::SMEM_BYTES);                                                                     \
    launch(c, "k32_ks_baby", hb32::k32_ks_baby<LOGN, WT>, batch * m, G::THREADS, 2 * G::SMEM_BYTES, 1, src, bstride, W, S, m, c->hyb_mmax, dec32, \
           2 * c->ht.WC + 1, chat, jstride, c->dt32);
    if (W == 1) { X(1) } else if (W == 2) { X(2) } else { X(0) }
#undef X
  }
  static void ks_limb_sum(hb_ctx *c, const uint64_t *src, size_t bstride, size_t gstride, int W, const hb32::Groups32 &S, unsigned ng, unsigned m,
                          const uint2 *dec32, uint32_t *chatX, unsigned batch) {
#define X(WT)                                                                                                                      \
    want_smem(hb32::k32_ks_limb_sum<LOGN, WT>, 3 * G::SMEM_BYTES);                                                                 \
    launch(c, "k32_ks_limb_sum", hb32::k32_ks_limb_sum<LOGN, WT>, batch * m, G::THREADS, 3 * G::SMEM_BYTES, 1, src, bstride, gstride, W, S, ng, m, \
           c->hyb_mmax, dec32, 2 * c->ht.WC + 1, chatX, c->dt32);
    if (W == 1) { X(1) } else if (W == 2) { X(2) } else { X(0) }
#undef X
  }
  static void ntt_fwd_plain(hb_ctx *c, const uint32_t *in, uint32_t *out, size_t nlimbs, unsigned D) {
    launch(c, "k32_ntt_fwd_plain", hb32::k32_ntt_fwd_plain<LOGN>, (unsigned)nlimbs, G::THREADS, G::SMEM_BYTES, 1, in, out, c->dt32, D);
  }
};
static void do32_ks_baby(hb_ctx *c, const uint64_t *src, size_t bstride, int W, const hb32::Baby32 &S, unsigned m, const uint2 *dec32, uint32_t *chat,
                         size_t jstride, unsigned batch) {
#define X(L) case L: LaunchHyb<L>::ks_baby(c, src, bstride, W, S, m, dec32, chat, jstride, batch); break;
  switch (c->ht.logn) { HB_LOGN32_CASES(X) default: hb_die("logn=%u has no 30-bit path", c->ht.logn); }
#undef X
}
static void do32_ks_limb_sum(hb_ctx *c, const uint64_t *src, size_t bstride, size_t gstride, int W, const hb32::Groups32 &S, unsigned ng, unsigned m,
                             const uint2 *dec32, uint32_t *chatX, unsigned batch) {
#define X(L) case L: LaunchHyb<L>::ks_limb_sum(c, src, bstride, gstride, W, S, ng, m, dec32, chatX, batch); break;
  switch (c->ht.logn) { HB_LOGN32_CASES(X) default: hb_die("logn=%u has no 30-bit path", c->ht.logn); }
#undef X
}
static void do32_ntt_fwd_plain(hb_ctx *c, const uint32_t *in, uint32_t *out, size_t nlimbs, unsigned D) {
#define X(L) case L: LaunchHyb<L>::ntt_fwd_plain(c, in, out, nlimbs, D); break;
  switch (c->ht.logn) { HB_LOGN32_CASES(X) default: hb_die("logn=%u has no 30-bit path", c->ht.logn); }
#undef X
}

static void do32_poly_ntt(hb_ctx *c, const uint64_t *src, size_t stride, int W, uint32_t *out, unsigned D, size_t npoly, bool canon) {
#define X(L) case L: Launch32<L>::poly_ntt(c, src, stride, W, out, D, npoly, canon); break;
  switch (c->ht.logn) { HB_LOGN32_CASES(X) default: hb_die("logn=%u has no 30-bit path", c->ht.logn); }
#undef X
}
// pthat_b != nullptr: the second half of the batch (instances >= batch / 2) multiplies the diagonals pthat_b
static void do32_mulacc_intt(hb_ctx *c, const uint32_t *hat, size_t jstride, const uint32_t *pthat, const uint32_t *pthat_b, size_t gstride,
                             const hbdev::GroupMap &GM, unsigned n1, unsigned ngroups, uint32_t *chat, unsigned D, unsigned batch) {
  const unsigned q_split = pthat_b ? batch : 2 * batch;        // in polynomials (two per instance)
  if (!pthat_b) pthat_b = pthat;
#define X(L) case L: Launch32<L>::mulacc_intt(c, hat, jstride, pthat, pthat_b, q_split, gstride, GM, n1, ngroups, chat, D, batch); break;
  switch (c->ht.logn) { HB_LOGN32_CASES(X) default: hb_die("logn=%u has no 30-bit path", c->ht.logn); }
#undef X
}
static void do32_recon_smod(hb_ctx *c, const uint32_t *chat, unsigned D, uint64_t *out, int Wout, int k, size_t npoly);

template <int LOGN>
struct Launch {
  typedef hbdev::Cfg<LOGN> G;
  static void ntt_fwd(hb_ctx *c, const uint64_t *in, uint64_t *out, size_t nlimbs, unsigned d0, unsigned D, bool strict) {
    if (strict) {
      want_smem(hbdev::k_ntt_fwd<LOGN, true>, G::SMEM_BYTES);
      launch(c, "k_ntt_fwd_strict", hbdev::k_ntt_fwd<LOGN, true>, (unsigned)(nlimbs * G::C), G::THREADS, G::SMEM_BYTES, G::C, in, out, c->dt, d0, D);
    } else {
      want_smem(hbdev::k_ntt_fwd<LOGN, false>, G::SMEM_BYTES);
      launch(c, "k_ntt_fwd", hbdev::k_ntt_fwd<LOGN, false>, (unsigned)(nlimbs * G::C), G::THREADS, G::SMEM_BYTES, G::C, in, out, c->dt, d0, D);
    }
  }
  static void ntt_inv(hb_ctx *c, const uint64_t *in, uint64_t *out, size_t nlimbs, unsigned d0, unsigned D) {
    want_smem(hbdev::k_ntt_inv<LOGN>, G::SMEM_BYTES);
    launch(c, "k_ntt_inv", hbdev::k_ntt_inv<LOGN>, (unsigned)(nlimbs * G::C), G::THREADS, G::SMEM_BYTES, G::C, in, out, c->dt, d0, D);
  }
  static void poly_ntt(hb_ctx *c, const uint64_t *src, size_t stride, int W, uint64_t *out, unsigned D, size_t npoly) {
    want_smem(hbdev::k_poly_ntt<LOGN>, G::SMEM_BYTES);
    launch(c, "k_poly_ntt", hbdev::k_poly_ntt<LOGN>, (unsigned)(npoly * D * G::C), G::THREADS, G::SMEM_BYTES, G::C, src, stride, W, out, c->dt, D);
  }
  static void ks_limb(hb_ctx *c, const uint64_t *src, size_t bstride, int W, const hbdev::KsGroups &S,
                      unsigned key_limbs, uint64_t *chat, unsigned D, unsigned batch, unsigned d0, unsigned Dn, const ulonglong2 *dec) {
    want_smem(hbdev::k_ks_limb<LOGN>, 2 * G::SMEM_BYTES);
    launch(c, "k_ks_limb", hbdev::k_ks_limb<LOGN>, (unsigned)((size_t)batch * Dn * G::C), G::THREADS, 2 * G::SMEM_BYTES, G::C, src, bstride, W,
           S, key_limbs, chat, c->dt, D, d0, Dn, dec);
  }
  static void ks_limb_sum(hb_ctx *c, const uint64_t *src, size_t bstride, size_t gstride, int W, const hbdev::KsGroups &S, unsigned ng,
                          unsigned key_limbs, uint64_t *chatX, unsigned D, unsigned batch, unsigned d0, unsigned Dn) {
    if constexpr (G::C == 1) {
      want_smem(hbdev::k_ks_limb_sum<LOGN>, 3 * G::SMEM_BYTES);
      launch(c, "k_ks_limb_sum", hbdev::k_ks_limb_sum<LOGN>, (unsigned)((size_t)batch * Dn), G::THREADS, 3 * G::SMEM_BYTES, 1, src, bstride, gstride,
             W, S, ng, key_limbs, chatX, c->dt, D, d0, Dn);
    } else {
      hb_die("summed giant steps are built for N <= 2^13 only");
    }
  }
  static void ks_baby(hb_ctx *c, const uint64_t *src, size_t bstride, int W, const hbdev::BabySet &S, uint64_t *chat, size_t jstride,
                      unsigned D, unsigned batch, unsigned Dn, const ulonglong2 *dec) {
    want_smem(hbdev::k_ks_baby<LOGN>, 2 * G::SMEM_BYTES);
    launch(c, "k_ks_baby", hbdev::k_ks_baby<LOGN>, (unsigned)((size_t)batch * Dn * G::C), G::THREADS, 2 * G::SMEM_BYTES, G::C, src, bstride, W,
           S, c->ht.dimevk, chat, jstride, c->dt, D, Dn, dec);
  }
  static void mulpt_limb(hb_ctx *c, const uint64_t *ct, size_t poly_stride, int W, const ulonglong2 *pthat, uint64_t *chat, unsigned D,
                         size_t npoly) {
    want_smem(hbdev::k_mulpt_limb<LOGN>, G::SMEM_BYTES);
    launch(c, "k_mulpt_limb", hbdev::k_mulpt_limb<LOGN>, (unsigned)(npoly * D * G::C), G::THREADS, G::SMEM_BYTES, G::C, ct, poly_stride, W, pthat,
           chat, c->dt, D);
  }
  static void mulacc_intt(hb_ctx *c, const uint64_t *hat, size_t jstride, const ulonglong2 *pthat, size_t pthat_gstride,
                          const hbdev::GroupMap &GM, unsigned n1, unsigned ngroups, uint64_t *chat, unsigned D, unsigned batch) {
    want_smem(hbdev::k_mulacc_intt<LOGN>, G::SMEM_BYTES);
    launch(c, "k_mulacc_intt", hbdev::k_mulacc_intt<LOGN>, (unsigned)((size_t)ngroups * batch * 2 * D * G::C), G::THREADS, G::SMEM_BYTES, G::C,
           hat, jstride, pthat, pthat_gstride, GM, n1, batch * 2, chat, c->dt, D);
  }
  static void tensor_intt(hb_ctx *c, const uint64_t *hatA, const uint64_t *hatB, uint64_t *dhat, unsigned D, unsigned batch) {
    want_smem(hbdev::k_tensor_intt<LOGN>, G::SMEM_BYTES);
    launch(c, "k_tensor_intt", hbdev::k_tensor_intt<LOGN>, (unsigned)((size_t)batch * 3 * D * G::C), G::THREADS, G::SMEM_BYTES, G::C, hatA, hatB, dhat, c->dt, D);
  }
};

static void do_ntt_fwd(hb_ctx *c, const uint64_t *in, uint64_t *out, size_t nlimbs, unsigned d0, unsigned D, bool strict) {
#define X(L) case L: Launch<L>::ntt_fwd(c, in, out, nlimbs, d0, D, strict); break;
  switch (c->ht.logn) { HB_LOGN_CASES(X) default: hb_die("logn=%u is not built into this library", c->ht.logn); }
#undef X
}
static void do_ntt_inv(hb_ctx *c, const uint64_t *in, uint64_t *out, size_t nlimbs, unsigned d0, unsigned D) {
#define X(L) case L: Launch<L>::ntt_inv(c, in, out, nlimbs, d0, D); break;
  switch (c->ht.logn) { HB_LOGN_CASES(X) default: hb_die("logn=%u is not built into this library", c->ht.logn); }
#undef X
}
static void do_poly_ntt(hb_ctx *c, const uint64_t *src, size_t stride, int W, uint64_t *out, unsigned D, size_t npoly) {
#define X(L) case L: Launch<L>::poly_ntt(c, src, stride, W, out, D, npoly); break;
  switch (c->ht.logn) { HB_LOGN_CASES(X) default: hb_die("logn=%u is not built into this library", c->ht.logn); }
#undef X
}
// key_limbs: limbs per half of the key array (dimevk for evaluation keys, dim for an uploaded public key)
// batch = instances in the launch (all groups); S.gbatch instances per group
// limbs [d0, d0 + Dn) of a D-limb key switch (Dn = 0: all of them), written compactly as chat[b][h][Dn][N]
static void do_ks_limb_g(hb_ctx *c, const uint64_t *src, size_t bstride, int W, const hbdev::KsGroups &S, unsigned key_limbs,
                         uint64_t *chat, unsigned D, unsigned batch, unsigned d0 = 0, unsigned Dn = 0, const ulonglong2 *dec = nullptr) {
  if (!Dn) Dn = D;
#define X(L) case L: Launch<L>::ks_limb(c, src, bstride, W, S, key_limbs, chat, D, batch, d0, Dn, dec); break;
  switch (c->ht.logn) { HB_LOGN_CASES(X) default: hb_die("logn=%u is not built into this library", c->ht.logn); }
#undef X
}
static void do_ks_limb_sum(hb_ctx *c, const uint64_t *src, size_t bstride, size_t gstride, int W, const hbdev::KsGroups &S, unsigned ng,
                           unsigned key_limbs, uint64_t *chatX, unsigned D, unsigned batch, unsigned d0, unsigned Dn) {
#define X(L) case L: Launch<L>::ks_limb_sum(c, src, bstride, gstride, W, S, ng, key_limbs, chatX, D, batch, d0, Dn); break;
  switch (c->ht.logn) { HB_LOGN_CASES(X) default: hb_die("logn=%u is not built into this library", c->ht.logn); }
#undef X
}
static void do_ks_limb(hb_ctx *c, const uint64_t *src, size_t bstride, int W, unsigned g, unsigned ginv, const ulonglong2 *evk,
                       unsigned key_limbs, uint64_t *chat, unsigned D, unsigned batch) {
  hbdev::KsGroups S;
  memset(&S, 0, sizeof S);
  S.evk[0] = evk; S.g[0] = g; S.ginv[0] = ginv; S.gbatch = batch;
  do_ks_limb_g(c, src, bstride, W, S, key_limbs, chat, D, batch);
}
static void do_ks_baby(hb_ctx *c, const uint64_t *src, size_t bstride, int W, const hbdev::BabySet &S, uint64_t *chat, size_t jstride,
                       unsigned D, unsigned batch, unsigned Dn = 0, const ulonglong2 *dec = nullptr) {
  if (!Dn) Dn = D;
#define X(L) case L: Launch<L>::ks_baby(c, src, bstride, W, S, chat, jstride, D, batch, Dn, dec); break;
  switch (c->ht.logn) { HB_LOGN_CASES(X) default: hb_die("logn=%u is not built into this library", c->ht.logn); }
#undef X
}
static void do_mulpt_limb(hb_ctx *c, const uint64_t *ct, size_t poly_stride, int W, const ulonglong2 *pthat, uint64_t *chat, unsigned D,
                          size_t npoly) {
#define X(L) case L: Launch<L>::mulpt_limb(c, ct, poly_stride, W, pthat, chat, D, npoly); break;
  switch (c->ht.logn) { HB_LOGN_CASES(X) default: hb_die("logn=%u is not built into this library", c->ht.logn); }
#undef X
}
// ngroups giant steps in one launch: group i uses the diagonals pthat + i * pthat_gstride and writes chat[i][batch][2][D][N]
static void do_mulacc_intt(hb_ctx *c, const uint64_t *hat, size_t jstride, const ulonglong2 *pthat, size_t pthat_gstride,
                           const hbdev::GroupMap &GM, unsigned n1, unsigned ngroups, uint64_t *chat, unsigned D, unsigned batch) {
#define X(L) case L: Launch<L>::mulacc_intt(c, hat, jstride, pthat, pthat_gstride, GM, n1, ngroups, chat, D, batch); break;
  switch (c->ht.logn) { HB_LOGN_CASES(X) default: hb_die("logn=%u is not built into this library", c->ht.logn); }
#undef X
}
static void do_tensor_intt(hb_ctx *c, const uint64_t *hatA, const uint64_t *hatB, uint64_t *dhat, unsigned D, unsigned batch) {
#define X(L) case L: Launch<L>::tensor_intt(c, hatA, hatB, dhat, D, batch); break;
  switch (c->ht.logn) { HB_LOGN_CASES(X) default: hb_die("logn=%u is not built into this library", c->ht.logn); }
#undef X
}

// canonical limbs[j][N] (prime j % D) -> Shoup pairs
// scale: 0 plain, 1 times N^-1, 2 times N^-1 (P_Dw/p_d)^-1 (see k_shoup_prep)
static void do_shoup_prep(hb_ctx *c, const uint64_t *in, ulonglong2 *out, unsigned D, size_t nlimbs, int scale, unsigned Dw) {
  const size_t total = nlimbs * c->ht.N;
  launch(c, "k_shoup_prep", hbdev::k_shoup_prep, (unsigned)((total + 255) / 256), 256, 0, 1, in, out, c->dt, D, nlimbs, scale, Dw);
}

static int wk_class(unsigned wk) { return wk <= 1 ? 1 : wk <= 2 ? 2 : wk <= 4 ? 4 : wk <= 8 ? 8 : 16; }
static int wp_class(unsigned wp) { return wp <= 4 ? 4 : wp <= 8 ? 8 : wp <= 12 ? 12 : 20; }

// pre: the limbs already carry the CRT weight (P_D/p_d)^-1 (k_mulpt_limb, k_mulacc_intt)
static void do_recon_smod(hb_ctx *c, const uint64_t *chat, unsigned D, uint64_t *out, int Wout, int k, size_t npoly, bool pre) {
  const unsigned threads = 128;
  const size_t total = npoly * c->ht.N;
  const unsigned grid = (unsigned)((total + threads - 1) / threads);
  const unsigned wk = (k + 63) / 64;
  switch (wk_class(wk)) {
#define X(WKC) case WKC:                                                                                                           \
    if (pre) launch(c, "k_recon_smod", hbdev::k_recon_smod<WKC, true>, grid, threads, 0, 1, chat, D, out, Wout, k, c->dt, (unsigned)npoly);  \
    else launch(c, "k_recon_smod", hbdev::k_recon_smod<WKC, false>, grid, threads, 0, 1, chat, D, out, Wout, k, c->dt, (unsigned)npoly);     \
    break;
    X(1) X(2) X(4) X(8) X(16)
#undef X
  }
}

template <int WKC>
static bool recon32_fixed(hb_ctx *c, const uint32_t *chat, unsigned D, uint64_t *out, int Wout, int k, size_t npoly) {
  const unsigned threads = 128;
  const unsigned grid = (unsigned)((npoly * (c->ht.N / 4) + threads - 1) / threads);
  switch (D) {
#define X(DD) case DD: launch(c, "k32_recon_smod4", hb32::k32_recon_smod4<WKC, DD>, grid, threads, 0, 1, chat, out, Wout, k, c->dt32, (unsigned)npoly); return true;
    X(2) X(3) X(4) X(5) X(6) X(7) X(8)
#undef X
  }
  return false;
}

static void do32_recon_smod(hb_ctx *c, const uint32_t *chat, unsigned D, uint64_t *out, int Wout, int k, size_t npoly) {
  const int wkc = wk_class((k + 63) / 64);
  if (c->ht.N >= 64 && wkc <= 4) {       // compile-time limb count, four coefficients per thread
    if (wkc == 1 && recon32_fixed<1>(c, chat, D, out, Wout, k, npoly)) return;
    if (wkc == 2 && recon32_fixed<2>(c, chat, D, out, Wout, k, npoly)) return;
    if (wkc == 4 && recon32_fixed<4>(c, chat, D, out, Wout, k, npoly)) return;
  }
  const unsigned threads = 128;
  const size_t total = npoly * c->ht.N;
  const unsigned grid = (unsigned)((total + threads - 1) / threads);
  switch (wkc) {
#define X(WKC) case WKC: launch(c, "k32_recon_smod", hb32::k32_recon_smod<WKC>, grid, threads, 0, 1, chat, D, out, Wout, k, c->dt32, (unsigned)npoly); break;
    X(1) X(2) X(4) X(8) X(16)
#undef X
  }
}

template <int WKC>
static void ks_finish_wp(hb_ctx *c, unsigned grid, unsigned threads, const uint64_t *chat, unsigned D, const uint64_t *add0,
                         const uint64_t *add1, size_t add_bstride, int Wadd, const hbdev::TailGroups &TG, uint64_t *out, int Wout, int k,
                         unsigned batch) {
  switch (wp_class(c->ht.WP)) {
#define X(WPC) case WPC: launch(c, "k_ks_finish", hbdev::k_ks_finish<WKC, WPC>, grid, threads, 0, 1, chat, D, add0, add1, add_bstride, Wadd, TG, out, Wout, k, c->dt, batch); break;
    X(4) X(8) X(12) X(20)
#undef X
  }
}

// the fixed-point tails (k_ks_finish_fast, k_ks_finish_sum) cover D <= 8 limbs, dim <= 4, q_l <= 2^128
static bool ks_fast_ok(const hb_ctx *c, unsigned D, unsigned wk) {
  const hb::HostTables &h = c->ht;
  return D <= 8 && h.dim <= 4 && h.dim <= D && wk <= 2 && h.WP <= 4 && !c->exact_tail;
}
static void ks_fast_consts(const hb_ctx *c, unsigned D, hbdev::KsFast &F) {
  const hb::HostTables &h = c->ht;
  memset(&F, 0, sizeof F);
  F.D = D; F.dim = h.dim;
  for (unsigned d = 0; d < D; d++) {
    F.p[d] = h.primes[d].p; F.mu[d] = h.primes[d].mu;
    for (unsigned w = 0; w < 2 && w < h.WC; w++) F.phat[d][w] = h.crt_phat[((size_t)D * h.dimub + d) * h.WC + w];
  }
  for (unsigned w = 0; w < 2 && w < h.WC; w++) { F.PD[w] = h.crt_P[(size_t)D * h.WC + w]; F.Pinv[w] = h.ks_Pinv[w]; }
  for (unsigned e = 0; e < h.dim; e++) {
    F.kinv[e] = h.ks_z[((size_t)D * h.dim + e) * 2]; F.kinvq[e] = h.ks_z[((size_t)D * h.dim + e) * 2 + 1];
    for (unsigned w = 0; w < 2 && w < h.WP; w++) F.kphat[e][w] = h.ks_phat[(size_t)e * h.WP + w];
  }
  for (unsigned w = 0; w < 2 && w < h.WP; w++) F.kP[w] = h.ks_P[w];
  // t = (x - R) P^-1 mod 2^128 is linear in every term of x and R: fold P^-1 into the constants once
  typedef unsigned __int128 u128;
  const u128 pinv = ((u128)F.Pinv[1] << 64) | F.Pinv[0];
  auto fold = [&](uint64_t (&w)[2]) { const u128 v = ((((u128)w[1]) << 64) | w[0]) * pinv; w[0] = (uint64_t)v; w[1] = (uint64_t)(v >> 64); };
  for (unsigned d = 0; d < D; d++) fold(F.phat[d]);
  for (unsigned e = 0; e < h.dim; e++) fold(F.kphat[e]);
  fold(F.PD);
  fold(F.kP);
}

// tail of ng summed giant-step key switches + he_rs (k_ks_finish_sum); false when this (D, dim) pair is not instantiated
static bool do_ks_finish_sum(hb_ctx *c, const uint64_t *chatP, size_t chatP_gstride, const uint64_t *chatX, unsigned D, const uint64_t *add0,
                             size_t add_gstride, size_t add_bstride, int Wadd, const hbdev::SumGroups &SG, uint64_t *out, int Wout, int k,
                             int shift, int kout, unsigned batch, bool probe) {
  const hb::HostTables &h = c->ht;
  const unsigned threads = 128;
  const size_t total = (size_t)batch * 2 * h.N;
  const unsigned grid = (unsigned)((total + threads - 1) / threads);
  hbdev::KsFast F;
  if (!probe) ks_fast_consts(c, D, F);
#define KSS(DD, MM)                                                                                                              \
  if (D == DD && h.dim == MM) {                                                                                                  \
    if (!probe) launch(c, "k_ks_finish_sum", hbdev::k_ks_finish_sum<4, DD, MM>, grid, threads, 0, 1, chatP, chatP_gstride, chatX, F, add0, \
                       add_gstride, add_bstride, Wadd, SG, out, Wout, k, shift, kout, c->dt, batch);                             \
    return true;                                                                                                                 \
  }
  KSS(5, 2) KSS(6, 2) KSS(5, 3) KSS(6, 3) KSS(7, 3) KSS(8, 3)
#undef KSS
  return false;
}

// hybrid-basis tails (k_ks_finish_hyb, k_ks_finish_sum_hyb); probe: is this (dim, m) pair instantiated?
#define HB_HYB_PAIRS(X) X(3, 8) X(3, 7) X(3, 9) X(2, 4) X(2, 5) X(2, 6)
static bool do_ks_finish_hyb(hb_ctx *c, const uint64_t *chat59, const uint32_t *chat32, const hbdev::KsHyb *F, const uint64_t *add0,
                             const uint64_t *add1, size_t add_bstride, int Wadd, const hbdev::TailGroups &TG, uint64_t *out, int Wout, int k,
                             unsigned batch, unsigned dim, unsigned m) {
  const unsigned threads = 128;
  const unsigned grid = (unsigned)(((size_t)batch * 2 * c->ht.N + threads - 1) / threads);
#define X(DD, MM)                                                                                                                \
  if (dim == DD && m == MM) {                                                                                                    \
    if (F) launch(c, "k_ks_finish_hyb", hbdev::k_ks_finish_hyb<4, DD, MM>, grid, threads, 0, 1, chat59, chat32, *F, add0, add1, add_bstride, \
                  Wadd, TG, out, Wout, k, c->dt, batch);                                                                         \
    return true;                                                                                                                 \
  }
  HB_HYB_PAIRS(X)
#undef X
  return false;
}
static bool do_ks_finish_sum_hyb(hb_ctx *c, const uint64_t *chatP, size_t chatP_gstride, const uint32_t *chatX32, const hbdev::KsHyb *F,
                                 const uint64_t *add0, size_t add_gstride, size_t add_bstride, int Wadd, const hbdev::SumGroups &SG, uint64_t *out,
                                 int Wout, int k, int shift, int kout, unsigned batch, unsigned dim, unsigned m) {
  const unsigned threads = 128;
  const unsigned grid = (unsigned)(((size_t)batch * 2 * c->ht.N + threads - 1) / threads);
#define X(DD, MM)                                                                                                                \
  if (dim == DD && m == MM) {                                                                                                    \
    if (F) {                                                                                                                     \
      switch (SG.ng) {      /* giant steps actually evaluated: 3 for the CSTR gains, 4 for a full 16-slot matrix */              \
        Y(DD, MM, 2) Y(DD, MM, 3) Y(DD, MM, 4)                                                                                   \
        default: Y0(DD, MM)                                                                                                      \
      }                                                                                                                          \
    }                                                                                                                            \
    return true;                                                                                                                 \
  }
#define Y(DD, MM, NG)                                                                                                            \
  case NG: launch(c, "k_ks_finish_sum_hyb", hbdev::k_ks_finish_sum_hyb<4, DD, MM, NG>, grid, threads, 0, 1, chatP, chatP_gstride, chatX32, *F, \
                  add0, add_gstride, add_bstride, Wadd, SG, out, Wout, k, shift, kout, c->dt, batch); break;
#define Y0(DD, MM)                                                                                                               \
  launch(c, "k_ks_finish_sum_hyb", hbdev::k_ks_finish_sum_hyb<4, DD, MM, 0>, grid, threads, 0, 1, chatP, chatP_gstride, chatX32, *F, add0, \
         add_gstride, add_bstride, Wadd, SG, out, Wout, k, shift, kout, c->dt, batch);
  HB_HYB_PAIRS(X)
#undef X
#undef Y
#undef Y0
  return false;
}

static void do_ks_finish_g(hb_ctx *c, const uint64_t *chat, unsigned D, const uint64_t *add0, const uint64_t *add1,
                           size_t add_bstride, int Wadd, const hbdev::TailGroups &TG, uint64_t *out, int Wout, int k, unsigned batch) {
  const unsigned threads = 128;
  const size_t total = (size_t)batch * 2 * c->ht.N;
  const unsigned grid = (unsigned)((total + threads - 1) / threads);
  const unsigned wk = (k + 63) / 64;
  const hb::HostTables &h = c->ht;
  if (ks_fast_ok(c, D, wk)) {
    hbdev::KsFast F;
    ks_fast_consts(c, D, F);
    bool done = true;
#define KSF(DD, MM)                                                                                                             \
    if (D == DD && h.dim == MM) {                                                                                               \
      if (wk == 1) launch(c, "k_ks_finish_fast", hbdev::k_ks_finish_fast<1, 4, DD, MM>, grid, threads, 0, 1, chat, F, add0, add1, add_bstride, Wadd, TG, out, Wout, k, c->dt, batch); \
      else launch(c, "k_ks_finish_fast", hbdev::k_ks_finish_fast<2, 4, DD, MM>, grid, threads, 0, 1, chat, F, add0, add1, add_bstride, Wadd, TG, out, Wout, k, c->dt, batch);        \
    } else
    KSF(5, 2) KSF(6, 2) KSF(5, 3) KSF(6, 3) KSF(7, 3) KSF(8, 3) done = false;
#undef KSF
    if (done) return;
  }
  switch (wk_class(wk)) {
#define X(WKC) case WKC: ks_finish_wp<WKC>(c, grid, threads, chat, D, add0, add1, add_bstride, Wadd, TG, out, Wout, k, batch); break;
    X(1) X(2) X(4) X(8) X(16)
#undef X
  }
}
static void do_ks_finish(hb_ctx *c, const uint64_t *chat, unsigned D, const uint64_t *add0, const uint64_t *add1,
                         size_t add_bstride, int Wadd, unsigned ginv, uint64_t *out, int Wout, int k, unsigned batch) {
  hbdev::TailGroups TG;
  memset(&TG, 0, sizeof TG);
  TG.ginv[0] = ginv; TG.gbatch = batch; TG.add_shared = 0;
  do_ks_finish_g(c, chat, D, add0, add1, add_bstride, Wadd, TG, out, Wout, k, batch);
}

// ---------------------------------------------------------------------------------------------------------
// context
// ---------------------------------------------------------------------------------------------------------
template <class T>
static T *upload_vec(hb_ctx *ctx, const std::vector<T> &v) {
  void *d = nullptr;
  size_t bytes = v.size() * sizeof(T);
  CUDA_OK(cudaMalloc(&d, bytes ? bytes : 8));
  if (bytes) CUDA_OK(cudaMemcpy(d, v.data(), bytes, cudaMemcpyHostToDevice));
  ctx->table_allocs.push_back(d);
  return (T *)d;
}

extern "C" hb_ctx *hb_ctx_create(unsigned logn, unsigned logq, unsigned slots, unsigned logDelta, int device) {
  hb_ctx *ctx = new hb_ctx();
  std::string err = hb::build_tables(ctx->ht, logn, logq, slots, logDelta);
  if (!err.empty()) hb_die("Invalid argument. %s", err.c_str());
  if (ctx->ht.WC > 16) hb_die("log2(q)=%u needs more than 16 words per coefficient", logq);
  if (ctx->ht.WP > 20) hb_die("special modulus P of %u bits is too wide for this build", ctx->ht.nbitsP);
  {
    bool ok = false;
#define X(L) if (logn == L) ok = true;
    HB_LOGN_CASES(X)
#undef X
    if (!ok) hb_die("logn=%u is not built into this library", logn);
  }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    hb_die("no CUDA device: this engine has no CPU fallback (%s)", e == cudaSuccess ? "0 devices" : cudaGetErrorString(e));
  if (device < 0 || device >= ndev) hb_die("device %d out of range (have %d)", device, ndev);
  ctx->device = device;
  CUDA_OK(cudaSetDevice(device));
  cudaDeviceProp prop;
  CUDA_OK(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) hb_die("device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
  ctx->sm_count = prop.multiProcessorCount;
  CUDA_OK(cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking));
  ctx->stream = ctx->own_stream;
  CUDA_OK(cudaEventCreate(&ctx->ev0));
  CUDA_OK(cudaEventCreate(&ctx->ev1));
  const hb::HostTables &h = ctx->ht;
  static_assert(sizeof(hb::PrimeConsts) == sizeof(hbdev::PrimeConsts), "PrimeConsts layout");
  Tables &t = ctx->dt;
  t.primes = (const hbdev::PrimeConsts *)upload_vec(ctx, h.primes);
  t.tw = (const ulonglong2 *)upload_vec(ctx, h.tw);
  t.itw = (const ulonglong2 *)upload_vec(ctx, h.itw);
  t.tw_tr = (const ulonglong2 *)upload_vec(ctx, h.tw_tr);
  t.itw_tr = (const ulonglong2 *)upload_vec(ctx, h.itw_tr);
  t.tr_stride = h.tr_stride;
  t.pow64 = upload_vec(ctx, h.pow64);
  t.crt_inv = upload_vec(ctx, h.crt_inv);
  t.crt_inv_q = upload_vec(ctx, h.crt_inv_q);
  t.crt_phat = upload_vec(ctx, h.crt_phat);
  t.crt_P = upload_vec(ctx, h.crt_P);
  t.ks_inv = upload_vec(ctx, h.ks_inv);
  t.ks_inv_q = upload_vec(ctx, h.ks_inv_q);
  t.ks_phat = upload_vec(ctx, h.ks_phat);
  t.ks_P = upload_vec(ctx, h.ks_P);
  t.ks_Ph = upload_vec(ctx, h.ks_Ph);
  t.ks_Pinv = upload_vec(ctx, h.ks_Pinv);
  t.dec_s = (const ulonglong2 *)upload_vec(ctx, h.dec_s);
  t.ks_z = (const ulonglong2 *)upload_vec(ctx, h.ks_z);
  t.N = h.N; t.logn = h.logn; t.dimub = h.dimub; t.WC = h.WC; t.WP = h.WP; t.dim = h.dim;
  {
    const char *e = getenv("HB_GEMV_STEPWISE");
    ctx->stepwise = e && *e && *e != '0';
    e = getenv("HB_KS_EXACT");
    ctx->exact_tail = e && *e && *e != '0';
    e = getenv("HB_GEMV_NOHOIST");
    ctx->no_hoist = e && *e && *e != '0';
    e = getenv("HB_HEMPC_NOPAIR");
    ctx->no_pair = e && *e && *e != '0';
    e = getenv("HB_GEMV_NOHYB");
    ctx->no_hyb = e && *e && *e != '0';
    e = getenv("HB_GEMV_NOSUM");
    ctx->no_sum = e && *e && *e != '0';
    e = getenv("HB_GEMV_NOSKIP");
    ctx->no_skip = e && *e && *e != '0';
  }
  ctx->ring = hb_ring_create(logn);
  ctx->enc_host.assign((size_t)h.WC * h.N, 0);
  {
    bool ok32 = false;
#define X(L) if (logn == L) ok32 = true;
    HB_LOGN32_CASES(X)
#undef X
    const char *e = getenv("HB_NO_PT30");
    if (ok32 && !(e && *e && *e != '0')) {
      std::string err32 = hb::build_tables32(ctx->ht32, logn, h.WC, h.Pfull[h.dimub].bits());
      if (!err32.empty()) hb_die("%s", err32.c_str());
      static_assert(sizeof(hb::Prime32) == sizeof(hb32::Prime32), "Prime32 layout");
      const hb::HostTables32 &g = ctx->ht32;
      hb32::Tables32 &d = ctx->dt32;
      d.primes = (const hb32::Prime32 *)upload_vec(ctx, g.primes);
      d.tw = (const uint2 *)upload_vec(ctx, g.tw);
      d.itw = (const uint2 *)upload_vec(ctx, g.itw);
      d.tw_tr = (const uint2 *)upload_vec(ctx, g.tw_tr);
      d.itw_tr = (const uint2 *)upload_vec(ctx, g.itw_tr);
      d.tr_stride = g.tr_stride;
      d.pow32 = (const uint2 *)upload_vec(ctx, g.pow32);
      d.crt_inv = (const uint2 *)upload_vec(ctx, g.crt_inv);
      d.crt_qhat = upload_vec(ctx, g.crt_qhat);
      d.crt_Q = upload_vec(ctx, g.crt_Q);
      d.N = g.N; d.logn = g.logn; d.nprimes = g.nprimes; d.WC = g.WC; d.ndig = g.ndig;
      ctx->have32 = true;
      // Hybrid key switch: limbs needed for the largest product this context can ask for (top level, 8 summed giant steps,
      // a key of P q_L / 2): |x| < 2^bound must stay below M / 8, M = P Q_m >= 2^(nbitsP + qbits[m] - 2).
      const unsigned bound = 3 + h.logn + (h.logq_l[h.L] - 1) + (h.nbitsPqL - 1);
      unsigned m = 0;
      for (unsigned t = 1; t <= g.nprimes; t++)
        if (h.nbitsP + g.qbits[t] - 2 >= bound + 3) { m = t; break; }
      if (m && m <= HB_HYB_MAX32 && h.dim <= 4 && h.dim < h.dimevk && h.WP <= 4 && !ctx->no_hyb) {
        ctx->hyb_mmax = m;
        std::vector<uint32_t> q(m);
        for (unsigned t = 0; t < m; t++) q[t] = g.primes[t].p;
        ctx->hyb_q32 = upload_vec(ctx, q);
      }
    }
  }
  return ctx;
}

extern "C" void hb_ctx_destroy(hb_ctx *ctx) { use_device(ctx);
  if (!ctx) return;
  CUDA_OK(cudaSetDevice(ctx->device));
  cudaStreamSynchronize(ctx->stream);
  for (void *p : ctx->table_allocs) cudaFree(p);
  for (auto &kv : ctx->pool) cudaFree(kv.second);
  for (auto &kv : ctx->live) cudaFree(kv.first);
  for (auto &pl : ctx->plans) { cudaFree(pl.pthat); cudaFree(pl.pthat32); }
  for (auto &kv : ctx->prof_recs)
    for (auto &pr : kv.second.pending) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
  for (cudaEvent_t e : ctx->prof_free) cudaEventDestroy(e);
  cudaFree(ctx->wsA);
  cudaFree(ctx->wsB);
  cudaFree(ctx->wsC);
  cudaFree(ctx->wsD);
  cudaFree(ctx->keycheck_consts);
  for (auto &kv : ctx->perms) cudaFree(kv.second);
  for (auto &kv : ctx->evk32) cudaFree(kv.second);
  for (auto &kv : ctx->hyb) { cudaFree(kv.second.dec59); cudaFree(kv.second.dec32); }
  if (ctx->up_stream) {
    cudaStreamDestroy(ctx->up_stream);
    cudaStreamDestroy(ctx->down_stream);
    for (int s2 = 0; s2 < 2; s2++) { cudaEventDestroy(ctx->ev_up[s2]); cudaEventDestroy(ctx->ev_comp[s2]); cudaEventDestroy(ctx->ev_down[s2]); }
  }
  for (int q = 0; q < 2; q++)
    if (ctx->job_done[q]) cudaEventDestroy(ctx->job_done[q]);
  cudaEventDestroy(ctx->ev0);
  cudaEventDestroy(ctx->ev1);
  cudaStreamDestroy(ctx->own_stream);
  hb_ring_destroy(ctx->ring);
  delete ctx;
}

static void fill_info(const hb::HostTables &h, hb_info *o);
extern "C" void hb_ctx_info(const hb_ctx *ctx, hb_info *o) {
  fill_info(ctx->ht, o);
  o->device = ctx->device; o->sm_count = ctx->sm_count;
}
static void check_level(const hb_ctx *ctx, unsigned l) {
  if (l > ctx->ht.L) hb_die("level %u out of range (L=%u)", l, ctx->ht.L);
}
extern "C" unsigned hb_ctx_words(const hb_ctx *ctx, unsigned l) { check_level(ctx, l); return hb::words_of_level(ctx->ht, l); }
extern "C" unsigned hb_ctx_logq(const hb_ctx *ctx, unsigned l) { check_level(ctx, l); return ctx->ht.logq_l[l]; }
extern "C" void hb_ctx_primes(const hb_ctx *ctx, uint64_t *out) {
  for (unsigned d = 0; d < ctx->ht.dimub; d++) out[d] = ctx->ht.primes[d].p;
}
extern "C" void hb_ctx_twiddles(const hb_ctx *ctx, unsigned d, int inverse, uint64_t *out) {
  if (d >= ctx->ht.dimub) hb_die("prime index %u out of range", d);
  const std::vector<uint64_t> &src = inverse ? ctx->ht.itw : ctx->ht.tw;
  for (unsigned i = 0; i < ctx->ht.N; i++) out[i] = src[((size_t)d * ctx->ht.N + i) * 2];
}
extern "C" unsigned hb_dim_keyswitch(const hb_ctx *ctx, unsigned l) { check_level(ctx, l); return hb::dim_keyswitch(ctx->ht, l); }
extern "C" unsigned hb_dim_mul(const hb_ctx *ctx, unsigned l) { check_level(ctx, l); return hb::dim_mul(ctx->ht, l); }
extern "C" unsigned hb_dim_mulpt(const hb_ctx *ctx, unsigned l, double nu) { check_level(ctx, l); return hb::dim_mulpt(ctx->ht, l, nu); }
extern "C" void hb_ctx_set_stream(hb_ctx *ctx, void *s) { use_device(ctx); ctx->stream = s ? (cudaStream_t)s : ctx->own_stream; }
extern "C" void hb_sync(hb_ctx *ctx) { use_device(ctx); CUDA_OK(cudaStreamSynchronize(ctx->stream)); }
extern "C" void hb_timer_start(hb_ctx *ctx) { use_device(ctx); CUDA_OK(cudaEventRecord(ctx->ev0, ctx->stream)); }
extern "C" float hb_timer_stop_ms(hb_ctx *ctx) { use_device(ctx);
  CUDA_OK(cudaEventRecord(ctx->ev1, ctx->stream));
  CUDA_OK(cudaEventSynchronize(ctx->ev1));
  float ms = 0;
  CUDA_OK(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
  return ms;
}
extern "C" uint64_t hb_launch_count(const hb_ctx *ctx) { return ctx->launches; }
extern "C" void hb_prof_enable(hb_ctx *ctx, int on) { use_device(ctx); ctx->prof = on != 0; }
extern "C" void hb_prof_reset(hb_ctx *ctx) { use_device(ctx);
  CUDA_OK(cudaStreamSynchronize(ctx->stream));
  for (auto &kv : ctx->prof_recs)
    for (auto &pr : kv.second.pending) { ctx->prof_free.push_back(pr.first); ctx->prof_free.push_back(pr.second); }
  ctx->prof_recs.clear();
}
extern "C" size_t hb_prof_report(hb_ctx *ctx, char *buf, size_t cap) { use_device(ctx);
  CUDA_OK(cudaStreamSynchronize(ctx->stream));
  std::string out = "{";
  bool first = true;
  for (auto &kv : ctx->prof_recs) {
    for (auto &pr : kv.second.pending) {
      float ms = 0;
      CUDA_OK(cudaEventElapsedTime(&ms, pr.first, pr.second));
      kv.second.ms += ms;
      ctx->prof_free.push_back(pr.first); ctx->prof_free.push_back(pr.second);
    }
    kv.second.pending.clear();
    char line[256];
    snprintf(line, sizeof line, "%s\"%s\": {\"launches\": %llu, \"ms\": %.6f}", first ? "" : ", ", kv.first.c_str(),
             (unsigned long long)kv.second.n, kv.second.ms);
    out += line;
    first = false;
  }
  out += "}";
  if (buf && cap) { size_t n = out.size() < cap - 1 ? out.size() : cap - 1; memcpy(buf, out.data(), n); buf[n] = 0; }
  return out.size();
}
extern "C" void hb_ct_export(hb_ctx *ctx, const hb_ct *ct, void *dst_device) { use_device(ctx);
  need_ct(ct, "hb_ct_export");
  CUDA_OK(cudaMemcpyAsync(dst_device, ct->words, ct_words(ctx, ct) * 8, cudaMemcpyDeviceToDevice, ctx->stream));
}
extern "C" void hb_ct_import(hb_ctx *ctx, hb_ct *ct, const void *src_device, unsigned l, double nu, double B) { use_device(ctx);
  need_ct(ct, "hb_ct_import"); check_level(ctx, l);
  ct->l = l; ct->nu = nu; ct->B = B;
  CUDA_OK(cudaMemcpyAsync(ct->words, src_device, ct_words(ctx, ct) * 8, cudaMemcpyDeviceToDevice, ctx->stream));
}

// ---------------------------------------------------------------------------------------------------------
// host-only tables view
// ---------------------------------------------------------------------------------------------------------
struct hb_tables {
  hb::HostTables ht;
  hb_ring *ring = nullptr;
};
extern "C" hb_tables *hb_tables_create(unsigned logn, unsigned logq, unsigned slots, unsigned logDelta) {
  hb_tables *t = new hb_tables();
  std::string err = hb::build_tables(t->ht, logn, logq, slots, logDelta);
  if (!err.empty()) {
    fprintf(stderr, "\033[1m\033[31merror:\033[0m \033[1mhectr_b200\033[0m. Invalid argument. %s\n", err.c_str());
    delete t;
    return nullptr;
  }
  t->ring = hb_ring_create(logn);
  return t;
}
extern "C" void hb_tables_destroy(hb_tables *t) { if (t) { hb_ring_destroy(t->ring); delete t; } }
static void fill_info(const hb::HostTables &h, hb_info *o) {
  o->logn = h.logn; o->n = h.N; o->logq = h.logq; o->logDelta = h.logDelta; o->slots = h.slots;
  o->L = h.L; o->dim = h.dim; o->dimevk = h.dimevk; o->dimub = h.dimub; o->words_top = h.WC;
  o->nbitsP = h.nbitsP; o->nbitsPqL = h.nbitsPqL; o->device = -1; o->sm_count = 0;
}
extern "C" void hb_tables_info(const hb_tables *t, hb_info *o) { fill_info(t->ht, o); }
extern "C" void hb_tables_primes(const hb_tables *t, uint64_t *out) {
  for (unsigned d = 0; d < t->ht.dimub; d++) out[d] = t->ht.primes[d].p;
}
extern "C" void hb_tables_twiddles(const hb_tables *t, unsigned d, int inverse, uint64_t *out) {
  if (d >= t->ht.dimub) hb_die("prime index %u out of range", d);
  const std::vector<uint64_t> &src = inverse ? t->ht.itw : t->ht.tw;
  for (unsigned i = 0; i < t->ht.N; i++) out[i] = src[((size_t)d * t->ht.N + i) * 2];
}
extern "C" void hb_tables_crt(const hb_tables *t, unsigned D, uint64_t *inv, uint64_t *phat_low, uint64_t *P_low) {
  const hb::HostTables &h = t->ht;
  if (D == 0 || D > h.dimub) hb_die("limb count %u outside the prime chain (%u)", D, h.dimub);
  for (unsigned d = 0; d < D; d++) {
    inv[d] = h.crt_inv[(size_t)D * h.dimub + d];
    for (unsigned w = 0; w < h.WC; w++) phat_low[(size_t)d * h.WC + w] = h.crt_phat[((size_t)D * h.dimub + d) * h.WC + w];
  }
  for (unsigned w = 0; w < h.WC; w++) P_low[w] = h.crt_P[(size_t)D * h.WC + w];
}
extern "C" unsigned hb_tables_dim(const hb_tables *t, int which, unsigned l, double nu) {
  if (l > t->ht.L) hb_die("level %u out of range (L=%u)", l, t->ht.L);
  return which == 0 ? hb::dim_keyswitch(t->ht, l) : which == 1 ? hb::dim_mul(t->ht, l) : hb::dim_mulpt(t->ht, l, nu);
}
extern "C" void hb_tables_encode(const hb_tables *t, const double *m, double Delta, uint64_t *words) {
  hb_ring_encode(t->ring, t->ht.slots, m, Delta, words, t->ht.WC);
}
extern "C" void hb_tables_decode(const hb_tables *t, const uint64_t *words, unsigned W, double Delta, double *m) {
  hb_ring_decode(t->ring, t->ht.slots, words, W, Delta, m);
}

// ---------------------------------------------------------------------------------------------------------
// raw memory
// ---------------------------------------------------------------------------------------------------------
extern "C" void *hb_dev_alloc(hb_ctx *ctx, size_t bytes) { use_device(ctx);
  CUDA_OK(cudaSetDevice(ctx->device));
  void *p = nullptr;
  CUDA_OK(cudaMalloc(&p, bytes ? bytes : 8));
  return p;
}
extern "C" void hb_dev_free(hb_ctx *ctx, void *p) { use_device(ctx);
  if (!p) return;
  CUDA_OK(cudaStreamSynchronize(ctx->stream));
  CUDA_OK(cudaFree(p));
}
extern "C" void *hb_host_alloc(hb_ctx *ctx, size_t bytes) { use_device(ctx);
  (void)ctx;
  void *p = nullptr;
  CUDA_OK(cudaMallocHost(&p, bytes ? bytes : 8));
  return p;
}
extern "C" void hb_host_free(hb_ctx *ctx, void *p) { use_device(ctx); (void)ctx; if (p) CUDA_OK(cudaFreeHost(p)); }
extern "C" void hb_h2d(hb_ctx *ctx, void *dst, const void *src, size_t bytes) { use_device(ctx);
  CUDA_OK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
}
extern "C" void hb_d2h(hb_ctx *ctx, void *dst, const void *src, size_t bytes) { use_device(ctx);
  CUDA_OK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
}
extern "C" void hb_dev_fill(hb_ctx *ctx, void *dst, int byte, size_t bytes) { use_device(ctx);
  CUDA_OK(cudaMemsetAsync(dst, byte, bytes, ctx->stream));
}

// the smallest pooled buffer of sufficient capacity (at most twice the request, so a small handle never pins a large
// buffer); a ragged last chunk of the host pipeline reuses the full-size buffers instead of a synchronising cudaMalloc
static uint64_t *pool_get(hb_ctx *ctx, size_t bytes) {
  auto it = ctx->pool.lower_bound(bytes);
  uint64_t *p;
  size_t cap = bytes;
  if (it != ctx->pool.end() && it->first <= 2 * bytes) { p = it->second; cap = it->first; ctx->pool.erase(it); }
  else { void *q = nullptr; CUDA_OK(cudaMalloc(&q, bytes)); p = (uint64_t *)q; }
  ctx->live[p] = cap;
  return p;
}
static void pool_put(hb_ctx *ctx, uint64_t *p) {
  if (!p) return;
  auto it = ctx->live.find(p);
  if (it == ctx->live.end()) hb_die("freeing a ciphertext buffer this context does not own");
  ctx->pool.insert({it->second, p});
  ctx->live.erase(it);
}
static uint64_t *ws_get(hb_ctx *ctx, uint64_t *&ws, size_t &cur, size_t bytes) {
  if (bytes > cur) {
    if (ws) { CUDA_OK(cudaStreamSynchronize(ctx->stream)); CUDA_OK(cudaFree(ws)); }
    void *q = nullptr;
    CUDA_OK(cudaMalloc(&q, bytes));
    ws = (uint64_t *)q;
    cur = bytes;
  }
  return ws;
}

// ---------------------------------------------------------------------------------------------------------
// containers
// ---------------------------------------------------------------------------------------------------------
static size_t ct_capacity(const hb_ctx *ctx, unsigned batch) { return (size_t)batch * 2 * ctx->ht.WC * ctx->ht.N * 8; }
static size_t ct_words(const hb_ctx *ctx, const hb_ct *ct) { return (size_t)ct->batch * 2 * hb::words_of_level(ctx->ht, ct->l) * ctx->ht.N; }

extern "C" void hb_ct_alloc(hb_ctx *ctx, hb_ct *ct, unsigned batch) { use_device(ctx);
  if (batch == 0) hb_die("a ciphertext handle must hold at least one instance");
  CUDA_OK(cudaSetDevice(ctx->device));
  ct->l = ctx->ht.L; ct->nu = 0; ct->B = 0; ct->batch = batch;
  ct->words = pool_get(ctx, ct_capacity(ctx, batch));
  CUDA_OK(cudaMemsetAsync(ct->words, 0, ct_capacity(ctx, batch), ctx->stream));   // gcry_mpi_new(0) == 0
}
// temporaries that are fully overwritten before they are read: no zero fill
static void ct_alloc_tmp(hb_ctx *ctx, hb_ct *ct, unsigned batch) {
  ct->l = ctx->ht.L; ct->nu = 0; ct->B = 0; ct->batch = batch;
  ct->words = pool_get(ctx, ct_capacity(ctx, batch));
}
extern "C" void hb_ct_free(hb_ctx *ctx, hb_ct *ct) { use_device(ctx);
  pool_put(ctx, ct->words);
  ct->words = nullptr; ct->batch = 0;
}
static void need_ct(const hb_ct *ct, const char *what) {
  if (!ct || !ct->words) hb_die("%s: ciphertext handle is not allocated", what);
}
extern "C" void hb_ct_copy(hb_ctx *ctx, hb_ct *dst, const hb_ct *src) { use_device(ctx);
  need_ct(dst, "hb_ct_copy"); need_ct(src, "hb_ct_copy");
  if (dst->batch != src->batch) hb_die("hb_ct_copy: batch mismatch (%u vs %u)", dst->batch, src->batch);
  dst->l = src->l; dst->nu = src->nu; dst->B = src->B;
  if (dst->words != src->words)
    CUDA_OK(cudaMemcpyAsync(dst->words, src->words, ct_words(ctx, src) * 8, cudaMemcpyDeviceToDevice, ctx->stream));
}
extern "C" void hb_ct_upload(hb_ctx *ctx, hb_ct *ct, const uint64_t *host, unsigned l, double nu, double B) { use_device(ctx);
  need_ct(ct, "hb_ct_upload"); check_level(ctx, l);
  ct->l = l; ct->nu = nu; ct->B = B;
  CUDA_OK(cudaMemcpyAsync(ct->words, host, ct_words(ctx, ct) * 8, cudaMemcpyHostToDevice, ctx->stream));
}
extern "C" void hb_ct_download(hb_ctx *ctx, const hb_ct *ct, uint64_t *host) { use_device(ctx);
  need_ct(ct, "hb_ct_download");
  CUDA_OK(cudaMemcpyAsync(host, ct->words, ct_words(ctx, ct) * 8, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_OK(cudaStreamSynchronize(ctx->stream));
}
extern "C" void hb_pt_alloc(hb_ctx *ctx, hb_pt *pt) { use_device(ctx);
  pt->nu = 0;
  pt->words = (uint64_t *)hb_dev_alloc(ctx, (size_t)ctx->ht.WC * ctx->ht.N * 8);
  CUDA_OK(cudaMemsetAsync(pt->words, 0, (size_t)ctx->ht.WC * ctx->ht.N * 8, ctx->stream));
}
extern "C" void hb_pt_free(hb_ctx *ctx, hb_pt *pt) { use_device(ctx); hb_dev_free(ctx, pt->words); pt->words = nullptr; }
extern "C" void hb_pt_upload(hb_ctx *ctx, hb_pt *pt, const uint64_t *host, double nu) { use_device(ctx);
  if (!pt->words) hb_die("hb_pt_upload: plaintext handle is not allocated");
  pt->nu = nu;
  CUDA_OK(cudaMemcpyAsync(pt->words, host, (size_t)ctx->ht.WC * ctx->ht.N * 8, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_OK(cudaStreamSynchronize(ctx->stream));
}
// largest |coefficient| (bit length) of a key given as canonical-or-[0,p] transform limbs raw[2][dimevk][N] on the device.
// key32_out (hybrid path available): the same key on the 30-bit primes, key32[2][hyb_mmax][HB32_KEY_PLANES][N], times N^-1.
static unsigned key_lift_bits(hb_ctx *ctx, const uint64_t *raw, uint32_t **key32_out) {
  const hb::HostTables &h = ctx->ht;
  const unsigned D = h.dimevk;
  const size_t N = h.N;
  if (!ctx->keycheck_consts) {
    const hb::BigU &PD = h.Pfull[D];
    const unsigned WX = (PD.bits() + 8 + 63) / 64 + 1;
    if (WX > (unsigned)hbkg::MAXW) return ~0u;                       // wider than the generic kernels: treated as unbounded
    std::vector<uint64_t> hc;
    auto put = [&](const std::vector<uint64_t> &v) { const size_t off = hc.size(); hc.insert(hc.end(), v.begin(), v.end()); return off; };
    std::vector<uint64_t> primes(D), inv(D), phat((size_t)D * WX);
    for (unsigned d = 0; d < D; d++) {
      primes[d] = h.primes[d].p;
      const hb::BigU ph = PD.divmod_u64(primes[d], nullptr);
      inv[d] = hb::powmod(ph.mod_u64(primes[d]), primes[d] - 2, primes[d]);
      const std::vector<uint64_t> lo = ph.low(WX);
      std::copy(lo.begin(), lo.end(), phat.begin() + (size_t)d * WX);
    }
    const size_t o_pr = put(primes), o_inv = put(inv), o_ph = put(phat), o_PD = put(PD.low(WX)), o_PDh = put(PD.shr1().low(WX));
    uint64_t *dc = (uint64_t *)hb_dev_alloc(ctx, hc.size() * 8 + 8);
    CUDA_OK(cudaMemcpyAsync(dc, hc.data(), hc.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
    CUDA_OK(cudaStreamSynchronize(ctx->stream));
    ctx->keycheck_consts = dc;
    ctx->keycheck = hbkg::Consts{dc + o_pr, dc + o_inv, dc + o_ph, dc + o_PD, dc + o_PDh, nullptr, nullptr, nullptr, nullptr, D, WX, 0, (unsigned)N};
  }
  const size_t limbs = (size_t)2 * D * N;
  uint64_t *coef = (uint64_t *)hb_dev_alloc(ctx, limbs * 8 + 8);
  unsigned *dmax = (unsigned *)(coef + limbs);
  CUDA_OK(cudaMemsetAsync(dmax, 0, 8, ctx->stream));
  do_ntt_inv(ctx, raw, coef, (size_t)2 * D, 0, D);
  const unsigned m = key32_out ? ctx->hyb_mmax : 0;
  uint32_t *res32 = nullptr, *hat32 = nullptr;
  if (m) {
    res32 = (uint32_t *)hb_dev_alloc(ctx, (size_t)2 * m * N * 4);
    hat32 = (uint32_t *)hb_dev_alloc(ctx, (size_t)2 * m * N * 4);
  }
  launch(ctx, "k_lift_bits", hbkg::k_lift_bits, (unsigned)((2 * N + 127) / 128), 128, 0, 1, (const uint64_t *)coef, ctx->keycheck, 2u, dmax,
         (const uint32_t *)ctx->hyb_q32, m, res32);
  if (m) {
    uint32_t *key32 = (uint32_t *)hb_dev_alloc(ctx, (size_t)2 * m * HB32_KEY_PLANES * N * 4);
    do32_ntt_fwd_plain(ctx, res32, hat32, (size_t)2 * m, m);
    const size_t total = (size_t)2 * m * N;
    launch(ctx, "k32_shoup_prep", hb32::k32_shoup_prep, (unsigned)((total + 255) / 256), 256, 0, 1, (const uint32_t *)hat32, key32, ctx->dt32, m, (size_t)2 * m, 0u,
           (unsigned)HB32_KEY_PLANES);
    *key32_out = key32;
  }
  unsigned bits = 0;
  CUDA_OK(cudaMemcpyAsync(&bits, dmax, 4, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_OK(cudaStreamSynchronize(ctx->stream));
  CUDA_OK(cudaFree(coef));
  if (res32) { CUDA_OK(cudaFree(res32)); CUDA_OK(cudaFree(hat32)); }
  return bits;
}

// constants of the hybrid basis M = P * Q_m (hb_dev32.cuh), built once per m
static const hb_ctx::HybBasis &hyb_basis(hb_ctx *ctx, unsigned m) {
  auto it = ctx->hyb.find(m);
  if (it != ctx->hyb.end()) return it->second;
  const hb::HostTables &h = ctx->ht;
  const hb::HostTables32 &g = ctx->ht32;
  hb_ctx::HybBasis B;
  B.m = m;
  hb::BigU Q(1);
  for (unsigned i = 0; i < m; i++) Q = Q.mul_u64(g.primes[i].p);
  const hb::BigU &P = h.Pfull[h.dim];
  const hb::BigU M = hb::BigU::mul(P, Q);
  B.Mbits = M.bits();
  hbdev::KsHyb &F = B.F;
  memset(&F, 0, sizeof F);
  F.dim = h.dim; F.m = m;
  typedef unsigned __int128 u128;
  u128 pinv = 0;
  for (unsigned w = 0; w < 2 && w < h.WC; w++) pinv |= (u128)h.ks_Pinv[w] << (64 * w);
  auto fold2 = [&](const hb::BigU &v, uint64_t (&o)[2]) {
    const std::vector<uint64_t> lo = v.low(2);
    const u128 x = ((((u128)lo[1]) << 64) | lo[0]) * pinv;
    o[0] = (uint64_t)x; o[1] = (uint64_t)(x >> 64);
  };
  std::vector<uint64_t> dec59((size_t)h.dim * (h.WC + 1) * 2), ztab((size_t)h.dim * 2);
  for (unsigned e = 0; e < h.dim; e++) {
    const uint64_t pe = h.primes[e].p;
    const hb::BigU Mhat = M.divmod_u64(pe, nullptr);
    const uint64_t mh = Mhat.mod_u64(pe), inv = hb::powmod(mh, pe - 2, pe);
    F.p[e] = pe; F.mu[e] = h.primes[e].mu;
    F.rp[e] = (float)(4294967296.0 / (double)pe);
    fold2(Mhat, F.phat[e]);
    for (unsigned w = 0; w <= h.WC; w++) {
      const uint64_t c = hb::mulmod(h.pow64[(size_t)e * (h.WC + 1) + w], inv, pe);
      dec59[((size_t)e * (h.WC + 1) + w) * 2] = c;
      dec59[((size_t)e * (h.WC + 1) + w) * 2 + 1] = hb::shoup_quotient(c, pe);
    }
    // y_e = a_e (M/p_e)^-1  ->  z_e = a_e (P/p_e)^-1 = y_e (M/p_e) (P/p_e)^-1
    const uint64_t z = hb::mulmod(h.ks_inv[e], mh, pe);
    F.kinv[e] = z; F.kinvq[e] = hb::shoup_quotient(z, pe);
    ztab[2 * e] = z; ztab[2 * e + 1] = F.kinvq[e];
    hb::BigU kph;
    kph.w.assign(h.ks_phat.begin() + (size_t)e * h.WP, h.ks_phat.begin() + (size_t)(e + 1) * h.WP);
    kph.trim();
    fold2(kph, F.kphat[e]);
  }
  fold2(P, F.kP);
  fold2(M, F.M);
  const unsigned nd = 2 * h.WC + 1;
  std::vector<uint32_t> dec32((size_t)m * nd * 2);
  for (unsigned i = 0; i < m; i++) {
    const uint64_t qi = g.primes[i].p;
    const hb::BigU Qhat = M.divmod_u64(qi, nullptr);
    const uint64_t inv = hb::powmod(Qhat.mod_u64(qi), qi - 2, qi);
    F.q[i] = (uint32_t)qi; F.mu32[i] = g.primes[i].mu;
    F.rq[i] = (float)(1.0 / (double)qi);
    fold2(Qhat, F.qhat[i]);
    for (unsigned w = 0; w < nd; w++) {
      const uint64_t c = hb::mulmod(g.pow32[((size_t)i * (g.ndig + 1) + w) * 2], inv, qi);
      dec32[((size_t)i * nd + w) * 2] = (uint32_t)c;
      dec32[((size_t)i * nd + w) * 2 + 1] = hb::shoup_quotient32((uint32_t)c, (uint32_t)qi);
    }
  }
  // one allocation: dec59 | ztab, and dec32
  std::vector<uint64_t> pack(dec59);
  pack.insert(pack.end(), ztab.begin(), ztab.end());
  B.dec59 = (ulonglong2 *)hb_dev_alloc(ctx, pack.size() * 8);
  CUDA_OK(cudaMemcpyAsync(B.dec59, pack.data(), pack.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
  B.dec32 = (uint2 *)hb_dev_alloc(ctx, dec32.size() * 4);
  CUDA_OK(cudaMemcpyAsync(B.dec32, dec32.data(), dec32.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_OK(cudaStreamSynchronize(ctx->stream));
  F.ztab = B.dec59 + (size_t)h.dim * (h.WC + 1);
  return ctx->hyb.emplace(m, B).first->second;
}

// 30-bit limbs needed so that |x| < 2^bound_bits stays below M / 8 (0: not available)
static unsigned hyb_limbs_for(const hb_ctx *ctx, unsigned bound_bits) {
  if (!ctx->hyb_mmax) return 0;
  for (unsigned m = 1; m <= ctx->hyb_mmax; m++)
    if (ctx->ht.nbitsP + ctx->ht32.qbits[m] - 2 >= bound_bits + 3) return m;
  return 0;
}

extern "C" void hb_evk_upload(hb_ctx *ctx, hb_evk *evk, const uint64_t *p0, const uint64_t *p1) { use_device(ctx);
  // stored as Shoup pairs (w, floor(w 2^64 / p_d)): the key is a fixed multiplicand of every key switch
  const size_t half = (size_t)ctx->ht.dimevk * ctx->ht.N;
  if (ctx->ht.dimevk > ctx->ht.dimub) hb_die("hb_evk_upload: keys of %u limbs but the prime chain has %u", ctx->ht.dimevk, ctx->ht.dimub);
  uint64_t *raw = (uint64_t *)hb_dev_alloc(ctx, 2 * half * 8);
  evk->p = (uint64_t *)hb_dev_alloc(ctx, 2 * half * 16);
  CUDA_OK(cudaMemcpyAsync(raw, p0, half * 8, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_OK(cudaMemcpyAsync(raw + half, p1, half * 8, cudaMemcpyHostToDevice, ctx->stream));
  do_shoup_prep(ctx, raw, (ulonglong2 *)evk->p, ctx->ht.dimevk, (size_t)2 * ctx->ht.dimevk, 1, 0);   // key * N^-1
  uint32_t *key32 = nullptr;
  ctx->evk_bits[evk->p] = key_lift_bits(ctx, raw, ctx->hyb_mmax ? &key32 : nullptr);
  if (key32) ctx->evk32[evk->p] = key32;
  hb_dev_free(ctx, raw);   // synchronises
}
extern "C" void hb_evk_free(hb_ctx *ctx, hb_evk *evk) { use_device(ctx);
  ctx->evk_bits.erase(evk->p);
  auto it32 = ctx->evk32.find(evk->p);
  if (it32 != ctx->evk32.end()) { hb_dev_free(ctx, it32->second); ctx->evk32.erase(it32); }
  hb_dev_free(ctx, evk->p);
  evk->p = nullptr;
}

// ---------------------------------------------------------------------------------------------------------
// encode / decode
// ---------------------------------------------------------------------------------------------------------
extern "C" void hb_encode(const hb_ctx *ctx, const double *m, double Delta, uint64_t *words) {
  hb_ring_encode(ctx->ring, ctx->ht.slots, m, Delta, words, ctx->ht.WC);
}
extern "C" void hb_decode(const hb_ctx *ctx, const uint64_t *words, unsigned W, double Delta, double *m) {
  hb_ring_decode(ctx->ring, ctx->ht.slots, words, W, Delta, m);
}
extern "C" void hb_he_ecd(hb_ctx *ctx, hb_pt *pt, const double *m) { use_device(ctx);
  if (!pt->words) hb_die("hb_he_ecd: plaintext handle is not allocated");
  // he_ecd writes 2*slots coefficients and keeps the rest (he-encode.c:58-63): read-modify-write
  const size_t bytes = (size_t)ctx->ht.WC * ctx->ht.N * 8;
  CUDA_OK(cudaMemcpyAsync(ctx->enc_host.data(), pt->words, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_OK(cudaStreamSynchronize(ctx->stream));
  const double Delta = std::ldexp(1.0, (int)ctx->ht.logDelta);
  hb_ring_encode(ctx->ring, ctx->ht.slots, m, Delta, ctx->enc_host.data(), ctx->ht.WC);
  hb_pt_upload(ctx, pt, ctx->enc_host.data(), Delta);
}

extern "C" void hb_he_const_pt(hb_ctx *ctx, hb_pt *pt, double re, double im) { use_device(ctx);
  if (!pt->words) hb_die("hb_he_const_pt: plaintext handle is not allocated");
  // he_const_pt writes coefficients 0 and N/2 and keeps the rest (he-encode.c:119-125): read-modify-write like hb_he_ecd
  const size_t bytes = (size_t)ctx->ht.WC * ctx->ht.N * 8;
  CUDA_OK(cudaMemcpyAsync(ctx->enc_host.data(), pt->words, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_OK(cudaStreamSynchronize(ctx->stream));
  const double Delta = std::ldexp(1.0, (int)ctx->ht.logDelta);
  hb_ring_const(ctx->ring, re, im, Delta, ctx->enc_host.data(), ctx->ht.WC);
  hb_pt_upload(ctx, pt, ctx->enc_host.data(), Delta);
}

// ---------------------------------------------------------------------------------------------------------
// evaluation
// ---------------------------------------------------------------------------------------------------------
static void same_shape(const hb_ct *a, const hb_ct *b, const char *what) {
  need_ct(a, what); need_ct(b, what);
  if (a->batch != b->batch) hb_die("%s: batch mismatch (%u vs %u)", what, a->batch, b->batch);
}

static void addsub(hb_ctx *ctx, hb_ct *ct, const hb_ct *a, const hb_ct *b, int op, const char *what) {
  same_shape(a, b, what); same_shape(ct, a, what);
  if (a->l != b->l) hb_die("%s: operands are at different levels (%u vs %u)", what, a->l, b->l);   // assert, he-add.c:35
  ct->l = a->l;
  ct->nu = (a->nu >= b->nu) ? a->nu : b->nu;
  ct->B = a->B + b->B;
  const int W = hb::words_of_level(ctx->ht, ct->l), k = ctx->ht.logq_l[ct->l];
  const size_t npoly = (size_t)ct->batch * 2, total = npoly * ctx->ht.N;
  launch(ctx, "k_ct_addsub", hbdev::k_ct_addsub, (unsigned)((total + 255) / 256), 256, 0, 1, a->words, b->words, ct->words, W, k, op, ctx->ht.N, npoly);
}
extern "C" void hb_he_add(hb_ctx *ctx, hb_ct *ct, const hb_ct *a, const hb_ct *b) { use_device(ctx); addsub(ctx, ct, a, b, 0, "hb_he_add"); }
extern "C" void hb_he_sub(hb_ctx *ctx, hb_ct *ct, const hb_ct *a, const hb_ct *b) { use_device(ctx); addsub(ctx, ct, a, b, 1, "hb_he_sub"); }
static void addsub_pt(hb_ctx *ctx, hb_ct *dst, const hb_ct *src, const hb_pt *pt, int op, const char *what) {
  same_shape(dst, src, what);
  if (!pt || !pt->words) hb_die("%s: plaintext handle is not allocated", what);
  dst->l = src->l;
  dst->nu = (src->nu >= pt->nu) ? src->nu : pt->nu;        // he-add.c:84
  dst->B = src->B;
  const int W = hb::words_of_level(ctx->ht, dst->l), k = ctx->ht.logq_l[dst->l];
  const size_t npoly = (size_t)dst->batch * 2, total = npoly * ctx->ht.N;
  launch(ctx, "k_ct_addpt", hbdev::k_ct_addpt, (unsigned)((total + 255) / 256), 256, 0, 1, src->words, pt->words, dst->words, W, k, op, ctx->ht.N, npoly);
}
extern "C" void hb_he_addpt(hb_ctx *ctx, hb_ct *dst, const hb_ct *src, const hb_pt *pt) { use_device(ctx); addsub_pt(ctx, dst, src, pt, 0, "hb_he_addpt"); }
extern "C" void hb_he_subpt(hb_ctx *ctx, hb_ct *dst, const hb_ct *src, const hb_pt *pt) { use_device(ctx); addsub_pt(ctx, dst, src, pt, 1, "hb_he_subpt"); }
extern "C" void hb_he_neg(hb_ctx *ctx, hb_ct *ct) { use_device(ctx);
  need_ct(ct, "hb_he_neg");
  const int W = hb::words_of_level(ctx->ht, ct->l), k = ctx->ht.logq_l[ct->l];
  const size_t npoly = (size_t)ct->batch * 2, total = npoly * ctx->ht.N;
  launch(ctx, "k_ct_addsub", hbdev::k_ct_addsub, (unsigned)((total + 255) / 256), 256, 0, 1, ct->words, (const uint64_t *)nullptr, ct->words, W, k, 2,
         ctx->ht.N, npoly);
}

static void rescale_like(hb_ctx *ctx, hb_ct *ct, int shift, const char *what) {
  need_ct(ct, what);
  if (ct->l == 0) hb_die("%s: ciphertext is already at level 0", what);
  const int Win = hb::words_of_level(ctx->ht, ct->l);
  ct->l -= 1;
  const int Wout = hb::words_of_level(ctx->ht, ct->l), kout = ctx->ht.logq_l[ct->l];
  uint64_t *out = pool_get(ctx, ct_capacity(ctx, ct->batch));
  const size_t npoly = (size_t)ct->batch * 2, total = npoly * ctx->ht.N;
  launch(ctx, "k_ct_rescale", hbdev::k_ct_rescale, (unsigned)((total + 255) / 256), 256, 0, 1, ct->words, out, Win, Wout, shift, kout, ctx->ht.N, npoly);
  pool_put(ctx, ct->words);
  ct->words = out;
}
extern "C" void hb_he_rs(hb_ctx *ctx, hb_ct *ct) { use_device(ctx);
  need_ct(ct, "hb_he_rs");
  const double Delta = std::ldexp(1.0, (int)ctx->ht.logDelta);
  const double n = ctx->ht.N, Brs = std::sqrt(n / 3.) * (3 + 8 * std::sqrt(64.0));   // precomp.c:417
  ct->nu /= Delta;
  ct->B = ct->B / Delta + Brs;
  rescale_like(ctx, ct, (int)ctx->ht.logDelta, "hb_he_rs");
}
extern "C" void hb_he_moddown(hb_ctx *ctx, hb_ct *ct) { use_device(ctx); rescale_like(ctx, ct, 0, "hb_he_moddown"); }

// pthat[D][N] -> chat -> dst   (shared by hb_he_mulpt and the gemv plan)
static void mulpt_with_hat(hb_ctx *ctx, hb_ct *dst, const hb_ct *src, const ulonglong2 *pthat, unsigned D, double ptnu) {
  const int W = hb::words_of_level(ctx->ht, src->l), k = ctx->ht.logq_l[src->l];
  uint64_t *chat = ws_get(ctx, ctx->wsA, ctx->wsA_bytes, (size_t)src->batch * 2 * D * ctx->ht.N * 8);
  do_mulpt_limb(ctx, src->words, (size_t)W * ctx->ht.N, W, pthat, chat, D, (size_t)src->batch * 2);
  dst->l = src->l;
  const double nu = src->nu, B = src->B;
  dst->nu = nu * ptnu;
  dst->B = B * ptnu;
  do_recon_smod(ctx, chat, D, dst->words, W, k, (size_t)src->batch * 2, true);
}

extern "C" void hb_he_mulpt(hb_ctx *ctx, hb_ct *dst, const hb_ct *src, const hb_pt *pt) { use_device(ctx);
  same_shape(dst, src, "hb_he_mulpt");
  if (!pt || !pt->words) hb_die("hb_he_mulpt: plaintext handle is not allocated");
  const unsigned D = hb::dim_mulpt(ctx->ht, src->l, pt->nu);
  if (D == 0 || D > ctx->ht.dimub) hb_die("hb_he_mulpt: limb count %u outside the prime chain (%u)", D, ctx->ht.dimub);
  const size_t hatw = (size_t)D * ctx->ht.N;
  uint64_t *raw = ws_get(ctx, ctx->wsB, ctx->wsB_bytes, hatw * 24);
  ulonglong2 *pthat = (ulonglong2 *)(raw + hatw);
  do_poly_ntt(ctx, pt->words, (size_t)ctx->ht.WC * ctx->ht.N, ctx->ht.WC, raw, D, 1);
  do_shoup_prep(ctx, raw, pthat, D, D, 2, D);
  mulpt_with_hat(ctx, dst, src, pthat, D, pt->nu);
}

// inverse of an odd number modulo 2N (Newton iteration)
static unsigned inv_odd_mod_2n(const hb_ctx *ctx, unsigned g) {
  unsigned x = g;
  for (int i = 0; i < 5; i++) x *= 2 - g * x;
  return x & (2 * ctx->ht.N - 1);
}

static unsigned pow5_inv_mod_2n(const hb_ctx *ctx, unsigned rot) {
  const unsigned m = 2 * ctx->ht.N;
  unsigned g = 1;
  for (unsigned j = 0; j < rot; j++) g = (unsigned)(((uint64_t)g * 5u) % m);   // poly.c:265-268
  // inverse of an odd number modulo a power of two (Newton)
  unsigned x = g;
  for (int i = 0; i < 5; i++) x *= 2 - g * x;
  return x & (m - 1);
}

// he_swk with the automorphism ginv; add1 = true adds sigma(c1) too (he_relin's d1), he_swk does not.
static void keyswitch_inplace(hb_ctx *ctx, hb_ct *ct, unsigned ginv, const hb_evk *key, const char *what) {
  need_ct(ct, what);
  if (!key || !key->p) hb_die("%s: evaluation key is not loaded", what);
  const hb::HostTables &h = ctx->ht;
  const unsigned D = hb::dim_keyswitch(h, ct->l);
  if (D > h.dimevk || D > h.dimub) hb_die("%s: needs %u limbs but keys hold %u and the prime chain %u", what, D, h.dimevk, h.dimub);
  const int W = hb::words_of_level(h, ct->l), k = h.logq_l[ct->l];
  const size_t bstride = (size_t)2 * W * h.N;
  uint64_t *chat = ws_get(ctx, ctx->wsA, ctx->wsA_bytes, (size_t)ct->batch * 2 * D * h.N * 8);
  do_ks_limb(ctx, ct->words + (size_t)W * h.N, bstride, W, inv_odd_mod_2n(ctx, ginv), ginv, (const ulonglong2 *)key->p, h.dimevk, chat, D, ct->batch);
  uint64_t *out = pool_get(ctx, ct_capacity(ctx, ct->batch));
  do_ks_finish(ctx, chat, D, ct->words, nullptr, bstride, W, ginv, out, W, k, ct->batch);
  pool_put(ctx, ct->words);
  ct->words = out;
}

extern "C" void hb_he_rot(hb_ctx *ctx, hb_ct *ct, int rot, const hb_evk *rk) { use_device(ctx);
  if (rot < 0 || (unsigned)rot >= ctx->ht.slots) hb_die("hb_he_rot: rotation %d outside [0,%u)", rot, ctx->ht.slots);
  keyswitch_inplace(ctx, ct, pow5_inv_mod_2n(ctx, (unsigned)rot), &rk[rot], "hb_he_rot");
}
extern "C" void hb_he_conj(hb_ctx *ctx, hb_ct *ct, const hb_evk *ck) { use_device(ctx);
  keyswitch_inplace(ctx, ct, 2 * ctx->ht.N - 1, ck, "hb_he_conj");
}

extern "C" void hb_he_mul(hb_ctx *ctx, hb_ct *ct, const hb_ct *a, const hb_ct *b, const hb_evk *rlk) { use_device(ctx);
  same_shape(a, b, "hb_he_mul"); same_shape(ct, a, "hb_he_mul");
  if (a->l != b->l) hb_die("hb_he_mul: operands are at different levels (%u vs %u)", a->l, b->l);   // assert, he-mult.c:90
  if (!rlk || !rlk->p) hb_die("hb_he_mul: relinearisation key is not loaded");
  const hb::HostTables &h = ctx->ht;
  const unsigned l = a->l, batch = a->batch;
  const unsigned Dm = hb::dim_mul(h, l), Dk = hb::dim_keyswitch(h, l);
  if (Dm > h.dimub || Dk > h.dimevk) hb_die("hb_he_mul: limb count outside the prime chain");
  const int W = hb::words_of_level(h, l), k = h.logq_l[l];
  const size_t N = h.N, polyw = (size_t)W * N;
  // workspace A: hat[b][4][Dm][N] then (reused) key-switch chat[b][2][Dk][N]; B: dhat[b][3][Dm][N] + d[b][3][W][N]
  const size_t hat_words = (size_t)batch * 4 * Dm * N, chat_words = (size_t)batch * 2 * Dk * N;
  uint64_t *wsA = ws_get(ctx, ctx->wsA, ctx->wsA_bytes, (hat_words > chat_words ? hat_words : chat_words) * 8);
  const size_t dhat_words = (size_t)batch * 3 * Dm * N, d_words = (size_t)batch * 3 * polyw;
  uint64_t *wsB = ws_get(ctx, ctx->wsB, ctx->wsB_bytes, (dhat_words + d_words) * 8);
  uint64_t *hat = wsA, *dhat = wsB, *dpoly = wsB + dhat_words;
  // 4 forward transforms per limb: hatA[b][0..1] from a, hatB[b][0..1] from b
  uint64_t *hatA = hat, *hatB = hat + hat_words / 2;
  do_poly_ntt(ctx, a->words, polyw, W, hatA, Dm, (size_t)batch * 2);
  do_poly_ntt(ctx, b->words, polyw, W, hatB, Dm, (size_t)batch * 2);
  do_tensor_intt(ctx, hatA, hatB, dhat, Dm, batch);
  do_recon_smod(ctx, dhat, Dm, dpoly, W, k, (size_t)batch * 3, false);      // d0,d1,d2 mod q_l  (he-mult.c:140-142)
  // relinearise d2 with rlk, add d0 and d1 (he-mult.c:40-85)
  uint64_t *chat = wsA;
  do_ks_limb(ctx, dpoly + 2 * polyw, 3 * polyw, W, 1u, 1u, (const ulonglong2 *)rlk->p, h.dimevk, chat, Dk, batch);
  const double nu1 = a->nu, nu2 = b->nu, B1 = a->B, B2 = b->B;
  ct->l = l;
  ct->nu = nu1 * nu2;
  {
    // bnd.Bmult[l] = Pinv*q_l*Bks + Brs (precomp.c:411-429) -- metadata only
    long double Pinv = 1;
    for (unsigned d = 0; d < h.dimub; d++) Pinv *= 1. / h.primes[d].p;
    const double sigma = 3.1915382432114616, n = h.N;
    const double Bks = 8 * sigma * n / std::sqrt(3.0), Brs = std::sqrt(n / 3.) * (3 + 8 * std::sqrt(64.0));
    long double Pinvql = Pinv * (long double)std::ldexp(1.0, (int)h.logq_l[0]);
    for (unsigned j = 1; j <= l; j++) Pinvql = Pinvql * std::ldexp(1.0, (int)h.logDelta);
    ct->B = nu1 * B2 + nu2 * B1 + B1 * B2 + (double)(Pinvql * Bks + Brs);
  }
  do_ks_finish(ctx, chat, Dk, dpoly, dpoly + polyw, 3 * polyw, W, 1u, ct->words, W, k, batch);
}

// ---------------------------------------------------------------------------------------------------------
// he_gemv (he-algo.c:47-93)
// ---------------------------------------------------------------------------------------------------------
static const GemvPlan &gemv_plan(hb_ctx *ctx, const double *A, unsigned l) {
  const hb::HostTables &h = ctx->ht;
  const unsigned S = h.slots;
  const size_t nA = (size_t)S * S * 2;
  for (const GemvPlan &p : ctx->plans)
    if (p.l == l && memcmp(p.A.data(), A, nA * sizeof(double)) == 0) return p;
  if (ctx->plans.size() >= 8) {   // small LRU: drop the oldest
    CUDA_OK(cudaStreamSynchronize(ctx->stream));
    cudaFree(ctx->plans.front().pthat);
    cudaFree(ctx->plans.front().pthat32);
    ctx->plans.erase(ctx->plans.begin());
  }
  GemvPlan plan;
  plan.A.assign(A, A + nA);
  plan.l = l;
  const double Delta = std::ldexp(1.0, (int)h.logDelta);
  plan.D = hb::dim_mulpt(h, l, Delta);
  if (plan.D == 0 || plan.D > h.dimub) hb_die("hb_he_gemv: limb count %u outside the prime chain", plan.D);
  unsigned n1 = (unsigned)std::sqrt((double)S);
  if (S != n1 * n1) n1 = (unsigned)std::sqrt((double)(2 * S));
  const unsigned n2 = S / n1;
  // encode all diagonals on the host in the reference's order; the single he_pt_t it reuses keeps stale
  // coefficients only at indices every encode overwrites (SURVEY hazard 11), so a zeroed buffer is equivalent.
  const size_t ptw = (size_t)h.WC * h.N;
  std::vector<uint64_t> host((size_t)S * ptw, 0);
  std::vector<double> diag(2 * S), rd(2 * S);
  for (unsigned i = 0; i < n2; i++) {
    const int shift = (int)(i * n1);
    for (unsigned j = 0; j < n1; j++) {
      const unsigned idx = shift + j;
      for (unsigned r = 0; r < S; r++) {             // zrotdiag, he-algo.c:29-43
        const size_t src = ((size_t)(r % S) * S + (idx + r) % S) * 2;
        diag[2 * r] = A[src]; diag[2 * r + 1] = A[src + 1];
      }
      for (unsigned r = 0; r < S; r++) {
        int ri = ((int)r - shift) % (int)S;
        if (ri < 0) ri += S;
        rd[2 * r] = diag[2 * ri]; rd[2 * r + 1] = diag[2 * ri + 1];
      }
      hb_ring_encode(ctx->ring, S, rd.data(), Delta, host.data() + (size_t)idx * ptw, h.WC);
    }
  }
  for (unsigned i = 0; i < n2; i++) {
    bool any = false;
    const uint64_t *w = host.data() + (size_t)i * n1 * ptw;
    for (size_t t = 0; t < (size_t)n1 * ptw && !any; t++) any = w[t] != 0;
    if (any) plan.giant_nz.push_back(i);
  }
  {
    // May the n1 products of a giant step be summed before reconstruction?  Every |ct coefficient| <= q_l/2, so a
    // coefficient of sum_j rot_j * pt_ij is bounded by 2^(k-1) * sum_j ||pt_ij||_1; it must stay below P_D/2 (> 2^(nbits-2)).
    long double worst = 0;
    for (unsigned i = 0; i < n2; i++) {
      long double s1 = 0;
      for (unsigned j = 0; j < n1; j++) {
        const uint64_t *w = host.data() + (size_t)(i * n1 + j) * ptw;
        for (unsigned c = 0; c < h.N; c++) {
          long double v = 0;
          const bool neg = (int64_t)w[(size_t)(h.WC - 1) * h.N + c] < 0;
          for (int ww = (int)h.WC - 1; ww >= 0; ww--) {
            uint64_t x = w[(size_t)ww * h.N + c];
            v = v * 18446744073709551616.0L + (long double)(neg ? ~x : x);
          }
          s1 += neg ? v + 1 : v;
        }
      }
      if (s1 > worst) worst = s1;
    }
    const long double lhs = (worst > 0 ? std::log2(worst) : 0) + (long double)(h.logq_l[l] - 1) + 0.01L;
    plan.fused = lhs < (long double)h.Pfull[plan.D].bits() - 2;
  }
  uint64_t *dpt = (uint64_t *)hb_dev_alloc(ctx, host.size() * 8);
  CUDA_OK(cudaMemcpyAsync(dpt, host.data(), host.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
  uint64_t *raw = (uint64_t *)hb_dev_alloc(ctx, (size_t)S * plan.D * h.N * 8);
  plan.pthat = (ulonglong2 *)hb_dev_alloc(ctx, (size_t)S * plan.D * h.N * 16);
  do_poly_ntt(ctx, dpt, ptw, h.WC, raw, plan.D, S);
  do_shoup_prep(ctx, raw, plan.pthat, plan.D, (size_t)S * plan.D, 2, plan.D);
  uint32_t *raw32 = nullptr;
  if (ctx->have32 && plan.fused) {
    // the exact product fits P_D (fused bound): any basis with Q >= P_D reconstructs the same integer
    plan.D32 = hb::limbs32_for_bits(ctx->ht32, h.Pfull[plan.D].bits());
    if (plan.D32) {
      raw32 = (uint32_t *)hb_dev_alloc(ctx, (size_t)S * plan.D32 * h.N * 4);
      plan.pthat32 = (uint32_t *)hb_dev_alloc(ctx, (size_t)S * plan.D32 * h.N * 4 * HB32_PT_PLANES);
      do32_poly_ntt(ctx, dpt, ptw, (int)h.WC, raw32, plan.D32, S, true);
      const size_t total = (size_t)S * plan.D32 * h.N;
      launch(ctx, "k32_shoup_prep", hb32::k32_shoup_prep, (unsigned)((total + 255) / 256), 256, 0, 1, (const uint32_t *)raw32, plan.pthat32, ctx->dt32,
             plan.D32, (size_t)S * plan.D32, plan.D32, (unsigned)HB32_PT_PLANES);
      plan.use32 = true;
    }
  }
  CUDA_OK(cudaStreamSynchronize(ctx->stream));
  CUDA_OK(cudaFree(dpt));
  CUDA_OK(cudaFree(raw));
  if (raw32) CUDA_OK(cudaFree(raw32));
  ctx->plans.push_back(std::move(plan));
  return ctx->plans.back();
}

// bookkeeping fields of he_gemv's result (level l -> l - 1) exactly as the reference's calls update them, for an input with (nu, B):
// he_mulpt, he_add over the n1 baby steps, he_rot, he_add over the n2 giant steps, he_rs (he-algo.c:57-90, precomp.c:417)
static void gemv_meta(const hb_ctx *ctx, double nu_in, double B_in, unsigned n1, unsigned n2, double *nu_out, double *B_out) {
  const hb::HostTables &h = ctx->ht;
  const double Delta = std::ldexp(1.0, (int)h.logDelta);
  double inu = 0, iB = 0, onu = 0, oB = 0;
  for (unsigned i = 0; i < n2; i++) {
    for (unsigned j = 0; j < n1; j++) {
      const double nu = nu_in * Delta, B = B_in * Delta;
      if (!j) { inu = nu; iB = B; }
      else { inu = inu >= nu ? inu : nu; iB = iB + B; }
    }
    if (!i) { onu = inu; oB = iB; }
    else { onu = onu >= inu ? onu : inu; oB = oB + iB; }
  }
  const double nn = h.N, Brs = std::sqrt(nn / 3.) * (3 + 8 * std::sqrt(64.0));
  *nu_out = onu / Delta;
  *B_out = oB / Delta + Brs;
}

// The reference schedule, one he_* call after the other (he-algo.c:57-87).
static void gemv_stepwise(hb_ctx *ctx, hb_ct *dst, const GemvPlan &plan, const hb_ct *ct, const hb_evk *rk, unsigned n1, unsigned n2) {
  const hb::HostTables &h = ctx->ht;
  const double Delta = std::ldexp(1.0, (int)h.logDelta);
  hb_ct inner, outer, rot;
  ct_alloc_tmp(ctx, &inner, ct->batch);
  ct_alloc_tmp(ctx, &outer, ct->batch);
  ct_alloc_tmp(ctx, &rot, ct->batch);
  for (unsigned i = 0; i < n2; i++) {
    const unsigned shift = i * n1;
    for (unsigned j = 0; j < n1; j++) {
      hb_ct_copy(ctx, &rot, ct);
      hb_he_rot(ctx, &rot, (int)j, rk);                          // key-switches even for j == 0 (he-algo.c:68)
      mulpt_with_hat(ctx, &rot, &rot, plan.pthat + (size_t)(shift + j) * plan.D * h.N, plan.D, Delta);
      if (!j) hb_ct_copy(ctx, &inner, &rot);
      else hb_he_add(ctx, &inner, &inner, &rot);
    }
    hb_he_rot(ctx, &inner, (int)shift, rk);
    if (!i) hb_ct_copy(ctx, &outer, &inner);
    else hb_he_add(ctx, &outer, &outer, &inner);
  }
  hb_ct_copy(ctx, dst, &outer);
  hb_he_rs(ctx, dst);
  hb_ct_free(ctx, &inner);
  hb_ct_free(ctx, &outer);
  hb_ct_free(ctx, &rot);
}

// perm_g[i] = brv(((2 brv(i) + 1) g mod 2N - 1) / 2): where ntt(sigma_g a) reads ntt(a) (k_ks_baby)
static const uint16_t *perm_table(hb_ctx *ctx, unsigned g) {
  auto it = ctx->perms.find(g);
  if (it != ctx->perms.end()) return it->second;
  const unsigned N = ctx->ht.N, logn = ctx->ht.logn;
  std::vector<uint16_t> h(N);
  auto brv = [&](unsigned x) { unsigned r = 0; for (unsigned b = 0; b < logn; b++) r |= ((x >> b) & 1u) << (logn - 1 - b); return r; };
  for (unsigned i = 0; i < N; i++) {
    const unsigned e = (unsigned)(((uint64_t)(2 * brv(i) + 1) * g) & (2 * N - 1));
    h[i] = (uint16_t)brv((e - 1) / 2);
  }
  uint16_t *d = (uint16_t *)hb_dev_alloc(ctx, (size_t)N * 2);
  CUDA_OK(cudaMemcpyAsync(d, h.data(), (size_t)N * 2, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_OK(cudaStreamSynchronize(ctx->stream));
  ctx->perms[g] = d;
  return d;
}

// Same ciphertext, fewer transforms.  The reference recomputes he_rot(ct, j) for every giant step i although it does not
// depend on i (he-algo.c:66-68): here the n1 baby rotations are key-switched once -- from ONE decomposition and forward
// transform of ct.c1 (k_ks_baby) --, transformed once (hat_j = ntt(decompose(rot_j))), and each giant step sums its n1
// diagonal products in the NTT domain before one inverse transform and one reconstruction (k_mulacc_intt).  Every
// he_swk rounding of the reference is still performed on the same integers, so the result is bit-identical.
// planB != nullptr: TWO matrix-vector products in one pass -- the first half of the batch (instances < batch / 2) is multiplied by
// plan's matrix, the second half by planB's (ctr_hempc's Ka (xhat - xr) and Kb (uhat - ur), src/hempc.c:257-259).  Only the
// diagonal products know which matrix an instance belongs to; rotations, key switches and the rescale are the same launches
// over twice the batch.  The caller sets the bookkeeping fields of the two halves (gemv_meta); dst->nu / dst->B are plan's.
static void gemv_fused(hb_ctx *ctx, hb_ct *dst, const GemvPlan &plan, const hb_ct *ct, const hb_evk *rk, unsigned n1, unsigned n2,
                       const GemvPlan *planB = nullptr) {
  const hb::HostTables &h = ctx->ht;
  if (planB && !(plan.use32 && planB->use32 && plan.D == planB->D && plan.D32 == planB->D32 && ct->batch % 2 == 0))
    hb_die("internal: paired matvec needs two plans on the 30-bit basis with equal limb counts");
  const double Delta = std::ldexp(1.0, (int)h.logDelta);
  const unsigned batch = ct->batch, D = plan.D, l = ct->l;
  const int W = hb::words_of_level(h, l), k = h.logq_l[l];
  const size_t N = h.N, polyw = (size_t)W * N, bstride = 2 * polyw;
  const size_t jstride = (size_t)batch * 2 * D * N;                 // one rotated batch in transform form
  const unsigned Dk = hb::dim_keyswitch(h, l);
  if (Dk > h.dimevk || Dk > h.dimub) hb_die("hb_he_gemv: needs %u limbs but keys hold %u and the prime chain %u", Dk, h.dimevk, h.dimub);
  const size_t cstride = (size_t)batch * 2 * Dk * N;                // one batch of key-switch limbs
  const bool grouped = !ctx->no_hoist && N <= 65536 && n1 <= HB_MAX_GROUPS && n2 <= HB_MAX_GROUPS;
  // 30-bit path: D32 limbs of 4 bytes per coefficient reuse the 59-bit workspaces; D32 may exceed 2 D (D = 1: two primes
  // below 2^30 do not cover a 60-bit P_1), so both buffers take the larger of the two layouts
  const size_t jstride32 = plan.use32 ? (size_t)batch * 2 * plan.D32 * N : 0;
  const size_t hat_bytes = std::max((size_t)n1 * jstride * 8, (size_t)n1 * jstride32 * 4);
  uint64_t *hat = ws_get(ctx, ctx->wsC, ctx->wsC_bytes, hat_bytes);
  if (planB && !grouped) hb_die("internal: paired matvec outside the grouped schedule");
  if (!grouped) {
    // one launch per rotation and per giant step (A/B timing, very large rings)
    hb_ct inner, outer, rot;
    ct_alloc_tmp(ctx, &inner, batch);
    ct_alloc_tmp(ctx, &outer, batch);
    ct_alloc_tmp(ctx, &rot, batch);
    for (unsigned j = 0; j < n1; j++) {
      hb_ct_copy(ctx, &rot, ct);
      hb_he_rot(ctx, &rot, (int)j, rk);
      do_poly_ntt(ctx, rot.words, polyw, W, hat + (size_t)j * jstride, D, (size_t)batch * 2);
    }
    for (unsigned i = 0; i < n2; i++) {
      const unsigned shift = i * n1;
      uint64_t *chat = ws_get(ctx, ctx->wsA, ctx->wsA_bytes, jstride * 8);
      hbdev::GroupMap G0;
      memset(&G0, 0, sizeof G0);
      do_mulacc_intt(ctx, hat, jstride, plan.pthat + (size_t)shift * D * N, 0, G0, n1, 1, chat, D, batch);
      do_recon_smod(ctx, chat, D, inner.words, W, k, (size_t)batch * 2, true);
      inner.l = l;
      for (unsigned j = 0; j < n1; j++) {          // metadata exactly as he_mulpt / he_add update it
        const double nu = ct->nu * Delta, B = ct->B * Delta;
        if (!j) { inner.nu = nu; inner.B = B; }
        else { inner.nu = inner.nu >= nu ? inner.nu : nu; inner.B = inner.B + B; }
      }
      hb_he_rot(ctx, &inner, (int)shift, rk);
      if (!i) hb_ct_copy(ctx, &outer, &inner);
      else hb_he_add(ctx, &outer, &outer, &inner);
    }
    hb_ct_copy(ctx, dst, &outer);
    hb_he_rs(ctx, dst);
    hb_ct_free(ctx, &inner);
    hb_ct_free(ctx, &outer);
    hb_ct_free(ctx, &rot);
    return;
  }
  // ---- grouped schedule: every phase of the matvec is ONE launch over (rotations or giant steps) x batch ----
  const unsigned ng = n1 > n2 ? n1 : n2;
  uint64_t *chat = ws_get(ctx, ctx->wsA, ctx->wsA_bytes, std::max((size_t)ng * cstride * 8, (size_t)n2 * jstride32 * 4));
  uint64_t *cts = ws_get(ctx, ctx->wsD, ctx->wsD_bytes, (size_t)(ng + n2) * batch * bstride * 8);   // rots / inner sums, then switched sums
  // 1. baby rotations: one decomposition + forward transform of ct.c1 per limb serves up to four key switches
  hbdev::TailGroups TB;
  memset(&TB, 0, sizeof TB);
  TB.gbatch = batch; TB.add_shared = 1;
  // Hybrid basis (hb_dev32.cuh): the limbs that do not divide P move to 30-bit primes when every key involved has its 30-bit
  // form and the product provably fits:  |x| < 2^(logn + k - 1 + keybits) <= M / 8.
  const unsigned wk = (unsigned)(k + 63) / 64;
  auto hybrid_m = [&](unsigned lg_terms, const std::vector<unsigned> &rots) -> unsigned {
    if (ctx->no_hyb || !ctx->hyb_mmax || W > 2 || wk > 2 || ctx->exact_tail || h.dim >= Dk) return 0;
    unsigned kb = 0;
    for (unsigned r : rots) {
      auto ib = ctx->evk_bits.find(rk[r].p);
      if (ib == ctx->evk_bits.end() || ib->second == ~0u || !ctx->evk32.count(rk[r].p)) return 0;
      kb = std::max(kb, ib->second);
    }
    const unsigned m = hyb_limbs_for(ctx, lg_terms + h.logn + (unsigned)(k - 1) + kb) ? ctx->hyb_mmax : 0;   // one limb count per context
    if (!m || !do_ks_finish_hyb(ctx, nullptr, nullptr, nullptr, nullptr, nullptr, 0, 0, hbdev::TailGroups(), nullptr, 0, 0, 0, h.dim, m)) return 0;
    if ((size_t)h.dim * 8 + (size_t)m * 4 > (size_t)Dk * 8) return 0;          // the shared key-switch workspace holds Dk words per coefficient
    return m;
  };
  std::vector<unsigned> baby_rots;
  for (unsigned j = 0; j < n1; j++) {
    if (!rk[j].p) hb_die("hb_he_gemv: rotation key %u is not loaded", j);
    baby_rots.push_back(j);
    TB.ginv[j] = pow5_inv_mod_2n(ctx, j);
  }
  const unsigned mb = hybrid_m(0, baby_rots);
  const hb_ctx::HybBasis *HB = mb ? &hyb_basis(ctx, mb) : nullptr;
  const size_t cstride59 = (size_t)batch * 2 * h.dim * N, jstride32h = (size_t)batch * 2 * mb * N;
  uint32_t *chat32b = (uint32_t *)(chat + (size_t)n1 * cstride59);             // after the n1 batches of P-limbs
  for (unsigned j0 = 0; j0 < n1; j0 += 4) {
    hbdev::BabySet S;
    hb32::Baby32 S32;
    memset(&S, 0, sizeof S);
    memset(&S32, 0, sizeof S32);
    S.nj = S32.nj = n1 - j0 < 4 ? n1 - j0 : 4;
    for (unsigned t = 0; t < S.nj; t++) {
      const unsigned j = j0 + t;
      S.evk[t] = (const ulonglong2 *)rk[j].p;
      S.perm[t] = S32.perm[t] = perm_table(ctx, inv_odd_mod_2n(ctx, TB.ginv[j]));
      if (mb) S32.key[t] = ctx->evk32[rk[j].p];
    }
    if (mb) {
      do_ks_baby(ctx, ct->words + polyw, bstride, W, S, chat + (size_t)j0 * cstride59, cstride59, Dk, batch, h.dim, HB->dec59);
      do32_ks_baby(ctx, ct->words + polyw, bstride, W, S32, mb, HB->dec32, chat32b + (size_t)j0 * jstride32h, jstride32h, batch);
    } else {
      do_ks_baby(ctx, ct->words + polyw, bstride, W, S, chat + (size_t)j0 * cstride, cstride, Dk, batch);
    }
  }
  //    tails of all n1 rotations: c0 = rdiv(.) + sigma_j(ct.c0), c1 = rdiv(.)   (he-automorphism.c:68-76)
  if (mb) do_ks_finish_hyb(ctx, chat, chat32b, &HB->F, ct->words, nullptr, bstride, W, TB, cts, W, k, n1 * batch, h.dim, mb);
  else do_ks_finish_g(ctx, chat, Dk, ct->words, nullptr, bstride, W, TB, cts, W, k, n1 * batch);
  // giant steps actually evaluated: all of them, or only those with a non-zero diagonal (GemvPlan::giant_nz)
  hbdev::GroupMap GM;
  memset(&GM, 0, sizeof GM);
  unsigned ngs = 0;
  {
    std::vector<bool> need(n2, false);                       // a giant step is skipped only if it is zero for every matrix involved
    for (unsigned i : plan.giant_nz) need[i] = true;
    if (planB) for (unsigned i : planB->giant_nz) need[i] = true;
    const bool any = std::find(need.begin(), need.end(), true) != need.end();
    for (unsigned i = 0; i < n2; i++)
      if (ctx->no_skip || !any || need[i]) GM.idx[ngs++] = i;
  }
  uint64_t *inner = cts;                                           // [ngs][batch][2][W][N]
  if (plan.use32) {
    // 2-3 on the internal 30-bit basis (twice the butterfly throughput per modulus bit; same integers, hb_tables32.hpp)
    const unsigned D32 = plan.D32;
    uint32_t *hat32 = (uint32_t *)hat, *mchat32 = (uint32_t *)chat;       // sized above for either layout
    do32_poly_ntt(ctx, cts, polyw, W, hat32, D32, (size_t)n1 * batch * 2, false);
    do32_mulacc_intt(ctx, hat32, jstride32, plan.pthat32, planB ? planB->pthat32 : nullptr, (size_t)n1 * D32, GM, n1, ngs, mchat32, D32, batch);
    do32_recon_smod(ctx, mchat32, D32, inner, W, k, (size_t)ngs * batch * 2);
  } else {
    // 2. transform the rotated ciphertexts once
    do_poly_ntt(ctx, cts, polyw, W, hat, D, (size_t)n1 * batch * 2);
    // 3. all giant steps: inner_i = sum_j rot_j * diag_{i n1 + j}, summed in the transform domain
    uint64_t *mchat = chat;                                          // [ngs][batch][2][D][N], D <= Dk
    do_mulacc_intt(ctx, hat, jstride, plan.pthat, (size_t)n1 * D * N, GM, n1, ngs, mchat, D, batch);
    do_recon_smod(ctx, mchat, D, inner, W, k, (size_t)ngs * batch * 2, true);
  }
  // 4. key-switch every inner sum with its own rotation i n1
  hbdev::KsGroups SG;
  hbdev::TailGroups TG;
  hbdev::SumGroups SS;
  memset(&SG, 0, sizeof SG);
  memset(&TG, 0, sizeof TG);
  memset(&SS, 0, sizeof SS);
  SG.gbatch = TG.gbatch = batch;
  SS.ng = ngs;
  unsigned keybits = 0;
  for (unsigned s = 0; s < ngs; s++) {
    const unsigned shift = GM.idx[s] * n1;
    if (shift >= h.slots) hb_die("hb_he_gemv: rotation %u outside [0,%u)", shift, h.slots);
    if (!rk[shift].p) hb_die("hb_he_gemv: rotation key %u is not loaded", shift);
    SS.ginv[s] = TG.ginv[s] = SG.ginv[s] = pow5_inv_mod_2n(ctx, shift);
    SG.g[s] = inv_odd_mod_2n(ctx, SG.ginv[s]);
    SG.evk[s] = (const ulonglong2 *)rk[shift].p;
    auto it = ctx->evk_bits.find(rk[shift].p);
    keybits = std::max(keybits, it == ctx->evk_bits.end() ? ~0u : it->second);
  }
  // metadata: the reference's own sequence of updates over ALL n1 x n2 steps (he_mulpt, he_add, he_rot, he_add, he_rs)
  dst->l = l;
  {
    double inu = 0, iB = 0, onu = 0, oB = 0;
    for (unsigned i = 0; i < n2; i++) {
      for (unsigned j = 0; j < n1; j++) {
        const double nu = ct->nu * Delta, B = ct->B * Delta;
        if (!j) { inu = nu; iB = B; }
        else { inu = inu >= nu ? inu : nu; iB = iB + B; }
      }
      if (!i) { onu = inu; oB = iB; }
      else { onu = onu >= inu ? onu : inu; oB = oB + iB; }
    }
    dst->nu = onu; dst->B = oB;
  }
  if (l == 0) hb_die("hb_he_rs: ciphertext is already at level 0");
  const double nn = h.N, Brs = std::sqrt(nn / 3.) * (3 + 8 * std::sqrt(64.0));   // precomp.c:417, as hb_he_rs
  const int Wout = hb::words_of_level(h, l - 1), kout = h.logq_l[l - 1];
  // 4'. summed giant steps (k_ks_limb_sum / k_ks_finish_sum): the giant-step key switches, their he_add chain and he_rs share one
  //     reconstruction.  Needs every giant key to be a real switching key (|coefficient| < 2^keybits from hb_evk_upload), so that
  //     X = sum_i x_i provably stays inside (-P_D/4, P_D/4):  log2 ngs + log2 N + (k - 1) + keybits < nbits(P_D) - 2.
  {
    unsigned lg = 0;
    while ((1u << lg) < ngs) lg++;
    const bool fits = keybits != ~0u && (uint64_t)lg + h.logn + (unsigned)(k - 1) + keybits + 2 < h.Pfull[Dk].bits();
    const unsigned wk = (unsigned)(k + 63) / 64;
    if (!ctx->no_sum && fits && h.logn <= 13 && ngs <= 8 && h.dim < Dk && ks_fast_ok(ctx, Dk, wk) && W <= 2 && Wout <= 2 &&
        do_ks_finish_sum(ctx, nullptr, 0, nullptr, Dk, nullptr, 0, 0, 0, SS, nullptr, 0, 0, 0, 0, 0, true)) {
      const unsigned dim = h.dim, Dx = Dk - dim;
      const size_t pstride = (size_t)batch * 2 * dim * N;          // limbs of the primes of P, per giant step
      uint64_t *chatP = chat, *chatX = chat + (size_t)ngs * pstride;   // fits: ngs (dim + Dx) <= ng Dk
      if ((size_t)ngs * pstride + (size_t)batch * 2 * Dx * N > (size_t)ng * cstride) hb_die("internal: key-switch workspace too small");
      std::vector<unsigned> giant_rots;
      for (unsigned s2 = 0; s2 < ngs; s2++) giant_rots.push_back(GM.idx[s2] * n1);
      const unsigned mg = hybrid_m(lg, giant_rots);
      dst->nu /= Delta;
      dst->B = dst->B / Delta + Brs;
      dst->l = l - 1;
      if (mg) {
        const hb_ctx::HybBasis &HG = hyb_basis(ctx, mg);
        hb32::Groups32 S32;
        memset(&S32, 0, sizeof S32);
        for (unsigned s2 = 0; s2 < ngs; s2++) { S32.key[s2] = ctx->evk32[rk[giant_rots[s2]].p]; S32.g[s2] = SG.g[s2]; }
        do_ks_limb_g(ctx, inner + polyw, bstride, W, SG, h.dimevk, chatP, Dk, ngs * batch, 0, dim, HG.dec59);
        do32_ks_limb_sum(ctx, inner + polyw, bstride, (size_t)batch * bstride, W, S32, ngs, mg, HG.dec32, (uint32_t *)chatX, batch);
        do_ks_finish_sum_hyb(ctx, chatP, pstride, (const uint32_t *)chatX, &HG.F, inner, (size_t)batch * bstride, bstride, W, SS, dst->words, Wout, k,
                             (int)h.logDelta, kout, batch, dim, mg);
        return;
      }
      do_ks_limb_g(ctx, inner + polyw, bstride, W, SG, h.dimevk, chatP, Dk, ngs * batch, 0, dim);
      do_ks_limb_sum(ctx, inner + polyw, bstride, (size_t)batch * bstride, W, SG, ngs, h.dimevk, chatX, Dk, batch, dim, Dx);
      do_ks_finish_sum(ctx, chatP, pstride, chatX, Dk, inner, (size_t)batch * bstride, bstride, W, SS, dst->words, Wout, k, (int)h.logDelta, kout,
                       batch, false);
      return;
    }
  }
  do_ks_limb_g(ctx, inner + polyw, bstride, W, SG, h.dimevk, chat, Dk, ngs * batch);
  uint64_t *switched = cts + (size_t)ng * batch * bstride;          // [ngs][batch][2][W][N]
  do_ks_finish_g(ctx, chat, Dk, inner, nullptr, bstride, W, TG, switched, W, k, ngs * batch);
  // 5. outer = sum_i, then he_rs: one pass (the centred sum never leaves the registers)
  if (W > 16) {                                                    // wider than the fused kernel's registers: two passes
    const size_t npoly = (size_t)batch * 2, total = npoly * N;
    launch(ctx, "k_ct_sum", hbdev::k_ct_sum, (unsigned)((total + 255) / 256), 256, 0, 1, (const uint64_t *)switched, (size_t)batch * bstride, ngs,
           dst->words, W, k, (unsigned)N, npoly);
    hb_he_rs(ctx, dst);
    return;
  }
  {
    dst->nu /= Delta;
    dst->B = dst->B / Delta + Brs;
    dst->l = l - 1;
    const size_t npoly = (size_t)batch * 2, total = npoly * N;
    launch(ctx, "k_ct_sum_rs", hbdev::k_ct_sum_rs, (unsigned)((total + 255) / 256), 256, 0, 1, (const uint64_t *)switched, (size_t)batch * bstride, ngs,
           dst->words, W, k, Wout, (int)h.logDelta, kout, (unsigned)N, npoly);
  }
}

extern "C" void hb_he_gemv(hb_ctx *ctx, hb_ct *dst, const double *A, const hb_ct *ct, const hb_evk *rk) { use_device(ctx);
  same_shape(dst, ct, "hb_he_gemv");
  const unsigned S = ctx->ht.slots;
  const GemvPlan &plan = gemv_plan(ctx, A, ct->l);
  unsigned n1 = (unsigned)std::sqrt((double)S);
  if (S != n1 * n1) n1 = (unsigned)std::sqrt((double)(2 * S));
  const unsigned n2 = S / n1;
  if (plan.fused && !ctx->stepwise) gemv_fused(ctx, dst, plan, ct, rk, n1, n2);
  else gemv_stepwise(ctx, dst, plan, ct, rk, n1, n2);
}

extern "C" void hb_ctr_hempc_eval(hb_ctx *ctx, hb_ct *up, const double *Ka, const double *Kb, const hb_ct *xhat,
                                  const hb_ct *xr, const hb_ct *uhat, const hb_ct *ur, const hb_evk *rk) { use_device(ctx);
  same_shape(xhat, xr, "hb_ctr_hempc_eval"); same_shape(uhat, ur, "hb_ctr_hempc_eval");
  same_shape(xhat, uhat, "hb_ctr_hempc_eval"); same_shape(up, xhat, "hb_ctr_hempc_eval");
  const hb::HostTables &h = ctx->ht;
  const unsigned batch = xhat->batch, S = h.slots;
  if (xhat->l != uhat->l || xhat->l == 0) hb_die("hb_ctr_hempc_eval: operands are at different levels");
  const unsigned l = xhat->l;
  const int W = hb::words_of_level(h, l), Wout = hb::words_of_level(h, l - 1), kout = h.logq_l[l - 1];
  const size_t N = h.N;
  // Both differences and both matrix-vector products live in ONE batch of 2 x batch instances: first half x, second half u.
  hb_ct diff2, out2;
  ct_alloc_tmp(ctx, &diff2, 2 * batch);
  ct_alloc_tmp(ctx, &out2, 2 * batch);
  hb_ct xdiff = diff2, udiff = diff2, KaX = out2, KbU = out2;
  xdiff.batch = udiff.batch = KaX.batch = KbU.batch = batch;
  udiff.words = diff2.words + (size_t)batch * 2 * W * N;
  KbU.words = out2.words + (size_t)batch * 2 * Wout * N;
  hb_he_sub(ctx, &xdiff, xhat, xr);              // hempc.c:253
  hb_he_sub(ctx, &udiff, uhat, ur);              // hempc.c:255
  unsigned n1 = (unsigned)std::sqrt((double)S);
  if (S != n1 * n1) n1 = (unsigned)std::sqrt((double)(2 * S));
  const unsigned n2 = S / n1;
  gemv_plan(ctx, Ka, l);
  gemv_plan(ctx, Kb, l);
  const GemvPlan *pa = &gemv_plan(ctx, Ka, l);    // Ka's plan may have been evicted by Kb's: this lookup settles the cache ...
  const GemvPlan *pb = &gemv_plan(ctx, Kb, l);    // ... and this one is a hit that moves nothing
  hb_ct sepA, sepB;
  sepA.words = sepB.words = nullptr;
  const bool pair = !ctx->no_pair && !ctx->stepwise && !ctx->no_hoist && pa->fused && pb->fused && pa->use32 && pb->use32 && pa->D == pb->D &&
                    pa->D32 == pb->D32 && n1 <= HB_MAX_GROUPS && n2 <= HB_MAX_GROUPS;
  if (pair) {
    diff2.l = l; diff2.nu = xdiff.nu; diff2.B = xdiff.B;
    gemv_fused(ctx, &out2, *pa, &diff2, rk, n1, n2, pb);      // hempc.c:257 and :259 in one pass
    KaX.l = KbU.l = out2.l;
    gemv_meta(ctx, xdiff.nu, xdiff.B, n1, n2, &KaX.nu, &KaX.B);
    gemv_meta(ctx, udiff.nu, udiff.B, n1, n2, &KbU.nu, &KbU.B);
  } else {
    // two passes into buffers of their own: he_rs inside the call-by-call schedule swaps its handle's buffer
    ct_alloc_tmp(ctx, &sepA, batch);
    ct_alloc_tmp(ctx, &sepB, batch);
    hb_he_gemv(ctx, &sepA, Ka, &xdiff, rk);      // hempc.c:257
    hb_he_gemv(ctx, &sepB, Kb, &udiff, rk);      // hempc.c:259
    KaX = sepA; KbU = sepB;
  }
  // hempc.c:261-266: he_add(du, KaX, KbU); he_neg(du); he_copy_ct(ucopy, uhat); he_moddown(ucopy); he_add(up, ucopy, du) -- all
  // of it arithmetic mod q_{l-1} followed by centring, so one pass computes smod(uhat - KaX - KbU, q_{l-1}); the bookkeeping
  // fields follow the five calls
  if (KaX.l != KbU.l || KaX.l != l - 1) hb_die("hb_ctr_hempc_eval: operands are at different levels");
  {
    const double du_nu = KaX.nu >= KbU.nu ? KaX.nu : KbU.nu, du_B = KaX.B + KbU.B;
    up->l = KaX.l;
    up->nu = uhat->nu >= du_nu ? uhat->nu : du_nu;
    up->B = uhat->B + du_B;
    const size_t npoly = (size_t)batch * 2, total = npoly * h.N;
    launch(ctx, "k_ct_combine", hbdev::k_ct_combine, (unsigned)((total + 255) / 256), 256, 0, 1, (const uint64_t *)uhat->words, W,
           (const uint64_t *)KaX.words, (const uint64_t *)KbU.words, up->words, Wout, kout, h.N, npoly);
  }
  hb_ct_free(ctx, &diff2);
  hb_ct_free(ctx, &out2);
  if (sepA.words) { hb_ct_free(ctx, &sepA); hb_ct_free(ctx, &sepB); }
}

// End to end from host buffers.  The batch is cut into chunks that flow through three streams -- upload (copy stream),
// evaluate (the context's stream), download (second copy stream) -- with two sets of device buffers, so the PCIe
// transfers of chunk c+1 and c-1 overlap the evaluation of chunk c.  HB_E2E_CHUNK sets the chunk size (instances).
extern "C" void hb_ctr_hempc_eval_host(hb_ctx *ctx, unsigned batch, unsigned l, const double *Ka, const double *Kb,
                                       const uint64_t *in_xhat, const uint64_t *in_xr, const uint64_t *in_uhat,
                                       const uint64_t *in_ur, const hb_evk *rk, uint64_t *out_up) { use_device(ctx);
  check_level(ctx, l);
  if (l == 0) hb_die("hb_ctr_hempc_eval_host: inputs must be above level 0");
  if (batch == 0) return;
  const double Delta = std::ldexp(1.0, (int)ctx->ht.logDelta);
  unsigned chunk = 512;
  if (batch < 4 * chunk) chunk = (batch + 7) / 8 > 64 ? (batch + 7) / 8 : 64;     // small batches (a shard of 8 GPUs) still pipeline
  if (const char *e = getenv("HB_E2E_CHUNK")) { const long v = atol(e); if (v > 0) chunk = (unsigned)v; }
  if (chunk > batch) chunk = batch;
  const uint64_t *host_in[4] = {in_xhat, in_xr, in_uhat, in_ur};
  const size_t in_words = (size_t)2 * hb::words_of_level(ctx->ht, l) * ctx->ht.N;         // per instance
  const size_t out_words = (size_t)2 * hb::words_of_level(ctx->ht, l - 1) * ctx->ht.N;
  pipeline_streams(ctx);
  const unsigned nsets = batch > chunk ? 2 : 1;
  hb_ct in[2][4], up[2];
  for (unsigned s = 0; s < nsets; s++) {
    for (int q = 0; q < 4; q++) ct_alloc_tmp(ctx, &in[s][q], chunk);
    ct_alloc_tmp(ctx, &up[s], chunk);
  }
  // pool buffers may still be read by work queued earlier on the context's stream: order the upload stream after it
  CUDA_OK(cudaEventRecord(ctx->ev_comp[0], ctx->stream));
  CUDA_OK(cudaStreamWaitEvent(ctx->up_stream, ctx->ev_comp[0], 0));
  // the first chunk is a quarter of the others: nothing overlaps its upload, so the evaluation starts sooner
  std::vector<unsigned> starts;
  for (unsigned b0 = 0, first = 1; b0 < batch; first = 0) {
    starts.push_back(b0);
    b0 += (first && batch > chunk && chunk >= 4) ? chunk / 4 : chunk;
  }
  starts.push_back(batch);
  const unsigned nchunks = (unsigned)starts.size() - 1;
  for (unsigned c = 0; c < nchunks; c++) {
    const unsigned s = c & (nsets - 1), b0 = starts[c], nb = (starts[c + 1] < batch ? starts[c + 1] : batch) - b0;
    if (c >= nsets) CUDA_OK(cudaStreamWaitEvent(ctx->up_stream, ctx->ev_comp[s], 0));      // chunk c-2 no longer reads in[s]
    for (int q = 0; q < 4; q++) {
      in[s][q].batch = nb; in[s][q].l = l; in[s][q].nu = Delta; in[s][q].B = 1.0;
      CUDA_OK(cudaMemcpyAsync(in[s][q].words, host_in[q] + (size_t)b0 * in_words, (size_t)nb * in_words * 8, cudaMemcpyHostToDevice,
                              ctx->up_stream));
    }
    CUDA_OK(cudaEventRecord(ctx->ev_up[s], ctx->up_stream));
    CUDA_OK(cudaStreamWaitEvent(ctx->stream, ctx->ev_up[s], 0));
    if (c >= nsets) CUDA_OK(cudaStreamWaitEvent(ctx->stream, ctx->ev_down[s], 0));         // chunk c-2's result has left up[s]
    up[s].batch = nb;
    hb_ctr_hempc_eval(ctx, &up[s], Ka, Kb, &in[s][0], &in[s][1], &in[s][2], &in[s][3], rk);
    CUDA_OK(cudaEventRecord(ctx->ev_comp[s], ctx->stream));
    CUDA_OK(cudaStreamWaitEvent(ctx->down_stream, ctx->ev_comp[s], 0));
    CUDA_OK(cudaMemcpyAsync(out_up + (size_t)b0 * out_words, up[s].words, (size_t)nb * out_words * 8, cudaMemcpyDeviceToHost, ctx->down_stream));
    CUDA_OK(cudaEventRecord(ctx->ev_down[s], ctx->down_stream));
  }
  CUDA_OK(cudaStreamSynchronize(ctx->down_stream));
  CUDA_OK(cudaStreamSynchronize(ctx->stream));
  for (unsigned s = 0; s < nsets; s++) {
    for (int q = 0; q < 4; q++) { in[s][q].batch = chunk; hb_ct_free(ctx, &in[s][q]); }
    up[s].batch = chunk;
    hb_ct_free(ctx, &up[s]);
  }
}

static void pipeline_streams(hb_ctx *ctx) {
  if (ctx->up_stream) return;
  CUDA_OK(cudaStreamCreateWithFlags(&ctx->up_stream, cudaStreamNonBlocking));
  CUDA_OK(cudaStreamCreateWithFlags(&ctx->down_stream, cudaStreamNonBlocking));
  for (int s = 0; s < 2; s++) {
    CUDA_OK(cudaEventCreateWithFlags(&ctx->ev_up[s], cudaEventDisableTiming));
    CUDA_OK(cudaEventCreateWithFlags(&ctx->ev_comp[s], cudaEventDisableTiming));
    CUDA_OK(cudaEventCreateWithFlags(&ctx->ev_down[s], cudaEventDisableTiming));
  }
}

// Streaming form of hb_ctr_hempc_eval_host for a server that evaluates batch after batch: enqueue the upload of this batch's
// four inputs (copy stream), its evaluation (the context's stream) and the download of its result (second copy stream), and
// return a ticket.  Two submissions may be in flight -- two persistent sets of device buffers --, so the transfers of batch
// k+1 and k-1 overlap the evaluation of batch k while every launch keeps the full batch size.  A third submission first
// waits for the first one.  hb_ctr_hempc_wait(ticket) returns once out_up of that submission is complete.
extern "C" uint64_t hb_ctr_hempc_submit_host(hb_ctx *ctx, unsigned batch, unsigned l, const double *Ka, const double *Kb,
                                             const uint64_t *in_xhat, const uint64_t *in_xr, const uint64_t *in_uhat,
                                             const uint64_t *in_ur, const hb_evk *rk, uint64_t *out_up) { use_device(ctx);
  check_level(ctx, l);
  if (l == 0) hb_die("hb_ctr_hempc_submit_host: inputs must be above level 0");
  if (batch == 0) hb_die("hb_ctr_hempc_submit_host: empty batch");
  pipeline_streams(ctx);
  const uint64_t ticket = ctx->job_next++;
  const unsigned s = (unsigned)(ticket & 1);
  if (!ctx->job_done[0])
    for (int q = 0; q < 2; q++) CUDA_OK(cudaEventCreateWithFlags(&ctx->job_done[q], cudaEventDisableTiming));
  if (ctx->job_busy[s]) { CUDA_OK(cudaEventSynchronize(ctx->job_done[s])); ctx->job_busy[s] = false; }
  if (batch > ctx->job_cap) {                       // (re)size both sets; nothing may be in flight
    for (int q = 0; q < 2; q++)
      if (ctx->job_busy[q]) { CUDA_OK(cudaEventSynchronize(ctx->job_done[q])); ctx->job_busy[q] = false; }
    CUDA_OK(cudaStreamSynchronize(ctx->stream));
    for (int t = 0; t < 2; t++) {
      if (ctx->job_cap) {
        for (int q = 0; q < 4; q++) { ctx->job_in[t][q].batch = ctx->job_cap; hb_ct_free(ctx, &ctx->job_in[t][q]); }
        ctx->job_up[t].batch = ctx->job_cap;
        hb_ct_free(ctx, &ctx->job_up[t]);
      }
      for (int q = 0; q < 4; q++) ct_alloc_tmp(ctx, &ctx->job_in[t][q], batch);
      ct_alloc_tmp(ctx, &ctx->job_up[t], batch);
    }
    ctx->job_cap = batch;
  }
  const double Delta = std::ldexp(1.0, (int)ctx->ht.logDelta);
  const uint64_t *host_in[4] = {in_xhat, in_xr, in_uhat, in_ur};
  const size_t in_words = (size_t)2 * hb::words_of_level(ctx->ht, l) * ctx->ht.N;
  const size_t out_words = (size_t)2 * hb::words_of_level(ctx->ht, l - 1) * ctx->ht.N;
  // set s was last read by the evaluation of ticket - 2 and its result left job_up[s] before job_done[s] fired (waited above)
  for (int q = 0; q < 4; q++) {
    hb_ct &c = ctx->job_in[s][q];
    c.batch = batch; c.l = l; c.nu = Delta; c.B = 1.0;
    CUDA_OK(cudaMemcpyAsync(c.words, host_in[q], (size_t)batch * in_words * 8, cudaMemcpyHostToDevice, ctx->up_stream));
  }
  CUDA_OK(cudaEventRecord(ctx->ev_up[s], ctx->up_stream));
  CUDA_OK(cudaStreamWaitEvent(ctx->stream, ctx->ev_up[s], 0));
  ctx->job_up[s].batch = batch;
  hb_ctr_hempc_eval(ctx, &ctx->job_up[s], Ka, Kb, &ctx->job_in[s][0], &ctx->job_in[s][1], &ctx->job_in[s][2], &ctx->job_in[s][3], rk);
  CUDA_OK(cudaEventRecord(ctx->ev_comp[s], ctx->stream));
  CUDA_OK(cudaStreamWaitEvent(ctx->down_stream, ctx->ev_comp[s], 0));
  CUDA_OK(cudaMemcpyAsync(out_up, ctx->job_up[s].words, (size_t)batch * out_words * 8, cudaMemcpyDeviceToHost, ctx->down_stream));
  CUDA_OK(cudaEventRecord(ctx->job_done[s], ctx->down_stream));
  ctx->job_busy[s] = true;
  return ticket;
}

extern "C" void hb_ctr_hempc_wait(hb_ctx *ctx, uint64_t ticket) { use_device(ctx);
  if (ticket >= ctx->job_next) hb_die("hb_ctr_hempc_wait: ticket %llu was never issued", (unsigned long long)ticket);
  if (ticket + 2 < ctx->job_next) return;           // its buffer set has been reused since: that submission had to finish first
  const unsigned s = (unsigned)(ticket & 1);
  if (ctx->job_busy[s]) { CUDA_OK(cudaEventSynchronize(ctx->job_done[s])); ctx->job_busy[s] = false; }
}

// ---------------------------------------------------------------------------------------------------------
// client side on the device: he_enc_pk / he_dec (he-encrypt.c:37-74,107-123)
// ---------------------------------------------------------------------------------------------------------
static void do_recon_add(hb_ctx *c, const uint64_t *chat, unsigned D, uint64_t *out, int Wout, int k, size_t npoly, unsigned H,
                         const uint64_t *addw0, const uint64_t *addw1, size_t addw_stride, int Waddw, const int64_t *adds0,
                         const int64_t *adds1) {
  const unsigned threads = 128;
  const size_t total = npoly * c->ht.N;
  const unsigned grid = (unsigned)((total + threads - 1) / threads);
  switch (wk_class((k + 63) / 64)) {
#define X(WKC) case WKC: launch(c, "k_recon_add", hbdev::k_recon_add<WKC>, grid, threads, 0, 1, chat, D, out, Wout, k, c->dt, (unsigned)npoly, H, addw0, addw1, addw_stride, Waddw, adds0, adds1); break;
    X(1) X(2) X(4) X(8) X(16)
#undef X
  }
}

extern "C" void hb_pk_upload(hb_ctx *ctx, hb_evk *pk, const uint64_t *p0_words, const uint64_t *p1_words) { use_device(ctx);
  const hb::HostTables &h = ctx->ht;
  const size_t polyw = (size_t)h.WC * h.N, half = (size_t)h.dim * h.N;
  uint64_t *words = (uint64_t *)hb_dev_alloc(ctx, 2 * polyw * 8);
  uint64_t *raw = (uint64_t *)hb_dev_alloc(ctx, 2 * half * 8);
  pk->p = (uint64_t *)hb_dev_alloc(ctx, 2 * half * 16);
  CUDA_OK(cudaMemcpyAsync(words, p0_words, polyw * 8, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_OK(cudaMemcpyAsync(words + polyw, p1_words, polyw * 8, cudaMemcpyHostToDevice, ctx->stream));
  do_poly_ntt(ctx, words, polyw, (int)h.WC, raw, h.dim, 2);                          // ntt(pk.p_h mod p_d), d < hectx.dim
  do_shoup_prep(ctx, raw, (ulonglong2 *)pk->p, h.dim, (size_t)2 * h.dim, 1, 0);     // times N^-1, like an evaluation key
  hb_dev_free(ctx, words);
  hb_dev_free(ctx, raw);
}

extern "C" void hb_sk_upload(hb_ctx *ctx, hb_sk *sk, const int64_t *s) { use_device(ctx);
  const hb::HostTables &h = ctx->ht;
  memset(sk, 0, sizeof *sk);
  uint64_t *words = (uint64_t *)hb_dev_alloc(ctx, (size_t)h.N * 8);
  sk->hat = (uint64_t *)hb_dev_alloc(ctx, (size_t)h.dimub * h.N * 8);
  CUDA_OK(cudaMemcpyAsync(words, s, (size_t)h.N * 8, cudaMemcpyHostToDevice, ctx->stream));
  do_poly_ntt(ctx, words, h.N, 1, sk->hat, h.dimub, 1);
  hb_dev_free(ctx, words);
}
extern "C" void hb_sk_free(hb_ctx *ctx, hb_sk *sk) { use_device(ctx);
  hb_dev_free(ctx, sk->hat);
  for (uint64_t *&p : sk->scaled) { hb_dev_free(ctx, p); p = nullptr; }
  sk->hat = nullptr;
}

extern "C" void hb_he_enc_pk(hb_ctx *ctx, hb_ct *ct, const uint64_t *pt_words, double pt_nu, const hb_evk *pk, const int64_t *v,
                             const int64_t *e0, const int64_t *e1) { use_device(ctx);
  need_ct(ct, "hb_he_enc_pk");
  if (!pk || !pk->p) hb_die("hb_he_enc_pk: public key is not loaded");
  const hb::HostTables &h = ctx->ht;
  const unsigned batch = ct->batch, D = h.dim, L = h.L;
  const size_t N = h.N, ptw = (size_t)h.WC * N;
  // staging: v, e0, e1 [batch][N] and the plaintexts [batch][WC][N]
  uint64_t *stage = ws_get(ctx, ctx->wsB, ctx->wsB_bytes, ((size_t)batch * (3 * N + ptw)) * 8);
  int64_t *dv = (int64_t *)stage, *de0 = dv + (size_t)batch * N, *de1 = de0 + (size_t)batch * N;
  uint64_t *dpt = (uint64_t *)(de1 + (size_t)batch * N);
  CUDA_OK(cudaMemcpyAsync(dv, v, (size_t)batch * N * 8, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_OK(cudaMemcpyAsync(de0, e0, (size_t)batch * N * 8, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_OK(cudaMemcpyAsync(de1, e1, (size_t)batch * N * 8, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_OK(cudaMemcpyAsync(dpt, pt_words, (size_t)batch * ptw * 8, cudaMemcpyHostToDevice, ctx->stream));
  uint64_t *chat = ws_get(ctx, ctx->wsA, ctx->wsA_bytes, (size_t)batch * 2 * D * N * 8);
  // both products v pk.p0, v pk.p1 share the transform of v: the key-switch limb kernel with the public key as "key"
  do_ks_limb(ctx, (const uint64_t *)dv, N, 1, 1u, 1u, (const ulonglong2 *)pk->p, D, chat, D, batch);
  ct->l = L;
  const double Delta = std::ldexp(1.0, (int)h.logDelta);
  ct->nu = pt_nu >= Delta ? pt_nu : Delta;
  {
    const double sigma = 3.1915382432114616, hw = 64, n = h.N;     // precomp.c:411-416
    ct->B = 8 * std::sqrt(2.0) * sigma * n + 6 * sigma * std::sqrt(n) + 16 * sigma * std::sqrt(hw * n);
  }
  do_recon_add(ctx, chat, D, ct->words, (int)h.WC, (int)h.logq_l[L], (size_t)batch * 2, 2, dpt, nullptr, ptw, (int)h.WC, de0, de1);
  CUDA_OK(cudaStreamSynchronize(ctx->stream));      // the host arrays may be reused by the caller
}

extern "C" void hb_he_dec(hb_ctx *ctx, uint64_t *pt_words, const hb_ct *ct, const hb_sk *sk) { use_device(ctx);
  need_ct(ct, "hb_he_dec");
  if (!sk || !sk->hat) hb_die("hb_he_dec: secret key is not loaded");
  const hb::HostTables &h = ctx->ht;
  const unsigned batch = ct->batch, l = ct->l;
  const unsigned D = (h.logq_l[l] + 1) / 59 + 1;                  // he-encrypt.c:115
  if (D > h.dimub) hb_die("hb_he_dec: limb count %u outside the prime chain (%u)", D, h.dimub);
  const int W = hb::words_of_level(h, l), k = h.logq_l[l];
  const size_t N = h.N;
  hb_sk *msk = const_cast<hb_sk *>(sk);
  ulonglong2 *pthat;
  uint64_t *tmp_scaled = nullptr;
  if (D < 64 && msk->scaled[D]) pthat = (ulonglong2 *)msk->scaled[D];
  else {
    pthat = (ulonglong2 *)hb_dev_alloc(ctx, (size_t)D * N * 16);
    do_shoup_prep(ctx, sk->hat, pthat, D, D, 2, D);
    if (D < 64) msk->scaled[D] = (uint64_t *)pthat; else tmp_scaled = (uint64_t *)pthat;
  }
  uint64_t *chat = ws_get(ctx, ctx->wsA, ctx->wsA_bytes, (size_t)batch * D * N * 8);
  uint64_t *out = ws_get(ctx, ctx->wsB, ctx->wsB_bytes, (size_t)batch * W * N * 8);
  do_mulpt_limb(ctx, ct->words + (size_t)W * N, (size_t)2 * W * N, W, pthat, chat, D, batch);      // c1 sk
  do_recon_add(ctx, chat, D, out, W, k, batch, 1, ct->words, nullptr, (size_t)2 * W * N, W, nullptr, nullptr);   // + c0
  CUDA_OK(cudaMemcpyAsync(pt_words, out, (size_t)batch * W * N * 8, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_OK(cudaStreamSynchronize(ctx->stream));
  if (tmp_scaled) hb_dev_free(ctx, tmp_scaled);
}

// ---------------------------------------------------------------------------------------------------------
// he_genswk on the device: he-kem.c:74-119.  Sampling stays with the caller.
// ---------------------------------------------------------------------------------------------------------
extern "C" unsigned hb_genswk_words(const hb_ctx *ctx) { return (ctx->ht.nbitsPqL + 1 + 63) / 64; }

extern "C" void hb_he_genswk(hb_ctx *ctx, unsigned nkeys, const uint64_t *a_words, const int64_t *e, const int64_t *sp, const hb_sk *sk,
                             uint64_t *out_p0, uint64_t *out_p1) { use_device(ctx);
  if (!sk || !sk->hat) hb_die("hb_he_genswk: secret key is not loaded");
  if (!nkeys) return;
  const hb::HostTables &h = ctx->ht;
  const size_t N = h.N;
  const hb::BigU &P = h.Pfull[h.dim];
  const hb::BigU M = P.shl(h.logq_l[h.L]);                          // P q_L
  const unsigned D = (M.bits() + h.logn) / 59 + 1;                  // he-kem.c:83
  if (D > h.dimub) hb_die("hb_he_genswk: needs %u primes but the chain has %u", D, h.dimub);
  const hb::BigU &PD = h.Pfull[D];
  const unsigned WA = hb_genswk_words(ctx), WX = (PD.bits() + 72 + 63) / 64;
  if (WX > (unsigned)hbkg::MAXW || WA > WX) hb_die("hb_he_genswk: %u-word intermediates exceed the built maximum %d", WX, hbkg::MAXW);
  if (M.bits() + 64 * hbkg::TOPW < 64 * WX + 128) hb_die("hb_he_genswk: P q_L is too narrow for the quotient estimate window");
  const unsigned Dmax = D > h.dimevk ? D : h.dimevk;
  // ---- constants (host big integers), one upload ----
  std::vector<uint64_t> hc;
  auto put = [&](const std::vector<uint64_t> &v) { const size_t off = hc.size(); hc.insert(hc.end(), v.begin(), v.end()); return off; };
  std::vector<uint64_t> primes(Dmax), inv(D), phat((size_t)D * WX);
  for (unsigned d = 0; d < Dmax; d++) primes[d] = h.primes[d].p;
  for (unsigned d = 0; d < D; d++) {
    const hb::BigU ph = PD.divmod_u64(primes[d], nullptr);
    inv[d] = hb::powmod(ph.mod_u64(primes[d]), primes[d] - 2, primes[d]);
    const std::vector<uint64_t> lo = ph.low(WX);
    std::copy(lo.begin(), lo.end(), phat.begin() + (size_t)d * WX);
  }
  const std::vector<uint64_t> Mw = M.low(WX), Mhw = M.shr1().low(WX);
  std::vector<uint64_t> Mlo(WX);                                    // floor(M/2) - M in two's complement
  {
    uint64_t br = 0;
    for (unsigned w = 0; w < WX; w++) {
      const uint64_t x = Mhw[w], y = Mw[w];
      Mlo[w] = x - y - br;
      br = (x < y) || (x == y && br);
    }
  }
  const size_t o_pr = put(primes), o_inv = put(inv), o_ph = put(phat), o_PD = put(PD.low(WX)), o_PDh = put(PD.shr1().low(WX)),
               o_M = put(Mw), o_Mh = put(Mhw), o_Mlo = put(Mlo), o_P = put(P.low(WX));
  uint64_t *dc = (uint64_t *)hb_dev_alloc(ctx, hc.size() * 8);
  CUDA_OK(cudaMemcpyAsync(dc, hc.data(), hc.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
  hbkg::Consts C{dc + o_pr, dc + o_inv, dc + o_ph, dc + o_PD, dc + o_PDh, dc + o_M, dc + o_Mh, dc + o_Mlo, dc + o_P, D, WX, WA, (unsigned)N};
  // ---- inputs ----
  const size_t aw = (size_t)nkeys * WA * N, kn = (size_t)nkeys * N;
  uint64_t *da = (uint64_t *)hb_dev_alloc(ctx, aw * 8);
  int64_t *de = (int64_t *)hb_dev_alloc(ctx, kn * 8), *dsp = (int64_t *)hb_dev_alloc(ctx, kn * 8);
  CUDA_OK(cudaMemcpyAsync(da, a_words, aw * 8, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_OK(cudaMemcpyAsync(de, e, kn * 8, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_OK(cudaMemcpyAsync(dsp, sp, kn * 8, cudaMemcpyHostToDevice, ctx->stream));
  // ---- a*sk on D primes: decompose, transform, pointwise, inverse transform (poly_mul, poly.c:84-107) ----
  const size_t limbsK = (size_t)nkeys * D * N;
  uint64_t *t0 = (uint64_t *)hb_dev_alloc(ctx, limbsK * 8), *t1 = (uint64_t *)hb_dev_alloc(ctx, limbsK * 8);
  launch(ctx, "k_decompose_wide", hbkg::k_decompose_wide, (unsigned)((limbsK + 255) / 256), 256, 0, 1, (const uint64_t *)da, (int)WA, (size_t)nkeys,
         C.primes, D, t0, (unsigned)N);
  do_ntt_fwd(ctx, t0, t1, (size_t)nkeys * D, 0, D, false);
  launch(ctx, "k_pointmul", hbkg::k_pointmul, (unsigned)((limbsK + 255) / 256), 256, 0, 1, (const uint64_t *)t1, (const uint64_t *)sk->hat, C.primes, D, t0,
         (unsigned)N, (size_t)nkeys);
  do_ntt_inv(ctx, t0, t1, (size_t)nkeys * D, 0, D);
  // ---- exact product, -x + e + P sp, centred mod P q_L; a centred mod P q_L ----
  uint64_t *dv = (uint64_t *)hb_dev_alloc(ctx, 2 * aw * 8), *dac = dv + aw;
  launch(ctx, "k_genswk_tail", hbkg::k_tail, (unsigned)((kn + 127) / 128), 128, 0, 1, (const uint64_t *)t1, C, (const int64_t *)de, (const int64_t *)dsp,
         (const uint64_t *)da, dv, dac, nkeys);
  // ---- both halves onto the dimevk key primes, transformed in the reference's [0,p] representation (he-kem.c:103-110) ----
  const size_t limbsE = (size_t)2 * nkeys * h.dimevk * N;
  uint64_t *r0 = (uint64_t *)hb_dev_alloc(ctx, limbsE * 8), *r1 = (uint64_t *)hb_dev_alloc(ctx, limbsE * 8);
  launch(ctx, "k_decompose_wide", hbkg::k_decompose_wide, (unsigned)((limbsE + 255) / 256), 256, 0, 1, (const uint64_t *)dv, (int)WA, (size_t)2 * nkeys,
         C.primes, h.dimevk, r0, (unsigned)N);
  do_ntt_fwd(ctx, r0, r1, (size_t)2 * nkeys * h.dimevk, 0, h.dimevk, true);
  CUDA_OK(cudaMemcpyAsync(out_p0, r1, limbsE * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_OK(cudaMemcpyAsync(out_p1, r1 + limbsE / 2, limbsE * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_OK(cudaStreamSynchronize(ctx->stream));
  for (void *q : {(void *)dc, (void *)da, (void *)de, (void *)dsp, (void *)t0, (void *)t1, (void *)dv, (void *)r0, (void *)r1}) hb_dev_free(ctx, q);
}

// ---------------------------------------------------------------------------------------------------------
// limb-level entry points
// ---------------------------------------------------------------------------------------------------------
static void check_primes(const hb_ctx *ctx, unsigned d0, unsigned D, const char *what) {
  if (D == 0 || d0 + D > ctx->ht.dimub) hb_die("%s: primes [%u,%u) outside the chain of %u", what, d0, d0 + D, ctx->ht.dimub);
}
extern "C" void hb_ntt_forward(hb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t nlimbs, unsigned d0, unsigned D, int strict) { use_device(ctx);
  check_primes(ctx, d0, D, "hb_ntt_forward");
  if (nlimbs) do_ntt_fwd(ctx, in, out, nlimbs, d0, D, strict != 0);
}
extern "C" void hb_ntt_inverse(hb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t nlimbs, unsigned d0, unsigned D) { use_device(ctx);
  check_primes(ctx, d0, D, "hb_ntt_inverse");
  if (nlimbs) do_ntt_inv(ctx, in, out, nlimbs, d0, D);
}
extern "C" void hb_ntt_forward_host(hb_ctx *ctx, const uint64_t *in, uint64_t *out, size_t nlimbs, unsigned d0, unsigned D, int strict) { use_device(ctx);
  check_primes(ctx, d0, D, "hb_ntt_forward_host");
  const size_t bytes = nlimbs * ctx->ht.N * 8;
  uint64_t *buf = ws_get(ctx, ctx->wsA, ctx->wsA_bytes, bytes);
  CUDA_OK(cudaMemcpyAsync(buf, in, bytes, cudaMemcpyHostToDevice, ctx->stream));
  if (nlimbs) do_ntt_fwd(ctx, buf, buf, nlimbs, d0, D, strict != 0);
  CUDA_OK(cudaMemcpyAsync(out, buf, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_OK(cudaStreamSynchronize(ctx->stream));
}
extern "C" void hb_poly_ntt(hb_ctx *ctx, const uint64_t *polys, unsigned W, size_t npoly, unsigned D, uint64_t *out) { use_device(ctx);
  check_primes(ctx, 0, D, "hb_poly_ntt");
  if (W == 0 || W > ctx->ht.WC) hb_die("hb_poly_ntt: %u words per coefficient outside [1,%u]", W, ctx->ht.WC);
  if (npoly) do_poly_ntt(ctx, polys, (size_t)W * ctx->ht.N, (int)W, out, D, npoly);
}
extern "C" void hb_rns2words(hb_ctx *ctx, const uint64_t *limbs, size_t npoly, unsigned D, unsigned k, unsigned Wout, uint64_t *out) { use_device(ctx);
  check_primes(ctx, 0, D, "hb_rns2words");
  if (k == 0 || k > 64 * ctx->ht.WC) hb_die("hb_rns2words: modulus 2^%u outside (0, 2^%u]", k, 64 * ctx->ht.WC);
  if (npoly) do_recon_smod(ctx, limbs, D, out, (int)Wout, (int)k, npoly, false);
}
extern "C" void hb_poly_rot(hb_ctx *ctx, const uint64_t *polys, unsigned W, size_t npoly, int rot, uint64_t *out) { use_device(ctx);
  const unsigned ginv = rot < 0 ? 2 * ctx->ht.N - 1 : pow5_inv_mod_2n(ctx, (unsigned)rot);
  const size_t total = npoly * ctx->ht.N;
  if (out == polys) hb_die("hb_poly_rot: the permutation is not in place (out must not alias polys)");
  if (total) launch(ctx, "k_poly_automorph", hbdev::k_poly_automorph, (unsigned)((total + 255) / 256), 256, 0, 1, polys, out, (int)W, ginv, ctx->ht.N, npoly);
}

// ---------------------------------------------------------------------------------------------------------
// hb_he_mulpt_static: he_mulpt for callers that may not allocate (CUDA graph capture, serving loops, harnesses
// that forbid cudaMalloc).  No counterpart in the reference.
//
// Same kernels as hb_he_mulpt (k_poly_ntt, k_shoup_prep, k_mulpt_limb, k_recon_smod); the three work buffers of
// the product live in ONE file-scope device array of HB_STATIC_SCRATCH_BYTES instead of the context's growing
// workspaces.  Every GPU that loads the library holds the array until the process exits.
//   words [0, hw)        raw   transform of the plaintext, D limbs          (written by k_poly_ntt)
//   words [hw, 3 hw)     pthat Shoup pairs of raw times N^-1 (P_D/p_d)^-1  (written by k_shoup_prep)
//   words [3 hw, ...)    chat  limbs of the 2 batch product polynomials     (written by k_mulpt_limb)
// hw = D N rounded up to 32 words, so every part keeps the 256-byte alignment of the array.  Each word that a
// kernel reads was written earlier in the same call, so nothing has to start at zero and the contents left by
// the previous call are never observed.  All calls on one GPU share the array: calls that overlap on the GPU
// (two streams, two contexts on the same device, two host threads) race.  Calls on one context's stream are
// ordered and safe.
//
// In host code `g_static_scratch` is only a handle for the runtime, not a device address: the address comes
// from cudaGetSymbolAddress, once per device, with that device current.
// ---------------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned long long g_static_scratch[HB_STATIC_SCRATCH_BYTES / 8];

static uint64_t *static_scratch(const hb_ctx *ctx) {
  enum { MAXDEV = 64 };
  static uint64_t *addr[MAXDEV];          // per device; filled on the first call with that device current
  const int dev = ctx->device;
  if (dev < 0 || dev >= MAXDEV) hb_die("hb_he_mulpt_static: device index %d outside [0,%d)", dev, (int)MAXDEV);
  if (!addr[dev]) {
    int cur = -1;
    CUDA_OK(cudaGetDevice(&cur));
    if (cur != dev) hb_die("hb_he_mulpt_static: device %d is current, the context lives on device %d", cur, dev);
    void *p = nullptr;
    CUDA_OK(cudaGetSymbolAddress(&p, g_static_scratch));
    if (!p || ((uintptr_t)p & 255)) hb_die("hb_he_mulpt_static: scratch array resolved to %p", p);
    addr[dev] = (uint64_t *)p;
  }
  return addr[dev];
}

extern "C" void hb_he_mulpt_static(hb_ctx *ctx, hb_ct *dst, const hb_ct *src, const hb_pt *pt) { use_device(ctx);
  same_shape(dst, src, "hb_he_mulpt_static");
  if (!pt || !pt->words) hb_die("hb_he_mulpt_static: plaintext handle is not allocated");
  if (ctx->prof) hb_die("hb_he_mulpt_static: per-kernel profiling creates events; switch it off (hb_prof_enable(ctx, 0))");
  const unsigned D = hb::dim_mulpt(ctx->ht, src->l, pt->nu);
  if (D == 0 || D > ctx->ht.dimub) hb_die("hb_he_mulpt_static: limb count %u outside the prime chain (%u)", D, ctx->ht.dimub);
  const size_t N = ctx->ht.N;
  const size_t hw = ((size_t)D * N + 31) & ~(size_t)31;
  const size_t chat_words = (size_t)src->batch * 2 * D * N;
  const size_t need = (3 * hw + chat_words) * 8;
  if (need > (size_t)HB_STATIC_SCRATCH_BYTES)
    hb_die("hb_he_mulpt_static: batch %u at level %u (N = %zu, %u limbs) needs %zu bytes of scratch, the static array holds %zu; "
           "use hb_he_mulpt or a smaller batch", src->batch, src->l, N, D, need, (size_t)HB_STATIC_SCRATCH_BYTES);
  uint64_t *raw = static_scratch(ctx);
  ulonglong2 *pthat = (ulonglong2 *)(raw + hw);
  uint64_t *chat = raw + 3 * hw;
  const int W = hb::words_of_level(ctx->ht, src->l), k = ctx->ht.logq_l[src->l];
  do_poly_ntt(ctx, pt->words, (size_t)ctx->ht.WC * N, ctx->ht.WC, raw, D, 1);
  do_shoup_prep(ctx, raw, pthat, D, D, 2, D);
  do_mulpt_limb(ctx, src->words, (size_t)W * N, W, pthat, chat, D, (size_t)src->batch * 2);
  const double nu = src->nu, B = src->B;
  dst->l = src->l;
  dst->nu = nu * pt->nu;
  dst->B = B * pt->nu;
  do_recon_smod(ctx, chat, D, dst->words, W, k, (size_t)src->batch * 2, true);
}
