/* gpqhe_b200_shim -- the reference-side binding of libhectr_b200.so.
 *
 * A maintainer of OChicken/HECTR links this ONE file (plus libhectr_b200.so) in front of libgpqhe.so:
 *
 *     cc tests/hectr.c -lgpqhe_b200 -lhectr -lgpqhe -lpmu -lgcrypt -lm          (or LD_PRELOAD=libgpqhe_b200.so)
 *
 * and the UNMODIFIED libhectr.so / test-hectr / test-gpqhe evaluate on the GPU: the file exports the evaluation half of
 * the GPQHE C API under the reference's own names, prototypes and struct layouts (GPQHE/src/gpqhe.h:84-97,99-175), so
 * ctr_hempc (src/hempc.c:253-266) and he_gemv's callers resolve to it.  Everything else (hectx_init's tables, key
 * generation, samplers, he_ecd/he_dcd, memory management) stays the reference's code in libgpqhe.so.
 *
 * It is compiled against the reference's headers where they lie (integration/build.sh, -I$REF/GPQHE/src) and is NOT part
 * of the product library: libhectr_b200.so has no libgcrypt dependency and knows nothing about MPIs.
 *
 *   reference symbol (file:line)                         forwarded to (include/hectr_b200.h)
 *   ---------------------------------------------------  ------------------------------------------
 *   hectx_init / hectx_exit  precomp.c:431-453,465-      reference's own, then hb_ctx_create / hb_ctx_destroy
 *   he_add, he_sub, he_neg   he-add.c:32-79,124-142      hb_he_add, hb_he_sub, hb_he_neg
 *   he_addpt, he_subpt       he-add.c:80-123             hb_he_addpt, hb_he_subpt
 *   he_rs, he_moddown        he-rescale.c:33-70          hb_he_rs, hb_he_moddown
 *   he_mulpt, he_mul         he-mult.c:88-196            hb_he_mulpt, hb_he_mul
 *   he_rot, he_conj          he-automorphism.c:87-115    hb_he_rot, hb_he_conj
 *   he_gemv                  he-algo.c:47-93             hb_he_gemv
 *   he_enc_pk, he_dec        he-encrypt.c:37-74,107-123  hb_he_enc_pk, hb_he_dec (the reference's samplers draw v, e0, e1
 *                                                        in the reference's order; HB_SHIM_CLIENT=0 leaves both to libgpqhe)
 *
 * Data: a he_ct_t holds N libgcrypt MPIs per polynomial (poly.h:69-72); the engine takes planar little-endian
 * two's-complement words [2][W_l][N].  Every call converts its operands, uploads, runs, downloads and converts back, so
 * the host structs stay the canonical objects the rest of the reference reads directly (ct->l, ct->nu, ct->B, coeffs).
 * Evaluation keys (he_evk_t: two flat uint64 arrays, gpqhe.h:78-82) are uploaded once and cached by address + fingerprint.
 *
 * Errors: like the reference -- message on stderr and abort() (precomp.c:344-350).  One context, one caller thread.
 */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "gpqhe.h"          /* the reference's own header: struct he_ct, he_pt, he_evk, he_pk, poly_ctx, he_ctx */
#include "hectr_b200.h"

extern struct poly_ctx polyctx;      /* precomp.c:41-47 */
extern struct he_ctx hectx;
extern void sample_error(poly_mpi_t *r);     /* sample.c:74-82   */
extern void sample_zo(poly_mpi_t *r);        /* sample.c:123-131 */

static hb_ctx *g_ctx;
static unsigned g_logn, g_logq, g_slots, g_logDelta, g_n, g_WC;
static int g_client = 1;
static uint64_t *g_stage;            /* pinned staging, [4][2][WC][N] words */
static uint64_t g_calls;             /* interposed evaluation calls (reported at hectx_exit when HB_SHIM_VERBOSE=1) */

static void shim_die(const char *what)
{
  fprintf(stderr, "\033[1m\033[31merror:\033[0m \033[1mgpqhe_b200_shim\033[0m. %s\n", what);
  abort();
}

/* ---- MPI <-> words -------------------------------------------------------------------------------------------- */
/* coeffs[i] -> words[w*n + i] (two's complement, W words); aborts when a coefficient does not fit */
static void mpi_to_words(MPI const *coeffs, unsigned n, unsigned W, uint64_t *words)
{
  unsigned char buf[2048];
  uint64_t tmp[256];
  if (W > 256) shim_die("coefficient wider than 256 words");
  for (unsigned i = 0; i < n; i++) {
    size_t nw = 0;
    if (gcry_mpi_print(GCRYMPI_FMT_USG, buf, sizeof buf, &nw, coeffs[i]) || nw > (size_t)W * 8)
      shim_die("ciphertext coefficient does not fit its level's word count");
    memset(tmp, 0, (size_t)W * 8);
    for (size_t b = 0; b < nw; b++) {                 /* big-endian magnitude */
      const size_t pos = nw - 1 - b;
      tmp[pos / 8] |= (uint64_t)buf[b] << (8 * (pos % 8));
    }
    if (gcry_mpi_is_neg(coeffs[i])) {
      unsigned carry = 1;
      for (unsigned w = 0; w < W; w++) { const uint64_t v = ~tmp[w] + carry; carry = (carry && v == 0); tmp[w] = v; }
    }
    for (unsigned w = 0; w < W; w++) words[(size_t)w * n + i] = tmp[w];
  }
}

static void words_to_mpi(MPI *coeffs, unsigned n, unsigned W, const uint64_t *words)
{
  unsigned char buf[2048];
  uint64_t tmp[256];
  for (unsigned i = 0; i < n; i++) {
    for (unsigned w = 0; w < W; w++) tmp[w] = words[(size_t)w * n + i];
    const int neg = (int)(tmp[W - 1] >> 63);
    if (neg) {
      unsigned carry = 1;
      for (unsigned w = 0; w < W; w++) { const uint64_t v = ~tmp[w] + carry; carry = (carry && v == 0); tmp[w] = v; }
    }
    const size_t nb = (size_t)W * 8;
    for (size_t b = 0; b < nb; b++) { const size_t pos = nb - 1 - b; buf[b] = (unsigned char)(tmp[pos / 8] >> (8 * (pos % 8))); }
    MPI t = NULL;
    gcry_mpi_scan(&t, GCRYMPI_FMT_USG, buf, nb, NULL);
    if (neg) gcry_mpi_neg(t, t);
    gcry_mpi_set(coeffs[i], t);
    gcry_mpi_release(t);
  }
}

/* ---- context: follows the reference's global one ------------------------------------------------------------------ */
static unsigned log2_exact(double v)
{
  int e = 0;
  const double m = frexp(v, &e);
  if (m != 0.5) shim_die("Delta is not a power of two (the engine takes q = 2^logq, Delta = 2^logDelta)");
  return (unsigned)(e - 1);
}

static void ctx_open(void)
{
  const char *e = getenv("HB_SHIM_CLIENT");
  g_client = !(e && *e == '0');
  g_logn = polyctx.logn;
  g_logq = gcry_mpi_get_nbits(hectx.q[hectx.L]) - 1;          /* q = 2^logq (src/ctr.c:514-516) */
  {
    MPI pw = gcry_mpi_set_ui(NULL, 1);
    gcry_mpi_lshift(pw, pw, g_logq);
    if (gcry_mpi_cmp(pw, hectx.q[hectx.L])) shim_die("q is not a power of two (the engine takes q = 2^logq)");
    gcry_mpi_release(pw);
  }
  g_slots = hectx.slots;
  g_logDelta = log2_exact(hectx.Delta);
  int dev = 0;
  if ((e = getenv("HB_SHIM_DEVICE"))) dev = atoi(e);
  g_ctx = hb_ctx_create(g_logn, g_logq, g_slots, g_logDelta, dev);
  hb_info info;
  hb_ctx_info(g_ctx, &info);
  if (info.L != hectx.L || info.dim != hectx.dim || info.dimevk != hectx.dimevk || info.dimub != polyctx.dimub)
    shim_die("engine context disagrees with hectx (L / dim / dimevk / dimub)");
  g_n = info.n;
  g_WC = info.words_top;
  g_stage = (uint64_t *)hb_host_alloc(g_ctx, (size_t)4 * 2 * g_WC * g_n * 8);
}

struct keyslot { const uint64_t *p0; uint64_t fp; hb_evk key; };
static struct keyslot *g_keys;
static size_t g_nkeys, g_capkeys;
struct pkslot { const MPI *p0; hb_evk key; int valid; } g_pk;
struct skslot { const MPI *s; hb_sk key; int valid; } g_sk;

static void ctx_close(void)
{
  if (!g_ctx) return;
  for (size_t i = 0; i < g_nkeys; i++) hb_evk_free(g_ctx, &g_keys[i].key);
  free(g_keys);
  g_keys = NULL; g_nkeys = g_capkeys = 0;
  if (g_pk.valid) { hb_evk_free(g_ctx, &g_pk.key); g_pk.valid = 0; }
  if (g_sk.valid) { hb_sk_free(g_ctx, &g_sk.key); g_sk.valid = 0; }
  hb_host_free(g_ctx, g_stage);
  g_stage = NULL;
  if (getenv("HB_SHIM_VERBOSE"))
    fprintf(stderr, "gpqhe_b200_shim: %llu evaluation calls forwarded, %llu GPU kernel launches\n", (unsigned long long)g_calls,
            (unsigned long long)hb_launch_count(g_ctx));
  hb_ctx_destroy(g_ctx);
  g_ctx = NULL;
}

void hectx_init(unsigned int logn, MPI q, unsigned int slots, uint64_t Delta)
{
  static void (*next)(unsigned int, MPI, unsigned int, uint64_t);
  if (!next) next = (void (*)(unsigned int, MPI, unsigned int, uint64_t))dlsym(RTLD_NEXT, "hectx_init");
  if (!next) shim_die("libgpqhe.so must follow this library in the link order (hectx_init not found)");
  ctx_close();
  next(logn, q, slots, Delta);
  ctx_open();
}

void hectx_exit(void)
{
  static void (*next)(void);
  if (!next) next = (void (*)(void))dlsym(RTLD_NEXT, "hectx_exit");
  ctx_close();
  if (next) next();
}

static void need_ctx(void)
{
  if (!g_ctx) {
    if (!polyctx.n) shim_die("hectx_init has not been called");
    ctx_open();            /* the library was preloaded after hectx_init ran */
  }
}

/* ---- evaluation keys: uploaded once, found again by address + fingerprint ------------------------------------------- */
static uint64_t fingerprint(const he_evk_t *k)
{
  const size_t half = (size_t)hectx.dimevk * g_n, step = half / 509 + 1;
  uint64_t h = 1469598103934665603ULL;
  for (size_t i = 0; i < half; i += step) {
    h = (h ^ k->p0.coeffs[i]) * 1099511628211ULL;
    h = (h ^ k->p1.coeffs[i]) * 1099511628211ULL;
  }
  return h;
}

static const hb_evk *key_for(const he_evk_t *k)
{
  if (!k || !k->p0.coeffs || !k->p1.coeffs) shim_die("evaluation key is not allocated");
  const uint64_t fp = fingerprint(k);
  for (size_t i = 0; i < g_nkeys; i++)
    if (g_keys[i].p0 == k->p0.coeffs) {
      if (g_keys[i].fp == fp) return &g_keys[i].key;
      hb_evk_free(g_ctx, &g_keys[i].key);                       /* regenerated in place */
      hb_evk_upload(g_ctx, &g_keys[i].key, k->p0.coeffs, k->p1.coeffs);
      g_keys[i].fp = fp;
      return &g_keys[i].key;
    }
  if (g_nkeys == g_capkeys) {
    g_capkeys = g_capkeys ? 2 * g_capkeys : 64;
    g_keys = (struct keyslot *)realloc(g_keys, g_capkeys * sizeof *g_keys);
    if (!g_keys) shim_die("out of memory");
  }
  struct keyslot *s = &g_keys[g_nkeys++];
  s->p0 = k->p0.coeffs; s->fp = fp;
  hb_evk_upload(g_ctx, &s->key, k->p0.coeffs, k->p1.coeffs);
  return &s->key;
}

/* the engine wants rk as an array indexed by rotation (like he_rot's rk[rot]): a scratch array of handles */
static hb_evk *g_rk;
static const hb_evk *keys_for_array(const he_evk_t *rk, unsigned count)
{
  g_rk = (hb_evk *)realloc(g_rk, (size_t)g_slots * sizeof *g_rk);
  memset(g_rk, 0, (size_t)g_slots * sizeof *g_rk);
  for (unsigned r = 0; r < count; r++) g_rk[r] = *key_for(&rk[r]);
  return g_rk;
}

/* ---- ciphertext marshalling ------------------------------------------------------------------------------------------ */
static void ct_up(hb_ct *d, const he_ct_t *s, unsigned slot)
{
  if (s->l > hectx.L) shim_die("ciphertext level out of range");
  const unsigned W = hb_ctx_words(g_ctx, s->l);
  uint64_t *st = g_stage + (size_t)slot * 2 * g_WC * g_n;
  mpi_to_words(s->c0.coeffs, g_n, W, st);
  mpi_to_words(s->c1.coeffs, g_n, W, st + (size_t)W * g_n);
  hb_ct_alloc(g_ctx, d, 1);
  hb_ct_upload(g_ctx, d, st, s->l, s->nu, s->B);
}

static void ct_down(he_ct_t *d, hb_ct *s)
{
  const unsigned W = hb_ctx_words(g_ctx, s->l);
  hb_ct_download(g_ctx, s, g_stage);                  /* synchronises */
  d->l = s->l; d->nu = s->nu; d->B = s->B;
  words_to_mpi(d->c0.coeffs, g_n, W, g_stage);
  words_to_mpi(d->c1.coeffs, g_n, W, g_stage + (size_t)W * g_n);
  hb_ct_free(g_ctx, s);
}

static void pt_up(hb_pt *d, const he_pt_t *s)
{
  mpi_to_words(s->m.coeffs, g_n, g_WC, g_stage + (size_t)3 * 2 * g_WC * g_n);
  hb_pt_alloc(g_ctx, d);
  hb_pt_upload(g_ctx, d, g_stage + (size_t)3 * 2 * g_WC * g_n, s->nu);
}

/* ---- he-add.c ------------------------------------------------------------------------------------------------------------ */
static void binary(he_ct_t *ct, const he_ct_t *a, const he_ct_t *b, int sub)
{
  need_ctx(); g_calls++;
  hb_ct x, y, r;
  ct_up(&x, a, 0); ct_up(&y, b, 1);
  hb_ct_alloc(g_ctx, &r, 1);
  if (sub) hb_he_sub(g_ctx, &r, &x, &y); else hb_he_add(g_ctx, &r, &x, &y);
  ct_down(ct, &r);
  hb_ct_free(g_ctx, &x); hb_ct_free(g_ctx, &y);
}
void he_add(he_ct_t *ct, const he_ct_t *ct1, const he_ct_t *ct2) { binary(ct, ct1, ct2, 0); }
void he_sub(he_ct_t *ct, const he_ct_t *ct1, const he_ct_t *ct2) { binary(ct, ct1, ct2, 1); }

void he_neg(he_ct_t *ct)
{
  need_ctx(); g_calls++;
  hb_ct x;
  ct_up(&x, ct, 0);
  hb_he_neg(g_ctx, &x);
  ct_down(ct, &x);
}

static void binary_pt(he_ct_t *dest, const he_ct_t *src, const he_pt_t *pt, int sub)
{
  need_ctx(); g_calls++;
  hb_ct x, r;
  hb_pt p;
  ct_up(&x, src, 0); pt_up(&p, pt);
  hb_ct_alloc(g_ctx, &r, 1);
  if (sub) hb_he_subpt(g_ctx, &r, &x, &p); else hb_he_addpt(g_ctx, &r, &x, &p);
  ct_down(dest, &r);
  hb_ct_free(g_ctx, &x); hb_pt_free(g_ctx, &p);
}
void he_addpt(he_ct_t *dest, const he_ct_t *src, const he_pt_t *pt) { binary_pt(dest, src, pt, 0); }
void he_subpt(he_ct_t *dest, const he_ct_t *src, const he_pt_t *pt) { binary_pt(dest, src, pt, 1); }

/* ---- he-rescale.c ------------------------------------------------------------------------------------------------------- */
void he_rs(struct he_ct *ct)
{
  need_ctx(); g_calls++;
  hb_ct x;
  ct_up(&x, ct, 0);
  hb_he_rs(g_ctx, &x);
  ct_down(ct, &x);
}
void he_moddown(he_ct_t *ct)
{
  need_ctx(); g_calls++;
  hb_ct x;
  ct_up(&x, ct, 0);
  hb_he_moddown(g_ctx, &x);
  ct_down(ct, &x);
}

/* ---- he-mult.c ------------------------------------------------------------------------------------------------------------ */
void he_mulpt(struct he_ct *dest, const struct he_ct *src, const struct he_pt *pt)
{
  need_ctx(); g_calls++;
  hb_ct x, r;
  hb_pt p;
  ct_up(&x, src, 0); pt_up(&p, pt);
  hb_ct_alloc(g_ctx, &r, 1);
  hb_he_mulpt(g_ctx, &r, &x, &p);
  ct_down(dest, &r);
  hb_ct_free(g_ctx, &x); hb_pt_free(g_ctx, &p);
}

void he_mul(he_ct_t *ct, const he_ct_t *ct1, const he_ct_t *ct2, const he_evk_t *rlk)
{
  need_ctx(); g_calls++;
  hb_ct x, y, r;
  ct_up(&x, ct1, 0); ct_up(&y, ct2, 1);
  hb_ct_alloc(g_ctx, &r, 1);
  hb_he_mul(g_ctx, &r, &x, &y, key_for(rlk));
  ct_down(ct, &r);
  hb_ct_free(g_ctx, &x); hb_ct_free(g_ctx, &y);
}

/* ---- he-automorphism.c -------------------------------------------------------------------------------------------------- */
void he_rot(he_ct_t *ct, const int rot, const he_evk_t *rk)
{
  need_ctx(); g_calls++;
  if (rot < 0 || (unsigned)rot >= g_slots) shim_die("he_rot: rotation outside [0, slots)");
  g_rk = (hb_evk *)realloc(g_rk, (size_t)g_slots * sizeof *g_rk);
  memset(g_rk, 0, (size_t)g_slots * sizeof *g_rk);
  g_rk[rot] = *key_for(&rk[rot]);                      /* he_rot reads rk[rot] only (he-automorphism.c:108) */
  hb_ct x;
  ct_up(&x, ct, 0);
  hb_he_rot(g_ctx, &x, rot, g_rk);
  ct_down(ct, &x);
}

void he_conj(he_ct_t *ct, const he_evk_t *ck)
{
  need_ctx(); g_calls++;
  hb_ct x;
  ct_up(&x, ct, 0);
  hb_he_conj(g_ctx, &x, key_for(ck));
  ct_down(ct, &x);
}

/* ---- he-algo.c ------------------------------------------------------------------------------------------------------------ */
void he_gemv(he_ct_t *ct_dest, const _Complex double *A, const he_ct_t *ct, const he_evk_t *rk)
{
  need_ctx(); g_calls++;
  const hb_evk *keys = keys_for_array(rk, g_slots);        /* he_genrk fills rk[0..slots) (he-kem.c:154-169) */
  hb_ct x, r;
  ct_up(&x, ct, 0);
  hb_ct_alloc(g_ctx, &r, 1);
  hb_he_gemv(g_ctx, &r, (const double *)A, &x, keys);     /* _Complex double == double[2] (C11 6.2.5p13) */
  ct_down(ct_dest, &r);
  hb_ct_free(g_ctx, &x);
}

/* ---- he-encrypt.c: client side on the device, sampling by the reference's own samplers in the reference's order ---------- */
static void (*next_enc_pk)(struct he_ct *, const struct he_pt *, const struct he_pk *);
static void (*next_dec)(struct he_pt *, const struct he_ct *, const poly_mpi_t *);

static void poly_to_i64(const poly_mpi_t *p, int64_t *out)
{
  mpi_to_words(p->coeffs, g_n, 1, (uint64_t *)out);
}

void he_enc_pk(struct he_ct *ct, const struct he_pt *pt, const struct he_pk *pk)
{
  need_ctx();
  if (!g_client) {
    if (!next_enc_pk) next_enc_pk = (void (*)(struct he_ct *, const struct he_pt *, const struct he_pk *))dlsym(RTLD_NEXT, "he_enc_pk");
    next_enc_pk(ct, pt, pk);
    return;
  }
  g_calls++;
  if (!g_pk.valid || g_pk.p0 != pk->p0.coeffs) {          /* one public key per context, as in src/ctr.c:519-526 */
    if (g_pk.valid) hb_evk_free(g_ctx, &g_pk.key);
    uint64_t *w = (uint64_t *)malloc((size_t)2 * g_WC * g_n * 8);
    mpi_to_words(pk->p0.coeffs, g_n, g_WC, w);
    mpi_to_words(pk->p1.coeffs, g_n, g_WC, w + (size_t)g_WC * g_n);
    hb_pk_upload(g_ctx, &g_pk.key, w, w + (size_t)g_WC * g_n);
    hb_sync(g_ctx);
    free(w);
    g_pk.p0 = pk->p0.coeffs; g_pk.valid = 1;
  }
  /* the reference's draws, in its order (he-encrypt.c:49-56): v, e0, e1 */
  poly_mpi_t v, e0, e1;
  poly_mpi_alloc(&v); poly_mpi_alloc(&e0); poly_mpi_alloc(&e1);
  sample_zo(&v); sample_error(&e0); sample_error(&e1);
  int64_t *s = (int64_t *)malloc((size_t)3 * g_n * 8);
  poly_to_i64(&v, s); poly_to_i64(&e0, s + g_n); poly_to_i64(&e1, s + 2 * (size_t)g_n);
  poly_mpi_free(&v); poly_mpi_free(&e0); poly_mpi_free(&e1);
  uint64_t *ptw = g_stage + (size_t)3 * 2 * g_WC * g_n;
  mpi_to_words(pt->m.coeffs, g_n, g_WC, ptw);
  hb_ct r;
  hb_ct_alloc(g_ctx, &r, 1);
  hb_he_enc_pk(g_ctx, &r, ptw, pt->nu, &g_pk.key, s, s + g_n, s + 2 * (size_t)g_n);
  free(s);
  ct_down(ct, &r);
  ct->B = hectx.bnd.Bclean;                              /* the reference's own constant (he-encrypt.c:42) */
}

void he_dec(struct he_pt *pt, const struct he_ct *ct, const poly_mpi_t *sk)
{
  need_ctx();
  if (!g_client) {
    if (!next_dec) next_dec = (void (*)(struct he_pt *, const struct he_ct *, const poly_mpi_t *))dlsym(RTLD_NEXT, "he_dec");
    next_dec(pt, ct, sk);
    return;
  }
  g_calls++;
  if (!g_sk.valid || g_sk.s != sk->coeffs) {
    if (g_sk.valid) hb_sk_free(g_ctx, &g_sk.key);
    int64_t *s = (int64_t *)malloc((size_t)g_n * 8);
    poly_to_i64(sk, s);
    hb_sk_upload(g_ctx, &g_sk.key, s);
    hb_sync(g_ctx);
    free(s);
    g_sk.s = sk->coeffs; g_sk.valid = 1;
  }
  hb_ct x;
  ct_up(&x, ct, 0);
  const unsigned W = hb_ctx_words(g_ctx, ct->l);
  hb_he_dec(g_ctx, g_stage, &x, &g_sk.key);              /* synchronises */
  pt->nu = ct->nu;                                       /* he-encrypt.c:110 */
  words_to_mpi(pt->m.coeffs, g_n, W, g_stage);
  hb_ct_free(g_ctx, &x);
}
