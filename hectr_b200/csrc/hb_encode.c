/* Host-side CKKS encode / decode for the engine (plain C so that double and long double
 * arithmetic follows the same C rules as the reference's host code).
 *
 * Restates, operation for operation in the same order (results must be bit-identical because
 * he_gemv encodes one diagonal per baby/giant step, GPQHE/src/he-algo.c:70-72):
 *   ring tables      GPQHE/src/precomp.c:295-312   cyc_group[i] = 5^i mod 2N, zeta_i = e^{2 pi i / 2N}
 *   special FFT      GPQHE/src/canemb.c:43-81      canemb / invcanemb over the 5^j orbit
 *   encode / decode  GPQHE/src/he-encode.c:53-75   round((long double)x*Delta) at stride gap
 *   double <-> int   GPQHE/src/types.c:77-106,225-245
 * Output integers are written as little-endian two's-complement words, planar [W][N].
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define HB_PI 3.141592653589793238462643383279502884

typedef struct hb_ring {
  unsigned n, m;
  unsigned *orbit;   /* [n/2] 5^i mod 2N */
  double *zre, *zim; /* [m+1] */
} hb_ring;

hb_ring *hb_ring_create(unsigned logn) {
  hb_ring *r = (hb_ring *)malloc(sizeof *r);
  r->n = 1u << logn;
  r->m = 2 * r->n;
  unsigned nh = r->n / 2;
  r->orbit = (unsigned *)malloc(sizeof(unsigned) * (nh ? nh : 1));
  r->orbit[0] = 1;
  for (unsigned i = 1; i < nh; i++) r->orbit[i] = (5u * r->orbit[i - 1]) % r->m;
  r->zre = (double *)malloc(sizeof(double) * (r->m + 1));
  r->zim = (double *)malloc(sizeof(double) * (r->m + 1));
  for (size_t i = 0; i < r->m; i++) {
    double theta = 2 * HB_PI * i / r->m;
    r->zre[i] = cos(theta);
    r->zim[i] = sin(theta);
  }
  r->zre[r->m] = r->zre[0];
  r->zim[r->m] = r->zim[0];
  return r;
}

void hb_ring_destroy(hb_ring *r) {
  if (!r) return;
  free(r->orbit); free(r->zre); free(r->zim); free(r);
}

static void swap_bitrev(double *re, double *im, unsigned n) {
  for (unsigned i = 1, j = 0; i < n; i++) {
    unsigned bit = n >> 1;
    while (j >= bit) { j -= bit; bit >>= 1; }
    j += bit;
    if (i < j) {
      double t = re[i]; re[i] = re[j]; re[j] = t;
      t = im[i]; im[i] = im[j]; im[j] = t;
    }
  }
}

/* slots -> coefficients direction (canemb.c:62-81) */
static void inv_embed(const hb_ring *r, double *re, double *im, unsigned slots) {
  for (unsigned len = slots; len >= 1; len >>= 1) {
    unsigned span = len << 2, step = r->m / span, mid = len >> 1;
    for (unsigned base = 0; base < slots; base += len) {
      for (unsigned j = 0; j < mid; j++) {
        unsigned k = (span - (r->orbit[j] % span)) * step;
        double ar = re[base + j], ai = im[base + j];
        double br = re[base + j + mid], bi = im[base + j + mid];
        double dr = ar - br, di = ai - bi;
        double wr = r->zre[k], wi = r->zim[k];
        re[base + j] = ar + br;
        im[base + j] = ai + bi;
        re[base + j + mid] = dr * wr - di * wi;
        im[base + j + mid] = dr * wi + di * wr;
      }
    }
  }
  swap_bitrev(re, im, slots);
  double s = (double)slots;
  for (unsigned i = 0; i < slots; i++) { re[i] /= s; im[i] /= s; }
}

/* coefficients -> slots direction (canemb.c:43-60) */
static void embed(const hb_ring *r, double *re, double *im, unsigned slots) {
  swap_bitrev(re, im, slots);
  for (unsigned len = 2; len <= slots; len <<= 1) {
    unsigned span = len << 2, step = r->m / span, mid = len >> 1;
    for (unsigned base = 0; base < slots; base += len) {
      for (unsigned j = 0; j < mid; j++) {
        unsigned k = (r->orbit[j] % span) * step;
        double ur = re[base + j], ui = im[base + j];
        double br = re[base + j + mid], bi = im[base + j + mid];
        double wr = r->zre[k], wi = r->zim[k];
        double vr = br * wr - bi * wi, vi = br * wi + bi * wr;
        re[base + j] = ur + vr;
        im[base + j] = ui + vi;
        re[base + j + mid] = ur - vr;
        im[base + j + mid] = ui - vi;
      }
    }
  }
}

/* types.c:225-245: a (an integer-valued long double) -> W-word two's complement at words[w*n + idx] */
static void store_rounded(uint64_t *words, unsigned n, unsigned W, unsigned idx, long double a) {
  int neg = !(a >= 0);
  if (neg) a = -a;
  long double u64max = UINT64_MAX;
  size_t count = (size_t)(a / u64max);
  a -= count * u64max;
  /* magnitude = (uint64_t)a + count*(2^64 - 1) */
  unsigned __int128 lowpart = (unsigned __int128)(uint64_t)a + ((unsigned __int128)count << 64) - count;
  uint64_t mag[2] = {(uint64_t)lowpart, (uint64_t)(lowpart >> 64)};
  uint64_t carry = 1;
  for (unsigned w = 0; w < W; w++) {
    uint64_t v = w < 2 ? mag[w] : 0;
    if (neg) { v = ~v + carry; carry = (carry && v == 0); }
    words[(size_t)w * n + idx] = v;
  }
}

/* types.c:77-106: |a| accumulated bit by bit in double, from the top bit down */
static double load_double(const uint64_t *words, unsigned n, unsigned W, unsigned idx) {
  uint64_t mag[64];
  if (W > 64) W = 64;
  int neg = (int)(words[(size_t)(W - 1) * n + idx] >> 63);
  uint64_t carry = 1;
  for (unsigned w = 0; w < W; w++) {
    uint64_t v = words[(size_t)w * n + idx];
    if (neg) { v = ~v + carry; carry = (carry && v == 0); }
    mag[w] = v;
  }
  int top = -1;
  for (int b = (int)W * 64 - 1; b >= 0; b--)
    if ((mag[b / 64] >> (b % 64)) & 1) { top = b; break; }
  if (top < 0) return 0;
  double num = 0;   /* the reference starts from test_bit(a, nbits) == 0 */
  for (int b = top; b >= 0; b--) {
    num *= 2.;
    num += (double)((mag[b / 64] >> (b % 64)) & 1);
  }
  return neg ? -num : num;
}

void hb_ring_encode(const hb_ring *r, unsigned slots, const double *m, double Delta, uint64_t *words, unsigned W) {
  double *re = (double *)malloc(sizeof(double) * slots), *im = (double *)malloc(sizeof(double) * slots);
  for (unsigned i = 0; i < slots; i++) { re[i] = m[2 * i]; im[i] = m[2 * i + 1]; }
  inv_embed(r, re, im, slots);
  unsigned nh = r->n / 2, gap = nh / slots;
  for (unsigned i = 0, j = 0; i < slots; i++, j += gap) {
    /* round() is the double function: the long double product is narrowed first (he-encode.c:61-62) */
    store_rounded(words, r->n, W, j, round((double)((long double)re[i] * Delta)));
    store_rounded(words, r->n, W, j + nh, round((double)((long double)im[i] * Delta)));
  }
  free(re); free(im);
}

void hb_ring_decode(const hb_ring *r, unsigned slots, const uint64_t *words, unsigned W, double Delta, double *m) {
  double *re = (double *)malloc(sizeof(double) * slots), *im = (double *)malloc(sizeof(double) * slots);
  unsigned nh = r->n / 2, gap = nh / slots;
  for (unsigned i = 0, j = 0; i < slots; i++, j += gap) {
    re[i] = load_double(words, r->n, W, j) / Delta;
    im[i] = load_double(words, r->n, W, j + nh) / Delta;
  }
  embed(r, re, im, slots);
  for (unsigned i = 0; i < slots; i++) { m[2 * i] = re[i]; m[2 * i + 1] = im[i]; }
  free(re); free(im);
}

/* he_const_pt (he-encode.c:119-125): only coefficients 0 and n/2 are written */
void hb_ring_const(const hb_ring *r, double re, double im, double Delta, uint64_t *words, unsigned W) {
  store_rounded(words, r->n, W, 0, round((double)((long double)re * Delta)));
  store_rounded(words, r->n, W, r->n / 2, round((double)((long double)im * Delta)));
}
