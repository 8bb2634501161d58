/* TEST INFRASTRUCTURE ONLY.  Marshals libgcrypt MPIs (the reference's coefficient type,
 * GPQHE/src/poly.h:69-72) to and from the engine's planar little-endian two's-complement
 * word layout words[w*n + i] (w = 0 is the least-significant word).  Our own code; used by
 * tests/ and the cpu_baseline leg to move data in and out of oracle/_ref/libgpqhe.so. */
#include <stdint.h>
#include <gcrypt.h>

typedef struct gcry_mpi *MPI;

/* coeffs[i] -> words[w*n+i], W words per coefficient.  Returns 0, or 1+i if coeffs[i] does not fit. */
int ref_mpi_to_words(const MPI *coeffs, unsigned n, unsigned W, uint64_t *words)
{
  unsigned char buf[4096];
  for (unsigned i = 0; i < n; i++) {
    size_t nw = 0;
    if (gcry_mpi_print(GCRYMPI_FMT_USG, buf, sizeof buf, &nw, coeffs[i])) return 1 + (int)i;
    if (nw > (size_t)W * 8) return 1 + (int)i;
    uint64_t tmp[512];
    for (unsigned w = 0; w < W; w++) tmp[w] = 0;
    for (size_t b = 0; b < nw; b++) {           /* buf is big-endian magnitude */
      size_t pos = nw - 1 - b;                  /* byte significance */
      tmp[pos / 8] |= (uint64_t)buf[b] << (8 * (pos % 8));
    }
    if (gcry_mpi_is_neg(coeffs[i])) {           /* two's complement negate */
      unsigned carry = 1;
      for (unsigned w = 0; w < W; w++) {
        uint64_t v = ~tmp[w] + carry;
        carry = (carry && v == 0) ? 1 : 0;
        tmp[w] = v;
      }
    }
    for (unsigned w = 0; w < W; w++) words[(size_t)w * n + i] = tmp[w];
  }
  return 0;
}

/* words -> coeffs[i] (MPIs must already exist, as poly_mpi_alloc leaves them). */
void ref_words_to_mpi(MPI *coeffs, unsigned n, unsigned W, const uint64_t *words)
{
  unsigned char buf[4096];
  for (unsigned i = 0; i < n; i++) {
    uint64_t tmp[512];
    for (unsigned w = 0; w < W; w++) tmp[w] = words[(size_t)w * n + i];
    int neg = (int)(tmp[W - 1] >> 63);
    if (neg) {
      unsigned carry = 1;
      for (unsigned w = 0; w < W; w++) {
        uint64_t v = ~tmp[w] + carry;
        carry = (carry && v == 0) ? 1 : 0;
        tmp[w] = v;
      }
    }
    size_t nb = (size_t)W * 8;
    for (size_t b = 0; b < nb; b++) {
      size_t pos = nb - 1 - b;
      buf[b] = (unsigned char)(tmp[pos / 8] >> (8 * (pos % 8)));
    }
    MPI t = NULL;
    gcry_mpi_scan(&t, GCRYMPI_FMT_USG, buf, nb, NULL);
    if (neg) gcry_mpi_neg(t, t);
    gcry_mpi_set(coeffs[i], t);
    gcry_mpi_release(t);
  }
}

/* q = 2^logq as a fresh MPI (what src/ctr.c:514-516 builds with mpi_set_ui + mpi_lshift). */
MPI ref_pow2(unsigned logq)
{
  MPI q = gcry_mpi_set_ui(NULL, 1);
  gcry_mpi_lshift(q, q, logq);
  return q;
}
