/* TEST INFRASTRUCTURE ONLY (oracle build).
 *
 * Declaration-only stand-in for libgcrypt's public <gcrypt.h>.  The reference
 * (GPQHE) includes <gcrypt.h> for its big-integer arithmetic, the image ships the
 * libgcrypt 1.10.3 *runtime* (libgcrypt.so.20) but not its headers.  This file
 * declares the handful of public-ABI functions GPQHE uses so the unmodified
 * reference sources compile where they lie; it contains no arithmetic and no
 * reference code.  Link with -l:libgcrypt.so.20.
 */
#ifndef HECTR_B200_ORACLE_GCRYPT_SHIM_H
#define HECTR_B200_ORACLE_GCRYPT_SHIM_H
#include <stddef.h>
#include <stdio.h>
#include <stdlib.h>
#include <stdarg.h>
#include <string.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef unsigned int gpg_error_t;
typedef gpg_error_t gcry_error_t;
#define GPG_ERR_NO_ERROR 0
const char *gpg_strerror(gpg_error_t err);
struct gcry_mpi;
typedef struct gcry_mpi *gcry_mpi_t;
struct gcry_sexp;
typedef struct gcry_sexp *gcry_sexp_t;
#define GCRYSEXP_FMT_ADVANCED 3
enum gcry_mpi_format { GCRYMPI_FMT_NONE = 0, GCRYMPI_FMT_STD = 1, GCRYMPI_FMT_PGP = 2,
                       GCRYMPI_FMT_SSH = 3, GCRYMPI_FMT_HEX = 4, GCRYMPI_FMT_USG = 5,
                       GCRYMPI_FMT_OPAQUE = 8 };
gcry_error_t gcry_sexp_build(gcry_sexp_t *retsexp, size_t *erroff, const char *format, ...);
size_t gcry_sexp_sprint(gcry_sexp_t sexp, int mode, void *buffer, size_t maxlength);
void gcry_sexp_release(gcry_sexp_t sexp);
void *gcry_malloc(size_t n);
void gcry_free(void *a);
gcry_mpi_t gcry_mpi_new(unsigned int nbits);
void gcry_mpi_release(gcry_mpi_t a);
gcry_mpi_t gcry_mpi_copy(const gcry_mpi_t a);
gcry_mpi_t gcry_mpi_set(gcry_mpi_t w, const gcry_mpi_t u);
gcry_mpi_t gcry_mpi_set_ui(gcry_mpi_t w, unsigned long u);
void gcry_mpi_neg(gcry_mpi_t w, gcry_mpi_t u);
int gcry_mpi_is_neg(gcry_mpi_t a);
int gcry_mpi_cmp(const gcry_mpi_t u, const gcry_mpi_t v);
int gcry_mpi_cmp_ui(const gcry_mpi_t u, unsigned long v);
void gcry_mpi_add(gcry_mpi_t w, gcry_mpi_t u, gcry_mpi_t v);
void gcry_mpi_add_ui(gcry_mpi_t w, gcry_mpi_t u, unsigned long v);
void gcry_mpi_addm(gcry_mpi_t w, gcry_mpi_t u, gcry_mpi_t v, gcry_mpi_t m);
void gcry_mpi_sub(gcry_mpi_t w, gcry_mpi_t u, gcry_mpi_t v);
void gcry_mpi_sub_ui(gcry_mpi_t w, gcry_mpi_t u, unsigned long v);
void gcry_mpi_subm(gcry_mpi_t w, gcry_mpi_t u, gcry_mpi_t v, gcry_mpi_t m);
void gcry_mpi_mul(gcry_mpi_t w, gcry_mpi_t u, gcry_mpi_t v);
void gcry_mpi_mul_ui(gcry_mpi_t w, gcry_mpi_t u, unsigned long v);
void gcry_mpi_mulm(gcry_mpi_t w, gcry_mpi_t u, gcry_mpi_t v, gcry_mpi_t m);
void gcry_mpi_div(gcry_mpi_t q, gcry_mpi_t r, gcry_mpi_t dividend, gcry_mpi_t divisor, int round);
void gcry_mpi_mod(gcry_mpi_t r, gcry_mpi_t dividend, gcry_mpi_t divisor);
unsigned int gcry_mpi_get_nbits(gcry_mpi_t a);
int gcry_mpi_test_bit(gcry_mpi_t a, unsigned int n);
void gcry_mpi_lshift(gcry_mpi_t x, gcry_mpi_t a, unsigned int n);
void gcry_mpi_rshift(gcry_mpi_t x, gcry_mpi_t a, unsigned int n);
gcry_error_t gcry_mpi_print(enum gcry_mpi_format format, unsigned char *buffer, size_t buflen,
                            size_t *nwritten, const gcry_mpi_t a);
gcry_error_t gcry_mpi_scan(gcry_mpi_t *ret_mpi, enum gcry_mpi_format format, const void *buffer,
                           size_t buflen, size_t *nscanned);
/* short aliases, as the public header spells them */
#define mpi_new(n) gcry_mpi_new((n))
#define mpi_release(a) do { gcry_mpi_release((a)); (a) = NULL; } while (0)
#define mpi_copy(a) gcry_mpi_copy((a))
#define mpi_set(w, u) gcry_mpi_set((w), (u))
#define mpi_set_ui(w, u) gcry_mpi_set_ui((w), (u))
#define mpi_neg(w, u) gcry_mpi_neg((w), (u))
#define mpi_is_neg(a) gcry_mpi_is_neg((a))
#define mpi_cmp(u, v) gcry_mpi_cmp((u), (v))
#define mpi_cmp_ui(u, v) gcry_mpi_cmp_ui((u), (v))
#define mpi_add_ui(w, u, v) gcry_mpi_add_ui((w), (u), (v))
#define mpi_add(w, u, v) gcry_mpi_add((w), (u), (v))
#define mpi_addm(w, u, v, m) gcry_mpi_addm((w), (u), (v), (m))
#define mpi_sub_ui(w, u, v) gcry_mpi_sub_ui((w), (u), (v))
#define mpi_sub(w, u, v) gcry_mpi_sub((w), (u), (v))
#define mpi_subm(w, u, v, m) gcry_mpi_subm((w), (u), (v), (m))
#define mpi_mul_ui(w, u, v) gcry_mpi_mul_ui((w), (u), (v))
#define mpi_mul(w, u, v) gcry_mpi_mul((w), (u), (v))
#define mpi_mulm(w, u, v, m) gcry_mpi_mulm((w), (u), (v), (m))
#define mpi_fdiv(q, r, a, m) gcry_mpi_div((q), (r), (a), (m), -1)
#define mpi_mod(r, a, m) gcry_mpi_mod((r), (a), (m))
#define mpi_get_nbits(a) gcry_mpi_get_nbits((a))
#define mpi_test_bit(a, b) gcry_mpi_test_bit((a), (b))
#define mpi_lshift(a, b, c) gcry_mpi_lshift((a), (b), (c))
#define mpi_rshift(a, b, c) gcry_mpi_rshift((a), (b), (c))
#ifdef __cplusplus
}
#endif
#endif
