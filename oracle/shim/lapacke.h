/* TEST INFRASTRUCTURE ONLY (oracle build).
 *
 * Declaration-only stand-in for <lapacke.h>: the image has no system LAPACKE, but
 * scipy's bundled OpenBLAS (LP64) exports the LAPACKE entry points with a
 * `scipy_` prefix.  HECTR's plaintext control code (src/matrices.c) needs six of
 * them.  No reference code here.
 */
#ifndef HECTR_B200_ORACLE_LAPACKE_SHIM_H
#define HECTR_B200_ORACLE_LAPACKE_SHIM_H
#define LAPACK_ROW_MAJOR 101
#define LAPACK_COL_MAJOR 102
typedef int lapack_int;
typedef double _Complex lapack_complex_double;
lapack_int scipy_LAPACKE_dgetrf(int layout, lapack_int m, lapack_int n, double *a, lapack_int lda, lapack_int *ipiv);
lapack_int scipy_LAPACKE_dgetri(int layout, lapack_int n, double *a, lapack_int lda, const lapack_int *ipiv);
lapack_int scipy_LAPACKE_zgetrf(int layout, lapack_int m, lapack_int n, lapack_complex_double *a, lapack_int lda, lapack_int *ipiv);
lapack_int scipy_LAPACKE_zgetri(int layout, lapack_int n, lapack_complex_double *a, lapack_int lda, const lapack_int *ipiv);
lapack_int scipy_LAPACKE_dgesvd(int layout, char jobu, char jobvt, lapack_int m, lapack_int n, double *a, lapack_int lda,
                                double *s, double *u, lapack_int ldu, double *vt, lapack_int ldvt, double *superb);
lapack_int scipy_LAPACKE_zgeev(int layout, char jobvl, char jobvr, lapack_int n, lapack_complex_double *a, lapack_int lda,
                               lapack_complex_double *w, lapack_complex_double *vl, lapack_int ldvl,
                               lapack_complex_double *vr, lapack_int ldvr);
#define LAPACKE_dgetrf scipy_LAPACKE_dgetrf
#define LAPACKE_dgetri scipy_LAPACKE_dgetri
#define LAPACKE_zgetrf scipy_LAPACKE_zgetrf
#define LAPACKE_zgetri scipy_LAPACKE_zgetri
#define LAPACKE_dgesvd scipy_LAPACKE_dgesvd
#define LAPACKE_zgeev scipy_LAPACKE_zgeev
#endif
