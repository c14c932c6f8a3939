/* TEST INFRASTRUCTURE ONLY -- declaration-only stand-in for <cblas.h>.
 * /root/reference/include/sms/auxiliary/math.hpp:4 includes cblas.h, but nothing from it is used on the
 * odd-cycle separation path; this image has no BLAS headers.  Nothing here is ever linked. */
#ifndef SMSGPU_ORACLE_STUB_CBLAS_H
#define SMSGPU_ORACLE_STUB_CBLAS_H
#ifdef __cplusplus
extern "C" {
#endif
enum CBLAS_ORDER { CblasRowMajor = 101, CblasColMajor = 102 };
enum CBLAS_TRANSPOSE { CblasNoTrans = 111, CblasTrans = 112, CblasConjTrans = 113 };
void cblas_dgemv(enum CBLAS_ORDER order, enum CBLAS_TRANSPOSE trans, int m, int n, double alpha, const double *a,
                 int lda, const double *x, int incx, double beta, double *y, int incy);
double cblas_ddot(int n, const double *x, int incx, const double *y, int incy);
#ifdef __cplusplus
}
#endif
#endif
