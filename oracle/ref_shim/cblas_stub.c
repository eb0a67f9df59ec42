/* Aborting definitions of the CBLAS prototypes in cblas.h (same directory): the reference's CPU
 * helpers link, the GPU path this build exists for never reaches them. */
#include <stdio.h>
#include <stdlib.h>
#include "cblas.h"
static void die(const char* f) { fprintf(stderr, "ref_shim: %s called -- no CBLAS behind the reference build\n", f); abort(); }
#define STUB(T, p) \
  T cblas_##p##asum(const int n, const T* x, const int incx) { (void)n; (void)x; (void)incx; die("asum"); return 0; } \
  void cblas_##p##axpy(const int n, const T a, const T* x, const int incx, T* y, const int incy) { (void)n; (void)a; (void)x; (void)incx; (void)y; (void)incy; die("axpy"); } \
  void cblas_##p##axpby(const int n, const T a, const T* x, const int incx, const T b, T* y, const int incy) { (void)n; (void)a; (void)x; (void)incx; (void)b; (void)y; (void)incy; die("axpby"); } \
  void cblas_##p##copy(const int n, const T* x, const int incx, T* y, const int incy) { (void)n; (void)x; (void)incx; (void)y; (void)incy; die("copy"); } \
  T cblas_##p##dot(const int n, const T* x, const int incx, const T* y, const int incy) { (void)n; (void)x; (void)incx; (void)y; (void)incy; die("dot"); return 0; } \
  void cblas_##p##scal(const int n, const T a, T* x, const int incx) { (void)n; (void)a; (void)x; (void)incx; die("scal"); } \
  void cblas_##p##gemm(const enum CBLAS_ORDER o, const enum CBLAS_TRANSPOSE ta, const enum CBLAS_TRANSPOSE tb, const int m, const int n, const int k, const T alpha, const T* a, const int lda, const T* b, const int ldb, const T beta, T* c, const int ldc) { (void)o; (void)ta; (void)tb; (void)m; (void)n; (void)k; (void)alpha; (void)a; (void)lda; (void)b; (void)ldb; (void)beta; (void)c; (void)ldc; die("gemm"); } \
  void cblas_##p##gemv(const enum CBLAS_ORDER o, const enum CBLAS_TRANSPOSE ta, const int m, const int n, const T alpha, const T* a, const int lda, const T* x, const int incx, const T beta, T* y, const int incy) { (void)o; (void)ta; (void)m; (void)n; (void)alpha; (void)a; (void)lda; (void)x; (void)incx; (void)beta; (void)y; (void)incy; die("gemv"); }
STUB(float, s)
STUB(double, d)
