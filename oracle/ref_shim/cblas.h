/* Build shim (oracle/Makefile.refcuda): prototypes of the CBLAS calls the reference's CPU helpers
 * name (util/math_functions.cpp, util/mkl_alternate.hpp).  No CBLAS is installed in this image and the
 * GPU path never calls them; oracle/ref_shim/cblas_stub.c defines them as aborting stubs. */
#ifndef DAU_REF_SHIM_CBLAS_H_
#define DAU_REF_SHIM_CBLAS_H_
enum CBLAS_ORDER { CblasRowMajor = 101, CblasColMajor = 102 };
enum CBLAS_TRANSPOSE { CblasNoTrans = 111, CblasTrans = 112, CblasConjTrans = 113 };
#define DAU_SHIM_BLAS(T, p) \
  T cblas_##p##asum(const int n, const T* x, const int incx); \
  void cblas_##p##axpy(const int n, const T a, const T* x, const int incx, T* y, const int incy); \
  void cblas_##p##axpby(const int n, const T a, const T* x, const int incx, const T b, T* y, const int incy); \
  void cblas_##p##copy(const int n, const T* x, const int incx, T* y, const int incy); \
  T cblas_##p##dot(const int n, const T* x, const int incx, const T* y, const int incy); \
  void cblas_##p##scal(const int n, const T a, T* x, const int incx); \
  void cblas_##p##gemm(const enum CBLAS_ORDER o, const enum CBLAS_TRANSPOSE ta, const enum CBLAS_TRANSPOSE tb, \
                       const int m, const int n, const int k, const T alpha, const T* a, const int lda, \
                       const T* b, const int ldb, const T beta, T* c, const int ldc); \
  void cblas_##p##gemv(const enum CBLAS_ORDER o, const enum CBLAS_TRANSPOSE ta, const int m, const int n, \
                       const T alpha, const T* a, const int lda, const T* x, const int incx, \
                       const T beta, T* y, const int incy);
DAU_SHIM_BLAS(float, s)
DAU_SHIM_BLAS(double, d)
#endif
