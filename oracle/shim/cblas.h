/* oracle/shim/cblas.h — TEST INFRASTRUCTURE. The three BLAS entry points the reference's native/compute/BatchDistance.cpp:9-12,24-96
 * calls on Linux (cblas_sgemm, cblas_sdot, cblas_sgemv), declared here because no BLAS is installed on this image; they are
 * implemented in oracle/ref_driver.cpp as the plain loops of the reference's own `#else` fallback branch (BatchDistance.cpp:36-46,
 * 62-70, 98-105). With this header on the include path the UNMODIFIED HNSWIndex.cpp / BatchDistance.cpp compile into oracle/_ref. */
#ifndef ORACLE_SHIM_CBLAS_H
#define ORACLE_SHIM_CBLAS_H
enum CBLAS_ORDER { CblasRowMajor = 101, CblasColMajor = 102 };
enum CBLAS_TRANSPOSE { CblasNoTrans = 111, CblasTrans = 112 };
void cblas_sgemm(enum CBLAS_ORDER order, enum CBLAS_TRANSPOSE ta, enum CBLAS_TRANSPOSE tb, int m, int n, int k, float alpha, const float* A,
                 int lda, const float* B, int ldb, float beta, float* C, int ldc);
float cblas_sdot(int n, const float* x, int incx, const float* y, int incy);
void cblas_sgemv(enum CBLAS_ORDER order, enum CBLAS_TRANSPOSE ta, int m, int n, float alpha, const float* A, int lda, const float* x, int incx,
                 float beta, float* y, int incy);
#endif
