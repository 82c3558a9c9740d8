// TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Stand-in for <cblas.h>: paddle/phi/kernels/funcs/blas/blas.h is pulled in by
// funcs/math_function.h and funcs/layer_norm_util.h, whose RowwiseMean and
// Sub/Div functors the oracle needs.  None of the CPU kernels on the hot path
// (reduce, argmax, ElementwiseCompute, softmax, layer_norm) calls BLAS, so the
// symbols only have to parse; they abort if ever reached.  The one exception is ?copy: the CPU add_grad
// kernel copies dout with blas.VCOPY (cpu/elementwise_grad.h:93-100) — a strided copy moves values without
// any arithmetic, so it is spelled out here.
#pragma once
#include <cstdlib>

typedef enum CBLAS_ORDER { CblasRowMajor = 101, CblasColMajor = 102 } CBLAS_ORDER;
typedef CBLAS_ORDER CBLAS_LAYOUT;
typedef enum CBLAS_TRANSPOSE { CblasNoTrans = 111, CblasTrans = 112, CblasConjTrans = 113, CblasConjNoTrans = 114 } CBLAS_TRANSPOSE;
typedef enum CBLAS_UPLO { CblasUpper = 121, CblasLower = 122 } CBLAS_UPLO;
typedef enum CBLAS_DIAG { CblasNonUnit = 131, CblasUnit = 132 } CBLAS_DIAG;
typedef enum CBLAS_SIDE { CblasLeft = 141, CblasRight = 142 } CBLAS_SIDE;

#define PDB200_CBLAS_VOID(name) \
  template <typename... A> inline void name(A...) { std::abort(); }
#define PDB200_CBLAS_RET(ret, name) \
  template <typename... A> inline ret name(A...) { std::abort(); }

PDB200_CBLAS_VOID(cblas_caxpy) PDB200_CBLAS_VOID(cblas_ccopy) PDB200_CBLAS_VOID(cblas_cgemm)
PDB200_CBLAS_VOID(cblas_cgemm_batch) PDB200_CBLAS_VOID(cblas_cgemv) PDB200_CBLAS_VOID(cblas_ctrsm)
PDB200_CBLAS_RET(double, cblas_dasum) PDB200_CBLAS_VOID(cblas_daxpy)
PDB200_CBLAS_RET(double, cblas_ddot) PDB200_CBLAS_VOID(cblas_dgemm) PDB200_CBLAS_VOID(cblas_dgemm_batch)
PDB200_CBLAS_VOID(cblas_dgemv) PDB200_CBLAS_VOID(cblas_dscal) PDB200_CBLAS_VOID(cblas_dtrsm)
PDB200_CBLAS_RET(float, cblas_sasum) PDB200_CBLAS_VOID(cblas_saxpy)
PDB200_CBLAS_RET(float, cblas_sdot) PDB200_CBLAS_VOID(cblas_sgemm) PDB200_CBLAS_VOID(cblas_sgemm_batch)
PDB200_CBLAS_VOID(cblas_sgemv) PDB200_CBLAS_VOID(cblas_sscal) PDB200_CBLAS_VOID(cblas_strsm)
PDB200_CBLAS_VOID(cblas_zaxpy) PDB200_CBLAS_VOID(cblas_zcopy) PDB200_CBLAS_VOID(cblas_zgemm)
PDB200_CBLAS_VOID(cblas_zgemm_batch) PDB200_CBLAS_VOID(cblas_zgemv) PDB200_CBLAS_VOID(cblas_ztrsm)

inline void cblas_scopy(int n, const float* x, int incx, float* y, int incy) {
  for (int i = 0; i < n; ++i) y[(long)i * incy] = x[(long)i * incx];
}
inline void cblas_dcopy(int n, const double* x, int incx, double* y, int incy) {
  for (int i = 0; i < n; ++i) y[(long)i * incy] = x[(long)i * incx];
}
