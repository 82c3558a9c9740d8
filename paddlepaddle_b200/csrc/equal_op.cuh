// equal_op.cuh — the reference's EqualFunctor as a device predicate, shared by the compare kernels
// (elementwise_pred.cu) and the max / min backward (reduce_grad.cu), which is "out_grad where x equals out".
#pragma once
#include "common.cuh"

namespace pdb200 {

template <typename T> __device__ __forceinline__ float as_float(T v) { return static_cast<float>(v); }
template <> __device__ __forceinline__ float as_float<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float as_float<__half>(__half v) { return __half2float(v); }


// EqualFunctor (compare_functors.h:37-51): float and double are "equal" when |a - b| < 1e-8 (difference taken
// in T, compared in double), infinities compare exactly, NaN equals nothing.  bf16 / fp16 take the same
// route (std::is_floating_point is specialised for them): the difference is one float subtraction rounded
// back to the 16-bit type.  Integers and bool use plain ==.
template <typename T> struct EqualOp {
  static __device__ __forceinline__ bool apply(T a, T b) { return a == b; }
};
// Written without branches (a select, not three early returns: divergent regions per element cost more than the
// arithmetic) and, for float, without the trip through double: for a float d, (double)|d| < 1e-8 holds exactly when
// |d| <= 9.99999993922529e-09f, the largest float below 1e-8.  A NaN operand makes the difference NaN and the comparison
// false; an infinite operand selects the exact comparison.
constexpr float kEqTol = 9.99999993922529e-09f;
template <> struct EqualOp<float> {
  static __device__ __forceinline__ bool apply(float a, float b) {
    const bool inf = (fabsf(a) == INFINITY) | (fabsf(b) == INFINITY);
    return inf ? (a == b) : (fabsf(__fsub_rn(a, b)) <= kEqTol);
  }
};
template <> struct EqualOp<double> {
  static __device__ __forceinline__ bool apply(double a, double b) {
    // the reference tests isinf / isnan on the value cast to float: a finite double beyond float range counts as inf
    const bool inf = (fabsf(static_cast<float>(a)) == INFINITY) | (fabsf(static_cast<float>(b)) == INFINITY);
    return inf ? (a == b) : (fabs(__dsub_rn(a, b)) < 1e-8);
  }
};
#define PDB_HALF_EQ(T)                                                                                   \
  template <> struct EqualOp<T> {                                                                        \
    static __device__ __forceinline__ bool apply(T a, T b) {                                             \
      const float fa = as_float(a), fb = as_float(b);                                                    \
      const bool inf = (fabsf(fa) == INFINITY) | (fabsf(fb) == INFINITY);                                \
      return inf ? (fa == fb) : (fabsf(as_float(cvt<T>(__fsub_rn(fa, fb)))) <= kEqTol);                  \
    }                                                                                                    \
  };
PDB_HALF_EQ(__nv_bfloat16)
PDB_HALF_EQ(__half)
#undef PDB_HALF_EQ

}  // namespace pdb200
