// elementwise_pred.cu — compare / logical / bitwise binary ops on the broadcast engine.
//
// Replaces (fresh design): the KPS compare / logical / bitwise kernels, which are thin callers of
// funcs::BroadcastKernel —
//   paddle/phi/kernels/kps/compare_kernel.cu:49-66,143-173   (functors: funcs/compare_functors.h:21-62)
//   paddle/phi/kernels/kps/logical_kernel.cu:112-134          (functors: funcs/logical_functor.h:20-40)
//   paddle/phi/kernels/kps/bitwise_kernel.cu:94-136           (functors: funcs/bitwise_functors.h:20-49)
// Same launcher and kernels as add/multiply (elementwise_impl.cuh); compare and logical write one bool per
// element, so a thread that loads 16 bytes of each operand stores a 2/4/8/16-byte pack of bools and a warp's
// store is still one contiguous run.
#include <type_traits>

#include "elementwise_impl.cuh"
#include "equal_op.cuh"

namespace pdb200 {

// ---- compare ------------------------------------------------------------------------------------------
template <typename T> struct LessThanOp {
  static __device__ __forceinline__ bool apply(T a, T b) { return a < b; }
};
template <typename T> struct LessEqualOp {
  static __device__ __forceinline__ bool apply(T a, T b) { return a <= b; }
};
template <typename T> struct NotEqualOp {
  static __device__ __forceinline__ bool apply(T a, T b) { return !EqualOp<T>::apply(a, b); }
};
#define PDB_HALF_CMP(T)                                                                                  \
  template <> struct LessThanOp<T> {                                                                     \
    static __device__ __forceinline__ bool apply(T a, T b) { return as_float(a) < as_float(b); }         \
  };                                                                                                     \
  template <> struct LessEqualOp<T> {                                                                    \
    static __device__ __forceinline__ bool apply(T a, T b) { return as_float(a) <= as_float(b); }        \
  };
PDB_HALF_CMP(__nv_bfloat16)
PDB_HALF_CMP(__half)
#undef PDB_HALF_CMP

// 16-bit less_than / less_equal, eight elements at a time: four packed compares (HSETP2 — ordered, so a NaN operand gives
// false exactly like the float comparison of the scalar functor) whose 0xFFFF / 0 lane masks are squeezed into one
// byte per element.  The scalar route costs ~10 instructions per element (unpack both operands, compare, select, pack).
template <typename T2>
__device__ __forceinline__ uint32_t lt2_mask(T2 a, T2 b) { return __hlt2_mask(a, b); }
template <typename T2>
__device__ __forceinline__ uint32_t le2_mask(T2 a, T2 b) { return __hle2_mask(a, b); }
template <typename T> struct Half2Of;
template <> struct Half2Of<__nv_bfloat16> { using type = __nv_bfloat162; };
template <> struct Half2Of<__half> { using type = __half2; };
template <typename T, bool LE>
__device__ __forceinline__ Pack<bool, 8> cmp8_packed(const Vec<T>& a, const Vec<T>& b) {
  using T2 = typename Half2Of<T>::type;
  const T2* pa = reinterpret_cast<const T2*>(&a);
  const T2* pb = reinterpret_cast<const T2*>(&b);
  uint32_t w[2];
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const uint32_t m0 = (LE ? le2_mask(pa[2 * h], pb[2 * h]) : lt2_mask(pa[2 * h], pb[2 * h])) & 0x00010001u;
    const uint32_t m1 = (LE ? le2_mask(pa[2 * h + 1], pb[2 * h + 1]) : lt2_mask(pa[2 * h + 1], pb[2 * h + 1])) & 0x00010001u;
    // m0 = elements (4h, 4h+1) in bits 0 and 16; m1 = elements (4h+2, 4h+3): bytes b0 b1 b2 b3 of the output word
    w[h] = (m0 & 0x1u) | ((m0 >> 8) & 0x100u) | ((m1 & 0x1u) << 16) | ((m1 << 8) & 0x01000000u);
  }
  Pack<bool, 8> r;
  *reinterpret_cast<uint2*>(&r) = make_uint2(w[0], w[1]);
  return r;
}
// 16-bit equal / not_equal, packed the same way.
//   float16 : the smallest positive value (2^-24) is above the 1e-8 tolerance and the difference of two float16 is a
//             multiple of 2^-24, so round16(a - b) is within the tolerance exactly when a == b: ONE ordered packed
//             compare (NaN equals nothing, inf == inf, +0 == -0 — what the scalar functor gives).
//   bfloat16: tiny values can differ by less than 1e-8, so equal = (a == b) | (|a - b| <= tol) with the difference in
//             packed bf16 arithmetic (one rounding) and tol = 0x322B, the largest bfloat16 not above 1e-8.  Same bits as
//             the reference's round_bf16(float(a) - float(b)): when the float subtraction is inexact the operands lie
//             more than 2^16 apart and both roundings return the larger one; a == b covers equal infinities, whose
//             difference is NaN.
__device__ __forceinline__ uint32_t eq2_mask(__half2 a, __half2 b) { return __heq2_mask(a, b); }
__device__ __forceinline__ uint32_t eq2_mask(__nv_bfloat162 a, __nv_bfloat162 b) {
  const unsigned int tol_bits = 0x322B322Bu;
  const __nv_bfloat162 tol = *reinterpret_cast<const __nv_bfloat162*>(&tol_bits);
  return __heq2_mask(a, b) | __hle2_mask(__habs2(__hsub2(a, b)), tol);
}
template <typename T, bool NEGATE>
__device__ __forceinline__ Pack<bool, 8> eq8_packed(const Vec<T>& a, const Vec<T>& b) {
  using T2 = typename Half2Of<T>::type;
  const T2* pa = reinterpret_cast<const T2*>(&a);
  const T2* pb = reinterpret_cast<const T2*>(&b);
  uint32_t w[2];
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    uint32_t m0 = eq2_mask(pa[2 * h], pb[2 * h]), m1 = eq2_mask(pa[2 * h + 1], pb[2 * h + 1]);
    if (NEGATE) {
      m0 = ~m0;
      m1 = ~m1;
    }
    m0 &= 0x00010001u;
    m1 &= 0x00010001u;
    w[h] = (m0 & 0x1u) | ((m0 >> 8) & 0x100u) | ((m1 & 0x1u) << 16) | ((m1 << 8) & 0x01000000u);
  }
  Pack<bool, 8> r;
  *reinterpret_cast<uint2*>(&r) = make_uint2(w[0], w[1]);
  return r;
}
#define PDB_PACKED_EQ(T)                                                                                      \
  template <> struct VecApply<T, EqualOp<T>> {                                                                \
    static __device__ __forceinline__ Pack<bool, 8> run(const Vec<T>& a, const Vec<T>& b) { return eq8_packed<T, false>(a, b); } \
  };                                                                                                          \
  template <> struct VecApply<T, NotEqualOp<T>> {                                                             \
    static __device__ __forceinline__ Pack<bool, 8> run(const Vec<T>& a, const Vec<T>& b) { return eq8_packed<T, true>(a, b); } \
  };
PDB_PACKED_EQ(__nv_bfloat16)
PDB_PACKED_EQ(__half)
#undef PDB_PACKED_EQ
#define PDB_PACKED_CMP(T)                                                                                     \
  template <> struct VecApply<T, LessThanOp<T>> {                                                             \
    static __device__ __forceinline__ Pack<bool, 8> run(const Vec<T>& a, const Vec<T>& b) { return cmp8_packed<T, false>(a, b); } \
  };                                                                                                          \
  template <> struct VecApply<T, LessEqualOp<T>> {                                                            \
    static __device__ __forceinline__ Pack<bool, 8> run(const Vec<T>& a, const Vec<T>& b) { return cmp8_packed<T, true>(a, b); } \
  };
PDB_PACKED_CMP(__nv_bfloat16)
PDB_PACKED_CMP(__half)
#undef PDB_PACKED_CMP

// ---- logical: static_cast<bool>(a) op static_cast<bool>(b) ------------------------------------------------
template <typename T> __device__ __forceinline__ bool truthy(T v) { return static_cast<bool>(v); }
template <> __device__ __forceinline__ bool truthy<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v) != 0.0f; }
template <> __device__ __forceinline__ bool truthy<__half>(__half v) { return __half2float(v) != 0.0f; }
template <typename T> struct LogicalAndOp {
  static __device__ __forceinline__ bool apply(T a, T b) { return truthy(a) && truthy(b); }
};
template <typename T> struct LogicalOrOp {
  static __device__ __forceinline__ bool apply(T a, T b) { return truthy(a) || truthy(b); }
};
template <typename T> struct LogicalXorOp {
  static __device__ __forceinline__ bool apply(T a, T b) { return truthy(a) != truthy(b); }
};

// ---- bitwise: a & b, a | b, a ^ b; bool uses &&, ||, != -----------------------------------------------------
template <typename T> struct BitAndOp {
  static __device__ __forceinline__ T apply(T a, T b) { return static_cast<T>(a & b); }
};
template <typename T> struct BitOrOp {
  static __device__ __forceinline__ T apply(T a, T b) { return static_cast<T>(a | b); }
};
template <typename T> struct BitXorOp {
  static __device__ __forceinline__ T apply(T a, T b) { return static_cast<T>(a ^ b); }
};
template <> struct BitAndOp<bool> {
  static __device__ __forceinline__ bool apply(bool a, bool b) { return a && b; }
};
template <> struct BitOrOp<bool> {
  static __device__ __forceinline__ bool apply(bool a, bool b) { return a || b; }
};
template <> struct BitXorOp<bool> {
  static __device__ __forceinline__ bool apply(bool a, bool b) { return a != b; }
};

// ---- shifts (funcs/bitwise_functors.h:51-200): a shift count outside [0, bits) gives 0 — or, for an arithmetic right
// shift of a signed value, the sign fill; a logical right shift of a signed value shifts its unsigned representation
template <typename T>
__device__ __forceinline__ bool bad_count(T b) {
  if constexpr (std::is_signed<T>::value) return b < T(0) || b >= T(sizeof(T) * 8);
  else return b >= T(sizeof(T) * 8);
}
template <typename T>
__device__ __forceinline__ bool is_negative(T a) {
  if constexpr (std::is_signed<T>::value) return a < T(0);
  else return false;
}
template <typename T> struct ShlOp {
  static __device__ __forceinline__ T apply(T a, T b) {
    using U = typename std::make_unsigned<T>::type;
    if (bad_count(b)) return T(0);
    return static_cast<T>(static_cast<U>(static_cast<U>(a) << b));
  }
};
template <typename T> struct ShrArithOp {
  static __device__ __forceinline__ T apply(T a, T b) {
    if (bad_count(b)) return is_negative(a) ? T(-1) : T(0);
    return static_cast<T>(a >> b);
  }
};
template <typename T> struct ShrLogicOp {
  static __device__ __forceinline__ T apply(T a, T b) {
    using U = typename std::make_unsigned<T>::type;
    if (bad_count(b)) return T(0);
    return static_cast<T>(static_cast<U>(static_cast<U>(a) >> b));
  }
};

#define B(dt) (1u << (dt))
constexpr uint32_t kShiftTypes = B(PDB200_UINT8) | B(PDB200_INT8) | B(PDB200_INT16) | B(PDB200_INT32) |
                                 B(PDB200_INT64);  // bitwise_kernel.cu:138-156
constexpr uint32_t kAllTypes = (1u << (PDB200_FLOAT64 + 1)) - 1;                      // compare_kernel.cu:157-173
constexpr uint32_t kLogicalTypes = kAllTypes & ~B(PDB200_UINT8);                      // logical_kernel.cu:112-129
constexpr uint32_t kBitwiseTypes = B(PDB200_BOOL) | B(PDB200_UINT8) | B(PDB200_INT8) | B(PDB200_INT16) |
                                   B(PDB200_INT32) | B(PDB200_INT64);                 // bitwise_kernel.cu:94-136
#undef B

}  // namespace pdb200

extern "C" {

int pdb200_compare(int op, const void* x, const int64_t* x_dims, int x_rank, const void* y, const int64_t* y_dims,
                   int y_rank, void* out, int dtype, void* stream) {
  using namespace pdb200;
  switch (op) {
    case PDB200_LESS_THAN:
      return elementwise_entry<LessThanOp, kAllTypes>("less_than", x, x_dims, x_rank, y, y_dims, y_rank, out, dtype, stream);
    case PDB200_LESS_EQUAL:
      return elementwise_entry<LessEqualOp, kAllTypes>("less_equal", x, x_dims, x_rank, y, y_dims, y_rank, out, dtype, stream);
    // a > b is b < a: the operands trade places on the host, the kernels are shared
    case PDB200_GREATER_THAN:
      return elementwise_entry<LessThanOp, kAllTypes>("greater_than", y, y_dims, y_rank, x, x_dims, x_rank, out, dtype, stream);
    case PDB200_GREATER_EQUAL:
      return elementwise_entry<LessEqualOp, kAllTypes>("greater_equal", y, y_dims, y_rank, x, x_dims, x_rank, out, dtype, stream);
    case PDB200_EQUAL:
      return elementwise_entry<EqualOp, kAllTypes>("equal", x, x_dims, x_rank, y, y_dims, y_rank, out, dtype, stream);
    case PDB200_NOT_EQUAL:
      return elementwise_entry<NotEqualOp, kAllTypes>("not_equal", x, x_dims, x_rank, y, y_dims, y_rank, out, dtype, stream);
    default:
      return fail(PDB200_INVALID_ARGUMENT, "compare: unknown op code " + std::to_string(op));
  }
}

int pdb200_logical(int op, const void* x, const int64_t* x_dims, int x_rank, const void* y, const int64_t* y_dims,
                   int y_rank, void* out, int dtype, void* stream) {
  using namespace pdb200;
  switch (op) {
    case PDB200_LOGICAL_AND:
      return elementwise_entry<LogicalAndOp, kLogicalTypes>("logical_and", x, x_dims, x_rank, y, y_dims, y_rank, out, dtype, stream);
    case PDB200_LOGICAL_OR:
      return elementwise_entry<LogicalOrOp, kLogicalTypes>("logical_or", x, x_dims, x_rank, y, y_dims, y_rank, out, dtype, stream);
    case PDB200_LOGICAL_XOR:
      return elementwise_entry<LogicalXorOp, kLogicalTypes>("logical_xor", x, x_dims, x_rank, y, y_dims, y_rank, out, dtype, stream);
    default:
      return fail(PDB200_INVALID_ARGUMENT, "logical: unknown op code " + std::to_string(op) +
                                               " (logical_not is unary: pdb200_logical_not)");
  }
}

int pdb200_bitwise(int op, const void* x, const int64_t* x_dims, int x_rank, const void* y, const int64_t* y_dims,
                   int y_rank, void* out, int dtype, void* stream) {
  using namespace pdb200;
  switch (op) {
    case PDB200_BITWISE_AND:
      return elementwise_entry<BitAndOp, kBitwiseTypes>("bitwise_and", x, x_dims, x_rank, y, y_dims, y_rank, out, dtype, stream);
    case PDB200_BITWISE_OR:
      return elementwise_entry<BitOrOp, kBitwiseTypes>("bitwise_or", x, x_dims, x_rank, y, y_dims, y_rank, out, dtype, stream);
    case PDB200_BITWISE_XOR:
      return elementwise_entry<BitXorOp, kBitwiseTypes>("bitwise_xor", x, x_dims, x_rank, y, y_dims, y_rank, out, dtype, stream);
    default:
      return fail(PDB200_INVALID_ARGUMENT, "bitwise: unknown op code " + std::to_string(op) +
                                               " (bitwise_not is unary: pdb200_bitwise_not)");
  }
}

int pdb200_bitwise_shift(int right, int is_arithmetic, const void* x, const int64_t* x_dims, int x_rank, const void* y,
                         const int64_t* y_dims, int y_rank, void* out, int dtype, void* stream) {
  using namespace pdb200;
  if (!right)  // the arithmetic and the logical left shift are the same function
    return elementwise_entry<ShlOp, kShiftTypes>("bitwise_left_shift", x, x_dims, x_rank, y, y_dims, y_rank, out, dtype, stream);
  if (is_arithmetic)
    return elementwise_entry<ShrArithOp, kShiftTypes>("bitwise_right_shift", x, x_dims, x_rank, y, y_dims, y_rank, out, dtype, stream);
  return elementwise_entry<ShrLogicOp, kShiftTypes>("bitwise_right_shift", x, x_dims, x_rank, y, y_dims, y_rank, out, dtype, stream);
}

}  // extern "C"
