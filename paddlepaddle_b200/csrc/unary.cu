// unary.cu — one-input (and the one three-input) maps of the elementwise family:
// logical_not, bitwise_not, scale, cast, where.
//
// Replaces (fresh design): the callers of funcs::ElementwiseKernel with a single tensor input —
//   paddle/phi/kernels/kps/logical_kernel.cu:71-91           LogicalNotKernel   (funcs/logical_functor.h:34-38)
//   paddle/phi/kernels/kps/bitwise_kernel.cu:69-80           BitwiseNotKernel   (funcs/bitwise_functors.h:38-49)
//   paddle/phi/kernels/gpu/scale_kernel.cu:24-65             ScaleKernel / ScaleFunctor
//   paddle/phi/kernels/gpu/cast_kernel.cu, gpu/cast_impl.h   CastKernel / CastFunctor (static_cast per element)
//   paddle/phi/kernels/gpu/where_kernel.cu:24-51             WhereKernel / CondFunctor (three same-shape inputs)
//
// Design: pure streams.  A thread moves N elements per step with N chosen so that the WIDER of the input and
// output element types moves 16 bytes (the narrower side moves 2..16 bytes; a warp's access is contiguous
// either way), four steps in flight per thread, 64-bit offsets.  Input and output may be the same buffer when
// the element sizes match: each pack is read and then written by the same thread.
#include <string>
#include <type_traits>

#include "common.cuh"

namespace pdb200 {

constexpr int kUThreads = 256;
constexpr int kUUnroll = 4;

template <typename TI, typename TO>
struct UnaryN {
  static constexpr int value = 16 / (sizeof(TI) > sizeof(TO) ? sizeof(TI) : sizeof(TO));
};

template <typename TI, typename TO, typename F>
__global__ void __launch_bounds__(kUThreads) unary_kernel(const TI* __restrict__ x, TO* __restrict__ out,
                                                          int64_t npack, int64_t numel, F f) {
  constexpr int N = UnaryN<TI, TO>::value;
  const int64_t base = (int64_t)blockIdx.x * (kUThreads * kUUnroll) + threadIdx.x;
  Pack<TI, N> a[kUUnroll];
  const bool full = base + (kUUnroll - 1) * kUThreads < npack;
#pragma unroll
  for (int u = 0; u < kUUnroll; ++u) {
    const int64_t v = base + u * kUThreads;
    if (full || v < npack) a[u] = ld_pack<TI, N>(x + v * N);
  }
#pragma unroll
  for (int u = 0; u < kUUnroll; ++u) {
    const int64_t v = base + u * kUThreads;
    if (full || v < npack) {
      Pack<TO, N> r;
#pragma unroll
      for (int i = 0; i < N; ++i) r.v[i] = f(a[u].v[i]);
      if constexpr (sizeof(TO) * N == 1) out[v] = r.v[0];
      else st_stream(out + v * N, r);
    }
  }
  if (blockIdx.x == gridDim.x - 1) {  // scalar tail: last CTA, first threads
    const int64_t t = npack * N + threadIdx.x;
    if (t < numel) out[t] = f(x[t]);
  }
}

template <typename TI, typename TO, typename F>
__global__ void __launch_bounds__(kUThreads) unary_scalar_kernel(const TI* x, TO* out, int64_t numel, F f) {
  int64_t i = (int64_t)blockIdx.x * kUThreads + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * kUThreads;
  for (; i < numel; i += stride) out[i] = f(x[i]);
}

template <typename TI, typename TO, typename F>
int launch_unary(const char* name, const TI* x, TO* out, int64_t numel, F f, cudaStream_t st) {
  if (numel == 0) return PDB200_OK;
  constexpr int N = UnaryN<TI, TO>::value;
  if (is_aligned_to(x, sizeof(TI) * N) && is_aligned_to(out, sizeof(TO) * N)) {
    const int64_t npack = numel / N;
    const int64_t tiles = std::max<int64_t>(1, (npack + kUThreads * kUUnroll - 1) / (kUThreads * kUUnroll));
    if (tiles > 0x7fffffffLL) return fail(PDB200_UNSUPPORTED, std::string(name) + ": tensor too large for one launch");
    unary_kernel<TI, TO, F><<<(unsigned)tiles, kUThreads, 0, st>>>(x, out, npack, numel, f);
  } else {
    const int64_t blocks = std::min<int64_t>((numel + kUThreads - 1) / kUThreads, (int64_t)sm_count() * 32);
    unary_scalar_kernel<TI, TO, F><<<(unsigned)blocks, kUThreads, 0, st>>>(x, out, numel, f);
  }
  return after_launch(name);
}

// ---- conversions: static_cast<TO>(v), bf16/fp16 through float (phi/common/bfloat16.h, float16.h) ---------------
template <typename T> struct IsHalf { static constexpr bool value = false; };
template <> struct IsHalf<__nv_bfloat16> { static constexpr bool value = true; };
template <> struct IsHalf<__half> { static constexpr bool value = true; };

template <typename TO, typename TI>
__device__ __forceinline__ TO convert(TI v) {
  if constexpr (std::is_same<TO, TI>::value) {
    return v;
  } else if constexpr (IsHalf<TI>::value) {
    return convert<TO, float>(cvt<float>(v));
  } else if constexpr (IsHalf<TO>::value) {
    return cvt<TO>(static_cast<float>(v));  // every source goes through float, as the reference's constructors do
  } else if constexpr (std::is_same<TO, bool>::value) {
    return v != TI(0);
  } else if constexpr (std::is_floating_point<TI>::value && std::is_integral<TO>::value && sizeof(TO) < 4) {
    // the host compiler truncates through a 32-bit integer and keeps the low bits; do the same
    return static_cast<TO>(static_cast<int32_t>(v));
  } else {
    return static_cast<TO>(v);
  }
}

template <typename TI, typename TO>
struct CastF {
  __device__ __forceinline__ TO operator()(TI v) const { return convert<TO, TI>(v); }
};
template <typename T>
struct LogicalNotF {
  __device__ __forceinline__ bool operator()(T v) const { return !convert<bool, T>(v); }
};
template <typename T>
struct BitNotF {
  __device__ __forceinline__ T operator()(T v) const {
    if constexpr (std::is_same<T, bool>::value) return !v;
    else return static_cast<T>(~v);
  }
};
// ScaleFunctor<DataT, ParamT> (gpu/scale_kernel.cu:24-43): ParamT = float for bf16/fp16, DataT otherwise.
template <typename T, typename P>
struct ScaleF {
  P scale, bias;
  bool bias_after_scale;
  __device__ __forceinline__ T operator()(T v) const {
    const P x = convert<P, T>(v);
    if constexpr (std::is_integral<P>::value && !std::is_same<P, bool>::value) {
      using U = typename std::make_unsigned<P>::type;  // integer overflow wraps
      const U r = bias_after_scale ? (U)((U)scale * (U)x + (U)bias) : (U)((U)scale * (U)((U)x + (U)bias));
      return static_cast<T>(static_cast<P>(r));
    } else {
      return convert<T, P>(bias_after_scale ? scale * x + bias : scale * (x + bias));
    }
  }
};

// ---- where: out = cond ? x : y on same-shape tensors; a pure bit select, so it is built per element SIZE -----
template <typename U>
__global__ void __launch_bounds__(kUThreads) where_kernel(const bool* __restrict__ cond, const U* __restrict__ x,
                                                          const U* __restrict__ y, U* __restrict__ out,
                                                          int64_t npack, int64_t numel) {
  constexpr int N = 16 / sizeof(U);
  const int64_t base = (int64_t)blockIdx.x * (kUThreads * kUUnroll) + threadIdx.x;
  Pack<bool, N> c[kUUnroll];
  Pack<U, N> a[kUUnroll], b[kUUnroll];
  const bool full = base + (kUUnroll - 1) * kUThreads < npack;
#pragma unroll
  for (int u = 0; u < kUUnroll; ++u) {
    const int64_t v = base + u * kUThreads;
    if (full || v < npack) {
      c[u] = ld_pack<bool, N>(cond + v * N);
      a[u] = ld_pack<U, N>(x + v * N);
      b[u] = ld_pack<U, N>(y + v * N);
    }
  }
#pragma unroll
  for (int u = 0; u < kUUnroll; ++u) {
    const int64_t v = base + u * kUThreads;
    if (full || v < npack) {
      Pack<U, N> r;
#pragma unroll
      for (int i = 0; i < N; ++i) r.v[i] = c[u].v[i] ? a[u].v[i] : b[u].v[i];
      st_stream(out + v * N, r);
    }
  }
  if (blockIdx.x == gridDim.x - 1) {
    const int64_t t = npack * N + threadIdx.x;
    if (t < numel) out[t] = cond[t] ? x[t] : y[t];
  }
}
template <typename U>
__global__ void __launch_bounds__(kUThreads) where_scalar_kernel(const bool* cond, const U* x, const U* y, U* out,
                                                                 int64_t numel) {
  int64_t i = (int64_t)blockIdx.x * kUThreads + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * kUThreads;
  for (; i < numel; i += stride) out[i] = cond[i] ? x[i] : y[i];
}
template <typename U>
int launch_where(const void* cond, const void* x, const void* y, void* out, int64_t numel, cudaStream_t st) {
  constexpr int N = 16 / sizeof(U);
  const bool* c = static_cast<const bool*>(cond);
  const U* xp = static_cast<const U*>(x);
  const U* yp = static_cast<const U*>(y);
  U* op = static_cast<U*>(out);
  if (is_aligned16(x) && is_aligned16(y) && is_aligned16(out) && is_aligned_to(cond, N)) {
    const int64_t npack = numel / N;
    const int64_t tiles = std::max<int64_t>(1, (npack + kUThreads * kUUnroll - 1) / (kUThreads * kUUnroll));
    if (tiles > 0x7fffffffLL) return fail(PDB200_UNSUPPORTED, "where: tensor too large for one launch");
    where_kernel<U><<<(unsigned)tiles, kUThreads, 0, st>>>(c, xp, yp, op, npack, numel);
  } else {
    const int64_t blocks = std::min<int64_t>((numel + kUThreads - 1) / kUThreads, (int64_t)sm_count() * 32);
    where_scalar_kernel<U><<<(unsigned)blocks, kUThreads, 0, st>>>(c, xp, yp, op, numel);
  }
  return after_launch("where");
}

// MPTypeTrait: the type ScaleFunctor computes in
template <typename T> struct ScaleParam { using type = T; };
template <> struct ScaleParam<__nv_bfloat16> { using type = float; };
template <> struct ScaleParam<__half> { using type = float; };

}  // namespace pdb200

extern "C" {

#define PDB_FOR_ALL_DTYPES(M)                                                                                       \
  M(PDB200_BOOL) M(PDB200_UINT8) M(PDB200_INT8) M(PDB200_INT16) M(PDB200_INT32) M(PDB200_INT64) M(PDB200_FLOAT16)   \
  M(PDB200_BFLOAT16) M(PDB200_FLOAT32) M(PDB200_FLOAT64)

int pdb200_logical_not(const void* x, int64_t numel, int dtype, void* out, void* stream) {
  using namespace pdb200;
  PDB_CHECK_ARG(numel >= 0, "logical_not: negative element count");
  if (dtype == PDB200_UINT8 || dtype < 0 || dtype > PDB200_FLOAT64)  // logical_kernel.cu:112-129 has no uint8
    return fail(PDB200_UNSUPPORTED, "logical_not: dtype code " + std::to_string(dtype) + " is not registered for this op");
  if (numel == 0) return PDB200_OK;
  PDB_CHECK_ARG(x && out, "NULL tensor pointer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (dtype) {
#define M(DT)                                                                                                \
  case DT: {                                                                                                 \
    using T = typename CType<DT>::type;                                                                      \
    return launch_unary<T, bool>("logical_not", static_cast<const T*>(x), static_cast<bool*>(out), numel,    \
                                 LogicalNotF<T>(), st);                                                      \
  }
    PDB_FOR_ALL_DTYPES(M)
#undef M
  }
  return fail(PDB200_UNSUPPORTED, "logical_not: dtype not built");
}

int pdb200_bitwise_not(const void* x, int64_t numel, int dtype, void* out, void* stream) {
  using namespace pdb200;
  PDB_CHECK_ARG(numel >= 0, "bitwise_not: negative element count");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (numel > 0) PDB_CHECK_ARG(x && out, "NULL tensor pointer");
  switch (dtype) {  // bitwise_kernel.cu:127-136
#define M(DT)                                                                                                \
  case DT: {                                                                                                 \
    using T = typename CType<DT>::type;                                                                      \
    return launch_unary<T, T>("bitwise_not", static_cast<const T*>(x), static_cast<T*>(out), numel,          \
                              BitNotF<T>(), st);                                                             \
  }
    M(PDB200_BOOL) M(PDB200_UINT8) M(PDB200_INT8) M(PDB200_INT16) M(PDB200_INT32) M(PDB200_INT64)
#undef M
    default:
      return fail(PDB200_UNSUPPORTED, "bitwise_not: dtype code " + std::to_string(dtype) + " is not registered for this op");
  }
}

int pdb200_scale(const void* x, int64_t numel, int dtype, double scale, double bias, int bias_after_scale, void* out,
                 void* stream) {
  using namespace pdb200;
  PDB_CHECK_ARG(numel >= 0, "scale: negative element count");
  if (dtype < 0 || dtype > PDB200_FLOAT64)
    return fail(PDB200_UNSUPPORTED, "scale: dtype code " + std::to_string(dtype) + " is not registered for this op");
  if (numel == 0) return PDB200_OK;  // scale_kernel.cu:55-57
  PDB_CHECK_ARG(x && out, "NULL tensor pointer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (dtype) {
#define M(DT)                                                                                                \
  case DT: {                                                                                                 \
    using T = typename CType<DT>::type;                                                                      \
    using P = typename ScaleParam<T>::type;                                                                  \
    /* Scalar::to<P>(): a static_cast of the attribute */                                                    \
    const ScaleF<T, P> f{static_cast<P>(scale), static_cast<P>(bias), bias_after_scale != 0};                \
    return launch_unary<T, T>("scale", static_cast<const T*>(x), static_cast<T*>(out), numel, f, st);        \
  }
    PDB_FOR_ALL_DTYPES(M)
#undef M
  }
  return fail(PDB200_UNSUPPORTED, "scale: dtype not built");
}

int pdb200_where(const void* condition, const void* x, const void* y, int64_t numel, int dtype, void* out,
                 void* stream) {
  using namespace pdb200;
  PDB_CHECK_ARG(numel >= 0, "where: negative element count");
  const size_t es = dtype_size(dtype);
  if (es == 0) return fail(PDB200_UNSUPPORTED, "where: dtype code " + std::to_string(dtype) + " is not registered for this op");
  if (numel == 0) return PDB200_OK;
  PDB_CHECK_ARG(condition && x && y && out, "NULL tensor pointer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (es) {
    case 1: return launch_where<uint8_t>(condition, x, y, out, numel, st);
    case 2: return launch_where<uint16_t>(condition, x, y, out, numel, st);
    case 4: return launch_where<uint32_t>(condition, x, y, out, numel, st);
    default: return launch_where<uint64_t>(condition, x, y, out, numel, st);
  }
}

}  // extern "C"

namespace pdb200 {
template <typename TI>
int cast_from(const void* x, int64_t numel, int out_dtype, void* out, cudaStream_t st) {
  switch (out_dtype) {
#define M(DT)                                                                                                \
  case DT: {                                                                                                 \
    using TO = typename CType<DT>::type;                                                                     \
    return launch_unary<TI, TO>("cast", static_cast<const TI*>(x), static_cast<TO*>(out), numel,             \
                                CastF<TI, TO>(), st);                                                        \
  }
    PDB_FOR_ALL_DTYPES(M)
#undef M
  }
  return fail(PDB200_UNSUPPORTED, "cast: output dtype code " + std::to_string(out_dtype) + " is not registered for this op");
}
}  // namespace pdb200

extern "C" int pdb200_cast(const void* x, int64_t numel, int x_dtype, int out_dtype, void* out, void* stream) {
  using namespace pdb200;
  PDB_CHECK_ARG(numel >= 0, "cast: negative element count");
  if (x_dtype < 0 || x_dtype > PDB200_FLOAT64 || out_dtype < 0 || out_dtype > PDB200_FLOAT64)
    return fail(PDB200_UNSUPPORTED, "cast: dtype code " + std::to_string(x_dtype) + " -> " + std::to_string(out_dtype) +
                                        " is not registered for this op");
  if (numel == 0) return PDB200_OK;
  PDB_CHECK_ARG(x && out, "NULL tensor pointer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (x_dtype) {
#define M(DT) \
  case DT: return cast_from<typename CType<DT>::type>(x, numel, out_dtype, out, st);
    PDB_FOR_ALL_DTYPES(M)
#undef M
  }
  return fail(PDB200_UNSUPPORTED, "cast: dtype not built");
}
