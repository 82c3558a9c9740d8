// reduce_impl.cuh — the reduction engine shared by sum/mean/max/min/prod and argmax.
//
// Replaces (fresh design): funcs::ReduceKernel / ReduceAnyKernel / ReduceHigherDimKernel /
// CubTensorReduce and ReduceConfig
//   paddle/phi/kernels/funcs/reduce_function.h:237-562,573-788,929-1198
// and the KPS block primitives it leans on (primitive/compute_primitives.h:52-129).
//
// Every reduction is executed as one or more passes over a 3-D view [outer][red][inner]:
//   inner == 1  -> row kernels   (reduce the contiguous axis)
//       rows_warp  : G lanes (1..32, power of two) per row, 16-byte loads, shuffle tree;
//       rows_block : one CTA per (row, split); `splits` > 1 cuts long rows / full-tensor
//                    reductions into partials that a second tiny pass folds — deterministic
//                    order, no atomics;
//   inner  > 1  -> cols kernel   (reduce a strided axis): threads run along `inner` with
//                  16-byte loads so every warp request is fully coalesced, TY thread-rows
//                  walk `red`, shared memory folds the TY partials.
// Differences from the reference that matter on B200: 16-byte loads for every dtype
// (bf16 gets 8 elements per load, the reference 4 ELEMENTS = 8 bytes), 256-512-thread CTAs
// instead of <= 128 (kReduceMaxThread, compute_primitives.h:35), kUnroll independent loads
// in flight per thread, 64-bit indexing (the reference's `int` dims overflow at 2^31).
#pragma once

#include <stdlib.h>

#include <limits>
#include <type_traits>

#include "common.cuh"
#include "shape.h"

namespace pdb200 {

// ---------------------------------------------------------------------------------------
// Operators.  Contract:
//   using Acc                      accumulator type
//   Acc  identity()
//   Acc  lift(T v, int64_t r)      element -> accumulator; r = index along the reduced axis
//   Acc  combine(Acc a, Acc b)     a's elements precede b's (matters for argmax ties only)
// identities follow primitive/functor_primitives.h:141-188.
// ---------------------------------------------------------------------------------------
template <typename A>
struct SumOp {
  using Acc = A;
  static __host__ __device__ __forceinline__ A identity() { return A(0); }
  template <typename T>
  static __device__ __forceinline__ A lift(T v, int64_t) { return cvt<A>(v); }
  static __device__ __forceinline__ A combine(A a, A b) { return a + b; }
};
template <> __device__ __forceinline__ int64_t SumOp<int64_t>::combine(int64_t a, int64_t b) {
  return (int64_t)((uint64_t)a + (uint64_t)b);  // wraps like the CPU build
}
template <> __device__ __forceinline__ int32_t SumOp<int32_t>::combine(int32_t a, int32_t b) {
  return (int32_t)((uint32_t)a + (uint32_t)b);
}

template <typename A>
struct ProdOp {
  using Acc = A;
  static __host__ __device__ __forceinline__ A identity() { return A(1); }
  template <typename T>
  static __device__ __forceinline__ A lift(T v, int64_t) { return cvt<A>(v); }
  static __device__ __forceinline__ A combine(A a, A b) { return a * b; }
};
template <> __device__ __forceinline__ int64_t ProdOp<int64_t>::combine(int64_t a, int64_t b) {
  return (int64_t)((uint64_t)a * (uint64_t)b);
}
template <> __device__ __forceinline__ int32_t ProdOp<int32_t>::combine(int32_t a, int32_t b) {
  return (int32_t)((uint32_t)a * (uint32_t)b);
}

template <typename A>
struct MaxOp {
  using Acc = A;
  static __host__ __device__ __forceinline__ A identity() { return std::numeric_limits<A>::lowest(); }
  template <typename T>
  static __device__ __forceinline__ A lift(T v, int64_t) { return cvt<A>(v); }
  static __device__ __forceinline__ A combine(A a, A b) { return (b > a) ? b : a; }  // functor_primitives.h:146-149
};
// The accumulator starts at lowest() and "(b > a) ? b : a" never lets a NaN in, so for floating point
// this is exactly fmax (which returns the non-NaN operand): one FMNMX instead of compare + select.
// (Only the sign of a zero result can differ, and that already depends on visiting order in the reference.)
template <> __device__ __forceinline__ float MaxOp<float>::combine(float a, float b) { return fmaxf(a, b); }
template <> __device__ __forceinline__ double MaxOp<double>::combine(double a, double b) { return fmax(a, b); }
template <typename A>
struct MinOp {
  using Acc = A;
  static __host__ __device__ __forceinline__ A identity() { return std::numeric_limits<A>::max(); }
  template <typename T>
  static __device__ __forceinline__ A lift(T v, int64_t) { return cvt<A>(v); }
  static __device__ __forceinline__ A combine(A a, A b) { return (b < a) ? b : a; }
};

template <> __device__ __forceinline__ float MinOp<float>::combine(float a, float b) { return fminf(a, b); }
template <> __device__ __forceinline__ double MinOp<double>::combine(double a, double b) { return fmin(a, b); }

// (value, index) pair for argmax.  16 bytes for every value type we instantiate.
template <typename V>
struct alignas(16) KV {
  V val;
  int64_t idx;
};
// Lowest index wins ties — Eigen's ArgMaxTupleReducer uses strict '>' (TensorFunctors.h:370-381)
// and so does CUB's ArgMax; combine() is order-independent because it compares indices too.
// MIN = true turns the same machinery into argmin: every value is mapped through an exactly order-REVERSING bijection
// first (floating point: negation; integers: bitwise NOT, which has no overflow corner) and the maximum of the mapped
// values is sought.  Ties, NaN (never wins a strict comparison) and rows on which nothing beats the start value
// (index 0) then behave as Eigen's ArgMinTupleReducer does (strict '<', TensorFunctors.h:383-394).
template <bool MIN, typename V>
__device__ __forceinline__ V arg_xform(V v) {
  if constexpr (!MIN) return v;
  else if constexpr (std::is_floating_point<V>::value) return -v;
  else return (V)~v;
}
template <typename V, bool MIN = false>
struct ArgMaxOp {
  using Acc = KV<V>;
  static constexpr bool kMin = MIN;
  static __host__ __device__ __forceinline__ Acc identity() {
    return Acc{std::numeric_limits<V>::lowest(), std::numeric_limits<int64_t>::max()};
  }
  template <typename T>
  static __device__ __forceinline__ Acc lift(T v, int64_t r) { return Acc{arg_xform<MIN>(cvt<V>(v)), r}; }
  static __device__ __forceinline__ Acc lift(Acc v, int64_t) { return v; }
  static __device__ __forceinline__ Acc combine(Acc a, Acc b) {
    return (b.val > a.val || (b.val == a.val && b.idx < a.idx)) ? b : a;
  }
};

// acc <- acc (+) element.  Every thread visits its elements in increasing index order, so argmax
// can use the cheap "replace on strictly greater" rule here (first occurrence wins) and keep the
// full index-aware combine() for merging accumulators.
template <typename Op, typename T>
__device__ __forceinline__ void accumulate(typename Op::Acc& a, T v, int64_t idx) {
  a = Op::combine(a, Op::template lift<T>(v, idx));
}
template <bool MIN, typename V, typename T>
__device__ __forceinline__ void accumulate_argmax(KV<V>& a, T v, int64_t idx) {
  const V x = arg_xform<MIN>(cvt<V>(v));
  if (x > a.val) {
    a.val = x;
    a.idx = idx;
  }
}

// Finalisers: accumulator -> output element.
template <typename A, typename O>
struct CastFin {
  __device__ __forceinline__ O operator()(A a) const { return cvt<O>(a); }
};
template <typename A, typename O>
struct MeanFin {  // sum / n at the very end (reduce_function.h:686-688); integer division truncates
  A n;
  __device__ __forceinline__ O operator()(A a) const { return cvt<O>(a / n); }
};
template <typename V, typename O>
struct IndexFin {
  // idx is still the identity when no element ever compared greater than lowest() (a row of -inf,
  // NaN or the type's minimum): Eigen's reducer starts at index 0 and never moves in that case.
  __device__ __forceinline__ O operator()(KV<V> a) const {
    return a.idx == std::numeric_limits<int64_t>::max() ? (O)0 : (O)a.idx;
  }
};

template <typename V> struct IsArgMax { static constexpr bool value = false; };
template <typename V, bool MIN> struct IsArgMax<ArgMaxOp<V, MIN>> { static constexpr bool value = true; };

template <typename Op, typename T>
__device__ __forceinline__ void acc_step(typename Op::Acc& a, T v, int64_t idx) {
  if constexpr (IsArgMax<Op>::value) {
    if constexpr (std::is_same<T, typename Op::Acc>::value) {
      a = Op::combine(a, v);  // folding partial (value, index) pairs: indices are not ordered
    } else {
      accumulate_argmax<Op::kMin>(a, v, idx);
    }
  } else {
    accumulate<Op, T>(a, v, idx);
  }
}

constexpr int kRedUnroll = 4;

}  // namespace pdb200
#include "reduce_cols_tma.cuh"  // needs acc_step / Vec; adds reduce_cols_tma_kernel
namespace pdb200 {

// ---- argmax over 16-bit values: track the best 16-BYTE VECTOR, not the best element ---------------
// Element-wise "replace on strictly greater" costs ~5 instructions per 2-byte element (unpack, compare,
// three selects) — more than the issue slots a 7 TB/s stream of bf16 leaves.  Instead: the maximum of a
// vector's 8 values takes 4 packed max instructions; only when it beats the running maximum (rare: the
// maximum of n random values rises ~ln n times) the vector and its position are remembered.  The element is
// picked from the remembered vector at the very end.  Result is the same as the element-wise rule: the
// first vector (in increasing index order) whose maximum equals the overall maximum, then the first equal
// element inside it; NaNs never win (packed max returns the other operand; '>' is false for NaN).
template <typename T> struct Pair16;
template <> struct Pair16<__nv_bfloat16> { using type = __nv_bfloat162; };
template <> struct Pair16<__half> { using type = __half2; };

template <typename T, typename Op>
__device__ __forceinline__ typename Op::Acc fold_span_argmax16(const T* __restrict__ row, int64_t begin, int64_t end,
                                                               int lane, int nlanes) {
  using A = typename Op::Acc;  // KV<float>
  using T2 = typename Pair16<T>::type;
  constexpr int V = Vec<T>::N;  // 8
  static_assert(V == 8, "16-bit vectors hold 8 values");
  float run[kRedUnroll];
  Vec<T> best[kRedUnroll];
  int64_t at[kRedUnroll];
#pragma unroll
  for (int u = 0; u < kRedUnroll; ++u) {
    run[u] = Op::identity().val;
    at[u] = Op::identity().idx;
    *reinterpret_cast<uint4*>(&best[u]) = make_uint4(0, 0, 0, 0);
  }
  auto visit = [&](int u, const Vec<T>& v, int64_t i) {
    const T2* p = reinterpret_cast<const T2*>(&v);
    // argmin: packed minimum, then negated — the running value lives in the mapped (negated) domain
    const T2 m = Op::kMin ? __hmin2(__hmin2(p[0], p[1]), __hmin2(p[2], p[3])) : __hmax2(__hmax2(p[0], p[1]), __hmax2(p[2], p[3]));
    const float f = Op::kMin ? -fminf(cvt<float>(m.x), cvt<float>(m.y)) : fmaxf(cvt<float>(m.x), cvt<float>(m.y));
    if (f > run[u]) {
      run[u] = f;
      best[u] = v;
      at[u] = i;
    }
  };
  const int64_t step = (int64_t)nlanes * V;
  int64_t i = begin + (int64_t)lane * V;
  for (; i + (kRedUnroll - 1) * step < end; i += kRedUnroll * step) {
    Vec<T> v[kRedUnroll];
#pragma unroll
    for (int u = 0; u < kRedUnroll; ++u) v[u] = ld_stream(row + i + u * step);
#pragma unroll
    for (int u = 0; u < kRedUnroll; ++u) visit(u, v[u], i + u * step);
  }
  for (; i < end; i += step) visit(0, ld_stream(row + i), i);
  // pick the element inside each remembered vector (last to first, so the first equal one sticks)
  A acc[kRedUnroll];
#pragma unroll
  for (int u = 0; u < kRedUnroll; ++u) {
    acc[u] = Op::identity();
    if (at[u] != Op::identity().idx) {
      int k_first = 0;
#pragma unroll
      for (int k = V - 1; k >= 0; --k)
        if (arg_xform<Op::kMin>(cvt<float>(best[u].v[k])) == run[u]) k_first = k;
      acc[u] = A{run[u], at[u] + k_first};
    }
  }
  return Op::combine(Op::combine(acc[0], acc[1]), Op::combine(acc[2], acc[3]));
}

// Per-thread streaming fold of elements [begin, end) of one row, this thread touching
// chunks begin + (lane + k*nlanes)*V.  VECTOR: 16-byte loads (caller guarantees alignment
// and that begin, end are multiples of V).
template <typename T, typename Op, bool VECTOR>
__device__ __forceinline__ typename Op::Acc fold_span(const T* __restrict__ row, int64_t begin,
                                                      int64_t end, int lane, int nlanes) {
  using A = typename Op::Acc;
  if constexpr (VECTOR && IsArgMax<Op>::value && sizeof(T) == 2 &&
                (std::is_same<T, __nv_bfloat16>::value || std::is_same<T, __half>::value)) {
    return fold_span_argmax16<T, Op>(row, begin, end, lane, nlanes);
  }
  constexpr int V = VECTOR ? Vec<T>::N : 1;
  A acc[kRedUnroll];
#pragma unroll
  for (int u = 0; u < kRedUnroll; ++u) acc[u] = Op::identity();
  const int64_t step = (int64_t)nlanes * V;
  int64_t i = begin + (int64_t)lane * V;
  if constexpr (VECTOR) {
    for (; i + (kRedUnroll - 1) * step < end; i += kRedUnroll * step) {
      Vec<T> v[kRedUnroll];
#pragma unroll
      for (int u = 0; u < kRedUnroll; ++u) v[u] = ld_stream(row + i + u * step);
#pragma unroll
      for (int u = 0; u < kRedUnroll; ++u)
#pragma unroll
        for (int k = 0; k < V; ++k) acc_step<Op, T>(acc[u], v[u].v[k], i + u * step + k);
    }
    for (; i < end; i += step) {
      Vec<T> v = ld_stream(row + i);
#pragma unroll
      for (int k = 0; k < V; ++k) acc_step<Op, T>(acc[0], v.v[k], i + k);
    }
  } else {
    for (; i + (kRedUnroll - 1) * step < end; i += kRedUnroll * step) {
      T v[kRedUnroll];
#pragma unroll
      for (int u = 0; u < kRedUnroll; ++u) v[u] = row[i + u * step];
#pragma unroll
      for (int u = 0; u < kRedUnroll; ++u) acc_step<Op, T>(acc[u], v[u], i + u * step);
    }
    for (; i < end; i += step) acc_step<Op, T>(acc[0], row[i], i);
  }
  // elements of acc[0..3] interleave in index order only loosely; combine() must be (and is)
  // insensitive to that: sums are re-associated, max/min/argmax are order-free by construction.
  return Op::combine(Op::combine(acc[0], acc[1]), Op::combine(acc[2], acc[3]));
}

// ---- rows_warp: G lanes per row -------------------------------------------------------
template <typename T, typename O, typename Op, typename Fin, bool VECTOR>
__global__ void __launch_bounds__(256) reduce_rows_warp_kernel(const T* __restrict__ x,
                                                               O* __restrict__ out, int64_t outer,
                                                               int64_t red, int G, Fin fin) {
  using A = typename Op::Acc;
  const int rows_per_cta = blockDim.x / G;
  const int g = threadIdx.x / G, lane = threadIdx.x - g * G;
  const int64_t row_stride = (int64_t)gridDim.x * rows_per_cta;
  // The trip count is CTA-uniform (every lane takes part in the full-mask shuffles); groups
  // past the last row fold nothing and keep the identity.
  for (int64_t row0 = (int64_t)blockIdx.x * rows_per_cta; row0 < outer; row0 += row_stride) {
    const int64_t row = row0 + g;
    A a = Op::identity();
    if (row < outer) a = fold_span<T, Op, VECTOR>(x + row * red, 0, red, lane, G);
    for (int m = G >> 1; m >= 1; m >>= 1) a = Op::combine(a, shfl_xor_any(a, m));
    if (lane == 0 && row < outer) out[row] = fin(a);
  }
}

// ---- rows_block: one CTA per (row, split) -----------------------------------------------
// out[row * splits + split]; with splits > 1 O is the accumulator type and Fin the identity.
// INTERLEAVE (vector path, splits > 1): the row is cut into tiles of blockDim * V * kRedUnroll elements and split s
// folds tiles s, s + splits, s + 2*splits, ...  All CTAs then sweep the row together, so at any moment the
// machine reads one ~20 MB window instead of ~600 far-apart streams: with contiguous segments a 17 GB row
// ran 8% slower than a 4 GB one (TLB reach / DRAM page locality; profiles/r1e_notes.txt).  The partial of a
// split is still a fixed, run-to-run identical sum.  The ragged end of the row goes to one split, after its tiles.
template <typename T, typename O, typename Op, typename Fin, bool VECTOR, bool INTERLEAVE>
__global__ void __launch_bounds__(512) reduce_rows_block_kernel(const T* __restrict__ x,
                                                                O* __restrict__ out, int64_t red,
                                                                int64_t seg, int splits, Fin fin) {
  using A = typename Op::Acc;
  __shared__ A smem[33];
  const int64_t row = (int64_t)blockIdx.x / splits;
  const int split = (int)((int64_t)blockIdx.x - row * splits);
  A a;
  if constexpr (VECTOR && INTERLEAVE) {
    constexpr int V = Vec<T>::N;
    const int64_t tile = (int64_t)blockDim.x * V * kRedUnroll;
    const int64_t full_tiles = red / tile;
    const T* base = x + row * red + (int64_t)threadIdx.x * V;
    const int64_t step = (int64_t)blockDim.x * V;
    A acc[kRedUnroll];
#pragma unroll
    for (int u = 0; u < kRedUnroll; ++u) acc[u] = Op::identity();
    if constexpr (IsArgMax<Op>::value && sizeof(T) == 2) {
      // 16-bit argmax keeps its vector-granular bookkeeping inside fold_span: go tile by tile through it
      for (int64_t t = split; t < full_tiles; t += splits)
        acc[0] = Op::combine(acc[0], fold_span<T, Op, true>(x + row * red, t * tile, (t + 1) * tile, threadIdx.x, blockDim.x));
    } else {
      for (int64_t t = split; t < full_tiles; t += splits) {
        const int64_t i0 = t * tile;
        Vec<T> v[kRedUnroll];
#pragma unroll
        for (int u = 0; u < kRedUnroll; ++u) v[u] = ld_stream(base + i0 + u * step);
#pragma unroll
        for (int u = 0; u < kRedUnroll; ++u)
#pragma unroll
          for (int k = 0; k < V; ++k) acc_step<Op, T>(acc[u], v[u].v[k], i0 + (int64_t)threadIdx.x * V + u * step + k);
      }
    }
    a = Op::combine(Op::combine(acc[0], acc[1]), Op::combine(acc[2], acc[3]));
    if (split == (int)(full_tiles % splits) && full_tiles * tile < red)
      a = Op::combine(a, fold_span<T, Op, true>(x + row * red, full_tiles * tile, red, threadIdx.x, blockDim.x));
  } else {
    const int64_t begin = (int64_t)split * seg;
    const int64_t end = begin + seg < red ? begin + seg : red;
    a = fold_span<T, Op, VECTOR>(x + row * red, begin, end, threadIdx.x, blockDim.x);
  }
  a = block_reduce(a, [](A p, A q) { return Op::combine(p, q); }, Op::identity(), smem);
  if (threadIdx.x == 0) out[blockIdx.x] = fin(a);
}

// ---- cols: reduce the middle axis of [outer][red][inner] --------------------------------
// blockDim = (TX, TY).  Thread (tx, ty) owns V adjacent columns and rows ty, ty+TY, ...
// of its split.  Partials land in out[((o*splits)+s)*inner + col].
template <typename T, typename O, typename Op, typename Fin, bool VECTOR>
__global__ void __launch_bounds__(256) reduce_cols_kernel(const T* __restrict__ x,
                                                          O* __restrict__ out, int64_t red,
                                                          int64_t inner, int64_t seg, int splits,
                                                          int64_t col_tiles, Fin fin) {
  using A = typename Op::Acc;
  constexpr int V = VECTOR ? Vec<T>::N : 1;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  A* smem = reinterpret_cast<A*>(smem_raw);  // [TY][TX*V]
  const int TX = blockDim.x, TY = blockDim.y;
  int64_t b = blockIdx.x;
  const int64_t ct = b % col_tiles;
  b /= col_tiles;
  const int split = (int)(b % splits);
  const int64_t o = b / splits;
  const int64_t col = (ct * TX + threadIdx.x) * V;
  const bool active = col < inner;
  const int64_t begin = (int64_t)split * seg;
  const int64_t end = begin + seg < red ? begin + seg : red;
  A acc[V];
#pragma unroll
  for (int k = 0; k < V; ++k) acc[k] = Op::identity();
  if (active) {
    const T* base = x + o * red * inner + col;
    int64_t r = begin + threadIdx.y;
    if constexpr (VECTOR) {
      for (; r + (kRedUnroll - 1) * TY < end; r += kRedUnroll * TY) {
        Vec<T> v[kRedUnroll];
#pragma unroll
        for (int u = 0; u < kRedUnroll; ++u) v[u] = ld_stream(base + (r + u * TY) * inner);
#pragma unroll
        for (int u = 0; u < kRedUnroll; ++u)
#pragma unroll
          for (int k = 0; k < V; ++k) acc_step<Op, T>(acc[k], v[u].v[k], r + u * TY);
      }
      for (; r < end; r += TY) {
        Vec<T> v = ld_stream(base + r * inner);
#pragma unroll
        for (int k = 0; k < V; ++k) acc_step<Op, T>(acc[k], v.v[k], r);
      }
    } else {
      for (; r + (kRedUnroll - 1) * TY < end; r += kRedUnroll * TY) {
        T v[kRedUnroll];
#pragma unroll
        for (int u = 0; u < kRedUnroll; ++u) v[u] = base[(r + u * TY) * inner];
#pragma unroll
        for (int u = 0; u < kRedUnroll; ++u) acc_step<Op, T>(acc[0], v[u], r + u * TY);
      }
      for (; r < end; r += TY) acc_step<Op, T>(acc[0], base[r * inner], r);
    }
  }
  // fold the TY partials of each column through shared memory (tree over ty)
#pragma unroll
  for (int k = 0; k < V; ++k) smem[(threadIdx.y * TX + threadIdx.x) * V + k] = acc[k];
  __syncthreads();
  for (int s = TY >> 1; s >= 1; s >>= 1) {
    if ((int)threadIdx.y < s) {
#pragma unroll
      for (int k = 0; k < V; ++k) {
        A& mine = smem[(threadIdx.y * TX + threadIdx.x) * V + k];
        mine = Op::combine(mine, smem[((threadIdx.y + s) * TX + threadIdx.x) * V + k]);
      }
    }
    __syncthreads();
  }
  if (threadIdx.y == 0 && active) {
    O* dst = out + (o * splits + split) * inner + col;
#pragma unroll
    for (int k = 0; k < V; ++k) dst[k] = fin(smem[threadIdx.x * V + k]);
  }
}

// ---------------------------------------------------------------------------------------
// Host side: launch geometry for one pass.
// ---------------------------------------------------------------------------------------
struct PassGeom {
  int kind;        // 0 rows_warp, 1 rows_block, 2 cols, 3 cols staged by TMA tensor-map boxes
  int G;           // rows_warp: lanes per row
  int threads;     // rows_block
  int splits;      // rows_block / cols: cuts of the reduced axis (1 = single pass)
  int64_t seg;     // elements of the reduced axis per split
  int TX, TY;      // cols
  int64_t col_tiles;
};

inline int pow2_ceil(int64_t v, int cap) {
  int p = 1;
  while (p < v && p < cap) p <<= 1;
  return p;
}

// elem_bytes: sizeof(T) of the pass input; V: vector width in elements (1 when unaligned).
// acc_bytes: sizeof the accumulator (16 for argmax's (value, index) pairs).
inline PassGeom plan_pass(const ReducePass& ps, size_t elem_bytes, int V, size_t acc_bytes = 8) {
  PassGeom g{};
  const int sms = sm_count();
  const int64_t target_ctas = (int64_t)sms * 4;
  if (ps.inner == 1) {
    const int64_t row_bytes = ps.red * (int64_t)elem_bytes;
    const bool many_rows = ps.outer >= target_ctas;
    // rows up to 16 KB go to the warp kernel; (value, index) accumulators up to 64 KB: folding 16-byte pairs across a CTA
    // through shared memory costs more than the longer per-lane loop (argmax fp32 [65536,8192]: 0.357 -> 0.295 ms, 7.3 TB/s;
    // sum / max of the same rows are equal or slightly better in the block kernel: profiles/r2_notes.txt)
    const int64_t warp_max = knobs().rows_warp_max_bytes > 0 ? knobs().rows_warp_max_bytes : (acc_bytes >= 16 ? 65536 : 16384);
    if (row_bytes <= warp_max && (many_rows || row_bytes <= 2048)) {
      g.kind = 0;
      // aim for >= 4 vectors per lane so kRedUnroll loads are in flight
      g.G = pow2_ceil((ps.red / V + 3) / 4, 32);
      g.splits = 1;
      g.seg = ps.red;
      return g;
    }
    g.kind = 1;
    g.threads = row_bytes >= 65536 ? 512 : 256;
    // cut rows only when there are too few of them to fill the machine
    int64_t splits = 1;
    if (ps.outer < target_ctas) {
      const int64_t per_iter = (int64_t)g.threads * V * kRedUnroll;  // elements one CTA folds per loop trip
      const int64_t max_by_work = std::max<int64_t>(1, ps.red / (per_iter * 2));
      splits = std::min<int64_t>(std::max<int64_t>(1, (target_ctas * 2 + ps.outer - 1) / ps.outer), max_by_work);
      splits = std::min<int64_t>(splits, 65535);
    }
    int64_t seg = (ps.red + splits - 1) / splits;
    const int64_t align = (int64_t)V * 1;  // keep split starts vector-aligned
    seg = (seg + align - 1) / align * align;
    splits = (ps.red + seg - 1) / seg;
    g.splits = (int)splits;
    g.seg = seg;
    return g;
  }
  g.kind = 2;
  const int64_t vec_cols = (ps.inner + V - 1) / V;
  // Wide enough, long enough, and an element size the tensor map knows: stage [32 rows][512 B] boxes with TMA.
  // The tile (32 vectors x 8 row-groups) is also a valid geometry for the plain kernel, which launch_pass
  // falls back to if the tensor map cannot be encoded.
  const bool tma = V > 1 && (elem_bytes == 2 || elem_bytes == 4 || elem_bytes == 8) && vec_cols >= 32 &&
                   ps.red >= 64 && !knobs().no_tma_cols;
  if (tma) {
    g.kind = 3;
    g.TX = 32;
    g.TY = 8;
  } else {
    g.TX = pow2_ceil(vec_cols, 64);
    if (g.TX < 1) g.TX = 1;
    g.TY = 256 / g.TX;
  }
  g.col_tiles = (vec_cols + g.TX - 1) / g.TX;
  const int64_t ctas = g.col_tiles * ps.outer;
  int64_t splits = 1;
  if (ctas < target_ctas) {
    const int64_t max_by_work = std::max<int64_t>(1, ps.red / ((int64_t)g.TY * kRedUnroll * 2));
    splits = std::min<int64_t>(std::max<int64_t>(1, (target_ctas * 2 + ctas - 1) / ctas), max_by_work);
    splits = std::min<int64_t>(splits, 65535);
  }
  int64_t seg = (ps.red + splits - 1) / splits;
  splits = (ps.red + seg - 1) / seg;
  g.splits = (int)splits;
  g.seg = seg;
  return g;
}

// Launch one pass: x[outer][red][inner] -> out[outer][splits][inner] (splits from g).
template <typename T, typename O, typename Op, typename Fin>
int launch_pass(const ReducePass& ps, const PassGeom& g, bool vec, const T* x, O* out, Fin fin,
                cudaStream_t st) {
  if (g.kind == 0) {
    const int rows_per_cta = 256 / g.G;
    int64_t blocks = (ps.outer + rows_per_cta - 1) / rows_per_cta;
    blocks = std::min<int64_t>(blocks, (int64_t)sm_count() * 64);
    if (vec)
      reduce_rows_warp_kernel<T, O, Op, Fin, true><<<(unsigned)blocks, 256, 0, st>>>(x, out, ps.outer, ps.red, g.G, fin);
    else
      reduce_rows_warp_kernel<T, O, Op, Fin, false><<<(unsigned)blocks, 256, 0, st>>>(x, out, ps.outer, ps.red, g.G, fin);
    return after_launch("reduce_rows_warp");
  }
  if (g.kind == 1) {
    const int64_t blocks = ps.outer * g.splits;
    if (blocks > 0x7fffffffLL) return fail(PDB200_UNSUPPORTED, "reduce: too many rows for one launch");
    const bool no_interleave = knobs().no_interleave;  // tuning knob
    if (vec && g.splits > 1 && !no_interleave)
      reduce_rows_block_kernel<T, O, Op, Fin, true, true><<<(unsigned)blocks, g.threads, 0, st>>>(x, out, ps.red, g.seg, g.splits, fin);
    else if (vec)
      reduce_rows_block_kernel<T, O, Op, Fin, true, false><<<(unsigned)blocks, g.threads, 0, st>>>(x, out, ps.red, g.seg, g.splits, fin);
    else
      reduce_rows_block_kernel<T, O, Op, Fin, false, false><<<(unsigned)blocks, g.threads, 0, st>>>(x, out, ps.red, g.seg, g.splits, fin);
    return after_launch("reduce_rows_block");
  }
  const int64_t blocks = g.col_tiles * g.splits * ps.outer;
  if (blocks > 0x7fffffffLL) return fail(PDB200_UNSUPPORTED, "reduce: too many tiles for one launch");
  if constexpr (sizeof(T) == 2 || sizeof(T) == 4 || sizeof(T) == 8) {
    if (g.kind == 3 && vec) {
      CUtensorMap map;
      if (make_cols_tensor_map(&map, x, ps.outer * ps.red, ps.inner, sizeof(T))) {
        using A = typename Op::Acc;
        constexpr int TI = 32 * Vec<T>::N;
        const size_t smem_tma = (size_t)kTmaStages * kTmaTileRows * TI * sizeof(T) + (size_t)8 * TI * sizeof(A);
        auto kern = reduce_cols_tma_kernel<T, O, Op, Fin>;
        // one size per instantiation: set once per (kernel, device) under the attribute lock, a map lookup afterwards
        FuncAttrHold hold;
        func_attr_apply(reinterpret_cast<const void*>(kern), (int)smem_tma, -1, &hold);
        const int ks = knobs().tma_cols_stages;  // tuning knob; the ring is allocated for kTmaStages
        const int nstages = ks >= 2 && ks <= kTmaStages ? ks : kTmaStages;
        kern<<<(unsigned)blocks, kTmaThreads, smem_tma, st>>>(map, out, ps.red, ps.inner, g.seg, g.splits, g.col_tiles, fin, nstages);
        return after_launch("reduce_cols_tma");
      }
    }
  }
  dim3 block(g.TX, g.TY);
  const int V = vec ? Vec<T>::N : 1;
  const size_t smem = sizeof(typename Op::Acc) * (size_t)g.TX * g.TY * V;
  if (vec) {
    // argmax over 1-byte values folds 16 columns x 16-byte (value, index) pairs per thread: 64 KB, above the 48 KB a
    // kernel gets without opting in
    FuncAttrHold hold;
    if (smem > 48 * 1024 &&
        !func_attr_apply(reinterpret_cast<const void*>(reduce_cols_kernel<T, O, Op, Fin, true>), (int)smem, -1, &hold))
      return fail(PDB200_CUDA_ERROR, "reduce_cols: cannot opt in to " + std::to_string(smem) + " bytes of shared memory");
    reduce_cols_kernel<T, O, Op, Fin, true><<<(unsigned)blocks, block, smem, st>>>(x, out, ps.red, ps.inner, g.seg, g.splits, g.col_tiles, fin);
  } else
    reduce_cols_kernel<T, O, Op, Fin, false><<<(unsigned)blocks, block, smem, st>>>(x, out, ps.red, ps.inner, g.seg, g.splits, g.col_tiles, fin);
  return after_launch("reduce_cols");
}

// ---------------------------------------------------------------------------------------
// A plan becomes a flat list of launches.  A pass whose reduced axis is cut into `splits`
// contributes two launches (partials, then the fold over [outer][splits][inner]).  Launch k
// reads x (k == 0) or scratch region (k-1)&1 and writes scratch region k&1, except the last
// launch, which writes `out` through the final finaliser.  Intermediates are kept in the
// accumulator type, so precision matches a single-pass reduction.
// ---------------------------------------------------------------------------------------
struct Stage {
  ReducePass ps;
  PassGeom g;
  bool vec;
};

// in_aligned: whether the first-stage input pointer is 16-byte aligned (scratch always is).
template <typename T, typename A>
inline std::vector<Stage> build_stages(const ReducePlan& plan, bool in_aligned) {
  std::vector<Stage> st;
  for (size_t i = 0; i < plan.passes.size(); ++i) {
    const ReducePass& ps = plan.passes[i];
    const bool first = st.empty();
    const size_t eb = first ? sizeof(T) : sizeof(A);
    const int Vfull = (int)(16 / eb);
    bool vec = (first ? in_aligned : true) && ((ps.inner == 1 ? ps.red : ps.inner) % Vfull == 0);
    Stage s{ps, plan_pass(ps, eb, vec ? Vfull : 1, sizeof(A)), vec};
    st.push_back(s);
    if (s.g.splits > 1) {
      ReducePass fold{ps.outer, s.g.splits, ps.inner};
      const int Va = (int)(16 / sizeof(A));
      const bool fvec = (fold.inner == 1 ? fold.red : fold.inner) % Va == 0;
      PassGeom fg = plan_pass(fold, sizeof(A), fvec ? Va : 1);
      fg.splits = 1;  // a few thousand partials per output never need cutting again
      fg.seg = fold.red;
      st.push_back(Stage{fold, fg, fvec});
    }
  }
  return st;
}

inline size_t stages_scratch_elems(const std::vector<Stage>& st) {
  size_t worst = 0;
  for (size_t k = 0; k + 1 < st.size(); ++k)
    worst = std::max(worst, (size_t)(st[k].ps.outer * st[k].g.splits * st[k].ps.inner));
  return worst;
}

// Bytes of scratch for a plan (both alignment cases considered: the query has no pointer).
template <typename T, typename A>
inline size_t reduce_scratch_bytes(const ReducePlan& plan) {
  size_t e = std::max(stages_scratch_elems(build_stages<T, A>(plan, true)),
                      stages_scratch_elems(build_stages<T, A>(plan, false)));
  if (e == 0) return 0;
  const size_t each = (e * sizeof(A) + 255) / 256 * 256;
  return each * 2;
}

// Run a whole plan.  T: input element, O: output element, Op::Acc: accumulator.
template <typename T, typename O, typename Op, typename FinLast>
int run_reduce(const ReducePlan& plan, const T* x, O* out, FinLast fin_last, void* workspace,
               size_t workspace_bytes, cudaStream_t stream) {
  using A = typename Op::Acc;
  std::vector<Stage> st = build_stages<T, A>(plan, is_aligned16(x));
  const size_t e = stages_scratch_elems(st);
  const size_t each = (e * sizeof(A) + 255) / 256 * 256;
  if (e > 0) {
    if (workspace == nullptr || each * 2 > workspace_bytes)
      return fail(PDB200_WORKSPACE_TOO_SMALL, "reduce: workspace has " + std::to_string(workspace_bytes) +
                                                  " bytes, this call needs " + std::to_string(each * 2) +
                                                  " (ask pdb200_reduce_workspace_bytes)");
    if (!is_aligned16(workspace)) return fail(PDB200_INVALID_ARGUMENT, "reduce: workspace must be 16-byte aligned");
  }
  A* buf[2] = {reinterpret_cast<A*>(workspace),
               reinterpret_cast<A*>(static_cast<char*>(workspace) + each)};
  for (size_t k = 0; k < st.size(); ++k) {
    const Stage& s = st[k];
    const bool last = (k + 1 == st.size());
    int rc;
    if (k == 0) {
      if (last) rc = launch_pass<T, O, Op, FinLast>(s.ps, s.g, s.vec, x, out, fin_last, stream);
      else rc = launch_pass<T, A, Op, CastFin<A, A>>(s.ps, s.g, s.vec, x, buf[0], CastFin<A, A>(), stream);
    } else {
      const A* in = buf[(k - 1) & 1];
      if (last) rc = launch_pass<A, O, Op, FinLast>(s.ps, s.g, s.vec, in, out, fin_last, stream);
      else rc = launch_pass<A, A, Op, CastFin<A, A>>(s.ps, s.g, s.vec, in, buf[k & 1], CastFin<A, A>(), stream);
    }
    if (rc != PDB200_OK) return rc;
  }
  return PDB200_OK;
}

}  // namespace pdb200
