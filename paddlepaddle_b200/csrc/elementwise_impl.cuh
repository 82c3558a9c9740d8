// elementwise_impl.cuh — binary elementwise ops with NumPy broadcasting, hand-written for sm_100a.
// (kernels, functors and the host launcher; elementwise.cu / elementwise_more.cu hold the entry points)
//
// Replaces (fresh design): funcs::BroadcastKernel + VectorizedBroadcastKernel
//   paddle/phi/kernels/funcs/broadcast_function.h:291-512,759-824
// and funcs::ElementwiseKernel / VectorizedElementwiseKernel
//   paddle/phi/kernels/funcs/elementwise_base.h:659-845
// with the functors of funcs/elementwise_functor.h:38-94.
//
// Design (memory-bound, no tensor cores — nothing here is a contraction):
//   * every dtype moves 16 bytes per load/store (8 x bf16, 4 x fp32, 2 x int64);
//   * the shape is classified on the host (shape.h) into four streams —
//       Same     : out[i] = x[i] (op) y[i]
//       Inner    : big[M,N] (op) row[N]   — the row index of each thread advances by a
//                  launch-time constant, so the loop carries one add+compare, no div/mod;
//       Outer    : big[M,N] (op) col[M]   — one scalar per row, CTA-uniform;
//       General  : merged N-D walk, one div/mod chain per 16-byte vector (not per element);
//   * each thread keeps UNROLL independent 16-byte loads per operand in flight;
//   * 64-bit element offsets everywhere (the reference slices >= 2^31 on the host,
//     broadcast_function.h:594-733).
#pragma once

#include "common.cuh"
#include "shape.h"

namespace pdb200 {

// ---- functors: funcs/elementwise_functor.h:38-94 --------------------------------------
// bf16/fp16: T(float(a) (op) float(b)) with one RNE rounding — phi/common/bfloat16.h:264-276.
template <typename T>
struct AddOp {
  static __device__ __forceinline__ T apply(T a, T b) { return a + b; }
};
template <typename T>
struct MulOp {
  static __device__ __forceinline__ T apply(T a, T b) { return a * b; }
};
template <> struct AddOp<bool> {  // bool + bool converts through int: true if either is set
  static __device__ __forceinline__ bool apply(bool a, bool b) { return a || b; }
};
template <> struct MulOp<bool> {  // MultiplyFunctor<bool> is logical AND (elementwise_functor.h:72-77)
  static __device__ __forceinline__ bool apply(bool a, bool b) { return a && b; }
};
template <> struct AddOp<__nv_bfloat16> {
  static __device__ __forceinline__ __nv_bfloat16 apply(__nv_bfloat16 a, __nv_bfloat16 b) {
    return __float2bfloat16_rn(__fadd_rn(__bfloat162float(a), __bfloat162float(b)));
  }
};
template <> struct MulOp<__nv_bfloat16> {
  static __device__ __forceinline__ __nv_bfloat16 apply(__nv_bfloat16 a, __nv_bfloat16 b) {
    return __float2bfloat16_rn(__fmul_rn(__bfloat162float(a), __bfloat162float(b)));
  }
};
template <> struct AddOp<__half> {
  static __device__ __forceinline__ __half apply(__half a, __half b) {
    return __float2half_rn(__fadd_rn(__half2float(a), __half2float(b)));
  }
};
template <> struct MulOp<__half> {
  static __device__ __forceinline__ __half apply(__half a, __half b) {
    return __float2half_rn(__fmul_rn(__half2float(a), __half2float(b)));
  }
};
// Signed integer overflow wraps (two's complement) exactly as the CPU build does; do the
// arithmetic unsigned so the compiler may not assume otherwise.
#define PDB_WRAP_INT(T, U)                                                                        \
  template <> struct AddOp<T> {                                                                   \
    static __device__ __forceinline__ T apply(T a, T b) { return (T)((U)a + (U)b); }              \
  };                                                                                              \
  template <> struct MulOp<T> {                                                                   \
    static __device__ __forceinline__ T apply(T a, T b) { return (T)((U)a * (U)b); }              \
  };
PDB_WRAP_INT(int8_t, uint8_t)
PDB_WRAP_INT(int16_t, uint16_t)
PDB_WRAP_INT(int32_t, uint32_t)
PDB_WRAP_INT(int64_t, uint64_t)
#undef PDB_WRAP_INT

// ---- siblings on the same engine: funcs/elementwise_functor.h:95-120,471-511 ----
// subtract / divide: a - b, a / b in T (bf16/fp16: one fp32 operation, one RNE rounding).
// maximum / minimum: a > b ? a : b, a < b ? a : b — NOT fmax/fmin: with a NaN operand the comparison is
// false and b is returned, exactly as the reference functor does.
template <typename T> struct SubOp {
  static __device__ __forceinline__ T apply(T a, T b) { return a - b; }
};
template <typename T> struct DivOp {
  static __device__ __forceinline__ T apply(T a, T b) { return a / b; }
};
template <typename T> struct MaximumOp {
  static __device__ __forceinline__ T apply(T a, T b) { return a > b ? a : b; }
};
template <typename T> struct MinimumOp {
  static __device__ __forceinline__ T apply(T a, T b) { return a < b ? a : b; }
};
#define PDB_HALF_OPS(T, TOF, FROMF)                                                                    \
  template <> struct SubOp<T> {                                                                        \
    static __device__ __forceinline__ T apply(T a, T b) { return FROMF(__fsub_rn(TOF(a), TOF(b))); }   \
  };                                                                                                   \
  template <> struct MaximumOp<T> {                                                                    \
    static __device__ __forceinline__ T apply(T a, T b) { return TOF(a) > TOF(b) ? a : b; }            \
  };                                                                                                   \
  template <> struct MinimumOp<T> {                                                                    \
    static __device__ __forceinline__ T apply(T a, T b) { return TOF(a) < TOF(b) ? a : b; }            \
  };
PDB_HALF_OPS(__nv_bfloat16, __bfloat162float, __float2bfloat16_rn)
PDB_HALF_OPS(__half, __half2float, __float2half_rn)
#undef PDB_HALF_OPS
template <> struct DivOp<__half> {  // 11-bit significands leave no such margin: the exact division
  static __device__ __forceinline__ __half apply(__half a, __half b) {
    return __float2half_rn(__fdiv_rn(__half2float(a), __half2float(b)));
  }
};
// bf16 divide without the IEEE division sequence, still bit-identical to round_bf16(a / b):
// the operands have 8-bit significands, a rounding boundary of the result (a midpoint between two bf16 values) has a
// 9-bit odd significand m, and a / b = m would need a = m * b, which has at least 9 significant bits — so the exact
// quotient is never a tie and sits at least 2^-17 (relative) from every boundary, while fa * rcp.approx(fb) is within ~2 ulp of
// fp32 (2^-22): the approximate quotient always rounds to the same bf16 as the exact one.  That holds for normal
// magnitudes; anything near the ends of the exponent range (and 0, inf, NaN) takes the exact division.
template <> struct DivOp<__nv_bfloat16> {
  static __device__ __forceinline__ __nv_bfloat16 apply(__nv_bfloat16 a, __nv_bfloat16 b) {
    const float fa = __bfloat162float(a), fb = __bfloat162float(b);
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(fb));  // 1 ulp; times fa: within ~2 ulp of fa / fb
    const float q = fa * r;
    const float ab = fabsf(fb), aq = fabsf(q);
    const bool safe = ab > 7.9e-31f && ab < 1.2e30f && aq > 7.9e-31f && aq < 1.2e30f;  // 2^-100 .. 2^100
    if (__builtin_expect(safe, 1)) return __float2bfloat16_rn(q);
    return __float2bfloat16_rn(__fdiv_rn(fa, fb));  // a real branch: the exact sequence must not run for every element
  }
};
// Integers: subtraction wraps; division truncates toward zero.  The reference's integer DivideFunctor
// raises on a zero divisor (PADDLE_ENFORCE, a device-side assert on GPU); a kernel cannot throw, so a zero
// divisor yields 0 here, and INT_MIN / -1 (undefined in C++) yields INT_MIN as two's-complement hardware does.
#define PDB_INT_SUBDIV(T, U)                                                                           \
  template <> struct SubOp<T> {                                                                        \
    static __device__ __forceinline__ T apply(T a, T b) { return (T)((U)a - (U)b); }                   \
  };                                                                                                   \
  template <> struct DivOp<T> {                                                                        \
    static __device__ __forceinline__ T apply(T a, T b) {                                              \
      if (b == 0) return (T)0;                                                                         \
      if (b == (T)-1) return (T)((U)0 - (U)a);                                                         \
      return (T)(a / b);                                                                               \
    }                                                                                                  \
  };
PDB_INT_SUBDIV(int8_t, uint8_t)
PDB_INT_SUBDIV(int16_t, uint16_t)
PDB_INT_SUBDIV(int32_t, uint32_t)
PDB_INT_SUBDIV(int64_t, uint64_t)
#undef PDB_INT_SUBDIV
template <> struct DivOp<uint8_t> {
  static __device__ __forceinline__ uint8_t apply(uint8_t a, uint8_t b) { return b == 0 ? (uint8_t)0 : (uint8_t)(a / b); }
};
template <> struct SubOp<bool> {
  static __device__ __forceinline__ bool apply(bool a, bool b) { return a != b; }
};
template <> struct DivOp<bool> {
  static __device__ __forceinline__ bool apply(bool a, bool b) { return a && b; }
};

// Result type of an operator on T: T itself for arithmetic, bool for compare / logical.
template <typename T, typename Op>
struct OpOut {
  using type = decltype(Op::apply(T(), T()));
};
// One 16-byte vector of each operand -> one pack of results.  The default applies the functor element by element;
// an operator may specialise VecApply for a dtype where a packed instruction does several elements at once
// (elementwise_pred.cu: 16-bit less_than / less_equal through the packed half compare).
template <typename T, typename Op>
struct VecApply {
  static __device__ __forceinline__ Pack<typename OpOut<T, Op>::type, Vec<T>::N> run(const Vec<T>& a, const Vec<T>& b) {
    Pack<typename OpOut<T, Op>::type, Vec<T>::N> r;
#pragma unroll
    for (int i = 0; i < Vec<T>::N; ++i) r.v[i] = Op::apply(a.v[i], b.v[i]);
    return r;
  }
};
template <typename T, typename Op>
__device__ __forceinline__ Pack<typename OpOut<T, Op>::type, Vec<T>::N> apply_vec(const Vec<T>& a, const Vec<T>& b) {
  return VecApply<T, Op>::run(a, b);
}
static __device__ __noinline__ float exact_div_f32(float a, float b) { return __fdiv_rn(a, b); }  // rare path, kept out of line

// bf16 divide, a vector at a time: eight approximate quotients and ONE branch (all eight in the safe range, or the whole
// vector through the exact division) instead of a divergent region per element
template <>
struct VecApply<__nv_bfloat16, DivOp<__nv_bfloat16>> {
  static __device__ __forceinline__ Pack<__nv_bfloat16, 8> run(const Vec<__nv_bfloat16>& a, const Vec<__nv_bfloat16>& b) {
    float fa[8], fb[8], q[8];
    unpack_to_float<__nv_bfloat16>(a, fa);
    unpack_to_float<__nv_bfloat16>(b, fb);
    // the range test on the smallest and the largest magnitude of the vector (two FMNMX with |.| operands per value instead
    // of four compares); a NaN slips through min / max, and may: the approximate quotient of a NaN is the same NaN
    float lo = INFINITY, hi = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float r;
      asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(fb[i]));
      q[i] = fa[i] * r;
      lo = fminf(lo, fminf(fabsf(fb[i]), fabsf(q[i])));
      hi = fmaxf(hi, fmaxf(fabsf(fb[i]), fabsf(q[i])));
    }
    const bool safe = lo > 7.9e-31f && hi < 1.2e30f;
    if (__builtin_expect(!safe, 0)) {
#pragma unroll
      for (int i = 0; i < 8; ++i) q[i] = exact_div_f32(fa[i], fb[i]);  // unrolled: the arrays must stay in registers
    }
    Vec<__nv_bfloat16> o;
    pack_from_float<__nv_bfloat16>(q, o);
    Pack<__nv_bfloat16, 8> r;
    *reinterpret_cast<uint4*>(&r) = *reinterpret_cast<const uint4*>(&o);
    return r;
  }
};


constexpr int kThreads = 256;
constexpr int kUnroll = 4;

// ---- Same shape ------------------------------------------------------------------------
// One CTA = one tile of kThreads*kUnroll vectors.  `nvec` full vectors, then a scalar tail.
template <typename T, typename Op>
__global__ void __launch_bounds__(kThreads, sizeof(T) >= 2 ? 4 : 1) ew_same_kernel(const T* __restrict__ x,
                                                           const T* __restrict__ y,
                                                           typename OpOut<T, Op>::type* __restrict__ out,
                                                           int64_t nvec, int64_t numel) {
  constexpr int V = Vec<T>::N;
  const int64_t base = (int64_t)blockIdx.x * (kThreads * kUnroll) + threadIdx.x;
  if (base + (kUnroll - 1) * kThreads < nvec) {  // full tile: all loads first, then math+stores
    Vec<T> a[kUnroll], b[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) a[u] = ld_stream(x + (base + u * kThreads) * V);
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) b[u] = ld_stream(y + (base + u * kThreads) * V);
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) st_stream(out + (base + u * kThreads) * V, apply_vec<T, Op>(a[u], b[u]));
  } else {
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int64_t v = base + u * kThreads;
      if (v < nvec) st_stream(out + v * V, apply_vec<T, Op>(ld_stream(x + v * V), ld_stream(y + v * V)));
    }
  }
  // scalar tail (numel not a multiple of the vector width): last CTA, first threads
  if (blockIdx.x == gridDim.x - 1) {
    const int64_t t = nvec * V + threadIdx.x;
    if (t < numel) out[t] = Op::apply(x[t], y[t]);
  }
}

// Unaligned / in-place-safe scalar fallback for Same.
template <typename T, typename Op>
__global__ void __launch_bounds__(kThreads) ew_same_scalar_kernel(const T* x, const T* y,
                                                                  typename OpOut<T, Op>::type* out, int64_t numel) {
  int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  for (; i < numel; i += stride) out[i] = Op::apply(x[i], y[i]);
}

// ---- Inner broadcast: big[M,N] (op) row[N] ---------------------------------------------
// The output is a 1-D stream of vectors; the row operand repeats with period P = N/V
// vectors.  Thread t of tile b handles vectors b*TILE + u*kThreads + t; its position in
// the row advances by (kThreads mod P) per step, so the modular index is carried with
// one add and one conditional subtract.  SWAP: the row operand is x (the left argument).
template <typename T, typename Op, bool SWAP>
__global__ void __launch_bounds__(kThreads, sizeof(T) >= 2 ? 4 : 1) ew_inner_kernel(const T* __restrict__ big,
                                                            const T* __restrict__ row,
                                                            typename OpOut<T, Op>::type* __restrict__ out,
                                                            int64_t nvec, uint32_t P, uint32_t step_mod) {
  constexpr int V = Vec<T>::N;
  const int64_t tile0 = (int64_t)blockIdx.x * (kThreads * kUnroll);
  const int64_t base = tile0 + threadIdx.x;
  uint32_t col = (uint32_t)(base % P);  // once per thread
  Vec<T> a[kUnroll], r[kUnroll];
  const bool full = base + (kUnroll - 1) * kThreads < nvec;
#pragma unroll
  for (int u = 0; u < kUnroll; ++u) {
    const int64_t v = base + u * kThreads;
    if (full || v < nvec) {
      a[u] = ld_stream(big + v * V);
      r[u] = ld_cached(row + (int64_t)col * V);
    }
    col += step_mod;
    if (col >= P) col -= P;
  }
#pragma unroll
  for (int u = 0; u < kUnroll; ++u) {
    const int64_t v = base + u * kThreads;
    if (full || v < nvec)
      st_stream(out + v * V, SWAP ? apply_vec<T, Op>(r[u], a[u]) : apply_vec<T, Op>(a[u], r[u]));
  }
}

// ---- Outer broadcast: big[M,N] (op) col[M] (also scalar: M == 1) ------------------------
// blockIdx.y = row, blockIdx.x tiles the row.  Rows need not be 16-byte multiples: VECTOR
// is chosen on the host only when N % V == 0 and the bases are aligned.
template <typename T, typename Op, bool SWAP, bool VECTOR>
__global__ void __launch_bounds__(kThreads) ew_outer_kernel(const T* __restrict__ big,
                                                            const T* __restrict__ colv,
                                                            typename OpOut<T, Op>::type* __restrict__ out,
                                                            int64_t rows, int64_t cols, int64_t tiles_per_row) {
  constexpr int V = VECTOR ? Vec<T>::N : 1;
  const int64_t row = (int64_t)blockIdx.x / tiles_per_row;
  const int64_t tile = (int64_t)blockIdx.x - row * tiles_per_row;
  const T s = colv[row];
  const int64_t n = cols / V;  // vectors (or elements) in this row
  const T* src = big + row * cols;
  typename OpOut<T, Op>::type* dst = out + row * cols;
  const int64_t base = tile * (kThreads * kUnroll) + threadIdx.x;
  if constexpr (VECTOR) {
    Vec<T> sv;
#pragma unroll
    for (int i = 0; i < V; ++i) sv.v[i] = s;
    Vec<T> a[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int64_t v = base + u * kThreads;
      if (v < n) a[u] = ld_stream(src + v * V);
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int64_t v = base + u * kThreads;
      if (v < n) st_stream(dst + v * V, SWAP ? apply_vec<T, Op>(sv, a[u]) : apply_vec<T, Op>(a[u], sv));
    }
  } else {
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int64_t v = base + u * kThreads;
      if (v < n) dst[v] = SWAP ? Op::apply(s, src[v]) : Op::apply(src[v], s);
    }
  }
}

// ---- General N-D ------------------------------------------------------------------------
struct GeneralParams {
  int rank;
  int64_t dims[PDB200_MAX_RANK];  // innermost first; dims[0] counts RUNS (vectors of V elements) in the kernel's view
  int64_t xs[PDB200_MAX_RANK];    // element strides per dim (0 = broadcast); xs[0] / ys[0] are per ELEMENT of the run
  int64_t ys[PDB200_MAX_RANK];
  int64_t inc[PDB200_MAX_RANK];   // the grid stride (gridDim * kThreads runs) written in the mixed radix of dims[]
  // The operand offsets move with the digits: one grid stride adds xinc = sum inc[d] * sx[d] (sx[0] = V * xs[0], sx[d] =
  // xs[d]), and a digit d that wraps (carry into d + 1) adds wx[d] = sx[d + 1] - dims[d] * sx[d] on top.
  int64_t xinc, yinc;
  int64_t wx[PDB200_MAX_RANK], wy[PDB200_MAX_RANK];
};

// One thread = one run of V output elements inside the innermost merged dim; inside the run each operand is either
// contiguous (stride 1: one 16-byte load when aligned) or a single broadcast value (stride 0).  Broadcast index math
// is resolved ONCE per thread: the div/mod chain runs for the thread's first run only; every further run is reached by
// adding the grid stride digit by digit with carry (`inc`, computed on the host) — adds and compares, no division in
// the loop.  (The first version ran a 64-bit div/mod chain per run: ~300 instructions per 16 bytes, 2.4 TB/s on a
// 4-D bf16 broadcast, profiles/r2_notes.txt.)
// RMAX: compile-time bound on the merged rank (4 covers almost every broadcast; 9 = DDim::kMaxRank).  IDX: type of the
// per-dim digits — uint32_t when every merged dim is below 2^31 (half the instructions of the 64-bit carry chain: the walk
// was issue-bound, 77 % issue slots at 45 % DRAM, profiles/r2_ew_general_bf16.txt), int64_t otherwise.
template <typename T, typename Op, bool VECTOR, int RMAX, typename IDX>
__global__ void __launch_bounds__(kThreads) ew_general_kernel(const T* __restrict__ x,
                                                              const T* __restrict__ y,
                                                              typename OpOut<T, Op>::type* __restrict__ out,
                                                              int64_t nruns, GeneralParams p) {
  constexpr int V = VECTOR ? Vec<T>::N : 1;
  int64_t run = (int64_t)blockIdx.x * kThreads + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  if (run >= nruns) return;
  // digits of the first run
  IDX idx[RMAX];
  {
    int64_t rest = run;
#pragma unroll
    for (int d = 0; d < RMAX; ++d) {
      idx[d] = 0;
      if (d < p.rank) {
        const int64_t q = rest / p.dims[d];
        idx[d] = (IDX)(rest - q * p.dims[d]);
        rest = q;
      }
    }
  }
  // offsets of the first run; afterwards they are carried along with the digits (adds and compares only: the first
  // carried-digit version still re-multiplied every digit by its 64-bit stride for both operands on every run — 130
  // instructions per 16-byte run, issue-bound)
  int64_t xo = (int64_t)idx[0] * V * p.xs[0], yo = (int64_t)idx[0] * V * p.ys[0];
#pragma unroll
  for (int d = 1; d < RMAX; ++d) {
    if (d < p.rank) {
      xo += (int64_t)idx[d] * p.xs[d];
      yo += (int64_t)idx[d] * p.ys[d];
    }
  }
  const bool xvec = p.xs[0] == 1, yvec = p.ys[0] == 1;
  for (; run < nruns; run += stride) {
    const int64_t oo = run * V;
    if constexpr (VECTOR) {
      Vec<T> a, b;
      if (xvec) {
        a = ld_stream(x + xo);
      } else {
        const T s = x[xo];
#pragma unroll
        for (int i = 0; i < V; ++i) a.v[i] = s;
      }
      if (yvec) {
        b = ld_stream(y + yo);
      } else {
        const T s = y[yo];
#pragma unroll
        for (int i = 0; i < V; ++i) b.v[i] = s;
      }
      st_stream(out + oo, apply_vec<T, Op>(a, b));
    } else {
      out[oo] = Op::apply(x[xo], y[yo]);
    }
    // advance by the grid stride: digit-wise add with carry, the offsets follow
    xo += p.xinc;
    yo += p.yinc;
    IDX carry = 0;
#pragma unroll
    for (int d = 0; d < RMAX; ++d) {
      if (d < p.rank) {
        IDX v = idx[d] + (IDX)p.inc[d] + carry;  // < 2 * dims[d]: no overflow below 2^31
        carry = 0;
        if (v >= (IDX)p.dims[d]) {
          v -= (IDX)p.dims[d];
          carry = 1;
          xo += p.wx[d];
          yo += p.wy[d];
        }
        idx[d] = v;
      }
    }
  }
}

// ---- host launcher ----------------------------------------------------------------------
template <typename T, typename Op>
int launch_elementwise(const BroadcastPlan& p, const T* x, const T* y, typename OpOut<T, Op>::type* out,
                       cudaStream_t st) {
  using O = typename OpOut<T, Op>::type;
  constexpr int V = Vec<T>::N;
  constexpr size_t kOutAlign = sizeof(O) * V;  // one output pack per input vector
  switch (p.kind) {
    case kBcastEmpty:
      return PDB200_OK;
    case kBcastSame: {
      const bool vec_ok = is_aligned16(x) && is_aligned16(y) && is_aligned_to(out, kOutAlign);
      if (vec_ok) {
        // In-place (out == x or y) is safe on the .nc path: each 16-byte chunk is read and
        // written by the same thread and never re-read.
        const int64_t nvec = p.numel / V;
        const int64_t tiles = std::max<int64_t>(1, (nvec + kThreads * kUnroll - 1) / (kThreads * kUnroll));
        ew_same_kernel<T, Op><<<(unsigned)tiles, kThreads, 0, st>>>(x, y, out, nvec, p.numel);
      } else {
        const int64_t blocks = std::min<int64_t>((p.numel + kThreads - 1) / kThreads, (int64_t)sm_count() * 32);
        ew_same_scalar_kernel<T, Op><<<(unsigned)blocks, kThreads, 0, st>>>(x, y, out, p.numel);
      }
      return after_launch("ew_same");
    }
    case kBcastInner: {
      const T* big = p.small_is_x ? y : x;
      const T* row = p.small_is_x ? x : y;
      const bool vec_ok = (p.cols % V == 0) && is_aligned16(big) && is_aligned16(row) &&
                          is_aligned_to(out, kOutAlign) && (p.cols / V) <= 0x7fffffffLL;
      if (!vec_ok) break;  // -> general
      const int64_t nvec = p.numel / V;
      const uint32_t P = (uint32_t)(p.cols / V);
      const uint32_t step_mod = (uint32_t)(kThreads % P);
      const int64_t tiles = (nvec + kThreads * kUnroll - 1) / (kThreads * kUnroll);
      if (p.small_is_x)
        ew_inner_kernel<T, Op, true><<<(unsigned)tiles, kThreads, 0, st>>>(big, row, out, nvec, P, step_mod);
      else
        ew_inner_kernel<T, Op, false><<<(unsigned)tiles, kThreads, 0, st>>>(big, row, out, nvec, P, step_mod);
      return after_launch("ew_inner");
    }
    case kBcastOuter: {
      const T* big = p.small_is_x ? y : x;
      const T* colv = p.small_is_x ? x : y;
      const bool vec_ok = (p.cols % V == 0) && is_aligned16(big) && is_aligned_to(out, kOutAlign);
      // rows shorter than one warp of vectors waste most of a CTA: use the general walk
      if (p.rows > 1 && p.cols < 32 * V) break;
      const int64_t n = vec_ok ? p.cols / V : p.cols;
      const int64_t tiles_per_row = (n + kThreads * kUnroll - 1) / (kThreads * kUnroll);
      const int64_t blocks = tiles_per_row * p.rows;
      if (blocks > 0x7fffffffLL) break;
#define PDB_OUTER(SW, VE) \
  ew_outer_kernel<T, Op, SW, VE><<<(unsigned)blocks, kThreads, 0, st>>>(big, colv, out, p.rows, p.cols, tiles_per_row)
      if (p.small_is_x) { if (vec_ok) PDB_OUTER(true, true); else PDB_OUTER(true, false); }
      else              { if (vec_ok) PDB_OUTER(false, true); else PDB_OUTER(false, false); }
#undef PDB_OUTER
      return after_launch("ew_outer");
    }
    default:
      break;
  }
  // General walk (also the fallback for unaligned Inner/Outer shapes).
  GeneralParams gp;
  gp.rank = p.rank;
  for (int i = 0; i < PDB200_MAX_RANK; ++i) {
    gp.dims[i] = i < p.rank ? p.dims[i] : 1;
    gp.xs[i] = i < p.rank ? p.xs[i] : 0;
    gp.ys[i] = i < p.rank ? p.ys[i] : 0;
    gp.inc[i] = 0;
  }
  // vector runs need: inner dim a multiple of V; every outer stride of a contiguous operand
  // is then a multiple of V automatically (strides are products of inner dims) as long as
  // that operand's innermost stride is 1; bases aligned.
  const bool vec_ok = (gp.dims[0] % V == 0) && is_aligned16(x) && is_aligned16(y) && is_aligned_to(out, kOutAlign);
  const int64_t nruns = vec_ok ? p.numel / V : p.numel;
  if (vec_ok) gp.dims[0] /= V;  // the kernel counts runs along the innermost dim
  if (gp.rank == 0) {           // a single element
    gp.rank = 1;
    gp.dims[0] = 1;
  }
  const int64_t blocks = std::min<int64_t>((nruns + kThreads - 1) / kThreads, (int64_t)sm_count() * 64);
  {
    int64_t rest = blocks * kThreads;  // the grid stride, in the mixed radix of dims[]
    for (int d = 0; d < gp.rank; ++d) {
      gp.inc[d] = rest % gp.dims[d];
      rest /= gp.dims[d];
    }
  }
  {
    const int64_t vv = vec_ok ? V : 1;
    gp.xinc = gp.yinc = 0;
    for (int d = 0; d < PDB200_MAX_RANK; ++d) {
      const int64_t sx = d == 0 ? gp.xs[0] * vv : gp.xs[d], sy = d == 0 ? gp.ys[0] * vv : gp.ys[d];
      const int64_t sx1 = d + 1 < gp.rank ? gp.xs[d + 1] : 0, sy1 = d + 1 < gp.rank ? gp.ys[d + 1] : 0;
      gp.wx[d] = gp.wy[d] = 0;
      if (d < gp.rank) {
        gp.xinc += gp.inc[d] * sx;
        gp.yinc += gp.inc[d] * sy;
        gp.wx[d] = sx1 - gp.dims[d] * sx;
        gp.wy[d] = sy1 - gp.dims[d] * sy;
      }
    }
  }
  bool small = gp.rank <= 4;
  for (int d = 0; d < gp.rank; ++d) small = small && gp.dims[d] < (int64_t(1) << 31);
  if (small) {
    if (vec_ok)
      ew_general_kernel<T, Op, true, 4, uint32_t><<<(unsigned)blocks, kThreads, 0, st>>>(x, y, out, nruns, gp);
    else
      ew_general_kernel<T, Op, false, 4, uint32_t><<<(unsigned)blocks, kThreads, 0, st>>>(x, y, out, nruns, gp);
  } else {
    if (vec_ok)
      ew_general_kernel<T, Op, true, PDB200_MAX_RANK, int64_t><<<(unsigned)blocks, kThreads, 0, st>>>(x, y, out, nruns, gp);
    else
      ew_general_kernel<T, Op, false, PDB200_MAX_RANK, int64_t><<<(unsigned)blocks, kThreads, 0, st>>>(x, y, out, nruns, gp);
  }
  return after_launch("ew_general");
}

// kMask: bit mask over pdb200_dtype of the dtypes the reference registers for this op on GPU
// (only those are instantiated).
template <template <typename> class OpT, uint32_t kMask>
int elementwise_entry(const char* name, const void* x, const int64_t* xd, int xr, const void* y,
                      const int64_t* yd, int yr, void* out, int dtype, void* stream) {
  PDB_CHECK_ARG(xr == 0 || xd != nullptr, "x_dims is NULL");
  PDB_CHECK_ARG(yr == 0 || yd != nullptr, "y_dims is NULL");
  BroadcastPlan p;
  std::string err = plan_broadcast(xd, xr, yd, yr, &p);
  if (!err.empty()) return fail(PDB200_INVALID_ARGUMENT, err);
  if (dtype < 0 || dtype > PDB200_FLOAT64 || !((kMask >> dtype) & 1u))
    return fail(PDB200_UNSUPPORTED, std::string(name) + ": dtype code " + std::to_string(dtype) +
                                        " is not registered for this op");
  if (p.kind == kBcastEmpty) return PDB200_OK;
  PDB_CHECK_ARG(x && y && out, "NULL tensor pointer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
#define PDB_CASE(DT)                                                                               \
  case DT: {                                                                                       \
    using T = typename CType<DT>::type;                                                            \
    if constexpr (((kMask >> DT) & 1u) != 0)                                                       \
      return launch_elementwise<T, OpT<T>>(p, static_cast<const T*>(x), static_cast<const T*>(y),  \
                                           static_cast<typename OpOut<T, OpT<T>>::type*>(out), st); \
    break;                                                                                         \
  }
  switch (dtype) {
    PDB_CASE(PDB200_BOOL)
    PDB_CASE(PDB200_UINT8)
    PDB_CASE(PDB200_INT8)
    PDB_CASE(PDB200_INT16)
    PDB_CASE(PDB200_INT32)
    PDB_CASE(PDB200_INT64)
    PDB_CASE(PDB200_FLOAT16)
    PDB_CASE(PDB200_BFLOAT16)
    PDB_CASE(PDB200_FLOAT32)
    PDB_CASE(PDB200_FLOAT64)
    default:
      break;
  }
#undef PDB_CASE
  return fail(PDB200_UNSUPPORTED, std::string(name) + ": dtype code " + std::to_string(dtype) + " is not built");
}

}  // namespace pdb200
