// reduce_grad.cu — backward of the reductions.  sum / mean: the gradient is out_grad broadcast
// back over the reduced dimensions (mean: times 1/n); max / min / prod (second half of the file): the same broadcast
// combined with x and the forward result.
//
// Replaces (fresh design): phi::ReduceSumGradKernel / ReduceMeanGradKernel on GPU,
//   paddle/phi/kernels/gpu/reduce_kernel.cu:40-76 (sum_grad), :122-155 (mean_grad) — a one-input
//   funcs::BroadcastKernel with kps::IdentityFunctor / kps::DivideFunctor (primitive/functor_primitives.h:89-105).
// A write-only stream: out_grad is 1/n-th of the traffic, so the kernels are organised around 16-byte streaming
// stores of x_grad — the shape is classified on the host exactly as for the binary ops (shape.h):
//   same  : no dimension was reduced (or all reduced dims have size 1): scaled copy
//   inner : leading dims reduced — x_grad[m][n] = f(out_grad[n]): each thread keeps its 16-byte piece of the row and
//           stores it down the rows;
//   outer : trailing dims reduced (incl. reduce_all) — x_grad[m][n] = f(out_grad[m]): one value per row, splat;
//   general: anything else, one div/mod chain per 16-byte run.
#include <string>

#include <type_traits>

#include "common.cuh"
#include "equal_op.cuh"
#include "shape.h"

namespace pdb200 {

constexpr int kGThreads = 256;
constexpr int kGUnroll = 4;

template <typename T>
struct IdentityF {
  __device__ __forceinline__ T operator()(T v) const { return v; }
};
// mean_grad: g / n.  The reference's GPU functor multiplies by MPType(1.0 / n) (kps::DivideFunctor), which for the
// integer dtypes is a multiplication by 0; its CPU kernel divides (funcs::MeanGradFunctor, reduce_functor.h:115-123).
// This kernel divides, in MPType: identical to the CPU kernel, within half an fp32 ulp (before the rounding to
// bf16 / fp16) of the GPU one — and integers get g / n.  The division runs once per out_grad element, not per output.
template <typename T>
struct DivByF {
  using M = typename AccType<T>::type;
  M n;
  __device__ __forceinline__ T operator()(T v) const { return cvt<T>(cvt<M>(v) / n); }
};

// x_grad[m][n] = f(g[n]); grid = column tiles (kGThreads vectors wide) x row slabs
template <typename T, typename F>
__global__ void __launch_bounds__(kGThreads) expand_inner_kernel(const T* __restrict__ g, T* __restrict__ out, int64_t rows,
                                                                 int64_t cols, int64_t rows_per_cta, F f) {
  constexpr int V = Vec<T>::N;
  const int64_t vcol = (int64_t)blockIdx.x * kGThreads + threadIdx.x;
  if (vcol * V >= cols) return;
  const int64_t z = blockIdx.z;  // batch: x_grad[z][m][n] = f(g[z][n]) (a middle dimension reduced)
  Vec<T> v = ld_cached(g + z * cols + vcol * V);
#pragma unroll
  for (int k = 0; k < V; ++k) v.v[k] = f(v.v[k]);
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_cta;
  const int64_t r1 = r0 + rows_per_cta < rows ? r0 + rows_per_cta : rows;
  T* dst = out + (z * rows + r0) * cols + vcol * V;
  for (int64_t r = r0; r < r1; ++r, dst += cols) st_stream(dst, v);
}

// the same for rows of up to K * kGThreads vectors: a thread keeps its K pieces of the row and a CTA writes WHOLE rows,
// so every store instruction of a CTA covers a contiguous run of a row
template <typename T, typename F, int K>
__global__ void __launch_bounds__(kGThreads) expand_inner_rows_kernel(const T* __restrict__ g, T* __restrict__ out, int64_t rows,
                                                                      int64_t cols, int64_t rows_per_cta, F f) {
  constexpr int V = Vec<T>::N;
  const int64_t nv = cols / V;
  Vec<T> v[K];
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const int64_t c = threadIdx.x + (int64_t)k * kGThreads;
    if (c < nv) {
      v[k] = ld_cached(g + c * V);
#pragma unroll
      for (int i = 0; i < V; ++i) v[k].v[i] = f(v[k].v[i]);
    }
  }
  // rows are dealt round-robin (CTA b writes rows b, b + grid, ...): all CTAs move through x_grad together, one window of
  // grid rows at a time — one slab per CTA means ~1200 write streams spread over the whole tensor (the same locality
  // effect as in the split reductions, profiles/r1e_notes.txt)
  (void)rows_per_cta;
  const int64_t step = (int64_t)gridDim.x * cols;
  T* dst = out + (int64_t)blockIdx.x * cols + (int64_t)threadIdx.x * V;
  for (int64_t r = blockIdx.x; r < rows; r += gridDim.x, dst += step) {
#pragma unroll
    for (int k = 0; k < K; ++k)
      if (threadIdx.x + (int64_t)k * kGThreads < nv) st_stream(dst + (int64_t)k * kGThreads * V, v[k]);
  }
}

// x_grad[m][n] = f(g[m]); one CTA per (row, tile of kGThreads*kGUnroll vectors)
template <typename T, typename F, bool VECTOR>
__global__ void __launch_bounds__(kGThreads) expand_outer_kernel(const T* __restrict__ g, T* __restrict__ out, int64_t cols,
                                                                 int64_t tiles_per_row, F f) {
  constexpr int V = VECTOR ? Vec<T>::N : 1;
  const int64_t row = (int64_t)blockIdx.x / tiles_per_row;
  const int64_t tile = (int64_t)blockIdx.x - row * tiles_per_row;
  const T s = f(g[row]);
  const int64_t n = cols / V;
  T* dst = out + row * cols;
  const int64_t base = tile * (kGThreads * kGUnroll) + threadIdx.x;
  if constexpr (VECTOR) {
    Vec<T> sv;
#pragma unroll
    for (int i = 0; i < V; ++i) sv.v[i] = s;
#pragma unroll
    for (int u = 0; u < kGUnroll; ++u) {
      const int64_t v = base + u * kGThreads;
      if (v < n) st_stream(dst + v * V, sv);
    }
  } else {
#pragma unroll
    for (int u = 0; u < kGUnroll; ++u) {
      const int64_t v = base + u * kGThreads;
      if (v < n) dst[v] = s;
    }
  }
}

struct ExpandParams {
  int rank;
  int64_t dims[PDB200_MAX_RANK];  // innermost first (merged)
  int64_t gs[PDB200_MAX_RANK];    // element strides of out_grad (0 on reduced dims)
};
// any other pattern: one thread = one run of V outputs inside the innermost merged dim
template <typename T, typename F, bool VECTOR>
__global__ void __launch_bounds__(kGThreads) expand_general_kernel(const T* __restrict__ g, T* __restrict__ out,
                                                                   int64_t nruns, ExpandParams p, F f) {
  constexpr int V = VECTOR ? Vec<T>::N : 1;
  const int64_t inner_runs = p.dims[0] / V;
  int64_t run = (int64_t)blockIdx.x * kGThreads + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * kGThreads;
  for (; run < nruns; run += stride) {
    int64_t rest = run / inner_runs;
    const int64_t i0 = (run - rest * inner_runs) * V;
    int64_t go = i0 * p.gs[0];
#pragma unroll 1
    for (int d = 1; d < p.rank; ++d) {
      const int64_t q = rest / p.dims[d];
      go += (rest - q * p.dims[d]) * p.gs[d];
      rest = q;
    }
    if constexpr (VECTOR) {
      Vec<T> v;
      if (p.gs[0] == 1) {
        v = ld_cached(g + go);
#pragma unroll
        for (int i = 0; i < V; ++i) v.v[i] = f(v.v[i]);
      } else {
        const T s = f(g[go]);
#pragma unroll
        for (int i = 0; i < V; ++i) v.v[i] = s;
      }
      st_stream(out + run * V, v);
    } else {
      out[run] = f(g[go]);
    }
  }
}

template <typename T, typename F>
__global__ void __launch_bounds__(kGThreads) expand_same_kernel(const T* g, T* out, int64_t numel, F f) {
  int64_t i = (int64_t)blockIdx.x * kGThreads + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * kGThreads;
  for (; i < numel; i += stride) out[i] = f(g[i]);
}

template <typename T, typename F>
int launch_expand(const BroadcastPlan& p, const T* g, T* out, F f, cudaStream_t st) {
  // the plan was made with out_grad's kept-dim shape as "x" and the input's shape as "y": small_is_x is always true here
  constexpr int V = Vec<T>::N;
  switch (p.kind) {
    case kBcastEmpty:
      return PDB200_OK;
    case kBcastSame: {
      const int64_t blocks = std::min<int64_t>((p.numel + kGThreads - 1) / kGThreads, (int64_t)sm_count() * 32);
      expand_same_kernel<T, F><<<(unsigned)std::max<int64_t>(1, blocks), kGThreads, 0, st>>>(g, out, p.numel, f);
      return after_launch("expand_same");
    }
    case kBcastInner: {
      if (!(p.cols % V == 0 && is_aligned16(g) && is_aligned16(out))) break;
      const int64_t col_tiles = (p.cols / V + kGThreads - 1) / kGThreads;
      if (col_tiles <= 8) {  // whole rows per CTA
        // many short-lived CTAs of ~4 rows each, not a persistent grid: bf16 [131072,4096] 0.182 ms with 1184
        // persistent CTAs, 0.151 ms (7.1 TB/s) with 4 rows per CTA, 0.158 with 16, 0.173 with 64 (profiles/r1e_notes.txt)
        int rows_each = 4;
        if (knobs().expand_rows_per_cta >= 1) rows_each = knobs().expand_rows_per_cta;  // tuning knob
        int64_t slabs = std::min<int64_t>(std::max<int64_t>(1, p.rows / rows_each), 0x7fffffffLL);
        const int64_t rpc = (p.rows + slabs - 1) / slabs;
        slabs = (p.rows + rpc - 1) / rpc;
        if (col_tiles == 1) expand_inner_rows_kernel<T, F, 1><<<(unsigned)slabs, kGThreads, 0, st>>>(g, out, p.rows, p.cols, rpc, f);
        else if (col_tiles == 2) expand_inner_rows_kernel<T, F, 2><<<(unsigned)slabs, kGThreads, 0, st>>>(g, out, p.rows, p.cols, rpc, f);
        else if (col_tiles <= 4) expand_inner_rows_kernel<T, F, 4><<<(unsigned)slabs, kGThreads, 0, st>>>(g, out, p.rows, p.cols, rpc, f);
        else expand_inner_rows_kernel<T, F, 8><<<(unsigned)slabs, kGThreads, 0, st>>>(g, out, p.rows, p.cols, rpc, f);
        return after_launch("expand_inner_rows");
      }
      // enough CTAs to fill the machine, but rows_per_cta large so that the one load per thread is amortised
      int64_t slabs = std::max<int64_t>(1, ((int64_t)sm_count() * 8 + col_tiles - 1) / col_tiles);
      slabs = std::min<int64_t>(std::min<int64_t>(slabs, p.rows), 65535);
      const int64_t rows_per_cta = (p.rows + slabs - 1) / slabs;
      slabs = (p.rows + rows_per_cta - 1) / rows_per_cta;
      if (col_tiles > 0x7fffffffLL) break;
      dim3 grid((unsigned)col_tiles, (unsigned)slabs);
      expand_inner_kernel<T, F><<<grid, kGThreads, 0, st>>>(g, out, p.rows, p.cols, rows_per_cta, f);
      return after_launch("expand_inner");
    }
    case kBcastOuter: {
      const bool vec_ok = (p.cols % V == 0) && is_aligned16(out);
      if (p.rows > 1 && p.cols < 32 * V) break;
      const int64_t n = vec_ok ? p.cols / V : p.cols;
      const int64_t tiles_per_row = (n + kGThreads * kGUnroll - 1) / (kGThreads * kGUnroll);
      const int64_t blocks = tiles_per_row * p.rows;
      if (blocks > 0x7fffffffLL) break;
      if (vec_ok) expand_outer_kernel<T, F, true><<<(unsigned)blocks, kGThreads, 0, st>>>(g, out, p.cols, tiles_per_row, f);
      else expand_outer_kernel<T, F, false><<<(unsigned)blocks, kGThreads, 0, st>>>(g, out, p.cols, tiles_per_row, f);
      return after_launch("expand_outer");
    }
    case kBcastGeneral: {
      // one middle dimension reduced: x_grad[b][m][n] = f(g[b][n]) — the tiled inner kernel once per b (grid z)
      if (!(p.rank == 3 && p.xs[0] == 1 && p.xs[1] == 0 && p.xs[2] == p.dims[0])) break;
      const int64_t cols = p.dims[0], rows = p.dims[1], batch = p.dims[2];
      if (!(cols % V == 0 && cols / V >= kGThreads / 2 && batch <= 65535 && is_aligned16(g) && is_aligned16(out))) break;
      const int64_t col_tiles = (cols / V + kGThreads - 1) / kGThreads;
      if (col_tiles > 0x7fffffffLL) break;
      int64_t slabs = std::max<int64_t>(1, ((int64_t)sm_count() * 8 + col_tiles * batch - 1) / (col_tiles * batch));
      slabs = std::min<int64_t>(std::min<int64_t>(slabs, rows), 65535);
      const int64_t rows_per_cta = (rows + slabs - 1) / slabs;
      slabs = (rows + rows_per_cta - 1) / rows_per_cta;
      dim3 grid((unsigned)col_tiles, (unsigned)slabs, (unsigned)batch);
      expand_inner_kernel<T, F><<<grid, kGThreads, 0, st>>>(g, out, rows, cols, rows_per_cta, f);
      return after_launch("expand_inner_batched");
    }
    default:
      break;
  }
  ExpandParams ep;
  ep.rank = p.rank;
  for (int i = 0; i < PDB200_MAX_RANK; ++i) {
    ep.dims[i] = i < p.rank ? p.dims[i] : 1;
    ep.gs[i] = i < p.rank ? p.xs[i] : 0;
  }
  const bool vec_ok = (ep.dims[0] % V == 0) && is_aligned16(g) && is_aligned16(out);
  const int64_t nruns = vec_ok ? p.numel / V : p.numel;
  const int64_t blocks = std::min<int64_t>((nruns + kGThreads - 1) / kGThreads, (int64_t)sm_count() * 64);
  if (vec_ok) expand_general_kernel<T, F, true><<<(unsigned)blocks, kGThreads, 0, st>>>(g, out, nruns, ep, f);
  else expand_general_kernel<T, F, false><<<(unsigned)blocks, kGThreads, 0, st>>>(g, out, nruns, ep, f);
  return after_launch("expand_general");
}

// Shared argument handling: which dims of x were reduced -> the broadcast plan of the kept-dim shape over x's shape.
// recompute_reduce_all (core/kernel_utils.h:407-417) + GetReduceDim (reduce_function.h:100-123)
int plan_reduce_grad(const char* name, const int64_t* x_dims, int x_rank, const int64_t* axes, int n_axes, int reduce_all,
                     BroadcastPlan* p, int64_t* reduce_num) {
  PDB_CHECK_ARG(x_rank >= 0 && x_rank <= PDB200_MAX_RANK, "tensor rank must be in [0, 9]");
  PDB_CHECK_ARG(x_rank == 0 || x_dims != nullptr, "x_dims is NULL");
  PDB_CHECK_ARG(n_axes == 0 || axes != nullptr, "axes is NULL");
  bool red[PDB200_MAX_RANK] = {false};
  if (x_rank == 0 || n_axes == 0) reduce_all = 1;
  for (int i = 0; i < n_axes; ++i) {
    const int64_t a = axes[i];
    if (a < -x_rank || a >= std::max(x_rank, 1))
      return fail(PDB200_INVALID_ARGUMENT, std::string(name) + ": reduce dim index " + std::to_string(a) +
                                               " is out of range for rank " + std::to_string(x_rank));
    if (x_rank > 0) red[a < 0 ? a + x_rank : a] = true;
  }
  int64_t gd[PDB200_MAX_RANK];
  *reduce_num = 1;
  for (int i = 0; i < x_rank; ++i) {
    if (x_dims[i] < 0) return fail(PDB200_INVALID_ARGUMENT, "negative dimension");
    const bool r = reduce_all || red[i];
    gd[i] = r ? 1 : x_dims[i];
    if (r) *reduce_num *= x_dims[i];
  }
  std::string err = plan_broadcast(gd, x_rank, x_dims, x_rank, p);
  if (!err.empty()) return fail(PDB200_INVALID_ARGUMENT, err);
  return PDB200_OK;
}

// op 0: sum_grad, 1: mean_grad
int reduce_grad_entry(int op, const void* out_grad, const int64_t* x_dims, int x_rank, const int64_t* axes, int n_axes,
                      int reduce_all, int dtype, void* x_grad, void* stream) {
  const char* name = op == 0 ? "sum_grad" : "mean_grad";
  BroadcastPlan p;
  int64_t reduce_num = 1;
  if (int rc = plan_reduce_grad(name, x_dims, x_rank, axes, n_axes, reduce_all, &p, &reduce_num)) return rc;
  if (p.kind == kBcastEmpty) return PDB200_OK;
  PDB_CHECK_ARG(out_grad && x_grad, "NULL tensor pointer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (dtype) {
#define M(DT)                                                                                                    \
  case DT: {                                                                                                     \
    using T = typename CType<DT>::type;                                                                          \
    const T* g = static_cast<const T*>(out_grad);                                                                \
    T* o = static_cast<T*>(x_grad);                                                                              \
    if (op == 0) return launch_expand<T, IdentityF<T>>(p, g, o, IdentityF<T>(), st);                             \
    DivByF<T> f{static_cast<typename DivByF<T>::M>(reduce_num)};                                                 \
    return launch_expand<T, DivByF<T>>(p, g, o, f, st);                                                          \
  }
    // gpu/reduce_kernel.cu:371-388 (sum_grad), :346-358 (mean_grad): bool / int8 / uint8 / int16 / complex stay in-tree
    M(PDB200_FLOAT32) M(PDB200_FLOAT64) M(PDB200_FLOAT16) M(PDB200_BFLOAT16) M(PDB200_INT32) M(PDB200_INT64)
#undef M
  }
  return fail(PDB200_UNSUPPORTED, std::string(name) + ": dtype code " + std::to_string(dtype) + " is not registered for this op");
}

// ======================================================================================================
// max_grad / min_grad / prod_grad: x_grad = F(x, out broadcast, out_grad broadcast).
//
// Replaces (fresh design):
//   phi::ReduceMaxGradKernel / ReduceMinGradKernel<T, GPUContext>, paddle/phi/kernels/gpu/reduce_kernel.cu:76-120,159-203:
//     equal_out = Equal(out, x)   (a BroadcastKernel writing an x-sized T tensor; funcs::EqualFunctor<T>, compare_functors.h:35-48)
//     x_grad    = out_grad * equal_out   (a second broadcast kernel)
//   phi::ProdGradKernel<T, GPUContext>, gpu/prod_grad_kernel.cu:21-30 -> impl/reduce_grad.h -> funcs::ProdGradFunctor
//     (funcs/reduce_functor.h:143-159): dx = dy.broadcast(dim) * y.broadcast(dim) * x.inverse(), an Eigen expression in T.
// One kernel reads x once and writes x_grad once (the reference's max/min form moves 4 x-sized streams, this one 2); out
// and out_grad are 1/n-th of the traffic and stay in registers (outer / inner) or in L1/L2 (general).  Arithmetic is done
// step by step in T, like the reference's chain of T-valued kernels / Eigen scalar ops: out_grad * {0,1} keeps
// inf * 0 = NaN; (dy * y) is rounded to T before it meets 1 / x, which is itself rounded to T.
// ======================================================================================================
template <typename T> struct TArith {  // integers: wrap-around products, truncating 1 / x
  static __device__ __forceinline__ T mul(T a, T b) {
    using U = typename std::make_unsigned<T>::type;
    return static_cast<T>(static_cast<U>(a) * static_cast<U>(b));
  }
  // 1 / x for an integer: 1, -1 or 0.  x == 0 is a division by zero in the reference (undefined); here it gives 0.
  static __device__ __forceinline__ T inv(T x) { return x == 1 ? T(1) : (x == T(-1) ? T(-1) : T(0)); }
  static __device__ __forceinline__ T flag(bool b) { return b ? T(1) : T(0); }
};
template <> struct TArith<float> {
  static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
  static __device__ __forceinline__ float inv(float x) { return __fdiv_rn(1.f, x); }
  static __device__ __forceinline__ float flag(bool b) { return b ? 1.f : 0.f; }
};
template <> struct TArith<double> {
  static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
  static __device__ __forceinline__ double inv(double x) { return __ddiv_rn(1.0, x); }
  static __device__ __forceinline__ double flag(bool b) { return b ? 1.0 : 0.0; }
};
// bf16 / fp16: every operation in float, rounded back to T (the product of two 16-bit values is exact in float, and a
// float quotient rounded to an 8- or 11-bit significand equals the correctly rounded quotient: 24 >= 2p + 2).
// 1 / x without the IEEE division sequence (which made prod_grad instruction-bound: 4.9 TB/s, profiles/r2_notes.txt), still
// the same bits: x = m * 2^e with a p-bit integer m (p = 8 or 11); a rounding boundary of the result is k / 2^(p+1) on the
// scale where 2^p / m lies in (1, 2], and |2^p / m - k / 2^(p+1)| = |2^(2p+1) - k m| / (m 2^(p+1)) is either 0 (m a power
// of two: the quotient is exact) or at least 2^-(2p+1) >= 2^-23 ... for p = 11 that is 2^-22 against rcp.approx's error of
// at most 1 ulp = 2^-23 on that scale, for p = 8 2^-17.  Ends of the float exponent range (bf16 only), 0, inf and NaN take
// the exact division.  tests/test_gpu_select_grad.py checks all 2^16 bit patterns of both types.
__device__ __forceinline__ float inv_for_16bit(float fx) {
  float r;
  asm("rcp.approx.f32 %0, %1;" : "=f"(r) : "f"(fx));
  const float ax = fabsf(fx);
  if (__builtin_expect(ax > 7.9e-31f && ax < 1.2e30f, 1)) return r;  // 2^-100 .. 2^100
  return __fdiv_rn(1.f, fx);
}
#define PDB_HALF_ARITH(T)                                                                                          \
  template <> struct TArith<T> {                                                                                   \
    static __device__ __forceinline__ T mul(T a, T b) { return cvt<T>(__fmul_rn(cvt<float>(a), cvt<float>(b))); }  \
    static __device__ __forceinline__ T inv(T x) { return cvt<T>(inv_for_16bit(cvt<float>(x))); }                  \
    static __device__ __forceinline__ T flag(bool b) { return cvt<T>(b ? 1.f : 0.f); }                             \
  };
PDB_HALF_ARITH(__nv_bfloat16)
PDB_HALF_ARITH(__half)
#undef PDB_HALF_ARITH

template <typename T> struct MaxMinGradF {
  static __device__ __forceinline__ T apply(T x, T y, T dy) { return TArith<T>::mul(dy, TArith<T>::flag(EqualOp<T>::apply(y, x))); }
};
template <typename T> struct ProdGradF {
  static __device__ __forceinline__ T apply(T x, T y, T dy) { return TArith<T>::mul(TArith<T>::mul(dy, y), TArith<T>::inv(x)); }
};

// trailing dims reduced (incl. reduce_all): x[m][n], out[m]; one CTA per (row, tile of kGThreads * kGUnroll vectors)
template <typename T, typename F, bool VECTOR>
__global__ void __launch_bounds__(kGThreads) select_outer_kernel(const T* __restrict__ x, const T* __restrict__ y,
                                                                 const T* __restrict__ dy, T* __restrict__ dx, int64_t cols,
                                                                 int64_t tiles_per_row) {
  constexpr int V = VECTOR ? Vec<T>::N : 1;
  const int64_t row = (int64_t)blockIdx.x / tiles_per_row;
  const int64_t tile = (int64_t)blockIdx.x - row * tiles_per_row;
  const T sy = y[row], sdy = dy[row];
  const int64_t n = cols / V;
  const T* src = x + row * cols;
  T* dst = dx + row * cols;
  const int64_t base = tile * (kGThreads * kGUnroll) + threadIdx.x;
  if constexpr (VECTOR) {
    Vec<T> xv[kGUnroll];
#pragma unroll
    for (int u = 0; u < kGUnroll; ++u) {
      const int64_t v = base + u * kGThreads;
      if (v < n) xv[u] = ld_stream(src + v * V);
    }
#pragma unroll
    for (int u = 0; u < kGUnroll; ++u) {
      const int64_t v = base + u * kGThreads;
      if (v < n) {
#pragma unroll
        for (int i = 0; i < V; ++i) xv[u].v[i] = F::apply(xv[u].v[i], sy, sdy);
        st_stream(dst + v * V, xv[u]);
      }
    }
  } else {
#pragma unroll
    for (int u = 0; u < kGUnroll; ++u) {
      const int64_t v = base + u * kGThreads;
      if (v < n) dst[v] = F::apply(src[v], sy, sdy);
    }
  }
}

// leading dims reduced: x[m][n], out[n]; a thread keeps its 16-byte pieces of out / out_grad and walks down a slab of rows
template <typename T, typename F>
__global__ void __launch_bounds__(kGThreads) select_inner_kernel(const T* __restrict__ x, const T* __restrict__ y,
                                                                 const T* __restrict__ dy, T* __restrict__ dx, int64_t rows,
                                                                 int64_t cols, int64_t rows_per_cta) {
  constexpr int V = Vec<T>::N;
  const int64_t vcol = (int64_t)blockIdx.x * kGThreads + threadIdx.x;
  if (vcol * V >= cols) return;
  const int64_t z = blockIdx.z;  // batch: x[z][m][n], out[z][n] (a middle dimension reduced)
  const Vec<T> yv = ld_cached(y + z * cols + vcol * V), dyv = ld_cached(dy + z * cols + vcol * V);
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_cta;
  const int64_t r1 = r0 + rows_per_cta < rows ? r0 + rows_per_cta : rows;
  const T* src = x + (z * rows + r0) * cols + vcol * V;
  T* dst = dx + (z * rows + r0) * cols + vcol * V;
  int64_t r = r0;
  for (; r + kGUnroll <= r1; r += kGUnroll, src += kGUnroll * cols, dst += kGUnroll * cols) {
    Vec<T> xv[kGUnroll];
#pragma unroll
    for (int u = 0; u < kGUnroll; ++u) xv[u] = ld_stream(src + u * cols);
#pragma unroll
    for (int u = 0; u < kGUnroll; ++u) {
#pragma unroll
      for (int i = 0; i < V; ++i) xv[u].v[i] = F::apply(xv[u].v[i], yv.v[i], dyv.v[i]);
      st_stream(dst + u * cols, xv[u]);
    }
  }
  for (; r < r1; ++r, src += cols, dst += cols) {
    Vec<T> xv = ld_stream(src);
#pragma unroll
    for (int i = 0; i < V; ++i) xv.v[i] = F::apply(xv.v[i], yv.v[i], dyv.v[i]);
    st_stream(dst, xv);
  }
}

// anything else (and the elementwise case): one thread = one run of V outputs inside the innermost merged dim
template <typename T, typename F, bool VECTOR>
__global__ void __launch_bounds__(kGThreads) select_general_kernel(const T* __restrict__ x, const T* __restrict__ y,
                                                                   const T* __restrict__ dy, T* __restrict__ dx, int64_t nruns,
                                                                   ExpandParams p) {
  constexpr int V = VECTOR ? Vec<T>::N : 1;
  const int64_t inner_runs = p.dims[0] / V;
  int64_t run = (int64_t)blockIdx.x * kGThreads + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * kGThreads;
  for (; run < nruns; run += stride) {
    int64_t rest = run / inner_runs;
    const int64_t i0 = (run - rest * inner_runs) * V;
    int64_t go = i0 * p.gs[0];
#pragma unroll 1
    for (int d = 1; d < p.rank; ++d) {
      const int64_t q = rest / p.dims[d];
      go += (rest - q * p.dims[d]) * p.gs[d];
      rest = q;
    }
    if constexpr (VECTOR) {
      Vec<T> xv = ld_stream(x + run * V);
      if (p.gs[0] == 1) {
        const Vec<T> yv = ld_cached(y + go), dyv = ld_cached(dy + go);
#pragma unroll
        for (int i = 0; i < V; ++i) xv.v[i] = F::apply(xv.v[i], yv.v[i], dyv.v[i]);
      } else {
        const T sy = y[go], sdy = dy[go];
#pragma unroll
        for (int i = 0; i < V; ++i) xv.v[i] = F::apply(xv.v[i], sy, sdy);
      }
      st_stream(dx + run * V, xv);
    } else {
      dx[run] = F::apply(x[run], y[go], dy[go]);
    }
  }
}

// x[batch][rows][cols] against out[batch][cols]; false when the shape does not suit the kernel (the caller falls back)
template <typename T, typename F>
bool launch_select_inner(const T* x, const T* y, const T* dy, T* dx, int64_t batch, int64_t rows, int64_t cols, cudaStream_t st) {
  constexpr int V = Vec<T>::N;
  if (!(cols % V == 0 && cols / V >= kGThreads / 2 && is_aligned16(x) && is_aligned16(y) && is_aligned16(dy) && is_aligned16(dx))) return false;
  const int64_t col_tiles = (cols / V + kGThreads - 1) / kGThreads;
  if (col_tiles > 0x7fffffffLL || batch > 65535) return false;
  // ~16 CTAs per SM over the whole grid, at least kGUnroll rows each
  int64_t slabs = std::max<int64_t>(1, ((int64_t)sm_count() * 16 + col_tiles * batch - 1) / (col_tiles * batch));
  slabs = std::min<int64_t>(std::min<int64_t>(slabs, std::max<int64_t>(1, rows / kGUnroll)), 65535);
  const int64_t rows_per_cta = (rows + slabs - 1) / slabs;
  slabs = (rows + rows_per_cta - 1) / rows_per_cta;
  dim3 grid((unsigned)col_tiles, (unsigned)slabs, (unsigned)batch);
  select_inner_kernel<T, F><<<grid, kGThreads, 0, st>>>(x, y, dy, dx, rows, cols, rows_per_cta);
  return true;
}

template <typename T, typename F>
int launch_select(const BroadcastPlan& p, const T* x, const T* y, const T* dy, T* dx, cudaStream_t st) {
  constexpr int V = Vec<T>::N;
  switch (p.kind) {
    case kBcastEmpty:
      return PDB200_OK;
    case kBcastInner:
      if (launch_select_inner<T, F>(x, y, dy, dx, 1, p.rows, p.cols, st)) return after_launch("select_inner");
      break;
    case kBcastGeneral:
      // one middle dimension reduced: x[b][m][n], out[b][n] — the inner form once per b (grid z)
      if (p.rank == 3 && p.xs[0] == 1 && p.xs[1] == 0 && p.xs[2] == p.dims[0] &&
          launch_select_inner<T, F>(x, y, dy, dx, p.dims[2], p.dims[1], p.dims[0], st))
        return after_launch("select_inner_batched");
      break;
    case kBcastOuter: {
      const bool vec_ok = (p.cols % V == 0) && is_aligned16(x) && is_aligned16(dx);
      if (p.rows > 1 && p.cols < 32 * V) break;
      const int64_t n = vec_ok ? p.cols / V : p.cols;
      const int64_t tiles_per_row = (n + kGThreads * kGUnroll - 1) / (kGThreads * kGUnroll);
      const int64_t blocks = tiles_per_row * p.rows;
      if (blocks > 0x7fffffffLL) break;
      if (vec_ok) select_outer_kernel<T, F, true><<<(unsigned)blocks, kGThreads, 0, st>>>(x, y, dy, dx, p.cols, tiles_per_row);
      else select_outer_kernel<T, F, false><<<(unsigned)blocks, kGThreads, 0, st>>>(x, y, dy, dx, p.cols, tiles_per_row);
      return after_launch("select_outer");
    }
    default:
      break;
  }
  ExpandParams ep;
  ep.rank = p.rank;
  for (int i = 0; i < PDB200_MAX_RANK; ++i) {
    ep.dims[i] = i < p.rank ? p.dims[i] : 1;
    ep.gs[i] = i < p.rank ? p.xs[i] : 0;
  }
  const bool vec_ok = (ep.dims[0] % V == 0) && is_aligned16(x) && is_aligned16(y) && is_aligned16(dy) && is_aligned16(dx);
  const int64_t nruns = vec_ok ? p.numel / V : p.numel;
  const int64_t blocks = std::max<int64_t>(1, std::min<int64_t>((nruns + kGThreads - 1) / kGThreads, (int64_t)sm_count() * 64));
  if (vec_ok) select_general_kernel<T, F, true><<<(unsigned)blocks, kGThreads, 0, st>>>(x, y, dy, dx, nruns, ep);
  else select_general_kernel<T, F, false><<<(unsigned)blocks, kGThreads, 0, st>>>(x, y, dy, dx, nruns, ep);
  return after_launch("select_general");
}

// op 0: max_grad, 1: min_grad, 2: prod_grad
int select_grad_entry(int op, const void* x, const int64_t* x_dims, int x_rank, const void* out, const void* out_grad,
                      const int64_t* axes, int n_axes, int reduce_all, int dtype, void* x_grad, void* stream) {
  const char* name = op == 0 ? "max_grad" : op == 1 ? "min_grad" : "prod_grad";
  BroadcastPlan p;
  int64_t reduce_num = 1;
  if (int rc = plan_reduce_grad(name, x_dims, x_rank, axes, n_axes, reduce_all, &p, &reduce_num)) return rc;
  if (p.kind == kBcastEmpty) return PDB200_OK;
  PDB_CHECK_ARG(x && out && out_grad && x_grad, "NULL tensor pointer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (dtype) {
#define M(DT)                                                                                                          \
  case DT: {                                                                                                           \
    using T = typename CType<DT>::type;                                                                                \
    const T* xp = static_cast<const T*>(x);                                                                            \
    const T* yp = static_cast<const T*>(out);                                                                          \
    const T* gp = static_cast<const T*>(out_grad);                                                                     \
    T* o = static_cast<T*>(x_grad);                                                                                    \
    if (op == 2) return launch_select<T, ProdGradF<T>>(p, xp, yp, gp, o, st);                                          \
    return launch_select<T, MaxMinGradF<T>>(p, xp, yp, gp, o, st);                                                     \
  }
    // gpu/reduce_kernel.cu:335-344,360-369 (max_grad, min_grad), gpu/prod_grad_kernel.cu:21-30: the same six types
    M(PDB200_FLOAT32) M(PDB200_FLOAT64) M(PDB200_FLOAT16) M(PDB200_BFLOAT16) M(PDB200_INT32) M(PDB200_INT64)
#undef M
  }
  return fail(PDB200_UNSUPPORTED, std::string(name) + ": dtype code " + std::to_string(dtype) + " is not registered for this op");
}

}  // namespace pdb200

extern "C" {
int pdb200_max_grad(const void* x, const int64_t* x_dims, int x_rank, const void* out, const void* out_grad,
                    const int64_t* axes, int n_axes, int reduce_all, int dtype, void* x_grad, void* stream) {
  return pdb200::select_grad_entry(0, x, x_dims, x_rank, out, out_grad, axes, n_axes, reduce_all, dtype, x_grad, stream);
}
int pdb200_min_grad(const void* x, const int64_t* x_dims, int x_rank, const void* out, const void* out_grad,
                    const int64_t* axes, int n_axes, int reduce_all, int dtype, void* x_grad, void* stream) {
  return pdb200::select_grad_entry(1, x, x_dims, x_rank, out, out_grad, axes, n_axes, reduce_all, dtype, x_grad, stream);
}
int pdb200_prod_grad(const void* x, const int64_t* x_dims, int x_rank, const void* out, const void* out_grad,
                     const int64_t* axes, int n_axes, int reduce_all, int dtype, void* x_grad, void* stream) {
  return pdb200::select_grad_entry(2, x, x_dims, x_rank, out, out_grad, axes, n_axes, reduce_all, dtype, x_grad, stream);
}
int pdb200_sum_grad(const void* out_grad, const int64_t* x_dims, int x_rank, const int64_t* axes, int n_axes,
                    int reduce_all, int dtype, void* x_grad, void* stream) {
  return pdb200::reduce_grad_entry(0, out_grad, x_dims, x_rank, axes, n_axes, reduce_all, dtype, x_grad, stream);
}
int pdb200_mean_grad(const void* out_grad, const int64_t* x_dims, int x_rank, const int64_t* axes, int n_axes,
                     int reduce_all, int dtype, void* x_grad, void* stream) {
  return pdb200::reduce_grad_entry(1, out_grad, x_dims, x_rank, axes, n_axes, reduce_all, dtype, x_grad, stream);
}
}
