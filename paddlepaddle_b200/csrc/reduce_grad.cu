// reduce_grad.cu — backward of sum and mean: the gradient of a reduction is out_grad
// broadcast back over the reduced dimensions (mean: times 1/n).
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

#include "common.cuh"
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
  Vec<T> v = ld_cached(g + vcol * V);
#pragma unroll
  for (int k = 0; k < V; ++k) v.v[k] = f(v.v[k]);
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_cta;
  const int64_t r1 = r0 + rows_per_cta < rows ? r0 + rows_per_cta : rows;
  T* dst = out + r0 * cols + vcol * V;
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

// op 0: sum_grad, 1: mean_grad
int reduce_grad_entry(int op, const void* out_grad, const int64_t* x_dims, int x_rank, const int64_t* axes, int n_axes,
                      int reduce_all, int dtype, void* x_grad, void* stream) {
  const char* name = op == 0 ? "sum_grad" : "mean_grad";
  PDB_CHECK_ARG(x_rank >= 0 && x_rank <= PDB200_MAX_RANK, "tensor rank must be in [0, 9]");
  PDB_CHECK_ARG(x_rank == 0 || x_dims != nullptr, "x_dims is NULL");
  PDB_CHECK_ARG(n_axes == 0 || axes != nullptr, "axes is NULL");
  // recompute_reduce_all (core/kernel_utils.h:407-417) + GetReduceDim (reduce_function.h:100-123)
  bool red[PDB200_MAX_RANK] = {false};
  if (x_rank == 0 || n_axes == 0) reduce_all = 1;
  for (int i = 0; i < n_axes; ++i) {
    const int64_t a = axes[i];
    if (a < -x_rank || a >= std::max(x_rank, 1))
      return fail(PDB200_INVALID_ARGUMENT, std::string(name) + ": reduce dim index " + std::to_string(a) +
                                               " is out of range for rank " + std::to_string(x_rank));
    if (x_rank > 0) red[a < 0 ? a + x_rank : a] = true;
  }
  int64_t gd[PDB200_MAX_RANK], reduce_num = 1;
  for (int i = 0; i < x_rank; ++i) {
    if (x_dims[i] < 0) return fail(PDB200_INVALID_ARGUMENT, "negative dimension");
    const bool r = reduce_all || red[i];
    gd[i] = r ? 1 : x_dims[i];
    if (r) reduce_num *= x_dims[i];
  }
  BroadcastPlan p;
  std::string err = plan_broadcast(gd, x_rank, x_dims, x_rank, &p);
  if (!err.empty()) return fail(PDB200_INVALID_ARGUMENT, err);
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

}  // namespace pdb200

extern "C" {
int pdb200_sum_grad(const void* out_grad, const int64_t* x_dims, int x_rank, const int64_t* axes, int n_axes,
                    int reduce_all, int dtype, void* x_grad, void* stream) {
  return pdb200::reduce_grad_entry(0, out_grad, x_dims, x_rank, axes, n_axes, reduce_all, dtype, x_grad, stream);
}
int pdb200_mean_grad(const void* out_grad, const int64_t* x_dims, int x_rank, const int64_t* axes, int n_axes,
                     int reduce_all, int dtype, void* x_grad, void* stream) {
  return pdb200::reduce_grad_entry(1, out_grad, x_dims, x_rank, axes, n_axes, reduce_all, dtype, x_grad, stream);
}
}
