// cross_entropy.cu — cross_entropy_with_softmax forward: softmax along `axis` plus the
// per-position loss, hard (index) or soft (distribution) labels.
//
// Replaces (fresh design): phi::CrossEntropyWithSoftmaxKernel<T, GPUContext>
//   paddle/phi/kernels/gpu/cross_entropy_kernel.cu (registration :1452-1476)
// with the numerics of the CPU kernel the parity oracle is built from
//   paddle/phi/kernels/cpu/cross_entropy_kernel.cc:27-104 -> funcs/cross_entropy.cc:29-134:
//     softmax = SoftmaxKernel(logits)                  (or a copy of the input when use_softmax is false)
//     hard:  loss[i, j] = label == ignore_index ? 0 : -TolerableValue(log(softmax[i, label, j]))
//     soft:  loss[i, j] = -sum_k label[i, k, j] * TolerableValue(log(softmax[i, k, j]))
//   with x viewed as [n][axis_dim][remain] and TolerableValue mapping +-inf to +-1e20 (funcs/cross_entropy.h:28-38).
// The softmax is this library's own kernel (every route of softmax.cu, one read + one write of the logits), which also
// hands out the two statistics of every row, max and log(sum) (softmax_with_stats).  The loss is built from those and
// the LOGITS in the accumulator type:  log(softmax[k]) = max(x[k] - max, -64) - log(sum)  — the same quantity the CPU
// kernel gets by taking the log of its stored probability, without the detour through a probability rounded to the
// tensor dtype (which makes a float16 loss useless: 11 bits, subnormal probabilities).  A hard-label loss costs one
// gathered logit per position, a soft-label loss one pass over logits and labels.  When `softmax` is the logits
// buffer itself (the op is declared in-place) the logits are gone by then and the loss falls back to log(softmax).
// A label outside [0, axis_dim) that is not ignore_index makes the reference raise (a device-side assert on GPU); a
// C-ABI kernel cannot, so the loss at that position is NaN.
#include <string>

#include "common.cuh"
#include "shape.h"

namespace pdb200 {

template <typename A> __device__ __forceinline__ A ce_log(A v);
template <> __device__ __forceinline__ float ce_log<float>(float v) { return logf(v); }
template <> __device__ __forceinline__ double ce_log<double>(double v) { return log(v); }
template <typename A>
__device__ __forceinline__ A tolerable(A x) {
  if (x == std::numeric_limits<A>::infinity()) return A(1e20);
  if (x == -std::numeric_limits<A>::infinity()) return A(-1e20);
  return x;
}

// log(softmax[k]) for one element: from the logits and the row statistics (FROM_LOGITS), or from the stored probability
template <typename T, bool FROM_LOGITS>
__device__ __forceinline__ typename AccType<T>::type log_prob(const T* __restrict__ src, int64_t idx,
                                                              const typename AccType<T>::type* __restrict__ stats,
                                                              int64_t pos) {
  using A = typename AccType<T>::type;
  if constexpr (FROM_LOGITS) {
    A d = cvt<A>(src[idx]) - stats[2 * pos];
    d = d < A(-64) ? A(-64) : d;  // ValueClip of the softmax (funcs/softmax_impl.h:36-42)
    return d - stats[2 * pos + 1];
  } else {
    return tolerable<A>(ce_log<A>(cvt<A>(src[idx])));
  }
}

// one thread per (i, j)
template <typename T, typename L, bool FROM_LOGITS>
__global__ void __launch_bounds__(256) ce_hard_kernel(const T* __restrict__ src, const typename AccType<T>::type* __restrict__ stats,
                                                      const L* __restrict__ label, T* __restrict__ loss, int64_t n_pos,
                                                      int64_t axis_dim, int64_t remain, int64_t ignore_index) {
  using A = typename AccType<T>::type;
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_pos) return;
  const int64_t i = p / remain, j = p - i * remain;
  const int64_t lbl = static_cast<int64_t>(label[p]);
  A v;
  if (lbl == ignore_index) v = A(0);
  else if (lbl < 0 || lbl >= axis_dim) v = std::numeric_limits<A>::quiet_NaN();
  else v = -log_prob<T, FROM_LOGITS>(src, (i * axis_dim + lbl) * remain + j, stats, p);
  loss[p] = cvt<T>(v);
}

// remain == 1: one CTA per row, threads stride over the classes (coalesced), one block reduction
template <typename T, bool FROM_LOGITS>
__global__ void __launch_bounds__(256) ce_soft_rows_kernel(const T* __restrict__ src, const typename AccType<T>::type* __restrict__ stats,
                                                           const T* __restrict__ label, T* __restrict__ loss, int64_t axis_dim) {
  using A = typename AccType<T>::type;
  __shared__ A red[33];
  const int64_t row = blockIdx.x, base = row * axis_dim;
  A acc[2] = {A(0), A(0)};
  int64_t k = threadIdx.x;
  for (; k + blockDim.x < axis_dim; k += 2 * blockDim.x) {
    acc[0] += cvt<A>(label[base + k]) * log_prob<T, FROM_LOGITS>(src, base + k, stats, row);
    acc[1] += cvt<A>(label[base + k + blockDim.x]) * log_prob<T, FROM_LOGITS>(src, base + k + blockDim.x, stats, row);
  }
  if (k < axis_dim) acc[0] += cvt<A>(label[base + k]) * log_prob<T, FROM_LOGITS>(src, base + k, stats, row);
  A s = block_reduce(acc[0] + acc[1], [](A a, A b) { return a + b; }, A(0), red);
  if (threadIdx.x == 0) loss[row] = cvt<T>(-s);
}
// remain > 1: one thread per (i, j), walking the axis with stride `remain` (neighbouring threads stay coalesced)
template <typename T, bool FROM_LOGITS>
__global__ void __launch_bounds__(256) ce_soft_cols_kernel(const T* __restrict__ src, const typename AccType<T>::type* __restrict__ stats,
                                                           const T* __restrict__ label, T* __restrict__ loss, int64_t n_pos,
                                                           int64_t axis_dim, int64_t remain) {
  using A = typename AccType<T>::type;
  const int64_t pos = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pos >= n_pos) return;
  const int64_t i = pos / remain, j = pos - i * remain;
  const int64_t base = i * axis_dim * remain + j;
  A s = A(0);
  for (int64_t k = 0; k < axis_dim; ++k)
    s += cvt<A>(label[base + k * remain]) * log_prob<T, FROM_LOGITS>(src, base + k * remain, stats, pos);
  loss[pos] = cvt<T>(-s);
}

template <typename T>
__global__ void __launch_bounds__(256) ce_copy_kernel(const T* src, T* dst, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) dst[i] = src[i];
}

template <typename T, bool FROM_LOGITS>
int ce_loss(const T* src, const typename AccType<T>::type* stats, const void* label, int label_dtype, bool soft, int64_t n,
            int64_t axis_dim, int64_t remain, int64_t ignore_index, T* loss, cudaStream_t st) {
  const int64_t n_pos = n * remain;
  const int64_t blocks = (n_pos + 255) / 256;
  if (blocks > 0x7fffffffLL || n > 0x7fffffffLL) return fail(PDB200_UNSUPPORTED, "cross_entropy: too many positions for one launch");
  if (soft) {
    const T* l = static_cast<const T*>(label);
    if (remain == 1) ce_soft_rows_kernel<T, FROM_LOGITS><<<(unsigned)n, 256, 0, st>>>(src, stats, l, loss, axis_dim);
    else ce_soft_cols_kernel<T, FROM_LOGITS><<<(unsigned)blocks, 256, 0, st>>>(src, stats, l, loss, n_pos, axis_dim, remain);
    return after_launch("ce_soft");
  }
  if (label_dtype == PDB200_INT64)
    ce_hard_kernel<T, int64_t, FROM_LOGITS><<<(unsigned)blocks, 256, 0, st>>>(src, stats, static_cast<const int64_t*>(label), loss, n_pos, axis_dim, remain, ignore_index);
  else if (label_dtype == PDB200_INT32)
    ce_hard_kernel<T, int32_t, FROM_LOGITS><<<(unsigned)blocks, 256, 0, st>>>(src, stats, static_cast<const int32_t*>(label), loss, n_pos, axis_dim, remain, ignore_index);
  else
    return fail(PDB200_UNSUPPORTED, "cross_entropy: hard labels must be int64 or int32, got dtype code " + std::to_string(label_dtype));
  return after_launch("ce_hard");
}

int softmax_with_stats(const void* x, const int64_t* dims, int rank, int axis, int dtype, void* out, void* stats, void* stream);

inline size_t ce_stats_bytes(int64_t n_pos, int dtype) {
  return (size_t)n_pos * 2 * (dtype == PDB200_FLOAT64 ? sizeof(double) : sizeof(float));
}

// ======================================================================================================
// Backward: logits_grad from the saved softmax, the labels and loss_grad (use_softmax = true).
// Replaces (fresh design): phi::CrossEntropyWithSoftmaxGradKernel<T, GPUContext>,
//   paddle/phi/kernels/gpu/cross_entropy_grad_kernel.cu:281-306 (SoftmaxWithCrossEntropyGradHardLabel / SoftCrossEntropyGradientKernel);
// arithmetic as the CPU kernel the oracle is built from (cpu/cross_entropy_grad_kernel.cc:143-179):
//   soft labels: dx = g * (p - y)
//   hard labels: dx = p * g everywhere, dx[label] = p * g - g (two roundings), a whole position zero when label == ignore_index
// with p the softmax, g = loss_grad of the position (one value per [n][remain] position), every step in T's accumulator
// type (16-bit types: float, one rounding at the end).  One pass: p (and y) read once, dx written once.
// ======================================================================================================
template <typename A> __device__ __forceinline__ A mul_rn(A a, A b);
template <> __device__ __forceinline__ float mul_rn<float>(float a, float b) { return __fmul_rn(a, b); }
template <> __device__ __forceinline__ double mul_rn<double>(double a, double b) { return __dmul_rn(a, b); }
template <typename A> __device__ __forceinline__ A sub_rn(A a, A b);
template <> __device__ __forceinline__ float sub_rn<float>(float a, float b) { return __fsub_rn(a, b); }
template <> __device__ __forceinline__ double sub_rn<double>(double a, double b) { return __dsub_rn(a, b); }

constexpr int kCeThreads = 256;
constexpr int kCeUnroll = 4;

// class axis last: one CTA per (row, tile of kCeThreads * kCeUnroll vectors)
template <typename T, typename LT, bool SOFT, bool VECTOR>
__global__ void __launch_bounds__(kCeThreads) ce_grad_rows_kernel(const T* __restrict__ p, const void* __restrict__ label,
                                                                  const T* __restrict__ lg, T* __restrict__ dx, int64_t cols,
                                                                  int64_t tiles_per_row, int64_t ignore_index) {
  using A = typename AccType<T>::type;
  constexpr int V = VECTOR ? Vec<T>::N : 1;
  const int64_t row = (int64_t)blockIdx.x / tiles_per_row;
  const int64_t tile = (int64_t)blockIdx.x - row * tiles_per_row;
  const A g = cvt<A>(lg[row]);
  const int64_t n = cols / V;
  const T* src = p + row * cols;
  T* dst = dx + row * cols;
  const int64_t base = tile * (kCeThreads * kCeUnroll) + threadIdx.x;
  int64_t lbl = -1;
  bool ignored = false;
  if constexpr (!SOFT) {
    lbl = static_cast<int64_t>(static_cast<const LT*>(label)[row]);
    ignored = lbl == ignore_index;
  }
  const T* ysrc = SOFT ? static_cast<const T*>(label) + row * cols : nullptr;
  if constexpr (VECTOR) {
    Vec<T> pv[kCeUnroll], yv[kCeUnroll];
#pragma unroll
    for (int u = 0; u < kCeUnroll; ++u) {
      const int64_t v = base + u * kCeThreads;
      if (v < n) {
        pv[u] = ld_stream(src + v * V);
        if constexpr (SOFT) yv[u] = ld_stream(ysrc + v * V);
      }
    }
#pragma unroll
    for (int u = 0; u < kCeUnroll; ++u) {
      const int64_t v = base + u * kCeThreads;
      if (v < n) {
#pragma unroll
        for (int i = 0; i < V; ++i) {
          const A pi = cvt<A>(pv[u].v[i]);
          A r;
          if constexpr (SOFT) {
            r = mul_rn<A>(g, sub_rn<A>(pi, cvt<A>(yv[u].v[i])));
          } else {
            r = mul_rn<A>(pi, g);
            if (v * V + i == lbl) r = sub_rn<A>(r, g);
            if (ignored) r = A(0);
          }
          pv[u].v[i] = cvt<T>(r);
        }
        st_stream(dst + v * V, pv[u]);
      }
    }
  } else {
#pragma unroll
    for (int u = 0; u < kCeUnroll; ++u) {
      const int64_t v = base + u * kCeThreads;
      if (v < n) {
        const A pi = cvt<A>(src[v]);
        A r;
        if constexpr (SOFT) {
          r = mul_rn<A>(g, sub_rn<A>(pi, cvt<A>(ysrc[v])));
        } else {
          r = mul_rn<A>(pi, g);
          if (v == lbl) r = sub_rn<A>(r, g);
          if (ignored) r = A(0);
        }
        dst[v] = cvt<T>(r);
      }
    }
  }
}

// class axis not last: [n][dim][remain], one thread per element
template <typename T, typename LT, bool SOFT>
__global__ void __launch_bounds__(256) ce_grad_general_kernel(const T* __restrict__ p, const void* __restrict__ label,
                                                              const T* __restrict__ lg, T* __restrict__ dx, int64_t numel,
                                                              int64_t dim, int64_t remain, int64_t ignore_index) {
  using A = typename AccType<T>::type;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (; i < numel; i += stride) {
    const int64_t b = i / (dim * remain), k = (i / remain) % dim, j = i % remain;
    const int64_t pos = b * remain + j;
    const A g = cvt<A>(lg[pos]);
    const A pi = cvt<A>(p[i]);
    A r;
    if constexpr (SOFT) {
      r = mul_rn<A>(g, sub_rn<A>(pi, cvt<A>(static_cast<const T*>(label)[i])));
    } else {
      const int64_t lbl = static_cast<int64_t>(static_cast<const LT*>(label)[pos]);
      r = mul_rn<A>(pi, g);
      if (k == lbl) r = sub_rn<A>(r, g);
      if (lbl == ignore_index) r = A(0);
    }
    dx[i] = cvt<T>(r);
  }
}

template <typename T, typename LT, bool SOFT>
int ce_grad_launch(const T* p, const void* label, const T* lg, T* dx, int64_t n, int64_t dim, int64_t remain, int64_t ignore_index,
                   cudaStream_t st) {
  constexpr int V = Vec<T>::N;
  if (remain == 1) {
    const bool vec = dim % V == 0 && is_aligned16(p) && is_aligned16(dx) && (!SOFT || is_aligned16(label));
    const int64_t nv = vec ? dim / V : dim;
    const int64_t tiles = (nv + kCeThreads * kCeUnroll - 1) / (kCeThreads * kCeUnroll);
    const int64_t blocks = tiles * n;
    if (blocks <= 0x7fffffffLL) {
      if (vec) ce_grad_rows_kernel<T, LT, SOFT, true><<<(unsigned)blocks, kCeThreads, 0, st>>>(p, label, lg, dx, dim, tiles, ignore_index);
      else ce_grad_rows_kernel<T, LT, SOFT, false><<<(unsigned)blocks, kCeThreads, 0, st>>>(p, label, lg, dx, dim, tiles, ignore_index);
      return after_launch("ce_grad_rows");
    }
  }
  const int64_t numel = n * dim * remain;
  const int64_t blocks = std::max<int64_t>(1, std::min<int64_t>((numel + 255) / 256, (int64_t)sm_count() * 32));
  ce_grad_general_kernel<T, LT, SOFT><<<(unsigned)blocks, 256, 0, st>>>(p, label, lg, dx, numel, dim, remain, ignore_index);
  return after_launch("ce_grad_general");
}

}  // namespace pdb200

extern "C" size_t pdb200_cross_entropy_workspace_bytes(const int64_t* dims, int rank, int axis, int dtype) {
  if (rank < 1 || rank > PDB200_MAX_RANK || dims == nullptr || axis < -rank || axis >= rank) return 0;
  if (axis < 0) axis += rank;
  int64_t n_pos = 1;
  for (int i = 0; i < rank; ++i)
    if (i != axis) n_pos *= dims[i] < 0 ? 0 : dims[i];
  return (pdb200::ce_stats_bytes(n_pos, dtype) + 255) / 256 * 256;
}

extern "C" int pdb200_cross_entropy_with_softmax(const void* logits, const int64_t* dims, int rank, const void* label,
                                                 int label_dtype, int soft_label, int use_softmax, int ignore_index,
                                                 int axis, int dtype, void* softmax, void* loss, void* workspace,
                                                 size_t workspace_bytes, void* stream) {
  using namespace pdb200;
  PDB_CHECK_ARG(rank >= 1 && rank <= PDB200_MAX_RANK, "cross_entropy_with_softmax: rank must be in [1, 9]");
  PDB_CHECK_ARG(dims != nullptr, "cross_entropy_with_softmax: dims is NULL");
  PDB_CHECK_ARG(axis >= -rank && axis < rank, "The axis is expected to be in range of [" + std::to_string(-rank) + ", " +
                                                  std::to_string(rank) + "), but got " + std::to_string(axis));
  if (axis < 0) axis += rank;
  int64_t n = 1, axis_dim = dims[axis], remain = 1;
  for (int i = 0; i < rank; ++i) {
    PDB_CHECK_ARG(dims[i] >= 0, "cross_entropy_with_softmax: negative dimension");
    if (i < axis) n *= dims[i];
    else if (i > axis) remain *= dims[i];
  }
  PDB_CHECK_ARG(axis_dim > 0, "The axis dimension should be larger than 0, but received axis dimension is " + std::to_string(axis_dim) + ".");
  PDB_CHECK_ARG(n > 0, "The size of axis should be larger than 0, but received SizeToAxis of softmax is " + std::to_string(n) + ".");
  if (remain == 0) return PDB200_OK;
  PDB_CHECK_ARG(logits && label && softmax && loss, "cross_entropy_with_softmax: NULL tensor pointer");
  if (soft_label && label_dtype != dtype)
    return fail(PDB200_INVALID_ARGUMENT, "cross_entropy_with_softmax: soft labels must have the dtype of the logits");
  if (!(dtype == PDB200_FLOAT32 || dtype == PDB200_FLOAT64 || dtype == PDB200_FLOAT16 || dtype == PDB200_BFLOAT16))
    return fail(PDB200_UNSUPPORTED, "cross_entropy_with_softmax: dtype must be float32/float64/float16/bfloat16, got code " + std::to_string(dtype));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  // the loss is built from logits + row statistics unless the logits are overwritten (in place) or are probabilities already
  const bool from_logits = use_softmax && softmax != logits;
  if (use_softmax) {
    const size_t need = pdb200_cross_entropy_workspace_bytes(dims, rank, axis, dtype);
    if (from_logits && (workspace == nullptr || workspace_bytes < need))
      return fail(PDB200_WORKSPACE_TOO_SMALL, "cross_entropy_with_softmax: workspace has " + std::to_string(workspace_bytes) +
                                                  " bytes, this call needs " + std::to_string(need) +
                                                  " (ask pdb200_cross_entropy_workspace_bytes)");
    const int rc = softmax_with_stats(logits, dims, rank, axis, dtype, softmax, from_logits ? workspace : nullptr, stream);
    if (rc != PDB200_OK) return rc;
  }
  switch (dtype) {
#define M(DT)                                                                                                       \
  case DT: {                                                                                                        \
    using T = typename CType<DT>::type;                                                                             \
    using A = typename AccType<T>::type;                                                                            \
    if (!use_softmax && softmax != logits) { /* the input already is a softmax: hand it on (cross_entropy_kernel.cc:83-88) */ \
      const int64_t total = n * axis_dim * remain;                                                                  \
      const int64_t blocks = std::min<int64_t>((total + 255) / 256, (int64_t)sm_count() * 32);                      \
      ce_copy_kernel<T><<<(unsigned)blocks, 256, 0, st>>>(static_cast<const T*>(logits), static_cast<T*>(softmax), total); \
      const int rc = after_launch("ce_copy");                                                                       \
      if (rc != PDB200_OK) return rc;                                                                               \
    }                                                                                                               \
    if (from_logits)                                                                                                \
      return ce_loss<T, true>(static_cast<const T*>(logits), static_cast<const A*>(workspace), label, label_dtype,  \
                              soft_label != 0, n, axis_dim, remain, ignore_index, static_cast<T*>(loss), st);       \
    return ce_loss<T, false>(static_cast<const T*>(softmax), nullptr, label, label_dtype, soft_label != 0, n,       \
                             axis_dim, remain, ignore_index, static_cast<T*>(loss), st);                            \
  }
    M(PDB200_FLOAT32) M(PDB200_FLOAT64) M(PDB200_FLOAT16) M(PDB200_BFLOAT16)
#undef M
  }
  return fail(PDB200_UNSUPPORTED, "cross_entropy_with_softmax: dtype not built");
}


extern "C" int pdb200_cross_entropy_with_softmax_grad(const void* label, int label_dtype, const void* softmax, const void* loss_grad,
                                                      const int64_t* dims, int rank, int soft_label, int use_softmax,
                                                      int ignore_index, int axis, int dtype, void* logits_grad, void* stream) {
  using namespace pdb200;
  PDB_CHECK_ARG(rank >= 1 && rank <= PDB200_MAX_RANK && dims != nullptr, "cross_entropy_with_softmax_grad: rank must be in [1, 9]");
  PDB_CHECK_ARG(axis >= -rank && axis < rank, "The axis is expected to be in range of [" + std::to_string(-rank) + ", " +
                                                  std::to_string(rank) + "), but got " + std::to_string(axis));
  if (axis < 0) axis += rank;
  if (!use_softmax)
    return fail(PDB200_UNSUPPORTED, "cross_entropy_with_softmax_grad: use_softmax = false (the input already is a distribution) is not on this path");
  int64_t n = 1, dim = dims[axis], remain = 1;
  for (int i = 0; i < rank; ++i) {
    PDB_CHECK_ARG(dims[i] >= 0, "cross_entropy_with_softmax_grad: negative dimension");
    if (i < axis) n *= dims[i];
    else if (i > axis) remain *= dims[i];
  }
  if (n * dim * remain == 0) return PDB200_OK;
  PDB_CHECK_ARG(label && softmax && loss_grad && logits_grad, "cross_entropy_with_softmax_grad: NULL tensor pointer");
  if (soft_label && label_dtype != dtype)
    return fail(PDB200_INVALID_ARGUMENT, "The Input(Label) should be with the same data type as kernel data type.");
  if (!soft_label && label_dtype != PDB200_INT64 && label_dtype != PDB200_INT32)
    return fail(PDB200_UNSUPPORTED, "cross_entropy_with_softmax_grad: hard labels must be int64 or int32, got dtype code " +
                                        std::to_string(label_dtype));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
#define PDB_CEG_CASE(DT)                                                                                                      \
  case DT: {                                                                                                                  \
    using T = typename CType<DT>::type;                                                                                       \
    const T* p = static_cast<const T*>(softmax);                                                                              \
    const T* g = static_cast<const T*>(loss_grad);                                                                            \
    T* dx = static_cast<T*>(logits_grad);                                                                                     \
    if (soft_label) return ce_grad_launch<T, int64_t, true>(p, label, g, dx, n, dim, remain, ignore_index, st);               \
    if (label_dtype == PDB200_INT64) return ce_grad_launch<T, int64_t, false>(p, label, g, dx, n, dim, remain, ignore_index, st); \
    return ce_grad_launch<T, int32_t, false>(p, label, g, dx, n, dim, remain, ignore_index, st);                              \
  }
  switch (dtype) {
    PDB_CEG_CASE(PDB200_FLOAT32)
    PDB_CEG_CASE(PDB200_FLOAT64)
    PDB_CEG_CASE(PDB200_FLOAT16)
    PDB_CEG_CASE(PDB200_BFLOAT16)
  }
#undef PDB_CEG_CASE
  return fail(PDB200_UNSUPPORTED, "cross_entropy_with_softmax_grad: dtype must be float32/float64/float16/bfloat16, got code " +
                                      std::to_string(dtype));
}
