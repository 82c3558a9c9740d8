// softmax_grad.cu — backward of softmax along one axis, hand-written for sm_100a.
//
// Replaces (fresh design): SoftmaxBackwardCUDAKernelDriver — WarpSoftmaxBackward / cudnnSoftmaxBackward /
// NormalSoftmaxBackward, paddle/phi/kernels/gpudnn/softmax_gpudnn.h:625-741,919-962,1331-1407.
// Numerics follow the CPU kernel the parity oracle is built from
// (paddle/phi/kernels/funcs/softmax_impl.h:275-308,406-445):
//     dot = sum_j y_j * dy_j ;  dx_i = (dy_i - dot) * y_i        (fp16/bf16: in float, one rounding)
// and, with LOG, the backward of log_softmax on the same kernels (phi::LogSoftmaxGradKernel, gpu/log_softmax_grad_kernel.cu;
// CPU functor cpu/log_softmax_grad_kernel.cc:30-62):
//     s = sum_j dy_j ;  dx_i = dy_i - exp(y_i) * s
//
// Routes (tensors viewed as [outer][dim][inner]):
//   inner == 1, many rows -> pipe : persistent CTAs on the TMA row pipeline of rowpipe.cuh; a stage holds the
//                            y row AND the dy row (two cp.async.bulk on one mbarrier); both are read from HBM once
//   inner == 1, otherwise -> rows : one CTA per row, two sweeps (the second hits L2)
//   inner  > 1            -> cols : threads along `inner` (coalesced), two sweeps over `dim`
// Algorithmic bytes: 3 * sizeof(T) per element.
#include "common.cuh"
#include "rowpipe.cuh"
#include "shape.h"

namespace pdb200 {

template <typename A> struct AddG { __device__ __forceinline__ A operator()(A a, A b) const { return a + b; } };

// ---- pipe ---------------------------------------------------------------------------------------------
template <typename T, int C, bool EXACT, bool LOG>
__global__ void __launch_bounds__(512) softmax_grad_pipe_kernel(const T* __restrict__ y, const T* __restrict__ dy,
                                                                T* __restrict__ dx, int64_t rows, int dim,
                                                                int stages) {
  constexpr int V = Vec<T>::N;
  extern __shared__ __align__(128) unsigned char ring_raw[];
  T* ring = reinterpret_cast<T*>(ring_raw);  // stage s: [y row | dy row]
  __shared__ uint64_t full[kPipeMaxStages];
  __shared__ float red[2][32];
  const int tid = threadIdx.x, nt = blockDim.x;
  const uint32_t row_bytes = (uint32_t)dim * sizeof(T);
  const int64_t row_step = gridDim.x;
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) mbar_init(&full[s], 1);
    mbar_fence_init();
  }
  block_reduce_1sync_init(red[0], 0.f);
  block_reduce_1sync_init(red[1], 0.f);
  __syncthreads();
  auto fill = [&](int s, int64_t r) {
    mbar_expect_tx(&full[s], 2 * row_bytes);
    bulk_load_row(ring + (size_t)s * 2 * dim, y + r * dim, row_bytes, &full[s]);
    bulk_load_row(ring + (size_t)s * 2 * dim + dim, dy + r * dim, row_bytes, &full[s]);
  };
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) {
      const int64_t r = blockIdx.x + s * row_step;
      if (r < rows) fill(s, r);
    }
  }
  int s = 0, flip = 0;
  uint32_t phase = 0;
  for (int64_t row = blockIdx.x; row < rows; row += row_step) {
    mbar_wait(&full[s], phase);
    const T* sy = ring + (size_t)s * 2 * dim;
    const T* sg = sy + dim;
    float yv[C * V], gv[C * V];
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + tid) * V;
      if (EXACT || off < dim) {
        unpack_to_float<T>(lds_vec(sy + off), &yv[c * V]);
        unpack_to_float<T>(lds_vec(sg + off), &gv[c * V]);
      } else {
#pragma unroll
        for (int k = 0; k < V; ++k) yv[c * V + k] = gv[c * V + k] = 0.f;
      }
#pragma unroll
      for (int k = 0; k < V; ++k) acc[k & 3] = LOG ? acc[k & 3] + gv[c * V + k] : fmaf(yv[c * V + k], gv[c * V + k], acc[k & 3]);
    }
    float dot = (acc[0] + acc[1]) + (acc[2] + acc[3]);
    dot = block_reduce_1sync(dot, AddG<float>(), red[flip]);  // its barrier also means: stage s fully read
    flip ^= 1;
    if (tid == 0) {
      const int64_t nrow = row + (int64_t)stages * row_step;
      if (nrow < rows) {
        fence_proxy_async_smem();
        fill(s, nrow);
      }
    }
    T* dst = dx + row * dim;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + tid) * V;
      if (EXACT || off < dim) {
        float r[V];
#pragma unroll
        for (int k = 0; k < V; ++k) {
          if constexpr (LOG) {
            float ey;
            asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ey) : "f"(yv[c * V + k] * 1.4426950408889634f));
            r[k] = fmaf(-ey, dot, gv[c * V + k]);
          } else {
            r[k] = __fmul_rn(gv[c * V + k] - dot, yv[c * V + k]);
          }
        }
        Vec<T> o;
        pack_from_float<T>(r, o);
        st_stream(dst + off, o);
      }
    }
    if (++s == stages) {
      s = 0;
      phase ^= 1;
    }
  }
}

// ---- rows: one CTA per row, any length / alignment --------------------------------------------------------
template <typename A> __device__ __forceinline__ A exp_acc(A v);
template <> __device__ __forceinline__ float exp_acc<float>(float v) { return expf(v); }
template <> __device__ __forceinline__ double exp_acc<double>(double v) { return exp(v); }

template <typename T, bool LOG>
__global__ void __launch_bounds__(1024) softmax_grad_rows_kernel(const T* y, const T* dy, T* dx, int64_t dim) {
  using A = typename AccType<T>::type;
  __shared__ A red[33];
  const T* ry = y + (int64_t)blockIdx.x * dim;
  const T* rg = dy + (int64_t)blockIdx.x * dim;
  T* rd = dx + (int64_t)blockIdx.x * dim;
  A dot = A(0);
  for (int64_t i = threadIdx.x; i < dim; i += blockDim.x) dot += LOG ? cvt<A>(rg[i]) : cvt<A>(ry[i]) * cvt<A>(rg[i]);
  dot = block_reduce(dot, AddG<A>(), A(0), red);
  // dx may alias dy (in-place grad): each element is read and written by the same thread
  for (int64_t i = threadIdx.x; i < dim; i += blockDim.x)
    rd[i] = cvt<T>(LOG ? cvt<A>(rg[i]) - exp_acc<A>(cvt<A>(ry[i])) * dot : (cvt<A>(rg[i]) - dot) * cvt<A>(ry[i]));
}

// ---- cols: softmax axis is not the last one ---------------------------------------------------------------
template <typename T, bool LOG>
__global__ void __launch_bounds__(256) softmax_grad_cols_kernel(const T* y, const T* dy, T* dx, int64_t dim,
                                                                int64_t inner, int64_t col_tiles) {
  using A = typename AccType<T>::type;
  const int64_t o = (int64_t)blockIdx.x / col_tiles;
  const int64_t ct = (int64_t)blockIdx.x - o * col_tiles;
  const int64_t col = ct * blockDim.x + threadIdx.x;
  if (col >= inner) return;
  const int64_t base = o * dim * inner + col;
  A dot = A(0);
  for (int64_t r = 0; r < dim; ++r) dot += LOG ? cvt<A>(dy[base + r * inner]) : cvt<A>(y[base + r * inner]) * cvt<A>(dy[base + r * inner]);
  for (int64_t r = 0; r < dim; ++r)
    dx[base + r * inner] = cvt<T>(LOG ? cvt<A>(dy[base + r * inner]) - exp_acc<A>(cvt<A>(y[base + r * inner])) * dot
                                      : (cvt<A>(dy[base + r * inner]) - dot) * cvt<A>(y[base + r * inner]));
}

template <typename T>
__global__ void softmax_grad_zero_kernel(T* dx) { dx[0] = cvt<T>(typename AccType<T>::type(0)); }

// ---- host ---------------------------------------------------------------------------------------------------
template <typename T, bool LOG>
int try_softmax_grad_pipe(const T* y, const T* dy, T* dx, int64_t outer, int64_t dim, cudaStream_t st) {
  if constexpr (std::is_same<T, double>::value) {
    return -1;
  } else {
  constexpr int VN = Vec<T>::N;
  const int64_t nchunk = dim / VN;
  int C = sizeof(T) == 2 ? (nchunk >= 256 ? 4 : (nchunk >= 128 ? 2 : 1)) : (nchunk >= 256 ? 2 : 1);
  while (C < 4 && (nchunk + C - 1) / C > 512) C <<= 1;
  int threads = (int)((nchunk + C - 1) / C);
  threads = (threads + 31) / 32 * 32;
  if (threads > 512) return -1;
  const size_t stage_bytes = 2 * (size_t)dim * sizeof(T);
  const bool exact = (int64_t)threads * C == nchunk;
  auto launch = [&](auto kern) -> int {
    int per_sm = 0, stages = 0;
    size_t smem = 0;
    int want = 768 / threads;
    want = want < 2 ? 2 : (want > 6 ? 6 : want);
    // ~64 KB of rows in flight per SM at two stages (profiles/r1e_notes.txt): bf16 [.,2048] 6 -> 4 CTAs 1.01 -> 0.93 ms
    const int by_bytes = (int)((64 * 1024) / (2 * stage_bytes));
    if (want > by_bytes) want = by_bytes < 1 ? 1 : by_bytes;
    FuncAttrHold hold;  // released when this lambda returns, i.e. after the launch
    if (!pipe_geometry(kern, threads, stage_bytes, want, 3, &per_sm, &stages, &smem, &hold)) return -1;
    int64_t grid = (int64_t)per_sm * sm_count();
    // 64-thread CTAs: short-lived CTAs of 4 rows instead of the persistent grid (bf16 [.,2048]: 0.935 -> 0.906 ms; the
    // 256-thread fp32 CTAs lose, 1.906 -> 2.05 ms, and stay persistent) — as in softmax.cu
    if (threads <= 64) grid = std::max<int64_t>(grid, outer / 4);
    if (const int rpc = knobs().pipe_rows_per_cta; rpc >= 1)  // tuning knob
      grid = std::max<int64_t>((int64_t)per_sm * sm_count(), outer / rpc);
    if (grid > outer) grid = outer;
    if (grid > 0x7fffffffLL) grid = 0x7fffffffLL;
    kern<<<(unsigned)grid, threads, smem, st>>>(y, dy, dx, outer, (int)dim, stages);
    return after_launch("softmax_grad_pipe");
  };
#define PDB_SMG(CC) return exact ? launch(softmax_grad_pipe_kernel<T, CC, true, LOG>) : launch(softmax_grad_pipe_kernel<T, CC, false, LOG>);
  if (C == 4) { PDB_SMG(4) }
  if (C == 2) { PDB_SMG(2) }
  PDB_SMG(1)
#undef PDB_SMG
  }
}

template <typename T, bool LOG>
int softmax_grad_typed(const T* y, const T* dy, T* dx, int64_t outer, int64_t dim, int64_t inner, cudaStream_t st) {
  constexpr int VN = Vec<T>::N;
  if (inner == 1) {
    const bool vec = (dim % VN == 0) && is_aligned16(y) && is_aligned16(dy) && is_aligned16(dx);
    const int64_t nchunk = dim / VN;
    if (vec && nchunk >= 64 && nchunk <= 512 * 4 && outer >= 4 * (int64_t)sm_count() && dx != y) {
      const int rc = try_softmax_grad_pipe<T, LOG>(y, dy, dx, outer, dim, st);
      if (rc >= 0) return rc;
    }
    if (outer > 0x7fffffffLL) return fail(PDB200_UNSUPPORTED, "softmax_grad: too many rows for one launch");
    int threads = 128;
    while (threads < 1024 && threads * 4 < dim) threads <<= 1;
    softmax_grad_rows_kernel<T, LOG><<<(unsigned)outer, threads, 0, st>>>(y, dy, dx, dim);
    return after_launch("softmax_grad_rows");
  }
  int threads = 32;
  while (threads < 256 && threads < inner) threads <<= 1;
  const int64_t col_tiles = (inner + threads - 1) / threads;
  const int64_t blocks = col_tiles * outer;
  if (blocks > 0x7fffffffLL) return fail(PDB200_UNSUPPORTED, "softmax_grad: too many tiles for one launch");
  softmax_grad_cols_kernel<T, LOG><<<(unsigned)blocks, threads, 0, st>>>(y, dy, dx, dim, inner, col_tiles);
  return after_launch("softmax_grad_cols");
}

}  // namespace pdb200

namespace pdb200 {
template <bool LOG>
int softmax_grad_entry(const void* out, const void* out_grad, const int64_t* dims, int rank, int axis, int dtype, void* x_grad,
                       void* stream) {
  PDB_CHECK_ARG(rank >= 0 && rank <= PDB200_MAX_RANK, "softmax_grad: rank must be in [0, 9]");
  PDB_CHECK_ARG(rank == 0 || dims != nullptr, "softmax_grad: dims is NULL");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (rank > 0) {
    PDB_CHECK_ARG(axis >= -rank && axis < rank, "The axis is expected to be in range of [" + std::to_string(-rank) +
                                                    ", " + std::to_string(rank) + "), but got " + std::to_string(axis));
    if (axis < 0) axis += rank;
  }
  int64_t outer = 1, dim = 1, inner = 1;
  for (int i = 0; i < rank; ++i) {
    PDB_CHECK_ARG(dims[i] >= 0, "softmax_grad: negative dimension");
    if (i < axis) outer *= dims[i];
    else if (i == axis) dim = dims[i];
    else inner *= dims[i];
  }
  if (rank > 0 && outer * dim * inner == 0) return PDB200_OK;
  PDB_CHECK_ARG(out && out_grad && x_grad, "softmax_grad: NULL tensor pointer");
#define PDB_SMG_CASE(DT)                                                                                   \
  case DT: {                                                                                               \
    using T = typename CType<DT>::type;                                                                    \
    if (rank == 0) { /* 0-D: the gradient of a constant 1 is 0 (impl/softmax_grad_kernel_impl.h:33-37) */  \
      softmax_grad_zero_kernel<T><<<1, 1, 0, st>>>(static_cast<T*>(x_grad));                               \
      return after_launch("softmax_grad_zero");                                                            \
    }                                                                                                      \
    return softmax_grad_typed<T, LOG>(static_cast<const T*>(out), static_cast<const T*>(out_grad),              \
                                 static_cast<T*>(x_grad), outer, dim, inner, st);                          \
  }
  switch (dtype) {
    PDB_SMG_CASE(PDB200_FLOAT32)
    PDB_SMG_CASE(PDB200_FLOAT64)
    PDB_SMG_CASE(PDB200_FLOAT16)
    PDB_SMG_CASE(PDB200_BFLOAT16)
    default:
      return fail(PDB200_UNSUPPORTED, "softmax_grad: dtype must be float32/float64/float16/bfloat16, got code " +
                                          std::to_string(dtype));
  }
#undef PDB_SMG_CASE
}
}  // namespace pdb200

extern "C" int pdb200_softmax_grad(const void* out, const void* out_grad, const int64_t* dims, int rank, int axis,
                                   int dtype, void* x_grad, void* stream) {
  return pdb200::softmax_grad_entry<false>(out, out_grad, dims, rank, axis, dtype, x_grad, stream);
}
extern "C" int pdb200_log_softmax_grad(const void* out, const void* out_grad, const int64_t* dims, int rank, int axis,
                                       int dtype, void* x_grad, void* stream) {
  return pdb200::softmax_grad_entry<true>(out, out_grad, dims, rank, axis, dtype, x_grad, stream);
}
