// layer_norm_grad.cu — backward of layer_norm, hand-written for sm_100a.
//
// Replaces (fresh design): phi::LayerNormGradKernel<T, GPUContext> and the LayerNormBackward* kernels of
// paddle/phi/kernels/funcs/layer_norm_impl.cu.h:545-2204 (paddle/phi/kernels/gpu/layer_norm_grad_kernel.cu).
// Numerics follow the CPU kernel the parity oracle is built from (paddle/phi/kernels/cpu/layer_norm_grad_kernel.cc:28-155):
//     xhat = (x - mean) / sqrt(var + eps);  g = dy * scale (or dy)
//     dx = (g - rowmean(g) - xhat * rowmean(g * xhat)) / sqrt(var + eps)
//     dscale[j] = sum_rows dy * xhat ;  dbias[j] = sum_rows dy
// all in float for fp16/bf16/fp32 inputs (double for fp64).
//
// Routes ([rows][D] view):
//   pipe     : many rows, D = 64..2048 16-byte chunks — persistent CTAs on the TMA row pipeline; a stage holds
//              the x row AND the dy row, both read from HBM ONCE.  Each thread owns fixed columns, so its share
//              of dscale / dbias accumulates in registers across all the rows of the CTA and is written once, as a
//              partial row of the scratch; a small second kernel folds the <= 592 partials per column.
//              Algorithmic bytes: 3 * sizeof(T) per element (+ the partials, ~D * 8 B per CTA).
//   fallback : any shape — dx by one CTA per row (two sweeps, the second from L2), dscale/dbias by a column kernel
//              over row slabs (x and dy are read a second time), same fold.
#include "common.cuh"
#include "rowpipe.cuh"
#include "shape.h"

namespace pdb200 {

constexpr int kLnGradMaxParts = 592;  // 4 CTAs x 148 SMs: upper bound of partial rows in the scratch

struct AddU2 { __device__ __forceinline__ float operator()(float a, float b) const { return a + b; } };
template <typename U> struct AddU1 { __device__ __forceinline__ U operator()(U a, U b) const { return a + b; } };

template <typename U> __device__ __forceinline__ U lng_rstd(U v);
template <> __device__ __forceinline__ float lng_rstd<float>(float v) { return 1.0f / sqrtf(v); }
template <> __device__ __forceinline__ double lng_rstd<double>(double v) { return 1.0 / sqrt(v); }

// ---- pipe ---------------------------------------------------------------------------------------------
template <typename T, typename P, int C, bool HAS_SCALE, bool PARAMS, bool EXACT>
__global__ void __launch_bounds__(256, 2)
ln_grad_pipe_kernel(const T* __restrict__ x, const T* __restrict__ dy, const P* __restrict__ scale,
                    const float* __restrict__ mean, const float* __restrict__ var, T* __restrict__ dx,
                    float* __restrict__ part, int64_t rows, int D, int stages, float eps, float inv_d) {
  constexpr int V = Vec<T>::N;
  constexpr int PV = Vec<P>::N;
  extern __shared__ __align__(128) unsigned char ring_raw[];
  T* ring = reinterpret_cast<T*>(ring_raw);  // stage s: [x row | dy row]
  __shared__ uint64_t full[kPipeMaxStages];
  __shared__ float red[2][2][32];
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t row_bytes = (uint32_t)D * sizeof(T);
  const int64_t row_step = gridDim.x;
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) mbar_init(&full[s], 1);
    mbar_fence_init();
  }
  if (tid < 32) red[0][0][tid] = red[0][1][tid] = red[1][0][tid] = red[1][1][tid] = 0.f;
  __syncthreads();
  auto fill = [&](int s, int64_t r) {
    mbar_expect_tx(&full[s], 2 * row_bytes);
    bulk_load_row(ring + (size_t)s * 2 * D, x + r * D, row_bytes, &full[s]);
    bulk_load_row(ring + (size_t)s * 2 * D + D, dy + r * D, row_bytes, &full[s]);
  };
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) {
      const int64_t r = blockIdx.x + s * row_step;
      if (r < rows) fill(s, r);
    }
  }
  // per-element arithmetic on fp32 pairs (FFMA2 / FADD2 / FMUL2, common.cuh): same results, half the issue slots
  constexpr int H = V / 2;
  F2 ds[PARAMS ? C * H : 1], db[PARAMS ? C * H : 1];
  if constexpr (PARAMS) {
#pragma unroll
    for (int i = 0; i < C * H; ++i) ds[i] = db[i] = f2_splat(0.f);
  }
  int s = 0, flip = 0;
  uint32_t phase = 0;
  for (int64_t row = blockIdx.x; row < rows; row += row_step) {
    const float mu = mean[row];
    const float rstd = lng_rstd<float>(var[row] + eps);
    const F2 rstd2 = f2_splat(rstd), nmr2 = f2_splat(-mu * rstd);
    mbar_wait(&full[s], phase);
    const T* sx = ring + (size_t)s * 2 * D;
    const T* sg = sx + D;
    // The row is read from shared memory twice (sums, then dx) instead of being parked in 32 registers:
    // with the dscale/dbias accumulators that keeps the kernel under 128 registers without spills.
    F2 s1p = f2_splat(0.f), s2p = f2_splat(0.f);
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + tid) * V;
      if (EXACT || off < D) {
        float xf[V], gf[V], w[V];
        unpack_to_float<T>(lds_vec(sx + off), xf);
        unpack_to_float<T>(lds_vec(sg + off), gf);
        if constexpr (HAS_SCALE) {
#pragma unroll
          for (int j = 0; j < V / PV; ++j) unpack_to_float<P>(ld_cached(scale + off + j * PV), &w[j * PV]);
        }
#pragma unroll
        for (int h = 0; h < H; ++h) {
          const F2 g2 = f2_make(gf[2 * h], gf[2 * h + 1]);
          const F2 xh = f2_fma(f2_make(xf[2 * h], xf[2 * h + 1]), rstd2, nmr2);
          const F2 gw = HAS_SCALE ? f2_mul(g2, f2_make(w[2 * h], w[2 * h + 1])) : g2;
          s1p = f2_add(s1p, gw);
          s2p = f2_fma(gw, xh, s2p);
          if constexpr (PARAMS) {
            db[c * H + h] = f2_add(db[c * H + h], g2);
            ds[c * H + h] = f2_fma(g2, xh, ds[c * H + h]);
          }
        }
      }
    }
    float s1 = f2_lo(s1p) + f2_hi(s1p), s2 = f2_lo(s2p) + f2_hi(s2p);
    // both row sums through ONE barrier
    s1 = group_reduce<32>(s1, AddU2());
    s2 = group_reduce<32>(s2, AddU2());
    if (lane == 0) {
      red[flip][0][warp] = s1;
      red[flip][1][warp] = s2;
    }
    __syncthreads();
    s1 = group_reduce<32>(red[flip][0][lane], AddU2());
    s2 = group_reduce<32>(red[flip][1][lane], AddU2());
    flip ^= 1;
    const float m1 = s1 * inv_d, m2 = s2 * inv_d;
    // dx = rstd * (gw - m1 - xhat * m2) with xhat = (x - mu) * rstd, regrouped as a*gw + (b*x + c): two FMAs per
    // element instead of five operations (b*x + c cancels to -rstd*(m1 + xhat*m2); its rounding error is
    // ~ulp(|mu| * rstd^2 * |m2|), far below the 1e-5 bar since m2 is a mean of zero-mean products)
    const float cb = -rstd * rstd * m2;
    const F2 ca2 = rstd2, cb2 = f2_splat(cb), cc2 = f2_splat(-rstd * m1 - mu * cb);
    T* dst = dx + row * D;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + tid) * V;
      if (EXACT || off < D) {
        float xf[V], gf[V], w[V], o[V];
        unpack_to_float<T>(lds_vec(sx + off), xf);
        unpack_to_float<T>(lds_vec(sg + off), gf);
        if constexpr (HAS_SCALE) {
#pragma unroll
          for (int j = 0; j < V / PV; ++j) unpack_to_float<P>(ld_cached(scale + off + j * PV), &w[j * PV]);
        }
#pragma unroll
        for (int h = 0; h < H; ++h) {
          const F2 g2 = f2_make(gf[2 * h], gf[2 * h + 1]);
          const F2 gw = HAS_SCALE ? f2_mul(g2, f2_make(w[2 * h], w[2 * h + 1])) : g2;
          const F2 r = f2_fma(ca2, gw, f2_fma(cb2, f2_make(xf[2 * h], xf[2 * h + 1]), cc2));
          o[2 * h] = f2_lo(r);
          o[2 * h + 1] = f2_hi(r);
        }
        Vec<T> ov;
        pack_from_float<T>(o, ov);
        st_stream(dst + off, ov);
      }
    }
    __syncthreads();  // every thread has finished with stage s: hand it back to the copy engine
    if (tid == 0) {
      const int64_t nrow = row + (int64_t)stages * row_step;
      if (nrow < rows) {
        fence_proxy_async_smem();
        fill(s, nrow);
      }
    }
    if (++s == stages) {
      s = 0;
      phase ^= 1;
    }
  }
  if constexpr (PARAMS) {
    float* pds = part + (size_t)blockIdx.x * 2 * D;
    float* pdb = pds + D;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + tid) * V;
      if (EXACT || off < D) {
#pragma unroll
        for (int h = 0; h < H; ++h) {
          pds[off + 2 * h] = f2_lo(ds[c * H + h]);
          pds[off + 2 * h + 1] = f2_hi(ds[c * H + h]);
          pdb[off + 2 * h] = f2_lo(db[c * H + h]);
          pdb[off + 2 * h + 1] = f2_hi(db[c * H + h]);
        }
      }
    }
  }
}

// ---- fold of the partial rows: part[nparts][2][D] -> dscale[D], dbias[D] ------------------------------------
template <typename P, typename U>
__global__ void __launch_bounds__(256) ln_grad_fold_kernel(const U* __restrict__ part, int nparts, int64_t D,
                                                           P* dscale, P* dbias) {
  const int64_t col = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int which = blockIdx.y;  // 0: dscale, 1: dbias
  P* dst = which == 0 ? dscale : dbias;
  if (col >= D || dst == nullptr) return;
  U a0 = U(0), a1 = U(0);
  int p = 0;
  for (; p + 1 < nparts; p += 2) {
    a0 += part[((size_t)p * 2 + which) * D + col];
    a1 += part[((size_t)(p + 1) * 2 + which) * D + col];
  }
  if (p < nparts) a0 += part[((size_t)p * 2 + which) * D + col];
  dst[col] = cvt<P>(a0 + a1);
}

// ---- fallback: dx, one CTA per row ----------------------------------------------------------------------------
template <typename T, typename P>
__global__ void __launch_bounds__(1024) ln_grad_dx_rows_kernel(const T* x, const T* dy, const P* scale,
                                                               const typename AccType<T>::type* mean,
                                                               const typename AccType<T>::type* var, T* dx, int64_t D,
                                                               typename AccType<T>::type eps, double inv_d) {
  using U = typename AccType<T>::type;
  __shared__ U red[33];
  const int64_t row = blockIdx.x;
  const T* rx = x + row * D;
  const T* rg = dy + row * D;
  T* rd = dx + row * D;
  const U mu = mean[row];
  const U rstd = lng_rstd<U>(var[row] + eps);
  U s1 = U(0), s2 = U(0);
  for (int64_t i = threadIdx.x; i < D; i += blockDim.x) {
    const U xh = (cvt<U>(rx[i]) - mu) * rstd;
    const U gw = scale ? cvt<U>(rg[i]) * cvt<U>(scale[i]) : cvt<U>(rg[i]);
    s1 += gw;
    s2 += gw * xh;
  }
  s1 = block_reduce(s1, AddU1<U>(), U(0), red);
  s2 = block_reduce(s2, AddU1<U>(), U(0), red);
  const U m1 = (U)((double)s1 * inv_d), m2 = (U)((double)s2 * inv_d);
  for (int64_t i = threadIdx.x; i < D; i += blockDim.x) {
    const U xh = (cvt<U>(rx[i]) - mu) * rstd;
    const U gw = scale ? cvt<U>(rg[i]) * cvt<U>(scale[i]) : cvt<U>(rg[i]);
    rd[i] = cvt<T>(((gw - m1) - xh * m2) * rstd);
  }
}

// ---- fallback: dscale / dbias partials over row slabs ---------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(128) ln_grad_params_kernel(const T* x, const T* dy,
                                                             const typename AccType<T>::type* mean,
                                                             const typename AccType<T>::type* var,
                                                             typename AccType<T>::type* part, int64_t rows, int64_t D,
                                                             int64_t slab, typename AccType<T>::type eps) {
  using U = typename AccType<T>::type;
  const int64_t col = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t r0 = (int64_t)blockIdx.y * slab;
  const int64_t r1 = r0 + slab < rows ? r0 + slab : rows;
  if (col >= D) return;
  U ds = U(0), db = U(0);
  for (int64_t r = r0; r < r1; ++r) {
    const U g = cvt<U>(dy[r * D + col]);
    const U xh = (cvt<U>(x[r * D + col]) - mean[r]) * lng_rstd<U>(var[r] + eps);
    db += g;
    ds += g * xh;
  }
  part[((size_t)blockIdx.y * 2 + 0) * D + col] = ds;
  part[((size_t)blockIdx.y * 2 + 1) * D + col] = db;
}

// ---- host ---------------------------------------------------------------------------------------------------
inline size_t ln_grad_scratch_bytes(int64_t D, size_t acc_size) {
  return (size_t)kLnGradMaxParts * 2 * (size_t)D * acc_size;
}

template <typename T, typename P>
int layer_norm_grad_typed(const T* x, const P* scale, const void* mean_v, const void* var_v, const T* dy, T* dx,
                          P* dscale, P* dbias, int64_t rows, int64_t D, float epsilon, void* ws, size_t ws_bytes,
                          cudaStream_t st) {
  using U = typename AccType<T>::type;
  constexpr int VN = Vec<T>::N;
  const U* mean = static_cast<const U*>(mean_v);
  const U* var = static_cast<const U*>(var_v);
  const U eps = static_cast<U>(epsilon);
  const bool params = dscale != nullptr || dbias != nullptr;
  if (params) {
    const size_t need = ln_grad_scratch_bytes(D, sizeof(U));
    if (ws == nullptr || ws_bytes < need)
      return fail(PDB200_WORKSPACE_TOO_SMALL, "layer_norm_grad: workspace has " + std::to_string(ws_bytes) +
                                                  " bytes, this call needs " + std::to_string(need) +
                                                  " (ask pdb200_layer_norm_grad_workspace_bytes)");
    if (!is_aligned16(ws)) return fail(PDB200_INVALID_ARGUMENT, "layer_norm_grad: workspace must be 16-byte aligned");
  }
  U* part = static_cast<U*>(ws);
  if (rows > 0x7fffffffLL) return fail(PDB200_UNSUPPORTED, "layer_norm_grad: too many rows for one launch");
  auto fold = [&](int nparts) -> int {
    dim3 grid((unsigned)((D + 255) / 256), 2);
    ln_grad_fold_kernel<P, U><<<grid, 256, 0, st>>>(part, nparts, D, dscale, dbias);
    return after_launch("ln_grad_fold");
  };
  // ---- pipe route (float accumulation types only; dx requested) ----
  if constexpr (!std::is_same<T, double>::value) {
    bool vec = (D % VN == 0) && is_aligned16(x) && is_aligned16(dy) && dx != nullptr && is_aligned16(dx);
    if (vec && scale) vec = is_aligned16(scale);
    const int64_t nchunk = D / VN;
    if (vec && nchunk >= 64 && nchunk <= 256 * 4 && rows >= 4 * (int64_t)sm_count()) {
      const int C = nchunk >= 512 ? 4 : (nchunk >= 256 ? 2 : 1);
      int threads = (int)((nchunk + C - 1) / C);
      threads = (threads + 31) / 32 * 32;
      const bool exact = (int64_t)threads * C == nchunk;
      const size_t stage_bytes = 2 * (size_t)D * sizeof(T);
      int nparts = 0;
      auto launch = [&](auto kern) -> int {
        int per_sm = 0, stages = 0;
        size_t smem = 0;
        FuncAttrHold hold;  // released when this lambda returns, i.e. after the launch
        // CTAs of up to 128 threads (rows of <= 4096 16-bit elements) run four to an SM, wider ones two: bf16 [.,1024]
        // 1.58 -> 0.92 ms, [.,2048] 0.91 -> 0.60, [.,4096] 0.67 -> 0.55 on 2^29-element tensors (profiles/r2_notes.txt section 14)
        if (!pipe_geometry(kern, threads, stage_bytes, threads <= 128 ? 4 : 2, 3, &per_sm, &stages, &smem, &hold)) return -1;
        int64_t grid = (int64_t)per_sm * sm_count();
        if (grid > rows) grid = rows;
        if (grid > kLnGradMaxParts) grid = kLnGradMaxParts;
        nparts = (int)grid;
        kern<<<(unsigned)grid, threads, smem, st>>>(x, dy, scale, mean, var, dx, part, rows, (int)D, stages, eps,
                                                    (float)(1.0 / (double)D));
        return after_launch("ln_grad_pipe");
      };
      int rc = -1;
#define PDB_LNG3(CC, HS, PA) rc = exact ? launch(ln_grad_pipe_kernel<T, P, CC, HS, PA, true>) \
                                        : launch(ln_grad_pipe_kernel<T, P, CC, HS, PA, false>);
#define PDB_LNG2(CC)                                                    \
  if (scale && params) { PDB_LNG3(CC, true, true) }                     \
  else if (scale) { PDB_LNG3(CC, true, false) }                         \
  else if (params) { PDB_LNG3(CC, false, true) }                        \
  else { PDB_LNG3(CC, false, false) }
      if (C == 4) { PDB_LNG2(4) } else if (C == 2) { PDB_LNG2(2) } else { PDB_LNG2(1) }
#undef PDB_LNG2
#undef PDB_LNG3
      if (rc == PDB200_OK) return params ? fold(nparts) : PDB200_OK;
      if (rc > 0) return rc;  // a real launch failure; rc < 0 means "does not fit": use the fallback
    }
  }
  // ---- fallback ----
  if (dx != nullptr) {
    int threads = 128;
    while (threads < 1024 && threads * 8 < D) threads <<= 1;
    ln_grad_dx_rows_kernel<T, P><<<(unsigned)rows, threads, 0, st>>>(x, dy, scale, mean, var, dx, D, eps,
                                                                     1.0 / (double)D);
    int rc = after_launch("ln_grad_dx_rows");
    if (rc != PDB200_OK) return rc;
  }
  if (params) {
    const int64_t col_blocks = (D + 127) / 128;
    int64_t nparts = (4 * (int64_t)sm_count() + col_blocks - 1) / col_blocks;
    nparts = std::max<int64_t>(1, std::min<int64_t>(std::min<int64_t>(nparts, rows), kLnGradMaxParts));
    const int64_t slab = (rows + nparts - 1) / nparts;
    nparts = (rows + slab - 1) / slab;
    if (col_blocks > 0x7fffffffLL) return fail(PDB200_UNSUPPORTED, "layer_norm_grad: D too large");
    dim3 grid((unsigned)col_blocks, (unsigned)nparts);
    ln_grad_params_kernel<T><<<grid, 128, 0, st>>>(x, dy, mean, var, part, rows, D, slab, eps);
    int rc = after_launch("ln_grad_params");
    if (rc != PDB200_OK) return rc;
    return fold((int)nparts);
  }
  return PDB200_OK;
}

}  // namespace pdb200

extern "C" {

size_t pdb200_layer_norm_grad_workspace_bytes(const int64_t* dims, int rank, int begin_norm_axis, int dtype) {
  if (dims == nullptr || rank < 1 || rank > PDB200_MAX_RANK || begin_norm_axis < 0 || begin_norm_axis > rank) return 0;
  int64_t D = 1;
  for (int i = begin_norm_axis; i < rank; ++i) D *= dims[i];
  return pdb200::ln_grad_scratch_bytes(D, dtype == PDB200_FLOAT64 ? 8 : 4);
}

int pdb200_layer_norm_grad(const void* x, const int64_t* dims, int rank, const void* scale, int param_dtype,
                           const void* mean, const void* variance, const void* out_grad, float epsilon,
                           int begin_norm_axis, int dtype, void* x_grad, void* scale_grad, void* bias_grad,
                           void* workspace, size_t workspace_bytes, void* stream) {
  using namespace pdb200;
  PDB_CHECK_ARG(rank >= 1 && rank <= PDB200_MAX_RANK && dims != nullptr, "layer_norm_grad: rank must be in [1, 9]");
  PDB_CHECK_ARG(begin_norm_axis >= 0 && begin_norm_axis <= rank,
                "'begin_norm_axis' must be less than the dimensions of X, but received 'begin_norm_axis' is " +
                    std::to_string(begin_norm_axis));
  int64_t rows = 1, D = 1;
  for (int i = 0; i < rank; ++i) {
    PDB_CHECK_ARG(dims[i] >= 0, "layer_norm_grad: negative dimension");
    (i < begin_norm_axis ? rows : D) *= dims[i];
  }
  if (rows * D == 0) return PDB200_OK;
  PDB_CHECK_ARG(x && mean && variance && out_grad, "layer_norm_grad: NULL tensor pointer");
  PDB_CHECK_ARG(x_grad || scale_grad || bias_grad, "layer_norm_grad: no output requested");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const bool has_param = scale != nullptr || scale_grad != nullptr || bias_grad != nullptr;
  const bool same = !has_param || param_dtype == dtype;
  if (has_param && !same && param_dtype != PDB200_FLOAT32)
    return fail(PDB200_INVALID_ARGUMENT,
                "Unsupported data type of Scale and Bias: the dtype must be the same as X's or float32");
#define PDB_LNG_CASE(DT)                                                                                        \
  case DT: {                                                                                                    \
    using T = typename CType<DT>::type;                                                                         \
    if (same)                                                                                                   \
      return layer_norm_grad_typed<T, T>(static_cast<const T*>(x), static_cast<const T*>(scale), mean, variance, \
                                         static_cast<const T*>(out_grad), static_cast<T*>(x_grad),              \
                                         static_cast<T*>(scale_grad), static_cast<T*>(bias_grad), rows, D,      \
                                         epsilon, workspace, workspace_bytes, st);                              \
    return layer_norm_grad_typed<T, float>(static_cast<const T*>(x), static_cast<const float*>(scale), mean,    \
                                           variance, static_cast<const T*>(out_grad), static_cast<T*>(x_grad),  \
                                           static_cast<float*>(scale_grad), static_cast<float*>(bias_grad),     \
                                           rows, D, epsilon, workspace, workspace_bytes, st);                   \
  }
  switch (dtype) {
    PDB_LNG_CASE(PDB200_FLOAT32)
    PDB_LNG_CASE(PDB200_BFLOAT16)
    PDB_LNG_CASE(PDB200_FLOAT16)
    case PDB200_FLOAT64: {
      if (!same) return fail(PDB200_INVALID_ARGUMENT, "layer_norm_grad: float64 input needs float64 scale");
      return layer_norm_grad_typed<double, double>(
          static_cast<const double*>(x), static_cast<const double*>(scale), mean, variance,
          static_cast<const double*>(out_grad), static_cast<double*>(x_grad), static_cast<double*>(scale_grad),
          static_cast<double*>(bias_grad), rows, D, epsilon, workspace, workspace_bytes, st);
    }
    default:
      return fail(PDB200_UNSUPPORTED, "layer_norm_grad: dtype must be float32/float64/float16/bfloat16, got code " +
                                          std::to_string(dtype));
  }
#undef PDB_LNG_CASE
}

}  // extern "C"
