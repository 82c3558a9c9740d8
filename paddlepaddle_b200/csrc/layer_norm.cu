// layer_norm.cu — layer_norm forward, hand-written for sm_100a.
//
// Replaces (fresh design): phi::LayerNormKernel<T, GPUContext> and its three forward kernels
//   fast_ln_fwd_kernel, LayerNormFwdWithWelford, LayerNormForward<T,U,512>
//   paddle/phi/kernels/gpu/layer_norm_kernel.cu:297-665,
//   paddle/phi/kernels/funcs/layer_norm_impl.cu.h:178-508.
// The reference's route for D = 8192 (config 3) is the generic kernel: one 512-thread CTA per
// row, scalar 2-byte loads, x read twice, variance as E[x^2]-E[x]^2.  Here every route reads
// x ONCE from HBM: the row lives in registers between the statistics and the normalisation.
//
// Numerics follow the CPU kernel the parity oracle is built from
//   paddle/phi/kernels/cpu/layer_norm_kernel.cc:68-102 (two-pass: mean, then mean of
//   (x-mean)^2, biased), computed in U = float (double for fp64 input);
//   y = ((x-mean) * (1/sqrt(var+eps))) * scale + bias, product and sum rounded separately.
//
// Routes (x viewed as [rows][D]):
//   rows_reg   : G lanes (<= 32) per row, C <= 8 16-byte chunks per lane     (D <= 256*V)
//   block_reg  : one CTA (128..1024 threads) per row, C <= 4 chunks a thread (D <= 4096*V)
//   cluster    : long rows — one thread-block CLUSTER (2..8 CTAs) per row, each CTA keeps its slice
//                in shared memory, the row sums are combined through distributed shared memory
//   block_strm : longer still — three streaming sweeps
//   pipe       : the fast path for many rows (config 3): persistent CTAs fed by the TMA row
//                pipeline of rowpipe.cuh; scale/bias stay in registers across rows
#include <cooperative_groups.h>

#include <map>
#include <tuple>

#include "common.cuh"
#include "rowpipe.cuh"
#include "shape.h"

namespace pdb200 {

template <typename U> __device__ __forceinline__ U rsqrt_exact(U v);
template <> __device__ __forceinline__ float rsqrt_exact<float>(float v) { return 1.0f / sqrtf(v); }
template <> __device__ __forceinline__ double rsqrt_exact<double>(double v) { return 1.0 / sqrt(v); }

// sum * (1.0 / n) with the double-precision reciprocal the CPU RowwiseMean uses
// (funcs/math_function_impl.h:180-189)
template <typename U>
__device__ __forceinline__ U mean_of(U sum, double inv_n) { return (U)((double)sum * inv_n); }

template <typename U> struct AddU { __device__ __forceinline__ U operator()(U a, U b) const { return a + b; } };

template <typename U> __device__ __forceinline__ U mul_rn(U a, U b);
template <> __device__ __forceinline__ float mul_rn<float>(float a, float b) { return __fmul_rn(a, b); }
template <> __device__ __forceinline__ double mul_rn<double>(double a, double b) { return __dmul_rn(a, b); }
template <typename U> __device__ __forceinline__ U add_rn(U a, U b);
template <> __device__ __forceinline__ float add_rn<float>(float a, float b) { return __fadd_rn(a, b); }
template <> __device__ __forceinline__ double add_rn<double>(double a, double b) { return __dadd_rn(a, b); }

// Normalise one chunk held in registers and store it.
template <typename T, typename P, typename U, int V, bool VECTOR>
__device__ __forceinline__ void ln_store_chunk(const U* e, U mean, U rstd, const P* __restrict__ scale,
                                               const P* __restrict__ bias, T* dst, int64_t off) {
  U o[V];
#pragma unroll
  for (int k = 0; k < V; ++k) o[k] = mul_rn<U>(e[k] - mean, rstd);
  if (scale != nullptr) {
    if constexpr (VECTOR && sizeof(P) == sizeof(T)) {
      Vec<P> s = ld_cached(scale + off);
#pragma unroll
      for (int k = 0; k < V; ++k) o[k] = mul_rn<U>(o[k], cvt<U>(s.v[k]));
    } else {
#pragma unroll
      for (int k = 0; k < V; ++k) o[k] = mul_rn<U>(o[k], cvt<U>(scale[off + k]));
    }
  }
  if (bias != nullptr) {
    if constexpr (VECTOR && sizeof(P) == sizeof(T)) {
      Vec<P> b = ld_cached(bias + off);
#pragma unroll
      for (int k = 0; k < V; ++k) o[k] = add_rn<U>(o[k], cvt<U>(b.v[k]));
    } else {
#pragma unroll
      for (int k = 0; k < V; ++k) o[k] = add_rn<U>(o[k], cvt<U>(bias[off + k]));
    }
  }
  if constexpr (VECTOR) {
    Vec<T> ov;
#pragma unroll
    for (int k = 0; k < V; ++k) ov.v[k] = cvt<T>(o[k]);
    st_stream(dst + off, ov);
  } else {
    dst[off] = cvt<T>(o[0]);
  }
}

// ---- rows_reg: G lanes per row ------------------------------------------------------------
template <typename T, typename P, int C, bool VECTOR>
__global__ void __launch_bounds__(256) ln_rows_reg_kernel(const T* __restrict__ x,
                                                          const P* __restrict__ scale,
                                                          const P* __restrict__ bias,
                                                          T* __restrict__ y, typename AccType<T>::type* mean_out,
                                                          typename AccType<T>::type* var_out, int64_t rows,
                                                          int D, int G, typename AccType<T>::type eps,
                                                          double inv_d) {
  using U = typename AccType<T>::type;
  constexpr int V = VECTOR ? Vec<T>::N : 1;
  const int rows_per_cta = blockDim.x / G;
  const int g = threadIdx.x / G, lane = threadIdx.x - g * G;
  const int64_t row = (int64_t)blockIdx.x * rows_per_cta + g;
  const bool valid = row < rows;
  const T* src = x + row * D;
  U e[C * V];
  U sum = U(0);
#pragma unroll
  for (int c = 0; c < C; ++c) {
    const int off = (c * G + lane) * V;
    if (valid && off < D) {
      if constexpr (VECTOR) {
        Vec<T> v = ld_stream(src + off);
#pragma unroll
        for (int k = 0; k < V; ++k) e[c * V + k] = cvt<U>(v.v[k]);
      } else {
        e[c] = cvt<U>(src[off]);
      }
#pragma unroll
      for (int k = 0; k < V; ++k) sum += e[c * V + k];
    } else {
#pragma unroll
      for (int k = 0; k < V; ++k) e[c * V + k] = U(0);
    }
  }
  for (int s = G >> 1; s >= 1; s >>= 1) sum += shfl_xor_any(sum, s);
  const U mean = mean_of<U>(sum, inv_d);
  U sq = U(0);
#pragma unroll
  for (int c = 0; c < C; ++c) {
    const int off = (c * G + lane) * V;
    if (off < D) {
#pragma unroll
      for (int k = 0; k < V; ++k) {
        const U d = e[c * V + k] - mean;
        sq += d * d;
      }
    }
  }
  for (int s = G >> 1; s >= 1; s >>= 1) sq += shfl_xor_any(sq, s);
  const U var = mean_of<U>(sq, inv_d);
  const U rstd = rsqrt_exact<U>(var + eps);
  if (valid && lane == 0) {
    if (mean_out) mean_out[row] = mean;
    if (var_out) var_out[row] = var;
  }
  if (!valid) return;
  T* dst = y + row * D;
#pragma unroll
  for (int c = 0; c < C; ++c) {
    const int off = (c * G + lane) * V;
    if (off < D) ln_store_chunk<T, P, U, V, VECTOR>(&e[c * V], mean, rstd, scale, bias, dst, off);
  }
}

// ---- block_reg: one CTA per row -----------------------------------------------------------
template <typename T, typename P, int C, bool VECTOR>
__global__ void __launch_bounds__(1024) ln_block_reg_kernel(const T* __restrict__ x,
                                                            const P* __restrict__ scale,
                                                            const P* __restrict__ bias,
                                                            T* __restrict__ y, typename AccType<T>::type* mean_out,
                                                            typename AccType<T>::type* var_out, int D,
                                                            typename AccType<T>::type eps, double inv_d) {
  using U = typename AccType<T>::type;
  constexpr int V = VECTOR ? Vec<T>::N : 1;
  __shared__ U red[33];
  const int64_t row = blockIdx.x;
  const T* src = x + row * D;
  const int nt = blockDim.x;
  U e[C * V];
  U sum = U(0);
  // all loads first (C independent 16-byte requests per thread), then the arithmetic
  if constexpr (VECTOR) {
    Vec<T> raw[C];
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + threadIdx.x) * V;
      if (off < D) raw[c] = ld_stream(src + off);
    }
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + threadIdx.x) * V;
#pragma unroll
      for (int k = 0; k < V; ++k) {
        e[c * V + k] = off < D ? cvt<U>(raw[c].v[k]) : U(0);
        sum += e[c * V + k];
      }
    }
  } else {
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = c * nt + threadIdx.x;
      e[c] = off < D ? cvt<U>(src[off]) : U(0);
      sum += e[c];
    }
  }
  sum = block_reduce(sum, AddU<U>(), U(0), red);
  const U mean = mean_of<U>(sum, inv_d);
  U sq = U(0);
#pragma unroll
  for (int c = 0; c < C; ++c) {
    const int off = (c * nt + threadIdx.x) * V;
    if (off < D) {
#pragma unroll
      for (int k = 0; k < V; ++k) {
        const U d = e[c * V + k] - mean;
        sq += d * d;
      }
    }
  }
  sq = block_reduce(sq, AddU<U>(), U(0), red);
  const U var = mean_of<U>(sq, inv_d);
  const U rstd = rsqrt_exact<U>(var + eps);
  if (threadIdx.x == 0) {
    if (mean_out) mean_out[row] = mean;
    if (var_out) var_out[row] = var;
  }
  T* dst = y + row * D;
#pragma unroll
  for (int c = 0; c < C; ++c) {
    const int off = (c * nt + threadIdx.x) * V;
    if (off < D) ln_store_chunk<T, P, U, V, VECTOR>(&e[c * V], mean, rstd, scale, bias, dst, off);
  }
}

// ---- pipe: persistent CTAs, rows staged in shared memory by cp.async.bulk -------------------
// blockDim.x threads per row, C chunks of V elements per thread (chunk c of thread t covers
// elements (c*blockDim.x + t)*V ..), `stages` rows in flight per CTA.  x is read ONCE from HBM by
// the copy engine.  T is float / bf16 / fp16 (U = float).  The first-round kernel was issue-bound
// (profiles/r1_ln_bf16.txt: 26 instructions per element): here AFFINE (scale and bias present) and
// EXACT (D == C*blockDim.x*V) are compile-time so no per-element predicates remain, bf16 unpacks
// in one instruction per element, sums run in four independent chains, the affine step is one
// FFMA, and each CTA reduction costs one barrier.
template <typename T, typename P, int C, bool AFFINE, bool EXACT, int MAXT, bool PSM = false>
__global__ void __launch_bounds__(MAXT, MAXT <= 256 ? 2 : 1)
ln_pipe_kernel(const T* __restrict__ x, const P* __restrict__ scale, const P* __restrict__ bias,
               T* __restrict__ y, float* mean_out, float* var_out, int64_t rows, int D, int stages, float eps,
               double inv_d, int contiguous) {
  constexpr int V = Vec<T>::N;
  constexpr int PV = Vec<P>::N;  // params per 16-byte load (== V, or V/2 when P is float and T 16-bit)
  extern __shared__ __align__(128) unsigned char ring_raw[];
  T* ring = reinterpret_cast<T*>(ring_raw);
  __shared__ uint64_t full[kPipeMaxStages];
  __shared__ float red[2][32];
  const int tid = threadIdx.x, nt = blockDim.x;
  const uint32_t row_bytes = (uint32_t)D * sizeof(T);
  // Row -> CTA mapping: strided (row = cta + k*grid) or one contiguous slab per CTA.
  int64_t row_first = blockIdx.x, row_step = gridDim.x, row_end = rows;
  if (contiguous) {
    const int64_t slab = (rows + gridDim.x - 1) / gridDim.x;
    row_first = (int64_t)blockIdx.x * slab;
    row_step = 1;
    row_end = row_first + slab < rows ? row_first + slab : rows;
  }
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) mbar_init(&full[s], 1);
    mbar_fence_init();
  }
  block_reduce_1sync_init(red[0], 0.f);
  block_reduce_1sync_init(red[1], 0.f);
  // PSM: scale and bias copied ONCE per (persistent) CTA into shared memory behind the ring and read from there for every
  // row.  Through L1 the parameters cost 2 * D * sizeof(P) bytes of global-load traffic per row — four times the row
  // itself for float parameters of a bf16 row — and that path sustains about half the bytes per clock of a shared-memory
  // read: bf16 [65536,8192] with float parameters ran 0.49 ms against 0.33 with bf16 parameters (profiles/r2_notes.txt).
  // Layout: thread t's parameters of chunk c arrive as V / PV 16-byte pieces; piece j of all threads is stored together
  // ([c][j][t][PV]) so that a warp's 16-byte reads are consecutive — in the tensor's own order a float parameter vector of a
  // 16-bit row sits at a 32-byte stride per thread: two-way bank conflicts on every read (shared-memory pipe 69 % busy,
  // profiles/r2h_ln_f32p.txt), and the same two passes on the L1 path.
  P* const sw = reinterpret_cast<P*>(ring_raw + (size_t)stages * row_bytes);
  P* const sb = sw + (size_t)C * nt * V;
  if constexpr (PSM) {
    constexpr int J = V / PV;
    for (int i = tid; i < D / PV; i += nt) {  // i: 16-byte piece of the parameter vectors
      const int vi = i / J, j = i - vi * J;   // vi: which 16-byte chunk of the row; j: which piece of its parameters
      const int c = vi / nt, t = vi - c * nt;
      const int dsti = ((c * J + j) * nt + t) * PV;
      *reinterpret_cast<uint4*>(sw + dsti) = *reinterpret_cast<const uint4*>(scale + (size_t)i * PV);
      *reinterpret_cast<uint4*>(sb + dsti) = *reinterpret_cast<const uint4*>(bias + (size_t)i * PV);
    }
  }
  __syncthreads();
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) {
      const int64_t r = row_first + s * row_step;
      if (r < row_end) {
        mbar_expect_tx(&full[s], row_bytes);
        bulk_load_row(ring + (size_t)s * D, x + r * D, row_bytes, &full[s]);
      }
    }
  }
  // scale / bias: re-read through L1 for every row (16-byte loads).  Keeping them packed in
  // registers instead (kParamRegs) costs 32 registers and measured slower on B200; the L1 path
  // needs the carve-out to leave L1 room for 2*D*sizeof(P) bytes, which pipe_geometry arranges.
  constexpr bool kParamRegs = false;  // measured slower than the L1 path (0.368 vs 0.339 ms); kept for reference
  Vec<P> wreg[kParamRegs ? C : 1], breg[kParamRegs ? C : 1];
  if constexpr (kParamRegs) {
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + tid) * V;
      if (EXACT || off < D) {
        wreg[c] = ld_cached(scale + off);
        breg[c] = ld_cached(bias + off);
      }
    }
  }
  int s = 0;
  uint32_t phase = 0;
  for (int64_t row = row_first; row < row_end; row += row_step) {
    mbar_wait(&full[s], phase);
    const T* st = ring + (size_t)s * D;
    Vec<T> raw[C];
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + tid) * V;
      if (EXACT || off < D) raw[c] = lds_vec(st + off);
    }
    float e[C * V];
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + tid) * V;
      if (EXACT || off < D) {
        unpack_to_float<T>(raw[c], &e[c * V]);
      } else {
#pragma unroll
        for (int k = 0; k < V; ++k) e[c * V + k] = 0.f;
      }
#pragma unroll
      for (int k = 0; k < V; ++k) acc[k & 3] += e[c * V + k];
    }
    float sum = (acc[0] + acc[1]) + (acc[2] + acc[3]);
    sum = block_reduce_1sync(sum, AddU<float>(), red[0]);  // its barrier also means: stage s fully read
    if (tid == 0) {
      const int64_t nrow = row + (int64_t)stages * row_step;
      if (nrow < row_end) {
        fence_proxy_async_smem();
        mbar_expect_tx(&full[s], row_bytes);
        bulk_load_row(ring + (size_t)s * D, x + nrow * D, row_bytes, &full[s]);
      }
    }
    const float mean = mean_of<float>(sum, inv_d);
    acc[0] = acc[1] = acc[2] = acc[3] = 0.f;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + tid) * V;
      if (EXACT || off < D) {
#pragma unroll
        for (int k = 0; k < V; ++k) {
          const float d = e[c * V + k] - mean;
          e[c * V + k] = d;  // keep the centred value: the normalisation needs exactly this
          acc[k & 3] = fmaf(d, d, acc[k & 3]);
        }
      }
    }
    float sq = (acc[0] + acc[1]) + (acc[2] + acc[3]);
    sq = block_reduce_1sync(sq, AddU<float>(), red[1]);
    const float var = mean_of<float>(sq, inv_d);
    const float rstd = rsqrt_exact<float>(var + eps);
    if (tid == 0) {
      if (mean_out) mean_out[row] = mean;
      if (var_out) var_out[row] = var;
    }
    T* dst = y + row * D;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + tid) * V;
      if (EXACT || off < D) {
        if constexpr (AFFINE) {
          float w[V], b[V];
          if constexpr (kParamRegs) {
            unpack_to_float<P>(wreg[c], w);
            unpack_to_float<P>(breg[c], b);
          } else {
#pragma unroll
            for (int j = 0; j < V / PV; ++j) {
              if constexpr (PSM) {
                unpack_to_float<P>(lds_vec(sw + ((c * (V / PV) + j) * nt + tid) * PV), &w[j * PV]);
                unpack_to_float<P>(lds_vec(sb + ((c * (V / PV) + j) * nt + tid) * PV), &b[j * PV]);
              } else {
                unpack_to_float<P>(ld_cached(scale + off + j * PV), &w[j * PV]);
                unpack_to_float<P>(ld_cached(bias + off + j * PV), &b[j * PV]);
              }
            }
          }
          float o[V];
#pragma unroll
          for (int k = 0; k < V; ++k) {
            // (d * rstd) * w + b with the product-sum fused: <= 0.5 ulp(fp32) from the CPU kernel's
            // separately rounded multiply and add
            o[k] = fmaf(__fmul_rn(e[c * V + k], rstd), w[k], b[k]);
          }
          Vec<T> ov;
          pack_from_float<T>(o, ov);
          st_stream(dst + off, ov);
        } else {
          // e[] already holds x - mean: hand ln_store_chunk a zero mean
          ln_store_chunk<T, P, float, V, true>(&e[c * V], 0.f, rstd, scale, bias, dst, off);
        }
      }
    }
    if (++s == stages) {
      s = 0;
      phase ^= 1;
    }
  }
}

// ---- cluster: one thread-block cluster per row, DSMEM combine -----------------------------------
// CTA `rank` owns elements [rank*slice, +slice) of the row, staged once in shared memory in their
// storage type (the only HBM read of x).  Sum and sum of squared deviations are produced per CTA,
// published in shared memory and folded by every CTA, in rank order, through DSMEM after a cluster
// barrier.  Vector path only (D % V == 0, 16-byte aligned bases).
template <typename T, typename P>
__global__ void __launch_bounds__(1024) ln_rows_cluster_kernel(const T* __restrict__ x, const P* __restrict__ scale,
                                                               const P* __restrict__ bias, T* __restrict__ y,
                                                               typename AccType<T>::type* mean_out,
                                                               typename AccType<T>::type* var_out, int64_t D,
                                                               int64_t slice, typename AccType<T>::type eps,
                                                               double inv_d) {
  namespace cg = cooperative_groups;
  using U = typename AccType<T>::type;
  constexpr int V = Vec<T>::N;
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned rank = cluster.block_rank(), cs = cluster.num_blocks();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* srow = reinterpret_cast<T*>(smem_raw);
  __shared__ U red[2][32];
  __shared__ U xchg[2];
  const int64_t row = (int64_t)blockIdx.x / cs;
  const int64_t begin = (int64_t)rank * slice;
  int64_t n = D - begin;
  n = n < 0 ? 0 : (n > slice ? slice : n);
  const T* src = x + row * D + begin;
  T* dst = y + row * D + begin;
  const int64_t step = (int64_t)blockDim.x * V;
  // HBM -> shared with ONE bulk async copy for the whole slice, then the local sum from shared memory
  __shared__ uint64_t bar;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    mbar_fence_init();
  }
  block_reduce_1sync_init(red[0], U(0));
  block_reduce_1sync_init(red[1], U(0));
  __syncthreads();
  if (n > 0) {
    if (threadIdx.x == 0) {
      mbar_expect_tx(&bar, (uint32_t)(n * sizeof(T)));
      bulk_load_row(srow, src, (uint32_t)(n * sizeof(T)), &bar);
    }
    mbar_wait(&bar, 0);
  }
  // Mean and sum of squared deviations OF THIS SLICE (two sweeps over shared memory, CTA reductions only), published as
  // a pair; after ONE cluster barrier every CTA folds the cs pairs with the pairwise-variance rule
  //     mean = sum_r n_r mean_r / D,     M2 = sum_r (M2_r + n_r (mean_r - mean)^2),     var = M2 / D
  // (as accurate as two passes over the whole row; the first version needed two cluster barriers).
  // (whole-vector unpack, four independent chains: the sweeps were issue-bound — 87 % issue slots at 44 % DRAM,
  // profiles/r2_ln_vocab_bf16.txt)
  // (one-barrier CTA reductions, the cluster's pairs read one per lane and folded with shuffles: the first version walked
  // the ranks in a loop of dependent DSMEM loads in every thread and used three-barrier reductions — with 2 to 8 vectors per
  // thread those fixed costs were most of the kernel: 1024 threads per CTA ran 3x slower than 128.)
  // float accumulators: the sweeps run on packed fp32 pairs (FADD2 / FFMA2: the same roundings, half the issue slots — this
  // kernel was issue-bound, 27 instructions per element at 82 % issue, profiles/r2g_ln_vocab_bf16.txt) with 32-bit indices
  U mean_loc, sq;
  const int ni = (int)n, stepi = (int)step, i0 = (int)threadIdx.x * V;
  if constexpr (std::is_same<U, float>::value) {
    F2 a2[2] = {f2_splat(0.f), f2_splat(0.f)};
#pragma unroll 2
    for (int i = i0; i < ni; i += stepi) {
      float e[V];
      unpack_to_float<T>(lds_vec(srow + i), e);
#pragma unroll
      for (int j = 0; j < V / 2; ++j) a2[j & 1] = f2_add(a2[j & 1], f2_make(e[2 * j], e[2 * j + 1]));
    }
    const F2 s2 = f2_add(a2[0], a2[1]);
    const float sum = block_reduce_1sync(f2_lo(s2) + f2_hi(s2), AddU<U>(), red[0]);
    mean_loc = n > 0 ? sum / (U)n : U(0);
    const F2 nm2 = f2_splat(-mean_loc);
    a2[0] = a2[1] = f2_splat(0.f);
#pragma unroll 2
    for (int i = i0; i < ni; i += stepi) {
      float e[V];
      unpack_to_float<T>(lds_vec(srow + i), e);
#pragma unroll
      for (int j = 0; j < V / 2; ++j) {
        const F2 d = f2_add(f2_make(e[2 * j], e[2 * j + 1]), nm2);
        a2[j & 1] = f2_fma(d, d, a2[j & 1]);
      }
    }
    const F2 q2 = f2_add(a2[0], a2[1]);
    sq = block_reduce_1sync(f2_lo(q2) + f2_hi(q2), AddU<U>(), red[1]);
  } else {
    U acc4[4] = {U(0), U(0), U(0), U(0)};
    for (int i = i0; i < ni; i += stepi) {
      const Vec<T> v = lds_vec(srow + i);
#pragma unroll
      for (int k = 0; k < V; ++k) acc4[k & 3] += cvt<U>(v.v[k]);
    }
    U sum = (acc4[0] + acc4[1]) + (acc4[2] + acc4[3]);
    sum = block_reduce_1sync(sum, AddU<U>(), red[0]);
    mean_loc = n > 0 ? sum / (U)n : U(0);
    acc4[0] = acc4[1] = acc4[2] = acc4[3] = U(0);
    for (int i = i0; i < ni; i += stepi) {
      const Vec<T> v = lds_vec(srow + i);
#pragma unroll
      for (int k = 0; k < V; ++k) {
        const U d = cvt<U>(v.v[k]) - mean_loc;
        acc4[k & 3] += d * d;
      }
    }
    sq = (acc4[0] + acc4[1]) + (acc4[2] + acc4[3]);
    sq = block_reduce_1sync(sq, AddU<U>(), red[1]);
  }
  if (threadIdx.x == 0) {
    xchg[0] = mean_loc;
    xchg[1] = sq;
  }
  cluster_sync_published_by(threadIdx.x == 0);
  // lane r < cs of every warp reads rank r's (mean, M2): one DSMEM round trip, then two shuffle folds
  const unsigned lane = threadIdx.x & 31;
  U mean_r = U(0), sq_r = U(0), n_r = U(0);
  if (lane < cs) {
    const U* q = cluster.map_shared_rank(xchg, lane);
    mean_r = q[0];
    sq_r = q[1];
    int64_t nr = D - (int64_t)lane * slice;
    nr = nr < 0 ? 0 : (nr > slice ? slice : nr);
    n_r = (U)nr;
  }
  const U total = group_reduce<8>(n_r * mean_r, AddU<U>());  // cs <= 8: lanes 0..7 hold the pairs; every lane of that group gets the sum
  const U mean = mean_of<U>(shfl_idx_any(total, 0), inv_d);
  const U dm = mean_r - mean;
  const U total_sq = group_reduce<8>(sq_r + n_r * dm * dm, AddU<U>());
  const U var = mean_of<U>(shfl_idx_any(total_sq, 0), inv_d);
  const U rstd = rsqrt_exact<U>(var + eps);
  if (rank == 0 && threadIdx.x == 0) {
    if (mean_out) mean_out[row] = mean;
    if (var_out) var_out[row] = var;
  }
  // every peer's pair has been read: arrive for the exit barrier now (relaxed — a release arrive after the stores would
  // make every thread wait for its stores to drain), wait for it at the very end
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  const P* sc = scale ? scale + begin : nullptr;
  const P* bi = bias ? bias + begin : nullptr;
  if constexpr (std::is_same<U, float>::value) {
    // o = fma((e - mean) * rstd, w, b) with the product rounded on its own, as ln_pipe_kernel: FADD2, FMUL2, FFMA2
    const F2 nm2 = f2_splat(-mean), r2 = f2_splat(rstd);
    auto sweep = [&](auto has_w, auto has_b) {
      constexpr int PV = Vec<P>::N;
#pragma unroll 2
      for (int i = i0; i < ni; i += stepi) {
        float e[V], w[V], b[V];
        unpack_to_float<T>(lds_vec(srow + i), e);
        if constexpr (decltype(has_w)::value) {
#pragma unroll
          for (int j = 0; j < V / PV; ++j) unpack_to_float<P>(ld_cached(sc + i + j * PV), &w[j * PV]);
        }
        if constexpr (decltype(has_b)::value) {
#pragma unroll
          for (int j = 0; j < V / PV; ++j) unpack_to_float<P>(ld_cached(bi + i + j * PV), &b[j * PV]);
        }
#pragma unroll
        for (int j = 0; j < V / 2; ++j) {
          F2 t = f2_mul(f2_add(f2_make(e[2 * j], e[2 * j + 1]), nm2), r2);
          if constexpr (decltype(has_w)::value && decltype(has_b)::value) t = f2_fma(t, f2_make(w[2 * j], w[2 * j + 1]), f2_make(b[2 * j], b[2 * j + 1]));
          else if constexpr (decltype(has_w)::value) t = f2_fma(t, f2_make(w[2 * j], w[2 * j + 1]), f2_splat(0.f));
          else if constexpr (decltype(has_b)::value) t = f2_fma(t, f2_splat(1.f), f2_make(b[2 * j], b[2 * j + 1]));
          else t = f2_fma(t, f2_splat(1.f), f2_splat(0.f));
          e[2 * j] = f2_lo(t);
          e[2 * j + 1] = f2_hi(t);
        }
        Vec<T> ov;
        pack_from_float<T>(e, ov);
        st_stream(dst + i, ov);
      }
    };
    // float parameters of a 16-bit row: a thread's 8 parameters are 32 bytes, i.e. two 16-byte loads at a 32-byte stride across
    // the warp (two passes per request).  That case walks the slice in units of FOUR elements per thread instead: 8-byte
    // reads of x, one 16-byte load each of scale and bias at a 16-byte stride, 8-byte stores (0.61 -> see r2_notes §11).
    if constexpr (sizeof(P) == 4 && sizeof(T) == 2) {
      if (sc != nullptr && bi != nullptr) {
        const int h0 = (int)threadIdx.x * 4, hstep = (int)blockDim.x * 4;
#pragma unroll 4
        for (int i = h0; i < ni; i += hstep) {
          const uint2 raw = *reinterpret_cast<const uint2*>(srow + i);
          const float4 w4 = __ldg(reinterpret_cast<const float4*>(sc + i));
          const float4 b4 = __ldg(reinterpret_cast<const float4*>(bi + i));
          float e0, e1, e2, e3;
          if constexpr (std::is_same<T, __nv_bfloat16>::value) {
            e0 = __uint_as_float(raw.x << 16);
            e1 = __uint_as_float(raw.x & 0xffff0000u);
            e2 = __uint_as_float(raw.y << 16);
            e3 = __uint_as_float(raw.y & 0xffff0000u);
          } else {
            const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&raw.x));
            const float2 c = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
            e0 = a.x; e1 = a.y; e2 = c.x; e3 = c.y;
          }
          const F2 t0 = f2_fma(f2_mul(f2_add(f2_make(e0, e1), nm2), r2), f2_make(w4.x, w4.y), f2_make(b4.x, b4.y));
          const F2 t1 = f2_fma(f2_mul(f2_add(f2_make(e2, e3), nm2), r2), f2_make(w4.z, w4.w), f2_make(b4.z, b4.w));
          uint2 o;
          if constexpr (std::is_same<T, __nv_bfloat16>::value) {
            const __nv_bfloat162 p0 = __floats2bfloat162_rn(f2_lo(t0), f2_hi(t0)), p1 = __floats2bfloat162_rn(f2_lo(t1), f2_hi(t1));
            o.x = *reinterpret_cast<const uint32_t*>(&p0);
            o.y = *reinterpret_cast<const uint32_t*>(&p1);
          } else {
            const __half2 p0 = __floats2half2_rn(f2_lo(t0), f2_hi(t0)), p1 = __floats2half2_rn(f2_lo(t1), f2_hi(t1));
            o.x = *reinterpret_cast<const uint32_t*>(&p0);
            o.y = *reinterpret_cast<const uint32_t*>(&p1);
          }
          asm volatile("st.global.L1::no_allocate.v2.u32 [%0], {%1,%2};" ::"l"(dst + i), "r"(o.x), "r"(o.y) : "memory");
        }
        asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
        return;
      }
    }
    if (sc != nullptr && bi != nullptr) sweep(std::true_type{}, std::true_type{});
    else if (sc != nullptr) sweep(std::true_type{}, std::false_type{});
    else if (bi != nullptr) sweep(std::false_type{}, std::true_type{});
    else sweep(std::false_type{}, std::false_type{});
  } else {
    for (int i = i0; i < ni; i += stepi) {
      const Vec<T> v = lds_vec(srow + i);
      U e[V];
#pragma unroll
      for (int k = 0; k < V; ++k) e[k] = cvt<U>(v.v[k]);
      ln_store_chunk<T, P, U, V, true>(e, mean, rstd, sc, bi, dst, i);
    }
  }
  asm volatile("barrier.cluster.wait.aligned;" ::: "memory");  // nobody leaves while a peer may still be reading its xchg[]
}

// ---- block_strm: rows too long for registers ------------------------------------------------
template <typename T, typename P>
__global__ void __launch_bounds__(1024) ln_block_stream_kernel(const T* x, const P* __restrict__ scale,
                                                               const P* __restrict__ bias, T* y,
                                                               typename AccType<T>::type* mean_out,
                                                               typename AccType<T>::type* var_out, int64_t D,
                                                               typename AccType<T>::type eps, double inv_d) {
  using U = typename AccType<T>::type;
  __shared__ U red[33];
  const int64_t row = blockIdx.x;
  const T* src = x + row * D;
  T* dst = y + row * D;
  U sum = U(0);
  for (int64_t i = threadIdx.x; i < D; i += blockDim.x) sum += cvt<U>(src[i]);
  sum = block_reduce(sum, AddU<U>(), U(0), red);
  const U mean = mean_of<U>(sum, inv_d);
  U sq = U(0);
  for (int64_t i = threadIdx.x; i < D; i += blockDim.x) {
    const U d = cvt<U>(src[i]) - mean;
    sq += d * d;
  }
  sq = block_reduce(sq, AddU<U>(), U(0), red);
  const U var = mean_of<U>(sq, inv_d);
  const U rstd = rsqrt_exact<U>(var + eps);
  if (threadIdx.x == 0) {
    if (mean_out) mean_out[row] = mean;
    if (var_out) var_out[row] = var;
  }
  for (int64_t i = threadIdx.x; i < D; i += blockDim.x) {
    U o = mul_rn<U>(cvt<U>(src[i]) - mean, rstd);
    if (scale) o = mul_rn<U>(o, cvt<U>(scale[i]));
    if (bias) o = add_rn<U>(o, cvt<U>(bias[i]));
    dst[i] = cvt<T>(o);
  }
}

// ---- host -------------------------------------------------------------------------------
template <typename T, typename P>
int layer_norm_typed(const T* x, const P* scale, const P* bias, T* y, void* mean, void* var,
                     int64_t rows, int64_t D, float epsilon, cudaStream_t st) {
  using U = typename AccType<T>::type;
  constexpr int VN = Vec<T>::N;
  U* m = static_cast<U*>(mean);
  U* v = static_cast<U*>(var);
  const U eps = static_cast<U>(epsilon);  // static_cast<T>(epsilon) on CPU (cpu/layer_norm_kernel.cc:88)
  const double inv_d = 1.0 / (double)D;
  bool vec = (D % VN == 0) && is_aligned16(x) && is_aligned16(y);
  if (vec && sizeof(P) == sizeof(T)) vec = (!scale || is_aligned16(scale)) && (!bias || is_aligned16(bias));
  const int V = vec ? VN : 1;
  const int64_t nchunk = D / V;
  if (rows > 0x7fffffffLL) return fail(PDB200_UNSUPPORTED, "layer_norm: too many rows for one launch");
#define PDB_LN_C(KERNEL, CC, ...)                                        \
  case CC:                                                               \
    if (vec) KERNEL<T, P, CC, true> __VA_ARGS__;                         \
    else KERNEL<T, P, CC, false> __VA_ARGS__;                            \
    break;
  // Fast path: many rows, 16-byte-multiple rows that fit <= 512 threads x 4 chunks -> TMA row pipeline.
  if (vec && sizeof(T) <= 4 && nchunk >= 64 && nchunk <= 512 * 4 && rows >= 4 * (int64_t)sm_count()) {
    // 32 values a thread (4 chunks of a 16-bit row) down to CTAs of one warp: short rows were the weak spot — bf16 [.,1024]
    // ran 0.737 ms with 128 threads x 1 chunk on 4 persistent CTAs per SM, 0.357 with 32 threads x 4 chunks, 8 CTAs per SM,
    // 16 rows per short-lived CTA; [.,2048] 0.445 -> 0.329 (profiles/r2_notes.txt section 13)
    int C = nchunk >= 128 ? 4 : (nchunk >= 64 ? 2 : 1);
    if (const int o = knobs().ln_pipe_c) C = o;
    int threads = (int)((nchunk + C - 1) / C);
    threads = (threads + 31) / 32 * 32;
    const size_t row_bytes = (size_t)D * sizeof(T);
    const bool affine = scale != nullptr && bias != nullptr;
    // scale and bias resident in shared memory (PSM) when both are given, 16-byte aligned and small enough to leave the ring its room
    const size_t param_bytes = 2 * (size_t)C * threads * VN * sizeof(P);  // the kernel's [C][pieces][threads] layout, D rounded up
    int psm_mode = knobs().ln_param_smem;  // tuning knob: 1 never, 2 always (when it fits); default: float parameters of 16-bit rows
    const bool psm = affine && is_aligned16(scale) && is_aligned16(bias) && D % (16 / (int)sizeof(P)) == 0 && param_bytes <= 64 * 1024 &&
                     (psm_mode == 2 || (psm_mode != 1 && sizeof(P) > sizeof(T)));
    auto launch = [&](auto kern, size_t extra) -> int {
      int per_sm = 0, stages = 0;
      size_t smem = 0;
      int want = 768 / threads;  // 2..8 CTAs per SM, but no more than keep ~64 KB of rows in flight at two stages each
      want = want < 2 ? 2 : (want > 8 ? 8 : want);
      const int by_bytes = (int)((64 * 1024) / (2 * row_bytes));
      if (want > by_bytes) want = by_bytes < 2 ? 2 : by_bytes;
      FuncAttrHold hold;  // released when this lambda returns, i.e. after the launch
      if (!pipe_geometry(kern, threads, row_bytes, want, 3, &per_sm, &stages, &smem, &hold, extra)) return -1;
      int64_t grid = (int64_t)per_sm * sm_count();
      if (threads <= 64) grid = std::max<int64_t>(grid, rows / 16);  // short rows: short-lived CTAs of 16 rows
      if (const int rpc = knobs().pipe_rows_per_cta; rpc >= 1)  // tuning knob: short-lived CTAs instead of a persistent grid
        grid = std::max<int64_t>(grid, rows / rpc);
      if (grid > rows) grid = rows;
      kern<<<(unsigned)grid, threads, smem, st>>>(x, scale, bias, y, m, v, rows, (int)D, stages, eps, inv_d, pipe_contiguous());
      return after_launch("ln_pipe");
    };
    const bool exact = (int64_t)threads * C == nchunk;
    int rc = -1;
#define PDB_LN_PIPE2(CC, MT)                                                                              \
  if (psm && exact) rc = launch(ln_pipe_kernel<T, P, CC, true, true, MT, true>, param_bytes);             \
  else if (psm) rc = launch(ln_pipe_kernel<T, P, CC, true, false, MT, true>, param_bytes);                \
  if (rc < 0) {                                                                                           \
    if (affine && exact) rc = launch(ln_pipe_kernel<T, P, CC, true, true, MT>, 0);                        \
    else if (affine) rc = launch(ln_pipe_kernel<T, P, CC, true, false, MT>, 0);                           \
    else rc = launch(ln_pipe_kernel<T, P, CC, false, false, MT>, 0);                                      \
  }
#define PDB_LN_PIPE(CC) if (threads <= 256) { PDB_LN_PIPE2(CC, 256) } else { PDB_LN_PIPE2(CC, 512) }
    if constexpr (!std::is_same<T, double>::value) {
      if (C == 4) { PDB_LN_PIPE(4) } else if (C == 2) { PDB_LN_PIPE(2) } else { PDB_LN_PIPE(1) }
    }
#undef PDB_LN_PIPE
#undef PDB_LN_PIPE2
    if (rc >= 0) return rc;  // rc < 0: the pipeline does not fit this device; fall through
  }
  if (nchunk <= 32 * 8) {
    int G = 1;
    while (G < 32 && G < nchunk) G <<= 1;
    int C = 1;
    while ((int64_t)C * G < nchunk) C <<= 1;
    const int rows_per_cta = 256 / G;
    const int64_t blocks = (rows + rows_per_cta - 1) / rows_per_cta;
    switch (C) {
      PDB_LN_C(ln_rows_reg_kernel, 1, <<<(unsigned)blocks, 256, 0, st>>>(x, scale, bias, y, m, v, rows, (int)D, G, eps, inv_d))
      PDB_LN_C(ln_rows_reg_kernel, 2, <<<(unsigned)blocks, 256, 0, st>>>(x, scale, bias, y, m, v, rows, (int)D, G, eps, inv_d))
      PDB_LN_C(ln_rows_reg_kernel, 4, <<<(unsigned)blocks, 256, 0, st>>>(x, scale, bias, y, m, v, rows, (int)D, G, eps, inv_d))
      PDB_LN_C(ln_rows_reg_kernel, 8, <<<(unsigned)blocks, 256, 0, st>>>(x, scale, bias, y, m, v, rows, (int)D, G, eps, inv_d))
    }
    return after_launch("ln_rows_reg");
  }
  if (nchunk <= 1024 * 4) {
    // up to 4 chunks per thread (more would push a 1024-thread CTA past 64 registers a thread);
    // widen the CTA first, deepen the chunks last
    int threads = 128;
    while (threads < 1024 && (int64_t)threads * 4 < nchunk) threads <<= 1;
    int C = 1;
    while ((int64_t)C * threads < nchunk) C <<= 1;
    switch (C) {
      PDB_LN_C(ln_block_reg_kernel, 1, <<<(unsigned)rows, threads, 0, st>>>(x, scale, bias, y, m, v, (int)D, eps, inv_d))
      PDB_LN_C(ln_block_reg_kernel, 2, <<<(unsigned)rows, threads, 0, st>>>(x, scale, bias, y, m, v, (int)D, eps, inv_d))
      PDB_LN_C(ln_block_reg_kernel, 4, <<<(unsigned)rows, threads, 0, st>>>(x, scale, bias, y, m, v, (int)D, eps, inv_d))
    }
    return after_launch("ln_block_reg");
  }
#undef PDB_LN_C
  // Long rows (beyond the register routes): thread-block cluster per row, slices in shared memory.
  if (vec && nchunk > 1024 * 4 && (!scale || is_aligned16(scale)) && (!bias || is_aligned16(bias))) {
    const size_t row_bytes = (size_t)D * sizeof(T);
    const size_t kMaxSlice = 200 * 1024;
    int cs = 1;
    size_t slice_target = 32 * 1024;
    if (knobs().cluster_slice_kb >= 4) slice_target = (size_t)knobs().cluster_slice_kb * 1024;  // tuning knob
    while (cs < 8 && (row_bytes / cs > slice_target || rows * cs < 2 * (int64_t)sm_count())) cs <<= 1;
    int64_t slice = (D + cs - 1) / cs;
    slice = (slice + VN - 1) / VN * VN;
    auto kern = ln_rows_cluster_kernel<T, P>;
    if ((size_t)slice * sizeof(T) <= kMaxSlice && rows * cs <= 0x7fffffffLL &&
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxSlice) == cudaSuccess) {
      int threads = 256;
      while (threads < 1024 && (int64_t)threads * VN * 16 < slice) threads <<= 1;
      if (knobs().cluster_threads >= 64) threads = knobs().cluster_threads;  // tuning knob
      cudaLaunchConfig_t cfg{};
      cfg.gridDim = dim3((unsigned)(rows * cs));
      cfg.blockDim = dim3(threads);
      cfg.dynamicSmemBytes = (size_t)slice * sizeof(T);
      cfg.stream = st;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = cs;
      attr[0].val.clusterDim.y = 1;
      attr[0].val.clusterDim.z = 1;
      cfg.attrs = attr;
      cfg.numAttrs = 1;
      if (cudaLaunchKernelEx(&cfg, kern, x, scale, bias, y, m, v, D, slice, eps, inv_d) == cudaSuccess)
        return after_launch("ln_rows_cluster");
    }
    cudaGetLastError();  // could not be scheduled as a cluster: fall through to the streaming route
  }
  ln_block_stream_kernel<T, P><<<(unsigned)rows, 1024, 0, st>>>(x, scale, bias, y, m, v, D, eps, inv_d);
  return after_launch("ln_block_stream");
}

}  // namespace pdb200

extern "C" int pdb200_layer_norm(const void* x, const int64_t* dims, int rank, const void* scale,
                                 const void* bias, int param_dtype, float epsilon, int begin_norm_axis,
                                 int dtype, void* y, void* mean, void* variance, void* stream) {
  using namespace pdb200;
  PDB_CHECK_ARG(rank >= 1 && rank <= PDB200_MAX_RANK && dims != nullptr, "layer_norm: rank must be in [1, 9]");
  // LayerNormInferMeta (infermeta/ternary.cc:1082): 0 < begin_norm_axis < rank (0 is tolerated by the kernels)
  PDB_CHECK_ARG(begin_norm_axis >= 0 && begin_norm_axis <= rank,
                "'begin_norm_axis' must be less than the dimensions of X, but received 'begin_norm_axis' is " +
                    std::to_string(begin_norm_axis));
  int64_t rows = 1, D = 1;
  for (int i = 0; i < rank; ++i) {
    PDB_CHECK_ARG(dims[i] >= 0, "layer_norm: negative dimension");
    (i < begin_norm_axis ? rows : D) *= dims[i];
  }
  if (rows * D == 0) return PDB200_OK;
  PDB_CHECK_ARG(x && y, "layer_norm: NULL tensor pointer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const bool has_param = scale != nullptr || bias != nullptr;
  // gpu/layer_norm_kernel.cu:516-538: scale/bias dtype == x dtype, or == float (U)
  const bool same = !has_param || param_dtype == dtype;
  if (has_param && !same && param_dtype != PDB200_FLOAT32)
    return fail(PDB200_INVALID_ARGUMENT,
                "Unsupported data type of Scale and Bias: the dtype must be the same as X's or float32");
#define PDB_LN_CASE(DT)                                                                               \
  case DT: {                                                                                          \
    using T = typename CType<DT>::type;                                                               \
    if (same)                                                                                         \
      return layer_norm_typed<T, T>(static_cast<const T*>(x), static_cast<const T*>(scale),           \
                                    static_cast<const T*>(bias), static_cast<T*>(y), mean, variance,  \
                                    rows, D, epsilon, st);                                            \
    return layer_norm_typed<T, float>(static_cast<const T*>(x), static_cast<const float*>(scale),     \
                                      static_cast<const float*>(bias), static_cast<T*>(y), mean,      \
                                      variance, rows, D, epsilon, st);                                \
  }
  switch (dtype) {
    PDB_LN_CASE(PDB200_FLOAT32)
    PDB_LN_CASE(PDB200_BFLOAT16)
    PDB_LN_CASE(PDB200_FLOAT16)
    case PDB200_FLOAT64: {
      if (!same) return fail(PDB200_INVALID_ARGUMENT, "layer_norm: float64 input needs float64 scale/bias");
      return layer_norm_typed<double, double>(static_cast<const double*>(x), static_cast<const double*>(scale),
                                              static_cast<const double*>(bias), static_cast<double*>(y), mean,
                                              variance, rows, D, epsilon, st);
    }
    default:
      return fail(PDB200_UNSUPPORTED, "layer_norm: dtype must be float32/float64/float16/bfloat16, got code " +
                                          std::to_string(dtype));
  }
#undef PDB_LN_CASE
}
