// softmax.cu — softmax forward along one axis, hand-written for sm_100a.
//
// Replaces (fresh design): SoftmaxForwardCUDAKernelDriver and its four routes
//   WarpSoftmaxForward (dim <= 512), cudnnSoftmaxForward (512 < dim < 100000),
//   KeMatrixSoftmaxForward (dim >= 100000), NormalSoftmaxForward (axis != last)
//   paddle/phi/kernels/gpudnn/softmax_gpudnn.h:442-611,870-912,1025-1069,1239-1329.
// Numerics follow the CPU kernel the parity oracle is built from
//   paddle/phi/kernels/funcs/softmax_impl.h:36-42,234-273:
//   e = exp(max(x - rowmax, -64)); y = e * (1 / sum(e)); fp16/bf16 compute in float and
//   round once at the end (softmax_gpudnn.h:257-269).
//
// Routes (x viewed as [outer][dim][inner]):
//   inner == 1, row fits registers  -> rows_reg   : G lanes per row hold the whole row
//        (C x 16-byte chunks per lane): ONE read + ONE write of HBM, max/sum by shuffles.
//        Covers dim <= 32*16*V (2048 fp32, 4096 bf16) — the reference needs cuDNN there.
//   inner == 1, row fits shared mem -> rows_smem  : one CTA per row, row staged once in
//        shared memory as float (<= 50 K elements): still one read + one write.
//   inner == 1, long rows           -> rows_cluster: one thread-block CLUSTER (2..8 CTAs) per row;
//        each CTA stages its slice of the row in shared memory, the per-CTA max and sum are
//        combined through distributed shared memory (DSMEM) after a cluster barrier: one read +
//        one write of HBM for rows up to 8 x 200 KB (vocabulary-sized rows), and 8x the CTAs when
//        there are only a few rows.
//   inner == 1, longer still        -> rows_stream: three streaming sweeps (as the reference
//        does from 100000 up).
//   inner  > 1                      -> cols       : threads along `inner` (coalesced),
//        sweeps over `dim`; later sweeps hit L2.
//   inner == 1, many rows (config 4) -> pipe      : persistent CTAs fed by the TMA row pipeline of
//        rowpipe.cuh (one cp.async.bulk per row into a shared-memory ring), 16 values per thread,
//        exp through ex2.approx on the pre-scaled difference.
#include <cooperative_groups.h>

#include <map>
#include <tuple>

#include "common.cuh"
#include "rowpipe.cuh"
#include "shape.h"

namespace pdb200 {

template <typename A> __device__ __forceinline__ A exp_acc(A v);
template <> __device__ __forceinline__ float exp_acc<float>(float v) { return expf(v); }
template <> __device__ __forceinline__ double exp_acc<double>(double v) { return exp(v); }

template <typename A>
__device__ __forceinline__ A clipped_exp(A x, A m) {
  A d = x - m;
  d = d < A(-64) ? A(-64) : d;  // ValueClip, softmax_impl.h:36-42
  return exp_acc<A>(d);
}

template <typename A> __device__ __forceinline__ A log_acc(A v);
template <> __device__ __forceinline__ float log_acc<float>(float v) { return logf(v); }
template <> __device__ __forceinline__ double log_acc<double>(double v) { return log(v); }
// max(x - m, -64): the shifted, clipped logit both softmax and log_softmax are built from
template <typename A>
__device__ __forceinline__ A clipped_diff(A x, A m) {
  const A d = x - m;
  return d < A(-64) ? A(-64) : d;
}
// LOG = false: softmax,      y = e * (1 / sum)          with e = exp(clipped_diff)
// LOG = true : log_softmax,  y = d - log(sum)           with d = clipped_diff
//   (phi::LogSoftmaxFunctor, paddle/phi/kernels/cpu/log_softmax_kernel.cc:40-100; GPU: gpu/log_softmax_kernel.cu)
// Every kernel can also hand out the two row statistics, stats[2*pos] = max and stats[2*pos + 1] = log(sum), in the
// accumulator type (pos = row, or outer*inner + col for a non-last axis): cross_entropy_with_softmax builds its loss from
// them and the logits, in float, instead of taking the log of an already rounded probability.
template <bool LOG, typename A>
__device__ __forceinline__ A row_constant(A sum) { return LOG ? log_acc<A>(sum) : A(1) / sum; }
template <bool LOG, typename A>
__device__ __forceinline__ A finish(A e_or_d, A k) { return LOG ? e_or_d - k : e_or_d * k; }

template <typename A> struct MaxF { __device__ __forceinline__ A operator()(A a, A b) const { return b > a ? b : a; } };
template <typename A> struct AddF { __device__ __forceinline__ A operator()(A a, A b) const { return a + b; } };

template <typename A>
__device__ __forceinline__ A neg_inf() { return -std::numeric_limits<A>::infinity(); }

// ---- rows_reg --------------------------------------------------------------------------
// G lanes (power of two <= 32) per row, C chunks of V elements per lane; chunk c of lane l
// covers elements (c*G + l)*V .. +V.  Out-of-row chunks hold -inf and are never stored.
template <typename T, int C, bool VECTOR, bool LOG>
__global__ void __launch_bounds__(256) softmax_rows_reg_kernel(const T* __restrict__ x,
                                                               T* __restrict__ y, int64_t outer,
                                                               int dim, int G, typename AccType<T>::type* __restrict__ stats) {
  using A = typename AccType<T>::type;
  constexpr int V = VECTOR ? Vec<T>::N : 1;
  const int rows_per_cta = blockDim.x / G;
  const int g = threadIdx.x / G, lane = threadIdx.x - g * G;
  const int64_t row = (int64_t)blockIdx.x * rows_per_cta + g;
  const bool valid = row < outer;  // shuffles below are full-mask: no early exit
  const T* src = x + row * dim;
  T* dst = y + row * dim;
  A e[C * V];
  A m = neg_inf<A>();
#pragma unroll
  for (int c = 0; c < C; ++c) {
    const int off = (c * G + lane) * V;
    if (valid && off < dim) {
      if constexpr (VECTOR) {
        Vec<T> v = ld_stream(src + off);  // y may alias x (softmax_): each chunk is read and written by the same thread
#pragma unroll
        for (int k = 0; k < V; ++k) e[c * V + k] = cvt<A>(v.v[k]);
      } else {
        e[c] = cvt<A>(src[off]);
      }
    } else {
#pragma unroll
      for (int k = 0; k < V; ++k) e[c * V + k] = neg_inf<A>();
    }
  }
#pragma unroll
  for (int i = 0; i < C * V; ++i) m = e[i] > m ? e[i] : m;
  for (int s = G >> 1; s >= 1; s >>= 1) {
    A o = shfl_xor_any(m, s);
    m = o > m ? o : m;
  }
  A sum = A(0);
#pragma unroll
  for (int c = 0; c < C; ++c) {
    const int off = (c * G + lane) * V;
    if (off < dim) {
#pragma unroll
      for (int k = 0; k < V; ++k) {
        if constexpr (LOG) {
          e[c * V + k] = clipped_diff<A>(e[c * V + k], m);
          sum += exp_acc<A>(e[c * V + k]);
        } else {
          e[c * V + k] = clipped_exp<A>(e[c * V + k], m);
          sum += e[c * V + k];
        }
      }
    }
  }
  for (int s = G >> 1; s >= 1; s >>= 1) sum += shfl_xor_any(sum, s);
  const A inv = row_constant<LOG, A>(sum);
  if (stats != nullptr && lane == 0 && valid) { stats[2 * row] = m; stats[2 * row + 1] = log_acc<A>(sum); }
#pragma unroll
  for (int c = 0; c < C; ++c) {
    const int off = (c * G + lane) * V;
    if (valid && off < dim) {
      if constexpr (VECTOR) {
        Vec<T> o;
#pragma unroll
        for (int k = 0; k < V; ++k) o.v[k] = cvt<T>(finish<LOG, A>(e[c * V + k], inv));
        st_stream(dst + off, o);
      } else {
        dst[off] = cvt<T>(finish<LOG, A>(e[c], inv));
      }
    }
  }
}

// exp(max(x - m, -64)) as 2^(x*log2e - m*log2e) with the clamp applied in the log2 domain.
// One FFMA + FMNMX + MUFU.EX2.  The rounding of m*log2e is a common factor of the whole row and
// cancels in the normalisation; per element the error is 0.5 ulp of the exponent argument plus
// ex2.approx's 2^-22: <= ~2e-6 relative for |x - m| <= 64, inside the 1e-5 bar.
// max(a, b) that returns NaN when either operand is NaN (fmaxf would drop it): one FMNMX.NAN.  A NaN logit, a +inf
// logit (inf - inf) and a row of -inf all reach the clamp as NaN and must come out as NaN, as they do from the CPU
// kernel's `x < -64 ? -64 : x` (softmax_impl.h:36-42) — never as a healthy-looking probability.
__device__ __forceinline__ float max_nan(float a, float b) {
  float r;
  asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ float fast_clipped_exp(float x, float neg_m_log2e) {
  constexpr float kLog2e = 1.4426950408889634f;
  float t = fmaf(x, kLog2e, neg_m_log2e);
  t = max_nan(t, -64.0f * kLog2e);
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(t));
  return r;
}

// float rows take the fast path, double rows the accurate one
template <typename A> __device__ __forceinline__ A row_exp(A x, A m, A neg_m_log2e);
template <> __device__ __forceinline__ float row_exp<float>(float x, float, float neg_m_log2e) {
  return fast_clipped_exp(x, neg_m_log2e);
}
template <> __device__ __forceinline__ double row_exp<double>(double x, double m, double) {
  return clipped_exp<double>(x, m);
}

// ---- pipe: persistent CTAs, rows staged in shared memory by cp.async.bulk -------------------
// blockDim.x threads per row, C chunks of V elements per thread.  T is float, bf16 or fp16.
template <typename T, int C, bool EXACT, bool LOG>
__global__ void __launch_bounds__(512) softmax_pipe_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                           int64_t rows, int dim, int stages, int contiguous,
                                                           float* __restrict__ stats) {
  constexpr int V = Vec<T>::N;
  extern __shared__ __align__(128) unsigned char ring_raw[];
  T* ring = reinterpret_cast<T*>(ring_raw);
  __shared__ uint64_t full[kPipeMaxStages];
  __shared__ float red[2][32];
  const int tid = threadIdx.x, nt = blockDim.x;
  const uint32_t row_bytes = (uint32_t)dim * sizeof(T);
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
  block_reduce_1sync_init(red[0], -INFINITY);
  block_reduce_1sync_init(red[1], 0.f);
  __syncthreads();
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) {
      const int64_t r = row_first + s * row_step;
      if (r < row_end) {
        mbar_expect_tx(&full[s], row_bytes);
        bulk_load_row(ring + (size_t)s * dim, x + r * dim, row_bytes, &full[s]);
      }
    }
  }
  int s = 0;
  uint32_t phase = 0;
  for (int64_t row = row_first; row < row_end; row += row_step) {
    mbar_wait(&full[s], phase);
    const T* st = ring + (size_t)s * dim;
    Vec<T> raw[C];
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + tid) * V;
      if (EXACT || off < dim) raw[c] = lds_vec(st + off);
    }
    float e[C * V];
    float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + tid) * V;
      if (EXACT || off < dim) {
        unpack_to_float<T>(raw[c], &e[c * V]);
      } else {
#pragma unroll
        for (int k = 0; k < V; ++k) e[c * V + k] = -INFINITY;
      }
#pragma unroll
      for (int k = 0; k < V; ++k) mx[k & 3] = fmaxf(mx[k & 3], e[c * V + k]);
    }
    float m = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
    m = block_reduce_1sync(m, MaxF<float>(), red[0]);  // its barrier also means: stage s fully read
    if (tid == 0) {
      const int64_t nrow = row + (int64_t)stages * row_step;
      if (nrow < row_end) {
        fence_proxy_async_smem();
        mbar_expect_tx(&full[s], row_bytes);
        bulk_load_row(ring + (size_t)s * dim, x + nrow * dim, row_bytes, &full[s]);
      }
    }
    const float nml = -m * 1.4426950408889634f;
    float sm[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + tid) * V;
#pragma unroll
      for (int k = 0; k < V; ++k) {
        float p;
        if constexpr (LOG) {
          const float d = max_nan(e[c * V + k] - m, -64.0f);  // the exact clipped difference is the output term
          e[c * V + k] = d;
          asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p) : "f"(d * 1.4426950408889634f));
        } else {
          p = e[c * V + k] = fast_clipped_exp(e[c * V + k], nml);
        }
        if (EXACT || off < dim) sm[k & 3] += p;  // padding chunks are not part of the row
      }
    }
    float sum = (sm[0] + sm[1]) + (sm[2] + sm[3]);
    sum = block_reduce_1sync(sum, AddF<float>(), red[1]);
    const float inv = row_constant<LOG, float>(sum);
    if (stats != nullptr && tid == 0) { stats[2 * row] = m; stats[2 * row + 1] = logf(sum); }
    T* dst = y + row * dim;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + tid) * V;
      if (EXACT || off < dim) {
        float r[V];
#pragma unroll
        for (int k = 0; k < V; ++k) r[k] = finish<LOG, float>(e[c * V + k], inv);
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

// ---- rows_smem -------------------------------------------------------------------------
template <typename T, bool VECTOR, bool LOG>
__global__ void __launch_bounds__(1024) softmax_rows_smem_kernel(const T* __restrict__ x,
                                                                 T* __restrict__ y, int64_t dim,
                                                                 typename AccType<T>::type* __restrict__ stats) {
  using A = typename AccType<T>::type;
  constexpr int V = VECTOR ? Vec<T>::N : 1;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  A* row = reinterpret_cast<A*>(smem_raw);  // [dim]
  __shared__ A red[33];
  const T* src = x + (int64_t)blockIdx.x * dim;
  T* dst = y + (int64_t)blockIdx.x * dim;
  A m = neg_inf<A>();
  for (int64_t i = (int64_t)threadIdx.x * V; i < dim; i += (int64_t)blockDim.x * V) {
    if constexpr (VECTOR) {
      Vec<T> v = ld_stream(src + i);
#pragma unroll
      for (int k = 0; k < V; ++k) {
        A a = cvt<A>(v.v[k]);
        row[i + k] = a;
        m = a > m ? a : m;
      }
    } else {
      A a = cvt<A>(src[i]);
      row[i] = a;
      m = a > m ? a : m;
    }
  }
  m = block_reduce(m, MaxF<A>(), neg_inf<A>(), red);  // also orders the smem writes
  A sum = A(0);
  for (int64_t i = threadIdx.x; i < dim; i += blockDim.x) {  // each thread revisits what any thread wrote: synced above
    if constexpr (LOG) {
      const A d = clipped_diff<A>(row[i], m);
      row[i] = d;
      sum += exp_acc<A>(d);
    } else {
      A ev = clipped_exp<A>(row[i], m);
      row[i] = ev;
      sum += ev;
    }
  }
  sum = block_reduce(sum, AddF<A>(), A(0), red);
  const A inv = row_constant<LOG, A>(sum);
  if (stats != nullptr && threadIdx.x == 0) { stats[2 * (int64_t)blockIdx.x] = m; stats[2 * (int64_t)blockIdx.x + 1] = log_acc<A>(sum); }
  for (int64_t i = (int64_t)threadIdx.x * V; i < dim; i += (int64_t)blockDim.x * V) {
    if constexpr (VECTOR) {
      Vec<T> o;
#pragma unroll
      for (int k = 0; k < V; ++k) o.v[k] = cvt<T>(finish<LOG, A>(row[i + k], inv));
      st_stream(dst + i, o);
    } else {
      dst[i] = cvt<T>(finish<LOG, A>(row[i], inv));
    }
  }
}

// Online (max, sum) bookkeeping shared by the long-row kernels.
struct MaxSum {
  float m, s, a;  // running max; sum of exp relative to it; clipped exp(-64) terms seen while the max was still -inf
};
__device__ __forceinline__ MaxSum ms_combine(MaxSum x, MaxSum y) {
  constexpr float kLog2e = 1.4426950408889634f;
  const float M = fmaxf(x.m, y.m);
  float fx = 0.f, fy = 0.f;  // a side whose max is -inf has s == 0: factor 0 avoids inf - inf
  if (x.m != -INFINITY) asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(fx) : "f"((x.m - M) * kLog2e));
  if (y.m != -INFINITY) asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(fy) : "f"((y.m - M) * kLog2e));
  return MaxSum{M, fmaf(x.s, fx, y.s * fy), x.a + y.a};
}
__device__ __forceinline__ MaxSum ms_warp_reduce(MaxSum v) {
#pragma unroll
  for (int k = 16; k >= 1; k >>= 1)
    v = ms_combine(v, MaxSum{__shfl_xor_sync(0xffffffffu, v.m, k), __shfl_xor_sync(0xffffffffu, v.s, k),
                             __shfl_xor_sync(0xffffffffu, v.a, k)});
  return v;
}
// fold V new values into a running (max, sum)
template <int N>
__device__ __forceinline__ void ms_accumulate(MaxSum& t, const float* v) {
  constexpr float kLog2e = 1.4426950408889634f;
  constexpr float kClip = 1.603810890548638e-28f;  // exp(-64)
  float nm = t.m;
#pragma unroll
  for (int i = 0; i < N; ++i) nm = fmaxf(nm, v[i]);
  if (nm == -INFINITY) {  // nothing finite yet: -inf clips to exp(-64) whatever the max turns out to be; NaN propagates
#pragma unroll
    for (int i = 0; i < N; ++i) t.a += (v[i] != v[i]) ? v[i] : kClip;
  } else {
    float resc;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(resc) : "f"((t.m - nm) * kLog2e));  // 0 when the old max was -inf (t.s is 0 then)
    float add[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < N; ++i) add[i & 3] += fast_clipped_exp(v[i], -nm * kLog2e);
    t.s = fmaf(t.s, resc, (add[0] + add[1]) + (add[2] + add[3]));
    t.m = nm;
  }
}

// ---- rows_cluster: one thread-block cluster per row, DSMEM combine ------------------------------
// CTA `rank` of the cluster owns elements [rank*slice, rank*slice + slice) of the row and keeps them in
// shared memory in their storage type.  The two row-wide quantities (max, sum of exp) are produced
// per CTA, published in that CTA's shared memory and read by every peer through DSMEM
// (cluster.map_shared_rank) after a cluster barrier; all CTAs fold the partials in rank order, so
// they agree bit for bit.  Requires dim % V == 0 and 16-byte aligned bases (vector path only).
template <typename T, bool LOG>
__global__ void __launch_bounds__(1024) softmax_rows_cluster_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                                    int64_t dim, int64_t slice,
                                                                    typename AccType<T>::type* __restrict__ stats) {
  namespace cg = cooperative_groups;
  using A = typename AccType<T>::type;
  constexpr int V = Vec<T>::N;
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned rank = cluster.block_rank(), cs = cluster.num_blocks();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* srow = reinterpret_cast<T*>(smem_raw);  // [slice]
  __shared__ A red[33];
  __shared__ A xchg[2];  // [0] this CTA's max, [1] this CTA's sum — read by peers through DSMEM
  const int64_t row = (int64_t)blockIdx.x / cs;
  const int64_t begin = (int64_t)rank * slice;
  int64_t n = dim - begin;
  n = n < 0 ? 0 : (n > slice ? slice : n);
  const T* src = x + row * dim + begin;
  T* dst = y + row * dim + begin;
  const int64_t step = (int64_t)blockDim.x * V;
  // sweep 1: HBM -> shared with ONE bulk async copy for the whole slice (the only read of x: all of
  // its bytes are in flight at once, no registers involved), then the local max from shared memory
  __shared__ uint64_t bar;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    mbar_fence_init();
  }
  __syncthreads();
  if (n > 0) {
    if (threadIdx.x == 0) {
      mbar_expect_tx(&bar, (uint32_t)(n * sizeof(T)));
      bulk_load_row(srow, src, (uint32_t)(n * sizeof(T)), &bar);
    }
    mbar_wait(&bar, 0);
  }
  // sweep 1 (shared memory): running (max, sum of exp) per thread — one pass, one rescale per vector — folded over the
  // CTA and, after ONE cluster barrier, over the cluster through DSMEM.  (The first version made a max sweep, a cluster
  // barrier, a sum sweep and a second cluster barrier.)
  A gm, gs;
  if constexpr (std::is_same<A, float>::value) {
    MaxSum t{-INFINITY, 0.f, 0.f};
    for (int64_t i = (int64_t)threadIdx.x * V; i < n; i += step) {
      float v[V];
      unpack_to_float<T>(lds_vec(srow + i), v);
      ms_accumulate<V>(t, v);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    __shared__ float red3[32][3];
    __shared__ float xchg3[3];
    t = ms_warp_reduce(t);
    if (lane == 0) { red3[warp][0] = t.m; red3[warp][1] = t.s; red3[warp][2] = t.a; }
    __syncthreads();
    MaxSum g{-INFINITY, 0.f, 0.f};
    if (lane < nwarps) g = MaxSum{red3[lane][0], red3[lane][1], red3[lane][2]};
    g = ms_warp_reduce(g);
    if (threadIdx.x == 0) { xchg3[0] = g.m; xchg3[1] = g.s; xchg3[2] = g.a; }
    cluster.sync();
    MaxSum p{-INFINITY, 0.f, 0.f};
    if (lane < (int)cs) {
      const float* r = cluster.map_shared_rank(xchg3, lane);
      p = MaxSum{r[0], r[1], r[2]};
    }
    g = ms_warp_reduce(p);
    gm = g.m;
    gs = g.s + g.a;
    if (g.m == -INFINITY) gs = __int_as_float(0x7fc00000);  // a row of -inf: NaN, as inf - inf gives in the CPU kernel
  } else {
    // double rows: exact exp, two sweeps and two barriers
    A m = neg_inf<A>();
    for (int64_t i = (int64_t)threadIdx.x * V; i < n; i += step) {
      const Vec<T> v = lds_vec(srow + i);
#pragma unroll
      for (int k = 0; k < V; ++k) {
        const A a = cvt<A>(v.v[k]);
        m = a > m ? a : m;
      }
    }
    m = block_reduce(m, MaxF<A>(), neg_inf<A>(), red);
    if (threadIdx.x == 0) xchg[0] = m;
    cluster.sync();
    gm = neg_inf<A>();
    for (unsigned r = 0; r < cs; ++r) {
      const A o = *cluster.map_shared_rank(&xchg[0], r);
      gm = o > gm ? o : gm;
    }
    A sum = A(0);
    for (int64_t i = (int64_t)threadIdx.x * V; i < n; i += step) {
      const Vec<T> v = lds_vec(srow + i);
#pragma unroll
      for (int k = 0; k < V; ++k) sum += row_exp<A>(cvt<A>(v.v[k]), gm, A(0));
    }
    sum = block_reduce(sum, AddF<A>(), A(0), red);
    if (threadIdx.x == 0) xchg[1] = sum;
    cluster.sync();
    gs = A(0);
    for (unsigned r = 0; r < cs; ++r) gs += *cluster.map_shared_rank(&xchg[1], r);
  }
  const A nml = -gm * A(1.4426950408889634);
  const A inv = row_constant<LOG, A>(gs);
  if (stats != nullptr && rank == 0 && threadIdx.x == 0) { stats[2 * row] = gm; stats[2 * row + 1] = log_acc<A>(gs); }
  // Every CTA has read its peers' pairs by now: arrive for the exit barrier here (relaxed: nothing published after this
  // point needs ordering, and a release arrive AFTER the stores made every thread wait for its stores to drain before
  // the CTA could leave — stall_membar 1.0, profiles/r2e_softmax_vocab_bf16.txt), wait for it at the very end.
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  // sweep 3: normalise and store
  for (int64_t i = (int64_t)threadIdx.x * V; i < n; i += step) {
    if constexpr (std::is_same<A, float>::value) {
      float v[V];
      unpack_to_float<T>(lds_vec(srow + i), v);
#pragma unroll
      for (int k = 0; k < V; ++k) v[k] = LOG ? max_nan(v[k] - gm, -64.0f) - inv : fast_clipped_exp(v[k], nml) * inv;
      Vec<T> o;
      pack_from_float<T>(v, o);
      st_stream(dst + i, o);
    } else {
      const Vec<T> v = lds_vec(srow + i);
      Vec<T> o;
#pragma unroll
      for (int k = 0; k < V; ++k)
        o.v[k] = cvt<T>(LOG ? clipped_diff<A>(cvt<A>(v.v[k]), gm) - inv : row_exp<A>(cvt<A>(v.v[k]), gm, nml) * inv);
      st_stream(dst + i, o);
    }
  }
  asm volatile("barrier.cluster.wait.aligned;" ::: "memory");  // nobody leaves while a peer may still be reading its xchg[]
}

// ---- rows_cluster_lean: the shared-memory cluster kernel for float accumulators, with cheap sweeps ----------------------
// Same data movement as softmax_rows_cluster_kernel (one bulk copy of the slice into shared memory, one cluster barrier),
// different arithmetic.  That kernel ran 20 instructions and 2.25 ex2 per element on [4096,128256] bf16 (issue 63 % busy,
// XU 63 %: profiles/r2e_softmax_vocab_bf16.txt) — its 1024 threads owned only 32 elements each, so the per-row reductions
// of (max, sum, clipped) triples (three shuffles and two ex2 per combine) cost as much as the sweeps.  Here:
//   sweep A  the slice maximum m_c alone: packed 16-bit max, half an instruction per element;
//   sweep B  s_c = sum exp(x - m_c) relative to the CTA's OWN maximum (no exchange needed before it), fp32 pairs;
//   one cluster barrier: (m_c, s_c) of every CTA through DSMEM -> M, S = sum s_c exp(m_c - M);
//   sweep C  y = max(exp(x - M), exp(-64)) / S, the clip applied to the packed 16-bit result.
// 256 or 512 threads own 64+ elements each; plain float reductions.  S ignores the clip (a clipped term adds at most
// 1.6e-28 to a sum >= 1).
template <typename T> struct PackedMax;  // running maximum of the raw vectors
template <> struct PackedMax<float> {
  float m = -INFINITY;
  __device__ __forceinline__ void add(const Vec<float>& v) { m = fmaxf(fmaxf(m, fmaxf(v.v[0], v.v[1])), fmaxf(v.v[2], v.v[3])); }
  __device__ __forceinline__ float get() const { return m; }
};
template <> struct PackedMax<__nv_bfloat16> {
  __nv_bfloat162 m = __nv_bfloat162(__ushort_as_bfloat16(0xFF80), __ushort_as_bfloat16(0xFF80));
  __device__ __forceinline__ void add(const Vec<__nv_bfloat16>& v) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
    m = __hmax2(m, __hmax2(__hmax2(h[0], h[1]), __hmax2(h[2], h[3])));
  }
  __device__ __forceinline__ float get() const { return fmaxf(__bfloat162float(m.x), __bfloat162float(m.y)); }
};
template <> struct PackedMax<__half> {
  __half2 m = __half2(__ushort_as_half(0xFC00), __ushort_as_half(0xFC00));
  __device__ __forceinline__ void add(const Vec<__half>& v) {
    const __half2* h = reinterpret_cast<const __half2*>(&v);
    m = __hmax2(m, __hmax2(__hmax2(h[0], h[1]), __hmax2(h[2], h[3])));
  }
  __device__ __forceinline__ float get() const { return fmaxf(__half2float(m.x), __half2float(m.y)); }
};
// the packed result never below `floor2` (NaN kept): the -64 clip on the rounded output
__device__ __forceinline__ void clip_packed(Vec<__nv_bfloat16>& o, float floor_v) {
  const __nv_bfloat162 f = __float2bfloat162_rn(floor_v);
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __hmax2_nan(h[i], f);
}
__device__ __forceinline__ void clip_packed(Vec<__half>& o, float floor_v) {
  const __half2 f = __float2half2_rn(floor_v);
  __half2* h = reinterpret_cast<__half2*>(&o);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __hmax2_nan(h[i], f);
}
__device__ __forceinline__ void clip_packed(Vec<float>& o, float floor_v) {
#pragma unroll
  for (int i = 0; i < 4; ++i) o.v[i] = max_nan(o.v[i], floor_v);
}
// elementwise maximum of two raw vectors (NaN operands ignored, like fmaxf) and the same with NaN kept
__device__ __forceinline__ void vmax_raw(Vec<__nv_bfloat16>& a, const Vec<__nv_bfloat16>& b) {
  __nv_bfloat162* pa = reinterpret_cast<__nv_bfloat162*>(&a);
  const __nv_bfloat162* pb = reinterpret_cast<const __nv_bfloat162*>(&b);
#pragma unroll
  for (int i = 0; i < 4; ++i) pa[i] = __hmax2(pa[i], pb[i]);
}
__device__ __forceinline__ void vmax_raw(Vec<__half>& a, const Vec<__half>& b) {
  __half2* pa = reinterpret_cast<__half2*>(&a);
  const __half2* pb = reinterpret_cast<const __half2*>(&b);
#pragma unroll
  for (int i = 0; i < 4; ++i) pa[i] = __hmax2(pa[i], pb[i]);
}
__device__ __forceinline__ void vmax_raw(Vec<float>& a, const Vec<float>& b) {
#pragma unroll
  for (int i = 0; i < 4; ++i) a.v[i] = fmaxf(a.v[i], b.v[i]);
}
__device__ __forceinline__ void vclip_raw(Vec<__nv_bfloat16>& a, const Vec<__nv_bfloat16>& floor_v) {
  __nv_bfloat162* pa = reinterpret_cast<__nv_bfloat162*>(&a);
  const __nv_bfloat162* pb = reinterpret_cast<const __nv_bfloat162*>(&floor_v);
#pragma unroll
  for (int i = 0; i < 4; ++i) pa[i] = __hmax2_nan(pa[i], pb[i]);
}
__device__ __forceinline__ void vclip_raw(Vec<__half>& a, const Vec<__half>& floor_v) {
  __half2* pa = reinterpret_cast<__half2*>(&a);
  const __half2* pb = reinterpret_cast<const __half2*>(&floor_v);
#pragma unroll
  for (int i = 0; i < 4; ++i) pa[i] = __hmax2_nan(pa[i], pb[i]);
}
__device__ __forceinline__ void vclip_raw(Vec<float>& a, const Vec<float>& floor_v) {
#pragma unroll
  for (int i = 0; i < 4; ++i) a.v[i] = max_nan(a.v[i], floor_v.v[i]);
}
__device__ __forceinline__ float ex2_fast(float t) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(t));
  return r;
}

template <typename T, bool LOG>
__global__ void __launch_bounds__(512) softmax_rows_cluster_lean_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t dim,
                                                                        int64_t slice, float* __restrict__ stats) {
  namespace cg = cooperative_groups;
  constexpr int V = Vec<T>::N;
  constexpr int P = V / 2;
  constexpr float kLog2e = 1.4426950408889634f;
  constexpr float kClip = 1.603810890548638e-28f;  // exp(-64)
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned rank = cluster.block_rank(), cs = cluster.num_blocks();
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* srow = reinterpret_cast<T*>(smem_raw);  // [slice]
  __shared__ float red[2][32];
  __shared__ float xchg[2];
  __shared__ uint64_t bar;
  const int tid = threadIdx.x;
  const int64_t row = (int64_t)blockIdx.x / cs;
  const int64_t begin = (int64_t)rank * slice;
  int64_t n64 = dim - begin;
  const int n = (int)(n64 < 0 ? 0 : (n64 > slice ? slice : n64));
  const T* src = x + row * dim + begin;
  T* dst = y + row * dim + begin;
  const int step = blockDim.x * V;
  if (tid == 0) {
    mbar_init(&bar, 1);
    mbar_fence_init();
  }
  if (tid < 32) { red[0][tid] = -INFINITY; red[1][tid] = 0.f; }
  __syncthreads();
  if (n > 0) {
    if (tid == 0) {
      mbar_expect_tx(&bar, (uint32_t)(n * sizeof(T)));
      bulk_load_row(srow, src, (uint32_t)(n * sizeof(T)), &bar);
    }
    mbar_wait(&bar, 0);
  }
  // sweep A
  PackedMax<T> pmx;
  int i = tid * V;
  for (; i + step < n; i += 2 * step) {
    const Vec<T> a = lds_vec(srow + i), b = lds_vec(srow + i + step);
    pmx.add(a);
    pmx.add(b);
  }
  if (i < n) pmx.add(lds_vec(srow + i));
  const float mc = block_reduce_1sync(pmx.get(), MaxF<float>(), red[0]);
  // sweep B
  const F2 l2e = f2_splat(kLog2e), nmc = f2_splat(mc == -INFINITY ? 0.f : -mc * kLog2e);  // a slice of -inf only: every term 0
  F2 acc[2] = {f2_splat(0.f), f2_splat(0.f)};
#pragma unroll 2
  for (i = tid * V; i < n; i += step) {
    float v[V];
    unpack_to_float<T>(lds_vec(srow + i), v);
#pragma unroll
    for (int j = 0; j < P; ++j) {
      const F2 t = f2_fma(f2_make(v[2 * j], v[2 * j + 1]), l2e, nmc);
      acc[j & 1] = f2_add(acc[j & 1], f2_make(ex2_fast(f2_lo(t)), ex2_fast(f2_hi(t))));
    }
  }
  const F2 a2 = f2_add(acc[0], acc[1]);
  const float sc = block_reduce_1sync(f2_lo(a2) + f2_hi(a2), AddF<float>(), red[1]);
  if (tid == 0) { xchg[0] = mc; xchg[1] = sc; }
  cluster_sync_published_by(tid == 0);
  float pm = -INFINITY, ps = 0.f;
  const int lane = tid & 31;
  if (lane < (int)cs) {
    const float* r = cluster.map_shared_rank(xchg, lane);
    pm = r[0];
    ps = r[1];
  }
  const float M = __shfl_sync(0xffffffffu, group_reduce<8>(pm, MaxF<float>()), 0);  // cs <= 8: lanes 0..7 hold the pairs
  float f = 0.f;
  if (pm != -INFINITY) f = ex2_fast((pm - M) * kLog2e);
  float S = __shfl_sync(0xffffffffu, group_reduce<8>(ps * f, AddF<float>()), 0);
  if (M == -INFINITY) S = __int_as_float(0x7fc00000);  // a row of -inf: NaN, as inf - inf gives in the CPU kernel
  if (stats != nullptr && rank == 0 && tid == 0) { stats[2 * row] = M; stats[2 * row + 1] = logf(S); }
  // every CTA has read its peers' pairs: arrive for the exit barrier now (relaxed), wait for it after the stores
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  const float inv = row_constant<LOG, float>(S);
  // sweep C
  const F2 nM = f2_splat(LOG ? -M : -M * kLog2e), k2 = f2_splat(LOG ? -inv : inv);
#pragma unroll 2
  for (i = tid * V; i < n; i += step) {
    float v[V];
    unpack_to_float<T>(lds_vec(srow + i), v);
    Vec<T> o;
    if constexpr (LOG) {
#pragma unroll
      for (int j = 0; j < P; ++j) {
        const F2 d = f2_add(f2_make(v[2 * j], v[2 * j + 1]), nM);
        const F2 r = f2_add(f2_make(max_nan(f2_lo(d), -64.0f), max_nan(f2_hi(d), -64.0f)), k2);
        v[2 * j] = f2_lo(r);
        v[2 * j + 1] = f2_hi(r);
      }
      pack_from_float<T>(v, o);
    } else {
#pragma unroll
      for (int j = 0; j < P; ++j) {
        const F2 t = f2_fma(f2_make(v[2 * j], v[2 * j + 1]), l2e, nM);
        const F2 r = f2_mul(f2_make(ex2_fast(f2_lo(t)), ex2_fast(f2_hi(t))), k2);
        v[2 * j] = f2_lo(r);
        v[2 * j + 1] = f2_hi(r);
      }
      pack_from_float<T>(v, o);
      clip_packed(o, kClip * inv);
    }
    st_stream(dst + i, o);
  }
  asm volatile("barrier.cluster.wait.aligned;" ::: "memory");  // nobody leaves while a peer may still be reading its xchg[]
}

// ---- rows_relay: long rows as a stream of chunks through persistent CTAs, one read, one write, one exp ----------------
// The cluster kernels above stop near 4.5-4.9 TB/s on vocabulary-sized rows whatever their instruction count (the lean
// kernel halves the instructions of the first one and gains 7 %): a CTA loads its slice, waits, computes, waits for
// the cluster, stores and leaves — the bytes in flight per SM follow that duty cycle.  This kernel keeps the copies
// running: a row is cut into chunks of THREADS x EPT elements, chunk g of the whole tensor goes to CTA g mod G of a
// persistent co-resident grid (several small CTAs per SM: independent chains whose phases overlap), and every CTA
// prefetches its next chunks into a ring of shared-memory stages with bulk copies.  Per chunk:
//   P1  stage -> registers (EPT floats per thread); every WARP works relative to its own maximum m_w: e = exp(x - m_w)
//       ONCE per element, kept in registers, s_w = sum e; one CTA barrier (which also frees the stage for its refill)
//       and every warp folds the warps' pairs into the chunk's (m_c, s_c), published as one 8-byte word in a global
//       record array (zeroed by the launcher; a published word is never all-zero);
//   P2  one iteration LATER (after P1 of the CTA's next chunk, so the CTAs working on the same row have had a whole
//       chunk's time to publish; the record loads are issued in the middle of that P1): M = max m_c,
//       S = sum s_c exp(m_c - M), and the chunk is finished from registers: y = max(e * exp(m_w - M), exp(-64)) / S.
// No cluster, no barrier between CTAs; the only cross-CTA wait is the poll, bounded by a timer (trap on expiry).  It cannot
// deadlock: P2 of chunk h needs P1 of chunks < h + chunks_per_row, whose CTAs need P2 of chunks 2 G earlier — strictly
// decreasing, and the cooperative launch guarantees every CTA is resident.
constexpr int kRelayMaxStages = 8;

__device__ __forceinline__ unsigned long long relay_peek(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long relay_poll(const unsigned long long* p, unsigned long long timeout_ns) {
  unsigned long long v = relay_peek(p);
  if (v != 0ull) return v;
  unsigned long long t0, t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
  for (;;) {
#pragma unroll 1
    for (int spin = 0; spin < 64; ++spin) {
      v = relay_peek(p);
      if (v != 0ull) return v;
    }
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    if (t - t0 > timeout_ns) __trap();  // a record that never arrives: fail loudly rather than hang
  }
}

template <typename T>
struct RelayCtx {
  const T* x;
  T* y;
  int64_t dim, total;
  int cpr, stages;
  unsigned long long* rec;
  float* stats;
  unsigned long long timeout_ns;
  T* ring;
  uint64_t* full;
  float2 (*wp)[32];  // [2][32]: the warps' (m_w, s_w), double-buffered by iteration parity
};

// Position of a chunk: g = row * cpr + ci.  Stepped by the grid size without a division (64-bit divisions cost ~100
// instructions each, three per chunk in the first version of this kernel: half of its instruction count).
struct RelayPos {
  int64_t g, row;
  int ci;
  __device__ __forceinline__ void step(int64_t dg, int64_t drow, int dci, int cpr) {
    g += dg;
    row += drow;
    ci += dci;
    if (ci >= cpr) {
      ci -= cpr;
      ++row;
    }
  }
};
// fold (m, s) pairs held one per lane (lanes without a pair: m = -inf, s = 0) over the warp
__device__ __forceinline__ void relay_fold(float m, float s, float& M, float& S) {
  constexpr float kLog2e = 1.4426950408889634f;
  M = group_reduce<32>(m, MaxF<float>());
  S = group_reduce<32>(m == -INFINITY ? 0.f : s * ex2_fast((m - M) * kLog2e), AddF<float>());
}

// `pend_rec` / `pre`: the records of the row whose P2 follows this P1 (or nullptr), and where the first 32 are prefetched to.
// FULL: the chunk fills the stage (every chunk of a row but its last): no bounds checks.
template <typename T, bool LOG, int THREADS, int EPT, bool FULL>
__device__ __forceinline__ void relay_p1(const RelayCtx<T>& c, const RelayPos& pos, const RelayPos& nxt, int n, int s, uint32_t phase,
                                         int par, float (&e)[EPT], float& mw_out, const unsigned long long* pend_rec,
                                         unsigned long long& pre) {
  constexpr int V = Vec<T>::N;
  constexpr int K = EPT / V;
  constexpr int NW = THREADS / 32;
  constexpr float kLog2e = 1.4426950408889634f;
  constexpr int kCap = THREADS * EPT;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  mbar_wait(&c.full[s], phase);
  const T* st = c.ring + (size_t)s * kCap + tid * V;
  PackedMax<T> pmx;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    if (FULL || (k * THREADS + tid) * V < n) {
      const Vec<T> raw = lds_vec(st + k * THREADS * V);
      pmx.add(raw);
      unpack_to_float<T>(raw, &e[k * V]);
    } else {
#pragma unroll
      for (int i = 0; i < V; ++i) e[k * V + i] = -INFINITY;  // exp -> 0, never stored
    }
  }
  pre = 0ull;
  if (pend_rec != nullptr && lane < c.cpr) pre = relay_peek(pend_rec + lane);  // back by the time P2 runs
  const float mw = group_reduce<32>(pmx.get(), MaxF<float>());
  const float nmw = mw == -INFINITY ? 0.f : -mw * kLog2e;  // -inf only: every term 0
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int j = 0; j < EPT; ++j) {
    const float pq = ex2_fast(fmaf(e[j], kLog2e, nmw));
    acc[j & 3] += pq;
    if (!LOG) e[j] = pq;
  }
  const float sw = group_reduce<32>((acc[0] + acc[1]) + (acc[2] + acc[3]), AddF<float>());
  if (lane == 0) c.wp[par][warp] = make_float2(mw, sw);
  __syncthreads();  // the warps' pairs are visible; stage s is fully read
  if (tid == 0 && nxt.g < c.total) {
    const int64_t nstart = (int64_t)nxt.ci * kCap;
    const int64_t nleft = c.dim - nstart;
    const uint32_t nbytes = (uint32_t)((nleft < kCap ? nleft : kCap) * sizeof(T));
    fence_proxy_async_smem();
    mbar_expect_tx(&c.full[s], nbytes);
    bulk_load_row(c.ring + (size_t)s * kCap, c.x + nxt.row * c.dim + nstart, nbytes, &c.full[s]);
  }
  if (warp == 0) {
    const float2 q = lane < NW ? c.wp[par][lane] : make_float2(-INFINITY, 0.f);
    float mc, sc;
    relay_fold(q.x, q.y, mc, sc);
    if (lane == 0) {
      const unsigned long long w = (unsigned long long)__float_as_uint(mc) | ((unsigned long long)__float_as_uint(sc) << 32);
      asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(c.rec + pos.g), "l"(w) : "memory");
    }
  }
  mw_out = mw;
}

template <typename T, bool LOG, int THREADS, int EPT, bool FULL>
__device__ __forceinline__ void relay_p2(const RelayCtx<T>& c, const RelayPos& ch, int n, const float (&e)[EPT], float mw,
                                         unsigned long long pre) {
  constexpr int V = Vec<T>::N;
  constexpr int K = EPT / V;
  constexpr int kCap = THREADS * EPT;
  constexpr float kLog2e = 1.4426950408889634f;
  constexpr float kClip = 1.603810890548638e-28f;  // exp(-64)
  const int tid = threadIdx.x, lane = tid & 31;
  // every warp gathers the row's records itself (no CTA barrier in P2)
  const unsigned long long* rrec = c.rec + (ch.g - ch.ci);
  float ml = -INFINITY, sl = 0.f;
  for (int j = lane; j < c.cpr; j += 32) {
    unsigned long long w;
    w = (j < 32 && pre != 0ull) ? pre : relay_poll(rrec + j, c.timeout_ns);
    const float pm = __uint_as_float((unsigned)(w & 0xffffffffull)), ps = __uint_as_float((unsigned)(w >> 32));
    const float nm = fmaxf(ml, pm);
    const float fa = ml == -INFINITY ? 0.f : ex2_fast((ml - nm) * kLog2e);
    const float fb = pm == -INFINITY ? 0.f : ex2_fast((pm - nm) * kLog2e);
    sl = fmaf(sl, fa, ps * fb);
    ml = nm;
  }
  float M, S;
  relay_fold(ml, sl, M, S);
  if (M == -INFINITY) S = __int_as_float(0x7fc00000);  // a row of -inf: NaN, as inf - inf gives in the CPU kernel
  if (c.stats != nullptr && ch.ci == 0 && tid == 0) { c.stats[2 * ch.row] = M; c.stats[2 * ch.row + 1] = logf(S); }
  const float inv = row_constant<LOG, float>(S);
  T* dst = c.y + ch.row * c.dim + (int64_t)ch.ci * kCap + tid * V;
  const float gf = (mw == -INFINITY ? 0.f : ex2_fast((mw - M) * kLog2e)) * inv;
  const float floor_v = kClip * inv;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    if (FULL || (k * THREADS + tid) * V < n) {
      float r[V];
#pragma unroll
      for (int j = 0; j < V; ++j) r[j] = LOG ? max_nan(e[k * V + j] - M, -64.0f) - inv : e[k * V + j] * gf;
      Vec<T> o;
      pack_from_float<T>(r, o);
      if constexpr (!LOG) clip_packed(o, floor_v);
      st_stream(dst + k * THREADS * V, o);
    }
  }
}

template <typename T, bool LOG, int THREADS, int EPT>
__global__ void __launch_bounds__(THREADS) softmax_rows_relay_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t dim,
                                                                     int64_t total, int cpr, int stages,
                                                                     unsigned long long* __restrict__ rec,
                                                                     float* __restrict__ stats, unsigned long long timeout_ns) {
  extern __shared__ __align__(128) unsigned char ring_raw[];
  __shared__ uint64_t full[kRelayMaxStages];
  __shared__ float2 wp[2][32];
  constexpr int kCap = THREADS * EPT;
  const int tid = threadIdx.x;
  RelayCtx<T> c{x, y, dim, total, cpr, stages, rec, stats, timeout_ns, reinterpret_cast<T*>(ring_raw), full, wp};
  // the grid step in (row, chunk-in-row) form, and this CTA's first chunk
  const int64_t G = gridDim.x;
  const int64_t drow = G / cpr;
  const int dci = (int)(G - drow * cpr);
  const int tail = (int)(dim - (int64_t)(cpr - 1) * kCap);  // elements of a row's last chunk
  RelayPos pos{(int64_t)blockIdx.x, (int64_t)blockIdx.x / cpr, (int)(blockIdx.x % cpr)};
  RelayPos nxt = pos;  // the chunk `stages` iterations ahead: what a refill loads
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) mbar_init(&full[s], 1);
    mbar_fence_init();
  }
  __syncthreads();
  for (int s = 0; s < stages; ++s) {
    if (tid == 0 && nxt.g < total) {
      const int64_t start = (int64_t)nxt.ci * kCap;
      const uint32_t bytes = (uint32_t)((nxt.ci == cpr - 1 ? tail : kCap) * sizeof(T));
      mbar_expect_tx(&full[s], bytes);
      bulk_load_row(c.ring + (size_t)s * kCap, x + nxt.row * dim + start, bytes, &full[s]);
    }
    nxt.step(G, drow, dci, cpr);
  }
  float eA[EPT], eB[EPT];
  float mwA = 0.f, mwB = 0.f;
  unsigned long long preA = 0ull, preB = 0ull;
  RelayPos chA{-1, 0, 0}, chB{-1, 0, 0};
  int s = 0;
  uint32_t phase = 0;
  // one P1 (chunk `pos` into e / mw, remembering it in `ch`) followed by the P2 of the chunk before it
#define RELAY_STEP(E1, MW1, CH1, PAR, E2, MW2, CH2, PRE2)                                                              \
  if (pos.g < total) {                                                                                                 \
    const unsigned long long* pend = CH2.g >= 0 ? rec + (CH2.g - CH2.ci) : nullptr;                                    \
    if (pos.ci != cpr - 1 || tail == kCap)                                                                             \
      relay_p1<T, LOG, THREADS, EPT, true>(c, pos, nxt, kCap, s, phase, PAR, E1, MW1, pend, PRE2);                     \
    else                                                                                                               \
      relay_p1<T, LOG, THREADS, EPT, false>(c, pos, nxt, tail, s, phase, PAR, E1, MW1, pend, PRE2);                    \
    CH1 = pos;                                                                                                         \
  } else {                                                                                                             \
    CH1.g = -1;                                                                                                        \
    PRE2 = 0ull;                                                                                                       \
  }                                                                                                                    \
  if (CH2.g >= 0) {                                                                                                    \
    if (CH2.ci != cpr - 1 || tail == kCap) relay_p2<T, LOG, THREADS, EPT, true>(c, CH2, kCap, E2, MW2, PRE2);          \
    else relay_p2<T, LOG, THREADS, EPT, false>(c, CH2, tail, E2, MW2, PRE2);                                           \
  }                                                                                                                    \
  if (CH1.g < 0) break;                                                                                                \
  pos.step(G, drow, dci, cpr);                                                                                         \
  nxt.step(G, drow, dci, cpr);                                                                                         \
  if (++s == stages) { s = 0; phase ^= 1; }
  for (;;) {
    RELAY_STEP(eA, mwA, chA, 0, eB, mwB, chB, preB)
    RELAY_STEP(eB, mwB, chB, 1, eA, mwA, chA, preA)
  }
#undef RELAY_STEP
}

// ---- rows_sweep2: long rows streamed twice, the second time out of L2 ------------------------------------------------
// For rows that do not fit the register / shared-memory routes (and as an alternative to them for vocabulary-sized rows):
// CTA `rank` of a cluster of cs (1..8) owns slice [rank*slice, +slice) of one row.  Sweep 1 streams the slice from HBM with
// 16-byte loads and keeps a per-thread running (max, sum) — four vectors at a time, one rescale per batch, 1.25 exp per
// element — folded over the CTA and, through DSMEM after ONE cluster barrier, over the cluster.  Sweep 2 re-reads the
// slice, normalises and stores; the launch keeps the slices of all resident CTAs within ~64 MB so that this second
// read comes out of the 126 MB L2: HBM sees one read and one write.  No shared-memory staging, no limit on the row
// length, cs x as many CTAs as rows when there are few rows.
template <typename T, bool LOG>
__global__ void __launch_bounds__(512) softmax_rows_sweep2_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t dim,
                                                                  int64_t slice, float* __restrict__ stats) {
  namespace cg = cooperative_groups;
  constexpr int V = Vec<T>::N;
  constexpr int U = 4;
  constexpr float kLog2e = 1.4426950408889634f;
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned rank = cluster.block_rank(), cs = cluster.num_blocks();
  __shared__ float red[32][3];
  __shared__ float xchg[3];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
  const int64_t row = (int64_t)blockIdx.x / cs;
  const int64_t begin = (int64_t)rank * slice;
  int64_t n = dim - begin;
  n = n < 0 ? 0 : (n > slice ? slice : n);
  const T* src = x + row * dim + begin;
  T* dst = y + row * dim + begin;
  const int64_t step = (int64_t)blockDim.x * V;
  MaxSum t{-INFINITY, 0.f, 0.f};
  int64_t i = (int64_t)tid * V;
  for (; i + (U - 1) * step < n; i += U * step) {
    Vec<T> raw[U];
#pragma unroll
    for (int u = 0; u < U; ++u) raw[u] = ld_stream(src + i + u * step);
    float v[U * V];
#pragma unroll
    for (int u = 0; u < U; ++u) unpack_to_float<T>(raw[u], &v[u * V]);
    ms_accumulate<U * V>(t, v);
  }
  for (; i < n; i += step) {
    float v[V];
    unpack_to_float<T>(ld_stream(src + i), v);
    ms_accumulate<V>(t, v);
  }
  t = ms_warp_reduce(t);
  if (lane == 0) { red[warp][0] = t.m; red[warp][1] = t.s; red[warp][2] = t.a; }
  __syncthreads();
  MaxSum g{-INFINITY, 0.f, 0.f};
  if (lane < nwarps) g = MaxSum{red[lane][0], red[lane][1], red[lane][2]};
  g = ms_warp_reduce(g);
  if (cs > 1) {
    if (tid == 0) { xchg[0] = g.m; xchg[1] = g.s; xchg[2] = g.a; }
    cluster.sync();
    MaxSum p{-INFINITY, 0.f, 0.f};
    if (lane < (int)cs) {
      const float* r = cluster.map_shared_rank(xchg, lane);
      p = MaxSum{r[0], r[1], r[2]};
    }
    g = ms_warp_reduce(p);
  }
  float gs = g.s + g.a;
  if (g.m == -INFINITY) gs = __int_as_float(0x7fc00000);  // a row of -inf: NaN, as inf - inf gives in the CPU kernel
  if (stats != nullptr && rank == 0 && tid == 0) { stats[2 * row] = g.m; stats[2 * row + 1] = logf(gs); }
  const float k = LOG ? logf(gs) : 1.0f / gs;
  const float nml = -g.m * kLog2e;
  // sweep 2: each chunk is read and written by the same thread (in place is fine)
  i = (int64_t)tid * V;
  for (; i + (U - 1) * step < n; i += U * step) {
    Vec<T> raw[U];
#pragma unroll
    for (int u = 0; u < U; ++u) raw[u] = ld_stream(src + i + u * step);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float v[V];
      unpack_to_float<T>(raw[u], v);
#pragma unroll
      for (int q = 0; q < V; ++q) v[q] = LOG ? max_nan(v[q] - g.m, -64.0f) - k : fast_clipped_exp(v[q], nml) * k;
      Vec<T> o;
      pack_from_float<T>(v, o);
      st_stream(dst + i + u * step, o);
    }
  }
  for (; i < n; i += step) {
    float v[V];
    unpack_to_float<T>(ld_stream(src + i), v);
#pragma unroll
    for (int q = 0; q < V; ++q) v[q] = LOG ? max_nan(v[q] - g.m, -64.0f) - k : fast_clipped_exp(v[q], nml) * k;
    Vec<T> o;
    pack_from_float<T>(v, o);
    st_stream(dst + i, o);
  }
  if (cs > 1) cluster.sync();  // nobody leaves while a peer may still be reading its xchg[]
}

// ---- rows_stream -----------------------------------------------------------------------
template <typename T, bool LOG>
__global__ void __launch_bounds__(1024) softmax_rows_stream_kernel(const T* x, T* y, int64_t dim,
                                                                   typename AccType<T>::type* stats) {
  using A = typename AccType<T>::type;
  __shared__ A red[33];
  const T* src = x + (int64_t)blockIdx.x * dim;
  T* dst = y + (int64_t)blockIdx.x * dim;
  A m = neg_inf<A>();
  for (int64_t i = threadIdx.x; i < dim; i += blockDim.x) {
    A a = cvt<A>(src[i]);
    m = a > m ? a : m;
  }
  m = block_reduce(m, MaxF<A>(), neg_inf<A>(), red);
  A sum = A(0);
  for (int64_t i = threadIdx.x; i < dim; i += blockDim.x) sum += clipped_exp<A>(cvt<A>(src[i]), m);
  sum = block_reduce(sum, AddF<A>(), A(0), red);
  const A inv = row_constant<LOG, A>(sum);
  if (stats != nullptr && threadIdx.x == 0) { stats[2 * (int64_t)blockIdx.x] = m; stats[2 * (int64_t)blockIdx.x + 1] = log_acc<A>(sum); }
  // in-place safe: each element is read and written by the same thread in this sweep
  for (int64_t i = threadIdx.x; i < dim; i += blockDim.x)
    dst[i] = cvt<T>(LOG ? clipped_diff<A>(cvt<A>(src[i]), m) - inv : clipped_exp<A>(cvt<A>(src[i]), m) * inv);
}

// ---- cols: softmax along the middle axis of [outer][dim][inner] ---------------------------
template <typename T, bool LOG>
__global__ void __launch_bounds__(256) softmax_cols_kernel(const T* x, T* y, int64_t dim,
                                                           int64_t inner, int64_t col_tiles,
                                                           typename AccType<T>::type* stats) {
  using A = typename AccType<T>::type;
  const int64_t o = (int64_t)blockIdx.x / col_tiles;
  const int64_t ct = (int64_t)blockIdx.x - o * col_tiles;
  const int64_t col = ct * blockDim.x + threadIdx.x;
  if (col >= inner) return;
  const T* src = x + o * dim * inner + col;
  T* dst = y + o * dim * inner + col;
  A m = neg_inf<A>();
  for (int64_t r = 0; r < dim; ++r) {
    A a = cvt<A>(src[r * inner]);
    m = a > m ? a : m;
  }
  A sum = A(0);
  for (int64_t r = 0; r < dim; ++r) sum += clipped_exp<A>(cvt<A>(src[r * inner]), m);
  const A inv = row_constant<LOG, A>(sum);
  if (stats != nullptr) { stats[2 * (o * inner + col)] = m; stats[2 * (o * inner + col) + 1] = log_acc<A>(sum); }
  for (int64_t r = 0; r < dim; ++r)
    dst[r * inner] = cvt<T>(LOG ? clipped_diff<A>(cvt<A>(src[r * inner]), m) - inv : clipped_exp<A>(cvt<A>(src[r * inner]), m) * inv);
}

// ---- cols_tile: softmax along the middle axis, vectorised along `inner` ------------------------------------------
// blockDim = (32, TY).  Thread (tx, ty) owns V adjacent columns (one 16-byte load per row: a warp reads 512
// contiguous bytes) and rows ty, ty + TY, ... of the reduced axis.  Sweep 1 keeps a running (max, sum) per column —
// four rows at a time, one rescale per batch, 1.25 exp per element — the TY partial pairs of a column are folded
// through shared memory; sweep 2 re-reads x (partly from L2), normalises and writes.  Two reads + one write of HBM at
// worst, against three sweeps of scalar, one-thread-per-column loads in softmax_cols_kernel (1.3 TB/s on bf16
// [64,2048,4096] axis 1, profiles/r2_notes.txt).  Elements seen while a column's running max is still -inf contribute
// their clipped exp(-64) to a separate, never rescaled sum (the CPU kernel's max(x - max, -64) does not depend on the
// max for them); NaN and +inf poison the column through the sum, a column of -inf only comes out NaN.
template <typename T, bool LOG>
__global__ void __launch_bounds__(256) softmax_cols_tile_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t dim,
                                                                int64_t inner, int64_t col_tiles, float* __restrict__ stats) {
  constexpr int V = Vec<T>::N;
  constexpr int U = 4;
  constexpr float kLog2e = 1.4426950408889634f;
  extern __shared__ __align__(16) float cols_smem[];  // [TY][32 * V][2]
  const int tx = threadIdx.x, ty = threadIdx.y, TY = blockDim.y;
  const int64_t o = (int64_t)blockIdx.x / col_tiles;
  const int64_t ct = (int64_t)blockIdx.x - o * col_tiles;
  const int64_t col = (ct * 32 + tx) * V;
  const bool active = col < inner;
  const T* src = x + o * dim * inner + col;
  T* dst = y + o * dim * inner + col;
  float m[V], sacc[V], sabs[V];
#pragma unroll
  for (int k = 0; k < V; ++k) { m[k] = -INFINITY; sacc[k] = 0.f; sabs[k] = 0.f; }
  const float kClip = 1.603810890548638e-28f;  // exp(-64)
  if (active) {
    int64_t r = ty;
    for (; r + (int64_t)(U - 1) * TY < dim; r += (int64_t)U * TY) {
      float v[U][V];
#pragma unroll
      for (int u = 0; u < U; ++u) unpack_to_float<T>(ld_stream(src + (r + (int64_t)u * TY) * inner), v[u]);
#pragma unroll
      for (int k = 0; k < V; ++k) {
        float nm = m[k];
#pragma unroll
        for (int u = 0; u < U; ++u) nm = fmaxf(nm, v[u][k]);
        if (nm == -INFINITY) {  // nothing finite yet: -inf clips to exp(-64) whatever the max turns out to be; NaN propagates
#pragma unroll
          for (int u = 0; u < U; ++u) sabs[k] += (v[u][k] != v[u][k]) ? v[u][k] : kClip;
        } else {
          float resc;
          asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(resc) : "f"((m[k] - nm) * kLog2e));  // 0 when m was -inf
          float add = 0.f;
#pragma unroll
          for (int u = 0; u < U; ++u) add += fast_clipped_exp(v[u][k], -nm * kLog2e);
          sacc[k] = fmaf(sacc[k], resc, add);
          m[k] = nm;
        }
      }
    }
    for (; r < dim; r += TY) {
      float v[V];
      unpack_to_float<T>(ld_stream(src + r * inner), v);
#pragma unroll
      for (int k = 0; k < V; ++k) {
        const float nm = fmaxf(m[k], v[k]);
        if (nm == -INFINITY) {
          sabs[k] += (v[k] != v[k]) ? v[k] : kClip;
        } else {
          float resc;
          asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(resc) : "f"((m[k] - nm) * kLog2e));
          sacc[k] = fmaf(sacc[k], resc, fast_clipped_exp(v[k], -nm * kLog2e));
          m[k] = nm;
        }
      }
    }
  }
  // fold the TY partials of every column (rank order, identical in every thread that needs them)
  float* mine = cols_smem + ((size_t)ty * 32 + tx) * V * 3;
#pragma unroll
  for (int k = 0; k < V; ++k) { mine[3 * k] = m[k]; mine[3 * k + 1] = sacc[k]; mine[3 * k + 2] = sabs[k]; }
  __syncthreads();
  float M[V], inv[V];
#pragma unroll
  for (int k = 0; k < V; ++k) {
    float gm = -INFINITY;
    for (int t = 0; t < TY; ++t) gm = fmaxf(gm, cols_smem[(((size_t)t * 32 + tx) * V + k) * 3]);
    float gs = 0.f;
    for (int t = 0; t < TY; ++t) {
      const float* p = cols_smem + (((size_t)t * 32 + tx) * V + k) * 3;
      float f = 0.f;
      if (p[0] != -INFINITY) asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(f) : "f"((p[0] - gm) * kLog2e));
      gs += fmaf(p[1], f, p[2]);
    }
    if (gm == -INFINITY) gs = __int_as_float(0x7fc00000);  // a column of -inf: NaN, as inf - inf gives in the CPU kernel
    M[k] = gm;
    inv[k] = row_constant<LOG, float>(gs);
    if (stats != nullptr && ty == 0 && active && col + k < inner) {
      stats[2 * (o * inner + col + k)] = gm;
      stats[2 * (o * inner + col + k) + 1] = logf(gs);
    }
  }
  if (!active) return;
  float nml[V];
#pragma unroll
  for (int k = 0; k < V; ++k) nml[k] = -M[k] * kLog2e;
  int64_t r = ty;
  for (; r + (int64_t)(U - 1) * TY < dim; r += (int64_t)U * TY) {  // U loads in flight per thread, as in sweep 1
    Vec<T> raw[U];
#pragma unroll
    for (int u = 0; u < U; ++u) raw[u] = ld_stream(src + (r + (int64_t)u * TY) * inner);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float v[V];
      unpack_to_float<T>(raw[u], v);
#pragma unroll
      for (int k = 0; k < V; ++k) v[k] = LOG ? max_nan(v[k] - M[k], -64.0f) - inv[k] : fast_clipped_exp(v[k], nml[k]) * inv[k];
      Vec<T> ov;
      pack_from_float<T>(v, ov);
      st_stream(dst + (r + (int64_t)u * TY) * inner, ov);
    }
  }
  for (; r < dim; r += TY) {
    float v[V];
    unpack_to_float<T>(ld_stream(src + r * inner), v);
#pragma unroll
    for (int k = 0; k < V; ++k) v[k] = LOG ? max_nan(v[k] - M[k], -64.0f) - inv[k] : fast_clipped_exp(v[k], nml[k]) * inv[k];
    Vec<T> ov;
    pack_from_float<T>(v, ov);
    st_stream(dst + r * inner, ov);
  }
}

template <typename T>
__global__ void softmax_fill_one_kernel(T* y, float v) { y[0] = cvt<T>(typename AccType<T>::type(v)); }

// ---- host -------------------------------------------------------------------------------
template <typename T, bool VECTOR, bool LOG>
int launch_rows_reg(const T* x, T* y, int64_t outer, int dim, int G, int C, typename AccType<T>::type* stats, cudaStream_t st) {
  const int rows_per_cta = 256 / G;
  const int64_t blocks = (outer + rows_per_cta - 1) / rows_per_cta;
  if (blocks > 0x7fffffffLL) return fail(PDB200_UNSUPPORTED, "softmax: too many rows for one launch");
#define PDB_SM_C(CC)                                                                                \
  case CC:                                                                                          \
    softmax_rows_reg_kernel<T, CC, VECTOR, LOG><<<(unsigned)blocks, 256, 0, st>>>(x, y, outer, dim, G, stats); \
    break;
  switch (C) {
    PDB_SM_C(1) PDB_SM_C(2) PDB_SM_C(4) PDB_SM_C(8) PDB_SM_C(16)
    default: return fail(PDB200_UNSUPPORTED, "softmax: bad chunk count");
  }
#undef PDB_SM_C
  return after_launch("softmax_rows_reg");
}

// TMA row pipeline (float / bf16 / fp16 rows of 64..2048 16-byte chunks, many rows).
template <typename T, bool LOG>
int try_softmax_pipe(const T* x, T* y, int64_t outer, int64_t dim, typename AccType<T>::type* stats, cudaStream_t st) {
  if constexpr (std::is_same<T, double>::value) {
    return -1;  // the pipeline computes in float
  } else {
  constexpr int VN = Vec<T>::N;
  const int64_t nchunk = dim / VN;
  // chunks per thread: 32 values a thread while that leaves a CTA of >= 64 threads — 4 chunks for 16-bit rows, 8 for fp32 rows
  // of up to 4096 elements (fp32 [.,2048]: 1.310 -> 1.248 ms on a quarter of config 4, [.,4096] 0.799 -> 0.708, [.,1024]
  // 0.656 -> 0.622; 8192-element fp32 rows stay at 4 chunks x 512 threads: 0.632 against 0.716) — profiles/r2_notes.txt
  int C;
  if (sizeof(T) == 2) {
    C = nchunk >= 256 ? 4 : (nchunk >= 128 ? 2 : 1);
  } else {
    C = nchunk <= 1024 ? 8 : 4;
    while (C > 1 && nchunk / C < 64) C >>= 1;
  }
  if (const int o = knobs().softmax_pipe_c) C = o;
  while (C < 8 && (nchunk + C - 1) / C > 512) C <<= 1;  // the kernel is built for <= 512 threads
  int threads = (int)((nchunk + C - 1) / C);
  if (threads > 512) return -1;
  threads = (threads + 31) / 32 * 32;
  const size_t row_bytes = (size_t)dim * sizeof(T);
  auto launch = [&](auto kern) -> int {
    int per_sm = 0, stages = 0;
    size_t smem = 0;
    int want = 768 / threads;  // ~768 threads per SM, 2..6 CTAs; a 512-thread CTA (rows of 16-32 KB) runs alone:
    const int lo = threads >= 512 ? 1 : 2;  // two of them put 128 KB of rows in flight, fp32 [65536,8192] 0.726 vs 0.634 ms
    want = want < lo ? lo : (want > 6 ? 6 : want);
    // ... and no more CTAs than keep ~64 KB of rows in flight per SM at two stages each (as in rms_norm / softmax_grad):
    // 16 KB rows run on 2 CTAs per SM, not 3 — bf16 [.,8192] 0.374 -> 0.325 ms, fp32 [.,4096] 0.701 -> 0.674
    const int by_bytes = (int)((64 * 1024) / (2 * row_bytes));
    if (want > by_bytes) want = by_bytes < lo ? lo : by_bytes;
    FuncAttrHold hold;  // released when this lambda returns, i.e. after the launch
    if (!pipe_geometry(kern, threads, row_bytes, want, 3, &per_sm, &stages, &smem, &hold)) return -1;
    int64_t grid = (int64_t)per_sm * sm_count();
    // 64-thread CTAs (16-bit rows of 2048): many short-lived CTAs of 4 rows beat the persistent grid, 0.660 -> 0.620 ms on
    // bf16 [8,32,2048,2048]; 256-thread CTAs lose (fp32: 1.305 -> 1.49 ms), so they stay persistent (profiles/r1e_notes.txt)
    if (threads <= 64) grid = std::max<int64_t>(grid, outer / 4);
    if (const int rpc = knobs().pipe_rows_per_cta; rpc >= 1)  // tuning knob
      grid = std::max<int64_t>((int64_t)per_sm * sm_count(), outer / rpc);
    if (grid > 0x7fffffffLL) grid = 0x7fffffffLL;
    if (grid > outer) grid = outer;
    kern<<<(unsigned)grid, threads, smem, st>>>(x, y, outer, (int)dim, stages, pipe_contiguous(), stats);
    return after_launch("softmax_pipe");
  };
  const bool exact = (int64_t)threads * C == nchunk;
#define PDB_SM_PIPE(CC) \
  return exact ? launch(softmax_pipe_kernel<T, CC, true, LOG>) : launch(softmax_pipe_kernel<T, CC, false, LOG>);
  if (C == 8) { PDB_SM_PIPE(8) }
  if (C == 4) { PDB_SM_PIPE(4) }
  if (C == 2) { PDB_SM_PIPE(2) }
  PDB_SM_PIPE(1)
#undef PDB_SM_PIPE
  }
}

// Two-sweep route (softmax_rows_sweep2_kernel): any row length.  < 0: could not launch.
template <typename T, bool LOG>
int try_softmax_sweep2(const T* x, T* y, int64_t outer, int64_t dim, typename AccType<T>::type* stats, cudaStream_t st) {
  if constexpr (std::is_same<T, double>::value) {
    return -1;  // computes in float
  } else {
    constexpr int VN = Vec<T>::N;
    int cs = 1;
    while (cs < 8 && outer * cs < 2 * (int64_t)sm_count()) cs <<= 1;
    int64_t slice = (dim + cs - 1) / cs;
    slice = (slice + VN - 1) / VN * VN;
    if (outer * cs > 0x7fffffffLL) return -1;
    const int threads = 512;
    // keep the slices of all resident CTAs within ~64 MB, so that sweep 2 is served by L2: cap the CTAs per SM with
    // dynamic shared memory the kernel never touches
    const size_t slice_bytes = (size_t)slice * sizeof(T);
    int per_sm = (int)((64u << 20) / ((size_t)sm_count() * std::max<size_t>(slice_bytes, 1)));
    per_sm = per_sm < 1 ? 1 : (per_sm > 4 ? 4 : per_sm);
    const size_t smem = per_sm >= 4 ? 0 : (size_t)(227 * 1024) / per_sm - 4096;
    auto kern = softmax_rows_sweep2_kernel<T, LOG>;
    FuncAttrHold hold;
    if (!func_attr_apply(reinterpret_cast<const void*>(kern), (int)smem, smem ? 100 : -1, &hold)) return -1;
    if (cs == 1) {
      kern<<<(unsigned)outer, threads, smem, st>>>(x, y, dim, slice, stats);
      return after_launch("softmax_rows_sweep2");
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)(outer * cs));
    cfg.blockDim = dim3(threads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cs;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (cudaLaunchKernelEx(&cfg, kern, x, y, dim, slice, stats) != cudaSuccess) {
      cudaGetLastError();
      return -1;
    }
    return after_launch("softmax_rows_sweep2");
  }
}

// The relay kernel's record array.  One buffer per (device, stream), kept for the life of the process and grown when a call
// needs more: calls on one stream are ordered, so the memset + kernel of the next call cannot overtake the previous one.
// (The first version took the array from cudaMallocAsync on every call: with the default pool's release threshold of 0 the
// memory goes back to the driver whenever the stream idles, and a bench.py run showed the call at 4.2 ms instead of 0.78.)
// While the stream is being captured into a CUDA graph a new buffer cannot be allocated the ordinary way: the capture then
// records a stream-ordered allocation (a graph memory node) instead.
struct RelayRecords {
  unsigned long long* p = nullptr;
  bool stream_ordered = false;  // from cudaMallocAsync: the caller frees it on the stream after the launch
};
inline RelayRecords relay_records(size_t bytes, cudaStream_t st) {
  struct Buf { void* p; size_t cap; };
  static std::mutex mu;
  static std::map<std::pair<int, cudaStream_t>, Buf> cache;
  RelayRecords r;
  int dev = 0;
  cudaGetDevice(&dev);
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cap) != cudaSuccess) {
    cudaGetLastError();
    return r;
  }
  std::lock_guard<std::mutex> lock(mu);
  auto it = cache.find({dev, st});
  if (it != cache.end() && it->second.cap >= bytes) {
    r.p = static_cast<unsigned long long*>(it->second.p);
    return r;
  }
  if (cap != cudaStreamCaptureStatusNone || (it == cache.end() && cache.size() >= 64)) {
    void* q = nullptr;
    if (cudaMallocAsync(&q, bytes, st) != cudaSuccess) {
      cudaGetLastError();
      return r;
    }
    r.p = static_cast<unsigned long long*>(q);
    r.stream_ordered = true;
    return r;
  }
  if (it != cache.end()) {  // grow: the old buffer may still be in use by work queued on this stream
    if (cudaStreamSynchronize(st) != cudaSuccess) {
      cudaGetLastError();
      return r;
    }
    cudaFree(it->second.p);
    cache.erase(it);
  }
  const size_t want = std::max<size_t>(bytes + bytes / 2, (size_t)1 << 20);
  void* q = nullptr;
  if (cudaMalloc(&q, want) != cudaSuccess) {
    cudaGetLastError();
    return r;
  }
  cache[{dev, st}] = Buf{q, want};
  r.p = static_cast<unsigned long long*>(q);
  return r;
}

// Relay route: returns < 0 when it cannot run (no cooperative launch, no memory for the records, launch failure).
template <typename T, bool LOG, int THREADS, int EPT>
int launch_softmax_relay(const T* x, T* y, int64_t outer, int64_t dim, float* stats, cudaStream_t st) {
  constexpr int VN = Vec<T>::N;
  constexpr int kCap = THREADS * EPT;
  const int64_t cpr64 = (dim + kCap - 1) / kCap;
  if (cpr64 > 16384 || outer <= 0) return -1;
  const int cpr = (int)cpr64;  // kCap elements per chunk, the row's last chunk takes what is left
  (void)VN;
  const int64_t total = outer * cpr;
  const size_t stage_bytes = (size_t)kCap * sizeof(T);
  auto kern = softmax_rows_relay_kernel<T, LOG, THREADS, EPT>;
  // CTAs per SM by registers / threads (no dynamic shared memory in the query), then the ring sized to share ~200 KB
  int occ = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, THREADS, 0) != cudaSuccess || occ < 1) {
    cudaGetLastError();
    return -1;
  }
  occ = std::min(occ, 8);
  int stages = 0;
  for (; occ >= 1; --occ) {
    stages = (int)std::min<size_t>(kRelayMaxStages, (200 * 1024 / occ - 1024) / stage_bytes);
    if (stages >= 3) break;
  }
  if (occ < 1) return -1;
  if (knobs().relay_stages >= 2 && knobs().relay_stages <= stages) stages = knobs().relay_stages;  // tuning knob
  const int64_t grid = std::min<int64_t>(total, (int64_t)sm_count() * occ);
  const size_t smem = stage_bytes * stages;
  FuncAttrHold hold;  // the attribute and the launch under one lock (another thread may launch this kernel with other sizes)
  if (!func_attr_apply(reinterpret_cast<const void*>(kern), (int)smem, 100, &hold)) {
    cudaGetLastError();
    return -1;
  }
  const RelayRecords records = relay_records((size_t)total * sizeof(unsigned long long), st);
  unsigned long long* rec = records.p;
  if (rec == nullptr) return -1;
  int rc = -1;
  if (cudaMemsetAsync(rec, 0, (size_t)total * sizeof(unsigned long long), st) == cudaSuccess) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;  // every CTA resident: the polls cannot wait for a CTA that never starts
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const unsigned long long timeout_ns = 10ull * 1000 * 1000 * 1000;
    if (cudaLaunchKernelEx(&cfg, kern, x, y, dim, total, cpr, stages, rec, stats, timeout_ns) == cudaSuccess)
      rc = after_launch("softmax_rows_relay");
  }
  if (rc < 0) cudaGetLastError();
  if (records.stream_ordered) cudaFreeAsync(rec, st);
  return rc;
}
template <typename T, bool LOG>
int try_softmax_relay(const T* x, T* y, int64_t outer, int64_t dim, float* stats, cudaStream_t st) {
  switch (knobs().relay_cfg) {  // tuning knob: threads x elements per thread
    case 1: return launch_softmax_relay<T, LOG, 256, 32>(x, y, outer, dim, stats, st);
    case 3: return launch_softmax_relay<T, LOG, 512, 32>(x, y, outer, dim, stats, st);
    case 6: return launch_softmax_relay<T, LOG, 128, 32>(x, y, outer, dim, stats, st);
    default: break;
  }
  // chunks that a row fills several times over: 256 x 32 elements (two CTAs per SM; 512 x 32 is no faster on vocabulary
  // rows and slower on 1M-element rows, 0.138 vs 0.119 ms), 128 x 32 for shorter rows
  if (dim >= 4 * 256 * 32) return launch_softmax_relay<T, LOG, 256, 32>(x, y, outer, dim, stats, st);
  return launch_softmax_relay<T, LOG, 128, 32>(x, y, outer, dim, stats, st);
}

// Cluster route: returns < 0 when the row does not fit 8 CTAs' shared memory.
template <typename T, bool LOG>
int try_softmax_cluster(const T* x, T* y, int64_t outer, int64_t dim, typename AccType<T>::type* stats, cudaStream_t st) {
  constexpr int VN = Vec<T>::N;
  const size_t row_bytes = (size_t)dim * sizeof(T);
  const size_t kMaxSlice = 200 * 1024;
  int cs = 1;
  // split until a slice is <= 32 KB (several CTAs then share an SM and overlap their phases), and
  // further while there are too few rows to fill the SMs
  size_t slice_target = 64 * 1024;  // 64 KB slices: bf16 [4096,128256] 0.562 -> 0.517 ms vs 32 KB (fewer cluster barriers per byte)
  if (knobs().cluster_slice_kb >= 4) slice_target = (size_t)knobs().cluster_slice_kb * 1024;  // tuning knob
  while (cs < 8 && (row_bytes / cs > slice_target || outer * cs < 2 * (int64_t)sm_count())) cs <<= 1;
  int64_t slice = (dim + cs - 1) / cs;
  slice = (slice + VN - 1) / VN * VN;
  if ((size_t)slice * sizeof(T) > kMaxSlice) return -1;
  if (outer * cs > 0x7fffffffLL) return -1;
  int threads = 256;
  while (threads < 1024 && (int64_t)threads * VN * 16 < slice) threads <<= 1;
  const size_t smem = (size_t)slice * sizeof(T);
  using A = typename AccType<T>::type;
  auto kern = softmax_rows_cluster_kernel<T, LOG>;
  if constexpr (std::is_same<A, float>::value) {
    if (knobs().softmax_long != 5) {  // 5: the first (online-rescale) kernel, kept for A/B runs and for double rows
      kern = softmax_rows_cluster_lean_kernel<T, LOG>;
      threads = 128;
      while (threads < 256 && (int64_t)threads * VN * 8 < slice) threads <<= 1;  // 64+ elements per thread; 256 beat 512 and 128
      if (knobs().cluster_threads >= 64) threads = knobs().cluster_threads;
    }
  }
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxSlice) != cudaSuccess) {
    cudaGetLastError();
    return -1;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)(outer * cs));
  cfg.blockDim = dim3(threads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cs;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, x, y, dim, slice, stats);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return -1;  // e.g. the cluster cannot be co-scheduled with this much shared memory: fall back
  }
  return after_launch("softmax_rows_cluster");
}

// ---- cols_cluster: softmax along a middle axis, one read: a column tile held in the shared memory of a cluster ---------
// softmax_cols_tile_kernel reads x twice, and on a tensor larger than L2 both reads come from HBM (DRAM read = 2 x input,
// profiles/r2_softmax_cols_bf16.txt).  (Tried first: the same two sweeps with narrow tiles and a capped number of resident
// CTAs so that the second sweep would hit L2 — 0.76-1.18 ms against 0.65: one or two CTAs per SM cannot overlap one tile's
// reads with another's writes.)  Here a cluster of cs <= 8 CTAs owns a tile of 16 vectors (256 bytes) x all `dim` rows; CTA
// `rank` keeps rows [rank*TR, +TR) of it in shared memory (<= 64 KB), forms the column maxima and the sums relative to ITS
// OWN maxima, publishes the two 16*V-float arrays, and after ONE cluster barrier every CTA folds the cs pairs per column
// through DSMEM and writes its rows: x is read once.  256 threads = 16 vector lanes x 16 row lanes.
template <typename T, bool LOG>
__global__ void __launch_bounds__(256) softmax_cols_cluster_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t dim,
                                                                   int64_t inner, int64_t col_tiles, int tr,
                                                                   float* __restrict__ stats) {
  namespace cg = cooperative_groups;
  constexpr int V = Vec<T>::N;
  constexpr int TCV = 16;            // vectors per row segment
  constexpr int RL = 16;             // row lanes
  constexpr float kLog2e = 1.4426950408889634f;
  constexpr float kClip = 1.603810890548638e-28f;  // exp(-64)
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned rank = cluster.block_rank(), cs = cluster.num_blocks();
  extern __shared__ __align__(16) unsigned char cc_raw[];
  T* tile = reinterpret_cast<T*>(cc_raw);                     // [tr][TCV * V]
  __shared__ __align__(16) float part[8][TCV * V];            // per-warp column partials; later [0]: the columns' maxima
  __shared__ __align__(16) float part2[1][TCV * V];           // the columns' sums
  __shared__ __align__(16) float xm[TCV * V], xs[TCV * V];    // this CTA's column (max, sum): read by the peers
  const int tid = threadIdx.x, vx = tid & 15, ry = tid >> 4, warp = tid >> 5;
  const int64_t cl = (int64_t)blockIdx.x / cs;
  const int64_t o = cl / col_tiles, ct = cl - o * col_tiles;
  const int64_t col = (ct * TCV + vx) * V;
  const bool active = col < inner;
  const int row0 = (int)rank * tr;
  int nrows = (int)dim - row0;
  nrows = nrows < 0 ? 0 : (nrows > tr ? tr : nrows);
  const T* src = x + (o * dim + row0) * inner + col;
  T* dst = y + (o * dim + row0) * inner + col;
  // phase 1: HBM -> shared (U loads in flight per thread), column maxima on the way (elementwise on the raw vectors)
  float m[V];
  {
    Vec<T> rmax;
    float ninf[V];
#pragma unroll
    for (int k = 0; k < V; ++k) ninf[k] = -INFINITY;
    pack_from_float<T>(ninf, rmax);
    if (active) {
      // every row segment of this thread as a 16-byte asynchronous copy straight into the tile (no registers: all of the
      // CTA's 64 KB are in flight at once; with 8 register-staged loads per thread the kernel ran 0.59 ms, not faster than
      // the two-sweep kernel), then the maxima from shared memory
      for (int r = ry; r < nrows; r += RL) {
        const unsigned dsts = smem_u32(tile + ((size_t)r * TCV + vx) * V);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dsts), "l"(src + (int64_t)r * inner) : "memory");
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      asm volatile("cp.async.wait_group 0;" ::: "memory");
#pragma unroll 4
      for (int r = ry; r < nrows; r += RL) vmax_raw(rmax, lds_vec(tile + ((size_t)r * TCV + vx) * V));
    }
    unpack_to_float<T>(rmax, m);
  }
  // column maxima of this CTA: the two row lanes of a warp by shuffle, the 8 warps through shared memory
#pragma unroll
  for (int k = 0; k < V; ++k) m[k] = fmaxf(m[k], __shfl_xor_sync(0xffffffffu, m[k], 16));
  if ((tid & 31) < 16) {
#pragma unroll
    for (int k = 0; k < V; ++k) part[warp][vx * V + k] = m[k];
  }
  __syncthreads();  // also: the tile is complete
  float nml[V];
#pragma unroll
  for (int k = 0; k < V; ++k) {
    float g = part[0][vx * V + k];
#pragma unroll
    for (int w = 1; w < 8; ++w) g = fmaxf(g, part[w][vx * V + k]);
    m[k] = g;
    nml[k] = g == -INFINITY ? 0.f : -g * kLog2e;  // a column slice of -inf only: every term 0
  }
  __syncthreads();  // part[] is reused for the sums
  // phase 2: sums relative to this CTA's own maxima
  float s[V];
  {
    const F2 l2e = f2_splat(kLog2e);
    F2 n2[V / 2], s2[V / 2];
#pragma unroll
    for (int j = 0; j < V / 2; ++j) {
      n2[j] = f2_make(nml[2 * j], nml[2 * j + 1]);
      s2[j] = f2_splat(0.f);
    }
    if (active) {
#pragma unroll 2
      for (int r = ry; r < nrows; r += RL) {
        float v[V];
        unpack_to_float<T>(lds_vec(tile + ((size_t)r * TCV + vx) * V), v);
#pragma unroll
        for (int j = 0; j < V / 2; ++j) {
          const F2 t = f2_fma(f2_make(v[2 * j], v[2 * j + 1]), l2e, n2[j]);
          s2[j] = f2_add(s2[j], f2_make(ex2_fast(f2_lo(t)), ex2_fast(f2_hi(t))));
        }
      }
    }
#pragma unroll
    for (int j = 0; j < V / 2; ++j) {
      s[2 * j] = f2_lo(s2[j]);
      s[2 * j + 1] = f2_hi(s2[j]);
    }
  }
#pragma unroll
  for (int k = 0; k < V; ++k) s[k] += __shfl_xor_sync(0xffffffffu, s[k], 16);
  if ((tid & 31) < 16) {
#pragma unroll
    for (int k = 0; k < V; ++k) part[warp][vx * V + k] = s[k];
  }
  __syncthreads();
  if (ry == 0) {
#pragma unroll
    for (int k = 0; k < V; ++k) {
      float g = part[0][vx * V + k];
#pragma unroll
      for (int w = 1; w < 8; ++w) g += part[w][vx * V + k];
      xm[vx * V + k] = m[k];
      xs[vx * V + k] = g;
    }
  }
  cluster.sync();
  // the cluster's pairs, folded ONCE per column: thread c < 16 * V reads column c's (max, sum) of every rank through DSMEM
  // and leaves the column's (M, S) in local shared memory for the threads that own it.  (Earlier versions: every thread
  // read all ranks with scalar loads — 32768 remote words per CTA, 5.1 ms; then every thread folded its 8 columns x 8 ranks
  // itself — 6 instructions and one extra ex2 per ELEMENT of its 128-element share, 0.72 ms.)
  if (tid < TCV * V) {
    float Mc = -INFINITY, Sc = 0.f;
    for (unsigned r = 0; r < cs; ++r) {
      const float a = cluster.map_shared_rank(xm, r)[tid], b = cluster.map_shared_rank(xs, r)[tid];
      const float nm = fmaxf(Mc, a);
      const float f1 = Mc == -INFINITY ? 0.f : ex2_fast((Mc - nm) * kLog2e);
      const float f2 = a == -INFINITY ? 0.f : ex2_fast((a - nm) * kLog2e);
      Sc = fmaf(Sc, f1, b * f2);
      Mc = nm;
    }
    part[0][tid] = Mc;
    part2[0][tid] = Sc;
  }
  __syncthreads();
  float M[V], S[V];
#pragma unroll
  for (int k = 0; k < V; ++k) {
    M[k] = part[0][vx * V + k];
    S[k] = part2[0][vx * V + k];
  }
  // every peer array has been read: arrive for the exit barrier (relaxed), wait for it after the stores
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  float inv[V], floor_v[V];
#pragma unroll
  for (int k = 0; k < V; ++k) {
    if (M[k] == -INFINITY) S[k] = __int_as_float(0x7fc00000);  // a column of -inf: NaN, as inf - inf gives in the CPU kernel
    inv[k] = row_constant<LOG, float>(S[k]);
    floor_v[k] = kClip * inv[k];
    nml[k] = -M[k] * kLog2e;
    if (stats != nullptr && rank == 0 && ry == 0 && active) {
      stats[2 * (o * inner + col + k)] = M[k];
      stats[2 * (o * inner + col + k) + 1] = logf(S[k]);
    }
  }
  if (active) {
    const F2 l2e = f2_splat(kLog2e);
    F2 n2[V / 2], i2[V / 2];
#pragma unroll
    for (int j = 0; j < V / 2; ++j) {
      n2[j] = LOG ? f2_make(-M[2 * j], -M[2 * j + 1]) : f2_make(nml[2 * j], nml[2 * j + 1]);
      i2[j] = LOG ? f2_make(-inv[2 * j], -inv[2 * j + 1]) : f2_make(inv[2 * j], inv[2 * j + 1]);
    }
    Vec<T> floor_raw;
    pack_from_float<T>(floor_v, floor_raw);  // the -64 clip on the rounded output (rounding is monotone)
#pragma unroll 2
    for (int r = ry; r < nrows; r += RL) {
      float v[V];
      unpack_to_float<T>(lds_vec(tile + ((size_t)r * TCV + vx) * V), v);
#pragma unroll
      for (int j = 0; j < V / 2; ++j) {
        F2 q;
        if constexpr (LOG) {
          const F2 d = f2_add(f2_make(v[2 * j], v[2 * j + 1]), n2[j]);
          q = f2_add(f2_make(max_nan(f2_lo(d), -64.0f), max_nan(f2_hi(d), -64.0f)), i2[j]);
        } else {
          const F2 t = f2_fma(f2_make(v[2 * j], v[2 * j + 1]), l2e, n2[j]);
          q = f2_mul(f2_make(ex2_fast(f2_lo(t)), ex2_fast(f2_hi(t))), i2[j]);
        }
        v[2 * j] = f2_lo(q);
        v[2 * j + 1] = f2_hi(q);
      }
      Vec<T> ov;
      pack_from_float<T>(v, ov);
      if constexpr (!LOG) vclip_raw(ov, floor_raw);
      st_stream(dst + (int64_t)r * inner, ov);
    }
  }
  asm volatile("barrier.cluster.wait.aligned;" ::: "memory");  // nobody leaves while a peer may still be reading its arrays
}

// Returns < 0 when the shape does not suit the kernel (the caller falls back to softmax_cols_tile_kernel).
template <typename T, bool LOG>
int try_softmax_cols_cluster(const T* x, T* y, int64_t outer, int64_t dim, int64_t inner, float* stats, cudaStream_t st) {
  constexpr int VN = Vec<T>::N;
  constexpr int TCV = 16;
  const size_t total = (size_t)outer * dim * inner * sizeof(T);
  // small tensors are served out of L2 by the two-sweep kernel; short axes leave the CTAs too little to do
  if (total <= ((size_t)96 << 20) || dim < 256 || dim > 8 * 256 || inner / VN < TCV) return -1;
  int cs = 1;
  while (cs < 8 && (dim + cs - 1) / cs > 256) cs <<= 1;
  const int tr = (int)((dim + cs - 1) / cs);
  if (tr > 256) return -1;
  const int64_t ctiles = (inner / VN + TCV - 1) / TCV;
  const int64_t nblk = ctiles * outer * cs;
  if (nblk > 0x7fffffffLL) return -1;
  const size_t smem = (size_t)tr * TCV * 16;
  auto kern = softmax_cols_cluster_kernel<T, LOG>;
  FuncAttrHold hold;
  if (!func_attr_apply(reinterpret_cast<const void*>(kern), (int)(256 * TCV * 16), 100, &hold)) {
    cudaGetLastError();
    return -1;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)nblk);
  cfg.blockDim = dim3(256);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cs;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (cudaLaunchKernelEx(&cfg, kern, x, y, dim, inner, ctiles, tr, stats) != cudaSuccess) {
    cudaGetLastError();
    return -1;
  }
  return after_launch("softmax_cols_cluster");
}

template <typename T, bool LOG>
int softmax_typed(const T* x, T* y, int64_t outer, int64_t dim, int64_t inner, typename AccType<T>::type* stats,
                  cudaStream_t st) {
  using A = typename AccType<T>::type;
  constexpr int VN = Vec<T>::N;
  if (inner == 1) {
    const bool vec = (dim % VN == 0) && is_aligned16(x) && is_aligned16(y);
    const int V = vec ? VN : 1;
    const int64_t nchunk = dim / V;
    if (vec && nchunk >= 64 && nchunk <= 512 * 4 && outer >= 4 * (int64_t)sm_count()) {
      const int rc = try_softmax_pipe<T, LOG>(x, y, outer, dim, stats, st);
      if (rc >= 0) return rc;  // < 0: the pipeline does not fit; use the register route
    }
    if (nchunk <= 32 * 16) {
      int G = 1;
      while (G < 32 && G < nchunk) G <<= 1;
      int C = 1;
      while ((int64_t)C * G < nchunk) C <<= 1;
      return vec ? launch_rows_reg<T, true, LOG>(x, y, outer, (int)dim, G, C, stats, st)
                 : launch_rows_reg<T, false, LOG>(x, y, outer, (int)dim, G, C, stats, st);
    }
    if (vec) {
      // long rows: shared-memory-resident slices across a cluster (one read, rows up to 8 x 200 KB) or two sweeps with
      // the second out of L2 (any length); PDB200_SOFTMAX_LONG = 1 / 2 forces the first / second (3: both, without the one-exp kernel)
      // PDB200_SOFTMAX_LONG forces a route: 1 cluster (lean kernel), 2 two-sweep, 5 cluster (first kernel), 6 relay
      const int pick = knobs().softmax_long;
      int rc = -1;
      if constexpr (std::is_same<A, float>::value) {
        // the relay kernel first for large tensors of 4-byte elements (measured: profiles/r2_notes.txt section 5b; its
        // record array, memset and cooperative launch cost ~15 us per call, so small tensors stay with the routes below);
        // 16-bit rows that fit a cluster take the lean cluster kernel, longer ones the two-sweep kernel
        const bool big = (size_t)outer * (size_t)dim * sizeof(T) >= ((size_t)128 << 20);
        if (pick == 6 || (pick == 0 && sizeof(T) == 4 && big)) rc = try_softmax_relay<T, LOG>(x, y, outer, dim, stats, st);
      }
      if (rc < 0 && pick != 2 && pick != 6) rc = try_softmax_cluster<T, LOG>(x, y, outer, dim, stats, st);
      if (rc < 0 && pick != 1 && pick != 5 && pick != 6) rc = try_softmax_sweep2<T, LOG>(x, y, outer, dim, stats, st);
      if (rc >= 0) return rc;
    }
    if (outer > 0x7fffffffLL) return fail(PDB200_UNSUPPORTED, "softmax: too many rows for one launch");
    const size_t smem = (size_t)dim * sizeof(A);
    if (smem <= 200 * 1024) {
      int threads = 256;
      while (threads < 1024 && (int64_t)threads * V * 8 < dim) threads <<= 1;
      if (vec) {
        cudaFuncSetAttribute(softmax_rows_smem_kernel<T, true, LOG>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        softmax_rows_smem_kernel<T, true, LOG><<<(unsigned)outer, threads, smem, st>>>(x, y, dim, stats);
      } else {
        cudaFuncSetAttribute(softmax_rows_smem_kernel<T, false, LOG>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        softmax_rows_smem_kernel<T, false, LOG><<<(unsigned)outer, threads, smem, st>>>(x, y, dim, stats);
      }
      return after_launch("softmax_rows_smem");
    }
    softmax_rows_stream_kernel<T, LOG><<<(unsigned)outer, 1024, 0, st>>>(x, y, dim, stats);
    return after_launch("softmax_rows_stream");
  }
  if constexpr (!std::is_same<T, double>::value) {
    // vector route: 16-byte loads along `inner`; in place is fine (sweep 2 reads and writes each chunk in one thread, after
    // sweep 1 of the whole CTA tile is done)
    if (inner % VN == 0 && is_aligned16(x) && is_aligned16(y) && dim >= 2) {
      if (!knobs().no_cols_cluster) {
        const int rc = try_softmax_cols_cluster<T, LOG>(x, y, outer, dim, inner, stats, st);
        if (rc >= 0) return rc;
      }
      const int64_t ctiles = (inner / VN + 31) / 32;
      const int64_t nblk = ctiles * outer;
      if (nblk <= 0x7fffffffLL) {
        // more thread-rows when there are few column tiles to fill the machine with
        int TY = 8;
        while (TY > 1 && TY * 2 > dim) TY >>= 1;
        const size_t smem = (size_t)TY * 32 * VN * 3 * sizeof(float);
        dim3 block(32, TY);
        softmax_cols_tile_kernel<T, LOG><<<(unsigned)nblk, block, smem, st>>>(x, y, dim, inner, ctiles, stats);
        return after_launch("softmax_cols_tile");
      }
    }
  }
  int threads = 32;
  while (threads < 256 && threads < inner) threads <<= 1;
  const int64_t col_tiles = (inner + threads - 1) / threads;
  const int64_t blocks = col_tiles * outer;
  if (blocks > 0x7fffffffLL) return fail(PDB200_UNSUPPORTED, "softmax: too many tiles for one launch");
  softmax_cols_kernel<T, LOG><<<(unsigned)blocks, threads, 0, st>>>(x, y, dim, inner, col_tiles, stats);
  return after_launch("softmax_cols");
}

template <bool LOG>
int softmax_entry(const void* x, const int64_t* dims, int rank, int axis, int dtype, void* out, void* stats, void* stream) {
  PDB_CHECK_ARG(rank >= 0 && rank <= PDB200_MAX_RANK, "softmax: rank must be in [0, 9]");
  PDB_CHECK_ARG(rank == 0 || dims != nullptr, "softmax: dims is NULL");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  // CanonicalAxis (funcs/axis_utils.h): axis in [-rank, rank)
  if (rank > 0) {
    PDB_CHECK_ARG(axis >= -rank && axis < rank,
                  "The axis is expected to be in range of [" + std::to_string(-rank) + ", " +
                      std::to_string(rank) + "), but got " + std::to_string(axis));
    if (axis < 0) axis += rank;
  }
  int64_t outer = 1, dim = 1, inner = 1;
  for (int i = 0; i < rank; ++i) {
    PDB_CHECK_ARG(dims[i] >= 0, "softmax: negative dimension");
    if (i < axis) outer *= dims[i];
    else if (i == axis) dim = dims[i];
    else inner *= dims[i];
  }
  if (rank > 0 && outer * dim * inner == 0) return PDB200_OK;  // 0-size tensor: no-op
  PDB_CHECK_ARG(x && out, "softmax: NULL tensor pointer");
#define PDB_SM_CASE(DT)                                                                       \
  case DT: {                                                                                  \
    using T = typename CType<DT>::type;                                                       \
    if (rank == 0) { /* 0-D tensor: softmax 1.0 (impl/softmax_kernel_impl.h:41-44), log_softmax 0.0 */ \
      softmax_fill_one_kernel<T><<<1, 1, 0, st>>>(static_cast<T*>(out), LOG ? 0.f : 1.f);     \
      return after_launch("softmax_fill_one");                                                \
    }                                                                                         \
    return softmax_typed<T, LOG>(static_cast<const T*>(x), static_cast<T*>(out), outer, dim, inner,          \
                                 static_cast<typename AccType<T>::type*>(stats), st);                       \
  }
  switch (dtype) {
    PDB_SM_CASE(PDB200_FLOAT32)
    PDB_SM_CASE(PDB200_FLOAT64)
    PDB_SM_CASE(PDB200_FLOAT16)
    PDB_SM_CASE(PDB200_BFLOAT16)
    default:
      return fail(PDB200_UNSUPPORTED, "softmax: dtype must be float32/float64/float16/bfloat16, got code " +
                                          std::to_string(dtype));
  }
#undef PDB_SM_CASE
}
}  // namespace pdb200

namespace pdb200 {
// softmax + row statistics for cross_entropy.cu (same library, not part of the C-ABI): stats = [positions][2] of the
// accumulator type (float, or double for float64 input)
int softmax_with_stats(const void* x, const int64_t* dims, int rank, int axis, int dtype, void* out, void* stats, void* stream) {
  return softmax_entry<false>(x, dims, rank, axis, dtype, out, stats, stream);
}
}  // namespace pdb200

extern "C" int pdb200_softmax(const void* x, const int64_t* dims, int rank, int axis, int dtype, void* out,
                              void* stream) {
  return pdb200::softmax_entry<false>(x, dims, rank, axis, dtype, out, nullptr, stream);
}
extern "C" int pdb200_log_softmax(const void* x, const int64_t* dims, int rank, int axis, int dtype, void* out,
                                  void* stream) {
  return pdb200::softmax_entry<true>(x, dims, rank, axis, dtype, out, nullptr, stream);
}
