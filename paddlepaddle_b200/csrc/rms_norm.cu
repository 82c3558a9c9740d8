// rms_norm.cu — rms_norm forward, with the optional fused "bias + residual" add in front.
//
// Replaces (fresh design): phi::RmsNormKernel<T, GPUContext> without quantisation
//   paddle/phi/kernels/gpu/rms_norm_kernel.cu:1080-1260 (ResidualAddBiasLoad :868-905, AffineStore :925-962,
//   RmsNormWarpImpl / BlockSMemImpl :276-700), python: paddle.incubate.nn.functional.fused_rms_norm.
//     v = x (+ residual + bias)            in float; residual_out = T(v) when a residual is given
//     inv_var = rsqrt(mean(v^2) + eps)     float, one value per row
//     out = T((v * inv_var) * norm_weight + norm_bias)
// The reference has no CPU kernel for this op; numerics are anchored on its own test's formula
// (test/legacy_test/test_rms_norm_op.py:37-47,74-79).
//
// Routes ([rows][D] view):
//   pipe     : many rows, D = 64..2048 16-byte chunks — persistent CTAs on the TMA row pipeline (rowpipe.cuh); a stage
//              holds the x row and, with a residual, the residual row: every input byte is read from HBM once; one
//              CTA reduction (sum of squares) per row.
//   fallback : any shape — one CTA per row, two sweeps (the second from L2).
#include <string>

#include "common.cuh"
#include "rowpipe.cuh"
#include "shape.h"

namespace pdb200 {

struct AddRms { __device__ __forceinline__ float operator()(float a, float b) const { return a + b; } };

template <typename T, int C, bool EXACT, bool RESID, int MAXT>
__global__ void __launch_bounds__(MAXT, MAXT <= 256 ? 2 : 1) rms_pipe_kernel(const T* __restrict__ x, const T* __restrict__ residual,
                                                       const T* __restrict__ bias, const T* __restrict__ weight,
                                                       const T* __restrict__ nbias, T* __restrict__ out,
                                                       T* __restrict__ residual_out, float* __restrict__ inv_var,
                                                       int64_t rows, int D, int stages, float eps, float inv_d) {
  constexpr int V = Vec<T>::N;
  constexpr int RS = RESID ? 2 : 1;  // rows of input per stage
  extern __shared__ __align__(128) unsigned char ring_raw[];
  T* ring = reinterpret_cast<T*>(ring_raw);
  __shared__ uint64_t full[kPipeMaxStages];
  __shared__ float red[2][32];
  const int tid = threadIdx.x, nt = blockDim.x;
  const uint32_t row_bytes = (uint32_t)D * sizeof(T);
  const int64_t row_step = gridDim.x;
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) mbar_init(&full[s], 1);
    mbar_fence_init();
  }
  block_reduce_1sync_init(red[0], 0.f);
  block_reduce_1sync_init(red[1], 0.f);
  __syncthreads();
  auto fill = [&](int s, int64_t r) {
    mbar_expect_tx(&full[s], RS * row_bytes);
    bulk_load_row(ring + (size_t)s * RS * D, x + r * D, row_bytes, &full[s]);
    if constexpr (RESID) bulk_load_row(ring + (size_t)s * RS * D + D, residual + r * D, row_bytes, &full[s]);
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
    const T* sx = ring + (size_t)s * RS * D;
    float v[C * V];
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    Vec<T> raw[C], wv[C];  // the staged row first (all loads in flight), and the weights: their L1 latency hides behind
#pragma unroll             // the CTA reduction instead of stalling the output loop
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + tid) * V;
      if (EXACT || off < D) {
        raw[c] = lds_vec(sx + off);
        wv[c] = ld_cached(weight + off);
      }
    }
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + tid) * V;
      if (EXACT || off < D) {
        unpack_to_float<T>(raw[c], &v[c * V]);
        if constexpr (RESID) {
          float r[V];
          unpack_to_float<T>(lds_vec(sx + D + off), r);
          if (bias != nullptr) {
            float b[V];
            unpack_to_float<T>(ld_cached(bias + off), b);
#pragma unroll
            for (int k = 0; k < V; ++k) v[c * V + k] = v[c * V + k] + r[k] + b[k];  // x + residual + bias, left to right
          } else {
#pragma unroll
            for (int k = 0; k < V; ++k) v[c * V + k] += r[k];
          }
          Vec<T> ro;
          pack_from_float<T>(&v[c * V], ro);
          st_stream(residual_out + row * D + off, ro);
        }
      } else {
#pragma unroll
        for (int k = 0; k < V; ++k) v[c * V + k] = 0.f;
      }
#pragma unroll
      for (int k = 0; k < V; ++k) acc[k & 3] = fmaf(v[c * V + k], v[c * V + k], acc[k & 3]);
    }
    float ss = (acc[0] + acc[1]) + (acc[2] + acc[3]);
    ss = block_reduce_1sync(ss, AddRms(), red[flip]);  // its barrier also means: stage s fully read
    flip ^= 1;
    if (tid == 0) {
      const int64_t nrow = row + (int64_t)stages * row_step;
      if (nrow < rows) {
        fence_proxy_async_smem();
        fill(s, nrow);
      }
    }
    const float inv = rsqrtf(ss * inv_d + eps);
    if (inv_var != nullptr && tid == 0) inv_var[row] = inv;
    T* dst = out + row * D;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + tid) * V;
      if (EXACT || off < D) {
        float w[V], o[V];
        unpack_to_float<T>(wv[c], w);
        if (nbias != nullptr) {
          float b[V];
          unpack_to_float<T>(ld_cached(nbias + off), b);
#pragma unroll
          for (int k = 0; k < V; ++k) o[k] = fmaf(v[c * V + k] * inv, w[k], b[k]);
        } else {
#pragma unroll
          for (int k = 0; k < V; ++k) o[k] = (v[c * V + k] * inv) * w[k];
        }
        Vec<T> ov;
        pack_from_float<T>(o, ov);
        st_stream(dst + off, ov);
      }
    }
    if (++s == stages) {
      s = 0;
      phase ^= 1;
    }
  }
}

// any shape: one CTA per row, two sweeps
template <typename T>
__global__ void __launch_bounds__(1024) rms_rows_kernel(const T* x, const T* residual, const T* bias, const T* weight,
                                                        const T* nbias, T* out, T* residual_out, float* inv_var,
                                                        int64_t D, float eps, double inv_d) {
  __shared__ float red[33];
  const int64_t base = (int64_t)blockIdx.x * D;
  auto value = [&](int64_t i) {
    float v = cvt<float>(x[base + i]);
    if (residual != nullptr) {
      v += cvt<float>(residual[base + i]);
      if (bias != nullptr) v += cvt<float>(bias[i]);
    }
    return v;
  };
  float ss = 0.f;
  for (int64_t i = threadIdx.x; i < D; i += blockDim.x) {
    const float v = value(i);
    ss = fmaf(v, v, ss);
  }
  ss = block_reduce(ss, AddRms(), 0.f, red);
  const float inv = rsqrtf((float)((double)ss * inv_d) + eps);
  if (inv_var != nullptr && threadIdx.x == 0) inv_var[blockIdx.x] = inv;
  for (int64_t i = threadIdx.x; i < D; i += blockDim.x) {
    const float v = value(i);
    if (residual != nullptr) residual_out[base + i] = cvt<T>(v);  // written after the last read of x / residual at i
    const float w = cvt<float>(weight[i]);
    out[base + i] = cvt<T>(nbias != nullptr ? fmaf(v * inv, w, cvt<float>(nbias[i])) : (v * inv) * w);
  }
}

}  // namespace pdb200

namespace pdb200 {

template <typename T>
int rms_typed(const T* x, const T* residual, const T* bias, const T* weight, const T* nbias, T* out, T* residual_out,
              float* inv_var, int64_t rows, int64_t D, float eps, cudaStream_t st) {
  constexpr int VN = Vec<T>::N;
  bool vec = (D % VN == 0) && is_aligned16(x) && is_aligned16(out) && is_aligned16(weight) && (!nbias || is_aligned16(nbias));
  if (residual) vec = vec && is_aligned16(residual) && is_aligned16(residual_out) && (!bias || is_aligned16(bias));
  const int64_t nchunk = D / VN;
  if (vec && nchunk >= 64 && nchunk <= 512 * 4 && rows >= 4 * (int64_t)sm_count()) {
    // 32 values a thread for 16-bit rows down to one-warp CTAs, as in layer_norm.cu (short rows: profiles/r2_notes.txt section 13)
    int C = sizeof(T) == 2 ? (nchunk >= 128 ? 4 : (nchunk >= 64 ? 2 : 1)) : (nchunk >= 512 ? 4 : (nchunk >= 256 ? 2 : 1));
    if (const int o = knobs().rms_pipe_c) C = o;  // tuning knob
    while (C < 4 && (nchunk + C - 1) / C > 512) C <<= 1;
    int threads = (int)((nchunk + C - 1) / C);
    threads = (threads + 31) / 32 * 32;
    const bool exact = (int64_t)threads * C == nchunk;
    const size_t stage_bytes = (residual ? 2 : 1) * (size_t)D * sizeof(T);
    auto launch = [&](auto kern) -> int {
      int per_sm = 0, stages = 0;
      size_t smem = 0;
      int want = 768 / threads;
      want = want < 2 ? 2 : (want > 8 ? 8 : want);
      // ~64 KB of rows in flight per SM at two stages: with a residual (32 KB stages) that is ONE CTA per SM
      // (bf16 [65536,8192] + residual: 0.698 ms with two CTAs, 0.640 ms with one)
      const int by_bytes = (int)((64 * 1024) / (2 * stage_bytes));
      if (want > by_bytes) want = by_bytes < 1 ? 1 : by_bytes;
      FuncAttrHold hold;  // released when this lambda returns, i.e. after the launch
      if (!pipe_geometry(kern, threads, stage_bytes, want, 3, &per_sm, &stages, &smem, &hold)) return -1;
      int64_t grid = (int64_t)per_sm * sm_count();
      if (threads <= 64) grid = std::max<int64_t>(grid, rows / 16);  // short rows: short-lived CTAs of 16 rows
      if (grid > rows) grid = rows;
      kern<<<(unsigned)grid, threads, smem, st>>>(x, residual, bias, weight, nbias, out, residual_out, inv_var, rows, (int)D,
                                                  stages, eps, (float)(1.0 / (double)D));
      return after_launch("rms_pipe");
    };
    int rc = -1;
#define PDB_RMS2(CC, MT)                                                                                               \
  if (residual) rc = exact ? launch(rms_pipe_kernel<T, CC, true, true, MT>) : launch(rms_pipe_kernel<T, CC, false, true, MT>); \
  else rc = exact ? launch(rms_pipe_kernel<T, CC, true, false, MT>) : launch(rms_pipe_kernel<T, CC, false, false, MT>);
#define PDB_RMS(CC) if (threads <= 256) { PDB_RMS2(CC, 256) } else { PDB_RMS2(CC, 512) }
    if (threads <= 512) {
      if (C == 4) { PDB_RMS(4) } else if (C == 2) { PDB_RMS(2) } else { PDB_RMS(1) }
    }
#undef PDB_RMS2
#undef PDB_RMS
    if (rc >= 0) return rc;  // < 0: the pipeline does not fit, use the fallback
  }
  if (rows > 0x7fffffffLL) return fail(PDB200_UNSUPPORTED, "rms_norm: too many rows for one launch");
  int threads = 128;
  while (threads < 1024 && threads * 8 < D) threads <<= 1;
  rms_rows_kernel<T><<<(unsigned)rows, threads, 0, st>>>(x, residual, bias, weight, nbias, out, residual_out, inv_var, D, eps,
                                                         1.0 / (double)D);
  return after_launch("rms_rows");
}

}  // namespace pdb200

extern "C" int pdb200_rms_norm(const void* x, const int64_t* dims, int rank, const void* bias, const void* residual,
                               const void* norm_weight, const void* norm_bias, float epsilon, int begin_norm_axis, int dtype,
                               void* out, void* residual_out, void* inv_var, void* stream) {
  using namespace pdb200;
  PDB_CHECK_ARG(rank >= 1 && rank <= PDB200_MAX_RANK, "rms_norm: rank must be in [1, 9]");
  PDB_CHECK_ARG(dims != nullptr, "rms_norm: dims is NULL");
  PDB_CHECK_ARG(begin_norm_axis >= 0 && begin_norm_axis <= rank, "rms_norm: begin_norm_axis out of range");
  int64_t rows = 1, D = 1;
  for (int i = 0; i < rank; ++i) {
    PDB_CHECK_ARG(dims[i] >= 0, "rms_norm: negative dimension");
    if (i < begin_norm_axis) rows *= dims[i];
    else D *= dims[i];
  }
  if (rows == 0 || D == 0) return PDB200_OK;
  PDB_CHECK_ARG(x && norm_weight && out, "rms_norm: NULL tensor pointer");
  PDB_CHECK_ARG(residual == nullptr || residual_out != nullptr, "rms_norm: a residual needs residual_out");
  PDB_CHECK_ARG(bias == nullptr || residual != nullptr, "rms_norm: bias is only added together with a residual (rms_norm_kernel.cu:1144-1150)");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (dtype) {
#define M(DT)                                                                                                          \
  case DT: {                                                                                                           \
    using T = typename CType<DT>::type;                                                                                \
    return rms_typed<T>(static_cast<const T*>(x), static_cast<const T*>(residual), static_cast<const T*>(bias),        \
                        static_cast<const T*>(norm_weight), static_cast<const T*>(norm_bias), static_cast<T*>(out),    \
                        static_cast<T*>(residual_out), static_cast<float*>(inv_var), rows, D, epsilon, st);            \
  }
    M(PDB200_FLOAT32) M(PDB200_FLOAT16) M(PDB200_BFLOAT16)  // rms_norm_kernel.cu:1333-1339
#undef M
  }
  return fail(PDB200_UNSUPPORTED, "rms_norm: dtype must be float32/float16/bfloat16, got code " + std::to_string(dtype));
}
