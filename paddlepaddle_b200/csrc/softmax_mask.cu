// softmax_mask.cu — the two masked softmaxes of attention, hand-written for sm_100a.
//
// Replaces (fresh design):
//   phi::fusion::FusedSoftmaxMaskKernel            paddle/phi/kernels/fusion/gpu/fused_softmax_mask_kernel.cu:26-147,149-316
//       out[b,h,q,:] = softmax(float(x[b,h,q,:]) + float(mask[b,0,q,:]))        x [B,H,Sq,Sk], mask [B,1,Sq,Sk]
//   phi::fusion::FusedSoftmaxMaskFuseUpperTriangleKernel
//                                                  fusion/gpu/fused_softmax_mask_upper_triangle_kernel.cu:27-253,255-400
//       out[b,h,q,k] = softmax over k <= q of x[b,h,q,k];  0 for k > q            x [B,H,S,S]
// python: paddle.incubate.softmax_mask_fuse / softmax_mask_fuse_upper_triangle.  Both are GPU-only in the reference:
// one warp per row, the row in registers, plain exp(d - max) with no clip, e / sum.  Numerics here are the same
// (float throughout, one rounding at the end); the reference's shape limits (32 <= Sk < 8192 resp. <= 16384, Sq a
// multiple of the rows a 128-thread block holds) come from its warp-per-row layout and do not apply.
//
// Routes ([rows][Sk] view, rows = B*H*Sq):
//   pipe : many rows, Sk a multiple of 16 bytes, 64..2048 chunks — persistent CTAs on the TMA row pipeline
//          (rowpipe.cuh).  A stage holds the x row and, for the additive mask, the mask row of the same (b, q): both
//          arrive by cp.async.bulk on one mbarrier.  The mask is shared by the H heads of a batch; consecutive heads
//          re-read it 2*Sq*Sk*sizeof(T) bytes of traffic later, i.e. out of the 126 MB L2, so HBM sees it about once.
//          Upper triangle: the bulk copy fetches only the first q+1 elements of row q (rounded up to 16 bytes) — the
//          bytes above the diagonal are never read — chunks wholly above the diagonal skip the arithmetic, and the
//          row is written whole (zeros above the diagonal): algorithmic bytes ~ (1/2 + 1) instead of 2 tensors.
//   rows : any other shape — one CTA per row, three sweeps (the later ones from L2).
#include <string>

#include "common.cuh"
#include "rowpipe.cuh"
#include "shape.h"

namespace pdb200 {

struct MaxSm { __device__ __forceinline__ float operator()(float a, float b) const { return fmaxf(a, b); } };
struct AddSm { __device__ __forceinline__ float operator()(float a, float b) const { return a + b; } };

// exp(x - m) = 2^(x*log2e - m*log2e): one FFMA + MUFU.EX2; no clip (the reference kernels have none)
__device__ __forceinline__ float exp_shifted(float x, float neg_m_log2e) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(fmaf(x, 1.4426950408889634f, neg_m_log2e)));
  return r;
}

// TRI == false: additive mask of type MT (row index (b*Sq + q): `rows_per_batch` = H*Sq rows share Sq mask rows)
// TRI == true : causal; `mask` unused, row r has q = r % Sq valid leading elements + 1
template <typename T, typename MT, int C, bool EXACT, bool TRI>
__global__ void __launch_bounds__(512) softmax_mask_pipe_kernel(const T* __restrict__ x, const MT* __restrict__ mask,
                                                                T* __restrict__ y, int64_t rows, int dim, int sq, int heads,
                                                                int stages, int contiguous) {
  constexpr int V = Vec<T>::N;
  constexpr int MV = Vec<MT>::N;  // mask elements per 16-byte load (V, or V/2 for a float mask of a 16-bit x)
  extern __shared__ __align__(128) unsigned char ring_raw[];
  __shared__ uint64_t full[kPipeMaxStages];
  __shared__ float red[2][32];
  const int tid = threadIdx.x, nt = blockDim.x;
  const uint32_t x_bytes = (uint32_t)dim * sizeof(T);
  const uint32_t m_bytes = TRI ? 0u : (uint32_t)dim * sizeof(MT);
  const size_t stage_bytes = (size_t)x_bytes + m_bytes;
  // Row -> CTA mapping: strided (row = cta + k*grid: a persistent grid then sweeps the rows in order) or, for short-lived
  // CTAs, one contiguous slab per CTA.  Either way rows are visited in roughly increasing order machine-wide, so the
  // mask rows of head h are re-read by head h+1 about 2*Sq*Sk*sizeof(T) bytes of traffic later — out of L2.  (Strided
  // short-lived CTAs touch four batches at once and push that distance past the L2 size: 1.24 ms instead of 0.66 ms
  // on bf16 [8,32,2048,2048].)
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
  auto fill = [&](int s, int64_t r) {
    unsigned char* st = ring_raw + (size_t)s * stage_bytes;
    if constexpr (TRI) {
      const int q = rows <= 0xffffffffLL ? (int)((uint32_t)r % (uint32_t)sq) : (int)(r % sq);  // 32-bit when it fits
      const uint32_t bytes = (((uint32_t)(q + 1) * sizeof(T)) + 15u) & ~15u;  // <= x_bytes: dim * sizeof(T) is a multiple of 16
      mbar_expect_tx(&full[s], bytes);
      bulk_load_row(st, x + r * dim, bytes, &full[s]);
    } else {
      int64_t mrow;
      if (rows <= 0xffffffffLL) {  // 32-bit index math when it fits (a 64-bit div / mod pair costs ~100 instructions)
        const uint32_t b = (uint32_t)r / ((uint32_t)heads * (uint32_t)sq);
        mrow = (int64_t)b * sq + (uint32_t)r % (uint32_t)sq;
      } else {
        mrow = (r / ((int64_t)heads * sq)) * sq + r % sq;
      }
      mbar_expect_tx(&full[s], x_bytes + m_bytes);
      bulk_load_row(st, x + r * dim, x_bytes, &full[s]);
      bulk_load_row(st + x_bytes, mask + mrow * dim, m_bytes, &full[s]);
    }
  };
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) {
      const int64_t r = row_first + s * row_step;
      if (r < row_end) fill(s, r);
    }
  }
  int s = 0;
  uint32_t phase = 0;
  for (int64_t row = row_first; row < row_end; row += row_step) {
    const int valid = TRI ? (rows <= 0xffffffffLL ? (int)((uint32_t)row % (uint32_t)sq) : (int)(row % sq)) + 1 : dim;  // elements that take part
    mbar_wait(&full[s], phase);
    const T* sx = reinterpret_cast<const T*>(ring_raw + (size_t)s * stage_bytes);
    const MT* sm = reinterpret_cast<const MT*>(ring_raw + (size_t)s * stage_bytes + x_bytes);
    float e[C * V];
    float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + tid) * V;
      if ((EXACT || off < dim) && off < valid) {
        unpack_to_float<T>(lds_vec(sx + off), &e[c * V]);
        if constexpr (!TRI) {
          float mk[V];
#pragma unroll
          for (int j = 0; j < V / MV; ++j) unpack_to_float<MT>(lds_vec(sm + off + j * MV), &mk[j * MV]);
#pragma unroll
          for (int j = 0; j < V / 2; ++j) {
            const F2 t = f2_add(f2_make(e[c * V + 2 * j], e[c * V + 2 * j + 1]), f2_make(mk[2 * j], mk[2 * j + 1]));
            e[c * V + 2 * j] = f2_lo(t);
            e[c * V + 2 * j + 1] = f2_hi(t);
          }
        } else if (off + V > valid) {  // only the vector the diagonal cuts pays for per-element selects
#pragma unroll
          for (int k = 0; k < V; ++k)
            if (off + k >= valid) e[c * V + k] = -INFINITY;  // the tail of the diagonal's vector (stale bytes of the stage)
        }
      } else {
#pragma unroll
        for (int k = 0; k < V; ++k) e[c * V + k] = -INFINITY;
      }
#pragma unroll
      for (int k = 0; k < V; ++k) mx[k & 3] = fmaxf(mx[k & 3], e[c * V + k]);
    }
    float m = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
    m = block_reduce_1sync(m, MaxSm(), red[0]);  // its barrier also means: stage s fully read
    if (tid == 0) {
      const int64_t nrow = row + (int64_t)stages * row_step;
      if (nrow < row_end) {
        fence_proxy_async_smem();
        fill(s, nrow);
      }
    }
    // exp, sum and the final scaling on packed fp32 pairs (FFMA2 / FADD2 / FMUL2: the same roundings, half the issue slots)
    const F2 nml2 = f2_splat(-m * 1.4426950408889634f), l2e2 = f2_splat(1.4426950408889634f);
    F2 sm2[2] = {f2_splat(0.f), f2_splat(0.f)};
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + tid) * V;
      if ((EXACT || off < dim) && off < valid) {  // CTA-uniform per chunk only in the mask form; warp-divergent at the diagonal
#pragma unroll
        for (int j = 0; j < V / 2; ++j) {
          const F2 t = f2_fma(f2_make(e[c * V + 2 * j], e[c * V + 2 * j + 1]), l2e2, nml2);  // exp(-inf - m) = 0 for the masked tail
          float p0, p1;
          asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p0) : "f"(f2_lo(t)));
          asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p1) : "f"(f2_hi(t)));
          e[c * V + 2 * j] = p0;
          e[c * V + 2 * j + 1] = p1;
          sm2[j & 1] = f2_add(sm2[j & 1], f2_make(p0, p1));
        }
      } else {
#pragma unroll
        for (int k = 0; k < V; ++k) e[c * V + k] = 0.f;
      }
    }
    const F2 s2 = f2_add(sm2[0], sm2[1]);
    float sum = f2_lo(s2) + f2_hi(s2);
    sum = block_reduce_1sync(sum, AddSm(), red[1]);
    const float inv = 1.0f / sum;
    const F2 inv2 = f2_splat(inv);
    T* dst = y + row * dim;
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int off = (c * nt + tid) * V;
      if (EXACT || off < dim) {
        float r[V];
#pragma unroll
        for (int j = 0; j < V / 2; ++j) {
          const F2 q = f2_mul(f2_make(e[c * V + 2 * j], e[c * V + 2 * j + 1]), inv2);
          r[2 * j] = f2_lo(q);
          r[2 * j + 1] = f2_hi(q);
        }
        if constexpr (TRI) {
          if (off + V > valid) {  // the diagonal's vector and everything after it
#pragma unroll
            for (int k = 0; k < V; ++k)
              if (off + k >= valid) r[k] = 0.f;  // exact zeros above the diagonal (0 * (1/sum) is NaN for a NaN / inf sum)
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

// any shape: one CTA per row, three sweeps over global memory (the row stays in L1/L2 between them)
template <typename T, typename MT, bool TRI>
__global__ void __launch_bounds__(1024) softmax_mask_rows_kernel(const T* x, const MT* mask, T* y, int64_t dim, int64_t sq,
                                                                 int64_t heads) {
  __shared__ float red[33];
  const int64_t row = blockIdx.x;
  const int64_t valid = TRI ? row % sq + 1 : dim;
  const T* src = x + row * dim;
  const MT* mrow = TRI ? nullptr : mask + ((row / (heads * sq)) * sq + row % sq) * dim;
  T* dst = y + row * dim;
  auto value = [&](int64_t i) {
    float v = cvt<float>(src[i]);
    if constexpr (!TRI) v += cvt<float>(mrow[i]);
    return v;
  };
  float m = -INFINITY;
  for (int64_t i = threadIdx.x; i < valid; i += blockDim.x) m = fmaxf(m, value(i));
  m = block_reduce(m, MaxSm(), -INFINITY, red);
  float sum = 0.f;
  for (int64_t i = threadIdx.x; i < valid; i += blockDim.x) sum += expf(value(i) - m);
  sum = block_reduce(sum, AddSm(), 0.f, red);
  // in-place safe: each element is read and written by the same thread in this sweep
  for (int64_t i = threadIdx.x; i < dim; i += blockDim.x) dst[i] = cvt<T>(i < valid ? expf(value(i) - m) / sum : 0.f);
}

template <typename T, typename MT, bool TRI>
int softmax_mask_typed(const T* x, const MT* mask, T* y, int64_t batch, int64_t heads, int64_t sq, int64_t sk, cudaStream_t st) {
  constexpr int VN = Vec<T>::N;
  const int64_t rows = batch * heads * sq;
  const int64_t nchunk = sk / VN;
  bool vec = (sk % VN == 0) && is_aligned16(x) && is_aligned16(y) && y != x;
  if (!TRI) vec = vec && is_aligned16(mask) && (sk * sizeof(MT)) % 16 == 0;
  if (vec && nchunk >= 32 && nchunk <= 512 * 4 && rows >= 4 * (int64_t)sm_count() && sq < (int64_t(1) << 31) &&
      heads < (int64_t(1) << 31)) {
    // chunks per thread as in softmax.cu: 16-bit rows 4 (32 values a thread), fp32 rows 2
    int C = sizeof(T) == 2 ? (nchunk >= 256 ? 4 : (nchunk >= 128 ? 2 : 1)) : (nchunk >= 256 ? 2 : 1);
    // (profiles/r2_notes.txt: bf16 [4,32,2048,2048] + mask, C x KB of rows in flight per SM -> ms: 1 x 64 0.78, 2 x 64 0.58,
    //  2 x 128 0.46, 4 x 64 0.65, 4 x 128 0.447: the x + mask stages want twice the bytes in flight of the plain softmax)
    if (const int o = knobs().softmax_pipe_c) C = o;  // tuning knob
    while (C < 8 && (nchunk + C - 1) / C > 512) C <<= 1;
    int threads = (int)((nchunk + C - 1) / C);
    threads = (threads + 31) / 32 * 32;
    const bool exact = (int64_t)threads * C == nchunk;
    const size_t stage_bytes = (size_t)sk * sizeof(T) + (TRI ? 0 : (size_t)sk * sizeof(MT));
    auto launch = [&](auto kern) -> int {
      int per_sm = 0, stages = 0;
      size_t smem = 0;
      int want = 768 / threads;
      want = want < 2 ? 2 : (want > 8 ? 8 : want);
      const int kb = knobs().pipe_inflight_kb >= 8 ? knobs().pipe_inflight_kb : (TRI ? 64 : 128);
      const int by_bytes = (int)(((size_t)kb * 1024) / (2 * stage_bytes));  // ~64 KB (x only) / ~128 KB (x + mask) in flight per SM
      if (want > by_bytes) want = by_bytes < 1 ? 1 : by_bytes;
      FuncAttrHold hold;  // released when this lambda returns, i.e. after the launch
      if (!pipe_geometry(kern, threads, stage_bytes, want, 3, &per_sm, &stages, &smem, &hold)) return -1;
      int64_t grid = (int64_t)per_sm * sm_count();
      int contiguous = 0;
      // rows per short-lived CTA: 4 (mask) / 16 (triangle: 0.379 -> 0.355 ms on bf16 [4,32,2048,2048]; 2 rows 0.408)
      int rows_each = knobs().pipe_rows_per_cta >= 1 ? knobs().pipe_rows_per_cta : (TRI ? 16 : 4);
      if (threads <= 64 || knobs().pipe_rows_per_cta >= 1) {  // short-lived CTAs for narrow rows (softmax.cu), each a contiguous slab
        grid = std::max<int64_t>(grid, rows / rows_each);
        contiguous = 1;
      }
      if (grid > rows) grid = rows;
      if (grid > 0x7fffffffLL) grid = 0x7fffffffLL;
      kern<<<(unsigned)grid, threads, smem, st>>>(x, mask, y, rows, (int)sk, (int)sq, (int)heads, stages, contiguous);
      return after_launch("softmax_mask_pipe");
    };
    int rc = -1;
#define PDB_SMM(CC) rc = exact ? launch(softmax_mask_pipe_kernel<T, MT, CC, true, TRI>) : launch(softmax_mask_pipe_kernel<T, MT, CC, false, TRI>);
    if (threads <= 512) {
      if (C == 8) { PDB_SMM(8) } else if (C == 4) { PDB_SMM(4) } else if (C == 2) { PDB_SMM(2) } else { PDB_SMM(1) }
    }
#undef PDB_SMM
    if (rc >= 0) return rc;  // < 0: the pipeline does not fit; use the row kernel
  }
  if (rows > 0x7fffffffLL) return fail(PDB200_UNSUPPORTED, "softmax_mask: too many rows for one launch");
  int threads = 128;
  while (threads < 1024 && threads * 8 < sk) threads <<= 1;
  softmax_mask_rows_kernel<T, MT, TRI><<<(unsigned)rows, threads, 0, st>>>(x, mask, y, sk, sq, heads);
  return after_launch("softmax_mask_rows");
}

}  // namespace pdb200

extern "C" int pdb200_fused_softmax_mask(const void* x, const int64_t* x_dims, const void* mask, const int64_t* mask_dims,
                                         int dtype, int mask_dtype, void* out, void* stream) {
  using namespace pdb200;
  PDB_CHECK_ARG(x_dims && mask_dims, "fused_softmax_mask: dims is NULL (x and mask are 4-D)");
  for (int i = 0; i < 4; ++i) PDB_CHECK_ARG(x_dims[i] >= 0 && mask_dims[i] >= 0, "fused_softmax_mask: negative dimension");
  // fused_softmax_mask_kernel.cu:177-202
  PDB_CHECK_ARG(mask_dims[1] == 1, "Input mask's second dim must be 1 received the second dimension of mask is " +
                                       std::to_string(mask_dims[1]));
  for (int i = 0; i < 4; ++i) {
    if (i == 1) continue;
    PDB_CHECK_ARG(x_dims[i] == mask_dims[i], "Input x's " + std::to_string(i) + "th dim should be equal with input mask's " +
                                                 std::to_string(i) + "th dim but the " + std::to_string(i) + "th dim of x is " +
                                                 std::to_string(x_dims[i]) + ", while the dim of mask is " + std::to_string(mask_dims[i]));
  }
  if (x_dims[0] * x_dims[1] * x_dims[2] * x_dims[3] == 0) return PDB200_OK;
  PDB_CHECK_ARG(x && mask && out, "fused_softmax_mask: NULL tensor pointer");
  PDB_CHECK_ARG(mask_dtype == dtype || mask_dtype == PDB200_FLOAT32,
                "fused_softmax_mask: the mask dtype must be x's or float32 (fused_softmax_mask_kernel.cu:228-270)");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int64_t B = x_dims[0], H = x_dims[1], Sq = x_dims[2], Sk = x_dims[3];
  switch (dtype) {
#define M(DT)                                                                                                              \
  case DT: {                                                                                                               \
    using T = typename CType<DT>::type;                                                                                    \
    if (mask_dtype == dtype)                                                                                               \
      return softmax_mask_typed<T, T, false>(static_cast<const T*>(x), static_cast<const T*>(mask), static_cast<T*>(out), B, H, Sq, \
                                             Sk, st);                                                                      \
    return softmax_mask_typed<T, float, false>(static_cast<const T*>(x), static_cast<const float*>(mask), static_cast<T*>(out), B, \
                                               H, Sq, Sk, st);                                                             \
  }
    M(PDB200_FLOAT32) M(PDB200_FLOAT16) M(PDB200_BFLOAT16)
#undef M
  }
  return fail(PDB200_UNSUPPORTED, "fused_softmax_mask: dtype must be float32/float16/bfloat16, got code " + std::to_string(dtype));
}

extern "C" int pdb200_fused_softmax_mask_upper_triangle(const void* x, const int64_t* x_dims, int dtype, void* out, void* stream) {
  using namespace pdb200;
  PDB_CHECK_ARG(x_dims, "fused_softmax_mask_upper_triangle: dims is NULL (x is 4-D)");
  for (int i = 0; i < 4; ++i) PDB_CHECK_ARG(x_dims[i] >= 0, "fused_softmax_mask_upper_triangle: negative dimension");
  // fused_softmax_mask_upper_triangle_kernel.cu:271-277
  PDB_CHECK_ARG(x_dims[2] == x_dims[3], "Key seq len must be equal with query seq len received key len: " +
                                            std::to_string(x_dims[3]) + ", query len: " + std::to_string(x_dims[2]));
  if (x_dims[0] * x_dims[1] * x_dims[2] * x_dims[3] == 0) return PDB200_OK;
  PDB_CHECK_ARG(x && out, "fused_softmax_mask_upper_triangle: NULL tensor pointer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (dtype) {
#define M(DT)                                                                                                          \
  case DT: {                                                                                                           \
    using T = typename CType<DT>::type;                                                                                \
    return softmax_mask_typed<T, T, true>(static_cast<const T*>(x), nullptr, static_cast<T*>(out), x_dims[0], x_dims[1], \
                                          x_dims[2], x_dims[3], st);                                                   \
  }
    M(PDB200_FLOAT32) M(PDB200_FLOAT16) M(PDB200_BFLOAT16)
#undef M
  }
  return fail(PDB200_UNSUPPORTED, "fused_softmax_mask_upper_triangle: dtype must be float32/float16/bfloat16, got code " +
                                      std::to_string(dtype));
}
