// reduce_cols_tma.cuh — strided-axis reduction with TMA bulk-TENSOR staging (cp.async.bulk.tensor.2d).
//
// x viewed as [outer][red][inner] is, for the copy engine, a 2-D tensor [outer*red rows][inner cols].
// A CTA owns one column tile (32 x 16-byte vectors = 512 B wide) of one (outer, split) and walks its
// rows in boxes of kTileRows rows: each box is ONE tensor-map copy into a shared-memory ring stage,
// signalled on an mbarrier, so the strided walk costs no address arithmetic and no registers, and
// kStages boxes stay in flight per CTA.  Threads read the staged box conflict-free (a warp reads one
// 512-byte box row per instruction), eight row-groups fold in parallel and meet in shared memory.
// Out-of-range columns of the last tile are zero-filled by the copy engine and never stored; rows past
// the split end are never read from shared memory.
//
// This is the tiled/strided-axis counterpart of the 1-D row pipeline in rowpipe.cuh; the plain
// 16-byte-load kernel (reduce_cols_kernel) remains for shapes the tensor map cannot describe
// (unaligned bases, inner not a multiple of 16 bytes, narrow inner, 1-byte types).
#pragma once

#include <cuda.h>

#include "rowpipe.cuh"

namespace pdb200 {

constexpr int kTmaTileRows = 32;
constexpr int kTmaStages = 4;
constexpr int kTmaThreads = 256;  // 32 vector-columns x 8 row-groups

__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int32_t c0, int32_t c1,
                                            uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
          smem_u32(smem_dst)),
      "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar))
      : "memory");
}

template <typename T, typename O, typename Op, typename Fin>
__global__ void __launch_bounds__(kTmaThreads) reduce_cols_tma_kernel(const __grid_constant__ CUtensorMap map,
                                                                      O* __restrict__ out, int64_t red, int64_t inner,
                                                                      int64_t seg, int splits, int64_t col_tiles,
                                                                      Fin fin, int nstages) {
  using A = typename Op::Acc;
  constexpr int V = Vec<T>::N;
  constexpr int TI = 32 * V;  // tile width in elements (512 bytes)
  constexpr uint32_t kTileBytes = (uint32_t)kTmaTileRows * TI * sizeof(T);
  extern __shared__ __align__(128) unsigned char tma_ring_raw[];
  T* ring = reinterpret_cast<T*>(tma_ring_raw);                             // [kTmaStages][kTmaTileRows][TI]
  A* fold = reinterpret_cast<A*>(tma_ring_raw + (size_t)kTmaStages * kTileBytes);  // [8][TI]
  __shared__ uint64_t full[kTmaStages];
  const int tid = threadIdx.x;
  const int vc = tid & 31, rg = tid >> 5;
  int64_t b = blockIdx.x;
  const int64_t ct = b % col_tiles;
  b /= col_tiles;
  const int split = (int)(b % splits);
  const int64_t o = b / splits;
  const int64_t begin = (int64_t)split * seg;
  const int64_t end = begin + seg < red ? begin + seg : red;
  const int32_t c0 = (int32_t)(ct * TI);
  const int64_t row_base = o * red;  // row of the 2-D view where this outer slice starts
  const int64_t ntiles = (end - begin + kTmaTileRows - 1) / kTmaTileRows;
  if (tid == 0) {
    for (int s = 0; s < kTmaStages; ++s) mbar_init(&full[s], 1);
    mbar_fence_init();
  }
  __syncthreads();
  if (tid == 0) {
    for (int s = 0; s < nstages && s < ntiles; ++s) {
      mbar_expect_tx(&full[s], kTileBytes);
      tma_load_2d(ring + (size_t)s * kTmaTileRows * TI, &map, c0, (int32_t)(row_base + begin + (int64_t)s * kTmaTileRows),
                  &full[s]);
    }
  }
  A acc[V];
#pragma unroll
  for (int k = 0; k < V; ++k) acc[k] = Op::identity();
  int s = 0;
  uint32_t phase = 0;
  for (int64_t t = 0; t < ntiles; ++t) {
    const int64_t r0 = begin + t * kTmaTileRows;
    const int nvalid = (int)(end - r0 < kTmaTileRows ? end - r0 : kTmaTileRows);
    mbar_wait(&full[s], phase);
    const T* tile = ring + (size_t)s * kTmaTileRows * TI;
#pragma unroll
    for (int i = 0; i < kTmaTileRows / 8; ++i) {
      const int r = rg + i * 8;
      if (r < nvalid) {
        const Vec<T> v = lds_vec(tile + (size_t)r * TI + vc * V);
#pragma unroll
        for (int k = 0; k < V; ++k) acc_step<Op, T>(acc[k], v.v[k], r0 + r);
      }
    }
    __syncthreads();  // every thread is done with stage s
    if (tid == 0 && t + nstages < ntiles) {
      fence_proxy_async_smem();
      mbar_expect_tx(&full[s], kTileBytes);
      tma_load_2d(ring + (size_t)s * kTmaTileRows * TI, &map, c0,
                  (int32_t)(row_base + begin + (t + nstages) * kTmaTileRows), &full[s]);
    }
    if (++s == nstages) {
      s = 0;
      phase ^= 1;
    }
  }
  // fold the 8 row-groups of each column
#pragma unroll
  for (int k = 0; k < V; ++k) fold[(size_t)rg * TI + vc * V + k] = acc[k];
  __syncthreads();
  if (rg == 0) {
    const int64_t col = (int64_t)c0 + vc * V;
    if (col < inner) {  // inner % V == 0: a vector is entirely inside or entirely outside
      O* dst = out + (o * splits + split) * inner + col;
#pragma unroll
      for (int k = 0; k < V; ++k) {
        A a = fold[vc * V + k];
#pragma unroll
        for (int g = 1; g < 8; ++g) a = Op::combine(a, fold[(size_t)g * TI + vc * V + k]);
        dst[k] = fin(a);
      }
    }
  }
}

// ---- host: tensor map over [rows2d][inner] ---------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline PFN_encodeTiled tma_encode_fn() {
  static PFN_encodeTiled fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess) {
      cudaGetLastError();
      p = nullptr;
    }
    return reinterpret_cast<PFN_encodeTiled>(p);
  }();
  return fn;
}

// Returns true and fills `map` when x[rows2d][inner] (element size es) can be described; the box is
// [kTmaTileRows][512 bytes].  The data type only tells the engine the element SIZE (no conversion happens).
inline bool make_cols_tensor_map(CUtensorMap* map, const void* x, int64_t rows2d, int64_t inner, size_t es) {
  PFN_encodeTiled enc = tma_encode_fn();
  if (enc == nullptr) return false;
  if (rows2d >= (int64_t(1) << 31) || inner >= (int64_t(1) << 31)) return false;  // int32 box coordinates
  CUtensorMapDataType dt;
  switch (es) {
    case 2: dt = CU_TENSOR_MAP_DATA_TYPE_UINT16; break;
    case 4: dt = CU_TENSOR_MAP_DATA_TYPE_UINT32; break;
    case 8: dt = CU_TENSOR_MAP_DATA_TYPE_UINT64; break;
    default: return false;
  }
  const cuuint64_t gdim[2] = {(cuuint64_t)inner, (cuuint64_t)rows2d};
  const cuuint64_t gstride[1] = {(cuuint64_t)inner * es};  // bytes between rows; must be a multiple of 16
  const cuuint32_t box[2] = {(cuuint32_t)(512 / es), (cuuint32_t)kTmaTileRows};
  const cuuint32_t estr[2] = {1, 1};
  return enc(map, dt, 2, const_cast<void*>(x), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace pdb200
