// rowpipe.cuh — TMA-fed row pipeline primitives for the row-wise kernels (softmax, layer_norm).
//
// Rows of a contiguous [rows][D] tensor are contiguous byte ranges, so each row is fetched with ONE
// 1-D bulk async copy (cp.async.bulk, the TMA engine; SASS UBLKCP) into a ring of shared-memory
// stages, completion signalled on an mbarrier by byte count.  A persistent CTA keeps STAGES rows in
// flight regardless of how many registers the arithmetic needs, so HBM latency is hidden by the
// copy engine instead of by occupancy: the register-resident kernels this replaces were either
// latency-bound at 16 warps/SM (softmax: 112 regs) or issue-bound re-loading scale/bias per row
// (layer_norm) — profiles/r1_*.txt.
//
// Protocol per stage s (one row):
//   producer (thread 0): mbarrier.arrive.expect_tx(full[s], row_bytes); cp.async.bulk -> stage s
//   consumers (all)    : mbarrier.try_wait.parity(full[s], phase); LDS the row into registers;
//                        CTA barrier; thread 0: fence.proxy.async (generic reads before async
//                        writes to the same bytes), then refills stage s with row k + STAGES.
// Results leave through ordinary streaming 16-byte global stores from registers.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace pdb200 {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
// make the inits visible to the async proxy (the TMA engine) before any copy targets them
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra WAIT_DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "WAIT_DONE:\n\t"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// order this thread's (and, after a CTA barrier, the CTA's) generic-proxy smem accesses before
// subsequent async-proxy accesses to the same addresses
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// 1-D bulk copy global -> shared, completion counted in bytes on `bar`.  16-byte aligned, size % 16 == 0.
__device__ __forceinline__ void bulk_load_row(void* smem_dst, const void* gmem_src, uint32_t bytes,
                                              uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(smem_dst)),
      "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// 16-byte shared-memory load
template <typename T>
__device__ __forceinline__ Vec<T> lds_vec(const T* p) {
  Vec<T> r;
  *reinterpret_cast<uint4*>(&r) = *reinterpret_cast<const uint4*>(p);
  return r;
}

// Host: ring depth for rows of `row_bytes` when `ctas_per_sm` CTAs share an SM.  Leaves ~27 KB of
// the 227 KB for static shared memory and the driver's reserve.
inline int pipe_stages(size_t row_bytes, int ctas_per_sm) {
  const size_t budget = 200 * 1024;
  int st = static_cast<int>(budget / (static_cast<size_t>(ctas_per_sm) * row_bytes));
  if (st > 8) st = 8;
  return st;
}

// Launch helper shared by the pipelined kernels: find how many CTAs fit an SM (registers,
// threads), give each an equal share of shared memory as ring stages, run a persistent grid.
// Returns false when fewer than 2 stages fit (caller falls back to a non-pipelined route).
template <typename Kern>
inline bool pipe_geometry(Kern kern, int threads, size_t row_bytes, int* grid_per_sm, int* stages,
                          size_t* smem_bytes) {
  int occ = 0;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * (int)row_bytes) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, threads, 2 * row_bytes) != cudaSuccess || occ < 1) {
    cudaGetLastError();
    return false;
  }
  if (occ > 16) occ = 16;
  int st = pipe_stages(row_bytes, occ);
  while (st < 3 && occ > 1) {  // prefer depth over CTA count
    --occ;
    st = pipe_stages(row_bytes, occ);
  }
  if (st < 2) return false;
  *grid_per_sm = occ;
  *stages = st;
  *smem_bytes = (size_t)st * row_bytes;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)*smem_bytes) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return true;
}

}  // namespace pdb200
