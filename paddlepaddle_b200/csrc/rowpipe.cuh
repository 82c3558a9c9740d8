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
#include <stdlib.h>

#include <map>
#include <mutex>
#include <tuple>

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

constexpr int kPipeMaxStages = 32;  // mbarriers per CTA

// Tuning knob: row -> CTA mapping of the pipelined kernels (1 = contiguous slab per CTA).
inline int pipe_contiguous() { return knobs().pipe_contig; }

// Launch helper shared by the pipelined kernels.  `want_ctas` CTAs per SM (capped by what registers
// and threads allow) each get `want_stages` ring stages (capped by shared memory); the result is
// VERIFIED with the occupancy calculator at the final shared-memory size — a persistent grid sized
// for N CTAs per SM runs in two waves if only N-1 fit.  The shared-memory carve-out is set to what
// the CTAs need and no more, so that the rest of the 256 KB stays L1 (layer_norm re-reads scale and
// bias through L1 for every row; with the carve-out at its maximum that kernel ran 30% slower).
// Geometry comes from sweeps on B200 (profiles/r1_sweep.txt): past ~512-768 threads and ~3 stages
// per SM more concurrency does not help and often hurts.
// Returns false when fewer than 2 stages fit (caller falls back to a non-pipelined route).
// On success `hold` owns the function-attribute lock (common.cuh): the caller launches while it is alive, so the
// (carve-out, max dynamic smem) pair this shape needs is the one in force at the launch, whatever other host
// threads are doing with other shapes of the same kernel.
template <typename Kern>
inline bool pipe_geometry(Kern kern, int threads, size_t row_bytes, int want_ctas, int want_stages,
                          int* grid_per_sm, int* stages, size_t* smem_bytes, FuncAttrHold* hold, size_t extra_bytes = 0) {
  auto ok = [](cudaError_t e) {
    if (e != cudaSuccess) cudaGetLastError();
    return e == cudaSuccess;
  };
  const Knobs& kn = knobs();
  if (kn.pipe_max_ctas >= 1) want_ctas = kn.pipe_max_ctas;
  if (kn.pipe_max_stages >= 2) want_stages = kn.pipe_max_stages;
  if (want_stages > kPipeMaxStages) want_stages = kPipeMaxStages;
  // The answer depends only on (kernel, device, threads, row size, wishes): remember it, so that a launch
  // costs one map lookup instead of a dozen attribute / occupancy calls.
  struct Geo { int per_sm, stages; size_t smem; int pct; bool ok; };
  static std::map<std::tuple<const void*, int, int, size_t, int, int, size_t>, Geo> cache;  // guarded by func_attr_mutex()
  int dev = 0;
  cudaGetDevice(&dev);
  const void* fn = reinterpret_cast<const void*>(kern);
  const auto key = std::make_tuple(fn, dev, threads, row_bytes, want_ctas, want_stages, extra_bytes);
  hold->lock = std::unique_lock<std::mutex>(func_attr_mutex());
  auto give_up = [&]() {
    hold->lock.unlock();
    return false;
  };
  {
    auto it = cache.find(key);
    if (it != cache.end()) {
      const Geo& g = it->second;
      if (!g.ok) return give_up();
      // another shape of the same kernel may have changed the function attributes since: re-apply if so
      if (!func_attr_apply_locked(fn, (int)g.smem, g.pct)) return give_up();
      *grid_per_sm = g.per_sm;
      *stages = g.stages;
      *smem_bytes = g.smem;
      return true;
    }
  }
  // extra_bytes: shared memory the kernel keeps next to its ring for the life of the CTA (layer_norm: scale and bias)
  if (!func_attr_apply_locked(fn, (int)(2 * row_bytes + extra_bytes), -1)) return give_up();
  int occ = 0;
  if (!ok(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, threads, 2 * row_bytes + extra_bytes)) || occ < 1) return give_up();
  if (occ > want_ctas) occ = want_ctas;
  const size_t sm_smem = 227 * 1024;
  for (; occ >= 1; --occ) {
    if (sm_smem / occ < 2048 + extra_bytes + 2 * row_bytes) continue;
    int st = static_cast<int>((sm_smem / occ - 2048 - extra_bytes) / row_bytes);  // 1 KB reserved + static barriers per CTA
    if (st > want_stages) st = want_stages;
    // Keep about 64 KB of rows in flight per SM and no more: 2 stages instead of 3 gained 3 % (layer_norm, 16 KB rows,
    // 2 CTAs), 10 % (layer_norm_grad), 8 % (rms_norm) and 8 % (softmax fp32, 8 KB rows, 3 CTAs); 4 stages were worse still
    // (profiles/r1e_notes.txt).
    // Never fewer than two stages: a CTA must be able to compute on one row while the next arrives.
    {
      const int kb = kn.pipe_inflight_kb >= 8 ? kn.pipe_inflight_kb : 64;
      int cap = static_cast<int>(((size_t)kb * 1024) / ((size_t)occ * row_bytes));
      if (cap < 2) cap = 2;
      if (st > cap) st = cap;
    }
    for (; st >= 2; --st) {
      const size_t smem = (size_t)st * row_bytes + extra_bytes;
      int pct = static_cast<int>((occ * (smem + 2048) * 100 + sm_smem - 1) / sm_smem);
      if (pct > 100) pct = 100;
      if (!func_attr_apply_locked(fn, (int)smem, pct)) continue;
      int real = 0;
      if (!ok(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&real, kern, threads, smem))) continue;
      if (real >= occ) {
        *grid_per_sm = occ;
        *stages = st;
        *smem_bytes = smem;
        cache[key] = Geo{occ, st, smem, pct, true};
        return true;
      }
    }
  }
  cache[key] = Geo{0, 0, 0, 0, false};
  return give_up();
}

}  // namespace pdb200
