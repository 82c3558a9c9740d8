// common.cuh — shared device/host helpers for the B200 (sm_100a) PHI hot path.
//
// Everything here is written for one target: compute_100a / sm_100a.  There is no
// multi-arch dispatch and no CPU fallback.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <stdlib.h>

#include <atomic>
#include <mutex>
#include <string>

#include "../../include/pdb200.h"

namespace pdb200 {

// ----------------------------------------------------------------------------
// error plumbing (host)
// ----------------------------------------------------------------------------
void set_error(const std::string& msg);
int fail(int code, const std::string& msg);
extern std::atomic<uint64_t> g_launch_count;

#define PDB_CHECK_ARG(cond, msg)                                          \
  do {                                                                    \
    if (!(cond)) return ::pdb200::fail(PDB200_INVALID_ARGUMENT, (msg));   \
  } while (0)

// Called after every <<<>>>: counts the launch and converts a launch error.
inline int after_launch(const char* what) {
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    return fail(PDB200_CUDA_ERROR, std::string(what) + ": " + cudaGetErrorString(e));
  }
  return PDB200_OK;
}

inline int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

// ----------------------------------------------------------------------------
// Tuning knobs (PDB200_* environment variables, profiles/sweep.py).  Read ONCE, at the first launch that asks:
// a launch costs no getenv() afterwards, and the values cannot change between two host threads' launches.
// 0 / false means "use the built-in choice".
// ----------------------------------------------------------------------------
struct Knobs {
  int pipe_contig, pipe_max_ctas, pipe_max_stages, pipe_inflight_kb, pipe_rows_per_cta;
  int softmax_pipe_c, ln_pipe_c, rms_pipe_c, cluster_slice_kb, cluster_rows_per_cluster, softmax_long, cluster_threads, relay_stages, relay_cfg, ln_param_smem;
  int tma_cols_stages, expand_rows_per_cta, mulgrad_stages;
  long long mulgrad_splits;
  bool no_tma_cols, no_interleave, no_fused_mulgrad, no_cluster_pipe, no_cols_cluster, no_fln_param_smem;
  long long peer_timeout_ms;
  long long rows_warp_max_bytes;
};
inline const Knobs& knobs() {
  static const Knobs k = [] {
    auto num = [](const char* name) -> long long {
      const char* v = getenv(name);
      return v ? atoll(v) : 0;
    };
    auto set = [](const char* name) { return getenv(name) != nullptr; };
    Knobs r{};
    r.pipe_contig = (int)num("PDB200_PIPE_CONTIG");
    r.pipe_max_ctas = (int)num("PDB200_PIPE_MAX_CTAS");
    r.pipe_max_stages = (int)num("PDB200_PIPE_MAX_STAGES");
    r.pipe_inflight_kb = (int)num("PDB200_PIPE_INFLIGHT_KB");
    r.pipe_rows_per_cta = (int)num("PDB200_PIPE_ROWS_PER_CTA");
    r.softmax_pipe_c = (int)num("PDB200_SOFTMAX_PIPE_C");
    r.ln_pipe_c = (int)num("PDB200_LN_PIPE_C");
    r.rms_pipe_c = (int)num("PDB200_RMS_PIPE_C");
    r.cluster_slice_kb = (int)num("PDB200_CLUSTER_SLICE_KB");
    r.cluster_threads = (int)num("PDB200_CLUSTER_THREADS");
    r.relay_stages = (int)num("PDB200_RELAY_STAGES");
    r.relay_cfg = (int)num("PDB200_RELAY_CFG");
    r.ln_param_smem = (int)num("PDB200_LN_PARAM_SMEM");
    r.cluster_rows_per_cluster = (int)num("PDB200_CLUSTER_ROWS");
    r.softmax_long = (int)num("PDB200_SOFTMAX_LONG");
    r.tma_cols_stages = (int)num("PDB200_TMA_COLS_STAGES");
    r.expand_rows_per_cta = (int)num("PDB200_EXPAND_ROWS_PER_CTA");
    r.mulgrad_stages = (int)num("PDB200_MULGRAD_STAGES");
    r.mulgrad_splits = num("PDB200_MULGRAD_SPLITS");
    r.no_tma_cols = set("PDB200_NO_TMA_COLS");
    r.no_interleave = set("PDB200_NO_INTERLEAVE");
    r.no_fused_mulgrad = set("PDB200_NO_FUSED_MULGRAD");
    r.no_cluster_pipe = set("PDB200_NO_CLUSTER_PIPE");
    r.no_cols_cluster = set("PDB200_NO_COLS_CLUSTER");
    r.no_fln_param_smem = set("PDB200_NO_FLN_PARAM_SMEM");
    r.peer_timeout_ms = num("PDB200_PEER_TIMEOUT_MS");
    r.rows_warp_max_bytes = num("PDB200_ROWS_WARP_MAX_BYTES");
    return r;
  }();
  return k;
}

// ----------------------------------------------------------------------------
// Per-function launch attributes (max dynamic shared memory, shared-memory carve-out) are state of the
// (kernel, device) pair inside the driver, not of a launch.  Every launch that depends on them takes this lock,
// applies its pair only if it differs from what the pair was last set to, and releases the lock AFTER the launch
// call has returned — two host threads launching different shapes of one instantiation cannot interleave a
// set-attribute of one with the launch of the other.
// ----------------------------------------------------------------------------
struct FuncAttrHold {
  std::unique_lock<std::mutex> lock;
};
std::mutex& func_attr_mutex();
// Apply (max dynamic smem, carve-out pct or -1) to `kern` on the current device; `hold` keeps the lock until it dies.
bool func_attr_apply(const void* kern, int smem_bytes, int carveout_pct, FuncAttrHold* hold);
bool func_attr_apply_locked(const void* kern, int smem_bytes, int carveout_pct);  // caller holds func_attr_mutex()

inline size_t dtype_size(int dt) {
  switch (dt) {
    case PDB200_BOOL: case PDB200_UINT8: case PDB200_INT8: return 1;
    case PDB200_INT16: case PDB200_FLOAT16: case PDB200_BFLOAT16: return 2;
    case PDB200_INT32: case PDB200_FLOAT32: return 4;
    case PDB200_INT64: case PDB200_FLOAT64: return 8;
  }
  return 0;
}

inline bool is_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ----------------------------------------------------------------------------
// Accumulation type: phi::dtype::MPTypeTrait (amp_type_traits.h:24-54) —
// float for fp16/bf16, the type itself otherwise.
// ----------------------------------------------------------------------------
template <typename T> struct AccType { using type = T; };
template <> struct AccType<__half> { using type = float; };
template <> struct AccType<__nv_bfloat16> { using type = float; };

// Conversions.  float -> bf16/fp16 is round-to-nearest-even with canonical NaN
// (0x7FFF for bf16), the same rule as phi/common/bfloat16.h:61-88.
template <typename To, typename From>
__device__ __forceinline__ To cvt(From v) { return static_cast<To>(v); }
template <> __device__ __forceinline__ float cvt<float, __nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float cvt<float, __half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ __nv_bfloat16 cvt<__nv_bfloat16, float>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half cvt<__half, float>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 cvt<__nv_bfloat16, __nv_bfloat16>(__nv_bfloat16 v) { return v; }
template <> __device__ __forceinline__ __half cvt<__half, __half>(__half v) { return v; }
template <> __device__ __forceinline__ double cvt<double, __nv_bfloat16>(__nv_bfloat16 v) { return (double)__bfloat162float(v); }
template <> __device__ __forceinline__ double cvt<double, __half>(__half v) { return (double)__half2float(v); }
template <> __device__ __forceinline__ int64_t cvt<int64_t, bool>(bool v) { return v ? 1 : 0; }

// ----------------------------------------------------------------------------
// 128-bit vectors.  Vec<T> is exactly 16 bytes: 8 x bf16/fp16/int16, 4 x fp32/int32,
// 2 x fp64/int64, 16 x 8-bit.  (The reference caps vectors at 4 ELEMENTS, i.e. 8-byte
// loads for bf16: funcs/aligned_vector.h:67-119.)
// ----------------------------------------------------------------------------
template <typename T>
struct alignas(16) Vec {
  static constexpr int N = 16 / sizeof(T);
  T v[N];
};

// Streaming 16-byte load: do not allocate in L1 (each byte is used once).  A COHERENT load (ld.global, not
// ld.global.nc): several entry points are called in place (add_, softmax_, scale_: out == x), where the non-coherent
// path is outside the PTX rules even though every chunk is read and written by the same thread.  The two forms run at
// the same speed on B200 (profiles/r2_notes.txt: config 2 6914.1 vs 6914.8 GB/s, the whole table within noise).
template <typename T>
__device__ __forceinline__ Vec<T> ld_stream(const T* p) {
  Vec<T> r;
  uint4 u;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w)
               : "l"(p)
               : "memory");
  *reinterpret_cast<uint4*>(&r) = u;
  return r;
}
// Cached 16-byte load (small operands that are re-read: broadcast rows, scale/bias).
template <typename T>
__device__ __forceinline__ Vec<T> ld_cached(const T* p) {
  Vec<T> r;
  *reinterpret_cast<uint4*>(&r) = __ldg(reinterpret_cast<const uint4*>(p));
  return r;
}
// Plain (coherent) 16-byte load — used when out may alias the input (in-place ops).
template <typename T>
__device__ __forceinline__ Vec<T> ld_plain(const T* p) {
  Vec<T> r;
  *reinterpret_cast<uint4*>(&r) = *reinterpret_cast<const uint4*>(p);
  return r;
}
// Streaming 16-byte store (evict-first in L1; written once, never re-read by this kernel).
template <typename T>
__device__ __forceinline__ void st_stream(T* p, const Vec<T>& r) {
  const uint4 u = *reinterpret_cast<const uint4*>(&r);
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(u.x), "r"(u.y),
               "r"(u.z), "r"(u.w)
               : "memory");
}

// N elements of T moved as one 2/4/8/16-byte access (N * sizeof(T) is a power of two <= 16): the output side
// of kernels whose result type is narrower than their input (compare/logical: bool out), and both sides of
// the unary kernels (cast).
template <typename T, int N>
struct alignas(sizeof(T) * N) Pack {
  T v[N];
};
template <typename T, int N>
__device__ __forceinline__ void st_stream(T* p, const Pack<T, N>& r) {
  constexpr int B = (int)sizeof(T) * N;
  static_assert(B == 2 || B == 4 || B == 8 || B == 16, "pack must be 2, 4, 8 or 16 bytes");
  if constexpr (B == 16) {
    const uint4 u = *reinterpret_cast<const uint4*>(&r);
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(u.x), "r"(u.y), "r"(u.z),
                 "r"(u.w)
                 : "memory");
  } else if constexpr (B == 8) {
    const uint2 u = *reinterpret_cast<const uint2*>(&r);
    asm volatile("st.global.L1::no_allocate.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(u.x), "r"(u.y) : "memory");
  } else if constexpr (B == 4) {
    const uint32_t u = *reinterpret_cast<const uint32_t*>(&r);
    asm volatile("st.global.L1::no_allocate.u32 [%0], %1;" ::"l"(p), "r"(u) : "memory");
  } else {
    const uint16_t u = *reinterpret_cast<const uint16_t*>(&r);
    asm volatile("st.global.L1::no_allocate.u16 [%0], %1;" ::"l"(p), "h"(u) : "memory");
  }
}
// Streaming load of a pack (coherent path: callers may run in place).
template <typename T, int N>
__device__ __forceinline__ Pack<T, N> ld_pack(const T* p) {
  constexpr int B = (int)sizeof(T) * N;
  static_assert(B == 1 || B == 2 || B == 4 || B == 8 || B == 16, "pack must be 1..16 bytes");
  Pack<T, N> r;
  if constexpr (B == 16) *reinterpret_cast<uint4*>(&r) = __ldcs(reinterpret_cast<const uint4*>(p));
  else if constexpr (B == 8) *reinterpret_cast<uint2*>(&r) = __ldcs(reinterpret_cast<const uint2*>(p));
  else if constexpr (B == 4) *reinterpret_cast<uint32_t*>(&r) = __ldcs(reinterpret_cast<const uint32_t*>(p));
  else if constexpr (B == 2) *reinterpret_cast<uint16_t*>(&r) = __ldcs(reinterpret_cast<const uint16_t*>(p));
  else *reinterpret_cast<uint8_t*>(&r) = __ldcs(reinterpret_cast<const uint8_t*>(p));
  return r;
}
inline bool is_aligned_to(const void* p, size_t bytes) { return (reinterpret_cast<uintptr_t>(p) % bytes) == 0; }

// Whole-vector conversion to float.  For bf16 the generic per-element path costs two
// instructions on odd elements (PRMT + SHL); working on the 32-bit words gives one each:
// low half = w << 16, high half = w & 0xffff0000.
template <typename T>
__device__ __forceinline__ void unpack_to_float(const Vec<T>& v, float* out) {
#pragma unroll
  for (int k = 0; k < Vec<T>::N; ++k) out[k] = cvt<float>(v.v[k]);
}
template <>
__device__ __forceinline__ void unpack_to_float<__nv_bfloat16>(const Vec<__nv_bfloat16>& v, float* out) {
  const uint4 u = *reinterpret_cast<const uint4*>(&v);
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    out[2 * i] = __uint_as_float(w[i] << 16);
    out[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
template <>
__device__ __forceinline__ void unpack_to_float<__half>(const Vec<__half>& v, float* out) {
  const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __half22float2(h[i]);
    out[2 * i] = f.x;
    out[2 * i + 1] = f.y;
  }
}
// float -> T, two at a time (one F2FP.PACK for the 16-bit types)
template <typename T>
__device__ __forceinline__ void pack_from_float(const float* in, Vec<T>& v) {
#pragma unroll
  for (int k = 0; k < Vec<T>::N; ++k) v.v[k] = cvt<T>(in[k]);
}
template <>
__device__ __forceinline__ void pack_from_float<__nv_bfloat16>(const float* in, Vec<__nv_bfloat16>& v) {
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(in[2 * i], in[2 * i + 1]);
}
template <>
__device__ __forceinline__ void pack_from_float<__half>(const float* in, Vec<__half>& v) {
  __half2* h = reinterpret_cast<__half2*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2half2_rn(in[2 * i], in[2 * i + 1]);
}

// ----------------------------------------------------------------------------
// Packed fp32 pairs.  sm_100 has FFMA2 / FADD2 / FMUL2: one instruction performs two independent IEEE
// round-to-nearest fp32 operations on a 64-bit register pair — the same results as two scalar instructions at
// the same FLOP rate (profiles/micro/f32x2.cu: 72 vs 74 TFLOP/s), for HALF the issue slots.  The row kernels
// are issue-bound, not pipe-bound, so their per-element arithmetic is written on pairs.
// ----------------------------------------------------------------------------
struct F2 {
  unsigned long long v;
};
__device__ __forceinline__ F2 f2_make(float lo, float hi) {
  F2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ F2 f2_splat(float x) { return f2_make(x, x); }
__device__ __forceinline__ float f2_lo(F2 a) {
  float lo;
  asm("{ .reg .b32 hi; mov.b64 {%0, hi}, %1; }" : "=f"(lo) : "l"(a.v));
  return lo;
}
__device__ __forceinline__ float f2_hi(F2 a) {
  float hi;
  asm("{ .reg .b32 lo; mov.b64 {lo, %0}, %1; }" : "=f"(hi) : "l"(a.v));
  return hi;
}
__device__ __forceinline__ F2 f2_fma(F2 a, F2 b, F2 c) {
  F2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v));
  return r;
}
__device__ __forceinline__ F2 f2_add(F2 a, F2 b) {
  F2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}
__device__ __forceinline__ F2 f2_mul(F2 a, F2 b) {
  F2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}

// ----------------------------------------------------------------------------
// warp / block reduction helpers
// ----------------------------------------------------------------------------
template <typename A>
__device__ __forceinline__ A shfl_xor_any(A v, int mask) {
  static_assert(sizeof(A) % 4 == 0, "shuffle payload must be a multiple of 4 bytes");
  constexpr int W = sizeof(A) / 4;
  union { A a; uint32_t w[W]; } in, out;
  in.a = v;
#pragma unroll
  for (int i = 0; i < W; ++i) out.w[i] = __shfl_xor_sync(0xffffffffu, in.w[i], mask);
  return out.a;
}

template <typename A>
__device__ __forceinline__ A shfl_idx_any(A v, int src_lane) {
  static_assert(sizeof(A) % 4 == 0, "shuffle payload must be a multiple of 4 bytes");
  constexpr int W = sizeof(A) / 4;
  union { A a; uint32_t w[W]; } in, out;
  in.a = v;
#pragma unroll
  for (int i = 0; i < W; ++i) out.w[i] = __shfl_sync(0xffffffffu, in.w[i], src_lane);
  return out.a;
}

// Cluster barrier whose only published data was written by ONE thread (which calls this with publisher = true right after
// its shared-memory stores): that thread fences at cluster scope, every thread arrives relaxed and waits with acquire.
// cg::cluster_group::sync() makes EVERY thread run MEMBAR.ALL.GPU before its arrive (stall_membar ~1.2 cycles per issue in
// the long-row kernels, profiles/r2g_softmax_vocab_bf16.txt).  All threads of a warp execute the same .aligned instructions.
__device__ __forceinline__ void cluster_sync_published_by(bool publisher) {
  if (publisher) asm volatile("fence.acq_rel.cluster;" ::: "memory");
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// Reduce across the G (power of two, <= 32) consecutive lanes that share a row.
template <int G, typename A, typename Combine>
__device__ __forceinline__ A group_reduce(A v, Combine comb) {
#pragma unroll
  for (int m = G / 2; m >= 1; m >>= 1) v = comb(v, shfl_xor_any(v, m));
  return v;
}

// Whole-CTA reduction; result valid in every thread.  `smem` must hold >= 33 A's.
template <typename A, typename Combine>
__device__ __forceinline__ A block_reduce(A v, Combine comb, A identity, A* smem) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nwarps = (blockDim.x + 31) >> 5;
  v = group_reduce<32>(v, comb);
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  if (warp == 0) {
    A w = lane < nwarps ? smem[lane] : identity;
    w = group_reduce<32>(w, comb);
    if (lane == 0) smem[32] = w;
  }
  __syncthreads();
  A r = smem[32];
  __syncthreads();  // smem may be reused by the caller right away
  return r;
}

// One-barrier CTA reduction for persistent row kernels: every warp leaves its partial in
// `buf[warp]`, one __syncthreads, then every warp folds the 32 slots with a shuffle tree.  `buf`
// has 32 slots; slots >= nwarps must hold the identity (block_reduce_1sync_init, once per kernel).
// The caller alternates between two `buf`s on consecutive calls, so the next write to a buffer is
// always separated from these reads by the barrier of the call in between.
template <typename A>
__device__ __forceinline__ void block_reduce_1sync_init(A* buf, A identity) {
  if (threadIdx.x < 32) buf[threadIdx.x] = identity;
}
template <typename A, typename Combine>
__device__ __forceinline__ A block_reduce_1sync(A v, Combine comb, A* buf) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = group_reduce<32>(v, comb);
  if (lane == 0) buf[warp] = v;
  __syncthreads();
  return group_reduce<32>(buf[lane], comb);
}

// ----------------------------------------------------------------------------
// dtype -> C++ type
// ----------------------------------------------------------------------------
template <int DT> struct CType;
template <> struct CType<PDB200_BOOL> { using type = bool; };
template <> struct CType<PDB200_UINT8> { using type = uint8_t; };
template <> struct CType<PDB200_INT8> { using type = int8_t; };
template <> struct CType<PDB200_INT16> { using type = int16_t; };
template <> struct CType<PDB200_INT32> { using type = int32_t; };
template <> struct CType<PDB200_INT64> { using type = int64_t; };
template <> struct CType<PDB200_FLOAT16> { using type = __half; };
template <> struct CType<PDB200_BFLOAT16> { using type = __nv_bfloat16; };
template <> struct CType<PDB200_FLOAT32> { using type = float; };
template <> struct CType<PDB200_FLOAT64> { using type = double; };

}  // namespace pdb200
