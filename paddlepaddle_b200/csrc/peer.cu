// peer.cu — the exchange step of sharded reductions over NVLink peer memory.
//
// Replaces (fresh design): paddle.distributed.all_reduce -> ProcessGroupNCCL::AllReduce -> ncclAllReduce
//   paddle/fluid/distributed/collective/process_group_nccl.cc:267-297 (AllReduce), :467-500 (ReduceScatter),
//   phi/core/distributed/nccl_comm_context.cc:199-205, and the bucketed gradient reduce of
//   paddle/fluid/distributed/collective/reducer.cc:1089-1111
// for what this path exchanges: a scalar per rank for a full reduce, a vector of column sums for a row-sharded
// sum(axis=0).  One process per GPU; every rank owns a "window" of device memory
//     [ control block | slot 0 | slot 1 | result 0 | result 1 ]
// that every other rank maps through CUDA IPC (NVLink / NVSwitch peer access).  The local reduction writes its partial
// result straight into this rank's slot (epoch & 1) — it is the `out` pointer of the reduction kernel, no copy — in the
// ACCUMULATOR type (float partials for bf16 / fp16 sums: the cross-rank sum is rounded once, at the very end).
// Then ONE kernel does the exchange, in one of two forms:
//
//   one-shot  (vectors up to 64 KB, scalars): every rank waits for all ranks' epochs and reads every rank's slot with
//             16-byte peer loads, combining in rank order.  Latency-optimal: one NVLink round trip.
//   two-shot  (longer vectors) = reduce-scatter + all-gather in one kernel: rank r reduces only columns
//             [r*n/N, (r+1)*n/N) from all ranks' slots (each remote byte crosses NVLink once), rounds, and stores its
//             chunk of the result into EVERY rank's result region with 16-byte peer stores; a second flag per rank says
//             "my chunk has landed"; every rank then copies its own result region to `out`.  Per rank and direction the
//             links carry (N-1)/N of the vector twice instead of (N-1) whole vectors.
//
// Every rank computes bit-identical results (fixed rank order).  No NCCL call, no host synchronisation, no staging
// copy.  Slots and result regions alternate by epoch: a rank publishes epoch e+1 only after its kernel of epoch e has
// finished (stream order), and a peer overwrites a region of epoch e only in epoch e+2, after it has seen every rank's
// e+1 — so nothing is overwritten while it may still be read.
//
// Waiting is bounded by TIME (%globaltimer; 30 s unless PDB200_PEER_TIMEOUT_MS says otherwise — NCCL would wait for
// ever, a data-loader stall on one rank must not trip it).  On expiry the kernel does not leave `out` untouched: it
// fills it with NaN (floating point) or the all-ones bit pattern, raises the window's error word in device memory AND
// in a host-mapped word, and every later pdb200_peer_all_reduce on that window fails immediately with a clear error:
// after a timeout the epochs of the ranks no longer agree and the group must be rebuilt.
//
// Launch-order contract: the kernel of epoch e on rank A waits for the kernel of epoch e on every other rank.  Each
// rank must therefore enqueue its exchanges in the same order, and nothing enqueued BEFORE rank B's exchange on B's
// stream may wait (event, stream dependency) on work that rank A only enqueues AFTER its own exchange.  One process
// per GPU: two ranks sharing one GPU would have kernels waiting on each other for SMs.
#include <map>
#include <mutex>
#include <string>
#include <type_traits>

#include "common.cuh"

namespace pdb200 {

constexpr int kPeerMaxWorld = 16;
constexpr size_t kPeerCtrlBytes = 1024;
constexpr size_t kPeerFlagAOffset = 0;    // 16 x u64: "the partial of epoch e is in my slot"
constexpr size_t kPeerErrOffset = 128;    // int: a wait timed out
constexpr size_t kPeerFlagBOffset = 256;  // 16 x u64: "my chunk of the result of epoch e is in your result region"
constexpr size_t kPeerTicketOffset = 512; // unsigned: CTAs of this rank that have finished their chunk stores
constexpr size_t kPeerOneShotBytes = 64 * 1024;

struct PeerPtrs {
  unsigned char* win[kPeerMaxWorld];  // window base of every rank, as mapped into THIS process
};

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// 16-byte load that never hits a (non-coherent) L1 line left from an earlier epoch
template <typename T>
__device__ __forceinline__ Vec<T> ld_peer(const T* p) {
  Vec<T> r;
  uint4 u;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w) : "l"(p) : "memory");
  *reinterpret_cast<uint4*>(&r) = u;
  return r;
}
template <typename T>
__device__ __forceinline__ void st_peer(T* p, const Vec<T>& r) {
  const uint4 u = *reinterpret_cast<const uint4*>(&r);
  asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(u.x), "r"(u.y), "r"(u.z), "r"(u.w) : "memory");
}

enum { kPeerSum = 0, kPeerMax = 1, kPeerMin = 2, kPeerProd = 3 };
template <typename A, int OP>
__device__ __forceinline__ A peer_combine(A a, A b) {
  if (OP == kPeerSum) return a + b;
  if (OP == kPeerProd) return a * b;
  if (OP == kPeerMax) return b > a ? b : a;
  return b < a ? b : a;
}

// What a timed-out exchange leaves in `out`: NaN for floating point, all-ones bits otherwise.
template <typename T>
__device__ __forceinline__ T peer_poison() {
  if constexpr (std::is_same<T, float>::value) return __int_as_float(0x7fc00000);
  else if constexpr (std::is_same<T, double>::value) return __longlong_as_double(0x7ff8000000000000ll);
  else if constexpr (std::is_same<T, __nv_bfloat16>::value) return __ushort_as_bfloat16((unsigned short)0x7fc0);
  else if constexpr (std::is_same<T, __half>::value) return __ushort_as_half((unsigned short)0x7e00);
  else return (T)(~(T)0);
}

// Lanes 0..world-1 of the first warp of every CTA wait until flag word r (at `flag_off` of the local window) has
// reached `epoch`.  Returns (CTA-uniformly) whether every flag arrived in time; on expiry the error words are raised.
__device__ __forceinline__ bool peer_wait_all(unsigned char* mine, size_t flag_off, int world, unsigned long long epoch,
                                              unsigned long long timeout_ns, int* host_err, int* timed_out_smem) {
  if (threadIdx.x < world) {
    const unsigned long long* flag = reinterpret_cast<const unsigned long long*>(mine + flag_off) + threadIdx.x;
    unsigned long long t0 = 0;
    int spins = 0;
    while (ld_acquire_sys(flag) < epoch) {
      if (++spins > 64) {
        __nanosleep(64);
        const unsigned long long now = global_timer_ns();
        if (t0 == 0) t0 = now;
        if (now - t0 > timeout_ns) {  // a peer never arrived: report loudly, do not hang the GPU
          *timed_out_smem = 1;
          *reinterpret_cast<volatile int*>(mine + kPeerErrOffset) = 1;
          if (host_err != nullptr) *reinterpret_cast<volatile int*>(host_err) = 1;
          __threadfence_system();
          break;
        }
      }
    }
  }
  __syncthreads();
  return *timed_out_smem == 0;
}

template <typename T>
__device__ __forceinline__ void peer_poison_out(T* out, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = peer_poison<T>();
}

// ---- one-shot ---------------------------------------------------------------------------------------------------
// n elements of S (the slot type) per rank in slot (epoch & 1); out[i] = T(op over ranks in rank order).
// VEC: out is 16-byte aligned (the slots always are); otherwise every element goes through the scalar loop.
template <typename S, typename T, int OP, bool VEC>
__device__ __forceinline__ void peer_allreduce_body(const PeerPtrs& ptrs, int world, int rank, size_t slot_bytes,
                                                    unsigned long long epoch, int64_t n, T* out,
                                                    unsigned long long timeout_ns, int* host_err) {
  __shared__ int timed_out;
  unsigned char* mine = ptrs.win[rank];
  if (threadIdx.x == 0) timed_out = 0;
  __syncthreads();
  // lane r of the first warp talks to rank r: all publishes and all waits are in flight at once
  if (blockIdx.x == 0 && threadIdx.x < world) {
    __threadfence_system();  // the slot was written by an earlier kernel on this stream: visible system-wide first
    st_release_sys(reinterpret_cast<unsigned long long*>(ptrs.win[threadIdx.x] + kPeerFlagAOffset) + rank, epoch);
  }
  if (!peer_wait_all(mine, kPeerFlagAOffset, world, epoch, timeout_ns, host_err, &timed_out)) {
    peer_poison_out(out, n);
    return;
  }
  const size_t off = kPeerCtrlBytes + (size_t)(epoch & 1ull) * slot_bytes;
  constexpr int VS = Vec<S>::N;  // slot elements per 16-byte load
  const int64_t nvec = VEC ? n / VS : 0;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    S acc[VS];
#pragma unroll
    for (int r = 0; r < kPeerMaxWorld; ++r) {
      if (r < world) {
        const Vec<S> v = ld_peer(reinterpret_cast<const S*>(ptrs.win[r] + off) + i * VS);
#pragma unroll
        for (int k = 0; k < VS; ++k) acc[k] = r == 0 ? v.v[k] : peer_combine<S, OP>(acc[k], v.v[k]);
      }
    }
    Pack<T, VS> o;
#pragma unroll
    for (int k = 0; k < VS; ++k) o.v[k] = cvt<T>(acc[k]);
    *reinterpret_cast<Pack<T, VS>*>(out + i * VS) = o;
  }
  if (blockIdx.x == 0) {  // ragged tail (everything when out is not 16-byte aligned)
    for (int64_t i = nvec * VS + threadIdx.x; i < n; i += blockDim.x) {
      S acc = reinterpret_cast<const volatile S*>(ptrs.win[0] + off)[i];
      for (int r = 1; r < world; ++r) acc = peer_combine<S, OP>(acc, reinterpret_cast<const volatile S*>(ptrs.win[r] + off)[i]);
      out[i] = cvt<T>(acc);
    }
  }
}

// ---- two-shot: reduce-scatter + all-gather ----------------------------------------------------------------------
// Requires n % Vec<S>::N == 0 handled by the caller through a ragged tail done one-shot style by CTA 0; `out` 16-byte
// aligned.  Grid <= SM count (every CTA must be resident: CTAs wait for the grid's last CTA to publish).
template <typename S, typename T, int OP>
__device__ __forceinline__ void peer_allreduce2_body(const PeerPtrs& ptrs, int world, int rank, size_t slot_bytes,
                                                     unsigned long long epoch, int64_t n, T* out,
                                                     unsigned long long timeout_ns, int* host_err) {
  __shared__ int timed_out;
  __shared__ int last_cta;
  unsigned char* mine = ptrs.win[rank];
  if (threadIdx.x == 0) timed_out = 0;
  __syncthreads();
  if (blockIdx.x == 0 && threadIdx.x < world) {
    __threadfence_system();
    st_release_sys(reinterpret_cast<unsigned long long*>(ptrs.win[threadIdx.x] + kPeerFlagAOffset) + rank, epoch);
  }
  if (!peer_wait_all(mine, kPeerFlagAOffset, world, epoch, timeout_ns, host_err, &timed_out)) {
    peer_poison_out(out, n);
    return;
  }
  const size_t slot_off = kPeerCtrlBytes + (size_t)(epoch & 1ull) * slot_bytes;
  const size_t res_off = kPeerCtrlBytes + (size_t)(2 + (epoch & 1ull)) * slot_bytes;
  constexpr int VS = Vec<S>::N;
  const int64_t nvec = n / VS;
  // this rank's chunk of vectors
  const int64_t chunk = (nvec + world - 1) / world;
  const int64_t v0 = (int64_t)rank * chunk;
  const int64_t v1 = v0 + chunk < nvec ? v0 + chunk : nvec;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  // kPeerUnroll vectors per thread and trip: world x kPeerUnroll 16-byte peer loads in flight per thread — a peer load
  // takes microseconds, and the link is only full with a few KB outstanding per SM
  constexpr int kPeerUnroll = 2;
  constexpr int kGroup = 4;  // ranks whose loads are issued together
  for (int64_t i0 = v0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < v1; i0 += kPeerUnroll * stride) {
    Vec<S> in[kPeerUnroll][kGroup];
    S acc[kPeerUnroll][VS];
    // ranks in groups of kGroup: the loads of a group are all issued before any is used
#pragma unroll
    for (int half = 0; half < kPeerMaxWorld / kGroup; ++half) {
      if (half * kGroup >= world) break;
#pragma unroll
      for (int u = 0; u < kPeerUnroll; ++u) {
        const int64_t i = i0 + u * stride;
#pragma unroll
        for (int r = 0; r < kGroup; ++r)
          if (half * kGroup + r < world && i < v1) in[u][r] = ld_peer(reinterpret_cast<const S*>(ptrs.win[half * kGroup + r] + slot_off) + i * VS);
      }
#pragma unroll
      for (int u = 0; u < kPeerUnroll; ++u) {
#pragma unroll
        for (int r = 0; r < kGroup; ++r) {
          if (half * kGroup + r < world) {
#pragma unroll
            for (int k = 0; k < VS; ++k)
              acc[u][k] = (half == 0 && r == 0) ? in[u][r].v[k] : peer_combine<S, OP>(acc[u][k], in[u][r].v[k]);
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kPeerUnroll; ++u) {
      const int64_t i = i0 + u * stride;
      if (i < v1) {
        Pack<T, VS> o;
#pragma unroll
        for (int k = 0; k < VS; ++k) o.v[k] = cvt<T>(acc[u][k]);
        // the chunk goes to every rank's result region (own included): sizeof(T) * VS bytes per store
#pragma unroll
        for (int r = 0; r < kPeerMaxWorld; ++r) {
          if (r < world) *reinterpret_cast<Pack<T, VS>*>(reinterpret_cast<T*>(ptrs.win[r] + res_off) + i * VS) = o;
        }
      }
    }
  }
  if (blockIdx.x == 0) {  // ragged tail (< VS elements): rank 0 reduces it for everyone
    if (rank == 0) {
      for (int64_t i = nvec * VS + threadIdx.x; i < n; i += blockDim.x) {
        S acc = reinterpret_cast<const volatile S*>(ptrs.win[0] + slot_off)[i];
        for (int r = 1; r < world; ++r) acc = peer_combine<S, OP>(acc, reinterpret_cast<const volatile S*>(ptrs.win[r] + slot_off)[i]);
        for (int r = 0; r < world; ++r) reinterpret_cast<volatile T*>(ptrs.win[r] + res_off)[i] = cvt<T>(acc);
      }
    }
  }
  // all of this rank's chunk stores are ordered before the publish: every CTA fences and takes a ticket, the last one
  // publishes "my chunk has landed" to every rank
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned* ticket = reinterpret_cast<unsigned*>(mine + kPeerTicketOffset);
    const unsigned t = atomicAdd(ticket, 1u);
    last_cta = (t == gridDim.x - 1);
    if (last_cta) *ticket = 0;  // everyone has arrived: ready for the next call on this stream
  }
  __syncthreads();
  if (last_cta && threadIdx.x < world) {
    __threadfence_system();
    st_release_sys(reinterpret_cast<unsigned long long*>(ptrs.win[threadIdx.x] + kPeerFlagBOffset) + rank, epoch);
  }
  if (!peer_wait_all(mine, kPeerFlagBOffset, world, epoch, timeout_ns, host_err, &timed_out)) {
    peer_poison_out(out, n);
    return;
  }
  // local copy: result region -> out
  const T* res = reinterpret_cast<const T*>(mine + res_off);
  constexpr int VT = Vec<T>::N;
  const int64_t nv_out = n / VT;
  for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < nv_out; i0 += kPeerUnroll * stride) {
    Vec<T> v[kPeerUnroll];
#pragma unroll
    for (int u = 0; u < kPeerUnroll; ++u)
      if (i0 + u * stride < nv_out) v[u] = ld_peer(res + (i0 + u * stride) * VT);
#pragma unroll
    for (int u = 0; u < kPeerUnroll; ++u)
      if (i0 + u * stride < nv_out) *reinterpret_cast<uint4*>(out + (i0 + u * stride) * VT) = *reinterpret_cast<const uint4*>(&v[u]);
  }
  if (blockIdx.x == 0)
    for (int64_t i = nv_out * VT + threadIdx.x; i < n; i += blockDim.x) out[i] = reinterpret_cast<const volatile T*>(res)[i];
}

template <typename S, typename T, int OP, bool VEC>
__global__ void __launch_bounds__(256) peer_allreduce_kernel(PeerPtrs ptrs, int world, int rank, size_t slot_bytes,
                                                             unsigned long long epoch, int64_t n, T* out,
                                                             unsigned long long timeout_ns, int* host_err) {
  peer_allreduce_body<S, T, OP, VEC>(ptrs, world, rank, slot_bytes, epoch, n, out, timeout_ns, host_err);
}
template <typename S, typename T, int OP>
__global__ void __launch_bounds__(256, 2) peer_allreduce2_kernel(PeerPtrs ptrs, int world, int rank, size_t slot_bytes,
                                                              unsigned long long epoch, int64_t n, T* out,
                                                              unsigned long long timeout_ns, int* host_err) {
  peer_allreduce2_body<S, T, OP>(ptrs, world, rank, slot_bytes, epoch, n, out, timeout_ns, host_err);
}

// ---- every rank of a group in ONE grid (self-test on a single GPU) --------------------------------------------
// The exchange kernels of different ranks wait for one another, so they may only run as separate launches when every
// rank has a GPU of its own.  For a box with fewer GPUs than ranks the same device code runs as ONE cooperative grid:
// blockIdx.y is the rank, its window is ptrs.win[rank] (all windows live on this GPU), its output outs.p[rank].  A
// cooperative launch guarantees that all ranks' CTAs are resident together, which is what the waits need.
struct PeerOuts {
  void* p[kPeerMaxWorld];
};
template <typename S, typename T, int OP, bool TWO_SHOT>
__global__ void __launch_bounds__(256, 2) peer_allreduce_all_ranks_kernel(PeerPtrs ptrs, PeerOuts outs, int world,
                                                                          size_t slot_bytes, unsigned long long epoch,
                                                                          int64_t n, unsigned long long timeout_ns) {
  const int rank = blockIdx.y;
  T* out = static_cast<T*>(outs.p[rank]);
  if constexpr (TWO_SHOT) peer_allreduce2_body<S, T, OP>(ptrs, world, rank, slot_bytes, epoch, n, out, timeout_ns, nullptr);
  else peer_allreduce_body<S, T, OP, true>(ptrs, world, rank, slot_bytes, epoch, n, out, timeout_ns, nullptr);
}

template <typename S, typename T>
int peer_launch_all_ranks(int op, const PeerPtrs& p, const PeerOuts& o, int world, size_t slot_bytes, unsigned long long epoch,
                          int64_t n, unsigned long long timeout_ns, cudaStream_t st) {
  if (op != kPeerSum && op != kPeerMax) return fail(PDB200_UNSUPPORTED, "peer_all_reduce_emulated: op must be sum (0) or max (1)");
  const bool two_shot = world > 1 && (size_t)n * sizeof(S) > kPeerOneShotBytes;
  const int64_t nvec = n / Vec<S>::N;
  const int64_t chunk = (nvec + world - 1) / world;
  int64_t blocks = two_shot ? (chunk + 255) / 256 : (nvec + 255) / 256;
  // all ranks' CTAs must be resident together: at most 2 CTAs per SM in total
  const int64_t cap = std::max<int64_t>(1, 2 * (int64_t)sm_count() / world);
  blocks = std::max<int64_t>(1, std::min<int64_t>(blocks, cap));
  PeerPtrs pp = p;
  PeerOuts oo = o;
  void* args[] = {&pp, &oo, &world, &slot_bytes, &epoch, &n, &timeout_ns};
  const void* kern = nullptr;
  if (op == kPeerSum) kern = two_shot ? (const void*)peer_allreduce_all_ranks_kernel<S, T, kPeerSum, true>
                                      : (const void*)peer_allreduce_all_ranks_kernel<S, T, kPeerSum, false>;
  else kern = two_shot ? (const void*)peer_allreduce_all_ranks_kernel<S, T, kPeerMax, true>
                       : (const void*)peer_allreduce_all_ranks_kernel<S, T, kPeerMax, false>;
  if (cudaLaunchCooperativeKernel(kern, dim3((unsigned)blocks, (unsigned)world), dim3(256), args, 0, st) != cudaSuccess)
    return fail(PDB200_CUDA_ERROR, std::string("peer_all_reduce_emulated: ") + cudaGetErrorString(cudaGetLastError()));
  return after_launch("peer_allreduce_all_ranks");
}

template <typename S, typename T>
int peer_launch(int op, const PeerPtrs& p, int world, int rank, size_t slot_bytes, unsigned long long epoch, int64_t n, T* out,
                unsigned long long timeout_ns, int* host_err, cudaStream_t st) {
  constexpr int VS = Vec<S>::N;
  const bool aligned = is_aligned16(out);
  const bool two_shot = aligned && world > 1 && (size_t)n * sizeof(S) > kPeerOneShotBytes;
  const int64_t nvec = n / VS;
  if (two_shot) {
    const int64_t chunk = (nvec + world - 1) / world;
    // the grid must be co-resident (its CTAs wait for its last CTA): never more CTAs than SMs
    // two CTAs per SM (256 threads, < 128 registers: both always fit) — twice the loads in flight
    const int64_t blocks = std::max<int64_t>(1, std::min<int64_t>((chunk + 255) / 256, 2 * (int64_t)sm_count()));
#define PDB_PEER2(OP) peer_allreduce2_kernel<S, T, OP><<<(unsigned)blocks, 256, 0, st>>>(p, world, rank, slot_bytes, epoch, n, out, timeout_ns, host_err)
    switch (op) {
      case kPeerSum: PDB_PEER2(kPeerSum); break;
      case kPeerMax: PDB_PEER2(kPeerMax); break;
      case kPeerMin: PDB_PEER2(kPeerMin); break;
      case kPeerProd: PDB_PEER2(kPeerProd); break;
      default: return fail(PDB200_INVALID_ARGUMENT, "peer_all_reduce: unknown op code " + std::to_string(op));
    }
#undef PDB_PEER2
    return after_launch("peer_allreduce2");
  }
  const int64_t blocks = std::max<int64_t>(1, std::min<int64_t>((nvec + 255) / 256, (int64_t)sm_count()));
#define PDB_PEER1(OP)                                                                                                        \
  if (aligned) peer_allreduce_kernel<S, T, OP, true><<<(unsigned)blocks, 256, 0, st>>>(p, world, rank, slot_bytes, epoch, n, out, timeout_ns, host_err); \
  else peer_allreduce_kernel<S, T, OP, false><<<1, 256, 0, st>>>(p, world, rank, slot_bytes, epoch, n, out, timeout_ns, host_err)
  switch (op) {
    case kPeerSum: PDB_PEER1(kPeerSum); break;
    case kPeerMax: PDB_PEER1(kPeerMax); break;
    case kPeerMin: PDB_PEER1(kPeerMin); break;
    case kPeerProd: PDB_PEER1(kPeerProd); break;
    default: return fail(PDB200_INVALID_ARGUMENT, "peer_all_reduce: unknown op code " + std::to_string(op));
  }
#undef PDB_PEER1
  return after_launch("peer_allreduce");
}

// host-mapped error word of each window created by this process
struct PeerHostState {
  int* host_err;      // host pointer (cudaHostAlloc, mapped)
  int* host_err_dev;  // the same word as the device sees it
};
static std::mutex g_peer_mu;
static std::map<void*, PeerHostState> g_peer_state;

}  // namespace pdb200

extern "C" {

size_t pdb200_peer_window_bytes(size_t slot_bytes) { return pdb200::kPeerCtrlBytes + 4 * ((slot_bytes + 255) / 256 * 256); }

int pdb200_peer_window_create(size_t slot_bytes, void** window, void* ipc_handle_64) {
  using namespace pdb200;
  PDB_CHECK_ARG(window && ipc_handle_64, "peer_window_create: NULL output");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
  void* p = nullptr;
  const size_t bytes = pdb200_peer_window_bytes(slot_bytes);
  if (cudaMalloc(&p, bytes) != cudaSuccess) return fail(PDB200_CUDA_ERROR, std::string("peer_window_create: cudaMalloc: ") + cudaGetErrorString(cudaGetLastError()));
  cudaMemset(p, 0, kPeerCtrlBytes);
  cudaIpcMemHandle_t h;
  if (cudaIpcGetMemHandle(&h, p) != cudaSuccess) {
    const std::string e = cudaGetErrorString(cudaGetLastError());
    cudaFree(p);
    return fail(PDB200_CUDA_ERROR, "peer_window_create: cudaIpcGetMemHandle: " + e);
  }
  PeerHostState hs{nullptr, nullptr};
  if (cudaHostAlloc(reinterpret_cast<void**>(&hs.host_err), sizeof(int), cudaHostAllocMapped) == cudaSuccess) {
    *hs.host_err = 0;
    if (cudaHostGetDevicePointer(reinterpret_cast<void**>(&hs.host_err_dev), hs.host_err, 0) != cudaSuccess) {
      cudaGetLastError();
      cudaFreeHost(hs.host_err);
      hs = PeerHostState{nullptr, nullptr};
    }
  } else {
    cudaGetLastError();  // no mapped host memory: the device-side error word is still there (pdb200_peer_error)
  }
  memcpy(ipc_handle_64, &h, 64);
  cudaDeviceSynchronize();  // the zeroed control block is in memory before any peer can see the handle
  {
    std::lock_guard<std::mutex> lock(g_peer_mu);
    g_peer_state[p] = hs;
  }
  *window = p;
  return PDB200_OK;
}

int pdb200_peer_window_open(const void* ipc_handle_64, void** mapped) {
  using namespace pdb200;
  PDB_CHECK_ARG(ipc_handle_64 && mapped, "peer_window_open: NULL argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, ipc_handle_64, 64);
  if (cudaIpcOpenMemHandle(mapped, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess)
    return fail(PDB200_CUDA_ERROR, std::string("peer_window_open: cudaIpcOpenMemHandle: ") + cudaGetErrorString(cudaGetLastError()));
  return PDB200_OK;
}

int pdb200_peer_window_close(void* mapped) {
  if (mapped) cudaIpcCloseMemHandle(mapped);
  cudaGetLastError();
  return PDB200_OK;
}

int pdb200_peer_window_destroy(void* window) {
  using namespace pdb200;
  if (window) {
    {
      std::lock_guard<std::mutex> lock(g_peer_mu);
      auto it = g_peer_state.find(window);
      if (it != g_peer_state.end()) {
        if (it->second.host_err) cudaFreeHost(it->second.host_err);
        g_peer_state.erase(it);
      }
    }
    cudaFree(window);
  }
  cudaGetLastError();
  return PDB200_OK;
}

void* pdb200_peer_slot(void* window, size_t slot_bytes, unsigned long long epoch) {
  return static_cast<unsigned char*>(window) + pdb200::kPeerCtrlBytes + (size_t)(epoch & 1ull) * ((slot_bytes + 255) / 256 * 256);
}

int pdb200_peer_error(void* window) {  // host read of the device-side error word (synchronises the device)
  int e = 0;
  if (cudaMemcpy(&e, static_cast<unsigned char*>(window) + pdb200::kPeerErrOffset, sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess) {
    cudaGetLastError();
    return -1;
  }
  return e;
}

int pdb200_peer_all_reduce(int op, void* const* windows, int world, int rank, size_t slot_bytes, unsigned long long epoch,
                           int64_t n, int slot_dtype, int out_dtype, void* out, void* stream) {
  using namespace pdb200;
  PDB_CHECK_ARG(windows && out, "peer_all_reduce: NULL argument");
  PDB_CHECK_ARG(world >= 1 && world <= kPeerMaxWorld && rank >= 0 && rank < world, "peer_all_reduce: bad world / rank");
  PDB_CHECK_ARG(epoch >= 1, "peer_all_reduce: epochs start at 1 (0 is the cleared state of the flags)");
  const size_t es = dtype_size(slot_dtype);
  if (es == 0 || dtype_size(out_dtype) == 0) return fail(PDB200_UNSUPPORTED, "peer_all_reduce: bad dtype code");
  const size_t slot = (slot_bytes + 255) / 256 * 256;
  if (n < 0 || (size_t)n * es > slot) return fail(PDB200_INVALID_ARGUMENT, "peer_all_reduce: " + std::to_string(n) + " elements do not fit the " + std::to_string(slot) + "-byte slot");
  PDB_CHECK_ARG(is_aligned_to(out, dtype_size(out_dtype)), "peer_all_reduce: out is not aligned to its element size");
  int* host_err_dev = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_peer_mu);
    auto it = g_peer_state.find(windows[rank]);
    if (it != g_peer_state.end() && it->second.host_err != nullptr) {
      // deferred, sync-free check of every earlier exchange on this window
      if (*reinterpret_cast<volatile int*>(it->second.host_err) != 0)
        return fail(PDB200_CUDA_ERROR, "peer_all_reduce: an earlier exchange on this window timed out waiting for a peer; its "
                                       "output was filled with NaN / all-ones and the ranks' epochs no longer agree — rebuild the group");
      host_err_dev = it->second.host_err_dev;
    }
  }
  const long long ms = knobs().peer_timeout_ms > 0 ? knobs().peer_timeout_ms : 30000;
  const unsigned long long timeout_ns = (unsigned long long)ms * 1000000ull;
  PeerPtrs p;
  for (int r = 0; r < kPeerMaxWorld; ++r) p.win[r] = r < world ? static_cast<unsigned char*>(windows[r]) : nullptr;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
#define M(SD, OD)                                                                                                     \
  if (slot_dtype == SD && out_dtype == OD)                                                                            \
    return peer_launch<typename CType<SD>::type, typename CType<OD>::type>(op, p, world, rank, slot, epoch, n,        \
                                                                            static_cast<typename CType<OD>::type*>(out), \
                                                                            timeout_ns, host_err_dev, st);
  M(PDB200_FLOAT32, PDB200_FLOAT32) M(PDB200_FLOAT64, PDB200_FLOAT64) M(PDB200_INT32, PDB200_INT32) M(PDB200_INT64, PDB200_INT64)
  M(PDB200_BFLOAT16, PDB200_BFLOAT16) M(PDB200_FLOAT16, PDB200_FLOAT16)        // max / min of 16-bit floats: exact as they are
  M(PDB200_FLOAT32, PDB200_BFLOAT16) M(PDB200_FLOAT32, PDB200_FLOAT16)         // float partials, rounded once after the sum
#undef M
  return fail(PDB200_UNSUPPORTED, "peer_all_reduce: (slot dtype " + std::to_string(slot_dtype) + ", out dtype " +
                                      std::to_string(out_dtype) + ") is not built");
}

int pdb200_peer_all_reduce_emulated(int op, void* const* windows, int world, size_t slot_bytes, unsigned long long epoch,
                                    int64_t n, int slot_dtype, int out_dtype, void* const* outs, void* stream) {
  using namespace pdb200;
  PDB_CHECK_ARG(windows && outs, "peer_all_reduce_emulated: NULL argument");
  PDB_CHECK_ARG(world >= 1 && world <= kPeerMaxWorld, "peer_all_reduce_emulated: bad world");
  PDB_CHECK_ARG(epoch >= 1, "peer_all_reduce_emulated: epochs start at 1");
  const size_t es = dtype_size(slot_dtype);
  if (es == 0) return fail(PDB200_UNSUPPORTED, "peer_all_reduce_emulated: bad dtype code");
  const size_t slot = (slot_bytes + 255) / 256 * 256;
  if (n < 0 || (size_t)n * es > slot) return fail(PDB200_INVALID_ARGUMENT, "peer_all_reduce_emulated: n does not fit the slot");
  PeerPtrs p;
  PeerOuts o;
  for (int r = 0; r < kPeerMaxWorld; ++r) {
    p.win[r] = r < world ? static_cast<unsigned char*>(windows[r]) : nullptr;
    o.p[r] = r < world ? outs[r] : nullptr;
    if (r < world) PDB_CHECK_ARG(is_aligned16(outs[r]), "peer_all_reduce_emulated: outputs must be 16-byte aligned");
  }
  const long long ms = knobs().peer_timeout_ms > 0 ? knobs().peer_timeout_ms : 30000;
  const unsigned long long timeout_ns = (unsigned long long)ms * 1000000ull;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (slot_dtype == PDB200_FLOAT32 && out_dtype == PDB200_FLOAT32) return peer_launch_all_ranks<float, float>(op, p, o, world, slot, epoch, n, timeout_ns, st);
  if (slot_dtype == PDB200_FLOAT32 && out_dtype == PDB200_BFLOAT16) return peer_launch_all_ranks<float, __nv_bfloat16>(op, p, o, world, slot, epoch, n, timeout_ns, st);
  if (slot_dtype == PDB200_INT64 && out_dtype == PDB200_INT64) return peer_launch_all_ranks<int64_t, int64_t>(op, p, o, world, slot, epoch, n, timeout_ns, st);
  return fail(PDB200_UNSUPPORTED, "peer_all_reduce_emulated: built for (float32, float32), (float32, bfloat16) and (int64, int64)");
}

}  // extern "C"
