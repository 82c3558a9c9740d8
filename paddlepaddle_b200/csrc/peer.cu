// peer.cu — the exchange step of sharded reductions over NVLink peer memory.
//
// Replaces (fresh design): paddle.distributed.all_reduce -> ProcessGroupNCCL::AllReduce -> ncclAllReduce
//   paddle/fluid/distributed/collective/process_group_nccl.cc:267-297, phi/core/distributed/nccl_comm_context.cc:199-205
// for the small results this path exchanges (a scalar per rank for a full reduce, a vector of column sums for a
// row-sharded sum(axis=0)).  One process per GPU; every rank owns a "window" of device memory
//     [ flags: one 64-bit epoch per rank | slot 0 | slot 1 ]
// that every other rank maps through CUDA IPC (NVLink / NVSwitch peer access).  An all-reduce is then:
//   1. the local reduction writes its result straight into this rank's slot (epoch & 1) — it is the `out` pointer of
//      the reduction kernel, no copy;
//   2. ONE kernel: lane r of CTA 0 publishes the epoch into this rank's flag word in rank r's window (system-scope
//      release store over NVLink); every CTA waits until all ranks' epochs have arrived in the LOCAL flags (acquire
//      loads), then reads every rank's slot with 16-byte peer loads and combines them in rank order — all ranks
//      compute bit-identical results.
// No NCCL call, no host synchronisation, no staging copy.  Two slots alternate by epoch: a rank can only be two calls
// ahead of the slowest reader after passing the intervening barrier, so a slot is never overwritten while it is read.
// The wait is bounded (~1 s); on expiry the kernel raises the window's error word instead of hanging the GPU.
#include <string>

#include "common.cuh"

namespace pdb200 {

constexpr int kPeerMaxWorld = 16;
constexpr size_t kPeerFlagBytes = 256;  // 16 ranks x 8 B epoch, + error word at byte 128, padded
constexpr size_t kPeerErrOffset = 128;

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

// 16-byte load that never hits a (non-coherent) L1 line left from an earlier epoch
template <typename T>
__device__ __forceinline__ Vec<T> ld_peer(const T* p) {
  Vec<T> r;
  uint4 u;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w) : "l"(p) : "memory");
  *reinterpret_cast<uint4*>(&r) = u;
  return r;
}

enum { kPeerSum = 0, kPeerMax = 1, kPeerMin = 2, kPeerProd = 3 };
template <typename A, int OP>
__device__ __forceinline__ A peer_combine(A a, A b) {
  if (OP == kPeerSum) return a + b;
  if (OP == kPeerProd) return a * b;
  if (OP == kPeerMax) return b > a ? b : a;
  return b < a ? b : a;
}

// n elements of T per rank in slot (epoch & 1); out[i] = op over ranks, accumulated in AccType<T>, rank order.
template <typename T, int OP>
__global__ void __launch_bounds__(256) peer_allreduce_kernel(PeerPtrs ptrs, int world, int rank, size_t slot_bytes,
                                                             unsigned long long epoch, int64_t n, T* __restrict__ out) {
  using A = typename AccType<T>::type;
  __shared__ int timed_out;
  unsigned char* mine = ptrs.win[rank];
  if (threadIdx.x == 0) timed_out = 0;
  __syncthreads();
  // lane r of the first warp talks to rank r: all publishes and all waits are in flight at once
  if (threadIdx.x < world) {
    const int r = threadIdx.x;
    if (blockIdx.x == 0) {
      __threadfence_system();  // the slot was written by an earlier kernel on this stream: visible system-wide first
      st_release_sys(reinterpret_cast<unsigned long long*>(ptrs.win[r]) + rank, epoch);
    }
    const unsigned long long* flag = reinterpret_cast<const unsigned long long*>(mine) + r;
    long long spins = 0;
    while (ld_acquire_sys(flag) < epoch) {
      if (++spins > (1ll << 24)) {  // ~1 s: a peer never arrived — report, do not hang
        timed_out = 1;
        *reinterpret_cast<volatile int*>(mine + kPeerErrOffset) = 1;
        break;
      }
      if (spins > 64) __nanosleep(32);
    }
  }
  __syncthreads();
  if (timed_out) return;
  const size_t off = kPeerFlagBytes + (size_t)(epoch & 1ull) * slot_bytes;
  constexpr int V = Vec<T>::N;
  const int64_t nvec = n / V;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    A acc[V];
#pragma unroll
    for (int r = 0; r < kPeerMaxWorld; ++r) {
      if (r < world) {
        const Vec<T> v = ld_peer(reinterpret_cast<const T*>(ptrs.win[r] + off) + i * V);
#pragma unroll
        for (int k = 0; k < V; ++k) acc[k] = r == 0 ? cvt<A>(v.v[k]) : peer_combine<A, OP>(acc[k], cvt<A>(v.v[k]));
      }
    }
    Vec<T> o;
#pragma unroll
    for (int k = 0; k < V; ++k) o.v[k] = cvt<T>(acc[k]);
    *reinterpret_cast<uint4*>(out + i * V) = *reinterpret_cast<const uint4*>(&o);
  }
  if (blockIdx.x == 0) {  // ragged tail
    for (int64_t i = nvec * V + threadIdx.x; i < n; i += blockDim.x) {
      A acc = cvt<A>(reinterpret_cast<const volatile T*>(ptrs.win[0] + off)[i]);
      for (int r = 1; r < world; ++r)
        acc = peer_combine<A, OP>(acc, cvt<A>(reinterpret_cast<const volatile T*>(ptrs.win[r] + off)[i]));
      out[i] = cvt<T>(acc);
    }
  }
}

template <typename T>
int peer_launch(int op, const PeerPtrs& p, int world, int rank, size_t slot_bytes, unsigned long long epoch, int64_t n, T* out,
                cudaStream_t st) {
  const int64_t nvec = n / Vec<T>::N;
  int64_t blocks = std::max<int64_t>(1, std::min<int64_t>((nvec + 255) / 256, (int64_t)sm_count()));
  switch (op) {
    case kPeerSum: peer_allreduce_kernel<T, kPeerSum><<<(unsigned)blocks, 256, 0, st>>>(p, world, rank, slot_bytes, epoch, n, out); break;
    case kPeerMax: peer_allreduce_kernel<T, kPeerMax><<<(unsigned)blocks, 256, 0, st>>>(p, world, rank, slot_bytes, epoch, n, out); break;
    case kPeerMin: peer_allreduce_kernel<T, kPeerMin><<<(unsigned)blocks, 256, 0, st>>>(p, world, rank, slot_bytes, epoch, n, out); break;
    case kPeerProd: peer_allreduce_kernel<T, kPeerProd><<<(unsigned)blocks, 256, 0, st>>>(p, world, rank, slot_bytes, epoch, n, out); break;
    default: return fail(PDB200_INVALID_ARGUMENT, "peer_all_reduce: unknown op code " + std::to_string(op));
  }
  return after_launch("peer_allreduce");
}

}  // namespace pdb200

extern "C" {

size_t pdb200_peer_window_bytes(size_t slot_bytes) { return pdb200::kPeerFlagBytes + 2 * ((slot_bytes + 255) / 256 * 256); }

int pdb200_peer_window_create(size_t slot_bytes, void** window, void* ipc_handle_64) {
  using namespace pdb200;
  PDB_CHECK_ARG(window && ipc_handle_64, "peer_window_create: NULL output");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
  void* p = nullptr;
  const size_t bytes = pdb200_peer_window_bytes(slot_bytes);
  if (cudaMalloc(&p, bytes) != cudaSuccess) return fail(PDB200_CUDA_ERROR, std::string("peer_window_create: cudaMalloc: ") + cudaGetErrorString(cudaGetLastError()));
  cudaMemset(p, 0, bytes);
  cudaIpcMemHandle_t h;
  if (cudaIpcGetMemHandle(&h, p) != cudaSuccess) {
    const std::string e = cudaGetErrorString(cudaGetLastError());
    cudaFree(p);
    return fail(PDB200_CUDA_ERROR, "peer_window_create: cudaIpcGetMemHandle: " + e);
  }
  memcpy(ipc_handle_64, &h, 64);
  cudaDeviceSynchronize();  // the zeroed flags are in memory before any peer can see the handle
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
  if (window) cudaFree(window);
  cudaGetLastError();
  return PDB200_OK;
}

void* pdb200_peer_slot(void* window, size_t slot_bytes, unsigned long long epoch) {
  return static_cast<unsigned char*>(window) + pdb200::kPeerFlagBytes + (size_t)(epoch & 1ull) * ((slot_bytes + 255) / 256 * 256);
}

int pdb200_peer_error(void* window) {  // host read of the error word (synchronises the device)
  int e = 0;
  if (cudaMemcpy(&e, static_cast<unsigned char*>(window) + pdb200::kPeerErrOffset, sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess) {
    cudaGetLastError();
    return -1;
  }
  return e;
}

int pdb200_peer_all_reduce(int op, void* const* windows, int world, int rank, size_t slot_bytes, unsigned long long epoch,
                           int64_t n, int dtype, void* out, void* stream) {
  using namespace pdb200;
  PDB_CHECK_ARG(windows && out, "peer_all_reduce: NULL argument");
  PDB_CHECK_ARG(world >= 1 && world <= kPeerMaxWorld && rank >= 0 && rank < world, "peer_all_reduce: bad world / rank");
  PDB_CHECK_ARG(epoch >= 1, "peer_all_reduce: epochs start at 1 (0 is the cleared state of the flags)");
  const size_t es = dtype_size(dtype);
  if (es == 0) return fail(PDB200_UNSUPPORTED, "peer_all_reduce: bad dtype code " + std::to_string(dtype));
  const size_t slot = (slot_bytes + 255) / 256 * 256;
  if (n < 0 || (size_t)n * es > slot) return fail(PDB200_INVALID_ARGUMENT, "peer_all_reduce: " + std::to_string(n) + " elements do not fit the " + std::to_string(slot) + "-byte slot");
  PeerPtrs p;
  for (int r = 0; r < kPeerMaxWorld; ++r) p.win[r] = r < world ? static_cast<unsigned char*>(windows[r]) : nullptr;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (dtype) {
#define M(DT) \
  case DT: return peer_launch<typename CType<DT>::type>(op, p, world, rank, slot, epoch, n, static_cast<typename CType<DT>::type*>(out), st);
    M(PDB200_FLOAT32) M(PDB200_FLOAT64) M(PDB200_INT32) M(PDB200_INT64) M(PDB200_BFLOAT16) M(PDB200_FLOAT16)
#undef M
  }
  return fail(PDB200_UNSUPPORTED, "peer_all_reduce: dtype code " + std::to_string(dtype) + " is not built");
}

}  // extern "C"
