// no_alloc.cu — column sums (sum over axis 0 of a [rows, cols] matrix) for callers that cannot
// allocate device memory.
//
// Who needs this: code that runs where cudaMalloc is off limits — a process whose device memory was
// sized and handed out at start-up (inference servers that pre-plan every byte), a real-time loop
// that must not take the allocator's lock or risk its implicit synchronisation, or a plugin host
// that gives a kernel library no allocator at all.  The regular entry point (pdb200_sum_raw) asks such a caller for a workspace buffer; a
// caller that has nowhere to get one from can use this variant instead, which brings its own
// scratch: a fixed 1.5 MiB array that lives in the library's device image.  (On a 148-SM B200 the
// split heuristic never asks for more than ~0.9 MB for float32/int64 input and ~1.8 MB for
// bfloat16 input accumulated in float; only the widest of the bfloat16 split shapes - roughly
// 131072 to 151552 columns - are refused.)
//
// Trade-offs the caller accepts:
//   * the array exists on every GPU that loads libpdb200.so, from load until the process exits;
//   * its size is the limit: a call whose partial sums do not fit is refused, not resized;
//   * there is ONE array per device, shared by every call and every stream.  Two calls that may run
//     on the GPU at the same time (different streams, or different host threads on different
//     streams) would write their partials over each other — run them on one stream, or order them
//     with events.  The array keeps whatever the previous call left in it; that is harmless here,
//     because the first pass overwrites every slot the second pass reads, so nothing has to start
//     at zero and nothing is cleared.
#include <mutex>

#include "reduce_impl.cuh"

namespace pdb200 {

constexpr size_t kStaticScratchBytes = size_t(3) << 19;  // 1.5 MiB per device, stated in pdb200.h

// File-scope device array.  In HOST code this name is only a handle for the runtime (passing it to
// a kernel would hand the kernel a host address); the device address comes from
// cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;

// Device address of g_static_scratch on the CURRENT device.  Looked up once per device, the first
// time this entry point is used there (so that first call should happen before any stream
// capture starts: symbol lookup is not a stream operation); later calls read the cached pointer.
static int static_scratch_address(void** out) {
  static std::mutex mu;
  static void* ptrs[kMaxDevices] = {nullptr};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(PDB200_CUDA_ERROR, std::string("cudaGetDevice: ") + cudaGetErrorString(e));
  if (dev < 0 || dev >= kMaxDevices) return fail(PDB200_UNSUPPORTED, "device ordinal beyond the static scratch table");
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_scratch);  // with device `dev` current
    if (e != cudaSuccess || p == nullptr)
      return fail(PDB200_CUDA_ERROR, std::string("cudaGetSymbolAddress(static scratch): ") + cudaGetErrorString(e));
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return PDB200_OK;
}

// x[rows][cols] -> out[cols].  Same kernels and the same launch geometry as pdb200_sum_raw with
// axes = {0}; only the home of the partial sums differs.
template <typename T, typename O>
int sum_axis0_static(const T* x, int64_t rows, int64_t cols, O* out, cudaStream_t st) {
  using A = typename AccType<O>::type;
  using Op = SumOp<A>;
  const ReducePass ps{1, rows, cols};
  const bool vec = is_aligned16(x) && ((cols == 1 ? rows : cols) % Vec<T>::N == 0);
  const PassGeom g = plan_pass(ps, sizeof(T), vec ? Vec<T>::N : 1);
  if (g.splits == 1)  // enough columns to fill the machine: one pass, no scratch touched
    return launch_pass<T, O, Op, CastFin<A, O>>(ps, g, vec, x, out, CastFin<A, O>(), st);
  const size_t need = (size_t)g.splits * (size_t)cols * sizeof(A);
  if (need > kStaticScratchBytes)
    return fail(PDB200_WORKSPACE_TOO_SMALL,
                "pdb200_sum_axis0_static: this shape needs " + std::to_string(need) +
                    " bytes of scratch (" + std::to_string(g.splits) + " partial rows of " + std::to_string(cols) +
                    " sums) but the built-in scratch holds " + std::to_string(kStaticScratchBytes) +
                    "; use pdb200_sum_raw with a caller-provided workspace");
  void* base = nullptr;
  int rc = static_scratch_address(&base);
  if (rc != PDB200_OK) return rc;
  A* part = static_cast<A*>(base);
  rc = launch_pass<T, A, Op, CastFin<A, A>>(ps, g, vec, x, part, CastFin<A, A>(), st);
  if (rc != PDB200_OK) return rc;
  const ReducePass fold{1, g.splits, cols};
  const bool fvec = ((cols == 1 ? (int64_t)g.splits : cols) % Vec<A>::N == 0);
  PassGeom fg = plan_pass(fold, sizeof(A), fvec ? Vec<A>::N : 1);
  fg.splits = 1;
  fg.seg = fold.red;
  return launch_pass<A, O, Op, CastFin<A, O>>(fold, fg, fvec, part, out, CastFin<A, O>(), st);
}

}  // namespace pdb200

extern "C" int pdb200_sum_axis0_static(const void* x, int64_t rows, int64_t cols, int x_dtype, int out_dtype,
                                       void* out, void* stream) {
  using namespace pdb200;
  PDB_CHECK_ARG(rows > 0 && cols > 0, "pdb200_sum_axis0_static: rows and cols must be positive");
  PDB_CHECK_ARG(x && out, "pdb200_sum_axis0_static: NULL tensor pointer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (x_dtype == PDB200_FLOAT32 && out_dtype == PDB200_FLOAT32)
    return sum_axis0_static<float, float>(static_cast<const float*>(x), rows, cols, static_cast<float*>(out), st);
  if (x_dtype == PDB200_BFLOAT16 && out_dtype == PDB200_BFLOAT16)
    return sum_axis0_static<__nv_bfloat16, __nv_bfloat16>(static_cast<const __nv_bfloat16*>(x), rows, cols,
                                                          static_cast<__nv_bfloat16*>(out), st);
  if (x_dtype == PDB200_BFLOAT16 && out_dtype == PDB200_FLOAT32)
    return sum_axis0_static<__nv_bfloat16, float>(static_cast<const __nv_bfloat16*>(x), rows, cols,
                                                  static_cast<float*>(out), st);
  if (x_dtype == PDB200_INT64 && out_dtype == PDB200_INT64)
    return sum_axis0_static<int64_t, int64_t>(static_cast<const int64_t*>(x), rows, cols,
                                              static_cast<int64_t*>(out), st);
  return fail(PDB200_UNSUPPORTED, "pdb200_sum_axis0_static: supported dtype pairs are float32->float32, "
                                  "bfloat16->bfloat16, bfloat16->float32, int64->int64");
}
