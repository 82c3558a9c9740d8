// argmax.cu — argmax along one axis or over the flattened tensor, on the shared reduction
// engine (reduce_impl.cuh) with a (value, index) accumulator.
//
// Replaces (fresh design): ArgCUDAKernel / ComputeFullArg / VisitDataCudaArgMinMaxFunctor
//   paddle/phi/kernels/gpu/arg_min_max_kernel.cu:59-204 (one CTA per row,
//   cub::BlockReduce<KeyValuePair<int, T>>, scalar strided loads, `int` keys).
// Here: 16-byte loads, 64-bit indices, long rows / flatten cut into partial (value, index)
// pairs folded by a second pass; lowest index wins ties in every pass.
#include "reduce_impl.cuh"

namespace pdb200 {

template <typename T, typename O, bool MIN>
int argmax_run(const ReducePlan& plan, const void* x, void* out, void* ws, size_t ws_bytes,
               cudaStream_t st) {
  using V = typename AccType<T>::type;  // compare bf16/fp16 as float: exact and order-preserving
  using Op = ArgMaxOp<V, MIN>;
  return run_reduce<T, O, Op>(plan, static_cast<const T*>(x), static_cast<O*>(out), IndexFin<V, O>(), ws,
                              ws_bytes, st);
}

template <typename T>
size_t argmax_ws(const ReducePlan& plan) {
  using V = typename AccType<T>::type;
  return reduce_scratch_bytes<T, KV<V>>(plan);
}

// [pre][n][post] view of the reduction (arg_min_max_kernel.cu:150-180)
inline int argmax_plan(const int64_t* dims, int rank, int64_t axis, int flatten, ReducePlan* plan) {
  if (rank < 0 || rank > PDB200_MAX_RANK) return fail(PDB200_INVALID_ARGUMENT, "argmax: rank must be in [0, 9]");
  if (rank > 0 && dims == nullptr) return fail(PDB200_INVALID_ARGUMENT, "argmax: x_dims is NULL");
  int64_t numel = 1;
  for (int i = 0; i < rank; ++i) {
    if (dims[i] < 0) return fail(PDB200_INVALID_ARGUMENT, "argmax: negative dimension");
    numel *= dims[i];
  }
  if (numel <= 0)
    return fail(PDB200_INVALID_ARGUMENT, "argmin/argmax input numel must > 0, bug got " + std::to_string(numel));
  plan->numel = numel;
  plan->passes.clear();
  if (flatten || rank == 0) {
    plan->reduce_num = numel;
    plan->out_numel = 1;
    plan->passes.push_back(ReducePass{1, numel, 1});
    return PDB200_OK;
  }
  if (axis < -rank || axis >= rank)
    return fail(PDB200_INVALID_ARGUMENT, "'axis'(" + std::to_string(axis) + ") must be greater than or equal to -Rank(X)(" +
                                             std::to_string(-rank) + ") and less than Rank(X)(" + std::to_string(rank) + ").");
  if (axis < 0) axis += rank;
  ReducePass ps{1, dims[axis], 1};
  for (int i = 0; i < axis; ++i) ps.outer *= dims[i];
  for (int i = (int)axis + 1; i < rank; ++i) ps.inner *= dims[i];
  plan->reduce_num = ps.red;
  plan->out_numel = ps.outer * ps.inner;
  plan->passes.push_back(ps);
  return PDB200_OK;
}

}  // namespace pdb200

#define PDB_ARG_TYPES(M) \
  M(PDB200_FLOAT32) M(PDB200_FLOAT64) M(PDB200_INT32) M(PDB200_INT64) M(PDB200_INT16) M(PDB200_UINT8) \
  M(PDB200_BFLOAT16) M(PDB200_FLOAT16)

namespace {
template <bool MIN>
int arg_entry(const void* x, const int64_t* x_dims, int x_rank, int64_t axis, int flatten, int x_dtype, int out_dtype,
              void* out, void* workspace, size_t workspace_bytes, void* stream) {
  using namespace pdb200;
  ReducePlan plan;
  int rc = argmax_plan(x_dims, x_rank, axis, flatten, &plan);
  if (rc != PDB200_OK) return rc;
  PDB_CHECK_ARG(x && out, "argmax: NULL tensor pointer");
  // arg_min_max_kernel.cu:221-230: the index dtype is int64 (default) or int32
  PDB_CHECK_ARG(out_dtype == PDB200_INT64 || out_dtype == PDB200_INT32,
                "The attribute of dtype in argmin/argmax must be [int32] or [int64], but received code " +
                    std::to_string(out_dtype));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
#define PDB_RUN(DT)                                                                              \
  case DT: {                                                                                     \
    using T = typename CType<DT>::type;                                                          \
    if (out_dtype == PDB200_INT64) return argmax_run<T, int64_t, MIN>(plan, x, out, workspace, workspace_bytes, st); \
    return argmax_run<T, int32_t, MIN>(plan, x, out, workspace, workspace_bytes, st);            \
  }
  switch (x_dtype) {
    PDB_ARG_TYPES(PDB_RUN)
    default:
      return fail(PDB200_UNSUPPORTED, std::string(MIN ? "argmin" : "argmax") + ": unsupported dtype code " + std::to_string(x_dtype));
  }
#undef PDB_RUN
}
}  // namespace

extern "C" {

size_t pdb200_argmax_workspace_bytes(const int64_t* x_dims, int x_rank, int64_t axis, int flatten,
                                     int x_dtype) {
  pdb200::ReducePlan plan;
  if (pdb200::argmax_plan(x_dims, x_rank, axis, flatten, &plan) != PDB200_OK) return 0;
#define PDB_WS(DT) \
  case DT: return pdb200::argmax_ws<typename pdb200::CType<DT>::type>(plan);
  switch (x_dtype) {
    PDB_ARG_TYPES(PDB_WS)
    default: return 0;
  }
#undef PDB_WS
}

int pdb200_argmax(const void* x, const int64_t* x_dims, int x_rank, int64_t axis, int flatten,
                  int x_dtype, int out_dtype, void* out, void* workspace, size_t workspace_bytes,
                  void* stream) {
  return arg_entry<false>(x, x_dims, x_rank, axis, flatten, x_dtype, out_dtype, out, workspace, workspace_bytes, stream);
}

int pdb200_argmin(const void* x, const int64_t* x_dims, int x_rank, int64_t axis, int flatten,
                  int x_dtype, int out_dtype, void* out, void* workspace, size_t workspace_bytes,
                  void* stream) {
  return arg_entry<true>(x, x_dims, x_rank, axis, flatten, x_dtype, out_dtype, out, workspace, workspace_bytes, stream);
}

}  // extern "C"
