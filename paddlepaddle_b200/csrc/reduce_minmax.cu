// reduce_minmax.cu — max_raw and min_raw entry points (see reduce_dispatch.cuh / reduce_impl.cuh).
#include "reduce_dispatch.cuh"

namespace pdb200 {

int minmax_dispatch(bool is_max, bool query, size_t* ws_out, const ReducePlan& plan, const void* x,
                    int x_dtype, int out_dtype, void* out, void* ws, size_t ws_bytes, cudaStream_t st) {
  // registered GPU dtypes: kps/reduce_kernel.cu:357-385 (fp8 max is out of scope)
  if (is_max) {
    PDB_REDUCE_PAIR(PDB200_FLOAT32, PDB200_FLOAT32, MaxOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_FLOAT64, PDB200_FLOAT64, MaxOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_INT32, PDB200_INT32, MaxOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_INT64, PDB200_INT64, MaxOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_BFLOAT16, PDB200_BFLOAT16, MaxOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_FLOAT16, PDB200_FLOAT16, MaxOp, kFinCast)
  } else {
    PDB_REDUCE_PAIR(PDB200_FLOAT32, PDB200_FLOAT32, MinOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_FLOAT64, PDB200_FLOAT64, MinOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_INT32, PDB200_INT32, MinOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_INT64, PDB200_INT64, MinOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_BFLOAT16, PDB200_BFLOAT16, MinOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_FLOAT16, PDB200_FLOAT16, MinOp, kFinCast)
  }
  return fail(PDB200_UNSUPPORTED, std::string(is_max ? "max" : "min") + ": unsupported dtype pair x=" +
                                      std::to_string(x_dtype) + " out=" + std::to_string(out_dtype));
}

}  // namespace pdb200

extern "C" {

int pdb200_max_raw(const void* x, const int64_t* x_dims, int x_rank, const int64_t* axes, int n_axes,
                   int reduce_all, int x_dtype, int out_dtype, void* out, void* workspace,
                   size_t workspace_bytes, void* stream) {
  pdb200::ReducePlan plan;
  int rc = pdb200::reduce_prepare(x_dims, x_rank, axes, n_axes, reduce_all, &plan);
  if (rc != PDB200_OK) return rc;
  return pdb200::minmax_dispatch(true, false, nullptr, plan, x, x_dtype, out_dtype, out, workspace,
                                 workspace_bytes, reinterpret_cast<cudaStream_t>(stream));
}

int pdb200_min_raw(const void* x, const int64_t* x_dims, int x_rank, const int64_t* axes, int n_axes,
                   int reduce_all, int x_dtype, int out_dtype, void* out, void* workspace,
                   size_t workspace_bytes, void* stream) {
  pdb200::ReducePlan plan;
  int rc = pdb200::reduce_prepare(x_dims, x_rank, axes, n_axes, reduce_all, &plan);
  if (rc != PDB200_OK) return rc;
  return pdb200::minmax_dispatch(false, false, nullptr, plan, x, x_dtype, out_dtype, out, workspace,
                                 workspace_bytes, reinterpret_cast<cudaStream_t>(stream));
}

}  // extern "C"
