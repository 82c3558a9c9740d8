// reduce_prod.cu — prod entry point and the shared workspace query
// (see reduce_dispatch.cuh / reduce_impl.cuh).
#include "reduce_dispatch.cuh"

namespace pdb200 {

int sum_mean_dispatch(bool mean, bool query, size_t* ws_out, const ReducePlan& plan, const void* x,
                      int x_dtype, int out_dtype, void* out, void* ws, size_t ws_bytes, cudaStream_t st);
int minmax_dispatch(bool is_max, bool query, size_t* ws_out, const ReducePlan& plan, const void* x,
                    int x_dtype, int out_dtype, void* out, void* ws, size_t ws_bytes, cudaStream_t st);

int prod_dispatch(bool query, size_t* ws_out, const ReducePlan& plan, const void* x, int x_dtype,
                  int out_dtype, void* out, void* ws, size_t ws_bytes, cudaStream_t st) {
  // registered GPU dtypes: kps/reduce_kernel.cu:399-411
  PDB_REDUCE_PAIR(PDB200_FLOAT32, PDB200_FLOAT32, ProdOp, kFinCast)
  PDB_REDUCE_PAIR(PDB200_FLOAT64, PDB200_FLOAT64, ProdOp, kFinCast)
  PDB_REDUCE_PAIR(PDB200_INT32, PDB200_INT32, ProdOp, kFinCast)
  PDB_REDUCE_PAIR(PDB200_INT64, PDB200_INT64, ProdOp, kFinCast)
  PDB_REDUCE_PAIR(PDB200_BFLOAT16, PDB200_BFLOAT16, ProdOp, kFinCast)
  PDB_REDUCE_PAIR(PDB200_FLOAT16, PDB200_FLOAT16, ProdOp, kFinCast)
  return fail(PDB200_UNSUPPORTED, "prod: unsupported dtype pair x=" + std::to_string(x_dtype) +
                                      " out=" + std::to_string(out_dtype));
}

}  // namespace pdb200

extern "C" {

int pdb200_prod(const void* x, const int64_t* x_dims, int x_rank, const int64_t* axes, int n_axes,
                int reduce_all, int x_dtype, int out_dtype, void* out, void* workspace,
                size_t workspace_bytes, void* stream) {
  pdb200::ReducePlan plan;
  int rc = pdb200::reduce_prepare(x_dims, x_rank, axes, n_axes, reduce_all, &plan);
  if (rc != PDB200_OK) return rc;
  return pdb200::prod_dispatch(false, nullptr, plan, x, x_dtype, out_dtype, out, workspace, workspace_bytes,
                               reinterpret_cast<cudaStream_t>(stream));
}

size_t pdb200_reduce_workspace_bytes(int op, const int64_t* x_dims, int x_rank, const int64_t* axes,
                                     int n_axes, int reduce_all, int x_dtype, int out_dtype) {
  pdb200::ReducePlan plan;
  if (pdb200::reduce_prepare(x_dims, x_rank, axes, n_axes, reduce_all, &plan) != PDB200_OK) return 0;
  size_t ws = 0;
  int rc;
  switch (op) {
    case PDB200_REDUCE_SUM:
      rc = pdb200::sum_mean_dispatch(false, true, &ws, plan, nullptr, x_dtype, out_dtype, nullptr, nullptr, 0, nullptr);
      break;
    case PDB200_REDUCE_MEAN:
      rc = pdb200::sum_mean_dispatch(true, true, &ws, plan, nullptr, x_dtype, out_dtype, nullptr, nullptr, 0, nullptr);
      break;
    case PDB200_REDUCE_MAX:
      rc = pdb200::minmax_dispatch(true, true, &ws, plan, nullptr, x_dtype, out_dtype, nullptr, nullptr, 0, nullptr);
      break;
    case PDB200_REDUCE_MIN:
      rc = pdb200::minmax_dispatch(false, true, &ws, plan, nullptr, x_dtype, out_dtype, nullptr, nullptr, 0, nullptr);
      break;
    case PDB200_REDUCE_PROD:
      rc = pdb200::prod_dispatch(true, &ws, plan, nullptr, x_dtype, out_dtype, nullptr, nullptr, 0, nullptr);
      break;
    default:
      rc = PDB200_INVALID_ARGUMENT;
  }
  return rc == PDB200_OK ? ws : 0;
}

}  // extern "C"
