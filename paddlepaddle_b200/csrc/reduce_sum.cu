// reduce_sum.cu — sum_raw and mean_raw entry points (see reduce_dispatch.cuh / reduce_impl.cuh).
#include "reduce_dispatch.cuh"

namespace pdb200 {

int sum_mean_dispatch(bool mean, bool query, size_t* ws_out, const ReducePlan& plan, const void* x,
                      int x_dtype, int out_dtype, void* out, void* ws, size_t ws_bytes, cudaStream_t st) {
  if (!mean) {
    // registered GPU dtypes: kps/reduce_kernel.cu:302-340; bool/int32 -> int64 by InferMeta
    // (infermeta/unary.cc:4743-4747); bf16/fp16 -> fp32 direct path (kps/reduce_kernel.cu:251-265)
    PDB_REDUCE_PAIR(PDB200_FLOAT32, PDB200_FLOAT32, SumOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_FLOAT64, PDB200_FLOAT64, SumOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_BFLOAT16, PDB200_BFLOAT16, SumOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_FLOAT16, PDB200_FLOAT16, SumOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_BFLOAT16, PDB200_FLOAT32, SumOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_FLOAT16, PDB200_FLOAT32, SumOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_INT64, PDB200_INT64, SumOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_INT32, PDB200_INT64, SumOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_INT32, PDB200_INT32, SumOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_BOOL, PDB200_INT64, SumOp, kFinCast)
    // the small integers keep their dtype and wrap (kps/reduce_kernel.cu:382-400 registers int8 / uint8 / int16 too)
    PDB_REDUCE_PAIR(PDB200_INT8, PDB200_INT8, SumOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_UINT8, PDB200_UINT8, SumOp, kFinCast)
    PDB_REDUCE_PAIR(PDB200_INT16, PDB200_INT16, SumOp, kFinCast)
    return fail(PDB200_UNSUPPORTED, "sum: unsupported dtype pair x=" + std::to_string(x_dtype) +
                                        " out=" + std::to_string(out_dtype));
  }
  // kps/reduce_kernel.cu:341-355
  PDB_REDUCE_PAIR(PDB200_FLOAT32, PDB200_FLOAT32, SumOp, kFinMean)
  PDB_REDUCE_PAIR(PDB200_FLOAT64, PDB200_FLOAT64, SumOp, kFinMean)
  PDB_REDUCE_PAIR(PDB200_BFLOAT16, PDB200_BFLOAT16, SumOp, kFinMean)
  PDB_REDUCE_PAIR(PDB200_FLOAT16, PDB200_FLOAT16, SumOp, kFinMean)
  PDB_REDUCE_PAIR(PDB200_INT32, PDB200_INT32, SumOp, kFinMean)
  PDB_REDUCE_PAIR(PDB200_INT64, PDB200_INT64, SumOp, kFinMean)
  return fail(PDB200_UNSUPPORTED, "mean: unsupported dtype pair x=" + std::to_string(x_dtype) +
                                      " out=" + std::to_string(out_dtype));
}

}  // namespace pdb200

extern "C" {

int pdb200_sum_raw(const void* x, const int64_t* x_dims, int x_rank, const int64_t* axes, int n_axes,
                   int reduce_all, int x_dtype, int out_dtype, void* out, void* workspace,
                   size_t workspace_bytes, void* stream) {
  pdb200::ReducePlan plan;
  int rc = pdb200::reduce_prepare(x_dims, x_rank, axes, n_axes, reduce_all, &plan);
  if (rc != PDB200_OK) return rc;
  return pdb200::sum_mean_dispatch(false, false, nullptr, plan, x, x_dtype, out_dtype, out, workspace,
                                   workspace_bytes, reinterpret_cast<cudaStream_t>(stream));
}

int pdb200_mean_raw(const void* x, const int64_t* x_dims, int x_rank, const int64_t* axes, int n_axes,
                    int reduce_all, int x_dtype, int out_dtype, void* out, void* workspace,
                    size_t workspace_bytes, void* stream) {
  pdb200::ReducePlan plan;
  int rc = pdb200::reduce_prepare(x_dims, x_rank, axes, n_axes, reduce_all, &plan);
  if (rc != PDB200_OK) return rc;
  return pdb200::sum_mean_dispatch(true, false, nullptr, plan, x, x_dtype, out_dtype, out, workspace,
                                   workspace_bytes, reinterpret_cast<cudaStream_t>(stream));
}

}  // extern "C"
