// elementwise.cu — add / multiply entry points (kernels in elementwise_impl.cuh).
#include "elementwise_impl.cuh"

namespace pdb200 {
constexpr uint32_t kAllTypes = 0x3ffu;  // every pdb200_dtype: kps/elementwise_kernel.cu:304-318 (add), multiply
}


extern "C" {

int pdb200_add(const void* x, const int64_t* x_dims, int x_rank, const void* y, const int64_t* y_dims,
               int y_rank, void* out, int dtype, void* stream) {
  return pdb200::elementwise_entry<pdb200::AddOp, pdb200::kAllTypes>("add", x, x_dims, x_rank, y, y_dims, y_rank, out, dtype, stream);
}

int pdb200_multiply(const void* x, const int64_t* x_dims, int x_rank, const void* y,
                    const int64_t* y_dims, int y_rank, void* out, int dtype, void* stream) {
  return pdb200::elementwise_entry<pdb200::MulOp, pdb200::kAllTypes>("multiply", x, x_dims, x_rank, y, y_dims, y_rank, out, dtype, stream);
}

int pdb200_broadcast_shape(const int64_t* x_dims, int x_rank, const int64_t* y_dims, int y_rank,
                           int64_t* out_dims, int* out_rank) {
  pdb200::BroadcastPlan p;
  std::string err = pdb200::plan_broadcast(x_dims, x_rank, y_dims, y_rank, &p);
  if (!err.empty()) return pdb200::fail(PDB200_INVALID_ARGUMENT, err);
  for (int i = 0; i < p.out_rank; ++i) out_dims[i] = p.out_dims[i];
  *out_rank = p.out_rank;
  return PDB200_OK;
}

}  // extern "C"
