// elementwise_more.cu — subtract / divide / maximum / minimum entry points: the other callers of
// funcs::BroadcastKernel on the same engine (kernels in elementwise_impl.cuh).
#include "elementwise_impl.cuh"

namespace pdb200 {
#define B(dt) (1u << (dt))
// kps/elementwise_kernel.cu:369-381 (subtract), :338-353 (divide), :183-202 (maximum, minimum); complex is out of scope
constexpr uint32_t kSubTypes = B(PDB200_FLOAT32) | B(PDB200_FLOAT64) | B(PDB200_INT16) | B(PDB200_INT32) |
                               B(PDB200_INT64) | B(PDB200_FLOAT16) | B(PDB200_BFLOAT16);
constexpr uint32_t kDivTypes = B(PDB200_FLOAT32) | B(PDB200_FLOAT64) | B(PDB200_INT8) | B(PDB200_UINT8) |
                               B(PDB200_INT16) | B(PDB200_INT32) | B(PDB200_INT64) | B(PDB200_FLOAT16) |
                               B(PDB200_BFLOAT16);
constexpr uint32_t kMaxMinTypes = B(PDB200_FLOAT32) | B(PDB200_FLOAT64) | B(PDB200_INT32) | B(PDB200_INT64) |
                                  B(PDB200_FLOAT16) | B(PDB200_BFLOAT16);
#undef B
}  // namespace pdb200

extern "C" {

int pdb200_subtract(const void* x, const int64_t* x_dims, int x_rank, const void* y, const int64_t* y_dims,
                    int y_rank, void* out, int dtype, void* stream) {
  return pdb200::elementwise_entry<pdb200::SubOp, pdb200::kSubTypes>("subtract", x, x_dims, x_rank, y, y_dims, y_rank,
                                                                      out, dtype, stream);
}
int pdb200_divide(const void* x, const int64_t* x_dims, int x_rank, const void* y, const int64_t* y_dims,
                  int y_rank, void* out, int dtype, void* stream) {
  return pdb200::elementwise_entry<pdb200::DivOp, pdb200::kDivTypes>("divide", x, x_dims, x_rank, y, y_dims, y_rank,
                                                                      out, dtype, stream);
}
int pdb200_maximum(const void* x, const int64_t* x_dims, int x_rank, const void* y, const int64_t* y_dims,
                   int y_rank, void* out, int dtype, void* stream) {
  return pdb200::elementwise_entry<pdb200::MaximumOp, pdb200::kMaxMinTypes>("maximum", x, x_dims, x_rank, y, y_dims,
                                                                             y_rank, out, dtype, stream);
}
int pdb200_minimum(const void* x, const int64_t* x_dims, int x_rank, const void* y, const int64_t* y_dims,
                   int y_rank, void* out, int dtype, void* stream) {
  return pdb200::elementwise_entry<pdb200::MinimumOp, pdb200::kMaxMinTypes>("minimum", x, x_dims, x_rank, y, y_dims,
                                                                             y_rank, out, dtype, stream);
}

}  // extern "C"
