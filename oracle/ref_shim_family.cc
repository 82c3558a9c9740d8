// TEST INFRASTRUCTURE ONLY — never linked into, imported by, or called from the product path.
//
// Second `extern "C"` doorway onto the reference's own CPU PHI kernels (see ref_shim.cc for the first):
// the rest of the elementwise family.  No restated algorithm here; every call lands in a
// kernel template compiled from /root/reference:
//   compare      phi::LessThan/LessEqual/GreaterThan/GreaterEqual/Equal/NotEqualKernel<T,CPUContext>
//                paddle/phi/kernels/impl/compare_kernel_impl.h:44-76, cpu/compare_kernel.cc:26-44
//   logical      phi::LogicalAnd/Or/Xor/NotKernel<T,CPUContext>  paddle/phi/kernels/cpu/logical_kernel.cc:27-91
//   bitwise      phi::BitwiseAnd/Or/Xor/NotKernel<T,CPUContext>  paddle/phi/kernels/cpu/bitwise_kernel.cc:27-92
//   scale        phi::ScaleKernel<T,CPUContext>                  paddle/phi/kernels/cpu/scale_kernel.cc:27-48
//   cast         phi::CastKernel<T,CPUContext>                   paddle/phi/kernels/cpu/cast_kernel.cc, cast_impl.h
//   where        phi::WhereKernel<T,CPUContext>                  paddle/phi/kernels/cpu/where_kernel.cc:20-36
//   add_grad     phi::AddGradKernel<T,CPUContext>                paddle/phi/kernels/cpu/elementwise_add_grad_kernel.cc:25-62
//   cross_entropy_with_softmax  phi::CrossEntropyWithSoftmaxKernel<T,CPUContext>
//                paddle/phi/kernels/cpu/cross_entropy_kernel.cc:73-104 (+ funcs/cross_entropy.cc, compiled as its own TU)
//   cross_entropy_with_softmax_grad  phi::CrossEntropyWithSoftmaxGradKernel<T,CPUContext>
//                paddle/phi/kernels/cpu/cross_entropy_grad_kernel.cc:26-215
//   log_softmax  phi::LogSoftmaxKernel<T,CPUContext>             paddle/phi/kernels/cpu/log_softmax_kernel.cc:40-127
//   log_softmax_grad  phi::LogSoftmaxGradKernel<T,CPUContext>    paddle/phi/kernels/cpu/log_softmax_grad_kernel.cc:30-84
//   sum_grad / mean_grad  phi::ReduceSumGradKernel / ReduceMeanGradKernel<T,CPUContext>
//                paddle/phi/kernels/cpu/reduce_sum_grad_kernel.cc:26-43, cpu/reduce_mean_grad_kernel.cc
//   max_grad / min_grad / prod_grad  phi::ReduceMaxGradKernel / ReduceMinGradKernel / ProdGradKernel<T,CPUContext>
//                paddle/phi/kernels/cpu/reduce_max_grad_kernel.cc, reduce_min_grad_kernel.cc, prod_grad_kernel.cc
//   multiply_grad phi::MultiplyGradKernel<T,CPUContext>          paddle/phi/kernels/cpu/elementwise_multiply_grad_kernel.cc:26-39
#include "paddle/common/flags.h"
#include "paddle/phi/common/bfloat16.h"
#include "paddle/phi/common/float16.h"
#include "paddle/phi/common/scalar.h"
#include "paddle/phi/core/kernel_registry.h"

// registrations at the end of the kernel TUs become unused functions (see ref_shim.cc)
#undef PD_REGISTER_KERNEL
#define PD_REGISTER_KERNEL(name, backend, layout, fn, ...)                \
  [[maybe_unused]] static void pdb200_unused_reg_##name##_##backend(      \
      const ::phi::KernelKey& kernel_key, ::phi::Kernel* kernel)

#include "paddle/phi/kernels/cpu/bitwise_kernel.cc"  // NOLINT (kernel templates live in the .cc)
#include "paddle/phi/kernels/cpu/cast_kernel.cc"     // NOLINT
#include "paddle/phi/kernels/cpu/compare_kernel.cc"  // NOLINT
#include "paddle/phi/kernels/cpu/cross_entropy_grad_kernel.cc"  // NOLINT
#include "paddle/phi/kernels/cpu/cross_entropy_kernel.cc"  // NOLINT
#include "paddle/phi/kernels/cpu/elementwise_add_grad_kernel.cc"       // NOLINT
#include "paddle/phi/kernels/cpu/elementwise_multiply_grad_kernel.cc"  // NOLINT
#include "paddle/phi/kernels/cpu/log_softmax_grad_kernel.cc"  // NOLINT
#include "paddle/phi/kernels/cpu/log_softmax_kernel.cc"  // NOLINT
#include "paddle/phi/kernels/cpu/logical_kernel.cc"  // NOLINT
#include "paddle/phi/kernels/cpu/prod_grad_kernel.cc"         // NOLINT
#include "paddle/phi/kernels/cpu/reduce_max_grad_kernel.cc"   // NOLINT
#include "paddle/phi/kernels/cpu/reduce_mean_grad_kernel.cc"  // NOLINT
#include "paddle/phi/kernels/cpu/reduce_min_grad_kernel.cc"   // NOLINT
#include "paddle/phi/kernels/cpu/reduce_sum_grad_kernel.cc"   // NOLINT
#include "paddle/phi/kernels/cpu/scale_kernel.cc"    // NOLINT
#include "paddle/phi/kernels/cpu/where_kernel.cc"    // NOLINT

#include "ref_shim_common.h"

namespace {

using phi::CPUContext;
using phi::DenseTensor;
using bf16 = phi::dtype::bfloat16;
using fp16 = phi::dtype::float16;

template <typename T>
void CompareT(int op, const DenseTensor& x, const DenseTensor& y, DenseTensor* out) {
  auto& ctx = GetCtx().ctx;
  switch (op) {
    case 0: phi::LessThanKernel<T, CPUContext>(ctx, x, y, out); break;
    case 1: phi::LessEqualKernel<T, CPUContext>(ctx, x, y, out); break;
    case 2: phi::GreaterThanKernel<T, CPUContext>(ctx, x, y, out); break;
    case 3: phi::GreaterEqualKernel<T, CPUContext>(ctx, x, y, out); break;
    case 4: phi::EqualKernel<T, CPUContext>(ctx, x, y, out); break;
    case 5: phi::NotEqualKernel<T, CPUContext>(ctx, x, y, out); break;
    default: throw std::runtime_error("bad compare op");
  }
}
template <typename T>
void LogicalT(int op, const DenseTensor& x, const DenseTensor& y, DenseTensor* out) {
  auto& ctx = GetCtx().ctx;
  switch (op) {
    case 0: phi::LogicalAndKernel<T, CPUContext>(ctx, x, y, out); break;
    case 1: phi::LogicalOrKernel<T, CPUContext>(ctx, x, y, out); break;
    case 2: phi::LogicalXorKernel<T, CPUContext>(ctx, x, y, out); break;
    default: throw std::runtime_error("bad logical op");
  }
}
template <typename T>
void BitwiseT(int op, const DenseTensor& x, const DenseTensor& y, DenseTensor* out) {
  auto& ctx = GetCtx().ctx;
  switch (op) {
    case 0: phi::BitwiseAndKernel<T, CPUContext>(ctx, x, y, out); break;
    case 1: phi::BitwiseOrKernel<T, CPUContext>(ctx, x, y, out); break;
    case 2: phi::BitwiseXorKernel<T, CPUContext>(ctx, x, y, out); break;
    default: throw std::runtime_error("bad bitwise op");
  }
}

}  // namespace

extern "C" {

// op: 0 less_than, 1 less_equal, 2 greater_than, 3 greater_equal, 4 equal, 5 not_equal.  out: bool.
int phi_ref_compare(int op, int dtype, const void* x, const int64_t* xdims, int xrank, const void* y,
                    const int64_t* ydims, int yrank, const int64_t* odims, int orank, void* out) {
  return Guard([&] {
    auto xt = Wrap(x, xdims, xrank, dtype);
    auto yt = Wrap(y, ydims, yrank, dtype);
    auto ot = Wrap(out, odims, orank, kBool);
    switch (dtype) {  // cpu/compare_kernel.cc:121-141
      case kBool: CompareT<bool>(op, xt, yt, &ot); break;
      case kUInt8: CompareT<uint8_t>(op, xt, yt, &ot); break;
      case kInt8: CompareT<int8_t>(op, xt, yt, &ot); break;
      case kInt16: CompareT<int16_t>(op, xt, yt, &ot); break;
      case kInt32: CompareT<int32_t>(op, xt, yt, &ot); break;
      case kInt64: CompareT<int64_t>(op, xt, yt, &ot); break;
      case kFloat16: CompareT<fp16>(op, xt, yt, &ot); break;
      case kBFloat16: CompareT<bf16>(op, xt, yt, &ot); break;
      case kFloat32: CompareT<float>(op, xt, yt, &ot); break;
      case kFloat64: CompareT<double>(op, xt, yt, &ot); break;
      default: throw std::runtime_error("phi_ref_compare: bad dtype");
    }
  });
}

// op: 0 and, 1 or, 2 xor.  out: bool.
int phi_ref_logical(int op, int dtype, const void* x, const int64_t* xdims, int xrank, const void* y,
                    const int64_t* ydims, int yrank, const int64_t* odims, int orank, void* out) {
  return Guard([&] {
    auto xt = Wrap(x, xdims, xrank, dtype);
    auto yt = Wrap(y, ydims, yrank, dtype);
    auto ot = Wrap(out, odims, orank, kBool);
    switch (dtype) {  // cpu/logical_kernel.cc:93-111 (no fp16 / bf16 / uint8 on CPU)
      case kBool: LogicalT<bool>(op, xt, yt, &ot); break;
      case kInt8: LogicalT<int8_t>(op, xt, yt, &ot); break;
      case kInt16: LogicalT<int16_t>(op, xt, yt, &ot); break;
      case kInt32: LogicalT<int32_t>(op, xt, yt, &ot); break;
      case kInt64: LogicalT<int64_t>(op, xt, yt, &ot); break;
      case kFloat32: LogicalT<float>(op, xt, yt, &ot); break;
      case kFloat64: LogicalT<double>(op, xt, yt, &ot); break;
      default: throw std::runtime_error("phi_ref_logical: dtype not registered on CPU");
    }
  });
}

// op: 0 and, 1 or, 2 xor.  out has the input dtype.
int phi_ref_bitwise(int op, int dtype, const void* x, const int64_t* xdims, int xrank, const void* y,
                    const int64_t* ydims, int yrank, const int64_t* odims, int orank, void* out) {
  return Guard([&] {
    auto xt = Wrap(x, xdims, xrank, dtype);
    auto yt = Wrap(y, ydims, yrank, dtype);
    auto ot = Wrap(out, odims, orank, dtype);
    switch (dtype) {
      case kBool: BitwiseT<bool>(op, xt, yt, &ot); break;
      case kUInt8: BitwiseT<uint8_t>(op, xt, yt, &ot); break;
      case kInt8: BitwiseT<int8_t>(op, xt, yt, &ot); break;
      case kInt16: BitwiseT<int16_t>(op, xt, yt, &ot); break;
      case kInt32: BitwiseT<int32_t>(op, xt, yt, &ot); break;
      case kInt64: BitwiseT<int64_t>(op, xt, yt, &ot); break;
      default: throw std::runtime_error("phi_ref_bitwise: dtype not registered on CPU");
    }
  });
}

// right: 0 bitwise_left_shift, 1 bitwise_right_shift (cpu/bitwise_kernel.cc:40-78)
int phi_ref_bitwise_shift(int right, int is_arithmetic, int dtype, const void* x, const int64_t* xdims, int xrank,
                          const void* y, const int64_t* ydims, int yrank, const int64_t* odims, int orank, void* out) {
  return Guard([&] {
    auto& ctx = GetCtx().ctx;
    auto xt = Wrap(x, xdims, xrank, dtype);
    auto yt = Wrap(y, ydims, yrank, dtype);
    auto ot = Wrap(out, odims, orank, dtype);
#define SH_CASE(T)                                                                                      \
  if (right) phi::BitwiseRightShiftKernel<T, CPUContext>(ctx, xt, yt, is_arithmetic != 0, &ot);          \
  else phi::BitwiseLeftShiftKernel<T, CPUContext>(ctx, xt, yt, is_arithmetic != 0, &ot);                 \
  break;
    switch (dtype) {
      case kUInt8: SH_CASE(uint8_t)
      case kInt8: SH_CASE(int8_t)
      case kInt16: SH_CASE(int16_t)
      case kInt32: SH_CASE(int32_t)
      case kInt64: SH_CASE(int64_t)
      default: throw std::runtime_error("phi_ref_bitwise_shift: dtype not registered on CPU");
    }
#undef SH_CASE
  });
}

// which: 0 logical_not (out bool), 1 bitwise_not (out dtype)
int phi_ref_not(int which, int dtype, const void* x, int64_t numel, void* out) {
  return Guard([&] {
    auto& ctx = GetCtx().ctx;
    auto xt = Wrap(x, &numel, 1, dtype);
    auto ot = Wrap(out, &numel, 1, which == 0 ? kBool : dtype);
    if (which == 0) {
      switch (dtype) {
        case kBool: phi::LogicalNotKernel<bool, CPUContext>(ctx, xt, &ot); break;
        case kInt8: phi::LogicalNotKernel<int8_t, CPUContext>(ctx, xt, &ot); break;
        case kInt16: phi::LogicalNotKernel<int16_t, CPUContext>(ctx, xt, &ot); break;
        case kInt32: phi::LogicalNotKernel<int32_t, CPUContext>(ctx, xt, &ot); break;
        case kInt64: phi::LogicalNotKernel<int64_t, CPUContext>(ctx, xt, &ot); break;
        case kFloat32: phi::LogicalNotKernel<float, CPUContext>(ctx, xt, &ot); break;
        case kFloat64: phi::LogicalNotKernel<double, CPUContext>(ctx, xt, &ot); break;
        default: throw std::runtime_error("phi_ref_not: dtype not registered on CPU");
      }
    } else {
      switch (dtype) {
        case kBool: phi::BitwiseNotKernel<bool, CPUContext>(ctx, xt, &ot); break;
        case kUInt8: phi::BitwiseNotKernel<uint8_t, CPUContext>(ctx, xt, &ot); break;
        case kInt8: phi::BitwiseNotKernel<int8_t, CPUContext>(ctx, xt, &ot); break;
        case kInt16: phi::BitwiseNotKernel<int16_t, CPUContext>(ctx, xt, &ot); break;
        case kInt32: phi::BitwiseNotKernel<int32_t, CPUContext>(ctx, xt, &ot); break;
        case kInt64: phi::BitwiseNotKernel<int64_t, CPUContext>(ctx, xt, &ot); break;
        default: throw std::runtime_error("phi_ref_not: dtype not registered on CPU");
      }
    }
  });
}

int phi_ref_scale(int dtype, const void* x, int64_t numel, double scale, double bias, int bias_after_scale, void* out) {
  return Guard([&] {
    auto& ctx = GetCtx().ctx;
    auto xt = Wrap(x, &numel, 1, dtype);
    auto ot = Wrap(out, &numel, 1, dtype);
    // paddle.scale's attributes are Python floats, which pybind turns into Scalar(double)
    // (paddle/fluid/pybind/eager_utils.cc:2297-2299); the kernel then takes scale.to<T>()
    phi::Scalar s(scale), b(bias);
    const bool after = bias_after_scale != 0;
    switch (dtype) {  // cpu/scale_kernel.cc:52-66
      case kBool: phi::ScaleKernel<bool, CPUContext>(ctx, xt, s, b, after, &ot); break;
      case kUInt8: phi::ScaleKernel<uint8_t, CPUContext>(ctx, xt, s, b, after, &ot); break;
      case kInt8: phi::ScaleKernel<int8_t, CPUContext>(ctx, xt, s, b, after, &ot); break;
      case kInt16: phi::ScaleKernel<int16_t, CPUContext>(ctx, xt, s, b, after, &ot); break;
      case kInt32: phi::ScaleKernel<int32_t, CPUContext>(ctx, xt, s, b, after, &ot); break;
      case kInt64: phi::ScaleKernel<int64_t, CPUContext>(ctx, xt, s, b, after, &ot); break;
      case kFloat16: phi::ScaleKernel<fp16, CPUContext>(ctx, xt, s, b, after, &ot); break;
      case kBFloat16: phi::ScaleKernel<bf16, CPUContext>(ctx, xt, s, b, after, &ot); break;
      case kFloat32: phi::ScaleKernel<float, CPUContext>(ctx, xt, s, b, after, &ot); break;
      case kFloat64: phi::ScaleKernel<double, CPUContext>(ctx, xt, s, b, after, &ot); break;
      default: throw std::runtime_error("phi_ref_scale: bad dtype");
    }
  });
}

int phi_ref_cast(int x_dtype, int out_dtype, const void* x, int64_t numel, void* out) {
  return Guard([&] {
    auto& ctx = GetCtx().ctx;
    auto xt = Wrap(x, &numel, 1, x_dtype);
    auto ot = Wrap(out, &numel, 1, out_dtype);
    const auto od = ToPhi(out_dtype);
    switch (x_dtype) {  // cpu/cast_kernel.cc registration list
      case kBool: phi::CastKernel<bool, CPUContext>(ctx, xt, od, &ot); break;
      case kUInt8: phi::CastKernel<uint8_t, CPUContext>(ctx, xt, od, &ot); break;
      case kInt8: phi::CastKernel<int8_t, CPUContext>(ctx, xt, od, &ot); break;
      case kInt16: phi::CastKernel<int16_t, CPUContext>(ctx, xt, od, &ot); break;
      case kInt32: phi::CastKernel<int32_t, CPUContext>(ctx, xt, od, &ot); break;
      case kInt64: phi::CastKernel<int64_t, CPUContext>(ctx, xt, od, &ot); break;
      case kFloat16: phi::CastKernel<fp16, CPUContext>(ctx, xt, od, &ot); break;
      case kBFloat16: phi::CastKernel<bf16, CPUContext>(ctx, xt, od, &ot); break;
      case kFloat32: phi::CastKernel<float, CPUContext>(ctx, xt, od, &ot); break;
      case kFloat64: phi::CastKernel<double, CPUContext>(ctx, xt, od, &ot); break;
      default: throw std::runtime_error("phi_ref_cast: bad dtype");
    }
  });
}

int phi_ref_where(int dtype, const void* cond, const void* x, const void* y, int64_t numel, void* out) {
  return Guard([&] {
    auto& ctx = GetCtx().ctx;
    auto ct = Wrap(cond, &numel, 1, kBool);
    auto xt = Wrap(x, &numel, 1, dtype);
    auto yt = Wrap(y, &numel, 1, dtype);
    auto ot = Wrap(out, &numel, 1, dtype);
    switch (dtype) {  // cpu/where_kernel.cc:40-41
      case kInt32: phi::WhereKernel<int32_t, CPUContext>(ctx, ct, xt, yt, &ot); break;
      case kInt64: phi::WhereKernel<int64_t, CPUContext>(ctx, ct, xt, yt, &ot); break;
      case kFloat32: phi::WhereKernel<float, CPUContext>(ctx, ct, xt, yt, &ot); break;
      case kFloat64: phi::WhereKernel<double, CPUContext>(ctx, ct, xt, yt, &ot); break;
      default: throw std::runtime_error("phi_ref_where: dtype not registered on CPU");
    }
  });
}

// op: 0 add_grad, 1 multiply_grad.  dx / dy may be NULL (gradient not wanted).  x / y are read by multiply_grad only,
// but must be valid buffers of their shapes for both (the kernels take the tensors for their dims).
int phi_ref_elementwise_grad(int op, int dtype, const void* x, const int64_t* xdims, int xrank, const void* y,
                             const int64_t* ydims, int yrank, const void* dout, const int64_t* odims, int orank,
                             void* dx, void* dy) {
  return Guard([&] {
    auto& ctx = GetCtx().ctx;
    auto xt = Wrap(x, xdims, xrank, dtype);
    auto yt = Wrap(y, ydims, yrank, dtype);
    auto gt = Wrap(dout, odims, orank, dtype);
    DenseTensor dxt, dyt;
    if (dx) dxt = Wrap(dx, xdims, xrank, dtype);
    if (dy) dyt = Wrap(dy, ydims, yrank, dtype);
    DenseTensor* pdx = dx ? &dxt : nullptr;
    DenseTensor* pdy = dy ? &dyt : nullptr;
#define GRAD_CASE(T)                                                                             \
  if (op == 0) phi::AddGradKernel<T, CPUContext>(ctx, xt, yt, gt, -1, pdx, pdy);                 \
  else phi::MultiplyGradKernel<T, CPUContext>(ctx, xt, yt, gt, -1, pdx, pdy);                    \
  break;
    switch (dtype) {
      case kFloat32: GRAD_CASE(float)
      case kFloat64: GRAD_CASE(double)
      case kInt32: GRAD_CASE(int32_t)
      case kInt64: GRAD_CASE(int64_t)
      case kBFloat16: GRAD_CASE(bf16)
      default: throw std::runtime_error("phi_ref_elementwise_grad: dtype not registered on CPU");
    }
#undef GRAD_CASE
    // the kernels may have re-allocated an output (Alloc on a borrowed buffer keeps it when the size fits)
    if (dx && dxt.data() != dx) std::memcpy(dx, dxt.data(), static_cast<size_t>(dxt.numel()) * phi::SizeOf(dxt.dtype()));
    if (dy && dyt.data() != dy) std::memcpy(dy, dyt.data(), static_cast<size_t>(dyt.numel()) * phi::SizeOf(dyt.dtype()));
  });
}

// label: int64 class indices (the logits shape with 1 along axis) or, with soft_label, distributions of the logits dtype
int phi_ref_cross_entropy_with_softmax(int dtype, const void* logits, const int64_t* dims, int rank, const void* label,
                                       const int64_t* ldims, int soft_label, int use_softmax, int ignore_index, int axis,
                                       void* softmax, void* loss) {
  return Guard([&] {
    auto& ctx = GetCtx().ctx;
    auto xt = Wrap(logits, dims, rank, dtype);
    auto lt = Wrap(label, ldims, rank, soft_label ? dtype : kInt64);
    auto st = Wrap(softmax, dims, rank, dtype);
    std::vector<int64_t> od(dims, dims + rank);
    od[axis < 0 ? axis + rank : axis] = 1;
    auto ot = Wrap(loss, od.data(), rank, dtype);
    switch (dtype) {
      case kFloat32:
        phi::CrossEntropyWithSoftmaxKernel<float, CPUContext>(ctx, xt, lt, soft_label != 0, use_softmax != 0, true, ignore_index, axis, &st, &ot);
        break;
      case kFloat64:
        phi::CrossEntropyWithSoftmaxKernel<double, CPUContext>(ctx, xt, lt, soft_label != 0, use_softmax != 0, true, ignore_index, axis, &st, &ot);
        break;
      default: throw std::runtime_error("phi_ref_cross_entropy_with_softmax: dtype not registered on CPU");
    }
    if (st.data() != softmax) std::memcpy(softmax, st.data(), static_cast<size_t>(st.numel()) * phi::SizeOf(st.dtype()));
    if (ot.data() != loss) std::memcpy(loss, ot.data(), static_cast<size_t>(ot.numel()) * phi::SizeOf(ot.dtype()));
  });
}

int phi_ref_log_softmax(int dtype, const void* x, const int64_t* dims, int rank, int axis, void* out) {
  return Guard([&] {
    auto& ctx = GetCtx().ctx;
    auto xt = Wrap(x, dims, rank, dtype);
    auto ot = Wrap(out, dims, rank, dtype);
    switch (dtype) {
      case kFloat32: phi::LogSoftmaxKernel<float, CPUContext>(ctx, xt, axis, &ot); break;
      case kFloat64: phi::LogSoftmaxKernel<double, CPUContext>(ctx, xt, axis, &ot); break;
      default: throw std::runtime_error("phi_ref_log_softmax: dtype not registered on CPU");
    }
  });
}

// op: 0 sum_grad, 1 mean_grad.  out_grad is passed with the kept-dim (or dropped-dim) shape the forward produced.
int phi_ref_reduce_grad(int op, int dtype, const void* x, const int64_t* xdims, int xrank, const void* out_grad,
                        const int64_t* gdims, int grank, const int64_t* axes, int n_axes, int keep_dim, int reduce_all,
                        void* x_grad) {
  return Guard([&] {
    auto& ctx = GetCtx().ctx;
    auto xt = Wrap(x, xdims, xrank, dtype);
    auto gt = Wrap(out_grad, gdims, grank, dtype);
    auto dxt = Wrap(x_grad, xdims, xrank, dtype);
    phi::IntArray ax(std::vector<int64_t>(axes, axes + n_axes));
#define RG_SUM(T) phi::ReduceSumGradKernel<T, CPUContext>(ctx, xt, gt, ax, keep_dim != 0, reduce_all != 0, &dxt); break;
#define RG_BOTH(T)                                                                                              \
  if (op == 0) phi::ReduceSumGradKernel<T, CPUContext>(ctx, xt, gt, ax, keep_dim != 0, reduce_all != 0, &dxt);   \
  else phi::ReduceMeanGradKernel<T, CPUContext>(ctx, xt, gt, ax, keep_dim != 0, reduce_all != 0, &dxt);          \
  break;
    if (op != 0 && (dtype == kBFloat16 || dtype == kFloat16))
      throw std::runtime_error("phi_ref_reduce_grad: mean_grad has no 16-bit kernel on CPU");
    switch (dtype) {
      case kFloat32: RG_BOTH(float)
      case kFloat64: RG_BOTH(double)
      case kInt32: RG_BOTH(int32_t)
      case kInt64: RG_BOTH(int64_t)
      case kBFloat16: RG_SUM(bf16)
      case kFloat16: RG_SUM(fp16)
      default: throw std::runtime_error("phi_ref_reduce_grad: dtype not registered on CPU");
    }
#undef RG_SUM
#undef RG_BOTH
    if (dxt.data() != x_grad) std::memcpy(x_grad, dxt.data(), static_cast<size_t>(dxt.numel()) * phi::SizeOf(dxt.dtype()));
  });
}

// op: 0 max_grad, 1 min_grad, 2 prod_grad (CPU registrations: float, double, int, int64_t).  out / out_grad carry the
// shape the forward produced.
int phi_ref_select_grad(int op, int dtype, const void* x, const int64_t* xdims, int xrank, const void* out,
                        const void* out_grad, const int64_t* gdims, int grank, const int64_t* axes, int n_axes,
                        int keep_dim, int reduce_all, void* x_grad) {
  return Guard([&] {
    auto& ctx = GetCtx().ctx;
    auto xt = Wrap(x, xdims, xrank, dtype);
    auto yt = Wrap(out, gdims, grank, dtype);
    auto gt = Wrap(out_grad, gdims, grank, dtype);
    auto dxt = Wrap(x_grad, xdims, xrank, dtype);
    phi::IntArray ax(std::vector<int64_t>(axes, axes + n_axes));
#define SG(T)                                                                                                        \
  if (op == 0) phi::ReduceMaxGradKernel<T, CPUContext>(ctx, xt, yt, gt, ax, keep_dim != 0, reduce_all != 0, &dxt);      \
  else if (op == 1) phi::ReduceMinGradKernel<T, CPUContext>(ctx, xt, yt, gt, ax, keep_dim != 0, reduce_all != 0, &dxt); \
  else phi::ProdGradKernel<T, CPUContext>(ctx, xt, yt, gt, ax, keep_dim != 0, reduce_all != 0, &dxt);                   \
  break;
    switch (dtype) {
      case kFloat32: SG(float)
      case kFloat64: SG(double)
      case kInt32: SG(int32_t)
      case kInt64: SG(int64_t)
      default: throw std::runtime_error("phi_ref_select_grad: dtype not registered on CPU");
    }
#undef SG
    if (dxt.data() != x_grad) std::memcpy(x_grad, dxt.data(), static_cast<size_t>(dxt.numel()) * phi::SizeOf(dxt.dtype()));
  });
}

int phi_ref_log_softmax_grad(int dtype, const void* out, const void* out_grad, const int64_t* dims, int rank, int axis,
                             void* x_grad) {
  return Guard([&] {
    auto yt = Wrap(out, dims, rank, dtype);
    auto gt = Wrap(out_grad, dims, rank, dtype);
    auto xt = Wrap(x_grad, dims, rank, dtype);
    auto& ctx = GetCtx().ctx;
    switch (dtype) {
      case kFloat32: phi::LogSoftmaxGradKernel<float, CPUContext>(ctx, yt, gt, axis, &xt); break;
      case kFloat64: phi::LogSoftmaxGradKernel<double, CPUContext>(ctx, yt, gt, axis, &xt); break;
      default: throw std::runtime_error("phi_ref_log_softmax_grad: dtype not registered on CPU");
    }
    if (xt.data() != x_grad) std::memcpy(x_grad, xt.data(), static_cast<size_t>(xt.numel()) * phi::SizeOf(xt.dtype()));
  });
}


// label as for the forward; loss_grad has the logits shape with 1 along axis
int phi_ref_cross_entropy_with_softmax_grad(int dtype, const void* label, const int64_t* ldims, const void* softmax, const void* loss_grad,
                                            const int64_t* dims, int rank, int soft_label, int use_softmax, int ignore_index, int axis,
                                            void* logits_grad) {
  return Guard([&] {
    auto& ctx = GetCtx().ctx;
    auto lt = Wrap(label, ldims, rank, soft_label ? dtype : kInt64);
    auto st = Wrap(softmax, dims, rank, dtype);
    std::vector<int64_t> od(dims, dims + rank);
    od[axis < 0 ? axis + rank : axis] = 1;
    auto gt = Wrap(loss_grad, od.data(), rank, dtype);
    auto dxt = Wrap(logits_grad, dims, rank, dtype);
    switch (dtype) {
      case kFloat32:
        phi::CrossEntropyWithSoftmaxGradKernel<float, CPUContext>(ctx, lt, st, gt, soft_label != 0, use_softmax != 0, true, ignore_index, axis, &dxt);
        break;
      case kFloat64:
        phi::CrossEntropyWithSoftmaxGradKernel<double, CPUContext>(ctx, lt, st, gt, soft_label != 0, use_softmax != 0, true, ignore_index, axis, &dxt);
        break;
      default: throw std::runtime_error("phi_ref_cross_entropy_with_softmax_grad: dtype not registered on CPU");
    }
    if (dxt.data() != logits_grad) std::memcpy(logits_grad, dxt.data(), static_cast<size_t>(dxt.numel()) * phi::SizeOf(dxt.dtype()));
  });
}

}  // extern "C"
