// reduce_dispatch.cuh — dtype dispatch shared by the per-op reduction translation units
// (reduce_sum.cu, reduce_minmax.cu, reduce_prod.cu).  Split across TUs only so that nvcc can
// build them in parallel.
//
// Mirrors the op wrappers of the reference: SumRawKernel / MeanRawKernel / MaxRawKernel /
// MinRawKernel / ProdKernel (paddle/phi/kernels/kps/reduce_kernel.cu:34-271,
// legacy/kps/reduce_max_kernel.cu:21-32) and phi::Reduce's cast-then-reduce rule
// (paddle/phi/kernels/gpu/reduce.h:26-80).
#pragma once

#include "reduce_impl.cuh"

namespace pdb200 {

enum FinKind { kFinCast = 0, kFinMean = 1 };

// Accumulator of a reduction that produces O: MPTypeTrait, except that 1- and 2-byte integers are summed in int32
// (shuffles move 4-byte words); the final cast back wraps, and modular sums do not depend on the width they were
// carried in, so the result equals the reference's sum in T.
template <typename O> struct RedAcc { using type = typename AccType<O>::type; };
template <> struct RedAcc<int8_t> { using type = int32_t; };
template <> struct RedAcc<uint8_t> { using type = int32_t; };
template <> struct RedAcc<int16_t> { using type = int32_t; };

// One (input T, output O) pair of one operator family.
template <typename T, typename O, template <typename> class OpT, int FIN>
struct ReduceRunner {
  using A = typename RedAcc<O>::type;  // MPTypeTrait of the dtype the reduction runs in
  using Op = OpT<A>;
  static size_t workspace(const ReducePlan& plan) { return reduce_scratch_bytes<T, A>(plan); }
  static int run(const ReducePlan& plan, const void* x, void* out, void* ws, size_t ws_bytes,
                 cudaStream_t st) {
    if (plan.passes.empty()) {
      // reduce_num == 1: every output is one input element (reduce_function.h:1065-1092).
      // Expressed as a [numel][1][1] row pass so the same kernels do the cast.
      ReducePlan p1 = plan;
      p1.passes.push_back(ReducePass{plan.numel, 1, 1});
      if constexpr (FIN == kFinMean)
        return run_reduce<T, O, Op>(p1, static_cast<const T*>(x), static_cast<O*>(out),
                                    MeanFin<A, O>{A(1)}, ws, ws_bytes, st);
      else
        return run_reduce<T, O, Op>(p1, static_cast<const T*>(x), static_cast<O*>(out), CastFin<A, O>(),
                                    ws, ws_bytes, st);
    }
    if constexpr (FIN == kFinMean)
      return run_reduce<T, O, Op>(plan, static_cast<const T*>(x), static_cast<O*>(out),
                                  MeanFin<A, O>{static_cast<A>(plan.reduce_num)}, ws, ws_bytes, st);
    else
      return run_reduce<T, O, Op>(plan, static_cast<const T*>(x), static_cast<O*>(out), CastFin<A, O>(), ws,
                                  ws_bytes, st);
  }
};

inline int reduce_prepare(const int64_t* x_dims, int x_rank, const int64_t* axes, int n_axes,
                          int reduce_all, ReducePlan* plan) {
  if (x_rank > 0 && x_dims == nullptr) return fail(PDB200_INVALID_ARGUMENT, "x_dims is NULL");
  if (n_axes > 0 && axes == nullptr) return fail(PDB200_INVALID_ARGUMENT, "axes is NULL");
  std::string err = plan_reduce(x_dims, x_rank, axes, n_axes, reduce_all != 0, plan);
  if (!err.empty()) return fail(PDB200_INVALID_ARGUMENT, err);
  return PDB200_OK;
}

// query == true: return the workspace size through *ws_out and launch nothing.
#define PDB_REDUCE_PAIR(XD, OD, OPT, FIN)                                                          \
  if (x_dtype == XD && out_dtype == OD) {                                                          \
    using R = ReduceRunner<typename CType<XD>::type, typename CType<OD>::type, OPT, FIN>;          \
    if (query) {                                                                                   \
      *ws_out = R::workspace(plan);                                                                \
      return PDB200_OK;                                                                            \
    }                                                                                              \
    if (!x || !out) return fail(PDB200_INVALID_ARGUMENT, "NULL tensor pointer");                   \
    return R::run(plan, x, out, ws, ws_bytes, st);                                                 \
  }

}  // namespace pdb200
