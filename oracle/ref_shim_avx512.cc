// TEST INFRASTRUCTURE ONLY — never linked into, imported by, or called from the product path.
//
// Third `extern "C"` doorway onto the reference's own CPU PHI kernels: the one kernel of the path that the reference wrote
// with AVX-512 intrinsics, compiled on its own with -mavx512f and entered only when the CPU has the instructions:
//   fused_bias_residual_layernorm  phi::fusion::FusedLayerNormAvxKernel<float, CPUContext>
//                paddle/phi/kernels/fusion/cpu/fused_layer_norm_avx_kernel.cc:27-238 (float only)
// Two things differ from the GPU kernel it stands beside and are handled by the caller of this shim (tests/oracle_lib.py):
// the CPU kernel forms x + alpha * residual + bias (the GPU kernel (x + bias) + alpha * residual), and its `variance`
// output holds 1 / sqrt(var + epsilon).
#include "paddle/phi/core/kernel_registry.h"

#undef PD_REGISTER_KERNEL
#define PD_REGISTER_KERNEL(name, backend, layout, fn, ...)                \
  [[maybe_unused]] static void pdb200_unused_reg_##name##_##backend(      \
      const ::phi::KernelKey& kernel_key, ::phi::Kernel* kernel)

#include "paddle/phi/kernels/fusion/cpu/fused_layer_norm_avx_kernel.cc"  // NOLINT (the kernel template lives in the .cc)

#include "ref_shim_common.h"

extern "C" {

int phi_ref_have_avx512() { return __builtin_cpu_supports("avx512f") ? 1 : 0; }

// x [rows][cols] float; bias / residual / norm_weight / norm_bias may be NULL.  out always; residual_out when residual is
// given; mean / rstd when a norm parameter is given (the kernel's own rule).
int phi_ref_fused_layer_norm(const float* x, int64_t rows, int64_t cols, const float* bias, const float* residual,
                             const float* norm_weight, const float* norm_bias, float epsilon, float residual_alpha, float* out,
                             float* residual_out, float* mean, float* rstd) {
  if (!phi_ref_have_avx512()) {
    g_last_error = "phi_ref_fused_layer_norm: this CPU has no AVX-512";
    return 3;
  }
  return Guard([&] {
    auto& ctx = GetCtx().ctx;
    const int64_t d2[2] = {rows, cols}, d1[1] = {cols}, dr[1] = {rows};
    auto xt = Wrap(x, d2, 2, kFloat32);
    auto ot = Wrap(out, d2, 2, kFloat32);
    auto rot = Wrap(residual_out ? residual_out : out, d2, 2, kFloat32);
    std::vector<float> scratch(static_cast<size_t>(rows) * 2);
    auto mt = Wrap(mean ? mean : scratch.data(), dr, 1, kFloat32);
    auto vt = Wrap(rstd ? rstd : scratch.data() + rows, dr, 1, kFloat32);
    paddle::optional<phi::DenseTensor> ob, ores, ow, onb;
    if (bias) ob = Wrap(bias, d1, 1, kFloat32);
    if (residual) ores = Wrap(residual, d2, 2, kFloat32);
    if (norm_weight) ow = Wrap(norm_weight, d1, 1, kFloat32);
    if (norm_bias) onb = Wrap(norm_bias, d1, 1, kFloat32);
    phi::DenseTensor o2 = ot, r2 = rot, m2 = mt, v2 = vt;
    phi::fusion::FusedLayerNormAvxKernel<float, phi::CPUContext>(ctx, xt, ob, ores, ow, onb, epsilon, residual_alpha, 1, -1.0f, 0, 0.0f,
                                                                 0.0f, &o2, &r2, &m2, &v2);
    auto back = [](const phi::DenseTensor& t, float* dst) {
      if (dst != nullptr && t.data() != dst) std::memcpy(dst, t.data(), static_cast<size_t>(t.numel()) * sizeof(float));
    };
    back(o2, out);
    if (residual && residual_out) back(r2, residual_out);
    back(m2, mean);
    back(v2, rstd);
  });
}

}  // extern "C"
