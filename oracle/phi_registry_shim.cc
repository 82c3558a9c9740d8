// TEST INFRASTRUCTURE ONLY.  A minimal host for Paddle's kernel REGISTRY and DISPATCH, built from the reference's
// own sources (oracle/build_registry.sh): KernelFactory, KernelRegistrar, DenseTensor, DeviceContext and what they link
// against — "phi_min" of SURVEY.md §7 step 2.  It exists so that
//   * tests/test_phi_registration.py can dlopen the Paddle-side binding (paddlepaddle_b200/plugin/phi_kernels.cc,
//     compiled against the reference headers) and check that its PD_REGISTER_KERNEL static initialisers really do put
//     the B200 kernels under Paddle's own (name, backend, layout, dtype) keys, replacing what was there before;
//   * tests/test_gpu_phi_dispatch.py can CALL those kernels on a B200 exactly as Paddle's generated eager API does
//     (paddle/phi/api/generator/api_base.py:1327-1354):
//         KernelFactory::Instance().SelectKernelOrThrowError(name, {backend, ALL_LAYOUT, dtype})
//         -> kernel.GetVariadicKernelFn<signature>() -> (*kernel_fn)(dev_ctx, inputs..., attrs..., outputs...)
//     with phi::DenseTensor arguments, outputs that arrive with meta set and NO allocation, and a phi::GPUContext whose
//     stream() is a real cudaStream_t, whose GetPlace() is a GPUPlace and whose Alloc<T>() is served by a stream-ordered
//     allocator (cudaMallocAsync on that stream) through the reference's own DeviceContext::Alloc.
// phi::GPUContext's member functions are defined HERE (the reference's gpu_context.cc drags in the whole handle /
// dynload / allocator-facade closure): the class layout is the header's, the behaviour is the three things a kernel
// body needs.  The few other heavy dependencies of the registry TUs that are irrelevant here are stubbed and abort if
// ever reached.
#include <cuda_runtime.h>

#include <cstdlib>
#include <cstring>
#include <string>

#include "paddle/phi/backends/context_pool.h"
#include "paddle/phi/backends/gpu/cuda/cuda_graph.h"
#include "paddle/phi/backends/gpu/gpu_context.h"
#include "paddle/phi/common/int_array.h"
#include "paddle/phi/common/scalar.h"
#include "paddle/phi/core/dense_tensor.h"
#include "paddle/phi/core/kernel_factory.h"
#include "paddle/phi/core/tensor_utils.h"
#include "paddle/utils/optional.h"

namespace phi {
DeviceContextPool& DeviceContextPool::Instance() { std::abort(); }
phi::DeviceContext* DeviceContextPool::Get(const phi::Place&) { std::abort(); }
template <typename Context>
void Copy(const Context&, const DenseTensor&, Place, bool, DenseTensor*) { std::abort(); }
template void Copy<DeviceContext>(const DeviceContext&, const DenseTensor&, Place, bool, DenseTensor*);
// phi::GPUContext: the state a kernel body reads.  Defining the key function here also emits the class's vtable and
// typeinfo, which the registrar needs (it stores typeid(GPUContext) for the context argument).
struct GPUContext::Impl {
  Place place;
  cudaStream_t stream{nullptr};
};
GPUContext::GPUContext(const GPUPlace& place, bool, int) : DeviceContext(), impl_(std::make_unique<Impl>()) {
  impl_->place = place;
  cudaSetDevice(place.GetDeviceId());
  cudaStreamCreateWithFlags(&impl_->stream, cudaStreamNonBlocking);
}
GPUContext::~GPUContext() {
  if (impl_ && impl_->stream) cudaStreamDestroy(impl_->stream);
}
const Place& GPUContext::GetPlace() const { return impl_->place; }
void GPUContext::Wait() const { cudaStreamSynchronize(impl_->stream); }
gpuStream_t GPUContext::stream() const { return impl_->stream; }
namespace backends { namespace gpu {
int GetCurrentDeviceId() {
  int d = 0;
  cudaGetDevice(&d);
  return d;
}
// statics of cuda_graph.cc (not part of this build): "not capturing"
std::unique_ptr<CUDAGraph> CUDAGraph::capturing_graph_{nullptr};
paddle::optional<std::thread::id> CUDAGraph::capturing_thread_id_{paddle::none};
void CUDAGraph::Reset() {}
}}  // namespace backends::gpu
}  // namespace phi

namespace {
phi::KernelKey MakeKey(const char* backend, int dtype) {
  return phi::KernelKey(paddle::experimental::StringToBackend(backend), phi::DataLayout::ALL_LAYOUT,
                        static_cast<phi::DataType>(dtype));
}
void DummyKernel(phi::KernelContext*) {}
}  // namespace

extern "C" {
// 1 if (name, backend, ALL_LAYOUT, dtype) is registered
int phi_min_has_kernel(const char* name, const char* backend, int dtype) {
  auto& m = phi::KernelFactory::Instance().kernels();
  auto it = m.find(name);
  if (it == m.end()) return 0;
  return it->second.count(MakeKey(backend, dtype)) ? 1 : 0;
}
// the variadic function pointer stored under the key (what the generated eager API calls), or NULL
void* phi_min_variadic_fn(const char* name, const char* backend, int dtype) {
  auto& m = phi::KernelFactory::Instance().kernels();
  auto it = m.find(name);
  if (it == m.end()) return nullptr;
  auto kt = it->second.find(MakeKey(backend, dtype));
  if (kt == it->second.end()) return nullptr;
  return kt->second.GetVariadicKernelFn<void*>();
}
// number of inputs / attributes / outputs the registrar parsed out of the C++ signature
int phi_min_arity(const char* name, const char* backend, int dtype, int* n_in, int* n_attr, int* n_out) {
  auto& m = phi::KernelFactory::Instance().kernels();
  auto it = m.find(name);
  if (it == m.end()) return 0;
  auto kt = it->second.find(MakeKey(backend, dtype));
  if (kt == it->second.end()) return 0;
  const auto& def = kt->second.args_def();
  *n_in = static_cast<int>(def.input_defs().size());
  *n_attr = static_cast<int>(def.attribute_defs().size());
  *n_out = static_cast<int>(def.output_defs().size());
  return 1;
}
// dtype recorded for output `i` (UNDEFINED = 0 means "decided at run time")
int phi_min_output_dtype(const char* name, const char* backend, int dtype, int i) {
  auto& m = phi::KernelFactory::Instance().kernels();
  const auto& def = m.at(name).at(MakeKey(backend, dtype)).args_def();
  return static_cast<int>(def.output_defs().at(i).dtype);
}
// stand-in for the in-tree kernel: put a dummy under the key so that the plugin has something to replace
void phi_min_register_dummy(const char* name, const char* backend, int dtype) {
  phi::KernelFactory::Instance().kernels()[name][MakeKey(backend, dtype)] = phi::Kernel(DummyKernel, nullptr);
}
int phi_min_kernel_count() {
  int n = 0;
  for (auto& kv : phi::KernelFactory::Instance().kernels()) n += static_cast<int>(kv.second.size());
  return n;
}
}

// ---------------------------------------------------------------------------------------------------------------
// GPU dispatch: call a registered kernel the way Paddle's generated API does.
// ---------------------------------------------------------------------------------------------------------------
namespace {

thread_local std::string g_pm_error;

// stream-ordered device allocator behind DeviceContext::Alloc (stands in for Paddle's stream-safe allocator)
cudaStream_t g_alloc_stream = nullptr;
int g_alloc_device = 0;
size_t g_alloc_calls = 0;
class AsyncGpuAllocator : public phi::Allocator {
 public:
  static void Delete(phi::Allocation* a) {
    if (a->ptr()) cudaFreeAsync(a->ptr(), g_alloc_stream);
    delete a;
  }
  AllocationPtr Allocate(size_t bytes) override {
    void* p = nullptr;
    ++g_alloc_calls;
    if (bytes > 0 && cudaMallocAsync(&p, bytes, g_alloc_stream) != cudaSuccess) {
      cudaGetLastError();
      throw std::runtime_error("phi_min: cudaMallocAsync failed");
    }
    return AllocationPtr(new phi::Allocation(p, bytes, phi::GPUPlace(g_alloc_device)), Delete);
  }
  bool IsAllocThreadSafe() const override { return true; }
};
class HostAllocator : public phi::Allocator {
 public:
  static void Delete(phi::Allocation* a) {
    std::free(a->ptr());
    delete a;
  }
  AllocationPtr Allocate(size_t bytes) override {
    void* p = std::malloc(bytes ? bytes : 64);
    return AllocationPtr(new phi::Allocation(p, bytes, phi::CPUPlace()), Delete);
  }
  bool IsAllocThreadSafe() const override { return true; }
};

struct GpuEnv {
  AsyncGpuAllocator dev;
  HostAllocator host;
  std::unique_ptr<phi::GPUContext> ctx;
};
GpuEnv* g_env = nullptr;

struct PmTensor {
  void* data;        // device pointer (inputs: borrowed; outputs: where the result is copied to) or NULL = absent
  int64_t dims[9];
  int rank;          // outputs: < 0 = "not wanted" (the kernel receives a NULL DenseTensor*)
  int dtype;         // phi::DataType as int
};

phi::DenseTensorMeta MetaOf(const PmTensor& t) {
  std::vector<int64_t> d(t.dims, t.dims + (t.rank > 0 ? t.rank : 0));
  return phi::DenseTensorMeta(static_cast<phi::DataType>(t.dtype), common::make_ddim(d));
}
// borrow caller device memory as a DenseTensor (no ownership)
phi::DenseTensor WrapIn(const PmTensor& t) {
  phi::DenseTensorMeta meta = MetaOf(t);
  const size_t bytes = static_cast<size_t>(common::product(meta.dims)) * phi::SizeOf(meta.dtype);
  auto holder = std::make_shared<phi::Allocation>(t.data, bytes, phi::GPUPlace(g_alloc_device));
  return phi::DenseTensor(holder, meta);
}
paddle::optional<phi::DenseTensor> WrapOpt(const PmTensor& t) {
  if (t.data == nullptr) return paddle::none;
  return WrapIn(t);
}
// what InferMeta leaves behind: dims + dtype, no allocation
phi::DenseTensor MetaOnly(const PmTensor& t) {
  phi::DenseTensor d;
  const phi::DenseTensorMeta meta = MetaOf(t);
  d.set_meta(meta);  // the const& overload: what InferMeta's MetaTensor::set_dims / set_dtype amount to
  return d;
}
// after the call: the kernel must have allocated the output on the GPU with the promised dtype and size; copy it out
void Export(const phi::DenseTensor& t, const PmTensor& want, cudaStream_t st) {
  if (want.rank < 0 || want.data == nullptr) return;
  if (!t.initialized()) throw std::runtime_error("phi_min: the kernel left an output unallocated");
  if (t.place().GetType() != phi::AllocationType::GPU) throw std::runtime_error("phi_min: output is not on the GPU");
  if (static_cast<int>(t.dtype()) != want.dtype)
    throw std::runtime_error("phi_min: output dtype " + std::to_string(static_cast<int>(t.dtype())) + " != expected " +
                             std::to_string(want.dtype));
  const size_t bytes = static_cast<size_t>(t.numel()) * phi::SizeOf(t.dtype());
  if (bytes && cudaMemcpyAsync(want.data, t.data(), bytes, cudaMemcpyDeviceToDevice, st) != cudaSuccess)
    throw std::runtime_error("phi_min: copy-out failed");
}

using DT = phi::DenseTensor;
using Ctx = phi::DeviceContext;  // the generated API's signatures name the base class (api_base.py:1349)
using OptT = paddle::optional<phi::DenseTensor>;

}  // namespace

extern "C" {

const char* phi_min_last_error() { return g_pm_error.c_str(); }
size_t phi_min_alloc_calls() { return g_alloc_calls; }

int phi_min_gpu_init(int device) {
  if (g_env) return 0;
  try {
    if (cudaSetDevice(device) != cudaSuccess) throw std::runtime_error("cudaSetDevice failed");
    auto* env = new GpuEnv();
    env->ctx = std::make_unique<phi::GPUContext>(phi::GPUPlace(device));
    g_alloc_stream = env->ctx->stream();
    g_alloc_device = device;
    env->ctx->SetAllocator(&env->dev);
    env->ctx->SetZeroAllocator(&env->dev);
    env->ctx->SetHostAllocator(&env->host);
    env->ctx->SetHostZeroAllocator(&env->host);
    g_env = env;
    return 0;
  } catch (const std::exception& e) {
    g_pm_error = e.what();
    return 1;
  }
}
void* phi_min_stream() { return g_env ? g_env->ctx->stream() : nullptr; }

// family: which C++ signature the kernel has (one per shape of argument list in plugin/phi_kernels.cc).
//   f = float/double attributes, iv = integer attributes (bools, ints, DataTypes, axes), in signature order.
int phi_min_run(const char* name, const char* backend, int key_dtype, const char* family, const PmTensor* in, int n_in,
                PmTensor* out, int n_out, const double* f, int nf, const int64_t* iv, int ni) {
  if (!g_env) {
    g_pm_error = "phi_min_gpu_init was not called";
    return 1;
  }
  try {
    const phi::GPUContext& gctx = *g_env->ctx;
    const Ctx& ctx = gctx;
    cudaStream_t st = gctx.stream();
    // 1. select, as the generated API does
    auto result = phi::KernelFactory::Instance().SelectKernelOrThrowError(
        name, {paddle::experimental::StringToBackend(backend), phi::DataLayout::ALL_LAYOUT, static_cast<phi::DataType>(key_dtype)}, true);
    const phi::Kernel& kernel = result.kernel;
    if (result.has_fallback_cpu) throw std::runtime_error("phi_min: selection fell back to CPU");
    const std::string fam = family;
    std::vector<DT> o;
    for (int i = 0; i < n_out; ++i) o.push_back(out[i].rank >= 0 ? MetaOnly(out[i]) : DT());
    auto op = [&](int i) -> DT* { return (i < n_out && out[i].rank >= 0) ? &o[i] : nullptr; };
    // 2. fetch the variadic function under the in-tree signature and 3. call it
    if (fam == "binary") {  // (ctx, x, y, out)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const DT&, DT*)>();
      (*fn)(ctx, WrapIn(in[0]), WrapIn(in[1]), op(0));
    } else if (fam == "binary_bool") {  // (ctx, x, y, bool, out): shifts
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const DT&, bool, DT*)>();
      (*fn)(ctx, WrapIn(in[0]), WrapIn(in[1]), iv[0] != 0, op(0));
    } else if (fam == "unary") {  // (ctx, x, out)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, DT*)>();
      (*fn)(ctx, WrapIn(in[0]), op(0));
    } else if (fam == "scale") {  // (ctx, x, Scalar scale, Scalar bias, bool bias_after_scale, out)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const phi::Scalar&, const phi::Scalar&, bool, DT*)>();
      (*fn)(ctx, WrapIn(in[0]), phi::Scalar(f[0]), phi::Scalar(f[1]), iv[0] != 0, op(0));
    } else if (fam == "cast") {  // (ctx, x, DataType, out)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, phi::DataType, DT*)>();
      (*fn)(ctx, WrapIn(in[0]), static_cast<phi::DataType>(iv[0]), op(0));
    } else if (fam == "ternary") {  // where(ctx, cond, x, y, out)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const DT&, const DT&, DT*)>();
      (*fn)(ctx, WrapIn(in[0]), WrapIn(in[1]), WrapIn(in[2]), op(0));
    } else if (fam == "binary_grad") {  // (ctx, x, y, dout, int axis, dx, dy)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const DT&, const DT&, int, DT*, DT*)>();
      (*fn)(ctx, WrapIn(in[0]), WrapIn(in[1]), WrapIn(in[2]), static_cast<int>(iv[0]), op(0), op(1));
    } else if (fam == "reduce_grad") {  // (ctx, x, out_grad, IntArray dims, bool keep_dim, bool reduce_all, x_grad)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const DT&, const phi::IntArray&, bool, bool, DT*)>();
      std::vector<int64_t> axes(iv + 2, iv + ni);
      (*fn)(ctx, WrapIn(in[0]), WrapIn(in[1]), phi::IntArray(axes), iv[0] != 0, iv[1] != 0, op(0));
    } else if (fam == "select_grad") {  // max/min/prod grad: (ctx, x, out, out_grad, IntArray dims, bool keep_dim, bool reduce_all, x_grad)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const DT&, const DT&, const phi::IntArray&, bool, bool, DT*)>();
      std::vector<int64_t> axes(iv + 2, iv + ni);
      (*fn)(ctx, WrapIn(in[0]), WrapIn(in[1]), WrapIn(in[2]), phi::IntArray(axes), iv[0] != 0, iv[1] != 0, op(0));
    } else if (fam == "sum_raw") {  // (ctx, x, IntArray, bool keep_dim, bool reduce_all, DataType out_dtype, out)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const phi::IntArray&, bool, bool, phi::DataType, DT*)>();
      std::vector<int64_t> axes(iv + 3, iv + ni);
      (*fn)(ctx, WrapIn(in[0]), phi::IntArray(axes), iv[0] != 0, iv[1] != 0, static_cast<phi::DataType>(iv[2]), op(0));
    } else if (fam == "sum") {  // (ctx, x, IntArray, DataType out_dtype, bool keep_dim, out)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const phi::IntArray&, phi::DataType, bool, DT*)>();
      std::vector<int64_t> axes(iv + 2, iv + ni);
      (*fn)(ctx, WrapIn(in[0]), phi::IntArray(axes), static_cast<phi::DataType>(iv[0]), iv[1] != 0, op(0));
    } else if (fam == "reduce_raw") {  // mean/max/min raw, prod: (ctx, x, IntArray, bool keep_dim, bool reduce_all, out)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const phi::IntArray&, bool, bool, DT*)>();
      std::vector<int64_t> axes(iv + 2, iv + ni);
      (*fn)(ctx, WrapIn(in[0]), phi::IntArray(axes), iv[0] != 0, iv[1] != 0, op(0));
    } else if (fam == "reduce") {  // mean/max/min: (ctx, x, IntArray, bool keep_dim, out)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const phi::IntArray&, bool, DT*)>();
      std::vector<int64_t> axes(iv + 1, iv + ni);
      (*fn)(ctx, WrapIn(in[0]), phi::IntArray(axes), iv[0] != 0, op(0));
    } else if (fam == "argmax") {  // (ctx, x, Scalar axis, bool keepdims, bool flatten, DataType, out)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const phi::Scalar&, bool, bool, phi::DataType, DT*)>();
      (*fn)(ctx, WrapIn(in[0]), phi::Scalar(static_cast<int64_t>(iv[0])), iv[1] != 0, iv[2] != 0, static_cast<phi::DataType>(iv[3]), op(0));
    } else if (fam == "axis") {  // softmax / log_softmax: (ctx, x, int axis, out)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, int, DT*)>();
      (*fn)(ctx, WrapIn(in[0]), static_cast<int>(iv[0]), op(0));
    } else if (fam == "axis_grad") {  // softmax_grad: (ctx, out, out_grad, int axis, x_grad)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const DT&, int, DT*)>();
      (*fn)(ctx, WrapIn(in[0]), WrapIn(in[1]), static_cast<int>(iv[0]), op(0));
    } else if (fam == "layer_norm") {  // (ctx, x, optional scale, optional bias, float eps, int begin_norm_axis, y, mean, var)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const OptT&, const OptT&, float, int, DT*, DT*, DT*)>();
      (*fn)(ctx, WrapIn(in[0]), WrapOpt(in[1]), WrapOpt(in[2]), static_cast<float>(f[0]), static_cast<int>(iv[0]), op(0), op(1), op(2));
    } else if (fam == "layer_norm_grad") {  // (ctx, x, scale?, bias?, mean, var, dy, eps, axis, dx, dscale, dbias)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const OptT&, const OptT&, const DT&, const DT&, const DT&,
                                                     float, int, DT*, DT*, DT*)>();
      (*fn)(ctx, WrapIn(in[0]), WrapOpt(in[1]), WrapOpt(in[2]), WrapIn(in[3]), WrapIn(in[4]), WrapIn(in[5]), static_cast<float>(f[0]),
            static_cast<int>(iv[0]), op(0), op(1), op(2));
    } else if (fam == "cross_entropy") {  // (ctx, logits, label, bool soft, bool use_softmax, bool stable, int ignore, int axis, softmax, loss)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const DT&, bool, bool, bool, int, int, DT*, DT*)>();
      (*fn)(ctx, WrapIn(in[0]), WrapIn(in[1]), iv[0] != 0, iv[1] != 0, iv[2] != 0, static_cast<int>(iv[3]), static_cast<int>(iv[4]), op(0),
            op(1));
    } else if (fam == "rms_norm") {  // (ctx, x, bias?, residual?, norm_weight, norm_bias?, eps, axis, quant x4, out, residual_out, inv_var)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const OptT&, const OptT&, const DT&, const OptT&, const float,
                                                     const int, const float, const int, const float, const float, DT*, DT*, DT*)>();
      (*fn)(ctx, WrapIn(in[0]), WrapOpt(in[1]), WrapOpt(in[2]), WrapIn(in[3]), WrapOpt(in[4]), static_cast<float>(f[0]),
            static_cast<int>(iv[0]), -1.0f, 0, 0.0f, 0.0f, op(0), op(1), op(2));
    } else if (fam == "softmax_mask") {  // fused_softmax_mask(ctx, x, mask, out)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const DT&, DT*)>();
      (*fn)(ctx, WrapIn(in[0]), WrapIn(in[1]), op(0));
    } else if (fam == "fused_ln") {  // (ctx, x, residual, bias?, ln_scale?, ln_bias?, rate, is_test, fix_seed, seed, impl, eps, y, dro, mask, mean, var)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const DT&, const OptT&, const OptT&, const OptT&, const float,
                                                     const bool, const bool, const int, const std::string&, const float, DT*, DT*, DT*,
                                                     DT*, DT*)>();
      const std::string impl = iv[1] != 0 ? "upscale_in_train" : "downscale_in_infer";
      (*fn)(ctx, WrapIn(in[0]), WrapIn(in[1]), WrapOpt(in[2]), WrapOpt(in[3]), WrapOpt(in[4]), static_cast<float>(f[0]), iv[0] != 0, false,
            0, impl, static_cast<float>(f[1]), op(0), op(1), op(2), op(3), op(4));
    } else if (fam == "cross_entropy_grad") {  // (ctx, label, softmax, loss_grad, bool soft, bool use_softmax, bool stable, int ignore, int axis, logits_grad)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const DT&, const DT&, bool, bool, bool, int, int, DT*)>();
      (*fn)(ctx, WrapIn(in[0]), WrapIn(in[1]), WrapIn(in[2]), iv[0] != 0, iv[1] != 0, true, static_cast<int>(iv[2]), static_cast<int>(iv[3]), op(0));
    } else if (fam == "fused_layernorm") {  // (ctx, x, bias?, residual?, norm_weight?, norm_bias?, eps, alpha, bna, qscale, qround, qmax, qmin, out, residual_out, mean, var)
      auto* fn = kernel.GetVariadicKernelFn<void (*)(const Ctx&, const DT&, const OptT&, const OptT&, const OptT&, const OptT&, const float,
                                                     const float, const int, const float, const int, const float, const float, DT*, DT*,
                                                     DT*, DT*)>();
      (*fn)(ctx, WrapIn(in[0]), WrapOpt(in[1]), WrapOpt(in[2]), WrapOpt(in[3]), WrapOpt(in[4]), static_cast<float>(f[0]),
            static_cast<float>(f[1]), static_cast<int>(iv[0]), -1.0f, 0, 0.0f, 0.0f, op(0), op(1), op(2), op(3));
    } else {
      throw std::runtime_error("phi_min: unknown signature family '" + fam + "'");
    }
    for (int i = 0; i < n_out; ++i)
      if (out[i].rank >= 0) Export(o[i], out[i], st);
    o.clear();  // frees are stream-ordered after the copies
    if (cudaStreamSynchronize(st) != cudaSuccess) throw std::runtime_error(std::string("phi_min: ") + cudaGetErrorString(cudaGetLastError()));
    return 0;
  } catch (const std::exception& e) {
    g_pm_error = e.what();
    return 1;
  } catch (...) {
    g_pm_error = "unknown exception";
    return 1;
  }
}

}  // extern "C"
