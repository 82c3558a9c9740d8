// TEST INFRASTRUCTURE ONLY.  A minimal host for Paddle's kernel REGISTRY, built from the reference's
// own sources (oracle/build_registry.sh): KernelFactory, KernelRegistrar and what they link against.
// It exists so that tests/test_phi_registration.py can dlopen the Paddle-side binding
// (paddlepaddle_b200/plugin/phi_kernels.cc, compiled against the reference headers) and check that
// its PD_REGISTER_KERNEL static initialisers really do put the B200 kernels under Paddle's own
// (name, backend, layout, dtype) keys — replacing an entry that was there before.
// No kernel is invoked through it: a GPUContext needs the rest of Paddle.  The few heavy
// dependencies of the registry TUs that are irrelevant to registration are stubbed below and abort
// if ever reached.
#include <cstdlib>
#include <cstring>
#include <string>

#include "paddle/phi/backends/context_pool.h"
#include "paddle/phi/backends/gpu/cuda/cuda_graph.h"
#include "paddle/phi/backends/gpu/gpu_context.h"
#include "paddle/phi/core/kernel_factory.h"
#include "paddle/phi/core/tensor_utils.h"

namespace phi {
DeviceContextPool& DeviceContextPool::Instance() { std::abort(); }
phi::DeviceContext* DeviceContextPool::Get(const phi::Place&) { std::abort(); }
template <typename Context>
void Copy(const Context&, const DenseTensor&, Place, bool, DenseTensor*) { std::abort(); }
template void Copy<DeviceContext>(const DeviceContext&, const DenseTensor&, Place, bool, DenseTensor*);
// phi::GPUContext's key function: defining it here emits the class's vtable and typeinfo, which the
// registrar needs (it stores typeid(GPUContext) for the context argument).  Never instantiated.
struct GPUContext::Impl {};
GPUContext::~GPUContext() {}
const Place& GPUContext::GetPlace() const { std::abort(); }
void GPUContext::Wait() const { std::abort(); }
gpuStream_t GPUContext::stream() const { std::abort(); }
namespace backends { namespace gpu {
int GetCurrentDeviceId() { return 0; }
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
