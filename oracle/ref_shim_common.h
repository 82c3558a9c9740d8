// TEST INFRASTRUCTURE ONLY — plumbing shared by the translation units of oracle/_ref/libphi_cpu_ref.so
// (ref_shim.cc, ref_shim_family.cc): a malloc-backed CPUContext, borrowing caller memory as a DenseTensor,
// and turning exceptions into a status + message.  No algorithm of any kind lives here.
#pragma once

#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "paddle/phi/backends/cpu/cpu_context.h"
#include "paddle/phi/core/dense_tensor.h"

inline thread_local std::string g_last_error;  // one per thread, shared by both shim TUs

namespace {

class MallocAllocator : public phi::Allocator {
 public:
  static void Delete(phi::Allocation* a) {
    std::free(a->ptr());
    delete a;
  }
  AllocationPtr Allocate(size_t bytes) override {
    void* p = nullptr;
    if (posix_memalign(&p, 64, bytes ? bytes : 64) != 0) p = nullptr;
    auto* a = new phi::Allocation(p, bytes, phi::CPUPlace());
    return AllocationPtr(a, Delete);
  }
  bool IsAllocThreadSafe() const override { return true; }
};

struct Ctx {
  MallocAllocator alloc;
  phi::CPUContext ctx;
  Ctx() {
    ctx.SetAllocator(&alloc);
    ctx.SetHostAllocator(&alloc);
    ctx.SetZeroAllocator(&alloc);
  }
};

Ctx& GetCtx() {
  static Ctx c;
  return c;
}

// dtype codes shared with include/pdb200.h
enum DType {
  kBool = 0, kUInt8 = 1, kInt8 = 2, kInt16 = 3, kInt32 = 4, kInt64 = 5,
  kFloat16 = 6, kBFloat16 = 7, kFloat32 = 8, kFloat64 = 9
};

phi::DataType ToPhi(int dt) {
  switch (dt) {
    case kBool: return phi::DataType::BOOL;
    case kUInt8: return phi::DataType::UINT8;
    case kInt8: return phi::DataType::INT8;
    case kInt16: return phi::DataType::INT16;
    case kInt32: return phi::DataType::INT32;
    case kInt64: return phi::DataType::INT64;
    case kFloat16: return phi::DataType::FLOAT16;
    case kBFloat16: return phi::DataType::BFLOAT16;
    case kFloat32: return phi::DataType::FLOAT32;
    case kFloat64: return phi::DataType::FLOAT64;
  }
  return phi::DataType::UNDEFINED;
}

// Borrow caller memory as a DenseTensor (no ownership).
phi::DenseTensor Wrap(const void* data, const int64_t* dims, int rank, int dt) {
  std::vector<int64_t> d(dims, dims + rank);
  phi::DDim ddim = common::make_ddim(d);
  phi::DenseTensorMeta meta(ToPhi(dt), ddim);
  size_t bytes = static_cast<size_t>(common::product(ddim)) * phi::SizeOf(ToPhi(dt));
  auto holder = std::make_shared<phi::Allocation>(const_cast<void*>(data), bytes, phi::CPUPlace());
  return phi::DenseTensor(holder, meta);
}

template <typename F>
int Guard(F&& f) {
  try {
    f();
    return 0;
  } catch (const std::exception& e) {
    g_last_error = e.what();
    return 1;
  } catch (...) {
    g_last_error = "unknown exception";
    return 1;
  }
}

}  // namespace
