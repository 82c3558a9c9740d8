// runtime.cu — error plumbing, version and launch accounting for libpdb200.so.
#include <map>
#include <utility>

#include "common.cuh"

namespace pdb200 {

std::mutex& func_attr_mutex() {
  static std::mutex mu;
  return mu;
}

bool func_attr_apply(const void* kern, int smem_bytes, int carveout_pct, FuncAttrHold* hold) {
  hold->lock = std::unique_lock<std::mutex>(func_attr_mutex());
  return func_attr_apply_locked(kern, smem_bytes, carveout_pct);
}

bool func_attr_apply_locked(const void* kern, int smem_bytes, int carveout_pct) {
  // what each (kernel, device) pair was last set to: an unchanged pair costs one map lookup, no driver call
  static std::map<std::pair<const void*, int>, std::pair<int, int>> applied;
  int dev = 0;
  cudaGetDevice(&dev);
  auto& cur = applied.emplace(std::make_pair(kern, dev), std::make_pair(-1, -2)).first->second;
  if (cur.first != smem_bytes) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes) != cudaSuccess) {
      cudaGetLastError();
      cur.first = -1;
      return false;
    }
    cur.first = smem_bytes;
  }
  if (carveout_pct >= 0 && cur.second != carveout_pct) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, carveout_pct) != cudaSuccess) {
      cudaGetLastError();
      cur.second = -2;
      return false;
    }
    cur.second = carveout_pct;
  }
  return true;
}

std::atomic<uint64_t> g_launch_count{0};
static thread_local std::string t_last_error;

void set_error(const std::string& msg) { t_last_error = msg; }
int fail(int code, const std::string& msg) {
  t_last_error = msg;
  return code;
}

}  // namespace pdb200

extern "C" {
const char* pdb200_last_error(void) { return pdb200::t_last_error.c_str(); }
const char* pdb200_version(void) { return "pdb200 0.1 (sm_100a)"; }
uint64_t pdb200_launch_count(void) { return pdb200::g_launch_count.load(); }
}
