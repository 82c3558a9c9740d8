// runtime.cu — error plumbing, version and launch accounting for libpdb200.so.
#include "common.cuh"

namespace pdb200 {

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
