// TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Minimal stand-in for <glog/logging.h> so that the reference's CPU PHI sources
// can be compiled in place for the parity oracle without the glog dependency.
// Every logging statement becomes a no-op stream; CHECKs keep their condition.
#pragma once
#include <cstdlib>
#include <iostream>
#include <sstream>

namespace pdb200_glog_stub {
struct NullStream {
  template <typename T>
  NullStream& operator<<(const T&) { return *this; }
  NullStream& operator<<(std::ostream& (*)(std::ostream&)) { return *this; }
};
struct Voidify {
  void operator&(const NullStream&) {}
};
struct FatalStream {
  std::ostringstream os;
  template <typename T>
  FatalStream& operator<<(const T& v) { os << v; return *this; }
  ~FatalStream() { std::cerr << "CHECK failed: " << os.str() << std::endl; std::abort(); }
};
struct VoidifyFatal {
  void operator&(const FatalStream&) {}
};
}  // namespace pdb200_glog_stub

#define PDB200_NULL_LOG true ? (void)0 : ::pdb200_glog_stub::Voidify() & ::pdb200_glog_stub::NullStream()
#define VLOG(n) PDB200_NULL_LOG
#define VLOG_IS_ON(n) false
#define LOG(sev) PDB200_NULL_LOG
#define LOG_IF(sev, c) PDB200_NULL_LOG
#define LOG_FIRST_N(sev, n) PDB200_NULL_LOG
#define LOG_EVERY_N(sev, n) PDB200_NULL_LOG
#define DLOG(sev) PDB200_NULL_LOG
#define PDB200_CHECK(c) (c) ? (void)0 : ::pdb200_glog_stub::VoidifyFatal() & ::pdb200_glog_stub::FatalStream() << #c << " "
#define CHECK(c) PDB200_CHECK(c)
#define CHECK_EQ(a, b) PDB200_CHECK((a) == (b))
#define CHECK_NE(a, b) PDB200_CHECK((a) != (b))
#define CHECK_LT(a, b) PDB200_CHECK((a) < (b))
#define CHECK_LE(a, b) PDB200_CHECK((a) <= (b))
#define CHECK_GT(a, b) PDB200_CHECK((a) > (b))
#define CHECK_GE(a, b) PDB200_CHECK((a) >= (b))
#define CHECK_NOTNULL(p) (p)
#define DCHECK(c) PDB200_CHECK(c)
#define DCHECK_EQ(a, b) CHECK_EQ(a, b)
#define DCHECK_NE(a, b) CHECK_NE(a, b)
#define DCHECK_LT(a, b) CHECK_LT(a, b)
#define DCHECK_LE(a, b) CHECK_LE(a, b)
#define DCHECK_GT(a, b) CHECK_GT(a, b)
#define DCHECK_GE(a, b) CHECK_GE(a, b)
#define PCHECK(c) PDB200_CHECK(c)

namespace google {
inline void InitGoogleLogging(const char*) {}
inline void InstallFailureSignalHandler() {}
inline void InstallFailureWriter(void (*)(const char*, size_t)) {}
}  // namespace google
