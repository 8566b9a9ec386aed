// TEST INFRASTRUCTURE — stand-in for gflags (not in this image): a flag's value comes from the environment variable
// VFLAG_<name>, else its default.  Only for compiling the reference's stand-alone GPU benchmark where it lies.
#pragma once
#include <cstdint>
#include <cstdlib>
#include <cstring>
namespace vflag {
inline long long geti(const char* n, long long d) { const char* e = getenv(n); return e ? atoll(e) : d; }
inline double getd(const char* n, double d) { const char* e = getenv(n); return e ? atof(e) : d; }
inline bool getb(const char* n, bool d) { const char* e = getenv(n); return e ? (e[0] == '1' || e[0] == 't' || e[0] == 'T') : d; }
} // namespace vflag
#define DEFINE_int32(name, dflt, doc) static int32_t FLAGS_##name = (int32_t)vflag::geti("VFLAG_" #name, dflt)
#define DEFINE_int64(name, dflt, doc) static int64_t FLAGS_##name = (int64_t)vflag::geti("VFLAG_" #name, dflt)
#define DEFINE_bool(name, dflt, doc) static bool FLAGS_##name = vflag::getb("VFLAG_" #name, dflt)
#define DEFINE_double(name, dflt, doc) static double FLAGS_##name = vflag::getd("VFLAG_" #name, dflt)
