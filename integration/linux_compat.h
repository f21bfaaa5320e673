// Force-included (-include) when the reference's UNCHANGED src/common sources are compiled on Linux: the only non-portable
// spot on the path is Codec::NumberOfCPUs (src/common/j2k_codec.cpp:430-456), which knows Mach and Win32 only.  This is the
// POSIX branch a maintainer would add there.
#pragma once
#if defined(__linux__) && defined(__cplusplus)
#include <unistd.h>
#include <cstdint>
#include <cstring>
#include <cstdlib>
#include <cmath>
struct SYSTEM_INFO { unsigned long dwNumberOfProcessors; };
static inline void GetSystemInfo(SYSTEM_INFO* s) { long n = sysconf(_SC_NPROCESSORS_ONLN); s->dwNumberOfProcessors = n > 0 ? (unsigned long)n : 1; }
#endif
