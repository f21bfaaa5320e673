// Error plumbing shared by the C-ABI translation units.
#pragma once
#include <cstddef>
#include <string>
#include <new>
#include <exception>

namespace j2k {
std::string& last_error();                       // thread-local
int set_error(int code, const std::string& msg); // stores msg, returns code
// every buffer the library hands out for j2kcuda_free(); large ones are recycled (common.cpp)
void* out_alloc(size_t n);
void out_free(void* p);
void out_release_all();                          // j2kcuda_shutdown
// j2kcuda_init's log callback: errors (level 0) and notices (level 1) of every entry point go to it
void set_log(void (*fn)(void*, int, const char*), void* user);
void log_msg(int level, const std::string& msg);
// the one-shot calls (j2kcuda_encode / j2kcuda_decode) keep a few contexts between calls, keyed by their parameters:
// a context for a 4K RGBA16 frame is about 1.5 GB of device and page-locked memory, far too much to allocate per frame
void encoder_pool_clear();
void decoder_pool_clear();
bool context_cache_enabled();
// No C++ exception may cross the C ABI (a std::bad_alloc from a crafted header used to reach std::terminate):
// every extern "C" entry point runs its body through this.
template <class F> inline int guarded(F&& f) noexcept {
    try { return f(); }
    catch (const std::bad_alloc&) { return set_error(5 /* J2KCUDA_ERR_NOMEM */, "out of host memory"); }
    catch (const std::exception& e) { return set_error(4 /* J2KCUDA_ERR_STREAM */, std::string("internal error: ") + e.what()); }
    catch (...) { return set_error(4, "internal error"); }
}
}  // namespace j2k

// Internal (not exported through include/j2kcuda.h): shape of the MQ kernel.  A lone frame wants few blocks per warp
// (the kernel ends when its longest chain ends); with many frames in flight the device is bound by issue slots and
// fuller warps do the same work with fewer of them.  api_batch.cpp tells its contexts which regime they are in.
struct j2kcuda_encoder;
namespace j2k { void encoder_set_throughput_mode(::j2kcuda_encoder* e, bool many_frames_in_flight); }
