// Error plumbing shared by the C-ABI translation units.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <cstring>
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
// Growable byte array living in out_alloc memory: the Tier-2 writer builds the codestream in it and the buffer itself is
// handed to the caller (release), so a 12-37 MB codestream is never copied a second time.  The subset of the
// std::vector<uint8_t> interface the writer uses.
class ByteBuf {
    uint8_t* p_ = nullptr; size_t n_ = 0, cap_ = 0;
    void grow(size_t need) {
        size_t c = cap_ ? cap_ * 2 : 4096;
        if (c < need) c = need;
        uint8_t* q = (uint8_t*)out_alloc(c);
        if (!q) throw std::bad_alloc();
        if (n_) memcpy(q, p_, n_);
        out_free(p_);
        p_ = q; cap_ = c;
    }
public:
    ByteBuf() {}
    ByteBuf(const ByteBuf&) = delete;
    ByteBuf& operator=(const ByteBuf&) = delete;
    ~ByteBuf() { out_free(p_); }
    void reserve(size_t c) { if (c > cap_) grow(c); }
    void clear() { n_ = 0; }
    size_t size() const { return n_; }
    uint8_t* data() { return p_; }
    uint8_t* begin() { return p_; }
    uint8_t* end() { return p_ + n_; }
    uint8_t& operator[](size_t i) { return p_[i]; }
    void push_back(uint8_t v) { if (n_ == cap_) grow(n_ + 1); p_[n_++] = v; }
    template <class It> void insert(uint8_t* /*at end()*/, It first, It last) {
        const size_t k = (size_t)(last - first);
        if (n_ + k > cap_) grow(n_ + k);
        for (size_t i = 0; i < k; ++i) p_[n_ + i] = (uint8_t)first[i];
        n_ += k;
    }
    void insert(uint8_t*, const uint8_t* first, const uint8_t* last) {
        const size_t k = (size_t)(last - first);
        if (n_ + k > cap_) grow(n_ + k);
        if (k) memcpy(p_ + n_, first, k);
        n_ += k;
    }
    void erase(uint8_t* first, uint8_t* last) { const size_t k = (size_t)(last - first); memmove(first, last, (size_t)(end() - last)); n_ -= k; }
    uint8_t* release() { if (!p_) grow(1); uint8_t* r = p_; p_ = nullptr; n_ = cap_ = 0; return r; }   // for j2kcuda_free
};
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
struct j2kcuda_decoder;
namespace j2k {
void encoder_set_throughput_mode(::j2kcuda_encoder* e, bool many_frames_in_flight);
// decode only tiles [first, first + count) of the frame (count 0 = all): the rest of the frame is another GPU's
void decoder_set_tile_range(::j2kcuda_decoder* d, uint32_t first, uint32_t count);
}
