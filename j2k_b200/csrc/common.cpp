#include "common.h"
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>
namespace j2k {
std::string& last_error() { static thread_local std::string e; return e; }
namespace {
std::mutex g_log_mu;
void (*g_log_fn)(void*, int, const char*) = nullptr;
void* g_log_user = nullptr;
}  // namespace
void set_log(void (*fn)(void*, int, const char*), void* user) { std::lock_guard<std::mutex> g(g_log_mu); g_log_fn = fn; g_log_user = user; }
void log_msg(int level, const std::string& msg) {
    void (*fn)(void*, int, const char*); void* user;
    { std::lock_guard<std::mutex> g(g_log_mu); fn = g_log_fn; user = g_log_user; }
    if (fn) fn(user, level, msg.c_str());
}
int set_error(int code, const std::string& msg) { last_error() = msg; if (code) log_msg(0, msg); return code; }
bool context_cache_enabled() { static const bool on = [] { const char* e = getenv("J2KCUDA_CONTEXT_CACHE"); return !(e && e[0] == '0'); }(); return on; }

// Output buffers handed to the caller (codestreams, parse tables) and returned through j2kcuda_free().  A frame's
// codestream is tens of megabytes; a fresh allocation of that size costs more in first-touch page faults than the
// Tier-2 writer spends filling it, so large buffers that come back through j2kcuda_free() are kept (a few, bounded) and
// handed out again.  Every buffer carries a 16-byte header with its capacity.
namespace {
struct OutHeader { uint64_t magic, cap; };
constexpr uint64_t OUT_MAGIC = 0x6a326b6f75746266ull;
constexpr size_t KEEP_MIN = 1u << 20, KEEP_BUFFERS = 48, KEEP_BYTES = 1024u << 20;
std::mutex g_out_mu;
std::vector<OutHeader*> g_out_free;
size_t g_out_bytes = 0;
}  // namespace

void* out_alloc(size_t n) {
    if (n == 0) n = 1;
    if (n >= KEEP_MIN) {
        std::lock_guard<std::mutex> g(g_out_mu);
        size_t best = g_out_free.size();
        for (size_t i = 0; i < g_out_free.size(); ++i)
            if (g_out_free[i]->cap >= n && g_out_free[i]->cap <= 2 * n && (best == g_out_free.size() || g_out_free[i]->cap < g_out_free[best]->cap)) best = i;
        if (best != g_out_free.size()) {
            OutHeader* h = g_out_free[best];
            g_out_free.erase(g_out_free.begin() + (long)best);
            g_out_bytes -= h->cap;
            return h + 1;
        }
    }
    const size_t cap = n >= KEEP_MIN ? n + n / 8 : n;          // a little slack: the next frame's stream is rarely the same size
    OutHeader* h = (OutHeader*)malloc(sizeof(OutHeader) + cap);
    if (!h) return nullptr;
    h->magic = OUT_MAGIC; h->cap = cap;
    return h + 1;
}

void out_free(void* p) {
    if (!p) return;
    OutHeader* h = (OutHeader*)p - 1;
    if (h->magic != OUT_MAGIC) return;                          // not one of ours: leave it alone rather than corrupt the heap
    if (h->cap >= KEEP_MIN) {
        std::lock_guard<std::mutex> g(g_out_mu);
        if (g_out_free.size() < KEEP_BUFFERS && g_out_bytes + h->cap <= KEEP_BYTES) {
            g_out_free.push_back(h);
            g_out_bytes += h->cap;
            return;
        }
    }
    h->magic = 0;
    free(h);
}

void out_release_all() {
    std::lock_guard<std::mutex> g(g_out_mu);
    for (OutHeader* h : g_out_free) { h->magic = 0; free(h); }
    g_out_free.clear();
    g_out_bytes = 0;
}
}  // namespace j2k
