// api.cu — library-wide state of libhfops: version, per-thread last error, launch counter.
#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "common.cuh"

namespace hf {
namespace {
thread_local char g_err[512] = "";
std::atomic<unsigned long long> g_launches{0};
}  // namespace

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
void count_launch(unsigned n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
}  // namespace hf

extern "C" const char *hf_version(void) { return "hfops-b200 0.1.0 (sm_100a)"; }
extern "C" const char *hf_last_error(void) { return hf::g_err; }
extern "C" unsigned long long hf_launch_count(void) { return hf::g_launches.load(std::memory_order_relaxed); }
