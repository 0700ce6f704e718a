#include <stdarg.h>
#include <atomic>
#include "common.cuh"
#include "../../include/maavss_b200.h"

static thread_local char g_err[512] = "";

void maavss_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* maavss_last_error(void) { return g_err; }
extern "C" int maavss_abi_version(void) { return MAAVSS_ABI_VERSION; }
extern "C" int maavss_device_sm_count(void) { return maavss_num_sms(); }

static std::atomic<long long> g_launches{0};
void maavss_count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
extern "C" long long maavss_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
