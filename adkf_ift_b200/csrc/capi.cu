// capi.cu -- version entry point and launch accounting of libadkf_b200.so (the other extern "C" symbols live
// beside their kernels).
#include <atomic>
#include "../../include/adkf_b200.h"

static std::atomic<long long> g_launches{0};

extern "C" int adkf_version(void) { return 100; }  // 0.1.0 -> major*10000 + minor*100 + patch
extern "C" void adkf_count_launches(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
extern "C" int64_t adkf_launch_count(void) { return (int64_t)g_launches.load(std::memory_order_relaxed); }
