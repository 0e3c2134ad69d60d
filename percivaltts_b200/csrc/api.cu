// Library-level entry points of libpctts.so: ABI version and device capability check.
#include "common.cuh"
#include "pctts.h"
#include <atomic>

static std::atomic<long> g_launches{0};
void pctts_note_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
extern "C" long pctts_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

extern "C" int pctts_abi_version(void) { return PCTTS_ABI_VERSION; }

// The kernels are compiled for sm_100a only (tcgen05 / TMEM / clusters): refuse anything else loudly.
extern "C" int pctts_device_check(int dev) {
    int count = 0;
    PCTTS_CUDA(cudaGetDeviceCount(&count));
    if (dev < 0 || dev >= count) return PCTTS_EINVAL;
    int major = 0, minor = 0;
    PCTTS_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    PCTTS_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
    (void)minor;
    return major == 10 ? PCTTS_OK : PCTTS_EUNSUPPORTED;
}
