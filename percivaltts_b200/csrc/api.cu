// Library-level entry points of libpctts.so: ABI version and device capability check.
#include "common.cuh"
#include "pctts.h"
#include <atomic>
#include <stdlib.h>

static std::atomic<long> g_launches{0};
void pctts_note_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
extern "C" long pctts_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

// programmatic dependent launch (common.cuh): off unless PCTTS_PDL=1 -- measured on B200 inside the captured step graphs it
// changes nothing (3.986 vs 3.972 ms per step), so the plain launch stays the default; pctts_set_pdl() flips it at run time
static int g_pdl = -1;
int pctts_pdl_enabled() {
    if (g_pdl < 0) {
        const char* e = getenv("PCTTS_PDL");
        g_pdl = (e && atoi(e) != 0) ? 1 : 0;
    }
    return g_pdl;
}
extern "C" int pctts_set_pdl(int on) {
    g_pdl = on ? 1 : 0;
    return PCTTS_OK;
}

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
