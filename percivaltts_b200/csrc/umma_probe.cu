// Developer probe: issue rate of tcgen05.mma (M = 128, K = 16, bf16) as a function of the shared-memory operand layout of A
// (no swizzle with the chain kernel's two-array descriptor, SWIZZLE_32B / 64B / 128B K-major) and of N.  One CTA, one warp issues
// `nmma` MMAs back to back, walking the A start address like the chain kernel does (row shifts), one commit at the end.
// Not on any product path (tools/umma_probe.py); it answers "is 49 cycles per MMA a property of the operand form?" (DESIGN.md 4.10).
#include "common.cuh"
#include "pctts.h"
#include "tc_common.cuh"

namespace {
using namespace tcx;

__device__ __forceinline__ uint64_t desc_any(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
    uint64_t d = 0;
    d |= (uint64_t)((addr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)(layout & 7) << 61;
    return d;
}

__global__ void __launch_bounds__(128, 1) umma_probe_kernel(int layout, int N, int nmma, int shift_rows, long long* out) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    for (int i = threadIdx.x; i < 160 * 1024 / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem_raw)[i] = make_uint4(0, 0, 0, 0);
    if (threadIdx.x == 0) {
        mbar_init(smem_u32(&bar), 1);
        fence_barrier_init();
    }
    const int warp = threadIdx.x >> 5;
    if (warp == 0) tmem_alloc<512>(smem_u32(&tmem_slot));
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_slot;
    if (warp == 0) {
        const uint32_t idesc = umma_idesc_bf16(128, N, 0, 0);
        // A region: first 96 KB; B region: at +128 KB (no swizzle, chunks of N rows x 16 B)
        uint32_t a_lbo, a_sbo, lt, rowb;
        if (layout == 0) { a_lbo = 40960u + 16u, a_sbo = 128u, lt = 0, rowb = 16u; }          // chain form: two parity arrays
        else if (layout == 1) { a_lbo = 16u, a_sbo = 256u, lt = 6, rowb = 32u; }                 // SWIZZLE_32B, rows of 32 B
        else if (layout == 2) { a_lbo = 16u, a_sbo = 512u, lt = 4, rowb = 64u; }                 // SWIZZLE_64B
        else if (layout == 3) { a_lbo = 16u, a_sbo = 1024u, lt = 2, rowb = 128u; }               // SWIZZLE_128B
        else { a_lbo = 2048u, a_sbo = 128u, lt = 0, rowb = 16u; }                                // plain no-swizzle, K chunks 2 KB apart
        const uint64_t bd = desc_any(base + 128u * 1024u, (uint32_t)N * 16u, 128u, 0);
        long long t0 = 0, t1 = 0, t2 = 0;
        if (elect_one()) {
            t0 = clock64();
            for (int i = 0; i < nmma; ++i) {
                const uint32_t off = (uint32_t)((i % 10) * shift_rows) * rowb + (uint32_t)((i / 10) % 8) * 128u * rowb / 8u;
                const uint64_t ad = desc_any(base + off, a_lbo, a_sbo, lt);
                umma_bf16(tmem + (uint32_t)((i & 1) * 128), ad, bd, idesc, i > 1 ? 1u : 0u);
            }
            t1 = clock64();
            umma_commit(smem_u32(&bar));
        }
        __syncwarp();
        mbar_wait(smem_u32(&bar), 0);
        t2 = clock64();
        if (elect_one()) out[0] = t1 - t0, out[1] = t2 - t0;
        __syncwarp();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tmem);
}
}  // namespace

/* developer probe: out[0] = cycles spent issuing, out[1] = cycles until the last of `nmma` tcgen05.mma (M = 128, N, K = 16) completed.
 * layout: 0 chain-kernel form (two parity arrays, no swizzle), 1 / 2 / 3 SWIZZLE_32B / 64B / 128B K-major, 4 plain no swizzle */
extern "C" int pctts_umma_probe(int layout, int N, int nmma, int shift_rows, long long* out, void* stream) {
    if (!out || N < 8 || N > 256 || (N & 7) || nmma < 1 || layout < 0 || layout > 4) return PCTTS_EINVAL;
    const int smem = 200 * 1024;
    PCTTS_CUDA(cudaFuncSetAttribute(umma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    umma_probe_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(layout, N, nmma, shift_rows, out);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}
