// PML vocoder band compression / decompression (bandwidth-bound).
// Reference call sites: vocoders.py:134-145 (compress_spectrum: sp.linbnd2fwbnd(log|SPEC|)), vocoders.py:147-166
// (decompress_spectrum: exp(sp.fwbnd2linbnd(., smooth=True))), vocoders.py:202 (noise mask, no smoothing),
// pulsemodel/analysis.py:336,366.  The interpolation tables (index + weight per output column) are inputs, built on
// the host by percivaltts_b200/sigproc.py, so the kernels carry no frequency-warping constants themselves.
#include "common.cuh"
#include "pctts.h"

namespace {

constexpr int WARPS = 8;

// one warp per frame: stage the 2049-bin row in shared memory with coalesced loads, then gather the K bands
__global__ void __launch_bounds__(WARPS * 32)
pml_compress_kernel(const float* __restrict__ in, const int* __restrict__ idx, const float* __restrict__ wgt,
                    float* __restrict__ out, long T, int nbins, int K, int log_abs) {
    pdl_entry();
    extern __shared__ float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* row = smem + (long)warp * nbins;
    for (long t = (long)blockIdx.x * WARPS + warp; t < T; t += (long)gridDim.x * WARPS) {
        const float* src = in + t * nbins;
        int i = lane;
        for (; i + 7 * 32 < nbins; i += 8 * 32) {
            float v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = __ldcs(src + i + u * 32);
#pragma unroll
            for (int u = 0; u < 8; ++u) row[i + u * 32] = v[u];
        }
        for (; i < nbins; i += 32) row[i] = __ldcs(src + i);
        __syncwarp();
        float* dst = out + t * K;
        for (int k = lane; k < K; k += 32) {
            const int j = __ldg(idx + k);
            const float w = __ldg(wgt + k);
            float a = row[j], b = row[min(j + 1, nbins - 1)];
            if (log_abs) {
                a = logf(fabsf(a));
                b = logf(fabsf(b));
            }
            dst[k] = (w == 0.f) ? a : fmaf(w, b - a, a);
        }
        __syncwarp();
    }
}

__global__ void __launch_bounds__(WARPS * 32)
pml_decompress_kernel(const float* __restrict__ in, const int* __restrict__ idx, const float* __restrict__ wgt,
                      float* __restrict__ out, long T, int K, int nbins, int smooth, int do_exp) {
    pdl_entry();
    extern __shared__ float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* row = smem + (long)warp * K;
    for (long t = (long)blockIdx.x * WARPS + warp; t < T; t += (long)gridDim.x * WARPS) {
        const float* src = in + t * K;
        for (int k = lane; k < K; k += 32) row[k] = __ldcs(src + k);
        __syncwarp();
        float* dst = out + t * nbins;
        for (int i = lane; i < nbins; i += 32) {
            const int j = __ldg(idx + i);
            const float w = __ldg(wgt + i);
            const float p1 = row[j], p2 = row[min(j + 1, K - 1)];
            float v;
            if (smooth) {
                const float p0 = row[max(j - 1, 0)], p3 = row[min(j + 2, K - 1)];
                const float w2 = w * w, w3 = w2 * w;
                v = 0.5f * (2.f * p1 + (p2 - p0) * w + (2.f * p0 - 5.f * p1 + 4.f * p2 - p3) * w2 +
                            (3.f * p1 - p0 - 3.f * p2 + p3) * w3);
            } else {
                v = (w == 0.f) ? p1 : fmaf(w, p2 - p1, p1);
            }
            if (do_exp) v = expf(v);
            __stcs(dst + i, v);
        }
        __syncwarp();
    }
}

}  // namespace

extern "C" int pctts_pml_linbnd2fwbnd(const float* in, const int* idx, const float* wgt, float* out, long T, int nbins, int K,
                                      int log_abs, void* stream) {
    if (!in || !idx || !wgt || !out || T < 0 || nbins <= 1 || K <= 0) return PCTTS_EINVAL;
    if (T == 0) return PCTTS_OK;
    const size_t smem = (size_t)WARPS * nbins * sizeof(float);
    if (smem > 227 * 1024) return PCTTS_EUNSUPPORTED;
    PCTTS_CUDA(cudaFuncSetAttribute(pml_compress_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long blocks = (T + WARPS - 1) / WARPS;
    if (blocks > 148 * 3) blocks = 148 * 3;
    pctts_launch(pml_compress_kernel, (int)blocks, WARPS * 32, smem, (cudaStream_t)stream, in, idx, wgt, out, T, nbins, K, log_abs);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

extern "C" int pctts_pml_fwbnd2linbnd(const float* in, const int* idx, const float* wgt, float* out, long T, int K, int nbins,
                                      int smooth, int do_exp, void* stream) {
    if (!in || !idx || !wgt || !out || T < 0 || nbins <= 1 || K <= 1) return PCTTS_EINVAL;
    if (T == 0) return PCTTS_OK;
    const size_t smem = (size_t)WARPS * K * sizeof(float);
    if (smem > 227 * 1024) return PCTTS_EUNSUPPORTED;
    if (smem > 48 * 1024)
        PCTTS_CUDA(cudaFuncSetAttribute(pml_decompress_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long blocks = (T + WARPS - 1) / WARPS;
    if (blocks > 148 * 8) blocks = 148 * 8;
    pctts_launch(pml_decompress_kernel, (int)blocks, WARPS * 32, smem, (cudaStream_t)stream, in, idx, wgt, out, T, K, nbins, smooth, do_exp);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}
