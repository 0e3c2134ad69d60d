// Weight gradient of the 5x5, 4 -> 4 channel Conv2D (and the bias gradient in the same pass):
//     dW[kt,kf,ci,co] = sum_{n,t,f} x[n, t+kt-2, f+kf-2, ci] * dy[n,t,f,co],      db[co] = sum_{n<nb_bias,t,f} dy[n,t,f,co]
// Reference ops: gradients of pCNN2D / Conv2D (networktts.py:122-126, networks_critic.py:66-68) as TensorFlow's
// Conv2DBackpropFilter / BiasAddGrad would produce them.
//
// The result is 400 + 4 numbers and the reduction runs over ~1.5 M pixels: CUDA-core work with register-resident accumulators.
// Persistent CTAs (one per SM, 15 warps) walk over tiles of 8 frame rows x F bins, staged in shared memory by cp.async (zero-filled
// halo, double-buffered).  Warp (kt, part) owns one kernel row; lane (row r, segment) slides along its run of SEG pixels with
// the 5-pixel input window held in registers, so that a pixel costs 2 LDS.128 (one new window pixel, one dy pixel) for
// 80 FMAs.  The 80 (+4) accumulators stay in registers over all the tiles of the CTA; one warp reduction at the end, then the
// deterministic partial sum over CTAs (reduce_partials_kernel in conv2d.cu).
#include "common.cuh"
#include "pctts.h"

namespace {

constexpr int W4_TT = 8;                 // output rows per tile
constexpr int W4_PARTS = 3;              // a kernel row is shared by 3 warps = 96 (row, segment) units
constexpr int W4_WARPS = 5 * W4_PARTS;   // 15 warps, one CTA per SM (112 registers per thread: 80 accumulators + window)
constexpr int W4_THREADS = W4_WARPS * 32;
constexpr int W4_NW = 400;

struct W4Params {
    const float* x;
    const float* dy;
    float* partial;      // [grid][400] then [grid][4]
    int NB, T, F, nb_bias;
    int tiles_per_sample, ntiles;
    long xpitch, dypitch, xbatch, dybatch;
};

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool ok) {
    const int sz = ok ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void fma_px(float (&acc)[5][4][4], const float4& w0, const float4& w1, const float4& w2, const float4& w3,
                                       const float4& w4, const float4& d) {
    const float4 win[5] = {w0, w1, w2, w3, w4};
#pragma unroll
    for (int kf = 0; kf < 5; ++kf) {
        const float xv[4] = {win[kf].x, win[kf].y, win[kf].z, win[kf].w};
#pragma unroll
        for (int ci = 0; ci < 4; ++ci) {
            acc[kf][ci][0] = fmaf(xv[ci], d.x, acc[kf][ci][0]);
            acc[kf][ci][1] = fmaf(xv[ci], d.y, acc[kf][ci][1]);
            acc[kf][ci][2] = fmaf(xv[ci], d.z, acc[kf][ci][2]);
            acc[kf][ci][3] = fmaf(xv[ci], d.w, acc[kf][ci][3]);
        }
    }
}

__global__ void __launch_bounds__(W4_THREADS, 1) conv2d_wgrad44_kernel(const W4Params p) {
    pdl_entry();
    extern __shared__ __align__(16) float4 smem4[];
    const int F = p.F, FW = F + 4;
    const int xs_n = (W4_TT + 4) * FW, dys_n = W4_TT * F;       // float4 (pixel) counts
    const int stage_n = xs_n + dys_n;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int kt = warp % 5, half = warp / 5;
    const int r = lane & 7, seg = (lane >> 3) + 4 * half;
    constexpr int NSEG = 4 * W4_PARTS;
    const int SEG = (F + NSEG - 1) / NSEG;
    const int f0 = seg * SEG;
    const int len = max(0, min(SEG, F - f0));

    auto stage_tile = [&](int ti, int s) {
        const int n = ti / p.tiles_per_sample;
        const int t0 = (ti - n * p.tiles_per_sample) * W4_TT;
        const float* xb = p.x + (long)n * p.xbatch;
        const float* db = p.dy + (long)n * p.dybatch;
        float4* xs = smem4 + (size_t)s * stage_n;
        float4* dys = xs + xs_n;
        // one warp per staged row (12 input rows, then 8 dy rows): no divisions, 16-byte cp.async with zero fill outside the map
        for (int q = warp; q < W4_TT + 4 + W4_TT; q += W4_WARPS) {
            if (q < W4_TT + 4) {
                const int t = t0 + q - 2;
                const bool rowok = t >= 0 && t < p.T;
                const float* src = xb + (long)(rowok ? t : 0) * p.xpitch;
                const uint32_t dst = (uint32_t)__cvta_generic_to_shared(xs + (size_t)q * FW);
                for (int c = lane; c < FW; c += 32) {
                    const int f = c - 2;
                    const bool ok = rowok && f >= 0 && f < F;
                    cp_async16(dst + 16u * c, ok ? src + (long)f * 4 : xb, ok);
                }
            } else {
                const int rr = q - (W4_TT + 4), t = t0 + rr;
                const bool rowok = t < p.T;
                const float* src = db + (long)(rowok ? t : 0) * p.dypitch;
                const uint32_t dst = (uint32_t)__cvta_generic_to_shared(dys + (size_t)rr * F);
                for (int c = lane; c < F; c += 32) cp_async16(dst + 16u * c, rowok ? src + (long)c * 4 : db, rowok);
            }
        }
        cp_async_commit();
    };

    float acc[5][4][4];
#pragma unroll
    for (int a = 0; a < 5; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b)
#pragma unroll
            for (int c = 0; c < 4; ++c) acc[a][b][c] = 0.f;
    float4 bsum = make_float4(0.f, 0.f, 0.f, 0.f);

    int it = 0;
    if ((int)blockIdx.x < p.ntiles) stage_tile(blockIdx.x, 0);
    for (int ti = blockIdx.x; ti < p.ntiles; ti += gridDim.x, ++it) {
        const int s = it & 1;
        const int nxt = ti + gridDim.x;
        if (nxt < p.ntiles) {
            stage_tile(nxt, s ^ 1);
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        const float4* xs = smem4 + (size_t)s * stage_n + (size_t)(r + kt) * FW + f0;
        const float4* dys = smem4 + (size_t)s * stage_n + xs_n + (size_t)r * F + f0;
        const bool with_bias = kt == 0 && (ti / p.tiles_per_sample) < p.nb_bias;
        if (len > 0) {
            float4 w0 = xs[0], w1 = xs[1], w2 = xs[2], w3 = xs[3], w4;
            int i = 0;
            // five pixels per trip: the window rotates through the registers without moves
            for (; i + 5 <= len; i += 5) {
                float4 d;
                w4 = xs[i + 4], d = dys[i];
                fma_px(acc, w0, w1, w2, w3, w4, d);
                if (with_bias) bsum.x += d.x, bsum.y += d.y, bsum.z += d.z, bsum.w += d.w;
                w0 = xs[i + 5], d = dys[i + 1];
                fma_px(acc, w1, w2, w3, w4, w0, d);
                if (with_bias) bsum.x += d.x, bsum.y += d.y, bsum.z += d.z, bsum.w += d.w;
                w1 = xs[i + 6], d = dys[i + 2];
                fma_px(acc, w2, w3, w4, w0, w1, d);
                if (with_bias) bsum.x += d.x, bsum.y += d.y, bsum.z += d.z, bsum.w += d.w;
                w2 = xs[i + 7], d = dys[i + 3];
                fma_px(acc, w3, w4, w0, w1, w2, d);
                if (with_bias) bsum.x += d.x, bsum.y += d.y, bsum.z += d.z, bsum.w += d.w;
                w3 = xs[i + 8], d = dys[i + 4];
                fma_px(acc, w4, w0, w1, w2, w3, d);
                if (with_bias) bsum.x += d.x, bsum.y += d.y, bsum.z += d.z, bsum.w += d.w;
            }
            for (; i < len; ++i) {
                w4 = xs[i + 4];
                const float4 d = dys[i];
                fma_px(acc, w0, w1, w2, w3, w4, d);
                if (with_bias) bsum.x += d.x, bsum.y += d.y, bsum.z += d.z, bsum.w += d.w;
                w0 = w1, w1 = w2, w2 = w3, w3 = w4;
            }
        }
        __syncthreads();      // every warp is done with stage s before the next trip refills it
    }

    // ---- per-CTA result: warp sums, then the two halves of a kernel row, in a fixed order
    float* red = reinterpret_cast<float*>(smem4);                // [W4_PARTS][404] (the stages are idle now)
#pragma unroll
    for (int kf = 0; kf < 5; ++kf)
#pragma unroll
        for (int ci = 0; ci < 4; ++ci)
#pragma unroll
            for (int co = 0; co < 4; ++co) {
                const float v = warp_sum(acc[kf][ci][co]);
                if (lane == 0) red[half * 404 + ((kt * 5 + kf) * 4 + ci) * 4 + co] = v;
            }
    if (kt == 0) {
        const float b0 = warp_sum(bsum.x), b1 = warp_sum(bsum.y), b2 = warp_sum(bsum.z), b3 = warp_sum(bsum.w);
        if (lane == 0) red[half * 404 + 400] = b0, red[half * 404 + 401] = b1, red[half * 404 + 402] = b2, red[half * 404 + 403] = b3;
    }
    __syncthreads();
    for (int i = tid; i < 404; i += W4_THREADS) {
        float v = red[i];
#pragma unroll
        for (int h = 1; h < W4_PARTS; ++h) v += red[h * 404 + i];
        if (i < 400) p.partial[(long)blockIdx.x * 400 + i] = v;
        else p.partial[(long)gridDim.x * 400 + (long)blockIdx.x * 4 + (i - 400)] = v;
    }
}

}  // namespace

// the last step (sum over CTAs) lives in conv2d.cu
int pctts_reduce_partials(const float* partial, float* out, int nblk, int NW, float scale, int accumulate, cudaStream_t st);

int pctts_conv2d_wgrad44_supported(int F, int CIN, int COUT, int KH, int KW, int pad_t, int pad_f, const void* x, const void* dy,
                                   long xpitch, long dypitch, long xbatch, long dybatch) {
    if (!(CIN == 4 && COUT == 4 && KH == 5 && KW == 5 && pad_t == 2 && pad_f == 2)) return 0;
    if ((((uintptr_t)x | (uintptr_t)dy) & 15) || ((xpitch | dypitch | xbatch | dybatch) & 3)) return 0;
    const size_t smem = 2 * (size_t)((W4_TT + 4) * (F + 4) + W4_TT * F) * 16;
    return smem <= 200 * 1024 ? 1 : 0;
}

int pctts_conv2d_wgrad44(const float* x, const float* dy, float* dw, float* dbias, int nb_bias, float* workspace, int NB, int T,
                         int F, long xpitch, long dypitch, long xbatch, long dybatch, float scale, int accumulate, cudaStream_t st) {
    W4Params p;
    p.x = x, p.dy = dy, p.partial = workspace;
    p.NB = NB, p.T = T, p.F = F, p.nb_bias = dbias ? nb_bias : 0;
    p.tiles_per_sample = (T + W4_TT - 1) / W4_TT;
    p.ntiles = NB * p.tiles_per_sample;
    p.xpitch = xpitch, p.dypitch = dypitch, p.xbatch = xbatch, p.dybatch = dybatch;
    const size_t smem = 2 * (size_t)((W4_TT + 4) * (F + 4) + W4_TT * F) * 16;
    static size_t attr_smem = 0;
    if (smem > attr_smem) {
        PCTTS_CUDA(cudaFuncSetAttribute(conv2d_wgrad44_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_smem = smem;
    }
    const int grid = p.ntiles < 148 ? p.ntiles : 148;
    pctts_launch(conv2d_wgrad44_kernel, grid, W4_THREADS, smem, st, p);
    PCTTS_LAUNCH_CHECK();
    int rc = pctts_reduce_partials(workspace, dw, grid, W4_NW, scale, accumulate, st);
    if (rc) return rc;
    if (dbias) rc = pctts_reduce_partials(workspace + (long)grid * 400, dbias, grid, 4, 1.f, 0, st);
    return rc;
}
