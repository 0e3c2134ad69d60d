// Weight gradient of the 5x5, 4 -> 4 (and 1 -> 4) channel Conv2D layers, with the bias gradient in the same pass, for ALL the layers
// of a stack in one launch:
//     dW[kt,kf,ci,co] = sum_{n,t,f} x[n, t+kt-2, f+kf-2, ci] * dy[n,t,f,co],      db[co] = sum_{n<nb_bias,t,f} dy[n,t,f,co]
// Reference ops: gradients of pCNN2D / Conv2D (networktts.py:122-126, networks_critic.py:66-68) as TensorFlow's
// Conv2DBackpropFilter / BiasAddGrad would produce them.
//
// The result is 400 + 4 numbers per layer and the reduction runs over ~1.5 M pixels: 50 flop per byte read, i.e. bound by the fp32
// FMA pipe, not by HBM; a tensor-core form would have M x N = 100 x 4, so this is CUDA-core work with
// register-resident accumulators.  The CTAs of the grid are divided between the layers (148 / nlayers each), so a stack of 7 layers
// is one launch with ~24 tiles per CTA instead of 7 launches of 3-7 tiles each with their fill, drain and partial-sum passes.
// A CTA (15 FMA warps + 1 producer warp) walks over tiles of 16 frame rows x F bins of its layer, staged in a two-stage shared-memory
// ring by the TMA engine (cp.async.bulk per row, mbarrier full / empty; conv2d_wgrad44_pipe_kernel below).  Warp (kt, part) owns one kernel row; lane (row r, segment) slides along its run of SEG = 22 pixels with
// the 5-pixel input window held in registers, so that a pixel costs 2 LDS.128 (one new window pixel, one dy pixel) for
// 80 FMAs.  The 80 (+4) accumulators stay in registers over all the tiles of the CTA; one warp reduction at the end, then the
// deterministic partial sum over the CTAs of the layer (wgrad_multi_reduce_kernel, one launch for all layers).
#include "common.cuh"
#include "pctts.h"
#include "tc_common.cuh"

namespace {

constexpr int W4_TT = 16;                // output rows per tile (one per lane & 15)
constexpr int W4_PARTS = 3;              // a kernel row is shared by 3 warps = 96 (row, segment) units: 16 rows x 6 segments
constexpr int W4_WARPS = 5 * W4_PARTS;   // 15 warps, one CTA per SM (112 registers per thread: 80 accumulators + window)
constexpr int W4_THREADS = W4_WARPS * 32;
constexpr int W4_NW = 400;
constexpr int W4_NP = 404;               // numbers per CTA partial: 400 kernel entries + 4 bias entries
constexpr int W4_MAXL = PCTTS_WGRAD_MAX_LAYERS;

struct W4Layer {
    const float* x;      // [NB,T,F,cin]
    const float* dy;     // [NB,T,F,4]
    float* dw;           // [5][5][cin][4]
    float* dbias;        // [4] or NULL
    long xpitch, dypitch, xbatch, dybatch;
    int cin;             // 4, or 1 (the pixel lands in channel 0 of the staged float4, the others stay zero)
};

struct W4Params {
    W4Layer L[W4_MAXL];
    float* partial;      // [nlayers][cpl][404]
    int nlayers, cpl;    // CTAs per layer
    int NB, T, F, nb_bias;
    int tiles_per_sample, ntiles;
    float scale;
    int accumulate;
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

__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src, bool ok) {
    const int sz = ok ? 4 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}

__global__ void __launch_bounds__(W4_THREADS, 1) conv2d_wgrad44_kernel(const W4Params p) {
    pdl_entry();
    extern __shared__ __align__(16) float4 smem4[];
    const int layer = (int)blockIdx.x / p.cpl, cta = (int)blockIdx.x - layer * p.cpl;
    const W4Layer& Ly = p.L[layer];
    const int F = p.F, FW = F + 4;
    const int xs_n = (W4_TT + 4) * FW, dys_n = W4_TT * F;       // float4 (pixel) counts
    const int stage_n = xs_n + dys_n;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int kt = warp % 5, half = warp / 5;
    // a lane owns one of the tile's 16 rows and one of 6 segments of the frequency axis (22 pixels at F = 129): long runs keep the
    // per-run set-up (addresses, the first four window pixels) and the run's tail small against the 80 FMAs per pixel
    const int r = lane & 15, seg = (lane >> 4) + 2 * half;
    constexpr int NSEG = 2 * W4_PARTS;
    const int SEG = (F + NSEG - 1) / NSEG;
    const int f0 = seg * SEG;
    const int len = max(0, min(SEG, F - f0));
    const bool one_ch = Ly.cin == 1;

    if (one_ch) {       // channels 1-3 of the staged input pixels are never written by the 4-byte copies: clear both stages once
        for (int i = tid; i < 2 * stage_n; i += W4_THREADS) smem4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        __syncthreads();
    }

    auto stage_tile = [&](int ti, int s) {
        const int n = ti / p.tiles_per_sample;
        const int t0 = (ti - n * p.tiles_per_sample) * W4_TT;
        const float* xb = Ly.x + (long)n * Ly.xbatch;
        const float* db = Ly.dy + (long)n * Ly.dybatch;
        float4* xs = smem4 + (size_t)s * stage_n;
        float4* dys = xs + xs_n;
        // one warp per staged row (20 input rows, then 16 dy rows): no divisions, cp.async with zero fill outside the map
        for (int q = warp; q < W4_TT + 4 + W4_TT; q += W4_WARPS) {
            if (q < W4_TT + 4) {
                const int t = t0 + q - 2;
                const bool rowok = t >= 0 && t < p.T;
                const float* src = xb + (long)(rowok ? t : 0) * Ly.xpitch;
                const uint32_t dst = (uint32_t)__cvta_generic_to_shared(xs + (size_t)q * FW);
                if (one_ch) {
                    for (int c = lane; c < FW; c += 32) {
                        const int f = c - 2;
                        const bool ok = rowok && f >= 0 && f < F;
                        cp_async4(dst + 16u * c, ok ? src + f : xb, ok);
                    }
                } else {
                    for (int c = lane; c < FW; c += 32) {
                        const int f = c - 2;
                        const bool ok = rowok && f >= 0 && f < F;
                        cp_async16(dst + 16u * c, ok ? src + (long)f * 4 : xb, ok);
                    }
                }
            } else {
                const int rr = q - (W4_TT + 4), t = t0 + rr;
                const bool rowok = t < p.T;
                const float* src = db + (long)(rowok ? t : 0) * Ly.dypitch;
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
    if (cta < p.ntiles) stage_tile(cta, 0);
    for (int ti = cta; ti < p.ntiles; ti += p.cpl, ++it) {
        const int s = it & 1;
        const int nxt = ti + p.cpl;
        if (nxt < p.ntiles) {
            stage_tile(nxt, s ^ 1);
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        const bool with_bias = kt == 0 && (ti / p.tiles_per_sample) < p.nb_bias;
        if (len > 0) {
            const float4* xs = smem4 + (size_t)s * stage_n + (size_t)(r + kt) * FW + f0;
            const float4* dys = smem4 + (size_t)s * stage_n + xs_n + (size_t)r * F + f0;
            float4 w0 = xs[0], w1 = xs[1], w2 = xs[2], w3 = xs[3], w4;
            float4 d;
#define W4_PX(WA, WB, WC, WD, WE, OFF)                                                     \
    WE = xs[i + (OFF) + 4], d = dys[i + (OFF)];                                            \
    fma_px(acc, WA, WB, WC, WD, WE, d);                                                    \
    if (with_bias) bsum.x += d.x, bsum.y += d.y, bsum.z += d.z, bsum.w += d.w;
            int i = 0;
            // five pixels per trip: the window rotates through the registers without moves
            for (; i + 5 <= len; i += 5) {
                W4_PX(w0, w1, w2, w3, w4, 0)
                W4_PX(w1, w2, w3, w4, w0, 1)
                W4_PX(w2, w3, w4, w0, w1, 2)
                W4_PX(w3, w4, w0, w1, w2, 3)
                W4_PX(w4, w0, w1, w2, w3, 4)
            }
            // the run's last len % 5 pixels: the same rotation, cut short (no register moves)
            const int rem = len - i;
            if (rem > 0) {
                W4_PX(w0, w1, w2, w3, w4, 0)
                if (rem > 1) {
                    W4_PX(w1, w2, w3, w4, w0, 1)
                    if (rem > 2) {
                        W4_PX(w2, w3, w4, w0, w1, 2)
                        if (rem > 3) {
                            W4_PX(w3, w4, w0, w1, w2, 3)
                        }
                    }
                }
            }
#undef W4_PX
        }
        __syncthreads();      // every warp is done with stage s before the next trip refills it
    }

    // ---- per-CTA result: warp sums, then the three parts of a kernel row, in a fixed order
    float* red = reinterpret_cast<float*>(smem4);                // [W4_PARTS][404] (the stages are idle now)
#pragma unroll
    for (int kf = 0; kf < 5; ++kf)
#pragma unroll
        for (int ci = 0; ci < 4; ++ci)
#pragma unroll
            for (int co = 0; co < 4; ++co) {
                const float v = warp_sum(acc[kf][ci][co]);
                if (lane == 0) red[half * W4_NP + ((kt * 5 + kf) * 4 + ci) * 4 + co] = v;
            }
    if (kt == 0) {
        const float b0 = warp_sum(bsum.x), b1 = warp_sum(bsum.y), b2 = warp_sum(bsum.z), b3 = warp_sum(bsum.w);
        if (lane == 0) red[half * W4_NP + 400] = b0, red[half * W4_NP + 401] = b1, red[half * W4_NP + 402] = b2, red[half * W4_NP + 403] = b3;
    }
    __syncthreads();
    float* const out = p.partial + ((long)layer * p.cpl + cta) * W4_NP;
    for (int i = tid; i < W4_NP; i += W4_THREADS) {
        float v = red[i];
#pragma unroll
        for (int h = 1; h < W4_PARTS; ++h) v += red[h * W4_NP + i];
        out[i] = v;
    }
}

// ---- the same contraction with a PRODUCER WARP and an mbarrier pipeline (the default) ---------------------------------------------
// Warp 15 feeds the two stages: one bulk copy of the TMA engine per staged row (cp.async.bulk, F x 16 bytes; rows outside the map
// are cleared by the lane instead; the single-channel first layer goes through 4-byte cp.async that arrive on the same mbarrier),
// every lane arriving once per tile on the stage's `full` barrier.  The 15 consumer warps wait for `full`, run the FMAs and arrive
// on `empty` -- no CTA-wide barrier and no staging instruction in the FMA warps: a warp that finishes a tile early (shorter
// segment, no bias sums) starts the next one, and the issue slots of the staging go to one warp that has nothing else to do.
constexpr int W4P_THREADS = W4_THREADS + 32;

__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
                 "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_arrive_noinc(uint32_t bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}

__global__ void __launch_bounds__(W4P_THREADS, 1) conv2d_wgrad44_pipe_kernel(const W4Params p) {
    pdl_entry();
    extern __shared__ __align__(16) float4 smem4[];
    __shared__ __align__(8) uint64_t full_bar[2], empty_bar[2];
    using namespace tcx;
    const int layer = (int)blockIdx.x / p.cpl, cta = (int)blockIdx.x - layer * p.cpl;
    const W4Layer& Ly = p.L[layer];
    const int F = p.F, FW = F + 4;
    const int xs_n = (W4_TT + 4) * FW, dys_n = W4_TT * F;       // float4 (pixel) counts
    const int stage_n = xs_n + dys_n;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const bool one_ch = Ly.cin == 1;

    // margins of the input rows (2 pixels on either side) are never written by the copies; with one input channel neither are the
    // channels 1-3: clear both stages once
    for (int i = tid; i < 2 * stage_n; i += W4P_THREADS) smem4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (tid == 0) {
        for (int s = 0; s < 2; ++s) {
            mbar_init(smem_u32(&full_bar[s]), 32);
            mbar_init(smem_u32(&empty_bar[s]), W4_WARPS);
        }
        fence_barrier_init();
    }
    fence_proxy_async();
    __syncthreads();

    if (warp == W4_WARPS) {
        // ------------------------------------------------ producer warp ------------------------------------------------
        const uint32_t rowbytes = (uint32_t)F * 16u;
        int it = 0;
        for (int ti = cta; ti < p.ntiles; ti += p.cpl, ++it) {
            const int s = it & 1;
            mbar_wait(smem_u32(&empty_bar[s]), (((uint32_t)it >> 1) & 1u) ^ 1u);
            const int n = ti / p.tiles_per_sample;
            const int t0 = (ti - n * p.tiles_per_sample) * W4_TT;
            const float* xb = Ly.x + (long)n * Ly.xbatch;
            const float* db = Ly.dy + (long)n * Ly.dybatch;
            float4* xs = smem4 + (size_t)s * stage_n;
            float4* dys = xs + xs_n;
            const uint32_t fb = smem_u32(&full_bar[s]);
            uint32_t tx = 0;
            bool used_cp = false;
            for (int q = lane; q < W4_TT + 4 + W4_TT; q += 32) {
                if (q < W4_TT + 4) {
                    const int t = t0 + q - 2;
                    float4* row = xs + (size_t)q * FW + 2;
                    if (t >= 0 && t < p.T) {
                        if (one_ch) {
                            const float* src = xb + (long)t * Ly.xpitch;
                            const uint32_t dst = smem_u32(row);
                            for (int f = 0; f < F; ++f) cp_async4(dst + 16u * f, src + f, true);
                            used_cp = true;
                        } else {
                            bulk_g2s(smem_u32(row), xb + (long)t * Ly.xpitch, rowbytes, fb);
                            tx += rowbytes;
                        }
                    } else {
                        for (int f = 0; f < F; ++f) row[f] = make_float4(0.f, 0.f, 0.f, 0.f);
                        fence_proxy_async();
                    }
                } else {
                    const int rr = q - (W4_TT + 4), t = t0 + rr;
                    float4* row = dys + (size_t)rr * F;
                    if (t < p.T) {
                        bulk_g2s(smem_u32(row), db + (long)t * Ly.dypitch, rowbytes, fb);
                        tx += rowbytes;
                    } else {
                        for (int f = 0; f < F; ++f) row[f] = make_float4(0.f, 0.f, 0.f, 0.f);
                        fence_proxy_async();
                    }
                }
            }
            // exactly one arrival per lane and tile
            if (used_cp) {
                if (tx) mbar_expect_tx(fb, tx);
                cp_async_arrive_noinc(fb);
            } else if (tx) {
                mbar_arrive_expect_tx(fb, tx);
            } else {
                mbar_arrive(fb);
            }
        }
        return;
    }

    // ---------------------------------------------------- consumer warps ----------------------------------------------------
    const int kt = warp % 5, half = warp / 5;
    const int r = lane & 15, seg = (lane >> 4) + 2 * half;
    constexpr int NSEG = 2 * W4_PARTS;
    const int SEG = (F + NSEG - 1) / NSEG;
    const int f0 = seg * SEG;
    const int len = max(0, min(SEG, F - f0));

    float acc[5][4][4];
#pragma unroll
    for (int a = 0; a < 5; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b)
#pragma unroll
            for (int c = 0; c < 4; ++c) acc[a][b][c] = 0.f;
    float4 bsum = make_float4(0.f, 0.f, 0.f, 0.f);

    int it = 0;
    for (int ti = cta; ti < p.ntiles; ti += p.cpl, ++it) {
        const int s = it & 1;
        mbar_wait(smem_u32(&full_bar[s]), ((uint32_t)it >> 1) & 1u);
        const bool with_bias = kt == 0 && (ti / p.tiles_per_sample) < p.nb_bias;
        if (len > 0) {
            const float4* xs = smem4 + (size_t)s * stage_n + (size_t)(r + kt) * FW + f0;
            const float4* dys = smem4 + (size_t)s * stage_n + xs_n + (size_t)r * F + f0;
            float4 w0 = xs[0], w1 = xs[1], w2 = xs[2], w3 = xs[3], w4;
            float4 d;
#define W4_PX(WA, WB, WC, WD, WE, OFF)                                                     \
    WE = xs[i + (OFF) + 4], d = dys[i + (OFF)];                                            \
    fma_px(acc, WA, WB, WC, WD, WE, d);                                                    \
    if (with_bias) bsum.x += d.x, bsum.y += d.y, bsum.z += d.z, bsum.w += d.w;
            int i = 0;
            for (; i + 5 <= len; i += 5) {
                W4_PX(w0, w1, w2, w3, w4, 0)
                W4_PX(w1, w2, w3, w4, w0, 1)
                W4_PX(w2, w3, w4, w0, w1, 2)
                W4_PX(w3, w4, w0, w1, w2, 3)
                W4_PX(w4, w0, w1, w2, w3, 4)
            }
            const int rem = len - i;
            if (rem > 0) {
                W4_PX(w0, w1, w2, w3, w4, 0)
                if (rem > 1) {
                    W4_PX(w1, w2, w3, w4, w0, 1)
                    if (rem > 2) {
                        W4_PX(w2, w3, w4, w0, w1, 2)
                        if (rem > 3) {
                            W4_PX(w3, w4, w0, w1, w2, 3)
                        }
                    }
                }
            }
#undef W4_PX
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&empty_bar[s]));
    }

    // ---- per-CTA result (consumer warps only: named barrier 1): warp sums, then the three parts of a kernel row, in a fixed order
    asm volatile("bar.sync 1, %0;" ::"n"(W4_THREADS) : "memory");      // every consumer is done with the stages the sums overlay
    float* red = reinterpret_cast<float*>(smem4);
#pragma unroll
    for (int kf = 0; kf < 5; ++kf)
#pragma unroll
        for (int ci = 0; ci < 4; ++ci)
#pragma unroll
            for (int co = 0; co < 4; ++co) {
                const float v = warp_sum(acc[kf][ci][co]);
                if (lane == 0) red[half * W4_NP + ((kt * 5 + kf) * 4 + ci) * 4 + co] = v;
            }
    if (kt == 0) {
        const float b0 = warp_sum(bsum.x), b1 = warp_sum(bsum.y), b2 = warp_sum(bsum.z), b3 = warp_sum(bsum.w);
        if (lane == 0) red[half * W4_NP + 400] = b0, red[half * W4_NP + 401] = b1, red[half * W4_NP + 402] = b2, red[half * W4_NP + 403] = b3;
    }
    asm volatile("bar.sync 1, %0;" ::"n"(W4_THREADS) : "memory");
    float* const out = p.partial + ((long)layer * p.cpl + cta) * W4_NP;
    for (int i = tid; i < W4_NP; i += W4_THREADS) {
        float v = red[i];
#pragma unroll
        for (int h = 1; h < W4_PARTS; ++h) v += red[h * W4_NP + i];
        out[i] = v;
    }
}

// second stage for all layers: one warp per number, lanes stride over the layer's CTA partials, butterfly sum in a fixed order
__global__ void wgrad_multi_reduce_kernel(const W4Params p) {
    pdl_entry();
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int layer = blockIdx.y, lane = threadIdx.x & 31;
    if (i >= W4_NP) return;
    const W4Layer& Ly = p.L[layer];
    const float* part = p.partial + (long)layer * p.cpl * W4_NP + i;
    const int nblk = p.cpl < p.ntiles ? p.cpl : p.ntiles;
    float s0 = 0.f, s1 = 0.f;
    int b = lane;
    for (; b + 32 < nblk; b += 64) {
        s0 += part[(long)b * W4_NP];
        s1 += part[(long)(b + 32) * W4_NP];
    }
    if (b < nblk) s0 += part[(long)b * W4_NP];
    const float v = warp_sum(s0 + s1);
    if (lane != 0) return;
    if (i < W4_NW) {
        const int ci = (i >> 2) & 3;
        if (ci >= Ly.cin) return;
        const int o = Ly.cin == 4 ? i : (i >> 4) * 4 + (i & 3);
        const float w = v * p.scale;
        Ly.dw[o] = p.accumulate ? Ly.dw[o] + w : w;
    } else if (Ly.dbias) {
        Ly.dbias[i - W4_NW] = v;
    }
}

}  // namespace

namespace {

size_t w4_smem(int F) { return 2 * (size_t)((W4_TT + 4) * (F + 4) + W4_TT * F) * 16; }

int w4_launch(W4Params& p, cudaStream_t st) {
    p.tiles_per_sample = (p.T + W4_TT - 1) / W4_TT;
    p.ntiles = p.NB * p.tiles_per_sample;
    p.cpl = 148 / p.nlayers;
    if (p.cpl > p.ntiles) p.cpl = p.ntiles;
    const size_t smem = w4_smem(p.F);
    static size_t attr_smem = 0;
    if (smem > attr_smem) {
        PCTTS_CUDA(cudaFuncSetAttribute(conv2d_wgrad44_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_smem = smem;
    }
    static const int pipe = getenv("PCTTS_WGRAD_PIPE") ? atoi(getenv("PCTTS_WGRAD_PIPE")) : 1;
    // the bulk copies need 16-byte rows: F x 16 bytes always is; bases / pitches are checked by the callers
    if (pipe) {
        static size_t attr_smem_p = 0;
        if (smem > attr_smem_p) {
            PCTTS_CUDA(cudaFuncSetAttribute(conv2d_wgrad44_pipe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr_smem_p = smem;
        }
        pctts_launch(conv2d_wgrad44_pipe_kernel, p.nlayers * p.cpl, W4P_THREADS, smem, st, p);
    } else {
        pctts_launch(conv2d_wgrad44_kernel, p.nlayers * p.cpl, W4_THREADS, smem, st, p);
    }
    PCTTS_LAUNCH_CHECK();
    pctts_launch(wgrad_multi_reduce_kernel, dim3(pctts_div_up(W4_NP, 8), p.nlayers), 256, 0, st, p);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

}  // namespace

int pctts_conv2d_wgrad44_supported(int F, int CIN, int COUT, int KH, int KW, int pad_t, int pad_f, const void* x, const void* dy,
                                   long xpitch, long dypitch, long xbatch, long dybatch) {
    if (!(CIN == 4 && COUT == 4 && KH == 5 && KW == 5 && pad_t == 2 && pad_f == 2)) return 0;
    if ((((uintptr_t)x | (uintptr_t)dy) & 15) || ((xpitch | dypitch | xbatch | dybatch) & 3)) return 0;
    return w4_smem(F) <= 200 * 1024 ? 1 : 0;
}

int pctts_conv2d_wgrad44(const float* x, const float* dy, float* dw, float* dbias, int nb_bias, float* workspace, int NB, int T,
                         int F, long xpitch, long dypitch, long xbatch, long dybatch, float scale, int accumulate, cudaStream_t st) {
    W4Params p;
    p.L[0].x = x, p.L[0].dy = dy, p.L[0].dw = dw, p.L[0].dbias = dbias, p.L[0].cin = 4;
    p.L[0].xpitch = xpitch, p.L[0].dypitch = dypitch, p.L[0].xbatch = xbatch, p.L[0].dybatch = dybatch;
    p.partial = workspace, p.nlayers = 1;
    p.NB = NB, p.T = T, p.F = F, p.nb_bias = dbias ? nb_bias : 0;
    p.scale = scale, p.accumulate = accumulate;
    return w4_launch(p, st);
}

/* floats of workspace for pctts_conv2d_wgrad_multi (host call) */
extern "C" long pctts_conv2d_wgrad_multi_workspace(int nlayers) {
    if (nlayers < 1 || nlayers > W4_MAXL) return 0;
    return (long)nlayers * (148 / nlayers) * W4_NP;
}

extern "C" int pctts_conv2d_wgrad_multi(const pctts_wgrad_layer* layers, int nlayers, int nb_bias, float* workspace, int NB, int T,
                                        int F, float scale, int accumulate, void* stream) {
    if (!layers || !workspace || nlayers < 1 || nlayers > W4_MAXL || NB <= 0 || T <= 0 || F <= 0 || nb_bias < 0 || nb_bias > NB)
        return PCTTS_EINVAL;
    if (w4_smem(F) > 200 * 1024) return PCTTS_EUNSUPPORTED;
    W4Params p;
    bool any_bias = false;
    for (int l = 0; l < nlayers; ++l) {
        const pctts_wgrad_layer& s = layers[l];
        if (!s.x || !s.dy || !s.dw) return PCTTS_EINVAL;
        if (s.cin != 1 && s.cin != 4) return PCTTS_EUNSUPPORTED;
        if ((((uintptr_t)s.dy) & 15) || ((s.dypitch | s.dybatch) & 3)) return PCTTS_EUNSUPPORTED;
        if (s.cin == 4 && ((((uintptr_t)s.x) & 15) || ((s.xpitch | s.xbatch) & 3))) return PCTTS_EUNSUPPORTED;
        W4Layer& d = p.L[l];
        d.x = s.x, d.dy = s.dy, d.dw = s.dw, d.dbias = s.dbias, d.cin = s.cin;
        d.xpitch = s.xpitch, d.dypitch = s.dypitch, d.xbatch = s.xbatch, d.dybatch = s.dybatch;
        any_bias = any_bias || s.dbias;
    }
    p.partial = workspace, p.nlayers = nlayers;
    p.NB = NB, p.T = T, p.F = F, p.nb_bias = any_bias ? nb_bias : 0;
    p.scale = scale, p.accumulate = accumulate;
    return w4_launch(p, (cudaStream_t)stream);
}
