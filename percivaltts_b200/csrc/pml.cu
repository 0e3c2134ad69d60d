// PML vocoder band compression / decompression (bandwidth-bound).
// Reference call sites: vocoders.py:134-145 (compress_spectrum: sp.linbnd2fwbnd(log|SPEC|)), vocoders.py:147-166
// (decompress_spectrum: exp(sp.fwbnd2linbnd(., smooth=True))), vocoders.py:202 (noise mask, no smoothing),
// pulsemodel/analysis.py:336,366.  The interpolation tables (index + weight per output column) are inputs, built on
// the host by percivaltts_b200/sigproc.py, so the kernels carry no frequency-warping constants themselves.
#include "common.cuh"
#include "pctts.h"
#include <stdint.h>
#include <stdlib.h>
#include "tc_common.cuh"

namespace {

constexpr int CWARPS = 8;      // compress: warps per block, one frame per warp
constexpr int CUNR = 8;        // compress: 128-byte warp loads in flight per warp
constexpr int DT = 256;        // decompress: threads per block
constexpr int DF = 16;         // decompress: frames per tile
constexpr int DFP = 20;        // decompress: padded frame pitch of the transposed tile (80 B rows: conflict-free LDS.128)

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// One warp per frame: the 2049-bin row (8196 B, only 4-byte aligned) is staged in shared memory with coalesced streaming
// loads, UNR x 128 B in flight per warp, then the K bands are gathered from it.  Measured alternatives on B200 (K=129):
// 4-byte cp.async with a double-buffered row 4.17 TB/s, UNR=16 4.35 TB/s, 4 warps x UNR=16 4.15 TB/s; this one 4.59 TB/s.
template <int UNR, int WARPS>
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
        for (; i + (UNR - 1) * 32 < nbins; i += UNR * 32) {
            float v[UNR];
#pragma unroll
            for (int u = 0; u < UNR; ++u) v[u] = __ldcs(src + i + u * 32);
#pragma unroll
            for (int u = 0; u < UNR; ++u) row[i + u * 32] = v[u];
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

// A block takes a tile of DF = 16 consecutive frames (their K-band rows are contiguous in memory) and stores it TRANSPOSED in
// shared memory, tile[k][f]: the 16 frames of one band are 64 contiguous bytes.  Every thread owns output bins i = tid,
// tid + DT, ... and keeps that bin's tap rows and interpolation coefficients in registers while it walks the frames four at a
// time: per 4 outputs 4 (2) LDS.128, 16 (4) FP32 ops, 4 MUFU and 4 coalesced streaming stores with immediate offsets -- the
// issue slots no longer limit it (the first version, 4 scalar LDS + address arithmetic per output, was issue-bound at
// 2.9 TB/s), the 8196 B written per frame do.  The coefficients depend on the bin only:
//   Catmull-Rom  c0 = (-w + 2w^2 - w^3)/2, c1 = (2 - 5w^2 + 3w^3)/2, c2 = (w + 4w^2 - 3w^3)/2, c3 = (-w^2 + w^3)/2
// and exp(v) is ex2(v * log2 e) with log2 e folded into them.  NB > 0 fixes the number of bins at compile time.
template <bool SMOOTH, bool EXP, bool FULL, int NB>
__device__ __forceinline__ void decompress_bin(const float* __restrict__ tile, int j, float w, int K, float* __restrict__ dst,
                                               int nbins_rt, int nf) {
    const long nbins = NB > 0 ? NB : nbins_rt;
    const float4* r1 = reinterpret_cast<const float4*>(tile + j * DFP);
    const float4* r2 = reinterpret_cast<const float4*>(tile + min(j + 1, K - 1) * DFP);
    auto put = [&](int f, float v) {
        if (FULL || f < nf) __stcs(dst + f * nbins, v);
    };
    if (SMOOTH) {
        const float4* r0 = reinterpret_cast<const float4*>(tile + max(j - 1, 0) * DFP);
        const float4* r3 = reinterpret_cast<const float4*>(tile + min(j + 2, K - 1) * DFP);
        const float w2 = w * w, w3 = w2 * w;
        float c0 = 0.5f * (-w + 2.f * w2 - w3), c1 = 0.5f * (2.f - 5.f * w2 + 3.f * w3);
        float c2 = 0.5f * (w + 4.f * w2 - 3.f * w3), c3 = 0.5f * (w3 - w2);
        if (EXP) {
            const float l2e = 1.4426950408889634f;
            c0 *= l2e, c1 *= l2e, c2 *= l2e, c3 *= l2e;
        }
        auto one = [&](float p0, float p1, float p2, float p3) {
            const float v = fmaf(c3, p3, fmaf(c0, p0, fmaf(c2, p2, c1 * p1)));
            return EXP ? ex2_approx(v) : v;
        };
#pragma unroll
        for (int q = 0; q < DF / 4; ++q) {
            if (!FULL && q * 4 >= nf) break;
            const float4 a0 = r0[q], a1 = r1[q], a2 = r2[q], a3 = r3[q];
            put(q * 4 + 0, one(a0.x, a1.x, a2.x, a3.x));
            put(q * 4 + 1, one(a0.y, a1.y, a2.y, a3.y));
            put(q * 4 + 2, one(a0.z, a1.z, a2.z, a3.z));
            put(q * 4 + 3, one(a0.w, a1.w, a2.w, a3.w));
        }
    } else {
        auto one = [&](float p1, float p2) {
            const float v = (w == 0.f) ? p1 : fmaf(w, p2 - p1, p1);      // exact at the knots, convex in between
            return EXP ? __expf(v) : v;
        };
#pragma unroll
        for (int q = 0; q < DF / 4; ++q) {
            if (!FULL && q * 4 >= nf) break;
            const float4 a1 = r1[q], a2 = r2[q];
            put(q * 4 + 0, one(a1.x, a2.x));
            put(q * 4 + 1, one(a1.y, a2.y));
            put(q * 4 + 2, one(a1.z, a2.z));
            put(q * 4 + 3, one(a1.w, a2.w));
        }
    }
}

template <bool SMOOTH, bool EXP, int NB>
__global__ void __launch_bounds__(DT)
pml_decompress_kernel(const float* __restrict__ in, const int* __restrict__ idx, const float* __restrict__ wgt,
                      float* __restrict__ out, long T, int K, int nbins_rt) {
    pdl_entry();
    extern __shared__ float4 tile4[];
    float* tile = reinterpret_cast<float*>(tile4);
    const int nbins = NB > 0 ? NB : nbins_rt;
    const int tid = threadIdx.x;
    const long ntiles = (T + DF - 1) / DF;
    for (long tl = blockIdx.x; tl < ntiles; tl += gridDim.x) {
        const long t0 = tl * DF;
        const int nf = (int)min((long)DF, T - t0);
        __syncthreads();
        const float* src = in + t0 * K;
        for (int e = tid; e < DF * K; e += DT) {
            const int f = e / K, k = e - f * K;
            tile[k * DFP + f] = f < nf ? __ldcs(src + e) : 0.f;
        }
        __syncthreads();
        float* dst0 = out + t0 * nbins;
        if (nf == DF) {
            for (int i = tid; i < nbins; i += DT)
                decompress_bin<SMOOTH, EXP, true, NB>(tile, __ldg(idx + i), __ldg(wgt + i), K, dst0 + i, nbins, nf);
        } else {
            for (int i = tid; i < nbins; i += DT)
                decompress_bin<SMOOTH, EXP, false, NB>(tile, __ldg(idx + i), __ldg(wgt + i), K, dst0 + i, nbins, nf);
        }
    }
}


// ---- 2049-bin fast path: full-line stores -------------------------------------------------------------------------------------
// A [T, 2049] fp32 row starts at a 4-byte-aligned address, so a warp's 128-byte store straddles two cache lines and ends in
// two partial sectors; the version above spends half of the L2 request rate on that (2.9 TB/s).  Four consecutive frames,
// however, are one contiguous 32 784-byte span whose start IS 16-byte aligned (4 x 8196 = 16 x 2049): the block computes a
// 4-frame group into a linear shared-memory stage and one thread hands it to the TMA engine as a single bulk store
// (cp.async.bulk.global.shared::cta), double-buffered so that the next group is computed while the previous one drains.
// Per-bin tap indices and coefficients live in registers for the life of the block (9 bins per thread).
constexpr int TG = 4;                      // frames per store group
constexpr int NB_FAST = 2049;
constexpr int BPT = (NB_FAST + DT - 1) / DT;   // bins per thread (9; the 9th only for tid 0)

__device__ __forceinline__ void bulk_store(float* gdst, const float* ssrc, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst),
                 "r"((uint32_t)__cvta_generic_to_shared(ssrc)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

template <bool SMOOTH, bool EXP>
__global__ void __launch_bounds__(DT)
pml_decompress_tma_kernel(const float* __restrict__ in, const int* __restrict__ idx, const float* __restrict__ wgt,
                          float* __restrict__ out, long T4, int K) {
    pdl_entry();
    extern __shared__ float4 smem4[];
    float* stage = reinterpret_cast<float*>(smem4);                  // [2][TG * 2049], 16-byte aligned
    float* tile = stage + 2 * TG * NB_FAST;                          // [K][DFP]  (2 * 4 * 2049 floats = 65 568 B: 16-byte multiple)
    const int tid = threadIdx.x;
    // per-bin constants
    int jj[BPT];
    float c[BPT][4];
#pragma unroll
    for (int m = 0; m < BPT; ++m) {
        const int i = tid + m * DT;
        jj[m] = 0;
        c[m][0] = c[m][1] = c[m][2] = c[m][3] = 0.f;
        if (i < NB_FAST) {
            const int j = __ldg(idx + i);
            const float w = __ldg(wgt + i);
            jj[m] = j;
            if (SMOOTH) {
                const float w2 = w * w, w3 = w2 * w;
                c[m][0] = 0.5f * (-w + 2.f * w2 - w3), c[m][1] = 0.5f * (2.f - 5.f * w2 + 3.f * w3);
                c[m][2] = 0.5f * (w + 4.f * w2 - 3.f * w3), c[m][3] = 0.5f * (w3 - w2);
                if (EXP) {
                    const float l2e = 1.4426950408889634f;
                    c[m][0] *= l2e, c[m][1] *= l2e, c[m][2] *= l2e, c[m][3] *= l2e;
                }
            } else {
                c[m][0] = w;
            }
        }
    }
    const long ntiles = (T4 + DF - 1) / DF;
    int buf = 0;
    for (long tl = blockIdx.x; tl < ntiles; tl += gridDim.x) {
        const long t0 = tl * DF;
        const int nf = (int)min((long)DF, T4 - t0);      // multiple of TG
        __syncthreads();                                  // the previous tile's LDS are done
        const float* src = in + t0 * K;
        for (int e = tid; e < DF * K; e += DT) {
            const int f = e / K, k = e - f * K;
            tile[k * DFP + f] = f < nf ? __ldcs(src + e) : 0.f;
        }
        for (int g = 0; g < nf / TG; ++g, buf ^= 1) {
            float* st = stage + buf * (TG * NB_FAST);
            if (tid == 0) bulk_wait_read<1>();            // the store that last read this stage buffer has drained it
            __syncthreads();                              // (also orders the tile fill before the first group)
#pragma unroll
            for (int m = 0; m < BPT; ++m) {
                const int i = tid + m * DT;
                if (m == BPT - 1 && i >= NB_FAST) break;
                const int j = jj[m];
                const float4 a1 = reinterpret_cast<const float4*>(tile + j * DFP)[g];
                const float4 a2 = reinterpret_cast<const float4*>(tile + min(j + 1, K - 1) * DFP)[g];
                float v0, v1, v2, v3;
                if (SMOOTH) {
                    const float4 a0 = reinterpret_cast<const float4*>(tile + max(j - 1, 0) * DFP)[g];
                    const float4 a3 = reinterpret_cast<const float4*>(tile + min(j + 2, K - 1) * DFP)[g];
                    v0 = fmaf(c[m][3], a3.x, fmaf(c[m][0], a0.x, fmaf(c[m][2], a2.x, c[m][1] * a1.x)));
                    v1 = fmaf(c[m][3], a3.y, fmaf(c[m][0], a0.y, fmaf(c[m][2], a2.y, c[m][1] * a1.y)));
                    v2 = fmaf(c[m][3], a3.z, fmaf(c[m][0], a0.z, fmaf(c[m][2], a2.z, c[m][1] * a1.z)));
                    v3 = fmaf(c[m][3], a3.w, fmaf(c[m][0], a0.w, fmaf(c[m][2], a2.w, c[m][1] * a1.w)));
                    if (EXP) v0 = ex2_approx(v0), v1 = ex2_approx(v1), v2 = ex2_approx(v2), v3 = ex2_approx(v3);
                } else {
                    const float w = c[m][0];
                    v0 = (w == 0.f) ? a1.x : fmaf(w, a2.x - a1.x, a1.x);
                    v1 = (w == 0.f) ? a1.y : fmaf(w, a2.y - a1.y, a1.y);
                    v2 = (w == 0.f) ? a1.z : fmaf(w, a2.z - a1.z, a1.z);
                    v3 = (w == 0.f) ? a1.w : fmaf(w, a2.w - a1.w, a1.w);
                    if (EXP) v0 = __expf(v0), v1 = __expf(v1), v2 = __expf(v2), v3 = __expf(v3);
                }
                st[0 * NB_FAST + i] = v0;
                st[1 * NB_FAST + i] = v1;
                st[2 * NB_FAST + i] = v2;
                st[3 * NB_FAST + i] = v3;
            }
            fence_async_smem();                           // generic-proxy writes -> visible to the bulk-copy engine
            __syncthreads();
            if (tid == 0) {
                bulk_store(out + (t0 + (long)g * TG) * NB_FAST, st, TG * NB_FAST * 4);
                bulk_commit();
            }
        }
    }
    if (tid == 0) bulk_wait_read<0>();                    // shared memory must outlive the reads of the last stores
}

// ---- compress fast path: full-line bulk loads ---------------------------------------------------------------------------------
// Same alignment argument as for the decompress stores: four consecutive [nbins] fp32 rows are one contiguous 16*nbins-byte
// span with a 16-byte-aligned start, which the TMA engine brings into shared memory as a single bulk copy
// (cp.async.bulk.shared::cluster.global, completion on an mbarrier) with no register staging and no split cache lines.
// A block owns a ring of CSTAGES such spans; warp w gathers the K bands of frame w of the current span.
constexpr int CSTAGES = 3;
constexpr int CG = 4;          // frames per span = warps per block

__device__ __forceinline__ void bulk_load(uint32_t sdst, const float* gsrc, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(sdst), "l"(gsrc),
                 "r"(bytes), "r"(bar)
                 : "memory");
}

__global__ void __launch_bounds__(CG * 32)
pml_compress_tma_kernel(const float* __restrict__ in, const int* __restrict__ idx, const float* __restrict__ wgt,
                        float* __restrict__ out, long ngroups, int nbins, int K, int log_abs) {
    pdl_entry();
    extern __shared__ float4 smem4[];
    __shared__ __align__(8) uint64_t full_bar[CSTAGES];
    float* stage = reinterpret_cast<float*>(smem4);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t span_bytes = (uint32_t)CG * nbins * 4u;
    const long span_floats = (long)CG * nbins;
    if (threadIdx.x == 0) {
        for (int s = 0; s < CSTAGES; ++s) tcx::mbar_init(tcx::smem_u32(&full_bar[s]), 1);
        tcx::fence_barrier_init();
    }
    __syncthreads();
    const long g0 = blockIdx.x, gstep = gridDim.x;
    auto issue = [&](long g, int s) {
        const uint32_t bar = tcx::smem_u32(&full_bar[s]);
        tcx::mbar_arrive_expect_tx(bar, span_bytes);
        bulk_load(tcx::smem_u32(stage + s * span_floats), in + g * span_floats, span_bytes, bar);
    };
    if (threadIdx.x == 0)
        for (int s = 0; s < CSTAGES; ++s)
            if (g0 + s * gstep < ngroups) issue(g0 + s * gstep, s);
    int s = 0;
    uint32_t ph = 0;
    for (long g = g0; g < ngroups; g += gstep) {
        tcx::mbar_wait(tcx::smem_u32(&full_bar[s]), ph);
        const float* row = stage + s * span_floats + (long)warp * nbins;
        float* dst = out + (g * CG + warp) * K;
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
        __syncthreads();                                   // every warp is done with stage s
        if (threadIdx.x == 0 && g + CSTAGES * gstep < ngroups) issue(g + CSTAGES * gstep, s);
        if (++s == CSTAGES) s = 0, ph ^= 1u;
    }
}

}  // namespace

extern "C" int pctts_pml_linbnd2fwbnd(const float* in, const int* idx, const float* wgt, float* out, long T, int nbins, int K,
                                      int log_abs, void* stream) {
    if (T == 0) return PCTTS_OK;
    if (!in || !idx || !wgt || !out || T < 0 || nbins <= 1 || K <= 0) return PCTTS_EINVAL;
    static const int no_tma = getenv("PCTTS_PML_NO_TMA") ? atoi(getenv("PCTTS_PML_NO_TMA")) : 0;
    const size_t span = (size_t)CG * nbins * sizeof(float);
    if (!no_tma && T >= CG && (((uintptr_t)in) & 15) == 0 && CSTAGES * span <= 220 * 1024) {
        // whole 4-frame spans through the bulk-load kernel, the (< 4 frames) tail through the generic one
        const long ngroups = T / CG;
        const size_t sm = CSTAGES * span;
        PCTTS_CUDA(cudaFuncSetAttribute(pml_compress_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
        const int per = (int)((227 * 1024) / (sm + 1024));
        long nb = ngroups < 148L * per ? ngroups : 148L * per;
        pctts_launch(pml_compress_tma_kernel, (int)nb, CG * 32, sm, (cudaStream_t)stream, in, idx, wgt, out, ngroups, nbins, K, log_abs);
        PCTTS_LAUNCH_CHECK();
        const long done = ngroups * CG;
        if (done == T) return PCTTS_OK;
        in += done * nbins, out += done * K, T -= done;
    }
    const size_t smem = (size_t)CWARPS * nbins * sizeof(float);
    if (smem > 227 * 1024) return PCTTS_EUNSUPPORTED;
    PCTTS_CUDA(cudaFuncSetAttribute(pml_compress_kernel<CUNR, CWARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int per_sm = (int)((227 * 1024) / (smem + 1024));
    long blocks = (T + CWARPS - 1) / CWARPS;
    if (blocks > 148L * per_sm) blocks = 148L * per_sm;
    pctts_launch(pml_compress_kernel<CUNR, CWARPS>, (int)blocks, CWARPS * 32, smem, (cudaStream_t)stream, in, idx, wgt, out, T, nbins,
                 K, log_abs);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

template <bool SMOOTH, bool EXP, int NB>
static int launch_decompress(const float* in, const int* idx, const float* wgt, float* out, long T, int K, int nbins, cudaStream_t st) {
    const size_t smem = (size_t)K * DFP * sizeof(float);
    if (smem > 227 * 1024) return PCTTS_EUNSUPPORTED;
    if (smem > 48 * 1024)
        PCTTS_CUDA(cudaFuncSetAttribute(pml_decompress_kernel<SMOOTH, EXP, NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long blocks = (T + DF - 1) / DF;
    if (blocks > 148 * 8) blocks = 148 * 8;
    pctts_launch(pml_decompress_kernel<SMOOTH, EXP, NB>, (int)blocks, DT, smem, st, in, idx, wgt, out, T, K, nbins);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

template <int NB>
static int dispatch_decompress(const float* in, const int* idx, const float* wgt, float* out, long T, int K, int nbins, int smooth,
                               int do_exp, cudaStream_t st) {
    if (smooth) return do_exp ? launch_decompress<true, true, NB>(in, idx, wgt, out, T, K, nbins, st)
                              : launch_decompress<true, false, NB>(in, idx, wgt, out, T, K, nbins, st);
    return do_exp ? launch_decompress<false, true, NB>(in, idx, wgt, out, T, K, nbins, st)
                  : launch_decompress<false, false, NB>(in, idx, wgt, out, T, K, nbins, st);
}

extern "C" int pctts_pml_fwbnd2linbnd(const float* in, const int* idx, const float* wgt, float* out, long T, int K, int nbins,
                                      int smooth, int do_exp, void* stream) {
    if (T == 0) return PCTTS_OK;
    if (!in || !idx || !wgt || !out || T < 0 || nbins <= 1 || K <= 1) return PCTTS_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    if (nbins == NB_FAST && (((uintptr_t)out) & 15) == 0 && T >= TG) {
        // full-line TMA stores for the whole 4-frame groups, the generic kernel for the (< 4 frames) tail
        const long T4 = T - T % TG;
        const size_t smem = (size_t)2 * TG * NB_FAST * sizeof(float) + (size_t)K * DFP * sizeof(float);
        if (smem <= 113 * 1024) {
            long blocks = (T4 + DF - 1) / DF;
            if (blocks > 148 * 2) blocks = 148 * 2;
#define PML_TMA(S, E)                                                                                                           \
    do {                                                                                                                        \
        PCTTS_CUDA(cudaFuncSetAttribute(pml_decompress_tma_kernel<S, E>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        pctts_launch(pml_decompress_tma_kernel<S, E>, (int)blocks, DT, smem, st, in, idx, wgt, out, T4, K);                     \
    } while (0)
            if (smooth && do_exp) PML_TMA(true, true);
            else if (smooth) PML_TMA(true, false);
            else if (do_exp) PML_TMA(false, true);
            else PML_TMA(false, false);
#undef PML_TMA
            PCTTS_LAUNCH_CHECK();
            if (T4 == T) return PCTTS_OK;
            return dispatch_decompress<2049>(in + T4 * K, idx, wgt, out + T4 * nbins, T - T4, K, nbins, smooth, do_exp, st);
        }
    }
    // dftlen = 4096 (2049 bins) is the reference's default (vocoders.py:172): its store offsets are compile-time immediates
    if (nbins == 2049) return dispatch_decompress<2049>(in, idx, wgt, out, T, K, nbins, smooth, do_exp, st);
    return dispatch_decompress<0>(in, idx, wgt, out, T, K, nbins, smooth, do_exp, st);
}
