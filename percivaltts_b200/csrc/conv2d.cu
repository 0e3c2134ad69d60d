// Small-channel NHWC 2-D convolution over (time x frequency) maps: forward, data-gradient (forward with
// flipped kernels) and weight-gradient.  Reference ops: networktts.py:122-126 (pCNN2D),
// modeltts_common.py:98,100 and networks_critic.py:67 (Conv2D 'same', stride 1, HWIO kernels).
//
// Per layer the maps are [N, T, F, C<=4]: ~25 FLOP/B, i.e. bandwidth/L2-bound, so the kernels stage a
// halo tile in shared memory once and keep every tap in registers (DESIGN.md "conv2d").
#include "common.cuh"
#include "pctts.h"

namespace {

constexpr int TT = 16;        // output rows (time) per CTA
constexpr int TP = 4;         // output rows per thread
constexpr int FT_MAX = 136;   // output columns (freq) per CTA
constexpr int NTHREADS = 192;

template <int CIN, int COUT, int KH, int KW>
__global__ void __launch_bounds__(NTHREADS)
conv2d_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                  const float* aux, float* y, int T, int F, int pad_t, int pad_f,
                  long xpitch, long ypitch, long auxpitch, long xbatch, long ybatch, long auxbatch, int flags) {
    pdl_entry();
    extern __shared__ __align__(16) float smem[];
    constexpr int NW = KH * KW * CIN * COUT;
    float* ws = smem;                                   // [KH][KW][CIN][COUT]
    float* xs = smem + ((NW + 3) & ~3);                 // [TT+KH-1][FW][CIN]
    const int t0 = blockIdx.x * TT;
    const int n = blockIdx.y;
    const int f0 = blockIdx.z * FT_MAX;
    const int fcur = min(FT_MAX, F - f0);
    const int FW = fcur + KW - 1;
    const int tid = threadIdx.x;

    for (int i = tid; i < NW; i += NTHREADS) ws[i] = __ldg(w + i);
    const float* xb = x + (long)n * xbatch;
    const int nin = (TT + KH - 1) * FW;
    for (int i = tid; i < nin; i += NTHREADS) {
        const int r = i / FW, c = i - r * FW;
        const int t = t0 + r - pad_t, f = f0 + c - pad_f;
        const bool ok = (t >= 0 && t < T && f >= 0 && f < F);
        const float* src = xb + (long)t * xpitch + (long)f * CIN;
#pragma unroll
        for (int ci = 0; ci < CIN; ++ci) xs[i * CIN + ci] = ok ? __ldg(src + ci) : 0.f;
    }
    __syncthreads();

    const int ngroups = (TT / TP) * fcur;
    for (int g = tid; g < ngroups; g += NTHREADS) {
        const int tq = g / fcur, fl = g - tq * fcur;
        float acc[TP][COUT];
#pragma unroll
        for (int p = 0; p < TP; ++p)
#pragma unroll
            for (int co = 0; co < COUT; ++co) acc[p][co] = (flags & PCTTS_EPI_BIAS) ? __ldg(bias + co) : 0.f;
#pragma unroll
        for (int kf = 0; kf < KW; ++kf) {
            float in[TP + KH - 1][CIN];
#pragma unroll
            for (int r = 0; r < TP + KH - 1; ++r)
#pragma unroll
                for (int ci = 0; ci < CIN; ++ci) in[r][ci] = xs[((tq * TP + r) * FW + fl + kf) * CIN + ci];
#pragma unroll
            for (int kt = 0; kt < KH; ++kt) {
                float wr[CIN][COUT];
#pragma unroll
                for (int ci = 0; ci < CIN; ++ci)
#pragma unroll
                    for (int co = 0; co < COUT; ++co) wr[ci][co] = ws[((kt * KW + kf) * CIN + ci) * COUT + co];
#pragma unroll
                for (int p = 0; p < TP; ++p)
#pragma unroll
                    for (int ci = 0; ci < CIN; ++ci)
#pragma unroll
                        for (int co = 0; co < COUT; ++co) acc[p][co] = fmaf(in[p + kt][ci], wr[ci][co], acc[p][co]);
            }
        }
#pragma unroll
        for (int p = 0; p < TP; ++p) {
            const int t = t0 + tq * TP + p;
            if (t >= T) continue;
            const long off = (long)t * ypitch + (long)(f0 + fl) * COUT;
            float* dst = y + (long)n * ybatch + off;
#pragma unroll
            for (int co = 0; co < COUT; ++co) {
                float v = acc[p][co];
                if (flags & PCTTS_EPI_LRELU) v = lrelu_f(v);
                if (flags & PCTTS_EPI_DLRELU_AUX)
                    v *= dlrelu_from_out(aux[(long)n * auxbatch + (long)t * auxpitch + (long)(f0 + fl) * COUT + co]);  // aux may alias y (tangent sweep)
                if (flags & PCTTS_EPI_ACCUM) v += dst[co];
                dst[co] = v;
            }
        }
    }
}

// generic (any CIN/COUT/KH/KW) fallback: one thread per output element, runtime loops
__global__ void conv2d_fwd_generic(const float* __restrict__ x, const float* __restrict__ w,
                                   const float* __restrict__ bias, const float* aux,
                                   float* y, int NB, int T, int F, int CIN, int COUT, int KH, int KW,
                                   int pad_t, int pad_f, long xpitch, long ypitch, long auxpitch, long xbatch,
                                   long ybatch, long auxbatch, int flags) {
    pdl_entry();
    const long total = (long)NB * T * F * COUT;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int co = (int)(i % COUT);
        long r = i / COUT;
        const int f = (int)(r % F); r /= F;
        const int t = (int)(r % T);
        const int n = (int)(r / T);
        float acc = (flags & PCTTS_EPI_BIAS) ? bias[co] : 0.f;
        for (int kt = 0; kt < KH; ++kt) {
            const int ti = t + kt - pad_t;
            if (ti < 0 || ti >= T) continue;
            for (int kf = 0; kf < KW; ++kf) {
                const int fi = f + kf - pad_f;
                if (fi < 0 || fi >= F) continue;
                const float* src = x + (long)n * xbatch + (long)ti * xpitch + (long)fi * CIN;
                const float* wp = w + ((long)(kt * KW + kf) * CIN) * COUT + co;
                for (int ci = 0; ci < CIN; ++ci) acc = fmaf(src[ci], wp[(long)ci * COUT], acc);
            }
        }
        if (flags & PCTTS_EPI_LRELU) acc = lrelu_f(acc);
        const long off = (long)t * ypitch + (long)f * COUT + co;
        if (flags & PCTTS_EPI_DLRELU_AUX) acc *= dlrelu_from_out(aux[(long)n * auxbatch + (long)t * auxpitch + (long)f * COUT + co]);
        float* dst = y + (long)n * ybatch + off;
        if (flags & PCTTS_EPI_ACCUM) acc += *dst;
        *dst = acc;
    }
}

__global__ void conv2d_flip_kernel(const float* __restrict__ w, float* __restrict__ wt, int KH, int KW, int CIN, int COUT) {
    pdl_entry();
    const int total = KH * KW * CIN * COUT;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        // i indexes wt[kt][kf][co][ci]
        const int ci = i % CIN;
        int r = i / CIN;
        const int co = r % COUT; r /= COUT;
        const int kf = r % KW;
        const int kt = r / KW;
        wt[i] = w[(((KH - 1 - kt) * KW + (KW - 1 - kf)) * CIN + ci) * COUT + co];
    }
}

// ---- weight gradient:  dW[kt,kf,ci,co] = sum_{n,t,f} x[n, t+kt-pad_t, f+kf-pad_f, ci] * dy[n,t,f,co] ----------
constexpr int WG_TT = 8;
constexpr int WG_PS = 2;

template <int CIN, int COUT, int KH, int KW>
__global__ void __launch_bounds__(KH * WG_PS * 32)
conv2d_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ partial, int T, int F,
                    int pad_t, int pad_f, long xpitch, long dypitch, long xbatch, long dybatch) {
    pdl_entry();
    extern __shared__ __align__(16) float smem[];
    constexpr int NW = KH * KW * CIN * COUT;
    constexpr int NTH = KH * WG_PS * 32;
    const int FW = F + KW - 1;
    float* xs = smem;                                           // [WG_TT+KH-1][FW][CIN]
    float* dys = xs + (((WG_TT + KH - 1) * FW * CIN + 3) & ~3); // [WG_TT][F][COUT]
    float* red = dys + ((WG_TT * F * COUT + 3) & ~3);           // [WG_PS][NW]
    const int t0 = blockIdx.x * WG_TT, n = blockIdx.y, tid = threadIdx.x;
    const float* xb = x + (long)n * xbatch;
    const float* db = dy + (long)n * dybatch;
    const int nin = (WG_TT + KH - 1) * FW;
    for (int i = tid; i < nin; i += NTH) {
        const int r = i / FW, c = i - r * FW;
        const int t = t0 + r - pad_t, f = c - pad_f;
        const bool ok = (t >= 0 && t < T && f >= 0 && f < F);
#pragma unroll
        for (int ci = 0; ci < CIN; ++ci) xs[i * CIN + ci] = ok ? __ldg(xb + (long)t * xpitch + (long)f * CIN + ci) : 0.f;
    }
    const int nout = WG_TT * F;
    for (int i = tid; i < nout; i += NTH) {
        const int r = i / F, f = i - r * F;
        const int t = t0 + r;
#pragma unroll
        for (int co = 0; co < COUT; ++co) dys[i * COUT + co] = (t < T) ? __ldg(db + (long)t * dypitch + (long)f * COUT + co) : 0.f;
    }
    __syncthreads();
    const int warp = tid >> 5, lane = tid & 31;
    const int kt = warp % KH, ps = warp / KH;
    float acc[KW][CIN][COUT];
#pragma unroll
    for (int a = 0; a < KW; ++a)
#pragma unroll
        for (int b = 0; b < CIN; ++b)
#pragma unroll
            for (int c = 0; c < COUT; ++c) acc[a][b][c] = 0.f;
    for (int p = ps * 32 + lane; p < nout; p += 32 * WG_PS) {
        const int r = p / F, f = p - r * F;
        float dv[COUT];
#pragma unroll
        for (int co = 0; co < COUT; ++co) dv[co] = dys[p * COUT + co];
#pragma unroll
        for (int kf = 0; kf < KW; ++kf) {
#pragma unroll
            for (int ci = 0; ci < CIN; ++ci) {
                const float xv = xs[((r + kt) * FW + f + kf) * CIN + ci];
#pragma unroll
                for (int co = 0; co < COUT; ++co) acc[kf][ci][co] = fmaf(xv, dv[co], acc[kf][ci][co]);
            }
        }
    }
#pragma unroll
    for (int kf = 0; kf < KW; ++kf)
#pragma unroll
        for (int ci = 0; ci < CIN; ++ci)
#pragma unroll
            for (int co = 0; co < COUT; ++co) {
                const float v = warp_sum(acc[kf][ci][co]);
                if (lane == 0) red[ps * NW + ((kt * KW + kf) * CIN + ci) * COUT + co] = v;
            }
    __syncthreads();
    const long blk = (long)blockIdx.y * gridDim.x + blockIdx.x;
    for (int i = tid; i < NW; i += NTH) {
        float v = 0.f;
#pragma unroll
        for (int q = 0; q < WG_PS; ++q) v += red[q * NW + i];
        partial[blk * NW + i] = v;
    }
}

__global__ void conv2d_wgrad_generic(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ partial,
                                     int T, int F, int CIN, int COUT, int KH, int KW, int pad_t, int pad_f, long xpitch,
                                     long dypitch, long xbatch, long dybatch) {
    pdl_entry();
    // one CTA per (time-slab, n); thread loops over weight elements; slow, exact, for unusual shapes only
    const int NW = KH * KW * CIN * COUT;
    const int t0 = blockIdx.x * WG_TT, n = blockIdx.y;
    const long blk = (long)blockIdx.y * gridDim.x + blockIdx.x;
    for (int i = threadIdx.x; i < NW; i += blockDim.x) {
        const int co = i % COUT;
        int r = i / COUT;
        const int ci = r % CIN; r /= CIN;
        const int kf = r % KW;
        const int kt = r / KW;
        float acc = 0.f;
        for (int t = t0; t < min(T, t0 + WG_TT); ++t) {
            const int ti = t + kt - pad_t;
            if (ti < 0 || ti >= T) continue;
            for (int f = 0; f < F; ++f) {
                const int fi = f + kf - pad_f;
                if (fi < 0 || fi >= F) continue;
                acc = fmaf(x[(long)n * xbatch + (long)ti * xpitch + (long)fi * CIN + ci],
                           dy[(long)n * dybatch + (long)t * dypitch + (long)f * COUT + co], acc);
            }
        }
        partial[blk * NW + i] = acc;
    }
}

// one warp per weight element: lanes stride over the per-CTA partials, butterfly sum in a fixed order
__global__ void reduce_partials_kernel(const float* __restrict__ partial, float* __restrict__ out, int nblk, int NW,
                                       float scale, int accumulate) {
    pdl_entry();
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (i >= NW) return;
    float s0 = 0.f, s1 = 0.f;
    int b = lane;
    for (; b + 32 < nblk; b += 64) {
        s0 += partial[(long)b * NW + i];
        s1 += partial[(long)(b + 32) * NW + i];
    }
    if (b < nblk) s0 += partial[(long)b * NW + i];
    const float v = warp_sum(s0 + s1) * scale;
    if (lane == 0) out[i] = accumulate ? out[i] + v : v;
}

template <int CIN, int COUT, int KH, int KW>
int launch_fwd(const float* x, const float* w, const float* bias, const float* aux, float* y, int NB, int T, int F,
               int pad_t, int pad_f, long xpitch, long ypitch, long auxpitch, long xbatch, long ybatch, long auxbatch,
               int flags, cudaStream_t st) {
    const int fcur = F < FT_MAX ? F : FT_MAX;
    const size_t smem = (size_t)(((KH * KW * CIN * COUT + 3) & ~3) + (TT + KH - 1) * (fcur + KW - 1) * CIN) * sizeof(float);
    auto kern = conv2d_fwd_kernel<CIN, COUT, KH, KW>;
    if (smem > 48 * 1024) PCTTS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid(pctts_div_up(T, TT), NB, pctts_div_up(F, FT_MAX));
    pctts_launch(kern, grid, NTHREADS, smem, st, x, w, bias, aux, y, T, F, pad_t, pad_f, xpitch, ypitch, auxpitch, xbatch, ybatch, auxbatch, flags);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

template <int CIN, int COUT, int KH, int KW>
int launch_wgrad(const float* x, const float* dy, float* partial, int NB, int T, int F, int pad_t, int pad_f, long xpitch,
                 long dypitch, long xbatch, long dybatch, cudaStream_t st) {
    constexpr int NW = KH * KW * CIN * COUT;
    const int FW = F + KW - 1;
    const size_t smem = (size_t)((((WG_TT + KH - 1) * FW * CIN + 3) & ~3) + ((WG_TT * F * COUT + 3) & ~3) + WG_PS * NW) * sizeof(float);
    auto kern = conv2d_wgrad_kernel<CIN, COUT, KH, KW>;
    if (smem > 200 * 1024) return PCTTS_EUNSUPPORTED;
    if (smem > 48 * 1024) PCTTS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid(pctts_div_up(T, WG_TT), NB);
    pctts_launch(kern, grid, KH * WG_PS * 32, smem, st, x, dy, partial, T, F, pad_t, pad_f, xpitch, dypitch, xbatch, dybatch);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

}  // namespace

#define PCTTS_CONV_DISPATCH(CI, CO, H, W) \
    if (CIN == CI && COUT == CO && KH == H && KW == W)

extern "C" int pctts_conv2d_fwd(const float* x, const float* w, const float* bias, const float* aux, float* y, int NB, int T,
                                int F, int CIN, int COUT, int KH, int KW, int pad_t, int pad_f, long xpitch, long ypitch,
                                long auxpitch, long xbatch, long ybatch, long auxbatch, int flags, void* stream) {
    if (!x || !w || !y || NB < 0 || T < 0 || F <= 0) return PCTTS_EINVAL;
    if ((flags & PCTTS_EPI_BIAS) && !bias) return PCTTS_EINVAL;
    if ((flags & PCTTS_EPI_DLRELU_AUX) && !aux) return PCTTS_EINVAL;
    if (NB == 0 || T == 0) return PCTTS_OK;
    cudaStream_t st = (cudaStream_t)stream;
#define FWD_ARGS x, w, bias, aux, y, NB, T, F, pad_t, pad_f, xpitch, ypitch, auxpitch, xbatch, ybatch, auxbatch, flags, st
    PCTTS_CONV_DISPATCH(1, 4, 5, 5) return launch_fwd<1, 4, 5, 5>(FWD_ARGS);
    PCTTS_CONV_DISPATCH(4, 4, 5, 5) return launch_fwd<4, 4, 5, 5>(FWD_ARGS);
    PCTTS_CONV_DISPATCH(4, 1, 5, 5) return launch_fwd<4, 1, 5, 5>(FWD_ARGS);
    PCTTS_CONV_DISPATCH(1, 2, 3, 3) return launch_fwd<1, 2, 3, 3>(FWD_ARGS);
    PCTTS_CONV_DISPATCH(2, 2, 3, 3) return launch_fwd<2, 2, 3, 3>(FWD_ARGS);
    PCTTS_CONV_DISPATCH(2, 1, 3, 3) return launch_fwd<2, 1, 3, 3>(FWD_ARGS);
#undef FWD_ARGS
    const long total = (long)NB * T * F * COUT;
    int blocks = pctts_div_up(total, 256);
    if (blocks > 148 * 16) blocks = 148 * 16;
    pctts_launch(conv2d_fwd_generic, blocks, 256, 0, st, x, w, bias, aux, y, NB, T, F, CIN, COUT, KH, KW, pad_t, pad_f, xpitch, ypitch, auxpitch, xbatch, ybatch, auxbatch, flags);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

extern "C" int pctts_conv2d_flip_weights(const float* w, float* wt, int KH, int KW, int CIN, int COUT, void* stream) {
    if (!w || !wt) return PCTTS_EINVAL;
    const int total = KH * KW * CIN * COUT;
    pctts_launch(conv2d_flip_kernel, pctts_div_up(total, 128), 128, 0, (cudaStream_t)stream, w, wt, KH, KW, CIN, COUT);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

extern "C" long pctts_conv2d_wgrad_workspace(int NB, int T, int CIN, int COUT, int KH, int KW) {
    return (long)NB * pctts_div_up(T, WG_TT) * (KH * KW * CIN * COUT + COUT);     // per-CTA partials of dW and of the bias gradient
}

// conv2d_wgrad44.cu
int pctts_conv2d_wgrad44_supported(int F, int CIN, int COUT, int KH, int KW, int pad_t, int pad_f, const void* x, const void* dy,
                                   long xpitch, long dypitch, long xbatch, long dybatch);
int pctts_conv2d_wgrad44(const float* x, const float* dy, float* dw, float* dbias, int nb_bias, float* workspace, int NB, int T,
                         int F, long xpitch, long dypitch, long xbatch, long dybatch, float scale, int accumulate, cudaStream_t st);

int pctts_reduce_partials(const float* partial, float* out, int nblk, int NW, float scale, int accumulate, cudaStream_t st) {
    pctts_launch(reduce_partials_kernel, pctts_div_up(NW, 8), 256, 0, st, partial, out, nblk, NW, scale, accumulate);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

extern "C" int pctts_conv2d_wgrad_bias(const float* x, const float* dy, float* dw, float* dbias, int nb_bias, float* workspace, int NB,
                                       int T, int F, int CIN, int COUT, int KH, int KW, int pad_t, int pad_f, long xpitch, long dypitch,
                                       long xbatch, long dybatch, float scale, int accumulate, void* stream) {
    if (!x || !dy || !dw || !workspace || NB <= 0 || T <= 0 || F <= 0 || nb_bias < 0 || nb_bias > NB) return PCTTS_EINVAL;
    if (!pctts_conv2d_wgrad44_supported(F, CIN, COUT, KH, KW, pad_t, pad_f, x, dy, xpitch, dypitch, xbatch, dybatch))
        return PCTTS_EUNSUPPORTED;
    return pctts_conv2d_wgrad44(x, dy, dw, dbias, nb_bias, workspace, NB, T, F, xpitch, dypitch, xbatch, dybatch, scale, accumulate,
                                (cudaStream_t)stream);
}

extern "C" int pctts_conv2d_wgrad(const float* x, const float* dy, float* dw, float* workspace, int NB, int T, int F, int CIN,
                                  int COUT, int KH, int KW, int pad_t, int pad_f, long xpitch, long dypitch, long xbatch,
                                  long dybatch, float scale, int accumulate, void* stream) {
    if (!x || !dy || !dw || !workspace || NB <= 0 || T <= 0 || F <= 0) return PCTTS_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    if (pctts_conv2d_wgrad44_supported(F, CIN, COUT, KH, KW, pad_t, pad_f, x, dy, xpitch, dypitch, xbatch, dybatch))
        return pctts_conv2d_wgrad44(x, dy, dw, nullptr, 0, workspace, NB, T, F, xpitch, dypitch, xbatch, dybatch, scale, accumulate, st);
    const int NW = KH * KW * CIN * COUT;
    const int nblk = NB * pctts_div_up(T, WG_TT);
    int rc = PCTTS_EUNSUPPORTED;
#define WG_ARGS x, dy, workspace, NB, T, F, pad_t, pad_f, xpitch, dypitch, xbatch, dybatch, st
    PCTTS_CONV_DISPATCH(1, 4, 5, 5) rc = launch_wgrad<1, 4, 5, 5>(WG_ARGS);
    PCTTS_CONV_DISPATCH(4, 4, 5, 5) rc = launch_wgrad<4, 4, 5, 5>(WG_ARGS);
    PCTTS_CONV_DISPATCH(4, 1, 5, 5) rc = launch_wgrad<4, 1, 5, 5>(WG_ARGS);
    PCTTS_CONV_DISPATCH(1, 2, 3, 3) rc = launch_wgrad<1, 2, 3, 3>(WG_ARGS);
    PCTTS_CONV_DISPATCH(2, 2, 3, 3) rc = launch_wgrad<2, 2, 3, 3>(WG_ARGS);
    PCTTS_CONV_DISPATCH(2, 1, 3, 3) rc = launch_wgrad<2, 1, 3, 3>(WG_ARGS);
#undef WG_ARGS
    if (rc == PCTTS_EUNSUPPORTED) {
        dim3 grid(pctts_div_up(T, WG_TT), NB);
        pctts_launch(conv2d_wgrad_generic, grid, 128, 0, st, x, dy, workspace, T, F, CIN, COUT, KH, KW, pad_t, pad_f, xpitch, dypitch, xbatch, dybatch);
        PCTTS_LAUNCH_CHECK();
        rc = PCTTS_OK;
    }
    if (rc != PCTTS_OK) return rc;
    pctts_launch(reduce_partials_kernel, pctts_div_up(NW, 8), 256, 0, st, workspace, dw, nblk, NW, scale, accumulate);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}
