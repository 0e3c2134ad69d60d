// Per-channel reductions and BatchNormalization(axis=-1) forward / backward with the LeakyReLU fused in.
// Reference: networktts.py:61,118,124 (BatchNormalization after Dense / Conv1D / Conv2D, training-mode batch
// statistics over every non-channel axis, momentum 0.99, eps 1e-3), networktts.py:62,119,125 (LeakyReLU 0.3).
//
// Data are [rows, C] with channels innermost.  Reductions are two-stage and deterministic: every CTA writes a
// double partial, the last CTA to finish (ticket counter) sums the partials in index order.
#include "common.cuh"
#include <cuda_bf16.h>
#include "pctts.h"
#include <mutex>

namespace {

constexpr int RT = 256;

enum { MODE_STATS = 0, MODE_BNBWD = 1, MODE_COLSUM = 2 };

struct RedArgs {
    const float* x;        // STATS: x; BNBWD: pre-BN x; COLSUM: dy
    const float* dy;       // BNBWD: grad wrt post-activation output
    const float* mean;     // BNBWD
    const float* invstd;   // BNBWD
    const float* gamma;    // BNBWD
    const float* beta;     // BNBWD
    long rows;
    int C;
    int lrelu;             // BNBWD: output went through LeakyReLU
};

template <int MODE>
__device__ __forceinline__ void red_terms(const RedArgs& a, long r, int c, double& q1, double& q2) {
    const float xv = a.x[r * a.C + c];
    if (MODE == MODE_STATS) {
        q1 += xv;
        q2 += (double)xv * xv;
    } else if (MODE == MODE_COLSUM) {
        q1 += xv;
    } else {
        const float xh = (xv - a.mean[c]) * a.invstd[c];
        float dz = a.dy[r * a.C + c];
        if (a.lrelu) {
            const float z = a.gamma[c] * xh + a.beta[c];
            dz *= (z > 0.f ? 1.f : PCTTS_LRELU_ALPHA);
        }
        q1 += dz;
        q2 += (double)dz * xh;
    }
}

// Stage 1: CTA b reduces its contiguous slab of rows and writes partial[b][2][C] (doubles).
template <int MODE>
__global__ void __launch_bounds__(RT) colreduce_kernel(RedArgs a, double* __restrict__ partial) {
    pdl_entry();
    __shared__ double s1[RT], s2[RT];
    const int C = a.C, tid = threadIdx.x;
    const int rpb = C >= RT ? 1 : RT / C;          // rows handled concurrently by one CTA
    const int cpt = C > RT ? (C + RT - 1) / RT : 1; // channels per thread (C > 256)
    const long rows_per_cta = (a.rows + gridDim.x - 1) / gridDim.x;
    const long r_begin = (long)blockIdx.x * rows_per_cta;
    const long r_end = min(a.rows, r_begin + rows_per_cta);
    for (int cc = 0; cc < cpt; ++cc) {
        double q1 = 0.0, q2 = 0.0;
        const int c = (C >= RT) ? tid + cc * RT : tid % C;
        const int rl = (C >= RT) ? 0 : tid / C;
        if (c < C && rl < rpb) {
            // four independent chains: the loop is bound by the latency of its 4-byte loads, not by arithmetic
            double p1 = 0.0, p2 = 0.0, u1 = 0.0, u2 = 0.0, v1 = 0.0, v2 = 0.0;
            long r = r_begin + rl;
            for (; r + 3L * rpb < r_end; r += 4L * rpb) {
                red_terms<MODE>(a, r, c, q1, q2);
                red_terms<MODE>(a, r + rpb, c, p1, p2);
                red_terms<MODE>(a, r + 2L * rpb, c, u1, u2);
                red_terms<MODE>(a, r + 3L * rpb, c, v1, v2);
            }
            for (; r < r_end; r += rpb) red_terms<MODE>(a, r, c, q1, q2);
            q1 = (q1 + p1) + (u1 + v1);
            q2 = (q2 + p2) + (u2 + v2);
        }
        s1[tid] = q1;
        s2[tid] = q2;
        __syncthreads();
        if (C < RT) {
            if (tid < C) {
                for (int j = 1; j < rpb; ++j) {
                    q1 += s1[tid + j * C];
                    q2 += s2[tid + j * C];
                }
            }
        }
        if (c < C && rl == 0) {
            partial[((long)blockIdx.x * 2 + 0) * C + c] = q1;
            partial[((long)blockIdx.x * 2 + 1) * C + c] = q2;
        }
        __syncthreads();
    }
}

// Stage 1 for C = 4 (the Conv2D maps: rows = pixels, 16 bytes each): one float4 per pixel and array, four pixels in flight per
// thread, fp64 accumulators per channel, fixed-order reduction (lanes: butterfly; warps: index order).
template <int MODE>
__global__ void __launch_bounds__(RT) colreduce_c4_kernel(RedArgs a, double* __restrict__ partial) {
    pdl_entry();
    __shared__ double red[RT / 32][8];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long rows_per_cta = (a.rows + gridDim.x - 1) / gridDim.x;
    const long r_begin = (long)blockIdx.x * rows_per_cta;
    const long r_end = min(a.rows, r_begin + rows_per_cta);
    const float4* const x4 = reinterpret_cast<const float4*>(a.x);
    const float4* const d4 = reinterpret_cast<const float4*>(a.dy);
    float mean[4] = {0.f, 0.f, 0.f, 0.f}, ivs[4] = {1.f, 1.f, 1.f, 1.f}, gam[4] = {1.f, 1.f, 1.f, 1.f}, bet[4] = {0.f, 0.f, 0.f, 0.f};
    if (MODE == MODE_BNBWD) {
#pragma unroll
        for (int c = 0; c < 4; ++c) mean[c] = a.mean[c], ivs[c] = a.invstd[c], gam[c] = a.gamma[c], bet[c] = a.beta[c];
    }
    double q1[4] = {0.0, 0.0, 0.0, 0.0}, q2[4] = {0.0, 0.0, 0.0, 0.0};
    auto terms = [&](const float4& xv4, const float4& dv4) {
        const float xv[4] = {xv4.x, xv4.y, xv4.z, xv4.w}, dv[4] = {dv4.x, dv4.y, dv4.z, dv4.w};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            if (MODE == MODE_STATS) {
                q1[c] += xv[c];
                q2[c] += (double)xv[c] * xv[c];
            } else if (MODE == MODE_COLSUM) {
                q1[c] += xv[c];
            } else {
                const float xh = (xv[c] - mean[c]) * ivs[c];
                float dz = dv[c];
                if (a.lrelu) {
                    const float z = gam[c] * xh + bet[c];
                    dz *= (z > 0.f ? 1.f : PCTTS_LRELU_ALPHA);
                }
                q1[c] += dz;
                q2[c] += (double)dz * xh;
            }
        }
    };
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    long r = r_begin + tid;
    for (; r + 3L * RT < r_end; r += 4L * RT) {
        float4 xv[4], dv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            xv[u] = x4[r + (long)u * RT];
            dv[u] = MODE == MODE_BNBWD ? d4[r + (long)u * RT] : zero4;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) terms(xv[u], dv[u]);
    }
    for (; r < r_end; r += RT) terms(x4[r], MODE == MODE_BNBWD ? d4[r] : zero4);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const double s1 = warp_sum_d(q1[c]), s2 = warp_sum_d(q2[c]);
        if (lane == 0) red[warp][c] = s1, red[warp][4 + c] = s2;
    }
    __syncthreads();
    if (tid < 8) {
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < RT / 32; ++w) t += red[w][tid];
        partial[(long)blockIdx.x * 8 + tid] = t;       // [2][4]: tid = moment * 4 + channel
    }
}

// Stage 2: one warp per output element e in [0, 2C): lanes stride over the CTA partials, butterfly sum (fixed order).
__global__ void __launch_bounds__(RT) colreduce_final_kernel(const double* __restrict__ partial, double* __restrict__ sums,
                                                             int nblk, int C) {
    pdl_entry();
    const int e = blockIdx.x * (RT / 32) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (e >= 2 * C) return;
    double q = 0.0;
    for (int b = lane; b < nblk; b += 32) q += partial[(long)b * 2 * C + e];
    q = warp_sum_d(q);
    if (lane == 0) sums[e] = q;
}

// Stage 2 of a plain column sum with the hand-over to the gradient buffer in the same launch (bias gradients): the order of
// colreduce_final_kernel, then out[c] (+)= (float)sum -- what colreduce_final_kernel + sums_to_grads_kernel did in two launches.
__global__ void __launch_bounds__(RT) colsum_final_kernel(const double* __restrict__ partial, double* __restrict__ sums, int nblk, int C,
                                                          float* __restrict__ out, int accumulate) {
    pdl_entry();
    const int c = blockIdx.x * (RT / 32) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (c >= C) return;
    double q = 0.0;
    for (int b = lane; b < nblk; b += 32) q += partial[(long)b * 2 * C + c];
    q = warp_sum_d(q);
    if (lane == 0) {
        sums[c] = q;
        out[c] = (accumulate ? out[c] : 0.f) + (float)(q * 1.0);
    }
}

// STATS finalize: mean / invstd from (possibly all-reduced) sums; optional moving-average update.
__global__ void bn_finalize_kernel(const double* __restrict__ sums, double count, int C, float eps, float momentum,
                                   float* __restrict__ mean, float* __restrict__ invstd, float* __restrict__ moving_mean,
                                   float* __restrict__ moving_var) {
    pdl_entry();
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const double m = sums[c] / count;
    double v = sums[C + c] / count - m * m;
    if (v < 0.0) v = 0.0;
    mean[c] = (float)m;
    invstd[c] = (float)(1.0 / sqrt(v + (double)eps));
    if (moving_mean) {
        moving_mean[c] = momentum * moving_mean[c] + (1.f - momentum) * (float)m;
        moving_var[c] = momentum * moving_var[c] + (1.f - momentum) * (float)v;
    }
}

// Stage 2 of the statistics and the finalize step in one launch (single replica: nothing to all-reduce in between).  One warp
// per channel sums both moments over the CTA partials exactly as colreduce_final_kernel does (lanes stride, butterfly), so the
// result is bit-identical to the two-launch path; sums[] is still written for readers of the workspace.
__global__ void __launch_bounds__(RT) bn_final_finalize_kernel(const double* __restrict__ partial, double* __restrict__ sums,
                                                               int nblk, double count, int C, float eps, float momentum,
                                                               float* __restrict__ mean, float* __restrict__ invstd,
                                                               float* __restrict__ moving_mean, float* __restrict__ moving_var) {
    pdl_entry();
    const int c = blockIdx.x * (RT / 32) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (c >= C) return;
    double q1 = 0.0, q2 = 0.0;
    for (int b = lane; b < nblk; b += 32) {
        q1 += partial[(long)b * 2 * C + c];
        q2 += partial[(long)b * 2 * C + C + c];
    }
    q1 = warp_sum_d(q1);
    q2 = warp_sum_d(q2);
    if (lane == 0) {
        sums[c] = q1, sums[C + c] = q2;
        const double m = q1 / count;
        double v = q2 / count - m * m;
        if (v < 0.0) v = 0.0;
        mean[c] = (float)m;
        invstd[c] = (float)(1.0 / sqrt(v + (double)eps));
        if (moving_mean) {
            moving_mean[c] = momentum * moving_mean[c] + (1.f - momentum) * (float)m;
            moving_var[c] = momentum * moving_var[c] + (1.f - momentum) * (float)v;
        }
    }
}

__global__ void bn_apply_kernel(const float* __restrict__ x, float* __restrict__ y, const float* __restrict__ mean,
                                const float* __restrict__ invstd_or_var, const float* __restrict__ gamma,
                                const float* __restrict__ beta, long total, int C, float eps, int use_var, int lrelu) {
    pdl_entry();
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        const float is = use_var ? rsqrtf(invstd_or_var[c] + eps) : invstd_or_var[c];
        float v = (x[i] - mean[c]) * is * gamma[c] + beta[c];
        if (lrelu) v = lrelu_f(v);
        y[i] = v;
    }
}

// The same for C % 4 == 0 and 16-byte aligned maps: one float4 per thread and iteration (the scalar loop above keeps ~1.2 MB
// in flight on the whole GPU and runs at load latency).  Per-element expression unchanged: bit-identical output.
constexpr int NORM_SMEM_C = 1024;

// bf16 operand planes of four consecutive values (the format and the rounding steps of pctts_split_bf16, gemm_tc.cu: bit-identical
// planes), written by the kernel that produces the values when their consumer is a tensor-core GEMM
struct PlaneOut {
    __nv_bfloat16* dst;      // NULL: none
    long ld, part_stride;    // elements
    int parts;
};
__device__ __forceinline__ void emit_planes4(const PlaneOut& po, long row, int c, float (&v)[4]) {
    __nv_bfloat16* d = po.dst + row * po.ld + c;
    for (int pl = 0; pl < po.parts; ++pl) {
        uint32_t w[2];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const __nv_bfloat16 ha = __float2bfloat16_rn(v[2 * j]), hb = __float2bfloat16_rn(v[2 * j + 1]);
            v[2 * j] -= __bfloat162float(ha), v[2 * j + 1] -= __bfloat162float(hb);
            w[j] = (uint32_t)__bfloat16_as_ushort(ha) | ((uint32_t)__bfloat16_as_ushort(hb) << 16);
        }
        *reinterpret_cast<uint2*>(d + (long)pl * po.part_stride) = make_uint2(w[0], w[1]);
    }
}

__global__ void __launch_bounds__(256) bn_apply_vec4_kernel(const float4* __restrict__ x, float4* __restrict__ y,
                                                            const float* __restrict__ mean, const float* __restrict__ invstd_or_var,
                                                            const float* __restrict__ gamma, const float* __restrict__ beta,
                                                            long total4, int C, float eps, int use_var, int lrelu, const PlaneOut po) {
    pdl_entry();
    __shared__ float s_m[NORM_SMEM_C], s_i[NORM_SMEM_C], s_g[NORM_SMEM_C], s_b[NORM_SMEM_C];
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        s_m[c] = mean[c], s_g[c] = gamma[c], s_b[c] = beta[c];
        s_i[c] = use_var ? rsqrtf(invstd_or_var[c] + eps) : invstd_or_var[c];
    }
    __syncthreads();
    const int C4 = C >> 2;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total4; i += (long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C4) * 4;
        const float4 v = x[i];
        float o[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float t = (o[j] - s_m[c + j]) * s_i[c + j] * s_g[c + j] + s_b[c + j];
            if (lrelu) t = lrelu_f(t);
            o[j] = t;
        }
        y[i] = make_float4(o[0], o[1], o[2], o[3]);
        if (po.dst) emit_planes4(po, i / C4, c, o);
    }
}

__global__ void __launch_bounds__(256) bn_bwd_apply_vec4_kernel(const float4* __restrict__ x, const float4* __restrict__ dy,
                                                                float4* __restrict__ dx, const float* __restrict__ mean,
                                                                const float* __restrict__ invstd, const float* __restrict__ gamma,
                                                                const float* __restrict__ beta, const double* __restrict__ sums,
                                                                double count, long total4, int C, int lrelu, const PlaneOut po) {
    pdl_entry();
    __shared__ float s_m[NORM_SMEM_C], s_i[NORM_SMEM_C], s_g[NORM_SMEM_C], s_b[NORM_SMEM_C], s_1[NORM_SMEM_C], s_2[NORM_SMEM_C];
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        s_m[c] = mean[c], s_i[c] = invstd[c], s_g[c] = gamma[c], s_b[c] = beta[c];
        s_1[c] = (float)(sums[c] / count), s_2[c] = (float)(sums[C + c] / count);     // once per CTA, not two fp64 divisions per element
    }
    __syncthreads();
    const int C4 = C >> 2;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total4; i += (long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C4) * 4;
        const float4 xv4 = x[i], dv4 = dy[i];
        const float xv[4] = {xv4.x, xv4.y, xv4.z, xv4.w}, dv[4] = {dv4.x, dv4.y, dv4.z, dv4.w};
        float o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float xh = (xv[j] - s_m[c + j]) * s_i[c + j];
            float dz = dv[j];
            if (lrelu) {
                const float z = s_g[c + j] * xh + s_b[c + j];
                dz *= (z > 0.f ? 1.f : PCTTS_LRELU_ALPHA);
            }
            o[j] = s_g[c + j] * s_i[c + j] * (dz - s_1[c + j] - xh * s_2[c + j]);
        }
        dx[i] = make_float4(o[0], o[1], o[2], o[3]);
        if (po.dst) emit_planes4(po, i / C4, c, o);
    }
}

// dx = gamma*invstd*(dz - mean(dz) - xhat*mean(dz*xhat)); dgamma = sum(dz*xhat); dbeta = sum(dz)
__global__ void bn_bwd_apply_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ dx,
                                    const float* __restrict__ mean, const float* __restrict__ invstd,
                                    const float* __restrict__ gamma, const float* __restrict__ beta,
                                    const double* __restrict__ sums, double count, long total, int C, int lrelu) {
    pdl_entry();
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        const float xh = (x[i] - mean[c]) * invstd[c];
        float dz = dy[i];
        if (lrelu) {
            const float z = gamma[c] * xh + beta[c];
            dz *= (z > 0.f ? 1.f : PCTTS_LRELU_ALPHA);
        }
        const float m1 = (float)(sums[c] / count), m2 = (float)(sums[C + c] / count);
        dx[i] = gamma[c] * invstd[c] * (dz - m1 - xh * m2);
    }
}

__global__ void sums_to_grads_kernel(const double* __restrict__ sums, int C, float* __restrict__ g1, float* __restrict__ g2,
                                     int accumulate, double scale) {
    pdl_entry();
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    if (g1) g1[c] = (accumulate ? g1[c] : 0.f) + (float)(sums[c] * scale);
    if (g2) g2[c] = (accumulate ? g2[c] : 0.f) + (float)(sums[C + c] * scale);
}

bool norm_vec() {      // PCTTS_NORM_VEC=0: the scalar kernels only (A/B measurements)
    static const bool on = !(getenv("PCTTS_NORM_VEC") && atoi(getenv("PCTTS_NORM_VEC")) == 0);
    return on;
}

int reduce_grid(long rows, int C) {
    const int rpb = C >= RT ? 1 : RT / C;
    long want = (rows + (long)rpb * 8 - 1) / ((long)rpb * 8);   // >= 8 row-iterations per CTA
    if (want < 1) want = 1;
    if (want > 148 * 2) want = 148 * 2;
    return (int)want;
}

}  // namespace

extern "C" long pctts_colreduce_workspace_bytes(long rows, int C) {
    // partial [grid][2][C] doubles + sums [2][C] doubles + counter (padded to 16 bytes)
    return ((long)reduce_grid(rows, C) * 2 * C + 2 * C) * (long)sizeof(double) + 16;
}

namespace {
struct Ws {
    double* partial;
    double* sums;
    unsigned int* counter;
};
Ws carve(void* workspace, long rows, int C) {
    Ws w;
    w.partial = (double*)workspace;
    w.sums = w.partial + (long)reduce_grid(rows, C) * 2 * C;
    w.counter = (unsigned int*)(w.sums + 2 * C);
    return w;
}
#define PCTTS_CUDA_RC(call)            \
    do {                               \
        int rc__ = (call);             \
        if (rc__ != PCTTS_OK) return rc__; \
    } while (0)

template <int MODE>
int run_colreduce(const RedArgs& a, const Ws& w, cudaStream_t st, bool final_stage = true) {
    const int grid = reduce_grid(a.rows, a.C);
    if (a.C == 4 && norm_vec() && !(((uintptr_t)a.x | (uintptr_t)a.dy) & 15))
        pctts_launch(colreduce_c4_kernel<MODE>, grid, RT, 0, st, a, w.partial);
    else
        pctts_launch(colreduce_kernel<MODE>, grid, RT, 0, st, a, w.partial);
    PCTTS_LAUNCH_CHECK();
    if (!final_stage) return PCTTS_OK;
    pctts_launch(colreduce_final_kernel, pctts_div_up(2 * a.C, RT / 32), RT, 0, st, w.partial, w.sums, grid, a.C);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}
}  // namespace

// Column sums of x and x^2 -> workspace sums (double [2][C]); the counter inside the workspace must start at 0
// (zero the workspace once at allocation time; the kernel re-arms it).
extern "C" int pctts_bn_stats(const float* x, long rows, int C, void* workspace, void* stream) {
    if (!x || !workspace || rows <= 0 || C <= 0) return PCTTS_EINVAL;
    Ws w = carve(workspace, rows, C);
    RedArgs a{x, nullptr, nullptr, nullptr, nullptr, nullptr, rows, C, 0};
    PCTTS_CUDA_RC(run_colreduce<MODE_STATS>(a, w, (cudaStream_t)stream));
    return PCTTS_OK;
}

// First stage only: nblk = pctts_colreduce_nblk(rows, C) partials [2][C] at the base of the workspace, to be consumed by
// pctts_bn_finalize_partials (single replica: statistics second stage + finalize in one launch).
extern "C" int pctts_bn_stats_partial(const float* x, long rows, int C, void* workspace, void* stream) {
    if (!x || !workspace || rows <= 0 || C <= 0) return PCTTS_EINVAL;
    Ws w = carve(workspace, rows, C);
    RedArgs a{x, nullptr, nullptr, nullptr, nullptr, nullptr, rows, C, 0};
    PCTTS_CUDA_RC(run_colreduce<MODE_STATS>(a, w, (cudaStream_t)stream, false));
    return PCTTS_OK;
}

extern "C" int pctts_colreduce_nblk(long rows, int C) { return (rows <= 0 || C <= 0) ? 0 : reduce_grid(rows, C); }

extern "C" int pctts_bn_finalize_partials(void* workspace, long rows, int C, int nblk, double count, float eps, float momentum,
                                          float* mean, float* invstd, float* moving_mean, float* moving_var, void* stream) {
    if (!workspace || !mean || !invstd || count <= 0 || rows <= 0 || C <= 0) return PCTTS_EINVAL;
    if (nblk < 1 || nblk > reduce_grid(rows, C)) return PCTTS_EINVAL;
    Ws w = carve(workspace, rows, C);
    pctts_launch(bn_final_finalize_kernel, pctts_div_up(C, RT / 32), RT, 0, (cudaStream_t)stream, w.partial, w.sums, nblk, count, C,
                 eps, momentum, mean, invstd, moving_mean, moving_var);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

// used by producers that fuse the first reduction stage into their epilogue (conv2d_tc.cu): nblk partials [2][C] at the base
// of the workspace -> sums
int pctts_colreduce_slots(long rows, int C) { return reduce_grid(rows, C); }
int pctts_colreduce_final_launch(void* workspace, long rows, int C, int nblk, cudaStream_t st) {
    if (nblk > reduce_grid(rows, C)) return PCTTS_EINVAL;
    Ws w = carve(workspace, rows, C);
    pctts_launch(colreduce_final_kernel, pctts_div_up(2 * C, RT / 32), RT, 0, st, w.partial, w.sums, nblk, C);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

extern "C" void* pctts_colreduce_sums_ptr(void* workspace, long rows, int C) { return (void*)carve(workspace, rows, C).sums; }

extern "C" int pctts_bn_finalize(void* workspace, long rows, int C, double count, float eps, float momentum, float* mean,
                                 float* invstd, float* moving_mean, float* moving_var, void* stream) {
    if (!workspace || !mean || !invstd || count <= 0) return PCTTS_EINVAL;
    Ws w = carve(workspace, rows, C);
    pctts_launch(bn_finalize_kernel, pctts_div_up(C, 128), 128, 0, (cudaStream_t)stream, w.sums, count, C, eps, momentum, mean, invstd, moving_mean, moving_var);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

static bool planes_ok(const void* planes, long ld_dst, long part_stride, int parts, int C) {
    return planes && parts >= 1 && parts <= 3 && ld_dst >= C && (ld_dst & 3) == 0 && (part_stride & 3) == 0 && !(((uintptr_t)planes) & 7);
}

/* y = LeakyReLU?(BN(x)) and, in the same pass, the bf16 operand planes of y (what pctts_split_bf16(y) would write): planes
 * [parts][rows][ld_dst] bf16 with plane stride part_stride (elements).  C % 4 == 0, 16-byte aligned x / y only, else PCTTS_EUNSUPPORTED. */
extern "C" int pctts_bn_apply_planes(const float* x, float* y, const float* mean, const float* invstd_or_var, const float* gamma,
                                     const float* beta, long rows, int C, float eps, int use_var, int lrelu, void* planes, long ld_dst,
                                     long part_stride, int parts, void* stream) {
    if (!x || !y || !mean || !invstd_or_var || !gamma || !beta) return PCTTS_EINVAL;
    if (!planes_ok(planes, ld_dst, part_stride, parts, C)) return PCTTS_EINVAL;
    if (!norm_vec() || (C & 3) || C > NORM_SMEM_C || (((uintptr_t)x | (uintptr_t)y) & 15)) return PCTTS_EUNSUPPORTED;
    const long total = rows * C;
    if (total == 0) return PCTTS_OK;
    int blocks = pctts_div_up(total / 4, 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    PlaneOut po{(__nv_bfloat16*)planes, ld_dst, part_stride, parts};
    pctts_launch(bn_apply_vec4_kernel, blocks, 256, 0, (cudaStream_t)stream, (const float4*)x, (float4*)y, mean, invstd_or_var, gamma,
                 beta, total / 4, C, eps, use_var, lrelu, po);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

extern "C" int pctts_bn_apply(const float* x, float* y, const float* mean, const float* invstd_or_var, const float* gamma,
                              const float* beta, long rows, int C, float eps, int use_var, int lrelu, void* stream) {
    if (!x || !y || !mean || !invstd_or_var || !gamma || !beta) return PCTTS_EINVAL;
    const long total = rows * C;
    if (total == 0) return PCTTS_OK;
    int blocks = pctts_div_up(total, 256 * 4);
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (norm_vec() && (C & 3) == 0 && C <= NORM_SMEM_C && !(((uintptr_t)x | (uintptr_t)y) & 15)) {
        blocks = pctts_div_up(total / 4, 256);
        if (blocks > 148 * 8) blocks = 148 * 8;
        pctts_launch(bn_apply_vec4_kernel, blocks, 256, 0, (cudaStream_t)stream, (const float4*)x, (float4*)y, mean, invstd_or_var, gamma,
                     beta, total / 4, C, eps, use_var, lrelu, PlaneOut{nullptr, 0, 0, 0});
    } else
        pctts_launch(bn_apply_kernel, blocks, 256, 0, (cudaStream_t)stream, x, y, mean, invstd_or_var, gamma, beta, total, C, eps, use_var, lrelu);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

// Backward, phase 1: sums[0]=sum(dz), sums[1]=sum(dz*xhat) with dz = dy*lrelu'(z).
extern "C" int pctts_bn_bwd_reduce(const float* x, const float* dy, const float* mean, const float* invstd, const float* gamma,
                                   const float* beta, long rows, int C, int lrelu, void* workspace, void* stream) {
    if (!x || !dy || !mean || !invstd || !gamma || !beta || !workspace || rows <= 0) return PCTTS_EINVAL;
    Ws w = carve(workspace, rows, C);
    RedArgs a{x, dy, mean, invstd, gamma, beta, rows, C, lrelu};
    PCTTS_CUDA_RC(run_colreduce<MODE_BNBWD>(a, w, (cudaStream_t)stream));
    return PCTTS_OK;
}

// Backward, phase 1, first stage only: pctts_colreduce_nblk(rows, C) partials [2][C] at the base of the workspace, consumed by
// pctts_bn_sync_bwd (second stage + replica exchange + dgamma / dbeta in one launch).
extern "C" int pctts_bn_bwd_reduce_partial(const float* x, const float* dy, const float* mean, const float* invstd,
                                           const float* gamma, const float* beta, long rows, int C, int lrelu, void* workspace,
                                           void* stream) {
    if (!x || !dy || !mean || !invstd || !gamma || !beta || !workspace || rows <= 0) return PCTTS_EINVAL;
    Ws w = carve(workspace, rows, C);
    RedArgs a{x, dy, mean, invstd, gamma, beta, rows, C, lrelu};
    PCTTS_CUDA_RC(run_colreduce<MODE_BNBWD>(a, w, (cudaStream_t)stream, false));
    return PCTTS_OK;
}

// Backward, phase 2 (count = global row count when the sums were all-reduced across ranks).
extern "C" int pctts_bn_bwd_apply(const float* x, const float* dy, float* dx, const float* mean, const float* invstd,
                                  const float* gamma, const float* beta, long rows, int C, double count, int lrelu,
                                  float* dgamma, float* dbeta, float grad_scale, int accumulate, void* workspace,
                                  void* stream) {
    if (!x || !dy || !dx || !workspace) return PCTTS_EINVAL;
    Ws w = carve(workspace, rows, C);
    cudaStream_t st = (cudaStream_t)stream;
    const long total = rows * C;
    int blocks = pctts_div_up(total, 256 * 4);
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (norm_vec() && (C & 3) == 0 && C <= NORM_SMEM_C && !(((uintptr_t)x | (uintptr_t)dy | (uintptr_t)dx) & 15)) {
        blocks = pctts_div_up(total / 4, 256);
        if (blocks > 148 * 8) blocks = 148 * 8;
        pctts_launch(bn_bwd_apply_vec4_kernel, blocks, 256, 0, st, (const float4*)x, (const float4*)dy, (float4*)dx, mean, invstd, gamma,
                     beta, w.sums, count, total / 4, C, lrelu, PlaneOut{nullptr, 0, 0, 0});
    } else
        pctts_launch(bn_bwd_apply_kernel, blocks, 256, 0, st, x, dy, dx, mean, invstd, gamma, beta, w.sums, count, total, C, lrelu);
    PCTTS_LAUNCH_CHECK();
    if (dgamma || dbeta) {
        pctts_launch(sums_to_grads_kernel, pctts_div_up(C, 128), 128, 0, st, w.sums, C, dbeta, dgamma, accumulate, (double)grad_scale);
        PCTTS_LAUNCH_CHECK();
    }
    return PCTTS_OK;
}

/* phase 2 with the bf16 operand planes of dx written in the same pass (see pctts_bn_apply_planes); no dgamma / dbeta here */
extern "C" int pctts_bn_bwd_apply_planes(const float* x, const float* dy, float* dx, const float* mean, const float* invstd,
                                         const float* gamma, const float* beta, long rows, int C, double count, int lrelu,
                                         void* workspace, void* planes, long ld_dst, long part_stride, int parts, void* stream) {
    if (!x || !dy || !dx || !workspace) return PCTTS_EINVAL;
    if (!planes_ok(planes, ld_dst, part_stride, parts, C)) return PCTTS_EINVAL;
    if (!norm_vec() || (C & 3) || C > NORM_SMEM_C || (((uintptr_t)x | (uintptr_t)dy | (uintptr_t)dx) & 15)) return PCTTS_EUNSUPPORTED;
    Ws w = carve(workspace, rows, C);
    const long total = rows * C;
    if (total == 0) return PCTTS_OK;
    int blocks = pctts_div_up(total / 4, 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    PlaneOut po{(__nv_bfloat16*)planes, ld_dst, part_stride, parts};
    pctts_launch(bn_bwd_apply_vec4_kernel, blocks, 256, 0, (cudaStream_t)stream, (const float4*)x, (const float4*)dy, (float4*)dx, mean,
                 invstd, gamma, beta, w.sums, count, total / 4, C, lrelu, po);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

// out[c] (+)= sum_r x[r, c]   (bias gradients)
extern "C" int pctts_colsum(const float* x, long rows, int C, float* out, int accumulate, void* workspace, void* stream) {
    if (!x || !out || !workspace || rows <= 0 || C <= 0) return PCTTS_EINVAL;
    Ws w = carve(workspace, rows, C);
    cudaStream_t st = (cudaStream_t)stream;
    RedArgs a{x, nullptr, nullptr, nullptr, nullptr, nullptr, rows, C, 0};
    PCTTS_CUDA_RC(run_colreduce<MODE_COLSUM>(a, w, st, false));
    pctts_launch(colsum_final_kernel, pctts_div_up(C, RT / 32), RT, 0, st, w.partial, w.sums, reduce_grid(rows, C), C, out, accumulate);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

// ---- pctts_colsum for callers that may not allocate device memory -------------------------------------------------------
// Same two launches as pctts_colsum; the per-CTA partials and the double sums live in a fixed array that is part of the
// library image instead of in a caller-provided workspace.  See include/pctts.h for who needs this and for the rules of use.
namespace {
constexpr int STATIC_MAX_C = PCTTS_COLSUM_STATIC_MAX_C;
// the largest layout carve() can ask for: reduce_grid() never exceeds 2 * 148 CTAs, each owning [2][C] doubles
constexpr long STATIC_SCRATCH_BYTES = ((long)(148 * 2) * 2 * STATIC_MAX_C + 2 * STATIC_MAX_C) * (long)sizeof(double) + 16;
constexpr int STATIC_MAX_DEVICES = 64;
}  // namespace

// One copy per GPU that loads the library (about 4.9 MB each), alive until the process exits.
static __device__ __align__(256) unsigned char g_colsum_static_scratch[STATIC_SCRATCH_BYTES];

namespace {
// In host code `g_colsum_static_scratch` is only a handle for the runtime's symbol table, never a device address: the address a
// kernel may use is looked up per device, once, with that device current.
std::mutex g_static_mu;
void* g_static_ptr[STATIC_MAX_DEVICES] = {};

int static_scratch_for_current_device(void** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= STATIC_MAX_DEVICES) return PCTTS_EUNSUPPORTED;
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_colsum_static_scratch);
        if (e != cudaSuccess) return (int)e;
        if (!p) return (int)cudaErrorInvalidSymbol;
        g_static_ptr[dev] = p;
    }
    *out = g_static_ptr[dev];
    return PCTTS_OK;
}
}  // namespace

extern "C" int pctts_colsum_static(const float* x, long rows, int C, float* out, int accumulate, void* stream) {
    if (!x || !out || rows <= 0 || C <= 0) return PCTTS_EINVAL;
    // the fixed array has no room for wider matrices (the second test guards the layout against a future change of reduce_grid)
    if (C > STATIC_MAX_C || pctts_colreduce_workspace_bytes(rows, C) > STATIC_SCRATCH_BYTES) return PCTTS_ELIMIT;
    void* scratch = nullptr;
    PCTTS_CUDA_RC(static_scratch_for_current_device(&scratch));
    // Stage 1 overwrites every partial that stage 2 reads, so nothing left in the array by an earlier call needs clearing.
    Ws w = carve(scratch, rows, C);
    cudaStream_t st = (cudaStream_t)stream;
    RedArgs a{x, nullptr, nullptr, nullptr, nullptr, nullptr, rows, C, 0};
    PCTTS_CUDA_RC(run_colreduce<MODE_COLSUM>(a, w, st, false));
    pctts_launch(colsum_final_kernel, pctts_div_up(C, RT / 32), RT, 0, st, w.partial, w.sums, reduce_grid(rows, C), C, out, accumulate);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}
