// Per-channel reductions and BatchNormalization(axis=-1) forward / backward with the LeakyReLU fused in.
// Reference: networktts.py:61,118,124 (BatchNormalization after Dense / Conv1D / Conv2D, training-mode batch
// statistics over every non-channel axis, momentum 0.99, eps 1e-3), networktts.py:62,119,125 (LeakyReLU 0.3).
//
// Data are [rows, C] with channels innermost.  Reductions are two-stage and deterministic: every CTA writes a
// double partial, the last CTA to finish (ticket counter) sums the partials in index order.
#include "common.cuh"
#include "pctts.h"

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
            // two independent chains for a little more memory-level parallelism
            double p1 = 0.0, p2 = 0.0;
            long r = r_begin + rl;
            for (; r + rpb < r_end; r += 2 * rpb) {
                red_terms<MODE>(a, r, c, q1, q2);
                red_terms<MODE>(a, r + rpb, c, p1, p2);
            }
            if (r < r_end) red_terms<MODE>(a, r, c, q1, q2);
            q1 += p1;
            q2 += p2;
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

extern "C" int pctts_bn_apply(const float* x, float* y, const float* mean, const float* invstd_or_var, const float* gamma,
                              const float* beta, long rows, int C, float eps, int use_var, int lrelu, void* stream) {
    if (!x || !y || !mean || !invstd_or_var || !gamma || !beta) return PCTTS_EINVAL;
    const long total = rows * C;
    if (total == 0) return PCTTS_OK;
    int blocks = pctts_div_up(total, 256 * 4);
    if (blocks > 148 * 8) blocks = 148 * 8;
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
    pctts_launch(bn_bwd_apply_kernel, blocks, 256, 0, st, x, dy, dx, mean, invstd, gamma, beta, w.sums, count, total, C, lrelu);
    PCTTS_LAUNCH_CHECK();
    if (dgamma || dbeta) {
        pctts_launch(sums_to_grads_kernel, pctts_div_up(C, 128), 128, 0, st, w.sums, C, dbeta, dgamma, accumulate, (double)grad_scale);
        PCTTS_LAUNCH_CHECK();
    }
    return PCTTS_OK;
}

// out[c] (+)= sum_r x[r, c]   (bias gradients)
extern "C" int pctts_colsum(const float* x, long rows, int C, float* out, int accumulate, void* workspace, void* stream) {
    if (!x || !out || !workspace || rows <= 0 || C <= 0) return PCTTS_EINVAL;
    Ws w = carve(workspace, rows, C);
    cudaStream_t st = (cudaStream_t)stream;
    RedArgs a{x, nullptr, nullptr, nullptr, nullptr, nullptr, rows, C, 0};
    PCTTS_CUDA_RC(run_colreduce<MODE_COLSUM>(a, w, st));
    pctts_launch(sums_to_grads_kernel, pctts_div_up(C, 128), 128, 0, st, w.sums, C, out, nullptr, accumulate, 1.0);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}
