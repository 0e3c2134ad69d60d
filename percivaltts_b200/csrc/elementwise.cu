// Bandwidth-bound pieces of the WGAN-GP step: random interpolation + critic input packing, gradient-penalty tail,
// Wasserstein / spectrally-weighted LSE losses with their gradients, fused Adam, and small data movers.
// Reference: optimizertts_wgan.py:44-51 (RandomWeightedAverage), :53-68 (gradient_penalty_loss), :70-71
// (wasserstein_loss), :73-79 (specweighted_lse_loss), :145,172 (Adam), networks_critic.py:58 (spec slice).
#include "common.cuh"
#include "pctts.h"

namespace {

constexpr int ET = 256;

inline int ew_blocks(long n, int per_thread = 4) {
    long b = (n + (long)ET * per_thread - 1) / ((long)ET * per_thread);
    if (b < 1) b = 1;
    if (b > 148 * 8) b = 148 * 8;
    return (int)b;
}

// S0[0:B] = real[:,:,off:off+S], S0[B:2B] = fake[...], S0[2B:3B] = alpha*real + (1-alpha)*fake   (spec columns only:
// the critic reads nothing else from its feature input, networks_critic.py:58)
__global__ void critic_pack_kernel(const float* __restrict__ real, const float* __restrict__ fake,
                                   const float* __restrict__ alpha, float* __restrict__ s0, int B, int T, int O, int S, int off,
                                   int skip_real) {
    pdl_entry();
    const long per = (long)B * T * S;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < per; i += (long)gridDim.x * blockDim.x) {
        const int f = (int)(i % S);
        const long bt = i / S;
        const int b = (int)(bt / T);
        const float r = real[bt * O + off + f], k = fake[bt * O + off + f];
        const float a = alpha[b];
        if (!skip_real) s0[i] = r;
        s0[per + i] = k;
        s0[2 * per + i] = a * r + (1.f - a) * k;
    }
}

__global__ void fill_segments_kernel(float* __restrict__ p, long n0, float v0, long n1, float v1, long n2, float v2) {
    pdl_entry();
    const long n = n0 + n1 + n2;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
        p[i] = i < n0 ? v0 : (i < n0 + n1 ? v1 : v2);
}

// losses[slot] = scale * sum(x[0:n])        (single CTA; n is at most a few thousand on this path)
__global__ void sum_scaled_kernel(const float* __restrict__ x, long n, double scale, float* __restrict__ losses, int slot) {
    pdl_entry();
    __shared__ double scratch[32];
    double s = 0.0;
    for (long i = threadIdx.x; i < n; i += blockDim.x) s += x[i];
    s = block_sum_d(s, scratch);
    if (threadIdx.x == 0) losses[slot] = (float)(s * scale);
}

// ss[b] = sum over the sample of g^2   (one CTA per sample, deterministic)
__global__ void gp_norm_kernel(const float* __restrict__ g, long per_sample, float* __restrict__ ss) {
    pdl_entry();
    __shared__ double scratch[32];
    const float* p = g + (long)blockIdx.x * per_sample;
    double s = 0.0;
    for (long i = threadIdx.x; i < per_sample; i += blockDim.x) {
        const float v = p[i];
        s += (double)v * v;
    }
    s = block_sum_d(s, scratch);
    if (threadIdx.x == 0) ss[blockIdx.x] = (float)s;
}

// The same with GPN_SLICES CTAs per sample (the norm sits on the critical chain of the critic step: gradient -> norm -> penalty
// -> tangent sweep, and 10 CTAs of strided 4-byte loads ran at load latency): every CTA leaves an fp64 partial, the last one to
// arrive (ticket) sums the slices in index order -- deterministic -- and re-arms the ticket for the next launch / graph replay.
constexpr int GPN_SLICES = 16, GPN_MAXB = 256;
__device__ double g_gpn_partial[GPN_MAXB * GPN_SLICES];
__device__ unsigned int g_gpn_ticket[GPN_MAXB];

__global__ void __launch_bounds__(256) gp_norm_sliced_kernel(const float* __restrict__ g, long per_sample, float* __restrict__ ss) {
    pdl_entry();
    __shared__ double scratch[32];
    __shared__ int s_last;
    const int b = blockIdx.y, sl = blockIdx.x;
    const long chunk = (per_sample + GPN_SLICES - 1) / GPN_SLICES;
    const long lo = min(per_sample, sl * chunk), hi = min(per_sample, lo + chunk);
    const float* p = g + (long)b * per_sample;
    double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
    long i = lo + threadIdx.x;
    for (; i + 3L * blockDim.x < hi; i += 4L * blockDim.x) {
        const float v0 = p[i], v1 = p[i + blockDim.x], v2 = p[i + 2L * blockDim.x], v3 = p[i + 3L * blockDim.x];
        s0 += (double)v0 * v0, s1 += (double)v1 * v1, s2 += (double)v2 * v2, s3 += (double)v3 * v3;
    }
    for (; i < hi; i += blockDim.x) {
        const float v = p[i];
        s0 += (double)v * v;
    }
    const double s = block_sum_d((s0 + s1) + (s2 + s3), scratch);
    if (threadIdx.x == 0) {
        g_gpn_partial[b * GPN_SLICES + sl] = s;
        __threadfence();
        s_last = atomicAdd(&g_gpn_ticket[b], 1u) == GPN_SLICES - 1;
        if (s_last) {
            __threadfence();
            double t = 0.0;
            for (int k = 0; k < GPN_SLICES; ++k) t += *(volatile double*)&g_gpn_partial[b * GPN_SLICES + k];
            ss[b] = (float)t;
            g_gpn_ticket[b] = 0u;
        }
    }
}

// r0[b,:] = lambda * 2 (n_b - 1) / (n_b * Bglobal) * g[b,:];  losses[gp_slot] = sum_b (1-n_b)^2 / Bglobal  (local share)
__global__ void gp_scale_kernel(const float* __restrict__ g, const float* __restrict__ ss, float* __restrict__ r0, int B,
                                long per_sample, float lambda, float inv_bglobal, float* __restrict__ losses, int gp_slot) {
    pdl_entry();
    const long n = (long)B * per_sample;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const int b = (int)(i / per_sample);
        const float nb = sqrtf(ss[b]);
        r0[i] = lambda * 2.f * (nb - 1.f) / nb * inv_bglobal * g[i];
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        float acc = 0.f;
        for (int b = 0; b < B; ++b) {
            const float nb = sqrtf(ss[b]);
            acc += (1.f - nb) * (1.f - nb);
        }
        losses[gp_slot] = acc * inv_bglobal;
    }
}

__global__ void loss_combine_kernel(float* __restrict__ losses, int out, int a, float ca, int b, float cb, int c, float cc) {
    pdl_entry();
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        float v = ca * losses[a];
        if (b >= 0) v += cb * losses[b];
        if (c >= 0) v += cc * losses[c];
        losses[out] = v;
    }
}

// spectrally weighted LSE: loss = scale_loss * sum((y-p)^2 * w[col]);  dp = scale_grad * 2 (p-y) w[col]
__global__ void lse_kernel(const float* __restrict__ pred, const float* __restrict__ real, const float* __restrict__ w,
                           float* __restrict__ dpred, long rows, int O, double scale_loss, float scale_grad,
                           double* __restrict__ partial, unsigned int* __restrict__ counter, float* __restrict__ losses, int slot) {
    pdl_entry();
    __shared__ double scratch[32];
    __shared__ bool is_last;
    const long n = rows * O;
    double s = 0.0;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const int c = (int)(i % O);
        const float wv = w ? w[c] : 1.f;
        const float d = pred[i] - real[i];
        s += (double)(d * d * wv);
        if (dpred) dpred[i] = scale_grad * 2.f * d * wv;
    }
    s = block_sum_d(s, scratch);
    if (threadIdx.x == 0) {
        partial[blockIdx.x] = s;
        __threadfence();
        is_last = (atomicAdd(counter, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last && threadIdx.x == 0) {
        __threadfence();
        double t = 0.0;
        for (unsigned int b = 0; b < gridDim.x; ++b) t += __ldcg(partial + b);
        losses[slot] = (float)(t * scale_loss);
        *counter = 0u;
    }
}

// the same over the column window [c0, c0 + nc) of the [rows, O] matrices (the generator step evaluates the F0 column on its own
// lane, behind the LSTM, and the other columns without waiting for it)
__global__ void lse_cols_kernel(const float* __restrict__ pred, const float* __restrict__ real, const float* __restrict__ w,
                                float* __restrict__ dpred, long rows, int O, int c0, int nc, double scale_loss, float scale_grad,
                                double* __restrict__ partial, unsigned int* __restrict__ counter, float* __restrict__ losses, int slot) {
    pdl_entry();
    __shared__ double scratch[32];
    __shared__ bool is_last;
    const long n = rows * nc;
    double s = 0.0;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const long r = i / nc;
        const int c = c0 + (int)(i - r * nc);
        const long k = r * O + c;
        const float wv = w ? w[c] : 1.f;
        const float d = pred[k] - real[k];
        s += (double)(d * d * wv);
        if (dpred) dpred[k] = scale_grad * 2.f * d * wv;
    }
    s = block_sum_d(s, scratch);
    if (threadIdx.x == 0) {
        partial[blockIdx.x] = s;
        __threadfence();
        is_last = (atomicAdd(counter, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last && threadIdx.x == 0) {
        __threadfence();
        double t = 0.0;
        for (unsigned int b = 0; b < gridDim.x; ++b) t += __ldcg(partial + b);
        losses[slot] = (float)(t * scale_loss);
        *counter = 0u;
    }
}

// Keras-2.2 Adam: t = step+1; lr_t = lr*sqrt(1-b2^t)/(1-b1^t); m,v updated; p -= lr_t*m/(sqrt(v)+eps)
__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                            long n, float lr, float b1, float b2, float eps, const int* __restrict__ step) {
    pdl_entry();
    __shared__ float lr_t_s;
    if (threadIdx.x == 0) {
        const double t = (double)(*step + 1);
        lr_t_s = (float)((double)lr * sqrt(1.0 - pow((double)b2, t)) / (1.0 - pow((double)b1, t)));
    }
    __syncthreads();
    const float lr_t = lr_t_s;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const float gi = g[i];
        const float mi = b1 * m[i] + (1.f - b1) * gi;
        const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
        m[i] = mi;
        v[i] = vi;
        p[i] -= lr_t * mi / (sqrtf(vi) + eps);
    }
}
__global__ void incr_kernel(int* step) {
    pdl_entry(); if (threadIdx.x == 0 && blockIdx.x == 0) *step += 1; }

__global__ void stamp_kernel(unsigned long long* slot) {
    if (threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        *slot = t;
    }
}

__global__ void add3_kernel(float* __restrict__ out, const float* __restrict__ a, const float* __restrict__ b,
                            const float* __restrict__ c, long n) {
    pdl_entry();
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) out[i] = (a[i] + b[i]) + c[i];
}

__global__ void add_kernel(float* __restrict__ out, const float* __restrict__ a, const float* __restrict__ b, long n) {
    pdl_entry();
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) out[i] = a[i] + b[i];
}

// dz = dy * y * (1-y) on a [rows, cols] window with independent leading dimensions
__global__ void sigmoid_bwd_kernel(const float* __restrict__ dy, long lddy, const float* __restrict__ y, long ldy,
                                   float* __restrict__ dz, long lddz, long rows, int cols) {
    pdl_entry();
    const long n = rows * cols;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const long r = i / cols;
        const int c = (int)(i - r * cols);
        const float yv = y[r * ldy + c];
        dz[r * lddz + c] = dy[r * lddy + c] * yv * (1.f - yv);
    }
}

__global__ void copy2d_kernel(float* __restrict__ dst, long ldd, const float* __restrict__ src, long lds, long rows, int cols,
                              int accumulate) {
    pdl_entry();
    const long n = rows * cols;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const long r = i / cols;
        const int c = (int)(i - r * cols);
        const float v = src[r * lds + c];
        float* d = dst + r * ldd + c;
        *d = accumulate ? *d + v : v;
    }
}

// dst [B, T+left+right, C] <- zero halo + src [B, T, C]
__global__ void pad_time_kernel(float* __restrict__ dst, const float* __restrict__ src, int B, int T, int C, int left, int right) {
    pdl_entry();
    const int Tp = T + left + right;
    const long n = (long)B * Tp * C;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        const long bt = i / C;
        const int tp = (int)(bt % Tp);
        const int b = (int)(bt / Tp);
        const int t = tp - left;
        dst[i] = (t >= 0 && t < T) ? src[((long)b * T + t) * C + c] : 0.f;
    }
}

// counter-based uniform [0,1): splitmix64 over (seed, counter, index); the counter lives on the device so that a
// captured graph draws fresh numbers on every replay
__device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
__global__ void uniform_kernel(float* __restrict__ out, int n, unsigned long long seed, unsigned long long* __restrict__ counter) {
    pdl_entry();
    const unsigned long long ctr = *counter;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        const unsigned long long r = splitmix64(splitmix64(seed ^ (ctr * 0xD1B54A32D192ED03ull)) + (unsigned long long)i);
        out[i] = (float)(r >> 40) * (1.0f / 16777216.0f);
    }
}
__global__ void incr64_kernel(unsigned long long* c) {
    pdl_entry(); if (threadIdx.x == 0 && blockIdx.x == 0) *c += 1ull; }

// Generation tail (modeltts.py:154-165 `denormalise`, vocoders.py:194-202 `synthesis`): CMP * std + mean per feature column,
// split into the vocoder's streams so that the band decompression kernels read contiguous [T,K] inputs:
//   out [T,O] (optional) the de-normalised features, f0 [T] = exp(out[:,0]), spec [T,S] = out[:,1:1+S], nm [T,NM] = out[:,1+S:1+S+NM]
__global__ void denorm_split_kernel(const float* __restrict__ cmp, const float* __restrict__ mean, const float* __restrict__ stdv,
                                    long T, int O, int S, int NM, float* __restrict__ out, float* __restrict__ f0,
                                    float* __restrict__ spec, float* __restrict__ nm) {
    pdl_entry();
    const long n = T * O;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const long t = i / O;
        const int c = (int)(i - t * O);
        float v = cmp[i];
        if (mean) v = v * __ldg(stdv + c) + __ldg(mean + c);
        if (out) out[i] = v;
        if (c == 0) {
            if (f0) f0[t] = expf(v);
        } else if (c < 1 + S) {
            if (spec) spec[t * S + (c - 1)] = v;
        } else if (c < 1 + S + NM) {
            if (nm) nm[t * NM + (c - 1 - S)] = v;
        }
    }
}

}  // namespace

extern "C" int pctts_critic_pack_inputs(const float* real, const float* fake, const float* alpha, float* s0, int B, int T, int O,
                                        int S, int off, int skip_real, void* stream) {
    if (!real || !fake || !alpha || !s0 || off < 0 || off + S > O) return PCTTS_EINVAL;
    const long per = (long)B * T * S;
    if (per == 0) return PCTTS_OK;
    pctts_launch(critic_pack_kernel, ew_blocks(per), ET, 0, (cudaStream_t)stream, real, fake, alpha, s0, B, T, O, S, off, skip_real);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

extern "C" int pctts_fill_segments(float* p, long n0, float v0, long n1, float v1, long n2, float v2, void* stream) {
    if (!p || n0 < 0 || n1 < 0 || n2 < 0) return PCTTS_EINVAL;
    if (n0 + n1 + n2 == 0) return PCTTS_OK;
    pctts_launch(fill_segments_kernel, ew_blocks(n0 + n1 + n2), ET, 0, (cudaStream_t)stream, p, n0, v0, n1, v1, n2, v2);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

extern "C" int pctts_sum_scaled(const float* x, long n, double scale, float* losses, int slot, void* stream) {
    if (!x || !losses || n < 0) return PCTTS_EINVAL;
    pctts_launch(sum_scaled_kernel, 1, 1024, 0, (cudaStream_t)stream, x, n, scale, losses, slot);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

extern "C" int pctts_gp_norms(const float* g, int B, long per_sample, float* ss, void* stream) {
    if (!g || !ss || B <= 0) return PCTTS_EINVAL;
    if (B <= GPN_MAXB && per_sample >= 4096)
        pctts_launch(gp_norm_sliced_kernel, dim3(GPN_SLICES, B), 256, 0, (cudaStream_t)stream, g, per_sample, ss);
    else
        pctts_launch(gp_norm_kernel, B, 1024, 0, (cudaStream_t)stream, g, per_sample, ss);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

extern "C" int pctts_gp_scale(const float* g, const float* ss, float* r0, int B, long per_sample, float lambda, int Bglobal,
                              float* losses, int gp_slot, void* stream) {
    if (!g || !ss || !r0 || !losses || B <= 0 || Bglobal <= 0) return PCTTS_EINVAL;
    pctts_launch(gp_scale_kernel, ew_blocks((long)B * per_sample), ET, 0, (cudaStream_t)stream, g, ss, r0, B, per_sample, lambda, 1.f / (float)Bglobal, losses, gp_slot);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

extern "C" int pctts_loss_combine(float* losses, int out, int a, float ca, int b, float cb, int c, float cc, void* stream) {
    if (!losses) return PCTTS_EINVAL;
    pctts_launch(loss_combine_kernel, 1, 32, 0, (cudaStream_t)stream, losses, out, a, ca, b, cb, c, cc);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

extern "C" long pctts_lse_workspace_bytes(void) { return 148 * 8 * (long)sizeof(double) + 16; }

extern "C" int pctts_lse_loss_grad(const float* pred, const float* real, const float* w, float* dpred, long rows, int O,
                                   double scale_loss, float scale_grad, float* losses, int slot, void* workspace, void* stream) {
    if (!pred || !real || !losses || !workspace || rows <= 0 || O <= 0) return PCTTS_EINVAL;
    double* partial = (double*)workspace;
    unsigned int* counter = (unsigned int*)(partial + 148 * 8);
    pctts_launch(lse_kernel, ew_blocks(rows * O), ET, 0, (cudaStream_t)stream, pred, real, w, dpred, rows, O, scale_loss, scale_grad, partial, counter, losses, slot);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

/* the same restricted to the columns [c0, c0 + nc) of the [rows, O] matrices (w indexed by the absolute column); the workspace must
 * not be shared with a call that may run concurrently on another stream */
extern "C" int pctts_lse_loss_grad_cols(const float* pred, const float* real, const float* w, float* dpred, long rows, int O, int c0,
                                        int nc, double scale_loss, float scale_grad, float* losses, int slot, void* workspace,
                                        void* stream) {
    if (!pred || !real || !losses || !workspace || rows <= 0 || O <= 0 || c0 < 0 || nc <= 0 || c0 + nc > O) return PCTTS_EINVAL;
    double* partial = (double*)workspace;
    unsigned int* counter = (unsigned int*)(partial + 148 * 8);
    pctts_launch(lse_cols_kernel, ew_blocks(rows * nc), ET, 0, (cudaStream_t)stream, pred, real, w, dpred, rows, O, c0, nc, scale_loss,
                 scale_grad, partial, counter, losses, slot);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

extern "C" int pctts_adam(float* p, const float* g, float* m, float* v, long n, float lr, float beta1, float beta2, float eps,
                          int* step, void* stream) {
    if (!p || !g || !m || !v || !step || n < 0) return PCTTS_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    if (n > 0) {
        pctts_launch(adam_kernel, ew_blocks(n), ET, 0, st, p, g, m, v, n, lr, beta1, beta2, eps, step);
        PCTTS_LAUNCH_CHECK();
    }
    pctts_launch(incr_kernel, 1, 32, 0, st, step);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

/* developer probe: writes %globaltimer (ns) into *slot when the stream reaches this point (tools/timeline.py) */
extern "C" int pctts_stamp(unsigned long long* slot, void* stream) {
    if (!slot) return PCTTS_EINVAL;
    stamp_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(slot);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

extern "C" int pctts_add(float* out, const float* a, const float* b, long n, void* stream) {
    if (!out || !a || !b || n < 0) return PCTTS_EINVAL;
    if (n == 0) return PCTTS_OK;
    pctts_launch(add_kernel, ew_blocks(n), ET, 0, (cudaStream_t)stream, out, a, b, n);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

/* out = (a + b) + c  (the generator's three branch shares of d(lctx)) */
extern "C" int pctts_add3(float* out, const float* a, const float* b, const float* c, long n, void* stream) {
    if (!out || !a || !b || !c || n < 0) return PCTTS_EINVAL;
    if (n == 0) return PCTTS_OK;
    pctts_launch(add3_kernel, ew_blocks(n), ET, 0, (cudaStream_t)stream, out, a, b, c, n);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

extern "C" int pctts_sigmoid_bwd(const float* dy, long lddy, const float* y, long ldy, float* dz, long lddz, long rows, int cols,
                                 void* stream) {
    if (!dy || !y || !dz || rows < 0 || cols < 0) return PCTTS_EINVAL;
    if (rows * cols == 0) return PCTTS_OK;
    pctts_launch(sigmoid_bwd_kernel, ew_blocks(rows * cols), ET, 0, (cudaStream_t)stream, dy, lddy, y, ldy, dz, lddz, rows, cols);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

extern "C" int pctts_copy2d(float* dst, long ldd, const float* src, long lds, long rows, int cols, int accumulate, void* stream) {
    if (!dst || !src || rows < 0 || cols < 0) return PCTTS_EINVAL;
    if (rows * cols == 0) return PCTTS_OK;
    pctts_launch(copy2d_kernel, ew_blocks(rows * cols), ET, 0, (cudaStream_t)stream, dst, ldd, src, lds, rows, cols, accumulate);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

extern "C" int pctts_pad_time(float* dst, const float* src, int B, int T, int C, int left, int right, void* stream) {
    if (!dst || !src || B < 0 || T < 0 || C <= 0 || left < 0 || right < 0) return PCTTS_EINVAL;
    const long n = (long)B * (T + left + right) * C;
    if (n == 0) return PCTTS_OK;
    pctts_launch(pad_time_kernel, ew_blocks(n), ET, 0, (cudaStream_t)stream, dst, src, B, T, C, left, right);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

extern "C" int pctts_denorm_split(const float* cmp, const float* mean, const float* stdv, long T, int O, int S, int NM, float* out,
                                  float* f0, float* spec, float* nm, void* stream) {
    if (!cmp || T < 0 || O <= 0 || S < 0 || NM < 0 || 1 + S + NM > O || ((mean == nullptr) != (stdv == nullptr))) return PCTTS_EINVAL;
    if (T == 0) return PCTTS_OK;
    pctts_launch(denorm_split_kernel, ew_blocks(T * O), ET, 0, (cudaStream_t)stream, cmp, mean, stdv, T, O, S, NM, out, f0, spec, nm);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

extern "C" int pctts_uniform(float* out, int n, unsigned long long seed, unsigned long long* counter, void* stream) {
    if (!out || !counter || n < 0) return PCTTS_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    if (n > 0) {
        pctts_launch(uniform_kernel, pctts_div_up(n, 128), 128, 0, st, out, n, seed, counter);
        PCTTS_LAUNCH_CHECK();
    }
    pctts_launch(incr64_kernel, 1, 32, 0, st, counter);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}
