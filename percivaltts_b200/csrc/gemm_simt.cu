// fp32 SIMT GEMM with segmented row addressing and fused epilogues.
//
// This is the exact-fp32 contraction used for every Dense / Conv1D product of the reference graph
// (networktts.py:59-63 pFC, :116-120 pCNN1D, LSTM input projections networktts.py:90-94) and for their
// data / weight gradients.  The tcgen05 path (gemm_tc.cu) replaces it for the large contractions.
//
// A stored matrix is described by (ptr, ld, seg, segstride): element (r, c) lives at
//     ptr[(r / seg) * segstride + (r % seg) * ld + c]          (seg <= 0: plain ptr[r*ld + c])
// which lets a zero-padded [B, T+W-1, C] buffer be read as the *virtual* im2col matrix of a 'same'
// Conv1D (row (b,t) = the W*C contiguous floats starting at xp[b, t, 0]; ld = C, seg = T,
// segstride = (T+W-1)*C) without materialising it.
#include "common.cuh"
#include "pctts.h"
#include <stdint.h>
#include <stdlib.h>

namespace {

struct Mat {
    const float* p;
    long ld;
    int seg;
    long segstride;
};

__device__ __forceinline__ long row_off(const Mat& m, int r) {
    if (m.seg > 0) {
        int s = r / m.seg;
        return (long)s * m.segstride + (long)(r - s * m.seg) * m.ld;
    }
    return (long)r * m.ld;
}

constexpr int BM = 128, BN = 64, BK = 16, NT = 256;
constexpr int TM = 8, TN = 4;

struct Epi {
    const float* bias;      // [N] or null
    const float* aux;       // lrelu-derivative source (same addressing as C) or null
    const float* addmod;    // [addrows, N] contiguous, added as addmod[m % addrows, n], or null
    int addrows;
    int flags;
    float alpha;            // scales the accumulator before everything else
};

template <bool TA, bool TB>
__global__ void __launch_bounds__(NT) gemm_f32_kernel(int M, int N, int K, Mat A, Mat B, Mat C, Epi e, int kchunk,
                                                       float* __restrict__ partial) {
    pdl_entry();
    __shared__ __align__(16) float As[BK][BM + 4];
    __shared__ __align__(16) float Bs[BK][BN + 4];
    const int tid = threadIdx.x;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int ty = tid / 16, tx = tid % 16;

    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

    // per-thread global row offsets that do not depend on the k-block
    long a_off[8];
    bool a_ok[8];
    if (!TA) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            int m = m0 + tid / 16 + 16 * i;
            a_ok[i] = m < M;
            a_off[i] = a_ok[i] ? row_off(A, m) : 0;
        }
    }
    long b_off[4];
    bool b_ok[4];
    if (TB) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int n = n0 + tid / 16 + 16 * i;
            b_ok[i] = n < N;
            b_off[i] = b_ok[i] ? row_off(B, n) : 0;
        }
    }

    // split-K: CTA z owns k in [z*kchunk, (z+1)*kchunk) and writes raw sums to partial[z][M][N] (kchunk % BK == 0)
    const int kbeg = partial ? blockIdx.z * kchunk : 0;
    const int kend = partial ? min(K, kbeg + kchunk) : K;
    for (int k0 = kbeg; k0 < kend; k0 += BK) {
        float ra[8], rb[4];
        if (!TA) {
            const int k = k0 + (tid % 16);
#pragma unroll
            for (int i = 0; i < 8; ++i) ra[i] = (a_ok[i] && k < kend) ? __ldg(A.p + a_off[i] + k) : 0.f;
        } else {
            const int m = m0 + (tid % 128);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                int k = k0 + tid / 128 + 2 * i;
                ra[i] = (m < M && k < kend) ? __ldg(A.p + row_off(A, k) + m) : 0.f;
            }
        }
        if (!TB) {
            const int n = n0 + (tid % 64);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                int k = k0 + tid / 64 + 4 * i;
                rb[i] = (n < N && k < kend) ? __ldg(B.p + row_off(B, k) + n) : 0.f;
            }
        } else {
            const int k = k0 + (tid % 16);
#pragma unroll
            for (int i = 0; i < 4; ++i) rb[i] = (b_ok[i] && k < kend) ? __ldg(B.p + b_off[i] + k) : 0.f;
        }
        __syncthreads();   // previous tile fully consumed
        if (!TA) {
#pragma unroll
            for (int i = 0; i < 8; ++i) As[tid % 16][tid / 16 + 16 * i] = ra[i];
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) As[tid / 128 + 2 * i][tid % 128] = ra[i];
        }
        if (!TB) {
#pragma unroll
            for (int i = 0; i < 4; ++i) Bs[tid / 64 + 4 * i][tid % 64] = rb[i];
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) Bs[tid % 16][tid / 16 + 16 * i] = rb[i];
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            float4 a0 = *reinterpret_cast<const float4*>(&As[kk][ty * TM]);
            float4 a1 = *reinterpret_cast<const float4*>(&As[kk][ty * TM + 4]);
            float4 b0 = *reinterpret_cast<const float4*>(&Bs[kk][tx * TN]);
            const float a[TM] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float b[TN] = {b0.x, b0.y, b0.z, b0.w};
#pragma unroll
            for (int i = 0; i < TM; ++i)
#pragma unroll
                for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
    }

    if (partial) {
        float* pz = partial + (long)blockIdx.z * M * N;
#pragma unroll
        for (int i = 0; i < TM; ++i) {
            const int m = m0 + ty * TM + i;
            if (m >= M) continue;
#pragma unroll
            for (int j = 0; j < TN; ++j) {
                const int n = n0 + tx * TN + j;
                if (n < N) pz[(long)m * N + n] = acc[i][j];
            }
        }
        return;
    }
    float* Cp = const_cast<float*>(C.p);
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        const int m = m0 + ty * TM + i;
        if (m >= M) continue;
        const long coff = row_off(C, m);
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            const int n = n0 + tx * TN + j;
            if (n >= N) continue;
            float v = acc[i][j] * e.alpha;
            if (e.flags & PCTTS_EPI_BIAS) v += __ldg(e.bias + n);
            if (e.flags & PCTTS_EPI_ADDMOD) v += __ldg(e.addmod + (long)(m % e.addrows) * N + n);
            if (e.flags & PCTTS_EPI_LRELU) v = lrelu_f(v);
            if (e.flags & PCTTS_EPI_SIGMOID) v = sigmoid_f(v);
            if (e.flags & PCTTS_EPI_DLRELU_AUX) v *= dlrelu_from_out(e.aux[coff + n]);   // aux may alias C (tangent sweep)
            if (e.flags & PCTTS_EPI_ATOMIC) {
                atomicAdd(Cp + coff + n, v);
            } else {
                if (e.flags & PCTTS_EPI_ACCUM) v += Cp[coff + n];
                Cp[coff + n] = v;
            }
        }
    }
}


// ---- degenerate shapes of the Dense(1) heads (networks_critic.py:96, modeltts_common.py:84): the 128 x 64 tiles above spend
// 17-37 us on what is a matrix-vector product, an outer product or a weighted column sum of a 12 MB matrix -----------------------
__device__ __forceinline__ void epi_store(const Mat& C, const Epi& e, int m, int n, int N, float v) {
    float* Cp = const_cast<float*>(C.p);
    const long off = row_off(C, m) + n;
    v *= e.alpha;
    if (e.flags & PCTTS_EPI_BIAS) v += __ldg(e.bias + n);
    if (e.flags & PCTTS_EPI_ADDMOD) v += __ldg(e.addmod + (long)(m % e.addrows) * N + n);
    if (e.flags & PCTTS_EPI_LRELU) v = lrelu_f(v);
    if (e.flags & PCTTS_EPI_SIGMOID) v = sigmoid_f(v);
    if (e.flags & PCTTS_EPI_DLRELU_AUX) v *= dlrelu_from_out(e.aux[off]);
    if (e.flags & PCTTS_EPI_ATOMIC) {
        atomicAdd(Cp + off, v);
    } else {
        if (e.flags & PCTTS_EPI_ACCUM) v += Cp[off];
        Cp[off] = v;
    }
}

// N == 1, A not transposed: y[m] = epilogue(sum_k A[m,k] b[k]); one warp per row, lanes stride k (coalesced), shuffle tree
__global__ void __launch_bounds__(256) gemv_rows_kernel(int M, int K, Mat A, Mat B, int tb, Mat C, Epi e) {
    pdl_entry();
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    for (int m = blockIdx.x * wpb + (threadIdx.x >> 5); m < M; m += gridDim.x * wpb) {
        const float* a = A.p + row_off(A, m);
        float s = 0.f;
        if (tb) {
            for (int k = lane; k < K; k += 32) s = fmaf(a[k], __ldg(B.p + k), s);
        } else {
            for (int k = lane; k < K; k += 32) s = fmaf(a[k], __ldg(B.p + row_off(B, k)), s);
        }
        s = warp_sum(s);
        if (lane == 0) epi_store(C, e, m, 0, 1, s);
    }
}

// K == 1: C[m,n] = epilogue(a[m] b[n])
__global__ void __launch_bounds__(256) outer_kernel(int M, int N, Mat A, int ta, Mat B, int tb, Mat C, Epi e) {
    pdl_entry();
    const long total = (long)M * N;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int m = (int)(i / N), n = (int)(i - (long)m * N);
        const float a = ta ? A.p[m] : A.p[row_off(A, m)];
        const float b = tb ? __ldg(B.p + row_off(B, n)) : __ldg(B.p + n);
        epi_store(C, e, m, n, N, a * b);
    }
}

// N == 1, A transposed (stored [K, M], 16-byte aligned rows), split over k: partial[z][m] = sum_{k in slab z} A[k,m] b[k].
// A warp covers 128 columns (one float4 per lane) and walks a sub-slab of the CTA's rows eight at a time (4 KB in flight per
// warp: the first version, one scalar column per thread, kept 184 KB in flight on the whole GPU and took 60 us for 12 MB);
// the sub-slabs of a CTA are summed in index order through shared memory.
__global__ void __launch_bounds__(256) wcolsum_kernel(int M, int K, Mat A, Mat B, int tb, int kchunk, int mg, int ksub,
                                                      float* __restrict__ partial) {
    pdl_entry();
    __shared__ float4 red[8][32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = warp % mg, ks = warp / mg;                    // column group, row sub-slab of this warp
    const int mgrp = blockIdx.x * mg + g;
    const int m = mgrp * 128 + lane * 4;
    const int k0 = blockIdx.y * kchunk, k1 = min(K, k0 + kchunk);
    const int per = (k1 - k0 + ksub - 1) / ksub;
    const int ka = k0 + ks * per, kb = min(k1, ka + per);
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ks < ksub && m < M) {
        int k = ka;
        for (; k + 7 < kb; k += 8) {
            float4 v[8];
            float bb[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                v[u] = *reinterpret_cast<const float4*>(A.p + (long)(k + u) * A.ld + m);
                bb[u] = tb ? __ldg(B.p + k + u) : __ldg(B.p + (long)(k + u) * B.ld);
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                s.x = fmaf(v[u].x, bb[u], s.x), s.y = fmaf(v[u].y, bb[u], s.y);
                s.z = fmaf(v[u].z, bb[u], s.z), s.w = fmaf(v[u].w, bb[u], s.w);
            }
        }
        for (; k < kb; ++k) {
            const float4 v = *reinterpret_cast<const float4*>(A.p + (long)k * A.ld + m);
            const float bk = tb ? __ldg(B.p + k) : __ldg(B.p + (long)k * B.ld);
            s.x = fmaf(v.x, bk, s.x), s.y = fmaf(v.y, bk, s.y), s.z = fmaf(v.z, bk, s.z), s.w = fmaf(v.w, bk, s.w);
        }
    }
    red[warp][lane] = s;
    __syncthreads();
    if (ks == 0 && m < M) {
        float4 t = red[g][lane];
        for (int j = 1; j < ksub; ++j) {
            const float4 o = red[j * mg + g][lane];
            t.x += o.x, t.y += o.y, t.z += o.z, t.w += o.w;
        }
        *reinterpret_cast<float4*>(partial + (long)blockIdx.y * M + m) = t;
    }
}

// C = alpha * sum_z partial[z] (+ C when accumulating), fixed summation order => deterministic
__global__ void splitk_reduce_kernel(const float* __restrict__ partial, int splits, int M, int N, Mat C, float alpha, int accum) {
    pdl_entry();
    const long total = (long)M * N;
    float* Cp = const_cast<float*>(C.p);
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        float s = 0.f;
        for (int z = 0; z < splits; ++z) s += partial[(long)z * total + i];
        const int m = (int)(i / N), n = (int)(i - (long)m * N);
        const long off = row_off(C, m) + n;
        s *= alpha;
        Cp[off] = accum ? Cp[off] + s : s;
    }
}

int launch(int ta, int tb, int M, int N, int K, Mat A, Mat B, Mat C, Epi e, cudaStream_t st, int splits = 1,
           float* partial = nullptr) {
    if (M <= 0 || N <= 0) return PCTTS_OK;
    if (K < 0) return PCTTS_EINVAL;
    int kchunk = K;
    if (splits > 1) {
        kchunk = ((K + splits - 1) / splits + BK - 1) / BK * BK;
        splits = (K + kchunk - 1) / kchunk;
    }
    if (splits <= 1) partial = nullptr;
    static const bool fast1 = !(getenv("PCTTS_NO_DENSE1") && atoi(getenv("PCTTS_NO_DENSE1")));
    if (fast1 && K == 1 && !partial) {
        int blocks = pctts_div_up((long)M * N, 256);
        if (blocks > 148 * 8) blocks = 148 * 8;
        pctts_launch(outer_kernel, blocks, 256, 0, st, M, N, A, ta, B, tb, C, e);
        PCTTS_LAUNCH_CHECK();
        return PCTTS_OK;
    }
    if (fast1 && N == 1 && !ta && !partial && K >= 1) {
        int blocks = pctts_div_up(M, 8);
        if (blocks > 148 * 8) blocks = 148 * 8;
        pctts_launch(gemv_rows_kernel, blocks, 256, 0, st, M, K, A, B, tb, C, e);
        PCTTS_LAUNCH_CHECK();
        return PCTTS_OK;
    }
    if (fast1 && N == 1 && ta && partial && A.seg <= 0 && (tb || B.seg <= 0) && (M & 3) == 0 && (A.ld & 3) == 0 &&
        (((uintptr_t)A.p | (uintptr_t)partial) & 15) == 0) {
        const int groups = pctts_div_up(M, 128);
        const int mg = groups < 8 ? groups : 8, ksub = 8 / mg;
        dim3 g1(pctts_div_up(groups, mg), splits);
        pctts_launch(wcolsum_kernel, g1, 256, 0, st, M, K, A, B, tb, kchunk, mg, ksub, partial);
        PCTTS_LAUNCH_CHECK();
        int blocks = pctts_div_up(M, 256);
        pctts_launch(splitk_reduce_kernel, blocks, 256, 0, st, partial, splits, M, N, C, e.alpha, (e.flags & PCTTS_EPI_ACCUM) ? 1 : 0);
        PCTTS_LAUNCH_CHECK();
        return PCTTS_OK;
    }
    dim3 grid(pctts_div_up(N, BN), pctts_div_up(M, BM), partial ? splits : 1);
    if (!ta && !tb) pctts_launch(gemm_f32_kernel<false, false>, grid, NT, 0, st, M, N, K, A, B, C, e, kchunk, partial);
    else if (!ta && tb) pctts_launch(gemm_f32_kernel<false, true>, grid, NT, 0, st, M, N, K, A, B, C, e, kchunk, partial);
    else if (ta && !tb) pctts_launch(gemm_f32_kernel<true, false>, grid, NT, 0, st, M, N, K, A, B, C, e, kchunk, partial);
    else pctts_launch(gemm_f32_kernel<true, true>, grid, NT, 0, st, M, N, K, A, B, C, e, kchunk, partial);
    PCTTS_LAUNCH_CHECK();
    if (partial) {
        const long total = (long)M * N;
        int blocks = pctts_div_up(total, 256);
        if (blocks > 148 * 8) blocks = 148 * 8;
        pctts_launch(splitk_reduce_kernel, blocks, 256, 0, st, partial, splits, M, N, C, e.alpha, (e.flags & PCTTS_EPI_ACCUM) ? 1 : 0);
        PCTTS_LAUNCH_CHECK();
    }
    return PCTTS_OK;
}

}  // namespace

extern "C" int pctts_gemm_f32(int transA, int transB, int M, int N, int K, float alpha,
                              const float* A, long lda, int segA, long segstrideA,
                              const float* B, long ldb, int segB, long segstrideB,
                              float* C, long ldc, int segC, long segstrideC,
                              const float* bias, const float* aux, const float* addmod, int addrows,
                              int flags, void* stream) {
    if (!A || !B || !C) return PCTTS_EINVAL;
    if ((flags & PCTTS_EPI_BIAS) && !bias) return PCTTS_EINVAL;
    if ((flags & PCTTS_EPI_DLRELU_AUX) && !aux) return PCTTS_EINVAL;
    if ((flags & PCTTS_EPI_ADDMOD) && (!addmod || addrows <= 0)) return PCTTS_EINVAL;
    Mat a{A, lda, segA, segstrideA}, b{B, ldb, segB, segstrideB}, c{C, ldc, segC, segstrideC};
    Epi e{bias, aux, addmod, addrows, flags, alpha};
    return launch(transA, transB, M, N, K, a, b, c, e, (cudaStream_t)stream);
}


// Split-K variant for long reductions with few output tiles (weight gradients: M,N ~ 256, K = B*T rows).
// workspace holds splits*M*N floats; only alpha and PCTTS_EPI_ACCUM are applied (by the deterministic reduce pass).
extern "C" int pctts_gemm_splitk_f32(int transA, int transB, int M, int N, int K, float alpha,
                                     const float* A, long lda, int segA, long segstrideA,
                                     const float* B, long ldb, int segB, long segstrideB,
                                     float* C, long ldc, int segC, long segstrideC,
                                     int flags, int splits, float* workspace, void* stream) {
    if (!A || !B || !C || !workspace || splits < 1) return PCTTS_EINVAL;
    if (flags & ~PCTTS_EPI_ACCUM) return PCTTS_EUNSUPPORTED;
    Mat a{A, lda, segA, segstrideA}, b{B, ldb, segB, segstrideB}, c{C, ldc, segC, segstrideC};
    Epi e{nullptr, nullptr, nullptr, 0, flags, alpha};
    return launch(transA, transB, M, N, K, a, b, c, e, (cudaStream_t)stream, splits, workspace);
}
