// GENERAL-SHAPE FALLBACK of the recurrence: used only where lstm_fast.cu does not apply (H not a multiple of 32 up to 256, the tiny
// test configurations) or with PCTTS_LSTM_FAST=0 for A/B runs.  The demo shape (H = 256) runs lstm_fast.cu.
//
// Persistent LSTM recurrence (forward and BPTT) for pLSTM / pBLSTM (networktts.py:72-96): Keras LSTM with tanh /
// sigmoid, gate order i,f,c,o, return_sequences; Bidirectional = two directions concatenated [fwd, bwd].
//
// One thread-block cluster per (direction, batch chunk).  CTA r of the cluster owns hidden units
// [r*HS, (r+1)*HS) and keeps the matching 4*HS gate columns of the recurrent matrix U resident in shared memory
// for all T steps; h_t slices are exchanged through distributed shared memory with one cluster barrier per step.
// The input projection x.W + b is hoisted out (one GEMM over all time steps) and read here as zx.
#include <cooperative_groups.h>
#include <stdlib.h>
#include "common.cuh"
#include "pctts.h"

namespace cg = cooperative_groups;

// lstm_fast.cu: register-resident, dataflow-synchronised kernels for H in {32, 64, 128, 256}
int pctts_lstm_fast_supported(int H);
int pctts_lstm_fast(bool bwd, const float* const* zx, const float* const* U, float* hout, float* const* gates, float* const* cstate,
                    const float* dhout, float* const* dz, int B, int T, int H, int ndir, int ldh, cudaStream_t st);

namespace {

constexpr int NTH = 256;

template <int BP>
__device__ __forceinline__ void load_vec(const float* p, float (&v)[BP]) {
    if constexpr (BP % 4 == 0) {
#pragma unroll
        for (int i = 0; i < BP / 4; ++i) {
            const float4 q = *reinterpret_cast<const float4*>(p + 4 * i);
            v[4 * i] = q.x; v[4 * i + 1] = q.y; v[4 * i + 2] = q.z; v[4 * i + 3] = q.w;
        }
    } else if constexpr (BP == 2) {
        const float2 q = *reinterpret_cast<const float2*>(p);
        v[0] = q.x; v[1] = q.y;
    } else {
#pragma unroll
        for (int i = 0; i < BP; ++i) v[i] = p[i];
    }
}

struct LstmArgs {
    const float* zx[2];      // [B,T,4H] per direction: x.W + b
    const float* U[2];       // [H,4H] per direction
    float* hout;             // [B,T,ldh]; direction d writes columns d*H .. d*H+H-1
    float* gates[2];         // [B,T,4H] post-activation i,f,g,o (written fwd when non-null, read bwd)
    float* cstate[2];        // [B,T,H]
    const float* dhout;      // bwd: [B,T,ldh]
    float* dz[2];            // bwd: [B,T,4H] pre-activation gradients
    int B, T, H, ldh, CS, HS, bc;
};

// shared-memory carve-up (floats): Us[H][LD] | hT or part [2][H][BP] | zpart / dzT | misc
template <int BP>
__global__ void __launch_bounds__(NTH) lstm_fwd_kernel(LstmArgs a) {
    cg::cluster_group cluster = cg::this_cluster();
    extern __shared__ __align__(16) float smem[];
    const int H = a.H, HS = a.HS, T = a.T;
    const int ncols = 4 * HS, LD = ncols + 1;
    const int rank = (int)cluster.block_rank();
    const int d = blockIdx.z;
    const int b0 = blockIdx.y * a.bc;
    const int bcur = min(a.bc, a.B - b0);
    const int j0 = rank * HS;
    const int nj = max(0, min(HS, H - j0));
    const int tid = threadIdx.x;
    const bool reverse = (d == 1);

    float* Us = smem;                                   // [H][LD]
    float* hT = Us + (((long)H * LD + 3) & ~3);         // [2][H][BP]
    float* zpart = hT + 2 * H * BP;                     // [KS][ncols][BP]
    const int KS = max(1, NTH / ncols);
    float* cst = zpart + KS * ncols * BP;               // [HS*BP]

    const float* U = a.U[d];
    for (int i = tid; i < H * ncols; i += NTH) {
        const int k = i / ncols, lc = i - k * ncols;
        const int q = lc / HS, jj = lc - q * HS;
        Us[k * LD + lc] = (jj < nj) ? __ldg(U + (long)k * 4 * H + q * H + j0 + jj) : 0.f;
    }
    for (int i = tid; i < 2 * H * BP; i += NTH) hT[i] = 0.f;
    for (int i = tid; i < HS * BP; i += NTH) cst[i] = 0.f;
    cluster.sync();

    const int lc = tid % ncols, ks = tid / ncols;
    const bool mv_active = ks < KS;
    const int kb = mv_active ? (int)((long)ks * H / KS) : 0;
    const int ke = mv_active ? (int)((long)(ks + 1) * H / KS) : 0;
    const float* zx = a.zx[d];
    float* gates = a.gates[d];
    float* cstate = a.cstate[d];

    for (int s = 0; s < T; ++s) {
        const int t = reverse ? T - 1 - s : s;
        const float* hcur = hT + (s & 1) * H * BP;
        float* hnext_local = hT + ((s + 1) & 1) * H * BP;
        // ---- recurrent mat-vec: zpart[ks][lc][b] = sum_{k in range} h[b][k] * U[k][lc]
        if (mv_active) {
            float acc[BP];
#pragma unroll
            for (int b = 0; b < BP; ++b) acc[b] = 0.f;
#pragma unroll 4
            for (int k = kb; k < ke; ++k) {
                const float u = Us[k * LD + lc];
                float hv[BP];
                load_vec<BP>(hcur + k * BP, hv);
#pragma unroll
                for (int b = 0; b < BP; ++b) acc[b] = fmaf(hv[b], u, acc[b]);
            }
#pragma unroll
            for (int b = 0; b < BP; ++b) zpart[(ks * ncols + lc) * BP + b] = acc[b];
        }
        __syncthreads();
        // ---- gates and state update for the local hidden slice
        for (int idx = tid; idx < HS * BP; idx += NTH) {
            const int jj = idx % HS, b = idx / HS;
            if (jj >= nj || b >= bcur) continue;
            const long row = (long)(b0 + b) * T + t;
            float z[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                float v = __ldg(zx + row * 4 * H + q * H + j0 + jj);
                for (int kk = 0; kk < KS; ++kk) v += zpart[(kk * ncols + q * HS + jj) * BP + b];
                z[q] = v;
            }
            const float ig = 1.f / (1.f + expf(-z[0]));
            const float fg = 1.f / (1.f + expf(-z[1]));
            const float gg = tanhf(z[2]);
            const float og = 1.f / (1.f + expf(-z[3]));
            const float c = fg * cst[idx] + ig * gg;
            cst[idx] = c;
            const float h = og * tanhf(c);
            a.hout[row * a.ldh + d * H + j0 + jj] = h;
            if (gates) {
                float* gp = gates + row * 4 * H + j0 + jj;
                gp[0] = ig; gp[H] = fg; gp[2 * H] = gg; gp[3 * H] = og;
                cstate[row * H + j0 + jj] = c;
            }
            const int off = (int)(hnext_local - smem) + (j0 + jj) * BP + b;
            for (int peer = 0; peer < a.CS; ++peer) cluster.map_shared_rank(smem, peer)[off] = h;
        }
        cluster.sync();
    }
}

template <int BP>
__global__ void __launch_bounds__(NTH) lstm_bwd_kernel(LstmArgs a) {
    cg::cluster_group cluster = cg::this_cluster();
    extern __shared__ __align__(16) float smem[];
    const int H = a.H, HS = a.HS, T = a.T;
    const int ncols = 4 * HS, LD = ncols + 1;
    const int rank = (int)cluster.block_rank();
    const int d = blockIdx.z;
    const int b0 = blockIdx.y * a.bc;
    const int bcur = min(a.bc, a.B - b0);
    const int j0 = rank * HS;
    const int nj = max(0, min(HS, H - j0));
    const int tid = threadIdx.x;
    const bool reverse = (d == 1);

    float* Us = smem;                                   // [H][LD]
    float* part = Us + (((long)H * LD + 3) & ~3);       // [2][H][BP]
    float* dzT = part + 2 * H * BP;                     // [ncols][BP]
    float* dhrec = dzT + ncols * BP;                    // [HS*BP]
    float* dccar = dhrec + HS * BP;                     // [HS*BP]

    const float* U = a.U[d];
    for (int i = tid; i < H * ncols; i += NTH) {
        const int k = i / ncols, lc = i - k * ncols;
        const int q = lc / HS, jj = lc - q * HS;
        Us[k * LD + lc] = (jj < nj) ? __ldg(U + (long)k * 4 * H + q * H + j0 + jj) : 0.f;
    }
    for (int i = tid; i < ncols * BP; i += NTH) dzT[i] = 0.f;
    for (int i = tid; i < HS * BP; i += NTH) { dhrec[i] = 0.f; dccar[i] = 0.f; }
    cluster.sync();

    const float* gates = a.gates[d];
    const float* cstate = a.cstate[d];
    float* dz = a.dz[d];

    for (int s = 0; s < T; ++s) {
        // BPTT visits the steps in the opposite order of the forward recurrence
        const int t = reverse ? s : T - 1 - s;
        const int tprev = reverse ? t + 1 : t - 1;      // forward-order predecessor of t
        for (int idx = tid; idx < HS * BP; idx += NTH) {
            const int jj = idx % HS, b = idx / HS;
            if (jj >= nj || b >= bcur) continue;
            const long row = (long)(b0 + b) * T + t;
            const int j = j0 + jj;
            const float* gp = gates + row * 4 * H + j;
            const float ig = gp[0], fg = gp[H], gg = gp[2 * H], og = gp[3 * H];
            const float c = cstate[row * H + j];
            const float cprev = (tprev >= 0 && tprev < T) ? cstate[((long)(b0 + b) * T + tprev) * H + j] : 0.f;
            const float dh = a.dhout[row * a.ldh + d * H + j] + dhrec[idx];
            const float tc = tanhf(c);
            const float dc = dccar[idx] + dh * og * (1.f - tc * tc);
            const float dzo = dh * tc * og * (1.f - og);
            const float dzi = dc * gg * ig * (1.f - ig);
            const float dzg = dc * ig * (1.f - gg * gg);
            const float dzf = dc * cprev * fg * (1.f - fg);
            dccar[idx] = dc * fg;
            float* dp = dz + row * 4 * H + j;
            dp[0] = dzi; dp[H] = dzf; dp[2 * H] = dzg; dp[3 * H] = dzo;
            dzT[(0 * HS + jj) * BP + b] = dzi;
            dzT[(1 * HS + jj) * BP + b] = dzf;
            dzT[(2 * HS + jj) * BP + b] = dzg;
            dzT[(3 * HS + jj) * BP + b] = dzo;
        }
        __syncthreads();
        // partial dh_{prev}[b][k] over the local gate columns, for every hidden unit k
        float* pcur = part + (s & 1) * H * BP;
        for (int k = tid; k < H; k += NTH) {
            float acc[BP];
#pragma unroll
            for (int b = 0; b < BP; ++b) acc[b] = 0.f;
#pragma unroll 4
            for (int lc = 0; lc < ncols; ++lc) {
                const float u = Us[k * LD + lc];
                float dv[BP];
                load_vec<BP>(dzT + lc * BP, dv);
#pragma unroll
                for (int b = 0; b < BP; ++b) acc[b] = fmaf(dv[b], u, acc[b]);
            }
#pragma unroll
            for (int b = 0; b < BP; ++b) pcur[k * BP + b] = acc[b];
        }
        cluster.sync();
        // reduce-scatter over the cluster: my slice of dh_{prev}
        const int poff = (int)(pcur - smem);
        for (int idx = tid; idx < HS * BP; idx += NTH) {
            const int jj = idx % HS, b = idx / HS;
            if (jj >= nj || b >= bcur) continue;
            float v = 0.f;
            for (int peer = 0; peer < a.CS; ++peer) v += cluster.map_shared_rank(smem, peer)[poff + (j0 + jj) * BP + b];
            dhrec[idx] = v;
        }
        // (the next step's first barrier orders these reads before part[] is overwritten two steps later)
    }
    cluster.sync();   // no CTA may exit while peers still read its shared memory
}

size_t smem_floats(int H, int HS, int BP, bool bwd) {
    const int ncols = 4 * HS, LD = ncols + 1;
    size_t n = (((size_t)H * LD + 3) & ~(size_t)3) + 2 * (size_t)H * BP;
    if (!bwd) {
        const int KS = NTH / ncols > 1 ? NTH / ncols : 1;
        n += (size_t)KS * ncols * BP + (size_t)HS * BP;
    } else {
        n += (size_t)ncols * BP + 2 * (size_t)HS * BP;
    }
    return n + 4;
}

template <int BP>
int launch(bool bwd, const LstmArgs& a, int ndir, cudaStream_t st) {
    const size_t smem = smem_floats(a.H, a.HS, BP, bwd) * sizeof(float);
    if (smem > 227 * 1024) return PCTTS_EUNSUPPORTED;
    auto kern = bwd ? lstm_bwd_kernel<BP> : lstm_fwd_kernel<BP>;
    PCTTS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (a.CS > 8) PCTTS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(a.CS, pctts_div_up(a.B, a.bc), ndir);
    cfg.blockDim = dim3(NTH);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = a.CS;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    PCTTS_CUDA(cudaLaunchKernelEx(&cfg, kern, a));
    pctts_note_launch();
    return PCTTS_OK;
}

int dispatch(bool bwd, LstmArgs& a, int ndir, int cluster_size, int batch_chunk, cudaStream_t st) {
    if (a.B <= 0 || a.T <= 0 || a.H <= 0 || ndir < 1 || ndir > 2) return PCTTS_EINVAL;
    int CS = cluster_size;
    if (CS < 1) CS = 1;
    if (CS > 16) CS = 16;
    while (CS > 1 && (a.H + CS - 1) / CS * (CS - 1) >= a.H) CS >>= 1;   // every CTA must own >= 1 hidden unit
    a.CS = CS;
    a.HS = (a.H + CS - 1) / CS;
    if (4 * a.HS > NTH) return PCTTS_EUNSUPPORTED;
    int bc = batch_chunk < 1 ? 1 : (batch_chunk > 8 ? 8 : batch_chunk);
    if (bc > a.B) bc = a.B;
    a.bc = bc;
    if (bc <= 1) return launch<1>(bwd, a, ndir, st);
    if (bc <= 2) return launch<2>(bwd, a, ndir, st);
    if (bc <= 4) return launch<4>(bwd, a, ndir, st);
    return launch<8>(bwd, a, ndir, st);
}

// PCTTS_LSTM_FAST=0 (environment) or a negative cluster_size argument (-n = generic kernels with clusters of n) selects the shared-memory-resident generic kernels
bool use_fast(int H) {
    static const bool on = !(getenv("PCTTS_LSTM_FAST") && atoi(getenv("PCTTS_LSTM_FAST")) == 0);
    return on && pctts_lstm_fast_supported(H);
}

}  // namespace

extern "C" int pctts_lstm_fwd(const float* zx0, const float* zx1, const float* U0, const float* U1, float* hout, float* gates0,
                              float* gates1, float* c0, float* c1, int B, int T, int H, int ndir, int ldh, int cluster_size,
                              int batch_chunk, void* stream) {
    if (!zx0 || !U0 || !hout || (ndir == 2 && (!zx1 || !U1))) return PCTTS_EINVAL;
    LstmArgs a = {};
    a.zx[0] = zx0; a.zx[1] = zx1; a.U[0] = U0; a.U[1] = U1; a.hout = hout;
    a.gates[0] = gates0; a.gates[1] = gates1; a.cstate[0] = c0; a.cstate[1] = c1;
    a.B = B; a.T = T; a.H = H; a.ldh = ldh;
    if (B > 0 && T > 0 && cluster_size >= 0 && use_fast(H))
        return pctts_lstm_fast(false, a.zx, a.U, hout, a.gates, a.cstate, nullptr, nullptr, B, T, H, ndir, ldh, (cudaStream_t)stream);
    return dispatch(false, a, ndir, cluster_size < 0 ? -cluster_size : cluster_size, batch_chunk, (cudaStream_t)stream);
}

extern "C" int pctts_lstm_bwd(const float* dhout, const float* U0, const float* U1, const float* gates0, const float* gates1,
                              const float* c0, const float* c1, float* dz0, float* dz1, int B, int T, int H, int ndir, int ldh,
                              int cluster_size, int batch_chunk, void* stream) {
    if (!dhout || !U0 || !gates0 || !c0 || !dz0 || (ndir == 2 && (!U1 || !gates1 || !c1 || !dz1))) return PCTTS_EINVAL;
    LstmArgs a = {};
    a.dhout = dhout; a.U[0] = U0; a.U[1] = U1;
    a.gates[0] = const_cast<float*>(gates0); a.gates[1] = const_cast<float*>(gates1);
    a.cstate[0] = const_cast<float*>(c0); a.cstate[1] = const_cast<float*>(c1);
    a.dz[0] = dz0; a.dz[1] = dz1;
    a.B = B; a.T = T; a.H = H; a.ldh = ldh;
    if (B > 0 && T > 0 && cluster_size >= 0 && use_fast(H))
        return pctts_lstm_fast(true, nullptr, a.U, nullptr, a.gates, a.cstate, dhout, a.dz, B, T, H, ndir, ldh, (cudaStream_t)stream);
    return dispatch(true, a, ndir, cluster_size < 0 ? -cluster_size : cluster_size, batch_chunk, (cudaStream_t)stream);
}
