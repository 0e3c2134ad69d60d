// One-shot all-reduce of a small fp64 vector over NVLink / NVSwitch peer memory (one process per GPU, one node).
//
// Use: the sync-BatchNorm exchange of the data-parallel step: per BatchNorm layer the fp64 column sums
// [sum x, sum x^2] (forward) and [sum dz, sum dz*xhat] (backward), 8 .. 512 doubles, have to be summed over the replicas before
// the layer can continue -- 11 such exchanges per critic step, 30 per generator step, each a latency-bound NCCL launch
// (~15 us inside the step graph).  Here every rank publishes its vector in a symmetric buffer (mapped into every peer's address
// space, torch.distributed._symmetric_memory), raises a flag in every peer's buffer, waits for the peers' flags in its own and
// sums the peers' vectors in RANK ORDER (bit-identical result on every rank, like a ring/tree all-reduce would not guarantee
// across sizes).  One CTA, no host involvement, capturable: the sequence number lives in device memory and advances per call,
// so a replayed graph keeps all ranks in lock step as long as they issue the same calls in the same order (they run the same
// program).
//
// Symmetric buffer layout per rank (bytes): [0] seq (u64, local) | [64] flags u64 [2][world] | [DATA] double [2][nmax]
// Both halves are indexed by the parity of seq: a slot is rewritten two calls later, and a rank reaches call seq+2 only after
// every peer has published seq+1, i.e. has finished reading seq.
#include "common.cuh"
#include "pctts.h"
#include <stdint.h>

namespace {

constexpr int PEER_MAX_WORLD = 16;
constexpr long long PEER_SPIN_LIMIT = 120000000000LL;   // ~60 s of SM clocks: ordinary rank skew (a rank capturing a graph, a slow batch
                                                         // load) is waited out like NCCL would; a peer that never comes ends in *err = 1 --
                                                         // which the trainer reads back with the losses of every step and raises on

__host__ __device__ inline long peer_data_offset(int world) { return 64 + ((2L * world * 8 + 127) / 128) * 128; }

__device__ __forceinline__ void st_release_sys(uint64_t* p, uint64_t v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ double ld_relaxed_sys_f64(const double* p) {
    double v;
    asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}

struct PeerBases {
    uint64_t b[PEER_MAX_WORLD];
};

__global__ void __launch_bounds__(256) peer_allreduce_kernel(double* __restrict__ x, int n, PeerBases bases, int rank, int world,
                                                             int nmax, int* __restrict__ err) {
    pdl_entry();
    __shared__ uint64_t s_seq;
    const int tid = threadIdx.x;
    uint8_t* const mine = reinterpret_cast<uint8_t*>(bases.b[rank]);
    if (tid == 0) s_seq = *reinterpret_cast<volatile uint64_t*>(mine) + 1;
    __syncthreads();
    const uint64_t seq = s_seq;
    const int par = (int)(seq & 1);
    const long doff = peer_data_offset(world);
    double* const mydata = reinterpret_cast<double*>(mine + doff) + (long)par * nmax;
    for (int i = tid; i < n; i += blockDim.x) mydata[i] = x[i];
    __threadfence_system();
    __syncthreads();
    if (tid < world) {
        // publish: my flag in peer `tid`'s buffer (my own included)
        st_release_sys(reinterpret_cast<uint64_t*>(bases.b[tid] + 64) + par * world + rank, seq);
        // wait for peer `tid`'s flag in mine
        const uint64_t* f = reinterpret_cast<const uint64_t*>(mine + 64) + par * world + tid;
        const long long t0 = clock64();
        while (ld_acquire_sys(f) < seq) {
            if (clock64() - t0 > PEER_SPIN_LIMIT) {
                *err = 1;
                break;
            }
        }
    }
    __syncthreads();
    for (int i = tid; i < n; i += blockDim.x) {
        double s = 0.0;
        for (int r = 0; r < world; ++r)
            s += ld_relaxed_sys_f64(reinterpret_cast<const double*>(bases.b[r] + doff) + (long)par * nmax + i);
        x[i] = s;
    }
    __syncthreads();
    if (tid == 0) *reinterpret_cast<volatile uint64_t*>(mine) = seq;
}

}  // namespace

/* bytes of the symmetric buffer every rank must allocate (zeroed, then a barrier, before the first call) */
extern "C" long pctts_peer_allreduce_bytes(int world, int nmax) {
    if (world < 1 || world > PEER_MAX_WORLD || nmax < 1) return -1;
    return peer_data_offset(world) + 2L * nmax * 8;
}

/* x[0:n] (fp64, device, in place) <- sum over the ranks, in rank order.  peer_bases: HOST array of `world` addresses, the
 * symmetric buffer of rank r as mapped into this process; err: device int set to 1 if a peer did not show up in ~60 s. */
extern "C" int pctts_peer_allreduce_f64(double* x, int n, const unsigned long long* peer_bases, int rank, int world, int nmax,
                                        int* err, void* stream) {
    if (!x || !peer_bases || !err || world < 1 || world > PEER_MAX_WORLD || rank < 0 || rank >= world || n < 0 || n > nmax)
        return PCTTS_EINVAL;
    if (n == 0) return PCTTS_OK;
    PeerBases b;
    for (int r = 0; r < PEER_MAX_WORLD; ++r) b.b[r] = r < world ? peer_bases[r] : 0;
    pctts_launch(peer_allreduce_kernel, 1, 256, 0, (cudaStream_t)stream, x, n, b, rank, world, nmax, err);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

// ---------------------------------------------------------------------------------------------------------------------------
// BatchNorm reductions with the replica exchange INSIDE the launch (sync-BN).
//
// Per BatchNorm layer the single-replica step runs  [first-stage partial sums] -> [second stage + finalize]  (forward) and
// [first-stage partials] -> [second stage] -> [backward apply] -> [sums -> dgamma / dbeta]  (backward).  On replicas the second
// stage, the peer-memory exchange and the finalize were three dependent launches per layer (41 layers' worth per critic +
// generator step).  The kernels below do  second stage -> publish -> wait -> rank-ordered sum -> finalize (or -> sums + dgamma /
// dbeta)  in ONE launch, one CTA per slice of 32 channels, each slice with its own sequence counter, flags and double-buffered
// slots in a second symmetric buffer.  Arithmetic and summation orders are those of colreduce_final_kernel / peer_allreduce_kernel /
// bn_finalize_kernel / sums_to_grads_kernel, so the results are bit-identical to the separate launches.  world == 1 (bases NULL)
// skips the exchange: the same launch then serves the single replica (backward: one launch fewer per layer).
//
// Symmetric buffer per rank: seq u64 [SL_MAX] | flags u64 [2][SL_MAX][world] | data double [2][SL_MAX][64]
namespace {

constexpr int SL_MAX = 32;           // slices of 32 channels: C <= 1024
constexpr int SL_COLS = 32;

__host__ __device__ inline long sl_flags_off() { return (long)SL_MAX * 8; }
__host__ __device__ inline long sl_data_off(int world) { return sl_flags_off() + ((2L * SL_MAX * world * 8 + 127) / 128) * 128; }
// Low-latency variant of the exchange (the default): every 32-bit half of a double travels as ONE 8-byte store {data, seq} straight
// into every peer's receive area, and a rank polls its OWN memory until the word carries the current sequence number -- one one-way
// NVLink write per word, no system-scope fence, no flag round, no remote load (the pull protocol above pays fence + flag + a remote
// read round trip, ~4 us more per exchange).  Receive area per rank: u64 [2][SL_MAX][world][128] behind the pull protocol's slots.
__host__ __device__ inline long sl_ll_off(int world) { return sl_data_off(world) + 2L * SL_MAX * 64 * 8; }
__host__ __device__ inline long sl_ll_bytes(int world) { return 2L * SL_MAX * world * 128 * 8; }
__device__ __forceinline__ void st_relaxed_sys_u64(uint64_t* p, uint64_t v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_relaxed_sys_u64(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

struct BnSyncArgs {
    const double* partial;   // [nblk][2][C]
    double* sums;            // [2][C] (global result, read by bn_bwd_apply / readers of the workspace)
    int nblk, C;
    int mode;                // 0: forward finalize, 1: backward (sums + dgamma / dbeta)
    double count;
    float eps, momentum;
    float *mean, *invstd, *moving_mean, *moving_var;      // mode 0
    float *g1, *g2;                                       // mode 1: g1 (+)= scale * sums[0] (dbeta), g2 (+)= scale * sums[1] (dgamma)
    double scale;
    int accumulate;
    int rank, world;
    int* err;
    int ll;                  // 1: low-latency push protocol, 0: publish / flag / pull
};

__global__ void __launch_bounds__(256) bn_sync_kernel(BnSyncArgs a, PeerBases bases) {
    pdl_entry();
    __shared__ double loc[2 * SL_COLS];
    __shared__ uint64_t s_seq;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int s = blockIdx.x, C = a.C;
    const int c0 = s * SL_COLS;
    const int ncol = min(SL_COLS, C - c0);
    // second stage: one warp per element, lanes stride over the CTA partials, butterfly (colreduce_final_kernel's order)
    for (int k = 0; k < 8; ++k) {
        const int e = warp * 8 + k, m = e >> 5, j = e & 31;
        double q = 0.0;
        if (j < ncol)
            for (int b = lane; b < a.nblk; b += 32) q += a.partial[(long)b * 2 * C + (long)m * C + c0 + j];
        q = warp_sum_d(q);
        if (lane == 0) loc[e] = q;
    }
    __syncthreads();
    if (a.world > 1) {
        uint8_t* const mine = reinterpret_cast<uint8_t*>(bases.b[a.rank]);
        if (tid == 0) s_seq = reinterpret_cast<volatile uint64_t*>(mine)[s] + 1;
        __syncthreads();
        const uint64_t seq = s_seq;
        const int par = (int)(seq & 1);
        if (a.ll) {
            const uint32_t tag = (uint32_t)seq;
            const long base = sl_ll_off(a.world) + (((long)par * SL_MAX + s) * a.world) * 128 * 8;
            if (tid < 128) {      // word tid = half (tid & 1) of double tid >> 1, to every peer's area for this rank
                const uint64_t bits = (uint64_t)__double_as_longlong(loc[tid >> 1]);
                const uint64_t word = ((uint64_t)tag << 32) | (uint32_t)(bits >> (32 * (tid & 1)));
                for (int r = 0; r < a.world; ++r)
                    if (r != a.rank) st_relaxed_sys_u64(reinterpret_cast<uint64_t*>(bases.b[r] + base) + (long)a.rank * 128 + tid, word);
            }
            double q = 0.0;
            if (tid < 64) {
                const long long t0 = clock64();
                for (int r = 0; r < a.world; ++r) {       // rank order: bit-identical sums on every rank
                    if (r == a.rank) {
                        q += loc[tid];
                        continue;
                    }
                    const uint64_t* w = reinterpret_cast<const uint64_t*>(mine + base) + (long)r * 128 + 2 * tid;
                    uint64_t lo, hi;
                    while ((uint32_t)((lo = ld_relaxed_sys_u64(w)) >> 32) != tag || (uint32_t)((hi = ld_relaxed_sys_u64(w + 1)) >> 32) != tag) {
                        if (clock64() - t0 > PEER_SPIN_LIMIT) {
                            *a.err = 1;
                            lo = hi = 0;
                            break;
                        }
                    }
                    q += __longlong_as_double((long long)(((hi & 0xffffffffull) << 32) | (lo & 0xffffffffull)));
                }
            }
            __syncthreads();
            if (tid < 64) loc[tid] = q;
            __syncthreads();
            if (tid == 0) reinterpret_cast<volatile uint64_t*>(mine)[s] = seq;
        } else {
        const long doff = sl_data_off(a.world) + ((long)par * SL_MAX + s) * 64 * 8;
        if (tid < 64) reinterpret_cast<double*>(mine + doff)[tid] = loc[tid];
        __threadfence_system();
        __syncthreads();
        if (tid < a.world) {
            const long foff = sl_flags_off() + (((long)par * SL_MAX + s) * a.world) * 8;
            st_release_sys(reinterpret_cast<uint64_t*>(bases.b[tid] + foff) + a.rank, seq);
            const uint64_t* f = reinterpret_cast<const uint64_t*>(mine + foff) + tid;
            const long long t0 = clock64();
            while (ld_acquire_sys(f) < seq) {
                if (clock64() - t0 > PEER_SPIN_LIMIT) {
                    *a.err = 1;
                    break;
                }
            }
        }
        __syncthreads();
        if (tid < 64) {
            double q = 0.0;
            for (int r = 0; r < a.world; ++r) q += ld_relaxed_sys_f64(reinterpret_cast<const double*>(bases.b[r] + doff) + tid);
            loc[tid] = q;
        }
        __syncthreads();
        if (tid == 0) reinterpret_cast<volatile uint64_t*>(mine)[s] = seq;
        }
    }
    if (tid < 64) {
        const int m = tid >> 5, j = tid & 31;
        if (j < ncol) a.sums[(long)m * C + c0 + j] = loc[tid];
    }
    if (tid < ncol) {
        const int c = c0 + tid;
        if (a.mode == 0) {
            const double mu = loc[tid] / a.count;
            double v = loc[SL_COLS + tid] / a.count - mu * mu;
            if (v < 0.0) v = 0.0;
            a.mean[c] = (float)mu;
            a.invstd[c] = (float)(1.0 / sqrt(v + (double)a.eps));
            if (a.moving_mean) {
                a.moving_mean[c] = a.momentum * a.moving_mean[c] + (1.f - a.momentum) * (float)mu;
                a.moving_var[c] = a.momentum * a.moving_var[c] + (1.f - a.momentum) * (float)v;
            }
        } else {
            if (a.g1) a.g1[c] = (a.accumulate ? a.g1[c] : 0.f) + (float)(loc[tid] * a.scale);
            if (a.g2) a.g2[c] = (a.accumulate ? a.g2[c] : 0.f) + (float)(loc[SL_COLS + tid] * a.scale);
        }
    }
}

int bn_sync_launch(BnSyncArgs& a, const unsigned long long* peer_bases, cudaStream_t st) {
    if (a.C < 1 || a.C > SL_MAX * SL_COLS || a.nblk < 1) return PCTTS_EINVAL;
    if (a.world > 1 && (!peer_bases || !a.err || a.world > PEER_MAX_WORLD || a.rank < 0 || a.rank >= a.world)) return PCTTS_EINVAL;
    PeerBases b;
    for (int r = 0; r < PEER_MAX_WORLD; ++r) b.b[r] = (a.world > 1 && r < a.world) ? peer_bases[r] : 0;
    static const int ll = getenv("PCTTS_BN_LL") ? atoi(getenv("PCTTS_BN_LL")) : 1;
    a.ll = ll;
    pctts_launch(bn_sync_kernel, pctts_div_up(a.C, SL_COLS), 256, 0, st, a, b);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

}  // namespace

/* bytes of the symmetric buffer of the sliced protocol (zeroed, then a barrier, before the first call) */
extern "C" long pctts_bn_sync_bytes(int world) {
    if (world < 1 || world > PEER_MAX_WORLD) return -1;
    return sl_ll_off(world) + sl_ll_bytes(world);
}

/* Forward BatchNorm statistics: second reduction stage over the nblk first-stage partials [nblk][2][C] (doubles) at `partial`,
 * all-reduce over the replicas (world > 1), sums -> `sums` [2][C], mean / invstd (and the moving averages) -- one launch. */
extern "C" int pctts_bn_sync_finalize(const double* partial, double* sums, int nblk, int C, double count, float eps, float momentum,
                                      float* mean, float* invstd, float* moving_mean, float* moving_var,
                                      const unsigned long long* peer_bases, int rank, int world, int* err, void* stream) {
    if (!partial || !sums || !mean || !invstd || count <= 0) return PCTTS_EINVAL;
    BnSyncArgs a{};
    a.partial = partial, a.sums = sums, a.nblk = nblk, a.C = C, a.mode = 0, a.count = count, a.eps = eps, a.momentum = momentum;
    a.mean = mean, a.invstd = invstd, a.moving_mean = moving_mean, a.moving_var = moving_var;
    a.rank = rank, a.world = world, a.err = err;
    return bn_sync_launch(a, peer_bases, (cudaStream_t)stream);
}

/* Backward BatchNorm sums [sum dz, sum dz*xhat]: second stage, all-reduce, sums -> `sums` (read by pctts_bn_bwd_apply), and
 * dbeta (+)= scale * sums[0], dgamma (+)= scale * sums[1] -- one launch. */
extern "C" int pctts_bn_sync_bwd(const double* partial, double* sums, int nblk, int C, float* dgamma, float* dbeta, float scale,
                                 int accumulate, const unsigned long long* peer_bases, int rank, int world, int* err, void* stream) {
    if (!partial || !sums) return PCTTS_EINVAL;
    BnSyncArgs a{};
    a.partial = partial, a.sums = sums, a.nblk = nblk, a.C = C, a.mode = 1, a.count = 1.0;
    a.g1 = dbeta, a.g2 = dgamma, a.scale = (double)scale, a.accumulate = accumulate;
    a.rank = rank, a.world = world, a.err = err;
    return bn_sync_launch(a, peer_bases, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------------------------------------------
// Gradient all-reduce FUSED with the Adam update ("gradients are reduced after each critic and generator step").
//
// The flat fp32 gradient buffer of a network lives in symmetric memory (mapped into every peer over NVLink).  One launch per
// network and step does a two-shot all-reduce and the optimiser update:
//   1. ready barrier    every rank raises its flag in every peer (its local gradients are complete: stream order);
//   2. reduce-scatter   rank r sums slice r of all the ranks' buffers in RANK ORDER (16-byte peer loads) ...
//   3. all-gather       ... and writes the sum back into slice r of every rank's buffer (16-byte peer stores), in place: slice r is
//                       touched by nobody else;
//   4. done barrier     last CTA of the grid raises the rank's second flag in every peer; everybody waits for all of them;
//   5. Adam             over the whole (now identical on every rank) gradient, Keras form as in adam_kernel (elementwise.cu),
//                       step counter advanced by the last CTA.
// No NCCL launch, no separate Adam launch; flags carry a sequence number that lives in device memory, so the launch is
// capturable and replays stay in lock step.  Control block per rank (symmetric, zeroed once):
//   [0] seq u64 | [8] cnt1 u32 | [12] cnt2 u32 | [64] ready u64[16] | [192] done u64[16] | [320] loss f32 [2][GS_NLOSS]
// The step's loss scalars (local shares of the global means, logging only) ride on the ready barrier: published in the control
// block in front of the flag, summed in rank order behind it -- no separate small all-reduce.
namespace {

constexpr int GS_THREADS = 256;      // 2 x 16 peer vectors in flight per thread: registers, not threads, carry the NVLink latency
constexpr int GS_NLOSS = 8;
constexpr long GS_READY = 64, GS_DONE = 192, GS_LOSS = 320, GS_BYTES = 512;

struct GradSyncArgs {
    uint64_t grad[PEER_MAX_WORLD];
    uint64_t ctl[PEER_MAX_WORLD];
    float *p, *m, *v;
    int* step;
    long n, n4;            // parameters; float4 vectors of the (zero-padded) gradient buffer
    float lr, b1, b2, eps;
    int rank, world;
    int* err;
    float* losses;         // nloss scalars summed over the ranks in place (or NULL)
    int nloss;
};
__device__ __forceinline__ float ld_relaxed_sys_f32(const float* p) {
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ float4 ld_relaxed_sys_v4(const float4* p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_sys_v4(float4* p, const float4& v) {
    asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

__device__ __forceinline__ void gs_wait(const uint64_t* flag, uint64_t seq, int* err) {
    const long long t0 = clock64();
    while (ld_acquire_sys(flag) < seq) {
        if (clock64() - t0 > PEER_SPIN_LIMIT) {
            *err = 1;
            break;
        }
    }
}

__global__ void __launch_bounds__(GS_THREADS) grad_sync_adam_kernel(const GradSyncArgs a) {
    pdl_entry();
    __shared__ uint64_t s_seq;
    __shared__ float s_lr;
    __shared__ int s_last;
    const int tid = threadIdx.x, W = a.world, R = a.rank;
    uint8_t* const ctl = reinterpret_cast<uint8_t*>(a.ctl[R]);
    if (tid == 0) {
        s_seq = *reinterpret_cast<volatile uint64_t*>(ctl) + 1;
        const double t = (double)(*a.step + 1);
        s_lr = (float)((double)a.lr * sqrt(1.0 - pow((double)a.b2, t)) / (1.0 - pow((double)a.b1, t)));
    }
    __syncthreads();
    const uint64_t seq = s_seq;
    if (W > 1) {
        // 1. ready: the local gradients were written by earlier launches of this stream; CTA 0 also publishes the loss scalars
        const int par = (int)(seq & 1);
        if (blockIdx.x == 0) {
            if (tid < a.nloss) reinterpret_cast<float*>(ctl + GS_LOSS)[par * GS_NLOSS + tid] = a.losses[tid];
            __threadfence_system();
            __syncthreads();
            if (tid < W) st_release_sys(reinterpret_cast<uint64_t*>(a.ctl[tid] + GS_READY) + R, seq);
        }
        if (tid < W) gs_wait(reinterpret_cast<const uint64_t*>(ctl + GS_READY) + tid, seq, a.err);
        __syncthreads();
        if (blockIdx.x == 0 && tid < a.nloss) {
            float s = 0.f;
            for (int q = 0; q < W; ++q) s += ld_relaxed_sys_f32(reinterpret_cast<const float*>(a.ctl[q] + GS_LOSS) + par * GS_NLOSS + tid);
            a.losses[tid] = s;
        }
        // 2 + 3. slice R: rank-ordered sum of every rank's values, written back to every rank
        const long lo = a.n4 * R / W, hi = a.n4 * (R + 1) / W;
        // (two vectors per thread and trip: the peer loads of both are in flight together -- the loop runs at NVLink latency)
        const long stride = (long)gridDim.x * GS_THREADS;
        for (long i = lo + (long)blockIdx.x * GS_THREADS + tid; i < hi; i += 2 * stride) {
            const long j = i + stride;
            const bool two = j < hi;
            float4 g0[PEER_MAX_WORLD], g1[PEER_MAX_WORLD];
#pragma unroll
            for (int q = 0; q < PEER_MAX_WORLD; ++q)
                if (q < W) {
                    g0[q] = ld_relaxed_sys_v4(reinterpret_cast<const float4*>(a.grad[q]) + i);
                    if (two) g1[q] = ld_relaxed_sys_v4(reinterpret_cast<const float4*>(a.grad[q]) + j);
                }
            float4 s0 = g0[0], s1 = two ? g1[0] : g0[0];
#pragma unroll
            for (int q = 1; q < PEER_MAX_WORLD; ++q)
                if (q < W) {
                    s0.x += g0[q].x, s0.y += g0[q].y, s0.z += g0[q].z, s0.w += g0[q].w;
                    if (two) s1.x += g1[q].x, s1.y += g1[q].y, s1.z += g1[q].z, s1.w += g1[q].w;
                }
#pragma unroll
            for (int q = 0; q < PEER_MAX_WORLD; ++q)
                if (q < W) {
                    st_relaxed_sys_v4(reinterpret_cast<float4*>(a.grad[q]) + i, s0);
                    if (two) st_relaxed_sys_v4(reinterpret_cast<float4*>(a.grad[q]) + j, s1);
                }
        }
        __threadfence_system();
        __syncthreads();
        // 4. done: the last CTA of this rank tells every peer; then wait for every peer's slice
        if (tid == 0) {
            const unsigned old = atomicAdd(reinterpret_cast<unsigned*>(ctl + 8), 1u);
            s_last = (old == gridDim.x - 1);
        }
        __syncthreads();
        if (s_last) {
            __threadfence_system();
            if (tid < W) st_release_sys(reinterpret_cast<uint64_t*>(a.ctl[tid] + GS_DONE) + R, seq);
            if (tid == 0) *reinterpret_cast<volatile unsigned*>(ctl + 8) = 0u;
        }
        if (tid < W) gs_wait(reinterpret_cast<const uint64_t*>(ctl + GS_DONE) + tid, seq, a.err);
        __syncthreads();
    }
    // 5. Adam (Keras 2.2 form, see adam_kernel) on the summed gradient; the sums were stored by peers: read at L2
    const float lr_t = s_lr;
    const float* g = reinterpret_cast<const float*>(a.grad[R]);
    const bool vec = !((reinterpret_cast<uintptr_t>(a.p) | reinterpret_cast<uintptr_t>(a.m) | reinterpret_cast<uintptr_t>(a.v)) & 15);
    const long nv = vec ? a.n / 4 : 0;
    for (long i = (long)blockIdx.x * GS_THREADS + tid; i < nv; i += (long)gridDim.x * GS_THREADS) {
        const float4 g4 = __ldcg(reinterpret_cast<const float4*>(g) + i);
        float4 m4 = reinterpret_cast<float4*>(a.m)[i], v4 = reinterpret_cast<float4*>(a.v)[i], p4 = reinterpret_cast<float4*>(a.p)[i];
        const float gg[4] = {g4.x, g4.y, g4.z, g4.w};
        float mm[4] = {m4.x, m4.y, m4.z, m4.w}, vv[4] = {v4.x, v4.y, v4.z, v4.w}, pp[4] = {p4.x, p4.y, p4.z, p4.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            mm[k] = a.b1 * mm[k] + (1.f - a.b1) * gg[k];
            vv[k] = a.b2 * vv[k] + (1.f - a.b2) * gg[k] * gg[k];
            pp[k] -= lr_t * mm[k] / (sqrtf(vv[k]) + a.eps);
        }
        reinterpret_cast<float4*>(a.m)[i] = make_float4(mm[0], mm[1], mm[2], mm[3]);
        reinterpret_cast<float4*>(a.v)[i] = make_float4(vv[0], vv[1], vv[2], vv[3]);
        reinterpret_cast<float4*>(a.p)[i] = make_float4(pp[0], pp[1], pp[2], pp[3]);
    }
    for (long i = nv * 4 + (long)blockIdx.x * GS_THREADS + tid; i < a.n; i += (long)gridDim.x * GS_THREADS) {
        const float gi = __ldcg(g + i);
        const float mi = a.b1 * a.m[i] + (1.f - a.b1) * gi;
        const float vi = a.b2 * a.v[i] + (1.f - a.b2) * gi * gi;
        a.m[i] = mi;
        a.v[i] = vi;
        a.p[i] -= lr_t * mi / (sqrtf(vi) + a.eps);
    }
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        const unsigned old = atomicAdd(reinterpret_cast<unsigned*>(ctl + 12), 1u);
        if (old == gridDim.x - 1) {
            *reinterpret_cast<volatile unsigned*>(ctl + 12) = 0u;
            *reinterpret_cast<volatile uint64_t*>(ctl) = seq;
            *a.step += 1;
        }
    }
}

}  // namespace

/* bytes of the symmetric control block of pctts_grad_sync_adam (zeroed once, then a barrier, before the first call) */
extern "C" long pctts_grad_sync_ctl_bytes(void) { return GS_BYTES; }

/* In place: grad (this rank's flat fp32 gradient buffer in symmetric memory, npad floats, npad % 4 == 0, zero padded beyond n)
 * <- sum over the ranks in rank order, then the Keras Adam update of p / m / v [n] with it and *step += 1 -- one launch.
 * grad_bases / ctl_bases: HOST arrays of `world` addresses (the buffers of rank r as mapped into this process).  world == 1:
 * plain Adam.  losses [nloss <= 8] (or NULL): scalars summed over the ranks in place on the way (the step's loss shares).
 * err: device int set to 1 if a peer did not show up. */
extern "C" int pctts_grad_sync_adam(const unsigned long long* grad_bases, const unsigned long long* ctl_bases, int rank, int world,
                                    float* p, float* m, float* v, long n, long npad, float lr, float beta1, float beta2, float eps,
                                    int* step, float* losses, int nloss, int* err, void* stream) {
    if (!grad_bases || !ctl_bases || !p || !m || !v || !step || !err || world < 1 || world > PEER_MAX_WORLD || rank < 0 || rank >= world ||
        n < 0 || npad < n || (npad & 3) || nloss < 0 || nloss > GS_NLOSS || (nloss > 0 && !losses))
        return PCTTS_EINVAL;
    GradSyncArgs a;
    for (int r = 0; r < PEER_MAX_WORLD; ++r) {
        a.grad[r] = r < world ? grad_bases[r] : 0;
        a.ctl[r] = r < world ? ctl_bases[r] : 0;
        if (r < world && ((a.grad[r] & 15) || (a.ctl[r] & 15))) return PCTTS_EINVAL;
    }
    a.p = p, a.m = m, a.v = v, a.step = step, a.n = n, a.n4 = npad / 4;
    a.lr = lr, a.b1 = beta1, a.b2 = beta2, a.eps = eps, a.rank = rank, a.world = world, a.err = err;
    a.losses = losses, a.nloss = nloss;
    // every CTA spins on peers' flags: the grid must be resident as a whole (one CTA per SM at most)
    long want = (n + GS_THREADS * 8 - 1) / (GS_THREADS * 8);
    int grid = (int)(want < 1 ? 1 : (want > 132 ? 132 : want));
    pctts_launch(grad_sync_adam_kernel, grid, GS_THREADS, 0, (cudaStream_t)stream, a);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}
