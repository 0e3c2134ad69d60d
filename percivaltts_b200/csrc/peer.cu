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
