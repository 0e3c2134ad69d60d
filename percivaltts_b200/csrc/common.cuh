// Shared device/host helpers for the percivaltts_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define PCTTS_LRELU_ALPHA 0.3f   // networktts.py:62,119,125 ; networks_critic.py:68

// status codes returned through the C ABI (0 = ok, >0 = cudaError_t, <0 = argument errors)
#define PCTTS_OK 0
#define PCTTS_EINVAL (-1)
#define PCTTS_EUNSUPPORTED (-2)

// every kernel launch of the library is counted (pctts_launch_count): evidence for bench.py's gpu_launches
void pctts_note_launch();

#define PCTTS_LAUNCH_CHECK()                          \
    do {                                              \
        pctts_note_launch();                          \
        cudaError_t e__ = cudaGetLastError();         \
        if (e__ != cudaSuccess) return (int)e__;      \
    } while (0)

#define PCTTS_CUDA(call)                              \
    do {                                              \
        cudaError_t e__ = (call);                     \
        if (e__ != cudaSuccess) return (int)e__;      \
    } while (0)

// Programmatic dependent launch: every kernel of the library can be launched with the programmatic-stream-serialization
// attribute and opens with pdl_entry(): `launch_dependents` lets the next kernel of the stream be scheduled (its CTAs become
// resident and run their prologue) while this one is still running; `wait` then blocks until the previous kernel has completed
// and flushed its writes.  No global memory is touched before the wait, so stream order semantics are unchanged; what
// would disappear is the launch latency + prologue between the ~200 dependent kernels of a step.  Opt-in (PCTTS_PDL=1):
// inside the captured step graphs the measured gain on B200 is nil (see api.cu).
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_entry() {
    pdl_trigger();
    pdl_wait();
}

int pctts_pdl_enabled();

template <typename... KA, typename... A>
static inline void pctts_launch(void (*kern)(KA...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, A&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pctts_pdl_enabled() ? 1 : 0;
    (void)cudaLaunchKernelEx(&cfg, kern, static_cast<KA>(args)...);   // the error, if any, is read by PCTTS_LAUNCH_CHECK
}

__device__ __forceinline__ float lrelu_f(float x) { return x > 0.f ? x : PCTTS_LRELU_ALPHA * x; }
// derivative of LeakyReLU written in terms of the *output* (sign-preserving, so y>0 <=> x>0)
__device__ __forceinline__ float dlrelu_from_out(float y) { return y > 0.f ? 1.f : PCTTS_LRELU_ALPHA; }
__device__ __forceinline__ float sigmoid_f(float x) { return 1.f / (1.f + __expf(-x)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// block-wide sum (all threads get the result); scratch must hold >= 32 doubles
__device__ __forceinline__ double block_sum_d(double v, double* scratch) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    v = warp_sum_d(v);
    __syncthreads();
    if (lane == 0) scratch[wid] = v;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    double r = (threadIdx.x < nw) ? scratch[threadIdx.x] : 0.0;
    if (wid == 0) {
        r = warp_sum_d(r);
        if (lane == 0) scratch[0] = r;
    }
    __syncthreads();
    r = scratch[0];
    return r;
}

// norm.cu: second stage of the deterministic column reductions, for producers that write the per-CTA partials themselves
int pctts_colreduce_slots(long rows, int C);
int pctts_colreduce_final_launch(void* workspace, long rows, int C, int nblk, cudaStream_t st);

static inline int pctts_div_up(long a, long b) { return (int)((a + b - 1) / b); }
