// Conv2D 5x5 'same', 4 -> 4 channels, NHWC, as an implicit GEMM on the 5th-generation tensor cores (tcgen05.mma, TMEM).
//
// Reference ops: the pCNN2D stack of the generator (networktts.py:122-126, modeltts_common.py:98) and the Conv2D stack of the
// critic (networks_critic.py:66-68), forward; with flipped kernels the data gradient; with a LeakyReLU' mask the
// forward-mode (tangent) sweep of the gradient penalty.
//
// Formulation.  One accumulator row = 4 consecutive output pixels of a frame row; its 16 columns are (pixel jo, channel co) --
// exactly the NHWC order of the 64 output bytes.  For one kernel row kt the 4 outputs read a window of 8 input pixels x 4
// channels = 32 values, so K = 5 x 32 = 160 and the weight operand is the block-Toeplitz matrix
//     Bmat[(kt, u, ci), (jo, co)] = w[kt, u - jo, ci, co]   (0 <= u - jo < 5, else 0),      u = 0..7.
// The activation operand is never expanded: the input tile lives in shared memory as bf16 NHWC pixel PAIRS (16 bytes =
// one UMMA core-matrix row), de-interleaved by pair parity into two arrays.  Accumulator rows r, r+1, .. are 4 pixels = 2 pairs
// apart, i.e. consecutive entries of one parity array: a K-major no-swizzle core matrix (8 rows x 16 B) is 8 consecutive
// entries, and the 4 chunks of a window alternate between the arrays.  The UMMA shared-memory descriptor expresses this with
// SBO = 128 B (next 8 rows) and LBO = (even array - odd array) + 16 B (next chunk): an im2col matrix that is 10x larger than
// the tile is read in place.
//
// Numerics (see gemm_tc.cu): fp32 operands are expanded into `PARTS` bf16 planes; the weight planes are concatenated along N
// ([B0 | B1 | B2], 48 columns), so one MMA per activation plane forms the products with every weight plane it needs
// (orders i + j <= PARTS-1).  hi.hi lands in its own 16 accumulator columns; the corrections share the two other blocks.
//
// Structure (persistent CTAs, one per SM; a tile = TT output frame rows of one sample):
//   warps 0-7  producers   fp32 global -> bf16 planes, pair arrays of a 2-stage ring, fence.proxy.async, mbarrier arrive
//   warp  8    MMA issuer  per 128-row block: 5 kt x 2 halves x PARTS tcgen05.mma; tcgen05.commit frees the stage / signals
//   warps 9-12 epilogue    tcgen05.ld (TMEM lane quarter = warp % 4), sum of the column blocks, bias / LeakyReLU / LeakyReLU'
//                          mask / accumulate, 64-byte stores; TMEM accumulators are double-buffered across tiles
#include "common.cuh"
#include "pctts.h"
#include "tc_common.cuh"
#include <cuda_bf16.h>

namespace {

using namespace tcx;

constexpr int C2_PROD = 256;             // producer threads (warps 0-7)
constexpr int C2_PW = C2_PROD / 32;      // index of the MMA warp; epilogue warps follow
constexpr int C2_THREADS = C2_PROD + 32 + 128;
constexpr int C2_UNR = 6;                // pixel pairs in flight per producer thread
constexpr int C2_MAXBLK = 5;             // 128-row blocks per tile (48 TMEM columns each, two buffers: 480 of 512 columns)
constexpr int C2_TMEM_COLS = 512;
constexpr int KT = 5, KW5 = 5;
constexpr int NMMA = KT * 2;             // (kt, half) k-steps of 16

struct C2Params {
    const float* x;
    const float* w;      // [5][5][4][4]
    const float* bias;
    const float* aux;
    float* y;
    int NB, T, F;
    int G;               // 4-pixel groups per padded row (Fp = 4G >= F + 2)
    int TT;              // output rows per tile
    int tiles_per_sample, ntiles;
    int nblk;            // 128-row blocks per tile
    int entries;         // 16-byte entries per parity array
    long xpitch, ypitch, auxpitch, xbatch, ybatch, auxbatch;
    int flags;
    long long* trace;    // developer probe (pctts_conv2d_tc_set_trace): clock64() stamps of CTA 0, 8 per tile
    const float *bn_mean, *bn_invstd, *bn_gamma, *bn_beta;   // optional: the input is a pre-BatchNorm map; BatchNorm + LeakyReLU
                                                             // are applied to every in-range pixel on load (fused bn_apply)
    int aux_prefetch;    // bit 0: the epilogue prefetches the tile's aux rows into L1 while the MMAs run; bit 1: the producers'
                         // input loads do not allocate in L1 (PCTTS_C2_AUX_PREFETCH, default 3)
    double* stats;       // optional: per-CTA partial [grid][2][4] of sum y, sum y^2 per output channel (fused BatchNorm statistics)
};

// streaming 16-byte load of the input map: read once per tile, kept out of L1 so that the epilogue's prefetched mask rows stay
__device__ __forceinline__ float4 ldg_stream4(const float* p, int no_alloc) {
    float4 v;
    if (no_alloc)
        asm("ld.global.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    else
        v = *reinterpret_cast<const float4*>(p);
    return v;
}

// shared-memory carve-up (bytes, 16-byte aligned):
//   Bmat   NMMA x [2 chunks][16*PARTS rows][16 B]
//   stage s, plane p: odd array [entries][16 B] | even array [entries][16 B]
template <int PARTS>
struct C2Smem {
    static constexpr uint32_t b_rows = 16 * PARTS;
    static constexpr uint32_t b_chunk = b_rows * 16;         // LBO of the weight operand
    static constexpr uint32_t b_mma = 2 * b_chunk;
    static constexpr uint32_t b_bytes = NMMA * b_mma;
    __host__ __device__ static uint32_t plane_bytes(int entries) { return 2u * (uint32_t)entries * 16u; }
    __host__ __device__ static uint32_t stage_bytes(int entries) { return PARTS * plane_bytes(entries); }
    __host__ __device__ static uint32_t total(int entries) { return b_bytes + 2 * stage_bytes(entries) + 16; }
};

template <int PARTS>
__global__ void __launch_bounds__(C2_THREADS, 1) conv2d_tc_kernel(const C2Params p) {
    pdl_trigger();   // the wait follows the barrier / TMEM set-up (nothing before it touches global memory)
    extern __shared__ __align__(16) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[2], empty_bar[2], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_slot;
    __shared__ double stat_red[4][8];
    using L = C2Smem<PARTS>;
    constexpr int BLKCOLS = 16 * PARTS;
    constexpr int BUFCOLS = C2_MAXBLK * BLKCOLS;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (p.trace && blockIdx.x == 0 && tid == 0) p.trace[31 * 8 + 0] = clock64();
    const uint32_t smem_base = (smem_u32(smem_raw) + 15u) & ~15u;
    uint8_t* const smem_al = smem_raw + (smem_base - smem_u32(smem_raw));
    const int G = p.G, TT = p.TT, entries = p.entries;
    const uint32_t plane_bytes = L::plane_bytes(entries), stage_bytes = L::stage_bytes(entries);
    const uint32_t even_off = (uint32_t)entries * 16u;

    if (tid == 0) {
        for (int s = 0; s < 2; ++s) {
            mbar_init(smem_u32(&full_bar[s]), C2_PROD);
            mbar_init(smem_u32(&empty_bar[s]), 1);
            mbar_init(smem_u32(&acc_full[s]), 1);
            mbar_init(smem_u32(&acc_empty[s]), 128);
        }
        fence_barrier_init();
    }
    if (warp == C2_PW) tmem_alloc<C2_TMEM_COLS>(smem_u32(&tmem_slot));

    // ---- producer set-up (all threads compute it; only the producer warps issue loads)
    const int row_pairs = 2 * G;
    const int npairs = (TT + 4) * row_pairs;
    constexpr int PASS = C2_UNR * C2_PROD;
    const int npass = (npairs + PASS - 1) / PASS;
    const int my_tiles = (p.ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    const int nitems = my_tiles * npass;
    // item k = (tile it = k / npass, pass k % npass): the loads of item k+1 are in flight while item k is converted
    auto issue = [&](int k, float4 (&va)[C2_UNR], float4 (&vb)[C2_UNR], uint32_t& inrange) {
        inrange = 0u;
        const int it = k / npass, pass = k - it * npass;
        const int ti = blockIdx.x + it * gridDim.x;
        const int n = ti / p.tiles_per_sample;
        const int t0 = (ti - n * p.tiles_per_sample) * TT;
        const float* xb = p.x + (long)n * p.xbatch;
#pragma unroll
        for (int u = 0; u < C2_UNR; ++u) {
            const int idx = pass * PASS + u * C2_PROD + tid;
            va[u] = vb[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (idx < npairs) {
                const int tt = idx / row_pairs, j = idx - tt * row_pairs;
                const int t = t0 - 2 + tt, f = 2 * j;
                if (t >= 0 && t < p.T && f < p.F) {
                    const float* src = xb + (long)t * p.xpitch + (long)f * 4;
                    va[u] = ldg_stream4(src, p.aux_prefetch & 2);
                    inrange |= 1u << (2 * u);
                    if (f + 1 < p.F) vb[u] = ldg_stream4(src + 4, p.aux_prefetch & 2), inrange |= 2u << (2 * u);
                }
            }
        }
    };
    if (p.trace && blockIdx.x == 0 && tid == 0) p.trace[31 * 8 + 3] = clock64();
    pdl_wait();
    __shared__ float wsm[KT * KW5 * 16];
    float4 ca[C2_UNR], cb[C2_UNR], na[C2_UNR], nb[C2_UNR];
    uint32_t cmask = 0u, nmask = 0u;
    // fused BatchNorm + LeakyReLU of the input (same expression as bn_apply_kernel); zero padding stays zero
    const bool bn_in = p.bn_mean != nullptr;
    float bm[4] = {0.f, 0.f, 0.f, 0.f}, bi[4] = {1.f, 1.f, 1.f, 1.f}, bg[4] = {1.f, 1.f, 1.f, 1.f}, bb[4] = {0.f, 0.f, 0.f, 0.f};
    if (warp < C2_PW) {
        if (nitems > 0) issue(0, ca, cb, cmask);      // the first tile's loads fly during the set-up below
        if (bn_in) {
#pragma unroll
            for (int c = 0; c < 4; ++c)
                bm[c] = __ldg(p.bn_mean + c), bi[c] = __ldg(p.bn_invstd + c), bg[c] = __ldg(p.bn_gamma + c), bb[c] = __ldg(p.bn_beta + c);
        }
    }
    auto bn_act = [&](float4& v) {
        v.x = lrelu_f((v.x - bm[0]) * bi[0] * bg[0] + bb[0]);
        v.y = lrelu_f((v.y - bm[1]) * bi[1] * bg[1] + bb[1]);
        v.z = lrelu_f((v.z - bm[2]) * bi[2] * bg[2] + bb[2]);
        v.w = lrelu_f((v.w - bm[3]) * bi[3] * bg[3] + bb[3]);
    };
    tc_fence_before();
    __syncthreads();      // barriers initialised, TMEM base published
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;
    if (p.trace && blockIdx.x == 0 && tid == 0) p.trace[31 * 8 + 4] = clock64();

    // The producers go straight to their first tile.  Meanwhile the MMA and epilogue warps (160 threads, idle until that tile is
    // filled) fetch the 400 kernel weights in ONE round trip and build the block-Toeplitz weight operand, all planes:
    // Bmat[mma j = (kt,h)][chunk c][row n = 16*plane + 4*jo + co][8 k], one 16-byte row per thread and plane.  (Built by all
    // 416 threads in front of the roles, the operand kept the first MMA 9.8 k cycles away from kernel entry: clock trace.)
    if (warp >= C2_PW) {
        constexpr int NB_THREADS = C2_THREADS - C2_PROD;
        const int bt = tid - C2_PROD;
        if (p.flags & PCTTS_EPI_WFLIP) {     // data-gradient kernel of the forward kernel w [5][5][co][ci]
            for (int i = bt; i < KT * KW5 * 16; i += NB_THREADS) {
                const int co = i & 3, ci = (i >> 2) & 3, tap = i >> 4;
                wsm[i] = __ldg(p.w + ((KT * KW5 - 1 - tap) * 4 + co) * 4 + ci);
            }
        } else {
            for (int i = bt; i < KT * KW5 * 16; i += NB_THREADS) wsm[i] = __ldg(p.w + i);
        }
        asm volatile("bar.sync 2, %0;" ::"n"(NB_THREADS) : "memory");
        for (int e = bt; e < NMMA * 2 * 16; e += NB_THREADS) {
            const int n16 = e & 15, c = (e >> 4) & 1, j = e >> 5;
            const int kt = j >> 1, h = j & 1, jo = n16 >> 2, co = n16 & 3;
            float v[8];
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {
                const int u = 4 * h + 2 * c + (kk >> 2), ci = kk & 3, kf = u - jo;
                v[kk] = (kf >= 0 && kf < KW5) ? wsm[((kt * KW5 + kf) * 4 + ci) * 4 + co] : 0.f;
            }
            uint8_t* const dst = smem_al + (size_t)j * L::b_mma + (size_t)c * L::b_chunk + (size_t)n16 * 16;
#pragma unroll
            for (int pl = 0; pl < PARTS; ++pl) {
                uint32_t w4[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const __nv_bfloat16 lo = __float2bfloat16_rn(v[2 * i]), hi = __float2bfloat16_rn(v[2 * i + 1]);
                    w4[i] = pack_bf16(lo, hi);
                    v[2 * i] -= __bfloat162float(lo), v[2 * i + 1] -= __bfloat162float(hi);
                }
                *reinterpret_cast<uint4*>(dst + (size_t)pl * 16 * 16) = make_uint4(w4[0], w4[1], w4[2], w4[3]);
            }
        }
        fence_proxy_async();
        asm volatile("bar.sync 2, %0;" ::"n"(NB_THREADS) : "memory");
        if (p.trace && blockIdx.x == 0 && tid == C2_PROD) p.trace[31 * 8 + 1] = clock64();
    }

    if (warp < C2_PW) {
        // ------------------------------------------------ producers ------------------------------------------------
        for (int k = 0; k < nitems; ++k) {
            const int it = k / npass, pass = k - it * npass;
            const int s = it & 1;
            const uint32_t ph = (uint32_t)(it >> 1) & 1u;
            uint8_t* const st = smem_al + L::b_bytes + (size_t)s * stage_bytes;
            if (k + 1 < nitems) issue(k + 1, na, nb, nmask);
            const bool tr = p.trace && blockIdx.x == 0 && tid == 0 && it < 32;
            if (pass == 0) {
                if (tr) p.trace[it * 8 + 0] = clock64();
                mbar_wait(smem_u32(&empty_bar[s]), ph ^ 1u);
                if (tr) p.trace[it * 8 + 1] = clock64();
                // pairs 0,1 (the zero prefix) and the two pairs after the last row
                if (tid < 4) {
                    const int m = tid < 2 ? tid : 2 + npairs + (tid - 2);
                    uint8_t* dst = st + ((m & 1) ? 0u : even_off) + (size_t)(m >> 1) * 16;
#pragma unroll
                    for (int pl = 0; pl < PARTS; ++pl) *reinterpret_cast<uint4*>(dst + (size_t)pl * plane_bytes) = make_uint4(0, 0, 0, 0);
                }
            }
#pragma unroll
            for (int u = 0; u < C2_UNR; ++u) {
                const int idx = pass * PASS + u * C2_PROD + tid;
                if (idx < npairs) {
                    if (bn_in) {
                        if (cmask & (1u << (2 * u))) bn_act(ca[u]);
                        if (cmask & (2u << (2 * u))) bn_act(cb[u]);
                    }
                    const float v[8] = {ca[u].x, ca[u].y, ca[u].z, ca[u].w, cb[u].x, cb[u].y, cb[u].z, cb[u].w};
                    uint4 e[PARTS];
                    split8<PARTS>(v, e);
                    const int m = 2 + idx;
                    uint8_t* dst = st + ((m & 1) ? 0u : even_off) + (size_t)(m >> 1) * 16;
#pragma unroll
                    for (int pl = 0; pl < PARTS; ++pl) *reinterpret_cast<uint4*>(dst + (size_t)pl * plane_bytes) = e[pl];
                }
            }
            if (pass == npass - 1) {
                fence_proxy_async();
                mbar_arrive(smem_u32(&full_bar[s]));
                if (tr) p.trace[it * 8 + 2] = clock64();
            }
#pragma unroll
            for (int u = 0; u < C2_UNR; ++u) ca[u] = na[u], cb[u] = nb[u];
            cmask = nmask;
        }
    } else if (warp == C2_PW) {
        // ------------------------------------------------ MMA issuer ------------------------------------------------
        uint32_t idesc[PARTS];
#pragma unroll
        for (int pa = 0; pa < PARTS; ++pa) idesc[pa] = umma_idesc_bf16(128, 16 * (PARTS - pa), 0, 0);
        const uint32_t a_lbo = even_off + 16u;
        const uint64_t bdesc0 = umma_desc_noswz(smem_base, L::b_chunk, 128u);
        int it = 0;
        for (int ti = blockIdx.x; ti < p.ntiles; ti += gridDim.x, ++it) {
            const int s = it & 1;
            const uint32_t ph = (uint32_t)(it >> 1) & 1u;
            mbar_wait(smem_u32(&acc_empty[s]), ph ^ 1u);
            mbar_wait(smem_u32(&full_bar[s]), ph);
            tc_fence_after();
            const bool tr = p.trace && blockIdx.x == 0 && lane == 0 && it < 32;
            if (tr) p.trace[it * 8 + 3] = clock64();
            const uint32_t st = smem_base + L::b_bytes + (uint32_t)s * stage_bytes;
            const uint64_t adesc0 = umma_desc_noswz(st, a_lbo, 128u);
            if (elect_one()) {
                for (int blk = 0; blk < p.nblk; ++blk) {
                    const uint32_t dcol = tmem_base + (uint32_t)(s * BUFCOLS + blk * BLKCOLS);
#pragma unroll
                    for (int kt = 0; kt < KT; ++kt) {
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            const uint32_t aoff = (uint32_t)(blk * 128 + kt * G + h);              // 16-byte units
                            const uint64_t bd = bdesc0 + (uint64_t)((kt * 2 + h) * (L::b_mma >> 4));
#pragma unroll
                            for (int pa = 0; pa < PARTS; ++pa) {
                                const uint64_t ad = adesc0 + (uint64_t)(aoff + (uint32_t)pa * (plane_bytes >> 4));
                                umma_bf16(dcol + 16u * pa, ad, bd, idesc[pa], (kt | h | pa) == 0 ? 0u : 1u);
                            }
                        }
                    }
                }
                umma_commit(smem_u32(&empty_bar[s]));
                umma_commit(smem_u32(&acc_full[s]));
            }
            __syncwarp();
            if (tr) p.trace[it * 8 + 4] = clock64();
        }
    } else {
        // ------------------------------------------------ epilogue ------------------------------------------------
        const int q = warp & 3;
        const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
        const int flags = p.flags;
        float bias4[4] = {0.f, 0.f, 0.f, 0.f};
        if (flags & PCTTS_EPI_BIAS) {
#pragma unroll
            for (int co = 0; co < 4; ++co) bias4[co] = __ldg(p.bias + co);
        }
        const int rows_valid = TT * G;
        double st1[4] = {0.0, 0.0, 0.0, 0.0}, st2[4] = {0.0, 0.0, 0.0, 0.0};
        int it = 0;
        for (int ti = blockIdx.x; ti < p.ntiles; ti += gridDim.x, ++it) {
            const int s = it & 1;
            const uint32_t ph = (uint32_t)(it >> 1) & 1u;
            const int n = ti / p.tiles_per_sample;
            const int t0 = (ti - n * p.tiles_per_sample) * TT;
            const bool tr = p.trace && blockIdx.x == 0 && tid == C2_PROD + 32 && it < 32;
            if (tr) p.trace[it * 8 + 5] = clock64();
            if ((flags & PCTTS_EPI_DLRELU_AUX) && (p.aux_prefetch & 1)) {
                // the mask rows this thread will need, requested while the tile's MMAs run: otherwise every 128-row block pays
                // one exposed global-load round trip between tcgen05.ld and its stores
                for (int blk = 0; blk < p.nblk; ++blk) {
                    const int r = blk * 128 + q * 32 + lane;
                    if (r < rows_valid) {
                        const int to = r / G, g = r - to * G;
                        if (t0 + to < p.T && 4 * g < p.F) {
                            const float* ax = p.aux + (long)n * p.auxbatch + (long)(t0 + to) * p.auxpitch + (long)g * 16;
                            asm volatile("prefetch.global.L1 [%0];" ::"l"(ax));
                        }
                    }
                }
            }
            mbar_wait(smem_u32(&acc_full[s]), ph);
            tc_fence_after();
            if (tr) p.trace[it * 8 + 6] = clock64();
            for (int blk = 0; blk < p.nblk; ++blk) {
                const uint32_t col = tmem_base + lane_addr + (uint32_t)(s * BUFCOLS + blk * BLKCOLS);
                uint32_t c0[16];
                float acc[16];
                tmem_ld_32x16(col, c0);
                if (PARTS > 1) {
                    uint32_t c1[16];
                    tmem_ld_32x16(col + 16, c1);
                    if (PARTS > 2) {
                        uint32_t c2[16];
                        tmem_ld_32x16(col + 32, c2);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            acc[i] = __uint_as_float(c0[i]) + (__uint_as_float(c1[i]) + __uint_as_float(c2[i]));
                    } else {
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 16; ++i) acc[i] = __uint_as_float(c0[i]) + __uint_as_float(c1[i]);
                    }
                } else {
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 16; ++i) acc[i] = __uint_as_float(c0[i]);
                }
                const int r = blk * 128 + q * 32 + lane;
                if (r < rows_valid) {
                    const int to = r / G, g = r - to * G;
                    const int t = t0 + to;
                    if (t < p.T) {
                        float* dst = p.y + (long)n * p.ybatch + (long)t * p.ypitch + (long)g * 16;
                        const float* ax = (flags & PCTTS_EPI_DLRELU_AUX)
                                              ? p.aux + (long)n * p.auxbatch + (long)t * p.auxpitch + (long)g * 16 : nullptr;
#pragma unroll
                        for (int jo = 0; jo < 4; ++jo) {
                            if (4 * g + jo < p.F) {
                                float o[4];
#pragma unroll
                                for (int co = 0; co < 4; ++co) {
                                    float v = acc[4 * jo + co] + bias4[co];
                                    if (flags & PCTTS_EPI_LRELU) v = lrelu_f(v);
                                    o[co] = v;
                                }
                                if (flags & PCTTS_EPI_DLRELU_AUX) {
                                    const float4 a4 = *reinterpret_cast<const float4*>(ax + 4 * jo);
                                    o[0] *= dlrelu_from_out(a4.x), o[1] *= dlrelu_from_out(a4.y);
                                    o[2] *= dlrelu_from_out(a4.z), o[3] *= dlrelu_from_out(a4.w);
                                }
                                if (flags & PCTTS_EPI_ACCUM) {
                                    const float4 d4 = *reinterpret_cast<const float4*>(dst + 4 * jo);
                                    o[0] += d4.x, o[1] += d4.y, o[2] += d4.z, o[3] += d4.w;
                                }
                                *reinterpret_cast<float4*>(dst + 4 * jo) = make_float4(o[0], o[1], o[2], o[3]);
                                if (p.stats) {
#pragma unroll
                                    for (int co = 0; co < 4; ++co) {
                                        st1[co] += (double)o[co];
                                        st2[co] += (double)o[co] * (double)o[co];
                                    }
                                }
                            }
                        }
                    }
                }
            }
            tc_fence_before();
            mbar_arrive(smem_u32(&acc_empty[s]));
            if (tr) p.trace[it * 8 + 7] = clock64();
        }
        if (p.stats) {      // fixed-order reduction: lanes (shuffle tree), then the four epilogue warps in index order
#pragma unroll
            for (int co = 0; co < 4; ++co) {
                const double a = warp_sum_d(st1[co]), b = warp_sum_d(st2[co]);
                if (lane == 0) stat_red[q][co] = a, stat_red[q][4 + co] = b;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (p.stats && tid < 8)
        p.stats[(long)blockIdx.x * 8 + tid] = ((stat_red[0][tid] + stat_red[1][tid]) + stat_red[2][tid]) + stat_red[3][tid];
    if (warp == C2_PW) tmem_dealloc<C2_TMEM_COLS>(tmem_base);
    if (p.trace && blockIdx.x == 0 && tid == 0) p.trace[31 * 8 + 2] = clock64();
}

long long* g_c2_trace = nullptr;

// rows per tile: fewest waves of the persistent grid, then fewest 128-row blocks per wave
int pick_tt(int NB, int T, int G, int* nblk_out) {
    int best_tt = 0, best_blk = 0;
    double best = 1e30;
    for (int tt = 1; tt <= T && tt <= 64; ++tt) {
        const int nblk = (tt * G + 127) / 128;
        if (nblk > C2_MAXBLK) break;
        const long tiles = (long)NB * ((T + tt - 1) / tt);
        const long waves = (tiles + 147) / 148;
        // in units of one 128-row block of MMAs (~1 k cycles, clock trace): per tile the blocks plus ~0.6 of hand-over gaps; per
        // CTA once the first tile's fill and the last tile's drain, which grow with the tile (~0.5 + ~0.75 per block) -- on the
        // B-sized maps (2 tiles per CTA) they weigh as much as a whole tile, so smaller tiles win there
        static const int old_cost = getenv("PCTTS_C2_OLD_COST") ? atoi(getenv("PCTTS_C2_OLD_COST")) : 0;
        const double cost = old_cost ? (double)waves * (nblk + 0.3 + 0.02 * (tt + 4))
                                     : (double)waves * (nblk + 0.6) + 1.25 * nblk - 1e-3 * tt;
        if (cost <= best) best = cost, best_tt = tt, best_blk = nblk;
    }
    *nblk_out = best_blk;
    return best_tt;
}

template <int PARTS>
int launch_c2(C2Params& p, cudaStream_t st) {
    const size_t smem = C2Smem<PARTS>::total(p.entries);
    if (smem > 226 * 1024) return PCTTS_EUNSUPPORTED;      // + the static barriers stays under the 227 KB opt-in limit
    static size_t attr_smem = 0;
    if (smem > attr_smem) {
        PCTTS_CUDA(cudaFuncSetAttribute(conv2d_tc_kernel<PARTS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_smem = smem;
    }
    const int grid = p.ntiles < 148 ? p.ntiles : 148;
    pctts_launch(conv2d_tc_kernel<PARTS>, grid, C2_THREADS, smem, st, p);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

}  // namespace

/* developer probe: device buffer of 32*8 int64 receiving clock64() stamps of CTA 0 (per tile: producer start / stage free /
 * stage filled, MMA operands ready / issued, epilogue start / accumulator ready / stored); NULL switches it off */
extern "C" int pctts_conv2d_tc_set_trace(void* buf) {
    g_c2_trace = (long long*)buf;
    return PCTTS_OK;
}

extern "C" int pctts_conv2d_tc_supported(int F, int CIN, int COUT, int KH, int KW, int pad_t, int pad_f) {
    const int G = (F + 2 + 3) / 4;
    return (CIN == 4 && COUT == 4 && KH == 5 && KW == 5 && pad_t == 2 && pad_f == 2 && F >= 1 && G <= 128) ? 1 : 0;
}

static int conv2d_tc_run(const float* x, const float* w, const float* bias, const float* aux, float* y, int NB, int T,
                         int F, int CIN, int COUT, int KH, int KW, int pad_t, int pad_f, long xpitch, long ypitch,
                         long auxpitch, long xbatch, long ybatch, long auxbatch, int flags, int nterms, const float* const* in_bn,
                         void* stats_workspace, int final_stage, void* stream) {
    if (!x || !w || !y || NB < 0 || T < 0 || F <= 0) return PCTTS_EINVAL;
    if (!pctts_conv2d_tc_supported(F, CIN, COUT, KH, KW, pad_t, pad_f)) return PCTTS_EUNSUPPORTED;
    if (nterms != 1 && nterms != 3 && nterms != 6) return PCTTS_EINVAL;
    if ((flags & PCTTS_EPI_BIAS) && !bias) return PCTTS_EINVAL;
    if ((flags & PCTTS_EPI_DLRELU_AUX) && !aux) return PCTTS_EINVAL;
    if (flags & ~(PCTTS_EPI_BIAS | PCTTS_EPI_LRELU | PCTTS_EPI_DLRELU_AUX | PCTTS_EPI_ACCUM | PCTTS_EPI_WFLIP)) return PCTTS_EUNSUPPORTED;
    // float4 accesses: 16-byte aligned bases and row / sample pitches
    if ((((uintptr_t)x | (uintptr_t)y | (uintptr_t)aux) & 15) || ((xpitch | ypitch | xbatch | ybatch) & 3) ||
        (aux && ((auxpitch | auxbatch) & 3)))
        return PCTTS_EUNSUPPORTED;
    if (NB == 0 || T == 0) return PCTTS_OK;
    C2Params p;
    p.x = x, p.w = w, p.bias = bias, p.aux = aux, p.y = y;
    p.NB = NB, p.T = T, p.F = F;
    p.G = (F + 2 + 3) / 4;
    p.TT = pick_tt(NB, T, p.G, &p.nblk);
    if (p.TT < 1) return PCTTS_EUNSUPPORTED;
    p.tiles_per_sample = (T + p.TT - 1) / p.TT;
    p.ntiles = NB * p.tiles_per_sample;
    p.entries = p.nblk * 128 + 4 * p.G + 8;
    p.xpitch = xpitch, p.ypitch = ypitch, p.auxpitch = auxpitch, p.xbatch = xbatch, p.ybatch = ybatch, p.auxbatch = auxbatch;
    p.flags = flags;
    p.trace = g_c2_trace;
    static const int aux_pf = getenv("PCTTS_C2_AUX_PREFETCH") ? atoi(getenv("PCTTS_C2_AUX_PREFETCH")) : 3;
    p.aux_prefetch = aux_pf;
    p.bn_mean = in_bn ? in_bn[0] : nullptr, p.bn_invstd = in_bn ? in_bn[1] : nullptr;
    p.bn_gamma = in_bn ? in_bn[2] : nullptr, p.bn_beta = in_bn ? in_bn[3] : nullptr;
    p.stats = (double*)stats_workspace;      // the partials sit at the base of a column-reduction workspace (norm.cu)
    cudaStream_t st = (cudaStream_t)stream;
    int rc;
    if (nterms == 6) rc = launch_c2<3>(p, st);
    else if (nterms == 3) rc = launch_c2<2>(p, st);
    else rc = launch_c2<1>(p, st);
    if (rc != PCTTS_OK || !stats_workspace || !final_stage) return rc;
    return pctts_colreduce_final_launch(stats_workspace, (long)NB * T * F, 4, p.ntiles < 148 ? p.ntiles : 148, st);
}

extern "C" int pctts_conv2d_tc_fwd(const float* x, const float* w, const float* bias, const float* aux, float* y, int NB, int T,
                                   int F, int CIN, int COUT, int KH, int KW, int pad_t, int pad_f, long xpitch, long ypitch,
                                   long auxpitch, long xbatch, long ybatch, long auxbatch, int flags, int nterms, void* stream) {
    return conv2d_tc_run(x, w, bias, aux, y, NB, T, F, CIN, COUT, KH, KW, pad_t, pad_f, xpitch, ypitch, auxpitch, xbatch, ybatch,
                         auxbatch, flags, nterms, nullptr, nullptr, 1, stream);
}

// The same convolution fused with the BatchNormalization layers around it (networktts.py:122-126, pCNN2D = Conv2D + BN + LReLU):
//  * in_mean / in_invstd / in_gamma / in_beta (all four or none, 4 floats each): x is the PREVIOUS layer's pre-BatchNorm map and
//    LeakyReLU(BN(x)) is formed on load, so the activation map is never written (forward-only uses: the frozen generator);
//  * workspace (or NULL): a column-reduction workspace for rows = NB*T*F, C = 4 (pctts_colreduce_workspace_bytes) that receives
//    sums = [sum y, sum y^2] of the output exactly as pctts_bn_stats(y) would, without re-reading y.
//    final_stage = 0 leaves the pctts_conv2d_tc_stat_partials(NB, T, F) per-CTA partials at the base of the workspace for
//    pctts_bn_finalize_partials (second stage + finalize in one launch) instead of running the second stage here.
// PCTTS_EUNSUPPORTED when a workspace is given but has fewer than 148 partial slots (small maps).
/* number of per-CTA statistic partials pctts_conv2d_tc_fwd_fused writes for a map of this size (host call) */
extern "C" int pctts_conv2d_tc_stat_partials(int NB, int T, int F) {
    if (NB <= 0 || T <= 0 || F <= 0) return 0;
    const int G = (F + 2 + 3) / 4;
    int nblk = 0;
    const int TT = pick_tt(NB, T, G, &nblk);
    if (TT < 1) return 0;
    const long ntiles = (long)NB * ((T + TT - 1) / TT);
    return ntiles < 148 ? (int)ntiles : 148;
}

extern "C" int pctts_conv2d_tc_fwd_fused(const float* x, const float* w, const float* bias, const float* aux, float* y, int NB,
                                         int T, int F, int CIN, int COUT, int KH, int KW, int pad_t, int pad_f, long xpitch,
                                         long ypitch, long auxpitch, long xbatch, long ybatch, long auxbatch, int flags, int nterms,
                                         const float* in_mean, const float* in_invstd, const float* in_gamma, const float* in_beta,
                                         void* workspace, int final_stage, void* stream) {
    const int nbn = (in_mean != nullptr) + (in_invstd != nullptr) + (in_gamma != nullptr) + (in_beta != nullptr);
    if (nbn != 0 && nbn != 4) return PCTTS_EINVAL;
    if (workspace && pctts_colreduce_slots((long)NB * T * F, 4) < 148) return PCTTS_EUNSUPPORTED;
    if (workspace && (NB == 0 || T == 0)) return PCTTS_EUNSUPPORTED;
    const float* in_bn[4] = {in_mean, in_invstd, in_gamma, in_beta};
    return conv2d_tc_run(x, w, bias, aux, y, NB, T, F, CIN, COUT, KH, KW, pad_t, pad_f, xpitch, ypitch, auxpitch, xbatch, ybatch,
                         auxbatch, flags, nterms, nbn ? in_bn : nullptr, workspace, final_stage, stream);
}
