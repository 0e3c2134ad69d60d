// A CHAIN of Conv2D 5x5 'same' layers over (time x frequency) maps with <= 4 channels in ONE launch: the map tile of a CTA
// stays in shared memory as bf16 operand planes from layer to layer, every layer is an implicit GEMM on the 5th-generation
// tensor cores (tcgen05.mma, accumulators in TMEM), and only what later passes need goes to HBM (the fp32 output of each layer).
//
// Reference ops: the critic's Conv2D stack, which has NO BatchNormalization between its layers (networks_critic.py:50,66-68:
// Conv2D(4,[5,5],bias) + LeakyReLU(0.3), 8 times), in its three uses on the WGAN-GP path:
//   forward           a_i = LeakyReLU(conv(a_{i-1}, w_i) + b_i)                                    (networks_critic.py:66-68)
//   data gradient     da_{i-1} = conv(da_i, flip(w_i)) * LeakyReLU'(a_{i-1})                        (K.gradients, optimizertts_wgan.py:57)
//   tangent sweep     t_i = conv(t_{i-1}, w_i) * LeakyReLU'(a_i), written over a_i                 (double backprop of the penalty, :57-68)
// Per layer these are HBM / launch-latency bound (25 flop/B): run layer by layer, each 8.26 MB map makes a
// round trip through L2 / HBM per layer and every layer pays a launch, an operand-build set-up, a pipeline fill and a drain.
// Here a sample is owned by a thread-block CLUSTER: CTA r of the cluster holds frames [r*Fc, (r+1)*Fc) plus a 2-frame halo on
// either side; after each layer the CTAs exchange only the halo frames (through the layer's fp32 output in L2, ordered by one
// cluster barrier), so nothing is recomputed and nothing but the halo is re-read.
//
// Formulation of one layer: exactly that of conv2d_tc.cu (accumulator row = 4 output pixels x 4 channels, K = 5 kernel rows x
// (8 input pixels x 4 channels), block-Toeplitz weight operand, activation operand read in place from de-interleaved pixel-pair
// arrays through a no-swizzle K-major UMMA descriptor; fp32 operands expanded into PARTS bf16 planes, products of order < PARTS).
//
// In-place layer hand-over.  The epilogue writes the next layer's operand planes into the SAME arrays the current layer's MMAs
// are still reading.  Even layers walk the 128-row blocks upwards and write their output D = 2G+2 entries BELOW the input
// origin; odd layers walk downwards and write D entries ABOVE: a block's output then only lands on entries that no later block of
// the layer reads (proof in DESIGN.md section 4.10), and the two origins alternate, so the slack is D entries, not L*D.
//
// Structure: 416 threads.  warps 0-11 = workers in three sets (initial tile load, block-Toeplitz build of the NEXT layer's weights, epilogue:
// tcgen05.ld -> bias / LeakyReLU / mask -> fp32 store + bf16 planes of the next layer, halo exchange); warp 12 = MMA issuer.
// TMEM holds 10 accumulator slots of one 128-row block each (48 columns at 3 planes), handed over one by one (an mbarrier pair per
// slot): the MMAs of up to 10 blocks run ahead of the epilogue, set s drains blocks s, s+3, .., so only the last block's epilogue
// is exposed at the end of a layer (clock trace: with two 5-block buffers the tail was ~5.5 k of a 27 k-cycle layer).
#include "common.cuh"
#include "pctts.h"
#include "tc_common.cuh"
#include <cuda_bf16.h>

namespace {

using namespace tcx;

constexpr int CH_MAXL = PCTTS_CHAIN_MAX_LAYERS;
constexpr int CH_SETS = 3;                   // worker sets: set s takes the blocks at positions s, s + 3, s + 6, .. of a layer
constexpr int CH_WORK = CH_SETS * 128;       // worker threads (warps 0-11; warp w reads TMEM lane quarter w & 3)
constexpr int CH_MW = CH_WORK / 32;          // index of the MMA warp
constexpr int CH_THREADS = CH_WORK + 32;
constexpr int CH_TMEM_COLS = 512;
constexpr int CH_MAXSLOT = 16;               // accumulator slots (one 128-row block each) in TMEM: 512 / (16 * PARTS), at most 16
constexpr int CKT = 5, CKW = 5, CNMMA = CKT * 2;

struct ChainLayer {
    const float* w;        // [5][5][cin][cout] (keras layout); flip: the data-gradient kernel is formed on the fly
    const float* bias;     // [cout] or NULL
    const float* aux;      // LeakyReLU' mask source [NB,T,F,4] or NULL
    float* y;              // fp32 output [NB,T,F,cout]
    long ybatch, ypitch, auxbatch, auxpitch;
    int cin, cout;         // channels of the convolution as executed (flip: already swapped)
    int flags;             // PCTTS_EPI_BIAS | LRELU | DLRELU_AUX | ACCUM
    int flip;              // 1: w is the FORWARD kernel [5][5][cout][cin]; use w[4-kt][4-kf][co][ci]
};

struct ChainParams {
    const float* x;
    long xbatch, xpitch;
    int cin0;
    int NB, T, F, G;
    int NC, Fc;            // CTAs per sample (cluster size), frames per CTA
    int nblk;              // 128-row blocks per CTA
    int E;                 // 16-byte entries per parity array
    int D;                 // slide between the two origins, in entries
    int nlayers;
    long long* trace;      // developer probe (pctts_conv2d_chain_set_trace): clock64() stamps of CTA 0, 16 per layer
    int debug;             // developer probe (PCTTS_CHAIN_DEBUG): 1 = epilogue without the next layer's operand stores, 2 = empty epilogue (results wrong)
    ChainLayer L[CH_MAXL];
};

template <int PARTS>
struct ChSmem {
    static constexpr uint32_t b_rows = 16 * PARTS;
    static constexpr uint32_t b_chunk = b_rows * 16;
    static constexpr uint32_t b_mma = 2 * b_chunk;
    static constexpr uint32_t b_bytes = CNMMA * b_mma;
    __host__ __device__ static uint32_t plane_bytes(int E) { return 2u * (uint32_t)E * 16u; }
    __host__ __device__ static uint32_t total(int E) { return 2 * b_bytes + PARTS * plane_bytes(E) + 16; }
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// halo frames come from the neighbour CTA's global stores of this very kernel: read at L2, never through this SM's L1 (a line
// that straddles the frame boundary may sit there with the previous layer's bytes)
__device__ __forceinline__ float4 ldg_cg4(const float* p) {
    float4 v;
    asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}

template <int PARTS>
__global__ void __launch_bounds__(CH_THREADS, 1) conv2d_chain_kernel(const ChainParams p) {
    pdl_entry();
    extern __shared__ __align__(16) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t acc_full[CH_MAXSLOT], acc_empty[CH_MAXSLOT];
    __shared__ uint32_t tmem_slot;
    __shared__ float wsm[CKT * CKW * 16];
    using S = ChSmem<PARTS>;
    constexpr int BLKCOLS = 16 * PARTS;
    constexpr int NSLOT = (CH_TMEM_COLS / BLKCOLS) < CH_MAXSLOT ? (CH_TMEM_COLS / BLKCOLS) : CH_MAXSLOT;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t smem_base = (smem_u32(smem_raw) + 15u) & ~15u;
    uint8_t* const smem_al = smem_raw + (smem_base - smem_u32(smem_raw));
    const int G = p.G, E = p.E, D = p.D;
    const uint32_t plane_bytes = S::plane_bytes(E);
    const uint32_t even_off = (uint32_t)E * 16u;
    uint8_t* const act = smem_al + 2 * S::b_bytes;            // plane pl: act + pl*plane_bytes; odd array, then even array
    const uint32_t act_u32 = smem_base + 2 * S::b_bytes;

    const int rank = p.NC > 1 ? (int)cluster_ctarank() : 0;
    const int n = (int)blockIdx.x / p.NC;
    const int t0 = rank * p.Fc;
    const int nfr = min(p.Fc, p.T - t0);                        // frames owned by this CTA (>= 1 by construction)
    const int rows_valid = nfr * G;
    const int nblk = (rows_valid + 127) / 128;
    const int row_pairs = 2 * G;

    if (tid == 0) {
        for (int s = 0; s < NSLOT; ++s) {
            mbar_init(smem_u32(&acc_full[s]), 1);
            mbar_init(smem_u32(&acc_empty[s]), 128);      // the four warps (TMEM lane quarters) of the set that drains the block
        }
        fence_barrier_init();
    }
    if (warp == CH_MW) tmem_alloc<CH_TMEM_COLS>(smem_u32(&tmem_slot));

    // entry (16-byte unit inside a parity array) and array of the pixel pair j of tile frame tt, for the layout with origin O
    auto pair_dst = [&](int O, int tt, int j) -> uint8_t* {
        const int m = 2 + tt * row_pairs + j;
        return act + ((m & 1) ? 0u : even_off) + (size_t)(O + (m >> 1)) * 16;
    };
    auto store_planes = [&](uint8_t* dst, const float (&v)[8]) {
        uint4 e[PARTS];
        split8<PARTS>(v, e);
#pragma unroll
        for (int pl = 0; pl < PARTS; ++pl) *reinterpret_cast<uint4*>(dst + (size_t)pl * plane_bytes) = e[pl];
    };
    auto zero_planes = [&](uint8_t* dst) {
#pragma unroll
        for (int pl = 0; pl < PARTS; ++pl) *reinterpret_cast<uint4*>(dst + (size_t)pl * plane_bytes) = make_uint4(0, 0, 0, 0);
    };
    // the zero prefix (pairs m = 0, 1) and the two pairs after the last tile frame of a layout
    auto zero_edges = [&](int O, int wt) {
        if (wt < 4) {
            const int m = wt < 2 ? wt : 2 + (nfr + 4) * row_pairs + (wt - 2);
            zero_planes(act + ((m & 1) ? 0u : even_off) + (size_t)(O + (m >> 1)) * 16);
        }
    };
    // block-Toeplitz weight operand of layer l (all planes) into Bmat buffer l & 1; wsm staged by the same threads
    auto build_bmat = [&](int l) {
        const ChainLayer& Ly = p.L[l];
        const int cin = Ly.cin, cout = Ly.cout;
        for (int i = tid; i < CKT * CKW * 16; i += CH_WORK) {
            const int co = i & 3, ci = (i >> 2) & 3, tap = i >> 4;
            float v = 0.f;
            if (ci < cin && co < cout) {
                if (Ly.flip) {      // data-gradient kernel of the forward kernel w [5][5][cout][cin]
                    const int kt = tap / CKW, kf = tap - kt * CKW;
                    v = __ldg(Ly.w + ((size_t)((CKT - 1 - kt) * CKW + (CKW - 1 - kf)) * cout + co) * cin + ci);
                } else {
                    v = __ldg(Ly.w + ((size_t)tap * cin + ci) * cout + co);
                }
            }
            wsm[i] = v;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(CH_WORK) : "memory");
        uint8_t* const bm = smem_al + (size_t)(l & 1) * S::b_bytes;
        for (int e = tid; e < CNMMA * 2 * 16; e += CH_WORK) {
            const int n16 = e & 15, c = (e >> 4) & 1, j = e >> 5;
            const int kt = j >> 1, h = j & 1, jo = n16 >> 2, co = n16 & 3;
            float v[8];
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {
                const int u = 4 * h + 2 * c + (kk >> 2), ci = kk & 3, kf = u - jo;
                v[kk] = (kf >= 0 && kf < CKW) ? wsm[((kt * CKW + kf) * 4 + ci) * 4 + co] : 0.f;
            }
            uint8_t* const dst = bm + (size_t)j * S::b_mma + (size_t)c * S::b_chunk + (size_t)n16 * 16;
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
        asm volatile("bar.sync 1, %0;" ::"n"(CH_WORK) : "memory");      // wsm may be restaged by the next build
    };

    // ---- set-up by the workers: the input tile (frames t0-2 .. t0+nfr+1) as operand planes at origin D, Bmat of layer 0 ------
    if (warp < CH_MW) {
        const float* xb = p.x + (long)n * p.xbatch;
        const int npairs = (nfr + 4) * row_pairs;
        for (int idx = tid; idx < npairs; idx += CH_WORK) {
            const int tt = idx / row_pairs, j = idx - tt * row_pairs;
            const int t = t0 - 2 + tt, f = 2 * j;
            float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            if (t >= 0 && t < p.T && f < p.F) {
                if (p.cin0 == 4) {
                    const float* src = xb + (long)t * p.xpitch + (long)f * 4;
                    const float4 a = *reinterpret_cast<const float4*>(src);
                    v[0] = a.x, v[1] = a.y, v[2] = a.z, v[3] = a.w;
                    if (f + 1 < p.F) {
                        const float4 b = *reinterpret_cast<const float4*>(src + 4);
                        v[4] = b.x, v[5] = b.y, v[6] = b.z, v[7] = b.w;
                    }
                } else {
                    const float* src = xb + (long)t * p.xpitch + f;
                    v[0] = __ldg(src);
                    if (f + 1 < p.F) v[4] = __ldg(src + 1);
                }
            }
            store_planes(pair_dst(D, tt, j), v);
        }
        zero_edges(D, tid);
        build_bmat(0);
        fence_proxy_async();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    // blocks are numbered through the whole launch: block `pos` of layer l is number bc = l * nblk + pos, its accumulator slot
    // bc % NSLOT, its mbarrier phase (bc / NSLOT) & 1 -- the same arithmetic in every role
    for (int l = 0; l < p.nlayers; ++l) {
        const ChainLayer& Ly = p.L[l];
        const int Oin = (l & 1) ? 0 : D, Oout = (l & 1) ? D : 0;
        const bool asc = (l & 1) == 0;
        const bool last = l + 1 == p.nlayers;
        const bool trm = p.trace && blockIdx.x == 0 && tid == CH_WORK, trw = p.trace && blockIdx.x == 0 && tid == 0;
        long long* const tr = p.trace ? p.trace + l * 16 : nullptr;
        if (warp == CH_MW) {
            // ------------------------------------------------ MMA issuer ------------------------------------------------
            if (trm) tr[0] = clock64();
            uint32_t idesc[PARTS];
#pragma unroll
            for (int pa = 0; pa < PARTS; ++pa) idesc[pa] = umma_idesc_bf16(128, 16 * (PARTS - pa), 0, 0);
            const uint32_t a_lbo = even_off + 16u;
            const uint64_t bdesc0 = umma_desc_noswz(smem_base + (uint32_t)(l & 1) * S::b_bytes, S::b_chunk, 128u);
            const uint64_t adesc0 = umma_desc_noswz(act_u32 + (uint32_t)Oin * 16u, a_lbo, 128u);
            for (int pos = 0; pos < nblk; ++pos) {
                const uint32_t bc = (uint32_t)(l * nblk + pos);
                const int s = (int)(bc % NSLOT);
                const uint32_t ph = (bc / NSLOT) & 1u;
                mbar_wait(smem_u32(&acc_empty[s]), ph ^ 1u);
                tc_fence_after();
                if (elect_one()) {
                    const int blk = asc ? pos : nblk - 1 - pos;
                    const uint32_t dcol = tmem_base + (uint32_t)(s * BLKCOLS);
#pragma unroll
                    for (int kt = 0; kt < CKT; ++kt) {
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            const uint32_t aoff = (uint32_t)(blk * 128 + kt * G + h);
                            const uint64_t bd = bdesc0 + (uint64_t)((kt * 2 + h) * (S::b_mma >> 4));
#pragma unroll
                            for (int pa = 0; pa < PARTS; ++pa) {
                                const uint64_t ad = adesc0 + (uint64_t)(aoff + (uint32_t)pa * (plane_bytes >> 4));
                                umma_bf16(dcol + 16u * pa, ad, bd, idesc[pa], (kt | h | pa) == 0 ? 0u : 1u);
                            }
                        }
                    }
                    umma_commit(smem_u32(&acc_full[s]));
                }
                __syncwarp();
            }
            if (trm) tr[1] = clock64();
        } else {
            // ------------------------------------------------ workers ------------------------------------------------
            if (!last) {        // the next layer's weight operand, while this layer's first chunk of MMAs runs
                build_bmat(l + 1);
                fence_proxy_async();
            }
            if (trw) tr[4] = clock64();
            const int q = warp & 3, set = warp >> 2;
            const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
            const int flags = Ly.flags;
            const bool has_aux = (flags & PCTTS_EPI_DLRELU_AUX) != 0;
            float bias4[4] = {0.f, 0.f, 0.f, 0.f};
            if (flags & PCTTS_EPI_BIAS) {
#pragma unroll
                for (int co = 0; co < 4; ++co)
                    if (co < Ly.cout) bias4[co] = __ldg(Ly.bias + co);
            }
            const float* const auxb = has_aux ? Ly.aux + (long)n * Ly.auxbatch : nullptr;
            float* const yb = Ly.y + (long)n * Ly.ybatch;
            // the mask rows of the block this thread handles next are fetched one block ahead (registers)
            float4 ax[4];
            auto fetch_aux = [&](int pos) {
                if (!has_aux || pos >= nblk) return;
                const int blk = asc ? pos : nblk - 1 - pos;
                const int r = blk * 128 + q * 32 + lane;
                if (r < rows_valid) {
                    const int to = r / G, g = r - to * G;
                    const float* a = auxb + (long)(t0 + to) * Ly.auxpitch + (long)g * 16;
#pragma unroll
                    for (int jo = 0; jo < 4; ++jo)
                        ax[jo] = (4 * g + jo < p.F) ? *reinterpret_cast<const float4*>(a + 4 * jo) : make_float4(1.f, 1.f, 1.f, 1.f);
                }
            };
            // positions handled by this worker set: set, set + CH_SETS, ..
            fetch_aux(set);
            for (int pos = set; pos < nblk; pos += CH_SETS) {
                {
                    const uint32_t bc = (uint32_t)(l * nblk + pos);
                    const int s = (int)(bc % NSLOT);
                    const uint32_t ph = (bc / NSLOT) & 1u;
                    mbar_wait(smem_u32(&acc_full[s]), ph);
                    tc_fence_after();
                    if (p.debug & 2) {
                        tc_fence_before();
                        mbar_arrive(smem_u32(&acc_empty[s]));
                        continue;
                    }
                    const int blk = asc ? pos : nblk - 1 - pos;
                    const uint32_t col = tmem_base + lane_addr + (uint32_t)(s * BLKCOLS);
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
                    float4 cur[4];
#pragma unroll
                    for (int jo = 0; jo < 4; ++jo) cur[jo] = ax[jo];
                    fetch_aux(pos + CH_SETS);          // the mask rows of this set's next block
                    const int r = blk * 128 + q * 32 + lane;
                    if (r < rows_valid) {
                        const int to = r / G, g = r - to * G;
                        const int t = t0 + to;
                        float o[16];
#pragma unroll
                        for (int jo = 0; jo < 4; ++jo) {
                            const bool pix = 4 * g + jo < p.F;
                            const float m4[4] = {cur[jo].x, cur[jo].y, cur[jo].z, cur[jo].w};
#pragma unroll
                            for (int co = 0; co < 4; ++co) {
                                float v = acc[4 * jo + co] + bias4[co];
                                if (flags & PCTTS_EPI_LRELU) v = lrelu_f(v);
                                if (has_aux) v *= dlrelu_from_out(m4[co]);
                                o[4 * jo + co] = pix ? v : 0.f;
                            }
                        }
                        if (Ly.cout == 4) {
                            float* dst = yb + (long)t * Ly.ypitch + (long)g * 16;
#pragma unroll
                            for (int jo = 0; jo < 4; ++jo) {
                                if (4 * g + jo < p.F) {
                                    float4 v4 = make_float4(o[4 * jo], o[4 * jo + 1], o[4 * jo + 2], o[4 * jo + 3]);
                                    if (flags & PCTTS_EPI_ACCUM) {
                                        const float4 d4 = *reinterpret_cast<const float4*>(dst + 4 * jo);
                                        v4.x += d4.x, v4.y += d4.y, v4.z += d4.z, v4.w += d4.w;
                                    }
                                    *reinterpret_cast<float4*>(dst + 4 * jo) = v4;
                                }
                            }
                        } else {      // single output channel (the input gradient of the stack): scalar stores, any pitch
                            float* dst = yb + (long)t * Ly.ypitch + (long)g * 4;
#pragma unroll
                            for (int jo = 0; jo < 4; ++jo) {
                                if (4 * g + jo < p.F) {
                                    float v = o[4 * jo];
                                    if (flags & PCTTS_EPI_ACCUM) v += dst[jo];
                                    dst[jo] = v;
                                }
                            }
                        }
                        if (!last && !(p.debug & 1)) {
                            // this row as input of the next layer: tile frame to + 2, pairs 2g and 2g + 1 -> entry 1 + r + 2G of
                            // the even and of the odd array of the layout with origin Oout
                            const float va[8] = {o[0], o[1], o[2], o[3], o[4], o[5], o[6], o[7]};
                            const float vb[8] = {o[8], o[9], o[10], o[11], o[12], o[13], o[14], o[15]};
                            uint8_t* const de = act + even_off + (size_t)(Oout + 1 + r + 2 * G) * 16;
                            store_planes(de, va);
                            store_planes(de - even_off, vb);
                        }
                    }
                    if (!last) fence_proxy_async();
                    tc_fence_before();
                    mbar_arrive(smem_u32(&acc_empty[s]));
                }
            }
            if (trw) tr[6] = clock64();
        }
        if (last) break;
        if (trw) tr[11] = clock64();
        // ---- layer hand-over: every CTA of the cluster has stored its fp32 output; the halo frames of the next layer's tile come
        // from the neighbours' rows of that output (zeros outside the sample), the layout's edge entries are cleared
        tc_fence_before();
        if (p.NC > 1) {
            __threadfence();
            cluster_sync_all();
        } else {
            __syncthreads();
        }
        if (trw) tr[12] = clock64();
        if (warp < CH_MW) {
            const float* ysrc = Ly.y + (long)n * Ly.ybatch;
            for (int idx = tid; idx < 4 * row_pairs; idx += CH_WORK) {
                const int hf = idx / row_pairs, j = idx - hf * row_pairs;
                const int tt = hf < 2 ? hf : nfr + hf;                // tile frames 0, 1, nfr+2, nfr+3
                const int t = t0 - 2 + tt, f = 2 * j;
                float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                if (t >= 0 && t < p.T && f < p.F) {
                    const float* src = ysrc + (long)t * Ly.ypitch + (long)f * 4;
                    const float4 a = ldg_cg4(src);
                    v[0] = a.x, v[1] = a.y, v[2] = a.z, v[3] = a.w;
                    if (f + 1 < p.F) {
                        const float4 b = ldg_cg4(src + 4);
                        v[4] = b.x, v[5] = b.y, v[6] = b.z, v[7] = b.w;
                    }
                }
                store_planes(pair_dst(Oout, tt, j), v);
            }
            zero_edges(Oout, tid);
            fence_proxy_async();
        }
        __syncthreads();
        tc_fence_after();
        if (trw) tr[13] = clock64();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == CH_MW) tmem_dealloc<CH_TMEM_COLS>(tmem_base);
}

// frames per CTA that fit the shared-memory budget for this F and precision
int chain_max_frames(int G, int parts) {
    const long budget = 226L * 1024 - 2L * (CNMMA * 2 * 16 * parts * 16) - 16;
    const long Emax = budget / (2L * 16 * parts);
    const long D = 2L * G + 2;
    const long rows = Emax - D - 4L * G - 8;            // nblk * 128 <= rows
    const long nblk = rows / 128;
    return nblk < 1 ? 0 : (int)((nblk * 128) / G);
}

// bf16 planes per fp32 operand: 3 / 2 / 1 for nterms = 6 / 3 / 1 (as in gemm_tc.cu); PCTTS_CHAIN_PARTS lowers it for A/B runs
int chain_parts(int nterms) {
    static const int cap = getenv("PCTTS_CHAIN_PARTS") ? atoi(getenv("PCTTS_CHAIN_PARTS")) : 3;
    const int parts = nterms == 6 ? 3 : nterms == 3 ? 2 : nterms == 1 ? 1 : 0;
    return (cap >= 1 && cap < parts) ? cap : parts;
}

int chain_pick_nc(int T, int fmax) {
    if (fmax < 1) return 0;
    for (int nc = 1; nc <= 16; nc *= 2)
        if ((long)nc * fmax >= T) {
            // every CTA of the cluster must own at least one frame: with Fc = ceil(T / nc) the last rank starts at (nc-1) * Fc,
            // which can reach or pass T when fmax is small against T (wide maps: F = 227, T = 225 -> nc = 16, Fc = 15); such
            // shapes are refused here (the callers then run the stack layer by layer) instead of launching an empty CTA
            const long fc = ((long)T + nc - 1) / nc;
            return (long)(nc - 1) * fc < T ? nc : 0;
        }
    return 0;
}

template <int PARTS>
int launch_chain(ChainParams& p, cudaStream_t st) {
    const size_t smem = ChSmem<PARTS>::total(p.E);
    if (smem > 227 * 1024 - 256) return PCTTS_EUNSUPPORTED;
    static size_t attr_smem = 0;
    static bool np_set = false;
    if (smem > attr_smem) {
        PCTTS_CUDA(cudaFuncSetAttribute(conv2d_chain_kernel<PARTS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_smem = smem;
    }
    if (p.NC > 8 && !np_set) {
        PCTTS_CUDA(cudaFuncSetAttribute(conv2d_chain_kernel<PARTS>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        np_set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(p.NB * p.NC), 1, 1);
    cfg.blockDim = dim3(CH_THREADS, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (p.NC > 1) {
        attr[na].id = cudaLaunchAttributeClusterDimension;
        attr[na].val.clusterDim.x = (unsigned)p.NC;
        attr[na].val.clusterDim.y = 1;
        attr[na].val.clusterDim.z = 1;
        ++na;
    }
    if (pctts_pdl_enabled()) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    (void)cudaLaunchKernelEx(&cfg, conv2d_chain_kernel<PARTS>, p);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

long long* g_chain_trace = nullptr;

}  // namespace

/* developer probe: device buffer of PCTTS_CHAIN_MAX_LAYERS*16 int64 receiving clock64() stamps of CTA 0, per layer: [0] MMA warp
 * starts, [1] last block issued, [4] next weight operand built, [6] this set's last epilogue done, [11] hand-over starts, [12] cluster barrier passed, [13] halo in place; NULL switches it off */
extern "C" int pctts_conv2d_chain_set_trace(void* buf) {
    g_chain_trace = (long long*)buf;
    return PCTTS_OK;
}

/* CTAs per sample (cluster size: 1, 2, 4, 8 or 16) the chain kernel would use for maps [*, T, F, <=4] at this precision, or 0 when
 * the shape is not supported (T too long for 16 CTAs of shared memory, F too wide).  Host call. */
extern "C" int pctts_conv2d_chain_supported(int T, int F, int nterms) {
    if (T < 1 || F < 1) return 0;
    const int parts = chain_parts(nterms);
    if (!parts) return 0;
    const int G = (F + 2 + 3) / 4;
    if (G > 128) return 0;
    return chain_pick_nc(T, chain_max_frames(G, parts));
}

extern "C" int pctts_conv2d_chain(const float* x, long xbatch, long xpitch, int cin0, int NB, int T, int F, int nlayers,
                                  const pctts_chain_layer* layers, int nterms, void* stream) {
    if (!x || !layers || NB < 0 || T < 0 || F <= 0 || nlayers < 1 || nlayers > CH_MAXL) return PCTTS_EINVAL;
    if (cin0 != 1 && cin0 != 4) return PCTTS_EUNSUPPORTED;
    const int parts = chain_parts(nterms);
    if (!parts) return PCTTS_EINVAL;
    if (NB == 0 || T == 0) return PCTTS_OK;
    ChainParams p;
    p.x = x, p.xbatch = xbatch, p.xpitch = xpitch, p.cin0 = cin0;
    p.NB = NB, p.T = T, p.F = F;
    p.G = (F + 2 + 3) / 4;
    if (p.G > 128) return PCTTS_EUNSUPPORTED;
    if (cin0 == 4 && ((((uintptr_t)x) & 15) || ((xbatch | xpitch) & 3))) return PCTTS_EUNSUPPORTED;
    const int fmax = chain_max_frames(p.G, parts);
    p.NC = chain_pick_nc(T, fmax);
    if (p.NC < 1) return PCTTS_EUNSUPPORTED;
    p.Fc = (T + p.NC - 1) / p.NC;
    p.nblk = (p.Fc * p.G + 127) / 128;
    p.D = 2 * p.G + 2;
    p.E = p.D + p.nblk * 128 + 4 * p.G + 8;
    p.nlayers = nlayers;
    p.trace = g_chain_trace;
    static const int dbg = getenv("PCTTS_CHAIN_DEBUG") ? atoi(getenv("PCTTS_CHAIN_DEBUG")) : 0;
    p.debug = dbg;
    int cin = cin0;
    for (int l = 0; l < nlayers; ++l) {
        const pctts_chain_layer& s = layers[l];
        ChainLayer& d = p.L[l];
        if (!s.w || !s.y) return PCTTS_EINVAL;
        if (s.cin != cin || (s.cout != 4 && !(s.cout == 1 && l == nlayers - 1))) return PCTTS_EUNSUPPORTED;
        if ((s.flags & PCTTS_EPI_BIAS) && !s.bias) return PCTTS_EINVAL;
        if ((s.flags & PCTTS_EPI_DLRELU_AUX) && !s.aux) return PCTTS_EINVAL;
        if (s.flags & ~(PCTTS_EPI_BIAS | PCTTS_EPI_LRELU | PCTTS_EPI_DLRELU_AUX | PCTTS_EPI_ACCUM)) return PCTTS_EUNSUPPORTED;
        if ((s.flags & PCTTS_EPI_ACCUM) && l != nlayers - 1) return PCTTS_EUNSUPPORTED;
        if (s.cout == 4 && ((((uintptr_t)s.y) & 15) || ((s.ybatch | s.ypitch) & 3))) return PCTTS_EUNSUPPORTED;
        if (s.aux && ((((uintptr_t)s.aux) & 15) || ((s.auxbatch | s.auxpitch) & 3))) return PCTTS_EUNSUPPORTED;
        d.w = s.w, d.bias = s.bias, d.aux = s.aux, d.y = s.y;
        d.ybatch = s.ybatch, d.ypitch = s.ypitch, d.auxbatch = s.auxbatch, d.auxpitch = s.auxpitch;
        d.cin = s.cin, d.cout = s.cout, d.flags = s.flags, d.flip = s.flip;
        cin = s.cout;
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (parts == 3) return launch_chain<3>(p, st);
    if (parts == 2) return launch_chain<2>(p, st);
    return launch_chain<1>(p, st);
}
