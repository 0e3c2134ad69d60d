// tcgen05 / TMEM / TMA GEMM family for the dense contractions of the WGAN-GP step.
//
// What it replaces in the reference: the cuBLAS / cuDNN calls behind keras Dense (networktts.py:59-63,196-197), Conv1D k=21
// (networktts.py:116-120; the dominant contraction, 18.3 GFLOP forward at the demo shape) and the LSTM input projections
// (networktts.py:90-94), forward, data gradient and weight gradient.
//
// Numerics.  The reference computes in fp32.  The 5th-generation tensor cores take bf16 (or tf32) inputs, so every fp32
// operand x is carried as a two-term bf16 expansion  x = hi + lo + O(2^-17 |x|)  (pctts_split_bf16) and a product is formed as
//     A.B  ~=  Ahi.Bhi + Ahi.Blo + Alo.Bhi          (three tcgen05.mma per k-slice, one fp32 accumulator in TMEM)
// which drops only the lo.lo term: relative error ~2^-16 per product, fp32 accumulation -- two orders of magnitude inside the
// 1e-3 parity budget of the fp32 configuration.  nterms = 1 (hi.hi only) is the bf16 configuration of BASELINE config #3.
//
// Structure (one 128 x BN output tile per CTA, 128 threads):
//   warp 0 / lane 0 : TMA producer.  Per 64-deep k-block one cp.async.bulk.tensor per operand plane into a 128B-swizzled
//                     stage (K-major operands: one [ROWS x 64] box; MN-major operands: ROWS/64 boxes of [64 k x 64 mn]).
//   warp 1 / lane 0 : MMA issuer.  4 k-slices (K=16) x nterms tcgen05.mma.cta_group::1.kind::f16 per stage, tcgen05.commit
//                     releases the stage back to the producer; a final commit signals the epilogue.
//   warp 2          : TMEM allocation (BN fp32 columns) / deallocation.
//   all 4 warps     : epilogue. tcgen05.ld 32x32b.x32 (thread = accumulator row), fused alpha / bias / per-row-block addend /
//                     LeakyReLU / sigmoid / LeakyReLU' mask / accumulate, fp32 stores; or raw split-K partials.
//
// Operand addressing is an affine map (tile index, k-block) -> TMA coordinates, which lets the same kernel run
//   * plain GEMMs in all four transpose combinations from row-major storage (K-major or MN-major UMMA descriptors),
//   * the 'same' Conv1D forward as a *virtual* im2col GEMM (k-block = (tap w, 64 channels); the A box is the activation
//     matrix shifted down by w rows; out-of-range channels / rows are zero-filled by TMA),
//   * the Conv1D weight gradient (M-tile = (tap w, 128 channels); k-block = (utterance b, 64 frames); 3-D tensor maps).
#include "common.cuh"
#include "pctts.h"
#include "tc_common.cuh"
#include <cuda_bf16.h>
#include <stdlib.h>

namespace {

using namespace tcx;

constexpr int BM = 128, BK = 64;
constexpr int MAX_STAGES = 8;
constexpr uint32_t A_PLANE_BYTES = BM * BK * 2;   // 16 KB

struct Geom {
    int mn_major;   // 0: K-major (TMA inner dim = k), 1: MN-major (TMA inner dim = m / n)
    int t_div;      // tile index t -> (t % t_div, t / t_div)
    int t_lo[3], t_hi[3], k_lo[3], k_hi[3];
};

struct Params {
    Geom a, b;
    int kb_div;     // k-block index kb -> (kb % kb_div, kb / kb_div)
    int num_kb, kb_per_split;
    int nterms, stages;
    long long* trace;   // developer probe: clock64() stamps of CTA (0,0,0), 8 per k-block (pctts_gemm_tc_set_trace)
    int debug;      // developer probe (PCTTS_TC_DEBUG): 1 = skip the TMA loads, 2 = skip the MMAs
    int chunk_kb;   // k-blocks accumulated in TMEM before the sum is promoted to fp32 registers (round-to-nearest adds)
    // output
    float* C;
    long ldc;
    int N;
    int seg_len, seg_valid, nseg;   // accumulator row r -> (seg = r / seg_len, t = r % seg_len); stored iff t < seg_valid,
                                    // at C row seg * seg_valid + t
    float alpha;
    const float* bias;
    const float* aux;
    const float* addmod;
    int addrows;
    int flags;
    float* partial;   // split-K: raw sums, plane z at partial + z * prows * N (dense rows of N floats, row = crow - prow0)
    long crows;
    // tail split (wave quantisation): the first tail_dp m-tiles run over the whole K range and store directly; every later
    // m-tile is cut into tail_splits K ranges whose raw sums go to `partial` (rows prow0.. of C, prows of them).  Plain
    // split-K is tail_splits = 0 with blockIdx.z as the split index, prow0 = 0, prows = crows.
    int tail_dp = 0, tail_splits = 0;
    long prow0 = 0, prows = -1;   // prows < 0: filled in by launch_tc (= crows)
};

__device__ __forceinline__ void coords(const Geom& g, int t, int kb_lo, int kb_hi, int (&c)[3]) {
    const int tl = t % g.t_div, th = t / g.t_div;
#pragma unroll
    for (int d = 0; d < 3; ++d) c[d] = tl * g.t_lo[d] + th * g.t_hi[d] + kb_lo * g.k_lo[d] + kb_hi * g.k_hi[d];
}

constexpr int NUM_THREADS = 192;   // warp 0: TMA producer, warp 1: MMA issuer, warps 2..5: accumulator drain + epilogue

template <int BN, int PARTS>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, const Params p) {
    pdl_trigger();   // the wait follows the barrier / TMEM set-up (nothing before it touches global memory)
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[MAX_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[MAX_STAGES];
    __shared__ __align__(8) uint64_t acc_full[2];
    __shared__ __align__(8) uint64_t acc_empty[2];
    __shared__ uint32_t tmem_slot;

    constexpr uint32_t B_PLANE_BYTES = BN * BK * 2;
    constexpr int TMEM_COLS = BN == 128 ? 512 : 256;   // main accumulator x2 (ping-pong) + correction accumulator
    constexpr int parts = PARTS;                       // bf16 planes per operand: 1 (bf16), 2 (fp32x3), 3 (fp32)
    constexpr uint32_t stage_bytes = parts * (A_PLANE_BYTES + B_PLANE_BYTES);
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;   // SWIZZLE_128B atoms need 1024-byte alignment

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool tr0 = p.trace && (blockIdx.x | blockIdx.y | blockIdx.z) == 0;
    if (tr0 && threadIdx.x == 0) p.trace[3] = clock64();
    const int nt = blockIdx.x;
    int mt = blockIdx.y, zsplit = blockIdx.z;
    float* part = p.partial;
    int kb_begin, kb_end;
    if (p.tail_splits > 0) {
        if (mt < p.tail_dp) {
            kb_begin = 0, kb_end = p.num_kb, part = nullptr;
        } else {
            const int j = mt - p.tail_dp;
            mt = p.tail_dp + j / p.tail_splits;
            zsplit = j - (mt - p.tail_dp) * p.tail_splits;
            kb_begin = zsplit * p.kb_per_split;
            kb_end = min(p.num_kb, kb_begin + p.kb_per_split);
        }
    } else {
        kb_begin = zsplit * p.kb_per_split;
        kb_end = min(p.num_kb, kb_begin + p.kb_per_split);
    }
    const int nkb = kb_end - kb_begin;
    const int CH = p.chunk_kb;
    const int nchunks = (nkb + CH - 1) / CH;

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&mapA);
        tma_prefetch_desc(&mapB);
        for (int s = 0; s < p.stages; ++s) {
            mbar_init(smem_u32(&full_bar[s]), 1);
            mbar_init(smem_u32(&empty_bar[s]), 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(smem_u32(&acc_full[b]), 1);
            mbar_init(smem_u32(&acc_empty[b]), 4);   // one arrival per drain warp
        }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc<TMEM_COLS>(smem_u32(&tmem_slot));
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    pdl_wait();
    const uint32_t tmem_base = tmem_slot;
    if (tr0 && threadIdx.x == 0) p.trace[7] = clock64();
    const uint32_t tmem_corr = tmem_base + 2 * BN;

    if (warp == 0) {
        // ---------------- TMA producer (whole warp walks the loop, one elected lane issues) ----------------
        int s = 0;
        uint32_t ph = 0;
        int kb_lo = kb_begin % p.kb_div, kb_hi = kb_begin / p.kb_div;
        int a0[3], b0[3];
        coords(p.a, mt, 0, 0, a0);      // tile part of the affine coordinate map; the k-block part is added per iteration
        coords(p.b, nt, 0, 0, b0);
        for (int i = 0; i < nkb; ++i) {
            const bool tr = p.trace && (blockIdx.x | blockIdx.y | blockIdx.z) == 0 && i < 64;
            if (tr && lane == 0) p.trace[i * 8 + 0] = clock64();
            mbar_wait(smem_u32(&empty_bar[s]), ph ^ 1u);
            if (tr && lane == 0) p.trace[i * 8 + 1] = clock64();
            const uint32_t fb = smem_u32(&full_bar[s]);
            int ca[3], cb[3];
#pragma unroll
            for (int d = 0; d < 3; ++d) {
                ca[d] = a0[d] + kb_lo * p.a.k_lo[d] + kb_hi * p.a.k_hi[d];
                cb[d] = b0[d] + kb_lo * p.b.k_lo[d] + kb_hi * p.b.k_hi[d];
            }
            const uint32_t sa = smem_base + s * stage_bytes;
            const uint32_t sb = sa + parts * A_PLANE_BYTES;
            if (elect_one()) {
                if (p.debug & 1) {
                    mbar_arrive(fb);
                } else {
                    mbar_arrive_expect_tx(fb, stage_bytes);
                    for (int part = 0; part < parts; ++part) {
                        if (!p.a.mn_major) {
                            tma_load_4d(sa + part * A_PLANE_BYTES, &mapA, fb, ca[0], ca[1], ca[2], part);
                        } else {
#pragma unroll
                            for (int h = 0; h < BM / 64; ++h)
                                tma_load_4d(sa + part * A_PLANE_BYTES + h * 8192, &mapA, fb, ca[0] + 64 * h, ca[1], ca[2], part);
                        }
                        if (!p.b.mn_major) {
                            tma_load_4d(sb + part * B_PLANE_BYTES, &mapB, fb, cb[0], cb[1], cb[2], part);
                        } else {
#pragma unroll
                            for (int h = 0; h < BN / 64; ++h)
                                tma_load_4d(sb + part * B_PLANE_BYTES + h * 8192, &mapB, fb, cb[0] + 64 * h, cb[1], cb[2], part);
                        }
                    }
                }
            }
            __syncwarp();
            if (tr && lane == 0) p.trace[i * 8 + 2] = clock64();
            if (++s == p.stages) s = 0, ph ^= 1u;
            if (++kb_lo == p.kb_div) kb_lo = 0, ++kb_hi;
        }
    } else if (warp == 1) {
        // ---------------- MMA issuer (whole warp walks the loop, one elected lane issues) ----------------
        const uint32_t idesc = umma_idesc_bf16(BM, BN, p.a.mn_major, p.b.mn_major);
        // K-major SW128 tile: rows of 128 B, 8-row groups 1024 B apart (SBO); a K=16 slice is 32 B further along the row.
        // MN-major SW128 tile: [k][64 mn] rows of 128 B; 8-k groups 1024 B apart (SBO); next 64 mn 8192 B apart (LBO);
        //                      a K=16 slice is 16 rows = 2048 B further.
        const uint32_t a_lbo = p.a.mn_major ? 8192u : 16u, a_step16 = p.a.mn_major ? 128u : 2u;
        const uint32_t b_lbo = p.b.mn_major ? 8192u : 16u, b_step16 = p.b.mn_major ? 128u : 2u;
        // descriptor of plane 0 / stage 0 / k-slice 0; the others differ only in the 14-bit start-address field (no carry out)
        const uint64_t adesc0 = umma_smem_desc(smem_base, a_lbo, 1024u);
        const uint64_t bdesc0 = umma_smem_desc(smem_base + parts * A_PLANE_BYTES, b_lbo, 1024u);
        int s = 0, c = 0, cpos = 0;
        uint32_t ph = 0;
        for (int i = 0; i < nkb; ++i) {
            const int buf = c & 1;
            const bool chunk_first = cpos == 0;
            if (chunk_first) {      // the drain warps must have emptied this ping-pong buffer
                mbar_wait(smem_u32(&acc_empty[buf]), (((uint32_t)c >> 1) & 1u) ^ 1u);
            }
            const bool tr = p.trace && (blockIdx.x | blockIdx.y | blockIdx.z) == 0 && i < 64;
            if (tr && lane == 0) p.trace[i * 8 + 4] = clock64();
            mbar_wait(smem_u32(&full_bar[s]), ph);
            tc_fence_after();
            if (tr && lane == 0) p.trace[i * 8 + 5] = clock64();
            const uint32_t soff = (uint32_t)s * (stage_bytes >> 4);     // descriptors address smem in 16-byte units
            const uint32_t tmem_main = tmem_base + buf * BN;
            if (elect_one()) {
                if (!(p.debug & 2)) {
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) {
                        // planes 0 (hi), 1, 2 (lo): every product A_i . B_j with i + j < parts (1, 3 or 6 tcgen05.mma per
                        // k-slice).  hi.hi goes to the chunk accumulator; the corrections (2^-8 and smaller) to their own
                        // accumulator, so that the tensor core's truncating accumulate costs them nothing.
#pragma unroll
                        for (int ia = 0; ia < parts; ++ia) {
                            const uint64_t ad = adesc0 + soff + (uint32_t)(ia * (A_PLANE_BYTES >> 4)) + (uint32_t)k * a_step16;
#pragma unroll
                            for (int ib = 0; ib < parts - ia; ++ib) {
                                const uint64_t bd = bdesc0 + soff + (uint32_t)(ib * (B_PLANE_BYTES >> 4)) + (uint32_t)k * b_step16;
                                if ((ia | ib) == 0)
                                    umma_bf16(tmem_main, ad, bd, idesc, (chunk_first && k == 0) ? 0u : 1u);
                                else
                                    umma_bf16(tmem_corr, ad, bd, idesc, (i == 0 && k == 0 && ia == 0 && ib == 1) ? 0u : 1u);
                            }
                        }
                    }
                }
                umma_commit(smem_u32(&empty_bar[s]));   // stage reusable once these MMAs have read it
                if (cpos == CH - 1 || i == nkb - 1) umma_commit(smem_u32(&acc_full[buf]));
            }
            __syncwarp();
            if (tr && lane == 0) p.trace[i * 8 + 6] = clock64();
            if (++s == p.stages) s = 0, ph ^= 1u;
            if (++cpos == CH) cpos = 0, ++c;
        }
    } else {
        // ---------------- accumulator drain + epilogue: warps 2..5, TMEM lane quarter (warp & 3) ----------------
        const int q = warp & 3;
        const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
        float tot[BN];
#pragma unroll
        for (int j = 0; j < BN; ++j) tot[j] = 0.f;
        // promote every finished chunk (chunk_kb k-blocks of hi.hi products) into fp32 registers with round-to-nearest adds
        for (int c = 0; c < nchunks; ++c) {
            const int buf = c & 1;
            mbar_wait(smem_u32(&acc_full[buf]), ((uint32_t)c >> 1) & 1u);
            tc_fence_after();
#pragma unroll
            for (int g = 0; g < BN / 32; ++g) {
                uint32_t v[32];
                tmem_ld_32x32(tmem_base + lane_addr + (uint32_t)(buf * BN + g * 32), v);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; ++j) tot[g * 32 + j] += __uint_as_float(v[j]);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&acc_empty[buf]));
        }
        if (tr0 && threadIdx.x == 64) p.trace[8 + 3] = clock64();
        // ---- epilogue, phase 1: add the correction accumulator and park the raw tile in (now idle) pipeline smem, row per thread
        constexpr int LDS = BN + 4;                     // floats per smem row: 16-byte phases of 8 lanes hit 8 distinct bank groups
        float* tile = reinterpret_cast<float*>(smem_raw + (smem_base - smem_u32(smem_raw)));
        long* crow_s = reinterpret_cast<long*>(tile + BM * LDS);
        const int rl = q * 32 + lane;                   // row inside the tile
        {
            const int r = mt * BM + rl;
            const int seg = r / p.seg_len, tt = r - seg * p.seg_len;
            crow_s[rl] = (tt < p.seg_valid && seg < p.nseg) ? (long)seg * p.seg_valid + tt : -1;
        }
#pragma unroll
        for (int g = 0; g < BN / 32; ++g) {
            if (parts > 1) {
                uint32_t v[32];
                tmem_ld_32x32(tmem_corr + lane_addr + (uint32_t)(g * 32), v);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; ++j) tot[g * 32 + j] += __uint_as_float(v[j]);
            }
#pragma unroll
            for (int j = 0; j < 32; j += 4)
                *reinterpret_cast<float4*>(tile + rl * LDS + g * 32 + j) =
                    make_float4(tot[g * 32 + j], tot[g * 32 + j + 1], tot[g * 32 + j + 2], tot[g * 32 + j + 3]);
        }
        if (tr0 && threadIdx.x == 64) p.trace[16 + 7] = clock64();
        asm volatile("bar.sync 1, 128;" ::: "memory");   // the four epilogue warps only
        if (tr0 && threadIdx.x == 64) p.trace[24 + 3] = clock64();
        // ---- phase 2: coalesced pass over the tile.  A thread keeps one group of 4 columns and walks down the rows, so that the
        //      column-only terms (bias, bounds) are hoisted; the body is kept short on purpose: with one epilogue warp per
        //      scheduler it runs at dependent-issue latency, and straight-line unrolled code ran at instruction-fetch speed.
        const int n0 = nt * BN;
        const int flags = p.flags;
        const int et = threadIdx.x - 64;                // 0..127
        constexpr int V = BN / 4;                       // float4 per row
        constexpr int RSTEP = 128 / V;                  // rows covered by the 128 threads per pass
        const int c4 = (et % V) * 4, nb = n0 + c4;
        const int nval = p.N - nb;                      // valid columns of this thread's group (<= 0: nothing to do)
        if (nval > 0) {
            const bool full4 = nval >= 4;
            float bias4[4] = {0.f, 0.f, 0.f, 0.f};
            if ((flags & PCTTS_EPI_BIAS) && !part) {     // raw sums go to the split-K planes; the reduce adds the bias once
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (j < nval) bias4[j] = __ldg(p.bias + nb + j);
            }
            const float alpha = p.alpha;
            const long ldo = part ? (long)p.N : p.ldc;
            float* const obase = part ? part + ((long)zsplit * p.prows - p.prow0) * p.N + nb : p.C + nb;
            const int eflags = part ? 0 : flags;
            const bool ovec = full4 && ((ldo & 3) == 0) && ((((uintptr_t)obase) & 15) == 0);
#pragma unroll 1
            for (int row = et / V; row < BM; row += RSTEP) {
                const long crow = crow_s[row];
                if (crow < 0) continue;
                const float4 t4 = *reinterpret_cast<const float4*>(tile + row * LDS + c4);
                float x[4] = {t4.x, t4.y, t4.z, t4.w};
                float* dst = obase + crow * ldo;
                if (eflags | (alpha != 1.f)) {
                    const float* auxp = (eflags & PCTTS_EPI_DLRELU_AUX) ? p.aux + crow * p.ldc + nb : nullptr;
                    const float* addp = (eflags & PCTTS_EPI_ADDMOD) ? p.addmod + (long)((int)crow % p.addrows) * p.N + nb : nullptr;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        if (j < nval) {
                            float y = x[j];
                            if (!part) y *= alpha;
                            y += bias4[j];
                            if (eflags & PCTTS_EPI_ADDMOD) y += __ldg(addp + j);
                            if (eflags & PCTTS_EPI_LRELU) y = lrelu_f(y);
                            if (eflags & PCTTS_EPI_SIGMOID) y = sigmoid_f(y);
                            if (eflags & PCTTS_EPI_DLRELU_AUX) y *= dlrelu_from_out(auxp[j]);
                            if (eflags & PCTTS_EPI_ACCUM) y += dst[j];
                            x[j] = y;
                        }
                    }
                }
                if (p.debug & 8) continue;
                if (ovec) {
                    *reinterpret_cast<float4*>(dst) = make_float4(x[0], x[1], x[2], x[3]);
                } else {
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if (j < nval) dst[j] = x[j];
                }
            }
        }
    }
    if (tr0 && threadIdx.x == 64) p.trace[8 + 7] = clock64();
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc<TMEM_COLS>(tmem_base);
    if (tr0 && threadIdx.x == 64) p.trace[16 + 3] = clock64();
}

// C (+)= epilogue(alpha * sum_z partial[z]), fixed order (deterministic); epilogue = bias, LeakyReLU, LeakyReLU' mask (aux is
// addressed like C)
__global__ void tc_splitk_reduce_kernel(const float* __restrict__ partial, int splits, long rows, int N, float* __restrict__ C,
                                        long ldc, float alpha, int accum, const float* __restrict__ bias,
                                        const float* __restrict__ aux, int flags) {
    pdl_entry();
    const long total = rows * N;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        float s = 0.f;
        for (int z = 0; z < splits; ++z) s += partial[(long)z * total + i];
        const long m = i / N;
        const int n = (int)(i - m * N);
        float* c = C + m * ldc + n;
        s *= alpha;
        if (flags & PCTTS_EPI_BIAS) s += __ldg(bias + n);
        if (flags & PCTTS_EPI_LRELU) s = lrelu_f(s);
        if (flags & PCTTS_EPI_DLRELU_AUX) s *= dlrelu_from_out(aux[m * ldc + n]);
        *c = accum ? *c + s : s;
    }
}

// The same, four columns per thread (N % 4 == 0, 16-byte aligned rows of C / aux / partial): same order of additions per element.
__global__ void __launch_bounds__(256) tc_splitk_reduce_vec4_kernel(const float4* __restrict__ partial, int splits, int rows, int N4,
                                                                    float* __restrict__ C, long ldc, float alpha, int accum,
                                                                    const float* __restrict__ bias, const float* __restrict__ aux,
                                                                    int flags) {
    pdl_entry();
    const int total = rows * N4;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int z = 0; z < splits; ++z) {
            const float4 v = partial[(long)z * total + i];
            s.x += v.x, s.y += v.y, s.z += v.z, s.w += v.w;
        }
        const int m = i / N4, n = (i - m * N4) * 4;
        float o[4] = {s.x * alpha, s.y * alpha, s.z * alpha, s.w * alpha};
        if (flags & PCTTS_EPI_BIAS) {
#pragma unroll
            for (int j = 0; j < 4; ++j) o[j] += __ldg(bias + n + j);
        }
        if (flags & PCTTS_EPI_LRELU) {
#pragma unroll
            for (int j = 0; j < 4; ++j) o[j] = lrelu_f(o[j]);
        }
        float4* c = reinterpret_cast<float4*>(C + (long)m * ldc + n);
        if (flags & PCTTS_EPI_DLRELU_AUX) {
            const float4 a = *reinterpret_cast<const float4*>(aux + (long)m * ldc + n);
            o[0] *= dlrelu_from_out(a.x), o[1] *= dlrelu_from_out(a.y), o[2] *= dlrelu_from_out(a.z), o[3] *= dlrelu_from_out(a.w);
        }
        if (accum) {
            const float4 d = *c;
            o[0] += d.x, o[1] += d.y, o[2] += d.z, o[3] += d.w;
        }
        *c = make_float4(o[0], o[1], o[2], o[3]);
    }
}

// x -> bf16 hi (+ lo) planes; one thread per pair of columns
__global__ void split_bf16_kernel(const float* __restrict__ src, long ld, int seg, long segstride, long rows, int cols,
                                  __nv_bfloat16* __restrict__ dst, long ld_dst, long part_stride, int parts) {
    pdl_entry();
    const long pairs_per_row = ld_dst / 2;
    const long total = rows * pairs_per_row;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const long r = i / pairs_per_row;
        const int c = (int)(i - r * pairs_per_row) * 2;
        long off;
        if (seg > 0) {
            const long s = r / seg;
            off = s * segstride + (r - s * seg) * ld;
        } else {
            off = r * ld;
        }
        const float x0 = c < cols ? src[off + c] : 0.f;
        const float x1 = c + 1 < cols ? src[off + c + 1] : 0.f;
        const __nv_bfloat16 h0 = __float2bfloat16_rn(x0), h1 = __float2bfloat16_rn(x1);
        __nv_bfloat162 hv;
        hv.x = h0;
        hv.y = h1;
        *reinterpret_cast<__nv_bfloat162*>(dst + r * ld_dst + c) = hv;
        if (parts > 1) {
            const float r0 = x0 - __bfloat162float(h0), r1 = x1 - __bfloat162float(h1);
            __nv_bfloat162 mv;
            mv.x = __float2bfloat16_rn(r0);
            mv.y = __float2bfloat16_rn(r1);
            *reinterpret_cast<__nv_bfloat162*>(dst + part_stride + r * ld_dst + c) = mv;
            if (parts > 2) {
                __nv_bfloat162 lv;
                lv.x = __float2bfloat16_rn(r0 - __bfloat162float(mv.x));
                lv.y = __float2bfloat16_rn(r1 - __bfloat162float(mv.y));
                *reinterpret_cast<__nv_bfloat162*>(dst + 2 * part_stride + r * ld_dst + c) = lv;
            }
        }
    }
}

// The same for cols % 4 == 0, 16-byte aligned rows (ld % 4 == 0): 8 columns per thread -- two float4 loads, one 16-byte store per
// plane -- and 32-bit index arithmetic.  Same rounding steps as above: bit-identical planes.
__device__ __forceinline__ uint32_t bf16x2_rn(float a, float b, float& ra, float& rb) {
    const __nv_bfloat16 ha = __float2bfloat16_rn(a), hb = __float2bfloat16_rn(b);
    ra = a - __bfloat162float(ha), rb = b - __bfloat162float(hb);
    return (uint32_t)__bfloat16_as_ushort(ha) | ((uint32_t)__bfloat16_as_ushort(hb) << 16);
}

__global__ void __launch_bounds__(256) split_bf16_vec8_kernel(const float* __restrict__ src, long ld, int seg, long segstride, int rows,
                                                              int cols, __nv_bfloat16* __restrict__ dst, long ld_dst, long part_stride,
                                                              int parts) {
    pdl_entry();
    const int groups = (int)(ld_dst >> 3);
    const int total = rows * groups;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int r = i / groups;
        const int c = (i - r * groups) * 8;
        long off;
        if (seg > 0) {
            const int sgm = r / seg;
            off = (long)sgm * segstride + (long)(r - sgm * seg) * ld;
        } else {
            off = (long)r * ld;
        }
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
        if (c < cols) a = *reinterpret_cast<const float4*>(src + off + c);
        if (c + 4 < cols) b = *reinterpret_cast<const float4*>(src + off + c + 4);
        float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        __nv_bfloat16* d = dst + (long)r * ld_dst + c;
        for (int pl = 0; pl < parts; ++pl) {
            uint32_t w[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) w[j] = bf16x2_rn(v[2 * j], v[2 * j + 1], v[2 * j], v[2 * j + 1]);
            *reinterpret_cast<uint4*>(d + (long)pl * part_stride) = make_uint4(w[0], w[1], w[2], w[3]);
        }
    }
}

// several matrices in ONE launch (the weight shadows of a network after its Adam update): blockIdx.y = job
constexpr int SPLIT_MAXJOBS = PCTTS_SPLIT_MAX_JOBS;
struct SplitJob {
    const float* src;
    __nv_bfloat16* dst;
    long ld, ld_dst, part_stride;
    int rows, cols, vec;
};
struct SplitJobs {
    SplitJob j[SPLIT_MAXJOBS];
    int parts;
};
__global__ void __launch_bounds__(256) split_bf16_multi_kernel(const SplitJobs P) {
    pdl_entry();
    const SplitJob& J = P.j[blockIdx.y];
    if (J.vec) {
        const int groups = (int)(J.ld_dst >> 3);
        const int total = J.rows * groups;
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
            const int r = i / groups;
            const int c = (i - r * groups) * 8;
            const long off = (long)r * J.ld;
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
            if (c < J.cols) a = *reinterpret_cast<const float4*>(J.src + off + c);
            if (c + 4 < J.cols) b = *reinterpret_cast<const float4*>(J.src + off + c + 4);
            float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
            __nv_bfloat16* d = J.dst + (long)r * J.ld_dst + c;
            for (int pl = 0; pl < P.parts; ++pl) {
                uint32_t w[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) w[q] = bf16x2_rn(v[2 * q], v[2 * q + 1], v[2 * q], v[2 * q + 1]);
                *reinterpret_cast<uint4*>(d + (long)pl * J.part_stride) = make_uint4(w[0], w[1], w[2], w[3]);
            }
        }
    } else {
        const long pairs = J.ld_dst / 2;
        const long total = (long)J.rows * pairs;
        for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
            const long r = i / pairs;
            const int c = (int)(i - r * pairs) * 2;
            float x0 = c < J.cols ? J.src[r * J.ld + c] : 0.f, x1 = c + 1 < J.cols ? J.src[r * J.ld + c + 1] : 0.f;
            for (int pl = 0; pl < P.parts; ++pl) {
                const uint32_t w = bf16x2_rn(x0, x1, x0, x1);
                *reinterpret_cast<uint32_t*>(J.dst + (long)pl * J.part_stride + r * J.ld_dst + c) = w;
            }
        }
    }
}

long long* g_trace = nullptr;

struct Launch {
    const void* a_base;
    uint64_t a_dims[4];
    uint64_t a_strides[3];
    const void* b_base;
    uint64_t b_dims[4];
    uint64_t b_strides[3];
    int m_tiles, n_tiles_hint;   // n tiles derive from N and BN
    int splits;
};

int pick_bn(int m_tiles, int N) {
    if (N <= 64) return 64;
    const int t128 = m_tiles * ((N + 127) / 128);
    return t128 >= 100 ? 128 : 64;
}

// Tail split against wave quantisation: with more output tiles than SMs, the tiles of the full waves run as they are and the
// m-tiles of the last, partly filled wave are cut along K so that their pieces fill the machine once more (e.g. the Conv1D
// weight gradient: 168 tiles of 70 k-blocks = 2 waves for 1.14 waves of work -> 148 tiles + 20 tiles x 7 ranges of 10).
// Returns the number of K ranges (0: leave the launch alone) and the count of whole-K m-tiles in *dp.
int pick_tail(int m_tiles, int n_tiles, int num_kb, int* dp) {
    *dp = 0;
    static const int off = getenv("PCTTS_NO_TAIL_SPLIT") ? atoi(getenv("PCTTS_NO_TAIL_SPLIT")) : 0;
    const int tiles = m_tiles * n_tiles;
    if (off || tiles <= 148 || n_tiles > 148 || num_kb < 12) return 0;
    const int dpm = (tiles / 148) * 148 / n_tiles;
    const int tail_tiles = (m_tiles - dpm) * n_tiles;
    if (tail_tiles <= 0 || tail_tiles > 74) return 0;      // nothing left over / the last wave is at least half full anyway
    int s = 148 / tail_tiles;
    if (s > num_kb / 4) s = num_kb / 4;
    if (s < 2) return 0;
    *dp = dpm;
    return s;
}

// first output row of m-tile mt under the (seg_len, seg_valid) row map of Params
long tile_first_row(int mt, int seg_len, int seg_valid) {
    const long r = (long)mt * BM;
    const long seg = r / seg_len, tt = r - seg * seg_len;
    return seg * seg_valid + (tt < seg_valid ? tt : seg_valid);
}

// fills the tail-split fields of p; `workspace` must hold splits * prows * N floats (see *_workspace host calls)
void apply_tail(const Launch& L, Params& p, int bn, float* workspace) {
    int dp = 0;
    const int s = workspace ? pick_tail(L.m_tiles, (p.N + bn - 1) / bn, p.num_kb, &dp) : 0;
    if (s < 2) return;
    p.tail_dp = dp, p.tail_splits = s;
    p.kb_per_split = (p.num_kb + s - 1) / s;
    p.prow0 = tile_first_row(dp, p.seg_len, p.seg_valid);
    p.prows = p.crows - p.prow0;
    p.partial = workspace;
}

long tail_workspace(int m_tiles, int N, int bn, int num_kb, int seg_len, int seg_valid, long crows) {
    int dp = 0;
    const int s = pick_tail(m_tiles, (N + bn - 1) / bn, num_kb, &dp);
    return s < 2 ? 0 : (long)s * (crows - tile_first_row(dp, seg_len, seg_valid)) * N;
}

template <int BN, int PARTS>
int launch_bn(const Launch& L, Params p, cudaStream_t st) {
    CUtensorMap mapA, mapB;
    const uint32_t boxA[4] = {64u, p.a.mn_major ? 64u : (uint32_t)BM, 1u, 1u};
    const uint32_t boxB[4] = {64u, p.b.mn_major ? 64u : (uint32_t)BN, 1u, 1u};
    int rc = make_map_bf16_4d(&mapA, L.a_base, L.a_dims, L.a_strides, boxA);
    if (rc) return rc < 0 ? rc : PCTTS_EINVAL;
    rc = make_map_bf16_4d(&mapB, L.b_base, L.b_dims, L.b_strides, boxB);
    if (rc) return rc < 0 ? rc : PCTTS_EINVAL;
    constexpr int parts = PARTS;
    const uint32_t stage_bytes = parts * (A_PLANE_BYTES + BN * BK * 2);
    int stages = (int)((219u * 1024u) / stage_bytes);
    if (stages > MAX_STAGES) stages = MAX_STAGES;
    if (stages > p.kb_per_split) stages = p.kb_per_split < 2 ? 2 : p.kb_per_split;
    p.stages = stages;
    p.chunk_kb = p.nterms == 1 ? (1 << 20) : 4;
    static const int dbg = getenv("PCTTS_TC_DEBUG") ? atoi(getenv("PCTTS_TC_DEBUG")) : 0;
    p.debug = dbg;
    p.trace = g_trace;
    size_t smem = (size_t)stages * stage_bytes;
    const size_t epi = (size_t)BM * (BN + 4) * 4 + BM * 8;      // the epilogue parks the output tile in the pipeline smem
    if (smem < epi) smem = epi;
    smem += 1024;
    static bool attr_done = false;
    if (!attr_done) {
        PCTTS_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<BN, PARTS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        attr_done = true;
    }
    dim3 grid((p.N + BN - 1) / BN, L.m_tiles, L.splits);
    if (p.tail_splits > 0) grid.y = p.tail_dp + (L.m_tiles - p.tail_dp) * p.tail_splits, grid.z = 1;
    pctts_launch(gemm_tc_kernel<BN, PARTS>, grid, NUM_THREADS, smem, st, mapA, mapB, p);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

int launch_tc(const Launch& L, Params p, int bn, float* C, long ldc, float alpha, int accum, cudaStream_t st) {
    int rc;
    if (p.prows < 0) p.prows = p.crows;
    if (p.nterms == 6) rc = bn == 128 ? launch_bn<128, 3>(L, p, st) : launch_bn<64, 3>(L, p, st);
    else if (p.nterms == 3) rc = bn == 128 ? launch_bn<128, 2>(L, p, st) : launch_bn<64, 2>(L, p, st);
    else rc = bn == 128 ? launch_bn<128, 1>(L, p, st) : launch_bn<64, 1>(L, p, st);
    if (rc) return rc;
    if (p.partial) {
        const long total = p.prows * p.N;
        int blocks = pctts_div_up(total, 256);
        if (blocks > 148 * 8) blocks = 148 * 8;
        const int nz = p.tail_splits > 0 ? p.tail_splits : L.splits;
        float* const c0 = C + p.prow0 * ldc;
        const float* const aux0 = p.aux ? p.aux + p.prow0 * ldc : nullptr;
        const int rflags = p.flags & (PCTTS_EPI_BIAS | PCTTS_EPI_LRELU | PCTTS_EPI_DLRELU_AUX);
        static const int no_vec = getenv("PCTTS_REDUCE_VEC") ? atoi(getenv("PCTTS_REDUCE_VEC")) == 0 : 0;
        if (!no_vec && (p.N & 3) == 0 && (ldc & 3) == 0 && !(((uintptr_t)c0 | (uintptr_t)aux0 | (uintptr_t)p.partial) & 15) && total < (1L << 31)) {
            blocks = pctts_div_up(total / 4, 256);
            if (blocks > 148 * 8) blocks = 148 * 8;
            pctts_launch(tc_splitk_reduce_vec4_kernel, blocks, 256, 0, st, (const float4*)p.partial, nz, (int)p.prows, p.N / 4, c0, ldc, alpha,
                         accum, p.bias, aux0, rflags);
        } else
            pctts_launch(tc_splitk_reduce_kernel, blocks, 256, 0, st, p.partial, nz, p.prows, p.N, c0, ldc, alpha, accum, p.bias, aux0, rflags);
        PCTTS_LAUNCH_CHECK();
    }
    return PCTTS_OK;
}

void zero3(int (&v)[3]) { v[0] = v[1] = v[2] = 0; }

// geometry of a plain row-major stored matrix used as the A (tile = 128 rows) or B (tile = bn cols) operand
//   K-major  (stored [mn, k]): coords (k, mn)      MN-major (stored [k, mn]): coords (mn, k)
void plain_geom(Geom& g, int mn_major, int tile) {
    g.mn_major = mn_major;
    g.t_div = 1 << 30;
    zero3(g.t_lo), zero3(g.t_hi), zero3(g.k_lo), zero3(g.k_hi);
    if (!mn_major) {
        g.k_lo[0] = BK;
        g.t_lo[1] = tile;
    } else {
        g.t_lo[0] = tile;
        g.k_lo[1] = BK;
    }
}

}  // namespace

extern "C" int pctts_split_bf16(const float* src, long ld, int seg, long segstride, long rows, int cols, void* dst, long ld_dst,
                                long part_stride, int parts, void* stream) {
    if (!src || !dst || rows < 0 || cols <= 0 || ld_dst < cols || (ld_dst & 7) || parts < 1 || parts > 3) return PCTTS_EINVAL;
    if (rows == 0) return PCTTS_OK;
    static const int no_vec = getenv("PCTTS_SPLIT_VEC") ? atoi(getenv("PCTTS_SPLIT_VEC")) == 0 : 0;
    if (!no_vec && (cols & 3) == 0 && (ld & 3) == 0 && (segstride & 3) == 0 && !((uintptr_t)src & 15) && !((uintptr_t)dst & 15) &&
        (part_stride & 7) == 0 && rows * (ld_dst / 8) < (1L << 31)) {
        const long groups = rows * (ld_dst / 8);
        int blocks = pctts_div_up(groups, 256);
        if (blocks > 148 * 16) blocks = 148 * 16;
        pctts_launch(split_bf16_vec8_kernel, blocks, 256, 0, (cudaStream_t)stream, src, ld, seg, segstride, (int)rows, cols,
                     (__nv_bfloat16*)dst, ld_dst, part_stride, parts);
        PCTTS_LAUNCH_CHECK();
        return PCTTS_OK;
    }
    const long total = rows * (ld_dst / 2);
    int blocks = pctts_div_up(total, 256);
    if (blocks > 148 * 16) blocks = 148 * 16;
    pctts_launch(split_bf16_kernel, blocks, 256, 0, (cudaStream_t)stream, src, ld, seg, segstride, rows, cols, (__nv_bfloat16*)dst, ld_dst, part_stride, parts);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

extern "C" int pctts_split_bf16_multi(const pctts_split_job* jobs, int njobs, int parts, void* stream) {
    if (!jobs || njobs < 1 || njobs > SPLIT_MAXJOBS || parts < 1 || parts > 3) return PCTTS_EINVAL;
    SplitJobs P;
    P.parts = parts;
    long most = 0;
    for (int i = 0; i < njobs; ++i) {
        const pctts_split_job& s = jobs[i];
        if (!s.src || !s.dst || s.rows < 0 || s.cols <= 0 || s.ld_dst < s.cols || (s.ld_dst & 7) || s.rows * (s.ld_dst / 2) >= (1L << 31))
            return PCTTS_EINVAL;
        SplitJob& d = P.j[i];
        d.src = s.src, d.dst = (__nv_bfloat16*)s.dst, d.ld = s.ld, d.ld_dst = s.ld_dst, d.part_stride = s.part_stride;
        d.rows = (int)s.rows, d.cols = s.cols;
        d.vec = ((s.cols & 3) == 0 && (s.ld & 3) == 0 && !((uintptr_t)s.src & 15) && !((uintptr_t)s.dst & 15) && (s.part_stride & 7) == 0) ? 1 : 0;
        const long work = d.vec ? s.rows * (s.ld_dst / 8) : s.rows * (s.ld_dst / 2);
        if (work > most) most = work;
    }
    if (most == 0) return PCTTS_OK;
    int bx = pctts_div_up(most, 256);
    if (bx > 148 * 4) bx = 148 * 4;
    pctts_launch(split_bf16_multi_kernel, dim3(bx, njobs), 256, 0, (cudaStream_t)stream, P);
    PCTTS_LAUNCH_CHECK();
    return PCTTS_OK;
}

/* developer probe: device buffer of 64*8 int64 receiving clock64() stamps of CTA (0,0,0); NULL switches it off */
extern "C" int pctts_gemm_tc_set_trace(void* buf) {
    g_trace = (long long*)buf;
    return PCTTS_OK;
}

extern "C" long pctts_gemm_tc_workspace(int M, int N, int splits) { return splits > 1 ? (long)splits * M * N : 0; }

extern "C" int pctts_gemm_tc_splits(int M, int N, int K) {
    const int m_tiles = (M + BM - 1) / BM;
    const int bn = pick_bn(m_tiles, N);
    const int tiles = m_tiles * ((N + bn - 1) / bn);
    const int num_kb = (K + BK - 1) / BK;
    if (tiles >= 74 || num_kb < 8) return 1;
    int splits = 148 / tiles;
    if (splits > num_kb / 2) splits = num_kb / 2;
    if (splits > 32) splits = 32;
    return splits < 1 ? 1 : splits;
}

/* floats of workspace with which pctts_gemm_tc (splits = 1) cuts the m-tiles of a partly filled last wave along K (0: no use) */
extern "C" long pctts_gemm_tc_tail_workspace(int M, int N, int K) {
    const int m_tiles = (M + BM - 1) / BM;
    return tail_workspace(m_tiles, N, pick_bn(m_tiles, N), (K + BK - 1) / BK, 1 << 30, M, M);
}

extern "C" int pctts_gemm_tc(int transA, int transB, int M, int N, int K, float alpha, const void* Ash, long lda, long a_part_stride,
                             const void* Bsh, long ldb, long b_part_stride, float* C, long ldc, const float* bias, const float* aux,
                             const float* addmod, int addrows, int flags, int nterms, int splits, float* workspace, void* stream) {
    if (!Ash || !Bsh || !C || M <= 0 || N <= 0 || K <= 0) return PCTTS_EINVAL;
    if (nterms != 1 && nterms != 3 && nterms != 6) return PCTTS_EINVAL;
    if (flags & PCTTS_EPI_ATOMIC) return PCTTS_EUNSUPPORTED;
    if ((flags & PCTTS_EPI_BIAS) && !bias) return PCTTS_EINVAL;
    if ((flags & PCTTS_EPI_DLRELU_AUX) && !aux) return PCTTS_EINVAL;
    if ((flags & PCTTS_EPI_ADDMOD) && (!addmod || addrows <= 0)) return PCTTS_EINVAL;
    if (splits > 1 && (!workspace || (flags & ~PCTTS_EPI_ACCUM))) return PCTTS_EINVAL;
    const int parts = nterms == 6 ? 3 : (nterms == 3 ? 2 : 1);
    Launch L;
    Params p;
    L.m_tiles = (M + BM - 1) / BM;
    const int bn = pick_bn(L.m_tiles, N);
    // A: stored [M,K] (K-major) or, transposed, [K,M] (MN-major);  B: stored [K,N] (MN-major) or, transposed, [N,K] (K-major)
    plain_geom(p.a, transA ? 1 : 0, BM);
    plain_geom(p.b, transB ? 0 : 1, bn);
    L.a_base = Ash;
    L.a_dims[0] = transA ? M : K, L.a_dims[1] = transA ? K : M, L.a_dims[2] = 1, L.a_dims[3] = parts;
    L.a_strides[0] = lda * 2, L.a_strides[1] = L.a_dims[1] * lda * 2, L.a_strides[2] = a_part_stride * 2;
    L.b_base = Bsh;
    L.b_dims[0] = transB ? K : N, L.b_dims[1] = transB ? N : K, L.b_dims[2] = 1, L.b_dims[3] = parts;
    L.b_strides[0] = ldb * 2, L.b_strides[1] = L.b_dims[1] * ldb * 2, L.b_strides[2] = b_part_stride * 2;
    p.kb_div = 1 << 30;
    p.num_kb = (K + BK - 1) / BK;
    if (splits < 1) splits = 1;
    if (splits > p.num_kb) splits = p.num_kb;
    p.kb_per_split = (p.num_kb + splits - 1) / splits;
    L.splits = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;
    p.nterms = nterms;
    p.C = C, p.ldc = ldc, p.N = N;
    p.seg_len = 1 << 30, p.seg_valid = M, p.nseg = 1;
    p.alpha = alpha, p.bias = bias, p.aux = aux, p.addmod = addmod, p.addrows = addrows, p.flags = flags;
    p.partial = L.splits > 1 ? workspace : nullptr;
    p.crows = M;
    if (L.splits == 1 && !(flags & ~(PCTTS_EPI_BIAS | PCTTS_EPI_LRELU | PCTTS_EPI_DLRELU_AUX | PCTTS_EPI_ACCUM)))
        apply_tail(L, p, bn, workspace);
    return launch_tc(L, p, bn, C, ldc, alpha, (flags & PCTTS_EPI_ACCUM) ? 1 : 0, (cudaStream_t)stream);
}

// 'same' Conv1D forward as a virtual-im2col GEMM.  xsh: bf16 planes of the time-padded input [B, T+W-1, CIN] (row pitch ldx);
// wsh: bf16 planes of the keras kernel [W*CIN, COUT] (row pitch ldw);  y [B*T, COUT] fp32 = epilogue(im2col(x) . w).
// Accumulator rows run over the flat padded row space r = b*(T+W-1) + t'; rows with t' >= T are computed and dropped.
// split of the reduction (taps x channel blocks) over two CTAs per output tile: with COUT = 256 and ~4000 frames the 128 x 128
// tiling gives 66 tiles for 148 SMs, and the 128 x 64 tiling that fills them is bound by the shared-memory operand reads
// (48 % of the tensor peak); two K halves of 128 x 128 tiles fill 132 SMs at the full MMA rate
static int conv1d_fwd_splits(long rows, int COUT, int num_kb) {
    const long m_tiles = (rows + BM - 1) / BM;
    const long t128 = m_tiles * ((COUT + 127) / 128);
    return (COUT >= 128 && t128 < 100 && 2 * t128 <= 148 && num_kb >= 32) ? 2 : 1;
}

/* floats of workspace pctts_conv1d_tc_fwd wants (0: none) */
extern "C" long pctts_conv1d_tc_fwd_workspace(int B, int T, int W, int CIN, int COUT) {
    const int num_kb = ((CIN + BK - 1) / BK) * W;
    return conv1d_fwd_splits((long)B * (T + W - 1), COUT, num_kb) > 1 ? 2L * B * T * COUT : 0;
}

extern "C" int pctts_conv1d_tc_fwd(const void* xsh, long ldx, long x_part_stride, const void* wsh, long ldw, long w_part_stride,
                                   float* y, long ldy, int B, int T, int W, int CIN, int COUT, const float* bias, const float* aux,
                                   int flags, int nterms, float* workspace, void* stream) {
    if (!xsh || !wsh || !y || B <= 0 || T <= 0 || W <= 0 || CIN <= 0 || COUT <= 0) return PCTTS_EINVAL;
    if (nterms != 1 && nterms != 3 && nterms != 6) return PCTTS_EINVAL;
    if (flags & (PCTTS_EPI_ATOMIC | PCTTS_EPI_ADDMOD)) return PCTTS_EUNSUPPORTED;
    if ((flags & PCTTS_EPI_BIAS) && !bias) return PCTTS_EINVAL;
    if ((flags & PCTTS_EPI_DLRELU_AUX) && !aux) return PCTTS_EINVAL;
    const int parts = nterms == 6 ? 3 : (nterms == 3 ? 2 : 1);
    const int Tp = T + W - 1;
    const long rows = (long)B * Tp;
    Launch L;
    Params p;
    L.m_tiles = (int)((rows + BM - 1) / BM);
    const int kb_seg = (CIN + BK - 1) / BK;
    int splits = workspace ? conv1d_fwd_splits(rows, COUT, kb_seg * W) : 1;
    if (splits > 1 && ((flags & ~(PCTTS_EPI_BIAS | PCTTS_EPI_LRELU | PCTTS_EPI_DLRELU_AUX)) || (aux && ldy != COUT))) splits = 1;
    const int bn = splits > 1 ? 128 : pick_bn(L.m_tiles, COUT);
    // A: K-major over the flat padded activation matrix; k-block (w, j): 64 channels from j*64, rows shifted down by w
    p.a.mn_major = 0, p.a.t_div = 1 << 30;
    zero3(p.a.t_lo), zero3(p.a.t_hi), zero3(p.a.k_lo), zero3(p.a.k_hi);
    p.a.t_lo[1] = BM, p.a.k_lo[0] = BK, p.a.k_hi[1] = 1;
    L.a_base = xsh;
    L.a_dims[0] = CIN, L.a_dims[1] = rows, L.a_dims[2] = 1, L.a_dims[3] = parts;
    L.a_strides[0] = ldx * 2, L.a_strides[1] = rows * ldx * 2, L.a_strides[2] = x_part_stride * 2;
    // B: MN-major kernel matrix [W*CIN, COUT]; k-block (w, j) starts at row w*CIN + j*64 (rows past the tap meet zero A columns)
    p.b.mn_major = 1, p.b.t_div = 1 << 30;
    zero3(p.b.t_lo), zero3(p.b.t_hi), zero3(p.b.k_lo), zero3(p.b.k_hi);
    p.b.t_lo[0] = bn, p.b.k_lo[1] = BK, p.b.k_hi[1] = CIN;
    L.b_base = wsh;
    L.b_dims[0] = COUT, L.b_dims[1] = (uint64_t)W * CIN, L.b_dims[2] = 1, L.b_dims[3] = parts;
    L.b_strides[0] = ldw * 2, L.b_strides[1] = L.b_dims[1] * ldw * 2, L.b_strides[2] = w_part_stride * 2;
    p.kb_div = kb_seg;
    p.num_kb = kb_seg * W;
    p.kb_per_split = (p.num_kb + splits - 1) / splits;
    L.splits = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;
    p.nterms = nterms;
    p.C = y, p.ldc = ldy, p.N = COUT;
    p.seg_len = Tp, p.seg_valid = T, p.nseg = B;
    p.alpha = 1.f, p.bias = bias, p.aux = aux, p.addmod = nullptr, p.addrows = 0, p.flags = flags;
    p.partial = L.splits > 1 ? workspace : nullptr;
    p.crows = (long)B * T;
    return launch_tc(L, p, bn, y, ldy, 1.f, 0, (cudaStream_t)stream);
}

// Conv1D weight gradient dw[w*CIN + c, n] (+)= sum_{b,t} xpad[b, t+w, c] dz[b, t, n].
// xsh as above; dzsh: bf16 planes of dz [B*T, COUT] (row pitch lddz).  M-tile = (tap w, 128 channels), k-block = (b, 64 frames).
extern "C" long pctts_conv1d_tc_wgrad_workspace(int B, int T, int W, int CIN, int COUT) {
    const int ct = (CIN + BM - 1) / BM, m_tiles = ct * W;
    return tail_workspace(m_tiles, COUT, pick_bn(m_tiles, COUT), ((T + BK - 1) / BK) * B, ct * BM, CIN, (long)W * CIN);
}

extern "C" int pctts_conv1d_tc_wgrad(const void* xsh, long ldx, long x_part_stride, const void* dzsh, long lddz, long dz_part_stride,
                                     float* dw, int B, int T, int W, int CIN, int COUT, int accumulate, int nterms,
                                     float* workspace, void* stream) {
    if (!xsh || !dzsh || !dw || B <= 0 || T <= 0 || W <= 0 || CIN <= 0 || COUT <= 0) return PCTTS_EINVAL;
    if (nterms != 1 && nterms != 3 && nterms != 6) return PCTTS_EINVAL;
    const int parts = nterms == 6 ? 3 : (nterms == 3 ? 2 : 1);
    const int Tp = T + W - 1;
    Launch L;
    Params p;
    const int ct = (CIN + BM - 1) / BM;          // channel tiles per tap
    L.m_tiles = ct * W;
    const int bn = pick_bn(L.m_tiles, COUT);
    const int kb_seg = (T + BK - 1) / BK;        // k-blocks per utterance
    // A: MN-major view of xpad as (c, t', b); tile (w, ci): channels ci*128.., frames shifted by w
    p.a.mn_major = 1, p.a.t_div = ct;
    zero3(p.a.t_lo), zero3(p.a.t_hi), zero3(p.a.k_lo), zero3(p.a.k_hi);
    p.a.t_lo[0] = BM, p.a.t_hi[1] = 1, p.a.k_lo[1] = BK, p.a.k_hi[2] = 1;
    L.a_base = xsh;
    L.a_dims[0] = CIN, L.a_dims[1] = Tp, L.a_dims[2] = B, L.a_dims[3] = parts;
    L.a_strides[0] = ldx * 2, L.a_strides[1] = (uint64_t)Tp * ldx * 2, L.a_strides[2] = x_part_stride * 2;
    // B: MN-major view of dz as (n, t, b); frames >= T are zero-filled by TMA, which also cancels the A rows read past them
    p.b.mn_major = 1, p.b.t_div = 1 << 30;
    zero3(p.b.t_lo), zero3(p.b.t_hi), zero3(p.b.k_lo), zero3(p.b.k_hi);
    p.b.t_lo[0] = bn, p.b.k_lo[1] = BK, p.b.k_hi[2] = 1;
    L.b_base = dzsh;
    L.b_dims[0] = COUT, L.b_dims[1] = T, L.b_dims[2] = B, L.b_dims[3] = parts;
    L.b_strides[0] = lddz * 2, L.b_strides[1] = (uint64_t)T * lddz * 2, L.b_strides[2] = dz_part_stride * 2;
    p.kb_div = kb_seg;
    p.num_kb = kb_seg * B;
    p.kb_per_split = p.num_kb;
    L.splits = 1;
    p.nterms = nterms;
    p.C = dw, p.ldc = COUT, p.N = COUT;
    p.seg_len = ct * BM, p.seg_valid = CIN, p.nseg = W;
    p.alpha = 1.f, p.bias = nullptr, p.aux = nullptr, p.addmod = nullptr, p.addrows = 0;
    p.flags = accumulate ? PCTTS_EPI_ACCUM : 0;
    p.partial = nullptr;
    p.crows = (long)W * CIN;
    apply_tail(L, p, bn, workspace);
    return launch_tc(L, p, bn, dw, COUT, 1.f, accumulate, (cudaStream_t)stream);
}
