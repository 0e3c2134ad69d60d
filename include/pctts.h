/*
 * pctts.h -- C ABI of libpctts.so, the B200 (sm_100a) backend of the PercivalTTS WGAN-GP training step.
 *
 * The reference (gillesdegottex/percivaltts) has no FFI of its own on this path: its only process->device boundary
 * is Keras `train_on_batch` / `predict` inside percivaltts/backend_tensorflow.py's global tf.Session
 * (backend_tensorflow.py:38-39; call sites optimizertts_wgan.py:220,236, modeltts.py:68-69).  libpctts.so replaces the
 * TensorFlow/cuDNN library calls that session dispatched; each entry point below names the reference op it stands for.
 * INTEGRATION.md shows the ctypes stub a maintainer would add to bind them.
 *
 * Conventions: plain pointers and sizes only; every pointer is a DEVICE pointer unless stated; tensors are fp32,
 * batch-major, channels innermost ([B,T,feat], NHWC maps [N,T,F,C], Keras kernel layouts [in,out] / [W,in,out] /
 * [KH,KW,in,out]); caller owns all buffers; calls are asynchronous on `stream` (a cudaStream_t passed as void*);
 * return value 0 = ok, >0 = cudaError_t, <0 = PCTTS_E* argument error.  Nothing allocates.
 */
#ifndef PCTTS_H
#define PCTTS_H

#ifdef __cplusplus
extern "C" {
#endif

#define PCTTS_ABI_VERSION 1

/* epilogue flags shared by the GEMM and conv2d entry points */
#define PCTTS_EPI_BIAS 1        /* + bias[n]                                   (Dense/Conv use_bias=True)            */
#define PCTTS_EPI_LRELU 2       /* LeakyReLU(alpha=0.3)                        networktts.py:62,119,125               */
#define PCTTS_EPI_SIGMOID 4     /* sigmoid                                     modeltts_common.py:121                 */
#define PCTTS_EPI_DLRELU_AUX 8  /* * LeakyReLU'(aux) with aux = the layer's stored OUTPUT (y>0 <=> x>0)               */
#define PCTTS_EPI_ACCUM 16      /* C += result                                                                        */
#define PCTTS_EPI_ATOMIC 32     /* atomicAdd(C, result) (overlapping virtual rows: Conv1D data gradient)              */
#define PCTTS_EPI_WFLIP 128     /* Conv2D (tensor-core kernel): w is the FORWARD kernel [KH][KW][cout][cin]; the data-gradient
                                   kernel w[KH-1-kt][KW-1-kf][co][ci] is formed on the fly                             */
#define PCTTS_EPI_ADDMOD 64     /* + addmod[m % addrows, n]  (critic: ctx-branch term shared by the 3 applications)   */

int pctts_abi_version(void);
/* 0 when device `dev` is compute capability 10.x (B200); else PCTTS_EUNSUPPORTED / cudaError. Host call. */
int pctts_device_check(int dev);
/* number of kernels this library has launched (or recorded into a CUDA graph under capture) so far. Host call. */
long pctts_launch_count(void);
/* programmatic dependent launch between the library's kernels (default off; env PCTTS_PDL=1 turns it on). Host call. */
int pctts_set_pdl(int on);

/* ---- dense contractions: keras Dense (networktts.py:59-63,196-197), Conv1D as virtual-im2col GEMM (networktts.py:116-120),
 *      LSTM input projections (networktts.py:90-94), and all their data / weight gradients ------------------------------- */
/* C = epilogue(alpha * op(A) op(B)); op(A) is MxK, op(B) is KxN.  A stored matrix X is addressed as
 * X[(r / segX) * segstrideX + (r % segX) * ldX + c] (segX <= 0: X[r*ldX + c]); `aux` uses C's addressing. */
int pctts_gemm_f32(int transA, int transB, int M, int N, int K, float alpha,
                   const float* A, long lda, int segA, long segstrideA,
                   const float* B, long ldb, int segB, long segstrideB,
                   float* C, long ldc, int segC, long segstrideC,
                   const float* bias, const float* aux, const float* addmod, int addrows,
                   int flags, void* stream);

/* Split-K form for long reductions with few output tiles (weight gradients dW = X^T dZ: M,N ~ 256, K = B*T rows).
 * `workspace` holds splits*M*N floats; flags: 0 or PCTTS_EPI_ACCUM; two-stage, deterministic. */
int pctts_gemm_splitk_f32(int transA, int transB, int M, int N, int K, float alpha,
                          const float* A, long lda, int segA, long segstrideA,
                          const float* B, long ldb, int segB, long segstrideB,
                          float* C, long ldc, int segC, long segstrideC,
                          int flags, int splits, float* workspace, void* stream);

/* ---- the same contractions on the 5th-generation tensor cores (tcgen05.mma, TMEM accumulators, TMA-fed) ----------------------
 * fp32 operands are carried as bf16 planes [hi | lo] with x = hi + lo + O(2^-17 |x|) (pctts_split_bf16); nterms = 3 forms
 * Ahi.Bhi + Ahi.Blo + Alo.Bhi in one fp32 accumulator (fp32-grade products), nterms = 1 is plain bf16 (hi planes only).
 * A shadow is `parts` planes of [rows, ld_dst] bf16 (ld_dst % 8 == 0, plane pitch part_stride elements, 16-byte aligned). */
int pctts_split_bf16(const float* src, long ld, int seg, long segstride, long rows, int cols, void* dst, long ld_dst,
                     long part_stride, int parts, void* stream);
/* up to PCTTS_SPLIT_MAX_JOBS plain matrices (no segments) in ONE launch: the weight shadows of a network after its Adam update */
#define PCTTS_SPLIT_MAX_JOBS 24
typedef struct pctts_split_job {
    const float* src;
    void* dst;
    long ld, rows, ld_dst, part_stride;
    int cols;
} pctts_split_job;
int pctts_split_bf16_multi(const pctts_split_job* jobs, int njobs, int parts, void* stream);
int pctts_gemm_tc_splits(int M, int N, int K);                 /* host call: recommended split-K factor (1 = none) */
long pctts_gemm_tc_workspace(int M, int N, int splits);        /* host call: floats of split-K workspace */
long pctts_gemm_tc_tail_workspace(int M, int N, int K);        /* host call: floats with which a splits = 1 call cuts the m-tiles of a
                                                                  partly filled last wave of CTAs along K (0: no use for it) */
int pctts_gemm_tc_set_trace(void* buf);  /* developer probe: 64*8 int64 clock stamps of CTA (0,0,0) of the next launches; NULL = off */
/* C[M,N] = epilogue(alpha * op(A) op(B)) from the shadows of the STORED matrices (A: [M,K], or [K,M] when transA; B: [K,N], or
 * [N,K] when transB); C / aux fp32 with row pitch ldc; flags as for pctts_gemm_f32 except ATOMIC; splits > 1 needs the workspace
 * and allows only PCTTS_EPI_ACCUM (deterministic two-stage sum). */
int pctts_gemm_tc(int transA, int transB, int M, int N, int K, float alpha, const void* Ash, long lda, long a_part_stride,
                  const void* Bsh, long ldb, long b_part_stride, float* C, long ldc, const float* bias, const float* aux,
                  const float* addmod, int addrows, int flags, int nterms, int splits, float* workspace, void* stream);
/* Conv1D 'same' (networktts.py:116-120) without materialising im2col: xsh = shadow of the time-padded input [B*(T+W-1), CIN],
 * wsh = shadow of the keras kernel [W*CIN, COUT]; y [B*T, COUT] (row pitch ldy) = epilogue(conv); aux uses y's addressing. */
long pctts_conv1d_tc_fwd_workspace(int B, int T, int W, int CIN, int COUT);   /* floats (0: none needed); host call */
int pctts_conv1d_tc_fwd(const void* xsh, long ldx, long x_part_stride, const void* wsh, long ldw, long w_part_stride, float* y,
                        long ldy, int B, int T, int W, int CIN, int COUT, const float* bias, const float* aux, int flags, int nterms,
                        float* workspace, void* stream);
/* dw[W*CIN, COUT] (+)= sum_{b,t} xpad[b,t+w,c] dz[b,t,n];  dzsh = shadow of dz [B*T, COUT]; workspace (or NULL): the floats
 * pctts_conv1d_tc_wgrad_workspace asks for (tail split of the last wave of tiles, see pctts_gemm_tc_tail_workspace) */
long pctts_conv1d_tc_wgrad_workspace(int B, int T, int W, int CIN, int COUT);
int pctts_conv1d_tc_wgrad(const void* xsh, long ldx, long x_part_stride, const void* dzsh, long lddz, long dz_part_stride, float* dw,
                          int B, int T, int W, int CIN, int COUT, int accumulate, int nterms, float* workspace, void* stream);

/* ---- Conv2D 'same', stride 1, NHWC, HWIO (networktts.py:122-126; modeltts_common.py:98,100; networks_critic.py:67) ------ */
/* y[n,t,f,:] = epilogue(sum_{kt,kf,ci} x[n,t+kt-pad_t,f+kf-pad_f,ci] w[kt,kf,ci,:]).  xpitch/ypitch/auxpitch = floats between
 * consecutive t rows, *batch = floats between consecutive n (lets the first / last layer read / write the spectral columns
 * of a [B,T,1+S+NM] feature tensor in place, networks_critic.py:58, modeltts_common.py:101,123).  The data gradient is this
 * same call with pctts_conv2d_flip_weights() kernels, swapped channel counts and pad = K-1-pad. */
int pctts_conv2d_fwd(const float* x, const float* w, const float* bias, const float* aux, float* y, int NB, int T, int F,
                     int CIN, int COUT, int KH, int KW, int pad_t, int pad_f, long xpitch, long ypitch, long auxpitch,
                     long xbatch, long ybatch, long auxbatch, int flags, void* stream);
/* The same convolution for CIN = COUT = 4, 5x5, pad 2 as an implicit GEMM on the tensor cores (tcgen05.mma + TMEM; the input tile
 * is staged once in shared memory as bf16 planes and read in place as the im2col operand).  nterms as for pctts_gemm_tc.
 * Bases must be 16-byte aligned and pitches multiples of 4 floats; pctts_conv2d_tc_supported() is a host call. */
int pctts_conv2d_tc_set_trace(void* buf);   /* developer probe: 32*8 int64 clock stamps of CTA 0 of the next launches; NULL = off */
int pctts_conv2d_tc_supported(int F, int CIN, int COUT, int KH, int KW, int pad_t, int pad_f);
int pctts_conv2d_tc_fwd(const float* x, const float* w, const float* bias, const float* aux, float* y, int NB, int T, int F,
                        int CIN, int COUT, int KH, int KW, int pad_t, int pad_f, long xpitch, long ypitch, long auxpitch,
                        long xbatch, long ybatch, long auxbatch, int flags, int nterms, void* stream);
/* the same, fused with the BatchNormalization layers around it (networktts.py:122-126): in_mean / in_invstd / in_gamma / in_beta
 * (all or none; 4 floats each) = apply LeakyReLU(BN(.)) of the previous layer to x on load (x is its pre-BN map; the activation
 * map is never written); workspace (or NULL) = a column-reduction workspace for rows = NB*T*F, C = 4 receiving what
 * pctts_bn_stats(y) would produce; final_stage = 0 stops after the per-CTA partials (pctts_conv2d_tc_stat_partials of them), which
 * pctts_bn_finalize_partials then consumes.  PCTTS_EUNSUPPORTED for maps too small to own 148 partial slots. */
int pctts_conv2d_tc_stat_partials(int NB, int T, int F);   /* host call */
int pctts_conv2d_tc_fwd_fused(const float* x, const float* w, const float* bias, const float* aux, float* y, int NB, int T, int F,
                              int CIN, int COUT, int KH, int KW, int pad_t, int pad_f, long xpitch, long ypitch, long auxpitch,
                              long xbatch, long ybatch, long auxbatch, int flags, int nterms, const float* in_mean,
                              const float* in_invstd, const float* in_gamma, const float* in_beta, void* workspace, int final_stage,
                              void* stream);
/* A chain of up to PCTTS_CHAIN_MAX_LAYERS Conv2D 5x5 'same' layers (<= 4 channels) in ONE launch (conv2d_chain.cu): a sample is owned
 * by a thread-block cluster, the map tile stays in shared memory as bf16 operand planes from layer to layer, every layer is a
 * tcgen05 implicit GEMM, the fp32 output of every layer is written once.  The critic's BatchNorm-free stack
 * (networks_critic.py:66-68) forward, its data-gradient chain and the forward-mode sweep of the gradient penalty.
 * x [NB,T,F,cin0] (cin0 = 1 or 4); layer l: w [5][5][cin][cout] (flip = 1: w is the forward kernel [5][5][cout][cin] and the data
 * gradient kernel is formed on the fly), y [NB,T,F,cout] with cout = 4 (cout = 1 allowed for the last layer), optional bias,
 * optional LeakyReLU' mask from aux [NB,T,F,4] (aux may alias y), flags PCTTS_EPI_BIAS | LRELU | DLRELU_AUX | ACCUM (last layer).
 * pitches / batch strides in floats.  PCTTS_EUNSUPPORTED when pctts_conv2d_chain_supported() is 0 or an alignment rule fails. */
#define PCTTS_CHAIN_MAX_LAYERS 9
typedef struct pctts_chain_layer {
    const float* w;
    const float* bias;
    const float* aux;
    float* y;
    long ybatch, ypitch, auxbatch, auxpitch;
    int cin, cout, flags, flip;
} pctts_chain_layer;
int pctts_conv2d_chain_set_trace(void* buf);  /* developer probe: PCTTS_CHAIN_MAX_LAYERS*16 int64 clock stamps of CTA 0; NULL = off */
int pctts_conv2d_chain_supported(int T, int F, int nterms);   /* host call: CTAs per sample (cluster size), 0 = unsupported */
int pctts_conv2d_chain(const float* x, long xbatch, long xpitch, int cin0, int NB, int T, int F, int nlayers,
                       const pctts_chain_layer* layers, int nterms, void* stream);
int pctts_conv2d_flip_weights(const float* w, float* wt, int KH, int KW, int CIN, int COUT, void* stream);
long pctts_conv2d_wgrad_workspace(int NB, int T, int CIN, int COUT, int KH, int KW); /* floats; host call */
/* dw[kt,kf,ci,co] (+)= scale * sum_{n,t,f} x[n,t+kt-pad_t,f+kf-pad_f,ci] dy[n,t,f,co]   (deterministic two-stage) */
int pctts_conv2d_wgrad(const float* x, const float* dy, float* dw, float* workspace, int NB, int T, int F, int CIN, int COUT,
                       int KH, int KW, int pad_t, int pad_f, long xpitch, long dypitch, long xbatch, long dybatch, float scale,
                       int accumulate, void* stream);

/* dw as above and, in the same pass, dbias[co] = sum_{n < nb_bias, t, f} dy[n,t,f,co] (dbias may be NULL).  4 -> 4 channels, 5x5,
 * pad 2, 16-byte aligned bases and pitches that are multiples of 4 floats only; otherwise PCTTS_EUNSUPPORTED. */
int pctts_conv2d_wgrad_bias(const float* x, const float* dy, float* dw, float* dbias, int nb_bias, float* workspace, int NB, int T,
                            int F, int CIN, int COUT, int KH, int KW, int pad_t, int pad_f, long xpitch, long dypitch, long xbatch,
                            long dybatch, float scale, int accumulate, void* stream);

/* The same gradients for up to PCTTS_WGRAD_MAX_LAYERS layers of a 5x5 'same' stack in ONE launch (+ one reduce launch): the CTAs of
 * the grid are divided between the layers.  Layer l: x [NB,T,F,cin] (cin = 4, or 1 for the first layer of the stack), dy [NB,T,F,4],
 * dw [5][5][cin][4] (+)= scale * sum, dbias [4] = sum over the first nb_bias samples (or NULL).  16-byte aligned dy (and x when
 * cin = 4), pitches / batch strides in floats and multiples of 4; workspace of pctts_conv2d_wgrad_multi_workspace(nlayers) floats. */
#define PCTTS_WGRAD_MAX_LAYERS 9
typedef struct pctts_wgrad_layer {
    const float* x;
    const float* dy;
    float* dw;
    float* dbias;
    long xbatch, xpitch, dybatch, dypitch;
    int cin;
} pctts_wgrad_layer;
long pctts_conv2d_wgrad_multi_workspace(int nlayers);   /* floats; host call */
int pctts_conv2d_wgrad_multi(const pctts_wgrad_layer* layers, int nlayers, int nb_bias, float* workspace, int NB, int T, int F,
                             float scale, int accumulate, void* stream);

/* ---- BatchNormalization(axis=-1) + LeakyReLU (networktts.py:61-62,118-119,124-125), data [rows, C] ------------------- */
long pctts_colreduce_workspace_bytes(long rows, int C);            /* host call; zero the workspace once after allocation */
void* pctts_colreduce_sums_ptr(void* workspace, long rows, int C); /* host call; double sums[2][C] inside the workspace
                                                                      (all-reduce these across ranks for sync-BN) */
int pctts_bn_stats(const float* x, long rows, int C, void* workspace, void* stream);  /* sums = [sum x, sum x^2] */
/* single replica (no all-reduce between the statistics and their use): first stage only (pctts_colreduce_nblk partials), then
 * second stage + finalize in ONE launch -- bit-identical to pctts_bn_stats + pctts_bn_finalize */
int pctts_bn_stats_partial(const float* x, long rows, int C, void* workspace, void* stream);
int pctts_colreduce_nblk(long rows, int C);   /* host call */
int pctts_bn_finalize_partials(void* workspace, long rows, int C, int nblk, double count, float eps, float momentum, float* mean,
                               float* invstd, float* moving_mean, float* moving_var, void* stream);
int pctts_bn_finalize(void* workspace, long rows, int C, double count, float eps, float momentum, float* mean, float* invstd,
                      float* moving_mean, float* moving_var, void* stream);          /* moving_* may be NULL (frozen) */
int pctts_bn_apply(const float* x, float* y, const float* mean, const float* invstd_or_var, const float* gamma,
                   const float* beta, long rows, int C, float eps, int use_var, int lrelu, void* stream);
/* pctts_bn_apply / pctts_bn_bwd_apply that also write the bf16 operand planes of their output (what pctts_split_bf16 would write:
 * [parts][rows][ld_dst] bf16, plane stride part_stride elements) when the consumer is a tensor-core GEMM.  C % 4 == 0, C <= 1024 and
 * 16-byte aligned maps only (PCTTS_EUNSUPPORTED otherwise). */
int pctts_bn_apply_planes(const float* x, float* y, const float* mean, const float* invstd_or_var, const float* gamma,
                          const float* beta, long rows, int C, float eps, int use_var, int lrelu, void* planes, long ld_dst,
                          long part_stride, int parts, void* stream);
int pctts_bn_bwd_apply_planes(const float* x, const float* dy, float* dx, const float* mean, const float* invstd,
                              const float* gamma, const float* beta, long rows, int C, double count, int lrelu, void* workspace,
                              void* planes, long ld_dst, long part_stride, int parts, void* stream);
int pctts_bn_bwd_reduce(const float* x, const float* dy, const float* mean, const float* invstd, const float* gamma,
                        const float* beta, long rows, int C, int lrelu, void* workspace, void* stream);
int pctts_bn_bwd_reduce_partial(const float* x, const float* dy, const float* mean, const float* invstd, const float* gamma,
                                const float* beta, long rows, int C, int lrelu, void* workspace, void* stream);  /* first stage only */
int pctts_bn_bwd_apply(const float* x, const float* dy, float* dx, const float* mean, const float* invstd, const float* gamma,
                       const float* beta, long rows, int C, double count, int lrelu, float* dgamma, float* dbeta,
                       float grad_scale, int accumulate, void* workspace, void* stream); /* dgamma/dbeta *= grad_scale
                       (1/world when the sums were all-reduced and the flat gradient is summed over ranks afterwards) */
int pctts_colsum(const float* x, long rows, int C, float* out, int accumulate, void* workspace, void* stream);

/* pctts_colsum without a workspace argument.  No counterpart in the reference; it exists for hosts that are not allowed to
 * create device buffers of their own beyond the tensors they were handed: a plugin running inside another framework's
 * allocator (which may forbid cudaMalloc while it owns the device or while a stream capture is open), a real-time or
 * embedded caller whose memory budget was fixed at start-up, or a binding from a language runtime that can pass pointers but
 * has no CUDA allocator.  The scratch pctts_colsum asks its caller for sits instead in a fixed array inside libpctts.so:
 * about 4.9 MB of device memory on every GPU that loads the library, held until the process exits.
 *   - x [rows, C] and out [C] are still the caller's device buffers; same result as pctts_colsum, bit for bit.
 *   - Limit: C <= PCTTS_COLSUM_STATIC_MAX_C (any number of rows).  A wider matrix returns PCTTS_ELIMIT (-3) and launches nothing.
 *   - The call runs on the CURRENT device; `stream` must belong to it.  The first call on each device looks the array's device
 *     address up (cudaGetSymbolAddress); make that first call outside any stream capture.  Neither it nor any later call
 *     calls cudaMalloc / cudaFree or any other allocating routine.
 *   - All calls on one GPU share the one array, and it keeps whatever the previous call left in it.  Nothing in it has to start
 *     at zero (each call overwrites every slot it later reads), but two calls whose kernels can overlap on the GPU -- different
 *     streams, or different host threads on different streams -- race on it: order them (same stream, or an event). */
#define PCTTS_COLSUM_STATIC_MAX_C 1024
#define PCTTS_ELIMIT (-3)
int pctts_colsum_static(const float* x, long rows, int C, float* out, int accumulate, void* stream);

/* ---- LSTM / Bidirectional(LSTM) recurrence (networktts.py:72-96): persistent cluster kernel, U resident in smem -------- */
/* zx* = x.W + b  [B,T,4H] (gate order i,f,c,o); U* [H,4H]; hout [B,T,ldh] (direction d -> columns d*H..); the gates and c buffers are
 * written when non-NULL (needed by the backward).  ndir = 1 (LSTM) or 2 (Bidirectional: direction 1 runs time-reversed).
 * H in {32,64,128,256}: register-resident weights, clusters of H/32 CTAs synchronised by DSMEM st.async + mbarrier transaction counts
 * (cluster_size / batch_chunk ignored); other H, or cluster_size = -n: shared-memory-resident weights, clusters of n, cluster barriers. */
int pctts_lstm_fwd(const float* zx0, const float* zx1, const float* U0, const float* U1, float* hout, float* gates0,
                   float* gates1, float* c0, float* c1, int B, int T, int H, int ndir, int ldh, int cluster_size,
                   int batch_chunk, void* stream);
/* BPTT: dz* [B,T,4H] = gradients wrt the gate pre-activations; weight gradients follow as GEMMs on dz. */
int pctts_lstm_bwd(const float* dhout, const float* U0, const float* U1, const float* gates0, const float* gates1,
                   const float* c0, const float* c1, float* dz0, float* dz1, int B, int T, int H, int ndir, int ldh,
                   int cluster_size, int batch_chunk, void* stream);

/* ---- WGAN-GP pieces (optimizertts_wgan.py:44-79) ------------------------------------------------------------------------ */
/* s0 [3B,T,S] = [real | fake | alpha*real+(1-alpha)*fake] restricted to feature columns off..off+S-1  (:44-51, critic :58) */
int pctts_critic_pack_inputs(const float* real, const float* fake, const float* alpha, float* s0, int B, int T, int O, int S,
                             int off, int skip_real, void* stream);   /* skip_real: the first block is already in place */
int pctts_fill_segments(float* p, long n0, float v0, long n1, float v1, long n2, float v2, void* stream);
int pctts_sum_scaled(const float* x, long n, double scale, float* losses, int slot, void* stream);   /* wasserstein_loss :70-71 */
int pctts_gp_norms(const float* g, int B, long per_sample, float* ss, void* stream);                 /* :59-64 */
int pctts_gp_scale(const float* g, const float* ss, float* r0, int B, long per_sample, float lambda, int Bglobal,
                   float* losses, int gp_slot, void* stream);                                        /* :66-68 and d/dg */
int pctts_loss_combine(float* losses, int out, int a, float ca, int b, float cb, int c, float cc, void* stream);
long pctts_lse_workspace_bytes(void);  /* host call; zero once */
int pctts_lse_loss_grad(const float* pred, const float* real, const float* w, float* dpred, long rows, int O,
                        double scale_loss, float scale_grad, float* losses, int slot, void* workspace, void* stream);
/* ... restricted to the columns [c0, c0 + nc) (w indexed by the absolute column): the generator step evaluates the F0 column on the
 * F0 lane, behind the LSTM; give concurrent calls their own workspace */
int pctts_lse_loss_grad_cols(const float* pred, const float* real, const float* w, float* dpred, long rows, int O, int c0, int nc,
                             double scale_loss, float scale_grad, float* losses, int slot, void* workspace, void* stream); /* :73-79 */
/* Keras Adam (:145,172): reads *step, uses t=*step+1, then increments *step (device int) */
int pctts_adam(float* p, const float* g, float* m, float* v, long n, float lr, float beta1, float beta2, float eps, int* step,
               void* stream);

/* ---- small movers -------------------------------------------------------------------------------------------------------- */
int pctts_umma_probe(int layout, int N, int nmma, int shift_rows, long long* out, void* stream);   /* developer probe: tcgen05.mma issue rate per operand layout (umma_probe.cu) */
int pctts_stamp(unsigned long long* slot, void* stream);   /* developer probe: %globaltimer (ns) at this point of the stream */
int pctts_add(float* out, const float* a, const float* b, long n, void* stream);
int pctts_add3(float* out, const float* a, const float* b, const float* c, long n, void* stream);   /* out = (a + b) + c */
int pctts_sigmoid_bwd(const float* dy, long lddy, const float* y, long ldy, float* dz, long lddz, long rows, int cols, void* stream);
int pctts_copy2d(float* dst, long ldd, const float* src, long lds, long rows, int cols, int accumulate, void* stream);
int pctts_pad_time(float* dst, const float* src, int B, int T, int C, int left, int right, void* stream);
int pctts_uniform(float* out, int n, unsigned long long seed, unsigned long long* counter, void* stream); /* K.random_uniform :50 */

/* ---- PML band compression (vocoders.py:134-166,202; pulsemodel/analysis.py:336,366; tables from sigproc.py) -------------- */
/* out[t,k] = (1-w[k]) f(in[t,idx[k]]) + w[k] f(in[t,idx[k]+1]),  f = log|.| when log_abs else identity */
int pctts_pml_linbnd2fwbnd(const float* in, const int* idx, const float* wgt, float* out, long T, int nbins, int K, int log_abs,
                           void* stream);
/* out[t,i] = interp(in[t,:], idx[i] + w[i]) linear (smooth=0) or Catmull-Rom (smooth=1); exp() applied when do_exp */
int pctts_pml_fwbnd2linbnd(const float* in, const int* idx, const float* wgt, float* out, long T, int K, int nbins, int smooth,
                           int do_exp, void* stream);

/* ---- generation tail (modeltts.py:154-165 denormalise; vocoders.py:194-202 synthesis front half) ---------------------------- */
/* v = cmp * std + mean per column (mean = std = NULL: none); out [T,O] = v (optional), f0 [T] = exp(v[:,0]), spec [T,S] = v[:,1:1+S],
 * nm [T,NM] = v[:,1+S:1+S+NM] (each optional): contiguous inputs for pctts_pml_fwbnd2linbnd */
int pctts_denorm_split(const float* cmp, const float* mean, const float* stdv, long T, int O, int S, int NM, float* out, float* f0,
                       float* spec, float* nm, void* stream);

/* ---- data-parallel exchange of the sync-BatchNorm sums over NVLink peer memory ------------------------------ */
/* One-shot all-reduce (sum, rank order => bit-identical on every rank) of n <= nmax doubles: every rank owns a zero-initialised
 * symmetric buffer of pctts_peer_allreduce_bytes(world, nmax) bytes mapped into all peers; peer_bases (HOST array) holds the
 * address of rank r's buffer in this process.  All ranks must issue the same calls in the same order.  *err (device) becomes 1
 * if a peer did not arrive within ~60 s (the caller must read it back: the sums are then invalid).  Asynchronous, capturable. */
long pctts_peer_allreduce_bytes(int world, int nmax);
int pctts_peer_allreduce_f64(double* x, int n, const unsigned long long* peer_bases, int rank, int world, int nmax, int* err,
                             void* stream);

/* sync-BatchNorm with the replica exchange inside the launch: second reduction stage over the nblk first-stage partials
 * [nblk][2][C] (doubles), all-reduce over the replicas through a second symmetric buffer of pctts_bn_sync_bytes(world) bytes
 * (peer_bases as above; world = 1 and peer_bases = NULL: no exchange), then mean / invstd / moving averages (finalize) or
 * sums + dbeta / dgamma (bwd).  Bit-identical to second stage + pctts_peer_allreduce_f64 + finalize as separate launches. */
long pctts_bn_sync_bytes(int world);
int pctts_bn_sync_finalize(const double* partial, double* sums, int nblk, int C, double count, float eps, float momentum,
                           float* mean, float* invstd, float* moving_mean, float* moving_var,
                           const unsigned long long* peer_bases, int rank, int world, int* err, void* stream);
int pctts_bn_sync_bwd(const double* partial, double* sums, int nblk, int C, float* dgamma, float* dbeta, float scale,
                      int accumulate, const unsigned long long* peer_bases, int rank, int world, int* err, void* stream);

/* Gradient all-reduce fused with the Adam update (peer.cu): the flat gradient buffer of a network lives in symmetric memory; one
 * launch = ready barrier, rank-ordered reduce-scatter over peer loads, all-gather over peer stores (in place), done barrier, Keras
 * Adam on the summed gradient, *step += 1; losses [nloss <= 8] (or NULL) are summed over the ranks on the way.  grad_bases / ctl_bases: HOST arrays of `world` mapped addresses (gradient buffers of
 * npad floats, npad % 4 == 0, zero beyond n; control blocks of pctts_grad_sync_ctl_bytes() bytes, zeroed once).  world == 1: Adam only. */
long pctts_grad_sync_ctl_bytes(void);
int pctts_grad_sync_adam(const unsigned long long* grad_bases, const unsigned long long* ctl_bases, int rank, int world, float* p,
                         float* m, float* v, long n, long npad, float lr, float beta1, float beta2, float eps, int* step,
                         float* losses, int nloss, int* err, void* stream);

#ifdef __cplusplus
}
#endif
#endif
