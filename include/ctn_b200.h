/*
 * ctn_b200.h -- C ABI of libctn_b200.so: the B200 (sm_100a) kernels behind the
 * Conv-TasNet separation path.
 *
 * The reference (jwr1995/Conv-TasNet) has no FFI: its hot path is three Python
 * modules (src/conv_tasnet.py, src/pit_criterion.py, src/utils.py).  The
 * drop-in modules of the same names under conv-tasnet_b200/ keep that Python
 * surface and call the functions below through ctypes.  Each entry point cites
 * the reference lines whose arithmetic it replaces.
 *
 * Conventions
 *  - Every pointer is a DEVICE pointer owned by the caller (PyTorch allocates);
 *    the library owns no device memory and keeps no mutable global state.
 *  - `stream` is a cudaStream_t passed as void*.  Calls are stream-ordered and
 *    never synchronise the host, so a whole forward/backward can be captured
 *    in a CUDA graph.
 *  - Return value: 0 on success, <0 on failure (CTN_ERR_*); never throws or
 *    exits.  ctn_last_error() returns a thread-local description.
 *  - Internal activation layout is frames-major ("channels-last"):
 *    act[(m*K + k)*C + c] for utterance m, frame k, channel c, where
 *    K = (T - L)/(L/2) + 1 (conv_tasnet.py:113).  Activations that feed a 1x1
 *    convolution are stored as a bf16 pair (hi, lo) with hi + lo ~= fp32 value
 *    (16 mantissa bits); the tensor-core GEMM computes hi*hi + lo*hi + hi*lo
 *    with fp32 accumulation.
 */
#ifndef CTN_B200_H
#define CTN_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define CTN_OK 0
#define CTN_ERR_INVALID (-1)     /* bad argument */
#define CTN_ERR_UNSUPPORTED (-2) /* shape the kernels do not cover */
#define CTN_ERR_CUDA (-3)        /* CUDA runtime / driver error */

#define CTN_NORM_NONE 0
#define CTN_NORM_GLN 1 /* GlobalLayerNorm, conv_tasnet.py:332-355 */
#define CTN_NORM_CLN 2 /* ChannelwiseLayerNorm, conv_tasnet.py:307-329 */

#define CTN_MASK_RELU 0    /* conv_tasnet.py:205-206 */
#define CTN_MASK_SOFTMAX 1 /* conv_tasnet.py:203-204 */

/* geometry of the statistics buffers the kernels exchange */
#define CTN_ROW_TILE 128 /* GEMM row tile (frames) */
#define CTN_COL_TILE 128 /* statistics column group */
#define CTN_DW_TILE 16   /* depthwise-kernel frame tile */

int ctn_version(void);
/* Operand format the library was built with: 3 = bf16 (hi, lo) pairs, three MMAs per product (the product);
 * 1 = bf16, one MMA and 101 = fp16, one MMA (experiment builds).  The Python layer
 * refuses anything but 3 unless CTN_ALLOW_EXPERIMENT_BUILD=1. */
int ctn_operand_format(void);
const char* ctn_last_error(void);

/* Number of (sum, sumsq) float2 partials per utterance written by
 * ctn_gemm1x1 (out_stat = GLN) and by ctn_dwconv_fwd (norm = GLN). */
int ctn_gemm_gln_partials(int K, int Cout);
int ctn_dwconv_gln_partials(int K, int H, int P, int dilation);
/* Number of per-row column-group partials written by ctn_gemm1x1 (out_stat = CLN). */
int ctn_gemm_cln_groups(int Cout);

/* Weight preparation for one 1x1 convolution W[rows, cols] (fp32, row-major;
 * nn.Conv1d weight [rows, cols, 1], conv_tasnet.py:169,185,217,256).
 *   Wg[r,c] = W[r,c] * gamma[c]   (gamma NULL -> 1)   -> (w_hi, w_lo) bf16 [rows, cols]
 *   col_g[r]   = sum_c (w_hi + w_lo)[r,c]              (NULL to skip)
 *   col_bias[r]= sum_c W[r,c] * beta[c]                (NULL to skip)
 * This folds the norm that precedes the convolution into its weights:
 *   W.(gamma*(y-mu)*rstd + beta) = rstd*(Wg.y) - rstd*mu*col_g + col_bias.
 * transpose != 0 writes (w_hi, w_lo) as [cols, rows] (operand of the dgrad GEMM). */
int ctn_weight_prep(const float* W, int rows, int cols, const float* gamma, const float* beta, int transpose,
                    void* w_hi, void* w_lo, float* col_g, float* col_bias, void* stream);

/* The same for a table of weights in ONE launch.  `items_dev` is a DEVICE array of n_items
 * entries (field meaning as the arguments of ctn_weight_prep); max_rows = largest `rows`. */
typedef struct ctn_prep_item {
    const float* W;
    const float* gamma;
    const float* beta;
    void* w_hi;
    void* w_lo;
    float* col_g;
    float* col_bias;
    int rows, cols, transpose, reserved;
} ctn_prep_item;
int ctn_weight_prep_batch(const ctn_prep_item* items_dev, int n_items, int max_rows, void* stream);

/* Encoder (conv_tasnet.py:97-117): w = relu(Conv1d(1,N,L,stride L/2)(mixture)).
 * mixture [M,T] fp32 -> w_f32 [M*K, N], (w_hi, w_lo) bf16 [M*K, N] and the
 * per-frame statistics of the leading cLN (conv_tasnet.py:167,326-327):
 * rowstat[M*K] = float2(mean, 1/sqrt(var + 1e-8)).  Any of the outputs may be NULL. */
int ctn_encoder_fwd(const float* mixture, const float* U, int M, int T, int N, int L, float* w_f32, void* w_hi,
                    void* w_lo, void* rowstat, void* stream);

typedef struct ctn_gemm_args {
    int M, K;                /* utterances, frames per utterance                         */
    int Cin, Cout;           /* out[r, n] = sum_k a[r, k] * w[n, k]                      */
    const void *a_hi, *a_lo; /* bf16 [M*K, Cin]                                           */
    const void *w_hi, *w_lo; /* bf16 [Cout, Cin] (from ctn_weight_prep)                   */
    float* out_f32;          /* [M*K, Cout] or NULL                                       */
    void *out_hi, *out_lo;   /* bf16 [M*K, Cout] or NULL                                  */
    const float* resid;      /* [M*K, Cout] added to the result, or NULL (may == out_f32) */
    const float* col_g;      /* [Cout] or NULL, see in_norm                               */
    const float* col_bias;   /* [Cout] or NULL                                            */
    /* Folded input normalisation: out = rstd*acc - rstd*mean*col_g + col_bias (+ resid).
     * GLN: in_stat = float2 partials [M, in_P] (sum, sumsq) over in_count elements per utterance.
     * CLN: in_stat = float2 [M*K] (mean, rstd) per frame. */
    int in_norm;
    const void* in_stat;
    int in_P;
    double in_count;
    /* Statistics of the OUTPUT for the next norm (taken after an optional PReLU
     * with slope *stat_prelu; the stored output itself is never activated):
     * GLN: out_stat_buf = float2 [M, ctn_gemm_gln_partials(K, Cout)]
     * CLN: out_stat_buf = float2 [M*K, ctn_gemm_cln_groups(Cout)] */
    int out_stat;
    void* out_stat_buf;
    const float* stat_prelu;
    int force_simt; /* testing: skip the tcgen05 path */
    /* Backward use (pointwise dgrad): with out = du, also emit the gLN backward sums
     * (sum g*du, sum g*du*qh), qh = (prelu(bw_pre) - mean) * rstd, as GLN-layout partials in
     * out_stat_buf (out_stat must be CTN_NORM_GLN).  bw_pre fp32 [M*K, Cout] = saved d,
     * bw_gamma [Cout], bw_slope device scalar, bw_stat float2 [M, bw_P] forward partials. */
    const float* bw_pre;
    const float* bw_gamma;
    const float* bw_slope;
    const void* bw_stat;
    int bw_P;
    /* residual given as a bf16 (hi, lo) pair [M*K, Cout] instead of fp32 (both or neither; exclusive with
     * resid; may alias out_hi / out_lo): inference keeps the block input only in that form */
    const void *resid_hi, *resid_lo;
    /* nonzero (needs stat_prelu): the STORED output is PReLU(out) too, not only the statistics.
     * Inference stores PReLU(conv1x1 output) so that the consumers skip the first PReLU. */
    int store_prelu;
} ctn_gemm_args;

/* 1x1 convolution as a GEMM over all frames of the batch (bottleneck
 * conv_tasnet.py:169; TemporalBlock conv1x1 :217; pointwise :256 with the
 * residual add of :237; mask conv :185).  tcgen05/TMEM/TMA path when
 * Cin % 8 == 0 && Cout % 8 == 0, a plain SIMT kernel otherwise. */
int ctn_gemm1x1(const ctn_gemm_args* args, void* stream);

typedef struct ctn_dwconv_args {
    int M, K, H, P, dilation, causal;
    const float* in;       /* a = conv1x1 output, fp32 [M*K, H]                        */
    const float* prelu_in; /* slope of the first PReLU (device scalar), :218           */
    const float *gamma_in, *beta_in; /* first norm affine [H], :219                    */
    int norm;              /* CTN_NORM_GLN or CTN_NORM_CLN                             */
    const void* in_stat;   /* GLN: float2 [M, in_P] partials; CLN: float2 [M*K, in_P]  */
    int in_P;
    const float* wd;        /* depthwise weight [H, P], :247-250                       */
    const float* prelu_out; /* slope of the second PReLU, :253                         */
    float* out_pre;         /* d = depthwise output before PReLU, fp32 [M*K,H] or NULL */
    void *out_hi, *out_lo;  /* q = prelu(d) as bf16 pair [M*K, H]                      */
    void* out_stat;         /* GLN: float2 [M, ctn_dwconv_gln_partials(K, H, P, dilation)] partials of q;
                               CLN: float2 [M*K] (mean, rstd) of q                     */
} ctn_dwconv_args;

/* First PReLU + norm applied on load, zero padding AFTER normalisation, dilated
 * depthwise cross-correlation (causal = left-only taps, i.e. pad+Chomp1d
 * :175-176,247-252,275-289), second PReLU, statistics for the second norm. */
int ctn_dwconv_fwd(const ctn_dwconv_args* args, void* stream);

typedef struct ctn_dwpw_args {
    int M, K, H, B;         /* a is [M*K, H], y is [M*K, B]; B == 256, H % 64 == 0              */
    int dilation, causal;   /* power-of-two dilation, P = 3                                    */
    const float* a;         /* conv1x1 output, fp32 [M*K, H]                                   */
    const float* prelu_in;  /* slope of the first PReLU (device scalar), conv_tasnet.py:218;
                               NULL: `a` is already PReLU(conv1x1 output) (ctn_gemm1x1 store_prelu) */
    const float *gamma_in, *beta_in; /* gLN-1 affine [H], :219                                 */
    const void* in_stat;    /* float2 [M, in_P]: gLN-1 partials (ctn_gemm1x1 out_stat)         */
    int in_P;
    const float* wd;        /* depthwise weight [H, 3], :247-250                               */
    const float* prelu_out; /* slope of the second PReLU, :253                                 */
    const void *w_hi, *w_lo;/* pointwise weight with gamma-2 folded (ctn_weight_prep), [B, H]  */
    float* y;               /* out: (W2 diag gamma2) q, fp32 [M*K, B]: NOT yet normalised      */
    void* out_stat;         /* out: float2 [M, ctn_dwpw_partials(K, dilation)] partials of q   */
    float* d_out;           /* out (training, NULL otherwise): the depthwise output BEFORE the
                               second PReLU, fp32 [M*K, H], which autograd saves for :253-254   */
} ctn_dwpw_args;

/* Inference fast path of one TemporalBlock's middle: first PReLU + gLN-1 + depthwise conv +
 * second PReLU (conv_tasnet.py:218-253) computed by producer warps straight into the tensor-core
 * operand of the pointwise convolution (:256); q never reaches memory.  gLN-2 (:254) needs the
 * statistics of the whole utterance, so its scale/shift and the residual add (:237) are applied
 * afterwards by ctn_resid_update.  ctn_dwpw_supported() != 0 says whether a configuration can
 * take this path (gLN, P == 3, B == 256, H % 64 == 0, power-of-two dilation); otherwise use
 * ctn_dwconv_fwd + ctn_gemm1x1. */
int ctn_dwpw_supported(int H, int B, int P, int dilation, int norm);
int ctn_dwpw_partials(int K, int dilation);
int ctn_dwpw_fwd(const ctn_dwpw_args* args, void* stream);

/* x += r * y - r * mu * col_g + col_bias on the bf16 (hi, lo) residual stream [M*K, B], in place:
 * (mu, r) = gLN-2 statistics from the P partials per utterance over `count` elements; col_g /
 * col_bias as produced by ctn_weight_prep (conv_tasnet.py:254-256 folded, residual :237). */
int ctn_resid_update(void* x_hi, void* x_lo, const float* y, const void* stat, int P, double count,
                     const float* col_g, const float* col_bias, int M, int K, int B, void* stream);

/* Mask non-linearity, mask * mixture_w, basis Linear(N -> L), overlap-add with
 * step L/2 and right zero-fill to T (conv_tasnet.py:202-209,137-142,56-59;
 * utils.py:9-46).  score fp32 [M*K, C*N] (channel = c*N + n), w fp32 [M*K, N],
 * V fp32 [L, N] -> est [M, C, T].  Scatter-free: every output sample is one
 * thread's sum of its (at most ceil(L/hop)) frame terms in ascending frame order. */
int ctn_decoder_fwd(const float* score, const float* w, const float* V, int M, int K, int N, int L, int C, int T,
                    int mask_nonlinear, float* est, void* stream);

/* utils.py:9-46 overlap_and_add for an arbitrary frame_step <= L:
 * frames [batch, K, L] -> out [batch, (K-1)*step + L]. */
int ctn_overlap_add(const float* frames, int batch, int K, int L, int step, float* out, void* stream);

/* ---- utterance-level PIT SI-SNR loss (pit_criterion.py) ------------------------------ */

/* Number of doubles of workspace ctn_pit_sisnr_fwd needs. */
int ctn_pit_workspace_doubles(int B, int C, int T);

/* cal_si_snr_with_pit (pit_criterion.py:27-76) + the loss of cal_loss (:22).
 * source, estimate fp32 [B, C, T]; lengths int64 [B]; perms int64 [nperm, C] in
 * itertools.permutations order (:66).  Side effect as in the reference (:38):
 * estimate[b, :, t >= lengths[b]] is zeroed IN PLACE.
 * Outputs: max_snr [B] (= [B,1]), max_idx int64 [B] (first maximum on ties),
 * loss [1] = 0 - mean(max_snr); optional pair_snr [B, C, C] (estimate i vs target j, dB)
 * and coef [B, C, 4], the closed-form gradient coefficients ctn_pit_sisnr_bwd consumes. */
int ctn_pit_sisnr_fwd(const float* source, float* estimate, const long long* lengths, int B, int C, int T,
                      const long long* perms, int nperm, double* workspace, float* max_snr, long long* max_idx,
                      float* loss, float* pair_snr, float* coef, void* stream);

/* ctn_pit_sisnr_fwd without the caller-provided workspace, for callers that cannot allocate device memory (code
 * under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  It has NO counterpart in the
 * reference.  Same arguments, outputs and in-place masking
 * side effect as ctn_pit_sisnr_fwd; the chunk partials go to a static __device__ array inside the library of
 * CTN_PIT_STATIC_SCRATCH_DOUBLES doubles (512 KiB, held by every GPU that loads the library until the process
 * exits).  Shapes with ctn_pit_workspace_doubles(B, C, T) above that limit return CTN_ERR_UNSUPPORTED with a
 * message naming the limit.  The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine).
 * All calls on a device share the array: calls that overlap on the GPU (different streams) race; calls ordered
 * on one stream are safe, and nothing needs clearing between them (every slot read is written first by the
 * same call).  The device of the pointers must be current (its copy of the array is resolved on first use). */
#define CTN_PIT_STATIC_SCRATCH_DOUBLES 65536
int ctn_pit_sisnr_fwd_static(const float* source, float* estimate, const long long* lengths, int B, int C, int T,
                             const long long* perms, int nperm, float* max_snr, long long* max_idx, float* loss,
                             float* pair_snr, float* coef, void* stream);

/* reorder_source (pit_criterion.py:79-98): out[b, c] = estimate[b, perms[max_idx[b]][c]]
 * (applies the permutation, not its inverse -- kept as in the reference). */
int ctn_pit_reorder(const float* estimate, const long long* perms, const long long* max_idx, int B, int C, int T,
                    float* out, void* stream);

/* d loss / d estimate for loss = 0 - mean(max_snr) (what autograd derives from
 * pit_criterion.py:37-76): one elementwise pass.  upstream = d(total)/d(loss) [1] or NULL (= 1).
 * estimate is the MASKED estimate that ctn_pit_sisnr_fwd left behind. */
int ctn_pit_sisnr_bwd(const float* source, const float* estimate, const long long* lengths, const long long* perms,
                      const long long* max_idx, const float* coef, const float* upstream, int B, int C, int T,
                      float* grad_estimate, void* stream);

/* ---- backward of the separator (autograd of conv_tasnet.py) ------- */

/* out = split_bf16(gamma * (act(in) - mean) * rstd + beta), act = PReLU(*prelu_slope) or
 * identity (NULL).  Re-materialises a normalised activation as the (hi, lo) operand of a
 * weight-gradient GEMM.  stat: GLN float2 [M, P] partials; CLN float2 [M*K] (mean, rstd). */
int ctn_norm_split(const float* in, const float* prelu_slope, const float* gamma, const float* beta, int norm,
                   const void* stat, int P, int M, int K, int C, void* out_hi, void* out_lo, void* stream);

/* Weight gradient of a 1x1 convolution: G[i, j] = sum_r x[r, i] * y[r, j] over all R rows,
 * x bf16 (hi, lo) [R, Ci], y bf16 (hi, lo) [R, Cj] -> fp32 partials [splits, Ci, Cj]
 * (one per row split; sum them with ctn_reduce_partials).  tcgen05 path with MN-major
 * operands when Ci % 8 == 0 && Cj % 8 == 0. */
int ctn_wgrad_splits(long long R, int Ci, int Cj);
int ctn_wgrad(const void* x_hi, const void* x_lo, const void* y_hi, const void* y_lo, long long R, int Ci, int Cj,
              float* partials, int force_simt, void* stream);

typedef struct ctn_bwd_dw_args {
    int M, K, H, P, dilation, causal;
    int P1, P2, Pb;        /* partial counts of stat1, stat2, bstat per utterance          */
    const float* du;       /* d loss / d u from the pointwise dgrad GEMM, fp32 [M*K, H]    */
    const float* d;        /* saved depthwise output (before PReLU)                        */
    const float* a;        /* saved conv1x1 output (before PReLU)                          */
    const void* stat1;     /* forward gLN partials of prelu(a)                             */
    const void* stat2;     /* forward gLN partials of prelu(d)                             */
    const void* bstat;     /* (sum g2*du, sum g2*du*qh) partials from the dgrad epilogue   */
    const float *g1, *b1, *g2, *wd, *slope1, *slope2;
    float* dn1;            /* out: d loss / d n1, fp32 [M*K, H]                            */
    float* chan_part;      /* out: [M*chan_rows][(4+P)*H]: dg2, db2, dg1, db1, dwd[j]      */
    float* scal_part;      /* out: [M*ctas][4]: sum g1*dn1, sum g1*dn1*ph, d slope2, 0     */
    int norm;              /* CTN_NORM_GLN (0 is read as gLN) or CTN_NORM_CLN.  cLN: stat1 / bstat are
                            * per-frame column-group partials [M*K, P1] / [M*K, Pb], stat2 is the
                            * per-frame (mean, rstd) [M*K]; scal_part[.][0..1] are not produced          */
    /* Optional: u = g2 * qh + b2 (the pointwise convolution's input, the operand of dW2 = g^T u) as bf16
     * (hi, lo) [M*K, H] -- what ctn_norm_split makes of d -- written while d is in registers anyway.  All three
     * NULL: not written.  Only where ctn_bwd_dw_emits_u() says so; an error elsewhere.                          */
    const float* b2;
    void* u_hi;
    void* u_lo;
} ctn_bwd_dw_args;

/* gLN-2 backward + PReLU-2 backward + depthwise backward (input and weight grads) +
 * gLN-1 parameter grads, one pass (gLN models). */
/* Partial-sum geometry per utterance for this (M, P, dilation): rows of scal_part (= CTAs) and
 * rows of chan_part. */
int ctn_bwd_dw_layout(int M, int K, int H, int P, int dilation, int norm, int* ctas, int* chan_rows);
int ctn_bwd_dwconv(const ctn_bwd_dw_args* args, void* stream);
/* 1 if ctn_bwd_dwconv can also write u (hi, lo) for this shape and norm (P = 3, H % 4 == 0, not cLN). */
int ctn_bwd_dw_emits_u(int H, int P, int norm);

/* gLN-1 input gradient + PReLU-1 backward: da (hi, lo) [M*K, H] and per-CTA d slope1 partials
 * [M * ctn_bwd_prelu_norm_tiles(K)]. */
int ctn_bwd_prelu_norm_tiles(int K);
int ctn_bwd_prelu_norm(const float* dn1, const float* a, const void* stat1, int P1, const float* scal_part,
                       int tiles_b, const float* g1, const float* slope1, int M, int K, int H, void* da_hi,
                       void* da_lo, float* alpha_part, void* stream);

/* The same for cLN blocks (per-frame statistics; stat1 = column-group partials [M*K, P1]). */
int ctn_bwd_prelu_norm_cln(const float* dn1, const float* a, const void* stat1, int P1, const float* g1,
                           const float* slope1, int M, int K, int H, void* da_hi, void* da_lo, float* alpha_part,
                           void* stream);

#define CTN_MAX_SEG 12
typedef struct ctn_reduce_args {
    const float* part;      /* [n_part] rows of partial sums                                */
    int n_part;
    long long part_stride;  /* floats between consecutive partial rows                      */
    int nseg;
    int accumulate;         /* 0: dst = sum, 1: dst += sum                                  */
    long long src_off[CTN_MAX_SEG];
    int len[CTN_MAX_SEG];
    int src_stride[CTN_MAX_SEG]; /* 0 -> 1 */
    float* dst[CTN_MAX_SEG];
    int dst_stride[CTN_MAX_SEG]; /* 0 -> 1 */
    /* optional per-segment partial buffers (NULL -> part / n_part / part_stride above), so that the
     * partials of several kernels are reduced by ONE launch */
    const float* seg_part[CTN_MAX_SEG];
    int seg_n_part[CTN_MAX_SEG];
    long long seg_part_stride[CTN_MAX_SEG];
} ctn_reduce_args;

/* dst[seg][i*dst_stride] (+)= sum_p part[p*part_stride + src_off[seg] + i*src_stride]; fixed order. */
int ctn_reduce_partials(const ctn_reduce_args* args, void* stream);

/* Backward of mask non-linearity, mask*w, basis Linear and overlap-add (conv_tasnet.py:137-142,
 * 202-209; utils.py:9-46): d score as bf16 (hi, lo) [M*K, C*N], the decoder's direct gradient
 * on the encoder output dw [M*K, N], and per-CTA partials of dV [M*tiles][L*N]. */
int ctn_decoder_bwd_tiles(int K);
int ctn_decoder_bwd(const float* grad_est, const float* score, const float* w, const float* V, int M, int K, int N,
                    int L, int C, int T, int mask_nonlinear, void* dscore_hi, void* dscore_lo, float* dw,
                    float* dV_part, void* stream);

/* Backward of the leading cLN (conv_tasnet.py:167) + encoder ReLU, merged with dw from the decoder:
 * dpre [M*K, N] and per-CTA partials [M*tiles][2*N] of (d gamma0, d beta0). */
int ctn_cln_bwd_tiles(int K);
int ctn_cln_bwd(const float* dn0, const float* w, const void* rowstat, const float* gamma0, const float* dw_dec,
                int M, int K, int N, float* dpre, float* chan_part, void* stream);

/* dU[n, l] = sum dpre[m,k,n] * mixture[m, k*hop + l] (conv_tasnet.py:106): per-CTA partials [M*tiles][N*L]. */
int ctn_encoder_wgrad_tiles(int K);
int ctn_encoder_wgrad(const float* dpre, const float* mixture, int M, int T, int N, int L, float* dU_part,
                      void* stream);

typedef struct ctn_updc1_args {
    int M, K, B, H;          /* x, y are [M*K, B] (B == 256); a is [M*K, H] (H == 256 or 512)         */
    void *x_hi, *x_lo;       /* in/out: residual stream as a bf16 (hi, lo) pair, updated in place     */
    const float* y;          /* raw pointwise product of the PREVIOUS block (ctn_dwpw_fwd)            */
    const void* in_stat;     /* float2 [M, in_P]: gLN-2 partials of the previous block                */
    int in_P;
    double in_count;         /* K * H of the previous block's depthwise output                        */
    const float *col_g, *col_bias; /* [B] of the previous block's pointwise weight (ctn_weight_prep)  */
    const void *w_hi, *w_lo; /* conv1x1 weight of THIS block, bf16 [H, B]                             */
    float* a;                /* out: PReLU(W1 x), fp32 [M*K, H]                                       */
    void* out_stat;          /* out: float2 [M, ctn_gemm_gln_partials(K, H)] gLN-1 partials of a      */
    const float* slope;      /* slope of this block's first PReLU (device scalar)                     */
    void *x_out_hi, *x_out_lo; /* training (NULL: in place): the updated residual stream goes HERE, the input
                                  pair stays intact (autograd saves every block's input for dW1)       */
    int store_raw;           /* training: `a` = W1 x BEFORE the PReLU (its backward needs the sign and
                                the value); the statistics are those of PReLU(a) either way            */
} ctn_updc1_args;

/* ctn_resid_update of block b-1 fused into the conv1x1 of block b (conv_tasnet.py:254-256, :237 then
 * :217-218): producer warps form x += r y - r mu rowsum + W2 beta2, write it back and feed it straight to
 * the tensor cores; the output is PReLU(conv1x1(x)) with the statistics for gLN-1, exactly what
 * ctn_gemm1x1 (store_prelu) would have produced.  ctn_updc1_supported() != 0: gLN, B == 256, H == 256 or 512. */
int ctn_updc1_supported(int B, int H, int norm);
int ctn_updc1_fwd(const ctn_updc1_args* args, void* stream);

/* ---- training-step tail: clip_grad_norm_ + Adam (solver.py:184-186, train.py:129-131) ---- */
typedef struct ctn_opt_chunk {
    float* param;       /* n parameters                                   */
    float* grad;        /* their gradient (written only if write_grad)    */
    float* exp_avg;     /* Adam first moment                              */
    float* exp_avg_sq;  /* Adam second moment                             */
    int n;
    int pad_;
} ctn_opt_chunk;

/* torch.nn.utils.clip_grad_norm_(params, max_norm) followed by torch.optim.Adam.step() (L2 weight
 * decay, amsgrad off) over a device table of chunks covering every parameter tensor, in two launches.
 * betas are doubles so that 1 - beta is rounded once, as torch does.  step_size = lr / (1 - beta1^t), inv_sqrt_bias2 = 1 / sqrt(1 - beta2^t) for the step count t (host).
 * partials: n_chunks doubles of workspace.  max_norm <= 0 disables clipping.  norm_out (may be NULL)
 * receives the total gradient norm before clipping, like clip_grad_norm_'s return value. */
int ctn_clip_adam(const ctn_opt_chunk* chunks_dev, int n_chunks, double* partials, float max_norm, float step_size,
                  double beta1, double beta2, float inv_sqrt_bias2, float eps, float weight_decay, int write_grad,
                  float* norm_out, void* stream);

/* nn.BatchNorm1d in inference mode (conv_tasnet.py:300-303, reached for any norm_type other than
 * "gLN" / "cLN"): folds (weight, bias, running_mean, running_var, eps) [C] into the per-channel affine
 *   scale = weight / sqrt(running_var + eps),  shift = bias - running_mean * scale
 * (weight / bias NULL -> 1 / 0).  The engine hands (scale, shift) to the kernels in place of a gLN's
 * (gamma, beta), with unit statistics. */
int ctn_bn_fold(const float* weight, const float* bias, const float* running_mean, const float* running_var, float eps,
                int C, float* scale, float* shift, void* stream);

/* ---- nn.BatchNorm1d in TRAINING mode (conv_tasnet.py:300-303; torch's batch_norm with training=True) --------
 * Per channel c over all R = M*K rows of the batch (every frame of every utterance, padded or not):
 *   mean[c], var[c] (biased) normalise;  running_mean / running_var move by `momentum` towards mean and the
 *   UNBIASED variance; num_batches_tracked += 1.  The block kernels then take the folded affine
 *   (scale = weight * rstd, shift = bias - mean * scale) in place of a gLN's (gamma, beta) with unit statistics.
 *
 * ctn_bn_colsums: column sums over the rows of x = PReLU(pre) (slope NULL = identity), pre fp32 [R, C]:
 *   dy == NULL (forward):  sums[0][c] = sum x,   sums[1][c] = sum x*x
 *   dy != NULL (backward): sums[0][c] = sum dy,  sums[1][c] = sum dy * xh,  xh = (x - mean[c]) * rstd[c]
 * Both rows are multiplied by `scale` (1 in the forward).  part: ctn_bn_tiles(R) * 2 * C floats of workspace; sums:
 * 2 * C doubles (fixed-order reduction).  A data-parallel job sums `sums` over its ranks (one small all-reduce)
 * before the next call, which makes utterance-sharded training equal to the single-process batch.  In the
 * backward each rank differentiates its LOCAL mean loss, which enters the global loss with weight
 * w = n_local / n_global (pit_criterion.py:22 over the whole batch): it passes scale = w, so that the all-reduced
 * sums are those of the global loss, and count * w to ctn_bn_bwd_center, which brings the batch-mean terms back
 * to local-loss units (the flat gradient is multiplied by w before its own all-reduce).
 * ctn_bn_train_fold: statistics from sums / count (or from the running buffers when use_running != 0: eval-mode
 *   BatchNorm inside a differentiated forward) -> scale, shift, mean, rstd [C]; updates the three buffers.
 * ctn_bn_bwd_center: dy -= sums[0]/count + xh * sums[1]/count in place (center != 0: the two batch-mean terms
 *   of the BatchNorm backward; the factor weight * rstd is applied by the consuming kernel as its "gamma"), and
 *   d weight[c] = sums[1][c], d bias[c] = sums[0][c] (either may be NULL). */
int ctn_bn_tiles(long long R);
int ctn_bn_colsums(const float* dy, const float* pre, const float* slope, const float* mean, const float* rstd,
                   long long R, int C, double scale, float* part, double* sums, void* stream);
int ctn_bn_train_fold(const double* sums, double count, const float* weight, const float* bias, float eps,
                      float momentum, int use_running, float* running_mean, float* running_var,
                      long long* num_batches_tracked, int C, float* scale, float* shift, float* mean, float* rstd,
                      void* stream);
int ctn_bn_bwd_center(float* dy, const float* pre, const float* slope, const float* mean, const float* rstd,
                      const double* sums, double count, int center, long long R, int C, float* dweight, float* dbias,
                      void* stream);

/* Table-driven device-to-device copies in one launch (state shift of the streaming separator,
 * ctn_b200/_stream.py: the frames a causal model's next chunk reaches back to).  items_dev: DEVICE array;
 * bytes must be a multiple of 4 and the pointers 4-byte aligned (16-byte aligned ranges move as 128-bit
 * words); the ranges of one table must not overlap one another.  max_bytes = largest item (grid sizing). */
typedef struct ctn_copy_item {
    void* dst;
    const void* src;
    long long bytes;
} ctn_copy_item;
int ctn_copy_table(const ctn_copy_item* items_dev, int n_items, long long max_bytes, void* stream);

/* SI-SNR improvement of a batch on the device (evaluate.py:108-144, called per utterance on host
 * numpy arrays at evaluate.py:66-80 after remove_pad): for utterance b with n = lengths[b] valid
 * samples (lengths NULL -> T) and source c,
 *   sisnr[b, c, 0] = SI-SNR(ref[b,c,:n], est[b,c,:n]),  sisnr[b, c, 1] = SI-SNR(ref[b,c,:n], mix[b,:n]),
 *   sisnri[b] = mean_c (sisnr[b,c,0] - sisnr[b,c,1])   (the reference hard-codes c = 0, 1)
 * with both signals made zero-mean over the n samples and eps = 1e-8 placed as in cal_SISNR (:128-144).
 * ref, est fp32 [B, C, T] (est already reordered by the best PIT permutation), mix fp32 [B, T],
 * lengths int64 [B]; outputs fp32, sisnri may be NULL.  fp64 accumulation, deterministic. */
int ctn_sisnri(const float* ref, const float* est, const float* mix, const long long* lengths, int B, int C, int T,
               float* sisnr, float* sisnri, void* stream);

/* Batch assembly on the device: ragged utterances arrive PACKED (valid samples only, one upload) and are
 * laid out padded as the model and the loss expect them.  Replaces pad_list (data.py:295-301) as called by
 * the collate functions (data.py:146-154, :222-225) plus the permute to [B, C, T] (data.py:153-154) and the
 * upload of the padded tensors (solver.py:173-176).
 *   offsets     int64 [B + 1] DEVICE: exclusive prefix sum of the utterance lengths, offsets[B] = total
 *   mix_packed  fp32 [total]          -> mix_pad fp32 [B, T]     (both NULL: skip the mixtures)
 *   src_packed  fp32 [total, C] (T x C per utterance, as load_mixtures_and_sources returns them, data.py:252)
 *                                     -> src_pad fp32 [B, C, T]  (both NULL: skip the sources)
 * Samples at t >= len_b get pad_value (the reference uses 0); lengths above T are cut at T. */
int ctn_pad_batch(const float* mix_packed, const float* src_packed, const long long* offsets, int B, int C, int T,
                  float pad_value, float* mix_pad, float* src_pad, void* stream);

/* Inverse for results (utils.py:49-66 remove_pad, which copies every utterance to the host separately):
 * packed[C * offsets[b] + c * len_b + t] = in[b, c, t] for t < len_b; in fp32 [B, C, T] (C = 1 for [B, T]),
 * packed fp32 [C * total] so that utterance b is one contiguous [C, len_b] block and the valid samples go
 * to the host in one copy. */
int ctn_remove_pad(const float* in, const long long* offsets, int B, int C, int T, float* packed, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CTN_B200_H */
