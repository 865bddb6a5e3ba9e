/* b3d -- C ABI of the B200 (sm_100a) kernels behind the X-UNet training step and the
 * stochastic-conditioning sampler.  Plain pointers and sizes only; no torch types.
 *
 * Conventions (all entry points):
 *   - device pointers unless the name says `host`; activations are bf16 NHWC ("pixels x channels"),
 *     images n = 2*b + frame (the reference flattens 'b f c h w -> (b f) c h w', xunet.py:494);
 *   - `stream` is a cudaStream_t passed as void*; nothing synchronises, nothing allocates and
 *     nothing is retained across calls; workspaces are supplied by the caller;
 *   - return 0 on success, <0 on error (-1 bad argument, -2 CUDA error, -3 unsupported);
 *     b3d_last_error() returns the message of the last failure on the calling thread.
 *
 * The reference (/root/reference) has no FFI: the interface each entry point replaces is a span
 * of Python/ATen calls, cited per function as file:line.
 */
#ifndef B3D_H
#define B3D_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* b3d_last_error(void);
int b3d_version(void);
/* number of kernels launched by this library since process start (bench.py's gpu_launches) */
int64_t b3d_launch_count(void);
/* Programmatic dependent launch of the frequent kernels (hides launch latency between short dependent launches; pays off
 * for small per-GPU batches).  On by default; B3D_NO_PDL=1 in the environment wins.  No reference counterpart. */
void b3d_set_pdl(int on);

/* ------------------------------------------------------------------------------------------
 * Tensor-core path: one tcgen05/TMEM/TMA kernel for every GEMM-shaped op of the model.
 * Replaces: torch.nn.Conv2d 3x3/1x1 (xunet.py:114-129,141,146,149,190,217,292-299,344,385,473),
 * FiLM.dense / Linear (xunet.py:80,84,273-277) and the bmm's inside nn.MultiheadAttention
 * (xunet.py:161,175), forward and backward (autograd of the same lines).
 * ------------------------------------------------------------------------------------------ */
typedef struct b3d_view4 {
  const void* ptr;    /* bf16 */
  int32_t dim[4];     /* extents: c (contiguous), x, y, n */
  int64_t stride[3];  /* element strides of x, y, n; multiples of 8 */
} b3d_view4;

typedef struct b3d_gemm_desc {
  /* D[rows, cols] = alpha * sum_k A[rows,k] * B[cols,k]  (+ epilogue)
   * a_major = 0: rows = pixels (x,y,n) of the OUTPUT grid (W,H,N); K = channels of up to three
   *              sources concatenated, times taps (3x3 taps shift the pixel by tap-pad, times
   *              conv_stride).  a_major = 1: rows = channels of a[0]; K = its pixels.
   * b_major = 0: cols = dim x of b, K = its channels.  b_major = 1: cols = channels of b,
   *              K = its pixels (shifted by the tap selected by grid z when b_taps = 9). */
  b3d_view4 a[3];
  int32_t a_taps[3];
  int32_t n_src;
  b3d_view4 b;
  int32_t a_major, b_major;
  int32_t conv_stride, pad;
  int32_t b_taps;
  int32_t W, H, N;
  int32_t m_rows, n_cols;
  int32_t batched;       /* (y,n) are batch coordinates shared by A, B and the output */
  int32_t b_batch_xor;   /* B is read from image n ^ b_batch_xor (cross-frame attention) */
  int32_t out_batch_xor; /* output is written to image n ^ out_batch_xor */
  int32_t split_k;       /* a_major = 1 only; > 1 needs out_atomic */
  void* out;
  int32_t out_fp32;
  int64_t out_stride[3]; /* element strides of (row or x), y, n */
  int64_t out_tap_stride;
  void* out2;            /* optional bf16 copy of silu(value) */
  const float* bias;     /* [n_cols] */
  const float* rowvec;   /* [N images, rowvec_ld]: added per image (logsnr embedding, xunet.py:502) */
  int64_t rowvec_ld;
  const void* resid;     /* bf16, addressed like out: value = (acc + bias + resid) * out_scale */
  float alpha, out_scale;
  int32_t act;           /* 0 none, 1 silu, 2: value = (acc + bias) * silu'(resid) * out_scale -- resid is then the pre-activation
                          * whose SiLU the forward took (autograd of F.silu fused into the data-gradient GEMM, xunet.py:84) */
  int32_t out_atomic;    /* fp32 atomicAdd into out */
  int32_t bn_tile;       /* 0 = auto */
  /* Optional GroupNorm statistics of the bf16 output, taken in the epilogue (saves the consumer's b3d_gn_stats pass):
   * gn_sums fp32 [images, 32, 2] (sum, sumsq) is ACCUMULATED (caller zeroes it); n_cols must be 32 * gn_cpg with
   * gn_cpg in {4, 8, 16, 32}.  Convolution / dense form with bf16 TMA-store output only; every tile must be full
   * (see b3d_gemm_stats_supported).  gn_rows_per_image: rows of one image for the dense form (H = N = 1), else 0. */
  float* gn_sums;
  int32_t gn_cpg;
  int32_t gn_rows_per_image;
} b3d_gemm_desc;

int b3d_gemm(const b3d_gemm_desc* d, void* stream);
/* Stride-1 3x3 convolutions in the halo-patch form (one (8+2) x (rows+2)-pixel input patch per 64-channel block in shared
 * memory, the nine taps read it through shifted UMMA descriptors) instead of nine shifted tile loads.  Same results.
 * 0 = off, 1 = patch rows at a 2048-byte pitch (one TMA load per row), 2 = packed rows (one load per patch; the default).
 * The environment's B3D_HALO sets the initial value.  A tuning switch of this library, no reference counterpart. */
void b3d_set_conv_halo(int on);
/* 1 when b3d_gemm can honour d->gn_sums for this descriptor (pure host code) */
int b3d_gemm_stats_supported(const b3d_gemm_desc* d);

/* Fused attention forward for L <= 256 tokens, head_dim <= 256 (all attention levels of CFG-64; the deepest of CFG-128):
 * softmax(Q K^T * scale) V per (image, head) with S and O in TMEM, P through shared memory, operands by TMA straight
 * from the packed in_proj output qkv [N, L, 3*heads*d] (bf16).  cross = 1 reads K, V of image n ^ 1 (xunet.py:206-211).
 * out: bf16 [N, L, heads*d].  probs: optional bf16 [N, heads, L, ld_probs] (normalised P, for a backward pass that reads P).
 * lse: optional fp32 [N, heads, L], per query row log2(sum_j 2^(s_ij*scale*log2e)): all a flash-style backward needs, it
 * recomputes P_ij = 2^(s_ij*scale*log2e - lse_i) from Q and K (8 B per row instead of 2*L).
 * Replaces the bmm / softmax / bmm inside nn.MultiheadAttention (xunet.py:161,175). */
int b3d_attn_fwd(const void* qkv, int N, int L, int heads, int d, int cross, float scale, void* out, void* probs, int64_t ld_probs,
                 float* lse, void* stream);
/* Fused backward of the same op (autograd of the MultiheadAttention core, xunet.py:161,175,202-211): from the packed qkv,
 * either the probabilities the forward saved (probs) or -- probs = NULL -- the row statistics lse it saved (P is then recomputed
 * tile by tile: one more QK^T on the tensor cores and one exponential per score, no L x L tensor in HBM at all),
 * the attention output o [N,L,C] and its gradient dout [N,L,C] it writes
 * dqkv [N,L,3C] (dQ | dK | dV; cross: dK, dV go to image n ^ 1).  dP, dS never leave the SM.  Covers
 * b3d_attn_bwd_supported(L, d) == 1 (L <= 256, head_dim <= 128 with the accumulators fitting 512 TMEM columns);
 * other shapes use the b3d_gemm + b3d_softmax_bwd path. */
int b3d_attn_bwd_supported(int L, int d);
int b3d_attn_bwd(const void* qkv, const void* probs, int64_t ld_probs, const float* lse, const void* o, const void* dout, void* dqkv, int N, int L,
                 int heads, int d, int cross, float scale, void* stream);

/* ------------------------------------------------------------------------------------------
 * HBM-bound kernels (elementwise / normalisation / reductions).  bf16 NHWC unless noted.
 * ------------------------------------------------------------------------------------------ */

/* K1: camera rays + NeRF positional encoding + cond mask + learned embeddings -> [2B,H,W,ld] (ch 0..143).
 * Replaces the host round trip through visu3d and the ATen chain at xunet.py:311-336 (posenc_nerf :49-59).
 * K,R,t: fp64 [B,3,3],[B,2,3,3],[B,2,3]; cond_mask: uint8 [B]; pos_emb fp32 [144,H,W]; first/other fp32 [144]. */
int b3d_pose_embedding(const double* K, const double* R, const double* t, const uint8_t* cond_mask, const float* pos_emb,
                       const float* first_emb, const float* other_emb, void* out, int out_fp32, int64_t ld, int B, int H, int W,
                       void* stream);
/* autograd of xunet.py:334-336 w.r.t. pos_emb / first_emb / other_emb; dP fp32 [2,H,W,ld] (already summed over b) */
int b3d_pose_embedding_bwd(const float* dP, int64_t ld, int H, int W, float* dpos_emb, float* dfirst, float* dother, void* stream);
/* posenc_ddpm(clip(logsnr)), xunet.py:32-46,305-307: fp32 [n] -> bf16 [n, emb_ch] */
int b3d_ddpm_embedding(const float* logsnr, void* out, int n, int emb_ch, void* stream);
/* torch.stack([x,z],1) + rearrange, xunet.py:493-494: fp32 NCHW x2 -> bf16 [2B,H,W,8]; and its adjoint */
int b3d_pack_input(const float* x, const float* z, void* out, int B, int HW, void* stream);
int b3d_unpack_input_grad(const void* g, float* dx, float* dz, int B, int HW, void* stream);
/* `[:, 1]` of the last conv, xunet.py:536: bf16 [2B,H,W,ld] -> fp32 [B,3,H,W]; and its adjoint */
int b3d_unpack_output(const void* y, int64_t ld, float* eps, int B, int HW, void* stream);
int b3d_pack_output_grad(const float* deps, void* dy, int64_t ld, int B, int HW, void* stream);
/* parameter layout: fp32 PyTorch [Co,Ci,kh,kw] (state_dict schema) <-> bf16 K-major GEMM operands.
 * mode 0 forward [Co, k_off + tap*Cpad + ci], mode 1 dgrad [Ci, tap'*Cpad + co] (flipped taps), mode 2 [(tap,ci), Co] */
int b3d_pack_weight(const float* w, int Co, int Ci, int taps, int mode, int ci0, int cn, int Cpad, int64_t k_off, void* out, int64_t out_ld,
                    void* stream);
int b3d_unpack_wgrad(const float* g, int64_t g_ld, int64_t k_off, int Cpad, int Co, int Ci, int taps, int ci0, int cn, float* dw,
                     int accumulate, void* stream);
int b3d_cast_bf16(const float* in, void* out, int64_t n, void* stream);
/* fp32 matrix-vector product on W [R, C]: y[R] = W v (+ add) (transpose 0) or y[C] = W^T v (+ add) (transpose 1); add may be
 * NULL.  Weight-space arithmetic of AttnBlock's two back-to-back Linear layers (xunet.py:211-219: out_proj inside
 * nn.MultiheadAttention, then `linear`), which the engine applies as ONE composed layer: bias W_lin b_out + b_lin in the
 * forward, d(b_out) = W_lin^T d(b_lin) in the backward. */
int b3d_matvec(const float* W, int R, int C, const float* v, const float* add, float* y, int transpose, void* stream);
/* A[r, c] += u[r] * v[c] on fp32 A [R, C]: the bias term of d W_lin = (dy^T o) W_out^T + d(b_lin) b_out^T of the same composed layer */
int b3d_rank1_add(float* A, int R, int C, const float* u, const float* v, void* stream);
/* The same conversions for MANY tensors in one launch (every parameter is repacked each step; one launch instead of
 * ~450).
 * pack: `groups` is a DEVICE array of b3d_pack_group -- one fp32 source [Co, Ci, taps] (taps <= 9) and the n_dst <= 8
 * destinations built from it, so that the source is read once: mode 0 = out[co, k_off + tap*Cpad + (ci - ci0)] for ci in
 * [ci0, ci0 + cn), mode 1 = out[ci, k_off + (taps-1-tap)*Cpad + co], mode 2 = out[tap*Ci + ci, co] (as b3d_pack_weight);
 * pad columns / rows are NOT written (zero the operands once).  A group whose d[0].mode is 3 is an fp32 copy of `total`
 * values.  Blocks: ceil(Co/32)*ceil(Ci/32) per group (copies: ceil(total/1024)); block0 = running sum, and `block_group`
 * (DEVICE array, total_blocks entries) names the group of every block.
 * unpack: `jobs` is a DEVICE array of b3d_pack_job, semantics of b3d_unpack_wgrad, ceil(Co/32)*ceil(cn/32) blocks per job,
 * block0 = running sum; `block_job` as above, or NULL (binary search over block0). */
typedef struct b3d_pack_job {
  const float* src; void* dst;
  int32_t Co, Ci, taps, mode, ci0, cn, Cpad, accumulate;
  int64_t k_off, ld, total, block0;
} b3d_pack_job;
typedef struct b3d_pack_dst {
  void* dst;
  int32_t mode, ci0, cn, Cpad;
  int64_t k_off, ld;
} b3d_pack_dst;
typedef struct b3d_pack_group {
  const float* src;
  int32_t Co, Ci, taps, n_dst;
  int64_t total, block0;
  b3d_pack_dst d[8];
} b3d_pack_group;
int b3d_pack_weights_multi(const void* groups, int n_groups, const int32_t* block_group, int64_t total_blocks, void* stream);
int b3d_unpack_wgrads_multi(const void* jobs, int n_jobs, const int32_t* block_job, int64_t total_blocks, void* stream);
/* out = scale*a (bf16 [rows,C]) and colsum[c] += column sums of out (bias gradients fused with the 1/sqrt2 of xunet.py:152,220) */
int b3d_scale_colsum(const void* a, void* out, float scale, int64_t rows, int C, float* colsum, void* stream);
/* avgpool_downsample / nearest_neighbor_upsample, xunet.py:17-28 (scale folds the adjoint's 1/4) */
int b3d_avgpool2(const void* in, void* out, int N, int H, int W, int C, float scale, void* stream);
int b3d_upsample2(const void* in, void* out, int N, int Ho, int Wo, int C, float scale, void* stream);
/* dx = dy * silu'(x)  (autograd of F.silu, xunet.py:84) ; out = scale*(a+b) */
int b3d_silu_bwd(const void* dy, const void* x, void* dx, int64_t n, void* stream);
/* (x = NULL and dx = NULL: the two column sums of dy alone.)
 * dx = dy * silu'(x) over contiguous [rows, C] plus, from the same pass, the column sums of dx per group of rows
 * (group_out[g*group_ld + c] +=, may be null) and over all rows (total_out[c] +=, may be null); fp32, pre-zeroed by the caller.
 * Autograd of F.silu(emb) (xunet.py:84) fused with d(logsnr_emb) (xunet.py:502) and the conditioning conv's bias gradient. */
int b3d_silu_bwd_colsum(const void* dy, const void* x, void* dx, int C, int64_t rows_per_group, int groups, float* group_out,
                        int64_t group_ld, float* total_out, void* stream);
int b3d_add_scale(const void* a, const void* b, void* out, float scale, int64_t n, void* stream);
/* out[g,c] += scale * sum of rows of group g (bias gradients -- scale folds the 1/sqrt2 of xunet.py:152,220; d logsnr_emb
 * per image, xunet.py:502); out fp32, pre-zeroed */
int b3d_colsum(const void* in, int64_t ld, int C, int64_t rows_per_group, int groups, float* out, int64_t out_ld, float scale, void* stream);
/* out[f,...] = sum_b in[b,f,...]  (gradient of the batch-shared learned embeddings, xunet.py:334-336) */
int b3d_batch_frame_sum(const void* in, void* out, int B, int64_t per_frame_elems, void* stream);
/* adjoint of the strided conditioning convs w.r.t. their input from per-tap GEMM products, xunet.py:340-349 */
int b3d_conv_tap_scatter(const float* G, int64_t ldg, int C, int h, int w, int s, int H, int W, float* dP, int64_t ldp, int accumulate,
                         void* stream);
/* softmax inside nn.MultiheadAttention (xunet.py:161,175) over rows of length L, and its adjoint.
 * Logits S / dP may be fp32 (the GEMM writes them straight from its fp32 accumulators) or bf16; P, dS are bf16. */
int b3d_softmax(const void* S, int s_fp32, int64_t ld_s, void* P, int64_t ld_p, int64_t rows, int L, void* stream);
int b3d_softmax_bwd(const void* P, int64_t ld_p, const void* dP, int dp_fp32, int64_t ld_dp, void* dS, int64_t ld_ds, int64_t rows, int L,
                    void* stream);

/* GroupNorm(32, C, eps 1e-5) per image (xunet.py:61-71) on the channel-concat of up to two tensors.
 * b3d_gn_stats accumulates raw (sum, sumsq) into ws_sums fp32 [N,32,2] (caller zeroes it).
 * b3d_gn_apply: mode 0 plain (AttnBlock :198), 1 +SiLU (:140,:535), 2 +FiLM(+dropout) (:87,:143-145) with
 *   scale = ss[row, c], shift = ss[row, F + c]; ss_image (optional int32 [N], device): image n takes the ss rows of image
 *   ss_image[n] -- the sampler's batch holds only four distinct conditionings, so FiLM is evaluated for four images.
 * b3d_gn_bwd: autograd of the same lines; writes d(scale,shift) to dss, accumulates dgamma/dbeta (fp32),
 *   uses ws_s12 fp32 [N,32,2] (caller zeroes it), writes dx0/dx1 (+ add0_scale * add0 + add1a/add1b when given: the
 *   residual branch's gradient joins here with its 1/sqrt2, xunet.py:152,220).  Optional fused
 *   reductions (fp32, accumulated): dss_sum [2F] = column sums of dss (FiLM dense bias gradient, xunet.py:80),
 *   dx_sum [C] = column sums of dx before the addends (bias gradient of the layer that produced x). */
int b3d_gn_stats(const void* x0, int C0, int64_t ld0, const void* x1, int C1, int64_t ld1, int N, int HW, float* ws_sums, void* stream);
int b3d_gn_apply(const void* x0, int C0, int64_t ld0, const void* x1, int C1, int64_t ld1, int N, int HW, const float* sums,
                 const float* gamma, const float* beta, int mode, const void* ss, int64_t ld_ss, int F, float drop_p, uint64_t seed, const uint64_t* seed_ptr,
                 void* out, int64_t ld_out, const int32_t* ss_image, void* stream);
/* b3d_gn_stats followed by b3d_gn_apply on the same input (ws_sums pre-zeroed by the caller, left holding the sums for the
 * backward pass).  Both passes run over a chunk of images at a time when a chunk budget is set, so the apply pass reads
 * the chunk from L2 rather than from HBM; b3d_gn_bwd chunks its reduce / apply passes the same way. */
int b3d_gn_stats_apply(const void* x0, int C0, int64_t ld0, const void* x1, int C1, int64_t ld1, int N, int HW, float* ws_sums,
                       const float* gamma, const float* beta, int mode, const void* ss, int64_t ld_ss, int F, float drop_p, uint64_t seed,
                       const uint64_t* seed_ptr, void* out, int64_t ld_out, const int32_t* ss_image, void* stream);
/* budget (MB) for the tensors both passes of a GroupNorm read per image chunk; 0 = no chunking.  Default: the
 * environment's B3D_GN_CHUNK_MB, else 0.  A tuning knob of this library, no reference counterpart. */
void b3d_set_gn_chunk_mb(int mb);
/* b3d_gn_stats_apply (and, with b3d_set_gn_fuse_bwd_mb, b3d_gn_bwd) run as ONE cluster-fused launch (the blocks of an image exchange their partial sums through
 * distributed shared memory; the second pass re-reads the image from L2) when the input tensor is at most this many MB --
 * small per-GPU batches and the sampler, where the launches, not the bytes, are what costs.  Default 40 (B3D_GN_FUSE_MB in the
 * environment overrides); 0 = always two launches.  The backward's threshold is separate and 0 by default
 * (B3D_GN_FUSE_BWD_MB): the fused backward measured slower than its two launches.  No reference counterpart. */
void b3d_set_gn_fuse_mb(int mb);
void b3d_set_gn_fuse_bwd_mb(int mb);
int b3d_gn_bwd(const void* x0, int C0, int64_t ld0, const void* x1, int C1, int64_t ld1, int N, int HW, const float* sums,
               const float* gamma, const float* beta, int mode, const void* dy, int64_t ld_dy, const void* ss, int64_t ld_ss, int F,
               void* dss, int64_t ld_dss, float drop_p, uint64_t seed, const uint64_t* seed_ptr, float* dgamma, float* dbeta, float* dss_sum,
               float* dx_sum, float* ws_s12, void* dx0,
               int64_t ld_dx0, void* dx1, int64_t ld_dx1, const void* add0, int64_t ld_add0, const void* add1a, int64_t ld_add1a,
               const void* add1b, int64_t ld_add1b, float add0_scale, void* stream);

/* ------------------------------------------------------------------------------------------
 * Diffusion glue (fp32 NCHW, as the reference holds these tensors)
 * ------------------------------------------------------------------------------------------ */
/* q_sample + conditioning-frame replacement: train.py:50-60,95-96 / lightning/diff3d.py:66-69,168-180 */
int b3d_q_sample(const float* z, const float* noise, const float* logsnr, const float* x, const float* xrand, const uint8_t* mask,
                 float* z_out, float* x_out, int B, int64_t per_sample, void* stream);
/* F.mse_loss / l1_loss / smooth_l1_loss (mean) + d loss/d pred: train.py:102-112, lightning/diff3d.py:87-94.
 * loss_type 0 l2, 1 l1, 2 huber.  `scratch`: b3d_loss_scratch_bytes() bytes of device memory whose last word
 * (the block ticket) is zero on entry; the kernel leaves it zero.  Deterministic two-stage reduction. */
int b3d_loss_scratch_bytes(void);
int b3d_loss(const float* noise, const float* pred, float* dpred, int64_t n, int loss_type, float grad_scale, void* scratch,
             float* loss_out, void* stream);
/* Same loss and gradient as b3d_loss, for callers that cannot hand in scratch memory (no counterpart in the
 * reference; it exists for hosts that run inside a fixed device-memory budget -- all HBM pre-carved into their
 * own arenas, or driver allocations forbidden after start-up because cudaMalloc/cudaFree stall the device).
 * The block partials and ticket live in a 4100-byte array inside the library image (held by every GPU that
 * loads the library until exit).  n must be <= b3d_loss_static_max_elems() (4 194 304); larger n returns -1
 * with a message.  The array is shared by all calls on a device: calls that overlap on the GPU race, so keep
 * them ordered.  The first call on each device resolves the array's address; make it before any stream
 * capture.  Allocates nothing. */
int64_t b3d_loss_static_max_elems(void);
int b3d_loss_static(const float* noise, const float* pred, float* dpred, int64_t n, int loss_type, float grad_scale, float* loss_out,
                    void* stream);
/* p_mean_variance + p_sample arithmetic given both UNet outputs: sampling.py:79-127 (noise skipped iff logsnr_next == 0) */
int b3d_sampler_update(const float* z, const float* eps_c, const float* eps_u, const float* w, const float* noise, float* out,
                       float logsnr, float logsnr_next, int B, int64_t per_sample, void* stream);
/* the same update with (logsnr, logsnr_next) read from two device floats: one captured graph serves all 256 steps */
int b3d_sampler_update_dev(const float* z, const float* eps_c, const float* eps_u, const float* w, const float* noise, float* out,
                           const float* sched_dev, int B, int64_t per_sample, void* stream);
/* Sampler-side caching of the conditioning path: conv_l(pose_emb) + bias of xunet.py:340-349 depends only on
 * K/R/t/cond_mask, not on the denoising step.  c: that cached part, fp32 [D, rows_per_image, E] for the D distinct
 * conditionings of a batch; lemb fp32 [*, E] the per-step logsnr embedding, of which conditioning j takes row lemb_row[j]
 * (device int32 [D]).  out bf16 [D, rows_per_image, E] = silu(c + lemb), the FiLM GEMM's operand (xunet.py:84,502). */
int b3d_cond_silu(const float* c, const float* lemb, const int32_t* lemb_row, void* out, int D, int64_t rows_per_image, int E, void* stream);
/* torch.optim.Adam(lr, betas) step over a flat fp32 arena: train.py:235,278 / lightning/diff3d.py:106 */
int b3d_adam(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2, float eps, int step,
             float grad_scale, void* stream);
/* the same step plus the exponential moving average of the parameters that lightning/diff3d.py:11-19 describes (and never
 * implements): ema = ema_decay * ema + (1 - ema_decay) * p_new, in the same pass (36 instead of 28 bytes per parameter) */
int b3d_adam_ema(float* p, const float* g, float* m, float* v, float* ema, int64_t n, float lr, float beta1, float beta2, float eps,
                 int step, float grad_scale, float ema_decay, void* stream);
/* the same step (ema may be NULL) reading a bf16 gradient: a data-parallel trainer all-reduces a bf16 copy of the gradient
 * range and steps straight from that buffer instead of casting it back into the fp32 arena */
int b3d_adam_bf16grad(float* p, const void* g_bf16, float* m, float* v, float* ema, int64_t n, float lr, float beta1, float beta2,
                      float eps, int step, float grad_scale, float ema_decay, void* stream);

#ifdef __cplusplus
}
#endif
#endif
