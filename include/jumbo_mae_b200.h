/*
 * jumbo_mae_b200.h -- C ABI of the B200 (sm_100a) kernel library for the Jumbo-CLS
 * MAE pre-training step of antofuller/jumbo_mae_tpu.
 *
 * The reference has no native code: every op below replaces what XLA executes for
 * the cited Python lines (paths relative to the reference checkout).  A JAX host
 * binds these through thin XLA-FFI handlers; the in-tree host
 * (jumbo_mae_tpu_b200/) binds them with ctypes.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless it says "host";
 *   - `stream` is a cudaStream_t passed as void*; calls only enqueue work: no
 *     allocation, no synchronisation, no host read-back;
 *   - workspaces are caller-owned; jm_*_workspace() returns the bytes needed;
 *   - return value 0 = success, non-zero = failure with a message available from
 *     jm_last_error() (thread-local).  Nothing throws across this boundary;
 *   - activations are row-major [rows, features]; "ld*" arguments are row strides
 *     in ELEMENTS; bf16 is the raw 16-bit pattern (uint16_t).
 */
#ifndef JUMBO_MAE_B200_H_
#define JUMBO_MAE_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef uint16_t jm_bf16;

const char* jm_last_error(void);
int jm_version(void);
/* 1 if the current device is sm_100 (B200); the library refuses to run elsewhere. */
int jm_device_ok(void);
/* Number of kernels this library has launched in this process (bench.py's gpu_launches). */
long long jm_launch_count(void);

/* ---------------------------------------------------------------- masking
 * random_masking, src/utils_mae.py:88-96, with the uniform noise as an input:
 * ids_shuffle = stable ascending argsort(noise); ids_restore = argsort(ids_shuffle);
 * mask_row[i] = 1 if ids_restore[i] >= keep else 0 (the reference's mask is this row
 * broadcast over the batch: one permutation per call).  length <= 1024. */
int jm_mask_sort(const float* noise, int length, int keep, int32_t* ids_shuffle, int32_t* ids_restore,
                 float* mask_row, void* stream);

/* Per-image masking, an opt-in extension (the reference draws ONE permutation per batch, src/utils_mae.py:88-96):
 * jm_mask_sort over a [batch, length] noise matrix, one row per image; and the index consumers with `ids_stride` =
 * elements between two images' index rows (jm_patchify_pi, jm_masked_targets_pi, jm_unshuffle_fwd_pi / _bwd_pi,
 * jm_posemb_grad_pi; jm_gemm takes pe_idx_stride).  The backward variants add with vector atomics. */
int jm_mask_sort_batched(const float* noise, int batch, int length, int keep, int32_t* ids_shuffle, int32_t* ids_restore,
                         float* mask, void* stream);
int jm_patchify_pi(const uint8_t* images, int batch, int image_size, int patch_size, const int32_t* ids, long ids_stride,
                   int n_sel, int rows_per_image, int row_off, jm_bf16* out, long ldo, void* stream);
int jm_masked_targets_pi(const uint8_t* images, int batch, int image_size, int patch_size, const int32_t* ids,
                         long ids_stride, int n_sel, int num_cls, int norm_pix, uint8_t* targets, long ldt, float* stats,
                         int32_t* row_map, void* stream);
int jm_unshuffle_fwd_pi(const float* y, const int32_t* ids_restore, long ids_stride, const float* mask_emb,
                        const float* dec_wpe, int batch, int N, int n, int J, int dim, float* out, void* stream);
int jm_unshuffle_bwd_pi(const float* dout, const int32_t* ids_restore, long ids_stride, int batch, int N, int n, int J,
                        int dim, jm_bf16* dy, float* dmask_emb, void* stream);
int jm_posemb_grad_pi(const float* dx, long ldx, int batch, int T, int J, int n, const int32_t* ids, long ids_stride,
                      int dim, float* dwpe, void* stream);

/* index_sequence, src/utils_mae.py:84-85: out[b, i, :] = x[b, ids[i], :] (f32). */
int jm_index_sequence(const float* x, const int32_t* ids, int batch, int length_in, int length_out, int dim,
                      float* out, void* stream);

/* ---------------------------------------------------------------- pixels
 * PretrainModule pre-processing (src/pretraining.py:90-91) fused with
 * extract_patches (src/utils_mae.py:67-73) and the token gather of random_masking:
 * out[b*rows_per_image + row_off + i, e] = normalised pixel e of patch ids[i] of image b
 * ((ph, pw, c) element order), bf16, zero in columns [P, ldo).  ids == NULL selects
 * patches 0..n_sel-1.  Rows [0,row_off) of each image block are zero-filled. */
int jm_patchify(const uint8_t* images, int batch, int image_size, int patch_size, const int32_t* ids, int n_sel,
                int rows_per_image, int row_off, jm_bf16* out, long ldo, void* stream);

/* jm_patchify with Mixup / CutMix on the normalised pixels (fine-tune input path, SURVEY.md 8(f) rank 1;
 * src/finetuning.py:87-88,95 and Mixup in src/utils.py:67-90).  perm [batch] int32 = the batch permutation.
 *   mode 1 (apply_mixup):  pixel = ratio * a + (1 - ratio) * b,  b = images[perm[image]]
 *   mode 2 (apply_cutmix): pixel = b inside the box, a outside; xmask [image_size], ymask [image_size] uint8 are the
 *                          box's column / row membership (random_bounding_box, src/utils.py:85-92, evaluated by the
 *                          caller: 2 x image_size bytes), ratio unused.
 * perm[i] must be in [0, batch).  Everything else as jm_patchify. */
int jm_patchify_mix(const uint8_t* images, int batch, int image_size, int patch_size, const int32_t* ids, int n_sel,
                    int rows_per_image, int row_off, jm_bf16* out, long ldo, const int32_t* perm, int mode,
                    float ratio, const uint8_t* xmask, const uint8_t* ymask, void* stream);
/* The same with the Mixup ratio read from device memory (one f32), so that a captured CUDA graph can be replayed with
 * a new draw every step (the reference draws a ratio per step, src/utils.py:92-111). */
int jm_patchify_mix_dev(const uint8_t* images, int batch, int image_size, int patch_size, const int32_t* ids, int n_sel,
                        int rows_per_image, int row_off, jm_bf16* out, long ldo, const int32_t* perm, int mode,
                        const float* ratio_dev, const uint8_t* xmask, const uint8_t* ymask, void* stream);

/* extract_patches on the normalised image as f32 (reference signature; tests only). */
int jm_extract_patches(const uint8_t* images, int batch, int image_size, int patch_size, float* out, void* stream);

/* patch_mse_loss (src/utils_mae.py:51-64) fused with the target computation
 * (src/pretraining.py:113-118).  pred is [batch*(J+N), P] f32 with leading J non-patch
 * rows per image.  Writes loss[0] (mean over images of mean over masked patches of the
 * per-patch MSE) and, if dpred != NULL, dpred (bf16, same shape, ld = lddp) = d loss /
 * d pred * grad_scale (zero on unmasked and CLS rows).
 * A mask_row without any masked patch gives 0 / 0 = NaN, as the reference's division by sum(valid) does.
 * workspace: jm_patch_loss_workspace(batch, N) bytes. */
size_t jm_patch_loss_workspace(int batch, int num_patches);
int jm_patch_loss(const float* pred, long ldp, const uint8_t* images, const float* mask_row, int batch,
                  int image_size, int patch_size, int num_cls, int norm_pix, float grad_scale, float* loss,
                  jm_bf16* dpred, long lddp, void* workspace, size_t workspace_bytes, void* stream);

/* RandomResizedCrop (bicubic) + horizontal flip of the pre-training recipe on the device (src/dataset.py:56-58,70-71:
 * T.RandomResizedCrop(S, scale=(0.2, 1.0), interpolation=3), T.RandomHorizontalFlip, T.PILToTensor).  Bit-exact
 * with PIL's crop(box).resize((S, S), BICUBIC) [+ FLIP_LEFT_RIGHT]: Pillow's two-pass fixed-point resampling.
 *   src      decoded images, uint8 HWC (RGB), packed in one buffer; image b starts at byte src_off[b]
 *   src_hw   [batch, 2] (height, width) of each source image
 *   boxes    [batch, 4] crop (top, left, height, width), drawn by the host (RandomResizedCrop.get_params);
 *            every height and width must be <= max_crop
 *   flips    [batch] 0 / 1 or NULL
 *   out      [batch, 3, S, S] uint8 (the batch the training step takes)
 * workspace: jm_rrc_flip_workspace(batch, max_crop, S) bytes (coefficient tables + the intermediate image). */
size_t jm_rrc_flip_workspace(int batch, int max_crop, int out_size);
int jm_rrc_flip(const uint8_t* src, const long* src_off, const int* src_hw, const int* boxes, const uint8_t* flips,
                int batch, int max_crop, int out_size, uint8_t* out, void* workspace, size_t workspace_bytes,
                void* stream);

/* Gradient of a learnable position embedding (--posemb learnable, src/modeling.py:115-118,123):
 * dwpe[ids[i], :] += sum over the batch of dx[b*T + J + i, :] for the n kept patches i (ids = ids_shuffle, or NULL for
 * the identity when every patch is kept).  dx is the f32 gradient of the encoder input, [batch*T, dim], stride ldx. */
int jm_posemb_grad(const float* dx, long ldx, int batch, int T, int J, int n, const int32_t* ids, int dim, float* dwpe,
                   void* stream);

/* In place x[r, :] *= (r % period < skip ? scale_cls : scale)[r / period] on a bf16 [rows, cols] matrix: the backward
 * of a DropPath-scaled residual branch (src/modeling.py:157,181-183) -- the gradient entering the branch is the block
 * output gradient times the image's keep-mask / (1 - p); drop_path3 of the Jumbo CLS rows has its own mask
 * (scale_cls, may be NULL). */
int jm_scale_rows_bf16(jm_bf16* x, long ld, long rows, int cols, const float* scale, const float* scale_cls, int period,
                       int skip, void* stream);

/* jm_scale_rows_bf16 with per-column factors as well (any of the four may be NULL = 1): col_scale [cols] on the rows
 * r % period >= skip, col_scale_cls [skip * cols] on the others (row t of a period uses entries t * cols ...).  The
 * gradient that enters a LayerScale-d branch is dy * scale (src/modeling.py:165-166,192-201). */
int jm_scale_bf16(jm_bf16* x, long ld, long rows, int cols, const float* scale, const float* scale_cls, int period,
                  int skip, const float* col_scale, const float* col_scale_cls, void* stream);

/* d scale (+=) of a LayerScale vector from the ACCUMULATED gradients of the Dense it scales (call once per optimizer
 * step, after the last micro-step / the all-reduce): dscale[n] += (sum_k W[k, n] dW[k, n] + b[n] db[n]) / scale[n]
 * with W, dW [K, N] f32 of row stride ld; see csrc/elementwise.cu for the identity.  jm_layerscale_cls_grad is the
 * per-layer form for scale3, whose Dense (the Jumbo MLP) is shared by all layers: dy, c_out (the layer's output CLS
 * rows) and c (its LayerNorm-ed input) are [batch, N] f32 views, dscale[n] += sum_b dy (c_out - c) / scale[n]. */
int jm_layerscale_grad(const float* W, const float* dW, long ld, int K, int N, const float* b, const float* db,
                       const float* scale, float* dscale, void* stream);
int jm_layerscale_cls_grad(const float* dy, long lddy, const float* c_out, long ldo, const float* c, long ldc, int batch,
                           int N, const float* scale, float* dscale, void* stream);

/* patch_mse_loss(patch_output, patch_target, valid=None) with explicit targets, exactly src/utils_mae.py:51-64:
 * mean over images of [ mean_L( valid > 0 ? mean_c (target - output)^2 : 0 ) / (sum(valid) / L) ].
 * patch_output / patch_target are [batch*L, P] f32 with row strides ldo / ldt; valid is [batch, L] f32 or NULL
 * (all ones).  An image without any valid patch gives 0/0 = NaN, as in the reference.
 * workspace: jm_patch_loss_workspace(batch, L) bytes. */
int jm_patch_mse_loss(const float* patch_output, long ldo, const float* patch_target, long ldt, const float* valid,
                      int batch, int num_patches, int patch_dim, float* loss, void* workspace, size_t workspace_bytes,
                      void* stream);

/* jm_patch_loss without a workspace argument and without any allocation.  It has no counterpart
 * in the reference; it exists for hosts that are not allowed to allocate device memory around the
 * call: an XLA-FFI handler (XLA owns every device buffer and a handler must not call cudaMalloc),
 * a service that runs under a fixed memory budget carved out at start-up, or an embedding whose
 * memory pool belongs to somebody else.
 *
 * The per-patch partial sums go to a fixed array inside the library instead
 * (JM_PATCH_LOSS_STATIC_MAX floats = 4 MiB, reserved on every GPU that loads the library until
 * the process exits).  Results are identical to jm_patch_loss.
 *   - batch * num_patches must not exceed JM_PATCH_LOSS_STATIC_MAX; larger problems are refused
 *     with an error (use jm_patch_loss and your own workspace);
 *   - the first call on each device looks the array's address up (one cudaGetSymbolAddress, no
 *     allocation); make that first call outside any stream capture;
 *   - all calls on a device share the one array: two calls whose kernels can overlap on the GPU
 *     (different streams, no dependency) race on it.  Issue them on one stream or order them
 *     with events.  Nothing needs clearing between calls: each call rewrites every entry it reads. */
#define JM_PATCH_LOSS_STATIC_MAX (1 << 20)
int jm_patch_loss_static(const float* pred, long ldp, const uint8_t* images, const float* mask_row, int batch,
                         int image_size, int patch_size, int num_cls, int norm_pix, float grad_scale, float* loss,
                         jm_bf16* dpred, long lddp, void* stream);

/* ---------------------------------------------------------------- LayerNorm
 * flax.linen.LayerNorm as used at src/modeling.py:155-156,177-179,238,282:
 * y = (x - mean) * rsqrt(E[x^2]-mean^2 + eps) * gamma + beta, f32 statistics.
 * y_bf16 and/or y_f32 may be NULL.  mean/rstd [rows] are saved for the backward. */
int jm_layernorm_fwd(const float* x, long ldx, int rows, int dim, const float* gamma, const float* beta, float eps,
                     jm_bf16* y_bf16, long ldy, float* y_f32, long ldyf, float* mean, float* rstd, void* stream);

/* Backward of the above.  dx = (dres ? dres : 0) + LN'(dy); rows with
 * (row % row_period) < row_skip are left untouched (used for the CLS rows of the
 * Jumbo layer, src/modeling.py:194-204).  dgamma/dbeta are ACCUMULATED (+=).
 * dcol (optional, [dim], +=) receives the column sums of dx over the rows with
 * (row % dcol_period) >= dcol_skip: dx is the gradient flowing into the Dense that
 * produced the residual branch, so this is that Dense's bias gradient for free.
 * workspace: jm_layernorm_bwd_workspace(rows, dim) bytes. */
size_t jm_layernorm_bwd_workspace(int rows, int dim);
int jm_layernorm_bwd(const jm_bf16* dy, long lddy, const float* x, long ldx, const float* mean, const float* rstd,
                     const float* gamma, int rows, int dim, const float* dres, long lddres, float* dx, long lddx,
                     jm_bf16* dx_bf16, long lddxb, int row_period, int row_skip, float* dgamma, float* dbeta,
                     float* dcol, int dcol_period, int dcol_skip, void* workspace, size_t workspace_bytes,
                     void* stream);

/* jm_layernorm_bwd with the incoming residual gradient in bf16 (dres_bf16, may be NULL) and dx optional (NULL: only the
 * bf16 copy dx_bf16 is written): the decoder's gradient stream kept in bf16 only, 10 instead of 16 bytes per element
 * (dim <= 1280).  Same arithmetic otherwise; no row skipping. */
int jm_layernorm_bwd_lowp(const jm_bf16* dy, long lddy, const float* x, long ldx, const float* mean, const float* rstd,
                          const float* gamma, int rows, int dim, const jm_bf16* dres_bf16, long lddres, float* dx,
                          long lddx, jm_bf16* dx_bf16, long lddxb, float* dgamma, float* dbeta, float* dcol,
                          void* workspace, size_t workspace_bytes, void* stream);

/* The general form of jm_layernorm_bwd_lowp (dim <= 1280): residual gradient read from bf16 (dres_bf16, may be NULL),
 * row skipping and column-sum selection as jm_layernorm_bwd, and the f32 copy dx (may be NULL) written only for the rows
 * with row % f32_period < f32_skip (f32_period == 0: every row) -- the encoder's gradient stream in bf16, with f32 kept
 * for the CLS rows whose Jumbo-MLP gradient is accumulated in f32 (src/modeling.py:191-206). */
int jm_layernorm_bwd_stream16(const jm_bf16* dy, long lddy, const float* x, long ldx, const float* mean, const float* rstd,
                              const float* gamma, int rows, int dim, const jm_bf16* dres_bf16, long lddres, float* dx,
                              long lddx, int f32_period, int f32_skip, jm_bf16* dx_bf16, long lddxb, int row_period,
                              int row_skip, float* dgamma, float* dbeta, float* dcol, int dcol_period, int dcol_skip,
                              void* workspace, size_t workspace_bytes, void* stream);

/* LayerNorm over a subset of rows: output row r (compact) is LayerNorm(x[row_map[r], :]); mean / rstd are compact too.
 * jm_layernorm_bwd_rows is its backward: dy, mean, rstd compact [rows], x and dx / dx_bf16 scattered through row_map;
 * dx / dx_bf16 [total_rows, dim] are zeroed first (rows outside row_map get no gradient through this LayerNorm).
 * dim <= 512.  Used for decoder_model/dec_norm on the masked rows only (src/modeling.py:296, src/pretraining.py:110). */
int jm_layernorm_fwd_rows(const float* x, long ldx, const int32_t* row_map, int rows, int dim, const float* gamma,
                          const float* beta, float eps, jm_bf16* y_bf16, long ldy, float* mean, float* rstd,
                          void* stream);
int jm_layernorm_bwd_rows(const jm_bf16* dy, long lddy, const float* x, long ldx, const int32_t* row_map,
                          const float* mean, const float* rstd, const float* gamma, int rows, int dim, long total_rows,
                          float* dx, long lddx, jm_bf16* dx_bf16, long lddxb, float* dgamma, float* dbeta, float* dcol,
                          void* workspace, size_t workspace_bytes, void* stream);

/* Targets of the selected (masked) patches for the loss epilogue of jm_gemm (act = 5): targets [batch * n_sel, ldt]
 * uint8 = the patch's pixels in extract_patches order (src/utils_mae.py:67-76), stats [batch * n_sel, 2] = mean and
 * 1 / sqrt(biased var + 1e-6) of the normalised patch (written when norm_pix; src/pretraining.py:115-118), row_map
 * (optional) [batch * n_sel] = image * (num_cls + N) + num_cls + ids[k], the decoder row of each target.
 * ids [n_sel] are shared by the batch (src/utils_mae.py:88-96). */
int jm_masked_targets(const uint8_t* images, int batch, int image_size, int patch_size, const int32_t* ids, int n_sel,
                      int num_cls, int norm_pix, uint8_t* targets, long ldt, float* stats, int32_t* row_map,
                      void* stream);

/* loss[0] = scale * sum(row_loss[0 .. count)), summed in a fixed order.  workspace: jm_row_loss_workspace() bytes. */
size_t jm_row_loss_workspace(void);
int jm_row_loss_finalize(const float* row_loss, long count, float scale, float* loss, void* workspace,
                         size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------- GEMM
 * D = epilogue(A * B), bf16 operands, f32 accumulation in tensor memory
 * (tcgen05.mma fed by TMA).  Replaces every Dense / DenseGeneral / Conv contraction
 * of src/modeling.py:108-148 and src/pretraining.py:84-85, forward and backward.
 *   A: a_mn_major == 0 -> stored [M, K] row-major (ld = lda);
 *      a_mn_major == 1 -> stored [K, M] row-major (the transposed view, for dW = X^T dY)
 *   B: b_mn_major == 0 -> stored [N, K] row-major; b_mn_major == 1 -> stored [K, N] row-major
 *      (a Flax kernel [in, out] is B with b_mn_major = 1 in the forward pass and
 *       b_mn_major = 0 in the input-gradient pass)
 * Epilogue, applied in this order to v = acc:
 *   v += bias[col]; if (col < col_scale_n) v *= col_scale;
 *   act == 1 (GELU-tanh, src/modeling.py:148): out_bf16 <- v (pre-activation), out2_bf16 <- gelu(v)
 *   act == 2 (GELU backward): v *= gelu'(aux[row, col])
 *   act == 3: as 1, but out_bf16 <- gelu'(v): the derivative is saved instead of the pre-activation ...
 *   act == 4: ... and the backward epilogue is v *= aux[row, col] (no tanh in the backward pass)
 *   v += pe[pe_idx[(row % row_period) - row_skip], col]   (positional embedding, modeling.py:123)
 *   v *= row_scale[row / row_scale_period]                (DropPath, f32 output only)
 *   v += residual[row, col]
 *   rows with (row % row_period) < row_skip: row_mask_mode 1 -> zeros are stored,
 *   2 -> nothing is stored.
 *   accumulate != 0 -> out_f32 += v (f32 reduce-add of whole tiles in L2 by TMA, or red.global.add on the
 *   generic path; also used for split-K), else out_f32 = v.
 */
typedef struct jm_gemm_args {
  int M, N, K;
  const jm_bf16* A; long lda; int a_mn_major;
  const jm_bf16* B; long ldb; int b_mn_major;
  float* out_f32; long ldo_f32;
  jm_bf16* out_bf16; long ldo_bf16;
  jm_bf16* out2_bf16; long ldo2_bf16;
  const float* bias;
  const float* residual; long ldr;
  const jm_bf16* aux; long ldaux;
  const float* pe; const int32_t* pe_idx; long ldpe;
  int act;
  float col_scale; int col_scale_n;
  int row_period, row_skip, row_mask_mode;
  int accumulate;
  int splits;      /* 0 = choose automatically */
  int block_n;     /* 0 = choose automatically (64, 128 or 256) */
  float* colsum;   /* optional [N]: += column sums of the stored output (fused bias gradient) */
  int cta_pair;    /* 0 = automatic, 1 = force the 2-SM (cta_group::2) 256x256 kernel, -1 = never use it;
                    * 2 = the plain pair kernel (B streamed, not resident), 5 = as 1 with the default pair count (the
                    * resident-B kernel then reloads B when a pair's column tile changes: a test hook) */
  /* optional per-image scale of the branch: out_f32 = row_scale[row / row_scale_period] * (acc + bias) + residual
   * (DropPath keep-mask / (1 - p) of a residual branch, src/modeling.py:157,181-183,192,199,201); f32 output only */
  const float* row_scale; int row_scale_period;
  /* act = 5, the loss epilogue of the pixel head (src/pretraining.py:110-120, src/utils_mae.py:51-64): with
   * d = acc + bias - target, out_bf16 (optional) = d * loss_scale and row_loss[row, col / 16] = sum of d^2 over the 16
   * columns (row_loss is [M, ceil(N / 16)] f32, every entry written once; jm_row_loss_finalize sums it).  target =
   * normalised pixel of byte loss_target[row, col] (channel col % 3, ImageNet statistics), then (t - stats[row, 0]) *
   * stats[row, 1] when loss_stats != NULL (norm_pix_loss).  jm_masked_targets produces loss_target / loss_stats.
   * A K-major, B MN-major, no other epilogue option; ld_target a multiple of 16 >= N, ldo_bf16 >= N rounded up to 8. */
  const uint8_t* loss_target; long ld_target; const float* loss_stats; float* row_loss; float loss_scale;
  /* optional [N] (out_f32 epilogue): out_f32 = row_scale * col_scale_vec[col] * (acc + bias) + residual -- LayerScale's
   * scale1 / scale2 / scale3 on a residual branch (src/modeling.py:159-166,185-201) */
  const float* col_scale_vec;
  /* elements between two images' rows of pe_idx: 0 = one index row shared by the batch (the reference), row_period -
   * row_skip entries per image otherwise (per-image masking) */
  long pe_idx_stride;
} jm_gemm_args;
int jm_gemm(const jm_gemm_args* args, void* stream);

/* ---------------------------------------------------------------- attention
 * Attention core of src/modeling.py:136-137 for short sequences: per (image, head)
 * softmax(q k^T) v with q already scaled.  qkv is [batch*T, 3*heads*hd] bf16 with
 * columns [q | k | v], each head-major.  lse [batch, heads, T] f32 is saved. */
int jm_attention_fwd(const jm_bf16* qkv, long ldqkv, int batch, int T, int heads, int head_dim, jm_bf16* out,
                     long ldo, float* lse, void* stream);
/* dqkv (same layout as qkv) from dout; the q part is multiplied by q_scale so that it
 * is the gradient w.r.t. the un-scaled projection.  delta_ws: unused (rowsum(dout * out)
 * is computed inside the kernel since round 1; kept in the signature, may be NULL).
 * dbias (optional, [3*heads*hd], +=)
 * receives the column sums of dqkv, i.e. the gradient of the fused q/k/v bias. */
int jm_attention_bwd(const jm_bf16* qkv, long ldqkv, const jm_bf16* out, long ldo, const jm_bf16* dout, long lddo,
                     const float* lse, float* delta_ws, int batch, int T, int heads, int head_dim, float q_scale,
                     jm_bf16* dqkv, long lddqkv, float* dbias, void* stream);

/* ---------------------------------------------------------------- token assembly
 * CLS prepend (src/modeling.py:249-250,257): x[b, j, :] = cls[j, :] for j < J. */
int jm_fill_cls(float* x, int batch, int T, int J, int dim, const float* cls, void* stream);

/* Decoder input assembly (src/pretraining.py:96-106 + src/modeling.py:289-292):
 * out[b, j<J] = y[b, j]; out[b, J+i] = (ids_restore[i] < n ? y[b, J+ids_restore[i]] : mask_emb) + dec_wpe[i]. */
int jm_unshuffle_fwd(const float* y, const int32_t* ids_restore, const float* mask_emb, const float* dec_wpe,
                     int batch, int N, int n, int J, int dim, float* out, void* stream);
/* Its backward: dy (bf16, [batch*(J+n), dim]) and dmask_emb (+=, f32 [dim]). */
int jm_unshuffle_bwd(const float* dout, const int32_t* ids_restore, int batch, int N, int n, int J, int dim,
                     jm_bf16* dy, float* dmask_emb, void* stream);

/* Column sums (bias gradients): out[c] += sum over rows r with (r % row_period) >= row_skip
 * (row_invert != 0: < row_skip) of x[r, c].  x is bf16 (is_f32 == 0) or f32. */
int jm_colsum(const void* x, int is_f32, long ldx, int rows, int cols, int row_period, int row_skip,
              int row_invert, float* out, void* stream);

/* ---------------------------------------------------------------- optimizer
 * optax.adamw leaf update (src/pretraining.py:223-234) over a flat parameter buffer:
 * g' = g * grad_scale; m = b1 m + (1-b1) g'; v = b2 v + (1-b2) g'^2;
 * u = (m / bc1) / (sqrt(v / bc2) + eps) + (i < n_decay ? wd * p : 0); p -= lr * u.
 * Also writes the bf16 compute copy of p. */
int jm_adamw(float* p, const float* g, float* m, float* v, jm_bf16* p_bf16, long n, long n_decay, float lr,
             float b1, float b2, float eps, float wd, float bc1, float bc2, float grad_scale, void* stream);

int jm_cast_bf16(const float* x, jm_bf16* y, long n, void* stream);

/* Zeroes `bytes` bytes on the stream (cudaMemsetAsync): the gradient accumulator at the first micro-step of an optimizer
 * step (jax.tree_map(jnp.zeros_like, ...), src/pretraining.py:154-156). */
int jm_zero(void* ptr, size_t bytes, void* stream);

/* --optimizer lamb: the reference's modified_lamb (src/utils.py:124-139) on the `kernel` leaves of the flat buffers:
 * Adam moments, u = m_hat / (sqrt(v_hat) + eps) + wd * p, p -= lr * (||p|| / ||u||) * u with the norms over one leaf of
 * the Flax tree (ratio 1 where a norm is 0).  groups [n_groups, 7] int64 = {offset, rows, ld, col0, ncols, first chunk,
 * chunk count}: a leaf is a column range of a stored [rows, ld] matrix (wq / wk / wv share one).  chunks [n_chunks, 3]
 * int64 = {group, row0, nrows}, the chunks of a group consecutive; one thread block per chunk, norms summed in a fixed
 * order.  Every other leaf takes jm_adamw with wd = 0.  workspace: jm_lamb_workspace(n_chunks) bytes. */
size_t jm_lamb_workspace(int n_chunks);
int jm_lamb(float* p, const float* g, float* m, float* v, jm_bf16* p_bf16, const int64_t* groups, int n_groups,
            const int64_t* chunks, int n_chunks, float lr, float b1, float b2, float eps, float wd, float bc1, float bc2,
            float grad_scale, void* workspace, size_t workspace_bytes, void* stream);

/* optax.clip_by_global_norm in front of the optimizer (src/pretraining.py:243-244, --clip-grad > 0): with
 * norm = pre_scale * ||g||_2 over the whole flat gradient buffer (pre_scale = 1/(accum*world), the factor the
 * caller folds into jm_adamw), g is left unchanged if norm < max_norm and multiplied by max_norm / norm
 * otherwise.  sumsq: one device float, receives sum(g^2) of the unscaled buffer. */
int jm_clip_by_global_norm(float* g, long n, float pre_scale, float max_norm, float* sumsq, void* stream);

/* dst = src + bias (bias optional) over a strided [rows, cols] f32 view (cols, lds, ldd multiples of 4), written
 * as f32 and/or bf16.  Seeds the output of a split-K GEMM with residual + bias (Jumbo MLP output,
 * src/modeling.py:199) and converts an f32 gradient for a bf16 consumer.  No reference counterpart: XLA fuses
 * these adds/converts into their neighbours. */
int jm_add_bias_cast(const float* src, long lds, int rows, int cols, const float* bias, float* dst_f32, long ldd,
                     jm_bf16* dst_bf16, long ldd16, void* stream);

/* Classification loss of the fine-tune step (src/finetuning.py:90-106), one launch:
 *   labels  = one_hot(labels_idx, C) or labels_soft                         (finetuning.py:90-91)
 *   labels  = (1 - label_smoothing) * labels + label_smoothing / C          (optax.smooth_labels, :94)
 *   labels  = ratio * labels + (1 - ratio) * labels[perm]   if perm != NULL (Mixup / CutMix label side,
 *                                                                            src/utils.py:70-71,82-83)
 *   loss[b] = -sum_c labels * log_softmax(logits)                           (optax.softmax_cross_entropy, :98)
 *   acc1[b] = labels[top1] == max(labels);  acc5[b] = any over top_k(logits, 5), ties to the lower index (:99-106)
 *   dlogits = grad_scale * (softmax(logits) * sum_c labels - labels)        (d sum_b loss[b] / d logits, scaled)
 * Exactly one of labels_soft [B, C] f32 and labels_idx [B] int32 is given.  loss_rows / acc1 / acc5 [B] f32,
 * metrics [3] f32 (batch means of the three, accumulated with atomics after being zeroed here) and
 * dlogits [B, C] f32 may each be NULL.  Evaluation (det=True) passes label_smoothing = 0 and perm = NULL.
 * 1 <= C <= 8192. */
int jm_softmax_ce(const float* logits, const float* labels_soft, const int* labels_idx, const int* perm,
                  float ratio, double label_smoothing, int B, int C, float grad_scale, float* loss_rows,
                  float* acc1, float* acc5, float* metrics, float* dlogits, void* stream);
/* jm_softmax_ce with the label ratio of Mixup / CutMix read from device memory (see jm_patchify_mix_dev). */
int jm_softmax_ce_dev(const float* logits, const float* labels_soft, const int* labels_idx, const int* perm,
                      const float* ratio_dev, double label_smoothing, int B, int C, float grad_scale, float* loss_rows,
                      float* acc1, float* acc5, float* metrics, float* dlogits, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* JUMBO_MAE_B200_H_ */
