// XLA FFI handlers over the C ABI of libjumbo_mae_b200.so (include/jumbo_mae_b200.h): what a JAX host -- the
// reference's pretraining.py / modeling.py under jax.jit or jax.pmap -- registers with
// jax.ffi.register_ffi_target and calls with jax.ffi.ffi_call (ffi/jm_ffi.py does both).
//
// Build (needs jaxlib's headers; this image has neither jax nor jaxlib, so the file is compile-gated and only
// pre-processed here -- tests/test_ffi_shim.py):
//   g++ -std=c++17 -shared -fPIC -I$(python -c 'import jax.ffi; print(jax.ffi.include_dir())') -Iinclude \
//       -I/usr/local/cuda/include ffi/jm_ffi.cc -Ljumbo_mae_tpu_b200 -ljumbo_mae_b200 -o ffi/libjm_ffi.so
//
// Conventions: XLA owns every buffer; results are pre-allocated by XLA; scratch is an
// extra result (the caller declares it with the jm_*_workspace size); state updated in place is passed through
// input_output_aliases (jm_adamw, jm_clip_by_global_norm, accumulating GEMMs); handlers only enqueue on the
// given stream, never allocate, never synchronise, never throw; errors come back as xla::ffi::Error.  Handlers
// are re-entrant: the library keeps its per-device caches behind atomics / mutexes.
#if defined(__has_include)
#if __has_include("xla/ffi/api/ffi.h")
#define JM_HAVE_XLA_FFI 1
#endif
#endif

#ifdef JM_HAVE_XLA_FFI
#include <cuda_runtime_api.h>

#include "jumbo_mae_b200.h"
#include "xla/ffi/api/ffi.h"

namespace ffi = xla::ffi;
using Stream = ffi::PlatformStream<cudaStream_t>;
using F32 = ffi::Buffer<ffi::F32>;
using BF16 = ffi::Buffer<ffi::BF16>;
using S32 = ffi::Buffer<ffi::S32>;
using U8 = ffi::Buffer<ffi::U8>;
using RF32 = ffi::ResultBuffer<ffi::F32>;
using RBF16 = ffi::ResultBuffer<ffi::BF16>;
using RS32 = ffi::ResultBuffer<ffi::S32>;
using RU8 = ffi::ResultBuffer<ffi::U8>;

static ffi::Error Rc(int rc) {
  return rc == 0 ? ffi::Error::Success() : ffi::Error(ffi::ErrorCode::kInternal, jm_last_error());
}
static jm_bf16* bf(BF16& b) { return reinterpret_cast<jm_bf16*>(b.untyped_data()); }
static jm_bf16* bf(RBF16& b) { return reinterpret_cast<jm_bf16*>(b->untyped_data()); }
static long last_dim(const ffi::AnyBuffer::Dimensions& d) { return static_cast<long>(d.back()); }

// ---- random_masking integer half (src/utils_mae.py:90-96): noise [L] -> ids_shuffle, ids_restore, mask_row
static ffi::Error MaskSort(cudaStream_t s, F32 noise, int32_t keep, RS32 ids_shuffle, RS32 ids_restore, RF32 mask_row) {
  return Rc(jm_mask_sort(noise.typed_data(), static_cast<int>(noise.element_count()), keep, ids_shuffle->typed_data(),
                         ids_restore->typed_data(), mask_row->typed_data(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmMaskSort, MaskSort,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Attr<int32_t>("keep").Ret<S32>().Ret<S32>().Ret<F32>());

// ---- index_sequence (src/utils_mae.py:84-85): x [B, Lin, D], ids [Lout] -> out [B, Lout, D]
static ffi::Error IndexSequence(cudaStream_t s, F32 x, S32 ids, RF32 out) {
  const auto d = x.dimensions();
  return Rc(jm_index_sequence(x.typed_data(), ids.typed_data(), static_cast<int>(d[0]), static_cast<int>(d[1]),
                              static_cast<int>(ids.element_count()), static_cast<int>(d[2]), out->typed_data(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmIndexSequence, IndexSequence, ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Arg<S32>().Ret<F32>());

// ---- pixel normalisation + extract_patches + gather (src/pretraining.py:90-91, src/utils_mae.py:67-73,93)
static ffi::Error Patchify(cudaStream_t s, U8 images, S32 ids, int32_t patch_size, int32_t rows_per_image,
                           int32_t row_off, RBF16 out) {
  const auto d = images.dimensions();
  return Rc(jm_patchify(images.typed_data(), static_cast<int>(d[0]), static_cast<int>(d[2]), patch_size,
                        ids.typed_data(), static_cast<int>(ids.element_count()), rows_per_image, row_off, bf(out),
                        last_dim(out->dimensions()), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmPatchify, Patchify,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<U8>().Arg<S32>().Attr<int32_t>("patch_size")
                                  .Attr<int32_t>("rows_per_image").Attr<int32_t>("row_off").Ret<BF16>());

// ---- extract_patches on the normalised image, f32 (src/utils_mae.py:67-73)
static ffi::Error ExtractPatches(cudaStream_t s, U8 images, int32_t patch_size, RF32 out) {
  const auto d = images.dimensions();
  return Rc(jm_extract_patches(images.typed_data(), static_cast<int>(d[0]), static_cast<int>(d[2]), patch_size,
                               out->typed_data(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmExtractPatches, ExtractPatches,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<U8>().Attr<int32_t>("patch_size").Ret<F32>());

// ---- fused target + patch_mse_loss + gradient (src/pretraining.py:113-120); scratch is a result
static ffi::Error PatchLoss(cudaStream_t s, F32 pred, U8 images, F32 mask_row, int32_t patch_size, int32_t num_cls,
                            int32_t norm_pix, float grad_scale, RF32 loss, RBF16 dpred, RU8 scratch) {
  const auto d = images.dimensions();
  return Rc(jm_patch_loss(pred.typed_data(), last_dim(pred.dimensions()), images.typed_data(), mask_row.typed_data(),
                          static_cast<int>(d[0]), static_cast<int>(d[2]), patch_size, num_cls, norm_pix, grad_scale,
                          loss->typed_data(), bf(dpred), last_dim(dpred->dimensions()), scratch->untyped_data(),
                          scratch->size_bytes(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmPatchLoss, PatchLoss,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Arg<U8>().Arg<F32>().Attr<int32_t>("patch_size")
                                  .Attr<int32_t>("num_cls").Attr<int32_t>("norm_pix").Attr<float>("grad_scale")
                                  .Ret<F32>().Ret<BF16>().Ret<U8>());

// ---- patch_mse_loss(patch_output, patch_target, valid) (src/utils_mae.py:51-64)
static ffi::Error PatchMseLoss(cudaStream_t s, F32 out, F32 tgt, F32 valid, RF32 loss, RU8 scratch) {
  const auto d = out.dimensions();
  return Rc(jm_patch_mse_loss(out.typed_data(), static_cast<long>(d[2]), tgt.typed_data(), static_cast<long>(d[2]),
                              valid.typed_data(), static_cast<int>(d[0]), static_cast<int>(d[1]), static_cast<int>(d[2]),
                              loss->typed_data(), scratch->untyped_data(), scratch->size_bytes(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmPatchMseLoss, PatchMseLoss,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Arg<F32>().Arg<F32>().Ret<F32>().Ret<U8>());

// ---- LayerNorm forward (src/modeling.py:155-156,177-179,238,282): bf16 copy for the next GEMM + statistics
static ffi::Error LayerNormFwd(cudaStream_t s, F32 x, F32 gamma, F32 beta, float eps, RBF16 y, RF32 mean, RF32 rstd) {
  const long dim = last_dim(x.dimensions());
  return Rc(jm_layernorm_fwd(x.typed_data(), dim, static_cast<int>(x.element_count() / dim), static_cast<int>(dim),
                             gamma.typed_data(), beta.typed_data(), eps, bf(y), dim, nullptr, 0, mean->typed_data(),
                             rstd->typed_data(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmLayerNormFwd, LayerNormFwd,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Arg<F32>().Arg<F32>().Attr<float>("eps")
                                  .Ret<BF16>().Ret<F32>().Ret<F32>());

// ---- LayerNorm backward: dx (f32 + bf16 copy), dgamma / dbeta accumulate in place (aliased operands)
static ffi::Error LayerNormBwd(cudaStream_t s, BF16 dy, F32 x, F32 mean, F32 rstd, F32 gamma, F32 dres, F32 dgamma_in,
                               F32 dbeta_in, RF32 dx, RBF16 dx16, RF32 dgamma, RF32 dbeta, RU8 scratch) {
  (void)dgamma_in, (void)dbeta_in;  // aliased to the results by input_output_aliases
  const long dim = last_dim(x.dimensions());
  return Rc(jm_layernorm_bwd(bf(dy), dim, x.typed_data(), dim, mean.typed_data(), rstd.typed_data(), gamma.typed_data(),
                             static_cast<int>(x.element_count() / dim), static_cast<int>(dim), dres.typed_data(), dim,
                             dx->typed_data(), dim, bf(dx16), dim, 0, 0, dgamma->typed_data(), dbeta->typed_data(),
                             nullptr, 0, 0, scratch->untyped_data(), scratch->size_bytes(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmLayerNormBwd, LayerNormBwd,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<BF16>().Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>()
                                  .Arg<F32>().Arg<F32>().Arg<F32>().Ret<F32>().Ret<BF16>().Ret<F32>().Ret<F32>().Ret<U8>());

// ---- Dense forward: out_bf16 = epilogue(A @ W + bias) (src/modeling.py:108-148); act 0 none / 3 GELU (two results)
static ffi::Error Dense(cudaStream_t s, BF16 a, BF16 w, F32 bias, int32_t act, float col_scale, int32_t col_scale_n,
                        RBF16 out, RBF16 out2) {
  jm_gemm_args g = {};
  const auto ad = a.dimensions(), wd = w.dimensions();
  g.K = static_cast<int>(ad.back()), g.M = static_cast<int>(a.element_count() / g.K), g.N = static_cast<int>(wd.back());
  g.A = bf(a), g.lda = g.K, g.B = bf(w), g.ldb = g.N, g.b_mn_major = 1;  // Flax kernel is [in, out]
  g.bias = bias.typed_data(), g.act = act, g.col_scale = col_scale, g.col_scale_n = col_scale_n;
  g.out_bf16 = bf(out), g.ldo_bf16 = g.N;
  if (act == 1 || act == 3) g.out2_bf16 = bf(out2), g.ldo2_bf16 = g.N;
  return Rc(jm_gemm(&g, s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmDense, Dense,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<BF16>().Arg<BF16>().Arg<F32>().Attr<int32_t>("act")
                                  .Attr<float>("col_scale").Attr<int32_t>("col_scale_n").Ret<BF16>().Ret<BF16>());

// ---- pixel head with the loss in its epilogue, on the masked rows only (src/pretraining.py:110-120): targets of the
// masked patches, dec_norm gathered through row_map, then act = 5 of jm_gemm and the ordered sum of its row partials
static ffi::Error MaskedTargets(cudaStream_t s, U8 images, S32 ids, int32_t patch_size, int32_t num_cls, int32_t norm_pix,
                                RU8 targets, RF32 stats, RS32 row_map) {
  const auto d = images.dimensions();  // [B, 3, S, S]
  return Rc(jm_masked_targets(images.typed_data(), static_cast<int>(d[0]), static_cast<int>(d[3]), patch_size,
                              ids.typed_data(), static_cast<int>(ids.element_count()), num_cls, norm_pix,
                              targets->typed_data(), last_dim(targets->dimensions()), stats->typed_data(),
                              row_map->typed_data(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmMaskedTargets, MaskedTargets,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<U8>().Arg<S32>().Attr<int32_t>("patch_size")
                                  .Attr<int32_t>("num_cls").Attr<int32_t>("norm_pix").Ret<U8>().Ret<F32>().Ret<S32>());

static ffi::Error LayerNormFwdRows(cudaStream_t s, F32 x, S32 row_map, F32 gamma, F32 beta, float eps, RBF16 y, RF32 mean,
                                   RF32 rstd) {
  const long dim = last_dim(x.dimensions());
  return Rc(jm_layernorm_fwd_rows(x.typed_data(), dim, row_map.typed_data(), static_cast<int>(row_map.element_count()),
                                  static_cast<int>(dim), gamma.typed_data(), beta.typed_data(), eps, bf(y), dim,
                                  mean->typed_data(), rstd->typed_data(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmLayerNormFwdRows, LayerNormFwdRows,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Arg<S32>().Arg<F32>().Arg<F32>().Attr<float>("eps")
                                  .Ret<BF16>().Ret<F32>().Ret<F32>());

static ffi::Error LayerNormBwdRows(cudaStream_t s, BF16 dy, F32 x, S32 row_map, F32 mean, F32 rstd, F32 gamma,
                                   F32 dgamma_in, F32 dbeta_in, RF32 dx, RBF16 dx16, RF32 dgamma, RF32 dbeta, RU8 scratch) {
  (void)dgamma_in, (void)dbeta_in;  // aliased to the results by input_output_aliases
  const long dim = last_dim(x.dimensions());
  return Rc(jm_layernorm_bwd_rows(bf(dy), dim, x.typed_data(), dim, row_map.typed_data(), mean.typed_data(),
                                  rstd.typed_data(), gamma.typed_data(), static_cast<int>(row_map.element_count()),
                                  static_cast<int>(dim), static_cast<long>(x.element_count() / dim), dx->typed_data(), dim,
                                  bf(dx16), dim, dgamma->typed_data(), dbeta->typed_data(), nullptr,
                                  scratch->untyped_data(), scratch->size_bytes(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmLayerNormBwdRows, LayerNormBwdRows,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<BF16>().Arg<F32>().Arg<S32>().Arg<F32>().Arg<F32>()
                                  .Arg<F32>().Arg<F32>().Arg<F32>().Ret<F32>().Ret<BF16>().Ret<F32>().Ret<F32>().Ret<U8>());

static ffi::Error HeadLoss(cudaStream_t s, BF16 a, BF16 w, F32 bias, U8 targets, F32 stats, int32_t norm_pix,
                           float loss_scale, float mean_scale, RF32 loss, RBF16 dpred, RF32 row_loss, RU8 scratch) {
  jm_gemm_args g = {};
  g.K = static_cast<int>(a.dimensions().back()), g.M = static_cast<int>(a.element_count() / g.K);
  g.N = static_cast<int>(w.dimensions().back());
  g.A = bf(a), g.lda = g.K, g.B = bf(w), g.ldb = g.N, g.b_mn_major = 1, g.bias = bias.typed_data(), g.act = 5;
  g.out_bf16 = bf(dpred), g.ldo_bf16 = last_dim(dpred->dimensions());
  g.loss_target = targets.typed_data(), g.ld_target = last_dim(targets.dimensions());
  g.loss_stats = norm_pix ? stats.typed_data() : nullptr, g.row_loss = row_loss->typed_data(), g.loss_scale = loss_scale;
  if (int rc = jm_gemm(&g, s)) return Rc(rc);
  return Rc(jm_row_loss_finalize(row_loss->typed_data(), static_cast<long>(row_loss->element_count()), mean_scale,
                                 loss->typed_data(), scratch->untyped_data(), scratch->size_bytes(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmHeadLoss, HeadLoss,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<BF16>().Arg<BF16>().Arg<F32>().Arg<U8>().Arg<F32>()
                                  .Attr<int32_t>("norm_pix").Attr<float>("loss_scale").Attr<float>("mean_scale")
                                  .Ret<F32>().Ret<BF16>().Ret<F32>().Ret<U8>());

// ---- Dense forward into the f32 residual stream: out = A @ W + bias + residual (src/modeling.py:165-166,192,201)
static ffi::Error DenseResidual(cudaStream_t s, BF16 a, BF16 w, F32 bias, F32 residual, RF32 out) {
  jm_gemm_args g = {};
  g.K = static_cast<int>(a.dimensions().back()), g.M = static_cast<int>(a.element_count() / g.K);
  g.N = static_cast<int>(w.dimensions().back());
  g.A = bf(a), g.lda = g.K, g.B = bf(w), g.ldb = g.N, g.b_mn_major = 1;
  g.bias = bias.typed_data(), g.residual = residual.typed_data(), g.ldr = g.N, g.out_f32 = out->typed_data(), g.ldo_f32 = g.N;
  return Rc(jm_gemm(&g, s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmDenseResidual, DenseResidual,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<BF16>().Arg<BF16>().Arg<F32>().Arg<F32>().Ret<F32>());

// ---- Dense backward: dX = dY @ W^T (bf16), dW += X^T dY (f32, aliased in place)
static ffi::Error DenseBwd(cudaStream_t s, BF16 x, BF16 dy, BF16 w, F32 dw_in, RBF16 dx, RF32 dw) {
  (void)dw_in;
  jm_gemm_args g = {};
  const int K = static_cast<int>(x.dimensions().back()), N = static_cast<int>(dy.dimensions().back());
  const int M = static_cast<int>(x.element_count() / K);
  g.M = K, g.N = N, g.K = M, g.A = bf(x), g.lda = K, g.a_mn_major = 1, g.B = bf(dy), g.ldb = N, g.b_mn_major = 1;
  g.out_f32 = dw->typed_data(), g.ldo_f32 = N, g.accumulate = 1;
  if (int rc = jm_gemm(&g, s)) return Rc(rc);
  jm_gemm_args h = {};
  h.M = M, h.N = K, h.K = N, h.A = bf(dy), h.lda = N, h.B = bf(w), h.ldb = N, h.b_mn_major = 0;
  h.out_bf16 = bf(dx), h.ldo_bf16 = K;
  return Rc(jm_gemm(&h, s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmDenseBwd, DenseBwd,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<BF16>().Arg<BF16>().Arg<BF16>().Arg<F32>().Ret<BF16>().Ret<F32>());

// ---- attention core softmax(q k^T) v (src/modeling.py:136-137), qkv [B*T, 3*heads*hd] with q pre-scaled
static ffi::Error AttentionFwd(cudaStream_t s, BF16 qkv, int32_t batch, int32_t heads, RBF16 out, RF32 lse) {
  const long ld = last_dim(qkv.dimensions());
  const int T = static_cast<int>(qkv.element_count() / ld / batch), hd = static_cast<int>(ld / 3 / heads);
  return Rc(jm_attention_fwd(bf(qkv), ld, batch, T, heads, hd, bf(out), ld / 3, lse->typed_data(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmAttentionFwd, AttentionFwd,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<BF16>().Attr<int32_t>("batch").Attr<int32_t>("heads")
                                  .Ret<BF16>().Ret<F32>());

static ffi::Error AttentionBwd(cudaStream_t s, BF16 qkv, BF16 out, BF16 dout, F32 lse, F32 dbias_in, int32_t batch,
                               int32_t heads, float q_scale, RBF16 dqkv, RF32 dbias) {
  (void)dbias_in;  // aliased: the fused q/k/v bias gradient accumulates in place
  const long ld = last_dim(qkv.dimensions());
  const int T = static_cast<int>(qkv.element_count() / ld / batch), hd = static_cast<int>(ld / 3 / heads);
  return Rc(jm_attention_bwd(bf(qkv), ld, bf(out), ld / 3, bf(dout), ld / 3, lse.typed_data(), nullptr, batch, T, heads,
                             hd, q_scale, bf(dqkv), ld, dbias->typed_data(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmAttentionBwd, AttentionBwd,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<BF16>().Arg<BF16>().Arg<BF16>().Arg<F32>().Arg<F32>()
                                  .Attr<int32_t>("batch").Attr<int32_t>("heads").Attr<float>("q_scale").Ret<BF16>().Ret<F32>());

// ---- CLS tile + concat (src/modeling.py:249-250,257): writes the J CLS rows of every image of x (aliased)
static ffi::Error FillCls(cudaStream_t s, F32 x_in, F32 cls, int32_t batch, RF32 x) {
  (void)x_in;
  const auto cd = cls.dimensions();  // [1, J, d]
  const int J = static_cast<int>(cd[cd.size() - 2]), d = static_cast<int>(cd.back());
  const int T = static_cast<int>(x->element_count() / d / batch);
  return Rc(jm_fill_cls(x->typed_data(), batch, T, J, d, cls.typed_data(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmFillCls, FillCls,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Arg<F32>().Attr<int32_t>("batch").Ret<F32>());

// ---- decoder input assembly (src/pretraining.py:96-106, src/modeling.py:289-292) and its transpose
static ffi::Error UnshuffleFwd(cudaStream_t s, F32 y, S32 ids_restore, F32 mask_emb, F32 dec_wpe, int32_t batch,
                               int32_t n, int32_t J, RF32 out) {
  const int N = static_cast<int>(ids_restore.element_count()), dd = static_cast<int>(last_dim(y.dimensions()));
  return Rc(jm_unshuffle_fwd(y.typed_data(), ids_restore.typed_data(), mask_emb.typed_data(), dec_wpe.typed_data(), batch,
                             N, n, J, dd, out->typed_data(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmUnshuffleFwd, UnshuffleFwd,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Arg<S32>().Arg<F32>().Arg<F32>().Attr<int32_t>("batch")
                                  .Attr<int32_t>("n").Attr<int32_t>("J").Ret<F32>());
static ffi::Error UnshuffleBwd(cudaStream_t s, F32 dout, S32 ids_restore, F32 dmask_in, int32_t batch, int32_t n, int32_t J,
                               RBF16 dy, RF32 dmask_emb) {
  (void)dmask_in;
  const int N = static_cast<int>(ids_restore.element_count()), dd = static_cast<int>(last_dim(dout.dimensions()));
  return Rc(jm_unshuffle_bwd(dout.typed_data(), ids_restore.typed_data(), batch, N, n, J, dd, bf(dy),
                             dmask_emb->typed_data(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmUnshuffleBwd, UnshuffleBwd,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Arg<S32>().Arg<F32>().Attr<int32_t>("batch")
                                  .Attr<int32_t>("n").Attr<int32_t>("J").Ret<BF16>().Ret<F32>());

// ---- bias gradients: out += column sums of x (aliased)
static ffi::Error Colsum(cudaStream_t s, BF16 x, F32 out_in, RF32 out) {
  (void)out_in;
  const long cols = last_dim(x.dimensions());
  return Rc(jm_colsum(bf(x), 0, cols, static_cast<int>(x.element_count() / cols), static_cast<int>(cols), 0, 0, 0,
                      out->typed_data(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmColsum, Colsum, ffi::Ffi::Bind().Ctx<Stream>().Arg<BF16>().Arg<F32>().Ret<F32>());

// ---- modified_lamb (src/utils.py:124-139) on the kernel leaves: group / chunk tables as int64 operands, p, m, v aliased
using S64 = ffi::Buffer<ffi::S64>;
static ffi::Error Lamb(cudaStream_t s, F32 p_in, F32 g, F32 m_in, F32 v_in, S64 groups, S64 chunks, float lr, float b1,
                       float b2, float eps, float wd, float bc1, float bc2, float grad_scale, RF32 p, RF32 m, RF32 v,
                       RBF16 p16, RU8 scratch) {
  (void)p_in, (void)m_in, (void)v_in;
  return Rc(jm_lamb(p->typed_data(), g.typed_data(), m->typed_data(), v->typed_data(), bf(p16),
                    reinterpret_cast<const int64_t*>(groups.typed_data()), static_cast<int>(groups.dimensions()[0]),
                    reinterpret_cast<const int64_t*>(chunks.typed_data()), static_cast<int>(chunks.dimensions()[0]), lr,
                    b1, b2, eps, wd, bc1, bc2, grad_scale, scratch->untyped_data(), scratch->size_bytes(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmLamb, Lamb,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>().Arg<S64>().Arg<S64>()
                                  .Attr<float>("lr").Attr<float>("b1").Attr<float>("b2").Attr<float>("eps").Attr<float>("wd")
                                  .Attr<float>("bc1").Attr<float>("bc2").Attr<float>("grad_scale")
                                  .Ret<F32>().Ret<F32>().Ret<F32>().Ret<BF16>().Ret<U8>());

// ---- LayerScale (src/modeling.py:159-166,185-201): d scale of one scaled Dense from its accumulated gradients
static ffi::Error LayerScaleGrad(cudaStream_t s, F32 w, F32 dw, F32 b, F32 db, F32 scale, F32 dscale_in, RF32 dscale) {
  (void)dscale_in;  // aliased to the result
  const auto d = w.dimensions();
  return Rc(jm_layerscale_grad(w.typed_data(), dw.typed_data(), d[1], static_cast<int>(d[0]), static_cast<int>(d[1]),
                               b.typed_data(), db.typed_data(), scale.typed_data(), dscale->typed_data(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmLayerScaleGrad, LayerScaleGrad,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>()
                                  .Ret<F32>());

// ---- optax.adamw + apply_updates over the flat buffers (src/pretraining.py:223-234,150); p, m, v aliased in place
static ffi::Error AdamW(cudaStream_t s, F32 p_in, F32 g, F32 m_in, F32 v_in, int64_t n_decay, float lr, float b1, float b2,
                        float eps, float wd, float bc1, float bc2, float grad_scale, RF32 p, RF32 m, RF32 v, RBF16 p16) {
  (void)p_in, (void)m_in, (void)v_in;
  return Rc(jm_adamw(p->typed_data(), g.typed_data(), m->typed_data(), v->typed_data(), bf(p16),
                     static_cast<long>(g.element_count()), static_cast<long>(n_decay), lr, b1, b2, eps, wd, bc1, bc2,
                     grad_scale, s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmAdamW, AdamW,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>()
                                  .Attr<int64_t>("n_decay").Attr<float>("lr").Attr<float>("b1").Attr<float>("b2")
                                  .Attr<float>("eps").Attr<float>("wd").Attr<float>("bc1").Attr<float>("bc2")
                                  .Attr<float>("grad_scale").Ret<F32>().Ret<F32>().Ret<F32>().Ret<BF16>());

// ---- optax.clip_by_global_norm (src/pretraining.py:243-244); g aliased in place, sumsq is scratch
static ffi::Error ClipByGlobalNorm(cudaStream_t s, F32 g_in, float pre_scale, float max_norm, RF32 g, RF32 sumsq) {
  (void)g_in;
  return Rc(jm_clip_by_global_norm(g->typed_data(), static_cast<long>(g->element_count()), pre_scale, max_norm,
                                   sumsq->typed_data(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmClipByGlobalNorm, ClipByGlobalNorm,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Attr<float>("pre_scale").Attr<float>("max_norm")
                                  .Ret<F32>().Ret<F32>());

// ---- f32 -> bf16 compute copy of the parameters
static ffi::Error CastBf16(cudaStream_t s, F32 x, RBF16 y) {
  return Rc(jm_cast_bf16(x.typed_data(), bf(y), static_cast<long>(x.element_count()), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmCastBf16, CastBf16, ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Ret<BF16>());

// ---- fine-tune loss end (src/finetuning.py:90-106): smoothed / mixed labels, CE, top-1 / top-5, logits gradient
static ffi::Error SoftmaxCe(cudaStream_t s, F32 logits, S32 labels, S32 perm, float ratio, float label_smoothing,
                            float grad_scale, RF32 loss_rows, RF32 acc1, RF32 acc5, RF32 metrics, RF32 dlogits) {
  const auto d = logits.dimensions();
  return Rc(jm_softmax_ce(logits.typed_data(), nullptr, labels.typed_data(), perm.typed_data(), ratio, label_smoothing,
                          static_cast<int>(d[0]), static_cast<int>(d[1]), grad_scale, loss_rows->typed_data(),
                          acc1->typed_data(), acc5->typed_data(), metrics->typed_data(), dlogits->typed_data(), s));
}
XLA_FFI_DEFINE_HANDLER_SYMBOL(JmSoftmaxCe, SoftmaxCe,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Arg<S32>().Arg<S32>().Attr<float>("ratio")
                                  .Attr<float>("label_smoothing").Attr<float>("grad_scale").Ret<F32>().Ret<F32>().Ret<F32>()
                                  .Ret<F32>().Ret<F32>());

#else  // !JM_HAVE_XLA_FFI
// jaxlib's headers are not on the include path: nothing to build (see the header comment).
extern "C" int jm_ffi_unavailable(void) { return 1; }
#endif
