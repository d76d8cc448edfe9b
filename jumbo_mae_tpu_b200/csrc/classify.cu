// Classification loss of the fine-tune step (first piece): label preparation,
// softmax cross-entropy, its gradient and the top-1 / top-5 hit flags in one launch over the [B, C] logits.
//   src/finetuning.py:90-91   one_hot(labels) when labels are class indices, f32
//   src/finetuning.py:94      optax.smooth_labels: (1 - a) * labels + a / C
//   src/utils.py:70-71,82-83  Mixup / CutMix label side: r * labels + (1 - r) * labels[randperm]  (r a scalar)
//   src/finetuning.py:98      optax.softmax_cross_entropy: -sum(labels * log_softmax(logits))
//   src/finetuning.py:99-106  labels == max(labels); top_k(logits, 5); acc1 = hit[0], acc5 = any(hit)
// HBM-bound and tiny (B x C x 4 bytes, 4 MB at B = 1024, C = 1000): one block per row, the row and its prepared
// labels live in shared memory, every reduction is a block reduction.
#include "../../include/jumbo_mae_b200.h"
#include "common.cuh"

namespace jm {
namespace {

constexpr int kCeThreads = 128;
constexpr int kCeMaxClasses = 8192;  // 2 x 32 KB of dynamic shared memory

struct CeArgs {
  const float* logits;
  const float* labels_soft;
  const int* labels_idx;
  const int* perm;
  float ratio, one_minus_alpha, alpha_over_c, grad_scale, inv_b;
  const float* ratio_dev;  // optional: the label ratio read from device memory (a captured graph replays with new draws)
  int B, C;
  float *loss_rows, *acc1, *acc5, *metrics, *dlogits;
};

__device__ __forceinline__ float ce_block_sum(float v, float* red) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = (lane < kCeThreads / 32) ? red[lane] : 0.f;
  return warp_sum(t);
}
__device__ __forceinline__ float ce_block_max(float v, float* red) {
  v = warp_max(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = (lane < kCeThreads / 32) ? red[lane] : -INFINITY;
  return warp_max(t);
}
// jax.lax.top_k order: larger value first, lower index first among equal values.
__device__ __forceinline__ bool ce_before(float v, int i, float bv, int bi) {
  return v > bv || (v == bv && i < bi);
}

// Smoothed label of class c in row b (no rounding difference from the numpy statement: separate mul and add).
__device__ __forceinline__ float ce_smoothed(const CeArgs& a, int b, int c) {
  const float base = a.labels_soft ? a.labels_soft[static_cast<long>(b) * a.C + c]
                                   : (a.labels_idx[b] == c ? 1.f : 0.f);
  return __fadd_rn(__fmul_rn(a.one_minus_alpha, base), a.alpha_over_c);
}

__global__ void __launch_bounds__(kCeThreads) softmax_ce_kernel(const CeArgs a) {
  extern __shared__ float ce_smem[];
  __shared__ float red[32];
  __shared__ int red_i[32];
  float* lg = ce_smem;
  float* lab = ce_smem + a.C;
  const int b = blockIdx.x, C = a.C;
  const int other = a.perm ? a.perm[b] : b;
  const float ratio = a.ratio_dev ? *a.ratio_dev : a.ratio;
  const float one_minus_r = __fsub_rn(1.f, ratio);

  float mx = -INFINITY, lmx = -INFINITY, lsum = 0.f;
  for (int c = threadIdx.x; c < C; c += kCeThreads) {
    const float x = a.logits[static_cast<long>(b) * C + c];
    float l = ce_smoothed(a, b, c);
    if (a.perm) l = __fadd_rn(__fmul_rn(ratio, l), __fmul_rn(one_minus_r, ce_smoothed(a, other, c)));
    lg[c] = x, lab[c] = l;
    mx = fmaxf(mx, x), lmx = fmaxf(lmx, l), lsum += l;
  }
  mx = ce_block_max(mx, red);
  lmx = ce_block_max(lmx, red);
  lsum = ce_block_sum(lsum, red);

  float se = 0.f;
  for (int c = threadIdx.x; c < C; c += kCeThreads) se += expf(lg[c] - mx);
  se = ce_block_sum(se, red);
  const float log_se = logf(se);

  float loss = 0.f;
  for (int c = threadIdx.x; c < C; c += kCeThreads) {
    const float logp = (lg[c] - mx) - log_se;  // jax.nn.log_softmax: shifted - log(sum(exp(shifted)))
    loss -= lab[c] * logp;
    if (a.dlogits) a.dlogits[static_cast<long>(b) * C + c] = a.grad_scale * (expf(logp) * lsum - lab[c]);
  }
  loss = ce_block_sum(loss, red);

  // top-5 of the logits, one block arg-max per rank; already chosen indices are skipped.
  int chosen[5] = {-1, -1, -1, -1, -1};
  bool hit1 = false, hit5 = false;
  const int k = C < 5 ? C : 5;
  for (int r = 0; r < k; ++r) {
    float bv = -INFINITY;
    int bi = 0x7fffffff;
    for (int c = threadIdx.x; c < C; c += kCeThreads) {
      bool taken = false;
#pragma unroll
      for (int q = 0; q < 5; ++q) taken |= (chosen[q] == c);
      const float v = lg[c];
      if (!taken && (bi == 0x7fffffff || ce_before(v, c, bv, bi))) bv = v, bi = c;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (oi != 0x7fffffff && (bi == 0x7fffffff || ce_before(ov, oi, bv, bi))) bv = ov, bi = oi;
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();
    if (lane == 0) red[warp] = bv, red_i[warp] = bi;
    __syncthreads();
    bv = red[0], bi = red_i[0];
#pragma unroll
    for (int w = 1; w < kCeThreads / 32; ++w) {
      const float ov = red[w];
      const int oi = red_i[w];
      if (oi != 0x7fffffff && (bi == 0x7fffffff || ce_before(ov, oi, bv, bi))) bv = ov, bi = oi;
    }
    chosen[r] = bi;
    const bool hit = lab[bi] == lmx;
    if (r == 0) hit1 = hit;
    hit5 |= hit;
  }

  if (threadIdx.x == 0) {
    if (a.loss_rows) a.loss_rows[b] = loss;
    if (a.acc1) a.acc1[b] = hit1 ? 1.f : 0.f;
    if (a.acc5) a.acc5[b] = hit5 ? 1.f : 0.f;
    if (a.metrics) {
      atomicAdd(a.metrics + 0, loss * a.inv_b);
      atomicAdd(a.metrics + 1, hit1 ? a.inv_b : 0.f);
      atomicAdd(a.metrics + 2, hit5 ? a.inv_b : 0.f);
    }
  }
}

}  // namespace
}  // namespace jm

static int softmax_ce_impl(const float* logits, const float* labels_soft, const int* labels_idx, const int* perm,
                           float ratio, const float* ratio_dev, double label_smoothing, int B, int C, float grad_scale,
                           float* loss_rows, float* acc1, float* acc5, float* metrics, float* dlogits, void* stream) {
  using namespace jm;
  JM_CHECK(logits, "jm_softmax_ce: logits is null");
  JM_CHECK((labels_soft != nullptr) != (labels_idx != nullptr),
           "jm_softmax_ce: give exactly one of labels_soft [B, C] and labels_idx [B]");
  JM_CHECK(B >= 0 && C >= 1 && C <= kCeMaxClasses, "jm_softmax_ce: need B >= 0 and 1 <= C <= %d (got B=%d C=%d)",
           kCeMaxClasses, B, C);
  JM_CHECK(label_smoothing >= 0.0 && label_smoothing < 1.0, "jm_softmax_ce: label_smoothing must be in [0, 1)");
  if (B == 0) return 0;
  CeArgs a;
  a.logits = logits, a.labels_soft = labels_soft, a.labels_idx = labels_idx, a.perm = perm;
  a.ratio = ratio, a.ratio_dev = ratio_dev;
  a.one_minus_alpha = static_cast<float>(1.0 - label_smoothing);  // Python-float arithmetic, then f32
  a.alpha_over_c = static_cast<float>(label_smoothing / C);
  a.grad_scale = grad_scale, a.inv_b = 1.f / static_cast<float>(B);
  a.B = B, a.C = C;
  a.loss_rows = loss_rows, a.acc1 = acc1, a.acc5 = acc5, a.metrics = metrics, a.dlogits = dlogits;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const size_t smem = 2 * sizeof(float) * static_cast<size_t>(C);
  if (smem > 48 * 1024)
    JM_CUDA(cudaFuncSetAttribute(softmax_ce_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  if (metrics) JM_CUDA(cudaMemsetAsync(metrics, 0, 3 * sizeof(float), s));
  softmax_ce_kernel<<<B, kCeThreads, smem, s>>>(a);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_softmax_ce(const float* logits, const float* labels_soft, const int* labels_idx, const int* perm,
                             float ratio, double label_smoothing, int B, int C, float grad_scale, float* loss_rows,
                             float* acc1, float* acc5, float* metrics, float* dlogits, void* stream) {
  return softmax_ce_impl(logits, labels_soft, labels_idx, perm, ratio, nullptr, label_smoothing, B, C, grad_scale,
                         loss_rows, acc1, acc5, metrics, dlogits, stream);
}

extern "C" int jm_softmax_ce_dev(const float* logits, const float* labels_soft, const int* labels_idx, const int* perm,
                                 const float* ratio_dev, double label_smoothing, int B, int C, float grad_scale,
                                 float* loss_rows, float* acc1, float* acc5, float* metrics, float* dlogits,
                                 void* stream) {
  JM_CHECK(ratio_dev != nullptr, "jm_softmax_ce_dev: ratio_dev is null");
  return softmax_ce_impl(logits, labels_soft, labels_idx, perm, 1.f, ratio_dev, label_smoothing, B, C, grad_scale,
                         loss_rows, acc1, acc5, metrics, dlogits, stream);
}
