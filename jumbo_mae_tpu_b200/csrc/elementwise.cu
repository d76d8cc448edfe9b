// HBM-bound kernels of the MAE step: masking sort, pixel pre-processing / patchify,
// masked-patch MSE, LayerNorm forward/backward, token assembly, bias-gradient column
// sums and the fused AdamW.  All are coalesced, vectorised (8/16-byte accesses) and
// sized from the problem, not from a tile shape: none of them has data reuse to stage.
#include "../../include/jumbo_mae_b200.h"
#include "common.cuh"

namespace jm {

// src/dataset.py:37-38, demoted to f32 exactly as JAX does at src/pretraining.py:91
__constant__ float c_mean[3] = {0.485f, 0.456f, 0.406f};
__constant__ float c_std[3] = {0.229f, 0.224f, 0.225f};

__device__ __forceinline__ float norm_pixel(uint8_t u, int c) {
  // (u / 255 - mean) / std with IEEE ops, no contraction: matches the f32 oracle bit for bit
  return __fdiv_rn(__fsub_rn(__fdiv_rn(static_cast<float>(u), 255.0f), c_mean[c]), c_std[c]);
}

// The 768 possible values of norm_pixel, computed once per block with the same IEEE operations (two divisions
// per element cost ~40 instructions; a row kernel that recomputes them per element is ALU-bound).
__device__ __forceinline__ void fill_norm_lut(float* lut /* [3][256] in shared memory */) {
  for (int i = threadIdx.x; i < 768; i += blockDim.x) lut[i] = norm_pixel(static_cast<uint8_t>(i & 255), i >> 8);
  __syncthreads();
}

__device__ __forceinline__ float block_sum(float v, float* red /* >= 32 floats */) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = (lane < nw) ? red[lane] : 0.f;
  t = warp_sum(t);
  return t;
}

// ---------------------------------------------------------------- masking sort
// Stable ascending argsort by rank counting in shared memory: rank(i) = #{j : key_j < key_i or
// (key_j == key_i and j < i)}.  length <= 1024, one thread per element, bit-exact and
// tie-stable like jnp.argsort (NaN sorts last).
__global__ void mask_sort_kernel(const float* __restrict__ noise, int L, int keep, int* __restrict__ ids_shuffle,
                                 int* __restrict__ ids_restore, float* __restrict__ mask_row) {
  extern __shared__ float s_noise[];
  const int i = threadIdx.x;
  // one block per row of a [batch, L] problem (per-image masking); the reference's shared permutation is batch = 1
  noise += static_cast<long>(blockIdx.x) * L, ids_shuffle += static_cast<long>(blockIdx.x) * L;
  ids_restore += static_cast<long>(blockIdx.x) * L, mask_row += static_cast<long>(blockIdx.x) * L;
  if (i < L) s_noise[i] = noise[i];
  __syncthreads();
  if (i >= L) return;
  const float me = s_noise[i];
  const bool me_nan = me != me;
  int rank = 0;
  for (int j = 0; j < L; ++j) {
    const float o = s_noise[j];
    const bool o_nan = o != o;
    bool less;
    if (me_nan) less = !o_nan || j < i;           // every non-NaN precedes a NaN; NaNs keep index order
    else less = !o_nan && (o < me || (o == me && j < i));
    rank += less ? 1 : 0;
  }
  ids_shuffle[rank] = i;
  ids_restore[i] = rank;
  mask_row[i] = rank >= keep ? 1.0f : 0.0f;
}

__global__ void index_sequence_kernel(const float* __restrict__ x, const int* __restrict__ ids, int length_in,
                                      int length_out, int dim, float* __restrict__ out) {
  const int i = blockIdx.x, b = blockIdx.y;
  const float* src = x + (static_cast<long>(b) * length_in + ids[i]) * dim;
  float* dst = out + (static_cast<long>(b) * length_out + i) * dim;
  for (int c = threadIdx.x; c < dim; c += blockDim.x) dst[c] = src[c];
}

// ---------------------------------------------------------------- patchify
// Copies one patch (3 planes of p x p uint8) into spix[c][ph][pw] with the lanes of one warp: a run of p
// contiguous bytes per (plane, row); 16-byte vector copies when p == 16 and the image rows keep them aligned.
__device__ __forceinline__ void stage_patch(uint8_t* spix, const uint8_t* img, int S, int p, int gy, int gx, int lane) {
  const int runs = 3 * p;
  if (p == 16 && (S & 15) == 0 && (reinterpret_cast<uintptr_t>(img) & 15) == 0) {  // 16-byte rows: vector loads
    for (int r = lane; r < runs; r += 32) {
      const int c = r >> 4, ph = r & 15;
      *reinterpret_cast<uint4*>(spix + r * 16) =
          *reinterpret_cast<const uint4*>(img + (static_cast<long>(c) * S + gy * 16 + ph) * S + gx * 16);
    }
  } else {
    for (int r = lane; r < runs; r += 32) {
      const int c = r / p, ph = r - c * p;
      const uint8_t* src = img + (static_cast<long>(c) * S + gy * p + ph) * S + gx * p;
      for (int i = 0; i < p; ++i) spix[r * p + i] = src[i];
    }
  }
}


// One warp per output row, eight rows in flight per block (see patch_loss_kernel for why).
__global__ void __launch_bounds__(256)
patchify_kernel(const uint8_t* __restrict__ images, int S, int p, const int* __restrict__ ids, int n_sel,
                int rows_per_image, int row_off, __nv_bfloat16* __restrict__ out, long ldo, int batch, long ids_stride) {
  __shared__ __align__(16) uint8_t spix_all[8][2048];  // per warp: the patch's 3 planes of p x p bytes (P <= 2048, checked)
  __shared__ float lut[768];
  fill_norm_lut(lut);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* spix = spix_all[warp];
  const int pp = p * p, P = pp * 3, g = S / p;
  const long rows = static_cast<long>(batch) * rows_per_image;
  for (long row = static_cast<long>(blockIdx.x) * 8 + warp; row < rows; row += static_cast<long>(gridDim.x) * 8) {
    const int b = static_cast<int>(row / rows_per_image), r = static_cast<int>(row - static_cast<long>(b) * rows_per_image);
    __nv_bfloat16* dst = out + row * ldo;
    if (r < row_off || r >= row_off + n_sel) {
      for (int e = lane * 4; e < ldo; e += 128) *reinterpret_cast<uint2*>(dst + e) = make_uint2(0u, 0u);
      continue;
    }
    const int patch = ids ? ids[b * ids_stride + r - row_off] : (r - row_off);
    const int gy = patch / g, gx = patch - gy * g;
    const uint8_t* img = images + static_cast<long>(b) * 3 * S * S;
    __syncwarp();
    stage_patch(spix, img, S, p, gy, gx, lane);
    __syncwarp();
    for (int e0 = lane * 4; e0 < ldo; e0 += 128) {  // ldo is a multiple of 4 (checked)
      float v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int e = e0 + j;
        v[j] = 0.f;
        if (e < P) {
          const int pix = e / 3, c = e - pix * 3;
          v[j] = lut[c * 256 + spix[c * pp + pix]];
        }
      }
      *reinterpret_cast<uint2*>(dst + e0) = make_uint2(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]));
    }
  }
}

// Targets of the masked patches for the loss epilogue of the pixel-head GEMM (src/pretraining.py:113-118): the raw
// bytes of each selected patch in patch-vector order (pixel-major, channel fastest: extract_patches'), one compact row
// per (image, selected patch), plus -- for norm_pix_loss -- the patch's mean and 1/sqrt(biased var + 1e-6) over its
// normalised pixels, and the decoder row each target belongs to.  One warp per row.
__global__ void __launch_bounds__(256)
masked_targets_kernel(const uint8_t* __restrict__ images, int S, int p, const int* __restrict__ ids, int n_sel, int J,
                      int batch, int norm_pix, uint8_t* __restrict__ tgt, long ldt, float* __restrict__ stats,
                      int* __restrict__ row_map, long ids_stride) {
  __shared__ __align__(16) uint8_t spix_all[8][2048];
  __shared__ float lut[768];
  fill_norm_lut(lut);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* spix = spix_all[warp];
  const int pp = p * p, P = pp * 3, g = S / p, N = g * g;
  const long rows = static_cast<long>(batch) * n_sel;
  for (long row = static_cast<long>(blockIdx.x) * 8 + warp; row < rows; row += static_cast<long>(gridDim.x) * 8) {
    const int b = static_cast<int>(row / n_sel), k = static_cast<int>(row - static_cast<long>(b) * n_sel);
    const int patch = ids[b * ids_stride + k];
    const int gy = patch / g, gx = patch - gy * g;
    if (row_map && lane == 0) row_map[row] = b * (J + N) + J + patch;
    __syncwarp();
    stage_patch(spix, images + static_cast<long>(b) * 3 * S * S, S, p, gy, gx, lane);
    __syncwarp();
    uint8_t* dst = tgt + row * ldt;
    for (int e0 = lane * 4; e0 < P; e0 += 128) {  // P and ldt are multiples of 4 (checked)
      uint32_t w = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int e = e0 + j, pix = e / 3, c = e - pix * 3;
        w |= static_cast<uint32_t>(spix[c * pp + pix]) << (8 * j);
      }
      *reinterpret_cast<uint32_t*>(dst + e0) = w;
    }
    if (norm_pix) {  // same two passes, in the same order, as patch_loss_kernel
      auto target = [&](int e) {
        const int pix = e / 3, c = e - pix * 3;
        return lut[c * 256 + spix[c * pp + pix]];
      };
      float s_ = 0.f;
      for (int e = lane; e < P; e += 32) s_ += target(e);
      const float mean = warp_sum(s_) / P;
      float ss = 0.f;
      for (int e = lane; e < P; e += 32) {
        const float d = target(e) - mean;
        ss += d * d;
      }
      const float inv = 1.0f / sqrtf(warp_sum(ss) / P + 1e-6f);
      if (lane == 0) stats[2 * row] = mean, stats[2 * row + 1] = inv;
    }
  }
}

// Sum of a [rows, parts] matrix of per-row loss partials, in a fixed order: per-block partial sums, then one block.
__global__ void __launch_bounds__(256) row_loss_partial_kernel(const float* __restrict__ row_loss, long total,
                                                               float* __restrict__ partial) {
  __shared__ float red[32];
  const long per = (total + gridDim.x - 1) / gridDim.x;
  const long lo = per * blockIdx.x, hi = lo + per < total ? lo + per : total;
  float s_ = 0.f;
  for (long i = lo + threadIdx.x; i < hi; i += blockDim.x) s_ += row_loss[i];
  const float t = block_sum(s_, red);
  if (threadIdx.x == 0) partial[blockIdx.x] = t;
}
__global__ void __launch_bounds__(256) row_loss_final_kernel(const float* __restrict__ partial, int n, float scale,
                                                             float* __restrict__ loss) {
  __shared__ float red[32];
  float s_ = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) s_ += partial[i];
  const float t = block_sum(s_, red);
  if (threadIdx.x == 0) loss[0] = t * scale;
}

// Fine-tune input path: patchify with Mixup / CutMix applied to the NORMALISED pixels
// (src/finetuning.py:87-88,95; src/utils.py:67-90), one warp per output row as above, both source patches staged.
//   mode 1 (Mixup.apply_mixup):  v = ratio * a + (1 - ratio) * b      (separate roundings, as the f32 array ops)
//   mode 2 (Mixup.apply_cutmix): v = inside_box(y, x) ? b : a         (image_mask * a + (1 - image_mask) * b with a
//                                                                     0/1 mask), inside = xmask[x] & ymask[y]
// with b = images[perm[image]].
__global__ void __launch_bounds__(256)
patchify_mix_kernel(const uint8_t* __restrict__ images, int S, int p, const int* __restrict__ ids, int n_sel,
                    int rows_per_image, int row_off, __nv_bfloat16* __restrict__ out, long ldo, int batch,
                    const int* __restrict__ perm, int mode, float ratio_arg, const uint8_t* __restrict__ xmask,
                    const uint8_t* __restrict__ ymask, const float* __restrict__ ratio_dev) {
  const float ratio = ratio_dev ? *ratio_dev : ratio_arg;  // device scalar: a captured graph replays with new draws
  __shared__ __align__(16) uint8_t spix_all[8][2][2048];
  __shared__ float lut[768];
  fill_norm_lut(lut);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* spa = spix_all[warp][0];
  uint8_t* spb = spix_all[warp][1];
  const int pp = p * p, P = pp * 3, g = S / p;
  const float one_minus_r = __fsub_rn(1.f, ratio);
  const long rows = static_cast<long>(batch) * rows_per_image;
  for (long row = static_cast<long>(blockIdx.x) * 8 + warp; row < rows; row += static_cast<long>(gridDim.x) * 8) {
    const int b = static_cast<int>(row / rows_per_image), r = static_cast<int>(row - static_cast<long>(b) * rows_per_image);
    __nv_bfloat16* dst = out + row * ldo;
    if (r < row_off || r >= row_off + n_sel) {
      for (int e = lane * 4; e < ldo; e += 128) *reinterpret_cast<uint2*>(dst + e) = make_uint2(0u, 0u);
      continue;
    }
    const int patch = ids ? ids[r - row_off] : (r - row_off);
    const int gy = patch / g, gx = patch - gy * g;
    __syncwarp();
    stage_patch(spa, images + static_cast<long>(b) * 3 * S * S, S, p, gy, gx, lane);
    stage_patch(spb, images + static_cast<long>(perm[b]) * 3 * S * S, S, p, gy, gx, lane);
    __syncwarp();
    for (int e0 = lane * 4; e0 < ldo; e0 += 128) {
      float v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int e = e0 + j;
        v[j] = 0.f;
        if (e < P) {
          const int pix = e / 3, c = e - pix * 3;
          const float a = lut[c * 256 + spa[c * pp + pix]], o = lut[c * 256 + spb[c * pp + pix]];
          if (mode == 1) {
            v[j] = __fadd_rn(__fmul_rn(ratio, a), __fmul_rn(one_minus_r, o));
          } else {
            const int ph = pix / p, pw = pix - ph * p;
            v[j] = (xmask[gx * p + pw] & ymask[gy * p + ph]) ? o : a;
          }
        }
      }
      *reinterpret_cast<uint2*>(dst + e0) = make_uint2(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]));
    }
  }
}

__global__ void extract_patches_kernel(const uint8_t* __restrict__ images, int S, int p, float* __restrict__ out) {
  const int patch = blockIdx.x, b = blockIdx.y;
  const int P = p * p * 3, g = S / p;
  const int gy = patch / g, gx = patch - gy * g;
  const uint8_t* img = images + static_cast<long>(b) * 3 * S * S;
  float* dst = out + (static_cast<long>(b) * g * g + patch) * P;
  for (int e = threadIdx.x; e < P; e += blockDim.x) {
    const int c = e % 3, pix = e / 3;
    const int ph = pix / p, pw = pix - ph * p;
    dst[e] = norm_pixel(img[(static_cast<long>(c) * S + gy * p + ph) * S + gx * p + pw], c);
  }
}

// ---------------------------------------------------------------- masked patch MSE
constexpr int kLossThreads = 256;
constexpr int kLossMaxPer = 8;  // P <= 2048

// One WARP per token row, eight rows in flight per block (a block per row leaves one dependent chain of four
// global round trips per 256 threads: latency-bound at a seventh of the HBM rate).  The patch's uint8 pixels are
// staged in the warp's slice of shared memory (3 planes of p x p bytes); every lane then owns groups of 4
// consecutive elements: one float4 read of the prediction, one 8-byte bf16 write of the gradient.
constexpr int kLossWarps = kLossThreads / 32;
static int loss_blocks(int batch, int tokens) {
  const long want = (static_cast<long>(batch) * tokens + kLossWarps - 1) / kLossWarps;
  const long cap = static_cast<long>(sm_count()) * 16;
  return static_cast<int>(want < cap ? want : cap);
}
constexpr int kLossMaxP = 2048;

template <int kLossGroups>  // 4-element groups per lane: P <= 128 * kLossGroups
__global__ void __launch_bounds__(kLossThreads)
patch_loss_kernel(const float* __restrict__ pred, long ldp, const uint8_t* __restrict__ images,
                  const float* __restrict__ mask_row, int S, int p, int J, int norm_pix, float grad_scale,
                  float* __restrict__ partial, __nv_bfloat16* __restrict__ dpred, long lddp, int batch) {
  __shared__ __align__(16) uint8_t spix_all[kLossWarps][kLossMaxP];
  __shared__ float lut[768];
  fill_norm_lut(lut);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* spix = spix_all[warp];
  const int g = S / p, N = g * g, pp = p * p, P = pp * 3, T = J + N;
  float cnt = 0.f;
  for (int i = lane; i < N; i += 32) cnt += mask_row[i] > 0.f ? 1.f : 0.f;
  const float num_masked = warp_sum(cnt);
  const float gscale = 2.0f * grad_scale / (static_cast<float>(P) * num_masked * static_cast<float>(batch));
  const long rows = static_cast<long>(batch) * T;
  for (long row = static_cast<long>(blockIdx.x) * kLossWarps + warp; row < rows;
       row += static_cast<long>(gridDim.x) * kLossWarps) {
    const int b = static_cast<int>(row / T), t = static_cast<int>(row - static_cast<long>(b) * T);
    const int patch = t - J;
    const bool masked = patch >= 0 && mask_row[patch] > 0.f;
    if (!masked) {
      if (patch >= 0 && lane == 0) partial[static_cast<long>(b) * N + patch] = 0.f;
      if (dpred)
        for (int e = lane * 4; e < P; e += 128) *reinterpret_cast<uint2*>(dpred + row * lddp + e) = make_uint2(0u, 0u);
      continue;
    }
    const int gy = patch / g, gx = patch - gy * g;
    const uint8_t* img = images + static_cast<long>(b) * 3 * S * S;
    __syncwarp();  // the previous row's readers are done with the staging slice
    stage_patch(spix, img, S, p, gy, gx, lane);
    __syncwarp();
    // target pixel e of the patch: element (pix = e / 3, channel = e % 3) of the staged planes
    auto target = [&](int e) {
      const int pix = e / 3, c = e - pix * 3;
      return lut[c * 256 + spix[c * pp + pix]];
    };
    float mean = 0.f, inv = 1.f;
    if (norm_pix) {  // src/pretraining.py:115-118, jnp.var is the biased variance: two cheap passes over the staged bytes
      float s = 0.f;
      for (int e = lane; e < P; e += 32) s += target(e);
      mean = warp_sum(s) / P;
      float ss = 0.f;
      for (int e = lane; e < P; e += 32) {
        const float d = target(e) - mean;
        ss += d * d;
      }
      inv = 1.0f / sqrtf(warp_sum(ss) / P + 1e-6f);
    }
    float sq = 0.f;
#pragma unroll 2
    for (int k = 0; k < kLossGroups; ++k) {  // one group in flight at a time: few registers, many resident warps
      const int e0 = (lane + 32 * k) * 4;
      if (e0 < P) {
        const float4 pv = *reinterpret_cast<const float4*>(pred + row * ldp + e0);
        float t0 = target(e0), t1 = target(e0 + 1), t2 = target(e0 + 2), t3 = target(e0 + 3);
        if (norm_pix) t0 = (t0 - mean) * inv, t1 = (t1 - mean) * inv, t2 = (t2 - mean) * inv, t3 = (t3 - mean) * inv;
        const float d0 = pv.x - t0, d1 = pv.y - t1, d2 = pv.z - t2, d3 = pv.w - t3;
        sq += d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3;
        if (dpred)
          *reinterpret_cast<uint2*>(dpred + row * lddp + e0) =
              make_uint2(pack_bf16(d0 * gscale, d1 * gscale), pack_bf16(d2 * gscale, d3 * gscale));
      }
    }
    const float tot = warp_sum(sq);
    if (lane == 0) partial[static_cast<long>(b) * N + patch] = tot / P;
  }
}

static void launch_patch_loss(const float* pred, long ldp, const uint8_t* images, const float* mask_row, int S, int p,
                              int J, int norm_pix, float grad_scale, float* partial, __nv_bfloat16* dpred, long lddp,
                              int batch, int N, int P, cudaStream_t stream) {
  const int grid = loss_blocks(batch, J + N);
#define JM_LOSS_ARGS pred, ldp, images, mask_row, S, p, J, norm_pix, grad_scale, partial, dpred, lddp, batch
  if (P <= 768) patch_loss_kernel<6><<<grid, kLossThreads, 0, stream>>>(JM_LOSS_ARGS);
  else if (P <= 1024) patch_loss_kernel<8><<<grid, kLossThreads, 0, stream>>>(JM_LOSS_ARGS);
  else patch_loss_kernel<16><<<grid, kLossThreads, 0, stream>>>(JM_LOSS_ARGS);
#undef JM_LOSS_ARGS
}

__global__ void patch_loss_finalize_kernel(const float* __restrict__ partial, const float* __restrict__ mask_row,
                                           int N, int batch, float* __restrict__ loss) {
  __shared__ float red[32];
  float cnt = 0.f;
  for (int i = threadIdx.x; i < N; i += blockDim.x) cnt += mask_row[i] > 0.f ? 1.f : 0.f;
  const float num_masked = block_sum(cnt, red);
  float s = 0.f;
  const long total = static_cast<long>(batch) * N;
  for (long i = threadIdx.x; i < total; i += blockDim.x) s += partial[i];
  const float tot = block_sum(s, red);
  if (threadIdx.x == 0) loss[0] = tot / (num_masked * static_cast<float>(batch));
}

// gradient of a learnable position embedding (src/modeling.py:115-118,123): dwpe[ids[i], :] += sum_b dx[b, J + i, :]
// (one block row per kept patch i, threads over 4-column groups, the batch summed in a fixed order)
__global__ void posemb_grad_kernel(const float* __restrict__ dx, long ldx, int batch, int T, int J, int n,
                                   const int* __restrict__ ids, int dim, float* __restrict__ dwpe) {
  const int i = blockIdx.x;
  const int c = (blockIdx.y * blockDim.x + threadIdx.x) * 4;
  if (c >= dim) return;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int b = 0; b < batch; ++b) {
    const float4 v = *reinterpret_cast<const float4*>(dx + (static_cast<long>(b) * T + J + i) * ldx + c);
    acc.x += v.x, acc.y += v.y, acc.z += v.z, acc.w += v.w;
  }
  float4* dst = reinterpret_cast<float4*>(dwpe + static_cast<long>(ids ? ids[i] : i) * dim + c);
  float4 o = *dst;
  o.x += acc.x, o.y += acc.y, o.z += acc.z, o.w += acc.w;
  *dst = o;
}

// the same with one index row per image: every (kept patch, image) adds its row into dwpe[ids[b, i]] (vector atomics)
__global__ void posemb_grad_pi_kernel(const float* __restrict__ dx, long ldx, int T, int J, int n,
                                      const int* __restrict__ ids, long ids_stride, int dim, float* __restrict__ dwpe) {
  const int i = blockIdx.x, b = blockIdx.y;
  const float* src = dx + (static_cast<long>(b) * T + J + i) * ldx;
  float* dst = dwpe + static_cast<long>(ids[b * ids_stride + i]) * dim;
  for (int c = threadIdx.x * 4; c < dim; c += blockDim.x * 4)
    atomicAdd(reinterpret_cast<float4*>(dst + c), *reinterpret_cast<const float4*>(src + c));
}

// x[r, :] *= (r % period < skip ? scale_cls : scale)[r / period]: the transpose of a DropPath-scaled residual branch
// (the gradient that enters the branch is the block-output gradient times the image's keep-mask / (1 - p))
__global__ void scale_rows_bf16_kernel(__nv_bfloat16* __restrict__ x, long ld, long rows, int cols,
                                       const float* __restrict__ scale, const float* __restrict__ scale_cls, int period,
                                       int skip, const float* __restrict__ col_scale,
                                       const float* __restrict__ col_scale_cls) {
  // LayerScale (src/modeling.py:159-166,185-201): besides the per-image factor, a per-column one -- col_scale [cols] on
  // the rows of the patch branch, col_scale_cls [skip * cols] on the CLS rows (scale3 spans the concatenated CLS tokens);
  // any of the four factors may be absent (= 1)
  const int chunks = cols >> 3;
  const long total = rows * chunks;
  for (long i = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const long r = i / chunks;
    const int c = static_cast<int>(i - r * chunks);
    const long img = r / period;
    const int t = static_cast<int>(r - img * period);
    const bool cls = t < skip;
    const float* rs = (cls && scale_cls != nullptr) ? scale_cls : scale;
    const float s_ = rs != nullptr ? rs[img] : 1.0f;
    const float* cs = cls ? (col_scale_cls != nullptr ? col_scale_cls + static_cast<long>(t) * cols : nullptr) : col_scale;
    uint4* p = reinterpret_cast<uint4*>(x + r * ld + c * 8);
    uint4 v = *p;
    uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 f = unpack_bf16(w[k]);
      const float c0 = cs != nullptr ? cs[c * 8 + 2 * k] : 1.0f, c1 = cs != nullptr ? cs[c * 8 + 2 * k + 1] : 1.0f;
      w[k] = pack_bf16(f.x * s_ * c0, f.y * s_ * c1);
    }
    *p = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

// d scale of a LayerScale vector from the accumulated parameter gradients of the Dense it scales: with y = scale * z,
// z = a W + b, the step accumulates dW' = a^T (dy * scale) = (a^T dy) * scale and db' = colsum(dy * scale), and
//   d scale[n] = sum_rows dy[., n] z[., n] = (sum_k W[k, n] dW'[k, n] + b[n] db'[n]) / scale[n]
// -- no activation is kept for it.  (scale[n] == 0 gives no contribution: the formula has no value there.)
__global__ void __launch_bounds__(256)
layerscale_grad_kernel(const float* __restrict__ W, const float* __restrict__ dW, long ld, int K, int N,
                       const float* __restrict__ b, const float* __restrict__ db, const float* __restrict__ gamma,
                       float* __restrict__ dgamma) {
  // block: 32 columns x 8 row lanes
  __shared__ float red[8][33];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int n = blockIdx.x * 32 + cx;
  float acc = 0.f;
  if (n < N)
    for (int k = ry; k < K; k += 8) acc = fmaf(W[static_cast<long>(k) * ld + n], dW[static_cast<long>(k) * ld + n], acc);
  red[ry][cx] = acc;
  __syncthreads();
  if (ry == 0 && n < N) {
    float t = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) t += red[j][cx];
    t = fmaf(b[n], db[n], t);
    const float g = gamma[n];
    if (g != 0.f) dgamma[n] += t / g;
  }
}

// The same for scale3 of a JumboLayer, whose Dense (the Jumbo MLP) is shared by all layers: from the layer's own CLS
// rows, d scale3[n] = sum_b dy[b, n] (c_out[b, n] - c[b, n]) / scale3[n], since c_out - c = drop3 * scale3 * z.
__global__ void __launch_bounds__(256)
layerscale_cls_grad_kernel(const float* __restrict__ dy, long lddy, const float* __restrict__ c_out, long ldo,
                           const float* __restrict__ c, long ldc, int batch, int N, const float* __restrict__ gamma,
                           float* __restrict__ dgamma) {
  __shared__ float red[8][33];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int n = blockIdx.x * 32 + cx;
  float acc = 0.f;
  if (n < N)
    for (int r = ry; r < batch; r += 8)
      acc = fmaf(dy[r * lddy + n], c_out[r * ldo + n] - c[r * ldc + n], acc);
  red[ry][cx] = acc;
  __syncthreads();
  if (ry == 0 && n < N) {
    float t = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) t += red[j][cx];
    const float g = gamma[n];
    if (g != 0.f) dgamma[n] += t / g;
  }
}

// patch_mse_loss with explicit targets (src/utils_mae.py:51-64): one warp per (image, patch) row -> mean_c (t - o)^2
__global__ void patch_mse_rows_kernel(const float* __restrict__ out, long ldo, const float* __restrict__ tgt, long ldt,
                                      long rows, int P, float* __restrict__ rowmse) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long row = static_cast<long>(blockIdx.x) * (blockDim.x >> 5) + warp;
  if (row >= rows) return;
  const float* o = out + row * ldo;
  const float* t = tgt + row * ldt;
  float acc = 0.f;
  for (int c = lane; c < P; c += 32) {
    const float dlt = t[c] - o[c];
    acc = fmaf(dlt, dlt, acc);
  }
  acc = warp_sum(acc);
  if (lane == 0) rowmse[row] = acc / static_cast<float>(P);
}
// ... then per image: mean_L(where(valid > 0, mse, 0)) / (sum(valid) / L), and the batch mean (one block, fixed order)
__global__ void patch_mse_finalize_kernel(const float* __restrict__ rowmse, const float* __restrict__ valid, int L,
                                          int batch, float* __restrict__ loss) {
  __shared__ float red[32];
  float total = 0.f;
  for (int b = 0; b < batch; ++b) {
    float num = 0.f, den = 0.f;
    for (int i = threadIdx.x; i < L; i += blockDim.x) {
      const float v = valid ? valid[static_cast<long>(b) * L + i] : 1.f;
      num += v > 0.f ? rowmse[static_cast<long>(b) * L + i] : 0.f;
      den += v;
    }
    num = block_sum(num, red);
    den = block_sum(den, red);
    total += (num / static_cast<float>(L)) / (den / static_cast<float>(L));
  }
  if (threadIdx.x == 0) loss[0] = total / static_cast<float>(batch);
}

// ---------------------------------------------------------------- LayerNorm
constexpr int kLnWarps = 8;

__global__ void __launch_bounds__(kLnWarps * 32)
layernorm_fwd_kernel(const float* __restrict__ x, long ldx, int rows, int D, const float* __restrict__ gamma,
                     const float* __restrict__ beta, float eps, __nv_bfloat16* __restrict__ yb, long ldy,
                     float* __restrict__ yf, long ldyf, float* __restrict__ mean_out, float* __restrict__ rstd_out,
                     const int* __restrict__ row_map) {
  pdl_wait();  // programmatic dependent launch: nothing above touches global memory
  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long row = static_cast<long>(blockIdx.x) * kLnWarps + warp;
  if (row >= rows) return;
  const float* xr = x + (row_map ? static_cast<long>(row_map[row]) : row) * ldx;  // gather: output rows stay compact
  float s = 0.f, ss = 0.f;
  for (int c = lane * 4; c < D; c += 128) {
    const float4 v = *reinterpret_cast<const float4*>(xr + c);
    s += v.x + v.y + v.z + v.w;
    ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  s = warp_sum(s), ss = warp_sum(ss);
  const float mean = s / D;
  const float var = fmaxf(ss / D - mean * mean, 0.f);  // flax "fast variance"
  const float rstd = 1.0f / sqrtf(var + eps);
  if (lane == 0) {
    if (mean_out) mean_out[row] = mean;
    if (rstd_out) rstd_out[row] = rstd;
  }
  for (int c = lane * 4; c < D; c += 128) {
    const float4 v = *reinterpret_cast<const float4*>(xr + c);
    const float4 g = *reinterpret_cast<const float4*>(gamma + c);
    const float4 b = *reinterpret_cast<const float4*>(beta + c);
    float4 y;
    y.x = (v.x - mean) * rstd * g.x + b.x;
    y.y = (v.y - mean) * rstd * g.y + b.y;
    y.z = (v.z - mean) * rstd * g.z + b.z;
    y.w = (v.w - mean) * rstd * g.w + b.w;
    if (yb) {
      uint2 u = {pack_bf16(y.x, y.y), pack_bf16(y.z, y.w)};
      *reinterpret_cast<uint2*>(yb + row * ldy + c) = u;
    }
    if (yf) *reinterpret_cast<float4*>(yf + row * ldyf + c) = y;
  }
}

constexpr int kLnBwdMaxBlocks = 444;  // 3 x 148 (3 resident CTAs per SM)

// dx = dres + rstd * (dy*g - mean(dy*g) - xhat * mean(dy*g*xhat)); per-block dgamma/dbeta partials
// are accumulated in shared memory and written to the workspace [block][2][D].
__global__ void __launch_bounds__(kLnWarps * 32)
layernorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, long lddy, const float* __restrict__ x, long ldx,
                     const float* __restrict__ mean, const float* __restrict__ rstd, const float* __restrict__ gamma,
                     int rows, int D, const float* __restrict__ dres, long lddres, float* __restrict__ dx, long lddx,
                     __nv_bfloat16* __restrict__ dxb, long lddxb, int row_period, int row_skip,
                     float* __restrict__ ws, int /*want_dcol*/, int /*dcol_period*/, int /*dcol_skip*/) {
  pdl_wait();  // programmatic dependent launch: nothing above touches global memory
  pdl_launch_dependents();
  extern __shared__ float s_acc[];  // [3][D] (the dcol part is not produced by this fallback)
  for (int c = threadIdx.x; c < 3 * D; c += blockDim.x) s_acc[c] = 0.f;
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (long row = static_cast<long>(blockIdx.x) * kLnWarps + warp; row < rows;
       row += static_cast<long>(gridDim.x) * kLnWarps) {
    if (row_period > 0 && (row % row_period) < row_skip) continue;
    const __nv_bfloat16* dyr = dy + row * lddy;
    const float* xr = x + row * ldx;
    const float mu = mean[row], rs = rstd[row];
    float c1 = 0.f, c2 = 0.f;
    for (int c = lane * 4; c < D; c += 128) {
      const uint2 u = *reinterpret_cast<const uint2*>(dyr + c);
      const float2 d01 = unpack_bf16(u.x), d23 = unpack_bf16(u.y);
      const float4 xv = *reinterpret_cast<const float4*>(xr + c);
      const float4 g = *reinterpret_cast<const float4*>(gamma + c);
      const float d[4] = {d01.x, d01.y, d23.x, d23.y};
      const float xh[4] = {(xv.x - mu) * rs, (xv.y - mu) * rs, (xv.z - mu) * rs, (xv.w - mu) * rs};
      const float gg[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        c1 += d[j] * gg[j];
        c2 += d[j] * gg[j] * xh[j];
        atomicAdd(&s_acc[c + j], d[j] * xh[j]);
        atomicAdd(&s_acc[D + c + j], d[j]);
      }
    }
    c1 = warp_sum(c1) / D, c2 = warp_sum(c2) / D;
    for (int c = lane * 4; c < D; c += 128) {
      const uint2 u = *reinterpret_cast<const uint2*>(dyr + c);
      const float2 d01 = unpack_bf16(u.x), d23 = unpack_bf16(u.y);
      const float4 xv = *reinterpret_cast<const float4*>(xr + c);
      const float4 g = *reinterpret_cast<const float4*>(gamma + c);
      float4 o;
      o.x = rs * (d01.x * g.x - c1 - (xv.x - mu) * rs * c2);
      o.y = rs * (d01.y * g.y - c1 - (xv.y - mu) * rs * c2);
      o.z = rs * (d23.x * g.z - c1 - (xv.z - mu) * rs * c2);
      o.w = rs * (d23.y * g.w - c1 - (xv.w - mu) * rs * c2);
      if (dres) {
        const float4 r = *reinterpret_cast<const float4*>(dres + row * lddres + c);
        o.x += r.x, o.y += r.y, o.z += r.z, o.w += r.w;
      }
      if (dx) *reinterpret_cast<float4*>(dx + row * lddx + c) = o;
      if (dxb) {
        uint2 w = {pack_bf16(o.x, o.y), pack_bf16(o.z, o.w)};
        *reinterpret_cast<uint2*>(dxb + row * lddxb + c) = w;
      }
    }
  }
  __syncthreads();
  float* out = ws + static_cast<long>(blockIdx.x) * 3 * D;
  for (int c = threadIdx.x; c < 3 * D; c += blockDim.x) out[c] = s_acc[c];
}

// Fast path for dim <= NV*128.  One warp per row, the row (dy, x-hat) stays in registers between the two
// passes (one global read per operand).  The dgamma / dbeta / column-sum partials live in a per-warp
// private shared-memory accumulator that each lane updates with plain float4 read-modify-writes on the
// columns it owns (no atomics, no conflicts): keeping them in registers costs 3*NV*4 registers per
// thread and leaves one resident CTA per SM, which starves a kernel that is pure HBM streaming.
constexpr int kLnBwdWarps = 4;

// Where a block's parameter-gradient partials go: straight into dgamma / dbeta / dcol with 16-byte vector
// reductions (on != 0), or into the per-block workspace that ln_param_reduce_kernel sums afterwards.
struct LnDirect {
  float* dgamma;
  float* dbeta;
  float* dcol;
  int on;
};

// The residual-gradient stream in bf16: the incoming residual gradient is read from a bf16 tensor (ld in bf16 elements,
// through the `dres` pointer), and the f32 copy of dx is written only for the rows with row % f32_period < f32_skip (the
// CLS rows of a Jumbo block, whose Jumbo-MLP gradient is accumulated in f32); f32_period == 0: every row.
struct LnStream {
  int res_bf16, f32_period, f32_skip;
};

template <int NV>
#ifndef JM_LN_REG_CTAS
#define JM_LN_REG_CTAS 4  // dim 768: 127 registers without spills, four resident CTAs: 81 -> 77 us at the encoder shape
#endif
__global__ void __launch_bounds__(kLnBwdWarps * 32, NV == 6 ? JM_LN_REG_CTAS : 3)
layernorm_bwd_reg_kernel(const __nv_bfloat16* __restrict__ dy, long lddy, const float* __restrict__ x, long ldx,
                         const float* __restrict__ mean, const float* __restrict__ rstd,
                         const float* __restrict__ gamma, int rows, int D, const float* __restrict__ dres,
                         long lddres, float* __restrict__ dx, long lddx, __nv_bfloat16* __restrict__ dxb, long lddxb,
                         int row_period, int row_skip, float* __restrict__ ws, int want_dcol, int dcol_period,
                         int dcol_skip, LnDirect direct, LnStream st) {
  pdl_wait();  // programmatic dependent launch: nothing above touches global memory
  pdl_launch_dependents();
  extern __shared__ __align__(16) float s_acc[];  // [warps][3][D]: dgamma, dbeta, column sums of dx
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* my = s_acc + static_cast<long>(warp) * 3 * D;
  for (int c = lane; c < 3 * D; c += 32) my[c] = 0.f;
  __syncwarp();
  for (long row = static_cast<long>(blockIdx.x) * kLnBwdWarps + warp; row < rows;
       row += static_cast<long>(gridDim.x) * kLnBwdWarps) {
    if (row_period > 0 && (row % row_period) < row_skip) continue;
    const bool col_row = want_dcol && !(dcol_period > 0 && (row % dcol_period) < dcol_skip);
    const __nv_bfloat16* dyr = dy + row * lddy;
    const float* xr = x + row * ldx;
    const float mu = mean[row], rs = rstd[row];
    float d[NV][4], xh[NV][4];
    float4 rres[NV];  // residual gradient, requested together with dy / x so that a row costs ONE memory round trip
    float c1 = 0.f, c2 = 0.f;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = lane * 4 + k * 128;
      rres[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (dres && c < D) {
        if (st.res_bf16) {
          const uint2 u = *reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(dres) + row * lddres + c);
          const float2 a = unpack_bf16(u.x), b = unpack_bf16(u.y);
          rres[k] = make_float4(a.x, a.y, b.x, b.y);
        } else {
          rres[k] = *reinterpret_cast<const float4*>(dres + row * lddres + c);
        }
      }
    }
    const bool f32_row = dx != nullptr && (st.f32_period == 0 || (row % st.f32_period) < st.f32_skip);
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = lane * 4 + k * 128;
      uint2 u = make_uint2(0u, 0u);
      float4 xv = make_float4(mu, mu, mu, mu);
      float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
      if (c < D) {
        u = *reinterpret_cast<const uint2*>(dyr + c);
        xv = *reinterpret_cast<const float4*>(xr + c);
        g = *reinterpret_cast<const float4*>(gamma + c);
      }
      const float2 d01 = unpack_bf16(u.x), d23 = unpack_bf16(u.y);
      d[k][0] = d01.x, d[k][1] = d01.y, d[k][2] = d23.x, d[k][3] = d23.y;
      xh[k][0] = (xv.x - mu) * rs, xh[k][1] = (xv.y - mu) * rs, xh[k][2] = (xv.z - mu) * rs,
      xh[k][3] = (xv.w - mu) * rs;
      const float gg[4] = {g.x, g.y, g.z, g.w};
      if (c < D) {  // parameter-gradient partials: this lane owns these columns of the warp's accumulator
        float4 ag = *reinterpret_cast<float4*>(my + c);
        float4 ab = *reinterpret_cast<float4*>(my + D + c);
        ag.x += d[k][0] * xh[k][0], ag.y += d[k][1] * xh[k][1], ag.z += d[k][2] * xh[k][2], ag.w += d[k][3] * xh[k][3];
        ab.x += d[k][0], ab.y += d[k][1], ab.z += d[k][2], ab.w += d[k][3];
        *reinterpret_cast<float4*>(my + c) = ag;
        *reinterpret_cast<float4*>(my + D + c) = ab;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        d[k][j] *= gg[j];  // from here on d holds dy * gamma
        c1 += d[k][j];
        c2 += d[k][j] * xh[k][j];
      }
    }
    c1 = warp_sum(c1) / D, c2 = warp_sum(c2) / D;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = lane * 4 + k * 128;
      if (c < D) {
        float o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) o[j] = rs * (d[k][j] - c1 - xh[k][j] * c2);
        o[0] += rres[k].x, o[1] += rres[k].y, o[2] += rres[k].z, o[3] += rres[k].w;
        if (f32_row) *reinterpret_cast<float4*>(dx + row * lddx + c) = make_float4(o[0], o[1], o[2], o[3]);
        if (dxb) {
          uint2 w = {pack_bf16(o[0], o[1]), pack_bf16(o[2], o[3])};
          *reinterpret_cast<uint2*>(dxb + row * lddxb + c) = w;
        }
        if (col_row) {
          float4 ac = *reinterpret_cast<float4*>(my + 2 * D + c);
          ac.x += o[0], ac.y += o[1], ac.z += o[2], ac.w += o[3];
          *reinterpret_cast<float4*>(my + 2 * D + c) = ac;
        }
      }
    }
  }
  __syncthreads();
  if (direct.on) {  // one 16-byte reduction per 4 columns and block (D is a multiple of 4, so is every section)
    for (int c = threadIdx.x * 4; c < 3 * D; c += blockDim.x * 4) {
      float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int w = 0; w < kLnBwdWarps; ++w) {
        const float4 a = *reinterpret_cast<const float4*>(s_acc + static_cast<long>(w) * 3 * D + c);
        t.x += a.x, t.y += a.y, t.z += a.z, t.w += a.w;
      }
      float* dst = c < D ? direct.dgamma : (c < 2 * D ? direct.dbeta : direct.dcol);
      const int cc = c < D ? c : (c < 2 * D ? c - D : c - 2 * D);
      if (dst) atomicAdd(reinterpret_cast<float4*>(dst + cc), t);
    }
    return;
  }
  float* out = ws + static_cast<long>(blockIdx.x) * 3 * D;
  for (int c = threadIdx.x; c < 3 * D; c += blockDim.x) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < kLnBwdWarps; ++w) t += s_acc[static_cast<long>(w) * 3 * D + c];
    out[c] = t;
  }
}

// Same, software-pipelined over rows (two rows of loads in flight per warp); used for dim <= 512, where the
// registers allow it: 193 -> 138 us at the decoder shape (93 % of the HBM rate).  Wider rows were measured
// slower with it and keep the kernel above.
template <int NV>
#ifndef JM_LN_PIPE_CTAS
#define JM_LN_PIPE_CTAS 3
#endif
__global__ void __launch_bounds__(kLnBwdWarps * 32, JM_LN_PIPE_CTAS)
layernorm_bwd_pipe_kernel(const __nv_bfloat16* __restrict__ dy, long lddy, const float* __restrict__ x, long ldx,
                         const float* __restrict__ mean, const float* __restrict__ rstd,
                         const float* __restrict__ gamma, int rows, int D, const float* __restrict__ dres,
                         long lddres, float* __restrict__ dx, long lddx, __nv_bfloat16* __restrict__ dxb, long lddxb,
                         int row_period, int row_skip, float* __restrict__ ws, int want_dcol, int dcol_period,
                         int dcol_skip, LnDirect direct, LnStream st, const int* __restrict__ row_map) {
  const int res_bf16 = st.res_bf16;
  pdl_wait();  // programmatic dependent launch: nothing above touches global memory
  pdl_launch_dependents();
  extern __shared__ __align__(16) float s_acc[];  // [warps][3][D]: dgamma, dbeta, column sums of dx
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* my = s_acc + static_cast<long>(warp) * 3 * D;
  for (int c = lane; c < 3 * D; c += 32) my[c] = 0.f;
  __syncwarp();
  // Software pipeline over the warp's rows: the loads of the NEXT row are issued before the arithmetic of the
  // current one, so a warp always has a row of HBM requests in flight (without it the kernel alternates
  // between a load phase and a compute phase and reaches ~60 % of the HBM rate).
  struct Raw {
    float4 res[NV], x[NV];
    uint2 dy[NV];
    float mu, rs;
    long xrow;  // row of x / dx (row_map: dy, mean and rstd are compact, x and dx are scattered)
  };
  const long stride = static_cast<long>(gridDim.x) * kLnBwdWarps;
  auto next_valid = [&](long r) {
    while (r < rows && row_period > 0 && (r % row_period) < row_skip) r += stride;
    return r;
  };
  auto load_row = [&](long r, Raw& w) {
    w.mu = mean[r], w.rs = rstd[r];
    w.xrow = row_map ? static_cast<long>(row_map[r]) : r;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = lane * 4 + k * 128;
      w.res[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      w.x[k] = make_float4(w.mu, w.mu, w.mu, w.mu);
      w.dy[k] = make_uint2(0u, 0u);
      if (c < D) {
        if (dres) {
          if (res_bf16) {  // the residual gradient stream kept in bf16 only (lddres in bf16 elements); the raw
                           // bits travel in .x/.y and are unpacked at the use, so that the load stays in flight
            const uint2 u = *reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(dres) + r * lddres + c);
            w.res[k].x = __uint_as_float(u.x), w.res[k].y = __uint_as_float(u.y);
          } else {
            w.res[k] = *reinterpret_cast<const float4*>(dres + r * lddres + c);
          }
        }
        w.dy[k] = *reinterpret_cast<const uint2*>(dy + r * lddy + c);
        w.x[k] = *reinterpret_cast<const float4*>(x + w.xrow * ldx + c);
      }
    }
  };
  constexpr bool kPipe = true;
  long row = next_valid(static_cast<long>(blockIdx.x) * kLnBwdWarps + warp);
  Raw cur, nxt;
  if (row < rows) load_row(row, cur);
  while (row < rows) {
    const long nrow = next_valid(row + stride);
    if constexpr (kPipe) {
      if (nrow < rows) load_row(nrow, nxt);
    }
    const bool col_row = want_dcol && !(dcol_period > 0 && (row % dcol_period) < dcol_skip);
    const float mu = cur.mu, rs = cur.rs;
    float d[NV][4], xh[NV][4];
    float4 rres[NV];
    float c1 = 0.f, c2 = 0.f;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = lane * 4 + k * 128;
      rres[k] = cur.res[k];
      if (res_bf16) {
        const float2 a = unpack_bf16(__float_as_uint(rres[k].x)), b = unpack_bf16(__float_as_uint(rres[k].y));
        rres[k] = make_float4(a.x, a.y, b.x, b.y);
      }
      const uint2 u = cur.dy[k];
      const float4 xv = cur.x[k];
      float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
      if (c < D) g = *reinterpret_cast<const float4*>(gamma + c);
      const float2 d01 = unpack_bf16(u.x), d23 = unpack_bf16(u.y);
      d[k][0] = d01.x, d[k][1] = d01.y, d[k][2] = d23.x, d[k][3] = d23.y;
      xh[k][0] = (xv.x - mu) * rs, xh[k][1] = (xv.y - mu) * rs, xh[k][2] = (xv.z - mu) * rs,
      xh[k][3] = (xv.w - mu) * rs;
      const float gg[4] = {g.x, g.y, g.z, g.w};
      if (c < D) {  // parameter-gradient partials: this lane owns these columns of the warp's accumulator
        float4 ag = *reinterpret_cast<float4*>(my + c);
        float4 ab = *reinterpret_cast<float4*>(my + D + c);
        ag.x += d[k][0] * xh[k][0], ag.y += d[k][1] * xh[k][1], ag.z += d[k][2] * xh[k][2], ag.w += d[k][3] * xh[k][3];
        ab.x += d[k][0], ab.y += d[k][1], ab.z += d[k][2], ab.w += d[k][3];
        *reinterpret_cast<float4*>(my + c) = ag;
        *reinterpret_cast<float4*>(my + D + c) = ab;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        d[k][j] *= gg[j];  // from here on d holds dy * gamma
        c1 += d[k][j];
        c2 += d[k][j] * xh[k][j];
      }
    }
    c1 = warp_sum(c1) / D, c2 = warp_sum(c2) / D;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = lane * 4 + k * 128;
      if (c < D) {
        float o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) o[j] = rs * (d[k][j] - c1 - xh[k][j] * c2);
        o[0] += rres[k].x, o[1] += rres[k].y, o[2] += rres[k].z, o[3] += rres[k].w;
        if (dx != nullptr && (st.f32_period == 0 || (row % st.f32_period) < st.f32_skip))
          *reinterpret_cast<float4*>(dx + cur.xrow * lddx + c) = make_float4(o[0], o[1], o[2], o[3]);
        if (dxb) {
          uint2 w = {pack_bf16(o[0], o[1]), pack_bf16(o[2], o[3])};
          *reinterpret_cast<uint2*>(dxb + cur.xrow * lddxb + c) = w;
        }
        if (col_row) {
          float4 ac = *reinterpret_cast<float4*>(my + 2 * D + c);
          ac.x += o[0], ac.y += o[1], ac.z += o[2], ac.w += o[3];
          *reinterpret_cast<float4*>(my + 2 * D + c) = ac;
        }
      }
    }
    if constexpr (kPipe) cur = nxt;
    else if (nrow < rows) load_row(nrow, cur);
    row = nrow;
  }
  __syncthreads();
  if (direct.on) {  // one 16-byte reduction per 4 columns and block (D is a multiple of 4, so is every section)
    for (int c = threadIdx.x * 4; c < 3 * D; c += blockDim.x * 4) {
      float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int w = 0; w < kLnBwdWarps; ++w) {
        const float4 a = *reinterpret_cast<const float4*>(s_acc + static_cast<long>(w) * 3 * D + c);
        t.x += a.x, t.y += a.y, t.z += a.z, t.w += a.w;
      }
      float* dst = c < D ? direct.dgamma : (c < 2 * D ? direct.dbeta : direct.dcol);
      const int cc = c < D ? c : (c < 2 * D ? c - D : c - 2 * D);
      if (dst) atomicAdd(reinterpret_cast<float4*>(dst + cc), t);
    }
    return;
  }
  float* out = ws + static_cast<long>(blockIdx.x) * 3 * D;
  for (int c = threadIdx.x; c < 3 * D; c += blockDim.x) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < kLnBwdWarps; ++w) t += s_acc[static_cast<long>(w) * 3 * D + c];
    out[c] = t;
  }
}

// out[c] += sum over blocks of ws[block][c]; 32 columns x 8 block-groups per CTA
__global__ void __launch_bounds__(256)
ln_param_reduce_kernel(const float* __restrict__ ws, int nblocks, int D, float* __restrict__ dgamma,
                       float* __restrict__ dbeta, float* __restrict__ dcol) {
  pdl_wait();  // programmatic dependent launch: nothing above touches global memory
  pdl_launch_dependents();
  __shared__ float red[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  float s = 0.f;
  if (c < 3 * D)
    for (int b = ty; b < nblocks; b += 8) s += ws[static_cast<long>(b) * 3 * D + c];
  red[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && c < 3 * D) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += red[k][tx];
    if (c < D) {
      if (dgamma) dgamma[c] += t;
    } else if (c < 2 * D) {
      if (dbeta) dbeta[c - D] += t;
    } else {
      if (dcol) dcol[c - 2 * D] += t;
    }
  }
}

// ---------------------------------------------------------------- token assembly
__global__ void fill_cls_kernel(float* __restrict__ x, int T, int J, int dim, const float* __restrict__ cls) {
  const int j = blockIdx.x, b = blockIdx.y;
  float* dst = x + (static_cast<long>(b) * T + j) * dim;
  for (int c = threadIdx.x; c < dim; c += blockDim.x) dst[c] = cls[static_cast<long>(j) * dim + c];
}

__global__ void unshuffle_fwd_kernel(const float* __restrict__ y, const int* __restrict__ ids_restore,
                                     const float* __restrict__ mask_emb, const float* __restrict__ wpe, int N, int n,
                                     int J, int dim, float* __restrict__ out, long ids_stride) {
  const int t = blockIdx.x, b = blockIdx.y;
  ids_restore += b * ids_stride;
  float* dst = out + (static_cast<long>(b) * (J + N) + t) * dim;
  const float* src;
  const float* pe = nullptr;
  if (t < J) {
    src = y + (static_cast<long>(b) * (J + n) + t) * dim;
  } else {
    const int i = t - J, r = ids_restore[i];
    src = r < n ? y + (static_cast<long>(b) * (J + n) + J + r) * dim : mask_emb;
    pe = wpe + static_cast<long>(i) * dim;
  }
  for (int c = threadIdx.x * 4; c < dim; c += blockDim.x * 4) {
    float4 v = *reinterpret_cast<const float4*>(src + c);
    if (pe) {
      const float4 q = *reinterpret_cast<const float4*>(pe + c);
      v.x += q.x, v.y += q.y, v.z += q.z, v.w += q.w;
    }
    *reinterpret_cast<float4*>(dst + c) = v;
  }
}

constexpr int kUnshufBwdImgs = 64;  // images per block: the mask-token rows reduce over them before touching L2
__global__ void unshuffle_bwd_kernel(const float* __restrict__ dout, const int* __restrict__ ids_restore, int batch,
                                     int N, int n, int J, int dim, __nv_bfloat16* __restrict__ dy,
                                     float* __restrict__ dmask) {
  const int t = blockIdx.x;
  const int b0 = blockIdx.y * kUnshufBwdImgs;
  const int b1 = min(batch, b0 + kUnshufBwdImgs);
  int dst_tok;  // -1: mask token
  if (t < J) dst_tok = t;
  else {
    const int r = ids_restore[t - J];
    dst_tok = r < n ? J + r : -1;
  }
  for (int c = threadIdx.x * 4; c < dim; c += blockDim.x * 4) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int bb = b0; bb < b1; bb += 8) {  // eight independent loads in flight per thread
      float4 v[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (bb + i < b1)
          v[i] = *reinterpret_cast<const float4*>(dout + (static_cast<long>(bb + i) * (J + N) + t) * dim + c);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (dst_tok >= 0) {
          if (bb + i < b1) {
            uint2 u = {pack_bf16(v[i].x, v[i].y), pack_bf16(v[i].z, v[i].w)};
            *reinterpret_cast<uint2*>(dy + (static_cast<long>(bb + i) * (J + n) + dst_tok) * dim + c) = u;
          }
        } else {
          acc.x += v[i].x, acc.y += v[i].y, acc.z += v[i].z, acc.w += v[i].w;
        }
      }
    }
    if (dst_tok < 0 && dmask) atomicAdd(reinterpret_cast<float4*>(dmask + c), acc);  // one 16-byte reduction
  }
}

// ---------------------------------------------------------------- column sums
template <bool F32>
__global__ void __launch_bounds__(256)
colsum_kernel(const void* __restrict__ xv, long ldx, int rows, int cols, int row_period, int row_skip,
              int row_invert, int rows_per_block, float* __restrict__ out) {
  pdl_wait();  // programmatic dependent launch: nothing above touches global memory
  pdl_launch_dependents();
  __shared__ float4 s_part[4][64];
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
  const int c = (blockIdx.x * 64 + tx) * 4;
  const int r0 = blockIdx.y * rows_per_block;
  const int r1 = min(rows, r0 + rows_per_block);
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (c < cols) {
    for (int r = r0 + ty; r < r1; r += 4) {
      if (row_period > 0 && (((r % row_period) < row_skip) != (row_invert != 0))) continue;
      if (F32) {
        const float4 v = *reinterpret_cast<const float4*>(static_cast<const float*>(xv) + r * ldx + c);
        acc.x += v.x, acc.y += v.y, acc.z += v.z, acc.w += v.w;
      } else {
        const uint2 u = *reinterpret_cast<const uint2*>(static_cast<const __nv_bfloat16*>(xv) + r * ldx + c);
        const float2 a = unpack_bf16(u.x), b = unpack_bf16(u.y);
        acc.x += a.x, acc.y += a.y, acc.z += b.x, acc.w += b.y;
      }
    }
  }
  s_part[ty][tx] = acc;
  __syncthreads();
  if (ty == 0 && c < cols) {
    float4 t = s_part[0][tx];
    for (int k = 1; k < 4; ++k) {
      const float4 o = s_part[k][tx];
      t.x += o.x, t.y += o.y, t.z += o.z, t.w += o.w;
    }
    atomicAdd(out + c + 0, t.x);
    atomicAdd(out + c + 1, t.y);
    atomicAdd(out + c + 2, t.z);
    atomicAdd(out + c + 3, t.w);
  }
}

// ---------------------------------------------------------------- AdamW
__global__ void __launch_bounds__(256)
adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
             __nv_bfloat16* __restrict__ pb, long n4, long n_decay4, float lr, float b1, float b2, float eps,
             float wd, float bc1, float bc2, float gscale) {
  const float omb1 = 1.0f - b1, omb2 = 1.0f - b2;
  for (long i = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    const float4 gv = reinterpret_cast<const float4*>(g)[i];
    float4 mv = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    const float decay = i < n_decay4 ? wd : 0.f;
    float* pp = reinterpret_cast<float*>(&pv);
    const float* gp = reinterpret_cast<const float*>(&gv);
    float* mp = reinterpret_cast<float*>(&mv);
    float* vp = reinterpret_cast<float*>(&vv);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float gg = gp[j] * gscale;
      mp[j] = b1 * mp[j] + omb1 * gg;
      vp[j] = b2 * vp[j] + omb2 * gg * gg;
      const float mh = mp[j] / bc1, vh = vp[j] / bc2;
      const float u = mh / (sqrtf(vh) + eps) + decay * pp[j];
      pp[j] = pp[j] - lr * u;
    }
    reinterpret_cast<float4*>(p)[i] = pv;
    reinterpret_cast<float4*>(m)[i] = mv;
    reinterpret_cast<float4*>(v)[i] = vv;
    if (pb) {
      uint2 u = {pack_bf16(pp[0], pp[1]), pack_bf16(pp[2], pp[3])};
      reinterpret_cast<uint2*>(pb)[i] = u;
    }
  }
}

__global__ void cast_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, long n) {
  for (long i = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long>(gridDim.x) * blockDim.x)
    y[i] = __float2bfloat16(x[i]);
}

}  // namespace jm

using namespace jm;
#define STREAM static_cast<cudaStream_t>(stream)

extern "C" int jm_mask_sort(const float* noise, int length, int keep, int32_t* ids_shuffle, int32_t* ids_restore,
                            float* mask_row, void* stream) {
  JM_CHECK(length > 0 && length <= 1024, "jm_mask_sort: length must be in [1, 1024] (got %d)", length);
  JM_CHECK(keep >= 0 && keep <= length, "jm_mask_sort: keep out of range");
  const int threads = (length + 31) / 32 * 32;
  mask_sort_kernel<<<1, threads, length * sizeof(float), STREAM>>>(noise, length, keep, ids_shuffle, ids_restore,
                                                                   mask_row);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_mask_sort_batched(const float* noise, int batch, int length, int keep, int32_t* ids_shuffle,
                                    int32_t* ids_restore, float* mask, void* stream) {
  JM_CHECK(length > 0 && length <= 1024, "jm_mask_sort_batched: length must be in [1, 1024] (got %d)", length);
  JM_CHECK(keep >= 0 && keep <= length && batch > 0, "jm_mask_sort_batched: keep / batch out of range");
  const int threads = (length + 31) / 32 * 32;
  mask_sort_kernel<<<batch, threads, length * sizeof(float), STREAM>>>(noise, length, keep, ids_shuffle, ids_restore,
                                                                       mask);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_index_sequence(const float* x, const int32_t* ids, int batch, int length_in, int length_out,
                                 int dim, float* out, void* stream) {
  if (batch == 0 || length_out == 0) return 0;
  index_sequence_kernel<<<dim3(length_out, batch), 128, 0, STREAM>>>(x, ids, length_in, length_out, dim, out);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_patchify(const uint8_t* images, int batch, int image_size, int patch_size, const int32_t* ids,
                           int n_sel, int rows_per_image, int row_off, jm_bf16* out, long ldo, void* stream) {
  JM_CHECK(image_size % patch_size == 0, "jm_patchify: image size must be a multiple of the patch size");
  JM_CHECK(ldo >= patch_size * patch_size * 3, "jm_patchify: ldo smaller than the patch dimension");
  JM_CHECK(row_off + n_sel <= rows_per_image, "jm_patchify: rows_per_image too small");
  JM_CHECK(patch_size * patch_size * 3 <= 2048, "jm_patchify: patch dimension %d too large", patch_size * patch_size * 3);
  JM_CHECK(ldo % 4 == 0 && (reinterpret_cast<uintptr_t>(out) & 7) == 0, "jm_patchify: ldo must be a multiple of 4");
  if (batch == 0) return 0;
  const long want = (static_cast<long>(batch) * rows_per_image + 7) / 8, cap = static_cast<long>(sm_count()) * 16;
  patchify_kernel<<<static_cast<int>(want < cap ? want : cap), 256, 0, STREAM>>>(
      images, image_size, patch_size, ids, n_sel, rows_per_image, row_off, reinterpret_cast<__nv_bfloat16*>(out), ldo,
      batch, 0L);
  JM_AFTER_LAUNCH();
  return 0;
}

// Per-image index tables (opt-in extension of random_masking: noise [batch, L], one permutation per image).  The
// entries below are the shared-permutation ones with `ids_stride` = elements between two images' index rows.
extern "C" int jm_patchify_pi(const uint8_t* images, int batch, int image_size, int patch_size, const int32_t* ids,
                              long ids_stride, int n_sel, int rows_per_image, int row_off, jm_bf16* out, long ldo,
                              void* stream) {
  JM_CHECK(images && out && ids, "jm_patchify_pi: null argument");
  JM_CHECK(image_size % patch_size == 0 && ldo >= patch_size * patch_size * 3 && row_off + n_sel <= rows_per_image &&
               patch_size * patch_size * 3 <= 2048 && ldo % 4 == 0 && (reinterpret_cast<uintptr_t>(out) & 7) == 0,
           "jm_patchify_pi: bad shape (see jm_patchify)");
  if (batch == 0) return 0;
  const long want = (static_cast<long>(batch) * rows_per_image + 7) / 8, cap = static_cast<long>(sm_count()) * 16;
  patchify_kernel<<<static_cast<int>(want < cap ? want : cap), 256, 0, STREAM>>>(
      images, image_size, patch_size, ids, n_sel, rows_per_image, row_off, reinterpret_cast<__nv_bfloat16*>(out), ldo,
      batch, ids_stride);
  JM_AFTER_LAUNCH();
  return 0;
}

static int patchify_mix_impl(const uint8_t* images, int batch, int image_size, int patch_size, const int32_t* ids,
                             int n_sel, int rows_per_image, int row_off, jm_bf16* out, long ldo, const int32_t* perm,
                             int mode, float ratio, const float* ratio_dev, const uint8_t* xmask, const uint8_t* ymask,
                             void* stream) {
  JM_CHECK(mode == 1 || mode == 2, "jm_patchify_mix: mode must be 1 (mixup) or 2 (cutmix); use jm_patchify without mixing");
  JM_CHECK(images && out && perm, "jm_patchify_mix: null argument");
  JM_CHECK(mode == 1 || (xmask && ymask), "jm_patchify_mix: cutmix needs xmask [image_size] and ymask [image_size]");
  JM_CHECK(image_size % patch_size == 0, "jm_patchify_mix: image size must be a multiple of the patch size");
  JM_CHECK(ldo >= patch_size * patch_size * 3, "jm_patchify_mix: ldo smaller than the patch dimension");
  JM_CHECK(row_off + n_sel <= rows_per_image, "jm_patchify_mix: rows_per_image too small");
  JM_CHECK(patch_size * patch_size * 3 <= 2048, "jm_patchify_mix: patch dimension %d too large", patch_size * patch_size * 3);
  JM_CHECK(ldo % 4 == 0 && (reinterpret_cast<uintptr_t>(out) & 7) == 0, "jm_patchify_mix: ldo must be a multiple of 4");
  if (batch == 0) return 0;
  const long want = (static_cast<long>(batch) * rows_per_image + 7) / 8, cap = static_cast<long>(sm_count()) * 16;
  patchify_mix_kernel<<<static_cast<int>(want < cap ? want : cap), 256, 0, STREAM>>>(
      images, image_size, patch_size, ids, n_sel, rows_per_image, row_off, reinterpret_cast<__nv_bfloat16*>(out), ldo,
      batch, perm, mode, ratio, xmask, ymask, ratio_dev);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_patchify_mix(const uint8_t* images, int batch, int image_size, int patch_size, const int32_t* ids,
                               int n_sel, int rows_per_image, int row_off, jm_bf16* out, long ldo, const int32_t* perm,
                               int mode, float ratio, const uint8_t* xmask, const uint8_t* ymask, void* stream) {
  return patchify_mix_impl(images, batch, image_size, patch_size, ids, n_sel, rows_per_image, row_off, out, ldo, perm,
                           mode, ratio, nullptr, xmask, ymask, stream);
}

extern "C" int jm_patchify_mix_dev(const uint8_t* images, int batch, int image_size, int patch_size, const int32_t* ids,
                                   int n_sel, int rows_per_image, int row_off, jm_bf16* out, long ldo,
                                   const int32_t* perm, int mode, const float* ratio_dev, const uint8_t* xmask,
                                   const uint8_t* ymask, void* stream) {
  JM_CHECK(ratio_dev != nullptr, "jm_patchify_mix_dev: ratio_dev is null");
  return patchify_mix_impl(images, batch, image_size, patch_size, ids, n_sel, rows_per_image, row_off, out, ldo, perm,
                           mode, 1.f, ratio_dev, xmask, ymask, stream);
}

extern "C" int jm_extract_patches(const uint8_t* images, int batch, int image_size, int patch_size, float* out,
                                  void* stream) {
  JM_CHECK(image_size % patch_size == 0, "jm_extract_patches: image size must be a multiple of the patch size");
  if (batch == 0) return 0;
  const int g = image_size / patch_size;
  extract_patches_kernel<<<dim3(g * g, batch), 256, 0, STREAM>>>(images, image_size, patch_size, out);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" size_t jm_patch_loss_workspace(int batch, int num_patches) {
  return static_cast<size_t>(batch) * num_patches * sizeof(float);
}

extern "C" int jm_patch_loss(const float* pred, long ldp, const uint8_t* images, const float* mask_row, int batch,
                             int image_size, int patch_size, int num_cls, int norm_pix, float grad_scale, float* loss,
                             jm_bf16* dpred, long lddp, void* workspace, size_t workspace_bytes, void* stream) {
  JM_CHECK(patch_size > 0 && image_size > 0 && image_size % patch_size == 0,
           "jm_patch_loss: image size %d must be a positive multiple of the patch size %d", image_size, patch_size);
  const int g = image_size / patch_size, N = g * g, P = patch_size * patch_size * 3;
  JM_CHECK(P <= kLossThreads * kLossMaxPer, "jm_patch_loss: patch dimension %d too large", P);
  JM_CHECK(P % 4 == 0 && ldp % 4 == 0 && (!dpred || lddp % 4 == 0) && (reinterpret_cast<uintptr_t>(pred) & 15) == 0 &&
               (reinterpret_cast<uintptr_t>(dpred) & 7) == 0,
           "jm_patch_loss: P, ldp, lddp must be multiples of 4 and the buffers 16/8-byte aligned");
  JM_CHECK(batch > 0, "jm_patch_loss: empty batch");
  JM_CHECK(workspace && workspace_bytes >= jm_patch_loss_workspace(batch, N), "jm_patch_loss: workspace too small");
  float* partial = static_cast<float*>(workspace);
  launch_patch_loss(pred, ldp, images, mask_row, image_size, patch_size, num_cls, norm_pix, grad_scale, partial,
                    reinterpret_cast<__nv_bfloat16*>(dpred), lddp, batch, N, P, STREAM);
  JM_AFTER_LAUNCH();
  patch_loss_finalize_kernel<<<1, 1024, 0, STREAM>>>(partial, mask_row, N, batch, loss);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_posemb_grad(const float* dx, long ldx, int batch, int T, int J, int n, const int32_t* ids, int dim,
                              float* dwpe, void* stream) {
  JM_CHECK(dx && dwpe && batch > 0 && n > 0 && T >= J + n && dim > 0, "jm_posemb_grad: bad arguments");
  JM_CHECK(dim % 4 == 0 && ldx % 4 == 0 && ((reinterpret_cast<uintptr_t>(dx) | reinterpret_cast<uintptr_t>(dwpe)) & 15) == 0,
           "jm_posemb_grad: dim and ldx must be multiples of 4 and the buffers 16-byte aligned");
  posemb_grad_kernel<<<dim3(n, (dim / 4 + 127) / 128), 128, 0, STREAM>>>(dx, ldx, batch, T, J, n, ids, dim, dwpe);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_posemb_grad_pi(const float* dx, long ldx, int batch, int T, int J, int n, const int32_t* ids,
                                 long ids_stride, int dim, float* dwpe, void* stream) {
  JM_CHECK(dx && dwpe && ids && batch > 0 && n > 0 && T >= J + n && dim > 0, "jm_posemb_grad_pi: bad arguments");
  JM_CHECK(dim % 4 == 0 && ldx % 4 == 0 && ((reinterpret_cast<uintptr_t>(dx) | reinterpret_cast<uintptr_t>(dwpe)) & 15) == 0,
           "jm_posemb_grad_pi: dim and ldx must be multiples of 4 and the buffers 16-byte aligned");
  posemb_grad_pi_kernel<<<dim3(n, batch), 128, 0, STREAM>>>(dx, ldx, T, J, n, ids, ids_stride, dim, dwpe);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_scale_bf16(jm_bf16* x, long ld, long rows, int cols, const float* scale, const float* scale_cls,
                             int period, int skip, const float* col_scale, const float* col_scale_cls, void* stream) {
  JM_CHECK(x && period > 0 && skip >= 0 && skip <= period, "jm_scale_bf16: bad argument");
  JM_CHECK(cols % 8 == 0 && ld % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0,
           "jm_scale_bf16: cols and ld must be multiples of 8 and x 16-byte aligned");
  if (rows == 0) return 0;
  const long total = rows * (cols / 8);
  const int blocks = static_cast<int>(total / 256 + 1 < 148L * 16 ? total / 256 + 1 : 148L * 16);
  scale_rows_bf16_kernel<<<blocks, 256, 0, STREAM>>>(reinterpret_cast<__nv_bfloat16*>(x), ld, rows, cols, scale,
                                                     scale_cls, period, skip, col_scale, col_scale_cls);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_scale_rows_bf16(jm_bf16* x, long ld, long rows, int cols, const float* scale, const float* scale_cls,
                                  int period, int skip, void* stream) {
  JM_CHECK(scale, "jm_scale_rows_bf16: null argument");
  return jm_scale_bf16(x, ld, rows, cols, scale, scale_cls, period, skip, nullptr, nullptr, stream);
}

extern "C" int jm_layerscale_grad(const float* W, const float* dW, long ld, int K, int N, const float* b,
                                  const float* db, const float* scale, float* dscale, void* stream) {
  JM_CHECK(W && dW && b && db && scale && dscale && K > 0 && N > 0, "jm_layerscale_grad: bad argument");
  layerscale_grad_kernel<<<(N + 31) / 32, 256, 0, STREAM>>>(W, dW, ld, K, N, b, db, scale, dscale);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_layerscale_cls_grad(const float* dy, long lddy, const float* c_out, long ldo, const float* c, long ldc,
                                      int batch, int N, const float* scale, float* dscale, void* stream) {
  JM_CHECK(dy && c_out && c && scale && dscale && batch > 0 && N > 0, "jm_layerscale_cls_grad: bad argument");
  layerscale_cls_grad_kernel<<<(N + 31) / 32, 256, 0, STREAM>>>(dy, lddy, c_out, ldo, c, ldc, batch, N, scale, dscale);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_patch_mse_loss(const float* patch_output, long ldo, const float* patch_target, long ldt,
                                 const float* valid, int batch, int num_patches, int patch_dim, float* loss,
                                 void* workspace, size_t workspace_bytes, void* stream) {
  JM_CHECK(batch > 0 && num_patches > 0 && patch_dim > 0, "jm_patch_mse_loss: empty problem");
  JM_CHECK(patch_output && patch_target && loss, "jm_patch_mse_loss: null argument");
  JM_CHECK(workspace && workspace_bytes >= jm_patch_loss_workspace(batch, num_patches),
           "jm_patch_mse_loss: workspace too small (jm_patch_loss_workspace(batch, num_patches) bytes)");
  const long rows = static_cast<long>(batch) * num_patches;
  float* rowmse = static_cast<float*>(workspace);
  patch_mse_rows_kernel<<<static_cast<unsigned>((rows + 7) / 8), 256, 0, STREAM>>>(patch_output, ldo, patch_target, ldt,
                                                                                  rows, patch_dim, rowmse);
  JM_AFTER_LAUNCH();
  patch_mse_finalize_kernel<<<1, 1024, 0, STREAM>>>(rowmse, valid, num_patches, batch, loss);
  JM_AFTER_LAUNCH();
  return 0;
}

// ---------------------------------------------------------------- allocation-free loss entry point
// Scratch for jm_patch_loss_static: one per-patch partial sum per (image, patch).  The array is
// part of the library image, so every GPU that loads libjumbo_mae_b200.so reserves
// JM_PATCH_LOSS_STATIC_MAX * 4 bytes (4 MiB) until the process exits.
namespace jm {
static __device__ __align__(256) float g_patch_loss_static_scratch[JM_PATCH_LOSS_STATIC_MAX];

// In host code the array's name is only a handle for the runtime, not a device address; the
// address is looked up once per device (with that device current) and cached here.
static int patch_loss_static_scratch(float** out) {
  static float* cached[64];
  static std::mutex mu;
  int dev = 0;
  JM_CUDA(cudaGetDevice(&dev));
  JM_CHECK(dev >= 0 && dev < 64, "jm_patch_loss_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (cached[dev] == nullptr) {
    float* p = nullptr;
    JM_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_patch_loss_static_scratch));
    JM_CHECK(p != nullptr, "jm_patch_loss_static: cudaGetSymbolAddress returned a null address");
    cached[dev] = p;
  }
  *out = cached[dev];
  return 0;
}
}  // namespace jm

extern "C" int jm_patch_loss_static(const float* pred, long ldp, const uint8_t* images, const float* mask_row,
                                    int batch, int image_size, int patch_size, int num_cls, int norm_pix,
                                    float grad_scale, float* loss, jm_bf16* dpred, long lddp, void* stream) {
  JM_CHECK(patch_size > 0 && image_size > 0 && image_size % patch_size == 0,
           "jm_patch_loss_static: image size must be a positive multiple of the patch size");
  const int g = image_size / patch_size, N = g * g, P = patch_size * patch_size * 3;
  JM_CHECK(P <= kLossThreads * kLossMaxPer, "jm_patch_loss_static: patch dimension %d too large", P);
  JM_CHECK(P % 4 == 0 && ldp % 4 == 0 && (!dpred || lddp % 4 == 0) && (reinterpret_cast<uintptr_t>(pred) & 15) == 0 &&
               (reinterpret_cast<uintptr_t>(dpred) & 7) == 0,
           "jm_patch_loss_static: P, ldp, lddp must be multiples of 4 and the buffers 16/8-byte aligned");
  JM_CHECK(batch > 0, "jm_patch_loss_static: empty batch");
  const long need = static_cast<long>(batch) * N;
  JM_CHECK(need <= JM_PATCH_LOSS_STATIC_MAX,
           "jm_patch_loss_static: batch * patches = %ld exceeds the built-in scratch of %d entries; "
           "use jm_patch_loss with a caller-owned workspace",
           need, JM_PATCH_LOSS_STATIC_MAX);
  float* partial = nullptr;
  if (patch_loss_static_scratch(&partial)) return 1;
  // Every entry [0, batch * N) is overwritten by the first kernel (zero for visible patches), so the
  // previous call's contents never leak into this one and nothing has to be cleared first.
  launch_patch_loss(pred, ldp, images, mask_row, image_size, patch_size, num_cls, norm_pix, grad_scale, partial,
                    reinterpret_cast<__nv_bfloat16*>(dpred), lddp, batch, N, P, STREAM);
  JM_AFTER_LAUNCH();
  patch_loss_finalize_kernel<<<1, 1024, 0, STREAM>>>(partial, mask_row, N, batch, loss);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_layernorm_fwd(const float* x, long ldx, int rows, int dim, const float* gamma, const float* beta,
                                float eps, jm_bf16* y_bf16, long ldy, float* y_f32, long ldyf, float* mean,
                                float* rstd, void* stream) {
  JM_CHECK(dim % 4 == 0 && ldx % 4 == 0, "jm_layernorm_fwd: dim and ldx must be multiples of 4");
  JM_CHECK(!y_bf16 || ldy % 4 == 0, "jm_layernorm_fwd: ldy must be a multiple of 4");
  JM_CHECK(!y_f32 || ldyf % 4 == 0, "jm_layernorm_fwd: ldyf must be a multiple of 4");
  if (rows == 0) return 0;
  const int grid = (rows + kLnWarps - 1) / kLnWarps;
  JM_LAUNCH_PDL(layernorm_fwd_kernel, grid, kLnWarps * 32, 0, STREAM, x, ldx, rows, dim, gamma, beta, eps,
                                                           reinterpret_cast<__nv_bfloat16*>(y_bf16), ldy, y_f32, ldyf,
                                                           mean, rstd, static_cast<const int*>(nullptr));
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_layernorm_fwd_rows(const float* x, long ldx, const int32_t* row_map, int rows, int dim,
                                     const float* gamma, const float* beta, float eps, jm_bf16* y_bf16, long ldy,
                                     float* mean, float* rstd, void* stream) {
  JM_CHECK(x && row_map && y_bf16 && gamma && beta, "jm_layernorm_fwd_rows: null argument");
  JM_CHECK(dim % 4 == 0 && ldx % 4 == 0 && ldy % 4 == 0, "jm_layernorm_fwd_rows: dim/ld must be multiples of 4");
  if (rows == 0) return 0;
  const int grid = (rows + kLnWarps - 1) / kLnWarps;
  JM_LAUNCH_PDL(layernorm_fwd_kernel, grid, kLnWarps * 32, 0, STREAM, x, ldx, rows, dim, gamma, beta, eps,
                reinterpret_cast<__nv_bfloat16*>(y_bf16), ldy, static_cast<float*>(nullptr), 0L, mean, rstd,
                static_cast<const int*>(row_map));
  JM_AFTER_LAUNCH();
  return 0;
}

static int ln_bwd_blocks(int rows, int dim) {
  const int need = (rows + 3) / 4;  // 4 rows (warps) per CTA in the register kernel, 8 in the fallback
  const int cap = (dim > 512 && dim <= 768) ? kLnBwdMaxBlocks / 3 * JM_LN_REG_CTAS : kLnBwdMaxBlocks;
  return need < cap ? need : cap;
}

extern "C" size_t jm_layernorm_bwd_workspace(int rows, int dim) {
  return static_cast<size_t>(ln_bwd_blocks(rows, dim)) * 3 * dim * sizeof(float);
}

static int layernorm_bwd_impl(const jm_bf16* dy, long lddy, const float* x, long ldx, const float* mean,
                              const float* rstd, const float* gamma, int rows, int dim, const float* dres,
                              long lddres, int res_bf16, float* dx, long lddx, jm_bf16* dx_bf16, long lddxb,
                              int row_period, int row_skip, float* dgamma, float* dbeta, float* dcol, int dcol_period,
                              int dcol_skip, void* workspace, size_t workspace_bytes, void* stream,
                              const int32_t* row_map = nullptr, int f32_period = 0, int f32_skip = 0) {
  JM_CHECK(dim % 4 == 0 && ldx % 4 == 0 && lddy % 4 == 0, "jm_layernorm_bwd: dim/ld must be multiples of 4");
  JM_CHECK(!row_map || dim <= 512, "jm_layernorm_bwd_rows: the scattered variant is built for dim <= 512");
  JM_CHECK(!res_bf16 || (dim <= 1280 && lddres % 4 == 0),
           "jm_layernorm_bwd_lowp: the bf16 residual-gradient input is built for dim <= 1280");
  const LnStream st{res_bf16, f32_period, f32_skip};
  JM_CHECK(dim * 3 * sizeof(float) <= 96 * 1024, "jm_layernorm_bwd: dim %d too large", dim);
  JM_CHECK(dcol == nullptr || dim <= 1280, "jm_layernorm_bwd: dcol is only produced for dim <= 1280");
  if (rows == 0) return 0;
  JM_CHECK(workspace && workspace_bytes >= jm_layernorm_bwd_workspace(rows, dim),
           "jm_layernorm_bwd: workspace too small");
  int grid = ln_bwd_blocks(rows, dim);
  const size_t smem = static_cast<size_t>(3) * dim * sizeof(float);
  if (smem > 48 * 1024) {
    JM_CUDA(cudaFuncSetAttribute(layernorm_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
  }
  const int want_dcol = dcol != nullptr ? 1 : 0;
#define JM_LN_BWD_ARGS                                                                                          \
  reinterpret_cast<const __nv_bfloat16*>(dy), lddy, x, ldx, mean, rstd, gamma, rows, dim, dres, lddres, dx, lddx, \
      reinterpret_cast<__nv_bfloat16*>(dx_bf16), lddxb, row_period, row_skip, static_cast<float*>(workspace), \
      want_dcol, dcol_period, dcol_skip
  const size_t smem_reg = static_cast<size_t>(kLnBwdWarps) * 3 * dim * sizeof(float);
  const int tr = kLnBwdWarps * 32;
  // register kernels: the blocks reduce their partials straight into the gradients (no second launch) when the
  // destinations allow 16-byte vector reductions; JM_LN_DIRECT=0 keeps the workspace + reduce-kernel path
  static const int direct_env = env_int("JM_LN_DIRECT", 1) != 0;  // process-wide, initialised once
  LnDirect direct{dgamma, dbeta, dcol, 0};
  direct.on = direct_env && dim <= 1280 &&
              ((reinterpret_cast<uintptr_t>(dgamma) | reinterpret_cast<uintptr_t>(dbeta) |
                reinterpret_cast<uintptr_t>(dcol)) & 15) == 0;
#define JM_LN_BWD_ARGS_D JM_LN_BWD_ARGS, direct
  if (dim <= 128) JM_LAUNCH_PDL(layernorm_bwd_pipe_kernel<1>, grid, tr, smem_reg, STREAM, JM_LN_BWD_ARGS_D, st, static_cast<const int*>(row_map));
  else if (dim <= 256) JM_LAUNCH_PDL(layernorm_bwd_pipe_kernel<2>, grid, tr, smem_reg, STREAM, JM_LN_BWD_ARGS_D, st, static_cast<const int*>(row_map));
  else if (dim <= 512) JM_LAUNCH_PDL(layernorm_bwd_pipe_kernel<4>, grid, tr, smem_reg, STREAM, JM_LN_BWD_ARGS_D, st, static_cast<const int*>(row_map));
#ifdef JM_LN_PIPE6
  else if (dim <= 768) JM_LAUNCH_PDL(layernorm_bwd_pipe_kernel<6>, grid, tr, smem_reg, STREAM, JM_LN_BWD_ARGS_D, st, static_cast<const int*>(nullptr));
#else
  else if (dim <= 768) JM_LAUNCH_PDL(layernorm_bwd_reg_kernel<6>, grid, tr, smem_reg, STREAM, JM_LN_BWD_ARGS_D, st);
#endif
  else if (dim <= 1024) {
    JM_CUDA(cudaFuncSetAttribute(layernorm_bwd_reg_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    JM_LAUNCH_PDL(layernorm_bwd_reg_kernel<8>, grid, tr, smem_reg, STREAM, JM_LN_BWD_ARGS_D, st);
  } else if (dim <= 1280) {
    JM_CUDA(cudaFuncSetAttribute(layernorm_bwd_reg_kernel<10>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    JM_LAUNCH_PDL(layernorm_bwd_reg_kernel<10>, grid, tr, smem_reg, STREAM, JM_LN_BWD_ARGS_D, st);
  } else {
    const int g8 = (rows + kLnWarps - 1) / kLnWarps < grid ? (rows + kLnWarps - 1) / kLnWarps : grid;
    grid = g8;
    JM_LAUNCH_PDL(layernorm_bwd_kernel, grid, kLnWarps * 32, smem, STREAM, JM_LN_BWD_ARGS);
  }
#undef JM_LN_BWD_ARGS_D
#undef JM_LN_BWD_ARGS
  JM_AFTER_LAUNCH();
  if (!direct.on) {
    JM_LAUNCH_PDL(ln_param_reduce_kernel, (3 * dim + 31) / 32, 256, 0, STREAM, static_cast<const float*>(workspace),
                  grid, dim, dgamma, dbeta, dcol);
    JM_AFTER_LAUNCH();
  }
  return 0;
}

extern "C" int jm_layernorm_bwd(const jm_bf16* dy, long lddy, const float* x, long ldx, const float* mean,
                                const float* rstd, const float* gamma, int rows, int dim, const float* dres,
                                long lddres, float* dx, long lddx, jm_bf16* dx_bf16, long lddxb, int row_period,
                                int row_skip, float* dgamma, float* dbeta, float* dcol, int dcol_period,
                                int dcol_skip, void* workspace, size_t workspace_bytes, void* stream) {
  return layernorm_bwd_impl(dy, lddy, x, ldx, mean, rstd, gamma, rows, dim, dres, lddres, 0, dx, lddx, dx_bf16, lddxb,
                            row_period, row_skip, dgamma, dbeta, dcol, dcol_period, dcol_skip, workspace,
                            workspace_bytes, stream);
}

extern "C" int jm_layernorm_bwd_lowp(const jm_bf16* dy, long lddy, const float* x, long ldx, const float* mean,
                                     const float* rstd, const float* gamma, int rows, int dim, const jm_bf16* dres_bf16,
                                     long lddres, float* dx, long lddx, jm_bf16* dx_bf16, long lddxb, float* dgamma,
                                     float* dbeta, float* dcol, void* workspace, size_t workspace_bytes, void* stream) {
  return layernorm_bwd_impl(dy, lddy, x, ldx, mean, rstd, gamma, rows, dim, reinterpret_cast<const float*>(dres_bf16),
                            lddres, dres_bf16 != nullptr ? 1 : 0, dx, lddx, dx_bf16, lddxb, 0, 0, dgamma, dbeta, dcol,
                            0, 0, workspace, workspace_bytes, stream);
}

extern "C" int jm_layernorm_bwd_stream16(const jm_bf16* dy, long lddy, const float* x, long ldx, const float* mean,
                                         const float* rstd, const float* gamma, int rows, int dim,
                                         const jm_bf16* dres_bf16, long lddres, float* dx, long lddx, int f32_period,
                                         int f32_skip, jm_bf16* dx_bf16, long lddxb, int row_period, int row_skip,
                                         float* dgamma, float* dbeta, float* dcol, int dcol_period, int dcol_skip,
                                         void* workspace, size_t workspace_bytes, void* stream) {
  JM_CHECK(dim <= 1280, "jm_layernorm_bwd_stream16: built for dim <= 1280");
  JM_CHECK(f32_period >= 0 && f32_skip >= 0 && (f32_period == 0 || f32_skip <= f32_period),
           "jm_layernorm_bwd_stream16: bad f32 row selection");
  return layernorm_bwd_impl(dy, lddy, x, ldx, mean, rstd, gamma, rows, dim, reinterpret_cast<const float*>(dres_bf16),
                            lddres, dres_bf16 != nullptr ? 1 : 0, dx, lddx, dx_bf16, lddxb, row_period, row_skip, dgamma,
                            dbeta, dcol, dcol_period, dcol_skip, workspace, workspace_bytes, stream, nullptr, f32_period,
                            f32_skip);
}

extern "C" int jm_layernorm_bwd_rows(const jm_bf16* dy, long lddy, const float* x, long ldx, const int32_t* row_map,
                                     const float* mean, const float* rstd, const float* gamma, int rows, int dim,
                                     long total_rows, float* dx, long lddx, jm_bf16* dx_bf16, long lddxb, float* dgamma,
                                     float* dbeta, float* dcol, void* workspace, size_t workspace_bytes, void* stream) {
  JM_CHECK(row_map && (dx || dx_bf16), "jm_layernorm_bwd_rows: null argument");
  // rows that are not in row_map receive no gradient through this LayerNorm: zero first, then scatter
  if (dx) JM_CUDA(cudaMemset2DAsync(dx, lddx * sizeof(float), 0, dim * sizeof(float), total_rows, STREAM));
  if (dx_bf16)
    JM_CUDA(cudaMemset2DAsync(dx_bf16, lddxb * sizeof(jm_bf16), 0, dim * sizeof(jm_bf16), total_rows, STREAM));
  return layernorm_bwd_impl(dy, lddy, x, ldx, mean, rstd, gamma, rows, dim, nullptr, 0, 0, dx, lddx, dx_bf16, lddxb, 0, 0,
                            dgamma, dbeta, dcol, 0, 0, workspace, workspace_bytes, stream, row_map);
}

extern "C" int jm_masked_targets(const uint8_t* images, int batch, int image_size, int patch_size, const int32_t* ids,
                                 int n_sel, int num_cls, int norm_pix, uint8_t* targets, long ldt, float* stats,
                                 int32_t* row_map, void* stream) {
  JM_CHECK(images && ids && targets, "jm_masked_targets: null argument");
  JM_CHECK(image_size % patch_size == 0, "jm_masked_targets: image size must be a multiple of the patch size");
  const int P = patch_size * patch_size * 3;
  JM_CHECK(P <= 2048 && P % 4 == 0, "jm_masked_targets: patch dimension %d must be a multiple of 4, at most 2048", P);
  JM_CHECK(ldt >= P && ldt % 16 == 0 && (reinterpret_cast<uintptr_t>(targets) & 15) == 0,
           "jm_masked_targets: ldt must be a multiple of 16 bytes, at least the patch dimension");
  JM_CHECK(!norm_pix || stats, "jm_masked_targets: norm_pix needs stats [batch * n_sel, 2]");
  if (batch == 0 || n_sel == 0) return 0;
  const long want = (static_cast<long>(batch) * n_sel + 7) / 8, cap = static_cast<long>(sm_count()) * 16;
  masked_targets_kernel<<<static_cast<int>(want < cap ? want : cap), 256, 0, STREAM>>>(
      images, image_size, patch_size, ids, n_sel, num_cls, batch, norm_pix, targets, ldt, stats, row_map, 0L);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_masked_targets_pi(const uint8_t* images, int batch, int image_size, int patch_size, const int32_t* ids,
                                    long ids_stride, int n_sel, int num_cls, int norm_pix, uint8_t* targets, long ldt,
                                    float* stats, int32_t* row_map, void* stream) {
  JM_CHECK(images && ids && targets, "jm_masked_targets_pi: null argument");
  const int P = patch_size * patch_size * 3;
  JM_CHECK(image_size % patch_size == 0 && P <= 2048 && P % 4 == 0 && ldt >= P && ldt % 16 == 0 &&
               (reinterpret_cast<uintptr_t>(targets) & 15) == 0 && (!norm_pix || stats),
           "jm_masked_targets_pi: bad shape (see jm_masked_targets)");
  if (batch == 0 || n_sel == 0) return 0;
  const long want = (static_cast<long>(batch) * n_sel + 7) / 8, cap = static_cast<long>(sm_count()) * 16;
  masked_targets_kernel<<<static_cast<int>(want < cap ? want : cap), 256, 0, STREAM>>>(
      images, image_size, patch_size, ids, n_sel, num_cls, batch, norm_pix, targets, ldt, stats, row_map, ids_stride);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" size_t jm_row_loss_workspace(void) { return 1024 * sizeof(float); }

extern "C" int jm_row_loss_finalize(const float* row_loss, long count, float scale, float* loss, void* workspace,
                                    size_t workspace_bytes, void* stream) {
  JM_CHECK(row_loss && loss && count > 0, "jm_row_loss_finalize: null or empty argument");
  JM_CHECK(workspace && workspace_bytes >= jm_row_loss_workspace(), "jm_row_loss_finalize: workspace too small");
  const int blocks = static_cast<int>(count / 4096 + 1 < 1024 ? count / 4096 + 1 : 1024);
  row_loss_partial_kernel<<<blocks, 256, 0, STREAM>>>(row_loss, count, static_cast<float*>(workspace));
  JM_AFTER_LAUNCH();
  row_loss_final_kernel<<<1, 256, 0, STREAM>>>(static_cast<const float*>(workspace), blocks, scale, loss);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_fill_cls(float* x, int batch, int T, int J, int dim, const float* cls, void* stream) {
  if (batch == 0 || J == 0) return 0;
  fill_cls_kernel<<<dim3(J, batch), 256, 0, STREAM>>>(x, T, J, dim, cls);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_unshuffle_fwd(const float* y, const int32_t* ids_restore, const float* mask_emb,
                                const float* dec_wpe, int batch, int N, int n, int J, int dim, float* out,
                                void* stream) {
  JM_CHECK(dim % 4 == 0, "jm_unshuffle_fwd: dim must be a multiple of 4");
  if (batch == 0) return 0;
  unshuffle_fwd_kernel<<<dim3(J + N, batch), 128, 0, STREAM>>>(y, ids_restore, mask_emb, dec_wpe, N, n, J, dim, out, 0L);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_unshuffle_fwd_pi(const float* y, const int32_t* ids_restore, long ids_stride, const float* mask_emb,
                                   const float* dec_wpe, int batch, int N, int n, int J, int dim, float* out,
                                   void* stream) {
  JM_CHECK(dim % 4 == 0, "jm_unshuffle_fwd_pi: dim must be a multiple of 4");
  if (batch == 0) return 0;
  unshuffle_fwd_kernel<<<dim3(J + N, batch), 128, 0, STREAM>>>(y, ids_restore, mask_emb, dec_wpe, N, n, J, dim, out,
                                                               ids_stride);
  JM_AFTER_LAUNCH();
  return 0;
}

// per-image permutation: every (token, image) moves on its own; mask-token rows add into dmask with vector atomics
__global__ void unshuffle_bwd_pi_kernel(const float* __restrict__ dout, const int* __restrict__ ids_restore,
                                        long ids_stride, int N, int n, int J, int dim, __nv_bfloat16* __restrict__ dy,
                                        float* __restrict__ dmask) {
  const int t = blockIdx.x, b = blockIdx.y;
  int dst_tok;
  if (t < J) dst_tok = t;
  else {
    const int r = ids_restore[b * ids_stride + t - J];
    dst_tok = r < n ? J + r : -1;
  }
  for (int c = threadIdx.x * 4; c < dim; c += blockDim.x * 4) {
    const float4 v = *reinterpret_cast<const float4*>(dout + (static_cast<long>(b) * (J + N) + t) * dim + c);
    if (dst_tok >= 0)
      *reinterpret_cast<uint2*>(dy + (static_cast<long>(b) * (J + n) + dst_tok) * dim + c) =
          make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
    else if (dmask) atomicAdd(reinterpret_cast<float4*>(dmask + c), v);
  }
}

extern "C" int jm_unshuffle_bwd_pi(const float* dout, const int32_t* ids_restore, long ids_stride, int batch, int N,
                                   int n, int J, int dim, jm_bf16* dy, float* dmask_emb, void* stream) {
  JM_CHECK(dim % 4 == 0 && (reinterpret_cast<uintptr_t>(dmask_emb) & 15) == 0,
           "jm_unshuffle_bwd_pi: dim must be a multiple of 4 and dmask_emb 16-byte aligned");
  if (batch == 0) return 0;
  unshuffle_bwd_pi_kernel<<<dim3(J + N, batch), 128, 0, STREAM>>>(dout, ids_restore, ids_stride, N, n, J, dim,
                                                                  reinterpret_cast<__nv_bfloat16*>(dy), dmask_emb);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_unshuffle_bwd(const float* dout, const int32_t* ids_restore, int batch, int N, int n, int J,
                                int dim, jm_bf16* dy, float* dmask_emb, void* stream) {
  JM_CHECK(dim % 4 == 0, "jm_unshuffle_bwd: dim must be a multiple of 4");
  JM_CHECK((reinterpret_cast<uintptr_t>(dmask_emb) & 15) == 0, "jm_unshuffle_bwd: dmask_emb must be 16-byte aligned");
  if (batch == 0) return 0;
  unshuffle_bwd_kernel<<<dim3(J + N, (batch + kUnshufBwdImgs - 1) / kUnshufBwdImgs), 128, 0, STREAM>>>(
      dout, ids_restore, batch, N, n, J, dim, reinterpret_cast<__nv_bfloat16*>(dy), dmask_emb);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_colsum(const void* x, int is_f32, long ldx, int rows, int cols, int row_period, int row_skip,
                         int row_invert, float* out, void* stream) {
  JM_CHECK(cols % 4 == 0 && ldx % 4 == 0, "jm_colsum: cols and ldx must be multiples of 4");
  if (rows == 0 || cols == 0) return 0;
  const int gx = (cols + 255) / 256;
  int gy = (4 * sm_count() + gx - 1) / gx;
  if (gy < 1) gy = 1;
  int rows_per_block = (rows + gy - 1) / gy;
  if (rows_per_block < 32) rows_per_block = 32;
  gy = (rows + rows_per_block - 1) / rows_per_block;
  if (is_f32)
    JM_LAUNCH_PDL(colsum_kernel<true>, dim3(gx, gy), 256, 0, STREAM, x, ldx, rows, cols, row_period, row_skip, row_invert,
                                                          rows_per_block, out);
  else
    JM_LAUNCH_PDL(colsum_kernel<false>, dim3(gx, gy), 256, 0, STREAM, x, ldx, rows, cols, row_period, row_skip, row_invert,
                                                           rows_per_block, out);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_adamw(float* p, const float* g, float* m, float* v, jm_bf16* p_bf16, long n, long n_decay,
                        float lr, float b1, float b2, float eps, float wd, float bc1, float bc2, float grad_scale,
                        void* stream) {
  JM_CHECK(n % 4 == 0 && n_decay % 4 == 0, "jm_adamw: n and n_decay must be multiples of 4");
  if (n == 0) return 0;
  const long n4 = n / 4;
  long blocks = (n4 + 255) / 256;
  const long cap = static_cast<long>(sm_count()) * 8;
  if (blocks > cap) blocks = cap;
  adamw_kernel<<<static_cast<int>(blocks), 256, 0, STREAM>>>(p, g, m, v, reinterpret_cast<__nv_bfloat16*>(p_bf16), n4,
                                                            n_decay / 4, lr, b1, b2, eps, wd, bc1, bc2, grad_scale);
  JM_AFTER_LAUNCH();
  return 0;
}

extern "C" int jm_cast_bf16(const float* x, jm_bf16* y, long n, void* stream) {
  if (n == 0) return 0;
  long blocks = (n + 255) / 256;
  const long cap = static_cast<long>(sm_count()) * 8;
  if (blocks > cap) blocks = cap;
  cast_bf16_kernel<<<static_cast<int>(blocks), 256, 0, STREAM>>>(x, reinterpret_cast<__nv_bfloat16*>(y), n);
  JM_AFTER_LAUNCH();
  return 0;
}

// dst = src + bias over a strided [rows, cols] f32 view, written as f32 and/or bf16: seeds an output that a
// split-K GEMM then accumulates into, and hands an f32 result to a bf16 consumer.
__global__ void add_bias_cast_kernel(const float* __restrict__ src, long lds, int rows, int cols,
                                     const float* __restrict__ bias, float* __restrict__ dst_f32, long ldd,
                                     __nv_bfloat16* __restrict__ dst_bf16, long ldd16) {
  pdl_wait();  // programmatic dependent launch: nothing above touches global memory
  pdl_launch_dependents();
  const int c4 = cols >> 2;
  const long total = static_cast<long>(rows) * c4;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const int r = static_cast<int>(i / c4), c = static_cast<int>(i - static_cast<long>(r) * c4) * 4;
    float4 v = *reinterpret_cast<const float4*>(src + r * lds + c);
    if (bias) {
      const float4 b = *reinterpret_cast<const float4*>(bias + c);
      v.x += b.x, v.y += b.y, v.z += b.z, v.w += b.w;
    }
    if (dst_f32) *reinterpret_cast<float4*>(dst_f32 + r * ldd + c) = v;
    if (dst_bf16) *reinterpret_cast<uint2*>(dst_bf16 + r * ldd16 + c) = make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
  }
}

extern "C" int jm_add_bias_cast(const float* src, long lds, int rows, int cols, const float* bias, float* dst_f32,
                                long ldd, jm_bf16* dst_bf16, long ldd16, void* stream) {
  JM_CHECK(src && (dst_f32 || dst_bf16), "jm_add_bias_cast: null argument");
  JM_CHECK(rows >= 0 && cols >= 0 && cols % 4 == 0 && lds % 4 == 0, "jm_add_bias_cast: cols and lds must be multiples of 4");
  JM_CHECK((!dst_f32 || ldd % 4 == 0) && (!dst_bf16 || ldd16 % 4 == 0), "jm_add_bias_cast: ldd must be a multiple of 4");
  if (rows == 0 || cols == 0) return 0;
  const long total = static_cast<long>(rows) * (cols / 4);
  long blocks = (total + 255) / 256;
  const long cap = static_cast<long>(sm_count()) * 8;
  if (blocks > cap) blocks = cap;
  JM_LAUNCH_PDL(add_bias_cast_kernel, static_cast<int>(blocks), 256, 0, STREAM, src, lds, rows, cols, bias, dst_f32, ldd,
                                                                     reinterpret_cast<__nv_bfloat16*>(dst_bf16), ldd16);
  JM_AFTER_LAUNCH();
  return 0;
}

// ---------------------------------------------------------------- gradient clipping
// optax.clip_by_global_norm (src/pretraining.py:243-244) over the flat gradient buffer: one pass for the sum
// of squares, one pass that rescales in place if the norm (of pre_scale * g) reaches max_norm.
namespace jm {
__global__ void __launch_bounds__(256)
sumsq_kernel(const float* __restrict__ g, long n4, float* __restrict__ out) {
  __shared__ float red[32];
  float acc = 0.f;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    const float4 v = reinterpret_cast<const float4*>(g)[i];
    acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  const float tot = block_sum(acc, red);
  if (threadIdx.x == 0) atomicAdd(out, tot);
}

__global__ void __launch_bounds__(256)
clip_scale_kernel(float* __restrict__ g, long n4, const float* __restrict__ sumsq, float pre_scale, float max_norm) {
  const float norm = sqrtf(*sumsq) * pre_scale;
  if (norm < max_norm) return;  // optax: unchanged below the threshold
  const float f = max_norm / norm;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    float4 v = reinterpret_cast<float4*>(g)[i];
    v.x *= f, v.y *= f, v.z *= f, v.w *= f;
    reinterpret_cast<float4*>(g)[i] = v;
  }
}
}  // namespace jm

extern "C" int jm_clip_by_global_norm(float* g, long n, float pre_scale, float max_norm, float* sumsq, void* stream) {
  using namespace jm;
  JM_CHECK(g && sumsq, "jm_clip_by_global_norm: null argument");
  JM_CHECK(n % 4 == 0 && (reinterpret_cast<uintptr_t>(g) & 15) == 0, "jm_clip_by_global_norm: n must be a multiple of 4, g 16-byte aligned");
  JM_CHECK(max_norm > 0.f && pre_scale > 0.f, "jm_clip_by_global_norm: max_norm and pre_scale must be positive");
  if (n == 0) return 0;
  const long n4 = n / 4;
  long blocks = (n4 + 255) / 256;
  const long cap = static_cast<long>(sm_count()) * 8;
  if (blocks > cap) blocks = cap;
  JM_CUDA(cudaMemsetAsync(sumsq, 0, sizeof(float), STREAM));
  sumsq_kernel<<<static_cast<int>(blocks), 256, 0, STREAM>>>(g, n4, sumsq);
  JM_AFTER_LAUNCH();
  clip_scale_kernel<<<static_cast<int>(blocks), 256, 0, STREAM>>>(g, n4, sumsq, pre_scale, max_norm);
  JM_AFTER_LAUNCH();
  return 0;
}
