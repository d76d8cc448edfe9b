// LAMB as the reference builds it (`modified_lamb`, src/utils.py:124-139; --optimizer lamb, src/pretraining.py:46,
// src/finetuning.py:45):
//   optax.chain(scale_by_adam, add_decayed_weights(mask), masked(scale_by_trust_ratio(), mask), scale_by_learning_rate)
// with mask = "the leaf is called kernel".  For a kernel leaf
//   m = b1 m + (1-b1) g;  v = b2 v + (1-b2) g^2;  u = (m / bc1) / (sqrt(v / bc2) + eps) + wd p;
//   p -= lr * ratio * u,  ratio = ||p|| / ||u||  (1 where either norm is 0),
// the norms taken over the WHOLE LEAF of the Flax tree.  The flat parameter buffer fuses wq / wk / wv into one matrix, so
// a "group" here is a column range of a stored matrix.  Every other leaf is a plain Adam step (jm_adamw with wd = 0).
//
// Two passes over the kernel leaves, both HBM-bound (pass 1: read p g m v, write m v = 24 B / element; pass 2: read p m v,
// write p + bf16 copy = 18 B / element).  The norms are summed in a fixed order (one partial pair per chunk, written by
// one block; pass 2 adds the partials of its group in index order), so that data-parallel ranks, which see identical
// gradients after the all-reduce, stay bit-identical.
#include "../../include/jumbo_mae_b200.h"
#include "common.cuh"

namespace jm {

struct LambGroup {  // a column range of a stored [rows, ld] matrix = one `kernel` leaf of the Flax tree
  long offset;      // of the matrix in the flat buffers
  long rows, ld, col0, ncols;
  long chunk0, nchunks;  // its chunks (consecutive)
};
struct LambChunk {
  long group, row0, nrows;
};

__device__ __forceinline__ float lamb_block_sum(float v, float* red) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = lane < (blockDim.x >> 5) ? red[lane] : 0.f;
  return warp_sum(t);
}

// pass 1: moments, and this chunk's partial sums of p^2 and u^2
__global__ void __launch_bounds__(256)
lamb_moments_kernel(const float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                    float* __restrict__ v, const LambGroup* __restrict__ groups, const LambChunk* __restrict__ chunks,
                    float b1, float b2, float eps, float wd, float bc1, float bc2, float gscale,
                    float* __restrict__ partial) {
  __shared__ float red[32];
  const LambChunk ch = chunks[blockIdx.x];
  const LambGroup gr = groups[ch.group];
  const float omb1 = 1.0f - b1, omb2 = 1.0f - b2;
  float sp = 0.f, su = 0.f;
  const long total = ch.nrows * gr.ncols;
  for (long i = threadIdx.x; i < total; i += blockDim.x) {
    const long r = i / gr.ncols, c = i - r * gr.ncols;
    const long idx = gr.offset + (ch.row0 + r) * gr.ld + gr.col0 + c;
    const float gg = g[idx] * gscale;
    const float mm = b1 * m[idx] + omb1 * gg;
    const float vv = b2 * v[idx] + omb2 * gg * gg;
    m[idx] = mm, v[idx] = vv;
    const float pp = p[idx];
    const float u = (mm / bc1) / (sqrtf(vv / bc2) + eps) + wd * pp;
    sp = fmaf(pp, pp, sp), su = fmaf(u, u, su);
  }
  sp = lamb_block_sum(sp, red);
  su = lamb_block_sum(su, red);
  if (threadIdx.x == 0) partial[2 * blockIdx.x] = sp, partial[2 * blockIdx.x + 1] = su;
}

// pass 2: trust ratio of the chunk's group from the partials (fixed order), then the step
__global__ void __launch_bounds__(256)
lamb_apply_kernel(float* __restrict__ p, const float* __restrict__ m, const float* __restrict__ v,
                  __nv_bfloat16* __restrict__ pb, const LambGroup* __restrict__ groups,
                  const LambChunk* __restrict__ chunks, float lr, float eps, float wd, float bc1, float bc2,
                  const float* __restrict__ partial) {
  __shared__ float red[32];
  __shared__ float s_ratio;
  const LambChunk ch = chunks[blockIdx.x];
  const LambGroup gr = groups[ch.group];
  float sp = 0.f, su = 0.f;
  for (long k = threadIdx.x; k < gr.nchunks; k += blockDim.x)
    sp += partial[2 * (gr.chunk0 + k)], su += partial[2 * (gr.chunk0 + k) + 1];
  sp = lamb_block_sum(sp, red);
  su = lamb_block_sum(su, red);
  if (threadIdx.x == 0) {
    const float pn = sqrtf(sp), un = sqrtf(su);
    s_ratio = (pn == 0.f || un == 0.f) ? 1.0f : pn / un;
  }
  __syncthreads();
  const float step = lr * s_ratio;
  const long total = ch.nrows * gr.ncols;
  for (long i = threadIdx.x; i < total; i += blockDim.x) {
    const long r = i / gr.ncols, c = i - r * gr.ncols;
    const long idx = gr.offset + (ch.row0 + r) * gr.ld + gr.col0 + c;
    const float pp = p[idx];
    const float u = (m[idx] / bc1) / (sqrtf(v[idx] / bc2) + eps) + wd * pp;
    const float np = pp - step * u;
    p[idx] = np;
    if (pb) pb[idx] = __float2bfloat16(np);
  }
}

}  // namespace jm

extern "C" size_t jm_lamb_workspace(int n_chunks) { return static_cast<size_t>(n_chunks) * 2 * sizeof(float); }

extern "C" int jm_lamb(float* p, const float* g, float* m, float* v, jm_bf16* p_bf16, const int64_t* groups,
                       int n_groups, const int64_t* chunks, int n_chunks, float lr, float b1, float b2, float eps,
                       float wd, float bc1, float bc2, float grad_scale, void* workspace, size_t workspace_bytes,
                       void* stream) {
  using namespace jm;
  static_assert(sizeof(LambGroup) == 7 * sizeof(int64_t) && sizeof(LambChunk) == 3 * sizeof(int64_t), "table layout");
  JM_CHECK(p && g && m && v && groups && chunks, "jm_lamb: null argument");
  JM_CHECK(n_groups > 0 && n_chunks > 0, "jm_lamb: empty group / chunk table");
  JM_CHECK(workspace && workspace_bytes >= jm_lamb_workspace(n_chunks), "jm_lamb: workspace too small");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* partial = static_cast<float*>(workspace);
  lamb_moments_kernel<<<n_chunks, 256, 0, st>>>(p, g, m, v, reinterpret_cast<const LambGroup*>(groups),
                                                reinterpret_cast<const LambChunk*>(chunks), b1, b2, eps, wd, bc1, bc2,
                                                grad_scale, partial);
  JM_AFTER_LAUNCH();
  lamb_apply_kernel<<<n_chunks, 256, 0, st>>>(p, m, v, reinterpret_cast<__nv_bfloat16*>(p_bf16),
                                              reinterpret_cast<const LambGroup*>(groups),
                                              reinterpret_cast<const LambChunk*>(chunks), lr, eps, wd, bc1, bc2, partial);
  JM_AFTER_LAUNCH();
  return 0;
}
