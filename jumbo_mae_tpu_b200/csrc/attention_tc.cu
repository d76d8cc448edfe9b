// Short-sequence attention on the Blackwell tensor path (tcgen05.mma, accumulators in tensor memory, operands
// staged by TMA), forward and backward, for head_dim 32 / 64 and up to 256 tokens: the 49+J / 196+J token counts of
// the MAE encoder / decoder (src/modeling.py:136-137: softmax(q k^T) v per image and head).
//
// One persistent CTA per SM walks over (image, head) problems.  Warp roles:
//   warp 0      TMA producer: Q, K, V (and dO in the backward) of the next problem into a ring of stages.  The
//               operands are boxes of a 3-D tensor map [image][token][column] over the qkv activation, so rows past
//               the sequence end are zero-filled by the TMA unit.
//   warp 1      MMA issuer (one elected thread) and owner of the tensor-memory allocation.
//   warps 4-11  two softmax groups of four warps; warp w reads the TMEM lanes 32 (w % 4) .. +31 (a hardware rule),
//               so a group covers the 128 rows of one accumulator tile, one row per thread.
//
// Forward, per (problem, 128-query tile), two tiles in flight (one per softmax group / TMEM slot):
//   S = Q K^T            tcgen05.mma  M = 128 queries, N = keys (padded to 16), K = head_dim       -> TMEM slot
//   P = exp2(S c - m c)  softmax group: two passes over the row with tcgen05.ld (row max, then exponentials and
//                        row sum), P stored as bf16 into a SWIZZLE_128B shared-memory tile (K-major A operand)
//   O = P V              tcgen05.mma  M = 128, N = head_dim, K = keys; V is the MN-major B operand, read in place
//                        from the [key][head_dim] rows TMA delivered; O reuses the first columns of the S slot
//   out = O / l, lse     softmax group: tcgen05.ld, scale, bf16 rows to global memory
// Scores and probabilities never touch global memory; the tensor core works on one slot while the other group's
// threads run the exponentials (the MUFU unit, 16 ex2 per clock per SM, is the bound of this kernel).
#include <cudaTypedefs.h>

#include "../../include/jumbo_mae_b200.h"
#include "common.cuh"

namespace jm {

namespace {

constexpr float kLog2eTc = 1.4426950408889634f;
constexpr int kTcThreads = 384;
constexpr int kTcSoftmaxWarp0 = 4;
constexpr int kSlotCols = 256;  // TMEM columns per slot: S (<= 256 keys); O aliases its first head_dim columns

struct AttnTcShape {
  int T;        // tokens
  int Nk;       // keys padded to a multiple of 16
  int nq;       // 128-query tiles per problem (1 or 2)
  int heads, batch;
  int stages;       // operand ring depth
  int q_bytes;      // nq * 128 rows
  int kv_bytes;     // Nk rows, rounded up to 1024
  int stage_bytes;  // Q + K + V (+ dO in the backward)
  int p_bytes;      // one P tile: ceil(Nk / 64) blocks of 128 rows x 128 bytes
  long ldo;
};

// Shared-memory descriptor pieces by head_dim: rows of HD bf16 = 64 bytes (SWIZZLE_64B) or 128 bytes (SWIZZLE_128B).
template <int HD>
struct RowCfg {
  static constexpr int ROWB = HD * 2;
  static constexpr uint32_t LT = HD == 32 ? 4u : 2u;  // UMMA layout type
  static constexpr uint32_t SBO = 8 * ROWB;           // 8-row swizzle atom
};

// K-major operand [rows][HD] (Q, K, dO as A or B with the contraction over head_dim): k-step ks covers 16 columns.
template <int HD>
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t base, int ks) {
  return make_smem_desc(base + ks * 32, 0, RowCfg<HD>::SBO, RowCfg<HD>::LT);
}
// MN-major B operand [k rows][HD] (V, K, Q, dO with the contraction over tokens): k-step kk covers 16 rows.
template <int HD>
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t base, int kk) {
  return make_smem_desc(base + kk * 16 * RowCfg<HD>::ROWB, 0, RowCfg<HD>::SBO, RowCfg<HD>::LT);
}
// K-major A operand from a P / dS tile: 64-key blocks of [128 rows][128 bytes], SWIZZLE_128B; k-step kk = 16 keys.
__device__ __forceinline__ uint64_t desc_ptile(uint32_t base, int kk) {
  return make_smem_desc(base + (kk >> 2) * 16384 + (kk & 3) * 32, 0, 1024, 2);
}
// Byte offset of the 16-byte chunk holding keys [8 kc, 8 kc + 8) of row `row` inside a P tile.
__device__ __forceinline__ uint32_t ptile_chunk(int row, int kc) {
  return static_cast<uint32_t>((kc >> 3) * 16384 + row * 128 + (((kc & 7) ^ (row & 7)) << 4));
}

// ------------------------------------------------------------------------------------------------ forward
template <int HD>
__global__ void __launch_bounds__(kTcThreads, 1)
attn_tc_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                   const AttnTcShape sh, __nv_bfloat16* __restrict__ out, float* __restrict__ lse) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sStage = smem;
  uint8_t* sP = smem + sh.stages * sh.stage_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 2 * sh.p_bytes);
  uint64_t* kv_full = bars;            // [stages]
  uint64_t* kv_empty = bars + 4;       // [stages]
  uint64_t* s_full = bars + 8;         // [2]
  uint64_t* p_full = bars + 10;        // [2]
  uint64_t* o_full = bars + 12;        // [2]
  uint64_t* o_empty = bars + 14;       // [2]
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 16);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmKV);
    for (int s = 0; s < sh.stages; ++s) {
      mbar_init(&kv_full[s], 1);
      mbar_init(&kv_empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&s_full[s], 1);
      mbar_init(&p_full[s], 128);
      mbar_init(&o_full[s], 1);
      mbar_init(&o_empty[s], 128);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<512>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_wait();
  pdl_launch_dependents();

  const int nprob = sh.batch * sh.heads;
  const int my_prob = (nprob - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const int n_items = my_prob * sh.nq;
  const int d = sh.heads * HD;

  if (warp == 0) {
    if (elect_one()) {
      for (int j = 0; j < my_prob; ++j) {
        const int prob = blockIdx.x + j * gridDim.x;
        const int b = prob / sh.heads, h = prob - b * sh.heads;
        const int st = j % sh.stages;
        mbar_wait(&kv_empty[st], ((j / sh.stages) & 1) ^ 1);
        uint8_t* dst = sStage + st * sh.stage_bytes;
        mbar_arrive_expect_tx(&kv_full[st], sh.nq * 128 * HD * 2 + 2 * sh.Nk * HD * 2);
        for (int qt = 0; qt < sh.nq; ++qt)
          tma_load_3d(&tmQ, &kv_full[st], dst + qt * 128 * HD * 2, h * HD, qt * 128, b);
        tma_load_3d(&tmKV, &kv_full[st], dst + sh.q_bytes, d + h * HD, 0, b);
        tma_load_3d(&tmKV, &kv_full[st], dst + sh.q_bytes + sh.kv_bytes, 2 * d + h * HD, 0, b);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      const uint32_t idesc_s = make_idesc_bf16(128, sh.Nk, 0, 0);
      const uint32_t idesc_o = make_idesc_bf16(128, HD, 0, 1);
      auto issue_pv = [&](int i) {
        const int slot = i & 1, j = i / sh.nq, st = j % sh.stages;
        mbar_wait(&p_full[slot], (i >> 1) & 1);
        tc_fence_after();
        const uint32_t p_addr = smem_u32(sP + slot * sh.p_bytes);
        const uint32_t v_addr = smem_u32(sStage + st * sh.stage_bytes + sh.q_bytes + sh.kv_bytes);
        const int ksteps = sh.Nk >> 4;
        for (int kk = 0; kk < ksteps; ++kk)
          umma_bf16(tmem_base + slot * kSlotCols, desc_ptile(p_addr, kk), desc_mnmajor<HD>(v_addr, kk), idesc_o,
                    kk > 0 ? 1u : 0u);
        umma_commit(&o_full[slot]);
        if (i % sh.nq == sh.nq - 1) umma_commit(&kv_empty[st]);  // every MMA that reads this stage has been issued
      };
      for (int i = 0; i < n_items; ++i) {
        const int slot = i & 1, j = i / sh.nq, qt = i - j * sh.nq, st = j % sh.stages;
        // With a single operand stage the previous problem's last P V must be issued before this problem's
        // operands can arrive (its commit is what frees the stage); otherwise P V lags one item behind S.
        const bool pv_first = sh.stages == 1 && qt == 0 && i >= 1;
        if (pv_first) issue_pv(i - 1);
        if (qt == 0) {
          mbar_wait(&kv_full[st], (j / sh.stages) & 1);
          tc_fence_after();
        }
        mbar_wait(&o_empty[slot], ((i >> 1) & 1) ^ 1);  // the slot's previous O has been read out
        tc_fence_after();
        const uint32_t q_addr = smem_u32(sStage + st * sh.stage_bytes + qt * 128 * HD * 2);
        const uint32_t k_addr = smem_u32(sStage + st * sh.stage_bytes + sh.q_bytes);
#pragma unroll
        for (int ks = 0; ks < HD / 16; ++ks)
          umma_bf16(tmem_base + slot * kSlotCols, desc_kmajor<HD>(q_addr, ks), desc_kmajor<HD>(k_addr, ks), idesc_s,
                    ks > 0 ? 1u : 0u);
        umma_commit(&s_full[slot]);
        if (i >= 1 && !pv_first) issue_pv(i - 1);
      }
      if (n_items >= 1) issue_pv(n_items - 1);
    }
  } else if (warp >= kTcSoftmaxWarp0) {
    const int grp = (warp - kTcSoftmaxWarp0) >> 2;  // softmax group == TMEM slot
    const int q = warp & 3;                         // TMEM lane quarter
    const int row = q * 32 + lane;                  // row of the 128-query tile
    const uint32_t t_slot = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + grp * kSlotCols;
    const uint32_t p_base = smem_u32(sP + grp * sh.p_bytes);
    const int nchunks = (sh.Nk + 31) >> 5;
    for (int i = grp; i < n_items; i += 2) {
      const int j = i / sh.nq, qt = i - j * sh.nq;
      const int prob = blockIdx.x + j * gridDim.x;
      const int b = prob / sh.heads, h = prob - b * sh.heads;
      const int tok = qt * 128 + row;
      const bool warp_live = qt * 128 + q * 32 < sh.T;  // warp-uniform: any valid row in this warp's 32
      const uint32_t ph = (i >> 1) & 1;
      mbar_wait(&s_full[grp], ph);
      tc_fence_after();
      float l = 0.f, mx = -INFINITY;
      if (warp_live) {
        // pass 1: row maximum (columns >= T hold q . 0 = 0 and must not take part)
        for (int c = 0; c < nchunks; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(t_slot + c * 32, r);
          tmem_ld_wait();
          if (c * 32 + 32 > sh.T) {
#pragma unroll
            for (int e = 0; e < 32; ++e)
              if (c * 32 + e >= sh.T) r[e] = 0xff800000u;  // -inf
          }
#pragma unroll
          for (int e = 0; e < 32; e += 2) mx = fmax3(mx, __uint_as_float(r[e]), __uint_as_float(r[e + 1]));
        }
        // pass 2: P = exp2(s log2e - max log2e), row sum, bf16 P into the swizzled shared-memory tile
        const float nm = -mx * kLog2eTc;
        const f32x2 k2 = pack2(kLog2eTc, kLog2eTc), nm2 = pack2(nm, nm);
        f32x2 l2 = pack2(0.f, 0.f);
        for (int c = 0; c < nchunks; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(t_slot + c * 32, r);
          tmem_ld_wait();
          if (c * 32 + 32 > sh.T) {
#pragma unroll
            for (int e = 0; e < 32; ++e)
              if (c * 32 + e >= sh.T) r[e] = 0xff800000u;
          }
          uint32_t w[16];
#pragma unroll
          for (int e = 0; e < 32; e += 2) {
            float x0, x1;
            unpack2(fma2(pack2(__uint_as_float(r[e]), __uint_as_float(r[e + 1])), k2, nm2), x0, x1);
            const float p0 = ex2_approx(x0), p1 = ex2_approx(x1);
            l2 = add2(l2, pack2(p0, p1));
            w[e >> 1] = pack_bf16(p0, p1);
          }
#pragma unroll
          for (int jj = 0; jj < 4; ++jj)
            sts128(p_base + ptile_chunk(row, c * 4 + jj), w[4 * jj], w[4 * jj + 1], w[4 * jj + 2], w[4 * jj + 3]);
        }
        float la, lb;
        unpack2(l2, la, lb);
        l = la + lb;
      }
      // P is complete and S has been consumed: the tensor core may overwrite the slot with O = P V
      tc_fence_before();
      fence_proxy_async();
      mbar_arrive(&p_full[grp]);
      mbar_wait(&o_full[grp], ph);
      tc_fence_after();
      if (warp_live) {
        const float inv = 1.0f / l;
        uint32_t r[32];
#pragma unroll
        for (int c = 0; c < HD / 32; ++c) {
          tmem_ld_32x32(t_slot + c * 32, r);
          tmem_ld_wait();
          if (tok < sh.T) {
            __nv_bfloat16* dst = out + (static_cast<long>(b) * sh.T + tok) * sh.ldo + h * HD + c * 32;
#pragma unroll
            for (int e = 0; e < 32; e += 8) {
              uint4 v;
              v.x = pack_bf16(__uint_as_float(r[e]) * inv, __uint_as_float(r[e + 1]) * inv);
              v.y = pack_bf16(__uint_as_float(r[e + 2]) * inv, __uint_as_float(r[e + 3]) * inv);
              v.z = pack_bf16(__uint_as_float(r[e + 4]) * inv, __uint_as_float(r[e + 5]) * inv);
              v.w = pack_bf16(__uint_as_float(r[e + 6]) * inv, __uint_as_float(r[e + 7]) * inv);
              *reinterpret_cast<uint4*>(dst + e) = v;
            }
          }
        }
        if (tok < sh.T) lse[(static_cast<long>(b) * sh.heads + h) * sh.T + tok] = mx + logf(l);
      }
      tc_fence_before();
      mbar_arrive(&o_empty[grp]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

// ------------------------------------------------------------------------------------------------ host side
constexpr int kMaxDynSmem = 227 * 1024;

// 3-D bf16 tensor map [batch][T][cols] over a [batch * T, ld] activation; box = [1][box_rows][hd].
int make_tmap_heads(CUtensorMap* map, const void* ptr, long cols, int T, int batch, long ld, int hd, int box_rows) {
  auto fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(tmap_encode_fn());
  JM_CHECK(fn != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  JM_CHECK((reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && ld % 8 == 0,
           "attention operands must be 16-byte aligned with row strides that are multiples of 8 elements");
  cuuint64_t dims[3] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(T), static_cast<cuuint64_t>(batch)};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(ld) * 2, static_cast<cuuint64_t>(ld) * 2 * T};
  cuuint32_t box[3] = {static_cast<cuuint32_t>(hd), static_cast<cuuint32_t>(box_rows), 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, hd == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  JM_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (attention) failed with CUresult %d (T=%d batch=%d ld=%ld hd=%d)",
           static_cast<int>(r), T, batch, ld, hd);
  return 0;
}

int fill_shape(AttnTcShape* sh, int batch, int T, int heads, int hd, int tiles_per_stage, int ptiles) {
  sh->T = T, sh->Nk = (T + 15) / 16 * 16, sh->nq = (T + 127) / 128;
  sh->heads = heads, sh->batch = batch;
  sh->q_bytes = sh->nq * 128 * hd * 2;
  sh->kv_bytes = (sh->Nk * hd * 2 + 1023) / 1024 * 1024;
  sh->p_bytes = (sh->Nk + 63) / 64 * 16384;
  (void)tiles_per_stage;
  (void)ptiles;
  return 0;
}

template <int HD>
int launch_tc_fwd(const jm_bf16* qkv, long ldqkv, int batch, int T, int heads, jm_bf16* out, long ldo, float* lse,
                  cudaStream_t stream) {
  AttnTcShape sh;
  fill_shape(&sh, batch, T, heads, HD, 0, 0);
  sh.stage_bytes = sh.q_bytes + 2 * sh.kv_bytes;
  sh.ldo = ldo;
  const int fixed = 1024 + 2 * sh.p_bytes + 256;
  sh.stages = (kMaxDynSmem - fixed) / sh.stage_bytes;
  if (sh.stages > 3) sh.stages = 3;
  JM_CHECK(sh.stages >= 1, "jm_attention_fwd: T=%d hd=%d does not fit shared memory", T, HD);
  const int smem = fixed + sh.stages * sh.stage_bytes;
  CUtensorMap tmQ, tmKV;
  const long d = static_cast<long>(heads) * HD;
  if (make_tmap_heads(&tmQ, qkv, 3 * d, T, batch, ldqkv, HD, 128)) return 1;
  if (make_tmap_heads(&tmKV, qkv, 3 * d, T, batch, ldqkv, HD, sh.Nk)) return 1;
  auto kern = attn_tc_fwd_kernel<HD>;
  static DeviceOnce attr_set;
  const int dev = current_device();
  JM_CHECK(dev >= 0, "jm_attention_fwd: no current CUDA device");
  if (!attr_set.done(dev)) {
    JM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem));
    attr_set.mark(dev);
  }
  const int nprob = batch * heads, sms = sm_count();
  const int grid = nprob < sms ? nprob : sms;
  JM_LAUNCH_PDL(kern, grid, kTcThreads, smem, stream, tmQ, tmKV, sh, reinterpret_cast<__nv_bfloat16*>(out), lse);
  JM_AFTER_LAUNCH();
  return 0;
}

}  // namespace

bool attn_tc_supported(int T, int hd) {
  static const int on = env_int("JM_ATTN_TC", 1);
  return on != 0 && (hd == 32 || hd == 64) && T >= 1 && T <= 256;
}

int attn_tc_fwd(const jm_bf16* qkv, long ldqkv, int batch, int T, int heads, int hd, jm_bf16* out, long ldo,
                float* lse, cudaStream_t stream) {
  JM_CHECK((reinterpret_cast<uintptr_t>(out) & 15) == 0 && ldo % 8 == 0 && ldqkv >= 3L * heads * hd,
           "jm_attention_fwd: out must be 16-byte aligned, ldo a multiple of 8, ldqkv >= 3 * heads * head_dim");
  if (hd == 32) return launch_tc_fwd<32>(qkv, ldqkv, batch, T, heads, out, ldo, lse, stream);
  return launch_tc_fwd<64>(qkv, ldqkv, batch, T, heads, out, ldo, lse, stream);
}

}  // namespace jm
