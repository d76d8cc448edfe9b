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
//                        row sum), P written back over S as packed bf16 pairs with tcgen05.st
//   O = P V              tcgen05.mma  M = 128, N = head_dim, K = keys; A = P read from tensor memory, V is the
//                        MN-major B operand, read in place from the [key][head_dim] rows TMA delivered; O goes to
//                        the columns of the slot behind P
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
  int col_o;        // TMEM column (inside a slot) of O: behind the bf16 P that overwrites S in place
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
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + sh.stages * sh.stage_bytes);
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
        const uint32_t v_addr = smem_u32(sStage + st * sh.stage_bytes + sh.q_bytes + sh.kv_bytes);
        const int ksteps = sh.Nk >> 4;
        const uint32_t t_p = tmem_base + slot * kSlotCols;  // P: bf16 pairs, 8 columns per 16-key step
        for (int kk = 0; kk < ksteps; ++kk)
          umma_bf16_ts(t_p + sh.col_o, t_p + kk * 8, desc_mnmajor<HD>(v_addr, kk), idesc_o, kk > 0 ? 1u : 0u);
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
          // P overwrites S in place: the 32 f32 columns just read become 16 columns of bf16 pairs at [16 c, 16 c + 16),
          // which lie inside columns this thread has already consumed
          tmem_st_32x16(t_slot + c * 16, w);
        }
        float la, lb;
        unpack2(l2, la, lb);
        l = la + lb;
        tmem_st_wait();
      }
      // P is complete (tensor memory, the A operand of O = P V) and S has been consumed
      tc_fence_before();
      mbar_arrive(&p_full[grp]);
      mbar_wait(&o_full[grp], ph);
      tc_fence_after();
      if (warp_live) {
        const float inv = 1.0f / l;
        uint32_t r[32];
#pragma unroll
        for (int c = 0; c < HD / 32; ++c) {
          tmem_ld_32x32(t_slot + sh.col_o + c * 32, r);
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

// ------------------------------------------------------------------------------------------------ backward
// Per (image, head) and 128-key tile kt, with TMEM lanes = keys (thread = one key row) and columns = queries:
//   S^T  = K_kt Q^T      tcgen05.mma  M = 128 keys, N = queries (padded to 16), K = head_dim   -> columns [0, Nq)
//   dP^T = V_kt dO^T     same shape                                                            -> columns [cP, cP + Nq)
//   P^T  = exp2(S^T c - lse c),  dS^T = P^T (dP^T - D)        both softmax groups (each half of the query columns);
//          lse and D = rowsum(dO . O) per query come from shared memory (broadcast reads); bf16 P^T and dS^T go to
//          two SWIZZLE_128B tiles [key][query]
//   dV_kt = P^T dO       A = P^T tile (K-major), B = dO rows in place (MN-major)       -> columns [0, hd)   (S is dead)
//   dK_kt = dS^T Q       A = dS^T tile (K-major), B = Q rows in place (MN-major)       -> columns [cP, cP + hd)
//   dQ   += dS K_kt      A = the same dS^T tile read MN-major (M = queries), B = K_kt rows (MN-major)
//                        accumulates over the key tiles in columns [cQ, cQ + nq hd)
// S and dP are computed once (5 matrix products per tile pair); nothing is recomputed and there are no atomics on
// the gradients.  The bias gradients (column sums of dq, dk, dv) are reduced in shared memory per problem and added
// to global memory with 3 hd atomics per problem, as the mma.sync kernel does.
struct AttnTcBwdShape {
  int T, Nq, nq, nkt, heads, batch, stages;
  int q_bytes;      // Q (and dO): Nq rows, rounded up to 1024
  int k_bytes;      // K (and V): nkt * 128 rows
  int stage_bytes;  // Q + dO + K + V
  int t_bytes;      // one P^T / dS^T tile
  int cP, cQ;       // TMEM columns of dP^T and dQ
  long ldo, lddo, lddqkv;
  float q_scale;
};

// Sum over the 32 lanes of each of 32 per-lane values: lane l returns sum_lanes v[l] (31 shuffles).
__device__ __forceinline__ float warp_colsum32(float (&v)[32], int lane) {
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    const bool up = (lane & s) != 0;
#pragma unroll
    for (int i = 0; i < s; ++i) {
      const float send = up ? v[i] : v[i + s];
      const float keep = up ? v[i + s] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
  return v[0];
}

template <int HD>
__global__ void __launch_bounds__(kTcThreads, 1)
attn_tc_bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                   const __grid_constant__ CUtensorMap tmdO, const AttnTcBwdShape sh,
                   const __nv_bfloat16* __restrict__ out, const __nv_bfloat16* __restrict__ dout,
                   const float* __restrict__ lse, __nv_bfloat16* __restrict__ dqkv, float* __restrict__ dbias) {
  constexpr int ROWB = HD * 2;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sStage = smem;
  uint8_t* sPt = smem + sh.stages * sh.stage_bytes;
  uint8_t* sDSt = sPt + sh.t_bytes;
  float* sL2 = reinterpret_cast<float*>(sDSt + sh.t_bytes);  // [256] lse * log2(e) per query (+big for padding)
  float* sD = sL2 + 256;                                     // [256] rowsum(dO . O) per query
  float* sBias = sD + 256;                                   // [3][HD]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sBias + 3 * HD);
  uint64_t* kv_full = bars;       // [stages]
  uint64_t* kv_empty = bars + 4;  // [stages]
  uint64_t* s_full = bars + 8;
  uint64_t* p_full = bars + 9;
  uint64_t* o_full = bars + 10;
  uint64_t* t_free = bars + 11;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 12);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmdO);
    for (int s = 0; s < sh.stages; ++s) {
      mbar_init(&kv_full[s], 1);
      mbar_init(&kv_empty[s], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(p_full, 256);
    mbar_init(o_full, 1);
    mbar_init(t_free, 256);
    fence_barrier_init();
  }
  for (int i = threadIdx.x; i < 3 * HD; i += blockDim.x) sBias[i] = 0.f;
  if (warp == 1) tmem_alloc<512>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_wait();
  pdl_launch_dependents();

  const int nprob = sh.batch * sh.heads;
  const int my_prob = (nprob - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const int d = sh.heads * HD;

  if (warp == 0) {
    if (elect_one()) {
      for (int j = 0; j < my_prob; ++j) {
        const int prob = blockIdx.x + j * gridDim.x;
        const int b = prob / sh.heads, h = prob - b * sh.heads;
        const int st = j % sh.stages;
        mbar_wait(&kv_empty[st], ((j / sh.stages) & 1) ^ 1);
        uint8_t* dst = sStage + st * sh.stage_bytes;
        mbar_arrive_expect_tx(&kv_full[st], 2 * sh.Nq * ROWB + 2 * sh.nkt * 128 * ROWB);
        tma_load_3d(&tmQ, &kv_full[st], dst, h * HD, 0, b);
        tma_load_3d(&tmdO, &kv_full[st], dst + sh.q_bytes, h * HD, 0, b);
        tma_load_3d(&tmK, &kv_full[st], dst + 2 * sh.q_bytes, d + h * HD, 0, b);
        tma_load_3d(&tmK, &kv_full[st], dst + 2 * sh.q_bytes + sh.k_bytes, 2 * d + h * HD, 0, b);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      const uint32_t idesc_s = make_idesc_bf16(128, sh.Nq, 0, 0);   // K-major A and B
      const uint32_t idesc_kv = make_idesc_bf16(128, HD, 0, 1);     // K-major A (tile), MN-major B
      const uint32_t idesc_q = make_idesc_bf16(128, HD, 1, 1);      // MN-major A (tile, M = queries), MN-major B
      const uint32_t pt_addr = smem_u32(sPt), dst_addr = smem_u32(sDSt);
      int t = 0;  // tile counter of this CTA (barrier phases)
      for (int j = 0; j < my_prob; ++j) {
        const int st = j % sh.stages;
        const uint32_t q_addr = smem_u32(sStage + st * sh.stage_bytes);
        const uint32_t do_addr = q_addr + sh.q_bytes;
        const uint32_t k_addr = q_addr + 2 * sh.q_bytes;
        const uint32_t v_addr = k_addr + sh.k_bytes;
        mbar_wait(&kv_full[st], (j / sh.stages) & 1);
        tc_fence_after();
        for (int kt = 0; kt < sh.nkt; ++kt, ++t) {
          mbar_wait(t_free, (t & 1) ^ 1);  // previous tile's dV / dK (and dQ) have left tensor memory
          tc_fence_after();
#pragma unroll
          for (int ks = 0; ks < HD / 16; ++ks)
            umma_bf16(tmem_base, desc_kmajor<HD>(k_addr + kt * 128 * ROWB, ks), desc_kmajor<HD>(q_addr, ks), idesc_s,
                      ks > 0 ? 1u : 0u);
#pragma unroll
          for (int ks = 0; ks < HD / 16; ++ks)
            umma_bf16(tmem_base + sh.cP, desc_kmajor<HD>(v_addr + kt * 128 * ROWB, ks), desc_kmajor<HD>(do_addr, ks),
                      idesc_s, ks > 0 ? 1u : 0u);
          umma_commit(s_full);
          mbar_wait(p_full, t & 1);
          tc_fence_after();
          const int qsteps = sh.Nq >> 4;
          for (int kk = 0; kk < qsteps; ++kk)  // dV = P^T dO
            umma_bf16(tmem_base, desc_ptile(pt_addr, kk), desc_mnmajor<HD>(do_addr, kk), idesc_kv, kk > 0 ? 1u : 0u);
          for (int kk = 0; kk < qsteps; ++kk)  // dK = dS^T Q
            umma_bf16(tmem_base + sh.cP, desc_ptile(dst_addr, kk), desc_mnmajor<HD>(q_addr, kk), idesc_kv,
                      kk > 0 ? 1u : 0u);
          for (int mt = 0; mt < sh.nq; ++mt)  // dQ += dS K_kt  (M = 128 queries of tile mt, K = 128 keys)
            for (int kk = 0; kk < 8; ++kk)
              umma_bf16(tmem_base + sh.cQ + mt * HD,
                        make_smem_desc(dst_addr + mt * 2 * 16384 + kk * 2048, 16384, 1024, 2),
                        desc_mnmajor<HD>(k_addr + kt * 128 * ROWB, kk), idesc_q, (kt > 0 || kk > 0) ? 1u : 0u);
          umma_commit(o_full);
          if (kt == sh.nkt - 1) umma_commit(&kv_empty[st]);
        }
      }
    }
  } else if (warp >= kTcSoftmaxWarp0) {
    const int ctid = threadIdx.x - kTcSoftmaxWarp0 * 32;  // 0..255
    const int grp = ctid >> 7;                            // half of the query columns
    const int q = warp & 3;                               // TMEM lane quarter
    const int krow = q * 32 + lane;                       // key row inside the 128-key tile
    const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const uint32_t pt_addr = smem_u32(sPt), dst_addr = smem_u32(sDSt);
    const int nch = (sh.Nq + 31) >> 5;
    const int ch0 = grp == 0 ? 0 : (nch + 1) / 2, ch1 = grp == 0 ? (nch + 1) / 2 : nch;
    int t = 0;
    for (int j = 0; j < my_prob; ++j) {
      const int prob = blockIdx.x + j * gridDim.x;
      const int b = prob / sh.heads, h = prob - b * sh.heads;
      // per-query statistics: lse * log2(e) and D = rowsum(dO . O), one query per thread, rows from global memory
      // (their latency overlaps the TMA fill and the first two matrix products of this problem)
      {
        float L2 = 1e30f, Dv = 0.f;
        if (ctid < sh.T) {
          const long tokr = static_cast<long>(b) * sh.T + ctid;
          const __nv_bfloat16* orow = out + tokr * sh.ldo + h * HD;
          const __nv_bfloat16* grow = dout + tokr * sh.lddo + h * HD;
          uint4 ov[HD / 8], gv[HD / 8];
#pragma unroll
          for (int c = 0; c < HD / 8; ++c) {
            ov[c] = *reinterpret_cast<const uint4*>(orow + c * 8);
            gv[c] = *reinterpret_cast<const uint4*>(grow + c * 8);
          }
          L2 = lse[(static_cast<long>(b) * sh.heads + h) * sh.T + ctid] * kLog2eTc;
#pragma unroll
          for (int c = 0; c < HD / 8; ++c) {
            const uint32_t ow[4] = {ov[c].x, ov[c].y, ov[c].z, ov[c].w}, gw[4] = {gv[c].x, gv[c].y, gv[c].z, gv[c].w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float2 a = unpack_bf16(ow[e]), g2 = unpack_bf16(gw[e]);
              Dv = fmaf(a.x, g2.x, Dv);
              Dv = fmaf(a.y, g2.y, Dv);
            }
          }
        }
        sL2[ctid] = L2;
        sD[ctid] = Dv;
      }
      named_bar_sync(1, 256);
      for (int kt = 0; kt < sh.nkt; ++kt, ++t) {
        const int key = kt * 128 + krow;
        const bool key_ok = key < sh.T;
        const bool warp_live = kt * 128 + q * 32 < sh.T;
        mbar_wait(s_full, t & 1);
        tc_fence_after();
        const f32x2 k2 = pack2(kLog2eTc, kLog2eTc);
        for (int c = ch0; c < ch1; ++c) {
          uint32_t wp[16], wd[16];
          if (warp_live) {
            uint32_t rs[32], rp[32];
            tmem_ld_32x32(t_lane + c * 32, rs);
            tmem_ld_32x32(t_lane + sh.cP + c * 32, rp);
            tmem_ld_wait();
#pragma unroll
            for (int e = 0; e < 32; e += 4) {
              const float4 l4 = *reinterpret_cast<const float4*>(sL2 + c * 32 + e);
              const float4 d4 = *reinterpret_cast<const float4*>(sD + c * 32 + e);
              float x0, x1, x2, x3;
              unpack2(fma2(pack2(__uint_as_float(rs[e]), __uint_as_float(rs[e + 1])), k2, pack2(-l4.x, -l4.y)), x0, x1);
              unpack2(fma2(pack2(__uint_as_float(rs[e + 2]), __uint_as_float(rs[e + 3])), k2, pack2(-l4.z, -l4.w)), x2, x3);
              float p0 = ex2_approx(x0), p1 = ex2_approx(x1), p2 = ex2_approx(x2), p3 = ex2_approx(x3);
              if (!key_ok) p0 = p1 = p2 = p3 = 0.f;
              float s0, s1, s2, s3;
              unpack2(mul2(pack2(p0, p1), add2(pack2(__uint_as_float(rp[e]), __uint_as_float(rp[e + 1])), pack2(-d4.x, -d4.y))), s0, s1);
              unpack2(mul2(pack2(p2, p3), add2(pack2(__uint_as_float(rp[e + 2]), __uint_as_float(rp[e + 3])), pack2(-d4.z, -d4.w))), s2, s3);
              wp[e >> 1] = pack_bf16(p0, p1), wp[(e >> 1) + 1] = pack_bf16(p2, p3);
              wd[e >> 1] = pack_bf16(s0, s1), wd[(e >> 1) + 1] = pack_bf16(s2, s3);
            }
          } else {
#pragma unroll
            for (int e = 0; e < 16; ++e) wp[e] = 0u, wd[e] = 0u;
          }
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) {
            const uint32_t off = ptile_chunk(krow, c * 4 + jj);
            sts128(pt_addr + off, wp[4 * jj], wp[4 * jj + 1], wp[4 * jj + 2], wp[4 * jj + 3]);
            sts128(dst_addr + off, wd[4 * jj], wd[4 * jj + 1], wd[4 * jj + 2], wd[4 * jj + 3]);
          }
        }
        tc_fence_before();
        fence_proxy_async();
        mbar_arrive(p_full);
        mbar_wait(o_full, t & 1);
        tc_fence_after();
        // drain: group 0 takes dV (columns [0, hd)), group 1 takes dK (columns [cP, cP + hd)); thread = key row
        if (warp_live) {
          const uint32_t col0 = grp == 0 ? 0u : static_cast<uint32_t>(sh.cP);
          __nv_bfloat16* dst = dqkv + (static_cast<long>(b) * sh.T + key) * sh.lddqkv + (grp == 0 ? 2 * d : d) + h * HD;
#pragma unroll
          for (int c = 0; c < HD / 32; ++c) {
            uint32_t r[32];
            tmem_ld_32x32(t_lane + col0 + c * 32, r);
            tmem_ld_wait();
            float v[32];
#pragma unroll
            for (int e = 0; e < 32; ++e) v[e] = key_ok ? __uint_as_float(r[e]) : 0.f;
            if (key_ok) {
#pragma unroll
              for (int e = 0; e < 32; e += 8) {
                uint4 w;
                w.x = pack_bf16(v[e], v[e + 1]), w.y = pack_bf16(v[e + 2], v[e + 3]);
                w.z = pack_bf16(v[e + 4], v[e + 5]), w.w = pack_bf16(v[e + 6], v[e + 7]);
                *reinterpret_cast<uint4*>(dst + c * 32 + e) = w;
              }
            }
            if (dbias != nullptr) {
              const float cs = warp_colsum32(v, lane);
              atomicAdd(&sBias[(grp == 0 ? 2 : 1) * HD + c * 32 + lane], cs);
            }
          }
        }
        if (kt == sh.nkt - 1 && grp < sh.nq) {
          // dQ of query tile `grp` is complete: thread = query row
          const int qrow = grp * 128 + krow;
          const bool q_ok = qrow < sh.T;
          if (grp * 128 + q * 32 < sh.T) {
            __nv_bfloat16* dst = dqkv + (static_cast<long>(b) * sh.T + qrow) * sh.lddqkv + h * HD;
#pragma unroll
            for (int c = 0; c < HD / 32; ++c) {
              uint32_t r[32];
              tmem_ld_32x32(t_lane + sh.cQ + grp * HD + c * 32, r);
              tmem_ld_wait();
              float v[32];
#pragma unroll
              for (int e = 0; e < 32; ++e) v[e] = q_ok ? __uint_as_float(r[e]) * sh.q_scale : 0.f;
              if (q_ok) {
#pragma unroll
                for (int e = 0; e < 32; e += 8) {
                  uint4 w;
                  w.x = pack_bf16(v[e], v[e + 1]), w.y = pack_bf16(v[e + 2], v[e + 3]);
                  w.z = pack_bf16(v[e + 4], v[e + 5]), w.w = pack_bf16(v[e + 6], v[e + 7]);
                  *reinterpret_cast<uint4*>(dst + c * 32 + e) = w;
                }
              }
              if (dbias != nullptr) {
                const float cs = warp_colsum32(v, lane);
                atomicAdd(&sBias[c * 32 + lane], cs);
              }
            }
          }
        }
        tc_fence_before();
        mbar_arrive(t_free);
      }
      // bias gradients of this (image, head): shared-memory sums -> global atomics, then clear for the next one
      named_bar_sync(1, 256);
      if (dbias != nullptr && ctid < 3 * HD) {
        const int part = ctid / HD, c = ctid - part * HD;
        atomicAdd(dbias + part * d + h * HD + c, sBias[ctid]);
        sBias[ctid] = 0.f;
      }
      named_bar_sync(1, 256);
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
  sh->col_o = (sh->Nk / 2 + 31) / 32 * 32;
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
  const int fixed = 1024 + 256;
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

template <int HD>
int launch_tc_bwd(const jm_bf16* qkv, long ldqkv, const jm_bf16* out, long ldo, const jm_bf16* dout, long lddo,
                  const float* lse, int batch, int T, int heads, float q_scale, jm_bf16* dqkv, long lddqkv,
                  float* dbias, cudaStream_t stream) {
  AttnTcBwdShape sh;
  sh.T = T, sh.Nq = (T + 15) / 16 * 16, sh.nq = (T + 127) / 128, sh.nkt = sh.nq;
  sh.heads = heads, sh.batch = batch;
  sh.q_bytes = (sh.Nq * HD * 2 + 1023) / 1024 * 1024;
  sh.k_bytes = sh.nkt * 128 * HD * 2;
  sh.stage_bytes = 2 * sh.q_bytes + 2 * sh.k_bytes;
  const int blocks = (sh.Nq + 63) / 64;
  sh.t_bytes = (blocks > 2 * sh.nq ? blocks : 2 * sh.nq) * 16384;  // dQ's M = 128 reads two 64-query blocks per tile
  sh.cP = (sh.Nq + 31) / 32 * 32;
  if (sh.cP < HD) sh.cP = HD;  // dV / dK reuse the first head_dim columns of the S^T / dP^T regions
  sh.cQ = 2 * sh.cP;
  sh.ldo = ldo, sh.lddo = lddo, sh.lddqkv = lddqkv, sh.q_scale = q_scale;
  const int fixed = 1024 + 2 * sh.t_bytes + 2 * 256 * 4 + 3 * HD * 4 + 128;
  sh.stages = (kMaxDynSmem - fixed) / sh.stage_bytes;
  if (sh.stages > 3) sh.stages = 3;
  JM_CHECK(sh.stages >= 1 && sh.cQ + sh.nq * HD <= 512, "jm_attention_bwd: T=%d hd=%d does not fit", T, HD);
  const int smem = fixed + sh.stages * sh.stage_bytes;
  CUtensorMap tmQ, tmK, tmdO;
  const long d = static_cast<long>(heads) * HD;
  if (make_tmap_heads(&tmQ, qkv, 3 * d, T, batch, ldqkv, HD, sh.Nq)) return 1;
  if (make_tmap_heads(&tmK, qkv, 3 * d, T, batch, ldqkv, HD, sh.nkt * 128)) return 1;
  if (make_tmap_heads(&tmdO, dout, d, T, batch, lddo, HD, sh.Nq)) return 1;
  auto kern = attn_tc_bwd_kernel<HD>;
  static DeviceOnce attr_set;
  const int dev = current_device();
  JM_CHECK(dev >= 0, "jm_attention_bwd: no current CUDA device");
  if (!attr_set.done(dev)) {
    JM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem));
    attr_set.mark(dev);
  }
  const int nprob = batch * heads, sms = sm_count();
  const int grid = nprob < sms ? nprob : sms;
  JM_LAUNCH_PDL(kern, grid, kTcThreads, smem, stream, tmQ, tmK, tmdO, sh, reinterpret_cast<const __nv_bfloat16*>(out),
                reinterpret_cast<const __nv_bfloat16*>(dout), lse, reinterpret_cast<__nv_bfloat16*>(dqkv), dbias);
  JM_AFTER_LAUNCH();
  return 0;
}

}  // namespace

bool attn_tc_supported(int T, int hd);

bool attn_tc_bwd_supported(int T, int hd) {
  static const int on = env_int("JM_ATTN_TC_BWD", 1);
  if (!on || !attn_tc_supported(T, hd)) return false;
  const int Nq = (T + 15) / 16 * 16, nq = (T + 127) / 128;
  int cP = (Nq + 31) / 32 * 32;
  if (cP < hd) cP = hd;
  return 2 * cP + nq * hd <= 512;  // S^T, dP^T and dQ must fit the 512 TMEM columns
}

int attn_tc_bwd(const jm_bf16* qkv, long ldqkv, const jm_bf16* out, long ldo, const jm_bf16* dout, long lddo,
                const float* lse, int batch, int T, int heads, int hd, float q_scale, jm_bf16* dqkv, long lddqkv,
                float* dbias, cudaStream_t stream) {
  JM_CHECK(((reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(dqkv)) & 15) == 0 && ldo % 8 == 0 &&
               lddqkv % 8 == 0 && ldqkv >= 3L * heads * hd && lddqkv >= 3L * heads * hd,
           "jm_attention_bwd: buffers must be 16-byte aligned, strides multiples of 8 and >= 3 * heads * head_dim");
  if (hd == 32)
    return launch_tc_bwd<32>(qkv, ldqkv, out, ldo, dout, lddo, lse, batch, T, heads, q_scale, dqkv, lddqkv, dbias, stream);
  return launch_tc_bwd<64>(qkv, ldqkv, out, ldo, dout, lddo, lse, batch, T, heads, q_scale, dqkv, lddqkv, dbias, stream);
}

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
