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
  int cs;           // key columns [0, cs) belong to softmax half 0, [cs, Nk) to half 1 (multiples of 16)
  int col_o;        // TMEM column (inside a slot) of O: a range that holds no P
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

// Optional timeline (build with -DJM_ATTN_TRACE): CTA 0 stamps clock64() at the hand-over points of its first items.
#ifdef JM_ATTN_TRACE
__device__ long long g_attn_trace[64 * 16];
#define JM_TRACE(item, id, cond) \
  do { if (blockIdx.x == 0 && (item) < 64 && (cond)) g_attn_trace[(item) * 16 + (id)] = clock64(); } while (0)
#else
#define JM_TRACE(item, id, cond) do { } while (0)
#endif

// ------------------------------------------------------------------------------------------------ forward
// Softmax work on one accumulator tile is split between two groups of four warps by key columns: half 0 owns
// columns [0, cs), half 1 owns [cs, Nk).  The halves exchange their row maxima (and later their row sums) through
// shared memory with a named barrier, so that 16 softmax warps (4 per scheduler) are in flight per SM.
constexpr int kFwdThreads = 640;  // warps 0-3 as above, warps 4-19: [slot 2][half 2][lane quarter 4]

template <int W>
__device__ __forceinline__ void tmem_ld_w(uint32_t taddr, uint32_t (&r)[W]) {
  if constexpr (W == 32) tmem_ld_32x32(taddr, r);
  else tmem_ld_32x16(taddr, r);
}

// running maximum over W score columns starting at column c0 (columns >= T are padding)
template <int W>
__device__ __forceinline__ void fwd_chunk_max(uint32_t t_row, int c0, int T, float& mx) {
  uint32_t r[W];
  tmem_ld_w<W>(t_row + c0, r);
  tmem_ld_wait();
  if (c0 + W > T) {
#pragma unroll
    for (int e = 0; e < W; ++e)
      if (c0 + e >= T) r[e] = 0xff800000u;  // -inf
  }
#pragma unroll
  for (int e = 0; e < W; e += 2) mx = fmax3(mx, __uint_as_float(r[e]), __uint_as_float(r[e + 1]));
}

// P = exp2(s log2e - max log2e) for W columns at c0; bf16 pairs go back to tensor memory at column `dst`
template <int W>
__device__ __forceinline__ void fwd_chunk_exp(uint32_t t_row, int c0, int dst, int T, f32x2 k2, f32x2 nm2, f32x2& l2) {
  uint32_t r[W];
  tmem_ld_w<W>(t_row + c0, r);
  tmem_ld_wait();
  if (c0 + W > T) {
#pragma unroll
    for (int e = 0; e < W; ++e)
      if (c0 + e >= T) r[e] = 0xff800000u;
  }
  uint32_t w[W / 2];
#pragma unroll
  for (int e = 0; e < W; e += 2) {
    float x0, x1;
    unpack2(fma2(pack2(__uint_as_float(r[e]), __uint_as_float(r[e + 1])), k2, nm2), x0, x1);
    const float p0 = ex2_approx(x0), p1 = ex2_approx(x1);
    l2 = add2(l2, pack2(p0, p1));
    w[e >> 1] = pack_bf16(p0, p1);
  }
  if constexpr (W == 32) tmem_st_32x16(t_row + dst, w);
  else tmem_st_32x8(t_row + dst, w);
}

template <int HD>
__global__ void __launch_bounds__(kFwdThreads, 1)
attn_tc_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                   const AttnTcShape sh, __nv_bfloat16* __restrict__ out, float* __restrict__ lse) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sStage = smem;
  float* sXm = reinterpret_cast<float*>(smem + sh.stages * sh.stage_bytes);  // [slot][half][128] row maxima
  float* sXl = sXm + 512;                                                    // [slot][half][128] row sums
  uint64_t* bars = reinterpret_cast<uint64_t*>(sXl + 512);
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
      mbar_init(&p_full[s], 256);
      mbar_init(&o_full[s], 1);
      mbar_init(&o_empty[s], 256);
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
        JM_TRACE(i, 8, true);
        const uint32_t v_addr = smem_u32(sStage + st * sh.stage_bytes + sh.q_bytes + sh.kv_bytes);
        const int ksteps = sh.Nk >> 4, k_half = sh.cs >> 4;
        const uint32_t t_p = tmem_base + slot * kSlotCols;
        // P: bf16 pairs, 8 columns per 16-key step; half 0 packed its keys at [0, cs / 2), half 1 at [cs, ...)
        for (int kk = 0; kk < ksteps; ++kk)
          umma_bf16_ts(t_p + sh.col_o, t_p + (kk < k_half ? kk * 8 : sh.cs + (kk - k_half) * 8),
                       desc_mnmajor<HD>(v_addr, kk), idesc_o, kk > 0 ? 1u : 0u);
        umma_commit(&o_full[slot]);
        JM_TRACE(i, 9, true);
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
        JM_TRACE(i, 10, true);
        mbar_wait(&o_empty[slot], ((i >> 1) & 1) ^ 1);  // the slot's previous O has been read out
        tc_fence_after();
        JM_TRACE(i, 11, true);
        const uint32_t q_addr = smem_u32(sStage + st * sh.stage_bytes + qt * 128 * HD * 2);
        const uint32_t k_addr = smem_u32(sStage + st * sh.stage_bytes + sh.q_bytes);
#pragma unroll
        for (int ks = 0; ks < HD / 16; ++ks)
          umma_bf16(tmem_base + slot * kSlotCols, desc_kmajor<HD>(q_addr, ks), desc_kmajor<HD>(k_addr, ks), idesc_s,
                    ks > 0 ? 1u : 0u);
        umma_commit(&s_full[slot]);
        JM_TRACE(i, 12, true);
        if (i >= 1 && !pv_first) issue_pv(i - 1);
      }
      if (n_items >= 1) issue_pv(n_items - 1);
    }
  } else if (warp >= kTcSoftmaxWarp0) {
    const int sw = warp - kTcSoftmaxWarp0;
    const int slot = sw >> 3;       // TMEM slot == parity of the item
    const int half = (sw >> 2) & 1; // which key columns of the tile
    const int q = warp & 3;         // TMEM lane quarter
    const int row = q * 32 + lane;  // row of the 128-query tile
    const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + slot * kSlotCols;
    const int h0 = half == 0 ? 0 : sh.cs, h1 = half == 0 ? sh.cs : sh.Nk;
    const int n32 = (h1 - h0) >> 5;
    const bool tail16 = ((h1 - h0) & 16) != 0;
    float* xm_mine = sXm + (slot * 2 + half) * 128 + row;
    float* xm_other = sXm + (slot * 2 + (half ^ 1)) * 128 + row;
    float* xl_mine = sXl + (slot * 2 + half) * 128 + row;
    float* xl_other = sXl + (slot * 2 + (half ^ 1)) * 128 + row;
    constexpr int OC = HD / 2;  // output columns drained by this half
    for (int i = slot; i < n_items; i += 2) {
      const int j = i / sh.nq, qt = i - j * sh.nq;
      const int prob = blockIdx.x + j * gridDim.x;
      const int b = prob / sh.heads, h = prob - b * sh.heads;
      const int tok = qt * 128 + row;
      const bool warp_live = qt * 128 + q * 32 < sh.T;  // warp-uniform: any valid row in this warp's 32
      const uint32_t ph = (i >> 1) & 1;
      const bool tr = half == 0 && q == 0 && lane == 0;
      JM_TRACE(i, 0, tr);
      mbar_wait(&s_full[slot], ph);
      tc_fence_after();
      JM_TRACE(i, 1, tr);
      float mx = -INFINITY;
      if (warp_live) {
        for (int c = 0; c < n32; ++c) fwd_chunk_max<32>(t_row, h0 + c * 32, sh.T, mx);
        if (tail16) fwd_chunk_max<16>(t_row, h0 + n32 * 32, sh.T, mx);
      }
      *xm_mine = mx;
      JM_TRACE(i, 2, tr);
      named_bar_sync(1 + slot, 256);
      mx = fmaxf(mx, *xm_other);
      JM_TRACE(i, 3, tr);
      float l = 0.f;
      if (warp_live) {
        const float nm = -mx * kLog2eTc;
        const f32x2 k2 = pack2(kLog2eTc, kLog2eTc), nm2 = pack2(nm, nm);
        f32x2 l2 = pack2(0.f, 0.f);
        // P overwrites S in place inside this half's own columns: chunk [c0, c0 + W) becomes W / 2 columns at
        // h0 + (c0 - h0) / 2, which this thread has already consumed
        for (int c = 0; c < n32; ++c) fwd_chunk_exp<32>(t_row, h0 + c * 32, h0 + c * 16, sh.T, k2, nm2, l2);
        if (tail16) fwd_chunk_exp<16>(t_row, h0 + n32 * 32, h0 + n32 * 16, sh.T, k2, nm2, l2);
        float la, lb;
        unpack2(l2, la, lb);
        l = la + lb;
        tmem_st_wait();
      }
      *xl_mine = l;
      // P is complete (tensor memory, the A operand of O = P V) and S has been consumed
      tc_fence_before();
      JM_TRACE(i, 4, tr);
      mbar_arrive(&p_full[slot]);
      mbar_wait(&o_full[slot], ph);
      tc_fence_after();
      JM_TRACE(i, 5, tr);
      named_bar_sync(1 + slot, 256);  // the other half's row sum is visible
      l += *xl_other;
      if (warp_live) {
        const float inv = 1.0f / l;
        uint32_t r[OC];
        tmem_ld_w<OC>(t_row + sh.col_o + half * OC, r);
        tmem_ld_wait();
        if (tok < sh.T) {
          __nv_bfloat16* dst = out + (static_cast<long>(b) * sh.T + tok) * sh.ldo + h * HD + half * OC;
#pragma unroll
          for (int e = 0; e < OC; e += 8) {
            uint4 v;
            v.x = pack_bf16(__uint_as_float(r[e]) * inv, __uint_as_float(r[e + 1]) * inv);
            v.y = pack_bf16(__uint_as_float(r[e + 2]) * inv, __uint_as_float(r[e + 3]) * inv);
            v.z = pack_bf16(__uint_as_float(r[e + 4]) * inv, __uint_as_float(r[e + 5]) * inv);
            v.w = pack_bf16(__uint_as_float(r[e + 6]) * inv, __uint_as_float(r[e + 7]) * inv);
            *reinterpret_cast<uint4*>(dst + e) = v;
          }
          if (half == 0) lse[(static_cast<long>(b) * sh.heads + h) * sh.T + tok] = mx + logf(l);
        }
      }
      tc_fence_before();
      JM_TRACE(i, 6, tr);
      mbar_arrive(&o_empty[slot]);
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
//   P^T  = exp2(S^T c - lse c),  dS^T = P^T (dP^T - D)        sixteen warps: four per TMEM lane quarter, each a
//          quarter of the query columns (in units of 16); lse and D = rowsum(dO . O) per query come from shared
//          memory (broadcast reads).  bf16 P^T and dS^T overwrite S^T and dP^T in place as packed pairs (inside
//          the column range of the warp that read them); dS^T also goes to a SWIZZLE_128B tile [key][query]
//   dV_kt = P^T dO       A = P^T from tensor memory, B = dO rows in place (MN-major)          -> columns cV
//   dK_kt = dS^T Q       A = dS^T from tensor memory, B = Q rows in place (MN-major)          -> columns cK
//   dQ   += dS K_kt      A = the dS^T tile read MN-major (M = queries), B = K_kt rows (MN-major)
//                        accumulates over the key tiles in columns [cQ, cQ + nq hd)
// S and dP are computed once (5 matrix products per tile); nothing is recomputed and there are no atomics on the
// gradients.  Bias gradients: sum_q dQ (column sums at the drain), sum_key dV = sum_q dO (rows of P sum to one),
// sum_key dK = 0 (rows of dS sum to zero); reduced in shared memory, 2 hd global atomics per problem.
constexpr int kBwdThreads = 640;  // warps 0-3 as in the forward, warps 4-19: [column group 4][lane quarter 4]

struct AttnTcBwdShape {
  int T, Nq, nq, nkt, heads, batch, stages;
  int q_bytes;      // Q (and dO): Nq rows, rounded up to 1024
  int k_bytes;      // K (and V): nkt * 128 rows
  int stage_bytes;  // Q + dO + K + V
  int t_bytes;      // the dS^T tile
  int cP, cQ, cV, cK;  // TMEM columns of dP^T, dQ, dV, dK
  int ustart[5];    // column group g owns the 16-query units [ustart[g], ustart[g + 1])
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
__global__ void __launch_bounds__(kBwdThreads, 1)
attn_tc_bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                   const __grid_constant__ CUtensorMap tmdO, const AttnTcBwdShape sh,
                   const __nv_bfloat16* __restrict__ out, const __nv_bfloat16* __restrict__ dout,
                   const float* __restrict__ lse, __nv_bfloat16* __restrict__ dqkv, float* __restrict__ dbias) {
  constexpr int ROWB = HD * 2;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sStage = smem;
  uint8_t* sDSt = smem + sh.stages * sh.stage_bytes;
  float* sL2 = reinterpret_cast<float*>(sDSt + sh.t_bytes);  // [256] lse * log2(e) per query (+big for padding)
  float* sD = sL2 + 256;                                     // [256] rowsum(dO . O) per query
  float* sBias = sD + 256;                                   // [2][HD]: dq, dv column sums
  uint64_t* bars = reinterpret_cast<uint64_t*>(sBias + 2 * HD);
  uint64_t* kv_full = bars;       // [stages]
  uint64_t* kv_empty = bars + 4;  // [stages]
  uint64_t* s_full = bars + 8;    // S^T and dP^T of the tile are in tensor memory
  uint64_t* p_full = bars + 9;    // P^T, dS^T written (512 arrivals)
  uint64_t* vk_full = bars + 10;  // dV, dK of the tile are complete
  uint64_t* q_full = bars + 11;   // dQ contribution of the tile is complete
  uint64_t* t_free = bars + 12;   // the tile's results have left tensor memory (512 arrivals)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 13);

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
    mbar_init(p_full, 512);
    mbar_init(vk_full, 1);
    mbar_init(q_full, 1);
    mbar_init(t_free, 512);
    fence_barrier_init();
  }
  for (int i = threadIdx.x; i < 2 * HD; i += blockDim.x) sBias[i] = 0.f;
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
      const uint32_t idesc_kv = make_idesc_bf16(128, HD, 0, 1);     // A from tensor memory, MN-major B
      const uint32_t idesc_q = make_idesc_bf16(128, HD, 1, 1);      // MN-major A (tile, M = queries), MN-major B
      const uint32_t dst_addr = smem_u32(sDSt);
      constexpr uint64_t kStepMn = (16 * ROWB) >> 4;  // descriptor start-address increment per 16-row k-step
      int t = 0;  // tile counter of this CTA (barrier phases)
      for (int j = 0; j < my_prob; ++j) {
        const int st = j % sh.stages;
        const uint32_t q_addr = smem_u32(sStage + st * sh.stage_bytes);
        const uint32_t do_addr = q_addr + sh.q_bytes;
        const uint32_t k_addr = q_addr + 2 * sh.q_bytes;
        const uint32_t v_addr = k_addr + sh.k_bytes;
        JM_TRACE(t, 8, true);
        mbar_wait(&kv_full[st], (j / sh.stages) & 1);
        tc_fence_after();
        for (int kt = 0; kt < sh.nkt; ++kt, ++t) {
          JM_TRACE(t, 9, true);
          mbar_wait(t_free, (t & 1) ^ 1);  // previous tile's dV / dK (and dQ) have left tensor memory
          tc_fence_after();
          JM_TRACE(t, 10, true);
#pragma unroll
          for (int ks = 0; ks < HD / 16; ++ks)
            umma_bf16(tmem_base, desc_kmajor<HD>(k_addr + kt * 128 * ROWB, ks), desc_kmajor<HD>(q_addr, ks), idesc_s,
                      ks > 0 ? 1u : 0u);
#pragma unroll
          for (int ks = 0; ks < HD / 16; ++ks)
            umma_bf16(tmem_base + sh.cP, desc_kmajor<HD>(v_addr + kt * 128 * ROWB, ks), desc_kmajor<HD>(do_addr, ks),
                      idesc_s, ks > 0 ? 1u : 0u);
          umma_commit(s_full);
          JM_TRACE(t, 11, true);
          mbar_wait(p_full, t & 1);
          tc_fence_after();
          JM_TRACE(t, 12, true);
          // dV = P^T dO and dK = dS^T Q: k-step kk = 16 queries; group g packed its unit kk at 8 (kk + ustart[g])
          {
            uint64_t d_do = desc_mnmajor<HD>(do_addr, 0), d_q = desc_mnmajor<HD>(q_addr, 0);
            int kk = 0;
            for (int g = 0; g < 4; ++g)
              for (; kk < sh.ustart[g + 1]; ++kk, d_do += kStepMn)
                umma_bf16_ts(tmem_base + sh.cV, tmem_base + 8 * (kk + sh.ustart[g]), d_do, idesc_kv, kk > 0 ? 1u : 0u);
            kk = 0;
            for (int g = 0; g < 4; ++g)
              for (; kk < sh.ustart[g + 1]; ++kk, d_q += kStepMn)
                umma_bf16_ts(tmem_base + sh.cK, tmem_base + sh.cP + 8 * (kk + sh.ustart[g]), d_q, idesc_kv,
                             kk > 0 ? 1u : 0u);
          }
          umma_commit(vk_full);
          JM_TRACE(t, 13, true);
          for (int mt = 0; mt < sh.nq; ++mt) {  // dQ += dS K_kt  (M = 128 queries of tile mt, K = 128 keys)
            uint64_t d_a = make_smem_desc(dst_addr + mt * 2 * 16384, 16384, 1024, 2);
            uint64_t d_k = desc_mnmajor<HD>(k_addr + kt * 128 * ROWB, 0);
#pragma unroll
            for (int kk = 0; kk < 8; ++kk, d_a += (2048 >> 4), d_k += kStepMn)
              umma_bf16(tmem_base + sh.cQ + mt * HD, d_a, d_k, idesc_q, (kt > 0 || kk > 0) ? 1u : 0u);
          }
          umma_commit(q_full);
          JM_TRACE(t, 14, true);
          if (kt == sh.nkt - 1) umma_commit(&kv_empty[st]);
        }
      }
    }
  } else if (warp >= kTcSoftmaxWarp0) {
    const int ctid = threadIdx.x - kTcSoftmaxWarp0 * 32;  // 0..511
    const int grp = ctid >> 7;                            // column group
    const int q = warp & 3;                               // TMEM lane quarter
    const int krow = q * 32 + lane;                       // key row inside the 128-key tile
    const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const uint32_t dst_addr = smem_u32(sDSt);
    const int u0 = sh.ustart[grp], u1 = sh.ustart[grp + 1];
    constexpr int DC = HD / 2;  // result columns drained per thread
    int t = 0;
    for (int j = 0; j < my_prob; ++j) {
      const int prob = blockIdx.x + j * gridDim.x;
      const int b = prob / sh.heads, h = prob - b * sh.heads;
      // per-query statistics: lse * log2(e) and D = rowsum(dO . O), one query per thread, rows from global memory
      // (their latency overlaps the TMA fill and the first two matrix products of this problem)
      const bool tr = ctid == 0;
      JM_TRACE(t, 0, tr);
      if (ctid < 256) {
        float L2 = 1e30f, Dv = 0.f;
        const bool live = ctid < sh.T;
        uint4 gv[HD / 8];
#pragma unroll
        for (int c = 0; c < HD / 8; ++c) gv[c] = make_uint4(0u, 0u, 0u, 0u);
        if (live) {
          const long tokr = static_cast<long>(b) * sh.T + ctid;
          const __nv_bfloat16* orow = out + tokr * sh.ldo + h * HD;
          const __nv_bfloat16* grow = dout + tokr * sh.lddo + h * HD;
          uint4 ov[HD / 8];
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
        if (dbias != nullptr && (ctid & ~31) < sh.T) {  // warp-uniform: this warp holds at least one live query
          // bias gradient of v: sum over keys of dV = sum over queries of dO (every row of P sums to one)
#pragma unroll
          for (int c = 0; c < HD / 32; ++c) {
            float v[32];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const uint4 u4 = gv[c * 4 + e];
              const uint32_t gw[4] = {u4.x, u4.y, u4.z, u4.w};
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const float2 g2 = unpack_bf16(gw[k]);
                v[e * 8 + 2 * k] = g2.x, v[e * 8 + 2 * k + 1] = g2.y;
              }
            }
            const float cs = warp_colsum32(v, lane);
            atomicAdd(&sBias[HD + c * 32 + lane], cs);
          }
        }
      }
      named_bar_sync(1, 512);
      JM_TRACE(t, 1, tr);
      for (int kt = 0; kt < sh.nkt; ++kt, ++t) {
        const int key = kt * 128 + krow;
        const bool key_ok = key < sh.T;
        const bool warp_live = kt * 128 + q * 32 < sh.T;
        JM_TRACE(t, 2, tr);
        mbar_wait(s_full, t & 1);
        tc_fence_after();
        JM_TRACE(t, 3, tr);
        const f32x2 k2 = pack2(kLog2eTc, kLog2eTc);
        for (int u = u0; u < u1; ++u) {
          const int c0 = u * 16;
          uint32_t wd[8];
          if (warp_live) {
            uint32_t rs[16], rp[16], wp[8];
            tmem_ld_32x16(t_lane + c0, rs);
            tmem_ld_32x16(t_lane + sh.cP + c0, rp);
            tmem_ld_wait();
#pragma unroll
            for (int e = 0; e < 16; e += 4) {
              const float4 l4 = *reinterpret_cast<const float4*>(sL2 + c0 + e);
              const float4 d4 = *reinterpret_cast<const float4*>(sD + c0 + e);
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
            // packed pairs go back in place, inside this warp's own column range (already consumed)
            tmem_st_32x8(t_lane + 8 * (u + u0), wp);
            tmem_st_32x8(t_lane + sh.cP + 8 * (u + u0), wd);
          } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) wd[e] = 0u;
          }
          // dS^T also goes to shared memory (dead key rows as zeros: they are contracted over in dQ)
          sts128(dst_addr + ptile_chunk(krow, 2 * u), wd[0], wd[1], wd[2], wd[3]);
          sts128(dst_addr + ptile_chunk(krow, 2 * u + 1), wd[4], wd[5], wd[6], wd[7]);
        }
        if (warp_live) tmem_st_wait();
        tc_fence_before();
        fence_proxy_async();
        JM_TRACE(t, 4, tr);
        mbar_arrive(p_full);
        mbar_wait(vk_full, t & 1);
        tc_fence_after();
        JM_TRACE(t, 5, tr);
        // drain: groups 0, 1 take the two halves of dV's columns, groups 2, 3 of dK's; thread = key row
        if (warp_live) {
          const uint32_t col0 = (grp < 2 ? sh.cV : sh.cK) + (grp & 1) * DC;
          uint32_t r[DC];
          tmem_ld_w<DC>(t_lane + col0, r);
          tmem_ld_wait();
          if (key_ok) {
            __nv_bfloat16* dst = dqkv + (static_cast<long>(b) * sh.T + key) * sh.lddqkv + (grp < 2 ? 2 * d : d) + h * HD +
                                 (grp & 1) * DC;
#pragma unroll
            for (int e = 0; e < DC; e += 8) {
              uint4 w;
              w.x = pack_bf16(__uint_as_float(r[e]), __uint_as_float(r[e + 1]));
              w.y = pack_bf16(__uint_as_float(r[e + 2]), __uint_as_float(r[e + 3]));
              w.z = pack_bf16(__uint_as_float(r[e + 4]), __uint_as_float(r[e + 5]));
              w.w = pack_bf16(__uint_as_float(r[e + 6]), __uint_as_float(r[e + 7]));
              *reinterpret_cast<uint4*>(dst + e) = w;
            }
          }
        }
        if (kt == sh.nkt - 1) {
          mbar_wait(q_full, t & 1);
          tc_fence_after();
          // dQ is complete: groups (2 mt, 2 mt + 1) drain the two column halves of query tile mt; thread = query row
          const int mt = grp >> 1;
          if (mt < sh.nq && mt * 128 + q * 32 < sh.T) {
            const int qrow = mt * 128 + krow;
            const bool q_ok = qrow < sh.T;
            uint32_t r[DC];
            tmem_ld_w<DC>(t_lane + sh.cQ + mt * HD + (grp & 1) * DC, r);
            tmem_ld_wait();
            float v[32];
#pragma unroll
            for (int e = 0; e < 32; ++e) v[e] = (e < DC && q_ok) ? __uint_as_float(r[e < DC ? e : 0]) * sh.q_scale : 0.f;
            if (q_ok) {
              __nv_bfloat16* dst = dqkv + (static_cast<long>(b) * sh.T + qrow) * sh.lddqkv + h * HD + (grp & 1) * DC;
#pragma unroll
              for (int e = 0; e < DC; e += 8) {
                uint4 w;
                w.x = pack_bf16(v[e], v[e + 1]), w.y = pack_bf16(v[e + 2], v[e + 3]);
                w.z = pack_bf16(v[e + 4], v[e + 5]), w.w = pack_bf16(v[e + 6], v[e + 7]);
                *reinterpret_cast<uint4*>(dst + e) = w;
              }
            }
            if (dbias != nullptr) {
              const float cs = warp_colsum32(v, lane);
              if (lane < DC) atomicAdd(&sBias[(grp & 1) * DC + lane], cs);
            }
          }
        }
        tc_fence_before();
        JM_TRACE(t, 6, tr);
        mbar_arrive(t_free);
      }
      // bias gradients of this (image, head): shared-memory sums -> global atomics, then clear for the next one
      named_bar_sync(1, 512);
      if (dbias != nullptr && ctid < 2 * HD) {
        const int part = ctid / HD, c = ctid - part * HD;  // 0: dq, 1: dv
        atomicAdd(dbias + (part == 0 ? 0 : 2 * d) + h * HD + c, sBias[ctid]);
        sBias[ctid] = 0.f;
      }
      named_bar_sync(1, 512);
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
  // P is packed in place at [0, cs / 2) and [cs, cs + (Nk - cs) / 2).  O goes behind the scores when the 256-column
  // slot has room, otherwise into the gap [cs / 2, cs) that half 0 leaves (cs is then at least 2 * head_dim).
  sh->cs = (sh->Nk / 2 + 15) / 16 * 16;
  const int behind = (sh->Nk + 31) / 32 * 32;
  if (behind + hd <= kSlotCols) {
    sh->col_o = behind;
  } else {
    if (sh->cs < 2 * hd) sh->cs = 2 * hd;
    sh->col_o = (sh->cs / 2 + 31) / 32 * 32;
  }
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
  const int fixed = 1024 + 256 + 2 * 512 * 4;
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
  JM_LAUNCH_PDL(kern, grid, kFwdThreads, smem, stream, tmQ, tmKV, sh, reinterpret_cast<__nv_bfloat16*>(out), lse);
  JM_AFTER_LAUNCH();
  return 0;
}

// Tensor-memory and column-group layout of the backward kernel; false when the problem does not fit.
bool bwd_layout(AttnTcBwdShape* sh, int T, int hd) {
  sh->T = T, sh->Nq = (T + 15) / 16 * 16, sh->nq = (T + 127) / 128, sh->nkt = sh->nq;
  const int U = sh->Nq / 16, base = U / 4, rem = U % 4;
  sh->ustart[0] = 0;
  for (int g = 0; g < 4; ++g) sh->ustart[g + 1] = sh->ustart[g] + base + (g < rem ? 1 : 0);
  sh->cP = (sh->Nq + 31) / 32 * 32;
  sh->cQ = 2 * sh->cP;
  if (sh->cQ + sh->nq * hd + 2 * hd <= 512) {  // room for dV and dK of their own
    sh->cV = sh->cQ + sh->nq * hd, sh->cK = sh->cV + hd;
    return true;
  }
  // otherwise dV / dK go into the gap that column group 0 leaves behind its packed pairs: [8 n0, 16 n0) of the
  // S^T region (and of the dP^T region), n0 = units of group 0
  const int n0 = sh->ustart[1];
  if (sh->cQ + sh->nq * hd > 512 || 8 * n0 < hd) return false;
  sh->cV = 8 * n0, sh->cK = sh->cP + 8 * n0;
  return true;
}

template <int HD>
int launch_tc_bwd(const jm_bf16* qkv, long ldqkv, const jm_bf16* out, long ldo, const jm_bf16* dout, long lddo,
                  const float* lse, int batch, int T, int heads, float q_scale, jm_bf16* dqkv, long lddqkv,
                  float* dbias, cudaStream_t stream) {
  AttnTcBwdShape sh;
  JM_CHECK(bwd_layout(&sh, T, HD), "jm_attention_bwd: T=%d hd=%d does not fit tensor memory", T, HD);
  sh.heads = heads, sh.batch = batch;
  sh.q_bytes = (sh.Nq * HD * 2 + 1023) / 1024 * 1024;
  sh.k_bytes = sh.nkt * 128 * HD * 2;
  sh.stage_bytes = 2 * sh.q_bytes + 2 * sh.k_bytes;
  const int blocks = (sh.Nq + 63) / 64;
  sh.t_bytes = (blocks > 2 * sh.nq ? blocks : 2 * sh.nq) * 16384;  // dQ's M = 128 reads two 64-query blocks per tile
  sh.ldo = ldo, sh.lddo = lddo, sh.lddqkv = lddqkv, sh.q_scale = q_scale;
  const int fixed = 1024 + sh.t_bytes + 2 * 256 * 4 + 2 * HD * 4 + 128;
  sh.stages = (kMaxDynSmem - fixed) / sh.stage_bytes;
  if (sh.stages > 3) sh.stages = 3;
  JM_CHECK(sh.stages >= 1, "jm_attention_bwd: T=%d hd=%d does not fit shared memory", T, HD);
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
  JM_LAUNCH_PDL(kern, grid, kBwdThreads, smem, stream, tmQ, tmK, tmdO, sh, reinterpret_cast<const __nv_bfloat16*>(out),
                reinterpret_cast<const __nv_bfloat16*>(dout), lse, reinterpret_cast<__nv_bfloat16*>(dqkv), dbias);
  JM_AFTER_LAUNCH();
  return 0;
}

}  // namespace

bool attn_tc_supported(int T, int hd);

#ifdef JM_ATTN_TRACE
extern "C" int jm_debug_attn_trace(long long* host_buf) {  // 64 items x 16 stamps
  JM_CUDA(cudaMemcpyFromSymbol(host_buf, g_attn_trace, sizeof(long long) * 64 * 16));
  return 0;
}
#endif

bool attn_tc_bwd_supported(int T, int hd) {
  static const int on = env_int("JM_ATTN_TC_BWD", 1);
  if (!on || !attn_tc_supported(T, hd)) return false;
  AttnTcBwdShape sh;
  return bwd_layout(&sh, T, hd);
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
