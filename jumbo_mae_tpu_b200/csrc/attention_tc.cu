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

// Byte offset of 16-byte chunk `chunk` of row `row` in a [rows][HD] bf16 tile laid out as TMA writes / reads it
// (SWIZZLE_64B: chunk ^ (row / 2 % 4), SWIZZLE_128B: chunk ^ (row % 8)).
template <int HD>
__device__ __forceinline__ uint32_t rowtile_chunk(int row, int chunk) {
  const int x = HD == 32 ? ((row >> 1) & 3) : (row & 7);
  return static_cast<uint32_t>(row * (HD * 2) + ((chunk ^ x) << 4));
}

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
                   const __grid_constant__ CUtensorMap tmO, const AttnTcShape sh, float* __restrict__ lse) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sStage = smem;
  uint8_t* sO = smem + sh.stages * sh.stage_bytes;                           // [slot][128 rows][HD] bf16 output tiles
  float* sXm = reinterpret_cast<float*>(sO + 2 * 128 * HD * 2);              // [slot][half][128] row maxima
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
    tma_prefetch_desc(&tmO);
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
#ifdef JM_EXP_NO_TMA
        if (j >= sh.stages) { mbar_arrive(&kv_full[st]); continue; }
#endif
        mbar_arrive_expect_tx(&kv_full[st], sh.nq * 128 * HD * 2 + 2 * sh.Nk * HD * 2);
        for (int qt = 0; qt < sh.nq; ++qt)
          tma_load_3d(&tmQ, &kv_full[st], dst + qt * 128 * HD * 2, h * HD, qt * 128, b);
        tma_load_3d(&tmKV, &kv_full[st], dst + sh.q_bytes, d + h * HD, 0, b);
        tma_load_3d(&tmKV, &kv_full[st], dst + sh.q_bytes + sh.kv_bytes, 2 * d + h * HD, 0, b);
      }
    }
  } else if (warp == 1) {
    // The whole warp runs the control flow (so that addresses and descriptors stay in uniform registers); one
    // elected lane issues.  Event-driven: whichever of "scores of the next item" and "P V of the oldest finished
    // softmax" has its inputs is issued first.
    const uint32_t idesc_s = make_idesc_bf16(128, sh.Nk, 0, 0);
    const uint32_t idesc_o = make_idesc_bf16(128, HD, 0, 1);
    // O behind the scores: the slot can take the next scores as soon as P V has READ P (its own commit), long before
    // the softmax warps have stored O; otherwise (O inside the score columns) only once O has been read out.
    const bool o_behind = sh.col_o >= sh.Nk;
    const int ksteps = sh.Nk >> 4, k_half = sh.cs >> 4;
    int s_i = 0, s_j = 0, s_qt = 0;  // next item to get its scores: (problem, query tile)
    int p_i = 0, p_j = 0, p_qt = 0;  // next item to get its P V
    const long long t_start = clock64();
    while (p_i < n_items) {
      bool progress = false;
      if (s_i < n_items) {
        const int slot = s_i & 1, st = s_j % sh.stages;
        const uint32_t use = (s_i >> 1) & 1;
        bool ready = s_qt != 0 || mbar_try_wait_warp(&kv_full[st], (s_j / sh.stages) & 1);
        // previous use of this slot: item s_i - 2
        if (ready) ready = o_behind ? mbar_try_wait_warp(&o_full[slot], use ^ 1) : mbar_try_wait_warp(&o_empty[slot], use ^ 1);
        if (ready) {
          tc_fence_after();
          const uint32_t q_addr = smem_u32(sStage + st * sh.stage_bytes + s_qt * 128 * HD * 2);
          const uint32_t k_addr = smem_u32(sStage + st * sh.stage_bytes + sh.q_bytes);
          if (elect_one()) {
#pragma unroll
            for (int ks = 0; ks < HD / 16; ++ks)
              umma_bf16(tmem_base + slot * kSlotCols, desc_kmajor<HD>(q_addr, ks), desc_kmajor<HD>(k_addr, ks), idesc_s,
                        ks > 0 ? 1u : 0u);
            umma_commit(&s_full[slot]);
          }
          __syncwarp();
          JM_TRACE(s_i, 12, lane == 0);
          ++s_i;
          if (++s_qt == sh.nq) s_qt = 0, ++s_j;
          progress = true;
        }
      }
      if (p_i < s_i) {
        const int slot = p_i & 1, st = p_j % sh.stages;
        const uint32_t use = (p_i >> 1) & 1;
        // P written, and (when the next scores did not already wait for it) the O columns of item p_i - 2 read out
        bool ready = mbar_try_wait_warp(&p_full[slot], use);
        if (ready && o_behind) ready = mbar_try_wait_warp(&o_empty[slot], use ^ 1);
        if (ready) {
          tc_fence_after();
          JM_TRACE(p_i, 8, lane == 0);
          const uint32_t v_addr = smem_u32(sStage + st * sh.stage_bytes + sh.q_bytes + sh.kv_bytes);
          const uint32_t t_p = tmem_base + slot * kSlotCols;
          const uint64_t d_v = desc_mnmajor<HD>(v_addr, 0);
          constexpr uint64_t kStepMn = (16 * HD * 2) >> 4;
          if (elect_one()) {
            // P: bf16 pairs, 8 columns per 16-key step; half 0 packed its keys at [0, cs / 2), half 1 at [cs, ...)
#pragma unroll
            for (int kk = 0; kk < 16; ++kk)
              if (kk < ksteps)
                umma_bf16_ts(t_p + sh.col_o, t_p + (kk < k_half ? kk * 8 : sh.cs + (kk - k_half) * 8),
                             d_v + kk * kStepMn, idesc_o, kk > 0 ? 1u : 0u);
            umma_commit(&o_full[slot]);
            if (p_qt == sh.nq - 1) umma_commit(&kv_empty[st]);  // every MMA that reads this stage has been issued
          }
          __syncwarp();
          JM_TRACE(p_i, 9, lane == 0);
          ++p_i;
          if (++p_qt == sh.nq) p_qt = 0, ++p_j;
          progress = true;
        }
      }
      if (!progress && clock64() - t_start > 20000000000LL) {
        if (lane == 0) printf("jm: attention forward issue loop timed out (block %d: %d %d of %d)\n", blockIdx.x, s_i, p_i, n_items);
        __trap();
      }
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
#ifdef JM_EXP_FWD_NO_PASS1
      mx = 8.f;
      if (false) {
#else
      if (warp_live) {
#endif
        for (int c = 0; c < n32; ++c) fwd_chunk_max<32>(t_row, h0 + c * 32, sh.T, mx);
        if (tail16) fwd_chunk_max<16>(t_row, h0 + n32 * 32, sh.T, mx);
      }
      *xm_mine = mx;
      JM_TRACE(i, 2, tr);
      named_bar_sync(1 + slot, 256);
      mx = fmaxf(mx, *xm_other);
      JM_TRACE(i, 3, tr);
      float l = 0.f;
#ifdef JM_EXP_FWD_NO_PASS2
      l = 1.f;
      if (false) {
#else
      if (warp_live) {
#endif
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
      // out = O / l as bf16 rows into this slot's shared-memory tile (the layout TMA expects), then ONE TMA store of
      // the 128 x head_dim tile: rows past the sequence end are clipped by the TMA unit.  (Per-thread 16-byte
      // stores of 64-byte row segments cost an L1 tag cycle per segment and were the slowest part of the kernel.)
      const uint32_t o_tile = smem_u32(sO + slot * 128 * HD * 2);
#ifdef JM_EXP_NO_DRAIN
      if (false) {
#else
      if (warp_live) {
#endif
        const float inv = 1.0f / l;
        uint32_t r[OC];
        tmem_ld_w<OC>(t_row + sh.col_o + half * OC, r);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < OC; e += 8)
          sts128(o_tile + rowtile_chunk<HD>(row, half * (OC / 8) + (e >> 3)),
                 pack_bf16(__uint_as_float(r[e]) * inv, __uint_as_float(r[e + 1]) * inv),
                 pack_bf16(__uint_as_float(r[e + 2]) * inv, __uint_as_float(r[e + 3]) * inv),
                 pack_bf16(__uint_as_float(r[e + 4]) * inv, __uint_as_float(r[e + 5]) * inv),
                 pack_bf16(__uint_as_float(r[e + 6]) * inv, __uint_as_float(r[e + 7]) * inv));
        if (half == 0 && tok < sh.T) lse[(static_cast<long>(b) * sh.heads + h) * sh.T + tok] = mx + logf(l);
      }
      tc_fence_before();
      mbar_arrive(&o_empty[slot]);  // O is in registers / shared memory: the slot may take the next scores
      fence_proxy_async();
      named_bar_sync(1 + slot, 256);
#ifdef JM_EXP_NO_DRAIN
      if (false) {
#else
      if (half == 0 && q == 0 && lane == 0) {
#endif
        tma_store_3d(&tmO, o_tile, h * HD, qt * 128, b);
        tma_store_commit();
        tma_store_wait_read<0>();  // the tile is rewritten only after this thread has passed the next barrier
      }
      JM_TRACE(i, 6, tr);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

// ------------------------------------------------------------------------------------------------ backward
// TMEM lanes = keys (thread = one key row), columns = queries.  The unit of work is a HALF-TILE: one 128-key tile kt
// against one block qh of at most 128 queries (the 199 + pad = 208 decoder queries are two blocks of 112 and 96; up
// to 128 queries are one block).  Half-tile i lives in TMEM slot g = i & 1 and is handled by compute group g: eight
// warps, two per TMEM lane quarter, each taking one half of the query columns (no exchange between the halves is
// needed in the backward pass).  The two groups are always in different phases: one runs its exponentials while the
// tensor core works for the other.
//   S^T  = K_kt Q_qh^T   tcgen05.mma  M = 128 keys, N = w queries, K = head_dim     -> slot columns [0, w)
//   dP^T = V_kt dO_qh^T  same shape                                                 -> slot columns [cw, cw + w)
//   P^T = exp2(S^T c - lse c), dS^T = P^T (dP^T - D): the thread walks along its key row in units of 16 queries;
//          lse and D = rowsum(dO . O) per query are broadcast reads from shared memory (produced one problem ahead
//          by warps 2-3).  The packed
//          bf16 pairs overwrite the scores in place (P^T over S^T, dS^T over dP^T, always inside columns the
//          thread itself has already consumed); dS^T also goes to a SWIZZLE_128B shared-memory tile [key][query]
//          (one per slot)
//   dV_kt += P^T dO_qh   A = P^T from tensor memory, B = dO rows in place (MN-major)   accumulates over qh in TMEM
//   dK_kt += dS^T Q_qh   A = dS^T from tensor memory, B = Q rows in place              accumulates over qh in TMEM
//   dQ part = dS K_kt    A = the dS^T tile read MN-major (M = queries), B = K_kt rows  one product per half-tile,
//                        summed over the key tiles in the registers of the thread that reads it (thread = query)
// S and dP are computed once (5 matrix products per half-tile), nothing is recomputed and no gradient needs
// atomics.  Bias gradients: sum_q dQ (column sums once per problem), sum_key dV = sum_q dO (rows of P sum to one),
// sum_key dK = 0 (rows of dS sum to zero).
constexpr int kBwdThreads = 640;  // warp 0 TMA + scores MMA, 1 gradient MMA, 2-3 statistics, 4-19: [group 2][half 2][quarter 4]
constexpr int kBwdComputeWarp0 = 4;

struct AttnTcBwdShape {
  int T, Nq, NH, nkt, heads, batch, stages;
  int w[2], qoff[2], U[2];  // query block: width (multiple of 16), first query, 16-query units
  int Uh[2];                // column half 0 owns the units [0, Uh), half 1 the units [Uh, U) of block qh
  int cw, slot;             // dP^T offset inside a slot; slot size
  int oQ;                   // dQ part: column offset (inside a slot if < slot, else absolute base for slot 0, + hd for 1)
  int cV, cK;               // TMEM columns of dV, dK
  int q_bytes;      // Q (and dO): Nq rows, rounded up to 1024
  int k_bytes;      // K (and V): nkt * 128 rows
  int stage_bytes;  // Q + dO + K + V
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

// One unit of 16 queries of a key row: P^T and dS^T as packed bf16 pairs from the loaded S^T (rs) and dP^T (rp).
// nL2 = -lse * log2(e), nD = -rowsum(dO . O).
__device__ __forceinline__ void bwd_unit(const uint32_t (&rs)[16], const uint32_t (&rp)[16], const float* nL2,
                                         const float* nD, bool key_ok, uint32_t (&wp)[8], uint32_t (&wd)[8]) {
  const f32x2 k2 = pack2(kLog2eTc, kLog2eTc);
#pragma unroll
  for (int e = 0; e < 16; e += 4) {
    const ulonglong2 l4 = *reinterpret_cast<const ulonglong2*>(nL2 + e);  // two f32x2 pairs
    const ulonglong2 d4 = *reinterpret_cast<const ulonglong2*>(nD + e);
    float x0, x1, x2, x3;
    unpack2(fma2(pack2(__uint_as_float(rs[e]), __uint_as_float(rs[e + 1])), k2, l4.x), x0, x1);
    unpack2(fma2(pack2(__uint_as_float(rs[e + 2]), __uint_as_float(rs[e + 3])), k2, l4.y), x2, x3);
    const float p0 = ex2_approx(x0), p1 = ex2_approx(x1), p2 = ex2_approx(x2), p3 = ex2_approx(x3);
    float s0, s1, s2, s3;
    unpack2(mul2(pack2(p0, p1), add2(pack2(__uint_as_float(rp[e]), __uint_as_float(rp[e + 1])), d4.x)), s0, s1);
    unpack2(mul2(pack2(p2, p3), add2(pack2(__uint_as_float(rp[e + 2]), __uint_as_float(rp[e + 3])), d4.y)), s2, s3);
    wp[e >> 1] = pack_bf16(p0, p1), wp[(e >> 1) + 1] = pack_bf16(p2, p3);
    wd[e >> 1] = pack_bf16(s0, s1), wd[(e >> 1) + 1] = pack_bf16(s2, s3);
  }
  if (!key_ok) {  // key rows past the sequence end (zero K / V rows): no probability, no gradient
#pragma unroll
    for (int e = 0; e < 8; ++e) wp[e] = 0u, wd[e] = 0u;
  }
}

template <int HD>
__global__ void __launch_bounds__(kBwdThreads, 1)
attn_tc_bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                   const __grid_constant__ CUtensorMap tmdO, const __grid_constant__ CUtensorMap tmStKV,
                   const __grid_constant__ CUtensorMap tmStQ0, const __grid_constant__ CUtensorMap tmStQ1,
                   const AttnTcBwdShape sh, const __nv_bfloat16* __restrict__ out,
                   const __nv_bfloat16* __restrict__ dout, const float* __restrict__ lse, float* __restrict__ dbias) {
  constexpr int ROWB = HD * 2;
  // Per slot: the dS^T tile [128 keys][128 queries] bf16 (two 64-query blocks, 32 KB).  Once the products of a
  // half-tile have completed the same memory stages the results on their way out: three [128 rows][HD] bf16 tiles
  // (dQ, dV, dK) that leave by TMA stores.
  constexpr int kRowTile = 128 * ROWB;
  constexpr int kTileBytes = 3 * kRowTile > 2 * 16384 ? 3 * kRowTile : 2 * 16384;
  constexpr int kStatBufs = 4;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sStage = smem;
  uint8_t* sDSt = smem + sh.stages * sh.stage_bytes;                 // [2 slots][kTileBytes]
  float* sL2 = reinterpret_cast<float*>(sDSt + 2 * kTileBytes);      // [4][256] -lse * log2(e) per query (-big: padding)
  float* sD = sL2 + kStatBufs * 256;                                 // [4][256] -rowsum(dO . O) per query
  uint64_t* bars = reinterpret_cast<uint64_t*>(sD + kStatBufs * 256);
  uint64_t* kv_full = bars;        // [stages]
  uint64_t* kv_empty = bars + 4;   // [stages]
  uint64_t* s_full = bars + 8;     // [2] S^T and dP^T of the half-tile are in the slot
  uint64_t* p_full = bars + 10;    // [2] P^T, dS^T written (256 arrivals)
  uint64_t* g_full = bars + 12;    // [2] the three gradient products of the half-tile have completed
  uint64_t* s_free = bars + 14;    // [2] the slot's dQ part has been read out (256 arrivals)
  uint64_t* vk_free = bars + 16;   // dV / dK of a key tile have been read out (256 arrivals)
  uint64_t* st_full = bars + 17;   // [4] per-query statistics of a problem are in shared memory (64 arrivals)
  uint64_t* st_empty = bars + 21;  // [4] ... and are no longer needed (one arrival per compute thread that read them)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 25);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmdO);
    tma_prefetch_desc(&tmStKV);
    tma_prefetch_desc(&tmStQ0);
    tma_prefetch_desc(&tmStQ1);
    for (int s = 0; s < sh.stages; ++s) {
      mbar_init(&kv_full[s], 1);
      mbar_init(&kv_empty[s], 1 + 64);  // the last product's commit + the 64 statistics threads (v-bias sums)
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&s_full[s], 1);
      mbar_init(&p_full[s], 256);
      mbar_init(&g_full[s], 1);
      mbar_init(&s_free[s], 256);
    }
    for (int s = 0; s < kStatBufs; ++s) {
      mbar_init(&st_full[s], 64);
      mbar_init(&st_empty[s], sh.nkt * sh.NH >= 2 ? 512 : 256);  // one arrival per group that touches a problem
    }
    mbar_init(vk_free, 256);
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
  const int d = sh.heads * HD;
  const int HT = sh.nkt * sh.NH;  // half-tiles per problem
  const int n_ht = my_prob * HT;
  // TMEM column of slot g's dQ part
  const int cq0 = sh.oQ;
  const int cq1 = sh.oQ < sh.slot ? sh.slot + sh.oQ : sh.oQ + HD;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer + MMA issuer 1 (scores)
    // Operand loads of the problems ahead (as their stages come free), and S^T / dP^T of the half-tiles in order.
    // The whole warp runs the control flow, one elected lane issues.
    int load_j = 0;
    auto load_problem = [&](int j) {
      const int prob = blockIdx.x + j * gridDim.x;
      const int b = prob / sh.heads, h = prob - b * sh.heads;
      const int st = j % sh.stages;
      uint8_t* dst = sStage + st * sh.stage_bytes;
      if (elect_one()) {
        mbar_arrive_expect_tx(&kv_full[st], 2 * sh.Nq * ROWB + 2 * sh.nkt * 128 * ROWB);
        tma_load_3d(&tmQ, &kv_full[st], dst, h * HD, 0, b);
        tma_load_3d(&tmdO, &kv_full[st], dst + sh.q_bytes, h * HD, 0, b);
        tma_load_3d(&tmK, &kv_full[st], dst + 2 * sh.q_bytes, d + h * HD, 0, b);
        tma_load_3d(&tmK, &kv_full[st], dst + 2 * sh.q_bytes + sh.k_bytes, 2 * d + h * HD, 0, b);
      }
      __syncwarp();
    };
    int s_j = 0, s_kt = 0, s_qh = 0;
    for (int s_i = 0; s_i < n_ht; ++s_i) {
      const int g = s_i & 1, st = s_j % sh.stages;
      // loads: every problem whose stage is free, up to `stages` problems ahead; this problem's at the latest now
      while (load_j < my_prob && load_j < s_j + sh.stages) {
        uint64_t* bar = &kv_empty[load_j % sh.stages];
        const uint32_t par = ((load_j / sh.stages) & 1) ^ 1;
        if (load_j <= s_j) mbar_wait(bar, par);
        else if (!mbar_try_wait_warp(bar, par)) break;
        load_problem(load_j++);
      }
      if (s_kt == 0 && s_qh == 0) mbar_wait(&kv_full[st], (s_j / sh.stages) & 1);
      mbar_wait(&s_free[g], ((s_i >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t q_addr = smem_u32(sStage + st * sh.stage_bytes);
      const uint32_t do_addr = q_addr + sh.q_bytes, k_addr = q_addr + 2 * sh.q_bytes, v_addr = k_addr + sh.k_bytes;
      const uint32_t tb = tmem_base + g * sh.slot;
      const uint32_t idesc_s = make_idesc_bf16(128, sh.w[s_qh], 0, 0);  // K-major A and B
      if (elect_one()) {
#pragma unroll
        for (int ks = 0; ks < HD / 16; ++ks)
          umma_bf16(tb, desc_kmajor<HD>(k_addr + s_kt * 128 * ROWB, ks),
                    desc_kmajor<HD>(q_addr + sh.qoff[s_qh] * ROWB, ks), idesc_s, ks > 0 ? 1u : 0u);
#pragma unroll
        for (int ks = 0; ks < HD / 16; ++ks)
          umma_bf16(tb + sh.cw, desc_kmajor<HD>(v_addr + s_kt * 128 * ROWB, ks),
                    desc_kmajor<HD>(do_addr + sh.qoff[s_qh] * ROWB, ks), idesc_s, ks > 0 ? 1u : 0u);
        umma_commit(&s_full[g]);
      }
      __syncwarp();
      JM_TRACE(s_i, 9, lane == 0);
      if (++s_qh == sh.NH) { s_qh = 0; if (++s_kt == sh.nkt) s_kt = 0, ++s_j; }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer 2: gradient products
    // dV, dK and the dQ part of the half-tiles, in order, each once its P^T / dS^T are written.
    const uint32_t idesc_kv = make_idesc_bf16(128, HD, 0, 1);  // A from tensor memory, MN-major B
    const uint32_t idesc_q = make_idesc_bf16(128, HD, 1, 1);   // MN-major A (tile, M = queries), MN-major B
    constexpr uint64_t kStepMn = (16 * ROWB) >> 4;  // descriptor start-address increment per 16-row k-step
    int g_j = 0, g_kt = 0, g_qh = 0, g_tile = 0;
    for (int g_i = 0; g_i < n_ht; ++g_i) {
      const int g = g_i & 1, st = g_j % sh.stages;
      mbar_wait(&p_full[g], (g_i >> 1) & 1);
      if (g_qh == 0) mbar_wait(vk_free, (g_tile & 1) ^ 1);  // previous tile's dV / dK have been read out
      tc_fence_after();
      JM_TRACE(g_i, 12, lane == 0);
      const uint32_t q_addr = smem_u32(sStage + st * sh.stage_bytes);
      const uint32_t do_addr = q_addr + sh.q_bytes, k_addr = q_addr + 2 * sh.q_bytes;
      const uint32_t tb = tmem_base + g * sh.slot;
      const int U = sh.U[g_qh], Uh = sh.Uh[g_qh];  // half 1 packed its unit kk at 8 (kk + Uh)
      const uint64_t d_do = desc_mnmajor<HD>(do_addr + sh.qoff[g_qh] * ROWB, 0);
      const uint64_t d_q = desc_mnmajor<HD>(q_addr + sh.qoff[g_qh] * ROWB, 0);
      const uint64_t d_a = make_smem_desc(smem_u32(sDSt + g * kTileBytes), 16384, 1024, 2);
      const uint64_t d_k = desc_mnmajor<HD>(k_addr + g_kt * 128 * ROWB, 0);
      const uint32_t acc0 = g_qh > 0 ? 1u : 0u;
      const bool last_of_problem = g_kt == sh.nkt - 1 && g_qh == sh.NH - 1;
      if (elect_one()) {
#pragma unroll
        for (int kk = 0; kk < 8; ++kk)  // dV_kt += P^T dO_qh
          if (kk < U)
            umma_bf16_ts(tmem_base + sh.cV, tb + 8 * (kk < Uh ? kk : kk + Uh), d_do + kk * kStepMn, idesc_kv,
                         kk > 0 ? 1u : acc0);
#pragma unroll
        for (int kk = 0; kk < 8; ++kk)  // dK_kt += dS^T Q_qh
          if (kk < U)
            umma_bf16_ts(tmem_base + sh.cK, tb + sh.cw + 8 * (kk < Uh ? kk : kk + Uh), d_q + kk * kStepMn, idesc_kv,
                         kk > 0 ? 1u : acc0);
#pragma unroll
        for (int kk = 0; kk < 8; ++kk)  // dQ part = dS K_kt  (M = the block's queries, K = 128 keys)
          umma_bf16(tmem_base + (g == 0 ? cq0 : cq1), d_a + kk * (2048 >> 4), d_k + kk * kStepMn, idesc_q,
                    kk > 0 ? 1u : 0u);
        umma_commit(&g_full[g]);
        // every MMA that reads this stage has been issued -- by THIS thread; the scores of the problem were issued
        // by warp 1 and completed long ago (their results were consumed before P^T could be written)
        if (last_of_problem) umma_commit(&kv_empty[st]);
      }
      __syncwarp();
      JM_TRACE(g_i, 14, lane == 0);
      if (++g_qh == sh.NH) {
        g_qh = 0, ++g_tile;
        if (++g_kt == sh.nkt) g_kt = 0, ++g_j;
      }
    }
  } else if (warp < kBwdComputeWarp0) {
    // ------------------------------------------------------------------ per-query statistics, up to three problems ahead
    // D = rowsum(dO . O) and -lse log2(e) per query, from global memory with all of a round's loads in flight together
    // (decoupled from the operand stages, so these warps can run several problems ahead of the tensor core).
    const int sid = threadIdx.x - 64;  // 0..63: queries sid + 64 r
    constexpr int CH = HD / 8;         // 16-byte chunks per row
    constexpr int RPR = HD == 32 ? 2 : 1;  // rows per round (register budget)
    int bias_j = 0;  // next problem whose v-bias gradient is to be summed (lags behind the statistics)
    // bias gradient of v: sum over keys of dV = sum over queries of dO (every row of P sums to one).  The dO rows are
    // read from the operand stage; the stage is released only after these threads have arrived too.
    auto bias_of = [&](int j) {
      const int prob = blockIdx.x + j * gridDim.x;
      const int h = prob % sh.heads;
      const int st = j % sh.stages;
      mbar_wait(&kv_full[st], (j / sh.stages) & 1);
      if (dbias != nullptr) {
        // thread = (column pair cp, row slice rs): HD / 2 column pairs x 128 / HD slices of the Nq rows
        const uint32_t do_tile = smem_u32(sStage + st * sh.stage_bytes + sh.q_bytes);
        const int cp = sid % (HD / 2), rs = sid / (HD / 2), nrs = 64 / (HD / 2);
        float s0 = 0.f, s1 = 0.f;
        for (int r = rs; r < sh.Nq; r += nrs) {
          uint32_t v;
          asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(do_tile + rowtile_chunk<HD>(r, cp >> 2) + (cp & 3) * 4));
          const float2 f = unpack_bf16(v);
          s0 += f.x, s1 += f.y;
        }
        atomicAdd(dbias + 2 * d + h * HD + 2 * cp, s0);
        atomicAdd(dbias + 2 * d + h * HD + 2 * cp + 1, s1);
      }
      mbar_arrive(&kv_empty[st]);
    };
    for (int j = 0; j < my_prob; ++j) {
      const int prob = blockIdx.x + j * gridDim.x;
      const int b = prob / sh.heads, h = prob - b * sh.heads;
      const int sb = j % kStatBufs;
      // never more than stages - 1 problems ahead of the bias sums: their arrival is what frees an operand stage
      while (bias_j < j && (bias_j + sh.stages <= j || mbar_test_wait(&kv_full[bias_j % sh.stages], (bias_j / sh.stages) & 1)))
        bias_of(bias_j++);
      mbar_wait(&st_empty[sb], ((j / kStatBufs) & 1) ^ 1);
#pragma unroll 1
      for (int r0 = 0; r0 < 4; r0 += RPR) {
        if (sid + 64 * r0 >= sh.T) {  // nothing but padding from here on
          for (int r = r0; r < 4; ++r) sL2[sb * 256 + sid + 64 * r] = -1e30f, sD[sb * 256 + sid + 64 * r] = 0.f;
          break;
        }
        uint4 ov[RPR][CH], gv[RPR][CH];
        float lv[RPR];
#pragma unroll
        for (int r = 0; r < RPR; ++r) {
          const int qi = sid + 64 * (r0 + r);
          lv[r] = 1e30f;
#pragma unroll
          for (int c = 0; c < CH; ++c) ov[r][c] = gv[r][c] = make_uint4(0u, 0u, 0u, 0u);
#ifndef JM_EXP_NO_STATS
          if (qi < sh.T) {
            const long tokr = static_cast<long>(b) * sh.T + qi;
            const __nv_bfloat16* orow = out + tokr * sh.ldo + h * HD;
            const __nv_bfloat16* grow = dout + tokr * sh.lddo + h * HD;
#pragma unroll
            for (int c = 0; c < CH; ++c) {
              ov[r][c] = *reinterpret_cast<const uint4*>(orow + c * 8);
              gv[r][c] = *reinterpret_cast<const uint4*>(grow + c * 8);
            }
            lv[r] = lse[(static_cast<long>(b) * sh.heads + h) * sh.T + qi] * kLog2eTc;
          }
#endif
        }
#pragma unroll
        for (int r = 0; r < RPR; ++r) {
          const int qi = sid + 64 * (r0 + r);
          float Dv = 0.f;
#pragma unroll
          for (int c = 0; c < CH; ++c) {
            const uint32_t ow[4] = {ov[r][c].x, ov[r][c].y, ov[r][c].z, ov[r][c].w};
            const uint32_t gw[4] = {gv[r][c].x, gv[r][c].y, gv[r][c].z, gv[r][c].w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float2 a = unpack_bf16(ow[e]), g2 = unpack_bf16(gw[e]);
              Dv = fmaf(a.x, g2.x, Dv);
              Dv = fmaf(a.y, g2.y, Dv);
            }
          }
          sL2[sb * 256 + qi] = -lv[r], sD[sb * 256 + qi] = -Dv;
        }
      }
      mbar_arrive(&st_full[sb]);
    }
    while (bias_j < my_prob) bias_of(bias_j++);
  } else {
    // ------------------------------------------------------------------ compute groups
    constexpr int DC = HD / 2;                    // dQ columns per thread
    const int sw = warp - kBwdComputeWarp0;
    const int g = sw >> 3;                        // group == TMEM slot == parity of the half-tile
    const int half = (sw >> 2) & 1;               // which half of the query columns / of the result columns
    const int q = warp & 3;                       // TMEM lane quarter
    const int krow = q * 32 + lane;               // key row (softmax) / query row (dQ) inside the 128-row tile
    const uint32_t t_lane = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const uint32_t tb = t_lane + g * sh.slot;
    const uint32_t t_dq = t_lane + (g == 0 ? cq0 : cq1) + half * DC;
    const uint32_t tile = smem_u32(sDSt + g * kTileBytes);
    const bool elected = half == 0 && q == 0 && lane == 0;  // issues this group's TMA stores
    bool store_pending = false;
    float dq[DC];  // this thread's half of the dQ row of its query, summed over the key tiles
#pragma unroll
    for (int c = 0; c < DC; ++c) dq[c] = 0.f;
    // (problem, key tile, query block) of this group's half-tiles i = g, g + 2, ...
    int j = 0, kt = 0, qh = 0;
    auto advance = [&](int steps) {
      for (int s = 0; s < steps; ++s)
        if (++qh == sh.NH) { qh = 0; if (++kt == sh.nkt) kt = 0, ++j; }
    };
    advance(g);
    for (int i = g; i < n_ht; i += 2) {
      const int sb = j % kStatBufs;
      const int prob = blockIdx.x + j * gridDim.x;
      const int b = prob / sh.heads, h = prob - b * sh.heads;
      const bool key_ok = kt * 128 + krow < sh.T;
      const bool warp_live = kt * 128 + q * 32 < sh.T;
      const bool tr = g == 0 && half == 0 && q == 0 && lane == 0;
      JM_TRACE(i, 0, tr);
      mbar_wait(&st_full[sb], (j / kStatBufs) & 1);
      mbar_wait(&s_full[g], (i >> 1) & 1);
      tc_fence_after();
      JM_TRACE(i, 1, tr);
      // the TMA stores of this group's previous results have finished reading the tile memory this half-tile's
      // dS^T is about to overwrite
      if (store_pending) {
        if (elected) tma_store_wait_read<0>();
        named_bar_sync(1 + g, 256);
        store_pending = false;
      }
      const float* l2p = sL2 + sb * 256 + sh.qoff[qh];
      const float* dp = sD + sb * 256 + sh.qoff[qh];
      const int ub = half == 0 ? 0 : sh.Uh[qh];                         // first unit of this half
      const int nu = half == 0 ? sh.Uh[qh] : sh.U[qh] - sh.Uh[qh];      // its units (at most 4)
#ifdef JM_EXP_NO_SOFTMAX
      if (false) {
#else
      if (warp_live) {
#endif
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (k < nu) {
            const int u = ub + k;
            uint32_t rs[16], rp[16];
            tmem_ld_32x16(tb + u * 16, rs);
            tmem_ld_32x16(tb + sh.cw + u * 16, rp);
            tmem_ld_wait();
            uint32_t wp[8], wd[8];
            bwd_unit(rs, rp, l2p + u * 16, dp + u * 16, key_ok, wp, wd);
            // packed pairs in place, inside this half's own columns: unit u -> [8 (u + ub), + 8), which this thread
            // has consumed
            tmem_st_32x8(tb + 8 * (u + ub), wp);
            tmem_st_32x8(tb + sh.cw + 8 * (u + ub), wd);
            sts128(tile + ptile_chunk(krow, 2 * u), wd[0], wd[1], wd[2], wd[3]);
            sts128(tile + ptile_chunk(krow, 2 * u + 1), wd[4], wd[5], wd[6], wd[7]);
          }
        }
        tmem_st_wait();
      } else {
        // dead key rows: zeros in the shared-memory tile (they are contracted over in dQ)
        for (int k = 0; k < nu; ++k) {
          sts128(tile + ptile_chunk(krow, 2 * (ub + k)), 0u, 0u, 0u, 0u);
          sts128(tile + ptile_chunk(krow, 2 * (ub + k) + 1), 0u, 0u, 0u, 0u);
        }
      }
      tc_fence_before();
      fence_proxy_async();
      mbar_arrive(&p_full[g]);
      JM_TRACE(i, 5, tr);
      // this problem's statistics are read for the last time by each group at ITS last half-tile of the problem
      if (HT == 1 || kt * sh.NH + qh == HT - 2 + g) mbar_arrive(&st_empty[sb]);
      // ---- the products behind this half-tile
      mbar_wait(&g_full[g], (i >> 1) & 1);
      tc_fence_after();
      JM_TRACE(i, 6, tr);
      // dQ part of query block qh: thread = query row, this half's DC columns; summed over the key tiles in registers.
      // Results leave through [128 rows][HD] tiles in this slot's (now idle) tile memory and TMA stores: per-thread
      // 16-byte stores of 64-byte row segments cost an L1 tag cycle each and were the slowest part of the kernel.
#ifdef JM_EXP_NO_DRAIN
      const bool q_live = false;
      const bool store_q = false, store_vk = false;
      if (qh == sh.NH - 1) mbar_arrive(vk_free);
#else
      const bool q_live = q * 32 < sh.w[qh];
      const bool store_q = kt == sh.nkt - 1, store_vk = qh == sh.NH - 1;
#endif
      if (q_live) {
        uint32_t r[DC];
        tmem_ld_w<DC>(t_dq, r);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < DC; ++e) dq[e] += __uint_as_float(r[e]);
        if (store_q) {
          const bool q_ok = krow < sh.w[qh] && sh.qoff[qh] + krow < sh.T;
          float v[32];
#pragma unroll
          for (int e = 0; e < 32; ++e) v[e] = (e < DC && q_ok) ? dq[e < DC ? e : 0] * sh.q_scale : 0.f;
#pragma unroll
          for (int e = 0; e < DC; e += 8)
            sts128(tile + rowtile_chunk<HD>(krow, half * (DC / 8) + (e >> 3)), pack_bf16(v[e], v[e + 1]),
                   pack_bf16(v[e + 2], v[e + 3]), pack_bf16(v[e + 4], v[e + 5]), pack_bf16(v[e + 6], v[e + 7]));
          if (dbias != nullptr) {
            const float cs = warp_colsum32(v, lane);
            if (lane < DC) atomicAdd(dbias + h * HD + half * DC + lane, cs);
          }
#pragma unroll
          for (int c = 0; c < DC; ++c) dq[c] = 0.f;
        }
      }
      if (store_vk && warp_live) {
        // dV / dK of key tile kt are complete (this half-tile's products were the last to add to them):
        // half 0 stages dV, half 1 dK; thread = key row
        const uint32_t vk_tile = tile + (1 + half) * kRowTile;
#pragma unroll
        for (int c = 0; c < HD / 32; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(t_lane + (half == 0 ? sh.cV : sh.cK) + c * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int e = 0; e < 32; e += 8)
            sts128(vk_tile + rowtile_chunk<HD>(krow, c * 4 + (e >> 3)),
                   pack_bf16(__uint_as_float(r[e]), __uint_as_float(r[e + 1])),
                   pack_bf16(__uint_as_float(r[e + 2]), __uint_as_float(r[e + 3])),
                   pack_bf16(__uint_as_float(r[e + 4]), __uint_as_float(r[e + 5])),
                   pack_bf16(__uint_as_float(r[e + 6]), __uint_as_float(r[e + 7])));
        }
      }
      tc_fence_before();
      if (store_vk) mbar_arrive(vk_free);
      mbar_arrive(&s_free[g]);  // everything this slot held is in registers / shared memory
      if (store_q || store_vk) {
        fence_proxy_async();
        named_bar_sync(1 + g, 256);
        if (elected) {
          // rows past the sequence end (and, for dQ, past this query block: the box is w rows) are clipped by TMA
          if (store_q) tma_store_3d(qh == 0 ? &tmStQ0 : &tmStQ1, tile, h * HD, sh.qoff[qh], b);
          if (store_vk) {
            tma_store_3d(&tmStKV, tile + kRowTile, 2 * d + h * HD, kt * 128, b);
            tma_store_3d(&tmStKV, tile + 2 * kRowTile, d + h * HD, kt * 128, b);
          }
          tma_store_commit();
        }
        store_pending = true;
      }
      advance(2);
    }
    if (store_pending && elected) tma_store_wait_read<0>();
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
  const int fixed = 1024 + 256 + 2 * 512 * 4 + 2 * 128 * HD * 2;
  sh.stages = (kMaxDynSmem - fixed) / sh.stage_bytes;
  if (sh.stages > 3) sh.stages = 3;
  JM_CHECK(sh.stages >= 1, "jm_attention_fwd: T=%d hd=%d does not fit shared memory", T, HD);
  const int smem = fixed + sh.stages * sh.stage_bytes;
  CUtensorMap tmQ, tmKV;
  const long d = static_cast<long>(heads) * HD;
  if (make_tmap_heads(&tmQ, qkv, 3 * d, T, batch, ldqkv, HD, 128)) return 1;
  if (make_tmap_heads(&tmKV, qkv, 3 * d, T, batch, ldqkv, HD, sh.Nk)) return 1;
  CUtensorMap tmO;
  if (make_tmap_heads(&tmO, out, d, T, batch, ldo, HD, 128)) return 1;
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
  JM_LAUNCH_PDL(kern, grid, kFwdThreads, smem, stream, tmQ, tmKV, tmO, sh, lse);
  JM_AFTER_LAUNCH();
  return 0;
}

// Tensor-memory and column-group layout of the backward kernel; false when the problem does not fit.
bool bwd_layout(AttnTcBwdShape* sh, int T, int hd) {
  sh->T = T, sh->Nq = (T + 15) / 16 * 16, sh->nkt = (T + 127) / 128;
  sh->NH = sh->Nq > 128 ? 2 : 1;
  sh->w[0] = sh->NH == 1 ? sh->Nq : (sh->Nq / 2 + 15) / 16 * 16;
  sh->w[1] = sh->Nq - sh->w[0];
  sh->qoff[0] = 0, sh->qoff[1] = sh->w[0];
  sh->U[0] = sh->w[0] / 16, sh->U[1] = sh->w[1] / 16;
  sh->cw = sh->w[0];       // dP^T behind the widest S^T
  sh->slot = 2 * sh->cw;
  sh->cV = 2 * sh->slot, sh->cK = sh->cV + hd;
  if (sh->cK + hd + 2 * hd <= 512) {
    sh->oQ = sh->cK + hd;  // a dQ part of its own per slot
    for (int qh = 0; qh < 2; ++qh) sh->Uh[qh] = (sh->U[qh] + 1) / 2;
    return true;
  }
  // otherwise the dQ part goes into the gap column half 0 leaves in the slot's (consumed) dP^T columns: half 0 owns
  // hd / 8 units = 2 hd columns, packs into the first hd of them, and the dQ part takes the second hd
  const int uh = hd / 8;
  for (int qh = 0; qh < sh->NH; ++qh) {
    if (sh->U[qh] < uh || sh->U[qh] - uh > 4) return false;
    sh->Uh[qh] = uh;
  }
  sh->oQ = sh->cw + hd;
  return uh <= 4 && sh->cK + hd <= 512 && sh->oQ + hd <= sh->slot;
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
  sh.ldo = ldo, sh.lddo = lddo, sh.lddqkv = lddqkv, sh.q_scale = q_scale;
  const int row_tile = 128 * HD * 2;
  const int tile_bytes = 3 * row_tile > 2 * 16384 ? 3 * row_tile : 2 * 16384;
  const int fixed = 1024 + 2 * tile_bytes + 2 * 4 * 256 * 4 + 256;
  sh.stages = (kMaxDynSmem - fixed) / sh.stage_bytes;
  if (sh.stages > 3) sh.stages = 3;
  JM_CHECK(sh.stages >= 1, "jm_attention_bwd: T=%d hd=%d does not fit shared memory", T, HD);
  const int smem = fixed + sh.stages * sh.stage_bytes;
  CUtensorMap tmQ, tmK, tmdO;
  const long d = static_cast<long>(heads) * HD;
  if (make_tmap_heads(&tmQ, qkv, 3 * d, T, batch, ldqkv, HD, sh.Nq)) return 1;
  if (make_tmap_heads(&tmK, qkv, 3 * d, T, batch, ldqkv, HD, sh.nkt * 128)) return 1;
  if (make_tmap_heads(&tmdO, dout, d, T, batch, lddo, HD, sh.Nq)) return 1;
  CUtensorMap tmStKV, tmStQ0, tmStQ1;  // result stores into dqkv: 128-key tiles, and one query block each
  if (make_tmap_heads(&tmStKV, dqkv, 3 * d, T, batch, lddqkv, HD, 128)) return 1;
  if (make_tmap_heads(&tmStQ0, dqkv, 3 * d, T, batch, lddqkv, HD, sh.w[0])) return 1;
  if (make_tmap_heads(&tmStQ1, dqkv, 3 * d, T, batch, lddqkv, HD, sh.NH == 2 ? sh.w[1] : sh.w[0])) return 1;
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
  JM_LAUNCH_PDL(kern, grid, kBwdThreads, smem, stream, tmQ, tmK, tmdO, tmStKV, tmStQ0, tmStQ1, sh,
                reinterpret_cast<const __nv_bfloat16*>(out), reinterpret_cast<const __nv_bfloat16*>(dout), lse, dbias);
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
