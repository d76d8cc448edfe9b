// tcgen05 / TMEM / TMA GEMM for sm_100a with fused epilogues.
//
// One persistent CTA per SM, 18 warps (576 threads: 112 registers each), warp-specialised:
//   warp 0       TMA producer   (one elected lane; cp.async.bulk.tensor into a SWIZZLE_128B ring)
//   warp 1       MMA issuer     (one elected lane; tcgen05.mma 128 x BN x 16, accumulators in TMEM) and TMEM
//                allocator (2 x BN f32 columns: the accumulator is double-buffered so the epilogue of
//                tile i overlaps the MMAs of tile i+1)
//   warps 2..17  epilogue       (tcgen05.ld -> registers -> fused epilogue -> staging transpose -> global)
// Pipelines: smem full/empty mbarriers (TMA <-> MMA), TMEM full/empty mbarriers (MMA <-> epilogue).
//
// Operand majors: both A and B may be K-major ([rows, K] row-major) or MN-major ([K, rows]
// row-major); this covers Y = X W (B MN-major: a Flax kernel is [in, out]), dX = dY W^T (B K-major)
// and dW = X^T dY (both MN-major) without any transposed copies.
//
// Replaces the XLA dot_general behind every Dense/DenseGeneral/Conv of the reference
// (src/modeling.py:108-148, src/pretraining.py:84-85) and their autodiff transposes.
#include <cudaTypedefs.h>
#include <stdlib.h>

#include "../../include/jumbo_mae_b200.h"
#include "common.cuh"

namespace jm {

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = 128 bytes = one SWIZZLE_128B row
constexpr int UMMA_K = 16;
constexpr int kEpiWarp0 = 4;  // warps 0..3 form the warpgroup that hands its registers to the epilogue warps
constexpr int kEpiWarps = 16;  // four per TMEM lane quarter, each takes a quarter of the tile's columns
constexpr int kGemmThreads = (kEpiWarp0 + kEpiWarps) * 32;
// 640 threads launch with 96 registers each (five warps per scheduler share 16 K registers).  Warps 0..3
// only move data: they shrink, and every scheduler's four epilogue warps grow by what its one warp released
// (the pool is the CTA's own allocation: 4 x 32 x (INC - 96) <= 32 x (96 - DEC)).
#ifndef JM_GEMM_REG_DEC
#define JM_GEMM_REG_DEC 32
#define JM_GEMM_REG_INC 112
#endif
static_assert(4 * (JM_GEMM_REG_INC - 96) <= 96 - JM_GEMM_REG_DEC, "setmaxnreg pool exceeded");
constexpr int kStageRowBytes = 64;                     // 16 f32 columns
constexpr int kStageBytesPerWarp = 32 * kStageRowBytes;  // 32 rows

struct GemmShape {
  int M, N, K;
  int num_m, num_n, splits, kblocks;
  int group_m;  // tile order: n fastest within bands of group_m tile rows, m fastest inside a band (1: plain n fastest)
};

// Tile index within one split -> tile coordinates.
__device__ __forceinline__ void tile_mn(const GemmShape& sh, int rem, int& m_blk, int& n_blk) {
  if (sh.group_m <= 1) {
    m_blk = rem / sh.num_n, n_blk = rem - m_blk * sh.num_n;
    return;
  }
  const int per = sh.group_m * sh.num_n;
  const int g = rem / per, r = rem - g * per;
  const int m0 = g * sh.group_m;
  const int gm = min(sh.group_m, sh.num_m - m0);
  n_blk = r / gm, m_blk = m0 + (r - n_blk * gm);
}

struct GemmMaps {
  CUtensorMap a, b, o, o2;  // operands; bf16 outputs of the row-register epilogues (TMA stores)
};

struct EpiParams {
  float* out_f32;
  long ldo_f32;
  __nv_bfloat16* out_bf16;
  long ldo_bf16;
  __nv_bfloat16* out2_bf16;
  long ldo2_bf16;
  const float* bias;
  const float* residual;
  long ldr;
  const __nv_bfloat16* aux;
  long ldaux;
  const float* pe;
  const int* pe_idx;
  long ldpe;
  int act;
  float col_scale;
  int col_scale_n;
  int row_period, row_skip, row_mask_mode;
  int accumulate;
  float* colsum;  // optional [N]: += column sums of the stored values (bias gradient of the producer Dense)
  int tma_store;  // row-register epilogues: write the bf16 outputs with TMA stores (maps o / o2)
  const float* row_scale;  // optional [M / row_scale_period]: (acc + bias) of a row is multiplied by its image's entry
  int row_scale_period;    // before the residual is added (DropPath keep-mask / (1 - p), src/modeling.py:157,181-183)
  // loss epilogue (act 5): d = acc + bias - target; out_bf16 = d * loss_scale; row_loss[row, col / 16] = sum d^2
  const uint8_t* loss_target;  // [M, ld] raw pixel bytes in patch-vector order (channel = col % 3)
  long ld_target;
  const float* loss_stats;     // optional [M, 2]: norm_pix mean and 1 / sqrt(var + eps) of the row's normalised target
  float* row_loss;             // [M, ceil(N / 16)]
  float loss_scale;
  long pe_idx_stride;          // elements between two images' rows of pe_idx (0: one index row shared by the batch)
  const float* col_vec;        // optional [N]: (acc + bias) is multiplied by col_vec[col] before row_scale and the residual
                               // (LayerScale, src/modeling.py:159-166,185-201); f32 TMA epilogue only
};

template <int BN>
struct GemmCfg {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  // staging tiles per epilogue warp: two where a pipeline stage can be spared for them (TMA stores of one
  // tile then overlap the arithmetic of the next), one for the 48 KB stages of the 128 x 256 tile
  static constexpr int EPI_BUFS = (BN == 256) ? 1 : 2;
  static constexpr int STAGES = (BN == 256) ? 4 : ((BN == 128) ? 5 : 6);
  static constexpr int TMEM_COLS = 2 * BN;
  static constexpr int BAR_BYTES = (2 * STAGES + 4) * 8 + 16 + kEpiWarps * 2 * 8;  // + residual-load barriers
  static constexpr int EPI_BYTES = kEpiWarps * kStageBytesPerWarp * EPI_BUFS;  // epilogue transpose staging
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + BAR_BYTES + 1024;  // +1024: alignment
};

__device__ __forceinline__ void store_f32x4(float* p, float4 v, bool atomic) {
  if (atomic) {
    atomicAdd(reinterpret_cast<float4*>(p), v);
  } else {
    *reinterpret_cast<float4*>(p) = v;
  }
}

__device__ __forceinline__ float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// GELU-tanh (src/modeling.py:148) and its derivative with the hardware tanh (abs. error ~5e-4,
// below the bf16 resolution of the values they are stored in).
__device__ __forceinline__ float gelu_fast(float x) {
  const float x2 = x * x;
  const float u = x * fmaf(x2, 0.0356774081f, 0.7978845608f);  // sqrt(2/pi) * (x + 0.044715 x^3)
  const float hx = 0.5f * x;
  return fmaf(hx, tanh_fast(u), hx);
}
__device__ __forceinline__ float gelu_grad_fast(float x) {
  const float x2 = x * x;
  const float t = tanh_fast(x * fmaf(x2, 0.0356774081f, 0.7978845608f));
  const float du = fmaf(x2, 0.1070322243f, 0.7978845608f);      // d/dx of the tanh argument
  const float a = (0.5f * x) * fmaf(-t, t, 1.0f);
  return fmaf(a, du, fmaf(0.5f, t, 0.5f));
}

// gelu_fast / gelu_grad_fast on a pair (identical arithmetic, element-wise)
__device__ __forceinline__ f32x2 gelu_fast2(f32x2 x) {
  const f32x2 c1 = pack2(0.0356774081f, 0.0356774081f), c0 = pack2(0.7978845608f, 0.7978845608f);
  const f32x2 half = pack2(0.5f, 0.5f);
  const f32x2 u = mul2(x, fma2(mul2(x, x), c1, c0));
  float u0, u1;
  unpack2(u, u0, u1);
  const f32x2 t = pack2(tanh_fast(u0), tanh_fast(u1));
  const f32x2 hx = mul2(x, half);
  return fma2(hx, t, hx);
}
__device__ __forceinline__ f32x2 gelu_grad_fast2(f32x2 x) {
  const f32x2 c1 = pack2(0.0356774081f, 0.0356774081f), c0 = pack2(0.7978845608f, 0.7978845608f);
  const f32x2 c3 = pack2(0.1070322243f, 0.1070322243f), half = pack2(0.5f, 0.5f), one = pack2(1.0f, 1.0f);
  const f32x2 x2 = mul2(x, x);
  const f32x2 u = mul2(x, fma2(x2, c1, c0));
  float u0, u1;
  unpack2(u, u0, u1);
  const float t0 = tanh_fast(u0), t1 = tanh_fast(u1);
  const f32x2 t = pack2(t0, t1), nt = pack2(-t0, -t1);
  const f32x2 du = fma2(x2, c3, c0);
  const f32x2 a = mul2(mul2(x, half), fma2(nt, t, one));
  return fma2(a, du, fma2(half, t, half));
}

// gelu(x) and gelu'(x) of a pair from one tanh: with s = (1 + t) / 2, gelu = x s and gelu' = s + x s (1 - t) du.
__device__ __forceinline__ void gelu_both2(f32x2 x, f32x2& g, f32x2& d) {
  const f32x2 c1 = pack2(0.0356774081f, 0.0356774081f), c0 = pack2(0.7978845608f, 0.7978845608f);
  const f32x2 c3 = pack2(0.1070322243f, 0.1070322243f), half = pack2(0.5f, 0.5f), one = pack2(1.0f, 1.0f);
  const f32x2 x2 = mul2(x, x);
  const f32x2 u = mul2(x, fma2(x2, c1, c0));
  float u0, u1;
  unpack2(u, u0, u1);
  const float t0 = tanh_fast(u0), t1 = tanh_fast(u1);
  const f32x2 t = pack2(t0, t1), nt = pack2(-t0, -t1);
  const f32x2 s = fma2(half, t, half);
  g = mul2(x, s);
  d = fma2(mul2(g, fma2(nt, one, one)), fma2(x2, c3, c0), s);
}
__device__ __forceinline__ float gelu_grad_of(float x) { return gelu_grad_fast(x); }

// Epilogue specialisations: straight-line code per kind (the generic one keeps every runtime switch).
enum : int {
  EPI_BF16 = 0,      // out_bf16 = (acc + bias) * scale
  EPI_F32 = 1,       // out_f32 (=|+=) acc + bias [+ pe] [+ residual]          (residual prefetched)
  EPI_GELU_FWD = 2,  // out_bf16 = h = acc + bias, out2_bf16 = gelu(h)
  EPI_GELU_BWD = 3,  // out_bf16 = acc * gelu'(aux)                             (aux prefetched)
  EPI_GENERIC = 4,
  EPI_LOSS = 5,      // out_bf16 = (acc + bias - target) * scale, row_loss = per-row partial sums of the squared error
};

template <int EPI>
struct EpiPrefetch {  // global operand of one lane for its 2 rows x 8 columns of a 32x16 sub-chunk
};
template <>
struct EpiPrefetch<EPI_F32> {
  float4 v[2][2];
};
template <>
struct EpiPrefetch<EPI_GELU_BWD> {
  uint4 v[2];
};

template <bool FULL>
__device__ __forceinline__ void store_bf16x8(__nv_bfloat16* p, const float (&v)[8], int ncols) {
  if (FULL || ncols >= 8) {
    *reinterpret_cast<uint4*>(p) =
        make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
  } else {
    *reinterpret_cast<uint2*>(p) = make_uint2(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]));
  }
}

// Fused epilogue for 8 consecutive columns [col, col+8) of one row (FULL: all 8 in range, else ncols is 8
// or 4 at a ragged edge), in the coalesced layout: 4 lanes cover the 32 columns of a chunk row.  sc/bs hold
// the per-column scale and the pre-scaled bias (v = acc * sc + bs), hoisted per chunk; `affine` says whether
// they are non-trivial.  `zero` rows store zeros.  On return v holds the stored values (for the column sums).
template <int EPI, bool FULL>
__device__ __forceinline__ void epilogue8(const EpiParams& ep, float (&v)[8], const float (&sc)[8],
                                          const float (&bs)[8], bool affine, long off_row, int row, int col,
                                          int ncols, bool zero, const EpiPrefetch<EPI>& pf, int it) {
  if (zero) {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = 0.f;
    if (ep.out_bf16) store_bf16x8<FULL>(ep.out_bf16 + off_row * ep.ldo_bf16 + col, v, ncols);
    if (ep.out2_bf16) store_bf16x8<FULL>(ep.out2_bf16 + off_row * ep.ldo2_bf16 + col, v, ncols);
    if (ep.out_f32 && !ep.accumulate) {
      float* o = ep.out_f32 + off_row * ep.ldo_f32 + col;
      *reinterpret_cast<float4*>(o) = make_float4(0.f, 0.f, 0.f, 0.f);
      if (FULL || ncols >= 8) *reinterpret_cast<float4*>(o + 4) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    return;
  }
  if (affine) {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = fmaf(v[j], sc[j], bs[j]);
  }
  if constexpr (EPI == EPI_BF16) {
    store_bf16x8<FULL>(ep.out_bf16 + off_row * ep.ldo_bf16 + col, v, ncols);
  } else if constexpr (EPI == EPI_F32) {
    if (ep.pe) {
      const float* pe = ep.pe + static_cast<long>(ep.pe_idx[(row / ep.row_period) * ep.pe_idx_stride + row % ep.row_period - ep.row_skip]) * ep.ldpe + col;
      const float4 a = *reinterpret_cast<const float4*>(pe);
      v[0] += a.x, v[1] += a.y, v[2] += a.z, v[3] += a.w;
      if (FULL || ncols >= 8) {
        const float4 b = *reinterpret_cast<const float4*>(pe + 4);
        v[4] += b.x, v[5] += b.y, v[6] += b.z, v[7] += b.w;
      }
    }
    const float4 r0 = pf.v[it][0], r1 = pf.v[it][1];
    v[0] += r0.x, v[1] += r0.y, v[2] += r0.z, v[3] += r0.w;
    v[4] += r1.x, v[5] += r1.y, v[6] += r1.z, v[7] += r1.w;
    float* o = ep.out_f32 + off_row * ep.ldo_f32 + col;
    store_f32x4(o, make_float4(v[0], v[1], v[2], v[3]), ep.accumulate != 0);
    if (FULL || ncols >= 8) store_f32x4(o + 4, make_float4(v[4], v[5], v[6], v[7]), ep.accumulate != 0);
  } else if constexpr (EPI == EPI_GELU_FWD) {
    float g[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) g[j] = gelu_fast(v[j]);
    store_bf16x8<FULL>(ep.out_bf16 + off_row * ep.ldo_bf16 + col, v, ncols);
    store_bf16x8<FULL>(ep.out2_bf16 + off_row * ep.ldo2_bf16 + col, g, ncols);
  } else if constexpr (EPI == EPI_GELU_BWD) {
    const uint4 u = pf.v[it];
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 a = unpack_bf16(w[j]);
      v[2 * j] *= ep.act == 4 ? a.x : gelu_grad_fast(a.x);
      v[2 * j + 1] *= ep.act == 4 ? a.y : gelu_grad_fast(a.y);
    }
    store_bf16x8<FULL>(ep.out_bf16 + off_row * ep.ldo_bf16 + col, v, ncols);
  } else {
    if (ep.act == 1 || ep.act == 3) {
      float g[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) g[j] = gelu_fast(v[j]);
      if (ep.act == 3) {  // the derivative is saved instead of the pre-activation
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = gelu_grad_fast(v[j]);
      }
      if (ep.out_bf16) store_bf16x8<FULL>(ep.out_bf16 + off_row * ep.ldo_bf16 + col, v, ncols);
      if (ep.out2_bf16) store_bf16x8<FULL>(ep.out2_bf16 + off_row * ep.ldo2_bf16 + col, g, ncols);
      return;
    }
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      if (hh * 4 >= ncols) break;
      const int c4 = col + hh * 4;
      float* x = v + hh * 4;
      if (ep.act == 2 || ep.act == 4) {
        const uint2 u = *reinterpret_cast<const uint2*>(ep.aux + off_row * ep.ldaux + c4);
        const float2 a = unpack_bf16(u.x), b = unpack_bf16(u.y);
        if (ep.act == 2) {
          x[0] *= gelu_grad_fast(a.x), x[1] *= gelu_grad_fast(a.y);
          x[2] *= gelu_grad_fast(b.x), x[3] *= gelu_grad_fast(b.y);
        } else {
          x[0] *= a.x, x[1] *= a.y, x[2] *= b.x, x[3] *= b.y;
        }
      }
      if (ep.pe) {
        const float4 b = *reinterpret_cast<const float4*>(
            ep.pe + static_cast<long>(ep.pe_idx[(row / ep.row_period) * ep.pe_idx_stride + row % ep.row_period - ep.row_skip]) * ep.ldpe + c4);
        x[0] += b.x, x[1] += b.y, x[2] += b.z, x[3] += b.w;
      }
      if (ep.residual) {
        const float4 b = *reinterpret_cast<const float4*>(ep.residual + off_row * ep.ldr + c4);
        x[0] += b.x, x[1] += b.y, x[2] += b.z, x[3] += b.w;
      }
      if (ep.out_f32)
        store_f32x4(ep.out_f32 + off_row * ep.ldo_f32 + c4, make_float4(x[0], x[1], x[2], x[3]), ep.accumulate != 0);
    }
    if (ep.out_bf16) store_bf16x8<FULL>(ep.out_bf16 + off_row * ep.ldo_bf16 + col, v, ncols);
  }
}

// Issues the global loads one lane needs for one 32x16 sub-chunk: the f32 residual (EPI_F32) or the
// bf16 pre-activation (EPI_GELU_BWD).  Issued before the TMEM load / shared-memory transpose of the
// sub-chunk (and, for a tile's first sub-chunk, before waiting for the accumulator), so that the HBM
// latency overlaps with them and with the other epilogue warps of the scheduler.
template <int EPI>
__device__ __forceinline__ void epi_prefetch(EpiPrefetch<EPI>& pf, const EpiParams& ep, const GemmShape& sh,
                                             int row0, int col) {
  if constexpr (EPI == EPI_F32 || EPI == EPI_GELU_BWD) {
    const int ncols = sh.N - col;
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const int row = row0 + it * 16;
      const bool ok = ncols >= 4 && row < sh.M;
      if constexpr (EPI == EPI_F32) {
        pf.v[it][0] = make_float4(0.f, 0.f, 0.f, 0.f);
        pf.v[it][1] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ok && ep.residual) {
          const float* r = ep.residual + static_cast<long>(row) * ep.ldr + col;
          pf.v[it][0] = *reinterpret_cast<const float4*>(r);
          if (ncols >= 8) pf.v[it][1] = *reinterpret_cast<const float4*>(r + 4);
        }
      } else {
        pf.v[it] = make_uint4(0u, 0u, 0u, 0u);
        if (ok) {
          const __nv_bfloat16* a = ep.aux + static_cast<long>(row) * ep.ldaux + col;
          if (ncols >= 8) {
            pf.v[it] = *reinterpret_cast<const uint4*>(a);
          } else {
            const uint2 u = *reinterpret_cast<const uint2*>(a);
            pf.v[it].x = u.x, pf.v[it].y = u.y;
          }
        }
      }
    }
  }
}

// Epilogue warps (16: four per TMEM lane quarter, each draining a quarter of the tile's columns).
// TMEM hands each thread one accumulator ROW; global memory wants lanes on consecutive COLUMNS.  Each warp
// transposes 32x16 f32 sub-chunks through a private, XOR-swizzled 2 KB staging tile (conflict-free both
// ways), then runs the fused epilogue with 2 lanes covering the 64 contiguous bytes of a sub-chunk row
// (8 columns per lane, 16 rows per warp instruction): every global access is a full 16-byte vector.
// Sixteen warps, not four or eight: one warp retires a long dependent instruction stream per sub-chunk at
// well under one instruction per cycle, so the drain rate scales with the number of warps per scheduler.
template <int EPI, int BN, int CG, int NBUF>
__device__ __forceinline__ void epilogue_role(const EpiParams& ep, const GemmShape& sh, uint8_t* sEpi,
                                              uint64_t* tfull, uint64_t* tempty, uint32_t tmem_base, int warp,
                                              int lane, int total_tiles, int tile0, int tile_step, int m_off) {
  const int e = warp - kEpiWarp0;
  const int q = warp & 3;   // TMEM lane quarter == warp % 4 (a hardware rule); (q, part) still enumerates the 16 warps
  const int part = e >> 2;  // which quarter of the tile's columns
  constexpr int PART_N = BN / 4;
  constexpr int NSUB = PART_N / 16;
  constexpr bool kColsum = (EPI == EPI_GELU_BWD || EPI == EPI_GENERIC);
  const uint32_t stage_addr = smem_u32(sEpi + e * (kStageBytesPerWarp * NBUF));
  const int rsub = lane >> 1, cu = lane & 1;
  const int tiles_mn = sh.num_m * sh.num_n;
  int iter = 0;
  for (int t = tile0; t < total_tiles; t += tile_step, ++iter) {
    const int rem = t % tiles_mn;
    int m_blk, n_blk;
    tile_mn(sh, rem, m_blk, n_blk);
    const int acc = iter & 1;
    const uint32_t acc_phase = (iter >> 1) & 1;
    // this lane's 2 rows are the same for every sub-chunk of the tile: resolve the row predicates once
    const int row0 = m_blk * (BM * CG) + m_off + q * 32 + rsub;
    uint32_t store_mask = 0, zero_mask = 0;
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const int row = row0 + it * 16;
      bool st = row < sh.M, zr = false;
      if (ep.row_period > 0 && (row % ep.row_period) < ep.row_skip) {
        if (ep.row_mask_mode == 2) st = false;
        if (ep.row_mask_mode == 1) zr = true;
      }
      store_mask |= (st ? 1u : 0u) << it;
      zero_mask |= (zr ? 1u : 0u) << it;
    }
    const int colp = n_blk * BN + part * PART_N;  // first column of this warp's part
    EpiPrefetch<EPI> pf;
    epi_prefetch<EPI>(pf, ep, sh, row0, colp + cu * 8);
    mbar_wait(&tfull[acc], acc_phase);
    tc_fence_after();
    const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN + part * PART_N;
#pragma unroll
    for (int c = 0; c < NSUB; ++c) {
      const int col0 = colp + c * 16;
      const bool live = col0 < sh.N;  // warp-uniform
      if (live) {
        uint32_t r[16];
        tmem_ld_32x16(t_addr + c * 16, r);
        tmem_ld_wait();
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint32_t a = stage_addr + lane * kStageRowBytes + ((u ^ ((lane >> 1) & 3)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(r[4 * u]), "r"(r[4 * u + 1]),
                       "r"(r[4 * u + 2]), "r"(r[4 * u + 3])
                       : "memory");
        }
      }
      if (c == NSUB - 1) {  // accumulator fully drained: hand the TMEM buffer back to the MMA warp
        tc_fence_before();
        if constexpr (CG == 1) mbar_arrive(&tempty[acc]);
        else mbar_arrive_cluster(mapa_shared(smem_u32(&tempty[acc]), cluster_ctarank() & ~1u));  // the leader CTA issues the MMAs
      }
      __syncwarp();
      const EpiPrefetch<EPI> cur = pf;
      if (c + 1 < NSUB) epi_prefetch<EPI>(pf, ep, sh, row0, col0 + 16 + cu * 8);
      const int col = col0 + cu * 8;
      const int ncols = sh.N - col;
      float cs[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      if (live && ncols >= 4) {
        float sc[8], bs[8];
        const bool scaled = col < ep.col_scale_n;
        const bool affine = ep.bias != nullptr || col0 < ep.col_scale_n;
        if (affine) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            sc[j] = (scaled && col + j < ep.col_scale_n) ? ep.col_scale : 1.f;
            bs[j] = 0.f;
          }
          if (ep.bias) {
            const float4 b0 = *reinterpret_cast<const float4*>(ep.bias + col);
            bs[0] = b0.x * sc[0], bs[1] = b0.y * sc[1], bs[2] = b0.z * sc[2], bs[3] = b0.w * sc[3];
            if (ncols >= 8) {
              const float4 b1 = *reinterpret_cast<const float4*>(ep.bias + col + 4);
              bs[4] = b1.x * sc[4], bs[5] = b1.y * sc[5], bs[6] = b1.z * sc[6], bs[7] = b1.w * sc[7];
            }
          }
        }
        const bool full = col0 + 16 <= sh.N;  // warp-uniform: the whole sub-chunk is inside the matrix
#pragma unroll
        for (int it = 0; it < 2; ++it) {
          const int rloc = it * 16 + rsub;
          float v[8];
          const uint32_t sw = (rloc >> 1) & 3;
          const uint32_t a0 = stage_addr + rloc * kStageRowBytes + (((2 * cu) ^ sw) << 4);
          const uint32_t a1 = stage_addr + rloc * kStageRowBytes + (((2 * cu + 1) ^ sw) << 4);
          asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                       : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3])
                       : "r"(a0));
          asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                       : "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
                       : "r"(a1));
          if (store_mask & (1u << it)) {
            const int row = row0 + it * 16;
            if (full)
              epilogue8<EPI, true>(ep, v, sc, bs, affine, static_cast<long>(row), row, col, 8,
                                   (zero_mask >> it) & 1u, cur, it);
            else
              epilogue8<EPI, false>(ep, v, sc, bs, affine, static_cast<long>(row), row, col, ncols,
                                    (zero_mask >> it) & 1u, cur, it);
            if constexpr (kColsum) {
#pragma unroll
              for (int j = 0; j < 8; ++j) cs[j] += v[j];
            }
          }
        }
      }
      if constexpr (kColsum) {
        if (ep.colsum != nullptr && live) {  // warp-uniform: fold the 16 row-lanes, one atomic per column
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            cs[j] += __shfl_xor_sync(0xffffffffu, cs[j], 2);
            cs[j] += __shfl_xor_sync(0xffffffffu, cs[j], 4);
            cs[j] += __shfl_xor_sync(0xffffffffu, cs[j], 8);
            cs[j] += __shfl_xor_sync(0xffffffffu, cs[j], 16);
          }
          if (rsub == 0 && ncols >= 4) {
#pragma unroll
            for (int j = 0; j < 8; ++j)
              if (j < ncols) atomicAdd(ep.colsum + col + j, cs[j]);
          }
        }
      }
      __syncwarp();
    }
  }
}

// Opaque copies: ptxas otherwise re-derives loop-invariant addresses (generic->shared conversion, 64-bit
// row * stride products) at every use instead of keeping them in a register.
__device__ __forceinline__ uint32_t opaque(uint32_t x) {
#ifdef JM_OPAQUE
  asm volatile("mov.b32 %0, %0;" : "+r"(x));
#endif
  return x;
}
__device__ __forceinline__ uint64_t opaque(uint64_t x) {
#ifdef JM_OPAQUE
  asm volatile("mov.b64 %0, %0;" : "+l"(x));
#endif
  return x;
}
// 16-byte global store under a predicate (no branch around it)
__device__ __forceinline__ void stg128_if(uint64_t addr, const uint4& v, uint32_t flag) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %5, 0;\n\t"
      "@p st.global.v4.b32 [%0], {%1, %2, %3, %4};\n\t"
      "}\n" ::"l"(addr),
      "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"(flag)
      : "memory");
}
// 16-byte shared store of v, or of zeros when `zero` is set (rows that store zeros), without a branch
__device__ __forceinline__ void sts128_or_zero(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d,
                                               uint32_t zero) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %5, 0;\n\t"
      "@!p st.shared.v4.b32 [%0], {%1, %2, %3, %4};\n\t"
      "@p st.shared.v4.b32 [%0], {%6, %6, %6, %6};\n\t"
      "}\n" ::"r"(addr),
      "r"(a), "r"(b), "r"(c), "r"(d), "r"(zero), "r"(0u)
      : "memory");
}

// Optional cycle accounting of the row-register epilogue (build with -DJM_EPI_PROF; block 0, first epilogue warp
// prints one line per launch): where an epilogue warp's time goes.
#ifdef JM_EPI_PROF
#define EPI_T(var) const long long var = clock64()
#define EPI_ACC(sum, t0) sum += clock64() - (t0)
#else
#define EPI_T(var)
#define EPI_ACC(sum, t0)
#endif

// Row-register epilogue for the kinds whose only per-element output is bf16 (EPI_BF16, EPI_GELU_FWD,
// EPI_GELU_BWD).  The arithmetic runs in the TMEM layout (one thread = one accumulator row, 32 columns at a
// time), the results are packed to bf16 and only then transposed through the warp's 2 KB staging tile
// (32 rows x 64 bytes, XOR-swizzled 16-byte chunks), so that
//   * a warp store instruction covers 8 rows x 64 contiguous bytes (8 L1 tags instead of the 32 of a
//     32 x 16 sub-chunk: the store rate, not DRAM, bounded the old path),
//   * the dependent chain per 32 columns is TMEM load -> FMA/pack -> STS -> LDS -> STG with no global load
//     in it: the bias of the warp's columns sits one value per lane in a register (loaded before the
//     accumulator wait) and is broadcast with shuffles; the GELU-backward pre-activations are loaded
//     coalesced before the wait and handed to the row threads through the same staging tile,
//   * the second TMEM load of a 64-column part is in flight while the first half is stored.
template <int EPI, int BN, int CG, int NBUF>
__device__ __forceinline__ void epilogue_rows(const EpiParams& ep, const GemmShape& sh, const CUtensorMap* tmO,
                                              const CUtensorMap* tmO2, uint8_t* sEpi, uint64_t* rbars,
                                              uint64_t* tfull, uint64_t* tempty, uint32_t tmem_base, int warp,
                                              int lane, int total_tiles, int tile0, int tile_step, int m_off) {
  constexpr int NH = (BN == 256) ? 2 : 1;   // 32-column halves per warp
  constexpr int PARTS = BN / (32 * NH);     // column parts per tile: 4, 4, 2 (BN = 64: half the warps idle)
  constexpr uint32_t kFull = 0xffffffffu;
  const int e = warp - kEpiWarp0;
  const int q = warp & 3;   // TMEM lane quarter == warp % 4 (a hardware rule); (q, part) still enumerates the 16 warps
  const int part = e >> 2;
  const uint32_t stage_addr = smem_u32(sEpi + e * (kStageBytesPerWarp * NBUF));
  // row view: lane = row, chunk j at ((j ^ sw) << 4); coalesced view: 4 lanes per row, 8 rows per instruction
  const uint32_t own_sw = (lane >> 1) & 3;
  const uint32_t own_addr = opaque(stage_addr + lane * 64);
  const int crow = lane >> 2, cch = lane & 3;
  const uint32_t co_addr = opaque(stage_addr + crow * 64 + ((cch ^ ((crow >> 1) & 3)) << 4));  // + i * 512: row 8 i + crow
  const int tiles_mn = sh.num_m * sh.num_n;
  const bool has_bias = ep.bias != nullptr;
  const bool tma = ep.tma_store != 0;  // kernel-uniform
  const bool saved_grad = ep.act >= 3;  // acts 3 / 4: the GELU derivative is stored / consumed directly
  uint32_t buf = 0;                     // byte offset of the staging tile in use (NBUF tiles, used in turn)
  // GELU backward with TMA stores: the pre-activation (or saved-derivative) tile of each 32-column half arrives in
  // the staging tile by a TMA load through map o2, issued one half ahead (no global-load instructions at all)
  const bool tma_aux = (EPI == EPI_GELU_BWD) && tma;
  uint64_t* rbar = rbars + e * 2;
  uint32_t rphase = 0, rissued = 0;     // per staging tile: parity to wait for / a load is in flight
  auto issue_aux = [&](uint32_t bi, int col, int row) {  // warp-uniform
    if (col < sh.N) {
      if (lane == 0) {
        mbar_arrive_expect_tx(&rbar[bi], kStageBytesPerWarp);
        tma_load_2d(tmO2, &rbar[bi], sEpi + e * (kStageBytesPerWarp * NBUF) + bi * kStageBytesPerWarp, col, row);
      }
      rissued |= 1u << bi;
    }
  };
  auto half_origin = [&](int t, int& row_base, int& colp) {
    const int rem = t % tiles_mn;
    int m_blk, n_blk;
    tile_mn(sh, rem, m_blk, n_blk);
    row_base = m_blk * (BM * CG) + m_off + q * 32;
    colp = n_blk * BN + part * (32 * NH);
  };
  if (tma_aux && part < PARTS && tile0 < total_tiles) {
    int rb, cp;
    half_origin(tile0, rb, cp);
    issue_aux(0, cp, rb);
  }
#ifdef JM_EPI_PROF
  long long pt_tfull = 0, pt_ld = 0, pt_math = 0, pt_flush = 0, pt_free = 0, pt_setup = 0, pt_sts = 0, pt_rel = 0, pt_ldi = 0, pt_half = 0;
  const long long pt_begin = clock64();
#endif
  int iter = 0;
  for (int t = tile0; t < total_tiles; t += tile_step, ++iter) {
    const int rem = t % tiles_mn;
    int m_blk, n_blk;
    tile_mn(sh, rem, m_blk, n_blk);
    const int acc = iter & 1;
    const uint32_t acc_phase = (iter >> 1) & 1;
    auto release = [&]() {
      EPI_T(pr0);
      tc_fence_before();
      if constexpr (CG == 1) mbar_arrive(&tempty[acc]);
      else mbar_arrive_cluster(mapa_shared(smem_u32(&tempty[acc]), cluster_ctarank() & ~1u));  // the leader CTA issues the MMAs
      EPI_ACC(pt_rel, pr0);
    };
    if constexpr (PARTS < 4) {
      if (part >= PARTS) {  // keep in lock-step with the accumulator hand-over
        mbar_wait(&tfull[acc], acc_phase);
        tc_fence_after();
        release();
        continue;
      }
    }
    EPI_T(psu);
    const int row_base = m_blk * (BM * CG) + m_off + q * 32;
    const int colp = n_blk * BN + part * (32 * NH);
    uint32_t zflag = 0;  // set for a row that stores zeros (row_mask_mode 1)
    uint32_t st_mask = 0;
    {
      const bool masked = ep.row_period > 0;
      if (masked && ep.row_mask_mode == 1 && ((row_base + lane) % ep.row_period) < ep.row_skip) zflag = 1u;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int row = row_base + 8 * i + crow;
        bool st = row < sh.M;
        if (masked && ep.row_mask_mode == 2 && (row % ep.row_period) < ep.row_skip) st = false;
        st_mask |= (st ? 1u : 0u) << i;
      }
    }
    // this lane's output addresses (row row_base + crow, first column of the part + its 16-byte chunk); row
    // 8 i + crow is i * step further, the second 32-column half 64 bytes further
    const long row0 = row_base + crow;
    const uint64_t o1 = opaque(reinterpret_cast<uint64_t>(ep.out_bf16 + row0 * ep.ldo_bf16 + colp + 8 * cch));
    const uint64_t step1 = opaque(static_cast<uint64_t>(ep.ldo_bf16) * 16);
    uint64_t o2 = 0, step2 = 0;
    if constexpr (EPI == EPI_GELU_FWD) {
      o2 = opaque(reinterpret_cast<uint64_t>(ep.out2_bf16 + row0 * ep.ldo2_bf16 + colp + 8 * cch));
      step2 = opaque(static_cast<uint64_t>(ep.ldo2_bf16) * 16);
    }
    float bs[NH], sc[NH];
    uint4 ax[4];
    auto load_aux = [&](int colh) {  // this lane's 4 coalesced 16-byte chunks of the pre-activation
      const int col = colh + 8 * cch;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        ax[i] = make_uint4(0u, 0u, 0u, 0u);
        if (((st_mask >> i) & 1u) && col < sh.N)
          ax[i] = *reinterpret_cast<const uint4*>(ep.aux + static_cast<long>(row_base + 8 * i + crow) * ep.ldaux + col);
      }
    };
#pragma unroll
    for (int h = 0; h < NH; ++h) {
      const int colh = colp + 32 * h;
      sc[h] = colh < ep.col_scale_n ? ep.col_scale : 1.f;
      bs[h] = (has_bias && colh + lane < sh.N) ? __ldg(ep.bias + colh + lane) * sc[h] : 0.f;
    }
    if constexpr (EPI == EPI_GELU_BWD) {
      if (!tma_aux) load_aux(colp);
    }
    EPI_ACC(pt_setup, psu);
    EPI_T(pt0);
    mbar_wait(&tfull[acc], acc_phase);
    EPI_ACC(pt_tfull, pt0);
    tc_fence_after();
    const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN + part * (32 * NH);
    uint32_t r[NH][32];
    EPI_T(pl0);
    if (colp < sh.N) tmem_ld_32x32(t_addr, r[0]);  // warp-uniform
    EPI_ACC(pt_ldi, pl0);
    EPI_T(pt1);
    tmem_ld_wait();
    EPI_ACC(pt_ld, pt1);
    if constexpr (NH == 1) release();
    EPI_T(ph0);
#pragma unroll
    for (int h = 0; h < NH; ++h) {
      const int colh = colp + 32 * h;
      const bool live = colh < sh.N;  // warp-uniform
      const int col = colh + 8 * cch;
      const uint32_t ok = col < sh.N ? st_mask : 0u;
      // The staging tile goes to global memory either as one TMA store (a SWIZZLE_64B 32 x 32 box: exactly
      // the staging layout; rows and columns beyond the matrix are clipped by the tensor map) or, when rows
      // must be skipped, with 16-byte stores in the coalesced view.
      auto flush = [&](int o) {
        EPI_T(pf0);
        if (tma) {
          fence_proxy_async();  // this lane's staging writes -> visible to the async proxy
          __syncwarp();
#ifdef JM_EXP_NO_STORE
          if (lane == 0 && colh < 0) {
#else
          if (lane == 0) {
#endif
            tma_store_2d(o == 0 ? tmO : tmO2, stage_addr + buf, colh, row_base);
            tma_store_commit();
          }
        } else {
          __syncwarp();
          const uint64_t obase = ((o == 0) ? o1 : o2) + 64 * h;
          const uint64_t ostep = (o == 0) ? step1 : step2;
          uint4 v[4];
#pragma unroll
          for (int i = 0; i < 4; ++i)
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                         : "=r"(v[i].x), "=r"(v[i].y), "=r"(v[i].z), "=r"(v[i].w)
                         : "r"(co_addr + buf + i * 512)
                         : "memory");
#pragma unroll
          for (int i = 0; i < 4; ++i) stg128_if(obase + i * ostep, v[i], ok & (1u << i));
        }
        EPI_ACC(pt_flush, pf0);
      };
      // next staging tile (with two tiles, the store issued from the other one may still be reading it)
      auto next_buf = [&]() {
        if constexpr (NBUF == 2) buf ^= kStageBytesPerWarp;
      };
      // the previous contents of the staging tile have left it (TMA: the store has read them)
      auto staging_free = [&]() {
        EPI_T(ps0);
        if (tma) {
          if (lane == 0) tma_store_wait_read<NBUF - 1>();
        }
        __syncwarp();
        EPI_ACC(pt_free, ps0);
      };
      uint32_t G[16];
      if constexpr (EPI == EPI_GELU_BWD) {
        if (tma_aux) {
          const uint32_t bi = buf / kStageBytesPerWarp;
          if (rissued & (1u << bi)) {  // warp-uniform: this half's tile has landed in staging tile bi
            mbar_wait(&rbar[bi], (rphase >> bi) & 1u);
            rphase ^= 1u << bi;
            rissued &= ~(1u << bi);
          }
          if constexpr (NBUF == 2) {
            // fetch the NEXT half's tile into the other staging tile, once the store that last used it has read it
            if (lane == 0) tma_store_wait_read<0>();
            __syncwarp();
            const uint32_t nb = live ? (bi ^ 1u) : bi;  // (a half beyond N uses no tile and does not toggle)
            if (NH == 2 && h == 0) {
              issue_aux(nb, colp + 32, row_base);
            } else if (t + tile_step < total_tiles) {
              int rb, cp;
              half_origin(t + tile_step, rb, cp);
              issue_aux(nb, cp, rb);
            }
          }
        }
      }
      if (live) {
        if constexpr (EPI == EPI_GELU_BWD) {
          // pre-activations: (coalesced registers ->) staging -> row threads; results overwrite them in place
          if (!tma_aux) {
            staging_free();
#pragma unroll
            for (int i = 0; i < 4; ++i)
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(co_addr + buf + i * 512), "r"(ax[i].x),
                           "r"(ax[i].y), "r"(ax[i].z), "r"(ax[i].w)
                           : "memory");
            __syncwarp();
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint32_t a = own_addr + buf + ((j ^ own_sw) << 4);
            uint32_t w[4], o[4];
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                         : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3])
                         : "r"(a)
                         : "memory");
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const float2 x = unpack_bf16(w[k]);
              const f32x2 a = pack2(__uint_as_float(r[h][8 * j + 2 * k]), __uint_as_float(r[h][8 * j + 2 * k + 1]));
              float v0, v1;
              // aux holds the pre-activation (act 2) or the saved derivative itself (act 4, kernel-uniform)
              unpack2(mul2(a, saved_grad ? pack2(x.x, x.y) : gelu_grad_fast2(pack2(x.x, x.y))), v0, v1);
              o[k] = pack_bf16(v0, v1);
            }
            sts128_or_zero(a, o[0], o[1], o[2], o[3], zflag);
          }
        } else {
          // branch-free (one basic block for the 32 columns); rows that store zeros are handled at the STS
          EPI_T(pm0);
          uint32_t P[16];
          f32x2 v[16];  // column pairs
          const f32x2 sc2 = pack2(sc[h], sc[h]);
          if (has_bias) {  // warp-uniform
#pragma unroll
            for (int j = 0; j < 16; ++j)
              v[j] = fma2(pack2(__uint_as_float(r[h][2 * j]), __uint_as_float(r[h][2 * j + 1])), sc2,
                          pack2(__shfl_sync(kFull, bs[h], 2 * j), __shfl_sync(kFull, bs[h], 2 * j + 1)));
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              v[j] = mul2(pack2(__uint_as_float(r[h][2 * j]), __uint_as_float(r[h][2 * j + 1])), sc2);
          }
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            float a, b;
            unpack2(v[j], a, b);
            P[j] = pack_bf16(a, b);
            if constexpr (EPI == EPI_GELU_FWD) {
              if (saved_grad) {  // act 3 (kernel-uniform): out <- gelu'(v) instead of v
                f32x2 gg, dd;
#ifdef JM_EXP_NO_GELU
                gg = v[j], dd = mul2(v[j], v[j]);
#else
                gelu_both2(v[j], gg, dd);
#endif
                unpack2(dd, a, b);
                P[j] = pack_bf16(a, b);
                unpack2(gg, a, b);
                G[j] = pack_bf16(a, b);
              } else {
                unpack2(gelu_fast2(v[j]), a, b);
                G[j] = pack_bf16(a, b);
              }
            }
          }
          EPI_ACC(pt_math, pm0);
          staging_free();
          EPI_T(pq0);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            sts128_or_zero(own_addr + buf + ((j ^ own_sw) << 4), P[4 * j], P[4 * j + 1], P[4 * j + 2], P[4 * j + 3], zflag);
          EPI_ACC(pt_sts, pq0);
          if constexpr (EPI == EPI_GELU_FWD) {
            flush(0);
            next_buf();
          }
        }
      }
      if constexpr (NH == 2) {
        if (h == 0 && colp + 32 < sh.N) {  // both in flight during the stores below
          if constexpr (EPI == EPI_GELU_BWD) {
            if (!tma_aux) load_aux(colp + 32);
          }
          EPI_T(pl1);
          tmem_ld_32x32(t_addr + 32, r[1]);
          EPI_ACC(pt_ldi, pl1);
        }
      }
      if (live) {
        if constexpr (EPI == EPI_GELU_FWD) {
          staging_free();
#pragma unroll
          for (int j = 0; j < 4; ++j)
            sts128_or_zero(own_addr + buf + ((j ^ own_sw) << 4), G[4 * j], G[4 * j + 1], G[4 * j + 2], G[4 * j + 3], zflag);
          flush(1);
        } else {
          flush(0);
        }
        const uint32_t sum_buf = buf;  // (GELU backward: the tile just stored, for the column sums)
        next_buf();
        if constexpr (EPI == EPI_GELU_BWD) {
          if (ep.colsum != nullptr) {  // warp-uniform: column sums of the stored tile, one vector reduction per 4
            float cs[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              uint4 v;
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                           : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                           : "r"(co_addr + sum_buf + i * 512)
                           : "memory");
              if (ok & (1u << i)) {
                const float2 a = unpack_bf16(v.x), b = unpack_bf16(v.y), c = unpack_bf16(v.z), d = unpack_bf16(v.w);
                cs[0] += a.x, cs[1] += a.y, cs[2] += b.x, cs[3] += b.y;
                cs[4] += c.x, cs[5] += c.y, cs[6] += d.x, cs[7] += d.y;
              }
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              cs[j] += __shfl_xor_sync(kFull, cs[j], 4);
              cs[j] += __shfl_xor_sync(kFull, cs[j], 8);
              cs[j] += __shfl_xor_sync(kFull, cs[j], 16);
            }
            if (crow == 0 && col < sh.N) {
              atomicAdd(reinterpret_cast<float4*>(ep.colsum + col), make_float4(cs[0], cs[1], cs[2], cs[3]));
              atomicAdd(reinterpret_cast<float4*>(ep.colsum + col + 4), make_float4(cs[4], cs[5], cs[6], cs[7]));
            }
          }
        }
      }
      if constexpr (EPI == EPI_GELU_BWD && NBUF == 1) {
        if (tma_aux) {  // one staging tile: the next half's tile can only be fetched once this store has read it
          if (lane == 0) tma_store_wait_read<0>();
          __syncwarp();
          if (NH == 2 && h == 0) {
            issue_aux(0, colp + 32, row_base);
          } else if (t + tile_step < total_tiles) {
            int rb, cp;
            half_origin(t + tile_step, rb, cp);
            issue_aux(0, cp, rb);
          }
        }
      }
      if constexpr (NH == 2) {
        if (h == 0) {
          EPI_T(pt2);
          tmem_ld_wait();
          EPI_ACC(pt_ld, pt2);
          release();
        }
      }
    }
    EPI_ACC(pt_half, ph0);
  }
  if (tma) {  // the last stores must have read the staging tile (and landed) before the CTA may go away
    if (lane == 0) tma_store_wait_all();
  }
  __syncwarp();
#ifdef JM_EPI_PROF
  if (blockIdx.x == 0 && e == 0 && lane == 0) {
    const long long tot = clock64() - pt_begin;
    printf("epi prof: tiles %d total %lld  setup %lld  tfull %lld  ld_issue %lld  tmem_ld %lld  release %lld  math %lld  sts %lld  flush %lld  staging_free %lld  half_loop %lld  rest %lld\n", iter, tot,
           pt_setup, pt_tfull, pt_ldi, pt_ld, pt_rel, pt_math, pt_sts, pt_flush, pt_free, pt_half,
           tot - pt_setup - pt_tfull - pt_ldi - pt_ld - pt_rel - pt_math - pt_sts - pt_flush - pt_free);
  }
#endif
}

// f32 epilogue through TMA (EPI_F32 without positional embedding / skipped rows): out (=|+=) acc + bias
// [+ residual].  Per warp and 16-column chunk: the residual tile (32 rows x 64 bytes) arrives in the staging
// tile by a TMA load that was issued one chunk earlier, every thread adds its accumulator row in place, and
// the tile leaves by a TMA store -- or a TMA reduce-add for gradient accumulation / split-K, which replaces
// the read-modify-write (and the atomics) by a reduction in L2.  No global load/store instruction is issued.
template <int BN, int CG, int NBUF>
__device__ __forceinline__ void epilogue_f32_tma(const EpiParams& ep, const GemmShape& sh, const CUtensorMap* tmO,
                                                 const CUtensorMap* tmR, uint8_t* sEpi, uint64_t* rbars,
                                                 uint64_t* tfull, uint64_t* tempty, uint32_t tmem_base, int warp,
                                                 int lane, int total_tiles, int tile0, int tile_step, int m_off) {
  constexpr uint32_t kFull = 0xffffffffu;
  constexpr int PART_N = BN / 4;
  constexpr int NC = PART_N / 16;
  constexpr int NB = (PART_N + 31) / 32;  // bias registers per lane
  const int e = warp - kEpiWarp0;
  const int q = warp & 3;
  const int part = e >> 2;
  const uint32_t stage_addr = smem_u32(sEpi + e * (kStageBytesPerWarp * NBUF));
  const uint32_t own_sw = (lane >> 1) & 3;
  const uint32_t own_addr = stage_addr + lane * 64;
  uint64_t* rbar = rbars + e * 2;
  const bool has_res = ep.residual != nullptr, has_bias = ep.bias != nullptr, reduce = ep.accumulate != 0;
  const int tiles_mn = sh.num_m * sh.num_n;
  uint32_t buf = 0;      // staging tile in use (index)
  uint32_t rphase = 0;   // bit b: parity to wait for on rbar[b]
  uint32_t rissued = 0;  // bit b: a residual load into tile b is in flight
  auto tile_origin = [&](int t, int& row_base, int& colp) {
    const int rem = t % tiles_mn;
    int m_blk, n_blk;
    tile_mn(sh, rem, m_blk, n_blk);
    row_base = m_blk * (BM * CG) + m_off + q * 32;
    colp = n_blk * BN + part * PART_N;
  };
  auto load_res = [&](uint32_t b, int col, int row) {  // warp-uniform call
    if (has_res && col < sh.N) {
      if (lane == 0) {
        mbar_arrive_expect_tx(&rbar[b], kStageBytesPerWarp);
        tma_load_2d(tmR, &rbar[b], sEpi + e * (kStageBytesPerWarp * NBUF) + b * kStageBytesPerWarp, col, row);
      }
      rissued |= 1u << b;
    }
  };
  if (tile0 < total_tiles) {
    int rb, cp;
    tile_origin(tile0, rb, cp);
    load_res(0, cp, rb);
  }
  int iter = 0;
  for (int t = tile0; t < total_tiles; t += tile_step, ++iter) {
    int row_base, colp;
    tile_origin(t, row_base, colp);
    const int acc = iter & 1;
    const uint32_t acc_phase = (iter >> 1) & 1;
    float bs[NB];
#pragma unroll
    for (int i = 0; i < NB; ++i) {
      const int col = colp + 32 * i + lane;
      bs[i] = (has_bias && col < sh.N && 32 * i + lane < PART_N) ? __ldg(ep.bias + col) : 0.f;
    }
    const bool has_cv = ep.col_vec != nullptr;
    float gs[NB];
#pragma unroll
    for (int i = 0; i < NB; ++i) {
      const int col = colp + 32 * i + lane;
      gs[i] = (has_cv && col < sh.N && 32 * i + lane < PART_N) ? __ldg(ep.col_vec + col) : 1.f;
    }
    // optional per-image scale of the branch output (this thread's row), applied before the residual is added
    const bool has_rs = ep.row_scale != nullptr;
    float rs = 1.f;
    if (has_rs) {
      const int row = row_base + lane;
      rs = row < sh.M ? __ldg(ep.row_scale + row / ep.row_scale_period) : 0.f;
    }
    mbar_wait(&tfull[acc], acc_phase);
    tc_fence_after();
    const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN + part * PART_N;
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      const int col0 = colp + 16 * c;
      const bool live = col0 < sh.N;  // warp-uniform
      uint32_t r[16];
      if (live) tmem_ld_32x16(t_addr + 16 * c, r);
      const uint32_t b = buf;
      if (rissued & (1u << b)) {  // warp-uniform: the residual tile has landed
        mbar_wait(&rbar[b], (rphase >> b) & 1u);
        rphase ^= 1u << b;
        rissued &= ~(1u << b);
      }
      tmem_ld_wait();
      if (c == NC - 1) {  // accumulator fully drained: hand the TMEM buffer back to the MMA warp
        tc_fence_before();
        if constexpr (CG == 1) mbar_arrive(&tempty[acc]);
        else mbar_arrive_cluster(mapa_shared(smem_u32(&tempty[acc]), cluster_ctarank() & ~1u));
      }
      if (live) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t a = own_addr + b * kStageBytesPerWarp + ((j ^ own_sw) << 4);
          float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
          if (has_res)
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                         : "=f"(x.x), "=f"(x.y), "=f"(x.z), "=f"(x.w)
                         : "r"(a)
                         : "memory");
          float4 v = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]), __uint_as_float(r[4 * j + 2]),
                                 __uint_as_float(r[4 * j + 3]));
          if (has_bias) {  // warp-uniform
            const float bsrc = bs[(16 * c) / 32];
            const int l0 = (16 * c) % 32 + 4 * j;
            v.x += __shfl_sync(kFull, bsrc, l0), v.y += __shfl_sync(kFull, bsrc, l0 + 1);
            v.z += __shfl_sync(kFull, bsrc, l0 + 2), v.w += __shfl_sync(kFull, bsrc, l0 + 3);
          }
          if (has_cv) {  // warp-uniform
            const float gsrc = gs[(16 * c) / 32];
            const int l0 = (16 * c) % 32 + 4 * j;
            v.x *= __shfl_sync(kFull, gsrc, l0), v.y *= __shfl_sync(kFull, gsrc, l0 + 1);
            v.z *= __shfl_sync(kFull, gsrc, l0 + 2), v.w *= __shfl_sync(kFull, gsrc, l0 + 3);
          }
          if (has_rs) v.x *= rs, v.y *= rs, v.z *= rs, v.w *= rs;  // warp-uniform branch
          x.x += v.x, x.y += v.y, x.z += v.z, x.w += v.w;
          asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(x.x), "f"(x.y), "f"(x.z), "f"(x.w)
                       : "memory");
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          if (reduce) tma_reduce_add_2d(tmO, stage_addr + b * kStageBytesPerWarp, col0, row_base);
          else tma_store_2d(tmO, stage_addr + b * kStageBytesPerWarp, col0, row_base);
          tma_store_commit();
        }
      }
      if constexpr (NBUF == 2) buf ^= 1u;
      // the tile used next: its previous store must have read it; then its residual can be fetched
      if (lane == 0) tma_store_wait_read<NBUF - 1>();
      __syncwarp();
      if (c + 1 < NC) {
        load_res(buf, col0 + 16, row_base);
      } else if (t + tile_step < total_tiles) {
        int rb, cp;
        tile_origin(t + tile_step, rb, cp);
        load_res(buf, cp, rb);
      }
    }
  }
  if (lane == 0) tma_store_wait_all();
  __syncwarp();
}

// Loss epilogue of the pixel head (act 5; src/pretraining.py:110-120, src/utils_mae.py:51-64): the accumulator row IS
// the predicted patch, so the squared error against the target patch and its gradient are formed here and the f32
// prediction never exists in memory.  Thread = row (TMEM layout), 16 columns at a time: 16 target bytes per thread
// (one 16-byte load from the compact byte targets jm_masked_targets wrote), normalised with one FMA per pixel
// ((u / 255 - mean_c) / std_c = u * a_c + b_c; channel = column % 3), optional norm_pix; d = acc + bias - t;
// out_bf16 = d * loss_scale straight from the row registers (32 contiguous bytes per thread), and the 16-column sum
// of d^2 goes to row_loss[row, col / 16]: every entry is written exactly once, so the loss is a deterministic sum.
// The pixel head is 0.9 % of the step's FLOPs; nothing here is tuned beyond "no shared-memory round trip".
template <int BN, int CG>
__device__ __forceinline__ void epilogue_loss(const EpiParams& ep, const GemmShape& sh, uint64_t* tfull,
                                              uint64_t* tempty, uint32_t tmem_base, int warp, int lane,
                                              int total_tiles, int tile0, int tile_step, int m_off) {
  constexpr int PART_N = BN / 4;
  constexpr int NC = PART_N / 16;
  const int e = warp - kEpiWarp0;
  const int q = warp & 3;
  const int part = e >> 2;
  const int tiles_mn = sh.num_m * sh.num_n;
  const int parts = (sh.N + 15) / 16;
  // ImageNet statistics of src/pretraining.py:90-91 as one FMA per pixel
  const float a3[3] = {1.0f / (255.0f * 0.229f), 1.0f / (255.0f * 0.224f), 1.0f / (255.0f * 0.225f)};
  const float b3[3] = {-0.485f / 0.229f, -0.456f / 0.224f, -0.406f / 0.225f};
  int iter = 0;
  for (int t = tile0; t < total_tiles; t += tile_step, ++iter) {
    const int rem = t % tiles_mn;
    int m_blk, n_blk;
    tile_mn(sh, rem, m_blk, n_blk);
    const int row = m_blk * (BM * CG) + m_off + q * 32 + lane;
    const int colp = n_blk * BN + part * PART_N;
    const bool row_ok = row < sh.M;
    const int acc = iter & 1;
    const uint32_t acc_phase = (iter >> 1) & 1;
    float mu = 0.f, inv = 1.f;
    if (ep.loss_stats != nullptr && row_ok) mu = ep.loss_stats[2 * static_cast<long>(row)], inv = ep.loss_stats[2 * static_cast<long>(row) + 1];
    // this row's target bytes of the warp's columns, fetched before the accumulator is waited for
    uint4 tb[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      const int col0 = colp + 16 * c;
      tb[c] = make_uint4(0u, 0u, 0u, 0u);
      if (row_ok && col0 < sh.N)  // (N is a multiple of 4 and ld_target of 16: a ragged last chunk stays inside the row)
        tb[c] = *reinterpret_cast<const uint4*>(ep.loss_target + static_cast<long>(row) * ep.ld_target + col0);
    }
    mbar_wait(&tfull[acc], acc_phase);
    tc_fence_after();
    const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN + part * PART_N;
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      const int col0 = colp + 16 * c;
      const bool live = col0 < sh.N;  // warp-uniform
      uint32_t r[16];
      if (live) tmem_ld_32x16(t_addr + 16 * c, r);
      tmem_ld_wait();
      if (c == NC - 1) {  // accumulator drained: hand the TMEM buffer back to the MMA warp
        tc_fence_before();
        if constexpr (CG == 1) mbar_arrive(&tempty[acc]);
        else mbar_arrive_cluster(mapa_shared(smem_u32(&tempty[acc]), cluster_ctarank() & ~1u));
      }
      if (!live) continue;
      const int ch0 = col0 % 3;
      const uint32_t words[4] = {tb[c].x, tb[c].y, tb[c].z, tb[c].w};
      float sq = 0.f;
      uint32_t o[8];
#pragma unroll
      for (int j = 0; j < 16; j += 2) {
        float d[2];
#pragma unroll
        for (int k = 0; k < 2; ++k) {
          const int jj = j + k;
          const int col = col0 + jj;
          int ch = ch0 + (jj % 3);
          ch = ch >= 3 ? ch - 3 : ch;
          const float u = static_cast<float>((words[jj >> 2] >> (8 * (jj & 3))) & 0xffu);
          float tv = fmaf(u, ch == 0 ? a3[0] : (ch == 1 ? a3[1] : a3[2]), ch == 0 ? b3[0] : (ch == 1 ? b3[1] : b3[2]));
          tv = (tv - mu) * inv;
          const float bias = (ep.bias != nullptr && col < sh.N) ? __ldg(ep.bias + col) : 0.f;
          d[k] = col < sh.N ? __uint_as_float(r[jj]) + bias - tv : 0.f;
          sq = fmaf(d[k], d[k], sq);
        }
        o[j >> 1] = pack_bf16(d[0] * ep.loss_scale, d[1] * ep.loss_scale);
      }
      if (row_ok) {
        ep.row_loss[static_cast<long>(row) * parts + (col0 >> 4)] = sq;
        if (ep.out_bf16 != nullptr) {  // (ldo_bf16 is a multiple of 8 and covers the padded last chunk: checked on the host)
          uint4* dst = reinterpret_cast<uint4*>(ep.out_bf16 + static_cast<long>(row) * ep.ldo_bf16 + col0);
          dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
          if (col0 + 8 < sh.N) dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
        }
      }
    }
  }
}

template <int EPI, int BN, int CG, int NBUF>
__device__ __forceinline__ void run_epilogue(const EpiParams& ep, const GemmShape& sh, const CUtensorMap* tmO,
                                             const CUtensorMap* tmO2, uint8_t* sEpi, uint64_t* rbars,
                                             uint64_t* tfull, uint64_t* tempty, uint32_t tmem_base, int warp,
                                             int lane, int total_tiles, int tile0, int tile_step, int m_off) {
  if constexpr (EPI == EPI_LOSS) {
    epilogue_loss<BN, CG>(ep, sh, tfull, tempty, tmem_base, warp, lane, total_tiles, tile0, tile_step, m_off);
  } else if constexpr (EPI == EPI_BF16 || EPI == EPI_GELU_FWD || EPI == EPI_GELU_BWD) {
    epilogue_rows<EPI, BN, CG, NBUF>(ep, sh, tmO, tmO2, sEpi, rbars, tfull, tempty, tmem_base, warp, lane, total_tiles, tile0,
                                     tile_step, m_off);
  } else {
    if constexpr (EPI == EPI_F32) {
      if (ep.tma_store) {  // kernel-uniform
        epilogue_f32_tma<BN, CG, NBUF>(ep, sh, tmO, tmO2, sEpi, rbars, tfull, tempty, tmem_base, warp, lane,
                                       total_tiles, tile0, tile_step, m_off);
        return;
      }
    }
    epilogue_role<EPI, BN, CG, NBUF>(ep, sh, sEpi, tfull, tempty, tmem_base, warp, lane, total_tiles, tile0, tile_step,
                                     m_off);
  }
}

template <int BN, bool A_MN, bool B_MN, int EPI>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmO2, const GemmShape sh,
            const EpiParams ep) {
  using Cfg = GemmCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B tiles need 1024-byte alignment; align by hand.
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * Cfg::A_BYTES;
  uint8_t* sEpi = smem + STAGES * Cfg::STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sEpi + Cfg::EPI_BYTES);
  uint64_t* full = bars;
  uint64_t* empty = bars + STAGES;
  uint64_t* tfull = bars + 2 * STAGES;
  uint64_t* tempty = bars + 2 * STAGES + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
  uint64_t* rbars = bars + 2 * STAGES + 6;  // two per epilogue warp: residual tiles arriving by TMA

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 0 && elect_one()) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], kEpiWarps * 32);
    }
    for (int i = 0; i < kEpiWarps * 2; ++i) mbar_init(&rbars[i], 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<Cfg::TMEM_COLS>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  // programmatic dependent launch: the prologue above (barriers, TMEM) ran while the previous kernel drained;
  // its results are needed from here on
  pdl_wait();
  pdl_launch_dependents();

  const int tiles_mn = sh.num_m * sh.num_n;
  const int total_tiles = tiles_mn * sh.splits;

  if (warp < kEpiWarp0) setmaxnreg_dec<JM_GEMM_REG_DEC>();
  if (warp == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int split = t / tiles_mn;
        const int rem = t - split * tiles_mn;
        int m_blk, n_blk;
    tile_mn(sh, rem, m_blk, n_blk);
        const int kb0 = static_cast<int>(static_cast<long>(sh.kblocks) * split / sh.splits);
        const int kb1 = static_cast<int>(static_cast<long>(sh.kblocks) * (split + 1) / sh.splits);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full[stage], Cfg::STAGE_BYTES);
          uint8_t* a_dst = sA + stage * Cfg::A_BYTES;
          uint8_t* b_dst = sB + stage * Cfg::B_BYTES;
          if (!A_MN) {
            tma_load_2d(&tmA, &full[stage], a_dst, kb * BK, m_blk * BM);
          } else {
#pragma unroll
            for (int i = 0; i < BM / 64; ++i)
              tma_load_2d(&tmA, &full[stage], a_dst + i * (BK * 128), m_blk * BM + i * 64, kb * BK);
          }
          if (!B_MN) {
            tma_load_2d(&tmB, &full[stage], b_dst, kb * BK, n_blk * BN);
          } else {
#pragma unroll
            for (int i = 0; i < BN / 64; ++i)
              tma_load_2d(&tmB, &full[stage], b_dst + i * (BK * 128), n_blk * BN + i * 64, kb * BK);
          }
          if (++stage == STAGES) stage = 0, phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = make_idesc_bf16(BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
    int stage = 0;
    uint32_t phase = 0;
    int iter = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++iter) {
      const int split = t / tiles_mn;
      const int kb0 = static_cast<int>(static_cast<long>(sh.kblocks) * split / sh.splits);
      const int kb1 = static_cast<int>(static_cast<long>(sh.kblocks) * (split + 1) / sh.splits);
      const int acc = iter & 1;
      const uint32_t acc_phase = (iter >> 1) & 1;
      mbar_wait(&tempty[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * BN;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t a_addr = smem_u32(sA + stage * Cfg::A_BYTES);
          const uint32_t b_addr = smem_u32(sB + stage * Cfg::B_BYTES);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            // K-major: 16 elements = 32 bytes inside the swizzled 128-byte row.
            // MN-major: 16 k-rows of 128 bytes; 64-wide MN atoms are BK*128 bytes apart (LBO).
            const uint64_t da = A_MN ? make_smem_desc_sw128(a_addr + k * (UMMA_K * 128), BK * 128, 1024)
                                     : make_smem_desc_sw128(a_addr + k * (UMMA_K * 2), 0, 1024);
            const uint64_t db = B_MN ? make_smem_desc_sw128(b_addr + k * (UMMA_K * 128), BK * 128, 1024)
                                     : make_smem_desc_sw128(b_addr + k * (UMMA_K * 2), 0, 1024);
            umma_bf16(d_tmem, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty[stage]);
          if (kb == kb1 - 1) umma_commit(&tfull[acc]);
        }
        __syncwarp();
        if (++stage == STAGES) stage = 0, phase ^= 1;
      }
    }
  } else if (warp >= kEpiWarp0) {
    setmaxnreg_inc<JM_GEMM_REG_INC>();
    run_epilogue<EPI, BN, 1, Cfg::EPI_BUFS>(ep, sh, &tmO, &tmO2, sEpi, rbars, tfull, tempty, tmem_base, warp, lane, total_tiles, blockIdx.x,
                             gridDim.x, 0);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}

// ---------------------------------------------------------------- CTA-pair (cta_group::2) kernel
// One 256 x 256 output tile per pair of CTAs on neighbouring SMs.  Each CTA stages its own 128 rows of A and
// only HALF of the B tile (128 columns); one tcgen05.mma.cta_group::2 issued by the leader CTA drives both
// tensor cores, each reading its local A and both halves of B.  Per CTA and k-block that is 32 KB of TMA
// writes and 32 KB of operand reads instead of 48 + 48: the single-CTA kernel is bound by exactly that
// shared-memory / L2-to-SM traffic (a 128 x 256 x 16 MMA consumes 96 B/clk of operands at full rate).
// Barriers: `full` lives in the leader (its producer arrives, both CTAs' TMA bytes complete on it);
// `empty` and `tfull` are signalled in both CTAs by multicast commits; `tempty` lives in the leader and is
// arrived on by the epilogue threads of both CTAs.
#ifndef JM_PAIR_BF16_BUFS
#define JM_PAIR_BF16_BUFS 1  // measured 1-3 % faster on the plain bf16 GEMMs than two tiles + 5 stages
#endif
#ifndef JM_BRES_BUFS
#define JM_BRES_BUFS 1
#endif
#ifndef JM_BRES_STAGES
#define JM_BRES_STAGES 4
#endif
// BRES ("B resident", K <= 512): the pair's B tile stays in shared memory over all its k-blocks (8 x 16 KB per CTA) and
// is loaded once per column tile; only A streams through the ring.  Measured on the long-K weight-gradient GEMM: a
// k-block costs 605 cycles with no operand loads at all, 734 with one operand's and 906 with both -- the TMA fills
// share the shared-memory port with the MMA's operand reads and add to them instead of hiding behind them -- so a
// K = 512 product whose pairs keep their column tile saves one of the two fills.
template <int EPI, bool BRES = false>
struct Gemm2Cfg {
  static constexpr int BN = 256;
  static constexpr int A_BYTES = BM * BK * 2;         // this CTA's 128 rows
  static constexpr int B_BYTES = (BN / 2) * BK * 2;   // this CTA's 128 of the 256 columns
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int RES_KB = 8;                    // k-blocks of B that stay resident (BRES)
  // Two staging tiles per epilogue warp (5 pipeline stages) where the epilogue is heavy; the light plain-bf16
  // epilogue can run with one tile and a sixth stage (JM_PAIR_BF16_BUFS).
  static constexpr int EPI_BUFS = BRES ? JM_BRES_BUFS : ((EPI == EPI_BF16) ? JM_PAIR_BF16_BUFS : 2);
  static constexpr int STAGES = BRES ? JM_BRES_STAGES : (EPI_BUFS == 2 ? 5 : 6);
  static constexpr int OPER_BYTES = BRES ? STAGES * A_BYTES + RES_KB * B_BYTES : STAGES * STAGE_BYTES;
  static constexpr int TMEM_COLS = 2 * BN;
  static constexpr int BAR_BYTES = (2 * STAGES + 4) * 8 + 16 + kEpiWarps * 2 * 8 + 16;  // + residual-load barriers, bfull
  static constexpr int EPI_BYTES = kEpiWarps * kStageBytesPerWarp * EPI_BUFS;
  static constexpr int SMEM_BYTES = OPER_BYTES + EPI_BYTES + BAR_BYTES + 1024;
  static_assert(SMEM_BYTES <= 232448, "shared memory budget");
};

template <bool A_MN, bool B_MN, int EPI, bool BRES = false>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
             const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmO2, const GemmShape sh,
             const EpiParams ep) {
  using Cfg = Gemm2Cfg<EPI, BRES>;
  constexpr int STAGES = Cfg::STAGES;
  constexpr int BN = Cfg::BN;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * Cfg::A_BYTES;
  uint8_t* sEpi = smem + Cfg::OPER_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sEpi + Cfg::EPI_BYTES);
  uint64_t* full = bars;
  uint64_t* empty = bars + STAGES;
  uint64_t* tfull = bars + 2 * STAGES;
  uint64_t* tempty = bars + 2 * STAGES + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
  uint64_t* rbars = bars + 2 * STAGES + 6;  // two per epilogue warp: residual tiles arriving by TMA
  uint64_t* bfull = rbars + kEpiWarps * 2;  // BRES: the resident B tile has landed (in the leader, like `full`)

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int rank = static_cast<int>(cluster_ctarank());  // within the pair
  const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;
#ifdef JM_DEBUG_SMID
  if (threadIdx.x == 0 && blockIdx.x < 8) {
    uint32_t smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    printf("block %d rank %d smid %u\n", blockIdx.x, rank, smid);
  }
#endif

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 0 && elect_one()) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full[s], 1);   // the leader's producer arrives with the byte count of BOTH CTAs' loads
      mbar_init(&empty[s], 1);  // one multicast commit
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], 2 * kEpiWarps * 32);
    }
    for (int i = 0; i < kEpiWarps * 2; ++i) mbar_init(&rbars[i], 1);
    mbar_init(bfull, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc2<Cfg::TMEM_COLS>(tmem_ptr);
  tc_fence_before();
  cluster_sync_all();  // peer barriers initialised, TMEM allocated in both CTAs
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_wait();  // (see gemm_kernel)
  pdl_launch_dependents();

  const int tiles_mn = sh.num_m * sh.num_n;  // num_m counts 256-row pair tiles here
  const int total_tiles = tiles_mn * sh.splits;

  if (warp < kEpiWarp0) setmaxnreg_dec<JM_GEMM_REG_DEC>();
  if (warp == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      int res_n = -1;                      // BRES: the column tile whose B is resident
      int last_stage = 0;
      uint32_t last_phase = 0;             // ring position of the last k-block issued
      for (int t = pair; t < total_tiles; t += npairs) {
        const int split = t / tiles_mn;
        const int rem = t - split * tiles_mn;
        int m_blk, n_blk;
        tile_mn(sh, rem, m_blk, n_blk);
        const int kb0 = static_cast<int>(static_cast<long>(sh.kblocks) * split / sh.splits);
        const int kb1 = static_cast<int>(static_cast<long>(sh.kblocks) * (split + 1) / sh.splits);
        const int m0 = m_blk * 2 * BM + rank * BM;
        const int n0 = n_blk * BN + rank * (BN / 2);
        if constexpr (BRES) {
          if (n_blk != res_n) {  // (host: splits == 1, kblocks <= RES_KB)
            // the MMAs of the previous tile were the last readers of the resident tile: they have retired when the
            // stage of its last k-block has been handed back
            if (res_n >= 0) mbar_wait(&empty[last_stage], last_phase);
            if (rank == 0) mbar_arrive_expect_tx(bfull, 2 * (kb1 - kb0) * Cfg::B_BYTES);
            for (int kb = kb0; kb < kb1; ++kb) {
              uint8_t* b_dst = sB + (kb - kb0) * Cfg::B_BYTES;
              if (!B_MN) {
                tma_load_2d_pair(&tmB, bfull, b_dst, kb * BK, n0);
              } else {
#pragma unroll
                for (int i = 0; i < (BN / 2) / 64; ++i)
                  tma_load_2d_pair(&tmB, bfull, b_dst + i * (BK * 128), n0 + i * 64, kb * BK);
              }
            }
            res_n = n_blk;
          }
        }
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          last_stage = stage, last_phase = phase;
          if constexpr (BRES) {
            if (rank == 0) mbar_arrive_expect_tx(&full[stage], 2 * Cfg::A_BYTES);
            tma_load_2d_pair(&tmA, &full[stage], sA + stage * Cfg::A_BYTES, kb * BK, m0);
            if (++stage == STAGES) stage = 0, phase ^= 1;
            continue;
          }
          // the peer's bytes may land before this arrive: the transaction count then runs negative until it
          // does, and the phase cannot complete without the arrival
#ifdef JM_EXP_TMA_MODE  // experiments: 1 no operand loads at all, 2 only A, 3 only B (results are garbage)
          {
          if (rank == 0)
            mbar_arrive_expect_tx(&full[stage], JM_EXP_TMA_MODE == 1 ? 0 : (JM_EXP_TMA_MODE == 2 ? 2 * Cfg::A_BYTES : 2 * Cfg::B_BYTES));
          uint8_t* a_dst = sA + stage * Cfg::A_BYTES;
          uint8_t* b_dst = sB + stage * Cfg::B_BYTES;
          if (JM_EXP_TMA_MODE == 2 && !A_MN) tma_load_2d_pair(&tmA, &full[stage], a_dst, kb * BK, m0);
          if (JM_EXP_TMA_MODE == 3 && !B_MN) tma_load_2d_pair(&tmB, &full[stage], b_dst, kb * BK, n0);
          if (JM_EXP_TMA_MODE == 3 && B_MN) {
            for (int i = 0; i < (BN / 2) / 64; ++i)
              tma_load_2d_pair(&tmB, &full[stage], b_dst + i * (BK * 128), n0 + i * 64, kb * BK);
          }
          if (++stage == STAGES) stage = 0, phase ^= 1;
          continue;
          }
#endif
          if (rank == 0) mbar_arrive_expect_tx(&full[stage], 2 * Cfg::STAGE_BYTES);
          uint8_t* a_dst = sA + stage * Cfg::A_BYTES;
          uint8_t* b_dst = sB + stage * Cfg::B_BYTES;
          if (!A_MN) {
            tma_load_2d_pair(&tmA, &full[stage], a_dst, kb * BK, m0);
          } else {
#pragma unroll
            for (int i = 0; i < BM / 64; ++i)
              tma_load_2d_pair(&tmA, &full[stage], a_dst + i * (BK * 128), m0 + i * 64, kb * BK);
          }
          if (!B_MN) {
            tma_load_2d_pair(&tmB, &full[stage], b_dst, kb * BK, n0);
          } else {
#pragma unroll
            for (int i = 0; i < (BN / 2) / 64; ++i)
              tma_load_2d_pair(&tmB, &full[stage], b_dst + i * (BK * 128), n0 + i * 64, kb * BK);
          }
          if (++stage == STAGES) stage = 0, phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    if (rank == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(2 * BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      int stage = 0;
      uint32_t phase = 0;
      int iter = 0;
      int res_n = -1;
      uint32_t res_phase = 0;  // BRES: resident column tile / parity of its next arrival
      for (int t = pair; t < total_tiles; t += npairs, ++iter) {
        const int split = t / tiles_mn;
        const int kb0 = static_cast<int>(static_cast<long>(sh.kblocks) * split / sh.splits);
        const int kb1 = static_cast<int>(static_cast<long>(sh.kblocks) * (split + 1) / sh.splits);
        const int acc = iter & 1;
        const uint32_t acc_phase = (iter >> 1) & 1;
        mbar_wait(&tempty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        if constexpr (BRES) {
          int m_blk, n_blk;
          tile_mn(sh, t - split * tiles_mn, m_blk, n_blk);
          if (n_blk != res_n) {
            mbar_wait(bfull, res_phase);
            res_phase ^= 1, res_n = n_blk;
          }
        }
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t a_addr = smem_u32(sA + stage * Cfg::A_BYTES);
            const uint32_t b_addr = smem_u32(sB + (BRES ? kb - kb0 : stage) * Cfg::B_BYTES);
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) {
              const uint64_t da = A_MN ? make_smem_desc_sw128(a_addr + k * (UMMA_K * 128), BK * 128, 1024)
                                       : make_smem_desc_sw128(a_addr + k * (UMMA_K * 2), 0, 1024);
              const uint64_t db = B_MN ? make_smem_desc_sw128(b_addr + k * (UMMA_K * 128), BK * 128, 1024)
                                       : make_smem_desc_sw128(b_addr + k * (UMMA_K * 2), 0, 1024);
              umma2_bf16(d_tmem, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            }
            umma2_commit_mc(&empty[stage]);
            if (kb == kb1 - 1) umma2_commit_mc(&tfull[acc]);
          }
          __syncwarp();
          if (++stage == STAGES) stage = 0, phase ^= 1;
        }
      }
    }
  } else if (warp >= kEpiWarp0) {
    setmaxnreg_inc<JM_GEMM_REG_INC>();
    run_epilogue<EPI, 256, 2, Cfg::EPI_BUFS>(ep, sh, &tmO, &tmO2, sEpi, rbars, tfull, tempty, tmem_base, warp, lane, total_tiles, pair, npairs,
                              rank * BM);
  }

  tc_fence_before();
  cluster_sync_all();  // no CTA may exit (or free TMEM) while its peer can still signal it or read its operands
  if (warp == 1) tmem_dealloc2<Cfg::TMEM_COLS>(tmem_base);
}

// ---------------------------------------------------------------- host side
static PFN_cuTensorMapEncodeTiled_v12000 get_encode_fn() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  });
  return fn;
}

}  // namespace jm
void* jm::tmap_encode_fn() { return reinterpret_cast<void*>(jm::get_encode_fn()); }
namespace jm {

// 2-D bf16 tensor [outer, inner] row-major with row stride ld elements; operand boxes are SWIZZLE_128B, the
// epilogue's 32 x 32 output boxes SWIZZLE_64B.
static int make_tmap(CUtensorMap* map, const void* ptr, long inner, long outer, long ld, int box_inner,
                     int box_outer, CUtensorMapSwizzle swizzle = CU_TENSOR_MAP_SWIZZLE_128B, bool f32 = false) {
  auto fn = get_encode_fn();
  const int esize = f32 ? 4 : 2;
  JM_CHECK(fn != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  JM_CHECK((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "GEMM operand pointer must be 16-byte aligned");
  JM_CHECK((ld * esize) % 16 == 0, "GEMM operand row stride must be a multiple of 16 bytes (got %ld elements)", ld);
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(inner), static_cast<cuuint64_t>(outer)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * esize};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(box_inner), static_cast<cuuint32_t>(box_outer)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                  const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  JM_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (inner=%ld outer=%ld ld=%ld)",
           static_cast<int>(r), inner, outer, ld);
  return 0;
}

template <int BN, bool A_MN, bool B_MN, int EPI>
static int launch_gemm(const GemmMaps& tm, const GemmShape& sh, const EpiParams& ep,
                       int grid, cudaStream_t stream) {
  auto kern = gemm_kernel<BN, A_MN, B_MN, EPI>;
  // function attributes are per device: one bit per device ordinal and template instantiation
  static DeviceOnce attr_set;
  const int dev = current_device();
  JM_CHECK(dev >= 0, "jm_gemm: no current CUDA device");
  if (!attr_set.done(dev)) {
    JM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, GemmCfg<BN>::SMEM_BYTES));
    attr_set.mark(dev);
  }
  JM_LAUNCH_PDL(kern, grid, kGemmThreads, GemmCfg<BN>::SMEM_BYTES, stream, tm.a, tm.b, tm.o, tm.o2, sh, ep);
  JM_AFTER_LAUNCH();
  return 0;
}

template <bool A_MN, bool B_MN, int EPI, bool BRES = false>
static int launch_gemm2(const GemmMaps& tm, const GemmShape& sh, const EpiParams& ep,
                        int grid, cudaStream_t stream) {
  auto kern = gemm2_kernel<A_MN, B_MN, EPI, BRES>;
  constexpr int kSmem = Gemm2Cfg<EPI, BRES>::SMEM_BYTES;
  static DeviceOnce attr_set;
  const int dev = current_device();
  JM_CHECK(dev >= 0, "jm_gemm: no current CUDA device");
  if (!attr_set.done(dev)) {
    JM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
    attr_set.mark(dev);
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = kSmem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[3];
  const int csize = 2;
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = csize;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  static const int policy = env_int("JM_PAIR_POLICY", 0);  // process-wide, initialised once (thread-safe)
  if (policy > 0) {
    attr[2].id = cudaLaunchAttributeClusterSchedulingPolicyPreference;
    attr[2].val.clusterSchedulingPolicyPreference =
        policy == 1 ? cudaClusterSchedulingPolicySpread : cudaClusterSchedulingPolicyLoadBalancing;
    cfg.numAttrs = 3;
  }
  // A persistent grid must not exceed the clusters the GPU can hold at once (GPCs with an odd SM left over
  // cannot host a pair there): surplus clusters would run as a second wave.
  static std::atomic<int> max_clusters_dev[2][kMaxDevices];  // per cluster size and device ordinal (zero-initialised)
  std::atomic<int>& cached = max_clusters_dev[0][dev];
  int max_clusters = cached.load(std::memory_order_acquire);
  if (max_clusters == 0) {
    int n = 0;
    JM_CUDA(cudaOccupancyMaxActiveClusters(&n, kern, &cfg));
    JM_CHECK(n > 0, "jm_gemm: the CTA-pair kernel cannot be scheduled on this device");
    if (getenv("JM_PAIR_VERBOSE")) fprintf(stderr, "jm: pair kernel max active clusters of %d = %d\n", csize, n);
    cached.store(n, std::memory_order_release);
    max_clusters = n;
  }
  if (grid > csize * max_clusters) {
    cfg.gridDim = dim3(csize * max_clusters);
  }
  JM_CUDA(cudaLaunchKernelEx(&cfg, kern, tm.a, tm.b, tm.o, tm.o2, sh, ep));
  count_launch();
  return 0;
}

// B-resident instantiations (K <= 512, A K-major): the four TMA-fed epilogues
template <bool B_MN>
static int dispatch_bres(int epi, const GemmMaps& tm, const GemmShape& sh, const EpiParams& ep, int grid,
                         cudaStream_t stream) {
  switch (epi) {
    case EPI_BF16: return launch_gemm2<false, B_MN, EPI_BF16, true>(tm, sh, ep, grid, stream);
    case EPI_F32: return launch_gemm2<false, B_MN, EPI_F32, true>(tm, sh, ep, grid, stream);
    case EPI_GELU_FWD: return launch_gemm2<false, B_MN, EPI_GELU_FWD, true>(tm, sh, ep, grid, stream);
    default: return launch_gemm2<false, B_MN, EPI_GELU_BWD, true>(tm, sh, ep, grid, stream);
  }
}

template <bool A_MN, bool B_MN>
static int dispatch_epi2(int epi, const GemmMaps& tm, const GemmShape& sh,
                         const EpiParams& ep, int grid, cudaStream_t stream) {
  switch (epi) {
    case EPI_BF16: return launch_gemm2<A_MN, B_MN, EPI_BF16>(tm, sh, ep, grid, stream);
    case EPI_F32: return launch_gemm2<A_MN, B_MN, EPI_F32>(tm, sh, ep, grid, stream);
    case EPI_GELU_FWD: return launch_gemm2<A_MN, B_MN, EPI_GELU_FWD>(tm, sh, ep, grid, stream);
    case EPI_GELU_BWD: return launch_gemm2<A_MN, B_MN, EPI_GELU_BWD>(tm, sh, ep, grid, stream);
    default: return launch_gemm2<A_MN, B_MN, EPI_GENERIC>(tm, sh, ep, grid, stream);
  }
}

// Pair kernel instantiations: forward (A K-major, B MN-major), input gradient (K, K), weight gradient (MN, MN).
static int dispatch_pair(bool a_mn, bool b_mn, int epi, const GemmMaps& tm, const GemmShape& sh, const EpiParams& ep, int grid, cudaStream_t stream, bool bres) {
  if (epi == EPI_LOSS) return launch_gemm2<false, true, EPI_LOSS>(tm, sh, ep, grid, stream);
  if (bres) return b_mn ? dispatch_bres<true>(epi, tm, sh, ep, grid, stream) : dispatch_bres<false>(epi, tm, sh, ep, grid, stream);
  if (!a_mn && !b_mn) return dispatch_epi2<false, false>(epi, tm, sh, ep, grid, stream);
  if (!a_mn && b_mn) return dispatch_epi2<false, true>(epi, tm, sh, ep, grid, stream);
  if (epi == EPI_F32) return launch_gemm2<true, true, EPI_F32>(tm, sh, ep, grid, stream);
  return launch_gemm2<true, true, EPI_GENERIC>(tm, sh, ep, grid, stream);
}

template <int BN, bool A_MN, bool B_MN>
static int dispatch_epi(int epi, const GemmMaps& tm, const GemmShape& sh,
                        const EpiParams& ep, int grid, cudaStream_t stream) {
  switch (epi) {
    case EPI_BF16: return launch_gemm<BN, A_MN, B_MN, EPI_BF16>(tm, sh, ep, grid, stream);
    case EPI_F32: return launch_gemm<BN, A_MN, B_MN, EPI_F32>(tm, sh, ep, grid, stream);
    case EPI_GELU_FWD: return launch_gemm<BN, A_MN, B_MN, EPI_GELU_FWD>(tm, sh, ep, grid, stream);
    case EPI_GELU_BWD: return launch_gemm<BN, A_MN, B_MN, EPI_GELU_BWD>(tm, sh, ep, grid, stream);
    default: return launch_gemm<BN, A_MN, B_MN, EPI_GENERIC>(tm, sh, ep, grid, stream);
  }
}

template <int BN>
static int dispatch_major(bool a_mn, bool b_mn, int epi, const GemmMaps& tm, const GemmShape& sh, const EpiParams& ep, int grid, cudaStream_t stream) {
  if (epi == EPI_LOSS) return launch_gemm<BN, false, true, EPI_LOSS>(tm, sh, ep, grid, stream);
  if (!a_mn && !b_mn) return dispatch_epi<BN, false, false>(epi, tm, sh, ep, grid, stream);
  if (!a_mn && b_mn) return dispatch_epi<BN, false, true>(epi, tm, sh, ep, grid, stream);
  // transposed-A products are weight gradients: f32 accumulate epilogue (anything else: generic)
  if (epi != EPI_F32) epi = EPI_GENERIC;
  if (a_mn && !b_mn) {
    if (epi == EPI_F32) return launch_gemm<BN, true, false, EPI_F32>(tm, sh, ep, grid, stream);
    return launch_gemm<BN, true, false, EPI_GENERIC>(tm, sh, ep, grid, stream);
  }
  if (epi == EPI_F32) return launch_gemm<BN, true, true, EPI_F32>(tm, sh, ep, grid, stream);
  return launch_gemm<BN, true, true, EPI_GENERIC>(tm, sh, ep, grid, stream);
}

}  // namespace jm

// Split-K factor for `tiles` output tiles on `slots` CTAs (or CTA pairs): the one that minimises
// waves(tiles * s) x (k-blocks per unit + a fixed per-unit charge), i.e. the time of the busiest CTA.
// (Filling the machine exactly once is not always best: 16 tiles on 74 pairs run 11 % faster as 9 splits
// in two waves than as 4 splits on 64 pairs.)
static int pick_splits(long tiles, long slots, int kblocks) {
  const int max_s = kblocks / 2 < 64 ? kblocks / 2 : 64;
  int best_s = 1;
  double best = 0.0;
  for (int s = 1; s <= (max_s < 1 ? 1 : max_s); ++s) {
    const long waves = (tiles * s + slots - 1) / slots;
    // per work unit: its share of the k-blocks plus ~4 k-blocks' worth of pipeline fill and epilogue drain
    const double cost = static_cast<double>(waves) * (static_cast<double>(kblocks) / s + 4.0);
    if (best == 0.0 || cost < best * 0.98) best = cost, best_s = s;
  }
  return best_s;
}

extern "C" int jm_gemm(const jm_gemm_args* a, void* stream_) {
  using namespace jm;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  JM_CHECK(a != nullptr, "jm_gemm: null args");
  JM_CHECK(a->M > 0 && a->N > 0 && a->K > 0, "jm_gemm: empty problem M=%d N=%d K=%d", a->M, a->N, a->K);
  JM_CHECK(a->N % 4 == 0, "jm_gemm: N must be a multiple of 4 (got %d)", a->N);
  JM_CHECK(a->A && a->B, "jm_gemm: null operand");
  JM_CHECK(a->out_f32 || a->out_bf16 || a->out2_bf16 || a->act == 5, "jm_gemm: no output");
  JM_CHECK(a->act >= 0 && a->act <= 5, "jm_gemm: act must be 0..5");
  if (a->act == 5) {
    JM_CHECK(a->loss_target && a->row_loss && !a->a_mn_major && a->b_mn_major,
             "jm_gemm: the loss epilogue needs loss_target, row_loss, A K-major and B MN-major");
    JM_CHECK(a->ld_target % 16 == 0 && a->ld_target >= a->N && (reinterpret_cast<uintptr_t>(a->loss_target) & 15) == 0,
             "jm_gemm: ld_target must be a multiple of 16 and at least N, loss_target 16-byte aligned");
    JM_CHECK(!a->out_f32 && !a->out2_bf16 && !a->residual && !a->pe && !a->colsum && !a->aux && !a->row_scale &&
                 a->row_period == 0 && a->col_scale_n == 0 && !a->accumulate && a->splits <= 1,
             "jm_gemm: the loss epilogue takes no other epilogue option");
    JM_CHECK(!a->out_bf16 || a->ldo_bf16 >= (a->N + 7) / 8 * 8, "jm_gemm: ldo_bf16 must cover N rounded up to 8");
  }
  JM_CHECK((a->act != 2 && a->act != 4) || a->aux, "jm_gemm: act=2/4 needs aux");
  JM_CHECK(!a->pe || (a->pe_idx && a->row_period > 0), "jm_gemm: pe needs pe_idx and row_period");
  JM_CHECK(!a->out_f32 || a->ldo_f32 % 4 == 0, "jm_gemm: ldo_f32 must be a multiple of 4");
  JM_CHECK(!a->out_bf16 || a->ldo_bf16 % 8 == 0, "jm_gemm: ldo_bf16 must be a multiple of 8");
  JM_CHECK(!a->out2_bf16 || a->ldo2_bf16 % 8 == 0, "jm_gemm: ldo2_bf16 must be a multiple of 8");
  JM_CHECK(((reinterpret_cast<uintptr_t>(a->out_bf16) | reinterpret_cast<uintptr_t>(a->out2_bf16) |
             reinterpret_cast<uintptr_t>(a->aux) | reinterpret_cast<uintptr_t>(a->out_f32) |
             reinterpret_cast<uintptr_t>(a->residual) | reinterpret_cast<uintptr_t>(a->bias)) & 15) == 0,
           "jm_gemm: epilogue pointers must be 16-byte aligned");
  JM_CHECK(!a->residual || a->ldr % 4 == 0, "jm_gemm: ldr must be a multiple of 4");
  JM_CHECK(!a->aux || a->ldaux % 8 == 0, "jm_gemm: ldaux must be a multiple of 8");

  const int sms = sm_count();
  JM_CHECK(sms > 0, "jm_gemm: no CUDA device");
  const int num_m = (a->M + BM - 1) / BM;
  const int kblocks = (a->K + BK - 1) / BK;
  const bool can_split = a->accumulate && a->out_f32 && !a->out_bf16 && !a->out2_bf16 && !a->bias &&
                         !a->residual && !a->pe && a->act == 0;
  const bool amn = a->a_mn_major != 0, bmn = a->b_mn_major != 0;
  static const int auto_pair = env_int("JM_GEMM_PAIR", 1) != 0;  // process-wide, initialised once (thread-safe)
  int bn = a->block_n;
  bool pair_pick = false;
  if (bn == 0) {
    // Pick the tile by a small cost model in SM cycles: waves x max(main loop, epilogue) per tile, plus the
    // last tile's exposed epilogue.  Cycles per 64-deep k-block per CTA are the measured steady-state rates
    // (the 256-wide tiles halve the operand traffic per MMA, the CTA pair halves the B traffic again).
    const bool pair_allowed = a->cta_pair >= 0 && (a->cta_pair > 0 || auto_pair) && !(amn && !bmn) && a->M > BM &&
                              a->N >= 256;
    double best = 0.0;
    bn = 64;
    for (int c = 0; c < 4; ++c) {  // pair 256, single 256, 128, 64: ties keep the earlier (larger) tile
      const bool pr = c == 0;
      const int w = c <= 1 ? 256 : (c == 2 ? 128 : 64);
      if (pr && !pair_allowed) continue;
      if (!pr && a->cta_pair > 0) continue;
      if (w > 64 && a->N < w) continue;
      const long slots = pr ? sms / 2 : sms;
      const long tiles = (pr ? (a->M + 2 * BM - 1) / (2 * BM) : num_m) * static_cast<long>((a->N + w - 1) / w);
      long sp = 1;
      if (a->splits > 0) sp = a->splits;
      else if (can_split) {
        sp = pick_splits(tiles, slots, kblocks);
      }
      const long waves = (tiles * sp + slots - 1) / slots;
      const double per_kb = pr ? 512.0 : (w == 256 ? 560.0 : (w == 128 ? 350.0 : 260.0));
      const double main_loop = per_kb * kblocks / sp, epi = 14.0 * w;
      const double cost = waves * (main_loop > epi ? main_loop : epi) + epi;
      if (best == 0.0 || cost < best) best = cost, bn = w, pair_pick = pr;
    }
  }
  JM_CHECK(bn == 64 || bn == 128 || bn == 256, "jm_gemm: block_n must be 64, 128 or 256");
  const int num_n = (a->N + bn - 1) / bn;
  int splits = a->splits;
  if (splits <= 0) {
    splits = 1;
    if (can_split) {
      splits = pick_splits(static_cast<long>(num_m) * num_n, sms, kblocks);
    }
  }
  JM_CHECK(splits == 1 || can_split, "jm_gemm: split-K needs a pure accumulate epilogue");
  JM_CHECK(splits <= kblocks, "jm_gemm: more splits (%d) than k-blocks (%d)", splits, kblocks);

  // CTA-pair kernel (256 x 256 tiles, cta_group::2): each CTA stages only half of B, which relieves the
  // shared-memory / L2-to-SM traffic the 128 x 256 single-CTA tile is bound by.
  const bool pair_ok = bn == 256 && !(amn && !bmn) && a->M > BM;
  bool pair = false;
  if (a->cta_pair > 0) {
    JM_CHECK(pair_ok, "jm_gemm: cta_pair needs block_n 256, M > 128 and a supported operand layout");
    pair = true;
  } else if (a->cta_pair == 0) {
    pair = a->block_n == 0 ? pair_pick : (pair_ok && auto_pair);
  }
  int num_m_eff = num_m;
  if (pair) {
    num_m_eff = (a->M + 2 * BM - 1) / (2 * BM);
    if (a->splits <= 0) {
      splits = 1;
      if (can_split) {
        splits = pick_splits(static_cast<long>(num_m_eff) * num_n, sms / 2, kblocks);
      }
    }
  }

  GemmMaps tm;
  memset(&tm, 0, sizeof(tm));
  CUtensorMap& tmA = tm.a;
  CUtensorMap& tmB = tm.b;
  if (!a->a_mn_major) {
    if (make_tmap(&tmA, a->A, a->K, a->M, a->lda, BK, BM)) return 1;
  } else {
    if (make_tmap(&tmA, a->A, a->M, a->K, a->lda, 64, BK)) return 1;
  }
  if (!a->b_mn_major) {
    if (make_tmap(&tmB, a->B, a->K, a->N, a->ldb, BK, pair ? bn / 2 : bn)) return 1;
  } else {
    if (make_tmap(&tmB, a->B, a->N, a->K, a->ldb, 64, BK)) return 1;
  }

  GemmShape sh{a->M, a->N, a->K, num_m_eff, num_n, splits, kblocks, 1};
  EpiParams ep;
  ep.out_f32 = a->out_f32, ep.ldo_f32 = a->ldo_f32;
  ep.out_bf16 = reinterpret_cast<__nv_bfloat16*>(a->out_bf16), ep.ldo_bf16 = a->ldo_bf16;
  ep.out2_bf16 = reinterpret_cast<__nv_bfloat16*>(a->out2_bf16), ep.ldo2_bf16 = a->ldo2_bf16;
  ep.bias = a->bias;
  ep.residual = a->residual, ep.ldr = a->ldr;
  ep.aux = reinterpret_cast<const __nv_bfloat16*>(a->aux), ep.ldaux = a->ldaux;
  ep.pe = a->pe, ep.pe_idx = a->pe_idx, ep.ldpe = a->ldpe;
  ep.act = a->act;
  ep.col_scale = a->col_scale, ep.col_scale_n = a->col_scale_n;
  ep.row_period = a->row_period, ep.row_skip = a->row_skip, ep.row_mask_mode = a->row_mask_mode;
  ep.accumulate = (a->accumulate || splits > 1) ? 1 : 0;
  ep.colsum = a->colsum;
  ep.row_scale = a->row_scale, ep.row_scale_period = a->row_scale_period;
  ep.col_vec = a->col_scale_vec;
  ep.pe_idx_stride = a->pe_idx_stride;
  ep.loss_target = a->loss_target, ep.ld_target = a->ld_target, ep.loss_stats = a->loss_stats;
  ep.row_loss = a->row_loss, ep.loss_scale = a->loss_scale;
  JM_CHECK(!a->row_scale || a->row_scale_period > 0, "jm_gemm: row_scale needs row_scale_period > 0");
  JM_CHECK(!a->colsum || (a->act != 1 && a->act != 3), "jm_gemm: colsum is not defined for the GELU-forward epilogue");

  const long total = static_cast<long>(num_m_eff) * num_n * splits;
  const int grid = pair ? 2 * static_cast<int>(total < sms / 2 ? total : sms / 2)
                        : static_cast<int>(total < sms ? total : sms);
  int epi = EPI_GENERIC;
  const bool only_bf16 = ep.out_bf16 && !ep.out_f32 && !ep.out2_bf16;
  const bool only_f32 = ep.out_f32 && !ep.out_bf16 && !ep.out2_bf16;
  // the row-register epilogues work on whole 16-byte column chunks and a per-32-column scale
  const bool rows_ok = a->N % 8 == 0 && ep.col_scale_n % 32 == 0;
  if ((ep.act == 1 || ep.act == 3) && ep.out_bf16 && ep.out2_bf16 && !ep.out_f32 && !ep.residual && !ep.pe && rows_ok)
    epi = EPI_GELU_FWD;
  else if ((ep.act == 2 || ep.act == 4) && only_bf16 && !ep.residual && !ep.pe && rows_ok && !ep.bias &&
           ep.col_scale_n == 0 &&
           (reinterpret_cast<uintptr_t>(ep.colsum) & 15) == 0)
    epi = EPI_GELU_BWD;
  else if (ep.act == 0 && only_f32 && !ep.colsum) epi = EPI_F32;
  else if (ep.act == 0 && only_bf16 && !ep.residual && !ep.pe && !ep.colsum && rows_ok) epi = EPI_BF16;
  if (ep.act == 5) epi = EPI_LOSS;
  if (amn && epi != EPI_F32) epi = EPI_GENERIC;  // transposed-A products are weight gradients
  ep.tma_store = 0;
  if ((epi == EPI_BF16 || epi == EPI_GELU_FWD || epi == EPI_GELU_BWD) &&
      !(ep.row_period > 0 && ep.row_mask_mode == 2)) {  // skipped rows need per-row predicates: plain stores
    if (make_tmap(&tm.o, ep.out_bf16, a->N, a->M, ep.ldo_bf16, 32, 32, CU_TENSOR_MAP_SWIZZLE_64B)) return 1;
    if (epi == EPI_GELU_FWD &&
        make_tmap(&tm.o2, ep.out2_bf16, a->N, a->M, ep.ldo2_bf16, 32, 32, CU_TENSOR_MAP_SWIZZLE_64B))
      return 1;
    if (epi == EPI_GELU_BWD &&  // the pre-activation / saved-derivative tiles come in by TMA as well
        make_tmap(&tm.o2, ep.aux, a->N, a->M, ep.ldaux, 32, 32, CU_TENSOR_MAP_SWIZZLE_64B))
      return 1;
    ep.tma_store = 1;
  }
  if (epi == EPI_F32 && !ep.pe && !(ep.row_period > 0 && ep.row_mask_mode != 0) && ep.col_scale_n == 0) {
    // f32 output (and residual) tiles of 32 rows x 16 columns move by TMA; accumulation is a TMA reduce-add
    if (make_tmap(&tm.o, ep.out_f32, a->N, a->M, ep.ldo_f32, 16, 32, CU_TENSOR_MAP_SWIZZLE_64B, true)) return 1;
    if (ep.residual &&
        make_tmap(&tm.o2, ep.residual, a->N, a->M, ep.ldr, 16, 32, CU_TENSOR_MAP_SWIZZLE_64B, true))
      return 1;
    ep.tma_store = 1;
  }
  JM_CHECK(!ep.col_vec || (epi == EPI_F32 && ep.tma_store),
           "jm_gemm: col_scale_vec is built for the f32 output epilogue (out_f32 only, no pe / row mask / col_scale)");
  JM_CHECK(!ep.row_scale || (epi == EPI_F32 && ep.tma_store),
           "jm_gemm: row_scale is built for the f32 output epilogue (out_f32 only, no pe / row mask / col_scale)");
  if (pair) {
    // K <= 512 with A K-major: B resident in shared memory.  The tile order is n fastest, so with a number of pairs
    // that is a multiple of the column-tile count every pair keeps its column tile (B is loaded once per CTA) while
    // the pairs of one tile row run side by side (A is read from HBM once).
    static const int auto_bres = env_int("JM_GEMM_BRES", 1);  // process-wide, initialised once (thread-safe)
    const bool bres = (auto_bres || a->cta_pair == 5) && !amn && splits == 1 && kblocks <= Gemm2Cfg<EPI_BF16, true>::RES_KB &&
                      a->cta_pair != 2 && num_n <= sms / 2 && total >= sms / 2 &&
                      (epi == EPI_BF16 || epi == EPI_GELU_FWD || epi == EPI_GELU_BWD || epi == EPI_F32) && ep.tma_store;
    int g = grid;
    if (bres && a->cta_pair != 5) g = 2 * ((sms / 2) / num_n) * num_n;  // (5: test hook, column tiles change)
    return dispatch_pair(amn, bmn, epi, tm, sh, ep, g, stream, bres);
  }
  if (bn == 256) return dispatch_major<256>(amn, bmn, epi, tm, sh, ep, grid, stream);
  if (bn == 128) return dispatch_major<128>(amn, bmn, epi, tm, sh, ep, grid, stream);
  return dispatch_major<64>(amn, bmn, epi, tm, sh, ep, grid, stream);
}
