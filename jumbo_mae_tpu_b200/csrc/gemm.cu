// tcgen05 / TMEM / TMA GEMM for sm_100a with fused epilogues.
//
// One persistent CTA per SM, 256 threads, warp-specialised:
//   warp 0      TMA producer   (one elected lane; cp.async.bulk.tensor into a SWIZZLE_128B ring)
//   warp 1      MMA issuer     (one elected lane; tcgen05.mma 128 x BN x 16, accumulators in TMEM)
//   warp 2      TMEM allocator (2 x BN f32 columns: the accumulator is double-buffered so the
//                               epilogue of tile i overlaps the MMAs of tile i+1)
//   warps 4..7  epilogue       (tcgen05.ld 32 lanes x 32 columns -> registers -> fused epilogue -> global)
// Pipelines: smem full/empty mbarriers (TMA <-> MMA), TMEM full/empty mbarriers (MMA <-> epilogue).
//
// Operand majors: both A and B may be K-major ([rows, K] row-major) or MN-major ([K, rows]
// row-major); this covers Y = X W (B MN-major: a Flax kernel is [in, out]), dX = dY W^T (B K-major)
// and dW = X^T dY (both MN-major) without any transposed copies.
//
// Replaces the XLA dot_general behind every Dense/DenseGeneral/Conv of the reference
// (src/modeling.py:108-148, src/pretraining.py:84-85) and their autodiff transposes.
#include <cudaTypedefs.h>

#include "../../include/jumbo_mae_b200.h"
#include "common.cuh"

namespace jm {

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = 128 bytes = one SWIZZLE_128B row
constexpr int UMMA_K = 16;
constexpr int kGemmThreads = 256;
constexpr int kEpiWarp0 = 4;

struct GemmShape {
  int M, N, K;
  int num_m, num_n, splits, kblocks;
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
};

template <int BN>
struct GemmCfg {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = (BN == 256) ? 4 : ((BN == 128) ? 6 : 8);
  static constexpr int TMEM_COLS = 2 * BN;
  static constexpr int BAR_BYTES = (2 * STAGES + 4) * 8 + 16;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + BAR_BYTES + 1024;  // +1024: manual alignment
};

__device__ __forceinline__ void store_f32x4(float* p, float4 v, bool atomic) {
  if (atomic) {
    atomicAdd(reinterpret_cast<float4*>(p), v);
  } else {
    *reinterpret_cast<float4*>(p) = v;
  }
}

// Epilogue for 8 consecutive columns [col, col+8) of one row; nvalid in {0, 4, 8} columns are in range.
__device__ __forceinline__ void epilogue8(const EpiParams& ep, float (&v)[8], long row, int col, int nvalid, int rp,
                                          bool masked) {
  if (ep.bias) {
#pragma unroll
    for (int h = 0; h < 2; ++h)
      if (4 * h < nvalid) {
        float4 b = *reinterpret_cast<const float4*>(ep.bias + col + 4 * h);
        v[4 * h + 0] += b.x, v[4 * h + 1] += b.y, v[4 * h + 2] += b.z, v[4 * h + 3] += b.w;
      }
  }
  if (col < ep.col_scale_n) {
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (col + j < ep.col_scale_n) v[j] *= ep.col_scale;
  }
  if (ep.act == 1) {
    // pre-activation -> out_bf16, gelu -> out2_bf16 (handled below through v / pre)
    float pre[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      pre[j] = masked ? 0.f : v[j];
      v[j] = masked ? 0.f : gelu_tanh_f(v[j]);
    }
    if (ep.out_bf16) {
      __nv_bfloat16* p = ep.out_bf16 + row * ep.ldo_bf16 + col;
      if (nvalid == 8) {
        uint4 u = {pack_bf16(pre[0], pre[1]), pack_bf16(pre[2], pre[3]), pack_bf16(pre[4], pre[5]),
                   pack_bf16(pre[6], pre[7])};
        *reinterpret_cast<uint4*>(p) = u;
      } else if (nvalid == 4) {
        uint2 u = {pack_bf16(pre[0], pre[1]), pack_bf16(pre[2], pre[3])};
        *reinterpret_cast<uint2*>(p) = u;
      }
    }
    if (ep.out2_bf16) {
      __nv_bfloat16* p = ep.out2_bf16 + row * ep.ldo2_bf16 + col;
      if (nvalid == 8) {
        uint4 u = {pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7])};
        *reinterpret_cast<uint4*>(p) = u;
      } else if (nvalid == 4) {
        uint2 u = {pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3])};
        *reinterpret_cast<uint2*>(p) = u;
      }
    }
    return;
  }
  if (ep.act == 2) {
    const __nv_bfloat16* a = ep.aux + row * ep.ldaux + col;
    if (nvalid == 8) {
      uint4 u = *reinterpret_cast<const uint4*>(a);
      float2 f0 = unpack_bf16(u.x), f1 = unpack_bf16(u.y), f2 = unpack_bf16(u.z), f3 = unpack_bf16(u.w);
      v[0] *= gelu_tanh_grad_f(f0.x), v[1] *= gelu_tanh_grad_f(f0.y);
      v[2] *= gelu_tanh_grad_f(f1.x), v[3] *= gelu_tanh_grad_f(f1.y);
      v[4] *= gelu_tanh_grad_f(f2.x), v[5] *= gelu_tanh_grad_f(f2.y);
      v[6] *= gelu_tanh_grad_f(f3.x), v[7] *= gelu_tanh_grad_f(f3.y);
    } else if (nvalid == 4) {
      uint2 u = *reinterpret_cast<const uint2*>(a);
      float2 f0 = unpack_bf16(u.x), f1 = unpack_bf16(u.y);
      v[0] *= gelu_tanh_grad_f(f0.x), v[1] *= gelu_tanh_grad_f(f0.y);
      v[2] *= gelu_tanh_grad_f(f1.x), v[3] *= gelu_tanh_grad_f(f1.y);
    }
  }
  if (ep.pe && !masked) {
    const float* pe = ep.pe + static_cast<long>(ep.pe_idx[rp - ep.row_skip]) * ep.ldpe + col;
#pragma unroll
    for (int h = 0; h < 2; ++h)
      if (4 * h < nvalid) {
        float4 b = *reinterpret_cast<const float4*>(pe + 4 * h);
        v[4 * h + 0] += b.x, v[4 * h + 1] += b.y, v[4 * h + 2] += b.z, v[4 * h + 3] += b.w;
      }
  }
  if (ep.residual) {
    const float* r = ep.residual + row * ep.ldr + col;
#pragma unroll
    for (int h = 0; h < 2; ++h)
      if (4 * h < nvalid) {
        float4 b = *reinterpret_cast<const float4*>(r + 4 * h);
        v[4 * h + 0] += b.x, v[4 * h + 1] += b.y, v[4 * h + 2] += b.z, v[4 * h + 3] += b.w;
      }
  }
  if (masked) {
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = 0.f;
  }
  if (ep.out_f32) {
    float* p = ep.out_f32 + row * ep.ldo_f32 + col;
#pragma unroll
    for (int h = 0; h < 2; ++h)
      if (4 * h < nvalid)
        store_f32x4(p + 4 * h, make_float4(v[4 * h], v[4 * h + 1], v[4 * h + 2], v[4 * h + 3]), ep.accumulate != 0);
  }
  if (ep.out_bf16) {
    __nv_bfloat16* p = ep.out_bf16 + row * ep.ldo_bf16 + col;
    if (nvalid == 8) {
      uint4 u = {pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7])};
      *reinterpret_cast<uint4*>(p) = u;
    } else if (nvalid == 4) {
      uint2 u = {pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3])};
      *reinterpret_cast<uint2*>(p) = u;
    }
  }
}

template <int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmShape sh,
            const EpiParams ep) {
  using Cfg = GemmCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B tiles need 1024-byte alignment; align by hand.
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * Cfg::A_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES);
  uint64_t* full = bars;
  uint64_t* empty = bars + STAGES;
  uint64_t* tfull = bars + 2 * STAGES;
  uint64_t* tempty = bars + 2 * STAGES + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && elect_one()) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && elect_one()) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], 128);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<Cfg::TMEM_COLS>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  const int tiles_mn = sh.num_m * sh.num_n;
  const int total_tiles = tiles_mn * sh.splits;

  if (warp == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int split = t / tiles_mn;
        const int rem = t - split * tiles_mn;
        const int m_blk = rem / sh.num_n, n_blk = rem - m_blk * sh.num_n;
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
    const int q = warp - kEpiWarp0;  // TMEM lane quarter == warp % 4
    int iter = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++iter) {
      const int split = t / tiles_mn;
      const int rem = t - split * tiles_mn;
      const int m_blk = rem / sh.num_n, n_blk = rem - m_blk * sh.num_n;
      const int acc = iter & 1;
      const uint32_t acc_phase = (iter >> 1) & 1;
      mbar_wait(&tfull[acc], acc_phase);
      tc_fence_after();
      const long row = static_cast<long>(m_blk) * BM + q * 32 + lane;
      const bool row_ok = row < sh.M;
      int rp = 0;
      bool masked = false;
      if (ep.row_period > 0) {
        rp = static_cast<int>(row % ep.row_period);
        masked = rp < ep.row_skip;
      }
      const bool do_store = row_ok && !(masked && ep.row_mask_mode == 2);
      const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN;
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        const int col0 = n_blk * BN + c0;
        if (col0 >= sh.N) break;  // warp-uniform
        uint32_t r[32];
        tmem_ld_32x32(t_addr + c0, r);
        tmem_ld_wait();
        if (do_store) {
#pragma unroll
          for (int j8 = 0; j8 < 4; ++j8) {
            const int col = col0 + j8 * 8;
            int nvalid = sh.N - col;
            nvalid = nvalid >= 8 ? 8 : (nvalid >= 4 ? 4 : 0);
            if (nvalid > 0) {
              float v[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[j8 * 8 + j]);
              epilogue8(ep, v, row, col, nvalid, rp, masked && ep.row_mask_mode == 1);
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&tempty[acc]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
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

// 2-D bf16 tensor [outer, inner] row-major with row stride ld elements, SWIZZLE_128B boxes.
static int make_tmap(CUtensorMap* map, const void* ptr, long inner, long outer, long ld, int box_inner,
                     int box_outer) {
  auto fn = get_encode_fn();
  JM_CHECK(fn != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  JM_CHECK((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "GEMM operand pointer must be 16-byte aligned");
  JM_CHECK((ld * 2) % 16 == 0, "GEMM operand row stride must be a multiple of 8 elements (got %ld)", ld);
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(inner), static_cast<cuuint64_t>(outer)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 2};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(box_inner), static_cast<cuuint32_t>(box_outer)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  JM_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (inner=%ld outer=%ld ld=%ld)",
           static_cast<int>(r), inner, outer, ld);
  return 0;
}

template <int BN, bool A_MN, bool B_MN>
static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmShape& sh, const EpiParams& ep,
                       int grid, cudaStream_t stream) {
  auto kern = gemm_kernel<BN, A_MN, B_MN>;
  static bool attr_set = false;  // per template instantiation; benign race (idempotent)
  if (!attr_set) {
    JM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, GemmCfg<BN>::SMEM_BYTES));
    attr_set = true;
  }
  kern<<<grid, kGemmThreads, GemmCfg<BN>::SMEM_BYTES, stream>>>(tmA, tmB, sh, ep);
  JM_AFTER_LAUNCH();
  return 0;
}

template <int BN>
static int dispatch_major(bool a_mn, bool b_mn, const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmShape& sh,
                          const EpiParams& ep, int grid, cudaStream_t stream) {
  if (!a_mn && !b_mn) return launch_gemm<BN, false, false>(tmA, tmB, sh, ep, grid, stream);
  if (!a_mn && b_mn) return launch_gemm<BN, false, true>(tmA, tmB, sh, ep, grid, stream);
  if (a_mn && !b_mn) return launch_gemm<BN, true, false>(tmA, tmB, sh, ep, grid, stream);
  return launch_gemm<BN, true, true>(tmA, tmB, sh, ep, grid, stream);
}

}  // namespace jm

extern "C" int jm_gemm(const jm_gemm_args* a, void* stream_) {
  using namespace jm;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  JM_CHECK(a != nullptr, "jm_gemm: null args");
  JM_CHECK(a->M > 0 && a->N > 0 && a->K > 0, "jm_gemm: empty problem M=%d N=%d K=%d", a->M, a->N, a->K);
  JM_CHECK(a->N % 4 == 0, "jm_gemm: N must be a multiple of 4 (got %d)", a->N);
  JM_CHECK(a->A && a->B, "jm_gemm: null operand");
  JM_CHECK(a->out_f32 || a->out_bf16 || a->out2_bf16, "jm_gemm: no output");
  JM_CHECK(a->act != 2 || a->aux, "jm_gemm: act=2 needs aux");
  JM_CHECK(!a->pe || (a->pe_idx && a->row_period > 0), "jm_gemm: pe needs pe_idx and row_period");
  JM_CHECK(!a->out_f32 || a->ldo_f32 % 4 == 0, "jm_gemm: ldo_f32 must be a multiple of 4");
  JM_CHECK(!a->out_bf16 || a->ldo_bf16 % 4 == 0, "jm_gemm: ldo_bf16 must be a multiple of 4");
  JM_CHECK(!a->out2_bf16 || a->ldo2_bf16 % 4 == 0, "jm_gemm: ldo2_bf16 must be a multiple of 4");
  JM_CHECK(!a->residual || a->ldr % 4 == 0, "jm_gemm: ldr must be a multiple of 4");
  JM_CHECK(!a->aux || a->ldaux % 4 == 0, "jm_gemm: ldaux must be a multiple of 4");

  const int sms = sm_count();
  JM_CHECK(sms > 0, "jm_gemm: no CUDA device");
  const int num_m = (a->M + BM - 1) / BM;
  const int kblocks = (a->K + BK - 1) / BK;
  const bool can_split = a->accumulate && a->out_f32 && !a->out_bf16 && !a->out2_bf16 && !a->bias &&
                         !a->residual && !a->pe && a->act == 0;
  int bn = a->block_n;
  if (bn == 0) {
    // Largest tile that still gives every SM work; 256-wide tiles halve the shared-memory
    // operand traffic per MMA, which is what a 1-CTA tcgen05.mma needs to run at full rate.
    bn = 64;
    const int cand[3] = {256, 128, 64};
    for (int i = 0; i < 3; ++i) {
      const int c = cand[i];
      if (c > 64 && a->N < c) continue;
      const long tiles = static_cast<long>(num_m) * ((a->N + c - 1) / c);
      const long reach = can_split ? tiles * kblocks : tiles;
      if (reach >= sms || c == 64) {
        bn = c;
        break;
      }
    }
  }
  JM_CHECK(bn == 64 || bn == 128 || bn == 256, "jm_gemm: block_n must be 64, 128 or 256");
  const int num_n = (a->N + bn - 1) / bn;
  int splits = a->splits;
  if (splits <= 0) {
    splits = 1;
    if (can_split) {
      const long tiles = static_cast<long>(num_m) * num_n;
      splits = static_cast<int>(sms / tiles);
      if (splits < 1) splits = 1;
      if (splits > kblocks / 2) splits = kblocks / 2 > 0 ? kblocks / 2 : 1;
    }
  }
  JM_CHECK(splits == 1 || can_split, "jm_gemm: split-K needs a pure accumulate epilogue");
  JM_CHECK(splits <= kblocks, "jm_gemm: more splits (%d) than k-blocks (%d)", splits, kblocks);

  CUtensorMap tmA, tmB;
  if (!a->a_mn_major) {
    if (make_tmap(&tmA, a->A, a->K, a->M, a->lda, BK, BM)) return 1;
  } else {
    if (make_tmap(&tmA, a->A, a->M, a->K, a->lda, 64, BK)) return 1;
  }
  if (!a->b_mn_major) {
    if (make_tmap(&tmB, a->B, a->K, a->N, a->ldb, BK, bn)) return 1;
  } else {
    if (make_tmap(&tmB, a->B, a->N, a->K, a->ldb, 64, BK)) return 1;
  }

  GemmShape sh{a->M, a->N, a->K, num_m, num_n, splits, kblocks};
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

  const long total = static_cast<long>(num_m) * num_n * splits;
  const int grid = static_cast<int>(total < sms ? total : sms);
  const bool amn = a->a_mn_major != 0, bmn = a->b_mn_major != 0;
  if (bn == 256) return dispatch_major<256>(amn, bmn, tmA, tmB, sh, ep, grid, stream);
  if (bn == 128) return dispatch_major<128>(amn, bmn, tmA, tmB, sh, ep, grid, stream);
  return dispatch_major<64>(amn, bmn, tmA, tmB, sh, ep, grid, stream);
}
