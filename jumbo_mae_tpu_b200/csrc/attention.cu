// Fused short-sequence attention, forward and backward, for the 49+J / 196+J token counts of
// the MAE encoder / decoder (src/modeling.py:136-137: softmax(q k^T) v per image and head).
//
// One CTA per (image, head); the whole K and V of that pair live in shared memory and every
// warp owns a 16-row strip (flash-attention-2 style: scores never leave registers, online
// softmax over 64-key chunks).  The backward recomputes the scores from the saved
// log-sum-exp: phase 1 (warp = 16 queries) produces dQ, phase 2 (warp = 16 keys) produces
// dK and dV, so no cross-warp reduction or atomics are needed.
//
// Tensor-core path: mma.sync.m16n8k16 bf16 with ldmatrix operands.  These tiles are 16..272
// rows by 32..80 columns -- far below a 128-row tcgen05 tile -- and the op is ~3 % of the
// step's FLOPs, so the legacy warp-level MMA is the better fit here.
#include "../../include/jumbo_mae_b200.h"
#include "common.cuh"

namespace jm {

// attention_tc.cu: the tcgen05 / TMEM kernels (head_dim 32 / 64, T <= 256); everything else runs below
bool attn_tc_supported(int T, int hd);
int attn_tc_fwd(const jm_bf16* qkv, long ldqkv, int batch, int T, int heads, int hd, jm_bf16* out, long ldo,
                float* lse, cudaStream_t stream);

bool attn_tc_bwd_supported(int T, int hd);
int attn_tc_bwd(const jm_bf16* qkv, long ldqkv, const jm_bf16* out, long ldo, const jm_bf16* dout, long lddo,
                const float* lse, int batch, int T, int heads, int hd, float q_scale, jm_bf16* dqkv, long lddqkv,
                float* dbias, cudaStream_t stream);

constexpr float kLog2e = 1.4426950408889634f;

__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, "
      "{%0, %1, %2, %3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// Address helpers (element offsets inside a [rows][LDS] bf16 tile), lane-dependent:
//   A operand 16x16 at (r0, k0), row-major                      -> a_off
//   B operand from [n][k] storage, two n-tiles (n0, n0+8), k0   -> bn_off   (plain ldmatrix)
//   B operand from [k][n] storage, k0 (16 rows), n-tiles n0,n0+8 -> bt_off  (ldmatrix.trans)
template <int LDS>
__device__ __forceinline__ int a_off(int lane, int r0, int k0) {
  return (r0 + (lane & 7) + ((lane >> 3) & 1) * 8) * LDS + k0 + (lane >> 4) * 8;
}
template <int LDS>
__device__ __forceinline__ int bn_off(int lane, int n0, int k0) {
  return (n0 + (lane & 7) + (lane >> 4) * 8) * LDS + k0 + ((lane >> 3) & 1) * 8;
}
template <int LDS>
__device__ __forceinline__ int bt_off(int lane, int k0, int n0) {
  return (k0 + (lane & 7) + ((lane >> 3) & 1) * 8) * LDS + n0 + (lane >> 4) * 8;
}

// Cooperative asynchronous load (cp.async, 16 bytes per request, zero-fill for rows >= T) of rows
// [0, rows_pad) x HD bf16 of one head into a padded smem tile.  All requests of a CTA are in flight
// together; cp_async_wait_all() + __syncthreads() publishes them.
template <int HD, int LDS>
__device__ __forceinline__ void load_head(const __nv_bfloat16* __restrict__ src, long ld, int T, int rows_pad,
                                          __nv_bfloat16* dst) {
  constexpr int CH = HD / 8;
  for (int idx = threadIdx.x; idx < rows_pad * CH; idx += blockDim.x) {
    const int row = idx / CH, ch = idx - row * CH;
    const bool ok = row < T;
    const __nv_bfloat16* g = src + (ok ? row : 0) * ld + ch * 8;
    const uint32_t sa = smem_u32(dst + row * LDS + ch * 8);
    const int bytes = ok ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(sa), "l"(g), "r"(bytes) : "memory");
  }
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
}

__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// One chunk of NT*8 keys for a warp's 16 query rows: S = Q K^T, online softmax (log2 domain), O += P V.
// TAIL: keys >= T exist in this chunk and must be masked (only ever the last chunk).
template <int HD, int NT, bool TAIL>
__device__ __forceinline__ void fwd_chunk(const uint32_t (&qf)[HD / 16][4], uint32_t sKa, uint32_t sVa, int kc, int T,
                                          int lane, float (&o)[HD / 8][4], float (&m)[2], float (&l)[2]) {
  constexpr int LDS = HD + 8;
  const int t4 = lane & 3;
  float s[NT][4];
#pragma unroll
  for (int i = 0; i < NT; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
#pragma unroll
  for (int ks = 0; ks < HD / 16; ++ks) {
#pragma unroll
    for (int np = 0; np < NT / 2; ++np) {
      uint32_t kb[4];
      ldsm_x4(kb, sKa + 2 * bn_off<LDS>(lane, kc + np * 16, ks * 16));
      mma_bf16(s[2 * np], qf[ks], kb[0], kb[1]);
      mma_bf16(s[2 * np + 1], qf[ks], kb[2], kb[3]);
    }
  }
  float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
  for (int nt = 0; nt < NT; ++nt)
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      if (TAIL) {
        const int key = kc + nt * 8 + 2 * t4 + (e & 1);
        if (key >= T) s[nt][e] = -INFINITY;
      }
      mx[e >> 1] = fmaxf(mx[e >> 1], s[nt][e]);
    }
  float alpha[2], mneg[2];
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
    mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
    const float mn = fmaxf(m[r], mx[r] * kLog2e);  // running max of s * log2(e)
    alpha[r] = ex2f(m[r] - mn);
    m[r] = mn;
    mneg[r] = -mn;
    l[r] *= alpha[r];
  }
  // packed f32x2 arithmetic on the (e0, e1) / (e2, e3) pairs, which share a row: the kernel is bound by
  // instruction issue, and one FFMA2 slot does two of the scale-and-shift / sum / rescale operations
  {
    const f32x2 k2 = pack2(kLog2e, kLog2e);
    const f32x2 mn2[2] = {pack2(mneg[0], mneg[0]), pack2(mneg[1], mneg[1])};
    f32x2 l2[2] = {pack2(0.f, 0.f), pack2(0.f, 0.f)};
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        float x0, x1;
        unpack2(fma2(pack2(s[nt][2 * r], s[nt][2 * r + 1]), k2, mn2[r]), x0, x1);
        const float p0 = ex2f(x0), p1 = ex2f(x1);
        s[nt][2 * r] = p0, s[nt][2 * r + 1] = p1;
        l2[r] = add2(l2[r], pack2(p0, p1));
      }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      float a, b;
      unpack2(l2[r], a, b);
      l[r] += a + b;
    }
    const f32x2 al2[2] = {pack2(alpha[0], alpha[0]), pack2(alpha[1], alpha[1])};
#pragma unroll
    for (int i = 0; i < HD / 8; ++i)
#pragma unroll
      for (int r = 0; r < 2; ++r) unpack2(mul2(pack2(o[i][2 * r], o[i][2 * r + 1]), al2[r]), o[i][2 * r], o[i][2 * r + 1]);
  }
#pragma unroll
  for (int kk = 0; kk < NT / 2; ++kk) {
    uint32_t a[4];
    a[0] = pack_bf16(s[2 * kk][0], s[2 * kk][1]);
    a[1] = pack_bf16(s[2 * kk][2], s[2 * kk][3]);
    a[2] = pack_bf16(s[2 * kk + 1][0], s[2 * kk + 1][1]);
    a[3] = pack_bf16(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
    for (int np = 0; np < HD / 16; ++np) {
      uint32_t vb[4];
      ldsm_x4_t(vb, sVa + 2 * bt_off<LDS>(lane, kc + kk * 16, np * 16));
      mma_bf16(o[2 * np], a, vb[0], vb[1]);
      mma_bf16(o[2 * np + 1], a, vb[2], vb[3]);
    }
  }
}

#ifndef JM_ATTN_FWD32_CTAS
#define JM_ATTN_FWD32_CTAS 3
#endif
#ifndef JM_ATTN_FWD64_CTAS
#define JM_ATTN_FWD64_CTAS 2  // measured: 79 -> 61 us at T=52 (128 registers, twice the resident CTAs)
#endif
#ifndef JM_ATTN_BWD64_CTAS
#define JM_ATTN_BWD64_CTAS 2  // 180 -> 170 us
#endif
template <int HD>
__global__ void __launch_bounds__(256, HD == 32 ? JM_ATTN_FWD32_CTAS : (HD == 64 ? JM_ATTN_FWD64_CTAS : 1))
attn_fwd_kernel(const __nv_bfloat16* __restrict__ qkv, long ldqkv, int T, int heads,
                                __nv_bfloat16* __restrict__ out, long ldo, float* __restrict__ lse) {
  pdl_wait();  // programmatic dependent launch: nothing above touches global memory
  pdl_launch_dependents();
  constexpr int LDS = HD + 8;
  extern __shared__ __align__(16) uint8_t smem_raw[];
  __nv_bfloat16* sK = reinterpret_cast<__nv_bfloat16*>(smem_raw);
  const int Tk = (T + 15) / 16 * 16;  // keys / queries padded to the 16-row MMA granule (zero rows)
  const int nstrips = Tk / 16, nwarps = blockDim.x >> 5;
  __nv_bfloat16* sV = sK + Tk * LDS;
  __nv_bfloat16* sQ = sV + Tk * LDS;
  const int b = blockIdx.x / heads, h = blockIdx.x - b * heads;
  const int d = heads * HD;
  const __nv_bfloat16* base = qkv + static_cast<long>(b) * T * ldqkv + h * HD;
  load_head<HD, LDS>(base, ldqkv, T, Tk, sQ);
  load_head<HD, LDS>(base + d, ldqkv, T, Tk, sK);
  load_head<HD, LDS>(base + 2 * d, ldqkv, T, Tk, sV);
  cp_async_wait_all();
  __syncthreads();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t4 = lane & 3;
  const uint32_t sQa = smem_u32(sQ), sKa = smem_u32(sK), sVa = smem_u32(sV);
  const int nfull = Tk / 64, rem = Tk - nfull * 64;
  const bool ragged = T != Tk;

  for (int strip = warp; strip < nstrips; strip += nwarps) {
    const int row0 = strip * 16;
    uint32_t qf[HD / 16][4];
#pragma unroll
    for (int ks = 0; ks < HD / 16; ++ks) ldsm_x4(qf[ks], sQa + 2 * a_off<LDS>(lane, row0, ks * 16));
    float o[HD / 8][4];
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m[2] = {-INFINITY, -INFINITY}, l[2] = {0.f, 0.f};

    for (int c = 0; c < nfull; ++c) {
      if (rem == 0 && c == nfull - 1 && ragged) fwd_chunk<HD, 8, true>(qf, sKa, sVa, c * 64, T, lane, o, m, l);
      else fwd_chunk<HD, 8, false>(qf, sKa, sVa, c * 64, T, lane, o, m, l);
    }
    if (rem == 16) fwd_chunk<HD, 2, true>(qf, sKa, sVa, nfull * 64, T, lane, o, m, l);
    else if (rem == 32) fwd_chunk<HD, 4, true>(qf, sKa, sVa, nfull * 64, T, lane, o, m, l);
    else if (rem == 48) fwd_chunk<HD, 6, true>(qf, sKa, sVa, nfull * 64, T, lane, o, m, l);

#pragma unroll
    for (int r = 0; r < 2; ++r) {
      l[r] += __shfl_xor_sync(0xffffffffu, l[r], 1);
      l[r] += __shfl_xor_sync(0xffffffffu, l[r], 2);
      const int row = row0 + g + 8 * r;
      if (row < T) {
        const float inv = 1.0f / l[r];
        __nv_bfloat16* dst = out + (static_cast<long>(b) * T + row) * ldo + h * HD + 2 * t4;
#pragma unroll
        for (int i = 0; i < HD / 8; ++i)
          *reinterpret_cast<uint32_t*>(dst + i * 8) = pack_bf16(o[i][2 * r] * inv, o[i][2 * r + 1] * inv);
        if (t4 == 0) lse[(static_cast<long>(b) * heads + h) * T + row] = m[r] * 0.6931471805599453f + logf(l[r]);
      }
    }
  }
}

// delta[b, h, t] = sum_c dO[b*T+t, h*hd+c] * O[b*T+t, h*hd+c]  (one thread per (token, head))
__global__ void attn_delta_kernel(const __nv_bfloat16* __restrict__ out, long ldo,
                                  const __nv_bfloat16* __restrict__ dout, long lddo, int batch, int T, int heads,
                                  int hd, float* __restrict__ delta) {
  const long idx = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long total = static_cast<long>(batch) * T * heads;
  if (idx >= total) return;
  const int h = static_cast<int>(idx % heads);
  const long tok = idx / heads;
  const __nv_bfloat16* o = out + tok * ldo + h * hd;
  const __nv_bfloat16* g = dout + tok * lddo + h * hd;
  float acc = 0.f;
  for (int c = 0; c < hd; c += 8) {
    const uint4 ov = *reinterpret_cast<const uint4*>(o + c);
    const uint4 dv = *reinterpret_cast<const uint4*>(g + c);
    const uint32_t oo[4] = {ov.x, ov.y, ov.z, ov.w}, dd[4] = {dv.x, dv.y, dv.z, dv.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 a = unpack_bf16(oo[j]), b = unpack_bf16(dd[j]);
      acc += a.x * b.x + a.y * b.y;
    }
  }
  const long b = tok / T;
  const int t = static_cast<int>(tok - b * T);
  delta[(b * heads + h) * T + t] = acc;
}

// Backward phase 1, one chunk of NT*8 keys for a warp's 16 queries: dQ += dS K with
// P = exp(S - lse), dS = P * (dP - D), dP = dO V^T.  Lr2 = lse * log2(e).
template <int HD, int NT, bool TAIL>
__device__ __forceinline__ void bwd_q_chunk(const uint32_t (&qf)[HD / 16][4], const uint32_t (&dof)[HD / 16][4],
                                            uint32_t sKa, uint32_t sVa, int kc, int T, int lane,
                                            const float (&Lr2)[2], const float (&Dr)[2], float (&dq)[HD / 8][4]) {
  constexpr int LDS = HD + 8;
  const int t4 = lane & 3;
  float s[NT][4], dp[NT][4];
#pragma unroll
  for (int i = 0; i < NT; ++i)
#pragma unroll
    for (int e = 0; e < 4; ++e) s[i][e] = 0.f, dp[i][e] = 0.f;
#pragma unroll
  for (int ks = 0; ks < HD / 16; ++ks) {
#pragma unroll
    for (int np = 0; np < NT / 2; ++np) {
      uint32_t kb[4], vb[4];
      ldsm_x4(kb, sKa + 2 * bn_off<LDS>(lane, kc + np * 16, ks * 16));
      mma_bf16(s[2 * np], qf[ks], kb[0], kb[1]);
      mma_bf16(s[2 * np + 1], qf[ks], kb[2], kb[3]);
      ldsm_x4(vb, sVa + 2 * bn_off<LDS>(lane, kc + np * 16, ks * 16));
      mma_bf16(dp[2 * np], dof[ks], vb[0], vb[1]);
      mma_bf16(dp[2 * np + 1], dof[ks], vb[2], vb[3]);
    }
  }
  {  // packed f32x2 arithmetic on the (e0, e1) / (e2, e3) pairs (same query row)
    const f32x2 k2 = pack2(kLog2e, kLog2e);
    const f32x2 nl2[2] = {pack2(-Lr2[0], -Lr2[0]), pack2(-Lr2[1], -Lr2[1])};
    const f32x2 nd2[2] = {pack2(-Dr[0], -Dr[0]), pack2(-Dr[1], -Dr[1])};
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        float x0, x1;
        unpack2(fma2(pack2(s[nt][2 * r], s[nt][2 * r + 1]), k2, nl2[r]), x0, x1);
        float p0 = ex2f(x0), p1 = ex2f(x1);
        if (TAIL) {
          const int key = kc + nt * 8 + 2 * t4;
          if (key >= T) p0 = 0.f;
          if (key + 1 >= T) p1 = 0.f;
        }
        const f32x2 ds = mul2(pack2(p0, p1), add2(pack2(dp[nt][2 * r], dp[nt][2 * r + 1]), nd2[r]));  // dS
        unpack2(ds, s[nt][2 * r], s[nt][2 * r + 1]);
      }
  }
#pragma unroll
  for (int kk = 0; kk < NT / 2; ++kk) {
    uint32_t a[4];
    a[0] = pack_bf16(s[2 * kk][0], s[2 * kk][1]);
    a[1] = pack_bf16(s[2 * kk][2], s[2 * kk][3]);
    a[2] = pack_bf16(s[2 * kk + 1][0], s[2 * kk + 1][1]);
    a[3] = pack_bf16(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
    for (int np = 0; np < HD / 16; ++np) {
      uint32_t kb[4];
      ldsm_x4_t(kb, sKa + 2 * bt_off<LDS>(lane, kc + kk * 16, np * 16));
      mma_bf16(dq[2 * np], a, kb[0], kb[1]);
      mma_bf16(dq[2 * np + 1], a, kb[2], kb[3]);
    }
  }
}

// Backward phase 2, one chunk of NT*8 queries for a warp's 16 keys (transposed problem):
// dV += P^T dO, dK += dS^T Q.  Padded queries carry lse = +big (P = 0) and zero dO rows.
template <int HD, int NT>
__device__ __forceinline__ void bwd_kv_chunk(const uint32_t (&kf)[HD / 16][4], const uint32_t (&vf)[HD / 16][4],
                                             uint32_t sQa, uint32_t sdOa, const float* sL2, const float* sD, int qc,
                                             int lane, float (&dk)[HD / 8][4], float (&dv)[HD / 8][4]) {
  constexpr int LDS = HD + 8;
  const int t4 = lane & 3;
  float st[NT][4], dpt[NT][4];
#pragma unroll
  for (int i = 0; i < NT; ++i)
#pragma unroll
    for (int e = 0; e < 4; ++e) st[i][e] = 0.f, dpt[i][e] = 0.f;
#pragma unroll
  for (int ks = 0; ks < HD / 16; ++ks) {
#pragma unroll
    for (int np = 0; np < NT / 2; ++np) {
      uint32_t qb[4], ob[4];
      ldsm_x4(qb, sQa + 2 * bn_off<LDS>(lane, qc + np * 16, ks * 16));
      mma_bf16(st[2 * np], kf[ks], qb[0], qb[1]);
      mma_bf16(st[2 * np + 1], kf[ks], qb[2], qb[3]);
      ldsm_x4(ob, sdOa + 2 * bn_off<LDS>(lane, qc + np * 16, ks * 16));
      mma_bf16(dpt[2 * np], vf[ks], ob[0], ob[1]);
      mma_bf16(dpt[2 * np + 1], vf[ks], ob[2], ob[3]);
    }
  }
  uint32_t pa[NT / 2][4], dsa[NT / 2][4];
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) {
    const int q0 = qc + nt * 8 + 2 * t4;
    const float2 L2 = *reinterpret_cast<const float2*>(sL2 + q0);
    const float2 D2 = *reinterpret_cast<const float2*>(sD + q0);
    float p[4], ds[4];
    {  // packed f32x2: the pair (e0, e1) (and (e2, e3)) is the query pair (q0, q0 + 1)
      const f32x2 k2 = pack2(kLog2e, kLog2e), nl = pack2(-L2.x, -L2.y), nd = pack2(-D2.x, -D2.y);
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        float x0, x1;
        unpack2(fma2(pack2(st[nt][2 * r], st[nt][2 * r + 1]), k2, nl), x0, x1);
        p[2 * r] = ex2f(x0), p[2 * r + 1] = ex2f(x1);
        unpack2(mul2(pack2(p[2 * r], p[2 * r + 1]), add2(pack2(dpt[nt][2 * r], dpt[nt][2 * r + 1]), nd)), ds[2 * r],
                ds[2 * r + 1]);
      }
    }
    pa[nt >> 1][(nt & 1) * 2 + 0] = pack_bf16(p[0], p[1]);
    pa[nt >> 1][(nt & 1) * 2 + 1] = pack_bf16(p[2], p[3]);
    dsa[nt >> 1][(nt & 1) * 2 + 0] = pack_bf16(ds[0], ds[1]);
    dsa[nt >> 1][(nt & 1) * 2 + 1] = pack_bf16(ds[2], ds[3]);
  }
#pragma unroll
  for (int kk = 0; kk < NT / 2; ++kk) {
#pragma unroll
    for (int np = 0; np < HD / 16; ++np) {
      uint32_t ob[4], qb[4];
      ldsm_x4_t(ob, sdOa + 2 * bt_off<LDS>(lane, qc + kk * 16, np * 16));
      mma_bf16(dv[2 * np], pa[kk], ob[0], ob[1]);
      mma_bf16(dv[2 * np + 1], pa[kk], ob[2], ob[3]);
      ldsm_x4_t(qb, sQa + 2 * bt_off<LDS>(lane, qc + kk * 16, np * 16));
      mma_bf16(dk[2 * np], dsa[kk], qb[0], qb[1]);
      mma_bf16(dk[2 * np + 1], dsa[kk], qb[2], qb[3]);
    }
  }
}

#ifndef JM_ATTN_BWD_THREADS
#define JM_ATTN_BWD_THREADS 256
#endif
#ifndef JM_ATTN_BWD_CTAS
#define JM_ATTN_BWD_CTAS 2
#endif
// MINCTAS = resident CTAs per SM the register allocation is bounded for.  head_dim 64 has two instantiations: 2 CTAs
// (128 registers, a few spills) wins for the pre-training encoder's short sequences (T = 52: 180 -> 170 us), 1 CTA
// (207 registers, no spills) for long ones (fine-tune step, T = 199: the step 22.7 -> 21.9 ms); launch_bwd picks by T.
#ifndef JM_ATTN_BWD64_LONG_T
#define JM_ATTN_BWD64_LONG_T 96
#endif
template <int HD, int MINCTAS>
__global__ void __launch_bounds__(JM_ATTN_BWD_THREADS, MINCTAS)
attn_bwd_kernel(const __nv_bfloat16* __restrict__ qkv, long ldqkv,
                                const __nv_bfloat16* __restrict__ dout, long lddo, const float* __restrict__ lse,
                                const __nv_bfloat16* __restrict__ out, long ldo, int T, int heads, float q_scale,
                                __nv_bfloat16* __restrict__ dqkv, long lddqkv, float* __restrict__ dbias) {
  pdl_wait();  // programmatic dependent launch: nothing above touches global memory
  pdl_launch_dependents();
  constexpr int LDS = HD + 8;
  extern __shared__ __align__(16) uint8_t smem_raw[];
  const int Tk = (T + 15) / 16 * 16;
  __nv_bfloat16* sQ = reinterpret_cast<__nv_bfloat16*>(smem_raw);
  __nv_bfloat16* sK = sQ + Tk * LDS;
  __nv_bfloat16* sV = sK + Tk * LDS;
  __nv_bfloat16* sdO = sV + Tk * LDS;
  float* sL2 = reinterpret_cast<float*>(sdO + Tk * LDS);  // lse * log2(e)
  float* sD = sL2 + Tk;
  float* sBias = sD + Tk;  // [3][HD] column sums of dq, dk, dv over this (image, head)
  for (int i = threadIdx.x; i < 3 * HD; i += blockDim.x) sBias[i] = 0.f;
  const int b = blockIdx.x / heads, h = blockIdx.x - b * heads;
  const int d = heads * HD;
  const __nv_bfloat16* base = qkv + static_cast<long>(b) * T * ldqkv + h * HD;
  load_head<HD, LDS>(base, ldqkv, T, Tk, sQ);
  load_head<HD, LDS>(base + d, ldqkv, T, Tk, sK);
  load_head<HD, LDS>(base + 2 * d, ldqkv, T, Tk, sV);
  load_head<HD, LDS>(dout + static_cast<long>(b) * T * lddo + h * HD, lddo, T, Tk, sdO);
  // saved log-sum-exp; padded rows get lse = +big so that their P is exactly 0
  for (int row = threadIdx.x; row < Tk; row += blockDim.x) {
    const long i = (static_cast<long>(b) * heads + h) * T + row;
    sL2[row] = row < T ? lse[i] * kLog2e : 1e30f;
  }
  // D = rowsum(dO * O), one thread per row: the O row is requested from global memory now (its latency
  // overlaps the cp.async fills above), the dO row is read from shared memory once it has landed
  for (int row0 = 0; row0 < Tk; row0 += blockDim.x) {
    const int row = row0 + threadIdx.x;
    uint4 ov[HD / 8];
#pragma unroll
    for (int c = 0; c < HD / 8; ++c) ov[c] = make_uint4(0u, 0u, 0u, 0u);
    if (row < T) {
      const __nv_bfloat16* orow = out + (static_cast<long>(b) * T + row) * ldo + h * HD;
#pragma unroll
      for (int c = 0; c < HD / 8; ++c) ov[c] = *reinterpret_cast<const uint4*>(orow + c * 8);
    }
    if (row0 == 0) {  // first pass: this is where everybody waits for the tiles
      cp_async_wait_all();
      __syncthreads();
    }
    if (row < Tk) {
      float acc = 0.f;
      const __nv_bfloat16* drow = sdO + row * LDS;
#pragma unroll
      for (int c = 0; c < HD / 8; ++c) {
        const uint4 dv = *reinterpret_cast<const uint4*>(drow + c * 8);
        const uint32_t ow[4] = {ov[c].x, ov[c].y, ov[c].z, ov[c].w}, dw[4] = {dv.x, dv.y, dv.z, dv.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 a = unpack_bf16(ow[j]), bb = unpack_bf16(dw[j]);
          acc = fmaf(a.x, bb.x, acc);
          acc = fmaf(a.y, bb.y, acc);
        }
      }
      sD[row] = acc;  // padded rows: O was not read, 0
    }
  }
  __syncthreads();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t4 = lane & 3;
  const int nstrips = Tk / 16, nwarps = blockDim.x >> 5;
  const uint32_t sQa = smem_u32(sQ), sKa = smem_u32(sK), sVa = smem_u32(sV), sdOa = smem_u32(sdO);
  const int nfull = Tk / 32, rem = Tk - nfull * 32;  // rem in {0, 16}
  const bool ragged = T != Tk;

  // ---------------- phase 1: a warp's 16 queries -> dQ
  for (int strip = warp; strip < nstrips; strip += nwarps) {
    const int row0 = strip * 16;
    uint32_t qf[HD / 16][4], dof[HD / 16][4];
#pragma unroll
    for (int ks = 0; ks < HD / 16; ++ks) {
      ldsm_x4(qf[ks], sQa + 2 * a_off<LDS>(lane, row0, ks * 16));
      ldsm_x4(dof[ks], sdOa + 2 * a_off<LDS>(lane, row0, ks * 16));
    }
    const float Lr2[2] = {sL2[row0 + g], sL2[row0 + g + 8]};
    const float Dr[2] = {sD[row0 + g], sD[row0 + g + 8]};
    float dq[HD / 8][4];
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) dq[i][0] = dq[i][1] = dq[i][2] = dq[i][3] = 0.f;
    for (int c = 0; c < nfull; ++c) {
      if (rem == 0 && c == nfull - 1 && ragged)
        bwd_q_chunk<HD, 4, true>(qf, dof, sKa, sVa, c * 32, T, lane, Lr2, Dr, dq);
      else
        bwd_q_chunk<HD, 4, false>(qf, dof, sKa, sVa, c * 32, T, lane, Lr2, Dr, dq);
    }
    if (rem == 16) bwd_q_chunk<HD, 2, true>(qf, dof, sKa, sVa, nfull * 32, T, lane, Lr2, Dr, dq);
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int row = row0 + g + 8 * r;
      if (row < T) {
        __nv_bfloat16* dst = dqkv + (static_cast<long>(b) * T + row) * lddqkv + h * HD + 2 * t4;
#pragma unroll
        for (int i = 0; i < HD / 8; ++i)
          *reinterpret_cast<uint32_t*>(dst + i * 8) = pack_bf16(dq[i][2 * r] * q_scale, dq[i][2 * r + 1] * q_scale);
      }
    }
    if (dbias != nullptr) {  // padded query rows have dS == 0, so they add nothing
#pragma unroll
      for (int i = 0; i < HD / 8; ++i) {
        float c0 = dq[i][0] + dq[i][2], c1 = dq[i][1] + dq[i][3];
#pragma unroll
        for (int o = 4; o < 32; o <<= 1) {
          c0 += __shfl_xor_sync(0xffffffffu, c0, o);
          c1 += __shfl_xor_sync(0xffffffffu, c1, o);
        }
        if (g == 0) {
          atomicAdd(&sBias[i * 8 + 2 * t4], c0 * q_scale);
          atomicAdd(&sBias[i * 8 + 2 * t4 + 1], c1 * q_scale);
        }
      }
    }
  }

  // ---------------- phase 2: a warp's 16 keys -> dK, dV (transposed problem)
  for (int strip = warp; strip < nstrips; strip += nwarps) {
    const int row0 = strip * 16;
    uint32_t kf[HD / 16][4], vf[HD / 16][4];
#pragma unroll
    for (int ks = 0; ks < HD / 16; ++ks) {
      ldsm_x4(kf[ks], sKa + 2 * a_off<LDS>(lane, row0, ks * 16));
      ldsm_x4(vf[ks], sVa + 2 * a_off<LDS>(lane, row0, ks * 16));
    }
    float dk[HD / 8][4], dv[HD / 8][4];
#pragma unroll
    for (int i = 0; i < HD / 8; ++i)
#pragma unroll
      for (int e = 0; e < 4; ++e) dk[i][e] = 0.f, dv[i][e] = 0.f;
    for (int c = 0; c < nfull; ++c) bwd_kv_chunk<HD, 4>(kf, vf, sQa, sdOa, sL2, sD, c * 32, lane, dk, dv);
    if (rem == 16) bwd_kv_chunk<HD, 2>(kf, vf, sQa, sdOa, sL2, sD, nfull * 32, lane, dk, dv);
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int row = row0 + g + 8 * r;
      if (row < T) {
        __nv_bfloat16* dst = dqkv + (static_cast<long>(b) * T + row) * lddqkv + h * HD + 2 * t4;
#pragma unroll
        for (int i = 0; i < HD / 8; ++i) {
          *reinterpret_cast<uint32_t*>(dst + d + i * 8) = pack_bf16(dk[i][2 * r], dk[i][2 * r + 1]);
          *reinterpret_cast<uint32_t*>(dst + 2 * d + i * 8) = pack_bf16(dv[i][2 * r], dv[i][2 * r + 1]);
        }
      }
    }
    if (dbias != nullptr) {
      const float m0 = (row0 + g) < T ? 1.f : 0.f, m1 = (row0 + g + 8) < T ? 1.f : 0.f;  // padded key rows
#pragma unroll
      for (int i = 0; i < HD / 8; ++i) {
        float k0 = dk[i][0] * m0 + dk[i][2] * m1, k1 = dk[i][1] * m0 + dk[i][3] * m1;
        float v0 = dv[i][0] * m0 + dv[i][2] * m1, v1 = dv[i][1] * m0 + dv[i][3] * m1;
#pragma unroll
        for (int o = 4; o < 32; o <<= 1) {
          k0 += __shfl_xor_sync(0xffffffffu, k0, o);
          k1 += __shfl_xor_sync(0xffffffffu, k1, o);
          v0 += __shfl_xor_sync(0xffffffffu, v0, o);
          v1 += __shfl_xor_sync(0xffffffffu, v1, o);
        }
        if (g == 0) {
          atomicAdd(&sBias[HD + i * 8 + 2 * t4], k0);
          atomicAdd(&sBias[HD + i * 8 + 2 * t4 + 1], k1);
          atomicAdd(&sBias[2 * HD + i * 8 + 2 * t4], v0);
          atomicAdd(&sBias[2 * HD + i * 8 + 2 * t4 + 1], v1);
        }
      }
    }
  }
  if (dbias != nullptr) {
    __syncthreads();
    for (int i = threadIdx.x; i < 3 * HD; i += blockDim.x) {
      const int part = i / HD, c = i - part * HD;
      atomicAdd(dbias + part * d + h * HD + c, sBias[i]);
    }
  }
}

static size_t fwd_smem(int T, int hd) {
  const int Tk = (T + 15) / 16 * 16;
  return static_cast<size_t>(3 * Tk) * (hd + 8) * 2;
}
static size_t bwd_smem(int T, int hd) {
  const int Tk = (T + 15) / 16 * 16;
  return static_cast<size_t>(4 * Tk) * (hd + 8) * 2 + static_cast<size_t>(2 * Tk) * 4 + 3 * hd * 4;
}

// Warps per CTA: one per 16-row strip when the register file allows it, otherwise the strips
// are dealt round-robin to fewer warps.  Each of the 4 SM sub-partitions holds 16384 registers
// and warps are spread evenly over them, so ceil(warps / 4) * 32 * regs must fit in 16384.
static int pick_warps(const void* kern, int T, int* warps) {
  cudaFuncAttributes fa;
  JM_CUDA(cudaFuncGetAttributes(&fa, kern));
  const int per_warp = (fa.numRegs + 7) / 8 * 8 * 32;
  int max_warps = (16384 / per_warp) * 4;
  if (max_warps > 8) max_warps = 8;  // small CTAs: several (image, head) pairs per SM overlap load and math
  JM_CHECK(max_warps >= 1, "attention kernel needs %d registers per thread", fa.numRegs);
  const int strips = (T + 15) / 16;
  const int rounds = (strips + max_warps - 1) / max_warps;
  *warps = (strips + rounds - 1) / rounds;
  return 0;
}

template <int HD>
static int launch_fwd(const jm_bf16* qkv, long ldqkv, int batch, int T, int heads, jm_bf16* out, long ldo,
                      float* lse, cudaStream_t stream) {
  const size_t smem = fwd_smem(T, HD);
  auto kern = attn_fwd_kernel<HD>;
  if (smem > 48 * 1024) JM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  JM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
  int warps = 0;
  if (pick_warps(reinterpret_cast<const void*>(kern), T, &warps)) return 1;
  const int threads = warps * 32;
  JM_LAUNCH_PDL(kern, batch * heads, threads, smem, stream, reinterpret_cast<const __nv_bfloat16*>(qkv), ldqkv, T, heads,
                                                 reinterpret_cast<__nv_bfloat16*>(out), ldo, lse);
  JM_AFTER_LAUNCH();
  return 0;
}

template <int HD, int MINCTAS>
static int launch_bwd_k(const jm_bf16* qkv, long ldqkv, const jm_bf16* out, long ldo, const jm_bf16* dout, long lddo,
                        const float* lse, float* delta, int batch, int T, int heads, float q_scale, jm_bf16* dqkv,
                        long lddqkv, float* dbias, cudaStream_t stream) {
  const size_t smem = bwd_smem(T, HD);
  auto kern = attn_bwd_kernel<HD, MINCTAS>;
  if (smem > 48 * 1024) JM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  JM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
  int warps = 0;
  if (pick_warps(reinterpret_cast<const void*>(kern), T, &warps)) return 1;
  const int threads = warps * 32;
  (void)delta;  // D = rowsum(dO * O) is computed inside the backward kernel
  JM_LAUNCH_PDL(kern, batch * heads, threads, smem, stream,
      reinterpret_cast<const __nv_bfloat16*>(qkv), ldqkv, reinterpret_cast<const __nv_bfloat16*>(dout), lddo, lse,
      reinterpret_cast<const __nv_bfloat16*>(out), ldo, T, heads, q_scale, reinterpret_cast<__nv_bfloat16*>(dqkv), lddqkv, dbias);
  JM_AFTER_LAUNCH();
  return 0;
}

template <int HD>
static int launch_bwd(const jm_bf16* qkv, long ldqkv, const jm_bf16* out, long ldo, const jm_bf16* dout, long lddo,
                      const float* lse, float* delta, int batch, int T, int heads, float q_scale, jm_bf16* dqkv,
                      long lddqkv, float* dbias, cudaStream_t stream) {
  if constexpr (HD == 32) {
    return launch_bwd_k<32, JM_ATTN_BWD_CTAS>(qkv, ldqkv, out, ldo, dout, lddo, lse, delta, batch, T, heads, q_scale,
                                              dqkv, lddqkv, dbias, stream);
  } else if constexpr (HD == 64) {
    if (T > JM_ATTN_BWD64_LONG_T)
      return launch_bwd_k<64, 1>(qkv, ldqkv, out, ldo, dout, lddo, lse, delta, batch, T, heads, q_scale, dqkv, lddqkv,
                                 dbias, stream);
    return launch_bwd_k<64, JM_ATTN_BWD64_CTAS>(qkv, ldqkv, out, ldo, dout, lddo, lse, delta, batch, T, heads, q_scale,
                                                dqkv, lddqkv, dbias, stream);
  } else {
    return launch_bwd_k<HD, 1>(qkv, ldqkv, out, ldo, dout, lddo, lse, delta, batch, T, heads, q_scale, dqkv, lddqkv,
                               dbias, stream);
  }
}

static int check_attn(const char* who, int batch, int T, int heads, int hd, long ld0, long ld1) {
  JM_CHECK(batch > 0 && heads > 0, "%s: empty problem", who);
  JM_CHECK(T >= 1 && T <= 512, "%s: sequence length %d outside [1, 512]", who, T);
  JM_CHECK(hd == 32 || hd == 64 || hd == 80, "%s: head_dim %d not in {32, 64, 80}", who, hd);
  JM_CHECK(ld0 % 8 == 0 && ld1 % 8 == 0, "%s: row strides must be multiples of 8 elements", who);
  return 0;
}

}  // namespace jm

extern "C" int jm_attention_fwd(const jm_bf16* qkv, long ldqkv, int batch, int T, int heads, int head_dim,
                                jm_bf16* out, long ldo, float* lse, void* stream_) {
  using namespace jm;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (check_attn("jm_attention_fwd", batch, T, heads, head_dim, ldqkv, ldo)) return 1;
  if (attn_tc_supported(T, head_dim)) return attn_tc_fwd(qkv, ldqkv, batch, T, heads, head_dim, out, ldo, lse, stream);
  JM_CHECK(fwd_smem(T, head_dim) <= 227 * 1024, "jm_attention_fwd: T=%d hd=%d does not fit shared memory", T,
           head_dim);
  if (head_dim == 32) return launch_fwd<32>(qkv, ldqkv, batch, T, heads, out, ldo, lse, stream);
  if (head_dim == 64) return launch_fwd<64>(qkv, ldqkv, batch, T, heads, out, ldo, lse, stream);
  return launch_fwd<80>(qkv, ldqkv, batch, T, heads, out, ldo, lse, stream);
}

extern "C" int jm_attention_bwd(const jm_bf16* qkv, long ldqkv, const jm_bf16* out, long ldo, const jm_bf16* dout,
                                long lddo, const float* lse, float* delta_ws, int batch, int T, int heads,
                                int head_dim, float q_scale, jm_bf16* dqkv, long lddqkv, float* dbias,
                                void* stream_) {
  using namespace jm;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (check_attn("jm_attention_bwd", batch, T, heads, head_dim, ldqkv, lddqkv)) return 1;
  JM_CHECK(ldo % 8 == 0 && lddo % 8 == 0, "jm_attention_bwd: row strides must be multiples of 8 elements");
  if (attn_tc_bwd_supported(T, head_dim))
    return attn_tc_bwd(qkv, ldqkv, out, ldo, dout, lddo, lse, batch, T, heads, head_dim, q_scale, dqkv, lddqkv, dbias,
                       stream);
  JM_CHECK(bwd_smem(T, head_dim) <= 227 * 1024, "jm_attention_bwd: T=%d hd=%d does not fit shared memory", T,
           head_dim);
  if (head_dim == 32)
    return launch_bwd<32>(qkv, ldqkv, out, ldo, dout, lddo, lse, delta_ws, batch, T, heads, q_scale, dqkv, lddqkv, dbias,
                          stream);
  if (head_dim == 64)
    return launch_bwd<64>(qkv, ldqkv, out, ldo, dout, lddo, lse, delta_ws, batch, T, heads, q_scale, dqkv, lddqkv, dbias,
                          stream);
  return launch_bwd<80>(qkv, ldqkv, out, ldo, dout, lddo, lse, delta_ws, batch, T, heads, q_scale, dqkv, lddqkv, dbias,
                          stream);
}
