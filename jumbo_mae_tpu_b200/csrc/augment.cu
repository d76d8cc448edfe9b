// GPU side of the pre-training recipe's augmentation: RandomResizedCrop (bicubic) + horizontal flip on decoded
// uint8 images (src/dataset.py:56-58,70-71: T.RandomResizedCrop(S, scale=(0.2, 1.0), interpolation=3) and
// T.RandomHorizontalFlip on PIL images, T.PILToTensor).  The crop boxes and flip bits are drawn on the host (the
// reference draws them in its DataLoader workers); this kernel family replaces the CPU resampling:
//   crop -> PIL.Image.resize((S, S), BICUBIC) -> optional FLIP_LEFT_RIGHT -> CHW
// bit for bit.  The arithmetic is Pillow's ImagingResample (src/libImaging/Resample.c): float64 bicubic coefficients
// with the filter support widened by the down-scaling factor (anti-aliasing), normalised, converted to 22-bit fixed
// point; a horizontal pass into a uint8 intermediate image and a vertical pass, each accumulating in int32 from
// 1 << 21 and clipping (acc >> 22) to [0, 255].  The float64 operations use the round-to-nearest intrinsics so that
// the compiler cannot contract them into FMAs (Pillow's x86-64 build has none).
//
// HBM-bound byte work: per image, reads the crop once (h w 3 bytes), writes and re-reads the intermediate
// (h S 3 bytes), writes S S 3 bytes.
#include "../../include/jumbo_mae_b200.h"
#include "common.cuh"

namespace jm {
namespace {

constexpr int kPrecisionBits = 32 - 8 - 2;

__device__ __forceinline__ double bicubic_filter_rn(double x) {
  const double a = -0.5;
  if (x < 0.0) x = -x;
  if (x < 1.0)  // ((a + 2) x - (a + 3)) x x + 1
    return __dadd_rn(__dmul_rn(__dmul_rn(__dadd_rn(__dmul_rn(a + 2.0, x), -(a + 3.0)), x), x), 1.0);
  if (x < 2.0)  // (((x - 5) x + 8) x - 4) a
    return __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(x, -5.0), x), 8.0), x), -4.0), a);
  return 0.0;
}

// Resample.c precompute_coeffs + normalize_coeffs_8bpc for box (0, in_size) -> out_size; one thread per output index.
// coeffs: [2 axes][batch][S][ksize] int32, bounds: [2 axes][batch][S][2] int32 (first source index, count).
__global__ void rrc_coeffs_kernel(const int* __restrict__ boxes, int batch, int S, int ksize, int* __restrict__ coeffs,
                                  int* __restrict__ bounds) {
  const int b = blockIdx.y, axis = blockIdx.z;  // axis 0: horizontal (crop width), 1: vertical (crop height)
  const int xx = blockIdx.x * blockDim.x + threadIdx.x;
  if (xx >= S) return;
  const int in_size = axis == 0 ? boxes[b * 4 + 3] : boxes[b * 4 + 2];
  const double scale = __ddiv_rn(static_cast<double>(in_size), static_cast<double>(S));
  const double filterscale = scale < 1.0 ? 1.0 : scale;
  const double support = __dmul_rn(2.0, filterscale);
  const double ss = __ddiv_rn(1.0, filterscale);
  const double center = __dadd_rn(0.0, __dmul_rn(static_cast<double>(xx) + 0.5, scale));
  int xmin = static_cast<int>(__dadd_rn(__dadd_rn(center, -support), 0.5));
  if (xmin < 0) xmin = 0;
  int xmax = static_cast<int>(__dadd_rn(__dadd_rn(center, support), 0.5));
  if (xmax > in_size) xmax = in_size;
  xmax -= xmin;
  if (xmax > ksize) xmax = ksize;  // cannot happen for the ksize the host computes; keeps the writes in bounds
  int* k = coeffs + ((static_cast<long>(axis) * batch + b) * S + xx) * ksize;
  double ww = 0.0;
  for (int x = 0; x < xmax; ++x)
    ww = __dadd_rn(ww, bicubic_filter_rn(__dmul_rn(__dadd_rn(__dadd_rn(static_cast<double>(x + xmin), -center), 0.5), ss)));
  for (int x = 0; x < xmax; ++x) {
    double w = bicubic_filter_rn(__dmul_rn(__dadd_rn(__dadd_rn(static_cast<double>(x + xmin), -center), 0.5), ss));
    if (ww != 0.0) w = __ddiv_rn(w, ww);
    const double f = __dmul_rn(w, static_cast<double>(1 << kPrecisionBits));
    k[x] = static_cast<int>(w < 0.0 ? __dadd_rn(-0.5, f) : __dadd_rn(0.5, f));
  }
  for (int x = xmax; x < ksize; ++x) k[x] = 0;
  int* bd = bounds + ((static_cast<long>(axis) * batch + b) * S + xx) * 2;
  bd[0] = xmin, bd[1] = xmax;
}

__device__ __forceinline__ uint8_t clip8(int v) {
  v >>= kPrecisionBits;
  return static_cast<uint8_t>(v < 0 ? 0 : (v > 255 ? 255 : v));
}

// horizontal pass: every row of the crop -> tmp [batch][max_h][S][3]; thread = (row, xx), three channels
__global__ void rrc_horizontal_kernel(const uint8_t* __restrict__ src, const long* __restrict__ src_off,
                                      const int* __restrict__ src_hw, const int* __restrict__ boxes, int batch, int S,
                                      int ksize, int max_h, const int* __restrict__ coeffs,
                                      const int* __restrict__ bounds, uint8_t* __restrict__ tmp) {
  const int b = blockIdx.z;
  const int top = boxes[b * 4], left = boxes[b * 4 + 1], h = boxes[b * 4 + 2];
  const int W = src_hw[b * 2 + 1];
  const int xx = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  if (xx >= S || y >= h) return;
  const int* k = coeffs + (static_cast<long>(b) * S + xx) * ksize;
  const int* bd = bounds + (static_cast<long>(b) * S + xx) * 2;
  const int x0 = bd[0], n = bd[1];
  const uint8_t* row = src + src_off[b] + (static_cast<long>(top + y) * W + left + x0) * 3;
  int a0 = 1 << (kPrecisionBits - 1), a1 = a0, a2 = a0;
  for (int x = 0; x < n; ++x) {
    const int c = __ldg(k + x);
    a0 += row[3 * x] * c, a1 += row[3 * x + 1] * c, a2 += row[3 * x + 2] * c;
  }
  uint8_t* o = tmp + ((static_cast<long>(b) * max_h + y) * S + xx) * 3;
  o[0] = clip8(a0), o[1] = clip8(a1), o[2] = clip8(a2);
}

// vertical pass + flip + HWC -> CHW; thread = (yy, xx)
__global__ void rrc_vertical_kernel(const uint8_t* __restrict__ tmp, const uint8_t* __restrict__ flips, int batch, int S,
                                    int ksize, int max_h, const int* __restrict__ coeffs,
                                    const int* __restrict__ bounds, uint8_t* __restrict__ out) {
  const int b = blockIdx.z;
  const int xx = blockIdx.x * blockDim.x + threadIdx.x;
  const int yy = blockIdx.y * blockDim.y + threadIdx.y;
  if (xx >= S || yy >= S) return;
  const int* k = coeffs + ((static_cast<long>(batch) + b) * S + yy) * ksize;   // axis 1
  const int* bd = bounds + ((static_cast<long>(batch) + b) * S + yy) * 2;
  const int y0 = bd[0], n = bd[1];
  const uint8_t* col = tmp + ((static_cast<long>(b) * max_h + y0) * S + xx) * 3;
  int a0 = 1 << (kPrecisionBits - 1), a1 = a0, a2 = a0;
  for (int y = 0; y < n; ++y) {
    const int c = __ldg(k + y);
    const uint8_t* p = col + static_cast<long>(y) * S * 3;
    a0 += p[0] * c, a1 += p[1] * c, a2 += p[2] * c;
  }
  const int ox = (flips != nullptr && flips[b]) ? S - 1 - xx : xx;
  uint8_t* o = out + static_cast<long>(b) * 3 * S * S + static_cast<long>(yy) * S + ox;
  o[0] = clip8(a0), o[static_cast<long>(S) * S] = clip8(a1), o[2L * S * S] = clip8(a2);
}

int rrc_ksize(int max_crop, int S) {
  const double scale = static_cast<double>(max_crop) / S;
  const double support = 2.0 * (scale < 1.0 ? 1.0 : scale);
  return static_cast<int>(ceil(support)) * 2 + 1;
}

}  // namespace
}  // namespace jm

extern "C" size_t jm_rrc_flip_workspace(int batch, int max_crop, int out_size) {
  if (batch <= 0 || max_crop <= 0 || out_size <= 0) return 0;
  const size_t k = static_cast<size_t>(jm::rrc_ksize(max_crop, out_size));
  const size_t tables = 2ull * batch * out_size * (k + 2) * sizeof(int);
  const size_t tmp = static_cast<size_t>(batch) * max_crop * out_size * 3;
  return (tables + 255) / 256 * 256 + tmp;
}

extern "C" int jm_rrc_flip(const uint8_t* src, const long* src_off, const int* src_hw, const int* boxes,
                           const uint8_t* flips, int batch, int max_crop, int out_size, uint8_t* out, void* workspace,
                           size_t workspace_bytes, void* stream_) {
  using namespace jm;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  JM_CHECK(src && src_off && src_hw && boxes && out, "jm_rrc_flip: null argument");
  JM_CHECK(batch > 0 && max_crop > 0 && out_size > 0, "jm_rrc_flip: empty problem");
  JM_CHECK(workspace && workspace_bytes >= jm_rrc_flip_workspace(batch, max_crop, out_size),
           "jm_rrc_flip: workspace too small (jm_rrc_flip_workspace(batch, max_crop, out_size) bytes)");
  const int S = out_size, ksize = rrc_ksize(max_crop, S);
  int* coeffs = static_cast<int*>(workspace);
  int* bounds = coeffs + 2L * batch * S * ksize;
  const size_t tables = 2ull * batch * S * (ksize + 2) * sizeof(int);
  uint8_t* tmp = static_cast<uint8_t*>(workspace) + (tables + 255) / 256 * 256;
  rrc_coeffs_kernel<<<dim3((S + 127) / 128, batch, 2), 128, 0, stream>>>(boxes, batch, S, ksize, coeffs, bounds);
  JM_AFTER_LAUNCH();
  rrc_horizontal_kernel<<<dim3((S + 31) / 32, (max_crop + 7) / 8, batch), dim3(32, 8), 0, stream>>>(
      src, src_off, src_hw, boxes, batch, S, ksize, max_crop, coeffs, bounds, tmp);
  JM_AFTER_LAUNCH();
  rrc_vertical_kernel<<<dim3((S + 31) / 32, (S + 7) / 8, batch), dim3(32, 8), 0, stream>>>(tmp, flips, batch, S, ksize,
                                                                                          max_crop, coeffs, bounds, out);
  JM_AFTER_LAUNCH();
  return 0;
}
