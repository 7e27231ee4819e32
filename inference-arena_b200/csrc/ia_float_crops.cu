// ia_float_crops.cu — the float32-input branch of MobileNetPreprocessor / imagenet_normalize
// (reference src/shared/processing/mobilenet_preprocess.py:204-206 accepts float32 crops,
// transforms.py:264-267 rescales them by 1/255 only when their maximum exceeds 1).
//
// Not a throughput path (the reference's pipelines only ever pass uint8 crops); it exists so
// the drop-in covers the whole signature on the GPU.  Two flavours of the float resize, both
// bit-identical to what cv2.resize returns in the corresponding build (oracle/cv_resize.py,
// tests/test_oracle_cv_resize.py):
//   IA_F32_RESIZE_IPP     the opencv-python wheels the reference pins: float resizes go to Intel
//                         IPP — double coordinates, weight = float(pos - floor(pos)), every
//                         interpolation fma(p1 - p0, w, p0), horizontal then vertical, no box
//                         average for exact 2x; one-pixel-wide/high images take the generic path;
//   IA_F32_RESIZE_OPENCV  OpenCV's own generic float INTER_LINEAR (resize.cpp HResizeLinear<float>
//                         / VResizeLinear<float>: float coefficients 1-f and f, every product
//                         and sum rounded separately; exact 2x = box average): cv2 with IPP off.
#include "ia_common.cuh"

namespace ia {

struct FloatAxis {
  int s0;
  float w0, w1;
};

// xofs/alpha (zero_frac_at_border) or yofs/beta entry in float form.
__device__ __forceinline__ FloatAxis axis_coef_f32(int i, double sc, int src, bool zero_frac_at_border) {
  const double pos = __dsub_rn(__dmul_rn(__dadd_rn((double)i, 0.5), sc), 0.5);
  float f = __double2float_rn(pos);
  int s0 = __float2int_rd(f);
  f = __fsub_rn(f, (float)s0);
  if (zero_frac_at_border) {
    if (s0 < 0) s0 = 0, f = 0.f;
    if (s0 >= src - 1) s0 = src - 1, f = 0.f;
  }
  FloatAxis a;
  a.s0 = s0;
  a.w0 = __fsub_rn(1.f, f);
  a.w1 = f;
  return a;
}

// IPP: tap and weight from the double coordinate (the fraction is not rounded to float first).
struct IppAxis {
  int s0, s1;
  float w;
};
__device__ __forceinline__ IppAxis axis_coef_ipp(int i, double sc, int src) {
  const double pos = __dsub_rn(__dmul_rn(__dadd_rn((double)i, 0.5), sc), 0.5);
  const double fl = floor(pos);
  IppAxis a;
  a.w = __double2float_rn(__dsub_rn(pos, fl));
  const int s0 = __double2int_rd(fl);
  a.s0 = clampi(s0, 0, src - 1);
  a.s1 = clampi(s0 + 1, 0, src - 1);
  return a;
}

__device__ __forceinline__ uint32_t ordered_bits(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// stats[0]: max of the values seen (ordered bits), stats[1]: 1 if a NaN was seen.
__device__ __forceinline__ void publish_max(float v, bool nan, uint32_t* stats) {
  uint32_t m = nan ? 0u : ordered_bits(v);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
  const unsigned any_nan = __ballot_sync(0xffffffffu, nan);
  if ((threadIdx.x & 31) == 0) {
    if (m) atomicMax(stats, m);
    if (any_nan) atomicExch(stats + 1, 1u);
  }
}

// One thread per output pixel (3 channels): src float [h,w,3] -> dst float [S,S,3] (HWC).
constexpr int kResizeIpp = 100;  // mode value: IPP's bilinear (see the file header)

__global__ void resize_f32_kernel(const float* __restrict__ src, int h, int w, int S, double sc_x,
                                  double sc_y, int mode, float* __restrict__ dst, uint32_t* stats) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  float out[3] = {0.f, 0.f, 0.f};
  const bool live = i < S * S;
  if (live) {
    const int y = i / S, x = i - y * S;
    if (mode == kResizeCopy) {
      const float* p = src + ((size_t)y * w + x) * 3;
      out[0] = p[0], out[1] = p[1], out[2] = p[2];
    } else if (mode == kResizeIpp) {
      const IppAxis ax = axis_coef_ipp(x, sc_x, w), ay = axis_coef_ipp(y, sc_y, h);
      const float* p0 = src + (size_t)ay.s0 * w * 3;
      const float* p1 = src + (size_t)ay.s1 * w * 3;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float a0 = p0[ax.s0 * 3 + c], a1 = p1[ax.s0 * 3 + c];
        const float h0 = __fmaf_rn(__fsub_rn(p0[ax.s1 * 3 + c], a0), ax.w, a0);
        const float h1 = __fmaf_rn(__fsub_rn(p1[ax.s1 * 3 + c], a1), ax.w, a1);
        out[c] = __fmaf_rn(__fsub_rn(h1, h0), ay.w, h0);
      }
    } else if (mode == kResizeBox2) {  // ResizeAreaFast: ((p00 + p01) + p10) + p11, times 0.25
      const float* a = src + ((size_t)(2 * y) * w + 2 * x) * 3;
      const float* b = a + (size_t)w * 3;
#pragma unroll
      for (int c = 0; c < 3; ++c)
        out[c] = __fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(a[c], a[c + 3]), b[c]), b[c + 3]), 0.25f);
    } else {
      const FloatAxis ax = axis_coef_f32(x, sc_x, w, true);
      const FloatAxis ay = axis_coef_f32(y, sc_y, h, false);
      const int x1 = min(ax.s0 + 1, w - 1);
      const int r0 = clampi(ay.s0, 0, h - 1), r1 = clampi(ay.s0 + 1, 0, h - 1);
      const float* p0 = src + (size_t)r0 * w * 3;
      const float* p1 = src + (size_t)r1 * w * 3;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float h0 = __fadd_rn(__fmul_rn(p0[ax.s0 * 3 + c], ax.w0), __fmul_rn(p0[x1 * 3 + c], ax.w1));
        const float h1 = __fadd_rn(__fmul_rn(p1[ax.s0 * 3 + c], ax.w0), __fmul_rn(p1[x1 * 3 + c], ax.w1));
        out[c] = __fadd_rn(__fmul_rn(h0, ay.w0), __fmul_rn(h1, ay.w1));
      }
    }
    float* d = dst + (size_t)i * 3;
    d[0] = out[0], d[1] = out[1], d[2] = out[2];
  }
  const bool nan = live && (out[0] != out[0] || out[1] != out[1] || out[2] != out[2]);
  const float m = live ? fmaxf(fmaxf(out[0], out[1]), out[2]) : -INFINITY;
  publish_max(m, nan, stats);
}

__global__ void max_f32_kernel(const float* __restrict__ in, size_t elems, uint32_t* stats) {
  float m = -INFINITY;
  bool nan = false;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < elems; i += (size_t)gridDim.x * blockDim.x) {
    const float v = in[i];
    nan |= v != v;
    m = fmaxf(m, v);
  }
  publish_max(m, nan, stats);
}

// (x [/255 when the image's maximum exceeds 1]) - mean) / std; HWC in, HWC or CHW out.
__global__ void normalize_f32_kernel(const float* __restrict__ in, size_t pixels, const uint32_t* stats,
                                     float m0, float m1, float m2, float s0, float s1, float s2,
                                     int chw, float* __restrict__ out) {
  const float mean[3] = {m0, m1, m2}, stdv[3] = {s0, s1, s2};
  // transforms.py:266: `if normalized.max() > 1.0` — false when the maximum is NaN
  const bool rescale = stats[1] == 0u && stats[0] > ordered_bits(1.0f);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < pixels; i += (size_t)gridDim.x * blockDim.x) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float v = in[i * 3 + c];
      if (rescale) v = __fdiv_rn(v, 255.0f);
      v = __fdiv_rn(__fsub_rn(v, mean[c]), stdv[c]);
      out[chw ? (size_t)c * pixels + i : i * 3 + c] = v;
    }
  }
}

static unsigned grid_for(size_t work, int block) {
  size_t g = (work + block - 1) / block;
  return (unsigned)(g > 148 * 16 ? 148 * 16 : (g ? g : 1));
}

}  // namespace ia

extern "C" size_t ia_f32_scratch_bytes(int out_size) {
  return out_size < 1 ? 256 : 256 + (size_t)out_size * out_size * 3 * sizeof(float);
}

extern "C" int ia_resize_normalize_f32(const float* crop_dev, int h, int w, int out_size,
                                       const float mean[3], const float std[3], float* out_chw_dev,
                                       void* scratch_dev, size_t scratch_bytes, void* stream) {
  return ia_resize_normalize_f32_as(crop_dev, h, w, out_size, IA_F32_RESIZE_OPENCV, mean, std, out_chw_dev,
                                    scratch_dev, scratch_bytes, stream);
}

extern "C" int ia_resize_normalize_f32_as(const float* crop_dev, int h, int w, int out_size, int flavour,
                                          const float mean[3], const float std[3], float* out_chw_dev,
                                          void* scratch_dev, size_t scratch_bytes, void* stream_) {
  using namespace ia;
  IA_RETURN_IF(flavour != IA_F32_RESIZE_OPENCV && flavour != IA_F32_RESIZE_IPP, IA_ERR_INVALID_ARGUMENT);
  cudaStream_t stream = (cudaStream_t)stream_;
  IA_RETURN_IF(!crop_dev || !mean || !std || !out_chw_dev || !scratch_dev, IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(h < 1 || w < 1 || out_size < 1 || ((uintptr_t)scratch_dev & 15) != 0, IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(scratch_bytes < ia_f32_scratch_bytes(out_size), IA_ERR_WORKSPACE_TOO_SMALL);
  uint32_t* stats = reinterpret_cast<uint32_t*>(scratch_dev);
  float* resized = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(scratch_dev) + 256);
  cudaError_t e = cudaMemsetAsync(stats, 0, 8, stream);
  if (e != cudaSuccess) return (int)e;
  const int pixels = out_size * out_size;
  int mode = resize_mode(h, w, out_size, out_size);
  double sc_x = axis_scale(out_size, w), sc_y = axis_scale(out_size, h);
  if (flavour == IA_F32_RESIZE_IPP && mode != kResizeCopy && h >= 2 && w >= 2) {
    mode = kResizeIpp;
    sc_x = (double)w / (double)out_size;
    sc_y = (double)h / (double)out_size;
  }
  resize_f32_kernel<<<(pixels + 255) / 256, 256, 0, stream>>>(crop_dev, h, w, out_size, sc_x, sc_y, mode, resized,
                                                              stats);
  int rc = launch_status();
  if (rc != IA_OK) return rc;
  normalize_f32_kernel<<<grid_for(pixels, 256), 256, 0, stream>>>(resized, (size_t)pixels, stats, mean[0], mean[1],
                                                                 mean[2], std[0], std[1], std[2], 1, out_chw_dev);
  return launch_status();
}

extern "C" int ia_imagenet_normalize_f32(const float* image_dev, size_t pixels, const float mean[3],
                                         const float std[3], float* out_dev, void* scratch_dev,
                                         size_t scratch_bytes, void* stream_) {
  using namespace ia;
  cudaStream_t stream = (cudaStream_t)stream_;
  if (pixels == 0) return IA_OK;
  IA_RETURN_IF(!image_dev || !mean || !std || !out_dev || !scratch_dev, IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(scratch_bytes < 8 || ((uintptr_t)scratch_dev & 3) != 0, IA_ERR_INVALID_ARGUMENT);
  uint32_t* stats = reinterpret_cast<uint32_t*>(scratch_dev);
  cudaError_t e = cudaMemsetAsync(stats, 0, 8, stream);
  if (e != cudaSuccess) return (int)e;
  max_f32_kernel<<<grid_for(pixels * 3, 256), 256, 0, stream>>>(image_dev, pixels * 3, stats);
  int rc = launch_status();
  if (rc != IA_OK) return rc;
  normalize_f32_kernel<<<grid_for(pixels, 256), 256, 0, stream>>>(image_dev, pixels, stats, mean[0], mean[1], mean[2],
                                                                 std[0], std[1], std[2], 0, out_dev);
  return launch_status();
}
