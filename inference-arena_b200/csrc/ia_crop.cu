// ia_crop.cu — K4: batched ROI crop -> bilinear resize -> ImageNet normalise -> CHW, fed by
// the NMS output on the device (no host round trip), plus the small box/normalise ops.
//
// Replaces the per-detection loop of the reference (architectures/monolithic/app/
// inference.py:188-193): scale_boxes (src/shared/processing/transforms.py:183-230),
// extract_crop (mobilenet_preprocess.py:236-269), MobileNetPreprocessor.preprocess
// (mobilenet_preprocess.py:114-160) and imagenet_normalize (transforms.py:238-272).
//
// Stage A (one CTA): exclusive scan of the per-image detection counts -> roi_offsets, and
// for every detection the float32 box arithmetic of scale_boxes followed by extract_crop's
// truncation/clamp -> an integer rectangle per ROI.
// Stage B (one CTA per ROI row-group): OpenCV's 8-bit INTER_LINEAR (copy / 2x2 box /
// 11-bit fixed-point bilinear, bit-identical to cv2.resize), a 3x256 LUT holding
// ((v/255) - mean) / std computed with the reference's float32 operations, and coalesced
// 128-bit plane stores.  HBM traffic: crop pixels + 3*S*S floats per ROI.
#include "ia_common.cuh"

namespace ia {

constexpr int kRoiThreads = 1024;


constexpr int kRoiInts = 8;

// ------------------------------------------------------------------------------ stage A
struct RoiParams {
  const float* __restrict__ dets;
  const int* __restrict__ count;
  int n, max_det;
  float scale;
  int pad_w, pad_h, src_h, src_w;
  int roi_capacity;
  int* __restrict__ offsets;  // [n+1]
  int* __restrict__ rois;     // [capacity, 8]
  const ia_frame_desc* __restrict__ descs;  // ragged: per-image size (device); geometry derived per image
  int target;                  // letterbox target used with descs
};

// Letterbox geometry of one frame on the device, in the reference's float64 arithmetic
// (transforms.py:153-157,174-175): same IEEE operations as ia_letterbox_geometry on the host.
__device__ __forceinline__ void device_geometry(int h, int w, int target, float& scale, int& pad_w, int& pad_h) {
  const double sh = __ddiv_rn((double)target, (double)h), sw = __ddiv_rn((double)target, (double)w);
  const double sc = sh < sw ? sh : sw;
  const int new_w = __double2int_rz(__dmul_rn((double)w, sc));
  const int new_h = __double2int_rz(__dmul_rn((double)h, sc));
  pad_w = (target - new_w) / 2;
  pad_h = (target - new_h) / 2;
  scale = __double2float_rn(sc);
}

// transforms.py:216-228 for one coordinate: (v - pad) / float32(scale), clip to [0, hi].
__device__ __forceinline__ float unletterbox(float v, int pad, float scale, int hi) {
  const float t = __fdiv_rn(__fsub_rn(v, (float)pad), scale);
  return np_min(np_max(t, 0.f), (float)hi);
}

__device__ __forceinline__ void write_roi(int* roi, int img, float fx1, float fy1, float fx2,
                                          float fy2, int src_h, int src_w) {
  // mobilenet_preprocess.py:256-267: int() truncates toward zero, then clamp.
  int x1 = max(0, __float2int_rz(fx1));
  int y1 = max(0, __float2int_rz(fy1));
  int x2 = min(src_w, __float2int_rz(fx2));
  int y2 = min(src_h, __float2int_rz(fy2));
  const int empty = (x2 <= x1 || y2 <= y1) ? 1 : 0;
  int4* out = reinterpret_cast<int4*>(roi);  // records are 32-byte aligned
  out[0] = make_int4(img, x1, y1, x2);
  out[1] = make_int4(y2, empty, 0, 0);
}

// scale_boxes + extract_crop for detection k of image i -> ROI record r.
__device__ __forceinline__ void emit_roi(const RoiParams& p, int i, int k, int r) {
  if (r >= p.roi_capacity) return;
  float scale = p.scale;
  int pad_w = p.pad_w, pad_h = p.pad_h, src_h = p.src_h, src_w = p.src_w;
  if (p.descs != nullptr) {
    src_h = p.descs[i].h;
    src_w = p.descs[i].w;
    device_geometry(src_h, src_w, p.target, scale, pad_w, pad_h);
  }
  const float* d = p.dets + ((size_t)i * p.max_det + k) * 6;
  write_roi(p.rois + (size_t)r * kRoiInts, i, unletterbox(d[0], pad_w, scale, src_w),
            unletterbox(d[1], pad_h, scale, src_h), unletterbox(d[2], pad_w, scale, src_w),
            unletterbox(d[3], pad_h, scale, src_h), src_h, src_w);
}

__global__ void __launch_bounds__(kRoiThreads) roi_rects_kernel(const RoiParams p) {
  __shared__ int warp_sums[32];
  __shared__ int carry;
  __shared__ int s_before[kRoiThreads];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < p.n; base += kRoiThreads) {
    // ---- exclusive scan of the clamped counts of up to 1024 images ------------------------
    const int i = base + tid;
    const int c = i < p.n ? min(max(p.count[i], 0), p.max_det) : 0;
    int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int w = warp_sums[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += t;
      }
      warp_sums[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    const int before = carry + (warp ? warp_sums[warp - 1] : 0) + incl - c;
    s_before[tid] = before;
    if (i < p.n) p.offsets[i] = before;
    __syncthreads();
    // ---- rectangles: one thread per ROI of the chunk (the image is found by a binary search
    // in the scanned offsets), so only real work is executed and all loads are independent
    const int chunk_n = min(kRoiThreads, p.n - base);
    const int chunk_total = warp_sums[31];
    for (int r = tid; r < chunk_total; r += kRoiThreads) {
      const int target = carry + r;  // global ROI index
      int lo = 0, hi = chunk_n - 1;  // last image whose first ROI index is <= target and that has ROIs
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (s_before[mid] <= target) lo = mid; else hi = mid - 1;
      }
      emit_roi(p, base + lo, target - s_before[lo], target);
    }
    __syncthreads();
    if (tid == 0) carry += warp_sums[31];
    __syncthreads();
  }
  if (tid == 0) p.offsets[p.n] = carry;
}

}  // namespace ia

#include "ia_resize_tiles.cuh"  // stage B: resize_tiles_kernel, launch_crops (+ K1's linear mode)

namespace ia {

// ------------------------------------------------------------------------- small kernels
__global__ void scale_boxes_kernel(const float* in, int m, int cols, float scale, int pad_w,
                                   int pad_h, int orig_h, int orig_w, float* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m * cols) return;
  const int c = i % cols;
  float v = in[i];
  if (c < 4) v = (c & 1) ? unletterbox(v, pad_h, scale, orig_h) : unletterbox(v, pad_w, scale, orig_w);
  out[i] = v;
}

// scale_boxes on float64 boxes: numpy computes `boxes[:, :4] /= scale` and the clips in the
// array's own precision (transforms.py:216-228), here binary64 with IEEE _rn operations.
__global__ void scale_boxes_f64_kernel(const double* in, int m, int cols, double scale, int pad_w,
                                       int pad_h, int orig_h, int orig_w, double* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m * cols) return;
  const int c = i % cols;
  double v = in[i];
  if (c < 4) {
    const double t = __ddiv_rn(__dsub_rn(v, (double)((c & 1) ? pad_h : pad_w)), scale);
    const double hi = (double)((c & 1) ? orig_h : orig_w);
    v = (t != t) ? t : fmin(fmax(t, 0.0), hi);
  }
  out[i] = v;
}

__global__ void crop_rects_kernel(const float* boxes, int m, int cols, int image_index, int src_h,
                                  int src_w, int* rois) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const float* b = boxes + (size_t)i * cols;
  write_roi(rois + (size_t)i * kRoiInts, image_index, b[0], b[1], b[2], b[3], src_h, src_w);
}

__global__ void normalize_u8_kernel(const uint8_t* in, size_t elems, float m0, float m1, float m2,
                                    float s0, float s1, float s2, float* out) {
  __shared__ float lut[3][256];
  const float mean[3] = {m0, m1, m2}, stdv[3] = {s0, s1, s2};
  for (int v = threadIdx.x; v < 256; v += blockDim.x) {
    const float u = __fdiv_rn((float)v, 255.0f);
#pragma unroll
    for (int c = 0; c < 3; ++c) lut[c][v] = __fdiv_rn(__fsub_rn(u, mean[c]), stdv[c]);
  }
  __syncthreads();
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < elems;
       i += (size_t)gridDim.x * blockDim.x)
    out[i] = lut[i % 3][in[i]];
}

}  // namespace ia

extern "C" int ia_scale_boxes(const float* boxes_dev, int m, int cols, float scale, int pad_w,
                              int pad_h, int orig_h, int orig_w, float* out_dev, void* stream) {
  IA_RETURN_IF(m < 0 || cols < 4, IA_ERR_INVALID_ARGUMENT);
  if (m == 0) return IA_OK;
  IA_RETURN_IF(!boxes_dev || !out_dev, IA_ERR_INVALID_ARGUMENT);
  const int total = m * cols;
  ia::scale_boxes_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
      boxes_dev, m, cols, scale, pad_w, pad_h, orig_h, orig_w, out_dev);
  return ia::launch_status();
}

extern "C" int ia_scale_boxes_f64(const double* boxes_dev, int m, int cols, double scale, int pad_w,
                                  int pad_h, int orig_h, int orig_w, double* out_dev, void* stream) {
  IA_RETURN_IF(m < 0 || cols < 4, IA_ERR_INVALID_ARGUMENT);
  if (m == 0) return IA_OK;
  IA_RETURN_IF(!boxes_dev || !out_dev, IA_ERR_INVALID_ARGUMENT);
  const int total = m * cols;
  ia::scale_boxes_f64_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
      boxes_dev, m, cols, scale, pad_w, pad_h, orig_h, orig_w, out_dev);
  return ia::launch_status();
}

extern "C" int ia_crop_resize_normalize(const uint8_t* frames_dev, int n, int src_h, int src_w,
                                        const float* dets_dev, const int32_t* count_dev, int max_det,
                                        float scale, int pad_w, int pad_h, int out_size,
                                        const float mean[3], const float std[3], float* crops_dev,
                                        int roi_capacity, int32_t* roi_offsets_dev, int32_t* rois_dev,
                                        void* stream_) {
  using namespace ia;
  cudaStream_t stream = (cudaStream_t)stream_;
  IA_RETURN_IF(n < 0 || src_h < 1 || src_w < 1 || max_det < 1 || out_size < 1 || roi_capacity < 0,
               IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(!roi_offsets_dev || !mean || !std, IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(n > 0 && (!frames_dev || !dets_dev || !count_dev), IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(roi_capacity > 0 && (!crops_dev || !rois_dev), IA_ERR_INVALID_ARGUMENT);
  RoiParams a;
  a.dets = dets_dev;
  a.count = count_dev;
  a.n = n;
  a.max_det = max_det;
  a.scale = scale;
  a.pad_w = pad_w;
  a.pad_h = pad_h;
  a.src_h = src_h;
  a.src_w = src_w;
  a.roi_capacity = roi_capacity;
  a.offsets = roi_offsets_dev;
  a.rois = rois_dev;
  a.descs = nullptr;
  a.target = 0;
  roi_rects_kernel<<<1, kRoiThreads, 0, stream>>>(a);
  int rc = launch_status();
  if (rc != IA_OK) return rc;
  return launch_crops(frames_dev, (size_t)n * src_h * src_w * 3, src_h, src_w, nullptr, rois_dev, roi_offsets_dev + n, 0, roi_capacity,
                      out_size, mean, std, crops_dev, rois_dev + 6, stream);
}

extern "C" int ia_crop_resize_normalize_ragged(const ia_frame_desc* frames_dev, int n, int max_src_w,
                                               const float* dets_dev, const int32_t* count_dev,
                                               int max_det, int letterbox_target, int out_size,
                                               const float mean[3], const float std[3], float* crops_dev,
                                               int roi_capacity, int32_t* roi_offsets_dev,
                                               int32_t* rois_dev, void* stream_) {
  using namespace ia;
  cudaStream_t stream = (cudaStream_t)stream_;
  IA_RETURN_IF(n < 0 || max_src_w < 1 || max_det < 1 || letterbox_target < 1 || out_size < 1 ||
                   roi_capacity < 0,
               IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(!roi_offsets_dev || !mean || !std, IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(n > 0 && (!frames_dev || !dets_dev || !count_dev), IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(roi_capacity > 0 && (!crops_dev || !rois_dev), IA_ERR_INVALID_ARGUMENT);
  RoiParams a;
  a.dets = dets_dev;
  a.count = count_dev;
  a.n = n;
  a.max_det = max_det;
  a.scale = 1.f;
  a.pad_w = a.pad_h = 0;
  a.src_h = a.src_w = 1;
  a.roi_capacity = roi_capacity;
  a.offsets = roi_offsets_dev;
  a.rois = rois_dev;
  a.descs = frames_dev;
  a.target = letterbox_target;
  roi_rects_kernel<<<1, kRoiThreads, 0, stream>>>(a);
  int rc = launch_status();
  if (rc != IA_OK) return rc;
  return launch_crops(nullptr, 0, 1, max_src_w, frames_dev, rois_dev, roi_offsets_dev + n, 0, roi_capacity,
                      out_size, mean, std, crops_dev, rois_dev + 6, stream);
}

extern "C" int ia_resize_normalize_rois(const uint8_t* frames_dev, int n, int src_h, int src_w,
                                        const int32_t* rois_dev, int num_rois, int out_size,
                                        const float mean[3], const float std[3], float* crops_dev,
                                        void* stream) {
  IA_RETURN_IF(n < 0 || src_h < 1 || src_w < 1 || num_rois < 0 || out_size < 1, IA_ERR_INVALID_ARGUMENT);
  if (num_rois == 0) return IA_OK;
  IA_RETURN_IF(!frames_dev || !rois_dev || !crops_dev || !mean || !std, IA_ERR_INVALID_ARGUMENT);
  return ia::launch_crops(frames_dev, (size_t)n * src_h * src_w * 3, src_h, src_w, nullptr, rois_dev, nullptr, num_rois, num_rois, out_size,
                          mean, std, crops_dev, nullptr, (cudaStream_t)stream);
}

extern "C" int ia_extract_crop_rects(const float* boxes_dev, int m, int cols, int image_index,
                                     int src_h, int src_w, int32_t* rois_dev, void* stream) {
  IA_RETURN_IF(m < 0 || cols < 4 || src_h < 1 || src_w < 1, IA_ERR_INVALID_ARGUMENT);
  if (m == 0) return IA_OK;
  IA_RETURN_IF(!boxes_dev || !rois_dev, IA_ERR_INVALID_ARGUMENT);
  ia::crop_rects_kernel<<<(m + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
      boxes_dev, m, cols, image_index, src_h, src_w, rois_dev);
  return ia::launch_status();
}

extern "C" int ia_imagenet_normalize_u8(const uint8_t* image_dev, size_t pixels, const float mean[3],
                                        const float std[3], float* out_dev, void* stream) {
  if (pixels == 0) return IA_OK;
  IA_RETURN_IF(!image_dev || !out_dev || !mean || !std, IA_ERR_INVALID_ARGUMENT);
  const size_t elems = pixels * 3;
  size_t blocks = (elems + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  ia::normalize_u8_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
      image_dev, elems, mean[0], mean[1], mean[2], std[0], std[1], std[2], out_dev);
  return ia::launch_status();
}

extern "C" const char* ia_version(void) { return "ia_b200 0.1.0 sm_100a"; }

extern "C" const char* ia_error_string(int code) {
  switch (code) {
    case IA_OK: return "success";
    case IA_ERR_INVALID_ARGUMENT: return "invalid argument";
    case IA_ERR_WORKSPACE_TOO_SMALL: return "workspace too small";
    case IA_ERR_UNSUPPORTED: return "unsupported shape or size";
    case IA_ERR_CAPACITY: return "capacity exceeded";
    default: return code > 0 ? cudaGetErrorString((cudaError_t)code) : "unknown error";
  }
}
