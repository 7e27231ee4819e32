/* A plain-C caller of the C-ABI (include/ia_b200.h): no Python, no torch — cudaMalloc'ed
 * buffers and raw pointers only, which is all a cgo/JNI/N-API/ctypes binding would pass.
 * Reads inputs written by tests/test_gpu_c_caller.py, runs K1 -> K2+K3 -> K4 and writes the
 * outputs back to files for the test to compare with the oracle. */
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "ia_b200.h"

#define CHECK_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); return 2; } } while (0)
#define CHECK_IA(x) do { int r_ = (x); if (r_ != IA_OK) { \
  fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, ia_error_string(r_)); return 3; } } while (0)

static void* read_file(const char* path, size_t bytes) {
  void* p = malloc(bytes);
  FILE* f = fopen(path, "rb");
  if (!f || fread(p, 1, bytes, f) != bytes) { fprintf(stderr, "cannot read %s\n", path); exit(4); }
  fclose(f);
  return p;
}
static void write_file(const char* path, const void* p, size_t bytes) {
  FILE* f = fopen(path, "wb");
  if (!f || fwrite(p, 1, bytes, f) != bytes) { fprintf(stderr, "cannot write %s\n", path); exit(5); }
  fclose(f);
}

int main(int argc, char** argv) {
  if (argc < 7) { fprintf(stderr, "usage: c_caller dir n h w anchors max_det\n"); return 1; }
  const char* dir = argv[1];
  const int n = atoi(argv[2]), h = atoi(argv[3]), w = atoi(argv[4]), anchors = atoi(argv[5]);
  const int max_det = atoi(argv[6]), classes = 80, target = 640, crop = 224, roi_cap = n * 8;
  char path[512];
  const size_t frame_bytes = (size_t)n * h * w * 3, head_bytes = (size_t)n * anchors * 85 * 4;
  const size_t tensor_bytes = (size_t)n * 3 * target * target * 4;
  const size_t crops_bytes = (size_t)roi_cap * 3 * crop * crop * 4;

  snprintf(path, sizeof path, "%s/frames.bin", dir);
  uint8_t* frames = (uint8_t*)read_file(path, frame_bytes);
  snprintf(path, sizeof path, "%s/heads.bin", dir);
  float* heads = (float*)read_file(path, head_bytes);

  uint8_t* d_frames; float *d_heads, *d_tensor, *d_dets, *d_crops; int32_t *d_keep, *d_count, *d_status, *d_offs, *d_rois;
  void* d_ws;
  const size_t ws_bytes = ia_decode_nms_workspace_bytes(n, anchors, 0);
  CHECK_CUDA(cudaMalloc((void**)&d_frames, frame_bytes));
  CHECK_CUDA(cudaMalloc((void**)&d_heads, head_bytes));
  CHECK_CUDA(cudaMalloc((void**)&d_tensor, tensor_bytes));
  CHECK_CUDA(cudaMalloc((void**)&d_dets, (size_t)n * max_det * 6 * 4));
  CHECK_CUDA(cudaMalloc((void**)&d_keep, (size_t)n * max_det * 4));
  CHECK_CUDA(cudaMalloc((void**)&d_count, (size_t)n * 4));
  CHECK_CUDA(cudaMalloc((void**)&d_status, (size_t)n * 4));
  CHECK_CUDA(cudaMalloc((void**)&d_offs, (size_t)(n + 1) * 4));
  CHECK_CUDA(cudaMalloc((void**)&d_rois, (size_t)roi_cap * 8 * 4));
  CHECK_CUDA(cudaMalloc((void**)&d_crops, crops_bytes));
  CHECK_CUDA(cudaMalloc(&d_ws, ws_bytes));
  cudaStream_t stream;
  CHECK_CUDA(cudaStreamCreate(&stream));
  CHECK_CUDA(cudaMemcpyAsync(d_frames, frames, frame_bytes, cudaMemcpyHostToDevice, stream));
  CHECK_CUDA(cudaMemcpyAsync(d_heads, heads, head_bytes, cudaMemcpyHostToDevice, stream));

  ia_letterbox_geom g;
  CHECK_IA(ia_letterbox_geometry(h, w, target, &g));
  const float mean[3] = {0.485f, 0.456f, 0.406f}, std[3] = {0.229f, 0.224f, 0.225f};
  CHECK_IA(ia_yolo_preprocess(d_frames, n, h, w, target, 255.0f, d_tensor, stream));
  CHECK_IA(ia_decode_nms(d_heads, IA_HEAD_LAYOUT_V5, n, anchors, classes, 0.25f, 0.45f, 0, max_det, d_dets,
                         d_keep, d_count, d_status, d_ws, ws_bytes, stream));
  CHECK_IA(ia_crop_resize_normalize(d_frames, n, h, w, d_dets, d_count, max_det, (float)g.scale, g.pad_w,
                                    g.pad_h, crop, mean, std, d_crops, roi_cap, d_offs, d_rois, stream));
  CHECK_CUDA(cudaStreamSynchronize(stream));

  float* tensor = (float*)malloc(tensor_bytes);
  float* dets = (float*)malloc((size_t)n * max_det * 6 * 4);
  int32_t* keep = (int32_t*)malloc((size_t)n * max_det * 4);
  int32_t* count = (int32_t*)malloc((size_t)n * 4);
  int32_t* offs = (int32_t*)malloc((size_t)(n + 1) * 4);
  float* crops = (float*)malloc(crops_bytes);
  CHECK_CUDA(cudaMemcpy(tensor, d_tensor, tensor_bytes, cudaMemcpyDeviceToHost));
  CHECK_CUDA(cudaMemcpy(dets, d_dets, (size_t)n * max_det * 6 * 4, cudaMemcpyDeviceToHost));
  CHECK_CUDA(cudaMemcpy(keep, d_keep, (size_t)n * max_det * 4, cudaMemcpyDeviceToHost));
  CHECK_CUDA(cudaMemcpy(count, d_count, (size_t)n * 4, cudaMemcpyDeviceToHost));
  CHECK_CUDA(cudaMemcpy(offs, d_offs, (size_t)(n + 1) * 4, cudaMemcpyDeviceToHost));
  CHECK_CUDA(cudaMemcpy(crops, d_crops, crops_bytes, cudaMemcpyDeviceToHost));
  snprintf(path, sizeof path, "%s/tensor.bin", dir); write_file(path, tensor, tensor_bytes);
  snprintf(path, sizeof path, "%s/dets.bin", dir);   write_file(path, dets, (size_t)n * max_det * 6 * 4);
  snprintf(path, sizeof path, "%s/keep.bin", dir);   write_file(path, keep, (size_t)n * max_det * 4);
  snprintf(path, sizeof path, "%s/count.bin", dir);  write_file(path, count, (size_t)n * 4);
  snprintf(path, sizeof path, "%s/offs.bin", dir);   write_file(path, offs, (size_t)(n + 1) * 4);
  snprintf(path, sizeof path, "%s/crops.bin", dir);  write_file(path, crops, crops_bytes);
  printf("%s scale=%.17g rois=%d\n", ia_version(), g.scale, offs[n]);
  return 0;
}
