/* Plain-C caller of the N3 / N4 entries of the C-ABI (include/ia_b200.h): reads a JPEG file and
 * a block of logits written by tests/test_gpu_c_caller.py, decodes the file on the device
 * (ia_jpeg_parse -> ia_jpeg_plan -> ia_jpeg_decode), runs ia_classify_topk, and writes the
 * results to files.  No Python, no torch: malloc, cudaMalloc, raw pointers. */
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "ia_b200.h"

#define CHECK_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); return 2; } } while (0)
#define CHECK_IA(x) do { int r_ = (x); if (r_ != IA_OK) { \
  fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, ia_error_string(r_)); return 3; } } while (0)

static void* read_all(const char* path, size_t* bytes) {
  FILE* f = fopen(path, "rb");
  if (!f) { fprintf(stderr, "cannot open %s\n", path); exit(4); }
  fseek(f, 0, SEEK_END);
  *bytes = (size_t)ftell(f);
  fseek(f, 0, SEEK_SET);
  void* p = malloc(*bytes + 64);
  if (fread(p, 1, *bytes, f) != *bytes) { fprintf(stderr, "cannot read %s\n", path); exit(4); }
  fclose(f);
  return p;
}
static void write_file(const char* path, const void* p, size_t bytes) {
  FILE* f = fopen(path, "wb");
  if (!f || fwrite(p, 1, bytes, f) != bytes) { fprintf(stderr, "cannot write %s\n", path); exit(5); }
  fclose(f);
}

int main(int argc, char** argv) {
  if (argc < 5) { fprintf(stderr, "usage: c_jpeg_caller dir rows classes k\n"); return 1; }
  const char* dir = argv[1];
  const int rows = atoi(argv[2]), classes = atoi(argv[3]), k = atoi(argv[4]);
  char path[512];
  cudaStream_t stream;
  CHECK_CUDA(cudaStreamCreate(&stream));

  /* ---- N3: one JPEG file ---- */
  size_t file_bytes;
  snprintf(path, sizeof path, "%s/image.jpg", dir);
  uint8_t* file = (uint8_t*)read_all(path, &file_bytes);
  ia_jpeg_job job;
  if (sizeof(job) != ia_jpeg_job_size()) { fprintf(stderr, "ia_jpeg_job layout mismatch\n"); return 6; }
  CHECK_IA(ia_jpeg_parse(file, file_bytes, &job));
  const size_t ws_bytes = ia_jpeg_workspace_bytes(&job, 1);
  const size_t out_bytes = (size_t)job.out_width * job.out_height * 3;
  uint8_t *d_file, *d_out;
  void* d_ws;
  ia_jpeg_job* d_job;
  int32_t* d_status;
  CHECK_CUDA(cudaMalloc((void**)&d_file, file_bytes + 64));
  CHECK_CUDA(cudaMalloc((void**)&d_out, out_bytes));
  CHECK_CUDA(cudaMalloc(&d_ws, ws_bytes));
  CHECK_CUDA(cudaMalloc((void**)&d_job, sizeof(job)));
  CHECK_CUDA(cudaMalloc((void**)&d_status, sizeof(int32_t)));
  const uint8_t* files[1] = {d_file};
  uint8_t* outs[1] = {d_out};
  CHECK_IA(ia_jpeg_plan(&job, 1, files, outs, d_ws, ws_bytes));
  CHECK_CUDA(cudaMemcpyAsync(d_file, file, file_bytes, cudaMemcpyHostToDevice, stream));
  CHECK_CUDA(cudaMemcpyAsync(d_job, &job, sizeof(job), cudaMemcpyHostToDevice, stream));
  CHECK_IA(ia_jpeg_decode(&job, d_job, 1, d_status, stream));
  uint8_t* out = (uint8_t*)malloc(out_bytes);
  int32_t status = -1;
  CHECK_CUDA(cudaMemcpyAsync(out, d_out, out_bytes, cudaMemcpyDeviceToHost, stream));
  CHECK_CUDA(cudaMemcpyAsync(&status, d_status, sizeof status, cudaMemcpyDeviceToHost, stream));
  CHECK_CUDA(cudaStreamSynchronize(stream));
  snprintf(path, sizeof path, "%s/rgb.bin", dir);
  write_file(path, out, out_bytes);

  /* ---- N4: softmax + top-k ---- */
  size_t logit_bytes;
  snprintf(path, sizeof path, "%s/logits.bin", dir);
  float* logits = (float*)read_all(path, &logit_bytes);
  if (logit_bytes != (size_t)rows * classes * 4) { fprintf(stderr, "logits.bin has the wrong size\n"); return 7; }
  float *d_logits, *d_values;
  int32_t *d_ids, *d_row_status;
  CHECK_CUDA(cudaMalloc((void**)&d_logits, logit_bytes));
  CHECK_CUDA(cudaMalloc((void**)&d_values, (size_t)rows * k * 4));
  CHECK_CUDA(cudaMalloc((void**)&d_ids, (size_t)rows * k * 4));
  CHECK_CUDA(cudaMalloc((void**)&d_row_status, (size_t)rows * 4));
  CHECK_CUDA(cudaMemcpyAsync(d_logits, logits, logit_bytes, cudaMemcpyHostToDevice, stream));
  CHECK_IA(ia_classify_topk(d_logits, rows, classes, k, 1, 0, d_ids, d_values, d_row_status, stream));
  int32_t* ids = (int32_t*)malloc((size_t)rows * k * 4);
  float* values = (float*)malloc((size_t)rows * k * 4);
  CHECK_CUDA(cudaMemcpyAsync(ids, d_ids, (size_t)rows * k * 4, cudaMemcpyDeviceToHost, stream));
  CHECK_CUDA(cudaMemcpyAsync(values, d_values, (size_t)rows * k * 4, cudaMemcpyDeviceToHost, stream));
  CHECK_CUDA(cudaStreamSynchronize(stream));
  snprintf(path, sizeof path, "%s/ids.bin", dir);
  write_file(path, ids, (size_t)rows * k * 4);
  snprintf(path, sizeof path, "%s/probs.bin", dir);
  write_file(path, values, (size_t)rows * k * 4);

  printf("%s jpeg %dx%d comps=%d status=%d\n", ia_version(), job.out_width, job.out_height, job.ncomp, status);
  return 0;
}
