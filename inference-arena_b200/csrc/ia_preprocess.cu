// ia_preprocess.cu — K1: fused letterbox + normalise + HWC->CHW (sm_100a).
//
// Replaces YOLOPreprocessor.preprocess (reference src/shared/processing/
// yolo_preprocess.py:128-174) and letterbox (transforms.py:118-180) for a batch of
// equally-sized uint8 HWC frames.  One CTA produces a group of output rows of one image:
//   1. warp 0 stages the source rows those output rows need into shared memory with 1-D
//      bulk async copies (cp.async.bulk -> TMA engine, completion on an mbarrier); the
//      <= 15 unaligned bytes at either end of a segment are copied by the same lanes;
//   2. meanwhile the other warps build the u8 -> float LUT (exact float32 division, same
//      as numpy's `astype(float32) / 255.0`) and the OpenCV x-coefficient table;
//   3. every thread then produces 4 consecutive output pixels: 8-bit fixed-point bilinear
//      (bit-identical to cv2.resize INTER_LINEAR), LUT, three coalesced 128-bit stores —
//      one per colour plane.
// HBM traffic = source rows actually needed + the output, each once.
#include "ia_common.cuh"

namespace ia {

constexpr int kPreThreads = 256;
constexpr int kPreMaxRows = 8;

struct PreParams {
  const uint8_t* frames;
  void* out;
  int n, src_h, src_w, target;
  int new_h, new_w, pad_h, pad_w;
  int rows_per_cta, groups;
  int slot_bytes;  // shared-memory bytes reserved per staged source row (16-byte multiple)
  float divisor;
  double sc_x, sc_y;
  uint32_t pad_rgb;  // r | g << 8 | b << 16
};

struct RowPlan {
  int image_row;  // 0: padding row
  int off0, off1; // shared-memory byte offsets of the two staged source rows
  int b0, b1;     // vertical coefficients
};

template <bool kOutU8, int kMode>
__global__ void __launch_bounds__(kPreThreads) letterbox_kernel(const PreParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ float lut[256];
  __shared__ RowPlan plan[kPreMaxRows];
  __shared__ int seg_row[2 * kPreMaxRows];  // source row per staging segment, -1 = none
  __shared__ __align__(8) uint64_t bar;

  const int tid = threadIdx.x;
  const int img = blockIdx.x / p.groups;
  const int y_begin = (blockIdx.x % p.groups) * p.rows_per_cta;
  const int rows = min(p.rows_per_cta, p.target - y_begin);
  const uint32_t row_bytes = (uint32_t)p.src_w * 3u;
  const uint8_t* frame = p.frames + (size_t)img * p.src_h * row_bytes;
  int2* xtab = reinterpret_cast<int2*>(smem + (size_t)2 * p.rows_per_cta * p.slot_bytes);

  if (tid == 0) {
    mbar_init(&bar, 1);
    mbar_fence_init();
  }
  // ---- per-row plan: which source rows, which vertical weights -------------------
  if (tid < rows) {
    const int dy = y_begin + tid - p.pad_h;
    RowPlan rp;
    rp.image_row = (dy >= 0 && dy < p.new_h) ? 1 : 0;
    rp.off0 = rp.off1 = 0;
    rp.b0 = kCoefOne;
    rp.b1 = 0;
    int r0 = -1, r1 = -1;
    if (rp.image_row) {
      if (kMode == kResizeCopy) {
        r0 = dy;
      } else if (kMode == kResizeBox2) {
        r0 = 2 * dy;
        r1 = 2 * dy + 1;
      } else {
        const AxisCoef c = axis_coef(dy, p.sc_y, p.src_h, false);
        rp.b0 = c.w0;
        rp.b1 = c.w1;
        r0 = clampi(c.s0, 0, p.src_h - 1);
        if (c.w1 != 0) r1 = clampi(c.s0 + 1, 0, p.src_h - 1);
      }
    }
    const uint32_t mis0 = r0 >= 0 ? (uint32_t)((uintptr_t)(frame + (size_t)r0 * row_bytes) & 15) : 0;
    const uint32_t mis1 = r1 >= 0 ? (uint32_t)((uintptr_t)(frame + (size_t)r1 * row_bytes) & 15) : 0;
    rp.off0 = (2 * tid) * p.slot_bytes + mis0;
    rp.off1 = (2 * tid + 1) * p.slot_bytes + mis1;
    plan[tid] = rp;
    seg_row[2 * tid] = r0;
    seg_row[2 * tid + 1] = r1;
  }
  __syncthreads();

  // ---- stage source rows: bulk async copy of the 16-byte aligned body ------------
  if (tid < 32) {
    const uint8_t* g = nullptr;
    uint8_t* s = nullptr;
    uint32_t head = 0, body = 0, tail = 0;
    if (tid < 2 * rows && seg_row[tid] >= 0) {
      g = frame + (size_t)seg_row[tid] * row_bytes;
      const uint32_t mis = (uint32_t)((uintptr_t)g & 15);
      s = smem + (size_t)tid * p.slot_bytes + mis;
      head = min(row_bytes, (16u - mis) & 15u);
      body = (row_bytes - head) & ~15u;
      tail = row_bytes - head - body;
    }
    uint32_t total = body;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
    if (tid == 0) mbar_expect_tx(&bar, total);
    __syncwarp();
    if (body) bulk_g2s(s + head, g + head, body, &bar);
    for (uint32_t i = 0; i < head; ++i) s[i] = g[i];
    for (uint32_t i = 0; i < tail; ++i) s[head + body + i] = g[head + body + i];
  }

  // ---- tables ---------------------------------------------------------------------
  if (!kOutU8) {
    for (int v = tid; v < 256; v += kPreThreads) lut[v] = __fdiv_rn((float)v, p.divisor);
  }
  if (kMode == kResizeLinear) {
    for (int dx = tid; dx < p.new_w; dx += kPreThreads) {
      const AxisCoef c = axis_coef(dx, p.sc_x, p.src_w, true);
      xtab[dx] = make_int2(c.s0 * 3, c.w0 | (c.w1 << 16));
    }
  }
  __syncthreads();
  mbar_wait(&bar, 0);

  // ---- compute: 4 pixels per item --------------------------------------------------
  const int T = p.target;
  const int xg = (T + 3) >> 2;
  const uint32_t pr = p.pad_rgb & 0xff, pg = (p.pad_rgb >> 8) & 0xff, pb = (p.pad_rgb >> 16) & 0xff;
  const bool vec_ok = (T & 3) == 0;

  for (int it = tid; it < rows * xg; it += kPreThreads) {
    const int j = it / xg;
    const int x = (it - j * xg) << 2;
    const int y = y_begin + j;
    const RowPlan rp = plan[j];
    const uint8_t* p0 = smem + rp.off0;
    const uint8_t* p1 = smem + rp.off1;
    uint32_t px[4][3];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int dx = x + k - p.pad_w;
      if (!rp.image_row || dx < 0 || dx >= p.new_w) {
        px[k][0] = pr;
        px[k][1] = pg;
        px[k][2] = pb;
      } else if (kMode == kResizeCopy) {
        const uint8_t* s = p0 + dx * 3;
        px[k][0] = s[0];
        px[k][1] = s[1];
        px[k][2] = s[2];
      } else if (kMode == kResizeBox2) {
        const uint8_t* a = p0 + dx * 6;
        const uint8_t* b = p1 + dx * 6;
#pragma unroll
        for (int c = 0; c < 3; ++c) px[k][c] = ((uint32_t)a[c] + a[c + 3] + b[c] + b[c + 3] + 2u) >> 2;
      } else {
        const int2 e = xtab[dx];
        const int w0 = e.y & 0xffff, w1 = e.y >> 16;
        const uint8_t* a = p0 + e.x;
        const uint8_t* b = p1 + e.x;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          int h0 = a[c] * w0;
          if (w1) h0 += a[c + 3] * w1;
          int h1 = 0;
          if (rp.b1) {
            h1 = b[c] * w0;
            if (w1) h1 += b[c + 3] * w1;
          }
          px[k][c] = (uint32_t)vresize(h0, h1, rp.b0, rp.b1);
        }
      }
    }
    if (kOutU8) {
      uint8_t* o = reinterpret_cast<uint8_t*>(p.out) + (((size_t)img * T + y) * T + x) * 3;
      if (vec_ok) {
        uint32_t* o32 = reinterpret_cast<uint32_t*>(o);
        o32[0] = px[0][0] | px[0][1] << 8 | px[0][2] << 16 | px[1][0] << 24;
        o32[1] = px[1][1] | px[1][2] << 8 | px[2][0] << 16 | px[2][1] << 24;
        o32[2] = px[2][2] | px[3][0] << 8 | px[3][1] << 16 | px[3][2] << 24;
      } else {
        for (int k = 0; k < 4 && x + k < T; ++k)
          for (int c = 0; c < 3; ++c) o[k * 3 + c] = (uint8_t)px[k][c];
      }
    } else {
      float* o = reinterpret_cast<float*>(p.out) + (((size_t)img * 3) * T + y) * T + x;
      const size_t plane = (size_t)T * T;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        if (vec_ok) {
          st_stream_f4(o + c * plane, lut[px[0][c]], lut[px[1][c]], lut[px[2][c]], lut[px[3][c]]);
        } else {
          for (int k = 0; k < 4 && x + k < T; ++k) st_stream_f1(o + c * plane + k, lut[px[k][c]]);
        }
      }
    }
  }
}

template <bool kOutU8>
static int launch_letterbox(const uint8_t* frames, int n, int src_h, int src_w, int target,
                            float divisor, uint32_t pad_rgb, void* out, cudaStream_t stream) {
  IA_RETURN_IF(n < 0 || src_h < 1 || src_w < 1 || target < 1, IA_ERR_INVALID_ARGUMENT);
  if (n == 0) return IA_OK;
  IA_RETURN_IF(!frames || !out, IA_ERR_INVALID_ARGUMENT);
  ia_letterbox_geom g;
  int rc = ia_letterbox_geometry(src_h, src_w, target, &g);
  if (rc != IA_OK) return rc;

  PreParams p;
  p.frames = frames;
  p.out = out;
  p.n = n;
  p.src_h = src_h;
  p.src_w = src_w;
  p.target = target;
  p.new_h = g.new_h;
  p.new_w = g.new_w;
  p.pad_h = g.pad_h;
  p.pad_w = g.pad_w;
  p.divisor = divisor;
  p.sc_x = axis_scale(g.new_w, src_w);
  p.sc_y = axis_scale(g.new_h, src_h);
  p.pad_rgb = pad_rgb;
  p.slot_bytes = (((src_w * 3 + 15) & ~15) + 16);
  const int mode = resize_mode(src_h, src_w, g.new_h, g.new_w);
  const size_t xtab_bytes = (size_t)g.new_w * sizeof(int2);
  const size_t budget = 96 * 1024;
  int rows = kPreMaxRows;
  while (rows > 1 && (size_t)2 * rows * p.slot_bytes + xtab_bytes > budget) rows >>= 1;
  const size_t smem = (size_t)2 * rows * p.slot_bytes + xtab_bytes;
  IA_RETURN_IF(smem > 220 * 1024, IA_ERR_UNSUPPORTED);
  IA_RETURN_IF((size_t)rows * 2 * src_w * 3 >= (1u << 20), IA_ERR_UNSUPPORTED);  // mbarrier tx range
  p.rows_per_cta = rows;
  p.groups = (target + rows - 1) / rows;
  const long long blocks = (long long)p.groups * n;
  IA_RETURN_IF(blocks > 0x7fffffffLL, IA_ERR_UNSUPPORTED);

  void (*kernel)(PreParams) = nullptr;
  switch (mode) {
    case kResizeCopy: kernel = letterbox_kernel<kOutU8, kResizeCopy>; break;
    case kResizeBox2: kernel = letterbox_kernel<kOutU8, kResizeBox2>; break;
    default: kernel = letterbox_kernel<kOutU8, kResizeLinear>; break;
  }
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  kernel<<<(unsigned)blocks, kPreThreads, smem, stream>>>(p);
  return launch_status();
}

}  // namespace ia

extern "C" int ia_letterbox_geometry(int src_h, int src_w, int target, ia_letterbox_geom* out) {
  if (!out || src_h < 1 || src_w < 1 || target < 1) return IA_ERR_INVALID_ARGUMENT;
  // transforms.py:153-157,174-175 — Python float (binary64) arithmetic, int() truncation.
  const double sh = (double)target / (double)src_h;
  const double sw = (double)target / (double)src_w;
  const double scale = sh < sw ? sh : sw;
  out->src_h = src_h;
  out->src_w = src_w;
  out->target = target;
  out->scale = scale;
  out->new_w = (int)((double)src_w * scale);
  out->new_h = (int)((double)src_h * scale);
  // cv2.resize rejects an empty destination; mirror it as an error.
  if (out->new_w < 1 || out->new_h < 1) return IA_ERR_UNSUPPORTED;
  out->pad_w = (target - out->new_w) / 2;
  out->pad_h = (target - out->new_h) / 2;
  return IA_OK;
}

extern "C" int ia_yolo_preprocess(const uint8_t* frames_dev, int n, int src_h, int src_w,
                                  int target, float divisor, float* out_dev, void* stream) {
  return ia::launch_letterbox<false>(frames_dev, n, src_h, src_w, target, divisor,
                                     114u | 114u << 8 | 114u << 16, out_dev, (cudaStream_t)stream);
}

extern "C" int ia_letterbox_u8(const uint8_t* frames_dev, int n, int src_h, int src_w, int target,
                               const uint8_t pad_rgb[3], uint8_t* out_dev, void* stream) {
  if (!pad_rgb) return IA_ERR_INVALID_ARGUMENT;
  const uint32_t rgb = (uint32_t)pad_rgb[0] | (uint32_t)pad_rgb[1] << 8 | (uint32_t)pad_rgb[2] << 16;
  return ia::launch_letterbox<true>(frames_dev, n, src_h, src_w, target, 255.0f, rgb, out_dev,
                                    (cudaStream_t)stream);
}
