// ia_common.cuh — shared device helpers for the sm_100a kernels of libia_b200.
//
//  * mbarrier + 1-D bulk async copy (TMA engine, `cp.async.bulk`, SASS UBLKCP) used to
//    stage frame rows / head tiles into shared memory;
//  * the OpenCV INTER_LINEAR coefficient tables reproduced in device code
//    (resize.cpp, 8-bit path — see oracle/cv_resize.py for the restated algorithm);
//  * IEEE float helpers that can never be contracted into FMAs (numpy parity).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <map>
#include <mutex>
#include <utility>

#include "../../include/ia_b200.h"

namespace ia {

// ----------------------------------------------------------------------------- errors
#define IA_RETURN_IF(cond, code) \
  do {                           \
    if (cond) return (code);     \
  } while (0)

inline int launch_status() {
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? IA_OK : (int)e;
}

// Lift a kernel's dynamic shared-memory cap to the device's opt-in maximum, once per (kernel,
// device).  The attribute belongs to the function and the device, not to a launch, and setting
// it again with a smaller value LOWERS it: a launcher that set it to its own launch's size
// could have it lowered by another host thread (another frame shape) between the call and the
// launch.  So it is only ever set to the one largest value, under a lock.
inline cudaError_t allow_max_dynamic_smem(const void* kernel) {
  static std::mutex mu;
  static std::map<std::pair<const void*, int>, bool> done;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  std::lock_guard<std::mutex> lock(mu);
  bool& flag = done[std::make_pair(kernel, dev)];
  if (flag) return cudaSuccess;
  int optin = 0;
  e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  if (e != cudaSuccess) return e;
  cudaFuncAttributes attr;  // static shared memory counts against the same limit
  e = cudaFuncGetAttributes(&attr, kernel);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, optin - (int)attr.sharedSizeBytes);
  if (e == cudaSuccess) flag = true;
  return e;
}

// ------------------------------------------------------------------- mbarrier / bulk copy
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t arrivals) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(arrivals));
}

// make the barrier initialisation visible to the async (TMA) proxy
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}

// raise the barrier's pending transaction count without arriving (the arrival comes separately)
__device__ __forceinline__ void mbar_add_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}

// plain arrival (release semantics at CTA scope): "this warp is done with the buffer"
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// named barrier among `threads` threads of the CTA (id 1..15; 0 is __syncthreads)
__device__ __forceinline__ void named_barrier_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(done)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return done != 0;
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

// global -> shared 1-D bulk copy; dst/src 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes,
                                         uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(smem_dst)),
      "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// ------------------------------------------------------------------ streaming load/store
__device__ __forceinline__ void st_stream_f4(float* p, float a, float b, float c, float d) {
  asm volatile("st.global.cs.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c),
               "f"(d)
               : "memory");
}
__device__ __forceinline__ void st_stream_f1(float* p, float a) {
  asm volatile("st.global.cs.f32 [%0], %1;" ::"l"(p), "f"(a) : "memory");
}

// ------------------------------------------------------------ OpenCV linear coefficients
constexpr int kCoefOne = 2048;  // INTER_RESIZE_COEF_SCALE (11 bits)

// kResizeGather: a linear resize whose every coefficient pair is (2048, 0) on both axes (an
// odd integer downscale lands on pixel centres): OpenCV's arithmetic reduces to dst = src[y0][x0].
enum ResizeMode : int { kResizeCopy = 0, kResizeBox2 = 1, kResizeLinear = 2, kResizeGather = 3 };

__host__ __device__ __forceinline__ int resize_mode(int src_h, int src_w, int dst_h, int dst_w) {
  if (src_h == dst_h && src_w == dst_w) return kResizeCopy;        // resize.cpp: dsize == ssize
  if (src_h == 2 * dst_h && src_w == 2 * dst_w) return kResizeBox2;  // is_area_fast, iscale == 2
  return kResizeLinear;
}

struct AxisCoef {
  int s0;  // first source index (clamped on X; raw floor on Y)
  int w0;  // coefficient of s0     (0..2048)
  int w1;  // coefficient of s0 + 1 (0..2048)
};

// OpenCV's scale_x / scale_y: 1. / ((double)dst / src)  (cv::resize, resize.cpp).
__host__ __device__ __forceinline__ double axis_scale(int dst, int src) {
#ifdef __CUDA_ARCH__
  return __ddiv_rn(1.0, __ddiv_rn((double)dst, (double)src));
#else
  volatile double inv = (double)dst / (double)src;
  return 1.0 / inv;
#endif
}

// One entry of the xofs/alpha (zero_frac_at_border = true) or yofs/beta table.
// Every operation is an explicit round-to-nearest intrinsic: nvcc must not fuse
// (i + 0.5) * sc - 0.5 into an FMA, OpenCV's x86 baseline build does not.
__device__ __forceinline__ AxisCoef axis_coef(int i, double sc, int src, bool zero_frac_at_border) {
  const double pos = __dsub_rn(__dmul_rn(__dadd_rn((double)i, 0.5), sc), 0.5);
  float f = __double2float_rn(pos);
  int s0 = __float2int_rd(f);
  f = __fsub_rn(f, (float)s0);
  if (zero_frac_at_border) {
    if (s0 < 0) {
      s0 = 0;
      f = 0.f;
    }
    if (s0 >= src - 1) {
      s0 = src - 1;
      f = 0.f;
    }
  }
  AxisCoef c;
  c.s0 = s0;
  c.w1 = __float2int_rn(__fmul_rn(f, (float)kCoefOne));
  c.w0 = __float2int_rn(__fmul_rn(__fsub_rn(1.f, f), (float)kCoefOne));
  return c;
}

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// VResizeLinear<uchar,int,short>: ((b0*(h0>>4))>>16) + ((b1*(h1>>4))>>16) + 2) >> 2
__device__ __forceinline__ int vresize(int h0, int h1, int b0, int b1) {
  return (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
}

// Six bytes starting at byte address `a` of the staging buffer -> (lo = bytes 0..3, hi = 4..7).
__device__ __forceinline__ void load6(const uint32_t* stage32, int a, uint32_t& lo, uint32_t& hi) {
  const uint32_t* w = stage32 + (a >> 2);
  const uint32_t w0 = w[0], w1 = w[1], w2 = w[2];
  const uint32_t sh = (uint32_t)(a & 3) << 3;
  lo = __funnelshift_r(w0, w1, sh);
  hi = __funnelshift_r(w1, w2, sh);
}

// PRMT selectors pairing (byte c, byte c+3) of the 8 bytes lo:hi — the two taps of channel c.
__device__ __forceinline__ uint32_t tap_pair(uint32_t lo, uint32_t hi, int c) {
  return __byte_perm(lo, hi, c == 0 ? 0x30u : (c == 1 ? 0x41u : 0x52u));
}

// K1's general linear resize runs in the tile kernel K4 uses (ia_resize_tiles.cuh, compiled in
// ia_crop.cu): n equally sized frames -> letterboxed float32 [n, 3, target, target] = v / divisor.
int launch_letterbox_tiles(const uint8_t* frames, int n, int src_h, int src_w, int target, int new_h, int new_w,
                           int pad_h, int pad_w, float divisor, uint32_t pad_rgb, float* out, cudaStream_t stream);

// numpy.maximum / numpy.minimum propagate NaN; fmaxf/fminf do not.
__device__ __forceinline__ float np_max(float a, float b) { return (a != a) ? a : ((b != b) ? b : fmaxf(a, b)); }
__device__ __forceinline__ float np_min(float a, float b) { return (a != a) ? a : ((b != b) ? b : fminf(a, b)); }

}  // namespace ia
