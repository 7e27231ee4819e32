// ia_decode_nms.cu — K2 (head decode + confidence filter + stream compaction) and
// K3 (sort + per-class greedy bitmask NMS + row assembly), sm_100a.
//
// Replaces, per image, the reference's v5 decode (src/shared/data/curator.py:289-307),
// the v8 parse (architectures/monolithic/app/postprocess.py:33-41), apply_nms
// (postprocess.py:76-160 == curator.py:323-400) and the row assembly (postprocess.py:
// 52-73).
//
// K2, v5 layout: the (A, 85) rows of one image are only 4-byte aligned (340-byte pitch),
// so a CTA pulls a tile of 128 anchors (43 520 B, a 16-byte multiple) into shared memory
// with one bulk async copy (TMA engine) and each thread then walks ONE anchor at a stride
// of 85 words (85 mod 32 = 21, coprime with 32: conflict-free).  Anchors that pass
// `obj * max(cls) >= conf` are compacted with warp ballot + prefix popcount + one atomic
// per warp into a per-image list of 64-bit sort keys
//      [ class : 10 | ~monotone(conf) : 32 | ~anchor : 22 ]
// so that an ascending sort yields class-ascending, confidence-descending order — the
// order apply_nms returns.  HBM traffic: the head, once.
//
// K3: one CTA per image sorts the keys (bitonic, shared memory up to 2048 candidates,
// global scratch above), then walks the sorted list 64 candidates at a time: a candidate
// is first tested against the boxes already kept for its class (warp per kept box, ballot
// over candidates), then the 64x64 intra-block suppression bitmask is built with ballots
// and resolved serially on 64-bit words.  All IoU arithmetic is float32 in numpy's
// operation order with no FMA contraction.
#include "ia_common.cuh"

namespace ia {

constexpr int kTileAnchors = 128;
constexpr int kDecodeThreads = 128;
constexpr int kNmsThreads = 256;
constexpr int kNmsWarps = kNmsThreads / 32;
constexpr int kSmemSortMax = 2048;
constexpr int kIdxBits = 22;
constexpr int kClsShift = 54;
constexpr int kMaxAnchors = 1 << kIdxBits;
constexpr int kMaxClasses = 1023;

__device__ __forceinline__ uint32_t mono_bits(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float mono_unbits(uint32_t m) {
  return __uint_as_float((m & 0x80000000u) ? (m ^ 0x80000000u) : ~m);
}
__device__ __forceinline__ uint64_t make_key(int cls, float conf, int anchor) {
  return ((uint64_t)cls << kClsShift) | ((uint64_t)(~mono_bits(conf)) << kIdxBits) |
         (uint64_t)((~(uint32_t)anchor) & (kMaxAnchors - 1));
}
__device__ __forceinline__ void split_key(uint64_t key, int& cls, float& conf, int& anchor) {
  cls = (int)(key >> kClsShift);
  conf = mono_unbits(~(uint32_t)(key >> kIdxBits));
  anchor = (int)((~(uint32_t)key) & (kMaxAnchors - 1));
}

// Append the passing lanes of a warp to the image's candidate list.
__device__ __forceinline__ void compact_append(bool pass, uint64_t key, int* counter,
                                               uint64_t* keys, int cap) {
  const unsigned lane = threadIdx.x & 31;
  const unsigned m = __ballot_sync(0xffffffffu, pass);
  if (m == 0) return;
  const int leader = __ffs(m) - 1;
  int base = 0;
  if ((int)lane == leader) base = atomicAdd(counter, __popc(m));
  base = __shfl_sync(0xffffffffu, base, leader);
  if (pass) {
    const int slot = base + __popc(m & ((1u << lane) - 1u));
    if (slot < cap) keys[slot] = key;
  }
}

struct DecodeParams {
  const float* head;
  int n, anchors, classes, row_floats;
  int tile_anchors, tiles;
  float conf_thr;
  int cap;           // max candidates kept per image
  size_t key_pitch;  // keys per image in the workspace (power of two >= cap)
  int* counts;
  uint64_t* keys;
};

// ------------------------------------------------------------------------------- K2 (v5)
__global__ void __launch_bounds__(kDecodeThreads) decode_filter_v5_kernel(const DecodeParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar;
  const int tid = threadIdx.x;
  const int img = blockIdx.x / p.tiles;
  const int tile = blockIdx.x - img * p.tiles;
  const int a0 = tile * p.tile_anchors;
  const int na = min(p.tile_anchors, p.anchors - a0);
  const float* g = p.head + ((size_t)img * p.anchors + a0) * p.row_floats;
  const uint32_t len = (uint32_t)na * p.row_floats * 4u;
  const uint32_t mis = (uint32_t)((uintptr_t)g & 15);

  if (tid == 0) {
    mbar_init(&bar, 1);
    mbar_fence_init();
  }
  __syncthreads();
  if (tid < 32) {
    const uint32_t head = min(len, (16u - mis) & 15u);
    const uint32_t body = (len - head) & ~15u;
    const uint32_t tail = len - head - body;
    const uint8_t* gb = reinterpret_cast<const uint8_t*>(g);
    uint8_t* s = smem + mis;
    if (tid == 0) {
      mbar_expect_tx(&bar, body);
      if (body) bulk_g2s(s + head, gb + head, body, &bar);
    }
    // unaligned ends (at most three 4-byte words each)
    if ((uint32_t)tid * 4u < head) *reinterpret_cast<float*>(s + tid * 4) = g[tid];
    if ((uint32_t)tid * 4u < tail)
      *reinterpret_cast<float*>(s + head + body + tid * 4) =
          *reinterpret_cast<const float*>(gb + head + body + tid * 4);
  }
  __syncthreads();
  mbar_wait(&bar, 0);

  bool pass = false;
  uint64_t key = 0;
  if (tid < na) {
    const float* row = reinterpret_cast<const float*>(smem + mis) + (size_t)tid * p.row_floats;
    const float obj = row[4];
    float best = row[5];
    int best_c = 0;
    bool has_nan = best != best;
#pragma unroll 8
    for (int c = 1; c < p.classes; ++c) {
      const float v = row[5 + c];
      has_nan |= (v != v);
      if (v > best) {  // strict: first maximum wins, like numpy argmax
        best = v;
        best_c = c;
      }
    }
    const float conf = __fmul_rn(obj, best);  // curator.py:303
    pass = !has_nan && (conf >= p.conf_thr);  // postprocess.py:99
    key = make_key(best_c, conf, a0 + tid);
  }
  compact_append(pass, key, p.counts + img, p.keys + (size_t)img * p.key_pitch, p.cap);
}

// ------------------------------------------------------------------------------- K2 (v8)
// Channel-major (4+C, A): consecutive threads take consecutive anchors, so every channel
// read is a fully coalesced 128-byte line; no staging needed.
__global__ void __launch_bounds__(kDecodeThreads) decode_filter_v8_kernel(const DecodeParams p) {
  const int img = blockIdx.x / p.tiles;
  const int tile = blockIdx.x - img * p.tiles;
  const int a = tile * kDecodeThreads + threadIdx.x;
  const float* g = p.head + (size_t)img * p.row_floats * p.anchors;
  bool pass = false;
  uint64_t key = 0;
  if (a < p.anchors) {
    const float* col = g + (size_t)4 * p.anchors + a;
    float best = __ldg(col);
    int best_c = 0;
    bool has_nan = best != best;
#pragma unroll 8
    for (int c = 1; c < p.classes; ++c) {
      const float v = __ldg(col + (size_t)c * p.anchors);
      has_nan |= (v != v);
      if (v > best) {
        best = v;
        best_c = c;
      }
    }
    pass = !has_nan && (best >= p.conf_thr);
    key = make_key(best_c, best, a);
  }
  compact_append(pass, key, p.counts + img, p.keys + (size_t)img * p.key_pitch, p.cap);
}

// Same, four consecutive anchors per thread: 128-bit loads (a warp reads 512 contiguous bytes
// of a channel row) and four independent max/arg-max chains.  Needs anchors % 4 == 0 and a
// 16-byte aligned head (the host falls back to the scalar kernel otherwise).
__global__ void __launch_bounds__(kDecodeThreads) decode_filter_v8x4_kernel(const DecodeParams p) {
  const int img = blockIdx.x / p.tiles;
  const int tile = blockIdx.x - img * p.tiles;
  const int a = (tile * kDecodeThreads + threadIdx.x) * 4;
  const float* g = p.head + (size_t)img * p.row_floats * p.anchors;
  bool pass[4] = {false, false, false, false};
  float best[4] = {0.f, 0.f, 0.f, 0.f};
  int best_c[4] = {0, 0, 0, 0};
  if (a < p.anchors) {
    const float4* col = reinterpret_cast<const float4*>(g + (size_t)4 * p.anchors + a);
    const size_t pitch4 = (size_t)p.anchors >> 2;
    const float4 v0 = __ldg(col);
    best[0] = v0.x, best[1] = v0.y, best[2] = v0.z, best[3] = v0.w;
    bool has_nan[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) has_nan[k] = best[k] != best[k];
#pragma unroll 8
    for (int c = 1; c < p.classes; ++c) {
      const float4 q = __ldg(col + (size_t)c * pitch4);
      const float v[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        has_nan[k] |= (v[k] != v[k]);
        if (v[k] > best[k]) {
          best[k] = v[k];
          best_c[k] = c;
        }
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) pass[k] = !has_nan[k] && (best[k] >= p.conf_thr);
  }
#pragma unroll
  for (int k = 0; k < 4; ++k)
    compact_append(pass[k], make_key(best_c[k], best[k], a + k), p.counts + img,
                   p.keys + (size_t)img * p.key_pitch, p.cap);
}

// -------------------------------------------------- filter for already-decoded arrays
struct FilterParams {
  const float* scores;
  const int* class_ids;
  int m;
  float conf_thr;
  int cap;
  int* counts;
  uint64_t* keys;
};
__global__ void __launch_bounds__(kDecodeThreads) filter_decoded_kernel(const FilterParams p) {
  const int i = blockIdx.x * kDecodeThreads + threadIdx.x;
  bool pass = false;
  uint64_t key = 0;
  if (i < p.m) {
    const float s = p.scores[i];
    pass = s >= p.conf_thr;
    key = make_key(p.class_ids[i], s, i);
  }
  compact_append(pass, key, p.counts, p.keys, p.cap);
}

// ------------------------------------------------------------------------------------ K3
struct NmsParams {
  const float* boxes;      // first box component of image 0
  size_t image_stride;     // floats between images
  int anchor_stride;       // floats between anchors
  int comp_stride;         // floats between cx,cy,w,h of one anchor
  int n, cap;
  size_t key_pitch;
  const int* counts;
  uint64_t* keys;
  float iou_thr;
  int max_det;
  float* dets;   // [n, max_det, 6] or null
  int* keep;     // [n, max_det]
  int* count_out;
  int* status;
};

__device__ __forceinline__ void bitonic_sort(uint64_t* a, int P) {
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < P; i += kNmsThreads) {
        const int l = i ^ j;
        if (l > i) {
          const uint64_t x = a[i], y = a[l];
          const bool up = (i & k) == 0;
          if ((x > y) == up) {
            a[i] = y;
            a[l] = x;
          }
        }
      }
      __syncthreads();
    }
  }
}

// !(iou <= thr) in numpy's float32 operation order (postprocess.py:139-157).
__device__ __forceinline__ bool suppresses(float ax1, float ay1, float ax2, float ay2, float aarea,
                                           float bx1, float by1, float bx2, float by2, float barea,
                                           float thr) {
  const float w = np_max(0.f, __fsub_rn(np_min(ax2, bx2), np_max(ax1, bx1)));
  const float h = np_max(0.f, __fsub_rn(np_min(ay2, by2), np_max(ay1, by1)));
  const float inter = __fmul_rn(w, h);
  const float uni = __fsub_rn(__fadd_rn(aarea, barea), inter);
  const float iou = __fdiv_rn(inter, __fadd_rn(uni, 1e-6f));
  return !(iou <= thr);
}

__global__ void __launch_bounds__(kNmsThreads) nms_kernel(const NmsParams p) {
  __shared__ uint64_t skeys[kSmemSortMax];
  __shared__ float bx1[64], by1[64], bx2[64], by2[64], barea[64], bconf[64];
  __shared__ int bcls[64], bidx[64];
  __shared__ uint64_t rowmask[64];
  __shared__ unsigned removed_lo, removed_hi;
  __shared__ int klist[64];
  __shared__ int s_kcount, s_K, s_seg_cls, s_seg_start, s_status;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int img = blockIdx.x;
  const int total = p.counts[img];
  const int M = min(total, p.cap);
  uint64_t* gkeys = p.keys + (size_t)img * p.key_pitch;

  // ---- sort -------------------------------------------------------------------------
  int P = 1;
  while (P < M) P <<= 1;
  uint64_t* sorted;
  if (P <= kSmemSortMax) {
    for (int i = tid; i < P; i += kNmsThreads) skeys[i] = i < M ? gkeys[i] : ~0ull;
    sorted = skeys;
  } else {
    for (int i = M + tid; i < P; i += kNmsThreads) gkeys[i] = ~0ull;
    sorted = gkeys;
  }
  if (tid == 0) {
    s_K = 0;
    s_seg_cls = -1;
    s_seg_start = 0;
    s_status = total > p.cap ? 1 : 0;
  }
  __syncthreads();
  bitonic_sort(sorted, P);

  const float* boxes = p.boxes + (size_t)img * p.image_stride;
  float* dets = p.dets ? p.dets + (size_t)img * p.max_det * 6 : nullptr;
  int* keep = p.keep + (size_t)img * p.max_det;

  for (int base = 0; base < M; base += 64) {
    const int nb = min(64, M - base);
    // ---- load 64 sorted candidates, gather their boxes --------------------------------
    if (tid < 64) {
      rowmask[tid] = 0;
      if (tid < nb) {
        int cls, anchor;
        float conf;
        split_key(sorted[base + tid], cls, conf, anchor);
        const float* b = boxes + (size_t)anchor * p.anchor_stride;
        const float cx = __ldg(b), cy = __ldg(b + p.comp_stride);
        const float hw = __fmul_rn(__ldg(b + 2 * (size_t)p.comp_stride), 0.5f);
        const float hh = __fmul_rn(__ldg(b + 3 * (size_t)p.comp_stride), 0.5f);
        const float x1 = __fsub_rn(cx, hw), y1 = __fsub_rn(cy, hh);  // postprocess.py:110-113
        const float x2 = __fadd_rn(cx, hw), y2 = __fadd_rn(cy, hh);
        bx1[tid] = x1;
        by1[tid] = y1;
        bx2[tid] = x2;
        by2[tid] = y2;
        barea[tid] = __fmul_rn(__fsub_rn(x2, x1), __fsub_rn(y2, y1));
        bconf[tid] = conf;
        bcls[tid] = cls;
        bidx[tid] = anchor;
      } else {
        bcls[tid] = -1;
      }
    }
    if (tid == 0) {
      removed_lo = 0;
      removed_hi = 0;
    }
    __syncthreads();

    const int K = s_K;
    const int first_cls = bcls[0];
    // ---- phase A: against boxes already kept for the block's first class ---------------
    if (dets && s_seg_cls == first_cls) {
      const int seg_start = s_seg_start;
      const bool v0 = lane < nb && bcls[lane] == first_cls;
      const bool v1 = lane + 32 < nb && bcls[lane + 32] == first_cls;
      unsigned acc_lo = 0, acc_hi = 0;
      for (int k = seg_start + warp; k < K; k += kNmsWarps) {
        const float* r = dets + (size_t)k * 6;
        const float kx1 = r[0], ky1 = r[1], kx2 = r[2], ky2 = r[3];
        const float karea = __fmul_rn(__fsub_rn(kx2, kx1), __fsub_rn(ky2, ky1));
        const bool s0 = v0 && suppresses(kx1, ky1, kx2, ky2, karea, bx1[lane], by1[lane], bx2[lane],
                                         by2[lane], barea[lane], p.iou_thr);
        const bool s1 = v1 && suppresses(kx1, ky1, kx2, ky2, karea, bx1[lane + 32], by1[lane + 32],
                                         bx2[lane + 32], by2[lane + 32], barea[lane + 32], p.iou_thr);
        acc_lo |= __ballot_sync(0xffffffffu, s0);
        acc_hi |= __ballot_sync(0xffffffffu, s1);
      }
      if (lane == 0) {
        if (acc_lo) atomicOr(&removed_lo, acc_lo);
        if (acc_hi) atomicOr(&removed_hi, acc_hi);
      }
    }
    __syncthreads();

    // ---- phase B: greedy resolve inside the block --------------------------------------
    // A row of the 64x64 suppression matrix is only needed for a box that is KEPT.  Warp 0
    // therefore starts on demand: take the first alive candidate, keep it, test it against
    // the alive candidates behind it (two per lane, one ballot each), clear what it
    // suppresses.  Images with duplicates around a few objects finish within a few rows.  If
    // more than kLazyRows boxes of the block are kept (little overlap), the remaining rows are
    // computed by all warps at once and resolved serially on 64-bit words.
    constexpr int kLazyRows = 8;
    if (warp == 0) {
      uint64_t alive = (nb == 64 ? ~0ull : ((1ull << nb) - 1ull)) &
                       ~((uint64_t)removed_lo | ((uint64_t)removed_hi << 32));
      int kc = 0;
      int seg_cls = s_seg_cls, seg_start = s_seg_start;
      const int j0 = lane, j1 = lane + 32;
      // Every class run of the block keeps at least its first live box: with more runs than lazy
      // rows the loop below cannot finish the block and would only delay the parallel path.
      const unsigned run0 = __ballot_sync(0xffffffffu, j0 < nb && (j0 == 0 || bcls[j0] != bcls[j0 - 1]));
      const unsigned run1 = __ballot_sync(0xffffffffu, j1 < nb && bcls[j1] != bcls[j1 - 1]);
      const int lazy_rows = __popc(run0) + __popc(run1) > kLazyRows ? 0 : kLazyRows;
      while (alive && kc < lazy_rows) {
        const int i = __ffsll((long long)alive) - 1;
        const int ci = bcls[i];
        if (lane == 0) klist[kc] = i;
        if (ci != seg_cls) {
          seg_cls = ci;
          seg_start = K + kc;
        }
        ++kc;
        const float ix1 = bx1[i], iy1 = by1[i], ix2 = bx2[i], iy2 = by2[i], ia_ = barea[i];
        const bool s0 = j0 > i && ((alive >> j0) & 1ull) && bcls[j0] == ci &&
                        suppresses(ix1, iy1, ix2, iy2, ia_, bx1[j0], by1[j0], bx2[j0], by2[j0], barea[j0], p.iou_thr);
        const bool s1 = j1 > i && ((alive >> j1) & 1ull) && bcls[j1] == ci &&
                        suppresses(ix1, iy1, ix2, iy2, ia_, bx1[j1], by1[j1], bx2[j1], by2[j1], barea[j1], p.iou_thr);
        const unsigned lo = __ballot_sync(0xffffffffu, s0);
        const unsigned hi = __ballot_sync(0xffffffffu, s1);
        alive &= ~((uint64_t)lo | ((uint64_t)hi << 32));
        alive &= ~(1ull << i);
      }
      if (lane == 0) {
        s_kcount = kc;
        s_seg_cls = seg_cls;
        s_seg_start = seg_start;
        removed_lo = (unsigned)alive;  // re-used: the candidates still undecided
        removed_hi = (unsigned)(alive >> 32);
      }
    }
    __syncthreads();
    const uint64_t pending = (uint64_t)removed_lo | ((uint64_t)removed_hi << 32);
    if (pending) {  // uniform: the rest of the block in parallel
      for (int i = warp; i < nb; i += kNmsWarps) {
        if (!((pending >> i) & 1ull)) continue;
        const int ci = bcls[i];
        const float ix1 = bx1[i], iy1 = by1[i], ix2 = bx2[i], iy2 = by2[i], ia_ = barea[i];
        const int j0 = lane, j1 = lane + 32;
        const bool s0 = j0 > i && ((pending >> j0) & 1ull) && bcls[j0] == ci &&
                        suppresses(ix1, iy1, ix2, iy2, ia_, bx1[j0], by1[j0], bx2[j0], by2[j0], barea[j0], p.iou_thr);
        const bool s1 = j1 > i && ((pending >> j1) & 1ull) && bcls[j1] == ci &&
                        suppresses(ix1, iy1, ix2, iy2, ia_, bx1[j1], by1[j1], bx2[j1], by2[j1], barea[j1], p.iou_thr);
        const unsigned lo = __ballot_sync(0xffffffffu, s0);
        const unsigned hi = __ballot_sync(0xffffffffu, s1);
        if (lane == 0) rowmask[i] = (uint64_t)lo | ((uint64_t)hi << 32);
      }
      __syncthreads();
      if (warp == 0) {
        // Greedy resolve.  Only a row that suppresses something can change the outcome, so the
        // serial walk (every lane, the same words) visits those rows alone, in order, each only
        // if its box is still alive when its turn comes; what is alive at the end is kept.
        const int j0 = lane, j1 = lane + 32;
        const bool nz0 = ((pending >> j0) & 1ull) && rowmask[j0] != 0;
        const bool nz1 = ((pending >> j1) & 1ull) && rowmask[j1] != 0;
        uint64_t todo = (uint64_t)__ballot_sync(0xffffffffu, nz0) | ((uint64_t)__ballot_sync(0xffffffffu, nz1) << 32);
        uint64_t alive = pending;
        while (todo) {
          const int i = __ffsll((long long)todo) - 1;
          todo &= todo - 1;
          if ((alive >> i) & 1ull) alive &= ~rowmask[i];
        }
        // the kept boxes in order: position = boxes kept before it
        const int kc0 = s_kcount;
        const bool k0 = (alive >> j0) & 1ull, k1 = (alive >> j1) & 1ull;
        if (k0) klist[kc0 + __popcll(alive & ((1ull << j0) - 1ull))] = j0;
        if (k1) klist[kc0 + __popcll(alive & ((1ull << j1) - 1ull))] = j1;
        // the class run the next block may continue: the last kept box's, from its first kept box
        // (classes ascend: if that class is the one already open, the run simply goes on)
        int seg_cls = s_seg_cls, seg_start = s_seg_start;
        if (alive) {
          const int cl = bcls[63 - __clzll((long long)alive)];
          if (cl != seg_cls) {
            const uint64_t same = (uint64_t)__ballot_sync(0xffffffffu, k0 && bcls[j0] == cl) |
                                  ((uint64_t)__ballot_sync(0xffffffffu, k1 && bcls[j1] == cl) << 32);
            const int f = __ffsll((long long)same) - 1;
            seg_cls = cl;
            seg_start = K + kc0 + __popcll(alive & ((1ull << f) - 1ull));
          }
        }
        __syncwarp();
        if (lane == 0) {
          s_kcount = kc0 + __popcll(alive);
          s_seg_cls = seg_cls;
          s_seg_start = seg_start;
        }
      }
      __syncthreads();
    }

    // ---- append the kept rows (postprocess.py:56-73) ----------------------------------
    const int kc = s_kcount;
    if (tid < kc) {
      const int i = klist[tid];
      const int k = K + tid;
      if (k < p.max_det) {
        keep[k] = bidx[i];
        if (dets) {
          float* r = dets + (size_t)k * 6;
          r[0] = bx1[i];
          r[1] = by1[i];
          r[2] = bx2[i];
          r[3] = by2[i];
          r[4] = bconf[i];
          r[5] = (float)bcls[i];
        }
      }
    }
    __syncthreads();
    if (tid == 0) s_K = min(K + kc, p.max_det);
    if (K + kc > p.max_det || (K + kc == p.max_det && base + 64 < M)) {
      // out of rows: later suppression tests could not see every kept box
      if (tid == 0 && (K + kc > p.max_det || base + 64 < M)) s_status |= 2;
      __syncthreads();
      break;
    }
    __syncthreads();
  }
  if (tid == 0) {
    p.count_out[img] = s_K;
    p.status[img] = s_status;
  }
}

// Without a dets buffer phase A has nothing to read kept boxes from; ia_nms therefore
// always hands the kernel a scratch dets buffer carved from the workspace.

static size_t pow2_ceil(size_t v) {
  size_t p = 64;
  while (p < v) p <<= 1;
  return p;
}
static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct Workspace {
  int* counts;
  uint64_t* keys;
  size_t key_pitch;
  size_t bytes;
};
static Workspace carve(void* base, int n, int cap) {
  Workspace w;
  w.key_pitch = pow2_ceil((size_t)cap);
  const size_t counts_bytes = align_up((size_t)n * sizeof(int), 256);
  w.counts = reinterpret_cast<int*>(base);
  w.keys = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(base) + counts_bytes);
  w.bytes = counts_bytes + (size_t)n * w.key_pitch * sizeof(uint64_t);
  return w;
}

}  // namespace ia

extern "C" size_t ia_decode_nms_workspace_bytes(int n, int anchors, int max_candidates) {
  if (n < 0 || anchors < 1) return 0;
  int cap = max_candidates < 1 || max_candidates > anchors ? anchors : max_candidates;
  return ia::carve(nullptr, n > 0 ? n : 1, cap).bytes;
}

extern "C" int ia_decode_nms(const float* head_dev, int layout, int n, int anchors, int num_classes,
                             float conf_threshold, float iou_threshold, int max_candidates,
                             int max_det, float* dets_dev, int32_t* keep_dev, int32_t* count_dev,
                             int32_t* status_dev, void* workspace_dev, size_t workspace_bytes,
                             void* stream_) {
  using namespace ia;
  cudaStream_t stream = (cudaStream_t)stream_;
  IA_RETURN_IF(n < 0 || anchors < 1 || anchors > kMaxAnchors || num_classes < 1 ||
                   num_classes > kMaxClasses || max_det < 1,
               IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(layout != IA_HEAD_LAYOUT_V5 && layout != IA_HEAD_LAYOUT_V8, IA_ERR_INVALID_ARGUMENT);
  if (n == 0) return IA_OK;
  IA_RETURN_IF(!head_dev || !dets_dev || !keep_dev || !count_dev || !status_dev || !workspace_dev,
               IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(((uintptr_t)workspace_dev & 255) != 0 || ((uintptr_t)head_dev & 3) != 0,
               IA_ERR_INVALID_ARGUMENT);
  const int cap = max_candidates < 1 || max_candidates > anchors ? anchors : max_candidates;
  Workspace ws = carve(workspace_dev, n, cap);
  IA_RETURN_IF(workspace_bytes < ws.bytes, IA_ERR_WORKSPACE_TOO_SMALL);

  cudaError_t e = cudaMemsetAsync(ws.counts, 0, (size_t)n * sizeof(int), stream);
  if (e != cudaSuccess) return (int)e;

  DecodeParams d;
  d.head = head_dev;
  d.n = n;
  d.anchors = anchors;
  d.classes = num_classes;
  d.conf_thr = conf_threshold;
  d.cap = cap;
  d.key_pitch = ws.key_pitch;
  d.counts = ws.counts;
  d.keys = ws.keys;

  NmsParams q;
  if (layout == IA_HEAD_LAYOUT_V5) {
    d.row_floats = 5 + num_classes;
    int tile = kTileAnchors;
    while (tile > 1 && (size_t)tile * d.row_floats * 4 + 16 > 200 * 1024) tile >>= 1;
    IA_RETURN_IF((size_t)tile * d.row_floats * 4 + 16 > 200 * 1024, IA_ERR_UNSUPPORTED);
    d.tile_anchors = tile;
    d.tiles = (anchors + tile - 1) / tile;
    const long long blocks = (long long)d.tiles * n;
    IA_RETURN_IF(blocks > 0x7fffffffLL, IA_ERR_UNSUPPORTED);
    const size_t smem = (size_t)tile * d.row_floats * 4 + 16;
    e = allow_max_dynamic_smem(reinterpret_cast<const void*>(decode_filter_v5_kernel));
    if (e != cudaSuccess) return (int)e;
    decode_filter_v5_kernel<<<(unsigned)blocks, kDecodeThreads, smem, stream>>>(d);
    q.image_stride = (size_t)anchors * d.row_floats;
    q.anchor_stride = d.row_floats;
    q.comp_stride = 1;
  } else {
    d.row_floats = 4 + num_classes;
    const bool vec4 = (anchors & 3) == 0 && ((uintptr_t)head_dev & 15) == 0;
    d.tile_anchors = vec4 ? 4 * kDecodeThreads : kDecodeThreads;
    d.tiles = (anchors + d.tile_anchors - 1) / d.tile_anchors;
    const long long blocks = (long long)d.tiles * n;
    IA_RETURN_IF(blocks > 0x7fffffffLL, IA_ERR_UNSUPPORTED);
    if (vec4) decode_filter_v8x4_kernel<<<(unsigned)blocks, kDecodeThreads, 0, stream>>>(d);
    else decode_filter_v8_kernel<<<(unsigned)blocks, kDecodeThreads, 0, stream>>>(d);
    q.image_stride = (size_t)anchors * d.row_floats;
    q.anchor_stride = 1;
    q.comp_stride = anchors;
  }
  int rc = launch_status();
  if (rc != IA_OK) return rc;

  q.boxes = head_dev;
  q.n = n;
  q.cap = cap;
  q.key_pitch = ws.key_pitch;
  q.counts = ws.counts;
  q.keys = ws.keys;
  q.iou_thr = iou_threshold;
  q.max_det = max_det;
  q.dets = dets_dev;
  q.keep = keep_dev;
  q.count_out = count_dev;
  q.status = status_dev;
  nms_kernel<<<n, kNmsThreads, 0, stream>>>(q);
  return launch_status();
}

extern "C" int ia_nms(const float* boxes_dev, const float* scores_dev, const int32_t* class_ids_dev,
                      int m, float conf_threshold, float iou_threshold, int max_det,
                      int32_t* keep_dev, int32_t* count_dev, int32_t* status_dev,
                      void* workspace_dev, size_t workspace_bytes, void* stream_) {
  using namespace ia;
  cudaStream_t stream = (cudaStream_t)stream_;
  IA_RETURN_IF(!keep_dev || !count_dev || !status_dev || !workspace_dev, IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(m < 0 || m > kMaxAnchors || max_det < 1, IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(((uintptr_t)workspace_dev & 255) != 0, IA_ERR_INVALID_ARGUMENT);
  if (m == 0) {
    cudaError_t e = cudaMemsetAsync(count_dev, 0, sizeof(int), stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(status_dev, 0, sizeof(int), stream);
    return e == cudaSuccess ? IA_OK : (int)e;
  }
  IA_RETURN_IF(!boxes_dev || !scores_dev || !class_ids_dev, IA_ERR_INVALID_ARGUMENT);
  Workspace ws = carve(workspace_dev, 1, m);
  // scratch rows for the kept boxes live behind the keys
  const size_t dets_bytes = (size_t)max_det * 6 * sizeof(float);
  IA_RETURN_IF(workspace_bytes < ws.bytes + dets_bytes, IA_ERR_WORKSPACE_TOO_SMALL);
  float* dets = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace_dev) + ws.bytes);

  cudaError_t e = cudaMemsetAsync(ws.counts, 0, sizeof(int), stream);
  if (e != cudaSuccess) return (int)e;
  FilterParams f;
  f.scores = scores_dev;
  f.class_ids = class_ids_dev;
  f.m = m;
  f.conf_thr = conf_threshold;
  f.cap = m;
  f.counts = ws.counts;
  f.keys = ws.keys;
  filter_decoded_kernel<<<(m + kDecodeThreads - 1) / kDecodeThreads, kDecodeThreads, 0, stream>>>(f);
  int rc = launch_status();
  if (rc != IA_OK) return rc;

  NmsParams q;
  q.boxes = boxes_dev;
  q.image_stride = 0;
  q.anchor_stride = 4;
  q.comp_stride = 1;
  q.n = 1;
  q.cap = m;
  q.key_pitch = ws.key_pitch;
  q.counts = ws.counts;
  q.keys = ws.keys;
  q.iou_thr = iou_threshold;
  q.max_det = max_det;
  q.dets = dets;
  q.keep = keep_dev;
  q.count_out = count_dev;
  q.status = status_dev;
  nms_kernel<<<1, kNmsThreads, 0, stream>>>(q);
  return launch_status();
}

// ------------------------------------------------------- allocation-free variant
// ia_decode_nms_static: the same decode + filter + NMS, but the candidate lists live in a
// fixed array inside the library image instead of a caller-supplied workspace.  It is meant
// for callers that have no device memory of their own to lend at call time and must not
// call an allocator: a serving loop whose HBM is entirely owned by another pool (an
// inference server's arena, a spent MPS/MIG memory quota), or a latency-critical thread
// that may never enter cudaMalloc/cudaFree because those serialise against the whole
// device.  Such a caller still passes the head and the output buffers it already owns.
#include <mutex>

namespace ia {

constexpr int kStaticMaxImages = IA_STATIC_MAX_IMAGES;
constexpr int kStaticMaxCandidates = IA_STATIC_MAX_CANDIDATES;
// counters (one 256-byte block) + one key list of kStaticMaxCandidates per image: 524 544 B
constexpr size_t kStaticScratchBytes =
    256 + (size_t)kStaticMaxImages * kStaticMaxCandidates * sizeof(uint64_t);

// Every GPU that loads the library holds these bytes until the process exits.
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

// In host code `g_static_scratch` is only a handle for the symbol, not an address a kernel
// can use.  The real device address differs per GPU, so it is looked up once per device,
// with that device current, and cached.
static std::mutex g_static_mutex;
static void* g_static_address[64] = {nullptr};

static int static_scratch_address(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= 64) return IA_ERR_UNSUPPORTED;
  std::lock_guard<std::mutex> lock(g_static_mutex);
  if (g_static_address[dev] == nullptr) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_static_scratch);
    if (e != cudaSuccess) return (int)e;
    if (ptr == nullptr) return IA_ERR_UNSUPPORTED;
    g_static_address[dev] = ptr;
  }
  *out = g_static_address[dev];
  return IA_OK;
}

}  // namespace ia

extern "C" int ia_decode_nms_static(const float* head_dev, int layout, int n, int anchors,
                                    int num_classes, float conf_threshold, float iou_threshold,
                                    int max_candidates, int max_det, float* dets_dev,
                                    int32_t* keep_dev, int32_t* count_dev, int32_t* status_dev,
                                    void* stream) {
  using namespace ia;
  // Sizes first: anything that would not fit the fixed array is refused, never truncated.
  IA_RETURN_IF(n < 0 || anchors < 1, IA_ERR_INVALID_ARGUMENT);
  IA_RETURN_IF(n > kStaticMaxImages, IA_ERR_CAPACITY);
  const int cap = max_candidates < 1 || max_candidates > anchors ? anchors : max_candidates;
  IA_RETURN_IF(cap > kStaticMaxCandidates, IA_ERR_CAPACITY);
  void* scratch = nullptr;
  const int rc = static_scratch_address(&scratch);  // first call on a device: one-time look-up
  if (rc != IA_OK) return rc;
  // The array still holds the previous call's keys and counters.  Only the counters must
  // start at zero; ia_decode_nms clears them with a stream-ordered memset before the filter
  // kernel runs, and every key slot that is read was written by this call first.
  return ia_decode_nms(head_dev, layout, n, anchors, num_classes, conf_threshold, iou_threshold, cap,
                       max_det, dets_dev, keep_dev, count_dev, status_dev, scratch,
                       kStaticScratchBytes, stream);
}
