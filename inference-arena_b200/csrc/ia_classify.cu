// ia_classify.cu — SURVEY §8f row N4: what the reference does to the classifier's logits.
//
//   softmax + top-k   architectures/microservices/classification/app/inference.py:112-132
//                     exp(s - max(s)) / sum(...), argsort(probs)[::-1][:k]
//   arg-max + value   architectures/monolithic/app/inference.py:201-202 and
//                     architectures/triton/gateway/app/pipeline.py:181-182
//
// One warp per row of logits.  The row is staged in shared memory once (the only DRAM
// traffic: classes * 4 bytes per row) and exponentiated in place with the float32 operations
// numpy performs (subtract, exp; the sum is accumulated in double and rounded once, which
// lies within one float32 ulp of numpy's pairwise sum).  Taking the k largest must not cost k
// passes over the row, so the warp first finds a bound: the k-th largest of the 32 per-lane
// maxima is <= the k-th largest element, hence every winner is >= it.  One more pass
// compacts the few elements above that bound (typically k..2k of them) with ballots, only
// those are divided by the sum, and the order among them is settled by k warp-wide maxima of
// a 64-bit key (ordered value bits : index).  Rows where that shortcut does not apply (NaN or
// infinite logits, masses of equal values, k > 32) take k plain passes instead.
//
// Rows of up to 1024 classes (MobileNetV2: 1000) live in 32 registers per lane and never touch
// shared memory except for the candidate list; their softmax sums a fast exponential and
// chooses the candidates on the logits (see classify_topk_reg_kernel); longer rows are staged
// in shared memory and follow the scheme above literally.
#include <math_constants.h>

#include "ia_common.cuh"

namespace ia {

constexpr int kClassifyMaxWarps = 8;
constexpr size_t kClassifySmemBudget = 96 * 1024;   // per CTA, keeps >= 2 CTAs per SM
constexpr size_t kClassifySmemLimit = 200 * 1024;   // one row must fit

struct ClassifyParams {
  const float* logits;
  int rows, classes, k;
  int softmax;          // 1: softmax probabilities; 0: raw values
  int low_index_first;  // tie rule: 0 = argsort()[::-1] of a stable sort, 1 = np.argmax
  int warps;
  int32_t* ids;
  float* values;
  int32_t* status;
};

// float -> uint32 whose unsigned order is numpy's sort order: -inf < ... < -0 == +0 < ... <
// +inf < NaN (NaNs sort last, so they come first in argsort()[::-1] and win np.argmax).
__device__ __forceinline__ uint32_t order_bits(float x) {
  if (x != x) return 0xffffffffu;
  if (x == 0.f) return 0x80000000u;
  const uint32_t b = __float_as_uint(x);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

__device__ __forceinline__ uint64_t warp_max_u64(uint64_t v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    const uint64_t o = __shfl_xor_sync(0xffffffffu, v, d);
    v = o > v ? o : v;
  }
  return v;
}

constexpr int kCandidates = 64;  // per row, fast path

template <bool kSoftmax>
__global__ void __launch_bounds__(kClassifyMaxWarps * 32) classify_topk_kernel(const ClassifyParams p) {
  extern __shared__ float rows_smem[];
  __shared__ uint64_t cand_smem[kClassifyMaxWarps][kCandidates];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * p.warps + warp;
  if (row >= p.rows) return;  // whole warps leave; no CTA-wide barrier below
  float* v = rows_smem + (size_t)warp * p.classes;
  uint64_t* cand = cand_smem[warp];
  const float* g = p.logits + (size_t)row * p.classes;
  const int C = p.classes, k = p.k;
  const uint32_t flip = p.low_index_first ? 0xffffffffu : 0u;
  constexpr uint32_t kFull = 0xffffffffu;

  // pass A: stage the row, row maximum (np.max: NaN if any NaN)
  float m = -CUDART_INF_F;
  bool has_nan = false;
#pragma unroll 4
  for (int i = lane; i < C; i += 32) {
    const float x = __ldg(g + i);
    v[i] = x;
    has_nan |= (x != x);
    m = fmaxf(m, x);  // ignores NaN; restored below
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) m = fmaxf(m, __shfl_xor_sync(kFull, m, d));
  has_nan = __any_sync(kFull, has_nan);

  // pass B: exp in place + sum (softmax) / nothing (raw); per-lane maximum of the ordered bits
  uint32_t lane_max = 0;
  float total = 0.f;
  bool slow = k > 32;
  if (kSoftmax) {
    if (has_nan) m = CUDART_NAN_F;
    double acc = 0.0;
    bool bad = false;
#pragma unroll 4
    for (int i = lane; i < C; i += 32) {
      const float e = expf(__fsub_rn(v[i], m));
      v[i] = e;
      acc += (double)e;
      bad |= (e != e);
      lane_max = max(lane_max, __float_as_uint(e));  // e >= 0: the raw bits are ordered
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(kFull, acc, d);
    total = (float)acc;
    slow |= __any_sync(kFull, bad);
  } else {
#pragma unroll 4
    for (int i = lane; i < C; i += 32) lane_max = max(lane_max, order_bits(v[i]));
  }
  __syncwarp();

  // bound: k-th largest lane maximum (lanes without elements hold 0, below every real value)
  uint32_t bound = 0;
  if (!slow) {
    bool alive = true;
    for (int j = 0; j < k; ++j) {
      bound = __reduce_max_sync(kFull, alive ? lane_max : 0u);
      const uint32_t owners = __ballot_sync(kFull, alive && lane_max == bound);
      if (owners && lane == __ffs(owners) - 1) alive = false;
    }
    if (kSoftmax) {
      // p = e / total can round two different e to one p only if they are neighbours, so a
      // relative margin keeps every element that could tie with a winner; where p would be
      // denormal the margin argument fails -> plain passes
      const float e_bound = __uint_as_float(bound);
      slow = !(e_bound >= 1e-30f);
      bound = __float_as_uint(__fmul_rn(e_bound, 0.999999f));
    }
  }

  int found = 0;
  if (!slow) {
    for (int base = 0; base < C; base += 32) {
      const int i = base + lane;
      bool is_cand = false;
      float x = 0.f;
      if (i < C) {
        x = v[i];
        is_cand = (kSoftmax ? __float_as_uint(x) : order_bits(x)) >= bound;
      }
      const uint32_t mask = __ballot_sync(kFull, is_cand);
      if (mask) {
        const int pos = found + __popc(mask & ((1u << lane) - 1u));
        if (is_cand && pos < kCandidates) {
          const float val = kSoftmax ? __fdiv_rn(x, total) : x;
          cand[pos] = ((uint64_t)order_bits(val) << 32) | ((uint32_t)i ^ flip);
        }
        found += __popc(mask);
      }
    }
    slow = found > kCandidates;
    __syncwarp();
  }

  if (!slow) {
    uint64_t c0 = lane < found ? cand[lane] : 0ull;
    uint64_t c1 = lane + 32 < found ? cand[lane + 32] : 0ull;
    uint64_t prev = ~0ull;
    for (int j = 0; j < k; ++j) {
      const uint64_t a = c0 < prev ? c0 : 0ull, b = c1 < prev ? c1 : 0ull;
      prev = warp_max_u64(a > b ? a : b);
      if (lane == 0) {
        const int idx = (int)((uint32_t)prev ^ flip);
        p.ids[(size_t)row * k + j] = idx;
        p.values[(size_t)row * k + j] = kSoftmax ? __fdiv_rn(v[idx], total) : v[idx];
      }
    }
  } else {
    if (kSoftmax) {
      for (int i = lane; i < C; i += 32) v[i] = __fdiv_rn(v[i], total);
      __syncwarp();
    }
    uint64_t prev = ~0ull;
    for (int j = 0; j < k; ++j) {
      uint64_t best = 0;
      for (int i = lane; i < C; i += 32) {
        const uint64_t key = ((uint64_t)order_bits(v[i]) << 32) | ((uint32_t)i ^ flip);
        if (key < prev && key > best) best = key;
      }
      best = warp_max_u64(best);
      prev = best;
      if (lane == 0) {
        const int idx = (int)((uint32_t)best ^ flip);
        p.ids[(size_t)row * k + j] = idx;
        p.values[(size_t)row * k + j] = v[idx];
      }
    }
  }
  if (lane == 0 && p.status) p.status[row] = has_nan ? 1 : 0;
}

// ---------------------------------------------------------------- rows of <= 1024 classes
constexpr int kRegPerLane = 32;
constexpr int kRegMaxClasses = kRegPerLane * 32;

__device__ __forceinline__ uint32_t warp_max_u32(uint32_t v) { return __reduce_max_sync(0xffffffffu, v); }

// Two float32 lanes per instruction (sm_100 f32x2): (d0, d1) = (a0, a1) * b + c and (s0, s1) += (e0, e1),
// each lane rounded to nearest exactly like the scalar instruction.
__device__ __forceinline__ void fma_f32x2(float& d0, float& d1, float a0, float a1, float b, float c) {
  uint64_t a, bb, cc, d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(a0), "f"(a1));
  asm("mov.b64 %0, {%1, %1};" : "=l"(bb) : "f"(b));
  asm("mov.b64 %0, {%1, %1};" : "=l"(cc) : "f"(c));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(bb), "l"(cc));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(d0), "=f"(d1) : "l"(d));
}
__device__ __forceinline__ void add_f32x2(float& s0, float& s1, float e0, float e1) {
  uint64_t a, b, d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(s0), "f"(s1));
  asm("mov.b64 %0, {%1, %2};" : "=l"(b) : "f"(e0), "f"(e1));
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(s0), "=f"(s1) : "l"(d));
}

// A lower bound of the k-th largest of the 32 lanes' ordered keys (k <= 32), within 32 steps of
// it.  The low five bits of every key are replaced by the lane number: the keys become distinct,
// so a round is one warp maximum and "was it mine" — no vote for who owns a tie; clearing those
// bits is monotone, so the k-th largest of the cleared keys is the cleared k-th largest.
__device__ __forceinline__ uint32_t kth_largest_lane_key(uint32_t key, int lane, int k) {
  uint32_t mine = (key & ~31u) | (uint32_t)lane, top = 0;
#pragma unroll 1  // k is a run-time value: the unrolled forms cost more in loop control than a round
  for (int j = 0; j < k; ++j) {
    top = warp_max_u32(mine);
    mine = mine == top ? 0u : mine;  // (a key of 0 is lane 0's at most: still distinct)
  }
  return top & ~31u;
}

// Order-preserving bits of a float that is not NaN (-0 sorts just below +0, which no bound minds),
// and back.
__device__ __forceinline__ uint32_t order_bits_no_nan(float x) {
  const uint32_t b = __float_as_uint(x);
  return b ^ ((uint32_t)((int32_t)b >> 31) | 0x80000000u);
}
__device__ __forceinline__ float from_order_bits(uint32_t u) {
  return __uint_as_float(u ^ ((u & 0x80000000u) ? 0x80000000u : 0xffffffffu));
}

// Resident CTAs per SM.  The softmax variant is bound by instruction issue with the warps it has
// resident (~70 % of the issue slots: chains of warp reductions and shuffles), not by a few
// spilled registers: 4 / 5 / 6 CTAs (64 / 48 / 40 registers) take 56 / 51 / 54 us for 65 536 x
// 1000 logits.  The raw-value variant keeps a second 32-entry array and stays at 3.  Measured
// and dropped: a persistent form, every warp walking many rows with the next row's bulk copy in
// flight in a staging buffer (the issue slots stay ~70 % busy either way, and its 80 extra
// instructions per row made it 12 % slower); listing candidates by class and taking their
// expf once, converged, after the loop (19 fewer instructions per row, 2 % slower).
constexpr int kSoftmaxCtasPerSm = 5;
template <bool kSoftmax, bool kVec>
__global__ void __launch_bounds__(kClassifyMaxWarps * 32, kSoftmax ? kSoftmaxCtasPerSm : 3) classify_topk_reg_kernel(const ClassifyParams p) {
  __shared__ float cand_raw[kClassifyMaxWarps][kCandidates];  // e (softmax) or the raw value
  __shared__ int cand_idx[kClassifyMaxWarps][kCandidates];
  __shared__ int cand_count[kClassifyMaxWarps];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * kClassifyMaxWarps + warp;
  if (row >= p.rows) return;
  const float* g = p.logits + (size_t)row * p.classes;
  const int C = p.classes, k = p.k;
  const uint32_t flip = p.low_index_first ? 0xffffffffu : 0u;
  constexpr uint32_t kFull = 0xffffffffu;

  // element j of this lane is class class_of(j); classes past the end read as -inf.  kVec
  // (classes % 4 == 0, 16-byte aligned rows): 128-bit loads, four consecutive classes per lane
  auto class_of = [lane](int j) { return kVec ? ((j >> 2) * 32 + lane) * 4 + (j & 3) : j * 32 + lane; };
  float x[kRegPerLane];
  if (kVec) {
    // group q of a lane is in range for every lane once classes >= (q + 1) * 128: rows of 897+
    // classes (MobileNetV2's 1000) need the check — and the -inf filler — for the last group only
    auto load_group = [&](int q, bool checked) {
      const int i = (q * 32 + lane) * 4;
      const float4 f = !checked || i < C ? __ldg(reinterpret_cast<const float4*>(g + i))
                                         : make_float4(-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F);
      x[4 * q] = f.x, x[4 * q + 1] = f.y, x[4 * q + 2] = f.z, x[4 * q + 3] = f.w;
    };
    if (C > (kRegPerLane / 4 - 1) * 128) {  // uniform: two straight-line copies of the loads
#pragma unroll
      for (int q = 0; q < kRegPerLane / 4; ++q) load_group(q, q == kRegPerLane / 4 - 1);
    } else {
#pragma unroll
      for (int q = 0; q < kRegPerLane / 4; ++q) load_group(q, true);
    }
  } else {
#pragma unroll
    for (int j = 0; j < kRegPerLane; ++j) x[j] = class_of(j) < C ? __ldg(g + class_of(j)) : -CUDART_INF_F;
  }

  float total = 0.f;
  float m = 0.f;
  bool slow = k > 32;
  bool has_nan = false;
  int found = 0;
  if (kSoftmax) {
    // Row maximum, then the sum of exp(s - max).  Only the k probabilities that are reported need
    // numpy's exp to the last bits; the sum takes the fast exponential (FMUL + MUFU.EX2, relative
    // error ~3e-7 on the terms that matter), which moves the reported values by as much — the
    // tolerance of this row of SURVEY 8c is 1e-5.  Which classes win is decided on the logits:
    // exp and the division are monotone, so every winner's logit is >= the k-th largest lane
    // maximum; two different logits can round to one probability only if they are neighbours,
    // so everything within 1e-5 below that bound is kept as a candidate, gets the accurate expf
    // and the division, and the order among the candidates — ties included — is settled on those
    // probabilities exactly as argsort(probs)[::-1] does.
    m = x[0];
#pragma unroll
    for (int j = 1; j < kRegPerLane; ++j) m = fmaxf(m, x[j]);  // a NaN is skipped here ...
    const float lane_top = m;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) m = fmaxf(m, __shfl_xor_sync(kFull, m, d));
    // exp(x - m) = 2^(x * log2(e) - m * log2(e)): one FFMA + one MUFU.EX2 per element
    constexpr float kLog2e = 1.4426950408889634f;
    const float shift = __fmul_rn(-m, kLog2e);
    // (the packed f32x2 forms, FFMA2 / FADD2: the same roundings, half the issue slots)
    float part[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int j = 0; j < kRegPerLane; j += 2) {
      float t0, t1, e0, e1;
      fma_f32x2(t0, t1, x[j], x[j + 1], kLog2e, shift);
      asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e0) : "f"(t0));  // ... and surfaces here
      asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e1) : "f"(t1));
      add_f32x2(part[j & 2], part[(j & 2) + 1], e0, e1);
    }
    // (an xor butterfly adds the same pairs in every lane: all lanes hold the same bits)
    total = __fadd_rn(__fadd_rn(part[0], part[1]), __fadd_rn(part[2], part[3]));
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) total = __fadd_rn(total, __shfl_xor_sync(kFull, total, d));
    // NaN logits, an infinite maximum or a row of -inf make the sum NaN, and then every
    // probability is NaN (numpy gives the same) -> plain passes
    slow |= !(total == total);
    float bound = 0.f;
    if (!slow) {
      // (the sum is not NaN, so no logit is: every lane maximum is a number or the padding's -inf)
      bound = from_order_bits(kth_largest_lane_key(order_bits_no_nan(lane_top), lane, k));
      // where the k-th probability would be denormal the neighbour argument fails -> plain passes
      slow = !(__fsub_rn(bound, m) >= -69.f);
    }
    if (!slow) {
      if (lane == 0) cand_count[warp] = 0;
      __syncwarp();
      // which of this lane's elements lie within 1e-5 of the bound (2e-5 here: the subtraction
      // rounds): one compare and one predicated bit per element, no branches
      const float thr = __fsub_rn(bound, 2e-5f);
      uint32_t hits = 0;
#pragma unroll
      for (int j = 0; j < kRegPerLane; ++j)  // padding is -inf
        asm("{ .reg .pred p; setp.ge.f32 p, %1, %2; @p or.b32 %0, %0, %3; }" : "+r"(hits) : "f"(x[j]), "f"(thr), "r"(1u << j));
      // the few hits are read back through L1 (a register array cannot be indexed by a variable)
      while (hits) {
        const int j = __ffs(hits) - 1;
        hits &= hits - 1;
        const int cls = class_of(j);
        const int pos = atomicAdd(&cand_count[warp], 1);
        if (pos < kCandidates) cand_raw[warp][pos] = expf(__fsub_rn(__ldg(g + cls), m)), cand_idx[warp][pos] = cls;
      }
      __syncwarp();
      found = cand_count[warp];
      slow = found > kCandidates;
    }
    if (slow) {  // what the plain passes below work on: numpy's e, and their sum
#pragma unroll
      for (int j = 0; j < kRegPerLane; ++j) x[j] = expf(__fsub_rn(x[j], m));
      float tree[kRegPerLane];
#pragma unroll
      for (int j = 0; j < kRegPerLane; ++j) tree[j] = x[j];
#pragma unroll
      for (int w = kRegPerLane / 2; w > 0; w >>= 1)
#pragma unroll
        for (int j = 0; j < w; ++j) tree[j] = __fadd_rn(tree[j], tree[j + w]);
      double full = (double)tree[0];
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) full += __shfl_xor_sync(kFull, full, d);
      total = (float)full;
    }
  } else {
    uint32_t u[kRegPerLane];  // ordered bits of the raw values
#pragma unroll
    for (int j = 0; j < kRegPerLane; ++j) u[j] = class_of(j) < C ? order_bits(x[j]) : 0u;
    uint32_t lane_max = u[0];
#pragma unroll
    for (int j = 1; j < kRegPerLane; ++j) lane_max = max(lane_max, u[j]);
    has_nan = warp_max_u32(lane_max) == 0xffffffffu;  // NaN orders above everything
    uint32_t bound = 0;
    if (!slow) {
      bound = kth_largest_lane_key(lane_max, lane, k);
      if (lane == 0) cand_count[warp] = 0;
      __syncwarp();
#pragma unroll
      for (int j = 0; j < kRegPerLane; ++j) {
        if (u[j] >= bound) {  // padding holds 0 < bound
          const int pos = atomicAdd(&cand_count[warp], 1);
          if (pos < kCandidates) cand_raw[warp][pos] = x[j], cand_idx[warp][pos] = class_of(j);
        }
      }
      __syncwarp();
      found = cand_count[warp];
      slow = found > kCandidates;
    }
  }

  const size_t out = (size_t)row * k;
  if (!slow && found <= 32) {
    // The usual case, k..2k candidates: one per lane.  A round finds the largest ordered value
    // (and, among equal values, the id the tie rule prefers); the lane that holds it takes rank j
    // and retires; every ranked lane stores its own result at the end.  (found >= k: the maximum
    // is never the 0 of a retired or empty lane.)
    float v = 0.f;
    uint32_t vb = 0, id = 0;
    if (lane < found) {
      v = kSoftmax ? __fdiv_rn(cand_raw[warp][lane], total) : cand_raw[warp][lane];
      // a probability is finite and >= 0: its raw bits are ordered already
      vb = kSoftmax ? (__float_as_uint(v) | 0x80000000u) : order_bits(v);
      id = (uint32_t)cand_idx[warp][lane] ^ flip;
    }
    int rank = -1;
#pragma unroll 1
    for (int j = 0; j < k; ++j) {
      // two maxima, no branch: the value, then — among the lanes that hold it, nearly always
      // one — the tie rule's id (candidates are distinct classes: exactly one lane matches both)
      const uint32_t top = warp_max_u32(vb);
      const uint32_t best = warp_max_u32(vb == top ? id : 0u);
      if (vb == top && id == best) rank = j, vb = 0;
    }
    if (rank >= 0) {
      p.ids[out + rank] = (int)(id ^ flip);
      p.values[out + rank] = v;
    }
  } else if (!slow) {
    // lane L owns candidates L and L+32: value to report, then its 64-bit key
    float v0 = 0.f, v1 = 0.f;
    uint64_t c0 = 0ull, c1 = 0ull;
    if (lane < found) {
      v0 = kSoftmax ? __fdiv_rn(cand_raw[warp][lane], total) : cand_raw[warp][lane];
      c0 = ((uint64_t)order_bits(v0) << 32) | ((uint32_t)cand_idx[warp][lane] ^ flip);
    }
    if (lane + 32 < found) {
      v1 = kSoftmax ? __fdiv_rn(cand_raw[warp][lane + 32], total) : cand_raw[warp][lane + 32];
      c1 = ((uint64_t)order_bits(v1) << 32) | ((uint32_t)cand_idx[warp][lane + 32] ^ flip);
    }
    uint64_t prev = ~0ull;
    for (int j = 0; j < k; ++j) {
      const uint64_t a = c0 < prev ? c0 : 0ull, b = c1 < prev ? c1 : 0ull;
      const uint64_t mine = a > b ? a : b;
      const uint32_t hi = warp_max_u32((uint32_t)(mine >> 32));
      const uint32_t lo = warp_max_u32((uint32_t)(mine >> 32) == hi ? (uint32_t)mine : 0u);
      prev = ((uint64_t)hi << 32) | lo;
      if (mine == prev) {  // keys are unique: exactly one lane
        p.ids[out + j] = (int)(lo ^ flip);
        p.values[out + j] = mine == c0 ? v0 : v1;
      }
    }
  } else {
    if (kSoftmax) {
#pragma unroll
      for (int j = 0; j < kRegPerLane; ++j) {
        x[j] = __fdiv_rn(x[j], total);
        has_nan |= class_of(j) < C && isnan(__ldg(g + class_of(j)));
      }
      has_nan = __any_sync(kFull, has_nan);
    }
    uint64_t prev = ~0ull;
    for (int r = 0; r < k; ++r) {
      uint64_t best = 0;
      float best_val = 0.f;
#pragma unroll
      for (int j = 0; j < kRegPerLane; ++j) {
        const int i = class_of(j);
        const uint64_t key = i < C ? (((uint64_t)order_bits(x[j]) << 32) | ((uint32_t)i ^ flip)) : 0ull;
        if (key < prev && key > best) best = key, best_val = x[j];
      }
      prev = warp_max_u64(best);
      if (best == prev) {
        p.ids[out + r] = (int)((uint32_t)prev ^ flip);
        p.values[out + r] = best_val;
      }
    }
  }
  if (lane == 0 && p.status) p.status[row] = has_nan ? 1 : 0;
}

}  // namespace ia

extern "C" int ia_classify_topk(const float* logits_dev, int rows, int classes, int k, int softmax,
                                int low_index_first, int32_t* ids_dev, float* values_dev,
                                int32_t* status_dev, void* stream_) {
  using namespace ia;
  cudaStream_t stream = (cudaStream_t)stream_;
  IA_RETURN_IF(rows < 0 || classes < 1 || k < 1 || k > classes, IA_ERR_INVALID_ARGUMENT);
  if (rows == 0) return IA_OK;
  IA_RETURN_IF(!logits_dev || !ids_dev || !values_dev, IA_ERR_INVALID_ARGUMENT);
  ClassifyParams p;
  p.logits = logits_dev;
  p.rows = rows, p.classes = classes, p.k = k;
  p.softmax = softmax != 0, p.low_index_first = low_index_first != 0;
  p.warps = kClassifyMaxWarps;
  p.ids = ids_dev, p.values = values_dev, p.status = status_dev;
  if (classes <= kRegMaxClasses) {
    const int blocks = (rows + kClassifyMaxWarps - 1) / kClassifyMaxWarps;
    const bool vec = (classes & 3) == 0 && ((uintptr_t)logits_dev & 15) == 0;
    const dim3 grid(blocks), block(kClassifyMaxWarps * 32);
    if (p.softmax && vec) classify_topk_reg_kernel<true, true><<<grid, block, 0, stream>>>(p);
    else if (p.softmax) classify_topk_reg_kernel<true, false><<<grid, block, 0, stream>>>(p);
    else if (vec) classify_topk_reg_kernel<false, true><<<grid, block, 0, stream>>>(p);
    else classify_topk_reg_kernel<false, false><<<grid, block, 0, stream>>>(p);
    return launch_status();
  }
  const size_t row_bytes = (size_t)classes * sizeof(float);
  IA_RETURN_IF(row_bytes > kClassifySmemLimit, IA_ERR_UNSUPPORTED);
  int warps = (int)(kClassifySmemBudget / row_bytes);
  warps = warps < 1 ? 1 : (warps > kClassifyMaxWarps ? kClassifyMaxWarps : warps);
  const size_t smem = row_bytes * warps;
  if (smem > 48 * 1024) {  // once per (kernel, device), always to the same largest value
    cudaError_t e = allow_max_dynamic_smem(reinterpret_cast<const void*>(classify_topk_kernel<true>));
    if (e == cudaSuccess) e = allow_max_dynamic_smem(reinterpret_cast<const void*>(classify_topk_kernel<false>));
    if (e != cudaSuccess) return (int)e;
  }
  p.warps = warps;
  const int blocks = (rows + warps - 1) / warps;
  if (p.softmax) classify_topk_kernel<true><<<blocks, warps * 32, smem, stream>>>(p);
  else classify_topk_kernel<false><<<blocks, warps * 32, smem, stream>>>(p);
  return launch_status();
}
