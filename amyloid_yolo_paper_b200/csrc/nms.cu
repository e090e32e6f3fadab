// Cluster-merging non-maximum suppression, bit-exact with the reference's
// utils/utils.py:235-273 (non_max_suppression) + 202-232 (bbox_iou) + 53-59 (xywh2xyxy).
//
// The reference is not plain NMS: walking candidates in descending score, every
// surviving box i absorbs all later surviving boxes of the same class with
// IoU(i, j) > nms_thres (itself included) and its coordinates become the
// confidence-weighted mean of that cluster, summed with torch.sum.  To return the
// same bits, every fp32 operation below is an explicit round-to-nearest intrinsic in
// the reference's order (no FMA contraction) and the two sums follow ATen's CPU
// cascade order (SumKernel.cpp: multi_row_sum / row_sum with 8 fp32 lanes), see
// oracle/restate.py aten_sum_outer / aten_sum_inner.
//
// Four kernels per call (all tiles of the batch in each):
//   prepare : xywh->xyxy in place, confidence filter, score, class arg-max, sort by
//             (score desc, row asc) with an in-shared-memory bitonic network
//   mask    : upper-triangular same-class IoU>thr bit matrix, 64x64 blocks
//   sweep   : one warp per tile, serial pass over sorted candidates: survivors become
//             keepers, their mask row AND NOT(removed) is the cluster membership
//   cluster : one warp per keeper: ordered member list, ATen-order weighted sums
#include "ay_common.h"

#include <cstdlib>
#include <mutex>

namespace ay {
namespace {

constexpr int kMaxRows = 16384;   // per-tile candidate capacity of the shared-memory sort
constexpr int kSortThreads = 1024;

struct NmsLayout {      // per-call scratch carve-up (all offsets in bytes)
  size_t count, nkeep, box, conf, cconf, cls, orig, keep_pos, keep_off, memb, mask, total;
  int words;            // 64-bit mask words per row = ceil(n_rows/64)
};

__host__ __device__ inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

__host__ __device__ inline NmsLayout make_layout(int batch, int n) {
  NmsLayout L;
  L.words = (n + 63) / 64;
  size_t off = 0;
  L.count = off;    off = align256(off + sizeof(int) * (size_t)batch);
  L.nkeep = off;    off = align256(off + sizeof(int) * (size_t)batch);
  L.box = off;      off = align256(off + sizeof(float) * 4 * (size_t)batch * n);
  L.conf = off;     off = align256(off + sizeof(float) * (size_t)batch * n);
  L.cconf = off;    off = align256(off + sizeof(float) * (size_t)batch * n);
  L.cls = off;      off = align256(off + sizeof(int) * (size_t)batch * n);
  L.orig = off;     off = align256(off + sizeof(int) * (size_t)batch * n);
  L.keep_pos = off; off = align256(off + sizeof(int) * (size_t)batch * n);
  L.keep_off = off; off = align256(off + sizeof(int) * (size_t)batch * n);
  L.memb = off;     off = align256(off + sizeof(int) * (size_t)batch * n);
  L.mask = off;     off = align256(off + sizeof(unsigned long long) * (size_t)batch * n * L.words);
  L.total = off;
  return L;
}

// utils/utils.py:218-230, one rounding per operation
__device__ __forceinline__ float iou_xyxy(const float4 a, const float4 b) {
  const float ix1 = fmaxf(a.x, b.x), iy1 = fmaxf(a.y, b.y);
  const float ix2 = fminf(a.z, b.z), iy2 = fminf(a.w, b.w);
  const float iw = fmaxf(__fadd_rn(__fsub_rn(ix2, ix1), 1.0f), 0.0f);
  const float ih = fmaxf(__fadd_rn(__fsub_rn(iy2, iy1), 1.0f), 0.0f);
  const float inter = __fmul_rn(iw, ih);
  const float aa = __fmul_rn(__fadd_rn(__fsub_rn(a.z, a.x), 1.0f), __fadd_rn(__fsub_rn(a.w, a.y), 1.0f));
  const float ab = __fmul_rn(__fadd_rn(__fsub_rn(b.z, b.x), 1.0f), __fadd_rn(__fsub_rn(b.w, b.y), 1.0f));
  return __fdiv_rn(inter, __fadd_rn(__fsub_rn(__fadd_rn(aa, ab), inter), 1e-16f));
}

// (b.z - b.x + 1) * (b.w - b.y + 1), the roundings of utils/utils.py:227-228
__device__ __forceinline__ float box_area(const float4 b) {
  return __fmul_rn(__fadd_rn(__fsub_rn(b.z, b.x), 1.0f), __fadd_rn(__fsub_rn(b.w, b.y), 1.0f));
}

// iou_xyxy(a, b) > thr with the SAME truth value as the rounded division of the reference, but the division is only
// carried out when a fast quotient (error below 1e-6 relative) lands within 1e-4 of the threshold: disjoint boxes have
// iou = 0 (never above a threshold >= 0), and for every other pair the approximate quotient decides unless it is that
// close.  The mask kernel evaluates n^2 / 2 pairs per tile (57 M at 10 647 candidates): this is its inner loop.
__device__ __forceinline__ bool iou_above(const float4 a, const float aa, const float4 b, const float ab, const float thr) {
  const float iw = __fadd_rn(__fsub_rn(fminf(a.z, b.z), fmaxf(a.x, b.x)), 1.0f);
  const float ih = __fadd_rn(__fsub_rn(fminf(a.w, b.w), fmaxf(a.y, b.y)), 1.0f);
  // disjoint (or NaN) boxes: clamp(min=0) makes one factor of the intersection 0, the quotient is 0 or NaN, and neither is
  // above a threshold >= 0 (the caller takes the plain path for a negative threshold)
  if (!(iw > 0.0f) || !(ih > 0.0f)) return false;
  const float inter = __fmul_rn(iw, ih);
  const float denom = __fadd_rn(__fsub_rn(__fadd_rn(aa, ab), inter), 1e-16f);
  if (denom < 1e30f) {                                   // (__fdividef is only specified below 2^126; NaN fails the test)
    const float q = __fdividef(inter, denom);            // at most 2 ulp from the true quotient
    if (q > thr * 1.0001f + 1e-30f) return true;
    if (q < thr * 0.9999f) return false;
  }
  return __fdiv_rn(inter, denom) > thr;
}

// ------------------------------------------------------------------ prepare
// Two entry modes.  Dense (cand_idx == nullptr, the reference-shaped ay_nms): every row is rewritten xywh -> xyxy IN PLACE
// (utils/utils.py:244 mutates its argument) and filtered by confidence here.  Candidates (ay_nms_candidates): the head
// convolutions' epilogue has already compacted the rows with conf >= conf_thres into per-tile index lists, the dense
// tensor is only read at those rows and stays xywh; a tile whose list overflowed (count > cand_cap) is scanned densely
// instead, without the in-place rewrite.
__global__ void __launch_bounds__(kSortThreads, 1)
nms_prepare_kernel(float* __restrict__ pred, int n, int n_classes, float conf_thres, uint8_t* scratch, NmsLayout L,
                   const int32_t* __restrict__ cand_idx, const int32_t* __restrict__ cand_count, int cand_cap) {
  extern __shared__ unsigned long long keys[];   // npow2 entries
  __shared__ int s_count;
  const int b = blockIdx.x;
  const int cols = 5 + n_classes;
  float* tile = pred + (size_t)b * n * cols;
  if (threadIdx.x == 0) s_count = 0;
  __syncthreads();
  const bool in_place = cand_idx == nullptr;
  const int listed = in_place ? -1 : cand_count[b];
  const bool use_list = listed >= 0 && listed <= cand_cap;
  auto score_key = [&](const float* row, int r) {
    float best = row[5];
    for (int c = 1; c < n_classes; ++c) {
      const float v = row[5 + c];
      if (v > best) best = v;
    }
    const float score = __fmul_rn(row[4], best);
    // descending score, ascending row: order-preserving float->uint map, inverted
    unsigned int u = __float_as_uint(score);
    u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    return ((unsigned long long)(~u) << 32) | (unsigned int)r;
  };
  if (use_list) {
    for (int i = threadIdx.x; i < listed; i += blockDim.x) {
      const int r = cand_idx[(size_t)b * cand_cap + i];
      keys[i] = score_key(tile + (size_t)r * cols, r);
    }
    if (threadIdx.x == 0) s_count = listed;
  } else
  // pass 1: xyxy in place + candidate keys (unordered append; the sort makes it deterministic)
  for (int r = threadIdx.x; r < n; r += blockDim.x) {
    float* row = tile + (size_t)r * cols;
    if (in_place) {
      const float cx = row[0], cy = row[1], w = row[2], h = row[3];
      const float hw = __fdiv_rn(w, 2.0f), hh = __fdiv_rn(h, 2.0f);
      row[0] = __fsub_rn(cx, hw);
      row[1] = __fsub_rn(cy, hh);
      row[2] = __fadd_rn(cx, hw);
      row[3] = __fadd_rn(cy, hh);
    }
    if (row[4] >= conf_thres) {
      const int slot = atomicAdd(&s_count, 1);
      keys[slot] = score_key(row, r);
    }
  }
  __syncthreads();
  const int cnt = s_count;
  int np2 = 1;
  while (np2 < cnt) np2 <<= 1;
  for (int i = cnt + threadIdx.x; i < np2; i += blockDim.x) keys[i] = ~0ull;
  __syncthreads();
  for (int k = 2; k <= np2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < np2; i += blockDim.x) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const unsigned long long a = keys[i], c = keys[ixj];
          const bool up = ((i & k) == 0);
          if ((a > c) == up) { keys[i] = c; keys[ixj] = a; }
        }
      }
      __syncthreads();
    }
  }
  int* count = reinterpret_cast<int*>(scratch + L.count);
  float4* box = reinterpret_cast<float4*>(scratch + L.box) + (size_t)b * n;
  float* conf = reinterpret_cast<float*>(scratch + L.conf) + (size_t)b * n;
  float* cconf = reinterpret_cast<float*>(scratch + L.cconf) + (size_t)b * n;
  int* cls = reinterpret_cast<int*>(scratch + L.cls) + (size_t)b * n;
  int* orig = reinterpret_cast<int*>(scratch + L.orig) + (size_t)b * n;
  if (threadIdx.x == 0) count[b] = cnt;
  for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
    const int r = (int)(keys[i] & 0xFFFFFFFFull);
    const float* row = tile + (size_t)r * cols;
    if (in_place) {
      box[i] = make_float4(row[0], row[1], row[2], row[3]);
    } else {                                             // xywh2xyxy of this candidate only (utils/utils.py:53-59)
      const float cx = row[0], cy = row[1];
      const float hw = __fdiv_rn(row[2], 2.0f), hh = __fdiv_rn(row[3], 2.0f);
      box[i] = make_float4(__fsub_rn(cx, hw), __fsub_rn(cy, hh), __fadd_rn(cx, hw), __fadd_rn(cy, hh));
    }
    conf[i] = row[4];
    float best = row[5];
    int arg = 0;
    for (int c = 1; c < n_classes; ++c) {
      const float v = row[5 + c];
      if (v > best) { best = v; arg = c; }
    }
    cconf[i] = best;
    cls[i] = arg;
    orig[i] = r;
  }
}

// ------------------------------------------------------------------ mask
__global__ void __launch_bounds__(64)
nms_mask_kernel(int n, float nms_thres, uint8_t* scratch, NmsLayout L, int plain) {
  const int b = blockIdx.y;
  const int cnt = reinterpret_cast<const int*>(scratch + L.count)[b];
  const int nb = (cnt + 63) >> 6;
  const long long pairs = (long long)nb * (nb + 1) / 2;
  const float4* box = reinterpret_cast<const float4*>(scratch + L.box) + (size_t)b * n;
  const int* cls = reinterpret_cast<const int*>(scratch + L.cls) + (size_t)b * n;
  unsigned long long* mask = reinterpret_cast<unsigned long long*>(scratch + L.mask) + (size_t)b * n * L.words;
  __shared__ float4 cbox[64];
  __shared__ float carea[64];
  __shared__ int ccls[64];
  for (long long pi = blockIdx.x; pi < pairs; pi += gridDim.x) {
    // linear index over the upper triangle (rb <= cb), row-major
    int rb = (int)((2.0 * nb + 1.0 - sqrt((2.0 * nb + 1.0) * (2.0 * nb + 1.0) - 8.0 * (double)pi)) * 0.5);
    while ((long long)rb * nb - (long long)rb * (rb - 1) / 2 > pi) --rb;
    while ((long long)(rb + 1) * nb - (long long)(rb + 1) * rb / 2 <= pi) ++rb;
    const int cb = rb + (int)(pi - ((long long)rb * nb - (long long)rb * (rb - 1) / 2));
    const int j0 = cb * 64;
    __syncthreads();
    if (j0 + (int)threadIdx.x < cnt) {
      const float4 bj = box[j0 + threadIdx.x];
      cbox[threadIdx.x] = bj;
      carea[threadIdx.x] = box_area(bj);
      ccls[threadIdx.x] = cls[j0 + threadIdx.x];
    }
    __syncthreads();
    const int i = rb * 64 + threadIdx.x;
    if (i < cnt) {
      const float4 bi = box[i];
      const float ai = box_area(bi);
      const int ci = cls[i];
      const int lim = min(64, cnt - j0);
      unsigned long long bits = 0;
      if (nms_thres >= 0.0f && !plain) {
        for (int j = 0; j < lim; ++j)
          if (ccls[j] == ci && iou_above(bi, ai, cbox[j], carea[j], nms_thres)) bits |= (1ull << j);
      } else {
        for (int j = 0; j < lim; ++j)
          if (ccls[j] == ci && iou_xyxy(bi, cbox[j]) > nms_thres) bits |= (1ull << j);
      }
      mask[(size_t)i * L.words + cb] = bits;
    }
  }
}

// ------------------------------------------------------------------ sweep
__global__ void __launch_bounds__(32)
nms_sweep_kernel(int n, uint8_t* scratch, NmsLayout L) {
  __shared__ unsigned long long removed[kMaxRows / 64];
  const int b = blockIdx.x;
  const int lane = threadIdx.x;
  const int cnt = reinterpret_cast<const int*>(scratch + L.count)[b];
  const int nw = (cnt + 63) >> 6;
  unsigned long long* mask = reinterpret_cast<unsigned long long*>(scratch + L.mask) + (size_t)b * n * L.words;
  int* keep_pos = reinterpret_cast<int*>(scratch + L.keep_pos) + (size_t)b * n;
  int* keep_off = reinterpret_cast<int*>(scratch + L.keep_off) + (size_t)b * n;
  for (int w = lane; w < nw; w += 32) removed[w] = 0;
  __syncwarp();
  int nkeep = 0, running = 0;
  for (int i = 0; i < cnt; ++i) {
    const int wi = i >> 6;
    if ((removed[wi] >> (i & 63)) & 1ull) continue;     // warp-uniform (shared memory broadcast)
    int pc = 0;
    // all words of the row first (up to kMaxRows / 64 / 32 = 8 per lane, independent loads: one memory round trip per
    // surviving row instead of one per 32 words - the walk is serial, so this latency is the kernel's time)
    unsigned long long mw[kMaxRows / 64 / 32];
    unsigned long long* mrow = mask + (size_t)i * L.words;
#pragma unroll
    for (int k = 0; k < kMaxRows / 64 / 32; ++k) {
      const int w = wi + lane + 32 * k;
      mw[k] = w < nw ? mrow[w] : 0ull;
    }
#pragma unroll
    for (int k = 0; k < kMaxRows / 64 / 32; ++k) {
      const int w = wi + lane + 32 * k;
      if (w < nw) {
        unsigned long long m = mw[k];
        if (w == wi) m &= ~((1ull << (i & 63)) - 1ull);    // only rows >= i are still in play
        m &= ~removed[w];
        removed[w] |= m;
        mrow[w] = m;                                        // row i now holds its cluster members
        pc += __popcll(m);
      }
    }
    for (int o = 16; o > 0; o >>= 1) pc += __shfl_xor_sync(0xffffffffu, pc, o);
    if (lane == 0) { keep_pos[nkeep] = i; keep_off[nkeep] = running; }
    ++nkeep;
    running += pc;
    __syncwarp();
  }
  if (lane == 0) reinterpret_cast<int*>(scratch + L.nkeep)[b] = nkeep;
}

// ------------------------------------------------------------------ cluster merge
// multi_row_sum for one column, streaming: 4 cascade levels, blocks of 16 rows.
struct Cascade {
  float acc[4];
  int i;
  __device__ __forceinline__ void init() { acc[0] = acc[1] = acc[2] = acc[3] = 0.f; i = 0; }
  // add one element of a full 16-row block or of the tail (caller knows which via `size`)
  __device__ __forceinline__ void push(float v, int size) {
    acc[0] = __fadd_rn(acc[0], v);
    ++i;
    const int full = size & ~15;
    if (i <= full && (i & 15) == 0) {
#pragma unroll
      for (int j = 1; j < 4; ++j) {
        acc[j] = __fadd_rn(acc[j], acc[j - 1]);
        acc[j - 1] = 0.f;
        if ((i & (15 << (4 * j))) != 0) break;
      }
    }
  }
  __device__ __forceinline__ float finish() {
    float r = acc[0];
    r = __fadd_rn(r, acc[1]);
    r = __fadd_rn(r, acc[2]);
    r = __fadd_rn(r, acc[3]);
    return r;
  }
};

__global__ void __launch_bounds__(128)
nms_cluster_kernel(int n, uint8_t* scratch, NmsLayout L, float* __restrict__ out_rows, int32_t* __restrict__ out_keep,
                   int32_t* __restrict__ out_counts, int cap, unsigned int* err_flag) {
  const int b = blockIdx.y;
  const int lane = threadIdx.x & 31;
  const int warps_per_tile = gridDim.x * (blockDim.x >> 5);
  const int wid = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int cnt = reinterpret_cast<const int*>(scratch + L.count)[b];
  const int nkeep = reinterpret_cast<const int*>(scratch + L.nkeep)[b];
  const int nw = (cnt + 63) >> 6;
  const unsigned long long* mask =
      reinterpret_cast<const unsigned long long*>(scratch + L.mask) + (size_t)b * n * L.words;
  const float4* box = reinterpret_cast<const float4*>(scratch + L.box) + (size_t)b * n;
  const float* conf = reinterpret_cast<const float*>(scratch + L.conf) + (size_t)b * n;
  const float* cconf = reinterpret_cast<const float*>(scratch + L.cconf) + (size_t)b * n;
  const int* cls = reinterpret_cast<const int*>(scratch + L.cls) + (size_t)b * n;
  const int* orig = reinterpret_cast<const int*>(scratch + L.orig) + (size_t)b * n;
  const int* keep_pos = reinterpret_cast<const int*>(scratch + L.keep_pos) + (size_t)b * n;
  const int* keep_off = reinterpret_cast<const int*>(scratch + L.keep_off) + (size_t)b * n;
  int* memb_all = reinterpret_cast<int*>(scratch + L.memb) + (size_t)b * n;

  if (wid == 0 && lane == 0) {
    if (nkeep > cap) { out_counts[b] = -1; atomicExch(err_flag, 0xCA9A0001u); }
    else out_counts[b] = nkeep;
  }
  for (int k = wid; k < nkeep && k < cap; k += warps_per_tile) {
    const int i = keep_pos[k];
    int* memb = memb_all + keep_off[k];
    // ---- ordered member list from the bit row
    int size = 0;
    for (int w0 = i >> 6; w0 < nw; w0 += 32) {
      const int w = w0 + lane;
      unsigned long long m = (w < nw) ? mask[(size_t)i * L.words + w] : 0ull;
      const int pc = __popcll(m);
      int incl = pc;
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      int pos = size + incl - pc;
      while (m) {
        const int bit = __ffsll((long long)m) - 1;
        memb[pos++] = w * 64 + bit;
        m &= m - 1;
      }
      size += __shfl_sync(0xffffffffu, incl, 31);
    }
    __syncwarp();
    // ---- numerator: sum over members of conf * box, per coordinate (lanes 0..3)
    float num = 0.f;
    if (lane < 4) {
      Cascade cs;
      cs.init();
      for (int t = 0; t < size; ++t) {
        const int j = memb[t];
        const float4 bj = box[j];
        const float c = (lane == 0) ? bj.x : (lane == 1) ? bj.y : (lane == 2) ? bj.z : bj.w;
        cs.push(__fmul_rn(conf[j], c), size);
      }
      num = cs.finish();
    }
    // ---- denominator: torch.sum over the k contiguous weights (8 fp32 lanes)
    float den;
    if (size < 8) {
      // scalar path: row_sum over [size][1]: 4 interleaved partials
      float ps0 = 0.f, ps1 = 0.f, ps2 = 0.f, ps3 = 0.f;
      const int q = size >> 2;
      if (q) {   // q == 1: multi_row_sum over one row of 4 (0 + x)
        ps0 = __fadd_rn(0.f, conf[memb[0]]);
        ps1 = __fadd_rn(0.f, conf[memb[1]]);
        ps2 = __fadd_rn(0.f, conf[memb[2]]);
        ps3 = __fadd_rn(0.f, conf[memb[3]]);
        // cascade finish of multi_row_sum adds three zero levels
        ps0 = __fadd_rn(__fadd_rn(__fadd_rn(ps0, 0.f), 0.f), 0.f);
        ps1 = __fadd_rn(__fadd_rn(__fadd_rn(ps1, 0.f), 0.f), 0.f);
        ps2 = __fadd_rn(__fadd_rn(__fadd_rn(ps2, 0.f), 0.f), 0.f);
        ps3 = __fadd_rn(__fadd_rn(__fadd_rn(ps3, 0.f), 0.f), 0.f);
      }
      for (int t = 4 * q; t < size; ++t) ps0 = __fadd_rn(ps0, conf[memb[t]]);
      ps0 = __fadd_rn(ps0, ps1);
      ps0 = __fadd_rn(ps0, ps2);
      ps0 = __fadd_rn(ps0, ps3);
      den = ps0;
    } else {
      const int nv = size >> 3;       // full 8-lane vectors
      const int q = nv >> 2;          // rows of 32 floats
      Cascade cs;
      cs.init();
      for (int r = 0; r < q; ++r) cs.push(conf[memb[32 * r + lane]], q);
      float ps = q ? cs.finish() : 0.f;          // lane = 8*j + l  <->  ps[j][l]
      for (int v = 4 * q; v < nv; ++v) {
        const float x = (lane < 8) ? conf[memb[8 * v + lane]] : 0.f;
        if (lane < 8) ps = __fadd_rn(ps, x);
      }
      const float p1 = __shfl_sync(0xffffffffu, ps, (lane & 7) + 8);
      const float p2 = __shfl_sync(0xffffffffu, ps, (lane & 7) + 16);
      const float p3 = __shfl_sync(0xffffffffu, ps, (lane & 7) + 24);
      float vacc = __fadd_rn(__fadd_rn(__fadd_rn(ps, p1), p2), p3);   // valid on lanes 0..7
      float t = 0.f;
      for (int e = nv * 8; e < size; ++e) t = __fadd_rn(t, conf[memb[e]]);
      for (int l = 0; l < 8; ++l) t = __fadd_rn(t, __shfl_sync(0xffffffffu, vacc, l));
      den = t;
    }
    if (lane < 4) {
      float* orow = out_rows + ((size_t)b * cap + k) * 7;
      orow[lane] = __fdiv_rn(num, den);
      if (lane == 0) {
        orow[4] = conf[i];
        orow[5] = cconf[i];
        orow[6] = (float)cls[i];
        out_keep[(size_t)b * cap + k] = orig[i];
      }
    }
    __syncwarp();
  }
}

__global__ void bbox_iou_kernel(const float4* b1, int n1, const float4* b2, int n2, float* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n2) out[i] = iou_xyxy(b1[n1 == 1 ? 0 : i], b2[i]);
}

}  // namespace

size_t nms_scratch_bytes(int batch, int n_rows) { return make_layout(batch, n_rows).total; }

int launch_nms(float* pred, int batch, int n_rows, int n_classes, float conf_thres, float nms_thres, float* out_rows,
               int32_t* out_keep, int32_t* out_counts, int cap, void* scratch, size_t scratch_bytes,
               unsigned int* err_flag, cudaStream_t s, cudaEvent_t* ev, const int32_t* cand_idx,
               const int32_t* cand_count, int cand_cap) {
  if (n_rows > kMaxRows) {
    set_error("nms: %d rows per tile exceed the on-chip sort capacity of %d", n_rows, kMaxRows);
    return AY_EINVAL;
  }
  if (n_classes < 1 || batch < 1 || cap < 1) {
    set_error("nms: bad shape (batch=%d classes=%d cap=%d)", batch, n_classes, cap);
    return AY_EINVAL;
  }
  const NmsLayout L = make_layout(batch, n_rows);
  if (scratch_bytes < L.total) {
    set_error("nms: scratch too small");
    return AY_EINVAL;
  }
  int np2 = 1;
  while (np2 < n_rows) np2 <<= 1;
  const size_t sort_smem = sizeof(unsigned long long) * (size_t)np2;
  // per launch, like conv_gemm: the attribute is per DEVICE (a process may drive several GPUs) and setting it is cheap
  AY_CUDA(cudaFuncSetAttribute(nms_prepare_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
  uint8_t* sc = static_cast<uint8_t*>(scratch);
  // ev (ay_nms_timed only): five events bracketing the four kernels
  if (ev) AY_CUDA(cudaEventRecord(ev[0], s));
  nms_prepare_kernel<<<batch, kSortThreads, sort_smem, s>>>(pred, n_rows, n_classes, conf_thres, sc, L, cand_idx, cand_count,
                                                            cand_cap);
  AY_CUDA(cudaGetLastError());
  if (ev) AY_CUDA(cudaEventRecord(ev[1], s));
  // AY_NMS_PLAIN=1 (tests): every pair through the rounded division instead of the fast quotient with exact fallback
  nms_mask_kernel<<<dim3(96, batch), 64, 0, s>>>(n_rows, nms_thres, sc, L, getenv("AY_NMS_PLAIN") ? 1 : 0);
  AY_CUDA(cudaGetLastError());
  if (ev) AY_CUDA(cudaEventRecord(ev[2], s));
  nms_sweep_kernel<<<batch, 32, 0, s>>>(n_rows, sc, L);
  AY_CUDA(cudaGetLastError());
  if (ev) AY_CUDA(cudaEventRecord(ev[3], s));
  nms_cluster_kernel<<<dim3(16, batch), 128, 0, s>>>(n_rows, sc, L, out_rows, out_keep, out_counts, cap, err_flag);
  AY_CUDA(cudaGetLastError());
  if (ev) AY_CUDA(cudaEventRecord(ev[4], s));
  return AY_OK;
}

int launch_bbox_iou(const float* b1, int n1, const float* b2, int n2, float* out, cudaStream_t s) {
  if (n2 <= 0) return AY_OK;
  bbox_iou_kernel<<<(n2 + 127) / 128, 128, 0, s>>>(reinterpret_cast<const float4*>(b1), n1,
                                                   reinterpret_cast<const float4*>(b2), n2, out);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

// ------------------------------------------------------------------ ay_nms_static
// The allocation-free variant of ay_nms: the scratch of the four kernels lives in one file-scope device array instead
// of a cudaMalloc'd, context-owned buffer.  Every GPU that loads the library holds these kStaticScratchBytes until the
// process exits.  The first 256 bytes hold the overflow flag of nms_cluster_kernel, the NMS carve-up follows.
constexpr size_t kStaticFlagBytes = 256;
constexpr size_t kStaticScratchBytes = (size_t)32 << 20;   // 32 MiB: 2 tiles x 10 647 rows (416^2), 36 x 2 535 (tiny)
constexpr int kStaticMaxDevices = 64;

static __device__ __align__(256) unsigned char g_nms_static_scratch[kStaticScratchBytes];

// In host code the array's name is only a handle: the device address is looked up once per device, with that device
// current, and it is this pointer that the kernels receive.
static int static_scratch_address(uint8_t** out) {
  static uint8_t* addr[kStaticMaxDevices] = {};
  static std::mutex mu;
  int dev = -1;
  AY_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kStaticMaxDevices) {
    set_error("ay_nms_static: device ordinal %d outside [0, %d)", dev, kStaticMaxDevices);
    return AY_EINVAL;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void* ptr = nullptr;
    AY_CUDA(cudaGetSymbolAddress(&ptr, g_nms_static_scratch));
    if (!ptr) {
      set_error("ay_nms_static: cudaGetSymbolAddress returned a null address on device %d", dev);
      return AY_ECUDA;
    }
    addr[dev] = static_cast<uint8_t*>(ptr);
  }
  *out = addr[dev];
  return AY_OK;
}

}  // namespace ay

extern "C" size_t ay_nms_static_capacity(void) { return ay::kStaticScratchBytes - ay::kStaticFlagBytes; }

extern "C" int ay_nms_static(float* pred, int batch, int n_rows, int n_classes, float conf_thres, float nms_thres,
                             float* out_rows, int32_t* out_keep, int32_t* out_counts, int cap, void* stream) {
  using namespace ay;
  if (!pred || !out_rows || !out_keep || !out_counts || batch < 1 || n_rows < 1) {
    set_error("ay_nms_static: bad arguments");
    return AY_EINVAL;
  }
  if (n_rows > kMaxRows) {
    set_error("ay_nms_static: %d rows per tile exceed the on-chip sort capacity of %d", n_rows, kMaxRows);
    return AY_EINVAL;
  }
  // in 64-bit arithmetic before make_layout: the IoU bit matrix grows with batch * n_rows^2
  const unsigned long long mask_bytes = 8ull * (unsigned long long)batch * (unsigned long long)n_rows *
                                        (unsigned long long)((n_rows + 63) / 64);
  const size_t room = kStaticScratchBytes - kStaticFlagBytes;
  const size_t need = mask_bytes > room ? (size_t)mask_bytes : nms_scratch_bytes(batch, n_rows);
  if (need > room) {
    set_error("ay_nms_static: %d tiles x %d rows need %zu bytes of scratch, the static array holds %zu "
              "(use ay_nms, whose scratch is allocated to size)", batch, n_rows, need, room);
    return AY_ECAPACITY;
  }
  uint8_t* base = nullptr;
  int rc = static_scratch_address(&base);
  if (rc != AY_OK) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  unsigned int* flag = reinterpret_cast<unsigned int*>(base);
  // the array keeps the previous call's contents: only the flag has to start at zero, the kernels write every other
  // word of the carve-up before they read it
  AY_CUDA(cudaMemsetAsync(flag, 0, kStaticFlagBytes, s));
  return launch_nms(pred, batch, n_rows, n_classes, conf_thres, nms_thres, out_rows, out_keep, out_counts, cap,
                    base + kStaticFlagBytes, room, flag, s);
}
