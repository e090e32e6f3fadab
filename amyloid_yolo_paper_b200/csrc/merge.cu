// rescale_boxes (utils/utils.py:36-50) + mergeDetections (core.py:366-423, with
// combineIfOverlapping core.py:326-364), bit-exact including output row order.
//
// The reference keeps the live boxes in a python `set` of 7-tuples and repeatedly
// replaces overlapping same-class pairs (classes 0 and 1 only) by their integer
// hull until a pass makes no change.  Both the merged VALUES (chains of >=3 boxes
// shrink by one pixel per re-merge, so order matters) and the output ROW ORDER are
// functions of CPython's set iteration order.  This kernel therefore carries a
// faithful model of CPython 3.12's float/tuple hashing (Python/pyhash.c,
// Objects/tupleobject.c) and of the set's open-addressing table
// (Objects/setobject.c: 9 linear probes, perturb>>5, last-dummy reuse, growth to
// the next power of two above 4*used) — see oracle/restate.py merge_emulated, which
// is the executable specification.  The overlap test is the closed form of the
// reference's pixel-set enumeration, sentinels (1000000 / -1) included.
//
// One single-warp block per tile, lane 0 working: the algorithm is a serial fixed point over a handful of
// boxes, tiles are independent, and a lane per tile inside one warp would serialise 32 divergent tiles (measured:
// 10.3 ms for 256 tiles against 0.x ms here).  Tiles with at most kSmallN detections keep pool, hash tables and
// work lists in shared memory; larger ones use the global scratch.
#include "ay_common.h"

namespace ay {
namespace {

constexpr int kMergeCap = 512;                 // detections per tile the merge step accepts
constexpr int kPool = 2 * kMergeCap;           // originals + one new entry per merge
constexpr int kTable = 8 * kMergeCap;          // >= smallest power of two above 4*used
constexpr unsigned long long kP61 = (1ull << 61) - 1ull;

struct Entry {
  double v[7];
  long long hash;
};

struct Slot {
  long long hash;
  int idx;       // >=0 live entry, -1 never used, -2 dummy (deleted)
  int pad;
};

struct TileScratch {
  Entry pool[kPool];
  Slot table[2][kTable];
  int gone[kPool];
  int snap[kPool];
  unsigned char gflag[kPool];
};

__device__ __noinline__ long long hash_double(double v) {
  int e;
  double m = frexp(v, &e);
  long long sign = 1;
  if (m < 0) { sign = -1; m = -m; }
  unsigned long long x = 0;
  while (m != 0.0) {
    x = ((x << 28) & kP61) | (x >> 33);
    m *= 268435456.0;
    e -= 28;
    const unsigned long long y = (unsigned long long)m;
    m -= (double)y;
    x += y;
    if (x >= kP61) x -= kP61;
  }
  e = e >= 0 ? e % 61 : 60 - ((-1 - e) % 61);
  x = ((x << e) & kP61) | (x >> (61 - e));
  long long r = (long long)x * sign;
  if (r == -1) r = -2;
  return r;
}

__device__ __noinline__ long long hash_tuple7(const double* v) {
  const unsigned long long P1 = 11400714785074694791ull, P2 = 14029467366897019727ull, P5 = 2870177450012600261ull;
  unsigned long long acc = P5;
  for (int i = 0; i < 7; ++i) {
    const unsigned long long lane = (unsigned long long)hash_double(v[i]);
    acc += lane * P2;
    acc = (acc << 31) | (acc >> 33);
    acc *= P1;
  }
  acc += 7ull ^ (P5 ^ 3527539ull);
  if (acc == ~0ull) return 1546275796ll;
  return (long long)acc;
}

__device__ __forceinline__ bool same_tuple(const Entry& a, const Entry& b) {
  for (int i = 0; i < 7; ++i)
    if (!(a.v[i] == b.v[i])) return false;
  return true;
}

constexpr int kSmallN = 48;                    // detections per tile handled out of shared memory
constexpr int kSmallPool = 2 * kSmallN;
constexpr int kSmallSlots = 512;               // the set never holds more than n <= 48 live keys: resize() asks for the
                                               // power of two above 4 * used <= 192, i.e. at most 256 slots
constexpr size_t kSmallBytes = sizeof(Entry) * kSmallPool + 2 * sizeof(Slot) * kSmallSlots + 2 * sizeof(int) * kSmallPool + kSmallPool;

struct Work {            // where this tile's working set lives
  Entry* pool;
  Slot* table[2];
  int* gone;
  int* snap;
  unsigned char* gflag;   // gflag[i] != 0: the VALUE of pool[i] is in the reference's `gone` list
  int pool_cap;
};

struct PySet {
  Work* ts;
  int cur;        // which table buffer is live
  int mask, fill, used;

  __device__ void init(Work* t) {
    ts = t; cur = 0; mask = 7; fill = 0; used = 0;
    for (int i = 0; i < 8; ++i) ts->table[0][i].idx = -1;
  }
  // slot of an equal live key, or -1
  __device__ int find(const Entry& key) const {
    const Slot* tab = ts->table[cur];
    unsigned long long perturb = (unsigned long long)key.hash;
    unsigned int i = (unsigned int)(perturb & (unsigned long long)mask);
    while (true) {
      const int span = (i + 9u <= (unsigned int)mask) ? 10 : 1;
      for (int s = 0; s < span; ++s) {
        const Slot& e = tab[i + s];
        if (e.idx == -1) return -1;
        if (e.idx >= 0 && e.hash == key.hash && same_tuple(ts->pool[e.idx], key)) return (int)(i + s);
      }
      perturb >>= 5;
      i = (unsigned int)(((unsigned long long)i * 5ull + 1ull + perturb) & (unsigned long long)mask);
    }
  }
  __device__ void insert_clean(Slot* tab, int m, long long h, int idx) {
    unsigned long long perturb = (unsigned long long)h;
    unsigned int i = (unsigned int)(perturb & (unsigned long long)m);
    while (true) {
      const int span = (i + 9u <= (unsigned int)m) ? 10 : 1;
      for (int s = 0; s < span; ++s) {
        if (tab[i + s].idx == -1) { tab[i + s].hash = h; tab[i + s].idx = idx; return; }
      }
      perturb >>= 5;
      i = (unsigned int)(((unsigned long long)i * 5ull + 1ull + perturb) & (unsigned long long)m);
    }
  }
  __device__ void resize(int minused) {
    int size = 8;
    while (size <= minused) size <<= 1;
    Slot* oldt = ts->table[cur];
    Slot* newt = ts->table[cur ^ 1];
    for (int i = 0; i < size; ++i) newt[i].idx = -1;
    for (int i = 0; i <= mask; ++i)
      if (oldt[i].idx >= 0) insert_clean(newt, size - 1, oldt[i].hash, oldt[i].idx);
    cur ^= 1;
    mask = size - 1;
    fill = used;
  }
  // add pool entry `idx` (no-op when an equal key is present)
  __device__ void add(int idx) {
    const Entry& key = ts->pool[idx];
    Slot* tab = ts->table[cur];
    unsigned long long perturb = (unsigned long long)key.hash;
    unsigned int i = (unsigned int)(perturb & (unsigned long long)mask);
    int freeslot = -1, target = -1;
    while (target < 0) {
      const int span = (i + 9u <= (unsigned int)mask) ? 10 : 1;
      for (int s = 0; s < span; ++s) {
        Slot& e = tab[i + s];
        if (e.idx == -1) { target = (int)(i + s); break; }
        if (e.idx == -2) freeslot = (int)(i + s);
        else if (e.hash == key.hash && same_tuple(ts->pool[e.idx], key)) return;
      }
      if (target >= 0) break;
      perturb >>= 5;
      i = (unsigned int)(((unsigned long long)i * 5ull + 1ull + perturb) & (unsigned long long)mask);
    }
    if (freeslot >= 0) {
      tab[freeslot].hash = key.hash; tab[freeslot].idx = idx; ++used;
      return;
    }
    tab[target].hash = key.hash; tab[target].idx = idx;
    ++fill; ++used;
    if ((long long)fill * 5 >= (long long)mask * 3) resize(used > 50000 ? used * 2 : used * 4);
  }
  __device__ void remove_slot(int s) {
    ts->table[cur][s].idx = -2;
    --used;
  }
};

__device__ __forceinline__ long long trunc_ll(double v) { return (long long)v; }
__device__ __forceinline__ long long min3(long long a, long long b, long long c) { return min(a, min(b, c)); }
__device__ __forceinline__ long long max3(long long a, long long b, long long c) { return max(a, max(b, c)); }

__global__ void rescale_merge_kernel(const float* __restrict__ rows, const int32_t* __restrict__ counts, int batch,
                                     int cap, float pad_x_half, float pad_y_half, float unpad_w, float unpad_h,
                                     float orig_w, float orig_h, int merge, float* __restrict__ out_rows,
                                     int32_t* __restrict__ out_counts, TileScratch* scratch, unsigned int* err_flag) {
  const int b = blockIdx.x;
  if (b >= batch) return;
  const int lane = threadIdx.x;
  const int n = counts[b];
  const float* in = rows + (size_t)b * cap * 7;
  float* out = out_rows + (size_t)b * cap * 7;
  if (n <= 0) { if (lane == 0) out_counts[b] = n < 0 ? n : 0; return; }
  if (!merge || n > kMergeCap) {
    if (lane != 0) return;
    if (merge) { atomicExch(err_flag, 0xCA9A0002u); out_counts[b] = -1; return; }
    for (int r = 0; r < n; ++r) {
      out[r * 7 + 0] = __fmul_rn(__fdiv_rn(__fsub_rn(in[r * 7 + 0], pad_x_half), unpad_w), orig_w);
      out[r * 7 + 1] = __fmul_rn(__fdiv_rn(__fsub_rn(in[r * 7 + 1], pad_y_half), unpad_h), orig_h);
      out[r * 7 + 2] = __fmul_rn(__fdiv_rn(__fsub_rn(in[r * 7 + 2], pad_x_half), unpad_w), orig_w);
      out[r * 7 + 3] = __fmul_rn(__fdiv_rn(__fsub_rn(in[r * 7 + 3], pad_y_half), unpad_h), orig_h);
      out[r * 7 + 4] = in[r * 7 + 4];
      out[r * 7 + 5] = in[r * 7 + 5];
      out[r * 7 + 6] = in[r * 7 + 6];
    }
    out_counts[b] = n;
    return;
  }
  extern __shared__ __align__(16) unsigned char merge_smem[];
  Work work;
  if (n <= kSmallN) {
    // a merge removes two live keys and adds one: at most n - 1 merges, so the pool holds <= 2n - 1 entries and
    // `gone` <= 2n - 2 indices
    work.pool = reinterpret_cast<Entry*>(merge_smem);
    work.table[0] = reinterpret_cast<Slot*>(merge_smem + sizeof(Entry) * kSmallPool);
    work.table[1] = work.table[0] + kSmallSlots;
    work.gone = reinterpret_cast<int*>(work.table[1] + kSmallSlots);
    work.snap = work.gone + kSmallPool;
    work.gflag = reinterpret_cast<unsigned char*>(work.snap + kSmallPool);
    work.pool_cap = kSmallPool;
  } else {
    TileScratch* g = scratch + b;
    work.pool = g->pool; work.table[0] = g->table[0]; work.table[1] = g->table[1];
    work.gone = g->gone; work.snap = g->snap; work.gflag = g->gflag; work.pool_cap = kPool;
  }
  Work* ts = &work;
  PySet live;
  live.init(ts);
  // rescale + tuple hash of the n detections: independent, one lane each (the FP64 hash is most of a small tile's time)
  bool finite = true;
  for (int r = lane; r < n; r += 32) {
    Entry& e = ts->pool[r];
    e.v[0] = (double)__fmul_rn(__fdiv_rn(__fsub_rn(in[r * 7 + 0], pad_x_half), unpad_w), orig_w);
    e.v[1] = (double)__fmul_rn(__fdiv_rn(__fsub_rn(in[r * 7 + 1], pad_y_half), unpad_h), orig_h);
    e.v[2] = (double)__fmul_rn(__fdiv_rn(__fsub_rn(in[r * 7 + 2], pad_x_half), unpad_w), orig_w);
    e.v[3] = (double)__fmul_rn(__fdiv_rn(__fsub_rn(in[r * 7 + 3], pad_y_half), unpad_h), orig_h);
    e.v[4] = (double)in[r * 7 + 4];
    e.v[5] = (double)in[r * 7 + 5];
    e.v[6] = (double)in[r * 7 + 6];
    // A non-finite field (exp overflow in the decode, an inf activation) has no tuple hash here: frexp(inf) = inf and
    // NaN != 0 forever, so hash_double would never return.  The reference fails loudly as well (int(inf) raises
    // OverflowError, int(nan) ValueError in combineIfOverlapping, core.py:330-337): report it instead of hanging.
    bool ok = true;
    for (int c = 0; c < 7; ++c) ok = ok && isfinite(e.v[c]);
    if (ok) e.hash = hash_tuple7(e.v);
    else finite = false;
    ts->gflag[r] = 0;
  }
  finite = __all_sync(0xffffffffu, finite);
  if (!finite) {
    if (lane == 0) { atomicExch(err_flag, 0xCA9A0004u); out_counts[b] = -1; }
    return;
  }
  __syncwarp();
  if (lane != 0) return;              // the fixed point itself is serial (set iteration order)
  int npool = 0, ngone = 0;
  for (int r = 0; r < n; ++r) {
    live.add(npool);
    ++npool;   // duplicates stay in the pool but not in the set
  }
  bool changed = true;
  while (changed) {
    changed = false;
    int ns = 0;
    {
      const Slot* tab = ts->table[live.cur];
      for (int s = 0; s <= live.mask; ++s)
        if (tab[s].idx >= 0) ts->snap[ns++] = tab[s].idx;
    }
    // The reference skips a pair when either tuple equals (by value) a member of `gone`, a list that persists across
    // passes.  Live keys are value-unique, so a snapshot entry can only equal a gone one if it IS that entry or was
    // created later with the same value: one flag per pool entry, set on removal and checked against the list once at
    // creation, replaces the O(n) list scan per pair.
    for (int i = 0; i < ns; ++i) {
      for (int j = i + 1; j < ns; ++j) {
        if (ts->gflag[ts->snap[i]]) break;
        if (ts->gflag[ts->snap[j]]) continue;
        const Entry& A = ts->pool[ts->snap[i]];
        const Entry& B = ts->pool[ts->snap[j]];
        const double la = A.v[6], lb = B.v[6];
        if (!((la == 1.0 && lb == 1.0) || (la == 0.0 && lb == 0.0))) continue;
        const long long x1 = trunc_ll(A.v[0]), y1 = trunc_ll(A.v[1]);
        const long long w1 = trunc_ll(A.v[2] - A.v[0]), h1 = trunc_ll(A.v[3] - A.v[1]);
        const long long x2 = trunc_ll(B.v[0]), y2 = trunc_ll(B.v[1]);
        const long long w2 = trunc_ll(B.v[2] - B.v[0]), h2 = trunc_ll(B.v[3] - B.v[1]);
        if (w1 <= 0 || h1 <= 0 || w2 <= 0 || h2 <= 0) continue;
        if (!(max(x1, x2) < min(x1 + w1, x2 + w2) && max(y1, y2) < min(y1 + h1, y2 + h2))) continue;
        if (npool >= ts->pool_cap) { atomicExch(err_flag, 0xCA9A0003u); out_counts[b] = -1; return; }
        Entry& N = ts->pool[npool];
        N.v[0] = (double)min3(1000000, x1, x2);
        N.v[1] = (double)min3(1000000, y1, y2);
        N.v[2] = (double)max3(-1, x1 + w1 - 1, x2 + w2 - 1);
        N.v[3] = (double)max3(-1, y1 + h1 - 1, y2 + h2 - 1);
        N.v[4] = (B.v[4] < A.v[4]) ? B.v[4] : A.v[4];
        N.v[5] = (B.v[5] < A.v[5]) ? B.v[5] : A.v[5];
        N.v[6] = la;
        N.hash = hash_tuple7(N.v);
        if (live.find(N) >= 0) continue;
        {
          unsigned char f = 0;
          for (int g = 0; g < ngone && !f; ++g) f = same_tuple(ts->pool[ts->gone[g]], N) ? 1 : 0;
          ts->gflag[npool] = f;
        }
        live.add(npool);
        ++npool;
        live.remove_slot(live.find(A));
        live.remove_slot(live.find(B));
        ts->gone[ngone++] = ts->snap[i];
        ts->gone[ngone++] = ts->snap[j];
        ts->gflag[ts->snap[i]] = 1;
        ts->gflag[ts->snap[j]] = 1;
        changed = true;
      }
    }
  }
  int k = 0;
  const Slot* tab = ts->table[live.cur];
  for (int s = 0; s <= live.mask; ++s) {
    if (tab[s].idx < 0) continue;
    const Entry& e = ts->pool[tab[s].idx];
    for (int c = 0; c < 7; ++c) out[k * 7 + c] = (float)e.v[c];
    ++k;
  }
  out_counts[b] = k;
}

}  // namespace

size_t merge_scratch_bytes(int batch, int cap) { (void)cap; return sizeof(TileScratch) * (size_t)batch; }

int launch_rescale_merge(const float* rows, const int32_t* counts, int batch, int cap, float pad_x_half,
                         float pad_y_half, float unpad_w, float unpad_h, float orig_w, float orig_h, int merge,
                         float* out_rows, int32_t* out_counts, void* scratch, size_t scratch_bytes,
                         unsigned int* err_flag, cudaStream_t s) {
  if (scratch_bytes < merge_scratch_bytes(batch, cap)) {
    set_error("rescale_merge: scratch too small");
    return AY_EINVAL;
  }
  rescale_merge_kernel<<<batch, 32, kSmallBytes, s>>>(rows, counts, batch, cap, pad_x_half, pad_y_half, unpad_w,
                                                        unpad_h, orig_w, orig_h, merge, out_rows, out_counts,
                                                        static_cast<TileScratch*>(scratch), err_flag);
  AY_CUDA(cudaGetLastError());
  return AY_OK;
}

}  // namespace ay
