// csrc/kdtree_build.cu -- GPU construction of a photon map: left-balanced kd-tree in heap layout (K5).
//
// Replaces `new R3Kdtree<Photon*>(photons, offset)` (src/photonmap.cpp:365,372), i.e. R3Kdtree::InsertPoints /
// PartitionPoints (src/R3Shapes/R3Kdtree.cpp:1624-1671, 1552-1620): a single-threaded recursive build with
// randomised quickselect and 31-photon leaf buckets (15 s for 1e7 photons on the CPU).
//
// Here: level-synchronous construction of Jensen's left-balanced tree from three PRESORTED lists (the idea of Brown's
// O(k n log n) build): the photon ids are sorted once along x, once along y and once along z (three 32-bit radix sorts);
// after that no level sorts anything.  The lists are kept in IN-ORDER layout: a subtree owns one contiguous range
// [begin, begin + size) of all three lists, its pivot sits at begin + (size of the left child subtree), the children own
// the ranges on either side (left-balanced => every size is closed-form in n).  For level l = 0..L-1:
//   choose    thread per subtree: the TIGHT bounding box of its photons is read off the two ends of its range in the three
//             lists (no reduction, no atomics); split axis = longest axis of that box; pivot = the entry at the median
//             position of that axis' list -> node s of the heap.  The reference splits on the longest axis of the
//             inherited node box (R3Kdtree.cpp:1640, R3Box::LongestAxis R3Box.cpp:258-274); for photons lying on planes
//             that rule keeps cutting along the plane normal, where every coordinate is equal and nothing is separated.
//             The choice of axis does not change any query result, only how much of the tree a query has to visit.
//   partition one segmented scan over the three lists of packed (left, right) counters whose OUTPUT ITERATOR moves every
//             entry to its place (single pass, cub's decoupled look-back).  A list entry is 16 bytes {photon id,
//             ordered float x, y, z}, so which side of the pivot an entry falls on is decided from the entry itself:
//             the lists are ordered by (coordinate, id), hence side = compare((entry's split-axis key, id), (pivot's)) --
//             no per-photon state, every access is coalesced.  Every range of every list is stably partitioned into left
//             entries, the pivot, right entries: the lists of the children are sorted again, in place in the in-order
//             layout; a per-position array says which subtree a position belongs to (the ranges coincide in the lists).
// Per level that is ~110 bytes of coalesced traffic per photon and three launches, against a full 64-bit radix sort of all unplaced
// photons plus a bounding-box pass of atomics in the first version of this file (10^7 photons: 41 ms -> see DESIGN.md);
// GI_KD_SORT_BUILD=1 selects the old sort-per-level build for A/B measurements.
// Splits are float32 (monotone rounding of the FP64 coordinate); the search widens them by one ulp, distances stay
// FP64, so the k-nearest SET is exact (see gather.cu).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <thrust/iterator/counting_iterator.h>
#include <thrust/iterator/transform_iterator.h>
#include <stdlib.h>
#include <iterator>
#include <cuda_runtime.h>
#include <float.h>
#include <math.h>
#include <string.h>
#include <stdint.h>
#include "context.h"
#include "kdtree_build.h"

namespace gi {

__host__ __device__ __forceinline__ int level_of(uint64_t s) {
#ifdef __CUDA_ARCH__
  return 63 - __clzll((long long)(s + 1));
#else
  int l = 0; while (((s + 1) >> (l + 1)) != 0) l++; return l;
#endif
}

// number of nodes in the subtree rooted at heap slot s of a left-balanced tree with n nodes and L levels
__host__ __device__ __forceinline__ uint64_t subtree_size(uint64_t s, uint64_t n, int L) {
  if (s >= n) return 0;
  const int h = L - 1 - level_of(s);
  const uint64_t inner = ((uint64_t)1 << h) - 1;
  const uint64_t first_last = ((s + 1) << h) - 1;
  uint64_t last = 0;
  if (n > first_last) { last = n - first_last; if (last > ((uint64_t)1 << h)) last = (uint64_t)1 << h; }
  return inner + last;
}

// array position where subtree s starts once the array is sorted by (tag) at the level of s
__host__ __device__ __forceinline__ uint64_t segment_begin(uint64_t s, uint64_t n, int L) {
  const int l = level_of(s);
  const int h = L - 1 - l;
  const uint64_t above = ((uint64_t)1 << l) - 1;
  const uint64_t k = s - above;
  const uint64_t full_above_last = ((uint64_t)1 << (L - 1)) - 1;
  const uint64_t total_last = n > full_above_last ? n - full_above_last : 0;
  uint64_t last = k << h; if (last > total_last) last = total_last;
  return above + k * (((uint64_t)1 << h) - 1) + last;
}

__device__ __forceinline__ uint32_t float_to_ordered(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ordered_to_float(uint32_t o) {
  return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}

__device__ __forceinline__ int longest_axis(const float *b) {          // R3Box::LongestAxis, R3Box.cpp:258-274
  const float dx = b[3] - b[0], dy = b[4] - b[1], dz = b[5] - b[2];
  if (dx > dy) return dx > dz ? 0 : 2;
  return dy > dz ? 1 : 2;
}

__global__ void init_kernel(uint64_t n, uint64_t *keys, uint32_t *vals) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { keys[i] = 0; vals[i] = (uint32_t)i; }
}

// bounds[s] = {min x,y,z ; max x,y,z} as ordered uints, for the subtrees s of this level
__global__ void reset_bounds_kernel(uint64_t n, int level, uint32_t *bounds) {
  const uint64_t s = (((uint64_t)1 << level) - 1) + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n || s >= ((uint64_t)2 << level) - 1) return;
  for (int k = 0; k < 3; k++) { bounds[6 * s + k] = 0xffffffffu; bounds[6 * s + 3 + k] = 0u; }
}

__global__ void bounds_kernel(const GiPhoton *__restrict__ ph, uint64_t first, uint64_t n, int level,
                              const uint64_t *__restrict__ keys, const uint32_t *__restrict__ vals, uint32_t *bounds) {
  const uint64_t p = first + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = p < n;
  uint64_t tag = live ? (keys[p] >> 32) : ~(uint64_t)0;
  const bool active = live && level_of(tag) == level;
  uint32_t lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u};
  if (active) {
    const GiPhoton &q = ph[vals[p]];
    for (int k = 0; k < 3; k++) lo[k] = hi[k] = float_to_ordered((float)q.position[k]);
  }
  // whole warp inside one subtree (the common case on the upper levels): reduce, one lane does the atomics
  const uint64_t tag0 = __shfl_sync(0xffffffffu, tag, 0);
  if (__all_sync(0xffffffffu, active && tag == tag0)) {
    for (int k = 0; k < 3; k++)
      for (int o = 16; o > 0; o >>= 1) {
        lo[k] = min(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
        hi[k] = max(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
      }
    if ((threadIdx.x & 31) == 0)
      for (int k = 0; k < 3; k++) { atomicMin(&bounds[6 * tag + k], lo[k]); atomicMax(&bounds[6 * tag + 3 + k], hi[k]); }
  } else if (active) {
    for (int k = 0; k < 3; k++) { atomicMin(&bounds[6 * tag + k], lo[k]); atomicMax(&bounds[6 * tag + 3 + k], hi[k]); }
  }
}

__global__ void axis_kernel(uint64_t n, int level, const uint32_t *__restrict__ bounds, uint8_t *axis) {
  const uint64_t s = (((uint64_t)1 << level) - 1) + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n || s >= ((uint64_t)2 << level) - 1) return;
  float b[6];
  for (int k = 0; k < 6; k++) b[k] = ordered_to_float(bounds[6 * s + k]);
  axis[s] = (uint8_t)longest_axis(b);
}

// key = tag << 32 | ordered float coordinate on the split axis of subtree `tag`
__global__ void make_keys_kernel(const GiPhoton *__restrict__ ph, uint64_t first, uint64_t n, int level,
                                 const uint8_t *__restrict__ axis, uint64_t *keys, const uint32_t *__restrict__ vals) {
  const uint64_t p = first + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  const uint64_t tag = keys[p] >> 32;
  if (level_of(tag) < level) { keys[p] = tag << 32; return; }          // placed on the previous level
  const float c = (float)ph[vals[p]].position[axis[tag]];
  keys[p] = (tag << 32) | float_to_ordered(c);
}

__global__ void assign_kernel(uint64_t first, uint64_t n, int L, int level, uint64_t *keys,
                              const uint32_t *__restrict__ vals, uint32_t *heap_photon, float *split) {
  const uint64_t p = first + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  const uint64_t key = keys[p];
  const uint64_t s = key >> 32;
  if (level_of(s) < level) return;
  const uint64_t pivot = segment_begin(s, n, L) + subtree_size(2 * s + 1, n, L);
  if (p < pivot) keys[p] = (2 * s + 1) << 32;
  else if (p > pivot) keys[p] = (2 * s + 2) << 32;
  else { heap_photon[s] = vals[p]; split[s] = ordered_to_float((uint32_t)key); }
}

__global__ void finalize_kernel(const GiPhoton *__restrict__ ph, uint64_t n, const uint32_t *__restrict__ heap_photon,
                                const float *__restrict__ split, const uint8_t *__restrict__ axis, double *px, double *py,
                                double *pz, uint2 *node, uint2 *shade, uint32_t *orig, const double *__restrict__ lut, double2 *srec) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t src = heap_photon[i];
  const uint4 *raw = (const uint4 *)(ph + src);                        // 32-byte record: 2 x LDG.128
  const uint4 a = raw[0], b = raw[1];
  px[i] = __hiloint2double(a.y, a.x); py[i] = __hiloint2double(a.w, a.z); pz[i] = __hiloint2double(b.y, b.x);
  shade[i] = make_uint2(b.z, b.w & 0xffffu);
  orig[i] = src;
  node[i] = make_uint2(__float_as_uint(split[i]), axis[i]);
  const double *dir = lut + 3 * (size_t)(b.w & 0xffffu);
  srec[2 * i] = make_double2(dir[0], dir[1]);
  srec[2 * i + 1] = make_double2(dir[2], __hiloint2double((int)(b.w & 0xffffu), (int)b.z));
}

// Per group of row g (photon_dev.h): the position block gpos (x[32], y[32], z[32], +inf where the group has no photon)
// and the tight FP64 box of the subtree below its root = min / max over the group's photons, over the boxes of its 32
// child groups (its own cbox block, filled by the previous launch for row g + 1) and, in a six-level bottom row, over
// its level-6 photons.  The box goes into the parent group's cbox block.  One warp per group.
__global__ void group_kernel(PhotonMapDev m, int g, double *gpos, double *cbox) {
  const uint64_t w = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const uint64_t first = ((uint64_t)1 << m.glevel[g]) - 1;
  const uint64_t r = first + w;
  if (r >= m.n || w >= ((uint64_t)1 << m.glevel[g])) return;
  const int h = m.groups == 1 ? (m.hb < GI_GROUP_LEVELS ? m.hb : GI_GROUP_LEVELS) : (g == 0 ? m.h0 : GI_GROUP_LEVELS);
  const double INF = __longlong_as_double(0x7ff0000000000000ll);
  double lo[3] = {INF, INF, INF}, hi[3] = {-INF, -INF, -INF};
  const int j = 31 - __clz(lane + 1);
  const uint64_t slot = (r << j) + (uint64_t)lane;
  double p[3] = {INF, INF, INF};
  if (lane < (1 << h) - 1 && slot < m.n) { p[0] = m.px[slot]; p[1] = m.py[slot]; p[2] = m.pz[slot]; for (int a = 0; a < 3; a++) lo[a] = hi[a] = p[a]; }
  const size_t gi = (size_t)m.goff[g] + w;
  for (int a = 0; a < 3; a++) gpos[gi * 96 + 32 * a + lane] = p[a];
  if (g + 1 < m.groups) {
    const double *b = cbox + gi * 192 + lane;                                  // boxes of my children (inf where none)
    for (int a = 0; a < 3; a++) { lo[a] = fmin(lo[a], b[64 * a]); hi[a] = fmax(hi[a], b[64 * a + 32]); }
  } else if (m.hb > GI_GROUP_LEVELS) {                                         // bottom row, six levels deep
    const uint64_t c = ((r + 1) << h) - 1 + (uint64_t)lane;
    if (c < m.n) {
      lo[0] = fmin(lo[0], m.px[c]); hi[0] = fmax(hi[0], m.px[c]); lo[1] = fmin(lo[1], m.py[c]); hi[1] = fmax(hi[1], m.py[c]);
      lo[2] = fmin(lo[2], m.pz[c]); hi[2] = fmax(hi[2], m.pz[c]);
    }
  }
  for (int a = 0; a < 3; a++)
    for (int o = 16; o > 0; o >>= 1) {
      lo[a] = fmin(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
      hi[a] = fmax(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
    }
  if (lane == 0 && g > 0) {
    const int hp = g == 1 ? m.h0 : GI_GROUP_LEVELS;                            // depth of the parent row
    const uint64_t pr = ((r + 1) >> hp) - 1, child = (r + 1) & (((uint64_t)1 << hp) - 1);
    double *b = cbox + ((size_t)m.goff[g - 1] + (pr - (((uint64_t)1 << m.glevel[g - 1]) - 1))) * 192 + child;
    for (int a = 0; a < 3; a++) { b[64 * a] = lo[a]; b[64 * a + 32] = hi[a]; }
  }
}
// level 6 of a six-level bottom row: runs of 32 consecutive heap slots -> position blocks x6_base + j
__global__ void level6_kernel(PhotonMapDev m, double *gpos) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (uint64_t)m.x6_blocks * 32) return;
  const uint64_t slot = (((uint64_t)1 << (m.levels - 1)) - 1) + t;
  const double INF = __longlong_as_double(0x7ff0000000000000ll);
  const bool v = slot < m.n;
  double *o = gpos + ((size_t)m.x6_base + (t >> 5)) * 96 + (t & 31);
  o[0] = v ? m.px[slot] : INF; o[32] = v ? m.py[slot] : INF; o[64] = v ? m.pz[slot] : INF;
}
// cbox starts as "no child anywhere": lo = +inf, hi = -inf
__global__ void cbox_init_kernel(double *cbox, uint64_t n_doubles) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_doubles) cbox[i] = ((i >> 5) & 1) ? -__longlong_as_double(0x7ff0000000000000ll) : __longlong_as_double(0x7ff0000000000000ll);
}

// ------------------------------------------------------------------------------------------------ presorted build
#define KD_DONE 0xffffffffu          /* pseg: the position holds a placed pivot */

__global__ void presort_keys_kernel(const GiPhoton *__restrict__ ph, uint64_t n, int a, uint32_t *keys, uint32_t *ids) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { keys[i] = float_to_ordered((float)ph[i].position[a]); ids[i] = (uint32_t)i; }
}
// list entry = {photon id, ordered float x, y, z}
__global__ void entries_kernel(const GiPhoton *__restrict__ ph, uint64_t n, const uint32_t *__restrict__ sorted_ids, uint4 *list, uint32_t *pseg) {
  const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const uint32_t id = sorted_ids[j];
  const GiPhoton &q = ph[id];
  list[j] = make_uint4(id, float_to_ordered((float)q.position[0]), float_to_ordered((float)q.position[1]), float_to_ordered((float)q.position[2]));
  if (pseg) pseg[j] = 0u;
}
__device__ __forceinline__ uint32_t entry_key(const uint4 &e, int a) { return a == 0 ? e.y : (a == 1 ? e.z : e.w); }

struct Lists { const uint4 *l[3]; };
struct ListsOut { uint4 *l[3]; };
struct Pivots {                      // per subtree (heap index) of the level being built
  const uint32_t *segbegin; const uint8_t *axis; const uint32_t *key; const uint32_t *id;
};

// thread per subtree of the level: tight box from the ends of its range in the three lists, axis, pivot, children's ranges
__global__ void choose_kernel(uint64_t n, int L, int level, Lists in, uint32_t *segbegin, uint32_t *heap_photon, uint32_t *pivot_key,
                              float *split, uint8_t *axis, uint32_t *root_bounds) {
  const uint64_t s = (((uint64_t)1 << level) - 1) + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n || s >= ((uint64_t)2 << level) - 1) return;
  const uint32_t begin = level == 0 ? 0u : segbegin[s];
  const uint64_t sz = subtree_size(s, n, L);
  uint32_t ob[6]; float b[6];
  for (int a = 0; a < 3; a++) {
    ob[a] = entry_key(in.l[a][begin], a); ob[3 + a] = entry_key(in.l[a][begin + sz - 1], a);
    b[a] = ordered_to_float(ob[a]); b[3 + a] = ordered_to_float(ob[3 + a]);
  }
  const int ax = longest_axis(b);
  const uint64_t m = subtree_size(2 * s + 1, n, L);
  const uint4 pv = in.l[ax][begin + m];
  heap_photon[s] = pv.x; pivot_key[s] = entry_key(pv, ax); split[s] = ordered_to_float(entry_key(pv, ax)); axis[s] = (uint8_t)ax;
  if (level == 0) segbegin[0] = 0u;
  if (2 * s + 1 < n) segbegin[2 * s + 1] = begin;
  if (2 * s + 2 < n) segbegin[2 * s + 2] = begin + (uint32_t)m + 1u;
  if (s == 0) for (int k = 0; k < 6; k++) root_bounds[k] = ob[k];
}

// 0: left of the pivot, 1: right, 2: the pivot itself -- in the order (split-axis key, photon id) of the split-axis list
__device__ __forceinline__ int entry_side(const uint4 &e, int ax, uint32_t pkey, uint32_t pid) {
  const uint32_t k = entry_key(e, ax);
  if (k < pkey || (k == pkey && e.x < pid)) return 0;
  return (k == pkey && e.x == pid) ? 2 : 1;
}

// The stable partition of every range of every list is ONE segmented inclusive scan whose output iterator does the
// scatter.  Scan value: low word = left entries, bits 32..62 = right entries, bit 63 = "reset" (the running counts start
// again here).  Neighbouring subtrees of a level are always separated by a placed pivot in the in-order layout, so the
// reset sits on the placed positions (and on position 0 of each list); cub's single-pass scan (decoupled look-back)
// carries the counts across thread blocks, and the assignment of the scan result to output element t moves entry t to
// its place -- no scan array, no second pass.
#define KD_RESET (1ull << 63)
struct SegScanOp {
  __host__ __device__ __forceinline__ unsigned long long operator()(unsigned long long a, unsigned long long b) const {
    return (b & KD_RESET) ? b : (((a & ~KD_RESET) + (b & ~KD_RESET)) | (a & KD_RESET));
  }
};
struct SideCount {
  Lists in; const uint32_t *pseg; Pivots pv; uint64_t n;
  __device__ __forceinline__ unsigned long long operator()(unsigned long long t) const {
    const int a = (t >= n) + (t >= 2 * n);
    const uint64_t j = t - (unsigned long long)a * n;
    const uint32_t s = pseg[j];
    if (s == KD_DONE) return KD_RESET;
    const int sd = entry_side(in.l[a][j], pv.axis[s], pv.key[s], pv.id[s]);
    return (sd == 0 ? 1ull : (sd == 1 ? (1ull << 32) : 0ull)) | (j == 0 ? KD_RESET : 0ull);
  }
};
struct ScatterCtx { Lists in; ListsOut out; const uint32_t *pseg; uint32_t *pseg_next; Pivots pv; uint64_t n; int L; };
struct ScatterRef {                  // proxy reference: assigning the inclusive scan value of element t performs its move
  const ScatterCtx *c; unsigned long long t;
  __device__ __forceinline__ const ScatterRef &operator=(unsigned long long v) const {
    const ScatterCtx &x = *c;
    const int a = (t >= x.n) + (t >= 2 * x.n);
    const uint64_t j = t - (unsigned long long)a * x.n;
    const uint32_t s = x.pseg[j];
    if (s == KD_DONE) { if (a == 0) x.pseg_next[j] = KD_DONE; return *this; }
    const uint4 e = x.in.l[a][j];
    const int sd = entry_side(e, x.pv.axis[s], x.pv.key[s], x.pv.id[s]);
    const uint64_t begin = x.pv.segbegin[s];
    const uint64_t m = subtree_size(2 * (uint64_t)s + 1, x.n, x.L);
    const unsigned long long incl = v & ~KD_RESET;
    uint64_t pos;
    if (sd == 0) pos = begin + ((incl & 0xffffffffull) - 1ull);
    else if (sd == 1) pos = begin + m + 1 + ((incl >> 32) - 1ull);
    else pos = begin + m;
    x.out.l[a][pos] = e;
    if (a == 0) x.pseg_next[pos] = sd == 0 ? 2u * s + 1u : (sd == 1 ? 2u * s + 2u : KD_DONE);
    return *this;
  }
};
struct ScatterIterator {
  using iterator_category = std::random_access_iterator_tag;
  using value_type = unsigned long long;
  using difference_type = long long;
  using pointer = void;
  using reference = ScatterRef;
  const ScatterCtx *c; unsigned long long t;
  __host__ __device__ __forceinline__ ScatterRef operator*() const { return ScatterRef{c, t}; }
  __host__ __device__ __forceinline__ ScatterRef operator[](difference_type i) const { return ScatterRef{c, t + (unsigned long long)i}; }
  __host__ __device__ __forceinline__ ScatterIterator operator+(difference_type i) const { return ScatterIterator{c, t + (unsigned long long)i}; }
  __host__ __device__ __forceinline__ ScatterIterator operator-(difference_type i) const { return ScatterIterator{c, t - (unsigned long long)i}; }
  __host__ __device__ __forceinline__ ScatterIterator &operator+=(difference_type i) { t += (unsigned long long)i; return *this; }
  __host__ __device__ __forceinline__ ScatterIterator &operator++() { ++t; return *this; }
  __host__ __device__ __forceinline__ difference_type operator-(const ScatterIterator &o) const { return (difference_type)(t - o.t); }
};

#define CKB(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { err = e_; goto done; } } while (0)

cudaError_t build_photon_map(GiContext *ctx, PhotonMapStore &m, uint64_t n, const double *d_lut) {
  cudaError_t err = cudaSuccess;
  cudaStream_t st = ctx->stream;
  const GiPhoton *ph = (const GiPhoton *)m.raw;
  int L = 0; while (((uint64_t)1 << L) <= n) L++;                       // floor(log2 n) + 1
  uint64_t *keys[2] = {nullptr, nullptr}; uint32_t *vals[2] = {nullptr, nullptr};
  uint32_t *heap_photon = nullptr, *bounds = nullptr; float *split = nullptr; uint8_t *axis = nullptr;
  void *tmp = nullptr; size_t tmp_bytes = 0;
  uint4 *lists[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr}; uint32_t *pseg[2] = {nullptr, nullptr}, *segbegin = nullptr, *pivot_key = nullptr;
  uint32_t *skeys[2] = {nullptr, nullptr}, *sids[2] = {nullptr, nullptr}; ScatterCtx *sctx_dev = nullptr;
  // A/B hook: the sort-per-level build.  It also takes maps of 2^30 photons and more (the packed counters of the scan
  // hold the left entries of all three lists in 32 bits).
  const bool sort_build = (getenv("GI_KD_SORT_BUILD") && atoi(getenv("GI_KD_SORT_BUILD")) != 0) || n >= ((uint64_t)1 << 30);
  double *px = nullptr, *py = nullptr, *pz = nullptr; uint2 *node = nullptr, *shade = nullptr; uint32_t *orig = nullptr;
  double *hint = nullptr;
  double *cbox = nullptr, *gpos = nullptr; double2 *srec = nullptr; uint64_t n_groups = 0, n_upper = 0;
  PhotonMapDev d{};
  const int B = 256;
  const unsigned G = (unsigned)((n + B - 1) / B);
  int cur = 0;
  // temporaries come from the device's stream-ordered pool (gi_create keeps freed pool memory): a frame rebuilds its
  // maps every time, and cudaMalloc / cudaFree of ~1 GB cost more than the build itself on a busy device
  CKB(cudaMallocAsync(&heap_photon, 4 * n, st)); CKB(cudaMallocAsync(&split, 4 * n, st)); CKB(cudaMallocAsync(&bounds, sort_build ? 24 * n : 256, st));
  CKB(cudaMallocAsync(&axis, n, st));
  if (sort_build) {
  for (int k = 0; k < 2; k++) { CKB(cudaMallocAsync(&keys[k], 8 * n, st)); CKB(cudaMallocAsync(&vals[k], 4 * n, st)); }
  CKB(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys[0], keys[1], vals[0], vals[1], (int)n, 0, 64, st));
  CKB(cudaMallocAsync(&tmp, tmp_bytes, st));
  init_kernel<<<G, B, 0, st>>>(n, keys[0], vals[0]);
  ctx->launches++;
  for (int level = 0; level < L; level++) {
    const uint64_t first = level == 0 ? 0 : ((uint64_t)1 << (level - 1)) - 1;    // placed two levels up: final
    const uint64_t cnt = n - first;
    const unsigned g = (unsigned)((cnt + B - 1) / B);
    const uint64_t level_nodes = (uint64_t)1 << level;
    const unsigned gn = (unsigned)((level_nodes + B - 1) / B);
    reset_bounds_kernel<<<gn, B, 0, st>>>(n, level, bounds);
    bounds_kernel<<<g, B, 0, st>>>(ph, first, n, level, keys[cur], vals[cur], bounds);
    axis_kernel<<<gn, B, 0, st>>>(n, level, bounds, axis);
    make_keys_kernel<<<g, B, 0, st>>>(ph, first, n, level, axis, keys[cur], vals[cur]);
    int end_bit = 33 + level; if (end_bit > 64) end_bit = 64;
    CKB(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys[cur] + first, keys[cur ^ 1] + first, vals[cur] + first,
                                        vals[cur ^ 1] + first, (int)cnt, 0, end_bit, st));
    cur ^= 1;
    assign_kernel<<<g, B, 0, st>>>(first, n, L, level, keys[cur], vals[cur], heap_photon, split);
    ctx->launches += 6;
    CKB(cudaGetLastError());
  }
  } else {
    // ---- presorted lists (see the header of this file)
    for (int k = 0; k < 6; k++) CKB(cudaMallocAsync(&lists[k], 16 * n, st));      // [0..2]: ping, [3..5]: pong
    for (int k = 0; k < 2; k++) CKB(cudaMallocAsync(&pseg[k], 4 * n, st));
    CKB(cudaMallocAsync(&segbegin, 4 * n, st)); CKB(cudaMallocAsync(&pivot_key, 4 * n, st));
    CKB(cudaMallocAsync(&skeys[0], 4 * n, st)); CKB(cudaMallocAsync(&skeys[1], 4 * n, st)); CKB(cudaMallocAsync(&sids[0], 4 * n, st));
    CKB(cudaMallocAsync(&sids[1], 4 * n, st));
    size_t sort_bytes = 0, scan_bytes = 0;
    CKB(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, skeys[0], skeys[1], sids[0], sids[1], (int)n, 0, 32, st));
    SideCount fn; fn.n = n; fn.pseg = pseg[0]; fn.pv.segbegin = segbegin; fn.pv.axis = axis; fn.pv.key = pivot_key; fn.pv.id = heap_photon;
    for (int a = 0; a < 3; a++) fn.in.l[a] = lists[a];
    CKB(cudaMallocAsync(&sctx_dev, 2 * sizeof(ScatterCtx), st));            // the output iterator reads its context from device memory
    {
      auto it = thrust::make_transform_iterator(thrust::counting_iterator<unsigned long long>(0ull), fn);
      CKB(cub::DeviceScan::InclusiveScan(nullptr, scan_bytes, it, ScatterIterator{sctx_dev, 0ull}, SegScanOp(), (size_t)(3 * n), st));
    }
    tmp_bytes = sort_bytes > scan_bytes ? sort_bytes : scan_bytes;
    CKB(cudaMallocAsync(&tmp, tmp_bytes, st));
    for (int a = 0; a < 3; a++) {
      presort_keys_kernel<<<G, B, 0, st>>>(ph, n, a, skeys[0], sids[0]);
      size_t bytes = tmp_bytes;
      CKB(cub::DeviceRadixSort::SortPairs(tmp, bytes, skeys[0], skeys[1], sids[0], sids[1], (int)n, 0, 32, st));
      entries_kernel<<<G, B, 0, st>>>(ph, n, sids[1], lists[a], a == 0 ? pseg[0] : nullptr);
      ctx->launches += 5;
    }
    int pp = 0;                                            // lists[3 pp ..] and pseg[pp] describe the level being built
    {
      ScatterCtx h[2];
      for (int q = 0; q < 2; q++) {
        for (int a = 0; a < 3; a++) { h[q].in.l[a] = lists[3 * q + a]; h[q].out.l[a] = lists[3 * (q ^ 1) + a]; }
        h[q].pseg = pseg[q]; h[q].pseg_next = pseg[q ^ 1]; h[q].pv = fn.pv; h[q].n = n; h[q].L = L;
      }
      CKB(cudaMemcpyAsync(sctx_dev, h, sizeof(h), cudaMemcpyHostToDevice, st));
      CKB(cudaStreamSynchronize(st));                       // h lives on this stack frame
    }
    for (int level = 0; level < L; level++) {
      const uint64_t level_nodes = (uint64_t)1 << level;
      const unsigned gn = (unsigned)((level_nodes + B - 1) / B);
      Lists in;
      for (int a = 0; a < 3; a++) in.l[a] = lists[3 * pp + a];
      choose_kernel<<<gn, B, 0, st>>>(n, L, level, in, segbegin, heap_photon, pivot_key, split, axis, bounds);
      if (level + 1 == L) { ctx->launches++; break; }      // the last level only places its pivots
      fn.in = in; fn.pseg = pseg[pp];
      auto it = thrust::make_transform_iterator(thrust::counting_iterator<unsigned long long>(0ull), fn);
      size_t bytes = tmp_bytes;
      CKB(cub::DeviceScan::InclusiveScan(tmp, bytes, it, ScatterIterator{sctx_dev + pp, 0ull}, SegScanOp(), (size_t)(3 * n), st));
      pp ^= 1;
      ctx->launches += 3;
      CKB(cudaGetLastError());
    }
  }
  // the map's own arrays: from the same pool (freed by free_map with cudaFreeAsync on this stream)
  m.stream = (void *)st;
  CKB(cudaMallocAsync(&px, 8 * n, st)); CKB(cudaMallocAsync(&py, 8 * n, st)); CKB(cudaMallocAsync(&pz, 8 * n, st));
  CKB(cudaMallocAsync(&node, 8 * n, st)); CKB(cudaMallocAsync(&shade, 8 * n, st)); CKB(cudaMallocAsync(&orig, 4 * n, st)); CKB(cudaMallocAsync(&srec, 32 * n, st));
  finalize_kernel<<<G, B, 0, st>>>(ph, n, heap_photon, split, axis, px, py, pz, node, shade, orig, d_lut, srec);
  ctx->launches++;
  CKB(cudaGetLastError());
  d.n = n; d.levels = L; d.px = px; d.py = py; d.pz = pz; d.node = node; d.shade = shade; d.orig = orig; d.srec = srec;
  // rows of 5-level groups (photon_dev.h) and their subtree boxes, bottom row first
  {
    const uint64_t last_cap = (uint64_t)1 << (L - 1), last = n - (last_cap - 1);      // slots / photons on the last level
    d.hb = L <= GI_GROUP_LEVELS ? L : (2 * last < last_cap ? GI_GROUP_LEVELS + 1 : GI_GROUP_LEVELS);
    const int rest = L - d.hb;                                                        // levels above the bottom row
    d.groups = 1 + (rest + GI_GROUP_LEVELS - 1) / GI_GROUP_LEVELS;
    d.h0 = d.groups == 1 ? d.hb : rest - GI_GROUP_LEVELS * (d.groups - 2);
  }
  for (int g = 0; g < d.groups; g++) {
    d.glevel[g] = g == 0 ? 0 : d.h0 + GI_GROUP_LEVELS * (g - 1);
    d.goff[g] = (uint32_t)n_groups;
    const uint64_t first = ((uint64_t)1 << d.glevel[g]) - 1;
    n_groups += n - first < first + 1 ? n - first : first + 1;
  }
  n_upper = d.groups > 1 ? d.goff[d.groups - 1] : 0;                    // groups above the bottom row
  CKB(cudaMallocAsync(&cbox, 1536 * (n_upper ? n_upper : 1), st));
  d.x6_base = (uint32_t)n_groups;
  d.x6_blocks = d.hb > GI_GROUP_LEVELS ? (uint32_t)((n - ((((uint64_t)1 << (L - 1)) - 1)) + 31) / 32) : 0u;
  CKB(cudaMallocAsync(&gpos, 768 * (n_groups + d.x6_blocks), st));
  d.cbox = cbox; d.gpos = gpos;
  CKB(cudaMallocAsync(&hint, 2 * sizeof(double), st));
  CKB(cudaMemsetAsync(hint, 0, 2 * sizeof(double), st));
  d.hint = hint;
  if (d.x6_blocks) { level6_kernel<<<(unsigned)(((uint64_t)d.x6_blocks * 32 + B - 1) / B), B, 0, st>>>(d, gpos); ctx->launches++; }
  if (n_upper) { cbox_init_kernel<<<(unsigned)((n_upper * 192 + B - 1) / B), B, 0, st>>>(cbox, n_upper * 192); ctx->launches++; }
  for (int g = d.groups - 1; g >= 0; g--) {
    const uint64_t first = ((uint64_t)1 << d.glevel[g]) - 1;
    const uint64_t rows = n - first < first + 1 ? n - first : first + 1;
    group_kernel<<<(unsigned)((rows * 32 + B - 1) / B), B, 0, st>>>(d, g, gpos, cbox);
    ctx->launches++;
  }
  CKB(cudaGetLastError());
  {
    // bounding box of the map (query_order.cu): the root's tight float bounds, one ulp outwards
    uint32_t hb[6];
    CKB(cudaMemcpyAsync(hb, bounds, sizeof(hb), cudaMemcpyDeviceToHost, st));
    CKB(cudaStreamSynchronize(st));
    for (int a = 0; a < 6; a++) {
      const uint32_t o = hb[a], u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
      float f; memcpy(&f, &u, 4);
      m.bbox[a] = (double)nextafterf(f, a < 3 ? -INFINITY : INFINITY);
    }
  }
  m.dev = d;
  m.built = true;
  px = py = pz = nullptr; node = shade = nullptr; orig = nullptr; cbox = nullptr; gpos = nullptr; srec = nullptr; hint = nullptr;
done:
  {
    void *temps[] = {keys[0], keys[1], vals[0], vals[1], heap_photon, split, bounds, axis, tmp, lists[0], lists[1], lists[2], lists[3], lists[4], lists[5],
                     pseg[0], pseg[1], segbegin, pivot_key, sctx_dev, skeys[0], skeys[1], sids[0], sids[1]};
    for (void *t : temps) if (t) cudaFreeAsync(t, st);
  }
  { void *own[] = {px, py, pz, node, shade, orig, cbox, gpos, srec, hint}; for (void *t : own) if (t) cudaFreeAsync(t, st); }      // only after a failure
  return err;
}

}  // namespace gi
