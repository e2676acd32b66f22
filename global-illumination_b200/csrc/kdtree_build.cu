// csrc/kdtree_build.cu -- GPU construction of a photon map: left-balanced kd-tree in heap layout (K5).
//
// Replaces `new R3Kdtree<Photon*>(photons, offset)` (src/photonmap.cpp:365,372), i.e. R3Kdtree::InsertPoints /
// PartitionPoints (src/R3Shapes/R3Kdtree.cpp:1624-1671, 1552-1620): a single-threaded recursive build with
// randomised quickselect and 31-photon leaf buckets (15 s for 1e7 photons on the CPU).
//
// Here: level-synchronous, sort-based construction of Jensen's left-balanced tree.  Every photon carries the id
// ("tag") of the subtree it currently belongs to.  For level l = 0..L-1:
//   1. key = (tag, float32 coordinate on the subtree's split axis); radix sort (key, photon id) -- after the sort
//      the photons of subtree s are contiguous and start at a closed-form offset (left-balanced => every subtree
//      size is known from n alone);
//   2. the photon at offset `size of the left child subtree` inside its segment becomes node s; photons before it
//      move to subtree 2s+1, after it to 2s+2;
//   3. the split axis of a subtree is the longest axis of the TIGHT bounding box of its photons (segmented min/max
//      with warp-aggregated atomics).  The reference splits on the longest axis of the inherited node box
//      (R3Kdtree.cpp:1640, R3Box::LongestAxis R3Box.cpp:258-274); for photons lying on planes that rule keeps
//      cutting along the plane normal, where every coordinate is equal and nothing is separated.  The choice of axis
//      does not change any query result, only how much of the tree a query has to visit.
// Splits are float32 (monotone rounding of the FP64 coordinate); the search widens them by one ulp, distances stay
// FP64, so the k-nearest SET is exact (see gather.cu).  Only photons not yet placed take part in a level's sort.
#include <cub/device/device_radix_sort.cuh>
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

#define CKB(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { err = e_; goto done; } } while (0)

cudaError_t build_photon_map(GiContext *ctx, PhotonMapStore &m, uint64_t n, const double *d_lut) {
  cudaError_t err = cudaSuccess;
  cudaStream_t st = ctx->stream;
  const GiPhoton *ph = (const GiPhoton *)m.raw;
  int L = 0; while (((uint64_t)1 << L) <= n) L++;                       // floor(log2 n) + 1
  uint64_t *keys[2] = {nullptr, nullptr}; uint32_t *vals[2] = {nullptr, nullptr};
  uint32_t *heap_photon = nullptr, *bounds = nullptr; float *split = nullptr; uint8_t *axis = nullptr;
  void *tmp = nullptr; size_t tmp_bytes = 0;
  double *px = nullptr, *py = nullptr, *pz = nullptr; uint2 *node = nullptr, *shade = nullptr; uint32_t *orig = nullptr;
  double *hint = nullptr;
  double *cbox = nullptr, *gpos = nullptr; double2 *srec = nullptr; uint64_t n_groups = 0, n_upper = 0;
  PhotonMapDev d{};
  const int B = 256;
  const unsigned G = (unsigned)((n + B - 1) / B);
  int cur = 0;
  // temporaries come from the device's stream-ordered pool (gi_create keeps freed pool memory): a frame rebuilds its
  // maps every time, and cudaMalloc / cudaFree of ~1 GB cost more than the build itself on a busy device
  for (int k = 0; k < 2; k++) { CKB(cudaMallocAsync(&keys[k], 8 * n, st)); CKB(cudaMallocAsync(&vals[k], 4 * n, st)); }
  CKB(cudaMallocAsync(&heap_photon, 4 * n, st)); CKB(cudaMallocAsync(&split, 4 * n, st)); CKB(cudaMallocAsync(&bounds, 24 * n, st));
  CKB(cudaMallocAsync(&axis, n, st));
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
    void *temps[] = {keys[0], keys[1], vals[0], vals[1], heap_photon, split, bounds, axis, tmp};
    for (void *t : temps) if (t) cudaFreeAsync(t, st);
  }
  { void *own[] = {px, py, pz, node, shade, orig, cbox, gpos, srec, hint}; for (void *t : own) if (t) cudaFreeAsync(t, st); }      // only after a failure
  return err;
}

}  // namespace gi
