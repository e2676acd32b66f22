// csrc/gather.cu -- photon k-nearest-neighbour gather + radiance estimate (K6), the bandwidth-bound hot kernel.
//
// Replaces EstimateRadiance (src/utils/photon_utils.cpp:72-162) and the search it calls,
// R3Kdtree::FindClosestQuick (src/R3Shapes/R3Kdtree.cpp:688-848): a recursive descent with a std:: max-heap in
// a std::vector per query.  Semantics kept: the k photons nearest to the query point with d^2 <= r^2 (both
// inclusive, :756-757), distances in FP64 with the reference's formula (R3SquaredDistance, R3Dist.cpp:41-46);
// estimate radius^2 = r^2 if fewer than k were found, else max(d^2, 1e-6) (photon_utils.cpp:85-96); side test,
// Phong weight, disk / cone / Gauss filters as photon_utils.cpp:109-158.  Documented tightening: exact d^2 ties at
// the k-th place go to the photon with the lower index (the reference's choice depends on its randomised tree).
//
// Design: one warp per query; the binary left-balanced tree is walked as a 32-ary tree of 5-level groups.
//   * A group = a node and the four levels below it: 31 photons in 5 contiguous runs of the heap layout, tested by
//     the 32 lanes in one round of FP64 distance tests.
//   * Its 32 child groups are one contiguous run of slots; the tight FP64 boxes of their subtrees (gbox, built with
//     the tree) are tested by the 32 lanes in one round, and the children whose box touches the search ball are
//     entered nearest first (warp min-reduction on the float bits of the box distance).  Box distances are lower
//     bounds of the FP64 photon distances, so nothing inside the ball is skipped.
//   * Candidates (d^2, slot, photon index) are appended to a per-warp buffer in shared memory with ballot/popc.
//     When the buffer runs full it is cut back with a 256-bin histogram of d^2 / bound (photons on a surface are
//     uniform in d^2, so the boundary bin holds ~cap/256 candidates): everything up to the bin that contains the
//     k-th candidate is kept whole, the few candidates inside that bin are ranked by shuffles, and the search bound
//     shrinks to the exact k-th smallest d^2.
#include <cub/device/device_radix_sort.cuh>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string>
#include "context.h"
#include "gather.h"

namespace gi {

#define FULL 0xffffffffu
#define NONE 0xffffffffu
#define GI_HIST_BINS 128          /* histogram of the exact cut (4 bins per lane) */
/* per-warp shared memory: cap candidates (8 B d2 + 4 B slot), the histogram, one row of walk keys per group row above gp */
#define GI_WARP_BYTES(cap, rows) ((size_t)(cap) * 12 + GI_HIST_BINS * 4 + (size_t)(rows) * 128)
/* STAGE: + a two-slot ring of bottom-row position blocks (x[32] y[32] z[32], and the level-6 block of a six-level row) */
#define GI_STAGE_SLOT_BYTES(deep) ((deep) ? 1536u : 768u)
#define GI_STAGE_BYTES(deep) (2u * GI_STAGE_SLOT_BYTES(deep))

// The per-warp buffers are addressed through 32-bit shared-window addresses and ld/st.shared: with C++ pointers the
// compiler re-derived the window base (S2R SR_CgaCtaId + 12 integer ops) in every iteration of the walk.
__device__ __forceinline__ double lds_f64(uint32_t a) { double v; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ void sts_f64(uint32_t a, double v) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v)); }
__device__ __forceinline__ void sts_u32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v)); }
__device__ __forceinline__ void red_u32(uint32_t a, uint32_t v) { asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(a), "r"(v)); }

// cp.async (LDGSTS): 8 bytes global -> shared without passing through a register; .ca keeps the line in L1, where the
// other warps of the SM (neighbouring queries) find it
__device__ __forceinline__ void cp_async8(uint32_t dst, const void *src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

struct WarpBuf {
  uint32_t a_d2, a_slot, a_hist, a_keys;          // shared memory: candidates (d2, heap slot), histogram bins, walk keys
  int limit;                                      // cut the buffer back to k once it holds more than this
  const uint32_t *orig;                           // global: heap slot -> photon index, read only to break exact d2 ties
  int cap;                                        // candidate slots (a power of two when idx_out is wanted: bitonic sort)
  __device__ __forceinline__ double d2(int i) const { return lds_f64(a_d2 + 8u * (uint32_t)i); }
  __device__ __forceinline__ uint32_t slot(int i) const { return lds_u32(a_slot + 4u * (uint32_t)i); }
  __device__ __forceinline__ uint32_t hist(int i) const { return lds_u32(a_hist + 4u * (uint32_t)i); }
  __device__ __forceinline__ void set_d2(int i, double v) const { sts_f64(a_d2 + 8u * (uint32_t)i, v); }
  __device__ __forceinline__ void set_slot(int i, uint32_t v) const { sts_u32(a_slot + 4u * (uint32_t)i, v); }
  __device__ __forceinline__ void set_hist(int i, uint32_t v) const { sts_u32(a_hist + 4u * (uint32_t)i, v); }
};

__device__ __forceinline__ bool lex_less(double da, uint32_t oa, double db, uint32_t ob) {
  return da < db || (da == db && oa < ob);
}

// warp-cooperative bitonic sort of buf[0..cap) ascending by (d2, photon index); cap is a power of two.  Padding
// entries carry d2 = +inf (and slot 0), real entries finite d2, so the index lookup happens only on real ties.
__device__ __noinline__ void bitonic_sort(const WarpBuf &b, int lane) {
  for (int k2 = 2; k2 <= b.cap; k2 <<= 1)
    for (int j = k2 >> 1; j > 0; j >>= 1) {
      for (int i = lane; i < b.cap; i += 32) {
        const int x = i ^ j;
        if (x > i) {
          const double di = b.d2(i), dx = b.d2(x);
          const uint32_t si = b.slot(i), sx = b.slot(x);
          const bool up = (i & k2) == 0;
          bool x_less_i = dx < di, i_less_x = di < dx;
          if (di == dx && di < INFINITY) { const uint32_t oi = b.orig[si], ox = b.orig[sx]; x_less_i = ox < oi; i_less_x = oi < ox; }
          if (up ? x_less_i : i_less_x) { b.set_d2(i, dx); b.set_d2(x, di); b.set_slot(i, sx); b.set_slot(x, si); }
        }
      }
      __syncwarp();
    }
}

// slow exact path (heavy exact ties, or a zero bound): every candidate counts the candidates that precede it in
// (d2, photon index) order -- O(count^2 / 32) -- and the k first are kept.  Keep flags go through the histogram words.
struct Sel { double bound; int count; int raw; };  // search state: squared search radius, candidates held, buffer entries they occupy

__device__ __noinline__ Sel rank_select(const WarpBuf b, int count, double bound, int k, int lane) {
  for (int base = 0; base < count; base += 32) {
    const int i = base + lane;
    const bool live = i < count;
    const double di = live ? b.d2(i) : 0.0;
    const uint32_t oi = live ? b.orig[b.slot(i)] : 0u;
    int rank = 0;
    for (int j = 0; j < count; j++) {
      const double dj = b.d2(j);
      if (dj < di) rank++;
      else if (dj == di && live && b.orig[b.slot(j)] < oi) rank++;
    }
    const unsigned keep = __ballot_sync(FULL, live && rank < k);
    if (lane == 0) b.set_hist(base >> 5, keep);
  }
  __syncwarp();
  int newcount = 0; double mx = 0.0;
  for (int base = 0; base < count; base += 32) {
    const int i = base + lane;
    const unsigned mk = b.hist(base >> 5);
    const bool keep = (mk >> lane) & 1u;
    double d = 0.0; uint32_t sl = 0;
    if (keep) { d = b.d2(i); sl = b.slot(i); }
    const int pos = newcount + __popc(mk & ((1u << lane) - 1u));
    __syncwarp();
    if (keep) { b.set_d2(pos, d); b.set_slot(pos, sl); mx = fmax(mx, d); }
    newcount += __popc(mk);
    __syncwarp();
  }
  for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(FULL, mx, o));
  Sel r; r.count = newcount; r.raw = newcount; r.bound = fmin(bound, mx);
  return r;
}

// Histogram bin of a candidate in the cut: monotone in d2, LOGARITHMIC in the gap to the bound -- 8 bins per octave of
// bound - d2, read off the exponent and the top three mantissa bits of the gap, relative to those of the bound (bin 0:
// gap = bound, bin 127: gap < bound / 65536).  Photons on a surface are uniform in d2 and put a handful of candidates
// into the bin of the k-th one; a query at some distance from a dense blob of a caustic map sees ALL its candidates in a
// thin shell of d2 just below the bound, which a linear histogram over [0, bound) cannot resolve (every cut then fell
// to the O(n^2) ranking: 30-50 ms launches in a cornell frame) -- the logarithmic bins resolve a shell of any thickness
// down to bound / 65536.  Dropped candidates (d2 = -1) land in bin 0; every caller tests d2 >= 0 as well.
__device__ __forceinline__ int cut_bin(double d, double bound, int base) {
  int v = (__double2hiint(bound - d) >> 17) - base;
  v = v < 0 ? 0 : (v > GI_HIST_BINS - 1 ? GI_HIST_BINS - 1 : v);
  return GI_HIST_BINS - 1 - v;
}

// Exact cut of the candidate buffer to the k smallest by (d2, orig).  Requires count > k.
// A 128-bin histogram (cut_bin) locates the bin B that holds the k-th candidate; bins below B are kept whole,
// and inside B (about a dozen candidates) each lane ranks one candidate
// against the others by shuffles.  The bound becomes the largest kept d2.  compact = false (the last cut of a search
// whose indices are not wanted): dropped candidates are only marked with d2 = -1 and stay where they are (raw = count).
__device__ __noinline__ Sel cut(const WarpBuf b, int count, double bound, int k, int lane, bool compact) {
  if (!(bound > 0.0)) return rank_select(b, count, bound, k, lane);
  for (int i = lane; i < GI_HIST_BINS; i += 32) b.set_hist(i, 0);
  __syncwarp();
  const int scale = (__double2hiint(bound) >> 17) - (GI_HIST_BINS - 1);      // cut_bin: code of the gap that maps to bin 0
  for (int i = lane; i < count; i += 32) {
    const int bin = cut_bin(b.d2(i), bound, scale);
    red_u32(b.a_hist + 4u * (uint32_t)bin, 1u);
  }
  __syncwarp();
  constexpr int PER = GI_HIST_BINS / 32;
  uint32_t h[PER]; uint32_t s = 0;
  for (int j = 0; j < PER; j++) { h[j] = b.hist(lane * PER + j); s += h[j]; }
  uint32_t incl = s;
  for (int o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += v; }
  const uint32_t excl = incl - s;
  const unsigned crossing = __ballot_sync(FULL, incl >= (uint32_t)k);
  const int owner = __ffs(crossing) - 1;                  // count > k  =>  some lane crosses
  int B = GI_HIST_BINS - 1, below = 0, m = 0;
  if (lane == owner) {
    uint32_t c = excl;
    for (int j = 0; j < PER; j++) { if (c + h[j] >= (uint32_t)k) { B = lane * PER + j; below = (int)c; m = (int)h[j]; break; } c += h[j]; }
  }
  B = __shfl_sync(FULL, B, owner); below = __shfl_sync(FULL, below, owner); m = __shfl_sync(FULL, m, owner);
  const int need = k - below;                             // candidates to keep from bin B, 1 <= need <= m
  __syncwarp();
  if (need < m) {
    if (m > 32) return rank_select(b, count, bound, k, lane);
    int c = 0;                                            // list the members of bin B in hist[0..m)
    for (int base = 0; base < count; base += 32) {
      const int i = base + lane;
      bool inb = false;
      if (i < count) { const int bin = cut_bin(b.d2(i), bound, scale); inb = bin == B; }
      const unsigned mk = __ballot_sync(FULL, inb);
      if (inb) b.set_hist(c + __popc(mk & ((1u << lane) - 1u)), (uint32_t)i);
      c += __popc(mk);
    }
    __syncwarp();
    double md = INFINITY; uint32_t mo = 0xffffffffu; int mi = -1;
    if (lane < m) { mi = (int)b.hist(lane); md = b.d2(mi); mo = b.orig[b.slot(mi)]; }
    int rank = 0;
    for (int l = 0; l < m; l++) {
      const double od = __shfl_sync(FULL, md, l); const uint32_t oo = __shfl_sync(FULL, mo, l);
      rank += lex_less(od, oo, md, mo) ? 1 : 0;
    }
    if (lane < m && rank >= need) b.set_d2(mi, -1.0);     // dropped
    __syncwarp();
  }
  if (!compact) {
    double mx = 0.0;
    for (int i = lane; i < count; i += 32) {
      const double d = b.d2(i);
      const int bin = cut_bin(d, bound, scale);
      if (bin > B) b.set_d2(i, -1.0);
      else if (d >= 0.0) mx = fmax(mx, d);
    }
    for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(FULL, mx, o));
    __syncwarp();
    Sel r; r.count = k; r.raw = count; r.bound = fmin(bound, mx);
    return r;
  }
  int newcount = 0; double mx = 0.0;
  for (int base = 0; base < count; base += 32) {
    const int i = base + lane;
    double d = 0.0; uint32_t sl = 0; bool keep = false;
    if (i < count) {
      d = b.d2(i); sl = b.slot(i);
      const int bin = cut_bin(d, bound, scale);
      keep = d >= 0.0 && bin <= B;
    }
    const unsigned mk = __ballot_sync(FULL, keep);
    const int pos = newcount + __popc(mk & ((1u << lane) - 1u));
    __syncwarp();
    if (keep) { b.set_d2(pos, d); b.set_slot(pos, sl); mx = fmax(mx, d); }
    newcount += __popc(mk);
    __syncwarp();
  }
  for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(FULL, mx, o));
  Sel r; r.count = newcount; r.raw = newcount; r.bound = fmin(bound, mx);
  return r;
}

// after the traversal: exactly the k nearest (or everything found, if fewer); bound = largest kept d2 when k are held
// (raw: buffer entries to look at -- more than count after a cut that did not compact; dropped ones have d2 < 0)
__device__ __forceinline__ void finish(const WarpBuf &b, int &count, int &raw, double &bound, int k, int lane, bool compact) {
  raw = count;
  if (count > k) { const Sel r = cut(b, count, bound, k, lane, compact); count = r.count; raw = r.raw; bound = r.bound; return; }
  if (count < k) return;
  double mx = 0.0;
  for (int i = lane; i < count; i += 32) mx = fmax(mx, b.d2(i));
  for (int s = 16; s > 0; s >>= 1) mx = fmax(mx, __shfl_xor_sync(FULL, mx, s));
  bound = fmin(bound, mx);
}

__device__ __forceinline__ void append(const WarpBuf &b, int &count, double &bound, int k, int lane, uint32_t lt_mask, bool pred,
                                       double d2, uint32_t slot) {
  const unsigned mask = __ballot_sync(FULL, pred);
  if (mask == 0) return;
  const int pos = count + __popc(mask & lt_mask);
  if (pred) { b.set_d2(pos, d2); b.set_slot(pos, slot); }
  count += __popc(mask);
  __syncwarp();
  if (count > b.limit) { const Sel r = cut(b, count, bound, k, lane, true); count = r.count; bound = r.bound; }
}

struct P3 { double x, y, z; };
// photons of a group: block gi of gpos, entry `lane` (+inf where the group has no photon: the distance test fails)
__device__ __forceinline__ P3 load_group(const double *__restrict__ gpos_lane, uint32_t gi) {
  const double *p = gpos_lane + (size_t)gi * 96;
  P3 r; r.x = p[0]; r.y = p[32]; r.z = p[64];
  return r;
}
__device__ __forceinline__ double dist2(const P3 &p, double qx, double qy, double qz) {
  const double dx = qx - p.x, dy = qy - p.y, dz = qz - p.z;                        // v = query - photon; v.Dot(v)
  return (dx * dx) + (dy * dy) + (dz * dz);
}
// lower bound (after rounding) of the squared distance from the query to anything inside a subtree box
// (box of child `lane` of group gi, from the group's cbox block; +inf when there is no such child)
__device__ __forceinline__ double box_dist2(const double *__restrict__ cbox_lane, uint32_t gi, double qx, double qy, double qz) {
  const double *b = cbox_lane + (size_t)gi * 192;
  const double lx = b[0], hx = b[32], ly = b[64], hy = b[96], lz = b[128], hz = b[160];
  const double dx = fmax(fmax(lx - qx, qx - hx), 0.0), dy = fmax(fmax(ly - qy, qy - hy), 0.0), dz = fmax(fmax(lz - qz, qz - hz), 0.0);
  return ((dx * dx) + (dy * dy) + (dz * dz)) * 0.99999999999999;
}

// The tail of a query, shared by the walking kernel and the small-map kernel: radiance estimate over the kept candidates
// (photon_utils.cpp:80-161) and the outputs.  `mine` = word `lane` of the query record (17 doubles), bound = the largest
// kept d^2 when k candidates are held, raw = buffer entries to look at (dropped ones carry d2 < 0).
__device__ __forceinline__ void estimate_and_store(const WarpBuf &b, const PhotonMapDev &m, const GatherArgs &ga, double mine, int count, int raw,
                                                   double bound, int k, int lane, uint64_t qi, int32_t *idx_out, int32_t *count_out,
                                                   double *rgb_out) {
  // ---- radiance estimate, photon_utils.cpp:80-161
  const int found = count;
  double est[3] = {0.0, 0.0, 0.0};
  if (found > 0 && rgb_out) {
    const double nx = __shfl_sync(FULL, mine, 3), ny = __shfl_sync(FULL, mine, 4), nz = __shfl_sync(FULL, mine, 5);
    const double bx = __shfl_sync(FULL, mine, 6), by = __shfl_sync(FULL, mine, 7), bz = __shfl_sync(FULL, mine, 8);
    const double cos_theta = __shfl_sync(FULL, mine, 9);
    double kd[3], ks[3];
    for (int c = 0; c < 3; c++) { kd[c] = __shfl_sync(FULL, mine, 10 + c); ks[c] = __shfl_sync(FULL, mine, 13 + c); }
    const double shin = __shfl_sync(FULL, mine, 16);
    const bool has_spec = ks[0] != 0.0 || ks[1] != 0.0 || ks[2] != 0.0;
    double max_d2 = 1.0e-6;                                                   // RN_EPSILON
    if (found < k) max_d2 = ga.r2;
    else max_d2 = fmax(max_d2, bound);                                        // bound == max kept d^2 after finish()
    double c1 = 1.0, c2 = 1.0, total_w = 0.0;
    if (ga.filter == GI_FILTER_CONE) c1 = 1.0 / (ga.fk * sqrt(max_d2));
    else if (ga.filter == GI_FILTER_GAUSS) { c1 = exp(-ga.fb); c2 = 1.0 / (2.0 * max_d2); }
    for (int i = lane; i < raw; i += 32) {
      const double d2i = b.d2(i);
      if (d2i < 0.0) continue;                                                // dropped by the last cut
      const double2 *rec = m.srec + 2 * (size_t)b.slot(i);
      const double2 ra = rec[0], rb = rec[1];
      const double ix = ra.x, iy = ra.y, iz = rb.x;
      const uint2 sh = make_uint2((uint32_t)__double2loint(rb.y), (uint32_t)__double2hiint(rb.y));
      const double perp = (nx * ix) + (ny * iy) + (nz * iz);
      if ((cos_theta < 0 && perp < 0) || (cos_theta > 0 && perp > 0)) continue;
      const uint32_t e = sh.x >> 24;
      if (e == 0) { if (ga.filter == GI_FILTER_GAUSS) {
          total_w += 1.0 - (1.0 - pow(c1, c2 * d2i)) / (1.0 - c1); } continue; }
      const double inv = __longlong_as_double((long long)(e - 136 + 1023) << 52);   // ldexp(1.0, e - 128 - 8)
      double rgb[3] = {(double)(sh.x & 0xff) * inv, (double)((sh.x >> 8) & 0xff) * inv, (double)((sh.x >> 16) & 0xff) * inv};
      double cos_alpha = (bx * -ix) + (by * -iy) + (bz * -iz);
      if (cos_alpha < 0) cos_alpha = 0;
      const double pw = has_spec ? pow(cos_alpha, shin) : 0.0;
      const double ap = fabs(perp);
      for (int c = 0; c < 3; c++) rgb[c] *= ap * kd[c] + pw * ks[c];
      if (ga.filter == GI_FILTER_CONE) {
        const double w = 1.0 - c1 * sqrt(d2i);
        for (int c = 0; c < 3; c++) rgb[c] *= w;
      } else if (ga.filter == GI_FILTER_GAUSS) {
        const double w = 1.0 - (1.0 - pow(c1, c2 * d2i)) / (1.0 - c1);
        for (int c = 0; c < 3; c++) rgb[c] *= w;
        total_w += w;
      }
      for (int c = 0; c < 3; c++) est[c] += rgb[c];
    }
    for (int o = 16; o > 0; o >>= 1) {
      for (int c = 0; c < 3; c++) est[c] += __shfl_xor_sync(FULL, est[c], o);
      total_w += __shfl_xor_sync(FULL, total_w, o);
    }
    const double PI = 3.14159265358979323846;
    if (ga.filter == GI_FILTER_DISK && max_d2 > 0) { for (int c = 0; c < 3; c++) est[c] /= PI * max_d2; }
    else if (ga.filter == GI_FILTER_CONE && max_d2 > 0) { for (int c = 0; c < 3; c++) est[c] /= (1.0 - 2.0 / 3.0 / ga.fk) * PI * max_d2; }
    else if (ga.filter == GI_FILTER_GAUSS && total_w > 0 && max_d2 > 0) {
      for (int c = 0; c < 3; c++) est[c] *= ga.fa * (found / total_w) / (PI * max_d2);
    } else { est[0] = est[1] = est[2] = 0.0; }
  }
  if (rgb_out && lane == 0) { rgb_out[3 * qi] = est[0]; rgb_out[3 * qi + 1] = est[1]; rgb_out[3 * qi + 2] = est[2]; }
  if (count_out && lane == 0) count_out[qi] = found;
  if (idx_out) {
    // ascending photon indices: put the index into the d2 column (exact in a double) and sort on it
    for (int i = lane; i < b.cap; i += 32) b.set_d2(i, i < found ? (double)b.orig[b.slot(i)] : INFINITY);
    __syncwarp();
    bitonic_sort(b, lane);
    for (int i = lane; i < k; i += 32) idx_out[qi * (uint64_t)k + i] = i < found ? (int32_t)b.d2(i) : -1;
  }
}

// ------------------------------------------------------------------------------------------------ small maps
// A default frame sends 97 % of its gathers to the GLOBAL map: ~2 000 photons (52 KB of positions: it lives in L1), k = 50,
// billions of queries.  Walking a kd-tree per query is the wrong shape for that.  Here a warp takes a CHUNK of 32 queries
// that are neighbours on the Morton curve (millimetres apart, against a k-th distance of decimetres) and
//   1. finds the exact k-th distance R of the chunk's first query c by brute force over the whole map,
//   2. collects ONCE the pool of photons within R + 2 rho of c (rho = the chunk's radius around c): every query q of the
//      chunk has its k nearest within R + |q - c| of itself (that ball contains c's), hence within R + 2 rho of c,
//   3. ranks every query against the pool only (a few dozen photons): exact FP64 distances, the same histogram cut and
//      the same estimate as the walking kernel -- identical kNN sets, ties by photon index.
// A chunk whose pool does not fit (scattered queries: small unsorted batches) scans the whole map per query instead.
// debug counters of the small-map kernel (GI_GATHER_DEBUG=1 prints them after every launch): chunks, chunks whose pool did
// not fit, chunks ranked by the slow (histogram) path, pool entries, centre searches that found fewer than k
__device__ unsigned long long g_small_dbg[8];
#define GI_SMALL_MAX_PHOTONS 8192
#define GI_SMALL_POOL 128            /* pool entries per chunk (u16 heap slots; the rank kernel also stages their positions) */
#define GI_SMALL_SUPER 256           /* super-pool entries collected during the centre search */
#define GI_SMALL_CHUNK 32
#define GI_SMALL_WARPS 24
#define GI_SMALL_RANK_BYTES (32u * GI_SMALL_SHELL * 10u + 24u * GI_SMALL_POOL + 2u * GI_SMALL_POOL)   /* rank kernel, per warp */
#define GI_SMALL_SHELL 16            /* shell entries per lane kept for the exact pick of the k-th (8 B d2 + 2 B slot each) */

__device__ __forceinline__ uint32_t lds_u16(uint32_t a) { uint16_t v; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a)); return v; }
__device__ __forceinline__ void sts_u16(uint32_t a, uint32_t v) { asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "h"((uint16_t)v)); }

// ---- thread-per-query ranking of a chunk (small-map kernel, radiance only).  Every lane owns one query of the chunk; the
// pool is walked by all lanes together, so every photon is one uniform (broadcast) load and the distance test, the
// selection and the estimate run for 32 queries per instruction.  The exact k-th (d2, photon index) of a lane is found
// without storing distances: up to three passes refine a 16-bin histogram (packed 16-bit counters in four registers, bins
// with exact edges) around the bin that holds the k-th candidate; what is left in that bin (usually one or two photons)
// is ordered exactly by repeated minimum extraction.
struct ThreadQuery { double nx, ny, nz, bx, by, bz, cos_theta, kd0, kd1, kd2, ks0, ks1, ks2, shin; };

__device__ __forceinline__ double bin_edge(double lo, double hi, double w, int j) { return j >= 16 ? hi : lo + (double)j * w; }

// one photon's term of the estimate (the loop body of estimate_and_store, photon_utils.cpp:96-150)
__device__ __forceinline__ void photon_term(const GatherArgs &ga, const ThreadQuery &q, bool has_spec, const double2 ra, const double2 rb,
                                            double d2i, double c1, double c2, double &e0, double &e1, double &e2, double &total_w) {
  const double ix = ra.x, iy = ra.y, iz = rb.x;
  const uint32_t shx = (uint32_t)__double2loint(rb.y);
  const double perp = (q.nx * ix) + (q.ny * iy) + (q.nz * iz);
  if ((q.cos_theta < 0 && perp < 0) || (q.cos_theta > 0 && perp > 0)) return;
  const uint32_t e = shx >> 24;
  if (e == 0) { if (ga.filter == GI_FILTER_GAUSS) total_w += 1.0 - (1.0 - pow(c1, c2 * d2i)) / (1.0 - c1); return; }
  const double inv = __longlong_as_double((long long)(e - 136 + 1023) << 52);
  double r0 = (double)(shx & 0xff) * inv, r1 = (double)((shx >> 8) & 0xff) * inv, r2 = (double)((shx >> 16) & 0xff) * inv;
  double cos_alpha = (q.bx * -ix) + (q.by * -iy) + (q.bz * -iz);
  if (cos_alpha < 0) cos_alpha = 0;
  const double pw = has_spec ? pow(cos_alpha, q.shin) : 0.0;
  const double ap = fabs(perp);
  r0 *= ap * q.kd0 + pw * q.ks0; r1 *= ap * q.kd1 + pw * q.ks1; r2 *= ap * q.kd2 + pw * q.ks2;
  if (ga.filter == GI_FILTER_CONE) { const double w = 1.0 - c1 * sqrt(d2i); r0 *= w; r1 *= w; r2 *= w; }
  else if (ga.filter == GI_FILTER_GAUSS) { const double w = 1.0 - (1.0 - pow(c1, c2 * d2i)) / (1.0 - c1); r0 *= w; r1 *= w; r2 *= w; total_w += w; }
  e0 += r0; e1 += r1; e2 += r2;
}

// np pool entries: positions at a_pos (x[POOL] y[POOL] z[POOL]) and heap slots at a_pool, both in shared memory.
// active: this lane owns a query; bound: its search bound (all of its k nearest lie within, d2 <= bound).
__device__ __forceinline__ void rank_chunk_per_thread(const PhotonMapDev &m, const GatherArgs &ga, uint32_t a_pool, uint32_t a_pos, uint32_t a_shell, int np, bool active,
                                                      const GiGatherQuery *__restrict__ queries, uint64_t qi, double qx, double qy, double qz,
                                                      double bound, double shell_lo, int32_t *count_out, double *rgb_out) {
  const int k = ga.k;
  const int lane = threadIdx.x & 31;
  // ---- selection: [lo, hi) narrows around the k-th smallest d2; `below` = candidates with d2 < lo
  double lo = 0.0, hi = __longlong_as_double(__double_as_longlong(bound) + 1);       // d2 <= bound  <=>  d2 < hi
  int below = 0, need = 0;
  bool done = !active;
  double thr_d2 = 0.0; uint32_t thr_idx = 0; bool thr_strict = true;                   // kept: d2 < thr_d2, or (!strict) d2 == thr_d2 and index <= thr_idx
  // Fast path.  The k-th distance of a query q lies in the shell [Rc - |q - c|, Rc + |q - c|] around the centre's k-th
  // distance Rc (fewer than k photons are nearer to c than Rc, at least k are within Rc): ONE pass counts the candidates
  // below the shell and keeps the handful inside it in registers, the k-th is then picked among those by rank.
  {
    // shell lists in shared memory, entry t of lane l: d2 at a_shell + 8 (32 t + l), slot at a_shell + 8 * 32 * S + 2 (32 t + l)
    constexpr int S = GI_SMALL_SHELL;
    const uint32_t a_sd2 = a_shell + 8u * (uint32_t)lane, a_ss = a_shell + 8u * 32u * S + 2u * (uint32_t)lane;
    int nb = 0, ns = 0;
#pragma unroll 4
    for (int i = 0; i < np; i++) {
      const P3 p = {lds_f64(a_pos + 8u * (uint32_t)i), lds_f64(a_pos + 8u * (GI_SMALL_POOL + (uint32_t)i)), lds_f64(a_pos + 8u * (2u * GI_SMALL_POOL + (uint32_t)i))};
      const double d2 = dist2(p, qx, qy, qz);
      if (d2 < shell_lo) nb++;
      else if (d2 < hi) {
        if (ns < S) { sts_f64(a_sd2 + 256u * (uint32_t)ns, d2); sts_u16(a_ss + 64u * (uint32_t)ns, lds_u16(a_pool + 2u * (uint32_t)i)); }
        ns++;
      }
    }
    const bool fits = ns <= S && nb < k;
    if (__all_sync(FULL, fits || !active)) {
      if (active) {
        if (nb + ns <= k) { thr_d2 = hi; thr_strict = true; }                          // the whole shell is kept
        else {
          const int want = k - nb - 1;                                                 // rank of the k-th inside the shell, 0 <= want < ns
          for (int t = 0; t < ns; t++) {
            const double dt = lds_f64(a_sd2 + 256u * (uint32_t)t);
            int rank = 0, ties = 0;
            for (int u = 0; u < ns; u++) { const double du = lds_f64(a_sd2 + 256u * (uint32_t)u); rank += du < dt ? 1 : 0; ties += du == dt ? 1 : 0; }
            if (rank <= want && want < rank + ties) {                                  // the k-th has this d2
              const uint32_t ot = m.orig[lds_u16(a_ss + 64u * (uint32_t)t)];
              int r2 = rank;
              if (ties > 1)
                for (int u = 0; u < ns; u++)
                  if (lds_f64(a_sd2 + 256u * (uint32_t)u) == dt && m.orig[lds_u16(a_ss + 64u * (uint32_t)u)] < ot) r2++;
              if (r2 == want) { thr_d2 = dt; thr_idx = ot; thr_strict = false; break; }
            }
          }
        }
      }
      done = true;
    } else if (lane == 0) atomicAdd(&g_small_dbg[2], 1ull);
    __syncwarp();
  }
  for (int pass = 0; pass < 3; pass++) {
    if (__all_sync(FULL, done)) break;
    const double w = (hi - lo) * (1.0 / 16.0);
    const double scale = w > 0.0 ? 1.0 / w : 0.0;
    unsigned long long c0 = 0ull, c1 = 0ull, c2 = 0ull, c3 = 0ull;                 // 16 counters of 16 bits (np <= 8192)
#pragma unroll 2
    for (int i = 0; i < np; i++) {
      const P3 p = {lds_f64(a_pos + 8u * (uint32_t)i), lds_f64(a_pos + 8u * (GI_SMALL_POOL + (uint32_t)i)), lds_f64(a_pos + 8u * (2u * GI_SMALL_POOL + (uint32_t)i))};
      const double d2 = dist2(p, qx, qy, qz);
      if (d2 >= lo && d2 < hi) {
        int bin = (int)((d2 - lo) * scale);
        bin = bin > 15 ? 15 : bin;
        if (bin > 0 && d2 < bin_edge(lo, hi, w, bin)) bin--;                           // exact edges: the next pass tests against them
        else if (bin < 15 && d2 >= bin_edge(lo, hi, w, bin + 1)) bin++;
        const unsigned long long inc = 1ull << (16 * (bin & 3));
        const int word = bin >> 2;
        c0 += word == 0 ? inc : 0ull; c1 += word == 1 ? inc : 0ull; c2 += word == 2 ? inc : 0ull; c3 += word == 3 ? inc : 0ull;
      }
    }
    if (!done) {
      const unsigned long long cw[4] = {c0, c1, c2, c3};
      int total = 0;
      for (int j = 0; j < 16; j++) total += (int)((cw[j >> 2] >> (16 * (j & 3))) & 0xffff);
      if (below + total <= k) { thr_d2 = hi; thr_strict = true; done = true; }      // everything in range is kept
      else {
        int cum = 0, B = 0, mB = 0;
        for (int j = 0; j < 16; j++) {
          const int c = (int)((cw[j >> 2] >> (16 * (j & 3))) & 0xffff);
          if (below + cum + c >= k) { B = j; mB = c; break; }
          cum += c;
        }
        below += cum; need = k - below;                                               // 1 <= need <= mB
        const double nlo = bin_edge(lo, hi, w, B), nhi = bin_edge(lo, hi, w, B + 1);
        lo = nlo; hi = nhi;
        if (need == mB) { thr_d2 = hi; thr_strict = true; done = true; }              // the whole bin is kept
      }
    }
  }
  // ---- what is left in [lo, hi): the `need` smallest by (d2, photon index), one minimum per pass
  if (!__all_sync(FULL, done)) {
    double last_d2 = -1.0; uint32_t last_idx = 0; int left = done ? 0 : need;
    while (__any_sync(FULL, left > 0)) {
      double best_d2 = INFINITY; uint32_t best_idx = 0xffffffffu;
      for (int i = 0; i < np; i++) {
        const P3 p = {lds_f64(a_pos + 8u * (uint32_t)i), lds_f64(a_pos + 8u * (GI_SMALL_POOL + (uint32_t)i)), lds_f64(a_pos + 8u * (2u * GI_SMALL_POOL + (uint32_t)i))};
        const double d2 = dist2(p, qx, qy, qz);
        if (left > 0 && d2 >= lo && d2 < hi && d2 >= last_d2 && d2 <= best_d2) {
          const uint32_t oi = m.orig[lds_u16(a_pool + 2u * (uint32_t)i)];
          const bool after_last = d2 > last_d2 || (last_d2 >= 0.0 && oi > last_idx) || last_d2 < 0.0;
          if (after_last && (d2 < best_d2 || oi < best_idx)) { best_d2 = d2; best_idx = oi; }
        }
      }
      if (left > 0) { last_d2 = best_d2; last_idx = best_idx; left--; }
    }
    if (!done) { thr_d2 = last_d2; thr_idx = last_idx; thr_strict = false; done = true; }
  }
  // ---- the query record (17 doubles per lane)
  ThreadQuery q;
  const double *qp = (const double *)(queries + qi);
  if (active) {
    q.nx = qp[3]; q.ny = qp[4]; q.nz = qp[5]; q.bx = qp[6]; q.by = qp[7]; q.bz = qp[8]; q.cos_theta = qp[9];
    q.kd0 = qp[10]; q.kd1 = qp[11]; q.kd2 = qp[12]; q.ks0 = qp[13]; q.ks1 = qp[14]; q.ks2 = qp[15]; q.shin = qp[16];
  } else { q.nx = q.ny = q.nz = q.bx = q.by = q.bz = q.cos_theta = q.kd0 = q.kd1 = q.kd2 = q.ks0 = q.ks1 = q.ks2 = q.shin = 0.0; }
  const bool has_spec = q.ks0 != 0.0 || q.ks1 != 0.0 || q.ks2 != 0.0;
  // ---- number found and the largest kept d2 (the filters other than the disk need them before the sum)
  int found = 0; double maxkept = 0.0;
  if (ga.filter != GI_FILTER_DISK) {
    for (int i = 0; i < np; i++) {
      const P3 p = {lds_f64(a_pos + 8u * (uint32_t)i), lds_f64(a_pos + 8u * (GI_SMALL_POOL + (uint32_t)i)), lds_f64(a_pos + 8u * (2u * GI_SMALL_POOL + (uint32_t)i))};
      const double d2 = dist2(p, qx, qy, qz);
      bool in = d2 < thr_d2;
      if (!thr_strict && d2 == thr_d2) in = m.orig[lds_u16(a_pool + 2u * (uint32_t)i)] <= thr_idx;
      if (in) { found++; maxkept = fmax(maxkept, d2); }
    }
  }
  double max_d2 = 1.0e-6, c1 = 1.0, c2 = 1.0;
  if (ga.filter != GI_FILTER_DISK) {
    max_d2 = found < k ? ga.r2 : fmax(1.0e-6, maxkept);
    if (ga.filter == GI_FILTER_CONE) c1 = 1.0 / (ga.fk * sqrt(max_d2));
    else { c1 = exp(-ga.fb); c2 = 1.0 / (2.0 * max_d2); }
    found = 0; maxkept = 0.0;
  }
  // ---- the estimate over the kept photons.  The shell lists are dead now: their region takes the pool's estimate records
  //      (32 B each: incident direction + RGBE), so the loop below reads shared memory only.
  __syncwarp();
  if (rgb_out) {
    for (int i = lane; i < np; i += 32) {
      const double2 *rec = m.srec + 2 * (size_t)lds_u16(a_pool + 2u * (uint32_t)i);
      const double2 ra = rec[0], rb = rec[1];
      sts_f64(a_shell + 32u * (uint32_t)i, ra.x); sts_f64(a_shell + 32u * (uint32_t)i + 8u, ra.y);
      sts_f64(a_shell + 32u * (uint32_t)i + 16u, rb.x); sts_f64(a_shell + 32u * (uint32_t)i + 24u, rb.y);
    }
  }
  __syncwarp();
  double e0 = 0.0, e1 = 0.0, e2 = 0.0, total_w = 0.0;
#pragma unroll 2
  for (int i = 0; i < np; i++) {
    const P3 p = {lds_f64(a_pos + 8u * (uint32_t)i), lds_f64(a_pos + 8u * (GI_SMALL_POOL + (uint32_t)i)), lds_f64(a_pos + 8u * (2u * GI_SMALL_POOL + (uint32_t)i))};
    const double d2 = dist2(p, qx, qy, qz);
    bool in = d2 < thr_d2;
    if (!thr_strict && d2 == thr_d2) in = m.orig[lds_u16(a_pool + 2u * (uint32_t)i)] <= thr_idx;
    if (in) {
      found++; maxkept = fmax(maxkept, d2);
      if (rgb_out) {
        const uint32_t ar = a_shell + 32u * (uint32_t)i;
        photon_term(ga, q, has_spec, make_double2(lds_f64(ar), lds_f64(ar + 8u)), make_double2(lds_f64(ar + 16u), lds_f64(ar + 24u)), d2, c1, c2, e0, e1, e2,
                    total_w);
      }
    }
  }
  if (!active) return;
  if (ga.filter == GI_FILTER_DISK) max_d2 = found < k ? ga.r2 : fmax(1.0e-6, maxkept);
  const double PI = 3.14159265358979323846;
  if (found > 0 && ga.filter == GI_FILTER_DISK && max_d2 > 0) { e0 /= PI * max_d2; e1 /= PI * max_d2; e2 /= PI * max_d2; }
  else if (found > 0 && ga.filter == GI_FILTER_CONE && max_d2 > 0) { const double dn = (1.0 - 2.0 / 3.0 / ga.fk) * PI * max_d2; e0 /= dn; e1 /= dn; e2 /= dn; }
  else if (found > 0 && ga.filter == GI_FILTER_GAUSS && total_w > 0 && max_d2 > 0) { const double f = ga.fa * (found / total_w) / (PI * max_d2); e0 *= f; e1 *= f; e2 *= f; }
  else { e0 = e1 = e2 = 0.0; }
  if (rgb_out) { rgb_out[3 * qi] = e0; rgb_out[3 * qi + 1] = e1; rgb_out[3 * qi + 2] = e2; }
  if (count_out) count_out[qi] = found;
}

// Two kernels, so that the warps of an SM run one phase's code at a time (one fused kernel spent 5 cycles per issued
// instruction waiting for instruction fetch): gather_small_kernel does the centre search and the pool of every chunk and
// leaves them in global memory (SmallChunk + 256 u16 slots per chunk); small_rank_kernel ranks a lane per query.  Chunks
// that cannot be pooled -- and every chunk when the index output is wanted -- are finished by the first kernel, a warp per
// query, and marked npool = -1.
struct SmallChunk { double cx, cy, cz, Rc; int32_t npool, centre_full; double pad; };       // 48 bytes

__global__ void __launch_bounds__(GI_SMALL_WARPS * 32, 1)
gather_small_kernel(PhotonMapDev m, const GiGatherQuery *__restrict__ queries, const uint32_t *__restrict__ order, uint64_t nq,
                    GatherArgs ga, unsigned long long *work_counter, SmallChunk *__restrict__ info, uint16_t *__restrict__ pools,
                    int32_t *idx_out, int32_t *count_out, double *rgb_out) {
  extern __shared__ double smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  WarpBuf b; b.cap = ga.cap; b.orig = m.orig;
  { const int room = ga.cap - 32; b.limit = room < ga.k + ga.slack ? room : ga.k + ga.slack; }
  uint32_t a_pool, a_super;
  {
    uint32_t base = (uint32_t)__cvta_generic_to_shared(smem) + (uint32_t)warp * ((uint32_t)GI_WARP_BYTES(ga.cap, 0) + GI_SMALL_POOL * 2u + GI_SMALL_SUPER * 2u);
    asm volatile("" : "+r"(base));
    b.a_d2 = base; b.a_slot = base + (uint32_t)ga.cap * 8u; b.a_hist = b.a_slot + (uint32_t)ga.cap * 4u; b.a_keys = 0;
    a_pool = base + (uint32_t)GI_WARP_BYTES(ga.cap, 0);
    a_super = a_pool + GI_SMALL_POOL * 2u;
  }
  const int k = ga.k;
  const uint32_t n = (uint32_t)m.n;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const uint64_t nchunks = (nq + GI_SMALL_CHUNK - 1) / GI_SMALL_CHUNK;
  // a warp takes RANGES of up to 8 consecutive chunks (neighbours on the Morton curve; fewer in small launches, for balance): the k-th distance of one centre, widened,
  // is the first search bound of the next -- the brute-force pass then keeps ~1.3 k candidates and needs one cut
  double guess2 = ga.r2;
  unsigned long long chunk = 0, range_end = 0;
  while (true) {
    if (chunk >= range_end) {
      if (lane == 0) chunk = atomicAdd(work_counter, (unsigned long long)ga.range);
      chunk = __shfl_sync(FULL, chunk, 0);
      range_end = chunk + (unsigned long long)ga.range;
      guess2 = ga.r2;
    }
    if (chunk >= nchunks) break;
    const unsigned long long this_chunk = chunk++;
    const uint64_t q0 = this_chunk * GI_SMALL_CHUNK;
    const int nqc = (int)(nq - q0 < GI_SMALL_CHUNK ? nq - q0 : GI_SMALL_CHUNK);
    // lane i holds query i of the chunk
    uint64_t my_qi = 0; double mx = 0.0, my = 0.0, mz = 0.0;
    if (lane < nqc) {
      my_qi = order ? (uint64_t)order[q0 + lane] : q0 + lane;
      const double *qp = (const double *)(queries + my_qi);
      mx = qp[0]; my = qp[1]; mz = qp[2];
    }
    // centre of the chunk: the centroid of its queries (any point would do; this one keeps the chunk radius small)
    double cx = lane < nqc ? mx : 0.0, cy = lane < nqc ? my : 0.0, cz = lane < nqc ? mz : 0.0;
    for (int o = 16; o > 0; o >>= 1) { cx += __shfl_xor_sync(FULL, cx, o); cy += __shfl_xor_sync(FULL, cy, o); cz += __shfl_xor_sync(FULL, cz, o); }
    cx /= (double)nqc; cy /= (double)nqc; cz /= (double)nqc;
    double my_rho2 = 0.0;
    if (lane < nqc) { const double dx = mx - cx, dy = my - cy, dz = mz - cz; my_rho2 = (dx * dx) + (dy * dy) + (dz * dz); }
    double rho2 = my_rho2;
    for (int o = 16; o > 0; o >>= 1) rho2 = fmax(rho2, __shfl_xor_sync(FULL, rho2, o));
    // ---- 1 + 2. ONE scan of the map (heap order; px / py / pz are L1-resident) finds the centre's exact k nearest -- first
    //      bound = the neighbouring chunk's k-th d2, widened -- and collects a SUPER-POOL: everything within
    //      sqrt(first bound) + 2 rho of the centre.  The pool proper (radius Rc + 2 rho, Rc the exact k-th distance) is then
    //      filtered out of the super-pool.  Only if the guess was too small, or the super-pool does not fit, the map is
    //      scanned again.
    int count = 0, raw = 0, npool = 0;
    double bound = fmin(ga.r2, guess2), Rc = 0.0, Rp2 = 0.0;
    bool centre_full = false, have_pool = false;
    const double rho = sqrt(rho2);
    while (true) {
      const double first = bound;
      const double Rs = (sqrt(first) + 2.0 * rho) * (1.0 + 1.0e-9), Rs2 = Rs * Rs;
      int nsuper = 0;
      count = 0;
      for (uint32_t base = 0; base < n; base += 32) {
        const uint32_t slot = base + (uint32_t)lane;
        bool pred = false, sp = false; double d2 = 0.0;
        if (slot < n) { const P3 p = {m.px[slot], m.py[slot], m.pz[slot]}; d2 = dist2(p, cx, cy, cz); pred = d2 <= bound; sp = d2 <= Rs2; }
        const unsigned smask = __ballot_sync(FULL, sp);
        if (smask) {
          const int at = nsuper + __popc(smask & lt_mask);
          if (sp && at < GI_SMALL_SUPER) sts_u16(a_super + 2u * (uint32_t)at, slot);
          nsuper += __popc(smask);
        }
        append(b, count, bound, k, lane, lt_mask, pred, d2, slot);
      }
      finish(b, count, raw, bound, k, lane, true);
      if (count >= k || first >= ga.r2) {
        centre_full = count >= k;
        Rc = sqrt(centre_full ? bound : ga.r2);
        guess2 = centre_full ? bound * 2.0 : ga.r2;
        const double Rp = (Rc + 2.0 * rho) * (1.0 + 1.0e-9);
        Rp2 = Rp * Rp;
        if (nsuper <= GI_SMALL_SUPER) {                       // the pool is a subset of the super-pool (Rc^2 <= first)
          __syncwarp();
          for (int base = 0; base < nsuper && npool <= GI_SMALL_POOL; base += 32) {
            const int i = base + lane;
            bool pred = false; uint32_t slot = 0;
            if (i < nsuper) { slot = lds_u16(a_super + 2u * (uint32_t)i); const P3 p = {m.px[slot], m.py[slot], m.pz[slot]}; pred = dist2(p, cx, cy, cz) <= Rp2; }
            const unsigned mask = __ballot_sync(FULL, pred);
            const int at = npool + __popc(mask & lt_mask);
            if (pred && at < GI_SMALL_POOL) sts_u16(a_pool + 2u * (uint32_t)at, slot);
            npool += __popc(mask);
          }
          have_pool = true;
        }
        break;
      }
      bound = ga.r2;                                          // the guess was too small: the full radius
    }
    if (!have_pool) {                                         // the pool by a scan of its own
      npool = 0;
      for (uint32_t base = 0; base < n && npool <= GI_SMALL_POOL; base += 32) {
        const uint32_t slot = base + (uint32_t)lane;
        bool pred = false;
        if (slot < n) { const P3 p = {m.px[slot], m.py[slot], m.pz[slot]}; pred = dist2(p, cx, cy, cz) <= Rp2; }
        const unsigned mask = __ballot_sync(FULL, pred);
        const int at = npool + __popc(mask & lt_mask);
        if (pred && at < GI_SMALL_POOL) sts_u16(a_pool + 2u * (uint32_t)at, slot);
        npool += __popc(mask);
      }
    }
    const bool pooled = npool <= GI_SMALL_POOL;
    if (lane == 0) {
      atomicAdd(&g_small_dbg[0], 1ull); if (!pooled) atomicAdd(&g_small_dbg[1], 1ull);
      atomicAdd(&g_small_dbg[3], (unsigned long long)npool); if (!centre_full) atomicAdd(&g_small_dbg[4], 1ull);
    }
    __syncwarp();
    // ---- 3. pooled chunks are ranked by small_rank_kernel (a lane per query); the others right here, a warp per query
    if (!idx_out && pooled) {
      if (lane == 0) { SmallChunk ci; ci.cx = cx; ci.cy = cy; ci.cz = cz; ci.Rc = Rc; ci.npool = npool; ci.centre_full = centre_full ? 1 : 0; ci.pad = 0.0; info[this_chunk] = ci; }
      for (int i = lane; i < npool; i += 32) pools[this_chunk * GI_SMALL_POOL + i] = (uint16_t)lds_u16(a_pool + 2u * (uint32_t)i);
      __syncwarp();
      continue;
    }
    if (!idx_out && lane == 0) info[this_chunk].npool = -1;
    for (int qq = 0; qq < nqc; qq++) {
      const uint64_t qi = __shfl_sync(FULL, my_qi, qq);
      const double qx = __shfl_sync(FULL, mx, qq), qy = __shfl_sync(FULL, my, qq), qz = __shfl_sync(FULL, mz, qq);
      const double *qp = (const double *)(queries + qi);
      const double mine = lane < 17 ? qp[lane] : 0.0;
      count = 0; raw = 0;
      bound = ga.r2;
      if (centre_full) {                                    // k photons lie within Rc of c, hence within Rc + |q - c| of q
        const double rq = (Rc + sqrt(__shfl_sync(FULL, my_rho2, qq))) * (1.0 + 1.0e-12);
        bound = fmin(ga.r2, rq * rq);
      }
      if (n > 0) {
        if (pooled) {
          for (int base = 0; base < npool; base += 32) {
            const int i = base + lane;
            bool pred = false; double d2 = 0.0; uint32_t slot = 0;
            if (i < npool) {
              slot = lds_u16(a_pool + 2u * (uint32_t)i);
              const P3 p = {m.px[slot], m.py[slot], m.pz[slot]};
              d2 = dist2(p, qx, qy, qz); pred = d2 <= bound;
            }
            append(b, count, bound, k, lane, lt_mask, pred, d2, slot);
          }
        } else {
          for (uint32_t base = 0; base < n; base += 32) {
            const uint32_t slot = base + (uint32_t)lane;
            bool pred = false; double d2 = 0.0;
            if (slot < n) { const P3 p = {m.px[slot], m.py[slot], m.pz[slot]}; d2 = dist2(p, qx, qy, qz); pred = d2 <= bound; }
            append(b, count, bound, k, lane, lt_mask, pred, d2, slot);
          }
        }
        finish(b, count, raw, bound, k, lane, idx_out != nullptr);
      }
      estimate_and_store(b, m, ga, mine, count, raw, bound, k, lane, qi, idx_out, count_out, rgb_out);
      __syncwarp();
    }
  }
}

__global__ void __launch_bounds__(GI_SMALL_WARPS * 32, 1)
small_rank_kernel(PhotonMapDev m, const GiGatherQuery *__restrict__ queries, const uint32_t *__restrict__ order, uint64_t nq, GatherArgs ga,
                  unsigned long long *work_counter, const SmallChunk *__restrict__ info, const uint16_t *__restrict__ pools, int32_t *count_out,
                  double *rgb_out) {
  extern __shared__ double smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t a_shell = (uint32_t)__cvta_generic_to_shared(smem) + (uint32_t)warp * GI_SMALL_RANK_BYTES;
  asm volatile("" : "+r"(a_shell));
  const uint32_t a_pos = a_shell + 32u * GI_SMALL_SHELL * 10u;             // x[POOL] y[POOL] z[POOL] of the pool's photons
  const uint32_t a_pool = a_pos + 24u * GI_SMALL_POOL;                      // their heap slots
  const uint64_t nchunks = (nq + GI_SMALL_CHUNK - 1) / GI_SMALL_CHUNK;
  while (true) {
    unsigned long long chunk = 0;
    if (lane == 0) chunk = atomicAdd(work_counter, 1ull);
    chunk = __shfl_sync(FULL, chunk, 0);
    if (chunk >= nchunks) break;
    const SmallChunk ci = info[chunk];
    if (ci.npool < 0) continue;
    for (int i = lane; i < ci.npool; i += 32) {
      const uint32_t slot = pools[chunk * GI_SMALL_POOL + i];
      sts_u16(a_pool + 2u * (uint32_t)i, slot);
      sts_f64(a_pos + 8u * (uint32_t)i, m.px[slot]); sts_f64(a_pos + 8u * (GI_SMALL_POOL + (uint32_t)i), m.py[slot]);
      sts_f64(a_pos + 8u * (2u * GI_SMALL_POOL + (uint32_t)i), m.pz[slot]);
    }
    const uint64_t q0 = chunk * GI_SMALL_CHUNK;
    const int nqc = (int)(nq - q0 < GI_SMALL_CHUNK ? nq - q0 : GI_SMALL_CHUNK);
    uint64_t my_qi = 0; double mx = 0.0, my = 0.0, mz = 0.0, my_rho2 = 0.0;
    if (lane < nqc) {
      my_qi = order ? (uint64_t)order[q0 + lane] : q0 + lane;
      const double *qp = (const double *)(queries + my_qi);
      mx = qp[0]; my = qp[1]; mz = qp[2];
      const double dx = mx - ci.cx, dy = my - ci.cy, dz = mz - ci.cz;
      my_rho2 = (dx * dx) + (dy * dy) + (dz * dz);                 // the same expression as in gather_small_kernel: the pool radius covers it
    }
    double bq = ga.r2, shell_lo = 0.0;
    if (ci.centre_full) {
      const double rq = sqrt(my_rho2);
      const double up = (ci.Rc + rq) * (1.0 + 1.0e-12);
      bq = fmin(ga.r2, up * up);
      if (ci.Rc > rq) { const double dn = (ci.Rc - rq) * (1.0 - 1.0e-12); shell_lo = dn * dn; }
    }
    __syncwarp();
    rank_chunk_per_thread(m, ga, a_pool, a_pos, a_shell, ci.npool, lane < nqc, queries, my_qi, mx, my, mz, bq, shell_lo, count_out, rgb_out);
    __syncwarp();
  }
}

// DEEP: bottom groups are 6 levels deep (m.hb == 6), a second round of 32 photons.
// STAGE: the position blocks of the bottom-row scan are staged in shared memory by cp.async (a two-slot ring per warp:
// the copy of the next touched group is in flight while the current one is tested) instead of being prefetched into
// 6-12 FP64 registers per lane.
// COUNT: the instrumented instantiation behind GI_GATHER_DEBUG=1 (searches per query, groups entered, photons tested /
// inside the bound): the figures DESIGN.md quotes come from it (profiles/r02_gather_counters.txt).
__device__ unsigned long long g_walk_dbg[8];
template <bool DEEP, bool STAGE, bool COUNT = false>
__global__ void __launch_bounds__(GI_GATHER_MAX_WARPS * 32, 1)
gather_kernel(PhotonMapDev m, const double *__restrict__ lut, const GiGatherQuery *__restrict__ queries,
              const uint32_t *__restrict__ order, uint64_t nq, GatherArgs ga, unsigned long long *work_counter, int32_t *idx_out,
              int32_t *count_out, double *rgb_out) {
  extern __shared__ double smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  WarpBuf b; b.cap = ga.cap; b.orig = m.orig;
  { const int room = ga.cap - (DEEP ? 64 : 32); b.limit = room < ga.k + ga.slack ? room : ga.k + ga.slack; }   // 32 (64) arrive at once
  {
    uint32_t base = (uint32_t)__cvta_generic_to_shared(smem) +
                    (uint32_t)warp * ((uint32_t)GI_WARP_BYTES(ga.cap, ga.key_rows) + (STAGE ? GI_STAGE_BYTES(DEEP) : 0u));
    asm volatile("" : "+r"(base));                          // opaque: keep it in a register instead of re-deriving it
    b.a_d2 = base; b.a_slot = base + (uint32_t)ga.cap * 8u; b.a_hist = b.a_slot + (uint32_t)ga.cap * 4u;
    b.a_keys = b.a_hist + GI_HIST_BINS * 4u;
  }
  // STAGE: this lane's words of ring slot 0 (x at +0, y at +256, z at +512; level 6 at +768 ...); slot 1 follows
  const uint32_t a_stage = b.a_keys + (uint32_t)ga.key_rows * 128u + 8u * (uint32_t)lane;
  const int k = ga.k;
  const uint64_t n = m.n;
  const int jl = 31 - __clz(lane + 1);                      // relative depth of this lane's photon inside a group
  const uint32_t lt_mask = (1u << lane) - 1u;
  const double *gpos_lane = m.gpos + lane, *cbox_lane = m.cbox + lane;
  const int G = m.groups, gp = m.groups - 2;                // rows; gp = the row whose children are the bottom groups
  const uint32_t bottom_first = (1u << m.glevel[G - 1]) - 1u;                                   // root slot of the first bottom group
  const uint32_t bottom_delta = m.goff[G - 1] - bottom_first;                                   // bottom group index - root slot
  // Work distribution: queries cost very different amounts (a caustic map is dense in a few places), so chunks are
  // handed out dynamically -- but in RANGES of consecutive chunks per CTA, so that the warps of an SM stay on
  // neighbouring queries (L1).  A warp takes a ticket from the CTA's counter; the warp whose ticket opens a new range
  // fetches it from the launch-wide counter and publishes it (generation-tagged slot), the others wait that moment.
  __shared__ unsigned int s_ticket;
  __shared__ volatile unsigned int s_gen[8];
  __shared__ volatile unsigned long long s_base[8];
  if (threadIdx.x == 0) s_ticket = 0;
  if (threadIdx.x < 8) s_gen[threadIdx.x] = 0;
  __syncthreads();
  const unsigned int range = (unsigned int)ga.range;
  const uint64_t nchunks = (nq + GI_GATHER_CHUNK - 1) / GI_GATHER_CHUNK;
  // Search radius carried from one query to the next: consecutive queries of a chunk are neighbouring pixels, so the
  // previous k-th distance (x guess_factor in area) is tried first and widened only if fewer than k photons are inside.
  // A warp's very first query has no predecessor: it starts from what the last launch on this map left behind
  // (m.hint, scaled to this estimate size, x4 for safety) -- without it every launch pays one search of the full
  // radius per warp, 10 % of a 10^6-query launch.
  double guess = ga.r2;
  if (m.hint) { const double hk = m.hint[0], hg = m.hint[1]; if (hk > 0.0 && hg > 0.0) guess = fmin(ga.r2, hg * ((double)k / hk) * 4.0); }
  while (true) {
  unsigned long long chunk = 0;
  if (lane == 0) {
    const unsigned int t = atomicAdd(&s_ticket, 1u);
    const unsigned int rn = t / range, off = t - rn * range, sl = rn & 7u;
    if (off == 0) {
      const unsigned long long base = atomicAdd(work_counter, (unsigned long long)range);
      s_base[sl] = base; __threadfence_block(); s_gen[sl] = rn + 1u;
      chunk = base;
    } else {
      // brief: the opener already holds ticket rn * range.  Exactly this range's generation: launch_gather keeps
      // range >= warps, so at most two ranges are open at a time and the 8 slots never alias
      while (s_gen[sl] != rn + 1u) { }
      chunk = s_base[sl] + off;
    }
  }
  chunk = __shfl_sync(FULL, chunk, 0);
  if (chunk >= nchunks) break;
  for (uint64_t qs = chunk * GI_GATHER_CHUNK; qs < nq && qs < (chunk + 1) * GI_GATHER_CHUNK; qs++) {
    // ---- query record: 17 doubles, one per lane, broadcast by shuffle.  order[] (optional) lists the queries along a
    //      Morton curve, so that the queries of a warp and of a CTA are neighbours in space; results go to slot qi.
    const uint64_t qi = order ? (uint64_t)order[qs] : qs;
    const double *qp = (const double *)(queries + qi);
    const double mine = lane < 17 ? qp[lane] : 0.0;
    const double qx = __shfl_sync(FULL, mine, 0), qy = __shfl_sync(FULL, mine, 1), qz = __shfl_sync(FULL, mine, 2);
    int count = 0, raw = 0;
    double bound = ga.r2;
    unsigned dbg_search = 0, dbg_upper = 0, dbg_bottom = 0, dbg_tested = 0, dbg_inside = 0, dbg_cuts = 0;
    if (n > 0) {
      double init = fmin(ga.r2, guess);
      while (true) {
        count = 0; bound = init;
        if (COUNT) dbg_search++;
        // The tree is walked as a 32-ary tree of groups (photon_dev.h), depth first.
        //   rows above gp: row g keeps the root of the group it is in (st_root, held by lane g) and, per lane, the
        //     lower bound of the squared distance from the query to child group `lane` (keys[g][lane], float bits
        //     rounded down, NONE once visited or culled); children are entered nearest first and re-tested against
        //     the bound, which shrinks when the candidate buffer is cut.
        //   row gp: the children are the bottom groups.  Those whose box touches the ball are scanned in slot order
        //     (= space order), the photons of the next one being loaded while the current one is tested.
        uint32_t st_root = 0;
        int g = 0;
        uint32_t s = 0;
        bool enter = true;
        if (G == 1) {                                                   // fewer than 64 photons: one group, no boxes
          const uint32_t s1 = (uint32_t)lane, s2 = 31u + (uint32_t)lane;
          const P3 a = load_group(gpos_lane, 0);
          double d2 = dist2(a, qx, qy, qz);
          append(b, count, bound, k, lane, lt_mask, d2 <= bound, d2, s1);
          if (DEEP) {
            const P3 c = load_group(gpos_lane, m.x6_base);
            d2 = dist2(c, qx, qy, qz); append(b, count, bound, k, lane, lt_mask, d2 <= bound, d2, s2);
          }
          g = -1;
        }
        while (g >= 0) {
          const int h = g == 0 ? m.h0 : GI_GROUP_LEVELS;
          if (enter) {
            // ---- the group's photons and the boxes of its 32 children (both blocks are loaded before either is used)
            const uint32_t cbase = ((s + 1) << h) - 1;
            const uint32_t gi = s + (m.goff[g] - ((1u << m.glevel[g]) - 1u));
            const P3 ph = load_group(gpos_lane, gi);
            const double bd = box_dist2(cbox_lane, gi, qx, qy, qz);
            {
              const double d2 = dist2(ph, qx, qy, qz);
              if (COUNT) { dbg_upper++; dbg_tested += __popc(__ballot_sync(FULL, d2 < INFINITY)); dbg_inside += __popc(__ballot_sync(FULL, d2 <= bound)); }
              append(b, count, bound, k, lane, lt_mask, d2 <= bound, d2, (s << jl) + (uint32_t)lane);
            }
            if (g == gp) {
              // ---- bottom groups under this one: branch-light loop, the next group's block is always in flight
              const float kf = __double2float_rd(bd);
              unsigned todo = __ballot_sync(FULL, (double)kf <= bound);
              if (todo && STAGE) {
                // ---- staged variant: the blocks go global -> shared by cp.async, one group ahead
                uint32_t ls = cbase + (uint32_t)(__ffs(todo) - 1); todo &= todo - 1;
                bool has6 = DEEP && ls - bottom_first < m.x6_blocks;
                uint32_t slot_a = a_stage;
                {
                  const double *p = gpos_lane + (size_t)(ls + bottom_delta) * 96;
                  cp_async8(slot_a, p); cp_async8(slot_a + 256u, p + 32); cp_async8(slot_a + 512u, p + 64);
                  if (has6) {
                    const double *q6 = gpos_lane + (size_t)(m.x6_base + (ls - bottom_first)) * 96;
                    cp_async8(slot_a + 768u, q6); cp_async8(slot_a + 1024u, q6 + 32); cp_async8(slot_a + 1280u, q6 + 64);
                  }
                  cp_async_commit();
                }
                while (true) {
                  const bool more = todo != 0;
                  const uint32_t ls_n = cbase + (more ? (uint32_t)(__ffs(todo) - 1) : 0u);
                  todo &= todo - 1;
                  const bool has6_n = DEEP && ls_n - bottom_first < m.x6_blocks;
                  const uint32_t other = a_stage + (slot_a == a_stage ? GI_STAGE_SLOT_BYTES(DEEP) : 0u);
                  if (more) {
                    const double *p = gpos_lane + (size_t)(ls_n + bottom_delta) * 96;
                    cp_async8(other, p); cp_async8(other + 256u, p + 32); cp_async8(other + 512u, p + 64);
                    if (has6_n) {
                      const double *q6 = gpos_lane + (size_t)(m.x6_base + (ls_n - bottom_first)) * 96;
                      cp_async8(other + 768u, q6); cp_async8(other + 1024u, q6 + 32); cp_async8(other + 1280u, q6 + 64);
                    }
                  }
                  cp_async_commit();
                  cp_async_wait<1>();                                   // everything but the group just committed has landed
                  P3 cur; cur.x = lds_f64(slot_a); cur.y = lds_f64(slot_a + 256u); cur.z = lds_f64(slot_a + 512u);
                  const double d2a = dist2(cur, qx, qy, qz);
                  double d2b = 0.0;
                  const bool pa = d2a <= bound; bool pb = false;
                  unsigned mb = 0u;
                  if (has6) {
                    P3 low; low.x = lds_f64(slot_a + 768u); low.y = lds_f64(slot_a + 1024u); low.z = lds_f64(slot_a + 1280u);
                    d2b = dist2(low, qx, qy, qz); pb = d2b <= bound; mb = __ballot_sync(FULL, pb);
                  }
                  const unsigned ma = __ballot_sync(FULL, pa);
                  if (ma | mb) {
                    const int na = __popc(ma);
                    if (pa) { const int pos = count + __popc(ma & lt_mask); b.set_d2(pos, d2a); b.set_slot(pos, (ls << jl) + (uint32_t)lane); }
                    if (pb) {
                      const int pos = count + na + __popc(mb & lt_mask);
                      b.set_d2(pos, d2b); b.set_slot(pos, (((ls + 1) << GI_GROUP_LEVELS) - 1) + (uint32_t)lane);
                    }
                    count += na + __popc(mb);
                    __syncwarp();
                    if (count > b.limit) {
                      const Sel r = cut(b, count, bound, k, lane, true); count = r.count; bound = r.bound;
                      todo &= __ballot_sync(FULL, (double)kf <= bound);
                    }
                  }
                  if (!more) break;
                  ls = ls_n; has6 = has6_n; slot_a = other;
                }
                cp_async_wait<0>();
              } else if (todo) {
                uint32_t ls = cbase + (uint32_t)(__ffs(todo) - 1); todo &= todo - 1;
                bool has6 = DEEP && ls - bottom_first < m.x6_blocks;                            // uniform: the group has a level 6
                P3 cur = load_group(gpos_lane, ls + bottom_delta);
                P3 low = {0.0, 0.0, 0.0};
                if (has6) low = load_group(gpos_lane, m.x6_base + (ls - bottom_first));
                while (true) {
                  const bool more = todo != 0;
                  const uint32_t ls_n = cbase + (more ? (uint32_t)(__ffs(todo) - 1) : 0u);      // child 0 always exists
                  todo &= todo - 1;
                  const bool has6_n = DEEP && ls_n - bottom_first < m.x6_blocks;
                  const P3 nxt = load_group(gpos_lane, ls_n + bottom_delta);
                  P3 low_n = {0.0, 0.0, 0.0};
                  if (has6_n) low_n = load_group(gpos_lane, m.x6_base + (ls_n - bottom_first));
                  const double d2a = dist2(cur, qx, qy, qz);
                  double d2b = 0.0;
                  const bool pa = d2a <= bound; bool pb = false;
                  unsigned mb = 0u;
                  if (has6) { d2b = dist2(low, qx, qy, qz); pb = d2b <= bound; mb = __ballot_sync(FULL, pb); }
                  const unsigned ma = __ballot_sync(FULL, pa);
                  if (COUNT) {
                    dbg_bottom++; dbg_tested += __popc(__ballot_sync(FULL, d2a < INFINITY)) + (has6 ? __popc(__ballot_sync(FULL, d2b < INFINITY)) : 0);
                    dbg_inside += __popc(ma) + __popc(mb);
                  }
                  if (ma | mb) {
                    const int na = __popc(ma);
                    if (pa) { const int pos = count + __popc(ma & lt_mask); b.set_d2(pos, d2a); b.set_slot(pos, (ls << jl) + (uint32_t)lane); }
                    if (pb) {
                      const int pos = count + na + __popc(mb & lt_mask);
                      b.set_d2(pos, d2b); b.set_slot(pos, (((ls + 1) << GI_GROUP_LEVELS) - 1) + (uint32_t)lane);
                    }
                    count += na + __popc(mb);
                    __syncwarp();
                    if (count > b.limit) {
                      if (COUNT) dbg_cuts++;
                      const Sel r = cut(b, count, bound, k, lane, true); count = r.count; bound = r.bound;
                      todo &= __ballot_sync(FULL, (double)kf <= bound);
                    }
                  }
                  if (!more) break;
                  ls = ls_n; cur = nxt; low = low_n; has6 = has6_n;
                }
              }
              --g; enter = false;
              continue;                                                 // h belongs to the row just left
            } else {
              sts_u32(b.a_keys + 4u * (uint32_t)(g * 32 + lane), bd <= bound ? __float_as_uint(__double2float_rd(bd)) : NONE);
              if (lane == g) st_root = s;
              enter = false;
            }
          }
          // ---- nearest not yet visited child of row g
          const uint32_t ka = b.a_keys + 4u * (uint32_t)(g * 32 + lane);
          const uint32_t key = lds_u32(ka);
          const uint32_t mn = __reduce_min_sync(FULL, key);
          if (mn == NONE || (double)__uint_as_float(mn) > bound) { --g; continue; }
          const int who = __ffs(__ballot_sync(FULL, key == mn)) - 1;
          if (lane == who) sts_u32(ka, NONE);
          const uint32_t root = __shfl_sync(FULL, st_root, g);
          s = (((root + 1) << h) - 1) + (uint32_t)who;
          g++; enter = true;
        }
        finish(b, count, raw, bound, k, lane, idx_out != nullptr);
        if (count >= k || init >= ga.r2) break;
        init = init > 0.0 ? fmin(ga.r2, init * 4.0) : ga.r2;            // the guess was too small: widen and retry
      }
      guess = count >= k ? bound * ga.guess_factor : ga.r2;
    }
    if (COUNT && lane == 0) {
      atomicAdd(&g_walk_dbg[0], 1ull); atomicAdd(&g_walk_dbg[1], (unsigned long long)dbg_search); atomicAdd(&g_walk_dbg[2], (unsigned long long)dbg_upper);
      atomicAdd(&g_walk_dbg[3], (unsigned long long)dbg_bottom); atomicAdd(&g_walk_dbg[4], (unsigned long long)dbg_tested);
      atomicAdd(&g_walk_dbg[5], (unsigned long long)dbg_inside); atomicAdd(&g_walk_dbg[6], (unsigned long long)dbg_cuts);
      atomicAdd(&g_walk_dbg[7], (unsigned long long)(count < k ? count : k));
    }
    estimate_and_store(b, m, ga, mine, count, raw, bound, k, lane, qi, idx_out, count_out, rgb_out);
    __syncwarp();
  }
  }   // chunks
  if (m.hint && blockIdx.x == 0 && threadIdx.x == 0 && guess < ga.r2) { m.hint[0] = (double)k; m.hint[1] = guess; }
}

// launch-wide chunk counter, one per stream slot of the context (pipe[0], pipe[1], any other stream)
static unsigned long long *work_counter(GiContext *ctx, cudaStream_t st) {
  if (!ctx->gather_counters && cudaMalloc(&ctx->gather_counters, 8 * sizeof(unsigned long long)) != cudaSuccess) {
    ctx->gather_counters = nullptr;
    return nullptr;
  }
  return (unsigned long long *)ctx->gather_counters + (st == ctx->pipe[0] ? 0 : (st == ctx->pipe[1] ? 1 : 2));
}

// One CTA per SM (persistent, warps take chunks of consecutive queries): the warps of an SM then work on
// neighbouring queries and share the kd-tree groups they read through L1.  As many warps as the candidate buffers
// leave room for: 24 at the default estimate sizes (132 KB of shared memory, the rest of the 256 KB stays L1).
cudaError_t launch_gather(GiContext *ctx, const PhotonMapDev &m, const double *d_lut, const GiGatherQuery *d_q,
                          const uint32_t *d_order, uint64_t nq, const GatherArgs &ga_in, int32_t *d_idx, int32_t *d_count, double *d_rgb,
                          cudaStream_t st) {
  GatherArgs ga = ga_in;
  // maps of a few thousand photons (the global map of a default frame): chunk-pooled brute force, no tree walk
  static const bool small_off = getenv("GI_GATHER_SMALL") && atoi(getenv("GI_GATHER_SMALL")) == 0;      // A/B hook
  // (the pool of 128 must hold the k nearest of every query of a chunk with room to spare: larger estimates walk the tree)
  if (m.n > 0 && m.n <= GI_SMALL_MAX_PHOTONS && ga.k * 5 <= GI_SMALL_POOL * 4 && !small_off) {
    const size_t per_warp_s = GI_WARP_BYTES(ga.cap, 0) + GI_SMALL_POOL * 2 + GI_SMALL_SUPER * 2;
    int warps_s = GI_SMALL_WARPS;
    while (warps_s > 4 && (size_t)warps_s * per_warp_s > 160 * 1024) warps_s -= 4;
    const size_t smem_s = (size_t)warps_s * per_warp_s;
    const size_t smem_r = (size_t)GI_SMALL_WARPS * GI_SMALL_RANK_BYTES;
    if (smem_s > ctx->gather_small_smem_configured) {
      cudaError_t e = cudaFuncSetAttribute(gather_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_s);
      if (e == cudaSuccess) e = cudaFuncSetAttribute(small_rank_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_r);
      if (e != cudaSuccess) return e;
      ctx->gather_small_smem_configured = smem_s;
    }
    unsigned long long *counter = work_counter(ctx, st);
    if (!counter) return cudaErrorMemoryAllocation;
    { cudaError_t e = cudaMemsetAsync(counter, 0, sizeof(unsigned long long), st); if (e != cudaSuccess) return e; }
    { cudaError_t e = cudaMemsetAsync(counter + 3, 0, sizeof(unsigned long long), st); if (e != cudaSuccess) return e; }
    const uint64_t nchunks_s = (nq + GI_SMALL_CHUNK - 1) / GI_SMALL_CHUNK;
    // pools and chunk records of this launch: grow-only scratch per stream slot (like the ordering scratch)
    const int sslot = st == ctx->pipe[0] ? 0 : (st == ctx->pipe[1] ? 1 : 2);
    SmallChunk *info = nullptr; uint16_t *pools = nullptr;
    if (!d_idx) {
      GiScratch &ws = ctx->small_ws[sslot];
      const size_t info_bytes = ((sizeof(SmallChunk) * nchunks_s + 255) / 256) * 256;
      if (!ws.fit(info_bytes + (size_t)nchunks_s * GI_SMALL_POOL * 2)) return cudaErrorMemoryAllocation;
      info = (SmallChunk *)ws.p; pools = (uint16_t *)((unsigned char *)ws.p + info_bytes);
    }
    const uint64_t want_s = (nchunks_s + warps_s - 1) / warps_s, cap_s = (uint64_t)ctx->sm_count;
    { const uint64_t r = nchunks_s / ((uint64_t)ctx->sm_count * warps_s * 8); ga.range = (int)(r < 1 ? 1 : (r > 8 ? 8 : r)); }
    gather_small_kernel<<<(unsigned)(want_s < cap_s ? want_s : cap_s), warps_s * 32, smem_s, st>>>(m, d_q, d_order, nq, ga, counter, info, pools, d_idx, d_count,
                                                                                                    d_rgb);
    ctx->launches++;
    if (!d_idx) {
      const uint64_t want_r = (nchunks_s + GI_SMALL_WARPS - 1) / GI_SMALL_WARPS;
      small_rank_kernel<<<(unsigned)(want_r < cap_s ? want_r : cap_s), GI_SMALL_WARPS * 32, smem_r, st>>>(m, d_q, d_order, nq, ga, counter + 3, info, pools, d_count,
                                                                                                         d_rgb);
      ctx->launches++;
    }
    static const bool dbg = getenv("GI_GATHER_DEBUG") && atoi(getenv("GI_GATHER_DEBUG")) != 0;
    if (dbg) {
      unsigned long long h[8] = {0}, z[8] = {0};
      cudaStreamSynchronize(st);
      cudaMemcpyFromSymbol(h, g_small_dbg, sizeof(h)); cudaMemcpyToSymbol(g_small_dbg, z, sizeof(z));
      fprintf(stderr, "[gather_small] queries %llu chunks %llu pool overflow %llu slow path %llu mean pool %.1f centre<k %llu\n", (unsigned long long)nq, h[0], h[1],
              h[2], h[0] ? (double)h[3] / h[0] : 0.0, h[4]);
    }
    return cudaGetLastError();
  }
  ga.key_rows = m.groups > 2 ? m.groups - 2 : 0;
  const bool deep = m.hb > GI_GROUP_LEVELS;
  // GI_GATHER_STAGE=1: the cp.async-staged bottom-row scan (measured slower than register prefetch, see
  // docs/experiments/r02_gather_cp_async_stage.md; kept selectable)
  static const bool stage_env = getenv("GI_GATHER_STAGE") && atoi(getenv("GI_GATHER_STAGE")) != 0;
  const bool stage = stage_env && m.groups >= 2;
  const size_t per_warp = GI_WARP_BYTES(ga.cap, ga.key_rows) + (stage ? GI_STAGE_BYTES(deep) : 0u);
  int warps = (int)((227 * 1024) / per_warp);
  if (warps > GI_GATHER_MAX_WARPS) warps = GI_GATHER_MAX_WARPS;
  if (ga.max_warps > 0 && warps > ga.max_warps) warps = ga.max_warps;
  if (warps < 1) return cudaErrorInvalidValue;
  const uint64_t nchunks = (nq + GI_GATHER_CHUNK - 1) / GI_GATHER_CHUNK;
  if ((uint64_t)warps > nchunks) warps = (int)nchunks;
  const size_t smem = (size_t)warps * per_warp;
  // the attribute is per DEVICE (and per kernel instantiation): remembered in the context, which is bound to one device
  const int variant = (deep ? 1 : 0) + (stage ? 2 : 0);
  if (smem > ctx->gather_smem_configured[variant]) {
    cudaError_t e = cudaSuccess;
    switch (variant) {
      case 0: e = cudaFuncSetAttribute(gather_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); break;
      case 1: e = cudaFuncSetAttribute(gather_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); break;
      case 2: e = cudaFuncSetAttribute(gather_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); break;
      default: e = cudaFuncSetAttribute(gather_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); break;
    }
    if (e != cudaSuccess) return e;
    ctx->gather_smem_configured[variant] = smem;
  }
  // range of consecutive chunks a CTA fetches at a time: 4 per warp, fewer when the launch is small (balance)
  {
    uint64_t r = nchunks / ((uint64_t)ctx->sm_count * 8);
    r = r / warps * warps;
    if (r < (uint64_t)warps) r = warps;
    if (r > (uint64_t)4 * warps) r = (uint64_t)4 * warps;
    ga.range = (int)r;
    if (ga_in.range > 0) ga.range = ga_in.range < warps ? warps : ga_in.range;     // (tuning hook) never below one chunk per warp
  }
  unsigned long long *counter = work_counter(ctx, st);
  if (!counter) return cudaErrorMemoryAllocation;
  { cudaError_t e = cudaMemsetAsync(counter, 0, sizeof(unsigned long long), st); if (e != cudaSuccess) return e; }
  const uint64_t want = (nchunks + warps - 1) / warps;
  const unsigned grid = (unsigned)(want < (uint64_t)ctx->sm_count ? want : (uint64_t)ctx->sm_count);
  static const bool walk_dbg = getenv("GI_GATHER_DEBUG") && atoi(getenv("GI_GATHER_DEBUG")) != 0;
  if (walk_dbg && !stage) {
    // the instrumented instantiation (its shared-memory attribute is set on every launch: a debugging path)
    if (deep) { cudaFuncSetAttribute(gather_kernel<true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
                gather_kernel<true, false, true><<<grid, warps * 32, smem, st>>>(m, d_lut, d_q, d_order, nq, ga, counter, d_idx, d_count, d_rgb); }
    else { cudaFuncSetAttribute(gather_kernel<false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
           gather_kernel<false, false, true><<<grid, warps * 32, smem, st>>>(m, d_lut, d_q, d_order, nq, ga, counter, d_idx, d_count, d_rgb); }
    unsigned long long h[8] = {0}, z[8] = {0};
    cudaStreamSynchronize(st);
    cudaMemcpyFromSymbol(h, g_walk_dbg, sizeof(h)); cudaMemcpyToSymbol(g_walk_dbg, z, sizeof(z));
    const double q = h[0] ? (double)h[0] : 1.0;
    fprintf(stderr, "[gather_kernel] queries %llu photons %llu k %d | per query: searches %.3f upper groups %.2f bottom groups %.2f photons tested %.1f "
            "inside the bound when tested %.1f kept %.1f mid-scan cuts %.3f\n", h[0], (unsigned long long)m.n, ga.k, h[1] / q, h[2] / q, h[3] / q, h[4] / q, h[5] / q,
            h[7] / q, h[6] / q);
    ctx->launches++;
    return cudaGetLastError();
  }
  switch (variant) {
    case 0: gather_kernel<false, false><<<grid, warps * 32, smem, st>>>(m, d_lut, d_q, d_order, nq, ga, counter, d_idx, d_count, d_rgb); break;
    case 1: gather_kernel<true, false><<<grid, warps * 32, smem, st>>>(m, d_lut, d_q, d_order, nq, ga, counter, d_idx, d_count, d_rgb); break;
    case 2: gather_kernel<false, true><<<grid, warps * 32, smem, st>>>(m, d_lut, d_q, d_order, nq, ga, counter, d_idx, d_count, d_rgb); break;
    default: gather_kernel<true, true><<<grid, warps * 32, smem, st>>>(m, d_lut, d_q, d_order, nq, ga, counter, d_idx, d_count, d_rgb); break;
  }
  ctx->launches++;
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ no-allocation variant
// gi_gather_batch_device_noalloc (gi_b200.h): the same kernel, with every piece of scratch memory in one file-scope device
// array instead of cudaMalloc'd buffers of the context.  Layout (offsets in bytes, Q = GI_NOALLOC_MAX_QUERIES):
//   [0, 256)            launch-wide chunk counter (u64) -- the only part that must start at zero: cleared per call
//   [256, 256 + 16 Q)   Morton keys in / out, permutation in / out (4 x u32[Q]) -- fully rewritten by every call
//   [.., + 1 MiB)       temporary storage of the radix sort (size checked against what the sort asks for)
// In HOST code `g_noalloc_scratch` is a host-side stand-in, not a device address: the address the kernels get comes from
// cudaGetSymbolAddress, taken once per context with the context's device current (noalloc_base).
#define GI_NOALLOC_SORT_TEMP (1u << 20)
#define GI_NOALLOC_BYTES (256u + 16u * GI_NOALLOC_MAX_QUERIES + GI_NOALLOC_SORT_TEMP)
static __device__ __align__(256) unsigned char g_noalloc_scratch[GI_NOALLOC_BYTES];

static unsigned char *noalloc_base(GiContext *ctx) {
  if (ctx->noalloc_scratch) return ctx->noalloc_scratch;
  unsigned char *p = nullptr;
  if (!cuda_ok(cudaGetSymbolAddress((void **)&p, g_noalloc_scratch), "cudaGetSymbolAddress(gather scratch)") || !p) return nullptr;
  ctx->noalloc_scratch = p;
  return p;
}

// 30-bit Morton key of a query position inside the map's bounding box (same keys as query_order.cu)
__device__ __forceinline__ uint32_t na_spread10(uint32_t v) {
  v &= 1023u; v = (v | (v << 16)) & 0x030000FFu; v = (v | (v << 8)) & 0x0300F00Fu; v = (v | (v << 4)) & 0x030C30C3u;
  return (v | (v << 2)) & 0x09249249u;
}
__global__ void __launch_bounds__(256) noalloc_keys_kernel(const GiGatherQuery *__restrict__ q, uint32_t n, double lx, double ly, double lz,
                                                           double ix, double iy, double iz, uint32_t *__restrict__ keys,
                                                           uint32_t *__restrict__ order) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const double x = (q[t].p[0] - lx) * ix, y = (q[t].p[1] - ly) * iy, z = (q[t].p[2] - lz) * iz;
  const uint32_t xi = (uint32_t)fmin(fmax(x * 1024.0, 0.0), 1023.0), yi = (uint32_t)fmin(fmax(y * 1024.0, 0.0), 1023.0),
                 zi = (uint32_t)fmin(fmax(z * 1024.0, 0.0), 1023.0);
  keys[t] = na_spread10(xi) | (na_spread10(yi) << 1) | (na_spread10(zi) << 2);
  order[t] = t;
}

// launch_gather with the chunk counter supplied by the caller (no work_counter(): that one allocates on first use)
static cudaError_t launch_gather_counter(GiContext *ctx, const PhotonMapDev &m, const double *d_lut, const GiGatherQuery *d_q,
                                         const uint32_t *d_order, uint64_t nq, const GatherArgs &ga_in, unsigned long long *counter,
                                         int32_t *d_idx, int32_t *d_count, double *d_rgb, cudaStream_t st) {
  GatherArgs ga = ga_in;
  ga.key_rows = m.groups > 2 ? m.groups - 2 : 0;
  const size_t per_warp = GI_WARP_BYTES(ga.cap, ga.key_rows);
  int warps = (int)((227 * 1024) / per_warp);
  if (warps > GI_GATHER_MAX_WARPS) warps = GI_GATHER_MAX_WARPS;
  if (warps < 1) return cudaErrorInvalidValue;
  const uint64_t nchunks = (nq + GI_GATHER_CHUNK - 1) / GI_GATHER_CHUNK;
  if ((uint64_t)warps > nchunks) warps = (int)nchunks;
  const size_t smem = (size_t)warps * per_warp;
  const bool deep = m.hb > GI_GROUP_LEVELS;
  if (smem > ctx->gather_smem_configured[deep]) {
    cudaError_t e = deep ? cudaFuncSetAttribute(gather_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)
                         : cudaFuncSetAttribute(gather_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    ctx->gather_smem_configured[deep] = smem;
  }
  uint64_t r = nchunks / ((uint64_t)ctx->sm_count * 8);
  r = r / warps * warps;
  if (r < (uint64_t)warps) r = warps;
  if (r > (uint64_t)4 * warps) r = (uint64_t)4 * warps;
  ga.range = (int)r;
  { cudaError_t e = cudaMemsetAsync(counter, 0, sizeof(unsigned long long), st); if (e != cudaSuccess) return e; }
  const uint64_t want = (nchunks + warps - 1) / warps;
  const unsigned grid = (unsigned)(want < (uint64_t)ctx->sm_count ? want : (uint64_t)ctx->sm_count);
  if (deep) gather_kernel<true, false><<<grid, warps * 32, smem, st>>>(m, d_lut, d_q, d_order, nq, ga, counter, d_idx, d_count, d_rgb);
  else gather_kernel<false, false><<<grid, warps * 32, smem, st>>>(m, d_lut, d_q, d_order, nq, ga, counter, d_idx, d_count, d_rgb);
  ctx->launches++;
  return cudaGetLastError();
}

}  // namespace gi

using namespace gi;

extern "C" int gi_gather_batch_device_noalloc(GiContext *ctx, int map, const GiGatherQuery *d_queries, uint64_t n, int k, double r,
                                              int filter, const GiParams *params, int32_t *d_idx, int32_t *d_count, double *d_rgb,
                                              void *stream) {
  if (!ctx || map < 0 || map > 1) { set_error("gi_gather_batch_device_noalloc: bad argument"); return 1; }
  if (k < 1 || k > GI_MAX_K) { set_error("gi_gather_batch_device_noalloc: estimate size must be in [1, 512]"); return 1; }
  if (filter < 0 || filter > 2) { set_error("gi_gather_batch_device_noalloc: unknown filter"); return 1; }
  if (n > (uint64_t)GI_NOALLOC_MAX_QUERIES) {
    set_error("gi_gather_batch_device_noalloc: " + std::to_string(n) + " queries exceed the static scratch of GI_NOALLOC_MAX_QUERIES = " +
              std::to_string(GI_NOALLOC_MAX_QUERIES) + " (split the batch, or use gi_gather_batch_device)");
    return 1;
  }
  PhotonMapStore &m = ctx->maps[map];
  if (m.dev.n > 0 && !m.built) { set_error("gi_gather_batch_device_noalloc: photon map not built (call gi_build_maps)"); return 1; }
  if (n == 0) return 0;
  if (!d_queries) { set_error("gi_gather_batch_device_noalloc: null query buffer"); return 1; }
  // the direction table is made by gi_build_maps (an allocation this call must not make)
  if (!ctx->d_lut) { set_error("gi_gather_batch_device_noalloc: no direction table on this context yet (call gi_build_maps once first)"); return 1; }
  GI_CUDA(cudaSetDevice(ctx->device));
  unsigned char *base = noalloc_base(ctx);             // device address of g_noalloc_scratch on ctx->device
  if (!base) return 1;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned long long *counter = (unsigned long long *)base;
  uint32_t *keys = (uint32_t *)(base + 256), *keys2 = keys + GI_NOALLOC_MAX_QUERIES, *ord = keys2 + GI_NOALLOC_MAX_QUERIES,
           *ord2 = ord + GI_NOALLOC_MAX_QUERIES;
  void *sort_tmp = base + 256 + 16u * GI_NOALLOC_MAX_QUERIES;

  GatherArgs ga;                                       // as gi_gather_batch_device sets them (capi_photon.cu), without the tuning hooks
  ga.k = k; ga.filter = filter; ga.r2 = r * r;
  int p2 = 1; while (p2 < 2 * k) p2 <<= 1;
  ga.cap = d_idx ? (p2 < 256 ? 256 : p2) : ((k + (k / 2 > 96 ? k / 2 : 96) + 31) / 32) * 32 + 32;
  ga.fa = params ? params->filter_const_a : 0.918; ga.fb = params ? params->filter_const_b : 1.953;
  ga.fk = params ? params->filter_const_k : 1.0;
  ga.guess_factor = 1.1; ga.slack = ga.cap; ga.max_warps = 0; ga.key_rows = 0; ga.range = 0;

  const uint32_t *order = nullptr;
  if (n >= 4096) {                                     // walk the batch along a Morton curve, like gi_gather_batch_device
    size_t need = 0;
    GI_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, need, keys, keys2, ord, ord2, (int)n, 0, 30, st));    // size query only
    if (need > GI_NOALLOC_SORT_TEMP) { set_error("gi_gather_batch_device_noalloc: the sort needs more temporary storage than the static scratch holds"); return 1; }
    const double *b = m.bbox;
    const double ix = 1.0 / fmax(b[3] - b[0], 1e-300), iy = 1.0 / fmax(b[4] - b[1], 1e-300), iz = 1.0 / fmax(b[5] - b[2], 1e-300);
    noalloc_keys_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d_queries, (uint32_t)n, b[0], b[1], b[2], ix, iy, iz, keys, ord);
    GI_CUDA(cudaGetLastError());
    size_t bytes = GI_NOALLOC_SORT_TEMP;
    GI_CUDA(cub::DeviceRadixSort::SortPairs(sort_tmp, bytes, keys, keys2, ord, ord2, (int)n, 0, 30, st));   // caller-provided storage: no allocation
    ctx->launches += 2;
    order = ord2;
  }
  GI_CUDA(launch_gather_counter(ctx, m.dev, ctx->d_lut, d_queries, order, n, ga, counter, d_idx, d_count, d_rgb, st));
  return 0;
}
