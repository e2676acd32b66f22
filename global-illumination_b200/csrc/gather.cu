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
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include "context.h"
#include "gather.h"

namespace gi {

#define FULL 0xffffffffu
#define GI_GATHER_WARP_EXTRA (1024 + GI_MAX_GROUPS * 32 * 4)   /* per-warp shared memory besides the candidates: histogram + walk keys */

struct WarpBuf {
  double *d2; uint32_t *slot; uint32_t *hist;     // shared memory: candidates (d2, heap slot) + 256 histogram bins
  int limit;                                      // cut the buffer back to k once it holds more than this
  const uint32_t *orig;                           // global: heap slot -> photon index, read only to break exact d2 ties
  int cap;
};

__device__ __forceinline__ bool lex_less(double da, uint32_t oa, double db, uint32_t ob) {
  return da < db || (da == db && oa < ob);
}

// warp-cooperative bitonic sort of buf[0..cap) ascending by (d2, photon index); cap is a power of two.  Padding
// entries carry d2 = +inf (and slot 0), real entries finite d2, so the index lookup happens only on real ties.
__device__ void bitonic_sort(WarpBuf &b, int lane) {
  for (int k2 = 2; k2 <= b.cap; k2 <<= 1)
    for (int j = k2 >> 1; j > 0; j >>= 1) {
      for (int i = lane; i < b.cap; i += 32) {
        const int x = i ^ j;
        if (x > i) {
          const double di = b.d2[i], dx = b.d2[x];
          const uint32_t si = b.slot[i], sx = b.slot[x];
          const bool up = (i & k2) == 0;
          bool x_less_i = dx < di, i_less_x = di < dx;
          if (di == dx && di < INFINITY) { const uint32_t oi = b.orig[si], ox = b.orig[sx]; x_less_i = ox < oi; i_less_x = oi < ox; }
          if (up ? x_less_i : i_less_x) { b.d2[i] = dx; b.d2[x] = di; b.slot[i] = sx; b.slot[x] = si; }
        }
      }
      __syncwarp();
    }
}

// slow exact path (heavy ties only): full sort, keep the first k
__device__ void sort_select(WarpBuf &b, int &count, double &bound, int k, int lane) {
  for (int i = count + lane; i < b.cap; i += 32) { b.d2[i] = INFINITY; b.slot[i] = 0; }
  __syncwarp();
  bitonic_sort(b, lane);
  count = k;
  bound = fmin(bound, b.d2[k - 1]);
  __syncwarp();
}

// Exact cut of the candidate buffer to the k smallest by (d2, orig).  Requires count > k.
// A 256-bin histogram of d2 / bound locates the bin B that holds the k-th candidate; bins below B are kept whole,
// and inside B (a handful of candidates, since surface photons are uniform in d2) each lane ranks one candidate
// against the others by shuffles.  The bound becomes the largest kept d2.
__device__ void cut(WarpBuf &b, int &count, double &bound, int k, int lane) {
  if (!(bound > 0.0)) { sort_select(b, count, bound, k, lane); return; }
  for (int i = lane; i < 256; i += 32) b.hist[i] = 0;
  __syncwarp();
  const double scale = 256.0 / bound;
  for (int i = lane; i < count; i += 32) {
    int bin = (int)(b.d2[i] * scale); bin = bin > 255 ? 255 : bin;
    atomicAdd(&b.hist[bin], 1u);
  }
  __syncwarp();
  uint32_t h[8]; uint32_t s = 0;
  for (int j = 0; j < 8; j++) { h[j] = b.hist[lane * 8 + j]; s += h[j]; }
  uint32_t incl = s;
  for (int o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += v; }
  const uint32_t excl = incl - s;
  const unsigned crossing = __ballot_sync(FULL, incl >= (uint32_t)k);
  const int owner = __ffs(crossing) - 1;                  // count > k  =>  some lane crosses
  int B = 255, below = 0, m = 0;
  if (lane == owner) {
    uint32_t c = excl;
    for (int j = 0; j < 8; j++) { if (c + h[j] >= (uint32_t)k) { B = lane * 8 + j; below = (int)c; m = (int)h[j]; break; } c += h[j]; }
  }
  B = __shfl_sync(FULL, B, owner); below = __shfl_sync(FULL, below, owner); m = __shfl_sync(FULL, m, owner);
  const int need = k - below;                             // candidates to keep from bin B, 1 <= need <= m
  __syncwarp();
  if (need < m) {
    if (m > 32) { sort_select(b, count, bound, k, lane); return; }
    int c = 0;                                            // list the members of bin B in hist[0..m)
    for (int base = 0; base < count; base += 32) {
      const int i = base + lane;
      bool inb = false;
      if (i < count) { int bin = (int)(b.d2[i] * scale); bin = bin > 255 ? 255 : bin; inb = bin == B; }
      const unsigned mk = __ballot_sync(FULL, inb);
      if (inb) b.hist[c + __popc(mk & ((1u << lane) - 1u))] = (uint32_t)i;
      c += __popc(mk);
    }
    __syncwarp();
    double md = INFINITY; uint32_t mo = 0xffffffffu; int mi = -1;
    if (lane < m) { mi = (int)b.hist[lane]; md = b.d2[mi]; mo = b.orig[b.slot[mi]]; }
    int rank = 0;
    for (int l = 0; l < m; l++) {
      const double od = __shfl_sync(FULL, md, l); const uint32_t oo = __shfl_sync(FULL, mo, l);
      rank += lex_less(od, oo, md, mo) ? 1 : 0;
    }
    if (lane < m && rank >= need) b.d2[mi] = -1.0;        // dropped
    __syncwarp();
  }
  int newcount = 0; double mx = 0.0;
  for (int base = 0; base < count; base += 32) {
    const int i = base + lane;
    double d = 0.0; uint32_t sl = 0; bool keep = false;
    if (i < count) {
      d = b.d2[i]; sl = b.slot[i];
      int bin = (int)(d * scale); bin = bin > 255 ? 255 : bin;
      keep = d >= 0.0 && bin <= B;
    }
    const unsigned mk = __ballot_sync(FULL, keep);
    const int pos = newcount + __popc(mk & ((1u << lane) - 1u));
    __syncwarp();
    if (keep) { b.d2[pos] = d; b.slot[pos] = sl; mx = fmax(mx, d); }
    newcount += __popc(mk);
    __syncwarp();
  }
  for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(FULL, mx, o));
  count = newcount; bound = fmin(bound, mx);
}

// after the traversal: exactly the k nearest (or everything found, if fewer); bound = largest kept d2 when k are held
__device__ void finish(WarpBuf &b, int &count, double &bound, int k, int lane) {
  if (count > k) { cut(b, count, bound, k, lane); return; }
  if (count < k) return;
  double mx = 0.0;
  for (int i = lane; i < count; i += 32) mx = fmax(mx, b.d2[i]);
  for (int s = 16; s > 0; s >>= 1) mx = fmax(mx, __shfl_xor_sync(FULL, mx, s));
  bound = fmin(bound, mx);
}

__device__ __forceinline__ void append(WarpBuf &b, const PhotonMapDev &m, int &count, double &bound, int k, int lane,
                                       bool pred, double d2, uint32_t slot) {
  const unsigned mask = __ballot_sync(FULL, pred);
  if (mask == 0) return;
  const int pos = count + __popc(mask & ((1u << lane) - 1u));
  if (pred) { b.d2[pos] = d2; b.slot[pos] = slot; }
  count += __popc(mask);
  __syncwarp();
  if (count > b.limit) cut(b, count, bound, k, lane);
}

__device__ __forceinline__ double dist2(const PhotonMapDev &m, uint64_t slot, double qx, double qy, double qz) {
  const double dx = qx - m.px[slot], dy = qy - m.py[slot], dz = qz - m.pz[slot];   // v = query - photon; v.Dot(v)
  return (dx * dx) + (dy * dy) + (dz * dz);
}

template <int MINB>
__global__ void __launch_bounds__(GI_GATHER_WARPS * 32, MINB)
gather_kernel(PhotonMapDev m, const double *__restrict__ lut, const GiGatherQuery *__restrict__ queries, uint64_t nq,
              GatherArgs ga, int32_t *idx_out, int32_t *count_out, double *rgb_out) {
  extern __shared__ double smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  WarpBuf b; b.cap = ga.cap; b.orig = m.orig;
  b.limit = ga.cap - 32 < ga.k + ga.slack ? ga.cap - 32 : ga.k + ga.slack;
  {
    char *base = (char *)smem + (size_t)warp * ((size_t)ga.cap * 12 + GI_GATHER_WARP_EXTRA);
    b.d2 = (double *)base; b.slot = (uint32_t *)(base + (size_t)ga.cap * 8);
    b.hist = b.slot + ga.cap;
  }
  uint32_t *keys = b.hist + 256;                            // GI_MAX_GROUPS x 32, each lane reads only its own column
  const uint32_t NONE = 0xffffffffu;
  const int k = ga.k;
  const uint64_t n = m.n;
  const uint32_t n32 = (uint32_t)n;
  const int jl = 31 - __clz(lane + 1);                      // relative depth of this lane's photon inside a group
  const uint64_t warps_total = (uint64_t)gridDim.x * GI_GATHER_WARPS;
  const uint64_t nchunks = (nq + GI_GATHER_CHUNK - 1) / GI_GATHER_CHUNK;
  // Search radius carried from one query to the next: consecutive queries of a chunk are neighbouring pixels, so the
  // previous k-th distance (x guess_factor in area) is tried first and widened only if fewer than k photons are inside.
  double guess = ga.r2;
  for (uint64_t chunk = (uint64_t)blockIdx.x * GI_GATHER_WARPS + warp; chunk < nchunks; chunk += warps_total)
  for (uint64_t qi = chunk * GI_GATHER_CHUNK; qi < nq && qi < (chunk + 1) * GI_GATHER_CHUNK; qi++) {
    // ---- query record: 17 doubles, one per lane, broadcast by shuffle
    const double *qp = (const double *)(queries + qi);
    const double mine = lane < 17 ? qp[lane] : 0.0;
    const double qx = __shfl_sync(FULL, mine, 0), qy = __shfl_sync(FULL, mine, 1), qz = __shfl_sync(FULL, mine, 2);
    int count = 0;
    double bound = ga.r2;
    if (n > 0) {
      double init = fmin(ga.r2, guess);
      while (true) {
        count = 0; bound = init;
        // The tree is walked as a 32-ary tree of 5-level groups (photon_dev.h), depth first.  Row g of the walk keeps
        // the root of the group it is in (st_root, held by lane g) and, per lane, the lower bound of the squared
        // distance from the query to child group `lane` (keys[g][lane], float bits rounded down, NONE once visited
        // or culled); children are entered nearest first and re-tested against the bound, which shrinks as the
        // candidate buffer is cut.
        uint32_t st_root = 0;
        int g = 0;
        uint32_t s = 0;
        bool enter = true;
        while (true) {
          const int h = g == 0 ? m.h0 : GI_GROUP_LEVELS;
          if (enter) {
            // ---- the group's photons: relative depth j occupies slots (s << j) + e, e in [2^j - 1, 2^(j+1) - 1)
            {
              const uint32_t slot = (s << jl) + (uint32_t)lane;
              bool pred = false; double d2 = 0.0;
              if (lane < (1 << h) - 1 && slot < n32) { d2 = dist2(m, slot, qx, qy, qz); pred = d2 <= bound; }
              append(b, m, count, bound, k, lane, pred, d2, slot);
            }
            if (g + 1 >= m.groups) { if (--g < 0) break; enter = false; continue; }      // bottom row: back to the parent
            // ---- the 32 child groups: one box test per lane
            const uint32_t c = (((s + 1) << h) - 1) + (uint32_t)lane;
            uint32_t key = NONE;
            if (lane < (1 << h) && c < n32) {
              const double2 *bx = m.gbox + 3 * ((size_t)m.gbox_off[g + 1] + (c - ((1u << m.glevel[g + 1]) - 1u)));
              const double2 X = bx[0], Y = bx[1], Z = bx[2];
              const double dx = fmax(fmax(X.x - qx, qx - X.y), 0.0), dy = fmax(fmax(Y.x - qy, qy - Y.y), 0.0),
                           dz = fmax(fmax(Z.x - qz, qz - Z.y), 0.0);
              const double bd = ((dx * dx) + (dy * dy) + (dz * dz)) * 0.99999999999999;   // rounding must not over-estimate
              if (bd <= bound) key = __float_as_uint(__double2float_rd(bd));
            }
            keys[g * 32 + lane] = key;
            if (lane == g) st_root = s;
            enter = false;
          }
          // ---- nearest not yet visited child of row g
          const uint32_t key = keys[g * 32 + lane];
          const uint32_t mn = __reduce_min_sync(FULL, key);
          if (mn == NONE || (double)__uint_as_float(mn) > bound) { if (--g < 0) break; continue; }
          const int who = __ffs(__ballot_sync(FULL, key == mn)) - 1;
          if (lane == who) keys[g * 32 + lane] = NONE;
          const uint32_t root = __shfl_sync(FULL, st_root, g);
          s = (((root + 1) << h) - 1) + (uint32_t)who;
          g++; enter = true;
        }
        finish(b, count, bound, k, lane);
        if (count >= k || init >= ga.r2) break;
        init = fmin(ga.r2, init * 4.0);                                 // the guess was too small: widen and retry
      }
      guess = count >= k ? bound * ga.guess_factor : ga.r2;
    }
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
      for (int i = lane; i < found; i += 32) {
        const uint2 sh = m.shade[b.slot[i]];
        const double *dir = lut + 3 * (size_t)(sh.y & 0xffffu);
        const double ix = dir[0], iy = dir[1], iz = dir[2];
        const double perp = (nx * ix) + (ny * iy) + (nz * iz);
        if ((cos_theta < 0 && perp < 0) || (cos_theta > 0 && perp > 0)) continue;
        const uint32_t e = sh.x >> 24;
        if (e == 0) { if (ga.filter == GI_FILTER_GAUSS) {
            total_w += 1.0 - (1.0 - pow(c1, c2 * b.d2[i])) / (1.0 - c1); } continue; }
        const double inv = __longlong_as_double((long long)(e - 136 + 1023) << 52);   // ldexp(1.0, e - 128 - 8)
        double rgb[3] = {(double)(sh.x & 0xff) * inv, (double)((sh.x >> 8) & 0xff) * inv, (double)((sh.x >> 16) & 0xff) * inv};
        double cos_alpha = (bx * -ix) + (by * -iy) + (bz * -iz);
        if (cos_alpha < 0) cos_alpha = 0;
        const double pw = has_spec ? pow(cos_alpha, shin) : 0.0;
        const double ap = fabs(perp);
        for (int c = 0; c < 3; c++) rgb[c] *= ap * kd[c] + pw * ks[c];
        if (ga.filter == GI_FILTER_CONE) {
          const double w = 1.0 - c1 * sqrt(b.d2[i]);
          for (int c = 0; c < 3; c++) rgb[c] *= w;
        } else if (ga.filter == GI_FILTER_GAUSS) {
          const double w = 1.0 - (1.0 - pow(c1, c2 * b.d2[i])) / (1.0 - c1);
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
      for (int i = lane; i < b.cap; i += 32) b.d2[i] = i < found ? (double)b.orig[b.slot[i]] : INFINITY;
      __syncwarp();
      bitonic_sort(b, lane);
      for (int i = lane; i < k; i += 32) idx_out[qi * (uint64_t)k + i] = i < found ? (int32_t)b.d2[i] : -1;
    }
    __syncwarp();
  }
}

template <int MINB>
static cudaError_t launch_t(GiContext *ctx, const PhotonMapDev &m, const double *d_lut, const GiGatherQuery *d_q, uint64_t nq,
                            const GatherArgs &ga, int32_t *d_idx, int32_t *d_count, double *d_rgb, cudaStream_t st) {
  const size_t smem = (size_t)GI_GATHER_WARPS * ((size_t)ga.cap * 12 + GI_GATHER_WARP_EXTRA);
  static size_t configured = 0;
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(gather_kernel<MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    configured = smem;
  }
  int per_sm = 1;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, gather_kernel<MINB>, GI_GATHER_WARPS * 32, smem);
  if (per_sm < 1) per_sm = 1;
  const uint64_t nchunks = (nq + GI_GATHER_CHUNK - 1) / GI_GATHER_CHUNK;
  uint64_t want = (nchunks + GI_GATHER_WARPS - 1) / GI_GATHER_WARPS;
  uint64_t cap = (uint64_t)ctx->sm_count * per_sm;         // persistent: one resident wave, warps stride over chunks
  unsigned grid = (unsigned)(want < cap ? want : cap);
  gather_kernel<MINB><<<grid, GI_GATHER_WARPS * 32, smem, st>>>(m, d_lut, d_q, nq, ga, d_idx, d_count, d_rgb);
  ctx->launches++;
  return cudaGetLastError();
}

cudaError_t launch_gather(GiContext *ctx, const PhotonMapDev &m, const double *d_lut, const GiGatherQuery *d_q,
                          uint64_t nq, const GatherArgs &ga, int32_t *d_idx, int32_t *d_count, double *d_rgb,
                          cudaStream_t st) {
  if (ga.min_blocks >= 4) return launch_t<4>(ctx, m, d_lut, d_q, nq, ga, d_idx, d_count, d_rgb, st);
  return launch_t<3>(ctx, m, d_lut, d_q, nq, ga, d_idx, d_count, d_rgb, st);
}

}  // namespace gi
