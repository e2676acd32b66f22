// csrc/irradiance_cache.cu -- the reference's irradiance cache (-cache).
//
// Build (MapPhotons, src/photonmap.cpp:380-411): after the kd-trees exist, every global photon's RGBE colour is
// replaced by   own power + sum of the powers of its GLOBAL_ESTIMATE_SIZE nearest photons / (pi * r^2)
// (EstimateIrradiance, src/utils/photon_utils.cpp:208-246), all estimates taken on the un-modified map.
// Lookup (EstimateCachedRadiance, photon_utils.cpp:164-205): the nearest photon (further than 1e-6) whose
// incident direction lies on the viewer's side of the surface, found by repeated FindClosest calls
// (R3Kdtree.cpp:319-412, 463-492) with min_distance = last distance + 1e-6; its colour times the Phong BRDF.
//
// Here the build is one launch of the kNN gather kernel at the photon positions (index output) plus a thread per
// photon that sums the neighbours' powers; the lookup is a thread per request walking the heap-ordered kd-tree
// with a register stack.  Two defined deviations from the reference: (1) exact distance ties go to the lower
// photon index (the reference's choice depends on its randomly pivoted tree); (2) when no photon qualifies inside
// GLOBAL_ESTIMATE_DIST the request contributes nothing -- the reference dereferences NULL there
// (photon_utils.cpp:181-183 reads closest_photon->direction before the NULL test).
#include <cuda_runtime.h>
#include <algorithm>
#include <chrono>
#include <cstdlib>
#include "context.h"
#include "irradiance_cache.h"
#include "shade.cuh"

extern "C" int gi_gather_batch_device(GiContext *ctx, int map, const GiGatherQuery *d_queries, uint64_t n, int k, double r,
                                      int filter, const GiParams *params, int32_t *d_idx, int32_t *d_count, double *d_rgb,
                                      void *stream);

namespace gi {

namespace {

__global__ void irradiance_queries_kernel(const GiPhoton *__restrict__ ph, uint64_t first, uint64_t n, GiGatherQuery *__restrict__ q) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  GiGatherQuery r = {};
  const GiPhoton p = ph[first + t];
  r.p[0] = p.position[0]; r.p[1] = p.position[1]; r.p[2] = p.position[2];
  q[t] = r;
}

// EstimateIrradiance for photon first + t from its kNN set (ascending photon indices, -1 padded)
__global__ void irradiance_kernel(const GiPhoton *__restrict__ ph, uint64_t first, uint64_t n, const int32_t *__restrict__ idx,
                                  const int32_t *__restrict__ count, int k, double r2, uint32_t *__restrict__ rgbe_out) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const GiPhoton self = ph[first + t];
  D3 irr = rgbe_decode(self.rgbe);                       // photonmap.cpp:390
  const int found = count[t];
  if (found > 0) {
    double max_d2 = 1.0e-6;                              // RN_EPSILON
    D3 est = mk3(0, 0, 0);
    for (int i = 0; i < found; i++) {
      const GiPhoton o = ph[idx[t * (uint64_t)k + i]];
      if (found >= k) {
        const double dx = self.position[0] - o.position[0], dy = self.position[1] - o.position[1], dz = self.position[2] - o.position[2];
        const double d2 = (dx * dx) + (dy * dy) + (dz * dz);
        if (d2 > max_d2) max_d2 = d2;
      }
      est = est + rgbe_decode(o.rgbe);
    }
    if (found < k) max_d2 = r2;
    const double PI = 3.14159265358979323846;
    const double den = PI * max_d2;
    est = mk3(est.x / den, est.y / den, est.z / den);
    irr = irr + est;
  }
  uint8_t e[4];
  rgbe_encode(irr, e);
  rgbe_out[first + t] = (uint32_t)e[0] | ((uint32_t)e[1] << 8) | ((uint32_t)e[2] << 16) | ((uint32_t)e[3] << 24);
}

// photonmap.cpp:400-407: the colours of the map's photons are overwritten in place (load order and heap order)
__global__ void apply_raw_kernel(GiPhoton *__restrict__ ph, uint64_t n, const uint32_t *__restrict__ rgbe) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const uint32_t v = rgbe[t];
  ph[t].rgbe[0] = (uint8_t)v; ph[t].rgbe[1] = (uint8_t)(v >> 8); ph[t].rgbe[2] = (uint8_t)(v >> 16); ph[t].rgbe[3] = (uint8_t)(v >> 24);
}
__global__ void apply_heap_kernel(uint2 *__restrict__ shade, double2 *__restrict__ srec, const uint32_t *__restrict__ orig, uint64_t n,
                                  const uint32_t *__restrict__ rgbe) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const uint32_t v = rgbe[orig[t]];
  shade[t].x = v;
  srec[2 * t + 1].y = __hiloint2double(__double2hiint(srec[2 * t + 1].y), (int)v);    // the estimate's copy (photon_dev.h)
}

// R3Kdtree::FindClosest(position, min_distance, max_distance): the photon with the smallest d^2 in
// [min2, max2]; ties: lower photon index.  Far subtrees are pruned on the float split plane widened by one ulp, so
// nothing the FP64 coordinates would admit is skipped.
GI_DEV bool nearest_beyond(const PhotonMapDev &m, double qx, double qy, double qz, double min2, double max2, uint32_t &slot_out,
                           double &d2_out) {
  const uint32_t n32 = (uint32_t)m.n, NONE = 0xffffffffu;
  uint32_t sk_node[32]; double sk_d[32];
  int sp = 0;
  double best = max2; uint32_t best_slot = NONE, best_orig = NONE;
  uint32_t s = 0;
  while (true) {
    while (s < n32) {
      const double dx = qx - m.px[s], dy = qy - m.py[s], dz = qz - m.pz[s];
      const double d2 = (dx * dx) + (dy * dy) + (dz * dz);
      if (d2 >= min2 && d2 <= best) {
        const uint32_t o = m.orig[s];
        if (d2 < best || best_slot == NONE || o < best_orig) { best = d2; best_slot = s; best_orig = o; }
      }
      const uint2 nd = m.node[s];
      const double sf = (double)__uint_as_float(nd.x);
      const int axis = (int)nd.y;
      const double wide = fabs(sf) * 1.1920929e-7 + 1.0e-44;
      const double qa = axis == 0 ? qx : (axis == 1 ? qy : qz);
      const bool left = qa < sf;
      const uint32_t near = 2 * s + (left ? 1 : 2), far = 2 * s + (left ? 2 : 1);
      double fd = left ? (sf - wide) - qa : qa - (sf + wide);
      fd = fd > 0 ? (fd * fd) * 0.99999999999999 : 0.0;
      if (far < n32 && fd <= best && sp < 32) { sk_node[sp] = far; sk_d[sp] = fd; sp++; }
      s = near;
    }
    bool got = false;
    while (sp > 0) { --sp; if (sk_d[sp] <= best) { s = sk_node[sp]; got = true; break; } }
    if (!got) break;
  }
  slot_out = best_slot; d2_out = best;
  return best_slot != NONE;
}

__global__ void __launch_bounds__(128) cached_radiance_kernel(PhotonMapDev m, const double *__restrict__ lut,
                                                              const GiGatherQuery *__restrict__ queries, uint64_t nq, double r,
                                                              double *__restrict__ rgb_out) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nq) return;
  const GiGatherQuery q = queries[t];
  D3 color = mk3(0, 0, 0);
  const D3 normal = mk3(q.n[0], q.n[1], q.n[2]);
  double closest_dist = 0.0;
  const double max2 = r * r;
  while (m.n > 0) {
    const double min_d = closest_dist + 1.0e-6;          // RN_EPSILON
    uint32_t slot; double d2;
    if (!nearest_beyond(m, q.p[0], q.p[1], q.p[2], min_d * min_d, max2, slot, d2)) break;
    closest_dist = sqrt(d2);
    const uint2 sh = m.shade[slot];
    const double *dir = lut + 3 * (size_t)(sh.y & 0xffffu);
    const D3 inc = mk3(dir[0], dir[1], dir[2]);
    const double perp = dot(normal, inc);
    if ((q.cos_theta < 0 && perp < 0) || (q.cos_theta > 0 && perp > 0)) continue;
    uint8_t e[4] = {(uint8_t)sh.x, (uint8_t)(sh.x >> 8), (uint8_t)(sh.x >> 16), (uint8_t)(sh.x >> 24)};
    const D3 pc = rgbe_decode(e);
    double cos_alpha = dot(mk3(q.bounce[0], q.bounce[1], q.bounce[2]), -inc);
    if (cos_alpha < 0) cos_alpha = 0;
    const bool has_spec = q.ks[0] != 0.0 || q.ks[1] != 0.0 || q.ks[2] != 0.0;
    const double pw = has_spec ? pow(cos_alpha, q.shininess) : 0.0;
    const double ap = fabs(perp);
    color = mk3(pc.x * (ap * q.kd[0] + pw * q.ks[0]), pc.y * (ap * q.kd[1] + pw * q.ks[1]), pc.z * (ap * q.kd[2] + pw * q.ks[2]));
    break;
  }
  rgb_out[3 * t] = color.x; rgb_out[3 * t + 1] = color.y; rgb_out[3 * t + 2] = color.z;
}

}  // namespace

int build_irradiance_cache(GiContext *ctx, const GiParams &p) {
  PhotonMapStore &m = ctx->maps[GI_MAP_GLOBAL];
  if (m.cached || m.dev.n == 0) return 0;
  const int k = p.global_estimate_size;
  if (k < 1 || k > GI_MAX_K) { set_error("gi_build_maps: estimate size must be in [1, 512]"); return 1; }
  const uint64_t n = m.dev.n;
  cudaStream_t st = ctx->stream;
  const auto t0 = std::chrono::steady_clock::now();
  const uint64_t chunk = std::min<uint64_t>(n, 1ull << 20);
  GiGatherQuery *d_q = nullptr; int32_t *d_idx = nullptr, *d_cnt = nullptr; uint32_t *d_rgbe = nullptr;
  auto cleanup = [&]() { cudaFree(d_q); cudaFree(d_idx); cudaFree(d_cnt); cudaFree(d_rgbe); };
  if (!cuda_ok(cudaMalloc(&d_q, sizeof(GiGatherQuery) * chunk), "cudaMalloc(cache queries)") ||
      !cuda_ok(cudaMalloc(&d_idx, sizeof(int32_t) * chunk * k), "cudaMalloc(cache idx)") ||
      !cuda_ok(cudaMalloc(&d_cnt, sizeof(int32_t) * chunk), "cudaMalloc(cache count)") ||
      !cuda_ok(cudaMalloc(&d_rgbe, sizeof(uint32_t) * n), "cudaMalloc(cache rgbe)")) { cleanup(); return 1; }
  const GiPhoton *raw = (const GiPhoton *)m.raw;
  for (uint64_t o = 0; o < n; o += chunk) {
    const uint64_t c = std::min(chunk, n - o);
    const unsigned blocks = (unsigned)((c + 255) / 256);
    irradiance_queries_kernel<<<blocks, 256, 0, st>>>(raw, o, c, d_q);
    if (gi_gather_batch_device(ctx, GI_MAP_GLOBAL, d_q, c, k, p.global_estimate_dist, GI_FILTER_DISK, &p, d_idx, d_cnt, nullptr, st)) {
      cleanup(); return 1;
    }
    irradiance_kernel<<<blocks, 256, 0, st>>>(raw, o, c, d_idx, d_cnt, k, p.global_estimate_dist * p.global_estimate_dist, d_rgbe);
    ctx->launches += 2;
  }
  const unsigned blocks = (unsigned)((n + 255) / 256);
  apply_raw_kernel<<<blocks, 256, 0, st>>>((GiPhoton *)m.raw, n, d_rgbe);
  apply_heap_kernel<<<blocks, 256, 0, st>>>((uint2 *)m.dev.shade, (double2 *)m.dev.srec, m.dev.orig, n, d_rgbe);
  ctx->launches += 2;
  const bool ok = cuda_ok(cudaGetLastError(), "irradiance cache") && cuda_ok(cudaStreamSynchronize(st), "irradiance cache sync");
  cleanup();
  if (!ok) return 1;
  m.cached = true;
  m.cache_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  return 0;
}

cudaError_t launch_cached_radiance(GiContext *ctx, const PhotonMapDev &m, const double *d_lut, const GiGatherQuery *d_q, uint64_t nq,
                                   double r, double *d_rgb, cudaStream_t st) {
  if (nq == 0) return cudaSuccess;
  cached_radiance_kernel<<<(unsigned)((nq + 127) / 128), 128, 0, st>>>(m, d_lut, d_q, nq, r, d_rgb);
  ctx->launches++;
  return cudaGetLastError();
}

}  // namespace gi
