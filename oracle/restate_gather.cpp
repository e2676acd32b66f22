// oracle/restate_gather.cpp -- TEST INFRASTRUCTURE: CPU restatement of the photon gather, i.e.
// EstimateRadiance (src/utils/photon_utils.cpp:72-162) on top of the k-nearest search
// R3Kdtree::FindClosestQuick (src/R3Shapes/R3Kdtree.cpp:688-848).  Checker for the CUDA path, never the product.
//
// The search is restated as what it computes, not how: the k photons with the smallest squared distance
// d^2 <= r^2 (inclusive, R3Kdtree.cpp:756-757), d^2 as R3SquaredDistance does it (R3Dist.cpp:41-46), found by a
// brute-force scan.  Exact d^2 ties at the k-th place are resolved towards the lower photon index (the
// reference's outcome there depends on its randomly pivoted tree; this is the canonical set the CUDA path
// is held to).  Pinned against the reference: tests/test_oracle_gather.py compares it with
// `oracle/_ref/ref_harness gather` (the reference's own R3Kdtree + EstimateRadiance) on reference-traced maps.
#include <algorithm>
#include <thread>
#include "restate_common.h"

namespace oracle {

// RGBE_to_RNRgb, src/utils/graphics_utils.cpp:63-77
static void rgbe_to_rgb(const uint8_t *rgbe, double *rgb) {
  if (!rgbe[3]) { rgb[0] = rgb[1] = rgb[2] = 0.0; return; }
  double inverse = ldexp(1.0, ((int)rgbe[3]) - 128 - 8);
  rgb[0] = rgbe[0] * inverse; rgb[1] = rgbe[1] * inverse; rgb[2] = rgbe[2] * inverse;
}

// BuildDirectionLookupTable, src/utils/photon_utils.cpp:253-272
struct DirLut {
  std::vector<double> x, y, z;
  DirLut() : x(65536), y(65536), z(65536) {
    const double PI = 3.14159265358979323846, TWO_PI = 2.0 * PI;        // RNScalar.cpp:31,35
    for (int phi = 0; phi < 256; phi++)
      for (int theta = 0; theta < 256; theta++) {
        double true_phi = (phi * TWO_PI / 255.0) - PI;
        double true_theta = (theta * PI / 255.0);
        V3 n = normalized(V3{sin(true_theta) * cos(true_phi), sin(true_theta) * sin(true_phi), cos(true_theta)});
        x[256 * phi + theta] = n.x; y[256 * phi + theta] = n.y; z[256 * phi + theta] = n.z;
      }
  }
};

struct Cand { double d2; int32_t idx; };

static void gather_one(const GiPhoton *ph, uint64_t n, const DirLut &lut, const double *q, int k, double r, int filter,
                       double fa, double fb, double fk, int32_t *idx_out, int32_t *count_out, double *rgb_out) {
  const V3 point = v3(q), normal = v3(q + 3), bounce = v3(q + 6);
  const double cos_theta = q[9];
  const double *kd = q + 10, *ks = q + 13; const double shininess = q[16];
  const double r2 = r * r;
  std::vector<Cand> c;
  for (uint64_t i = 0; i < n; i++) {
    V3 v = point - v3(ph[i].position);
    double d2 = dot(v, v);
    if (d2 >= 0.0 && d2 <= r2) c.push_back({d2, (int32_t)i});
  }
  auto less = [](const Cand &a, const Cand &b) { return a.d2 < b.d2 || (a.d2 == b.d2 && a.idx < b.idx); };
  if ((int)c.size() > k) { std::nth_element(c.begin(), c.begin() + k, c.end(), less); c.resize(k); }
  const int num = (int)c.size();
  if (count_out) *count_out = num;
  if (idx_out) {
    std::vector<int32_t> ids; for (auto &e : c) ids.push_back(e.idx);
    std::sort(ids.begin(), ids.end());
    for (int i = 0; i < k; i++) idx_out[i] = i < num ? ids[i] : -1;
  }
  double color[3] = {0, 0, 0};
  if (num > 0) {
    double max_d2 = EPS;
    if (num < k) max_d2 = r2;
    else for (auto &e : c) if (e.d2 > max_d2) max_d2 = e.d2;
    double est[3] = {0, 0, 0}, c1 = 1.0, c2 = 1.0, total_w = 0;
    const double PI = 3.14159265358979323846, E = 2.7182818284590452354;
    if (filter == GI_FILTER_CONE) c1 = 1.0 / (fk * sqrt(max_d2));
    else if (filter == GI_FILTER_GAUSS) { c1 = pow(E, -fb); c2 = 1.0 / (2.0 * max_d2); }
    for (auto &e : c) {
      const GiPhoton &p = ph[e.idx];
      V3 inc = {lut.x[p.direction], lut.y[p.direction], lut.z[p.direction]};
      double perp = dot(normal, inc);
      if ((cos_theta < 0 && perp < 0) || (cos_theta > 0 && perp > 0)) continue;
      double rgb[3]; rgbe_to_rgb(p.rgbe, rgb);
      double cos_alpha = dot(bounce, -inc);
      if (cos_alpha < 0) cos_alpha = 0;
      double pw = pow(cos_alpha, shininess);
      for (int ch = 0; ch < 3; ch++) rgb[ch] *= fabs(perp) * kd[ch] + pw * ks[ch];
      if (filter == GI_FILTER_CONE) {
        double w = 1.0 - c1 * sqrt(e.d2);
        for (int ch = 0; ch < 3; ch++) rgb[ch] *= w;
      } else if (filter == GI_FILTER_GAUSS) {
        double w = (1.0 - (1.0 - pow(c1, c2 * e.d2)) / (1.0 - c1));
        for (int ch = 0; ch < 3; ch++) rgb[ch] *= w;
        total_w += w;
      }
      for (int ch = 0; ch < 3; ch++) est[ch] += rgb[ch];
    }
    bool ok = true;
    if (filter == GI_FILTER_DISK && max_d2 > 0) { for (int ch = 0; ch < 3; ch++) est[ch] /= PI * max_d2; }
    else if (filter == GI_FILTER_CONE && max_d2 > 0) { for (int ch = 0; ch < 3; ch++) est[ch] /= (1.0 - 2.0 / 3.0 / fk) * PI * max_d2; }
    else if (filter == GI_FILTER_GAUSS && total_w > 0 && max_d2 > 0) {
      for (int ch = 0; ch < 3; ch++) est[ch] *= fa * (num / total_w) / (PI * max_d2);
    } else ok = false;
    if (ok) for (int ch = 0; ch < 3; ch++) color[ch] += est[ch];
  }
  if (rgb_out) { rgb_out[0] = color[0]; rgb_out[1] = color[1]; rgb_out[2] = color[2]; }
}

}  // namespace oracle

extern "C" int oracle_gather(const GiPhoton *photons, uint64_t n, const double *queries, uint64_t nq, int k, double r,
                             int filter, double fa, double fb, double fk, int32_t *idx_out, int32_t *count_out,
                             double *rgb_out, int threads) {
  static oracle::DirLut lut;
  if (threads < 1) threads = 1;
  auto work = [&](int id) {
    for (uint64_t i = id; i < nq; i += threads)
      oracle::gather_one(photons, n, lut, queries + 17 * i, k, r, filter, fa, fb, fk,
                         idx_out ? idx_out + i * (uint64_t)k : nullptr, count_out ? count_out + i : nullptr,
                         rgb_out ? rgb_out + 3 * i : nullptr);
  };
  std::vector<std::thread> th;
  for (int i = 1; i < threads; i++) th.emplace_back(work, i);
  work(0);
  for (auto &t : th) t.join();
  return 0;
}
