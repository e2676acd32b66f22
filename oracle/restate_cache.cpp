// oracle/restate_cache.cpp -- TEST INFRASTRUCTURE: CPU restatement of the reference's irradiance cache (-cache).
// Checker for the CUDA path (csrc/irradiance_cache.cu), never the product.
//
//   pre-pass  src/photonmap.cpp:380-411 around EstimateIrradiance (src/utils/photon_utils.cpp:208-246):
//             colour_i <- RGBE( RGB(colour_i) + sum_{j in kNN(position_i)} RGB(colour_j) / (pi * R2) ),
//             R2 = r^2 if fewer than k photons lie within r, else max(1e-6, largest kNN d^2); every estimate reads
//             the ORIGINAL colours (the reference estimates into copies and overwrites afterwards, :387-407).
//   lookup    EstimateCachedRadiance (photon_utils.cpp:164-205): repeated R3Kdtree::FindClosest
//             (R3Kdtree.cpp:319-412,463-492: smallest d^2 with min^2 <= d^2 <= max^2) with
//             min = previous distance + 1e-6, until the photon's incident direction is on the viewer's side.
//
// Searches are restated as what they compute (brute-force scans).  Ties in d^2 go to the lower photon index (the
// reference's pick depends on its randomly pivoted tree).  Where no photon qualifies the reference dereferences NULL
// (photon_utils.cpp:181-183); the restatement returns black and index -1, which is what the CUDA path is held to.
// Pinned against the reference: tests/test_oracle_cache.py compares with `oracle/_ref/ref_harness cache` output
// stored in tests/golden/cache.npz.
#include <algorithm>
#include <thread>
#include "restate_common.h"

namespace oracle {
namespace {

void rgbe_to_rgb(const uint8_t *rgbe, double *rgb) {                     // graphics_utils.cpp:63-77
  if (!rgbe[3]) { rgb[0] = rgb[1] = rgb[2] = 0.0; return; }
  const double inverse = ldexp(1.0, ((int)rgbe[3]) - 128 - 8);
  rgb[0] = rgbe[0] * inverse; rgb[1] = rgbe[1] * inverse; rgb[2] = rgbe[2] * inverse;
}

void rgb_to_rgbe(const double *rgb, uint8_t *rgbe) {                     // graphics_utils.cpp:48-60
  double mx = 0;
  for (int i = 0; i < 3; i++) if (rgb[i] > mx) mx = rgb[i];
  int exponent;
  const double mantissa = frexp(mx, &exponent);
  if (!(mx > 0)) { rgbe[0] = rgbe[1] = rgbe[2] = 0; rgbe[3] = (uint8_t)(exponent + 128); return; }   // 0/0 in the reference
  for (int i = 0; i < 3; i++) rgbe[i] = (uint8_t)(int)(256.0 * rgb[i] / mx * mantissa);
  rgbe[3] = (uint8_t)(exponent + 128);
}

struct Lut {                                                             // photon_utils.cpp:253-272
  std::vector<double> v;
  Lut() : v(65536 * 3) {
    const double PI = 3.14159265358979323846, TWO_PI = 2.0 * PI;
    for (int phi = 0; phi < 256; phi++)
      for (int theta = 0; theta < 256; theta++) {
        const double tp = (phi * TWO_PI / 255.0) - PI, tt = (theta * PI / 255.0);
        const V3 n = normalized(V3{sin(tt) * cos(tp), sin(tt) * sin(tp), cos(tt)});
        double *o = &v[3 * (256 * phi + theta)];
        o[0] = n.x; o[1] = n.y; o[2] = n.z;
      }
  }
};

double dist2(const double *a, const double *b) { const V3 v = v3(a) - v3(b); return dot(v, v); }

void prepass_one(const GiPhoton *ph, uint64_t n, uint64_t i, int k, double r, uint8_t *out) {
  struct C { double d2; uint64_t j; };
  std::vector<C> c;
  const double r2 = r * r;
  for (uint64_t j = 0; j < n; j++) {
    const double d2 = dist2(ph[i].position, ph[j].position);
    if (d2 <= r2) c.push_back({d2, j});
  }
  auto less = [](const C &a, const C &b) { return a.d2 < b.d2 || (a.d2 == b.d2 && a.j < b.j); };
  if ((int)c.size() > k) { std::nth_element(c.begin(), c.begin() + k, c.end(), less); c.resize(k); }
  std::sort(c.begin(), c.end(), [](const C &a, const C &b) { return a.j < b.j; });
  double irr[3]; rgbe_to_rgb(ph[i].rgbe, irr);
  if (!c.empty()) {
    double max_d2 = EPS;
    if ((int)c.size() < k) max_d2 = r2;
    else for (auto &e : c) if (e.d2 > max_d2) max_d2 = e.d2;
    double est[3] = {0, 0, 0};
    for (auto &e : c) { double rgb[3]; rgbe_to_rgb(ph[e.j].rgbe, rgb); for (int ch = 0; ch < 3; ch++) est[ch] += rgb[ch]; }
    const double PI = 3.14159265358979323846;
    for (int ch = 0; ch < 3; ch++) irr[ch] += est[ch] / (PI * max_d2);
  }
  rgb_to_rgbe(irr, out);
}

void lookup_one(const GiPhoton *ph, uint64_t n, const Lut &lut, const double *q, double r, int32_t *idx_out, double *rgb_out) {
  const V3 normal = v3(q + 3), bounce = v3(q + 6);
  const double cos_theta = q[9];
  const double *kd = q + 10, *ks = q + 13; const double shininess = q[16];
  rgb_out[0] = rgb_out[1] = rgb_out[2] = 0.0; *idx_out = -1;
  double closest = 0.0;
  while (true) {
    const double mn = closest + EPS, min2 = mn * mn;
    double best = r * r; int64_t bi = -1;
    for (uint64_t j = 0; j < n; j++) {
      const double d2 = dist2(q, ph[j].position);
      if (d2 >= min2 && (bi < 0 ? d2 <= best : d2 < best)) { best = d2; bi = (int64_t)j; }
    }
    if (bi < 0) return;
    closest = sqrt(best);
    const double *d = &lut.v[3 * ph[bi].direction];
    const V3 inc = {d[0], d[1], d[2]};
    const double perp = dot(normal, inc);
    if ((cos_theta < 0 && perp < 0) || (cos_theta > 0 && perp > 0)) continue;
    double rgb[3]; rgbe_to_rgb(ph[bi].rgbe, rgb);
    double cos_alpha = dot(bounce, -inc);
    if (cos_alpha < 0) cos_alpha = 0;
    const double pw = pow(cos_alpha, shininess);
    for (int ch = 0; ch < 3; ch++) rgb_out[ch] = rgb[ch] * (fabs(perp) * kd[ch] + pw * ks[ch]);
    *idx_out = (int32_t)bi;
    return;
  }
}

template <class F> void parallel(uint64_t n, int threads, F f) {
  if (threads < 1) threads = 1;
  auto work = [&](int id) { for (uint64_t i = id; i < n; i += threads) f(i); };
  std::vector<std::thread> th;
  for (int i = 1; i < threads; i++) th.emplace_back(work, i);
  work(0);
  for (auto &t : th) t.join();
}

}  // namespace
}  // namespace oracle

// rgbe_out: n x 4 bytes, the photons' colours after the pre-pass
extern "C" int oracle_irradiance_prepass(const GiPhoton *photons, uint64_t n, int k, double r, uint8_t *rgbe_out, int threads) {
  oracle::parallel(n, threads, [&](uint64_t i) { oracle::prepass_one(photons, n, i, k, r, rgbe_out + 4 * i); });
  return 0;
}

// photons: the map AFTER the pre-pass.  idx_out[i] = chosen photon or -1, rgb_out[3i..] = what EstimateCachedRadiance adds
extern "C" int oracle_cached_radiance(const GiPhoton *photons, uint64_t n, const double *queries, uint64_t nq, double r,
                                      int32_t *idx_out, double *rgb_out, int threads) {
  static oracle::Lut lut;
  oracle::parallel(nq, threads, [&](uint64_t i) { oracle::lookup_one(photons, n, lut, queries + 17 * i, r, idx_out + i, rgb_out + 3 * i); });
  return 0;
}
