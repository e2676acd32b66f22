// tests/host_sim/isect_sim.cpp -- TEST INFRASTRUCTURE.  Compiles the DEVICE code of csrc/isect.cuh for the CPU
// (GI_HD makes it __host__ __device__) so that the traversal + exact-test + tie-rule logic can be checked
// against the oracle in this GPU-less container.  The BVH here is a simple CPU median-split builder that emits
// the same 64-byte node records as csrc/bvh_build.cu; it exists only for this simulation.  Never shipped,
// never used as a fallback: the product library has no CPU path.
#include <algorithm>
#include <cstdio>
#include <numeric>
#include <thread>
#include "isect.cuh"
#include "scene.h"
#include "scene_prep.h"

using namespace gi;

struct Sim {
  SceneData data; PreparedScene ps; std::vector<float4> bvh; DevScene dev;
};

static int build_rec(Sim &s, std::vector<int> &idx, int lo, int hi, float *box_out) {
  // returns child reference (>= 0 internal node, < 0 ~prim); writes the subtree box
  const float *pb = s.ps.prim_boxes.data();
  if (hi - lo == 1) { for (int k = 0; k < 6; k++) box_out[k] = pb[6 * (size_t)idx[lo] + k]; return ~idx[lo]; }
  float cb[6] = {1e30f, 1e30f, 1e30f, -1e30f, -1e30f, -1e30f};
  for (int i = lo; i < hi; i++) for (int k = 0; k < 3; k++) {
    float c = 0.5f * (pb[6 * (size_t)idx[i] + k] + pb[6 * (size_t)idx[i] + 3 + k]);
    cb[k] = std::min(cb[k], c); cb[3 + k] = std::max(cb[3 + k], c);
  }
  int ax = 0; float ext = cb[3] - cb[0];
  for (int k = 1; k < 3; k++) if (cb[3 + k] - cb[k] > ext) { ext = cb[3 + k] - cb[k]; ax = k; }
  int mid = (lo + hi) / 2;
  std::nth_element(idx.begin() + lo, idx.begin() + mid, idx.begin() + hi, [&](int a, int b) {
    return pb[6 * (size_t)a + ax] + pb[6 * (size_t)a + 3 + ax] < pb[6 * (size_t)b + ax] + pb[6 * (size_t)b + 3 + ax];
  });
  int me = (int)(s.bvh.size() / 4);
  s.bvh.resize(s.bvh.size() + 4);
  float b0[6], b1[6];
  int c0 = build_rec(s, idx, lo, mid, b0), c1 = build_rec(s, idx, mid, hi, b1);
  s.bvh[4 * (size_t)me + 0] = float4{b0[0], b0[1], b0[2], b0[3]};
  s.bvh[4 * (size_t)me + 1] = float4{b0[4], b0[5], b1[0], b1[1]};
  s.bvh[4 * (size_t)me + 2] = float4{b1[2], b1[3], b1[4], b1[5]};
  s.bvh[4 * (size_t)me + 3] = float4{__int_as_float(c0), __int_as_float(c1), 0, 0};
  for (int k = 0; k < 3; k++) { box_out[k] = std::min(b0[k], b1[k]); box_out[3 + k] = std::max(b0[3 + k], b1[3 + k]); }
  return me;
}

#include "rng.cuh"

extern "C" {

// the renderer's Philox4x32-10 (csrc/rng.cuh) compiled for the host: known-answer + stream checks
void sim_philox_block(const uint32_t *ctr, const uint32_t *key, uint32_t *out) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
  for (int i = 0; i < 10; i++) { gi::philox_round(c0, c1, c2, c3, k0, k1); k0 += 0x9E3779B9u; k1 += 0xBB67AE85u; }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
void sim_rng_doubles(uint64_t seed, uint64_t index, uint32_t stage, uint32_t sample, int n, double *out) {
  gi::Rng r; gi::rng_init(r, seed, index, stage, sample);
  for (int i = 0; i < n; i++) out[i] = gi::rng_double(r);
}

void *sim_scene_load(const char *gisc) {
  Sim *s = new Sim(); std::string err;
  if (!load_scene_binary(gisc, s->data, err)) { fprintf(stderr, "%s\n", err.c_str()); delete s; return nullptr; }
  GiScene v = s->data.view();
  if (!prepare_scene(v, s->ps, err)) { fprintf(stderr, "%s\n", err.c_str()); delete s; return nullptr; }
  int n = v.n_prims;
  if (n == 1) {
    const float *b = s->ps.prim_boxes.data();
    s->bvh = {float4{b[0], b[1], b[2], b[3]}, float4{b[4], b[5], 0, 0}, float4{0, 0, 0, 0},
              float4{__int_as_float(~0), __int_as_float(0x7fffffff), 0, 0}};
  } else if (n > 1) {
    std::vector<int> idx(n); std::iota(idx.begin(), idx.end(), 0);
    float box[6]; build_rec(*s, idx, 0, n, box);
  }
  DevScene &d = s->dev; memset(&d, 0, sizeof(d));
  d.bvh = s->bvh.data(); d.n_bvh_nodes = (int)(s->bvh.size() / 4); d.n_prims = n;
  d.prim_a = s->ps.prim_a.data(); d.prim_b = s->ps.prim_b.data(); d.prim_meta = s->ps.prim_meta.data();
  d.shapes = s->ps.shapes.data(); d.elements = s->ps.elements.data(); d.nodes = s->ps.nodes.data();
  d.materials = s->ps.materials.data();
  d.n_shapes = v.n_shapes; d.n_elements = v.n_elements; d.n_nodes = v.n_nodes; d.n_materials = v.n_materials;
  d.has_mesh = s->ps.has_mesh; d.abs_max = s->ps.abs_max;
  return s;
}
void sim_scene_free(void *p) { delete (Sim *)p; }

int sim_intersect(void *p, const double *rays, uint64_t n, GiHit *hits, int threads) {
  Sim *s = (Sim *)p;
  if (threads < 1) threads = 1;
  auto work = [&](int id) {
    for (uint64_t i = id; i < n; i += threads) {
      RayD r; r.o = ld3(rays + 6 * i); r.d = ld3(rays + 6 * i + 3);
      HitRec h; closest_hit(s->dev, r, h);
      GiHit g; memset(&g, 0, sizeof(g));
      g.hit = h.prim >= 0; g.shape = h.shape; g.prim = h.prim; g.element = h.element; g.t = h.t;
      g.p[0] = h.p.x; g.p[1] = h.p.y; g.p[2] = h.p.z; g.n[0] = h.n.x; g.n[1] = h.n.y; g.n[2] = h.n.z;
      hits[i] = g;
    }
  };
  std::vector<std::thread> th;
  for (int i = 1; i < threads; i++) th.emplace_back(work, i);
  work(0);
  for (auto &t : th) t.join();
  return 0;
}
}
