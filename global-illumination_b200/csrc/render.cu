// csrc/render.cu -- gi_render / gi_render_float: the reference's RenderImage (src/render.cpp:155-259) as batched
// kernels.
//
// The reference walks the image with one CPU thread per column stripe and, per primary hit, RECURSES: RayTrace
// (raytracer.cpp:174-233) calls the direct-light loops, then up to TRANSMISSIVE_TEST + SPECULAR_TEST Monte Carlo
// paths (montecarlo.cpp:16-171) and up to INDIRECT_TEST indirect paths (:177-305), every one of which ends in
// kd-tree gathers.  Here the recursion is unrolled into flat batches with no per-pixel state on a stack:
//
//   primary_kernel   thread per (supersample, depth-of-field sample): eye ray -> closest hit -> PrimHit record
//   direct_kernel    WARP per primary hit: ambient + direct lighting; the 128+256+128 samples of a soft-shadowed
//                    area light are split over the 32 lanes (shade.cuh); writes the number of secondary paths the
//                    hit owns and queues the hit's own caustic / global gathers
//   (cub scan)       secondary paths of all hits get consecutive item numbers
//   path_kernel      thread per secondary path (one transmissive / specular / indirect sample of one hit): the whole
//                    MonteCarlo_PathTrace or MonteCarlo_IndirectSample loop; its local colour goes to item_rgb[],
//                    every EstimateRadiance it would call becomes a request {query, weight, pixel} in a queue
//   gather kernel    (gather.cu) evaluates each queue in one launch, requests ordered by a Morton key for locality
//   commit kernels   item colours and weight x estimate are added into the FP64 accumulator
//   resolve_kernel   clamp, box-filter the 2^aa x 2^aa supersamples, quantise like R2Image::SetPixelRGB
//
// Random numbers: Philox streams keyed by (seed, primary sample, stage, item) -- see shade.cuh; an image does not
// depend on batch size, launch geometry or the number of GPUs sharing the frame.
#include <cub/cub.cuh>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "comm.h"
#include "context.h"
#include "gather.h"
#include "shade.cuh"
#include "wave.cuh"

namespace gi {

enum { STAGE_DOF = 1, STAGE_PATH = 2, STAGE_LIGHT = 16 };     // Philox stage ids (STAGE_LIGHT + 4*light + pass)

struct PrimHit { double p[3], n[3]; int32_t material; uint32_t pix; uint32_t k, pad; };   // material < 0: the eye ray missed
                                                                                          // k: depth-of-field sample of the pixel

struct Queue {                       // gather requests of one photon map, filled by direct_kernel / path_kernel
  GiGatherQuery *q;
  double *w;                         // 3 per request: the colour the estimate is multiplied by
  uint32_t *pix;
  unsigned long long *count;         // may exceed cap: the batch is then repeated with fewer items
  unsigned long long cap;
};

struct Frame2 {                      // per-frame constants of the eye-ray generator (render.cpp:64-77)
  D3 eye, far_org, far_right, far_up, u, v;
  double xc, yc, w, h;               // viewport centre and size in supersamples
  int sw, sh, dof;
  int aaf, rank, world;              // this rank renders the pixel columns x = rank (mod world): local column l is x = l * world + rank
};

GI_DEV void add3(double *a, D3 c) { atomicAdd(a, c.x); atomicAdd(a + 1, c.y); atomicAdd(a + 2, c.z); }

GI_DEV void flush_stats(RayStats *g, const RayStats &s) {
  const unsigned m = __activemask();
  const unsigned long long v[7] = {s.screen, s.shadow, s.monte, s.transmissive, s.specular, s.indirect, s.caustic};
  unsigned long long *out = &g->screen;
  const int leader = __ffs(m) - 1;
#pragma unroll
  for (int i = 0; i < 7; i++) {
    const unsigned t = __reduce_add_sync(m, (unsigned)v[i]);      // per-thread counts stay far below 2^32 / 32
    if ((threadIdx.x & 31) == leader && t) atomicAdd(out + i, (unsigned long long)t);
  }
}

GI_DEV RayStats zero_stats() { RayStats s; s.screen = s.shadow = s.monte = s.transmissive = s.specular = s.indirect = s.caustic = 0; return s; }

// One queue slot per calling lane with ONE atomicAdd per warp: the lanes that are at this call together (and push to
// the same queue) are found with a match on the counter's address, the lowest of them claims slots for all.
GI_DEV unsigned long long claim_slot(unsigned long long *count) {
  const unsigned act = __activemask();
  const unsigned grp = __match_any_sync(act, (unsigned long long)count);
  const int lane = threadIdx.x & 31, leader = __ffs(grp) - 1;
  unsigned long long base = 0;
  if (lane == leader) base = atomicAdd(count, (unsigned long long)__popc(grp));
  base = __shfl_sync(grp, base, leader);
  return base + __popc(grp & ((1u << lane) - 1u));
}

GI_DEV void push_request(const Queue &qu, D3 p, D3 n, D3 bounce, double cos_theta, const Brdf &b, D3 w, uint32_t pix) {
  const unsigned long long i = claim_slot(qu.count);
  if (i >= qu.cap) return;
  GiGatherQuery q;
  q.p[0] = p.x; q.p[1] = p.y; q.p[2] = p.z; q.n[0] = n.x; q.n[1] = n.y; q.n[2] = n.z;
  q.bounce[0] = bounce.x; q.bounce[1] = bounce.y; q.bounce[2] = bounce.z; q.cos_theta = cos_theta;
  q.kd[0] = b.kd.x; q.kd[1] = b.kd.y; q.kd[2] = b.kd.z; q.ks[0] = b.ks.x; q.ks[1] = b.ks.y; q.ks[2] = b.ks.z;
  q.shininess = b.n;
  qu.q[i] = q;
  qu.w[3 * i] = w.x; qu.w[3 * i + 1] = w.y; qu.w[3 * i + 2] = w.z;
  qu.pix[i] = pix;
}

// ------------------------------------------------------------------------------------------------ primary rays
// Threadable_RayTracer, render.cpp:91-131.  Sample s = pix * dof + k, pix = i * sh + j (image_buffer[i][j]).
__global__ void __launch_bounds__(128) primary_kernel(ShadeEnv env, Frame2 f, uint64_t first, uint64_t n,
                                                      PrimHit *__restrict__ hits, double *__restrict__ accum,
                                                      RayStats *stats) {
  RayStats st = zero_stats();
  for (uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t s = first + t;                       // local sample number: (local supersample column, row, dof sample)
    const uint64_t lp = s / f.dof;
    const int k = (int)(s % f.dof), li = (int)(lp / f.sh), j = (int)(lp % f.sh);
    const int i = ((li / f.aaf) * f.world + f.rank) * f.aaf + (li % f.aaf);
    const uint32_t pix = (uint32_t)i * (uint32_t)f.sh + (uint32_t)j;
    const double dx = (double)(2 * (i - f.xc)) / f.w, dy = (double)(2 * (j - f.yc)) / f.h;
    const D3 far_point = f.far_org + (f.far_right * dx) + (f.far_up * dy);
    D3 org = f.eye;
    if (env.p.depth_of_field) {
      Rng rng; rng_init(rng, env.p.seed, pix, STAGE_DOF, (uint32_t)k);
      double r1, r2;
      do { r1 = (rng_double(rng) * 2.0) - 1.0; r2 = (rng_double(rng) * 2.0) - 1.0; } while (r1 * r1 + r2 * r2 > 1.0);
      org = f.eye + f.u * r1 + f.v * r2;
    }
    RayD ray; ray.o = org; ray.d = normalized(far_point - org);       // R3Ray(point, point), R3Ray.cpp:52-58
    HitRec h;
    closest_hit(env.sc, ray, h);
    PrimHit ph;
    ph.pix = pix; ph.k = (uint32_t)k; ph.pad = 0;
    if (h.prim >= 0) {
      ph.p[0] = h.p.x; ph.p[1] = h.p.y; ph.p[2] = h.p.z; ph.n[0] = h.n.x; ph.n[1] = h.n.y; ph.n[2] = h.n.z;
      ph.material = env.sc.elements[h.element].material;
      st.screen++;
    } else {
      ph.p[0] = ph.p[1] = ph.p[2] = ph.n[0] = ph.n[1] = ph.n[2] = 0; ph.material = -1;
      add3(accum + 3 * (size_t)pix, env.background);
    }
    hits[t] = ph;
  }
  flush_stats(stats, st);
}

// The number of secondary paths RayTrace starts at a primary hit (raytracer.cpp:55-57, 87-89, 117-120, 204-223)
struct SubPaths { int n_t, n_s, n_i; double r_coeff; };

GI_DEV SubPaths sub_paths(const GiParams &p, const Brdf &b, double cos_theta) {
  SubPaths sp; sp.n_t = sp.n_s = sp.n_i = 0; sp.r_coeff = 0;
  if (p.transmissive_illum && b.transparent) {
    if (p.specular_illum && p.fresnel) sp.r_coeff = reflection_coeff(cos_theta, b.ir, p.ir_air);
    if (sp.r_coeff < 1.0) {
      const double hw = max_channel(b.kt * (1.0 - sp.r_coeff));
      sp.n_t = (int)ceil((p.transmissive_test * hw + p.transmissive_test) / 2.0);
    }
  }
  if (p.specular_illum && (b.specular || sp.r_coeff > 0)) {
    const double hw = max_channel(b.kt * sp.r_coeff + b.ks);
    sp.n_s = (int)ceil((p.specular_test * hw + p.specular_test) / 2.0);
  }
  if (p.indirect_illum && b.diffuse) {
    const double hw = max_channel(b.kd);
    sp.n_i = (int)ceil((p.indirect_test * hw + p.indirect_test) / 2.0);
  }
  return sp;
}

// ------------------------------------------------------------------------------------------------ primary hits
// The non-recursive part of RayTrace (raytracer.cpp:174-203, 224-231) for one primary hit per WARP.
__global__ void __launch_bounds__(128) direct_kernel(ShadeEnv env, uint64_t n,
                                                     const PrimHit *__restrict__ hits, double *__restrict__ accum,
                                                     uint32_t *__restrict__ n_items, Queue qglobal, Queue qcaustic,
                                                     RayStats *stats) {
  RayStats st = zero_stats();
  const int lane = threadIdx.x & 31;
  const uint64_t warp0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  for (uint64_t t = warp0; t < n; t += nwarps) {
    const PrimHit ph = hits[t];
    if (ph.material < 0) { if (lane == 0) n_items[t] = 0; continue; }
    const Brdf b = load_brdf(env.materials[ph.material]);
    D3 point = ld3(ph.p), normal = ld3(ph.n);
    const D3 cam = env.eye;
    D3 color = mk3(0, 0, 0);
    if (env.p.ambient) color = color + env.ambient;
    const D3 view = normalized(point - cam);
    const double cos_theta = dot(normal, -view);
    if (env.p.ambient && b.ambient) color = color + b.ka;
    if (env.p.direct_illum && (b.diffuse || b.specular)) {
      Rng unused; unused.left = 0;
      SampleKey key; key.seed = env.p.seed; key.index = (uint64_t)ph.pix * (uint64_t)env.p.dof_test + ph.k; key.stage = STAGE_LIGHT;
      direct_illumination(env, point, normal, cam, color, b, cos_theta, false, unused, &key, st);
    }
    if (lane == 0) {
      const SubPaths sp = sub_paths(env.p, b, cos_theta);
      n_items[t] = (uint32_t)(sp.n_t + sp.n_s + sp.n_i);
      const D3 one = mk3(1, 1, 1);
      if (env.p.caustic_illum && b.diffuse) {                                    // CausticIllumination, :138-149
        push_request(qcaustic, point, normal, reflective_bounce(normal, view, cos_theta), cos_theta, b, one, ph.pix);
        st.caustic++;
      }
      if (env.p.direct_photon_illum && b.diffuse) {                              // EstimateGlobalIllumination, :151-167
        push_request(qglobal, point, normal, reflective_bounce(normal, view, cos_theta), cos_theta, b, one, ph.pix);
        if (!env.p.irradiance_cache) st.indirect++;                              // raytracer.cpp:159-166
      }
      add3(accum + 3 * (size_t)ph.pix, color);
    }
  }
  flush_stats(stats, st);
}

// ------------------------------------------------------------------------------------------------ secondary paths
enum { MODE_PATH = 0, MODE_INDIRECT = 1 };


// MonteCarlo_PathTrace (montecarlo.cpp:16-171, mode PATH) and MonteCarlo_IndirectSample (:177-305, mode INDIRECT)
// share one loop body.  `scale` is what the caller multiplies the path's colour with before adding it to the pixel; a
// diffuse pick in mode PATH with INDIRECT_ILLUM on is the nested IndirectIllumination(inMonteCarlo = true) call
// (raytracer.cpp:112-135 with one sample): the walk goes on in mode INDIRECT with the product of the weights and a
// fresh depth count.
struct PathState {
  RayD ray;                          // ray.o doubles as the reference's ray_start
  D3 scale, tw, out;
  Rng rng;
  uint32_t pix, item;
  int mode, iter;
};

// One iteration of the loop; returns false when the walk ends.  MODE = MODE_PATH / MODE_INDIRECT compiles the body for
// that mode only (the wavefront kernels below: the indirect variant carries no direct-light code at all); MODE < 0
// follows s.mode.  The buffered Philox outputs are dropped at the end of every iteration, so that the stream of a path
// is a function of its counter alone and a path can be parked in a queue between two bounces.
// `hit_element` < 0: the ray missed; otherwise the element, point and normal of its closest hit.
template <int MODE>
GI_DEV bool path_shade(const ShadeEnv &env, PathState &s, int hit_element, D3 point, D3 normal, const Queue &qglobal,
                       const Queue &qcaustic, RayStats &st) {
  const GiParams &p = env.p;
  const int mode = MODE < 0 ? s.mode : MODE;
  if (hit_element < 0) { s.out = s.out + mul3(s.scale, mul3(s.tw, env.background)); return false; }
  st.monte++;
  const Brdf b = load_brdf(env.materials[env.sc.elements[hit_element].material]);
  const D3 view = normalized(point - s.ray.o);
  const double cos_theta = dot(normal, -view);
  double r_coeff = 0;
  if (mode == MODE_PATH) {
    D3 cb = mk3(0, 0, 0);
    if (p.ambient) cb = cb + env.ambient;
    if (b.diffuse || b.specular) direct_illumination(env, point, normal, s.ray.o, cb, b, cos_theta, true, s.rng, nullptr, st);
    if (p.caustic_illum && b.diffuse) {
      push_request(qcaustic, point, normal, reflective_bounce(normal, view, cos_theta), cos_theta, b, mul3(s.scale, s.tw), s.pix);
      st.caustic++;
    }
    s.out = s.out + mul3(s.scale, mul3(cb, s.tw));
    if (p.specular_illum && p.transmissive_illum && p.fresnel && b.transparent) r_coeff = reflection_coeff(cos_theta, b.ir, p.ir_air);
  } else if (p.fresnel && b.transparent) r_coeff = reflection_coeff(cos_theta, b.ir, p.ir_air);

  const double prob_diffuse = max_channel(b.kd);
  double prob_transmission = max_channel(b.kt);
  const double prob_specular = max_channel(b.ks) + r_coeff * prob_transmission;
  prob_transmission *= (1.0 - r_coeff);
  const double prob_terminate = max_channel(b.ke) + p.prob_absorb;
  const double prob_total = prob_diffuse + prob_transmission + prob_specular + prob_terminate;
  double rnd = rng_double(s.rng);
  if (prob_total > 1.0) rnd *= prob_total;

  D3 bounce;
  if (rnd < prob_diffuse) {
    const D3 w = mul3(b.kd, s.tw) / prob_diffuse;
    if (mode == MODE_INDIRECT || (!p.indirect_illum && p.fast_global)) {
      push_request(qglobal, point, normal, reflective_bounce(normal, view, cos_theta), cos_theta, b, mul3(s.scale, w), s.pix);
      if (mode == MODE_PATH && !p.irradiance_cache) st.indirect++;   // EstimateGlobalIllumination counts (not with -cache), EstimateRadiance does not
      return false;
    }
    if (!p.indirect_illum) return false;
    s.scale = mul3(s.scale, mul3(b.kd, w));                    // (inner * kd) * kd * tw / prob_diffuse
    bounce = diffuse_sample(normal, cos_theta, s.rng);
    st.indirect++;
    s.mode = MODE_INDIRECT; s.tw = mk3(1, 1, 1); s.iter = -1;
  } else if (rnd < prob_diffuse + prob_transmission) {
    if (mode == MODE_PATH && !p.transmissive_illum) return false;
    const D3 exact = transmissive_bounce(normal, view, cos_theta, b.ir, p.ir_air);
    bounce = p.distrib_transmissive ? specular_sample(exact, b.n, cos_theta, s.rng) : exact;
    st.transmissive++;
    s.tw = mul3(s.tw, (b.kt * (1.0 - r_coeff)) / prob_transmission);
  } else if (rnd < prob_diffuse + prob_transmission + prob_specular) {
    if (mode == MODE_PATH && !p.specular_illum) return false;
    const D3 exact = reflective_bounce(normal, view, cos_theta);
    bounce = p.distrib_specular ? specular_sample(exact, b.n, cos_theta, s.rng) : exact;
    st.specular++;
    s.tw = mul3(s.tw, (b.ks + b.kt * r_coeff) / prob_specular);
  } else return false;
  s.ray.o = point + bounce * GI_EPS; s.ray.d = bounce;
  s.iter++;
  s.rng.left = 0;
  return s.iter < p.max_monte_depth;
}

// closest hit + shading of one bounce in one thread (the finishing pass and the thread-per-path kernel)
template <int MODE>
GI_DEV bool path_step(const ShadeEnv &env, PathState &s, const Queue &qglobal, const Queue &qcaustic, RayStats &st) {
  HitRec h;
  closest_hit(env.sc, s.ray, h);
  return path_shade<MODE>(env, s, h.prim < 0 ? -1 : h.element, h.p, h.n, qglobal, qcaustic, st);
}

// Start of secondary path `g`: item g belongs to the primary hit h with offs[h] <= g < offs[h+1]; inside the hit the
// items are its transmissive samples, then its specular samples, then its indirect samples
// (TransmissiveIllumination / SpecularIllumination / IndirectIllumination, raytracer.cpp:47-135).
// Returns false if the path contributes nothing (MONTE_CARLO off).
GI_DEV bool path_begin(const ShadeEnv &env, const PrimHit *__restrict__ hits,
                       const uint64_t *__restrict__ offs, uint64_t g, uint32_t hidx, PathState &s, RayStats &st) {
  const uint32_t local = (uint32_t)(g - offs[hidx]);
  const PrimHit ph = hits[hidx];
  const Brdf b = load_brdf(env.materials[ph.material]);
  const D3 point = ld3(ph.p), normal = ld3(ph.n);
  const D3 view = normalized(point - env.eye);
  const double cos_theta = dot(normal, -view);
  const SubPaths sp = sub_paths(env.p, b, cos_theta);
  rng_init(s.rng, env.p.seed, (uint64_t)ph.pix * (uint64_t)env.p.dof_test + ph.k, STAGE_PATH, local);
  D3 total_weight, bounce; int count;
  if (local < (uint32_t)sp.n_t) {
    const D3 exact = transmissive_bounce(normal, view, cos_theta, b.ir, env.p.ir_air);
    total_weight = b.kt * (1.0 - sp.r_coeff); count = sp.n_t; s.mode = MODE_PATH;
    bounce = env.p.distrib_transmissive ? specular_sample(exact, b.n, cos_theta, s.rng) : exact;
    st.transmissive++;
  } else if (local < (uint32_t)(sp.n_t + sp.n_s)) {
    const D3 exact = reflective_bounce(normal, view, cos_theta);
    total_weight = b.kt * sp.r_coeff + b.ks; count = sp.n_s; s.mode = MODE_PATH;
    bounce = env.p.distrib_specular ? specular_sample(exact, b.n, cos_theta, s.rng) : exact;
    st.specular++;
  } else {
    total_weight = b.kd; count = sp.n_i; s.mode = MODE_INDIRECT;
    bounce = diffuse_sample(normal, cos_theta, s.rng);
    st.indirect++;
  }
  s.ray.o = point + bounce * GI_EPS; s.ray.d = bounce;
  s.scale = total_weight / (double)count; s.tw = mk3(1, 1, 1); s.out = mk3(0, 0, 0);
  s.pix = ph.pix; s.iter = 0;
  s.rng.left = 0;
  return !(s.mode == MODE_PATH && !env.p.monte_carlo) && env.p.max_monte_depth > 0;
}

// ------------------------------------------------------------------------------------------------ wavefront path stage
// The reference recurses per pixel (RayTrace -> Transmissive/Specular/IndirectIllumination -> MonteCarlo_PathTrace /
// MonteCarlo_IndirectSample, raytracer.cpp:47-135, montecarlo.cpp:16-305).  A thread that walks one path from start to
// end leaves most of its warp idle -- nearly every walk ends at its first hit, a few go on through glass and mirrors,
// and the warp waits for them with 1-4 live lanes (ncu, round 1: 11 of 32 lanes active) -- and it runs 38 000
// instructions of code (600 KB: sampling, FP64 traversal and primitive tests three times over, light loops) against an
// instruction cache of 32 KB per SM: 13 cycles of instruction-fetch stall per issued instruction.  Here the walk is cut
// into BOUNCE PASSES of small kernels.  Between kernels the live paths are parked in queues, one per RAY CLASS:
//   class SPEC (MODE_PATH):     specular / refractive / Fresnel Monte Carlo rays; their hits are shaded with direct light
//                               (shadow rays) and queue a caustic gather,
//   class DIFF (MODE_INDIRECT): diffuse-bounce rays of the indirect estimate; their hits only play roulette and end
//                               in a global-map gather request.
// A pass =  wave_hit_kernel   closest hit for every parked ray of both classes (traversal + primitive tests only),
//           wave_shade_kernel one launch per class: material, direct light (SPEC), roulette, next bounce; gather
//                             requests go to the map queues, survivors are compacted into the next pass's class queues
//                             with a warp ballot (one atomicAdd per warp and class).
// Every kernel runs full warps of ONE class through ONE stage.  The pass count is fixed (GI_WAVE_PASSES) and the kernels
// read the queue lengths on the device, so the host never waits between passes; the few paths still alive after the
// last pass (walks of dozens of bounces inside glass) are finished by a thread each.
GI_DEV void park(PathRec *r, const PathState &s) {
  double4 *o = (double4 *)r;
  o[0] = make_double4(s.ray.o.x, s.ray.o.y, s.ray.o.z, s.ray.d.x);
  o[1] = make_double4(s.ray.d.y, s.ray.d.z, s.scale.x, s.scale.y);
  o[2] = make_double4(s.scale.z, s.tw.x, s.tw.y, s.tw.z);
  uint4 *u = (uint4 *)(r) + 6;
  u[0] = make_uint4(s.rng.c0, s.rng.c1, s.rng.c2, s.rng.c3);
  u[1] = make_uint4(s.pix, s.item, (uint32_t)s.iter, 0u);
}
GI_DEV void unpark(const PathRec *r, uint64_t seed, PathState &s) {
  const double4 *i = (const double4 *)r;
  const double4 a = i[0], b = i[1], c = i[2];
  s.ray.o = mk3(a.x, a.y, a.z); s.ray.d = mk3(a.w, b.x, b.y); s.scale = mk3(b.z, b.w, c.x); s.tw = mk3(c.y, c.z, c.w);
  const uint4 *u = (const uint4 *)(r) + 6;
  const uint4 x = u[0], y = u[1];
  s.rng.k0 = (uint32_t)seed; s.rng.k1 = (uint32_t)(seed >> 32);
  s.rng.c0 = x.x; s.rng.c1 = x.y; s.rng.c2 = x.z; s.rng.c3 = x.w; s.rng.left = 0;
  s.pix = y.x; s.item = y.y; s.iter = (int)y.z;
  s.out = mk3(0, 0, 0);
}

// Every lane of the warp calls this; the lanes with `live` set park their path in the queue of class `cls`.
GI_DEV void wave_push(const PathWave &w, int cls, bool live, const PathState &s) {
  const unsigned m = __ballot_sync(0xffffffffu, live);
  if (!m) return;
  const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
  unsigned base = 0;
  if (lane == leader) base = atomicAdd(w.count + cls, (unsigned)__popc(m));
  base = __shfl_sync(0xffffffffu, base, leader);
  if (live) {
    const unsigned i = base + __popc(m & ((1u << lane) - 1u));
    park(w.rec + (cls == CLASS_SPEC ? i : w.cap - 1u - i), s);
  }
}

// Pass 0, a thread per secondary path of the batch: the path's first ray, parked by class.  Clears every item's colour
// slot (the shade passes add to it).
__global__ void __launch_bounds__(128, 6) wave_begin_kernel(ShadeEnv env, const PrimHit *__restrict__ hits,
                                                            const uint64_t *__restrict__ offs, const uint32_t *__restrict__ item_hit,
                                                            uint64_t item0, uint32_t n, double *__restrict__ item_rgb,
                                                            uint32_t *__restrict__ item_pix, PathWave next, RayStats *stats) {
  RayStats st = zero_stats();
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;       // the grid covers n rounded up to whole warps
  PathState s;
  s.mode = -1; s.pix = 0;
  bool live = false;
  if (t < n) {
    s.item = t;
    live = path_begin(env, hits, offs, item0 + t, item_hit[t] - 1, s, st);
    item_rgb[3 * (size_t)t] = 0.0; item_rgb[3 * (size_t)t + 1] = 0.0; item_rgb[3 * (size_t)t + 2] = 0.0;
    item_pix[t] = s.pix;
  }
  wave_push(next, CLASS_SPEC, live && s.mode == MODE_PATH, s);
  wave_push(next, CLASS_DIFF, live && s.mode == MODE_INDIRECT, s);
  flush_stats(stats, st);
}

// Shading of one class after wave_hit_kernel: a thread per parked path (persistent grid, warp-uniform loop); survivors
// are parked in `next`.  FINISH: the last pass walks every path it gets to its end instead (closest hits inline).
template <int MODE, bool FINISH>
__global__ void __launch_bounds__(128, 6) wave_shade_kernel(ShadeEnv env, PathWave cur, PathWave next, double *__restrict__ item_rgb,
                                                            Queue qglobal, Queue qcaustic, RayStats *stats) {
  RayStats st = zero_stats();
  const int cls = MODE == MODE_PATH ? CLASS_SPEC : CLASS_DIFF;
  const unsigned n = cur.count[cls];
  const unsigned lane = threadIdx.x & 31u;
  for (unsigned base = (blockIdx.x * blockDim.x + threadIdx.x) - lane; base < n; base += gridDim.x * blockDim.x) {
    const unsigned t = base + lane;
    bool live = t < n;
    PathState s;
    s.mode = MODE;
    if (live) {
      const unsigned slot = cls == CLASS_SPEC ? t : cur.cap - 1u - t;
      unpark(cur.rec + slot, env.p.seed, s);
      s.mode = MODE;
      if (FINISH) {
        live = path_step<MODE>(env, s, qglobal, qcaustic, st);
        while (live) live = path_step<-1>(env, s, qglobal, qcaustic, st);
      } else {
        const double4 *hi = (const double4 *)(cur.hit + slot);
        const double4 a = hi[0], b = hi[1];
        live = path_shade<MODE>(env, s, __double2loint(b.z), mk3(a.x, a.y, a.z), mk3(a.w, b.x, b.y), qglobal, qcaustic, st);
      }
      if (s.out.x != 0.0 || s.out.y != 0.0 || s.out.z != 0.0) {     // an item is owned by one thread at a time
        double *o = item_rgb + 3 * (size_t)s.item;
        o[0] += s.out.x; o[1] += s.out.y; o[2] += s.out.z;
      }
    }
    if (!FINISH) {
      if (MODE == MODE_PATH) wave_push(next, CLASS_SPEC, live && s.mode == MODE_PATH, s);
      wave_push(next, CLASS_DIFF, live && s.mode == MODE_INDIRECT, s);
    }
  }
  flush_stats(stats, st);
}

// The round-1 layout, kept for A/B measurements (GI_PATH_MEGAKERNEL=1): one thread per secondary path, start to end.
// MINB = resident CTAs per SM the kernel is compiled for: 6 -> 80 registers, 24 warps per SM.
template <int MINB>
__global__ void __launch_bounds__(128, MINB) path_kernel(ShadeEnv env, const PrimHit *__restrict__ hits,
                                                         const uint64_t *__restrict__ offs, const uint32_t *__restrict__ item_hit,
                                                         uint64_t item0, uint32_t n, double *__restrict__ item_rgb,
                                                         uint32_t *__restrict__ item_pix, Queue qglobal, Queue qcaustic, RayStats *stats) {
  RayStats st = zero_stats();
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) {
    PathState s;
    s.item = t;
    bool active = path_begin(env, hits, offs, item0 + t, item_hit[t] - 1, s, st);
    while (active) active = path_step<-1>(env, s, qglobal, qcaustic, st);
    item_rgb[3 * (size_t)t] = s.out.x; item_rgb[3 * (size_t)t + 1] = s.out.y; item_rgb[3 * (size_t)t + 2] = s.out.z;
    item_pix[t] = s.pix;
  }
  flush_stats(stats, st);
}

// item -> primary hit for a batch of items [item0, item0 + m): every hit that owns items marks its first item with its
// index, the rest is filled by an inclusive max-scan (item_hit[] is zero-initialised; indices are stored + 1).
__global__ void __launch_bounds__(256) mark_items_kernel(const uint64_t *__restrict__ offs, const uint32_t *__restrict__ n_items,
                                                         uint64_t n_hits, uint64_t item0, uint32_t m, uint32_t *__restrict__ item_hit) {
  const uint64_t h = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (h >= n_hits || n_items[h] == 0) return;
  const uint64_t o = offs[h];
  if (o >= item0 && o < item0 + m) item_hit[o - item0] = (uint32_t)h + 1;
  else if (o < item0 && o + n_items[h] > item0) item_hit[0] = (uint32_t)h + 1;     // the hit the batch starts inside
}

// Adds the item colours of a finished batch into the accumulator.  Items of one pixel are consecutive, so a warp
// first folds equal-pixel runs with shuffles and only run heads touch memory.
__global__ void __launch_bounds__(256) commit_items_kernel(const double *__restrict__ item_rgb,
                                                           const uint32_t *__restrict__ item_pix, uint32_t n,
                                                           double *__restrict__ accum) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  const bool live = t < n;
  uint32_t key = live ? item_pix[t] : 0xffffffffu;
  D3 v = live ? mk3(item_rgb[3 * (size_t)t], item_rgb[3 * (size_t)t + 1], item_rgb[3 * (size_t)t + 2]) : mk3(0, 0, 0);
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t k2 = __shfl_down_sync(0xffffffffu, key, o);
    const double x = __shfl_down_sync(0xffffffffu, v.x, o), y = __shfl_down_sync(0xffffffffu, v.y, o),
                 z = __shfl_down_sync(0xffffffffu, v.z, o);
    // runs are contiguous: lane + o has my key only if every lane between has it too
    if (lane + o < 32 && k2 == key) { v.x += x; v.y += y; v.z += z; }
  }
  const uint32_t prev = __shfl_up_sync(0xffffffffu, key, 1);
  if (live && (lane == 0 || prev != key)) add3(accum + 3 * (size_t)key, v);
}

// accum[pix] += weight * estimate for every request of a queue
__global__ void __launch_bounds__(256) commit_requests_kernel(const double *__restrict__ w, const uint32_t *__restrict__ pix,
                                                              const double *__restrict__ rgb, uint64_t n,
                                                              double *__restrict__ accum) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  add3(accum + 3 * (size_t)pix[t], mk3(w[3 * t] * rgb[3 * t], w[3 * t + 1] * rgb[3 * t + 1], w[3 * t + 2] * rgb[3 * t + 2]));
}

// Requests reach a queue in the order the path threads happened to push them, i.e. scattered over the scene.  The
// gather orders its batch along a Morton curve itself (query_order.cu); the irradiance-cache lookups (a thread per
// request walking the kd-tree) are sorted here for the same reason: neighbouring threads then take the same branches.
GI_DEV uint32_t spread10(uint32_t v) {
  v &= 1023u; v = (v | (v << 16)) & 0x030000FFu; v = (v | (v << 8)) & 0x0300F00Fu; v = (v | (v << 4)) & 0x030C30C3u;
  return (v | (v << 2)) & 0x09249249u;
}
__global__ void __launch_bounds__(256) request_keys_kernel(const GiGatherQuery *__restrict__ q, uint64_t n, D3 lo, D3 inv,
                                                           uint32_t *__restrict__ keys, uint32_t *__restrict__ order) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const double x = (q[t].p[0] - lo.x) * inv.x, y = (q[t].p[1] - lo.y) * inv.y, z = (q[t].p[2] - lo.z) * inv.z;
  const uint32_t xi = (uint32_t)fmin(fmax(x * 1024.0, 0.0), 1023.0), yi = (uint32_t)fmin(fmax(y * 1024.0, 0.0), 1023.0),
                 zi = (uint32_t)fmin(fmax(z * 1024.0, 0.0), 1023.0);
  keys[t] = spread10(xi) | (spread10(yi) << 1) | (spread10(zi) << 2);
  order[t] = (uint32_t)t;
}
// 136-byte records moved as 17 coalesced 8-byte words per request
__global__ void __launch_bounds__(256) permute_requests_kernel(const double *__restrict__ in, const uint32_t *__restrict__ order,
                                                               uint64_t n, double *__restrict__ out) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * 17) return;
  const uint64_t r = t / 17, c = t % 17;
  out[t] = in[(uint64_t)order[r] * 17 + c];
}
// accum[pix] += weight * estimate; request order[t] owns estimate t
__global__ void __launch_bounds__(256) commit_sorted_kernel(const double *__restrict__ w, const uint32_t *__restrict__ pix,
                                                            const uint32_t *__restrict__ order, const double *__restrict__ rgb,
                                                            uint64_t n, double *__restrict__ accum) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const uint64_t r = order[t];
  add3(accum + 3 * (size_t)pix[r], mk3(w[3 * r] * rgb[3 * t], w[3 * r + 1] * rgb[3 * t + 1], w[3 * r + 2] * rgb[3 * t + 2]));
}

// image_buffer[i][j] /= DOF_TEST (render.cpp:135)
__global__ void scale_kernel(double *a, uint64_t n, double dof) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) a[t] = a[t] / dof;
}

// RenderImage's down-sampling loop (render.cpp:205-221) + R2Image::SetPixelRGB (R2Image.cpp:194-212): clamp every
// supersample to [0,1], sum each pixel's block in the reference's (row, column) order, scale by the box weight,
// truncate to a byte.  Output pixel (x, y) at rgb[(y * width + x) * 3], y = 0 the bottom row.
__global__ void resolve_kernel(const double *__restrict__ accum, int width, int height, int aaf, int sh, uint8_t *rgb) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= width || y >= height) return;
  const double box_weight = 1.0 / aaf / aaf;
  double s[3] = {0, 0, 0};
  for (int j = y * aaf; j < (y + 1) * aaf; j++)
    for (int i = x * aaf; i < (x + 1) * aaf; i++)
      for (int c = 0; c < 3; c++) {
        double v = accum[3 * ((size_t)i * sh + j) + c];
        if (v < 0) v = 0;
        if (v > 1.0) v = 1.0;
        s[c] += v;
      }
  for (int c = 0; c < 3; c++) rgb[3 * ((size_t)y * width + x) + c] = (uint8_t)(int)(255 * (box_weight * s[c]));
}

// Multi-GPU frame (comm.cu): rank r owns the pixel columns x = r (mod world).  resolve_strip_kernel is resolve_kernel for
// those columns only, written as a compact strip [y][l][3] with l = x / world; on rank 0 assemble_kernel interleaves the
// gathered strips (strip of rank r starts at byte off[r]) into the frame.
__global__ void resolve_strip_kernel(const double *__restrict__ accum, int width, int height, int aaf, int sh, int rank, int world,
                                     uint8_t *strip) {
  const int l = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  const int ncols = (width - rank + world - 1) / world;
  if (l >= ncols || y >= height) return;
  const int x = l * world + rank;
  const double box_weight = 1.0 / aaf / aaf;
  double s[3] = {0, 0, 0};
  for (int j = y * aaf; j < (y + 1) * aaf; j++)
    for (int i = x * aaf; i < (x + 1) * aaf; i++)
      for (int c = 0; c < 3; c++) {
        double v = accum[3 * ((size_t)i * sh + j) + c];
        if (v < 0) v = 0;
        if (v > 1.0) v = 1.0;
        s[c] += v;
      }
  for (int c = 0; c < 3; c++) strip[3 * ((size_t)y * ncols + l) + c] = (uint8_t)(int)(255 * (box_weight * s[c]));
}
__global__ void assemble_kernel(const uint8_t *__restrict__ strips, const uint64_t *__restrict__ off, int width, int height, int world,
                                uint8_t *rgb) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= width || y >= height) return;
  const int r = x % world, l = x / world, ncols = (width - r + world - 1) / world;
  const uint8_t *src = strips + off[r] + 3 * ((size_t)y * ncols + l);
  uint8_t *dst = rgb + 3 * ((size_t)y * width + x);
  dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2];
}

// ------------------------------------------------------------------------------------------------ host driver
namespace {

struct DevBuf {                      // frees on scope exit
  void *p = nullptr;
  ~DevBuf() { if (p) cudaFree(p); }
  bool alloc(size_t bytes) {
    if (p) { cudaFree(p); p = nullptr; }
    return cuda_ok(cudaMalloc(&p, bytes ? bytes : 1), "cudaMalloc(render)");
  }
  template <class T> T *as() const { return (T *)p; }
};

struct QueueStore {
  DevBuf q, w, pix, count, rgb;
  DevBuf sorted_q, keys, keys2, order, order2, sort_tmp;      // Morton ordering of the requests
  size_t sort_bytes = 0;
  Queue view{};
  bool alloc(unsigned long long cap) {
    if (!q.alloc(sizeof(GiGatherQuery) * cap) || !w.alloc(sizeof(double) * 3 * cap) || !pix.alloc(sizeof(uint32_t) * cap) ||
        !count.alloc(sizeof(unsigned long long)) || !rgb.alloc(sizeof(double) * 3 * cap))
      return false;
    view.q = q.as<GiGatherQuery>(); view.w = w.as<double>(); view.pix = pix.as<uint32_t>();
    view.count = count.as<unsigned long long>(); view.cap = cap;
    if (!keys.alloc(4 * cap) || !keys2.alloc(4 * cap) || !order.alloc(4 * cap) ||
        !order2.alloc(4 * cap))
      return false;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, keys.as<uint32_t>(), keys2.as<uint32_t>(), order.as<uint32_t>(),
                                    order2.as<uint32_t>(), (int)cap, 0, 30);
    return sort_tmp.alloc(sort_bytes);
  }
};

// Device buffers of the render stages, allocated on the first gi_render of a context and kept until gi_destroy
// (cudaMalloc / cudaFree of a few GB per frame costs tens of milliseconds and jitters).
struct Workspace {
  static constexpr uint64_t chunk = 1ull << 22;          // primary samples per pass
  static constexpr uint32_t batch_max = 1u << 23;        // secondary paths per pass (the serial tail of a batch -- a few walks of dozens of bounces -- is paid once per batch)
  static constexpr unsigned long long qcap = 1ull << 23; // >= chunk (direct_kernel pushes at most one request per hit)
  DevBuf hits, n_items, offs, item_rgb, item_pix, scan_tmp, item_hit, hit_scan_tmp, lights, stats, accum, image, strips, comm_tmp;
  DevBuf wave_rec[2], wave_hit[2], wave_count;                        // parked paths of the wavefront path stage: ping / pong, 2 x 2 class counters
  size_t image_bytes = 0, strips_bytes = 0;
  QueueStore qg, qc;
  size_t scan_bytes = 0, hit_scan_bytes = 0, accum_doubles = 0, n_lights = 0;
  cudaStream_t side = nullptr;                           // the finishing pass of a path batch runs here, beside the gather drains
  cudaEvent_t ev_passes = nullptr, ev_tail = nullptr;
  bool ready = false;
  ~Workspace() {
    if (side) cudaStreamDestroy(side);
    if (ev_passes) cudaEventDestroy(ev_passes);
    if (ev_tail) cudaEventDestroy(ev_tail);
  }
  bool init(cudaStream_t st) {
    if (ready) return true;
    if (!hits.alloc(sizeof(PrimHit) * chunk) || !n_items.alloc(4 * (chunk + 1)) || !offs.alloc(8 * (chunk + 1)) ||
        !item_rgb.alloc(sizeof(double) * 3 * batch_max) || !item_pix.alloc(4 * (size_t)batch_max) || !item_hit.alloc(4 * (size_t)batch_max) ||
        !stats.alloc(2 * sizeof(RayStats)) || !qg.alloc(qcap) || !qc.alloc(qcap) ||
        !wave_rec[0].alloc(sizeof(PathRec) * (size_t)batch_max) || !wave_rec[1].alloc(sizeof(PathRec) * (size_t)batch_max) ||
        !wave_hit[0].alloc(sizeof(WaveHit) * (size_t)batch_max) || !wave_hit[1].alloc(sizeof(WaveHit) * (size_t)batch_max) || !wave_count.alloc(16))
      return false;
    cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, n_items.as<uint32_t>(), offs.as<uint64_t>(), (int)(chunk + 1), st);
    cub::DeviceScan::InclusiveScan(nullptr, hit_scan_bytes, item_hit.as<uint32_t>(), item_hit.as<uint32_t>(), cub::Max(), (int)batch_max, st);
    if (!scan_tmp.alloc(scan_bytes) || !hit_scan_tmp.alloc(hit_scan_bytes)) return false;
    if (!cuda_ok(cudaStreamCreateWithFlags(&side, cudaStreamNonBlocking), "cudaStreamCreate(side)") ||
        !cuda_ok(cudaEventCreateWithFlags(&ev_passes, cudaEventDisableTiming), "cudaEventCreate") ||
        !cuda_ok(cudaEventCreateWithFlags(&ev_tail, cudaEventDisableTiming), "cudaEventCreate"))
      return false;
    ready = true;
    return true;
  }
  bool fit_accum(size_t doubles) {
    if (doubles <= accum_doubles) return true;
    if (!accum.alloc(sizeof(double) * doubles)) { accum_doubles = 0; return false; }
    accum_doubles = doubles;
    return true;
  }
  bool fit_image(size_t bytes) {
    if (bytes <= image_bytes) return true;
    if (!image.alloc(bytes)) { image_bytes = 0; return false; }
    image_bytes = bytes;
    return true;
  }
  bool fit_strips(size_t bytes) {
    if (bytes <= strips_bytes) return true;
    if (!strips.alloc(bytes)) { strips_bytes = 0; return false; }
    strips_bytes = bytes;
    return comm_tmp.p || comm_tmp.alloc(8 * 64 + 8 * 8);          // strip offsets of up to 64 ranks + the 7 counters
  }
  bool fit_lights(size_t n) {
    if (n <= n_lights && lights.p) return true;
    if (!lights.alloc(sizeof(GiLight) * (n ? n : 1))) { n_lights = 0; return false; }
    n_lights = n;
    return true;
  }
};
void free_workspace(void *p) { delete (Workspace *)p; }

Workspace *workspace(GiContext *ctx) {
  if (!ctx->render_ws) { ctx->render_ws = new Workspace(); ctx->render_ws_free = free_workspace; }
  Workspace *w = (Workspace *)ctx->render_ws;
  const bool first = !w->ready;
  if (!w->init(ctx->stream)) return nullptr;
  if (first && !gather_reserve(ctx, Workspace::qcap, ctx->stream)) { set_error("gi_render: out of device memory (gather scratch)"); return nullptr; }
  return w;
}

inline int grid_for(uint64_t n, int block, int sm_count, int per_sm) {
  const uint64_t want = (n + block - 1) / block, cap = (uint64_t)sm_count * per_sm;
  return (int)std::max<uint64_t>(1, std::min(want, cap));
}

struct Renderer {
  GiContext *ctx;
  ShadeEnv env;
  cudaStream_t st;
  double *accum;
  RayStats *d_stats;
  QueueStore &qg, &qc;
  bool sort_requests = true;
  Renderer(Workspace &w) : qg(w.qg), qc(w.qc) {}

  // d_stats[0]: primary + direct kernels; d_stats[1]: the path batch in flight (counted once the batch commits)
  bool reset_queues() {
    return cuda_ok(cudaMemsetAsync(d_stats + 1, 0, sizeof(RayStats), st), "memset(stats)") &&
           cuda_ok(cudaMemsetAsync(qg.view.count, 0, sizeof(unsigned long long), st), "memset(queue)") &&
           cuda_ok(cudaMemsetAsync(qc.view.count, 0, sizeof(unsigned long long), st), "memset(queue)");
  }
  // reads both queue counters; false on a CUDA error
  bool queue_counts(unsigned long long &ng, unsigned long long &nc, RayStats *batch_stats = nullptr) {
    unsigned long long h[2] = {0, 0};
    if ((batch_stats && !cuda_ok(cudaMemcpyAsync(batch_stats, d_stats + 1, sizeof(RayStats), cudaMemcpyDeviceToHost, st), "D2H stats")) ||
        !cuda_ok(cudaMemcpyAsync(&h[0], qg.view.count, 8, cudaMemcpyDeviceToHost, st), "D2H queue count") ||
        !cuda_ok(cudaMemcpyAsync(&h[1], qc.view.count, 8, cudaMemcpyDeviceToHost, st), "D2H queue count") ||
        !cuda_ok(cudaStreamSynchronize(st), "render sync"))
      return false;
    ctx->clock.collect();
    ng = h[0]; nc = h[1];
    return true;
  }
  // The two halves of drain() for requests [off, off + n) of a queue (photon gathers only, not the -cache lookups): the
  // estimates can be computed while the batch's finishing pass still runs; the commit waits until the batch is known good.
  bool estimate_range(QueueStore &qs, unsigned long long off, unsigned long long n, int map) {
    if (n == 0 || ctx->maps[map].dev.n == 0) return true;
    const GiParams &p = env.p;
    const int k = map == GI_MAP_GLOBAL ? p.global_estimate_size : p.caustic_estimate_size;
    const double r = map == GI_MAP_GLOBAL ? p.global_estimate_dist : p.caustic_estimate_dist;
    const int filter = map == GI_MAP_GLOBAL ? p.global_filter : p.caustic_filter;
    ctx->clock.begin(map == GI_MAP_GLOBAL ? ST_GATHER_GLOBAL : ST_GATHER_CAUSTIC, st);
    const int rc = gi_gather_batch_device(ctx, map, qs.view.q + off, n, k, r, filter, &p, nullptr, nullptr, qs.rgb.as<double>() + 3 * off, st);
    ctx->clock.end(st);
    return rc == 0;
  }
  bool commit_range(QueueStore &qs, unsigned long long n, int map) {
    if (n == 0 || ctx->maps[map].dev.n == 0) return true;
    ctx->clock.begin(ST_COMMIT, st);
    commit_requests_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(qs.view.w, qs.view.pix, qs.rgb.as<double>(), n, accum);
    ctx->clock.end(st);
    ctx->launches++;
    return cuda_ok(cudaGetLastError(), "commit_requests_kernel");
  }
  // EstimateRadiance for every request of a queue, then accum += weight * estimate
  bool drain(QueueStore &qs, unsigned long long n, int map) {
    if (n == 0) return true;
    const GiParams &p = env.p;
    const int k = map == GI_MAP_GLOBAL ? p.global_estimate_size : p.caustic_estimate_size;
    const double r = map == GI_MAP_GLOBAL ? p.global_estimate_dist : p.caustic_estimate_dist;
    const int filter = map == GI_MAP_GLOBAL ? p.global_filter : p.caustic_filter;
    if (ctx->maps[map].dev.n == 0) return true;          // empty map: EstimateRadiance adds nothing
    // -cache: requests to the global map are EstimateCachedRadiance lookups (raytracer.cpp:159-162, montecarlo.cpp:251-254)
    const bool cached = map == GI_MAP_GLOBAL && p.irradiance_cache;
    auto estimate = [&](const GiGatherQuery *q) {
      return cached ? gi_cached_radiance_batch_device(ctx, q, n, r, qs.rgb.as<double>(), st)
                    : gi_gather_batch_device(ctx, map, q, n, k, r, filter, &p, nullptr, nullptr, qs.rgb.as<double>(), st);
    };
    // photon gathers order their batch themselves (query_order.cu); only the cache lookups are sorted here
    StageClock &clk = ctx->clock;
    const int gstage = map == GI_MAP_GLOBAL ? ST_GATHER_GLOBAL : ST_GATHER_CAUSTIC;
    if (n < 4096 || !sort_requests || !cached) {
      clk.begin(gstage, st);
      if (estimate(qs.view.q)) return false;
      clk.end(st);
      clk.begin(ST_COMMIT, st);
      commit_requests_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(qs.view.w, qs.view.pix, qs.rgb.as<double>(), n, accum);
      clk.end(st);
      ctx->launches++;
      return cuda_ok(cudaGetLastError(), "commit_requests_kernel");
    }
    clk.begin(gstage, st);
    const D3 lo = ld3(ctx->bbox_min), hi = ld3(ctx->bbox_max);
    const D3 inv = mk3(1.0 / fmax(hi.x - lo.x, 1e-300), 1.0 / fmax(hi.y - lo.y, 1e-300), 1.0 / fmax(hi.z - lo.z, 1e-300));
    const unsigned blocks = (unsigned)((n + 255) / 256);
    request_keys_kernel<<<blocks, 256, 0, st>>>(qs.view.q, n, lo, inv, qs.keys.as<uint32_t>(), qs.order.as<uint32_t>());
    size_t bytes = qs.sort_bytes;
    if (!cuda_ok(cub::DeviceRadixSort::SortPairs(qs.sort_tmp.p, bytes, qs.keys.as<uint32_t>(), qs.keys2.as<uint32_t>(),
                                                 qs.order.as<uint32_t>(), qs.order2.as<uint32_t>(), (int)n, 0, 30, st), "sort requests"))
      return false;
    if (!qs.sorted_q.p && !qs.sorted_q.alloc(sizeof(GiGatherQuery) * qs.view.cap)) return false;      // -cache only: allocated on first use
    permute_requests_kernel<<<(unsigned)((n * 17 + 255) / 256), 256, 0, st>>>((const double *)qs.view.q, qs.order2.as<uint32_t>(), n,
                                                                             qs.sorted_q.as<double>());
    if (estimate(qs.sorted_q.as<GiGatherQuery>())) return false;
    clk.end(st);
    clk.begin(ST_COMMIT, st);
    commit_sorted_kernel<<<blocks, 256, 0, st>>>(qs.view.w, qs.view.pix, qs.order2.as<uint32_t>(), qs.rgb.as<double>(), n, accum);
    clk.end(st);
    ctx->launches += 7;
    return cuda_ok(cudaGetLastError(), "sorted gather");
  }
};

}  // namespace

// Renders this rank's pixel columns (x = rank mod world; all 2^aa x 2^aa supersamples of a pixel stay on one rank) of
// the sw x sh supersampled frame into `accum` (3 doubles per supersample, zero-initialised by the caller), already
// divided by DOF_TEST.  Columns are interleaved like the reference's `i % THREADS == id` (render.cpp:90): neighbouring
// columns cost about the same, so the ranks finish together.
static int render_accum(GiContext *ctx, const GiParams &params, int sw, int sh, int aaf, int rank, int world, double *accum,
                        RayStats *d_stats, RayStats &host_paths) {
  Workspace *wsp = workspace(ctx);
  if (!wsp) return 1;
  Workspace &ws = *wsp;
  Renderer R(ws);
  R.ctx = ctx; R.st = ctx->stream; R.accum = accum; R.d_stats = d_stats;
  if (const char *e = getenv("GI_SORT_REQUESTS")) R.sort_requests = atoi(e) != 0;
  ShadeEnv &env = R.env;
  env.sc = ctx->scene; env.materials = ctx->scene.materials; env.p = params;
  env.ambient = ld3(ctx->ambient); env.background = ld3(ctx->background); env.eye = ld3(ctx->camera.eye);
  env.radius = ctx->radius; env.spec_boxes = nullptr; env.n_spec_boxes = -1;
  GiParams &p = env.p;
  if (p.dof_test < 1) p.dof_test = 1;
  // a map without photons contributes nothing: same switches MapPhotons flips (photonmap.cpp:339-361)
  if (ctx->maps[GI_MAP_GLOBAL].dev.n == 0) { p.indirect_illum = 0; p.direct_photon_illum = 0; p.fast_global = 0; }
  if (ctx->maps[GI_MAP_CAUSTIC].dev.n == 0) p.caustic_illum = 0;

  env.n_lights = (int)ctx->lights.size();
  if (!ws.fit_lights(ctx->lights.size())) return 1;
  if (env.n_lights)
    GI_CUDA(cudaMemcpyAsync(ws.lights.p, ctx->lights.data(), sizeof(GiLight) * ctx->lights.size(), cudaMemcpyHostToDevice, R.st));
  env.lights = ws.lights.as<GiLight>();

  Frame2 f;                                               // render.cpp:64-77
  const GiCamera &cam = ctx->camera;
  f.eye = ld3(cam.eye);
  f.far_org = f.eye + ld3(cam.towards) * p.focus_depth;
  f.far_right = ld3(cam.right) * tan(cam.xfov) * p.focus_depth;
  // SCENE->SetViewport() re-derives the vertical field of view from the frame's aspect ratio (R3Viewer.cpp:76-84)
  const double yfov = atan(((double)sh / (double)sw) * tan(cam.xfov));
  f.far_up = ld3(cam.up) * tan(yfov) * p.focus_depth;
  f.u = normalized(ld3(cam.up)) * p.aperture_radius; f.v = normalized(ld3(cam.right)) * p.aperture_radius;
  f.xc = sw / 2; f.yc = sh / 2; f.w = sw; f.h = sh; f.sw = sw; f.sh = sh; f.dof = p.dof_test;
  f.aaf = aaf; f.rank = rank; f.world = world;
  const int width = sw / aaf;
  const uint64_t local_cols = (uint64_t)((width - rank + world - 1) / world) * aaf;       // supersample columns of this rank

  const uint64_t chunk = Workspace::chunk;
  uint32_t batch = Workspace::batch_max;
  unsigned long long qcap = Workspace::qcap;
  // test hooks: a smaller path batch / request capacity of the path stage, to exercise the overflow-and-repeat branch
  if (const char *e = getenv("GI_BATCH_MAX")) batch = (uint32_t)std::min<long long>(std::max<long long>(atoll(e), 1024), Workspace::batch_max);
  if (const char *e = getenv("GI_QCAP")) qcap = (unsigned long long)std::min<long long>(std::max<long long>(atoll(e), 1024), (long long)Workspace::qcap);
  Queue path_qg = R.qg.view, path_qc = R.qc.view;
  path_qg.cap = path_qc.cap = qcap;
  DevBuf &hits = ws.hits, &n_items = ws.n_items, &offs = ws.offs, &item_rgb = ws.item_rgb, &item_pix = ws.item_pix, &scan_tmp = ws.scan_tmp,
         &item_hit = ws.item_hit, &hit_scan_tmp = ws.hit_scan_tmp;
  size_t scan_bytes = ws.scan_bytes, hit_scan_bytes = ws.hit_scan_bytes;
  int path_occ = 6;                                       // resident CTAs per SM: 4 (128 registers) or 6 (80 registers)
  if (const char *e = getenv("GI_PATH_OCC")) path_occ = atoi(e);
  const bool megakernel = getenv("GI_PATH_MEGAKERNEL") && atoi(getenv("GI_PATH_MEGAKERNEL")) != 0;      // A/B hook: the round-1 thread-per-path kernel
  int hit_occ = 6;                                        // resident CTAs per SM wave_hit_kernel is compiled for (registers: 6 -> 80, 4 -> 128, 3 -> 168)
  if (const char *e = getenv("GI_HIT_OCC")) hit_occ = atoi(e);
  int wave_passes = 8;                                    // wavefront bounce passes per class before the finishing pass (sweep 2-16 on one B200:
                                                          // cornell 512x512 aa 1 render 501 / 476 / 467 / 456 / 451 / 454 / 454 ms at 2 / 3 / 4 / 6 / 8 / 12 / 16)
  if (const char *e = getenv("GI_WAVE_PASSES")) wave_passes = std::min(std::max(atoi(e), 0), 64);

  const uint64_t s_begin = 0, s_end = local_cols * sh * p.dof_test;
  for (uint64_t s0 = s_begin; s0 < s_end; s0 += chunk) {
    const uint64_t n = std::min<uint64_t>(chunk, s_end - s0);
    if (!R.reset_queues()) return 1;
    StageClock &clk = ctx->clock;
    clk.begin(ST_PRIMARY, R.st);
    primary_kernel<<<grid_for(n, 128, ctx->sm_count, 64), 128, 0, R.st>>>(env, f, s0, n, hits.as<PrimHit>(), accum, d_stats);
    clk.end(R.st);
    GI_CUDA(cudaMemsetAsync(n_items.p, 0, 4 * (chunk + 1), R.st));
    clk.begin(ST_DIRECT, R.st);
    direct_kernel<<<grid_for(n * 32, 128, ctx->sm_count, 16), 128, 0, R.st>>>(env, n, hits.as<PrimHit>(), accum,
                                                                             n_items.as<uint32_t>(), R.qg.view, R.qc.view, d_stats);
    clk.end(R.st);
    GI_CUDA(cudaGetLastError());
    GI_CUDA(cub::DeviceScan::ExclusiveSum(scan_tmp.p, scan_bytes, n_items.as<uint32_t>(), offs.as<uint64_t>(), (int)(n + 1), R.st));
    ctx->launches += 3;
    uint64_t total_items = 0;
    GI_CUDA(cudaMemcpyAsync(&total_items, offs.as<uint64_t>() + n, 8, cudaMemcpyDeviceToHost, R.st));
    unsigned long long ng = 0, nc = 0;
    if (!R.queue_counts(ng, nc)) return 1;
    if (!R.drain(R.qg, ng, GI_MAP_GLOBAL) || !R.drain(R.qc, nc, GI_MAP_CAUSTIC)) return 1;

    for (uint64_t item0 = 0; item0 < total_items;) {
      const uint32_t m = (uint32_t)std::min<uint64_t>(batch, total_items - item0);
      if (!R.reset_queues()) return 1;
      GI_CUDA(cudaMemsetAsync(item_hit.p, 0, 4 * (size_t)m, R.st));
      mark_items_kernel<<<(unsigned)((n + 255) / 256), 256, 0, R.st>>>(offs.as<uint64_t>(), n_items.as<uint32_t>(), n, item0, m,
                                                                      item_hit.as<uint32_t>());
      GI_CUDA(cub::DeviceScan::InclusiveScan(hit_scan_tmp.p, hit_scan_bytes, item_hit.as<uint32_t>(), item_hit.as<uint32_t>(), cub::Max(),
                                             (int)m, R.st));
      bool overlap_tail = false;
      if (megakernel) {
        clk.begin(ST_PATH_SHADE, R.st);
        auto launch = [&](auto kernel) {
          kernel<<<(m + 127) / 128, 128, 0, R.st>>>(env, hits.as<PrimHit>(), offs.as<uint64_t>(), item_hit.as<uint32_t>(), item0, m,
                                                    item_rgb.as<double>(), item_pix.as<uint32_t>(), path_qg, path_qc, d_stats + 1);
        };
        if (path_occ <= 4) launch(path_kernel<4>); else launch(path_kernel<6>);
        clk.end(R.st);
      } else {
        // wavefront: pass 0 parks every path's first ray by class; then bounce passes (closest-hit kernel + one shade kernel per
        // class) over ballot-compacted queues.  The pass count is fixed: the host never waits for a queue length.
        PathWave wv[2];
        for (int i = 0; i < 2; i++) {
          wv[i].rec = ws.wave_rec[i].as<PathRec>(); wv[i].hit = ws.wave_hit[i].as<WaveHit>();
          wv[i].count = ws.wave_count.as<unsigned int>() + 2 * i; wv[i].cap = Workspace::batch_max;
        }
        GI_CUDA(cudaMemsetAsync(ws.wave_count.p, 0, 16, R.st));
        clk.begin(ST_PATH_BEGIN, R.st);
        wave_begin_kernel<<<(m + 127) / 128, 128, 0, R.st>>>(env, hits.as<PrimHit>(), offs.as<uint64_t>(), item_hit.as<uint32_t>(), item0, m,
                                                             item_rgb.as<double>(), item_pix.as<uint32_t>(), wv[0], d_stats + 1);
        clk.end(R.st);
        const unsigned wgrid = (unsigned)std::min<uint64_t>((m + 127) / 128, (uint64_t)ctx->sm_count * 6);
        int cur = 0;
        for (int pass = 0; pass < wave_passes; pass++, cur ^= 1) {
          GI_CUDA(cudaMemsetAsync(wv[cur ^ 1].count, 0, 8, R.st));
          clk.begin(ST_PATH_HIT, R.st);
          if (hit_occ <= 3) wave_hit_kernel<3><<<wgrid, 128, 0, R.st>>>(env.sc, wv[cur]);
          else if (hit_occ == 4) wave_hit_kernel<4><<<wgrid, 128, 0, R.st>>>(env.sc, wv[cur]);
          else if (hit_occ == 5) wave_hit_kernel<5><<<wgrid, 128, 0, R.st>>>(env.sc, wv[cur]);
          else wave_hit_kernel<6><<<wgrid, 128, 0, R.st>>>(env.sc, wv[cur]);
          clk.end(R.st);
          clk.begin(ST_PATH_SHADE, R.st);
          wave_shade_kernel<MODE_PATH, false><<<wgrid, 128, 0, R.st>>>(env, wv[cur], wv[cur ^ 1], item_rgb.as<double>(), path_qg, path_qc, d_stats + 1);
          wave_shade_kernel<MODE_INDIRECT, false><<<wgrid, 128, 0, R.st>>>(env, wv[cur], wv[cur ^ 1], item_rgb.as<double>(), path_qg, path_qc, d_stats + 1);
          clk.end(R.st);
        }
        // The finishing pass -- a few walks of dozens of bounces, a thread each: ~1 ms with the GPU idle -- runs on a side
        // stream while this stream already computes the estimates of the requests queued so far.  Nothing is committed
        // before the final counts are known (a queue overflow repeats the batch).
        overlap_tail = !p.irradiance_cache && !getenv("GI_NO_TAIL_OVERLAP");
        cudaStream_t ts = overlap_tail ? ws.side : R.st;
        if (overlap_tail) { GI_CUDA(cudaEventRecord(ws.ev_passes, R.st)); GI_CUDA(cudaStreamWaitEvent(ws.side, ws.ev_passes, 0)); }
        clk.begin(ST_PATH_FINISH, ts);
        wave_shade_kernel<MODE_PATH, true><<<wgrid, 128, 0, ts>>>(env, wv[cur], wv[cur ^ 1], item_rgb.as<double>(), path_qg, path_qc, d_stats + 1);
        wave_shade_kernel<MODE_INDIRECT, true><<<wgrid, 128, 0, ts>>>(env, wv[cur], wv[cur ^ 1], item_rgb.as<double>(), path_qg, path_qc, d_stats + 1);
        clk.end(ts);
        if (overlap_tail) GI_CUDA(cudaEventRecord(ws.ev_tail, ws.side));
        ctx->launches += 2 + 3 * (uint64_t)wave_passes;
      }
      ctx->launches += 2;
      GI_CUDA(cudaGetLastError());
      ctx->launches++;
      RayStats bs;
      unsigned long long ng1 = 0, nc1 = 0;
      if (overlap_tail) {
        if (!R.queue_counts(ng1, nc1)) return 1;              // counts before the tail's requests
        if (ng1 <= qcap && nc1 <= qcap) { if (!R.estimate_range(R.qg, 0, ng1, GI_MAP_GLOBAL) || !R.estimate_range(R.qc, 0, nc1, GI_MAP_CAUSTIC)) return 1; }
        GI_CUDA(cudaStreamWaitEvent(R.st, ws.ev_tail, 0));
      }
      if (!R.queue_counts(ng, nc, &bs)) return 1;
      if (ng > qcap || nc > qcap) {                       // a queue overflowed: nothing was committed, repeat with fewer paths
        if (batch <= 1024) { set_error("gi_render: gather queue overflow"); return 1; }
        batch >>= 1;
        ctx->queue_overflows++;
        continue;
      }
      clk.begin(ST_COMMIT, R.st);
      commit_items_kernel<<<(m + 255) / 256, 256, 0, R.st>>>(item_rgb.as<double>(), item_pix.as<uint32_t>(), m, accum);
      clk.end(R.st);
      GI_CUDA(cudaGetLastError());
      ctx->launches++;
      if (overlap_tail) {
        if (!R.estimate_range(R.qg, ng1, ng - ng1, GI_MAP_GLOBAL) || !R.estimate_range(R.qc, nc1, nc - nc1, GI_MAP_CAUSTIC)) return 1;
        if (!R.commit_range(R.qg, ng, GI_MAP_GLOBAL) || !R.commit_range(R.qc, nc, GI_MAP_CAUSTIC)) return 1;
      } else if (!R.drain(R.qg, ng, GI_MAP_GLOBAL) || !R.drain(R.qc, nc, GI_MAP_CAUSTIC)) return 1;
      host_paths.shadow += bs.shadow; host_paths.monte += bs.monte; host_paths.transmissive += bs.transmissive;
      host_paths.specular += bs.specular; host_paths.indirect += bs.indirect; host_paths.caustic += bs.caustic;
      item0 += m;
    }
  }
  if (p.dof_test > 1) {                                   // (zeros of the other ranks' columns stay zero)
    const uint64_t cnt = 3 * (uint64_t)sw * sh;
    scale_kernel<<<(unsigned)((cnt + 255) / 256), 256, 0, R.st>>>(accum, cnt, (double)p.dof_test);
    ctx->launches++;
  }
  GI_CUDA(cudaStreamSynchronize(R.st));
  ctx->clock.collect();
  return 0;
}

}  // namespace gi

using namespace gi;

// shared by gi_render and gi_render_float: returns the device accumulator of the supersampled frame
static int render_common(GiContext *ctx, const GiParams *params, int width, int height, int aa, double **accum_out,
                         int *sw_out, int *sh_out, GiRenderStats *stats) {
  if (!ctx || !params) { set_error("gi_render: null argument"); return 1; }
  if (!ctx->has_scene) { set_error("gi_render: no scene uploaded"); return 1; }
  if (width < 1 || height < 1 || aa < 0 || aa > 6) { set_error("gi_render: bad image size or anti-aliasing level"); return 1; }
  {
    const PhotonMapStore &gm = ctx->maps[GI_MAP_GLOBAL];
    const bool uses_global = gm.dev.n > 0 && (params->indirect_illum || params->direct_photon_illum || params->fast_global);
    if (uses_global && (params->irradiance_cache != 0) != gm.cached) {
      set_error(gm.cached ? "gi_render: the global map holds cached irradiance (built with -cache) but irradiance_cache is off"
                          : "gi_render: irradiance_cache is set but the global map was built without it (call gi_build_maps with the same params)");
      return 1;
    }
  }
  if (params->global_estimate_size < 1 || params->global_estimate_size > GI_MAX_K || params->caustic_estimate_size < 1 ||
      params->caustic_estimate_size > GI_MAX_K) { set_error("gi_render: estimate size must be in [1, 512]"); return 1; }
  for (int m = 0; m < 2; m++)
    if (ctx->maps[m].dev.n > 0 && !ctx->maps[m].built) { set_error("gi_render: photon map not built (call gi_build_maps)"); return 1; }
  const int aaf = 1 << aa;                                  // pow(2.0, aa), render.cpp:174
  const uint64_t sw = (uint64_t)width * aaf, sh = (uint64_t)height * aaf, npix = sw * sh;
  if (npix >= (1ull << 32)) { set_error("gi_render: supersampled frame too large"); return 1; }
  GI_CUDA(cudaSetDevice(ctx->device));
  Workspace *ws = workspace(ctx);
  if (!ws || !ws->fit_accum(3 * npix)) return 1;
  double *accum = ws->accum.as<double>(); RayStats *d_stats = ws->stats.as<RayStats>();
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  int rc = 1;
  RayStats hs[2]; RayStats paths; memset(&paths, 0, sizeof(paths));
  do {
    ctx->clock.reset_render();
    if (!cuda_ok(cudaMemsetAsync(accum, 0, sizeof(double) * 3 * npix, ctx->stream), "memset(accum)")) break;
    if (!cuda_ok(cudaMemsetAsync(d_stats, 0, 2 * sizeof(RayStats), ctx->stream), "memset(stats)")) break;
    cudaEventRecord(e0, ctx->stream);
    // sharding across GPUs: rank r of w renders the pixel columns x = r (mod w), the rest of the frame stays zero.
    // The communicator of the context decides; without one params->rank / world_size select the columns (test hook).
    const bool comm = comm_attached(ctx);
    const int world = comm ? comm_world(ctx) : (params->world_size > 1 ? params->world_size : 1);
    const int rank = comm ? comm_rank(ctx) : (params->world_size > 1 ? params->rank : 0);
    if (world > 64) { set_error("gi_render: at most 64 ranks"); break; }
    if (rank < 0 || rank >= world) { set_error("gi_render: rank outside [0, world_size)"); break; }
    if (render_accum(ctx, *params, (int)sw, (int)sh, aaf, rank, world, accum, d_stats, paths)) break;
    cudaEventRecord(e1, ctx->stream);
    if (!cuda_ok(cudaMemcpyAsync(hs, d_stats, 2 * sizeof(RayStats), cudaMemcpyDeviceToHost, ctx->stream), "D2H stats")) break;
    if (!cuda_ok(cudaStreamSynchronize(ctx->stream), "render sync")) break;
    rc = 0;
  } while (0);
  if (rc == 0 && stats) {
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    stats->screen_rays = hs[0].screen; stats->shadow_rays = hs[0].shadow + paths.shadow; stats->monte_rays = paths.monte;
    stats->transmissive_samples = paths.transmissive; stats->specular_samples = paths.specular;
    stats->indirect_samples = hs[0].indirect + paths.indirect; stats->caustic_samples = hs[0].caustic + paths.caustic;
    stats->render_ms = ms;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  if (rc) return 1;
  *accum_out = accum; *sw_out = (int)sw; *sh_out = (int)sh;
  return 0;
}

extern "C" {

int gi_render_float(GiContext *ctx, const GiParams *params, int width, int height, int aa, double *rgb_out,
                    GiRenderStats *stats) {
  if (!rgb_out) { set_error("gi_render_float: null output buffer"); return 1; }
  double *accum = nullptr; int sw = 0, sh = 0;
  if (render_common(ctx, params, width, height, aa, &accum, &sw, &sh, stats)) return 1;
  return cuda_ok(cudaMemcpy(rgb_out, accum, sizeof(double) * 3 * (size_t)sw * sh, cudaMemcpyDeviceToHost), "D2H frame") ? 0 : 1;
}

int gi_render(GiContext *ctx, const GiParams *params, int width, int height, int aa, uint8_t *rgb_out,
              GiRenderStats *stats) {
  const bool sharded = ctx && comm_attached(ctx) && comm_world(ctx) > 1;
  if (!rgb_out && !(sharded && comm_rank(ctx) != 0)) { set_error("gi_render: null output buffer"); return 1; }
  double *accum = nullptr; int sw = 0, sh = 0;
  if (render_common(ctx, params, width, height, aa, &accum, &sw, &sh, stats)) return 1;
  Workspace *ws = workspace(ctx);
  const size_t frame_bytes = (size_t)width * height * 3;
  if (!ws || !ws->fit_image(frame_bytes)) return 1;
  uint8_t *d_rgb = ws->image.as<uint8_t>();
  cudaStream_t st = ctx->stream;
  dim3 block(16, 16);
  if (!sharded) {
    dim3 grid((width + 15) / 16, (height + 15) / 16);
    ctx->clock.begin(ST_RESOLVE, st);
    resolve_kernel<<<grid, block, 0, st>>>(accum, width, height, 1 << aa, sh, d_rgb);
    ctx->clock.end(st);
    ctx->launches++;
    GI_CUDA(cudaGetLastError());
    GI_CUDA(cudaMemcpyAsync(rgb_out, d_rgb, frame_bytes, cudaMemcpyDeviceToHost, st));
    GI_CUDA(cudaStreamSynchronize(st));
    ctx->clock.collect();
    return 0;
  }
  // ---- multi-GPU: own column strip -> gather to rank 0 -> interleave; ray counters summed over the ranks
  const int world = comm_world(ctx), rank = comm_rank(ctx);
  std::vector<uint64_t> bytes_of(world), off(world);
  uint64_t total = 0;
  for (int r = 0; r < world; r++) { bytes_of[r] = 3ull * height * ((width - r + world - 1) / world); off[r] = total; total += bytes_of[r]; }
  if (!ws->fit_strips(frame_bytes + bytes_of[0])) return 1;              // [all strips (rank 0) | my strip]
  uint8_t *d_all = ws->strips.as<uint8_t>(), *d_mine = d_all + frame_bytes;
  uint64_t *d_off = ws->comm_tmp.as<uint64_t>(), *d_cnt = d_off + 64;
  const int ncols = (width - rank + world - 1) / world;
  if (ncols > 0) {
    dim3 grid((ncols + 15) / 16, (height + 15) / 16);
    ctx->clock.begin(ST_RESOLVE, st);
    resolve_strip_kernel<<<grid, block, 0, st>>>(accum, width, height, 1 << aa, sh, rank, world, d_mine);
    ctx->clock.end(st);
    ctx->launches++;
    GI_CUDA(cudaGetLastError());
  }
  ctx->clock.begin(ST_FRAME_EXCHANGE, st);
  if (!comm_gatherv_root0(ctx, d_mine, d_all, bytes_of.data(), off.data(), st)) return 1;
  ctx->clock.end(st);
  if (stats) {
    uint64_t h[7] = {stats->screen_rays, stats->shadow_rays, stats->monte_rays, stats->transmissive_samples, stats->specular_samples,
                     stats->indirect_samples, stats->caustic_samples};
    GI_CUDA(cudaMemcpyAsync(d_cnt, h, sizeof(h), cudaMemcpyHostToDevice, st));
    if (!comm_allreduce_sum_u64(ctx, d_cnt, 7, st)) return 1;
    GI_CUDA(cudaMemcpyAsync(h, d_cnt, sizeof(h), cudaMemcpyDeviceToHost, st));
    GI_CUDA(cudaStreamSynchronize(st));
    stats->screen_rays = h[0]; stats->shadow_rays = h[1]; stats->monte_rays = h[2]; stats->transmissive_samples = h[3];
    stats->specular_samples = h[4]; stats->indirect_samples = h[5]; stats->caustic_samples = h[6];
  } else {
    // every rank makes the same collective calls
    GI_CUDA(cudaMemsetAsync(d_cnt, 0, 7 * 8, st));
    if (!comm_allreduce_sum_u64(ctx, d_cnt, 7, st)) return 1;
  }
  if (rank == 0) {
    GI_CUDA(cudaMemcpyAsync(d_off, off.data(), 8 * (size_t)world, cudaMemcpyHostToDevice, st));
    dim3 grid((width + 15) / 16, (height + 15) / 16);
    assemble_kernel<<<grid, block, 0, st>>>(d_all, d_off, width, height, world, d_rgb);
    ctx->launches++;
    GI_CUDA(cudaGetLastError());
    GI_CUDA(cudaMemcpyAsync(rgb_out, d_rgb, frame_bytes, cudaMemcpyDeviceToHost, st));
  }
  GI_CUDA(cudaStreamSynchronize(st));
  ctx->clock.collect();
  return 0;
}

}  // extern "C"
