// csrc/photon_trace.cu -- gi_trace_photons: the reference's MapPhotons (src/photonmap.cpp:260-436) on the GPU.
//
// The reference gives every CPU thread a share of the photon budget and lets it emit photons light by light in
// adaptive rounds (Threadable_PhotonTracer, photonmap.cpp:144-257): each round emits
// (target - stored) / average_stores_per_photon / slowdown + 1 photons, split over the lights by power, and
// PhotonTrace (photontracer.cpp:28-176) follows every photon through its bounces with Russian roulette, storing it
// at diffuse surfaces.  Here one round is ONE kernel launch with a thread per emitted photon; stored photons are
// appended through an atomic cursor together with the key (emitted photon number, bounce), and once the rounds
// are over the photons are sorted by that key, so a map depends on the seed only -- not on scheduling, and not on the
// number of GPUs: with a communicator attached (comm.cu) rank r traces the photons t = r (mod world) of every round, the
// stored counts are all-reduced per round, and one ragged all-gather over NVLink replicates the photons at the end.
#include <cub/cub.cuh>
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "comm.h"
#include "context.h"
#include "kdtree_build.h"
#include "shade.cuh"
#include "wave.cuh"

namespace gi {

enum { STAGE_PHOTON = 8 };           // Philox stage = STAGE_PHOTON + map
enum { MAP_GLOBAL = 0, MAP_CAUSTIC = 1 };

struct PhotonSink { GiPhoton *photons; uint64_t *keys; unsigned long long *count; unsigned long long cap; };

// StorePhoton, photon_utils.cpp:40-66
GI_DEV void store_photon(const PhotonSink &sink, D3 power, D3 incident, D3 point, uint64_t key) {
  const unsigned long long i = atomicAdd(sink.count, 1ull);
  if (i >= sink.cap) return;
  const double PI = 3.14159265358979323846, TWO_PI = 6.28318530717958647692;
  GiPhoton ph;
  ph.position[0] = point.x; ph.position[1] = point.y; ph.position[2] = point.z;
  rgbe_encode(power, ph.rgbe);
  const int phi = (uint8_t)(int)(255.0 * (atan2(incident.y, incident.x) + PI) / (TWO_PI));
  const int theta = (uint8_t)(int)(255.0 * acos(fmin(1.0, fmax(-1.0, incident.z))) / PI);
  ph.direction = (uint16_t)(phi * 256 + theta);
  ph.pad = 0;
  sink.photons[i] = ph;
  sink.keys[i] = key;
}

// One iteration of PhotonTrace's bounce loop (photontracer.cpp:64-158) after its closest-hit query: `hit_element` < 0 =
// the ray left the scene.  Returns false when the photon's walk ends.  The buffered Philox outputs are dropped at the end,
// so that a photon parked between two bounces is described by its counter alone.
GI_DEV bool photon_shade(const ShadeEnv &env, RayD &ray, D3 &photon, bool &store, int iter, int map_type, uint64_t id, Rng &rng,
                         const PhotonSink &sink, int hit_element, D3 point, D3 normal) {
  const GiParams &p = env.p;
  if (hit_element < 0) return false;
  const Brdf b = load_brdf(env.materials[env.sc.elements[hit_element].material]);
  const D3 view = normalized(point - ray.o);                 // ray.o is the reference's ray_start
  const double cos_theta = dot(normal, -view);
  if (b.diffuse && store) store_photon(sink, photon, view, point, (id << 8) | (uint64_t)(iter & 255));
  double r_coeff = 0;
  if (p.fresnel && b.transparent) r_coeff = reflection_coeff(cos_theta, b.ir, p.ir_air);
  const double mc = max_channel(photon);
  const double prob_diffuse = max_channel(mul3(b.kd, photon)) / mc;
  double prob_transmission = max_channel(mul3(b.kt, photon)) / mc;
  const double prob_specular = (max_channel(mul3(b.ks, photon)) / mc) + r_coeff * prob_transmission;
  prob_transmission *= (1.0 - r_coeff);
  const double prob_total = prob_diffuse + prob_transmission + prob_specular + p.prob_absorb;
  double rnd = rng_double(rng);
  if (prob_total > 1.0) rnd *= prob_total;
  D3 bounce;
  if (rnd < prob_diffuse) {
    if (map_type == MAP_CAUSTIC) return false;
    store = true;
    bounce = diffuse_sample(normal, cos_theta, rng);
    photon = mul3(photon, b.kd / prob_diffuse);
  } else if (rnd < prob_diffuse + prob_transmission) {
    if (map_type == MAP_CAUSTIC) store = true;
    const D3 exact = transmissive_bounce(normal, view, cos_theta, b.ir, p.ir_air);
    bounce = p.distrib_transmissive ? specular_sample(exact, b.n, cos_theta, rng) : exact;
    photon = mul3(photon, (b.kt * (1.0 - r_coeff)) / prob_transmission);
  } else if (rnd < prob_diffuse + prob_transmission + prob_specular) {
    if (map_type == MAP_CAUSTIC) store = true;
    const D3 exact = reflective_bounce(normal, view, cos_theta);
    bounce = p.distrib_specular ? specular_sample(exact, b.n, cos_theta, rng) : exact;
    photon = mul3(photon, (b.ks + b.kt * r_coeff) / prob_specular);
  } else return false;
  ray.o = point + bounce * GI_EPS; ray.d = bounce;
  rng.left = 0;
  return true;
}

// closest hit + shading of one bounce in one thread (the finishing pass and the lane-refill kernel)
GI_DEV bool photon_bounce(const ShadeEnv &env, RayD &ray, D3 &photon, bool &store, int iter, int map_type, uint64_t id, Rng &rng,
                          const PhotonSink &sink) {
  HitRec h;
  closest_hit(env.sc, ray, h);
  return photon_shade(env, ray, photon, store, iter, map_type, id, rng, sink, h.prim < 0 ? -1 : h.element, h.p, h.n);
}

// basis of a disk with normal n: EmitPhotons, photontracer.cpp:204-215 / 296-307
GI_DEV void disk_basis(D3 n, double radius, D3 &u, D3 &v) {
  u = mk3(n.y, -n.x, 0);
  if (1.0 - fabs(n.z) < 0.1) u = mk3(n.z, 0, -n.x);
  v = cross(u, n);
  u = normalized(u) * radius; v = normalized(v) * radius;
}

struct Round {                       // one emission round (photonmap.cpp:165-173): photon t of the round comes from the
  int map_type;                      // light whose range [first[l], first[l+1]) holds t and has the number id0 + t
  uint64_t id0;
  const uint64_t *first;
  D3 centroid;
  uint32_t rank, world;              // this GPU takes the photons t = rank (mod world) of the round: work item w is t = w * world + rank
};

// The emission part of EmitPhotons (photontracer.cpp:182-360) for photon t of a round: seeds the photon's stream and
// samples its first ray.
GI_DEV void emit_photon(const ShadeEnv &env, const Round &rd, uint64_t t, Rng &rng, RayD &ray, D3 &photon) {
  int li = 0;
  while (li + 1 < env.n_lights && t >= rd.first[li + 1]) li++;
  const GiLight &l = env.lights[li];
  photon = ld3(l.color);                                              // NormalizeColor, graphics_utils.cpp:25-36
  const double total = photon.x + photon.y + photon.z;
  if (total > 0) photon = photon / total;
  rng_init(rng, env.p.seed, rd.id0 + t, STAGE_PHOTON + rd.map_type, 0);
  const D3 ln = ld3(l.direction);
  if (l.type == GI_LIGHT_DIR) {
    const D3 center = rd.centroid - ln * env.radius * 3.0;
    D3 u, v; disk_basis(ln, env.radius, u, v);
    double r1, r2;
    do { r1 = rng_double(rng) * 2.0 - 1.0; r2 = rng_double(rng) * 2.0 - 1.0; } while (r1 * r1 + r2 * r2 > 1.0);
    ray.o = u * r1 + v * r2 + center + ln * GI_EPS; ray.d = ln;
  } else if (l.type == GI_LIGHT_POINT) {
    double x, y, z;
    do { x = rng_double(rng) * 2.0 - 1.0; y = rng_double(rng) * 2.0 - 1.0; z = rng_double(rng) * 2.0 - 1.0; } while (x * x + y * y + z * z > 1.0);
    ray.o = ld3(l.position); ray.d = normalized(mk3(x, y, z));
  } else if (l.type == GI_LIGHT_SPOT) {
    const double cutoff = fabs(cos(l.cutoff));
    int attempts_left = 20;
    D3 dir;
    do { dir = specular_sample(ln, l.dropoff, 1.0, rng); } while (dot(dir, ln) < cutoff && attempts_left-- > 0);
    if (attempts_left == 0) dir = specular_sample(ln, l.dropoff, cutoff, rng);     // sic (photontracer.cpp:279-282)
    ray.o = ld3(l.position); ray.d = dir;
  } else if (l.type == GI_LIGHT_AREA) {
    D3 u, v; disk_basis(ln, l.radius, u, v);
    double r1, r2;
    do { r1 = rng_double(rng) * 2.0 - 1.0; r2 = rng_double(rng) * 2.0 - 1.0; } while (r1 * r1 + r2 * r2 > 1.0);
    ray.o = u * r1 + v * r2 + ld3(l.position) + ln * GI_EPS;
    ray.d = diffuse_sample(ln, 1.0, rng);
  } else {
    const D3 a1 = ld3(l.axis1) * l.len1, a2 = ld3(l.axis2) * l.len2;
    const double r1 = rng_double(rng) - 0.5, r2 = rng_double(rng) - 0.5;
    ray.o = a1 * r1 + a2 * r2 + ld3(l.position) + ln * GI_EPS;
    ray.d = diffuse_sample(ln, 1.0, rng);
  }
}

// Conservative test of a ray against the world boxes of every shape whose material can reflect or refract specularly
// (ks or kt not black).  A CAUSTIC photon stores nothing before its first specular / transmissive bounce and ends at
// a diffuse pick (photontracer.cpp:56-60, 106-110); a photon whose first ray misses all of these boxes can only meet
// surfaces with kt = ks = 0, where both non-diffuse probabilities are zero -- its walk ends at the first hit without
// side effects, whatever its random numbers.  Skipping it changes nothing but the time.
GI_DEV bool may_hit_specular(const ShadeEnv &env, const RayD &ray) {
  for (int i = 0; i < env.n_spec_boxes; i++) {
    const double *b = env.spec_boxes + 6 * i;
    double tmin = -1.0e-3, tmax = 1.0e30;
    bool ok = true;
    const double o[3] = {ray.o.x, ray.o.y, ray.o.z}, d[3] = {ray.d.x, ray.d.y, ray.d.z};
#pragma unroll
    for (int a = 0; a < 3; a++) {
      if (fabs(d[a]) < 1.0e-12) { ok = ok && o[a] >= b[a] && o[a] <= b[a + 3]; continue; }
      const double inv = 1.0 / d[a];
      const double t1 = (b[a] - o[a]) * inv, t2 = (b[a + 3] - o[a]) * inv;
      tmin = fmax(tmin, fmin(t1, t2)); tmax = fmin(tmax, fmax(t1, t2));
    }
    if (ok && tmin <= tmax) return true;
  }
  return false;
}

// CAUSTIC rounds, pass 1: thread per emitted photon; photons that may reach a specular surface are listed.
__global__ void __launch_bounds__(256) emit_filter_kernel(ShadeEnv env, Round rd, uint64_t w0, uint32_t n, uint32_t *__restrict__ list,
                                                          unsigned int *__restrict__ count) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  bool keep = false;
  if (i < n) {
    Rng rng; RayD ray; D3 photon;
    emit_photon(env, rd, (w0 + i) * rd.world + rd.rank, rng, ray, photon);
    keep = may_hit_specular(env, ray);
  }
  const unsigned m = __ballot_sync(0xffffffffu, keep);
  if (m) {
    const int lane = threadIdx.x & 31;
    unsigned base = 0;
    if (lane == __ffs(m) - 1) base = atomicAdd(count, (unsigned)__popc(m));
    base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
    if (keep) list[base + __popc(m & ((1u << lane) - 1u))] = i;
  }
}

// Persistent photon tracer.  Every lane walks one photon; a lane whose photon ends takes the next one at once, so a
// warp is never held up by its longest walk (most walks end at the first hit, a few go on for dozens of bounces
// inside glass).  Work is handed out in blocks of 256 photons per warp from one global cursor.  `list` (may be NULL)
// maps work item -> photon of the batch.  Compiled for 6 CTAs per SM (80 registers, 24 warps; 168 registers unbounded):
// the walks are latency-bound like the render's path kernel -- jensen's photon pass 129 -> 110 ms, cornell unchanged.
// n_items_dev (may be NULL): the number of work items, left in device memory by emit_filter_kernel -- the host does not
// wait for the filter to know it.
__global__ void __launch_bounds__(128, 6) trace_kernel(ShadeEnv env, Round rd, uint64_t w0, uint64_t n_items_host,
                                                    const unsigned int *__restrict__ n_items_dev,
                                                    const uint32_t *__restrict__ list, unsigned long long *cursor, PhotonSink sink) {
  const int lane = threadIdx.x & 31;
  const uint64_t n_items = n_items_dev ? (uint64_t)*n_items_dev : n_items_host;
  bool active = false, store = false;
  int iter = 0;
  uint64_t id = 0;
  RayD ray; D3 photon; Rng rng;
  unsigned long long wbase = 0; int wleft = 0;
  bool exhausted = false;
  for (;;) {
    unsigned idle = __ballot_sync(0xffffffffu, !active);
    while (idle && !(exhausted && wleft == 0)) {
      if (wleft == 0) {
        unsigned long long b = 0;
        if (lane == 0) b = atomicAdd(cursor, 256ull);
        b = __shfl_sync(0xffffffffu, b, 0);
        if (b >= n_items) { exhausted = true; break; }
        wbase = b; wleft = (int)((n_items - b) < 256ull ? (n_items - b) : 256ull);
      }
      const int mine = __popc(idle & ((1u << lane) - 1u));
      if (!active && mine < wleft) {
        const unsigned long long w = wbase + mine;
        const uint64_t t = (w0 + (list ? (uint64_t)list[w] : (uint64_t)w)) * rd.world + rd.rank;
        emit_photon(env, rd, t, rng, ray, photon);
        rng.left = 0;                                          // as the wavefront kernels: a photon's stream is a function of its counter
        id = rd.id0 + t; iter = 0; active = true;
        store = rd.map_type == MAP_GLOBAL && !env.p.fast_global;
      }
      const int taken = min(__popc(idle), wleft);
      wbase += taken; wleft -= taken;
      idle = __ballot_sync(0xffffffffu, !active);
    }
    if (!__any_sync(0xffffffffu, active)) break;
    if (active) {
      active = photon_bounce(env, ray, photon, store, iter, rd.map_type, id, rng, sink);
      iter++;
      if (iter >= env.p.max_photon_depth) active = false;
    }
  }
}

// ------------------------------------------------------------------------------------------------ wavefront photon walks
// The same cut into bounce passes as the render's secondary paths (render.cu): emission parks every photon's first ray,
// then passes of  wave_hit_kernel (closest hit only)  +  photon_shade_kernel (store, roulette, next ray; survivors are
// compacted into the next pass's queue with a warp ballot), and a finishing pass that walks what is still alive a thread
// each.  A parked photon reuses the 128-byte PathRec: scale[] = power, tw[0] = photon number (bits), pad = store flag.
GI_DEV void park_photon(PathRec *r, const RayD &ray, D3 power, uint64_t id, const Rng &rng, int iter, bool store) {
  double4 *o = (double4 *)r;
  o[0] = make_double4(ray.o.x, ray.o.y, ray.o.z, ray.d.x);
  o[1] = make_double4(ray.d.y, ray.d.z, power.x, power.y);
  o[2] = make_double4(power.z, __longlong_as_double((long long)id), 0.0, 0.0);
  uint4 *u = (uint4 *)(r) + 6;
  u[0] = make_uint4(rng.c0, rng.c1, rng.c2, rng.c3);
  u[1] = make_uint4(0u, 0u, (uint32_t)iter, store ? 1u : 0u);
}
GI_DEV void unpark_photon(const PathRec *r, uint64_t seed, RayD &ray, D3 &power, uint64_t &id, Rng &rng, int &iter, bool &store) {
  const double4 *i = (const double4 *)r;
  const double4 a = i[0], b = i[1], c = i[2];
  ray.o = mk3(a.x, a.y, a.z); ray.d = mk3(a.w, b.x, b.y); power = mk3(b.z, b.w, c.x);
  id = (uint64_t)__double_as_longlong(c.y);
  const uint4 *u = (const uint4 *)(r) + 6;
  const uint4 x = u[0], y = u[1];
  rng.k0 = (uint32_t)seed; rng.k1 = (uint32_t)(seed >> 32);
  rng.c0 = x.x; rng.c1 = x.y; rng.c2 = x.z; rng.c3 = x.w; rng.left = 0;
  iter = (int)y.z; store = y.w != 0u;
}

// thread per work item [first, first + count) of the batch: emission; every photon is parked (slot = its index)
__global__ void __launch_bounds__(128, 6) photon_emit_kernel(ShadeEnv env, Round rd, uint64_t w0, uint32_t first, uint32_t count,
                                                             const uint32_t *__restrict__ list, PathWave next) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t == 0) { next.count[CLASS_SPEC] = count; next.count[CLASS_DIFF] = 0u; }
  if (t >= count) return;
  const uint32_t w = first + t;
  const uint64_t tt = (w0 + (list ? (uint64_t)list[w] : (uint64_t)w)) * rd.world + rd.rank;
  Rng rng; RayD ray; D3 photon;
  emit_photon(env, rd, tt, rng, ray, photon);
  rng.left = 0;
  park_photon(next.rec + t, ray, photon, rd.id0 + tt, rng, 0, rd.map_type == MAP_GLOBAL && !env.p.fast_global);
}

template <bool FINISH>
__global__ void __launch_bounds__(128, 6) photon_shade_kernel(ShadeEnv env, int map_type, PathWave cur, PathWave next, PhotonSink sink) {
  const unsigned n = cur.count[CLASS_SPEC];
  const unsigned lane = threadIdx.x & 31u;
  for (unsigned base = (blockIdx.x * blockDim.x + threadIdx.x) - lane; base < n; base += gridDim.x * blockDim.x) {
    const unsigned t = base + lane;
    bool live = t < n, store = false;
    RayD ray; D3 photon; uint64_t id = 0; Rng rng; int iter = 0;
    if (live) {
      unpark_photon(cur.rec + t, env.p.seed, ray, photon, id, rng, iter, store);
      const double4 *hi = (const double4 *)(cur.hit + t);
      const double4 a = hi[0], b = hi[1];
      live = photon_shade(env, ray, photon, store, iter, map_type, id, rng, sink, __double2loint(b.z), mk3(a.x, a.y, a.z), mk3(a.w, b.x, b.y));
      iter++;
      if (iter >= env.p.max_photon_depth) live = false;
      if (FINISH)
        while (live) {
          live = photon_bounce(env, ray, photon, store, iter, map_type, id, rng, sink);
          iter++;
          if (iter >= env.p.max_photon_depth) live = false;
        }
    }
    if (!FINISH) {
      const unsigned m = __ballot_sync(0xffffffffu, live);
      if (m) {
        const int leader = __ffs(m) - 1;
        unsigned at = 0;
        if ((int)lane == leader) at = atomicAdd(next.count + CLASS_SPEC, (unsigned)__popc(m));
        at = __shfl_sync(0xffffffffu, at, leader);
        if (live) park_photon(next.rec + at + __popc(m & ((1u << lane) - 1u)), ray, photon, id, rng, iter, store);
      }
    }
  }
}

// the power scale of MapPhotons (photonmap.cpp:339-361): decode, multiply, re-encode
__global__ void scale_power_kernel(GiPhoton *ph, uint64_t n, double factor) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  uint8_t e[4] = {ph[t].rgbe[0], ph[t].rgbe[1], ph[t].rgbe[2], ph[t].rgbe[3]};
  const D3 c = rgbe_decode(e) * factor;
  rgbe_encode(c, e);
  ph[t].rgbe[0] = e[0]; ph[t].rgbe[1] = e[1]; ph[t].rgbe[2] = e[2]; ph[t].rgbe[3] = e[3];
}

__global__ void permute_kernel(const GiPhoton *__restrict__ in, const uint32_t *__restrict__ order, uint64_t n, GiPhoton *__restrict__ out) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) out[t] = in[order[t]];
}
__global__ void iota_kernel(uint32_t *a, uint64_t n) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) a[t] = (uint32_t)t;
}

// LightPower, graphics_utils.cpp:223-258
static double light_power(const GiLight &l, double scene_radius) {
  const double PI = 3.14159265358979323846, TWO_PI = 6.28318530717958647692;
  double area = 1.0, flux = 4.0 * PI;
  if (l.type == GI_LIGHT_DIR) { area = PI * pow(scene_radius, 2.0); flux = 1.0; }
  else if (l.type == GI_LIGHT_AREA) { area = PI * pow(l.radius, 2.0); flux /= 2.0; }
  else if (l.type == GI_LIGHT_RECT) {
    const D3 a1 = ld3(l.axis1) * l.len1, a2 = ld3(l.axis2) * l.len2;
    area = length(cross(a1, a2)); flux /= 2.0;
  } else if (l.type == GI_LIGHT_SPOT) flux = TWO_PI / (l.dropoff + 1.0) * (1.0 - pow(cos(l.cutoff), l.dropoff + 1.0));
  return (l.color[0] + l.color[1] + l.color[2]) * area * flux;
}

namespace {
// stream-ordered pool allocations (gi_create keeps freed pool memory): a frame traces its maps every time, and
// cudaMalloc / cudaFree of a few hundred MB per map cost more than the kernels on a busy device
struct Buf {
  void *p = nullptr; cudaStream_t st = nullptr;
  ~Buf() { release(); }
  void release() { if (p) cudaFreeAsync(p, st); p = nullptr; }
  bool alloc(size_t bytes, cudaStream_t s) { release(); st = s; return cuda_ok(cudaMallocAsync(&p, bytes ? bytes : 1, s), "cudaMallocAsync(photon trace)"); }
  template <class T> T *as() const { return (T *)p; }
};
}  // namespace

// One map: the adaptive rounds of Threadable_PhotonTracer, shared by all ranks of the context's communicator.
// A round emits n photons numbered id0 .. id0 + n - 1 over ALL ranks; rank r traces the photons t = r (mod world).  The
// stored counts are all-reduced after every round, so every rank computes the same next goal and the union of the
// photons is the same for any number of ranks.  At the end the photons (with their keys) are all-gathered and sorted
// by (photon number, bounce) on every rank: the map in ctx->maps[map].raw is byte-identical everywhere.
static int trace_map(GiContext *ctx, const ShadeEnv &env_in, int map_type, int target, const std::vector<double> &light_powers,
                     double total_power, uint64_t &emitted_out, int &rounds_out) {
  ShadeEnv env = env_in;
  cudaStream_t st = ctx->stream;
  const int nl = env.n_lights;
  const uint32_t world = (uint32_t)comm_world(ctx), rank = (uint32_t)comm_rank(ctx);
  unsigned long long cap = ((unsigned long long)target + (unsigned long long)target / 2) / world + 65536;     // this rank's sink
  const bool filter = map_type == MAP_CAUSTIC && env.n_spec_boxes >= 0;
  const uint64_t batch = 1ull << 26;                          // work items per launch (bounds the survivor list)
  Buf photons, keys, count, first, cursor, list;
  if (!photons.alloc(sizeof(GiPhoton) * cap, st) || !keys.alloc(8 * cap, st) || !count.alloc(16, st) || !first.alloc(8 * (size_t)(nl + 1), st) ||
      !cursor.alloc(16, st) || !list.alloc(filter ? 4 * batch : 4, st))
    return 1;
  GI_CUDA(cudaMemsetAsync(count.p, 0, 16, st));
  // wavefront photon walks (GI_PHOTON_MEGAKERNEL=1: the lane-refill kernel of round 1, for A/B): parked photons of one
  // sub-batch, ping / pong, with the closest hits of the current pass
  const bool megakernel = getenv("GI_PHOTON_MEGAKERNEL") && atoi(getenv("GI_PHOTON_MEGAKERNEL")) != 0;
  const uint32_t wcap = 1u << 21;
  int wave_passes = 4;
  if (const char *e = getenv("GI_PHOTON_WAVE_PASSES")) wave_passes = std::min(std::max(atoi(e), 0), 64);
  Buf wrec[2], whit[2], wcount;
  PathWave wv[2];
  if (!megakernel) {
    if (!wrec[0].alloc(sizeof(PathRec) * (size_t)wcap, st) || !wrec[1].alloc(sizeof(PathRec) * (size_t)wcap, st) ||
        !whit[0].alloc(sizeof(WaveHit) * (size_t)wcap, st) || !whit[1].alloc(sizeof(WaveHit) * (size_t)wcap, st) || !wcount.alloc(16, st))
      return 1;
    for (int i = 0; i < 2; i++) { wv[i].rec = wrec[i].as<PathRec>(); wv[i].hit = whit[i].as<WaveHit>(); wv[i].count = wcount.as<unsigned int>() + 2 * i; wv[i].cap = wcap; }
    GI_CUDA(cudaMemsetAsync(wcount.p, 0, 16, st));
  }
  unsigned long long stored = 0, stored_local = 0;            // over all ranks / in this rank's sink
  uint64_t emitted = 0;
  // photonmap.cpp:157-196 (global) / :213-250 (caustic)
  double average_bounce_rate = map_type == MAP_GLOBAL ? 4.0 : (double)env.p.max_photon_depth;
  double slowdown_factor = 1.0;
  int attempts_left = 10;
  const D3 centroid = ld3(ctx->centroid);
  while (stored < (unsigned long long)target && attempts_left > 0) {
    const int emit_goal = (int)((int)(target - (long long)stored) / average_bounce_rate / slowdown_factor + 1);
    std::vector<uint64_t> h_first(nl + 1, 0);
    for (int i = 0; i < nl; i++) {
      int num = (int)ceil(emit_goal * (light_powers[i] / total_power));
      if (!ctx->lights[i].active || num < 0) num = 0;          // EmitPhotons returns at once for an inactive light,
      h_first[i + 1] = h_first[i] + (uint64_t)num;             // but photons_assigned still counts its share (:169-172)
    }
    uint64_t assigned = 0;
    for (int i = 0; i < nl; i++) assigned += (uint64_t)std::max(0, (int)ceil(emit_goal * (light_powers[i] / total_power)));
    const uint64_t n = h_first[nl];
    const uint64_t n_local = n > rank ? (n - rank + world - 1) / world : 0;      // photons t = rank (mod world), t < n
    for (;;) {                                               // repeated only if this rank's sink was too small
      GI_CUDA(cudaMemcpyAsync(first.p, h_first.data(), 8 * (size_t)(nl + 1), cudaMemcpyHostToDevice, st));
      PhotonSink sink{photons.as<GiPhoton>(), keys.as<uint64_t>(), count.as<unsigned long long>(), cap};
      Round rd; rd.map_type = map_type; rd.id0 = emitted; rd.first = first.as<uint64_t>(); rd.centroid = centroid;
      rd.rank = rank; rd.world = world;
      for (uint64_t w0 = 0; w0 < n_local; w0 += batch) {
        const uint64_t m = std::min(batch, n_local - w0);
        GI_CUDA(cudaMemsetAsync(cursor.p, 0, 16, st));
        unsigned int *kept = (unsigned int *)(cursor.as<unsigned long long>() + 1);
        if (filter) {
          ctx->clock.begin(ST_EMIT_FILTER, st);
          emit_filter_kernel<<<(unsigned)((m + 255) / 256), 256, 0, st>>>(env, rd, w0, (uint32_t)m, list.as<uint32_t>(), kept);
          ctx->clock.end(st);
          ctx->launches++;
        }
        if (megakernel) {
          // persistent: 8 CTAs of 4 warps per SM; with the filter the item count stays on the device (no host round trip)
          ctx->clock.begin(ST_TRACE, st);
          const uint64_t want = (m + 127) / 128, gcap = (uint64_t)ctx->sm_count * 8;
          trace_kernel<<<(unsigned)std::max<uint64_t>(1, std::min(want, gcap)), 128, 0, st>>>(
              env, rd, w0, m, filter ? kept : nullptr, filter ? list.as<uint32_t>() : nullptr, cursor.as<unsigned long long>(), sink);
          ctx->clock.end(st);
          ctx->launches++;
          GI_CUDA(cudaGetLastError());
          continue;
        }
        // wavefront: the survivors (or all work items) in sub-batches of `wcap` photons
        uint64_t items = m;
        if (filter) {
          unsigned int h_kept = 0;
          GI_CUDA(cudaMemcpyAsync(&h_kept, kept, 4, cudaMemcpyDeviceToHost, st));
          GI_CUDA(cudaStreamSynchronize(st));
          ctx->clock.collect();
          items = h_kept;
        }
        ctx->clock.begin(ST_TRACE, st);
        for (uint64_t s0 = 0; s0 < items; s0 += wcap) {
          const uint32_t c = (uint32_t)std::min<uint64_t>(wcap, items - s0);
          photon_emit_kernel<<<(c + 127) / 128, 128, 0, st>>>(env, rd, w0, (uint32_t)s0, c, filter ? list.as<uint32_t>() : nullptr, wv[0]);
          const unsigned wgrid = (unsigned)std::min<uint64_t>((c + 127) / 128, (uint64_t)ctx->sm_count * 6);
          int cur = 0;
          for (int pass = 0; pass < wave_passes; pass++, cur ^= 1) {
            GI_CUDA(cudaMemsetAsync(wv[cur ^ 1].count, 0, 8, st));
            wave_hit_kernel<6><<<wgrid, 128, 0, st>>>(env.sc, wv[cur]);
            photon_shade_kernel<false><<<wgrid, 128, 0, st>>>(env, map_type, wv[cur], wv[cur ^ 1], sink);
          }
          wave_hit_kernel<6><<<wgrid, 128, 0, st>>>(env.sc, wv[cur]);
          photon_shade_kernel<true><<<wgrid, 128, 0, st>>>(env, map_type, wv[cur], wv[cur ^ 1], sink);
          ctx->launches += 3 + 2 * (uint64_t)wave_passes;
          GI_CUDA(cudaGetLastError());
        }
        ctx->clock.end(st);
      }
      unsigned long long now = 0;
      GI_CUDA(cudaMemcpyAsync(&now, count.p, 8, cudaMemcpyDeviceToHost, st));
      GI_CUDA(cudaStreamSynchronize(st));
      ctx->clock.collect();
      if (now <= cap) { stored_local = now; break; }
      // grow the sink, keep what earlier rounds stored, and run this round again (same streams -> same photons)
      const unsigned long long ncap = std::max(now + now / 2, cap * 2);
      Buf np, nk;
      if (!np.alloc(sizeof(GiPhoton) * ncap, st) || !nk.alloc(8 * ncap, st)) return 1;
      GI_CUDA(cudaMemcpyAsync(np.p, photons.p, sizeof(GiPhoton) * stored_local, cudaMemcpyDeviceToDevice, st));
      GI_CUDA(cudaMemcpyAsync(nk.p, keys.p, 8 * stored_local, cudaMemcpyDeviceToDevice, st));
      GI_CUDA(cudaMemcpyAsync(count.p, &stored_local, 8, cudaMemcpyHostToDevice, st));
      GI_CUDA(cudaStreamSynchronize(st));
      std::swap(photons.p, np.p); std::swap(keys.p, nk.p);
      cap = ncap;
    }
    // the round's stored count over all ranks drives the next goal (photonmap.cpp:175-192)
    if (world > 1) {
      uint64_t *tot = count.as<uint64_t>() + 1;
      GI_CUDA(cudaMemcpyAsync(tot, count.p, 8, cudaMemcpyDeviceToDevice, st));
      if (!comm_allreduce_sum_u64(ctx, tot, 1, st)) return 1;
      GI_CUDA(cudaMemcpyAsync(&stored, tot, 8, cudaMemcpyDeviceToHost, st));
      GI_CUDA(cudaStreamSynchronize(st));
    } else stored = stored_local;
    emitted += assigned;
    rounds_out++;
    if (stored > 0 && emitted > 0) {
      average_bounce_rate = ((double)stored) / emitted;
      const double progress = map_type == MAP_GLOBAL ? (double)stored / emitted : (double)stored / target;
      slowdown_factor = progress < 0.75 ? 2.0 : 1.0;
    } else {
      average_bounce_rate /= 2.0;
      attempts_left--;
    }
  }
  emitted_out = emitted;
  PhotonMapStore &m = ctx->maps[map_type];
  free_map(m);
  m.dev.n = stored;
  if (stored == 0) return 0;
  if (stored >= (1ull << 31)) { set_error("gi_trace_photons: more than 2^31 - 1 photons stored"); return 1; }
  // replicate: counts, then ONE ragged all-gather of the photons and one of their keys (comm.cu)
  Buf all_photons, all_keys, counts_dev;
  const GiPhoton *src_photons = photons.as<GiPhoton>(); uint64_t *src_keys = keys.as<uint64_t>();
  if (world > 1) {
    ctx->clock.begin(ST_PHOTON_EXCHANGE, st);
    std::vector<uint64_t> h_counts(world, 0), bytes32(world), off32(world), bytes8(world), off8(world);
    if (!counts_dev.alloc(8 * (size_t)(world + 1), st)) return 1;
    GI_CUDA(cudaMemcpyAsync(counts_dev.as<uint64_t>() + world, count.p, 8, cudaMemcpyDeviceToDevice, st));
    if (!comm_allgather_u64(ctx, counts_dev.as<uint64_t>() + world, counts_dev.as<uint64_t>(), 1, st)) return 1;
    GI_CUDA(cudaMemcpyAsync(h_counts.data(), counts_dev.p, 8 * (size_t)world, cudaMemcpyDeviceToHost, st));
    GI_CUDA(cudaStreamSynchronize(st));
    uint64_t total = 0;
    for (uint32_t r = 0; r < world; r++) { off32[r] = 32 * total; off8[r] = 8 * total; bytes32[r] = 32 * h_counts[r]; bytes8[r] = 8 * h_counts[r]; total += h_counts[r]; }
    if (total != stored) { set_error("gi_trace_photons: ranks disagree on the stored photon count"); return 1; }
    if (!all_photons.alloc(32 * total, st) || !all_keys.alloc(8 * total, st)) return 1;
    if (!comm_allgatherv(ctx, photons.p, all_photons.p, bytes32.data(), off32.data(), st) ||
        !comm_allgatherv(ctx, keys.p, all_keys.p, bytes8.data(), off8.data(), st)) return 1;
    src_photons = all_photons.as<GiPhoton>(); src_keys = all_keys.as<uint64_t>();
    ctx->clock.end(st);
  }
  ctx->clock.begin(ST_PHOTON_SORT, st);
  // deterministic order: sort by (emitted photon number, bounce) -- only the key bits that can be set take part
  Buf keys2, order, order2, tmp;
  if (!keys2.alloc(8 * stored, st) || !order.alloc(4 * stored, st) || !order2.alloc(4 * stored, st)) return 1;
  const unsigned blocks = (unsigned)((stored + 255) / 256);
  iota_kernel<<<blocks, 256, 0, st>>>(order.as<uint32_t>(), stored);
  int end_bit = 8; while (end_bit < 64 && (emitted >> (end_bit - 8)) != 0) end_bit++;
  size_t tmp_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, src_keys, keys2.as<uint64_t>(), order.as<uint32_t>(),
                                  order2.as<uint32_t>(), (int)stored, 0, end_bit, st);
  if (!tmp.alloc(tmp_bytes, st)) return 1;
  GI_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, src_keys, keys2.as<uint64_t>(), order.as<uint32_t>(),
                                          order2.as<uint32_t>(), (int)stored, 0, end_bit, st));
  if (!alloc_raw(ctx, m, stored)) return 1;
  permute_kernel<<<blocks, 256, 0, st>>>(src_photons, order2.as<uint32_t>(), stored, (GiPhoton *)m.raw);
  ctx->clock.end(st);
  ctx->launches += 4;
  GI_CUDA(cudaGetLastError());
  GI_CUDA(cudaStreamSynchronize(st));
  ctx->clock.collect();
  return 0;
}

}  // namespace gi

using namespace gi;

extern "C" {

int gi_scale_photon_power(GiContext *ctx, int map, double factor) {
  if (!ctx || map < 0 || map > 1) { set_error("gi_scale_photon_power: bad argument"); return 1; }
  PhotonMapStore &m = ctx->maps[map];
  if (m.built) { set_error("gi_scale_photon_power: the map is already built"); return 1; }
  if (m.dev.n == 0) return 0;
  GI_CUDA(cudaSetDevice(ctx->device));
  scale_power_kernel<<<(unsigned)((m.dev.n + 255) / 256), 256, 0, ctx->stream>>>((GiPhoton *)m.raw, m.dev.n, factor);
  ctx->launches++;
  GI_CUDA(cudaGetLastError());
  GI_CUDA(cudaStreamSynchronize(ctx->stream));
  return 0;
}

int gi_trace_photons(GiContext *ctx, const GiParams *params, GiPhotonStats *stats) {
  if (!ctx || !params) { set_error("gi_trace_photons: null argument"); return 1; }
  if (!ctx->has_scene) { set_error("gi_trace_photons: no scene uploaded"); return 1; }
  GiPhotonStats s; memset(&s, 0, sizeof(s));
  s.indirect_illum = params->indirect_illum; s.caustic_illum = params->caustic_illum; s.direct_photon_illum = params->direct_photon_illum;
  if (stats) *stats = s;
  GI_CUDA(cudaSetDevice(ctx->device));
  free_map(ctx->maps[0]); free_map(ctx->maps[1]);
  const int nl = (int)ctx->lights.size();
  if (nl <= 0) return 0;                                       // MapPhotons corner case (photonmap.cpp:263-265)
  std::vector<double> light_powers(nl, 0.0);
  double total_power = 0;
  for (int i = 0; i < nl; i++) {
    if (!ctx->lights[i].active) continue;
    light_powers[i] = light_power(ctx->lights[i], ctx->radius);
    total_power += light_powers[i];
  }
  s.total_power = total_power;
  if (total_power <= 0) { if (stats) *stats = s; return 0; }   // (:285-287)

  ShadeEnv env;
  env.sc = ctx->scene; env.materials = ctx->scene.materials; env.p = *params;
  env.ambient = ld3(ctx->ambient); env.background = ld3(ctx->background); env.eye = ld3(ctx->camera.eye);
  env.radius = ctx->radius; env.n_lights = nl;
  env.spec_boxes = ctx->d_spec_boxes; env.n_spec_boxes = ctx->n_spec_boxes;
  if (getenv("GI_NO_SPEC_FILTER")) env.n_spec_boxes = -1;
  GiLight *d_lights = nullptr;
  GI_CUDA(cudaMalloc(&d_lights, sizeof(GiLight) * nl));
  env.lights = d_lights;
  int rc = 1;
  cudaEvent_t e0, e1, e2; cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&e2);
  do {
    if (!cuda_ok(cudaMemcpyAsync(d_lights, ctx->lights.data(), sizeof(GiLight) * nl, cudaMemcpyHostToDevice, ctx->stream), "H2D lights")) break;
    // The budget split of MapPhotons (:289-321) -- COUNT / THREADS photons per worker -- becomes a split of every
    // emission round over the ranks of the communicator (trace_map); without one this context traces everything.
    if (params->world_size > 1 && !comm_attached(ctx)) {
      set_error("gi_trace_photons: params->world_size > 1 needs a communicator (gi_comm_init_rank / gi_comm_init_all)");
      break;
    }
    const bool want_global = params->indirect_illum || params->direct_photon_illum, want_caustic = params->caustic_illum != 0;
    ctx->clock.reset_photon();
    cudaEventRecord(e0, ctx->stream);
    int rounds = 0;
    if (want_global && trace_map(ctx, env, MAP_GLOBAL, params->global_photon_count, light_powers, total_power, s.global_emitted, rounds)) break;
    if (want_caustic && trace_map(ctx, env, MAP_CAUSTIC, params->caustic_photon_count, light_powers, total_power, s.caustic_emitted, rounds)) break;
    cudaEventRecord(e1, ctx->stream);
    s.rounds = rounds;
    s.global_stored = ctx->maps[0].dev.n; s.caustic_stored = ctx->maps[1].dev.n;
    // power scale, flag updates and kd-tree builds of MapPhotons (:339-377); counts are global, every rank does the same
    if (want_global && s.global_stored == 0) { s.indirect_illum = 0; s.direct_photon_illum = 0; }
    if (want_caustic && s.caustic_stored == 0) s.caustic_illum = 0;
    if (s.global_stored && gi_scale_photon_power(ctx, 0, total_power / (double)s.global_emitted)) break;
    if (s.caustic_stored && gi_scale_photon_power(ctx, 1, total_power / (double)s.caustic_emitted)) break;
    if (gi_build_maps(ctx, params)) break;
    cudaEventRecord(e2, ctx->stream);
    if (!cuda_ok(cudaStreamSynchronize(ctx->stream), "trace sync")) break;
    float a = 0, b = 0; cudaEventElapsedTime(&a, e0, e1); cudaEventElapsedTime(&b, e1, e2);
    s.trace_ms = a; s.build_ms = b;
    ctx->clock.collect();
    if (ctx->clock.on) ctx->clock.ms[ST_KD_BUILD] = b;
    s.cache_ms = ctx->maps[0].cached ? ctx->maps[0].cache_ms : 0.0;
    s.build_ms -= s.cache_ms < s.build_ms ? s.cache_ms : 0.0;          // the reference times the two separately (:378,:410)
    rc = 0;
  } while (0);
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2);
  cudaFree(d_lights);
  if (rc == 0 && stats) *stats = s;
  return rc;
}

}  // extern "C"
