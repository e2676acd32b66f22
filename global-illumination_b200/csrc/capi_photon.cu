// csrc/capi_photon.cu -- photon-map entry points of include/gi_b200.h: load / dump / build / gather.
//
// gi_load_photons + gi_build_maps stand in for GLOBAL_PHOTONS / CAUSTIC_PHOTONS and the two
// `new R3Kdtree<Photon*>` of MapPhotons (src/photonmap.cpp:364-377); gi_gather_batch stands in for a loop of
// EstimateRadiance calls (src/utils/photon_utils.cpp:72-162).  The photon file format ("GIPM": u32 magic,
// u32 version, u64 n, n x 32-byte Photon) is new -- the reference never writes its maps -- and is shared with
// oracle/ref_harness so that both sides can be fed the same map.
#include <cuda_runtime.h>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <vector>
#include "comm.h"
#include "context.h"
#include "gather.h"
#include "gi_math.cuh"
#include "irradiance_cache.h"
#include "kdtree_build.h"
#include "shade.cuh"

namespace gi {

static const uint32_t PMAP_MAGIC = 0x4d504947u;   // "GIPM"
static const uint32_t PMAP_VERSION = 1u;

// Every array of a map comes from the device's stream-ordered pool (cudaMallocAsync on the context's stream; gi_create
// keeps freed pool memory): a frame -- and a bench step -- loads and rebuilds its maps every time, and cudaFree +
// cudaMalloc of the 1.7 GB a 10^7-photon map holds cost 20 ms, half as much again as the kd-tree build itself.
// The device is synchronised first, like cudaFree does implicitly: a gather launched on a caller's stream may still read
// the arrays.
void free_map(PhotonMapStore &m) {
  void *ptrs[] = {(void *)m.dev.px, (void *)m.dev.py, (void *)m.dev.pz, (void *)m.dev.node, (void *)m.dev.shade, (void *)m.dev.orig,
                  (void *)m.dev.cbox, (void *)m.dev.gpos, (void *)m.dev.srec, (void *)m.dev.hint, m.raw};
  bool any = false;
  for (void *p : ptrs) any = any || p;
  if (any) {
    cudaDeviceSynchronize();
    for (void *p : ptrs) if (p) cudaFreeAsync(p, (cudaStream_t)m.stream);
  }
  m = PhotonMapStore();
}

bool alloc_raw(GiContext *ctx, PhotonMapStore &m, uint64_t n) {
  m.stream = (void *)ctx->stream;
  return cuda_ok(cudaMallocAsync(&m.raw, sizeof(GiPhoton) * (n ? n : 1), ctx->stream), "cudaMallocAsync(photons)");
}

// BuildDirectionLookupTable (src/utils/photon_utils.cpp:253-272), built on the host with the same libm calls and
// kept in HBM (1.5 MB, L2-resident): entry [phi * 256 + theta] = normalised (x, y, z)
const double *direction_lut(GiContext *ctx) {
  if (ctx->d_lut) return ctx->d_lut;
  std::vector<double> h(65536 * 3);
  const double PI = 3.14159265358979323846, TWO_PI = 6.28318530717958647692;   // RN_PI / RN_TWO_PI (RNScalar.h)
  for (int phi = 0; phi < 256; phi++)
    for (int theta = 0; theta < 256; theta++) {
      const double tp = (phi * TWO_PI / 255.0) - PI, tt = (theta * PI / 255.0);
      D3 v = normalized(mk3(sin(tt) * cos(tp), sin(tt) * sin(tp), cos(tt)));
      double *o = &h[3 * (size_t)(256 * phi + theta)];
      o[0] = v.x; o[1] = v.y; o[2] = v.z;
    }
  double *d = nullptr;
  if (!cuda_ok(cudaMalloc(&d, h.size() * 8), "cudaMalloc(lut)")) return nullptr;
  if (!cuda_ok(cudaMemcpy(d, h.data(), h.size() * 8, cudaMemcpyHostToDevice), "cudaMemcpy(lut)")) { cudaFree(d); return nullptr; }
  ctx->d_lut = d;
  return d;
}

static int pow2ceil(int v) { int p = 1; while (p < v) p <<= 1; return p; }

}  // namespace gi

using namespace gi;

namespace gi {
// GiGatherQueryCompact -> GiGatherQuery as RayTrace sets up its EstimateRadiance calls (raytracer.cpp:138-167, 174-190)
__global__ void expand_queries_kernel(const GiGatherQueryCompact *__restrict__ c, uint64_t n, const DevMaterial *__restrict__ mats, int n_mats,
                                      GiGatherQuery *__restrict__ out) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const GiGatherQueryCompact q = c[t];
  const D3 normal = ld3(q.n), view = ld3(q.view);
  const double cos_theta = dot(normal, -view);
  const D3 bounce = reflective_bounce(normal, view, cos_theta);
  GiGatherQuery o;
  for (int a = 0; a < 3; a++) { o.p[a] = q.p[a]; o.n[a] = q.n[a]; }
  o.bounce[0] = bounce.x; o.bounce[1] = bounce.y; o.bounce[2] = bounce.z; o.cos_theta = cos_theta;
  const int mi = q.material >= 0 && q.material < n_mats ? q.material : 0;
  for (int a = 0; a < 3; a++) { o.kd[a] = mats[mi].kd[a]; o.ks[a] = mats[mi].ks[a]; }
  o.shininess = mats[mi].shininess;
  out[t] = o;
}
}  // namespace gi

extern "C" {

int gi_load_photons(GiContext *ctx, int map, const GiPhoton *photons, uint64_t n) {
  if (!ctx || map < 0 || map > 1) { set_error("gi_load_photons: bad argument"); return 1; }
  if (n > 0 && !photons) { set_error("gi_load_photons: null photon buffer"); return 1; }
  if (n >= ((uint64_t)1 << 31)) { set_error("gi_load_photons: at most 2^31 - 1 photons per map"); return 1; }
  GI_CUDA(cudaSetDevice(ctx->device));
  PhotonMapStore &m = ctx->maps[map];
  free_map(m);
  m.dev.n = n;
  if (n == 0) return 0;
  if (!alloc_raw(ctx, m, n)) return 1;
  GI_CUDA(cudaMemcpyAsync(m.raw, photons, sizeof(GiPhoton) * n, cudaMemcpyHostToDevice, ctx->stream));
  GI_CUDA(cudaStreamSynchronize(ctx->stream));
  return 0;
}

// device-resident variant used by the multi-GPU path (photons all-gathered by NCCL into d_photons)
int gi_load_photons_device(GiContext *ctx, int map, const GiPhoton *d_photons, uint64_t n) {
  if (!ctx || map < 0 || map > 1) { set_error("gi_load_photons_device: bad argument"); return 1; }
  if (n >= ((uint64_t)1 << 31)) { set_error("gi_load_photons_device: at most 2^31 - 1 photons per map"); return 1; }
  GI_CUDA(cudaSetDevice(ctx->device));
  PhotonMapStore &m = ctx->maps[map];
  free_map(m);
  m.dev.n = n;
  if (n == 0) return 0;
  if (!alloc_raw(ctx, m, n)) return 1;
  GI_CUDA(cudaMemcpyAsync(m.raw, d_photons, sizeof(GiPhoton) * n, cudaMemcpyDeviceToDevice, ctx->stream));
  GI_CUDA(cudaStreamSynchronize(ctx->stream));
  return 0;
}

// Multi-GPU load: every rank of the context's communicator passes ITS share of the photons (device memory); the shares
// are concatenated in rank order on every rank by one ragged all-gather over NVLink straight into the map's photon array
// (no padding, no staging copy).  Without a communicator this is gi_load_photons_device.
int gi_load_photons_sharded_device(GiContext *ctx, int map, const GiPhoton *d_share, uint64_t n_share) {
  if (!ctx || map < 0 || map > 1) { set_error("gi_load_photons_sharded_device: bad argument"); return 1; }
  if (n_share > 0 && !d_share) { set_error("gi_load_photons_sharded_device: null photon buffer"); return 1; }
  const int world = comm_world(ctx);
  if (world <= 1) return gi_load_photons_device(ctx, map, d_share, n_share);
  GI_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  PhotonMapStore &m = ctx->maps[map];
  free_map(m);
  uint64_t *d_counts = nullptr;                           // [world] gathered + [1] mine
  GI_CUDA(cudaMallocAsync((void **)&d_counts, 8 * (size_t)(world + 1), st));
  int rc = 1;
  do {
    std::vector<uint64_t> h(world, 0), bytes_of(world), off(world);
    if (!cuda_ok(cudaMemcpyAsync(d_counts + world, &n_share, 8, cudaMemcpyHostToDevice, st), "H2D share size")) break;
    if (!comm_allgather_u64(ctx, d_counts + world, d_counts, 1, st)) break;
    if (!cuda_ok(cudaMemcpyAsync(h.data(), d_counts, 8 * (size_t)world, cudaMemcpyDeviceToHost, st), "D2H share sizes")) break;
    if (!cuda_ok(cudaStreamSynchronize(st), "share sizes sync")) break;
    uint64_t total = 0;
    for (int r = 0; r < world; r++) { off[r] = sizeof(GiPhoton) * total; bytes_of[r] = sizeof(GiPhoton) * h[r]; total += h[r]; }
    if (total >= ((uint64_t)1 << 31)) { set_error("gi_load_photons_sharded_device: at most 2^31 - 1 photons per map"); break; }
    m.dev.n = total;
    if (total == 0) { rc = 0; break; }
    if (!alloc_raw(ctx, m, total)) { m.dev.n = 0; break; }
    if (!comm_allgatherv(ctx, d_share, m.raw, bytes_of.data(), off.data(), st)) break;
    if (!cuda_ok(cudaStreamSynchronize(st), "photon all-gather sync")) break;
    rc = 0;
  } while (0);
  cudaFreeAsync(d_counts, st);
  return rc;
}

int gi_photon_count(GiContext *ctx, int map, uint64_t *n_out) {
  if (!ctx || map < 0 || map > 1 || !n_out) { set_error("gi_photon_count: bad argument"); return 1; }
  *n_out = ctx->maps[map].dev.n;
  return 0;
}

int gi_get_photons(GiContext *ctx, int map, GiPhoton *out, uint64_t capacity) {
  if (!ctx || map < 0 || map > 1) { set_error("gi_get_photons: bad argument"); return 1; }
  PhotonMapStore &m = ctx->maps[map];
  if (capacity < m.dev.n) { set_error("gi_get_photons: buffer too small"); return 1; }
  if (m.dev.n == 0) return 0;
  GI_CUDA(cudaSetDevice(ctx->device));
  GI_CUDA(cudaMemcpy(out, m.raw, sizeof(GiPhoton) * m.dev.n, cudaMemcpyDeviceToHost));
  return 0;
}

// device-to-device variant for the multi-GPU path: the rank's photons go straight into the NCCL send buffer
int gi_get_photons_device(GiContext *ctx, int map, GiPhoton *d_out, uint64_t capacity, void *stream) {
  if (!ctx || map < 0 || map > 1) { set_error("gi_get_photons_device: bad argument"); return 1; }
  PhotonMapStore &m = ctx->maps[map];
  if (capacity < m.dev.n) { set_error("gi_get_photons_device: buffer too small"); return 1; }
  if (m.dev.n == 0) return 0;
  if (!d_out) { set_error("gi_get_photons_device: null buffer"); return 1; }
  GI_CUDA(cudaSetDevice(ctx->device));
  GI_CUDA(cudaMemcpyAsync(d_out, m.raw, sizeof(GiPhoton) * m.dev.n, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return 0;
}

// The header is not trusted: magic, version and the photon count against the bytes that are really in the file are
// checked before anything is allocated, and no C++ exception crosses the extern "C" boundary.
int gi_load_photons_file(GiContext *ctx, int map, const char *path) {
  if (!ctx || map < 0 || map > 1) { set_error("gi_load_photons_file: bad argument"); return 1; }
  FILE *f = path ? fopen(path, "rb") : nullptr;
  if (!f) { set_error(std::string("Unable to open photon file ") + (path ? path : "(null)")); return 1; }
  int rc = 1;
  try {
    uint32_t hdr[2]; uint64_t n = 0;
    if (fread(hdr, 4, 2, f) != 2 || hdr[0] != PMAP_MAGIC || fread(&n, 8, 1, f) != 1) throw std::string("Bad photon file ");
    if (hdr[1] != PMAP_VERSION) throw std::string("Unsupported photon file version in ");
    if (n >= ((uint64_t)1 << 31)) throw std::string("Bad photon count in photon file ");
    const long body = ftell(f);
    if (body < 0 || fseek(f, 0, SEEK_END) != 0) throw std::string("Bad photon file ");
    const long end = ftell(f);
    if (end < body || (uint64_t)(end - body) < n * sizeof(GiPhoton) || fseek(f, body, SEEK_SET) != 0) throw std::string("Truncated photon file ");
    std::vector<GiPhoton> ph(n);
    if (n && fread(ph.data(), sizeof(GiPhoton), n, f) != n) throw std::string("Truncated photon file ");
    rc = gi_load_photons(ctx, map, ph.data(), n);
  } catch (const std::string &what) { set_error(what + path);
  } catch (const std::exception &e) { set_error(std::string("gi_load_photons_file: ") + e.what()); }
  fclose(f);
  return rc;
}

int gi_dump_photons_file(GiContext *ctx, int map, const char *path) {
  if (!ctx || map < 0 || map > 1 || !path) { set_error("gi_dump_photons_file: bad argument"); return 1; }
  try {
    const uint64_t n = ctx->maps[map].dev.n;
    std::vector<GiPhoton> ph(n);
    if (gi_get_photons(ctx, map, ph.data(), n)) return 1;
    FILE *f = fopen(path, "wb");
    if (!f) { set_error(std::string("Unable to open photon file ") + path); return 1; }
    uint32_t hdr[2] = {PMAP_MAGIC, PMAP_VERSION};
    bool ok = fwrite(hdr, 4, 2, f) == 2 && fwrite(&n, 8, 1, f) == 1 && (n == 0 || fwrite(ph.data(), sizeof(GiPhoton), n, f) == n);
    ok = (fclose(f) == 0) && ok;
    if (!ok) { set_error(std::string("Unable to write photon file ") + path); return 1; }
  } catch (const std::exception &e) { set_error(std::string("gi_dump_photons_file: ") + e.what()); return 1; }
  return 0;
}

int gi_build_maps(GiContext *ctx, const GiParams *params) {
  if (!ctx) { set_error("gi_build_maps: null context"); return 1; }
  GI_CUDA(cudaSetDevice(ctx->device));
  const double *lut = direction_lut(ctx);
  if (!lut) return 1;
  for (int k = 0; k < 2; k++) {
    PhotonMapStore &m = ctx->maps[k];
    if (m.built || m.dev.n == 0) continue;
    GI_CUDA(build_photon_map(ctx, m, m.dev.n, lut));
  }
  // photonmap.cpp:380-411: with -cache the global photons' colours become irradiance estimates
  if (params && params->irradiance_cache && (params->indirect_illum || params->direct_photon_illum) &&
      ctx->maps[GI_MAP_GLOBAL].dev.n > 0 && build_irradiance_cache(ctx, *params))
    return 1;
  return 0;
}

int gi_cached_radiance_batch_device(GiContext *ctx, const GiGatherQuery *d_queries, uint64_t n, double r, double *d_rgb, void *stream) {
  if (!ctx) { set_error("gi_cached_radiance_batch: null context"); return 1; }
  PhotonMapStore &m = ctx->maps[GI_MAP_GLOBAL];
  if (m.dev.n > 0 && !m.built) { set_error("gi_cached_radiance_batch: photon map not built (call gi_build_maps)"); return 1; }
  if (m.dev.n > 0 && !m.cached) {
    set_error("gi_cached_radiance_batch: the global map holds no irradiance cache (call gi_build_maps with irradiance_cache set)");
    return 1;
  }
  if (n == 0) return 0;
  if (!d_queries || !d_rgb) { set_error("gi_cached_radiance_batch: null buffer"); return 1; }
  GI_CUDA(cudaSetDevice(ctx->device));
  const double *lut = direction_lut(ctx);
  if (!lut) return 1;
  GI_CUDA(launch_cached_radiance(ctx, m.dev, lut, d_queries, n, r, d_rgb, (cudaStream_t)stream));
  return 0;
}

int gi_cached_radiance_batch(GiContext *ctx, const GiGatherQuery *queries, uint64_t n, double r, double *rgb_out) {
  if (!ctx) { set_error("gi_cached_radiance_batch: null context"); return 1; }
  if (n == 0) return 0;
  if (!queries || !rgb_out) { set_error("gi_cached_radiance_batch: null buffer"); return 1; }
  GI_CUDA(cudaSetDevice(ctx->device));
  const uint64_t chunk = 1ull << 20;
  int rc = 0;
  for (uint64_t o = 0, c = 0; o < n && rc == 0; o += chunk, c++) {
    const uint64_t m = n - o < chunk ? n - o : chunk;
    const int slot = (int)(c & 1);
    cudaStream_t st = ctx->pipe[slot];
    GiScratch *sc = ctx->stage[slot];
    if (!sc[0].fit(sizeof(GiGatherQuery) * chunk) || !sc[1].fit(sizeof(double) * 3 * chunk)) {
      set_error("gi_cached_radiance_batch: out of device memory"); rc = 1; break;
    }
    if (!cuda_ok(cudaMemcpyAsync(sc[0].p, queries + o, sizeof(GiGatherQuery) * m, cudaMemcpyHostToDevice, st), "H2D queries") ||
        gi_cached_radiance_batch_device(ctx, (const GiGatherQuery *)sc[0].p, m, r, (double *)sc[1].p, st) != 0 ||
        !cuda_ok(cudaMemcpyAsync(rgb_out + 3 * o, sc[1].p, sizeof(double) * 3 * m, cudaMemcpyDeviceToHost, st), "D2H rgb"))
      rc = 1;
  }
  if (!cuda_ok(cudaStreamSynchronize(ctx->pipe[0]), "cache sync") || !cuda_ok(cudaStreamSynchronize(ctx->pipe[1]), "cache sync")) rc = 1;
  return rc;
}

int gi_gather_batch_device(GiContext *ctx, int map, const GiGatherQuery *d_queries, uint64_t n, int k, double r,
                           int filter, const GiParams *params, int32_t *d_idx, int32_t *d_count, double *d_rgb,
                           void *stream) {
  if (!ctx || map < 0 || map > 1) { set_error("gi_gather_batch: bad argument"); return 1; }
  if (k < 1 || k > GI_MAX_K) { set_error("gi_gather_batch: estimate size must be in [1, 512]"); return 1; }
  if (filter < 0 || filter > 2) { set_error("gi_gather_batch: unknown filter"); return 1; }
  PhotonMapStore &m = ctx->maps[map];
  if (m.dev.n > 0 && !m.built) { set_error("gi_gather_batch: photon map not built (call gi_build_maps)"); return 1; }
  if (n == 0) return 0;
  GI_CUDA(cudaSetDevice(ctx->device));
  const double *lut = direction_lut(ctx);
  if (!lut) return 1;
  GatherArgs ga;
  ga.k = k; ga.filter = filter; ga.r2 = r * r;
  // candidate slots: k + half as many again (at least 96) + 32; the index output sorts the
  // buffer with a bitonic network and needs a power of two
  ga.cap = d_idx ? (pow2ceil(2 * k) < 256 ? 256 : pow2ceil(2 * k)) : ((k + (k / 2 > 96 ? k / 2 : 96) + 31) / 32) * 32 + 32;
  ga.fa = params ? params->filter_const_a : 0.918; ga.fb = params ? params->filter_const_b : 1.953;
  ga.fk = params ? params->filter_const_k : 1.0;
  ga.guess_factor = 1.1;
  ga.slack = ga.cap;
  ga.max_warps = 0; ga.key_rows = 0; ga.range = 0;
  if (const char *e = getenv("GI_GATHER_RANGE")) ga.range = atoi(e);
  if (const char *e = getenv("GI_GATHER_WARPS")) ga.max_warps = atoi(e);
  if (const char *e = getenv("GI_CUT_SLACK")) ga.slack = atoi(e);
  if (ga.slack < 32) ga.slack = 32;
  if (const char *e = getenv("GI_GUESS_FACTOR")) ga.guess_factor = atof(e);
  cudaStream_t st = (cudaStream_t)stream;      // NULL = the CUDA default stream, as documented
  // batches are walked along a Morton curve (query_order.cu); small ones are not worth the sort
  const uint32_t *order = nullptr;
  bool sort = n >= 4096;
  if (const char *e = getenv("GI_SORT_QUERIES")) sort = atoi(e) != 0;
  if (sort) {
    bool failed = false;
    order = morton_order(ctx, m, d_queries, n, st, &failed);
    if (failed) return 1;
  }
  GI_CUDA(launch_gather(ctx, m.dev, lut, d_queries, order, n, ga, d_idx, d_count, d_rgb, st));
  return 0;
}

int gi_gather_batch(GiContext *ctx, int map, const GiGatherQuery *queries, uint64_t n, int k, double r, int filter,
                    const GiParams *params, int32_t *idx_out, int32_t *count_out, double *rgb_out) {
  if (!ctx) { set_error("gi_gather_batch: null context"); return 1; }
  if (n == 0) return 0;
  if (!queries) { set_error("gi_gather_batch: null query buffer"); return 1; }
  if (k < 1 || k > GI_MAX_K) { set_error("gi_gather_batch: estimate size must be in [1, 512]"); return 1; }
  GI_CUDA(cudaSetDevice(ctx->device));
  // two-slot pipeline like gi_intersect_batch.  Chunks grow from 2^18 to 2^21 queries (2^20 with index output: its
  // staging buffer is k times larger): a small first chunk starts the kernels early, large later ones keep the
  // launches efficient (a persistent launch pays a ramp and a tail of a few hundred queries per warp).
  uint64_t cmax = idx_out ? 1ull << 20 : 1ull << 21;
  if (const char *e = getenv("GI_PIPE_CHUNK_LOG2")) { const int l = atoi(e); if (l >= 12 && l <= 24) cmax = 1ull << l; }
  uint64_t chunk = cmax < (1ull << 18) ? cmax : 1ull << 18;
  int rc = 0;
  for (uint64_t o = 0, c = 0; o < n && rc == 0; o += chunk, chunk = chunk * 2 < cmax ? chunk * 2 : cmax, c++) {
    const uint64_t m = n - o < chunk ? n - o : chunk;
    const int slot = (int)(c & 1);
    cudaStream_t st = ctx->pipe[slot];
    GiScratch *sc = ctx->stage[slot];
    const uint64_t room = n < cmax ? n : cmax;             // sized once for the largest chunk: growing means cudaFree + cudaMalloc
    if (!sc[0].fit(sizeof(GiGatherQuery) * room) || (rgb_out && !sc[1].fit(sizeof(double) * 3 * room)) ||
        (count_out && !sc[2].fit(sizeof(int32_t) * room)) || (idx_out && !sc[3].fit(sizeof(int32_t) * room * k))) {
      set_error("gi_gather_batch: out of device memory"); rc = 1; break;
    }
    GiGatherQuery *d_q = (GiGatherQuery *)sc[0].p;
    double *d_rgb = rgb_out ? (double *)sc[1].p : nullptr;
    int32_t *d_cnt = count_out ? (int32_t *)sc[2].p : nullptr, *d_idx = idx_out ? (int32_t *)sc[3].p : nullptr;
    if (!cuda_ok(cudaMemcpyAsync(d_q, queries + o, sizeof(GiGatherQuery) * m, cudaMemcpyHostToDevice, st), "H2D queries") ||
        gi_gather_batch_device(ctx, map, d_q, m, k, r, filter, params, d_idx, d_cnt, d_rgb, st) != 0 ||
        (idx_out && !cuda_ok(cudaMemcpyAsync(idx_out + o * k, d_idx, sizeof(int32_t) * m * k, cudaMemcpyDeviceToHost, st), "D2H idx")) ||
        (count_out && !cuda_ok(cudaMemcpyAsync(count_out + o, d_cnt, sizeof(int32_t) * m, cudaMemcpyDeviceToHost, st), "D2H count")) ||
        (rgb_out && !cuda_ok(cudaMemcpyAsync(rgb_out + 3 * o, d_rgb, sizeof(double) * 3 * m, cudaMemcpyDeviceToHost, st), "D2H rgb")))
      rc = 1;
  }
  if (!cuda_ok(cudaStreamSynchronize(ctx->pipe[0]), "gather sync") || !cuda_ok(cudaStreamSynchronize(ctx->pipe[1]), "gather sync")) rc = 1;
  return rc;
}

int gi_gather_batch_compact(GiContext *ctx, int map, const GiGatherQueryCompact *queries, uint64_t n, int k, double r, int filter,
                            const GiParams *params, int32_t *count_out, double *rgb_out) {
  if (!ctx) { set_error("gi_gather_batch_compact: null context"); return 1; }
  if (!ctx->has_scene) { set_error("gi_gather_batch_compact: no scene uploaded (the material indices refer to it)"); return 1; }
  if (n == 0) return 0;
  if (!queries) { set_error("gi_gather_batch_compact: null query buffer"); return 1; }
  if (k < 1 || k > GI_MAX_K) { set_error("gi_gather_batch_compact: estimate size must be in [1, 512]"); return 1; }
  GI_CUDA(cudaSetDevice(ctx->device));
  const uint64_t cmax = 1ull << 21;                        // the pipeline of gi_gather_batch, 80 instead of 136 bytes per query in
  uint64_t chunk = 1ull << 18;
  int rc = 0;
  for (uint64_t o = 0, c = 0; o < n && rc == 0; o += chunk, chunk = chunk * 2 < cmax ? chunk * 2 : cmax, c++) {
    const uint64_t m = n - o < chunk ? n - o : chunk;
    const int slot = (int)(c & 1);
    cudaStream_t st = ctx->pipe[slot];
    GiScratch *sc = ctx->stage[slot];
    const uint64_t room = n < cmax ? n : cmax;
    if (!sc[0].fit(sizeof(GiGatherQuery) * room) || (rgb_out && !sc[1].fit(sizeof(double) * 3 * room)) ||
        (count_out && !sc[2].fit(sizeof(int32_t) * room)) || !sc[3].fit(sizeof(GiGatherQueryCompact) * room)) {
      set_error("gi_gather_batch_compact: out of device memory"); rc = 1; break;
    }
    GiGatherQuery *d_q = (GiGatherQuery *)sc[0].p; GiGatherQueryCompact *d_c = (GiGatherQueryCompact *)sc[3].p;
    double *d_rgb = rgb_out ? (double *)sc[1].p : nullptr;
    int32_t *d_cnt = count_out ? (int32_t *)sc[2].p : nullptr;
    if (!cuda_ok(cudaMemcpyAsync(d_c, queries + o, sizeof(GiGatherQueryCompact) * m, cudaMemcpyHostToDevice, st), "H2D queries")) { rc = 1; break; }
    expand_queries_kernel<<<(unsigned)((m + 255) / 256), 256, 0, st>>>(d_c, m, ctx->scene.materials, ctx->scene.n_materials, d_q);
    ctx->launches++;
    if (!cuda_ok(cudaGetLastError(), "expand_queries_kernel") ||
        gi_gather_batch_device(ctx, map, d_q, m, k, r, filter, params, nullptr, d_cnt, d_rgb, st) != 0 ||
        (count_out && !cuda_ok(cudaMemcpyAsync(count_out + o, d_cnt, sizeof(int32_t) * m, cudaMemcpyDeviceToHost, st), "D2H count")) ||
        (rgb_out && !cuda_ok(cudaMemcpyAsync(rgb_out + 3 * o, d_rgb, sizeof(double) * 3 * m, cudaMemcpyDeviceToHost, st), "D2H rgb")))
      rc = 1;
  }
  if (!cuda_ok(cudaStreamSynchronize(ctx->pipe[0]), "gather sync") || !cuda_ok(cudaStreamSynchronize(ctx->pipe[1]), "gather sync")) rc = 1;
  return rc;
}

}  // extern "C"
