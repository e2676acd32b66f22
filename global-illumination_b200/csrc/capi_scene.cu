// csrc/capi_scene.cu -- context, scene upload and the batched closest-hit entry points of include/gi_b200.h.
//
// gi_upload_scene stands in for building the R3Scene graph the reference traverses (R3Scene::ReadFile leaves
// the graph in host memory, R3Scene.cpp:514-587); gi_intersect_batch stands in for a loop of
// R3Scene::Intersects calls (R3Scene.cpp:471-479), e.g. the pixel loop of render.cpp:90-125.
#include <cuda_runtime.h>
#include <cstdio>
#include <algorithm>
#include <cstring>
#include <mutex>
#include "bvh_build.h"
#include "comm.h"
#include "context.h"
#include "isect.cuh"
#include "scene_prep.h"

namespace gi {

static thread_local std::string g_error;
void set_error(const std::string &msg) { g_error = msg; }
bool cuda_ok(cudaError_t e, const char *what) {
  if (e == cudaSuccess) return true;
  set_error(std::string("CUDA error: ") + cudaGetErrorString(e) + " in " + what);
  return false;
}

void free_scene(GiContext *ctx) {
  for (void *p : ctx->scene_allocs) cudaFree(p);
  ctx->scene_allocs.clear();
  ctx->has_scene = false;
  ctx->d_spec_boxes = nullptr; ctx->n_spec_boxes = -1;
  memset(&ctx->scene, 0, sizeof(ctx->scene));
}

template <class T> static bool upload(GiContext *ctx, const std::vector<T> &h, const T **d) {
  *d = nullptr;
  size_t bytes = sizeof(T) * (h.empty() ? 1 : h.size());
  void *p = nullptr;
  if (!cuda_ok(cudaMalloc(&p, bytes), "cudaMalloc(scene)")) return false;
  ctx->scene_allocs.push_back(p);
  if (!h.empty() && !cuda_ok(cudaMemcpyAsync(p, h.data(), sizeof(T) * h.size(), cudaMemcpyHostToDevice, ctx->stream),
                             "cudaMemcpy(scene)"))
    return false;
  *d = (const T *)p;
  return true;
}

// one thread per ray; rays and hits are the AoS records of the C ABI (48 B in, 72 B out)
__global__ void __launch_bounds__(128) intersect_kernel(DevScene sc, const GiRay *__restrict__ rays, uint64_t n,
                                                        GiHit *__restrict__ hits) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    const double2 *rp = (const double2 *)(rays + i);        // 48-byte records are 16-byte aligned
    const double2 r0 = rp[0], r1 = rp[1], r2 = rp[2];
    RayD ray; ray.o = mk3(r0.x, r0.y, r1.x); ray.d = mk3(r1.y, r2.x, r2.y);
    HitRec h;
    closest_hit(sc, ray, h);
    GiHit g;
    g.hit = h.prim >= 0 ? 1 : 0; g.shape = h.shape; g.prim = h.prim; g.element = h.element; g.t = h.t;
    g.p[0] = h.p.x; g.p[1] = h.p.y; g.p[2] = h.p.z; g.n[0] = h.n.x; g.n[1] = h.n.y; g.n[2] = h.n.z;
    hits[i] = g;
  }
}

}  // namespace gi

using namespace gi;

namespace gi {
__global__ void compact_hits_kernel(const GiHit *__restrict__ h, uint64_t n, GiHitCompact *__restrict__ out) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  GiHitCompact c; c.prim = h[t].hit ? h[t].prim : -1; c.shape = h[t].hit ? h[t].shape : -1; c.t = h[t].t;
  out[t] = c;
}
}  // namespace gi

extern "C" {

int gi_abi_version(void) { return GI_ABI_VERSION; }
const char *gi_last_error(void) { return g_error.c_str(); }

int gi_create(int device, GiContext **out) {
  if (!out) { set_error("gi_create: null output pointer"); return 1; }
  *out = nullptr;
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count <= 0) {
    set_error(std::string("gi_create: no CUDA device (") + cudaGetErrorString(e) + "); this library has no CPU path");
    return 1;
  }
  if (device < 0 || device >= count) { set_error("gi_create: device index out of range"); return 1; }
  GI_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  GI_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) { set_error("gi_create: this library is built for sm_100a (B200) only"); return 1; }
  GiContext *ctx = new GiContext();
  ctx->device = device; ctx->sm_count = prop.multiProcessorCount;
  {
    // keep memory freed to the device's stream-ordered pool (kd-tree build temporaries) instead of returning it to the driver
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, ctx->device) == cudaSuccess) {
      unsigned long long keep = ~0ull;
      cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
  }
  if (!cuda_ok(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking), "cudaStreamCreate") ||
      !cuda_ok(cudaStreamCreateWithFlags(&ctx->pipe[0], cudaStreamNonBlocking), "cudaStreamCreate") ||
      !cuda_ok(cudaStreamCreateWithFlags(&ctx->pipe[1], cudaStreamNonBlocking), "cudaStreamCreate")) { delete ctx; return 1; }
  *out = ctx;
  return 0;
}

void gi_destroy(GiContext *ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  comm_release(ctx);
  free_scene(ctx);
  free_map(ctx->maps[0]); free_map(ctx->maps[1]);
  if (ctx->render_ws && ctx->render_ws_free) ctx->render_ws_free(ctx->render_ws);
  for (int i = 0; i < 2; i++) {
    for (int j = 0; j < 4; j++) ctx->stage[i][j].release();
    for (int j = 0; j < 5; j++) ctx->order_ws[i][j].release();
    if (ctx->pipe[i]) cudaStreamDestroy(ctx->pipe[i]);
  }
  for (int j = 0; j < 5; j++) ctx->order_ws[2][j].release();
  for (int j = 0; j < 3; j++) ctx->small_ws[j].release();
  ctx->clock.release();
  cudaFree(ctx->gather_counters);
  cudaFree(ctx->d_lut);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

int gi_upload_scene(GiContext *ctx, const GiScene *scene) {
  if (!ctx || !scene) { set_error("gi_upload_scene: null argument"); return 1; }
  GI_CUDA(cudaSetDevice(ctx->device));
  free_scene(ctx);
  PreparedScene ps; std::string err;
  if (!prepare_scene(*scene, ps, err)) { set_error("gi_upload_scene: " + err); return 1; }
  DevScene d; memset(&d, 0, sizeof(d));
  const float *d_boxes = nullptr;
  bool ok = upload(ctx, ps.nodes, &d.nodes) && upload(ctx, ps.elements, &d.elements) && upload(ctx, ps.shapes, &d.shapes) &&
            upload(ctx, ps.materials, &d.materials) && upload(ctx, ps.prim_a, &d.prim_a) &&
            upload(ctx, ps.prim_b, &d.prim_b) && upload(ctx, ps.prim_meta, &d.prim_meta) &&
            upload(ctx, ps.prim_boxes, &d_boxes);
  if (!ok) { free_scene(ctx); return 1; }
  d.n_prims = scene->n_prims; d.n_shapes = scene->n_shapes; d.n_elements = scene->n_elements;
  d.n_nodes = scene->n_nodes; d.n_materials = scene->n_materials; d.has_mesh = ps.has_mesh; d.abs_max = ps.abs_max;
  d.n_bvh_nodes = scene->n_prims > 1 ? scene->n_prims - 1 : 1;
  float4 *bvh = nullptr;
  if (!cuda_ok(cudaMalloc(&bvh, sizeof(float4) * 4 * (size_t)d.n_bvh_nodes), "cudaMalloc(bvh)")) { free_scene(ctx); return 1; }
  ctx->scene_allocs.push_back(bvh);
  int bvh_depth = 0;
  if (!cuda_ok(cudaMemsetAsync(bvh, 0, sizeof(float4) * 4 * (size_t)d.n_bvh_nodes, ctx->stream), "memset(bvh)") ||
      !cuda_ok(build_bvh(d_boxes, scene->n_prims, ps.lo, ps.hi, bvh, ctx->stream, &bvh_depth), "build_bvh") ||
      !cuda_ok(cudaStreamSynchronize(ctx->stream), "upload sync")) { free_scene(ctx); return 1; }
  // closest_hit / find_poisoned_meshes keep at most one pending sibling per level: a deeper tree would drop subtrees
  if (bvh_depth > GI_STACK) {
    free_scene(ctx);
    set_error("gi_upload_scene: BVH depth " + std::to_string(bvh_depth) + " exceeds the traversal stack (" + std::to_string(GI_STACK) + ")");
    return 1;
  }
  ctx->bvh_depth = bvh_depth;
  ctx->launches += scene->n_prims > 1 ? 7 : 1;
  d.bvh = bvh;
  ctx->scene = d;
  // world boxes of the shapes that can bounce a photon specularly (caustic-photon prefilter, photon_trace.cu)
  {
    std::vector<double> boxes;
    const double margin = 1.0e-4 * (1.0 + (double)ps.abs_max);
    for (int si = 0; si < scene->n_shapes; si++) {
      const DevShape &sh = ps.shapes[si];
      const int mat = ps.elements[sh.element].material;
      if (mat < 0 || mat >= scene->n_materials) continue;
      const GiMaterial &m = scene->materials[mat];
      const bool spec = m.ks[0] != 0 || m.ks[1] != 0 || m.ks[2] != 0 || m.kt[0] != 0 || m.kt[1] != 0 || m.kt[2] != 0;
      if (!spec || sh.n_prims <= 0) continue;
      double b[6] = {1e300, 1e300, 1e300, -1e300, -1e300, -1e300};
      for (int pi = sh.first_prim; pi < sh.first_prim + sh.n_prims; pi++)
        for (int a = 0; a < 3; a++) {
          b[a] = std::min(b[a], (double)ps.prim_boxes[(size_t)pi * 6 + a] - margin);
          b[a + 3] = std::max(b[a + 3], (double)ps.prim_boxes[(size_t)pi * 6 + 3 + a] + margin);
        }
      boxes.insert(boxes.end(), b, b + 6);
    }
    ctx->n_spec_boxes = (int)(boxes.size() / 6);
    ctx->d_spec_boxes = nullptr;
    if (ctx->n_spec_boxes > 64) ctx->n_spec_boxes = -1;          // too many to test one by one: no prefilter
    else if (ctx->n_spec_boxes > 0) {
      const double *dp = nullptr;
      if (!upload(ctx, boxes, &dp) || !cuda_ok(cudaStreamSynchronize(ctx->stream), "upload sync")) { free_scene(ctx); return 1; }
      ctx->d_spec_boxes = const_cast<double *>(dp);
    }
  }
  ctx->lights.assign(scene->lights, scene->lights + scene->n_lights);
  ctx->materials.assign(scene->materials, scene->materials + scene->n_materials);
  ctx->camera = scene->camera;
  for (int k = 0; k < 3; k++) {
    ctx->ambient[k] = scene->ambient[k]; ctx->background[k] = scene->background[k];
    ctx->bbox_min[k] = scene->bbox_min[k]; ctx->bbox_max[k] = scene->bbox_max[k]; ctx->centroid[k] = scene->centroid[k];
  }
  ctx->radius = scene->radius;
  ctx->has_scene = true;
  return 0;
}

int gi_intersect_batch_device(GiContext *ctx, const GiRay *d_rays, uint64_t n, GiHit *d_hits, void *stream) {
  if (!ctx || !ctx->has_scene) { set_error("gi_intersect_batch: no scene uploaded"); return 1; }
  if (n == 0) return 0;
  GI_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = (cudaStream_t)stream;      // NULL = the CUDA default stream, as documented
  const int block = 128;
  uint64_t want = (n + block - 1) / block;
  uint64_t cap = (uint64_t)ctx->sm_count * 64;         // grid-stride beyond 64 CTAs per SM
  int grid = (int)(want < cap ? want : cap);
  intersect_kernel<<<grid, block, 0, st>>>(ctx->scene, d_rays, n, d_hits);
  ctx->launches++;
  GI_CUDA(cudaGetLastError());
  return 0;
}

int gi_intersect_batch(GiContext *ctx, const GiRay *rays, uint64_t n, GiHit *hits_out) {
  if (!ctx || !ctx->has_scene) { set_error("gi_intersect_batch: no scene uploaded"); return 1; }
  if (n == 0) return 0;
  if (!rays || !hits_out) { set_error("gi_intersect_batch: null buffer"); return 1; }
  GI_CUDA(cudaSetDevice(ctx->device));
  // chunks of 2^21 rays alternate between the two pipeline slots: copy-in, kernel and copy-out of a chunk are ordered
  // on its slot's stream, the other slot's copies overlap them (pinned host buffers make the copies asynchronous)
  const uint64_t chunk = 1ull << 21;
  int rc = 0;
  for (uint64_t o = 0, c = 0; o < n && rc == 0; o += chunk, c++) {
    const uint64_t m = n - o < chunk ? n - o : chunk;
    const int slot = (int)(c & 1);
    cudaStream_t st = ctx->pipe[slot];
    if (!ctx->stage[slot][0].fit(sizeof(GiRay) * chunk) || !ctx->stage[slot][1].fit(sizeof(GiHit) * chunk)) {
      set_error("gi_intersect_batch: out of device memory"); rc = 1; break;
    }
    GiRay *d_rays = (GiRay *)ctx->stage[slot][0].p; GiHit *d_hits = (GiHit *)ctx->stage[slot][1].p;
    if (!cuda_ok(cudaMemcpyAsync(d_rays, rays + o, sizeof(GiRay) * m, cudaMemcpyHostToDevice, st), "H2D rays") ||
        gi_intersect_batch_device(ctx, d_rays, m, d_hits, st) != 0 ||
        !cuda_ok(cudaMemcpyAsync(hits_out + o, d_hits, sizeof(GiHit) * m, cudaMemcpyDeviceToHost, st), "D2H hits"))
      rc = 1;
  }
  if (!cuda_ok(cudaStreamSynchronize(ctx->pipe[0]), "intersect sync") || !cuda_ok(cudaStreamSynchronize(ctx->pipe[1]), "intersect sync")) rc = 1;
  return rc;
}

int gi_intersect_batch_compact(GiContext *ctx, const GiRay *rays, uint64_t n, GiHitCompact *hits_out) {
  if (!ctx || !ctx->has_scene) { set_error("gi_intersect_batch_compact: no scene uploaded"); return 1; }
  if (n == 0) return 0;
  if (!rays || !hits_out) { set_error("gi_intersect_batch_compact: null buffer"); return 1; }
  GI_CUDA(cudaSetDevice(ctx->device));
  const uint64_t chunk = 1ull << 21;                       // the two-slot pipeline of gi_intersect_batch; only 16 bytes per hit go back
  int rc = 0;
  for (uint64_t o = 0, c = 0; o < n && rc == 0; o += chunk, c++) {
    const uint64_t m = n - o < chunk ? n - o : chunk;
    const int slot = (int)(c & 1);
    cudaStream_t st = ctx->pipe[slot];
    if (!ctx->stage[slot][0].fit(sizeof(GiRay) * chunk) || !ctx->stage[slot][1].fit(sizeof(GiHit) * chunk) ||
        !ctx->stage[slot][2].fit(sizeof(GiHitCompact) * chunk)) {
      set_error("gi_intersect_batch_compact: out of device memory"); rc = 1; break;
    }
    GiRay *d_rays = (GiRay *)ctx->stage[slot][0].p; GiHit *d_hits = (GiHit *)ctx->stage[slot][1].p;
    GiHitCompact *d_c = (GiHitCompact *)ctx->stage[slot][2].p;
    if (!cuda_ok(cudaMemcpyAsync(d_rays, rays + o, sizeof(GiRay) * m, cudaMemcpyHostToDevice, st), "H2D rays") ||
        gi_intersect_batch_device(ctx, d_rays, m, d_hits, st) != 0) { rc = 1; break; }
    compact_hits_kernel<<<(unsigned)((m + 255) / 256), 256, 0, st>>>(d_hits, m, d_c);
    ctx->launches++;
    if (!cuda_ok(cudaGetLastError(), "compact_hits_kernel") ||
        !cuda_ok(cudaMemcpyAsync(hits_out + o, d_c, sizeof(GiHitCompact) * m, cudaMemcpyDeviceToHost, st), "D2H hits"))
      rc = 1;
  }
  if (!cuda_ok(cudaStreamSynchronize(ctx->pipe[0]), "intersect sync") || !cuda_ok(cudaStreamSynchronize(ctx->pipe[1]), "intersect sync")) rc = 1;
  return rc;
}

int gi_launch_count(GiContext *ctx, uint64_t *out) {
  if (!ctx || !out) { set_error("gi_launch_count: null argument"); return 1; }
  *out = ctx->launches;
  return 0;
}

int gi_stage_timing(GiContext *ctx, int enable) {
  if (!ctx) { set_error("gi_stage_timing: null argument"); return 1; }
  ctx->clock.on = enable != 0;
  return 0;
}

int gi_stage_times(GiContext *ctx, GiStageTimes *out) {
  if (!ctx || !out) { set_error("gi_stage_times: null argument"); return 1; }
  const double *m = ctx->clock.ms;
  out->emit_filter_ms = m[ST_EMIT_FILTER]; out->trace_ms = m[ST_TRACE]; out->photon_exchange_ms = m[ST_PHOTON_EXCHANGE];
  out->photon_sort_ms = m[ST_PHOTON_SORT]; out->kd_build_ms = m[ST_KD_BUILD];
  out->primary_ms = m[ST_PRIMARY]; out->direct_ms = m[ST_DIRECT]; out->path_begin_ms = m[ST_PATH_BEGIN]; out->path_hit_ms = m[ST_PATH_HIT];
  out->path_shade_ms = m[ST_PATH_SHADE]; out->path_finish_ms = m[ST_PATH_FINISH]; out->gather_global_ms = m[ST_GATHER_GLOBAL];
  out->gather_caustic_ms = m[ST_GATHER_CAUSTIC]; out->commit_ms = m[ST_COMMIT]; out->resolve_ms = m[ST_RESOLVE];
  out->frame_exchange_ms = m[ST_FRAME_EXCHANGE];
  return 0;
}

int gi_queue_overflow_count(GiContext *ctx, uint64_t *out) {
  if (!ctx || !out) { set_error("gi_queue_overflow_count: null argument"); return 1; }
  *out = ctx->queue_overflows;
  return 0;
}

}  // extern "C"
