/* gi_b200.h -- C ABI of the B200-native hot path of the photon-mapping renderer.
 *
 * Drop-in boundary.  The reference (ReillyBova/Global-Illumination) has no FFI: its "boundary" is two
 * C++ calls inside main() -- MapPhotons() (src/photonmap.cpp:461) and RenderImage(aa, W, H)
 * (src/photonmap.cpp:471) -- which talk through ~60 mutable globals (src/render.h:27-110) and the
 * R3Scene object graph.  This header replaces exactly those two call sites with a plain-C ABI:
 * plain pointers and sizes, POD structs, no C++/torch types.  Each entry point cites the reference
 * interface it stands in for.  Host code (our `photonmap` CLI, tests, bench) stays C++/Python above
 * this line; everything below it is hand-written sm_100a CUDA.
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on failure; gi_last_error() gives the message.
 *     The library never calls exit() (the reference prints to stderr and exit(-1)s; main() maps our
 *     status to the same text/exit code).
 *   - the caller owns every host buffer passed in or out; the library owns all device memory behind
 *     the opaque GiContext.
 *   - there is NO CPU fallback: without a CUDA device gi_create fails.
 *   - all geometry is double precision like the reference (RNScalar = double, RNCompat.h:82-88).
 */
#ifndef GI_B200_H
#define GI_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GI_ABI_VERSION 7

/* ------------------------------------------------------------------------------------------------
 * Scene description (flattened R3Scene; src/R3Graphics/R3Scene.h).  Order is the order in which
 * R3SceneNode::Intersects (R3SceneNode.cpp:420-510) visits things: a node's elements in insertion
 * order, each element's shapes in insertion order, then the node's children, depth first.
 * ---------------------------------------------------------------------------------------------- */

typedef struct GiMaterial {          /* R3Brdf (R3Brdf.h:50-58) */
  double ka[3], kd[3], ks[3], kt[3], ke[3];
  double shininess;
  double ir;
} GiMaterial;

typedef struct GiNode {              /* R3SceneNode: transformation to the parent's coordinates */
  int32_t parent;                    /* -1 for the root */
  int32_t reserved;
  double m[16];                      /* row-major 4x4, R3Affine::Matrix() */
  double minv[16];                   /* R3Affine::InverseMatrix() (R4Matrix::Invert, R4Matrix.cpp:252) */
  double bbox[6];                    /* min(3) max(3) in the PARENT's coordinates (R3SceneNode::UpdateBBox,  */
                                     /* R3SceneNode.cpp:539-560); min > max = empty                          */
} GiNode;

typedef struct GiElement {           /* R3SceneElement: (node, material) bucket of shapes */
  int32_t node;
  int32_t material;                  /* index into GiScene.materials */
  double bbox[6];                    /* node-local union of the shapes' boxes (R3SceneElement.cpp:294-302) */
} GiElement;

typedef struct GiShape {             /* one R3Shape of the reference, in DFS visiting order */
  int32_t is_mesh;                   /* 1: R3TriangleArray (brute-force loop, R3Isect.cpp:800-833) */
  int32_t element;
  int32_t first_prim, n_prims;       /* run inside GiScene.prims */
  double bbox[6];                    /* node-local; a mesh's box covers ALL its vertices (R3TriangleArray.cpp:385-393) */
} GiShape;

enum GiShapeType {
  GI_SHAPE_TRIANGLE = 0,             /* d = v0 v1 v2                                (R3Triangle)      */
  GI_SHAPE_SPHERE   = 1,             /* d = centre(3) radius                        (R3Sphere)        */
  GI_SHAPE_BOX      = 2,             /* d = min(3) max(3)                           (R3Box)           */
  GI_SHAPE_CYLINDER = 3,             /* d = axis start(3) axis end(3) radius        (R3Cylinder)      */
  GI_SHAPE_CONE     = 4,             /* d = axis start(3) axis end(3) radius        (R3Cone)          */
  GI_SHAPE_CIRCLE   = 5              /* d = centre(3) unit normal(3) radius         (R3Circle)        */
};

typedef struct GiPrim {              /* one intersectable primitive; a mesh is a run of triangles */
  int32_t type;                      /* GiShapeType */
  int32_t element;                   /* index into GiScene.elements (gives node + material) */
  int32_t shape;                     /* reference shape id (DFS order); all triangles of one       */
                                     /* R3TriangleArray share it                                   */
  int32_t index_in_shape;            /* triangle index inside its mesh, 0 otherwise                */
  double d[9];
} GiPrim;

enum GiLightType { GI_LIGHT_DIR = 0, GI_LIGHT_POINT = 1, GI_LIGHT_SPOT = 2, GI_LIGHT_AREA = 3, GI_LIGHT_RECT = 4 };

typedef struct GiLight {             /* R3{Directional,Point,Spot,Area,Rect}Light */
  int32_t type;
  int32_t active;
  double color[3];
  double intensity;
  double position[3];
  double direction[3];               /* unit; area/rect: emitting normal */
  double axis1[3], axis2[3];         /* rect light unit axes */
  double len1, len2;                 /* rect light side lengths */
  double radius;                     /* area light */
  double ca, la, qa;                 /* constant / linear / quadratic attenuation */
  double dropoff, cutoff;            /* spot light */
} GiLight;

typedef struct GiCamera {            /* R3Camera */
  double eye[3], towards[3], up[3], right[3];
  double xfov, yfov, neardist, fardist;
} GiCamera;

typedef struct GiScene {
  int32_t n_nodes, n_elements, n_materials, n_prims, n_lights, n_shapes;
  const GiNode *nodes;
  const GiElement *elements;
  const GiMaterial *materials;
  const GiPrim *prims;
  const GiShape *shapes;
  const GiLight *lights;
  GiCamera camera;
  double ambient[3];
  double background[3];
  double bbox_min[3], bbox_max[3];   /* R3Scene::BBox() */
  double centroid[3];                /* R3Scene::Centroid() */
  double radius;                     /* SCENE_RADIUS = BBox().DiagonalRadius() (photonmap.cpp:455) */
} GiScene;

/* ------------------------------------------------------------------------------------------------
 * Batched ray queries: R3Scene::Intersects(ray, &node,&element,&shape,&point,&normal,&t)
 * (R3Scene.cpp:471-479) with min_t = 0, max_t = RN_INFINITY = 1e6 (R3Scene.h:73-76).
 * ---------------------------------------------------------------------------------------------- */

typedef struct GiRay { double o[3]; double d[3]; } GiRay;   /* d must be unit (R3Ray(p, v, TRUE)) */

typedef struct GiHit {
  int32_t hit;                       /* 0 = miss (all other fields zero, prim/shape = -1) */
  int32_t shape;                     /* reference shape id */
  int32_t prim;                      /* index into GiScene.prims */
  int32_t element;
  double t;
  double p[3];                       /* world hit point */
  double n[3];                       /* world geometric normal as the reference returns it */
} GiHit;

/* ------------------------------------------------------------------------------------------------
 * Photon maps: struct Photon (src/render.h:17-21), 32 bytes.
 * ---------------------------------------------------------------------------------------------- */

typedef struct GiPhoton {
  double position[3];
  uint8_t rgbe[4];                   /* Ward RGBE (graphics_utils.cpp:50-77) */
  uint16_t direction;                /* phi*256 + theta (photon_utils.cpp:56-60) */
  uint16_t pad;
} GiPhoton;

enum GiMap { GI_MAP_GLOBAL = 0, GI_MAP_CAUSTIC = 1 };
enum GiFilter { GI_FILTER_DISK = 0, GI_FILTER_CONE = 1, GI_FILTER_GAUSS = 2 };   /* Filter_Type, render.h:63 */

typedef struct GiGatherQuery {       /* arguments of EstimateRadiance (photon_utils.cpp:72-75) */
  double p[3];                       /* point */
  double n[3];                       /* normal */
  double bounce[3];                  /* exact_bounce */
  double cos_theta;
  double kd[3], ks[3];               /* brdf->Diffuse(), brdf->Specular() */
  double shininess;
} GiGatherQuery;

/* ------------------------------------------------------------------------------------------------
 * Parameters: one field per reference global (src/render.h:27-110, defaults src/photonmap.cpp:41-106)
 * ---------------------------------------------------------------------------------------------- */

typedef struct GiParams {
  int32_t verbose, threads, fresnel;
  double ir_air;
  int32_t ambient, direct_illum, transmissive_illum, specular_illum, indirect_illum, caustic_illum;
  int32_t direct_photon_illum, fast_global, irradiance_cache;
  int32_t shadows, soft_shadows, light_test, shadow_test;
  int32_t monte_carlo, max_monte_depth;
  double prob_absorb;
  int32_t recursive_shadows, distrib_transmissive, transmissive_test, distrib_specular, specular_test;
  int32_t depth_of_field, dof_test;
  double focus_depth, aperture_radius;
  int32_t global_photon_count, caustic_photon_count, max_photon_depth;
  int32_t indirect_test, global_estimate_size;
  double global_estimate_dist;
  int32_t global_filter, caustic_estimate_size;
  double caustic_estimate_dist;
  int32_t caustic_filter;
  double filter_const_a, filter_const_b, filter_const_k;
  uint64_t seed;                     /* new: counter-based RNG seed (reference is unseeded) */
  int32_t rank, world_size;          /* new: gi_render* without a communicator renders only the pixel columns x = rank (mod world_size) */
} GiParams;

typedef struct GiPhotonStats {       /* side effects of MapPhotons (photonmap.cpp:339-361) */
  uint64_t global_stored, caustic_stored;
  uint64_t global_emitted, caustic_emitted;
  int32_t indirect_illum, caustic_illum, direct_photon_illum;   /* flags after the pass */
  int32_t rounds;
  double trace_ms, build_ms;
  double total_power;                /* sum of LightPower() over the active lights (graphics_utils.cpp:223-258) */
  double cache_ms;                   /* irradiance-cache pre-pass (photonmap.cpp:380-411), 0 without -cache */
} GiPhotonStats;

typedef struct GiRenderStats {       /* the -v block of RenderImage (render.cpp:224-255) */
  uint64_t screen_rays, shadow_rays, monte_rays, transmissive_samples, specular_samples,
           indirect_samples, caustic_samples;
  double render_ms;
} GiRenderStats;

typedef struct GiContext GiContext;

/* library / device */
int gi_abi_version(void);
const char *gi_last_error(void);
int gi_create(int device, GiContext **out);                       /* fails if no CUDA device */
void gi_destroy(GiContext *ctx);
void gi_default_params(GiParams *p);                              /* photonmap.cpp:41-106 */

/* the program arguments that are not render globals (photonmap.cpp:22-34) */
typedef struct GiCliOptions {
  const char *input_scene_name, *output_image_name;   /* point into argv */
  int32_t width, height, aa, real_material;
  int32_t device;                                     /* new: CUDA device index (-device N); with -gpus: the first of N */
  int32_t gpus;                                       /* new: -gpus N, devices device .. device + N - 1 share the frame (default 1) */
} GiCliOptions;
void gi_default_cli_options(GiCliOptions *o);
/* ParseArgs (src/utils/io_utils.cpp:16-214); argv[0] is the program name.  New flags: -seed N, -device N, -gpus N. */
int gi_parse_args(int argc, const char *const *argv, GiParams *params, GiCliOptions *options);

int gi_launch_count(GiContext *ctx, uint64_t *out);                /* kernels launched so far by this context */
/* gi_render path batches that were repeated with fewer paths because a gather-request queue ran full (diagnostic;
 * the environment variables GI_QCAP / GI_BATCH_MAX shrink the queue / the batch to provoke it in tests) */
int gi_queue_overflow_count(GiContext *ctx, uint64_t *out);

/* Per-stage device times of the last gi_trace_photons and the last gi_render* of a context, in milliseconds (CUDA events on
 * the context's stream; the reference prints wall-clock totals only: the -v blocks of photonmap.cpp:331-377 and
 * render.cpp:224-255).  Off by default; gi_stage_timing(ctx, 1) switches the event pairs on (a few microseconds per
 * stage launch, no extra synchronisation). */
typedef struct GiStageTimes {
  /* gi_trace_photons */
  double emit_filter_ms, trace_ms, photon_exchange_ms, photon_sort_ms, kd_build_ms;
  /* gi_render / gi_render_float */
  double primary_ms, direct_ms, path_begin_ms, path_hit_ms, path_shade_ms, path_finish_ms, gather_global_ms, gather_caustic_ms,
         commit_ms, resolve_ms, frame_exchange_ms;
} GiStageTimes;
int gi_stage_timing(GiContext *ctx, int enable);
int gi_stage_times(GiContext *ctx, GiStageTimes *out);

/* host-side .scn/.off loader (no device needed): ReadScene (src/utils/io_utils.cpp:219-250) ->
 * R3Scene::ReadFile (R3Scene.cpp:514-587) -> ReadPrinceton (:1463-1953) / R3Mesh::ReadOffFile
 * (R3Mesh.cpp:4075-4215).  `path` ending in .gisc reads a flattened scene saved by gi_host_scene_save. */
typedef struct GiHostScene GiHostScene;
int gi_host_scene_load(const char *path, int real_material, GiHostScene **out);
int gi_host_scene_view(const GiHostScene *h, GiScene *out);       /* pointers stay owned by `h` */
int gi_host_scene_save(const GiHostScene *h, const char *gisc_path);
void gi_host_scene_free(GiHostScene *h);

/* replaces the pointer-chasing R3Scene graph: flatten + upload + GPU BVH build */
int gi_upload_scene(GiContext *ctx, const GiScene *scene);

/* R3Scene::Intersects over a batch.  Host buffers; copies are part of the call. */
int gi_intersect_batch(GiContext *ctx, const GiRay *rays, uint64_t n, GiHit *hits_out);
/* same with device-resident buffers on `stream` (a cudaStream_t, NULL = default stream) */
int gi_intersect_batch_device(GiContext *ctx, const GiRay *d_rays, uint64_t n, GiHit *d_hits, void *stream);

/* photon maps: GLOBAL_PHOTONS / CAUSTIC_PHOTONS + R3Kdtree build (photonmap.cpp:364-377) */
int gi_load_photons(GiContext *ctx, int map, const GiPhoton *photons, uint64_t n);
int gi_load_photons_device(GiContext *ctx, int map, const GiPhoton *d_photons, uint64_t n);   /* device buffer */
int gi_load_photons_file(GiContext *ctx, int map, const char *path);
int gi_dump_photons_file(GiContext *ctx, int map, const char *path);
int gi_photon_count(GiContext *ctx, int map, uint64_t *n_out);
int gi_get_photons(GiContext *ctx, int map, GiPhoton *out, uint64_t capacity);   /* map order */
/* same into a device buffer, asynchronously on `stream` (NULL = default stream): feeds the photon all-gather */
int gi_get_photons_device(GiContext *ctx, int map, GiPhoton *d_out, uint64_t capacity, void *stream);
int gi_build_maps(GiContext *ctx, const GiParams *params);

/* EstimateRadiance over a batch (photon_utils.cpp:72-162): kNN index sets + radiance.
 * idx_out: n*k int32, ascending photon indices of the k nearest within r (ties broken by lower index),
 * padded with -1; rgb_out: n*3 doubles (the estimate that the reference adds into `color`);
 * count_out: n int32.  Any output pointer may be NULL. */
int gi_gather_batch(GiContext *ctx, int map, const GiGatherQuery *queries, uint64_t n, int k, double r,
                    int filter, const GiParams *params, int32_t *idx_out, int32_t *count_out, double *rgb_out);
int gi_gather_batch_device(GiContext *ctx, int map, const GiGatherQuery *d_queries, uint64_t n, int k, double r,
                           int filter, const GiParams *params, int32_t *d_idx, int32_t *d_count, double *d_rgb,
                           void *stream);
/* The _device variant enqueues on `stream` and returns without waiting.  Batches of >= 4096 queries are processed along a
 * Morton curve (results stay at the caller's index); the ordering scratch, the chunk counter and the radius hint live
 * in the context, so calls on ONE context must not run concurrently on different caller streams (order them with
 * events, or use one context per stream).  n < 2^31 per device batch; the host-buffer call has no such limit. */

/* gi_gather_batch_device WITHOUT any allocation.  This entry point has no counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arguments, same results as gi_gather_batch_device (identical kNN sets
 * and counts; radiance equal up to the summation order, <= 1e-9 relative), but everything the gather needs besides the
 * caller's buffers -- the launch-wide chunk counter, the Morton keys / permutation of the batch and the sort's temporary
 * storage -- lives in ONE file-scope `static __device__` array of csrc/gather.cu (256 B + 16 B x GI_NOALLOC_MAX_QUERIES
 * + 1 MiB = 5.25 MB, held by every GPU that loads the library until the process exits).  The call makes no cudaMalloc /
 * cudaFree / allocating library call; it fails with a clear error if n > GI_NOALLOC_MAX_QUERIES, and if the map was not
 * built (gi_build_maps creates the direction table the estimate reads).
 * The array is shared by EVERY call on the device, whatever the context or stream, and keeps the last call's contents
 * (the call clears the counter it needs at zero itself): two calls that overlap on the GPU race -- order them on one
 * stream or with events.  Enqueues on `stream` and returns without waiting, like gi_gather_batch_device. */
#define GI_NOALLOC_MAX_QUERIES 262144
int gi_gather_batch_device_noalloc(GiContext *ctx, int map, const GiGatherQuery *d_queries, uint64_t n, int k, double r,
                                   int filter, const GiParams *params, int32_t *d_idx, int32_t *d_count, double *d_rgb,
                                   void *stream);

/* Compact host-buffer variants: the same results with fewer bytes over PCIe (the 72-byte GiHit and the 136-byte
 * GiGatherQuery mirror the reference's argument lists; a host that only needs ids and distances, or whose queries are
 * surface hits of the uploaded scene, moves 16 and 80 bytes instead).
 *   gi_intersect_batch_compact   R3Scene::Intersects over a batch (R3Scene.cpp:471-479) returning {prim, shape, t}; prim < 0:
 *                                miss.  Point and normal are recomputable from the ray and t / the primitive.
 *   gi_gather_batch_compact      EstimateRadiance over a batch (photon_utils.cpp:72-162) for queries given as RayTrace has
 *                                them at a hit (raytracer.cpp:138-167, 174-190): point, normal, the unit direction `view` of
 *                                the ray that arrived there and the hit element's material (index into GiScene.materials of
 *                                the uploaded scene); cos_theta = normal . (-view), exact_bounce = ReflectiveBounce(normal,
 *                                view, cos_theta) (graphics_utils.cpp:104-117) and kd / ks / shininess are derived on the
 *                                device.  count_out / rgb_out as gi_gather_batch (either may be NULL). */
typedef struct GiHitCompact { int32_t prim, shape; double t; } GiHitCompact;
typedef struct GiGatherQueryCompact { double p[3], n[3], view[3]; int32_t material, pad; } GiGatherQueryCompact;
int gi_intersect_batch_compact(GiContext *ctx, const GiRay *rays, uint64_t n, GiHitCompact *hits_out);
int gi_gather_batch_compact(GiContext *ctx, int map, const GiGatherQueryCompact *queries, uint64_t n, int k, double r, int filter,
                            const GiParams *params, int32_t *count_out, double *rgb_out);

/* The irradiance cache (-cache).  gi_build_maps with params->irradiance_cache (and indirect_illum or
 * direct_photon_illum) set runs the pre-pass of MapPhotons (photonmap.cpp:380-411) once per loaded global map: every
 * global photon's colour becomes its power + EstimateIrradiance at its position (photon_utils.cpp:208-246);
 * gi_get_photons then returns the converted photons.  gi_cached_radiance_batch is a loop of
 * EstimateCachedRadiance(point, normal, color, brdf, exact_bounce, cos_theta, GLOBAL_PMAP, r)
 * (photon_utils.cpp:164-205) on that map; rgb_out: n*3 doubles.  Where no photon qualifies within r the result is
 * black (the reference reads through a NULL pointer there). */
int gi_cached_radiance_batch(GiContext *ctx, const GiGatherQuery *queries, uint64_t n, double r, double *rgb_out);
int gi_cached_radiance_batch_device(GiContext *ctx, const GiGatherQuery *d_queries, uint64_t n, double r, double *d_rgb,
                                    void *stream);

/* MapPhotons() (photonmap.cpp:260-436): emission + tracing + power scaling; maps stay on the device */
int gi_trace_photons(GiContext *ctx, const GiParams *params, GiPhotonStats *stats);
/* The power scale of MapPhotons on a loaded, not yet built map (the decode - multiply - encode loop of
 * photonmap.cpp:339-361); gi_trace_photons applies it itself. */
int gi_scale_photon_power(GiContext *ctx, int map, double factor);

/* ------------------------------------------------------------------------------------------------
 * Multi-GPU (one 8 x B200 box): one GiContext per GPU, tied together by an NCCL communicator over NVLink / NVSwitch.
 * A rank takes the place of one worker thread of the reference: MapPhotons splits the photon budget over its threads
 * (photonmap.cpp:294-329) and RenderImage strides the pixel columns over them (render.cpp:80-90,188-198).  Scene and
 * photon maps are replicated (1e8 photons = 3.2 GB); nothing is exchanged while rendering.
 *   one process, N devices:  gi_create x N, gi_upload_scene x N, gi_comm_init_all(ctxs, N), then one host thread per
 *                            context calls gi_trace_photons / gi_render (`photonmap -gpus N` does exactly this);
 *   one process per GPU:     rank 0 calls gi_comm_unique_id, the host hands the 128 bytes to every rank (MPI,
 *                            torch.distributed, a file ...), every rank calls gi_comm_init_rank.
 * With a communicator attached:
 *   gi_trace_photons  traces this rank's share of every emission round: photon t of a round belongs to rank
 *                     t mod world, random streams are keyed by the GLOBAL photon number and the stored counts are
 *                     all-reduced after every round, so the adaptive schedule (photonmap.cpp:163-192) and the final map
 *                     are byte-identical for ANY number of ranks; one ragged all-gather replicates the photons, every
 *                     rank applies the power scale and builds its kd-trees.  GiPhotonStats holds global totals.
 *   gi_render         renders the pixel columns x = rank (mod world) and gathers the 8-bit column strips to rank 0:
 *                     rgb_out is written on rank 0 only (may be NULL elsewhere); GiRenderStats counters are summed
 *                     over the ranks.  gi_render_float returns this rank's columns (zeros elsewhere), no exchange.
 * Every rank must make the same calls with the same parameters.  params->rank / world_size are ignored when a
 * communicator is attached.  NCCL is bound at run time (dlopen libnccl.so.2); without it only these calls fail.
 * ---------------------------------------------------------------------------------------------- */
typedef struct GiUniqueId { char bytes[128]; } GiUniqueId;        /* an ncclUniqueId */
int gi_comm_unique_id(GiUniqueId *out);
int gi_comm_init_rank(GiContext *ctx, const GiUniqueId *id, int rank, int world);
int gi_comm_init_all(GiContext **ctxs, int n);                    /* rank i = ctxs[i]; contexts on distinct devices */
int gi_comm_info(GiContext *ctx, int *rank, int *world);          /* 0 / 1 without a communicator */
int gi_comm_destroy(GiContext *ctx);                              /* also done by gi_destroy */

/* Loading a map whose photons sit sharded over the ranks (device memory): stands in for the ONE GLOBAL_PHOTONS /
 * CAUSTIC_PHOTONS array that all worker threads of the reference insert into (photonmap.cpp:117-118, StorePhoton,
 * photon_utils.cpp:40-66).  Every rank passes its share; one ragged all-gather over NVLink concatenates the shares in
 * rank order straight into the map's photon array on every rank (no padding, no staging copy); gi_build_maps then builds
 * the kd-tree locally.  Collective: every rank must call it.  Without a communicator: gi_load_photons_device. */
int gi_load_photons_sharded_device(GiContext *ctx, int map, const GiPhoton *d_share, uint64_t n_share);

/* RenderImage(aa, W, H) (render.cpp:155-259).  rgb_out: W*H*3 bytes, row 0 = bottom row like R2Image. */
int gi_render(GiContext *ctx, const GiParams *params, int width, int height, int aa, uint8_t *rgb_out,
              GiRenderStats *stats);
/* same, but returns the un-quantised, un-clamped supersampled buffer (W*2^aa x H*2^aa x 3 doubles,
 * column-major like image_buffer[i][j]) for parity tests */
int gi_render_float(GiContext *ctx, const GiParams *params, int width, int height, int aa, double *rgb_out,
                    GiRenderStats *stats);

/* The debug views of the reference's `visualize` program (src/visualize.cpp), headless: the pictures it draws with
 * OpenGL through the scene camera, rasterised by CUDA splat / line kernels into a z-buffered frame (no graphics stack).
 *   GI_VIEW_GLOBAL_PHOTONS / GI_VIEW_CAUSTIC_PHOTONS  DrawStoredGlobal / DrawStoredCaustic (visualize.cpp:1366-1424):
 *        per stored photon a sphere of radius 0.01 in the hue of its power and a span along its incident direction;
 *   GI_VIEW_RAYS    DrawRays (:268-298): hit point and normal of a 20 x 20 grid of eye rays;
 *   GI_VIEW_PATHS   DrawPaths + MonteCarlo_PathTrace (:436-530, :589-700): the specular / refractive walks (<= 10 bounces)
 *        of a 25 x 25 grid of eye rays, one span per segment coloured by the pixel (params->seed drives the roulette).
 * draw_shapes != 0 also draws the scene's surfaces (grey head-light shade) so that geometry hides what lies behind it, and
 * fills empty pixels with the scene's background colour (black otherwise).  rgb_out: W*H*3 bytes, row 0 = bottom row.
 * params may be NULL (defaults). */
enum GiDebugView { GI_VIEW_GLOBAL_PHOTONS = 0, GI_VIEW_CAUSTIC_PHOTONS = 1, GI_VIEW_RAYS = 2, GI_VIEW_PATHS = 3 };
int gi_debug_view(GiContext *ctx, int kind, const GiParams *params, int width, int height, int draw_shapes, uint8_t *rgb_out);

/* R2Image::Write (R2Image.cpp:317-340) by file extension: .png, .ppm (ASCII, the reference's layout), .bmp.
 * rgb is the gi_render frame (row 0 = bottom row). */
int gi_write_image(const char *path, const uint8_t *rgb, int width, int height);

#ifdef __cplusplus
}
#endif
#endif /* GI_B200_H */
