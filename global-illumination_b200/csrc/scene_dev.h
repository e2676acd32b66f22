// csrc/scene_dev.h -- device-resident scene layout (what gi_upload_scene builds in HBM).
//
// The reference walks a pointer graph R3Scene -> R3SceneNode -> R3SceneElement -> R3Shape with a virtual
// call per shape (R3SceneNode.cpp:420-510).  Here the scene is flat:
//   * bvh[]       one LBVH over world-space, outward-padded FP32 boxes of ALL primitives (a mesh is a run of
//                 triangle primitives), 64-byte nodes = the two child boxes + child ids, read as 4 x LDG.128;
//   * prim_a/b[]  per-primitive FP64 data in the primitive's node-local frame, precomputed once with the
//                 reference's arithmetic (triangle plane, box, edge planes), read as 128-bit loads;
//   * shapes/elements/nodes[]  the (tiny) tables that carry the reference's per-level bounding-box culls and
//                 per-node matrices, needed to reproduce its decisions and tie rules exactly.
#pragma once
#include <stdint.h>
#if defined(__CUDACC__)
#include <cuda_runtime.h>
#else
// host_sim build (tests only): minimal stand-ins for the CUDA vector types
struct float4 { float x, y, z, w; };
struct int4 { int x, y, z, w; };
struct alignas(16) double2 { double x, y; };
#include <string.h>
static inline int __float_as_int(float f) { int i; memcpy(&i, &f, 4); return i; }
static inline float __int_as_float(int i) { float f; memcpy(&f, &i, 4); return f; }
#endif

namespace gi {

#define GI_MAX_NODE_DEPTH 8     /* root + 7 nested `begin` groups (deepest input scene: 6 levels) */
#define GI_PRIM_B_STRIDE 18     /* doubles per primitive in prim_b */

struct DevNode {                /* R3SceneNode */
  double m[16];                 /* node -> parent */
  double minv[16];              /* parent -> node */
  double bbox[6];               /* in the parent's frame */
  int32_t depth;                /* number of entries of path[] (root: 1) */
  int32_t path[GI_MAX_NODE_DEPTH];   /* node ids from the root down to this node */
  int32_t pad;
};

struct DevElement { double bbox[6]; int32_t node; int32_t material; };            /* R3SceneElement */
struct DevShape { double bbox[6]; int32_t element; int32_t is_mesh; int32_t first_prim; int32_t n_prims; };

struct DevMaterial { double ka[3], kd[3], ks[3], kt[3], ke[3]; double shininess, ir; };

/* prim_a[i] (2 x double2):  triangle: plane a,b,c,d | sphere: centre xyz, radius | others: unused
 * prim_b[i*18 ..]:          triangle: bbox min(3) max(3), edge planes e01(4) e12(4) e20(4)
 *                           box: min(3) max(3) | cylinder: axis start(3) end(3) radius | circle: centre(3) normal(3) radius
 * prim_meta[i]:             x = GiShapeType, y = shape id, z = element id, w = index inside the shape            */
struct DevScene {
  const float4 *bvh;            /* 4 float4 per node; root = node 0 */
  int32_t n_bvh_nodes;
  int32_t n_prims;
  const double2 *prim_a;        /* 2 double2 per primitive */
  const double *prim_b;
  const int4 *prim_meta;
  const DevShape *shapes;
  const DevElement *elements;
  const DevNode *nodes;
  const DevMaterial *materials;
  int32_t n_shapes, n_elements, n_nodes, n_materials;
  int32_t has_mesh;             /* any R3TriangleArray shape: enables the negative-t pre-pass (isect.cuh) */
  float abs_max;                /* max |coordinate| of the world bbox: scales the FP32 traversal padding */
};

}  // namespace gi
