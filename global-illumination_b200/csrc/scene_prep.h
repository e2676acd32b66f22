// csrc/scene_prep.h -- host-side preparation of the device scene tables from a GiScene (header-only, plain C++).
//
// Everything the kernels can reuse across rays is computed once here with the reference's arithmetic:
// triangle planes (R3Triangle::Update, R3Triangle.cpp:475-497), the three edge planes that
// R3Contains(triangle, point) rebuilds on every call (R3Cont.cpp:499-510), node paths, and the padded
// world-space FP32 boxes the BVH is built over.  Must be compiled without FMA contraction.
#pragma once
#include <cfloat>
#include <cmath>
#include <cstring>
#include <string>
#include <vector>
#include "gi_b200.h"
#include "gi_math.cuh"
#include "scene_dev.h"

namespace gi {

struct PreparedScene {
  std::vector<DevNode> nodes;
  std::vector<DevElement> elements;
  std::vector<DevShape> shapes;
  std::vector<DevMaterial> materials;
  std::vector<double2> prim_a;          // 2 per primitive
  std::vector<double> prim_b;           // GI_PRIM_B_STRIDE per primitive
  std::vector<int4> prim_meta;
  std::vector<float> prim_boxes;        // 6 per primitive, padded, world space
  float lo[3], hi[3];                   // union of prim_boxes
  float abs_max;
  int has_mesh;
};

inline float round_down(double v) { float f = (float)v; if ((double)f > v) f = nextafterf(f, -INFINITY); return f; }
inline float round_up(double v) { float f = (float)v; if ((double)f < v) f = nextafterf(f, INFINITY); return f; }

inline bool prepare_scene(const GiScene &s, PreparedScene &out, std::string &err) {
  if (s.n_nodes <= 0) { err = "scene has no root node"; return false; }
  out = PreparedScene();
  out.nodes.resize(s.n_nodes);
  for (int i = 0; i < s.n_nodes; i++) {
    DevNode &d = out.nodes[i]; memset(&d, 0, sizeof(d));
    memcpy(d.m, s.nodes[i].m, sizeof(d.m)); memcpy(d.minv, s.nodes[i].minv, sizeof(d.minv));
    memcpy(d.bbox, s.nodes[i].bbox, sizeof(d.bbox));
    int chain[64], n = 0;
    for (int c = i; c >= 0; c = s.nodes[c].parent) {
      if (n >= GI_MAX_NODE_DEPTH) { err = "scene groups nested deeper than GI_MAX_NODE_DEPTH"; return false; }
      chain[n++] = c;
    }
    d.depth = n;
    for (int k = 0; k < n; k++) d.path[k] = chain[n - 1 - k];
  }
  out.elements.resize(s.n_elements);
  for (int i = 0; i < s.n_elements; i++) {
    DevElement &e = out.elements[i];
    memcpy(e.bbox, s.elements[i].bbox, sizeof(e.bbox)); e.node = s.elements[i].node; e.material = s.elements[i].material;
  }
  out.shapes.resize(s.n_shapes);
  out.has_mesh = 0;
  for (int i = 0; i < s.n_shapes; i++) {
    DevShape &d = out.shapes[i];
    memcpy(d.bbox, s.shapes[i].bbox, sizeof(d.bbox)); d.element = s.shapes[i].element; d.is_mesh = s.shapes[i].is_mesh;
    d.first_prim = s.shapes[i].first_prim; d.n_prims = s.shapes[i].n_prims;
    if (d.is_mesh) out.has_mesh = 1;
  }
  out.materials.resize(s.n_materials);
  for (int i = 0; i < s.n_materials; i++) {
    static_assert(sizeof(DevMaterial) == sizeof(GiMaterial), "material layout");
    memcpy(&out.materials[i], &s.materials[i], sizeof(GiMaterial));
  }
  const int np = s.n_prims;
  out.prim_a.assign((size_t)np * 2, double2{0, 0});
  out.prim_b.assign((size_t)np * GI_PRIM_B_STRIDE, 0.0);
  out.prim_meta.resize(np);
  out.prim_boxes.resize((size_t)np * 6);
  for (int k = 0; k < 3; k++) { out.lo[k] = FLT_MAX; out.hi[k] = -FLT_MAX; }
  for (int i = 0; i < np; i++) {
    const GiPrim &p = s.prims[i];
    double2 *a = &out.prim_a[(size_t)i * 2];
    double *b = &out.prim_b[(size_t)i * GI_PRIM_B_STRIDE];
    out.prim_meta[i] = int4{p.type, p.shape, p.element, p.index_in_shape};
    double lb[6];                                     // node-local bounds of the primitive
    switch (p.type) {
      case GI_SHAPE_TRIANGLE: {
        D3 p0 = ld3(p.d), p1 = ld3(p.d + 3), p2 = ld3(p.d + 6);
        Plane4 pl = plane_3_points(p0, p1, p2);
        a[0] = double2{pl.a, pl.b}; a[1] = double2{pl.c, pl.d};
        for (int k = 0; k < 3; k++) {                 // R3Box::Union(point), R3Box.cpp:455-471
          double lo = p.d[k], hi = p.d[k];
          for (int v = 1; v < 3; v++) { double c = p.d[3 * v + k]; if (lo > c) lo = c; if (hi < c) hi = c; }
          b[k] = lo; b[3 + k] = hi;
        }
        Plane4 e[3] = {plane_point_2_vectors(p1, pn(pl), p1 - p0), plane_point_2_vectors(p2, pn(pl), p2 - p1),
                       plane_point_2_vectors(p0, pn(pl), p0 - p2)};
        for (int k = 0; k < 3; k++) { b[6 + 4 * k] = e[k].a; b[7 + 4 * k] = e[k].b; b[8 + 4 * k] = e[k].c; b[9 + 4 * k] = e[k].d; }
        memcpy(lb, b, sizeof(lb));
        break;
      }
      case GI_SHAPE_SPHERE:
        a[0] = double2{p.d[0], p.d[1]}; a[1] = double2{p.d[2], p.d[3]};
        for (int k = 0; k < 3; k++) { lb[k] = p.d[k] - p.d[3]; lb[3 + k] = p.d[k] + p.d[3]; }
        break;
      case GI_SHAPE_BOX:
        memcpy(b, p.d, 6 * sizeof(double)); memcpy(lb, p.d, sizeof(lb));
        break;
      case GI_SHAPE_CONE:
        // SURVEY.md §8 A6 "flagged unsupported": the reference's own intersector is unfinished (no base cap, ZERO normal,
        // R3Isect.cpp:1201-1275), so a cone cannot be shaded there either; refusing by name beats rendering it as nothing
        err = "cone primitives are not supported (the reference's R3Intersects(ray, cone) is unfinished: no cap, zero normal, "
              "R3Isect.cpp:1201-1275); primitive " + std::to_string(i);
        return false;
      case GI_SHAPE_CYLINDER:
        memcpy(b, p.d, 7 * sizeof(double));
        for (int k = 0; k < 3; k++) {
          lb[k] = fmin(p.d[k], p.d[3 + k]) - p.d[6]; lb[3 + k] = fmax(p.d[k], p.d[3 + k]) + p.d[6];
        }
        break;
      case GI_SHAPE_CIRCLE:
        memcpy(b, p.d, 7 * sizeof(double));
        for (int k = 0; k < 3; k++) { lb[k] = p.d[k] - p.d[6]; lb[3 + k] = p.d[k] + p.d[6]; }
        break;
      default: err = "unknown primitive type"; return false;
    }
    // world box: the local box, grown by the intersectors' 1e-6 tolerances, pushed up the node path
    const DevNode &nd = out.nodes[s.elements[p.element].node];
    double wb[6];
    for (int k = 0; k < 3; k++) { wb[k] = lb[k] - 4.0e-6; wb[3 + k] = lb[3 + k] + 4.0e-6; }
    for (int lvl = nd.depth - 1; lvl >= 0; lvl--) {
      const double *m = out.nodes[nd.path[lvl]].m;
      double nb[6] = {DBL_MAX, DBL_MAX, DBL_MAX, -DBL_MAX, -DBL_MAX, -DBL_MAX};
      for (int o = 0; o < 8; o++) {
        D3 c = mk3((o & 4) ? wb[3] : wb[0], (o & 2) ? wb[4] : wb[1], (o & 1) ? wb[5] : wb[2]);
        D3 q = xf_point(m, c);
        nb[0] = fmin(nb[0], q.x); nb[1] = fmin(nb[1], q.y); nb[2] = fmin(nb[2], q.z);
        nb[3] = fmax(nb[3], q.x); nb[4] = fmax(nb[4], q.y); nb[5] = fmax(nb[5], q.z);
      }
      memcpy(wb, nb, sizeof(wb));
    }
    float *fb = &out.prim_boxes[(size_t)i * 6];
    for (int k = 0; k < 3; k++) {
      double padw = 4.0e-6 + 1.0e-9 * fmax(fabs(wb[k]), fabs(wb[3 + k]));
      fb[k] = round_down(wb[k] - padw); fb[3 + k] = round_up(wb[3 + k] + padw);
      if (fb[k] < out.lo[k]) out.lo[k] = fb[k];
      if (fb[3 + k] > out.hi[k]) out.hi[k] = fb[3 + k];
    }
  }
  out.abs_max = 0.0f;
  if (np > 0) for (int k = 0; k < 3; k++) out.abs_max = fmaxf(out.abs_max, fmaxf(fabsf(out.lo[k]), fabsf(out.hi[k])));
  return true;
}

}  // namespace gi
