// csrc/isect.cuh -- closest-hit query of one ray against the flattened scene (device code).
//
// Replaces R3Scene::Intersects (src/R3Graphics/R3Scene.cpp:471-479) and everything below it:
// R3SceneNode::Intersects (R3SceneNode.cpp:420-510), R3SceneElement::Intersects
// (R3SceneElement.cpp:209-243) and the ray/primitive routines of src/R3Shapes/R3Isect.cpp:700-1197.
//
// Design (not the reference's): candidates come from an FP32 BVH traversal with conservatively padded
// boxes; each candidate primitive is then decided in FP64 in its node-local frame with the reference's
// exact sequence of epsilon tests, so the winner is the primitive the reference would return:
//   * the local ray of a node is obtained by walking the node path from the root (inverse matrix, then
//     re-normalisation, per level -- R3Line.cpp:139-146), with the node / element / mesh bounding-box
//     culls of the reference applied on the way; the result is cached while candidates stay in the same
//     node / element / shape;
//   * candidates are compared on the world-space t the reference would report (scale factors applied
//     innermost first, R3SceneNode.cpp:499); exact ties follow the reference's visiting order:
//     same mesh -> lower triangle index (R3Isect.cpp:819), same element -> later shape
//     (R3SceneElement.cpp:231), different elements -> earlier element (an element reports a hit only if it
//     strictly improves on max_t, R3SceneElement.cpp:242);
//   * a mesh whose nearest accepted triangle lies at t in [-1e-6, 0) yields NO hit at all in the reference
//     (R3Isect.cpp:813-829 keeps the minimum t, the element then rejects t < min_t); a point query at the
//     ray origin finds such meshes first and the main traversal skips them.
#pragma once
#include "gi_math.cuh"
#include "scene_dev.h"

#if defined(__CUDACC__)
#define GI_DEV __device__ __forceinline__
#else
#define GI_DEV inline
#endif

namespace gi {

struct RayD { D3 o, d; };

// ------------------------------------------------------------------------------------------------
// primitive routines (node-local frame, FP64)
// ------------------------------------------------------------------------------------------------

// R3Contains(box, point), R3Cont.cpp:776-787
GI_DEV bool box_contains(const double *b, D3 p) {
  if ((b[0] > b[3]) || (b[1] > b[4]) || (b[2] > b[5])) return false;
  if (is_neg(p.x - b[0])) return false;
  if (is_neg(p.y - b[1])) return false;
  if (is_neg(p.z - b[2])) return false;
  if (is_pos(p.x - b[3])) return false;
  if (is_pos(p.y - b[4])) return false;
  if (is_pos(p.z - b[5])) return false;
  return true;
}

// R3Intersects(ray, box), R3Isect.cpp:883-942 -- first face hit; far faces when the start is inside
GI_DEV bool ray_box(const RayD &r, const double *b, D3 *hp, D3 *hn, double *ht) {
  if ((b[0] > b[3]) || (b[1] > b[4]) || (b[2] > b[5])) return false;
  const bool start_inside = box_contains(b, r.o);
#pragma unroll
  for (int dim = 0; dim < 3; dim++) {
    const double od = comp(r.o, dim), dd = comp(r.d, dim);
    double tval;
    if (is_pos(dd)) {
      double delta = (start_inside ? b[3 + dim] : b[dim]) - od;
      if (delta < 0.0) continue;
      tval = delta / dd;
    } else if (is_neg(dd)) {
      double delta = (start_inside ? b[dim] : b[3 + dim]) - od;
      if (delta > 0.0) continue;
      tval = delta / dd;
    } else continue;
    const int d1 = (dim + 1) % 3, d2 = (dim + 2) % 3;
    D3 p = mk3(r.o.x + r.d.x * tval, r.o.y + r.d.y * tval, r.o.z + r.d.z * tval);
    const double p1 = comp(p, d1), p2 = comp(p, d2);
    if (is_neg0(p1 - b[3 + d1]) && is_pos0(p1 - b[d1]) && is_neg0(p2 - b[3 + d2]) && is_pos0(p2 - b[d2])) {
      if (ht) *ht = tval;
      if (hp) *hp = p;
      if (hn) {
        const double s = is_neg(dd) ? 1.0 : -1.0;
        *hn = mk3(dim == 0 ? s : 0.0, dim == 1 ? s : 0.0, dim == 2 ? s : 0.0);
      }
      return true;
    }
  }
  return false;
}

// R3Intersects(ray, plane), R3Isect.cpp:700-732.  0 miss, 1 point, 2 ray lies in the plane
GI_DEV int ray_plane(const RayD &r, const Plane4 &pl, D3 *hp, double *ht) {
  const double denom = dot(pn(pl), r.d);
  if (is_zero(denom)) return is_zero(signed_distance(pl, r.o)) ? 2 : 0;
  const double s = -(signed_distance(pl, r.o)) / denom;
  if (is_neg(s)) return 0;
  if (hp) *hp = r.o + r.d * s;
  if (ht) *ht = s;
  return 1;
}

// R3Intersects(ray, triangle) R3Isect.cpp:761-796 + R3Contains(triangle, point) R3Cont.cpp:491-514.
// `a` = cached plane (R3Triangle::Update, R3Triangle.cpp:475-497); b = bbox + the three edge planes, which
// the reference rebuilds on every call from the same vertices (same bits, so they are precomputed here).
GI_DEV bool ray_triangle(const RayD &r, const double2 *a, const double *b, D3 *hp, D3 *hn, double *ht) {
  const double2 a0 = a[0], a1 = a[1];
  Plane4 pl; pl.a = a0.x; pl.b = a0.y; pl.c = a1.x; pl.d = a1.y;
  D3 p; double t;
  if (ray_plane(r, pl, &p, &t) != 1) return false;
  if (!box_contains(b, p)) return false;
  if (!is_zero(signed_distance(pl, p))) return false;
#pragma unroll
  for (int e = 0; e < 3; e++) {
    Plane4 ep; ep.a = b[6 + 4 * e]; ep.b = b[7 + 4 * e]; ep.c = b[8 + 4 * e]; ep.d = b[9 + 4 * e];
    if (is_neg(signed_distance(ep, p))) return false;
  }
  *hp = p; *hn = pn(pl); *ht = t;
  return true;
}

// R3Intersects(ray, sphere) R3Isect.cpp:975-1021; R3Contains(sphere, point) R3Cont.cpp:1056-1063
GI_DEV bool ray_sphere(const RayD &r, const double2 *a, D3 *hp, D3 *hn, double *ht) {
  const double2 a0 = a[0], a1 = a[1];
  const D3 c = mk3(a0.x, a0.y, a1.x);
  const double rad = a1.y;
  const double r2 = rad * rad;
  const D3 EO = c - r.o;
  const double d2 = EO.x * EO.x + EO.y * EO.y + EO.z * EO.z;
  const bool start_inside = is_neg0(d2 - r2);
  const double v = dot(EO, r.d);
  if (!start_inside && is_neg0(v)) return false;
  const double disc = rad * rad - (dot(EO, EO) - v * v);
  if (is_neg(disc)) return false;
  const double d = sqrt(disc);           // NaN for disc in [-1e-6, 0): rejected by the caller's t >= 0 test
  const double t = start_inside ? v + d : v - d;
  const D3 p = r.o + r.d * t;
  *ht = t; *hp = p; *hn = (p - c) / rad;
  return true;
}

// R3Intersects(ray, circle) R3Isect.cpp:837-879 (the in-plane case returns garbage there; a miss here)
GI_DEV bool ray_circle(const RayD &r, const double *b, D3 *hp, D3 *hn, double *ht) {
  const D3 c = ld3(b), n = ld3(b + 3);
  const double rad = b[6];
  if (rad < 0.0) return false;
  const Plane4 pl = plane_point_normal(c, n);
  D3 p; double t;
  if (ray_plane(r, pl, &p, &t) != 1) return false;
  const D3 v = c - p;
  const double d2 = v.x * v.x + v.y * v.y + v.z * v.z;
  if (is_pos(d2 - rad * rad)) return false;
  *hp = p; *hn = n; *ht = t;
  return true;
}

// R3Intersects(ray, cylinder) R3Isect.cpp:1025-1197 (Graphics Gems IV p.356); b = axis start, end, radius
GI_DEV bool ray_cylinder(const RayD &ray, const double *b, D3 *hp, D3 *hn, double *ht) {
  const D3 a0 = ld3(b), a1 = ld3(b + 3);
  const double radius = b[6];
  D3 A = a1 - a0;                                                     // R3Span ctor, R3Span.cpp:63-70
  const double alen = length(A);
  if (!is_zero(alen)) A = A / alen;
  const D3 R = ray.d;
  const Plane4 base = plane_point_normal(a0, -A), top = plane_point_normal(a1, A);   // R3Cylinder.cpp:72-78
  double cyl_t1, cyl_t2;
  D3 D = cross(R, A);
  const double a = length(D);
  if (is_zero(a)) {
    const double d = length(cross(A, a0 - ray.o));                    // R3Distance(point, line), R3Dist.cpp:59-65
    if (is_neg0(d - radius)) { cyl_t1 = 0.0; cyl_t2 = GI_INFINITY; }
    else return false;
  } else {
    D = D / a;
    const D3 RC = ray.o - a0;
    double d = dot(RC, D);
    if (d < 0.0) d = -d;
    if (is_pos(d - radius)) return false;
    const double t = dot(cross(RC, A), D) / -a;
    double s;
    const D3 O = normalized(cross(D, A));
    const double bb = radius * radius - d * d;
    if (is_pos(bb)) {
      const double e = dot(R, O);
      s = sqrt(bb) / e;
      if (s < 0.0) s = -s;
    } else s = 0.0;
    cyl_t2 = t + s;
    if (is_neg(cyl_t2)) return false;
    cyl_t1 = t - s;
    if (cyl_t1 < 0.0) cyl_t1 = 0.0;
  }
  double cap_t1 = 0, cap_t2 = 0;
  const double dt = dot(ray.d, A);
  const double d_top = signed_distance(top, ray.o);
  const double d_base = signed_distance(base, ray.o);
  if (is_pos(d_top)) {
    if (is_pos0(dt)) return false;
    if (ray_plane(ray, top, nullptr, &cap_t1) && is_pos(cap_t1 - cyl_t2)) return false;
    else if (ray_plane(ray, plane_neg(base), nullptr, &cap_t2) && is_neg(cap_t2 - cyl_t1)) return false;
    else if (is_pos(cap_t1 - cyl_t1)) { *ht = cap_t1; *hp = ray.o + ray.d * cap_t1; *hn = pn(top); return true; }
  } else if (is_pos(d_base)) {
    if (is_neg0(dt)) return false;
    if (ray_plane(ray, base, nullptr, &cap_t1) && is_pos(cap_t1 - cyl_t2)) return false;
    else if (ray_plane(ray, plane_neg(top), nullptr, &cap_t2) && is_neg(cap_t2 - cyl_t1)) return false;
    else if (is_pos(cap_t1 - cyl_t1)) { *ht = cap_t1; *hp = ray.o + ray.d * cap_t1; *hn = pn(base); return true; }
  } else {
    if (is_pos(dt)) {
      if (ray_plane(ray, plane_neg(top), nullptr, &cap_t1)) { if (is_neg(cap_t1 - cyl_t1)) return false; }
    } else if (is_neg(dt)) {
      if (ray_plane(ray, plane_neg(base), nullptr, &cap_t1)) { if (is_neg(cap_t1 - cyl_t1)) return false; }
    }
  }
  *ht = cyl_t1;
  const D3 H = ray.o + ray.d * cyl_t1;
  *hp = H;
  const D3 HB = H - a0;
  *hn = (HB - A * dot(HB, A)) / radius;
  return true;
}

// dispatch on GiShapeType (0 triangle, 1 sphere, 2 box, 3 cylinder, 4 cone, 5 circle).  A cone never gets here:
// gi_upload_scene refuses scenes that hold one (scene_prep.h; the reference's version is unfinished -- no cap, zero
// normal, R3Isect.cpp:1259-1274 -- and no input scene uses it).
GI_DEV bool ray_prim(const DevScene &sc, int prim, int type, const RayD &r, D3 *hp, D3 *hn, double *ht) {
  const double2 *a = sc.prim_a + 2 * (size_t)prim;
  const double *b = sc.prim_b + GI_PRIM_B_STRIDE * (size_t)prim;
  switch (type) {
    case 0: return ray_triangle(r, a, b, hp, hn, ht);
    case 1: return ray_sphere(r, a, hp, hn, ht);
    case 2: return ray_box(r, b, hp, hn, ht);
    case 3: return ray_cylinder(r, b, hp, hn, ht);
    case 5: return ray_circle(r, b, hp, hn, ht);
    default: return false;
  }
}

// ------------------------------------------------------------------------------------------------
// node-local frames
// ------------------------------------------------------------------------------------------------

struct Frame {
  int node;                         // node id this frame belongs to (-1: none)
  bool ok;                          // false: a node bounding-box cull rejected the ray on the way down
  RayD ray;                         // the ray in the node's frame
  double scale[GI_MAX_NODE_DEPTH];  // per level, root first; 1.0 where the reference leaves t alone
  int depth;
};

// Walk root -> node as R3SceneNode::Intersects does (R3SceneNode.cpp:438-458)
GI_DEV void eval_frame(const DevScene &sc, int node_id, const RayD &world, Frame &f) {
  const DevNode &target = sc.nodes[node_id];
  f.node = node_id; f.depth = target.depth; f.ok = false;
  RayD r = world;
  for (int lvl = 0; lvl < target.depth; lvl++) {
    const DevNode &n = sc.nodes[target.path[lvl]];
    if (!box_contains(n.bbox, r.o)) {
      if (!ray_box(r, n.bbox, nullptr, nullptr, nullptr)) return;
    }
    const double len = length(xf_vector(n.m, r.d));          // sic: forward matrix (R3SceneNode.cpp:450-452)
    if (is_neg0(len)) return;
    f.scale[lvl] = is_zero(len - 1.0) ? 1.0 : len;
    RayD q;
    q.o = xf_point(n.minv, r.o);
    q.d = normalized(xf_vector(n.minv, r.d));                // R3Line::InverseTransform, R3Line.cpp:139-146
    r = q;
  }
  f.ray = r; f.ok = true;
}

GI_DEV double frame_t_to_world(const Frame &f, double t) {
  for (int lvl = f.depth - 1; lvl >= 0; lvl--) t = f.scale[lvl] * t;       // R3SceneNode.cpp:499
  return t;
}

// ------------------------------------------------------------------------------------------------
// closest hit
// ------------------------------------------------------------------------------------------------

struct HitRec {
  int prim;                 // -1: miss
  int shape, element, index_in_shape, node;
  double t;                 // world
  D3 p, n;                  // world
};

struct Cand { double tw; int prim, shape, element, idx, node; D3 pl, nl; double tl; };

// exact-tie rule, see the header of this file
GI_DEV bool tie_wins(const Cand &c, const Cand &best) {
  if (c.element == best.element) {
    if (c.shape == best.shape) return c.idx < best.idx;
    return c.shape > best.shape;
  }
  return c.element < best.element;
}

#define GI_MAX_POISON 4
// Traversal stack: one pending sibling per level of the descent.  gi_upload_scene measures the depth of the built BVH
// (bvh_build.cu: depth_kernel) and refuses a scene deeper than this, so the `sp < GI_STACK` guards below never drop a
// subtree.  (The LBVH over unique 62-bit keys is at most 62 levels deep by construction.)
#define GI_STACK 64

struct LeafCache {            // culls already decided for the current candidate's element / shape
  int element; bool element_ok;
  int shape; bool shape_ok;
};

// Decide one candidate primitive.  Returns true and fills `c` when the reference would accept it.
GI_DEV bool test_candidate(const DevScene &sc, int prim, const RayD &world, Frame &frame, LeafCache &lc, Cand &c) {
  const int4 meta = sc.prim_meta[prim];
  const int shape = meta.y, element = meta.z;
  if (lc.element != element) {
    const DevElement &e = sc.elements[element];
    if (frame.node != e.node) eval_frame(sc, e.node, world, frame);
    lc.element = element; lc.shape = -1;
    lc.element_ok = frame.ok;
    if (frame.ok && !box_contains(e.bbox, frame.ray.o))                  // R3SceneElement.cpp:221-225
      lc.element_ok = ray_box(frame.ray, e.bbox, nullptr, nullptr, nullptr);
  }
  if (!lc.element_ok) return false;
  if (lc.shape != shape) {
    const DevShape &s = sc.shapes[shape];
    lc.shape = shape;
    lc.shape_ok = s.is_mesh ? ray_box(frame.ray, s.bbox, nullptr, nullptr, nullptr) : true;   // R3Isect.cpp:807-809
  }
  if (!lc.shape_ok) return false;
  D3 p, n; double t;
  if (!ray_prim(sc, prim, meta.x, frame.ray, &p, &n, &t)) return false;
  c.tl = t; c.pl = p; c.nl = n; c.prim = prim; c.shape = shape; c.element = element; c.idx = meta.w;
  c.node = frame.node;
  return true;
}

// FP32 slab test of one padded child box; returns entry distance or a negative "miss" marker via `hit`
GI_DEV bool slab(float bx0, float by0, float bz0, float bx1, float by1, float bz1, float ox, float oy, float oz,
                float ix, float iy, float iz, float pad, float tmax, float *tnear) {
  float t0x = ((bx0 - pad) - ox) * ix, t1x = ((bx1 + pad) - ox) * ix;
  float t0y = ((by0 - pad) - oy) * iy, t1y = ((by1 + pad) - oy) * iy;
  float t0z = ((bz0 - pad) - oz) * iz, t1z = ((bz1 + pad) - oz) * iz;
  float tn = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fminf(t0z, t1z));
  float tf = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fmaxf(t0z, t1z));
  *tnear = tn;
  // the two sides are relaxed by a relative ulp margin so that rounding never culls a box the exact ray touches
  return (tn <= tf * 1.0000005f + 1e-30f) && (tf >= -1e-5f) && (tn <= tmax);
}

// Point query at the ray origin: meshes with an accepted triangle at local t < 0 (see file header)
GI_DEV int find_poisoned_meshes(const DevScene &sc, const RayD &world, float pad, int *poison) {
  int np = 0;
  Frame frame; frame.node = -1;
  LeafCache lc; lc.element = -1; lc.shape = -1;
  const float ox = (float)world.o.x, oy = (float)world.o.y, oz = (float)world.o.z;
  const float q = pad * 2.0f + 4e-6f;
  int stack[GI_STACK]; int sp = 0;
  int cur = 0;
  while (true) {
    const float4 *nd = sc.bvh + 4 * (size_t)cur;
    const float4 A = nd[0], B = nd[1], C = nd[2], Dd = nd[3];
    const int c0 = __float_as_int(Dd.x), c1 = __float_as_int(Dd.y);
    const bool in0 = ox >= A.x - q && oy >= A.y - q && oz >= A.z - q && ox <= A.w + q && oy <= B.x + q && oz <= B.y + q;
    const bool in1 = ox >= B.z - q && oy >= B.w - q && oz >= C.x - q && ox <= C.y + q && oy <= C.z + q && oz <= C.w + q;
    int next = -1;
    const int ch[2] = {c0, c1}; const bool in[2] = {in0, in1};
    for (int k = 0; k < 2; k++) {
      if (!in[k] || ch[k] == 0x7fffffff) continue;
      if (ch[k] < 0) {
        const int prim = ~ch[k];
        const int4 meta = sc.prim_meta[prim];
        if (meta.x == 0 && sc.shapes[meta.y].is_mesh) {
          bool known = false;
          for (int i = 0; i < np; i++) known = known || (poison[i] == meta.y);
          Cand c;
          if (!known && test_candidate(sc, prim, world, frame, lc, c) && c.tl < 0.0 && np < GI_MAX_POISON)
            poison[np++] = meta.y;
        }
      } else if (next < 0) next = ch[k];
      else if (sp < GI_STACK) stack[sp++] = ch[k];
    }
    if (next >= 0) { cur = next; continue; }
    if (sp == 0) break;
    cur = stack[--sp];
  }
  return np;
}

GI_DEV void closest_hit(const DevScene &sc, const RayD &world, HitRec &out) {
  out.prim = -1; out.shape = -1; out.element = -1; out.index_in_shape = 0; out.node = -1; out.t = 0.0;
  out.p = mk3(0, 0, 0); out.n = mk3(0, 0, 0);
  if (sc.n_prims <= 0) return;

  const float ox = (float)world.o.x, oy = (float)world.o.y, oz = (float)world.o.z;
  const float ix = 1.0f / (float)world.d.x, iy = 1.0f / (float)world.d.y, iz = 1.0f / (float)world.d.z;
  // FP32 padding: the float ray deviates from the double ray by <= 2^-24 (|o| + t) per axis over the part of
  // the ray that can touch the scene, and the slab arithmetic adds a few ulps of the same magnitude
  const float K = fmaxf(fmaxf(fabsf(ox), fabsf(oy)), fmaxf(fabsf(oz), sc.abs_max));
  const float pad = K * 1.9073486e-6f;     // 2^-19 * K

  int poison[GI_MAX_POISON]; int np = 0;
  if (sc.has_mesh) np = find_poisoned_meshes(sc, world, pad, poison);

  Frame frame; frame.node = -1;
  LeafCache lc; lc.element = -1; lc.shape = -1;
  Cand best; best.prim = -1; best.tw = GI_INFINITY; best.shape = -1; best.element = -1; best.idx = 0; best.node = -1;
  best.tl = 0; best.pl = mk3(0, 0, 0); best.nl = mk3(0, 0, 0);
  float tmax = 1.0e6f * 1.00001f;

  // Traversal in two alternating phases ("while-while"): a lane walks inner nodes (FP32 slab tests, near child first)
  // until the next thing on its path is a LEAF, then decides that primitive in FP64.  The lanes of a warp reconverge
  // between the phases, so the expensive exact tests run with every lane that has a candidate instead of with the one
  // or two lanes that happen to reach a leaf in the same step.  Stack entries may be leaves (negative ids).
  const int NONE = 0x7fffffff;
  int stack[GI_STACK]; int sp = 0;
  int cur = 0;
  while (true) {
    while (cur >= 0 && cur != NONE) {
      const float4 *nd = sc.bvh + 4 * (size_t)cur;
      const float4 A = nd[0], B = nd[1], C = nd[2], Dd = nd[3];
      int c0 = __float_as_int(Dd.x), c1 = __float_as_int(Dd.y);
      float tn0, tn1;
      bool h0 = c0 != NONE && slab(A.x, A.y, A.z, A.w, B.x, B.y, ox, oy, oz, ix, iy, iz, pad, tmax, &tn0);
      bool h1 = c1 != NONE && slab(B.z, B.w, C.x, C.y, C.z, C.w, ox, oy, oz, ix, iy, iz, pad, tmax, &tn1);
      if (h0 && h1) {
        if (tn1 < tn0) { const int ti = c0; c0 = c1; c1 = ti; }              // near child first
        if (sp < GI_STACK) stack[sp++] = c1;
        cur = c0;
      } else if (h0) cur = c0;
      else if (h1) cur = c1;
      else cur = sp > 0 ? stack[--sp] : NONE;
    }
    if (cur == NONE) break;
    {
      const int prim = ~cur;
      bool skip = false;
      if (np > 0) { const int sh = sc.prim_meta[prim].y; for (int i = 0; i < np; i++) skip = skip || (poison[i] == sh); }
      Cand c;
      if (!skip && test_candidate(sc, prim, world, frame, lc, c) && c.tl >= 0.0) {
        c.tw = frame_t_to_world(frame, c.tl);
        if (c.tw <= GI_INFINITY && ((c.tw < best.tw && c.tw < GI_INFINITY) ||
                                    (best.prim >= 0 && c.tw == best.tw && tie_wins(c, best)))) {
          best = c;
          tmax = (float)best.tw * 1.00001f + 1e-5f;
        }
      }
    }
    cur = sp > 0 ? stack[--sp] : NONE;
  }
  if (best.prim < 0) return;

  // map the winner back to the world frame, level by level (R3SceneNode.cpp:495-506)
  if (frame.node != best.node) eval_frame(sc, best.node, world, frame);
  const DevNode &target = sc.nodes[best.node];
  D3 p = best.pl, n = best.nl;
  for (int lvl = target.depth - 1; lvl >= 0; lvl--) {
    const DevNode &nd = sc.nodes[target.path[lvl]];
    p = xf_point(nd.m, p);
    n = normalized(xf_vector(nd.m, n));        // normals go through M, not M^-T, as in the reference
  }
  out.prim = best.prim; out.shape = best.shape; out.element = best.element; out.index_in_shape = best.idx;
  out.node = best.node; out.t = best.tw; out.p = p; out.n = n;
}

}  // namespace gi
