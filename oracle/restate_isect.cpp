// oracle/restate_isect.cpp -- TEST INFRASTRUCTURE: CPU restatement of the reference's ray-scene
// intersection, R3Scene::Intersects (src/R3Graphics/R3Scene.cpp:471-479), written from the reference's
// behaviour in our own flat data model.  It is the checker for the CUDA path, never the product.
// Pinned against the reference itself: tests/test_oracle_isect.py compares it hit-for-hit with
// oracle/_ref/ref_harness (the unmodified reference sources) on every input scene.
//
// Structure mirrors the reference's recursion on purpose (node -> elements -> shapes -> children), because
// the order of visits defines which candidate wins a tie; the CUDA path uses a BVH instead.
#include <thread>
#include "restate_common.h"

namespace oracle {

struct Ray { V3 o, d; };
struct Hit { bool hit; int shape; int prim; double t; V3 p, n; };

// R3Contains(box, point), R3Cont.cpp:776-787
static bool box_contains(const double *b, V3 p) {
  if ((b[0] > b[3]) || (b[1] > b[4]) || (b[2] > b[5])) return false;       // IsEmpty, R3Box.h:238-242
  if (is_less(p.x, b[0])) return false;
  if (is_less(p.y, b[1])) return false;
  if (is_less(p.z, b[2])) return false;
  if (is_greater(p.x, b[3])) return false;
  if (is_greater(p.y, b[4])) return false;
  if (is_greater(p.z, b[5])) return false;
  return true;
}

// R3Intersects(ray, box), R3Isect.cpp:883-942: first face hit, far faces when the start is inside
static bool ray_box(const Ray &r, const double *b, V3 *hp, V3 *hn, double *ht) {
  if ((b[0] > b[3]) || (b[1] > b[4]) || (b[2] > b[5])) return false;
  bool start_inside = box_contains(b, r.o);
  const double o[3] = {r.o.x, r.o.y, r.o.z}, d[3] = {r.d.x, r.d.y, r.d.z};
  for (int dim = 0; dim < 3; dim++) {
    double tval;
    if (is_pos(d[dim])) {
      double c = start_inside ? b[3 + dim] : b[dim];
      double delta = c - o[dim];
      if (delta < 0.0) continue;
      tval = delta / d[dim];
    } else if (is_neg(d[dim])) {
      double c = start_inside ? b[dim] : b[3 + dim];
      double delta = c - o[dim];
      if (delta > 0.0) continue;
      tval = delta / d[dim];
    } else continue;
    int d1 = (dim + 1) % 3, d2 = (dim + 2) % 3;
    double p[3];
    p[dim] = o[dim] + d[dim] * tval;
    p[d1] = o[d1] + d[d1] * tval;
    p[d2] = o[d2] + d[d2] * tval;
    if (is_le(p[d1], b[3 + d1]) && is_ge(p[d1], b[d1]) && is_le(p[d2], b[3 + d2]) && is_ge(p[d2], b[d2])) {
      if (ht) *ht = tval;
      if (hp) *hp = {p[0], p[1], p[2]};
      if (hn) {
        double n[3] = {0, 0, 0};
        n[dim] = is_neg(d[dim]) ? 1.0 : -1.0;
        *hn = {n[0], n[1], n[2]};
      }
      return true;
    }
  }
  return false;
}

// R3Intersects(ray, plane), R3Isect.cpp:700-732.  0 = miss, 1 = point, 2 = ray lies in the plane
static int ray_plane(const Ray &r, const Plane &pl, V3 *hp, double *ht) {
  double denom = dot(pl.n, r.d);
  if (is_zero(denom)) return is_zero(signed_distance(pl, r.o)) ? 2 : 0;
  double s = -(signed_distance(pl, r.o)) / denom;
  if (is_neg(s)) return 0;
  if (hp) *hp = r.o + r.d * s;
  if (ht) *ht = s;
  return 1;
}

// R3Intersects(ray, triangle) R3Isect.cpp:761-796 + R3Contains(triangle, point) R3Cont.cpp:491-514
static bool ray_triangle(const Ray &r, const double *v, V3 *hp, V3 *hn, double *ht) {
  V3 p0 = v3(v), p1 = v3(v + 3), p2 = v3(v + 6);
  Plane pl = plane_from_3_points(p0, p1, p2);                       // R3Triangle::Update, R3Triangle.cpp:475-497
  V3 p; double t;
  if (ray_plane(r, pl, &p, &t) != 1) return false;
  double bb[6] = {p0.x, p0.y, p0.z, p0.x, p0.y, p0.z};
  const V3 q[2] = {p1, p2};
  for (const V3 &c : q) {                                             // R3Box::Union(point), R3Box.cpp:455-471
    if (bb[0] > c.x) bb[0] = c.x; if (bb[1] > c.y) bb[1] = c.y; if (bb[2] > c.z) bb[2] = c.z;
    if (bb[3] < c.x) bb[3] = c.x; if (bb[4] < c.y) bb[4] = c.y; if (bb[5] < c.z) bb[5] = c.z;
  }
  if (!box_contains(bb, p)) return false;
  if (!is_zero(signed_distance(pl, p))) return false;
  Plane e01 = plane_from_point_2_vectors(p1, pl.n, p1 - p0);
  if (is_neg(signed_distance(e01, p))) return false;
  Plane e12 = plane_from_point_2_vectors(p2, pl.n, p2 - p1);
  if (is_neg(signed_distance(e12, p))) return false;
  Plane e20 = plane_from_point_2_vectors(p0, pl.n, p0 - p2);
  if (is_neg(signed_distance(e20, p))) return false;
  *hp = p; *hn = pl.n; *ht = t;
  return true;
}

// R3Intersects(ray, sphere) R3Isect.cpp:975-1021, R3Contains(sphere, point) R3Cont.cpp:1056-1063
static bool ray_sphere(const Ray &r, const double *s, V3 *hp, V3 *hn, double *ht) {
  V3 c = v3(s); double rad = s[3];
  double r2 = rad * rad;
  V3 vv = c - r.o;
  double d2 = vv.x * vv.x + vv.y * vv.y + vv.z * vv.z;
  bool start_inside = is_le(d2, r2);
  V3 EO = c - r.o;
  double v = dot(EO, r.d);
  if (!start_inside && is_neg0(v)) return false;
  double disc = rad * rad - (dot(EO, EO) - v * v);
  if (is_neg(disc)) return false;
  double d = sqrt(disc);                                              // NaN for disc in [-1e-6, 0): rejected upstream
  double t = start_inside ? v + d : v - d;
  V3 p = r.o + r.d * t;
  *ht = t; *hp = p; *hn = (p - c) / rad;
  return true;
}

// R3Intersects(ray, circle) R3Isect.cpp:837-879
static bool ray_circle(const Ray &r, const double *s, V3 *hp, V3 *hn, double *ht) {
  V3 c = v3(s), n = v3(s + 3); double rad = s[6];
  if (rad < 0.0) return false;                                        // IsEmpty
  Plane pl = plane_from_point_normal(c, n);
  V3 p; double t;
  int k = ray_plane(r, pl, &p, &t);
  if (k == 0) return false;
  if (k == 2) return false;   // reference returns R3_SPAN_CLASS_ID with t/point unset (undefined); we call it a miss
  double r2 = rad * rad;
  V3 v = c - p;
  double d2 = v.x * v.x + v.y * v.y + v.z * v.z;
  if (is_greater(d2, r2)) return false;
  *hp = p; *hn = n; *ht = t;
  return true;
}

// R3Intersects(ray, cylinder) R3Isect.cpp:1025-1197 (Graphics Gems IV p.356)
static bool ray_cylinder(const Ray &ray, const double *s, V3 *hp, V3 *hn, double *ht) {
  V3 a0 = v3(s), a1 = v3(s + 3); double radius = s[6];
  V3 av = a1 - a0;                                                    // R3Span ctor, R3Span.cpp:63-70
  double alen = length(av);
  if (!is_zero(alen)) av = av / alen;
  V3 A = av, R = ray.d;
  Plane base = plane_from_point_normal(a0, -A), top = plane_from_point_normal(a1, A);   // R3Cylinder.cpp:72-78
  double cyl_t1, cyl_t2;
  V3 D = cross(R, A);
  double a = length(D);
  if (is_zero(a)) {
    V3 w = cross(A, a0 - ray.o);                                      // R3Distance(point, line) R3Dist.cpp:59-65
    double d = length(w);
    if (is_le(d, radius)) { cyl_t1 = 0.0; cyl_t2 = RN_INF; }
    else return false;
  } else {
    D = D / a;
    V3 RC = ray.o - a0;
    double d = dot(RC, D);
    if (d < 0.0) d = -d;
    if (is_greater(d, radius)) return false;
    double t = dot(cross(RC, A), D) / -a;
    double sdelta;
    V3 O = normalized(cross(D, A));
    double b = radius * radius - d * d;
    if (is_pos(b)) {
      double e = dot(R, O);
      sdelta = sqrt(b) / e;
      if (sdelta < 0.0) sdelta = -sdelta;
    } else sdelta = 0.0;
    cyl_t2 = t + sdelta;
    if (is_neg(cyl_t2)) return false;
    cyl_t1 = t - sdelta;
    if (cyl_t1 < 0.0) cyl_t1 = 0.0;
  }
  double cap_t1 = 0, cap_t2 = 0;
  double dt = dot(ray.d, A);
  double d_top = signed_distance(top, ray.o);
  double d_base = signed_distance(base, ray.o);
  auto neg = [](const Plane &p) { return Plane{-p.n, -p.d}; };
  if (is_pos(d_top)) {
    if (is_pos0(dt)) return false;
    if (ray_plane(ray, top, nullptr, &cap_t1) && is_greater(cap_t1, cyl_t2)) return false;
    else if (ray_plane(ray, neg(base), nullptr, &cap_t2) && is_less(cap_t2, cyl_t1)) return false;
    else if (is_greater(cap_t1, cyl_t1)) { *ht = cap_t1; *hp = ray.o + ray.d * cap_t1; *hn = top.n; return true; }
  } else if (is_pos(d_base)) {
    if (is_neg0(dt)) return false;
    if (ray_plane(ray, base, nullptr, &cap_t1) && is_greater(cap_t1, cyl_t2)) return false;
    else if (ray_plane(ray, neg(top), nullptr, &cap_t2) && is_less(cap_t2, cyl_t1)) return false;
    else if (is_greater(cap_t1, cyl_t1)) { *ht = cap_t1; *hp = ray.o + ray.d * cap_t1; *hn = base.n; return true; }
  } else {
    if (is_pos(dt)) {
      if (ray_plane(ray, neg(top), nullptr, &cap_t1)) { if (is_less(cap_t1, cyl_t1)) return false; }
    } else if (is_neg(dt)) {
      if (ray_plane(ray, neg(base), nullptr, &cap_t1)) { if (is_less(cap_t1, cyl_t1)) return false; }
    }
  }
  *ht = cyl_t1;
  V3 H = ray.o + ray.d * cyl_t1;
  *hp = H;
  V3 HB = H - a0;
  *hn = (HB - A * dot(HB, A)) / radius;
  return true;
}

// one R3Shape::Intersects.  Mesh = R3Intersects(ray, R3TriangleArray) R3Isect.cpp:800-833
static bool shape_intersects(const Scene &s, int sid, const Ray &r, V3 *hp, V3 *hn, double *ht, int *hprim) {
  const GiShape &sh = s.shapes[sid];
  if (sh.is_mesh) {
    if (!ray_box(r, sh.bbox, nullptr, nullptr, nullptr)) return false;
    bool status = false;
    double min_t = FLT_MAX;
    for (int i = 0; i < sh.n_prims; i++) {
      V3 p, n; double t;
      if (ray_triangle(r, s.prims[sh.first_prim + i].d, &p, &n, &t)) {
        if (t < min_t) { status = true; *hp = p; *hn = n; min_t = t; *hprim = sh.first_prim + i; }
      }
    }
    *ht = min_t;
    return status;
  }
  const GiPrim &pr = s.prims[sh.first_prim];
  *hprim = sh.first_prim;
  switch (pr.type) {
    case GI_SHAPE_TRIANGLE: return ray_triangle(r, pr.d, hp, hn, ht);
    case GI_SHAPE_SPHERE: return ray_sphere(r, pr.d, hp, hn, ht);
    case GI_SHAPE_BOX: return ray_box(r, pr.d, hp, hn, ht);
    case GI_SHAPE_CYLINDER: return ray_cylinder(r, pr.d, hp, hn, ht);
    case GI_SHAPE_CIRCLE: return ray_circle(r, pr.d, hp, hn, ht);
    default: return false;     // cone: incomplete in the reference (R3Isect.cpp:1259-1274), unsupported
  }
}

// R3SceneElement::Intersects, R3SceneElement.cpp:209-243
static bool element_intersects(const Scene &s, int eid, const Ray &r, Hit *h, double min_t, double max_t) {
  const GiElement &e = s.elements[eid];
  double bbox_t, closest_t = max_t;
  if (!box_contains(e.bbox, r.o)) {
    if (!ray_box(r, e.bbox, nullptr, nullptr, &bbox_t)) return false;
    if (is_greater(bbox_t, max_t)) return false;
  }
  for (int sid : s.element_shapes[eid]) {
    V3 p, n; double t; int prim = -1;
    if (shape_intersects(s, sid, r, &p, &n, &t, &prim)) {
      if ((t >= min_t) && (t <= closest_t)) {
        h->shape = sid; h->prim = prim; h->p = p; h->n = n; h->t = t;
        closest_t = t;
      }
    }
  }
  return !(closest_t == max_t);
}

// R3SceneNode::Intersects, R3SceneNode.cpp:420-510
static bool node_intersects(const Scene &s, int nid, const Ray &ray, Hit *out, double min_t, double max_t) {
  const GiNode &node = s.nodes[nid];
  bool found = false;
  V3 closest_point = {0, 0, 0};
  double closest_t = max_t, bbox_t;
  if (!box_contains(node.bbox, ray.o)) {
    if (!ray_box(ray, node.bbox, nullptr, nullptr, &bbox_t)) return false;
    if (is_greater(bbox_t, max_t)) return false;
  }
  Ray nr;                                                             // R3Line::InverseTransform, R3Line.cpp:139-146
  nr.o = xf_point(node.minv, ray.o);
  nr.d = normalized(xf_vector(node.minv, ray.d));
  double scale = 1.0;
  V3 v = xf_vector(node.m, ray.d);                                    // sic: the FORWARD matrix (R3SceneNode.cpp:450-451)
  double len = length(v);
  if (is_neg0(len)) return false;
  if (!is_zero(len - 1.0)) { scale = len; min_t /= scale; closest_t /= scale; }
  Hit h;
  for (int eid : s.node_elements[nid]) {
    if (element_intersects(s, eid, nr, &h, min_t, closest_t)) {
      if ((h.t >= min_t) && (h.t <= closest_t)) {
        out->shape = h.shape; out->prim = h.prim; out->n = h.n;
        found = true; closest_point = h.p; closest_t = h.t;
      }
    }
  }
  for (int cid : s.node_children[nid]) {
    if (node_intersects(s, cid, nr, &h, min_t, closest_t)) {
      if ((h.t >= min_t) && (h.t <= closest_t)) {
        out->shape = h.shape; out->prim = h.prim; out->n = h.n;
        found = true; closest_point = h.p; closest_t = h.t;
      }
    }
  }
  if (!found) return false;
  out->p = xf_point(node.m, closest_point);
  out->t = scale * closest_t;
  out->n = normalized(xf_vector(node.m, out->n));                     // normals go through M, not M^-T (:503-505)
  return true;
}

bool scene_intersects(const Scene &s, const double *ray6, Hit *h) {
  Ray r = {v3(ray6), v3(ray6 + 3)};
  h->hit = node_intersects(s, 0, r, h, 0.0, RN_INF);                  // min_t = 0, max_t = RN_INFINITY (R3Scene.h:73-76)
  return h->hit;
}

}  // namespace oracle

// ------------------------------------------------------------------------------------------ C entry points
extern "C" {

void *oracle_scene_load(const char *gisc_path) {
  oracle::Scene *s = new oracle::Scene();
  std::string err;
  if (!oracle::read_gisc(gisc_path, *s, err)) { fprintf(stderr, "oracle: %s\n", err.c_str()); delete s; return nullptr; }
  return s;
}
void oracle_scene_free(void *scene) { delete (oracle::Scene *)scene; }

// rays: n x 6 doubles (origin, unit direction); hits: n x GiHit
int oracle_intersect(void *scene, const double *rays, uint64_t n, GiHit *hits, int threads) {
  if (!scene) return 1;
  const oracle::Scene &s = *(const oracle::Scene *)scene;
  if (threads < 1) threads = 1;
  auto work = [&](int id) {
    for (uint64_t i = id; i < n; i += threads) {
      oracle::Hit h; memset(&h, 0, sizeof(h));
      GiHit g; memset(&g, 0, sizeof(g)); g.shape = -1; g.prim = -1; g.element = -1;
      if (oracle::scene_intersects(s, rays + 6 * i, &h)) {
        g.hit = 1; g.shape = h.shape; g.prim = h.prim; g.element = s.shapes[h.shape].element; g.t = h.t;
        g.p[0] = h.p.x; g.p[1] = h.p.y; g.p[2] = h.p.z; g.n[0] = h.n.x; g.n[1] = h.n.y; g.n[2] = h.n.z;
      }
      hits[i] = g;
    }
  };
  std::vector<std::thread> th;
  for (int i = 1; i < threads; i++) th.emplace_back(work, i);
  work(0);
  for (auto &t : th) t.join();
  return 0;
}

}  // extern "C"
