// oracle/restate_common.h -- TEST INFRASTRUCTURE (never linked into, loaded by, or timed as the product,
// except as bench.py's cpu_baseline / --impl reference leg).
//
// Shared pieces of the CPU restatement of the reference's hot path: epsilon compares, small vector
// algebra in the reference's operation order, and a reader for the flattened-scene (GISC) files.
// Everything is double precision like the reference (RNScalar = double, RNBasics/RNCompat.h:82-88) and
// is compiled with -ffp-contract=off so that no multiply-add is fused (the reference's x86-64 build
// has no FMA either).
#pragma once
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include "gi_b200.h"

namespace oracle {

// RNBasics/RNScalar.cpp:19-24
static const double EPS = 1.0E-6;       // RN_EPSILON
static const double RN_INF = 1.0E6;     // RN_INFINITY

// RNBasics/RNScalar.h:224-330
inline bool is_pos(double s) { return s > EPS; }
inline bool is_neg(double s) { return s < -EPS; }
inline bool is_pos0(double s) { return s >= -EPS; }
inline bool is_neg0(double s) { return s <= EPS; }
inline bool is_zero(double s) { return is_pos0(s) && is_neg0(s); }
inline bool is_greater(double a, double b) { return is_pos(a - b); }
inline bool is_less(double a, double b) { return is_neg(a - b); }
inline bool is_ge(double a, double b) { return is_pos0(a - b); }
inline bool is_le(double a, double b) { return is_neg0(a - b); }

struct V3 {
  double x, y, z;
  double operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
inline V3 v3(const double *p) { return {p[0], p[1], p[2]}; }
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator*(V3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator/(V3 a, double s) { return {a.x / s, a.y / s, a.z / s}; }
inline V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
inline double dot(V3 a, V3 b) { return (a.x * b.x) + (a.y * b.y) + (a.z * b.z); }   // R3Vector.cpp:206-209
inline V3 cross(V3 a, V3 b) {                                                         // R3Vector.cpp:255-261
  return {(a.y * b.z) - (a.z * b.y), (a.z * b.x) - (a.x * b.z), (a.x * b.y) - (a.y * b.x)};
}
inline double length(V3 a) { return sqrt((a.x * a.x) + (a.y * a.y) + (a.z * a.z)); }  // R3Vector.cpp:101-104
inline V3 normalized(V3 a) {                                                          // R3Vector.cpp:232-239
  double l = length(a);
  if (l == 0.0) return a;
  return {a.x / l, a.y / l, a.z / l};
}
// R4Matrix * R3Point / R3Vector (R4Matrix.cpp:774-793), row-major m[16]
inline V3 xf_point(const double *m, V3 p) {
  return {m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3], m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7],
          m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]};
}
inline V3 xf_vector(const double *m, V3 v) {
  return {m[0] * v.x + m[1] * v.y + m[2] * v.z, m[4] * v.x + m[5] * v.y + m[6] * v.z,
          m[8] * v.x + m[9] * v.y + m[10] * v.z};
}

struct Plane { V3 n; double d; };
inline Plane plane_from_point_normal(V3 p, V3 n) {                    // R3Plane.cpp:84-90
  return {n, -(n.x * p.x + n.y * p.y + n.z * p.z)};
}
inline Plane plane_from_3_points(V3 p1, V3 p2, V3 p3) {               // R3Plane.cpp:117-126
  V3 v = normalized(cross(p2 - p1, p3 - p1));
  return {v, -(v.x * p1.x + v.y * p1.y + v.z * p1.z)};
}
inline Plane plane_from_point_2_vectors(V3 p, V3 v1, V3 v2) {         // R3Plane.cpp:106-113
  V3 v = normalized(cross(v1, v2));
  return {v, -(v.x * p.x + v.y * p.y + v.z * p.z)};
}
inline double signed_distance(const Plane &pl, V3 p) {                // R3Dist.cpp:603-607
  return (p.x * pl.n.x + p.y * pl.n.y + p.z * pl.n.z + pl.d);
}

// flattened scene read from a GISC file (format: global-illumination_b200/host/scene_io.cpp)
struct Scene {
  std::vector<GiNode> nodes;
  std::vector<GiElement> elements;
  std::vector<GiMaterial> materials;
  std::vector<GiPrim> prims;
  std::vector<GiShape> shapes;
  std::vector<GiLight> lights;
  GiCamera camera;
  double ambient[3], background[3], bbox_min[3], bbox_max[3], centroid[3], radius;
  // derived: children / elements / shapes per node in insertion order
  std::vector<std::vector<int>> node_children, node_elements, element_shapes;
};

template <class T> static bool rd_vec(FILE *f, std::vector<T> &v, int n) {
  if (n < 0) return false;
  v.resize((size_t)n);
  return n == 0 || fread(v.data(), sizeof(T), (size_t)n, f) == (size_t)n;
}

inline bool read_gisc(const char *path, Scene &s, std::string &err) {
  FILE *f = fopen(path, "rb");
  if (!f) { err = std::string("cannot open ") + path; return false; }
  uint32_t hdr[2]; int32_t c[6]; double tail[16];
  bool ok = fread(hdr, 4, 2, f) == 2 && hdr[0] == 0x43534947u && hdr[1] == 2 && fread(c, 4, 6, f) == 6;
  ok = ok && rd_vec(f, s.nodes, c[0]) && rd_vec(f, s.elements, c[1]) && rd_vec(f, s.materials, c[2]) &&
       rd_vec(f, s.prims, c[3]) && rd_vec(f, s.shapes, c[4]) && rd_vec(f, s.lights, c[5]) &&
       fread(&s.camera, sizeof(GiCamera), 1, f) == 1 && fread(tail, 8, 16, f) == 16;
  fclose(f);
  if (!ok) { err = std::string("bad GISC file ") + path; return false; }
  memcpy(s.ambient, tail, 24); memcpy(s.background, tail + 3, 24); memcpy(s.bbox_min, tail + 6, 24);
  memcpy(s.bbox_max, tail + 9, 24); memcpy(s.centroid, tail + 12, 24); s.radius = tail[15];
  s.node_children.assign(s.nodes.size(), {});
  s.node_elements.assign(s.nodes.size(), {});
  s.element_shapes.assign(s.elements.size(), {});
  for (size_t i = 0; i < s.nodes.size(); i++)
    if (s.nodes[i].parent >= 0) s.node_children[s.nodes[i].parent].push_back((int)i);
  for (size_t i = 0; i < s.elements.size(); i++) s.node_elements[s.elements[i].node].push_back((int)i);
  for (size_t i = 0; i < s.shapes.size(); i++) s.element_shapes[s.shapes[i].element].push_back((int)i);
  return true;
}

}  // namespace oracle
