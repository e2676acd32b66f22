// csrc/gi_math.cuh -- double-precision vector algebra for the parity-critical code.
//
// Every decision of the reference's intersectors is an epsilon compare on doubles (RN_EPSILON = 1e-6,
// src/RNBasics/RNScalar.h:224-330), so the hit/miss result and the hit id depend on the last bits of
// the arithmetic.  These helpers keep the reference's operation ORDER (cited per function) and the
// library is compiled with -fmad=false, so that sm_100a's IEEE add/mul/div/sqrt reproduce the
// reference's x86-64 results bit for bit.  GI_HD lets tests/host_sim compile the same code for the
// CPU (a debugging aid used only by the CPU test-suite -- the product has no CPU path).
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define GI_HD __host__ __device__ __forceinline__
#else
#define GI_HD inline
#endif

namespace gi {

#define GI_EPS 1.0E-6          /* RN_EPSILON,  RNScalar.cpp:19 */
#define GI_INFINITY 1.0E6      /* RN_INFINITY, RNScalar.cpp:21 */

GI_HD bool is_pos(double s) { return s > GI_EPS; }
GI_HD bool is_neg(double s) { return s < -GI_EPS; }
GI_HD bool is_pos0(double s) { return s >= -GI_EPS; }
GI_HD bool is_neg0(double s) { return s <= GI_EPS; }
GI_HD bool is_zero(double s) { return (s >= -GI_EPS) && (s <= GI_EPS); }

struct D3 { double x, y, z; };

GI_HD D3 mk3(double x, double y, double z) { D3 r; r.x = x; r.y = y; r.z = z; return r; }
GI_HD D3 ld3(const double *p) { return mk3(p[0], p[1], p[2]); }
GI_HD D3 operator+(D3 a, D3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
GI_HD D3 operator-(D3 a, D3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
GI_HD D3 operator-(D3 a) { return mk3(-a.x, -a.y, -a.z); }
GI_HD D3 operator*(D3 a, double s) { return mk3(a.x * s, a.y * s, a.z * s); }
GI_HD D3 operator/(D3 a, double s) { return mk3(a.x / s, a.y / s, a.z / s); }
GI_HD D3 mul3(D3 a, D3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
GI_HD double comp(D3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }
GI_HD double dot(D3 a, D3 b) { return (a.x * b.x) + (a.y * b.y) + (a.z * b.z); }          // R3Vector.cpp:206-209
GI_HD D3 cross(D3 a, D3 b) {                                                                // R3Vector.cpp:255-261
  return mk3((a.y * b.z) - (a.z * b.y), (a.z * b.x) - (a.x * b.z), (a.x * b.y) - (a.y * b.x));
}
GI_HD double length(D3 a) { return sqrt((a.x * a.x) + (a.y * a.y) + (a.z * a.z)); }        // R3Vector.cpp:101-104
GI_HD D3 normalized(D3 a) {                                                                 // R3Vector.cpp:232-239
  double l = length(a);
  if (l == 0.0) return a;
  return mk3(a.x / l, a.y / l, a.z / l);
}
// R4Matrix * R3Point / R3Vector (R4Matrix.cpp:774-793); m is row-major 4x4
GI_HD D3 xf_point(const double *m, D3 p) {
  return mk3(m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3], m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7],
             m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]);
}
GI_HD D3 xf_vector(const double *m, D3 v) {
  return mk3(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[4] * v.x + m[5] * v.y + m[6] * v.z,
             m[8] * v.x + m[9] * v.y + m[10] * v.z);
}

struct Plane4 { double a, b, c, d; };                                                       // R3Plane: normal + d
GI_HD D3 pn(const Plane4 &p) { return mk3(p.a, p.b, p.c); }
GI_HD Plane4 plane_point_normal(D3 p, D3 n) {                                               // R3Plane.cpp:84-90
  Plane4 r; r.a = n.x; r.b = n.y; r.c = n.z; r.d = -(n.x * p.x + n.y * p.y + n.z * p.z); return r;
}
GI_HD Plane4 plane_3_points(D3 p1, D3 p2, D3 p3) {                                          // R3Plane.cpp:117-126
  return plane_point_normal(p1, normalized(cross(p2 - p1, p3 - p1)));
}
GI_HD Plane4 plane_point_2_vectors(D3 p, D3 v1, D3 v2) {                                    // R3Plane.cpp:106-113
  return plane_point_normal(p, normalized(cross(v1, v2)));
}
GI_HD Plane4 plane_neg(const Plane4 &p) { Plane4 r; r.a = -p.a; r.b = -p.b; r.c = -p.c; r.d = -p.d; return r; }
GI_HD double signed_distance(const Plane4 &pl, D3 p) {                                      // R3Dist.cpp:603-607
  return (p.x * pl.a + p.y * pl.b + p.z * pl.c + pl.d);
}

}  // namespace gi
