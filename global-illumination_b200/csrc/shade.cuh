// csrc/shade.cuh -- per-ray shading arithmetic of the renderer (device code), shared by the render stages and the
// photon tracer.  Follows, function by function:
//   src/utils/graphics_utils.cpp:95-216  Schlick Fresnel, reflective / transmissive bounce, cosine-hemisphere and
//                                        Phong-lobe importance sampling
//   src/utils/illumination_utils.cpp     shadow test, point/spot/directional reflection, sampled area / rect lights
//   src/R3Graphics/R3{Point,Spot,Directional}Light.cpp  Reflection()
// All arithmetic is FP64 like the reference.  The random numbers are NOT the reference's (a thread-local mt19937
// seeded from random_device, RNScalar.cpp:99-131): every stochastic decision here draws from a counter-based
// Philox4x32-10 stream keyed by (seed, pixel or photon index, stage, sample), so results do not depend on launch
// geometry, batch size or GPU count.
#pragma once
#include "gi_b200.h"
#include "isect.cuh"
#include "rng.cuh"

namespace gi {

// ------------------------------------------------------------------------------------------------ colours
GI_DEV double max_channel(D3 c) {                       // MaxChannelVal, graphics_utils.cpp:38-46
  double m = 0;
  if (c.x > m) m = c.x;
  if (c.y > m) m = c.y;
  if (c.z > m) m = c.z;
  return m;
}

// RNRgb_to_RGBE, graphics_utils.cpp:50-62
GI_DEV void rgbe_encode(D3 c, uint8_t *out) {
  const double mx = max_channel(c);
  if (!(mx > 0)) { out[0] = out[1] = out[2] = 0; int e; frexp(0.0, &e); out[3] = (uint8_t)(e + 128); return; }
  int exponent;
  const double mantissa = frexp(mx, &exponent);
  out[0] = (uint8_t)(int)(256.0 * c.x / mx * mantissa);
  out[1] = (uint8_t)(int)(256.0 * c.y / mx * mantissa);
  out[2] = (uint8_t)(int)(256.0 * c.z / mx * mantissa);
  out[3] = (uint8_t)(exponent + 128);
}
// RGBE_to_RNRgb, graphics_utils.cpp:65-79
GI_DEV D3 rgbe_decode(const uint8_t *in) {
  if (!in[3]) return mk3(0, 0, 0);
  const double inverse = ldexp(1.0, ((int)in[3]) - 128 - 8);
  return mk3(in[0], in[1], in[2]) * inverse;
}

GI_DEV bool is_black(const double *c) { return c[0] == 0.0 && c[1] == 0.0 && c[2] == 0.0; }   // RNRgb::IsBlack

struct Brdf {                                           // R3Brdf values + the flags RayTrace tests
  D3 ka, kd, ks, kt, ke;
  double n, ir;
  bool ambient, diffuse, specular, transparent;
};
GI_DEV Brdf load_brdf(const DevMaterial &m) {
  Brdf b;
  b.ka = ld3(m.ka); b.kd = ld3(m.kd); b.ks = ld3(m.ks); b.kt = ld3(m.kt); b.ke = ld3(m.ke);
  b.n = m.shininess; b.ir = m.ir;
  b.ambient = !is_black(m.ka); b.diffuse = !is_black(m.kd); b.specular = !is_black(m.ks); b.transparent = !is_black(m.kt);
  return b;
}

// ------------------------------------------------------------------------------------------------ optics
struct Optics { int fresnel; double ir_air; };

GI_DEV double reflection_coeff(double cos_theta, double ir_mat, double ir_air) {             // graphics_utils.cpp:95-101
  const double q = (ir_air - ir_mat) / (ir_air + ir_mat);
  const double R_o = q * q;
  const double w = 1.0 - fabs(cos_theta);
  return (R_o + (1.0 - R_o) * (w * w * w * w * w));
}

GI_DEV D3 reflective_bounce(D3 normal, D3 view, double cos_theta) {                         // graphics_utils.cpp:104-117
  if (cos_theta < 0) { normal = -normal; cos_theta *= -1.0; }
  return normalized(view + (normal * cos_theta) * 2.0);
}

GI_DEV D3 transmissive_bounce(D3 normal, D3 view, double cos_theta, double ir_mat, double ir_air) {   // :121-154
  double eta;
  if (cos_theta < 0) { eta = ir_mat / ir_air; normal = -normal; cos_theta *= -1.0; }
  else eta = ir_air / ir_mat;
  const double theta = acos(cos_theta);
  const double sin_phi = eta * sin(theta);
  if (sin_phi < -1.0 || 1.0 < sin_phi) return reflective_bounce(normal, view, cos_theta);    // total internal reflection
  const double phi = asin(sin_phi);
  const D3 par = normalized(view + normal * cos_theta);
  return normalized(par * tan(phi) - normal);
}

GI_DEV D3 rotate_about(D3 v, D3 axis, double theta) {                                       // R3Vector::Rotate, R3Vector.cpp:352-363
  const double c = cos(theta), d = dot(v, axis);
  const D3 cr = cross(v, axis);
  D3 r = v * c;
  r = r + (axis * d) * (1.0 - c);
  r = r - cr * sin(theta);
  return r;
}

GI_DEV D3 perpendicular_to(D3 a) {                                                          // graphics_utils.cpp:174-178
  D3 p = mk3(a.y, -a.x, 0);
  if (1.0 - fabs(a.z) < 0.1) p = mk3(a.z, 0, -a.x);
  return normalized(p);
}

GI_DEV D3 diffuse_sample(D3 normal, double cos_theta, Rng &rng) {                           // graphics_utils.cpp:162-185
  if (cos_theta < 0) normal = -normal;
  const double theta = acos(sqrt(rng_double(rng)));
  const double phi = 2 * 3.14159265358979323846 * rng_double(rng);
  const D3 perp = perpendicular_to(normal);
  D3 r = perp * sin(theta) + normal * cos(theta);
  return normalized(rotate_about(r, normal, phi));
}

GI_DEV D3 specular_sample(D3 exact, double n, double cos_theta, Rng &rng) {                 // graphics_utils.cpp:189-216
  const double PI = 3.14159265358979323846;
  const double angle_limit = (1.0 - acos(fabs(cos_theta)) * 2.0 / PI);
  const double alpha = acos(pow(rng_double(rng), 1.0 / (n + 1.0))) * angle_limit;
  const double phi = 2.0 * PI * rng_double(rng);
  const D3 perp = perpendicular_to(exact);
  D3 r = perp * sin(alpha) + exact * cos(alpha);
  return normalized(rotate_about(r, exact, phi));
}

// ------------------------------------------------------------------------------------------------ lights
struct RayStats { unsigned long long screen, shadow, monte, transmissive, specular, indirect, caustic; };

struct ShadeEnv {                 // everything a shading thread needs besides the hit itself
  DevScene sc;
  const GiLight *lights; int n_lights;
  const DevMaterial *materials;
  D3 ambient, background, eye;    // eye = camera origin (RayTrace's `eye`, render.cpp:187)
  double radius;                  // SCENE_RADIUS
  const double *spec_boxes;       // world boxes (min xyz, max xyz) of the shapes with ks or kt not black; photon tracing only
  int n_spec_boxes;               // -1: list not available (too many shapes)
  GiParams p;
};

// RayIlluminationTest, illumination_utils.cpp:16-31: first hit of the ray light -> surface must lie at the surface
GI_DEV bool illumination_test(const ShadeEnv &env, D3 point_in_scene, D3 point_on_light, RayStats &st) {
  const double unoccluded = length(point_on_light - point_in_scene);
  RayD ray; ray.o = point_on_light; ray.d = normalized(point_in_scene - point_on_light);     // R3Line(p1, p2), R3Line.cpp:67-73
  HitRec h;
  closest_hit(env.sc, ray, h);
  const double len = h.prim >= 0 ? length(point_on_light - h.p) : GI_INFINITY;               // IntersectionDist, graphics_utils.cpp:84-92
  st.shadow++;
  return fabs(len - unoccluded) < GI_EPS;
}

// TestLightIntersection, illumination_utils.cpp:35-84
GI_DEV int test_light_intersection(const GiLight &l, D3 point, D3 eye) {
  if (l.type == GI_LIGHT_AREA) {
    D3 v = point - ld3(l.position);
    const double vlen = length(v);
    v = normalized(v);
    const D3 norm = ld3(l.direction);
    if (fabs(dot(v, norm)) < GI_EPS && vlen <= l.radius) return dot(norm, eye - point) <= 0 ? -1 : 1;
  } else if (l.type == GI_LIGHT_RECT) {
    D3 v = point - ld3(l.position);
    const double a1 = dot(v, ld3(l.axis1)), a2 = dot(v, ld3(l.axis2));
    v = normalized(v);
    const D3 norm = ld3(l.direction);
    if (fabs(dot(v, norm)) < GI_EPS && fabs(a1 * 2.0) <= l.len1 && fabs(a2 * 2.0) <= l.len2)
      return dot(norm, eye - point) <= 0 ? -1 : 1;
  }
  return 0;
}

GI_DEV double point_intensity(const GiLight &l, D3 point) {                                  // R3PointLight.cpp:112-121
  const double d = length(point - ld3(l.position));
  double denom = l.ca;
  denom += d * l.la;
  denom += d * d * l.qa;
  if (is_zero(denom)) return l.intensity;
  return l.intensity / denom;
}

// R3{Point,Spot,Directional}Light::Reflection (R3PointLight.cpp:213-244, R3SpotLight.cpp:105-115,
// R3DirectionalLight.cpp:134-166)
GI_DEV D3 simple_light_reflection(const GiLight &l, const Brdf &b, D3 eye, D3 point, D3 normal) {
  if (!l.active) return mk3(0, 0, 0);
  double I; D3 L;
  if (l.type == GI_LIGHT_DIR) { I = l.intensity; L = -ld3(l.direction); }
  else {
    I = point_intensity(l, point);
    if (l.type == GI_LIGHT_SPOT) {
      const D3 ml = normalized(point - ld3(l.position));
      const double cos_alpha = dot(ml, ld3(l.direction));
      if (cos(l.cutoff) > cos_alpha) I = 0.0; else I = I * pow(cos_alpha, l.dropoff);
    }
    L = normalized(ld3(l.position) - point);
  }
  const D3 Ic = ld3(l.color);
  const double NL = dot(normal, L);
  const D3 R = normal * (2.0 * NL) - L;
  const D3 V = normalized(eye - point);
  const double VR = dot(V, R);
  D3 rgb = mul3(b.kd, Ic) * (I * fabs(NL));
  if (is_pos(VR)) rgb = rgb + mul3(b.ks, Ic) * (I * pow(VR, b.n));
  return rgb;
}

// Where the random numbers of one light-sampling loop come from.  Sequential (key == nullptr): the calling path's own
// stream, sample after sample, as the reference draws them.  Cooperative (key != nullptr): the 32 lanes of a warp
// share one shading point, lane l takes samples l, l+32, ... and sample i of pass p draws from its own Philox
// stream (seed, key->index, key->stage + p, i), so the result does not depend on which lane evaluated it.
struct SampleKey { uint64_t seed, index; uint32_t stage; };

// one sample point on an area (disk) or rect light, offset by 1e-6 along the emitting normal
GI_DEV D3 sample_light_point(const GiLight &l, D3 u, D3 v, Rng &rng) {
  double r1, r2;
  if (l.type == GI_LIGHT_AREA) {
    do { r1 = (rng_double(rng) * 2.0) - 1.0; r2 = (rng_double(rng) * 2.0) - 1.0; } while (r1 * r1 + r2 * r2 > 1.0);
  } else { r1 = rng_double(rng) - 0.5; r2 = rng_double(rng) - 0.5; }
  return u * r1 + v * r2 + ld3(l.position) + ld3(l.direction) * GI_EPS;
}

GI_DEV double warp_sum(double v) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
GI_DEV int warp_sum(int v) { return (int)__reduce_add_sync(0xffffffffu, (unsigned)v); }

// ComputeAreaLightReflection / ComputeRectLightReflection, illumination_utils.cpp:91-262 / 265-417.
// `occlusion` = false stands in for the sampled R3{Area,Rect}Light::Reflection used when shadows are off
// (R3AreaLight.cpp:118-319: the same uniform-on-the-light estimator, mean x area, without the shadow rays).
// NOTE the last line: the hit rate rescales the WHOLE colour accumulated so far, as in the reference (:259-261).
// With key != nullptr all 32 lanes must call this with identical arguments; every lane returns the same colour.
GI_DEV void sampled_light_reflection(const ShadeEnv &env, const GiLight &l, D3 &color, const Brdf &b, D3 eye, D3 point,
                                     D3 normal, int n_light, int n_shadow, bool occlusion, Rng &rng,
                                     const SampleKey *key, RayStats &st) {
  if (!l.active) return;
  const D3 center = ld3(l.position), ln = ld3(l.direction);
  if (dot(ln, point - center) < 0) return;
  D3 u, v; double area;
  if (l.type == GI_LIGHT_AREA) {
    u = mk3(ln.y, -ln.x, 0);
    if (1.0 - fabs(ln.z) < 0.1) u = mk3(ln.z, 0, -ln.x);
    v = cross(u, ln);
    u = normalized(u) * l.radius; v = normalized(v) * l.radius;
    area = 3.14159265358979323846 * (l.radius * l.radius);
  } else {
    u = ld3(l.axis1) * l.len1; v = ld3(l.axis2) * l.len2;
    area = length(cross(u, v));
  }
  const int first = key ? (int)(threadIdx.x & 31) : 0, stride = key ? 32 : 1;
  const D3 Ic = ld3(l.color);
  int total_samples = 0, total_hits = 0;
  const D3 V = normalized(eye - point);
  for (int pass = 0; pass < 3; pass++) {                 // 0: diffuse, 1: specular (twice the samples), 2: shadow only
    if (pass == 0 && !b.diffuse) continue;
    if (pass == 1 && !b.specular) continue;
    if (pass == 2 && !occlusion) continue;
    if (pass == 1) n_light *= 2;
    const int n = pass == 2 ? n_shadow : n_light;
    double weight = 0; int hits = 0;
    for (int i = first; i < n; i += stride) {
      Rng own;
      if (key) rng_init(own, key->seed, key->index, key->stage + pass, (uint32_t)i);
      const D3 sp = sample_light_point(l, u, v, key ? own : rng);
      if (occlusion && !illumination_test(env, point, sp, st)) continue;
      hits++;
      if (pass == 2) continue;
      double I = l.intensity;
      const double d = length(point - sp);
      double denom = l.ca; denom += d * l.la; denom += d * d * l.qa;
      if (is_pos(denom)) I /= denom;
      const D3 L = normalized(sp - point);
      I *= dot(ln, -L) * 2.0;
      if (pass == 0) weight += I * fabs(dot(normal, L));
      else {
        const double NL = dot(normal, L);
        const D3 R = normal * (2.0 * NL) - L;
        const double VR = dot(V, R);
        if (is_neg0(VR)) continue;
        weight += (I * pow(VR, b.n));
      }
    }
    if (key) { __syncwarp(); weight = warp_sum(weight); hits = warp_sum(hits); }
    if (pass < 2 && hits > 0) color = color + mul3(pass == 0 ? b.kd : b.ks, Ic) * (weight * area / hits);
    total_hits += hits; total_samples += n;
  }
  if (occlusion && total_samples > 0) color = color * (((double)total_hits) / total_samples);
}

// ComputeIllumination, illumination_utils.cpp:425-494
GI_DEV void compute_illumination(const ShadeEnv &env, D3 &color, const GiLight &l, const Brdf &b, D3 eye, D3 point,
                                 D3 normal, double cos_theta, bool in_mc, Rng &rng, const SampleKey *key, RayStats &st) {
  const GiParams &p = env.p;
  const bool shadows = p.shadows && (!in_mc || p.recursive_shadows);
  int n_light = p.light_test, n_shadow = p.shadow_test;
  if (in_mc) { n_light = 2; n_shadow = 0; }
  const bool two_d = l.type == GI_LIGHT_AREA || l.type == GI_LIGHT_RECT;
  if (!shadows) {
    if (two_d) sampled_light_reflection(env, l, color, b, eye, point, normal, n_light, 0, false, rng, key, st);
    else color = color + simple_light_reflection(l, b, eye, point, normal);
    return;
  }
  D3 pl;
  if (l.type == GI_LIGHT_DIR) pl = point - ld3(l.direction) * env.radius * 3.0;
  else if (two_d) {
    if (p.soft_shadows) { sampled_light_reflection(env, l, color, b, eye, point, normal, n_light, n_shadow, true, rng, key, st); return; }
    pl = ld3(l.position) + ld3(l.direction) * GI_EPS;
  } else pl = ld3(l.position);
  const double side = dot(normal, pl - point);
  if ((side > 0 && cos_theta < 0) || (side < 0 && cos_theta > 0)) return;
  const bool lit = illumination_test(env, point, pl, st);
  if (key && (threadIdx.x & 31) != 0) st.shadow--;        // cooperative call: the 32 lanes repeat ONE shadow ray
  if (lit) {
    if (two_d) sampled_light_reflection(env, l, color, b, eye, point, normal, n_light, 0, false, rng, key, st);
    else color = color + simple_light_reflection(l, b, eye, point, normal);
  }
}

// DirectIllumination, raytracer.cpp:18-44
GI_DEV void direct_illumination(const ShadeEnv &env, D3 point, D3 normal, D3 eye, D3 &color, const Brdf &b,
                                double cos_theta, bool in_mc, Rng &rng, const SampleKey *key, RayStats &st) {
  bool should_emit = true;
  for (int k = 0; k < env.n_lights; k++) {
    const GiLight &l = env.lights[k];
    const int li = test_light_intersection(l, point, eye);
    if (li != 0) { if (li == -1) should_emit = false; continue; }
    SampleKey lk;
    if (key) { lk = *key; lk.stage += 4u * (uint32_t)k; }
    compute_illumination(env, color, l, b, eye, point, normal, cos_theta, in_mc, rng, key ? &lk : nullptr, st);
  }
  if (should_emit) color = color + b.ke;
}

}  // namespace gi
