// csrc/debug_view.cu -- gi_debug_view: the reference's debug views (src/visualize.cpp) as headless CUDA raster kernels.
//
// `visualize` draws, with immediate-mode OpenGL through the scene camera: the stored photons of the global and the caustic
// map (DrawStoredGlobal / DrawStoredCaustic, visualize.cpp:1366-1424: a sphere of radius 0.01 in the photon's decoded
// power colour + a span along its incident direction), a 20 x 20 grid of eye rays with their hit points and normals
// (DrawRays, :268-298), and the Monte Carlo paths of a 25 x 25 grid of eye rays that meet a specular or transparent
// surface (DrawPaths + the drawing MonteCarlo_PathTrace, :436-530, :589-700).  There is no graphics stack on this
// machine and none is needed: the same pictures are splatted here by compute kernels into a z-buffered frame --
//   shapes_kernel       thread per pixel: eye ray -> closest hit -> head-light grey shade (the GL shape pass, optional),
//   photon_view_kernel  thread per photon: projects the sphere to a disc and the span to a line,
//   ray_view_kernel     thread per grid ray: closest hit, disc + normal span,
//   path_view_kernel    thread per grid ray: the specular / refractive walk (<= 10 bounces), a line per segment,
//   unpack_kernel       z-buffer word -> rgb bytes (row 0 = bottom row, like gi_render).
// A pixel is one 64-bit word {float depth bits : 32 | 0 : 8 | r g b : 24}; nearer fragments win by atomicMin (positive
// float bits order like integers).  Projection is the inverse of the eye-ray generator of render.cu (render.cpp:64-117).
#include <cstring>
#include <vector>
#include "context.h"
#include "shade.cuh"

namespace gi {

struct ViewCam {
  D3 eye, towards, up, right;
  double tan_x, tan_y;
  int w, h;
};

GI_DEV unsigned long long pack_frag(float depth, D3 c) {
  const unsigned R = (unsigned)(fmin(fmax(c.x, 0.0), 1.0) * 255.0), G = (unsigned)(fmin(fmax(c.y, 0.0), 1.0) * 255.0),
                 B = (unsigned)(fmin(fmax(c.z, 0.0), 1.0) * 255.0);
  return ((unsigned long long)__float_as_uint(depth) << 32) | (unsigned long long)(R | (G << 8) | (B << 16));
}

// world point -> continuous pixel coordinates (sx, sy) and depth along the view direction; false behind the near plane
GI_DEV bool project(const ViewCam &cam, D3 p, double &sx, double &sy, double &depth) {
  const D3 v = p - cam.eye;
  depth = dot(v, cam.towards);
  if (depth < 1.0e-4) return false;
  const double x = dot(v, cam.right) / (depth * cam.tan_x), y = dot(v, cam.up) / (depth * cam.tan_y);
  sx = (x * cam.w + 2 * (cam.w / 2)) * 0.5;             // inverse of dx = 2 (i - W/2) / W
  sy = (y * cam.h + 2 * (cam.h / 2)) * 0.5;
  return true;
}

GI_DEV void plot(unsigned long long *zbuf, const ViewCam &cam, int x, int y, float depth, D3 c) {
  if (x < 0 || y < 0 || x >= cam.w || y >= cam.h) return;
  atomicMin(zbuf + (size_t)y * cam.w + x, pack_frag(depth, c));
}

// R3Sphere(p, radius).Draw(): a filled disc of the projected radius (at least one pixel)
GI_DEV void splat_sphere(unsigned long long *zbuf, const ViewCam &cam, D3 p, double radius, D3 c) {
  double sx, sy, depth;
  if (!project(cam, p, sx, sy, depth)) return;
  const double pr = radius / (depth * cam.tan_x) * cam.w * 0.5;
  const int r = (int)fmin(fmax(pr, 0.0), 64.0);
  const int cx = (int)floor(sx), cy = (int)floor(sy);
  if (cx + r < 0 || cy + r < 0 || cx - r >= cam.w || cy - r >= cam.h) return;
  const float z = (float)fmax(depth - radius, 1.0e-4);
  for (int dy = -r; dy <= r; dy++)
    for (int dx = -r; dx <= r; dx++)
      if (dx * dx + dy * dy <= r * r) plot(zbuf, cam, cx + dx, cy + dy, z, c);
}

// R3Span(a, b).Draw(): a one-pixel line, depth interpolated along it; clipped against the near plane
GI_DEV void splat_span(unsigned long long *zbuf, const ViewCam &cam, D3 a, D3 b, D3 c) {
  const double near_d = 1.0e-3;
  double da = dot(a - cam.eye, cam.towards), db = dot(b - cam.eye, cam.towards);
  if (da < near_d && db < near_d) return;
  if (da < near_d) { const double t = (near_d - da) / (db - da); a = a + (b - a) * t; }
  else if (db < near_d) { const double t = (near_d - db) / (da - db); b = b + (a - b) * t; }
  double ax, ay, bx, by;
  if (!project(cam, a, ax, ay, da) || !project(cam, b, bx, by, db)) return;
  // keep the rasterised part inside a guard band around the frame
  const double span = fmax(fabs(bx - ax), fabs(by - ay));
  int steps = (int)fmin(span, 16384.0) + 1;
  for (int i = 0; i <= steps; i++) {
    const double t = (double)i / steps;
    const double x = ax + (bx - ax) * t, y = ay + (by - ay) * t;
    if (x < 0 || y < 0 || x >= cam.w || y >= cam.h) continue;
    plot(zbuf, cam, (int)floor(x), (int)floor(y), (float)(da + (db - da) * t), c);
  }
}

GI_DEV RayD eye_ray(const ViewCam &cam, double i, double j) {            // render.cpp:64-117 without depth of field
  const double dx = (2.0 * (i - (cam.w / 2))) / cam.w, dy = (2.0 * (j - (cam.h / 2))) / cam.h;
  const D3 far_point = cam.eye + cam.towards * 100.0 + cam.right * (cam.tan_x * 100.0 * dx) + cam.up * (cam.tan_y * 100.0 * dy);
  RayD r; r.o = cam.eye; r.d = normalized(far_point - cam.eye);
  return r;
}

__global__ void clear_kernel(unsigned long long *zbuf, size_t n, unsigned long long v) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) zbuf[t] = v;
}

// the shape pass: GL draws the scene's surfaces lit by the viewer's head light; here a grey N.V shade
__global__ void __launch_bounds__(128) shapes_kernel(DevScene sc, ViewCam cam, unsigned long long *zbuf) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= cam.w) return;
  const RayD ray = eye_ray(cam, x, y);
  HitRec h;
  closest_hit(sc, ray, h);
  if (h.prim < 0) return;
  const double s = 0.15 + 0.45 * fabs(dot(h.n, ray.d));
  plot(zbuf, cam, x, y, (float)dot(h.p - cam.eye, cam.towards), mk3(s, s, s));
}

// DrawStoredGlobal / DrawStoredCaustic (visualize.cpp:1366-1424)
__global__ void __launch_bounds__(128) photon_view_kernel(const GiPhoton *__restrict__ ph, uint64_t n, const double *__restrict__ lut,
                                                          double span_len, ViewCam cam, unsigned long long *zbuf) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const GiPhoton p = ph[t];
  const D3 pos = ld3(p.position);
  D3 c = rgbe_decode(p.rgbe);
  // photon powers are tiny (total power / emitted count): the view shows their hue at full value, as glColor3d clamps
  const double mx = max_channel(c);
  if (mx > 0) c = c / mx;
  const D3 incident = ld3(lut + 3 * (size_t)p.direction);
  splat_sphere(zbuf, cam, pos, 0.01, c);
  splat_span(zbuf, cam, pos, pos + incident * span_len, c);
}

// DrawRays (visualize.cpp:268-298): hit point and normal of a grid of eye rays
__global__ void __launch_bounds__(64) ray_view_kernel(DevScene sc, ViewCam cam, int grid, double radius, unsigned long long *zbuf) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= grid * grid) return;
  int istep = cam.w / grid, jstep = cam.h / grid;
  if (istep == 0) istep = 1;
  if (jstep == 0) jstep = 1;
  const int i = istep / 2 + (t / grid) * istep, j = jstep / 2 + (t % grid) * jstep;
  if (i >= cam.w || j >= cam.h) return;
  HitRec h;
  closest_hit(sc, eye_ray(cam, i, j), h);
  if (h.prim < 0) return;
  const D3 white = mk3(1, 1, 1);
  splat_sphere(zbuf, cam, h.p, radius, white);
  splat_span(zbuf, cam, h.p, h.p + h.n * (2.0 * radius), white);
}

// DrawPaths + the drawing MonteCarlo_PathTrace (visualize.cpp:436-530, 589-700): eye rays that meet a specular or
// transparent surface are followed through their specular / refractive bounces (roulette as in montecarlo.cpp:16-171,
// diffuse picks end the walk), one coloured span per segment, a white sphere per vertex.
__global__ void __launch_bounds__(64) path_view_kernel(ShadeEnv env, ViewCam cam, int grid, unsigned long long *zbuf) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= grid * grid) return;
  int istep = cam.w / grid, jstep = cam.h / grid;
  if (istep == 0) istep = 1;
  if (jstep == 0) jstep = 1;
  const int i = (t / grid) * istep, j = (t % grid) * jstep;
  if (i >= cam.w || j >= cam.h) return;
  const D3 colour = mk3((double)i / cam.w, ((double)j * i) / ((double)cam.h * cam.w), (double)j / cam.h);
  const D3 white = mk3(1, 1, 1);
  RayD ray = eye_ray(cam, i, j);
  Rng rng; rng_init(rng, env.p.seed, (uint64_t)t, 30, 0);
  for (int iter = 0; iter < 11; iter++) {
    HitRec h;
    closest_hit(env.sc, ray, h);
    if (h.prim < 0) break;
    const Brdf b = load_brdf(env.materials[env.sc.elements[h.element].material]);
    if (iter == 0 && !b.transparent && !b.specular) break;
    splat_sphere(zbuf, cam, h.p, 0.01, white);
    splat_span(zbuf, cam, ray.o, h.p, colour);
    const D3 view = normalized(h.p - ray.o);
    const double cos_theta = dot(h.n, -view);
    double r_coeff = 0;
    if (env.p.fresnel && b.transparent) r_coeff = reflection_coeff(cos_theta, b.ir, env.p.ir_air);
    const double prob_diffuse = max_channel(b.kd);
    double prob_transmission = max_channel(b.kt);
    const double prob_specular = max_channel(b.ks) + r_coeff * prob_transmission;
    prob_transmission *= (1.0 - r_coeff);
    const double prob_total = prob_diffuse + prob_transmission + prob_specular + max_channel(b.ke) + env.p.prob_absorb;
    double rnd = rng_double(rng);
    if (prob_total > 1.0) rnd *= prob_total;
    D3 bounce;
    if (rnd < prob_diffuse) break;
    else if (rnd < prob_diffuse + prob_transmission) bounce = transmissive_bounce(h.n, view, cos_theta, b.ir, env.p.ir_air);
    else if (rnd < prob_diffuse + prob_transmission + prob_specular) bounce = reflective_bounce(h.n, view, cos_theta);
    else break;
    ray.o = h.p + bounce * GI_EPS; ray.d = bounce;
  }
}

__global__ void unpack_kernel(const unsigned long long *__restrict__ zbuf, int w, int h, D3 background, uint8_t *rgb) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= w) return;
  const unsigned long long v = zbuf[(size_t)y * w + x];
  uint8_t *o = rgb + 3 * ((size_t)y * w + x);
  if ((v >> 32) == 0x7f800000ull) {
    o[0] = (uint8_t)(fmin(fmax(background.x, 0.0), 1.0) * 255.0); o[1] = (uint8_t)(fmin(fmax(background.y, 0.0), 1.0) * 255.0);
    o[2] = (uint8_t)(fmin(fmax(background.z, 0.0), 1.0) * 255.0);
  } else { o[0] = (uint8_t)(v & 0xff); o[1] = (uint8_t)((v >> 8) & 0xff); o[2] = (uint8_t)((v >> 16) & 0xff); }
}

const double *direction_lut(GiContext *ctx);      // capi_photon.cu

}  // namespace gi

using namespace gi;

extern "C" int gi_debug_view(GiContext *ctx, int kind, const GiParams *params, int width, int height, int draw_shapes, uint8_t *rgb_out) {
  if (!ctx || !rgb_out) { set_error("gi_debug_view: null argument"); return 1; }
  if (!ctx->has_scene) { set_error("gi_debug_view: no scene uploaded"); return 1; }
  if (kind < GI_VIEW_GLOBAL_PHOTONS || kind > GI_VIEW_PATHS) { set_error("gi_debug_view: unknown view"); return 1; }
  if (width < 1 || height < 1 || (uint64_t)width * height > (1ull << 28)) { set_error("gi_debug_view: bad image size"); return 1; }
  GI_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  GiParams prm; if (params) prm = *params; else gi_default_params(&prm);
  ViewCam cam;
  const GiCamera &c = ctx->camera;
  cam.eye = ld3(c.eye); cam.towards = ld3(c.towards); cam.up = ld3(c.up); cam.right = ld3(c.right);
  cam.tan_x = tan(c.xfov);
  cam.tan_y = ((double)height / (double)width) * tan(c.xfov);      // SetViewport re-derives yfov from the aspect ratio (R3Viewer.cpp:76-84)
  cam.w = width; cam.h = height;
  const size_t npix = (size_t)width * height;
  unsigned long long *zbuf = nullptr; uint8_t *d_rgb = nullptr;
  GI_CUDA(cudaMallocAsync((void **)&zbuf, 8 * npix, st));
  int rc = 1;
  do {
    if (!cuda_ok(cudaMallocAsync((void **)&d_rgb, 3 * npix, st), "cudaMallocAsync(view)")) break;
    clear_kernel<<<(unsigned)((npix + 255) / 256), 256, 0, st>>>(zbuf, npix, 0x7f800000ull << 32);
    ctx->launches++;
    if (draw_shapes) {
      shapes_kernel<<<dim3((width + 127) / 128, height), 128, 0, st>>>(ctx->scene, cam, zbuf);
      ctx->launches++;
    }
    if (kind == GI_VIEW_GLOBAL_PHOTONS || kind == GI_VIEW_CAUSTIC_PHOTONS) {
      const PhotonMapStore &m = ctx->maps[kind == GI_VIEW_GLOBAL_PHOTONS ? GI_MAP_GLOBAL : GI_MAP_CAUSTIC];
      if (m.dev.n > 0) {
        const double *lut = direction_lut(ctx);
        if (!lut) break;
        const double radius = 0.01;
        // DrawStoredGlobal: -2 radius along the incident direction; DrawStoredCaustic: +10 radius (visualize.cpp:1372,1393,1406,1421)
        const double span = kind == GI_VIEW_GLOBAL_PHOTONS ? -2.0 * radius : 10.0 * radius;
        photon_view_kernel<<<(unsigned)((m.dev.n + 127) / 128), 128, 0, st>>>((const GiPhoton *)m.raw, m.dev.n, lut, span, cam, zbuf);
        ctx->launches++;
      }
    } else if (kind == GI_VIEW_RAYS) {
      ray_view_kernel<<<(20 * 20 + 63) / 64, 64, 0, st>>>(ctx->scene, cam, 20, 0.025 * ctx->radius, zbuf);
      ctx->launches++;
    } else {
      ShadeEnv env;
      env.sc = ctx->scene; env.materials = ctx->scene.materials; env.p = prm;
      env.ambient = ld3(ctx->ambient); env.background = ld3(ctx->background); env.eye = cam.eye;
      env.radius = ctx->radius; env.spec_boxes = nullptr; env.n_spec_boxes = -1; env.lights = nullptr; env.n_lights = 0;
      path_view_kernel<<<(25 * 25 + 63) / 64, 64, 0, st>>>(env, cam, 25, zbuf);
      ctx->launches++;
    }
    unpack_kernel<<<dim3((width + 127) / 128, height), 128, 0, st>>>(zbuf, width, height, draw_shapes ? ld3(ctx->background) : mk3(0, 0, 0), d_rgb);
    ctx->launches++;
    if (!cuda_ok(cudaGetLastError(), "debug view kernels")) break;
    if (!cuda_ok(cudaMemcpyAsync(rgb_out, d_rgb, 3 * npix, cudaMemcpyDeviceToHost, st), "D2H view")) break;
    if (!cuda_ok(cudaStreamSynchronize(st), "debug view sync")) break;
    rc = 0;
  } while (0);
  if (zbuf) cudaFreeAsync(zbuf, st);
  if (d_rgb) cudaFreeAsync(d_rgb, st);
  return rc;
}
