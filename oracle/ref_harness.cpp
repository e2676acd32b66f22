// oracle/ref_harness.cpp -- TEST INFRASTRUCTURE (never linked into, loaded by, or timed as the product).
//
// Harness around the UNMODIFIED reference sources, compiled where they lie under /root/reference/src
// (see oracle/Makefile).  It reaches the reference's file-static MapPhotons() and globals by
// #including src/photonmap.cpp with `main` renamed, and exposes seeded, batched entry points whose
// binary outputs become the golden fixtures under tests/golden/ and the "reference" CPU baseline:
//
//   ref_harness dump_scene <in.scn> <out.txt> [-real]
//   ref_harness camrays    <in.scn> <W> <H> <rays.bin>
//   ref_harness isect      <in.scn> <rays.bin> <hits.bin> [threads]      (R3Scene::Intersects, R3Scene.cpp:471)
//   ref_harness photons    <in.scn> <out_prefix> <seed> [photonmap flags] (MapPhotons, photonmap.cpp:260)
//   ref_harness gather     <photons.bin> <queries.bin> <out.bin> <k> <r> <filter 0|1|2> [threads]
//                                                     (EstimateRadiance, photon_utils.cpp:72)
//   ref_harness cache      <photons.bin> <queries.bin> <out.bin> <k> <r> <out_photons.bin>
//                                                     (irradiance pre-pass photonmap.cpp:380-411 + EstimateCachedRadiance)
//   ref_harness parse      [photonmap flags]                               (ParseArgs, io_utils.cpp:16)
//   ref_harness shade      <in.scn> <rays.bin> <out.bin> [-load_global f] [-load_caustic f] [photonmap flags]
//                                                     (Intersects + RayTrace, render.cpp:119-124)
// With -DGI_REF_MAIN the same translation unit is just the reference program (`photonmap_ref`).
//
// The only deviation from the reference text: `offsetof` is redefined to yield an int, because
// `new R3Kdtree<Photon*>(array, offsetof(...))` (photonmap.cpp:365,372) is ambiguous under g++
// (int offset vs function-pointer overload, R3Kdtree.h:35-36).  Same value, same overload the
// author's compiler picked.

#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdint>
#include <cmath>
#include <map>
#include <vector>
#include <string>
#include <thread>
#include <algorithm>
#include <chrono>
#include "R3Graphics/R3Graphics.h"
#undef offsetof
#define offsetof(t, m) ((int)__builtin_offsetof(t, m))
#define main ref_main
#include "photonmap.cpp"
#undef main
#include "raytracer.h"

#ifdef GI_REF_MAIN
int main(int argc, char **argv) { return ref_main(argc, argv); }
#else

// ---------------------------------------------------------------- file formats (shared with tests/)
static const uint32_t RAYS_MAGIC = 0x59415247;  // "GRAY": u32 magic, u32 n, n x {f64 o[3], f64 d[3]}
static const uint32_t HITS_MAGIC = 0x54494847;  // "GHIT": u32 magic, u32 n, n x HitRec
static const uint32_t PMAP_MAGIC = 0x4d504947;  // "GIPM": u32 magic, u32 version, u64 n, n x 32-byte Photon
struct HitRec { int32_t hit; int32_t shape_id; double t; double p[3]; double n[3]; };
static_assert(sizeof(HitRec) == 64, "HitRec layout");
static_assert(sizeof(Photon) == 32, "reference Photon layout (render.h:17-21)");
struct QueryRec { double p[3], n[3], bounce[3], cos_theta, kd[3], ks[3], shininess; };  // 17 doubles
static_assert(sizeof(QueryRec) == 136, "QueryRec layout");

static void die(const char *msg) { fprintf(stderr, "ref_harness: %s\n", msg); exit(2); }

static std::vector<double> read_rays(const char *path, uint32_t &n) {
  FILE *f = fopen(path, "rb"); if (!f) die("cannot open rays file");
  uint32_t hdr[2]; if (fread(hdr, 4, 2, f) != 2 || hdr[0] != RAYS_MAGIC) die("bad rays file");
  n = hdr[1]; std::vector<double> r((size_t)n * 6);
  if (fread(r.data(), 48, n, f) != n) die("short rays file");
  fclose(f); return r;
}

// ---------------------------------------------------------------- scene walk (reference DFS order)
// Order = order in which R3SceneNode::Intersects visits shapes: a node's elements (insertion order),
// each element's shapes (insertion order), then the node's children (R3SceneNode.cpp:461-490).
struct Walk {
  std::map<const R3Shape *, int> shape_id;
  std::vector<const R3Shape *> shapes;
  std::vector<int> shape_node, shape_elem;
  std::vector<const R3SceneNode *> nodes;
  std::vector<int> node_parent;
  std::vector<const R3SceneElement *> elems;
  void visit(const R3SceneNode *node, int parent) {
    int me = (int)nodes.size(); nodes.push_back(node); node_parent.push_back(parent);
    for (int i = 0; i < node->NElements(); i++) {
      const R3SceneElement *e = node->Element(i);
      int eid = (int)elems.size(); elems.push_back(e);
      for (int j = 0; j < e->NShapes(); j++) {
        const R3Shape *s = e->Shape(j);
        shape_id[s] = (int)shapes.size(); shapes.push_back(s); shape_node.push_back(me); shape_elem.push_back(eid);
      }
    }
    for (int i = 0; i < node->NChildren(); i++) visit(node->Child(i), me);
  }
};

static R3Scene *load_scene(const char *path, bool real) {
  R3Scene *s = ReadScene((char *)path, real);
  if (!s) die("cannot read scene");
  SCENE = s;
  SCENE_RADIUS = s->BBox().DiagonalRadius();
  SCENE_AMBIENT = s->Ambient();
  SCENE_NLIGHTS = s->NLights();
  return s;
}

// ---------------------------------------------------------------- dump_scene
static void p3(FILE *f, const double *v) { fprintf(f, " %.17g %.17g %.17g", v[0], v[1], v[2]); }
static void prgb(FILE *f, const RNRgb &c) { fprintf(f, " %.17g %.17g %.17g", c.R(), c.G(), c.B()); }
static void ppt(FILE *f, const R3Point &p) { fprintf(f, " %.17g %.17g %.17g", p.X(), p.Y(), p.Z()); }
static void pvec(FILE *f, const R3Vector &p) { fprintf(f, " %.17g %.17g %.17g", p.X(), p.Y(), p.Z()); }

static int cmd_dump_scene(int argc, char **argv) {
  if (argc < 4) die("usage: dump_scene in.scn out.txt [-real]");
  bool real = (argc > 4 && !strcmp(argv[4], "-real"));
  R3Scene *scene = load_scene(argv[2], real);
  Walk w; w.visit(scene->Root(), -1);
  FILE *f = fopen(argv[3], "w"); if (!f) die("cannot open output");
  // materials are identified by their brdf values per element (pointer identity is not portable)
  fprintf(f, "nodes %d\n", (int)w.nodes.size());
  for (size_t i = 0; i < w.nodes.size(); i++) {
    const R4Matrix &m = w.nodes[i]->Transformation().Matrix();
    const R4Matrix &mi = w.nodes[i]->Transformation().InverseMatrix();
    fprintf(f, "node %d parent %d\n m", (int)i, w.node_parent[i]);
    for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) fprintf(f, " %.17g", m[r][c]);
    fprintf(f, "\n minv");
    for (int r = 0; r < 4; r++) for (int c = 0; c < 4; c++) fprintf(f, " %.17g", mi[r][c]);
    fprintf(f, "\n");
  }
  fprintf(f, "elements %d\n", (int)w.elems.size());
  for (size_t i = 0; i < w.elems.size(); i++) {
    const R3Material *mat = w.elems[i]->Material();
    const R3Brdf *b = mat ? mat->Brdf() : &R3default_brdf;
    if (!b) b = &R3default_brdf;
    fprintf(f, "element %d\n ka", (int)i); prgb(f, b->Ambient());
    fprintf(f, "\n kd"); prgb(f, b->Diffuse());
    fprintf(f, "\n ks"); prgb(f, b->Specular());
    fprintf(f, "\n kt"); prgb(f, b->Transmission());
    fprintf(f, "\n ke"); prgb(f, b->Emission());
    fprintf(f, "\n n %.17g ir %.17g\n", b->Shininess(), b->IndexOfRefraction());
  }
  fprintf(f, "shapes %d\n", (int)w.shapes.size());
  for (size_t i = 0; i < w.shapes.size(); i++) {
    const R3Shape *s = w.shapes[i];
    fprintf(f, "shape %d node %d element %d ", (int)i, w.shape_node[i], w.shape_elem[i]);
    if (s->ClassID() == R3Sphere::CLASS_ID()) {
      const R3Sphere *q = (const R3Sphere *)s;
      fprintf(f, "sphere"); ppt(f, q->Center()); fprintf(f, " %.17g\n", q->Radius());
    } else if (s->ClassID() == R3Box::CLASS_ID()) {
      const R3Box *q = (const R3Box *)s;
      fprintf(f, "box"); ppt(f, q->Min()); ppt(f, q->Max()); fprintf(f, "\n");
    } else if (s->ClassID() == R3Cylinder::CLASS_ID()) {
      const R3Cylinder *q = (const R3Cylinder *)s;
      fprintf(f, "cylinder"); ppt(f, q->Axis().Start()); ppt(f, q->Axis().End()); fprintf(f, " %.17g\n", q->Radius());
    } else if (s->ClassID() == R3Cone::CLASS_ID()) {
      const R3Cone *q = (const R3Cone *)s;
      fprintf(f, "cone"); ppt(f, q->Axis().Start()); ppt(f, q->Axis().End()); fprintf(f, " %.17g\n", q->Radius());
    } else if (s->ClassID() == R3Circle::CLASS_ID()) {
      const R3Circle *q = (const R3Circle *)s;
      fprintf(f, "circle"); ppt(f, q->Center()); pvec(f, q->Normal()); fprintf(f, " %.17g\n", q->Radius());
    } else if (s->ClassID() == R3Triangle::CLASS_ID()) {
      const R3Triangle *q = (const R3Triangle *)s;
      fprintf(f, "tri");
      for (int k = 0; k < 3; k++) ppt(f, q->Vertex(k)->Position());
      fprintf(f, "\n");
    } else if (s->ClassID() == R3TriangleArray::CLASS_ID()) {
      const R3TriangleArray *q = (const R3TriangleArray *)s;
      fprintf(f, "mesh %d\n", q->NTriangles());
      for (int t = 0; t < q->NTriangles(); t++) {
        fprintf(f, " t");
        for (int k = 0; k < 3; k++) ppt(f, q->Triangle(t)->Vertex(k)->Position());
        fprintf(f, "\n");
      }
    } else {
      fprintf(f, "unknown %d\n", (int)s->ClassID());
    }
  }
  fprintf(f, "lights %d\n", scene->NLights());
  for (int i = 0; i < scene->NLights(); i++) {
    R3Light *l = scene->Light(i);
    fprintf(f, "light %d ", i);
    if (l->ClassID() == R3DirectionalLight::CLASS_ID()) {
      R3DirectionalLight *q = (R3DirectionalLight *)l;
      fprintf(f, "dir"); prgb(f, q->Color()); fprintf(f, " %.17g", q->Intensity()); pvec(f, q->Direction());
    } else if (l->ClassID() == R3PointLight::CLASS_ID()) {
      R3PointLight *q = (R3PointLight *)l;
      fprintf(f, "point"); prgb(f, q->Color()); fprintf(f, " %.17g", q->Intensity()); ppt(f, q->Position());
      fprintf(f, " %.17g %.17g %.17g", q->ConstantAttenuation(), q->LinearAttenuation(), q->QuadraticAttenuation());
    } else if (l->ClassID() == R3SpotLight::CLASS_ID()) {
      R3SpotLight *q = (R3SpotLight *)l;
      fprintf(f, "spot"); prgb(f, q->Color()); fprintf(f, " %.17g", q->Intensity()); ppt(f, q->Position());
      pvec(f, q->Direction());
      fprintf(f, " %.17g %.17g %.17g", q->ConstantAttenuation(), q->LinearAttenuation(), q->QuadraticAttenuation());
      fprintf(f, " %.17g %.17g", q->DropOffRate(), q->CutOffAngle());
    } else if (l->ClassID() == R3AreaLight::CLASS_ID()) {
      R3AreaLight *q = (R3AreaLight *)l;
      fprintf(f, "area"); prgb(f, q->Color()); fprintf(f, " %.17g", q->Intensity()); ppt(f, q->Position());
      pvec(f, q->Direction()); fprintf(f, " %.17g", q->Radius());
      fprintf(f, " %.17g %.17g %.17g", q->ConstantAttenuation(), q->LinearAttenuation(), q->QuadraticAttenuation());
    } else if (l->ClassID() == R3RectLight::CLASS_ID()) {
      R3RectLight *q = (R3RectLight *)l;
      fprintf(f, "rect"); prgb(f, q->Color()); fprintf(f, " %.17g", q->Intensity()); ppt(f, q->Position());
      pvec(f, q->Direction()); pvec(f, q->PrimaryAxis()); pvec(f, q->SecondaryAxis());
      fprintf(f, " %.17g %.17g", q->PrimaryLength(), q->SecondaryLength());
      fprintf(f, " %.17g %.17g %.17g", q->ConstantAttenuation(), q->LinearAttenuation(), q->QuadraticAttenuation());
    } else fprintf(f, "unknown");
    fprintf(f, "\n");
  }
  const R3Camera &c = scene->Camera();
  fprintf(f, "camera"); ppt(f, c.Origin()); pvec(f, c.Towards()); pvec(f, c.Up()); pvec(f, c.Right());
  fprintf(f, " %.17g %.17g %.17g %.17g\n", c.XFOV(), c.YFOV(), c.Near(), c.Far());
  fprintf(f, "ambient"); prgb(f, scene->Ambient()); fprintf(f, "\nbackground"); prgb(f, scene->Background());
  fprintf(f, "\nbbox"); ppt(f, scene->BBox().Min()); ppt(f, scene->BBox().Max());
  fprintf(f, "\ncentroid"); ppt(f, scene->Centroid());
  fprintf(f, "\nradius %.17g\n", SCENE_RADIUS);
  fclose(f);
  return 0;
}

// ---------------------------------------------------------------- camrays (render.cpp:64-117, no DOF)
static int cmd_camrays(int argc, char **argv) {
  if (argc < 6) die("usage: camrays in.scn W H rays.bin");
  R3Scene *scene = load_scene(argv[2], false);
  int W = atoi(argv[3]), H = atoi(argv[4]);
  scene->SetViewport(R2Viewport(0, 0, W, H));
  R3Camera camera = scene->Camera();
  R2Viewport viewport = scene->Viewport();
  R3Point far_org = camera.Origin() + camera.Towards() * FOCUS_DEPTH;
  R3Vector far_right = camera.Right() * tan(camera.XFOV()) * FOCUS_DEPTH;
  R3Vector far_up = camera.Up() * tan(camera.YFOV()) * FOCUS_DEPTH;
  FILE *f = fopen(argv[5], "wb"); if (!f) die("cannot open output");
  uint32_t hdr[2] = {RAYS_MAGIC, (uint32_t)(W * H)};
  fwrite(hdr, 4, 2, f);
  for (int i = 0; i < W; i++) for (int j = 0; j < H; j++) {   // column-major like image_buffer[i][j]
    RNScalar dx = (RNScalar)(2 * (i - viewport.XCenter())) / (RNScalar)viewport.Width();
    RNScalar dy = (RNScalar)(2 * (j - viewport.YCenter())) / (RNScalar)viewport.Height();
    R3Point far_point = far_org + (far_right * dx) + (far_up * dy);
    R3Ray ray(camera.Origin(), far_point);
    double r[6] = {ray.Start().X(), ray.Start().Y(), ray.Start().Z(), ray.Vector().X(), ray.Vector().Y(), ray.Vector().Z()};
    fwrite(r, 8, 6, f);
  }
  fclose(f);
  return 0;
}

// ---------------------------------------------------------------- isect
static int cmd_isect(int argc, char **argv) {
  if (argc < 5) die("usage: isect in.scn rays.bin hits.bin [threads]");
  R3Scene *scene = load_scene(argv[2], false);
  Walk w; w.visit(scene->Root(), -1);
  uint32_t n; std::vector<double> rays = read_rays(argv[3], n);
  int threads = argc > 5 ? atoi(argv[5]) : 1; if (threads < 1) threads = 1;
  std::vector<HitRec> hits(n);
  auto t0 = std::chrono::steady_clock::now();
  auto work = [&](int id) {
    for (uint32_t i = id; i < n; i += threads) {
      const double *r = &rays[(size_t)i * 6];
      R3Ray ray(R3Point(r[0], r[1], r[2]), R3Vector(r[3], r[4], r[5]), TRUE);
      R3SceneNode *node; R3SceneElement *element; R3Shape *shape; R3Point point; R3Vector normal; RNScalar t;
      HitRec h; memset(&h, 0, sizeof(h)); h.shape_id = -1;
      if (scene->Intersects(ray, &node, &element, &shape, &point, &normal, &t)) {
        h.hit = 1; h.shape_id = w.shape_id[shape]; h.t = t;
        h.p[0] = point.X(); h.p[1] = point.Y(); h.p[2] = point.Z();
        h.n[0] = normal.X(); h.n[1] = normal.Y(); h.n[2] = normal.Z();
      }
      hits[i] = h;
    }
  };
  std::vector<std::thread> th;
  for (int i = 1; i < threads; i++) th.emplace_back(work, i);
  work(0);
  for (auto &x : th) x.join();
  double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  FILE *f = fopen(argv[4], "wb"); if (!f) die("cannot open output");
  uint32_t hdr[2] = {HITS_MAGIC, n};
  fwrite(hdr, 4, 2, f); fwrite(hits.data(), sizeof(HitRec), n, f); fclose(f);
  printf("{\"cmd\": \"isect\", \"rays\": %u, \"threads\": %d, \"seconds\": %.6f, \"mrays_per_s\": %.6f}\n",
         n, threads, sec, n / sec / 1e6);
  return 0;
}

// ---------------------------------------------------------------- photon files
static std::vector<Photon> read_photons(const char *path) {
  FILE *f = fopen(path, "rb"); if (!f) die("cannot open photon file");
  uint32_t hdr[2]; uint64_t n;
  if (fread(hdr, 4, 2, f) != 2 || hdr[0] != PMAP_MAGIC || fread(&n, 8, 1, f) != 1) die("bad photon file");
  std::vector<Photon> ph(n);
  if (n && fread(ph.data(), sizeof(Photon), n, f) != n) die("short photon file");
  fclose(f); return ph;
}
static void write_photons(const std::string &path, RNArray<Photon *> &arr) {
  FILE *f = fopen(path.c_str(), "wb"); if (!f) die("cannot open photon output");
  uint32_t hdr[2] = {PMAP_MAGIC, 1}; uint64_t n = arr.NEntries();
  fwrite(hdr, 4, 2, f); fwrite(&n, 8, 1, f);
  for (uint64_t i = 0; i < n; i++) {
    Photon p; memset(&p, 0, sizeof(p)); p.position = arr[i]->position; memcpy(p.rgbe, arr[i]->rgbe, 4);
    p.direction = arr[i]->direction;
    fwrite(&p, sizeof(Photon), 1, f);
  }
  fclose(f);
}

// ---------------------------------------------------------------- photons (MapPhotons, photonmap.cpp:260-436)
static int cmd_photons(int argc, char **argv) {
  if (argc < 4) die("usage: photons in.scn out_prefix [-global N] [-caustic N] [-threads N] [-pd N]");
  for (int i = 4; i + 1 < argc; i += 2) {
    if (!strcmp(argv[i], "-fast_global")) { FAST_GLOBAL = atoi(argv[i + 1]) != 0; if (FAST_GLOBAL) { DIRECT_PHOTON_ILLUM = true; INDIRECT_ILLUM = false; } }
    else if (!strcmp(argv[i], "-global")) GLOBAL_PHOTON_COUNT = atoi(argv[i + 1]);
    else if (!strcmp(argv[i], "-caustic")) CAUSTIC_PHOTON_COUNT = atoi(argv[i + 1]);
    else if (!strcmp(argv[i], "-threads")) THREADS = atoi(argv[i + 1]);
    else if (!strcmp(argv[i], "-pd")) MAX_PHOTON_DEPTH = atoi(argv[i + 1]);
    else die("unknown photons flag");
  }
  load_scene(argv[2], false);
  auto t0 = std::chrono::steady_clock::now();
  MapPhotons();
  double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  write_photons(std::string(argv[3]) + ".global.gipm", GLOBAL_PHOTONS);
  write_photons(std::string(argv[3]) + ".caustic.gipm", CAUSTIC_PHOTONS);
  printf("{\"cmd\": \"photons\", \"global\": %d, \"caustic\": %d, \"threads\": %d, \"seconds\": %.6f}\n",
         GLOBAL_PHOTONS.NEntries(), CAUSTIC_PHOTONS.NEntries(), THREADS, sec);
  return 0;
}

// ---------------------------------------------------------------- gather (EstimateRadiance, photon_utils.cpp:72-162)
// queries.bin: u32 magic "GQRY", u32 n, n x QueryRec.  out.bin: u32 magic "GGAT", u32 n, u32 k, u32 0,
// then n x {i32 count, f64 rgb[3] (packed: 4 + 24 = 28 -> stored as i32 count, i32 pad, 3 f64)}, then n*k i32 ascending
// photon indices (padded with -1) from FindClosestQuick (R3Kdtree.cpp:829-848) on the same tree.
static int cmd_gather(int argc, char **argv) {
  if (argc < 8) die("usage: gather photons.gipm queries.bin out.bin k r filter [threads]");
  std::vector<Photon> ph = read_photons(argv[2]);
  FILE *f = fopen(argv[3], "rb"); if (!f) die("cannot open queries");
  uint32_t hdr[2]; if (fread(hdr, 4, 2, f) != 2 || hdr[0] != 0x59525147) die("bad queries file");
  uint32_t n = hdr[1]; std::vector<QueryRec> q(n);
  if (n && fread(q.data(), sizeof(QueryRec), n, f) != n) die("short queries file");
  fclose(f);
  int k = atoi(argv[5]); double r = atof(argv[6]); int filter = atoi(argv[7]);
  int threads = argc > 8 ? atoi(argv[8]) : 1; if (threads < 1) threads = 1;
  RNArray<Photon *> arr;
  for (size_t i = 0; i < ph.size(); i++) arr.Insert(&ph[i]);
  auto tb = std::chrono::steady_clock::now();
  R3Kdtree<Photon *> *tree = ph.empty() ? NULL : new R3Kdtree<Photon *>(arr, offsetof(struct Photon, position));
  double build_sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - tb).count();
  BuildDirectionLookupTable();
  struct Out { int32_t count, pad; double rgb[3]; };
  std::vector<Out> out(n); std::vector<int32_t> idx((size_t)n * k, -1);
  auto t0 = std::chrono::steady_clock::now();
  auto work = [&](int id) {
    for (uint32_t i = id; i < n; i += threads) {
      const QueryRec &qq = q[i];
      R3Point point(qq.p[0], qq.p[1], qq.p[2]); R3Vector normal(qq.n[0], qq.n[1], qq.n[2]);
      R3Vector bounce(qq.bounce[0], qq.bounce[1], qq.bounce[2]);
      R3Brdf brdf(RNblack_rgb, RNRgb(qq.kd[0], qq.kd[1], qq.kd[2]), RNRgb(qq.ks[0], qq.ks[1], qq.ks[2]), RNblack_rgb,
                  RNblack_rgb, qq.shininess, 1.0);
      RNRgb color = RNblack_rgb;
      Out o; memset(&o, 0, sizeof(o));
      if (tree) {
        EstimateRadiance(point, normal, color, &brdf, bounce, qq.cos_theta, tree, k, r, (Filter_Type)filter);
        std::vector<PointAndDistanceSqd<Photon *> > near;
        tree->FindClosestQuick(point, 0, r, k, near);
        o.count = (int32_t)near.size();
        std::vector<int32_t> ids;
        for (auto &e : near) ids.push_back((int32_t)(e.point - ph.data()));
        std::sort(ids.begin(), ids.end());
        for (size_t j = 0; j < ids.size(); j++) idx[(size_t)i * k + j] = ids[j];
      }
      o.rgb[0] = color.R(); o.rgb[1] = color.G(); o.rgb[2] = color.B();
      out[i] = o;
    }
  };
  std::vector<std::thread> th;
  for (int i = 1; i < threads; i++) th.emplace_back(work, i);
  work(0);
  for (auto &x : th) x.join();
  double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  f = fopen(argv[4], "wb"); if (!f) die("cannot open output");
  uint32_t oh[4] = {0x54414747, n, (uint32_t)k, 0};
  fwrite(oh, 4, 4, f); fwrite(out.data(), sizeof(Out), n, f); fwrite(idx.data(), 4, idx.size(), f); fclose(f);
  printf("{\"cmd\": \"gather\", \"photons\": %zu, \"queries\": %u, \"k\": %d, \"threads\": %d, \"build_seconds\": %.6f, "
         "\"seconds\": %.6f, \"gathers_per_s\": %.3f}\n", ph.size(), n, k, threads, build_sec, sec, 2.0 * n / sec);
  return 0;
}

// ---------------------------------------------------------------- irradiance cache (-cache)
// The cache pre-pass is inlined in MapPhotons (photonmap.cpp:380-411), which also traces the photons; for a LOADED map
// the same steps are restated here around the reference's own EstimateIrradiance (photon_utils.cpp:208-246): every
// photon's colour becomes its own power + the irradiance estimate at its position, all estimates taken before any
// colour is overwritten.
static void irradiance_prepass(RNArray<Photon *> &photons, R3Kdtree<Photon *> *tree, int estimate_size, RNScalar estimate_dist) {
  const int n = photons.NEntries();
  std::vector<Photon> converted(n);
  for (int i = 0; i < n; i++) {
    converted[i] = *photons[i];
    RNRgb irradiance = RGBE_to_RNRgb(converted[i].rgbe);
    EstimateIrradiance(converted[i].position, irradiance, tree, estimate_size, estimate_dist);
    RNRgb_to_RGBE(irradiance, converted[i].rgbe);
  }
  for (int i = 0; i < n; i++)
    for (int j = 0; j < 4; j++) photons[i]->rgbe[j] = converted[i].rgbe[j];
}

// usage: cache photons.gipm queries.bin out.bin k r out_photons.gipm
// Pre-pass on the loaded map (k = GLOBAL_ESTIMATE_SIZE, r = GLOBAL_ESTIMATE_DIST), converted photons written to
// out_photons.gipm; then the reference's EstimateCachedRadiance (photon_utils.cpp:164-205) for every query.
// out.bin has the `gather` layout with k = 1: count = 1 and the chosen photon's index, or count = -1 where no
// photon qualifies inside r -- the reference would dereference NULL there (photon_utils.cpp:181-183), so the call
// is skipped and the colour left black.
static int cmd_cache(int argc, char **argv) {
  if (argc < 8) die("usage: cache photons.gipm queries.bin out.bin k r out_photons.gipm");
  static std::vector<Photon> ph; ph = read_photons(argv[2]);
  FILE *f = fopen(argv[3], "rb"); if (!f) die("cannot open queries");
  uint32_t hdr[2]; if (fread(hdr, 4, 2, f) != 2 || hdr[0] != 0x59525147) die("bad queries file");
  uint32_t n = hdr[1]; std::vector<QueryRec> q(n);
  if (n && fread(q.data(), sizeof(QueryRec), n, f) != n) die("short queries file");
  fclose(f);
  int k = atoi(argv[5]); double r = atof(argv[6]);
  RNArray<Photon *> arr;
  for (size_t i = 0; i < ph.size(); i++) arr.Insert(&ph[i]);
  if (ph.empty()) die("empty photon map");
  R3Kdtree<Photon *> *tree = new R3Kdtree<Photon *>(arr, offsetof(struct Photon, position));
  BuildDirectionLookupTable();
  irradiance_prepass(arr, tree, k, r);
  write_photons(argv[7], arr);
  struct Out { int32_t count, pad; double rgb[3]; };
  std::vector<Out> out(n); std::vector<int32_t> idx(n, -1);
  for (uint32_t i = 0; i < n; i++) {
    const QueryRec &qq = q[i];
    R3Point point(qq.p[0], qq.p[1], qq.p[2]); R3Vector normal(qq.n[0], qq.n[1], qq.n[2]);
    R3Vector bounce(qq.bounce[0], qq.bounce[1], qq.bounce[2]);
    R3Brdf brdf(RNblack_rgb, RNRgb(qq.kd[0], qq.kd[1], qq.kd[2]), RNRgb(qq.ks[0], qq.ks[1], qq.ks[2]), RNblack_rgb,
                RNblack_rgb, qq.shininess, 1.0);
    // which photon will the reference's loop settle on, if any?
    Photon *p = NULL; RNLength d = 0;
    while (true) {
      p = tree->FindClosest(point, d + RN_EPSILON, r, &d);
      if (!p) break;
      const int dir = p->direction;
      const RNScalar perp = normal.Dot(R3Vector(PHOTON_X_LOOKUP[dir], PHOTON_Y_LOOKUP[dir], PHOTON_Z_LOOKUP[dir]));
      if (!((qq.cos_theta < 0 && perp < 0) || (qq.cos_theta > 0 && perp > 0))) break;
    }
    Out o; memset(&o, 0, sizeof(o));
    RNRgb color = RNblack_rgb;
    if (p) {
      EstimateCachedRadiance(point, normal, color, &brdf, bounce, qq.cos_theta, tree, r);
      o.count = 1; idx[i] = (int32_t)(p - ph.data());
    } else o.count = -1;
    o.rgb[0] = color.R(); o.rgb[1] = color.G(); o.rgb[2] = color.B();
    out[i] = o;
  }
  f = fopen(argv[4], "wb"); if (!f) die("cannot open output");
  uint32_t oh[4] = {0x54414747, n, 1u, 0};
  fwrite(oh, 4, 4, f); fwrite(out.data(), sizeof(Out), n, f); fwrite(idx.data(), 4, idx.size(), f); fclose(f);
  printf("{\"cmd\": \"cache\", \"photons\": %zu, \"queries\": %u, \"k\": %d}\n", ph.size(), n, k);
  return 0;
}

// ---------------------------------------------------------------- bench: the hot path on all host threads
// One "step" = R3Scene::Intersects over every ray of rays.bin + EstimateRadiance over every query of queries.bin
// (kd-tree built once, outside the timed steps).  Prints one JSON line with the per-step seconds.
static int cmd_bench(int argc, char **argv) {
  if (argc < 12) die("usage: bench in.scn rays.bin photons.gipm queries.bin k r filter threads steps warmup");
  R3Scene *scene = load_scene(argv[2], false);
  uint32_t nr; std::vector<double> rays = read_rays(argv[3], nr);
  std::vector<Photon> ph = read_photons(argv[4]);
  FILE *f = fopen(argv[5], "rb"); if (!f) die("cannot open queries");
  uint32_t hdr[2]; if (fread(hdr, 4, 2, f) != 2 || hdr[0] != 0x59525147) die("bad queries file");
  uint32_t nq = hdr[1]; std::vector<QueryRec> q(nq);
  if (nq && fread(q.data(), sizeof(QueryRec), nq, f) != nq) die("short queries file");
  fclose(f);
  int k = atoi(argv[6]); double r = atof(argv[7]); int filter = atoi(argv[8]);
  int threads = atoi(argv[9]), steps = atoi(argv[10]), warmup = atoi(argv[11]); if (threads < 1) threads = 1;
  RNArray<Photon *> arr;
  for (size_t i = 0; i < ph.size(); i++) arr.Insert(&ph[i]);
  auto tb = std::chrono::steady_clock::now();
  R3Kdtree<Photon *> *tree = ph.empty() ? NULL : new R3Kdtree<Photon *>(arr, offsetof(struct Photon, position));
  double build_sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - tb).count();
  BuildDirectionLookupTable();
  std::vector<double> sink(threads, 0.0), isect_s, gather_s;
  for (int step = 0; step < warmup + steps; step++) {
    auto t0 = std::chrono::steady_clock::now();
    {
      auto work = [&](int id) {
        double acc = 0;
        for (uint32_t i = id; i < nr; i += threads) {
          const double *rr = &rays[(size_t)i * 6];
          R3Ray ray(R3Point(rr[0], rr[1], rr[2]), R3Vector(rr[3], rr[4], rr[5]), TRUE);
          R3SceneNode *node; R3SceneElement *element; R3Shape *shape; R3Point point; R3Vector normal; RNScalar t;
          if (scene->Intersects(ray, &node, &element, &shape, &point, &normal, &t)) acc += t;
        }
        sink[id] += acc;
      };
      std::vector<std::thread> th;
      for (int i = 1; i < threads; i++) th.emplace_back(work, i);
      work(0);
      for (auto &x : th) x.join();
    }
    auto t1 = std::chrono::steady_clock::now();
    if (tree) {
      auto work = [&](int id) {
        double acc = 0;
        for (uint32_t i = id; i < nq; i += threads) {
          const QueryRec &qq = q[i];
          R3Point point(qq.p[0], qq.p[1], qq.p[2]); R3Vector normal(qq.n[0], qq.n[1], qq.n[2]);
          R3Vector bounce(qq.bounce[0], qq.bounce[1], qq.bounce[2]);
          R3Brdf brdf(RNblack_rgb, RNRgb(qq.kd[0], qq.kd[1], qq.kd[2]), RNRgb(qq.ks[0], qq.ks[1], qq.ks[2]), RNblack_rgb,
                      RNblack_rgb, qq.shininess, 1.0);
          RNRgb color = RNblack_rgb;
          EstimateRadiance(point, normal, color, &brdf, bounce, qq.cos_theta, tree, k, r, (Filter_Type)filter);
          acc += color.R();
        }
        sink[id] += acc;
      };
      std::vector<std::thread> th;
      for (int i = 1; i < threads; i++) th.emplace_back(work, i);
      work(0);
      for (auto &x : th) x.join();
    }
    auto t2 = std::chrono::steady_clock::now();
    if (step >= warmup) {
      isect_s.push_back(std::chrono::duration<double>(t1 - t0).count());
      gather_s.push_back(std::chrono::duration<double>(t2 - t1).count());
    }
  }
  double si = 0, sg = 0; for (double v : isect_s) si += v; for (double v : gather_s) sg += v;
  double chk = 0; for (double v : sink) chk += v;
  printf("{\"cmd\": \"bench\", \"rays\": %u, \"queries\": %u, \"photons\": %zu, \"k\": %d, \"threads\": %d, \"steps\": %d, "
         "\"kd_build_seconds\": %.6f, \"isect_seconds_per_step\": %.6f, \"gather_seconds_per_step\": %.6f, \"checksum\": %.6g}\n",
         nr, nq, ph.size(), k, threads, steps, build_sec, si / steps, sg / steps, chk);
  return 0;
}

// ---------------------------------------------------------------- shade (render.cpp:119-131: Intersects + RayTrace)
// usage: shade in.scn rays.bin out.bin [-load_global f.gipm] [-load_caustic f.gipm] [photonmap flags]
// The flags go through the reference's own ParseArgs (io_utils.cpp:16-214).  A loaded photon file replaces
// MapPhotons(): the photons are used as stored (already power-scaled) and the kd-tree is built like
// photonmap.cpp:364-377; a map that is not loaded switches its illumination off, like an empty map (:349-361).
// out.bin: u32 magic "GSHD", u32 n, n x f64 rgb[3] (what render.cpp:125/130 adds into image_buffer[i][j]).
static void irradiance_prepass(RNArray<Photon *> &photons, R3Kdtree<Photon *> *tree, int estimate_size, RNScalar estimate_dist);
static int cmd_shade(int argc, char **argv) {
  if (argc < 5) die("usage: shade in.scn rays.bin out.bin [-load_global f] [-load_caustic f] [photonmap flags]");
  const char *gfile = NULL, *cfile = NULL;
  std::vector<char *> av;
  av.push_back((char *)"photonmap"); av.push_back(argv[2]); av.push_back((char *)"unused.png");
  for (int i = 5; i < argc; i++) {
    if (!strcmp(argv[i], "-load_global") && i + 1 < argc) gfile = argv[++i];
    else if (!strcmp(argv[i], "-load_caustic") && i + 1 < argc) cfile = argv[++i];
    else av.push_back(argv[i]);
  }
  char *in_name = NULL, *out_name = NULL; int w = 0, h = 0, aa_ = 0; bool real = false;
  if (!ParseArgs((int)av.size(), av.data(), in_name, out_name, w, h, aa_, real)) die("bad flags");
  load_scene(argv[2], real);
  BuildDirectionLookupTable();
  static std::vector<Photon> gph, cph;
  if (gfile) gph = read_photons(gfile);
  if (cfile) cph = read_photons(cfile);
  for (size_t i = 0; i < gph.size(); i++) GLOBAL_PHOTONS.Insert(&gph[i]);
  for (size_t i = 0; i < cph.size(); i++) CAUSTIC_PHOTONS.Insert(&cph[i]);
  if (gph.empty()) { INDIRECT_ILLUM = false; DIRECT_PHOTON_ILLUM = false; FAST_GLOBAL = false; }
  else GLOBAL_PMAP = new R3Kdtree<Photon *>(GLOBAL_PHOTONS, offsetof(struct Photon, position));
  if (cph.empty()) CAUSTIC_ILLUM = false;
  else CAUSTIC_PMAP = new R3Kdtree<Photon *>(CAUSTIC_PHOTONS, offsetof(struct Photon, position));
  if (IRRADIANCE_CACHE && (INDIRECT_ILLUM || DIRECT_PHOTON_ILLUM))
    irradiance_prepass(GLOBAL_PHOTONS, GLOBAL_PMAP, GLOBAL_ESTIMATE_SIZE, GLOBAL_ESTIMATE_DIST);
  uint32_t n; std::vector<double> rays = read_rays(argv[3], n);
  std::vector<double> out((size_t)n * 3, 0.0);
  const R3Point eye = SCENE->Camera().Origin();
  int threads = THREADS;
  auto t0 = std::chrono::steady_clock::now();
  auto work = [&](int id) {
    RNInitThreadRandomness();
    for (uint32_t i = id; i < n; i += threads) {
      const double *r = &rays[(size_t)i * 6];
      R3Ray ray(R3Point(r[0], r[1], r[2]), R3Vector(r[3], r[4], r[5]), TRUE);
      R3SceneNode *node; R3SceneElement *element; R3Shape *shape; R3Point point; R3Vector normal; RNScalar t;
      RNRgb color = RNblack_rgb;
      if (SCENE->Intersects(ray, &node, &element, &shape, &point, &normal, &t)) RayTrace(element, point, normal, ray, eye, color);
      else color = SCENE->Background();
      out[(size_t)i * 3] = color.R(); out[(size_t)i * 3 + 1] = color.G(); out[(size_t)i * 3 + 2] = color.B();
    }
    RNClearThreadRandomness();
  };
  std::vector<std::thread> th;
  for (int i = 1; i < threads; i++) th.emplace_back(work, i);
  work(0);
  for (auto &x : th) x.join();
  double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  FILE *f = fopen(argv[4], "wb"); if (!f) die("cannot open output");
  uint32_t hdr[2] = {0x44485347, n};
  fwrite(hdr, 4, 2, f); fwrite(out.data(), 8, out.size(), f); fclose(f);
  printf("{\"cmd\": \"shade\", \"rays\": %u, \"threads\": %d, \"seconds\": %.6f}\n", n, threads, sec);
  return 0;
}

// ---------------------------------------------------------------- parse (ParseArgs, io_utils.cpp:16-214)
// usage: parse [photonmap flags]   -> one JSON object with every global the flags can touch
static int cmd_parse(int argc, char **argv) {
  std::vector<char *> av;
  av.push_back((char *)"photonmap"); av.push_back((char *)"in.scn"); av.push_back((char *)"out.png");
  for (int i = 2; i < argc; i++) av.push_back(argv[i]);
  char *in_name = NULL, *out_name = NULL; int w = 1024, h = 1024, aa_ = 2; bool real = false;
  if (!ParseArgs((int)av.size(), av.data(), in_name, out_name, w, h, aa_, real)) die("bad flags");
  printf("{\"verbose\": %d, \"threads\": %d, \"fresnel\": %d, \"ir_air\": %.17g, \"ambient\": %d, \"direct_illum\": %d, "
         "\"transmissive_illum\": %d, \"specular_illum\": %d, \"indirect_illum\": %d, \"caustic_illum\": %d, "
         "\"direct_photon_illum\": %d, \"fast_global\": %d, \"irradiance_cache\": %d, \"shadows\": %d, \"soft_shadows\": %d, "
         "\"light_test\": %d, \"shadow_test\": %d, \"monte_carlo\": %d, \"max_monte_depth\": %d, \"prob_absorb\": %.17g, "
         "\"recursive_shadows\": %d, \"distrib_transmissive\": %d, \"transmissive_test\": %d, \"distrib_specular\": %d, "
         "\"specular_test\": %d, \"depth_of_field\": %d, \"dof_test\": %d, \"focus_depth\": %.17g, \"aperture_radius\": %.17g, "
         "\"global_photon_count\": %d, \"caustic_photon_count\": %d, \"max_photon_depth\": %d, \"indirect_test\": %d, "
         "\"global_estimate_size\": %d, \"global_estimate_dist\": %.17g, \"global_filter\": %d, \"caustic_estimate_size\": %d, "
         "\"caustic_estimate_dist\": %.17g, \"caustic_filter\": %d, \"filter_const_a\": %.17g, \"filter_const_b\": %.17g, "
         "\"filter_const_k\": %.17g, \"width\": %d, \"height\": %d, \"aa\": %d, \"real_material\": %d}\n",
         (int)VERBOSE, THREADS, (int)FRESNEL, IR_AIR, (int)AMBIENT, (int)DIRECT_ILLUM, (int)TRANSMISSIVE_ILLUM, (int)SPECULAR_ILLUM,
         (int)INDIRECT_ILLUM, (int)CAUSTIC_ILLUM, (int)DIRECT_PHOTON_ILLUM, (int)FAST_GLOBAL, (int)IRRADIANCE_CACHE, (int)SHADOWS,
         (int)SOFT_SHADOWS, LIGHT_TEST, SHADOW_TEST, (int)MONTE_CARLO, MAX_MONTE_DEPTH, PROB_ABSORB, (int)RECURSIVE_SHADOWS,
         (int)DISTRIB_TRANSMISSIVE, TRANSMISSIVE_TEST, (int)DISTRIB_SPECULAR, SPECULAR_TEST, (int)DEPTH_OF_FIELD, DOF_TEST, FOCUS_DEPTH,
         APERTURE_RADIUS, GLOBAL_PHOTON_COUNT, CAUSTIC_PHOTON_COUNT, MAX_PHOTON_DEPTH, INDIRECT_TEST, GLOBAL_ESTIMATE_SIZE,
         GLOBAL_ESTIMATE_DIST, (int)GLOBAL_FILTER, CAUSTIC_ESTIMATE_SIZE, CAUSTIC_ESTIMATE_DIST, (int)CAUSTIC_FILTER, FILTER_CONST_A,
         FILTER_CONST_B, FILTER_CONST_K, w, h, aa_, (int)real);
  return 0;
}

int main(int argc, char **argv) {
  if (argc < 2) die("usage: ref_harness <dump_scene|camrays|isect|photons|gather|cache|shade> ...");
  std::string c = argv[1];
  if (c == "dump_scene") return cmd_dump_scene(argc, argv);
  if (c == "camrays") return cmd_camrays(argc, argv);
  if (c == "isect") return cmd_isect(argc, argv);
  if (c == "photons") return cmd_photons(argc, argv);
  if (c == "gather") return cmd_gather(argc, argv);
  if (c == "bench") return cmd_bench(argc, argv);
  if (c == "shade") return cmd_shade(argc, argv);
  if (c == "cache") return cmd_cache(argc, argv);
  if (c == "parse") return cmd_parse(argc, argv);
  die("unknown command");
  return 2;
}
#endif
