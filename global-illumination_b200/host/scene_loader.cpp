// host/scene_loader.cpp -- .scn / .off loader producing the flat GiScene (see scene.h).
//
// Semantics follow the reference's ReadPrinceton (src/R3Graphics/R3Scene.cpp:1463-1953),
// FindPrincetonMaterialAndElement (:1420-1459), ReadMesh (:1359-1394), R3Mesh::ReadOffFile
// (src/R3Shapes/R3Mesh.cpp:4075-4215) and the defaults added by R3Scene::ReadFile (:557-583).
// The data structure is ours: a temporary tree flattened into DFS-ordered POD arrays.
#include "scene.h"

#include <cctype>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <unordered_map>

namespace gi {
namespace {

struct V3 { double x, y, z; };
inline V3 sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 cross(V3 a, V3 b) {           // R3Vector::Cross (R3Vector.cpp:255-261)
  return {(a.y * b.z) - (a.z * b.y), (a.z * b.x) - (a.x * b.z), (a.x * b.y) - (a.y * b.x)};
}
inline double length(V3 a) { return std::sqrt((a.x * a.x) + (a.y * a.y) + (a.z * a.z)); }
inline V3 normalized(V3 a) {             // R3Vector::Normalize (R3Vector.cpp:232-239)
  double l = length(a);
  if (l == 0.0) return a;
  return {a.x / l, a.y / l, a.z / l};
}

struct Box {                             // R3Box with R3null_box = (FLT_MAX.. , -FLT_MAX..)
  double lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
  double hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  bool empty() const { return lo[0] > hi[0] || lo[1] > hi[1] || lo[2] > hi[2]; }
  void add(const double p[3]) {
    for (int i = 0; i < 3; i++) { if (p[i] < lo[i]) lo[i] = p[i]; if (p[i] > hi[i]) hi[i] = p[i]; }
  }
  void add(const Box &b) {
    if (b.empty()) return;
    add(b.lo); add(b.hi);
  }
};

struct TmpShape {
  int type;                              // GiShapeType, or -1 for a mesh
  double d[9];
  std::vector<double> tris;              // mesh: 9 doubles per triangle
  std::vector<double> verts;             // mesh: every vertex (the mesh bbox covers unused vertices too)
};
struct TmpElement { int material; std::vector<TmpShape> shapes; };
struct TmpNode {
  double m[16];
  int parent = -1;
  std::vector<TmpElement> elements;
  std::vector<int> children;
};

struct Builder {
  std::vector<TmpNode> nodes;
  std::vector<GiMaterial> materials;
  std::vector<GiLight> lights;
  GiCamera camera{};
  bool has_camera = false;
  double ambient[3] = {0, 0, 0}, background[3] = {0, 0, 0};
  bool real_material = false;
  std::string err;
};

// ---------------------------------------------------------------------------------------------
// .off reader: fan triangulation + the half-edge manifold rule that defers "degenerate" faces to
// the end of the face list (R3Mesh.cpp:4163-4208, CreateFace :1166-1193, UpdateFaceRefs :5825-5845)
// ---------------------------------------------------------------------------------------------
struct OffMesh {
  std::vector<V3> verts;
  std::vector<int> faces;                // 3 vertex ids per face
  struct Edge { int v0; bool face0 = false, face1 = false; };
  std::unordered_map<uint64_t, Edge> edges;

  Edge &edge(int a, int b) {             // EdgeBetweenVertices / CreateEdge(a, b): vertex[0] = a on creation
    uint64_t key = a < b ? ((uint64_t)a << 32) | (uint32_t)b : ((uint64_t)b << 32) | (uint32_t)a;
    auto it = edges.find(key);
    if (it == edges.end()) { Edge e; e.v0 = a; it = edges.emplace(key, e).first; }
    return it->second;
  }
  bool create_face(int v1, int v2, int v3) {
    Edge &e1 = edge(v1, v2); Edge &e2 = edge(v2, v3); Edge &e3 = edge(v3, v1);
    // "two faces share same side of same edge" rejections
    if ((e1.v0 == v1) && e1.face0) return false;
    if ((e1.v0 == v2) && e1.face1) return false;
    if ((e2.v0 == v2) && e2.face0) return false;
    if ((e2.v0 == v3) && e2.face1) return false;
    if ((e3.v0 == v3) && e3.face0) return false;
    if ((e3.v0 == v1) && e3.face1) return false;
    if (e1.v0 == v1) e1.face0 = true; else e1.face1 = true;
    if (e2.v0 == v2) e2.face0 = true; else e2.face1 = true;
    if (e3.v0 == v3) e3.face0 = true; else e3.face1 = true;
    faces.push_back(v1); faces.push_back(v2); faces.push_back(v3);
    return true;
  }
};

bool read_off(const std::string &path, TmpShape &shape, std::string &err) {
  FILE *fp = fopen(path.c_str(), "r");
  if (!fp) { err = "Unable to open file " + path + "\nUnable to read mesh " + path; return false; }
  OffMesh mesh;
  int nverts = 0, nfaces = 0, nedges = 0, line_count = 0, vertex_count = 0, face_count = 0;
  char buffer[1024], header[64];
  std::vector<int> degenerate;
  bool ok = true;
  while (fgets(buffer, 1023, fp)) {
    line_count++;
    char *bp = buffer;
    while (isspace((unsigned char)*bp)) bp++;
    if (*bp == '#' || *bp == '\0') continue;
    if (nverts == 0) {
      if (strstr(bp, "OFF")) {
        int tmp;
        if (sscanf(bp, "%63s%d%d%d", header, &tmp, &nfaces, &nedges) == 4) nverts = tmp;
      } else if ((sscanf(bp, "%d%d%d", &nverts, &nfaces, &nedges) != 3) || (nverts == 0)) {
        err = "Syntax error reading header on line " + std::to_string(line_count) + " in file " + path;
        ok = false; break;
      }
    } else if (vertex_count < nverts) {
      double x, y, z;
      if (sscanf(bp, "%lf%lf%lf", &x, &y, &z) != 3) {
        err = "Syntax error with vertex coordinates on line " + std::to_string(line_count) + " in file " + path;
        ok = false; break;
      }
      mesh.verts.push_back({x, y, z});
      vertex_count++;
    } else if (face_count < nfaces) {
      int face_nverts = 0;
      char *tok = strtok(bp, " \t");
      if (tok) face_nverts = atoi(tok);
      else { err = "Syntax error with face on line " + std::to_string(line_count) + " in file " + path; ok = false; break; }
      int v1 = -1, v2 = -1, v3 = -1;
      for (int i = 0; i < face_nverts; i++) {
        tok = strtok(NULL, " \t");
        if (!tok) { err = "Syntax error with face on line " + std::to_string(line_count) + " in file " + path; ok = false; break; }
        int v = atoi(tok);
        if (v < 0 || v >= (int)mesh.verts.size()) { err = "Bad vertex index on line " + std::to_string(line_count) + " in file " + path; ok = false; break; }
        if (v1 < 0) v1 = v; else v3 = v;
        if (v1 >= 0 && v2 >= 0 && v3 >= 0 && (v1 != v2) && (v2 != v3) && (v1 != v3)) {
          if (!mesh.create_face(v1, v2, v3)) { degenerate.push_back(v1); degenerate.push_back(v2); degenerate.push_back(v3); }
        }
        v2 = v3;
      }
      if (!ok) break;
      face_count++;
    } else {
      break;                             // "Found extra text": the reference warns and stops reading
    }
  }
  fclose(fp);
  if (!ok) { err += "\nUnable to read mesh " + path; return false; }
  for (size_t i = 0; i + 3 <= degenerate.size(); i += 3) {
    int v1 = degenerate[i], v2 = degenerate[i + 1], v3 = degenerate[i + 2];
    if (!mesh.create_face(v1, v2, v3)) {
      if (!mesh.create_face(v1, v3, v2)) {
        int a = (int)mesh.verts.size();
        mesh.verts.push_back(mesh.verts[v1]); mesh.verts.push_back(mesh.verts[v2]); mesh.verts.push_back(mesh.verts[v3]);
        mesh.create_face(a, a + 1, a + 2);
      }
    }
  }
  shape.type = -1;
  shape.tris.reserve(mesh.faces.size() * 3);
  for (int v : mesh.faces) { shape.tris.push_back(mesh.verts[v].x); shape.tris.push_back(mesh.verts[v].y); shape.tris.push_back(mesh.verts[v].z); }
  shape.verts.reserve(mesh.verts.size() * 3);
  for (const V3 &v : mesh.verts) { shape.verts.push_back(v.x); shape.verts.push_back(v.y); shape.verts.push_back(v.z); }
  return true;
}

// ---------------------------------------------------------------------------------------------
// .scn reader (R3Scene.cpp:1463-1953)
// ---------------------------------------------------------------------------------------------
static const GiMaterial kDefaultBrdf = {   // R3default_brdf (R3Brdf.cpp:13-15): opacity 1 => kt = 0
    {0.2, 0.2, 0.2}, {0.8, 0.8, 0.8}, {0, 0, 0}, {0, 0, 0}, {0, 0, 0}, 0.2, 1.0};

std::string dir_of(const std::string &filename) {
  size_t p = filename.rfind('/');
  return p == std::string::npos ? std::string() : filename.substr(0, p + 1);
}

// FindPrincetonMaterialAndElement: material index -> element of `node` with that material
bool find_element(Builder &b, int node, const std::vector<int> &parsed, int m, int &group_material,
                  TmpElement *&element) {
  int material;
  if (m >= 0) {
    if (m < (int)parsed.size()) material = parsed[m];
    else return false;
  } else {
    material = group_material;
    if (material < 0) {                    // lazily created copy of the default brdf, one per group
      b.materials.push_back(kDefaultBrdf);
      material = (int)b.materials.size() - 1;
      group_material = material;
    }
  }
  for (TmpElement &e : b.nodes[node].elements)
    if (e.material == material) { element = &e; return true; }
  b.nodes[node].elements.push_back(TmpElement{material, {}});
  element = &b.nodes[node].elements.back();
  return true;
}

bool read_princeton(Builder &b, int root_node, const std::string &filename) {
  FILE *fp = fopen(filename.c_str(), "r");
  if (!fp) { b.err = "Unable to open file " + filename; return false; }
  std::vector<int> parsed;                 // parsed_materials: local to this file
  int material = -1;                       // the `material` variable (sticky across commands, :1474)
  const int max_depth = 1024;
  std::vector<int> group_nodes(max_depth, -1), group_materials(max_depth, -1);
  group_nodes[0] = root_node;
  int depth = 0;
  char cmd[128];
  int command_number = 1;
  auto fail = [&](const char *what) {
    b.err = std::string(what) + " at command " + std::to_string(command_number) + " in file " + filename;
    fclose(fp);
    return false;
  };
  auto add_shape = [&](int m, const TmpShape &s) {
    TmpElement *element = nullptr;
    if (!find_element(b, group_nodes[depth], parsed, m, group_materials[depth], element)) return false;
    // note: like the reference, `material` is overwritten by every shape command
    material = element->material;
    element->shapes.push_back(s);
    return true;
  };
  while (fscanf(fp, "%127s", cmd) == 1) {
    if (cmd[0] == '#') {
      int c;
      do { c = fgetc(fp); } while ((c >= 0) && (c != '\n'));
    } else if (!strcmp(cmd, "tri")) {
      int m; TmpShape s; s.type = GI_SHAPE_TRIANGLE;
      if (fscanf(fp, "%d%lf%lf%lf%lf%lf%lf%lf%lf%lf", &m, &s.d[0], &s.d[1], &s.d[2], &s.d[3], &s.d[4], &s.d[5],
                 &s.d[6], &s.d[7], &s.d[8]) != 10) return fail("Unable to read triangle");
      if (!add_shape(m, s)) return fail("Invalid material id");
    } else if (!strcmp(cmd, "box")) {
      int m; TmpShape s; s.type = GI_SHAPE_BOX; memset(s.d, 0, sizeof(s.d));
      if (fscanf(fp, "%d%lf%lf%lf%lf%lf%lf", &m, &s.d[0], &s.d[1], &s.d[2], &s.d[3], &s.d[4], &s.d[5]) != 7)
        return fail("Unable to read box");
      for (int i = 0; i < 3; i++) if (s.d[i] > s.d[3 + i]) std::swap(s.d[i], s.d[3 + i]);
      if (!add_shape(m, s)) return fail("Invalid material id");
    } else if (!strcmp(cmd, "sphere")) {
      int m; TmpShape s; s.type = GI_SHAPE_SPHERE; memset(s.d, 0, sizeof(s.d));
      if (fscanf(fp, "%d%lf%lf%lf%lf", &m, &s.d[0], &s.d[1], &s.d[2], &s.d[3]) != 5) return fail("Unable to read sphere");
      if (!add_shape(m, s)) return fail("Invalid material id");
    } else if (!strcmp(cmd, "circle")) {
      int m; TmpShape s; s.type = GI_SHAPE_CIRCLE; memset(s.d, 0, sizeof(s.d));
      V3 d;
      if (fscanf(fp, "%d%lf%lf%lf%lf%lf%lf%lf", &m, &s.d[0], &s.d[1], &s.d[2], &d.x, &d.y, &d.z, &s.d[6]) != 8)
        return fail("Unable to read circle");
      d = normalized(d);
      s.d[3] = d.x; s.d[4] = d.y; s.d[5] = d.z;
      if (!add_shape(m, s)) return fail("Invalid material id");
    } else if (!strcmp(cmd, "cylinder") || !strcmp(cmd, "cone")) {
      bool cyl = !strcmp(cmd, "cylinder");
      int m; double c[3], r, h;
      if (fscanf(fp, "%d%lf%lf%lf%lf%lf", &m, &c[0], &c[1], &c[2], &r, &h) != 6)
        return fail(cyl ? "Unable to read cylinder" : "Unable to read cone");
      TmpShape s; s.type = cyl ? GI_SHAPE_CYLINDER : GI_SHAPE_CONE; memset(s.d, 0, sizeof(s.d));
      // p1 = c - 0.5*h*(0,1,0); p2 = c + 0.5*h*(0,1,0)   (operator order of R3Scene.cpp:1603-1604)
      double hh = 0.5 * h;
      s.d[0] = c[0] - hh * 0.0; s.d[1] = c[1] - hh * 1.0; s.d[2] = c[2] - hh * 0.0;
      s.d[3] = c[0] + hh * 0.0; s.d[4] = c[1] + hh * 1.0; s.d[5] = c[2] + hh * 0.0;
      s.d[6] = r;
      if (!add_shape(m, s)) return fail("Invalid material id");
    } else if (!strcmp(cmd, "mesh")) {
      int m; char meshname[256];
      if (fscanf(fp, "%d%255s", &m, meshname) != 2) return fail("Unable to parse mesh command");
      TmpShape s;
      if (!read_off(dir_of(filename) + meshname, s, b.err)) { fclose(fp); return false; }
      if (!add_shape(m, s)) return fail("Invalid material id");
    } else if (!strcmp(cmd, "line")) {
      int m; TmpShape s; s.type = GI_SHAPE_CYLINDER; memset(s.d, 0, sizeof(s.d));
      if (fscanf(fp, "%d%lf%lf%lf%lf%lf%lf", &m, &s.d[0], &s.d[1], &s.d[2], &s.d[3], &s.d[4], &s.d[5]) != 7)
        return fail("Unable to read line");
      s.d[6] = 1.0E-3;                     // RN_BIG_EPSILON
      if (!add_shape(m, s)) return fail("Invalid material id");
    } else if (!strcmp(cmd, "begin")) {
      int m; double mat[16];
      if (fscanf(fp, "%d%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf", &m, &mat[0], &mat[1], &mat[2], &mat[3],
                 &mat[4], &mat[5], &mat[6], &mat[7], &mat[8], &mat[9], &mat[10], &mat[11], &mat[12], &mat[13],
                 &mat[14], &mat[15]) != 17) return fail("Unable to read begin");
      if (m >= 0) material = (m < (int)parsed.size()) ? parsed[m] : -1;
      if (depth + 1 >= max_depth) return fail("Group nesting too deep");
      TmpNode node; memcpy(node.m, mat, sizeof(mat)); node.parent = group_nodes[depth];
      b.nodes.push_back(node);
      int id = (int)b.nodes.size() - 1;
      b.nodes[group_nodes[depth]].children.push_back(id);
      depth++;
      group_nodes[depth] = id;
      group_materials[depth] = material;
    } else if (!strcmp(cmd, "end")) {
      if (depth <= 0) return fail("Extra end statement");
      depth--;
    } else if (!strcmp(cmd, "material")) {
      GiMaterial g; char texture_name[256];
      if (fscanf(fp, "%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%255s", &g.ka[0], &g.ka[1], &g.ka[2],
                 &g.kd[0], &g.kd[1], &g.kd[2], &g.ks[0], &g.ks[1], &g.ks[2], &g.kt[0], &g.kt[1], &g.kt[2],
                 &g.ke[0], &g.ke[1], &g.ke[2], &g.shininess, &g.ir, texture_name) != 18)
        return fail("Unable to read material");
      if (b.real_material) {               // -real (R3Scene.cpp:1742-1755)
        double max_val = 1.0;
        for (int i = 0; i < 3; i++) { double t = g.kd[i] + g.ks[i] + g.kt[i]; if (t > max_val) max_val = t; }
        if (max_val > 1.0) for (int i = 0; i < 3; i++) { g.kd[i] /= max_val; g.ks[i] /= max_val; g.kt[i] /= max_val; }
      }
      // textures are loaded by the reference but never sampled (README.md:127): ignored here
      b.materials.push_back(g);
      parsed.push_back((int)b.materials.size() - 1);
    } else if (!strcmp(cmd, "dir_light")) {
      GiLight l; memset(&l, 0, sizeof(l)); l.type = GI_LIGHT_DIR; l.active = 1; l.intensity = 1;
      V3 d;
      if (fscanf(fp, "%lf%lf%lf%lf%lf%lf", &l.color[0], &l.color[1], &l.color[2], &d.x, &d.y, &d.z) != 6)
        return fail("Unable to read directional light");
      d = normalized(normalized(d));       // normalised by the parser and again by the light ctor
      l.direction[0] = d.x; l.direction[1] = d.y; l.direction[2] = d.z;
      b.lights.push_back(l);
    } else if (!strcmp(cmd, "point_light")) {
      GiLight l; memset(&l, 0, sizeof(l)); l.type = GI_LIGHT_POINT; l.active = 1; l.intensity = 1;
      if (fscanf(fp, "%lf%lf%lf%lf%lf%lf%lf%lf%lf", &l.color[0], &l.color[1], &l.color[2], &l.position[0],
                 &l.position[1], &l.position[2], &l.ca, &l.la, &l.qa) != 9) return fail("Unable to read point light");
      b.lights.push_back(l);
    } else if (!strcmp(cmd, "spot_light")) {
      GiLight l; memset(&l, 0, sizeof(l)); l.type = GI_LIGHT_SPOT; l.active = 1; l.intensity = 1;
      V3 d; double sc, sd;
      if (fscanf(fp, "%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf", &l.color[0], &l.color[1], &l.color[2],
                 &l.position[0], &l.position[1], &l.position[2], &d.x, &d.y, &d.z, &l.ca, &l.la, &l.qa, &sc, &sd) != 14)
        return fail("Unable to read point light");
      d = normalized(normalized(d));       // parser + R3SpotLight ctor (R3SpotLight.cpp:72)
      l.direction[0] = d.x; l.direction[1] = d.y; l.direction[2] = d.z;
      l.dropoff = sd; l.cutoff = sc;       // R3SpotLight(p, d, c, sd, sc, ...) = (dropoff, cutoff) (:1836)
      b.lights.push_back(l);
    } else if (!strcmp(cmd, "area_light")) {
      GiLight l; memset(&l, 0, sizeof(l)); l.type = GI_LIGHT_AREA; l.active = 1; l.intensity = 1;
      V3 d;
      if (fscanf(fp, "%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf", &l.color[0], &l.color[1], &l.color[2], &l.position[0],
                 &l.position[1], &l.position[2], &d.x, &d.y, &d.z, &l.radius, &l.ca, &l.la, &l.qa) != 13)
        return fail("Unable to read area light");
      d = normalized(d); l.direction[0] = d.x; l.direction[1] = d.y; l.direction[2] = d.z;
      b.lights.push_back(l);
    } else if (!strcmp(cmd, "rect_light")) {
      GiLight l; memset(&l, 0, sizeof(l)); l.type = GI_LIGHT_RECT; l.active = 1; l.intensity = 1;
      V3 a1, a2;
      if (fscanf(fp, "%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf", &l.color[0], &l.color[1], &l.color[2],
                 &l.position[0], &l.position[1], &l.position[2], &a1.x, &a1.y, &a1.z, &a2.x, &a2.y, &a2.z, &l.len1,
                 &l.len2, &l.ca, &l.la, &l.qa) != 17) return fail("Unable to read rect light");
      a1 = normalized(normalized(a1)); a2 = normalized(normalized(a2));   // parser + ctor (R3RectLight.cpp:63-66)
      l.axis1[0] = a1.x; l.axis1[1] = a1.y; l.axis1[2] = a1.z;
      l.axis2[0] = a2.x; l.axis2[1] = a2.y; l.axis2[2] = a2.z;
      V3 n = normalized(cross(a1, a2));    // R3RectLight ctor: norm = a1 % a2, normalised
      l.direction[0] = n.x; l.direction[1] = n.y; l.direction[2] = n.z;
      b.lights.push_back(l);
    } else if (!strcmp(cmd, "camera")) {
      double e[3]; V3 t, u; double xfov, nd, fd;
      if (fscanf(fp, "%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf%lf", &e[0], &e[1], &e[2], &t.x, &t.y, &t.z, &u.x, &u.y, &u.z,
                 &xfov, &nd, &fd) != 12) return fail("Unable to read camera");
      // R3Triad(towards, up) (R3Triad.cpp:72-79): z = normalize(-t); x = normalize(up % z); y = z % x
      V3 az = normalized(V3{-t.x, -t.y, -t.z});
      V3 ax = normalized(cross(u, az));
      V3 ay = cross(az, ax);
      GiCamera &c = b.camera;
      memcpy(c.eye, e, sizeof(e));
      c.towards[0] = -az.x; c.towards[1] = -az.y; c.towards[2] = -az.z;
      c.up[0] = ay.x; c.up[1] = ay.y; c.up[2] = ay.z;
      c.right[0] = ax.x; c.right[1] = ax.y; c.right[2] = ax.z;
      c.xfov = xfov; c.yfov = xfov; c.neardist = nd; c.fardist = fd;
      b.has_camera = true;
    } else if (!strcmp(cmd, "include")) {
      char scenename[256];
      if (fscanf(fp, "%255s", scenename) != 1) return fail("Unable to read include command");
      std::string inc = dir_of(filename) + scenename;
      if (!read_princeton(b, group_nodes[depth], inc)) {
        b.err += "\nUnable to read included scene: " + inc;
        fclose(fp);
        return false;
      }
    } else if (!strcmp(cmd, "background")) {
      if (fscanf(fp, "%lf%lf%lf", &b.background[0], &b.background[1], &b.background[2]) != 3)
        return fail("Unable to read background");
    } else if (!strcmp(cmd, "ambient")) {
      if (fscanf(fp, "%lf%lf%lf", &b.ambient[0], &b.ambient[1], &b.ambient[2]) != 3)
        return fail("Unable to read ambient");
    } else {
      b.err = "Unrecognized command " + std::to_string(command_number) + " in file " + filename + ": " + cmd;
      fclose(fp);
      return false;
    }
    command_number++;
  }
  fclose(fp);
  return true;
}

// ---------------------------------------------------------------------------------------------
// matrices, bounding boxes, flattening
// ---------------------------------------------------------------------------------------------
inline double det3(double a, double b, double c, double d, double e, double f, double g, double h, double i) {
  return (a * (e * i - h * f) - b * (d * i - g * f) + c * (d * h - g * e));   // R4MatrixDet3 (R4Matrix.cpp:807-814)
}

void invert4(const double m[16], double out[16]) {      // R4Matrix::Invert (R4Matrix.cpp:252-293)
  double Ma = m[0], Mb = m[1], Mc = m[2], Md = m[3], Me = m[4], Mf = m[5], Mg = m[6], Mh = m[7];
  double Mi = m[8], Mj = m[9], Mk = m[10], Ml = m[11], Mm = m[12], Mn = m[13], Mo = m[14], Mp = m[15];
  double a1 = det3(Mf, Mg, Mh, Mj, Mk, Ml, Mn, Mo, Mp);
  double a2 = det3(Me, Mg, Mh, Mi, Mk, Ml, Mm, Mo, Mp);
  double a3 = det3(Me, Mf, Mh, Mi, Mj, Ml, Mm, Mn, Mp);
  double a4 = det3(Me, Mf, Mg, Mi, Mj, Mk, Mm, Mn, Mo);
  double det = Ma * a1 - Mb * a2 + Mc * a3 - Md * a4;
  if (det >= -1E-8 && det <= 1E-8) { memcpy(out, m, 16 * sizeof(double)); return; }   // left unchanged + warning
  out[0] = (a1) / det;  out[4] = -(a2) / det;  out[8] = (a3) / det;  out[12] = -(a4) / det;
  out[1] = -(det3(Mb, Mc, Md, Mj, Mk, Ml, Mn, Mo, Mp)) / det;
  out[5] = (det3(Ma, Mc, Md, Mi, Mk, Ml, Mm, Mo, Mp)) / det;
  out[9] = -(det3(Ma, Mb, Md, Mi, Mj, Ml, Mm, Mn, Mp)) / det;
  out[13] = (det3(Ma, Mb, Mc, Mi, Mj, Mk, Mm, Mn, Mo)) / det;
  out[2] = (det3(Mb, Mc, Md, Mf, Mg, Mh, Mn, Mo, Mp)) / det;
  out[6] = -(det3(Ma, Mc, Md, Me, Mg, Mh, Mm, Mo, Mp)) / det;
  out[10] = (det3(Ma, Mb, Md, Me, Mf, Mh, Mm, Mn, Mp)) / det;
  out[14] = -(det3(Ma, Mb, Mc, Me, Mf, Mg, Mm, Mn, Mo)) / det;
  out[3] = -(det3(Mb, Mc, Md, Mf, Mg, Mh, Mj, Mk, Ml)) / det;
  out[7] = (det3(Ma, Mc, Md, Me, Mg, Mh, Mi, Mk, Ml)) / det;
  out[11] = -(det3(Ma, Mb, Md, Me, Mf, Mh, Mi, Mj, Ml)) / det;
  out[15] = (det3(Ma, Mb, Mc, Me, Mf, Mg, Mi, Mj, Mk)) / det;
}

void xform_point(const double m[16], const double p[3], double out[3]) {   // R4Matrix * R3Point (R4Matrix.cpp:786-793)
  out[0] = m[0] * p[0] + m[1] * p[1] + m[2] * p[2] + m[3];
  out[1] = m[4] * p[0] + m[5] * p[1] + m[6] * p[2] + m[7];
  out[2] = m[8] * p[0] + m[9] * p[1] + m[10] * p[2] + m[11];
}

Box xform_box(const Box &b, const double m[16]) {        // R3Box::Transform (R3Box.cpp:525-538)
  if (b.empty()) return b;
  Box t;
  for (int o = 0; o < 8; o++) {
    double c[3] = {(o & 4) ? b.hi[0] : b.lo[0], (o & 2) ? b.hi[1] : b.lo[1], (o & 1) ? b.hi[2] : b.lo[2]};
    double q[3]; xform_point(m, c, q); t.add(q);
  }
  return t;
}

Box shape_bbox(const TmpShape &s) {
  Box b;
  switch (s.type) {
    case -1: for (size_t i = 0; i + 3 <= s.verts.size(); i += 3) b.add(&s.verts[i]); break;
    case GI_SHAPE_TRIANGLE: b.add(&s.d[0]); b.add(&s.d[3]); b.add(&s.d[6]); break;
    case GI_SHAPE_BOX: b.add(&s.d[0]); b.add(&s.d[3]); break;
    case GI_SHAPE_SPHERE: case GI_SHAPE_CIRCLE: {
      double r = s.type == GI_SHAPE_SPHERE ? s.d[3] : s.d[6];
      double lo[3] = {s.d[0] - r, s.d[1] - r, s.d[2] - r}, hi[3] = {s.d[0] + r, s.d[1] + r, s.d[2] + r};
      b.add(lo); b.add(hi); break;
    }
    case GI_SHAPE_CYLINDER: case GI_SHAPE_CONE: {       // union of the cap circles' cubes (R3Cylinder.cpp:175-182)
      double r = s.d[6];
      for (int e = 0; e < 2; e++) {
        const double *c = &s.d[3 * e];
        double lo[3] = {c[0] - r, c[1] - r, c[2] - r}, hi[3] = {c[0] + r, c[1] + r, c[2] + r};
        b.add(lo); b.add(hi);
      }
      break;
    }
  }
  return b;
}

Box node_bbox(const Builder &b, int n) {                 // R3SceneNode::UpdateBBox (R3SceneNode.cpp:539-560)
  const TmpNode &node = b.nodes[n];
  Box box;
  for (const TmpElement &e : node.elements) {
    Box eb;
    for (const TmpShape &s : e.shapes) eb.add(shape_bbox(s));
    box.add(xform_box(eb, node.m));
  }
  for (int c : node.children) box.add(xform_box(node_bbox(b, c), node.m));
  return box;
}

void flatten(const Builder &b, int n, int parent_flat, SceneData &out) {
  const TmpNode &node = b.nodes[n];
  GiNode g; memset(&g, 0, sizeof(g));
  g.parent = parent_flat;
  memcpy(g.m, node.m, sizeof(g.m));
  invert4(node.m, g.minv);
  Box nb = node_bbox(b, n);
  for (int i = 0; i < 3; i++) { g.bbox[i] = nb.lo[i]; g.bbox[3 + i] = nb.hi[i]; }
  int me = (int)out.nodes.size();
  out.nodes.push_back(g);
  for (const TmpElement &e : node.elements) {
    int eid = (int)out.elements.size();
    GiElement ge; memset(&ge, 0, sizeof(ge));
    ge.node = me; ge.material = e.material;
    Box eb;
    for (const TmpShape &s : e.shapes) eb.add(shape_bbox(s));
    for (int i = 0; i < 3; i++) { ge.bbox[i] = eb.lo[i]; ge.bbox[3 + i] = eb.hi[i]; }
    out.elements.push_back(ge);
    for (const TmpShape &s : e.shapes) {
      int sid = out.n_shapes++;
      out.shape_is_mesh.push_back(s.type == -1 ? 1 : 0);
      GiShape gs; memset(&gs, 0, sizeof(gs));
      gs.is_mesh = s.type == -1 ? 1 : 0; gs.element = eid; gs.first_prim = (int)out.prims.size();
      gs.n_prims = s.type == -1 ? (int)(s.tris.size() / 9) : 1;
      Box sb = shape_bbox(s);
      for (int i = 0; i < 3; i++) { gs.bbox[i] = sb.lo[i]; gs.bbox[3 + i] = sb.hi[i]; }
      out.shapes.push_back(gs);
      if (s.type == -1) {
        for (size_t t = 0; t * 9 < s.tris.size(); t++) {
          GiPrim p; p.type = GI_SHAPE_TRIANGLE; p.element = eid; p.shape = sid; p.index_in_shape = (int)t;
          memcpy(p.d, &s.tris[t * 9], 9 * sizeof(double));
          out.prims.push_back(p);
        }
      } else {
        GiPrim p; p.type = s.type; p.element = eid; p.shape = sid; p.index_in_shape = 0;
        memcpy(p.d, s.d, sizeof(p.d));
        out.prims.push_back(p);
      }
    }
  }
  for (int c : node.children) flatten(b, c, me, out);
}

}  // namespace

GiScene SceneData::view() const {
  GiScene s; memset(&s, 0, sizeof(s));
  s.n_nodes = (int)nodes.size(); s.n_elements = (int)elements.size(); s.n_materials = (int)materials.size();
  s.n_prims = (int)prims.size(); s.n_lights = (int)lights.size(); s.n_shapes = n_shapes;
  s.nodes = nodes.data(); s.elements = elements.data(); s.materials = materials.data();
  s.prims = prims.data(); s.shapes = shapes.data(); s.lights = lights.data();
  s.camera = camera;
  for (int i = 0; i < 3; i++) {
    s.ambient[i] = ambient[i]; s.background[i] = background[i];
    s.bbox_min[i] = bbox_min[i]; s.bbox_max[i] = bbox_max[i]; s.centroid[i] = centroid[i];
  }
  s.radius = radius;
  return s;
}

bool load_scene(const std::string &path, bool real_material, SceneData &out, std::string &err) {
  const char *ext = strrchr(path.c_str(), '.');
  if (!ext) { err = "Filename " + path + " has no extension (e.g., .txt)"; return false; }
  if (strncmp(ext, ".scn", 4)) { err = "Unable to read file " + path + " (unrecognized extension: " + ext + ")"; return false; }
  Builder b;
  b.real_material = real_material;
  TmpNode root;
  static const double ident[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
  memcpy(root.m, ident, sizeof(ident));
  b.nodes.push_back(root);
  if (!read_princeton(b, 0, path)) { err = b.err; return false; }

  out = SceneData();
  flatten(b, 0, -1, out);
  out.materials = b.materials;
  out.lights = b.lights;
  for (int i = 0; i < 3; i++) { out.ambient[i] = b.ambient[i]; out.background[i] = b.background[i]; }

  Box box = node_bbox(b, 0);
  for (int i = 0; i < 3; i++) {
    out.bbox_min[i] = box.lo[i]; out.bbox_max[i] = box.hi[i];
    out.centroid[i] = 0.5 * (box.lo[i] + box.hi[i]);
  }
  V3 diag = sub(V3{box.lo[0], box.lo[1], box.lo[2]}, V3{box.hi[0], box.hi[1], box.hi[2]});
  out.radius = 0.5 * length(diag);

  if (b.has_camera) {
    out.camera = b.camera; out.has_camera = true;
  } else {
    // default camera (R3Scene.cpp:557-566): eye = centre + 3 r z, towards -z, up y, fov 0.25
    GiCamera &c = out.camera; memset(&c, 0, sizeof(c));
    V3 towards{0, 0, -1}, up{0, 1, 0};
    for (int i = 0; i < 3; i++) c.eye[i] = out.centroid[i];
    double k = 3 * out.radius;
    c.eye[0] = out.centroid[0] - k * towards.x; c.eye[1] = out.centroid[1] - k * towards.y; c.eye[2] = out.centroid[2] - k * towards.z;
    V3 az = normalized(V3{-towards.x, -towards.y, -towards.z});
    V3 ax = normalized(cross(up, az));
    V3 ay = cross(az, ax);
    c.towards[0] = -az.x; c.towards[1] = -az.y; c.towards[2] = -az.z;
    c.up[0] = ay.x; c.up[1] = ay.y; c.up[2] = ay.z;
    c.right[0] = ax.x; c.right[1] = ax.y; c.right[2] = ax.z;
    c.xfov = 0.25; c.yfov = 0.25; c.neardist = 0.01 * out.radius; c.fardist = 100 * out.radius;
  }
  if (out.lights.empty()) {              // two default directional lights (R3Scene.cpp:568-583)
    const double col[2][3] = {{1, 1, 1}, {0.5, 0.5, 0.5}};
    const V3 dir[2] = {{-3, -4, -5}, {3, 2, 3}};
    for (int k = 0; k < 2; k++) {
      GiLight l; memset(&l, 0, sizeof(l)); l.type = GI_LIGHT_DIR; l.active = 1; l.intensity = 1;
      V3 d = normalized(normalized(dir[k]));
      l.direction[0] = d.x; l.direction[1] = d.y; l.direction[2] = d.z;
      for (int i = 0; i < 3; i++) l.color[i] = col[k][i];
      out.lights.push_back(l);
    }
  }
  return true;
}

static void p3(FILE *f, const double *v) { fprintf(f, " %.17g %.17g %.17g", v[0], v[1], v[2]); }

bool dump_scene(const SceneData &s, const std::string &path) {
  FILE *f = fopen(path.c_str(), "w");
  if (!f) return false;
  fprintf(f, "nodes %d\n", (int)s.nodes.size());
  for (size_t i = 0; i < s.nodes.size(); i++) {
    fprintf(f, "node %d parent %d\n m", (int)i, s.nodes[i].parent);
    for (int k = 0; k < 16; k++) fprintf(f, " %.17g", s.nodes[i].m[k]);
    fprintf(f, "\n minv");
    for (int k = 0; k < 16; k++) fprintf(f, " %.17g", s.nodes[i].minv[k]);
    fprintf(f, "\n");
  }
  fprintf(f, "elements %d\n", (int)s.elements.size());
  for (size_t i = 0; i < s.elements.size(); i++) {
    const GiMaterial &m = s.materials[s.elements[i].material];
    fprintf(f, "element %d\n ka", (int)i); p3(f, m.ka);
    fprintf(f, "\n kd"); p3(f, m.kd); fprintf(f, "\n ks"); p3(f, m.ks);
    fprintf(f, "\n kt"); p3(f, m.kt); fprintf(f, "\n ke"); p3(f, m.ke);
    fprintf(f, "\n n %.17g ir %.17g\n", m.shininess, m.ir);
  }
  fprintf(f, "shapes %d\n", s.n_shapes);
  // shapes that produced no primitive (empty meshes) still own an id
  size_t pi = 0;
  for (int sid = 0; sid < s.n_shapes; sid++) {
    if (pi >= s.prims.size() || s.prims[pi].shape != sid) { fprintf(f, "shape %d empty\n", sid); continue; }
    const GiPrim &p = s.prims[pi];
    fprintf(f, "shape %d node %d element %d ", sid, s.elements[p.element].node, p.element);
    size_t end = pi;
    while (end < s.prims.size() && s.prims[end].shape == sid) end++;
    switch (p.type) {
      case GI_SHAPE_SPHERE: fprintf(f, "sphere"); p3(f, p.d); fprintf(f, " %.17g\n", p.d[3]); break;
      case GI_SHAPE_BOX: fprintf(f, "box"); p3(f, p.d); p3(f, p.d + 3); fprintf(f, "\n"); break;
      case GI_SHAPE_CYLINDER: fprintf(f, "cylinder"); p3(f, p.d); p3(f, p.d + 3); fprintf(f, " %.17g\n", p.d[6]); break;
      case GI_SHAPE_CONE: fprintf(f, "cone"); p3(f, p.d); p3(f, p.d + 3); fprintf(f, " %.17g\n", p.d[6]); break;
      case GI_SHAPE_CIRCLE: fprintf(f, "circle"); p3(f, p.d); p3(f, p.d + 3); fprintf(f, " %.17g\n", p.d[6]); break;
      case GI_SHAPE_TRIANGLE:
        if (!s.shape_is_mesh[sid]) { fprintf(f, "tri"); p3(f, p.d); p3(f, p.d + 3); p3(f, p.d + 6); fprintf(f, "\n"); }
        else {
          fprintf(f, "mesh %d\n", (int)(end - pi));
          for (size_t t = pi; t < end; t++) { fprintf(f, " t"); p3(f, s.prims[t].d); p3(f, s.prims[t].d + 3); p3(f, s.prims[t].d + 6); fprintf(f, "\n"); }
        }
        break;
    }
    pi = end;
  }
  fprintf(f, "lights %d\n", (int)s.lights.size());
  for (size_t i = 0; i < s.lights.size(); i++) {
    const GiLight &l = s.lights[i];
    fprintf(f, "light %d ", (int)i);
    switch (l.type) {
      case GI_LIGHT_DIR: fprintf(f, "dir"); p3(f, l.color); fprintf(f, " %.17g", l.intensity); p3(f, l.direction); break;
      case GI_LIGHT_POINT:
        fprintf(f, "point"); p3(f, l.color); fprintf(f, " %.17g", l.intensity); p3(f, l.position);
        fprintf(f, " %.17g %.17g %.17g", l.ca, l.la, l.qa); break;
      case GI_LIGHT_SPOT:
        fprintf(f, "spot"); p3(f, l.color); fprintf(f, " %.17g", l.intensity); p3(f, l.position); p3(f, l.direction);
        fprintf(f, " %.17g %.17g %.17g %.17g %.17g", l.ca, l.la, l.qa, l.dropoff, l.cutoff); break;
      case GI_LIGHT_AREA:
        fprintf(f, "area"); p3(f, l.color); fprintf(f, " %.17g", l.intensity); p3(f, l.position); p3(f, l.direction);
        fprintf(f, " %.17g %.17g %.17g %.17g", l.radius, l.ca, l.la, l.qa); break;
      case GI_LIGHT_RECT:
        fprintf(f, "rect"); p3(f, l.color); fprintf(f, " %.17g", l.intensity); p3(f, l.position); p3(f, l.direction);
        p3(f, l.axis1); p3(f, l.axis2);
        fprintf(f, " %.17g %.17g %.17g %.17g %.17g", l.len1, l.len2, l.ca, l.la, l.qa); break;
    }
    fprintf(f, "\n");
  }
  const GiCamera &c = s.camera;
  fprintf(f, "camera"); p3(f, c.eye); p3(f, c.towards); p3(f, c.up); p3(f, c.right);
  fprintf(f, " %.17g %.17g %.17g %.17g\n", c.xfov, c.yfov, c.neardist, c.fardist);
  fprintf(f, "ambient"); p3(f, s.ambient); fprintf(f, "\nbackground"); p3(f, s.background);
  fprintf(f, "\nbbox"); p3(f, s.bbox_min); p3(f, s.bbox_max);
  fprintf(f, "\ncentroid"); p3(f, s.centroid);
  fprintf(f, "\nradius %.17g\n", s.radius);
  fclose(f);
  return true;
}

}  // namespace gi
