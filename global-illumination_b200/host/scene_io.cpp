// host/scene_io.cpp -- binary serialisation of the flattened scene ("GISC" files).
//
// Not a reference format: the reference keeps its scene only as an in-memory R3Scene graph.  A GISC
// file is the flattened GiScene of include/gi_b200.h written array by array, so that tests, the oracle
// restatement and the bench can share one scene without re-parsing .scn/.off (and without the
// reference tree, which does not exist on the GPU box).
//   u32 magic "GISC", u32 version, i32 counts[6] = nodes, elements, materials, prims, shapes, lights
//   GiNode[n], GiElement[n], GiMaterial[n], GiPrim[n], GiShape[n], GiLight[n]
//   GiCamera, f64 ambient[3], background[3], bbox_min[3], bbox_max[3], centroid[3], radius
#include <cstdio>
#include <cstring>
#include "scene.h"

namespace gi {

static const uint32_t kMagic = 0x43534947u;  // "GISC"
static const uint32_t kVersion = 2;

template <class T> static bool wr(FILE *f, const std::vector<T> &v) {
  return v.empty() || fwrite(v.data(), sizeof(T), v.size(), f) == v.size();
}
template <class T> static bool rd(FILE *f, std::vector<T> &v, int n) {
  if (n < 0) return false;
  v.resize((size_t)n);
  return n == 0 || fread(v.data(), sizeof(T), (size_t)n, f) == (size_t)n;
}

bool save_scene_binary(const SceneData &s, const std::string &path) {
  FILE *f = fopen(path.c_str(), "wb");
  if (!f) return false;
  uint32_t hdr[2] = {kMagic, kVersion};
  int32_t counts[6] = {(int32_t)s.nodes.size(), (int32_t)s.elements.size(), (int32_t)s.materials.size(),
                       (int32_t)s.prims.size(), (int32_t)s.shapes.size(), (int32_t)s.lights.size()};
  bool ok = fwrite(hdr, 4, 2, f) == 2 && fwrite(counts, 4, 6, f) == 6;
  ok = ok && wr(f, s.nodes) && wr(f, s.elements) && wr(f, s.materials) && wr(f, s.prims) && wr(f, s.shapes) &&
       wr(f, s.lights);
  double tail[16];
  memcpy(tail, s.ambient, 24); memcpy(tail + 3, s.background, 24); memcpy(tail + 6, s.bbox_min, 24);
  memcpy(tail + 9, s.bbox_max, 24); memcpy(tail + 12, s.centroid, 24); tail[15] = s.radius;
  ok = ok && fwrite(&s.camera, sizeof(GiCamera), 1, f) == 1 && fwrite(tail, 8, 16, f) == 16;
  fclose(f);
  return ok;
}

bool load_scene_binary(const std::string &path, SceneData &s, std::string &err) {
  FILE *f = fopen(path.c_str(), "rb");
  if (!f) { err = "Unable to open file " + path; return false; }
  uint32_t hdr[2]; int32_t c[6];
  if (fread(hdr, 4, 2, f) != 2 || hdr[0] != kMagic || hdr[1] != kVersion || fread(c, 4, 6, f) != 6) {
    fclose(f); err = "Bad GISC header in " + path; return false;
  }
  s = SceneData();
  double tail[16];
  bool ok = rd(f, s.nodes, c[0]) && rd(f, s.elements, c[1]) && rd(f, s.materials, c[2]) && rd(f, s.prims, c[3]) &&
            rd(f, s.shapes, c[4]) && rd(f, s.lights, c[5]) && fread(&s.camera, sizeof(GiCamera), 1, f) == 1 &&
            fread(tail, 8, 16, f) == 16;
  fclose(f);
  if (!ok) { err = "Truncated GISC file " + path; return false; }
  memcpy(s.ambient, tail, 24); memcpy(s.background, tail + 3, 24); memcpy(s.bbox_min, tail + 6, 24);
  memcpy(s.bbox_max, tail + 9, 24); memcpy(s.centroid, tail + 12, 24); s.radius = tail[15];
  s.n_shapes = (int)s.shapes.size();
  s.shape_is_mesh.resize(s.shapes.size());
  for (size_t i = 0; i < s.shapes.size(); i++) s.shape_is_mesh[i] = (uint8_t)s.shapes[i].is_mesh;
  s.has_camera = true;
  return true;
}

}  // namespace gi
