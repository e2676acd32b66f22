// host/scene_tool.cpp -- small host-only utility around the .scn/.off loader (no CUDA needed):
//   gi_scene_tool dump   in.scn out.txt [-real]    text dump in the format of `ref_harness dump_scene`
//   (in.scn may also be a .gisc file written by `flatten`)
//   gi_scene_tool export in.gisc out.scn           Princeton .scn text of a group-free flattened scene
//   gi_scene_tool flatten in.scn out.gisc [-real]  binary flattened scene (scene_io.cpp)
#include <cstdio>
#include <cstring>
#include "scene.h"

// Write a Princeton .scn equivalent to a flattened scene whose only node is the root (cornell, jensen, ...):
// used to hand the SAME scene to the reference binary on machines where the reference's input/ tree is absent.
// Numbers are printed with %.17g, which fscanf("%lf") reads back exactly.
static bool export_scn(const gi::SceneData &s, const char *path) {
  if (s.nodes.size() != 1) { fprintf(stderr, "export: only scenes without groups are supported\n"); return false; }
  for (const GiShape &sh : s.shapes) if (sh.is_mesh) { fprintf(stderr, "export: meshes are not supported\n"); return false; }
  FILE *f = fopen(path, "w");
  if (!f) return false;
  const GiCamera &c = s.camera;
  fprintf(f, "camera %.17g %.17g %.17g  %.17g %.17g %.17g  %.17g %.17g %.17g  %.17g %.17g %.17g\n", c.eye[0], c.eye[1],
          c.eye[2], c.towards[0], c.towards[1], c.towards[2], c.up[0], c.up[1], c.up[2], c.xfov, c.neardist, c.fardist);
  fprintf(f, "ambient %.17g %.17g %.17g\nbackground %.17g %.17g %.17g\n", s.ambient[0], s.ambient[1], s.ambient[2],
          s.background[0], s.background[1], s.background[2]);
  for (const GiLight &l : s.lights) {
    switch (l.type) {
      case GI_LIGHT_DIR: fprintf(f, "dir_light %.17g %.17g %.17g  %.17g %.17g %.17g\n", l.color[0], l.color[1], l.color[2],
                                 l.direction[0], l.direction[1], l.direction[2]); break;
      case GI_LIGHT_POINT: fprintf(f, "point_light %.17g %.17g %.17g  %.17g %.17g %.17g  %.17g %.17g %.17g\n", l.color[0], l.color[1],
                                   l.color[2], l.position[0], l.position[1], l.position[2], l.ca, l.la, l.qa); break;
      case GI_LIGHT_SPOT: fprintf(f, "spot_light %.17g %.17g %.17g  %.17g %.17g %.17g  %.17g %.17g %.17g  %.17g %.17g %.17g  %.17g %.17g\n",
                                  l.color[0], l.color[1], l.color[2], l.position[0], l.position[1], l.position[2], l.direction[0],
                                  l.direction[1], l.direction[2], l.ca, l.la, l.qa, l.cutoff, l.dropoff); break;
      case GI_LIGHT_AREA: fprintf(f, "area_light %.17g %.17g %.17g  %.17g %.17g %.17g  %.17g %.17g %.17g  %.17g  %.17g %.17g %.17g\n",
                                  l.color[0], l.color[1], l.color[2], l.position[0], l.position[1], l.position[2], l.direction[0],
                                  l.direction[1], l.direction[2], l.radius, l.ca, l.la, l.qa); break;
      case GI_LIGHT_RECT: fprintf(f, "rect_light %.17g %.17g %.17g  %.17g %.17g %.17g  %.17g %.17g %.17g  %.17g %.17g %.17g  %.17g %.17g  %.17g %.17g %.17g\n",
                                  l.color[0], l.color[1], l.color[2], l.position[0], l.position[1], l.position[2], l.axis1[0],
                                  l.axis1[1], l.axis1[2], l.axis2[0], l.axis2[1], l.axis2[2], l.len1, l.len2, l.ca, l.la, l.qa); break;
    }
  }
  for (const GiMaterial &m : s.materials)
    fprintf(f, "material %.17g %.17g %.17g  %.17g %.17g %.17g  %.17g %.17g %.17g  %.17g %.17g %.17g  %.17g %.17g %.17g  %.17g %.17g 0\n",
            m.ka[0], m.ka[1], m.ka[2], m.kd[0], m.kd[1], m.kd[2], m.ks[0], m.ks[1], m.ks[2], m.kt[0], m.kt[1], m.kt[2],
            m.ke[0], m.ke[1], m.ke[2], m.shininess, m.ir);
  for (const GiPrim &p : s.prims) {
    const int m = s.elements[p.element].material;
    const double *d = p.d;
    switch (p.type) {
      case GI_SHAPE_TRIANGLE: fprintf(f, "tri %d  %.17g %.17g %.17g  %.17g %.17g %.17g  %.17g %.17g %.17g\n", m, d[0], d[1], d[2], d[3],
                                      d[4], d[5], d[6], d[7], d[8]); break;
      case GI_SHAPE_SPHERE: fprintf(f, "sphere %d  %.17g %.17g %.17g  %.17g\n", m, d[0], d[1], d[2], d[3]); break;
      case GI_SHAPE_BOX: fprintf(f, "box %d  %.17g %.17g %.17g  %.17g %.17g %.17g\n", m, d[0], d[1], d[2], d[3], d[4], d[5]); break;
      default: fclose(f); fprintf(stderr, "export: unsupported primitive type\n"); return false;
    }
  }
  fclose(f);
  return true;
}

int main(int argc, char **argv) {
  if (argc < 4) { fprintf(stderr, "usage: gi_scene_tool <dump|flatten> in.scn out [-real]\n"); return 2; }
  bool real = argc > 4 && !strcmp(argv[4], "-real");
  gi::SceneData s; std::string err;
  size_t len = strlen(argv[2]);
  bool is_gisc = len > 5 && !strcmp(argv[2] + len - 5, ".gisc");
  if (!(is_gisc ? gi::load_scene_binary(argv[2], s, err) : gi::load_scene(argv[2], real, s, err))) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
  bool ok = !strcmp(argv[1], "dump") ? gi::dump_scene(s, argv[3])
            : !strcmp(argv[1], "export") ? export_scn(s, argv[3]) : gi::save_scene_binary(s, argv[3]);
  if (!ok) { fprintf(stderr, "Unable to write %s\n", argv[3]); return 1; }
  return 0;
}
