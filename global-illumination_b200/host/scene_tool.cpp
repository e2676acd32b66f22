// host/scene_tool.cpp -- small host-only utility around the .scn/.off loader (no CUDA needed):
//   gi_scene_tool dump   in.scn out.txt [-real]    text dump in the format of `ref_harness dump_scene`
//   gi_scene_tool flatten in.scn out.gisc [-real]  binary flattened scene (scene_io.cpp)
#include <cstdio>
#include <cstring>
#include "scene.h"

int main(int argc, char **argv) {
  if (argc < 4) { fprintf(stderr, "usage: gi_scene_tool <dump|flatten> in.scn out [-real]\n"); return 2; }
  bool real = argc > 4 && !strcmp(argv[4], "-real");
  gi::SceneData s; std::string err;
  if (!gi::load_scene(argv[2], real, s, err)) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
  bool ok = !strcmp(argv[1], "dump") ? gi::dump_scene(s, argv[3]) : gi::save_scene_binary(s, argv[3]);
  if (!ok) { fprintf(stderr, "Unable to write %s\n", argv[3]); return 1; }
  return 0;
}
