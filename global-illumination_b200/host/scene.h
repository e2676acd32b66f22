// host/scene.h -- owning container for the flattened scene handed to gi_upload_scene().
//
// Host side of the drop-in boundary: our own .scn/.off loader reproducing the semantics of the
// reference's ReadPrinceton (src/R3Graphics/R3Scene.cpp:1463-1953) and R3Mesh::ReadOffFile
// (src/R3Shapes/R3Mesh.cpp:4075-4215), but producing flat POD arrays instead of an object graph.
#pragma once
#include <string>
#include <vector>
#include "gi_b200.h"

namespace gi {

struct SceneData {
  std::vector<GiNode> nodes;
  std::vector<GiElement> elements;
  std::vector<GiMaterial> materials;
  std::vector<GiPrim> prims;
  std::vector<GiShape> shapes;
  std::vector<GiLight> lights;
  int n_shapes = 0;
  std::vector<uint8_t> shape_is_mesh;  // per reference shape id: came from a `mesh` command
  GiCamera camera{};
  bool has_camera = false;
  double ambient[3] = {0, 0, 0};
  double background[3] = {0, 0, 0};
  double bbox_min[3] = {0, 0, 0}, bbox_max[3] = {0, 0, 0};
  double centroid[3] = {0, 0, 0};
  double radius = 0;

  GiScene view() const;                 // non-owning GiScene pointing into the vectors
};

// Reads a Princeton .scn file (and the .off meshes / included .scn files it names).  On failure
// returns false and sets `err` to the message the reference would print to stderr.
bool load_scene(const std::string &path, bool real_material, SceneData &out, std::string &err);

// Binary (GISC) serialisation of the flattened scene, see scene_io.cpp.
bool save_scene_binary(const SceneData &s, const std::string &path);
bool load_scene_binary(const std::string &path, SceneData &out, std::string &err);

// Text dump in the format of `oracle/_ref/ref_harness dump_scene` (loader parity tests diff the two).
bool dump_scene(const SceneData &s, const std::string &path);

}  // namespace gi
