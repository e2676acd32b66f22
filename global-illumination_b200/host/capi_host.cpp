// host/capi_host.cpp -- C-ABI wrappers around the host-side scene loader, so that non-C++ callers (the Python
// tests and bench through ctypes) reach the same .scn/.off loader as the `photonmap` CLI.  Stands in for
// ReadScene (src/utils/io_utils.cpp:219-250) -> R3Scene::ReadFile (R3Scene.cpp:514-587).
#include <cstring>
#include <string>
#include "scene.h"

namespace gi { void set_error(const std::string &msg); }

struct GiHostScene { gi::SceneData data; };

extern "C" {

int gi_host_scene_load(const char *path, int real_material, GiHostScene **out) {
  if (!path || !out) { gi::set_error("gi_host_scene_load: null argument"); return 1; }
  *out = nullptr;
  GiHostScene *h = new GiHostScene();
  std::string err;
  size_t len = strlen(path);
  bool ok = (len > 5 && !strcmp(path + len - 5, ".gisc")) ? gi::load_scene_binary(path, h->data, err)
                                                           : gi::load_scene(path, real_material != 0, h->data, err);
  if (!ok) { gi::set_error(err); delete h; return 1; }
  *out = h;
  return 0;
}

void gi_host_scene_free(GiHostScene *h) { delete h; }

int gi_host_scene_view(const GiHostScene *h, GiScene *out) {
  if (!h || !out) { gi::set_error("gi_host_scene_view: null argument"); return 1; }
  *out = h->data.view();
  return 0;
}

int gi_host_scene_save(const GiHostScene *h, const char *gisc_path) {
  if (!h || !gisc_path) { gi::set_error("gi_host_scene_save: null argument"); return 1; }
  if (!gi::save_scene_binary(h->data, gisc_path)) { gi::set_error(std::string("Unable to write ") + gisc_path); return 1; }
  return 0;
}

// defaults of src/photonmap.cpp:41-106
void gi_default_params(GiParams *p) {
  if (!p) return;
  memset(p, 0, sizeof(*p));
  p->verbose = 0; p->threads = 1; p->fresnel = 1; p->ir_air = 1.0;
  p->ambient = 1; p->direct_illum = 1; p->transmissive_illum = 1; p->specular_illum = 1; p->indirect_illum = 1;
  p->caustic_illum = 1; p->direct_photon_illum = 0; p->fast_global = 0; p->irradiance_cache = 0;
  p->shadows = 1; p->soft_shadows = 1; p->light_test = 128; p->shadow_test = 128;
  p->monte_carlo = 1; p->max_monte_depth = 128; p->prob_absorb = 0.005;
  p->recursive_shadows = 1; p->distrib_transmissive = 1; p->transmissive_test = 128; p->distrib_specular = 1;
  p->specular_test = 128;
  p->depth_of_field = 0; p->dof_test = 1; p->focus_depth = 100.0; p->aperture_radius = 0.025;
  p->global_photon_count = 2176; p->caustic_photon_count = 10000000; p->max_photon_depth = 128;
  p->indirect_test = 256; p->global_estimate_size = 50; p->global_estimate_dist = 2.5; p->global_filter = GI_FILTER_DISK;
  p->caustic_estimate_size = 225; p->caustic_estimate_dist = 0.225; p->caustic_filter = GI_FILTER_DISK;
  p->filter_const_a = 0.918; p->filter_const_b = 1.953; p->filter_const_k = 1.0;
  p->seed = 1; p->rank = 0; p->world_size = 1;
}

}  // extern "C"
