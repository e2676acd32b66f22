// host/capi_host.cpp -- C-ABI wrappers around the host-side scene loader, so that non-C++ callers (the Python
// tests and bench through ctypes) reach the same .scn/.off loader as the `photonmap` CLI.  Stands in for
// ReadScene (src/utils/io_utils.cpp:219-250) -> R3Scene::ReadFile (R3Scene.cpp:514-587).
#include <cstdlib>
#include <cstring>
#include <string>
#include "image_io.h"
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

// ParseArgs (src/utils/io_utils.cpp:16-214): same flags, same clamping, same order dependence.  Extra flags of this
// implementation: -seed N, -device N, -gpus N.  Returns 0 on success; an unknown flag or a missing file name is an error.
int gi_parse_args(int argc, const char *const *argv, GiParams *p, GiCliOptions *o) {
  if (!p || !o || (argc > 0 && !argv)) { gi::set_error("gi_parse_args: null argument"); return 1; }
  int i = 1;
  auto next = [&](const char *&out) { if (i + 1 >= argc) { gi::set_error(std::string("Missing value after ") + argv[i]); return false; } out = argv[++i]; return true; };
  for (; i < argc; i++) {
    const char *a = argv[i], *v = nullptr;
    if (a[0] != '-') {
      if (!o->input_scene_name) o->input_scene_name = a;
      else if (!o->output_image_name) o->output_image_name = a;
      else { gi::set_error(std::string("Invalid program argument: ") + a); return 1; }
      continue;
    }
    auto is = [&](const char *f) { return !strcmp(a, f); };
    if (is("-v")) p->verbose = 1;
    else if (is("-threads")) { if (!next(v)) return 1; p->threads = (int)atof(v); if (p->threads <= 0) p->threads = 1; }
    else if (is("-aa")) { if (!next(v)) return 1; o->aa = atoi(v); if (o->aa < 0) o->aa *= -1; }
    else if (is("-real")) o->real_material = 1;
    else if (is("-no_fresnel")) p->fresnel = 0;
    else if (is("-ir")) { if (!next(v)) return 1; p->ir_air = atof(v); if (p->ir_air <= 0) p->ir_air = 1.0E-6; }
    else if (is("-no_ambient")) p->ambient = 0;
    else if (is("-no_direct")) p->direct_illum = 0;
    else if (is("-no_transmissive")) p->transmissive_illum = 0;
    else if (is("-no_specular")) p->specular_illum = 0;
    else if (is("-no_indirect")) p->indirect_illum = 0;
    else if (is("-no_caustic")) p->caustic_illum = 0;
    else if (is("-photon_viz")) p->direct_photon_illum = 1;
    else if (is("-fast_global")) { p->fast_global = 1; p->direct_photon_illum = 1; p->indirect_illum = 0; }
    else if (is("-cache")) p->irradiance_cache = 1;
    else if (is("-no_monte")) p->monte_carlo = 0;
    else if (is("-md")) { if (!next(v)) return 1; p->max_monte_depth = atoi(v); if (p->max_monte_depth < 1) p->max_monte_depth = 1; }
    else if (is("-absorb")) { if (!next(v)) return 1; p->prob_absorb = atof(v); if (p->prob_absorb < 0) p->prob_absorb = 0; }
    else if (is("-no_rs")) p->recursive_shadows = 0;
    else if (is("-no_dt")) p->distrib_transmissive = 0;
    else if (is("-tt")) { if (!next(v)) return 1; p->transmissive_test = atoi(v); if (p->transmissive_test < 1) p->transmissive_test = 1; }
    else if (is("-no_ds")) p->distrib_specular = 0;
    else if (is("-st")) { if (!next(v)) return 1; p->specular_test = atoi(v); if (p->specular_test < 1) p->specular_test = 1; }
    else if (is("-global")) { if (!next(v)) return 1; p->global_photon_count = atoi(v); if (p->global_photon_count < 1) p->global_photon_count = 1; }
    else if (is("-caustic")) { if (!next(v)) return 1; p->caustic_photon_count = atoi(v); if (p->caustic_photon_count < 1) p->caustic_photon_count = 1; }
    else if (is("-pd")) { if (!next(v)) return 1; p->max_photon_depth = atoi(v); if (p->max_photon_depth < 1) p->max_photon_depth = 1; }
    else if (is("-it")) { if (!next(v)) return 1; p->indirect_test = atoi(v); if (p->indirect_test < 1) p->indirect_test = 1; }
    else if (is("-gs")) { if (!next(v)) return 1; p->global_estimate_size = atoi(v); if (p->global_estimate_size < 1) p->global_estimate_size = 1; }
    else if (is("-gd")) { if (!next(v)) return 1; p->global_estimate_dist = atof(v); if (p->global_estimate_dist < 0.0) p->global_estimate_dist = 1.0E-6; }
    else if (is("-cs")) { if (!next(v)) return 1; p->caustic_estimate_size = atoi(v); if (p->caustic_estimate_size < 1) p->caustic_estimate_size = 1; }
    else if (is("-cd")) { if (!next(v)) return 1; p->caustic_estimate_dist = atof(v); if (p->caustic_estimate_dist < 0.0) p->caustic_estimate_dist = 1.0E-6; }
    else if (is("-gf") || is("-cf")) {
      const bool global = is("-gf");
      if (!next(v)) return 1;
      if (!strcmp(v, "cone")) {
        (global ? p->global_filter : p->caustic_filter) = GI_FILTER_CONE;
        if (!next(v)) return 1;
        p->filter_const_k = atof(v);
        if (p->filter_const_k < 1) p->filter_const_k = 1;
      } else if (!strcmp(v, "gauss")) (global ? p->global_filter : p->caustic_filter) = GI_FILTER_GAUSS;
    }
    else if (is("-no_shadow")) p->shadows = 0;
    else if (is("-no_ss")) p->soft_shadows = 0;
    else if (is("-lt")) { if (!next(v)) return 1; p->light_test = atoi(v); if (p->light_test < 1) p->light_test = 1; }
    else if (is("-ss")) { if (!next(v)) return 1; p->shadow_test = atoi(v); if (p->shadow_test < 0) p->shadow_test = 0; }
    else if (is("-dof")) {
      p->depth_of_field = 1;
      if (!next(v)) return 1; p->dof_test = atoi(v);
      if (!next(v)) return 1; p->focus_depth = atof(v);
      if (!next(v)) return 1; p->aperture_radius = atof(v);
      if (p->dof_test < 1) p->dof_test = 1;
      if (p->focus_depth < 1.0E-6) p->focus_depth = 1.0E-6;
      if (p->aperture_radius <= 0) p->aperture_radius = 1.0E-6;
    }
    else if (is("-resolution")) {
      if (!next(v)) return 1; o->width = atoi(v);
      if (!next(v)) return 1; o->height = atoi(v);
      if (o->width < 0) o->width *= -1;
      if (o->height < 0) o->height *= -1;
    }
    else if (is("-seed")) { if (!next(v)) return 1; p->seed = strtoull(v, nullptr, 10); }
    else if (is("-device")) { if (!next(v)) return 1; o->device = atoi(v); }
    else if (is("-gpus")) { if (!next(v)) return 1; o->gpus = atoi(v); if (o->gpus < 1) o->gpus = 1; }
    else { gi::set_error(std::string("Invalid program argument: ") + a); return 1; }
  }
  if (!o->input_scene_name || !o->output_image_name) { gi::set_error("Usage: photonmap inputscenefile outputimagefile [-FLAGS]"); return 1; }
  return 0;
}

// R2Image::Write (R2Image.cpp:317-340) for the frame gi_render returns
int gi_write_image(const char *path, const uint8_t *rgb, int width, int height) {
  if (!path || !rgb || width < 1 || height < 1) { gi::set_error("gi_write_image: bad argument"); return 1; }
  std::string err;
  if (!gi::write_image(path, rgb, width, height, err)) { gi::set_error(err); return 1; }
  return 0;
}

void gi_default_cli_options(GiCliOptions *o) {       // src/photonmap.cpp:22-34
  if (!o) return;
  memset(o, 0, sizeof(*o));
  o->width = 1024; o->height = 1024; o->aa = 2; o->real_material = 0; o->device = 0; o->gpus = 1;
}

}  // extern "C"
