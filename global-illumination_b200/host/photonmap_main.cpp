// host/photonmap_main.cpp -- `photonmap inputscenefile outputimagefile [-FLAGS]`: the reference's program
// (src/photonmap.cpp:442-499) on top of the C ABI of include/gi_b200.h.  Same flags (gi_parse_args = ParseArgs),
// same stages in the same order -- read scene, MapPhotons, RenderImage, write image -- and the same -v report
// blocks; every stage runs on the GPU, and the program fails if there is none.
//
// Extra flags: -seed N (the reference RNG is unseeded), -device N, -gpus N (N GPUs of this box share the photon pass and
// the frame: one context + one host thread per GPU, NCCL between them), -save_global f / -save_caustic f and
// -load_global f / -load_caustic f (photon maps as "GIPM" files, shared with oracle/ref_harness), and the debug views of the
// reference's `visualize` program as image files: -viz_global f / -viz_caustic f / -viz_rays f / -viz_paths f.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <unistd.h>
#include <vector>
#include "gi_b200.h"

static double seconds_since(std::chrono::steady_clock::time_point t0) {
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}
static int fail(const char *what) { fprintf(stderr, "%s: %s\n", what, gi_last_error()); return -1; }

int main(int argc, char **argv) {
  GiParams params; gi_default_params(&params);
  GiCliOptions opt; gi_default_cli_options(&opt);
  const char *save[2] = {nullptr, nullptr}, *load[2] = {nullptr, nullptr};
  const char *viz[4] = {nullptr, nullptr, nullptr, nullptr};        // GiDebugView order
  std::vector<const char *> av;
  for (int i = 0; i < argc; i++) {
    const bool has_value = i + 1 < argc;
    if (has_value && !strcmp(argv[i], "-save_global")) save[0] = argv[++i];
    else if (has_value && !strcmp(argv[i], "-save_caustic")) save[1] = argv[++i];
    else if (has_value && !strcmp(argv[i], "-load_global")) load[0] = argv[++i];
    else if (has_value && !strcmp(argv[i], "-load_caustic")) load[1] = argv[++i];
    else if (has_value && !strcmp(argv[i], "-viz_global")) viz[GI_VIEW_GLOBAL_PHOTONS] = argv[++i];
    else if (has_value && !strcmp(argv[i], "-viz_caustic")) viz[GI_VIEW_CAUSTIC_PHOTONS] = argv[++i];
    else if (has_value && !strcmp(argv[i], "-viz_rays")) viz[GI_VIEW_RAYS] = argv[++i];
    else if (has_value && !strcmp(argv[i], "-viz_paths")) viz[GI_VIEW_PATHS] = argv[++i];
    else av.push_back(argv[i]);
  }
  if (gi_parse_args((int)av.size(), av.data(), &params, &opt)) { fprintf(stderr, "%s\n", gi_last_error()); return 1; }

  // ReadScene (io_utils.cpp:219-250)
  auto t0 = std::chrono::steady_clock::now();
  GiHostScene *host = nullptr;
  if (gi_host_scene_load(opt.input_scene_name, opt.real_material, &host)) return fail("Unable to read scene");
  GiScene scene;
  gi_host_scene_view(host, &scene);
  if (params.verbose) {
    printf("Read scene from %s ...\n", opt.input_scene_name);
    printf("  Time = %.2f seconds\n", seconds_since(t0));
    printf("  # Nodes = %d\n", scene.n_nodes);
    printf("  # Lights = %d\n", scene.n_lights);
    fflush(stdout);
  }

  // one context per GPU; -gpus N: devices device .. device + N - 1 share the photon pass and the frame (gi_b200.h, multi-GPU)
  const int ngpu = opt.gpus > 1 ? opt.gpus : 1;
  std::vector<GiContext *> ctxs(ngpu, nullptr);
  t0 = std::chrono::steady_clock::now();
  for (int r = 0; r < ngpu; r++) {
    if (gi_create(opt.device + r, &ctxs[r])) return fail("photonmap");
    if (gi_upload_scene(ctxs[r], &scene)) return fail("Unable to upload scene");
  }
  if (ngpu > 1 && gi_comm_init_all(ctxs.data(), ngpu)) return fail("Unable to set up the GPU communicator");
  if (params.verbose) {
    printf("Built GPU scene ...\n  Time = %.2f seconds\n  # Primitives = %d\n", seconds_since(t0), scene.n_prims);
    if (ngpu > 1) printf("  # GPUs = %d\n", ngpu);
    fflush(stdout);
  }
  std::vector<uint8_t> rgb((size_t)opt.width * opt.height * 3);
  std::vector<int> status(ngpu, 0);
  // what the reference's main() does between ReadScene and WriteImage, once per GPU; rank 0 reports and owns the image
  auto run_rank = [&](int r) -> int {
    GiContext *ctx = ctxs[r];
    GiParams prm = params;
    const bool talk = prm.verbose && r == 0;
    // MapPhotons (photonmap.cpp:260-436)
    if (prm.indirect_illum || prm.caustic_illum || prm.direct_photon_illum) {
      const auto t0 = std::chrono::steady_clock::now();
      GiPhotonStats ps; memset(&ps, 0, sizeof(ps));
      if (load[0] || load[1]) {
        for (int m = 0; m < 2; m++)
          if (load[m] && gi_load_photons_file(ctx, m, load[m])) return fail("Unable to load photon map");
        const auto tb = std::chrono::steady_clock::now();
        if (gi_build_maps(ctx, &prm)) return fail("Unable to build photon maps");
        ps.build_ms = seconds_since(tb) * 1000.0;
        gi_photon_count(ctx, 0, &ps.global_stored); gi_photon_count(ctx, 1, &ps.caustic_stored);
        if (!ps.global_stored) { prm.indirect_illum = 0; prm.direct_photon_illum = 0; }
        if (!ps.caustic_stored) prm.caustic_illum = 0;
      } else {
        if (talk) {
          if (prm.indirect_illum || prm.direct_photon_illum) printf("Building global photon map ...\n");
          if (prm.caustic_illum) printf("Building caustic photon map ...\n");
        }
        if (gi_trace_photons(ctx, &prm, &ps)) return fail("Unable to trace photons");
        if (scene.n_lights > 0 && ps.total_power > 0) {
          prm.indirect_illum = ps.indirect_illum; prm.caustic_illum = ps.caustic_illum;
          prm.direct_photon_illum = ps.direct_photon_illum;
        }
      }
      if (talk) {
        const bool cache = prm.irradiance_cache && (prm.indirect_illum || prm.direct_photon_illum);
        printf("Building kdtrees ...\n");
        if (cache) printf("Building irradiance cache ...\n");
        printf("Built photon map ...\n");
        printf("  Total Time = %.2f seconds\n", seconds_since(t0));
        printf("  Photon Tracing = %.2f seconds\n", ps.trace_ms / 1000.0);
        printf("  KdTree Construction = %.2f seconds\n", ps.build_ms / 1000.0);
        if (cache) printf("  Irradiance Cache Computation = %.2f seconds\n", ps.cache_ms / 1000.0);
        unsigned long long total = 0;
        if (prm.indirect_illum || prm.direct_photon_illum) { printf("  # Global Photons Stored = %llu\n", (unsigned long long)ps.global_stored); total += ps.global_stored; }
        if (prm.caustic_illum) { printf("  # Caustic Photons Stored = %llu\n", (unsigned long long)ps.caustic_stored); total += ps.caustic_stored; }
        printf("Total Photons Stored: %llu\n", total);
        fflush(stdout);
      }
      for (int m = 0; m < 2; m++)
        if (r == 0 && save[m] && gi_dump_photons_file(ctx, m, save[m])) return fail("Unable to save photon map");
    }

    // the debug views of `visualize` (visualize.cpp:268-298, 589-700, 1366-1424), from rank 0's copy of the maps
    for (int v = 0; v < 4; v++) {
      if (r != 0 || !viz[v]) continue;
      std::vector<uint8_t> view((size_t)opt.width * opt.height * 3);
      if (gi_debug_view(ctx, v, &prm, opt.width, opt.height, 1, view.data())) return fail("Unable to draw debug view");
      if (gi_write_image(viz[v], view.data(), opt.width, opt.height)) return fail("Unable to write debug view");
      if (talk) printf("Wrote debug view to %s ...\n", viz[v]);
    }

    // RenderImage (render.cpp:155-259)
    if (talk) printf("Rendering image ...\n");
    GiRenderStats rs; memset(&rs, 0, sizeof(rs));
    if (gi_render(ctx, &prm, opt.width, opt.height, opt.aa, r == 0 ? rgb.data() : nullptr, &rs)) return fail("Unable to render");
    if (talk) {
      unsigned long long total = rs.screen_rays;
      printf("Rendered image ...\n");
      printf("  Time = %.2f seconds\n", rs.render_ms / 1000.0);
      printf("  # Screen Rays = %llu\n", (unsigned long long)rs.screen_rays);
      if (prm.shadows) { printf("  # Shadow Rays = %llu\n", (unsigned long long)rs.shadow_rays); total += rs.shadow_rays; }
      if (prm.monte_carlo) { printf("  # Monte Carlo Rays = %llu\n", (unsigned long long)rs.monte_rays); total += rs.monte_rays; }
      if (prm.transmissive_illum) { printf("  # Transmissive Samples = %llu\n", (unsigned long long)rs.transmissive_samples); total += rs.transmissive_samples; }
      if (prm.specular_illum) { printf("  # Specular Samples = %llu\n", (unsigned long long)rs.specular_samples); total += rs.specular_samples; }
      if (prm.indirect_illum) { printf("  # Indirect Samples = %llu\n", (unsigned long long)rs.indirect_samples); total += rs.indirect_samples; }
      if (prm.caustic_illum) { printf("  # Caustic Samples = %llu\n", (unsigned long long)rs.caustic_samples); total += rs.caustic_samples; }
      printf("Total Rays: %llu\n", total);
      fflush(stdout);
    }

    return 0;
  };
  if (ngpu == 1) { if (run_rank(0)) return -1; }
  else {
    // NCCL collectives inside gi_trace_photons / gi_render need all ranks in flight: one host thread per GPU
    std::vector<std::thread> th;
    for (int r = 0; r < ngpu; r++)
      th.emplace_back([&, r] { status[r] = run_rank(r); if (status[r]) { fflush(stdout); fflush(stderr); _exit(255); } });   // a dead rank would hang the others
    for (auto &t : th) t.join();
  }

  // WriteImage (io_utils.cpp:271-291)
  t0 = std::chrono::steady_clock::now();
  if (gi_write_image(opt.output_image_name, rgb.data(), opt.width, opt.height)) { fprintf(stderr, "%s\n", gi_last_error()); return -1; }
  if (params.verbose) {
    printf("Wrote image to %s ...\n", opt.output_image_name);
    printf("  Time = %.2f seconds\n", seconds_since(t0));
    printf("  Width = %d\n", opt.width);
    printf("  Height = %d\n", opt.height);
    fflush(stdout);
  }
  for (GiContext *c : ctxs) gi_destroy(c);
  gi_host_scene_free(host);
  return 0;
}
