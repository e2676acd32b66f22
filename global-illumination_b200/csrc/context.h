// csrc/context.h -- the opaque GiContext behind include/gi_b200.h: one CUDA device, one stream, and every
// device allocation of the uploaded scene and photon maps.  The library owns all device memory; callers own
// every host buffer they pass in (see the ownership rules in gi_b200.h).
#pragma once
#include <cuda_runtime.h>
#include <string>
#include <vector>
#include "gi_b200.h"
#include "scene_dev.h"
#include "photon_dev.h"
#include "stage_clock.h"

// grow-only device buffer used by the host-buffer entry points to stage chunks (freed by gi_destroy)
struct GiScratch {
  void *p = nullptr; size_t cap = 0;
  // grows with 50 % headroom: batch sizes drift upwards during a frame, and every cudaFree / cudaMalloc synchronises the
  // whole device (including the render's side stream)
  bool fit(size_t bytes) {
    if (bytes <= cap) return true;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    const size_t want = bytes + bytes / 2;
    if (cudaMalloc(&p, want) != cudaSuccess) {
      p = nullptr;
      if (cudaMalloc(&p, bytes) != cudaSuccess) { p = nullptr; return false; }
      cap = bytes;
      return true;
    }
    cap = want;
    return true;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

struct GiContext {
  int device = 0;
  cudaStream_t stream = nullptr;
  int sm_count = 0;
  bool has_scene = false;
  gi::DevScene scene{};                       // device pointers, passed to kernels by value
  std::vector<void *> scene_allocs;
  // host copy of the flattened scene (lights, camera, materials are needed by the render stages)
  std::vector<GiLight> lights;
  std::vector<GiMaterial> materials;
  GiCamera camera{};
  double ambient[3] = {0, 0, 0}, background[3] = {0, 0, 0}, bbox_min[3] = {0, 0, 0}, bbox_max[3] = {0, 0, 0};
  double centroid[3] = {0, 0, 0}, radius = 0;
  double *d_spec_boxes = nullptr;             // scene_allocs entry: boxes of the specular / transmissive shapes (photon_trace.cu)
  int n_spec_boxes = -1;
  int bvh_depth = 0;                          // longest root-to-leaf chain of the scene BVH (<= GI_STACK, checked at upload)
  gi::PhotonMapStore maps[2];                 // GI_MAP_GLOBAL, GI_MAP_CAUSTIC
  // Host-buffer calls (gi_intersect_batch, gi_gather_batch) run as a two-slot pipeline: while one chunk computes on
  // pipe[0] the next chunk's inputs are copied in and the previous chunk's results copied out on pipe[1].
  cudaStream_t pipe[2] = {nullptr, nullptr};
  GiScratch stage[2][4];                      // per slot: input, and up to three outputs
  GiScratch order_ws[3][5];                   // query_order.cu, per stream slot (pipe[0], pipe[1], any other): keys x2, order x2, sort temp
  size_t gather_smem_configured[4] = {0, 0, 0, 0};  // gather.cu: dynamic shared memory the kernel variants (deep x staged) are configured for on this device
  GiScratch small_ws[3];                      // gather.cu: chunk records + pools of the small-map kernels, per stream slot
  size_t gather_small_smem_configured = 0;     // gather.cu: the same for the small-map kernel
  unsigned char *noalloc_scratch = nullptr;   // gather.cu: device address of its static scratch array on this context's device (cudaGetSymbolAddress, once)
  double *d_lut = nullptr;                    // capi_photon.cu: the 65536 x 3 incident-direction table (photon_utils.cpp:253-272)
  void *gather_counters = nullptr;            // gather.cu: launch-wide chunk counters, one per stream slot
  void *render_ws = nullptr;                  // render.cu's device workspace, created by the first gi_render
  void (*render_ws_free)(void *) = nullptr;
  void *comm = nullptr;                       // comm.cu: NCCL communicator + rank / world of this context (nullptr = single GPU)
  gi::StageClock clock;                        // per-stage device times (gi_stage_timing / gi_stage_times), off by default
  uint64_t queue_overflows = 0;               // render path batches repeated because a gather-request queue ran full
  uint64_t launches = 0;                      // kernels launched by this context (bench.py: gpu_launches)
};

namespace gi {
void set_error(const std::string &msg);
bool cuda_ok(cudaError_t e, const char *what);
void free_scene(GiContext *ctx);
void free_map(PhotonMapStore &m);
bool alloc_raw(GiContext *ctx, PhotonMapStore &m, uint64_t n);   // m.raw = GiPhoton[n] from the stream-ordered pool
}  // namespace gi

#define GI_CUDA(call) do { if (!gi::cuda_ok((call), #call)) return 1; } while (0)
