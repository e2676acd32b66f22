// csrc/stage_clock.h -- per-stage device times of gi_trace_photons / gi_render (gi_stage_times, gi_b200.h).
// Off by default.  When enabled, every stage of a frame is bracketed by a pair of CUDA events on the context's stream;
// pairs are resolved (cudaEventElapsedTime) at the host synchronisations the drivers make anyway, so the clock adds no
// synchronisation of its own.  Events are pooled and reused.
#pragma once
#include <cuda_runtime.h>
#include <vector>
#include "gi_b200.h"

namespace gi {

enum Stage {
  ST_EMIT_FILTER = 0, ST_TRACE, ST_PHOTON_EXCHANGE, ST_PHOTON_SORT, ST_KD_BUILD,
  ST_PRIMARY, ST_DIRECT, ST_PATH_BEGIN, ST_PATH_HIT, ST_PATH_SHADE, ST_PATH_FINISH, ST_GATHER_GLOBAL, ST_GATHER_CAUSTIC, ST_COMMIT,
  ST_RESOLVE, ST_FRAME_EXCHANGE, ST_COUNT
};

struct StageClock {
  bool on = false;
  double ms[ST_COUNT] = {};
  struct Pair { int stage; cudaEvent_t a, b; };
  std::vector<Pair> open;                     // recorded, not yet resolved
  std::vector<cudaEvent_t> pool;
  cudaEvent_t cur_a = nullptr; int cur_stage = -1;
  cudaEvent_t get() {
    if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
    cudaEvent_t e = nullptr; cudaEventCreate(&e); return e;
  }
  void begin(int stage, cudaStream_t st) {
    if (!on) return;
    cur_a = get(); cur_stage = stage; cudaEventRecord(cur_a, st);
  }
  void end(cudaStream_t st) {
    if (!on || cur_stage < 0) return;
    cudaEvent_t b = get(); cudaEventRecord(b, st);
    open.push_back(Pair{cur_stage, cur_a, b}); cur_stage = -1;
  }
  // after a stream synchronisation: the pairs that have completed are resolved (a pair recorded on another stream of the
  // context -- the render's side stream -- may still be running: it stays open until the next call)
  void collect() {
    std::vector<Pair> still;
    for (const Pair &p : open) {
      if (cudaEventQuery(p.b) != cudaSuccess) { still.push_back(p); continue; }
      float t = 0; if (cudaEventElapsedTime(&t, p.a, p.b) == cudaSuccess) ms[p.stage] += t;
      pool.push_back(p.a); pool.push_back(p.b);
    }
    open.swap(still);
  }
  void reset_photon() { for (int i = ST_EMIT_FILTER; i <= ST_KD_BUILD; i++) ms[i] = 0; }
  void reset_render() { for (int i = ST_PRIMARY; i < ST_COUNT; i++) ms[i] = 0; }
  void release() {
    collect();
    for (const Pair &p : open) { pool.push_back(p.a); pool.push_back(p.b); }
    open.clear();
    for (cudaEvent_t e : pool) cudaEventDestroy(e);
    pool.clear();
  }
};

}  // namespace gi
