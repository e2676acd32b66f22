// csrc/wave.cuh -- the parked-ray queues of the wavefront stages and the closest-hit kernel that serves them: shared by
// the render's secondary paths (render.cu) and the photon walks (photon_trace.cu).
#pragma once
#include "isect.cuh"

namespace gi {

struct __align__(16) PathRec {       // 128 B: one parked path
  double o[3], d[3], scale[3], tw[3];
  uint32_t c0, c1, c2, c3;           // Philox counter of the path's stream (key = seed)
  uint32_t pix, item;
  int32_t iter; uint32_t pad;
};
struct __align__(16) WaveHit {       // 64 B: closest hit of the parked ray in the same slot
  double p[3], n[3];
  int32_t element, prim;             // element < 0: miss
  double t;
};
enum { CLASS_SPEC = 0, CLASS_DIFF = 1 };
struct PathWave {                    // one buffer, two queues: class SPEC grows from the front, class DIFF from the back
  PathRec *rec;
  WaveHit *hit;
  unsigned int *count;               // [2], per class; never exceeds cap in total (a path parks at most once per pass)
  uint32_t cap;
};

// slot of entry t of the two queues taken as one sequence: the SPEC entries, then the DIFF entries
GI_DEV unsigned wave_slot(unsigned t, unsigned n_spec, unsigned cap) { return t < n_spec ? t : cap - 1u - (t - n_spec); }

// Closest hit of every parked ray (both classes).  Nothing but BVH traversal and the FP64 primitive tests: the code all
// warps of the SM are in fits the instruction cache.  Persistent grid.
template <int MINB>
__global__ void __launch_bounds__(128, MINB) wave_hit_kernel(DevScene sc, PathWave cur) {
  const unsigned n_spec = cur.count[CLASS_SPEC], n = n_spec + cur.count[CLASS_DIFF];
  for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
    const unsigned slot = wave_slot(t, n_spec, cur.cap);
    const double4 *i = (const double4 *)(cur.rec + slot);
    const double4 a = i[0]; const double2 b = *(const double2 *)(i + 1);
    RayD ray; ray.o = mk3(a.x, a.y, a.z); ray.d = mk3(a.w, b.x, b.y);
    HitRec h;
    closest_hit(sc, ray, h);
    double4 *o = (double4 *)(cur.hit + slot);
    o[0] = make_double4(h.p.x, h.p.y, h.p.z, h.n.x);
    o[1] = make_double4(h.n.y, h.n.z, __hiloint2double(h.prim, h.prim < 0 ? -1 : h.element), h.t);
  }
}

}  // namespace gi
