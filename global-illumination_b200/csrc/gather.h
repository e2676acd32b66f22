// csrc/gather.h -- see gather.cu
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "gi_b200.h"
#include "photon_dev.h"
struct GiContext;
namespace gi {
#define GI_GATHER_WARPS 8          /* warps (= concurrent queries) per CTA */
#define GI_GATHER_CHUNK 16         /* consecutive queries handled by one warp (coherent: neighbouring pixels) */
struct GatherArgs {
  int k;                           /* estimate size (<= GI_MAX_K) */
  int cap;                         /* candidate slots per warp: power of two, >= max(256, 2k) */
  int filter;                      /* GiFilter */
  int min_blocks;                  /* 3: 80 registers / 24 warps per SM; 4: 64 registers / 32 warps per SM */
  int slack;                       /* the candidate buffer is cut back to k once it holds more than k + slack */
  double guess_factor;             /* next query's first search radius^2 = this factor x the last k-th d^2 */
  double r2;                       /* estimate_dist^2 */
  double fa, fb, fk;               /* FILTER_CONST_A / _B / _K */
};
cudaError_t launch_gather(GiContext *ctx, const PhotonMapDev &m, const double *d_lut, const GiGatherQuery *d_q,
                          uint64_t nq, const GatherArgs &ga, int32_t *d_idx, int32_t *d_count, double *d_rgb,
                          cudaStream_t st);
}  // namespace gi
