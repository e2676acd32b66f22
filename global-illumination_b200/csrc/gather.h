// csrc/gather.h -- see gather.cu
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "gi_b200.h"
#include "photon_dev.h"
struct GiContext;
namespace gi {
#define GI_GATHER_MAX_WARPS 24     /* warps (= concurrent queries) per CTA, one CTA per SM (85 registers) */
#define GI_GATHER_CHUNK 16         /* consecutive queries handled by one warp (coherent: neighbouring pixels) */
struct GatherArgs {
  int k;                           /* estimate size (<= GI_MAX_K) */
  int cap;                         /* candidate slots per warp: a multiple of 32, > k + 32; a power of two if indices are wanted */
  int filter;                      /* GiFilter */
  int max_warps;                   /* test / tuning hook: at most this many warps per CTA (0 = as many as fit) */
  int range;                       /* chunks a CTA fetches at a time from the launch-wide counter (0 = chosen by launch_gather) */
  int key_rows;                    /* set by launch_gather: group rows above the bottom-parent row */
  int slack;                       /* the candidate buffer is cut back to k once it holds more than k + slack */
  double guess_factor;             /* next query's first search radius^2 = this factor x the last k-th d^2 */
  double r2;                       /* estimate_dist^2 */
  double fa, fb, fk;               /* FILTER_CONST_A / _B / _K */
};
/* d_order: optional permutation of [0, nq) in which the queries are processed (results stay at their own index) */
cudaError_t launch_gather(GiContext *ctx, const PhotonMapDev &m, const double *d_lut, const GiGatherQuery *d_q,
                          const uint32_t *d_order, uint64_t nq, const GatherArgs &ga, int32_t *d_idx, int32_t *d_count, double *d_rgb,
                          cudaStream_t st);
/* query_order.cu: Morton order of a device batch of queries through the map's bounding box */
struct PhotonMapStore;
const uint32_t *morton_order(GiContext *ctx, const PhotonMapStore &m, const GiGatherQuery *d_q, uint64_t n, cudaStream_t st, bool *failed);
/* pre-sizes the per-stream scratch of gi_gather_batch_device for batches of up to n queries (query_order.cu) */
bool gather_reserve(GiContext *ctx, uint64_t n, cudaStream_t st);
}  // namespace gi
