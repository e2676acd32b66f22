// csrc/irradiance_cache.h -- see irradiance_cache.cu
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "gi_b200.h"
#include "photon_dev.h"
struct GiContext;
namespace gi {
// photonmap.cpp:380-411: converts the colours of the (built) global map to cached irradiance; no-op if already done
int build_irradiance_cache(GiContext *ctx, const GiParams &p);
// EstimateCachedRadiance for nq requests: d_rgb[3 * i ..] = the colour it adds
cudaError_t launch_cached_radiance(GiContext *ctx, const PhotonMapDev &m, const double *d_lut, const GiGatherQuery *d_q, uint64_t nq,
                                   double r, double *d_rgb, cudaStream_t st);
}  // namespace gi
