// csrc/kdtree_build.h -- see kdtree_build.cu
#pragma once
#include <cuda_runtime.h>
#include "photon_dev.h"
struct GiContext;
namespace gi {
// builds m.dev from m.raw (GiPhoton[n] on the device); synchronises ctx->stream
cudaError_t build_photon_map(GiContext *ctx, PhotonMapStore &m, uint64_t n, const double *d_lut);   /* d_lut: the 65536 x 3 direction table */
}  // namespace gi
