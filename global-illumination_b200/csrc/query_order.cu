// csrc/query_order.cu -- Morton order of a batch of gather queries.
//
// The reference estimates radiance wherever its recursion happens to be (EstimateRadiance, photon_utils.cpp:72);
// a batch handed to gi_gather_batch can be in any order (scanlines, path-tracing requests).  The gather kernel
// (gather.cu) is bound by the latency of its photon loads, and what hides that latency is L1: the 8 warps of a CTA
// and the 16 consecutive queries of a warp must touch the same groups of the kd-tree.  So every batch is processed
// along a Morton curve through the photon map's bounding box: 30-bit key per query, one radix sort of (key, index);
// the kernel follows the permutation, the records themselves are not moved and results stay at the caller's index.
#include <cub/device/device_radix_sort.cuh>
#include <cuda_runtime.h>
#include "context.h"
#include "gather.h"

namespace gi {

namespace {
__device__ __forceinline__ uint32_t spread10(uint32_t v) {
  v &= 1023u; v = (v | (v << 16)) & 0x030000FFu; v = (v | (v << 8)) & 0x0300F00Fu; v = (v | (v << 4)) & 0x030C30C3u;
  return (v | (v << 2)) & 0x09249249u;
}
__global__ void __launch_bounds__(256) query_keys_kernel(const GiGatherQuery *__restrict__ q, uint64_t n, double lx, double ly, double lz,
                                                         double ix, double iy, double iz, uint32_t *__restrict__ keys,
                                                         uint32_t *__restrict__ order) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const double x = (q[t].p[0] - lx) * ix, y = (q[t].p[1] - ly) * iy, z = (q[t].p[2] - lz) * iz;
  const uint32_t xi = (uint32_t)fmin(fmax(x * 1024.0, 0.0), 1023.0), yi = (uint32_t)fmin(fmax(y * 1024.0, 0.0), 1023.0),
                 zi = (uint32_t)fmin(fmax(z * 1024.0, 0.0), 1023.0);                 // NaN -> 0
  keys[t] = spread10(xi) | (spread10(yi) << 1) | (spread10(zi) << 2);
  order[t] = (uint32_t)t;
}
}  // namespace

// Returns the device permutation (owned by the context, valid until the next call on the same stream slot), or
// nullptr with *failed = true on an allocation / launch error.
const uint32_t *morton_order(GiContext *ctx, const PhotonMapStore &m, const GiGatherQuery *d_q, uint64_t n, cudaStream_t st,
                             bool *failed) {
  *failed = false;
  if (n > 0x7fffffffull) { *failed = true; set_error("gi_gather_batch: more than 2^31 - 1 queries in one device batch"); return nullptr; }
  const int slot = st == ctx->pipe[0] ? 0 : (st == ctx->pipe[1] ? 1 : 2);
  GiScratch *sc = ctx->order_ws[slot];
  size_t tmp_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, (uint32_t *)nullptr, (uint32_t *)nullptr, (uint32_t *)nullptr, (uint32_t *)nullptr,
                                  (int)n, 0, 30, st);
  for (int i = 0; i < 4; i++)
    if (!sc[i].fit(4 * n)) { *failed = true; set_error("gi_gather_batch: out of device memory (query order)"); return nullptr; }
  if (!sc[4].fit(tmp_bytes)) { *failed = true; set_error("gi_gather_batch: out of device memory (query order)"); return nullptr; }
  uint32_t *keys = (uint32_t *)sc[0].p, *keys2 = (uint32_t *)sc[1].p, *ord = (uint32_t *)sc[2].p, *ord2 = (uint32_t *)sc[3].p;
  const double *b = m.bbox;
  const double ix = 1.0 / fmax(b[3] - b[0], 1e-300), iy = 1.0 / fmax(b[4] - b[1], 1e-300), iz = 1.0 / fmax(b[5] - b[2], 1e-300);
  query_keys_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d_q, n, b[0], b[1], b[2], ix, iy, iz, keys, ord);
  size_t bytes = sc[4].cap;
  if (!cuda_ok(cub::DeviceRadixSort::SortPairs(sc[4].p, bytes, keys, keys2, ord, ord2, (int)n, 0, 30, st), "sort queries")) {
    *failed = true; return nullptr;
  }
  ctx->launches += 2;            // the key kernel + the sort (cub's own passes are not counted one by one)
  return ord2;
}

// Sizes the ordering scratch of stream `st` and the small-map scratch for device batches of up to n queries, so that a
// frame never reallocates (cudaFree / cudaMalloc synchronise the device) in the middle of its batches.
bool gather_reserve(GiContext *ctx, uint64_t n, cudaStream_t st) {
  const int slot = st == ctx->pipe[0] ? 0 : (st == ctx->pipe[1] ? 1 : 2);
  GiScratch *sc = ctx->order_ws[slot];
  size_t tmp_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, (uint32_t *)nullptr, (uint32_t *)nullptr, (uint32_t *)nullptr, (uint32_t *)nullptr,
                                  (int)n, 0, 30, st);
  for (int i = 0; i < 4; i++) if (!sc[i].fit(4 * n)) return false;
  if (!sc[4].fit(tmp_bytes)) return false;
  const uint64_t chunks = (n + 31) / 32;
  return ctx->small_ws[slot].fit(((48 * chunks + 255) / 256) * 256 + chunks * 256);
}

}  // namespace gi
