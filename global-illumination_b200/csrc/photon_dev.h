// csrc/photon_dev.h -- device layout of one photon map (GLOBAL_PMAP / CAUSTIC_PMAP of the reference).
//
// The reference keeps `Photon*` arrays and a pointer kd-tree with 31-photon leaf buckets
// (R3Kdtree.cpp:1624-1671).  Here a map is ONE left-balanced kd-tree in Jensen's heap layout (node i has
// children 2i+1 and 2i+2, every node is a photon), stored as structure-of-arrays and split in two streams:
//   search stream : px/py/pz (FP64, exact distances) + node[] = {float split, axis}  (what traversal touches)
//   shade stream  : rgbe + direction code + original index                            (read for the k winners only)
// In heap layout the subtree below node s occupies, at relative depth j, the contiguous run
// [(s << j) + 2^j - 1, ...) of 2^j slots, so the bottom GI_TILE_LEVELS levels of the tree are read by a warp as a
// handful of coalesced runs ("node tiles") instead of node by node.
#pragma once
#include <stdint.h>
#if defined(__CUDACC__)
#include <cuda_runtime.h>
#endif

namespace gi {

#define GI_TILE_LEVELS_MIN 6      /* a tile = the bottom t levels below a node, t chosen per launch in [6, 9]: */
#define GI_TILE_LEVELS_MAX 9      /* a subtree of up to 2^t - 1 photons read in (2^t - 1) / 32 warp-wide rounds  */
#define GI_MAX_K 512              /* largest estimate size (-gs/-cs) the gather kernel supports */

struct PhotonMapDev {
  uint64_t n;                     /* photons */
  int32_t levels;                 /* floor(log2 n) + 1 */
  const double *px, *py, *pz;     /* heap order */
  const uint2 *node;              /* heap order: x = float bits of the split coordinate, y = split axis (0..2) */
  const double2 *inode;           /* first n_inode heap slots again, packed for the warp-uniform walk of the upper    */
  uint64_t n_inode;               /* levels: 2 x double2 = {x, y}, {z, split float bits | axis << 32}: one 32 B sector */
  const uint2 *shade;             /* heap order: x = rgbe bytes (r | g<<8 | b<<16 | e<<24), y = direction code */
  const uint32_t *orig;           /* heap order -> index in the order the photons were loaded / traced */
};

struct PhotonMapStore {
  PhotonMapDev dev{};             /* device pointers (owned) */
  void *raw = nullptr;            /* GiPhoton[n] in load order (device), kept for dump / get */
  bool built = false;
  bool cached = false;             /* colours hold cached irradiance (-cache pre-pass, irradiance_cache.cu) */
  double cache_ms = 0;             /* wall time of that pre-pass */
  double bbox[6] = {0, 0, 0, 0, 0, 0};
};

}  // namespace gi
