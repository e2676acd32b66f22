// csrc/photon_dev.h -- device layout of one photon map (GLOBAL_PMAP / CAUSTIC_PMAP of the reference).
//
// The reference keeps `Photon*` arrays and a pointer kd-tree with 31-photon leaf buckets
// (R3Kdtree.cpp:1624-1671).  Here a map is ONE left-balanced kd-tree in Jensen's heap layout (node i has
// children 2i+1 and 2i+2, every node is a photon), stored as structure-of-arrays and split in two streams:
//   search stream : px/py/pz (FP64, exact distances) + node[] = {float split, axis}  (what traversal touches)
//   shade stream  : rgbe + direction code + original index                            (read for the k winners only)
// In heap layout the subtree below node s occupies, at relative depth j, the contiguous run
// [(s << j) + 2^j - 1, ...) of 2^j slots.  The search (gather.cu) therefore reads the tree in GROUPS of five levels:
// the 31 photons of a group are 5 contiguous runs (one warp-wide round of distance tests) and its 32 child groups are
// one contiguous run of slots whose tight FP64 bounding boxes (gbox) are tested by the 32 lanes at once -- the
// binary tree is walked as a 32-ary one.  Rows of group roots sit at tree levels glevel[g].  The bottom row is hb = 5
// or 6 levels deep: the last level of a left-balanced tree is only partly filled, and when it is less than half full
// most 5-level bottom groups would hold 15 photons instead of 31, so the last level is then attached to the row above
// it as a second round of at most 32 photons.  The top group takes the remainder (h0 in [1, 5]).
#pragma once
#include <stdint.h>
#if defined(__CUDACC__)
#include <cuda_runtime.h>
#endif

namespace gi {

#define GI_GROUP_LEVELS 5         /* 2^5 - 1 = 31 photons and 32 children per group: one warp-wide round each */
#define GI_MAX_GROUPS 8           /* rows of groups: 32 tree levels at most (n < 2^32) */
#define GI_MAX_K 512              /* largest estimate size (-gs/-cs) the gather kernel supports */

struct PhotonMapDev {
  uint64_t n;                     /* photons */
  int32_t levels;                 /* floor(log2 n) + 1 */
  const double *px, *py, *pz;     /* heap order */
  const uint2 *node;              /* heap order: x = float bits of the split coordinate, y = split axis (0..2) */
  /* groups (see above), numbered row by row: group of root slot s in row g has index goff[g] + s - (2^glevel[g] - 1) */
  const double *gpos;             /* per group 96 doubles: x[32], y[32], z[32] of its (up to) 31 photons, lane order = slot    */
                                  /* order within the group; unused entries hold +inf, so their distance test fails by itself */
  uint32_t x6_base, x6_blocks;    /* six-level bottom row: level 6 of bottom group number j (j < x6_blocks: left-balanced) is block   */
                                  /* x6_base + j of gpos, again +inf padded                                                     */
  const double *cbox;             /* per group above the bottom row 192 doubles: lo_x[32], hi_x[32], lo_y[32], hi_y[32], lo_z[32],  */
                                  /* hi_z[32] = tight boxes of the subtrees of its 32 child groups (lo = +inf, hi = -inf: no child) */
  int32_t groups, h0, hb;         /* rows of groups; depth of the top group; depth of the bottom-row groups (one row: groups = 1, hb = levels) */
  int32_t glevel[GI_MAX_GROUPS];  /* tree level of the group roots of row g (glevel[0] = 0) */
  uint32_t goff[GI_MAX_GROUPS];   /* index of the first group of row g */
  double *hint;                   /* {estimate size, k-th d2 x guess factor} left behind by the last gather launch on this map: */
                                  /* the first search radius of the next launch's warps (zeroed = none yet)                */
  const uint2 *shade;             /* heap order: x = rgbe bytes (r | g<<8 | b<<16 | e<<24), y = direction code */
  const double2 *srec;            /* heap order, 32 B: {dir.x, dir.y}, {dir.z, bits: rgbe | code << 32}: the estimate's view of a */
                                  /* photon -- the incident direction already looked up (photon_utils.cpp:253-272), one sector  */
  const uint32_t *orig;           /* heap order -> index in the order the photons were loaded / traced */
};

struct PhotonMapStore {
  PhotonMapDev dev{};             /* device pointers (owned) */
  void *raw = nullptr;            /* GiPhoton[n] in load order (device), kept for dump / get */
  void *stream = nullptr;         /* the cudaStream_t whose stream-ordered pool every array of this map came from (free_map) */
  bool built = false;
  bool cached = false;             /* colours hold cached irradiance (-cache pre-pass, irradiance_cache.cu) */
  double cache_ms = 0;             /* wall time of that pre-pass */
  double bbox[6] = {0, 0, 0, 0, 0, 0};
};

}  // namespace gi
