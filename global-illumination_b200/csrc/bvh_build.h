// csrc/bvh_build.h -- see bvh_build.cu
#pragma once
#include <cuda_runtime.h>
namespace gi {
// d_prim_boxes: n x {min(3), max(3)} padded FP32 world boxes; d_bvh: max(n - 1, 1) nodes of 4 float4.
// lo / hi: scene bounds used to quantise the Morton codes.  Synchronises `stream` before returning.
// *depth_out = internal nodes on the longest root-to-leaf chain (what bounds the traversal stack of isect.cuh).
cudaError_t build_bvh(const float *d_prim_boxes, int n, const float lo[3], const float hi[3], float4 *d_bvh,
                      cudaStream_t stream, int *depth_out);
}  // namespace gi
