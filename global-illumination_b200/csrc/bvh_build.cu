// csrc/bvh_build.cu -- GPU construction of the scene BVH (K1 of SURVEY.md §2b).
//
// Replaces the reference's implicit hierarchy -- the scene-graph bounding boxes (R3SceneNode.cpp:539-560) plus a
// brute-force loop over every triangle of a mesh (R3Isect.cpp:814-826) -- by one linear BVH over all
// primitives: 30-bit Morton codes of the box centroids, one radix sort, Karras' parallel radix tree
// (one thread per internal node), and a bottom-up refit that emits the 64-byte traversal nodes
// (both child boxes stored in the parent, so a visit is 4 x LDG.128).  Scenes are small (<= ~1.1e5 primitives),
// so the build is latency-bound and off the hot path; cub's device radix sort is used for the one sort.
#include <cub/device/device_radix_sort.cuh>
#include <cuda_runtime.h>
#include <stdint.h>
#include "bvh_build.h"

namespace gi {

__device__ __forceinline__ uint32_t expand_bits10(uint32_t v) {
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}

__global__ void morton_kernel(const float *__restrict__ boxes, int n, float3 lo, float3 inv_ext, uint64_t *keys) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float *b = boxes + 6 * (size_t)i;
  float cx = ((b[0] + b[3]) * 0.5f - lo.x) * inv_ext.x;
  float cy = ((b[1] + b[4]) * 0.5f - lo.y) * inv_ext.y;
  float cz = ((b[2] + b[5]) * 0.5f - lo.z) * inv_ext.z;
  uint32_t x = (uint32_t)fminf(fmaxf(cx * 1024.0f, 0.0f), 1023.0f);
  uint32_t y = (uint32_t)fminf(fmaxf(cy * 1024.0f, 0.0f), 1023.0f);
  uint32_t z = (uint32_t)fminf(fmaxf(cz * 1024.0f, 0.0f), 1023.0f);
  uint32_t code = (expand_bits10(x) << 2) | (expand_bits10(y) << 1) | expand_bits10(z);
  keys[i] = ((uint64_t)code << 32) | (uint32_t)i;      // primitive index in the low word makes every key unique
}

__device__ __forceinline__ int delta(const uint64_t *keys, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  return __clzll((long long)(keys[i] ^ keys[j]));
}

// Karras 2012: one thread per internal node; children >= 0 are internal nodes, < 0 are ~(sorted leaf index)
__global__ void radix_tree_kernel(const uint64_t *__restrict__ keys, int n, int2 *children, int *parent_internal,
                                  int *parent_leaf) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
  int dmin = delta(keys, n, i, i - d);
  int lmax = 2;
  while (delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
  int l = 0;
  for (int t = lmax / 2; t >= 1; t /= 2)
    if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
  int j = i + l * d;
  int dnode = delta(keys, n, i, j);
  int s = 0;
  int t = l;
  do {
    t = (t + 1) >> 1;
    if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
  } while (t > 1);
  int gamma = i + s * d + min(d, 0);
  int lo = min(i, j), hi = max(i, j);
  int left = (lo == gamma) ? ~gamma : gamma;
  int right = (hi == gamma + 1) ? ~(gamma + 1) : gamma + 1;
  children[i] = make_int2(left, right);
  if (left >= 0) parent_internal[left] = i; else parent_leaf[~left] = i;
  if (right >= 0) parent_internal[right] = i; else parent_leaf[~right] = i;
  if (i == 0) parent_internal[0] = -1;
}

__device__ __forceinline__ void load_box(const float *boxes, size_t i, float *b) {
  for (int k = 0; k < 6; k++) b[k] = boxes[6 * i + k];
}

// bottom-up refit: the second thread to reach a node owns it, writes its box and its traversal record
__global__ void refit_kernel(const uint64_t *__restrict__ keys, int n, const int2 *__restrict__ children,
                             const int *__restrict__ parent_internal, const int *__restrict__ parent_leaf,
                             const float *__restrict__ prim_boxes, float *node_boxes, int *flags, float4 *bvh) {
  int leaf = blockIdx.x * blockDim.x + threadIdx.x;
  if (leaf >= n) return;
  int cur = parent_leaf[leaf];
  while (cur >= 0) {
    __threadfence();
    if (atomicAdd(&flags[cur], 1) == 0) return;
    __threadfence();
    int2 ch = children[cur];
    float b0[6], b1[6];
    int id0, id1;
    if (ch.x < 0) { uint32_t p = (uint32_t)keys[~ch.x]; load_box(prim_boxes, p, b0); id0 = ~(int)p; }
    else { load_box(node_boxes, ch.x, b0); id0 = ch.x; }
    if (ch.y < 0) { uint32_t p = (uint32_t)keys[~ch.y]; load_box(prim_boxes, p, b1); id1 = ~(int)p; }
    else { load_box(node_boxes, ch.y, b1); id1 = ch.y; }
    float *nb = node_boxes + 6 * (size_t)cur;
    for (int k = 0; k < 3; k++) { nb[k] = fminf(b0[k], b1[k]); nb[3 + k] = fmaxf(b0[3 + k], b1[3 + k]); }
    float4 *o = bvh + 4 * (size_t)cur;
    o[0] = make_float4(b0[0], b0[1], b0[2], b0[3]);
    o[1] = make_float4(b0[4], b0[5], b1[0], b1[1]);
    o[2] = make_float4(b1[2], b1[3], b1[4], b1[5]);
    o[3] = make_float4(__int_as_float(id0), __int_as_float(id1), 0.0f, 0.0f);
    cur = parent_internal[cur];
  }
}

// Depth of the tree = the longest leaf-to-root chain.  The traversal stacks of isect.cuh hold at most one pending sibling
// per level of the descent, so depth <= GI_STACK means they cannot overflow.  (A radix tree over unique 62-bit keys
// -- 30-bit Morton code + 32-bit primitive index -- splits on a strictly increasing bit position per level, so the depth is
// at most 62 by construction; the check makes the guarantee explicit and gi_upload_scene fails if it were ever broken.)
__global__ void depth_kernel(int n, const int *__restrict__ parent_internal, const int *__restrict__ parent_leaf, int *max_depth) {
  int leaf = blockIdx.x * blockDim.x + threadIdx.x;
  if (leaf >= n) return;
  int d = 0;
  for (int cur = parent_leaf[leaf]; cur >= 0; cur = parent_internal[cur]) d++;
  atomicMax(max_depth, d);
}

__global__ void single_prim_kernel(const float *prim_boxes, float4 *bvh) {
  const float *b = prim_boxes;
  bvh[0] = make_float4(b[0], b[1], b[2], b[3]);
  bvh[1] = make_float4(b[4], b[5], 0.0f, 0.0f);
  bvh[2] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
  bvh[3] = make_float4(__int_as_float(~0), __int_as_float(0x7fffffff), 0.0f, 0.0f);
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { err = e_; goto done; } } while (0)

cudaError_t build_bvh(const float *d_prim_boxes, int n, const float lo[3], const float hi[3], float4 *d_bvh,
                      cudaStream_t stream, int *depth_out) {
  cudaError_t err = cudaSuccess;
  uint64_t *keys_in = nullptr, *keys_out = nullptr;
  int2 *children = nullptr; int *parent_internal = nullptr, *parent_leaf = nullptr, *flags = nullptr;
  float *node_boxes = nullptr; void *tmp = nullptr; size_t tmp_bytes = 0;
  if (depth_out) *depth_out = n > 0 ? 1 : 0;
  if (n <= 0) return cudaSuccess;
  if (n == 1) {
    single_prim_kernel<<<1, 1, 0, stream>>>(d_prim_boxes, d_bvh);
    return cudaGetLastError();
  }
  {
    float3 l = make_float3(lo[0], lo[1], lo[2]);
    float3 ie = make_float3(hi[0] > lo[0] ? 1.0f / (hi[0] - lo[0]) : 0.0f, hi[1] > lo[1] ? 1.0f / (hi[1] - lo[1]) : 0.0f,
                            hi[2] > lo[2] ? 1.0f / (hi[2] - lo[2]) : 0.0f);
    const int B = 256, G = (n + B - 1) / B;
    CK(cudaMalloc(&keys_in, sizeof(uint64_t) * n));
    CK(cudaMalloc(&keys_out, sizeof(uint64_t) * n));
    CK(cudaMalloc(&children, sizeof(int2) * n));
    CK(cudaMalloc(&parent_internal, sizeof(int) * n));
    CK(cudaMalloc(&parent_leaf, sizeof(int) * n));
    CK(cudaMalloc(&flags, sizeof(int) * n));
    CK(cudaMalloc(&node_boxes, sizeof(float) * 6 * n));
    CK(cudaMemsetAsync(flags, 0, sizeof(int) * n, stream));
    morton_kernel<<<G, B, 0, stream>>>(d_prim_boxes, n, l, ie, keys_in);
    CK(cudaGetLastError());
    CK(cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, keys_in, keys_out, n, 0, 62, stream));
    CK(cudaMalloc(&tmp, tmp_bytes));
    CK(cub::DeviceRadixSort::SortKeys(tmp, tmp_bytes, keys_in, keys_out, n, 0, 62, stream));
    radix_tree_kernel<<<G, B, 0, stream>>>(keys_out, n, children, parent_internal, parent_leaf);
    CK(cudaGetLastError());
    refit_kernel<<<G, B, 0, stream>>>(keys_out, n, children, parent_internal, parent_leaf, d_prim_boxes, node_boxes,
                                      flags, d_bvh);
    CK(cudaGetLastError());
    CK(cudaMemsetAsync(flags, 0, sizeof(int), stream));                 // reused as the depth cell
    depth_kernel<<<G, B, 0, stream>>>(n, parent_internal, parent_leaf, flags);
    CK(cudaGetLastError());
    int depth = 0;
    CK(cudaMemcpyAsync(&depth, flags, sizeof(int), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    if (depth_out) *depth_out = depth;
  }
done:
  cudaFree(keys_in); cudaFree(keys_out); cudaFree(children); cudaFree(parent_internal); cudaFree(parent_leaf);
  cudaFree(flags); cudaFree(node_boxes); cudaFree(tmp);
  return err;
}

}  // namespace gi
