// dev check: Philox streams on the device vs the host
#include <cstdio>
#include "rng.cuh"
using namespace gi;
__global__ void k(uint32_t *out) {
  const int s = threadIdx.x;
  Rng r; rng_init(r, 1000, 5, 2, s);
  out[s * 16 + 0] = r.c0; out[s * 16 + 1] = r.c1; out[s * 16 + 2] = r.c2; out[s * 16 + 3] = r.c3;
  for (int j = 0; j < 8; j++) out[s * 16 + 4 + j] = rng_u32(r);
}
int main() {
  uint32_t *d; cudaMalloc(&d, 4 * 16 * 32);
  k<<<1, 32>>>(d);
  uint32_t h[16 * 32]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  for (int s = 0; s < 3; s++) { for (int j = 0; j < 12; j++) printf("%08x ", h[s * 16 + j]); printf("\n"); }
}
