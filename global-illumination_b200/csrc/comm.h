// csrc/comm.h -- the NCCL communicator of a context (see comm.cu): where the renderer shards across GPUs.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>
struct GiContext;
namespace gi {
// All functions return false (and set the error string) on failure.  Collectives are enqueued on `st`; every rank of
// the communicator must make the same calls in the same order.
bool comm_attached(const GiContext *ctx);
int comm_rank(const GiContext *ctx);                   // 0 without a communicator
int comm_world(const GiContext *ctx);                  // 1 without a communicator
bool comm_allreduce_sum_u64(GiContext *ctx, uint64_t *d_buf, size_t count, cudaStream_t st);         // in place
bool comm_allgather_u64(GiContext *ctx, const uint64_t *d_send, uint64_t *d_recv, size_t count_per_rank, cudaStream_t st);
// ragged all-gather: rank r contributes bytes_of[r] bytes (d_send on this rank); they land at d_recv + offset_of[r] on
// every rank.  One grouped exchange of point-to-point transfers over NVLink -- no padding to the largest block, no
// staging copy: the caller passes the final buffer (e.g. the photon map's raw array).
bool comm_allgatherv(GiContext *ctx, const void *d_send, void *d_recv, const uint64_t *bytes_of, const uint64_t *offset_of, cudaStream_t st);
// gather to rank 0: rank r's block (bytes_of[r] bytes at d_send) lands at d_recv + offset_of[r] on rank 0 only
bool comm_gatherv_root0(GiContext *ctx, const void *d_send, void *d_recv, const uint64_t *bytes_of, const uint64_t *offset_of, cudaStream_t st);
void comm_release(GiContext *ctx);
}  // namespace gi
