// csrc/comm.cu -- multi-GPU plumbing of the C ABI: one GiContext per B200, tied together by an NCCL communicator.
//
// The reference is a single process whose worker threads split the photon budget (MapPhotons, photonmap.cpp:294-329) and
// the pixel columns (RenderImage / Threadable_RayTracer, render.cpp:80-90,188-198).  Here a rank = one GPU takes the place
// of a worker thread, and the two exchanges the threads did through shared memory become collectives over NVLink 5 /
// NVSwitch:
//   photon pass : per emission round ONE all-reduce of the stored-photon counts (so every rank follows the same adaptive
//                 schedule, photonmap.cpp:163-192), and at the end ONE ragged all-gather of the photons (+ their sort keys)
//                 straight into the map buffers; the kd-trees are then built on every rank.
//   render      : no communication while rendering (scene and maps are replicated); the finished 8-bit column strips are
//                 gathered to rank 0.
// Two ways to form the communicator, both ending in the same per-context state:
//   gi_comm_init_all(ctxs, n)            one process driving n devices (ncclCommInitAll) -- `photonmap -gpus N`
//   gi_comm_init_rank(ctx, id, r, n)     one process (or thread) per GPU; rank 0 makes the id with gi_comm_unique_id and the
//                                        host hands it to the others (bench.py: torch.distributed broadcast)
// NCCL is bound at run time (dlopen of libnccl.so.2): the library loads and every single-GPU entry point works on a
// machine without NCCL; only the gi_comm_* calls fail there, loudly.
#include <dlfcn.h>
#include <nccl.h>
#include <mutex>
#include <string>
#include <vector>
#include "comm.h"
#include "context.h"

namespace gi {

namespace {

struct NcclApi {
  void *handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
  std::string error;
};

NcclApi *nccl() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    // a process that already holds an NCCL (torch's bundled copy) gets that one: same soname
    for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
      api.handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
      if (api.handle) break;
    }
    if (!api.handle) { api.error = std::string("cannot load libnccl.so.2: ") + dlerror(); return; }
    auto sym = [&](const char *n) { void *p = dlsym(api.handle, n); if (!p && api.error.empty()) api.error = std::string("libnccl misses ") + n; return p; };
    api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
    api.CommInitAll = (decltype(api.CommInitAll))sym("ncclCommInitAll");
    api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
    api.AllReduce = (decltype(api.AllReduce))sym("ncclAllReduce");
    api.AllGather = (decltype(api.AllGather))sym("ncclAllGather");
    api.Send = (decltype(api.Send))sym("ncclSend");
    api.Recv = (decltype(api.Recv))sym("ncclRecv");
    api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
    api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
    api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
  });
  if (!api.error.empty()) { set_error("gi_comm: " + api.error); return nullptr; }
  return &api;
}

bool nccl_ok(NcclApi *a, ncclResult_t r, const char *what) {
  if (r == ncclSuccess) return true;
  set_error(std::string("NCCL error: ") + (a->GetErrorString ? a->GetErrorString(r) : "?") + " in " + what);
  return false;
}

struct Comm { ncclComm_t comm = nullptr; int rank = 0, world = 1; };
Comm *comm_of(const GiContext *ctx) { return (Comm *)ctx->comm; }

}  // namespace

bool comm_attached(const GiContext *ctx) { return ctx->comm != nullptr; }
int comm_rank(const GiContext *ctx) { return ctx->comm ? comm_of(ctx)->rank : 0; }
int comm_world(const GiContext *ctx) { return ctx->comm ? comm_of(ctx)->world : 1; }

void comm_release(GiContext *ctx) {
  if (!ctx->comm) return;
  Comm *c = comm_of(ctx);
  NcclApi *a = nccl();
  if (a && c->comm) a->CommDestroy(c->comm);
  delete c;
  ctx->comm = nullptr;
}

bool comm_allreduce_sum_u64(GiContext *ctx, uint64_t *d_buf, size_t count, cudaStream_t st) {
  if (!ctx->comm) return true;
  NcclApi *a = nccl();
  return a && nccl_ok(a, a->AllReduce(d_buf, d_buf, count, ncclUint64, ncclSum, comm_of(ctx)->comm, st), "ncclAllReduce");
}

bool comm_allgather_u64(GiContext *ctx, const uint64_t *d_send, uint64_t *d_recv, size_t count_per_rank, cudaStream_t st) {
  if (!ctx->comm) return cuda_ok(cudaMemcpyAsync(d_recv, d_send, 8 * count_per_rank, cudaMemcpyDeviceToDevice, st), "allgather(1 rank)");
  NcclApi *a = nccl();
  return a && nccl_ok(a, a->AllGather(d_send, d_recv, count_per_rank, ncclUint64, comm_of(ctx)->comm, st), "ncclAllGather");
}

static bool exchange(GiContext *ctx, const void *d_send, void *d_recv, const uint64_t *bytes_of, const uint64_t *offset_of, cudaStream_t st,
                     bool everyone) {
  const int rank = comm_rank(ctx), world = comm_world(ctx);
  const bool receiver = everyone || rank == 0;
  // my own block: a plain device copy
  if (receiver && bytes_of[rank] &&
      !cuda_ok(cudaMemcpyAsync((char *)d_recv + offset_of[rank], d_send, bytes_of[rank], cudaMemcpyDeviceToDevice, st), "gather(self)"))
    return false;
  if (world == 1) return true;
  NcclApi *a = nccl();
  if (!a) return false;
  Comm *c = comm_of(ctx);
  if (!nccl_ok(a, a->GroupStart(), "ncclGroupStart")) return false;
  bool ok = true;
  for (int peer = 0; peer < world && ok; peer++) {
    if (peer == rank) continue;
    if (bytes_of[rank] && (everyone || peer == 0)) ok = nccl_ok(a, a->Send(d_send, bytes_of[rank], ncclUint8, peer, c->comm, st), "ncclSend");
    if (ok && receiver && bytes_of[peer]) ok = nccl_ok(a, a->Recv((char *)d_recv + offset_of[peer], bytes_of[peer], ncclUint8, peer, c->comm, st), "ncclRecv");
  }
  const bool ended = nccl_ok(a, a->GroupEnd(), "ncclGroupEnd");
  return ok && ended;
}

bool comm_allgatherv(GiContext *ctx, const void *d_send, void *d_recv, const uint64_t *bytes_of, const uint64_t *offset_of, cudaStream_t st) {
  return exchange(ctx, d_send, d_recv, bytes_of, offset_of, st, true);
}
bool comm_gatherv_root0(GiContext *ctx, const void *d_send, void *d_recv, const uint64_t *bytes_of, const uint64_t *offset_of, cudaStream_t st) {
  return exchange(ctx, d_send, d_recv, bytes_of, offset_of, st, false);
}

}  // namespace gi

using namespace gi;

extern "C" {

int gi_comm_unique_id(GiUniqueId *out) {
  if (!out) { set_error("gi_comm_unique_id: null argument"); return 1; }
  static_assert(sizeof(GiUniqueId) == sizeof(ncclUniqueId), "GiUniqueId must hold an ncclUniqueId");
  NcclApi *a = nccl();
  if (!a) return 1;
  ncclUniqueId id;
  if (!nccl_ok(a, a->GetUniqueId(&id), "ncclGetUniqueId")) return 1;
  memcpy(out->bytes, id.internal, sizeof(id.internal));
  return 0;
}

int gi_comm_init_rank(GiContext *ctx, const GiUniqueId *id, int rank, int world) {
  if (!ctx || !id || world < 1 || rank < 0 || rank >= world) { set_error("gi_comm_init_rank: bad argument"); return 1; }
  NcclApi *a = nccl();
  if (!a) return 1;
  GI_CUDA(cudaSetDevice(ctx->device));
  comm_release(ctx);
  ncclUniqueId nid; memcpy(nid.internal, id->bytes, sizeof(nid.internal));
  Comm *c = new Comm(); c->rank = rank; c->world = world;
  if (!nccl_ok(a, a->CommInitRank(&c->comm, world, nid, rank), "ncclCommInitRank")) { delete c; return 1; }
  ctx->comm = c;
  return 0;
}

int gi_comm_init_all(GiContext **ctxs, int n) {
  if (!ctxs || n < 1) { set_error("gi_comm_init_all: bad argument"); return 1; }
  NcclApi *a = nccl();
  if (!a) return 1;
  std::vector<int> devs(n);
  for (int i = 0; i < n; i++) {
    if (!ctxs[i]) { set_error("gi_comm_init_all: null context"); return 1; }
    devs[i] = ctxs[i]->device;
    for (int j = 0; j < i; j++) if (devs[j] == devs[i]) { set_error("gi_comm_init_all: two contexts on one device"); return 1; }
    comm_release(ctxs[i]);
  }
  std::vector<ncclComm_t> comms(n);
  if (!nccl_ok(a, a->CommInitAll(comms.data(), n, devs.data()), "ncclCommInitAll")) return 1;
  for (int i = 0; i < n; i++) { Comm *c = new Comm(); c->comm = comms[i]; c->rank = i; c->world = n; ctxs[i]->comm = c; }
  return 0;
}

int gi_comm_info(GiContext *ctx, int *rank, int *world) {
  if (!ctx) { set_error("gi_comm_info: null context"); return 1; }
  if (rank) *rank = comm_rank(ctx);
  if (world) *world = comm_world(ctx);
  return 0;
}

int gi_comm_destroy(GiContext *ctx) {
  if (!ctx) { set_error("gi_comm_destroy: null context"); return 1; }
  cudaSetDevice(ctx->device);
  comm_release(ctx);
  return 0;
}

}  // extern "C"
