// comm.cu -- NCCL behind the Comm interface (comm.cuh), opened with dlopen.
#include <dlfcn.h>
#include <nccl.h>
#include <mutex>
#include "comm.cuh"

namespace psc {

static_assert(sizeof(ncclUniqueId) == 128, "psc_dist_unique_id hands out 128 bytes");

namespace {
struct NcclApi {
  void *handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*CommAbort)(ncclComm_t) = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
  ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*ReduceScatter)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
};
NcclApi g_nccl;
std::mutex g_nccl_mu;

template <typename F>
bool sym(F &fn, const char *name) {
  fn = reinterpret_cast<F>(dlsym(g_nccl.handle, name));
  if (!fn) set_error("libnccl has no symbol %s", name);
  return fn != nullptr;
}

void check(ncclResult_t r, const char *what) {
  if (r == ncclSuccess) return;
  set_error("NCCL: %s failed: %s", what, g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?");
  throw CudaError{PSC_ECUDA};
}
}  // namespace

bool nccl_load() {
  std::lock_guard<std::mutex> lk(g_nccl_mu);
  if (g_nccl.handle) return true;
  // a copy that is already part of the process (torch ships its own) wins over the system one
  const char *names[] = {getenv("PSC_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
  for (const char *nm : names) {
    if (!nm || !*nm) continue;
    g_nccl.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
    if (g_nccl.handle) break;
  }
  if (!g_nccl.handle) { set_error("cannot open libnccl: %s", dlerror()); return false; }
  bool ok = sym(g_nccl.GetUniqueId, "ncclGetUniqueId") && sym(g_nccl.CommInitRank, "ncclCommInitRank") &&
            sym(g_nccl.CommDestroy, "ncclCommDestroy") && sym(g_nccl.CommAbort, "ncclCommAbort") && sym(g_nccl.GetErrorString, "ncclGetErrorString") &&
            sym(g_nccl.Broadcast, "ncclBroadcast") && sym(g_nccl.ReduceScatter, "ncclReduceScatter") &&
            sym(g_nccl.AllGather, "ncclAllGather") && sym(g_nccl.AllReduce, "ncclAllReduce") &&
            sym(g_nccl.Send, "ncclSend") && sym(g_nccl.Recv, "ncclRecv") && sym(g_nccl.GroupStart, "ncclGroupStart") &&
            sym(g_nccl.GroupEnd, "ncclGroupEnd");
  if (!ok) { dlclose(g_nccl.handle); g_nccl.handle = nullptr; }
  return ok;
}

int nccl_unique_id(uint8_t id[128]) {
  if (!nccl_load()) return PSC_ECUDA;
  ncclUniqueId u;
  ncclResult_t r = g_nccl.GetUniqueId(&u);
  if (r != ncclSuccess) { set_error("ncclGetUniqueId: %s", g_nccl.GetErrorString(r)); return PSC_ECUDA; }
  memcpy(id, &u, 128);
  return PSC_OK;
}

int nccl_comm_init(Comm *c, const uint8_t id[128], int rank, int world) {
  c->rank = rank;
  c->world = world;
  c->nccl = nullptr;
  if (world == 1) return PSC_OK;
  if (!nccl_load()) return PSC_ECUDA;
  ncclUniqueId u;
  memcpy(&u, id, 128);
  ncclComm_t comm = nullptr;
  ncclResult_t r = g_nccl.CommInitRank(&comm, world, u, rank);
  if (r != ncclSuccess) { set_error("ncclCommInitRank(rank %d of %d): %s", rank, world, g_nccl.GetErrorString(r)); return PSC_ECUDA; }
  c->nccl = comm;
  return PSC_OK;
}

// any thread may call this: outstanding collectives of the communicator end with an error instead of waiting
// for a rank that has failed
void nccl_comm_abort(Comm *c) {
  void *h = c->nccl;
  c->nccl = nullptr;
  if (h && g_nccl.CommAbort) g_nccl.CommAbort((ncclComm_t)h);
}

void nccl_comm_destroy(Comm *c) {
  if (c->nccl && g_nccl.CommDestroy) g_nccl.CommDestroy((ncclComm_t)c->nccl);
  c->nccl = nullptr;
}

void Comm::bcast(psc_context *ctx, void *buf_d, size_t nbytes, int root) {
  if (world == 1 || nbytes == 0) return;
  check(g_nccl.Broadcast(buf_d, buf_d, nbytes, ncclUint8, root, (ncclComm_t)nccl, ctx->stream), "ncclBroadcast");
  n_calls++; bytes += (double)nbytes;
}

void Comm::reduce_scatter_u32(psc_context *ctx, const uint32_t *send_d, uint32_t *recv_d, size_t per) {
  if (per == 0) return;
  if (world == 1) {
    if (recv_d != send_d) PSC_CUDA(cudaMemcpyAsync(recv_d, send_d, per * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    return;
  }
  check(g_nccl.ReduceScatter(send_d, recv_d, per, ncclUint32, ncclSum, (ncclComm_t)nccl, ctx->stream), "ncclReduceScatter");
  n_calls++; bytes += (double)per * 4 * world;
}

void Comm::reduce_scatter_u8(psc_context *ctx, const uint8_t *send_d, uint8_t *recv_d, size_t per) {
  if (per == 0) return;
  if (world == 1) {
    if (recv_d != send_d) PSC_CUDA(cudaMemcpyAsync(recv_d, send_d, per, cudaMemcpyDeviceToDevice, ctx->stream));
    return;
  }
  check(g_nccl.ReduceScatter(send_d, recv_d, per, ncclUint8, ncclSum, (ncclComm_t)nccl, ctx->stream), "ncclReduceScatter(u8)");
  n_calls++; bytes += (double)per * world;
}

void Comm::allreduce_max_u32(psc_context *ctx, uint32_t *buf_d, size_t count) {
  if (world == 1 || count == 0) return;
  check(g_nccl.AllReduce(buf_d, buf_d, count, ncclUint32, ncclMax, (ncclComm_t)nccl, ctx->stream), "ncclAllReduce(max)");
  n_calls++; bytes += (double)count * 4;
}

void Comm::allgather(psc_context *ctx, const void *send_d, void *recv_d, size_t bytes_per) {
  if (bytes_per == 0) return;
  if (world == 1) {
    if (recv_d != send_d) PSC_CUDA(cudaMemcpyAsync(recv_d, send_d, bytes_per, cudaMemcpyDeviceToDevice, ctx->stream));
    return;
  }
  check(g_nccl.AllGather(send_d, recv_d, bytes_per, ncclUint8, (ncclComm_t)nccl, ctx->stream), "ncclAllGather");
  n_calls++; bytes += (double)bytes_per * world;
}

void Comm::allreduce_sum_u64(psc_context *ctx, uint64_t *buf_d, size_t count) {
  if (world == 1 || count == 0) return;
  check(g_nccl.AllReduce(buf_d, buf_d, count, ncclUint64, ncclSum, (ncclComm_t)nccl, ctx->stream), "ncclAllReduce");
  n_calls++; bytes += (double)count * 8;
}

void Comm::group_begin() { if (world > 1) check(g_nccl.GroupStart(), "ncclGroupStart"); }
void Comm::group_end() { if (world > 1) check(g_nccl.GroupEnd(), "ncclGroupEnd"); }

void Comm::send(psc_context *ctx, const void *buf_d, size_t nbytes, int peer) {
  if (nbytes == 0) return;
  check(g_nccl.Send(buf_d, nbytes, ncclUint8, peer, (ncclComm_t)nccl, ctx->stream), "ncclSend");
  n_calls++; bytes += (double)nbytes;
}

void Comm::recv(psc_context *ctx, void *buf_d, size_t nbytes, int peer) {
  if (nbytes == 0) return;
  check(g_nccl.Recv(buf_d, nbytes, ncclUint8, peer, (ncclComm_t)nccl, ctx->stream), "ncclRecv");
}

}  // namespace psc
