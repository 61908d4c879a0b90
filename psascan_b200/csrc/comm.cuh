// comm.cuh -- the collectives of the multi-GPU schedule (dist.cu), NCCL over NVLink / NVSwitch.
// libnccl is opened at run time (dlopen), so libpsascan_cuda.so has no link-time dependency on it: a
// single-GPU host needs no NCCL, and inside a torch process the already loaded copy is the one used.
// Every collective is enqueued on the context's stream: no host synchronisation, no callback.
#pragma once
#include "common.cuh"

namespace psc {

struct Comm {
  int rank = 0, world = 1;
  void *nccl = nullptr;  // ncclComm_t
  // counters of one construction (reported through psc_dist_stats)
  long n_calls = 0;
  double bytes = 0;

  void bcast(psc_context *ctx, void *buf_d, size_t bytes, int root);
  // recv_d[0 .. per) = sum over ranks of their send_d[rank * per .. (rank + 1) * per)
  void reduce_scatter_u32(psc_context *ctx, const uint32_t *send_d, uint32_t *recv_d, size_t per);
  void reduce_scatter_u8(psc_context *ctx, const uint8_t *send_d, uint8_t *recv_d, size_t per);  // sums must stay < 256
  void allreduce_max_u32(psc_context *ctx, uint32_t *buf_d, size_t count);
  // recv_d[g * bytes_per .. ) = rank g's send_d; in place allowed when send_d == recv_d + rank * bytes_per
  void allgather(psc_context *ctx, const void *send_d, void *recv_d, size_t bytes_per);
  void allreduce_sum_u64(psc_context *ctx, uint64_t *buf_d, size_t count);
  // point-to-point exchange: all sends and receives between group_begin and group_end form one NCCL group;
  // a pair of ranks must issue its transfers in the same order on both sides
  void group_begin();
  void group_end();
  void send(psc_context *ctx, const void *buf_d, size_t bytes, int peer);
  void recv(psc_context *ctx, void *buf_d, size_t bytes, int peer);
};

bool nccl_load();            // false + set_error when libnccl cannot be opened
int nccl_unique_id(uint8_t id[128]);
int nccl_comm_init(Comm *c, const uint8_t id[128], int rank, int world);
void nccl_comm_destroy(Comm *c);
void nccl_comm_abort(Comm *c);

}  // namespace psc
