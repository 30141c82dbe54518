// halo.cu -- neighbour exchange of one patch: the replacement of the reference's MPI ghost exchanges
// (interface/wrf_pmc_trans_aero.F90 C1 :150-273, C2 :1256-1362, C3 :381-473, C4 :543-675; MPI_Cart_shift on
// WRF's periodic Cartesian communicator, :153-159).  One process per GPU: between ranks the buffers travel
// as grouped ncclSend / ncclRecv on the context's stream (NVLink 5 / NVSwitch); when the neighbour on a side
// is this rank itself (a periodic direction that is not decomposed) the "message" is a device copy.
//
// Convention: side s in {0 y-, 1 y+, 2 x-, 3 x+}; send[s] goes to the neighbour on side s, and recv[s]
// receives what the neighbour on side s sent towards its opposite side (its send[s ^ 1]).
#include "pmcgpu_internal.cuh"
#include <nccl.h>
#include <cstring>

#define NCCL_OK(ctx, call)                                                  \
  do {                                                                      \
    ncclResult_t r_ = (call);                                               \
    if (r_ != ncclSuccess) {                                                \
      (ctx)->err = std::string(#call) + ": " + ncclGetErrorString(r_);      \
      return PMCGPU_ERR_NCCL;                                               \
    }                                                                       \
  } while (0)

extern "C" int pmcgpu_comm_unique_id(char* out, int len) {
  if (!out || len < (int)sizeof(ncclUniqueId)) return PMCGPU_ERR_ARG;
  ncclUniqueId id;
  if (ncclGetUniqueId(&id) != ncclSuccess) return PMCGPU_ERR_NCCL;
  memcpy(out, &id, sizeof(id));
  return PMCGPU_OK;
}

// patches tile the periodic domain px x py, rank = rx + px * ry (WRF's 2-D x-y decomposition,
// WRFV3/main/module_wrf_top.F:505-510)
extern "C" int pmcgpu_comm_init(pmcgpu_ctx* ctx, const char* unique_id, int n_ranks, int rank, int px, int py) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx || n_ranks < 1 || rank < 0 || rank >= n_ranks || px * py != n_ranks) return PMCGPU_ERR_ARG;
  ctx->rank = rank;
  ctx->n_ranks = n_ranks;
  ctx->px = px;
  ctx->py = py;
  const int rx = rank % px, ry = rank / px;
  ctx->nb[0] = rx + px * ((ry - 1 + py) % py);
  ctx->nb[1] = rx + px * ((ry + 1) % py);
  ctx->nb[2] = (rx - 1 + px) % px + px * ry;
  ctx->nb[3] = (rx + 1) % px + px * ry;
  if (ctx->v.kind) {  // open domain: no neighbour beyond the domain faces (MPI_PROC_NULL in the reference)
    if (ry == 0) ctx->nb[0] = -1;
    if (ry == py - 1) ctx->nb[1] = -1;
    if (rx == 0) ctx->nb[2] = -1;
    if (rx == px - 1) ctx->nb[3] = -1;
  }
  if (n_ranks > 1) {
    if (!unique_id) return PMCGPU_ERR_ARG;
    ncclUniqueId id;
    memcpy(&id, unique_id, sizeof(id));
    ncclComm_t comm;
    CUDA_OK(ctx, cudaSetDevice(ctx->cfg.device));
    NCCL_OK(ctx, ncclCommInitRank(&comm, n_ranks, id, rank));
    ctx->nccl_comm = comm;
  }
  ctx->tot_valid = false;  // the neighbours changed: the next C4 messages use the full slab pitch
  return PMCGPU_OK;
}
extern "C" int pmcgpu_comm_destroy(pmcgpu_ctx* ctx) {
  if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PMCGPU_ERR_CUDA;  // contexts may live on different devices
  if (!ctx) return PMCGPU_ERR_ARG;
  if (ctx->nccl_comm) {
    ncclCommDestroy((ncclComm_t)ctx->nccl_comm);
    ctx->nccl_comm = nullptr;
  }
  return PMCGPU_OK;
}

int pmc_exchange(pmcgpu_ctx* ctx, void* const send[4], void* const recv[4], const size_t bytes[4]) {
  cudaStream_t st = ctx->stream;
  bool any_remote = false;
  for (int s = 0; s < 4; s++) {
    if (ctx->nb[s] < 0) continue;  // open domain face: nothing arrives (the receive buffer keeps its zeros)
    if (ctx->nb[s] == ctx->rank) {
      // the periodic neighbour is this patch: what left through side s ^ 1 arrives on side s
      if (bytes[s ^ 1]) CUDA_OK(ctx, cudaMemcpyAsync(recv[s], send[s ^ 1], bytes[s ^ 1], cudaMemcpyDeviceToDevice, st));
    } else {
      any_remote = true;
    }
  }
  if (!any_remote) return PMCGPU_OK;
  ncclComm_t comm = (ncclComm_t)ctx->nccl_comm;
  if (!comm) {
    ctx->err = "neighbour patches on other ranks need pmcgpu_comm_init";
    return PMCGPU_ERR_NCCL;
  }
  NCCL_OK(ctx, ncclGroupStart());
  for (int s = 0; s < 4; s++)
    if (ctx->nb[s] >= 0 && ctx->nb[s] != ctx->rank && bytes[s]) NCCL_OK(ctx, ncclSend(send[s], bytes[s], ncclChar, ctx->nb[s], comm, st));
  // receives are posted in the order 1, 0, 3, 2: when both neighbours of a direction are the same rank,
  // its first send (side 0 / 2) is the message for this rank's side 1 / 3
  const int order[4] = {1, 0, 3, 2};
  for (int q = 0; q < 4; q++) {
    const int s = order[q];
    if (ctx->nb[s] >= 0 && ctx->nb[s] != ctx->rank && bytes[s ^ 1])
      NCCL_OK(ctx, ncclRecv(recv[s], bytes[s ^ 1], ncclChar, ctx->nb[s], comm, st));
  }
  NCCL_OK(ctx, ncclGroupEnd());
  ctx->launches++;
  return PMCGPU_OK;
}

// like pmc_exchange, with the size of every incoming message given explicitly: the C4 particle messages
// (one per neighbour: header + SoA records, trans.cu)
int pmc_exchange_sized(pmcgpu_ctx* ctx, void* const send[4], void* const recv[4], const size_t send_bytes[4],
                       const size_t recv_bytes[4]) {
  cudaStream_t st = ctx->stream;
  bool any_remote = false;
  for (int s = 0; s < 4; s++) {
    if (ctx->nb[s] < 0) continue;  // open domain face
    if (ctx->nb[s] == ctx->rank) {
      if (send_bytes[s ^ 1])
        CUDA_OK(ctx, cudaMemcpyAsync(recv[s], send[s ^ 1], send_bytes[s ^ 1], cudaMemcpyDeviceToDevice, st));
    } else {
      any_remote = true;
    }
  }
  if (!any_remote) return PMCGPU_OK;
  ncclComm_t comm = (ncclComm_t)ctx->nccl_comm;
  if (!comm) {
    ctx->err = "neighbour patches on other ranks need pmcgpu_comm_init";
    return PMCGPU_ERR_NCCL;
  }
  NCCL_OK(ctx, ncclGroupStart());
  for (int s = 0; s < 4; s++)
    if (ctx->nb[s] >= 0 && ctx->nb[s] != ctx->rank && send_bytes[s])
      NCCL_OK(ctx, ncclSend(send[s], send_bytes[s], ncclChar, ctx->nb[s], comm, st));
  const int order[4] = {1, 0, 3, 2};
  for (int q = 0; q < 4; q++) {
    const int s = order[q];
    if (ctx->nb[s] >= 0 && ctx->nb[s] != ctx->rank && recv_bytes[s])
      NCCL_OK(ctx, ncclRecv(recv[s], recv_bytes[s], ncclChar, ctx->nb[s], comm, st));
  }
  NCCL_OK(ctx, ncclGroupEnd());
  ctx->launches++;
  return PMCGPU_OK;
}
