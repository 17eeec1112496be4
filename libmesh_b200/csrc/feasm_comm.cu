// feasm_comm.cu — the multi-rank exchange below the C ABI: NCCL communicator handling and the per-step ghost-row exchange.
//
// Mirrors what DistributedVector::close (src/numerics/distributed_vector.C:399-470) and PETSc's MatAssemblyBegin/End do at
// the end of a parallel assembly: contributions to rows another rank owns are shipped to the owner and added there.  Here
// the ghost rows were already reduced and packed on the device (feasm_assemble_begin / feasm_pack_ghost); this file issues
// the grouped ncclSend / ncclRecv of the exchange plan on the context's own high-priority communication stream, ordered by
// events against the assembly stream, and adds the received values (feasm_unpack_ghost, fixed peer order).
//
// NCCL is bound at run time (dlopen of libnccl.so.2, reusing a copy the process already loaded — e.g. the one torch
// brought): libfeasm.so itself loads on a box without NCCL, and a host application's NCCL is never shadowed by another.
#include <dlfcn.h>
#include <nccl.h>

#include <cstdlib>
#include <mutex>
#include <vector>

#include "feasm_internal.h"

namespace {

struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  std::string err;
};

NcclApi* nccl_api() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);  // the process's own NCCL, if it has one
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { api.err = std::string("NCCL not found: ") + dlerror(); return; }
    api.lib = h;
    auto sym = [&](const char* name) { void* p = dlsym(h, name); if (!p) api.err = std::string("NCCL symbol missing: ") + name; return p; };
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
    api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
    api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
    api.Send = reinterpret_cast<decltype(api.Send)>(sym("ncclSend"));
    api.Recv = reinterpret_cast<decltype(api.Recv)>(sym("ncclRecv"));
    api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
  });
  return &api;
}

int nccl_fail(feasm_ctx* ctx, NcclApi* a, const char* what, ncclResult_t r) {
  return feasm_fail(ctx, FEASM_ERR_COMM, std::string(what) + ": " + (a->GetErrorString ? a->GetErrorString(r) : "NCCL error"));
}
#define FEASM_NCCL(ctx, a, call)                                   \
  do {                                                             \
    ncclResult_t r__ = (call);                                     \
    if (r__ != ncclSuccess) return nccl_fail(ctx, a, #call, r__);  \
  } while (0)

int ensure_comm_stream(feasm_ctx* ctx) {
  if (ctx->comm_stream) return FEASM_OK;
  int lo = 0, hi = 0;
  FEASM_CUDA(ctx, cudaDeviceGetStreamPriorityRange(&lo, &hi));
  // high priority: the transfer's CTAs should take SM slots as they free up while the owned rows are being reduced
  FEASM_CUDA(ctx, cudaStreamCreateWithPriority(&ctx->comm_stream, cudaStreamNonBlocking, hi));
  FEASM_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_packed, cudaEventDisableTiming));
  FEASM_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_exchanged, cudaEventDisableTiming));
  return FEASM_OK;
}

}  // namespace

void feasm_comm_release(feasm_ctx* ctx) {
  if (ctx->comm && ctx->comm_owned) {
    NcclApi* a = nccl_api();
    if (a->CommDestroy) a->CommDestroy(static_cast<ncclComm_t>(ctx->comm));
  }
  ctx->comm = nullptr;
  ctx->comm_owned = false;
  if (ctx->comm_stream) {
    cudaStreamSynchronize(ctx->comm_stream);
    cudaStreamDestroy(ctx->comm_stream);
    cudaEventDestroy(ctx->ev_packed);
    cudaEventDestroy(ctx->ev_exchanged);
    ctx->comm_stream = nullptr;
  }
}

// Grouped point-to-point exchange with the peers of a plan on `stream` (elem_bytes 4: uint32 words, 8: doubles): the
// building block of the halo exchange of a distributed vector (feasm_halo.cu).
int feasm_comm_p2p(feasm_ctx* ctx, int n_peers, const int32_t* peer, const uint64_t* send_ptr, const void* d_send,
                   const uint64_t* recv_ptr, void* d_recv, int elem_bytes, cudaStream_t stream) {
  if (!ctx->comm) return feasm_fail(ctx, FEASM_ERR_STATE, "no communicator: feasm_comm_init / feasm_set_communicator first");
  NcclApi* a = nccl_api();
  ncclComm_t comm = static_cast<ncclComm_t>(ctx->comm);
  const ncclDataType_t dt = elem_bytes == 8 ? ncclDouble : ncclUint32;
  FEASM_NCCL(ctx, a, a->GroupStart());
  for (int p = 0; p < n_peers; ++p) {
    const uint64_t r0 = recv_ptr[p], nr = recv_ptr[p + 1] - r0;
    if (nr) FEASM_NCCL(ctx, a, a->Recv(static_cast<char*>(d_recv) + r0 * elem_bytes, nr, dt, peer[p], comm, stream));
  }
  for (int p = 0; p < n_peers; ++p) {
    const uint64_t s0 = send_ptr[p], ns = send_ptr[p + 1] - s0;
    if (ns) FEASM_NCCL(ctx, a, a->Send(static_cast<const char*>(d_send) + s0 * elem_bytes, ns, dt, peer[p], comm, stream));
  }
  FEASM_NCCL(ctx, a, a->GroupEnd());
  return FEASM_OK;
}

// n doubles from every rank to every rank, in rank order, device to device on `stream`: the partial sums of a distributed
// dot product, which every rank then adds in the same fixed order (feasm_solve.cu).
int feasm_comm_allgather_f64(feasm_ctx* ctx, const double* d_send, double* d_recv, int n, cudaStream_t stream) {
  if (!ctx->comm) return feasm_fail(ctx, FEASM_ERR_STATE, "no communicator");
  NcclApi* a = nccl_api();
  FEASM_NCCL(ctx, a, a->AllGather(d_send, d_recv, (size_t)n, ncclDouble, static_cast<ncclComm_t>(ctx->comm), stream));
  return FEASM_OK;
}

extern "C" {

int feasm_comm_unique_id(void* id_bytes) {
  static_assert(sizeof(ncclUniqueId) == FEASM_COMM_ID_BYTES, "include/feasm.h states the size of the bootstrap id");
  if (!id_bytes) return feasm_fail(nullptr, FEASM_ERR_ARG, "id_bytes is NULL");
  NcclApi* a = nccl_api();
  if (!a->lib || !a->err.empty()) return feasm_fail(nullptr, FEASM_ERR_COMM, a->err);
  ncclUniqueId id;
  FEASM_NCCL(nullptr, a, a->GetUniqueId(&id));
  std::memcpy(id_bytes, &id, sizeof(id));
  return FEASM_OK;
}

int feasm_comm_init(feasm_ctx* ctx, int n_ranks, int rank, const void* id_bytes) {
  if (!ctx || !id_bytes || n_ranks < 1 || rank < 0 || rank >= n_ranks) return feasm_fail(ctx, FEASM_ERR_ARG, "bad communicator arguments");
  NcclApi* a = nccl_api();
  if (!a->lib || !a->err.empty()) return feasm_fail(ctx, FEASM_ERR_COMM, a->err);
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  feasm_comm_release(ctx);
  ncclUniqueId id;
  std::memcpy(&id, id_bytes, sizeof(id));
  ncclComm_t comm = nullptr;
  FEASM_NCCL(ctx, a, a->CommInitRank(&comm, n_ranks, id, rank));
  ctx->comm = comm;
  ctx->comm_owned = true;
  ctx->comm_rank = rank;
  ctx->comm_size = n_ranks;
  return ensure_comm_stream(ctx);
}

int feasm_set_communicator(feasm_ctx* ctx, void* nccl_comm, int n_ranks, int rank) {
  if (!ctx || n_ranks < 1 || rank < 0 || rank >= n_ranks) return feasm_fail(ctx, FEASM_ERR_ARG, "bad communicator arguments");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  feasm_comm_release(ctx);
  if (!nccl_comm) return FEASM_OK;  // back to one rank
  NcclApi* a = nccl_api();
  if (!a->lib || !a->err.empty()) return feasm_fail(ctx, FEASM_ERR_COMM, a->err);
  ctx->comm = nccl_comm;
  ctx->comm_owned = false;
  ctx->comm_rank = rank;
  ctx->comm_size = n_ranks;
  return ensure_comm_stream(ctx);
}

int feasm_exchange_begin(feasm_ctx* ctx) {
  if (!ctx) return FEASM_ERR_ARG;
  if (!ctx->comm) return feasm_fail(ctx, FEASM_ERR_STATE, "no communicator: feasm_comm_init / feasm_set_communicator first");
  if (!ctx->d_send_buf) return feasm_fail(ctx, FEASM_ERR_STATE, "no exchange plan");
  if (ctx->exchange_open) return feasm_fail(ctx, FEASM_ERR_STATE, "feasm_exchange_begin twice without feasm_exchange_end");
  NcclApi* a = nccl_api();
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  ncclComm_t comm = static_cast<ncclComm_t>(ctx->comm);
  // the transfer starts when the send buffer is packed (and the previous unpack has left the receive buffer: same stream)
  FEASM_CUDA(ctx, cudaEventRecord(ctx->ev_packed, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamWaitEvent(ctx->comm_stream, ctx->ev_packed, 0));
  FEASM_NCCL(ctx, a, a->GroupStart());
  for (int p = 0; p < ctx->n_peers; ++p) {
    const uint64_t r0 = ctx->recv_ptr[p], nr = ctx->recv_ptr[p + 1] - r0;
    if (nr) FEASM_NCCL(ctx, a, a->Recv(ctx->d_recv_buf + r0, nr, ncclDouble, ctx->peer_rank[p], comm, ctx->comm_stream));
  }
  for (int p = 0; p < ctx->n_peers; ++p) {
    const uint64_t s0 = ctx->send_ptr[p], ns = ctx->send_ptr[p + 1] - s0;
    if (ns) FEASM_NCCL(ctx, a, a->Send(ctx->d_send_buf + s0, ns, ncclDouble, ctx->peer_rank[p], comm, ctx->comm_stream));
  }
  FEASM_NCCL(ctx, a, a->GroupEnd());
  FEASM_CUDA(ctx, cudaEventRecord(ctx->ev_exchanged, ctx->comm_stream));
  ctx->exchange_open = true;
  return FEASM_OK;
}

int feasm_exchange_end(feasm_ctx* ctx) {
  if (!ctx) return FEASM_ERR_ARG;
  if (!ctx->exchange_open) return feasm_fail(ctx, FEASM_ERR_STATE, "feasm_exchange_end without feasm_exchange_begin");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  ctx->exchange_open = false;
  // the assembly stream waits for the transfer; the next step's pack into the send buffer is ordered behind it as well
  FEASM_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_exchanged, 0));
  return feasm_launch_unpack(ctx);
}

int feasm_exchange_ghost_rows(feasm_ctx* ctx) {
  const int rc = feasm_exchange_begin(ctx);
  return rc ? rc : feasm_exchange_end(ctx);
}

// Setup helper: every rank contributes n values, every rank receives all of them in rank order (ncclAllGather on device
// scratch) — the ownership ranges / counts a host without MPI needs to build the exchange plan.
int feasm_comm_allgather_u64(feasm_ctx* ctx, const uint64_t* send, int n, uint64_t* recv) {
  if (!ctx || !send || !recv || n <= 0) return feasm_fail(ctx, FEASM_ERR_ARG, "bad allgather arguments");
  if (!ctx->comm) return feasm_fail(ctx, FEASM_ERR_STATE, "no communicator");
  NcclApi* a = nccl_api();
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  uint64_t* d = nullptr;
  const size_t total = (size_t)n * (ctx->comm_size + 1);
  FEASM_CUDA(ctx, cudaMalloc(&d, sizeof(uint64_t) * total));
  int rc = FEASM_OK;
  auto body = [&]() -> int {
    FEASM_CUDA(ctx, cudaMemcpyAsync(d, send, sizeof(uint64_t) * n, cudaMemcpyHostToDevice, ctx->stream));
    FEASM_NCCL(ctx, a, a->AllGather(d, d + n, (size_t)n, ncclUint64, static_cast<ncclComm_t>(ctx->comm), ctx->stream));
    FEASM_CUDA(ctx, cudaMemcpyAsync(recv, d + n, sizeof(uint64_t) * n * ctx->comm_size, cudaMemcpyDeviceToHost, ctx->stream));
    FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return FEASM_OK;
  };
  rc = body();
  cudaFree(d);
  return rc;
}

// The setup of a multi-rank assembly over the context's own communicator: pattern of this rank (device) -> ghost-row
// messages -> owners (sizes by ncclAllGather, the lists by one grouped ncclSend / ncclRecv of uint32 words) -> merged
// pattern, slot plan, scatter maps (feasm_plan.cpp).  Collective: every rank of the communicator calls it.
int feasm_build_exchange(feasm_ctx* ctx, const uint64_t* first_dof) {
  if (!ctx || !first_dof) return feasm_fail(ctx, FEASM_ERR_ARG, "first_dof is NULL");
  if (!ctx->comm) return feasm_fail(ctx, FEASM_ERR_STATE, "no communicator: feasm_comm_init / feasm_set_communicator first");
  if (!ctx->d_row_ptr) return feasm_fail(ctx, FEASM_ERR_STATE, "pattern not set: call feasm_build_pattern first");
  if (ctx->n_owned < ctx->n_rows && !ctx->d_gid) return feasm_fail(ctx, FEASM_ERR_STATE, "ghost rows need global ids (feasm_set_dofs)");
  NcclApi* a = nccl_api();
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  const int world = ctx->comm_size, rank = ctx->comm_rank;
  std::vector<uint64_t> rp((size_t)ctx->n_rows + 1);
  std::vector<uint32_t> ci((size_t)ctx->nnz), gid((size_t)ctx->n_rows);
  FEASM_CUDA(ctx, cudaMemcpyAsync(rp.data(), ctx->d_row_ptr, sizeof(uint64_t) * rp.size(), cudaMemcpyDeviceToHost, ctx->stream));
  FEASM_CUDA(ctx, cudaMemcpyAsync(ci.data(), ctx->d_col_idx, sizeof(uint32_t) * ci.size(), cudaMemcpyDeviceToHost, ctx->stream));
  if (ctx->d_gid) FEASM_CUDA(ctx, cudaMemcpyAsync(gid.data(), ctx->d_gid, sizeof(uint32_t) * gid.size(), cudaMemcpyDeviceToHost, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  feasm_planner* pl = feasm_planner_create(rank, world, first_dof, (uint64_t)ctx->n_owned, (uint64_t)ctx->n_rows, gid.data(), rp.data(), ci.data());
  if (!pl) return feasm_fail(ctx, FEASM_ERR_ARG, "dof ranges / global ids do not describe owned rows first, then ghost rows in ascending global id");
  std::vector<uint32_t>().swap(ci);  // the planner holds its own copy
  uint32_t* d_out = nullptr;
  uint32_t* d_in = nullptr;
  auto body = [&]() -> int {
    // message sizes: row q of the gathered matrix = what rank q sends to every rank
    std::vector<uint64_t> mine(world, 0), all((size_t)world * world, 0);
    uint64_t n_out = 0, n_in = 0;
    for (int q = 0; q < world; ++q) { mine[q] = feasm_planner_message(pl, q, nullptr); n_out += mine[q]; }
    int rc = feasm_comm_allgather_u64(ctx, mine.data(), world, all.data());
    if (rc) return rc;
    for (int q = 0; q < world; ++q) n_in += all[(size_t)q * world + rank];
    FEASM_CUDA(ctx, cudaMalloc(&d_out, sizeof(uint32_t) * (n_out + 1)));
    FEASM_CUDA(ctx, cudaMalloc(&d_in, sizeof(uint32_t) * (n_in + 1)));
    uint64_t o = 0;
    for (int q = 0; q < world; ++q) {
      const uint32_t* w = nullptr;
      const uint64_t n = feasm_planner_message(pl, q, &w);
      if (n) FEASM_CUDA(ctx, cudaMemcpyAsync(d_out + o, w, sizeof(uint32_t) * n, cudaMemcpyHostToDevice, ctx->stream));
      o += n;
    }
    ncclComm_t comm = static_cast<ncclComm_t>(ctx->comm);
    FEASM_NCCL(ctx, a, a->GroupStart());
    uint64_t i = 0;
    for (int q = 0; q < world; ++q) {
      const uint64_t n = all[(size_t)q * world + rank];
      if (n) FEASM_NCCL(ctx, a, a->Recv(d_in + i, n, ncclUint32, q, comm, ctx->stream));
      i += n;
    }
    o = 0;
    for (int q = 0; q < world; ++q) {
      if (mine[q]) FEASM_NCCL(ctx, a, a->Send(d_out + o, mine[q], ncclUint32, q, comm, ctx->stream));
      o += mine[q];
    }
    FEASM_NCCL(ctx, a, a->GroupEnd());
    std::vector<uint32_t> in(n_in + 1);
    FEASM_CUDA(ctx, cudaMemcpyAsync(in.data(), d_in, sizeof(uint32_t) * n_in, cudaMemcpyDeviceToHost, ctx->stream));
    FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    i = 0;
    for (int q = 0; q < world; ++q) {
      const uint64_t n = all[(size_t)q * world + rank];
      if (n && (rc = feasm_planner_receive(pl, q, in.data() + i, n)) != FEASM_OK) return feasm_fail(ctx, rc, feasm_planner_error(pl));
      i += n;
    }
    if ((rc = feasm_planner_finish(pl)) != FEASM_OK) return feasm_fail(ctx, rc, feasm_planner_error(pl));
    return feasm_apply_plan(ctx, pl);
  };
  int rc = body();
  cudaFree(d_out);
  cudaFree(d_in);
  feasm_planner_destroy(pl);
  if (rc) return rc;
  // interface elements first (feasm_set_interface_prefix): the leading elements that cover every element touching a ghost
  // row — with LinearPartitioner slabs the first element layer.  Used when that prefix is at most a third of the elements.
  int64_t prefix = 0;
  if (ctx->n_owned < ctx->n_rows && ctx->d_elem_dofs && !(getenv("FEASM_IFACE") && atoi(getenv("FEASM_IFACE")) == 0)) {
    std::vector<uint32_t> ed((size_t)ctx->n_elem * ctx->nd);
    FEASM_CUDA(ctx, cudaMemcpy(ed.data(), ctx->d_elem_dofs, sizeof(uint32_t) * ed.size(), cudaMemcpyDeviceToHost));
    for (int64_t e = ctx->n_elem - 1; e >= 0 && !prefix; --e)
      for (int k = 0; k < ctx->nd; ++k)
        if (ed[(size_t)e * ctx->nd + k] >= (uint64_t)ctx->n_owned) { prefix = e + 1; break; }
    if (prefix > ctx->n_elem / 3) prefix = 0;
  }
  return feasm_set_interface_prefix(ctx, prefix);
}

}  // extern "C"
