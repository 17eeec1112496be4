// feasm_halo.cu — the matrix of a multi-rank assembly used on the device: the halo of a distributed vector.
//
// After a parallel assembly every rank holds its owned rows with GLOBAL column ids (the layout of a PETSc MPIAIJ matrix
// behind libMesh's PetscMatrix, or of DistributedVector + the Graph's rows).  y = A x on those rows
// (SparseMatrix::vector_mult, include/numerics/sparse_matrix.h:579-587) needs the entries of x that other ranks own —
// what NumericVector::localize(v_local, send_list) (include/numerics/numeric_vector.h:478-487, DistributedVector::localize,
// src/numerics/distributed_vector.C:474-560) fetches in libMesh with the DofMap's send_list.  Here:
//
//   setup     feasm_build_halo: the columns of the owned rows outside the owned range (plus the ids of the ghost rows, for
//             feasm_residual) are found on the device, split by owner, and the request lists are exchanged once; every rank
//             then knows which of its owned entries each peer needs.
//   per use   pack (one gather kernel) -> one grouped ncclSend / ncclRecv per neighbour -> unpack (one scatter kernel) into
//             a column-space copy of x of GLOBAL length (n_global doubles: 57 MB at 96^3 — the price of keeping the global
//             column ids of the assembled pattern as they are), in which the owned part is a plain copy.
//
// Hosts that carry messages themselves (and the one-process emulation of the GPU tests) use feasm_halo_columns +
// feasm_set_halo_plan and feasm_halo_pack / _recv_buffer / _unpack.
#include <algorithm>
#include <vector>

#include "feasm_internal.h"

struct feasm_halo {
  uint64_t n_global = 0, first = 0;
  int n_peers = 0;
  std::vector<int32_t> peer;
  std::vector<uint64_t> send_ptr, recv_ptr;
  uint32_t* d_send_rows = nullptr;  // owned local rows whose entries the peers need, peer-major
  uint32_t* d_recv_gids = nullptr;  // global ids of the entries received, peer-major
  double* d_send = nullptr;
  double* d_recv = nullptr;
  double* d_xfull = nullptr;        // [n_global]
  std::vector<uint32_t> cols;       // feasm_halo_columns' last answer
  bool cols_valid = false;
  // feasm_residual on several ranks: local copy of u (owned + ghost rows), local row sums, and the vector part of the
  // assembly's exchange plan (the right-hand-side slots) as row lists
  double* d_uloc = nullptr;
  double* d_rloc = nullptr;
  bool vplan = false;
  std::vector<uint64_t> vsend_ptr, vrecv_ptr;
  uint32_t* d_vsend_rows = nullptr;
  uint32_t* d_vrecv_rows = nullptr;
  double* d_vsend = nullptr;
  double* d_vrecv = nullptr;
};

namespace {

template <typename T>
void drop(T*& p) {
  if (p) cudaFree(p);
  p = nullptr;
}
template <typename T>
int put(feasm_ctx* ctx, T** dst, const T* src, size_t n) {
  drop(*dst);
  FEASM_CUDA(ctx, cudaMalloc(dst, sizeof(T) * (n ? n : 1)));
  if (n) FEASM_CUDA(ctx, cudaMemcpyAsync(*dst, src, sizeof(T) * n, cudaMemcpyHostToDevice, ctx->stream));
  return FEASM_OK;
}

// flag[c] = 1 for every column c of the first `n_entries` pattern entries (the owned rows) outside [first, end)
__global__ void k_halo_mark_cols(const uint32_t* __restrict__ col_idx, uint64_t n_entries, uint32_t first, uint32_t end, uint64_t n_global,
                                 uint8_t* __restrict__ flag) {
  for (uint64_t k = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; k < n_entries; k += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t c = col_idx[k];
    if ((c < first || c >= end) && c < n_global) flag[c] = 1;
  }
}
__global__ void k_halo_mark_rows(const uint32_t* __restrict__ gid, int64_t r0, int64_t r1, uint64_t n_global, uint8_t* __restrict__ flag) {
  const int64_t r = r0 + blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r < r1 && gid[r] < n_global) flag[gid[r]] = 1;
}
// `lim`: length of the indexed vector (checked by the FEASM_DEBUG_BOUNDS build only)
__global__ void k_halo_pack(const double* __restrict__ x, const uint32_t* __restrict__ rows, uint64_t n, double* __restrict__ out, uint64_t lim) {
  const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (i < n) { FEASM_DEV_ASSERT(rows[i] < lim); out[i] = x[rows[i]]; }
}
__global__ void k_halo_unpack(const double* __restrict__ in, const uint32_t* __restrict__ gids, uint64_t n, double* __restrict__ xfull, uint64_t lim) {
  const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (i < n) { FEASM_DEV_ASSERT(gids[i] < lim); xfull[gids[i]] = in[i]; }
}
// u_loc[r] = x_full[global id of local row r]
__global__ void k_halo_localize(const double* __restrict__ xfull, const uint32_t* __restrict__ gid, int64_t n_rows, double* __restrict__ uloc, uint64_t lim) {
  const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r < n_rows) { FEASM_DEV_ASSERT(gid[r] < lim); uloc[r] = xfull[gid[r]]; }
}
// r[rows[i]] += in[i]; the rows of one peer's message are distinct, peers are added one launch after the other (fixed order)
__global__ void k_halo_add_rows(const double* __restrict__ in, const uint32_t* __restrict__ rows, uint64_t n, double* __restrict__ r, uint64_t lim) {
  const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (i < n) { FEASM_DEV_ASSERT(rows[i] < lim); r[rows[i]] += in[i]; }
}

feasm_halo* halo_of(feasm_ctx* ctx) {
  if (!ctx->halo) ctx->halo = new feasm_halo;
  return ctx->halo;
}

unsigned blocks(uint64_t n, unsigned per = 256) { return (unsigned)((n + per - 1) / per); }

// the right-hand-side part of the assembly's exchange plan as row lists (slot >= nnz: row = slot - nnz)
int ensure_vplan(feasm_ctx* ctx, feasm_halo* h) {
  if (h->vplan) return FEASM_OK;
  const uint64_t ns = ctx->send_ptr.empty() ? 0 : ctx->send_ptr.back(), nr = ctx->recv_ptr.empty() ? 0 : ctx->recv_ptr.back();
  std::vector<uint64_t> ss(ns), rs(nr);
  if (ns) FEASM_CUDA(ctx, cudaMemcpyAsync(ss.data(), ctx->d_send_slots, sizeof(uint64_t) * ns, cudaMemcpyDeviceToHost, ctx->stream));
  if (nr) FEASM_CUDA(ctx, cudaMemcpyAsync(rs.data(), ctx->d_recv_slots, sizeof(uint64_t) * nr, cudaMemcpyDeviceToHost, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  std::vector<uint32_t> srows, rrows;
  h->vsend_ptr.assign(1, 0);
  h->vrecv_ptr.assign(1, 0);
  for (int p = 0; p < ctx->n_peers; ++p) {
    for (uint64_t i = ctx->send_ptr[p]; i < ctx->send_ptr[p + 1]; ++i)
      if (ss[i] >= ctx->nnz) srows.push_back((uint32_t)(ss[i] - ctx->nnz));
    for (uint64_t i = ctx->recv_ptr[p]; i < ctx->recv_ptr[p + 1]; ++i)
      if (rs[i] >= ctx->nnz) rrows.push_back((uint32_t)(rs[i] - ctx->nnz));
    h->vsend_ptr.push_back(srows.size());
    h->vrecv_ptr.push_back(rrows.size());
  }
  int rc = put(ctx, &h->d_vsend_rows, srows.data(), srows.size());
  if (!rc) rc = put(ctx, &h->d_vrecv_rows, rrows.data(), rrows.size());
  if (rc) return rc;
  drop(h->d_vsend); drop(h->d_vrecv);
  FEASM_CUDA(ctx, cudaMalloc(&h->d_vsend, sizeof(double) * (srows.size() + 1)));
  FEASM_CUDA(ctx, cudaMalloc(&h->d_vrecv, sizeof(double) * (rrows.size() + 1)));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  h->vplan = true;
  return FEASM_OK;
}

}  // namespace

void feasm_halo_release(feasm_ctx* ctx) {
  feasm_halo* h = ctx->halo;
  if (!h) return;
  drop(h->d_send_rows); drop(h->d_recv_gids); drop(h->d_send); drop(h->d_recv); drop(h->d_xfull);
  drop(h->d_uloc); drop(h->d_rloc); drop(h->d_vsend_rows); drop(h->d_vrecv_rows); drop(h->d_vsend); drop(h->d_vrecv);
  delete h;
  ctx->halo = nullptr;
}

bool feasm_halo_ready(const feasm_ctx* ctx) { return ctx->halo && ctx->halo->d_xfull; }
uint32_t feasm_halo_first(const feasm_ctx* ctx) { return feasm_halo_ready(ctx) ? (uint32_t)ctx->halo->first : 0u; }

int feasm_halo_exchange(feasm_ctx* ctx, const double* d_x, const double** d_xfull) {
  feasm_halo* h = ctx->halo;
  if (!h || !h->d_xfull) {
    if (!ctx->d_gid && ctx->n_owned == ctx->n_rows) { *d_xfull = d_x; return FEASM_OK; }  // one rank: column ids are row ids
    return feasm_fail(ctx, FEASM_ERR_STATE, "rows of this matrix are spread over several ranks: feasm_build_halo (or feasm_set_halo_plan) first");
  }
  if (ctx->n_owned) FEASM_CUDA(ctx, cudaMemcpyAsync(h->d_xfull + h->first, d_x, sizeof(double) * ctx->n_owned, cudaMemcpyDeviceToDevice, ctx->stream));
  if (ctx->comm && h->n_peers) {
    const uint64_t ns = h->send_ptr.back(), nr = h->recv_ptr.back();
    if (ns) k_halo_pack<<<blocks(ns), 256, 0, ctx->stream>>>(d_x, h->d_send_rows, ns, h->d_send, (uint64_t)ctx->n_owned);
    int rc = feasm_comm_p2p(ctx, h->n_peers, h->peer.data(), h->send_ptr.data(), h->d_send, h->recv_ptr.data(), h->d_recv, 8, ctx->stream);
    if (rc) return rc;
    if (nr) k_halo_unpack<<<blocks(nr), 256, 0, ctx->stream>>>(h->d_recv, h->d_recv_gids, nr, h->d_xfull, h->n_global);
    FEASM_CUDA(ctx, cudaGetLastError());
  }
  *d_xfull = h->d_xfull;
  return FEASM_OK;
}

// feasm_residual on several ranks, after the element phase: u's ghost entries from their owners, the row sums over owned
// AND ghost rows, the ghost-row sums to their owners (added there peer by peer in the plan's order), owned part to d_r.
int feasm_residual_distributed(feasm_ctx* ctx, const double* d_u, double* d_r) {
  feasm_halo* h = ctx->halo;
  if (!h || !h->d_xfull) return feasm_fail(ctx, FEASM_ERR_STATE, "feasm_residual on several ranks: feasm_build_halo first");
  if (ctx->n_owned < ctx->n_rows && !ctx->comm) return feasm_fail(ctx, FEASM_ERR_STATE, "feasm_residual on several ranks needs the context's communicator");
  if (!ctx->d_gid) return feasm_fail(ctx, FEASM_ERR_STATE, "feasm_residual on several ranks needs the global ids of feasm_set_dofs");
  int rc = ensure_vplan(ctx, h);
  if (rc) return rc;
  const size_t nb = sizeof(double) * (size_t)(ctx->n_rows ? ctx->n_rows : 1);
  if (!h->d_uloc) FEASM_CUDA(ctx, cudaMalloc(&h->d_uloc, nb));
  if (!h->d_rloc) FEASM_CUDA(ctx, cudaMalloc(&h->d_rloc, nb));
  const double* xf = nullptr;
  rc = feasm_halo_exchange(ctx, d_u, &xf);
  if (rc) return rc;
  if (ctx->n_rows) k_halo_localize<<<blocks(ctx->n_rows), 256, 0, ctx->stream>>>(xf, ctx->d_gid, ctx->n_rows, h->d_uloc, h->n_global);
  rc = feasm_launch_row_residual(ctx, h->d_uloc, h->d_rloc);
  if (rc) return rc;
  if (ctx->n_peers) {
    const uint64_t ns = h->vsend_ptr.back();
    if (ns) k_halo_pack<<<blocks(ns), 256, 0, ctx->stream>>>(h->d_rloc, h->d_vsend_rows, ns, h->d_vsend, (uint64_t)ctx->n_rows);
    rc = feasm_comm_p2p(ctx, ctx->n_peers, ctx->peer_rank.data(), h->vsend_ptr.data(), h->d_vsend, h->vrecv_ptr.data(), h->d_vrecv, 8, ctx->stream);
    if (rc) return rc;
    for (int p = 0; p < ctx->n_peers; ++p) {
      const uint64_t r0 = h->vrecv_ptr[p], n = h->vrecv_ptr[p + 1] - r0;
      if (n) k_halo_add_rows<<<blocks(n), 256, 0, ctx->stream>>>(h->d_vrecv + r0, h->d_vrecv_rows + r0, n, h->d_rloc, (uint64_t)ctx->n_owned);
    }
  }
  FEASM_CUDA(ctx, cudaGetLastError());
  if (ctx->n_owned) FEASM_CUDA(ctx, cudaMemcpyAsync(d_r, h->d_rloc, sizeof(double) * ctx->n_owned, cudaMemcpyDeviceToDevice, ctx->stream));
  return FEASM_OK;
}

extern "C" {

int64_t feasm_halo_columns(feasm_ctx* ctx, uint64_t n_global, uint64_t first, uint32_t* gids) {
  if (!ctx || !ctx->d_row_ptr) return feasm_fail(ctx, FEASM_ERR_STATE, "no pattern");
  if (first + (uint64_t)ctx->n_owned > n_global || n_global > 0xffffffffull) return feasm_fail(ctx, FEASM_ERR_ARG, "owned range outside [0, n_global)");
  if (cudaSetDevice(ctx->device) != cudaSuccess) return feasm_fail(ctx, FEASM_ERR_CUDA, "cudaSetDevice");
  feasm_halo* h = halo_of(ctx);
  if (!h->cols_valid || h->n_global != n_global || h->first != first) {
    uint8_t* d_flag = nullptr;
    uint64_t n_entries = 0;
    FEASM_CUDA(ctx, cudaMalloc(&d_flag, n_global ? n_global : 1));
    auto body = [&]() -> int {
      FEASM_CUDA(ctx, cudaMemsetAsync(d_flag, 0, n_global, ctx->stream));
      FEASM_CUDA(ctx, cudaMemcpyAsync(&n_entries, ctx->d_row_ptr + ctx->n_owned, sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream));
      FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
      if (n_entries) {
        const uint64_t want = (n_entries + 255) / 256, cap = (uint64_t)ctx->sm_count * 16;
        k_halo_mark_cols<<<(unsigned)(want < cap ? want : cap), 256, 0, ctx->stream>>>(ctx->d_col_idx, n_entries, (uint32_t)first,
                                                                                      (uint32_t)(first + ctx->n_owned), n_global, d_flag);
      }
      if (ctx->d_gid && ctx->n_rows > ctx->n_owned)
        k_halo_mark_rows<<<blocks(ctx->n_rows - ctx->n_owned), 256, 0, ctx->stream>>>(ctx->d_gid, ctx->n_owned, ctx->n_rows, n_global, d_flag);
      FEASM_CUDA(ctx, cudaGetLastError());
      std::vector<uint8_t> flag(n_global);
      if (n_global) FEASM_CUDA(ctx, cudaMemcpyAsync(flag.data(), d_flag, n_global, cudaMemcpyDeviceToHost, ctx->stream));
      FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
      h->cols.clear();
      for (uint64_t c = 0; c < n_global; ++c)
        if (flag[c]) h->cols.push_back((uint32_t)c);
      return FEASM_OK;
    };
    const int rc = body();
    cudaFree(d_flag);
    if (rc) return rc;
    h->cols_valid = true;
    h->n_global = n_global;
    h->first = first;
  }
  if (gids && !h->cols.empty()) std::memcpy(gids, h->cols.data(), sizeof(uint32_t) * h->cols.size());
  return (int64_t)h->cols.size();
}

int feasm_set_halo_plan(feasm_ctx* ctx, uint64_t n_global, uint64_t first, int n_peers, const int32_t* peer_rank, const uint64_t* send_ptr,
                        const uint32_t* send_rows, const uint64_t* recv_ptr, const uint32_t* recv_gids) {
  if (!ctx || !ctx->d_row_ptr) return feasm_fail(ctx, FEASM_ERR_STATE, "no pattern");
  if (n_peers < 0 || (n_peers && (!peer_rank || !send_ptr || !recv_ptr))) return feasm_fail(ctx, FEASM_ERR_ARG, "bad halo plan");
  if (first + (uint64_t)ctx->n_owned > n_global || n_global > 0xffffffffull) return feasm_fail(ctx, FEASM_ERR_ARG, "owned range outside [0, n_global)");
  const uint64_t ns = n_peers ? send_ptr[n_peers] : 0, nr = n_peers ? recv_ptr[n_peers] : 0;
  if ((ns && !send_rows) || (nr && !recv_gids)) return feasm_fail(ctx, FEASM_ERR_ARG, "bad halo plan");
  for (uint64_t i = 0; i < ns; ++i)
    if (send_rows[i] >= (uint64_t)ctx->n_owned) return feasm_fail(ctx, FEASM_ERR_ARG, "halo plan: a requested entry is not an owned row");
  for (uint64_t i = 0; i < nr; ++i)
    if (recv_gids[i] >= n_global || (recv_gids[i] >= first && recv_gids[i] < first + (uint64_t)ctx->n_owned))
      return feasm_fail(ctx, FEASM_ERR_ARG, "halo plan: a received entry is owned by this rank or outside the vector");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  feasm_halo* h = halo_of(ctx);
  h->n_global = n_global;
  h->first = first;
  h->n_peers = n_peers;
  h->peer.assign(peer_rank, peer_rank + n_peers);
  h->send_ptr.assign(1, 0);
  h->recv_ptr.assign(1, 0);
  if (n_peers) { h->send_ptr.assign(send_ptr, send_ptr + n_peers + 1); h->recv_ptr.assign(recv_ptr, recv_ptr + n_peers + 1); }
  int rc = put(ctx, &h->d_send_rows, send_rows, ns);
  if (!rc) rc = put(ctx, &h->d_recv_gids, recv_gids, nr);
  if (rc) return rc;
  drop(h->d_send); drop(h->d_recv); drop(h->d_xfull);
  FEASM_CUDA(ctx, cudaMalloc(&h->d_send, sizeof(double) * (ns + 1)));
  FEASM_CUDA(ctx, cudaMalloc(&h->d_recv, sizeof(double) * (nr + 1)));
  FEASM_CUDA(ctx, cudaMalloc(&h->d_xfull, sizeof(double) * (n_global ? n_global : 1)));
  FEASM_CUDA(ctx, cudaMemsetAsync(h->d_xfull, 0, sizeof(double) * n_global, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return FEASM_OK;
}

// Collective over the context's communicator: request lists to the owners (sizes by all-gather, ids by one grouped
// send / recv of uint32 words), then feasm_set_halo_plan.
int feasm_build_halo(feasm_ctx* ctx, const uint64_t* first_dof) {
  if (!ctx || !first_dof) return feasm_fail(ctx, FEASM_ERR_ARG, "first_dof is NULL");
  if (!ctx->comm) return feasm_fail(ctx, FEASM_ERR_STATE, "no communicator: feasm_comm_init / feasm_set_communicator first");
  const int world = ctx->comm_size, rank = ctx->comm_rank;
  const uint64_t n_global = first_dof[world], first = first_dof[rank];
  if (first_dof[rank + 1] - first != (uint64_t)ctx->n_owned) return feasm_fail(ctx, FEASM_ERR_ARG, "first_dof does not match the owned rows of this rank");
  const int64_t nc = feasm_halo_columns(ctx, n_global, first, nullptr);
  if (nc < 0) return (int)nc;
  feasm_halo* h = ctx->halo;
  const std::vector<uint32_t>& cols = h->cols;  // ascending: the requests to one owner are a contiguous run
  std::vector<uint64_t> mine(world, 0), all((size_t)world * world, 0), req_ptr(world + 1, 0);
  for (int q = 0; q < world; ++q) {
    req_ptr[q + 1] = std::lower_bound(cols.begin(), cols.end(), first_dof[q + 1], [](uint32_t c, uint64_t v) { return (uint64_t)c < v; }) - cols.begin();
    mine[q] = req_ptr[q + 1] - req_ptr[q];
  }
  if (mine[rank]) return feasm_fail(ctx, FEASM_ERR_STATE, "halo: an owned column was listed as a ghost");
  int rc = feasm_comm_allgather_u64(ctx, mine.data(), world, all.data());
  if (rc) return rc;
  std::vector<int32_t> peers;
  std::vector<uint64_t> sp(1, 0), rp(1, 0), req_send_ptr(1, 0);  // sp: what peers want from me; rp: what I get
  std::vector<uint32_t> recv_gids;
  for (int q = 0; q < world; ++q) {
    const uint64_t want = all[(size_t)q * world + rank], get = mine[q];
    if (q == rank || (!want && !get)) continue;
    peers.push_back(q);
    sp.push_back(sp.back() + want);
    rp.push_back(rp.back() + get);
    recv_gids.insert(recv_gids.end(), cols.begin() + req_ptr[q], cols.begin() + req_ptr[q + 1]);
  }
  // ship the request lists: I SEND my requests (recv_gids, per peer rp) and RECEIVE the peers' requests (per peer sp)
  uint32_t *d_out = nullptr, *d_in = nullptr;
  std::vector<uint32_t> wanted(sp.back());
  auto body = [&]() -> int {
    FEASM_CUDA(ctx, cudaMalloc(&d_out, sizeof(uint32_t) * (recv_gids.size() + 1)));
    FEASM_CUDA(ctx, cudaMalloc(&d_in, sizeof(uint32_t) * (wanted.size() + 1)));
    if (!recv_gids.empty()) FEASM_CUDA(ctx, cudaMemcpyAsync(d_out, recv_gids.data(), sizeof(uint32_t) * recv_gids.size(), cudaMemcpyHostToDevice, ctx->stream));
    int r2 = feasm_comm_p2p(ctx, (int)peers.size(), peers.data(), rp.data(), d_out, sp.data(), d_in, 4, ctx->stream);
    if (r2) return r2;
    if (!wanted.empty()) FEASM_CUDA(ctx, cudaMemcpyAsync(wanted.data(), d_in, sizeof(uint32_t) * wanted.size(), cudaMemcpyDeviceToHost, ctx->stream));
    FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return FEASM_OK;
  };
  rc = body();
  cudaFree(d_out);
  cudaFree(d_in);
  if (rc) return rc;
  for (auto& g : wanted) {
    if (g < first || g >= first + (uint64_t)ctx->n_owned) return feasm_fail(ctx, FEASM_ERR_COMM, "halo: a peer asked for an entry this rank does not own");
    g -= (uint32_t)first;
  }
  return feasm_set_halo_plan(ctx, n_global, first, (int)peers.size(), peers.data(), sp.data(), wanted.data(), rp.data(), recv_gids.data());
}

double* feasm_halo_pack(feasm_ctx* ctx, const double* d_x, uint64_t* n_doubles) {
  if (n_doubles) *n_doubles = 0;
  if (!ctx || !ctx->halo || !ctx->halo->d_send || !d_x) return nullptr;
  feasm_halo* h = ctx->halo;
  if (cudaSetDevice(ctx->device) != cudaSuccess) return nullptr;
  const uint64_t ns = h->send_ptr.back();
  if (ns) k_halo_pack<<<blocks(ns), 256, 0, ctx->stream>>>(d_x, h->d_send_rows, ns, h->d_send, (uint64_t)ctx->n_owned);
  if (n_doubles) *n_doubles = ns;
  return h->d_send;
}

double* feasm_halo_recv_buffer(feasm_ctx* ctx, uint64_t* n_doubles) {
  if (n_doubles) *n_doubles = 0;
  if (!ctx || !ctx->halo || !ctx->halo->d_recv) return nullptr;
  if (n_doubles) *n_doubles = ctx->halo->recv_ptr.back();
  return ctx->halo->d_recv;
}

int feasm_halo_unpack(feasm_ctx* ctx) {
  if (!ctx || !ctx->halo || !ctx->halo->d_xfull) return feasm_fail(ctx, FEASM_ERR_STATE, "no halo plan");
  feasm_halo* h = ctx->halo;
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  const uint64_t nr = h->recv_ptr.back();
  if (nr) k_halo_unpack<<<blocks(nr), 256, 0, ctx->stream>>>(h->d_recv, h->d_recv_gids, nr, h->d_xfull, h->n_global);
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}

int feasm_get_halo_plan(feasm_ctx* ctx, int32_t* peers, uint64_t* send_ptr, uint32_t* send_rows, uint64_t* recv_ptr, uint32_t* recv_gids) {
  if (!ctx || !ctx->halo || !ctx->halo->d_xfull) return feasm_fail(ctx, FEASM_ERR_STATE, "no halo plan");
  feasm_halo* h = ctx->halo;
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  if (peers && h->n_peers) std::memcpy(peers, h->peer.data(), sizeof(int32_t) * h->n_peers);
  if (send_ptr) std::memcpy(send_ptr, h->send_ptr.data(), sizeof(uint64_t) * h->send_ptr.size());
  if (recv_ptr) std::memcpy(recv_ptr, h->recv_ptr.data(), sizeof(uint64_t) * h->recv_ptr.size());
  if (send_rows && h->send_ptr.back()) FEASM_CUDA(ctx, cudaMemcpy(send_rows, h->d_send_rows, sizeof(uint32_t) * h->send_ptr.back(), cudaMemcpyDeviceToHost));
  if (recv_gids && h->recv_ptr.back()) FEASM_CUDA(ctx, cudaMemcpy(recv_gids, h->d_recv_gids, sizeof(uint32_t) * h->recv_ptr.back(), cudaMemcpyDeviceToHost));
  return h->n_peers;
}

}  // extern "C"
