// feasm_plan.cpp — the setup of a multi-rank assembly, below the C ABI: which rows of this rank belong to another rank,
// what their owners have to add to their own sparsity pattern, and where every exchanged value goes.
//
// Mirrors:
//   * SparsityPattern::Build::parallel_sync (src/base/sparsity_pattern.C:446-575): the column lists of the rows a rank
//     touches but does not own are shipped to the owner, which unites them with its own;
//   * DistributedVector::close / PETSc's MatAssemblyBegin-End stash (src/numerics/distributed_vector.C:399-470): the
//     values of those rows follow the same route every assembly.  Here the route is fixed once as a slot plan
//     (feasm_set_exchange_plan), so that a step is pack -> one grouped send / recv per neighbour -> unpack.
//
// Two layers:
//   feasm_planner_*       host only (no device, no communicator): messages out, messages in, merged pattern and slot plan.
//                         The host carries the messages with whatever it has (libMesh: Parallel::push_parallel_vector_data;
//                         the CPU tests: gloo).
//   feasm_build_exchange  the same over the context's NCCL communicator, from the pattern on the device to a prepared
//                         context: what a multi-GPU host calls after feasm_build_pattern.
// Local numbering (the host's job, gpu_assemble.C / multi.py): owned rows first (global id - first_dof[rank]), then the
// ghost rows in ascending global id; column ids of the pattern are GLOBAL dof ids.
#include <algorithm>
#include <cstdlib>
#include <map>
#include <string>
#include <vector>

#include "feasm_internal.h"

struct feasm_planner {
  int rank = 0, world = 1;
  std::vector<uint64_t> first_dof;
  uint64_t n_owned = 0, n_rows = 0;
  std::vector<uint32_t> gid;                      // [n_rows - n_owned] global ids of the ghost rows, ascending
  std::vector<uint64_t> rp;                       // pattern: local at first, merged after finish
  std::vector<uint32_t> ci;
  std::map<int, std::vector<uint32_t>> ghost_rows;  // owner -> local ghost rows it owns (ascending)
  std::map<int, std::vector<uint32_t>> out_msg, in_msg;
  bool finished = false, merged = false;
  std::vector<int32_t> peers;
  std::vector<uint64_t> send_ptr, recv_ptr, send_slots, recv_slots;
  std::string err;
};

namespace {

int plan_fail(feasm_planner* p, int code, const std::string& msg) {
  if (p) p->err = msg;
  return code;
}

// message to an owner: n | global row ids [n] | row lengths [n] | columns (global ids), rows in ascending global id
void build_messages(feasm_planner* p) {
  const uint64_t f0 = p->first_dof[p->rank];
  (void)f0;
  for (uint64_t g = 0; g < p->n_rows - p->n_owned; ++g) {
    const uint64_t id = p->gid[g];
    const int owner = (int)(std::upper_bound(p->first_dof.begin(), p->first_dof.end(), id) - p->first_dof.begin()) - 1;
    p->ghost_rows[owner].push_back((uint32_t)(p->n_owned + g));
  }
  for (auto& kv : p->ghost_rows) {
    const std::vector<uint32_t>& rows = kv.second;
    uint64_t tot = 0;
    for (uint32_t r : rows) tot += p->rp[r + 1] - p->rp[r];
    std::vector<uint32_t>& m = p->out_msg[kv.first];
    m.reserve(1 + 2 * rows.size() + tot);
    m.push_back((uint32_t)rows.size());
    for (uint32_t r : rows) m.push_back(p->gid[r - p->n_owned]);
    for (uint32_t r : rows) m.push_back((uint32_t)(p->rp[r + 1] - p->rp[r]));
    for (uint32_t r : rows) m.insert(m.end(), p->ci.begin() + p->rp[r], p->ci.begin() + p->rp[r + 1]);
  }
}

}  // namespace

extern "C" {

feasm_planner* feasm_planner_create(int rank, int world, const uint64_t* first_dof, uint64_t n_owned, uint64_t n_rows,
                                    const uint32_t* global_ids, const uint64_t* row_ptr, const uint32_t* col_idx) {
  if (world < 1 || rank < 0 || rank >= world || !first_dof || !row_ptr || n_owned > n_rows || (n_rows > n_owned && !global_ids)) return nullptr;
  if (first_dof[rank + 1] - first_dof[rank] != n_owned) return nullptr;
  for (int q = 0; q < world; ++q) if (first_dof[q + 1] < first_dof[q]) return nullptr;
  feasm_planner* p = new feasm_planner;
  p->rank = rank; p->world = world;
  p->first_dof.assign(first_dof, first_dof + world + 1);
  p->n_owned = n_owned; p->n_rows = n_rows;
  for (uint64_t r = n_owned; r < n_rows; ++r) {
    const uint32_t g = global_ids[r];
    const bool mine = g >= first_dof[rank] && g < first_dof[rank + 1];
    if (mine || g >= first_dof[world] || (r > n_owned && g <= global_ids[r - 1])) { delete p; return nullptr; }
    p->gid.push_back(g);
  }
  p->rp.assign(row_ptr, row_ptr + n_rows + 1);
  p->ci.assign(col_idx, col_idx + row_ptr[n_rows]);
  build_messages(p);
  return p;
}

void feasm_planner_destroy(feasm_planner* p) { delete p; }
const char* feasm_planner_error(const feasm_planner* p) { return p ? p->err.c_str() : "planner is NULL (bad arguments to feasm_planner_create)"; }

// Length (uint32 words) of the message for `peer` (0: none) and its words.
uint64_t feasm_planner_message(feasm_planner* p, int peer, const uint32_t** words) {
  if (!p) return 0;
  auto it = p->out_msg.find(peer);
  if (it == p->out_msg.end()) { if (words) *words = nullptr; return 0; }
  if (words) *words = it->second.data();
  return it->second.size();
}

int feasm_planner_receive(feasm_planner* p, int peer, const uint32_t* words, uint64_t n) {
  if (!p) return FEASM_ERR_ARG;
  if (p->finished) return plan_fail(p, FEASM_ERR_STATE, "planner already finished");
  if (peer < 0 || peer >= p->world || peer == p->rank) return plan_fail(p, FEASM_ERR_ARG, "bad peer");
  if (n == 0) return FEASM_OK;
  if (!words || n < 1 || n < 1 + 2 * (uint64_t)words[0]) return plan_fail(p, FEASM_ERR_ARG, "truncated message");
  const uint64_t nr = words[0], f0 = p->first_dof[p->rank];
  uint64_t tot = 0;
  for (uint64_t i = 0; i < nr; ++i) {
    const uint64_t g = words[1 + i];
    if (g < f0 || g >= f0 + p->n_owned) return plan_fail(p, FEASM_ERR_ARG, "peer " + std::to_string(peer) + " sent a row this rank does not own");
    tot += words[1 + nr + i];
  }
  if (n != 1 + 2 * nr + tot) return plan_fail(p, FEASM_ERR_ARG, "message length does not match its row lengths");
  p->in_msg[peer].assign(words, words + n);
  return FEASM_OK;
}

// Unites the received column lists with the owned rows and fixes the slot plan: per peer (ascending rank), the value slots
// of its rows (row-major) followed by the rows' right-hand-side slots (nnz + row) — send side: this rank's ghost rows the
// peer owns; receive side: the peer's rows in the order of its message, each column at its place in the merged owned row.
int feasm_planner_finish(feasm_planner* p) {
  if (!p) return FEASM_ERR_ARG;
  if (p->finished) return FEASM_OK;
  const uint64_t f0 = p->first_dof[p->rank];
  struct Seg { uint32_t row; const uint32_t* cols; uint32_t len; };
  std::vector<Seg> segs;
  for (auto& kv : p->in_msg) {
    const uint32_t* w = kv.second.data();
    const uint64_t nr = w[0];
    const uint32_t* cols = w + 1 + 2 * nr;
    for (uint64_t i = 0; i < nr; ++i) {
      const uint32_t len = w[1 + nr + i];
      segs.push_back({(uint32_t)(w[1 + i] - f0), cols, len});
      cols += len;
    }
  }
  if (!segs.empty()) {
    std::stable_sort(segs.begin(), segs.end(), [](const Seg& a, const Seg& b) { return a.row < b.row; });
    // united column list of every affected row
    std::vector<uint64_t> new_len(p->n_rows);
    for (uint64_t r = 0; r < p->n_rows; ++r) new_len[r] = p->rp[r + 1] - p->rp[r];
    std::vector<uint32_t> pool;              // united lists back to back
    std::vector<uint64_t> pool_off;          // per affected row (in segs order): offset into pool
    std::vector<uint32_t> aff_rows;
    std::vector<uint32_t> tmp;
    bool grew = false;
    for (size_t i = 0; i < segs.size();) {
      const uint32_t r = segs[i].row;
      tmp.assign(p->ci.begin() + p->rp[r], p->ci.begin() + p->rp[r + 1]);
      size_t j = i;
      for (; j < segs.size() && segs[j].row == r; ++j) tmp.insert(tmp.end(), segs[j].cols, segs[j].cols + segs[j].len);
      std::sort(tmp.begin(), tmp.end());
      tmp.erase(std::unique(tmp.begin(), tmp.end()), tmp.end());
      if (tmp.size() != new_len[r]) grew = true;
      new_len[r] = tmp.size();
      aff_rows.push_back(r);
      pool_off.push_back(pool.size());
      pool.insert(pool.end(), tmp.begin(), tmp.end());
      i = j;
    }
    pool_off.push_back(pool.size());
    if (grew) {
      std::vector<uint64_t> nrp(p->n_rows + 1, 0);
      for (uint64_t r = 0; r < p->n_rows; ++r) nrp[r + 1] = nrp[r] + new_len[r];
      std::vector<uint32_t> nci(nrp[p->n_rows]);
      size_t a = 0;
      for (uint64_t r = 0; r < p->n_rows; ++r) {
        if (a < aff_rows.size() && aff_rows[a] == r) {
          std::copy(pool.begin() + pool_off[a], pool.begin() + pool_off[a + 1], nci.begin() + nrp[r]);
          ++a;
        } else {
          std::copy(p->ci.begin() + p->rp[r], p->ci.begin() + p->rp[r + 1], nci.begin() + nrp[r]);
        }
      }
      p->rp.swap(nrp);
      p->ci.swap(nci);
      p->merged = true;
    }
  }
  const uint64_t nnz = p->rp[p->n_rows];
  std::vector<int> peers;
  for (auto& kv : p->ghost_rows) peers.push_back(kv.first);
  for (auto& kv : p->in_msg) peers.push_back(kv.first);
  std::sort(peers.begin(), peers.end());
  peers.erase(std::unique(peers.begin(), peers.end()), peers.end());
  p->peers.assign(peers.begin(), peers.end());
  p->send_ptr.assign(1, 0); p->recv_ptr.assign(1, 0);
  p->send_slots.clear(); p->recv_slots.clear();
  for (int q : peers) {
    auto gs = p->ghost_rows.find(q);
    if (gs != p->ghost_rows.end()) {
      for (uint32_t r : gs->second) for (uint64_t k = p->rp[r]; k < p->rp[r + 1]; ++k) p->send_slots.push_back(k);
      for (uint32_t r : gs->second) p->send_slots.push_back(nnz + r);
    }
    p->send_ptr.push_back(p->send_slots.size());
    auto im = p->in_msg.find(q);
    if (im != p->in_msg.end()) {
      const uint32_t* w = im->second.data();
      const uint64_t nr = w[0];
      const uint32_t* cols = w + 1 + 2 * nr;
      for (uint64_t i = 0; i < nr; ++i) {
        const uint64_t r = w[1 + i] - f0;
        const uint32_t* b = p->ci.data() + p->rp[r];
        const uint32_t* e = p->ci.data() + p->rp[r + 1];
        for (uint32_t k = 0; k < w[1 + nr + i]; ++k) {
          const uint32_t* at = std::lower_bound(b, e, cols[k]);
          if (at == e || *at != cols[k]) return plan_fail(p, FEASM_ERR_PATTERN, "received column missing from the merged pattern");
          p->recv_slots.push_back((uint64_t)(at - p->ci.data()));
        }
        cols += w[1 + nr + i];
      }
      for (uint64_t i = 0; i < nr; ++i) p->recv_slots.push_back(nnz + (w[1 + i] - f0));
    }
    p->recv_ptr.push_back(p->recv_slots.size());
  }
  p->finished = true;
  return FEASM_OK;
}

// sizes[0..4] = nnz of the (merged) pattern, number of peers, send slots, receive slots, 1 if the pattern grew
int feasm_planner_sizes(const feasm_planner* p, uint64_t sizes[5]) {
  if (!p || !sizes || !p->finished) return FEASM_ERR_STATE;
  sizes[0] = p->rp[p->n_rows]; sizes[1] = p->peers.size(); sizes[2] = p->send_slots.size(); sizes[3] = p->recv_slots.size();
  sizes[4] = p->merged ? 1 : 0;
  return FEASM_OK;
}

// any pointer may be NULL
int feasm_planner_get(const feasm_planner* p, uint64_t* row_ptr, uint32_t* col_idx, int32_t* peers, uint64_t* send_ptr,
                      uint64_t* send_slots, uint64_t* recv_ptr, uint64_t* recv_slots) {
  if (!p || !p->finished) return FEASM_ERR_STATE;
  if (row_ptr) std::copy(p->rp.begin(), p->rp.end(), row_ptr);
  if (col_idx) std::copy(p->ci.begin(), p->ci.end(), col_idx);
  if (peers) std::copy(p->peers.begin(), p->peers.end(), peers);
  if (send_ptr) std::copy(p->send_ptr.begin(), p->send_ptr.end(), send_ptr);
  if (send_slots) std::copy(p->send_slots.begin(), p->send_slots.end(), send_slots);
  if (recv_ptr) std::copy(p->recv_ptr.begin(), p->recv_ptr.end(), recv_ptr);
  if (recv_slots) std::copy(p->recv_slots.begin(), p->recv_slots.end(), recv_slots);
  return FEASM_OK;
}

// Hands a finished plan to the context: merged pattern (SparseMatrix::update_sparsity_pattern route), slot plan, scatter maps.
int feasm_apply_plan(feasm_ctx* ctx, const feasm_planner* p) {
  if (!ctx || !p) return FEASM_ERR_ARG;
  if (!p->finished) return feasm_fail(ctx, FEASM_ERR_STATE, "feasm_planner_finish first");
  if ((uint64_t)ctx->n_rows != p->n_rows || (uint64_t)ctx->n_owned != p->n_owned) return feasm_fail(ctx, FEASM_ERR_ARG, "plan is for other row counts");
  int rc = FEASM_OK;
  if (p->merged || !ctx->d_row_ptr || ctx->nnz != p->rp[p->n_rows]) {
    rc = feasm_set_pattern(ctx, p->rp.data(), p->ci.data());
    if (rc) return rc;
  }
  rc = feasm_set_exchange_plan(ctx, (int)p->peers.size(), p->peers.data(), p->send_ptr.data(), p->send_slots.data(), p->recv_ptr.data(),
                               p->recv_slots.data());
  if (rc) return rc;
  return feasm_prepare(ctx);
}

// The plan the context holds: returns the number of peers; peers[n], send_ptr[n+1], send_slots, recv_ptr[n+1], recv_slots
// are copied out when non-NULL.
int feasm_get_exchange_plan(feasm_ctx* ctx, int32_t* peers, uint64_t* send_ptr, uint64_t* send_slots, uint64_t* recv_ptr, uint64_t* recv_slots) {
  if (!ctx) return FEASM_ERR_ARG;
  if (peers) std::copy(ctx->peer_rank.begin(), ctx->peer_rank.end(), peers);
  if (send_ptr) std::copy(ctx->send_ptr.begin(), ctx->send_ptr.end(), send_ptr);
  if (recv_ptr) std::copy(ctx->recv_ptr.begin(), ctx->recv_ptr.end(), recv_ptr);
  const uint64_t ns = ctx->n_peers ? ctx->send_ptr[ctx->n_peers] : 0, nr = ctx->n_peers ? ctx->recv_ptr[ctx->n_peers] : 0;
  if (send_slots && ns) FEASM_CUDA(ctx, cudaMemcpy(send_slots, ctx->d_send_slots, sizeof(uint64_t) * ns, cudaMemcpyDeviceToHost));
  if (recv_slots && nr) FEASM_CUDA(ctx, cudaMemcpy(recv_slots, ctx->d_recv_slots, sizeof(uint64_t) * nr, cudaMemcpyDeviceToHost));
  return ctx->n_peers;
}

}  // extern "C"
