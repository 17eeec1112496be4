// feasm_general.cu — general linear dof constraints (hanging nodes) at element level.
//
//     u_i = sum_j coef_ij u_j + g_i        one DofConstraintRow per constrained dof (include/base/dof_map.h:97-123)
//
// What the reference does for every element before add_matrix (DofMap::heterogeneously_constrain_element_matrix_and_vector,
// src/base/dof_map_constraints.C:2519-2614, and constrain_element_matrix_and_vector, :2439-2515):
//     C, H <- build_constraint_matrix_and_vector(elem_dofs)      (:3543-3673; elem_dofs grows by the constraining dofs)
//     F <- C^T (F - K H),  K <- C^T K C,  then every constrained row: zero, 1 on the diagonal, -coef at its constraining dofs,
//     rhs = g_i
// Elements whose constrained dofs all have EMPTY rows (Dirichlet) keep the fast path: nothing grows, the element kernel
// constrains them in its shared-memory tile.  Elements that touch a dof with a non-empty row ("widened" elements, the few
// around a refinement interface) take this file's path:
//   setup (host)   the widened dof list, C, H and the constrained-row data of each such element (the recursion of
//                  build_constraint_matrix_and_vector, restated in build_cv below); the sparsity pattern grows by the
//                  couplings DofMap::find_connected_dofs adds (src/base/dof_map.C:2915-2975, sparsity_pattern.C:139);
//                  every entry of every widened element matrix gets its CSR slot, grouped by slot in element order
//   per assembly   the element kernel leaves these elements' Ke / Fe unconstrained in their records;
//                  k_general_transform (one CTA per element) forms C^T K C and C^T (F - K H), fixes the constrained rows,
//                  writes the widened block to a side buffer and zeroes the element's records (the row reduction then adds
//                  zeros for it); after the row reduction the side buffer is added into the CSR values / rhs slot by slot,
//                  each slot in element order (the unpack kernel's fixed-order sum): bitwise reproducible.
#include <algorithm>
#include <map>
#include <set>

#include "feasm_internal.h"

struct feasm_general {
  // the rows as given (host)
  std::map<uint32_t, std::vector<std::pair<uint32_t, double>>> rows;  // non-empty and empty rows alike
  std::map<uint32_t, double> values;
  // widened elements
  std::vector<uint32_t> elems;              // element ids, ascending
  std::vector<uint64_t> dof_ptr;            // [n + 1] into exp_dofs
  std::vector<uint32_t> exp_dofs;
  std::vector<uint64_t> c_ptr, b_ptr;       // [n + 1] into C (nd x n_exp, row-major) and the block buffer (n_exp^2 + n_exp)
  std::vector<double> C, H;                 // H: [n][nd]
  std::vector<uint8_t> conmask;             // per widened dof: constrained
  std::vector<double> gval;                 // per widened dof: g (constrained ones)
  std::vector<uint64_t> fix_ptr;            // [n + 1] into fix_*: entries (row p, column k, value -coef)
  std::vector<uint32_t> fix_p, fix_k;
  std::vector<double> fix_v;
  int max_exp = 0;
  // device copies
  uint32_t* d_elems = nullptr; uint64_t* d_dof_ptr = nullptr; uint64_t* d_c_ptr = nullptr; uint64_t* d_b_ptr = nullptr;
  double* d_C = nullptr; double* d_H = nullptr; uint8_t* d_conmask = nullptr; double* d_gval = nullptr;
  uint64_t* d_fix_ptr = nullptr; uint32_t* d_fix_p = nullptr; uint32_t* d_fix_k = nullptr; double* d_fix_v = nullptr;
  double* d_B = nullptr;
  // scatter plan (feasm_general_prepare): destination slots, groups, permutation into d_B
  bool plan_valid = false;
  uint64_t n_slots = 0;
  uint64_t* d_slot = nullptr; uint64_t* d_seg = nullptr; uint64_t* d_perm = nullptr;
};

namespace {

template <typename T>
void freep(T*& p) { if (p) cudaFree(p); p = nullptr; }

void release_device(feasm_general* g) {
  freep(g->d_elems); freep(g->d_dof_ptr); freep(g->d_c_ptr); freep(g->d_b_ptr); freep(g->d_C); freep(g->d_H); freep(g->d_conmask);
  freep(g->d_gval); freep(g->d_fix_ptr); freep(g->d_fix_p); freep(g->d_fix_k); freep(g->d_fix_v); freep(g->d_B);
  freep(g->d_slot); freep(g->d_seg); freep(g->d_perm);
  g->plan_valid = false;
}

template <typename T>
int up(feasm_ctx* ctx, T** d, const std::vector<T>& h) {
  freep(*d);
  FEASM_CUDA(ctx, cudaMalloc(d, sizeof(T) * (h.size() ? h.size() : 1)));
  if (!h.empty()) FEASM_CUDA(ctx, cudaMemcpyAsync(*d, h.data(), sizeof(T) * h.size(), cudaMemcpyHostToDevice, ctx->stream));
  return FEASM_OK;
}

// build_constraint_matrix_and_vector (dof_map_constraints.C:3543-3673).  dofs grows; Cm is (old size) x (new size), row-major.
// Returns false when the element is not constrained at this level (C left empty).
bool build_cv(const feasm_general& g, std::vector<uint32_t>& dofs, std::vector<double>& Cm, std::vector<double>& H, bool recursive) {
  std::set<uint32_t> dep;
  bool have = false;
  for (uint32_t d : dofs) {
    auto it = g.rows.find(d);
    if (it == g.rows.end()) continue;
    have = true;
    for (auto& e : it->second) dep.insert(e.first);
  }
  if (!have) return false;
  for (uint32_t d : dofs) dep.erase(d);
  if (dep.empty() && recursive) return false;  // recursion goes on only while new dofs appear (:3595-3600)
  const size_t old = dofs.size();
  dofs.insert(dofs.end(), dep.begin(), dep.end());
  const size_t n = dofs.size();
  Cm.assign(old * n, 0.);
  H.assign(old, 0.);
  for (size_t i = 0; i < old; ++i) {
    auto it = g.rows.find(dofs[i]);
    if (it == g.rows.end()) { Cm[i * n + i] = 1.; continue; }
    for (auto& e : it->second)
      for (size_t k = 0; k < n; ++k)
        if (dofs[k] == e.first) Cm[i * n + k] = e.second;
    auto v = g.values.find(dofs[i]);
    if (v != g.values.end()) H[i] = v->second;
  }
  std::vector<double> Cn, Hn;
  if (build_cv(g, dofs, Cn, Hn, true) && Cn.size() == n * dofs.size()) {
    // x = C y + h, y = D z + g  ->  x = (C D) z + (C g + h)   (:3662-3669)
    const size_t n2 = dofs.size();
    for (size_t i = 0; i < old; ++i) {
      double acc = 0.;
      for (size_t k = 0; k < n; ++k) acc += Cm[i * n + k] * Hn[k];
      H[i] += acc;
    }
    std::vector<double> CD(old * n2, 0.);
    for (size_t i = 0; i < old; ++i)
      for (size_t k = 0; k < n; ++k) {
        const double c = Cm[i * n + k];
        if (c == 0.) continue;
        for (size_t q = 0; q < n2; ++q) CD[i * n2 + q] += c * Cn[k * n2 + q];
      }
    Cm.swap(CD);
  }
  return true;
}

__global__ void k_mark_general(const uint32_t* __restrict__ dofs, int64_t n, uint8_t* __restrict__ con_flag) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) con_flag[dofs[i]] = 2;  // constrained with a non-empty row (1: Dirichlet)
}

// One CTA per widened element.  Shared memory: T [nd][n_exp] | Fm [nd].
__global__ void __launch_bounds__(128)
k_general_transform(const uint32_t* __restrict__ elems, const uint64_t* __restrict__ dof_ptr, const uint64_t* __restrict__ c_ptr,
                    const uint64_t* __restrict__ b_ptr, const double* __restrict__ Cg, const double* __restrict__ Hg,
                    const uint8_t* __restrict__ conmask, const double* __restrict__ gval, const uint64_t* __restrict__ fix_ptr,
                    const uint32_t* __restrict__ fix_p, const uint32_t* __restrict__ fix_k, const double* __restrict__ fix_v,
                    double* __restrict__ kf, int rs, int nd, int hetero, double* __restrict__ B) {
  extern __shared__ double sm[];
  const int g = blockIdx.x;
  const uint64_t d0 = dof_ptr[g];
  const int ne = (int)(dof_ptr[g + 1] - d0);
  double* K = kf + (uint64_t)elems[g] * nd * rs;  // record row a: Ke(a, 0..nd-1), Fe(a) at column nd
  const double* C = Cg + c_ptr[g];
  const double* H = Hg + (uint64_t)g * nd;
  double* T = sm;
  double* Fm = sm + (size_t)nd * ne;
  double* out = B + b_ptr[g];
  // F - K H (heterogeneous) or F
  for (int i = threadIdx.x; i < nd; i += blockDim.x) {
    double kh = 0.;
    if (hetero)
      for (int j = 0; j < nd; ++j) kh += K[(size_t)i * rs + j] * H[j];
    Fm[i] = K[(size_t)i * rs + nd] - kh;
  }
  // T = K C
  for (int t = threadIdx.x; t < nd * ne; t += blockDim.x) {
    const int i = t / ne, q = t % ne;
    double acc = 0.;
    for (int j = 0; j < nd; ++j) acc += K[(size_t)i * rs + j] * C[(size_t)j * ne + q];
    T[t] = acc;
  }
  __syncthreads();
  // K' = C^T T, F' = C^T Fm, then the constrained rows (:2570-2612)
  for (int t = threadIdx.x; t < ne * ne; t += blockDim.x) {
    const int p = t / ne, q = t % ne;
    double acc = 0.;
    for (int i = 0; i < nd; ++i) acc += C[(size_t)i * ne + p] * T[i * ne + q];
    if (conmask[d0 + p]) acc = p == q ? 1. : 0.;
    out[t] = acc;
  }
  for (int p = threadIdx.x; p < ne; p += blockDim.x) {
    double acc = 0.;
    for (int i = 0; i < nd; ++i) acc += C[(size_t)i * ne + p] * Fm[i];
    if (conmask[d0 + p]) acc = hetero ? gval[d0 + p] : 0.;
    out[(size_t)ne * ne + p] = acc;
  }
  __syncthreads();
  for (uint64_t f = fix_ptr[g] + threadIdx.x; f < fix_ptr[g + 1]; f += blockDim.x) out[(size_t)fix_p[f] * ne + fix_k[f]] = fix_v[f];
  // the records of this element now hold nothing for the row reduction to add
  for (int t = threadIdx.x; t < nd * (nd + 1); t += blockDim.x) K[(size_t)(t / (nd + 1)) * rs + t % (nd + 1)] = 0.;
}

// the fixed-order add of feasm_scatter.cu's k_unpack, here with the block buffer as the source
__global__ void k_general_add(const uint64_t* __restrict__ uslot, const uint64_t* __restrict__ seg, const uint64_t* __restrict__ perm,
                              uint64_t n_slots, uint64_t nnz, const double* __restrict__ in, double* __restrict__ values,
                              double* __restrict__ rhs) {
  const uint64_t k = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  if (k >= n_slots) return;
  const uint64_t s = uslot[k];
  double* dst = s < nnz ? values + s : rhs + (s - nnz);
  double acc = *dst;
  for (uint64_t i = seg[k]; i < seg[k + 1]; ++i) acc += in[perm[i]];
  *dst = acc;
}

}  // namespace

void feasm_general_release(feasm_ctx* ctx) {
  if (!ctx->gen) return;
  release_device(ctx->gen);
  delete ctx->gen;
  ctx->gen = nullptr;
}

void feasm_general_invalidate_plan(feasm_ctx* ctx) { if (ctx->gen) ctx->gen->plan_valid = false; }

bool feasm_general_active(const feasm_ctx* ctx) { return ctx->gen && !ctx->gen->elems.empty(); }

extern "C" int feasm_set_constraint_rows(feasm_ctx* ctx, int64_t n, const uint32_t* dofs, const uint64_t* row_ptr, const uint32_t* cols,
                                         const double* coefs, const double* values) {
  if (!ctx || !ctx->d_con_flag) return feasm_fail(ctx, FEASM_ERR_STATE, "set_dofs first");
  if (n < 0 || (n && (!dofs || !row_ptr || !values)) || (n && row_ptr[0] != 0)) return feasm_fail(ctx, FEASM_ERR_ARG, "bad constraint rows");
  const uint64_t n_ent = n ? row_ptr[n] : 0;
  if (n_ent && (!cols || !coefs)) return feasm_fail(ctx, FEASM_ERR_ARG, "bad constraint rows");
  for (int64_t i = 0; i < n; ++i) {
    if (dofs[i] >= (uint64_t)ctx->n_rows || row_ptr[i + 1] < row_ptr[i]) return feasm_fail(ctx, FEASM_ERR_ARG, "constraint row out of range");
    for (uint64_t k = row_ptr[i]; k < row_ptr[i + 1]; ++k)
      if (cols[k] >= (uint64_t)ctx->n_rows || cols[k] == dofs[i]) return feasm_fail(ctx, FEASM_ERR_ARG, "constraining dof out of range or equal to its own row");
  }
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  int rc = feasm_set_constraints(ctx, n, dofs, values);  // flags + values of every constrained dof; drops the previous rows
  if (rc) return rc;
  if (n_ent == 0) return FEASM_OK;  // Dirichlet rows only: nothing is widened
  if (ctx->n_owned < ctx->n_rows) return feasm_fail(ctx, FEASM_ERR_STATE, "general constraint rows: one rank per context in this version");
  feasm_general* g = new feasm_general;
  ctx->gen = g;
  std::vector<uint32_t> general;
  for (int64_t i = 0; i < n; ++i) {
    auto& row = g->rows[dofs[i]];
    for (uint64_t k = row_ptr[i]; k < row_ptr[i + 1]; ++k) row.emplace_back(cols[k], coefs[k]);
    if (values[i] != 0.) g->values[dofs[i]] = values[i];
    if (!row.empty()) general.push_back(dofs[i]);
  }
  // flag 2 on the dofs with a non-empty row: k_elem_flags keeps their elements off the Dirichlet fast path
  uint32_t* d_gen = nullptr;
  FEASM_CUDA(ctx, cudaMalloc(&d_gen, sizeof(uint32_t) * general.size()));
  FEASM_CUDA(ctx, cudaMemcpyAsync(d_gen, general.data(), sizeof(uint32_t) * general.size(), cudaMemcpyHostToDevice, ctx->stream));
  k_mark_general<<<(unsigned)((general.size() + 255) / 256), 256, 0, ctx->stream>>>(d_gen, (int64_t)general.size(), ctx->d_con_flag);
  FEASM_CUDA(ctx, cudaGetLastError());
  ctx->elem_flag_valid = false;
  ctx->con_flags_untracked = true;  // the next feasm_set_constraints clears the flags wholesale
  // the elements that touch such a dof, widened
  const int nd = ctx->nd;
  std::vector<uint32_t> ed((size_t)ctx->n_elem * nd);
  FEASM_CUDA(ctx, cudaMemcpyAsync(ed.data(), ctx->d_elem_dofs, sizeof(uint32_t) * ed.size(), cudaMemcpyDeviceToHost, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  cudaFree(d_gen);
  std::vector<uint8_t> is_general((size_t)ctx->n_rows, 0);
  for (uint32_t d : general) is_general[d] = 1;
  g->dof_ptr.assign(1, 0); g->c_ptr.assign(1, 0); g->b_ptr.assign(1, 0); g->fix_ptr.assign(1, 0);
  std::vector<uint32_t> dl;
  std::vector<double> Cm, H;
  for (int64_t e = 0; e < ctx->n_elem; ++e) {
    bool any = false;
    for (int a = 0; a < nd; ++a) any |= is_general[ed[(size_t)e * nd + a]] != 0;
    if (!any) continue;
    dl.assign(ed.begin() + (size_t)e * nd, ed.begin() + (size_t)(e + 1) * nd);
    if (!build_cv(*g, dl, Cm, H, false)) continue;
    const size_t ne = dl.size();
    g->elems.push_back((uint32_t)e);
    const uint64_t d0 = g->exp_dofs.size();
    g->exp_dofs.insert(g->exp_dofs.end(), dl.begin(), dl.end());
    g->C.insert(g->C.end(), Cm.begin(), Cm.end());
    g->H.insert(g->H.end(), H.begin(), H.begin() + nd);
    for (size_t p = 0; p < ne; ++p) {
      auto it = g->rows.find(dl[p]);
      g->conmask.push_back(it != g->rows.end());
      auto v = g->values.find(dl[p]);
      g->gval.push_back(v == g->values.end() ? 0. : v->second);
      if (it != g->rows.end())  // asymmetric_constraint_rows (the default): -coef at the constraining dofs (:2586-2596)
        for (auto& en : it->second)
          for (size_t k = 0; k < ne; ++k)
            if (dl[k] == en.first) { g->fix_p.push_back((uint32_t)p); g->fix_k.push_back((uint32_t)k); g->fix_v.push_back(-en.second); }
    }
    (void)d0;
    g->dof_ptr.push_back(g->exp_dofs.size());
    g->c_ptr.push_back(g->C.size());
    g->b_ptr.push_back(g->b_ptr.back() + ne * ne + ne);
    g->fix_ptr.push_back(g->fix_p.size());
    g->max_exp = std::max(g->max_exp, (int)ne);
  }
  if ((size_t)nd * g->max_exp + nd > 24 * 1024) return feasm_fail(ctx, FEASM_ERR_ARG, "a widened element is too large for the transform kernel");
  if ((rc = up(ctx, &g->d_elems, g->elems)) || (rc = up(ctx, &g->d_dof_ptr, g->dof_ptr)) || (rc = up(ctx, &g->d_c_ptr, g->c_ptr)) ||
      (rc = up(ctx, &g->d_b_ptr, g->b_ptr)) || (rc = up(ctx, &g->d_C, g->C)) || (rc = up(ctx, &g->d_H, g->H)) ||
      (rc = up(ctx, &g->d_conmask, g->conmask)) || (rc = up(ctx, &g->d_gval, g->gval)) || (rc = up(ctx, &g->d_fix_ptr, g->fix_ptr)) ||
      (rc = up(ctx, &g->d_fix_p, g->fix_p)) || (rc = up(ctx, &g->d_fix_k, g->fix_k)) || (rc = up(ctx, &g->d_fix_v, g->fix_v)))
    return rc;
  FEASM_CUDA(ctx, cudaMalloc(&g->d_B, sizeof(double) * (g->b_ptr.back() ? g->b_ptr.back() : 1)));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->prepared = false;  // the scatter plan of the widened blocks is built by feasm_prepare
  return FEASM_OK;
}

// After the device built the pattern from the element dof lists: add the couplings of the widened elements
// (DofMap::find_connected_dofs: every dof of the widened list couples with every other) and hand the union back.
int feasm_general_extend_pattern(feasm_ctx* ctx) {
  if (!feasm_general_active(ctx)) return FEASM_OK;
  feasm_general* g = ctx->gen;
  std::vector<uint64_t> rp((size_t)ctx->n_rows + 1);
  std::vector<uint32_t> ci((size_t)ctx->nnz);
  int rc = feasm_get_pattern(ctx, rp.data(), ci.data());
  if (rc) return rc;
  std::map<uint32_t, std::set<uint32_t>> extra;
  for (size_t k = 0; k + 1 < g->dof_ptr.size(); ++k) {
    const uint32_t* d = g->exp_dofs.data() + g->dof_ptr[k];
    const size_t ne = g->dof_ptr[k + 1] - g->dof_ptr[k];
    for (size_t p = 0; p < ne; ++p) extra[d[p]].insert(d, d + ne);
  }
  std::vector<uint64_t> nrp((size_t)ctx->n_rows + 1, 0);
  std::vector<uint32_t> nci;
  nci.reserve(ci.size() + 64 * extra.size());
  std::vector<uint32_t> tmp;
  bool grew = false;
  for (int64_t r = 0; r < ctx->n_rows; ++r) {
    auto it = extra.find((uint32_t)r);
    if (it == extra.end()) {
      nci.insert(nci.end(), ci.begin() + rp[r], ci.begin() + rp[r + 1]);
    } else {
      tmp.assign(ci.begin() + rp[r], ci.begin() + rp[r + 1]);
      tmp.insert(tmp.end(), it->second.begin(), it->second.end());
      std::sort(tmp.begin(), tmp.end());
      tmp.erase(std::unique(tmp.begin(), tmp.end()), tmp.end());
      grew |= tmp.size() != rp[r + 1] - rp[r];
      nci.insert(nci.end(), tmp.begin(), tmp.end());
    }
    nrp[r + 1] = nci.size();
  }
  if (!grew) return FEASM_OK;
  return feasm_set_pattern(ctx, nrp.data(), nci.data());
}

// The CSR slot of every entry of every widened block, grouped by slot, each group in element order.
int feasm_general_prepare(feasm_ctx* ctx) {
  if (!feasm_general_active(ctx) || ctx->gen->plan_valid) return FEASM_OK;
  feasm_general* g = ctx->gen;
  std::vector<uint64_t> rp((size_t)ctx->n_rows + 1);
  std::vector<uint32_t> ci((size_t)ctx->nnz);
  int rc = feasm_get_pattern(ctx, rp.data(), ci.data());
  if (rc) return rc;
  const uint64_t total = g->b_ptr.back();
  std::vector<uint64_t> dest(total);
  for (size_t k = 0; k + 1 < g->dof_ptr.size(); ++k) {
    const uint32_t* d = g->exp_dofs.data() + g->dof_ptr[k];
    const size_t ne = g->dof_ptr[k + 1] - g->dof_ptr[k];
    const uint64_t b0 = g->b_ptr[k];
    for (size_t p = 0; p < ne; ++p) {
      const uint32_t* rb = ci.data() + rp[d[p]];
      const uint32_t* re = ci.data() + rp[d[p] + 1];
      for (size_t q = 0; q < ne; ++q) {
        const uint32_t* at = std::lower_bound(rb, re, d[q]);
        if (at == re || *at != d[q])
          return feasm_fail(ctx, FEASM_ERR_PATTERN, "the sparsity pattern lacks a coupling of a constrained element (pass libMesh's Graph, or call feasm_build_pattern after feasm_set_constraint_rows)");
        dest[b0 + p * ne + q] = (uint64_t)(at - ci.data());
      }
      dest[b0 + ne * ne + p] = ctx->nnz + d[p];
    }
  }
  std::vector<uint64_t> perm(total);
  for (uint64_t i = 0; i < total; ++i) perm[i] = i;
  std::stable_sort(perm.begin(), perm.end(), [&](uint64_t a, uint64_t b) { return dest[a] < dest[b]; });
  std::vector<uint64_t> slot, seg;
  for (uint64_t i = 0; i < total; ++i)
    if (i == 0 || dest[perm[i]] != dest[perm[i - 1]]) { slot.push_back(dest[perm[i]]); seg.push_back(i); }
  seg.push_back(total);
  g->n_slots = slot.size();
  if ((rc = up(ctx, &g->d_slot, slot)) || (rc = up(ctx, &g->d_seg, seg)) || (rc = up(ctx, &g->d_perm, perm))) return rc;
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  g->plan_valid = true;
  return FEASM_OK;
}

// after the element kernel: widened blocks into the side buffer, their records zeroed
int feasm_general_transform(feasm_ctx* ctx, int rs, int constrain_mode) {
  if (!feasm_general_active(ctx)) return FEASM_OK;
  feasm_general* g = ctx->gen;
  const size_t smem = sizeof(double) * ((size_t)ctx->nd * g->max_exp + ctx->nd);
  if (smem > 48 * 1024) FEASM_CUDA(ctx, cudaFuncSetAttribute(k_general_transform, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  k_general_transform<<<(unsigned)g->elems.size(), 128, smem, ctx->stream>>>(
      g->d_elems, g->d_dof_ptr, g->d_c_ptr, g->d_b_ptr, g->d_C, g->d_H, g->d_conmask, g->d_gval, g->d_fix_ptr, g->d_fix_p, g->d_fix_k,
      g->d_fix_v, ctx->d_kf, rs, ctx->nd, constrain_mode == FEASM_CONSTRAIN_HETEROGENEOUS ? 1 : 0, g->d_B);
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}

// after the row reduction: the side buffer into the CSR values and the rhs
int feasm_general_add(feasm_ctx* ctx) {
  if (!feasm_general_active(ctx)) return FEASM_OK;
  feasm_general* g = ctx->gen;
  if (!g->plan_valid) return feasm_fail(ctx, FEASM_ERR_STATE, "general constraints: feasm_prepare has not run");
  if (g->n_slots)
    k_general_add<<<(unsigned)((g->n_slots + 255) / 256), 256, 0, ctx->stream>>>(g->d_slot, g->d_seg, g->d_perm, g->n_slots, ctx->nnz, g->d_B,
                                                                                  ctx->d_values, ctx->d_rhs);
  FEASM_CUDA(ctx, cudaGetLastError());
  return FEASM_OK;
}
