// feasm_side.cu — side (face) terms of the assemble loop: FE::reinit(elem, side) and the boundary integral
//   Fe_var[i](a) += JxW_face[qp] * (g_i * phi_face[a][qp])
// of systems_of_equations_ex6.C:240-257 (a traction; with one variable, a Neumann flux), added to the Fe column of the
// element's records after the volume kernel and before the element-level constraint pass — the place the reference
// adds it.  SURVEY 8f rank 3; HEX8 / HEX27 (QUAD4 / QUAD9 sides) and TET10 (TRI6 sides, tables for four sides).
//
// What FE::reinit(elem, side) computes (src/fe/fe_boundary.C:109-240) splits into a part that depends on the element
// TYPE and the side number only — the side's Gauss rule, its map functions psi and the parent's shape functions at the
// side points carried into the parent's reference space (FE::side_map :347-400) — built on the host once
// (feasm_make_side_tables) — and the part that depends on the element's coordinates: FEMap::compute_face_map
// (:885-1017), JxW = sqrt(g11 g22 - g12 g21) w with g the first fundamental form of the side map.  The kernel does the
// second part.  Loaded sides are few (a boundary face), so one warp per loaded element is plenty: lane = side
// quadrature point for the map, lane = shape function for the sum, which runs over qp in the reference's order.
#include <algorithm>

#include "feasm_internal.h"

namespace {

constexpr int SIDE_WARPS = 4;

__global__ void k_side_loads(const uint32_t* __restrict__ elems, const uint8_t* __restrict__ masks, int n_list, int64_t e0, int64_t e1,
                             const double* __restrict__ xyz, const uint32_t* __restrict__ conn, int npe, int n_vars, int ns, int nqf,
                             const double* __restrict__ tab, const int* __restrict__ side_nodes, double g0, double g1, double g2,
                             double* __restrict__ kf, int rs, int* flags) {
  __shared__ double s_x[SIDE_WARPS][27 * 3];
  __shared__ double s_jw[SIDE_WARPS][9];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int w = blockIdx.x * SIDE_WARPS + warp;
  if (w >= n_list) return;
  const int64_t e = elems[w];
  if (e < e0 || e >= e1) return;
  const unsigned mask = masks[w];
  double* X = s_x[warp];
  double* jw = s_jw[warp];
  if (lane < npe) {
    const uint64_t node = conn[e * npe + lane];
    X[3 * lane] = xyz[3 * node]; X[3 * lane + 1] = xyz[3 * node + 1]; X[3 * lane + 2] = xyz[3 * node + 2];
  }
  __syncwarp();
  const double* qw = tab;
  const double* dxi = tab + nqf;
  const double* deta = dxi + ns * nqf;
  const double* phi = deta + ns * nqf;
  const int nd = npe * n_vars;
  double* rec = kf + (uint64_t)(e - e0) * nd * rs;
  const double g[3] = {g0, g1, g2};
  for (int s = 0; s < 6; ++s) {
    if (!(mask >> s & 1)) continue;
    if (lane < nqf) {  // FEMap::compute_face_map, 3D case: tangents of the side map at side point `lane`
      double a0 = 0., a1 = 0., a2 = 0., b0 = 0., b1 = 0., b2 = 0.;
      for (int i = 0; i < ns; ++i) {
        const double* P = X + 3 * side_nodes[s * ns + i];
        const double dx = dxi[i * nqf + lane], de = deta[i * nqf + lane];
        a0 += dx * P[0]; a1 += dx * P[1]; a2 += dx * P[2];
        b0 += de * P[0]; b1 += de * P[1]; b2 += de * P[2];
      }
      const double g11 = a0 * a0 + a1 * a1 + a2 * a2, g12 = a0 * b0 + a1 * b1 + a2 * b2, g22 = b0 * b0 + b1 * b1 + b2 * b2;
      const double jac = sqrt(g11 * g22 - g12 * g12);
      if (!(jac > 0.)) atomicAdd(&flags[0], 1);  // libmesh_assert_greater(the_jac, 0.)
      jw[lane] = jac * qw[lane];
    }
    __syncwarp();
    if (lane < npe) {
      const double* ph = phi + ((size_t)s * npe + lane) * nqf;
      for (int v = 0; v < n_vars; ++v) {
        double* dst = rec + (size_t)(v * npe + lane) * rs + nd;
        double acc = *dst;
        for (int qp = 0; qp < nqf; ++qp) acc += jw[qp] * (g[v] * ph[qp]);
        *dst = acc;
      }
    }
    __syncwarp();
  }
}

// introduction_ex3/exact_solution.C:40-47
__device__ __forceinline__ double exact_solution_dev(double x, double y, double z) {
  const double hp = .5 * 3.141592653589793;
  return __dmul_rn(__dmul_rn(cos(__dmul_rn(hp, x)), sin(__dmul_rn(hp, y))), cos(__dmul_rn(hp, z)));
}

// Dirichlet data by the penalty method: introduction_ex3.C:437-453 on the listed sides,
//   Ke(i,j) += JxW_face[qp]*penalty*phi_face[i][qp]*phi_face[j][qp];   Fe(i) += JxW_face[qp]*penalty*value*phi_face[i][qp];
// added to the element's records where the reference adds it (after the volume terms, before any constraint).  One warp
// per listed element: lane = side point for the face map and the boundary value, then lane = column j (rows one after
// the other) for Ke and lane = i for Fe; the sums run over qp in the reference's order with its left-to-right products.
__global__ void k_side_penalty(const uint32_t* __restrict__ elems, const uint8_t* __restrict__ masks, int n_list, int64_t e0, int64_t e1,
                               const double* __restrict__ xyz, const uint32_t* __restrict__ conn, int npe, int n_vars, int ns, int nqf,
                               const double* __restrict__ tab, const int* __restrict__ side_nodes, double penalty, int kind, double value,
                               double* __restrict__ kf, int rs, int* flags) {
  __shared__ double s_x[SIDE_WARPS][27 * 3];
  __shared__ double s_jw[SIDE_WARPS][9];
  __shared__ double s_val[SIDE_WARPS][9];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int w = blockIdx.x * SIDE_WARPS + warp;
  if (w >= n_list) return;
  const int64_t e = elems[w];
  if (e < e0 || e >= e1) return;
  const unsigned mask = masks[w];
  double* X = s_x[warp];
  double* jw = s_jw[warp];
  double* val = s_val[warp];
  if (lane < npe) {
    const uint64_t node = conn[e * npe + lane];
    X[3 * lane] = xyz[3 * node]; X[3 * lane + 1] = xyz[3 * node + 1]; X[3 * lane + 2] = xyz[3 * node + 2];
  }
  __syncwarp();
  const double* qw = tab;
  const double* dxi = tab + nqf;
  const double* deta = dxi + ns * nqf;
  const double* phi = deta + ns * nqf;
  const double* psi = phi + (size_t)6 * npe * nqf;
  const int nd = npe * n_vars;
  double* rec = kf + (uint64_t)(e - e0) * nd * rs;
  for (int s = 0; s < 6; ++s) {
    if (!(mask >> s & 1)) continue;
    if (lane < nqf) {  // FEMap::compute_face_map (fe_boundary.C:885-1017): side point, tangents, JxW
      double a0 = 0., a1 = 0., a2 = 0., b0 = 0., b1 = 0., b2 = 0., x = 0., y = 0., z = 0.;
      for (int i = 0; i < ns; ++i) {
        const double* P = X + 3 * side_nodes[s * ns + i];
        const double dx = dxi[i * nqf + lane], de = deta[i * nqf + lane], ps = psi[i * nqf + lane];
        x += ps * P[0]; y += ps * P[1]; z += ps * P[2];
        a0 += dx * P[0]; a1 += dx * P[1]; a2 += dx * P[2];
        b0 += de * P[0]; b1 += de * P[1]; b2 += de * P[2];
      }
      const double g11 = a0 * a0 + a1 * a1 + a2 * a2, g12 = a0 * b0 + a1 * b1 + a2 * b2, g22 = b0 * b0 + b1 * b1 + b2 * b2;
      const double jac = sqrt(g11 * g22 - g12 * g12);
      if (!(jac > 0.)) atomicAdd(&flags[0], 1);
      jw[lane] = jac * qw[lane];
      val[lane] = kind == FEASM_PENALTY_EX4_EXACT ? exact_solution_dev(x, y, z) : value;
    }
    __syncwarp();
    if (lane < npe) {
      const double* ph_l = phi + ((size_t)s * npe + lane) * nqf;
      for (int v = 0; v < n_vars; ++v) {
        for (int i = 0; i < npe; ++i) {
          const double* ph_i = phi + ((size_t)s * npe + i) * nqf;
          double* dst = rec + (size_t)(v * npe + i) * rs + v * npe + lane;
          double acc = *dst;
          for (int qp = 0; qp < nqf; ++qp)
            acc = __dadd_rn(acc, __dmul_rn(__dmul_rn(__dmul_rn(jw[qp], penalty), ph_i[qp]), ph_l[qp]));
          *dst = acc;
        }
        double* dst = rec + (size_t)(v * npe + lane) * rs + nd;
        double acc = *dst;
        for (int qp = 0; qp < nqf; ++qp) acc = __dadd_rn(acc, __dmul_rn(__dmul_rn(__dmul_rn(jw[qp], penalty), val[qp]), ph_l[qp]));
        *dst = acc;
      }
    }
    __syncwarp();
  }
}

}  // namespace

static int ensure_side_tables(feasm_ctx* ctx, int qorder) {
  if (ctx->d_side_tab && ctx->side_qorder == qorder) return FEASM_OK;
  FeasmSideTables t;
  std::string err;
  if (!feasm_make_side_tables(ctx->elem_type, qorder, t, err)) return feasm_fail(ctx, FEASM_ERR_ARG, err);
  std::vector<double> flat;
  flat.insert(flat.end(), t.qw.begin(), t.qw.end());
  flat.insert(flat.end(), t.dxi.begin(), t.dxi.end());
  flat.insert(flat.end(), t.deta.begin(), t.deta.end());
  flat.insert(flat.end(), t.phi.begin(), t.phi.end());
  flat.insert(flat.end(), t.psi.begin(), t.psi.end());
  if (ctx->d_side_tab) { cudaFree(ctx->d_side_tab); ctx->d_side_tab = nullptr; }
  if (ctx->d_side_nodes) { cudaFree(ctx->d_side_nodes); ctx->d_side_nodes = nullptr; }
  FEASM_CUDA(ctx, cudaMalloc(&ctx->d_side_tab, sizeof(double) * flat.size()));
  FEASM_CUDA(ctx, cudaMalloc(&ctx->d_side_nodes, sizeof(int) * t.side_nodes.size()));
  FEASM_CUDA(ctx, cudaMemcpyAsync(ctx->d_side_tab, flat.data(), sizeof(double) * flat.size(), cudaMemcpyHostToDevice, ctx->stream));
  FEASM_CUDA(ctx, cudaMemcpyAsync(ctx->d_side_nodes, t.side_nodes.data(), sizeof(int) * t.side_nodes.size(), cudaMemcpyHostToDevice, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // the host vectors go away
  ctx->side_qorder = qorder;
  ctx->side_ns = t.ns;
  ctx->side_nqf = t.nqf;
  return FEASM_OK;
}

int feasm_launch_side_loads(feasm_ctx* ctx, int qorder, int64_t e0, int64_t e1, double* d_kf, int rs) {
  if (ctx->n_side_elems == 0 && ctx->n_pen_elems == 0) return FEASM_OK;
  const int rc = ensure_side_tables(ctx, qorder);
  if (rc) return rc;
  if (ctx->n_side_elems > 0) {
    const unsigned grid = (unsigned)((ctx->n_side_elems + SIDE_WARPS - 1) / SIDE_WARPS);
    k_side_loads<<<grid, SIDE_WARPS * 32, 0, ctx->stream>>>(ctx->d_side_elems, ctx->d_side_mask, (int)ctx->n_side_elems, e0, e1, ctx->d_xyz,
                                                            ctx->d_conn, ctx->npe, ctx->n_vars, ctx->side_ns, ctx->side_nqf, ctx->d_side_tab,
                                                            ctx->d_side_nodes, ctx->side_g[0], ctx->side_g[1], ctx->side_g[2], d_kf, rs,
                                                            ctx->d_flags);
    FEASM_CUDA(ctx, cudaGetLastError());
  }
  if (ctx->n_pen_elems > 0) {
    const unsigned grid = (unsigned)((ctx->n_pen_elems + SIDE_WARPS - 1) / SIDE_WARPS);
    k_side_penalty<<<grid, SIDE_WARPS * 32, 0, ctx->stream>>>(ctx->d_pen_elems, ctx->d_pen_mask, (int)ctx->n_pen_elems, e0, e1, ctx->d_xyz,
                                                              ctx->d_conn, ctx->npe, ctx->n_vars, ctx->side_ns, ctx->side_nqf, ctx->d_side_tab,
                                                              ctx->d_side_nodes, ctx->penalty, ctx->pen_kind, ctx->pen_value, d_kf, rs,
                                                              ctx->d_flags);
    FEASM_CUDA(ctx, cudaGetLastError());
  }
  return FEASM_OK;
}

// (element, side) list -> one entry per element with the side bits merged (two listed sides of one element must not race
// on its records), ascending element ids
static int upload_side_list(feasm_ctx* ctx, int64_t n, const uint32_t* elems, const uint8_t* sides, uint32_t** d_elems, uint8_t** d_mask,
                            int64_t* n_out) {
  std::vector<std::pair<uint32_t, uint8_t>> es((size_t)n);
  for (int64_t k = 0; k < n; ++k) {
    if ((int64_t)elems[k] >= ctx->n_elem || sides[k] >= (ctx->elem_type == FEASM_TET10 ? 4 : 6))
      return feasm_fail(ctx, FEASM_ERR_ARG, "side list: element or side out of range");
    es[(size_t)k] = {elems[k], (uint8_t)(1u << sides[k])};
  }
  std::sort(es.begin(), es.end());
  std::vector<uint32_t> ue;
  std::vector<uint8_t> um;
  for (const auto& p : es) {
    if (!ue.empty() && ue.back() == p.first) um.back() |= p.second;
    else { ue.push_back(p.first); um.push_back(p.second); }
  }
  FEASM_CUDA(ctx, cudaMalloc(d_elems, sizeof(uint32_t) * ue.size()));
  FEASM_CUDA(ctx, cudaMalloc(d_mask, ue.size()));
  FEASM_CUDA(ctx, cudaMemcpy(*d_elems, ue.data(), sizeof(uint32_t) * ue.size(), cudaMemcpyHostToDevice));
  FEASM_CUDA(ctx, cudaMemcpy(*d_mask, um.data(), um.size(), cudaMemcpyHostToDevice));
  *n_out = (int64_t)ue.size();
  return FEASM_OK;
}

extern "C" int feasm_set_side_penalty(feasm_ctx* ctx, int64_t n, const uint32_t* elems, const uint8_t* sides, double penalty, int kind,
                                      double value) {
  if (!ctx) return FEASM_ERR_ARG;
  if (!ctx->d_conn) return feasm_fail(ctx, FEASM_ERR_STATE, "set_mesh first");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // a running assembly may still read the old list
  if (ctx->d_pen_elems) { cudaFree(ctx->d_pen_elems); ctx->d_pen_elems = nullptr; }
  if (ctx->d_pen_mask) { cudaFree(ctx->d_pen_mask); ctx->d_pen_mask = nullptr; }
  ctx->n_pen_elems = 0;
  if (n <= 0) return FEASM_OK;
  if (!elems || !sides || (kind != FEASM_PENALTY_CONSTANT && kind != FEASM_PENALTY_EX4_EXACT)) return feasm_fail(ctx, FEASM_ERR_ARG, "bad side penalty arguments");
  const int rc = upload_side_list(ctx, n, elems, sides, &ctx->d_pen_elems, &ctx->d_pen_mask, &ctx->n_pen_elems);
  if (rc) return rc;
  ctx->penalty = penalty; ctx->pen_kind = kind; ctx->pen_value = value;
  return FEASM_OK;
}

extern "C" int feasm_set_side_loads(feasm_ctx* ctx, int64_t n, const uint32_t* elems, const uint8_t* sides, const double* g, int n_g) {
  if (!ctx) return FEASM_ERR_ARG;
  if (!ctx->d_conn) return feasm_fail(ctx, FEASM_ERR_STATE, "set_mesh first");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // a running assembly may still read the old list
  if (ctx->d_side_elems) { cudaFree(ctx->d_side_elems); ctx->d_side_elems = nullptr; }
  if (ctx->d_side_mask) { cudaFree(ctx->d_side_mask); ctx->d_side_mask = nullptr; }
  ctx->n_side_elems = 0;
  if (n <= 0) return FEASM_OK;
  if (!elems || !sides || !g || n_g < 1 || n_g > 3) return feasm_fail(ctx, FEASM_ERR_ARG, "bad side load arguments");
  const int rc = upload_side_list(ctx, n, elems, sides, &ctx->d_side_elems, &ctx->d_side_mask, &ctx->n_side_elems);
  if (rc) return rc;
  for (int i = 0; i < 3; ++i) ctx->side_g[i] = i < n_g ? g[i] : 0.;
  return FEASM_OK;
}
