// orc_system.cpp — DOF numbering, Dirichlet constraints, sparsity pattern, element matrices and the serial
// scatter-add assembly loop of the CPU oracle.  TEST INFRASTRUCTURE (see fe_oracle.h).  Follows
//   src/base/dof_map.C:1417-1535 (distribute_local_dofs_var_major), :999-1015 (per-rank ranges),
//   src/base/dof_map.C:2201-2314 + include/base/dof_map.h:2618-2772 + include/base/dof_object.h:1028-1056 (dof_indices),
//   src/base/dof_map_constraints.C:600-740 (Lagrange Dirichlet = nodal interpolation), :2439-2614, :3543-3680,
//   src/base/sparsity_pattern.C:96-363, src/numerics/laspack_matrix.C:259-275/:626-640/:713-732,
//   src/numerics/numeric_vector.C:690-697,
//   examples/introduction/introduction_ex4/introduction_ex4.C:407-558,
//   examples/systems_of_equations/systems_of_equations_ex6/systems_of_equations_ex6.C:136-264.
#include <omp.h>

#include <algorithm>
#include <chrono>
#include <cstring>

#include "orc_internal.h"

namespace orc {

struct System {
  const Mesh* mesh = nullptr;
  int n_vars = 1, nd = 0;
  int64_t n_dofs = 0;
  std::vector<uint32_t> node_base;  // n_nodes
  std::vector<uint32_t> elem_dofs;  // n_elem*nd, variable-major then element-local node order
  std::vector<uint32_t> first_dof;  // n_parts+1
  std::vector<uint8_t> is_con;      // n_dofs
  std::vector<double> con_val;      // n_dofs
  int64_t n_con = 0;
  std::vector<uint64_t> row_ptr;
  std::vector<uint32_t> col_idx;
  // side loads (systems_of_equations_ex6.C:240-257): bit s of side_mask[e] = side s of element e carries the
  // traction side_g (one component per variable)
  std::vector<uint8_t> side_mask;
  double side_g[3] = {0., 0., 0.};
  // penalty sides (introduction_ex3.C:386-455): bit s of pen_mask[e] = side s of element e carries the penalty term
  std::vector<uint8_t> pen_mask;
  double penalty = 0., pen_value = 0.;
  int pen_kind = 0;  // 0: value = pen_value, 1: value = exact_solution at the side point
};

static void distribute_dofs(System& s) {
  const Mesh& m = *s.mesh;
  const uint32_t INVALID = 0xffffffffu;
  s.node_base.assign(m.n_nodes, INVALID);
  s.first_dof.assign(m.n_parts + 1, 0);
  uint32_t next_free = 0;
  // One variable group holding n_vars same-type variables; rank p numbers, walking its local elements in id
  // order, every node it owns at first touch with n_vars consecutive ids (dof_map.C:1445-1507).
  for (int p = 0; p < m.n_parts; ++p) {
    s.first_dof[p] = next_free;
    for (int64_t e = 0; e < m.n_elem; ++e) {
      if (m.elem_pid[e] != p) continue;
      for (int a = 0; a < m.npe; ++a) {
        const uint32_t node = m.conn[e * m.npe + a];
        if (m.node_pid[node] == p && s.node_base[node] == INVALID) {
          s.node_base[node] = next_free;
          next_free += (uint32_t)s.n_vars;
        }
      }
    }
  }
  s.first_dof[m.n_parts] = next_free;
  s.n_dofs = next_free;
  s.nd = m.npe * s.n_vars;
  s.elem_dofs.resize((size_t)m.n_elem * s.nd);
  // DofMap::dof_indices: for each variable, for each element node: base + var (dof_map.C:2280-2308)
  for (int64_t e = 0; e < m.n_elem; ++e)
    for (int v = 0; v < s.n_vars; ++v)
      for (int a = 0; a < m.npe; ++a) s.elem_dofs[e * s.nd + v * m.npe + a] = s.node_base[m.conn[e * m.npe + a]] + v;
  s.is_con.assign(s.n_dofs, 0);
  s.con_val.assign(s.n_dofs, 0.);
}

static int64_t build_sparsity(System& s) {
  const Mesh& m = *s.mesh;
  // node -> elements
  std::vector<uint64_t> nptr(m.n_nodes + 1, 0);
  for (uint32_t c : m.conn) nptr[c + 1]++;
  for (int64_t i = 0; i < m.n_nodes; ++i) nptr[i + 1] += nptr[i];
  std::vector<uint32_t> nelem(nptr[m.n_nodes]);
  {
    std::vector<uint64_t> fill(nptr.begin(), nptr.end() - 1);
    for (int64_t e = 0; e < m.n_elem; ++e)
      for (int a = 0; a < m.npe; ++a) nelem[fill[m.conn[e * m.npe + a]]++] = (uint32_t)e;
  }
  // Row of dof (node, v) = sorted unique union of the dof sets of the elements touching the node
  // (SparsityPattern::Build::operator(), sparsity_pattern.C:279-363; all variables couple).
  std::vector<uint32_t> dof_node(s.n_dofs);
  for (int64_t i = 0; i < m.n_nodes; ++i)
    for (int v = 0; v < s.n_vars; ++v) dof_node[s.node_base[i] + v] = (uint32_t)i;
  s.row_ptr.assign(s.n_dofs + 1, 0);
  std::vector<uint32_t> counts(s.n_dofs);
  for (int pass = 0; pass < 2; ++pass) {
#pragma omp parallel
    {
      std::vector<uint32_t> tmp;
#pragma omp for schedule(dynamic, 1024)
      for (int64_t r = 0; r < s.n_dofs; ++r) {
        const uint32_t node = dof_node[r];
        tmp.clear();
        for (uint64_t k = nptr[node]; k < nptr[node + 1]; ++k) {
          const uint32_t* d = &s.elem_dofs[(size_t)nelem[k] * s.nd];
          tmp.insert(tmp.end(), d, d + s.nd);
        }
        std::sort(tmp.begin(), tmp.end());
        tmp.erase(std::unique(tmp.begin(), tmp.end()), tmp.end());
        if (pass == 0)
          counts[r] = (uint32_t)tmp.size();
        else
          std::memcpy(&s.col_idx[s.row_ptr[r]], tmp.data(), tmp.size() * sizeof(uint32_t));
      }
    }
    if (pass == 0) {
      for (int64_t r = 0; r < s.n_dofs; ++r) s.row_ptr[r + 1] = s.row_ptr[r] + counts[r];
      s.col_idx.resize(s.row_ptr[s.n_dofs]);
    }
  }
  return (int64_t)s.row_ptr[s.n_dofs];
}

// ---- element matrices ------------------------------------------------------------------------------
struct ElemWork {
  FEData fe;
  SideData sd;
  std::vector<double> Ke, Fe, KH, X;
  std::vector<uint32_t> dofs;
};

struct Problem {
  int problem, forcing, qorder;
  const double* f_tab;
  QRule q;
  ShapeTable tab;
};

static bool setup_problem(const System& s, int problem, int qorder, int forcing, const double* f_tab, Problem& P) {
  const Mesh& m = *s.mesh;
  if (qorder == 0) {
    // ex4 hard-codes FIFTH (introduction_ex4.C:352); ex6 uses default_quadrature_order() = 2*order+1
    if (problem == ORC_ELASTICITY_EX6) qorder = (m.type == ORC_HEX8) ? 3 : 5;
    else qorder = 5;
  }
  if ((problem == ORC_ELASTICITY_EX6) != (s.n_vars == 3) && problem == ORC_ELASTICITY_EX6) { set_error("elasticity needs 3 variables"); return false; }
  if (problem != ORC_ELASTICITY_EX6 && s.n_vars != 1) { set_error("Poisson needs 1 variable"); return false; }
  P.problem = problem;
  P.forcing = forcing;
  P.qorder = qorder;
  P.f_tab = f_tab;
  if (!gauss_rule(m.type, qorder, P.q)) return false;
  shape_table(m.type, P.q.nq, P.q.pts.data(), P.tab);
  return true;
}

// elasticity_tensor, systems_of_equations_ex6.C:136-151
static inline double kd(int i, int j) { return i == j ? 1. : 0.; }
static inline double elasticity_tensor(int i, int j, int k, int l) {
  const double poisson_ratio = 0.3;
  const double young_modulus = 1.;
  const double lambda_1 = (young_modulus * poisson_ratio) / ((1. + poisson_ratio) * (1. - 2. * poisson_ratio));
  const double lambda_2 = young_modulus / (2. * (1. + poisson_ratio));
  return lambda_1 * kd(i, j) * kd(k, l) + lambda_2 * (kd(i, k) * kd(j, l) + kd(i, l) * kd(j, k));
}

// Computes unconstrained Ke, Fe of element e.  Returns 0 or -1 (degenerate).  t[0..2] accumulate phase seconds.
static int element_matrix(const System& s, const Problem& P, int64_t e, ElemWork& w, double* t) {
  using clk = std::chrono::steady_clock;
  const Mesh& m = *s.mesh;
  const int n = m.npe, nq = P.q.nq, nd = s.nd;
  auto t0 = clk::now();
  w.dofs.assign(&s.elem_dofs[e * nd], &s.elem_dofs[e * nd] + nd);
  w.X.resize(3 * n);
  for (int a = 0; a < n; ++a)
    for (int d = 0; d < 3; ++d) w.X[3 * a + d] = m.xyz[3 * (size_t)m.conn[e * n + a] + d];
  if (fe_reinit(m.type, w.X.data(), P.tab, P.q.w.data(), w.fe) < 0) return -1;
  w.Ke.assign((size_t)nd * nd, 0.);
  w.Fe.assign(nd, 0.);
  auto t1 = clk::now();
  const double* JxW = w.fe.JxW.data();
  const double* dphi = w.fe.dphi.data();
  const double* phi = P.tab.phi.data();
  if (P.problem != ORC_ELASTICITY_EX6) {
    // introduction_ex4.C:464-467
    for (int qp = 0; qp < nq; qp++)
      for (int i = 0; i != nd; i++) {
        const double* gi = &dphi[((size_t)i * nq + qp) * 3];
        for (int j = 0; j != nd; j++) {
          const double* gj = &dphi[((size_t)j * nq + qp) * 3];
          w.Ke[(size_t)i * nd + j] += JxW[qp] * (gi[0] * gj[0] + gi[1] * gj[1] + gi[2] * gj[2]);
        }
      }
    auto t2 = clk::now();
    if (P.problem == ORC_POISSON_EX4) {
      for (int qp = 0; qp < nq; qp++) {  // introduction_ex4.C:480-537
        double fxy;
        if (P.forcing == ORC_FORCING_TABULATED)
          fxy = P.f_tab[e * nq + qp];
        else
          fxy = ex4_forcing(w.fe.xyz[3 * qp], w.fe.xyz[3 * qp + 1], w.fe.xyz[3 * qp + 2]);
        for (int i = 0; i != nd; i++) w.Fe[i] += JxW[qp] * fxy * phi[i * nq + qp];
      }
    }
    auto t3 = clk::now();
    if (t) {
      t[0] += std::chrono::duration<double>(t1 - t0).count();
      t[1] += std::chrono::duration<double>(t2 - t1).count();
      t[2] += std::chrono::duration<double>(t3 - t2).count();
    }
  } else {
    // systems_of_equations_ex6.C:220-237: Ke_var[i][k](dof_i,dof_j) += JxW*C_ijkl*dphi[dof_j](l)*dphi[dof_i](j)
    static double C[3][3][3][3];
    static bool init = false;
    if (!init) {
#pragma omp critical
      {
        for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) for (int k = 0; k < 3; k++) for (int l = 0; l < 3; l++)
          C[i][j][k][l] = elasticity_tensor(i, j, k, l);
        init = true;
      }
    }
    for (int qp = 0; qp < nq; qp++) {
      for (int di = 0; di < n; di++)
        for (int dj = 0; dj < n; dj++) {
          const double* gi = &dphi[((size_t)di * nq + qp) * 3];
          const double* gj = &dphi[((size_t)dj * nq + qp) * 3];
          for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++)
              for (int k = 0; k < 3; k++)
                for (int l = 0; l < 3; l++) {
                  const double c = C[i][j][k][l];
                  if (c == 0.) continue;  // the reference adds an exact zero here
                  w.Ke[(size_t)(i * n + di) * nd + (k * n + dj)] += JxW[qp] * c * gj[l] * gi[j];
                }
        }
      const double f_vec[3] = {0., 0., -1.};
      for (int di = 0; di < n; di++)
        for (int i = 0; i < 3; i++) w.Fe[i * n + di] += JxW[qp] * (f_vec[i] * phi[di * nq + qp]);
    }
    auto t2 = clk::now();
    if (t) {
      t[0] += std::chrono::duration<double>(t1 - t0).count();
      t[1] += std::chrono::duration<double>(t2 - t1).count();
    }
  }
  // assemble \int_\Gamma g_i v_i ds (systems_of_equations_ex6.C:240-257): fe_face->reinit(elem, side) with
  // QGauss(dim-1, same order), then Fe_var[i](dof_i) += JxW_face[qp] * (g_vec(i) * phi_face[dof_i][qp])
  if (!s.side_mask.empty() && s.side_mask[e]) {
    for (int side = 0; side < 6; ++side) {
      if (!(s.side_mask[e] >> side & 1)) continue;
      if (side_reinit(m.type, w.X.data(), side, P.qorder, w.sd) < 0) return -1;
      const int nqf = w.sd.nqf;
      for (int qp = 0; qp < nqf; qp++)
        for (int di = 0; di < n; di++)
          for (int i = 0; i < s.n_vars; i++) w.Fe[i * n + di] += w.sd.JxW[qp] * (s.side_g[i] * w.sd.phi[di * nqf + qp]);
    }
  }
  // Dirichlet data by the penalty method, introduction_ex3.C:386-455 (there in 2D; the same loop on the sides of a 3D
  // element): for every side without a neighbour
  //   Ke(i,j) += JxW_face[qp]*penalty*phi_face[i][qp]*phi_face[j][qp];   Fe(i) += JxW_face[qp]*penalty*value*phi_face[i][qp];
  // with several variables the term is applied to every variable's diagonal block (this restatement's extension).
  if (!s.pen_mask.empty() && s.pen_mask[e]) {
    for (int side = 0; side < 6; ++side) {
      if (!(s.pen_mask[e] >> side & 1)) continue;
      if (side_reinit(m.type, w.X.data(), side, P.qorder, w.sd) < 0) return -1;
      const int nqf = w.sd.nqf;
      for (int qp = 0; qp < nqf; qp++) {
        const double value = s.pen_kind == 1 ? exact_solution(w.sd.xyz[3 * qp], w.sd.xyz[3 * qp + 1], w.sd.xyz[3 * qp + 2]) : s.pen_value;
        for (int v = 0; v < s.n_vars; v++) {
          for (int i = 0; i < n; i++)
            for (int j = 0; j < n; j++)
              w.Ke[(size_t)(v * n + i) * nd + (v * n + j)] += w.sd.JxW[qp] * s.penalty * w.sd.phi[i * nqf + qp] * w.sd.phi[j * nqf + qp];
          for (int i = 0; i < n; i++) w.Fe[v * n + i] += w.sd.JxW[qp] * s.penalty * value * w.sd.phi[i * nqf + qp];
        }
      }
    }
  }
  return 0;
}

// Dirichlet-only restatement of (heterogeneously_)constrain_element_matrix_and_vector
// (dof_map_constraints.C:2439-2614): C = diag(0 on constrained, 1 else), H = g on constrained.
static void constrain_element(const System& s, int problem, ElemWork& w) {
  const int nd = s.nd;
  bool any = false;
  for (int i = 0; i < nd; ++i) any |= (s.is_con[w.dofs[i]] != 0);
  if (!any) return;  // build_constraint_matrix returns an empty C (:3575)
  const bool hetero = (problem != ORC_ELASTICITY_EX6);
  if (hetero) {
    // F <- C^T (F - K H)
    w.KH.assign(nd, 0.);
    for (int i = 0; i < nd; ++i) {
      double acc = 0.;
      for (int j = 0; j < nd; ++j)
        if (s.is_con[w.dofs[j]]) acc += w.Ke[(size_t)i * nd + j] * s.con_val[w.dofs[j]];
      w.KH[i] = acc;
    }
    for (int i = 0; i < nd; ++i) w.Fe[i] = s.is_con[w.dofs[i]] ? 0. : (w.Fe[i] - w.KH[i]);
  } else {
    for (int i = 0; i < nd; ++i)
      if (s.is_con[w.dofs[i]]) w.Fe[i] = 0.;  // C^T F (:2508-2512)
  }
  // K <- C^T K C, then constrained rows: zero, diag = 1, rhs = g (hetero) (:2570-2612)
  for (int i = 0; i < nd; ++i)
    for (int j = 0; j < nd; ++j)
      if (s.is_con[w.dofs[i]] || s.is_con[w.dofs[j]]) w.Ke[(size_t)i * nd + j] = 0.;
  for (int i = 0; i < nd; ++i)
    if (s.is_con[w.dofs[i]]) {
      w.Ke[(size_t)i * nd + i] = 1.;
      if (hetero) w.Fe[i] = s.con_val[w.dofs[i]];
    }
}

static int assemble(const System& s, int problem, int qorder, int forcing, const double* f_tab, int n_threads,
                    int64_t e0, int64_t e1, bool zero, double* values, double* rhs, double* phases) {
  using clk = std::chrono::steady_clock;
  Problem P;
  if (!setup_problem(s, problem, qorder, forcing, f_tab, P)) return -1;
  if (s.row_ptr.empty()) { set_error("assemble: sparsity not built"); return -2; }
  const int nd = s.nd;
  if (zero) {
    std::memset(values, 0, sizeof(double) * s.row_ptr[s.n_dofs]);
    std::memset(rhs, 0, sizeof(double) * s.n_dofs);
  }
  int err = 0;
  if (n_threads <= 1) {
    ElemWork w;
    double t[4] = {0, 0, 0, 0};
    for (int64_t e = e0; e < e1; ++e) {
      if (element_matrix(s, P, e, w, t) < 0) return -3;
      auto t0 = clk::now();
      constrain_element(s, problem, w);
      // SparseMatrix::add_matrix: row-major walk, binary search of the column in the sorted row
      // (laspack_matrix.C:713-732); NumericVector::add_vector (numeric_vector.C:690-697)
      for (int i = 0; i < nd; ++i) {
        const uint32_t r = w.dofs[i];
        const uint32_t* cb = &s.col_idx[s.row_ptr[r]];
        const uint32_t* ce = &s.col_idx[s.row_ptr[r + 1]];
        for (int j = 0; j < nd; ++j) {
          const uint32_t* pos = std::lower_bound(cb, ce, w.dofs[j]);
          if (pos == ce || *pos != w.dofs[j]) { set_error("add_matrix: entry not in sparsity pattern"); return -4; }
          values[s.row_ptr[r] + (pos - cb)] += w.Ke[(size_t)i * nd + j];
        }
      }
      for (int i = 0; i < nd; ++i) rhs[w.dofs[i]] += w.Fe[i];
      t[3] += std::chrono::duration<double>(clk::now() - t0).count();
    }
    if (phases) std::memcpy(phases, t, sizeof(t));
  } else {
#pragma omp parallel num_threads(n_threads)
    {
      ElemWork w;
#pragma omp for schedule(static)
      for (int64_t e = e0; e < e1; ++e) {
        if (err) continue;
        if (element_matrix(s, P, e, w, nullptr) < 0) {
#pragma omp atomic write
          err = -3;
          continue;
        }
        constrain_element(s, problem, w);
        for (int i = 0; i < nd; ++i) {
          const uint32_t r = w.dofs[i];
          const uint32_t* cb = &s.col_idx[s.row_ptr[r]];
          const uint32_t* ce = &s.col_idx[s.row_ptr[r + 1]];
          for (int j = 0; j < nd; ++j) {
            const uint32_t* pos = std::lower_bound(cb, ce, w.dofs[j]);
            double& dst = values[s.row_ptr[r] + (pos - cb)];
            const double v = w.Ke[(size_t)i * nd + j];
#pragma omp atomic
            dst += v;
          }
        }
        for (int i = 0; i < nd; ++i) {
          double& dst = rhs[w.dofs[i]];
          const double v = w.Fe[i];
#pragma omp atomic
          dst += v;
        }
      }
    }
  }
  return err;
}

}  // namespace orc

struct orc_system {
  orc::System s;
};

extern "C" {

orc_system* orc_system_create(const orc_mesh* mesh, int n_vars) {
  orc_system* h = new orc_system;
  h->s.mesh = &mesh->m;
  h->s.n_vars = n_vars;
  orc::distribute_dofs(h->s);
  return h;
}
void orc_system_free(orc_system* h) { delete h; }
int64_t orc_system_n_dofs(const orc_system* h) { return h->s.n_dofs; }
int orc_system_dofs_per_elem(const orc_system* h) { return h->s.nd; }
void orc_system_get_dofs(const orc_system* h, uint32_t* elem_dofs, uint32_t* node_dof_base, uint32_t* first_dof) {
  const orc::System& s = h->s;
  if (elem_dofs) std::memcpy(elem_dofs, s.elem_dofs.data(), sizeof(uint32_t) * s.elem_dofs.size());
  if (node_dof_base) std::memcpy(node_dof_base, s.node_base.data(), sizeof(uint32_t) * s.node_base.size());
  if (first_dof) std::memcpy(first_dof, s.first_dof.data(), sizeof(uint32_t) * s.first_dof.size());
}

int64_t orc_system_add_dirichlet(orc_system* h, uint32_t id_mask, int kind) {
  orc::System& s = h->s;
  const orc::Mesh& m = *s.mesh;
  // apply_lagrange_dirichlet_impl (dof_map_constraints.C:600-740): every node on a side carrying one of the
  // boundary ids gets an empty constraint row with value f(node)
  for (int64_t i = 0; i < m.n_nodes; ++i) {
    if (!(m.bmask[i] & id_mask)) continue;
    for (int v = 0; v < s.n_vars; ++v) {
      const uint32_t d = s.node_base[i] + v;
      if (!s.is_con[d]) s.n_con++;
      s.is_con[d] = 1;
      s.con_val[d] = (kind == 0) ? orc::exact_solution(m.xyz[3 * i], m.xyz[3 * i + 1], m.xyz[3 * i + 2]) : 0.;
    }
  }
  return s.n_con;
}

void orc_system_get_constraints(const orc_system* h, uint32_t* dofs, double* values) {
  const orc::System& s = h->s;
  int64_t k = 0;
  for (int64_t d = 0; d < s.n_dofs; ++d)
    if (s.is_con[d]) {
      if (dofs) dofs[k] = (uint32_t)d;
      if (values) values[k] = s.con_val[d];
      k++;
    }
}

void orc_system_set_side_loads(orc_system* h, int64_t n, const uint32_t* elems, const uint8_t* sides, const double* g) {
  orc::System& s = h->s;
  s.side_mask.clear();
  if (n <= 0) return;
  s.side_mask.assign(s.mesh->n_elem, 0);
  for (int64_t k = 0; k < n; ++k) s.side_mask[elems[k]] |= (uint8_t)(1u << sides[k]);
  for (int i = 0; i < 3; ++i) s.side_g[i] = i < s.n_vars ? g[i] : 0.;
}

void orc_system_set_side_penalty(orc_system* h, int64_t n, const uint32_t* elems, const uint8_t* sides, double penalty, int kind, double value) {
  orc::System& s = h->s;
  s.pen_mask.clear();
  if (n <= 0) return;
  s.pen_mask.assign(s.mesh->n_elem, 0);
  for (int64_t k = 0; k < n; ++k) s.pen_mask[elems[k]] |= (uint8_t)(1u << sides[k]);
  s.penalty = penalty; s.pen_kind = kind; s.pen_value = value;
}

int64_t orc_system_build_sparsity(orc_system* h) { return orc::build_sparsity(h->s); }
void orc_system_get_sparsity(const orc_system* h, uint64_t* row_ptr, uint32_t* col_idx) {
  const orc::System& s = h->s;
  if (row_ptr) std::memcpy(row_ptr, s.row_ptr.data(), sizeof(uint64_t) * s.row_ptr.size());
  if (col_idx) std::memcpy(col_idx, s.col_idx.data(), sizeof(uint32_t) * s.col_idx.size());
}

int orc_element_matrices(const orc_system* h, int problem, int qorder, int forcing, const double* f_tab, int64_t e0,
                         int64_t e1, int constrained, double* Ke, double* Fe) {
  const orc::System& s = h->s;
  orc::Problem P;
  if (!orc::setup_problem(s, problem, qorder, forcing, f_tab, P)) return -1;
  orc::ElemWork w;
  const size_t nd = s.nd;
  for (int64_t e = e0; e < e1; ++e) {
    if (orc::element_matrix(s, P, e, w, nullptr) < 0) return -3;
    if (constrained) orc::constrain_element(s, problem, w);
    if (Ke) std::memcpy(Ke + (size_t)(e - e0) * nd * nd, w.Ke.data(), sizeof(double) * nd * nd);
    if (Fe) std::memcpy(Fe + (size_t)(e - e0) * nd, w.Fe.data(), sizeof(double) * nd);
  }
  return 0;
}

int orc_forcing_at_qps(const orc_system* h, int qorder, double* f, double* xyz_q) {
  const orc::System& s = h->s;
  const orc::Mesh& m = *s.mesh;
  orc::Problem P;
  if (!orc::setup_problem(s, ORC_POISSON_EX4, qorder, 0, nullptr, P)) return -1;
  const int nq = P.q.nq, n = m.npe;
  int err = 0;
#pragma omp parallel
  {
    orc::FEData fe;
    std::vector<double> X(3 * n);
#pragma omp for schedule(static)
    for (int64_t e = 0; e < m.n_elem; ++e) {
      for (int a = 0; a < n; ++a)
        for (int d = 0; d < 3; ++d) X[3 * a + d] = m.xyz[3 * (size_t)m.conn[e * n + a] + d];
      if (orc::fe_reinit(m.type, X.data(), P.tab, P.q.w.data(), fe) < 0) { err = -3; continue; }
      for (int q = 0; q < nq; ++q) {
        if (f) f[e * nq + q] = orc::ex4_forcing(fe.xyz[3 * q], fe.xyz[3 * q + 1], fe.xyz[3 * q + 2]);
        if (xyz_q) std::memcpy(&xyz_q[(e * nq + q) * 3], &fe.xyz[3 * q], 3 * sizeof(double));
      }
    }
  }
  return err;
}

int orc_assemble(const orc_system* h, int problem, int qorder, int forcing, const double* f_tab, int n_threads,
                 double* values, double* rhs, double* phase_seconds) {
  return orc::assemble(h->s, problem, qorder, forcing, f_tab, n_threads, 0, h->s.mesh->n_elem, true, values, rhs,
                       phase_seconds);
}

int orc_assemble_range(const orc_system* h, int problem, int qorder, int forcing, const double* f_tab, int n_threads,
                       int64_t e0, int64_t e1, double* values, double* rhs) {
  return orc::assemble(h->s, problem, qorder, forcing, f_tab, n_threads, e0, e1, false, values, rhs, nullptr);
}

}  // extern "C"
