// orc_internal.h — shared declarations of the CPU oracle (TEST INFRASTRUCTURE, see fe_oracle.h).
#ifndef ORC_INTERNAL_H
#define ORC_INTERNAL_H

#include <cstdint>
#include <string>
#include <vector>

#include "fe_oracle.h"

namespace orc {

void set_error(const std::string& msg);

// ---- quadrature + reference shape tables ---------------------------------------------------
struct QRule {
  int nq = 0;
  std::vector<double> pts;  // nq*3
  std::vector<double> w;    // nq
};
// src/quadrature/quadrature_gauss_1D.C:28-70 + quadrature.C:283-305 (hex), quadrature_gauss_3D.C:233-290 (tet)
bool gauss_rule(int elem_type, int qorder, QRule& q);

// Reference shapes, layout [i][p] like libMesh's std::vector<std::vector<Real>>
struct ShapeTable {
  int n = 0, np = 0;
  std::vector<double> phi, dxi, deta, dzeta;  // n*np each
};
void shape_table(int elem_type, int np, const double* ref_pts, ShapeTable& t);

// What FE::reinit leaves behind for one element
struct FEData {
  int n = 0, np = 0;
  std::vector<double> xyz;   // np*3
  std::vector<double> JxW;   // np
  std::vector<double> jinv;  // np*9: dxidx dxidy dxidz detadx detady detadz dzetadx dzetady dzetadz
  std::vector<double> dphi;  // n*np*3   dphi[i][p](d)
  const ShapeTable* tab = nullptr;
  void resize(int n_, int np_);
};
bool has_affine_map(int elem_type, const double* X /*n*3*/);
// returns 1 affine, 0 general, -1 degenerate
int fe_reinit(int elem_type, const double* X, const ShapeTable& tab, const double* qw, FEData& fe);

// What FE::reinit(elem, side) leaves behind (fe_boundary.C:109-240)
struct SideData {
  int nqf = 0;
  std::vector<double> ref;     // nqf*3: the side quadrature points in the parent's reference space (side_map)
  std::vector<double> xyz;     // nqf*3
  std::vector<double> JxW;     // nqf
  std::vector<double> normal;  // nqf*3
  std::vector<double> phi;     // n*nqf: parent shapes at ref
};
int side_nodes(int elem_type, int side, int* nodes);  // Elem::local_side_node; returns nodes per side (0: unsupported)
int side_reinit(int elem_type, const double* X, int side, int qorder, SideData& sd);  // returns nqf or -1

double exact_solution(double x, double y, double z);
double ex4_forcing(double x, double y, double z);

// ---- mesh ---------------------------------------------------------------------------------
struct Mesh {
  int type = 0, npe = 0;
  int64_t n_elem = 0, n_nodes = 0;
  std::vector<double> xyz;
  std::vector<uint32_t> conn;
  std::vector<uint8_t> bmask;
  int n_parts = 1;
  std::vector<int32_t> elem_pid, node_pid;
  void compute_node_pids();
};

}  // namespace orc

struct orc_mesh {
  orc::Mesh m;
};

#endif
