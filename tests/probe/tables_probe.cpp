// tests/probe/tables_probe.cpp — test-only: the PRODUCT's host-side table generator (libmesh_b200/csrc/feasm_tables.cpp: the
// quadrature rule and the shape tables every element kernel reads, and the side tables of feasm_side.cu) compiled for the host
// alone, so that tests/test_product_tables.py can hold it to the reference's own tables without a GPU.
#include <cstring>

#include "feasm_internal.h"

extern "C" {

// npe, nq of (elem_type, qorder); -1 with the message when the product rejects the pair
int probe_table_sizes(int elem_type, int qorder, int* npe, int* nq, char* err, int err_len) {
  FeasmTables t;
  std::string e;
  if (!feasm_make_tables(elem_type, qorder, t, e)) {
    std::strncpy(err, e.c_str(), err_len - 1);
    err[err_len - 1] = 0;
    return -1;
  }
  *npe = t.npe;
  *nq = t.nq;
  return 0;
}
int probe_tables(int elem_type, int qorder, double* qp, double* qw, double* phi, double* dxi, double* deta, double* dzeta) {
  FeasmTables t;
  std::string e;
  if (!feasm_make_tables(elem_type, qorder, t, e)) return -1;
  std::memcpy(qp, t.qp.data(), sizeof(double) * t.qp.size());
  std::memcpy(qw, t.qw.data(), sizeof(double) * t.qw.size());
  const size_t b = sizeof(double) * (size_t)t.npe * t.nq;
  std::memcpy(phi, t.phi.data(), b);
  std::memcpy(dxi, t.dxi.data(), b);
  std::memcpy(deta, t.deta.data(), b);
  std::memcpy(dzeta, t.dzeta.data(), b);
  return 0;
}
// side tables: sizes (npe, ns, nqf), then side_nodes [n_sides][ns], qw [nqf], psi / dxi / deta [ns][nqf], phi [n_sides][npe][nqf]
int probe_side_sizes(int elem_type, int qorder, int* npe, int* ns, int* nqf, int* n_sides) {
  FeasmSideTables t;
  std::string e;
  if (!feasm_make_side_tables(elem_type, qorder, t, e)) return -1;
  *npe = t.npe; *ns = t.ns; *nqf = t.nqf;
  *n_sides = t.ns ? (int)(t.side_nodes.size() / t.ns) : 0;
  return 0;
}
int probe_side_tables(int elem_type, int qorder, int* side_nodes, double* qw, double* psi, double* dxi, double* deta, double* phi) {
  FeasmSideTables t;
  std::string e;
  if (!feasm_make_side_tables(elem_type, qorder, t, e)) return -1;
  std::memcpy(side_nodes, t.side_nodes.data(), sizeof(int) * t.side_nodes.size());
  std::memcpy(qw, t.qw.data(), sizeof(double) * t.qw.size());
  std::memcpy(psi, t.psi.data(), sizeof(double) * t.psi.size());
  std::memcpy(dxi, t.dxi.data(), sizeof(double) * t.dxi.size());
  std::memcpy(deta, t.deta.data(), sizeof(double) * t.deta.size());
  std::memcpy(phi, t.phi.data(), sizeof(double) * t.phi.size());
  return 0;
}

}  // extern "C"
