// orc_mesh.cpp — MeshTools::Generation::build_cube restated for HEX8 / HEX27 / TET10, plus the
// LinearPartitioner and libMesh's node-ownership rule.  TEST INFRASTRUCTURE (see fe_oracle.h).
// Follows src/mesh/mesh_generation.C:117-150 (idx), :1285-1340 (nodes), :1399-1432 (HEX8),
// :1575-1626 (HEX27), :1779-1785 (rescale), :1795-1907 (24-tet split + all_second_order),
// src/mesh/unstructured_mesh.C:60-210 (mid-edge node creation order),
// src/mesh/replicated_mesh.C:700-850 (renumber_nodes_and_elements),
// src/partitioning/linear_partitioner.C:36-80, src/geom/node.C:78-103.
#include <algorithm>
#include <cstring>
#include <map>
#include <random>

#include "orc_internal.h"

namespace orc {

// src/geom/cell_hex27.C:37-45
static const int HEX27_SIDE_NODES[6][9] = {{0, 3, 2, 1, 11, 10, 9, 8, 20},   {0, 1, 5, 4, 8, 13, 16, 12, 21},
                                           {1, 2, 6, 5, 9, 14, 17, 13, 22},  {2, 3, 7, 6, 10, 15, 18, 14, 23},
                                           {3, 0, 4, 7, 11, 12, 19, 15, 24}, {4, 5, 6, 7, 16, 17, 18, 19, 25}};
// TET10 mid-edge node e (local 4+e) sits between these vertices (src/geom/cell_tet10.C edge_nodes_map)
static const int TET_EDGE[6][2] = {{0, 1}, {1, 2}, {0, 2}, {0, 3}, {1, 3}, {2, 3}};

void Mesh::compute_node_pids() {
  // Node owner = min processor id over touching elements (src/geom/node.C:99-102)
  node_pid.assign(n_nodes, INT32_MAX);
  for (int64_t e = 0; e < n_elem; ++e)
    for (int a = 0; a < npe; ++a) {
      int32_t& p = node_pid[conn[e * npe + a]];
      p = std::min(p, elem_pid[e]);
    }
}

// ReplicatedMesh::renumber_nodes_and_elements (replicated_mesh.C:700-850), node part: ids are handed out in
// element-walk first-touch order by swapping slots; nodes left beyond next_free_node are orphans and deleted.
static void renumber_nodes(Mesh& m, bool libmesh_order) {
  const int64_t nn = m.n_nodes;
  std::vector<int64_t> new_id(nn, -1);
  int64_t next_free = 0;
  if (libmesh_order) {
    std::vector<int64_t> slot_of(nn), node_at(nn);  // node -> slot (= current id), slot -> node
    for (int64_t i = 0; i < nn; ++i) slot_of[i] = node_at[i] = i;
    for (int64_t e = 0; e < m.n_elem; ++e)
      for (int a = 0; a < m.npe; ++a) {
        const int64_t node = m.conn[e * m.npe + a];
        const int64_t id = slot_of[node];
        if (id == next_free)
          next_free++;
        else if (id > next_free) {
          const int64_t dst = next_free++;
          const int64_t other = node_at[dst];
          node_at[dst] = node; slot_of[node] = dst;
          node_at[id] = other; slot_of[other] = id;
        }
      }
    for (int64_t i = 0; i < nn; ++i) new_id[i] = slot_of[i] < next_free ? slot_of[i] : -1;
  } else {
    std::vector<uint8_t> used(nn, 0);
    for (uint32_t c : m.conn) used[c] = 1;
    for (int64_t i = 0; i < nn; ++i)
      if (used[i]) new_id[i] = next_free++;
  }
  std::vector<double> xyz((size_t)next_free * 3);
  std::vector<uint8_t> bm(next_free);
  for (int64_t i = 0; i < nn; ++i)
    if (new_id[i] >= 0) {
      for (int d = 0; d < 3; ++d) xyz[3 * new_id[i] + d] = m.xyz[3 * i + d];
      bm[new_id[i]] = m.bmask[i];
    }
  for (auto& c : m.conn) c = (uint32_t)new_id[c];
  m.xyz.swap(xyz);
  m.bmask.swap(bm);
  m.n_nodes = next_free;
}

static bool build_cube(Mesh& m, int type, int nx, int ny, int nz, double xmin, double xmax, double ymin, double ymax,
                       double zmin, double zmax, bool libmesh_renumber) {
  if (type != ORC_HEX8 && type != ORC_HEX27 && type != ORC_TET10 && type != ORC_HEX20 && type != ORC_PRISM18) { set_error("build_cube: bad element type"); return false; }
  if (nx < 1 || ny < 1 || nz < 1) { set_error("build_cube: n < 1"); return false; }
  m.type = type;
  m.npe = type;
  const bool lin = (type == ORC_HEX8);
  const int gx = lin ? nx : 2 * nx, gy = lin ? ny : 2 * ny, gz = lin ? nz : 2 * nz;  // grid intervals
  m.n_nodes = (int64_t)(gx + 1) * (gy + 1) * (gz + 1);
  m.xyz.resize((size_t)m.n_nodes * 3);
  m.bmask.assign(m.n_nodes, 0);
  int64_t id = 0;
  for (int k = 0; k <= gz; k++)
    for (int j = 0; j <= gy; j++)
      for (int i = 0; i <= gx; i++) {
        // mesh_generation.C:1289-1296 / :1329-1336 — unit-cube coordinates first ...
        double x = static_cast<double>(i) / static_cast<double>(gx);
        double y = static_cast<double>(j) / static_cast<double>(gy);
        double z = static_cast<double>(k) / static_cast<double>(gz);
        // ... then rescaled in place (:1779-1785)
        m.xyz[3 * id + 0] = x * (xmax - xmin) + xmin;
        m.xyz[3 * id + 1] = y * (ymax - ymin) + ymin;
        m.xyz[3 * id + 2] = z * (zmax - zmin) + zmin;
        uint8_t b = 0;
        if (k == 0) b |= 1u << 0;
        if (k == gz) b |= 1u << 5;
        if (j == 0) b |= 1u << 1;
        if (j == gy) b |= 1u << 3;
        if (i == 0) b |= 1u << 4;
        if (i == gx) b |= 1u << 2;
        m.bmask[id] = b;
        id++;
      }
  auto idx = [&](int i, int j, int k) -> uint32_t { return (uint32_t)(i + (int64_t)(gx + 1) * (j + (int64_t)k * (gy + 1))); };
  const int64_t nhex = (int64_t)nx * ny * nz;
  std::vector<uint32_t> hex;  // HEX8 or HEX27 connectivity
  const int hn = lin ? 8 : 27;
  hex.resize((size_t)nhex * hn);
  int64_t e = 0;
  const int st = lin ? 1 : 2;
  for (int k = 0; k < gz; k += st)
    for (int j = 0; j < gy; j += st)
      for (int i = 0; i < gx; i += st) {
        uint32_t* c = &hex[e * hn];
        const int s = st;
        c[0] = idx(i, j, k);         c[1] = idx(i + s, j, k);
        c[2] = idx(i + s, j + s, k); c[3] = idx(i, j + s, k);
        c[4] = idx(i, j, k + s);     c[5] = idx(i + s, j, k + s);
        c[6] = idx(i + s, j + s, k + s); c[7] = idx(i, j + s, k + s);
        if (!lin) {
          c[8] = idx(i + 1, j, k);       c[9] = idx(i + 2, j + 1, k);
          c[10] = idx(i + 1, j + 2, k);  c[11] = idx(i, j + 1, k);
          c[12] = idx(i, j, k + 1);      c[13] = idx(i + 2, j, k + 1);
          c[14] = idx(i + 2, j + 2, k + 1); c[15] = idx(i, j + 2, k + 1);
          c[16] = idx(i + 1, j, k + 2);  c[17] = idx(i + 2, j + 1, k + 2);
          c[18] = idx(i + 1, j + 2, k + 2); c[19] = idx(i, j + 1, k + 2);
          c[20] = idx(i + 1, j + 1, k);  c[21] = idx(i + 1, j, k + 1);
          c[22] = idx(i + 2, j + 1, k + 1); c[23] = idx(i + 1, j + 2, k + 1);
          c[24] = idx(i, j + 1, k + 1);  c[25] = idx(i + 1, j + 1, k + 2);
          c[26] = idx(i + 1, j + 1, k + 1);
        }
        e++;
      }
  if (type == ORC_HEX20) {
    // mesh_generation.C:1559-1592: a HEX20 takes nodes 0-19 of the HEX27 lattice cell; the face and cell centres stay
    // without an element and are deleted by prepare_for_use (below)
    m.n_elem = nhex;
    m.conn.resize((size_t)nhex * 20);
    for (int64_t h = 0; h < nhex; ++h)
      for (int a = 0; a < 20; ++a) m.conn[h * 20 + a] = hex[h * 27 + a];
  } else if (type == ORC_PRISM18) {
    // mesh_generation.C:1634-1756: every lattice cell gives two prisms split along the diagonal (i+2, j) - (i, j+2); with
    // c = the HEX27 connectivity of the cell, the node picks of the two set_node lists are
    //   first : (i,j) (i+2,j) (i,j+2) | top | mid-edges (i+1,j) (i+1,j+1) (i,j+1) | verticals | top mid-edges | mid-layer
    static const int P1[18] = {0, 1, 3, 4, 5, 7, 8, 20, 11, 12, 13, 15, 16, 25, 19, 21, 26, 24};
    static const int P2[18] = {1, 2, 3, 5, 6, 7, 9, 10, 20, 13, 14, 15, 17, 18, 25, 22, 23, 26};
    m.n_elem = nhex * 2;
    m.conn.resize((size_t)m.n_elem * 18);
    for (int64_t h = 0; h < nhex; ++h)
      for (int a = 0; a < 18; ++a) {
        m.conn[(2 * h) * 18 + a] = hex[h * 27 + P1[a]];
        m.conn[(2 * h + 1) * 18 + a] = hex[h * 27 + P2[a]];
      }
  } else if (type != ORC_TET10) {
    m.n_elem = nhex;
    m.conn.swap(hex);
  } else {
    // 24 sub-tets per hex (mesh_generation.C:1795-1860): side s, sub-tet t -> (side[t], side[8], side[(t+1)%4], hex[26])
    m.n_elem = nhex * 24;
    m.conn.assign((size_t)m.n_elem * 10, 0);
    int64_t t = 0;
    for (int64_t h = 0; h < nhex; ++h) {
      const uint32_t* c = &hex[h * 27];
      for (int s = 0; s < 6; ++s)
        for (int sub = 0; sub < 4; ++sub) {
          uint32_t* v = &m.conn[t * 10];
          v[0] = c[HEX27_SIDE_NODES[s][sub]];
          v[1] = c[HEX27_SIDE_NODES[s][8]];
          v[2] = c[HEX27_SIDE_NODES[s][sub == 3 ? 0 : sub + 1]];
          v[3] = c[26];
          t++;
        }
    }
    // all_second_order (unstructured_mesh.C:60-210): walk tets in id order, local mid-edge nodes 4..9 in order;
    // a node is created at first sight of its (sorted) vertex pair, at the average of the two vertices, with
    // the next free id.
    std::map<std::pair<uint32_t, uint32_t>, uint32_t> edge_node;
    for (int64_t q = 0; q < m.n_elem; ++q) {
      uint32_t* v = &m.conn[q * 10];
      for (int ed = 0; ed < 6; ++ed) {
        uint32_t a = v[TET_EDGE[ed][0]], b = v[TET_EDGE[ed][1]];
        if (a > b) std::swap(a, b);
        auto it = edge_node.find({a, b});
        if (it == edge_node.end()) {
          const uint32_t nid = (uint32_t)m.n_nodes++;
          for (int d = 0; d < 3; ++d) {
            double loc = 0;
            loc += m.xyz[3 * (size_t)a + d];
            loc += m.xyz[3 * (size_t)b + d];
            loc /= 2.0;
            m.xyz.push_back(loc);
          }
          m.bmask.push_back(m.bmask[a] & m.bmask[b]);
          it = edge_node.emplace(std::make_pair(a, b), nid).first;
        }
        v[4 + ed] = it->second;
      }
    }
  }
  // prepare_for_use(): orphan (unused HEX27 mid-edge) nodes are deleted and, by default, nodes renumbered
  if (libmesh_renumber || type == ORC_TET10 || type == ORC_HEX20) renumber_nodes(m, libmesh_renumber);
  m.n_parts = 1;
  m.elem_pid.assign(m.n_elem, 0);
  m.node_pid.assign(m.n_nodes, 0);
  return true;
}

}  // namespace orc

extern "C" {

orc_mesh* orc_mesh_build_cube(int elem_type, int nx, int ny, int nz, double xmin, double xmax, double ymin, double ymax,
                              double zmin, double zmax, int renumber_nodes) {
  orc_mesh* h = new orc_mesh;
  if (!orc::build_cube(h->m, elem_type, nx, ny, nz, xmin, xmax, ymin, ymax, zmin, zmax, renumber_nodes != 0)) {
    delete h;
    return nullptr;
  }
  return h;
}
// A mesh given by its arrays (tests of meshes build_cube cannot make: locally refined meshes with hanging nodes)
orc_mesh* orc_mesh_from_arrays(int elem_type, int64_t n_elem, int64_t n_nodes, const double* xyz, const uint32_t* conn, const uint8_t* mask) {
  if ((elem_type != ORC_HEX8 && elem_type != ORC_HEX27 && elem_type != ORC_TET10 && elem_type != ORC_HEX20 && elem_type != ORC_PRISM18) || n_elem < 1 || n_nodes < 1 || !xyz || !conn) {
    orc::set_error("mesh_from_arrays: bad arguments");
    return nullptr;
  }
  orc_mesh* h = new orc_mesh;
  orc::Mesh& m = h->m;
  m.type = m.npe = elem_type;
  m.n_elem = n_elem; m.n_nodes = n_nodes;
  m.xyz.assign(xyz, xyz + 3 * n_nodes);
  m.conn.assign(conn, conn + n_elem * elem_type);
  if (mask) m.bmask.assign(mask, mask + n_nodes); else m.bmask.assign(n_nodes, 0);
  for (uint32_t c : m.conn) if (c >= (uint64_t)n_nodes) { orc::set_error("mesh_from_arrays: node id out of range"); delete h; return nullptr; }
  m.n_parts = 1;
  m.elem_pid.assign(n_elem, 0);
  m.node_pid.assign(n_nodes, 0);
  return h;
}
void orc_mesh_free(orc_mesh* h) { delete h; }
int64_t orc_mesh_n_elem(const orc_mesh* h) { return h->m.n_elem; }
int64_t orc_mesh_n_nodes(const orc_mesh* h) { return h->m.n_nodes; }
int orc_mesh_nodes_per_elem(const orc_mesh* h) { return h->m.npe; }
void orc_mesh_get(const orc_mesh* h, double* xyz, uint32_t* conn, uint8_t* mask) {
  const orc::Mesh& m = h->m;
  if (xyz) std::memcpy(xyz, m.xyz.data(), sizeof(double) * 3 * m.n_nodes);
  if (conn) std::memcpy(conn, m.conn.data(), sizeof(uint32_t) * m.n_elem * m.npe);
  if (mask) std::memcpy(mask, m.bmask.data(), m.n_nodes);
}

void orc_mesh_skew(orc_mesh* h, int kind, double amplitude, uint64_t seed) {
  orc::Mesh& m = h->m;
  if (kind == 0) {  // tests/fe/fe_test.h:40-64 SkewFunc (an affine shear: elements stay affine)
    for (int64_t i = 0; i < m.n_nodes; ++i) {
      double* p = &m.xyz[3 * i];
      const double x = p[0], y = p[1], z = p[2];
      p[0] = x + .1 * y;
      p[1] = y + .2 * z;
      p[2] = z;
    }
  } else {  // seeded perturbation of interior nodes: exercises the general (non-affine) Jacobian branch
    std::mt19937_64 rng(seed);
    std::uniform_real_distribution<double> U(-1.0, 1.0);
    for (int64_t i = 0; i < m.n_nodes; ++i) {
      const double dx = U(rng), dy = U(rng), dz = U(rng);
      if (m.bmask[i]) continue;
      m.xyz[3 * i] += amplitude * dx;
      m.xyz[3 * i + 1] += amplitude * dy;
      m.xyz[3 * i + 2] += amplitude * dz;
    }
  }
}

// The (element, side) pairs BoundaryInfo holds for one boundary id after build_cube (mesh_generation.C:1416-1432,
// :1913-1926: side s of a lattice-edge element gets id s' iff the side lies in that face of the box), recovered from
// the per-node masks: a side lies in face b iff all its nodes do.  Element-major, side-minor order.
int64_t orc_mesh_boundary_sides(const orc_mesh* h, int boundary_id, int64_t cap, uint32_t* elems, uint8_t* sides) {
  const orc::Mesh& m = h->m;
  int sn[9];
  if (!orc::side_nodes(m.type, 0, sn)) return -1;
  int64_t k = 0;
  for (int64_t e = 0; e < m.n_elem; ++e)
    for (int s = 0; s < 6; ++s) {
      const int ns = orc::side_nodes(m.type, s, sn);
      bool all = true;
      for (int i = 0; i < ns && all; ++i) all = (m.bmask[m.conn[e * m.npe + sn[i]]] >> boundary_id) & 1;
      if (!all) continue;
      if (k < cap) {
        if (elems) elems[k] = (uint32_t)e;
        if (sides) sides[k] = (uint8_t)s;
      }
      ++k;
    }
  return k;
}

void orc_mesh_partition_linear(orc_mesh* h, int n_parts) {
  orc::Mesh& m = h->m;
  m.n_parts = n_parts;
  m.elem_pid.assign(m.n_elem, 0);
  if (n_parts > 1) {  // linear_partitioner.C:60-75: remainder elements fall back to processor 0
    const int64_t blksize = m.n_elem / n_parts;
    for (int64_t e = 0; e < m.n_elem; ++e) m.elem_pid[e] = (e / blksize) < n_parts ? (int32_t)(e / blksize) : 0;
  }
  m.compute_node_pids();
}

void orc_mesh_set_partition(orc_mesh* h, int n_parts, const int32_t* elem_pid) {
  orc::Mesh& m = h->m;
  m.n_parts = n_parts;
  m.elem_pid.assign(elem_pid, elem_pid + m.n_elem);
  m.compute_node_pids();
}

void orc_mesh_get_partition(const orc_mesh* h, int32_t* elem_pid, int32_t* node_pid) {
  const orc::Mesh& m = h->m;
  if (elem_pid) std::memcpy(elem_pid, m.elem_pid.data(), sizeof(int32_t) * m.n_elem);
  if (node_pid) std::memcpy(node_pid, m.node_pid.data(), sizeof(int32_t) * m.n_nodes);
}

}  // extern "C"
