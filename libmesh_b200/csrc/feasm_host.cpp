// feasm_host.cpp — host-side input generators for stand-alone runs (bench.py, smoke) where no libMesh
// process supplies the mesh and the DofMap.  Product code: nothing here touches oracle/.
//
// feasm_build_cube: the HEX8 / HEX27 lattice of MeshTools::Generation::build_cube
// (src/mesh/mesh_generation.C:1285-1340 nodes, :1399-1432 / :1575-1626 elements, :1779-1785 rescale), written as
// a closed form over the lattice: element (ei,ej,ek) owns the lattice cell whose corner is stride*(ei,ej,ek) and
// local node a sits at corner + offset[a], offset[a] = (reference coordinate of node a + 1) * stride / 2.
// Node ids are the natural lattice ids (libMesh's post-prepare_for_use renumbering only relabels nodes; the
// assembled system does not depend on it).
// feasm_distribute_dofs: first-touch numbering of DofMap::distribute_local_dofs_var_major
// (src/base/dof_map.C:1417-1535) with libMesh's node-ownership rule (src/geom/node.C:99-102).
#include <algorithm>
#include <cstring>
#include <limits>
#include <unordered_map>
#include <vector>

#include "feasm.h"

namespace {
// reference coordinates (-1,0,1) of HEX27 nodes in libMesh element-local order; the first 8 are HEX8
const signed char kRef[27][3] = {
    {-1, -1, -1}, {1, -1, -1}, {1, 1, -1}, {-1, 1, -1}, {-1, -1, 1}, {1, -1, 1}, {1, 1, 1},  {-1, 1, 1},  {0, -1, -1},
    {1, 0, -1},   {0, 1, -1},  {-1, 0, -1}, {-1, -1, 0}, {1, -1, 0},  {1, 1, 0},  {-1, 1, 0}, {0, -1, 1},  {1, 0, 1},
    {0, 1, 1},    {-1, 0, 1},  {0, 0, -1}, {0, -1, 0},  {1, 0, 0},   {0, 1, 0},  {-1, 0, 0}, {0, 0, 1},   {0, 0, 0}};
// HEX27 local nodes of each side: 4 corners in side order, then the side's centre node (src/geom/cell_hex27.C:37-45,
// columns 0-3 and 8) — all the 24-tet split needs.
const signed char kSide[6][5] = {{0, 3, 2, 1, 20}, {0, 1, 5, 4, 21}, {1, 2, 6, 5, 22}, {2, 3, 7, 6, 23}, {3, 0, 4, 7, 24}, {4, 5, 6, 7, 25}};
// TET10 mid-edge node 4+e lies between these two vertices (src/geom/cell_tet10.C edge_nodes_map)
const signed char kTetEdge[6][2] = {{0, 1}, {1, 2}, {0, 2}, {0, 3}, {1, 3}, {2, 3}};

// build_cube(TET10) (src/mesh/mesh_generation.C:1795-1907): every HEX27 cell is cut into 24 tets
// (side vertex t, side centre, side vertex t+1, cell centre), then all_second_order() adds one new node per tet
// edge at first sight while walking the tets in order.  The HEX27 mid-edge lattice nodes end up unused and are
// dropped (prepare_for_use); ids are compacted in ascending old-id order.
int build_tet10_cube(int nx, int ny, int nz, const double b[6], double* xyz, uint32_t* conn, uint8_t* mask) {
  const int64_t gx = 2 * (int64_t)nx, gy = 2 * (int64_t)ny, gz = 2 * (int64_t)nz, sx = gx + 1, sy = gy + 1;
  const int64_t n_lat = sx * sy * (gz + 1), n_cell = (int64_t)nx * ny * nz;
  if (n_lat + 40 * n_cell >= 0xffffffffll) return FEASM_ERR_ARG;
  std::vector<uint32_t> lat_new(n_lat, 0xffffffffu);  // lattice id -> compact id (vertices only)
  uint32_t next = 0;
  for (int64_t id = 0; id < n_lat; ++id) {
    const int odd = (int)((id % sx) & 1) + (int)(((id / sx) % sy) & 1) + (int)((id / (sx * sy)) & 1);
    if (odd != 1) lat_new[id] = next++;  // corners (0 odd), side centres (2), cell centres (3); mid-edges (1) are orphans
  }
  const uint32_t n_vert = next;
  auto put_vertex = [&](int64_t id) {
    const int64_t i = id % sx, j = (id / sx) % sy, k = id / (sx * sy);
    const uint32_t n = lat_new[id];
    if (xyz) {
      xyz[3 * (size_t)n + 0] = (double)i / (double)gx * (b[1] - b[0]) + b[0];
      xyz[3 * (size_t)n + 1] = (double)j / (double)gy * (b[3] - b[2]) + b[2];
      xyz[3 * (size_t)n + 2] = (double)k / (double)gz * (b[5] - b[4]) + b[4];
    }
    if (mask) mask[n] = (uint8_t)((k == 0) << 0 | (j == 0) << 1 | (i == gx) << 2 | (j == gy) << 3 | (i == 0) << 4 | (k == gz) << 5);
  };
  if (xyz || mask)
    for (int64_t id = 0; id < n_lat; ++id)
      if (lat_new[id] != 0xffffffffu) put_vertex(id);
  if (!conn) return FEASM_OK;
  if (!xyz || !mask) return FEASM_ERR_ARG;  // mid-edge nodes are averages of vertex data
  int64_t off[27];
  for (int a = 0; a < 27; ++a) off[a] = (kRef[a][0] + 1) + sx * ((kRef[a][1] + 1) + sy * (int64_t)(kRef[a][2] + 1));
  std::unordered_map<uint64_t, uint32_t> mid;
  mid.reserve((size_t)(n_cell * 40));
  (void)n_vert;
  int64_t t = 0;
  for (int64_t c = 0; c < n_cell; ++c) {
    const int64_t ci = c % nx, cj = (c / nx) % ny, ck = c / ((int64_t)nx * ny);
    const int64_t corner = 2 * ci + sx * (2 * cj + sy * (2 * ck));
    for (int s = 0; s < 6; ++s)
      for (int sub = 0; sub < 4; ++sub, ++t) {
        uint32_t* v = conn + t * 10;
        v[0] = lat_new[corner + off[kSide[s][sub]]];
        v[1] = lat_new[corner + off[kSide[s][4]]];
        v[2] = lat_new[corner + off[kSide[s][(sub + 1) & 3]]];
        v[3] = lat_new[corner + off[26]];
        for (int e = 0; e < 6; ++e) {
          uint32_t p = v[kTetEdge[e][0]], q = v[kTetEdge[e][1]];
          if (p > q) std::swap(p, q);
          auto ins = mid.emplace(((uint64_t)p << 32) | q, next);
          if (ins.second) {
            for (int d = 0; d < 3; ++d) xyz[3 * (size_t)next + d] = (xyz[3 * (size_t)p + d] + xyz[3 * (size_t)q + d]) / 2.0;
            mask[next] = mask[p] & mask[q];
            ++next;
          }
          v[4 + e] = ins.first->second;
        }
      }
  }
  return FEASM_OK;
}
}  // namespace

extern "C" {

int feasm_cube_sizes(int elem_type, int nx, int ny, int nz, int64_t* n_elem, int64_t* n_nodes) {
  if (elem_type == FEASM_TET10 && nx >= 1 && ny >= 1 && nz >= 1) {
    const int64_t X = nx, Y = ny, Z = nz, cells = X * Y * Z;
    const int64_t corners = (X + 1) * (Y + 1) * (Z + 1), faces = X * Y * (Z + 1) + X * (Y + 1) * Z + (X + 1) * Y * Z;
    const int64_t hex_edges = X * (Y + 1) * (Z + 1) + (X + 1) * Y * (Z + 1) + (X + 1) * (Y + 1) * Z;
    if (n_elem) *n_elem = 24 * cells;
    // vertices (corners, side centres, cell centres) + one node per tet edge (hex edges, 4 per side, 14 per cell)
    if (n_nodes) *n_nodes = corners + faces + cells + hex_edges + 4 * faces + 14 * cells;
    return FEASM_OK;
  }
  if ((elem_type != FEASM_HEX8 && elem_type != FEASM_HEX27 && elem_type != FEASM_HEX20 && elem_type != FEASM_PRISM18) || nx < 1 || ny < 1 || nz < 1) return FEASM_ERR_ARG;
  const int s = elem_type == FEASM_HEX8 ? 1 : 2;
  if (n_elem) *n_elem = (int64_t)nx * ny * nz * (elem_type == FEASM_PRISM18 ? 2 : 1);  // two prisms per lattice cell
  if (n_nodes) *n_nodes = (int64_t)(s * nx + 1) * (s * ny + 1) * (s * nz + 1);
  if (n_nodes && elem_type == FEASM_HEX20) {  // the HEX27 lattice without its face and cell centres: corners + one node per edge
    const int64_t X = nx, Y = ny, Z = nz;
    *n_nodes = (X + 1) * (Y + 1) * (Z + 1) + X * (Y + 1) * (Z + 1) + (X + 1) * Y * (Z + 1) + (X + 1) * (Y + 1) * Z;
  }
  return FEASM_OK;
}

int feasm_build_cube(int elem_type, int nx, int ny, int nz, const double bounds[6], double* xyz, uint32_t* conn,
                     uint8_t* mask) {
  int64_t ne, nn;
  if (feasm_cube_sizes(elem_type, nx, ny, nz, &ne, &nn) || !bounds) return FEASM_ERR_ARG;
  if (nn >= 0xffffffffll) return FEASM_ERR_ARG;
  if (elem_type == FEASM_TET10) return build_tet10_cube(nx, ny, nz, bounds, xyz, conn, mask);
  const int s = elem_type == FEASM_HEX8 ? 1 : 2;
  const int64_t gx = (int64_t)s * nx, gy = (int64_t)s * ny, gz = (int64_t)s * nz;
  const int64_t sx = gx + 1, sy = gy + 1;
  if (elem_type == FEASM_HEX20) {
    // build_cube(HEX20) (src/mesh/mesh_generation.C:1315-1345 nodes, :1559-1592 elements): the HEX27 lattice is laid down,
    // every element takes the 20 lattice points of its cell with at most one odd index (corners, edge mid-points); the
    // lattice points no element uses (two or three odd indices: face and cell centres) are deleted by prepare_for_use and
    // the ids compacted in ascending lattice order.
    const int64_t n_lat = sx * sy * (gz + 1);
    std::vector<uint32_t> compact(n_lat, 0xffffffffu);
    uint32_t next = 0;
    for (int64_t id = 0; id < n_lat; ++id) {
      const int64_t i = id % sx, j = (id / sx) % sy, k = id / (sx * sy);
      if ((i & 1) + (j & 1) + (k & 1) > 1) continue;
      const uint32_t n = compact[id] = next++;
      if (xyz) {
        xyz[3 * (size_t)n + 0] = (double)i / (double)gx * (bounds[1] - bounds[0]) + bounds[0];
        xyz[3 * (size_t)n + 1] = (double)j / (double)gy * (bounds[3] - bounds[2]) + bounds[2];
        xyz[3 * (size_t)n + 2] = (double)k / (double)gz * (bounds[5] - bounds[4]) + bounds[4];
      }
      if (mask) mask[n] = (uint8_t)((k == 0) << 0 | (j == 0) << 1 | (i == gx) << 2 | (j == gy) << 3 | (i == 0) << 4 | (k == gz) << 5);
    }
    if ((int64_t)next != nn) return FEASM_ERR_ARG;
    if (conn) {
      int64_t off[20];
      for (int a = 0; a < 20; ++a) off[a] = (kRef[a][0] + 1) + sx * ((kRef[a][1] + 1) + sy * (int64_t)(kRef[a][2] + 1));
      for (int64_t e = 0; e < ne; ++e) {
        const int64_t ei = e % nx, ej = (e / nx) % ny, ek = e / ((int64_t)nx * ny);
        const int64_t corner = 2 * ei + sx * (2 * ej + sy * (2 * ek));
        for (int a = 0; a < 20; ++a) conn[e * 20 + a] = compact[corner + off[a]];
      }
    }
    return FEASM_OK;
  }
  if (xyz || mask)
    for (int64_t id = 0; id < nn; ++id) {
      const int64_t i = id % sx, j = (id / sx) % sy, k = id / (sx * sy);
      if (xyz) {
        // unit-cube coordinate first, then the in-place affine rescale — the two-step arithmetic of the reference
        xyz[3 * id + 0] = (double)i / (double)gx * (bounds[1] - bounds[0]) + bounds[0];
        xyz[3 * id + 1] = (double)j / (double)gy * (bounds[3] - bounds[2]) + bounds[2];
        xyz[3 * id + 2] = (double)k / (double)gz * (bounds[5] - bounds[4]) + bounds[4];
      }
      if (mask)  // boundary ids: 0 z-min, 1 y-min, 2 x-max, 3 y-max, 4 x-min, 5 z-max
        mask[id] = (uint8_t)((k == 0) << 0 | (j == 0) << 1 | (i == gx) << 2 | (j == gy) << 3 | (i == 0) << 4 | (k == gz) << 5);
    }
  if (conn && elem_type == FEASM_PRISM18) {
    // build_cube(PRISM18) (src/mesh/mesh_generation.C:1634-1756): every cell of the HEX27 lattice is cut along the diagonal
    // (i+2, j) - (i, j+2) into two prisms, first the one at the cell's corner (i, j), then the one at (i+2, j+2).  With the
    // reference prism's nodes at (xi, eta, zeta) (triangle (0,0) (1,0) (0,1) times [-1, 1]), the lattice offset of a node is
    // the affine image of 2 (xi, eta): first prism (2 xi, 2 eta), second (2 - 2 eta, 2 xi + 2 eta); zeta + 1 along k.
    static const signed char kPrism2[18][3] = {{0, 0, -1}, {2, 0, -1}, {0, 2, -1}, {0, 0, 1}, {2, 0, 1}, {0, 2, 1}, {1, 0, -1}, {1, 1, -1}, {0, 1, -1},
                                               {0, 0, 0},  {2, 0, 0},  {0, 2, 0},  {1, 0, 1}, {1, 1, 1}, {0, 1, 1}, {1, 0, 0},  {1, 1, 0},  {0, 1, 0}};  // (2 xi, 2 eta, zeta)
    int64_t off[2][18];
    for (int a = 0; a < 18; ++a) {
      const int x2 = kPrism2[a][0], e2 = kPrism2[a][1], dk = kPrism2[a][2] + 1;
      off[0][a] = x2 + sx * (e2 + sy * (int64_t)dk);
      off[1][a] = (2 - e2) + sx * ((x2 + e2) + sy * (int64_t)dk);
    }
    const int64_t ncell = (int64_t)nx * ny * nz;
    for (int64_t c = 0; c < ncell; ++c) {
      const int64_t ei = c % nx, ej = (c / nx) % ny, ek = c / ((int64_t)nx * ny);
      const int64_t corner = 2 * ei + sx * (2 * ej + sy * (2 * ek));
      for (int h = 0; h < 2; ++h)
        for (int a = 0; a < 18; ++a) conn[(2 * c + h) * 18 + a] = (uint32_t)(corner + off[h][a]);
    }
    return FEASM_OK;
  }
  if (conn) {
    const int npe = elem_type;
    int64_t off[27];
    for (int a = 0; a < npe; ++a)
      off[a] = ((kRef[a][0] + 1) * s / 2) + sx * (((kRef[a][1] + 1) * s / 2) + sy * (int64_t)((kRef[a][2] + 1) * s / 2));
    for (int64_t e = 0; e < ne; ++e) {
      const int64_t ei = e % nx, ej = (e / nx) % ny, ek = e / ((int64_t)nx * ny);
      const int64_t corner = s * ei + sx * (s * ej + sy * (s * ek));
      for (int a = 0; a < npe; ++a) conn[e * npe + a] = (uint32_t)(corner + off[a]);
    }
  }
  return FEASM_OK;
}

int64_t feasm_distribute_dofs(int npe, int64_t n_elem, int64_t n_nodes, const uint32_t* conn, int n_vars, int n_parts,
                              const int32_t* elem_pid, uint32_t* node_dof_base, uint32_t* first_dof, int32_t* node_pid_out) {
  if (npe < 1 || n_elem < 0 || n_nodes < 1 || !conn || n_vars < 1 || n_parts < 1 || !node_dof_base) return FEASM_ERR_ARG;
  const uint32_t none = std::numeric_limits<uint32_t>::max();
  std::fill(node_dof_base, node_dof_base + n_nodes, none);
  // a node belongs to the lowest rank among the elements that touch it
  std::vector<int32_t> owner;
  if (n_parts > 1) {
    if (!elem_pid) return FEASM_ERR_ARG;
    owner.assign(n_nodes, std::numeric_limits<int32_t>::max());
    for (int64_t e = 0; e < n_elem; ++e)
      for (int a = 0; a < npe; ++a) owner[conn[e * npe + a]] = std::min(owner[conn[e * npe + a]], elem_pid[e]);
  }
  uint64_t next = 0;
  // per-rank element lists keep the walk linear in n_elem
  std::vector<int64_t> start(n_parts + 1, 0);
  std::vector<int64_t> order;
  if (n_parts > 1) {
    for (int64_t e = 0; e < n_elem; ++e) {
      if (elem_pid[e] < 0 || elem_pid[e] >= n_parts) return FEASM_ERR_ARG;
      start[elem_pid[e] + 1]++;
    }
    for (int p = 0; p < n_parts; ++p) start[p + 1] += start[p];
    order.resize(n_elem);
    std::vector<int64_t> cur(start.begin(), start.end() - 1);
    for (int64_t e = 0; e < n_elem; ++e) order[cur[elem_pid[e]]++] = e;
  } else {
    start[1] = n_elem;
  }
  for (int p = 0; p < n_parts; ++p) {
    if (first_dof) first_dof[p] = (uint32_t)next;
    for (int64_t k = start[p]; k < start[p + 1]; ++k) {
      const int64_t e = n_parts > 1 ? order[k] : k;
      for (int a = 0; a < npe; ++a) {
        const uint32_t node = conn[e * npe + a];
        if (node_dof_base[node] != none) continue;
        if (n_parts > 1 && owner[node] != p) continue;
        node_dof_base[node] = (uint32_t)next;
        next += (uint64_t)n_vars;
      }
    }
  }
  if (next >= none) return FEASM_ERR_ARG;
  if (first_dof) first_dof[n_parts] = (uint32_t)next;
  if (node_pid_out)
    for (int64_t i = 0; i < n_nodes; ++i) node_pid_out[i] = n_parts > 1 ? owner[i] : 0;
  return (int64_t)next;
}

}  // extern "C"
