// feasm_ingest.cu — device-side mesh ingest: the step immediately BEFORE the assembly path.
//
//   feasm_generate_cube          MeshTools::Generation::build_cube for HEX8 / HEX27 / TET10 (src/mesh/mesh_generation.C:
//                                1285-1340 nodes, :1399-1432 / :1575-1626 elements, :1779-1785 rescale, :1416-1432
//                                boundary ids) written straight into the context's device arrays: one thread per
//                                node / per (element, local node), closed form over the lattice.
//   feasm_distribute_dofs_device DofMap::distribute_local_dofs_var_major (src/base/dof_map.C:1445-1507) + dof_indices
//                                (:2280-2308) for one variable group on one rank: "first touch while walking the
//                                elements in order" = the minimum incidence id e*npe+a per node (atomicMin: the result
//                                does not depend on the order the atomics land in), an exclusive scan over the
//                                first-touch flags gives the node's place in the walk, dof = n_vars * place + var.
//                                Works for ANY mesh in the context (generated here or handed over by feasm_set_mesh).
//   feasm_generate_dirichlet     DirichletBoundary on boundary ids (DofMap::create_dof_constraints,
//                                src/base/dof_map_constraints.C:1736: for LAGRANGE the projection is nodal
//                                interpolation): constrained rows = dofs of the nodes whose boundary
//                                mask meets id_mask, value = 0 or the ex4 exact solution at the node.
//
// Product code: never includes oracle/.  Setup path, HBM-bound and tiny next to one assembly (96^3 HEX27: 7.2 M
// nodes, 24 M incidences — well under a millisecond of traffic); kernels are plain grid-stride loops with coalesced
// stores.
#include <algorithm>
#include <cub/cub.cuh>

#include <vector>

#include "feasm_internal.h"

namespace {

// reference coordinates (-1,0,1) of the HEX27 nodes in libMesh element-local order (src/geom/cell_hex27.C); the first
// eight are HEX8
__constant__ signed char c_ref[27][3] = {
    {-1, -1, -1}, {1, -1, -1}, {1, 1, -1}, {-1, 1, -1}, {-1, -1, 1}, {1, -1, 1}, {1, 1, 1},  {-1, 1, 1},  {0, -1, -1},
    {1, 0, -1},   {0, 1, -1},  {-1, 0, -1}, {-1, -1, 0}, {1, -1, 0},  {1, 1, 0},  {-1, 1, 0}, {0, -1, 1},  {1, 0, 1},
    {0, 1, 1},    {-1, 0, 1},  {0, 0, -1}, {0, -1, 0},  {1, 0, 0},   {0, 1, 0},  {-1, 0, 0}, {0, 0, 1},   {0, 0, 0}};

struct Lattice {
  int s;                // lattice steps per element edge: 1 (HEX8) or 2 (HEX27)
  int64_t nx, ny, nz;   // elements
  int64_t gx, gy, gz;   // lattice intervals
  int64_t sx, sy;       // lattice strides
  double b[6];
};

__global__ void k_cube_nodes(double* __restrict__ xyz, uint8_t* __restrict__ mask, const Lattice L, const int64_t n_nodes) {
  for (int64_t id = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; id < n_nodes; id += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = id % L.sx, j = (id / L.sx) % L.sy, k = id / (L.sx * L.sy);
    // unit-cube coordinate first, then the in-place rescale: the two-step arithmetic of the reference (App. A.1);
    // explicit _rn so that no FMA contraction changes the last bit
    xyz[3 * id + 0] = __dadd_rn(__dmul_rn(__ddiv_rn((double)i, (double)L.gx), __dadd_rn(L.b[1], -L.b[0])), L.b[0]);
    xyz[3 * id + 1] = __dadd_rn(__dmul_rn(__ddiv_rn((double)j, (double)L.gy), __dadd_rn(L.b[3], -L.b[2])), L.b[2]);
    xyz[3 * id + 2] = __dadd_rn(__dmul_rn(__ddiv_rn((double)k, (double)L.gz), __dadd_rn(L.b[5], -L.b[4])), L.b[4]);
    // boundary ids: 0 z-min, 1 y-min, 2 x-max, 3 y-max, 4 x-min, 5 z-max
    mask[id] = (uint8_t)((k == 0) << 0 | (j == 0) << 1 | (i == L.gx) << 2 | (j == L.gy) << 3 | (i == 0) << 4 | (k == L.gz) << 5);
  }
}

__global__ void k_cube_conn(uint32_t* __restrict__ conn, const Lattice L, const int npe, const int64_t n_inc) {
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < n_inc; t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = t / npe;
    const int a = (int)(t - e * npe);
    const int64_t ei = e % L.nx, ej = (e / L.nx) % L.ny, ek = e / (L.nx * L.ny);
    const int64_t i = L.s * ei + (c_ref[a][0] + 1) * L.s / 2, j = L.s * ej + (c_ref[a][1] + 1) * L.s / 2,
                  k = L.s * ek + (c_ref[a][2] + 1) * L.s / 2;
    conn[t] = (uint32_t)(i + L.sx * (j + L.sy * k));
  }
}

__global__ void k_first_touch(uint32_t* __restrict__ first, const uint32_t* __restrict__ conn, const int64_t n_inc) {
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < n_inc; t += (int64_t)gridDim.x * blockDim.x)
    atomicMin(&first[conn[t]], (uint32_t)t);
}

__global__ void k_touch_flags(uint32_t* __restrict__ flag, const uint32_t* __restrict__ first, const uint32_t* __restrict__ conn,
                              const int64_t n_inc) {
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < n_inc; t += (int64_t)gridDim.x * blockDim.x)
    flag[t] = first[conn[t]] == (uint32_t)t;
}

// place[t] = number of nodes first touched before incidence t; a node's base dof is n_vars * place[first[node]]
__global__ void k_node_base(uint32_t* __restrict__ node_base, const uint32_t* __restrict__ first, const uint32_t* __restrict__ place,
                            const int n_vars, const int64_t n_nodes) {
  for (int64_t n = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; n < n_nodes; n += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t f = first[n];
    node_base[n] = f == 0xffffffffu ? 0xffffffffu : (uint32_t)n_vars * place[f];  // orphan nodes carry no dofs
  }
}

// DofMap::dof_indices: variable-major, element-local node order
__global__ void k_elem_dofs(uint32_t* __restrict__ elem_dofs, const uint32_t* __restrict__ conn, const uint32_t* __restrict__ node_base,
                            const int npe, const int n_vars, const int64_t n_out) {
  const int nd = npe * n_vars;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < n_out; t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = t / nd;
    const int r = (int)(t - e * nd), v = r / npe, a = r - v * npe;
    elem_dofs[t] = node_base[conn[e * npe + a]] + (uint32_t)v;
  }
}

// introduction_ex3/exact_solution.C:40-47
__device__ __forceinline__ double ex4_exact(double x, double y, double z) {
  const double hp = .5 * 3.141592653589793;
  return __dmul_rn(__dmul_rn(cos(__dmul_rn(hp, x)), sin(__dmul_rn(hp, y))), cos(__dmul_rn(hp, z)));
}

__global__ void k_dirichlet(uint8_t* __restrict__ con_flag, double* __restrict__ con_val, unsigned long long* __restrict__ count,
                            const uint8_t* __restrict__ mask, const uint32_t* __restrict__ node_base, const double* __restrict__ xyz,
                            const unsigned id_mask, const unsigned var_mask, const int kind, const int n_vars, const int64_t n_nodes) {
  unsigned long long mine = 0;
  for (int64_t n = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; n < n_nodes; n += (int64_t)gridDim.x * blockDim.x) {
    if (!(mask[n] & id_mask) || node_base[n] == 0xffffffffu) continue;
    const double g = kind == FEASM_DIRICHLET_EX4_EXACT ? ex4_exact(xyz[3 * n], xyz[3 * n + 1], xyz[3 * n + 2]) : 0.;
    for (int v = 0; v < n_vars; ++v)
      if (var_mask >> v & 1u) {
        con_flag[node_base[n] + v] = 1;
        con_val[node_base[n] + v] = g;
        ++mine;
      }
  }
  if (mine) atomicAdd(count, mine);
}

// ---- build_cube(TET10) (src/mesh/mesh_generation.C:1795-1907) ------------------------------------------------------------
// Every HEX27 cell is cut into 24 tets (side vertex `sub`, side centre, side vertex `sub+1`, cell centre), then
// all_second_order() gives every tet edge one new node at FIRST SIGHT while walking the tets in order.  On the device:
// vertices = lattice nodes that are not HEX27 mid-edge nodes, compacted by a scan; "first sight" = the smallest
// (tet, edge) sequence number among the occurrences of an edge key, found by a stable radix sort of the keys; the new
// node ids are the ranks of those first sequence numbers (a second sort).
__constant__ signed char c_side5[6][5] = {{0, 3, 2, 1, 20}, {0, 1, 5, 4, 21}, {1, 2, 6, 5, 22}, {2, 3, 7, 6, 23}, {3, 0, 4, 7, 24}, {4, 5, 6, 7, 25}};
__constant__ signed char c_tet_edge[6][2] = {{0, 1}, {1, 2}, {0, 2}, {0, 3}, {1, 3}, {2, 3}};  // cell_tet10.C edge_nodes_map

__global__ void k_tet_vertex_flags(uint32_t* __restrict__ flag, const Lattice L, const int64_t n_lat) {
  for (int64_t id = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; id < n_lat; id += (int64_t)gridDim.x * blockDim.x) {
    const int odd = (int)((id % L.sx) & 1) + (int)(((id / L.sx) % L.sy) & 1) + (int)((id / (L.sx * L.sy)) & 1);
    flag[id] = odd != 1;  // corners (0 odd), side centres (2), cell centres (3); mid-edges (1) are orphans
  }
}

__global__ void k_tet_vertices(double* __restrict__ xyz, uint8_t* __restrict__ mask, const uint32_t* __restrict__ flag,
                               const uint32_t* __restrict__ lat_new, const Lattice L, const int64_t n_lat) {
  for (int64_t id = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; id < n_lat; id += (int64_t)gridDim.x * blockDim.x) {
    if (!flag[id]) continue;
    const int64_t i = id % L.sx, j = (id / L.sx) % L.sy, k = id / (L.sx * L.sy);
    const uint32_t n = lat_new[id];
    xyz[3 * (size_t)n + 0] = __dadd_rn(__dmul_rn(__ddiv_rn((double)i, (double)L.gx), __dadd_rn(L.b[1], -L.b[0])), L.b[0]);
    xyz[3 * (size_t)n + 1] = __dadd_rn(__dmul_rn(__ddiv_rn((double)j, (double)L.gy), __dadd_rn(L.b[3], -L.b[2])), L.b[2]);
    xyz[3 * (size_t)n + 2] = __dadd_rn(__dmul_rn(__ddiv_rn((double)k, (double)L.gz), __dadd_rn(L.b[5], -L.b[4])), L.b[4]);
    mask[n] = (uint8_t)((k == 0) << 0 | (j == 0) << 1 | (i == L.gx) << 2 | (j == L.gy) << 3 | (i == 0) << 4 | (k == L.gz) << 5);
  }
}

// vertices of tet t into conn[t][0..3]; the six edge keys (min vertex << 32 | max vertex) with their sequence numbers
__global__ void k_tet_vertices_and_keys(uint32_t* __restrict__ conn, uint64_t* __restrict__ key, uint32_t* __restrict__ seq,
                                        const uint32_t* __restrict__ lat_new, const Lattice L, const int64_t n_tet) {
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < n_tet; t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t c = t / 24;
    const int s = (int)((t % 24) / 4), sub = (int)(t % 4);
    const int64_t ci = c % L.nx, cj = (c / L.nx) % L.ny, ck = c / (L.nx * L.ny);
    const int64_t corner = 2 * ci + L.sx * (2 * cj + L.sy * (2 * ck));
    auto lat = [&](int a) { return corner + (c_ref[a][0] + 1) + L.sx * ((c_ref[a][1] + 1) + L.sy * (int64_t)(c_ref[a][2] + 1)); };
    uint32_t v[4];
    v[0] = lat_new[lat(c_side5[s][sub])];
    v[1] = lat_new[lat(c_side5[s][4])];
    v[2] = lat_new[lat(c_side5[s][(sub + 1) & 3])];
    v[3] = lat_new[lat(26)];
#pragma unroll
    for (int a = 0; a < 4; ++a) conn[t * 10 + a] = v[a];
#pragma unroll
    for (int e = 0; e < 6; ++e) {
      uint32_t p = v[c_tet_edge[e][0]], q = v[c_tet_edge[e][1]];
      if (p > q) { const uint32_t w = p; p = q; q = w; }
      key[t * 6 + e] = ((uint64_t)p << 32) | q;
      seq[t * 6 + e] = (uint32_t)(t * 6 + e);
    }
  }
}

// sorted by key (stable): head[k] = 1 where a new key starts; the head's seq is the edge's first sight
__global__ void k_edge_heads(uint32_t* __restrict__ head, const uint64_t* __restrict__ key_s, const int64_t n) {
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x)
    head[k] = k == 0 || key_s[k] != key_s[k - 1];
}

// compact the heads: first_seq[u] = seq of the head of segment u, seg_id[u] = u
__global__ void k_edge_first(uint32_t* __restrict__ first_seq, uint32_t* __restrict__ seg_id, const uint32_t* __restrict__ head,
                             const uint32_t* __restrict__ seg_of, const uint32_t* __restrict__ seq_s, const int64_t n) {
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x)
    if (head[k]) { first_seq[seg_of[k]] = seq_s[k]; seg_id[seg_of[k]] = seg_of[k]; }
}

// segments sorted by first sight: the r-th one gets node n_vert + r
__global__ void k_edge_rank(uint32_t* __restrict__ rank_of_seg, const uint32_t* __restrict__ seg_sorted, const int64_t n_seg) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n_seg; r += (int64_t)gridDim.x * blockDim.x)
    rank_of_seg[seg_sorted[r]] = (uint32_t)r;
}

// every (tet, edge) occurrence learns its node; the head of a segment also creates it (midpoint, boundary bits)
__global__ void k_edge_nodes(uint32_t* __restrict__ conn, double* __restrict__ xyz, uint8_t* __restrict__ mask,
                             const uint64_t* __restrict__ key_s, const uint32_t* __restrict__ seq_s, const uint32_t* __restrict__ head,
                             const uint32_t* __restrict__ seg_of, const uint32_t* __restrict__ rank_of_seg, const uint32_t n_vert,
                             const int64_t n) {
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x) {
    // seg_of is the exclusive scan of head: the segment of entry k is seg_of[k] + head[k] - 1
    const uint32_t seg = seg_of[k] + head[k] - 1u;
    const uint32_t node = n_vert + rank_of_seg[seg];
    const uint32_t sq = seq_s[k];
    conn[(uint64_t)(sq / 6) * 10 + 4 + sq % 6] = node;
    if (head[k]) {
      const uint32_t p = (uint32_t)(key_s[k] >> 32), q = (uint32_t)key_s[k];
#pragma unroll
      for (int d = 0; d < 3; ++d) xyz[3 * (size_t)node + d] = __ddiv_rn(__dadd_rn(xyz[3 * (size_t)p + d], xyz[3 * (size_t)q + d]), 2.0);
      mask[node] = mask[p] & mask[q];
    }
  }
}

int grid_for(const feasm_ctx* ctx, int64_t n) {
  const int64_t want = (n + 255) / 256, cap = (int64_t)ctx->sm_count * 16;
  return (int)(want < 1 ? 1 : (want < cap ? want : cap));
}

template <typename T>
int dev_alloc(feasm_ctx* ctx, T** p, size_t n) {
  if (*p) { cudaFree(*p); *p = nullptr; }
  FEASM_CUDA(ctx, cudaMalloc(p, sizeof(T) * (n ? n : 1)));
  return FEASM_OK;
}

}  // namespace

extern "C" {

// build_cube(TET10) on the device; see the kernels above
static int generate_tet10_cube(feasm_ctx* ctx, const Lattice& L, int64_t n_elem, int64_t n_nodes) {
  const int64_t n_lat = L.sx * L.sy * (L.gz + 1), n_occ = n_elem * 6;
  if (n_occ >= 0x7fffffffll) return feasm_fail(ctx, FEASM_ERR_ARG, "mesh too large for 32-bit edge sequence numbers");
  uint32_t *d_flag = nullptr, *d_lat_new = nullptr, *d_seq = nullptr, *d_seq_s = nullptr, *d_head = nullptr, *d_seg_of = nullptr,
           *d_first = nullptr, *d_first_s = nullptr, *d_seg = nullptr, *d_seg_s = nullptr, *d_rank = nullptr;
  uint64_t *d_key = nullptr, *d_key_s = nullptr;
  void* d_tmp = nullptr;
  auto cleanup = [&]() {
    cudaFree(d_flag); cudaFree(d_lat_new); cudaFree(d_seq); cudaFree(d_seq_s); cudaFree(d_head); cudaFree(d_seg_of); cudaFree(d_first);
    cudaFree(d_first_s); cudaFree(d_seg); cudaFree(d_seg_s); cudaFree(d_rank); cudaFree(d_key); cudaFree(d_key_s); cudaFree(d_tmp);
  };
  auto body = [&]() -> int {
    cudaStream_t st = ctx->stream;
    FEASM_CUDA(ctx, cudaMalloc(&d_flag, sizeof(uint32_t) * (n_lat + 1)));
    FEASM_CUDA(ctx, cudaMalloc(&d_lat_new, sizeof(uint32_t) * (n_lat + 1)));
    FEASM_CUDA(ctx, cudaMalloc(&d_key, sizeof(uint64_t) * n_occ));
    FEASM_CUDA(ctx, cudaMalloc(&d_key_s, sizeof(uint64_t) * n_occ));
    FEASM_CUDA(ctx, cudaMalloc(&d_seq, sizeof(uint32_t) * n_occ));
    FEASM_CUDA(ctx, cudaMalloc(&d_seq_s, sizeof(uint32_t) * n_occ));
    FEASM_CUDA(ctx, cudaMalloc(&d_head, sizeof(uint32_t) * (n_occ + 1)));
    FEASM_CUDA(ctx, cudaMalloc(&d_seg_of, sizeof(uint32_t) * (n_occ + 1)));
    size_t b0 = 0, b1 = 0, b2 = 0, b3 = 0;
    FEASM_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, b0, d_flag, d_lat_new, n_lat + 1, st));
    FEASM_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, b1, d_key, d_key_s, d_seq, d_seq_s, n_occ, 0, 64, st));
    FEASM_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, b2, d_head, d_seg_of, n_occ + 1, st));
    FEASM_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, b3, d_first, d_first_s, d_seg, d_seg_s, n_occ, 0, 32, st));
    const size_t tmp_bytes = std::max(std::max(b0, b1), std::max(b2, b3));
    FEASM_CUDA(ctx, cudaMalloc(&d_tmp, tmp_bytes));
    size_t tb = tmp_bytes;
    // vertices: lattice nodes that are not HEX27 mid-edge nodes, ids compacted in ascending lattice order
    FEASM_CUDA(ctx, cudaMemsetAsync(d_flag + n_lat, 0, sizeof(uint32_t), st));
    k_tet_vertex_flags<<<grid_for(ctx, n_lat), 256, 0, st>>>(d_flag, L, n_lat);
    FEASM_CUDA(ctx, cub::DeviceScan::ExclusiveSum(d_tmp, tb, d_flag, d_lat_new, n_lat + 1, st));
    uint32_t n_vert = 0;
    FEASM_CUDA(ctx, cudaMemcpyAsync(&n_vert, d_lat_new + n_lat, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    k_tet_vertices<<<grid_for(ctx, n_lat), 256, 0, st>>>(ctx->d_xyz, ctx->d_node_mask, d_flag, d_lat_new, L, n_lat);
    k_tet_vertices_and_keys<<<grid_for(ctx, n_elem), 256, 0, st>>>(ctx->d_conn, d_key, d_seq, d_lat_new, L, n_elem);
    // first sight of every edge: stable sort by key, the head of each run carries the smallest sequence number
    tb = tmp_bytes;
    FEASM_CUDA(ctx, cub::DeviceRadixSort::SortPairs(d_tmp, tb, d_key, d_key_s, d_seq, d_seq_s, n_occ, 0, 64, st));
    FEASM_CUDA(ctx, cudaMemsetAsync(d_head + n_occ, 0, sizeof(uint32_t), st));
    k_edge_heads<<<grid_for(ctx, n_occ), 256, 0, st>>>(d_head, d_key_s, n_occ);
    tb = tmp_bytes;
    FEASM_CUDA(ctx, cub::DeviceScan::ExclusiveSum(d_tmp, tb, d_head, d_seg_of, n_occ + 1, st));
    uint32_t n_seg = 0;
    FEASM_CUDA(ctx, cudaMemcpyAsync(&n_seg, d_seg_of + n_occ, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    FEASM_CUDA(ctx, cudaStreamSynchronize(st));
    if ((int64_t)n_vert + n_seg != n_nodes) return feasm_fail(ctx, FEASM_ERR_STATE, "tet split: node count differs from feasm_cube_sizes");
    FEASM_CUDA(ctx, cudaMalloc(&d_first, sizeof(uint32_t) * n_seg));
    FEASM_CUDA(ctx, cudaMalloc(&d_first_s, sizeof(uint32_t) * n_seg));
    FEASM_CUDA(ctx, cudaMalloc(&d_seg, sizeof(uint32_t) * n_seg));
    FEASM_CUDA(ctx, cudaMalloc(&d_seg_s, sizeof(uint32_t) * n_seg));
    FEASM_CUDA(ctx, cudaMalloc(&d_rank, sizeof(uint32_t) * n_seg));
    k_edge_first<<<grid_for(ctx, n_occ), 256, 0, st>>>(d_first, d_seg, d_head, d_seg_of, d_seq_s, n_occ);
    tb = tmp_bytes;
    FEASM_CUDA(ctx, cub::DeviceRadixSort::SortPairs(d_tmp, tb, d_first, d_first_s, d_seg, d_seg_s, (int64_t)n_seg, 0, 32, st));
    k_edge_rank<<<grid_for(ctx, n_seg), 256, 0, st>>>(d_rank, d_seg_s, n_seg);
    k_edge_nodes<<<grid_for(ctx, n_occ), 256, 0, st>>>(ctx->d_conn, ctx->d_xyz, ctx->d_node_mask, d_key_s, d_seq_s, d_head, d_seg_of, d_rank,
                                                        n_vert, n_occ);
    FEASM_CUDA(ctx, cudaGetLastError());
    FEASM_CUDA(ctx, cudaStreamSynchronize(st));
    return FEASM_OK;
  };
  const int rc = body();
  cudaStreamSynchronize(ctx->stream);
  cleanup();
  return rc;
}

int feasm_generate_cube(feasm_ctx* ctx, int elem_type, int nx, int ny, int nz, const double bounds[6]) {
  if (!ctx) return FEASM_ERR_ARG;
  if (elem_type != FEASM_HEX8 && elem_type != FEASM_HEX27 && elem_type != FEASM_TET10 && elem_type != FEASM_HEX20 && elem_type != FEASM_PRISM18)
    return feasm_fail(ctx, FEASM_ERR_ARG, "feasm_generate_cube: HEX8 / HEX20 / HEX27 / TET10 / PRISM18");
  int64_t n_elem = 0, n_nodes = 0;
  if (!bounds || feasm_cube_sizes(elem_type, nx, ny, nz, &n_elem, &n_nodes)) return feasm_fail(ctx, FEASM_ERR_ARG, "bad build_cube arguments");
  if (n_nodes >= 0xffffffffll || n_elem * elem_type >= 0xffffffffll) return feasm_fail(ctx, FEASM_ERR_ARG, "mesh too large for 32-bit ids");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  if (elem_type == FEASM_HEX20 || elem_type == FEASM_PRISM18) {
    // the compacted HEX20 numbering (lattice points with at most one odd index, ascending) and the two-prisms-per-cell
    // connectivity are made by the host generator and uploaded; dof numbering, Dirichlet data and the pattern are then
    // built on the device as for the other types
    std::vector<double> xyz((size_t)n_nodes * 3);
    std::vector<uint32_t> conn((size_t)n_elem * elem_type);
    std::vector<uint8_t> mask((size_t)n_nodes);
    if (feasm_build_cube(elem_type, nx, ny, nz, bounds, xyz.data(), conn.data(), mask.data())) return feasm_fail(ctx, FEASM_ERR_ARG, "bad build_cube arguments");
    int rc = feasm_set_mesh(ctx, elem_type, n_elem, n_nodes, xyz.data(), conn.data());
    if (rc) return rc;
    if ((rc = dev_alloc(ctx, &ctx->d_node_mask, (size_t)n_nodes))) return rc;
    FEASM_CUDA(ctx, cudaMemcpy(ctx->d_node_mask, mask.data(), mask.size(), cudaMemcpyHostToDevice));
    return FEASM_OK;
  }
  feasm_reset_mesh(ctx, elem_type, n_elem, n_nodes);
  Lattice L;
  L.s = elem_type == FEASM_HEX8 ? 1 : 2;
  L.nx = nx; L.ny = ny; L.nz = nz;
  L.gx = (int64_t)L.s * nx; L.gy = (int64_t)L.s * ny; L.gz = (int64_t)L.s * nz;
  L.sx = L.gx + 1; L.sy = L.gy + 1;
  for (int i = 0; i < 6; ++i) L.b[i] = bounds[i];
  int rc;
  if ((rc = dev_alloc(ctx, &ctx->d_xyz, (size_t)n_nodes * 3))) return rc;
  if ((rc = dev_alloc(ctx, &ctx->d_conn, (size_t)n_elem * elem_type))) return rc;
  if ((rc = dev_alloc(ctx, &ctx->d_node_mask, (size_t)n_nodes))) return rc;
  if (elem_type == FEASM_TET10) return generate_tet10_cube(ctx, L, n_elem, n_nodes);
  k_cube_nodes<<<grid_for(ctx, n_nodes), 256, 0, ctx->stream>>>(ctx->d_xyz, ctx->d_node_mask, L, n_nodes);
  k_cube_conn<<<grid_for(ctx, n_elem * elem_type), 256, 0, ctx->stream>>>(ctx->d_conn, L, elem_type, n_elem * elem_type);
  FEASM_CUDA(ctx, cudaGetLastError());
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return FEASM_OK;
}

int64_t feasm_distribute_dofs_device(feasm_ctx* ctx, int n_vars) {
  if (!ctx || !ctx->d_conn) return feasm_fail(ctx, FEASM_ERR_STATE, "set_mesh or generate_cube first");
  if (n_vars < 1 || n_vars > 8) return feasm_fail(ctx, FEASM_ERR_ARG, "bad n_vars");
  const int64_t n_inc = ctx->n_elem * ctx->npe, n_nodes = ctx->n_nodes;
  if (n_inc >= 0xffffffffll) return feasm_fail(ctx, FEASM_ERR_ARG, "mesh too large for 32-bit incidence ids");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  uint32_t *d_first = nullptr, *d_flag = nullptr, *d_place = nullptr;
  void* d_tmp = nullptr;
  size_t tmp_bytes = 0;
  auto cleanup = [&]() { cudaFree(d_first); cudaFree(d_flag); cudaFree(d_place); cudaFree(d_tmp); };
  int rc = FEASM_OK;
  auto body = [&]() -> int {
    FEASM_CUDA(ctx, cudaMalloc(&d_first, sizeof(uint32_t) * n_nodes));
    FEASM_CUDA(ctx, cudaMalloc(&d_flag, sizeof(uint32_t) * (n_inc + 1)));
    FEASM_CUDA(ctx, cudaMalloc(&d_place, sizeof(uint32_t) * (n_inc + 1)));
    FEASM_CUDA(ctx, cudaMemsetAsync(d_first, 0xff, sizeof(uint32_t) * n_nodes, ctx->stream));
    FEASM_CUDA(ctx, cudaMemsetAsync(d_flag + n_inc, 0, sizeof(uint32_t), ctx->stream));
    k_first_touch<<<grid_for(ctx, n_inc), 256, 0, ctx->stream>>>(d_first, ctx->d_conn, n_inc);
    k_touch_flags<<<grid_for(ctx, n_inc), 256, 0, ctx->stream>>>(d_flag, d_first, ctx->d_conn, n_inc);
    FEASM_CUDA(ctx, cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, d_flag, d_place, n_inc + 1, ctx->stream));
    FEASM_CUDA(ctx, cudaMalloc(&d_tmp, tmp_bytes));
    FEASM_CUDA(ctx, cub::DeviceScan::ExclusiveSum(d_tmp, tmp_bytes, d_flag, d_place, n_inc + 1, ctx->stream));
    uint32_t n_numbered = 0;  // = place[n_inc]: nodes that belong to at least one element
    FEASM_CUDA(ctx, cudaMemcpyAsync(&n_numbered, d_place + n_inc, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const int64_t n_rows = (int64_t)n_numbered * n_vars;
    if (n_rows <= 0 || n_rows >= 0xffffffffll) return feasm_fail(ctx, FEASM_ERR_ARG, "dof ids are 32-bit (dof_id_type)");
    int r2;
    if ((r2 = dev_alloc(ctx, &ctx->d_node_base, (size_t)n_nodes))) return r2;
    if ((r2 = dev_alloc(ctx, &ctx->d_elem_dofs, (size_t)n_inc * n_vars))) return r2;
    if (ctx->d_gid) { cudaFree(ctx->d_gid); ctx->d_gid = nullptr; }
    k_node_base<<<grid_for(ctx, n_nodes), 256, 0, ctx->stream>>>(ctx->d_node_base, d_first, d_place, n_vars, n_nodes);
    k_elem_dofs<<<grid_for(ctx, n_inc * n_vars), 256, 0, ctx->stream>>>(ctx->d_elem_dofs, ctx->d_conn, ctx->d_node_base, ctx->npe, n_vars,
                                                                         n_inc * n_vars);
    FEASM_CUDA(ctx, cudaGetLastError());
    return feasm_reset_dofs(ctx, n_vars, n_rows, n_rows);
  };
  rc = body();
  cudaStreamSynchronize(ctx->stream);
  cleanup();
  return rc ? rc : ctx->n_rows;
}

int64_t feasm_generate_dirichlet(feasm_ctx* ctx, unsigned id_mask, int kind, unsigned var_mask) {
  if (!ctx || !ctx->d_con_flag) return feasm_fail(ctx, FEASM_ERR_STATE, "distribute dofs first");
  if (!ctx->d_node_mask || !ctx->d_node_base)
    return feasm_fail(ctx, FEASM_ERR_STATE, "feasm_generate_dirichlet needs feasm_generate_cube + feasm_distribute_dofs_device (boundary ids and the node -> dof table live on the device)");
  if (kind != FEASM_DIRICHLET_ZERO && kind != FEASM_DIRICHLET_EX4_EXACT) return feasm_fail(ctx, FEASM_ERR_ARG, "bad Dirichlet kind");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  unsigned long long* d_count = nullptr;
  FEASM_CUDA(ctx, cudaMalloc(&d_count, sizeof(unsigned long long)));
  cudaMemsetAsync(d_count, 0, sizeof(unsigned long long), ctx->stream);
  cudaMemsetAsync(ctx->d_con_flag, 0, ctx->n_rows, ctx->stream);
  cudaMemsetAsync(ctx->d_con_val, 0, sizeof(double) * ctx->n_rows, ctx->stream);
  k_dirichlet<<<grid_for(ctx, ctx->n_nodes), 256, 0, ctx->stream>>>(ctx->d_con_flag, ctx->d_con_val, d_count, ctx->d_node_mask, ctx->d_node_base,
                                                                     ctx->d_xyz, id_mask, var_mask, kind, ctx->n_vars, ctx->n_nodes);
  unsigned long long n = 0;
  cudaMemcpyAsync(&n, d_count, sizeof(n), cudaMemcpyDeviceToHost, ctx->stream);
  // the flags on the device no longer mirror a host-given list: a later feasm_set_constraints clears them wholesale
  ctx->h_con_rows.clear();
  ctx->n_con_dev_prev = 0;
  ctx->con_flags_untracked = true;
  cudaError_t e = cudaStreamSynchronize(ctx->stream);
  cudaFree(d_count);
  if (e != cudaSuccess) return feasm_fail(ctx, FEASM_ERR_CUDA, cudaGetErrorString(e));
  ctx->elem_flag_valid = false;
  ctx->n_con = (int64_t)n;
  return (int64_t)n;
}

int feasm_get_mesh(feasm_ctx* ctx, double* xyz, uint32_t* conn, uint8_t* node_mask) {
  if (!ctx || !ctx->d_conn) return feasm_fail(ctx, FEASM_ERR_STATE, "no mesh");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  if (node_mask && !ctx->d_node_mask) return feasm_fail(ctx, FEASM_ERR_STATE, "boundary masks exist for generated meshes only");
  if (xyz) FEASM_CUDA(ctx, cudaMemcpyAsync(xyz, ctx->d_xyz, sizeof(double) * 3 * ctx->n_nodes, cudaMemcpyDeviceToHost, ctx->stream));
  if (conn) FEASM_CUDA(ctx, cudaMemcpyAsync(conn, ctx->d_conn, sizeof(uint32_t) * ctx->n_elem * ctx->npe, cudaMemcpyDeviceToHost, ctx->stream));
  if (node_mask) FEASM_CUDA(ctx, cudaMemcpyAsync(node_mask, ctx->d_node_mask, (size_t)ctx->n_nodes, cudaMemcpyDeviceToHost, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return FEASM_OK;
}

int feasm_get_dofs(feasm_ctx* ctx, uint32_t* elem_dofs, uint32_t* node_base) {
  if (!ctx || !ctx->d_elem_dofs) return feasm_fail(ctx, FEASM_ERR_STATE, "no dofs");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  if (node_base && !ctx->d_node_base) return feasm_fail(ctx, FEASM_ERR_STATE, "the node -> dof table exists after feasm_distribute_dofs_device only");
  if (elem_dofs) FEASM_CUDA(ctx, cudaMemcpyAsync(elem_dofs, ctx->d_elem_dofs, sizeof(uint32_t) * ctx->n_elem * ctx->nd, cudaMemcpyDeviceToHost, ctx->stream));
  if (node_base) FEASM_CUDA(ctx, cudaMemcpyAsync(node_base, ctx->d_node_base, sizeof(uint32_t) * ctx->n_nodes, cudaMemcpyDeviceToHost, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return FEASM_OK;
}

int feasm_get_constraints(feasm_ctx* ctx, uint8_t* flag, double* value) {
  if (!ctx || !ctx->d_con_flag) return feasm_fail(ctx, FEASM_ERR_STATE, "no dofs");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  if (flag) FEASM_CUDA(ctx, cudaMemcpyAsync(flag, ctx->d_con_flag, (size_t)ctx->n_rows, cudaMemcpyDeviceToHost, ctx->stream));
  if (value) FEASM_CUDA(ctx, cudaMemcpyAsync(value, ctx->d_con_val, sizeof(double) * ctx->n_rows, cudaMemcpyDeviceToHost, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return FEASM_OK;
}

}  // extern "C"
