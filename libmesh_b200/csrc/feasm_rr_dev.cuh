// feasm_rr_dev.cuh — device code of the row reduction (feasm_scatter.cu): shared-memory helpers and the per-batch
// routine one warp runs.
#ifndef FEASM_RR_DEV_CUH
#define FEASM_RR_DEV_CUH
#include "feasm_internal.h"

namespace {

// Shared memory is addressed through explicit 32-bit shared-window addresses: every access below is an immediate
// offset from a register, no generic-pointer conversion in the loops.
__device__ __forceinline__ void cp_async16(uint32_t smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(smem_dst), "l"(gsrc) : "memory");  // SASS LDGSTS, L1 bypass
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }
__device__ __forceinline__ double lds_f64(uint32_t a) { double v; asm volatile("ld.shared.f64 %0, [%1];\n" : "=d"(v) : "r"(a) : "memory"); return v; }
__device__ __forceinline__ void sts_f64(uint32_t a, double v) { asm volatile("st.shared.f64 [%0], %1;\n" ::"r"(a), "d"(v) : "memory"); }
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];\n" : "=r"(v) : "r"(a) : "memory"); return v; }
__device__ __forceinline__ void sts_u32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;\n" ::"r"(a), "r"(v) : "memory"); }
template <typename PosT>
__device__ __forceinline__ unsigned lds_pos(uint32_t a) {
  unsigned v;
  if (sizeof(PosT) == 1) asm volatile("ld.shared.u8 %0, [%1];\n" : "=r"(v) : "r"(a) : "memory");
  else asm volatile("ld.shared.u16 %0, [%1];\n" : "=r"(v) : "r"(a) : "memory");
  return v;
}

// ---- bulk-copy (TMA) ring fill: one cp.async.bulk per element record, completion counted in bytes by an mbarrier ----
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() {
  asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t smem_dst, const void* gsrc, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(smem_dst), "l"(gsrc),
               "r"(bytes), "r"(bar)
               : "memory");  // SASS UBLKCP
}

constexpr int RR_NG = 4;
// ring geometry: RR_NG stages of G contributors, ring about 4 KB per warp (occupancy beats deeper stages)
__host__ __device__ constexpr int rr_stage(int recb) {
  int g = 8;
  while (g > 2 && RR_NG * g * recb > 4096) --g;
  return g;
}
constexpr unsigned RR_LAST = 1u;  // flag entry of a record: (e,a) is the last contributor of its CSR row

// A warp owns a batch of up to 32 consecutive entries of the row list (or consecutive rows); lane l holds the
// pointers of row l of the batch.  The incidence ids of the whole batch are loaded into shared memory in one round
// trip; then the element records REC[e][a] (values + position row, recb bytes, feasm_internal.h) stream through a
// per-warp ring of RR_NG x G slots filled by cp.async (LDGSTS, 16 B per lane, no register staging): RR_NG-1 stages
// are in flight while one is accumulated into the CSR row image in shared memory.
//
// One contributor costs the warp a handful of instructions: lane b <= nd loads its position entry and its value
// (lane nd: the element's Fe, whose "position" is the image cell behind the longest row), adds the value into that
// image cell and stores it back.  The one thing the control flow needs — "this was the last contributor of the CSR
// row, write it out" — is static per (e,a) and rides in the record itself.  Position, value and flag of the next
// contributor are loaded before the current one is accumulated.  Dirichlet constraints were already applied to the
// records by the element kernel, so there is no other case.  Rows are finished strictly in order and every row is
// summed in element order by one warp, so the result is bitwise reproducible.
// ND = dofs per element when known at compile time (strides and lane loops fold away), 0 = generic (runtime nd).
// BULK: the ring is filled by bulk copies (one per record, issued by lane c for contributor c of the stage; needs G <= 32,
// RR_NG mbarriers at bar_a initialised to count 1, and `phase` = the caller's running parity bits, one per ring group);
// else by cp.async (LDGSTS).
template <typename PosT, int ND, bool BULK>
__device__ __forceinline__ void rr_process_batch(const int64_t base, const int nrows, const int lane, const uint32_t ring_a,
                                                 const uint32_t img_a, const uint32_t list_a, const int G, const int nd_rt,
                                                 const int fe_cell, const uint64_t* __restrict__ row_ptr,
                                                 const uint64_t* __restrict__ inc_ptr, const uint32_t* __restrict__ inc,
                                                 const unsigned char* __restrict__ rec, const uint32_t* __restrict__ row_list,
                                                 double* __restrict__ values, double* __restrict__ rhs, const uint32_t bar_a,
                                                 uint32_t& phase) {
  // the row image at img_a (cells 0..fe_cell) must be all zero on entry; it is all zero again on exit
  constexpr unsigned FULL = 0xffffffffu;
  constexpr int PB = (int)sizeof(PosT);
  const int nd = ND ? ND : nd_rt;
  const int ldrb = feasm_ldr(nd) * 8, recb = feasm_recb(nd, PB);
  constexpr int NDC = ND ? (ND + 2 + 31) / 32 : 5;  // lane passes over the nd+2 entries of one record (nd <= 128)
  const uint32_t fe_a = img_a + 8 * fe_cell;
  // ---- per-lane row descriptors ---------------------------------------------------------------------------
  uint32_t my_r = 0, my_ip0 = 0;  // incidence offsets fit 32 bits (n_elem*nd < 2^31 is checked at setup)
  uint64_t my_rp0 = 0;
  int my_cnt = 0, my_len = 0;
  if (lane < nrows) {
    my_r = row_list ? row_list[base + lane] : (uint32_t)(base + lane);
    const uint64_t i0 = inc_ptr[my_r], i1 = inc_ptr[my_r + 1], p0 = row_ptr[my_r], p1 = row_ptr[my_r + 1];
    my_ip0 = (uint32_t)i0; my_cnt = (int)(i1 - i0);
    my_rp0 = p0; my_len = (int)(p1 - p0);
    if (my_cnt == 0) {  // a row no local element touches: zero it here, it never shows up in the contributor stream
      for (int t = 0; t < my_len; ++t) values[p0 + t] = 0.;
      rhs[my_r] = 0.;
    }
  }
  uint32_t todo = __ballot_sync(FULL, my_cnt > 0);  // rows still to be written, lowest bit = current row
  // exclusive prefix of the contributor counts -> offset of row l inside the batch list
  int my_off = my_cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(FULL, my_off, o);
    if (lane >= o) my_off += v;
  }
  const int T = __shfl_sync(FULL, my_off, 31);  // <= list capacity by the host's choice of rows per batch
  my_off -= my_cnt;
  if (row_list == nullptr) {
    // consecutive rows: the batch's incidence ids are one contiguous run
    const uint32_t i0 = __shfl_sync(FULL, my_ip0, 0);
    for (int t = lane; t < T; t += 32) sts_u32(list_a + 4 * t, inc[i0 + t]);
  } else {
    for (int j = 0; j < nrows; ++j) {
      const int cj = __shfl_sync(FULL, my_cnt, j), oj = __shfl_sync(FULL, my_off, j);
      const uint32_t ij = __shfl_sync(FULL, my_ip0, j);
      for (int t = lane; t < cj; t += 32) sts_u32(list_a + 4 * (oj + t), inc[ij + t]);
    }
  }
  __syncwarp();

  // lane -> (contributor within an issue instruction, 16 B chunk of its record)
  const int ct = recb >> 4;
  const int per = ct <= 32 ? 32 / ct : 1;
  const int my_c = ct <= 32 ? lane / ct : 0, my_ch = ct <= 32 ? lane % ct : lane;
  const unsigned char* my_src = rec + (my_ch << 4);
  auto issue = [&](int k, int rg) {  // stage k: contributors [k*G, k*G+G) -> ring group rg
    const int t0 = k * G;
    if (BULK) {
      const int n = T - t0 < G ? T - t0 : G;
      if (n > 0) {
        if (lane == 0) mbar_expect_tx(bar_a + 8 * rg, (uint32_t)(n * recb));
        __syncwarp();
        if (lane < n)
          bulk_g2s(ring_a + (rg * G + lane) * recb, rec + (uint64_t)lds_u32(list_a + 4 * (t0 + lane)) * (uint32_t)recb, (uint32_t)recb,
                   bar_a + 8 * rg);
      }
      return;
    }
    const uint32_t sbase = ring_a + rg * G * recb + (my_ch << 4);
    for (int c0 = 0; c0 < G; c0 += per) {
      const int c = c0 + my_c, t = t0 + c;
      if (my_c < per && c < G && t < T) {
        const unsigned char* src = my_src + (uint64_t)lds_u32(list_a + 4 * t) * (uint32_t)recb;
        cp_async16(sbase + c * recb, src);
        if (ct > 32) {  // wide records: further 16 B chunks of the same slot
          for (int ch = 32; my_ch + ch < ct; ch += 32) cp_async16(sbase + c * recb + (ch << 4), src + (ch << 4));
        }
      }
    }
    cp_async_commit();
  };
  const int nstages = (T + G - 1) / G;
#pragma unroll
  for (int k = 0; k < RR_NG - 1; ++k) issue(k, k);  // empty groups past the end keep the group count uniform

  // what one contributor needs: position and value per lane pass, and the flag entry
  struct Slot { unsigned p[NDC]; double v[NDC]; };
  auto load = [&](uint32_t slot, Slot& s) {
#pragma unroll
    for (int bb = 0; bb < NDC; ++bb) {
      const int b = lane + 32 * bb;
      s.p[bb] = 0; s.v[bb] = 0.;
      if (b <= nd + 1) s.p[bb] = lds_pos<PosT>(slot + ldrb + b * PB);  // entry nd+1: the flag, read by its own lane
      if (b <= nd) s.v[bb] = lds_f64(slot + 8 * b);
    }
  };
  const int fl_lane = (nd + 1) & 31, fl_pass = (nd + 1) >> 5;
  int rg = 0;  // ring group of stage k
  for (int k = 0; k < nstages; ++k) {
    issue(k + RR_NG - 1, rg == 0 ? RR_NG - 1 : rg - 1);
    if (BULK) {
      mbar_wait(bar_a + 8 * rg, (phase >> rg) & 1u);
      phase ^= 1u << rg;
    } else {
      cp_async_wait<RR_NG - 1>();
    }
    __syncwarp();
    uint32_t slot = ring_a + rg * G * recb;
    const int cend = T - k * G < G ? T - k * G : G;
    Slot cur, nxt;
    load(slot, cur);
    for (int c = 0; c < cend; ++c) {
      slot += recb;
      load(slot, nxt);  // past the stage's last contributor this reads stale ring / image bytes that are never used
#pragma unroll
      for (int bb = 0; bb < NDC; ++bb) {
        if (lane + 32 * bb <= nd) {  // positions within one record are distinct
          FEASM_DEV_ASSERT((int)cur.p[bb] <= fe_cell);  // a position names a cell of the row image (Fe: the cell behind it)
          const uint32_t ia = img_a + (cur.p[bb] << 3);
          sts_f64(ia, lds_f64(ia) + cur.v[bb]);
        }
      }
      __syncwarp();
      unsigned flp = 0;
#pragma unroll
      for (int bb = 0; bb < NDC; ++bb) flp = bb == fl_pass ? cur.p[bb] : flp;
      if (__ballot_sync(FULL, lane == fl_lane && flp != 0)) {  // RR_LAST: write the finished row out and clear the image
        FEASM_DEV_ASSERT(todo != 0);  // a RR_LAST flag for every row of the batch, no more
        const int jl = __ffs(todo) - 1;
        todo &= todo - 1;
        const uint64_t rp0 = __shfl_sync(FULL, my_rp0, jl);
        const int len = __shfl_sync(FULL, my_len, jl);
        FEASM_DEV_ASSERT(len >= 0 && len <= fe_cell);
        const uint32_t r = row_list ? __shfl_sync(FULL, my_r, jl) : (uint32_t)(base + jl);
        double* dst = values + rp0 + lane;
        const uint32_t src = img_a + 8 * lane;
        if (sizeof(PosT) == 1) {  // rows of at most 255 entries: eight predicated steps, immediate offsets
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            if (32 * i >= len) break;
            if (lane + 32 * i < len) {
              __stcs(dst + 32 * i, lds_f64(src + 256 * i));
              sts_f64(src + 256 * i, 0.);
            }
          }
        } else {
          for (int t = lane; t < len; t += 32) {
            __stcs(dst + (t - lane), lds_f64(img_a + 8 * t));
            sts_f64(img_a + 8 * t, 0.);
          }
        }
        if (lane == 0) {
          rhs[r] = lds_f64(fe_a);
          sts_f64(fe_a, 0.);
        }
        __syncwarp();
      }
      cur = nxt;
    }
    __syncwarp();  // the slots of this stage are refilled by the next issue
    rg = rg == RR_NG - 1 ? 0 : rg + 1;
  }
  if (!BULK) cp_async_wait<0>();
  __syncwarp();
}

}  // namespace
#endif
