// feasm_rr_dev.cuh — device code of the row reduction shared by feasm_scatter.cu (stand-alone kernel) and
// feasm_fused.cu (element + row-reduction warps in one persistent kernel).  See feasm_scatter.cu for the design.
#ifndef FEASM_RR_DEV_CUH
#define FEASM_RR_DEV_CUH
#include "feasm_internal.h"

namespace {

template <typename PosT> struct PosBits;
template <> struct PosBits<uint8_t> { static constexpr unsigned FLAG = 0x80u, MASK = 0x7fu; };
template <> struct PosBits<uint16_t> { static constexpr unsigned FLAG = 0x8000u, MASK = 0x7fffu; };

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

constexpr int RR_NG = 4;
// ring geometry: about 12 contributors in flight per warp, ring <= 8 KB
__host__ __device__ constexpr int rr_stage(int slotb) {
  int g = 4;
  while (g > 2 && RR_NG * g * slotb > 8192) --g;
  return g;
}

// The uncommon contributor: its CSR row is a Dirichlet row, or some of its columns are.  Kept out of line so the
// common path stays short (the compiler would otherwise predicate all of this into it).
template <typename PosT>
__device__ __noinline__ double2 rr_constrained_contributor(uint32_t slot, uint32_t img_a, int lane, int nd, int ldrb, int row_con,
                                                            double row_g, uint32_t ea, int mode, const double* __restrict__ con_val,
                                                            const uint32_t* __restrict__ elem_dofs) {
  // returns (addend of the row's RHS sum, addend of this lane's K(a,b) g_b correction)
  if (row_con) {
    // this element contributes K(a,a) = 1, F(a) = g
    const int a = (int)(ea % (uint32_t)nd);
    if (lane == 0) {
      const uint32_t ia = img_a + ((lds_pos<PosT>(slot + ldrb + a * (int)sizeof(PosT)) & PosBits<PosT>::MASK) << 3);
      sts_f64(ia, lds_f64(ia) + 1.);
    }
    return make_double2(row_g, 0.);
  }
  double corr = 0.;
  const uint64_t e = ea / (uint32_t)nd;
  for (int b = lane; b < nd; b += 32) {
    double val = lds_f64(slot + 8 * b);
    const unsigned pp = lds_pos<PosT>(slot + ldrb + b * (int)sizeof(PosT));
    if (pp & PosBits<PosT>::FLAG) {
      if (mode == FEASM_CONSTRAIN_HETEROGENEOUS) corr += val * con_val[elem_dofs[e * nd + b]];
      val = 0.;
    }
    const uint32_t ia = img_a + ((pp & PosBits<PosT>::MASK) << 3);
    sts_f64(ia, lds_f64(ia) + val);
  }
  return make_double2(lds_f64(slot + 8 * nd), corr);
}

// A warp owns a batch of RB consecutive entries of the row list (or RB consecutive rows); lane l holds the pointers
// of row l of the batch.  The incidence ids of the whole batch are loaded into shared memory in one round trip;
// then the element rows KF[e][a][:] (ldrb bytes) and their position-map rows (pldb bytes) stream through a
// per-warp ring of NG x G slots filled by cp.async (LDGSTS, 16 B per lane, no register staging): NG-1 stages —
// several KB per warp — are in flight while one is accumulated into the CSR row image in shared memory.  Rows are
// finished strictly in order and every row is summed in element order, so the result is bitwise reproducible.
// ND = dofs per element when known at compile time (strides and lane loops fold away), 0 = generic (runtime nd).
template <typename PosT, int ND>
__device__ __forceinline__ void rr_process_batch(const int64_t base, const int nrows, const int lane, const uint32_t ring_a,
                                                 const uint32_t img_a, const uint32_t list_a, const int G, const int nd_rt,
                                                 const uint64_t* __restrict__ row_ptr, const uint64_t* __restrict__ inc_ptr,
                                                 const uint32_t* __restrict__ inc, const unsigned char* __restrict__ pos,
                                                 const unsigned char* __restrict__ kf, const uint32_t* __restrict__ row_list,
                                                 const uint8_t* __restrict__ con_flag, const double* __restrict__ con_val,
                                                 const uint32_t* __restrict__ elem_dofs, const int mode,
                                                 double* __restrict__ values, double* __restrict__ rhs) {
  // the row image at img_a must be all zero on entry; it is all zero again on exit
  constexpr unsigned FULL = 0xffffffffu;
  const int nd = ND ? ND : nd_rt;
  const int ldrb = feasm_ldr(nd) * 8, pldb = feasm_pld(nd, (int)sizeof(PosT)) * (int)sizeof(PosT), slotb = ldrb + pldb;
  constexpr int NDC = ND ? (ND + 31) / 32 : 4;  // lane passes over one element row
  // ---- per-lane row descriptors ---------------------------------------------------------------------------
  uint32_t my_r = 0, my_ip0 = 0;  // incidence offsets fit 32 bits (n_elem*nd < 2^32 is checked at setup)
  uint64_t my_rp0 = 0;
  int my_cnt = 0, my_len = 0, my_cf = 0;
  double my_g = 0.;
  if (lane < nrows) {
    my_r = row_list ? row_list[base + lane] : (uint32_t)(base + lane);
    const uint64_t i0 = inc_ptr[my_r], i1 = inc_ptr[my_r + 1], p0 = row_ptr[my_r], p1 = row_ptr[my_r + 1];
    my_ip0 = (uint32_t)i0; my_cnt = (int)(i1 - i0);
    my_rp0 = p0; my_len = (int)(p1 - p0);
    my_cf = con_flag[my_r];
    if (my_cf && mode == FEASM_CONSTRAIN_HETEROGENEOUS) my_g = con_val[my_r];
  }
  // exclusive prefix of the contributor counts -> offset of row l inside the batch list
  int my_off = my_cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(FULL, my_off, o);
    if (lane >= o) my_off += v;
  }
  const int T = __shfl_sync(FULL, my_off, 31);  // <= lcap by the host's choice of RB
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

  // lane -> (contributor within an issue instruction, 16 B chunk of its slot); a lane always copies the same chunk
  // of a slot, so its source is one base pointer plus (incidence id) * (row stride of the array it reads)
  const int ct = slotb >> 4, ck = ldrb >> 4;
  const int per = ct <= 32 ? 32 / ct : 1;
  const int my_c = ct <= 32 ? lane / ct : 0, my_ch = ct <= 32 ? lane % ct : lane;
  const unsigned char* my_src = my_ch < ck ? kf + (my_ch << 4) : pos + ((my_ch - ck) << 4);
  const uint32_t my_stride = my_ch < ck ? ldrb : pldb;
  auto issue = [&](int k, int rg) {  // stage k: contributors [k*G, k*G+G) -> ring group rg
    const int t0 = k * G;
    const uint32_t sbase = ring_a + rg * G * slotb + (my_ch << 4);
#pragma unroll
    for (int c0 = 0; c0 < G; c0 += per) {
      const int c = c0 + my_c, t = t0 + c;
      if (my_c < per && c < G && t < T) {
        const uint64_t ea = lds_u32(list_a + 4 * t) & FEASM_INC_MASK;
        cp_async16(sbase + c * slotb, my_src + ea * my_stride);
        if (ct > 32) {  // wide element rows: further 16 B chunks of the same slot
          for (int ch = my_ch + 32; ch < ct; ch += 32)
            cp_async16(sbase + c * slotb + ((ch - my_ch) << 4),
                       ch < ck ? kf + ea * ldrb + (ch << 4) : pos + ea * pldb + ((ch - ck) << 4));
        }
      }
    }
    cp_async_commit();
  };
  const int nstages = (T + G - 1) / G;
#pragma unroll
  for (int k = 0; k < RR_NG - 1; ++k) issue(k, k);  // empty groups past the end keep the group count uniform

  int j = 0, left = 0, row_con = 0;
  double row_g = 0., bsum = 0., corr = 0.;
  auto flush = [&]() {  // write row j out and clear the image
    const uint64_t rp0 = __shfl_sync(FULL, my_rp0, j);
    const int len = __shfl_sync(FULL, my_len, j);
    const uint32_t r = __shfl_sync(FULL, my_r, j);
    if (__any_sync(FULL, corr != 0.)) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) corr += __shfl_xor_sync(FULL, corr, o);
      bsum -= corr;
    }
    double* dst = values + rp0;
    for (int t = lane; t < len; t += 32) {
      __stcs(dst + t, lds_f64(img_a + 8 * t));
      sts_f64(img_a + 8 * t, 0.);
    }
    if (lane == 0) rhs[r] = bsum;
    bsum = 0.; corr = 0.;
    __syncwarp();
  };
  auto next_row = [&]() {  // advance to the next row that has contributors, flushing empty rows on the way
    for (;;) {
      left = __shfl_sync(FULL, my_cnt, j);
      row_con = __shfl_sync(FULL, my_cf, j);
      row_g = __shfl_sync(FULL, my_g, j);
      if (left > 0) return;
      flush();
      if (++j >= nrows) return;
    }
  };
  next_row();
  int rg = 0;  // ring group of stage k
  for (int k = 0; k < nstages; ++k) {
    issue(k + RR_NG - 1, rg == 0 ? RR_NG - 1 : rg - 1);
    cp_async_wait<RR_NG - 1>();
    __syncwarp();
    const uint32_t sbase = ring_a + rg * G * slotb;
    const int cend = T - k * G < G ? T - k * G : G;
#pragma unroll
    for (int c = 0; c < G; ++c) {
      if (c >= cend) break;
      const uint32_t slot = sbase + c * slotb;
      // bit 31 of the incidence id: this element touches a constrained dof (set by feasm_launch_pos_flags)
      const uint32_t ea = lds_u32(list_a + 4 * (k * G + c));
      if (ea & FEASM_INC_SLOW) {
        const double2 add = rr_constrained_contributor<PosT>(slot, img_a, lane, nd, ldrb, row_con, row_g, ea & FEASM_INC_MASK,
                                                             mode, con_val, elem_dofs);
        bsum += add.x; corr += add.y;
      } else {
#pragma unroll
        for (int bb = 0; bb < NDC; ++bb) {
          const int b = lane + 32 * bb;
          if (b < nd) {  // positions within one element row are distinct
            const uint32_t ia = img_a + (lds_pos<PosT>(slot + ldrb + b * (int)sizeof(PosT)) << 3);
            const double val = lds_f64(slot + 8 * b);
            sts_f64(ia, lds_f64(ia) + val);
          }
        }
        bsum += lds_f64(slot + 8 * nd);
      }
      __syncwarp();
      if (--left == 0) {
        flush();
        if (++j < nrows) next_row();
      }
    }
    __syncwarp();  // the slots of this stage are refilled by the next issue
    rg = rg == RR_NG - 1 ? 0 : rg + 1;
  }
  cp_async_wait<0>();
  __syncwarp();
}

}  // namespace
#endif
