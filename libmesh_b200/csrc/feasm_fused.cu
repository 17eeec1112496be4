// feasm_fused.cu — element kernel and row reduction as ONE persistent kernel (Poisson).
//
// The element phase is bound by the FP64 pipe, the row reduction by HBM and issue slots; run back to back they leave
// each other's resource idle.  Here every CTA (one per SM, cooperative launch: all CTAs are resident) has EW element
// warps and RW reduction warps:
//   element warps   walk the elements in waves of gridDim.x*EW consecutive ids (wave k, CTA b, warp w -> element
//                   (k*gridDim.x + b)*EW + w), write their KF block, then publish: __threadfence + atomicAdd(done[k]).
//   reduction warps take batches of 32 consecutive CSR rows from a global ticket counter, wait until every wave that
//                   holds a contributor of the batch is complete (batch_need[b] = 1 + the largest contributing element
//                   id, computed at setup; waves complete almost in order, so a warp only ever polls its next
//                   unverified wave), then run the same cp.async-fed reduction as feasm_scatter.cu.
// Element warps never wait, so there is no deadlock; a row's sum is still taken in element order by one warp, so the
// result is bitwise identical to the two-kernel path.  KF rows are consumed a few waves (tens of MB) after they were
// written, i.e. mostly out of L2.
#include <cstdlib>

#include "feasm_elem_dev.cuh"
#include "feasm_rr_dev.cuh"

namespace {

__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}

constexpr int FUSED_RR_G = 12;  // contributors per ring stage of a reduction warp (3 stages = 36 element rows in flight)

template <int NPE, int NQ, typename PosT, int EW, int RW, int ER, int RR>
__global__ void __launch_bounds__((EW + RW) * 32, 1)
k_fused_poisson(int diag, const double* __restrict__ xyz, const uint32_t* __restrict__ conn, const double* __restrict__ tab, int forcing,
                const double* __restrict__ f_tab, int64_t n_elem, double* __restrict__ kf, int* flags,
                int rs, ElemConArgs con, const uint64_t* __restrict__ row_ptr, const uint64_t* __restrict__ inc_ptr,
                const uint32_t* __restrict__ inc, int64_t n_rows, int cap, int fe_cell, double* __restrict__ values, double* __restrict__ rhs, const uint32_t* __restrict__ batch_need,
                uint32_t* done /* [n_waves] */, uint32_t* ticket) {
  using C = ElemCfg<NPE, NQ>;
  extern __shared__ __align__(16) double smem[];
  double* s_tab = smem;
  double* s_tw = smem + C::TAB_SZ;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  elem_load_tables<NPE, NQ>(s_tab, s_tw, tab, threadIdx.x, (EW + RW) * 32);
  __syncthreads();
  const int64_t W = (int64_t)gridDim.x * EW;  // elements per wave

  // the launch gives every warp the same register count; the reduction warpgroups hand theirs to the element warpgroups
  if (warp < EW) {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;\n" ::"n"(ER));
    // ---- element role --------------------------------------------------------------------------------------------
    double* X = s_tw + C::TW_SZ + (size_t)warp * C::WARP_DBL;
    double* fw = X + C::X_SZ;
    const double* twp = s_tw + (lane >> 2) * C::TS + (lane & 3);
    int64_t e = (int64_t)blockIdx.x * EW + warp;
    double nx = 0., ny = 0., nz = 0.;
    uint32_t nnode = 0;
    if (e < n_elem && lane < NPE) {
      const uint64_t node = conn[e * NPE + lane];
      nx = xyz[node * 3]; ny = xyz[node * 3 + 1]; nz = xyz[node * 3 + 2];
      if (e + W < n_elem) nnode = conn[(e + W) * NPE + lane];
    }
    for (int64_t k = 0; e < n_elem; e += W, ++k) {
      if (lane < NPE) {
        X[3 * lane] = nx; X[3 * lane + 1] = ny; X[3 * lane + 2] = nz;
        if (e + W < n_elem) {
          const uint64_t node = nnode;
          nx = xyz[node * 3]; ny = xyz[node * 3 + 1]; nz = xyz[node * 3 + 2];
          if (e + 2 * W < n_elem) nnode = conn[(e + 2 * W) * NPE + lane];
        }
      }
      __syncwarp();
      if (diag != 2)
      {
        double* out = kf + (uint64_t)e * (NPE * rs);
        elem_process<NPE, NQ>(X, fw, s_tab, twp, lane, forcing, f_tab ? f_tab + e * NQ : nullptr, flags, out, rs);
        if (con.elem_flag && con.elem_flag[e])
          constrain_element(out, rs, NPE, con.elem_dofs + (uint64_t)e * NPE, con.con_flag, con.con_val, con.mode, lane);
      }
      __threadfence();  // every lane: its KF stores are visible device-wide before the wave counter moves
      __syncwarp();
      if (lane == 0) atomicAdd(&done[k], 1u);
    }
  } else {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" ::"n"(RR));
    if (diag == 1) return;
    // ---- reduction role --------------------------------------------------------------------------------------------
    constexpr int slotb = feasm_recb(NPE, (int)sizeof(PosT));
    constexpr int G = FUSED_RR_G, LCAP = 256;
    const uint32_t warp_bytes = (uint32_t)(RR_NG * G * slotb + cap * 8 + LCAP * 4 + 8 * RR_NG);
    const uint32_t rr0 = (uint32_t)__cvta_generic_to_shared(smem) + (uint32_t)((C::TAB_SZ + C::TW_SZ + EW * C::WARP_DBL) * 8);
    const uint32_t ring_a = rr0 + (warp - EW) * warp_bytes;
    const uint32_t img_a = ring_a + RR_NG * G * slotb;
    const uint32_t list_a = img_a + cap * 8;
    const uint32_t bar_a = list_a + LCAP * 4;
    uint32_t phase = 0;
    for (int t = lane; t < cap; t += 32) sts_f64(img_a + 8 * t, 0.);
    if (lane < RR_NG) mbar_init(bar_a + 8 * lane, 1);
    mbar_init_fence();
    __syncwarp();
    const int64_t n_batches = (n_rows + 31) / 32;
    const int64_t n_waves = (n_elem + W - 1) / W;
    int64_t verified = 0;  // waves [0, verified) are known complete
    for (;;) {
      uint32_t b = 0;
      if (lane == 0) b = atomicAdd(ticket, 1u);
      b = __shfl_sync(0xffffffffu, b, 0);
      if ((int64_t)b >= n_batches) break;
      if (lane == 0) {
        const int64_t need = batch_need[b];           // elements [0, need) must be done
        const int64_t kn = (need + W - 1) / W;        // = waves [0, kn)
        while (verified < kn) {
          const int64_t lo = verified * W;
          const uint32_t full = (uint32_t)(n_elem - lo < W ? n_elem - lo : W);
          while (ld_acquire_u32(done + verified) < full) __nanosleep(200);
          ++verified;
        }
      }
      verified = __shfl_sync(0xffffffffu, verified, 0);
      (void)n_waves;
      const int64_t base = (int64_t)b * 32;
      const int nrows = (int)(n_rows - base < 32 ? n_rows - base : 32);
      rr_process_batch<PosT, NPE, true>(base, nrows, lane, ring_a, img_a, list_a, G, NPE, fe_cell, row_ptr, inc_ptr, inc,
                                        reinterpret_cast<const unsigned char*>(kf), nullptr, values, rhs, bar_a, phase);
    }
  }
}

template <int NPE, int NQ, typename PosT, int EW, int RW, int ER, int RR>
int launch_fused(feasm_ctx* ctx, const feasm_params& p, const double* d_f) {
  using C = ElemCfg<NPE, NQ>;
  static_assert(EW % 4 == 0 && RW % 4 == 0, "setmaxnreg works on whole warpgroups");
  constexpr int slotb = feasm_recb(NPE, (int)sizeof(PosT));
  int cap = ((int)ctx->max_row_len + 2) & ~1, fe_cell = (int)ctx->max_row_len;
  const size_t smem = sizeof(double) * (size_t)(C::TAB_SZ + C::TW_SZ + EW * C::WARP_DBL) +
                      (size_t)RW * (RR_NG * FUSED_RR_G * slotb + cap * 8 + 256 * 4 + 8 * RR_NG);
  if (smem > ctx->smem_optin || ctx->max_inc * 32u > 256u) return 1;  // not eligible: caller falls back to two kernels
  auto kern = k_fused_poisson<NPE, NQ, PosT, EW, RW, ER, RR>;
  static const int diag_env = getenv("FEASM_FUSED_DIAG") ? atoi(getenv("FEASM_FUSED_DIAG")) : 0;
  int diag = diag_env;
  FEASM_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int grid = ctx->sm_count;
  const int64_t W = (int64_t)grid * EW, n_waves = (ctx->n_elem + W - 1) / W;
  if ((size_t)n_waves + 1 > ctx->sync_cap) {
    if (ctx->d_sync) cudaFree(ctx->d_sync);
    ctx->d_sync = nullptr;
    FEASM_CUDA(ctx, cudaMalloc(&ctx->d_sync, sizeof(uint32_t) * (n_waves + 1)));
    ctx->sync_cap = n_waves + 1;
  }
  FEASM_CUDA(ctx, cudaMemsetAsync(ctx->d_sync, 0, sizeof(uint32_t) * (n_waves + 1), ctx->stream));
  const double* xyz = ctx->d_xyz; const uint32_t* conn = ctx->d_conn; const double* tab = ctx->d_tab;
  int forcing = p.forcing; int64_t n_elem = ctx->n_elem; double* kf = ctx->d_kf; int* flags = ctx->d_flags;
  int rs = slotb / 8;
  ElemConArgs con{ctx->n_con > 0 ? ctx->d_elem_flag : nullptr, ctx->d_elem_dofs, ctx->d_con_flag, ctx->d_con_val, p.constrain};
  const uint64_t* row_ptr = ctx->d_row_ptr; const uint64_t* inc_ptr = ctx->d_inc_ptr; const uint32_t* inc = ctx->d_inc;
  int64_t n_rows = ctx->n_rows; double* values = ctx->d_values; double* rhs = ctx->d_rhs; const uint32_t* need = ctx->d_batch_need;
  uint32_t* done = ctx->d_sync + 1; uint32_t* ticket = ctx->d_sync;
  void* args[] = {&diag, &xyz, &conn, &tab, &forcing, &d_f, &n_elem, &kf, &flags, &rs, &con, &row_ptr, &inc_ptr, &inc, &n_rows, &cap,
                  &fe_cell, &values, &rhs, &need, &done, &ticket};
  // cooperative launch: fails instead of deadlocking if the grid cannot be co-resident
  FEASM_CUDA(ctx, cudaLaunchCooperativeKernel((void*)kern, dim3(grid), dim3((EW + RW) * 32), args, smem, ctx->stream));
  return FEASM_OK;
}

}  // namespace

// returns FEASM_OK when the fused kernel was launched, 1 when this configuration has no fused kernel (the caller runs
// the two-kernel path), a negative error code otherwise
int feasm_launch_fused(feasm_ctx* ctx, const feasm_params& p, const double* d_f) {
  if (p.problem != FEASM_POISSON || !ctx->d_batch_need) return 1;
  if (ctx->elem_type == FEASM_HEX27 && ctx->tab_nq == 27 && ctx->nd == 27 && ctx->pos_bytes == 1)
  {
    static const int cfg = getenv("FEASM_FUSED") ? atoi(getenv("FEASM_FUSED")) : 1;
    if (cfg == 2) return launch_fused<27, 27, uint8_t, 12, 4, 144, 64>(ctx, p, d_f);
    if (cfg == 3) return launch_fused<27, 27, uint8_t, 8, 12, 152, 56>(ctx, p, d_f);
    return launch_fused<27, 27, uint8_t, 8, 8, 184, 64>(ctx, p, d_f);
  }
  return 1;
}
