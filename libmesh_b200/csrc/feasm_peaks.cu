// feasm_peaks.cu — live FP64 roofline denominators (instrumentation, not on the assembly path).
// MEASURED_PEAKS.json carries HBM bandwidth and bf16 GEMM throughput only; the assembly kernels run on the FP64
// pipe, so bench.py measures that pipe on the same box, in the same process: a register-resident DFMA chain and a
// DMMA (mma.sync.m8n8k4.f64) chain, 8 independent accumulators per thread, 2 CTAs of 256 threads per SM.
#include "feasm_internal.h"

namespace {

__global__ void __launch_bounds__(256) k_peak_dfma(double* out, int iters, double a, double b) {
  double r[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) r[i] = threadIdx.x * 1e-3 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int i = 0; i < 8; ++i) r[i] = fma(r[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += r[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) k_peak_dmma(double* out, int iters, double a, double b) {
  double c[8][2];
#pragma unroll
  for (int i = 0; i < 8; ++i) { c[i][0] = threadIdx.x * 1e-3; c[i][1] = i; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int i = 0; i < 8; ++i)
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                     : "+d"(c[i][0]), "+d"(c[i][1])
                     : "d"(a), "d"(b));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

}  // namespace

extern "C" int feasm_measure_fp64_peaks(feasm_ctx* ctx, double* dfma_tflops, double* dmma_tflops) {
  if (!ctx) return FEASM_ERR_ARG;
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  const int blocks = ctx->sm_count * 2, threads = 256, iters = 4096;
  double* d_out = nullptr;
  FEASM_CUDA(ctx, cudaMalloc(&d_out, sizeof(double) * blocks * threads));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  double best[2] = {0, 0};
  for (int which = 0; which < 2; ++which)
    for (int rep = 0; rep < 4; ++rep) {  // rep 0 warms up
      cudaEventRecord(e0, ctx->stream);
      if (which == 0) k_peak_dfma<<<blocks, threads, 0, ctx->stream>>>(d_out, iters, 1.0000001, 1e-9);
      else k_peak_dmma<<<blocks, threads, 0, ctx->stream>>>(d_out, iters, 1.0000001, 1e-9);
      cudaEventRecord(e1, ctx->stream);
      FEASM_CUDA(ctx, cudaEventSynchronize(e1));
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      // flops: DFMA = 2 per thread-op; DMMA m8n8k4 = 2*8*8*4 per warp-op
      const double ops = (double)iters * 32.0;
      const double flops = which == 0 ? ops * 2.0 * blocks * threads : ops * 512.0 * blocks * (threads / 32);
      if (rep > 0) best[which] = std::max(best[which], flops / (ms * 1e-3) / 1e12);
    }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d_out);
  if (dfma_tflops) *dfma_tflops = best[0];
  if (dmma_tflops) *dmma_tflops = best[1];
  return FEASM_OK;
}
