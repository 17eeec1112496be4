// FP64 peak microbenchmarks for B200 (sm_100a): DFMA chain, DMMA (mma.sync f64) shapes,
// mixed DFMA+DMMA, shared-memory LDS.64 bandwidth and a global copy.  The numbers printed here
// are the FP64 denominators of the roofline in bench.py (MEASURED_PEAKS.json has no FP64 entry).
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -o build/fp64_peaks fp64_peaks.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__global__ void __launch_bounds__(256) k_dfma(double* out, int iters, double a, double b) {
  double r[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) r[i] = threadIdx.x * 1e-3 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
#pragma unroll
      for (int i = 0; i < 8; ++i) r[i] = fma(r[i], a, b);
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += r[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma1688(double (&c)[4], const double (&a)[4], const double (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
}
__device__ __forceinline__ void dmma16816(double (&c)[4], const double (&a)[8], const double (&b)[4]) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};\n"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                 "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
}

template <int NACC>
__global__ void __launch_bounds__(256) k_dmma884(double* out, int iters) {
  double c[NACC][2];
#pragma unroll
  for (int i = 0; i < NACC; ++i) { c[i][0] = 0; c[i][1] = 0; }
  double a = threadIdx.x * 1e-6, b = 1.0 + threadIdx.x * 1e-7;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) dmma884(c[i][0], c[i][1], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NACC>
__global__ void __launch_bounds__(256) k_dmma1688(double* out, int iters) {
  double c[NACC][4];
#pragma unroll
  for (int i = 0; i < NACC; ++i) for (int j = 0; j < 4; ++j) c[i][j] = 0;
  double a[4], b[2];
  for (int j = 0; j < 4; ++j) a[j] = threadIdx.x * 1e-6 + j;
  for (int j = 0; j < 2; ++j) b[j] = 1.0 + threadIdx.x * 1e-7 + j;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) dmma1688(c[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) for (int j = 0; j < 4; ++j) s += c[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NACC>
__global__ void __launch_bounds__(256) k_dmma16816(double* out, int iters) {
  double c[NACC][4];
#pragma unroll
  for (int i = 0; i < NACC; ++i) for (int j = 0; j < 4; ++j) c[i][j] = 0;
  double a[8], b[4];
  for (int j = 0; j < 8; ++j) a[j] = threadIdx.x * 1e-6 + j;
  for (int j = 0; j < 4; ++j) b[j] = 1.0 + threadIdx.x * 1e-7 + j;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) dmma16816(c[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) for (int j = 0; j < 4; ++j) s += c[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// Mixed: each warp interleaves DMMA (8 accumulators) with NF independent DFMA per DMMA.
template <int NF>
__global__ void __launch_bounds__(256) k_mixed(double* out, int iters, double fa, double fb) {
  double c[8][2];
  double r[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { c[i][0] = 0; c[i][1] = 0; r[i] = threadIdx.x + i; }
  double a = threadIdx.x * 1e-6, b = 1.0 + threadIdx.x * 1e-7;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      dmma884(c[i][0], c[i][1], a, b);
#pragma unroll
      for (int f = 0; f < NF; ++f) r[(i + f) & 7] = fma(r[(i + f) & 7], fa, fb);
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1] + r[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// Shared memory LDS.64 bandwidth: conflict-free distinct addresses.
__global__ void __launch_bounds__(256) k_lds64(double* out, int iters) {
  __shared__ double sm[4096];
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = i;
  __syncthreads();
  double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
  int idx = threadIdx.x;
  for (int it = 0; it < iters; ++it) {
    s0 += sm[(idx) & 4095];
    s1 += sm[(idx + 256) & 4095];
    s2 += sm[(idx + 512) & 4095];
    s3 += sm[(idx + 768) & 4095];
    idx += 1024;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s0 + s1 + s2 + s3;
}

__global__ void k_copy(const double2* __restrict__ in, double2* __restrict__ out, size_t n) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) out[i] = in[i];
}

template <typename F>
float time_ms(F f, int reps = 5) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  f(); f();
  CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < reps; ++r) {
    CK(cudaEventRecord(e0));
    f();
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (ms < best) best = ms;
  }
  return best;
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  int nsm = p.multiProcessorCount;
  printf("{\"device\": \"%s\", \"sms\": %d, \"clock_khz\": %d}\n", p.name, nsm, p.clockRate);
  double* out; CK(cudaMalloc(&out, sizeof(double) * nsm * 8 * 256));
  const int iters = 20000;
  for (int bps : {1, 2, 4, 8}) {
    int grid = nsm * bps;
    float ms = time_ms([&] { k_dfma<<<grid, 256>>>(out, iters, 1.0000001, 1e-9); });
    double flops = 2.0 * grid * 256.0 * iters * 32;
    printf("{\"bench\": \"dfma\", \"blocks_per_sm\": %d, \"ms\": %.3f, \"tflops\": %.2f}\n", bps, ms, flops / ms * 1e-9);
  }
  for (int bps : {1, 2, 4}) {
    int grid = nsm * bps;
    float ms = time_ms([&] { k_dmma884<8><<<grid, 256>>>(out, iters); });
    double flops = 2.0 * 256 * (double)grid * 8 * iters * 8;
    printf("{\"bench\": \"dmma_m8n8k4_acc8\", \"blocks_per_sm\": %d, \"ms\": %.3f, \"tflops\": %.2f}\n", bps, ms, flops / ms * 1e-9);
    ms = time_ms([&] { k_dmma884<16><<<grid, 256>>>(out, iters); });
    flops = 2.0 * 256 * (double)grid * 8 * iters * 16;
    printf("{\"bench\": \"dmma_m8n8k4_acc16\", \"blocks_per_sm\": %d, \"ms\": %.3f, \"tflops\": %.2f}\n", bps, ms, flops / ms * 1e-9);
    ms = time_ms([&] { k_dmma1688<8><<<grid, 256>>>(out, iters); });
    flops = 2.0 * 1024 * (double)grid * 8 * iters * 8;
    printf("{\"bench\": \"dmma_m16n8k8_acc8\", \"blocks_per_sm\": %d, \"ms\": %.3f, \"tflops\": %.2f}\n", bps, ms, flops / ms * 1e-9);
    ms = time_ms([&] { k_dmma16816<8><<<grid, 256>>>(out, iters); });
    flops = 2.0 * 2048 * (double)grid * 8 * iters * 8;
    printf("{\"bench\": \"dmma_m16n8k16_acc8\", \"blocks_per_sm\": %d, \"ms\": %.3f, \"tflops\": %.2f}\n", bps, ms, flops / ms * 1e-9);
  }
  {
    int grid = nsm * 2;
    float ms = time_ms([&] { k_mixed<1><<<grid, 256>>>(out, iters, 1.0000001, 1e-9); });
    double mm = 2.0 * 256 * (double)grid * 8 * iters * 8, ff = 2.0 * grid * 256.0 * iters * 8 * 1;
    printf("{\"bench\": \"mixed_dmma+1dfma\", \"ms\": %.3f, \"dmma_tflops\": %.2f, \"dfma_tflops\": %.2f}\n", ms, mm / ms * 1e-9, ff / ms * 1e-9);
    ms = time_ms([&] { k_mixed<4><<<grid, 256>>>(out, iters, 1.0000001, 1e-9); });
    ff = 2.0 * grid * 256.0 * iters * 8 * 4;
    printf("{\"bench\": \"mixed_dmma+4dfma\", \"ms\": %.3f, \"dmma_tflops\": %.2f, \"dfma_tflops\": %.2f}\n", ms, mm / ms * 1e-9, ff / ms * 1e-9);
    ms = time_ms([&] { k_mixed<8><<<grid, 256>>>(out, iters, 1.0000001, 1e-9); });
    ff = 2.0 * grid * 256.0 * iters * 8 * 8;
    printf("{\"bench\": \"mixed_dmma+8dfma\", \"ms\": %.3f, \"dmma_tflops\": %.2f, \"dfma_tflops\": %.2f}\n", ms, mm / ms * 1e-9, ff / ms * 1e-9);
  }
  {
    int grid = nsm * 4;
    float ms = time_ms([&] { k_lds64<<<grid, 256>>>(out, iters); });
    double bytes = (double)grid * 256 * iters * 4 * 8;
    printf("{\"bench\": \"lds64\", \"ms\": %.3f, \"bytes_per_clk_per_sm_at_1965\": %.1f, \"TBps\": %.2f}\n", ms,
           bytes / (ms * 1e-3) / nsm / 1.965e9, bytes / ms * 1e-9);
  }
  {
    size_t n = (size_t)1 << 27;  // 2 GiB of double2 per buffer
    double2 *a, *b; CK(cudaMalloc(&a, n * 16)); CK(cudaMalloc(&b, n * 16));
    CK(cudaMemset(a, 1, n * 16));
    float ms = time_ms([&] { k_copy<<<nsm * 16, 512>>>(a, b, n); });
    printf("{\"bench\": \"copy\", \"ms\": %.3f, \"GBps_rw\": %.1f}\n", ms, 2.0 * n * 16 / ms * 1e-6);
    ms = time_ms([&] { CK(cudaMemsetAsync(b, 0, n * 16)); });
    printf("{\"bench\": \"memset\", \"ms\": %.3f, \"GBps_w\": %.1f}\n", ms, 1.0 * n * 16 / ms * 1e-6);
  }
  return 0;
}
