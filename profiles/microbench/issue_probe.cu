// Does a non-FP64 instruction issue in the shadow of an FP64 instruction on sm_100a?  The element kernel's time model
// depends on it: a warp-wide DFMA keeps the 16-lane FP64 unit of a sub-partition busy for 2 cycles and a
// DMMA.8x8x4 for 16; this probe interleaves K independent integer (or FP32) instructions per FP64 instruction and prints
// cycles per FP64 instruction per sub-partition.  Flat in K = the other work hides; rising by K = it is added time.
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o build/issue_probe profiles/microbench/issue_probe.cu
// Result on B200 (profiles/r2/issue_probe.md): DFMA 2.07 cycles alone, 2.70 with one LOP3, 4.18 / 6.20 / 8.18 with 2 / 3 / 4;
// DMMA 16.06 alone, 17.2 / 17.7 with 2 / 4 LOP3, then 22.4 / 29.9 / 36.3 / 56.6 with 8 / 12 / 16 / 24.
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "CUDA %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void iop(int& x, int y) { asm volatile("lop3.b32 %0, %0, %1, %1, 0x96;\n" : "+r"(x) : "r"(y)); }
__device__ __forceinline__ void fop(float& x, float y) { asm volatile("fma.rn.f32 %0, %0, %1, %1;\n" : "+f"(x) : "f"(y)); }
__device__ __forceinline__ void dfma(double& x, double y) { asm volatile("fma.rn.f64 %0, %0, %1, %1;\n" : "+d"(x) : "d"(y)); }

// MODE 0: 6 independent DMMA chains + K integer ops per DMMA; MODE 1: 8 independent DFMA chains + K integer ops per DFMA;
// MODE 2 / 3: the same with FP32 FMA as the filler
template <int MODE, int K>
__global__ void __launch_bounds__(128) k_probe(double* out, int iters) {
  double c[8][2];
  for (int i = 0; i < 8; ++i) { c[i][0] = threadIdx.x * 1e-9; c[i][1] = i; }
  const double a = 1.0 + threadIdx.x * 1e-12, b = 1e-3;
  int xi[4] = {(int)threadIdx.x, 1, 2, 3};
  float xf[4] = {1.f, 2.f, 3.f, 4.f};
  const int yi = blockIdx.x | 1;
  const float yf = 1e-3f;
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0 || MODE == 2) {
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        dmma884(c[j][0], c[j][1], a, b);
#pragma unroll
        for (int k = 0; k < K; ++k) { if (MODE == 0) iop(xi[k & 3], yi); else fop(xf[k & 3], yf); }
      }
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        dfma(c[j][0], b);
#pragma unroll
        for (int k = 0; k < K; ++k) { if (MODE == 1) iop(xi[k & 3], yi); else fop(xf[k & 3], yf); }
      }
    }
  }
  double s = 0.;
  for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + xi[0] + xi[1] + xi[2] + xi[3] + xf[0] + xf[1] + xf[2] + xf[3];
}

template <int MODE, int K>
void run(int warps_per_smsp, double* d_out) {
  int sm = 0, khz = 0;
  CK(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, 0));
  CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0));
  const int iters = 20000;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  // one CTA of 128 threads = one warp per sub-partition; warps_per_smsp CTAs per SM
  k_probe<MODE, K><<<sm * warps_per_smsp, 128>>>(d_out, 100);
  CK(cudaEventRecord(e0));
  k_probe<MODE, K><<<sm * warps_per_smsp, 128>>>(d_out, iters);
  CK(cudaEventRecord(e1));
  CK(cudaEventSynchronize(e1));
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  const int per_iter = (MODE == 0 || MODE == 2) ? 6 : 8;
  const double cyc = ms * 1e-3 * khz * 1e3 / ((double)iters * per_iter * warps_per_smsp);
  const char* names[4] = {"DMMA + K lop3", "DFMA + K lop3", "DMMA + K ffma", "DFMA + K ffma"};
  printf("%-14s K=%2d  %d warps/SMSP: %7.2f cycles per FP64 instruction per sub-partition (nominal clock %d MHz)\n", names[MODE], K,
         warps_per_smsp, cyc, khz / 1000);
}

int main() {
  double* d_out;
  CK(cudaMalloc(&d_out, sizeof(double) * 148 * 8 * 128));
  for (int w : {1, 2, 4}) {
    run<0, 0>(w, d_out); run<0, 2>(w, d_out); run<0, 4>(w, d_out); run<0, 8>(w, d_out); run<0, 12>(w, d_out); run<0, 16>(w, d_out); run<0, 24>(w, d_out);
    run<2, 4>(w, d_out); run<2, 12>(w, d_out);
    run<1, 0>(w, d_out); run<1, 1>(w, d_out); run<1, 2>(w, d_out); run<1, 3>(w, d_out); run<1, 4>(w, d_out);
    run<3, 1>(w, d_out); run<3, 2>(w, d_out);
  }
  return 0;
}
