// Where does the ex4 forcing's time go?  Times ex4_forcing_t<true> (feasm_elem_dev.cuh) and variants in a register-resident
// loop at several occupancies; prints cycles per evaluation per SM sub-partition.
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I include -I libmesh_b200/csrc -o build/forcing_probe profiles/microbench/forcing_probe.cu
#include <cstdio>
#include <cstdlib>
#include "feasm_elem_dev.cuh"

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "CUDA %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

// variant 1: the polynomial without the per-lane coefficient selection (always the sin set): isolates FSEL
__device__ __forceinline__ double trig_nosel(double a, int add) {
  const double MAGIC = 6755399441055744.0;
  const double q = __fma_rn(a, 0x1.45f306dc9c883p-1, MAGIC);
  const int n = __double2loint(q) + add;
  const double j = __dadd_rn(q, -MAGIC);
  double r = __fma_rn(-j, 0x1.921fb54442d18p+0, a);
  r = __fma_rn(-j, 0x1.1a62633145c07p-54, r);
  r = __fma_rn(-j, -0x1.f1976b7ed8fbcp-110, r);
  const double x2 = __dmul_rn(r, r);
  double z = 1.58969099521155010221e-10;
  z = __fma_rn(z, x2, -2.50507602534068634195e-08);
  z = __fma_rn(z, x2, 2.75573137070700676789e-06);
  z = __fma_rn(z, x2, -1.98412698298579493134e-04);
  z = __fma_rn(z, x2, 8.33333333332248946124e-03);
  z = __fma_rn(z, x2, -1.66666666666666324348e-01);
  const double w = __fma_rn(z, x2, 0.0);
  const double res = __fma_rn(w, r, r);
  return (n & 2) ? -res : res;
}

template <int MODE>
__global__ void __launch_bounds__(256) k_probe(double* out, int iters, double dx) {
  double x = threadIdx.x * 1e-3 - 0.4, y = blockIdx.x * 1e-4 + 0.1, z = 0.3;
  double acc = 0.;
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) acc += ex4_forcing_t<true>(x, y, z);
    if (MODE == 1) {  // nine trig evaluations only, no selection
      const double hp = .5 * 3.141592653589793, eps = 1e-3;
      acc += trig_nosel(hp * x, 1) + trig_nosel(hp * (x - eps), 1) + trig_nosel(hp * (x + eps), 1) + trig_nosel(hp * y, 0) +
             trig_nosel(hp * (y - eps), 0) + trig_nosel(hp * (y + eps), 0) + trig_nosel(hp * z, 1) + trig_nosel(hp * (z - eps), 1) +
             trig_nosel(hp * (z + eps), 1);
    }
    if (MODE == 2) {  // nine trig evaluations with selection
      const double hp = .5 * 3.141592653589793, eps = 1e-3;
      acc += trig_fast(hp * x, 1) + trig_fast(hp * (x - eps), 1) + trig_fast(hp * (x + eps), 1) + trig_fast(hp * y, 0) +
             trig_fast(hp * (y - eps), 0) + trig_fast(hp * (y + eps), 0) + trig_fast(hp * z, 1) + trig_fast(hp * (z - eps), 1) +
             trig_fast(hp * (z + eps), 1);
    }
    x += dx; y += dx; z -= dx;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int MODE>
void run(const char* name, int blocks_per_sm, double* d_out) {
  int sm = 0;
  CK(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, 0));
  const int iters = 2000;
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  k_probe<MODE><<<sm * blocks_per_sm, 256>>>(d_out, 10, 1e-6);
  CK(cudaEventRecord(a));
  k_probe<MODE><<<sm * blocks_per_sm, 256>>>(d_out, iters, 1e-6);
  CK(cudaEventRecord(b));
  CK(cudaEventSynchronize(b));
  float ms = 0;
  cudaEventElapsedTime(&ms, a, b);
  int khz = 0;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const double warps_per_smsp = blocks_per_sm * 8 / 4.0;
  const double cyc = ms * 1e-3 * khz * 1e3 / (iters * warps_per_smsp);
  printf("%-28s %2d warps/SM: %.3f ms, %.0f cycles per warp-evaluation per SMSP\n", name, blocks_per_sm * 8, ms, cyc);
}

int main() {
  double* d_out;
  CK(cudaMalloc(&d_out, sizeof(double) * 148 * 8 * 256 * 2));
  for (int b : {1, 2, 3, 4}) {
    run<0>("ex4_forcing_t<true>", b, d_out);
    run<2>("9 trig_fast", b, d_out);
    run<1>("9 trig, no coefficient select", b, d_out);
  }
  return 0;
}
