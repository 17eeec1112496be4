// feasm_noalloc.cu — feasm_add_matrix_noalloc: SparseMatrix::add_matrix for one host element matrix WITHOUT any device
// allocation.  No counterpart in the reference; added for callers that cannot
// call cudaMalloc (tight serving loops, harnesses that forbid it, code that will later run under CUDA graph capture).
//
// feasm_add_matrix (feasm_api.cu) stages the dense block and its row / column ids in three cudaMalloc'ed scratch buffers
// per call.  Here that scratch is ONE file-scope __device__ array of a fixed, modest size (FEASM_NOALLOC_SCRATCH_BYTES =
// 132 096 B: a 128 x 128 block of doubles + 128 row ids + 128 column ids), resident on every GPU that loads the library
// until the process exits.  Blocks larger than FEASM_NOALLOC_MAX_DIM in either direction are rejected.
//
// The array's NAME is only a host-side stand-in in host code; the device address comes from cudaGetSymbolAddress, taken
// once per device with that device current, checked, cached, and passed to the kernel.
//
// Every call on a device shares the array.  Nothing in it has to start at zero (a call overwrites exactly the part it
// reads), but two calls that overlap on the GPU — two contexts on one device driven from two threads — race on it: callers
// serialise per device.  Calls on one context are serialised anyway (include/feasm.h).
#include <mutex>

#include "feasm_internal.h"

namespace {

constexpr int MAXD = FEASM_NOALLOC_MAX_DIM;
constexpr size_t DM_BYTES = sizeof(double) * MAXD * MAXD;
constexpr size_t IDS_BYTES = sizeof(uint32_t) * MAXD;
static_assert(DM_BYTES + 2 * IDS_BYTES == FEASM_NOALLOC_SCRATCH_BYTES, "include/feasm.h states the size of the scratch");

// dense block | row ids | column ids
static __device__ __align__(256) unsigned char g_noalloc_scratch[FEASM_NOALLOC_SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
std::mutex g_mutex;
unsigned char* g_scratch_addr[MAX_DEVICES] = {};  // device address of g_noalloc_scratch on device d, or null = not resolved yet

// The device address of the scratch on ctx's device (which must be current).  Resolved on the first call per device.
int scratch_address(feasm_ctx* ctx, unsigned char** out) {
  if (ctx->device < 0 || ctx->device >= MAX_DEVICES) return feasm_fail(ctx, FEASM_ERR_ARG, "feasm_add_matrix_noalloc: device ordinal beyond 63");
  std::lock_guard<std::mutex> lock(g_mutex);
  if (!g_scratch_addr[ctx->device]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p)
      return feasm_fail(ctx, FEASM_ERR_CUDA, std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e));
    g_scratch_addr[ctx->device] = static_cast<unsigned char*>(p);
  }
  *out = g_scratch_addr[ctx->device];
  return FEASM_OK;
}

}  // namespace

extern "C" int feasm_add_matrix_noalloc(feasm_ctx* ctx, const double* dm, const uint32_t* rows, int n_rows, const uint32_t* cols,
                                        int n_cols) {
  if (!ctx || !ctx->d_values) return feasm_fail(ctx, FEASM_ERR_STATE, "no matrix");
  if (!dm || !rows || !cols || n_rows <= 0 || n_cols <= 0) return feasm_fail(ctx, FEASM_ERR_ARG, "bad add_matrix arguments");
  if (n_rows > MAXD || n_cols > MAXD)
    return feasm_fail(ctx, FEASM_ERR_ARG,
                      "feasm_add_matrix_noalloc: block of " + std::to_string(n_rows) + " x " + std::to_string(n_cols) +
                          " exceeds the fixed scratch (FEASM_NOALLOC_MAX_DIM = " + std::to_string(MAXD) + " per side); use feasm_add_matrix");
  for (int i = 0; i < n_rows; ++i)
    if (rows[i] >= (uint64_t)ctx->n_rows) return feasm_fail(ctx, FEASM_ERR_ARG, "row out of range");
  FEASM_CUDA(ctx, cudaSetDevice(ctx->device));
  unsigned char* s = nullptr;
  int rc = scratch_address(ctx, &s);
  if (rc) return rc;
  double* d_dm = reinterpret_cast<double*>(s);
  uint32_t* d_r = reinterpret_cast<uint32_t*>(s + DM_BYTES);
  uint32_t* d_c = reinterpret_cast<uint32_t*>(s + DM_BYTES + IDS_BYTES);
  // the part of the scratch this call reads is overwritten here; the rest keeps an earlier call's bytes and is not read
  FEASM_CUDA(ctx, cudaMemcpyAsync(d_dm, dm, sizeof(double) * n_rows * n_cols, cudaMemcpyHostToDevice, ctx->stream));
  FEASM_CUDA(ctx, cudaMemcpyAsync(d_r, rows, sizeof(uint32_t) * n_rows, cudaMemcpyHostToDevice, ctx->stream));
  FEASM_CUDA(ctx, cudaMemcpyAsync(d_c, cols, sizeof(uint32_t) * n_cols, cudaMemcpyHostToDevice, ctx->stream));
  FEASM_CUDA(ctx, cudaMemsetAsync(ctx->d_flags, 0, 2 * sizeof(int), ctx->stream));
  rc = feasm_launch_add_matrix(ctx, d_dm, d_r, n_rows, d_c, n_cols);  // k_add_matrix (feasm_scatter.cu) on the real device address
  if (rc) return rc;
  int flags[2] = {0, 0};
  FEASM_CUDA(ctx, cudaMemcpyAsync(flags, ctx->d_flags, sizeof(flags), cudaMemcpyDeviceToHost, ctx->stream));
  FEASM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // also: the scratch is free for the next call
  if (flags[1]) {
    FEASM_CUDA(ctx, cudaMemsetAsync(ctx->d_flags, 0, sizeof(flags), ctx->stream));
    return feasm_fail(ctx, FEASM_ERR_PATTERN, "add to an entry outside the sparsity pattern");
  }
  return FEASM_OK;
}
