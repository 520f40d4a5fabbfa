// ctx.cu -- context, error reporting, scratch.
#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include <new>
#include "pmx_common.cuh"

static thread_local char g_err[512] = "";

int pmx_fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

int pmx_cuda_fail(cudaError_t e, const char* what) {
  snprintf(g_err, sizeof(g_err), "CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return PMX_E_CUDA;
}

void* pmx_scratch(pmx_ctx* ctx, int slot, size_t bytes) {
  if (bytes <= ctx->scratch_bytes[slot]) return ctx->scratch[slot];
  // the block belongs to the ctx's device whatever device the caller has current; a caller on
  // another device gets its launch refused below instead of scratch on the wrong GPU
  int cur = -1;
  cudaGetDevice(&cur);
  if (cur != ctx->device) {
    pmx_fail(PMX_E_INVALID, "ctx belongs to device %d but device %d is current: make the ctx's device current "
             "(cudaSetDevice / torch.cuda.device) around calls that use it", ctx->device, cur);
    return nullptr;
  }
  // cudaFree synchronises the device, so work still using the old block has finished.
  if (ctx->scratch[slot]) cudaFree(ctx->scratch[slot]);
  ctx->scratch[slot] = nullptr;
  ctx->scratch_bytes[slot] = 0;
  size_t want = (bytes + (size_t(1) << 20) - 1) & ~((size_t(1) << 20) - 1);
  cudaError_t e = cudaMalloc(&ctx->scratch[slot], want);
  if (e != cudaSuccess) { pmx_cuda_fail(e, "cudaMalloc(scratch)"); return nullptr; }
  ctx->scratch_bytes[slot] = want;
  return ctx->scratch[slot];
}

extern "C" {

int pmx_version(void) { return PMX_VERSION; }

const char* pmx_last_error(void) { return g_err; }

int pmx_ctx_create(int device, pmx_ctx** out) {
  if (!out) return pmx_fail(PMX_E_INVALID, "pmx_ctx_create: out is NULL");
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0)
    return pmx_fail(PMX_E_CUDA, "pmx_ctx_create: no CUDA device available (%s); this engine has no CPU path",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
  if (device < 0 || device >= n) return pmx_fail(PMX_E_INVALID, "pmx_ctx_create: device %d out of range [0,%d)", device, n);
  cudaDeviceProp prop;
  PMX_CUDA(cudaGetDeviceProperties(&prop, device));        // (does not change the caller's current device)
  if (prop.major < 10)
    return pmx_fail(PMX_E_CUDA, "pmx_ctx_create: device %d is sm_%d%d; libpmx is built for sm_100a only", device, prop.major, prop.minor);
  pmx_ctx* c = new (std::nothrow) pmx_ctx();
  if (!c) return pmx_fail(PMX_E_INVALID, "pmx_ctx_create: out of host memory");
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  for (int i = 0; i < PMX_SCR_SLOTS; ++i) { c->scratch[i] = nullptr; c->scratch_bytes[i] = 0; }
  *out = c;
  return PMX_OK;
}

void pmx_ctx_destroy(pmx_ctx* ctx) {
  if (!ctx) return;
  for (int i = 0; i < PMX_SCR_SLOTS; ++i) if (ctx->scratch[i]) cudaFree(ctx->scratch[i]);
  delete ctx;
}

size_t pmx_ctx_scratch_bytes(const pmx_ctx* ctx) {
  return ctx ? ctx->scratch_bytes[0] + ctx->scratch_bytes[1] + ctx->scratch_bytes[2] + ctx->scratch_bytes[3] : 0;
}

}  // extern "C"
