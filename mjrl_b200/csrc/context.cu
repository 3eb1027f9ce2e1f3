// Context, error reporting and scratch management of the C-ABI (include/mjrl_b200.h).
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

static thread_local char g_err[512] = "";

void mjrl_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int mjrl_check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    mjrl_set_error("%s: %s", what, cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

static void* grow(void** buf, size_t* have, size_t want) {
  if (*have >= want) return *buf;
  // scratch is only touched by kernels already enqueued; drain before freeing
  cudaDeviceSynchronize();
  if (*buf) cudaFree(*buf);
  *buf = nullptr;
  *have = 0;
  size_t sz = want < (1u << 20) ? (1u << 20) : want;
  cudaError_t e = cudaMalloc(buf, sz);
  if (e != cudaSuccess) {
    mjrl_set_error("scratch cudaMalloc(%zu): %s", sz, cudaGetErrorString(e));
    *buf = nullptr;
    return nullptr;
  }
  *have = sz;
  return *buf;
}

void* mjrl_scratch(mjrl_ctx* ctx, size_t bytes) { return grow(&ctx->scratch, &ctx->scratch_bytes, bytes); }
void* mjrl_scratch2(mjrl_ctx* ctx, size_t bytes) { return grow(&ctx->scratch2, &ctx->scratch2_bytes, bytes); }

extern "C" {

int mjrl_abi_version(void) { return MJRL_ABI_VERSION; }

const char* mjrl_last_error(void) { return g_err; }

int mjrl_ctx_create(mjrl_ctx** out, int device) {
  MJRL_REQUIRE(out != nullptr, MJRL_E_ARG, "ctx_create: out is null");
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    mjrl_set_error("ctx_create: no CUDA device (%s)", cudaGetErrorString(e));
    return MJRL_E_NOGPU;
  }
  MJRL_REQUIRE(device >= 0 && device < count, MJRL_E_ARG, "ctx_create: bad device %d", device);
  e = cudaSetDevice(device);
  if (e != cudaSuccess) {
    mjrl_set_error("cudaSetDevice: %s", cudaGetErrorString(e));
    return (int)e;
  }
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, device);
  if (prop.major != 10) {
    mjrl_set_error("ctx_create: device is sm_%d%d; this library is built for sm_100a only",
                   prop.major, prop.minor);
    return MJRL_E_NOGPU;
  }
  mjrl_ctx* c = new mjrl_ctx();
  memset(c, 0, sizeof(*c));
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  c->max_smem_optin = (int)prop.sharedMemPerBlockOptin;
  *out = c;
  return 0;
}

int mjrl_ctx_destroy(mjrl_ctx* ctx) {
  if (!ctx) return 0;
  cudaSetDevice(ctx->device);
  if (ctx->scratch) cudaFree(ctx->scratch);
  if (ctx->scratch2) cudaFree(ctx->scratch2);
  delete ctx;
  return 0;
}

int mjrl_ctx_sm_count(const mjrl_ctx* ctx) { return ctx ? ctx->sm_count : 0; }

}  // extern "C"
