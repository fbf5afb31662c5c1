// Error plumbing and device queries of the C-ABI (include/b200_paged_ops.h).
#include "common.cuh"

namespace b200 {

static thread_local char g_err[1024] = {0};

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

void clear_error() { g_err[0] = 0; }

static int g_pdl = 1;
bool pdl_enabled() { return g_pdl != 0; }

int sm_count() {
  static int cached = 0;
  if (cached > 0) return cached;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  int n = 0;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
  cached = n;
  return n;
}

}  // namespace b200

extern "C" {

const char* b200_last_error(void) { return b200::g_err; }

// Programmatic dependent launch for every kernel of the step (default on; 0 = classic stream order).
int b200_set_pdl(int enabled) {
  b200::g_pdl = enabled ? 1 : 0;
  return B200_OK;
}

int b200_abi_version(void) { return B200_ABI_VERSION; }

// Reference: PARTITION_SIZE (vllm_metal/metal/constants.py:4).
int b200_partition_size(void) { return 512; }

int b200_sm_count(void) {
  int n = b200::sm_count();
  if (n < 0) return b200::set_error(B200_ERR_CUDA, "no CUDA device: %s",
                                    cudaGetErrorString(cudaGetLastError()));
  return n;
}

// Reference: min_decode_grid() = 8 threadgroups per GPU core
// (paged_ops.cpp:102-115).  One B200 SM keeps 8 of our 128-thread decode CTAs
// resident at most, so the same multiplier is the "GPU is already full" bound.
int b200_min_decode_grid(void) {
  int n = b200_sm_count();
  return n < 0 ? n : 8 * n;
}

}  // extern "C"
