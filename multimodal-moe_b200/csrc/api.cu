// libb200moe: error plumbing and the small host-only entry points of include/b200moe.h.
#include <stdarg.h>
#include <stdio.h>

#include "common.cuh"

namespace b200moe {

static thread_local char g_err[512] = "";
thread_local bool g_pdl = false;

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static std::atomic<long long> g_launches{0};   // kernels launched by this library (every launch site ends in check_launch)

int current_device() {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return -1; }
  return dev;
}

int num_sms() {
  static std::atomic<int> cached[64];
  const int dev = current_device();
  if (dev < 0 || dev >= 64) return 148;
  int n = cached[dev].load(std::memory_order_relaxed);
  if (n > 0) return n;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) {
    cudaGetLastError();
    return 148;
  }
  cached[dev].store(n, std::memory_order_relaxed);
  return n;
}

int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return B200MOE_ECUDA;
  }
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return B200MOE_OK;
}

}  // namespace b200moe

extern "C" {

int b200moe_version(void) { return 100; }

int64_t b200moe_launch_count(int reset) {
  return reset ? b200moe::g_launches.exchange(0) : b200moe::g_launches.load();
}

const char* b200moe_last_error(void) { return b200moe::g_err; }

int b200moe_check_device(void) {
  int dev = 0, major = 0;
  if (cudaGetDevice(&dev) != cudaSuccess ||
      cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) {
    cudaGetLastError();
    b200moe::set_error("no CUDA device available (libb200moe has no CPU fallback)");
    return B200MOE_EARCH;
  }
  if (major != 10) {
    b200moe::set_error("device compute capability %d.x is not sm_100 (libb200moe is B200-only)", major);
    return B200MOE_EARCH;
  }
  return B200MOE_OK;
}

int b200moe_rows_upper_bound(int T, int k, int E, int capacity) {
  if (T < 0 || k <= 0 || E <= 0) return B200MOE_EINVAL;
  long long kept = 1LL * T * k;
  if (capacity >= 0 && 1LL * E * capacity < kept) kept = 1LL * E * capacity;
  long long rows = kept + 1LL * E * (B200MOE_ROW_PAD - 1);
  rows = (rows + B200MOE_ROW_PAD - 1) / B200MOE_ROW_PAD * B200MOE_ROW_PAD;
  if (rows < B200MOE_ROW_PAD) rows = B200MOE_ROW_PAD;
  if (rows > 0x7fffff00LL) return B200MOE_EINVAL;
  return static_cast<int>(rows);
}

int b200moe_layer_row_pad(int T, int k, int E, int capacity, int d, int F) {
  if (T <= 0 || k <= 0 || E <= 0 || d % 256 != 0 || F % 256 != 0) return B200MOE_ROW_PAD;
  long long kept = 1LL * T * k;
  if (capacity >= 0 && 1LL * E * capacity < kept) kept = 1LL * E * capacity;
  // CTA pairs pay off once the narrower GEMM (N = d) still has two waves of 256 x 256 tiles
  return (kept / 256) * (d / 256) >= 2LL * b200moe::num_sms() ? 256 : B200MOE_ROW_PAD;
}

int b200moe_layer_rows_cap(int T, int k, int E, int capacity, int d, int F) {
  const int pad = b200moe_layer_row_pad(T, k, E, capacity, d, F);
  if (pad == B200MOE_ROW_PAD) return b200moe_rows_upper_bound(T, k, E, capacity);
  if (T < 0 || k <= 0 || E <= 0) return B200MOE_EINVAL;
  long long kept = 1LL * T * k;
  if (capacity >= 0 && 1LL * E * capacity < kept) kept = 1LL * E * capacity;
  long long rows = kept + 1LL * E * (pad - 1);
  rows = (rows + pad - 1) / pad * pad;
  if (rows > 0x7fffff00LL) return B200MOE_EINVAL;
  return static_cast<int>(rows);
}

int b200moe_router_block(int T) { return b200moe::router_block_tokens(T); }

int b200moe_num_router_blocks(int T) {
  const int tb = b200moe::router_block_tokens(T);
  return T <= 0 ? 0 : (T + tb - 1) / tb;
}

}  // extern "C"
