#include "runtime.h"

#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <atomic>

#include "../include/attackfl_b200.h"

namespace afl_rt {
namespace {
thread_local char g_err[512] = "";
std::atomic<long long> g_launches{0};
int g_sms[64];
int g_smem[64];
bool g_have[64];

int cur_dev() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
  return dev;
}

void query(int dev) {
  if (g_have[dev]) return;
  int v = 148, s = 227 * 1024;
  if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) v = 148;
  if (cudaDeviceGetAttribute(&s, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev) != cudaSuccess) s = 227 * 1024;
  g_sms[dev] = v;
  g_smem[dev] = s;
  g_have[dev] = true;
}
}  // namespace

int num_sms() { int d = cur_dev(); query(d); return g_sms[d]; }
int max_smem_optin() { int d = cur_dev(); query(d); return g_smem[d]; }

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

const char* last_error() { return g_err; }

int check_last(const char* where) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", where, cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
long long launches() { return g_launches.load(); }
void reset_launches() { g_launches.store(0); }
}  // namespace afl_rt

namespace afl_rt {
const char* last_error();
long long launches();
void reset_launches();
}  // namespace afl_rt

extern "C" {
const char* afl_last_error(void) { return afl_rt::last_error(); }
long long afl_launch_count(void) { return afl_rt::launches(); }
void afl_reset_launch_count(void) { afl_rt::reset_launches(); }
int afl_is_cuda_build(void) { return 1; }
int afl_num_sms(void) { return afl_rt::num_sms(); }
}
