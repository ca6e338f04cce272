// Host-side runtime helpers of the native library: error reporting, device properties, launch counter.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace afl_rt {
int num_sms();
int max_smem_optin();
void set_error(const char* fmt, ...);
int check_last(const char* where);          // returns 0 or a non-zero code and records the message
void count_launch(int n = 1);
}  // namespace afl_rt

#define AFL_CHECK_ARG(cond, msg) do { if (!(cond)) { afl_rt::set_error("%s", msg); return -1; } } while (0)
#define AFL_CUDA(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { \
  afl_rt::set_error("%s failed: %s", #expr, cudaGetErrorString(_e)); return (int)_e; } } while (0)
