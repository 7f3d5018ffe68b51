// Shared helpers for every translation unit of libpmdist: error handling, bf16 packing, small
// device utilities.  sm_100a only (B200); there is no other code path.
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>

#define PMD_API __attribute__((visibility("default")))

namespace pmd {

// Last error message, readable through pmd_last_error() (thread-local so ranks-as-threads tests work).
void set_error(const std::string& msg);
const char* last_error_cstr();

#define PMD_CHECK_CUDA(expr)                                                                   \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess) {                                                                   \
      ::pmd::set_error(std::string(#expr) + " failed: " + cudaGetErrorString(_e) + " at " +    \
                       __FILE__ + ":" + std::to_string(__LINE__));                             \
      return -1;                                                                               \
    }                                                                                          \
  } while (0)

#define PMD_REQUIRE(cond, msg)                                                                 \
  do {                                                                                         \
    if (!(cond)) {                                                                             \
      ::pmd::set_error(std::string(msg) + " (" #cond ") at " + __FILE__ + ":" +                \
                       std::to_string(__LINE__));                                              \
      return -1;                                                                               \
    }                                                                                          \
  } while (0)

#define PMD_LAUNCH_CHECK()                                                                     \
  do {                                                                                         \
    cudaError_t _e = cudaGetLastError();                                                       \
    if (_e != cudaSuccess) {                                                                   \
      ::pmd::set_error(std::string("kernel launch failed: ") + cudaGetErrorString(_e) +        \
                       " at " + __FILE__ + ":" + std::to_string(__LINE__));                    \
      return -1;                                                                               \
    }                                                                                          \
  } while (0)

// Counts every kernel this library launches (bench.py reports it as "gpu_launches").
extern unsigned long long g_launch_count;
inline void count_launch(int n = 1) { g_launch_count += (unsigned long long)n; }

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
static inline long long ceil_div_ll(long long a, long long b) { return (a + b - 1) / b; }

#ifdef __CUDACC__
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(v);
}
__device__ __forceinline__ float bf16_round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ void red_add_f32(float* addr, float v) {
  asm volatile("red.global.add.f32 [%0], %1;" ::"l"(addr), "f"(v) : "memory");
}
#endif

}  // namespace pmd
