// Shared helpers for every translation unit of libpmdist: error handling, bf16 packing, small
// device utilities.  sm_100a only (B200); there is no other code path.
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <utility>

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

// Host-side per-device state (function attributes already set, cached scratch pointers, SM counts): the C ABI may be
// driven on several GPUs from one process, so "did this once" has to be keyed by the current device.
// get() returns the zero-initialised slot of the current device (nullptr if the device cannot be determined).
template <typename T, int kMaxDevices = 64>
struct PerDevice {
  T slot[kMaxDevices] = {};
  T* get() {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
      set_error("cannot determine the current CUDA device (or ordinal >= 64)");
      return nullptr;
    }
    return &slot[dev];
  }
};

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

// Programmatic dependent launch.  A kernel launched with launch_pdl() may become resident while its predecessor
// in the stream is still draining; it runs its private prologue (barrier init, TMEM allocation, tensor-map
// prefetch) and then blocks in pdl_wait() until the predecessor has completed and its writes are visible.  The
// trigger sits right behind the wait, so at most one dependent grid is ever staged ahead.  Both instructions are
// no-ops for a kernel launched the ordinary way.  RULE: every kernel launched through launch_pdl() must execute
// pdl_wait() before its first access to global memory that another kernel produces or consumes.
__device__ __forceinline__ void pdl_wait() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
#endif

#ifdef __CUDACC__
// PMD_PDL=0 disables programmatic dependent launch (every kernel then waits for its predecessor to drain).
inline bool pdl_enabled() {
  static const bool on = getenv("PMD_PDL") == nullptr || atoi(getenv("PMD_PDL")) != 0;
  return on;
}
// cluster_x > 1: thread-block clusters of cluster_x CTAs along grid.x (CTA pairs for tcgen05 cta_group::2 kernels).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_cluster(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                      int cluster_x, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (cluster_x > 1) {
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = static_cast<unsigned>(cluster_x);
    attr[1].val.clusterDim.y = 1;
    attr[1].val.clusterDim.z = 1;
    cfg.numAttrs = 2;
  }
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(std::forward<Args>(args))...);
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  return launch_pdl_cluster(kernel, grid, block, smem, stream, 1, std::forward<Args>(args)...);
}
#endif

}  // namespace pmd
