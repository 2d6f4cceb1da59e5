// Common definitions for the cosyvoice_b200 native library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

namespace cvb {

// ---------------------------------------------------------------- error reporting (C-ABI: int status + last_error)
void set_last_error(const std::string& msg);
const char* get_last_error();

#define CVB_OK 0
#define CVB_ERR_INVALID 1
#define CVB_ERR_CUDA 2
#define CVB_ERR_WORKSPACE 3
#define CVB_ERR_NOTFOUND 4

#define CVB_CUDA(expr)                                                                          \
  do {                                                                                          \
    cudaError_t _e = (expr);                                                                    \
    if (_e != cudaSuccess) {                                                                    \
      ::cvb::set_last_error(std::string(#expr) + " failed: " + cudaGetErrorString(_e) + " at " + \
                            __FILE__ + ":" + std::to_string(__LINE__));                         \
      return CVB_ERR_CUDA;                                                                      \
    }                                                                                           \
  } while (0)

#define CVB_REQUIRE(cond, msg)                                                              \
  do {                                                                                      \
    if (!(cond)) {                                                                          \
      ::cvb::set_last_error(std::string("invalid argument: ") + (msg) + " (" #cond ") at " + \
                            __FILE__ + ":" + std::to_string(__LINE__));                     \
      return CVB_ERR_INVALID;                                                               \
    }                                                                                       \
  } while (0)

#define CVB_TRY(expr)            \
  do {                           \
    int _s = (expr);             \
    if (_s != CVB_OK) return _s; \
  } while (0)

void count_launch();
#define CVB_LAUNCH_CHECK()       \
  do {                           \
    ::cvb::count_launch();       \
    CVB_CUDA(cudaGetLastError()); \
  } while (0)

// ---------------------------------------------------------------- programmatic dependent launch
// Every kernel of the library is launched with the programmatic-stream-serialization attribute and begins with
// pdl_sync(): the next kernel's CTAs may be scheduled (and run their prologue: barrier init, TMEM allocation,
// tensor-map prefetch) while the previous grid drains, but touch global memory only after it has fully completed.
// The estimator is ~540 dependent launches per call, so the launch gap matters for chunk latency.
bool pdl_enabled();
void set_pdl_enabled(bool on);

#ifdef __CUDACC__
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_sync() {
  pdl_trigger();
  pdl_wait();
}

template <class... KArgs, class... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
#endif

// ---------------------------------------------------------------- dtypes
// Activation/weight storage type of a run: fp32 (parity mode, CUDA cores) or bf16 (tensor cores).
enum DType : int { DT_F32 = 0, DT_BF16 = 1 };

__host__ __device__ inline int dtype_size(int dt) { return dt == DT_BF16 ? 2 : 4; }

__device__ __forceinline__ float ld_as_float(const void* p, size_t idx, int dt) {
  return dt == DT_BF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[idx])
                       : reinterpret_cast<const float*>(p)[idx];
}
__device__ __forceinline__ void st_from_float(void* p, size_t idx, int dt, float v) {
  if (dt == DT_BF16) reinterpret_cast<__nv_bfloat16*>(p)[idx] = __float2bfloat16_rn(v);
  else reinterpret_cast<float*>(p)[idx] = v;
}

template <typename T> __device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

constexpr int TILE_M = 128;     // rows per M tile everywhere (one tile never spans two sequences)
constexpr int NUM_SMS = 148;

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline int round_up(int a, int b) { return ceil_div(a, b) * b; }

// ---------------------------------------------------------------- math helpers
__device__ __forceinline__ float mishf(float x) {
  // x * tanh(softplus(x)); softplus with torch's threshold 20
  float sp = x > 20.f ? x : log1pf(__expf(x));
  return x * tanhf(sp);
}
__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752f)); }
__device__ __forceinline__ float eluf(float x) { return x > 0.f ? x : expm1f(x); }
__device__ __forceinline__ float siluf(float x) { return x / (1.f + __expf(-x)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace cvb
