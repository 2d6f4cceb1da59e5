// Micro-benchmark: MUFU.EX2 issue rate per scheduler on sm_100a (how many cycles a warp-wide ex2.approx occupies the XU pipe),
// alone and inside the softmax instruction mix (FFMA -> EX2 -> FADD, F2FP pack per pair).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mufu_rate mufu_rate.cu && ./mufu_rate
#include <cstdio>
#include <cuda_bf16.h>
__device__ __forceinline__ float ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

template <int MODE>
__global__ void k(float* out, long long* cyc, int iters) {
  float v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = threadIdx.x * 1e-3f + i;
  float acc = 0.f; unsigned pk = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; i += 2) {
      if (MODE == 0) { v[i] = ex2(v[i]); v[i + 1] = ex2(v[i + 1]); }
      else {
        const float a = ex2(fmaf(v[i], 0.99f, -0.5f)), b = ex2(fmaf(v[i + 1], 0.99f, -0.5f));
        acc += a + b;
        __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
        pk ^= *reinterpret_cast<unsigned*>(&t);
        v[i] = a * 1e-3f; v[i + 1] = b * 1e-3f;
      }
    }
  }
  const long long t1 = clock64();
  float s = acc + __uint_as_float(pk);
#pragma unroll
  for (int i = 0; i < 16; ++i) s += v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
  float* out; long long* cyc; cudaMalloc(&out, 1 << 20); cudaMallocManaged(&cyc, 8);
  const int iters = 2000;
  for (int warps = 4; warps <= 32; warps *= 2) {
    k<0><<<1, warps * 32>>>(out, cyc, iters); cudaDeviceSynchronize();
    k<0><<<1, warps * 32>>>(out, cyc, iters); cudaDeviceSynchronize();
    const double per = (double)*cyc / (iters * 16.0);
    k<1><<<1, warps * 32>>>(out, cyc, iters); cudaDeviceSynchronize();
    const double per1 = (double)*cyc / (iters * 16.0);
    printf("warps/SM %2d (per scheduler %d): pure EX2 %.2f cycles per warp-EX2 per warp => %.2f cycles per EX2 per scheduler; softmax mix %.2f => %.2f\n",
           warps, warps / 4, per, per / (warps / 4), per1, per1 / (warps / 4));
  }
  return 0;
}
