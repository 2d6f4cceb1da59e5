// Micro-benchmark: does ex2.approx.ftz.bf16x2 (two bf16 exponentials per MUFU instruction) double the exp throughput of the
// softmax inner loop on sm_100a?  Compares, per scheduler, cycles per PAIR of exponentials for (a) two ex2.approx.ftz.f32 and
// (b) one ex2.approx.ftz.bf16x2, each alone and inside a softmax-like mix (scale/offset FFMA, row sum, packed bf16 result).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mufu_bf16x2 mufu_bf16x2.cu && ./mufu_bf16x2
#include <cstdio>
#include <cuda_bf16.h>
__device__ __forceinline__ float ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ unsigned ex2x2(unsigned x) { unsigned y; asm volatile("ex2.approx.ftz.bf16x2 %0, %1;" : "=r"(y) : "r"(x)); return y; }
__device__ __forceinline__ unsigned pack(float a, float b) { __nv_bfloat162 t = __floats2bfloat162_rn(a, b); return *reinterpret_cast<unsigned*>(&t); }

template <int MODE>
__global__ void k(float* out, long long* cyc, int iters) {
  float v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = -(threadIdx.x * 1e-3f + i * 0.25f);
  float acc = 0.f; unsigned pk = 0;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; i += 2) {
      if (MODE == 0) {              // f32 pair, softmax mix
        const float a = ex2(fmaf(v[i], 0.99f, -0.5f)), b = ex2(fmaf(v[i + 1], 0.99f, -0.5f));
        acc += a + b;
        pk ^= pack(a, b);
        v[i] = -a; v[i + 1] = -b;
      } else if (MODE == 1) {       // bf16x2, softmax mix: the result IS the packed P; the row sum unpacks it
        const unsigned p = ex2x2(pack(fmaf(v[i], 0.99f, -0.5f), fmaf(v[i + 1], 0.99f, -0.5f)));
        const float a = __uint_as_float(p << 16), b = __uint_as_float(p & 0xffff0000u);
        acc += a + b;
        pk ^= p;
        v[i] = -a; v[i + 1] = -b;
      } else if (MODE == 2) {       // pure f32 pair
        v[i] = ex2(v[i]) - 2.f; v[i + 1] = ex2(v[i + 1]) - 2.f;
      } else {                      // pure bf16x2
        const unsigned p = ex2x2(__float_as_uint(v[i]));
        v[i] = __uint_as_float(p) ;
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
template <int MODE> double run(int warps, float* out, long long* cyc) {
  const int iters = 2000;
  k<MODE><<<1, warps * 32>>>(out, cyc, iters); cudaDeviceSynchronize();
  k<MODE><<<1, warps * 32>>>(out, cyc, iters); cudaDeviceSynchronize();
  return (double)*cyc / (iters * 8.0) / (warps / 4);      // cycles per pair per scheduler
}
int main() {
  float* out; long long* cyc; cudaMalloc(&out, 1 << 20); cudaMallocManaged(&cyc, 8);
  for (int warps = 4; warps <= 16; warps *= 2)
    printf("warps per scheduler %d: cycles per PAIR of exps per scheduler: f32 mix %.2f | bf16x2 mix %.2f | pure f32 %.2f | pure bf16x2 %.2f\n",
           warps / 4, run<0>(warps, out, cyc), run<1>(warps, out, cyc), run<2>(warps, out, cyc), run<3>(warps, out, cyc));
  return 0;
}
