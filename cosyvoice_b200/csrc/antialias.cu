// Anti-aliased SnakeBeta activation (BigVGAN "Activation1d"): 2x Kaiser-sinc upsample, SnakeBeta, 2x downsample.
//   reference math   cosyvoice/BigVGAN/alias_free_activation/torch/act.py:26-31, resample.py:27-58, filter.py
//   reference kernel cosyvoice/BigVGAN/alias_free_activation/cuda/anti_alias_activation_cuda.cu:44-181
//                    (one thread per 32 outputs with uncoalesced scalar loads, register arrays of 2*12+2*32+10)
//
//   up[2a]   = 2 * sum_{m<6} fu[1+2m] * x[cl(a+2-m)]      up[2a+1] = 2 * sum_{m<6} fu[2m] * x[cl(a+3-m)]
//   z[u]     = up[u] + sin^2(up[u] * e^alpha) / (e^beta + 1e-9)
//   y[t]     = sum_{q<12} fd[q] * z[clamp(2t + q - 5, 0, 2T-1)]
//
// This file: a register-only streaming formulation.  A thread walks along time carrying a 7-sample x window and 7
// partial output sums; every step consumes one x sample, produces the two z samples of that position and scatters
// them into the partial sums, so each x is loaded once, each z computed once and nothing goes through shared
// memory in the inner loop.  All arithmetic is packed fp32x2 (FFMA2, sm_100): the two lanes are two adjacent
// channels (rows layout) or two adjacent (b, c) rows (NCT layout).  HBM-bound by design: 1 load + 1 store per
// output against 24 packed FMAs per output pair.
#include <cuda_bf16.h>

#include <type_traits>

#include "common.cuh"
#include "../../include/cvb200.h"

namespace cvb {

using Layout = cvb200_layout;

namespace {

struct AAFilt {
  float2 u[12];     // upsample taps (fu), each duplicated into both lanes
  float2 d[12];     // downsample taps (fd)
};

__device__ __forceinline__ float2 f2(float a) { return make_float2(a, a); }

#ifdef AA_SCALAR_MATH
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) { return make_float2(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y)); }
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) { return make_float2(a.x * b.x, a.y * b.y); }
#else
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) { return __fmul2_rn(a, b); }
#endif

// sin for the fp32-output kernels: two-term Cody-Waite reduction to [-pi, pi] and the SFU on the remainder.  Absolute
// error <= 2^-21.4 at any magnitude the activations reach (plain __sinf loses accuracy with |x|; libm sinf costs ~4x
// the instructions, which made the fp32 kernels compute-bound at half the bf16 rate).  -DAA_LIBM_SIN restores sinf.
__device__ __forceinline__ float sin_reduced(float t) {
#ifdef AA_LIBM_SIN
  return sinf(t);
#else
  const float k = rintf(t * 0.15915494309189535f);
  float r = fmaf(k, -6.2831854820251465f, t);
  r = fmaf(k, 1.7484555e-7f, r);
  return __sinf(r);
#endif
}

template <bool FAST>
__device__ __forceinline__ float2 snake_beta2(float2 up, float2 ea, float2 ib) {
  // z = up + sin^2(up e^alpha) ib
  const float2 t = fmul2(up, ea);
  float2 s;
  if (FAST) {
    s.x = __sinf(t.x);
    s.y = __sinf(t.y);
  } else {
    s.x = sin_reduced(t.x);
    s.y = sin_reduced(t.y);
  }
  return ffma2(fmul2(s, ib), s, up);
}

// One streaming pass over positions a = a0 .. a0 + n_steps - 1.  load(a) returns x[clamp(a)] for both lanes,
// emit(t, y) receives the finished output of position t = a - 3 (callers drop t before their strip).
// zc = z[0] (stands for every z[u < 0]); only needed when the strip touches the sequence start.
// F.u holds 2 * fu (the upsampler's gain folded into its taps: exact).
template <bool FAST, bool SAME, class Load, class Emit>
__device__ __forceinline__ void aa_stream(const AAFilt& F, float2 ea, float2 ib, int a0, int n_steps, int T, float2 zc,
                                          Load load, Emit emit) {
  float2 xw[7], acc[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) xw[i + 1] = load(a0 - 3 + i);      // window of a0 - 1, shifted in by the first step
#pragma unroll
  for (int i = 0; i < 6; ++i) acc[i] = f2(0.f);
  xw[0] = xw[1];
  float2 zlast = zc;
  int a = a0;
  // MODE 0: a < 0 (z = z[0]); 1: 0 <= a < T; 2: a >= T (z = z[2T-1])
  auto step = [&](auto mode) {
    constexpr int MODE = decltype(mode)::value;
#pragma unroll
    for (int i = 0; i < 6; ++i) xw[i] = xw[i + 1];
    xw[6] = load(a + 3);
    float2 z0, z1;
    if (MODE == 0) {
      z0 = z1 = zc;
    } else if (MODE == 2) {
      z0 = z1 = zlast;
    } else {
      float2 u0 = fmul2(F.u[11], xw[0]);
      u0 = ffma2(F.u[9], xw[1], u0);
      u0 = ffma2(F.u[7], xw[2], u0);
      u0 = ffma2(F.u[5], xw[3], u0);
      u0 = ffma2(F.u[3], xw[4], u0);
      u0 = ffma2(F.u[1], xw[5], u0);
      float2 u1 = fmul2(F.u[10], xw[1]);
      u1 = ffma2(F.u[8], xw[2], u1);
      u1 = ffma2(F.u[6], xw[3], u1);
      u1 = ffma2(F.u[4], xw[4], u1);
      u1 = ffma2(F.u[2], xw[5], u1);
      u1 = ffma2(F.u[0], xw[6], u1);
      z0 = snake_beta2<FAST>(u0, ea, ib);
      z1 = snake_beta2<FAST>(u1, ea, ib);
      zlast = z1;
    }
    // scatter: y[a+k] += fd[5-2k] z0 + fd[6-2k] z1, k = -3..3; y[a-3] is complete after this position.
    // SAME (fd == fu): the doubled up-taps serve as down-taps and the finished sum is halved (exact).
    const float2* D = SAME ? F.u : F.d;
    const float2 y = ffma2(D[11], z0, acc[0]);
    emit(a - 3, SAME ? fmul2(y, f2(0.5f)) : y);
    acc[0] = ffma2(D[10], z1, ffma2(D[9], z0, acc[1]));
    acc[1] = ffma2(D[8], z1, ffma2(D[7], z0, acc[2]));
    acc[2] = ffma2(D[6], z1, ffma2(D[5], z0, acc[3]));
    acc[3] = ffma2(D[4], z1, ffma2(D[3], z0, acc[4]));
    acc[4] = ffma2(D[2], z1, ffma2(D[1], z0, acc[5]));
    acc[5] = fmul2(D[0], z1);
    ++a;
  };
  using M0 = std::integral_constant<int, 0>;
  using M1 = std::integral_constant<int, 1>;
  using M2 = std::integral_constant<int, 2>;
  const int a_end = a0 + n_steps;
  const int main_end = min(a_end, T);
#pragma unroll 1
  while (a < 0 && a < a_end) step(M0{});
#pragma unroll 1
  while (a + 7 <= main_end) {             // 7 = window length: the shifts become register renaming
#pragma unroll
    for (int s = 0; s < 7; ++s) step(M1{});
  }
#pragma unroll 1
  while (a < main_end) step(M1{});
#pragma unroll 1
  while (a < a_end) step(M2{});
}

// z[0] of a sequence: the window of a = 0 is x0,x0,x0,x0,x1,x2,x3
template <bool FAST>
__device__ __forceinline__ float2 aa_z0(const AAFilt& F, float2 ea, float2 ib, float2 x0, float2 x1, float2 x2) {
  float2 u0 = fmul2(F.u[11], x0);
  u0 = ffma2(F.u[9], x0, u0);
  u0 = ffma2(F.u[7], x0, u0);
  u0 = ffma2(F.u[5], x0, u0);
  u0 = ffma2(F.u[3], x1, u0);
  u0 = ffma2(F.u[1], x2, u0);
  return snake_beta2<FAST>(u0, ea, ib);
}

template <bool SAME>
__device__ __forceinline__ void load_filters(AAFilt& F, const float* fu, const float* fd) {
#pragma unroll
  for (int i = 0; i < 12; ++i) {
    F.u[i] = f2(2.f * __ldg(fu + i));
    if (!SAME) F.d[i] = f2(__ldg(fd + i));
  }
}

__device__ __forceinline__ bool row_info(const Layout& lay, int row, int& b, int& lrow, int& len) {
  const int tile = row >> 7;
  b = __ldg(lay.tile_seq + tile);
  lrow = __ldg(lay.tile_row0 + tile) + (row & 127);
  len = __ldg(lay.seq_len + b);
  return lrow < len;
}

template <int DT>
__device__ __forceinline__ float2 ld2(const void* p, size_t idx) {
  if (DT == CVB200_BF16) {
    const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(reinterpret_cast<const __nv_bfloat16*>(p) + idx);
    return __bfloat1622float2(v);
  } else {
    return *reinterpret_cast<const float2*>(reinterpret_cast<const float*>(p) + idx);
  }
}

template <int DT>
__device__ __forceinline__ void st2(void* p, size_t idx, float2 v) {
  if (DT == CVB200_BF16) {
    *reinterpret_cast<__nv_bfloat162*>(reinterpret_cast<__nv_bfloat16*>(p) + idx) = __float22bfloat162_rn(v);
  } else {
    *reinterpret_cast<float2*>(reinterpret_cast<float*>(p) + idx) = v;
  }
}

// ---------------------------------------------------------------------------------------------- rows layout
// Time-major varlen rows [rows, ld] (the trunk's activation layout).  One thread = one (128-row strip, channel
// pair): consecutive threads take consecutive channel pairs, so every row access of a warp is one contiguous
// segment.  Pad rows (t >= len) are written as zeros (the layout invariant every conv relies on).
constexpr int AA_TS = 128;

template <int XDT, int ODT>
__global__ void __launch_bounds__(128) aa_rows_kernel(const void* __restrict__ x, int ldx, void* __restrict__ out, int ldo,
                                                      const float* __restrict__ filt, const float* __restrict__ alpha,
                                                      const float* __restrict__ beta, int C, const Layout lay,
                                                      long long n_items) {
  pdl_sync();
  constexpr bool FAST = ODT == CVB200_BF16;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n_items) return;
  const int hc = C >> 1;
  const int cp = (int)(idx % hc);
  const int row0 = (int)(idx / hc) * AA_TS;
  const int ch = cp * 2;
  int b, t0, T;
  row_info(lay, row0, b, t0, T);
  const size_t obase = (size_t)row0 * ldo + ch;
  if (t0 >= T) {
    for (int j = 0; j < AA_TS; ++j) st2<ODT>(out, obase + (size_t)j * ldo, f2(0.f));
    return;
  }
  AAFilt F;
  load_filters<true>(F, filt, filt);
  const float2 al = *reinterpret_cast<const float2*>(alpha + ch);
  const float2 be = *reinterpret_cast<const float2*>(beta + ch);
  const float2 ea = make_float2(expf(al.x), expf(al.y));
  const float2 ib = make_float2(1.f / (expf(be.x) + 1e-9f), 1.f / (expf(be.y) + 1e-9f));
  constexpr int XB = XDT == CVB200_BF16 ? 2 : 4, OB = ODT == CVB200_BF16 ? 2 : 4;
  const char* xseq = reinterpret_cast<const char*>(x) + ((size_t)(row0 - t0) * ldx + ch) * XB;   // frame 0 of the sequence
  const int ldxb = ldx * XB, ldob = ldo * OB;
  const int Tm1 = T - 1;
  // the streaming loads have only one 7-step iteration of lookahead: a row-ahead L2 prefetch (no destination
  // register, so any depth is free) turns their DRAM latency into L2-hit latency
  constexpr int PF_AHEAD = 16;
  auto load = [&](int a) -> float2 {
    const int t = min(max(a, 0), Tm1);
    const int tp = min(a + PF_AHEAD, Tm1);
    if (a >= 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(xseq + (long long)tp * ldxb));
    return ld2<XDT>(xseq + (long long)t * ldxb, 0);
  };
  float2 zc = f2(0.f);
  if (t0 < 3) zc = aa_z0<FAST>(F, ea, ib, load(0), load(1), load(2));
  char* op = reinterpret_cast<char*>(out) + obase * OB - 6LL * ldob;      // the first position emitted is t0 - 6
  auto emit = [&](int t, float2 y) {
    if (t >= t0) st2<ODT>(op, 0, y);
    op += ldob;
  };
  const int n_valid = min(AA_TS, T - t0);
  aa_stream<FAST, true>(F, ea, ib, t0 - 3, n_valid + 6, T, zc, load, emit);
  for (int t = t0 + n_valid; t < t0 + AA_TS; ++t)      // pad rows of the strip
    st2<ODT>(out, obase + (size_t)(t - t0) * ldo, f2(0.f));
}

// ---------------------------------------------------------------------------------------------- [B, C, T] layout
// Drop-in for the reference extension's fwd_cuda (time contiguous).  A CTA takes two adjacent (b, c) rows and
// NC_N = 128 x 32 consecutive samples of them: the samples (+ 6 halo each side, replicate-clamped) are staged
// in shared memory with coalesced loads, every thread streams over its own 32-sample strip (index j stored at
// j + j/32: conflict-free for the per-thread strided walk), outputs go back through shared memory so that the
// global stores are coalesced too.
constexpr int NC_THREADS = 128;
constexpr int NC_S = 32;
constexpr int NC_N = NC_THREADS * NC_S;
constexpr int NC_XN = NC_N + 12;
__host__ __device__ constexpr int nc_pad(int j) { return j + (j >> 5); }
constexpr int NC_XP = nc_pad(NC_XN) + 1, NC_OP = nc_pad(NC_N) + 1;
constexpr int NC_SMEM = (2 * NC_XP + 2 * NC_OP) * (int)sizeof(float);

template <int DT>
__global__ void __launch_bounds__(NC_THREADS) aa_nct_kernel(const void* __restrict__ x, void* __restrict__ y, int C, int T,
                                                            long long n_rows, const float* __restrict__ fu,
                                                            const float* __restrict__ fd, const float* __restrict__ alpha,
                                                            const float* __restrict__ beta) {
  pdl_sync();
  constexpr bool FAST = DT == CVB200_BF16;
  extern __shared__ float nc_smem[];
  float* sx0 = nc_smem;
  float* sx1 = sx0 + NC_XP;
  float* so0 = sx1 + NC_XP;
  float* so1 = so0 + NC_OP;
  const int s0 = blockIdx.x * NC_N;
  const long long r0 = 2LL * blockIdx.y;
  const bool two = r0 + 1 < n_rows;
  const long long r1 = two ? r0 + 1 : r0;
  const size_t base0 = (size_t)r0 * T, base1 = (size_t)r1 * T;
  const int n_here = min(NC_N, T - s0);
  const int Tm1 = T - 1;
  for (int j = threadIdx.x; j < n_here + 12; j += NC_THREADS) {
    const int t = min(max(s0 - 6 + j, 0), Tm1);
    float v0, v1;
    if (DT == CVB200_BF16) {
      v0 = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(x)[base0 + t]);
      v1 = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(x)[base1 + t]);
    } else {
      v0 = reinterpret_cast<const float*>(x)[base0 + t];
      v1 = reinterpret_cast<const float*>(x)[base1 + t];
    }
    sx0[nc_pad(j)] = v0;
    sx1[nc_pad(j)] = v1;
  }
  __syncthreads();
  const int tl = threadIdx.x * NC_S;            // strip start inside the CTA's window
  if (tl < n_here) {
    AAFilt F;
    load_filters<false>(F, fu, fd);
    const int c0 = (int)(r0 % C), c1 = (int)(r1 % C);
    const float2 ea = make_float2(expf(__ldg(alpha + c0)), expf(__ldg(alpha + c1)));
    const float2 ib = make_float2(1.f / (expf(__ldg(beta + c0)) + 1e-9f), 1.f / (expf(__ldg(beta + c1)) + 1e-9f));
    const int t0 = s0 + tl;
    const int n_valid = min(NC_S, T - t0);
    // position a <-> staged index a - (s0 - 6); positions beyond the staged window only occur past T (clamped)
    const int jmax = n_here + 11;
    auto load = [&](int a) -> float2 {
      const int j = min(max(a - (s0 - 6), 0), jmax);
      const int p = nc_pad(j);
      return make_float2(sx0[p], sx1[p]);
    };
    float2 zc = f2(0.f);
    if (t0 < 3) zc = aa_z0<FAST>(F, ea, ib, load(0), load(1), load(2));
    auto emit = [&](int t, float2 v) {
      const int j = t - s0;
      if (t >= t0) {
        const int p = nc_pad(j);
        so0[p] = v.x;
        so1[p] = v.y;
      }
    };
    aa_stream<FAST, false>(F, ea, ib, t0 - 3, n_valid + 6, T, zc, load, emit);
  }
  __syncthreads();
  for (int j = threadIdx.x; j < n_here; j += NC_THREADS) {
    const int p = nc_pad(j);
    if (DT == CVB200_BF16) {
      reinterpret_cast<__nv_bfloat16*>(y)[base0 + s0 + j] = __float2bfloat16_rn(so0[p]);
      if (two) reinterpret_cast<__nv_bfloat16*>(y)[base1 + s0 + j] = __float2bfloat16_rn(so1[p]);
    } else {
      reinterpret_cast<float*>(y)[base0 + s0 + j] = so0[p];
      if (two) reinterpret_cast<float*>(y)[base1 + s0 + j] = so1[p];
    }
  }
}

}  // namespace

// ---------------------------------------------------------------------------------------------- launchers
int launch_anti_alias_rows(const void* x, int ldx, int x_dt, void* out, int ldo, int out_dt, const float* filt,
                           const float* alpha, const float* beta, int C, const Layout& lay, cudaStream_t st) {
  if (lay.rows <= 0 || C <= 0) return CVB_OK;
  CVB_REQUIRE(C % 2 == 0 && ldx % 2 == 0 && ldo % 2 == 0, "anti_alias_act_rows: C, ldx, ldo must be even");
  CVB_REQUIRE((reinterpret_cast<uintptr_t>(x) & 7) == 0 && (reinterpret_cast<uintptr_t>(out) & 7) == 0,
              "anti_alias_act_rows: x / out must be 8-byte aligned");
  const long long n_items = (long long)(lay.rows / AA_TS) * (C / 2);
  const unsigned grid = (unsigned)((n_items + 127) / 128);
#define AA_ROWS(XD, OD)                                                                                        \
  CVB_CUDA(launch_pdl(aa_rows_kernel<XD, OD>, dim3(grid), dim3(128), 0, st, x, ldx, out, ldo, filt, alpha, beta, C, lay, n_items))
  if (x_dt == CVB200_F32 && out_dt == CVB200_BF16) AA_ROWS(CVB200_F32, CVB200_BF16);
  else if (x_dt == CVB200_F32 && out_dt == CVB200_F32) AA_ROWS(CVB200_F32, CVB200_F32);
  else if (x_dt == CVB200_BF16 && out_dt == CVB200_BF16) AA_ROWS(CVB200_BF16, CVB200_BF16);
  else AA_ROWS(CVB200_BF16, CVB200_F32);
#undef AA_ROWS
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_anti_alias_nct(const void* x, void* y, int dt, int B, int C, int T, const float* fu, const float* fd,
                          const float* alpha, const float* beta, cudaStream_t st) {
  if (B <= 0 || C <= 0 || T <= 0) return CVB_OK;      // the reference kernel is a silent no-op on seq_len == 0 (.cu:195-198)
  const long long n_rows = (long long)B * C;
  CVB_REQUIRE((n_rows + 1) / 2 <= 65535, "anti_alias_act_nct: B*C/2 exceeds grid.y");
  static bool attr_done = false;
  if (!attr_done) {
    CVB_CUDA(cudaFuncSetAttribute(aa_nct_kernel<CVB200_F32>, cudaFuncAttributeMaxDynamicSharedMemorySize, NC_SMEM));
    CVB_CUDA(cudaFuncSetAttribute(aa_nct_kernel<CVB200_BF16>, cudaFuncAttributeMaxDynamicSharedMemorySize, NC_SMEM));
    attr_done = true;
  }
  dim3 grid((unsigned)ceil_div(T, NC_N), (unsigned)((n_rows + 1) / 2));
  if (dt == CVB200_BF16)
    CVB_CUDA(launch_pdl(aa_nct_kernel<CVB200_BF16>, dim3(grid), dim3(NC_THREADS), NC_SMEM, st, x, y, C, T, n_rows, fu, fd, alpha, beta));
  else
    CVB_CUDA(launch_pdl(aa_nct_kernel<CVB200_F32>, dim3(grid), dim3(NC_THREADS), NC_SMEM, st, x, y, C, T, n_rows, fu, fd, alpha, beta));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

}  // namespace cvb
