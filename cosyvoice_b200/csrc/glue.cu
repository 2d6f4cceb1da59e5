// Memory-bound glue of the flow estimator / CFM solver: LayerNorm, GroupNorm|LayerNorm + Mish
// (+ time bias, mask), timestep embedding, tiny fp32 linears, CFG input packing, fused CFG + Euler
// update, and the [C,T] <-> time-major transposes.  Warp-shuffle reductions, 128-bit accesses.
#include "common.cuh"
#include "tapgemm.cuh"

namespace cvb {

using Layout = cvb200_layout;

// ---------------------------------------------------------------------------- row helpers
// 8 consecutive channels of one row, as floats.
template <typename T> __device__ __forceinline__ void load8(const T* p, float (&v)[8]);
template <> __device__ __forceinline__ void load8<float>(const float* p, float (&v)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
template <> __device__ __forceinline__ void load8<__nv_bfloat16>(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __bfloat1622float2(h[i]);
    v[2 * i] = f.x; v[2 * i + 1] = f.y;
  }
}
__device__ __forceinline__ void store8(void* base, size_t idx, int dt, const float (&v)[8]) {
  if (dt == DT_BF16) {
    uint4 pk;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(base) + idx) = pk;
  } else {
    float* o = reinterpret_cast<float*>(base) + idx;
    *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(o + 4) = make_float4(v[4], v[5], v[6], v[7]);
  }
}
__device__ __forceinline__ void load8_dt(const void* base, size_t idx, int dt, float (&v)[8]) {
  if (dt == DT_BF16) load8<__nv_bfloat16>(reinterpret_cast<const __nv_bfloat16*>(base) + idx, v);
  else load8<float>(reinterpret_cast<const float*>(base) + idx, v);
}

// ---------------------------------------------------------------------------- LayerNorm(256)
// one warp per row; lane owns channels [8*lane, 8*lane+8)
__global__ void __launch_bounds__(256) layernorm256_kernel(const float* __restrict__ x, int ldx, void* __restrict__ out,
                                                           int ldo, int out_dt, const float* __restrict__ gamma,
                                                           const float* __restrict__ beta, int rows, float eps) {
  pdl_sync();
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  float v[8];
  load8<float>(x + (size_t)row * ldx + lane * 8, v);
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += v[i];
  const float mean = warp_sum(s) * (1.f / 256.f);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) { v[i] -= mean; q += v[i] * v[i]; }
  const float rstd = rsqrtf(warp_sum(q) * (1.f / 256.f) + eps);
  float g[8], b[8];
  load8<float>(gamma + lane * 8, g);
  load8<float>(beta + lane * 8, b);
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = v[i] * rstd * g[i] + b[i];
  store8(out, (size_t)row * ldo + lane * 8, out_dt, v);
}

// same, reading a residual stream kept in the operand dtype (bf16 engine: cvb200_layernorm_act).  Four rows per warp
// with all four loads issued before the first reduction: the kernel is a pure HBM stream (one 16-byte load per lane
// per row), and one row per warp left too few bytes in flight per SM (13.2 us per launch against 7.3 at HBM speed).
// Small problems (streaming chunks: ~1400 rows) keep one row per warp: there the grid, not the bytes in flight, is short
// (measured 6.8 vs 7.6 us).
template <int LN_ROWS_PER_WARP>
__global__ void __launch_bounds__(256) layernorm256_act_kernel(const void* __restrict__ x, int ldx, int x_dt,
                                                               void* __restrict__ out, int ldo, int out_dt,
                                                               const float* __restrict__ gamma,
                                                               const float* __restrict__ beta, int rows, float eps) {
  pdl_sync();
  const int row0 = (blockIdx.x * 8 + (threadIdx.x >> 5)) * LN_ROWS_PER_WARP;
  const int lane = threadIdx.x & 31;
  if (row0 >= rows) return;
  float v[LN_ROWS_PER_WARP][8];
#pragma unroll
  for (int r = 0; r < LN_ROWS_PER_WARP; ++r) {
    const int row = min(row0 + r, rows - 1);                   // tail rows re-read the last row, never stored
    load8_dt(x, (size_t)row * ldx + lane * 8, x_dt, v[r]);
  }
  float g[8], b[8];
  load8<float>(gamma + lane * 8, g);
  load8<float>(beta + lane * 8, b);
#pragma unroll
  for (int r = 0; r < LN_ROWS_PER_WARP; ++r) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += v[r][i];
    const float mean = warp_sum(s) * (1.f / 256.f);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { v[r][i] -= mean; q += v[r][i] * v[r][i]; }
    const float rstd = rsqrtf(warp_sum(q) * (1.f / 256.f) + eps);
#pragma unroll
    for (int i = 0; i < 8; ++i) v[r][i] = v[r][i] * rstd * g[i] + b[i];
    if (row0 + r < rows) store8(out, (size_t)(row0 + r) * ldo + lane * 8, out_dt, v[r]);
  }
}

// ---------------------------------------------------------------------------- Norm + Mish (Block1D tail)
// Mish for the bf16 engine: tanh(softplus(x)) = (u^2 + 2u) / (u^2 + 2u + 2) with u = e^x -- one MUFU.EX2 and one MUFU.RCP
// instead of expf + log1pf + tanhf (relative error ~1e-6, far below the bf16 rounding of the output it feeds).
__device__ __forceinline__ float mish_fast(float x) {
  const float u = __expf(fminf(x, 20.f));
  const float n = u * (u + 2.f);
  return x > 20.f ? x : x * __fdividef(n, n + 2.f);
}

// ROWS rows per warp (all inside one 128-row tile = one sequence): the GroupNorm statistics are turned into (mean, rstd)
// in fp64 ONCE per warp instead of once per row -- fp64 division and square root run at 1/64 rate on this part, and were
// most of the 30 us this kernel took at 40 960 rows (round-1 launch list) for 42 MB of traffic.
// FASTGN: the bf16 engine's GroupNorm case as its own instantiation (bf16 output, fast Mish): the affine part collapses to
// one FFMA per element (a = rstd * gamma, c = beta - mean * a, per lane, once) and no runtime branch sits in the element
// loop; the generic instantiation carried both Mish variants and both normalisations behind branches in every element.
template <int ROWS, bool FASTGN>
__global__ void __launch_bounds__(256) norm_mish256_kernel(const void* __restrict__ h, int ldh, int h_dt,
                                                           void* __restrict__ out, int ldo, int out_dt,
                                                           const float* __restrict__ gamma, const float* __restrict__ beta,
                                                           const double* __restrict__ gn_stats,
                                                           const float* __restrict__ seq_bias, int seq_bias_ld,
                                                           const Layout lay, float eps) {
  pdl_sync();
  const int row0 = (blockIdx.x * 8 + (threadIdx.x >> 5)) * ROWS;
  const int lane = threadIdx.x & 31;
  if (row0 >= lay.rows) return;
  const int tile = row0 >> 7;
  // the rows are requested before the tile -> sequence -> length lookup (two dependent L2 round trips) has come back: pad
  // rows lie inside the buffer and are discarded below
  float v[ROWS][8];
#pragma unroll
  for (int r = 0; r < ROWS; ++r) load8_dt(h, (size_t)(row0 + r) * ldh + lane * 8, h_dt, v[r]);
  const int b = __ldg(lay.tile_seq + tile);
  const int lrow0 = __ldg(lay.tile_row0 + tile) + (row0 & 127);
  const int len = __ldg(lay.seq_len + b);
  float gmean = 0.f, grstd = 0.f;
  __shared__ float2 s_stat[8];
  if (gn_stats) {
    // GroupNorm(8, 256): group = 32 channels = 4 lanes; statistics over 32 channels x len frames.  The block's rows lie in
    // one 128-row tile = one sequence, so the eight (mean, rstd) pairs are computed by eight threads and shared.
    if (threadIdx.x < 8) {
      // fp64 WITHOUT fp64 division or square root (each expands to ~30 dependent fp64 instructions and the fp64 pipe of
      // this part is narrow: two divisions and a sqrt per warp were a fifth of the kernel's 20 us at 40 960 rows): fp32
      // reciprocal / rsqrt seeds refined by one Newton step in fp64 (relative error ~1e-14).
      const int g = threadIdx.x;
      const double n = (double)len * 32.0;
      const double s = gn_stats[((size_t)b * 8 + g) * 2], q = gn_stats[((size_t)b * 8 + g) * 2 + 1];
      double rn = (double)(1.f / (float)n);
      rn = rn * (2.0 - n * rn);
      const double m = s * rn;
      double var = q * rn - m * m;
      if (var < 0.0) var = 0.0;
      const double ve = var + (double)eps;
      double y = (double)rsqrtf((float)ve);
      y = y * (1.5 - 0.5 * ve * y * y);
      s_stat[g] = make_float2((float)m, (float)y);
    }
    __syncthreads();
    gmean = s_stat[lane >> 2].x;
    grstd = s_stat[lane >> 2].y;
  }
  float g8[8], b8[8], t8[8];
  load8<float>(gamma + lane * 8, g8);
  load8<float>(beta + lane * 8, b8);
  if (seq_bias) load8<float>(seq_bias + (size_t)b * seq_bias_ld + lane * 8, t8);
  const bool fast = out_dt == DT_BF16;
  if constexpr (FASTGN) {
    float a8[8], c8[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      a8[i] = grstd * g8[i];
      c8[i] = fmaf(-gmean, a8[i], b8[i]);
    }
#pragma unroll
    for (int r = 0; r < ROWS; ++r) {
      const bool ok = lrow0 + r < len;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        // Mish without the large-x select: for y > 20 the ratio below is exactly 1 in fp32
        const float y = fmaf(v[r][i], a8[i], c8[i]);
        const float u = __expf(fminf(y, 20.f));
        const float n = u * (u + 2.f);
        float o = y * __fdividef(n, n + 2.f);
        if (seq_bias) o += t8[i];
        v[r][i] = ok ? o : 0.f;
      }
      store8(out, (size_t)(row0 + r) * ldo + lane * 8, out_dt, v[r]);
    }
    return;
  }
#pragma unroll
  for (int r = 0; r < ROWS; ++r) {
    if (lrow0 + r >= len) {                  // pad rows are written as zeros
#pragma unroll
      for (int i = 0; i < 8; ++i) v[r][i] = 0.f;
      store8(out, (size_t)(row0 + r) * ldo + lane * 8, out_dt, v[r]);
      continue;
    }
    float mean = gmean, rstd = grstd;
    if (gn_stats) {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[r][i] -= mean;
    } else {
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) s += v[r][i];
      mean = warp_sum(s) * (1.f / 256.f);
      float q = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) { v[r][i] -= mean; q += v[r][i] * v[r][i]; }
      rstd = rsqrtf(warp_sum(q) * (1.f / 256.f) + eps);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float y = v[r][i] * rstd * g8[i] + b8[i];
      v[r][i] = fast ? mish_fast(y) : mishf(y);
    }
    if (seq_bias) {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[r][i] += t8[i];
    }
    store8(out, (size_t)(row0 + r) * ldo + lane * 8, out_dt, v[r]);
  }
}

// ---------------------------------------------------------------------------- time embedding
__global__ void time_sinusoid_kernel(const float* __restrict__ t, int S, int dim, float scale, float* __restrict__ out) {
  pdl_sync();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int half = dim / 2;
  if (i >= S * half) return;
  const int s = i / half, j = i - s * half;
  const float k = logf(10000.f) / (float)(half - 1);
  const float e = scale * t[s] * expf((float)j * -k);
  out[(size_t)s * dim + j] = sinf(e);
  out[(size_t)s * dim + half + j] = cosf(e);
}

__device__ __forceinline__ float small_act(int act, float v) {
  switch (act) {
    case ACT_MISH: return mishf(v);
    case ACT_SILU: return v / (1.f + expf(-v));
    case ACT_GELU: return gelu_erf(v);
    case ACT_TANH: return tanhf(v);
    default: return v;
  }
}

// one warp per output column n; loops over the S rows.  K % 32 == 0, K <= 1024.
__global__ void __launch_bounds__(256) small_linear_kernel(const float* __restrict__ x, int S, int K,
                                                           const float* __restrict__ W, const float* __restrict__ b,
                                                           int N, int pre_act, int post_act, float* __restrict__ y) {
  pdl_sync();
  const int n = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (n >= N) return;
  float w[32];
  const int per = K / 32;
#pragma unroll
  for (int i = 0; i < 32; ++i) w[i] = i < per ? __ldg(W + (size_t)n * K + i * 32 + lane) : 0.f;
  const float bias = b ? __ldg(b + n) : 0.f;
  for (int s = 0; s < S; ++s) {
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i)
      if (i < per) acc = fmaf(small_act(pre_act, __ldg(x + (size_t)s * K + i * 32 + lane)), w[i], acc);
    acc = warp_sum(acc);
    if (lane == 0) y[(size_t)s * N + n] = small_act(post_act, acc + bias);
  }
}

// ---------------------------------------------------------------------------- CFG input packing
// out row r of the 2B-sequence layout: r < rows1 -> [x | mu | spks[b] | cond], else [x | 0 | 0 | 0]
__global__ void __launch_bounds__(256) pack_cfg_kernel(const float* __restrict__ x, const float* __restrict__ mu,
                                                       const float* __restrict__ cond, const float* __restrict__ spks,
                                                       void* __restrict__ out, int out_dt, const Layout lay) {
  pdl_sync();
  // one thread per (row of the pair layout, 8-channel chunk): 320 / 8 = 40 chunks per row
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = 2LL * lay.rows * 40;
  if (idx >= total) return;
  const int chunk = (int)(idx % 40);
  const long long r2 = idx / 40;
  const bool uncond = r2 >= lay.rows;
  const int r = (int)(uncond ? r2 - lay.rows : r2);
  const int tile = r >> 7;
  const int b = __ldg(lay.tile_seq + tile);
  const int lrow = __ldg(lay.tile_row0 + tile) + (r & 127);
  const bool valid = lrow < __ldg(lay.seq_len + b);
  float v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = 0.f;
  if (valid) {
    const int sect = chunk / 10, c0 = (chunk % 10) * 8;
    if (sect == 0) load8<float>(x + (size_t)r * 80 + c0, v);
    else if (!uncond) {
      if (sect == 1) load8<float>(mu + (size_t)r * 80 + c0, v);
      else if (sect == 2) load8<float>(spks + (size_t)b * 80 + c0, v);
      else load8<float>(cond + (size_t)r * 80 + c0, v);
    }
  }
  store8(out, (size_t)r2 * 320 + chunk * 8, out_dt, v);
}

__global__ void __launch_bounds__(256) euler_cfg_kernel(float* __restrict__ x, const float* __restrict__ v, long long n4,
                                                        float dt, float cfg) {
  pdl_sync();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n4) return;
  float4 xv = reinterpret_cast<float4*>(x)[i];
  const float4 vc = reinterpret_cast<const float4*>(v)[i];
  const float4 vu = reinterpret_cast<const float4*>(v)[n4 + i];
  const float a = 1.f + cfg;
  xv.x += dt * (a * vc.x - cfg * vu.x);
  xv.y += dt * (a * vc.y - cfg * vu.y);
  xv.z += dt * (a * vc.z - cfg * vu.z);
  xv.w += dt * (a * vc.w - cfg * vu.w);
  reinterpret_cast<float4*>(x)[i] = xv;
}

// ---------------------------------------------------------------------------- fp32 -> 3 x bf16 split
// x = x1 + x2 + x3 with x1 = bf16(x), x2 = bf16(x - x1), x3 = bf16(x - x1 - x2): 24 mantissa bits in three bf16
// operands, so that a product with 16-24 bit operand mantissas runs on the bf16 tensor cores (hift.py f0 path).
// out[row, i*cpad + c] = x_i[row, c]; columns c >= C of each block are written as zeros.
__global__ void __launch_bounds__(256) split3_bf16_kernel(const float* __restrict__ x, int ldx, int C, int cpad,
                                                          __nv_bfloat16* __restrict__ out, int ldo, long long total) {
  pdl_sync();
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int c = (int)(idx % cpad);
  const long long row = idx / cpad;
  float v = c < C ? x[row * ldx + c] : 0.f;
  const __nv_bfloat16 h1 = __float2bfloat16_rn(v);
  v -= __bfloat162float(h1);
  const __nv_bfloat16 h2 = __float2bfloat16_rn(v);
  v -= __bfloat162float(h2);
  const __nv_bfloat16 h3 = __float2bfloat16_rn(v);
  __nv_bfloat16* o = out + row * ldo + c;
  o[0] = h1;
  o[cpad] = h2;
  o[2 * cpad] = h3;
}

// ---------------------------------------------------------------------------- token2wav glue
// Cross-fade of a streamed chunk with the tail kept from the previous one (cosyvoice/utils/common.py:149-157, called
// from cli/model.py:150,163,174): x[r, i] = x[r, i] * w[i] + prev[r, prev_len - n + i] * w[n + i] for i < n.  The
// reference multiplies fp32 tensors by a float64 numpy window, so the arithmetic is double and the result is rounded
// to fp32 once; explicit _rn intrinsics keep the compiler from contracting the two products into an FMA.
__global__ void __launch_bounds__(256) fade_in_out_kernel(float* __restrict__ x, long long ldx,
                                                          const float* __restrict__ prev, long long ldp, int prev_len,
                                                          const double* __restrict__ window, int n, long long total) {
  pdl_sync();
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int i = (int)(idx % n);
  const long long r = idx / n;
  const double a = __dmul_rn((double)x[r * ldx + i], window[i]);
  const double b = __dmul_rn((double)prev[r * ldp + prev_len - n + i], window[n + i]);
  x[r * ldx + i] = __double2float_rn(__dadd_rn(a, b));
}

// F.interpolate(mel, size=t_out, mode='linear') with align_corners=False (cli/model.py:171, speed != 1.0):
// src = (j + 0.5) * t_in / t_out - 0.5 clamped at 0, out = (1 - l) * x[i0] + l * x[min(i0 + 1, t_in - 1)], with the
// FMA placement of ATen's CPU kernel so that the result is bit-identical to the reference's (tests/golden/token2wav.pt).
__global__ void __launch_bounds__(256) interp_linear_kernel(const float* __restrict__ x, long long ldx, int t_in,
                                                            float* __restrict__ out, long long ldo, int t_out,
                                                            long long total) {
  pdl_sync();
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int j = (int)(idx % t_out);
  const long long r = idx / t_out;
  const float scale = (float)t_in / (float)t_out;
  float src = __fmaf_rn(scale, (float)j + 0.5f, -0.5f);
  src = src < 0.f ? 0.f : src;
  const int i0 = min((int)src, t_in - 1);
  const int i1 = i0 + (i0 < t_in - 1 ? 1 : 0);
  const float l1 = src - (float)i0, l0 = 1.f - l1;
  out[r * ldo + j] = __fmaf_rn(l0, x[r * ldx + i0], __fmul_rn(l1, x[r * ldx + i1]));
}

// int16 PCM packing of a finished waveform (BigVGAN/utils.py:97-101 save_audio: audio * MAX_WAV_VALUE, .astype("int16")):
// truncation toward zero like the numpy cast; values outside int16 saturate (the numpy cast is undefined there).
__global__ void __launch_bounds__(256) pcm16_kernel(const float* __restrict__ wav, short* __restrict__ out, float scale,
                                                    long long n) {
  pdl_sync();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float v = fminf(fmaxf(__fmul_rn(wav[i], scale), -32768.f), 32767.f);
  out[i] = (short)__float2int_rz(v);
}

// ---------------------------------------------------------------------------- layout transposes
// [B, C, T] (fp32) -> time-major rows.  32x32 smem tile transpose; grid (row tiles of 32, C tiles of 32).
__global__ void __launch_bounds__(256) nct_to_rows_kernel(const float* __restrict__ src, long long seq_stride,
                                                          long long c_stride, int C, void* __restrict__ out, int ldo,
                                                          int out_dt, const Layout lay) {
  pdl_sync();
  __shared__ float tile[32][33];
  const int r0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int t128 = r0 >> 7;
  const int b = __ldg(lay.tile_seq + t128);
  const int l0 = __ldg(lay.tile_row0 + t128) + (r0 & 127);
  const int len = __ldg(lay.seq_len + b);
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int j = ty; j < 32; j += 8) {      // j: channel within tile, tx: time within tile (coalesced along T)
    const int c = c0 + j, l = l0 + tx;
    tile[j][tx] = (c < C && l < len) ? src[(size_t)b * seq_stride + (size_t)c * c_stride + l] : 0.f;
  }
  __syncthreads();
  for (int j = ty; j < 32; j += 8) {      // j: time, tx: channel (coalesced along C)
    const int c = c0 + tx;
    if (c < C) st_from_float(out, (size_t)(r0 + j) * ldo + c, out_dt, tile[tx][j]);
  }
}

__global__ void __launch_bounds__(256) rows_to_nct_kernel(const float* __restrict__ src, int lds, int C,
                                                          float* __restrict__ out, long long seq_stride,
                                                          long long c_stride, const Layout lay) {
  pdl_sync();
  __shared__ float tile[32][33];
  const int r0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int t128 = r0 >> 7;
  const int b = __ldg(lay.tile_seq + t128);
  const int l0 = __ldg(lay.tile_row0 + t128) + (r0 & 127);
  const int len = __ldg(lay.seq_len + b);
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int j = ty; j < 32; j += 8) {      // j: time, tx: channel
    const int c = c0 + tx;
    tile[j][tx] = c < C ? src[(size_t)(r0 + j) * lds + c] : 0.f;
  }
  __syncthreads();
  for (int j = ty; j < 32; j += 8) {      // j: channel, tx: time
    const int c = c0 + j, l = l0 + tx;
    if (c < C && l < len) out[(size_t)b * seq_stride + (size_t)c * c_stride + l] = tile[tx][j];
  }
}

// ---------------------------------------------------------------------------- launchers
int launch_layernorm(const float* x, int ldx, void* out, int ldo, int out_dt, const float* gamma, const float* beta,
                     int rows, int C, float eps, cudaStream_t st) {
  CVB_REQUIRE(C == 256, "LayerNorm kernel is specialised for 256 channels");
  CVB_REQUIRE(ldx % 4 == 0 && ldo % 8 == 0, "leading dimensions must keep 16-byte alignment");
  if (rows <= 0) return CVB_OK;
  CVB_CUDA(launch_pdl(layernorm256_kernel, dim3(ceil_div(rows, 8)), dim3(256), 0, st, x, ldx, out, ldo, out_dt, gamma, beta, rows, eps));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_layernorm_act(const void* x, int ldx, int x_dt, void* out, int ldo, int out_dt, const float* gamma,
                         const float* beta, int rows, int C, float eps, cudaStream_t st) {
  CVB_REQUIRE(C == 256, "LayerNorm kernel is specialised for 256 channels");
  CVB_REQUIRE(ldx % 8 == 0 && ldo % 8 == 0, "leading dimensions must keep 16-byte alignment");
  if (rows <= 0) return CVB_OK;
  if (rows >= 16384)
    CVB_CUDA(launch_pdl(layernorm256_act_kernel<4>, dim3(ceil_div(rows, 32)), dim3(256), 0, st, x, ldx, x_dt, out, ldo, out_dt, gamma, beta, rows, eps));
  else
    CVB_CUDA(launch_pdl(layernorm256_act_kernel<1>, dim3(ceil_div(rows, 8)), dim3(256), 0, st, x, ldx, x_dt, out, ldo, out_dt, gamma, beta, rows, eps));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_norm_mish(const void* h, int ldh, int h_dt, void* out, int ldo, int out_dt, const float* gamma,
                     const float* beta, const double* gn_stats, const float* seq_bias, int seq_bias_ld,
                     const Layout& lay, int C, float eps, cudaStream_t st) {
  CVB_REQUIRE(C == 256, "Norm+Mish kernel is specialised for 256 channels");
  CVB_REQUIRE(ldh % 8 == 0 && ldo % 8 == 0, "leading dimensions must keep 16-byte alignment");
  if (lay.rows <= 0) return CVB_OK;
  const bool fastgn = gn_stats != nullptr && out_dt == DT_BF16;
#define CVB_NM(R_, F_)                                                                                                         \
  CVB_CUDA(launch_pdl(norm_mish256_kernel<R_, F_>, dim3(ceil_div(lay.rows, 8 * R_)), dim3(256), 0, st, h, ldh, h_dt, out, ldo, \
                      out_dt, gamma, beta, gn_stats, seq_bias, seq_bias_ld, lay, eps))
  if (lay.rows >= 16384) {
    if (fastgn) CVB_NM(4, true);
    else CVB_NM(4, false);
  } else {
    if (fastgn) CVB_NM(1, true);
    else CVB_NM(1, false);
  }
#undef CVB_NM
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_time_sinusoid(const float* t, int S, int dim, float scale, float* out, cudaStream_t st) {
  CVB_REQUIRE(dim % 2 == 0 && dim >= 4, "dim must be even");
  if (S <= 0) return CVB_OK;
  const int n = S * dim / 2;
  CVB_CUDA(launch_pdl(time_sinusoid_kernel, dim3(ceil_div(n, 128)), dim3(128), 0, st, t, S, dim, scale, out));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_small_linear(const float* x, int S, int K, const float* W, const float* b, int N, int pre_act, int post_act,
                        float* y, cudaStream_t st) {
  CVB_REQUIRE(K % 32 == 0 && K <= 1024, "K must be a multiple of 32 and <= 1024");
  if (S <= 0 || N <= 0) return CVB_OK;
  CVB_CUDA(launch_pdl(small_linear_kernel, dim3(ceil_div(N, 8)), dim3(256), 0, st, x, S, K, W, b, N, pre_act, post_act, y));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_pack_cfg(const float* x, const float* mu, const float* cond, const float* spks, void* out, int out_dt,
                    const Layout& lay, cudaStream_t st) {
  const long long total = 2LL * lay.rows * 40;
  if (total <= 0) return CVB_OK;
  CVB_CUDA(launch_pdl(pack_cfg_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, st, x, mu, cond, spks, out, out_dt, lay));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_euler_cfg(float* x, const float* v, long long n_half, float dt, float cfg, cudaStream_t st) {
  CVB_REQUIRE(n_half % 4 == 0, "element count must be a multiple of 4");
  const long long n4 = n_half / 4;
  if (n4 <= 0) return CVB_OK;
  CVB_CUDA(launch_pdl(euler_cfg_kernel, dim3((unsigned)((n4 + 255) / 256)), dim3(256), 0, st, x, v, n4, dt, cfg));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_split3_bf16(const float* x, int ldx, int C, int cpad, void* out, int ldo, long long rows, cudaStream_t st) {
  CVB_REQUIRE(x && out && C > 0 && cpad >= C && ldo >= 3 * cpad, "split3: bad shapes");
  const long long total = rows * cpad;
  if (total <= 0) return CVB_OK;
  CVB_CUDA(launch_pdl(split3_bf16_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, st, x, ldx, C, cpad,
                      reinterpret_cast<__nv_bfloat16*>(out), ldo, total));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_fade_in_out(float* x, long long ldx, int x_len, const float* prev, long long ldp, int prev_len,
                       const double* window, int window_len, int rows, cudaStream_t st) {
  const int n = window_len / 2;
  CVB_REQUIRE(x && prev && window && rows >= 0 && window_len >= 0, "fade_in_out: bad arguments");
  // the reference's slice assignment raises on a shape mismatch when either side is shorter than the overlap
  CVB_REQUIRE(x_len >= n && prev_len >= n, "fade_in_out: chunk or kept tail shorter than the overlap");
  const long long total = (long long)rows * n;
  if (total <= 0) return CVB_OK;
  CVB_CUDA(launch_pdl(fade_in_out_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, st, x, ldx, prev, ldp,
                      prev_len, window, n, total));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_interp_linear(const float* x, long long ldx, int t_in, float* out, long long ldo, int t_out, int rows,
                         cudaStream_t st) {
  CVB_REQUIRE(x && out && t_in > 0 && t_out >= 0 && rows >= 0, "interp_linear: bad arguments");
  const long long total = (long long)rows * t_out;
  if (total <= 0) return CVB_OK;
  CVB_CUDA(launch_pdl(interp_linear_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, st, x, ldx, t_in, out,
                      ldo, t_out, total));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_pcm16(const float* wav, short* out, float scale, long long n, cudaStream_t st) {
  if (n == 0) return CVB_OK;
  CVB_REQUIRE(wav && out && n > 0, "pcm16: bad arguments");
  CVB_CUDA(launch_pdl(pcm16_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, st, wav, out, scale, n));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_nct_to_rows(const float* src, long long seq_stride, long long c_stride, int C, void* out, int ldo, int out_dt,
                       const Layout& lay, cudaStream_t st) {
  if (lay.rows <= 0) return CVB_OK;
  dim3 grid(lay.rows / 32, ceil_div(C, 32));
  CVB_CUDA(launch_pdl(nct_to_rows_kernel, dim3(grid), dim3(256), 0, st, src, seq_stride, c_stride, C, out, ldo, out_dt, lay));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_rows_to_nct(const float* src, int lds, int C, float* out, long long seq_stride, long long c_stride,
                       const Layout& lay, cudaStream_t st) {
  if (lay.rows <= 0) return CVB_OK;
  dim3 grid(lay.rows / 32, ceil_div(C, 32));
  CVB_CUDA(launch_pdl(rows_to_nct_kernel, dim3(grid), dim3(256), 0, st, src, lds, C, out, seq_stride, c_stride, lay));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

}  // namespace cvb
