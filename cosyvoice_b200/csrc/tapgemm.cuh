// "Tap GEMM": the one contraction kernel interface of the library.
//
//   D[m, phase, n] = epilogue( sum_{seg} sum_{c < seg_len} A[m + seg_row[phase][seg], seg_col[phase][seg] + c]
//                                                          * W[phase][n][seg * seg_len + c] )
//
// A is a row-major 2-D *view* of a time-major activation buffer ([rows, lda]); every Linear, Conv1d
// (any k / dilation / causal pad), strided Conv1d (view [rows/s, s*C]) and ConvTranspose1d (one
// phase per output residue, output view [rows, s*Cout]) of the reference's hot path is an instance:
//   cosyvoice/flow/components/decoder.py:34,52,65,142   cosyvoice/flow/decoder.py:83-84,205
//   cosyvoice/flow/components/transformer.py:110,126 + diffusers Attention to_q/k/v/out
//   cosyvoice/hifigan/generator.py:58-78,268-312        cosyvoice/hifigan/f0_predictor.py:28-50
//   cosyvoice/BigVGAN/bigvgan.py:61-85,303-377
// Rows are grouped in tiles of 128; a tile belongs to exactly one sequence (utterance), rows past
// the sequence's valid length are written as zeros (they are the conv zero padding of the next op).
#pragma once
#include "common.cuh"
#include "../../include/cvb200.h"

namespace cvb {

constexpr int MAX_SEGS = CVB200_MAX_SEGS;   // n_phase * n_seg <= MAX_SEGS
constexpr int GEMM_BK = 64;    // K chunk (elements) of the tcgen05 engine; seg_len must be a multiple

enum Act : int { ACT_NONE = 0, ACT_GELU = 1, ACT_ELU = 2, ACT_SNAKE = 3, ACT_LRELU = 4, ACT_MISH = 5, ACT_TANH = 6, ACT_SILU = 7 };
enum ResKind : int { RES_NONE = 0, RES_F32 = 1, RES_ACT = 2 };
enum Out1Mode : int { OUT1_NONE = 0, OUT1_COPY = 1, OUT1_SNAKE = 2, OUT1_LRELU = 3 };

using Epi = cvb200_epilogue;       // include/cvb200.h
using GemmParams = cvb200_gemm;

int launch_tapgemm_simt(const GemmParams& p, cudaStream_t stream);
int launch_tapgemm_tc(const GemmParams& p, cudaStream_t stream);   // dt must be DT_BF16

// ------------------------------------------------------------------------------------ epilogue
// Layout contract (checked by the launchers): every vector the epilogue touches (bias, seq_bias rows,
// alphas, res, out0, out1) is 16-byte aligned at every chunk start and covers n_pad (>= N, multiple of
// 16) columns, so each NV-column chunk is handled with 128-bit accesses and no per-element bounds test.
// All branches below are warp-uniform and sit OUTSIDE the per-element loops: the first version of this
// epilogue branched per element and its unrolled code overflowed the instruction cache
// (profiles/r01_tapgemm_v0.md: stall_no_inst dominant, 3 % of tensor peak).
static __device__ __noinline__ float rare_act(int act, float v, float slope) {
  switch (act) {
    case ACT_ELU: return eluf(v);
    case ACT_LRELU: return v > 0.f ? v : v * slope;
    case ACT_MISH: return mishf(v);
    case ACT_TANH: return tanhf(v);
    case ACT_SILU: return siluf(v);
    default: return v;
  }
}

template <int NV> __device__ __forceinline__ void vec_load_f32(const float* __restrict__ p, float (&o)[NV]) {
#pragma unroll
  for (int j = 0; j < NV; j += 4) {
    const float4 t = *reinterpret_cast<const float4*>(p + j);
    o[j] = t.x; o[j + 1] = t.y; o[j + 2] = t.z; o[j + 3] = t.w;
  }
}
template <int NV> __device__ __forceinline__ void vec_add_f32(const float* p, float (&v)[NV]) {
#pragma unroll
  for (int j = 0; j < NV; j += 4) {
    const float4 t = *reinterpret_cast<const float4*>(p + j);
    v[j] += t.x; v[j + 1] += t.y; v[j + 2] += t.z; v[j + 3] += t.w;
  }
}
// 256-bit global load (sm_100+): a whole 32-byte sector per thread per instruction, so that the residual read of the
// epilogue (each thread owns 64-128 contiguous bytes of a row) does not depend on L1 to merge half-sector accesses
// (the GEMM CTA leaves ~25 KB of L1).  p must be 32-byte aligned.
template <int NV> __device__ __forceinline__ void vec_add_f32_wide(const float* p, float (&v)[NV]) {
  static_assert(NV % 8 == 0, "wide loads take 8 floats at a time");
#pragma unroll
  for (int j = 0; j < NV; j += 8) {
    float t[8];
    asm volatile("ld.global.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=f"(t[0]), "=f"(t[1]), "=f"(t[2]), "=f"(t[3]), "=f"(t[4]), "=f"(t[5]), "=f"(t[6]), "=f"(t[7])
                 : "l"(p + j));
#pragma unroll
    for (int i = 0; i < 8; ++i) v[j + i] += t[i];
  }
}
template <int NV> __device__ __forceinline__ void vec_add_bf16(const __nv_bfloat16* p, float (&v)[NV]) {
  if constexpr (NV % 8 == 0) {
#pragma unroll
    for (int j = 0; j < NV; j += 8) {
      const uint4 u = *reinterpret_cast<const uint4*>(p + j);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 f = __bfloat1622float2(h[i]);
        v[j + 2 * i] += f.x; v[j + 2 * i + 1] += f.y;
      }
    }
  } else {
#pragma unroll
    for (int j = 0; j < NV; j += 4) {
      const uint2 u = *reinterpret_cast<const uint2*>(p + j);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
      const float2 f0 = __bfloat1622float2(h[0]), f1 = __bfloat1622float2(h[1]);
      v[j] += f0.x; v[j + 1] += f0.y; v[j + 2] += f1.x; v[j + 3] += f1.y;
    }
  }
}
template <int NV> __device__ __forceinline__ void vec_add_dt(const void* base, size_t idx, int dt, float (&v)[NV]) {
  if (dt == DT_BF16) vec_add_bf16<NV>(reinterpret_cast<const __nv_bfloat16*>(base) + idx, v);
  else vec_add_f32<NV>(reinterpret_cast<const float*>(base) + idx, v);
}
// 256-bit global store (sm_100+): one full 32-byte sector per thread per instruction
__device__ __forceinline__ void st_global_v8(void* p, const uint32_t (&r)[8]) {
#ifdef CVB_EXPERIMENT_NO_STORE
  if (r[0] != 0x7fc12345u) return;       // experiment: how much of the kernel is the store path
#endif
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(r[0]), "r"(r[1]), "r"(r[2]),
               "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}

template <int NV> __device__ __forceinline__ void vec_store_dt(void* base, size_t idx, int dt, const float (&v)[NV]) {
  if constexpr (NV % 16 == 0) {
    // chunks start at multiples of 16 columns of 64-element-aligned rows: 32-byte aligned in both dtypes
    if (dt == DT_BF16) {
      __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(base) + idx;
#pragma unroll
      for (int j = 0; j < NV; j += 16) {
        uint32_t r[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          __nv_bfloat162 t = __floats2bfloat162_rn(v[j + 2 * i], v[j + 2 * i + 1]);
          r[i] = *reinterpret_cast<uint32_t*>(&t);
        }
        st_global_v8(o + j, r);
      }
    } else {
      float* o = reinterpret_cast<float*>(base) + idx;
#pragma unroll
      for (int j = 0; j < NV; j += 8) {
        uint32_t r[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) r[i] = __float_as_uint(v[j + i]);
        st_global_v8(o + j, r);
      }
    }
    return;
  }
  if (dt == DT_BF16) {
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(base) + idx;
    if constexpr (NV % 8 == 0) {
#pragma unroll
      for (int j = 0; j < NV; j += 8) {
        uint4 pk;
        __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
        for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[j + 2 * i], v[j + 2 * i + 1]);
        *reinterpret_cast<uint4*>(o + j) = pk;
      }
    } else {
#pragma unroll
      for (int j = 0; j < NV; j += 4) {
        uint2 pk;
        __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
        h[0] = __floats2bfloat162_rn(v[j], v[j + 1]);
        h[1] = __floats2bfloat162_rn(v[j + 2], v[j + 3]);
        *reinterpret_cast<uint2*>(o + j) = pk;
      }
    }
  } else {
    float* o = reinterpret_cast<float*>(base) + idx;
#pragma unroll
    for (int j = 0; j < NV; j += 4) *reinterpret_cast<float4*>(o + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
  }
}

// erf-GELU for the bf16 engine: x Phi(x) with Phi(x) = (1 + tanh(x (a + b x^2 + c x^4))) / 2, coefficients fitted to
// the exact erf form on [-8, 8] (max abs error 2.5e-5, below the 4.9e-4 relative error of MUFU.TANH and far below
// the bf16 rounding of the FFN hidden activations it feeds).  1 MUFU + 6 FMA-pipe instructions; the FFN-in
// epilogue is instruction-bound (ncu: 27 -> 15 instructions per output).  The fp32 parity engine uses erff.
__device__ __forceinline__ float gelu_erf_fast(float v) {
  const float xc = fminf(fmaxf(v, -6.f), 6.f);        // the fitted odd polynomial turns over beyond |x| ~ 10
  const float x2 = xc * xc;
  float p = fmaf(-3.51516916e-4f, x2, 3.70056467e-2f);
  p = fmaf(p, x2, 7.97507884e-1f);
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(xc * p));
  const float h = 0.5f * v;
  return fmaf(h, t, h);
}

template <int NV, bool PRECISE>
__device__ __forceinline__ void snake_vec(const float* __restrict__ alpha, const float (&x)[NV], float (&y)[NV]) {
  float al[NV];
  vec_load_f32<NV>(alpha, al);
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const float s = PRECISE ? sinf(x[j] * al[j]) : __sinf(x[j] * al[j]);
    y[j] = x[j] + s * s * (PRECISE ? 1.f / (al[j] + 1e-9f) : __fdividef(1.f, al[j] + 1e-9f));
  }
}

// Applies the epilogue to the NV consecutive columns [n0, n0+NV) of output row m (phase `phase`) and stores them.
//   b: sequence of the row, valid: row (and phase) inside the sequence's valid length.
// GroupNorm partial sums of this chunk are ADDED to (gsum, gsq).
// RP: the row-partials variant (row_parts / row_parts_out of include/cvb200.h).  A separate instantiation of the tcgen05
// kernel, so that the common GEMMs do not carry its code (the epilogue sits close to the instruction-cache limit).
template <int NV, bool PRECISE, bool RP = false>
__device__ __forceinline__ void epi_apply(const Epi& e, size_t m, int b, bool valid, int phase, int n0, float (&v)[NV],
                                          float& gsum, float& gsq) {
  static_assert(NV % 4 == 0, "chunk must be a multiple of 4 columns");
  if (RP && e.row_parts) {             // folded LayerNorm from the producer's partial sums (8 x (sum, sum of squares) per row)
    const float4* pp = reinterpret_cast<const float4*>(e.row_parts + m * 16);
    float sx = 0.f, sq = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float4 t = __ldg(pp + i);
      sx += t.x + t.z;
      sq += t.y + t.w;
    }
    const float mean = sx * (1.f / 256.f);
    const float rstd = rsqrtf(fmaxf(sq * (1.f / 256.f) - mean * mean, 0.f) + e.row_eps);
    const float nb = -mean * rstd;
    float u[NV];
    vec_load_f32<NV>(e.row_u + n0, u);
#pragma unroll
    for (int j = 0; j < NV; ++j) v[j] = fmaf(v[j], rstd, nb * u[j]);
  } else if (e.row_stats) {                   // folded LayerNorm: per-row scale and rank-1 mean correction (include/cvb200.h)
    const float2 rs = __ldg(reinterpret_cast<const float2*>(e.row_stats) + m);
    float u[NV];
    vec_load_f32<NV>(e.row_u + n0, u);
#pragma unroll
    for (int j = 0; j < NV; ++j) v[j] = fmaf(v[j], rs.x, rs.y * u[j]);
  }
  if (e.bias) vec_add_f32<NV>(e.bias + n0, v);
  if (e.seq_bias) vec_add_f32<NV>(e.seq_bias + (size_t)b * e.seq_bias_ld + n0, v);
  if (e.act != ACT_NONE) {
    if (e.act == ACT_GELU) {
#pragma unroll
      for (int j = 0; j < NV; ++j) v[j] = PRECISE ? gelu_erf(v[j]) : gelu_erf_fast(v[j]);
    } else if (e.act == ACT_SNAKE) {
      snake_vec<NV, PRECISE>(e.act_alpha + n0, v, v);
    } else {
#pragma unroll
      for (int j = 0; j < NV; ++j) v[j] = rare_act(e.act, v[j], e.act_slope);
    }
  }
  const size_t col = (size_t)phase * e.phase_col + n0;
  if (e.res_kind == RES_F32) {
    if constexpr (NV % 16 == 0) vec_add_f32_wide<NV>(reinterpret_cast<const float*>(e.res) + m * e.ldr + col, v);
    else vec_add_f32<NV>(reinterpret_cast<const float*>(e.res) + m * e.ldr + col, v);
  }
  else if (e.res_kind == RES_ACT) vec_add_dt<NV>(e.res, m * e.ldr + col, e.res_dt, v);
  if (e.out_scale != 1.f) {
#pragma unroll
    for (int j = 0; j < NV; ++j) v[j] *= e.out_scale;
  }
  if (e.accumulate) vec_add_dt<NV>(e.out0, m * e.ldo0 + col, e.out0_dt, v);
  if (!valid) {
#pragma unroll
    for (int j = 0; j < NV; ++j) v[j] = 0.f;
  }
  if (e.gn_stats) {
#pragma unroll
    for (int j = 0; j < NV; ++j) { gsum += v[j]; gsq = fmaf(v[j], v[j], gsq); }
  }
  if (RP && e.row_parts_out) {         // statistics of the values AS STORED (what the consumer's MMA will read)
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const float r = e.out0_dt == DT_BF16 ? __bfloat162float(__float2bfloat16(v[j])) : v[j];
      gsum += r;
      gsq = fmaf(r, r, gsq);
    }
  }
  vec_store_dt<NV>(e.out0, m * e.ldo0 + col, e.out0_dt, v);
  if (e.out1_mode != OUT1_NONE) {
    float y[NV];
    if (e.out1_mode == OUT1_SNAKE) {
      snake_vec<NV, PRECISE>(e.out1_alpha + n0, v, y);
    } else if (e.out1_mode == OUT1_LRELU) {
#pragma unroll
      for (int j = 0; j < NV; ++j) y[j] = v[j] > 0.f ? v[j] : v[j] * e.out1_slope;
    } else {
#pragma unroll
      for (int j = 0; j < NV; ++j) y[j] = v[j];
    }
    vec_store_dt<NV>(e.out1, m * e.ldo1 + col, e.out1_dt, y);
  }
}

// host-side check of the alignment contract above
inline int check_epilogue_layout(const GemmParams& p) {
  const Epi& e = p.epi;
  auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  CVB_REQUIRE(p.n_pad % 16 == 0 && p.N <= p.n_pad, "n_pad must be a multiple of 16 covering N");
  auto al32 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 31) == 0; };
  CVB_REQUIRE(e.out0 && al32(e.out0), "out0 must be 32-byte aligned");
  const int s0 = dtype_size(e.out0_dt);
  CVB_REQUIRE((e.ldo0 * s0) % 32 == 0 && (e.phase_col * s0) % 32 == 0, "out0 pitch must keep 32-byte alignment");
  if (e.out1_mode != OUT1_NONE) {
    const int s1 = dtype_size(e.out1_dt);
    CVB_REQUIRE(e.out1 && al32(e.out1) && (e.ldo1 * s1) % 32 == 0 && (e.phase_col * s1) % 32 == 0, "out1 alignment");
    CVB_REQUIRE(e.out1_mode != OUT1_SNAKE || (e.out1_alpha && al16(e.out1_alpha)), "out1_alpha alignment");
  }
  if (e.res_kind != RES_NONE) {
    const int sr = e.res_kind == RES_F32 ? 4 : dtype_size(e.res_dt);
    CVB_REQUIRE(e.res && al16(e.res) && (e.ldr * sr) % 16 == 0 && (e.phase_col * sr) % 16 == 0, "res alignment");
    CVB_REQUIRE(e.res_kind != RES_F32 || (al32(e.res) && (e.ldr * 4) % 32 == 0 && (e.phase_col * 4) % 32 == 0),
                "fp32 residual must keep 32-byte alignment");
  }
  CVB_REQUIRE(!e.bias || al16(e.bias), "bias must be 16-byte aligned (and padded to n_pad)");
  CVB_REQUIRE(!e.seq_bias || (al16(e.seq_bias) && e.seq_bias_ld % 4 == 0), "seq_bias alignment");
  CVB_REQUIRE(e.act != ACT_SNAKE || (e.act_alpha && al16(e.act_alpha)), "act_alpha alignment");
  CVB_REQUIRE(!e.gn_stats || (p.N % 32 == 0 && e.gn_groups == p.N / 32), "GroupNorm statistics need N % 32 == 0");
  CVB_REQUIRE(!e.row_stats || (e.row_u && al16(e.row_u) && (reinterpret_cast<uintptr_t>(e.row_stats) & 7) == 0 && p.n_phase == 1),
              "row_stats needs row_u (16-byte aligned, n_pad long), 8-byte aligned statistics and a single phase");
  return CVB_OK;
}

}  // namespace cvb
