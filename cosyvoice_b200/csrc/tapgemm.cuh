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
__device__ __forceinline__ float apply_act(int act, float v, float alpha, float slope) {
  switch (act) {
    case ACT_GELU: return gelu_erf(v);
    case ACT_ELU: return eluf(v);
    case ACT_SNAKE: { float s = __sinf(v * alpha); return v + s * s / (alpha + 1e-9f); }
    case ACT_LRELU: return v > 0.f ? v : v * slope;
    case ACT_MISH: return mishf(v);
    case ACT_TANH: return tanhf(v);
    case ACT_SILU: return siluf(v);
    default: return v;
  }
}

// precise snake (sinf) for the fp32 parity path
__device__ __forceinline__ float snake_precise(float v, float alpha) {
  float s = sinf(v * alpha);
  return v + s * s / (alpha + 1e-9f);
}

// Applies the epilogue to NV consecutive columns [n0, n0+NV) of one output row.
//   m: global view row, b: sequence, valid: row within valid length (for this phase)
//   precise: use accurate transcendental functions (fp32 parity mode)
// Accumulates GroupNorm partial sums for this row into (gsum, gsq) -- all NV columns must lie in
// one 32-channel group (NV <= 32, n0 % NV == 0).
template <int NV>
__device__ __forceinline__ void epi_apply(const Epi& e, int m, int b, bool valid, int phase, int n0, int N,
                                          float (&v)[NV], bool precise, float& gsum, float& gsq) {
  const size_t col0 = (size_t)phase * e.phase_col + n0;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int n = n0 + j;
    float x = v[j];
    if (n < N && valid) {
      if (e.bias) x += __ldg(e.bias + n);
      if (e.seq_bias) x += __ldg(e.seq_bias + (size_t)b * e.seq_bias_ld + n);
      if (e.act != ACT_NONE) {
        const float al = e.act == ACT_SNAKE ? __ldg(e.act_alpha + n) : 0.f;
        x = (precise && e.act == ACT_SNAKE) ? snake_precise(x, al) : apply_act(e.act, x, al, e.act_slope);
      }
      if (e.res_kind == RES_F32) x += reinterpret_cast<const float*>(e.res)[(size_t)m * e.ldr + col0 + j];
      else if (e.res_kind == RES_ACT) x += ld_as_float(e.res, (size_t)m * e.ldr + col0 + j, e.res_dt);
      x *= e.out_scale;
      if (e.accumulate) x += ld_as_float(e.out0, (size_t)m * e.ldo0 + col0 + j, e.out0_dt);
      gsum += x;
      gsq += x * x;
    } else {
      x = 0.f;
    }
    v[j] = x;
  }
  // ---- stores (vectorised when the full NV chunk is in range and aligned)
  const bool full = (n0 + NV <= N);
  if (e.out0) {
    if (e.out0_dt == DT_BF16) {
      __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(e.out0) + (size_t)m * e.ldo0 + col0;
      if (full && NV % 8 == 0 && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
#pragma unroll
        for (int j = 0; j < NV; j += 8) {
          uint4 pk;
          __nv_bfloat162 t0 = __floats2bfloat162_rn(v[j], v[j + 1]), t1 = __floats2bfloat162_rn(v[j + 2], v[j + 3]);
          __nv_bfloat162 t2 = __floats2bfloat162_rn(v[j + 4], v[j + 5]), t3 = __floats2bfloat162_rn(v[j + 6], v[j + 7]);
          pk.x = *reinterpret_cast<uint32_t*>(&t0); pk.y = *reinterpret_cast<uint32_t*>(&t1);
          pk.z = *reinterpret_cast<uint32_t*>(&t2); pk.w = *reinterpret_cast<uint32_t*>(&t3);
          *reinterpret_cast<uint4*>(o + j) = pk;
        }
      } else {
#pragma unroll
        for (int j = 0; j < NV; ++j) if (n0 + j < N) o[j] = __float2bfloat16_rn(v[j]);
      }
    } else {
      float* o = reinterpret_cast<float*>(e.out0) + (size_t)m * e.ldo0 + col0;
      if (full && NV % 4 == 0 && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
#pragma unroll
        for (int j = 0; j < NV; j += 4) *reinterpret_cast<float4*>(o + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
      } else {
#pragma unroll
        for (int j = 0; j < NV; ++j) if (n0 + j < N) o[j] = v[j];
      }
    }
  }
  if (e.out1_mode != OUT1_NONE) {
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int n = n0 + j;
      if (n < N) {
        float y = v[j];
        if (valid) {
          if (e.out1_mode == OUT1_SNAKE) {
            const float al = __ldg(e.out1_alpha + n);
            y = precise ? snake_precise(y, al) : apply_act(ACT_SNAKE, y, al, 0.f);
          } else if (e.out1_mode == OUT1_LRELU) {
            y = y > 0.f ? y : y * e.out1_slope;
          }
        }
        st_from_float(e.out1, (size_t)m * e.ldo1 + col0 + j, e.out1_dt, y);
      }
    }
  }
}

}  // namespace cvb
