// Flow front-end kernels (everything flow.inference does before the CFM): token embedding gather,
// LayerNorm(512) over a varlen layout, ESPnet relative-position self-attention (fp32 flash kernel with the chunk mask of
// utils/mask.py), GroupNorm(1, C) + Mish of the length regulator, its piecewise-linear interpolation, and the speaker
// vector normalisation.  The Linear / Conv1d layers of the encoders and the regulator run on the tap GEMM (tapgemm_*.cu).
// The front-end is < 1 % of the flow FLOPs (it runs once per call at token rate); these kernels are sized for latency and
// launch count, the tensor-core budget stays with the estimator.
#include "common.cuh"
#include "tapgemm.cuh"

namespace cvb {

using Layout = cvb200_layout;

__device__ __forceinline__ void seq_of_row(const Layout& lay, int row, int& b, int& lrow, int& len) {
  const int tile = row >> 7;
  b = __ldg(lay.tile_seq + tile);
  lrow = __ldg(lay.tile_row0 + tile) + (row & 127);
  len = __ldg(lay.seq_len + b);
}

// ---------------------------------------------------------------------------- token embedding
// out[row] = table[clamp(token)] on valid rows, 0 elsewhere  (flow/flow.py:131-135: input_embedding(clamp(token, min=0)) * mask).
// One thread per 4 channels.
__global__ void __launch_bounds__(256) embed_tokens_kernel(const int* __restrict__ tokens, int tok_ld,
                                                           const float* __restrict__ table, int vocab, int C,
                                                           void* __restrict__ out, int ldo, int out_dt, const Layout lay) {
  pdl_sync();
  const int per_row = C >> 2;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)lay.rows * per_row) return;
  const int row = (int)(idx / per_row), c = (int)(idx % per_row) * 4;
  int b, lrow, len;
  seq_of_row(lay, row, b, lrow, len);
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (lrow < len) {
    int tok = __ldg(tokens + (size_t)b * tok_ld + lrow);
    tok = min(max(tok, 0), vocab - 1);
    v = *reinterpret_cast<const float4*>(table + (size_t)tok * C + c);
  }
  if (out_dt == DT_BF16) {
    __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(reinterpret_cast<__nv_bfloat16*>(out) + (size_t)row * ldo + c);
    o[0] = __floats2bfloat162_rn(v.x, v.y);
    o[1] = __floats2bfloat162_rn(v.z, v.w);
  } else {
    *reinterpret_cast<float4*>(reinterpret_cast<float*>(out) + (size_t)row * ldo + c) = v;
  }
}

// ---------------------------------------------------------------------------- LayerNorm over a varlen layout
// One warp per row, C = 128 * NV channels (lane owns NV groups of 4 consecutive channels); invalid rows are written as
// zeros so that a following convolution sees the reference's zero padding (transformer/subsampling.py:83-86,
// encoder_layer.py:145-146, encoder.py:104).  Two-pass variance like ATen's.
template <int NV>
__global__ void __launch_bounds__(256) layernorm_seq_kernel(const float* __restrict__ x, int ldx, void* __restrict__ out,
                                                            int ldo, int out_dt, const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, const Layout lay, float eps) {
  pdl_sync();
  constexpr int C = 128 * NV;
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= lay.rows) return;
  int b, lrow, len;
  seq_of_row(lay, row, b, lrow, len);
  float v[NV][4];
  const bool valid = lrow < len;
  if (valid) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const float4 a = *reinterpret_cast<const float4*>(x + (size_t)row * ldx + i * 128 + lane * 4);
      v[i][0] = a.x; v[i][1] = a.y; v[i][2] = a.z; v[i][3] = a.w;
      s += (a.x + a.y) + (a.z + a.w);
    }
    const float mean = warp_sum(s) * (1.f / C);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) { v[i][j] -= mean; q += v[i][j] * v[i][j]; }
    const float rstd = rsqrtf(warp_sum(q) * (1.f / C) + eps);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const float4 g = *reinterpret_cast<const float4*>(gamma + i * 128 + lane * 4);
      const float4 be = *reinterpret_cast<const float4*>(beta + i * 128 + lane * 4);
      v[i][0] = v[i][0] * rstd * g.x + be.x; v[i][1] = v[i][1] * rstd * g.y + be.y;
      v[i][2] = v[i][2] * rstd * g.z + be.z; v[i][3] = v[i][3] * rstd * g.w + be.w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < NV; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) v[i][j] = 0.f;
  }
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const size_t o = (size_t)row * ldo + i * 128 + lane * 4;
    if (out_dt == DT_BF16) {
      __nv_bfloat162* p = reinterpret_cast<__nv_bfloat162*>(reinterpret_cast<__nv_bfloat16*>(out) + o);
      p[0] = __floats2bfloat162_rn(v[i][0], v[i][1]);
      p[1] = __floats2bfloat162_rn(v[i][2], v[i][3]);
    } else {
      *reinterpret_cast<float4*>(reinterpret_cast<float*>(out) + o) = make_float4(v[i][0], v[i][1], v[i][2], v[i][3]);
    }
  }
}

// ---------------------------------------------------------------------------- relative-position self-attention
// RelPositionMultiHeadedAttention (transformer/attention.py:243-330) for one (64-query tile, head, sequence) per CTA:
//   s[i,j] = ((q_i + u_h) . k_j + (q_i + v_h) . P[i - j]) * scale,  P[r] = linear_pos(PE(r)) (rel_shift is the index i - j),
//   key j visible iff j < len and (chunk == 0 or j < (i / chunk + 1) * chunk)   (utils/mask.py:116-124, all left chunks),
//   out = softmax_j(s) V.
// fp32 on the CUDA cores (parity mode and bf16 mode share it: the front-end is token-rate work).  Flash-style over 64-key
// tiles with running max / sum per query row.  The first version (32-query tiles, one query x 8 keys per thread, tiles
// loaded synchronously) ran at 12 TFLOP/s: ncu showed the tile fill (STS waiting on its LDG) and the LDS queue as the
// stalls, FMA pipe 25 % (profiles/r01_ncu_relpos.md).  This version
//   * gives every thread a 4 x 4 block of (query, key) pairs: the 16 pairs share 4 K rows, 8 Q rows and only 7 rows of P
//     (those with equal i - j), 19 shared loads per 128 FMAs instead of 18 per 64;
//   * stores K and P with an XOR swizzle of the 16-byte chunk index by (row / 4), so that the 16 threads that read rows
//     4 apart hit 16 different chunks (conflict-free LDS.128);
//   * double-buffers the K / V / P tiles with cp.async (zero-fill for keys past the sequence end): the next tile is in
//     flight during the score, softmax and PV phases of the current one.
constexpr int RA_BQ = 64, RA_BK = 64, RA_D = 64, RA_LD = 68, RA_NP = RA_BQ + RA_BK - 1;
constexpr int RA_STAGE_FLOATS = RA_BK * RA_D * 2 + RA_NP * RA_D;          // K, V, P of one key tile
constexpr int RA_SMEM_FLOATS = 2 * RA_BQ * RA_LD + 2 * RA_STAGE_FLOATS + RA_BQ * RA_LD;

__device__ __forceinline__ float dot4(const float4 a, const float4 b, float acc) {
  acc = fmaf(a.x, b.x, acc); acc = fmaf(a.y, b.y, acc); acc = fmaf(a.z, b.z, acc); return fmaf(a.w, b.w, acc);
}
__device__ __forceinline__ void cp_async16(float* smem_dst, const float* gsrc, bool valid) {
  const unsigned dst = (unsigned)__cvta_generic_to_shared(smem_dst);
  const int bytes = valid ? 16 : 0;                 // src-size 0: the 16 bytes are zero-filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(gsrc), "r"(bytes) : "memory");
}
// row-major [rows][64] floats with the 16-byte chunk index XOR-swizzled by (row / 4)
__device__ __forceinline__ int ra_swz(int row, int chunk) { return row * RA_D + ((chunk ^ ((row >> 2) & 15)) << 2); }

__global__ void __launch_bounds__(256, 1) relpos_attention_kernel(const float* __restrict__ qkv, int ld,
                                                                  const float* __restrict__ pos, int ldp, int center,
                                                                  int n_rel, const float* __restrict__ bias_u,
                                                                  const float* __restrict__ bias_v, void* __restrict__ out,
                                                                  int ldo, int out_dt, const Layout lay, int heads,
                                                                  float scale, int chunk) {
  extern __shared__ __align__(16) float ra_smem[];
  float* Qu = ra_smem;
  float* Qv = Qu + RA_BQ * RA_LD;
  float* stage0 = Qv + RA_BQ * RA_LD;
  float* Ss = stage0 + 2 * RA_STAGE_FLOATS;
  pdl_sync();
  const int b = blockIdx.z, h = blockIdx.y, i0 = blockIdx.x * RA_BQ;
  const int len = __ldg(lay.seq_len + b);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const size_t row0 = (size_t)__ldg(lay.seq_row0 + b);
  const int dm = heads * RA_D;
  const int stride_rows = (b + 1 < lay.n_seq ? __ldg(lay.seq_row0 + b + 1) : lay.rows) - (int)row0;
  if (i0 >= stride_rows) return;
  const float* qbase = qkv + row0 * ld + h * RA_D;
  const int i_last = min(i0 + RA_BQ - 1, len - 1);
  const int kmax = i0 < len ? (chunk > 0 ? min((i_last / chunk + 1) * chunk, len) : len) : 0;
  const int n_tiles = (kmax + RA_BK - 1) / RA_BK;

  auto load_tile = [&](int t) {                    // cp.async of key tile t into stage t & 1
    float* Ks = stage0 + (t & 1) * RA_STAGE_FLOATS;
    float* Vs = Ks + RA_BK * RA_D;
    float* Ps = Vs + RA_BK * RA_D;
    const int j0 = t * RA_BK;
    for (int e = tid; e < RA_BK * (RA_D / 4); e += 256) {
      const int r = e >> 4, c = e & 15;
      const bool ok = j0 + r < len;
      const float* src = qbase + (size_t)(ok ? j0 + r : 0) * ld + c * 4;
      cp_async16(Ks + ra_swz(r, c), src + dm, ok);
      cp_async16(Vs + r * RA_D + c * 4, src + 2 * dm, ok);
    }
    const int rel0 = i0 - j0 - (RA_BK - 1);        // relative position of P row 0 of this tile
    for (int e = tid; e < RA_NP * (RA_D / 4); e += 256) {
      const int r = e >> 4, c = e & 15;
      const int trow = min(max(center - (rel0 + r), 0), n_rel - 1);
      cp_async16(Ps + ra_swz(r, c), pos + (size_t)trow * ldp + h * RA_D + c * 4, true);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  if (n_tiles > 0) load_tile(0);
  // ---- Q tile (+ biases, * scale)
  for (int e = tid; e < RA_BQ * (RA_D / 4); e += 256) {
    const int r = e >> 4, c = (e & 15) * 4;
    float4 qu = make_float4(0.f, 0.f, 0.f, 0.f), qv = qu;
    if (i0 + r < len) {
      const float4 q = *reinterpret_cast<const float4*>(qbase + (size_t)(i0 + r) * ld + c);
      const float4 u = *reinterpret_cast<const float4*>(bias_u + h * RA_D + c);
      const float4 v = *reinterpret_cast<const float4*>(bias_v + h * RA_D + c);
      qu = make_float4((q.x + u.x) * scale, (q.y + u.y) * scale, (q.z + u.z) * scale, (q.w + u.w) * scale);
      qv = make_float4((q.x + v.x) * scale, (q.y + v.y) * scale, (q.z + v.z) * scale, (q.w + v.w) * scale);
    }
    *reinterpret_cast<float4*>(Qu + r * RA_LD + c) = qu;
    *reinterpret_cast<float4*>(Qv + r * RA_LD + c) = qv;
  }
  // score phase: thread (tq4, tk4) owns queries 4*tq4 .. +3 and keys 4*tk4 .. +3
  const int tq4 = tid >> 4, tk4 = tid & 15;
  int lim_q[4];
#pragma unroll
  for (int a = 0; a < 4; ++a) {
    const int iq = i0 + 4 * tq4 + a;
    lim_q[a] = iq < len ? (chunk > 0 ? min((iq / chunk + 1) * chunk, len) : len) : 0;
  }
  // softmax / PV phase: warp w owns query rows 8w .. 8w+7, lane owns output dims 2*lane, 2*lane+1
  float m_run[8], l_run[8], o[8][2];
#pragma unroll
  for (int r = 0; r < 8; ++r) { m_run[r] = -INFINITY; l_run[r] = 0.f; o[r][0] = o[r][1] = 0.f; }

  for (int t = 0; t < n_tiles; ++t) {
    const int j0 = t * RA_BK;
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();                 // tile t has landed; everyone is done with tile t-1 (stage (t+1)&1 and Ss are free)
    if (t + 1 < n_tiles) load_tile(t + 1);
    const float* Ks = stage0 + (t & 1) * RA_STAGE_FLOATS;
    const float* Vs = Ks + RA_BK * RA_D;
    const float* Ps = Vs + RA_BK * RA_D;
    float acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[a][c] = 0.f;
    const int pr0 = 4 * tq4 - 4 * tk4 + (RA_BK - 1) - 3;      // P row of (a = 0, c = 3); (a, c) uses row pr0 + a - c + 3
#ifndef RA_UNROLL
#define RA_UNROLL 2
#endif
    constexpr int kUnroll = RA_UNROLL;
#pragma unroll kUnroll
    for (int ch = 0; ch < RA_D / 4; ++ch) {
      float4 qu[4], qv[4], kk[4], pp[7];
#pragma unroll
      for (int a = 0; a < 4; ++a) {
        qu[a] = *reinterpret_cast<const float4*>(Qu + (4 * tq4 + a) * RA_LD + ch * 4);
        qv[a] = *reinterpret_cast<const float4*>(Qv + (4 * tq4 + a) * RA_LD + ch * 4);
      }
#pragma unroll
      for (int c = 0; c < 4; ++c) kk[c] = *reinterpret_cast<const float4*>(Ks + ra_swz(4 * tk4 + c, ch));
#pragma unroll
      for (int e = 0; e < 7; ++e) pp[e] = *reinterpret_cast<const float4*>(Ps + ra_swz(pr0 + e, ch));
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          acc[a][c] = dot4(qu[a], kk[c], acc[a][c]);
          acc[a][c] = dot4(qv[a], pp[a - c + 3], acc[a][c]);
        }
    }
#pragma unroll
    for (int a = 0; a < 4; ++a) {
      float4 sv;
      sv.x = (j0 + 4 * tk4 + 0 < lim_q[a]) ? acc[a][0] : -INFINITY;
      sv.y = (j0 + 4 * tk4 + 1 < lim_q[a]) ? acc[a][1] : -INFINITY;
      sv.z = (j0 + 4 * tk4 + 2 < lim_q[a]) ? acc[a][2] : -INFINITY;
      sv.w = (j0 + 4 * tk4 + 3 < lim_q[a]) ? acc[a][3] : -INFINITY;
      *reinterpret_cast<float4*>(Ss + (4 * tq4 + a) * RA_LD + 4 * tk4) = sv;
    }
    __syncthreads();
    // ---- online softmax
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const int rr = warp * 8 + r;
      const float s0 = Ss[rr * RA_LD + lane], s1 = Ss[rr * RA_LD + lane + 32];
      float mx = fmaxf(s0, s1);
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
      const float m_new = fmaxf(m_run[r], mx);
      float p0 = 0.f, p1 = 0.f, alpha = 1.f;
      if (m_new > -INFINITY) {
        alpha = expf(m_run[r] - m_new);
        p0 = expf(s0 - m_new);
        p1 = expf(s1 - m_new);
      }
      Ss[rr * RA_LD + lane] = p0;
      Ss[rr * RA_LD + lane + 32] = p1;
      l_run[r] = l_run[r] * alpha + warp_sum(p0 + p1);
      m_run[r] = m_new;
      o[r][0] *= alpha;
      o[r][1] *= alpha;
    }
    __syncwarp();
    // ---- O += P V
#pragma unroll 2
    for (int j = 0; j < RA_BK; j += 4) {
      float2 v[4];
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) v[jj] = *reinterpret_cast<const float2*>(Vs + (j + jj) * RA_D + 2 * lane);
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const float4 p = *reinterpret_cast<const float4*>(Ss + (warp * 8 + r) * RA_LD + j);
        o[r][0] = fmaf(p.x, v[0].x, o[r][0]); o[r][1] = fmaf(p.x, v[0].y, o[r][1]);
        o[r][0] = fmaf(p.y, v[1].x, o[r][0]); o[r][1] = fmaf(p.y, v[1].y, o[r][1]);
        o[r][0] = fmaf(p.z, v[2].x, o[r][0]); o[r][1] = fmaf(p.z, v[2].y, o[r][1]);
        o[r][0] = fmaf(p.w, v[3].x, o[r][0]); o[r][1] = fmaf(p.w, v[3].y, o[r][1]);
      }
    }
  }
  // ---- write (zeros on rows past the valid length)
#pragma unroll
  for (int r = 0; r < 8; ++r) {
    const int i = i0 + warp * 8 + r;
    if (i >= stride_rows) continue;
    float a = 0.f, c = 0.f;
    if (i < len && l_run[r] > 0.f) { a = o[r][0] / l_run[r]; c = o[r][1] / l_run[r]; }
    const size_t idx = (row0 + i) * ldo + h * RA_D + 2 * lane;
    if (out_dt == DT_BF16) *reinterpret_cast<__nv_bfloat162*>(reinterpret_cast<__nv_bfloat16*>(out) + idx) = __floats2bfloat162_rn(a, c);
    else *reinterpret_cast<float2*>(reinterpret_cast<float*>(out) + idx) = make_float2(a, c);
  }
}

// ---------------------------------------------------------------------------- GroupNorm(1, C) + Mish (length regulator)
// InterpolateRegulator.model: Conv1d k3 -> GroupNorm(1, 80) -> Mish (flow/length_regulator.py:33-38).  The statistics run over
// all C channels x all valid frames of ONE utterance; one CTA per sequence (a 30 s utterance is 225 k elements), sums in
// double.  Invalid rows are written as zeros (the next conv's padding).
__global__ void __launch_bounds__(1024) groupnorm1_mish_kernel(const float* __restrict__ h, int ldh, void* __restrict__ out,
                                                               int ldo, int out_dt, const float* __restrict__ gamma,
                                                               const float* __restrict__ beta, int C, const Layout lay,
                                                               float eps) {
  __shared__ double red[2][32];
  __shared__ float stat[2];
  pdl_sync();
  const int b = blockIdx.x;
  const int len = __ldg(lay.seq_len + b);
  const size_t row0 = (size_t)__ldg(lay.seq_row0 + b);
  const int stride_rows = (b + 1 < lay.n_seq ? __ldg(lay.seq_row0 + b + 1) : lay.rows) - (int)row0;
  const int n = len * C;
  float s = 0.f, q = 0.f;
  for (int e = threadIdx.x; e < n; e += blockDim.x) {
    const float v = h[(row0 + e / C) * ldh + e % C];
    s += v;
    q = fmaf(v, v, q);
  }
  double ds = s, dq = q;
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    ds += __shfl_xor_sync(0xffffffffu, ds, off);
    dq += __shfl_xor_sync(0xffffffffu, dq, off);
  }
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = ds; red[1][threadIdx.x >> 5] = dq; }
  __syncthreads();
  if (threadIdx.x < 32) {
    ds = red[0][threadIdx.x];
    dq = red[1][threadIdx.x];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      ds += __shfl_xor_sync(0xffffffffu, ds, off);
      dq += __shfl_xor_sync(0xffffffffu, dq, off);
    }
    if (threadIdx.x == 0) {
      const double mean = n > 0 ? ds / n : 0.0;
      const double var = n > 0 ? fmax(dq / n - mean * mean, 0.0) : 0.0;
      stat[0] = (float)mean;
      stat[1] = (float)(1.0 / sqrt(var + (double)eps));
    }
  }
  __syncthreads();
  const float mean = stat[0], rstd = stat[1];
  const int total = stride_rows * C;
  for (int e = threadIdx.x; e < total; e += blockDim.x) {
    const int r = e / C, c = e % C;
    float v = 0.f;
    if (r < len) v = mishf((h[(row0 + r) * ldh + c] - mean) * rstd * __ldg(gamma + c) + __ldg(beta + c));
    st_from_float(out, (row0 + r) * ldo + c, out_dt, v);
  }
}

// ---------------------------------------------------------------------------- piecewise linear interpolation (rows)
// InterpolateRegulator.inference (flow/length_regulator.py:52-70): up to 4 pieces per utterance, each
// F.interpolate(mode='linear', align_corners=False) of src rows [s, s+n) onto dst rows [d, d+m), with ATen's arithmetic
// (scale = n/m in fp32, src = scale*(j+0.5)-0.5 clamped at 0, (1-l)*x0 + l*x1).  segs: int32 [n_seq][4][4] = (s, n, d, m),
// m = 0 marks an unused piece.  One thread per (dst row, channel); invalid dst rows are written as zeros.
__global__ void __launch_bounds__(256) interp_rows_kernel(const float* __restrict__ x, int ldx, const Layout src_lay,
                                                          void* __restrict__ out, int ldo, int out_dt, const Layout dst_lay,
                                                          const int* __restrict__ segs, int C) {
  pdl_sync();
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)dst_lay.rows * C) return;
  const int row = (int)(idx / C), c = (int)(idx % C);
  int b, t, len;
  seq_of_row(dst_lay, row, b, t, len);
  float v = 0.f;
  if (t < len) {
    const int* sg = segs + b * 16;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int s0 = __ldg(sg + 4 * k), n = __ldg(sg + 4 * k + 1), d0 = __ldg(sg + 4 * k + 2), m = __ldg(sg + 4 * k + 3);
      if (m > 0 && t >= d0 && t < d0 + m) {
        const int j = t - d0;
        const float scale = (float)n / (float)m;
        float src = __fmaf_rn(scale, (float)j + 0.5f, -0.5f);
        src = src < 0.f ? 0.f : src;
        const int i0 = min((int)src, n - 1);
        const int i1 = i0 + (i0 < n - 1 ? 1 : 0);
        const float l1 = src - (float)i0, l0 = 1.f - l1;
        const size_t base = (size_t)__ldg(src_lay.seq_row0 + b) + s0;
        v = __fmaf_rn(l0, x[(base + i0) * ldx + c], __fmul_rn(l1, x[(base + i1) * ldx + c]));
      }
    }
  }
  st_from_float(out, (size_t)row * ldo + c, out_dt, v);
}

// ---------------------------------------------------------------------------- speaker vector normalisation
// F.normalize(embedding, dim=1) (flow/flow.py:127): x / max(||x||_2, 1e-12); one warp per row.
__global__ void __launch_bounds__(32) l2_normalize_kernel(const float* __restrict__ x, float* __restrict__ out, int K) {
  pdl_sync();
  const float* xr = x + (size_t)blockIdx.x * K;
  float q = 0.f;
  for (int k = threadIdx.x; k < K; k += 32) q = fmaf(xr[k], xr[k], q);
  const float inv = 1.f / fmaxf(sqrtf(warp_sum(q)), 1e-12f);
  for (int k = threadIdx.x; k < K; k += 32) out[(size_t)blockIdx.x * K + k] = xr[k] * inv;
}

// ---------------------------------------------------------------------------- launchers
int launch_embed_tokens(const int* tokens, int tok_ld, const float* table, int vocab, int C, void* out, int ldo, int out_dt,
                        const Layout& lay, cudaStream_t st) {
  CVB_REQUIRE(C % 4 == 0 && ldo % 4 == 0 && vocab > 0, "embed_tokens: C and ldo must be multiples of 4");
  const long long total = (long long)lay.rows * (C / 4);
  if (total <= 0) return CVB_OK;
  CVB_CUDA(launch_pdl(embed_tokens_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, st, tokens, tok_ld, table, vocab,
                      C, out, ldo, out_dt, lay));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_layernorm_seq(const float* x, int ldx, void* out, int ldo, int out_dt, const float* gamma, const float* beta,
                         const Layout& lay, int C, float eps, cudaStream_t st) {
  CVB_REQUIRE(C == 256 || C == 512, "layernorm_seq: C must be 256 or 512");
  CVB_REQUIRE(ldx % 4 == 0 && ldo % 4 == 0, "layernorm_seq: leading dimensions must keep 16-byte alignment");
  if (lay.rows <= 0) return CVB_OK;
  if (C == 256)
    CVB_CUDA(launch_pdl(layernorm_seq_kernel<2>, dim3(ceil_div(lay.rows, 8)), dim3(256), 0, st, x, ldx, out, ldo, out_dt, gamma, beta, lay, eps));
  else
    CVB_CUDA(launch_pdl(layernorm_seq_kernel<4>, dim3(ceil_div(lay.rows, 8)), dim3(256), 0, st, x, ldx, out, ldo, out_dt, gamma, beta, lay, eps));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_relpos_attention(const float* qkv, int ld, const float* pos, int ldp, int center, int n_rel, const float* bias_u,
                            const float* bias_v, void* out, int ldo, int out_dt, const Layout& lay, int heads, float scale,
                            int chunk, cudaStream_t st) {
  CVB_REQUIRE(heads > 0 && ld >= 3 * heads * RA_D && ld % 4 == 0 && ldp % 4 == 0 && ldo % 2 == 0, "relpos_attention: bad strides");
  CVB_REQUIRE(center - (lay.max_len - 1) >= 0 && center + (lay.max_len - 1) < n_rel,
              "relpos_attention: the position table is shorter than the longest sequence");
  CVB_REQUIRE(chunk >= 0, "relpos_attention: chunk must be >= 0");
  if (lay.rows <= 0 || lay.max_len <= 0) return CVB_OK;
  static bool attr_set = false;
  const size_t smem = RA_SMEM_FLOATS * sizeof(float);
  if (!attr_set) {
    CVB_CUDA(cudaFuncSetAttribute(relpos_attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set = true;
  }
  const dim3 grid(ceil_div(lay.max_len, RA_BQ), heads, lay.n_seq);
  CVB_CUDA(launch_pdl(relpos_attention_kernel, grid, dim3(256), smem, st, qkv, ld, pos, ldp, center, n_rel, bias_u, bias_v, out,
                      ldo, out_dt, lay, heads, scale, chunk));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_groupnorm1_mish(const float* h, int ldh, void* out, int ldo, int out_dt, const float* gamma, const float* beta,
                           int C, const Layout& lay, float eps, cudaStream_t st) {
  CVB_REQUIRE(C > 0 && C <= ldh && C <= ldo, "groupnorm1_mish: C must fit the row pitch");
  if (lay.n_seq <= 0) return CVB_OK;
  CVB_CUDA(launch_pdl(groupnorm1_mish_kernel, dim3(lay.n_seq), dim3(1024), 0, st, h, ldh, out, ldo, out_dt, gamma, beta, C, lay, eps));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_interp_rows(const float* x, int ldx, const Layout& src_lay, void* out, int ldo, int out_dt, const Layout& dst_lay,
                       const int* segs, int C, cudaStream_t st) {
  CVB_REQUIRE(src_lay.n_seq == dst_lay.n_seq && C > 0 && C <= ldx && C <= ldo, "interp_rows: bad arguments");
  const long long total = (long long)dst_lay.rows * C;
  if (total <= 0) return CVB_OK;
  CVB_CUDA(launch_pdl(interp_rows_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, st, x, ldx, src_lay, out, ldo, out_dt,
                      dst_lay, segs, C));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_l2_normalize(const float* x, float* out, int S, int K, cudaStream_t st) {
  CVB_REQUIRE(K > 0, "l2_normalize: K must be positive");
  if (S <= 0) return CVB_OK;
  CVB_CUDA(launch_pdl(l2_normalize_kernel, dim3(S), dim3(32), 0, st, x, out, K));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

}  // namespace cvb
