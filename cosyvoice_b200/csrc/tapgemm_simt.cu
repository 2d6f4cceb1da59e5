// CUDA-core (FFMA) engine for the tap GEMM: fp32 parity mode and the on-device cross-check for
// the tcgen05 engine (same operands, fp32 accumulate).  Not the performance path.
#include "tapgemm.cuh"

namespace cvb {

namespace {

constexpr int BM = TILE_M, BN = 64, BK = 16, NT = 256;

template <typename T>
__global__ void __launch_bounds__(NT) tapgemm_simt_kernel(const GemmParams p) {
  pdl_sync();
  __shared__ float As[BK][BM + 4];
  __shared__ float Ws[BK][BN + 4];
  const int tile = blockIdx.x, nt = blockIdx.y, phase = blockIdx.z;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int m0 = tile * BM;
  const int b = p.tile_seq[tile];
  const int lrow0 = p.tile_row0[tile];
  const T* __restrict__ A = reinterpret_cast<const T*>(p.A);
  const T* __restrict__ W = reinterpret_cast<const T*>(p.W) + (size_t)phase * p.n_pad * (size_t)(p.n_seg * p.seg_len);
  const int K = p.n_seg * p.seg_len;

  float acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  const int a_r = tid >> 1, a_k = (tid & 1) * 8;     // A loader: row, first k
  const int w_n = tid >> 2, w_k = (tid & 3) * 4;     // W loader
  for (int k0 = 0; k0 < K; k0 += BK) {
    const int seg = k0 / p.seg_len, c0 = k0 - seg * p.seg_len;
    const int srow = p.seg_row[phase * p.n_seg + seg], scol = p.seg_col[phase * p.n_seg + seg];
    {
      const int r = m0 + a_r + srow;
      const bool rin = (r >= 0 && r < p.a_rows);
      const int col = scol + c0 + a_k;
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        float v = 0.f;
        if (rin && col + q < p.a_cols) v = to_f<T>(A[(size_t)r * p.lda + col + q]);
        As[a_k + q][a_r] = v;
      }
    }
    {
      const int n = nt * BN + w_n;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float v = 0.f;
        if (n < p.n_pad) v = to_f<T>(W[(size_t)n * K + k0 + w_k + q]);
        Ws[w_k + q][w_n] = v;
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      const float4 a0 = *reinterpret_cast<const float4*>(&As[k][ty * 8]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[k][ty * 8 + 4]);
      const float4 w = *reinterpret_cast<const float4*>(&Ws[k][tx * 4]);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], ww[j], acc[i][j]);
    }
    __syncthreads();
  }

  // ---- epilogue
  const int n0 = nt * BN + tx * 4;
  const int slen = p.seq_len[b];
  const int olim = p.out_limit ? p.out_limit[b] : slen * p.n_phase;
  float gsum = 0.f, gsq = 0.f;
  if (n0 < p.n_pad) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int lrow = lrow0 + ty * 8 + i;
      const bool valid = lrow < slen && (lrow * p.n_phase + phase) < olim;
      if (p.dt == DT_F32) epi_apply<4, true>(p.epi, (size_t)(m0 + ty * 8 + i), b, valid, phase, n0, acc[i], gsum, gsq);
      else epi_apply<4, false>(p.epi, (size_t)(m0 + ty * 8 + i), b, valid, phase, n0, acc[i], gsum, gsq);
    }
  }
  if (p.epi.gn_stats) {
    // lanes with equal (tx >> 3) share a 32-channel group: reduce over tx&7 (xor 1,2,4) and ty parity (xor 16)
#pragma unroll
    for (int o : {1, 2, 4, 16}) {
      gsum += __shfl_xor_sync(0xffffffffu, gsum, o);
      gsq += __shfl_xor_sync(0xffffffffu, gsq, o);
    }
    const int lane = tid & 31;
    if ((lane & 0x17) == 0 && n0 < p.N) {
      const int g = n0 >> 5;
      double* s = p.epi.gn_stats + ((size_t)b * p.epi.gn_groups + g) * 2;
      atomicAdd(s, (double)gsum);
      atomicAdd(s + 1, (double)gsq);
    }
  }
}

}  // namespace

int launch_tapgemm_simt(const GemmParams& p, cudaStream_t stream) {
  CVB_REQUIRE(p.seg_len % BK == 0, "seg_len must be a multiple of 16");
  CVB_REQUIRE(p.n_phase * p.n_seg <= MAX_SEGS, "too many segments");
  CVB_REQUIRE(p.m_tiles > 0, "empty problem");
  CVB_REQUIRE(!p.epi.row_parts && !p.epi.row_parts_out, "row_parts / row_parts_out: tcgen05 engine only");
  CVB_TRY(check_epilogue_layout(p));
  dim3 grid(p.m_tiles, ceil_div(p.n_pad, BN), p.n_phase);
  if (p.dt == DT_BF16) CVB_CUDA(launch_pdl(tapgemm_simt_kernel<__nv_bfloat16>, dim3(grid), dim3(NT), 0, stream, p));
  else CVB_CUDA(launch_pdl(tapgemm_simt_kernel<float>, dim3(grid), dim3(NT), 0, stream, p));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

}  // namespace cvb
