// Fused FeedForward of the flow estimator's BasicTransformerBlock on tcgen05 / TMEM:
//     X[m, :] += W2 . gelu(W1 . a[m, :] + b1) + b2          (cosyvoice/flow/components/transformer.py:83-136, 309-314)
// with a = LayerNorm(X) in bf16 [rows, 256], W1 [1024, 256], W2 [256, 1024] (nn.Linear layouts), X the fp32
// residual stream.  The two GEMMs of the unfused form are HBM-write/read bound on the [rows, 1024] hidden tensor
// (168 MB out + 168 MB back in at rows = 81920); here the hidden activations never leave the SM:
//
//   per 128-row tile, for each of the 8 hidden chunks c (128 units):
//     G1(c): H[c%2] (TMEM, fp32 128x128)  = A(smem) . W1[128c:128c+128, :]^T           16 x UMMA 128x128x16 (SS)
//     epilogue warps: H -> + b1 -> GELU -> bf16, written back IN PLACE into the first half of each warp's own
//                     32-column slice of H (each warp only overwrites columns it has already pulled into registers)
//     G2(c): O (TMEM, fp32 128x256)      += Hbf16(TMEM) . W2[:, 128c:128c+128]^T         8 x UMMA 128x256x16 (TS)
//   G1(c+1) is issued before G2(c) so the tensor pipe works while the GELU of chunk c runs.
//   final: O + b2 + X -> X (fp32, 256-bit accesses), optional bf16 copy; pad rows are written as zeros.
//
// TMEM: H0 [0,128) H1 [128,256) O [256,512).  smem: A tile 64 KB + 4 x 32 KB weight ring + b1/b2.
// Weights stream from L2 (1 MB per tile); activations make one HBM round trip (a in, X in/out).
#include <cuda.h>
#include <cuda_bf16.h>

#include <mutex>

#include "common.cuh"
#include "ptx.cuh"
#include "../../include/cvb200.h"

namespace cvb {

using Layout = cvb200_layout;
int make_map_2d_bf16(CUtensorMap* map, const void* base, uint64_t cols, uint64_t rows, uint64_t row_stride_elems,
                     uint32_t box_cols, uint32_t box_rows);       // tapgemm_tc.cu
int device_sm_count();

namespace {

constexpr int FF_C = 256;            // model width
constexpr int FF_H = 1024;           // hidden width
constexpr int FF_CH = 128;           // hidden chunk
constexpr int FF_NCH = FF_H / FF_CH;
constexpr int FF_M = 128;
constexpr int FF_BK = 64;
constexpr int FF_A_BYTES = FF_M * FF_C * 2;          // 64 KiB: 4 k-blocks of [128 x 64]
constexpr int FF_STAGE_BYTES = 32 * 1024;            // G1: two [128 x 64] W1 boxes; G2: one [256 x 64] W2 box
constexpr int FF_STAGES = 4;
constexpr int FF_EPI_WARPS = 16;
constexpr int FF_THREADS = 64 + 32 * FF_EPI_WARPS;
constexpr int FF_TM_H = 0, FF_TM_O = 256, FF_TM_COLS = 512;
constexpr int FF_VEC_BYTES = (FF_H + FF_C) * 4;
constexpr int FF_BAR_BYTES = 256;
constexpr int FF_SMEM = 1024 + FF_A_BYTES + FF_STAGES * FF_STAGE_BYTES + FF_BAR_BYTES + FF_VEC_BYTES;

#ifdef FFN_PROFILE
#define FF_WAIT(acc, bar, par)             \
  do {                                     \
    const long long _t0 = clock64();       \
    ptx::mbar_wait(bar, par);              \
    acc += clock64() - _t0;                \
  } while (0)
#else
#define FF_WAIT(acc, bar, par) ptx::mbar_wait(bar, par)
#endif

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}

// same fitted tanh-form erf-GELU as the tap GEMM's bf16 epilogue (tapgemm.cuh gelu_erf_fast)
__device__ __forceinline__ float gelu_fast(float v) {
  const float xc = fminf(fmaxf(v, -6.f), 6.f);
  const float x2 = xc * xc;
  float p = fmaf(-3.51516916e-4f, x2, 3.70056467e-2f);
  p = fmaf(p, x2, 7.97507884e-1f);
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(xc * p));
  const float h = 0.5f * v;
  return fmaf(h, t, h);
}

__global__ void __launch_bounds__(FF_THREADS, 1)
ffn_fused_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW1,
                 const __grid_constant__ CUtensorMap tmW2, const float* __restrict__ b1, const float* __restrict__ b2,
                 float* __restrict__ X, int ldx, __nv_bfloat16* __restrict__ out1, int ldo1, const Layout lay,
                 int m_tiles) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sW = smem + FF_A_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sW + FF_STAGES * FF_STAGE_BYTES);
  uint64_t* w_full = bars;                   // [STAGES]
  uint64_t* w_empty = bars + FF_STAGES;      // [STAGES]
  uint64_t* a_full = bars + 2 * FF_STAGES;   // A tile landed
  uint64_t* a_free = a_full + 1;             // the tile's last G1 retired: A may be overwritten
  uint64_t* h_full = a_free + 1;             // [2] G1(c) complete
  uint64_t* p_full = h_full + 2;             // [2] bf16 H(c) written back (16 warp arrivals)
  uint64_t* o_full = p_full + 2;             // the tile's last G2 retired
  uint64_t* o_free = o_full + 1;             // O drained by the epilogue (16 warp arrivals)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_free + 1);
  float* sb1 = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + FF_BAR_BYTES);
  float* sb2 = sb1 + FF_H;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < FF_STAGES; ++i) {
      ptx::mbar_init(&w_full[i], 1);
      ptx::mbar_init(&w_empty[i], 1);
    }
    ptx::mbar_init(a_full, 1);
    ptx::mbar_init(a_free, 1);
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&h_full[i], 1);
      ptx::mbar_init(&p_full[i], FF_EPI_WARPS);
    }
    ptx::mbar_init(o_full, 1);
    ptx::mbar_init(o_free, FF_EPI_WARPS);
    ptx::fence_barrier_init();
    ptx::prefetch_tensormap(&tmA);
    ptx::prefetch_tensormap(&tmW1);
    ptx::prefetch_tensormap(&tmW2);
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, FF_TM_COLS);
  for (int i = threadIdx.x; i < FF_H; i += FF_THREADS) sb1[i] = __ldg(b1 + i);       // model constants
  for (int i = threadIdx.x; i < FF_C; i += FF_THREADS) sb2[i] = __ldg(b2 + i);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  pdl_trigger();
  pdl_wait();

  const int t_first = blockIdx.x, t_step = gridDim.x;
#ifdef FFN_STAGGER
  {   // experiment: spread the CTAs' phases so that their HBM-heavy drains do not coincide
    const long long t0 = clock64();
    const long long d = (long long)(blockIdx.x & 3) * FFN_STAGGER;
    while (clock64() - t0 < d) {}
  }
#endif
#ifdef FFN_PROFILE
  long long prof[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  const long long t_start = clock64();
#endif

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t ph = 0;
      auto next_stage = [&]() { if (++stage == FF_STAGES) { stage = 0; ph ^= 1; } };
      auto load_g1 = [&](int c) {            // W1 rows [128c, +128): two stages of two [128 x 64] boxes
        for (int h = 0; h < 2; ++h) {
          FF_WAIT(prof[0], &w_empty[stage], ph ^ 1);
          ptx::mbar_arrive_expect_tx(&w_full[stage], FF_STAGE_BYTES);
          uint8_t* dst = sW + stage * FF_STAGE_BYTES;
          ptx::tma_load_2d(dst, &tmW1, &w_full[stage], (2 * h) * FF_BK, c * FF_CH);
          ptx::tma_load_2d(dst + 16384, &tmW1, &w_full[stage], (2 * h + 1) * FF_BK, c * FF_CH);
          next_stage();
        }
      };
      auto load_g2 = [&](int c) {            // W2[:, 128c + 64j, +64): two stages of one [256 x 64] box
        for (int j = 0; j < 2; ++j) {
          FF_WAIT(prof[0], &w_empty[stage], ph ^ 1);
          ptx::mbar_arrive_expect_tx(&w_full[stage], FF_STAGE_BYTES);
          ptx::tma_load_2d(sW + stage * FF_STAGE_BYTES, &tmW2, &w_full[stage], c * FF_CH + j * FF_BK, 0);
          next_stage();
        }
      };
      int n = 0;
      for (int t = t_first; t < m_tiles; t += t_step, ++n) {
        if (n > 0) FF_WAIT(prof[1], a_free, (n - 1) & 1);
        ptx::mbar_arrive_expect_tx(a_full, FF_A_BYTES);
        for (int kb = 0; kb < FF_C / FF_BK; ++kb)
          ptx::tma_load_2d(sA + kb * 16384, &tmA, a_full, kb * FF_BK, t * FF_M);
        // weight boxes in the order the MMA warp consumes them: G1(0), then G1(c+1), G2(c)
        load_g1(0);
        for (int c = 0; c < FF_NCH; ++c) {
          if (c + 1 < FF_NCH) load_g1(c + 1);
          load_g2(c);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      const uint32_t idesc1 = ptx::make_idesc_bf16(FF_M, FF_CH);       // 128 x 128
      const uint32_t idesc2 = ptx::make_idesc_bf16(FF_M, FF_C);        // 128 x 256
      const uint32_t a_base = ptx::smem_u32(sA);
      int stage = 0;
      uint32_t ph = 0;
      auto next_stage = [&]() { if (++stage == FF_STAGES) { stage = 0; ph ^= 1; } };
      uint32_t hcount[2] = {0, 0};           // completed uses of each H buffer (for the p_full parity)
      auto issue_g1 = [&](int c) {
        const uint32_t d = tmem + FF_TM_H + (c & 1) * FF_CH;
        for (int h = 0; h < 2; ++h) {
          FF_WAIT(prof[2], &w_full[stage], ph);
          ptx::tc_fence_after();
          const uint32_t w_base = ptx::smem_u32(sW + stage * FF_STAGE_BYTES);
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {       // 2 k-blocks x 4 K16 steps
            const int kb = 2 * h + (kk >> 2), k = kk & 3;
            ptx::umma_ss(d, ptx::make_sw128_kmajor_desc(a_base + kb * 16384 + k * 32),
                         ptx::make_sw128_kmajor_desc(w_base + (kk >> 2) * 16384 + k * 32), idesc1, (h | kk) != 0);
          }
          ptx::umma_commit(&w_empty[stage]);
          next_stage();
        }
        ptx::umma_commit(&h_full[c & 1]);
      };
      int n = 0;
      for (int t = t_first; t < m_tiles; t += t_step, ++n) {
        FF_WAIT(prof[3], a_full, n & 1);
        ptx::tc_fence_after();
        issue_g1(0);
        for (int c = 0; c < FF_NCH; ++c) {
          if (c + 1 < FF_NCH) issue_g1(c + 1);
          else ptx::umma_commit(a_free);                         // all G1 of this tile are issued: A is free when they retire
          const int hb = c & 1;
          FF_WAIT(prof[4], &p_full[hb], hcount[hb] & 1);
          ++hcount[hb];
          if (c == 0 && n > 0) FF_WAIT(prof[5], o_free, (n - 1) & 1);   // the previous tile's O has been drained
          ptx::tc_fence_after();
          const uint32_t hbase = tmem + FF_TM_H + hb * FF_CH;
          for (int j = 0; j < 2; ++j) {
            FF_WAIT(prof[6], &w_full[stage], ph);
            ptx::tc_fence_after();
            const uint32_t w_base = ptx::smem_u32(sW + stage * FF_STAGE_BYTES);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              // K16 step u = 4j + k of the chunk: hidden units [16u, 16u+16) live in packed columns
              // 32*(u/2) + 8*(u%2) (each epilogue warp packs its 32 units into the first 16 columns of its slice)
              const int u = 4 * j + k;
              ptx::umma_ts(tmem + FF_TM_O, hbase + 32 * (u >> 1) + 8 * (u & 1),
                           ptx::make_sw128_kmajor_desc(w_base + k * 32), idesc2, (c | u) != 0);
            }
            ptx::umma_commit(&w_empty[stage]);
            next_stage();
          }
        }
        ptx::umma_commit(o_full);
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps
    const int q = warp & 3;                    // TMEM lane quadrant this warp may access
    const int sub = (warp - 2) >> 2;           // 0..3: 32-column slice of an H chunk / 64-column slice of O
    const uint32_t tbase = tmem + ((uint32_t)(q * 32) << 16);
    const uint32_t sb1_addr = ptx::smem_u32(sb1), sb2_addr = ptx::smem_u32(sb2);
    uint32_t hcount[2] = {0, 0};
    int n = 0;
    for (int t = t_first; t < m_tiles; t += t_step, ++n) {
      for (int c = 0; c < FF_NCH; ++c) {
        const int hb = c & 1;
        FF_WAIT(prof[7], &h_full[hb], hcount[hb] & 1);
        ++hcount[hb];
        ptx::tc_fence_after();
        const uint32_t haddr = tbase + FF_TM_H + hb * FF_CH + sub * 32;
        float v[32];
        ptx::tmem_ld32(haddr, v);
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          float bb[4];
          asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                       : "=f"(bb[0]), "=f"(bb[1]), "=f"(bb[2]), "=f"(bb[3])
                       : "r"(sb1_addr + (c * FF_CH + sub * 32 + i) * 4));
          pk[i >> 1] = pack_bf16x2(gelu_fast(v[i] + bb[0]), gelu_fast(v[i + 1] + bb[1]));
          pk[(i >> 1) + 1] = pack_bf16x2(gelu_fast(v[i + 2] + bb[2]), gelu_fast(v[i + 3] + bb[3]));
        }
        ptx::tmem_st16(haddr, pk);
        ptx::tmem_wait_st();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&p_full[hb]);
        if (c == FF_NCH / 2) {                     // pull this thread's residual slice into L2 ahead of the drain
          const float* xp = X + (size_t)(t * FF_M + q * 32 + lane) * ldx + sub * 64;
          asm volatile("prefetch.global.L2 [%0];" ::"l"(xp));
          asm volatile("prefetch.global.L2 [%0];" ::"l"(xp + 32));
        }
      }
      // ---- O + b2 + X -> X.  The residual slice (64 fp32 per thread) is fetched with a two-deep software pipeline
      // that starts BEFORE the wait on the last GEMM, so its latency overlaps the tail of the tile's MMAs.
      const int row = t * FF_M + q * 32 + lane;
      const int b = __ldg(lay.tile_seq + t);
      const bool valid = __ldg(lay.tile_row0 + t) + q * 32 + lane < __ldg(lay.seq_len + b);
      float* xr = X + (size_t)row * ldx + sub * 64;
      float res[8][8];
      auto ld_res = [&](int i) {
#ifdef FFN_NO_RES
#pragma unroll
        for (int e = 0; e < 8; ++e) res[i][e] = 0.f;
        return;
#endif
        asm volatile("ld.global.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                     : "=f"(res[i][0]), "=f"(res[i][1]), "=f"(res[i][2]), "=f"(res[i][3]), "=f"(res[i][4]),
                       "=f"(res[i][5]), "=f"(res[i][6]), "=f"(res[i][7])
                     : "l"(xr + i * 8));
      };
      ld_res(0);
      ld_res(1);
      ld_res(2);
      FF_WAIT(prof[8], o_full, n & 1);
      ptx::tc_fence_after();
#ifdef FFN_PROFILE
      const long long _td = clock64();
#endif
#pragma unroll
      for (int hx = 0; hx < 2; ++hx) {
        float v[32];
        ptx::tmem_ld32(tbase + FF_TM_O + sub * 64 + hx * 32, v);
        if (hx == 1) {                       // both halves are in registers: the next tile's G2 may overwrite O
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(o_free);
        }
#pragma unroll
        for (int i4 = 0; i4 < 4; ++i4) {
          const int i = hx * 4 + i4;           // 8-column group of the thread's 64 columns
          if (i + 3 < 8) ld_res(i + 3);
          float bb[8];
          asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                       : "=f"(bb[0]), "=f"(bb[1]), "=f"(bb[2]), "=f"(bb[3])
                       : "r"(sb2_addr + (sub * 64 + i * 8) * 4));
          asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                       : "=f"(bb[4]), "=f"(bb[5]), "=f"(bb[6]), "=f"(bb[7])
                       : "r"(sb2_addr + (sub * 64 + i * 8 + 4) * 4));
          uint32_t o[8];
          float y[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            y[e] = valid ? v[i4 * 8 + e] + bb[e] + res[i][e] : 0.f;
            o[e] = __float_as_uint(y[e]);
          }
#ifdef FFN_NO_ST
          if (o[0] == 0x12345678u)
#endif
          asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(xr + i * 8), "r"(o[0]),
                       "r"(o[1]), "r"(o[2]), "r"(o[3]), "r"(o[4]), "r"(o[5]), "r"(o[6]), "r"(o[7])
                       : "memory");
          if (out1) {
            uint4 p4;
            p4.x = pack_bf16x2(y[0], y[1]);
            p4.y = pack_bf16x2(y[2], y[3]);
            p4.z = pack_bf16x2(y[4], y[5]);
            p4.w = pack_bf16x2(y[6], y[7]);
            *reinterpret_cast<uint4*>(out1 + (size_t)row * ldo1 + sub * 64 + i * 8) = p4;
          }
        }
      }
#ifdef FFN_PROFILE
      prof[9] += clock64() - _td;
#endif
    }
  }
#ifdef FFN_PROFILE
  if (blockIdx.x == 0 && lane == 0 && (warp < 3))
    printf("ffn prof warp %d total %lld: w_empty %lld a_free %lld | g1 w_full %lld a_full %lld p_full %lld o_free %lld g2 w_full %lld | "
           "h_full %lld o_full %lld drain %lld\n", warp, clock64() - t_start, prof[0], prof[1], prof[2], prof[3], prof[4], prof[5],
           prof[6], prof[7], prof[8], prof[9]);
#endif
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem, FF_TM_COLS);
}

}  // namespace

int launch_ffn_fused(const void* a, int lda, const void* w1, const float* b1, const void* w2, const float* b2, float* x,
                     int ldx, void* out1, int ldo1, const Layout& lay, cudaStream_t st) {
  CVB_REQUIRE(a && w1 && w2 && b1 && b2 && x, "null operand");
  CVB_REQUIRE(lay.rows > 0 && lay.rows % FF_M == 0, "rows must be a positive multiple of 128");
  CVB_REQUIRE(lda >= FF_C && lda % 8 == 0, "lda must cover 256 bf16 columns (16-byte pitch)");
  CVB_REQUIRE(ldx >= FF_C && ldx % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 31) == 0,
              "the residual stream must be 32-byte aligned with a 32-byte row pitch");
  CVB_REQUIRE(!out1 || (ldo1 >= FF_C && ldo1 % 8 == 0 && (reinterpret_cast<uintptr_t>(out1) & 15) == 0), "out1 alignment");
  CVB_REQUIRE((reinterpret_cast<uintptr_t>(b1) & 15) == 0 && (reinterpret_cast<uintptr_t>(b2) & 15) == 0, "bias alignment");
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [] {
    attr_err = cudaFuncSetAttribute(ffn_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FF_SMEM);
  });
  CVB_CUDA(attr_err);
  CUtensorMap tmA, tmW1, tmW2;
  CVB_TRY(make_map_2d_bf16(&tmA, a, FF_C, (uint64_t)lay.rows, (uint64_t)lda, FF_BK, FF_M));
  CVB_TRY(make_map_2d_bf16(&tmW1, w1, FF_C, FF_H, FF_C, FF_BK, FF_CH));
  CVB_TRY(make_map_2d_bf16(&tmW2, w2, FF_H, FF_C, FF_H, FF_BK, FF_C));
  const int m_tiles = lay.rows / FF_M;
  const int sms = device_sm_count();
  const int grid = m_tiles < sms ? m_tiles : sms;
  CVB_CUDA(launch_pdl(ffn_fused_kernel, dim3(grid), dim3(FF_THREADS), FF_SMEM, st, tmA, tmW1, tmW2, b1, b2, x, ldx,
                      reinterpret_cast<__nv_bfloat16*>(out1), ldo1, lay, m_tiles));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

}  // namespace cvb
