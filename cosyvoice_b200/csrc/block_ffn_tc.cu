// Second half of the estimator's BasicTransformerBlock as ONE kernel on the bf16 residual stream (tcgen05 / TMEM):
//     X     += W2 . gelu(W1 . LayerNorm(X; g3, be3) + b1) + b2          norm3 + ff + residual
//                                                   cosyvoice/flow/components/transformer.py:300, 83-136, 309-314
//     stats  = (rstd, -mean rstd) of every new row (optional)           what norm1 of the NEXT block (:262) needs
// X is the bf16 residual stream [rows, 256], read once by TMA and written once.  The unfused form (LayerNorm launch, FFN-in
// GEMM, FFN-out GEMM, next LayerNorm launch) moves 273 + 42 MB through HBM per block at 40 960 rows where this kernel moves
// 42: the [rows, 1024] hidden tensor never leaves the SM and no normalised copy of X is ever written.
//
// LayerNorm is folded into the GEMM algebraically: with W1' = W1 diag(g3) (rounded to bf16 by the host), u = rowsum(W1'),
// b1' = b1 + W1 be3,
//     W1 . LN(x) + b1 = rstd (W1' . x) - rstd mean u + b1'
// so G1 runs on the RAW tile as the TMA delivered it (no in-place normalisation, no proxy fence, nothing between the load and
// the first MMA) and the GELU warps apply the per-row scale and the rank-1 correction while they read H.  W1' . x contains
// mean u exactly (same rounded weights), so the subtraction cancels it to fp32 accumulation error; x itself is not
// re-rounded, which the unfused path does when it writes LN(x) as bf16.
//
// Warp roles (576 threads = 2 + BF_GELU_WARPS_N + BF_ROW_WARPS_N warps, one CTA per SM, persistent over 128-row tiles):
//   warp 0       TMA producer: X tile -> smem (4 k-blocks of [128 x 64], 128B swizzle), W1' / W2 boxes -> 4 x 32 KB ring
//   warp 1       tcgen05.mma issuer (whole warp, one elected lane: ptx.cuh "lean issue path"); per hidden chunk c of 128 units:
//                  G1(c): H[c%2] (TMEM fp32 128x128) = X(smem) . W1'[128c:128c+128, :]^T     16 x UMMA 128x128x16 (SS)
//                  G2(c): O (TMEM fp32 128x256)     += Hbf16(TMEM) . W2[:, 128c:128c+128]^T   8 x UMMA 128x256x16 (TS)
//                G1(c+1) is issued before G2(c)
//   GELU warps   (8 by default: thread = row x 64 hidden units, in 32-column slices): H -> rstd H - rstd mean u + b1' -> GELU ->
//                bf16 written back in place into TMEM.  They do nothing else, so the chunk pipeline never waits for
//                tile-boundary work.
//   row warps    (8 by default: thread = half a row, halves merged through smem): row statistics of tile n+1 from the smem
//                tile while G2(7) of tile n / G1(0) of tile n+1 run; then the drain of tile n:
//                O + b2 + X (bf16, re-read from L2) -> bf16 X (+ copy) and the new row's statistics.
// The elementwise work is what bounds this kernel (131 072 GELUs per tile against 16 384 cycles of MMA): it is written
// in packed fp32x2 arithmetic (FFMA2 / FMUL2 / FADD2), 7 issue slots per hidden unit.
//
// With OUTP (cvb200_block_out_ffn) the block's attention out-projection runs first, in the same kernel:
//     X' = bf16((o . W_out^T + b_out) + X)            attn1.to_out[0] + residual, components/transformer.py:266-276
// o [rows, 512] and W_out [256, 512] stream through the weight ring in 8 K steps of 64 (W_out slab, then o slab) into an
// accumulator D0 that borrows the H0 | H1 columns (free until G1(0); the issuer first waits for the previous tile's last G2,
// which still reads H).  The row warps turn D0 into the tile: X' replaces X_old in shared memory (generic-proxy writes +
// fence.proxy.async before the MMA reads them), goes to global X (this thread reads it back as the drain's residual), and its
// LayerNorm statistics are computed on the way; st_full then releases G1(0) and the GELU warps.  Arithmetic and order are those
// of the separate out-projection GEMM's epilogue: the two forms are bit-identical (tests/test_gpu_ops.py).  Measured no faster
// than the two launches, hence opt-in.
//
// TMEM: H0 [0,128) H1 [128,256) O [256,512).  smem: X tile 64 KB + 4 x 32 KB weight ring + vectors = 207 KB.
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

constexpr int BF_C = 256;            // model width
constexpr int BF_H = 1024;           // hidden width
constexpr int BF_CH = 128;           // hidden chunk
constexpr int BF_NCH = BF_H / BF_CH;
constexpr int BF_M = 128;
constexpr int BF_BK = 64;
constexpr int BF_A_BYTES = BF_M * BF_C * 2;          // 64 KiB: 4 k-blocks of [128 x 64]
constexpr int BF_STAGE_BYTES = 32 * 1024;            // G1: two [128 x 64] W1 boxes; G2: one [256 x 64] W2 box
constexpr int BF_STAGES = 4;
#ifndef BF_GELU_WARPS_N
#define BF_GELU_WARPS_N 8
#endif
#ifndef BF_ROW_WARPS_N
#define BF_ROW_WARPS_N 8
#endif
constexpr int BF_GELU_WARPS = BF_GELU_WARPS_N;        // 8 or 16: a thread owns 64 or 32 hidden units of a chunk
constexpr int BF_ROW_WARPS = BF_ROW_WARPS_N;          // 4 or 8: a thread owns a whole row or half of it
constexpr int BF_GELU_COLS = 128 / (BF_GELU_WARPS / 4);
constexpr int BF_ROW_SPLIT = BF_ROW_WARPS / 4;
constexpr int BF_ROW_COLS = 256 / BF_ROW_SPLIT;
constexpr int BF_THREADS = 32 * (2 + BF_GELU_WARPS + BF_ROW_WARPS);
constexpr int BF_TM_H = 0, BF_TM_O = 256, BF_TM_COLS = 512;
constexpr int BF_BAR_BYTES = 256;
constexpr int BF_VEC_FLOATS = 2 * BF_H + BF_C;       // b1' | u | b2
constexpr int BF_XCH_BYTES = 2 * BF_M * 8 + 2 * 2 * BF_M * 8;   // row statistics [tile parity][row] float2 + half-row exchange [use parity][half][row]
constexpr int BF_SMEM = 1024 + BF_A_BYTES + BF_STAGES * BF_STAGE_BYTES + BF_BAR_BYTES + BF_VEC_FLOATS * 4 + BF_XCH_BYTES + BF_C * 4;

#ifdef BF_PROFILE
#define BF_WAIT(acc, bar, par)             \
  do {                                     \
    const long long _t0 = clock64();       \
    ptx::mbar_wait(bar, par);              \
    acc += clock64() - _t0;                \
  } while (0)
#define BF_T0() const long long _tt = clock64()
#define BF_T1(acc) acc += clock64() - _tt
#else
#define BF_WAIT(acc, bar, par) ptx::mbar_wait(bar, par)
#define BF_T0()
#define BF_T1(acc)
#endif

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
  return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u));
}

__device__ __forceinline__ float2 f2(float a) { return make_float2(a, a); }
__device__ __forceinline__ float tanh_approx(float x) {
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(x));
  return t;
}
// The fitted tanh-form erf-GELU of the tap GEMM's bf16 epilogue (tapgemm.cuh gelu_erf_fast: x Phi(x), Phi = (1 + tanh(x (a +
// b x^2 + c x^4))) / 2, max abs error 2.5e-5) on a PAIR of values in packed fp32x2 arithmetic.  x^2 is clamped at 36 (one
// FMNMX per value instead of the two of a clamp on x): the polynomial then stays at p(36) = 1.674 and tanh saturates, which
// gives the same values as the clamp on x for |x| > 6.  11 issue slots per pair (2 MUFU) after the pre-activation.
__device__ __forceinline__ uint32_t gelu_pair_bf16(float2 x) {
  float2 x2 = __fmul2_rn(x, x);
  x2.x = fminf(x2.x, 36.f);
  x2.y = fminf(x2.y, 36.f);
  float2 p = __ffma2_rn(f2(-3.51516916e-4f), x2, f2(3.70056467e-2f));
  p = __ffma2_rn(p, x2, f2(7.97507884e-1f));
  const float2 u = __fmul2_rn(x, p);
  const float2 t = make_float2(tanh_approx(u.x), tanh_approx(u.y));
  const float2 h = __fmul2_rn(x, f2(0.5f));
  const float2 g = __ffma2_rn(h, t, h);
  return pack_bf16x2(g.x, g.y);
}

__device__ __forceinline__ void lds128(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void sts128(uint32_t addr, const uint32_t (&r)[4]) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]) : "memory");
}
__device__ __forceinline__ void lds128f(uint32_t addr, float (&r)[4]) {
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r[0]), "=f"(r[1]), "=f"(r[2]), "=f"(r[3]) : "r"(addr));
}

// OUTP: the block's attention out-projection runs first, inside this kernel (header comment, "With OUTP").
template <bool OUTP>
__global__ void __launch_bounds__(BF_THREADS, 1)
block_ffn_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW1,
                 const __grid_constant__ CUtensorMap tmW2, const __grid_constant__ CUtensorMap tmO,
                 const __grid_constant__ CUtensorMap tmWo, const float* __restrict__ bo, float eps, const float* __restrict__ u1,
                 const float* __restrict__ b1, const float* __restrict__ b2, __nv_bfloat16* __restrict__ X, int ldx,
                 __nv_bfloat16* __restrict__ out1, int ldo1, float2* __restrict__ stats_out, const Layout lay, int m_tiles) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sW = smem + BF_A_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sW + BF_STAGES * BF_STAGE_BYTES);
  uint64_t* w_full = bars;                   // [STAGES]
  uint64_t* w_empty = bars + BF_STAGES;      // [STAGES]
  uint64_t* a_full = bars + 2 * BF_STAGES;   // X tile landed (TMA)
  uint64_t* st_full = a_full + 1;            // row statistics of the tile are in smem (8 row-warp arrivals)
  uint64_t* a_free = st_full + 1;            // the tile's last G1 retired: A may be overwritten
  uint64_t* h_full = a_free + 1;             // [2] G1(c) complete
  uint64_t* p_full = h_full + 2;             // [2] bf16 H(c) written back (8 GELU-warp arrivals)
  uint64_t* o_full = p_full + 2;             // the tile's last G2 retired
  uint64_t* o_free = o_full + 1;             // O read out of TMEM by the row warps (4 arrivals)
  uint64_t* d0_full = o_free + 1;            // OUTP: the tile's out-projection accumulator is complete in TMEM
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(d0_full + 1);
  float* sb1 = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + BF_BAR_BYTES);
  float* su1 = sb1 + BF_H;
  float* sb2 = su1 + BF_H;
  float2* srow = reinterpret_cast<float2*>(sb2 + BF_C);          // [2][128] (rstd, -mean rstd) of the tile's rows
  float2* xch = srow + 2 * BF_M;                                 // [2][2][128] partial statistics of half rows
  float* sbo = reinterpret_cast<float*>(xch + 2 * 2 * BF_M);     // [256] out-projection bias (OUTP)

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < BF_STAGES; ++i) {
      ptx::mbar_init(&w_full[i], 1);
      ptx::mbar_init(&w_empty[i], 1);
    }
    ptx::mbar_init(a_full, 1);
    ptx::mbar_init(st_full, BF_ROW_WARPS);
    ptx::mbar_init(a_free, 1);
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&h_full[i], 1);
      ptx::mbar_init(&p_full[i], BF_GELU_WARPS);
    }
    ptx::mbar_init(o_full, 1);
    ptx::mbar_init(o_free, BF_ROW_WARPS);
    ptx::mbar_init(d0_full, 1);
    ptx::fence_barrier_init();
    ptx::prefetch_tensormap(&tmX);
    ptx::prefetch_tensormap(&tmW1);
    ptx::prefetch_tensormap(&tmW2);
    if (OUTP) {
      ptx::prefetch_tensormap(&tmO);
      ptx::prefetch_tensormap(&tmWo);
    }
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, BF_TM_COLS);
  for (int i = threadIdx.x; i < BF_H; i += BF_THREADS) {                              // model constants
    sb1[i] = __ldg(b1 + i);
    su1[i] = __ldg(u1 + i);
  }
  for (int i = threadIdx.x; i < BF_C; i += BF_THREADS) {
    sb2[i] = __ldg(b2 + i);
    if (OUTP) sbo[i] = __ldg(bo + i);
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  pdl_trigger();
  pdl_wait();

  const int t_first = blockIdx.x, t_step = gridDim.x;
#ifdef BF_PROFILE
  long long prof[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  const long long t_start = clock64();
#endif

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t ph = 0;
      auto next_stage = [&]() { if (++stage == BF_STAGES) { stage = 0; ph ^= 1; } };
      auto load_g1 = [&](int c) {            // W1 rows [128c, +128): two stages of two [128 x 64] boxes
        for (int h = 0; h < 2; ++h) {
          ptx::mbar_wait(&w_empty[stage], ph ^ 1);
          ptx::mbar_arrive_expect_tx(&w_full[stage], BF_STAGE_BYTES);
          uint8_t* dst = sW + stage * BF_STAGE_BYTES;
          ptx::tma_load_2d(dst, &tmW1, &w_full[stage], (2 * h) * BF_BK, c * BF_CH);
          ptx::tma_load_2d(dst + 16384, &tmW1, &w_full[stage], (2 * h + 1) * BF_BK, c * BF_CH);
          next_stage();
        }
      };
      auto load_g2 = [&](int c) {            // W2[:, 128c + 64j, +64): two stages of one [256 x 64] box
        for (int j = 0; j < 2; ++j) {
          ptx::mbar_wait(&w_empty[stage], ph ^ 1);
          ptx::mbar_arrive_expect_tx(&w_full[stage], BF_STAGE_BYTES);
          ptx::tma_load_2d(sW + stage * BF_STAGE_BYTES, &tmW2, &w_full[stage], c * BF_CH + j * BF_BK, 0);
          next_stage();
        }
      };
      auto load_x = [&](int t) {
        ptx::mbar_arrive_expect_tx(a_full, BF_A_BYTES);
        for (int kb = 0; kb < BF_C / BF_BK; ++kb)
          ptx::tma_load_2d(sA + kb * 16384, &tmX, a_full, kb * BF_BK, t * BF_M);
      };
      // Weight boxes in the order the MMA warp consumes them: G1(0), then G1(c+1), G2(c).  The NEXT tile's X is requested
      // as soon as this tile's last G1 has retired (a_free, committed right after G1(7) is issued), behind the boxes of
      // G2(6): the ring is full at that point, so the wait costs the weight stream nothing, and X lands -- and is
      // normalised -- while GELU(7) / G2(7) run.
      auto load_g0 = [&](int t) {            // out-projection operands, 8 K steps of 64: W_out[:, 64k, +64) then o[tile, 64k, +64)
        for (int ks = 0; ks < 512 / BF_BK; ++ks) {
          ptx::mbar_wait(&w_empty[stage], ph ^ 1);
          ptx::mbar_arrive_expect_tx(&w_full[stage], BF_STAGE_BYTES);
          ptx::tma_load_2d(sW + stage * BF_STAGE_BYTES, &tmWo, &w_full[stage], ks * BF_BK, 0);
          next_stage();
          ptx::mbar_wait(&w_empty[stage], ph ^ 1);
          ptx::mbar_arrive_expect_tx(&w_full[stage], BF_STAGE_BYTES / 2);
          ptx::tma_load_2d(sW + stage * BF_STAGE_BYTES, &tmO, &w_full[stage], ks * BF_BK, t * BF_M);
          next_stage();
        }
      };
      int n = 0;
      load_x(t_first);
      for (int t = t_first; t < m_tiles; t += t_step, ++n) {
        if (OUTP) load_g0(t);
        load_g1(0);
        for (int c = 0; c < BF_NCH; ++c) {
          if (c + 1 < BF_NCH) load_g1(c + 1);
          load_g2(c);
          if (c == BF_NCH - 2 && t + t_step < m_tiles) {
            ptx::mbar_wait(a_free, n & 1);
            load_x(t + t_step);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    // whole warp, uniform state, one elected lane issues (ptx.cuh "lean issue path": with the loop inside `if (lane == 0)`
    // every UTCHMMA carried a ~20-instruction uniform-register conversion loop and G1's 64-cycle MMAs were issue-bound)
    {
      const uint32_t idesc1 = ptx::make_idesc_bf16(BF_M, BF_CH);       // 128 x 128
      const uint32_t idesc2 = ptx::make_idesc_bf16(BF_M, BF_C);        // 128 x 256
      const uint32_t desc_hi = (uint32_t)(ptx::make_sw128_kmajor_desc(0) >> 32);
      const uint32_t a_lo0 = (uint32_t)ptx::make_sw128_kmajor_desc(ptx::smem_u32(sA));
      const uint32_t w_lo0 = (uint32_t)ptx::make_sw128_kmajor_desc(ptx::smem_u32(sW));
      int stage = 0;
      uint32_t ph = 0;
      auto next_stage = [&]() { if (++stage == BF_STAGES) { stage = 0; ph ^= 1; } };
      uint32_t hcount[2] = {0, 0};           // completed uses of each H buffer (for the p_full parity)
      auto issue_g1 = [&](int c, bool free_a) {
        const uint32_t d = tmem + BF_TM_H + (c & 1) * BF_CH;
        for (int h = 0; h < 2; ++h) {
          BF_WAIT(prof[4], &w_full[stage], ph);
          ptx::tc_fence_after();
          const uint32_t w_lo = ptx::bcast0(w_lo0 + stage * (BF_STAGE_BYTES >> 4));
          const uint32_t a_lo = ptx::bcast0(a_lo0 + 2 * h * (16384 >> 4));
          const uint32_t hu = ptx::bcast0(h), fa = ptx::bcast0(free_a);
          if (ptx::elect_one()) {
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {       // 2 k-blocks x 4 K16 steps
              const uint32_t off = (kk >> 2) * (16384 >> 4) + (kk & 3) * 2;
              ptx::umma_ss_lo(d, a_lo + off, desc_hi, w_lo + off, desc_hi, idesc1, kk == 0 ? hu : 1u);
            }
            ptx::umma_commit(&w_empty[stage]);
            if (hu) {
              ptx::umma_commit(&h_full[c & 1]);
              if (fa) ptx::umma_commit(a_free);    // all G1 of this tile are issued: A is free when they retire
            }
          }
          __syncwarp();
          next_stage();
        }
      };
      auto issue_g0 = [&]() {                 // D0 (TMEM columns [0, 256) = H0 | H1) = o . W_out^T, K = 512
        for (int ks = 0; ks < 512 / BF_BK; ++ks) {
          BF_WAIT(prof[4], &w_full[stage], ph);
          const int st_w = stage;
          next_stage();
          BF_WAIT(prof[4], &w_full[stage], ph);
          const int st_o = stage;
          next_stage();
          ptx::tc_fence_after();
          const uint32_t wo_lo = ptx::bcast0(w_lo0 + st_w * (BF_STAGE_BYTES >> 4));
          const uint32_t o_lo = ptx::bcast0(w_lo0 + st_o * (BF_STAGE_BYTES >> 4));
          const uint32_t first = ptx::bcast0(ks == 0), last = ptx::bcast0(ks + 1 == 512 / BF_BK);
          if (ptx::elect_one()) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
              ptx::umma_ss_lo(tmem + BF_TM_H, o_lo + kk * 2, desc_hi, wo_lo + kk * 2, desc_hi, idesc2, (kk == 0 && first) ? 0u : 1u);
            ptx::umma_commit(&w_empty[st_w]);
            ptx::umma_commit(&w_empty[st_o]);
            if (last) ptx::umma_commit(d0_full);
          }
          __syncwarp();
        }
      };
      int n = 0;
      for (int t = t_first; t < m_tiles; t += t_step, ++n) {
        if (OUTP) {
          // the accumulator shares H0 | H1 with the GELU outputs the previous tile's last G2 still reads
          if (n > 0) BF_WAIT(prof[0], o_full, (n - 1) & 1);
          ptx::tc_fence_after();
          issue_g0();
          BF_WAIT(prof[0], st_full, n & 1);       // the row warps have turned D0 into the bf16 X tile in shared memory
        } else {
          BF_WAIT(prof[0], a_full, n & 1);        // the raw tile is the operand: nothing sits between the load and G1(0)
        }
        ptx::tc_fence_after();
        issue_g1(0, false);
        for (int c = 0; c < BF_NCH; ++c) {
          if (c + 1 < BF_NCH) issue_g1(c + 1, c + 2 == BF_NCH);
          const int hb = c & 1;
          BF_WAIT(prof[1], &p_full[hb], hcount[hb] & 1);
          ++hcount[hb];
          if (c == 0 && n > 0) BF_WAIT(prof[2], o_free, (n - 1) & 1);   // the previous tile's O has been read out
          ptx::tc_fence_after();
          const uint32_t hbase = tmem + BF_TM_H + hb * BF_CH;
          for (int j = 0; j < 2; ++j) {
            BF_WAIT(prof[3], &w_full[stage], ph);
            ptx::tc_fence_after();
            const uint32_t w_lo = ptx::bcast0(w_lo0 + stage * (BF_STAGE_BYTES >> 4));
            const uint32_t cj = ptx::bcast0((c | j) != 0), last = ptx::bcast0(c + 1 == BF_NCH && j == 1);
            const uint32_t hj = ptx::bcast0(hbase + 64 * j);
            if (ptx::elect_one()) {
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                // K16 step u = 4j + k of the chunk: hidden units [16u, 16u+16) live in packed columns
                // 32*(u/2) + 8*(u%2) (the 32 units of a 32-column slice are packed into its first 16 columns)
                ptx::umma_ts_lo(tmem + BF_TM_O, hj + 32 * (k >> 1) + 8 * (k & 1), w_lo + k * 2, desc_hi, idesc2, k == 0 ? cj : 1u);
              }
              ptx::umma_commit(&w_empty[stage]);
              if (last) ptx::umma_commit(o_full);
            }
            __syncwarp();
            next_stage();
          }
        }
      }
    }
  } else if (warp < 2 + BF_GELU_WARPS) {
    // ------------------------------------------------------------------ GELU warps
    const int q = warp & 3;                    // TMEM lane quadrant this warp may access
    const int sub = (warp - 2) >> 2;           // this thread's BF_GELU_COLS-column slice of an H chunk
    const uint32_t tbase = tmem + ((uint32_t)(q * 32) << 16);
    const uint32_t sb1_addr = ptx::smem_u32(sb1), su1_addr = ptx::smem_u32(su1);
    const int trow = q * 32 + lane;
    uint32_t hcount[2] = {0, 0};
    int n = 0;
    for (int t = t_first; t < m_tiles; t += t_step, ++n) {
      BF_WAIT(prof[10], st_full, n & 1);
      const float2 st = srow[(n & 1) * BF_M + trow];
      const float2 rs = f2(st.x), rb = f2(st.y);         // pre-activation = rs * H + rb * u + b1'
      for (int c = 0; c < BF_NCH; ++c) {
        const int hb = c & 1;
        BF_WAIT(prof[7], &h_full[hb], hcount[hb] & 1);
        ++hcount[hb];
        ptx::tc_fence_after();
#pragma unroll
        for (int s2 = 0; s2 < BF_GELU_COLS / 32; ++s2) {
          // one 32-column slice at a time (its 32 units go back packed into its first 16 columns): only 32 raw values are
          // live, which leaves the registers the compiler needs to keep several GELU chains in flight
          const int col = sub * BF_GELU_COLS + s2 * 32;
          const uint32_t haddr = tbase + BF_TM_H + hb * BF_CH + col;
          uint32_t r[32];
          ptx::tmem_ld32_nowait(haddr, r);
          ptx::tmem_wait_ld();
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            float bb[4], uu[4];
            const uint32_t off = (c * BF_CH + col + i) * 4;
            lds128f(sb1_addr + off, bb);
            lds128f(su1_addr + off, uu);
            const float2 x0 = __ffma2_rn(make_float2(__uint_as_float(r[i]), __uint_as_float(r[i + 1])), rs,
                                         __ffma2_rn(rb, make_float2(uu[0], uu[1]), make_float2(bb[0], bb[1])));
            const float2 x1 = __ffma2_rn(make_float2(__uint_as_float(r[i + 2]), __uint_as_float(r[i + 3])), rs,
                                         __ffma2_rn(rb, make_float2(uu[2], uu[3]), make_float2(bb[2], bb[3])));
            pk[i >> 1] = gelu_pair_bf16(x0);
            pk[(i >> 1) + 1] = gelu_pair_bf16(x1);
          }
          ptx::tmem_st16(haddr, pk);
        }
        ptx::tmem_wait_st();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&p_full[hb]);
      }
    }
  } else {
    // ------------------------------------------------------------------ row warps: row statistics + drain
    // thread = row trow (its TMEM lane) x BF_ROW_COLS columns.  With eight row warps the two halves of a row live in the two
    // warps of a lane quadrant and merge their partial statistics through smem (named barrier per quadrant, 64 threads).
    const int q = warp & 3;
    const int hf = (warp - 2 - BF_GELU_WARPS) >> 2;                          // 0 .. BF_ROW_SPLIT-1
    const int trow = q * 32 + lane;
    const int c0 = hf * BF_ROW_COLS;                                         // first column of this thread's part
    const uint32_t tbase = tmem + ((uint32_t)(q * 32) << 16);
    const uint32_t a_row = ptx::smem_u32(sA) + (c0 / 64) * 16384 + trow * 128;
    const int swz = trow & 7;
    const uint32_t sb2_addr = ptx::smem_u32(sb2) + c0 * 4;
    int xuse = 0;

    // (pivot, sum of (x - pivot), sum of (x - pivot)^2) over this thread's columns -> (rstd, -mean rstd) of the whole row.
    // The shift by the first value removes the cancellation of a common offset; halves are merged with Chan's formula.
    auto finish = [&](float pivot, float2 s, float2 ss) {
      constexpr float inv_n = 1.f / BF_ROW_COLS;
      const float sd = s.x + s.y;
      float mean = pivot + sd * inv_n;
      float m2 = fmaxf((ss.x + ss.y) - sd * sd * inv_n, 0.f);
      if (BF_ROW_SPLIT == 2) {
        float2* xp = xch + (xuse & 1) * 2 * BF_M;
        ++xuse;
        xp[hf * BF_M + trow] = make_float2(mean, m2);
        ptx::named_bar_sync(1 + q, 64);
        const float2 o = xp[(hf ^ 1) * BF_M + trow];
        const float d = mean - o.x;
        m2 = m2 + o.y + (0.5f * BF_ROW_COLS) * d * d;
        mean = 0.5f * (mean + o.x);
      }
      const float rstd = rsqrtf(m2 * (1.f / 256.f) + eps);
      return make_float2(rstd, -mean * rstd);
    };

    // norm3 statistics of row trow of the X tile (read-only pass over the swizzled smem tile)
    auto stats_tile = [&](int n) {
      BF_WAIT(prof[5], a_full, n & 1);
      BF_T0();
      float2 s = f2(0.f), ss = f2(0.f);
      float pivot = 0.f;
#pragma unroll
      for (int kb = 0; kb < BF_ROW_COLS / 64; ++kb) {
        uint32_t r[8][4];
#pragma unroll
        for (int j = 0; j < 8; ++j) lds128(a_row + kb * 16384 + ((j ^ swz) << 4), r[j]);
        if (kb == 0) pivot = unpack_bf16x2(r[0][0]).x;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 d = __fadd2_rn(unpack_bf16x2(r[j][e]), f2(-pivot));
            s = __fadd2_rn(s, d);
            ss = __ffma2_rn(d, d, ss);
          }
        }
      }
      const float2 st = finish(pivot, s, ss);
      if (hf == 0) srow[(n & 1) * BF_M + trow] = st;
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(st_full);
      BF_T1(prof[6]);
    };

    auto ldg256 = [](const __nv_bfloat16* p, uint32_t (&r)[8]) {
      asm volatile("ld.global.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                   : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                   : "l"(p));
    };
    auto stg256 = [](__nv_bfloat16* p, const uint32_t (&r)[8]) {
      asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(r[0]), "r"(r[1]), "r"(r[2]),
                   "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
                   : "memory");
    };

    constexpr int NG = BF_ROW_COLS / 32;       // 32-column groups per thread
    int n = 0;
    if (!OUTP) stats_tile(0);
    const uint32_t sbo_addr = ptx::smem_u32(sbo) + c0 * 4;
    for (int t = t_first; t < m_tiles; t += t_step, ++n) {
      const int row = t * BF_M + trow;
      const int b = __ldg(lay.tile_seq + t);
      const bool valid = __ldg(lay.tile_row0 + t) + trow < __ldg(lay.seq_len + b);
      __nv_bfloat16* xr = X + (size_t)row * ldx + c0;
      uint32_t res[2][2][8];                   // [group parity][16-column half][8 packed pairs]
      if constexpr (OUTP) {
        // ---- out-projection epilogue: X' = bf16((D0 + b_out) + X_old), written over the X_old tile in shared memory (the
        // A operand of every G1 of this tile) and to X in global memory (L2), where this thread reads it back as the residual
        // of the drain below (keeping it in 64 registers spilled), plus its LayerNorm statistics.  Same arithmetic, in the
        // same order, as the separate out-projection GEMM's epilogue.
        BF_WAIT(prof[5], a_full, n & 1);
        BF_WAIT(prof[5], d0_full, n & 1);
        ptx::tc_fence_after();
        BF_T0();
        float2 s = f2(0.f), ss = f2(0.f);
        float pivot = 0.f;
#pragma unroll
        for (int k = 0; k < NG; ++k) {
          float v[32];
          ptx::tmem_ld32(tbase + BF_TM_H + c0 + k * 32, v);
          const uint32_t slab = a_row + (k >> 1) * 16384;
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) {
            const uint32_t addr = slab + (((((k & 1) << 2) + jj) ^ swz) << 4);     // 16-byte chunk = 8 columns of the 64-wide slab
            uint32_t r[4], y[4];
            float bb[8];
            lds128(addr, r);
            lds128f(sbo_addr + (k * 32 + jj * 8) * 4, reinterpret_cast<float(&)[4]>(bb[0]));
            lds128f(sbo_addr + (k * 32 + jj * 8 + 4) * 4, reinterpret_cast<float(&)[4]>(bb[4]));
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float2 yv = __fadd2_rn(__fadd2_rn(make_float2(v[jj * 8 + 2 * e], v[jj * 8 + 2 * e + 1]),
                                                      make_float2(bb[2 * e], bb[2 * e + 1])), unpack_bf16x2(r[e]));
              y[e] = valid ? pack_bf16x2(yv.x, yv.y) : 0u;
            }
            if (k == 0 && jj == 0) pivot = unpack_bf16x2(y[0]).x;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float2 d = __fadd2_rn(unpack_bf16x2(y[e]), f2(-pivot));
              s = __fadd2_rn(s, d);
              ss = __ffma2_rn(d, d, ss);
            }
            sts128(addr, y);
            *reinterpret_cast<uint4*>(xr + k * 32 + jj * 8) = make_uint4(y[0], y[1], y[2], y[3]);
          }
        }
        const float2 st = finish(pivot, s, ss);
        if (hf == 0) srow[(n & 1) * BF_M + trow] = st;
        ptx::fence_proxy_async();              // the tile was written by threads, the MMA reads it through the async proxy
        ptx::tc_fence_before();                // D0 is in registers: G1(0) may overwrite H0
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(st_full);
        BF_T1(prof[6]);
      } else {
        // ---- the next tile's row statistics first: the GELU warps need them at its chunk 0
        if (t + t_step < m_tiles) stats_tile(n + 1);
      }
      // ---- drain: O + b2 + X -> X (bf16) in 32-column groups; the residual (L2-resident: the TMA read it a tile ago, or
      // this thread has just written it) is requested two groups ahead, the first two before the wait on the last GEMM
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        ldg256(xr + k * 32, res[k][0]);
        ldg256(xr + k * 32 + 16, res[k][1]);
      }
      BF_WAIT(prof[8], o_full, n & 1);
      ptx::tc_fence_after();
#ifdef BF_PROFILE
      const long long _td = clock64();
#endif
      float2 s = f2(0.f), ss = f2(0.f);
      float pivot = 0.f;
#pragma unroll
      for (int k = 0; k < NG; ++k) {
        float v[32];
        ptx::tmem_ld32(tbase + BF_TM_O + c0 + k * 32, v);
        if (k == NG - 1) {                     // this thread's part of O is in registers: the next tile's G2 may overwrite it
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(o_free);
        }
#pragma unroll
        for (int g16 = 0; g16 < 2; ++g16) {
          uint32_t y[8];
#pragma unroll
          for (int i = 0; i < 16; i += 4) {
            float bb[4];
            lds128f(sb2_addr + (k * 32 + g16 * 16 + i) * 4, bb);
            const uint32_t r0 = res[k & 1][g16][i >> 1];
            const uint32_t r1 = res[k & 1][g16][(i >> 1) + 1];
            const float2 y0 = __fadd2_rn(__fadd2_rn(make_float2(v[g16 * 16 + i], v[g16 * 16 + i + 1]), make_float2(bb[0], bb[1])),
                                         unpack_bf16x2(r0));
            const float2 y1 = __fadd2_rn(__fadd2_rn(make_float2(v[g16 * 16 + i + 2], v[g16 * 16 + i + 3]), make_float2(bb[2], bb[3])),
                                         unpack_bf16x2(r1));
            y[i >> 1] = valid ? pack_bf16x2(y0.x, y0.y) : 0u;
            y[(i >> 1) + 1] = valid ? pack_bf16x2(y1.x, y1.y) : 0u;
          }
          if (stats_out) {                     // statistics of the ROUNDED stream (what norm1 of the next block reads)
            if (k == 0 && g16 == 0) pivot = unpack_bf16x2(y[0]).x;
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const float2 d = __fadd2_rn(unpack_bf16x2(y[e]), f2(-pivot));
              s = __fadd2_rn(s, d);
              ss = __ffma2_rn(d, d, ss);
            }
          }
          stg256(xr + k * 32 + g16 * 16, y);
          if (out1) stg256(out1 + (size_t)row * ldo1 + c0 + k * 32 + g16 * 16, y);
        }
        if (k + 2 < NG) {
          ldg256(xr + (k + 2) * 32, res[k & 1][0]);
          ldg256(xr + (k + 2) * 32 + 16, res[k & 1][1]);
        }
      }
      if (stats_out) {
        const float2 st = finish(pivot, s, ss);
        if (hf == 0) stats_out[row] = valid ? st : make_float2(0.f, 0.f);
      }
#ifdef BF_PROFILE
      prof[9] += clock64() - _td;
#endif
    }
  }
#ifdef BF_PROFILE
  if (blockIdx.x == 0 && lane == 0 && (warp == 1 || warp == 2 || warp == 2 + BF_GELU_WARPS))
    printf("block_ffn prof warp %d total %lld | mma: a_full %lld p_full %lld o_free %lld g2 w_full %lld g1 w_full %lld | "
           "row: a_full %lld stats %lld | gelu: h_full %lld st_full %lld | row: o_full %lld drain %lld\n", warp, clock64() - t_start,
           prof[0], prof[1], prof[2], prof[3], prof[4], prof[5], prof[6], prof[7], prof[10], prof[8], prof[9]);
#endif
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem, BF_TM_COLS);
}

}  // namespace

// o / wo / bo == nullptr: norm3 + FeedForward + residual only (cvb200_block_ffn); otherwise the attention out-projection of the
// block runs first inside the same kernel (cvb200_block_out_ffn).
static int launch_block_ffn_impl(void* x, int ldx, const void* o, int ldo, const void* wo, const float* bo, float eps,
                                 const void* w1f, const float* u1, const float* b1f, const void* w2, const float* b2, void* out1,
                                 int ldo1, float* stats_out, const Layout& lay, cudaStream_t st) {
  CVB_REQUIRE(x && w1f && u1 && b1f && w2 && b2, "null operand");
  CVB_REQUIRE(lay.rows > 0 && lay.rows % BF_M == 0, "rows must be a positive multiple of 128");
  CVB_REQUIRE(ldx >= BF_C && ldx % 16 == 0 && (reinterpret_cast<uintptr_t>(x) & 31) == 0,
              "the residual stream must be 32-byte aligned with a 32-byte row pitch");
  CVB_REQUIRE(!out1 || (ldo1 >= BF_C && ldo1 % 16 == 0 && (reinterpret_cast<uintptr_t>(out1) & 31) == 0), "out1 alignment");
  CVB_REQUIRE((reinterpret_cast<uintptr_t>(stats_out) & 7) == 0, "stats_out must be 8-byte aligned");
  const bool outp = o != nullptr;
  CVB_REQUIRE(!outp || (wo && bo && ldo >= 512 && ldo % 8 == 0 && (reinterpret_cast<uintptr_t>(o) & 15) == 0),
              "out-projection operands: o bf16 [rows, ldo >= 512] 16-byte aligned, W_out [256, 512], b_out [256]");
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [] {
    attr_err = cudaFuncSetAttribute(block_ffn_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, BF_SMEM);
    if (attr_err == cudaSuccess)
      attr_err = cudaFuncSetAttribute(block_ffn_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, BF_SMEM);
  });
  CVB_CUDA(attr_err);
  CUtensorMap tmX, tmW1, tmW2, tmO, tmWo;
  CVB_TRY(make_map_2d_bf16(&tmX, x, BF_C, (uint64_t)lay.rows, (uint64_t)ldx, BF_BK, BF_M));
  CVB_TRY(make_map_2d_bf16(&tmW1, w1f, BF_C, BF_H, BF_C, BF_BK, BF_CH));
  CVB_TRY(make_map_2d_bf16(&tmW2, w2, BF_H, BF_C, BF_H, BF_BK, BF_C));
  if (outp) {
    CVB_TRY(make_map_2d_bf16(&tmO, o, 512, (uint64_t)lay.rows, (uint64_t)ldo, BF_BK, BF_M));
    CVB_TRY(make_map_2d_bf16(&tmWo, wo, 512, BF_C, 512, BF_BK, BF_C));
  } else {
    tmO = tmX;
    tmWo = tmX;
  }
  const int m_tiles = lay.rows / BF_M;
  const int sms = device_sm_count();
  const int grid = m_tiles < sms ? m_tiles : sms;
  if (outp)
    CVB_CUDA(launch_pdl(block_ffn_kernel<true>, dim3(grid), dim3(BF_THREADS), BF_SMEM, st, tmX, tmW1, tmW2, tmO, tmWo, bo, eps, u1,
                        b1f, b2, reinterpret_cast<__nv_bfloat16*>(x), ldx, reinterpret_cast<__nv_bfloat16*>(out1), ldo1,
                        reinterpret_cast<float2*>(stats_out), lay, m_tiles));
  else
    CVB_CUDA(launch_pdl(block_ffn_kernel<false>, dim3(grid), dim3(BF_THREADS), BF_SMEM, st, tmX, tmW1, tmW2, tmO, tmWo, bo, eps, u1,
                        b1f, b2, reinterpret_cast<__nv_bfloat16*>(x), ldx, reinterpret_cast<__nv_bfloat16*>(out1), ldo1,
                        reinterpret_cast<float2*>(stats_out), lay, m_tiles));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_block_ffn(void* x, int ldx, float eps, const void* w1f, const float* u1, const float* b1f, const void* w2,
                     const float* b2, void* out1, int ldo1, float* stats_out, const Layout& lay, cudaStream_t st) {
  return launch_block_ffn_impl(x, ldx, nullptr, 0, nullptr, nullptr, eps, w1f, u1, b1f, w2, b2, out1, ldo1, stats_out, lay, st);
}

int launch_block_out_ffn(void* x, int ldx, const void* o, int ldo, const void* wo, const float* bo, float eps, const void* w1f,
                         const float* u1, const float* b1f, const void* w2, const float* b2, void* out1, int ldo1,
                         float* stats_out, const Layout& lay, cudaStream_t st) {
  CVB_REQUIRE(o && wo && bo, "null out-projection operand");
  return launch_block_ffn_impl(x, ldx, o, ldo, wo, bo, eps, w1f, u1, b1f, w2, b2, out1, ldo1, stats_out, lay, st);
}

}  // namespace cvb
