// tcgen05 engine of the tap GEMM (bf16 operands, fp32 accumulation in TMEM).
//
// Persistent, warp-specialised, one CTA per SM:
//   warp 0   : TMA producer. One 128x64 A box per (tap segment, 64-column chunk) at row coordinate
//              m0 + seg_row (the conv tap shift; rows outside the view are zero-filled by TMA = the
//              conv padding) and one bn x 64 box of the packed weights, into a 4-stage smem ring.
//   warp 1   : TMEM allocator + single-thread tcgen05.mma issuer (UMMA 128 x bn x 16, SS operands,
//              128B-swizzled K-major descriptors), double-buffered accumulators (2 x 256 columns).
//   warps 2-9: epilogue (two warps per TMEM lane group, column chunks interleaved). tcgen05.ld the accumulator rows, fused bias / time-bias / activation /
//              residual / scale / accumulate / dual store / GroupNorm statistics (tapgemm.cuh).
// The epilogue of tile i overlaps the MMAs of tile i+1 through the two TMEM buffers.
#include <cuda.h>
#include <mutex>
#include "ptx.cuh"
#include "tapgemm.cuh"

namespace cvb {

namespace {

#ifndef CVB_GEMM_STAGES
#define CVB_GEMM_STAGES 4
#endif
constexpr int STAGES = CVB_GEMM_STAGES;
constexpr int A_STAGE_BYTES = TILE_M * GEMM_BK * 2;        // 16 KiB
constexpr int B_STAGE_BYTES = 256 * GEMM_BK * 2;           // 32 KiB (bn <= 256)
constexpr int BAR_BYTES = 256;
constexpr int SMEM_BYTES = 1024 + STAGES * (A_STAGE_BYTES + B_STAGE_BYTES) + BAR_BYTES;
constexpr int EPI_WARPS = 16;                                // 2 per TMEM lane group: chunks interleaved between them
constexpr int NUM_THREADS = 64 + 32 * EPI_WARPS;
constexpr int TMEM_COLS = 512;

// One accumulator row (this thread's TMEM lane) of a tile: NV columns at a time.
template <int NV, bool RP>
__device__ __forceinline__ void epilogue_chunks(const GemmParams& p, uint32_t taddr, int bn, int n_base, size_t m, int b,
                                                bool valid, int phase, int lane, int half) {
  float gs = 0.f, gq = 0.f;
  for (int c = half * NV; c < bn; c += (EPI_WARPS / 4) * NV) {
    float v[NV];
    if constexpr (NV == 32) ptx::tmem_ld32(taddr + c, v);
    else ptx::tmem_ld16(taddr + c, v);
    epi_apply<NV, false, RP>(p.epi, m, b, valid, phase, n_base + c, v, gs, gq);
    if (RP && p.epi.row_parts_out && ((c + NV) & 31) == 0) {
      // this thread's row, one finished 32-column chunk: (sum, sum of squares) for the consumer's folded LayerNorm
      reinterpret_cast<float2*>(p.epi.row_parts_out)[m * 8 + ((n_base + c) >> 5)] = make_float2(gs, gq);
      gs = gq = 0.f;
    }
    if (p.epi.gn_stats && ((c + NV) & 31) == 0) {
      // one 32-channel GroupNorm group finished: reduce over the warp's 32 rows, one atomic pair per warp
      const float s = warp_sum(gs), q = warp_sum(gq);
      if (lane == 0) {
        double* dst = p.epi.gn_stats + ((size_t)b * p.epi.gn_groups + ((n_base + c) >> 5)) * 2;
        atomicAdd(dst, (double)s);
        atomicAdd(dst + 1, (double)q);
      }
      gs = gq = 0.f;
    }
  }
}

template <bool RP>
__global__ void __launch_bounds__(NUM_THREADS, 1)
tapgemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                  const __grid_constant__ GemmParams p, const int resident) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + STAGES * A_STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sB + STAGES * B_STAGE_BYTES);
  uint64_t* full = bars;                 // [STAGES]
  uint64_t* empty = bars + STAGES;       // [STAGES]
  uint64_t* tfull = bars + 2 * STAGES;   // [2]
  uint64_t* tempty = tfull + 2;          // [2]
  uint64_t* w_full = tempty + 2;         // [1] resident-W mode: the whole weight tile has landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(w_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bn = p.bn;
  const int n_tiles = p.n_pad / bn;
  const int chunks_per_seg = p.seg_len / GEMM_BK;
  // Work items.  Streaming mode: item -> (m tile, phase, n tile), round-robin over CTAs, A and W boxes stream
  // through the ring.  Resident mode (the [bn x K] weight tile fits the 128 KiB B region): the CTA owns ONE
  // (phase, n tile), loads its weights once and streams only A tiles -- 3x less L2->SM traffic for the K=256
  // projections (QKV, FFN-in), which are otherwise bound by re-fetching W per tile.
  const int keys = p.n_phase * n_tiles;
  const int it0 = resident ? (int)blockIdx.x / keys : (int)blockIdx.x;
  const int it_step = resident ? (int)gridDim.x / keys : (int)gridDim.x;
  const int it_end = resident ? p.m_tiles : p.m_tiles * keys;
  const int key_phase = ((int)blockIdx.x % keys) / n_tiles, key_nt = ((int)blockIdx.x % keys) % n_tiles;
  auto decode = [&](int it, int& mt, int& phase, int& nt) {
    if (resident) { mt = it; phase = key_phase; nt = key_nt; }
    else { nt = it % n_tiles; const int rest = it / n_tiles; phase = rest % p.n_phase; mt = rest / p.n_phase; }
  };

  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) {
      ptx::mbar_init(&full[i], 1);
      ptx::mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&tfull[i], 1);
      ptx::mbar_init(&tempty[i], 32 * EPI_WARPS);
    }
    ptx::mbar_init(w_full, 1);
    ptx::fence_barrier_init();
    ptx::prefetch_tensormap(&tmA);
    ptx::prefetch_tensormap(&tmW);
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, TMEM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // Programmatic dependent launch: everything above (and the weight preload below: W is a constant of the model,
  // never written by a kernel) overlaps the tail of the previous kernel; activations are touched after pdl_wait().
  pdl_trigger();
  if (warp != 0) pdl_wait();

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t ph = 0;
      const uint32_t w_box_bytes = bn * GEMM_BK * 2;
      if (resident && it0 < it_end) {
        const int kblocks = p.n_seg * chunks_per_seg;
        ptx::mbar_arrive_expect_tx(w_full, kblocks * w_box_bytes);
        for (int kb = 0; kb < kblocks; ++kb)
          ptx::tma_load_2d(sB + kb * w_box_bytes, &tmW, w_full, kb * GEMM_BK, key_phase * p.n_pad + key_nt * bn);
      }
      pdl_wait();
      for (int it = it0; it < it_end; it += it_step) {
        int mt, phase, nt;
        decode(it, mt, phase, nt);
        const int m0 = mt * TILE_M;
        const int wrow = phase * p.n_pad + nt * bn;
        const uint32_t tx_bytes = A_STAGE_BYTES + (resident ? 0u : w_box_bytes);
        for (int seg = 0; seg < p.n_seg; ++seg) {
          const int srow = p.seg_row[phase * p.n_seg + seg], scol = p.seg_col[phase * p.n_seg + seg];
          for (int c = 0; c < chunks_per_seg; ++c) {
            ptx::mbar_wait(&empty[stage], ph ^ 1);
            ptx::mbar_arrive_expect_tx(&full[stage], tx_bytes);
            ptx::tma_load_2d(sA + stage * A_STAGE_BYTES, &tmA, &full[stage], scol + c * GEMM_BK, m0 + srow);
            if (!resident)
              ptx::tma_load_2d(sB + stage * B_STAGE_BYTES, &tmW, &full[stage], seg * p.seg_len + c * GEMM_BK, wrow);
            if (++stage == STAGES) { stage = 0; ph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    // whole warp, uniform state, one elected lane issues (ptx.cuh "lean issue path")
    {
      const uint32_t idesc = ptx::make_idesc_bf16(TILE_M, bn);
      const int kblocks = p.n_seg * chunks_per_seg;
      const uint32_t desc_hi = (uint32_t)(ptx::make_sw128_kmajor_desc(0) >> 32);
      const uint32_t a_lo0 = (uint32_t)ptx::make_sw128_kmajor_desc(ptx::smem_u32(sA));
      const uint32_t b_lo0 = (uint32_t)ptx::make_sw128_kmajor_desc(ptx::smem_u32(sB));
      const uint32_t b_step = (resident ? bn * GEMM_BK * 2 : B_STAGE_BYTES) >> 4;
      int stage = 0;
      uint32_t ph = 0;
      int acc = 0;
      uint32_t acc_ph = 0;
      if (resident && it0 < it_end) ptx::mbar_wait(w_full, 0);
      for (int it = it0; it < it_end; it += it_step) {
        ptx::mbar_wait(&tempty[acc], acc_ph ^ 1);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * 256;
        for (int kb = 0; kb < kblocks; ++kb) {
          ptx::mbar_wait(&full[stage], ph);
          ptx::tc_fence_after();
          const uint32_t a_lo = ptx::bcast0(a_lo0 + stage * (A_STAGE_BYTES >> 4));
          const uint32_t b_lo = ptx::bcast0(b_lo0 + (resident ? kb : stage) * b_step);
          const uint32_t first = ptx::bcast0(kb != 0);
          if (ptx::elect_one()) {
#pragma unroll
            for (int k = 0; k < GEMM_BK / 16; ++k)
              ptx::umma_ss_lo(d_tmem, a_lo + k * 2, desc_hi, b_lo + k * 2, desc_hi, idesc, k == 0 ? first : 1u);
            ptx::umma_commit(&empty[stage]);      // frees the smem slot when these MMAs retire
            if (kb + 1 == kblocks) ptx::umma_commit(&tfull[acc]);   // accumulator complete -> epilogue
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; ph ^= 1; }
        }
        acc ^= 1;
        if (acc == 0) acc_ph ^= 1;
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..9)
    const int lane_grp = warp & 3;               // TMEM lanes [32*lane_grp, +32) are accessible to this warp
    const int half = (warp - 2) >> 2;            // which of the interleaved column chunks this warp takes
    const int row = lane_grp * 32 + lane;
    int acc = 0;
    uint32_t acc_ph = 0;
    for (int it = it0; it < it_end; it += it_step) {
      int mt, phase, nt;
      decode(it, mt, phase, nt);
      const int b = __ldg(p.tile_seq + mt);
      const int lrow = __ldg(p.tile_row0 + mt) + row;
      const int slen = __ldg(p.seq_len + b);
      const int olim = p.out_limit ? __ldg(p.out_limit + b) : slen * p.n_phase;
      const bool valid = lrow < slen && (lrow * p.n_phase + phase) < olim;
      const int m = mt * TILE_M + row;

      ptx::mbar_wait(&tfull[acc], acc_ph);
      ptx::tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(lane_grp * 32) << 16) + acc * 256;
      if ((bn & 31) == 0) epilogue_chunks<32, RP>(p, taddr, bn, nt * bn, (size_t)m, b, valid, phase, lane, half);
      else epilogue_chunks<16, RP>(p, taddr, bn, nt * bn, (size_t)m, b, valid, phase, lane, half);
      ptx::tc_fence_before();
      ptx::mbar_arrive(&tempty[acc]);
      acc ^= 1;
      if (acc == 0) acc_ph ^= 1;
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_base, TMEM_COLS);
}

// ------------------------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  });
  return fn;
}

int make_map_2d(CUtensorMap* map, const void* base, uint64_t cols, uint64_t rows, uint64_t row_stride_elems,
                uint32_t box_cols, uint32_t box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  CVB_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  CVB_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base must be 16-byte aligned");
  CVB_REQUIRE((row_stride_elems * 2) % 16 == 0, "TMA row stride must be a multiple of 16 bytes");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {row_stride_elems * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r) + " (cols " +
                   std::to_string(cols) + ", rows " + std::to_string(rows) + ", stride " +
                   std::to_string(row_stride_elems) + ", box " + std::to_string(box_cols) + "x" +
                   std::to_string(box_rows) + ")");
    return CVB_ERR_CUDA;
  }
  return CVB_OK;
}

int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = NUM_SMS;
  }
  return n;
}

}  // namespace

// shared with the other tcgen05 kernels of the library (ffn_tc.cu)
int make_map_2d_bf16(CUtensorMap* map, const void* base, uint64_t cols, uint64_t rows, uint64_t row_stride_elems,
                     uint32_t box_cols, uint32_t box_rows) {
  return make_map_2d(map, base, cols, rows, row_stride_elems, box_cols, box_rows);
}
int device_sm_count() { return num_sms(); }

int launch_tapgemm_tc(const GemmParams& p_in, cudaStream_t stream) {
  GemmParams p = p_in;
  CVB_REQUIRE(p.dt == DT_BF16, "tcgen05 engine needs bf16 operands");
  CVB_REQUIRE(p.seg_len % GEMM_BK == 0, "seg_len must be a multiple of 64");
  CVB_REQUIRE(p.n_phase * p.n_seg <= MAX_SEGS, "too many segments");
  CVB_REQUIRE(p.m_tiles > 0, "empty problem");
  CVB_REQUIRE(p.bn >= 16 && p.bn <= 256 && p.bn % 16 == 0 && p.n_pad % p.bn == 0, "bad N tile");
  CVB_REQUIRE(!p.epi.gn_stats || p.bn % 32 == 0, "GroupNorm statistics need bn % 32 == 0");
  const bool rp = p.epi.row_parts != nullptr || p.epi.row_parts_out != nullptr;
  CVB_REQUIRE(!p.epi.row_parts_out || (p.N == 256 && p.n_pad == 256 && p.n_phase == 1 && p.bn % 32 == 0 && !p.epi.gn_stats &&
                                       (reinterpret_cast<uintptr_t>(p.epi.row_parts_out) & 7) == 0),
              "row_parts_out: N == 256, one phase, no GroupNorm statistics, 8-byte aligned");
  CVB_REQUIRE(!p.epi.row_parts || (!p.epi.row_stats && p.epi.row_u && (reinterpret_cast<uintptr_t>(p.epi.row_parts) & 15) == 0),
              "row_parts: needs row_u, excludes row_stats, 16-byte aligned");
  CVB_TRY(check_epilogue_layout(p));
  const int sms = num_sms();
  // Latency mode (streaming chunks: a handful of 128-row tiles): narrower N tiles put the problem on more SMs.
  // Each CTA then streams a quarter of the weights; the A tile is re-read from L2 by its N-tile siblings.
  while (p.bn % 128 == 0 && 2LL * p.m_tiles * p.n_phase * (p.n_pad / p.bn) <= sms) p.bn /= 2;
  // A 256-wide tile whose weights do not fit the B region but whose two 128-wide halves do (K = 512: the attention
  // out-projection) runs as two resident-W keys: W is fetched once per CTA instead of once per tile and the tail
  // is 9 of 8.65 rounds instead of 5 of 4.3 (profiles/r01_gemm_bench.txt: 66.4 -> 62.6 us; K = 768 / 1024 / 1536
  // stay non-resident at 128 and measured slower there).
  {
    const uint64_t Kfull = (uint64_t)p.n_seg * p.seg_len;
    const uint64_t cap = (uint64_t)STAGES * B_STAGE_BYTES;
    if (p.bn == 256 && (uint64_t)p.bn * Kfull * 2 > cap && 128ull * Kfull * 2 <= cap &&
        p.m_tiles * p.n_phase * (p.n_pad / p.bn) > sms)
      p.bn = 128;
  }
  static std::once_flag attr_once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(attr_once, [] {
    attr_err = cudaFuncSetAttribute(tapgemm_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (attr_err == cudaSuccess)
      attr_err = cudaFuncSetAttribute(tapgemm_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
  });
  CVB_CUDA(attr_err);
  CUtensorMap tmA, tmW;
  CVB_TRY(make_map_2d(&tmA, p.A, (uint64_t)p.a_cols, (uint64_t)p.a_rows, (uint64_t)p.lda, GEMM_BK, TILE_M));
  const uint64_t K = (uint64_t)p.n_seg * p.seg_len;
  CVB_TRY(make_map_2d(&tmW, p.W, K, (uint64_t)p.n_phase * p.n_pad, K, GEMM_BK, (uint32_t)p.bn));
  const int keys = p.n_phase * (p.n_pad / p.bn);
  const int total = p.m_tiles * keys;
  const bool resident = (uint64_t)p.bn * K * 2 <= (uint64_t)STAGES * B_STAGE_BYTES && keys <= sms &&
                        (p.bn * GEMM_BK * 2) % 1024 == 0;
  int grid;
  if (resident) {
    int per_key = sms / keys;
    if (per_key > p.m_tiles) per_key = p.m_tiles;
    grid = per_key * keys;
  } else {
    grid = total < sms ? total : sms;
  }
  if (rp)
    CVB_CUDA(launch_pdl(tapgemm_tc_kernel<true>, dim3(grid), dim3(NUM_THREADS), SMEM_BYTES, stream, tmA, tmW, p, resident ? 1 : 0));
  else
    CVB_CUDA(launch_pdl(tapgemm_tc_kernel<false>, dim3(grid), dim3(NUM_THREADS), SMEM_BYTES, stream, tmA, tmW, p, resident ? 1 : 0));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

}  // namespace cvb
