// tcgen05 / TMEM flash attention for the estimator's transformer blocks (8 heads x 64, bf16 operands).
//   out[q, h*64+d] = sum_j softmax_j(Q[q,h,:].K[j,h,:] * scale) V[j,h,d],  j over the valid rows of q's sequence
// Reference: cosyvoice/flow/components/transformer.py:266-271 -> diffusers Attention / AttnProcessor2_0
// (F.scaled_dot_product_attention with the all-ones additive mask of flow/decoder.py:258 = no mask at B=1).
//
// Work item = (128-query tile, head), items strided over a persistent grid; 2 CTAs co-resident per SM (256 TMEM
// columns, 80 KiB smem each) so that one CTA's softmax overlaps the other's MMAs.
//   warp 0      : TMA producer -- Q tile once, then (K, V) 128-key tiles into a 2-stage ring; Q, K and V are
//                 128x64 boxes of the SAME time-major qkv buffer (one tensor map, 128B swizzle).
//   warp 1      : TMEM allocator + single-thread tcgen05.mma issuer
//                   S[128x128] = Q K^T        (SS, both operands K-major)
//                   O[128x64] += P V          (A = P from TMEM as packed bf16, B = V tile read MN-major in place:
//                                              no transposed copy of V is ever made)
//   warps 2..9  : softmax, two threads per query row (64 scores each, kept in registers): tcgen05.ld S, max
//                 (halves exchanged through smem), exp2, pack to bf16, tcgen05.st P; O is rescaled in TMEM only
//                 when some row max of the warp moved.
// TMEM columns: S [0,128) fp32, P [128,192) packed bf16, O [192,256) fp32.
#include <cuda.h>
#include <mutex>
#include "common.cuh"
#include "ptx.cuh"
#include "../../include/cvb200.h"

namespace cvb {

using Layout = cvb200_layout;
int device_sm_count();        // tapgemm_tc.cu

namespace {

constexpr int HD = 64, BQ = 128, BK = 128;
constexpr int TILE_BYTES = 128 * 64 * 2;            // 16 KiB
constexpr int KV_STAGES = 2;
constexpr int SMEM_BYTES = 1024 + TILE_BYTES * (1 + 2 * KV_STAGES) + 256 + 2 * 2 * 128 * 4;   // barriers + smax
constexpr int TM_S = 0, TM_P = 128, TM_O = 192, TM_COLS = 256;

// MN-major 128B-swizzled shared-memory operand (V tile: rows = keys (K dim), 64 d (N dim) contiguous = 128 B).
// 8 K-rows form one 1024-byte swizzle atom: SBO = 1024; a single 64-wide atom along N: LBO unused.
__device__ __forceinline__ uint64_t make_sw128_mnmajor_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)(1024 >> 4) << 16;       // LBO
  d |= (uint64_t)(1024 >> 4) << 32;       // SBO
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// exp2 on the FMA/ALU pipes (Cody-Waite split + degree-3 minimax polynomial of 2^f on [-0.5, 0.5], relative error
// 7.5e-5, far below the bf16 rounding of P) for every ATTN_POLY_EVERY-th pair of probabilities.  x <= RESCALE_LOG2;
// -inf (masked key) gives 2^-126.  MEASURED SLOWER and therefore OFF by default (profiles/r01_attn_ab.txt: T = 609
// 401 TFLOP/s without, 394 / 392 / 372 with every 4th / 3rd / 2nd pair): the MUFU pipe is the busiest unit (ncu 54 %)
// but not the limiter -- the per-block chain S -> max -> exp -> P -> PV of a CTA is latency bound, and the extra
// FMA/ALU instructions lengthen it.
#ifndef ATTN_POLY_EVERY
#define ATTN_POLY_EVERY 0
#endif
__device__ __forceinline__ float ex2_poly(float x) {
  x = fmaxf(x, -126.f);
  const float t = x + 12582912.f;                 // 1.5 * 2^23: the integer part lands in the low mantissa bits
  const float f = x - (t - 12582912.f);
  const float p = fmaf(fmaf(fmaf(0.05517164617776871f, f, 0.2426111251115799f), f, 0.6932609677314758f), f, 0.9999280571937561f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}
// O is rescaled only when a row's running max has grown by more than 2^RESCALE_LOG2 since the offset in use was
// chosen (probabilities then stay below 2^8, exact in fp32 sums and harmless in bf16): with random scores the max
// moves in almost every early block, so the unconditional variant paid a TMEM round trip of O per block.
constexpr float RESCALE_LOG2 = 8.f;

__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}

// Work item = (128-query tile, head), items strided over a persistent grid of 2 CTAs per SM: barriers, TMEM and the
// tensor map are set up once per CTA, and the K/V ring keeps streaming across item boundaries (at T = 609 an item is
// only 5 key blocks, so a per-item prologue was a quarter of its time).  All block-level barrier parities run on a
// global block counter g; q_full / q_free / o_free run on the item counter n.
struct AttnItem {
  int head, b, q0, len, seq0, nblk;
  bool valid;
};
__device__ __forceinline__ AttnItem attn_item(const Layout& lay, int it, int heads) {
  AttnItem w;
  w.head = it % heads;
  const int tile = it / heads;
  w.b = __ldg(lay.tile_seq + tile);
  w.q0 = __ldg(lay.tile_row0 + tile);
  w.len = __ldg(lay.seq_len + w.b);
  w.seq0 = __ldg(lay.seq_row0 + w.b);
  w.nblk = (w.len + BK - 1) / BK;
  w.valid = w.q0 < w.len;                                  // padding-only tiles are skipped by every role
  return w;
}

__global__ void __launch_bounds__(320, 2)
attention_tc_kernel(const __grid_constant__ CUtensorMap tm, __nv_bfloat16* __restrict__ out, int ldo, const Layout lay,
                    int heads, float scale_log2, int n_items) {
  const int inner = heads * HD;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = smem + TILE_BYTES;                         // [KV_STAGES]
  uint8_t* sV = sK + KV_STAGES * TILE_BYTES;               // [KV_STAGES]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + KV_STAGES * TILE_BYTES);
  uint64_t* q_full = bars;                                 // 1
  uint64_t* k_full = bars + 1;                             // [2]  K and V run on separate rings: a K stage is free as soon
  uint64_t* k_empty = bars + 3;                            // [2]  as its QK^T has retired, a V stage only after its PV
  uint64_t* v_full = bars + 5;                             // [2]
  uint64_t* v_empty = bars + 7;                            // [2]
  uint64_t* s_full = bars + 9;
  uint64_t* s_free = bars + 10;
  uint64_t* p_full = bars + 11;
  uint64_t* pv_done = bars + 12;
  uint64_t* q_free = bars + 13;                            // every QK^T of the item has retired: Q may be overwritten
  uint64_t* o_free = bars + 14;                            // the item's O has been read out (256 arrivals)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 15);
  float* smax = reinterpret_cast<float*>(bars + 32);      // [2 parities][2 halves][128 rows]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    ptx::mbar_init(q_full, 1);
    for (int i = 0; i < KV_STAGES; ++i) {
      ptx::mbar_init(&k_full[i], 1); ptx::mbar_init(&k_empty[i], 1);
      ptx::mbar_init(&v_full[i], 1); ptx::mbar_init(&v_empty[i], 1);
    }
    ptx::mbar_init(s_full, 1);
    ptx::mbar_init(s_free, 256);
    ptx::mbar_init(p_full, 256);
    ptx::mbar_init(pv_done, 1);
    ptx::mbar_init(q_free, 1);
    ptx::mbar_init(o_free, 256);
    ptx::fence_barrier_init();
    ptx::prefetch_tensormap(&tm);
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, TM_COLS);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  pdl_trigger();      // programmatic dependent launch: the prologue above overlapped the previous kernel's tail
  pdl_wait();

  if (warp == 0) {
    if (lane == 0) {
      // Two cursors over the same flattened (item, key block) sequence: the K stream (which also brings each item's Q
      // tile) runs one block ahead of the V stream.  K(g+1)'s stage frees when QK(g-1) retires, which precedes the
      // retirement of PV(g-2) that frees V(g)'s stage (the MMA thread issues QK(g+1) before PV(g)), so the single
      // producer thread never waits on the later event first.
      struct Cur { int it, j; uint32_t g, n; AttnItem w; bool ok; };
      auto seek = [&](Cur& c) {
        for (; c.it < n_items; c.it += gridDim.x) {
          c.w = attn_item(lay, c.it, heads);
          if (c.w.valid) { c.ok = true; return; }
        }
        c.ok = false;
      };
      auto adv = [&](Cur& c) {
        ++c.g;
        if (++c.j == c.w.nblk) { c.j = 0; ++c.n; c.it += gridDim.x; seek(c); }
      };
      auto load_k = [&](const Cur& c) {
        if (c.j == 0) {
          if (c.n > 0) ptx::mbar_wait(q_free, (c.n - 1) & 1);
          ptx::mbar_arrive_expect_tx(q_full, TILE_BYTES);
          ptx::tma_load_2d(sQ, &tm, q_full, c.w.head * HD, c.w.seq0 + c.w.q0);
        }
        const int st = c.g % KV_STAGES;
        ptx::mbar_wait(&k_empty[st], ((c.g / KV_STAGES) & 1) ^ 1);
        ptx::mbar_arrive_expect_tx(&k_full[st], TILE_BYTES);
        ptx::tma_load_2d(sK + st * TILE_BYTES, &tm, &k_full[st], inner + c.w.head * HD, c.w.seq0 + c.j * BK);
      };
      auto load_v = [&](const Cur& c) {
        const int st = c.g % KV_STAGES;
        ptx::mbar_wait(&v_empty[st], ((c.g / KV_STAGES) & 1) ^ 1);
        ptx::mbar_arrive_expect_tx(&v_full[st], TILE_BYTES);
        ptx::tma_load_2d(sV + st * TILE_BYTES, &tm, &v_full[st], 2 * inner + c.w.head * HD, c.w.seq0 + c.j * BK);
      };
      Cur kc;
      kc.it = blockIdx.x; kc.j = 0; kc.g = 0; kc.n = 0; kc.ok = false;
      seek(kc);
      Cur vc = kc;
      if (kc.ok) { load_k(kc); adv(kc); }
      while (vc.ok) {
        if (kc.ok) { load_k(kc); adv(kc); }
        load_v(vc);
        adv(vc);
      }
    }
  } else if (warp == 1) {
    // whole warp, uniform state, one elected lane issues (ptx.cuh "lean issue path")
    {
      const uint32_t idesc_qk = ptx::make_idesc_bf16(BQ, BK);                       // 128 x 128, A/B K-major
      const uint32_t idesc_pv = ptx::make_idesc_bf16(BQ, HD) | (1u << 16);          // 128 x 64, B MN-major
      const uint32_t kdesc_hi = (uint32_t)(ptx::make_sw128_kmajor_desc(0) >> 32);
      const uint32_t vdesc_hi = (uint32_t)(make_sw128_mnmajor_desc(0) >> 32);
      const uint32_t q_lo = (uint32_t)ptx::make_sw128_kmajor_desc(ptx::smem_u32(sQ));
      const uint32_t k_lo0 = (uint32_t)ptx::make_sw128_kmajor_desc(ptx::smem_u32(sK));
      const uint32_t v_lo0 = (uint32_t)make_sw128_mnmajor_desc(ptx::smem_u32(sV));
      auto issue_qk = [&](uint32_t gg) {
        const int st = gg % KV_STAGES;
        ptx::mbar_wait(&k_full[st], (gg / KV_STAGES) & 1);
        if (gg > 0) ptx::mbar_wait(s_free, (gg - 1) & 1);     // softmax has pulled S(g-1) out of TMEM
        ptx::tc_fence_after();
        const uint32_t k_lo = ptx::bcast0(k_lo0 + st * (TILE_BYTES >> 4));
        if (ptx::elect_one()) {
#pragma unroll
          for (int k = 0; k < HD / 16; ++k)
            ptx::umma_ss_lo(tmem + TM_S, q_lo + k * 2, kdesc_hi, k_lo + k * 2, kdesc_hi, idesc_qk, k != 0);
          ptx::umma_commit(&k_empty[st]);                       // the K stage is dead once this QK^T has retired
          ptx::umma_commit(s_full);
        }
        __syncwarp();
      };
      uint32_t n = 0, g = 0;
      for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
        const AttnItem w = attn_item(lay, it, heads);
        if (!w.valid) continue;
        ptx::mbar_wait(q_full, n & 1);
        issue_qk(g);
        for (int j = 0; j < w.nblk; ++j, ++g) {
          if (j + 1 < w.nblk) issue_qk(g + 1);
          else {
            if (ptx::elect_one()) ptx::umma_commit(q_free);      // the item's last QK^T is in flight
            __syncwarp();
          }
          const int st = g % KV_STAGES;
          ptx::mbar_wait(p_full, g & 1);                       // P(g) is in TMEM, O has been rescaled
          if (j == 0 && n > 0) ptx::mbar_wait(o_free, (n - 1) & 1);   // previous item's O has been read out
          ptx::mbar_wait(&v_full[st], (g / KV_STAGES) & 1);
          ptx::tc_fence_after();
          const uint32_t v_lo = ptx::bcast0(v_lo0 + st * (TILE_BYTES >> 4));
          const uint32_t acc0 = ptx::bcast0(j != 0);
          if (ptx::elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k)
              ptx::umma_ts_lo(tmem + TM_O, tmem + TM_P + k * 8, v_lo + k * (2048 >> 4), vdesc_hi, idesc_pv, k == 0 ? acc0 : 1u);
            ptx::umma_commit(&v_empty[st]);
            ptx::umma_commit(pv_done);
          }
          __syncwarp();
        }
        ++n;
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax + output (warps 2..9)
    // Two warps per TMEM lane group: thread (row, half) owns 64 of the row's 128 scores, which stay in
    // registers between the max and the exp pass (one batch of tcgen05.ld per block); the two halves of a
    // row exchange their maxima through shared memory.
    const int lane_grp = warp & 3;
    const int half = (warp - 2) >> 2;
    const int row = lane_grp * 32 + lane;
    const uint32_t tbase = tmem + ((uint32_t)(lane_grp * 32) << 16);
    const uint32_t smax_addr = ptx::smem_u32(smax);
    uint32_t g = 0;
    for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
      const AttnItem w = attn_item(lay, it, heads);
      if (!w.valid) continue;
      const int len = w.len;
      float m_run = -INFINITY, m_use = -INFINITY, l_half = 0.f;   // running max; the max the exponent offset uses
      for (int j = 0; j < w.nblk; ++j, ++g) {
        ptx::mbar_wait(s_full, g & 1);
        ptx::tc_fence_after();
        const int nvalid = len - j * BK - half * 64;            // valid keys among this thread's 64 columns
        uint32_t raw[2][32];
        ptx::tmem_ld32_nowait(tbase + TM_S + half * 64, raw[0]);
        ptx::tmem_ld32_nowait(tbase + TM_S + half * 64 + 32, raw[1]);
        ptx::tmem_wait_ld();
        ptx::tc_fence_before();
        ptx::mbar_arrive(s_free);                               // S(g) is in registers: QK(g+1) may overwrite it
        float v[64];
#pragma unroll
        for (int i = 0; i < 64; ++i) v[i] = __uint_as_float(raw[i >> 5][i & 31]);
        if (nvalid < 64) {
#pragma unroll
          for (int i = 0; i < 64; ++i) if (i >= nvalid) v[i] = -INFINITY;
        }
        float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int i = 0; i < 64; i += 4) {
          mx4[0] = fmaxf(mx4[0], v[i]);
          mx4[1] = fmaxf(mx4[1], v[i + 1]);
          mx4[2] = fmaxf(mx4[2], v[i + 2]);
          mx4[3] = fmaxf(mx4[3], v[i + 3]);
        }
        const float m_loc = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
        // the two halves of a row live in the two warps of a TMEM lane group: explicit shared-space accesses (the generic
        // LD the compiler emitted for this exchange was the kernel's top stall) and a barrier of just those 64 threads
        const uint32_t xch = smax_addr + (g & 1) * 1024;
        asm volatile("st.shared.f32 [%0], %1;" ::"r"(xch + (half * 128 + row) * 4), "f"(m_loc) : "memory");
        ptx::named_bar_sync(1 + lane_grp, 64);
        float m_other;
        asm volatile("ld.shared.f32 %0, [%1];" : "=f"(m_other) : "r"(xch + ((half ^ 1) * 128 + row) * 4) : "memory");
        const float m_new = fmaxf(m_run, fmaxf(m_loc, m_other));
        const bool need = j == 0 || (m_new - m_use) * scale_log2 > RESCALE_LOG2;
        float corr = 1.f;
        if (need) {
          corr = j == 0 ? 0.f : ex2_approx((m_use - m_new) * scale_log2);
          m_use = m_new;
        }
        const bool moved = __any_sync(0xffffffffu, need);
        m_run = m_new;
        const float moff = m_use * scale_log2;
        if (g > 0) ptx::mbar_wait(pv_done, (g - 1) & 1);               // PV(g-1) retired: P and O may be touched
        ptx::tc_fence_after();
        float ls4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int c = 0; c < 64; c += 32) {
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            float p0, p1;
            if (ATTN_POLY_EVERY > 0 && ((i >> 1) % (ATTN_POLY_EVERY > 0 ? ATTN_POLY_EVERY : 1)) == ATTN_POLY_EVERY - 1) {
              p0 = ex2_poly(fmaf(v[c + i], scale_log2, -moff));                   // masked columns: 2^-126
              p1 = ex2_poly(fmaf(v[c + i + 1], scale_log2, -moff));
            } else {
              p0 = ex2_approx(fmaf(v[c + i], scale_log2, -moff));                 // masked columns: exp2(-inf) = 0
              p1 = ex2_approx(fmaf(v[c + i + 1], scale_log2, -moff));
            }
            ls4[(i >> 1) & 3] += p0 + p1;
            pk[i >> 1] = pack2(p0, p1);
          }
          ptx::tmem_st16(tbase + TM_P + half * 32 + (c >> 1), pk);
        }
        l_half = l_half * corr + ((ls4[0] + ls4[1]) + (ls4[2] + ls4[3]));
        if (j > 0 && moved) {                                          // rescale this thread's 32 output columns
          float o[32];
          ptx::tmem_ld32(tbase + TM_O + half * 32, o);
          uint32_t lo[16], hi[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) { lo[i] = __float_as_uint(o[i] * corr); hi[i] = __float_as_uint(o[16 + i] * corr); }
          ptx::tmem_st16(tbase + TM_O + half * 32, lo);
          ptx::tmem_st16(tbase + TM_O + half * 32 + 16, hi);
        }
        ptx::tmem_wait_st();
        ptx::tc_fence_before();
        ptx::mbar_arrive(p_full);
      }
      // ---- epilogue: O / l -> bf16 (each thread writes 32 of the row's 64 output columns)
      const uint32_t xch = smax_addr + (g & 1) * 1024;
      asm volatile("st.shared.f32 [%0], %1;" ::"r"(xch + (half * 128 + row) * 4), "f"(l_half) : "memory");
      ptx::named_bar_sync(1 + lane_grp, 64);
      float l_other;
      asm volatile("ld.shared.f32 %0, [%1];" : "=f"(l_other) : "r"(xch + ((half ^ 1) * 128 + row) * 4) : "memory");
      const float inv = 1.f / (l_half + l_other);
      ptx::mbar_wait(pv_done, (g - 1) & 1);
      ptx::tc_fence_after();
      const int lrow = w.q0 + row;
      float o[32];
      ptx::tmem_ld32(tbase + TM_O + half * 32, o);
      ptx::tc_fence_before();
      ptx::mbar_arrive(o_free);                                        // the next item's first PV may overwrite O
      if (lrow < len) {
        __nv_bfloat16* dst = out + (size_t)(w.seq0 + lrow) * ldo + w.head * HD + half * 32;
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          uint4 pk;
          pk.x = pack2(o[i] * inv, o[i + 1] * inv);
          pk.y = pack2(o[i + 2] * inv, o[i + 3] * inv);
          pk.z = pack2(o[i + 4] * inv, o[i + 5] * inv);
          pk.w = pack2(o[i + 6] * inv, o[i + 7] * inv);
          *reinterpret_cast<uint4*>(dst + i) = pk;
        }
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem, TM_COLS);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
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

}  // namespace

int launch_attention_tc(const void* qkv, int ld, void* out, int ldo, const Layout& lay, int heads, float scale,
                        cudaStream_t st) {
  CVB_REQUIRE(heads > 0 && ld >= 3 * heads * HD && ldo >= heads * HD, "bad attention shapes");
  CVB_REQUIRE(ld % 8 == 0 && ldo % 8 == 0, "leading dimensions must keep 16-byte alignment");
  CVB_REQUIRE((reinterpret_cast<uintptr_t>(qkv) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0, "alignment");
  if (lay.rows <= 0) return CVB_OK;
  EncodeTiledFn fn = encode_fn();
  CVB_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  CUtensorMap tm;
  cuuint64_t dims[2] = {(cuuint64_t)(3 * heads * HD), (cuuint64_t)lay.rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {64, 128};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(qkv), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled (attention) failed with CUresult " + std::to_string((int)r));
    return CVB_ERR_CUDA;
  }
  static std::once_flag attr_once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(attr_once, [] {
    attr_err = cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
  });
  CVB_CUDA(attr_err);
  const int n_items = (lay.rows / 128) * heads;
  const int slots = 2 * device_sm_count();                 // 2 co-resident CTAs per SM
  const int grid = n_items < slots ? n_items : slots;
  CVB_CUDA(launch_pdl(attention_tc_kernel, dim3(grid), dim3(320), SMEM_BYTES, st, tm, reinterpret_cast<__nv_bfloat16*>(out), ldo,
                      lay, heads, scale * 1.4426950408889634f, n_items));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

}  // namespace cvb
