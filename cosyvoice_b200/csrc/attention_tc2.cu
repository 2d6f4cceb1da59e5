// tcgen05 / TMEM flash attention, second generation: ONE CTA per SM running TWO independent (128-query tile, head) work
// streams ("slots"), one softmax thread per query row.
//   out[q, h*64+d] = sum_j softmax_j(Q[q,h,:].K[j,h,:] * scale) V[j,h,d],  j over the valid rows of q's sequence
// Reference: cosyvoice/flow/components/transformer.py:266-271 -> diffusers Attention / AttnProcessor2_0
// (F.scaled_dot_product_attention with the all-ones additive mask of flow/decoder.py:258 = no mask at B=1).
//
// The first-generation kernel (attention_tc.cu) ran two CTAs per SM with two threads per query row; its top stall was the
// exchange of the two half-row maxima through shared memory (profiles/r01_ncu_attn_v2.md).  Here a thread owns a whole row
// of S (128 scores in registers between the max and the exp pass), so there is no exchange, no named barrier and no vote on
// the critical path of a block.  Each stream has its own loader and MMA-issuing warp with blocking mbarrier waits (a shared
// control thread polling both streams was tried: the polling loop starved the softmax warps of its scheduler).
//
//   warps 8, 10 : TMA loader of slot 0 / 1 (one thread, two cursors: the K stream -- which also brings each item's Q tile -- and
//                 the V stream); Q, K and V are 128x64 boxes of the SAME time-major qkv buffer (one tensor map, 128B swizzle)
//   warps 9, 11 : tcgen05.mma issuer of slot 0 / 1 (warp 9 also allocates TMEM)
//                   S[128x128] = Q K^T        (SS, both operands K-major)
//                   O[128x64] += P V          (A = P from TMEM as packed bf16, B = V tile read MN-major in place)
//                 QK^T(g+1) of a slot is issued as soon as its softmax threads have pulled S(g) into registers
//   warps 0-3   : softmax of slot 0 (thread = query row = TMEM lane): tcgen05.ld S, running max, exp2, bf16 P -> TMEM,
//                 lazy O rescale (only when the max grew by more than 2^8), final O / l -> global
//   warps 4-7   : softmax of slot 1
// TMEM columns per slot: S [0,128) fp32, P [128,192) packed bf16, O [192,256) fp32; slot 1 at +256.
#include <cuda.h>
#include <mutex>
#include "common.cuh"
#include "ptx.cuh"
#include "../../include/cvb200.h"

namespace cvb {

using Layout = cvb200_layout;
int device_sm_count();        // tapgemm_tc.cu

namespace {

#ifndef ATTN2_SLOTS
#define ATTN2_SLOTS 2
#endif
#ifndef ATTN2_BK
#define ATTN2_BK 128
#endif
constexpr int HD = 64, BQ = 128, BK = ATTN2_BK;     // BK: keys per block (64 or 128)
constexpr int SLOTS = ATTN2_SLOTS;
static_assert(BK == 64 || BK == 128, "key block = 64 or 128");
constexpr int CH = BK / 32;                          // 32-column chunks of a row of S
constexpr int Q_BYTES = BQ * HD * 2;                 // 16 KiB
constexpr int KV_BYTES = BK * HD * 2;                // 8 / 16 KiB
constexpr int KV_STAGES = 2;
constexpr int Q_BUFS = 2;                                            // the next item's Q tile is fetched during the current item
constexpr int SLOT_SMEM = Q_BYTES * Q_BUFS + 2 * KV_BYTES * KV_STAGES;     // Q x 2 + K ring + V ring
constexpr int SLOT_BARS = 20;                                        // mbarriers per slot
#ifndef ATTN2_MAX_ITEMS
#define ATTN2_MAX_ITEMS 256
#endif
constexpr int MAX_ITEMS = ATTN2_MAX_ITEMS;                           // per slot: the CTA's item table lives in shared memory
constexpr int SMEM_BYTES = 1024 + SLOTS * SLOT_SMEM + SLOTS * SLOT_BARS * 8 + 64 + SLOTS * MAX_ITEMS * 16;
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory");
// TMEM columns of a slot: S [0, BK) fp32, P [BK, BK + BK/2) packed bf16, O [.., +64) fp32
constexpr int TM_S = 0, TM_P = BK, TM_O = BK + BK / 2, TM_SLOT = BK + BK / 2 + HD, TM_COLS = 512;
static_assert(SLOTS * TM_SLOT <= TM_COLS, "TMEM columns");
// warps 0 .. 4*SLOTS-1: softmax (one warpgroup per slot); then the control warps, rounded up to whole warpgroups: per slot a
// loader warp and an MMA warp; the rest idle
constexpr int CTRL_WARP0 = SLOTS * 4, MMA_WARP = CTRL_WARP0 + 1;
constexpr int CTRL_WARPS = (2 * SLOTS + 3) / 4 * 4;
constexpr int NUM_THREADS = 32 * (CTRL_WARP0 + CTRL_WARPS);
// Register budget (setmaxnreg, per warpgroup): the kernel launches with the average budget, the control warpgroups drop to
// CTRL_REGS and every softmax warpgroup takes SOFTMAX_REGS (a softmax thread keeps a whole row of S in registers).
// setmaxnreg only moves registers inside the CTA's own pool (what the control warps give back), so
//   softmax threads x (SOFTMAX_REGS - LAUNCH_REGS) <= control threads x (LAUNCH_REGS - CTRL_REGS);
// asking for more faults the launch (seen with 3 slots: "unspecified launch failure").
constexpr int LAUNCH_REGS = 65536 / NUM_THREADS / 8 * 8;            // what __launch_bounds__(NUM_THREADS, 1) compiles to
constexpr int CTRL_REGS = 56;
constexpr int SOFTMAX_REGS_ = LAUNCH_REGS + (32 * CTRL_WARPS * (LAUNCH_REGS - CTRL_REGS)) / (128 * SLOTS) / 8 * 8;
constexpr int SOFTMAX_REGS = SOFTMAX_REGS_ > 232 ? 232 : SOFTMAX_REGS_;

// MN-major 128B-swizzled shared-memory operand (V tile: rows = keys (K dim), 64 d (N dim) contiguous = 128 B).
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
// exp2 on the FMA/ALU pipes (Cody-Waite split + degree-3 minimax polynomial of 2^f on [-0.5, 0.5], relative error 7.5e-5, far
// below the bf16 rounding of P) for every ATTN2_POLY_EVERY-th pair of probabilities: the MUFU pipe is the bound of this kernel
// (16 exp2 per cycle per SM against 128 x 128 scores per 512 cycles of MMA).
#ifndef ATTN2_POLY_EVERY
#define ATTN2_POLY_EVERY 0
#endif
__device__ __forceinline__ float ex2_poly(float x) {
  x = fmaxf(x, -126.f);
  const float t = x + 12582912.f;                 // 1.5 * 2^23: the integer part lands in the low mantissa bits
  const float f = x - (t - 12582912.f);
  const float p = fmaf(fmaf(fmaf(0.05517164617776871f, f, 0.2426111251115799f), f, 0.6932609677314758f), f, 0.9999280571937561f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}
constexpr float RESCALE_LOG2 = 8.f;               // O is rescaled only when a row max grew by more than 2^8 (see attention_tc.cu)

__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}

// A work item = (128-query tile, head).  Every role walks the same item list; reading it from global memory costs two
// dependent L2 round trips per item in EVERY role (tile -> sequence -> length), which the in-kernel profile showed as ~1400
// stalled cycles per item in the MMA issuer alone.  The CTA therefore resolves its items ONCE, in parallel, into shared memory.
struct __align__(16) AttnItem {
  int head, q0, len, seq0;      // len = 0 marks a padding-only tile (skipped by every role)
};
__device__ __forceinline__ AttnItem attn_item(const Layout& lay, int it, int heads) {
  AttnItem w;
  w.head = it % heads;
  const int tile = it / heads;
  const int b = __ldg(lay.tile_seq + tile);
  w.q0 = __ldg(lay.tile_row0 + tile);
  w.len = __ldg(lay.seq_len + b);
  w.seq0 = __ldg(lay.seq_row0 + b);
  if (w.q0 >= w.len) w.len = 0;
  return w;
}

// Per-slot barriers.  Block-level parities run on the slot's global block counter g, item-level ones on its item counter n.
struct SlotBars {
  uint64_t* q_full;    // [2] Q tile landed (buffer = item parity)
  uint64_t* q_free;    // [2] the item's last QK^T retired: its Q buffer may be overwritten
  uint64_t* k_full;    // [2]
  uint64_t* k_empty;   // [2] QK^T(g) retired
  uint64_t* v_full;    // [2]
  uint64_t* v_empty;   // [2] PV(g) retired
  uint64_t* s_full;    // QK^T(g) retired: S is in TMEM
  uint64_t* s_free;    // S(g) is in the softmax threads' registers (4 warp arrivals)
  uint64_t* p_full;    // P(g) is in TMEM, O rescaled if needed (4 warp arrivals)
  uint64_t* pv_done;   // PV(g) retired: P and O may be touched
  uint64_t* o_free;    // the item's O has been read out (4 warp arrivals)
};
__device__ __forceinline__ SlotBars slot_bars(uint64_t* base) {
  SlotBars b;
  b.q_full = base; b.q_free = base + 2; b.k_full = base + 4; b.k_empty = base + 6; b.v_full = base + 8; b.v_empty = base + 10;
  b.s_full = base + 12; b.s_free = base + 13; b.p_full = base + 14; b.pv_done = base + 15; b.o_free = base + 16;
  return b;
}

// cursor over the flattened (item, key block) sequence of one slot (items from the shared-memory table)
struct Cursor {
  int k, j, nblk;       // table index, key block within the item, blocks of the item
  uint32_t g, n;        // global block counter, item counter (barrier parities)
  AttnItem w;
  bool ok;
};

__global__ void __launch_bounds__(NUM_THREADS, 1)
attention_tc2_kernel(const __grid_constant__ CUtensorMap tm, const __grid_constant__ CUtensorMap tmkv,
                     __nv_bfloat16* __restrict__ out, int ldo, const Layout lay,
                     int heads, float scale_log2, int item0, int n_items) {
  const int inner = heads * HD;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + SLOTS * SLOT_SMEM);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + SLOTS * SLOT_BARS);
  AttnItem* table = reinterpret_cast<AttnItem*>(reinterpret_cast<uint8_t*>(tmem_slot) + 64);      // [SLOTS][MAX_ITEMS]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < SLOTS; ++s) {
      const SlotBars b = slot_bars(bars + s * SLOT_BARS);
      for (int i = 0; i < Q_BUFS; ++i) { ptx::mbar_init(&b.q_full[i], 1); ptx::mbar_init(&b.q_free[i], 1); }
      for (int i = 0; i < KV_STAGES; ++i) {
        ptx::mbar_init(&b.k_full[i], 1); ptx::mbar_init(&b.k_empty[i], 1);
        ptx::mbar_init(&b.v_full[i], 1); ptx::mbar_init(&b.v_empty[i], 1);
      }
      ptx::mbar_init(b.s_full, 1);
      ptx::mbar_init(b.s_free, 4);
      ptx::mbar_init(b.p_full, 4);
      ptx::mbar_init(b.pv_done, 1);
      ptx::mbar_init(b.o_free, 4);
    }
    ptx::fence_barrier_init();
    ptx::prefetch_tensormap(&tm);
    ptx::prefetch_tensormap(&tmkv);
  }
  if (warp == MMA_WARP) ptx::tmem_alloc(tmem_slot, TM_COLS);
  // Items are dealt round-robin to the 2 * gridDim.x slots of the grid: slot s of CTA c takes items 2c + s, + 2 gridDim.x, ...
  // The table is built BEFORE the programmatic-dependency wait: the layout arrays are not written by the preceding kernel
  // (lengths are re-bound by a copy that the stream orders before it), so the two dependent L2 round trips per item
  // overlap the previous kernel's tail -- what a streaming chunk's 560 small attention launches are mostly made of.
  const int it_step = SLOTS * (int)gridDim.x;
  const int per_slot = (n_items - SLOTS * (int)blockIdx.x + it_step - 1) / it_step;     // items of slot 0 (slot 1: <= this)
  for (int idx = threadIdx.x; idx < SLOTS * per_slot; idx += NUM_THREADS) {
    const int s = idx % SLOTS, k = idx / SLOTS;
    const int it = SLOTS * (int)blockIdx.x + s + k * it_step;
    AttnItem w;
    w.head = 0; w.q0 = 0; w.len = 0; w.seq0 = 0;
    if (it < n_items) w = attn_item(lay, item0 + it, heads);
    table[s * MAX_ITEMS + k] = w;
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  pdl_trigger();      // programmatic dependent launch: the prologue above overlapped the previous kernel's tail
  pdl_wait();

  auto seek = [&](Cursor& c, int s) {
    for (; c.k < per_slot; ++c.k) {
      c.w = table[s * MAX_ITEMS + c.k];
      if (c.w.len > 0) { c.nblk = (c.w.len + BK - 1) / BK; c.ok = true; return; }
    }
    c.ok = false;
  };
  auto adv = [&](Cursor& c, int s) {
    ++c.g;
    if (++c.j == c.nblk) { c.j = 0; ++c.n; ++c.k; seek(c, s); }
  };
  auto first = [&](int s) {
    Cursor c;
    c.k = 0; c.j = 0; c.nblk = 0; c.g = 0; c.n = 0; c.ok = false;
    seek(c, s);
    return c;
  };

  // (setmaxnreg sits at the top of each role branch: after a merge point ptxas would apply the smaller budget to all.)
  if (warp >= CTRL_WARP0 + 2 * SLOTS) {
    ptx::reg_dealloc<CTRL_REGS>();                  // padding warps of the last control warpgroup: nothing to do
  } else if (warp >= CTRL_WARP0 && ((warp - CTRL_WARP0) & 1) == 0) {
    // ------------------------------------------------------------------ TMA loader of slot s (blocking waits: a control warp
    // that polls steals issue slots from the softmax warp of its scheduler -- measured: the exp pass of a warp took 1700-2400
    // cycles next to a polling loop against 1280 in isolation)
    ptx::reg_dealloc<CTRL_REGS>();
    const int s = (warp - CTRL_WARP0) >> 1;
    if (lane == 0) {
      const SlotBars b = slot_bars(bars + s * SLOT_BARS);
      uint8_t* sQ = smem + s * SLOT_SMEM;
      uint8_t* sK = sQ + Q_BUFS * Q_BYTES;
      uint8_t* sV = sK + KV_STAGES * KV_BYTES;
      // Two cursors over the same (item, key block) sequence: the K stream (which also brings each item's Q tile) runs one
      // block ahead of the V stream.  K(g+1)'s stage frees when QK(g-1) retires, which precedes the retirement of PV(g-2)
      // that frees V(g)'s stage (the MMA thread issues QK(g+1) before PV(g)), so this thread never waits on the later event first.
      auto load_k = [&](const Cursor& c) {
        if (c.j == 0) {
          const int qb = c.n % Q_BUFS;
          if (c.n >= Q_BUFS) ptx::mbar_wait(&b.q_free[qb], ((c.n / Q_BUFS) - 1) & 1);
          ptx::mbar_arrive_expect_tx(&b.q_full[qb], Q_BYTES);
          ptx::tma_load_2d(sQ + qb * Q_BYTES, &tm, &b.q_full[qb], c.w.head * HD, c.w.seq0 + c.w.q0);
        }
        const int st = c.g % KV_STAGES;
        ptx::mbar_wait(&b.k_empty[st], ((c.g / KV_STAGES) & 1) ^ 1);
        ptx::mbar_arrive_expect_tx(&b.k_full[st], KV_BYTES);
        ptx::tma_load_2d(sK + st * KV_BYTES, &tmkv, &b.k_full[st], inner + c.w.head * HD, c.w.seq0 + c.j * BK);
      };
      auto load_v = [&](const Cursor& c) {
        const int st = c.g % KV_STAGES;
        ptx::mbar_wait(&b.v_empty[st], ((c.g / KV_STAGES) & 1) ^ 1);
        ptx::mbar_arrive_expect_tx(&b.v_full[st], KV_BYTES);
        ptx::tma_load_2d(sV + st * KV_BYTES, &tmkv, &b.v_full[st], 2 * inner + c.w.head * HD, c.w.seq0 + c.j * BK);
      };
      Cursor kc = first(s);
      Cursor vc = kc;
      if (kc.ok) { load_k(kc); adv(kc, s); }
      while (vc.ok) {
        if (kc.ok) { load_k(kc); adv(kc, s); }
        load_v(vc);
        adv(vc, s);
      }
    }
  } else if (warp >= CTRL_WARP0) {
    // ------------------------------------------------------------------ MMA issuer of slot s
    // All 32 lanes run the loop (identical state in every lane); one elected lane issues.  With the loop inside
    // `if (lane == 0)` the compiler could not prove the operands uniform and emitted a register -> uniform-register
    // conversion loop (ELECT / R2UR.BROADCAST / BRA.U.ANY) in front of every tcgen05.mma: ~100 cycles per MMA.
    // Two cursors over the slot's flattened (item, key block) sequence: QK^T runs one block ahead of PV ACROSS item
    // boundaries, so the first S of the next item is already in TMEM when the softmax threads finish the current item
    // (the first version issued it after the item's last PV: 2400 idle cycles per item in the softmax warps).
    ptx::reg_dealloc<CTRL_REGS>();
    const int s = (warp - CTRL_WARP0) >> 1;
    const SlotBars b = slot_bars(bars + s * SLOT_BARS);
    const uint32_t idesc_qk = ptx::make_idesc_bf16(BQ, BK);                       // 128 x BK, A/B K-major
    const uint32_t idesc_pv = ptx::make_idesc_bf16(BQ, HD) | (1u << 16);          // 128 x 64, B MN-major
    // descriptor words: hi is the same for every K-major (resp. MN-major) operand, lo = (address & 0x3FFFF) >> 4
    const uint32_t kdesc_hi = (uint32_t)(ptx::make_sw128_kmajor_desc(0) >> 32);
    const uint32_t vdesc_hi = (uint32_t)(make_sw128_mnmajor_desc(0) >> 32);
    const uint32_t a0 = ptx::smem_u32(smem + s * SLOT_SMEM);
    const uint32_t q_lo0 = (uint32_t)ptx::make_sw128_kmajor_desc(a0);
    const uint32_t k_lo0 = (uint32_t)ptx::make_sw128_kmajor_desc(a0 + Q_BYTES * Q_BUFS);
    const uint32_t v_lo0 = (uint32_t)make_sw128_mnmajor_desc(a0 + Q_BYTES * Q_BUFS + KV_BYTES * KV_STAGES);
    const uint32_t tm_s = tmem + s * TM_SLOT;
    auto issue_qk = [&](const Cursor& c) {
      const uint32_t gg = c.g, nn = c.n;
      const int st = gg % KV_STAGES;
      const int qb = nn % Q_BUFS;
      if (c.j == 0) ptx::mbar_wait(&b.q_full[qb], (nn / Q_BUFS) & 1);
      const uint32_t q_lo = ptx::bcast0(q_lo0 + qb * (Q_BYTES >> 4));
      ptx::mbar_wait(&b.k_full[st], (gg / KV_STAGES) & 1);
      if (gg > 0) ptx::mbar_wait(b.s_free, (gg - 1) & 1);        // the softmax threads hold S(g-1) in registers
      ptx::tc_fence_after();
      const uint32_t k_lo = ptx::bcast0(k_lo0 + st * (KV_BYTES >> 4));
      const uint32_t lastu = ptx::bcast0(c.j + 1 == c.nblk);
      if (ptx::elect_one()) {
#pragma unroll
        for (int k = 0; k < HD / 16; ++k)
          ptx::umma_ss_lo(tm_s + TM_S, q_lo + k * 2, kdesc_hi, k_lo + k * 2, kdesc_hi, idesc_qk, k != 0);
        ptx::umma_commit(&b.k_empty[st]);                        // the K stage is dead once this QK^T has retired
        ptx::umma_commit(b.s_full);
        if (lastu) ptx::umma_commit(&b.q_free[qb]);              // the item's last QK^T is in flight
      }
      __syncwarp();
    };
    Cursor qc = first(s);
    Cursor pc = qc;
    if (qc.ok) { issue_qk(qc); adv(qc, s); }
    while (pc.ok) {
      if (qc.ok) { issue_qk(qc); adv(qc, s); }
      const uint32_t g = pc.g;
      const int st = g % KV_STAGES;
      ptx::mbar_wait(b.p_full, g & 1);                           // P(g) is in TMEM, O has been rescaled
      if (pc.j == 0 && pc.n > 0) ptx::mbar_wait(b.o_free, (pc.n - 1) & 1);   // previous item's O has been read out
      ptx::mbar_wait(&b.v_full[st], (g / KV_STAGES) & 1);
      ptx::tc_fence_after();
      const uint32_t v_lo = ptx::bcast0(v_lo0 + st * (KV_BYTES >> 4));
      const uint32_t acc0 = ptx::bcast0(pc.j != 0);
      if (ptx::elect_one()) {
#pragma unroll
        for (int kk = 0; kk < BK / 16; ++kk)
          ptx::umma_ts_lo(tm_s + TM_O, tm_s + TM_P + kk * 8, v_lo + kk * (2048 >> 4), vdesc_hi, idesc_pv, kk == 0 ? acc0 : 1u);
        ptx::umma_commit(&b.v_empty[st]);
        ptx::umma_commit(b.pv_done);
      }
      __syncwarp();
      adv(pc, s);
    }
  } else {
    // ------------------------------------------------------------------ softmax + output: thread = one query row of a slot
    ptx::reg_alloc<SOFTMAX_REGS>();
    const int s = warp >> 2;
    const int lane_grp = warp & 3;
    const int row = lane_grp * 32 + lane;
    const SlotBars b = slot_bars(bars + s * SLOT_BARS);
    const uint32_t tbase = tmem + s * TM_SLOT + ((uint32_t)(lane_grp * 32) << 16);
    uint32_t g = 0, n = 0;
#ifdef ATTN2_PROFILE
    long long pr[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const long long t_begin = clock64();
#define AP(i, t0) pr[i] += clock64() - (t0)
#define AP_NOW() clock64()
#else
#define AP(i, t0)
#define AP_NOW() 0
#endif
    // The output of an item (O / l -> bf16) is written at the NEXT item's first "PV retired" point: the last PV of an item
    // then retires under the next item's exp pass instead of being waited for (3500 cycles per item in the first version).
    bool pending = false;
    float pend_inv = 0.f;
    __nv_bfloat16* pend_dst = nullptr;             // nullptr: padding row, nothing to store
    auto flush_output = [&]() {
      long long t0 = AP_NOW();
#pragma unroll
      for (int h2 = 0; h2 < 2; ++h2) {
        float o[32];
        ptx::tmem_ld32(tbase + TM_O + h2 * 32, o);
        if (h2 == 1) {                             // O is in registers: the next item's first PV may overwrite it
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(b.o_free);
        }
        if (pend_dst) {
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            uint4 pk;
            pk.x = pack2(o[i] * pend_inv, o[i + 1] * pend_inv);
            pk.y = pack2(o[i + 2] * pend_inv, o[i + 3] * pend_inv);
            pk.z = pack2(o[i + 4] * pend_inv, o[i + 5] * pend_inv);
            pk.w = pack2(o[i + 6] * pend_inv, o[i + 7] * pend_inv);
            *reinterpret_cast<uint4*>(pend_dst + h2 * 32 + i) = pk;
          }
        }
      }
      pending = false;
      AP(6, t0);
    };
    for (int k = 0; k < per_slot; ++k) {
      const AttnItem w = table[s * MAX_ITEMS + k];
      if (w.len == 0) continue;
      const int len = w.len, nblk = (len + BK - 1) / BK;
      float m_run = -INFINITY, m_use = -INFINITY, l_row = 0.f;   // running max; the max the exponent offset uses
      for (int j = 0; j < nblk; ++j, ++g) {
        long long t0 = AP_NOW();
        ptx::mbar_wait(b.s_full, g & 1);
        AP(0, t0); t0 = AP_NOW();
        ptx::tc_fence_after();
        uint32_t raw[CH][32];
#pragma unroll
        for (int c = 0; c < CH; ++c) ptx::tmem_ld32_nowait(tbase + TM_S + c * 32, raw[c]);
        ptx::tmem_wait_ld();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(b.s_free);               // S(g) is in registers: QK(g+1) may overwrite it
        AP(1, t0); t0 = AP_NOW();
        const int nvalid = len - j * BK;                         // valid keys among the block's BK columns
        if (nvalid < BK) {
#pragma unroll
          for (int c = 0; c < CH; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (c * 32 + i >= nvalid) raw[c][i] = 0xff800000u;  // -inf
        }
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int c = 0; c < CH; ++c)
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            mx[0] = fmaxf(mx[0], __uint_as_float(raw[c][i]));
            mx[1] = fmaxf(mx[1], __uint_as_float(raw[c][i + 1]));
            mx[2] = fmaxf(mx[2], __uint_as_float(raw[c][i + 2]));
            mx[3] = fmaxf(mx[3], __uint_as_float(raw[c][i + 3]));
          }
        const float m_new = fmaxf(m_run, fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])));
        const bool need = j == 0 || (m_new - m_use) * scale_log2 > RESCALE_LOG2;
        float corr = 1.f;
        if (need) {
          corr = j == 0 ? 0.f : ex2_approx((m_use - m_new) * scale_log2);
          m_use = m_new;
        }
        m_run = m_new;
        const float moff = m_use * scale_log2;
        AP(2, t0); t0 = AP_NOW();
        // exp2 of the whole row into packed bf16 registers FIRST; only the TMEM stores wait for PV(g-1) (which still reads
        // P(g-1)), so that wait is off the critical path.
        // (Tried and dropped: a per-scheduler lock that makes the slots' exp passes take turns -- the lock hand-over cost more
        // than the MUFU sharing it avoided.)
        float ls[4] = {0.f, 0.f, 0.f, 0.f};
        uint32_t pk[CH][16];
#pragma unroll
        for (int c = 0; c < CH; ++c) {
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            float p0, p1;
            const float x0 = fmaf(__uint_as_float(raw[c][i]), scale_log2, -moff);
            const float x1 = fmaf(__uint_as_float(raw[c][i + 1]), scale_log2, -moff);
            if (ATTN2_POLY_EVERY > 0 && ((i >> 1) % (ATTN2_POLY_EVERY > 0 ? ATTN2_POLY_EVERY : 1)) == ATTN2_POLY_EVERY - 1) {
              p0 = ex2_poly(x0);                                 // masked columns: 2^-126
              p1 = ex2_poly(x1);
            } else {
              p0 = ex2_approx(x0);                               // masked columns: exp2(-inf) = 0
              p1 = ex2_approx(x1);
            }
            ls[(i >> 1) & 3] += p0 + p1;
            pk[c][i >> 1] = pack2(p0, p1);
          }
        }
        AP(4, t0); t0 = AP_NOW();
        if (g > 0) ptx::mbar_wait(b.pv_done, (g - 1) & 1);       // PV(g-1) retired: P and O may be touched
        AP(3, t0); t0 = AP_NOW();
        ptx::tc_fence_after();
        if (pending) flush_output();                             // previous item's O (its last PV has just retired)
#pragma unroll
        for (int c = 0; c < CH; ++c) ptx::tmem_st16(tbase + TM_P + c * 16, pk[c]);
        l_row = l_row * corr + ((ls[0] + ls[1]) + (ls[2] + ls[3]));
        if (j > 0 && __any_sync(0xffffffffu, need)) {            // rare: some row of this warp moved its offset
          float o[32];
#pragma unroll
          for (int h2 = 0; h2 < 2; ++h2) {
            ptx::tmem_ld32(tbase + TM_O + h2 * 32, o);
            uint32_t lo[16], hi[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) { lo[i] = __float_as_uint(o[i] * corr); hi[i] = __float_as_uint(o[16 + i] * corr); }
            ptx::tmem_st16(tbase + TM_O + h2 * 32, lo);
            ptx::tmem_st16(tbase + TM_O + h2 * 32 + 16, hi);
          }
        }
        ptx::tmem_wait_st();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(b.p_full);
        AP(5, t0);
      }
      pending = true;
      pend_inv = 1.f / l_row;
      pend_dst = (w.q0 + row < len) ? out + (size_t)(w.seq0 + w.q0 + row) * ldo + w.head * HD : nullptr;
      ++n;
    }
    if (pending) {
      ptx::mbar_wait(b.pv_done, (g - 1) & 1);
      ptx::tc_fence_after();
      flush_output();
    }
    (void)n;
#ifdef ATTN2_PROFILE
    if (blockIdx.x == 0 && lane == 0 && (warp == 0 || warp == 4))
      printf("attn2 prof warp %d: total %lld blocks %u items %u | s_full %lld ld %lld max %lld pv_done %lld exp+st %lld rescale+arrive %lld epilogue %lld\n",
             warp, clock64() - t_begin, g, n, pr[0], pr[1], pr[2], pr[3], pr[4], pr[5], pr[6]);
#endif
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == MMA_WARP) ptx::tmem_dealloc(tmem, TM_COLS);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn2() {
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

int launch_attention_tc2(const void* qkv, int ld, void* out, int ldo, const Layout& lay, int heads, float scale,
                         cudaStream_t st) {
  CVB_REQUIRE(heads > 0 && ld >= 3 * heads * HD && ldo >= heads * HD, "bad attention shapes");
  CVB_REQUIRE(ld % 8 == 0 && ldo % 8 == 0, "leading dimensions must keep 16-byte alignment");
  CVB_REQUIRE((reinterpret_cast<uintptr_t>(qkv) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0, "alignment");
  if (lay.rows <= 0) return CVB_OK;
  EncodeTiledFn fn = encode_fn2();
  CVB_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  CUtensorMap tm;
  cuuint64_t dims[2] = {(cuuint64_t)(3 * heads * HD), (cuuint64_t)lay.rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {64, 128};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(qkv), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  CUtensorMap tmkv;
  cuuint32_t boxkv[2] = {64, (cuuint32_t)BK};
  if (r == CUDA_SUCCESS)
    r = fn(&tmkv, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(qkv), dims, strides, boxkv, estr,
           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled (attention) failed with CUresult " + std::to_string((int)r));
    return CVB_ERR_CUDA;
  }
  static std::once_flag attr_once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(attr_once, [] {
    attr_err = cudaFuncSetAttribute(attention_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
  });
  CVB_CUDA(attr_err);
  const int n_items = (lay.rows / 128) * heads;
  const int sms = device_sm_count();
  const int want = (n_items + SLOTS - 1) / SLOTS;
  const int grid = want < sms ? want : sms;
  // a CTA's item table holds MAX_ITEMS entries per work stream: larger problems go out as several launches
  const int per_launch = MAX_ITEMS * SLOTS * grid;
  for (int item0 = 0; item0 < n_items; item0 += per_launch) {
    const int n = n_items - item0 < per_launch ? n_items - item0 : per_launch;
    CVB_CUDA(launch_pdl(attention_tc2_kernel, dim3(grid), dim3(NUM_THREADS), SMEM_BYTES, st, tm, tmkv,
                        reinterpret_cast<__nv_bfloat16*>(out), ldo, lay, heads, scale * 1.4426950408889634f, item0, n));
    CVB_LAUNCH_CHECK();
  }
  return CVB_OK;
}

}  // namespace cvb
