// Self-attention of the estimator's BasicTransformerBlock (8 heads x 64):
//   out[q, h*64 + d] = sum_j softmax_j(Q[q,h,:] . K[j,h,:] * scale) V[j,h,d],  j over the valid rows of q's sequence
// (cosyvoice/flow/components/transformer.py:266-271 -> diffusers Attention/AttnProcessor2_0; the reference's
// additive {0,1} mask of flow/decoder.py:258 is a uniform +1 inside an utterance, i.e. a softmax no-op).
//
// qkv is time-major [rows, ld] with row = [Q(512) | K(512) | V(512)].  One CTA per (128-row tile, head);
// tiles never span sequences, so the varlen batch needs no masking beyond key < len.
//   * bf16: flash-style single pass, warp-level tensor-core MMA (m16n8k16, fp32 accumulate), K/V
//     tiles double-buffered through shared memory with cp.async, online softmax in registers.
//   * fp32: CUDA-core parity path (one query per thread, keys broadcast from shared memory).
#include "common.cuh"
#include "../../include/cvb200.h"

namespace cvb {

using Layout = cvb200_layout;

namespace {

constexpr int HD = 64;      // head dim
constexpr int BQ = 128;     // queries per CTA (= layout tile)
constexpr int BKEY = 64;    // keys per step

// ------------------------------------------------------------------ small PTX helpers
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool pred) {
  const int bytes = pred ? 16 : 0;   // src-size 0 -> zero fill
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
// tile of [rows][64 bf16] = 128 B per row, 16-byte chunks XOR-swizzled by (row & 7)
__device__ __forceinline__ uint32_t swz(int row, int chunk) { return (uint32_t)(row * 128 + ((chunk ^ (row & 7)) << 4)); }

// Loads `nrows` rows x 64 columns (bf16) starting at global row g0 / column col0 into a swizzled tile.
// rows >= limit (relative to g0) are zero-filled.
__device__ __forceinline__ void load_tile(uint32_t tile_base, const __nv_bfloat16* __restrict__ src, int ld, long long g0,
                                          int col0, int nrows, int limit, int tid, int nthreads) {
  for (int i = tid; i < nrows * 8; i += nthreads) {
    const int r = i >> 3, c = i & 7;
    const bool ok = r < limit;
    const __nv_bfloat16* p = src + (size_t)(g0 + (ok ? r : 0)) * ld + col0 + c * 8;
    cp_async16(tile_base + swz(r, c), p, ok);
  }
}

// ------------------------------------------------------------------ bf16 flash kernel
// 4 warps; warp w owns query rows [32w, 32w+32) of the tile as two m16 blocks.
__global__ void __launch_bounds__(128) attention_bf16_kernel(const __nv_bfloat16* __restrict__ qkv, int ld,
                                                             __nv_bfloat16* __restrict__ out, int ldo, const Layout lay,
                                                             int heads, float scale_log2) {
  pdl_sync();
  extern __shared__ __align__(128) uint8_t smem[];
  const int tile = blockIdx.x, head = blockIdx.y;
  const int b = __ldg(lay.tile_seq + tile);
  const int q0 = __ldg(lay.tile_row0 + tile);
  const int len = __ldg(lay.seq_len + b);
  if (q0 >= len) return;                                    // padding-only tile
  const long long seq0 = __ldg(lay.seq_row0 + b);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int inner = heads * HD;

  const uint32_t sQ = smem_addr(smem);                      // 128 x 64 bf16 = 16 KiB
  const uint32_t sK = sQ + BQ * 128;                        // 2 stages x 8 KiB
  const uint32_t sV = sK + 2 * BKEY * 128;                  // 2 stages x 8 KiB

  load_tile(sQ, qkv, ld, seq0 + q0, head * HD, BQ, len - q0, tid, 128);
  const int nblk = (len + BKEY - 1) / BKEY;
  load_tile(sK, qkv, ld, seq0, inner + head * HD, BKEY, len, tid, 128);
  load_tile(sV, qkv, ld, seq0, 2 * inner + head * HD, BKEY, len, tid, 128);
  cp_async_commit();

  uint32_t qf[2][4][4];                                     // [m block][k step over d][A fragment]
  float o[2][8][4];                                         // [m block][n tile over d][C fragment]
  float row_max[2][2], row_sum[2][2];
#pragma unroll
  for (int mi = 0; mi < 2; ++mi) {
    row_max[mi][0] = row_max[mi][1] = -INFINITY;
    row_sum[mi][0] = row_sum[mi][1] = 0.f;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt)
#pragma unroll
      for (int i = 0; i < 4; ++i) o[mi][nt][i] = 0.f;
  }

  for (int kb = 0; kb < nblk; ++kb) {
    const int st = kb & 1;
    if (kb + 1 < nblk) {                                    // prefetch next K/V block into the other stage
      const int k0n = (kb + 1) * BKEY;
      load_tile(sK + (st ^ 1) * BKEY * 128, qkv, ld, seq0 + k0n, inner + head * HD, BKEY, len - k0n, tid, 128);
      load_tile(sV + (st ^ 1) * BKEY * 128, qkv, ld, seq0 + k0n, 2 * inner + head * HD, BKEY, len - k0n, tid, 128);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    if (kb == 0) {
#pragma unroll
      for (int mi = 0; mi < 2; ++mi)
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const int r = warp * 32 + mi * 16 + (lane & 15), c = ks * 2 + (lane >> 4);
          ldsm_x4(sQ + swz(r, c), qf[mi][ks]);
        }
    }
    const uint32_t kbase = sK + st * BKEY * 128, vbase = sV + st * BKEY * 128;

    // ---- S = Q K^T  (16 x 64 per m block)
    float s[2][8][4];
#pragma unroll
    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
      for (int nt = 0; nt < 8; ++nt)
#pragma unroll
        for (int i = 0; i < 4; ++i) s[mi][nt][i] = 0.f;
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
#pragma unroll
      for (int np = 0; np < 4; ++np) {                      // pairs of key n-tiles (16 keys)
        uint32_t kf[4];
        const int r = np * 16 + (lane & 7) + ((lane >> 4) << 3), c = ks * 2 + ((lane >> 3) & 1);
        ldsm_x4(kbase + swz(r, c), kf);
#pragma unroll
        for (int mi = 0; mi < 2; ++mi) {
          mma_bf16(s[mi][2 * np], qf[mi][ks], kf[0], kf[1]);
          mma_bf16(s[mi][2 * np + 1], qf[mi][ks], kf[2], kf[3]);
        }
      }
    }

    // ---- online softmax (exp2 domain)
    const int key0 = kb * BKEY;
    const bool tail = key0 + BKEY > len;
#pragma unroll
    for (int mi = 0; mi < 2; ++mi) {
      float mx[2] = {row_max[mi][0], row_max[mi][1]};
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float v = s[mi][nt][i] * scale_log2;
          if (tail) {
            const int key = key0 + nt * 8 + (lane & 3) * 2 + (i & 1);
            if (key >= len) v = -INFINITY;
          }
          s[mi][nt][i] = v;
          mx[i >> 1] = fmaxf(mx[i >> 1], v);
        }
      }
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 1));
        mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 2));
      }
      float corr[2];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        corr[h] = exp2f(row_max[mi][h] - mx[h]);            // first block: exp2(-inf) = 0
        row_max[mi][h] = mx[h];
        row_sum[mi][h] *= corr[h];
      }
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float p = exp2f(s[mi][nt][i] - mx[i >> 1]);
          s[mi][nt][i] = p;
          row_sum[mi][i >> 1] += p;
          o[mi][nt][i] *= corr[i >> 1];
        }
      }
    }

    // ---- O += P V
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {                        // 16 keys per step
      uint32_t pf[2][4];
#pragma unroll
      for (int mi = 0; mi < 2; ++mi) {
        pf[mi][0] = pack_bf16(s[mi][2 * ks][0], s[mi][2 * ks][1]);
        pf[mi][1] = pack_bf16(s[mi][2 * ks][2], s[mi][2 * ks][3]);
        pf[mi][2] = pack_bf16(s[mi][2 * ks + 1][0], s[mi][2 * ks + 1][1]);
        pf[mi][3] = pack_bf16(s[mi][2 * ks + 1][2], s[mi][2 * ks + 1][3]);
      }
#pragma unroll
      for (int np = 0; np < 4; ++np) {                      // pairs of d n-tiles (16 columns)
        uint32_t vf[4];
        const int r = ks * 16 + (lane & 7) + (((lane >> 3) & 1) << 3), c = np * 2 + (lane >> 4);
        ldsm_x4_trans(vbase + swz(r, c), vf);
#pragma unroll
        for (int mi = 0; mi < 2; ++mi) {
          mma_bf16(o[mi][2 * np], pf[mi], vf[0], vf[1]);
          mma_bf16(o[mi][2 * np + 1], pf[mi], vf[2], vf[3]);
        }
      }
    }
    __syncthreads();                                        // stage st may be overwritten by the next prefetch
  }

  // ---- normalise and store
#pragma unroll
  for (int mi = 0; mi < 2; ++mi) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float l = row_sum[mi][h];
      l += __shfl_xor_sync(0xffffffffu, l, 1);
      l += __shfl_xor_sync(0xffffffffu, l, 2);
      const float inv = 1.f / l;
      const int lrow = q0 + warp * 32 + mi * 16 + h * 8 + (lane >> 2);
      if (lrow < len) {
        __nv_bfloat16* dst = out + (size_t)(seq0 + lrow) * ldo + head * HD + (lane & 3) * 2;
#pragma unroll
        for (int nt = 0; nt < 8; ++nt)
          *reinterpret_cast<uint32_t*>(dst + nt * 8) = pack_bf16(o[mi][nt][h * 2] * inv, o[mi][nt][h * 2 + 1] * inv);
      }
    }
  }
}

// ------------------------------------------------------------------ fp32 parity kernel
// 128 threads = 128 queries of one tile; K/V blocks of 32 keys staged in shared memory.
__global__ void __launch_bounds__(128) attention_f32_kernel(const float* __restrict__ qkv, int ld, float* __restrict__ out,
                                                            int ldo, const Layout lay, int heads, float scale) {
  pdl_sync();
  __shared__ float sK[32][HD];
  __shared__ float sV[32][HD];
  const int tile = blockIdx.x, head = blockIdx.y;
  const int b = __ldg(lay.tile_seq + tile);
  const int q0 = __ldg(lay.tile_row0 + tile);
  const int len = __ldg(lay.seq_len + b);
  if (q0 >= len) return;
  const long long seq0 = __ldg(lay.seq_row0 + b);
  const int inner = heads * HD;
  const int tid = threadIdx.x;
  const int lrow = q0 + tid;
  const bool active = lrow < len;
  float q[HD], acc[HD];
#pragma unroll
  for (int d = 0; d < HD; ++d) { q[d] = 0.f; acc[d] = 0.f; }
  if (active) {
    const float* qp = qkv + (size_t)(seq0 + lrow) * ld + head * HD;
#pragma unroll
    for (int d = 0; d < HD; d += 4) {
      const float4 t = *reinterpret_cast<const float4*>(qp + d);
      q[d] = t.x * scale; q[d + 1] = t.y * scale; q[d + 2] = t.z * scale; q[d + 3] = t.w * scale;
    }
  }
  float m = -INFINITY, l = 0.f;
  for (int k0 = 0; k0 < len; k0 += 32) {
    __syncthreads();
    for (int i = tid; i < 32 * (HD / 4); i += 128) {
      const int r = i / (HD / 4), c = (i % (HD / 4)) * 4;
      float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), vv = kv;
      if (k0 + r < len) {
        const float* base = qkv + (size_t)(seq0 + k0 + r) * ld + head * HD + c;
        kv = *reinterpret_cast<const float4*>(base + inner);
        vv = *reinterpret_cast<const float4*>(base + 2 * inner);
      }
      *reinterpret_cast<float4*>(&sK[r][c]) = kv;
      *reinterpret_cast<float4*>(&sV[r][c]) = vv;
    }
    __syncthreads();
    const int nk = min(32, len - k0);
    float sc[32];
    float bm = m;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      float d0 = 0.f;
#pragma unroll
      for (int d = 0; d < HD; ++d) d0 = fmaf(q[d], sK[j][d], d0);
      sc[j] = j < nk ? d0 : -INFINITY;
      bm = fmaxf(bm, sc[j]);
    }
    const float corr = expf(m - bm);
    m = bm;
    l *= corr;
#pragma unroll
    for (int d = 0; d < HD; ++d) acc[d] *= corr;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const float p = expf(sc[j] - m);
      l += p;
#pragma unroll
      for (int d = 0; d < HD; ++d) acc[d] = fmaf(p, sV[j][d], acc[d]);
    }
  }
  if (active) {
    const float inv = 1.f / l;
    float* op = out + (size_t)(seq0 + lrow) * ldo + head * HD;
#pragma unroll
    for (int d = 0; d < HD; d += 4)
      *reinterpret_cast<float4*>(op + d) = make_float4(acc[d] * inv, acc[d + 1] * inv, acc[d + 2] * inv, acc[d + 3] * inv);
  }
}

}  // namespace

int launch_attention(const void* qkv, int ld, void* out, int ldo, int dt, const Layout& lay, int heads, float scale,
                     cudaStream_t st) {
  CVB_REQUIRE(heads > 0 && ld >= 3 * heads * HD && ldo >= heads * HD, "bad attention shapes");
  CVB_REQUIRE(ld % 8 == 0 && ldo % 8 == 0, "leading dimensions must keep 16-byte alignment");
  if (lay.rows <= 0) return CVB_OK;
  dim3 grid(lay.rows / 128, heads);
  if (dt == DT_BF16) {
    constexpr int SMEM = BQ * 128 + 4 * BKEY * 128;   // 48 KiB
    static bool attr_done = false;
    if (!attr_done) {
      CVB_CUDA(cudaFuncSetAttribute(attention_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
      attr_done = true;
    }
    CVB_CUDA(launch_pdl(attention_bf16_kernel, dim3(grid), dim3(128), SMEM, st, reinterpret_cast<const __nv_bfloat16*>(qkv), ld,
                                                   reinterpret_cast<__nv_bfloat16*>(out), ldo, lay, heads,
                                                   scale * 1.4426950408889634f));
  } else {
    CVB_CUDA(launch_pdl(attention_f32_kernel, dim3(grid), dim3(128), 0, st, reinterpret_cast<const float*>(qkv), ld, reinterpret_cast<float*>(out), ldo,
                                               lay, heads, scale));
  }
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

}  // namespace cvb
