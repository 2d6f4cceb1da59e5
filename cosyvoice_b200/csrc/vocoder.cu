// Memory-bound pieces of the HiFT / BigVGAN vocoders:
//   rowdot_abs        f0 classifier + abs                       hifigan/f0_predictor.py:53-55
//   sine_source       SineGen + SourceModuleHnNSF merge         hifigan/generator.py:138-168,204-220
//   stft16 / istft16  16-point (i)STFT, hop 4, Hann             hifigan/generator.py:333-347,376-381
//   snake             x + sin^2(a x)/(a + 1e-9)                 transformer/activation.py:79-82
//   reflect_row0      ReflectionPad1d((1,0)) fix-up             hifigan/generator.py:358-359
//   (the anti-aliased SnakeBeta activation lives in antialias.cu)
#include <mutex>
#include "common.cuh"
#include "../../include/cvb200.h"

// Scratch of cvb200_sine_source_noalloc (the frame prefix of the NSF source, fp64): a fixed array in the library's
// device image instead of a caller buffer, so the entry point never allocates.  One copy per GPU that loads the
// library, held until the process exits (CVB200_NOALLOC_MAX_ROWS * 8 bytes = 512 KiB).  In HOST code this name is only
// a stand-in for the symbol: its device address comes from cudaGetSymbolAddress (noalloc_scratch() below).
static __device__ __align__(256) double g_noalloc_prefix[CVB200_NOALLOC_MAX_ROWS];

namespace cvb {

using Layout = cvb200_layout;

namespace {

__device__ __forceinline__ bool row_info(const Layout& lay, int row, int& b, int& lrow, int& len) {
  const int tile = row >> 7;
  b = __ldg(lay.tile_seq + tile);
  lrow = __ldg(lay.tile_row0 + tile) + (row & 127);
  len = __ldg(lay.seq_len + b);
  return lrow < len;
}

// ------------------------------------------------------------------ f0 classifier
__global__ void __launch_bounds__(256) rowdot_abs_kernel(const void* __restrict__ x, int ldx, int x_dt,
                                                         const float* __restrict__ w, const float* __restrict__ bias, int K,
                                                         float* __restrict__ out, int rows) {
  pdl_sync();
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  float acc = 0.f;
  for (int k = lane; k < K; k += 32) acc = fmaf(ld_as_float(x, (size_t)row * ldx + k, x_dt), __ldg(w + k), acc);
  acc = warp_sum(acc);
  if (lane == 0) out[row] = fabsf(acc + __ldg(bias));
}

// ------------------------------------------------------------------ NSF sine source
// exclusive prefix (in cycles, fp64) of hop * f0 / sr over the frames of each sequence; one warp per sequence.
__global__ void __launch_bounds__(32) f0_prefix_kernel(const float* __restrict__ f0, const Layout lay, double cycles_per_hz,
                                                       double* __restrict__ prefix) {
  pdl_sync();
  const int b = blockIdx.x, lane = threadIdx.x;
  const int len = lay.seq_len[b];
  const long long r0 = lay.seq_row0[b];
  double carry = 0.0;
  for (int f0i = 0; f0i < len; f0i += 32) {
    const int f = f0i + lane;
    const double v = f < len ? (double)f0[r0 + f] * cycles_per_hz : 0.0;
    double inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const double t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (f < len) prefix[r0 + f] = carry + inc - v;
    carry += __shfl_sync(0xffffffffu, inc, 31);
  }
}

// one thread per output sample.  The reference integrates f0_up/sr with an fp32 cumsum over the whole
// utterance (generator.py:148); the closed form below (frame prefix + in-frame ramp, fp64) is the same
// sum without the fp32 rounding drift.
__global__ void __launch_bounds__(256) sine_source_kernel(const float* __restrict__ f0, const double* __restrict__ prefix,
                                                          const Layout lay, int hop, double inv_sr,
                                                          const float* __restrict__ noise, const float* __restrict__ phase,
                                                          const float* __restrict__ lin_w, const float* __restrict__ lin_b,
                                                          float* __restrict__ s, long long l_max) {
  pdl_sync();
  const int b = blockIdx.y;
  const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int len = __ldg(lay.seq_len + b);
  if (n >= (long long)len * hop) return;
  const long long r0 = __ldg(lay.seq_row0 + b);
  const int f = (int)(n / hop), i = (int)(n - (long long)f * hop);
  const float f0v = __ldg(f0 + r0 + f);
  const double c = prefix[r0 + f] + (double)(i + 1) * (double)f0v * inv_sr;     // cumsum(f0_up / sr) at sample n
  const float uv = f0v > 10.f ? 1.f : 0.f;
  const float namp = uv * 0.003f + (1.f - uv) * (0.1f / 3.f);
  float acc = __ldg(lin_b);
#pragma unroll
  for (int h = 0; h < 9; ++h) {
    const double ch = c * (double)(h + 1);
    const float theta = (float)(6.283185307179586 * (ch - floor(ch)));
    const float sine = 0.1f * sinf(theta + __ldg(phase + b * 9 + h));
    const float nz = __ldg(noise + ((size_t)b * 9 + h) * l_max + n);
    acc = fmaf(__ldg(lin_w + h), sine * uv + namp * nz, acc);
  }
  s[(size_t)b * l_max + n] = tanhf(acc);
}

// ------------------------------------------------------------------ STFT (n_fft 16, hop 4)
__global__ void __launch_bounds__(256) stft16_kernel(const float* __restrict__ s, long long l_max, void* __restrict__ out,
                                                     int ldo, int out_dt, const Layout lay) {
  pdl_sync();
  __shared__ float cs[16], sn[16], win[16];
  if (threadIdx.x < 16) {
    float sv, cv;
    sincospif((float)threadIdx.x * 0.125f, &sv, &cv);
    cs[threadIdx.x] = cv;
    sn[threadIdx.x] = sv;
    win[threadIdx.x] = 0.5f - 0.5f * cv;      // periodic Hann: 0.5 - 0.5 cos(2 pi n / 16)
  }
  __syncthreads();
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= lay.rows) return;
  int b, f, F;
  float re[9], im[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) re[k] = im[k] = 0.f;
  if (row_info(lay, row, b, f, F)) {
    const int L = (F - 1) * 4;
    const float* sp = s + (size_t)b * l_max;
    float x[16];
#pragma unroll
    for (int n = 0; n < 16; ++n) {
      int i = f * 4 + n - 8;
      if (i < 0) i = -i;
      if (i >= L) i = 2 * (L - 1) - i;
      x[n] = __ldg(sp + i) * win[n];
    }
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      float a = 0.f, c = 0.f;
#pragma unroll
      for (int n = 0; n < 16; ++n) {
        a = fmaf(x[n], cs[(k * n) & 15], a);
        c = fmaf(x[n], sn[(k * n) & 15], c);
      }
      re[k] = a;
      im[k] = -c;
    }
  }
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    st_from_float(out, (size_t)row * ldo + k, out_dt, re[k]);
    st_from_float(out, (size_t)row * ldo + 9 + k, out_dt, im[k]);
  }
}

// ------------------------------------------------------------------ iSTFT head
// one thread per hop (4 output samples); frames q-1 .. q+2 overlap it.
__global__ void __launch_bounds__(128) istft16_kernel(const float* __restrict__ x, int ldx, const Layout lay,
                                                      float* __restrict__ wav, long long l_max, float limit) {
  pdl_sync();
  __shared__ float cs[16], sn[16], win[16];
  if (threadIdx.x < 16) {
    float sv, cv;
    sincospif((float)threadIdx.x * 0.125f, &sv, &cv);
    cs[threadIdx.x] = cv;
    sn[threadIdx.x] = sv;
    win[threadIdx.x] = 0.5f - 0.5f * cv;
  }
  __syncthreads();
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= lay.rows) return;
  int b, q, F;
  if (!row_info(lay, row, b, q, F)) return;
  if (q >= F - 1) return;                         // L = 4 (F - 1) samples = F - 1 hops
  float acc[4] = {0.f, 0.f, 0.f, 0.f}, env[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int df = -1; df <= 2; ++df) {
    const int f = q + df;
    if (f < 0 || f >= F) continue;
    const float* xp = x + (size_t)(row + df) * ldx;
    float re[9], im[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const float mag = fminf(expf(__ldg(xp + k)), 100.f);
      const float ph = sinf(__ldg(xp + 9 + k));
      float sp, cp;
      sincosf(ph, &sp, &cp);
      re[k] = mag * cp;
      im[k] = mag * sp;
    }
    const int n0 = 8 - 4 * df;                    // sample offset inside frame f of this hop's first sample
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int n = n0 + r;
      float y = re[0] + ((n & 1) ? -re[8] : re[8]);            // DC and Nyquist (their imaginary parts are ignored)
#pragma unroll
      for (int k = 1; k < 8; ++k) y += 2.f * (re[k] * cs[(k * n) & 15] - im[k] * sn[(k * n) & 15]);
      const float w = win[n];
      acc[r] = fmaf(y * (1.f / 16.f), w, acc[r]);
      env[r] = fmaf(w, w, env[r]);
    }
  }
  float4 o;
  o.x = fminf(fmaxf(acc[0] / env[0], -limit), limit);
  o.y = fminf(fmaxf(acc[1] / env[1], -limit), limit);
  o.z = fminf(fmaxf(acc[2] / env[2], -limit), limit);
  o.w = fminf(fmaxf(acc[3] / env[3], -limit), limit);
  *reinterpret_cast<float4*>(wav + (size_t)b * l_max + (size_t)q * 4) = o;
}

// ------------------------------------------------------------------ snake / reflect
__global__ void __launch_bounds__(256) snake_kernel(const void* __restrict__ x, int ldx, int x_dt, void* __restrict__ out,
                                                    int ldo, int out_dt, const float* __restrict__ alpha, int C,
                                                    const Layout lay) {
  pdl_sync();
  const int cpr = C / 4;                           // 4-channel chunks per row
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)lay.rows * cpr) return;
  const int row = (int)(idx / cpr), c0 = (int)(idx % cpr) * 4;
  int b, lrow, len;
  const bool valid = row_info(lay, row, b, lrow, len);
  float v[4] = {0.f, 0.f, 0.f, 0.f};
  if (valid) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float xv = ld_as_float(x, (size_t)row * ldx + c0 + i, x_dt);
      const float a = __ldg(alpha + c0 + i);
      const float sv = sinf(xv * a);
      v[i] = xv + sv * sv / (a + 1e-9f);
    }
  }
  if (out_dt == DT_BF16) {
    __nv_bfloat162 p0 = __floats2bfloat162_rn(v[0], v[1]), p1 = __floats2bfloat162_rn(v[2], v[3]);
    uint2 pk;
    pk.x = *reinterpret_cast<uint32_t*>(&p0);
    pk.y = *reinterpret_cast<uint32_t*>(&p1);
    *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(out) + (size_t)row * ldo + c0) = pk;
  } else {
    *reinterpret_cast<float4*>(reinterpret_cast<float*>(out) + (size_t)row * ldo + c0) = make_float4(v[0], v[1], v[2], v[3]);
  }
}

__global__ void reflect_row0_kernel(void* __restrict__ x, int ldx, int dt, int C, const Layout lay) {
  pdl_sync();
  const int b = blockIdx.x;
  if (__ldg(lay.seq_len + b) < 3) return;
  const size_t r0 = (size_t)__ldg(lay.seq_row0 + b);
  for (int c = threadIdx.x; c < C; c += blockDim.x)
    st_from_float(x, r0 * ldx + c, dt, ld_as_float(x, (r0 + 2) * ldx + c, dt));
}

}  // namespace

// ------------------------------------------------------------------ launchers
int launch_rowdot_abs(const void* x, int ldx, int x_dt, const float* w, const float* b, int K, float* out, int rows,
                      cudaStream_t st) {
  if (rows <= 0) return CVB_OK;
  CVB_CUDA(launch_pdl(rowdot_abs_kernel, dim3(ceil_div(rows, 8)), dim3(256), 0, st, x, ldx, x_dt, w, b, K, out, rows));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_sine_source(const float* f0, const Layout& lay0, int hop, float sr, const float* noise, const float* phase,
                       const float* lin_w, const float* lin_b, double* prefix_ws, float* s, long long l_max,
                       cudaStream_t st) {
  CVB_REQUIRE(hop > 0 && sr > 0.f, "bad hop / sample rate");
  CVB_REQUIRE((long long)lay0.max_len * hop <= l_max, "l_max smaller than the longest utterance");
  if (lay0.n_seq <= 0) return CVB_OK;
  CVB_CUDA(launch_pdl(f0_prefix_kernel, dim3(lay0.n_seq), dim3(32), 0, st, f0, lay0, (double)hop / (double)sr, prefix_ws));
  CVB_LAUNCH_CHECK();
  const long long nmax = (long long)lay0.max_len * hop;
  dim3 grid((unsigned)((nmax + 255) / 256), lay0.n_seq);
  CVB_CUDA(launch_pdl(sine_source_kernel, dim3(grid), dim3(256), 0, st, f0, prefix_ws, lay0, hop, 1.0 / (double)sr, noise, phase, lin_w, lin_b, s, l_max));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

// Device address of g_noalloc_prefix on the CURRENT device, resolved once per device with that device current.
static int noalloc_scratch(double** out) {
  constexpr int kMaxDev = 64;
  static std::mutex mu;
  static double* ptr[kMaxDev] = {};
  int dev = -1;
  CVB_CUDA(cudaGetDevice(&dev));
  CVB_REQUIRE(dev >= 0 && dev < kMaxDev, "device ordinal beyond the no-alloc scratch table");
  std::lock_guard<std::mutex> lock(mu);
  if (ptr[dev] == nullptr) {
    double* p = nullptr;
    CVB_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_prefix));
    CVB_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null scratch address");
    ptr[dev] = p;
  }
  *out = ptr[dev];
  return CVB_OK;
}

int launch_sine_source_noalloc(const float* f0, const Layout& lay0, int hop, float sr, const float* noise,
                               const float* phase, const float* lin_w, const float* lin_b, float* s, long long l_max,
                               cudaStream_t st) {
  if (lay0.rows > CVB200_NOALLOC_MAX_ROWS) {
    set_last_error("invalid argument: cvb200_sine_source_noalloc: layout has " + std::to_string(lay0.rows) +
                   " frame rows, the built-in scratch holds CVB200_NOALLOC_MAX_ROWS = " +
                   std::to_string(CVB200_NOALLOC_MAX_ROWS) + " (use cvb200_sine_source with a caller workspace)");
    return CVB_ERR_INVALID;
  }
  double* scratch = nullptr;
  CVB_TRY(noalloc_scratch(&scratch));
  // Nothing in the scratch has to start at zero: f0_prefix_kernel writes prefix[r] for every valid frame row r and
  // sine_source_kernel reads only those rows, so the previous call's contents are never observed.
  return launch_sine_source(f0, lay0, hop, sr, noise, phase, lin_w, lin_b, scratch, s, l_max, st);
}

int launch_stft16(const float* s, long long l_max, void* out, int ldo, int out_dt, const Layout& lay, cudaStream_t st) {
  CVB_REQUIRE(ldo >= 18, "stft output needs 18 columns");
  if (lay.rows <= 0) return CVB_OK;
  CVB_CUDA(launch_pdl(stft16_kernel, dim3(ceil_div(lay.rows, 256)), dim3(256), 0, st, s, l_max, out, ldo, out_dt, lay));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_istft16(const float* x, int ldx, const Layout& lay, float* wav, long long l_max, float limit, cudaStream_t st) {
  CVB_REQUIRE(ldx >= 18 && l_max % 4 == 0, "istft needs 18 columns and l_max % 4 == 0");
  if (lay.rows <= 0) return CVB_OK;
  CVB_CUDA(launch_pdl(istft16_kernel, dim3(ceil_div(lay.rows, 128)), dim3(128), 0, st, x, ldx, lay, wav, l_max, limit));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_snake(const void* x, int ldx, int x_dt, void* out, int ldo, int out_dt, const float* alpha, int C,
                 const Layout& lay, cudaStream_t st) {
  CVB_REQUIRE(C % 4 == 0 && ldo % 4 == 0, "C and ldo must be multiples of 4");
  const long long total = (long long)lay.rows * (C / 4);
  if (total <= 0) return CVB_OK;
  CVB_CUDA(launch_pdl(snake_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, st, x, ldx, x_dt, out, ldo, out_dt, alpha, C, lay));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

int launch_reflect_row0(void* x, int ldx, int dt, int C, const Layout& lay, cudaStream_t st) {
  if (lay.n_seq <= 0) return CVB_OK;
  CVB_CUDA(launch_pdl(reflect_row0_kernel, dim3(lay.n_seq), dim3(128), 0, st, x, ldx, dt, C, lay));
  CVB_LAUNCH_CHECK();
  return CVB_OK;
}

}  // namespace cvb

// The one entry point of the library that owns its scratch (declared in include/cvb200.h next to cvb200_sine_source).
// It lives in this file because the scratch array's device address can only be taken where the symbol is visible.
extern "C" int cvb200_sine_source_noalloc(const float* f0, const cvb200_layout* lay0, int hop, float sr, const float* noise,
                                          const float* phase, const float* lin_w, const float* lin_b, float* s,
                                          int64_t l_max, void* stream) {
  CVB_REQUIRE(f0 && lay0 && noise && phase && lin_w && lin_b && s, "null argument");
  return cvb::launch_sine_source_noalloc(f0, *lay0, hop, sr, noise, phase, lin_w, lin_b, s, (long long)l_max,
                                         reinterpret_cast<cudaStream_t>(stream));
}
