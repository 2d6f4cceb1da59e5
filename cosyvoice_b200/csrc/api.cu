// extern "C" surface of libcvb200.so (declared in include/cvb200.h).  No torch types cross this boundary.
#include <atomic>
#include "common.cuh"
#include "tapgemm.cuh"

namespace cvb {

namespace {
thread_local std::string g_last_error;
std::atomic<long long> g_launches{0};
}  // namespace

void set_last_error(const std::string& msg) { g_last_error = msg; }
const char* get_last_error() { return g_last_error.c_str(); }
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

static std::atomic<int> g_pdl{-1};
bool pdl_enabled() {
  int v = g_pdl.load(std::memory_order_relaxed);
  if (v < 0) {
    const char* e = getenv("CVB200_PDL");
    v = (e && e[0] == '0') ? 0 : 1;
    g_pdl.store(v, std::memory_order_relaxed);
  }
  return v != 0;
}
void set_pdl_enabled(bool on) { g_pdl.store(on ? 1 : 0, std::memory_order_relaxed); }

using Layout = cvb200_layout;
int launch_layernorm(const float*, int, void*, int, int, const float*, const float*, int, int, float, cudaStream_t);
int launch_layernorm_act(const void*, int, int, void*, int, int, const float*, const float*, int, int, float, cudaStream_t);
int launch_norm_mish(const void*, int, int, void*, int, int, const float*, const float*, const double*, const float*, int,
                     const Layout&, int, float, cudaStream_t);
int launch_time_sinusoid(const float*, int, int, float, float*, cudaStream_t);
int launch_small_linear(const float*, int, int, const float*, const float*, int, int, int, float*, cudaStream_t);
int launch_pack_cfg(const float*, const float*, const float*, const float*, void*, int, const Layout&, cudaStream_t);
int launch_euler_cfg(float*, const float*, long long, float, float, cudaStream_t);
int launch_nct_to_rows(const float*, long long, long long, int, void*, int, int, const Layout&, cudaStream_t);
int launch_rows_to_nct(const float*, int, int, float*, long long, long long, const Layout&, cudaStream_t);
int launch_attention(const void*, int, void*, int, int, const Layout&, int, float, cudaStream_t);
int launch_attention_tc(const void*, int, void*, int, const Layout&, int, float, cudaStream_t);
int launch_attention_tc2(const void*, int, void*, int, const Layout&, int, float, cudaStream_t);
int launch_rowdot_abs(const void*, int, int, const float*, const float*, int, float*, int, cudaStream_t);
int launch_sine_source(const float*, const Layout&, int, float, const float*, const float*, const float*, const float*,
                       double*, float*, long long, cudaStream_t);
int launch_stft16(const float*, long long, void*, int, int, const Layout&, cudaStream_t);
int launch_istft16(const float*, int, const Layout&, float*, long long, float, cudaStream_t);
int launch_snake(const void*, int, int, void*, int, int, const float*, int, const Layout&, cudaStream_t);
int launch_reflect_row0(void*, int, int, int, const Layout&, cudaStream_t);
int launch_ffn_fused(const void*, int, const void*, const float*, const void*, const float*, float*, int, void*, int,
                     const cvb200_layout&, cudaStream_t);
int launch_block_ffn(void*, int, float, const void*, const float*, const float*, const void*, const float*, void*, int, float*,
                     const cvb200_layout&, cudaStream_t);
int launch_block_out_ffn(void*, int, const void*, int, const void*, const float*, float, const void*, const float*, const float*,
                         const void*, const float*, void*, int, float*, const cvb200_layout&, cudaStream_t);
int launch_split3_bf16(const float*, int, int, int, void*, int, long long, cudaStream_t);
int launch_fade_in_out(float*, long long, int, const float*, long long, int, const double*, int, int, cudaStream_t);
int launch_interp_linear(const float*, long long, int, float*, long long, int, int, cudaStream_t);
int launch_pcm16(const float*, short*, float, long long, cudaStream_t);
int launch_anti_alias_rows(const void*, int, int, void*, int, int, const float*, const float*, const float*, int,
                           const Layout&, cudaStream_t);
int launch_anti_alias_nct(const void*, void*, int, int, int, int, const float*, const float*, const float*, const float*,
                          cudaStream_t);

int launch_embed_tokens(const int*, int, const float*, int, int, void*, int, int, const Layout&, cudaStream_t);
int launch_layernorm_seq(const float*, int, void*, int, int, const float*, const float*, const Layout&, int, float, cudaStream_t);
int launch_relpos_attention(const float*, int, const float*, int, int, int, const float*, const float*, void*, int, int,
                            const Layout&, int, float, int, cudaStream_t);
int launch_groupnorm1_mish(const float*, int, void*, int, int, const float*, const float*, int, const Layout&, float, cudaStream_t);
int launch_interp_rows(const float*, int, const Layout&, void*, int, int, const Layout&, const int*, int, cudaStream_t);
int launch_l2_normalize(const float*, float*, int, int, cudaStream_t);

}  // namespace cvb

using namespace cvb;
// the ctypes mirror in cosyvoice_b200/_lib.py relies on this layout (tests/test_lib_cpu.py)
static_assert(sizeof(cvb200_epilogue) == 168, "cvb200_epilogue layout");
static_assert(sizeof(cvb200_gemm) == 392, "cvb200_gemm layout");
static_assert(sizeof(cvb200_layout) == 48, "cvb200_layout layout");
#define ST(s) reinterpret_cast<cudaStream_t>(s)
#define NEED(p, what)                                                                    \
  do {                                                                                   \
    if ((p) == nullptr) {                                                                \
      set_last_error(std::string("invalid argument: null ") + what);                     \
      return CVB_ERR_INVALID;                                                            \
    }                                                                                    \
  } while (0)

extern "C" {

const char* cvb200_last_error(void) { return get_last_error(); }
int cvb200_version(void) { return 100; }
int64_t cvb200_launch_count(void) { return (int64_t)g_launches.load(); }
int cvb200_set_pdl(int on) {
  const int prev = pdl_enabled() ? 1 : 0;
  set_pdl_enabled(on != 0);
  return prev;
}

int cvb200_tapgemm(const cvb200_gemm* p, int engine, void* stream) {
  NEED(p, "gemm descriptor");
  NEED(p->A, "A"); NEED(p->W, "W"); NEED(p->tile_seq, "tile_seq"); NEED(p->tile_row0, "tile_row0"); NEED(p->seq_len, "seq_len");
  if (p->m_tiles == 0) return CVB_OK;
  if (engine == CVB200_ENGINE_TCGEN05) return launch_tapgemm_tc(*p, ST(stream));
  if (engine == CVB200_ENGINE_SIMT) return launch_tapgemm_simt(*p, ST(stream));
  set_last_error("invalid argument: unknown engine");
  return CVB_ERR_INVALID;
}

int cvb200_ffn_fused(const void* a, int lda, const void* w1, const float* b1, const void* w2, const float* b2, float* x,
                     int ldx, void* out1, int ldo1, const cvb200_layout* lay, void* stream) {
  CVB_REQUIRE(lay != nullptr, "null layout");
  return launch_ffn_fused(a, lda, w1, b1, w2, b2, x, ldx, out1, ldo1, *lay, ST(stream));
}

int cvb200_block_ffn(void* x, int ldx, float eps, const void* w1f, const float* u1, const float* b1f, const void* w2,
                     const float* b2, void* out1, int ldo1, float* stats_out, const cvb200_layout* lay, void* stream) {
  CVB_REQUIRE(lay != nullptr, "null layout");
  return launch_block_ffn(x, ldx, eps, w1f, u1, b1f, w2, b2, out1, ldo1, stats_out, *lay, ST(stream));
}

int cvb200_block_out_ffn(void* x, int ldx, const void* o, int ldo, const void* wo, const float* bo, float eps, const void* w1f,
                         const float* u1, const float* b1f, const void* w2, const float* b2, void* out1, int ldo1,
                         float* stats_out, const cvb200_layout* lay, void* stream) {
  CVB_REQUIRE(lay != nullptr, "null layout");
  return launch_block_out_ffn(x, ldx, o, ldo, wo, bo, eps, w1f, u1, b1f, w2, b2, out1, ldo1, stats_out, *lay, ST(stream));
}

int cvb200_split3_bf16(const float* x, int ldx, int C, int cpad, void* out, int ldo, int64_t rows, void* stream) {
  return launch_split3_bf16(x, ldx, C, cpad, out, ldo, (long long)rows, ST(stream));
}

int cvb200_fade_in_out(float* x, int64_t ldx, int x_len, const float* prev, int64_t ldp, int prev_len,
                       const double* window, int window_len, int rows, void* stream) {
  return launch_fade_in_out(x, (long long)ldx, x_len, prev, (long long)ldp, prev_len, window, window_len, rows, ST(stream));
}

int cvb200_interp_linear(const float* x, int64_t ldx, int t_in, float* out, int64_t ldo, int t_out, int rows,
                         void* stream) {
  return launch_interp_linear(x, (long long)ldx, t_in, out, (long long)ldo, t_out, rows, ST(stream));
}

int cvb200_pcm16(const float* wav, int16_t* out, float scale, int64_t n, void* stream) {
  return launch_pcm16(wav, reinterpret_cast<short*>(out), scale, (long long)n, ST(stream));
}

int cvb200_layernorm(const float* x, int ldx, void* out, int ldo, int out_dt, const float* gamma, const float* beta,
                     int rows, int C, float eps, void* stream) {
  NEED(x, "x"); NEED(out, "out"); NEED(gamma, "gamma"); NEED(beta, "beta");
  return launch_layernorm(x, ldx, out, ldo, out_dt, gamma, beta, rows, C, eps, ST(stream));
}

int cvb200_layernorm_act(const void* x, int ldx, int x_dt, void* out, int ldo, int out_dt, const float* gamma,
                         const float* beta, int rows, int C, float eps, void* stream) {
  NEED(x, "x"); NEED(out, "out"); NEED(gamma, "gamma"); NEED(beta, "beta");
  return launch_layernorm_act(x, ldx, x_dt, out, ldo, out_dt, gamma, beta, rows, C, eps, ST(stream));
}

int cvb200_norm_mish(const void* h, int ldh, int h_dt, void* out, int ldo, int out_dt, const float* gamma,
                     const float* beta, const double* gn_stats, const float* seq_bias, int seq_bias_ld,
                     const cvb200_layout* lay, int C, float eps, void* stream) {
  NEED(h, "h"); NEED(out, "out"); NEED(gamma, "gamma"); NEED(beta, "beta"); NEED(lay, "layout");
  return launch_norm_mish(h, ldh, h_dt, out, ldo, out_dt, gamma, beta, gn_stats, seq_bias, seq_bias_ld, *lay, C, eps,
                          ST(stream));
}

int cvb200_attention_ex(const void* qkv, int ld, void* out, int ldo, int dt, const cvb200_layout* lay, int heads,
                        float scale, int impl, void* stream) {
  NEED(qkv, "qkv"); NEED(out, "out"); NEED(lay, "layout");
  if (dt == CVB200_BF16 && impl == CVB200_ATTN_TCGEN05)
    return launch_attention_tc(qkv, ld, out, ldo, *lay, heads, scale, ST(stream));
  if (dt == CVB200_BF16 && impl == CVB200_ATTN_TCGEN05_V2)
    return launch_attention_tc2(qkv, ld, out, ldo, *lay, heads, scale, ST(stream));
  return launch_attention(qkv, ld, out, ldo, dt, *lay, heads, scale, ST(stream));
}

int cvb200_attention(const void* qkv, int ld, void* out, int ldo, int dt, const cvb200_layout* lay, int heads,
                     float scale, void* stream) {
  return cvb200_attention_ex(qkv, ld, out, ldo, dt, lay, heads, scale, CVB200_ATTN_TCGEN05_V2, stream);
}

int cvb200_time_sinusoid(const float* t, int S, int dim, float scale, float* out, void* stream) {
  NEED(t, "t"); NEED(out, "out");
  return launch_time_sinusoid(t, S, dim, scale, out, ST(stream));
}

int cvb200_small_linear(const float* x, int S, int K, const float* W, const float* b, int N, int pre_act,
                        int post_act, float* y, void* stream) {
  NEED(x, "x"); NEED(W, "W"); NEED(y, "y");
  return launch_small_linear(x, S, K, W, b, N, pre_act, post_act, y, ST(stream));
}

int cvb200_pack_cfg_input(const float* x, const float* mu, const float* cond, const float* spks, void* out,
                          int out_dt, const cvb200_layout* lay1, void* stream) {
  NEED(x, "x"); NEED(mu, "mu"); NEED(cond, "cond"); NEED(spks, "spks"); NEED(out, "out"); NEED(lay1, "layout");
  return launch_pack_cfg(x, mu, cond, spks, out, out_dt, *lay1, ST(stream));
}

int cvb200_euler_cfg(float* x, const float* v, int64_t n_half, float dt, float cfg_rate, void* stream) {
  NEED(x, "x"); NEED(v, "v");
  return launch_euler_cfg(x, v, (long long)n_half, dt, cfg_rate, ST(stream));
}

int cvb200_rowdot_abs(const void* x, int ldx, int x_dt, const float* w, const float* b, int K, float* out, int rows,
                      void* stream) {
  NEED(x, "x"); NEED(w, "w"); NEED(b, "b"); NEED(out, "out");
  return launch_rowdot_abs(x, ldx, x_dt, w, b, K, out, rows, ST(stream));
}

int cvb200_sine_source(const float* f0, const cvb200_layout* lay0, int hop, float sr, const float* noise,
                       const float* phase, const float* lin_w, const float* lin_b, double* prefix_ws, float* s,
                       int64_t l_max, void* stream) {
  NEED(f0, "f0"); NEED(lay0, "layout"); NEED(noise, "noise"); NEED(phase, "phase"); NEED(lin_w, "lin_w");
  NEED(lin_b, "lin_b"); NEED(prefix_ws, "prefix_ws"); NEED(s, "s");
  return launch_sine_source(f0, *lay0, hop, sr, noise, phase, lin_w, lin_b, prefix_ws, s, (long long)l_max, ST(stream));
}

int cvb200_stft16(const float* s, int64_t l_max, void* out, int ldo, int out_dt, const cvb200_layout* lay,
                  void* stream) {
  NEED(s, "s"); NEED(out, "out"); NEED(lay, "layout");
  return launch_stft16(s, (long long)l_max, out, ldo, out_dt, *lay, ST(stream));
}

int cvb200_istft16(const float* x, int ldx, const cvb200_layout* lay, float* wav, int64_t l_max, float audio_limit,
                   void* stream) {
  NEED(x, "x"); NEED(lay, "layout"); NEED(wav, "wav");
  return launch_istft16(x, ldx, *lay, wav, (long long)l_max, audio_limit, ST(stream));
}

int cvb200_snake(const void* x, int ldx, int x_dt, void* out, int ldo, int out_dt, const float* alpha, int C,
                 const cvb200_layout* lay, void* stream) {
  NEED(x, "x"); NEED(out, "out"); NEED(alpha, "alpha"); NEED(lay, "layout");
  return launch_snake(x, ldx, x_dt, out, ldo, out_dt, alpha, C, *lay, ST(stream));
}

int cvb200_reflect_row0(void* x, int ldx, int dt, int C, const cvb200_layout* lay, void* stream) {
  NEED(x, "x"); NEED(lay, "layout");
  return launch_reflect_row0(x, ldx, dt, C, *lay, ST(stream));
}

int cvb200_anti_alias_act_nct(const void* x, void* y, int dt, int B, int C, int T, const float* up_filter,
                              const float* down_filter, const float* alpha, const float* beta, void* stream) {
  if (B > 0 && C > 0 && T > 0) {
    NEED(x, "x"); NEED(y, "y"); NEED(up_filter, "up_filter"); NEED(down_filter, "down_filter");
    NEED(alpha, "alpha"); NEED(beta, "beta");
  }
  if (dt != CVB200_F32 && dt != CVB200_BF16) {
    set_last_error("invalid argument: unsupported dtype (the reference raises AT_ERROR, type_shim.h:41-42)");
    return CVB_ERR_INVALID;
  }
  return launch_anti_alias_nct(x, y, dt, B, C, T, up_filter, down_filter, alpha, beta, ST(stream));
}

int cvb200_anti_alias_act_rows(const void* x, int ldx, int x_dt, void* out, int ldo, int out_dt, const float* filt,
                               const float* alpha, const float* beta, int C, const cvb200_layout* lay, void* stream) {
  NEED(x, "x"); NEED(out, "out"); NEED(filt, "filt"); NEED(alpha, "alpha"); NEED(beta, "beta"); NEED(lay, "layout");
  return launch_anti_alias_rows(x, ldx, x_dt, out, ldo, out_dt, filt, alpha, beta, C, *lay, ST(stream));
}

int cvb200_nct_to_rows(const float* src, int64_t seq_stride, int64_t c_stride, int C, void* out, int ldo, int out_dt,
                       const cvb200_layout* lay, void* stream) {
  NEED(src, "src"); NEED(out, "out"); NEED(lay, "layout");
  return launch_nct_to_rows(src, (long long)seq_stride, (long long)c_stride, C, out, ldo, out_dt, *lay, ST(stream));
}

int cvb200_rows_to_nct(const float* src, int lds, int C, float* out, int64_t seq_stride, int64_t c_stride,
                       const cvb200_layout* lay, void* stream) {
  NEED(src, "src"); NEED(out, "out"); NEED(lay, "layout");
  return launch_rows_to_nct(src, lds, C, out, (long long)seq_stride, (long long)c_stride, *lay, ST(stream));
}

int cvb200_embed_tokens(const int32_t* tokens, int tok_ld, const float* table, int vocab, int C, void* out, int ldo, int out_dt,
                        const cvb200_layout* lay, void* stream) {
  NEED(tokens, "tokens"); NEED(table, "table"); NEED(out, "out"); NEED(lay, "layout");
  return launch_embed_tokens(tokens, tok_ld, table, vocab, C, out, ldo, out_dt, *lay, ST(stream));
}

int cvb200_layernorm_seq(const float* x, int ldx, void* out, int ldo, int out_dt, const float* gamma, const float* beta,
                         const cvb200_layout* lay, int C, float eps, void* stream) {
  NEED(x, "x"); NEED(out, "out"); NEED(gamma, "gamma"); NEED(beta, "beta"); NEED(lay, "layout");
  return launch_layernorm_seq(x, ldx, out, ldo, out_dt, gamma, beta, *lay, C, eps, ST(stream));
}

int cvb200_relpos_attention(const float* qkv, int ld, const float* pos, int ldp, int center, int n_rel, const float* bias_u,
                            const float* bias_v, void* out, int ldo, int out_dt, const cvb200_layout* lay, int heads,
                            float scale, int chunk, void* stream) {
  NEED(qkv, "qkv"); NEED(pos, "pos"); NEED(bias_u, "pos_bias_u"); NEED(bias_v, "pos_bias_v"); NEED(out, "out"); NEED(lay, "layout");
  return launch_relpos_attention(qkv, ld, pos, ldp, center, n_rel, bias_u, bias_v, out, ldo, out_dt, *lay, heads, scale, chunk,
                                 ST(stream));
}

int cvb200_groupnorm1_mish(const float* h, int ldh, void* out, int ldo, int out_dt, const float* gamma, const float* beta,
                           int C, const cvb200_layout* lay, float eps, void* stream) {
  NEED(h, "h"); NEED(out, "out"); NEED(gamma, "gamma"); NEED(beta, "beta"); NEED(lay, "layout");
  return launch_groupnorm1_mish(h, ldh, out, ldo, out_dt, gamma, beta, C, *lay, eps, ST(stream));
}

int cvb200_interp_rows(const float* x, int ldx, const cvb200_layout* src_lay, void* out, int ldo, int out_dt,
                       const cvb200_layout* dst_lay, const int32_t* segs, int C, void* stream) {
  NEED(x, "x"); NEED(out, "out"); NEED(src_lay, "src layout"); NEED(dst_lay, "dst layout"); NEED(segs, "segs");
  return launch_interp_rows(x, ldx, *src_lay, out, ldo, out_dt, *dst_lay, segs, C, ST(stream));
}

int cvb200_l2_normalize(const float* x, float* out, int S, int K, void* stream) {
  NEED(x, "x"); NEED(out, "out");
  return launch_l2_normalize(x, out, S, K, ST(stream));
}

}  // extern "C"
