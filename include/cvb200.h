/* cvb200 -- C ABI of the B200-native (sm_100a) acoustic back-end kernels for CosyVoice.
 *
 * Drop-in boundary.  The reference (duj12/CosyVoice) composes its hot
 * path from PyTorch/cuDNN/cuBLAS calls plus ONE native operator,
 *   extern "C" torch::Tensor fwd_cuda(input, up_filter, down_filter, alpha, beta)
 *   (cosyvoice/BigVGAN/alias_free_activation/cuda/anti_alias_activation_cuda.cu:214-256, bound at
 *    anti_alias_activation.cpp:19-23, called from activation1d.py:24-28).
 * This library replaces those calls with hand-written kernels behind plain-C entry points:
 * device pointers + sizes + a cudaStream_t (passed as void*), caller-owned buffers, int status
 * (0 = ok) and cvb200_last_error() for the message -- the raw-pointer convention the reference's
 * own TensorRT estimator branch documents (cosyvoice/flow/flow_matching.py:151-163).
 *
 * All activations inside the library are TIME-MAJOR: a buffer is [rows, ld] with one row per
 * frame/sample and channels contiguous.  A batch is varlen-packed: sequence b occupies rows
 * [seq_row0[b], seq_row0[b] + seq_stride[b]) (a multiple of 128), its first seq_len[b] rows are
 * valid, the rest are ZERO (they are the conv zero padding between utterances).  Every op writes
 * all rows of its layout and keeps the invalid rows zero.
 */
#ifndef CVB200_H
#define CVB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CVB200_OK 0
#define CVB200_ERR_INVALID 1
#define CVB200_ERR_CUDA 2

#define CVB200_F32 0
#define CVB200_BF16 1

#define CVB200_ENGINE_SIMT 0    /* CUDA-core FFMA (fp32 parity mode, cross-check) */
#define CVB200_ENGINE_TCGEN05 1 /* tcgen05.mma + TMEM + TMA (bf16 operands, fp32 accumulate) */

#define CVB200_MAX_SEGS 32

/* activation codes (epilogue `act`, cvb200_small_linear pre/post) */
#define CVB200_ACT_NONE 0
#define CVB200_ACT_GELU 1  /* exact erf GELU: diffusers GELU(approximate="none"), components/transformer.py:109-110 */
#define CVB200_ACT_ELU 2   /* hifigan/f0_predictor.py:30-48 */
#define CVB200_ACT_SNAKE 3 /* x + sin^2(a x)/(a+1e-9): transformer/activation.py:79-82 */
#define CVB200_ACT_LRELU 4 /* hifigan/generator.py:355,374 */
#define CVB200_ACT_MISH 5  /* flow/components/decoder.py:36,47 */
#define CVB200_ACT_TANH 6  /* BigVGAN/bigvgan.py:433-434 */
#define CVB200_ACT_SILU 7  /* flow/components/decoder.py:90 (TimestepEmbedding act_fn) */

#define CVB200_RES_NONE 0
#define CVB200_RES_F32 1
#define CVB200_RES_ACT 2

#define CVB200_OUT1_NONE 0
#define CVB200_OUT1_COPY 1
#define CVB200_OUT1_SNAKE 2
#define CVB200_OUT1_LRELU 3

/* Fused epilogue of the tap GEMM.  Per output element (row m, phase, column n):
 *   acc = row_stats[m][0] * acc + row_stats[m][1] * row_u[n]   (only when row_stats != NULL: folded LayerNorm)
 *   v = acc + bias[n] + seq_bias[seq][n];  v = act(v);  v += res[m][phase*phase_col + n];
 *   v *= out_scale;  if (accumulate) v += out0[m][...];  rows beyond the valid length -> 0;
 *   out0[m][phase*phase_col + n] = v;  out1[...] = out1_fn(v);
 *   gn_stats[seq][n/32] += (v, v*v)      (GroupNorm(8,256) statistics, components/decoder.py:35) */
typedef struct cvb200_epilogue {
  const float* bias;       /* [N] or NULL */
  const float* seq_bias;   /* [n_seq][seq_bias_ld] or NULL (time-embedding / speaker conditioning) */
  const float* act_alpha;  /* ACT_SNAKE: per-channel alpha [N] */
  const void* res;         /* residual [rows, ldr] or NULL */
  void* out0;              /* primary output [rows, ldo0] */
  void* out1;              /* optional second output [rows, ldo1] */
  const float* out1_alpha; /* OUT1_SNAKE alpha [N] */
  double* gn_stats;        /* [n_seq][gn_groups][2] or NULL */
  int32_t seq_bias_ld;
  int32_t act;
  float act_slope;
  int32_t res_kind;
  int32_t ldr;
  int32_t res_dt;
  float out_scale;
  int32_t accumulate;
  int32_t ldo0;
  int32_t out0_dt;
  int32_t phase_col;
  int32_t out1_mode;
  int32_t ldo1;
  int32_t out1_dt;
  float out1_slope;
  int32_t gn_groups;
  /* LayerNorm folded into the GEMM (A = the RAW rows, W = W * gamma): acc = row_stats[m][0] * acc + row_stats[m][1] * row_u[n]
   * before the bias, with row_stats[m] = (rstd, -mean * rstd) of input row m, row_u[n] = sum_k W[n][k], bias = b + W . beta */
  const float* row_stats;  /* [rows][2] or NULL */
  const float* row_u;      /* [n_pad] */
  /* The same fold with the statistics carried as PARTIAL sums, so that the GEMM that PRODUCES a 256-wide row stream can hand
   * them to the GEMM that consumes it and no LayerNorm launch is left (small layouts, where launches are the cost):
   *   row_parts_out[m][n/32] = (sum, sum of squares) of the stored (rounded) values of columns [32 (n/32), +32) of output row m
   *                            -- written by a GEMM with N == 256, n_phase == 1;
   *   row_parts[m][0..8)     -- read by the consumer: mean = sum/256, rstd = rsqrt(sumsq/256 - mean^2 + row_eps), then as
   *                            row_stats.  tcgen05 engine only. */
  const float* row_parts;  /* [rows][8][2] or NULL (excludes row_stats) */
  float* row_parts_out;    /* [rows][8][2] or NULL */
  float row_eps;
  int32_t reserved0;
} cvb200_epilogue;

/* The one contraction of the library ("tap GEMM"):
 *   D[m, phase, n] = sum_seg sum_{c<seg_len} A[m + seg_row[phase][seg], seg_col[phase][seg] + c]
 *                                            * W[phase][n][seg*seg_len + c]
 * A is a 2-D row-major view [a_rows, a_cols] (leading dimension lda) of a time-major buffer; reads
 * outside the view return zero.  Linear, Conv1d (any k/dilation/causal pad), strided Conv1d (view
 * [rows/s, s*C]) and ConvTranspose1d (one phase per output residue) are all instances:
 *   flow/components/decoder.py:34,52,65,142  flow/decoder.py:83-84,205
 *   flow/components/transformer.py:110,126 + diffusers Attention to_q/k/v/out
 *   hifigan/generator.py:58-78,268-312  hifigan/f0_predictor.py:28-50  BigVGAN/bigvgan.py:61-85,303-377 */
typedef struct cvb200_gemm {
  const void* A;
  const void* W;            /* packed [n_phase][n_pad][n_seg*seg_len], same dtype as A */
  const int32_t* tile_seq;  /* [m_tiles] sequence of each 128-row tile */
  const int32_t* tile_row0; /* [m_tiles] row of the tile's first row within its sequence */
  const int32_t* seq_len;   /* [n_seq] valid rows (M space) */
  const int32_t* out_limit; /* [n_seq] or NULL: output (row*n_phase+phase) valid iff < out_limit */
  int32_t lda, a_rows, a_cols;
  int32_t n_pad, N, n_phase, n_seg, seg_len;
  int32_t dt;
  int32_t m_tiles;
  int32_t bn;               /* N tile of the tcgen05 engine: multiple of 16, <= 256, divides n_pad */
  int16_t seg_row[CVB200_MAX_SEGS];
  int16_t seg_col[CVB200_MAX_SEGS];
  cvb200_epilogue epi;
} cvb200_gemm;

/* varlen layout handed to the non-GEMM kernels */
typedef struct cvb200_layout {
  const int32_t* tile_seq;
  const int32_t* tile_row0;
  const int32_t* seq_len;
  const int32_t* seq_row0; /* [n_seq] first row of each sequence */
  int32_t n_seq;
  int32_t rows;            /* total rows (multiple of 128) */
  int32_t max_len;
} cvb200_layout;

const char* cvb200_last_error(void);
int cvb200_version(void);
/* number of kernels this library has launched in this process (bench.py's gpu_launches) */
int64_t cvb200_launch_count(void);
/* programmatic dependent launch of the library's kernels (default on; env CVB200_PDL=0 turns it off).  Returns the
 * previous setting.  Kernels launched while a CUDA graph is being captured keep the setting of capture time. */
int cvb200_set_pdl(int on);

int cvb200_tapgemm(const cvb200_gemm* p, int engine, void* stream);

/* Fused FeedForward + residual of the estimator's BasicTransformerBlock (bf16 operands, tcgen05):
 *   x[m, 0:256] += W2 . gelu(W1 . a[m, 0:256] + b1) + b2        cosyvoice/flow/components/transformer.py:83-136
 *   (FeedForward: GELU(Linear 256->1024), Linear 1024->256), :309-314 (ff_output + hidden_states).
 * a: bf16 [rows, lda] (the LayerNorm output), w1: bf16 [1024, 256], w2: bf16 [256, 1024] (nn.Linear layouts),
 * b1 [1024], b2 [256] fp32, x: fp32 [rows, ldx] read and written in place, out1: optional bf16 [rows, ldo1] copy of
 * the new x.  Rows past a sequence's valid length are written as zeros.  The hidden activations stay in TMEM. */
int cvb200_ffn_fused(const void* a, int lda, const void* w1, const float* b1, const void* w2, const float* b2, float* x,
                     int ldx, void* out1, int ldo1, const cvb200_layout* lay, void* stream);

/* Second half of a BasicTransformerBlock on the bf16 residual stream, one kernel (csrc/block_ffn_tc.cu):
 *   x += W2 . gelu(W1 . LayerNorm(x; g3, be3) + b1) + b2   (in place)
 *        norm3 + ff + residual, cosyvoice/flow/components/transformer.py:300, :83-136, :309-314
 *   stats_out[row] = (rstd, -mean * rstd) of the new row (optional): what norm1 of the NEXT block (:262) needs, consumed
 *        by a tap GEMM through epilogue.row_stats
 *   out1 = x_new (optional bf16 copy)
 * The LayerNorm is folded into the first GEMM by the caller: w1f = bf16(W1 * g3[None, :]) [1024, 256],
 * u1[j] = sum_k w1f[j, k] (fp32, of the ROUNDED weights), b1f = b1 + W1 . be3; then W1 . LN(x) + b1 = rstd (w1f . x) - rstd mean u1
 * + b1f, with the row statistics computed inside the kernel.  x: bf16 [rows, ldx >= 256]; w2 bf16 [256, 1024]
 * (nn.Linear layout); b2 [256] fp32.  Rows past a sequence's valid length are written as zeros (x, out1; stats 0, 0).
 * The [rows, 1024] hidden activations never leave the SM (TMEM) and no normalised copy of x is written. */
int cvb200_block_ffn(void* x, int ldx, float eps, const void* w1f, const float* u1, const float* b1f, const void* w2,
                     const float* b2, void* out1, int ldo1, float* stats_out, const cvb200_layout* lay, void* stream);

/* The same with the block's attention out-projection in front, still one kernel (csrc/block_ffn_tc.cu, OUTP):
 *   x  = bf16((o . W_out^T + b_out) + x)                      attn1's to_out[0] + residual add,
 *                                                             cosyvoice/flow/components/transformer.py:266-276 (diffusers
 *                                                             Attention.to_out[0], `hidden_states = attn_output + hidden_states`)
 *   x += W2 . gelu(W1 . LayerNorm(x; g3, be3) + b1) + b2      as cvb200_block_ffn
 * o: the attention output, bf16 [rows, ldo >= 512] (heads concatenated); wo: bf16 [256, 512] (nn.Linear layout); bo [256] fp32.
 * The intermediate x is rounded to bf16 exactly as the separate out-projection GEMM (cvb200_tapgemm with a RES_ACT residual)
 * writes it, lives only in shared memory / registers, and the two forms give bit-identical results. */
int cvb200_block_out_ffn(void* x, int ldx, const void* o, int ldo, const void* wo, const float* bo, float eps, const void* w1f,
                         const float* u1, const float* b1f, const void* w2, const float* b2, void* out1, int ldo1,
                         float* stats_out, const cvb200_layout* lay, void* stream);

/* fp32 -> three bf16 operands with x = x1 + x2 + x3 (24 mantissa bits): out[row, i*cpad + c] = x_i[row, c], i = 0..2.
 * Lets the f0 predictor (cosyvoice/hifigan/f0_predictor.py:52-55, whose output is integrated into the source phase)
 * run on the bf16 tensor cores with split operands: one tap GEMM over the cross terms x_i * w_k of lowest order. */
int cvb200_split3_bf16(const float* x, int ldx, int C, int cpad, void* out, int ldo, int64_t rows, void* stream);

/* Streaming cross-fade, in place on x [rows, x_len] (row pitch ldx floats): with n = window_len / 2,
 * x[r, i] = x[r, i] * window[i] + prev[r, prev_len - n + i] * window[n + i] for i < n, computed in double and rounded
 * to fp32 once -- fade_in_out, cosyvoice/utils/common.py:149-157 (the reference does it on the host after two .cpu()
 * copies; cli/model.py:150 for the mel overlap, :163/:174 and :354/:365 for the speech overlap).  window is the float64
 * np.hamming(2 * n) table in device memory.  CVB200_ERR_INVALID when x_len or prev_len < n (the reference raises). */
int cvb200_fade_in_out(float* x, int64_t ldx, int x_len, const float* prev, int64_t ldp, int prev_len,
                       const double* window, int window_len, int rows, void* stream);

/* Speed change on the generated mel: F.interpolate(tts_mel, size=t_out, mode='linear') of cli/model.py:171 / :362
 * (align_corners=False) over rows = 80 mel bins; x [rows, t_in] -> out [rows, t_out]. */
int cvb200_interp_linear(const float* x, int64_t ldx, int t_in, float* out, int64_t ldo, int t_out, int rows,
                         void* stream);

/* int16 PCM wire format of a finished waveform: out[i] = (int16) trunc(wav[i] * scale), saturated to [-32768, 32767].
 * save_audio, cosyvoice/BigVGAN/utils.py:97-101 (scale = MAX_WAV_VALUE = 32768, numpy .astype("int16")). */
int cvb200_pcm16(const float* wav, int16_t* out, float scale, int64_t n, void* stream);

/* LayerNorm(256) over channels of each row: flow/components/transformer.py:195,231,262,300 */
int cvb200_layernorm(const float* x, int ldx, void* out, int ldo, int out_dt, const float* gamma,
                     const float* beta, int rows, int C, float eps, void* stream);

/* The same LayerNorm reading a residual stream kept in the operand dtype x_dt (the bf16 engine's stream; the reference's
 * flow.half() path, cli/model.py:43-45, keeps it in half precision too). */
int cvb200_layernorm_act(const void* x, int ldx, int x_dt, void* out, int ldo, int out_dt, const float* gamma,
                         const float* beta, int rows, int C, float eps, void* stream);

/* Block1D tail: out = Mish(Norm(h)) on valid rows (+ seq_bias[seq]), 0 elsewhere.
 * gn_stats != NULL: GroupNorm(8, C) with the statistics accumulated by the producing GEMM
 * (flow/components/decoder.py:35-36,39-41,56); NULL: per-row LayerNorm(C) (flow/decoder.py:41-44). */
int cvb200_norm_mish(const void* h, int ldh, int h_dt, void* out, int ldo, int out_dt,
                     const float* gamma, const float* beta, const double* gn_stats,
                     const float* seq_bias, int seq_bias_ld, const cvb200_layout* lay, int C, float eps,
                     void* stream);

/* softmax(Q K^T * scale) V per (sequence, head); qkv row = [Q(heads*64) | K | V]; keys are the
 * valid rows of the query's own sequence (diffusers AttnProcessor2_0 at batch 1 with the all-ones
 * additive mask of flow/decoder.py:258). */
int cvb200_attention(const void* qkv, int ld, void* out, int ldo, int dt, const cvb200_layout* lay,
                     int heads, float scale, void* stream);
/* same, with an explicit kernel choice for bf16: 0 = first tcgen05/TMEM flash kernel (two CTAs per SM, two threads per
 * query row), 1 = warp-level mma.sync flash kernel (kept as the on-device cross-check), 2 = second tcgen05/TMEM kernel
 * (one CTA per SM running two independent work streams, one softmax thread per query row; default of cvb200_attention) */
#define CVB200_ATTN_TCGEN05 0
#define CVB200_ATTN_MMA_SYNC 1
#define CVB200_ATTN_TCGEN05_V2 2
int cvb200_attention_ex(const void* qkv, int ld, void* out, int ldo, int dt, const cvb200_layout* lay,
                        int heads, float scale, int impl, void* stream);

/* SinusoidalPosEmb(dim, scale 1000): flow/components/decoder.py:18-27.  t [S] -> out [S, dim] */
int cvb200_time_sinusoid(const float* t, int S, int dim, float scale, float* out, void* stream);
/* y[s][n] = post(b[n] + sum_k pre(x[s][k]) W[n][k]), fp32 (time MLP, components/decoder.py:47,103-115) */
int cvb200_small_linear(const float* x, int S, int K, const float* W, const float* b, int N, int pre_act,
                        int post_act, float* y, void* stream);

/* estimator input concat [x | mu | spks | cond] for the CFG pair (flow/decoder.py:243-249 +
 * flow_matching.py:103-108): sequences [0,B) conditional, [B,2B) unconditional (mu=spks=cond=0).
 * x, mu, cond: [rows1, 80] fp32 time-major over the B-sequence layout `lay1`; out [2*rows1, 320]. */
int cvb200_pack_cfg_input(const float* x, const float* mu, const float* cond, const float* spks,
                          void* out, int out_dt, const cvb200_layout* lay1, void* stream);
/* x += dt * ((1+cfg) * v[r] - cfg * v[rows1 + r])   (flow_matching.py:116-119) */
int cvb200_euler_cfg(float* x, const float* v, int64_t n_half, float dt, float cfg_rate, void* stream);

/* f0 = |x[r] . w + b|  (hifigan/f0_predictor.py:53-55 classifier + abs) */
int cvb200_rowdot_abs(const void* x, int ldx, int x_dt, const float* w, const float* b, int K, float* out,
                      int rows, void* stream);
/* SineGen + SourceModuleHnNSF merge (hifigan/generator.py:138-168,204-220) for B utterances.
 * f0: [rows0] per-frame f0 in layout lay0 (frame rate); noise [B][9][L_max] fp32; phase [B][9];
 * lin_w[9], lin_b[1]; s out [B][L_max] fp32 (sample n valid for n < hop*len).  */
int cvb200_sine_source(const float* f0, const cvb200_layout* lay0, int hop, float sr, const float* noise,
                       const float* phase, const float* lin_w, const float* lin_b, double* prefix_ws /*[rows0]*/,
                       float* s, int64_t l_max, void* stream);
/* cvb200_sine_source without a caller workspace.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  The fp64 frame prefix lives in a fixed array inside the library's device image
 * (CVB200_NOALLOC_MAX_ROWS doubles = 512 KiB on every GPU that loads the library, until the process exits); the call
 * makes no cudaMalloc / cudaFree / allocating library call.  Inputs and outputs are the caller's, as above.
 * lay0->rows > CVB200_NOALLOC_MAX_ROWS -> CVB200_ERR_INVALID with a message naming the limit.  Every call on a device
 * shares the one array: calls that may overlap on the GPU (different streams, no ordering) RACE -- serialise them.
 * Nothing in it needs clearing between calls (each call writes every element it reads).  The first call on a device
 * resolves the array's address (cudaGetSymbolAddress): make it outside stream capture. */
#define CVB200_NOALLOC_MAX_ROWS 65536
int cvb200_sine_source_noalloc(const float* f0, const cvb200_layout* lay0, int hop, float sr, const float* noise,
                               const float* phase, const float* lin_w, const float* lin_b, float* s, int64_t l_max,
                               void* stream);
/* torch.stft(n_fft 16, hop 4, periodic hann, center/reflect) -> [real(9) | imag(9) | 0...] per frame
 * row (hifigan/generator.py:333-339).  s [B][l_max]; out rows follow layout `lay` (frame rate),
 * frame f of sequence b valid for f < seq_len[b] (= L_b/4 + 1). */
int cvb200_stft16(const float* s, int64_t l_max, void* out, int ldo, int out_dt, const cvb200_layout* lay,
                  void* stream);
/* exp / sin / inverse 16-point DFT / windowed overlap-add / clamp (generator.py:341-347,376-381).
 * x [rows, ldx] fp32 (18 used columns) in layout `lay` (frames); wav [B][l_max], L_b = 4*(F_b-1). */
int cvb200_istft16(const float* x, int ldx, const cvb200_layout* lay, float* wav, int64_t l_max,
                   float audio_limit, void* stream);
/* y = snake_alpha(x) on valid rows (0 elsewhere); x fp32 or bf16 */
int cvb200_snake(const void* x, int ldx, int x_dt, void* out, int ldo, int out_dt, const float* alpha,
                 int C, const cvb200_layout* lay, void* stream);
/* ReflectionPad1d((1,0)) fix-up (generator.py:358-359): row 0 of every sequence := row 2. */
int cvb200_reflect_row0(void* x, int ldx, int dt, int C, const cvb200_layout* lay, void* stream);

/* Drop-in for the reference's native operator fwd_cuda (anti_alias_activation_cuda.cu:214-256):
 * x, y [B, C, T] contiguous (fp32 or bf16), filters fp32 [12], alpha/beta fp32 [C] (log scale). */
int cvb200_anti_alias_act_nct(const void* x, void* y, int dt, int B, int C, int T, const float* up_filter,
                              const float* down_filter, const float* alpha, const float* beta, void* stream);
/* same operator on a time-major varlen buffer (used inside the BigVGAN trunk) */
int cvb200_anti_alias_act_rows(const void* x, int ldx, int x_dt, void* out, int ldo, int out_dt,
                               const float* filt, const float* alpha, const float* beta, int C,
                               const cvb200_layout* lay, void* stream);
/* out[r][c] += seq_bias[seq][c] on valid rows (BigVGAN speaker conditioning, bigvgan.py:407,417-418)
 * -- also available fused as epilogue.seq_bias */

/* layout transposes between the reference's [C, T] tensors and time-major rows */
/* element (b, c, t) of the [B, C, T] side lives at base + b*seq_stride + c*c_stride + t */
int cvb200_nct_to_rows(const float* src, int64_t seq_stride, int64_t c_stride, int C, void* out, int ldo,
                       int out_dt, const cvb200_layout* lay, void* stream);
int cvb200_rows_to_nct(const float* src, int lds, int C, float* out, int64_t seq_stride, int64_t c_stride,
                       const cvb200_layout* lay, void* stream);

/* ---- flow front-end: what flow.inference does before the CFM ---- */

/* nn.Embedding gather: out[row] = table[clamp(token, 0, vocab-1)] on valid rows, 0 elsewhere.
 * flow/flow.py:131-135 (input_embedding(torch.clamp(token, min=0)) * mask), :288-290.  tokens int32 [n_seq][tok_ld]
 * (token t of sequence b at tokens[b*tok_ld + t]); table fp32 [vocab][C]. */
int cvb200_embed_tokens(const int32_t* tokens, int tok_ld, const float* table, int vocab, int C, void* out, int ldo, int out_dt,
                        const cvb200_layout* lay, void* stream);

/* LayerNorm(C = 256 | 512) per valid row of a varlen layout, zeros on the invalid rows (they are the next conv's padding):
 * transformer/subsampling.py:83-86 (eps 1e-5), transformer/encoder_layer.py:145-146 (eps 1e-12), transformer/encoder.py:104. */
int cvb200_layernorm_seq(const float* x, int ldx, void* out, int ldo, int out_dt, const float* gamma, const float* beta,
                         const cvb200_layout* lay, int C, float eps, void* stream);

/* RelPositionMultiHeadedAttention core, transformer/attention.py:243-330 (ESPnet rel-pos, rel_shift as an index):
 *   s[i,j] = ((q_i + u_h) . k_j + (q_i + v_h) . pos[center - (i - j)][h]) * scale,  out = softmax_j(s) V,
 *   key j visible iff j < len and (chunk == 0 or j < (i / chunk + 1) * chunk)   (utils/mask.py:116-124,191-197).
 * qkv fp32 [rows, ld] = [Q(heads*64) | K | V] per row; pos fp32 [n_rel, ldp] = linear_pos(PE(r)) at row center - r
 * (transformer/embedding.py:224-254); bias_u / bias_v fp32 [heads*64]; out [rows, ldo] fp32 or bf16. */
int cvb200_relpos_attention(const float* qkv, int ld, const float* pos, int ldp, int center, int n_rel, const float* bias_u,
                            const float* bias_v, void* out, int ldo, int out_dt, const cvb200_layout* lay, int heads,
                            float scale, int chunk, void* stream);

/* GroupNorm(1, C) + Mish over each utterance (statistics over C channels x valid frames), zeros on invalid rows:
 * flow/length_regulator.py:33-38. */
int cvb200_groupnorm1_mish(const float* h, int ldh, void* out, int ldo, int out_dt, const float* gamma, const float* beta,
                           int C, const cvb200_layout* lay, float eps, void* stream);

/* InterpolateRegulator.inference's piecewise F.interpolate(mode='linear') (flow/length_regulator.py:52-70) on time-major
 * rows: segs int32 [n_seq][4][4] = (src_start, src_len, dst_start, dst_len) per piece, dst_len = 0 for an unused piece;
 * x fp32 rows in src_lay (token rate) -> out rows in dst_lay (mel rate). */
int cvb200_interp_rows(const float* x, int ldx, const cvb200_layout* src_lay, void* out, int ldo, int out_dt,
                       const cvb200_layout* dst_lay, const int32_t* segs, int C, void* stream);

/* F.normalize(x, dim=1) of the speaker embedding (flow/flow.py:127): x [S][K] -> x / max(||x||_2, 1e-12). */
int cvb200_l2_normalize(const float* x, float* out, int S, int K, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CVB200_H */
