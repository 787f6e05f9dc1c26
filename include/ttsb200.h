/*
 * ttsb200 — C ABI of the B200 (sm_100a) kernels behind the FastSpeech2 mel-generation path.
 *
 * The reference (syoamakase/Transformer_TTS) has no FFI layer: its hot path is Python `nn.Module`s that
 * call PyTorch library ops.  This header is the boundary a maintainer binds instead (ctypes stub in
 * INTEGRATION.md).  Each entry point names the reference code it replaces.  All pointers are DEVICE
 * pointers owned by the caller (PyTorch), all launches are asynchronous on `stream`, nothing is
 * allocated or freed here, and every function returns 0 or a negative code with
 * ttsb200_last_error_string() describing it.  No CPU fallback exists.
 *
 * Layout conventions: activations are dense (B, T, C) row-major ("time-major"), never (B, C, T).
 * bf16 = raw uint16 storage of __nv_bfloat16.  Masks are uint8 (B, T): 1 = valid, 0 = padded.
 */
#ifndef TTSB200_H_
#define TTSB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* ttsb200_stream_t; /* cudaStream_t */

#define TTSB200_OK 0
#define TTSB200_ERR_ARG (-1)
#define TTSB200_ERR_CUDA (-2)
#define TTSB200_ERR_TMAP (-3)

#define TTSB200_ABI_VERSION 200 /* round 2: gemm_args.tile_map/tile_count, the fp32-accurate chain, dropout */
int ttsb200_version(void);
const char* ttsb200_last_error_string(void);
/* hits / misses of the library's TMA-descriptor cache (keyed by pointer, shape, strides, box, swizzle) since load */
void ttsb200_debug_tmap_cache(unsigned long long* hits, unsigned long long* misses);

/* ------------------------------------------------------------------------------------------------
 * Implicit-GEMM conv1d / linear on tcgen05 tensor cores (TMA-fed, TMEM accumulators).
 *
 *   acc[b,t,n] = sum_{j<taps} sum_{c<C} x[b, t + j - pad_left, c] * w[j, n, c]      (zero outside [0,T))
 *
 * Replaces nn.Linear (taps=1) and nn.Conv1d in Models/modules.py:49-51,68,76-78,
 * Models/varianceadaptor.py:205-214, Models/postnets.py:26-42, Models/encoder.py:57, each together
 * with the elementwise / normalisation ops that follow it in the reference (epilogue modes below).
 * ---------------------------------------------------------------------------------------------- */
enum {
  TTSB200_EPI_AFFINE_ACT = 0,  /* v = act(acc*scale[n] + bias[n]) (+ res_f32)                      */
  TTSB200_EPI_RES_LN = 1,      /* v = acc + bias + res_scale*res_f32 ; out_f32=v ; out_bf16=LN1(v)  */
  TTSB200_EPI_FFN2 = 2,        /* u = LN1(acc+bias+res_bf16); z = res_f32+u; out_f32=z;
                                  out_bf16 = LN2(z) (out2_f32 = LN2(z) too if given)
                                  — Models/modules.py:84-87 + Models/layers.py:40 + next norm_1      */
  TTSB200_EPI_RELU_LN = 3,     /* out_bf16 = LN1(relu(acc+bias)) — Models/varianceadaptor.py:217-219 */
  TTSB200_EPI_RELU_LN_DOT = 4  /* out_row = mask ? dot(LN1(relu(acc+bias)), dot_w) + dot_b : 0
                                  — Models/varianceadaptor.py:220-229                               */
};
enum {
  TTSB200_ACT_NONE = 0,
  TTSB200_ACT_RELU = 1,
  TTSB200_ACT_TANH = 2,
  TTSB200_ACT_RELU_GRAD = 3 /* backward of ReLU: v = gate_bf16[b,t,n] > 0 ? v : 0 (dgrad GEMMs) */
};

typedef struct ttsb200_gemm_args {
  /* problem */
  int32_t B, T, N, C, taps, pad_left;
  int32_t block_n; /* AFFINE_ACT: 80, 128, 192 or 256; LN modes need block_n == N (384; 256 for RELU_LN*) */
  int32_t epi, act;
  const void* x; /* bf16 (B,T,ldx) */
  int64_t ldx;   /* elements, >= C, multiple of 8 */
  const void* w; /* bf16 [taps][N][C] */
  /* epilogue operands (NULL when unused) */
  const float* bias;      /* [N] */
  const float* scale;     /* [N] (AFFINE_ACT only) */
  const float* res_f32;   /* element (b,t,n) at res_f32[b*res_bstride + t*res_tstride + n] */
  int64_t res_bstride, res_tstride;
  const float* res_scale; /* device scalar multiplying res_f32 (RES_LN), NULL = 1 */
  const void* res_bf16;   /* bf16 (B,T,N) */
  const float *ln1_g, *ln1_b, *ln2_g, *ln2_b;
  float eps;
  const float* dot_w;      /* [N] */
  const float* dot_b;      /* device scalar */
  const uint8_t* row_mask; /* (B,T) */
  /* outputs (NULL when unused), dense (B,T,N) */
  float* out_f32;
  void* out_bf16;
  float* out2_f32;
  float* out_row; /* (B,T) */
  /* training extras (NULL in inference) */
  float* pre_f32;        /* LN epilogues: pre-LayerNorm value (B,T,N) saved for the backward pass */
  const void* gate_bf16; /* ACT_RELU_GRAD: saved forward activation, bf16 (B,T,N) */
  /* ragged batches (NULL = every 128-row tile of every utterance): a device list of the row tiles to compute,
   * entry = b * ceil(T/128) + tile, and its device length, from ttsb200_build_tile_map.  Rows of tiles that are
   * not listed are neither read nor written (the reference computes its padding rows, Models/layers.py:29-41;
   * nothing valid reads them in a k=1 block with masked keys). */
  const int32_t* tile_map;
  const int32_t* tile_count;
} ttsb200_gemm_args;

int ttsb200_conv_gemm(const ttsb200_gemm_args* args, ttsb200_stream_t stream);

/* The FeedForward of a k=1 FFT block as ONE kernel (Models/modules.py:81-88 + Models/layers.py:40,31):
 *   z = y + LN(relu(x W1^T + b1) W2^T + b2 + x; ln1);  zn = LN(z; ln2)
 * x bf16 (B,T,384) = norm_2(y); w1 bf16 [1536][384]; w2 bf16 [384][1536]; y / z fp32 (B,T,384); zn bf16.
 * The 1536-wide hidden activation stays in tensor memory (TMEM): it is neither written to nor read from HBM.
 * tile_map / tile_count: the ragged row-tile list of the ttsb200_gemm_args struct; NULL = all tiles. */
int ttsb200_ffn_fused(const void* x_bf16, const void* w1_bf16, const float* b1, const void* w2_bf16, const float* b2,
                      const float* y_f32, const float* ln1_g, const float* ln1_b, const float* ln2_g, const float* ln2_b,
                      float eps, float* z_f32, void* zn_bf16, int32_t B, int32_t T, int32_t D, int32_t hidden,
                      const int32_t* tile_map, const int32_t* tile_count, ttsb200_stream_t stream);

/* Row-tile list of a ragged batch for ttsb200_gemm_args.tile_map: utterance b contributes its first
 * min(m_tiles, ceil((len[b*len_stride] + halo) / 128)) tiles; map has room for B*m_tiles entries, *count receives
 * the number written.  len may be kv_info of ttsb200_mask_lengths (len_stride 2).  No host synchronisation. */
int ttsb200_build_tile_map(const int32_t* len, int32_t len_stride, int32_t B, int32_t m_tiles, int32_t halo,
                           int32_t* map, int32_t* count, ttsb200_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Fused masked multi-head self-attention (Models/modules.py:7-21 inside :43-70).
 *   qkv: bf16 (B,T,3*H*Dh) as [q | k | v] along the last axis (the fused q/k/v linears' output)
 *   ctx: bf16 (B,T,H*Dh)  = concat_h softmax(mask(q k^T / sqrt(Dh), -1e4)) v
 * mask (B,T) uint8 marks valid keys; kv_info (B,2) int32 from ttsb200_mask_lengths:
 * kv_info[b][0] = 1 + index of the last valid key (0 = none) lets the kernel skip fully padded key
 * tiles (their probabilities are exactly 0 in fp32); kv_info[b][1] = length of the all-valid prefix,
 * inside which key tiles never read the mask.
 * Dh must be 96 (H*Dh = 384) in this build.
 * ---------------------------------------------------------------------------------------------- */
/* As ttsb200_attention_fwd, but query tiles (128 rows) that start at or beyond kv_info[b][0] — pure padding of a
 * ragged batch — are skipped: their ctx rows are left unwritten (pair with ttsb200_gemm_args.tile_map). */
int ttsb200_attention_fwd_ragged(const void* qkv_bf16, const uint8_t* mask, const int32_t* kv_info, void* ctx_bf16,
                                 int32_t B, int32_t T, int32_t H, int32_t Dh, ttsb200_stream_t stream);
int ttsb200_attention_fwd(const void* qkv, const uint8_t* mask, const int32_t* kv_info, void* ctx,
                          int32_t B, int32_t T, int32_t H, int32_t Dh, ttsb200_stream_t stream);
/* The same under the CAUSAL and key-padding mask of Transformer-TTS (train.py:26-58; attn_1 of Models/layers.py:111):
 * key j attends to query i only for j <= i; hidden keys get the reference's -1e4 fill.  Key tiles above the diagonal are
 * never loaded.  lse2 (optional, (B,H,T) fp32) as in ttsb200_attention_fwd_lse, for ttsb200_attention_bwd_causal. */
int ttsb200_attention_fwd_causal(const void* qkv, const uint8_t* mask, const int32_t* kv_info, void* ctx, float* lse2,
                                 int32_t B, int32_t T, int32_t H, int32_t Dh, ttsb200_stream_t stream);
/* ... with dropout on the probabilities (train.py builds the model with dropout 0.1; Models/modules.py:18-19): the mask of
 * element (b, h, i, j) as in the dropout variant of the plain kernel. */
int ttsb200_attention_fwd_causal_drop(const void* qkv, const uint8_t* mask, const int32_t* kv_info, void* ctx, float* lse2,
                                      int32_t B, int32_t T, int32_t H, int32_t Dh, float drop_p, uint64_t drop_seed,
                                      int64_t drop_site, ttsb200_stream_t stream);
/* Training forward: same kernel, also writes lse2 (B,H,T) fp32 = log2-domain log-sum-exp of every
 * query row, so that p = exp2(s * log2(e)/sqrt(Dh) - lse2) in the backward kernels. */
int ttsb200_attention_fwd_lse(const void* qkv, const uint8_t* mask, const int32_t* kv_info, void* ctx,
                              float* lse2, int32_t B, int32_t T, int32_t H, int32_t Dh,
                              ttsb200_stream_t stream);
/* Debug/parity only: also materialise the (B,H,T,T) fp32 probabilities the reference returns
 * (Models/modules.py:21, Models/encoder.py:97,105).  Plain CUDA-core kernel, not on the timed path. */
int ttsb200_attention_probs(const void* qkv, const uint8_t* mask, float* probs, int32_t B, int32_t T,
                            int32_t H, int32_t Dh, ttsb200_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * HBM-bound row kernels
 * ---------------------------------------------------------------------------------------------- */
/* kv_info (B,2) int32: [b][0] = 1 + last index with mask != 0 (0 if none), [b][1] = index of the first
 * zero (T if none).  (mask contract: train_fastspeech2.py:55-82) */
int ttsb200_mask_lengths(const uint8_t* mask, int32_t* kv_info, int32_t B, int32_t T,
                         ttsb200_stream_t stream);

/* x0 = table[ids] + alpha*pe[t]  (Models/encoder.py:84,94; Models/modules.py:107-111) and
 * xn = LayerNorm(x0; g, b) (Models/layers.py:31).  ids int64 (B,T); table fp32 (V,D); pe fp32 (>=T,D). */
int ttsb200_embed_pe_ln(const int64_t* ids, const float* table, const float* pe, const float* alpha,
                        const float* g, const float* b, float eps, float* x0, void* xn_bf16,
                        int32_t B, int32_t T, int32_t D, int32_t V, ttsb200_stream_t stream);

/* Standalone LayerNorm, fp32 in -> bf16 (and optional fp32) out (Models/encoder.py:112 etc.). */
int ttsb200_layernorm(const float* x, const float* g, const float* b, float eps, void* y_bf16,
                      float* y_f32, int64_t rows, int32_t D, ttsb200_stream_t stream);

/* Tail of an FFT block for short batches (Models/modules.py:84-87, Models/layers.py:40,31): given
 * u = conv2(h) + b2 (fp32, from conv_gemm AFFINE), yn = norm_2(y) (bf16) and the residual stream y (fp32):
 * upre = u + yn; z = y + LayerNorm(upre; ln1); zn = LayerNorm(z; ln2).  Same arithmetic as the
 * TTSB200_EPI_FFN2 epilogue; zn_f32 / pre_f32 (the saved pre-LayerNorm value) are optional. */
int ttsb200_ffn2_tail(const float* u, const void* yn_bf16, const float* y, const float* ln1_g,
                      const float* ln1_b, const float* ln2_g, const float* ln2_b, float eps, float* z,
                      void* zn_bf16, float* zn_f32, float* pre_f32, int64_t rows, int32_t D,
                      ttsb200_stream_t stream);

/* duration = clamp(round_half_even(exp(log_d) - 1), 0)  (Models/varianceadaptor.py:75). int32 out. */
int ttsb200_round_durations(const float* log_d, int32_t* dur, int64_t n, ttsb200_stream_t stream);

/* LengthRegulator (Models/varianceadaptor.py:141-177,233-249).
 *   pass 1 (scan): per utterance inclusive prefix sum of durations -> cum (B,L) int32, mel_len (B).
 *   pass 2 (expand): out[b,t,:] = x[b, idx(b,t), :] for t < mel_len[b], else 0, where idx is the
 *   phoneme whose [cum-d, cum) interval holds t; frame_idx (B,Tmax) int32 (-1 on padding),
 *   mel_pos (B,Tmax) int64 = t+1 on valid frames else 0, mel_mask (B,Tmax) uint8 by the
 *   reference's `arange <= mel_pos` rule (varianceadaptor.py:251-259).  out_f32 (optional, both kernels below): the
 *   same output widened to fp32 — slots 6 / 5 of the reference's return tuple (Models/fastspeech2.py:241) are fp32.
 * dur is int32 (B,L), negative values count as 0. */
int ttsb200_lr_scan(const int32_t* dur, int32_t* cum, int32_t* mel_len, int32_t B, int32_t L,
                    ttsb200_stream_t stream);
int ttsb200_lr_expand(const void* x_bf16, const int32_t* cum, const int32_t* mel_len, void* out_bf16,
                      int32_t* frame_idx, int64_t* mel_pos, uint8_t* mel_mask, float* out_f32, int32_t B, int32_t L,
                      int32_t Tmax, int32_t D, ttsb200_stream_t stream);

/* out = x + pitch_table[bucketize(p, pitch_bins)] + energy_table[bucketize(e, energy_bins)]
 * (Models/varianceadaptor.py:100,109,116-118,123-126; torch.bucketize right=False: first index i
 * with bins[i] >= v).  x/out bf16 (rows,D); tables fp32 (nb+1, D); bins fp32 [nb]; also writes the
 * int32 bucket indices when the pointers are non-NULL. */
int ttsb200_bucket_embed_add(const void* x_bf16, const float* p, const float* e,
                             const float* pitch_bins, const float* energy_bins,
                             const float* pitch_table, const float* energy_table, void* out_bf16,
                             int32_t* p_idx, int32_t* e_idx, float* out_f32, int64_t rows, int32_t D, int32_t nb,
                             ttsb200_stream_t stream);

/* LengthRegulator that allocates nothing.  It has no counterpart in the reference and was added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  = ttsb200_lr_scan followed by ttsb200_lr_expand on the same stream, with the scratch between
 * them (cum (B,L) int32 and mel_len (B) int32) held in a file-scope device array of the library (260 KB on every GPU
 * that loads it, until exit) instead of caller memory.  The caller still passes inputs and outputs; arguments as in
 * ttsb200_lr_expand, plus dur int32 (B,L) and the optional mel_len_out int32 (B).
 * Limits: B <= TTSB200_LR_STATIC_MAX_BATCH and B*L <= TTSB200_LR_STATIC_MAX_PHONEMES; larger sizes return
 * TTSB200 error -1 with a message naming the limit (use the two entry points above with your own scratch).
 * Calls no cudaMalloc / cudaFree / allocating routine.  The first call on each device resolves the array's device
 * address (cudaGetSymbolAddress), so make one call per device before stream capture.  The array is shared by every
 * call on a device and keeps the last call's contents; the scan rewrites every entry the expansion reads, so nothing
 * needs clearing, but calls that OVERLAP on the GPU (different streams without a dependency) race on it. */
#define TTSB200_LR_STATIC_MAX_BATCH 1024
#define TTSB200_LR_STATIC_MAX_PHONEMES 65536
int ttsb200_length_regulate_static(const void* x_bf16, const int32_t* dur, void* out_bf16, int32_t* frame_idx,
                                   int64_t* mel_pos, uint8_t* mel_mask, float* out_f32, int32_t* mel_len_out,
                                   int32_t B, int32_t L, int32_t Tmax, int32_t D, ttsb200_stream_t stream);

/* fp32 -> bf16 cast of a dense buffer (weight/activation staging). */
int ttsb200_cast_bf16(const float* x, void* y_bf16, int64_t n, ttsb200_stream_t stream);

/* ================================================================================================
 * fp32-ACCURATE ENCODER -> DURATION CHAIN (free-running inference).
 * The reference rounds an fp32 log_d to integer durations (Models/varianceadaptor.py:69,75), so the chain that
 * produces it — Models/encoder.py:83-112, Models/layers.py:29-41, Models/modules.py:7-88 and the duration
 * predictor Models/varianceadaptor.py:216-229 — has to reproduce fp32 arithmetic for the durations, frame indices
 * and mel lengths to come out bit-exact.  These entry points evaluate it on the tensor cores with every operand
 * split into two fp16 values ("split" tensors: fp16 (B,T,2C) = [hi | lo] along the last axis, v = hi + lo) and
 * three products per term (hi*hi + lo*hi + hi*lo); partial sums leave the truncating tensor-core accumulator
 * after every 64-wide k-block and are added in fp32 round-to-nearest registers.
 * ============================================================================================== */

/* acc[b,t,n] = sum_j sum_c x[b,t+j-pad_left,c] * w[j,n,c];  v = acc*inv_scale + bias[n] (then ReLU if relu).
 * x_split fp16 (B,T,2C); w_split fp16 [taps][N][2C] = split of w*2^e (inv_scale = 2^-e keeps the lo parts of the
 * weights normal); N % 128 == 0, C % 64 == 0.  Exactly one output: out_f32 (B,T,N) or out_split fp16 (B,T,2N). */
int ttsb200_gemm_x3(const void* x_split, const void* w_split, const float* bias, float inv_scale, int32_t B,
                    int32_t T, int32_t N, int32_t C, int32_t taps, int32_t pad_left, int32_t relu, float* out_f32,
                    void* out_split, ttsb200_stream_t stream);

/* Models/modules.py:7-21 on the split q|k|v (fp16 (B,T,2*3*H*Dh) = [q k v hi | q k v lo], ttsb200_gemm_x3's split
 * output): ctx_split fp16 (B,T,2*H*Dh).  mask / kv_info as in ttsb200_attention_fwd.  Dh must be 96. */
int ttsb200_attention_x3(const void* qkv_split, const uint8_t* mask, const int32_t* kv_info, void* ctx_split,
                         int32_t B, int32_t T, int32_t H, int32_t Dh, ttsb200_stream_t stream);

/* fp32 row operations of the chain (one warp per row, D = 256 or 384):
 *   s = a (+ b0 fp32 | + b0_bf16 | + b0_split hi+lo);  s_out = s (optional);  t = LayerNorm(s; g1, b1)
 *   training with dropout (DROPOUT below): b0 := b0 * m(site_b) before the add, s := s * m(site_s) before s_out and
 *   the LayerNorm, t := t * m(site_t) after it
 *   mode 0 ADD_LN   (Models/layers.py:31,33-35):                 out = t
 *   mode 1 FFN_TAIL (Models/modules.py:85-87, layers.py:40,31):  z = c + t; z_out = z; out = LayerNorm(z; g2, b2)
 *   mode 2 LN_DOT   (Models/varianceadaptor.py:222-229):         out_row = row_mask ? dot(t, dot_w) + dot_b : 0
 * out goes to any of out_split (fp16 (rows,2D)), out_bf16, out_f32. */
int ttsb200_xrow(int32_t mode, const float* a, const float* b0, const void* b0_bf16, const void* b0_split,
                 const float* c, const float* g1, const float* b1, const float* g2, const float* b2, const float* dot_w,
                 const float* dot_b, const uint8_t* row_mask, float eps, float* s_out, float* z_out, void* out_split,
                 void* out_bf16, float* out_f32, float* out_row, int64_t rows, int32_t D, float drop_p,
                 uint64_t drop_seed, int64_t site_b, int64_t site_s, int64_t site_t, ttsb200_stream_t stream);

/* out_split fp16 (rows,2D) = split of x*scale, x fp32 (rows,D) (weight / activation staging). */
int ttsb200_split_f16(const float* x, void* out_split, int64_t rows, int32_t D, float scale, ttsb200_stream_t stream);

/* ================================================================================================
 * TRAINING STEP (forward extras above: pre_f32, gate_bf16, ttsb200_attention_fwd_lse).
 * The reference has no explicit backward code: it calls loss.backward() / optimizer.step()
 * (train_fastspeech2.py:304-315) and PyTorch autograd differentiates the modules cited below.
 * Input-gradient ("dgrad") GEMMs reuse ttsb200_conv_gemm with transposed, tap-flipped weights.
 * ============================================================================================== */

/* Weight gradient of nn.Linear / nn.Conv1d (Models/modules.py:49-51,68,83-84; varianceadaptor.py:217,220;
 * postnets.py:67,71-74):  dW[j,n,c] += sum_{b,t} dy[b,t,n] * x[b,t+j-pad_left,c].
 * dy bf16 (B,T,ld_dy) (N columns used), x bf16 (B,T,ld_x) (C columns used); dW fp32, element (j,n,c) at
 * dw[j*stride_tap + n*stride_n + c*stride_c], ACCUMULATED with atomics (zero it first). tcgen05, both
 * operands MN-major, split-K over time. */
int ttsb200_conv_wgrad(const void* dy, int64_t ld_dy, const void* x, int64_t ld_x, float* dw,
                       int64_t stride_tap, int64_t stride_n, int64_t stride_c, int32_t B, int32_t T,
                       int32_t N, int32_t C, int32_t taps, int32_t pad_left, ttsb200_stream_t stream);
/* The same plus the bias gradient in the same launch: db[n] += sum_{b,t} dy[b,t,n] (fp32 (N), accumulated; NULL = none).  The
 * tiles of x-channel block 0 multiply dy by a constant operand of ones, so dy is not read a second time by a column-sum pass. */
int ttsb200_conv_wgrad_bias(const void* dy, int64_t ld_dy, const void* x, int64_t ld_x, float* dw, float* db,
                            int64_t stride_tap, int64_t stride_n, int64_t stride_c, int32_t B, int32_t T,
                            int32_t N, int32_t C, int32_t taps, int32_t pad_left, ttsb200_stream_t stream);

/* Backward of ttsb200_attention_fwd (autograd of Models/modules.py:7-21): dqkv bf16 (B,T,3*H*Dh) =
 * [dq | dk | dv] from qkv, dctx bf16 (B,T,H*Dh), the saved lse2 (B,H,T) and delta (B,H,T) =
 * rowsum(dctx * ctx) per head (ttsb200_attn_delta). */
int ttsb200_attention_bwd(const void* qkv, const void* dctx, const uint8_t* mask, const int32_t* kv_info,
                          const float* lse2, const float* delta, void* dqkv, int32_t B, int32_t T,
                          int32_t H, int32_t Dh, ttsb200_stream_t stream);
int ttsb200_attn_delta(const void* dctx, const void* ctx, float* delta, int32_t B, int32_t T, int32_t H,
                       int32_t Dh, ttsb200_stream_t stream);
/* Backward of ttsb200_attention_fwd_causal (causal AND key-padding mask): a key tile starts at the query tile that holds
 * its diagonal, a query tile stops at it; only the diagonal tiles evaluate the per-row predicate. */
int ttsb200_attention_bwd_causal(const void* qkv, const void* dctx, const uint8_t* mask, const int32_t* kv_info,
                                 const float* lse2, const float* delta, void* dqkv, int32_t B, int32_t T, int32_t H,
                                 int32_t Dh, ttsb200_stream_t stream);
int ttsb200_attention_bwd_causal_drop(const void* qkv, const void* dctx, const uint8_t* mask, const int32_t* kv_info,
                                      const float* lse2, const float* delta, void* dqkv, int32_t B, int32_t T, int32_t H,
                                      int32_t Dh, float drop_p, uint64_t drop_seed, int64_t drop_site,
                                      ttsb200_stream_t stream);

/* LayerNorm backward (Models/layers.py:31,35; modules.py:87; varianceadaptor.py:219,222; encoder.py:112).
 * x = saved pre-LN input (rows,D) fp32, statistics are recomputed.  Exactly one gradient source:
 * g_f32 (rows,D) | g_bf16 (rows,D) | g_row[rows] * g_vec[D] (the predictor's 256->1 linear head,
 * varianceadaptor.py:223-229; rows with row_mask == 0 contribute 0).  dx (+ add_f32) goes to dx_f32 and/or
 * dx_bf16; relu_gate zeroes dx where x <= 0 (x is then a ReLU output).  dgamma/dbeta/dvec (D) and dscal (1)
 * are ACCUMULATED (dvec = sum g_row * LN(x), dscal = sum g_row: the linear head's weight/bias grads). */
int ttsb200_layernorm_bwd(const float* x, const float* gamma, const float* beta, const float* g_f32,
                          const void* g_bf16, const float* g_row, const float* g_vec,
                          const uint8_t* row_mask, const float* add_f32, int32_t relu_gate, float* dx_f32,
                          void* dx_bf16, float* dgamma, float* dbeta, float* dvec, float* dscal,
                          int64_t rows, int32_t D, float eps, ttsb200_stream_t stream);
/* The same with the dropout masks of the forward pass regenerated (see DROPOUT below): site_g >= 0: the LayerNorm
 * OUTPUT was dropped (g := g * m, the head's dvec uses y * m); site_dx >= 0 with dx_mode 1: the LayerNorm INPUT was a
 * dropout's output (dx := dx * m on both outputs), dx_mode 2: only dx_bf16 is masked (the copy that feeds the GEMM
 * gradients of a dropped branch, Models/layers.py:33; dx_f32 continues down the residual path). */
int ttsb200_layernorm_bwd_drop(const float* x, const float* gamma, const float* beta, const float* g_f32,
                               const void* g_bf16, const float* g_row, const float* g_vec, const uint8_t* row_mask,
                               const float* add_f32, int32_t relu_gate, float* dx_f32, void* dx_bf16, float* dgamma,
                               float* dbeta, float* dvec, float* dscal, int64_t rows, int32_t D, float eps,
                               float drop_p, uint64_t drop_seed, int64_t site_g, int64_t site_dx, int32_t dx_mode,
                               ttsb200_stream_t stream);

/* out[c] += sum_rows x[row*ld + c]  (bias gradients); x bf16 or fp32. */
int ttsb200_colsum(const void* x, int32_t is_bf16, int64_t ld, float* out, int64_t rows, int32_t cols,
                   ttsb200_stream_t stream);

/* BatchNorm1d (training mode) + tanh of the PostNet (Models/postnets.py:71-73).  x fp32 (rows,384) is the
 * conv output; stats (2,384) accumulates [sum x | sum x^2]; mean_rstd (2,384) = [mean | rstd] is derived by
 * the caller (after the cross-rank all-reduce of SyncBatchNorm). */
int ttsb200_bn_stats(const float* x, float* stats, int64_t rows, int32_t C, ttsb200_stream_t stream);
int ttsb200_bn_tanh_fwd(const float* x, const float* mean_rstd, const float* gamma, const float* beta,
                        void* h_bf16, int64_t rows, int32_t C, float drop_p, uint64_t drop_seed, int64_t drop_site,
                        ttsb200_stream_t stream);
/* apply == 0: red (2,384) += [sum dt | sum dt*xhat] with dt = dh*(1-h^2)  (= dbeta | dgamma);
 * apply == 1: dx bf16 = gamma*rstd*(dt - red[0]*inv_count - xhat*red[1]*inv_count). */
int ttsb200_bn_tanh_bwd(const void* dh_bf16, const void* h_bf16, const float* x, const float* mean_rstd,
                        const float* gamma, float* red, void* dx_bf16, int64_t rows, int32_t C,
                        float inv_count, int32_t apply, float drop_p, uint64_t drop_seed, int64_t drop_site,
                        ttsb200_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * DROPOUT (training; the reference constructs the model with dropout 0.1 / 0.5, train_fastspeech2.py:381-389).
 * Masks are counter-based: keep(i) = hash32(i, key(seed, site)) >= p * 2^32, kept elements scaled by 1 / (1 - p) as
 * F.dropout does; `site` names one dropout call site of one module instance, i is the element's offset in the dropped
 * tensor.  Forward kernels apply the mask, backward kernels regenerate it from the same (p, seed, site): no mask
 * tensor exists.  drop_p == 0 or site < 0 disables a dropout point.  Call sites in the reference:
 * Models/modules.py:18-19 (attention probabilities), :86 (FeedForward), :111 (PositionalEncoder),
 * Models/layers.py:33,40 (dropout_1 / dropout_2), Models/varianceadaptor.py:219,222, Models/postnets.py:71-73.
 * ---------------------------------------------------------------------------------------------- */
/* dst[i] = src[i] * m(i); is_bf16 selects the element type; dst may alias src. */
int ttsb200_dropout(const void* src, void* dst, int32_t is_bf16, int64_t n, float p, uint64_t seed, int64_t site,
                    ttsb200_stream_t stream);
/* out[i] = m(i) (0 or 1/(1-p)): the mask itself, for parity tests. */
int ttsb200_dropout_mask(float* out, int64_t n, float p, uint64_t seed, int64_t site, ttsb200_stream_t stream);
/* ttsb200_attention_fwd_lse / ttsb200_attention_bwd with dropout on the softmax probabilities (Models/modules.py:
 * 18-19): ctx = (P * m) V; element index of m = ((b*H + h)*T + query)*T + key. */
int ttsb200_attention_fwd_drop(const void* qkv, const uint8_t* mask, const int32_t* kv_info, void* ctx, float* lse2,
                               int32_t B, int32_t T, int32_t H, int32_t Dh, float drop_p, uint64_t drop_seed,
                               int64_t drop_site, ttsb200_stream_t stream);
int ttsb200_attention_bwd_drop(const void* qkv, const void* dctx, const uint8_t* mask, const int32_t* kv_info,
                               const float* lse2, const float* delta, void* dqkv, int32_t B, int32_t T, int32_t H,
                               int32_t Dh, float drop_p, uint64_t drop_seed, int64_t drop_site,
                               ttsb200_stream_t stream);

/* LengthRegulator backward (Models/varianceadaptor.py:167-177): dx[b,l,:] = add[b,l,:] + sum of g[b,t,:]
 * over the frames of phoneme l (cum from ttsb200_lr_scan, frames beyond Tmax were cropped). fp32. */
int ttsb200_lr_bwd(const float* g, const int32_t* cum, float* dx, const float* add_f32, int32_t B, int32_t L,
                   int32_t Tmax, int32_t D, ttsb200_stream_t stream);

/* Embedding weight gradient, table[idx[row],:] += g[row,:] (skip_idx = padding_idx or -1; encoder.py:55,
 * varianceadaptor.py:123-126), and/or dalpha += sum_rows g[row,:].pe[row % T,:] (modules.py:110). */
int ttsb200_embed_bwd(const float* g, const int64_t* idx64, const int32_t* idx32, float* table,
                      int32_t skip_idx, const float* pe, float* dalpha, int64_t rows, int32_t T, int32_t D,
                      int32_t V, ttsb200_stream_t stream);

/* nn.L1Loss (mean) and its gradient (train_fastspeech2.py:212-217,245-259): loss += scale*sum|pred-target|,
 * grad = scale*sign(pred-target); log_target: target := log(target+1) (duration loss). */
int ttsb200_l1_loss(const float* pred, const float* target, float* grad, void* grad_bf16, float* loss,
                    int64_t n, float scale, int32_t log_target, ttsb200_stream_t stream);

/* F.binary_cross_entropy_with_logits(x, y, reduction='mean', pos_weight) (train.py:217, the stop-token loss of
 * Transformer-TTS): loss += scale * sum[(1-y) x + (1 + (pos_weight-1) y) softplus(-x)], grad (optional) = its derivative. */
int ttsb200_bce_logits_loss(const float* x, const float* y, float* grad, float* loss, int64_t n, float scale,
                            float pos_weight, ttsb200_stream_t stream);
/* out_bf16[i] = gate_bf16[i] > 0 ? g[i] : 0: backward of a ReLU whose (bf16) output is `gate` (Models/prenets.py:36-37). */
int ttsb200_relu_gate(const float* g, const void* gate_bf16, void* out_bf16, int64_t n, ttsb200_stream_t stream);

/* Input-gradient operands of a whole backward pass in one launch.  The reference gets input gradients from autograd
 * (loss.backward(), train_fastspeech2.py:304 / train.py:252); here they are ttsb200_conv_gemm calls on the transposed,
 * tap-flipped weights, and this entry builds all of them: for every entry dst[taps-1-j][c][n] = src[j][n][c] (bf16, device
 * pointers; src = the forward operand (taps, N, C)).  tile0 = index of the entry's first 64 x 64 tile: entries in ascending
 * tile0, taps * ceil(N/64) * ceil(C/64) tiles each, total_tiles their sum.  The table itself is in device memory. */
typedef struct {
  const void* src;
  void* dst;
  int32_t taps, N, C, tile0;
} ttsb200_wt_entry;
int ttsb200_weight_transpose(const ttsb200_wt_entry* table_dev, int32_t n_entries, int32_t total_tiles,
                             ttsb200_stream_t stream);

/* Optimizer over one flat fp32 buffer (train_fastspeech2.py:117-120,304-315): out += sum g^2; then
 * g' = g*gscale*min(1, clip/(sqrt(sumsq)*gscale + 1e-6)) (clip_grad_norm_; clip <= 0 disables) and one
 * torch.optim.Adam step (step >= 1 for the bias corrections). */
int ttsb200_sumsq(const float* g, float* out, int64_t n, ttsb200_stream_t stream);
int ttsb200_adam_step(float* p, const float* g, float* m, float* v, const float* sumsq, int64_t n,
                      float gscale, float clip, float lr, float beta1, float beta2, float eps, int32_t step,
                      ttsb200_stream_t stream);

/* radam.RAdam.step (radam.py:30-92, the optimizer train_fastspeech2.py:412-414 builds for hp.optimizer ==
 * 'radam': lr fixed at max_lr, betas (0.9, 0.98), eps 1e-9, weight_decay 0, degenerated_to_sgd=True) over the
 * same flat buffers, with the same gradient scaling / clipping prologue as ttsb200_adam_step.  The rectification
 * term N_sma and the step size are computed on the host in double, as the reference does in Python:
 * N_sma >= 5 -> p -= lr*step_size*m/(sqrt(v)+eps), else p -= lr*m/(1-beta1^t). */
int ttsb200_radam_step(float* p, const float* g, float* m, float* v, const float* sumsq, int64_t n,
                       float gscale, float clip, float lr, float beta1, float beta2, float eps, int32_t step,
                       ttsb200_stream_t stream);

/* ================================================================================================
 * AUTOREGRESSIVE DECODE STEP of Transformer-TTS (BASELINE.json configs[4]).
 * The reference decodes by re-running the whole prefix through the model at every step (test.py:128-191); these entry
 * points process ONE new decoder position per utterance against cached keys / values (Models/decoder.py:43-56,
 * Models/layers.py:109-124, Models/modules.py:7-21,43-68).  R = rows = utterances in the batch.  Position-dependent
 * addresses read the device scalar *pos, so a captured CUDA graph of one step replays for every step.
 * ============================================================================================== */

/* y[r,n] = act( sum_k xin[r,k] * w[n,k] + bias[n] ) (+ res[r,n]) (+ add_scale[0] * add_row[pos*N + n])
 *   xin = x_bf16 rows, or x_f32 rows, optionally LayerNorm(x_f32 row; ln_g, ln_b, eps) (nn.LayerNorm, Models/layers.py:110,
 *   114,118; Models/decoder.py:56), rounded to bf16 (the operand precision of every GEMM of this library);
 *   input row r starts at x + r*ldx + (pos + x_pos_bias)*x_pos_stride (x_pos_stride != 0; a negative step index reads
 *   zeros: the all-zero first decoder input of test.py:139);
 *   w: bf16 [N][ldw], rows zero-padded to a multiple of 32 columns (ldw >= roundup(K,32), 16-byte aligned rows);
 *   act = ReLU if relu (Models/prenets.py:32-37, Models/modules.py:83); add_row = positional-encoding table (>= pos+1, N)
 *   with add_scale = alpha (Models/modules.py:107-111);
 *   out_f32[r*ldo + pos*out_pos_stride + n] and / or out_bf16[r*ldob + n];
 *   xn_bf16 (optional, (R,K) dense) receives the bf16 input rows after the LayerNorm (the FeedForward's inner residual,
 *   Models/modules.py:81,85). */
typedef struct {
  const float* x_f32;
  const void* x_bf16;
  int64_t ldx;
  const float* ln_g;
  const float* ln_b;
  float eps;
  const void* w_bf16;
  int64_t ldw;
  const float* bias;
  int32_t relu;
  const float* res_f32;
  int64_t ldres;
  const float* add_row;
  const float* add_scale;
  float* out_f32;
  int64_t ldo;
  void* out_bf16;
  int64_t ldob;
  void* xn_bf16;
  const int32_t* pos;
  int64_t x_pos_stride;
  int32_t x_pos_bias;
  int64_t out_pos_stride;
  int32_t R, K, N;
} ttsb200_small_linear_args;
int ttsb200_small_linear(const ttsb200_small_linear_args* args, ttsb200_stream_t stream);

/* One query row per (utterance, head) against cached keys / values (Models/modules.py:7-21 for the newest position):
 *   ctx[r, h*Dh + d] = sum_j softmax_j(q[r,h,:] . K[r,h,j,:] / sqrt(Dh)) V[r,h,j,d]
 * q bf16, head h at columns [h*Dh, (h+1)*Dh) of row r (row stride ldq).  Key j of (r,h) is the Dh bf16 values at
 * k_cache + r*kv_stride_b + h*kv_stride_h + j*kv_stride_j (v_cache likewise; strides in elements): [R][H][cap][Dh] caches
 * and the (R, L, 2*H*Dh) output of a k|v projection are both expressible.  cap = number of key slots.
 * Self-attention (attn_1, Models/layers.py:111): pos given; knew / vnew (row stride ldkv, head h at column h*Dh) are
 * written to slot *pos first, then keys 0..*pos attend (the causal mask of train.py:26-58 at the last row), clamped to
 * kv_len[r] keys when kv_len is given too (the padding half of that mask, for teacher-forced prefixes).
 * Cross-attention (attn_2, Models/layers.py:115): kv_len only; keys 0..kv_len[r]-1 attend (the reference fills padded
 * keys with -1e4, whose probabilities are exactly 0 in fp32 next to any valid key).  Dh must be 96. */
int ttsb200_attn_decode(const void* q_bf16, int64_t ldq, const void* knew_bf16, const void* vnew_bf16, int64_t ldkv,
                        void* k_cache, void* v_cache, int64_t kv_stride_b, int64_t kv_stride_h, int64_t kv_stride_j,
                        const int32_t* pos, const int32_t* kv_len, void* ctx_bf16, int64_t ldc, int32_t R, int32_t H,
                        int32_t Dh, int32_t cap, ttsb200_stream_t stream);

/* FeedForward tail of a decoder block for a few rows (Models/modules.py:85-87, Models/layers.py:122) and the NEXT
 * LayerNorm (norm_1 of the following block, Models/layers.py:110, or Models/decoder.py:56):
 *   z = y + LayerNorm(u + yn; ln1);  zn = bf16(LayerNorm(z; ln2))
 * u = conv2 output + bias (fp32), yn = the FeedForward's input rows (bf16), y = the residual stream (fp32); (R, D) dense,
 * D = 384.  The arithmetic of ttsb200_ffn2_tail, as a member of the step's programmatic-dependent-launch chain. */
int ttsb200_ffn_tail_rows(const float* u, const void* yn_bf16, const float* y, const float* ln1_g, const float* ln1_b,
                          const float* ln2_g, const float* ln2_b, float eps, float* z, void* zn_bf16, int32_t R,
                          int32_t D, ttsb200_stream_t stream);

/* Fused sub-blocks of a decoder block for up to 16 rows per thread-block CLUSTER of 16 CTAs (csrc/decode_cluster.cu):
 * every CTA owns 1/16 of each linear's output columns, keeps its weight slices in shared memory and pushes its result
 * slices into the shared memory of all CTAs of the cluster; cluster barriers stand where kernel boundaries were.
 * Weights bf16 row-major as for ttsb200_small_linear (wo, wq [384][384]; w1 [1536][384]; w2 [384][1536]); D = 384.
 *   attn_out_q: x_out = x_in + ctx Wo^T + bo  (Models/layers.py:112);  q = bf16(LN(x_out; ln) Wq^T + bq)  (layers.py:114-115)
 *   ffn_block:  x' = x_in + ctx Wo^T + bo (layers.py:116; kept on chip);  h = relu(LN3(x') W1^T + b1);  u = h W2^T + b2;
 *               z = x' + LN_ff(u + LN3(x'))  (layers.py:118-122, Models/modules.py:81-87);  zn = bf16(LN_next(z)) */
int ttsb200_decode_attn_out_q(const void* ctx_bf16, const float* x_in, const void* wo_bf16, const float* bo, const float* ln_g,
                              const float* ln_b, const void* wq_bf16, const float* bq, float eps, float* x_out, void* q_bf16,
                              int32_t R, int32_t D, ttsb200_stream_t stream);
int ttsb200_decode_ffn_block(const void* ctx_bf16, const float* x_in, const void* wo_bf16, const float* bo, const float* ln3_g,
                             const float* ln3_b, const void* w1_bf16, const float* b1, const void* w2_bf16, const float* b2,
                             const float* lnff_g, const float* lnff_b, const float* lnn_g, const float* lnn_b, float eps,
                             float* z_out, void* zn_bf16, int32_t R, int32_t D, int32_t hidden, ttsb200_stream_t stream);

/* Experiments only: phase timeline (clock64 of one thread) of the last ttsb200_decode_ffn_block; mode 1 on, 0 off, 2 read. */
int ttsb200_debug_dcl_trace(long long* dst, int mode);

/* *pos += delta (the last node of a captured decode step). */
int ttsb200_advance_pos(int32_t* pos, int32_t delta, ttsb200_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* TTSB200_H_ */
