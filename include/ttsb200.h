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

int ttsb200_version(void);
const char* ttsb200_last_error_string(void);

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
enum { TTSB200_ACT_NONE = 0, TTSB200_ACT_RELU = 1, TTSB200_ACT_TANH = 2 };

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
} ttsb200_gemm_args;

int ttsb200_conv_gemm(const ttsb200_gemm_args* args, ttsb200_stream_t stream);

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
int ttsb200_attention_fwd(const void* qkv, const uint8_t* mask, const int32_t* kv_info, void* ctx,
                          int32_t B, int32_t T, int32_t H, int32_t Dh, ttsb200_stream_t stream);
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

/* duration = clamp(round_half_even(exp(log_d) - 1), 0)  (Models/varianceadaptor.py:75). int32 out. */
int ttsb200_round_durations(const float* log_d, int32_t* dur, int64_t n, ttsb200_stream_t stream);

/* LengthRegulator (Models/varianceadaptor.py:141-177,233-249).
 *   pass 1 (scan): per utterance inclusive prefix sum of durations -> cum (B,L) int32, mel_len (B).
 *   pass 2 (expand): out[b,t,:] = x[b, idx(b,t), :] for t < mel_len[b], else 0, where idx is the
 *   phoneme whose [cum-d, cum) interval holds t; frame_idx (B,Tmax) int32 (-1 on padding),
 *   mel_pos (B,Tmax) int64 = t+1 on valid frames else 0, mel_mask (B,Tmax) uint8 by the
 *   reference's `arange <= mel_pos` rule (varianceadaptor.py:251-259).
 * dur is int32 (B,L), negative values count as 0. */
int ttsb200_lr_scan(const int32_t* dur, int32_t* cum, int32_t* mel_len, int32_t B, int32_t L,
                    ttsb200_stream_t stream);
int ttsb200_lr_expand(const void* x_bf16, const int32_t* cum, const int32_t* mel_len, void* out_bf16,
                      int32_t* frame_idx, int64_t* mel_pos, uint8_t* mel_mask, int32_t B, int32_t L,
                      int32_t Tmax, int32_t D, ttsb200_stream_t stream);

/* out = x + pitch_table[bucketize(p, pitch_bins)] + energy_table[bucketize(e, energy_bins)]
 * (Models/varianceadaptor.py:100,109,116-118,123-126; torch.bucketize right=False: first index i
 * with bins[i] >= v).  x/out bf16 (rows,D); tables fp32 (nb+1, D); bins fp32 [nb]; also writes the
 * int32 bucket indices when the pointers are non-NULL. */
int ttsb200_bucket_embed_add(const void* x_bf16, const float* p, const float* e,
                             const float* pitch_bins, const float* energy_bins,
                             const float* pitch_table, const float* energy_table, void* out_bf16,
                             int32_t* p_idx, int32_t* e_idx, int64_t rows, int32_t D, int32_t nb,
                             ttsb200_stream_t stream);

/* fp32 -> bf16 cast of a dense buffer (weight/activation staging). */
int ttsb200_cast_bf16(const float* x, void* y_bf16, int64_t n, ttsb200_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* TTSB200_H_ */
