// HBM-bound row kernels of the FastSpeech2 path: embedding + positional encoding + LayerNorm,
// standalone LayerNorm, duration rounding, LengthRegulator (warp-scan prefix sum + vectorised
// frame gather), bucketize + embedding add, mask -> length, fp32 -> bf16 cast.
// All of them are sized so every warp moves 16-byte (or 8-byte for bf16x4) vectors, fully coalesced.
#include <mutex>

#include "common.cuh"
#include "../../include/ttsb200.h"

namespace ttsb {

// ------------------------------------------------------------------------------------------------
// LayerNorm helpers: one warp per row, NV float4 per lane (D = 128*NV)
// ------------------------------------------------------------------------------------------------
template <int NV>
__device__ __forceinline__ void warp_ln(float4 (&v)[NV], const float* __restrict__ g,
                                        const float* __restrict__ b, float eps, int lane) {
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) s += v[i].x + v[i].y + v[i].z + v[i].w;
  const float mean = warp_sum(s) * (1.f / (128.f * NV));
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const float a = v[i].x - mean, c = v[i].y - mean, d = v[i].z - mean, e = v[i].w - mean;
    q += a * a + c * c + d * d + e * e;
  }
  const float rstd = rsqrtf(warp_sum(q) * (1.f / (128.f * NV)) + eps);
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const float4 gg = __ldg(reinterpret_cast<const float4*>(g) + i * 32 + lane);
    const float4 bb = __ldg(reinterpret_cast<const float4*>(b) + i * 32 + lane);
    v[i].x = fmaf((v[i].x - mean) * rstd, gg.x, bb.x);
    v[i].y = fmaf((v[i].y - mean) * rstd, gg.y, bb.y);
    v[i].z = fmaf((v[i].z - mean) * rstd, gg.z, bb.z);
    v[i].w = fmaf((v[i].w - mean) * rstd, gg.w, bb.w);
  }
}

__device__ __forceinline__ void store_bf16x4(__nv_bfloat16* dst, const float4& v) {
  uint2 u;
  u.x = pack_bf16(v.x, v.y);
  u.y = pack_bf16(v.z, v.w);
  *reinterpret_cast<uint2*>(dst) = u;
}

template <int NV>
__global__ void __launch_bounds__(256)
embed_pe_ln_kernel(const int64_t* __restrict__ ids, const float* __restrict__ table,
                   const float* __restrict__ pe, const float* __restrict__ alpha,
                   const float* __restrict__ g, const float* __restrict__ b, float eps,
                   float* __restrict__ x0, __nv_bfloat16* __restrict__ xn, int64_t rows, int T, int V) {
  constexpr int D = 128 * NV;
  const int lane = threadIdx.x & 31;
  const int64_t row = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int t = static_cast<int>(row % T);
  int64_t id = ids[row];
  if (id < 0 || id >= V) id = 0;  // the reference would raise; ids are validated on the host too
  const float a = __ldg(alpha);
  float4 v[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const float4 e = __ldg(reinterpret_cast<const float4*>(table + id * D) + i * 32 + lane);
    const float4 p = __ldg(reinterpret_cast<const float4*>(pe + static_cast<int64_t>(t) * D) + i * 32 + lane);
    // separate multiply and add (no FMA contraction): bit-identical to `x + alpha * pe` (modules.py:110)
    v[i] = make_float4(__fadd_rn(e.x, __fmul_rn(a, p.x)), __fadd_rn(e.y, __fmul_rn(a, p.y)),
                       __fadd_rn(e.z, __fmul_rn(a, p.z)), __fadd_rn(e.w, __fmul_rn(a, p.w)));
    reinterpret_cast<float4*>(x0 + row * D)[i * 32 + lane] = v[i];
  }
  warp_ln<NV>(v, g, b, eps, lane);
#pragma unroll
  for (int i = 0; i < NV; ++i) store_bf16x4(xn + row * D + (i * 32 + lane) * 4, v[i]);
}

template <int NV>
__global__ void __launch_bounds__(256)
layernorm_kernel(const float* __restrict__ x, const float* __restrict__ g, const float* __restrict__ b,
                 float eps, __nv_bfloat16* __restrict__ y_bf16, float* __restrict__ y_f32, int64_t rows) {
  constexpr int D = 128 * NV;
  const int lane = threadIdx.x & 31;
  const int64_t row = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  float4 v[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) v[i] = __ldcs(reinterpret_cast<const float4*>(x + row * D) + i * 32 + lane);
  warp_ln<NV>(v, g, b, eps, lane);
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    if (y_bf16) store_bf16x4(y_bf16 + row * D + (i * 32 + lane) * 4, v[i]);
    if (y_f32) reinterpret_cast<float4*>(y_f32 + row * D)[i * 32 + lane] = v[i];
  }
}

// ------------------------------------------------------------------------------------------------
// FFN tail for SHORT batches (modules.py:84-87 + layers.py:40 + the next norm_1): when the batch holds
// fewer 128-row tiles than half the SMs, the whole-row LayerNorm GEMM epilogue (one CTA per tile)
// leaves most of the GPU idle.  conv2 then runs as a 192-wide AFFINE GEMM over twice the CTAs and
// this row kernel finishes the block:
//   upre = u + yn ; z = y + LN_ff(upre) ; zn = LN_next(z)
// ------------------------------------------------------------------------------------------------
template <int NV>
__global__ void __launch_bounds__(256)
ffn2_tail_kernel(const float* __restrict__ u, const __nv_bfloat16* __restrict__ yn, const float* __restrict__ y,
                 const float* __restrict__ g1, const float* __restrict__ b1, const float* __restrict__ g2,
                 const float* __restrict__ b2, float eps, float* __restrict__ z, __nv_bfloat16* __restrict__ zn,
                 float* __restrict__ zn32, float* __restrict__ pre, int64_t rows) {
  constexpr int D = 128 * NV;
  const int lane = threadIdx.x & 31;
  const int64_t row = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  float4 v[NV], r[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    v[i] = __ldcs(reinterpret_cast<const float4*>(u + row * D) + i * 32 + lane);
    const uint2 h = __ldg(reinterpret_cast<const uint2*>(yn + row * D) + i * 32 + lane);
    const float2 lo = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&h.x));
    const float2 hi = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&h.y));
    v[i].x += lo.x; v[i].y += lo.y; v[i].z += hi.x; v[i].w += hi.y;
    r[i] = __ldcs(reinterpret_cast<const float4*>(y + row * D) + i * 32 + lane);
    if (pre) reinterpret_cast<float4*>(pre + row * D)[i * 32 + lane] = v[i];
  }
  warp_ln<NV>(v, g1, b1, eps, lane);
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    v[i].x += r[i].x; v[i].y += r[i].y; v[i].z += r[i].z; v[i].w += r[i].w;
    reinterpret_cast<float4*>(z + row * D)[i * 32 + lane] = v[i];
  }
  warp_ln<NV>(v, g2, b2, eps, lane);
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    store_bf16x4(zn + row * D + (i * 32 + lane) * 4, v[i]);
    if (zn32) reinterpret_cast<float4*>(zn32 + row * D)[i * 32 + lane] = v[i];
  }
}

// ------------------------------------------------------------------------------------------------
// duration rounding: clamp(round_half_even(exp(x) - 1), 0)
// exp is evaluated in double and rounded once to fp32 (a correctly rounded fp32 exp), then the
// reference's fp32 subtract and round-half-even (torch.round) follow.
// ------------------------------------------------------------------------------------------------
__global__ void round_durations_kernel(const float* __restrict__ log_d, int32_t* __restrict__ dur, int64_t n) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float e = static_cast<float>(exp(static_cast<double>(log_d[i])));
  const float d = fmaxf(rintf(e - 1.0f), 0.f);
  dur[i] = d > 2147483520.f ? 2147483647 : static_cast<int32_t>(d);
}

// ------------------------------------------------------------------------------------------------
// LengthRegulator
// ------------------------------------------------------------------------------------------------
// One warp per utterance: inclusive prefix sum of max(dur,0) with a shuffle scan, 32 phonemes a step.
__global__ void lr_scan_kernel(const int32_t* __restrict__ dur, int32_t* __restrict__ cum,
                               int32_t* __restrict__ mel_len, int B, int L) {
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (b >= B) return;
  int carry = 0;
  for (int base = 0; base < L; base += 32) {
    const int i = base + lane;
    int v = (i < L) ? max(dur[static_cast<int64_t>(b) * L + i], 0) : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += u;
    }
    v += carry;
    if (i < L) cum[static_cast<int64_t>(b) * L + i] = v;
    carry = __shfl_sync(0xffffffffu, v, 31);
  }
  if (lane == 0) mel_len[b] = carry;
}

constexpr int kLrFrames = 64;  // frames per CTA
__global__ void __launch_bounds__(256)
lr_expand_kernel(const __nv_bfloat16* __restrict__ x, const int32_t* __restrict__ cum,
                 const int32_t* __restrict__ mel_len, __nv_bfloat16* __restrict__ out,
                 int32_t* __restrict__ frame_idx, int64_t* __restrict__ mel_pos,
                 uint8_t* __restrict__ mel_mask, float* __restrict__ out32, int L, int Tmax, int D) {
  __shared__ int s_src[kLrFrames];
  const int b = blockIdx.y;
  const int t0 = blockIdx.x * kLrFrames;
  const int len = min(mel_len[b], Tmax);
  if (threadIdx.x < kLrFrames) {
    const int t = t0 + threadIdx.x;
    int src = -1;
    if (t < len) {
      // first phoneme whose inclusive cumulative duration exceeds t
      const int32_t* c = cum + static_cast<int64_t>(b) * L;
      int lo = 0, hi = L - 1;
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(c + mid) > t) hi = mid; else lo = mid + 1;
      }
      src = lo;
    }
    s_src[threadIdx.x] = src;
    if (t < Tmax) {
      const int64_t o = static_cast<int64_t>(b) * Tmax + t;
      if (frame_idx) frame_idx[o] = src;
      if (mel_pos) mel_pos[o] = (t < len) ? t + 1 : 0;
      if (mel_mask) mel_mask[o] = (t < len || t == 0) ? 1 : 0;  // `arange <= mel_pos` (varianceadaptor.py:257)
    }
  }
  __syncthreads();
  const int chunks = D >> 3;  // 16-byte chunks per row
  const int total = kLrFrames * chunks;
  const uint4* xb = reinterpret_cast<const uint4*>(x + static_cast<int64_t>(b) * L * D);
  uint4* ob = reinterpret_cast<uint4*>(out + (static_cast<int64_t>(b) * Tmax + t0) * D);
  // four independent 16-byte gathers in flight per thread before the first store
  constexpr int U = 4;
  for (int c0 = threadIdx.x; c0 < total; c0 += U * 256) {
    uint4 v[U];
    bool ok[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int c = c0 + u * 256;
      const int f = c / chunks;
      ok[u] = c < total && t0 + f < Tmax;
      v[u] = make_uint4(0, 0, 0, 0);
      if (ok[u]) {
        const int src = s_src[f];
        if (src >= 0) v[u] = __ldg(xb + static_cast<int64_t>(src) * chunks + (c - f * chunks));
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (ok[u]) __stcs(ob + c0 + u * 256, v[u]);
    if (out32) {  // the same rows in fp32 (slot 6 of the reference's return tuple, Models/fastspeech2.py:241)
      float4* o32 = reinterpret_cast<float4*>(out32 + (static_cast<int64_t>(b) * Tmax + t0) * D);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!ok[u]) continue;
        const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&v[u]);
        const float2 a = __bfloat1622float2(h2[0]), c = __bfloat1622float2(h2[1]);
        const float2 d = __bfloat1622float2(h2[2]), e = __bfloat1622float2(h2[3]);
        __stcs(o32 + 2 * (c0 + u * 256), make_float4(a.x, a.y, c.x, c.y));
        __stcs(o32 + 2 * (c0 + u * 256) + 1, make_float4(d.x, d.y, e.x, e.y));
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// bucketize (right=False -> lower_bound) + two embedding gathers + add; warp per row
// ------------------------------------------------------------------------------------------------
// Each lane keeps 8 consecutive bin edges of both tables in registers; lower_bound(v) = number of edges
// that are not >= v, summed over the warp (edges ascend, so this is the first index with bins[i] >= v,
// and a NaN lands in the last bucket exactly like the binary search / torch.bucketize).
template <int NV>
__global__ void __launch_bounds__(256)
bucket_embed_add_kernel(const __nv_bfloat16* __restrict__ x, const float* __restrict__ p,
                        const float* __restrict__ e, const float* __restrict__ pbins,
                        const float* __restrict__ ebins, const float* __restrict__ ptab,
                        const float* __restrict__ etab, __nv_bfloat16* __restrict__ out,
                        int32_t* __restrict__ p_idx, int32_t* __restrict__ e_idx, float* __restrict__ out32,
                        int64_t rows, int nb) {
  constexpr int D = 128 * NV;
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nwarps = static_cast<int64_t>(gridDim.x) * (blockDim.x >> 5);
  float pb[8], eb[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int k = lane * 8 + i;
    pb[i] = k < nb ? __ldg(pbins + k) : INFINITY;
    eb[i] = k < nb ? __ldg(ebins + k) : INFINITY;
  }
  for (int64_t row = warp0; row < rows; row += nwarps) {
    const float pv = __ldg(p + row), ev = __ldg(e + row);
    // issue the activation loads before the bucket reduction
    uint2 xv[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) xv[i] = __ldcs(reinterpret_cast<const uint2*>(x + row * D) + i * 32 + lane);
    int pc = 0, ec = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      pc += !(pb[i] >= pv);
      ec += !(eb[i] >= ev);
    }
    int both = pc | (ec << 16);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) both += __shfl_xor_sync(0xffffffffu, both, o);
    const int pi = min(both & 0xffff, nb), ei = min(both >> 16, nb);
    if (lane == 0) {
      if (p_idx) p_idx[row] = pi;
      if (e_idx) e_idx[row] = ei;
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = i * 32 + lane;  // 4-element chunk
      const float2 x01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&xv[i].x));
      const float2 x23 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&xv[i].y));
      const float4 a = __ldg(reinterpret_cast<const float4*>(ptab + static_cast<int64_t>(pi) * D) + c);
      const float4 bq = __ldg(reinterpret_cast<const float4*>(etab + static_cast<int64_t>(ei) * D) + c);
      // the reference adds pitch first, then energy (varianceadaptor.py:123-126)
      const float4 r = make_float4((x01.x + a.x) + bq.x, (x01.y + a.y) + bq.y, (x23.x + a.z) + bq.z,
                                   (x23.y + a.w) + bq.w);
      store_bf16x4(out + row * D + c * 4, r);
      if (out32) {  // slot 5 of the reference's return tuple in its dtype: the bf16-rounded value, widened
        const uint2 u = *reinterpret_cast<const uint2*>(out + row * D + c * 4);
        const float2 lo = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.x));
        const float2 hi = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.y));
        __stcs(reinterpret_cast<float4*>(out32 + row * D) + c, make_float4(lo.x, lo.y, hi.x, hi.y));
      }
    }
  }
}

// kv_info[b] = (1 + index of the last valid key, length of the all-valid prefix)
__global__ void mask_lengths_kernel(const uint8_t* __restrict__ mask, int32_t* __restrict__ kv_info, int B, int T) {
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (b >= B) return;
  int last = 0, first_zero = T;
  for (int t = lane; t < T; t += 32) {
    if (mask[static_cast<int64_t>(b) * T + t]) last = t + 1;
    else first_zero = min(first_zero, t);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    last = max(last, __shfl_xor_sync(0xffffffffu, last, o));
    first_zero = min(first_zero, __shfl_xor_sync(0xffffffffu, first_zero, o));
  }
  if (lane == 0) {
    kv_info[2 * b] = last;
    kv_info[2 * b + 1] = first_zero;
  }
}

__global__ void cast_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, int64_t n) {
  const int64_t i = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
  if (i + 3 < n) {
    const float4 v = *reinterpret_cast<const float4*>(x + i);
    store_bf16x4(y + i, v);
  } else {
    for (int64_t j = i; j < n; ++j) y[j] = __float2bfloat16(x[j]);
  }
}

}  // namespace ttsb

using namespace ttsb;

#define TTSB_LAUNCH_CHECK() TTSB_CHECK_CUDA(cudaGetLastError())

extern "C" int ttsb200_mask_lengths(const uint8_t* mask, int32_t* kv_info, int32_t B, int32_t T,
                                    ttsb200_stream_t s) {
  TTSB_CHECK_ARG(mask && kv_info && B > 0 && T > 0, "mask_lengths: bad args");
  mask_lengths_kernel<<<(B + 7) / 8, 256, 0, static_cast<cudaStream_t>(s)>>>(mask, kv_info, B, T);
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

extern "C" int ttsb200_embed_pe_ln(const int64_t* ids, const float* table, const float* pe,
                                   const float* alpha, const float* g, const float* b, float eps,
                                   float* x0, void* xn, int32_t B, int32_t T, int32_t D, int32_t V,
                                   ttsb200_stream_t s) {
  TTSB_CHECK_ARG(ids && table && pe && alpha && g && b && x0 && xn, "embed_pe_ln: null pointer");
  TTSB_CHECK_ARG(B > 0 && T > 0 && V > 0, "embed_pe_ln: bad dims");
  const int64_t rows = static_cast<int64_t>(B) * T;
  const unsigned grid = static_cast<unsigned>((rows + 7) / 8);
  cudaStream_t st = static_cast<cudaStream_t>(s);
  if (D == 384)
    embed_pe_ln_kernel<3><<<grid, 256, 0, st>>>(ids, table, pe, alpha, g, b, eps, x0,
                                                static_cast<__nv_bfloat16*>(xn), rows, T, V);
  else if (D == 256)
    embed_pe_ln_kernel<2><<<grid, 256, 0, st>>>(ids, table, pe, alpha, g, b, eps, x0,
                                                static_cast<__nv_bfloat16*>(xn), rows, T, V);
  else
    TTSB_CHECK_ARG(false, "embed_pe_ln: D=%d unsupported (256 or 384)", D);
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

extern "C" int ttsb200_layernorm(const float* x, const float* g, const float* b, float eps, void* y_bf16,
                                 float* y_f32, int64_t rows, int32_t D, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(x && g && b && (y_bf16 || y_f32) && rows > 0, "layernorm: bad args");
  const unsigned grid = static_cast<unsigned>((rows + 7) / 8);
  cudaStream_t st = static_cast<cudaStream_t>(s);
  if (D == 384)
    layernorm_kernel<3><<<grid, 256, 0, st>>>(x, g, b, eps, static_cast<__nv_bfloat16*>(y_bf16), y_f32, rows);
  else if (D == 256)
    layernorm_kernel<2><<<grid, 256, 0, st>>>(x, g, b, eps, static_cast<__nv_bfloat16*>(y_bf16), y_f32, rows);
  else
    TTSB_CHECK_ARG(false, "layernorm: D=%d unsupported (256 or 384)", D);
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

extern "C" int ttsb200_ffn2_tail(const float* u, const void* yn_bf16, const float* y, const float* ln1_g,
                                 const float* ln1_b, const float* ln2_g, const float* ln2_b, float eps, float* z,
                                 void* zn_bf16, float* zn_f32, float* pre_f32, int64_t rows, int32_t D,
                                 ttsb200_stream_t s) {
  TTSB_CHECK_ARG(u && yn_bf16 && y && ln1_g && ln1_b && ln2_g && ln2_b && z && zn_bf16 && rows > 0,
                 "ffn2_tail: bad args");
  TTSB_CHECK_ARG(D == 384, "ffn2_tail: D=%d unsupported (384)", D);
  const unsigned grid = static_cast<unsigned>((rows + 7) / 8);
  ffn2_tail_kernel<3><<<grid, 256, 0, static_cast<cudaStream_t>(s)>>>(
      u, static_cast<const __nv_bfloat16*>(yn_bf16), y, ln1_g, ln1_b, ln2_g, ln2_b, eps, z,
      static_cast<__nv_bfloat16*>(zn_bf16), zn_f32, pre_f32, rows);
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

// Row-tile list of a ragged batch (ttsb200_gemm_args.tile_map): one block, a shuffle-free serial scan is enough for the
// few hundred utterances of a batch.
__global__ void build_tile_map_kernel(const int32_t* __restrict__ len, int len_stride, int B, int m_tiles, int halo,
                                      int32_t* __restrict__ map, int32_t* __restrict__ count) {
  __shared__ int s_off[1025];
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    const int n = len[static_cast<int64_t>(b) * len_stride];
    int nt = n > 0 ? (n + halo + 127) / 128 : 0;
    s_off[b + 1] = nt < m_tiles ? nt : m_tiles;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    s_off[0] = 0;
    for (int b = 0; b < B; ++b) s_off[b + 1] += s_off[b];
    *count = s_off[B];
  }
  __syncthreads();
  for (int b = threadIdx.x; b < B; b += blockDim.x)
    for (int i = s_off[b]; i < s_off[b + 1]; ++i) map[i] = b * m_tiles + (i - s_off[b]);
}

extern "C" int ttsb200_build_tile_map(const int32_t* len, int32_t len_stride, int32_t B, int32_t m_tiles,
                                      int32_t halo, int32_t* map, int32_t* count, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(len && map && count && B > 0 && B <= 1024 && m_tiles > 0 && len_stride > 0 && halo >= 0,
                 "build_tile_map: bad args (B <= 1024)");
  build_tile_map_kernel<<<1, 256, 0, static_cast<cudaStream_t>(s)>>>(len, len_stride, B, m_tiles, halo, map, count);
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

extern "C" int ttsb200_round_durations(const float* log_d, int32_t* dur, int64_t n, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(log_d && dur && n > 0, "round_durations: bad args");
  round_durations_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(s)>>>(
      log_d, dur, n);
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

extern "C" int ttsb200_lr_scan(const int32_t* dur, int32_t* cum, int32_t* mel_len, int32_t B, int32_t L,
                               ttsb200_stream_t s) {
  TTSB_CHECK_ARG(dur && cum && mel_len && B > 0 && L > 0, "lr_scan: bad args");
  lr_scan_kernel<<<(B + 3) / 4, 128, 0, static_cast<cudaStream_t>(s)>>>(dur, cum, mel_len, B, L);
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

extern "C" int ttsb200_lr_expand(const void* x, const int32_t* cum, const int32_t* mel_len, void* out,
                                 int32_t* frame_idx, int64_t* mel_pos, uint8_t* mel_mask, float* out_f32, int32_t B,
                                 int32_t L, int32_t Tmax, int32_t D, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(x && cum && mel_len && out && B > 0 && L > 0 && Tmax > 0, "lr_expand: bad args");
  TTSB_CHECK_ARG(D % 8 == 0, "lr_expand: D (%d) must be a multiple of 8", D);
  dim3 grid((Tmax + kLrFrames - 1) / kLrFrames, B);
  lr_expand_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(s)>>>(
      static_cast<const __nv_bfloat16*>(x), cum, mel_len, static_cast<__nv_bfloat16*>(out), frame_idx,
      mel_pos, mel_mask, out_f32, L, Tmax, D);
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

extern "C" int ttsb200_bucket_embed_add(const void* x, const float* p, const float* e,
                                        const float* pbins, const float* ebins, const float* ptab,
                                        const float* etab, void* out, int32_t* p_idx, int32_t* e_idx,
                                        float* out_f32, int64_t rows, int32_t D, int32_t nb, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(x && p && e && pbins && ebins && ptab && etab && out && rows > 0 && nb > 0,
                 "bucket_embed_add: bad args");
  TTSB_CHECK_ARG(nb <= 256, "bucket_embed_add: at most 256 bin edges (got %d)", nb);
  const int64_t want = (rows + 7) / 8;
  const unsigned grid = static_cast<unsigned>(want < 148 * 16 ? want : 148 * 16);  // persistent warps, 4+ rows each
  cudaStream_t st = static_cast<cudaStream_t>(s);
  if (D == 384)
    bucket_embed_add_kernel<3><<<grid, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x), p, e, pbins,
                                                     ebins, ptab, etab, static_cast<__nv_bfloat16*>(out),
                                                     p_idx, e_idx, out_f32, rows, nb);
  else if (D == 256)
    bucket_embed_add_kernel<2><<<grid, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x), p, e, pbins,
                                                     ebins, ptab, etab, static_cast<__nv_bfloat16*>(out),
                                                     p_idx, e_idx, out_f32, rows, nb);
  else
    TTSB_CHECK_ARG(false, "bucket_embed_add: D=%d unsupported (256 or 384)", D);
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

extern "C" int ttsb200_cast_bf16(const float* x, void* y, int64_t n, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(x && y && n > 0, "cast_bf16: bad args");
  TTSB_CHECK_ARG((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(y) & 7) == 0,
                 "cast_bf16: alignment");
  const int64_t thr = (n + 3) / 4;
  cast_bf16_kernel<<<static_cast<unsigned>((thr + 255) / 256), 256, 0, static_cast<cudaStream_t>(s)>>>(
      x, static_cast<__nv_bfloat16*>(y), n);
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

// ------------------------------------------------------------------------------------------------
// LengthRegulator without allocations: ttsb200_length_regulate_static.
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  ttsb200_lr_scan /
// ttsb200_lr_expand take their scratch (cum (B,L) and mel_len (B)) from the caller; here it lives in a file-scope
// device array of fixed size that every GPU loading the library holds until exit (260 KB).
//
// g_lr_scratch is a HOST-SIDE STAND-IN in host code, not a device address: the entry point resolves the device
// address once per device with cudaGetSymbolAddress (that device current) and passes the pointer to the kernels.
// Nothing has to start at zero: lr_scan_kernel writes every cum[b][0..L) and mel_len[b] that lr_expand_kernel
// reads, in the same call and stream, so the previous call's contents are never observed.  All calls on one
// device share the array: calls that overlap on the GPU (different streams, no dependency) race.
// ------------------------------------------------------------------------------------------------
namespace ttsb {
constexpr int kLrStaticMaxCum = TTSB200_LR_STATIC_MAX_PHONEMES;  // B*L entries of cum
constexpr int kLrStaticMaxB = TTSB200_LR_STATIC_MAX_BATCH;       // mel_len entries
static __device__ __align__(256) int32_t g_lr_scratch[kLrStaticMaxCum + kLrStaticMaxB];

constexpr int kLrMaxDevices = 64;
static int32_t* g_lr_scratch_ptr[kLrMaxDevices];  // device address of g_lr_scratch, per device ordinal
static std::mutex g_lr_scratch_mu;

// Device address of g_lr_scratch on the CURRENT device (resolved once per device, return code checked).
static int lr_static_scratch(int32_t** out) {
  int dev = 0;
  TTSB_CHECK_CUDA(cudaGetDevice(&dev));
  TTSB_CHECK_ARG(dev >= 0 && dev < kLrMaxDevices, "length_regulate_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_lr_scratch_mu);
  if (!g_lr_scratch_ptr[dev]) {
    void* ptr = nullptr;
    TTSB_CHECK_CUDA(cudaGetSymbolAddress(&ptr, g_lr_scratch));
    TTSB_CHECK_ARG(ptr != nullptr, "length_regulate_static: cudaGetSymbolAddress returned a null address");
    g_lr_scratch_ptr[dev] = static_cast<int32_t*>(ptr);
  }
  *out = g_lr_scratch_ptr[dev];
  return TTSB_OK;
}
}  // namespace ttsb

extern "C" int ttsb200_length_regulate_static(const void* x, const int32_t* dur, void* out, int32_t* frame_idx,
                                              int64_t* mel_pos, uint8_t* mel_mask, float* out_f32,
                                              int32_t* mel_len_out, int32_t B, int32_t L, int32_t Tmax, int32_t D,
                                              ttsb200_stream_t s) {
  TTSB_CHECK_ARG(x && dur && out && B > 0 && L > 0 && Tmax > 0, "length_regulate_static: bad args");
  TTSB_CHECK_ARG(D % 8 == 0, "length_regulate_static: D (%d) must be a multiple of 8", D);
  TTSB_CHECK_ARG(B <= kLrStaticMaxB && static_cast<int64_t>(B) * L <= kLrStaticMaxCum,
                 "length_regulate_static: B=%d, B*L=%lld exceed the static scratch (B <= %d, B*L <= %d); use "
                 "ttsb200_lr_scan + ttsb200_lr_expand with caller-owned scratch",
                 B, static_cast<long long>(B) * L, kLrStaticMaxB, kLrStaticMaxCum);
  int32_t* scratch = nullptr;
  const int rc = lr_static_scratch(&scratch);
  if (rc != TTSB_OK) return rc;
  int32_t* cum = scratch;
  int32_t* mel_len = scratch + kLrStaticMaxCum;
  cudaStream_t st = static_cast<cudaStream_t>(s);
  lr_scan_kernel<<<(B + 3) / 4, 128, 0, st>>>(dur, cum, mel_len, B, L);
  TTSB_LAUNCH_CHECK();
  dim3 grid((Tmax + kLrFrames - 1) / kLrFrames, B);
  lr_expand_kernel<<<grid, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x), cum, mel_len,
                                         static_cast<__nv_bfloat16*>(out), frame_idx, mel_pos, mel_mask, out_f32, L,
                                         Tmax, D);
  TTSB_LAUNCH_CHECK();
  if (mel_len_out)  // device-to-device copy of B int32: no allocation, legal under stream capture
    TTSB_CHECK_CUDA(cudaMemcpyAsync(mel_len_out, mel_len, sizeof(int32_t) * B, cudaMemcpyDeviceToDevice, st));
  return TTSB_OK;
}
