// Autoregressive decode step of Transformer-TTS (BASELINE.json configs[4]): the per-position work of
// Models/decoder.py:43-56 / Models/layers.py:109-124 for ONE new decoder position of every utterance, with the keys and
// values of the earlier positions kept in a cache instead of being recomputed (the reference's test.py:128-191 re-runs
// the whole prefix through the model at every step).
//
// A step has R = batch rows (16 at cfg5), so every contraction is [R x K] . [K x N] with R <= 16: one
// mma.sync.m16n8k16 tile in M.  The work is streaming the layer weights (28 MB bf16 for six layers, L2-resident) and
// the cached keys / values; there is no tile big enough for tcgen05 (M = 128 minimum), so these are warp-level kernels:
//
//   small_linear : y = act(LN?(x) W^T + b) (+ residual) (+ alpha * pe[pos]); 32 output columns per CTA (8 warps = 4 column
//                  groups x 2 k-slices); the CTA's W rows and the R input rows travel to shared memory by cp.async, all in
//                  flight at once; fp32 rows are normalised from shared memory; the contraction index is permuted so that
//                  a 16-byte vector of a W row is two k-steps' worth of B fragments;
//   attn_decode  : one query per (utterance, head) against the cached keys/values (appending this step's first):
//                  thread-per-key dot products with 16-byte loads, block softmax, 12-thread-per-key P.V;
//   ffn_tail_rows: FeedForward tail + the next LayerNorm for a few rows.
//
// What costs time here is the number of DEPENDENT memory round trips and launches, not bytes: kernels
// launch with programmatic dependent launch and load what is immutable before waiting for their predecessor; the
// stretches between a block's attentions run as thread-block-cluster kernels (decode_cluster.cu).
// Position-dependent addressing (which cache slot, which positional-encoding row, where the step's output goes) reads a
// device scalar, so one captured CUDA graph replays for every step.  TTSB_NO_PDL=1 in the environment launches without
// the dependent-launch attribute (A/B: B = 1, 378 -> 327 us per position with it).
#include <cuda_bf16.h>
#include <stdlib.h>

#include "common.cuh"
#include "../../include/ttsb200.h"

namespace ttsb {
namespace dec {

constexpr int kSlColGroups = 4;      // 8-column groups per CTA: 32 output columns
constexpr int kSlKSlices = 2;        // warps that share one column group, each taking every 2nd 32-wide k-block
constexpr int kSlWarps = kSlColGroups * kSlKSlices;
constexpr int kSlThreads = kSlWarps * 32;
constexpr int kSlCols = kSlColGroups * 8;
constexpr int kSlRows = 16;          // rows per CTA = M of the MMA tile
constexpr int kSlPadBytes = 16;      // row padding in shared memory: rows g, g+1, .. land in different 16-byte bank groups

__device__ __forceinline__ void mma_bf16_16816(float* c, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                               uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// Programmatic dependent launch: a decode step is ~58 dependent kernels of a few microseconds, so the gap between two
// of them matters as much as the kernels.  Each kernel lets its successor start at once (launch_dependents) and does
// whatever touches only immutable data (weight fragments) before it waits for its predecessor's results (wait returns
// when the preceding kernel has completed and its writes are visible).  Pre-wait code must not read anything a kernel
// of the step writes: two kernels back may still be running.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// 16-byte asynchronous global -> shared copy (src_bytes = 0 writes zeros): every copy of a stage is issued before the first
// one is waited for, so a stage costs ONE trip to L2 / HBM however many rows it has.
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

template <typename... Params, typename... Args>
static cudaError_t launch_pdl(void (*kern)(Params...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  static const bool no_pdl = getenv("TTSB_NO_PDL") != nullptr;
  cfg.numAttrs = no_pdl ? 0 : 1;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<Params>(args)...);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// A decode step is a chain of ~58 dependent kernels with a few microseconds of work each: what matters inside a kernel is
// the number of DEPENDENT memory round trips, not bytes.  Hence: every global -> shared transfer of a stage is a cp.async
// issued before any is waited for (one round trip per stage); 8 warps per CTA, so the LayerNorm pass has 2 rows per warp
// and the contraction is cut in 4 k-slices per column group (partial sums meet in shared memory).
__global__ void __launch_bounds__(kSlThreads)
small_linear_kernel(const ttsb200_small_linear_args p) {
  extern __shared__ __align__(16) uint8_t sl_smem[];
  __shared__ float s_red[kSlKSlices - 1][kSlColGroups][32][4];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cgp = warp % kSlColGroups, ks = warp / kSlColGroups;
  const int Kp = (p.K + 31) & ~31;
  const int row_bytes = Kp * 2 + kSlPadBytes;
  const int r0 = blockIdx.y * kSlRows;
  pdl_launch_dependents();
  // weights are immutable: this CTA's 16 W rows leave for shared memory before the predecessor has finished
  uint8_t* sW = sl_smem + kSlRows * row_bytes;
  {
    const int nvp = Kp >> 3;   // 16-byte vectors per (zero-padded) row
    const int ncta0 = blockIdx.x * kSlCols;
    for (int i = threadIdx.x; i < kSlCols * nvp; i += kSlThreads) {
      const int r = i / nvp, c = i - r * nvp;
      const int wr = min(ncta0 + r, p.N - 1);
      cp_async16(sW + r * row_bytes + c * 16,
                 static_cast<const __nv_bfloat16*>(p.w_bf16) + static_cast<int64_t>(wr) * p.ldw + c * 8, 16);
    }
  }
  const int n0 = blockIdx.x * kSlCols + cgp * 8;
  const int g = lane >> 2, tq = lane & 3;
  const int nkb = Kp >> 5;
  // epilogue operands of this lane's two columns: bias is immutable (loaded now), the residual after the wait
  const int ncol = n0 + tq * 2;
  float bias_v[2] = {0.f, 0.f};
  if (ks == 0 && p.bias) {
    if (ncol < p.N) bias_v[0] = __ldg(p.bias + ncol);
    if (ncol + 1 < p.N) bias_v[1] = __ldg(p.bias + ncol + 1);
  }
  pdl_wait();
  const int pos = p.pos ? *reinterpret_cast<const volatile int32_t*>(p.pos) : 0;
  const int xpos = pos + p.x_pos_bias;   // index of the step the input rows were produced at (negative: zeros)

  // ---- stage the input rows as bf16 (LayerNorm applied on the way) ----
  const bool pos_ok = p.x_pos_stride == 0 || xpos >= 0;
  const int64_t pos_off = p.x_pos_stride ? static_cast<int64_t>(xpos) * p.x_pos_stride : 0;
  float* sF = reinterpret_cast<float*>(sW + kSlCols * row_bytes);   // fp32 staging rows [16][Kp] (fp32 inputs only)
  const bool vec_bf16 = p.x_bf16 && (p.K & 7) == 0 && (p.ldx & 7) == 0 && (p.x_pos_stride & 7) == 0 &&
                        (reinterpret_cast<uintptr_t>(p.x_bf16) & 15) == 0;
  const bool vec_f32 = p.x_f32 && (p.K & 3) == 0 && (p.ldx & 3) == 0 && (p.x_pos_stride & 3) == 0 &&
                       (reinterpret_cast<uintptr_t>(p.x_f32) & 15) == 0;
  if (vec_bf16) {
    const int nv = p.K >> 3, nvp = Kp >> 3;
    for (int i = threadIdx.x; i < kSlRows * nvp; i += kSlThreads) {
      const int rr = i / nvp, c = i - rr * nvp;
      const int row = r0 + rr;
      const bool valid = row < p.R && pos_ok && c < nv;
      cp_async16(sl_smem + rr * row_bytes + c * 16,
                 static_cast<const __nv_bfloat16*>(p.x_bf16) + static_cast<int64_t>(valid ? row : 0) * p.ldx + pos_off + (valid ? c * 8 : 0),
                 valid ? 16 : 0);
    }
  } else if (vec_f32) {
    const int nv = p.K >> 2, nvp = Kp >> 2;
    for (int i = threadIdx.x; i < kSlRows * nvp; i += kSlThreads) {
      const int rr = i / nvp, c = i - rr * nvp;
      const int row = r0 + rr;
      const bool valid = row < p.R && pos_ok && c < nv;
      cp_async16(sF + rr * Kp + c * 4, p.x_f32 + static_cast<int64_t>(valid ? row : 0) * p.ldx + pos_off + (valid ? c * 4 : 0),
                 valid ? 16 : 0);
    }
  } else {   // unaligned rows: element-wise
    for (int i = threadIdx.x; i < kSlRows * Kp; i += kSlThreads) {
      const int rr = i / Kp, k = i - rr * Kp;
      const int row = r0 + rr;
      const bool valid = row < p.R && pos_ok && k < p.K;
      const int64_t off = static_cast<int64_t>(valid ? row : 0) * p.ldx + pos_off + (valid ? k : 0);
      if (p.x_bf16)
        reinterpret_cast<__nv_bfloat16*>(sl_smem + rr * row_bytes)[k] =
            valid ? static_cast<const __nv_bfloat16*>(p.x_bf16)[off] : __float2bfloat16(0.f);
      else
        sF[rr * Kp + k] = valid ? p.x_f32[off] : 0.f;
    }
  }
  float res_v[4] = {0.f, 0.f, 0.f, 0.f};   // in flight together with the input rows
  if (ks == 0 && p.res_f32) {
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int row = r0 + g + half * 8;
#pragma unroll
      for (int e = 0; e < 2; ++e)
        if (row < p.R && ncol + e < p.N) res_v[half * 2 + e] = p.res_f32[static_cast<int64_t>(row) * p.ldres + ncol + e];
    }
  }
  cp_async_wait_all();
  __syncthreads();
  if (p.x_f32) {   // fp32 rows in shared memory -> (LayerNorm) -> bf16 operand rows
    __shared__ float2 s_stat[kSlRows];   // (mean, rstd) per row
    if (p.ln_g) {   // two-pass statistics, as torch.nn.functional.layer_norm: one warp per row, 2 rows per warp
      const float inv_k = 1.f / static_cast<float>(p.K);
      for (int rr = warp; rr < kSlRows; rr += kSlWarps) {
        const float* src = sF + rr * Kp;
        float sum = 0.f;
        for (int k = lane; k < p.K; k += 32) sum += src[k];
        const float mean = warp_sum(sum) * inv_k;
        float q = 0.f;
        for (int k = lane; k < p.K; k += 32) {
          const float d = src[k] - mean;
          q = fmaf(d, d, q);
        }
        const float rstd = rsqrtf(warp_sum(q) * inv_k + p.eps);
        if (lane == 0) s_stat[rr] = make_float2(mean, rstd);
      }
      __syncthreads();
    }
    // element pairs, flattened over the tile: thread t takes pair columns t % (Kp/2) of rows t / (Kp/2), ...
    const int hp = Kp >> 1;
    int rr = threadIdx.x / hp, c = threadIdx.x - rr * hp;
    const int drow = kSlThreads / hp, dcol = kSlThreads - drow * hp;
    while (rr < kSlRows) {
      const int k = 2 * c;
      float2 v = *reinterpret_cast<const float2*>(sF + rr * Kp + k);
      if (p.ln_g && k < p.K) {   // (K is even: checked on the host for LayerNorm inputs)
        const float2 st = s_stat[rr];
        const float2 gg = __ldg(reinterpret_cast<const float2*>(p.ln_g + k)), bb = __ldg(reinterpret_cast<const float2*>(p.ln_b + k));
        v.x = fmaf((v.x - st.x) * st.y, gg.x, bb.x);
        v.y = fmaf((v.y - st.x) * st.y, gg.y, bb.y);
      }
      const __nv_bfloat162 h = __floats2bfloat162_rn(v.x, v.y);
      *reinterpret_cast<__nv_bfloat162*>(sl_smem + rr * row_bytes + k * 2) = h;
      const int row = r0 + rr;
      if (p.xn_bf16 && blockIdx.x == 0 && row < p.R && k < p.K)
        *reinterpret_cast<__nv_bfloat162*>(static_cast<__nv_bfloat16*>(p.xn_bf16) + static_cast<int64_t>(row) * p.K + k) = h;
      rr += drow;
      c += dcol;
      if (c >= hp) {
        c -= hp;
        ++rr;
      }
    }
    __syncthreads();
  }

  // ---- 16 x 8 outputs per column group; contraction index permuted: lane (g, tq) takes physical k = 32 kb + 8 tq + [0, 8)
  // of W row n0 + g as the B fragments of two k-steps, and the same physical k of x rows g, g + 8 as the A fragments;
  // k-slice ks takes k-blocks ks, ks + 4, ... ----
  const uint8_t* xa = sl_smem + g * row_bytes + tq * 16;
  const uint8_t* xb = xa + 8 * row_bytes;
  const uint8_t* wa = sW + (cgp * 8 + g) * row_bytes + tq * 16;
  // mma.sync has a long latency on this part (a chain of 12 dependent ones measured ~1.8 k clk): four independent
  // accumulator chains, summed at the end
  float accs[4][4] = {};
  {
    int kb = ks;
    for (; kb + kSlKSlices < nkb; kb += 2 * kSlKSlices) {
      const uint4 a0 = *reinterpret_cast<const uint4*>(xa + kb * 64), a1 = *reinterpret_cast<const uint4*>(xa + (kb + kSlKSlices) * 64);
      const uint4 b0 = *reinterpret_cast<const uint4*>(xb + kb * 64), b1 = *reinterpret_cast<const uint4*>(xb + (kb + kSlKSlices) * 64);
      const uint4 w0 = *reinterpret_cast<const uint4*>(wa + kb * 64), w1 = *reinterpret_cast<const uint4*>(wa + (kb + kSlKSlices) * 64);
      mma_bf16_16816(accs[0], a0.x, b0.x, a0.y, b0.y, w0.x, w0.y);
      mma_bf16_16816(accs[1], a0.z, b0.z, a0.w, b0.w, w0.z, w0.w);
      mma_bf16_16816(accs[2], a1.x, b1.x, a1.y, b1.y, w1.x, w1.y);
      mma_bf16_16816(accs[3], a1.z, b1.z, a1.w, b1.w, w1.z, w1.w);
    }
    if (kb < nkb) {
      const uint4 a = *reinterpret_cast<const uint4*>(xa + kb * 64);
      const uint4 b = *reinterpret_cast<const uint4*>(xb + kb * 64);
      const uint4 w = *reinterpret_cast<const uint4*>(wa + kb * 64);
      mma_bf16_16816(accs[0], a.x, b.x, a.y, b.y, w.x, w.y);
      mma_bf16_16816(accs[1], a.z, b.z, a.w, b.w, w.z, w.w);
    }
  }
  float acc[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) acc[e] = (accs[0][e] + accs[1][e]) + (accs[2][e] + accs[3][e]);
  if (ks > 0) {
#pragma unroll
    for (int e = 0; e < 4; ++e) s_red[ks - 1][cgp][lane][e] = acc[e];
  }
  __syncthreads();
  if (ks > 0 || n0 >= p.N) return;
#pragma unroll
  for (int q = 0; q < kSlKSlices - 1; ++q)
#pragma unroll
    for (int e = 0; e < 4; ++e) acc[e] += s_red[q][cgp][lane][e];

  // ---- epilogue: c0,c1 = (row g, cols 2 tq, 2 tq + 1); c2,c3 = (row g + 8, same cols) ----
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    const int row = r0 + g + half * 8;
    if (row >= p.R) continue;
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int n = n0 + tq * 2 + e;
      if (n >= p.N) continue;
      float v = acc[half * 2 + e] + bias_v[e];
      if (p.relu) v = fmaxf(v, 0.f);
      v += res_v[half * 2 + e];
      if (p.add_row) v = fmaf(__ldg(p.add_scale), __ldg(p.add_row + static_cast<int64_t>(pos) * p.N + n), v);
      const int64_t opos = p.out_pos_stride ? static_cast<int64_t>(pos) * p.out_pos_stride : 0;
      if (p.out_f32) p.out_f32[static_cast<int64_t>(row) * p.ldo + opos + n] = v;
      if (p.out_bf16) static_cast<__nv_bfloat16*>(p.out_bf16)[static_cast<int64_t>(row) * p.ldob + n] = __float2bfloat16(v);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// attn_decode: grid (H, R), 512 threads.  Models/modules.py:7-21 for one query row.
// ------------------------------------------------------------------------------------------------
constexpr int kAdThreads = 512;

template <int DH>
__global__ void __launch_bounds__(kAdThreads)
attn_decode_kernel(const __nv_bfloat16* __restrict__ q, int64_t ldq, const __nv_bfloat16* __restrict__ knew,
                   const __nv_bfloat16* __restrict__ vnew, int64_t ldkv, __nv_bfloat16* __restrict__ kc,
                   __nv_bfloat16* __restrict__ vc, const int32_t* __restrict__ pos_ptr,
                   const int32_t* __restrict__ kv_len, __nv_bfloat16* __restrict__ ctx, int64_t ldc, int cap, int64_t sb,
                   int64_t sh, int64_t sj) {
  static_assert(DH % 8 == 0 && DH <= 128, "head dimension");
  constexpr int VEC = DH / 8;                  // 16-byte vectors per key row
  constexpr int GROUPS = kAdThreads / VEC;     // key groups of the P.V phase
  extern __shared__ __align__(16) uint8_t ad_smem[];
  float* s_p = reinterpret_cast<float*>(ad_smem);       // [cap] scores, then probabilities
  __shared__ float s_q[DH];
  __shared__ float s_red[kAdThreads / 32];
  __shared__ float s_part[GROUPS][DH];
  const int h = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
  pdl_launch_dependents();
  pdl_wait();
  const int pos = pos_ptr ? *reinterpret_cast<const volatile int32_t*>(pos_ptr) : 0;
  int n = pos_ptr ? min(pos + 1, cap) : cap;
  if (kv_len) n = min(n, __ldg(kv_len + b));
  __nv_bfloat16* kb = kc + b * sb + h * sh;
  __nv_bfloat16* vb = vc + b * sb + h * sh;
  if (knew && pos < cap && tid < DH) {          // this step's key / value join the cache first
    kb[pos * sj + tid] = knew[static_cast<int64_t>(b) * ldkv + h * DH + tid];
    vb[pos * sj + tid] = vnew[static_cast<int64_t>(b) * ldkv + h * DH + tid];
  }
  if (tid < DH) s_q[tid] = __bfloat162float(q[static_cast<int64_t>(b) * ldq + h * DH + tid]) * rsqrtf(static_cast<float>(DH));
  __syncthreads();

  // scores: one key per thread
  float mx = -INFINITY;
  for (int j = tid; j < n; j += kAdThreads) {
    const uint4* kr = reinterpret_cast<const uint4*>(kb + j * sj);
    uint4 kv[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) kv[i] = kr[i];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
      const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&kv[i]);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f = __bfloat1622float2(h2[e]);
        s = fmaf(f.x, s_q[i * 8 + e * 2], s);
        s = fmaf(f.y, s_q[i * 8 + e * 2 + 1], s);
      }
    }
    s_p[j] = s;
    mx = fmaxf(mx, s);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if ((tid & 31) == 0) s_red[tid >> 5] = mx;
  __syncthreads();
  mx = s_red[0];
#pragma unroll
  for (int i = 1; i < kAdThreads / 32; ++i) mx = fmaxf(mx, s_red[i]);
  __syncthreads();
  float sum = 0.f;
  for (int j = tid; j < n; j += kAdThreads) {
    const float e = expf(s_p[j] - mx);
    s_p[j] = e;
    sum += e;
  }
  sum = warp_sum(sum);
  if ((tid & 31) == 0) s_red[tid >> 5] = sum;
  __syncthreads();
  sum = 0.f;
#pragma unroll
  for (int i = 0; i < kAdThreads / 32; ++i) sum += s_red[i];

  // P.V: VEC threads per key (16 bytes of the value row each), GROUPS keys in flight
  const int grp = tid / VEC, vi = tid - grp * VEC;
  float o[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (grp < GROUPS) {
    constexpr int UV = 8;   // value rows in flight per thread
    for (int j0 = grp; j0 < n; j0 += GROUPS * UV) {
      uint4 vv[UV];
      float pj[UV];
#pragma unroll
      for (int u = 0; u < UV; ++u) {
        const int j = j0 + u * GROUPS;
        const bool ok = j < n;
        vv[u] = ok ? *reinterpret_cast<const uint4*>(vb + j * sj + vi * 8) : make_uint4(0, 0, 0, 0);
        pj[u] = ok ? s_p[j] : 0.f;
      }
#pragma unroll
      for (int u = 0; u < UV; ++u) {
        const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&vv[u]);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 f = __bfloat1622float2(h2[e]);
          o[e * 2] = fmaf(pj[u], f.x, o[e * 2]);
          o[e * 2 + 1] = fmaf(pj[u], f.y, o[e * 2 + 1]);
        }
      }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) s_part[grp][vi * 8 + e] = o[e];
  }
  __syncthreads();
  if (tid < DH) {
    float a = 0.f;
#pragma unroll 1
    for (int gI = 0; gI < GROUPS; ++gI) a += s_part[gI][tid];
    ctx[static_cast<int64_t>(b) * ldc + h * DH + tid] = __float2bfloat16(a / sum);
  }
}

// FeedForward tail of a decoder block for a few rows (Models/modules.py:85-87, Models/layers.py:122, then the NEXT
// LayerNorm: norm_1 of the following block, Models/layers.py:110, or the decoder's final norm, Models/decoder.py:56):
//   z = y + LN(u + yn; ln1);  zn = bf16(LN(z; ln2))      one warp per row, D = 384
// Same arithmetic as ttsb200_ffn2_tail; this one takes part in the programmatic-dependent-launch chain of the step.
__global__ void __launch_bounds__(512)
ffn_tail_rows_kernel(const float* __restrict__ u, const __nv_bfloat16* __restrict__ yn, const float* __restrict__ y,
                     const float* __restrict__ g1, const float* __restrict__ b1, const float* __restrict__ g2,
                     const float* __restrict__ b2, float eps, float* __restrict__ z, __nv_bfloat16* __restrict__ zn, int R) {
  constexpr int D = 384, NV = D / 32;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 16 + warp;
  pdl_launch_dependents();
  float ga[NV], ba[NV], gb[NV], bb[NV];
#pragma unroll
  for (int j = 0; j < NV; ++j) {   // LayerNorm parameters are immutable: loaded before the wait
    const int k = lane + 32 * j;
    ga[j] = __ldg(g1 + k); ba[j] = __ldg(b1 + k); gb[j] = __ldg(g2 + k); bb[j] = __ldg(b2 + k);
  }
  pdl_wait();
  if (row >= R) return;
  float v[NV], yy[NV];
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int64_t o = static_cast<int64_t>(row) * D + lane + 32 * j;
    v[j] = u[o] + __bfloat162float(yn[o]);
    yy[j] = y[o];
  }
  auto stats = [&](const float* a, float& mean, float& rstd) {   // two-pass, as torch
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) s += a[j];
    mean = warp_sum(s) * (1.f / D);
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) q = fmaf(a[j] - mean, a[j] - mean, q);
    rstd = rsqrtf(warp_sum(q) * (1.f / D) + eps);
  };
  float mean, rstd;
  stats(v, mean, rstd);
#pragma unroll
  for (int j = 0; j < NV; ++j) v[j] = yy[j] + fmaf((v[j] - mean) * rstd, ga[j], ba[j]);
  stats(v, mean, rstd);
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int64_t o = static_cast<int64_t>(row) * D + lane + 32 * j;
    z[o] = v[j];
    zn[o] = __float2bfloat16(fmaf((v[j] - mean) * rstd, gb[j], bb[j]));
  }
}

__global__ void advance_pos_kernel(int32_t* pos, int32_t delta) {
  pdl_launch_dependents();
  pdl_wait();
  if (threadIdx.x == 0 && blockIdx.x == 0) *pos += delta;
}

}  // namespace dec
}  // namespace ttsb

using namespace ttsb;

extern "C" int ttsb200_small_linear(const ttsb200_small_linear_args* a, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(a != nullptr, "small_linear: null args");
  const ttsb200_small_linear_args p = *a;
  TTSB_CHECK_ARG((p.x_f32 != nullptr) != (p.x_bf16 != nullptr), "small_linear: exactly one of x_f32 / x_bf16");
  TTSB_CHECK_ARG(p.w_bf16 && p.R > 0 && p.K > 0 && p.N > 0, "small_linear: bad args");
  TTSB_CHECK_ARG(p.out_f32 || p.out_bf16, "small_linear: no output");
  const int Kp = (p.K + 31) & ~31;
  TTSB_CHECK_ARG(p.ldw >= Kp && p.ldw % 8 == 0 && (reinterpret_cast<uintptr_t>(p.w_bf16) & 15) == 0,
                 "small_linear: weight rows must be 16-byte aligned and zero-padded to a multiple of 32 columns (ldw %lld, K %d)",
                 static_cast<long long>(p.ldw), p.K);
  TTSB_CHECK_ARG((p.ln_g == nullptr) == (p.ln_b == nullptr) && (!p.ln_g || p.x_f32), "small_linear: LayerNorm needs ln_g, ln_b and x_f32");
  TTSB_CHECK_ARG(!p.xn_bf16 || p.x_f32, "small_linear: xn_bf16 needs x_f32");
  TTSB_CHECK_ARG(!(p.ln_g || p.xn_bf16) || (p.K % 2 == 0 && (reinterpret_cast<uintptr_t>(p.ln_g) & 7) == 0 &&
                                            (reinterpret_cast<uintptr_t>(p.ln_b) & 7) == 0 &&
                                            (reinterpret_cast<uintptr_t>(p.xn_bf16) & 3) == 0),
                 "small_linear: LayerNorm / xn rows need an even K and 8-byte aligned ln_g, ln_b");
  TTSB_CHECK_ARG((p.add_row == nullptr) == (p.add_scale == nullptr), "small_linear: add_row and add_scale go together");
  TTSB_CHECK_ARG((!p.x_pos_stride && !p.out_pos_stride && !p.add_row) || p.pos, "small_linear: position-dependent addressing needs pos");
  TTSB_CHECK_ARG(!p.out_f32 || p.ldo >= p.N, "small_linear: ldo < N");
  TTSB_CHECK_ARG(!p.out_bf16 || p.ldob >= p.N, "small_linear: ldob < N");
  // x rows + this CTA's W rows (bf16, padded) + fp32 staging rows for fp32 inputs
  const int smem = (dec::kSlRows + dec::kSlCols) * (Kp * 2 + dec::kSlPadBytes) + (p.x_f32 ? dec::kSlRows * Kp * 4 : 0);
  TTSB_CHECK_ARG(smem <= 200 * 1024, "small_linear: K = %d is too large", p.K);
  auto kern = dec::small_linear_kernel;
  TTSB_SMEM_OPT_IN(kern, 200 * 1024);
  dim3 grid((p.N + dec::kSlCols - 1) / dec::kSlCols, (p.R + dec::kSlRows - 1) / dec::kSlRows);
  TTSB_CHECK_CUDA(dec::launch_pdl(kern, grid, dim3(dec::kSlThreads), smem, static_cast<cudaStream_t>(s), p));
  return TTSB_OK;
}

extern "C" int ttsb200_attn_decode(const void* q_bf16, int64_t ldq, const void* knew_bf16, const void* vnew_bf16,
                                   int64_t ldkv, void* k_cache, void* v_cache, int64_t kv_stride_b, int64_t kv_stride_h,
                                   int64_t kv_stride_j, const int32_t* pos, const int32_t* kv_len, void* ctx_bf16,
                                   int64_t ldc, int32_t R, int32_t H, int32_t Dh, int32_t cap, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(q_bf16 && k_cache && v_cache && ctx_bf16 && R > 0 && H > 0 && cap > 0, "attn_decode: bad args");
  TTSB_CHECK_ARG(Dh == 96, "attn_decode: head dimension %d unsupported (96)", Dh);
  TTSB_CHECK_ARG((knew_bf16 == nullptr) == (vnew_bf16 == nullptr), "attn_decode: knew and vnew go together");
  TTSB_CHECK_ARG(pos || kv_len, "attn_decode: pos (self-attention) and / or kv_len (cross-attention) is required");
  TTSB_CHECK_ARG(!knew_bf16 || pos, "attn_decode: appending needs pos");
  TTSB_CHECK_ARG(cap <= 16384, "attn_decode: %d keys above the 16384 supported", cap);
  TTSB_CHECK_ARG((reinterpret_cast<uintptr_t>(k_cache) & 15) == 0 && (reinterpret_cast<uintptr_t>(v_cache) & 15) == 0 &&
                     kv_stride_b % 8 == 0 && kv_stride_h % 8 == 0 && kv_stride_j % 8 == 0 && kv_stride_j >= Dh,
                 "attn_decode: key / value rows must be 16-byte aligned");
  auto kern = dec::attn_decode_kernel<96>;
  const int smem = cap * 4;
  TTSB_SMEM_OPT_IN(kern, 16384 * 4);
  TTSB_CHECK_CUDA(dec::launch_pdl(kern, dim3(H, R), dim3(dec::kAdThreads), smem, static_cast<cudaStream_t>(s),
                                  static_cast<const __nv_bfloat16*>(q_bf16), ldq, static_cast<const __nv_bfloat16*>(knew_bf16),
                                  static_cast<const __nv_bfloat16*>(vnew_bf16), ldkv, static_cast<__nv_bfloat16*>(k_cache),
                                  static_cast<__nv_bfloat16*>(v_cache), pos, kv_len, static_cast<__nv_bfloat16*>(ctx_bf16),
                                  ldc, static_cast<int>(cap), kv_stride_b, kv_stride_h, kv_stride_j));
  return TTSB_OK;
}

extern "C" int ttsb200_ffn_tail_rows(const float* u, const void* yn_bf16, const float* y, const float* ln1_g,
                                     const float* ln1_b, const float* ln2_g, const float* ln2_b, float eps, float* z,
                                     void* zn_bf16, int32_t R, int32_t D, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(u && yn_bf16 && y && ln1_g && ln1_b && ln2_g && ln2_b && z && zn_bf16 && R > 0, "ffn_tail_rows: bad args");
  TTSB_CHECK_ARG(D == 384, "ffn_tail_rows: D = %d unsupported (384)", D);
  TTSB_CHECK_CUDA(dec::launch_pdl(dec::ffn_tail_rows_kernel, dim3((R + 15) / 16), dim3(512), 0, static_cast<cudaStream_t>(s),
                                  u, static_cast<const __nv_bfloat16*>(yn_bf16), y, ln1_g, ln1_b, ln2_g, ln2_b, eps, z,
                                  static_cast<__nv_bfloat16*>(zn_bf16), static_cast<int>(R)));
  return TTSB_OK;
}

extern "C" int ttsb200_advance_pos(int32_t* pos, int32_t delta, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(pos != nullptr, "advance_pos: null pointer");
  TTSB_CHECK_CUDA(dec::launch_pdl(dec::advance_pos_kernel, dim3(1), dim3(32), 0, static_cast<cudaStream_t>(s), pos, delta));
  return TTSB_OK;
}
