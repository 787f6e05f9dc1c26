// HBM-bound row kernels of the FastSpeech2 TRAINING step (backward passes of the row-wise ops, losses,
// BatchNorm batch statistics, optimizer).  The reference gets all of these from PyTorch autograd /
// torch.optim (train_fastspeech2.py:212-217,245-259,304-315); here every one is a hand-written kernel.
// Conventions as in rowwise.cu: one warp per row, 16-byte vectors, persistent warps (grid-stride over
// rows) so that per-column reductions (d gamma, d beta, d bias) are accumulated in registers and leave
// the CTA as one atomicAdd per column.
#include "common.cuh"
#include "../../include/ttsb200.h"

namespace ttsb {

__device__ __forceinline__ float4 ld_bf16x4(const __nv_bfloat16* p) {
  const uint2 u = *reinterpret_cast<const uint2*>(p);
  const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.x));
  const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.y));
  return make_float4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ void st_bf16x4(__nv_bfloat16* dst, const float4& v) {
  uint2 u;
  u.x = pack_bf16(v.x, v.y);
  u.y = pack_bf16(v.z, v.w);
  *reinterpret_cast<uint2*>(dst) = u;
}

constexpr int kRowThreads = 256;
constexpr int kRowWarps = kRowThreads / 32;

static inline unsigned persistent_grid(int64_t rows) {
  const int64_t want = (rows + kRowWarps - 1) / kRowWarps;
  const int64_t cap = 148 * 4;
  return static_cast<unsigned>(want < cap ? (want > 0 ? want : 1) : cap);
}

// Block-level reduction of per-lane column partials: acc[NV*4] per lane covers columns (i*32+lane)*4..+3.
// smem: [kRowWarps][D].  Result is atomically added to out[D].
template <int NV>
__device__ __forceinline__ void block_col_atomic(float (&acc)[NV * 4], float* __restrict__ out, float* sbuf,
                                                 int lane, int warp) {
  constexpr int D = 128 * NV;
  __syncthreads();
#pragma unroll
  for (int i = 0; i < NV; ++i)
    *reinterpret_cast<float4*>(sbuf + warp * D + (i * 32 + lane) * 4) =
        make_float4(acc[4 * i], acc[4 * i + 1], acc[4 * i + 2], acc[4 * i + 3]);
  __syncthreads();
  for (int c = threadIdx.x; c < D; c += kRowThreads) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < kRowWarps; ++w) s += sbuf[w * D + c];
    atomicAdd(out + c, s);
  }
}

// ------------------------------------------------------------------------------------------------
// LayerNorm backward (autograd of F.layer_norm, Models/layers.py:31,35, Models/modules.py:87,
// Models/varianceadaptor.py:219,222, Models/encoder.py:112).
//   y = gamma * xhat + beta, xhat = (x - mean) * rstd            (statistics recomputed from the saved x)
//   g = dL/dy comes from ONE of: g_f32 (rows,D) | g_bf16 (rows,D) | g_row[row] * g_vec[:]  (the 256->1 linear
//       head of a VariancePredictor, varianceadaptor.py:223-229, masked rows give 0)
//   dx = rstd * (g*gamma - mean(g*gamma) - xhat * mean(g*gamma*xhat))  (+ add_f32)   ; relu_gate: dx = x > 0 ? dx : 0
//       (x is then the ReLU output that was normalised, varianceadaptor.py:218-219, and dx the gradient
//        w.r.t. the conv output)
//   dgamma += sum_rows g*xhat, dbeta += sum_rows g, dvec += sum_rows g_row*y, dscal += sum_rows g_row
// ------------------------------------------------------------------------------------------------
// HEAD = the predictor-head variant (g = g_row * g_vec, dvec / dscal outputs): its 36 extra registers per thread
// cost the plain variant (19 launches per training step) its second resident CTA, so it is a template switch.
// DROP (training with dropout > 0, its own instantiation so the plain variant keeps its register count):
//   dg: the LayerNorm OUTPUT was dropped in the forward (y' = y * m): g := g * m before anything else, and the head's
//       dvec uses y * m  (Models/modules.py:86-87 is followed by layers.py:40's dropout; varianceadaptor.py:219,222)
//   dx: the LayerNorm INPUT was the output of a dropout (x = x0 * m): dx := dx * m   — mode 1: both outputs;
//       mode 2: only dx_bf16, the copy that feeds the GEMM gradients of a dropped branch (layers.py:33), while
//       dx_f32 continues down the residual path unmasked
struct LnDrop {
  DropCfg dg, dx;
  int dx_mode;
};

template <int NV, bool HEAD, bool DROP = false>
__global__ void __launch_bounds__(kRowThreads, (HEAD || DROP) ? 1 : 2)
ln_bwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
              const float* __restrict__ g_f32, const __nv_bfloat16* __restrict__ g_bf16,
              const float* __restrict__ g_row, const float* __restrict__ g_vec,
              const uint8_t* __restrict__ row_mask, const float* __restrict__ add_f32, int relu_gate,
              float* __restrict__ dx_f32, __nv_bfloat16* __restrict__ dx_bf16, float* __restrict__ dgamma,
              float* __restrict__ dbeta, float* __restrict__ dvec, float* __restrict__ dscal, int64_t rows,
              float eps, const LnDrop drop = LnDrop()) {
  constexpr int D = 128 * NV;
  __shared__ float sbuf[kRowWarps * D];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t warp0 = static_cast<int64_t>(blockIdx.x) * kRowWarps + warp;
  const int64_t nwarps = static_cast<int64_t>(gridDim.x) * kRowWarps;
  float4 gm[NV], bt[NV], gv[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    gm[i] = __ldg(reinterpret_cast<const float4*>(gamma) + i * 32 + lane);
    bt[i] = (HEAD && beta) ? __ldg(reinterpret_cast<const float4*>(beta) + i * 32 + lane) : make_float4(0, 0, 0, 0);
    gv[i] = (HEAD && g_vec) ? __ldg(reinterpret_cast<const float4*>(g_vec) + i * 32 + lane) : make_float4(0, 0, 0, 0);
  }
  float a_dg[NV * 4], a_db[NV * 4], a_dv[NV * 4];
#pragma unroll
  for (int i = 0; i < NV * 4; ++i) a_dg[i] = a_db[i] = a_dv[i] = 0.f;
  float a_ds = 0.f;
  for (int64_t row = warp0; row < rows; row += nwarps) {
    float4 v[NV], g[NV];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      v[i] = __ldcs(reinterpret_cast<const float4*>(x + row * D) + i * 32 + lane);
      s += v[i].x + v[i].y + v[i].z + v[i].w;
    }
    float gr = 0.f;
    if (HEAD && g_row) {
      gr = __ldg(g_row + row);
      if (row_mask && row_mask[row] == 0) gr = 0.f;
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      if (g_f32) g[i] = __ldcs(reinterpret_cast<const float4*>(g_f32 + row * D) + i * 32 + lane);
      else if (g_bf16) g[i] = ld_bf16x4(g_bf16 + row * D + (i * 32 + lane) * 4);
      else g[i] = make_float4(gr * gv[i].x, gr * gv[i].y, gr * gv[i].z, gr * gv[i].w);
    }
    [[maybe_unused]] float4 mk[NV];
    if constexpr (DROP) {
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        mk[i] = make_float4(1.f, 1.f, 1.f, 1.f);
        if (drop.dg.thresh) {
          const uint64_t e0 = static_cast<uint64_t>(row) * D + (i * 32 + lane) * 4;
          mk[i] = make_float4(drop_scale(drop.dg, e0), drop_scale(drop.dg, e0 + 1), drop_scale(drop.dg, e0 + 2),
                              drop_scale(drop.dg, e0 + 3));
          g[i].x *= mk[i].x; g[i].y *= mk[i].y; g[i].z *= mk[i].z; g[i].w *= mk[i].w;
        }
      }
    }
    const float mean = warp_sum(s) * (1.f / D);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      v[i].x -= mean; v[i].y -= mean; v[i].z -= mean; v[i].w -= mean;
      q += v[i].x * v[i].x + v[i].y * v[i].y + v[i].z * v[i].z + v[i].w * v[i].w;
    }
    const float rstd = rsqrtf(warp_sum(q) * (1.f / D) + eps);
    float m1 = 0.f, m2 = 0.f;
    float4 xh[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      xh[i] = make_float4(v[i].x * rstd, v[i].y * rstd, v[i].z * rstd, v[i].w * rstd);
      const float4 d = make_float4(g[i].x * gm[i].x, g[i].y * gm[i].y, g[i].z * gm[i].z, g[i].w * gm[i].w);
      m1 += d.x + d.y + d.z + d.w;
      m2 += d.x * xh[i].x + d.y * xh[i].y + d.z * xh[i].z + d.w * xh[i].w;
      a_dg[4 * i] += g[i].x * xh[i].x; a_dg[4 * i + 1] += g[i].y * xh[i].y;
      a_dg[4 * i + 2] += g[i].z * xh[i].z; a_dg[4 * i + 3] += g[i].w * xh[i].w;
      a_db[4 * i] += g[i].x; a_db[4 * i + 1] += g[i].y; a_db[4 * i + 2] += g[i].z; a_db[4 * i + 3] += g[i].w;
      if (HEAD && g_row) {
        float4 y = make_float4(fmaf(gm[i].x, xh[i].x, bt[i].x), fmaf(gm[i].y, xh[i].y, bt[i].y),
                               fmaf(gm[i].z, xh[i].z, bt[i].z), fmaf(gm[i].w, xh[i].w, bt[i].w));
        if constexpr (DROP) { y.x *= mk[i].x; y.y *= mk[i].y; y.z *= mk[i].z; y.w *= mk[i].w; }
        a_dv[4 * i] += gr * y.x; a_dv[4 * i + 1] += gr * y.y; a_dv[4 * i + 2] += gr * y.z; a_dv[4 * i + 3] += gr * y.w;
      }
    }
    a_ds += gr;
    m1 = warp_sum(m1) * (1.f / D);
    m2 = warp_sum(m2) * (1.f / D);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      float4 d;
      d.x = rstd * (g[i].x * gm[i].x - m1 - xh[i].x * m2);
      d.y = rstd * (g[i].y * gm[i].y - m1 - xh[i].y * m2);
      d.z = rstd * (g[i].z * gm[i].z - m1 - xh[i].z * m2);
      d.w = rstd * (g[i].w * gm[i].w - m1 - xh[i].w * m2);
      if (add_f32) {
        const float4 a = __ldcs(reinterpret_cast<const float4*>(add_f32 + row * D) + i * 32 + lane);
        d.x += a.x; d.y += a.y; d.z += a.z; d.w += a.w;
      }
      if (relu_gate) {  // v = x - mean, so x > 0 <=> v > -mean
        if (!(v[i].x > -mean)) d.x = 0.f;
        if (!(v[i].y > -mean)) d.y = 0.f;
        if (!(v[i].z > -mean)) d.z = 0.f;
        if (!(v[i].w > -mean)) d.w = 0.f;
      }
      float4 dm = d;
      if constexpr (DROP) {
        if (drop.dx.thresh) {
          const uint64_t e0 = static_cast<uint64_t>(row) * D + (i * 32 + lane) * 4;
          dm.x *= drop_scale(drop.dx, e0); dm.y *= drop_scale(drop.dx, e0 + 1);
          dm.z *= drop_scale(drop.dx, e0 + 2); dm.w *= drop_scale(drop.dx, e0 + 3);
          if (drop.dx_mode == 1) d = dm;
        }
      }
      if (dx_f32) __stcs(reinterpret_cast<float4*>(dx_f32 + row * D) + i * 32 + lane, d);
      if (dx_bf16) st_bf16x4(dx_bf16 + row * D + (i * 32 + lane) * 4, dm);
    }
  }
  if (dgamma) block_col_atomic<NV>(a_dg, dgamma, sbuf, lane, warp);
  if (dbeta) block_col_atomic<NV>(a_db, dbeta, sbuf, lane, warp);
  if (HEAD && dvec) block_col_atomic<NV>(a_dv, dvec, sbuf, lane, warp);
  if (HEAD && dscal) {
    a_ds = warp_sum(a_ds);  // every lane of the warp carried the same gr: count it once
    if (lane == 0) atomicAdd(dscal, a_ds * (1.f / 32.f));
  }
}

// ------------------------------------------------------------------------------------------------
// Column sums (bias gradients): out[c] += sum_rows x[row, c]; x bf16 or fp32 with leading dimension ld.
// A thread owns one 16-byte chunk of columns (8 bf16 / 4 fp32) and walks rows; 256/chunks rows are read per
// pass so every warp issues full 16-byte coalesced loads; partials meet in smem, one atomicAdd per column per CTA.
// ------------------------------------------------------------------------------------------------
template <typename TIn>
__global__ void __launch_bounds__(256)
colsum_kernel(const TIn* __restrict__ x, int64_t ld, float* __restrict__ out, int64_t rows, int cols) {
  constexpr int VEC = 16 / sizeof(TIn);
  __shared__ float sbuf[256 * VEC];
  const int cpr = cols / VEC;                       // chunks per row
  const int cb0 = blockIdx.x * 256;                 // first chunk of this column block
  const int cpb = min(cpr - cb0, 256);              // chunks in this block
  const int rpp = 256 / cpb;                        // rows per pass
  const int c = threadIdx.x % cpb, r0 = threadIdx.x / cpb;
  float acc[VEC];
#pragma unroll
  for (int i = 0; i < VEC; ++i) acc[i] = 0.f;
  if (r0 < rpp) {
    const TIn* base = x + static_cast<int64_t>(cb0 + c) * VEC;
    for (int64_t r = static_cast<int64_t>(blockIdx.y) * rpp + r0; r < rows; r += static_cast<int64_t>(gridDim.y) * rpp) {
      const uint4 u = __ldcs(reinterpret_cast<const uint4*>(base + r * ld));
      if constexpr (sizeof(TIn) == 2) {
        const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 f = __bfloat1622float2(h2[i]);
          acc[2 * i] += f.x;
          acc[2 * i + 1] += f.y;
        }
      } else {
        const float* f = reinterpret_cast<const float*>(&u);
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[i] += f[i];
      }
    }
  }
#pragma unroll
  for (int i = 0; i < VEC; ++i) sbuf[threadIdx.x * VEC + i] = acc[i];
  __syncthreads();
  for (int i = threadIdx.x; i < cpb * VEC; i += 256) {
    float sum = 0.f;
    for (int r = 0; r < rpp; ++r) sum += sbuf[(r * cpb) * VEC + i];
    atomicAdd(out + cb0 * VEC + i, sum);
  }
}

// ------------------------------------------------------------------------------------------------
// BatchNorm1d in training mode + tanh (Models/postnets.py:71-73 with nn.BatchNorm1d batch statistics over
// all B*T positions, padding included).  C = 128*NV channels, x fp32 (rows, C) = conv output incl. bias.
//   stats[0][c] = sum x, stats[1][c] = sum x^2   (fp32 partials per warp, combined by atomics; the caller
//   all-reduces them across ranks for SyncBatchNorm and derives mean / rstd)
// ------------------------------------------------------------------------------------------------
template <int NV>
__global__ void __launch_bounds__(kRowThreads)
bn_stats_kernel(const float* __restrict__ x, float* __restrict__ stats, int64_t rows) {
  constexpr int D = 128 * NV;
  __shared__ float sbuf[kRowWarps * D];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float a1[NV * 4], a2[NV * 4];
#pragma unroll
  for (int i = 0; i < NV * 4; ++i) a1[i] = a2[i] = 0.f;
  for (int64_t row = static_cast<int64_t>(blockIdx.x) * kRowWarps + warp; row < rows;
       row += static_cast<int64_t>(gridDim.x) * kRowWarps) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const float4 v = __ldcs(reinterpret_cast<const float4*>(x + row * D) + i * 32 + lane);
      a1[4 * i] += v.x; a1[4 * i + 1] += v.y; a1[4 * i + 2] += v.z; a1[4 * i + 3] += v.w;
      a2[4 * i] = fmaf(v.x, v.x, a2[4 * i]); a2[4 * i + 1] = fmaf(v.y, v.y, a2[4 * i + 1]);
      a2[4 * i + 2] = fmaf(v.z, v.z, a2[4 * i + 2]); a2[4 * i + 3] = fmaf(v.w, v.w, a2[4 * i + 3]);
    }
  }
  block_col_atomic<NV>(a1, stats, sbuf, lane, warp);
  block_col_atomic<NV>(a2, stats + D, sbuf, lane, warp);
}

// h = tanh(gamma * (x - mean) * rstd + beta) -> bf16; ms = [mean | rstd] (2, C)
template <int NV>
__global__ void __launch_bounds__(kRowThreads)
bn_tanh_fwd_kernel(const float* __restrict__ x, const float* __restrict__ ms, const float* __restrict__ gamma,
                   const float* __restrict__ beta, __nv_bfloat16* __restrict__ h, int64_t rows, const DropCfg drop) {
  constexpr int D = 128 * NV;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float4 sc[NV], sh[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const float4 m = __ldg(reinterpret_cast<const float4*>(ms) + i * 32 + lane);
    const float4 r = __ldg(reinterpret_cast<const float4*>(ms + D) + i * 32 + lane);
    const float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + i * 32 + lane);
    const float4 b = __ldg(reinterpret_cast<const float4*>(beta) + i * 32 + lane);
    sc[i] = make_float4(g.x * r.x, g.y * r.y, g.z * r.z, g.w * r.w);
    sh[i] = make_float4(b.x - m.x * sc[i].x, b.y - m.y * sc[i].y, b.z - m.z * sc[i].z, b.w - m.w * sc[i].w);
  }
  for (int64_t row = static_cast<int64_t>(blockIdx.x) * kRowWarps + warp; row < rows;
       row += static_cast<int64_t>(gridDim.x) * kRowWarps) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const float4 v = __ldcs(reinterpret_cast<const float4*>(x + row * D) + i * 32 + lane);
      float4 o = make_float4(tanhf(fmaf(v.x, sc[i].x, sh[i].x)), tanhf(fmaf(v.y, sc[i].y, sh[i].y)),
                             tanhf(fmaf(v.z, sc[i].z, sh[i].z)), tanhf(fmaf(v.w, sc[i].w, sh[i].w)));
      if (drop.thresh) {  // Models/postnets.py:71-73: dropout(tanh(batch_norm(conv(x))))
        const uint64_t e0 = static_cast<uint64_t>(row) * D + (i * 32 + lane) * 4;
        o.x *= drop_scale(drop, e0); o.y *= drop_scale(drop, e0 + 1);
        o.z *= drop_scale(drop, e0 + 2); o.w *= drop_scale(drop, e0 + 3);
      }
      st_bf16x4(h + row * D + (i * 32 + lane) * 4, o);
    }
  }
}

// Backward of h = tanh(BN(x)): dt = dh * (1 - h^2).
//   APPLY == false: red[0][c] += sum dt, red[1][c] += sum dt * xhat          (= d beta, d gamma)
//   APPLY == true : dx = gamma * rstd * (dt - red[0]/M - xhat * red[1]/M) -> bf16     (M = inv_count^-1 rows, all ranks)
template <int NV, bool APPLY>
__global__ void __launch_bounds__(kRowThreads)
bn_tanh_bwd_kernel(const __nv_bfloat16* __restrict__ dh, const __nv_bfloat16* __restrict__ h,
                   const float* __restrict__ x, const float* __restrict__ ms, const float* __restrict__ gamma,
                   float* __restrict__ red, __nv_bfloat16* __restrict__ dx, int64_t rows, float inv_count,
                   const DropCfg drop) {
  constexpr int D = 128 * NV;
  __shared__ float sbuf[APPLY ? 1 : kRowWarps * D];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float4 mean[NV], rstd[NV], k0[NV], k1[NV], gr[NV];
  float a1[NV * 4], a2[NV * 4];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    mean[i] = __ldg(reinterpret_cast<const float4*>(ms) + i * 32 + lane);
    rstd[i] = __ldg(reinterpret_cast<const float4*>(ms + D) + i * 32 + lane);
    if (APPLY) {
      const float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + i * 32 + lane);
      const float4 r0 = __ldg(reinterpret_cast<const float4*>(red) + i * 32 + lane);
      const float4 r1 = __ldg(reinterpret_cast<const float4*>(red + D) + i * 32 + lane);
      gr[i] = make_float4(g.x * rstd[i].x, g.y * rstd[i].y, g.z * rstd[i].z, g.w * rstd[i].w);
      k0[i] = make_float4(r0.x * inv_count, r0.y * inv_count, r0.z * inv_count, r0.w * inv_count);
      k1[i] = make_float4(r1.x * inv_count, r1.y * inv_count, r1.z * inv_count, r1.w * inv_count);
    }
  }
#pragma unroll
  for (int i = 0; i < NV * 4; ++i) a1[i] = a2[i] = 0.f;
  for (int64_t row = static_cast<int64_t>(blockIdx.x) * kRowWarps + warp; row < rows;
       row += static_cast<int64_t>(gridDim.x) * kRowWarps) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int64_t off = row * D + (i * 32 + lane) * 4;
      float4 g = ld_bf16x4(dh + off);
      float4 t = ld_bf16x4(h + off);
      if (drop.thresh) {
        // h = tanh(.) * m was saved: kept elements give tanh back as h * (1 - p), dropped ones carry no gradient
        const uint64_t e0 = static_cast<uint64_t>(off);
        const float4 m = make_float4(drop_scale(drop, e0), drop_scale(drop, e0 + 1), drop_scale(drop, e0 + 2),
                                     drop_scale(drop, e0 + 3));
        const float keep = 1.f / drop.inv_keep;
        g.x *= m.x; g.y *= m.y; g.z *= m.z; g.w *= m.w;
        t.x *= keep; t.y *= keep; t.z *= keep; t.w *= keep;
      }
      const float4 v = __ldcs(reinterpret_cast<const float4*>(x + row * D) + i * 32 + lane);
      const float4 dt = make_float4(g.x * (1.f - t.x * t.x), g.y * (1.f - t.y * t.y), g.z * (1.f - t.z * t.z),
                                    g.w * (1.f - t.w * t.w));
      const float4 xh = make_float4((v.x - mean[i].x) * rstd[i].x, (v.y - mean[i].y) * rstd[i].y,
                                    (v.z - mean[i].z) * rstd[i].z, (v.w - mean[i].w) * rstd[i].w);
      if (APPLY) {
        const float4 o = make_float4(gr[i].x * (dt.x - k0[i].x - xh.x * k1[i].x), gr[i].y * (dt.y - k0[i].y - xh.y * k1[i].y),
                                     gr[i].z * (dt.z - k0[i].z - xh.z * k1[i].z), gr[i].w * (dt.w - k0[i].w - xh.w * k1[i].w));
        st_bf16x4(dx + off, o);
      } else {
        a1[4 * i] += dt.x; a1[4 * i + 1] += dt.y; a1[4 * i + 2] += dt.z; a1[4 * i + 3] += dt.w;
        a2[4 * i] = fmaf(dt.x, xh.x, a2[4 * i]); a2[4 * i + 1] = fmaf(dt.y, xh.y, a2[4 * i + 1]);
        a2[4 * i + 2] = fmaf(dt.z, xh.z, a2[4 * i + 2]); a2[4 * i + 3] = fmaf(dt.w, xh.w, a2[4 * i + 3]);
      }
    }
  }
  if (!APPLY) {
    block_col_atomic<NV>(a1, red, sbuf, lane, warp);
    block_col_atomic<NV>(a2, red + D, sbuf, lane, warp);
  }
}


// ------------------------------------------------------------------------------------------------
// LengthRegulator backward (autograd of expand + cat, Models/varianceadaptor.py:167-177):
//   dx[b, l, :] = sum_{t in [cum[l]-d[l], cum[l]) , t < Tmax} g[b, t, :]        one warp per phoneme
// ------------------------------------------------------------------------------------------------
template <int NV>
__global__ void __launch_bounds__(kRowThreads)
lr_bwd_kernel(const float* __restrict__ g, const int32_t* __restrict__ cum, float* __restrict__ dx,
              const float* __restrict__ add_f32, int B, int L, int Tmax) {
  constexpr int D = 128 * NV;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t idx = static_cast<int64_t>(blockIdx.x) * kRowWarps + warp;
  if (idx >= static_cast<int64_t>(B) * L) return;
  const int b = static_cast<int>(idx / L), l = static_cast<int>(idx - static_cast<int64_t>(b) * L);
  int lo = l > 0 ? cum[idx - 1] : 0, hi = cum[idx];
  if (hi > Tmax) hi = Tmax;
  float4 acc[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i)
    acc[i] = add_f32 ? __ldg(reinterpret_cast<const float4*>(add_f32 + idx * D) + i * 32 + lane) : make_float4(0, 0, 0, 0);
  for (int t = lo; t < hi; ++t) {
    const float4* gp = reinterpret_cast<const float4*>(g + (static_cast<int64_t>(b) * Tmax + t) * D);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const float4 v = __ldcs(gp + i * 32 + lane);
      acc[i].x += v.x; acc[i].y += v.y; acc[i].z += v.z; acc[i].w += v.w;
    }
  }
#pragma unroll
  for (int i = 0; i < NV; ++i) *(reinterpret_cast<float4*>(dx + idx * D) + i * 32 + lane) = acc[i];
}

// ------------------------------------------------------------------------------------------------
// Embedding backward (nn.Embedding weight gradient: encoder.embed with padding_idx=0, encoder.py:55;
// pitch / energy embeddings, varianceadaptor.py:123-126): table[idx[row], :] += g[row, :]   (fp32 atomics)
// Also the positional-encoding scale: dalpha += sum g[row,:] . pe[row % T, :]  (modules.py:110)
// ------------------------------------------------------------------------------------------------
template <int NV>
__global__ void __launch_bounds__(kRowThreads)
embed_bwd_kernel(const float* __restrict__ g, const int64_t* __restrict__ idx64, const int32_t* __restrict__ idx32,
                 float* __restrict__ table, int skip_idx, const float* __restrict__ pe, float* __restrict__ dalpha,
                 int64_t rows, int T, int V) {
  constexpr int D = 128 * NV;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float dot = 0.f;
  for (int64_t row = static_cast<int64_t>(blockIdx.x) * kRowWarps + warp; row < rows;
       row += static_cast<int64_t>(gridDim.x) * kRowWarps) {
    int64_t id = -1;
    if (idx64) id = idx64[row];
    else if (idx32) id = idx32[row];
    const bool scatter = table != nullptr && id >= 0 && id < V && id != skip_idx;
    const float4* gp = reinterpret_cast<const float4*>(g + row * D);
    const float4* pp = pe ? reinterpret_cast<const float4*>(pe + static_cast<int64_t>(row % T) * D) : nullptr;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const float4 v = __ldcs(gp + i * 32 + lane);
      if (scatter) {
        float* tp = table + id * D + (i * 32 + lane) * 4;
        atomicAdd(tp, v.x); atomicAdd(tp + 1, v.y); atomicAdd(tp + 2, v.z); atomicAdd(tp + 3, v.w);
      }
      if (pp) {
        const float4 q = __ldg(pp + i * 32 + lane);
        dot += v.x * q.x + v.y * q.y + v.z * q.z + v.w * q.w;
      }
    }
  }
  if (dalpha) {
    dot = warp_sum(dot);
    if (lane == 0) atomicAdd(dalpha, dot);
  }
}

// Embedding weight gradient for small vocabularies (V <= 256): a CTA privatises a 32-column slab of the table
// in shared memory (V x 32 fp32 <= 32 KB), walks its share of the rows with one warp per row (128-byte coalesced
// reads), accumulates with shared-memory atomics and flushes the slab once with global atomics.
// grid = (D/32 slabs, row splits).
__global__ void __launch_bounds__(kRowThreads)
embed_scatter_kernel(const float* __restrict__ g, const int64_t* __restrict__ idx64, const int32_t* __restrict__ idx32,
                     float* __restrict__ table, int skip_idx, int64_t rows, int D, int V) {
  extern __shared__ float stab[];  // [V][32]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int col = blockIdx.x * 32 + lane;
  for (int i = threadIdx.x; i < V * 32; i += kRowThreads) stab[i] = 0.f;
  __syncthreads();
  for (int64_t row = static_cast<int64_t>(blockIdx.y) * kRowWarps + warp; row < rows;
       row += static_cast<int64_t>(gridDim.y) * kRowWarps) {
    const int64_t id = idx64 ? idx64[row] : static_cast<int64_t>(idx32[row]);
    if (id < 0 || id >= V || id == skip_idx) continue;
    atomicAdd(&stab[id * 32 + lane], __ldcs(g + row * D + col));
  }
  __syncthreads();
  for (int i = threadIdx.x; i < V * 32; i += kRowThreads) {
    const float v = stab[i];
    if (v != 0.f) atomicAdd(table + static_cast<int64_t>(i >> 5) * D + blockIdx.x * 32 + (i & 31), v);
  }
}

// ------------------------------------------------------------------------------------------------
// L1 loss, mean reduction, with its gradient (nn.L1Loss, train_fastspeech2.py:212-217,245-259):
//   loss += scale * sum |pred - target| ; grad = scale * sign(pred - target)       (scale = 1/numel)
// log_target: target := log(target + 1) (duration loss, train_fastspeech2.py:252).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
l1_loss_kernel(const float* __restrict__ pred, const float* __restrict__ target, float* __restrict__ grad,
               __nv_bfloat16* __restrict__ grad_bf16, float* __restrict__ loss, int64_t n, float scale,
               int log_target) {
  float acc = 0.f;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * 256 + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * 256) {
    float t = target[i];
    if (log_target) t = logf(t + 1.f);
    const float d = pred[i] - t;
    acc += fabsf(d);
    const float gval = d > 0.f ? scale : (d < 0.f ? -scale : 0.f);
    if (grad) grad[i] = gval;
    if (grad_bf16) grad_bf16[i] = __float2bfloat16(gval);
  }
  acc = warp_sum(acc);
  __shared__ float sw[8];
  if ((threadIdx.x & 31) == 0) sw[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int w = 0; w < 8; ++w) s += sw[w];
    atomicAdd(loss, s * scale);
  }
}

// ------------------------------------------------------------------------------------------------
// F.binary_cross_entropy_with_logits(x, y, reduction='mean', pos_weight=pw) (train.py:217, the stop-token loss of
// Transformer-TTS): l = (1 - y) x + (1 + (pw - 1) y) softplus(-x);  loss += scale * sum l;
// grad = scale * ((1 - y) - (1 + (pw - 1) y) sigmoid(-x)).  softplus(-x) = max(-x, 0) + log1p(exp(-|x|)) as torch.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
bce_logits_loss_kernel(const float* __restrict__ x, const float* __restrict__ y, float* __restrict__ grad,
                       float* __restrict__ loss, int64_t n, float scale, float pos_weight) {
  float acc = 0.f;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * 256 + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * 256) {
    const float xv = x[i], yv = y[i];
    const float w = 1.f + (pos_weight - 1.f) * yv;
    const float sp = fmaxf(-xv, 0.f) + log1pf(expf(-fabsf(xv)));   // softplus(-x)
    acc += (1.f - yv) * xv + w * sp;
    if (grad) grad[i] = scale * ((1.f - yv) - w / (1.f + expf(xv)));   // sigmoid(-x) = 1 / (1 + e^x)
  }
  acc = warp_sum(acc);
  __shared__ float sw[8];
  if ((threadIdx.x & 31) == 0) sw[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int w = 0; w < 8; ++w) s += sw[w];
    atomicAdd(loss, s * scale);
  }
}

// out_bf16[i] = gate_bf16[i] > 0 ? g[i] : 0   (backward of a ReLU whose output is `gate`; Models/prenets.py:36-37)
__global__ void __launch_bounds__(256)
relu_gate_kernel(const float* __restrict__ g, const __nv_bfloat16* __restrict__ gate, __nv_bfloat16* __restrict__ out, int64_t n) {
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * 256 + threadIdx.x; i < n; i += static_cast<int64_t>(gridDim.x) * 256)
    out[i] = __float2bfloat16(__bfloat162float(gate[i]) > 0.f ? g[i] : 0.f);
}

// ------------------------------------------------------------------------------------------------
// dgrad operands of a whole backward pass in ONE launch: for every table entry, dst[taps-1-j][c][n] = src[j][n][c]
// (bf16; the forward operand (taps, N, C) of a Linear / Conv1d -> the transposed, tap-flipped operand the input-gradient
// GEMM reads).  One CTA per 64 x 64 tile through shared memory; entries carry the index of their first tile.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
weight_transpose_kernel(const ttsb200_wt_entry* __restrict__ tab, int n_entries) {
  __shared__ __nv_bfloat16 tile[64][66];
  __shared__ ttsb200_wt_entry ent;
  if (threadIdx.x == 0) {
    int lo = 0, hi = n_entries - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (tab[mid].tile0 <= static_cast<int>(blockIdx.x)) lo = mid; else hi = mid - 1;
    }
    ent = tab[lo];
  }
  __syncthreads();
  const int N = ent.N, C = ent.C;
  const int tn = (N + 63) / 64, tc = (C + 63) / 64;
  const int local = blockIdx.x - ent.tile0;
  const int j = local / (tn * tc);
  const int r = local - j * (tn * tc);
  const int n0 = (r / tc) * 64, c0 = (r % tc) * 64;
  const __nv_bfloat16* src = static_cast<const __nv_bfloat16*>(ent.src) + static_cast<int64_t>(j) * N * C;
  __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(ent.dst) + static_cast<int64_t>(ent.taps - 1 - j) * C * N;
  const int ty = threadIdx.x >> 5, tx = threadIdx.x & 31;
  const __nv_bfloat16 zero = __float2bfloat16(0.f);
#pragma unroll
  for (int rr = 0; rr < 8; ++rr) {
    const int row = ty + rr * 8, n = n0 + row, c = c0 + 2 * tx;
    tile[row][2 * tx] = (n < N && c < C) ? src[static_cast<int64_t>(n) * C + c] : zero;
    tile[row][2 * tx + 1] = (n < N && c + 1 < C) ? src[static_cast<int64_t>(n) * C + c + 1] : zero;
  }
  __syncthreads();
#pragma unroll
  for (int rr = 0; rr < 8; ++rr) {
    const int crow = ty + rr * 8, c = c0 + crow, n = n0 + 2 * tx;
    if (c < C) {
      if (n < N) dst[static_cast<int64_t>(c) * N + n] = tile[2 * tx][crow];
      if (n + 1 < N) dst[static_cast<int64_t>(c) * N + n + 1] = tile[2 * tx + 1][crow];
    }
  }
}

// ------------------------------------------------------------------------------------------------
// delta[b,h,t] = sum_d dctx[b,t,h*Dh+d] * ctx[b,t,h*Dh+d]   (softmax backward row term of attention)
// one warp per (b,t); lanes own 4-column chunks; Dh = 96 = 24 chunks, so chunk c belongs to head c/24.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kRowThreads)
attn_delta_kernel(const __nv_bfloat16* __restrict__ dctx, const __nv_bfloat16* __restrict__ ctx,
                  float* __restrict__ delta, int B, int T, int H) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t row = static_cast<int64_t>(blockIdx.x) * kRowWarps + warp;
  if (row >= static_cast<int64_t>(B) * T) return;
  const int b = static_cast<int>(row / T), t = static_cast<int>(row - static_cast<int64_t>(b) * T);
  const int D = H * 96;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int c = lane; c < D / 4; c += 32) {
    const float4 a = ld_bf16x4(dctx + row * D + c * 4);
    const float4 o = ld_bf16x4(ctx + row * D + c * 4);
    const float v = a.x * o.x + a.y * o.y + a.z * o.z + a.w * o.w;
    const int hh = c / 24;
#pragma unroll
    for (int k = 0; k < 4; ++k) acc[k] += hh == k ? v : 0.f;
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float s = warp_sum(acc[k]);
    if (lane == 0 && k < H) delta[(static_cast<int64_t>(b) * H + k) * T + t] = s;
  }
}

// ------------------------------------------------------------------------------------------------
// Optimizer (train_fastspeech2.py:117-120,304-315; utils/utils.py:204-215): global-norm clip + Adam over
// ONE flat fp32 parameter buffer (all 289 tensors are views into it).
//   sumsq_kernel: out += sum g^2
//   adam_kernel : g *= gscale * min(1, clip / (sqrt(sumsq)*gscale + 1e-6)) ; Adam(beta1, beta2, eps) step
//                 with the bias corrections folded into step_size / bc2 (torch.optim.Adam semantics)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
sumsq_kernel(const float* __restrict__ g, float* __restrict__ out, int64_t n) {
  float acc = 0.f;
  const int64_t n4 = n >> 2;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * 256 + threadIdx.x; i < n4;
       i += static_cast<int64_t>(gridDim.x) * 256) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(g) + i);
    acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const float v = g[(n4 << 2) + threadIdx.x];
    acc += v * v;
  }
  acc = warp_sum(acc);
  __shared__ float sw[8];
  if ((threadIdx.x & 31) == 0) sw[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int w = 0; w < 8; ++w) s += sw[w];
    atomicAdd(out, s);
  }
}

// MODE 0: torch.optim.Adam (denominator sqrt(v)/sqrt(bc2) + eps, step lr/bc1)
// MODE 1: radam.RAdam rectified branch (radam.py:83-88): p -= step * m / (sqrt(v) + eps), step = lr*step_size
// MODE 2: radam.RAdam degenerated-to-SGD branch (radam.py:89-93, N_sma < 5): p -= step * m
template <int MODE>
__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
            const float* __restrict__ sumsq, int64_t n, float gscale, float clip, float step, float beta1,
            float beta2, float eps, float bc2_sqrt) {
  float coef = gscale;
  if (clip > 0.f) {
    const float norm = sqrtf(__ldg(sumsq)) * gscale;
    const float c = clip / (norm + 1e-6f);  // torch.nn.utils.clip_grad_norm_
    coef *= c < 1.f ? c : 1.f;
  }
  const float inv_bc2 = 1.f / bc2_sqrt;
  auto upd = [&](float& pi, float gi, float& mi, float& vi) {
    gi *= coef;
    mi = beta1 * mi + (1.f - beta1) * gi;
    vi = beta2 * vi + (1.f - beta2) * gi * gi;
    if (MODE == 0) pi -= step * mi / (sqrtf(vi) * inv_bc2 + eps);
    else if (MODE == 1) pi -= step * mi / (sqrtf(vi) + eps);
    else pi -= step * mi;
  };
  const int64_t n4 = n >> 2;  // the flat buffers are 16-byte aligned (every tensor starts on a multiple of 4)
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * 256 + threadIdx.x; i < n4;
       i += static_cast<int64_t>(gridDim.x) * 256) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    const float4 gg = __ldcs(reinterpret_cast<const float4*>(g) + i);
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    upd(pp.x, gg.x, mm.x, vv.x);
    upd(pp.y, gg.y, mm.y, vv.y);
    upd(pp.z, gg.z, mm.z, vv.z);
    upd(pp.w, gg.w, mm.w, vv.w);
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const int64_t i = (n4 << 2) + threadIdx.x;
    upd(p[i], g[i], m[i], v[i]);
  }
}

}  // namespace ttsb

using namespace ttsb;

#define TTSB_LAUNCH_CHECK() TTSB_CHECK_CUDA(cudaGetLastError())

extern "C" int ttsb200_layernorm_bwd(const float* x, const float* gamma, const float* beta, const float* g_f32,
                                     const void* g_bf16, const float* g_row, const float* g_vec,
                                     const uint8_t* row_mask, const float* add_f32, int32_t relu_gate,
                                     float* dx_f32, void* dx_bf16, float* dgamma, float* dbeta, float* dvec,
                                     float* dscal, int64_t rows, int32_t D, float eps, ttsb200_stream_t s) {
  return ttsb200_layernorm_bwd_drop(x, gamma, beta, g_f32, g_bf16, g_row, g_vec, row_mask, add_f32, relu_gate, dx_f32,
                                    dx_bf16, dgamma, dbeta, dvec, dscal, rows, D, eps, 0.f, 0, -1, -1, 0, s);
}

extern "C" int ttsb200_layernorm_bwd_drop(const float* x, const float* gamma, const float* beta, const float* g_f32,
                                          const void* g_bf16, const float* g_row, const float* g_vec,
                                          const uint8_t* row_mask, const float* add_f32, int32_t relu_gate,
                                          float* dx_f32, void* dx_bf16, float* dgamma, float* dbeta, float* dvec,
                                          float* dscal, int64_t rows, int32_t D, float eps, float drop_p,
                                          uint64_t drop_seed, int64_t site_g, int64_t site_dx, int32_t dx_mode,
                                          ttsb200_stream_t s) {
  TTSB_CHECK_ARG(x && gamma && rows > 0, "layernorm_bwd: bad args");
  TTSB_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f && dx_mode >= 0 && dx_mode <= 2, "layernorm_bwd: bad dropout arguments");
  LnDrop drop;
  drop.dg = make_drop(drop_p, drop_seed, site_g);
  drop.dx = make_drop(drop_p, drop_seed, dx_mode ? site_dx : -1);
  drop.dx_mode = dx_mode;
  const bool dropping = drop.dg.thresh != 0 || drop.dx.thresh != 0;
  TTSB_CHECK_ARG((g_f32 != nullptr) + (g_bf16 != nullptr) + (g_row != nullptr) == 1,
                 "layernorm_bwd: exactly one gradient source");
  TTSB_CHECK_ARG(!g_row || g_vec, "layernorm_bwd: g_row needs g_vec");
  cudaStream_t st = static_cast<cudaStream_t>(s);
  const unsigned grid = persistent_grid(rows);
  auto gb = static_cast<const __nv_bfloat16*>(g_bf16);
  auto db = static_cast<__nv_bfloat16*>(dx_bf16);
  const bool head = g_row != nullptr || g_vec != nullptr || dvec != nullptr || dscal != nullptr;
#define TTSB_LN_BWD(NV, HEAD, DROP)                                                                                 \
  ln_bwd_kernel<NV, HEAD, DROP><<<grid, kRowThreads, 0, st>>>(x, gamma, beta, g_f32, gb, g_row, g_vec, row_mask,       \
                                                              add_f32, relu_gate, dx_f32, db, dgamma, dbeta, dvec,      \
                                                              dscal, rows, eps, drop)
  if (D == 384) {
    if (dropping) { if (head) TTSB_LN_BWD(3, true, true); else TTSB_LN_BWD(3, false, true); }
    else if (head) TTSB_LN_BWD(3, true, false);
    else TTSB_LN_BWD(3, false, false);
  } else if (D == 256) {
    if (dropping) { if (head) TTSB_LN_BWD(2, true, true); else TTSB_LN_BWD(2, false, true); }
    else if (head) TTSB_LN_BWD(2, true, false);
    else TTSB_LN_BWD(2, false, false);
  }
#undef TTSB_LN_BWD
  else {
    set_last_error("layernorm_bwd: D=%d unsupported (384 or 256)", D);
    return TTSB_ERR_ARG;
  }
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

extern "C" int ttsb200_colsum(const void* x, int32_t is_bf16, int64_t ld, float* out, int64_t rows, int32_t cols,
                              ttsb200_stream_t s) {
  const int vec = is_bf16 ? 8 : 4;
  TTSB_CHECK_ARG(x && out && rows > 0 && cols > 0 && cols % vec == 0 && ld % vec == 0 &&
                     (reinterpret_cast<uintptr_t>(x) & 15) == 0,
                 "colsum: cols and ld must be multiples of %d, x 16-byte aligned", vec);
  const int cpr = cols / vec;
  const int cpb = cpr < 256 ? cpr : 256;
  const int rpp = 256 / cpb;
  int64_t gy = (rows + rpp * 8 - 1) / (rpp * 8);
  if (gy > 592) gy = 592;
  if (gy < 1) gy = 1;
  dim3 grid((cpr + 255) / 256, static_cast<unsigned>(gy));
  if (is_bf16)
    colsum_kernel<__nv_bfloat16><<<grid, 256, 0, static_cast<cudaStream_t>(s)>>>(
        static_cast<const __nv_bfloat16*>(x), ld, out, rows, cols);
  else
    colsum_kernel<float><<<grid, 256, 0, static_cast<cudaStream_t>(s)>>>(static_cast<const float*>(x), ld, out,
                                                                         rows, cols);
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

extern "C" int ttsb200_bn_stats(const float* x, float* stats, int64_t rows, int32_t C, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(x && stats && rows > 0 && C == 384, "bn_stats: bad args (C must be 384)");
  bn_stats_kernel<3><<<persistent_grid(rows), kRowThreads, 0, static_cast<cudaStream_t>(s)>>>(x, stats, rows);
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

extern "C" int ttsb200_bn_tanh_fwd(const float* x, const float* mean_rstd, const float* gamma, const float* beta,
                                   void* h_bf16, int64_t rows, int32_t C, float drop_p, uint64_t drop_seed,
                                   int64_t drop_site, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(x && mean_rstd && gamma && beta && h_bf16 && rows > 0 && C == 384, "bn_tanh_fwd: bad args");
  TTSB_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f, "bn_tanh_fwd: dropout probability must be in [0, 1)");
  bn_tanh_fwd_kernel<3><<<persistent_grid(rows), kRowThreads, 0, static_cast<cudaStream_t>(s)>>>(
      x, mean_rstd, gamma, beta, static_cast<__nv_bfloat16*>(h_bf16), rows, make_drop(drop_p, drop_seed, drop_site));
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

extern "C" int ttsb200_bn_tanh_bwd(const void* dh_bf16, const void* h_bf16, const float* x, const float* mean_rstd,
                                   const float* gamma, float* red, void* dx_bf16, int64_t rows, int32_t C,
                                   float inv_count, int32_t apply, float drop_p, uint64_t drop_seed,
                                   int64_t drop_site, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(dh_bf16 && h_bf16 && x && mean_rstd && red && rows > 0 && C == 384, "bn_tanh_bwd: bad args");
  TTSB_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f, "bn_tanh_bwd: dropout probability must be in [0, 1)");
  const DropCfg drop = make_drop(drop_p, drop_seed, drop_site);
  auto dh = static_cast<const __nv_bfloat16*>(dh_bf16);
  auto h = static_cast<const __nv_bfloat16*>(h_bf16);
  cudaStream_t st = static_cast<cudaStream_t>(s);
  if (apply) {
    TTSB_CHECK_ARG(gamma && dx_bf16, "bn_tanh_bwd: apply needs gamma and dx");
    bn_tanh_bwd_kernel<3, true><<<persistent_grid(rows), kRowThreads, 0, st>>>(
        dh, h, x, mean_rstd, gamma, red, static_cast<__nv_bfloat16*>(dx_bf16), rows, inv_count, drop);
  } else {
    bn_tanh_bwd_kernel<3, false><<<persistent_grid(rows), kRowThreads, 0, st>>>(dh, h, x, mean_rstd, gamma, red,
                                                                                nullptr, rows, inv_count, drop);
  }
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

namespace ttsb {
// ------------------------------------------------------------------------------------------------
// Standalone dropout over a dense buffer (Models/modules.py:111 on embedding + positional encoding, and the same mask
// on the gradient in the backward pass): dst[i] = src[i] * m(i).  dtype 0 = fp32, 1 = bf16; dst may alias src.
// ttsb200_dropout_mask writes m itself (0 or 1 / (1 - p)) for tests.
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) dropout_kernel(const T* __restrict__ src, T* __restrict__ dst, int64_t n,
                                                      const DropCfg drop) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float m = drop_scale(drop, static_cast<uint64_t>(i));
  if constexpr (sizeof(T) == 4) dst[i] = src[i] * m;
  else dst[i] = __float2bfloat16_rn(__bfloat162float(src[i]) * m);
}
__global__ void __launch_bounds__(256) dropout_mask_kernel(float* __restrict__ out, int64_t n, const DropCfg drop) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) out[i] = drop.thresh ? drop_scale(drop, static_cast<uint64_t>(i)) : 1.f;
}

}  // namespace ttsb
extern "C" int ttsb200_dropout(const void* src, void* dst, int32_t is_bf16, int64_t n, float p, uint64_t seed,
                               int64_t site, ttsb200_stream_t s) {
  using namespace ttsb;
  TTSB_CHECK_ARG(src && dst && n > 0 && p > 0.f && p < 1.f && site >= 0, "dropout: bad args (0 < p < 1)");
  const DropCfg d = make_drop(p, seed, site);
  const unsigned grid = static_cast<unsigned>((n + 255) / 256);
  if (is_bf16)
    dropout_kernel<__nv_bfloat16><<<grid, 256, 0, static_cast<cudaStream_t>(s)>>>(
        static_cast<const __nv_bfloat16*>(src), static_cast<__nv_bfloat16*>(dst), n, d);
  else
    dropout_kernel<float><<<grid, 256, 0, static_cast<cudaStream_t>(s)>>>(static_cast<const float*>(src),
                                                                         static_cast<float*>(dst), n, d);
  TTSB_CHECK_CUDA(cudaGetLastError());
  return TTSB_OK;
}
extern "C" int ttsb200_dropout_mask(float* out, int64_t n, float p, uint64_t seed, int64_t site, ttsb200_stream_t s) {
  using namespace ttsb;
  TTSB_CHECK_ARG(out && n > 0 && p >= 0.f && p < 1.f, "dropout_mask: bad args");
  dropout_mask_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(s)>>>(
      out, n, make_drop(p, seed, site));
  TTSB_CHECK_CUDA(cudaGetLastError());
  return TTSB_OK;
}
extern "C" int ttsb200_lr_bwd(const float* g, const int32_t* cum, float* dx, const float* add_f32, int32_t B,
                              int32_t L, int32_t Tmax, int32_t D, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(g && cum && dx && B > 0 && L > 0 && Tmax > 0 && D == 384, "lr_bwd: bad args (D must be 384)");
  const int64_t n = static_cast<int64_t>(B) * L;
  lr_bwd_kernel<3><<<static_cast<unsigned>((n + kRowWarps - 1) / kRowWarps), kRowThreads, 0,
                     static_cast<cudaStream_t>(s)>>>(g, cum, dx, add_f32, B, L, Tmax);
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

extern "C" int ttsb200_embed_bwd(const float* g, const int64_t* idx64, const int32_t* idx32, float* table,
                                 int32_t skip_idx, const float* pe, float* dalpha, int64_t rows, int32_t T,
                                 int32_t D, int32_t V, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(g && rows > 0 && D == 384 && T > 0, "embed_bwd: bad args (D must be 384)");
  TTSB_CHECK_ARG((pe == nullptr) == (dalpha == nullptr), "embed_bwd: pe and dalpha go together");
  TTSB_CHECK_ARG(table == nullptr || idx64 || idx32, "embed_bwd: a table needs indices");
  cudaStream_t st = static_cast<cudaStream_t>(s);
  float* table_generic = table;
  if (table && V <= 256) {
    // small vocabulary: shared-memory privatised scatter (one pass over g per table)
    TTSB_SMEM_OPT_IN(embed_scatter_kernel, 256 * 32 * 4);
    int64_t gy = (rows + kRowWarps * 16 - 1) / (kRowWarps * 16);
    if (gy > 48) gy = 48;
    if (gy < 1) gy = 1;
    dim3 grid(D / 32, static_cast<unsigned>(gy));
    embed_scatter_kernel<<<grid, kRowThreads, V * 32 * sizeof(float), st>>>(g, idx64, idx32, table, skip_idx, rows, D, V);
    TTSB_LAUNCH_CHECK();
    table_generic = nullptr;
  }
  if (table_generic || dalpha) {
    embed_bwd_kernel<3><<<persistent_grid(rows), kRowThreads, 0, st>>>(g, idx64, idx32, table_generic, skip_idx, pe,
                                                                       dalpha, rows, T, V);
    TTSB_LAUNCH_CHECK();
  }
  return TTSB_OK;
}

extern "C" int ttsb200_l1_loss(const float* pred, const float* target, float* grad, void* grad_bf16, float* loss,
                               int64_t n, float scale, int32_t log_target, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(pred && target && loss && n > 0, "l1_loss: bad args");
  const int64_t want = (n + 255) / 256;
  l1_loss_kernel<<<static_cast<unsigned>(want < 1184 ? want : 1184), 256, 0, static_cast<cudaStream_t>(s)>>>(
      pred, target, grad, static_cast<__nv_bfloat16*>(grad_bf16), loss, n, scale, log_target);
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

extern "C" int ttsb200_bce_logits_loss(const float* x, const float* y, float* grad, float* loss, int64_t n, float scale,
                                       float pos_weight, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(x && y && loss && n > 0, "bce_logits_loss: bad args");
  const int64_t want = (n + 255) / 256;
  bce_logits_loss_kernel<<<static_cast<unsigned>(want < 1184 ? want : 1184), 256, 0, static_cast<cudaStream_t>(s)>>>(
      x, y, grad, loss, n, scale, pos_weight);
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

extern "C" int ttsb200_relu_gate(const float* g, const void* gate_bf16, void* out_bf16, int64_t n, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(g && gate_bf16 && out_bf16 && n > 0, "relu_gate: bad args");
  const int64_t want = (n + 255) / 256;
  relu_gate_kernel<<<static_cast<unsigned>(want < 1184 ? want : 1184), 256, 0, static_cast<cudaStream_t>(s)>>>(
      g, static_cast<const __nv_bfloat16*>(gate_bf16), static_cast<__nv_bfloat16*>(out_bf16), n);
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

extern "C" int ttsb200_weight_transpose(const ttsb200_wt_entry* table_dev, int32_t n_entries, int32_t total_tiles,
                                        ttsb200_stream_t s) {
  TTSB_CHECK_ARG(table_dev && n_entries > 0 && total_tiles > 0, "weight_transpose: bad args");
  weight_transpose_kernel<<<static_cast<unsigned>(total_tiles), 256, 0, static_cast<cudaStream_t>(s)>>>(table_dev, n_entries);
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

extern "C" int ttsb200_attn_delta(const void* dctx, const void* ctx, float* delta, int32_t B, int32_t T, int32_t H,
                                  int32_t Dh, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(dctx && ctx && delta && B > 0 && T > 0 && H > 0 && H <= 4 && Dh == 96, "attn_delta: bad args");
  const int64_t rows = static_cast<int64_t>(B) * T;
  attn_delta_kernel<<<static_cast<unsigned>((rows + kRowWarps - 1) / kRowWarps), kRowThreads, 0,
                      static_cast<cudaStream_t>(s)>>>(static_cast<const __nv_bfloat16*>(dctx),
                                                      static_cast<const __nv_bfloat16*>(ctx), delta, B, T, H);
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

extern "C" int ttsb200_sumsq(const float* g, float* out, int64_t n, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(g && out && n > 0 && (reinterpret_cast<uintptr_t>(g) & 15) == 0, "sumsq: bad args");
  const int64_t want = (n / 4 + 255) / 256;
  sumsq_kernel<<<static_cast<unsigned>(want < 1184 ? (want > 0 ? want : 1) : 1184), 256, 0,
                 static_cast<cudaStream_t>(s)>>>(g, out, n);
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

static inline bool opt_aligned(const void* a, const void* b, const void* c, const void* d) {
  return ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(c) |
           reinterpret_cast<uintptr_t>(d)) & 15) == 0;
}
static inline unsigned opt_grid(int64_t n) {
  const int64_t want = ((n >> 2) + 255) / 256;
  return static_cast<unsigned>(want < 1 ? 1 : (want < 2368 ? want : 2368));
}

extern "C" int ttsb200_adam_step(float* p, const float* g, float* m, float* v, const float* sumsq, int64_t n,
                                 float gscale, float clip, float lr, float beta1, float beta2, float eps,
                                 int32_t step, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(p && g && m && v && n > 0 && step > 0, "adam_step: bad args");
  TTSB_CHECK_ARG(clip <= 0.f || sumsq, "adam_step: clipping needs the squared gradient norm");
  TTSB_CHECK_ARG(opt_aligned(p, g, m, v), "adam_step: buffers must be 16-byte aligned");
  const float bc1 = 1.f - powf(beta1, static_cast<float>(step));
  const float bc2_sqrt = sqrtf(1.f - powf(beta2, static_cast<float>(step)));
  adam_kernel<0><<<opt_grid(n), 256, 0, static_cast<cudaStream_t>(s)>>>(p, g, m, v, sumsq, n, gscale, clip,
                                                                       lr / bc1, beta1, beta2, eps, bc2_sqrt);
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}

extern "C" int ttsb200_radam_step(float* p, const float* g, float* m, float* v, const float* sumsq, int64_t n,
                                  float gscale, float clip, float lr, float beta1, float beta2, float eps,
                                  int32_t step, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(p && g && m && v && n > 0 && step > 0, "radam_step: bad args");
  TTSB_CHECK_ARG(clip <= 0.f || sumsq, "radam_step: clipping needs the squared gradient norm");
  TTSB_CHECK_ARG(opt_aligned(p, g, m, v), "radam_step: buffers must be 16-byte aligned");
  // radam.py:66-80, in double as the reference's Python floats
  const double b1 = beta1, b2 = beta2, t = step;
  const double beta2_t = pow(b2, t);
  const double n_sma_max = 2.0 / (1.0 - b2) - 1.0;
  const double n_sma = n_sma_max - 2.0 * t * beta2_t / (1.0 - beta2_t);
  cudaStream_t st = static_cast<cudaStream_t>(s);
  if (n_sma >= 5.0) {
    const double step_size = sqrt((1.0 - beta2_t) * (n_sma - 4.0) / (n_sma_max - 4.0) * (n_sma - 2.0) / n_sma *
                                  n_sma_max / (n_sma_max - 2.0)) / (1.0 - pow(b1, t));
    adam_kernel<1><<<opt_grid(n), 256, 0, st>>>(p, g, m, v, sumsq, n, gscale, clip,
                                                static_cast<float>(step_size * lr), beta1, beta2, eps, 1.f);
  } else {
    const double step_size = 1.0 / (1.0 - pow(b1, t));
    adam_kernel<2><<<opt_grid(n), 256, 0, st>>>(p, g, m, v, sumsq, n, gscale, clip,
                                                static_cast<float>(step_size * lr), beta1, beta2, eps, 1.f);
  }
  TTSB_LAUNCH_CHECK();
  return TTSB_OK;
}
