// Transformer-TTS decode step, fused per sub-block with thread-block CLUSTERS.
//
// A decode step is a chain of dependent launches at a ~4.6 us floor each (csrc/decode.cu); the pieces between the two
// attentions of a decoder block need full rows between them (LayerNorm, the 1536-wide hidden activation), which is what
// forces a kernel boundary — unless the CTAs that hold the column slices can exchange them on chip.  Here ONE cluster of
// 16 CTAs processes up to 16 rows (the batch): every CTA owns 1/16 of the output columns of each linear, keeps its slice
// of every weight matrix in shared memory (cp.async, issued before the programmatic-dependent-launch wait: weights are
// immutable), and PUSHES its slice of each result into the shared memory of all 16 CTAs (st.shared::cluster); a cluster
// barrier replaces the kernel boundary.
//
//   attn_out_q : x' = x + ctx Wo^T + bo (Models/layers.py:112);  q = LN2(x') Wq^T + bq (layers.py:114-115, modules.py:50)
//   ffn_block  : x' = x + ctx Wo^T + bo (layers.py:116);  h = relu(LN3(x') W1^T + b1);  u = h W2^T + b2 (split over K);
//                z = x' + LN_ff(u + LN3(x')) (layers.py:118-122, modules.py:81-87);  zn = bf16(LN_next(z))
//
// Arithmetic is that of the unfused kernels (bf16 operands, fp32 accumulation, fp32 residual stream, two-pass LayerNorm).
#include <cuda_bf16.h>

#include "common.cuh"
#include "../../include/ttsb200.h"

namespace ttsb {
namespace dcl {

constexpr int kCluster = 16;
constexpr int kD = 384, kHid = 1536;
constexpr int kRows = 16;
constexpr int kThreads = 512, kWarps = 16;
constexpr int kPad = 16;
constexpr int kRbD = kD * 2 + kPad;        // padded bf16 row of 384: 784 bytes
constexpr int kColsD = kD / kCluster;      // 24 output columns per CTA of a 384-wide linear
constexpr int kColsH = kHid / kCluster;    // 96 hidden columns per CTA

__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ void cluster_barrier() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t local_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_cluster_v4(uint32_t addr, uint4 v) {
  asm volatile("st.shared::cluster.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void st_cluster_v2(uint32_t addr, uint2 v) {
  asm volatile("st.shared::cluster.v2.u32 [%0], {%1, %2};" ::"r"(addr), "r"(v.x), "r"(v.y) : "memory");
}
__device__ __forceinline__ void mma_bf16_16816(float* c, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                               uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// this CTA's `rows` rows of a bf16 weight matrix [*][K] (K a multiple of 32) -> padded shared-memory rows
__device__ __forceinline__ void load_w_slice(uint8_t* dst, int rb, const __nv_bfloat16* w, int64_t ldw, int row0, int rows, int K) {
  const int nv = K >> 3;
  for (int i = threadIdx.x; i < rows * nv; i += kThreads) {
    const int r = i / nv, c = i - r * nv;
    cp_async16(dst + r * rb + c * 16, w + static_cast<int64_t>(row0 + r) * ldw + c * 8, 16);
  }
}
// R rows of a dense bf16 (R, 384) tensor -> the padded operand tile (rows >= R: zeros)
__device__ __forceinline__ void load_rows_bf16(uint8_t* dst, const __nv_bfloat16* x, int r0, int R) {
  constexpr int nv = kD >> 3;
  for (int i = threadIdx.x; i < kRows * nv; i += kThreads) {
    const int rr = i / nv, c = i - rr * nv;
    const bool valid = r0 + rr < R;
    cp_async16(dst + rr * kRbD + c * 16, x + static_cast<int64_t>(valid ? r0 + rr : 0) * kD + (valid ? c * 8 : 0), valid ? 16 : 0);
  }
}

// n floats (n % 4 == 0, 16-byte aligned source) -> shared memory
__device__ __forceinline__ void load_f32(float* dst, const float* src, int n) {
  for (int i = threadIdx.x; i < (n >> 2); i += kThreads) cp_async16(dst + i * 4, src + i * 4, 16);
}
// this CTA's 16 x 24 slice of fp32 rows (R, 384): row r at dst + 24 r
__device__ __forceinline__ void load_res_slice(float* dst, const float* x, int r0, int R, int c0) {
  for (int i = threadIdx.x; i < kRows * (kColsD / 4); i += kThreads) {
    const int rr = i / (kColsD / 4), q4 = i - rr * (kColsD / 4);
    const bool valid = r0 + rr < R;
    cp_async16(dst + rr * kColsD + q4 * 4, x + static_cast<int64_t>(valid ? r0 + rr : 0) * kD + c0 + q4 * 4, valid ? 16 : 0);
  }
}

// Partial products of a [16 x K] . [K x 8 n_tiles] slice: task (n-tile, k-slice) -> red[ks][nt][lane][4].
// Fragment convention and the permuted contraction index as in small_linear (csrc/decode.cu).
template <int KS>
__device__ __forceinline__ void slice_mma(const uint8_t* A, int a_rb, const uint8_t* W, int w_rb, int n_tiles, int nkb, float* red) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, tq = lane & 3;
  for (int t = warp; t < n_tiles * KS; t += kWarps) {
    const int nt = t % n_tiles, ks = t / n_tiles;
    const uint8_t* xa = A + g * a_rb + tq * 16;
    const uint8_t* xb = xa + 8 * a_rb;
    const uint8_t* wa = W + (nt * 8 + g) * w_rb + tq * 16;
    // mma.sync has a long latency on this part: four independent accumulator chains instead of one of 2 * nkb / KS links
    float acc[4][4] = {};
    int kb = ks;
    for (; kb + KS < nkb; kb += 2 * KS) {
      const uint4 a0 = *reinterpret_cast<const uint4*>(xa + kb * 64), a1 = *reinterpret_cast<const uint4*>(xa + (kb + KS) * 64);
      const uint4 b0 = *reinterpret_cast<const uint4*>(xb + kb * 64), b1 = *reinterpret_cast<const uint4*>(xb + (kb + KS) * 64);
      const uint4 w0 = *reinterpret_cast<const uint4*>(wa + kb * 64), w1 = *reinterpret_cast<const uint4*>(wa + (kb + KS) * 64);
      mma_bf16_16816(acc[0], a0.x, b0.x, a0.y, b0.y, w0.x, w0.y);
      mma_bf16_16816(acc[1], a0.z, b0.z, a0.w, b0.w, w0.z, w0.w);
      mma_bf16_16816(acc[2], a1.x, b1.x, a1.y, b1.y, w1.x, w1.y);
      mma_bf16_16816(acc[3], a1.z, b1.z, a1.w, b1.w, w1.z, w1.w);
    }
    if (kb < nkb) {
      const uint4 a = *reinterpret_cast<const uint4*>(xa + kb * 64);
      const uint4 b = *reinterpret_cast<const uint4*>(xb + kb * 64);
      const uint4 w = *reinterpret_cast<const uint4*>(wa + kb * 64);
      mma_bf16_16816(acc[0], a.x, b.x, a.y, b.y, w.x, w.y);
      mma_bf16_16816(acc[1], a.z, b.z, a.w, b.w, w.z, w.w);
    }
    *reinterpret_cast<float4*>(red + ((ks * n_tiles + nt) * 32 + lane) * 4) =
        make_float4((acc[0][0] + acc[1][0]) + (acc[2][0] + acc[3][0]), (acc[0][1] + acc[1][1]) + (acc[2][1] + acc[3][1]),
                    (acc[0][2] + acc[1][2]) + (acc[2][2] + acc[3][2]), (acc[0][3] + acc[1][3]) + (acc[2][3] + acc[3][3]));
  }
}
// `cols` columns [k0, k0 + cols) of this CTA's `rows` rows of a bf16 matrix [*][ldw] -> padded shared-memory rows
__device__ __forceinline__ void load_w_cols(uint8_t* dst, int rb, const __nv_bfloat16* w, int64_t ldw, int row0, int rows, int k0, int cols) {
  const int nv = cols >> 3;
  for (int i = threadIdx.x; i < rows * nv; i += kThreads) {
    const int r = i / nv, c = i - r * nv;
    cp_async16(dst + r * rb + c * 16, w + static_cast<int64_t>(row0 + r) * ldw + k0 + c * 8, 16);
  }
}
// value of output element (row, col) of the slice from the KS partials
template <int KS>
__device__ __forceinline__ float slice_value(const float* red, int n_tiles, int row, int col) {
  const int nt = col >> 3, c8 = col & 7;
  const int lane = (row & 7) * 4 + (c8 >> 1), e = (row >> 3) * 2 + (c8 & 1);
  float v = 0.f;
#pragma unroll
  for (int ks = 0; ks < KS; ++ks) v += red[((ks * n_tiles + nt) * 32 + lane) * 4 + e];
  return v;
}

// LayerNorm of the 16 fp32 rows in sX (two-pass, as torch) -> bf16 operand tile; 2 rows per warp
constexpr int kNV = kD / 32;   // columns per lane of a 384-wide row
struct LnParams {
  float g[kNV], b[kNV];
};
__device__ __forceinline__ LnParams load_ln(const float* g, const float* b) {   // immutable: may run before pdl_wait
  LnParams r;
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int j = 0; j < kNV; ++j) {
    r.g[j] = __ldg(g + lane + 32 * j);
    r.b[j] = __ldg(b + lane + 32 * j);
  }
  return r;
}
__device__ __forceinline__ void layernorm_rows(const float* sX, uint8_t* sA, const LnParams& ln, float eps) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int NV = kNV;
  const float* gg = ln.g;
  const float* bb = ln.b;
  for (int rr = warp; rr < kRows; rr += kWarps) {
    float v[NV];
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      v[j] = sX[rr * kD + lane + 32 * j];
      s += v[j];
    }
    const float mean = warp_sum(s) * (1.f / kD);
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) q = fmaf(v[j] - mean, v[j] - mean, q);
    const float rstd = rsqrtf(warp_sum(q) * (1.f / kD) + eps);
    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(sA + rr * kRbD);
#pragma unroll
    for (int j = 0; j < NV; ++j) dst[lane + 32 * j] = __float2bfloat16(fmaf((v[j] - mean) * rstd, gg[j], bb[j]));
  }
}

// out-projection slice: y[row][c0 + col] = ctx . Wo^T + bo + x  for this CTA's 24 columns; written to x_out (global) and
// pushed into every CTA's sX.  `red` / `stage` are scratch (stage: 16 x 24 floats).
// sBo: this CTA's 24 bias values; sRes: its 16 x 24 slice of the residual rows (both prefetched into shared memory)
__device__ __forceinline__ void out_proj_push(const uint8_t* sA0, const uint8_t* sWo, const float* sBo, const float* sRes,
                                              float* x_out, float* sX, float* red, float* stage, int r0, int R, uint32_t rank) {
  slice_mma<2>(sA0, kRbD, sWo, kRbD, kColsD / 8, kD / 32, red);
  __syncthreads();
  const int c0 = rank * kColsD;
  for (int i = threadIdx.x; i < kRows * kColsD; i += kThreads) {
    const int row = i / kColsD, col = i - row * kColsD;
    float v = 0.f;
    if (r0 + row < R) {
      v = slice_value<2>(red, kColsD / 8, row, col) + sBo[col] + sRes[i];
      if (x_out) x_out[static_cast<int64_t>(r0 + row) * kD + c0 + col] = v;
    }
    stage[i] = v;
  }
  __syncthreads();
  // push: 16 rows x 24 floats = 16 x 6 float4 per target CTA
  const uint32_t sx_addr = smem_u32(sX);
  for (int i = threadIdx.x; i < kCluster * kRows * (kColsD / 4); i += kThreads) {
    const int tgt = i / (kRows * (kColsD / 4)), j = i - tgt * (kRows * (kColsD / 4));
    const int row = j / (kColsD / 4), q4 = j - row * (kColsD / 4);
    const uint4 v = *reinterpret_cast<const uint4*>(stage + row * kColsD + q4 * 4);
    st_cluster_v4(map_to_cta(sx_addr + (row * kD + c0 + q4 * 4) * 4, tgt), v);
  }
}

struct AttnOutQ {
  const __nv_bfloat16* ctx;
  const float* x_in;
  const __nv_bfloat16 *wo, *wq;
  const float *bo, *bq, *ln_g, *ln_b;
  float* x_out;
  __nv_bfloat16* q_out;
  float eps;
  int R;
};

constexpr int kSmemA = 2 * kColsD * kRbD + 2 * kRows * kRbD + kRows * kD * 4 + 4096 + kRows * kColsD * 4 * 2 + 256;

__global__ void __launch_bounds__(kThreads, 1) attn_out_q_kernel(const AttnOutQ p) {
  extern __shared__ __align__(16) uint8_t sm[];
  uint8_t* sWo = sm;
  uint8_t* sWq = sWo + kColsD * kRbD;
  uint8_t* sA0 = sWq + kColsD * kRbD;
  uint8_t* sA1 = sA0 + kRows * kRbD;
  float* sX = reinterpret_cast<float*>(sA1 + kRows * kRbD);
  float* red = sX + kRows * kD;                 // 2 k-slices x 3 n-tiles x 128 floats = 3 KB (4 KB reserved)
  float* stage = red + 1024;
  float* sRes = stage + kRows * kColsD;
  float* sBo = sRes + kRows * kColsD;           // 24 + 24 floats: bo | bq slices
  const uint32_t rank = cluster_rank();
  const int r0 = (blockIdx.x / kCluster) * kRows;
  const int c0 = rank * kColsD;
  pdl_launch_dependents();
  load_w_slice(sWo, kRbD, p.wo, kD, c0, kColsD, kD);
  load_w_slice(sWq, kRbD, p.wq, kD, c0, kColsD, kD);
  load_f32(sBo, p.bo + c0, kColsD);
  load_f32(sBo + kColsD, p.bq + c0, kColsD);
  const LnParams ln = load_ln(p.ln_g, p.ln_b);
  pdl_wait();
  load_rows_bf16(sA0, p.ctx, r0, p.R);
  load_res_slice(sRes, p.x_in, r0, p.R, c0);
  cp_async_wait_all();
  __syncthreads();
  out_proj_push(sA0, sWo, sBo, sRes, p.x_out, sX, red, stage, r0, p.R, rank);
  cluster_barrier();                            // every CTA holds the 16 full rows of x'
  layernorm_rows(sX, sA1, ln, p.eps);
  __syncthreads();
  slice_mma<2>(sA1, kRbD, sWq, kRbD, kColsD / 8, kD / 32, red);
  __syncthreads();
  for (int i = threadIdx.x; i < kRows * kColsD; i += kThreads) {
    const int row = i / kColsD, col = i - row * kColsD;
    if (r0 + row < p.R)
      p.q_out[static_cast<int64_t>(r0 + row) * kD + c0 + col] =
          __float2bfloat16(slice_value<2>(red, kColsD / 8, row, col) + sBo[kColsD + col]);
  }
  // (every push into this CTA's shared memory happened before the barrier above: it may exit on its own)
}

__device__ long long g_dcl_trace[16];
__device__ int g_dcl_trace_on = 0;
#define DCL_T(i) do { if (tr) tt[i] = clock64(); } while (0)

struct FfnBlock {
  const __nv_bfloat16* ctx;
  const float* x_in;
  const __nv_bfloat16 *wo, *w1, *w2;
  const float *bo, *b1, *b2, *ln3_g, *ln3_b, *lnff_g, *lnff_b, *lnn_g, *lnn_b;
  float* z_out;
  __nv_bfloat16* zn_out;
  float eps;
  int R;
};

// FFN2 is split over K, not N: every CTA contracts its OWN 96 hidden columns against the matching 96 K-columns of all 384
// W2 rows, and the 16 partial sums of output row r meet on CTA r (24 KB out per CTA, no exchange of the hidden activation,
// no barrier between FFN1 and FFN2).
// shared memory: Wo slice | W1 slice (later: rows 192..383 of the W2 K-slice) | rows 0..191 of the W2 K-slice |
// R0 (ctx tile + out-projection scratch; after barrier #1: the 16 partial u rows this CTA receives) |
// X fp32 [16][384] (after LayerNorm 3: FFN1 scratch) | A1 = LN3(x') bf16 | hidden slice bf16 [16][96] | x' row | u row | vectors
constexpr int kRbK = kColsH * 2 + kPad;                         // 208 bytes: one row of the W2 K-slice / of the hidden slice
constexpr int kHalfN = kD / 2;                                  // 192 W2 rows per half
constexpr int kOffW1 = kColsD * kRbD;
constexpr int kOffW2a = kOffW1 + kColsH * kRbD;
constexpr int kOffR0 = kOffW2a + kHalfN * kRbK;
constexpr int kOffX = kOffR0 + kRows * kD * 4;
constexpr int kOffA1 = kOffX + kRows * kD * 4;
constexpr int kOffHs = kOffA1 + kRows * kRbD;
constexpr int kOffXrow = kOffHs + kRows * kRbK;
constexpr int kOffU = kOffXrow + kD * 4;
constexpr int kOffVec = kOffU + kD * 4;                         // bo[24] | b1[96] | b2[384] | residual slice [16][24]
constexpr int kVecBytes = (kColsD + kColsH + kD + kRows * kColsD) * 4;
constexpr int kSmemB = kOffVec + kVecBytes;
static_assert(kHalfN * kRbK <= kColsH * kRbD, "the second half of the W2 K-slice must fit where the W1 slice was");
static_assert(4 * (kColsH / 8) * 128 * 4 <= kRows * kD * 4, "FFN1 scratch (4 k-slices x 12 n-tiles) must fit in the X tile");
static_assert(kRows * kRbD + (1024 + kRows * kColsD) * 4 <= kRows * kD * 4, "ctx tile + out-projection scratch must fit in R0");
static_assert(kSmemB <= 231424, "shared memory budget (1 KB is static)");

__global__ void __launch_bounds__(kThreads, 1) ffn_block_kernel(const FfnBlock p) {
  extern __shared__ __align__(16) uint8_t sm[];
  uint8_t* sWo = sm;
  uint8_t* sW1 = sm + kOffW1;
  uint8_t* sW2a = sm + kOffW2a;
  uint8_t* sR0 = sm + kOffR0;
  float* sPart = reinterpret_cast<float*>(sR0);                 // [16 source CTAs][384]
  float* sX = reinterpret_cast<float*>(sm + kOffX);
  uint8_t* sA1 = sm + kOffA1;
  uint8_t* sHs = sm + kOffHs;
  float* sXrow = reinterpret_cast<float*>(sm + kOffXrow);
  float* sU = reinterpret_cast<float*>(sm + kOffU);
  float* sBo = reinterpret_cast<float*>(sm + kOffVec);
  float* sB1 = sBo + kColsD;
  float* sB2 = sB1 + kColsH;
  float* sRes = sB2 + kD;
  float* red0 = reinterpret_cast<float*>(sR0 + kRows * kRbD);   // out-projection scratch, behind the ctx tile
  float* stage0 = red0 + 1024;
  float* red = sX;                                              // FFN1 scratch: the X tile after LayerNorm 3
  const uint32_t rank = cluster_rank();
  const int r0 = (blockIdx.x / kCluster) * kRows;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool tr = g_dcl_trace_on && blockIdx.x == 0 && threadIdx.x == 0;   // phase timeline of one thread (experiments)
  long long tt[14];
  DCL_T(0);
  pdl_launch_dependents();
  // everything immutable leaves for shared memory / registers before the predecessor has finished
  load_w_slice(sWo, kRbD, p.wo, kD, rank * kColsD, kColsD, kD);
  load_w_slice(sW1, kRbD, p.w1, kD, rank * kColsH, kColsH, kD);
  load_w_cols(sW2a, kRbK, p.w2, kHid, 0, kHalfN, rank * kColsH, kColsH);
  load_f32(sBo, p.bo + rank * kColsD, kColsD);
  load_f32(sB1, p.b1 + rank * kColsH, kColsH);
  load_f32(sB2, p.b2, kD);
  const LnParams ln3 = load_ln(p.ln3_g, p.ln3_b);
  LnParams lnff, lnn;
  if (warp == 0) {
    lnff = load_ln(p.lnff_g, p.lnff_b);
    lnn = load_ln(p.lnn_g, p.lnn_b);
  }
  DCL_T(1);
  pdl_wait();
  DCL_T(2);
  load_rows_bf16(sR0, p.ctx, r0, p.R);
  load_res_slice(sRes, p.x_in, r0, p.R, rank * kColsD);
  cp_async_wait_all();
  __syncthreads();
  DCL_T(3);
  out_proj_push(sR0, sWo, sBo, sRes, nullptr, sX, red0, stage0, r0, p.R, rank);
  DCL_T(4);
  cluster_barrier();                            // #1: full rows of x' everywhere; R0 is free to receive partial sums
  DCL_T(5);
  layernorm_rows(sX, sA1, ln3, p.eps);
  for (int i = threadIdx.x; i < kD; i += kThreads) sXrow[i] = sX[rank * kD + i];   // the tail's residual row
  __syncthreads();                              // the X tile is scratch from here on
  DCL_T(6);
  // ---- FFN1 slice: this CTA's 96 hidden columns = 12 n-tiles x 4 k-slices of 3 k-blocks; warp w takes k-slice w & 3 and
  // n-tiles (w >> 2) + 4 i: the activation fragments of its k-slice are loaded once and used for its three n-tiles (a task
  // per (n-tile, k-slice) re-read them every time: 4 k clk, shared-memory bandwidth) ----
  {
    constexpr int KS = 4, PER = kD / 32 / KS, NT = kColsH / 8;
    const int g = lane >> 2, tq = lane & 3;
    const int ks = warp & (KS - 1);
    uint4 a[PER], b[PER];
#pragma unroll
    for (int i = 0; i < PER; ++i) {
      a[i] = *reinterpret_cast<const uint4*>(sA1 + g * kRbD + tq * 16 + (ks * PER + i) * 64);
      b[i] = *reinterpret_cast<const uint4*>(sA1 + (g + 8) * kRbD + tq * 16 + (ks * PER + i) * 64);
    }
    for (int nt = warp >> 2; nt < NT; nt += kWarps / KS) {
      const uint8_t* wa = sW1 + (nt * 8 + g) * kRbD + tq * 16 + ks * PER * 64;
      float c0[4] = {0.f, 0.f, 0.f, 0.f}, c1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int i = 0; i < PER; ++i) {
        const uint4 w = *reinterpret_cast<const uint4*>(wa + i * 64);
        mma_bf16_16816(c0, a[i].x, b[i].x, a[i].y, b[i].y, w.x, w.y);
        mma_bf16_16816(c1, a[i].z, b[i].z, a[i].w, b[i].w, w.z, w.w);
      }
      *reinterpret_cast<float4*>(red + ((ks * NT + nt) * 32 + lane) * 4) =
          make_float4(c0[0] + c1[0], c0[1] + c1[1], c0[2] + c1[2], c0[3] + c1[3]);
    }
  }
  __syncthreads();
  DCL_T(7);
  load_w_cols(sW1, kRbK, p.w2, kHid, kHalfN, kHalfN, rank * kColsH, kColsH);   // W2 rows 192..383 replace the W1 slice
  for (int i = threadIdx.x; i < kRows * kColsH; i += kThreads) {
    const int row = i / kColsH, col = i - row * kColsH;
    reinterpret_cast<__nv_bfloat16*>(sHs + row * kRbK)[col] =
        __float2bfloat16(fmaxf(slice_value<4>(red, kColsH / 8, row, col) + sB1[col], 0.f));
  }
  __syncthreads();
  DCL_T(8);
  // ---- FFN2 over this CTA's K-slice: partial u[16][384]; row r's partial goes to CTA r ----
  {
    const int g = lane >> 2, tq = lane & 3;
    uint4 a[kColsH / 32], b[kColsH / 32];       // the hidden slice's fragments: loaded once, used for every n-tile
#pragma unroll
    for (int kb = 0; kb < kColsH / 32; ++kb) {
      a[kb] = *reinterpret_cast<const uint4*>(sHs + g * kRbK + tq * 16 + kb * 64);
      b[kb] = *reinterpret_cast<const uint4*>(sHs + (g + 8) * kRbK + tq * 16 + kb * 64);
    }
    const uint32_t part_addr = smem_u32(sPart);
    auto tiles = [&](const uint8_t* W, int nt0) {
      for (int t = warp; t < kHalfN / 8; t += kWarps) {
        const uint8_t* wa = W + (t * 8 + g) * kRbK + tq * 16;
        float c0[4] = {0.f, 0.f, 0.f, 0.f}, c1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int kb = 0; kb < kColsH / 32; ++kb) {
          const uint4 w = *reinterpret_cast<const uint4*>(wa + kb * 64);
          mma_bf16_16816(c0, a[kb].x, b[kb].x, a[kb].y, b[kb].y, w.x, w.y);
          mma_bf16_16816(c1, a[kb].z, b[kb].z, a[kb].w, b[kb].w, w.z, w.w);
        }
        const uint32_t off = part_addr + (rank * kD + (nt0 + t) * 8 + tq * 2) * 4;
        st_cluster_v2(map_to_cta(off, g), make_uint2(__float_as_uint(c0[0] + c1[0]), __float_as_uint(c0[1] + c1[1])));
        st_cluster_v2(map_to_cta(off, g + 8), make_uint2(__float_as_uint(c0[2] + c1[2]), __float_as_uint(c0[3] + c1[3])));
      }
    };
    tiles(sW2a, 0);
    DCL_T(9);
    cp_async_wait_all();
    __syncthreads();
    DCL_T(10);
    tiles(sW1, kHalfN / 8);
  }
  DCL_T(11);
  cluster_barrier();                            // #2: CTA r holds the 16 partial sums of row r
  DCL_T(12);
  // ---- tail of row `rank`: u = sum of the partials + b2;  z = x' + LN_ff(u + LN3(x'));  zn = bf16(LN_next(z)) ----
  const int row = r0 + static_cast<int>(rank);
  if (row < p.R) {
    for (int k = threadIdx.x; k < kD; k += kThreads) {
      float u = sB2[k];
#pragma unroll
      for (int src = 0; src < kCluster; ++src) u += sPart[src * kD + k];
      sU[k] = u;
    }
  }
  __syncthreads();
  if (warp == 0 && row < p.R) {
    constexpr int NV = kNV;
    float v[NV];
    const __nv_bfloat16* xn = reinterpret_cast<const __nv_bfloat16*>(sA1 + rank * kRbD);
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int k = lane + 32 * j;
      v[j] = sU[k] + __bfloat162float(xn[k]);
    }
    auto stats = [&](const float* a, float& mean, float& rstd) {
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < NV; ++j) s += a[j];
      mean = warp_sum(s) * (1.f / kD);
      float q = 0.f;
#pragma unroll
      for (int j = 0; j < NV; ++j) q = fmaf(a[j] - mean, a[j] - mean, q);
      rstd = rsqrtf(warp_sum(q) * (1.f / kD) + p.eps);
    };
    float mean, rstd;
    stats(v, mean, rstd);
#pragma unroll
    for (int j = 0; j < NV; ++j) v[j] = sXrow[lane + 32 * j] + fmaf((v[j] - mean) * rstd, lnff.g[j], lnff.b[j]);
    stats(v, mean, rstd);
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int k = lane + 32 * j;
      p.z_out[static_cast<int64_t>(row) * kD + k] = v[j];
      p.zn_out[static_cast<int64_t>(row) * kD + k] = __float2bfloat16(fmaf((v[j] - mean) * rstd, lnn.g[j], lnn.b[j]));
    }
  }
  if (tr) {
    tt[13] = clock64();
    for (int i = 0; i < 14; ++i) g_dcl_trace[i] = tt[i] - tt[0];
  }
}

template <typename P>
static cudaError_t launch_cluster(void (*kern)(P), int row_tiles, size_t smem, cudaStream_t st, const P& p) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(kCluster * row_tiles);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kCluster;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  return cudaLaunchKernelEx(&cfg, kern, p);
}

// a 16-CTA cluster is above the portable size of 8: opt in once per device, together with the shared-memory size
template <typename K>
static int opt_in(K kern, int smem) {
  static unsigned long long done = 0;
  int dev = 0;
  TTSB_CHECK_CUDA(cudaGetDevice(&dev));
  if (!((done >> (dev & 63)) & 1ull)) {
    TTSB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    TTSB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    done |= 1ull << (dev & 63);
  }
  return TTSB_OK;
}

}  // namespace dcl
}  // namespace ttsb

using namespace ttsb;

static bool dcl_aligned(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

extern "C" int ttsb200_decode_attn_out_q(const void* ctx_bf16, const float* x_in, const void* wo_bf16, const float* bo,
                                         const float* ln_g, const float* ln_b, const void* wq_bf16, const float* bq, float eps,
                                         float* x_out, void* q_bf16, int32_t R, int32_t D, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(ctx_bf16 && x_in && wo_bf16 && bo && ln_g && ln_b && wq_bf16 && bq && x_out && q_bf16 && R > 0,
                 "decode_attn_out_q: bad args");
  TTSB_CHECK_ARG(D == dcl::kD, "decode_attn_out_q: D = %d unsupported (384)", D);
  TTSB_CHECK_ARG(dcl_aligned(ctx_bf16) && dcl_aligned(wo_bf16) && dcl_aligned(wq_bf16), "decode_attn_out_q: 16-byte alignment");
  dcl::AttnOutQ p;
  p.ctx = static_cast<const __nv_bfloat16*>(ctx_bf16);
  p.x_in = x_in;
  p.wo = static_cast<const __nv_bfloat16*>(wo_bf16);
  p.wq = static_cast<const __nv_bfloat16*>(wq_bf16);
  p.bo = bo; p.bq = bq; p.ln_g = ln_g; p.ln_b = ln_b;
  p.x_out = x_out;
  p.q_out = static_cast<__nv_bfloat16*>(q_bf16);
  p.eps = eps;
  p.R = R;
  int rc = dcl::opt_in(dcl::attn_out_q_kernel, dcl::kSmemA);
  if (rc) return rc;
  TTSB_CHECK_CUDA(dcl::launch_cluster(dcl::attn_out_q_kernel, (R + dcl::kRows - 1) / dcl::kRows, dcl::kSmemA,
                                      static_cast<cudaStream_t>(s), p));
  return TTSB_OK;
}

extern "C" int ttsb200_decode_ffn_block(const void* ctx_bf16, const float* x_in, const void* wo_bf16, const float* bo,
                                        const float* ln3_g, const float* ln3_b, const void* w1_bf16, const float* b1,
                                        const void* w2_bf16, const float* b2, const float* lnff_g, const float* lnff_b,
                                        const float* lnn_g, const float* lnn_b, float eps, float* z_out, void* zn_bf16,
                                        int32_t R, int32_t D, int32_t hidden, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(ctx_bf16 && x_in && wo_bf16 && bo && ln3_g && ln3_b && w1_bf16 && b1 && w2_bf16 && b2 && lnff_g && lnff_b &&
                     lnn_g && lnn_b && z_out && zn_bf16 && R > 0, "decode_ffn_block: bad args");
  TTSB_CHECK_ARG(D == dcl::kD && hidden == dcl::kHid, "decode_ffn_block: this build is for D=384, hidden=1536");
  TTSB_CHECK_ARG(dcl_aligned(ctx_bf16) && dcl_aligned(wo_bf16) && dcl_aligned(w1_bf16) && dcl_aligned(w2_bf16),
                 "decode_ffn_block: 16-byte alignment");
  dcl::FfnBlock p;
  p.ctx = static_cast<const __nv_bfloat16*>(ctx_bf16);
  p.x_in = x_in;
  p.wo = static_cast<const __nv_bfloat16*>(wo_bf16);
  p.w1 = static_cast<const __nv_bfloat16*>(w1_bf16);
  p.w2 = static_cast<const __nv_bfloat16*>(w2_bf16);
  p.bo = bo; p.b1 = b1; p.b2 = b2;
  p.ln3_g = ln3_g; p.ln3_b = ln3_b; p.lnff_g = lnff_g; p.lnff_b = lnff_b; p.lnn_g = lnn_g; p.lnn_b = lnn_b;
  p.z_out = z_out;
  p.zn_out = static_cast<__nv_bfloat16*>(zn_bf16);
  p.eps = eps;
  p.R = R;
  int rc = dcl::opt_in(dcl::ffn_block_kernel, dcl::kSmemB);
  if (rc) return rc;
  TTSB_CHECK_CUDA(dcl::launch_cluster(dcl::ffn_block_kernel, (R + dcl::kRows - 1) / dcl::kRows, dcl::kSmemB,
                                      static_cast<cudaStream_t>(s), p));
  return TTSB_OK;
}

// experiments: mode 1 = record the phase timeline of ffn_block (CTA 0, thread 0), 0 = off, 2 = read 16 values
extern "C" int ttsb200_debug_dcl_trace(long long* dst, int mode) {
  if (mode == 0 || mode == 1) return cudaMemcpyToSymbol(ttsb::dcl::g_dcl_trace_on, &mode, sizeof(int)) == cudaSuccess ? 0 : -1;
  return cudaMemcpyFromSymbol(dst, ttsb::dcl::g_dcl_trace, 16 * sizeof(long long)) == cudaSuccess ? 0 : -1;
}
