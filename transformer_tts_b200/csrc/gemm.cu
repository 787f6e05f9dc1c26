// Implicit-GEMM conv1d / linear for sm_100a: TMA tile loads -> tcgen05.mma (bf16 x bf16 -> fp32 in TMEM)
// -> fused epilogues (bias/affine + activation, residual + LayerNorm, FFN tail with two LayerNorms,
// predictor ReLU+LayerNorm(+dot)).  One CTA computes a 128(time) x BLOCK_N(channel) output tile of ONE
// utterance, so conv taps are plain time-shifted TMA boxes and TMA's out-of-bounds zero fill is the
// conv zero padding.  Warp roles: warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer,
// warps 2..5 = epilogue (thread == output row, TMEM lane == row).
//
// Reference ops replaced: nn.Linear / nn.Conv1d call sites listed in include/ttsb200.h.
#include "common.cuh"
#include "../../include/ttsb200.h"

namespace ttsb {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;  // bf16 elements = one 128-byte swizzle row
constexpr int kUmmaK = 16;
constexpr int kGemmThreads = 192;

template <int BLOCK_N>
struct GemmCfg {
  static constexpr int kWBoxN = BLOCK_N <= 256 ? BLOCK_N : BLOCK_N / 2;  // TMA box / UMMA N
  static constexpr int kNumWBoxes = BLOCK_N / kWBoxN;
  static constexpr int kABytes = kBlockM * kBlockK * 2;
  static constexpr int kWBoxBytes = kWBoxN * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kNumWBoxes * kWBoxBytes;
  static constexpr int kStages = BLOCK_N == 384 ? 3 : (BLOCK_N == 256 ? 2 : (BLOCK_N == 128 ? 3 : 4));
  static constexpr int kTmemCols = BLOCK_N <= 128 ? 128 : (BLOCK_N <= 256 ? 256 : 512);
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
  static_assert(kWBoxBytes % 1024 == 0, "W box must keep 1024B swizzle-atom alignment");
  static_assert(kWBoxN % 16 == 0 && kWBoxN <= 256, "UMMA N");
};

struct GemmDev {
  int B, T, N, C, taps, pad_left, epi, act;
  const float* bias;
  const float* scale;
  const float* res_f32;
  long long res_bstride, res_tstride;
  const float* res_scale;
  const __nv_bfloat16* res_bf16;
  const float *ln1_g, *ln1_b, *ln2_g, *ln2_b;
  float eps;
  const float* dot_w;
  const float* dot_b;
  const uint8_t* row_mask;
  float* out_f32;
  __nv_bfloat16* out_bf16;
  float* out2_f32;
  float* out_row;
};

__device__ __forceinline__ float apply_act(float v, int act) {
  if (act == TTSB200_ACT_RELU) return fmaxf(v, 0.f);
  if (act == TTSB200_ACT_TANH) return tanhf(v);
  return v;
}

__device__ __forceinline__ void store_bf16x8(__nv_bfloat16* dst, const float* v) {
  uint4 u;
  u.x = pack_bf16(v[0], v[1]);
  u.y = pack_bf16(v[2], v[3]);
  u.z = pack_bf16(v[4], v[5]);
  u.w = pack_bf16(v[6], v[7]);
  *reinterpret_cast<uint4*>(dst) = u;
}
__device__ __forceinline__ void load_bf16x8(const __nv_bfloat16* src, float* v) {
  uint4 u = *reinterpret_cast<const uint4*>(src);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 f = __bfloat1622float2(h[i]);
    v[2 * i] = f.x;
    v[2 * i + 1] = f.y;
  }
}

// Row LayerNorm pass over values living in TMEM (thread == row): y = (v-mean)*rstd*g + b.
// The caller has already written v to TMEM columns [0, N) of its lane and knows sum / sumsq.
template <int N, class F>
__device__ __forceinline__ void for_each_chunk32(uint32_t trow, F&& f) {
#pragma unroll 1
  for (int c = 0; c < N; c += 32) {
    float v[32];
    tmem_ld32(trow + c, v);
    tmem_ld_wait();
    f(c, v);
  }
}

template <int BLOCK_N, int EPI>
__global__ void __launch_bounds__(kGemmThreads, 1)
conv_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                 const GemmDev p) {
  using Cfg = GemmCfg<BLOCK_N>;
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment for SWIZZLE_128B atoms
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  // bars[0..S) full, [S..2S) empty, [2S] accumulator ready; then the TMEM base address slot
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * Cfg::kStages + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int t0 = blockIdx.x * kBlockM;
  const int n0 = blockIdx.y * BLOCK_N;
  const int b = blockIdx.z;
  const int kblocks = (p.C + kBlockK - 1) / kBlockK;
  const int num_iters = p.taps * kblocks;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmW);
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(smem_u32(&bars[s]), 1);
      mbar_init(smem_u32(&bars[Cfg::kStages + s]), 1);
    }
    mbar_init(smem_u32(&bars[2 * Cfg::kStages]), 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(tmem_slot), Cfg::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      for (int it = 0; it < num_iters; ++it) {
        const int s = it % Cfg::kStages;
        const uint32_t ph = (it / Cfg::kStages) & 1;
        mbar_wait(smem_u32(&bars[Cfg::kStages + s]), ph ^ 1);
        const uint32_t full = smem_u32(&bars[s]);
        mbar_arrive_expect_tx(full, Cfg::kStageBytes);
        const int tap = it / kblocks;
        const int c0 = (it - tap * kblocks) * kBlockK;
        uint8_t* st = smem + s * Cfg::kStageBytes;
        tma_load_3d(smem_u32(st), &tmA, full, c0, t0 + tap - p.pad_left, b);
#pragma unroll
        for (int wb = 0; wb < Cfg::kNumWBoxes; ++wb)
          tma_load_2d(smem_u32(st + Cfg::kABytes + wb * Cfg::kWBoxBytes), &tmW, full, c0,
                      tap * p.N + n0 + wb * Cfg::kWBoxN);
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(kBlockM, Cfg::kWBoxN, 0, 0);
      for (int it = 0; it < num_iters; ++it) {
        const int s = it % Cfg::kStages;
        const uint32_t ph = (it / Cfg::kStages) & 1;
        mbar_wait(smem_u32(&bars[s]), ph);
        tc_fence_after();
        const uint32_t a_base = smem_u32(smem + s * Cfg::kStageBytes);
        const uint32_t w_base = a_base + Cfg::kABytes;
#pragma unroll
        for (int k = 0; k < kBlockK / kUmmaK; ++k) {
          const uint64_t adesc = umma_smem_desc(a_base + k * kUmmaK * 2, 16, 1024, kLayoutSW128);
#pragma unroll
          for (int wb = 0; wb < Cfg::kNumWBoxes; ++wb) {
            const uint64_t bdesc = umma_smem_desc(w_base + wb * Cfg::kWBoxBytes + k * kUmmaK * 2,
                                                  16, 1024, kLayoutSW128);
            umma_bf16(tmem_base + wb * Cfg::kWBoxN, adesc, bdesc, idesc, (it | k) != 0 ? 1u : 0u);
          }
        }
        umma_commit(smem_u32(&bars[Cfg::kStages + s]));  // frees the smem stage when the MMAs retire
      }
      umma_commit(smem_u32(&bars[2 * Cfg::kStages]));  // accumulator complete
    }
  } else {
    // ===================== epilogue: thread == output row =====================
    const int q = warp & 3;  // TMEM lane quarter this warp may touch
    const int row = q * 32 + lane;
    const int t = t0 + row;
    const bool valid = t < p.T;
    const long long r = static_cast<long long>(b) * p.T + (valid ? t : 0);
    const uint32_t trow = tmem_base + (static_cast<uint32_t>(q * 32) << 16);

    mbar_wait(smem_u32(&bars[2 * Cfg::kStages]), 0);
    tc_fence_after();

    if constexpr (EPI == TTSB200_EPI_AFFINE_ACT) {
      constexpr int CW = (BLOCK_N % 32 == 0) ? 32 : 16;
#pragma unroll 1
      for (int c = 0; c < BLOCK_N; c += CW) {
        float v[CW];
        if constexpr (CW == 32) tmem_ld32(trow + c, v);
        else tmem_ld16(trow + c, v);
        tmem_ld_wait();
        if (valid) {
          const int n = n0 + c;
#pragma unroll
          for (int i = 0; i < CW; ++i) {
            const float sc = p.scale ? __ldg(p.scale + n + i) : 1.f;
            v[i] = apply_act(fmaf(v[i], sc, __ldg(p.bias + n + i)), p.act);
          }
          if (p.res_f32) {
            const float* rp = p.res_f32 + b * p.res_bstride + t * p.res_tstride + n;
#pragma unroll
            for (int i = 0; i < CW; i += 4) {
              const float4 f = *reinterpret_cast<const float4*>(rp + i);
              v[i] += f.x; v[i + 1] += f.y; v[i + 2] += f.z; v[i + 3] += f.w;
            }
          }
          if (p.out_f32) {
            float* op = p.out_f32 + r * p.N + n;
#pragma unroll
            for (int i = 0; i < CW; i += 4)
              *reinterpret_cast<float4*>(op + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
          }
          if (p.out_bf16) {
            __nv_bfloat16* op = p.out_bf16 + r * p.N + n;
#pragma unroll
            for (int i = 0; i < CW; i += 8) store_bf16x8(op + i, v + i);
          }
        }
      }
    } else {
      // ---- LayerNorm family: the CTA owns whole rows (BLOCK_N == N, n0 == 0) ----
      const float inv_n = 1.f / static_cast<float>(BLOCK_N);
      float sum = 0.f, sq = 0.f;
      // pass 1: v = pre-LN value; written back to TMEM in place
      for_each_chunk32<BLOCK_N>(trow, [&](int c, float* v) {
        if constexpr (EPI == TTSB200_EPI_RES_LN) {
          const float rs = p.res_scale ? __ldg(p.res_scale) : 1.f;
          const float* rp = p.res_f32 + b * p.res_bstride + (valid ? t : 0) * p.res_tstride + c;
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            const float4 f = *reinterpret_cast<const float4*>(rp + i);
            v[i] += __ldg(p.bias + c + i) + rs * f.x;
            v[i + 1] += __ldg(p.bias + c + i + 1) + rs * f.y;
            v[i + 2] += __ldg(p.bias + c + i + 2) + rs * f.z;
            v[i + 3] += __ldg(p.bias + c + i + 3) + rs * f.w;
          }
          if (valid && p.out_f32) {
            float* op = p.out_f32 + r * p.N + c;
#pragma unroll
            for (int i = 0; i < 32; i += 4)
              *reinterpret_cast<float4*>(op + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
          }
        } else if constexpr (EPI == TTSB200_EPI_FFN2) {
          const __nv_bfloat16* rp = p.res_bf16 + r * p.N + c;
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            float f[8];
            load_bf16x8(rp + i, f);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[i + j] += __ldg(p.bias + c + i + j) + f[j];
          }
        } else {  // RELU_LN / RELU_LN_DOT
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i] + __ldg(p.bias + c + i), 0.f);
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          sum += v[i];
          sq = fmaf(v[i], v[i], sq);
        }
        tmem_st32(trow + c, v);
      });
      tmem_st_wait();
      float mean = sum * inv_n;
      float rstd = rsqrtf(fmaxf(sq * inv_n - mean * mean, 0.f) + p.eps);

      if constexpr (EPI == TTSB200_EPI_RES_LN || EPI == TTSB200_EPI_RELU_LN) {
        for_each_chunk32<BLOCK_N>(trow, [&](int c, float* v) {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            v[i] = fmaf((v[i] - mean) * rstd, __ldg(p.ln1_g + c + i), __ldg(p.ln1_b + c + i));
          if (valid) {
            __nv_bfloat16* op = p.out_bf16 + r * p.N + c;
#pragma unroll
            for (int i = 0; i < 32; i += 8) store_bf16x8(op + i, v + i);
          }
        });
      } else if constexpr (EPI == TTSB200_EPI_RELU_LN_DOT) {
        float dot = 0.f;
        for_each_chunk32<BLOCK_N>(trow, [&](int c, float* v) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const float u = fmaf((v[i] - mean) * rstd, __ldg(p.ln1_g + c + i), __ldg(p.ln1_b + c + i));
            dot = fmaf(u, __ldg(p.dot_w + c + i), dot);
          }
        });
        if (valid) {
          float o = dot + __ldg(p.dot_b);
          if (p.row_mask && p.row_mask[r] == 0) o = 0.f;
          p.out_row[r] = o;
        }
      } else {  // FFN2: z = y + LN1(v); then LN2(z)
        float sum2 = 0.f, sq2 = 0.f;
        for_each_chunk32<BLOCK_N>(trow, [&](int c, float* v) {
          const float* yp = p.res_f32 + b * p.res_bstride + (valid ? t : 0) * p.res_tstride + c;
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            const float4 f = *reinterpret_cast<const float4*>(yp + i);
            const float y4[4] = {f.x, f.y, f.z, f.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float u = fmaf((v[i + j] - mean) * rstd, __ldg(p.ln1_g + c + i + j),
                                   __ldg(p.ln1_b + c + i + j));
              v[i + j] = y4[j] + u;
            }
          }
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            sum2 += v[i];
            sq2 = fmaf(v[i], v[i], sq2);
          }
          if (valid && p.out_f32) {
            float* op = p.out_f32 + r * p.N + c;
#pragma unroll
            for (int i = 0; i < 32; i += 4)
              *reinterpret_cast<float4*>(op + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
          }
          tmem_st32(trow + c, v);
        });
        tmem_st_wait();
        mean = sum2 * inv_n;
        rstd = rsqrtf(fmaxf(sq2 * inv_n - mean * mean, 0.f) + p.eps);
        for_each_chunk32<BLOCK_N>(trow, [&](int c, float* v) {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            v[i] = fmaf((v[i] - mean) * rstd, __ldg(p.ln2_g + c + i), __ldg(p.ln2_b + c + i));
          if (valid) {
            if (p.out_bf16) {
              __nv_bfloat16* op = p.out_bf16 + r * p.N + c;
#pragma unroll
              for (int i = 0; i < 32; i += 8) store_bf16x8(op + i, v + i);
            }
            if (p.out2_f32) {
              float* op = p.out2_f32 + r * p.N + c;
#pragma unroll
              for (int i = 0; i < 32; i += 4)
                *reinterpret_cast<float4*>(op + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
            }
          }
        });
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::kTmemCols);
}

template <int BLOCK_N, int EPI>
static int launch_gemm(const ttsb200_gemm_args& a, const GemmDev& d, cudaStream_t stream) {
  using Cfg = GemmCfg<BLOCK_N>;
  CUtensorMap tmA, tmW;
  {
    const uint64_t dims[3] = {static_cast<uint64_t>(a.C), static_cast<uint64_t>(a.T),
                              static_cast<uint64_t>(a.B)};
    const uint64_t strides[2] = {static_cast<uint64_t>(a.ldx) * 2,
                                 static_cast<uint64_t>(a.ldx) * 2 * static_cast<uint64_t>(a.T)};
    const uint32_t box[3] = {kBlockK, kBlockM, 1};
    int rc = make_tmap_bf16(&tmA, a.x, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  }
  {
    const uint64_t dims[2] = {static_cast<uint64_t>(a.C),
                              static_cast<uint64_t>(a.taps) * static_cast<uint64_t>(a.N)};
    const uint64_t strides[1] = {static_cast<uint64_t>(a.C) * 2};
    const uint32_t box[2] = {kBlockK, static_cast<uint32_t>(Cfg::kWBoxN)};
    int rc = make_tmap_bf16(&tmW, a.w, 2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  }
  auto kern = conv_gemm_kernel<BLOCK_N, EPI>;
  static bool attr_done = false;  // per instantiation
  if (!attr_done) {
    TTSB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Cfg::kSmemBytes));
    attr_done = true;
  }
  dim3 grid((a.T + kBlockM - 1) / kBlockM, a.N / BLOCK_N, a.B);
  kern<<<grid, kGemmThreads, Cfg::kSmemBytes, stream>>>(tmA, tmW, d);
  TTSB_CHECK_CUDA(cudaGetLastError());
  return TTSB_OK;
}

}  // namespace ttsb

using namespace ttsb;

extern "C" int ttsb200_conv_gemm(const ttsb200_gemm_args* a, ttsb200_stream_t stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  TTSB_CHECK_ARG(a != nullptr, "conv_gemm: null args");
  TTSB_CHECK_ARG(a->B > 0 && a->T > 0 && a->N > 0 && a->C > 0 && a->taps > 0, "conv_gemm: bad dims");
  TTSB_CHECK_ARG(a->x && a->w, "conv_gemm: null x/w");
  TTSB_CHECK_ARG(a->C % 8 == 0 && a->ldx % 8 == 0 && a->ldx >= a->C,
                 "conv_gemm: C (%d) and ldx (%lld) must be multiples of 8, ldx >= C", a->C,
                 (long long)a->ldx);
  TTSB_CHECK_ARG((reinterpret_cast<uintptr_t>(a->x) & 15) == 0 &&
                     (reinterpret_cast<uintptr_t>(a->w) & 15) == 0,
                 "conv_gemm: x/w must be 16-byte aligned");
  TTSB_CHECK_ARG(a->N % a->block_n == 0, "conv_gemm: N (%d) not a multiple of block_n (%d)", a->N,
                 a->block_n);
  TTSB_CHECK_ARG(a->bias != nullptr, "conv_gemm: bias required");
  GemmDev d;
  d.B = a->B; d.T = a->T; d.N = a->N; d.C = a->C; d.taps = a->taps; d.pad_left = a->pad_left;
  d.epi = a->epi; d.act = a->act;
  d.bias = a->bias; d.scale = a->scale; d.res_f32 = a->res_f32;
  d.res_bstride = a->res_bstride; d.res_tstride = a->res_tstride; d.res_scale = a->res_scale;
  d.res_bf16 = static_cast<const __nv_bfloat16*>(a->res_bf16);
  d.ln1_g = a->ln1_g; d.ln1_b = a->ln1_b; d.ln2_g = a->ln2_g; d.ln2_b = a->ln2_b; d.eps = a->eps;
  d.dot_w = a->dot_w; d.dot_b = a->dot_b; d.row_mask = a->row_mask;
  d.out_f32 = a->out_f32; d.out_bf16 = static_cast<__nv_bfloat16*>(a->out_bf16);
  d.out2_f32 = a->out2_f32; d.out_row = a->out_row;

  const int e = a->epi, bn = a->block_n;
  if (e == TTSB200_EPI_AFFINE_ACT) {
    TTSB_CHECK_ARG(a->out_f32 || a->out_bf16, "conv_gemm: AFFINE_ACT needs an output");
    if (a->res_f32)
      TTSB_CHECK_ARG(a->res_tstride % 4 == 0 && a->res_bstride % 4 == 0, "conv_gemm: res strides % 4");
    if (bn == 80) return launch_gemm<80, TTSB200_EPI_AFFINE_ACT>(*a, d, stream);
    if (bn == 128) return launch_gemm<128, TTSB200_EPI_AFFINE_ACT>(*a, d, stream);
    if (bn == 256) return launch_gemm<256, TTSB200_EPI_AFFINE_ACT>(*a, d, stream);
    if (bn == 384) return launch_gemm<384, TTSB200_EPI_AFFINE_ACT>(*a, d, stream);
  } else if (e == TTSB200_EPI_RES_LN) {
    TTSB_CHECK_ARG(bn == a->N, "conv_gemm: LN epilogue needs block_n == N");
    TTSB_CHECK_ARG(a->res_f32 && a->ln1_g && a->ln1_b && a->out_bf16, "conv_gemm: RES_LN operands");
    TTSB_CHECK_ARG(a->res_tstride % 4 == 0 && a->res_bstride % 4 == 0, "conv_gemm: res strides % 4");
    if (bn == 384) return launch_gemm<384, TTSB200_EPI_RES_LN>(*a, d, stream);
  } else if (e == TTSB200_EPI_FFN2) {
    TTSB_CHECK_ARG(bn == a->N, "conv_gemm: LN epilogue needs block_n == N");
    TTSB_CHECK_ARG(a->res_f32 && a->res_bf16 && a->ln1_g && a->ln1_b && a->ln2_g && a->ln2_b,
                   "conv_gemm: FFN2 operands");
    TTSB_CHECK_ARG(a->res_tstride % 4 == 0 && a->res_bstride % 4 == 0, "conv_gemm: res strides % 4");
    if (bn == 384) return launch_gemm<384, TTSB200_EPI_FFN2>(*a, d, stream);
  } else if (e == TTSB200_EPI_RELU_LN) {
    TTSB_CHECK_ARG(bn == a->N, "conv_gemm: LN epilogue needs block_n == N");
    TTSB_CHECK_ARG(a->ln1_g && a->ln1_b && a->out_bf16, "conv_gemm: RELU_LN operands");
    if (bn == 256) return launch_gemm<256, TTSB200_EPI_RELU_LN>(*a, d, stream);
  } else if (e == TTSB200_EPI_RELU_LN_DOT) {
    TTSB_CHECK_ARG(bn == a->N, "conv_gemm: LN epilogue needs block_n == N");
    TTSB_CHECK_ARG(a->ln1_g && a->ln1_b && a->dot_w && a->dot_b && a->out_row,
                   "conv_gemm: RELU_LN_DOT operands");
    if (bn == 256) return launch_gemm<256, TTSB200_EPI_RELU_LN_DOT>(*a, d, stream);
  }
  set_last_error("conv_gemm: unsupported (epi=%d, block_n=%d)", e, bn);
  return TTSB_ERR_ARG;
}
