// Fused masked multi-head self-attention for sm_100a (Models/modules.py:7-21 inside :43-70).
//
// One CTA = one (utterance, head, 128-query tile).  Per 128-key tile j:
//   TMA (SWIZZLE_64B boxes of 32 head-dims x 128 rows) -> smem K, V
//   tcgen05.mma  S_j = Q K_j^T        (M128 N128 K96, both operands K-major)        -> TMEM cols [0,128)
//   softmax warps: S_j TMEM -> 128 registers per thread (thread == query row), columns released at once so
//                  S_{j+1} is computed under the exponentials of tile j; scale, key-padding mask
//                  (-1e4 like the reference), p = exp2(s - m) with a lazily moved reference point m
//                  (O and l are rescaled only when the row maximum grew by more than 2^8),
//                  P (bf16) -> smem in the SWIZZLE_128B K-major layout
//   tcgen05.mma  O += P_j V_j         (M128 N96 K128, V is the MN-major B operand)  -> TMEM cols [128,224)
// O never leaves TMEM until the final O / l.  Two CTAs fit per SM (107 KB smem, 256 TMEM columns each)
// so one CTA's softmax overlaps the other's MMAs.  Fully padded key tiles are skipped: with the
// reference's -1e4 fill their probabilities underflow to exactly 0 in fp32 whenever the utterance has
// one valid key.  Key tiles inside the all-valid prefix of the mask (kv_info[b][1]) never touch the mask.
#include "common.cuh"
#include "../../include/ttsb200.h"

namespace ttsb {

constexpr int kDh = 96;
constexpr int kBQ = 128;
constexpr int kBKV = 128;
constexpr int kAttThreads = 192;
constexpr int kBoxBytes = 32 * 2 * 128;       // 32 dims x 128 rows bf16 = 8192
constexpr int kTileBytes = 3 * kBoxBytes;     // 96 dims
constexpr int kPBytes = kBQ * kBKV * 2;       // 32768
constexpr int kAttSmem = 3 * kTileBytes + kPBytes + 1024 /*align*/ + 128 /*barriers*/ + 512 /*flags*/;
constexpr int kAttTmemCols = 256;
constexpr float kLog2e = 1.4426950408889634f;

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

enum { BAR_Q = 0, BAR_KF, BAR_KE, BAR_VF, BAR_VE, BAR_SF, BAR_SE, BAR_PF, BAR_OF, BAR_COUNT };

__global__ void __launch_bounds__(kAttThreads, 2)
attention_fwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const uint8_t* __restrict__ mask,
                     const int32_t* __restrict__ kv_info, __nv_bfloat16* __restrict__ ctx,
                     float* __restrict__ lse2, int T, int H, float scale_log2) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + kTileBytes;
  uint8_t* sV = sK + kTileBytes;
  uint8_t* sP = sV + kTileBytes;  // offset 73728 = 72 * 1024
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + kPBytes);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + BAR_COUNT);
  uint8_t* sflag = reinterpret_cast<uint8_t*>(tmem_slot + 4);  // [2][128] key flags of a tile outside the clean prefix

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * kBQ;
  const int h = blockIdx.y;
  const int b = blockIdx.z;
  const int D = H * kDh;
  const int klen = kv_info[2 * b];          // 1 + index of the last valid key
  const int kprefix = kv_info[2 * b + 1];   // keys [0, kprefix) are all valid
  const int kv_eff = klen > 0 ? klen : T;   // no valid key: the reference degenerates to uniform over T
  const int nkv = (kv_eff + kBKV - 1) / kBKV;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQKV);
    for (int i = 0; i < BAR_COUNT; ++i) mbar_init(smem_u32(&bars[i]), (i == BAR_PF || i == BAR_SE) ? 128 : 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(tmem_slot), kAttTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_S = tmem_base;
  const uint32_t tmem_O = tmem_base + 128;

  if (warp == 0) {
    if (lane == 0) {
      const uint32_t qb = smem_u32(&bars[BAR_Q]);
      mbar_arrive_expect_tx(qb, kTileBytes);
#pragma unroll
      for (int i = 0; i < 3; ++i) tma_load_3d(smem_u32(sQ + i * kBoxBytes), &tmQKV, qb, h * kDh + i * 32, q0, b);
      for (int j = 0; j < nkv; ++j) {
        const uint32_t ph = j & 1;
        mbar_wait(smem_u32(&bars[BAR_KE]), ph ^ 1);
        const uint32_t kf = smem_u32(&bars[BAR_KF]);
        mbar_arrive_expect_tx(kf, kTileBytes);
#pragma unroll
        for (int i = 0; i < 3; ++i)
          tma_load_3d(smem_u32(sK + i * kBoxBytes), &tmQKV, kf, D + h * kDh + i * 32, j * kBKV, b);
        mbar_wait(smem_u32(&bars[BAR_VE]), ph ^ 1);
        const uint32_t vf = smem_u32(&bars[BAR_VF]);
        mbar_arrive_expect_tx(vf, kTileBytes);
#pragma unroll
        for (int i = 0; i < 3; ++i)
          tma_load_3d(smem_u32(sV + i * kBoxBytes), &tmQKV, vf, 2 * D + h * kDh + i * 32, j * kBKV, b);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(kBQ, kBKV, 0, 0);
      constexpr uint32_t idesc_o = umma_idesc_bf16(kBQ, kDh, 0, 1);  // B = V is MN-major
      const uint32_t q_base = smem_u32(sQ), k_base = smem_u32(sK), v_base = smem_u32(sV),
                     p_base = smem_u32(sP);
      // S_j = Q K_j^T is issued one tile AHEAD of O += P_j V_j: the softmax warps copy S_j into
      // registers and release the TMEM columns at once, so S_{j+1} runs under the exponentials of tile j.
      auto issue_s = [&](int j) {
        mbar_wait(smem_u32(&bars[BAR_KF]), j & 1);
        if (j > 0) mbar_wait(smem_u32(&bars[BAR_SE]), (j - 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int ks = 0; ks < kDh / 16; ++ks) {
          const uint32_t off = (ks >> 1) * kBoxBytes + (ks & 1) * 32;
          umma_bf16(tmem_S, umma_smem_desc(q_base + off, 16, 512, kLayoutSW64),
                    umma_smem_desc(k_base + off, 16, 512, kLayoutSW64), idesc_s, ks != 0);
        }
        umma_commit(smem_u32(&bars[BAR_KE]));
        umma_commit(smem_u32(&bars[BAR_SF]));
      };
      mbar_wait(smem_u32(&bars[BAR_Q]), 0);
      issue_s(0);
      for (int j = 0; j < nkv; ++j) {
        const uint32_t ph = j & 1;
        if (j + 1 < nkv) issue_s(j + 1);
        // ---- O += P_j V_j ----
        mbar_wait(smem_u32(&bars[BAR_PF]), ph);
        mbar_wait(smem_u32(&bars[BAR_VF]), ph);
        tc_fence_after();
#pragma unroll
        for (int ks = 0; ks < kBKV / 16; ++ks) {
          const uint64_t adesc =
              umma_smem_desc(p_base + (ks >> 2) * (kBQ * 128) + (ks & 3) * 32, 16, 1024, kLayoutSW128);
          const uint64_t bdesc = umma_smem_desc(v_base + ks * 1024, kBoxBytes, 512, kLayoutSW64);
          umma_bf16(tmem_O, adesc, bdesc, idesc_o, (j | ks) != 0);
        }
        umma_commit(smem_u32(&bars[BAR_VE]));
        umma_commit(smem_u32(&bars[BAR_OF]));
      }
    }
  } else {
    // ===================== softmax / output warps: thread == query row =====================
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int st = threadIdx.x - 64;  // 0..127 among softmax threads
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    const float masked = -1e4f * kLog2e;  // reference fill value (modules.py:14), in log2 units
    // m is the reference point of the exponentials, not necessarily the running maximum: it is only
    // moved (and O, l rescaled) when the maximum has grown by more than 2^8, so p <= 256 stays well
    // inside bf16/fp32 range and the final O / l is unchanged.
    float m = -INFINITY, l = 0.f;
    uint8_t* prow = sP + row * 128;

    for (int j = 0; j < nkv; ++j) {
      const uint32_t ph = j & 1;
      // a tile inside the all-valid prefix of the mask needs no flags (CTA-uniform decision)
      const bool clean = (j + 1) * kBKV <= kprefix;
      uint8_t* flg = sflag + ph * 128;
      if (!clean) {
        const int key = j * kBKV + st;
        flg[st] = key < T ? (mask[static_cast<int64_t>(b) * T + key] ? 1 : 0) : 2;
        asm volatile("bar.sync 1, 128;" ::: "memory");
      }
      mbar_wait(smem_u32(&bars[BAR_SF]), ph);
      tc_fence_after();
      float s[kBKV];
#pragma unroll
      for (int c = 0; c < kBKV; c += 32) tmem_ld32(tmem_S + lane_off + c, s + c);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(smem_u32(&bars[BAR_SE]));  // S columns may be overwritten by the next QK^T
      // scores in log2 units, masked like the reference (-1e4 fill; keys beyond T do not exist)
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
      if (clean) {
#pragma unroll
        for (int i = 0; i < kBKV; i += 4) {
          mx4[0] = fmaxf(mx4[0], s[i]); mx4[1] = fmaxf(mx4[1], s[i + 1]);
          mx4[2] = fmaxf(mx4[2], s[i + 2]); mx4[3] = fmaxf(mx4[3], s[i + 3]);
        }
      } else {
#pragma unroll
        for (int i = 0; i < kBKV; ++i) {
          const uint8_t f = flg[i];
          s[i] = f == 1 ? s[i] * scale_log2 : (f == 0 ? masked : -INFINITY);
          mx4[i & 3] = fmaxf(mx4[i & 3], s[i]);
        }
      }
      float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
      if (clean) mx *= scale_log2;  // the positive scale commutes with max
      // P_{j-1} V_{j-1} must have retired before sP is overwritten (and before O is rescaled)
      if (j > 0) {
        mbar_wait(smem_u32(&bars[BAR_OF]), (j - 1) & 1);
        tc_fence_after();
      }
      const bool grow = mx > m + 8.f;  // also true on the first tile (m = -inf)
      if (__any_sync(0xffffffffu, grow)) {
        const float m_new = fmaxf(m, mx);
        const float alpha = fast_exp2(m - m_new);  // m = -inf -> 0
        l *= alpha;
        m = m_new;
        if (j > 0) {
#pragma unroll
          for (int c = 0; c < kDh; c += 32) {
            float o[32];
            tmem_ld32(tmem_O + lane_off + c, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] *= alpha;
            tmem_st32(tmem_O + lane_off + c, o);
          }
          tmem_st_wait();
        }
      }
      // p = exp2(s - m) -> bf16 -> swizzled smem; fp32 row sum (four independent partial sums)
      float ls4[4] = {0.f, 0.f, 0.f, 0.f};
      const float neg_m = -m;
      const float sc = clean ? scale_log2 : 1.f;  // masked tiles were scaled while masking
#pragma unroll
      for (int c = 0; c < kBKV; c += 32) {
        float* v = s + c;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          v[i] = fast_exp2(fmaf(v[i], sc, neg_m));
          ls4[i & 3] += v[i];
        }
        uint8_t* pblk = prow + (c >> 6) * (kBQ * 128);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int ch = ((c & 63) >> 3) + g;  // 16-byte chunk index inside the 128-byte row
          uint4 u;
          u.x = pack_bf16(v[g * 8 + 0], v[g * 8 + 1]);
          u.y = pack_bf16(v[g * 8 + 2], v[g * 8 + 3]);
          u.z = pack_bf16(v[g * 8 + 4], v[g * 8 + 5]);
          u.w = pack_bf16(v[g * 8 + 6], v[g * 8 + 7]);
          *reinterpret_cast<uint4*>(pblk + ((ch ^ (row & 7)) << 4)) = u;
        }
      }
      l += (ls4[0] + ls4[1]) + (ls4[2] + ls4[3]);
      fence_async_smem();  // generic-proxy smem writes -> visible to the tensor core (async proxy)
      tc_fence_before();   // orders the O rescale (tcgen05.st) before the MMA that accumulates into O
      mbar_arrive(smem_u32(&bars[BAR_PF]));
    }
    // ---- O / l ----
    mbar_wait(smem_u32(&bars[BAR_OF]), (nkv - 1) & 1);
    tc_fence_after();
    const int t = q0 + row;
    const float inv = 1.f / l;
    // training: log2-domain log-sum-exp of the row, p = exp2(s*scale_log2 - lse2) in the backward kernels
    if (lse2 != nullptr && t < T) lse2[(static_cast<int64_t>(b) * H + h) * T + t] = m + log2f(l);
    __nv_bfloat16* op = ctx + (static_cast<int64_t>(b) * T + (t < T ? t : 0)) * D + h * kDh;
#pragma unroll
    for (int c = 0; c < kDh; c += 32) {
      float o[32];
      tmem_ld32(tmem_O + lane_off + c, o);
      tmem_ld_wait();
      if (t < T) {
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          uint4 u;
          u.x = pack_bf16(o[i] * inv, o[i + 1] * inv);
          u.y = pack_bf16(o[i + 2] * inv, o[i + 3] * inv);
          u.z = pack_bf16(o[i + 4] * inv, o[i + 5] * inv);
          u.w = pack_bf16(o[i + 6] * inv, o[i + 7] * inv);
          *reinterpret_cast<uint4*>(op + c + i) = u;
        }
      }
    }
    tc_fence_before();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, kAttTmemCols);
}

// ------------------------------------------------------------------------------------------------
// Debug / parity: materialise the (B,H,T,T) probabilities the reference returns.  CUDA cores only.
// One warp per (b,h,query); lanes stride over keys.
// ------------------------------------------------------------------------------------------------
__global__ void attention_probs_kernel(const __nv_bfloat16* __restrict__ qkv, const uint8_t* __restrict__ mask,
                                       float* __restrict__ probs, int B, int T, int H, int Dh) {
  extern __shared__ float sq[];  // [warps][Dh]
  const int lane = threadIdx.x & 31;
  const int wl = threadIdx.x >> 5;
  const int64_t gw = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + wl;
  const int64_t total = static_cast<int64_t>(B) * H * T;
  if (gw >= total) return;
  const int tq = static_cast<int>(gw % T);
  const int h = static_cast<int>((gw / T) % H);
  const int b = static_cast<int>(gw / (static_cast<int64_t>(T) * H));
  const int D = H * Dh;
  const int64_t ld = 3LL * D;
  float* q = sq + wl * Dh;
  const __nv_bfloat16* qp = qkv + (static_cast<int64_t>(b) * T + tq) * ld + h * Dh;
  for (int i = lane; i < Dh; i += 32) q[i] = __bfloat162float(qp[i]);
  __syncwarp();
  float* pr = probs + ((static_cast<int64_t>(b) * H + h) * T + tq) * T;
  const float scale = rsqrtf(static_cast<float>(Dh));
  float mx = -INFINITY;
  for (int k = lane; k < T; k += 32) {
    const __nv_bfloat16* kp = qkv + (static_cast<int64_t>(b) * T + k) * ld + D + h * Dh;
    float s = 0.f;
    for (int i = 0; i < Dh; ++i) s = fmaf(q[i], __bfloat162float(kp[i]), s);
    s *= scale;
    if (mask[static_cast<int64_t>(b) * T + k] == 0) s = -1e4f;
    pr[k] = s;
    mx = fmaxf(mx, s);
  }
  mx = warp_max(mx);
  float sum = 0.f;
  for (int k = lane; k < T; k += 32) {
    const float e = expf(pr[k] - mx);
    pr[k] = e;
    sum += e;
  }
  sum = warp_sum(sum);
  const float inv = 1.f / sum;
  for (int k = lane; k < T; k += 32) pr[k] *= inv;
}

}  // namespace ttsb

using namespace ttsb;

extern "C" int ttsb200_attention_fwd_lse(const void* qkv, const uint8_t* mask, const int32_t* kv_info, void* ctx,
                                         float* lse2, int32_t B, int32_t T, int32_t H, int32_t Dh,
                                         ttsb200_stream_t s) {
  TTSB_CHECK_ARG(qkv && mask && kv_info && ctx, "attention_fwd: null pointer");
  TTSB_CHECK_ARG(B > 0 && T > 0 && H > 0, "attention_fwd: bad dims");
  TTSB_CHECK_ARG(Dh == kDh, "attention_fwd: head dim %d unsupported (this build: %d)", Dh, kDh);
  TTSB_CHECK_ARG((reinterpret_cast<uintptr_t>(qkv) & 15) == 0 && (reinterpret_cast<uintptr_t>(ctx) & 15) == 0,
                 "attention_fwd: qkv/ctx must be 16-byte aligned");
  CUtensorMap tm;
  const uint64_t ld = 3ull * H * Dh;
  const uint64_t dims[3] = {ld, static_cast<uint64_t>(T), static_cast<uint64_t>(B)};
  const uint64_t strides[2] = {ld * 2, ld * 2 * static_cast<uint64_t>(T)};
  const uint32_t box[3] = {32, 128, 1};
  int rc = make_tmap_bf16(&tm, qkv, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_64B);
  if (rc) return rc;
  static bool attr_done = false;
  if (!attr_done) {
    TTSB_CHECK_CUDA(cudaFuncSetAttribute(attention_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kAttSmem));
    attr_done = true;
  }
  dim3 grid((T + kBQ - 1) / kBQ, H, B);
  const float scale_log2 = kLog2e / sqrtf(static_cast<float>(Dh));
  attention_fwd_kernel<<<grid, kAttThreads, kAttSmem, static_cast<cudaStream_t>(s)>>>(
      tm, mask, kv_info, static_cast<__nv_bfloat16*>(ctx), lse2, T, H, scale_log2);
  TTSB_CHECK_CUDA(cudaGetLastError());
  return TTSB_OK;
}

extern "C" int ttsb200_attention_fwd(const void* qkv, const uint8_t* mask, const int32_t* kv_info, void* ctx,
                                     int32_t B, int32_t T, int32_t H, int32_t Dh, ttsb200_stream_t s) {
  return ttsb200_attention_fwd_lse(qkv, mask, kv_info, ctx, nullptr, B, T, H, Dh, s);
}

extern "C" int ttsb200_attention_probs(const void* qkv, const uint8_t* mask, float* probs, int32_t B,
                                       int32_t T, int32_t H, int32_t Dh, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(qkv && mask && probs && B > 0 && T > 0 && H > 0 && Dh > 0, "attention_probs: bad args");
  const int64_t total = static_cast<int64_t>(B) * H * T;
  const int warps = 8;
  attention_probs_kernel<<<static_cast<unsigned>((total + warps - 1) / warps), warps * 32,
                           warps * Dh * sizeof(float), static_cast<cudaStream_t>(s)>>>(
      static_cast<const __nv_bfloat16*>(qkv), mask, probs, B, T, H, Dh);
  TTSB_CHECK_CUDA(cudaGetLastError());
  return TTSB_OK;
}
