// Backward of the fused masked multi-head self-attention for sm_100a (training path).
// The reference obtains it from PyTorch autograd through Models/modules.py:7-21 (matmul, masked_fill,
// softmax, matmul); here it is two tcgen05 kernels that recompute the probabilities from the saved
// log-sum-exp (flash-attention style), so the (B,H,T,T) score tensors never exist in HBM:
//
//   p   = exp2(s * log2(e)/sqrt(Dh) - lse2)         s = q.k, masked keys use the reference's -1e4 fill
//   dp  = dO . v
//   ds  = p * (dp - delta) / sqrt(Dh)               delta = rowsum(dO * O); masked keys: ds = 0 (masked_fill)
//   dV  = P^T dO      dK = dS^T Q      dQ = dS K
//
// attn_bwd_dkv_kernel: one CTA = (utterance, head, 128-key tile); K, V stay in smem, the CTA walks all query
//   tiles; S and dP are formed in TMEM, the softmax warps (thread == query row) write P and dS (bf16) to smem in
//   the SWIZZLE_128B [query][key] layout, which the tensor core then reads TRANSPOSED through MN-major
//   descriptors (dV += P^T dO, dK += dS^T Q; dO and Q are the MN-major B operands).  dV, dK accumulate in TMEM.
// attn_bwd_dq_kernel: one CTA = (utterance, head, 128-query tile); walks the key tiles like the forward kernel,
//   dQ += dS K_j accumulates in TMEM (dS K-major A operand, K_j MN-major B operand).
// Two kernels (7 GEMMs instead of 5) keep every accumulation on chip and deterministic: no fp32 atomics on dQ.
#include "common.cuh"
#include "../../include/ttsb200.h"

namespace ttsb {

namespace ab {
constexpr int kDh = 96;
constexpr int kBQ = 128;
constexpr int kBKV = 128;
constexpr int kThreads = 320;  // TMA warp, MMA warp, 8 softmax warps (two per TMEM lane quarter: 64 keys each)
constexpr int kSoftmaxThreads = 256;
constexpr int kBoxBytes = 32 * 2 * 128;    // 32 dims x 128 rows bf16
constexpr int kTileBytes = 3 * kBoxBytes;  // 96 dims = 24 KB
constexpr int kPBytes = kBQ * kBKV * 2;    // 32 KB
constexpr float kLog2e = 1.4426950408889634f;
// dkv: K, V, 2 x (Q, dO), P, dS
constexpr int kSmemDkv = 6 * kTileBytes + 2 * kPBytes + 1024 + 256 + 128;
// dq : Q, dO, 2 x (K, V), dS
constexpr int kSmemDq = 6 * kTileBytes + kPBytes + 1024 + 256 + 256;

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// 32 bf16 values -> this thread's row of the SWIZZLE_128B [row][key] tile, columns [c, c+32)
__device__ __forceinline__ void store_row32(uint8_t* tile, int row, int c, const float* v) {
  uint8_t* pblk = tile + row * 128 + (c >> 6) * (kBQ * 128);
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    const int ch = ((c & 63) >> 3) + g;
    uint4 u;
    u.x = pack_bf16(v[g * 8 + 0], v[g * 8 + 1]);
    u.y = pack_bf16(v[g * 8 + 2], v[g * 8 + 3]);
    u.z = pack_bf16(v[g * 8 + 4], v[g * 8 + 5]);
    u.w = pack_bf16(v[g * 8 + 6], v[g * 8 + 7]);
    *reinterpret_cast<uint4*>(pblk + ((ch ^ (row & 7)) << 4)) = u;
  }
}

// p and ds of 32 keys for one query row.  flg: 1 valid key, 0 masked key (< T), 2 key beyond T.
// vis: keys [0, vis] of these 32 are visible to the row (causal mask of Transformer-TTS: a key beyond the query's own index
// is filled like a padded one); pass a large value when there is no causal mask.
__device__ __forceinline__ void softmax_bwd32(float* s /* in: s, out: p */, float* dp /* in: dp, out: ds */,
                                              const uint8_t* flg, bool row_ok, float lse2, float delta,
                                              float scale_log2, float scale, int vis = 1 << 30) {
  const float masked = -1e4f * kLog2e;
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    const uint8_t f = flg[i];
    const bool v = f == 1 && i <= vis;
    const float sl = v ? s[i] * scale_log2 : masked;
    float p = ex2(sl - lse2);
    if (f == 2 || !row_ok) p = 0.f;
    s[i] = p;
    dp[i] = v ? p * (dp[i] - delta) * scale : 0.f;
  }
}

// The same for a tile whose keys are all valid (the common case): packed f32x2 arithmetic, no flags.
//   neg_l2 = -lse2 (+inf -> p = 0 for query rows beyond T), neg_dsc = -delta * scale
__device__ __forceinline__ void softmax_bwd32_clean(float* s, float* dp, float neg_l2, float neg_dsc, float scale_log2,
                                                    float scale) {
#pragma unroll
  for (int i = 0; i < 32; i += 2) {
    fma2(s[i], s[i + 1], scale_log2, scale_log2, neg_l2, neg_l2);
    s[i] = ex2(s[i]);
    s[i + 1] = ex2(s[i + 1]);
    fma2(dp[i], dp[i + 1], scale, scale, neg_dsc, neg_dsc);
    mul2(dp[i], dp[i + 1], s[i], s[i + 1]);
  }
}

enum { D_KV = 0, D_QF0, D_QF1, D_QE0, D_QE1, D_SF, D_PF, D_DONE, D_COUNT };

// CAUSAL (its own instantiation): key k reaches query t only for k <= t, so a key tile starts at the query tile that
// holds its diagonal (i0) and only that tile needs the per-row predicate.
template <bool CAUSAL>
__global__ void __launch_bounds__(kThreads, 1)
attn_bwd_dkv_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
                    const uint8_t* __restrict__ mask, const int32_t* __restrict__ kv_info,
                    const float* __restrict__ lse2, const float* __restrict__ delta,
                    __nv_bfloat16* __restrict__ dqkv, int T, int H, float scale_log2, float scale,
                    const DropCfg drop) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* sK = smem;
  uint8_t* sV = sK + kTileBytes;
  uint8_t* sQ = sV + kTileBytes;        // [2]
  uint8_t* sDO = sQ + 2 * kTileBytes;   // [2]
  uint8_t* sP = sDO + 2 * kTileBytes;   // 6 * 24 KB = 144 KB, 1024-aligned
  uint8_t* sDS = sP + kPBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDS + kPBytes);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + D_COUNT);
  uint8_t* sflag = reinterpret_cast<uint8_t*>(tmem_slot + 4);  // [128]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k0 = blockIdx.x * kBKV;
  const int h = blockIdx.y, b = blockIdx.z;
  const int D = H * kDh;
  const int klen = kv_info[2 * b];
  const int kv_eff = klen > 0 ? klen : T;
  const bool clean = k0 + kBKV <= kv_info[2 * b + 1];  // every key of this tile is valid: no flags needed
  const int nq_all = (T + kBQ - 1) / kBQ;
  const int i0 = CAUSAL ? k0 / kBQ : 0;     // first query tile that sees this key tile
  const int nq = nq_all - i0;               // query tiles processed (local index n, tile index i0 + n)
  const int64_t ld = 3LL * D;

  if (k0 >= kv_eff) {
    // fully padded key tile: probabilities are exactly 0 -> zero gradients (uniform over the CTA, no barriers yet)
    for (int i = threadIdx.x; i < kBKV * 24; i += kThreads) {
      const int r = i / 24, c = i - r * 24;
      if (k0 + r < T) {
        __nv_bfloat16* base = dqkv + (static_cast<int64_t>(b) * T + k0 + r) * ld + h * kDh + c * 4;
        *reinterpret_cast<uint2*>(base + D) = make_uint2(0, 0);
        *reinterpret_cast<uint2*>(base + 2 * D) = make_uint2(0, 0);
      }
    }
    return;
  }

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQKV);
    tma_prefetch_desc(&tmDO);
    for (int i = 0; i < D_COUNT; ++i) mbar_init(smem_u32(&bars[i]), i == D_PF ? kSoftmaxThreads : 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(tmem_slot), 512);
    tmem_relinquish();
  }
  if (threadIdx.x >= 64 && threadIdx.x < 64 + kBKV) {
    const int st = threadIdx.x - 64;
    const int key = k0 + st;
    sflag[st] = key < T ? (mask[static_cast<int64_t>(b) * T + key] ? 1 : 0) : 2;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_S = tmem_base, tmem_dP = tmem_base + 128, tmem_dV = tmem_base + 256,
                 tmem_dK = tmem_base + 384;

  if (warp == 0) {
    if (lane == 0) {
      const uint32_t kvb = smem_u32(&bars[D_KV]);
      mbar_arrive_expect_tx(kvb, 2 * kTileBytes);
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        tma_load_3d(smem_u32(sK + i * kBoxBytes), &tmQKV, kvb, D + h * kDh + i * 32, k0, b);
        tma_load_3d(smem_u32(sV + i * kBoxBytes), &tmQKV, kvb, 2 * D + h * kDh + i * 32, k0, b);
      }
      for (int i = 0; i < nq; ++i) {
        const int buf = i & 1;
        mbar_wait(smem_u32(&bars[D_QE0 + buf]), ((i >> 1) & 1) ^ 1);
        const uint32_t qf = smem_u32(&bars[D_QF0 + buf]);
        mbar_arrive_expect_tx(qf, 2 * kTileBytes);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          tma_load_3d(smem_u32(sQ + buf * kTileBytes + c * kBoxBytes), &tmQKV, qf, h * kDh + c * 32, (i0 + i) * kBQ, b);
          tma_load_3d(smem_u32(sDO + buf * kTileBytes + c * kBoxBytes), &tmDO, qf, h * kDh + c * 32, (i0 + i) * kBQ, b);
        }
      }
    }
  } else if (warp == 1) {
    // whole warp converged, tcgen05 instructions predicated on one elected lane: operands stay in uniform registers
    // (under `if (lane == 0)` every UTCHMMA is wrapped in an ELECT / BRA.U.ANY waterfall loop)
    constexpr uint32_t idesc_s = umma_idesc_bf16(kBQ, kBKV, 0, 0);
    constexpr uint32_t idesc_t = umma_idesc_bf16(kBKV, kDh, 1, 1);  // A = P^T / dS^T (MN-major), B = dO / Q (MN-major)
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t k_base = smem_u32(sK), v_base = smem_u32(sV), p_base = smem_u32(sP), ds_base = smem_u32(sDS);
    const uint32_t leader = elect_one();
    mbar_wait(smem_u32(&bars[D_KV]), 0);
    for (int i = 0; i < nq; ++i) {
      const int buf = i & 1;
      const uint32_t q_base = smem_u32(sQ + buf * kTileBytes), do_base = smem_u32(sDO + buf * kTileBytes);
      mbar_wait(smem_u32(&bars[D_QF0 + buf]), (i >> 1) & 1);
      tc_fence_after();
      // S = Q K^T, dP = dO V^T  (the softmax warps released these columns with PF of the previous tile)
      if (leader) {
#pragma unroll
        for (int ks = 0; ks < kDh / 16; ++ks) {
          const uint32_t off = (ks >> 1) * kBoxBytes + (ks & 1) * 32;
          umma_bf16(tb, umma_smem_desc(q_base + off, 16, 512, kLayoutSW64),
                    umma_smem_desc(k_base + off, 16, 512, kLayoutSW64), idesc_s, ks != 0);
        }
#pragma unroll
        for (int ks = 0; ks < kDh / 16; ++ks) {
          const uint32_t off = (ks >> 1) * kBoxBytes + (ks & 1) * 32;
          umma_bf16(tb + 128, umma_smem_desc(do_base + off, 16, 512, kLayoutSW64),
                    umma_smem_desc(v_base + off, 16, 512, kLayoutSW64), idesc_s, ks != 0);
        }
        umma_commit(smem_u32(&bars[D_SF]));
      }
      __syncwarp();
      mbar_wait(smem_u32(&bars[D_PF]), i & 1);
      tc_fence_after();
      // dV += P^T dO ; dK += dS^T Q   (reduction over the 128 queries, 16 per MMA)
      if (leader) {
#pragma unroll
        for (int ks = 0; ks < kBQ / 16; ++ks) {
          const uint64_t a_p = umma_smem_desc(p_base + ks * 2048, kBQ * 128, 1024, kLayoutSW128);
          const uint64_t b_do = umma_smem_desc(do_base + ks * 1024, kBoxBytes, 512, kLayoutSW64);
          umma_bf16(tb + 256, a_p, b_do, idesc_t, (i | ks) != 0);
        }
#pragma unroll
        for (int ks = 0; ks < kBQ / 16; ++ks) {
          const uint64_t a_ds = umma_smem_desc(ds_base + ks * 2048, kBQ * 128, 1024, kLayoutSW128);
          const uint64_t b_q = umma_smem_desc(q_base + ks * 1024, kBoxBytes, 512, kLayoutSW64);
          umma_bf16(tb + 384, a_ds, b_q, idesc_t, (i | ks) != 0);
        }
        umma_commit(smem_u32(&bars[D_QE0 + buf]));
      }
      __syncwarp();
    }
    if (leader) umma_commit(smem_u32(&bars[D_DONE]));
    __syncwarp();
  } else {
    // eight softmax warps: warp % 4 is the TMEM lane quarter (query rows), `half` the 64-key half of the tile.  With
    // one warp per scheduler the recompute was latency-bound; two per scheduler hide each other's stalls.
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    const int row = q * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    const int64_t stat_base = (static_cast<int64_t>(b) * H + h) * T;
    for (int i = 0; i < nq; ++i) {
      const int t = (i0 + i) * kBQ + row;
      const bool row_ok = t < T;
      const bool diag = CAUSAL && i == 0;   // the tile that holds this key tile's diagonal: per-row visibility
      const float l2 = row_ok ? __ldg(lse2 + stat_base + t) : 0.f;
      const float dl = row_ok ? __ldg(delta + stat_base + t) : 0.f;
      const float neg_l2 = row_ok ? -l2 : -INFINITY, neg_dsc = -dl * scale;
      // SF(i) also tells that the dV/dK MMAs of tile i-1 have retired (same in-order pipe): sP/sDS are free
      mbar_wait(smem_u32(&bars[D_SF]), i & 1);
      tc_fence_after();
#pragma unroll 1
      for (int c = half * 64; c < half * 64 + 64; c += 32) {
        float s[32], dp[32];
        tmem_ld32(tmem_S + lane_off + c, s);
        tmem_ld32(tmem_dP + lane_off + c, dp);
        tmem_ld_wait();
        // dropout on the probabilities (Models/modules.py:18-19): ctx = (P * m) V, so dP := (dO V^T) * m before
        // dS = P * (dP - delta), and dV += (P * m)^T dO
        float mk[32];
        if (drop.thresh) {
          const uint64_t e0 = (static_cast<uint64_t>(stat_base) + static_cast<uint64_t>(row_ok ? t : 0)) * T + k0 + c;
#pragma unroll
          for (int e = 0; e < 32; ++e) {
            mk[e] = drop_scale(drop, e0 + e);
            dp[e] *= mk[e];
          }
        }
        if (clean && !diag) softmax_bwd32_clean(s, dp, neg_l2, neg_dsc, scale_log2, scale);
        else softmax_bwd32(s, dp, sflag + c, row_ok, l2, dl, scale_log2, scale, diag ? t - k0 - c : (1 << 30));
        if (drop.thresh) {
#pragma unroll
          for (int e = 0; e < 32; ++e) s[e] *= mk[e];
        }
        store_row32(sP, row, c, s);
        store_row32(sDS, row, c, dp);
      }
      fence_async_smem();
      tc_fence_before();
      mbar_arrive(smem_u32(&bars[D_PF]));
    }
    // ---- epilogue: thread == key row; dK at [.., D + h*Dh), dV at [.., 2D + h*Dh) ----
    mbar_wait(smem_u32(&bars[D_DONE]), 0);
    tc_fence_after();
    const int key = k0 + row;
    __nv_bfloat16* base = dqkv + (static_cast<int64_t>(b) * T + (key < T ? key : 0)) * ld + h * kDh;
    {
      const int which = half;  // the two warps of a lane quarter write dK and dV side by side
      const uint32_t tm = which == 0 ? tmem_dK : tmem_dV;
      __nv_bfloat16* op = base + (which == 0 ? D : 2 * D);
#pragma unroll
      for (int c = 0; c < kDh; c += 32) {
        float o[32];
        tmem_ld32(tm + lane_off + c, o);
        tmem_ld_wait();
        if (key < T) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            uint4 u;
            u.x = pack_bf16(o[j], o[j + 1]); u.y = pack_bf16(o[j + 2], o[j + 3]);
            u.z = pack_bf16(o[j + 4], o[j + 5]); u.w = pack_bf16(o[j + 6], o[j + 7]);
            *reinterpret_cast<uint4*>(op + c + j) = u;
          }
        }
      }
    }
    tc_fence_before();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

enum { Q_Q = 0, Q_KF0, Q_KF1, Q_KE0, Q_KE1, Q_SF, Q_PF, Q_DONE, Q_COUNT };

template <bool CAUSAL>
__global__ void __launch_bounds__(kThreads, 1)
attn_bwd_dq_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
                   const uint8_t* __restrict__ mask, const int32_t* __restrict__ kv_info,
                   const float* __restrict__ lse2, const float* __restrict__ delta,
                   __nv_bfloat16* __restrict__ dqkv, int T, int H, float scale_log2, float scale,
                   const DropCfg drop) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* sQ = smem;
  uint8_t* sDO = sQ + kTileBytes;
  uint8_t* sK = sDO + kTileBytes;      // [2]
  uint8_t* sV = sK + 2 * kTileBytes;   // [2]
  uint8_t* sDS = sV + 2 * kTileBytes;  // 144 KB offset, 1024-aligned
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDS + kPBytes);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + Q_COUNT);
  uint8_t* sflag = reinterpret_cast<uint8_t*>(tmem_slot + 4);  // [2][128]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * kBQ;
  const int h = blockIdx.y, b = blockIdx.z;
  const int D = H * kDh;
  const int klen = kv_info[2 * b];
  const int kv_eff = klen > 0 ? klen : T;
  // CAUSAL: key tiles above this query tile's diagonal contribute nothing and are never loaded
  const int nkv = CAUSAL ? min((kv_eff + kBKV - 1) / kBKV, (q0 + kBQ - 1) / kBKV + 1) : (kv_eff + kBKV - 1) / kBKV;
  const int64_t ld = 3LL * D;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQKV);
    tma_prefetch_desc(&tmDO);
    for (int i = 0; i < Q_COUNT; ++i) mbar_init(smem_u32(&bars[i]), i == Q_PF ? kSoftmaxThreads : 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(tmem_slot), 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_S = tmem_base, tmem_dP = tmem_base + 128, tmem_dQ = tmem_base + 256;

  if (warp == 0) {
    if (lane == 0) {
      const uint32_t qb = smem_u32(&bars[Q_Q]);
      mbar_arrive_expect_tx(qb, 2 * kTileBytes);
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        tma_load_3d(smem_u32(sQ + c * kBoxBytes), &tmQKV, qb, h * kDh + c * 32, q0, b);
        tma_load_3d(smem_u32(sDO + c * kBoxBytes), &tmDO, qb, h * kDh + c * 32, q0, b);
      }
      for (int j = 0; j < nkv; ++j) {
        const int buf = j & 1;
        mbar_wait(smem_u32(&bars[Q_KE0 + buf]), ((j >> 1) & 1) ^ 1);
        const uint32_t kf = smem_u32(&bars[Q_KF0 + buf]);
        mbar_arrive_expect_tx(kf, 2 * kTileBytes);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          tma_load_3d(smem_u32(sK + buf * kTileBytes + c * kBoxBytes), &tmQKV, kf, D + h * kDh + c * 32, j * kBKV, b);
          tma_load_3d(smem_u32(sV + buf * kTileBytes + c * kBoxBytes), &tmQKV, kf, 2 * D + h * kDh + c * 32,
                      j * kBKV, b);
        }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc_s = umma_idesc_bf16(kBQ, kBKV, 0, 0);
    constexpr uint32_t idesc_q = umma_idesc_bf16(kBQ, kDh, 0, 1);  // A = dS (K-major), B = K_j (MN-major)
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t q_base = smem_u32(sQ), do_base = smem_u32(sDO), ds_base = smem_u32(sDS);
    const uint32_t leader = elect_one();
    mbar_wait(smem_u32(&bars[Q_Q]), 0);
    for (int j = 0; j < nkv; ++j) {
      const int buf = j & 1;
      const uint32_t k_base = smem_u32(sK + buf * kTileBytes), v_base = smem_u32(sV + buf * kTileBytes);
      mbar_wait(smem_u32(&bars[Q_KF0 + buf]), (j >> 1) & 1);
      tc_fence_after();
      if (leader) {
#pragma unroll
        for (int ks = 0; ks < kDh / 16; ++ks) {
          const uint32_t off = (ks >> 1) * kBoxBytes + (ks & 1) * 32;
          umma_bf16(tb, umma_smem_desc(q_base + off, 16, 512, kLayoutSW64),
                    umma_smem_desc(k_base + off, 16, 512, kLayoutSW64), idesc_s, ks != 0);
        }
#pragma unroll
        for (int ks = 0; ks < kDh / 16; ++ks) {
          const uint32_t off = (ks >> 1) * kBoxBytes + (ks & 1) * 32;
          umma_bf16(tb + 128, umma_smem_desc(do_base + off, 16, 512, kLayoutSW64),
                    umma_smem_desc(v_base + off, 16, 512, kLayoutSW64), idesc_s, ks != 0);
        }
        umma_commit(smem_u32(&bars[Q_SF]));
      }
      __syncwarp();
      mbar_wait(smem_u32(&bars[Q_PF]), j & 1);
      tc_fence_after();
      if (leader) {
#pragma unroll
        for (int ks = 0; ks < kBKV / 16; ++ks) {
          const uint64_t adesc =
              umma_smem_desc(ds_base + (ks >> 2) * (kBQ * 128) + (ks & 3) * 32, 16, 1024, kLayoutSW128);
          const uint64_t bdesc = umma_smem_desc(k_base + ks * 1024, kBoxBytes, 512, kLayoutSW64);
          umma_bf16(tb + 256, adesc, bdesc, idesc_q, (j | ks) != 0);
        }
        umma_commit(smem_u32(&bars[Q_KE0 + buf]));
      }
      __syncwarp();
    }
    if (leader) umma_commit(smem_u32(&bars[Q_DONE]));
    __syncwarp();
  } else {
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;  // as in the dKV kernel: two warps per lane quarter, 64 keys each
    const int row = q * 32 + lane;
    const int st = threadIdx.x - 64;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    const int t = q0 + row;
    const bool row_ok = t < T;
    const int64_t stat_base = (static_cast<int64_t>(b) * H + h) * T;
    const float l2 = row_ok ? __ldg(lse2 + stat_base + t) : 0.f;
    const float dl = row_ok ? __ldg(delta + stat_base + t) : 0.f;
    const float neg_l2 = row_ok ? -l2 : -INFINITY, neg_dsc = -dl * scale;
    const int kprefix = kv_info[2 * b + 1];
    for (int j = 0; j < nkv; ++j) {
      const bool diag = CAUSAL && (j + 1) * kBKV > q0 + 1;   // the tile crosses the diagonal: per-row visibility
      const bool clean = (j + 1) * kBKV <= kprefix && !diag;    // CTA-uniform
      uint8_t* flg = sflag + (j & 1) * 128;
      if (!clean) {
        if (st < kBKV) {
          const int key = j * kBKV + st;
          flg[st] = key < T ? (mask[static_cast<int64_t>(b) * T + key] ? 1 : 0) : 2;
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");
      }
      mbar_wait(smem_u32(&bars[Q_SF]), j & 1);  // also: the dQ MMAs of tile j-1 have retired, sDS is free
      tc_fence_after();
#pragma unroll 1
      for (int c = half * 64; c < half * 64 + 64; c += 32) {
        float s[32], dp[32];
        tmem_ld32(tmem_S + lane_off + c, s);
        tmem_ld32(tmem_dP + lane_off + c, dp);
        tmem_ld_wait();
        if (drop.thresh) {  // dP := (dO V^T) * m, as in the dKV kernel
          const uint64_t e0 = (static_cast<uint64_t>(stat_base) + static_cast<uint64_t>(row_ok ? t : 0)) * T + j * kBKV + c;
#pragma unroll
          for (int e = 0; e < 32; ++e) dp[e] *= drop_scale(drop, e0 + e);
        }
        if (clean) softmax_bwd32_clean(s, dp, neg_l2, neg_dsc, scale_log2, scale);
        else softmax_bwd32(s, dp, flg + c, row_ok, l2, dl, scale_log2, scale, diag ? t - j * kBKV - c : (1 << 30));
        store_row32(sDS, row, c, dp);
      }
      fence_async_smem();
      tc_fence_before();
      mbar_arrive(smem_u32(&bars[Q_PF]));
    }
    mbar_wait(smem_u32(&bars[Q_DONE]), 0);
    tc_fence_after();
    __nv_bfloat16* op = dqkv + (static_cast<int64_t>(b) * T + (row_ok ? t : 0)) * ld + h * kDh;
#pragma unroll
    for (int c = 0; c < kDh; c += 32) {
      if ((c == 32 ? 1 : 0) != half) continue;  // warp-uniform split: half 0 writes columns 0-31 and 64-95, half 1 32-63
      float o[32];
      tmem_ld32(tmem_dQ + lane_off + c, o);
      tmem_ld_wait();
      if (row_ok) {
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
          uint4 u;
          u.x = pack_bf16(o[j], o[j + 1]); u.y = pack_bf16(o[j + 2], o[j + 3]);
          u.z = pack_bf16(o[j + 4], o[j + 5]); u.w = pack_bf16(o[j + 6], o[j + 7]);
          *reinterpret_cast<uint4*>(op + c + j) = u;
        }
      }
    }
    tc_fence_before();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 512);
}

}  // namespace ab
}  // namespace ttsb

using namespace ttsb;

extern "C" int ttsb200_attention_bwd(const void* qkv, const void* dctx, const uint8_t* mask, const int32_t* kv_info,
                                     const float* lse2, const float* delta, void* dqkv, int32_t B, int32_t T,
                                     int32_t H, int32_t Dh, ttsb200_stream_t s) {
  return ttsb200_attention_bwd_drop(qkv, dctx, mask, kv_info, lse2, delta, dqkv, B, T, H, Dh, 0.f, 0, -1, s);
}

static int attention_bwd_launch(const void* qkv, const void* dctx, const uint8_t* mask, const int32_t* kv_info,
                                const float* lse2, const float* delta, void* dqkv, int32_t B, int32_t T, int32_t H,
                                int32_t Dh, const DropCfg drop, bool causal, ttsb200_stream_t s);

extern "C" int ttsb200_attention_bwd_drop(const void* qkv, const void* dctx, const uint8_t* mask,
                                          const int32_t* kv_info, const float* lse2, const float* delta, void* dqkv,
                                          int32_t B, int32_t T, int32_t H, int32_t Dh, float drop_p, uint64_t drop_seed,
                                          int64_t drop_site, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f, "attention_bwd: dropout probability must be in [0, 1)");
  return attention_bwd_launch(qkv, dctx, mask, kv_info, lse2, delta, dqkv, B, T, H, Dh, make_drop(drop_p, drop_seed, drop_site),
                              false, s);
}

extern "C" int ttsb200_attention_bwd_causal(const void* qkv, const void* dctx, const uint8_t* mask, const int32_t* kv_info,
                                            const float* lse2, const float* delta, void* dqkv, int32_t B, int32_t T,
                                            int32_t H, int32_t Dh, ttsb200_stream_t s) {
  return attention_bwd_launch(qkv, dctx, mask, kv_info, lse2, delta, dqkv, B, T, H, Dh, make_drop(0.f, 0, -1), true, s);
}

extern "C" int ttsb200_attention_bwd_causal_drop(const void* qkv, const void* dctx, const uint8_t* mask,
                                                 const int32_t* kv_info, const float* lse2, const float* delta, void* dqkv,
                                                 int32_t B, int32_t T, int32_t H, int32_t Dh, float drop_p, uint64_t drop_seed,
                                                 int64_t drop_site, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f, "attention_bwd: dropout probability must be in [0, 1)");
  return attention_bwd_launch(qkv, dctx, mask, kv_info, lse2, delta, dqkv, B, T, H, Dh, make_drop(drop_p, drop_seed, drop_site),
                              true, s);
}


static int attention_bwd_launch(const void* qkv, const void* dctx, const uint8_t* mask, const int32_t* kv_info,
                                const float* lse2, const float* delta, void* dqkv, int32_t B, int32_t T, int32_t H,
                                int32_t Dh, const DropCfg drop, bool causal, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(qkv && dctx && mask && kv_info && lse2 && delta && dqkv, "attention_bwd: null pointer");
  TTSB_CHECK_ARG(B > 0 && T > 0 && H > 0 && Dh == ab::kDh, "attention_bwd: bad dims (Dh must be 96)");
  TTSB_CHECK_ARG(((reinterpret_cast<uintptr_t>(qkv) | reinterpret_cast<uintptr_t>(dctx) |
                   reinterpret_cast<uintptr_t>(dqkv)) & 15) == 0, "attention_bwd: 16-byte alignment");
  CUtensorMap tmQKV, tmDO;
  const uint32_t box[3] = {32, 128, 1};
  {
    const uint64_t ld = 3ull * H * Dh;
    const uint64_t dims[3] = {ld, static_cast<uint64_t>(T), static_cast<uint64_t>(B)};
    const uint64_t strides[2] = {ld * 2, ld * 2 * static_cast<uint64_t>(T)};
    int rc = make_tmap_bf16(&tmQKV, qkv, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_64B);
    if (rc) return rc;
  }
  {
    const uint64_t ld = 1ull * H * Dh;
    const uint64_t dims[3] = {ld, static_cast<uint64_t>(T), static_cast<uint64_t>(B)};
    const uint64_t strides[2] = {ld * 2, ld * 2 * static_cast<uint64_t>(T)};
    int rc = make_tmap_bf16(&tmDO, dctx, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_64B);
    if (rc) return rc;
  }
  const float scale = 1.f / sqrtf(static_cast<float>(Dh));
  const float scale_log2 = ab::kLog2e * scale;
  dim3 grid((T + 127) / 128, H, B);
  cudaStream_t st = static_cast<cudaStream_t>(s);
  if (causal) {
    auto kdkv = ab::attn_bwd_dkv_kernel<true>;
    auto kdq = ab::attn_bwd_dq_kernel<true>;
    TTSB_SMEM_OPT_IN(kdkv, ab::kSmemDkv);
    TTSB_SMEM_OPT_IN(kdq, ab::kSmemDq);
    kdkv<<<grid, ab::kThreads, ab::kSmemDkv, st>>>(tmQKV, tmDO, mask, kv_info, lse2, delta, static_cast<__nv_bfloat16*>(dqkv), T,
                                                   H, scale_log2, scale, drop);
    TTSB_CHECK_CUDA(cudaGetLastError());
    kdq<<<grid, ab::kThreads, ab::kSmemDq, st>>>(tmQKV, tmDO, mask, kv_info, lse2, delta, static_cast<__nv_bfloat16*>(dqkv), T, H,
                                                 scale_log2, scale, drop);
  } else {
    auto kdkv = ab::attn_bwd_dkv_kernel<false>;
    auto kdq = ab::attn_bwd_dq_kernel<false>;
    TTSB_SMEM_OPT_IN(kdkv, ab::kSmemDkv);
    TTSB_SMEM_OPT_IN(kdq, ab::kSmemDq);
    kdkv<<<grid, ab::kThreads, ab::kSmemDkv, st>>>(tmQKV, tmDO, mask, kv_info, lse2, delta, static_cast<__nv_bfloat16*>(dqkv), T,
                                                   H, scale_log2, scale, drop);
    TTSB_CHECK_CUDA(cudaGetLastError());
    kdq<<<grid, ab::kThreads, ab::kSmemDq, st>>>(tmQKV, tmDO, mask, kv_info, lse2, delta, static_cast<__nv_bfloat16*>(dqkv), T, H,
                                                 scale_log2, scale, drop);
  }
  TTSB_CHECK_CUDA(cudaGetLastError());
  return TTSB_OK;
}
