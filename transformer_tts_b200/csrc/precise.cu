// fp32-accurate encoder -> duration-predictor chain on the tensor cores (SURVEY §7 "Parity conditioning", row N1).
//
// The reference rounds an fp32 log_d (Models/varianceadaptor.py:69,75); a bf16 encoder chain flips ~5 % of the
// durations.  This file holds the kernels of a chain that reproduces fp32 arithmetic to ~fp32 noise
// while still running its contractions on tcgen05 / mma tensor cores:
//
//   * every operand is split into two fp16 values,  v = hi + lo  (hi = fp16(v), lo = fp16(v - hi): 22 mantissa
//     bits; weights are pre-scaled by a power of two so their lo parts stay normal), and a product is evaluated as
//     hi*hi + lo*hi + hi*lo (the lo*lo term is below 2^-22);
//   * the tensor core's own accumulator TRUNCATES (measured: tools/exp/tc_accum_probe.py — one round-toward-zero
//     per K=16 MMA step, a drift of 2^-24 per step, 2.8e-5 at K = 7680), so an accumulation chain never runs
//     longer than one 64-wide k-block (8 tiny + 4 large MMA steps): each k-block's partial sum is read back from
//     TMEM and added in fp32 round-to-nearest REGISTERS by the epilogue warps, overlapped with the next k-block's
//     MMAs through four TMEM accumulator buffers.
//
// Kernels: gemm_x3 (implicit-GEMM conv1d / linear, TMA -> tcgen05 -> TMEM -> register accumulation),
// attention_x3 (flash-style masked attention on mma.sync m16n8k16 with the same splitting and short chains; the
// encoder's sequences are a few hundred phonemes, <2 % of the chain's FLOPs), xrow (fp32 residual / LayerNorm /
// FFN tail / predictor head, one warp per row) and split_f16.
//
// Reference ops replaced: Models/encoder.py:83-112, Models/layers.py:29-41, Models/modules.py:7-88,
// Models/varianceadaptor.py:216-229 (the duration predictor), evaluated to fp32 accuracy.
#include <cuda_fp16.h>

#include "common.cuh"
#include "../../include/ttsb200.h"

namespace ttsb {

// ------------------------------------------------------------------------------------------------
// fp16 hi/lo split
// ------------------------------------------------------------------------------------------------
// {hi(a), hi(b)} and {lo(a), lo(b)} as packed half2 words (a in the low half).  satfinite: |v| > 65504 saturates
// instead of producing inf/NaN (the chain is specified for |activations| < 6e4).
__device__ __forceinline__ void split2(float a, float b, uint32_t& hi, uint32_t& lo) {
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(hi) : "f"(b), "f"(a));
  const float2 h = __half22float2(*reinterpret_cast<const __half2*>(&hi));
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(lo) : "f"(b - h.y), "f"(a - h.x));
}

// ------------------------------------------------------------------------------------------------
// gemm_x3: CTA pairs (cta_group::2).  One unit = 256 rows (two 128-row tiles, one per CTA of the cluster) x BN
// output channels; each CTA stages its own rows of A (hi, lo) and HALF of the BN weight rows (hi, lo).
//
// The kernel is bound by L2 -> SM operand traffic, not by the tensor pipe: with every operand present twice (hi and
// lo planes) a k-block moves 32 KB + BN*128 B into each SM for 12 MMAs, and with the MMAs AND the TMEM read-back
// deleted the k=5 convolutions ran no faster (ttsb200_debug_x3, tools/exp/precise_times.py: 132 of 140 us).  So a
// conv's taps share ONE activation box: A is loaded once per 64-channel block as 128 + 8 rows (time steps
// t0 - pad_left ...), and tap j reads rows [j, j + 128) of it through a descriptor whose start address is offset by
// j rows.  (The SWIZZLE_128B XOR is a function of the shared-memory ADDRESS bits, so a start address inside a
// 1024-byte atom reads the rows TMA wrote there; the descriptor's "matrix base offset" field stays 0 — setting it
// to j & 7 gave wrong products.)  Activations then cost 1/taps of the traffic; only the weights stream per tap,
// through their own ring.  Cross-CTA "buffer drained" arrivals are RELAXED: see mbar_arrive_leader_relaxed.
// ------------------------------------------------------------------------------------------------
constexpr int kXM = 128;            // rows per CTA (time steps of one utterance)
constexpr int kXK = 64;             // k-block: 64 fp16 = one 128-byte swizzle row
constexpr int kXBox = 16384;        // 128 rows x 128 bytes
constexpr int kXThreads = 64 + 256; // warp 0 TMA producer, warp 1 MMA issuer, warps 2..9 accumulate / epilogue
constexpr int kXStoreBytes = 2 * kXBox;  // one staging box per epilogue warpgroup
constexpr int kXAStages = 2;

template <int BN, int HALO>
struct X3Cfg {
  static constexpr int kABox = (kXM + HALO) * 128;           // one plane of this CTA's activation rows (+ conv halo)
  static constexpr int kAStage = 2 * kABox;                  // hi | lo
  static constexpr int kWBox = (BN / 2) * 128;               // this CTA's half of the weight rows, one plane
  static constexpr int kWStage = 2 * kWBox;                  // hi | lo
  static constexpr int kWFit = (232448 - 1024 - kXStoreBytes - kXAStages * kAStage) / kWStage;
  static constexpr int kWStages = kWFit > 6 ? 6 : kWFit;
  static constexpr int kNumAcc = 512 / BN;                   // TMEM accumulator buffers (BN columns each)
  static constexpr int kHalf = BN / 2;                       // columns per epilogue warpgroup
  static constexpr int kSmem = kXAStages * kAStage + kWStages * kWStage + kXStoreBytes + 1024;
  static_assert(BN == 128 || BN == 192, "tile widths built");
  static_assert(kWBox % 1024 == 0 && kABox % 1024 == 0, "swizzle atoms");
  static_assert(kWStages >= 3, "weight ring too shallow");
  static_assert(kSmem <= 232448, "shared memory budget");
};

struct X3Dev {
  int B, T, N, C, taps, pad_left, relu, out_split;
  int m_tiles, n_tiles;
  float inv_scale;
  const float* bias;
  int debug;  // timing experiments only (ttsb200_debug_x3): 1 = no TMEM read-back, 2 = no MMAs
};

__device__ long long g_x3_trace[148 * 16];

// "this warp has read its part of the accumulator buffer" -> the leader CTA's barrier.  RELAXED: the values are in
// registers (tcgen05.wait::ld), nothing this thread wrote has to become visible to the issuer, and the default
// release.cluster form of a remote arrive cost ~1500 clk per k-block (tools/exp/precise_trace.py).
__device__ __forceinline__ void mbar_arrive_leader_relaxed(uint32_t bar) {
  asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(bar & kPeerBitMask) : "memory");
}

// kind::f16 instruction descriptor with fp16 (not bf16) operands: D fp32 (1<<4), A/B format 0
__host__ __device__ constexpr uint32_t umma_idesc_f16(uint32_t M, uint32_t N) {
  return (1u << 4) | ((N >> 3) << 17) | ((M >> 4) << 24);
}
__device__ __forceinline__ void x_tma_store_3d(const CUtensorMap* m, uint32_t src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(m)), "r"(src), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
__device__ __forceinline__ void x_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void x_store_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void x_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void x_wg_bar(int wg) { asm volatile("bar.sync %0, 128;" ::"r"(2 + wg) : "memory"); }

template <int BN, int HALO>
__global__ void __launch_bounds__(kXThreads, 1)
gemm_x3_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
               const __grid_constant__ CUtensorMap tmOut, const X3Dev p) {
  using Cfg = X3Cfg<BN, HALO>;
  constexpr int SA = kXAStages, SW = Cfg::kWStages, NA = Cfg::kNumAcc, HALF = Cfg::kHalf;
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) {
    if (threadIdx.x == 0) printf("ttsb200: gemm_x3 dynamic shared memory base is not 1024-byte aligned\n");
    __trap();
  }
  uint8_t* sA = smem;
  uint8_t* sW = sA + SA * Cfg::kAStage;
  uint8_t* sStore = sW + SW * Cfg::kWStage;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sStore + kXStoreBytes);
  uint64_t* bar_afull = bars;                      // [SA] (the leader's full barriers are used: both CTAs' loads signal them)
  uint64_t* bar_aempty = bars + SA;                // [SA] multicast commit
  uint64_t* bar_wfull = bars + 2 * SA;             // [SW]
  uint64_t* bar_wempty = bars + 2 * SA + SW;       // [SW] multicast commit
  uint64_t* bar_accf = bars + 2 * SA + 2 * SW;     // [NA] one (k-block, tap) partial sum complete in TMEM (multicast commit)
  uint64_t* bar_acce = bar_accf + NA;              // [NA] ... and read back: one arrival per epilogue warp of both CTAs, on the leader's
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_acce + NA);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = cluster_ctarank();
  const bool is_leader = cta_rank == 0;
  const int kblocks = p.C / kXK;
  const int taps = p.taps;
  const int total_mt = p.B * p.m_tiles;
  const int num_units = ((total_mt + 1) / 2) * p.n_tiles;
  const int unit0 = static_cast<int>(blockIdx.x >> 1);
  const int ustride = static_cast<int>(gridDim.x >> 1);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmOut);
    for (int s = 0; s < SA; ++s) {
      mbar_init(smem_u32(&bar_afull[s]), 1);
      mbar_init(smem_u32(&bar_aempty[s]), 1);
    }
    for (int s = 0; s < SW; ++s) {
      mbar_init(smem_u32(&bar_wfull[s]), 1);
      mbar_init(smem_u32(&bar_wempty[s]), 1);
    }
    for (int i = 0; i < NA; ++i) {
      mbar_init(smem_u32(&bar_accf[i]), 1);
      mbar_init(smem_u32(&bar_acce[i]), 16);
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc_2sm(smem_u32(tmem_slot), 512);
    tmem_relinquish_2sm();
  }
  tc_fence_before();
  cluster_sync_all();  // the peer's barriers are initialised before anything arrives on them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // unit -> this CTA's (utterance, first row, n-tile); n fastest so concurrently running pairs share activations in L2
  auto decode = [&](int unit, int& b, int& t0, int& nt) {
    const int mtp = unit / p.n_tiles;
    nt = unit - mtp * p.n_tiles;
    const int mt = 2 * mtp + static_cast<int>(cta_rank);
    if (mt >= total_mt) {  // odd tile count: the last pair's second tile is empty (loads zero-fill, stores clip)
      b = 0;
      t0 = (p.m_tiles + 1) * kXM;
      return;
    }
    b = mt / p.m_tiles;
    t0 = (mt - b * p.m_tiles) * kXM;
  };

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (lane == 0) {
      uint32_t ga = 0, gw = 0;
      for (int unit = unit0; unit < num_units; unit += ustride) {
        int b, t0, nt;
        decode(unit, b, t0, nt);
        for (int kb = 0; kb < kblocks; ++kb, ++ga) {
          const int sa = ga % SA;
          mbar_wait(smem_u32(&bar_aempty[sa]), ((ga / SA) & 1) ^ 1);
          const uint32_t afull = smem_u32(&bar_afull[sa]);
          if (is_leader) mbar_arrive_expect_tx(afull, 2 * Cfg::kAStage);
          const uint32_t ast = smem_u32(sA + sa * Cfg::kAStage);
          // rows t0 - pad_left .. + 128 + HALO: every tap's time-shifted window; out-of-range rows are zero-filled
          tma_load_3d_2sm(ast, &tmA, afull, kb * kXK, t0 - p.pad_left, b);
          tma_load_3d_2sm(ast + Cfg::kABox, &tmA, afull, p.C + kb * kXK, t0 - p.pad_left, b);
          for (int tap = 0; tap < taps; ++tap, ++gw) {
            const int sw = gw % SW;
            mbar_wait(smem_u32(&bar_wempty[sw]), ((gw / SW) & 1) ^ 1);
            const uint32_t wfull = smem_u32(&bar_wfull[sw]);
            if (is_leader) mbar_arrive_expect_tx(wfull, 2 * Cfg::kWStage);
            const uint32_t wst = smem_u32(sW + sw * Cfg::kWStage);
            const int wrow = tap * p.N + nt * BN + static_cast<int>(cta_rank) * (BN / 2);
            tma_load_2d_2sm(wst, &tmW, wfull, kb * kXK, wrow);
            tma_load_2d_2sm(wst + Cfg::kWBox, &tmW, wfull, p.C + kb * kXK, wrow);
          }
        }
      }
    }
  } else if (warp == 1 && is_leader) {
    // ===================== MMA issuer (leader CTA; converged warp, one elected lane issues) =====================
    constexpr uint32_t idesc = umma_idesc_f16(2 * kXM, BN);
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t leader = elect_one();
    uint32_t ga = 0, gw = 0;
    const bool tr = (p.debug & 8) != 0;
    long long t_a = 0, t_e = 0, t_w = 0, t_i = 0, t0c = 0, t_start = tr ? clock64() : 0;
    for (int unit = unit0; unit < num_units; unit += ustride) {
      for (int kb = 0; kb < kblocks; ++kb, ++ga) {
        const int sa = ga % SA;
        if (tr) t0c = clock64();
        mbar_wait(smem_u32(&bar_afull[sa]), (ga / SA) & 1);
        if (tr) t_a += clock64() - t0c;
        const uint32_t ast = smem_u32(sA + sa * Cfg::kAStage);
        for (int tap = 0; tap < taps; ++tap, ++gw) {
          const int sw = gw % SW;
          const uint32_t ab = gw % NA;
          if (tr) t0c = clock64();
          mbar_wait(smem_u32(&bar_acce[ab]), ((gw / NA) & 1) ^ 1);  // buffer read back by the epilogue warps of both CTAs
          if (tr) { const long long t1 = clock64(); t_e += t1 - t0c; t0c = t1; }
          mbar_wait(smem_u32(&bar_wfull[sw]), (gw / SW) & 1);
          if (tr) { const long long t1 = clock64(); t_w += t1 - t0c; t0c = t1; }
          tc_fence_after();
          const uint32_t wst = smem_u32(sW + sw * Cfg::kWStage);
          const uint32_t d_tmem = tb + ab * BN;
          const uint32_t arow = ast + tap * 128;  // tap j = rows [j, j + 128) of the staged window
          if (leader && (p.debug & 2)) {
            umma_commit_2sm(smem_u32(&bar_wempty[sw]));
            if (tap == taps - 1) umma_commit_2sm(smem_u32(&bar_aempty[sa]));
            umma_commit_2sm(smem_u32(&bar_accf[ab]));
          } else if (leader) {
            // the two small cross terms first, the large hi*hi term last: the accumulator truncates once per MMA
            // step relative to its current magnitude, so only the last four steps cost 2^-24 each
#pragma unroll
            for (int k = 0; k < kXK / 16; ++k) {
              const uint64_t a_hi = umma_smem_desc(arow + k * 32, 16, 1024, kLayoutSW128);
              const uint64_t a_lo = umma_smem_desc(arow + Cfg::kABox + k * 32, 16, 1024, kLayoutSW128);
              const uint64_t w_hi = umma_smem_desc(wst + k * 32, 16, 1024, kLayoutSW128);
              const uint64_t w_lo = umma_smem_desc(wst + Cfg::kWBox + k * 32, 16, 1024, kLayoutSW128);
              umma_bf16_2sm(d_tmem, a_lo, w_hi, idesc, k != 0 ? 1u : 0u);
              umma_bf16_2sm(d_tmem, a_hi, w_lo, idesc, 1u);
            }
#pragma unroll
            for (int k = 0; k < kXK / 16; ++k) {
              const uint64_t a_hi = umma_smem_desc(arow + k * 32, 16, 1024, kLayoutSW128);
              const uint64_t w_hi = umma_smem_desc(wst + k * 32, 16, 1024, kLayoutSW128);
              umma_bf16_2sm(d_tmem, a_hi, w_hi, idesc, 1u);
            }
            umma_commit_2sm(smem_u32(&bar_wempty[sw]));
            if (tap == taps - 1) umma_commit_2sm(smem_u32(&bar_aempty[sa]));
            umma_commit_2sm(smem_u32(&bar_accf[ab]));
          }
          __syncwarp();
          if (tr) t_i += clock64() - t0c;
        }
      }
    }
    if (tr && lane == 0 && blockIdx.x < 296) {
      long long* o = g_x3_trace + (blockIdx.x >> 1) * 16;
      o[0] = clock64() - t_start; o[1] = t_a; o[2] = t_e; o[3] = t_w; o[4] = t_i; o[5] = gw;
    }
  } else if (warp >= 2) {
    // ===================== accumulate + epilogue warps: thread == output row == TMEM lane =====================
    const int e = warp - 2;
    const int wg = e >> 2;          // column half of the tile
    const int q = warp & 3;         // TMEM lane quarter this warp may touch
    const int row = q * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    uint8_t* my_box = sStore + wg * kXBox;
    const int num_iters = kblocks * taps;
    uint32_t git = 0;
    for (int unit = unit0; unit < num_units; unit += ustride) {
      int b, t0, nt;
      decode(unit, b, t0, nt);
      float acc[HALF];
#pragma unroll
      for (int i = 0; i < HALF; ++i) acc[i] = 0.f;
      for (int it = 0; it < num_iters; ++it, ++git) {
        const uint32_t ab = git % NA;
        const long long te0 = (p.debug & 8) ? clock64() : 0;
        mbar_wait(smem_u32(&bar_accf[ab]), (git / NA) & 1);
        if ((p.debug & 8) && e == 0 && lane == 0 && is_leader) g_x3_trace[(blockIdx.x >> 1) * 16 + 8] += clock64() - te0;
        tc_fence_after();
        const uint32_t ta = tmem_base + ab * BN + wg * HALF + lane_off;
        constexpr int CW = 32;
#pragma unroll
        for (int c = 0; c < HALF; c += CW) {
          float v[CW];
          if (!(p.debug & 1)) {
            if constexpr (CW == 32) tmem_ld32(ta + c, v);
            else tmem_ld16(ta + c, v);
            tmem_ld_wait();
          } else {
#pragma unroll
            for (int i = 0; i < CW; ++i) v[i] = 1.f;
          }
          if (c + CW >= HALF) {  // the last chunk is in registers: the buffer is free again
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_leader_relaxed(smem_u32(&bar_acce[ab]));
          }
#pragma unroll
          for (int i = 0; i < CW; i += 2) add2(acc[c + i], acc[c + i + 1], v[i], v[i + 1]);  // fp32 round-to-nearest
        }
      }
      // ---- tile epilogue: v = acc * 2^-e + bias (+ ReLU) -> fp32 boxes or fp16 hi | lo boxes -> TMA store ----
      const int n0 = nt * BN + wg * HALF;
#pragma unroll
      for (int i = 0; i < HALF; ++i) {
        acc[i] = fmaf(acc[i], p.inv_scale, __ldg(p.bias + n0 + i));
        if (p.relu) acc[i] = fmaxf(acc[i], 0.f);
      }
      if (!p.out_split) {
#pragma unroll
        for (int h = 0; h < HALF / 32; ++h) {
          if (row == 0) x_store_wait_read0();  // the previous store from this box has read it
          x_wg_bar(wg);
          uint8_t* rp = my_box + row * 128;
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(rp + ((j ^ (row & 7)) << 4)) =
                make_float4(acc[h * 32 + 4 * j], acc[h * 32 + 4 * j + 1], acc[h * 32 + 4 * j + 2], acc[h * 32 + 4 * j + 3]);
          fence_async_smem();
          x_wg_bar(wg);
          if (row == 0) {
            x_tma_store_3d(&tmOut, smem_u32(my_box), n0 + h * 32, t0, b);
            x_store_commit();
          }
        }
      } else {
#pragma unroll
        for (int h = 0; h < HALF / 32; ++h) {
#pragma unroll
          for (int pl = 0; pl < 2; ++pl) {  // hi plane, then lo plane (the split is recomputed: registers are scarce)
            if (row == 0) x_store_wait_read0();
            x_wg_bar(wg);
            uint8_t* rp = my_box + row * 64;  // 128 rows x 32 fp16, SWIZZLE_64B: chunk j of row r at (j ^ ((r >> 1) & 3))
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              uint32_t w[4], o[4];
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const int c = h * 32 + 8 * j + 2 * i;
                if (pl == 0) split2(acc[c], acc[c + 1], w[i], o[i]);
                else split2(acc[c], acc[c + 1], o[i], w[i]);
              }
              *reinterpret_cast<uint4*>(rp + ((j ^ ((row >> 1) & 3)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
            }
            fence_async_smem();
            x_wg_bar(wg);
            if (row == 0) {
              x_tma_store_3d(&tmOut, smem_u32(my_box), pl * p.N + n0 + h * 32, t0, b);
              x_store_commit();
            }
          }
        }
      }
    }
    if (row == 0) x_store_wait_all();
  }

  tc_fence_before();
  cluster_sync_all();  // every multicast commit has been consumed by a wait in both CTAs before either one leaves
  if (warp == 1) tmem_dealloc_2sm(tmem_base, 512);
}

// ------------------------------------------------------------------------------------------------
// attention_x3: softmax(mask(q k^T / sqrt(dk), -1e4)) v on mma.sync with split operands and short chains.
//   qkv fp16 hi | lo (B,T,2*3*H*96) (the q|k|v GEMM's split output) -> ctx fp16 hi | lo (B,T,2*H*96)
// One CTA = 64 queries of one (utterance, head), 4 warps x 16 rows; the hi / lo planes of a 64-key tile of K and V
// are copied to shared memory with cp.async ([key][dim], 208-byte rows: conflict-free for the fragment loads and
// ldmatrix).
// ------------------------------------------------------------------------------------------------
constexpr int kADh = 96;
constexpr int kABQ = 64;
constexpr int kABK = 64;
constexpr int kAPitch = 104;  // halfs per smem row (208 bytes)
constexpr int kAPlane = kABK * kAPitch * 2;                   // bytes of one plane
constexpr int kASmem = 4 * kAPlane + 64;                      // K hi, K lo, V hi, V lo, key flags
constexpr float kAPScale = 1024.f;    // p <= 1 is scaled by 2^10 before the split

__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldmatrix_x2_trans(uint32_t& r0, uint32_t& r1, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0, %1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(addr));
}

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
  const int n = valid ? 16 : 0;  // src-size 0: the 16 bytes are zero-filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(n) : "memory");
}

__global__ void __launch_bounds__(128)
attention_x3_kernel(const __half* __restrict__ qkv, const uint8_t* __restrict__ mask, const int32_t* __restrict__ kv_info,
                    __half* __restrict__ ctx, int T, int H) {
  extern __shared__ __align__(16) uint8_t asmem[];
  __half* sKh = reinterpret_cast<__half*>(asmem);
  __half* sKl = reinterpret_cast<__half*>(asmem + kAPlane);
  __half* sVh = reinterpret_cast<__half*>(asmem + 2 * kAPlane);
  __half* sVl = reinterpret_cast<__half*>(asmem + 3 * kAPlane);
  uint8_t* sflag = asmem + 4 * kAPlane;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t4 = lane & 3;
  const int q0 = blockIdx.x * kABQ, h = blockIdx.y, b = blockIdx.z;
  const int D = H * kADh;
  const int64_t ld = 6LL * D;                 // [q k v hi | q k v lo]
  const int64_t lo_off = 3LL * D;
  const __half* base = qkv + static_cast<int64_t>(b) * T * ld;
  const int klen = kv_info[2 * b];
  const int kv_eff = klen > 0 ? klen : T;  // no valid key: uniform over all T keys like the reference
  const int nkv = (kv_eff + kABK - 1) / kABK;

  // ---- Q fragments (rows g and g + 8 of this warp's 16): hi and lo planes as they are ----
  uint32_t qh[6][4], ql[6][4];
  {
    const int r0 = q0 + warp * 16 + g, r1 = r0 + 8;
    const __half* p0 = base + static_cast<int64_t>(r0 < T ? r0 : 0) * ld + h * kADh + 2 * t4;
    const __half* p1 = base + static_cast<int64_t>(r1 < T ? r1 : 0) * ld + h * kADh + 2 * t4;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      qh[j][0] = *reinterpret_cast<const uint32_t*>(p0 + 16 * j);
      qh[j][1] = *reinterpret_cast<const uint32_t*>(p1 + 16 * j);
      qh[j][2] = *reinterpret_cast<const uint32_t*>(p0 + 16 * j + 8);
      qh[j][3] = *reinterpret_cast<const uint32_t*>(p1 + 16 * j + 8);
      ql[j][0] = *reinterpret_cast<const uint32_t*>(p0 + lo_off + 16 * j);
      ql[j][1] = *reinterpret_cast<const uint32_t*>(p1 + lo_off + 16 * j);
      ql[j][2] = *reinterpret_cast<const uint32_t*>(p0 + lo_off + 16 * j + 8);
      ql[j][3] = *reinterpret_cast<const uint32_t*>(p1 + lo_off + 16 * j + 8);
    }
  }
  const float s_scale = 1.f / sqrtf(static_cast<float>(kADh));
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
  float o[12][4];
#pragma unroll
  for (int i = 0; i < 12; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;

  for (int jt = 0; jt < nkv; ++jt) {
    __syncthreads();  // the previous tile's fragments have been read
    // ---- K, V tile: four planes of 64 keys x 96 dims, 12 x 16 bytes per row ----
    for (int i = threadIdx.x; i < kABK * 12; i += 128) {
      const int key = i / 12, c8 = i - key * 12;
      const int kg = jt * kABK + key;
      const bool ok = kg < T;
      const __half* rp = base + static_cast<int64_t>(ok ? kg : 0) * ld + h * kADh + 8 * c8;
      const uint32_t so = static_cast<uint32_t>((key * kAPitch + 8 * c8) * 2);
      cp_async16(smem_u32(sKh) + so, rp + D, ok);
      cp_async16(smem_u32(sKl) + so, rp + lo_off + D, ok);
      cp_async16(smem_u32(sVh) + so, rp + 2 * D, ok);
      cp_async16(smem_u32(sVl) + so, rp + lo_off + 2 * D, ok);
    }
    if (threadIdx.x < kABK) {
      const int kg = jt * kABK + threadIdx.x;
      sflag[threadIdx.x] = kg < T ? (mask[static_cast<int64_t>(b) * T + kg] ? 1 : 0) : 2;
    }
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
    __syncthreads();

    // ---- S = Q K^T (16 x 64 per warp).  mma.sync has a long latency on this part and a warp issues in order: the eight key
    // blocks' accumulator chains are interleaved (k-step outer, key block inner) instead of run one after the other; every
    // chain still sees its k-steps in the same order, so the result is bit-identical ----
    float s[8][4];
    {
      float c_lo[8][4], c_hi[8][4];
#pragma unroll
      for (int nb = 0; nb < 8; ++nb)
#pragma unroll
        for (int i = 0; i < 4; ++i) c_lo[nb][i] = c_hi[nb][i] = 0.f;
#pragma unroll
      for (int j = 0; j < 6; ++j) {
#pragma unroll
        for (int nb = 0; nb < 8; ++nb) {
          const __half* kh = sKh + (nb * 8 + g) * kAPitch + 2 * t4;
          const __half* kl = sKl + (nb * 8 + g) * kAPitch + 2 * t4;
          const uint32_t bh0 = *reinterpret_cast<const uint32_t*>(kh + 16 * j);
          const uint32_t bh1 = *reinterpret_cast<const uint32_t*>(kh + 16 * j + 8);
          const uint32_t bl0 = *reinterpret_cast<const uint32_t*>(kl + 16 * j);
          const uint32_t bl1 = *reinterpret_cast<const uint32_t*>(kl + 16 * j + 8);
          mma16816(c_lo[nb], ql[j], bh0, bh1);
          mma16816(c_lo[nb], qh[j], bl0, bl1);
          mma16816(c_hi[nb], qh[j], bh0, bh1);
        }
      }
#pragma unroll
      for (int nb = 0; nb < 8; ++nb) {
#pragma unroll
        for (int i = 0; i < 4; ++i) s[nb][i] = (c_hi[nb][i] + c_lo[nb][i]) * s_scale;
        // mask like the reference: -1e4 on padded keys (modules.py:14); keys beyond T do not exist
        const uint8_t f0 = sflag[nb * 8 + 2 * t4], f1 = sflag[nb * 8 + 2 * t4 + 1];
        if (f0 != 1) { const float mv = f0 == 0 ? -1e4f : -INFINITY; s[nb][0] = mv; s[nb][2] = mv; }
        if (f1 != 1) { const float mv = f1 == 0 ? -1e4f : -INFINITY; s[nb][1] = mv; s[nb][3] = mv; }
      }
    }
    // ---- online softmax (rows g: elements 0,1; rows g + 8: elements 2,3) ----
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int nb = 0; nb < 8; ++nb) {
      mx0 = fmaxf(mx0, fmaxf(s[nb][0], s[nb][1]));
      mx1 = fmaxf(mx1, fmaxf(s[nb][2], s[nb][3]));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
    const float al0 = expf(m0 - mn0), al1 = expf(m1 - mn1);  // m = -inf on the first tile -> 0
    m0 = mn0; m1 = mn1;
    l0 *= al0; l1 *= al1;
#pragma unroll
    for (int i = 0; i < 12; ++i) { o[i][0] *= al0; o[i][1] *= al0; o[i][2] *= al1; o[i][3] *= al1; }
    uint32_t ph[4][4], pl[4][4];
#pragma unroll
    for (int nb = 0; nb < 8; ++nb) {
      const float p0 = expf(s[nb][0] - m0), p1 = expf(s[nb][1] - m0);
      const float p2 = expf(s[nb][2] - m1), p3 = expf(s[nb][3] - m1);
      l0 += p0 + p1;
      l1 += p2 + p3;
      // C fragment of key blocks 2jj, 2jj+1 == A fragment of k-step jj
      const int jj = nb >> 1, o2 = (nb & 1) * 2;
      split2(p0 * kAPScale, p1 * kAPScale, ph[jj][o2], pl[jj][o2]);
      split2(p2 * kAPScale, p3 * kAPScale, ph[jj][o2 + 1], pl[jj][o2 + 1]);
    }
    // ---- O += P V (16 x 96 per warp); V fragments through ldmatrix.trans from the [key][dim] planes.  Chains interleaved
    // as above: six dim blocks at a time, k-step outer ----
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      float c_lo[6][4], c_hi[6][4];
#pragma unroll
      for (int d = 0; d < 6; ++d)
#pragma unroll
        for (int i = 0; i < 4; ++i) c_lo[d][i] = c_hi[d][i] = 0.f;
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        // lanes 0..15 supply the row addresses: keys 16jj + (lane & 15), dims [8db, 8db + 8)
        const int krow = 16 * jj + (lane & 15);
#pragma unroll
        for (int d = 0; d < 6; ++d) {
          const int db = half * 6 + d;
          uint32_t vh0, vh1, vl0, vl1;
          ldmatrix_x2_trans(vh0, vh1, smem_u32(sVh + krow * kAPitch + 8 * db));
          ldmatrix_x2_trans(vl0, vl1, smem_u32(sVl + krow * kAPitch + 8 * db));
          mma16816(c_lo[d], pl[jj], vh0, vh1);
          mma16816(c_lo[d], ph[jj], vl0, vl1);
          mma16816(c_hi[d], ph[jj], vh0, vh1);
        }
      }
#pragma unroll
      for (int d = 0; d < 6; ++d)
#pragma unroll
        for (int i = 0; i < 4; ++i) o[half * 6 + d][i] += c_hi[d][i] + c_lo[d][i];
    }
  }
  // ---- O / l -> fp16 hi | lo ----
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
  l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float inv0 = 1.f / (l0 * kAPScale), inv1 = 1.f / (l1 * kAPScale);
  const int r0 = q0 + warp * 16 + g, r1 = r0 + 8;
#pragma unroll
  for (int db = 0; db < 12; ++db) {
    const int col = h * kADh + 8 * db + 2 * t4;
    uint32_t hi, lo;
    if (r0 < T) {
      split2(o[db][0] * inv0, o[db][1] * inv0, hi, lo);
      __half* cp = ctx + (static_cast<int64_t>(b) * T + r0) * (2 * D);
      *reinterpret_cast<uint32_t*>(cp + col) = hi;
      *reinterpret_cast<uint32_t*>(cp + D + col) = lo;
    }
    if (r1 < T) {
      split2(o[db][2] * inv1, o[db][3] * inv1, hi, lo);
      __half* cp = ctx + (static_cast<int64_t>(b) * T + r1) * (2 * D);
      *reinterpret_cast<uint32_t*>(cp + col) = hi;
      *reinterpret_cast<uint32_t*>(cp + D + col) = lo;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// xrow: fp32 row operations of the chain, one warp per row, D = 128 * NV (256 or 384)
//   s  = a (+ b0 | + b0_bf16 | + hi/lo pair b0_split);  s_out = s
//   t  = LN(s; g1, b1)
//   training with dropout: b0 := drop_b(b0) before the add (Models/layers.py:33), s := drop_s(s) before s_out and the
//   LayerNorm (Models/modules.py:86), t := drop_t(t) after it (layers.py:40, varianceadaptor.py:219,222)
//   mode ADD_LN:   out = t
//   mode FFN_TAIL: z = c + t ; z_out = z ; out = LN(z; g2, b2)
//   mode LN_DOT:   out_row = mask ? dot(t, w) + wb : 0
//   out -> fp16 hi | lo (rows, 2D) and/or bf16 (rows, D) and/or fp32
// ------------------------------------------------------------------------------------------------
struct XRowArgs {
  const float* a;
  const float* b0;
  const __nv_bfloat16* b0_bf16;
  const __half* b0_split;
  DropCfg drop_b, drop_s, drop_t;
  const float* c;
  const float *g1, *b1, *g2, *b2;
  const float* dot_w;
  const float* dot_b;
  const uint8_t* row_mask;
  float* s_out;
  float* z_out;
  __half* out_split;
  __nv_bfloat16* out_bf16;
  float* out_f32;
  float* out_row;
  float eps;
  int64_t rows;
  int mode;
};
enum { XROW_ADD_LN = 0, XROW_FFN_TAIL = 1, XROW_LN_DOT = 2 };

template <int NV>
__device__ __forceinline__ void xrow_ln(float4 (&v)[NV], const float* __restrict__ g, const float* __restrict__ b,
                                        float eps, int lane) {
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
  const float mean = warp_sum(s) * (1.f / (128.f * NV));
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const float a = v[i].x - mean, c = v[i].y - mean, d = v[i].z - mean, e = v[i].w - mean;
    q += (a * a + c * c) + (d * d + e * e);
  }
  // 1 / sqrt(var + eps) with an IEEE square root and division, as the reference's layer_norm does
  const float rstd = 1.f / sqrtf(warp_sum(q) * (1.f / (128.f * NV)) + eps);
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

template <int NV>
__global__ void __launch_bounds__(256) xrow_kernel(const XRowArgs p) {
  constexpr int D = 128 * NV;
  const int lane = threadIdx.x & 31;
  const int64_t row = static_cast<int64_t>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= p.rows) return;
  float4 v[NV];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    v[i] = *(reinterpret_cast<const float4*>(p.a + row * D) + i * 32 + lane);
    const uint64_t e0 = static_cast<uint64_t>(row) * D + (i * 32 + lane) * 4;
    if (p.b0 || p.b0_bf16) {
      float4 r;
      if (p.b0) {
        r = *(reinterpret_cast<const float4*>(p.b0 + row * D) + i * 32 + lane);
      } else {
        const uint2 u = *(reinterpret_cast<const uint2*>(p.b0_bf16 + row * D) + i * 32 + lane);
        const float2 lo = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.x));
        const float2 hi = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.y));
        r = make_float4(lo.x, lo.y, hi.x, hi.y);
      }
      if (p.drop_b.thresh) {
        r.x *= drop_scale(p.drop_b, e0); r.y *= drop_scale(p.drop_b, e0 + 1);
        r.z *= drop_scale(p.drop_b, e0 + 2); r.w *= drop_scale(p.drop_b, e0 + 3);
      }
      v[i].x += r.x; v[i].y += r.y; v[i].z += r.z; v[i].w += r.w;
    }
    if (p.b0_split) {
      const uint2 hh = *(reinterpret_cast<const uint2*>(p.b0_split + row * 2 * D) + i * 32 + lane);
      const uint2 ll = *(reinterpret_cast<const uint2*>(p.b0_split + row * 2 * D + D) + i * 32 + lane);
      const float2 h0 = __half22float2(*reinterpret_cast<const __half2*>(&hh.x));
      const float2 h1 = __half22float2(*reinterpret_cast<const __half2*>(&hh.y));
      const float2 e0 = __half22float2(*reinterpret_cast<const __half2*>(&ll.x));
      const float2 e1 = __half22float2(*reinterpret_cast<const __half2*>(&ll.y));
      v[i].x += h0.x + e0.x; v[i].y += h0.y + e0.y; v[i].z += h1.x + e1.x; v[i].w += h1.y + e1.y;
    }
    if (p.drop_s.thresh) {
      v[i].x *= drop_scale(p.drop_s, e0); v[i].y *= drop_scale(p.drop_s, e0 + 1);
      v[i].z *= drop_scale(p.drop_s, e0 + 2); v[i].w *= drop_scale(p.drop_s, e0 + 3);
    }
    if (p.s_out) reinterpret_cast<float4*>(p.s_out + row * D)[i * 32 + lane] = v[i];
  }
  xrow_ln<NV>(v, p.g1, p.b1, p.eps, lane);
  if (p.drop_t.thresh) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const uint64_t e0 = static_cast<uint64_t>(row) * D + (i * 32 + lane) * 4;
      v[i].x *= drop_scale(p.drop_t, e0); v[i].y *= drop_scale(p.drop_t, e0 + 1);
      v[i].z *= drop_scale(p.drop_t, e0 + 2); v[i].w *= drop_scale(p.drop_t, e0 + 3);
    }
  }
  if (p.mode == XROW_LN_DOT) {
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const float4 w = __ldg(reinterpret_cast<const float4*>(p.dot_w) + i * 32 + lane);
      dot = fmaf(v[i].x, w.x, dot); dot = fmaf(v[i].y, w.y, dot);
      dot = fmaf(v[i].z, w.z, dot); dot = fmaf(v[i].w, w.w, dot);
    }
    dot = warp_sum(dot);
    if (lane == 0) {
      float o = dot + __ldg(p.dot_b);
      if (p.row_mask && p.row_mask[row] == 0) o = 0.f;
      p.out_row[row] = o;
    }
    return;
  }
  if (p.mode == XROW_FFN_TAIL) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const float4 r = *(reinterpret_cast<const float4*>(p.c + row * D) + i * 32 + lane);
      v[i].x += r.x; v[i].y += r.y; v[i].z += r.z; v[i].w += r.w;
      if (p.z_out) reinterpret_cast<float4*>(p.z_out + row * D)[i * 32 + lane] = v[i];
    }
    xrow_ln<NV>(v, p.g2, p.b2, p.eps, lane);
  }
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    if (p.out_split) {
      uint2 hh, ll;
      split2(v[i].x, v[i].y, hh.x, ll.x);
      split2(v[i].z, v[i].w, hh.y, ll.y);
      reinterpret_cast<uint2*>(p.out_split + row * 2 * D)[i * 32 + lane] = hh;
      reinterpret_cast<uint2*>(p.out_split + row * 2 * D + D)[i * 32 + lane] = ll;
    }
    if (p.out_bf16) {
      uint2 u;
      u.x = pack_bf16(v[i].x, v[i].y);
      u.y = pack_bf16(v[i].z, v[i].w);
      reinterpret_cast<uint2*>(p.out_bf16 + row * D)[i * 32 + lane] = u;
    }
    if (p.out_f32) reinterpret_cast<float4*>(p.out_f32 + row * D)[i * 32 + lane] = v[i];
  }
}

// fp32 (rows, D) * scale -> fp16 hi | lo (rows, 2D); D % 4 == 0
__global__ void split_f16_kernel(const float* __restrict__ x, __half* __restrict__ out, int64_t rows, int D, float scale) {
  const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  const int d4 = D / 4;
  if (i >= rows * d4) return;
  const int64_t r = i / d4;
  const int c = static_cast<int>(i - r * d4);
  const float4 v = *(reinterpret_cast<const float4*>(x + r * D) + c);
  uint2 hh, ll;
  split2(v.x * scale, v.y * scale, hh.x, ll.x);
  split2(v.z * scale, v.w * scale, hh.y, ll.y);
  *(reinterpret_cast<uint2*>(out + r * 2 * D) + c) = hh;
  *(reinterpret_cast<uint2*>(out + r * 2 * D + D) + c) = ll;
}

}  // namespace ttsb

using namespace ttsb;

static int g_x3_debug = 0;
extern "C" void ttsb200_debug_x3(int flags) { g_x3_debug = flags; }
extern "C" int ttsb200_debug_x3_trace(long long* dst, int n, int clear) {
  if (clear) {
    static long long zeros[148 * 16] = {0};
    return cudaMemcpyToSymbol(g_x3_trace, zeros, sizeof(zeros)) == cudaSuccess ? 0 : -1;
  }
  return cudaMemcpyFromSymbol(dst, g_x3_trace, n * sizeof(long long)) == cudaSuccess ? 0 : -1;
}

static inline bool x_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <int BN, int HALO>
static int launch_gemm_x3(const void* x_split, const void* w_split, float* out_f32, void* out_split, X3Dev d,
                          cudaStream_t stream) {
  using Cfg = X3Cfg<BN, HALO>;
  CUtensorMap tmA, tmW, tmOut;
  const uint64_t uT = static_cast<uint64_t>(d.T), uB = static_cast<uint64_t>(d.B), uN = static_cast<uint64_t>(d.N);
  const uint64_t uC = static_cast<uint64_t>(d.C);
  {
    const uint64_t dims[3] = {2 * uC, uT, uB};
    const uint64_t strides[2] = {4 * uC, 4 * uC * uT};
    const uint32_t box[3] = {kXK, kXM + HALO, 1};
    int rc = make_tmap(&tmA, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, x_split, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  }
  {
    const uint64_t dims[2] = {2 * uC, static_cast<uint64_t>(d.taps) * uN};
    const uint64_t strides[1] = {4 * uC};
    const uint32_t box[2] = {kXK, BN / 2};
    int rc = make_tmap(&tmW, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, w_split, 2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  }
  if (out_f32) {
    const uint64_t dims[3] = {uN, uT, uB};
    const uint64_t strides[2] = {uN * 4, uN * 4 * uT};
    const uint32_t box[3] = {32, kXM, 1};
    int rc = make_tmap(&tmOut, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, out_f32, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  } else {
    const uint64_t dims[3] = {2 * uN, uT, uB};
    const uint64_t strides[2] = {uN * 4, uN * 4 * uT};
    const uint32_t box[3] = {32, kXM, 1};
    int rc = make_tmap(&tmOut, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, out_split, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_64B);
    if (rc) return rc;
  }
  auto kern = gemm_x3_kernel<BN, HALO>;
  TTSB_SMEM_OPT_IN(kern, Cfg::kSmem);
  d.n_tiles = d.N / BN;
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int units = ((d.B * d.m_tiles + 1) / 2) * d.n_tiles;
  const int pairs = units < sms / 2 ? units : sms / 2;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * pairs);
  cfg.blockDim = dim3(kXThreads);
  cfg.dynamicSmemBytes = Cfg::kSmem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  TTSB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, tmA, tmW, tmOut, d));
  return TTSB_OK;
}

extern "C" int ttsb200_gemm_x3(const void* x_split, const void* w_split, const float* bias, float inv_scale,
                               int32_t B, int32_t T, int32_t N, int32_t C, int32_t taps, int32_t pad_left,
                               int32_t relu, float* out_f32, void* out_split, ttsb200_stream_t stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  TTSB_CHECK_ARG(x_split && w_split && bias, "gemm_x3: null operand");
  TTSB_CHECK_ARG(B > 0 && T > 0 && taps > 0 && taps <= 8, "gemm_x3: bad dims (taps <= 8)");
  TTSB_CHECK_ARG(N % 128 == 0 && N > 0, "gemm_x3: N (%d) must be a multiple of 128", N);
  TTSB_CHECK_ARG(C % kXK == 0 && C > 0, "gemm_x3: C (%d) must be a multiple of %d", C, kXK);
  TTSB_CHECK_ARG((out_f32 != nullptr) != (out_split != nullptr), "gemm_x3: exactly one of out_f32 / out_split");
  TTSB_CHECK_ARG(x_aligned16(x_split) && x_aligned16(w_split) && x_aligned16(out_f32) && x_aligned16(out_split),
                 "gemm_x3: operands must be 16-byte aligned");
  X3Dev d;
  d.B = B; d.T = T; d.N = N; d.C = C; d.taps = taps; d.pad_left = pad_left; d.relu = relu; d.out_split = out_split ? 1 : 0;
  d.m_tiles = (T + kXM - 1) / kXM;
  d.n_tiles = 0;
  d.inv_scale = inv_scale;
  d.bias = bias;
  d.debug = g_x3_debug;
  // tile width: the candidate that minimises waves x MMA clocks per unit over the 74 CTA pairs
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const long long mtp = (static_cast<long long>(B) * d.m_tiles + 1) / 2;
  int best = 0;
  long long best_cost = 0;
  const int cands[2] = {192, 128};
  for (int i = 0; i < 2; ++i) {
    const int bn = cands[i];
    if (N % bn) continue;
    const long long units = mtp * (N / bn);
    const long long waves = (units + sms / 2 - 1) / (sms / 2);
    const long long cost = waves * (bn + 24);  // + a per-k-block overhead that does not scale with the width
    if (!best || cost < best_cost) { best = bn; best_cost = cost; }
  }
  if (taps > 1) {  // conv: one activation window of 128 + 8 rows serves every tap
    if (best == 192) return launch_gemm_x3<192, 8>(x_split, w_split, out_f32, out_split, d, stream);
    if (best == 128) return launch_gemm_x3<128, 8>(x_split, w_split, out_f32, out_split, d, stream);
  } else {
    if (best == 192) return launch_gemm_x3<192, 0>(x_split, w_split, out_f32, out_split, d, stream);
    if (best == 128) return launch_gemm_x3<128, 0>(x_split, w_split, out_f32, out_split, d, stream);
  }
  set_last_error("gemm_x3: no tile width for N=%d (split output: %d)", N, out_split ? 1 : 0);
  return TTSB_ERR_ARG;
}

extern "C" int ttsb200_attention_x3(const void* qkv_split, const uint8_t* mask, const int32_t* kv_info, void* ctx_split,
                                    int32_t B, int32_t T, int32_t H, int32_t Dh, ttsb200_stream_t stream) {
  TTSB_CHECK_ARG(qkv_split && mask && kv_info && ctx_split, "attention_x3: null pointer");
  TTSB_CHECK_ARG(B > 0 && T > 0 && H > 0, "attention_x3: bad dims");
  TTSB_CHECK_ARG(Dh == kADh, "attention_x3: head dim %d unsupported (this build: %d)", Dh, kADh);
  TTSB_CHECK_ARG(x_aligned16(qkv_split) && x_aligned16(ctx_split), "attention_x3: qkv/ctx must be 16-byte aligned");
  TTSB_SMEM_OPT_IN(attention_x3_kernel, kASmem);
  dim3 grid((T + kABQ - 1) / kABQ, H, B);
  attention_x3_kernel<<<grid, 128, kASmem, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const __half*>(qkv_split), mask, kv_info, static_cast<__half*>(ctx_split), T, H);
  TTSB_CHECK_CUDA(cudaGetLastError());
  return TTSB_OK;
}

extern "C" int ttsb200_xrow(int32_t mode, const float* a, const float* b0, const void* b0_bf16, const void* b0_split,
                            const float* c, const float* g1, const float* b1, const float* g2, const float* b2,
                            const float* dot_w, const float* dot_b, const uint8_t* row_mask, float eps, float* s_out,
                            float* z_out, void* out_split, void* out_bf16, float* out_f32, float* out_row, int64_t rows,
                            int32_t D, float drop_p, uint64_t drop_seed, int64_t site_b, int64_t site_s, int64_t site_t,
                            ttsb200_stream_t stream) {
  TTSB_CHECK_ARG(a && g1 && b1 && rows > 0, "xrow: bad args");
  TTSB_CHECK_ARG(mode >= XROW_ADD_LN && mode <= XROW_LN_DOT, "xrow: bad mode %d", mode);
  TTSB_CHECK_ARG((b0 != nullptr) + (b0_bf16 != nullptr) + (b0_split != nullptr) <= 1, "xrow: b0, b0_bf16 and b0_split are exclusive");
  TTSB_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f, "xrow: dropout probability must be in [0, 1)");
  if (mode == XROW_FFN_TAIL) TTSB_CHECK_ARG(c && g2 && b2, "xrow: FFN_TAIL needs c, g2, b2");
  if (mode == XROW_LN_DOT) TTSB_CHECK_ARG(dot_w && dot_b && out_row, "xrow: LN_DOT needs dot_w, dot_b, out_row");
  else TTSB_CHECK_ARG(out_split || out_bf16 || out_f32, "xrow: no output");
  XRowArgs p;
  p.a = a; p.b0 = b0; p.b0_bf16 = static_cast<const __nv_bfloat16*>(b0_bf16);
  p.b0_split = static_cast<const __half*>(b0_split); p.c = c;
  p.drop_b = make_drop(drop_p, drop_seed, site_b);
  p.drop_s = make_drop(drop_p, drop_seed, site_s);
  p.drop_t = make_drop(drop_p, drop_seed, site_t);
  p.g1 = g1; p.b1 = b1; p.g2 = g2; p.b2 = b2; p.dot_w = dot_w; p.dot_b = dot_b; p.row_mask = row_mask;
  p.s_out = s_out; p.z_out = z_out; p.out_split = static_cast<__half*>(out_split);
  p.out_bf16 = static_cast<__nv_bfloat16*>(out_bf16); p.out_f32 = out_f32; p.out_row = out_row;
  p.eps = eps; p.rows = rows; p.mode = mode;
  const unsigned grid = static_cast<unsigned>((rows + 7) / 8);
  if (D == 384) xrow_kernel<3><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  else if (D == 256) xrow_kernel<2><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  else TTSB_CHECK_ARG(false, "xrow: D=%d unsupported (256 or 384)", D);
  TTSB_CHECK_CUDA(cudaGetLastError());
  return TTSB_OK;
}

extern "C" int ttsb200_split_f16(const float* x, void* out_split, int64_t rows, int32_t D, float scale,
                                 ttsb200_stream_t stream) {
  TTSB_CHECK_ARG(x && out_split && rows > 0 && D > 0 && D % 4 == 0, "split_f16: bad args");
  const int64_t n = rows * (D / 4);
  split_f16_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      x, static_cast<__half*>(out_split), rows, D, scale);
  TTSB_CHECK_CUDA(cudaGetLastError());
  return TTSB_OK;
}
