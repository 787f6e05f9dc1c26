// conv_gemm_kernel — included by gemm.cu (see the header comment there for the design).
#pragma once

namespace ttsb {

#if defined(TTSB_GEMM_TRACE)
__device__ long long g_gemm_trace[148 * 32];
#define GT_NOW() clock64()
#else
#define GT_NOW() 0ll
#endif

template <int BLOCK_N, int EPI, bool A_STAT, bool PAIR>
__global__ void __launch_bounds__(kGemmThreads, 1)
conv_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                 const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmOut32,
                 const __grid_constant__ CUtensorMap tmRes32, const __grid_constant__ CUtensorMap tmResBf,
                 const GemmDev p) {
  using Cfg = GemmCfg<BLOCK_N, EPI, A_STAT, PAIR>;
  constexpr int S = Cfg::kStages;
  constexpr int RS = Cfg::kRingSlots > 0 ? Cfg::kRingSlots : 1;
  constexpr int HALF = BLOCK_N / 2;  // LN epilogues: columns per warpgroup
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem) & 1023u) != 0) {
    if (threadIdx.x == 0) printf("ttsb200: dynamic shared memory base is not 1024-byte aligned\n");
    __trap();
  }
  uint8_t* sARes = smem;                                  // A_STAT only: 6 x 16 KB
  uint8_t* sStages = smem + Cfg::kAResBytes;
  uint8_t* sStore = sStages + S * Cfg::kStageBytes;       // one bf16 staging box per warpgroup, 1024-aligned
  uint8_t* sRing = sStore + Cfg::kStoreBytes;             // kRing only: 4 x 16 KB
  float* sParam = reinterpret_cast<float*>(sRing + Cfg::kRingBytes);
  float2* sStat = reinterpret_cast<float2*>(sParam + Cfg::kParamFloats);
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(sStat) + Cfg::kStatBytes);
  uint64_t* bar_full = bars;             // [S]
  uint64_t* bar_empty = bars + S;        // [S]
  uint64_t* bar_accf = bars + 2 * S;     // [2]
  uint64_t* bar_acce = bars + 2 * S + 2; // [2]
  uint64_t* bar_rf = bars + 2 * S + 4;   // [8] ring slot filled
  uint64_t* bar_re = bars + 2 * S + 12;  // [8] ring slot drained
  uint64_t* bar_af = bars + 2 * S + 20;  // [6] resident A k-block filled
  uint64_t* bar_ae = bars + 2 * S + 26;  // [6] resident A k-block free
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * S + 32);
  static_assert(RS <= 8 && (2 * S + 32) * 8 + 16 <= 512, "barrier block");

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int kblocks = (p.C + kBlockK - 1) / kBlockK;
  const int num_iters = A_STAT ? kblocks : p.taps * kblocks;
  const int n_inner = A_STAT ? p.n_tiles : 1;
  // CTA pair: the two CTAs of a cluster take the m-tiles 2u and 2u+1 of unit u and share every W tile
  const uint32_t cta_rank = PAIR ? cluster_ctarank() : 0u;
  const bool is_leader = cta_rank == 0;
  const int total_mt = p.mt_count ? __ldg(p.mt_count) : p.B * p.m_tiles;
  const int mt_units = PAIR ? (total_mt + 1) / 2 : total_mt;
  const int num_units = A_STAT ? mt_units : mt_units * p.n_tiles;
  const int unit0 = PAIR ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
  const int ustride = PAIR ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmOut);
    for (int s = 0; s < S; ++s) {
      mbar_init(smem_u32(&bar_full[s]), 1);
      mbar_init(smem_u32(&bar_empty[s]), 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&bar_accf[i]), 1);
      // every epilogue thread; pair: one arrival per epilogue warp of either CTA, all on the leader's barrier
      mbar_init(smem_u32(&bar_acce[i]), PAIR ? 16 : kEpiThreads);
    }
    for (int i = 0; i < 6; ++i) {
      mbar_init(smem_u32(&bar_af[i]), 1);
      mbar_init(smem_u32(&bar_ae[i]), 1);
    }
    for (int i = 0; i < 8; ++i) {
      mbar_init(smem_u32(&bar_rf[i]), 1);
      mbar_init(smem_u32(&bar_re[i]), kWgThreads);
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    if (PAIR) {
      tmem_alloc_2sm(smem_u32(tmem_slot), Cfg::kTmemCols);
      tmem_relinquish_2sm();
    } else {
      tmem_alloc(smem_u32(tmem_slot), Cfg::kTmemCols);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all();  // the peer's barriers are initialised before anything arrives on them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // unit -> (b, t0, first n-tile); generic: unit = tile id with n fastest; A_STAT: unit = m-tile id
  auto decode = [&](int unit, int& b, int& t0, int& nt0) {
    int mt;
    if (A_STAT) {
      mt = unit;
      nt0 = 0;
    } else {
      mt = unit / p.n_tiles;
      nt0 = unit - mt * p.n_tiles;
    }
    if (PAIR) mt = 2 * mt + static_cast<int>(cta_rank);
    if (PAIR && mt >= total_mt) {  // odd tile count: the last pair's second tile is empty (loads zero-fill, stores clip)
      b = 0;
      t0 = p.m_tiles * kBlockM;
      return;
    }
    if (p.mt_map) mt = __ldg(p.mt_map + mt);  // ragged batch: the mt-th listed row tile
    b = mt / p.m_tiles;
    t0 = (mt - b * p.m_tiles) * kBlockM;
  };

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      uint32_t git = 0, ui = 0;
      for (int unit = unit0; unit < num_units; unit += ustride, ++ui) {
        int b, t0, nt0;
        decode(unit, b, t0, nt0);
        for (int nt = nt0; nt < nt0 + n_inner; ++nt) {
          for (int it = 0; it < num_iters; ++it, ++git) {
            if (A_STAT && nt == nt0) {
              // resident A, k-block `it`: free once the previous unit's last n-tile has read it; it is refilled
              // just ahead of the W stage that first needs it, so the A fetch pipelines with the weight stream
              mbar_wait(smem_u32(&bar_ae[it]), (ui & 1) ^ 1);
              const uint32_t af = smem_u32(&bar_af[it]);
              if (is_leader) mbar_arrive_expect_tx(af, (PAIR ? 2 : 1) * Cfg::kABytes);
              if (PAIR) tma_load_3d_2sm(smem_u32(sARes + it * Cfg::kABytes), &tmA, af, it * kBlockK, t0, b);
              else tma_load_3d(smem_u32(sARes + it * Cfg::kABytes), &tmA, af, it * kBlockK, t0, b);
            }
            const int s = git % S;
            const uint32_t ph = (git / S) & 1;
            mbar_wait(smem_u32(&bar_empty[s]), ph ^ 1);
            const uint32_t full = smem_u32(&bar_full[s]);
            if (is_leader) mbar_arrive_expect_tx(full, (PAIR ? 2 : 1) * Cfg::kStageBytes);
            const int tap = A_STAT ? 0 : it / kblocks;
            const int c0 = (it - tap * kblocks) * kBlockK;
            uint8_t* st = sStages + s * Cfg::kStageBytes;
            uint8_t* sw = st;
            if (!A_STAT) {
              if (PAIR) tma_load_3d_2sm(smem_u32(st), &tmA, full, c0, t0 + tap - p.pad_left, b);
              else tma_load_3d(smem_u32(st), &tmA, full, c0, t0 + tap - p.pad_left, b);
              sw = st + Cfg::kABytes;
            }
            if (PAIR) {
#pragma unroll
              for (int wb = 0; wb < Cfg::kNumWBoxes; ++wb)
                tma_load_2d_2sm(smem_u32(sw + wb * Cfg::kWBoxBytes), &tmW, full, c0,
                                tap * p.N + nt * BLOCK_N + wb * Cfg::kWBoxN + static_cast<int>(cta_rank) * Cfg::kWLoadN);
            } else {
#pragma unroll
              for (int wb = 0; wb < Cfg::kNumWBoxes; ++wb)
                tma_load_2d(smem_u32(sw + wb * Cfg::kWBoxBytes), &tmW, full, c0,
                            tap * p.N + nt * BLOCK_N + wb * Cfg::kWBoxN);
            }
          }
        }
      }
    }
  } else if (warp == 1 && is_leader) {
    // ===================== MMA issuer (pair: the leader CTA issues for both) =====================
    // The whole warp runs the loop converged (waits included); only the tcgen05 instructions are predicated on one
    // elected lane, so their operands are warp-uniform and live in uniform registers.  Under `if (lane == 0)` the
    // compiler wraps every UTCHMMA in an ELECT / R2UR / BRA.U.ANY waterfall loop, which made MMA issue the limiter
    // for the narrower tiles.
    {
      constexpr uint32_t idesc = umma_idesc_bf16(PAIR ? 2 * kBlockM : kBlockM, Cfg::kWBoxN, 0, 0);
      auto commit = [&](uint32_t bar) {
        if (PAIR) umma_commit_2sm(bar);
        else umma_commit(bar);
      };
      const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
      const uint32_t leader = elect_one();
      uint32_t git = 0, ti = 0, ui = 0;
      [[maybe_unused]] long long t_af = 0, t_acce = 0, t_full = 0, t_start = GT_NOW(), t0;
      for (int unit = unit0; unit < num_units; unit += ustride, ++ui) {
        for (int nt = 0; nt < n_inner; ++nt, ++ti) {
          const uint32_t ab = ti % Cfg::kNumAcc;
          const uint32_t aph = (ti / Cfg::kNumAcc) & 1;
          t0 = GT_NOW();
          mbar_wait(smem_u32(&bar_acce[ab]), aph ^ 1);  // epilogue drained this accumulator
          tc_fence_after();
          t_acce += GT_NOW() - t0;
          const uint32_t d_tmem = tb + ab * BLOCK_N;
          for (int it = 0; it < num_iters; ++it, ++git) {
            const int s = git % S;
            const uint32_t ph = (git / S) & 1;
            if (A_STAT && nt == 0) {
              t0 = GT_NOW();
              mbar_wait(smem_u32(&bar_af[it]), ui & 1);
              t_af += GT_NOW() - t0;
            }
            t0 = GT_NOW();
            mbar_wait(smem_u32(&bar_full[s]), ph);
            tc_fence_after();
            t_full += GT_NOW() - t0;
            const uint32_t st = smem_u32(sStages + s * Cfg::kStageBytes);
            const uint32_t a_base = A_STAT ? smem_u32(sARes + it * Cfg::kABytes) : st;
            const uint32_t w_base = A_STAT ? st : st + Cfg::kABytes;
            if (leader) {
#pragma unroll
              for (int k = 0; k < kBlockK / kUmmaK; ++k) {
                const uint64_t adesc = umma_smem_desc(a_base + k * kUmmaK * 2, 16, 1024, kLayoutSW128);
#pragma unroll
                for (int wb = 0; wb < Cfg::kNumWBoxes; ++wb) {
                  const uint64_t bdesc =
                      umma_smem_desc(w_base + wb * Cfg::kWBoxBytes + k * kUmmaK * 2, 16, 1024, kLayoutSW128);
                  if (PAIR) umma_bf16_2sm(d_tmem + wb * Cfg::kWBoxN, adesc, bdesc, idesc, (it | k) != 0 ? 1u : 0u);
                  else umma_bf16(d_tmem + wb * Cfg::kWBoxN, adesc, bdesc, idesc, (it | k) != 0 ? 1u : 0u);
                }
              }
              commit(smem_u32(&bar_empty[s]));  // frees the smem stage when these MMAs retire
              if (A_STAT && nt == n_inner - 1) commit(smem_u32(&bar_ae[it]));  // ... and this k-block of resident A
            }
            __syncwarp();
          }
          if (leader) commit(smem_u32(&bar_accf[ab]));  // accumulator complete
          __syncwarp();
        }
      }
      if (PAIR && A_STAT && ui > 0) mbar_wait(smem_u32(&bar_ae[kblocks - 1]), (ui - 1) & 1);  // last multicast commit has fired
#if defined(TTSB_GEMM_TRACE)
      if (lane == 0 && blockIdx.x < 148) {
        long long* o = g_gemm_trace + blockIdx.x * 32;
        o[0] = GT_NOW() - t_start; o[1] = t_af; o[2] = t_acce; o[3] = t_full; o[4] = git; o[5] = ti;
      }
#endif
    }
  } else if (Cfg::kRing && warp == 2) {
    // ===================== residual-ring producer =====================
    // Items are pushed in the order the two warpgroups consume them: item i belongs to half (i & 1).
    if (Cfg::kRing && lane == 0) {
      tma_prefetch_desc(&tmRes32);
      tma_prefetch_desc(&tmResBf);
      uint32_t i = 0;
      auto push = [&](const CUtensorMap* tm, int c0, int t0, int b) {
        const int slot = i % RS;
        mbar_wait(smem_u32(&bar_re[slot]), ((i / RS) & 1) ^ 1);
        const uint32_t full = smem_u32(&bar_rf[slot]);
        mbar_arrive_expect_tx(full, kBoxBytes16K);
        tma_load_3d(smem_u32(sRing + slot * kBoxBytes16K), tm, full, c0, t0, b);
        ++i;
      };
      for (int unit = unit0; unit < num_units; unit += ustride) {
        int b, t0, nt0;
        decode(unit, b, t0, nt0);
        if (EPI == TTSB200_EPI_FFN2)
          for (int g = 0; g < HALF / 64; ++g)
            for (int hh = 0; hh < 2; ++hh) push(&tmResBf, hh * HALF + g * 64, t0, b);  // bf16 inner residual
        for (int c = 0; c < HALF / 32; ++c)
          for (int hh = 0; hh < 2; ++hh) push(&tmRes32, hh * HALF + c * 32, t0, p.res_bcast ? 0 : b);  // fp32 stream
      }
    }
  } else if (!Cfg::kIsLN && (warp == 2 || warp == 3)) {
    // ===================== AFFINE store warps: warp 2 + g issues the TMA stores of epilogue warpgroup g =====================
    // The epilogue threads only fill staging boxes; issuing the bulk store (~190 clk for the issuing thread) and
    // waiting for it to have read the box happen here, off the epilogue's critical path.
    const int wg = warp - 2;
    const bool tma_out = (BLOCK_N % 64 == 0) && p.out_bf16 && !p.out_f32 && !p.res_f32;
    if (tma_out) {
      constexpr uint32_t NB = Cfg::kStoreBoxes;
      constexpr int kSplit = ((BLOCK_N / 2 + 31) / 32) * 32;  // as in the epilogue: warpgroup g's column range
      const int c_lo = wg == 0 ? 0 : kSplit;
      const int c_hi = wg == 0 ? kSplit : BLOCK_N;
      uint8_t* my_box = sStore + wg * (Cfg::kStoreBytes / 2);
      uint32_t sg = 0;
      for (int unit = unit0; unit < num_units; unit += ustride) {
        int b, t0, nt0;
        decode(unit, b, t0, nt0);
        for (int nt = nt0; nt < nt0 + n_inner; ++nt) {
          for (int c = c_lo; c < c_hi; c += 32, ++sg) {
            const uint32_t bx = sg % NB;
            nbar_sync(kBarFull0 + wg * 3 + bx, kWgThreads + 32);
            if (lane == 0) {
              tma_store_3d(&tmOut, smem_u32(my_box + bx * 8192), nt * BLOCK_N + c, t0, b);
              tma_store_commit();
              // at most NB - 1 stores may still be reading: the oldest box is free again
              if (NB == 3) tma_store_wait_read2();
              else if (NB == 2) tma_store_wait_read1();
              else tma_store_wait_read0();
            }
            __syncwarp();
            if (sg + 1 >= NB) nbar_arrive(kBarFree0 + wg * 3 + (sg + 1) % NB, kWgThreads + 32);
          }
        }
      }
      if (lane == 0) tma_store_wait_all();  // global writes complete (and every box read) before the CTA may exit
      __syncwarp();
      // the stores not yet reported free: the last NB-1 of them
      for (uint32_t g = (sg + 1 >= NB ? sg + 1 - NB : 0); g < sg; ++g)
        nbar_arrive(kBarFree0 + wg * 3 + g % NB, kWgThreads + 32);
    }
  } else if (warp >= kEpiWarp0) {
#include "gemm_epilogue.inc"
  }

  tc_fence_before();
  if (PAIR) {
    // every multicast commit has been consumed by a wait in both CTAs before either one leaves
    cluster_sync_all();
    if (warp == 1) tmem_dealloc_2sm(tmem_base, Cfg::kTmemCols);
  } else {
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

}  // namespace ttsb
