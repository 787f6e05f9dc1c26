// conv_gemm_kernel — included by gemm.cu (see the header comment there for the design).
#pragma once

namespace ttsb {

template <int BLOCK_N, int EPI, bool A_STAT>
__global__ void __launch_bounds__(kGemmThreads, 1)
conv_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                 const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmOut32,
                 const __grid_constant__ CUtensorMap tmRes32, const __grid_constant__ CUtensorMap tmResBf,
                 const GemmDev p) {
  using Cfg = GemmCfg<BLOCK_N, EPI, A_STAT>;
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
  uint64_t* bar_af = bars + 2 * S + 4;   // resident A full
  uint64_t* bar_ae = bars + 2 * S + 5;   // resident A free
  uint64_t* bar_rf = bars + 2 * S + 6;   // [4] ring slot filled
  uint64_t* bar_re = bars + 2 * S + 10;  // [4] ring slot drained
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * S + 14);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int kblocks = (p.C + kBlockK - 1) / kBlockK;
  const int num_iters = A_STAT ? kblocks : p.taps * kblocks;
  const int num_units = A_STAT ? p.B * p.m_tiles : p.num_tiles;
  const int n_inner = A_STAT ? p.n_tiles : 1;

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
      mbar_init(smem_u32(&bar_acce[i]), Cfg::kIsLN ? kEpiThreads : kWgThreads);
    }
    mbar_init(smem_u32(bar_af), 1);
    mbar_init(smem_u32(bar_ae), 1);
    for (int i = 0; i < 4; ++i) {
      mbar_init(smem_u32(&bar_rf[i]), 1);
      mbar_init(smem_u32(&bar_re[i]), kWgThreads);
    }
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
    b = mt / p.m_tiles;
    t0 = (mt - b * p.m_tiles) * kBlockM;
  };

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      uint32_t git = 0, ui = 0;
      for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x, ++ui) {
        int b, t0, nt0;
        decode(unit, b, t0, nt0);
        if (A_STAT) {
          mbar_wait(smem_u32(bar_ae), (ui & 1) ^ 1);
          mbar_arrive_expect_tx(smem_u32(bar_af), kblocks * Cfg::kABytes);
          for (int kb = 0; kb < kblocks; ++kb)
            tma_load_3d(smem_u32(sARes + kb * Cfg::kABytes), &tmA, smem_u32(bar_af), kb * kBlockK, t0, b);
        }
        for (int nt = nt0; nt < nt0 + n_inner; ++nt) {
          for (int it = 0; it < num_iters; ++it, ++git) {
            const int s = git % S;
            const uint32_t ph = (git / S) & 1;
            mbar_wait(smem_u32(&bar_empty[s]), ph ^ 1);
            const uint32_t full = smem_u32(&bar_full[s]);
            mbar_arrive_expect_tx(full, Cfg::kStageBytes);
            const int tap = A_STAT ? 0 : it / kblocks;
            const int c0 = (it - tap * kblocks) * kBlockK;
            uint8_t* st = sStages + s * Cfg::kStageBytes;
            uint8_t* sw = st;
            if (!A_STAT) {
              tma_load_3d(smem_u32(st), &tmA, full, c0, t0 + tap - p.pad_left, b);
              sw = st + Cfg::kABytes;
            }
#pragma unroll
            for (int wb = 0; wb < Cfg::kNumWBoxes; ++wb)
              tma_load_2d(smem_u32(sw + wb * Cfg::kWBoxBytes), &tmW, full, c0,
                          tap * p.N + nt * BLOCK_N + wb * Cfg::kWBoxN);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // The whole warp runs the loop converged (waits included); only the tcgen05 instructions are predicated on one
    // elected lane, so their operands are warp-uniform and live in uniform registers.  Under `if (lane == 0)` the
    // compiler wraps every UTCHMMA in an ELECT / R2UR / BRA.U.ANY waterfall loop, which made MMA issue the limiter
    // for the narrower tiles.
    {
      constexpr uint32_t idesc = umma_idesc_bf16(kBlockM, Cfg::kWBoxN, 0, 0);
      const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
      const uint32_t leader = elect_one();
      uint32_t git = 0, ti = 0, ui = 0;
      for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x, ++ui) {
        if (A_STAT) {
          mbar_wait(smem_u32(bar_af), ui & 1);
          tc_fence_after();
        }
        for (int nt = 0; nt < n_inner; ++nt, ++ti) {
          const uint32_t ab = ti % Cfg::kNumAcc;
          const uint32_t aph = (ti / Cfg::kNumAcc) & 1;
          mbar_wait(smem_u32(&bar_acce[ab]), aph ^ 1);  // epilogue drained this accumulator
          tc_fence_after();
          const uint32_t d_tmem = tb + ab * BLOCK_N;
          for (int it = 0; it < num_iters; ++it, ++git) {
            const int s = git % S;
            const uint32_t ph = (git / S) & 1;
            mbar_wait(smem_u32(&bar_full[s]), ph);
            tc_fence_after();
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
                  umma_bf16(d_tmem + wb * Cfg::kWBoxN, adesc, bdesc, idesc, (it | k) != 0 ? 1u : 0u);
                }
              }
              umma_commit(smem_u32(&bar_empty[s]));  // frees the smem stage when these MMAs retire
            }
            __syncwarp();
          }
          if (leader) umma_commit(smem_u32(&bar_accf[ab]));  // accumulator complete
          __syncwarp();
        }
        if (A_STAT) {
          if (leader) umma_commit(smem_u32(bar_ae));  // resident A may be overwritten
          __syncwarp();
        }
      }
    }
  } else if (warp == 2) {
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
      for (int unit = blockIdx.x; unit < num_units; unit += gridDim.x) {
        int b, t0, nt0;
        decode(unit, b, t0, nt0);
        if (EPI == TTSB200_EPI_FFN2)
          for (int g = 0; g < HALF / 64; ++g)
            for (int hh = 0; hh < 2; ++hh) push(&tmResBf, hh * HALF + g * 64, t0, b);  // bf16 inner residual
        for (int c = 0; c < HALF / 32; ++c)
          for (int hh = 0; hh < 2; ++hh) push(&tmRes32, hh * HALF + c * 32, t0, p.res_bcast ? 0 : b);  // fp32 stream
      }
    }
  } else if (warp >= kEpiWarp0) {
#include "gemm_epilogue.inc"
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::kTmemCols);
}

}  // namespace ttsb
