// Fused FeedForward of a k=1 FFT block — the reference's ONE function (Models/modules.py:81-88) plus the outer
// residual and the next norm_1 (Models/layers.py:40,31) — as one kernel; included by gemm.cu.
//
//   h = relu(x W1^T + b1)            x: norm_2(y) bf16 (B,T,384); W1 (1536,384); the hidden activation h (B,T,1536)
//   u = h W2^T + b2 + x               W2 (384,1536)
//   z = y + LN_ff(u);  zn = LN_next(z)
//
// The two-kernel path writes h to HBM and reads it back: 434 MB of the pair's 812 MB per decoder layer at cfg2, 2.3x
// the op's algorithmic traffic (x in, y in, z out, zn out: 325 MB).  Here h never leaves the SM:
//
//   * one unit = a CTA PAIR's 2 x 128 rows; the 128 x 384 activation tile of each CTA stays resident in shared memory;
//   * the hidden dimension is walked in 24 chunks of 64: GEMM1 (M=256 pair, N=64, K=384) puts chunk j into one of two
//     64-column TMEM buffers; the epilogue warps read it, add b1, apply ReLU and write it back to the SAME TMEM buffer
//     as packed bf16 pairs; GEMM2 (M=256, N=2x192, K=64) takes that as its A operand FROM TMEM (tcgen05 TS form) and
//     accumulates u in the 384-column TMEM accumulator.  GEMM1 runs two chunks ahead of GEMM2 (the tensor pipe
//     executes one thread's MMAs in order, so a buffer is only overwritten after the GEMM2 that read it), which hides
//     the convert step.  h touches neither shared memory nor HBM;
//   * each CTA stages its own rows of x and HALF of every weight chunk (cta_group::2): 48 KB per chunk per SM for
//     1536 MMA clocks, under the ~42 B/clk L2 -> SM limit that a single-CTA version (96 KB per chunk) would hit;
//   * the LayerNorm epilogue is the FFN2 epilogue of gemm_epilogue.inc (two warpgroups, statistics through smem, fp32
//     residual stream through a TMA ring, outputs by TMA store); its ring and staging boxes ALIAS the weight stages,
//     which are idle between the last GEMM2 of a unit and the first weight load of the next.
#pragma once

namespace ttsb {
namespace ff {

constexpr int kD = 384;
constexpr int kHid = 1536;
constexpr int kChunk = 64;
constexpr int kChunks = kHid / kChunk;          // 24
constexpr int kKB = kD / kBlockK;               // 6 k-blocks of x
constexpr int kThreads = 384;                   // warp 0 TMA, warp 1 MMA issuer, warp 2 residual ring, warps 4..11 epilogue
constexpr int kXBytes = kKB * kBoxBytes16K;     // 96 KB
constexpr int kW1Box = 32 * 128;                // this CTA's 32 of the chunk's 64 hidden rows, one k-block
constexpr int kW1Stage = kKB * kW1Box;          // 24 KB
constexpr int kW2Box = 96 * 128;                // this CTA's 96 of 192 output rows
constexpr int kW2Stage = 2 * kW2Box;            // 24 KB
constexpr int kWStages = 2;
constexpr int kWBytes = kWStages * (kW1Stage + kW2Stage);  // 96 KB; aliased by the epilogue's fp32 residual ring
constexpr int kRing = 6;                        // residual boxes (16 KB) that alias the weight stages; the other 6 alias the x tile
constexpr int kResBoxes = kD / 32;              // 12 boxes of 128 rows x 32 fp32 columns: the whole residual tile, one slot each
constexpr int kStoreBytes = 0;                  // (the bf16 outputs are staged in the drained ring slots)
constexpr int kParamFloats = 5 * kD;            // b2, g1, b1, g2, b2
constexpr int kStatBytes = 2 * 2 * kBlockM * 8;
constexpr int kSmem = kXBytes + kWBytes + kStoreBytes + kParamFloats * 4 + kStatBytes + 512;
static_assert(kRing * kBoxBytes16K <= kWBytes, "the residual ring must fit in the weight stages it aliases");
static_assert(kSmem <= kMaxSmem, "shared memory budget");

__device__ long long g_ffn_trace[148 * 16];

struct Dev {
  int B, T, m_tiles, debug;
  const float* b1;       // [1536]
  const float* b2;       // [384]
  const float *ln1_g, *ln1_b, *ln2_g, *ln2_b;
  float eps;
  const int* mt_map;
  const int* mt_count;
};

enum {
  B_XF = 0,               // x tile landed (leader's; both CTAs' loads signal it)
  B_XE,                   // x tile free (multicast commit after the unit's last GEMM1)
  B_W1F, B_W1E = B_W1F + kWStages,
  B_W2F = B_W1E + kWStages, B_W2E = B_W2F + kWStages,
  B_A1F = B_W2E + kWStages,   // [2] hidden chunk in TMEM (multicast commit)
  B_HF = B_A1F + 2,           // [2] hidden chunk converted in BOTH CTAs (leader's; one arrival per epilogue warp)
  B_A2F = B_HF + 2,           // u complete (multicast commit)
  B_A2E,                      // u read back by both CTAs' epilogue warps (leader's)
  B_EPI,                      // this CTA's epilogue is done with the ring, which aliases the weight stages (local, 256 arrivals)
  B_XR,                       // [6] this CTA's epilogue has read the inner residual out of k-block i of the x tile (local, 128 arrivals)
  B_XS = B_XR + kKB,                       // the fp32 stores out of the residual boxes that alias the x tile have read them (local, 2 arrivals)
  B_RF,                       // [12] residual box landed
  B_COUNT = B_RF + kResBoxes
};

__device__ __forceinline__ void umma_bf16_ts_2sm(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc,
                                                 uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t"
      "}" ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_leader_rlx(uint32_t bar) {
  asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(bar & kPeerBitMask) : "memory");
}

__global__ void __launch_bounds__(kThreads, 1)
ffn_fused_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW1,
                 const __grid_constant__ CUtensorMap tmW2, const __grid_constant__ CUtensorMap tmOut,
                 const __grid_constant__ CUtensorMap tmOut32, const __grid_constant__ CUtensorMap tmRes32,
                 const __grid_constant__ CUtensorMap tmResBf, const Dev p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) {
    if (threadIdx.x == 0) printf("ttsb200: ffn_fused dynamic shared memory base is not 1024-byte aligned\n");
    __trap();
  }
  uint8_t* sX = smem;
  uint8_t* sW = smem + kXBytes;
  uint8_t* sW1 = sW;                               // [2] x 24 KB
  uint8_t* sW2 = sW + kWStages * kW1Stage;         // [2] x 24 KB
  uint8_t* sRing = sW;                             // aliases the weight stages: 6 x 16 KB
  float* sParam = reinterpret_cast<float*>(sW + kWBytes);
  float2* sStat = reinterpret_cast<float2*>(sParam + kParamFloats);
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(sStat) + kStatBytes);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + B_COUNT);
  auto bar = [&](int i) { return smem_u32(&bars[i]); };

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = cluster_ctarank();
  const bool is_leader = cta_rank == 0;
  const int total_mt = p.mt_count ? __ldg(p.mt_count) : p.B * p.m_tiles;
  const int num_units = (total_mt + 1) / 2;
  const int unit0 = static_cast<int>(blockIdx.x >> 1);
  const int ustride = static_cast<int>(gridDim.x >> 1);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmW1);
    tma_prefetch_desc(&tmW2);
    for (int i = 0; i < B_COUNT; ++i) {
      uint32_t cnt = 1;
      if (i == B_HF || i == B_HF + 1 || i == B_A2E) cnt = 16;      // 8 epilogue warps x 2 CTAs
      if (i == B_EPI) cnt = kEpiThreads;
      if (i >= B_XR && i < B_XR + kKB) cnt = kEpiThreads / 2;
      if (i == B_XS) cnt = 2;
      mbar_init(bar(i), cnt);
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc_2sm(smem_u32(tmem_slot), 512);
    tmem_relinquish_2sm();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_u = tmem_base;               // u: columns [0, 384)
  const uint32_t tmem_h = tmem_base + kD;          // hidden chunks: [384, 448) and [448, 512)

  // residual box idx = 2 * chunk + column half: 0..5 in the weight stages; 6.. in the x tile: the k-block that the box's
  // own warpgroup reads (and releases) during chunks 2 * (chunk - 3), 2 * (chunk - 3) + 1 of pass 1
  auto res_box = [&](int idx) -> uint8_t* {
    if (idx < kRing) return sRing + idx * kBoxBytes16K;
    return sX + ((idx & 1) * 3 + ((idx - kRing) >> 1)) * kBoxBytes16K;
  };
  auto decode = [&](int unit, int& b, int& t0) {
    int mt = 2 * unit + static_cast<int>(cta_rank);
    if (mt >= total_mt) {  // odd tile count: the last pair's second tile is empty (loads zero-fill, stores clip)
      b = 0;
      t0 = p.m_tiles * kBlockM;
      return;
    }
    if (p.mt_map) mt = __ldg(p.mt_map + mt);
    b = mt / p.m_tiles;
    t0 = (mt - b * p.m_tiles) * kBlockM;
  };

  if (warp == 0) {
    // ===================== TMA producer (both CTAs): x tile, then the weight chunks in consumption order =============
    if (lane == 0) {
      uint32_t g1 = 0, g2 = 0, ui = 0;
      auto load_w1 = [&](int j) {
        const int s = g1 % kWStages;
        mbar_wait(bar(B_W1E + s), ((g1 / kWStages) & 1) ^ 1);
        const uint32_t full = bar(B_W1F + s);
        if (is_leader) mbar_arrive_expect_tx(full, 2 * kW1Stage);
        const uint32_t st = smem_u32(sW1 + s * kW1Stage);
        const int row = j * kChunk + static_cast<int>(cta_rank) * 32;
#pragma unroll
        for (int kb = 0; kb < kKB; ++kb) tma_load_2d_2sm(st + kb * kW1Box, &tmW1, full, kb * kBlockK, row);
        ++g1;
      };
      auto load_w2 = [&](int j) {
        const int s = g2 % kWStages;
        mbar_wait(bar(B_W2E + s), ((g2 / kWStages) & 1) ^ 1);
        const uint32_t full = bar(B_W2F + s);
        if (is_leader) mbar_arrive_expect_tx(full, 2 * kW2Stage);
        const uint32_t st = smem_u32(sW2 + s * kW2Stage);
#pragma unroll
        for (int wb = 0; wb < 2; ++wb)
          tma_load_2d_2sm(st + wb * kW2Box, &tmW2, full, j * kChunk, wb * 192 + static_cast<int>(cta_rank) * 96);
        ++g2;
      };
      for (int unit = unit0; unit < num_units; unit += ustride, ++ui) {
        int b, t0;
        decode(unit, b, t0);
        if (ui > 0) {
          mbar_wait(bar(B_XE), (ui - 1) & 1);   // the previous unit's last GEMM1 has read the x tile ...
          mbar_wait(bar(B_XS), (ui - 1) & 1);   // ... its epilogue has read the inner residual (x itself) and the second
                                                // half of the residual boxes, which borrow the tile's memory, has left
        }
        const uint32_t xf = bar(B_XF);
        if (is_leader) mbar_arrive_expect_tx(xf, 2 * kXBytes);
#pragma unroll
        for (int kb = 0; kb < kKB; ++kb) tma_load_3d_2sm(smem_u32(sX + kb * kBoxBytes16K), &tmX, xf, kb * kBlockK, t0, b);
        // the W1 stages alias residual boxes 0..2, drained with pass 2 (B_XS above): the first two GEMM1s of this unit run
        // while pass 3 of the previous one is still storing; the W2 stages hold pass 3's staging boxes until its end
        load_w1(0);
        load_w1(1);
        for (int j = 0; j < kChunks; ++j) {
          if (j == 0 && ui > 0) mbar_wait(bar(B_EPI), (ui - 1) & 1);
          load_w2(j);
          if (j + 2 < kChunks) load_w1(j + 2);
        }
      }
    }
  } else if (warp == 1 && is_leader) {
    // ===================== MMA issuer (leader CTA; converged warp, one elected lane issues) =====================
    constexpr uint32_t idesc1 = umma_idesc_bf16(2 * kBlockM, kChunk, 0, 0);
    constexpr uint32_t idesc2 = umma_idesc_bf16(2 * kBlockM, 192, 0, 0);
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t leader = elect_one();
    const uint32_t x_base = smem_u32(sX);
    uint32_t g1 = 0, g2 = 0, ui = 0;
    const bool tr = (p.debug & 1) != 0;
    long long t_w1 = 0, t_w2 = 0, t_hf = 0, t_xf = 0, t_a2e = 0, tc0 = 0;
    const long long t_start = tr ? clock64() : 0;
    auto gemm1 = [&](int j, bool last) {   // hidden chunk j -> TMEM buffer j & 1
      const int s = g1 % kWStages;
      if (tr) tc0 = clock64();
      mbar_wait(bar(B_W1F + s), (g1 / kWStages) & 1);
      if (tr) t_w1 += clock64() - tc0;
      tc_fence_after();
      const uint32_t st = smem_u32(sW1 + s * kW1Stage);
      const uint32_t d = tb + kD + (j & 1) * kChunk;
      if (leader) {
#pragma unroll
        for (int kb = 0; kb < kKB; ++kb) {
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k)
            umma_bf16_2sm(d, umma_smem_desc(x_base + kb * kBoxBytes16K + k * 32, 16, 1024, kLayoutSW128),
                          umma_smem_desc(st + kb * kW1Box + k * 32, 16, 1024, kLayoutSW128), idesc1, (kb | k) != 0 ? 1u : 0u);
        }
        umma_commit_2sm(bar(B_W1E + s));
        if (last) umma_commit_2sm(bar(B_XE));
        umma_commit_2sm(bar(B_A1F + (j & 1)));
      }
      __syncwarp();
      ++g1;
    };
    for (int unit = unit0; unit < num_units; unit += ustride, ++ui) {
      if (tr) tc0 = clock64();
      mbar_wait(bar(B_XF), ui & 1);
      if (tr) t_xf += clock64() - tc0;
      gemm1(0, false);
      gemm1(1, false);
      for (int j = 0; j < kChunks; ++j) {
        const int s = g2 % kWStages;
        if (tr) tc0 = clock64();
        mbar_wait(bar(B_HF + (j & 1)), ((ui * (kChunks / 2) + (j >> 1))) & 1);   // relu(chunk j) sits in TMEM in both CTAs
        if (tr) { const long long t1 = clock64(); t_hf += t1 - tc0; tc0 = t1; }
        mbar_wait(bar(B_W2F + s), (g2 / kWStages) & 1);
        if (tr) { const long long t1 = clock64(); t_w2 += t1 - tc0; tc0 = t1; }
        if (j == 0 && ui > 0) mbar_wait(bar(B_A2E), (ui - 1) & 1);               // the previous unit's u has been read back
        if (tr) t_a2e += clock64() - tc0;
        tc_fence_after();
        const uint32_t st = smem_u32(sW2 + s * kW2Stage);
        if (leader) {
#pragma unroll
          for (int k = 0; k < kChunk / kUmmaK; ++k) {
            const uint32_t a_tmem = tb + kD + (j & 1) * kChunk + k * 8;   // 16 hidden values = 8 packed columns
#pragma unroll
            for (int wb = 0; wb < 2; ++wb)
              umma_bf16_ts_2sm(tb + wb * 192, a_tmem, umma_smem_desc(st + wb * kW2Box + k * 32, 16, 1024, kLayoutSW128),
                               idesc2, (j | k) != 0 ? 1u : 0u);
          }
          umma_commit_2sm(bar(B_W2E + s));
          if (j == kChunks - 1) umma_commit_2sm(bar(B_A2F));
        }
        __syncwarp();
        ++g2;
        if (j + 2 < kChunks) gemm1(j + 2, j + 2 == kChunks - 1);
      }
    }
    if (tr && lane == 0) {
      long long* o = g_ffn_trace + (blockIdx.x >> 1) * 16;
      o[0] = clock64() - t_start; o[1] = t_xf; o[2] = t_w1; o[3] = t_w2; o[4] = t_hf; o[5] = t_a2e; o[6] = ui;
    }
  } else if (warp == 2) {
    // ===================== residual producer: the fp32 residual tile y, 12 boxes, EVERY box in its own slot ============
    // Boxes 0..5 (chunks 0..2 of each column half) alias the weight stages: 0..2 the W1 stages, idle after the unit's last
    // GEMM1, 3..5 the W2 stages, idle after its last GEMM2.  Boxes 6..11 alias the x tile, k-block by k-block as pass 1
    // of the epilogue has read the inner residual out of it (the next unit's x tile is loaded after pass 2).  No slot is
    // recycled inside a unit, so no load waits for a store, and the loads leave in four groups instead of one burst
    // (every pair reaches its epilogue at the same time).
    if (lane == 0) {
      tma_prefetch_desc(&tmRes32);
      constexpr int HALF = kD / 2;
      uint32_t ui = 0;
      for (int unit = unit0; unit < num_units; unit += ustride, ++ui) {
        int b, t0;
        decode(unit, b, t0);
        auto push = [&](int idx) {   // idx = 2 * chunk + column half
          const uint32_t full = bar(B_RF + idx);
          mbar_arrive_expect_tx(full, kBoxBytes16K);
          tma_load_3d(smem_u32(res_box(idx)), &tmRes32, full, (idx & 1) * HALF + (idx >> 1) * 32, t0, b);
        };
        mbar_wait(bar(B_XE), ui & 1);
        for (int idx = 0; idx < 3; ++idx) push(idx);
        mbar_wait(bar(B_A2F), ui & 1);
        for (int idx = 3; idx < kRing; ++idx) push(idx);
        for (int k = 0; k < 3; ++k) {
          mbar_wait(bar(B_XR + k), ui & 1);
          push(kRing + 2 * k);
          mbar_wait(bar(B_XR + 3 + k), ui & 1);
          push(kRing + 2 * k + 1);
        }
      }
    }
  } else if (warp >= kEpiWarp0) {
    // ===================== epilogue warps 4..11: thread == row == TMEM lane; wg = column half =====================
    constexpr int BLOCK_N = kD;
    constexpr int HALF = kD / 2;
    constexpr int NCH = HALF / 32;
    const int q = warp & 3;
    const int wg = (warp - kEpiWarp0) >> 2;
    const int row = q * 32 + lane;
    const int et = threadIdx.x - kEpiWarp0 * 32;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    float* sBias = sParam;
    float* sG1 = sParam + BLOCK_N;
    float* sB1 = sParam + 2 * BLOCK_N;
    float* sG2 = sParam + 3 * BLOCK_N;
    float* sB2 = sParam + 4 * BLOCK_N;
    for (int i = et; i < BLOCK_N; i += kEpiThreads) {
      sBias[i] = __ldg(p.b2 + i);
      sG1[i] = __ldg(p.ln1_g + i);
      sB1[i] = __ldg(p.ln1_b + i);
      sG2[i] = __ldg(p.ln2_g + i);
      sB2[i] = __ldg(p.ln2_b + i);
    }
    epi_bar_all();
    const int c0h = wg * HALF;
    const float inv_n = 1.f / static_cast<float>(BLOCK_N);
    uint32_t xc = 0, ui = 0;
    long long t_ring = 0;
    auto exchange = [&](float& a, float& c) {
      float2* buf = sStat + (xc & 1) * 2 * kBlockM;
      buf[wg * kBlockM + row] = make_float2(a, c);
      epi_bar_all();
      const float2 o = buf[(wg ^ 1) * kBlockM + row];
      a += o.x;
      c += o.y;
      ++xc;
    };
    auto layernorm32 = [&](float* v, int c, float mean, float rstd, const float* sG, const float* sB) {
      const float nm = -mean * rstd;
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        const float4 g4 = *reinterpret_cast<const float4*>(sG + c + i);
        const float4 b4 = *reinterpret_cast<const float4*>(sB + c + i);
        v[i] = fmaf(fmaf(v[i], rstd, nm), g4.x, b4.x);
        v[i + 1] = fmaf(fmaf(v[i + 1], rstd, nm), g4.y, b4.y);
        v[i + 2] = fmaf(fmaf(v[i + 2], rstd, nm), g4.z, b4.z);
        v[i + 3] = fmaf(fmaf(v[i + 3], rstd, nm), g4.w, b4.w);
      }
    };

    for (int unit = unit0; unit < num_units; unit += ustride, ++ui) {
      int b, t0;
      decode(unit, b, t0);
      // ---- hidden stage: chunk j: fp32 accumulator -> + b1 -> ReLU -> packed bf16 back into the same TMEM buffer ----
      for (int j = 0; j < kChunks; ++j) {
        const uint32_t hb = j & 1;
        mbar_wait(bar(B_A1F + hb), (ui * (kChunks / 2) + (j >> 1)) & 1);
        tc_fence_after();
        float v[32];
        tmem_ld32(tmem_h + hb * kChunk + wg * 32 + lane_off, v);
        const float* bp = p.b1 + j * kChunk + wg * 32;
        float bv[32];
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          const float4 t = __ldg(reinterpret_cast<const float4*>(bp + i));
          bv[i] = t.x; bv[i + 1] = t.y; bv[i + 2] = t.z; bv[i + 3] = t.w;
        }
        tmem_ld_wait();
        // the packed write of column half 1 lands on fp32 columns [16, 32) of half 0: both warps of this lane quarter
        // must have their values in registers first
        asm volatile("bar.sync %0, 64;" ::"r"(4 + q) : "memory");
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) pk[i] = pack_bf16_relu(v[2 * i] + bv[2 * i], v[2 * i + 1] + bv[2 * i + 1]);
        tmem_st16(tmem_h + hb * kChunk + wg * 16 + lane_off, pk);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_leader_rlx(bar(B_HF + hb));
      }

      // ---- LayerNorm epilogue on u (the FFN2 epilogue of gemm_epilogue.inc) ----
      const int t = t0 + row;
      const bool valid = t < p.T;
      mbar_wait(bar(B_A2F), ui & 1);
      tc_fence_after();
      const long long te0 = (p.debug & 1) ? clock64() : 0;
      const uint32_t trow = tmem_u + lane_off + c0h;
      float sum = 0.f, sq = 0.f;
#pragma unroll 1
      for (int ch = 0; ch < NCH; ++ch) {
        const int cc = ch * 32;
        float v[32];
        tmem_ld32(trow + cc, v);
        tmem_ld_wait();
        // the FeedForward's inner residual is its own input: still resident in shared memory as the GEMM1 operand
        // (k-block = 64 columns, the layout of a SWIZZLE_128B bf16 box)
        float f[32];
        box_row_ld_bf16x32(sX + ((c0h + cc) >> 6) * kBoxBytes16K, row, (ch & 1) * 4, f);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = (v[i] + sBias[c0h + cc + i]) + f[i];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          sum += v[i];
          sq = fmaf(v[i], v[i], sq);
        }
        tmem_st32(trow + cc, v);
        if (ch & 1) mbar_arrive(bar(B_XR + ((c0h + cc) >> 6)));   // this k-block of the x tile may receive a residual box
      }
      tmem_st_wait();
      const long long te1 = (p.debug & 1) ? clock64() : 0;
      exchange(sum, sq);
      float mean = sum * inv_n;
      float rstd = rsqrtf(fmaxf(sq * inv_n - mean * mean, 0.f) + p.eps);
      float sum2 = 0.f, sq2 = 0.f;
#pragma unroll 1
      for (int ch = 0; ch < NCH; ++ch) {
        const int cc = ch * 32;
        float v[32];
        tmem_ld32(trow + cc, v);
        tmem_ld_wait();
        layernorm32(v, c0h + cc, mean, rstd, sG1, sB1);
        const int idx = 2 * ch + wg;
        uint8_t* box = res_box(idx);
        {
          const long long tr0 = (p.debug & 1) ? clock64() : 0;
          mbar_wait(bar(B_RF + idx), ui & 1);
          if (p.debug & 1) t_ring += clock64() - tr0;
        }
        float y[32];
        box_row_ld_f32(box, row, y);
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          v[i] += y[i];
          sum2 += v[i];
          sq2 = fmaf(v[i], v[i], sq2);
        }
        tmem_st32(trow + cc, v);
        box_row_st_f32(box, row, v);   // z formed in place in the residual box leaves by TMA store
        fence_async_smem();
        wg_bar(wg);
        if (row == 0) {
          tma_store_3d(&tmOut32, smem_u32(box), c0h + cc, t0, b);
          tma_store_commit();
        }
      }
      if (row == 0) {             // every fp32 store of this warpgroup has read its box: the x tile's memory is free
        tma_store_wait_read0();   // again, and pass 3 may stage its output in this warpgroup's first three boxes
        mbar_arrive(bar(B_XS));
      }
      tmem_st_wait();
      const long long te2 = (p.debug & 1) ? clock64() : 0;
      exchange(sum2, sq2);      // (a barrier over all epilogue threads: every slot of the ring is drained past this point)
      mean = sum2 * inv_n;
      rstd = rsqrtf(fmaxf(sq2 * inv_n - mean * mean, 0.f) + p.eps);
#pragma unroll 1
      for (int ch = 0; ch < NCH; ++ch) {
        const int cc = ch * 32;
        float v[32];
        tmem_ld32(trow + cc, v);
        tmem_ld_wait();
        if (ch == NCH - 1) {  // the last read of u: the next unit's first GEMM2 may overwrite it
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_leader_rlx(bar(B_A2E));
        }
        layernorm32(v, c0h + cc, mean, rstd, sG2, sB2);
        // staging: three 8 KB boxes per warpgroup in the W2 stages' memory (the W1 stages already receive the next unit's
        // weights), box ch % 3.  Its previous store (chunk ch - 3) has been read: the issuing thread waited for all but
        // the latest store before the warpgroup barrier of chunk ch - 1.
        uint8_t* sbox = sW2 + (wg * 3 + ch % 3) * 8192;
        box_row_st_bf16x32_sw64(sbox, row, v);
        fence_async_smem();
        if (row == 0) tma_store_wait_read1();
        wg_bar(wg);
        if (row == 0) {
          tma_store_3d(&tmOut, smem_u32(sbox), c0h + cc, t0, b);
          tma_store_commit();
        }
      }
      (void)valid;
      if (row == 0) tma_store_wait_read0();   // the boxes alias the W2 stages: the stores have read them
      if ((p.debug & 1) && et == 0 && is_leader) {
        long long* o = g_ffn_trace + (blockIdx.x >> 1) * 16;
        const long long te3 = clock64();
        o[8] += te3 - te0; o[9] += te1 - te0; o[10] += te2 - te1; o[11] += te3 - te2; o[12] += t_ring;
        t_ring = 0;
      }
      mbar_arrive(bar(B_EPI));                // the next unit's weights may stream in
    }
    if (row == 0) tma_store_wait_all();
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) tmem_dealloc_2sm(tmem_base, 512);
}

}  // namespace ff
}  // namespace ttsb

static int g_ffn_debug = 0;
extern "C" int ttsb200_debug_ffn_trace(long long* dst, int n, int mode) {  // mode 1: enable + clear, 2: read, 0: disable
  if (mode == 1 || mode == 0) {
    g_ffn_debug = mode;
    static long long zeros[148 * 16] = {0};
    return cudaMemcpyToSymbol(ttsb::ff::g_ffn_trace, zeros, sizeof(zeros)) == cudaSuccess ? 0 : -1;
  }
  return cudaMemcpyFromSymbol(dst, ttsb::ff::g_ffn_trace, n * sizeof(long long)) == cudaSuccess ? 0 : -1;
}

// z = y + LN_ff(relu(x W1^T + b1) W2^T + b2 + x);  zn = LN_next(z)   (Models/modules.py:81-88, Models/layers.py:40,31)
extern "C" int ttsb200_ffn_fused(const void* x_bf16, const void* w1_bf16, const float* b1, const void* w2_bf16,
                                 const float* b2, const float* y_f32, const float* ln1_g, const float* ln1_b,
                                 const float* ln2_g, const float* ln2_b, float eps, float* z_f32, void* zn_bf16,
                                 int32_t B, int32_t T, int32_t D, int32_t hidden, const int32_t* tile_map,
                                 const int32_t* tile_count, ttsb200_stream_t stream_) {
  using namespace ttsb;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  TTSB_CHECK_ARG(x_bf16 && w1_bf16 && b1 && w2_bf16 && b2 && y_f32 && ln1_g && ln1_b && ln2_g && ln2_b && z_f32 && zn_bf16,
                 "ffn_fused: null pointer");
  TTSB_CHECK_ARG(B > 0 && T > 0 && D == ff::kD && hidden == ff::kHid, "ffn_fused: this build is for D=384, hidden=1536");
  TTSB_CHECK_ARG((tile_map == nullptr) == (tile_count == nullptr), "ffn_fused: tile_map and tile_count go together");
  TTSB_CHECK_ARG(aligned16(x_bf16) && aligned16(w1_bf16) && aligned16(w2_bf16) && aligned16(y_f32) && aligned16(z_f32) &&
                     aligned16(zn_bf16), "ffn_fused: operands must be 16-byte aligned");
  CUtensorMap tmX, tmW1, tmW2, tmOut, tmOut32, tmRes32, tmResBf;
  const uint64_t uT = static_cast<uint64_t>(T), uB = static_cast<uint64_t>(B), uD = ff::kD, uH = ff::kHid;
  const uint64_t dims3[3] = {uD, uT, uB};
  const uint64_t str_bf[2] = {uD * 2, uD * 2 * uT};
  const uint64_t str_32[2] = {uD * 4, uD * 4 * uT};
  int rc;
  {
    const uint32_t box[3] = {kBlockK, kBlockM, 1};
    if ((rc = make_tmap_bf16(&tmX, x_bf16, 3, dims3, str_bf, box, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
    if ((rc = make_tmap_bf16(&tmResBf, x_bf16, 3, dims3, str_bf, box, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  }
  {
    const uint64_t dims[2] = {uD, uH};
    const uint64_t strides[1] = {uD * 2};
    const uint32_t box[2] = {kBlockK, 32};
    if ((rc = make_tmap_bf16(&tmW1, w1_bf16, 2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  }
  {
    const uint64_t dims[2] = {uH, uD};
    const uint64_t strides[1] = {uH * 2};
    const uint32_t box[2] = {kBlockK, 96};
    if ((rc = make_tmap_bf16(&tmW2, w2_bf16, 2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  }
  {
    const uint32_t box[3] = {32, kBlockM, 1};
    if ((rc = make_tmap_bf16(&tmOut, zn_bf16, 3, dims3, str_bf, box, CU_TENSOR_MAP_SWIZZLE_64B))) return rc;
    if ((rc = make_tmap(&tmOut32, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, z_f32, 3, dims3, str_32, box, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
    if ((rc = make_tmap(&tmRes32, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, y_f32, 3, dims3, str_32, box, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  }
  auto kern = ff::ffn_fused_kernel;
  TTSB_SMEM_OPT_IN(kern, ff::kSmem);
  ff::Dev d;
  d.B = B; d.T = T; d.m_tiles = (T + kBlockM - 1) / kBlockM;
  d.b1 = b1; d.b2 = b2; d.ln1_g = ln1_g; d.ln1_b = ln1_b; d.ln2_g = ln2_g; d.ln2_b = ln2_b; d.eps = eps;
  d.mt_map = tile_map; d.mt_count = tile_count;
  d.debug = g_ffn_debug;
  const int units = (B * d.m_tiles + 1) / 2;
  const int pairs = units < num_sms() / 2 ? units : num_sms() / 2;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * pairs);
  cfg.blockDim = dim3(ff::kThreads);
  cfg.dynamicSmemBytes = ff::kSmem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  TTSB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, tmX, tmW1, tmW2, tmOut, tmOut32, tmRes32, tmResBf, d));
  return TTSB_OK;
}
