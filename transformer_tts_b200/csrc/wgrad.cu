// Weight-gradient GEMM of a conv1d / linear for sm_100a (training path).
//
//   dW[j, n, c] += sum_{b,t} dy[b, t, n] * x[b, t + j - pad_left, c]          (x is zero outside [0,T))
//
// This is the autograd backward of nn.Linear / nn.Conv1d w.r.t. the weight (the reference gets it from
// PyTorch autograd: train_fastspeech2.py:304 `loss.backward()` through Models/modules.py:49-51,68,83-84,
// Models/varianceadaptor.py:217,220, Models/postnets.py:67,71-74).
//
// The reduction runs over TIME, and both operands live in HBM time-major (B,T,channels) with the channels
// contiguous, i.e. both are "MN-major" for the tensor core: TMA brings [64 time rows x 64 channels]
// SWIZZLE_128B boxes and tcgen05.mma reads them through MN-major shared-memory descriptors (a_major =
// b_major = 1), no transpose pass ever touches HBM.  A conv tap is a time-shifted box of x; TMA's
// out-of-bounds zero fill is the conv padding and clips the ragged last chunk of each utterance.
//
// One CTA = one (tap, 128 output channels n, 128 input channels c) tile of dW and one contiguous range of
// the B * ceil(T/64) time chunks (split-K across blockIdx.y so that ~2 waves of CTAs cover the 148 SMs);
// the fp32 accumulator stays in TMEM over the whole range and is added to dW with red.global.add.f32
// (dW is the caller's zeroed / accumulating gradient buffer, arbitrary element strides so that the
// reference's Conv1d (out, in, k) layout is written directly).
//
// Bias gradient (db[n] = sum_{b,t} dy[b,t,n], the autograd backward of the bias): the tiles of x-channel block 0 / tap 0
// issue one more MMA per 16 time rows against a constant B operand of ones (N = 16, a 2 KB shared-memory region filled with
// bf16 1.0 — every element equal, so the operand layout does not matter) into 16 more TMEM columns; column 0 of that
// accumulator is the bias gradient of the tile's 128 channels.  It replaces a separate column-sum pass over dy.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer, warps 2..5 = epilogue.
#include "common.cuh"
#include "../../include/ttsb200.h"

namespace ttsb {

constexpr int kWgM = 128;       // dy channels per tile (UMMA M)
constexpr int kWgN = 128;       // x channels per tile (UMMA N)
constexpr int kWgK = 64;        // time rows per stage
constexpr int kWgStages = 6;
constexpr int kWgBox = 64 * 64 * 2;                  // 64 rows x 64 channels bf16 = 8 KB
constexpr int kWgStageBytes = 4 * kWgBox;            // A: 2 boxes, B: 2 boxes
constexpr int kWgThreads = 192;
constexpr int kWgOnes = 2048;                        // the ones operand: 16 time rows x 128 B
constexpr int kWgSmem = kWgStages * kWgStageBytes + kWgOnes + 1024 + 256;

struct WgradDev {
  int B, T, N, C, taps, pad_left;
  int n_tiles, c_tiles, chunks_per_utt, total_chunks, chunks_per_split;
  float* dw;
  float* db;             // optional bias gradient (N), accumulated
  long long sj, sn, sc;  // element strides of dW[j][n][c]
  int vec_ok;            // dW base and row strides keep 16-byte alignment for v4 reductions
};

__global__ void __launch_bounds__(kWgThreads, 1)
wgrad_kernel(const __grid_constant__ CUtensorMap tmDy, const __grid_constant__ CUtensorMap tmX, const WgradDev p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* ones = smem + kWgStages * kWgStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(ones + kWgOnes);
  uint64_t* bar_full = bars;
  uint64_t* bar_empty = bars + kWgStages;
  uint64_t* bar_acc = bars + 2 * kWgStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kWgStages + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  int tile = blockIdx.x;
  const int ct = tile % p.c_tiles;
  tile /= p.c_tiles;
  const int nt = tile % p.n_tiles;
  const int tap = tile / p.n_tiles;
  const int n0 = nt * kWgM, c0 = ct * kWgN;
  const int shift = tap - p.pad_left;
  const int ch_lo = blockIdx.y * p.chunks_per_split;
  int ch_hi = ch_lo + p.chunks_per_split;
  if (ch_hi > p.total_chunks) ch_hi = p.total_chunks;
  const int nch = ch_hi - ch_lo;  // may be <= 0 for the last split
  const bool do_bias = p.db != nullptr && ct == 0 && tap == 0;
  if (do_bias) {
    for (int i = threadIdx.x; i < kWgOnes / 4; i += kWgThreads) reinterpret_cast<uint32_t*>(ones)[i] = 0x3F803F80u;  // bf16 1.0 x 2
    fence_async_smem();
  }

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmDy);
    tma_prefetch_desc(&tmX);
    for (int s = 0; s < kWgStages; ++s) {
      mbar_init(smem_u32(&bar_full[s]), 1);
      mbar_init(smem_u32(&bar_empty[s]), 1);
    }
    mbar_init(smem_u32(bar_acc), 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(tmem_slot), 256);   // 128 columns of dW + 16 of the bias-gradient accumulator
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (nch > 0) {
    if (warp == 0) {
      if (lane == 0) {
        for (int i = 0; i < nch; ++i) {
          const int s = i % kWgStages;
          const uint32_t ph = (i / kWgStages) & 1;
          mbar_wait(smem_u32(&bar_empty[s]), ph ^ 1);
          const uint32_t full = smem_u32(&bar_full[s]);
          mbar_arrive_expect_tx(full, kWgStageBytes);
          const int ch = ch_lo + i;
          const int b = ch / p.chunks_per_utt;
          const int t0 = (ch - b * p.chunks_per_utt) * kWgK;
          uint8_t* st = smem + s * kWgStageBytes;
          tma_load_3d(smem_u32(st), &tmDy, full, n0, t0, b);
          tma_load_3d(smem_u32(st + kWgBox), &tmDy, full, n0 + 64, t0, b);
          tma_load_3d(smem_u32(st + 2 * kWgBox), &tmX, full, c0, t0 + shift, b);
          tma_load_3d(smem_u32(st + 3 * kWgBox), &tmX, full, c0 + 64, t0 + shift, b);
        }
      }
    } else if (warp == 1) {
      // whole warp converged, MMAs predicated on one elected lane (uniform-register operands, no waterfall loops)
      constexpr uint32_t idesc = umma_idesc_bf16(kWgM, kWgN, 1, 1);  // both operands MN-major
      constexpr uint32_t idesc_b = umma_idesc_bf16(kWgM, 16, 1, 1);
      const uint64_t odesc = umma_smem_desc(smem_u32(ones), kWgBox, 1024, kLayoutSW128);
      const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
      const uint32_t leader = elect_one();
      for (int i = 0; i < nch; ++i) {
        const int s = i % kWgStages;
        const uint32_t ph = (i / kWgStages) & 1;
        mbar_wait(smem_u32(&bar_full[s]), ph);
        tc_fence_after();
        const uint32_t a_base = smem_u32(smem + s * kWgStageBytes);
        const uint32_t b_base = a_base + 2 * kWgBox;
        if (leader) {
#pragma unroll
          for (int k = 0; k < kWgK / 16; ++k) {
            // 16 time rows per MMA: start at row 16k (128 B per row); LBO = next 64-channel box, SBO = 8 rows
            const uint64_t adesc = umma_smem_desc(a_base + k * 2048, kWgBox, 1024, kLayoutSW128);
            const uint64_t bdesc = umma_smem_desc(b_base + k * 2048, kWgBox, 1024, kLayoutSW128);
            umma_bf16(tb, adesc, bdesc, idesc, (i | k) != 0 ? 1u : 0u);
            if (do_bias) umma_bf16(tb + kWgN, adesc, odesc, idesc_b, (i | k) != 0 ? 1u : 0u);
          }
          umma_commit(smem_u32(&bar_empty[s]));
        }
        __syncwarp();
      }
      if (leader) umma_commit(smem_u32(bar_acc));
      __syncwarp();
    } else {
      // ---- epilogue: TMEM lane = n (dy channel), column = c (x channel) ----
      const int q = warp & 3;
      const int n = n0 + q * 32 + lane;
      const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
      mbar_wait(smem_u32(bar_acc), 0);
      tc_fence_after();
      float* base = p.dw + tap * p.sj + static_cast<long long>(n) * p.sn;
      if (do_bias) {
        float v[32];
        tmem_ld32(tmem_base + lane_off + kWgN, v);
        tmem_ld_wait();
        if (n < p.N) atomicAdd(p.db + n, v[0]);
      }
#pragma unroll 1
      for (int c = 0; c < kWgN; c += 32) {
        float v[32];
        tmem_ld32(tmem_base + lane_off + c, v);
        tmem_ld_wait();
        if (n < p.N) {
          if (p.sc == 1 && c0 + c + 32 <= p.C && p.vec_ok) {
            // contiguous x-channels: 16-byte vector reductions (red.global.add.v4.f32, sm_90+)
            float* dst = base + c0 + c;
#pragma unroll
            for (int i = 0; i < 32; i += 4)
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + i), "f"(v[i]), "f"(v[i + 1]),
                           "f"(v[i + 2]), "f"(v[i + 3])
                           : "memory");
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const int cc = c0 + c + i;
              if (cc < p.C) atomicAdd(base + static_cast<long long>(cc) * p.sc, v[i]);
            }
          }
        }
      }
      tc_fence_before();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, 256);
}

}  // namespace ttsb

using namespace ttsb;

extern "C" int ttsb200_conv_wgrad_bias(const void* dy, int64_t ld_dy, const void* x, int64_t ld_x, float* dw, float* db,
                                       int64_t stride_tap, int64_t stride_n, int64_t stride_c, int32_t B, int32_t T,
                                       int32_t N, int32_t C, int32_t taps, int32_t pad_left, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(dy && x && dw, "conv_wgrad: null pointer");
  TTSB_CHECK_ARG(B > 0 && T > 0 && N > 0 && C > 0 && taps > 0, "conv_wgrad: bad dims");
  TTSB_CHECK_ARG(ld_dy % 8 == 0 && ld_x % 8 == 0 && ld_dy >= N && ld_x >= C && N % 8 == 0 && C % 8 == 0,
                 "conv_wgrad: channel counts and leading dimensions must be multiples of 8");
  TTSB_CHECK_ARG((reinterpret_cast<uintptr_t>(dy) & 15) == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0,
                 "conv_wgrad: dy/x must be 16-byte aligned");
  CUtensorMap tmDy, tmX;
  const uint32_t box[3] = {64, kWgK, 1};
  {
    const uint64_t dims[3] = {static_cast<uint64_t>(N), static_cast<uint64_t>(T), static_cast<uint64_t>(B)};
    const uint64_t strides[2] = {static_cast<uint64_t>(ld_dy) * 2, static_cast<uint64_t>(ld_dy) * 2 * T};
    int rc = make_tmap_bf16(&tmDy, dy, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  }
  {
    const uint64_t dims[3] = {static_cast<uint64_t>(C), static_cast<uint64_t>(T), static_cast<uint64_t>(B)};
    const uint64_t strides[2] = {static_cast<uint64_t>(ld_x) * 2, static_cast<uint64_t>(ld_x) * 2 * T};
    int rc = make_tmap_bf16(&tmX, x, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  }
  TTSB_SMEM_OPT_IN(wgrad_kernel, kWgSmem);
  WgradDev d;
  d.B = B; d.T = T; d.N = N; d.C = C; d.taps = taps; d.pad_left = pad_left;
  d.n_tiles = (N + kWgM - 1) / kWgM;
  d.c_tiles = (C + kWgN - 1) / kWgN;
  d.chunks_per_utt = (T + kWgK - 1) / kWgK;
  d.total_chunks = B * d.chunks_per_utt;
  const int tiles = d.n_tiles * d.c_tiles * taps;
  // split-K over the time chunks: one CTA per SM is resident, so the launch runs in ceil(tiles * splits / 148) waves of
  // (chunks / splits) main-loop steps plus a fixed prologue + 128 x 128 reduction epilogue worth ~8 steps.  Take the cheapest
  // split count (e.g. 180 tiles of the encoder's k = 5 convs: 1 split = 2 waves x 72, 3 splits = 4 waves x 29).
  int splits = 1;
  {
    double best = 1e30;
    const int max_s = d.total_chunks < 32 ? d.total_chunks : 32;
    for (int sp = 1; sp <= max_s; ++sp) {
      const int waves = (tiles * sp + 147) / 148;
      const double cost = waves * (static_cast<double>((d.total_chunks + sp - 1) / sp) + 8.0);
      if (cost < best - 1e-9) { best = cost; splits = sp; }
    }
  }
  d.chunks_per_split = (d.total_chunks + splits - 1) / splits;
  splits = (d.total_chunks + d.chunks_per_split - 1) / d.chunks_per_split;
  d.dw = dw; d.db = db; d.sj = stride_tap; d.sn = stride_n; d.sc = stride_c;
  d.vec_ok = ((reinterpret_cast<uintptr_t>(dw) & 15) == 0 && stride_n % 4 == 0 && stride_tap % 4 == 0) ? 1 : 0;
  dim3 grid(tiles, splits);
  wgrad_kernel<<<grid, kWgThreads, kWgSmem, static_cast<cudaStream_t>(s)>>>(tmDy, tmX, d);
  TTSB_CHECK_CUDA(cudaGetLastError());
  return TTSB_OK;
}

extern "C" int ttsb200_conv_wgrad(const void* dy, int64_t ld_dy, const void* x, int64_t ld_x, float* dw,
                                  int64_t stride_tap, int64_t stride_n, int64_t stride_c, int32_t B, int32_t T,
                                  int32_t N, int32_t C, int32_t taps, int32_t pad_left, ttsb200_stream_t s) {
  return ttsb200_conv_wgrad_bias(dy, ld_dy, x, ld_x, dw, nullptr, stride_tap, stride_n, stride_c, B, T, N, C, taps, pad_left, s);
}
