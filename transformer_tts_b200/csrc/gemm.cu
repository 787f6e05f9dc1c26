// Implicit-GEMM conv1d / linear for sm_100a, v2: persistent CTAs, TMA -> tcgen05.mma -> TMEM with
// double-buffered accumulators, fused epilogues, bf16 outputs staged through swizzled smem and
// written with TMA stores.
//
//   acc[b,t,n] = sum_{j<taps} sum_{c<C} x[b, t+j-pad_left, c] * w[j, n, c]
//
// One tile = 128 time steps of ONE utterance x BLOCK_N output channels, so conv taps are plain
// time-shifted TMA boxes and TMA's out-of-bounds zero fill is the conv zero padding (and clips the
// ragged last tile on the store side).  Grid = min(#tiles, #SMs * CTAs/SM); each CTA walks tiles
// tile = blockIdx.x + i*gridDim.x with n fastest, so concurrently running CTAs share activations
// and weights in L2.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer (one elected
// thread), warps 2..5 = epilogue (thread == output row == TMEM lane).  With BLOCK_N <= 256 the
// accumulator is double buffered in TMEM so the epilogue of tile i overlaps the main loop of i+1.
//
// A-stationary mode (taps == 1, C <= 384, AFFINE epilogue): the 128 x C activation tile is loaded
// once and stays in smem while the CTA sweeps all N/BLOCK_N weight tiles; only weights stream through
// the stage ring.  That cuts L2->SM traffic, which bounds K=384 GEMMs at this tile size.
//
// Reference ops replaced: nn.Linear / nn.Conv1d call sites listed in include/ttsb200.h.
#include "common.cuh"
#include "../../include/ttsb200.h"

namespace ttsb {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;  // bf16 elements = one 128-byte swizzle row
constexpr int kUmmaK = 16;
constexpr int kGemmThreads = 224;  // 7 warps: TMA, MMA, 4 x epilogue, residual-ring producer
constexpr int kEpiThreads = 128;
constexpr int kMaxSmem = 232448;  // 227 KB opt-in limit per CTA

template <int BLOCK_N, bool A_STAT, bool RING = false>
struct GemmCfg {
  static constexpr int kWBoxN = BLOCK_N <= 256 ? BLOCK_N : BLOCK_N / 2;  // TMA box rows / UMMA N
  static constexpr int kNumWBoxes = BLOCK_N / kWBoxN;
  static constexpr int kABytes = kBlockM * kBlockK * 2;                  // 16 KB
  static constexpr int kWBytes = BLOCK_N * kBlockK * 2;
  static constexpr int kWBoxBytes = kWBoxN * kBlockK * 2;
  static constexpr int kStageBytes = A_STAT ? kWBytes : kABytes + kWBytes;
  static constexpr int kAResBytes = A_STAT ? 6 * kABytes : 0;            // resident A: up to C = 384
  static constexpr int kStoreBytes = 2 * kBlockM * 128;                  // two 128x64 bf16 staging boxes
  static constexpr bool kScaleInSmem = BLOCK_N <= 256;                   // 384-wide tiles read scale from L1
  static constexpr int kParamBytes = (kScaleInSmem ? 2 : 1) * BLOCK_N * 4;  // bias (+ scale) of the tile
  // no alignment slack: the kernel has no static smem, so the dynamic window starts 1024-aligned
  // (checked at run time; a misaligned base traps instead of corrupting swizzled tiles)
  // RING: residual operands of the LayerNorm epilogues (fp32 128x32 / bf16 128x64 boxes, 16 KB each) are
  // prefetched by TMA into a 4-slot ring instead of being read row-by-row from global memory
  static constexpr int kRingSlots = RING ? 4 : 0;
  static constexpr int kRingSlotBytes = 16384;
  static constexpr int kRingBytes = kRingSlots * kRingSlotBytes;
  static constexpr int kFixed = kAResBytes + kStoreBytes + kRingBytes + kParamBytes + 256 /*barriers*/;
  static constexpr int kStagesFit = (kMaxSmem - kFixed) / kStageBytes;
  static constexpr int kStages = kStagesFit > 6 ? 6 : kStagesFit;
  static constexpr int kNumAcc = BLOCK_N <= 256 ? 2 : 1;
  static constexpr int kTmemCols = (kNumAcc * BLOCK_N) <= 128 ? 128 : ((kNumAcc * BLOCK_N) <= 256 ? 256 : 512);
  static constexpr int kSmemBytes = kFixed + kStages * kStageBytes;
  static_assert(kStages >= 2, "not enough shared memory for a pipeline");
  static_assert(kWBoxBytes % 1024 == 0, "W box must keep 1024B swizzle-atom alignment");
  static_assert(kWBoxN % 16 == 0 && kWBoxN <= 256, "UMMA N");
};

struct GemmDev {
  int B, T, N, C, taps, pad_left, epi, act;
  int m_tiles, n_tiles, num_tiles;
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

// This thread's row of a TMA-loaded, 128B-swizzled ring box: 32 fp32 or 64 bf16 values.
__device__ __forceinline__ void ring_row_f32(const uint8_t* box, int row, float* v) {
  const uint8_t* rp = box + row * 128;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float4 f = *reinterpret_cast<const float4*>(rp + ((j ^ (row & 7)) << 4));
    v[4 * j] = f.x; v[4 * j + 1] = f.y; v[4 * j + 2] = f.z; v[4 * j + 3] = f.w;
  }
}
__device__ __forceinline__ void ring_row_bf16(const uint8_t* box, int row, float* v) {
  const uint8_t* rp = box + row * 128;
#pragma unroll
  for (int j = 0; j < 8; ++j)
    load_bf16x8(reinterpret_cast<const __nv_bfloat16*>(rp + ((j ^ (row & 7)) << 4)), v + 8 * j);
}

__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// TMA store of one 64-column x 128-row bf16 box from swizzled smem; rows beyond T are clipped.
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* m, uint32_t src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(m)), "r"(src), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// Writes 32 consecutive bf16 columns [c, c+32) of this thread's row into the 128x64 staging box
// (SWIZZLE_128B: 16-byte chunk j of row r lives at r*128 + ((j ^ (r&7)) << 4)).
__device__ __forceinline__ void stage_row32(uint8_t* box, int row, int c_in_box, const float* v) {
  uint8_t* rp = box + row * 128;
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    const int ch = (c_in_box >> 3) + g;
    uint4 u;
    u.x = pack_bf16(v[g * 8 + 0], v[g * 8 + 1]);
    u.y = pack_bf16(v[g * 8 + 2], v[g * 8 + 3]);
    u.z = pack_bf16(v[g * 8 + 4], v[g * 8 + 5]);
    u.w = pack_bf16(v[g * 8 + 6], v[g * 8 + 7]);
    *reinterpret_cast<uint4*>(rp + ((ch ^ (row & 7)) << 4)) = u;
  }
}

// Streams a [128 x 64*k] bf16 result through the two staging boxes.  `fill(c, v)` must produce
// 32 output values for columns [c, c+32) of this thread's row.  All 128 epilogue threads call it.
template <int NCOLS, class Fill>
__device__ __forceinline__ void store_tile_bf16(const CUtensorMap* tmOut, uint8_t* sStore, uint32_t& gc, int row,
                                                int et, int n0, int t0, int b, Fill&& fill) {
#pragma unroll 1
  for (int g = 0; g < NCOLS / 64; ++g, ++gc) {
    uint8_t* box = sStore + (gc & 1) * (kBlockM * 128);  // gc counts store groups across tiles
    if (et == 0) tma_store_wait_read<1>();  // the store issued from this box two groups ago has read it
    epi_bar();
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float v[32];
      fill(g * 64 + h * 32, v);
      stage_row32(box, row, h * 32, v);
    }
    fence_async_smem();
    epi_bar();
    if (et == 0) {
      tma_store_3d(tmOut, smem_u32(box), n0 + g * 64, t0, b);
      tma_store_commit();
    }
  }
}

}  // namespace ttsb

#include "gemm_kernel.cuh"  // the kernel (producer / MMA / epilogue roles)
#include "gemm_host.cuh"    // tensor maps, launch, extern "C" entry point
