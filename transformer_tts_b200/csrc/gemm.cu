// Implicit-GEMM conv1d / linear for sm_100a, v3: persistent CTAs, TMA -> tcgen05.mma -> TMEM,
// two epilogue warpgroups, every large output staged through swizzled smem and written by TMA.
//
//   acc[b,t,n] = sum_{j<taps} sum_{c<C} x[b, t+j-pad_left, c] * w[j, n, c]
//
// One tile = 128 time steps of ONE utterance x BLOCK_N output channels, so conv taps are plain
// time-shifted TMA boxes and TMA's out-of-bounds zero fill is the conv zero padding (and clips the
// ragged last tile on the store side).  Grid = min(#units, #SMs); each CTA walks units
// unit = blockIdx.x + i*gridDim.x with n fastest, so concurrently running CTAs share activations
// and weights in L2.
//
// Warp roles (384 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer (one elected
// thread), warp 2 = residual-ring producer (LayerNorm epilogues with a residual), warp 3 idle,
// warps 4..11 = two epilogue warpgroups (thread == output row == TMEM lane).
//   * AFFINE epilogue (BLOCK_N <= 256, double-buffered accumulators): warpgroup g owns accumulator g,
//     i.e. alternate tiles, so each epilogue has two main loops of time to drain.
//   * LayerNorm epilogues (the CTA owns whole rows, BLOCK_N == N): the two warpgroups split the
//     columns in halves and exchange the per-row (sum, sum of squares) through smem.
//     Residual operands arrive through a TMA-fed ring of 16 KB boxes; the fp32 residual-stream output
//     is formed IN PLACE in its ring box and leaves by TMA store, so no thread ever issues a strided
//     global load or store.
//
// A-stationary mode (taps == 1, C <= 384, AFFINE epilogue): the 128 x C activation tile is loaded
// once and stays in smem while the CTA sweeps all N/BLOCK_N weight tiles; only weights stream through
// the stage ring.
//
// Reference ops replaced: nn.Linear / nn.Conv1d call sites listed in include/ttsb200.h.
#include "common.cuh"
#include "../../include/ttsb200.h"

namespace ttsb {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;  // bf16 elements = one 128-byte swizzle row
constexpr int kUmmaK = 16;
constexpr int kGemmThreads = 384;  // 12 warps
constexpr int kEpiWarp0 = 4;
constexpr int kEpiThreads = 256;
constexpr int kWgThreads = 128;
#ifndef TTSB_STORE_BOXES
#define TTSB_STORE_BOXES 2   // bf16 staging boxes per epilogue warpgroup of the CTA-pair AFFINE kernels (1..3)
#endif
constexpr int kMaxSmem = 232448;  // 227 KB opt-in limit per CTA
constexpr int kBoxBytes16K = 16384;

template <int BLOCK_N, int EPI, bool A_STAT, bool PAIR = false>
struct GemmCfg {
  static constexpr bool kIsLN = EPI != TTSB200_EPI_AFFINE_ACT;
  static constexpr bool kRing = EPI == TTSB200_EPI_RES_LN || EPI == TTSB200_EPI_FFN2;
  static constexpr int kWBoxN = BLOCK_N <= 256 ? BLOCK_N : BLOCK_N / 2;  // TMA box rows / UMMA N
  static constexpr int kNumWBoxes = BLOCK_N / kWBoxN;
  static constexpr int kABytes = kBlockM * kBlockK * 2;                  // 16 KB
  // CTA pair: each CTA of the pair stages half of the W tile's rows (the M=256 MMA reads both halves)
  static constexpr int kWLoadN = PAIR ? kWBoxN / 2 : kWBoxN;             // rows per TMA box in this CTA
  static constexpr int kWBoxBytes = kWLoadN * kBlockK * 2;
  static constexpr int kWBytes = kNumWBoxes * kWBoxBytes;
  static constexpr int kStageBytes = A_STAT ? kWBytes : kABytes + kWBytes;
  static constexpr int kAResBytes = A_STAT ? 6 * kABytes : 0;            // resident A: up to C = 384
  // one bf16 staging box per warpgroup: 128 rows x 32 columns (SWIZZLE_64B)
  // (CTA pairs have the room for two per warpgroup: the TMA store of one chunk overlaps the next chunk's math)
  static constexpr int kStoreBoxes = (PAIR && !kIsLN) ? TTSB_STORE_BOXES : 1;
  static constexpr int kStoreBytes = 2 * kStoreBoxes * 8192;
#ifndef TTSB_RING_SLOTS
#define TTSB_RING_SLOTS 4
#endif
  static constexpr int kRingSlots = kRing ? (PAIR ? TTSB_RING_SLOTS : 4) : 0;  // even: each half owns its own slots
  static constexpr int kRingBytes = kRingSlots * kBoxBytes16K;
  // parameter vectors in smem: AFFINE: (bias, scale) per warpgroup; LN: bias, g1, b1, g2, b2, dot_w
  static constexpr int kParamFloats = kIsLN ? 6 * BLOCK_N : 4 * BLOCK_N;
  static constexpr int kStatBytes = kIsLN ? 2 * 2 * kBlockM * 8 : 0;     // [buf][half][row] float2
  static constexpr int kFixed = kAResBytes + kStoreBytes + kRingBytes + kParamFloats * 4 + kStatBytes + 512;
  static constexpr int kStagesFit = (kMaxSmem - kFixed) / kStageBytes;
  static constexpr int kMaxStages = PAIR ? 7 : 6;
  static constexpr int kStages = kStagesFit > kMaxStages ? kMaxStages : kStagesFit;
  static constexpr int kNumAcc = BLOCK_N <= 256 ? 2 : 1;
  static constexpr int kTmemCols = (kNumAcc * BLOCK_N) <= 128 ? 128 : ((kNumAcc * BLOCK_N) <= 256 ? 256 : 512);
  static constexpr int kSmemBytes = kFixed + kStages * kStageBytes;
  static_assert(kStages >= 2, "not enough shared memory for a pipeline");
  static_assert(kWBoxBytes % 1024 == 0, "W box must keep 1024B swizzle-atom alignment");
  static_assert(kWBoxN % 16 == 0 && kWBoxN <= 256, "UMMA N");
  static_assert(!kIsLN || (BLOCK_N % 128 == 0), "LN epilogues split the row in two halves of 64-column groups");
  static_assert(kIsLN || kNumAcc == 2, "AFFINE epilogue: one accumulator per epilogue warpgroup");
  static_assert(!PAIR || kWBoxN % 32 == 0, "CTA pairs: each CTA stages half of every W box");
};

struct GemmDev {
  int B, T, N, C, taps, pad_left, epi, act;
  int m_tiles, n_tiles, num_tiles;
  int res_bcast;  // residual has no batch dimension (positional table)
  const float* bias;
  const float* scale;
  const float* res_f32;
  long long res_bstride, res_tstride;
  const float* res_scale;
  const float *ln1_g, *ln1_b, *ln2_g, *ln2_b;
  float eps;
  const float* dot_w;
  const float* dot_b;
  const uint8_t* row_mask;
  float* out_f32;
  __nv_bfloat16* out_bf16;
  float* out2_f32;
  float* out_row;
  float* pre_f32;                 // LN epilogues: the pre-LayerNorm value (saved for the backward pass)
  const __nv_bfloat16* gate_bf16; // AFFINE + ACT_RELU_GRAD: v = gate > 0 ? v : 0
  const int* mt_map;              // ragged batches: compacted list of row tiles (b * m_tiles + tile) ...
  const int* mt_count;            // ... and its length, both on the device (NULL = all B * m_tiles tiles)
};

__device__ __forceinline__ float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// This thread's row of a TMA-loaded, 128B-swizzled box (128 rows x 128 bytes).
__device__ __forceinline__ void box_row_ld_f32(const uint8_t* box, int row, float* v) {
  const uint8_t* rp = box + row * 128;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float4 f = *reinterpret_cast<const float4*>(rp + ((j ^ (row & 7)) << 4));
    v[4 * j] = f.x; v[4 * j + 1] = f.y; v[4 * j + 2] = f.z; v[4 * j + 3] = f.w;
  }
}
__device__ __forceinline__ void box_row_st_f32(uint8_t* box, int row, const float* v) {
  uint8_t* rp = box + row * 128;
#pragma unroll
  for (int j = 0; j < 8; ++j)
    *reinterpret_cast<float4*>(rp + ((j ^ (row & 7)) << 4)) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
}
// 32 bf16 values = 16-byte chunks [j0, j0+4) of the row
__device__ __forceinline__ void box_row_ld_bf16x32(const uint8_t* box, int row, int j0, float* v) {
  const uint8_t* rp = box + row * 128;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const uint4 u = *reinterpret_cast<const uint4*>(rp + (((j0 + j) ^ (row & 7)) << 4));
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __bfloat1622float2(h[i]);
      v[8 * j + 2 * i] = f.x;
      v[8 * j + 2 * i + 1] = f.y;
    }
  }
}
// 32 bf16 values -> 16-byte chunks [j0, j0+4) of this thread's row of a 128x64 bf16 staging box
__device__ __forceinline__ void box_row_st_bf16x32(uint8_t* box, int row, int j0, const float* v) {
  uint8_t* rp = box + row * 128;
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    uint4 u;
    u.x = pack_bf16(v[g * 8 + 0], v[g * 8 + 1]);
    u.y = pack_bf16(v[g * 8 + 2], v[g * 8 + 3]);
    u.z = pack_bf16(v[g * 8 + 4], v[g * 8 + 5]);
    u.w = pack_bf16(v[g * 8 + 6], v[g * 8 + 7]);
    *reinterpret_cast<uint4*>(rp + (((j0 + g) ^ (row & 7)) << 4)) = u;
  }
}

// 32 bf16 values = this thread's whole 64-byte row of a 128x32 bf16 staging box (SWIZZLE_64B:
// 16-byte chunk j of row r lives at r*64 + ((j ^ ((r >> 1) & 3)) << 4))
__device__ __forceinline__ void box_row_st_bf16x32_sw64(uint8_t* box, int row, const float* v) {
  uint8_t* rp = box + row * 64;
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    uint4 u;
    u.x = pack_bf16(v[g * 8 + 0], v[g * 8 + 1]);
    u.y = pack_bf16(v[g * 8 + 2], v[g * 8 + 3]);
    u.z = pack_bf16(v[g * 8 + 4], v[g * 8 + 5]);
    u.w = pack_bf16(v[g * 8 + 6], v[g * 8 + 7]);
    *reinterpret_cast<uint4*>(rp + ((g ^ ((row >> 1) & 3)) << 4)) = u;
  }
}

// max(x, 0) and round-to-nearest bf16 of a pair in one instruction (NaN stays NaN, as torch.relu keeps it)
__device__ __forceinline__ uint32_t pack_bf16_relu(float lo, float hi) {
  uint32_t d;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}
template <bool RELU>
__device__ __forceinline__ void box_row_st_bf16x32_sw64_t(uint8_t* box, int row, const float* v) {
  uint8_t* rp = box + row * 64;
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    uint4 u;
    if (RELU) {
      u.x = pack_bf16_relu(v[g * 8 + 0], v[g * 8 + 1]);
      u.y = pack_bf16_relu(v[g * 8 + 2], v[g * 8 + 3]);
      u.z = pack_bf16_relu(v[g * 8 + 4], v[g * 8 + 5]);
      u.w = pack_bf16_relu(v[g * 8 + 6], v[g * 8 + 7]);
    } else {
      u.x = pack_bf16(v[g * 8 + 0], v[g * 8 + 1]);
      u.y = pack_bf16(v[g * 8 + 2], v[g * 8 + 3]);
      u.z = pack_bf16(v[g * 8 + 4], v[g * 8 + 5]);
      u.w = pack_bf16(v[g * 8 + 6], v[g * 8 + 7]);
    }
    *reinterpret_cast<uint4*>(rp + ((g ^ ((row >> 1) & 3)) << 4)) = u;
  }
}

__device__ __forceinline__ void epi_bar_all() { asm volatile("bar.sync 1, 256;" ::: "memory"); }
// named barriers 4..7 / 8..11: staging box (warpgroup, box) full / free, between an epilogue warpgroup and its store warp
constexpr int kBarFull0 = 4;                          // [2 warpgroups][kStoreBoxes <= 3]
constexpr int kBarFree0 = 4 + 2 * 3;                  // ids 10..15
__device__ __forceinline__ void nbar_sync(int id, int count) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void nbar_arrive(int id, int count) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void wg_bar(int wg) { asm volatile("bar.sync %0, 128;" ::"r"(2 + wg) : "memory"); }

// TMA store of one 128-row box from swizzled smem; rows beyond T are clipped by the tensor map.
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* m, uint32_t src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(m)), "r"(src), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read2() { asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

}  // namespace ttsb

#include "gemm_kernel.cuh"  // the kernel (producer / MMA / epilogue roles)
#include "gemm_host.cuh"    // tensor maps, launch, extern "C" entry point
#include "ffn_fused.cuh"    // fused FeedForward of the k=1 decoder blocks
