// Fused masked multi-head self-attention for sm_100a (Models/modules.py:7-21 inside :43-70).
//
// One CTA = one (utterance, head, 128-query tile).  Per 64-key tile j:
//   TMA (a SWIZZLE_128B box of 64 head-dims + a SWIZZLE_64B box of 32, x 64 rows) -> smem K (2 stages), V (2 stages)
//   tcgen05.mma  S_j = Q K_j^T        (M128 N64 K96, both operands K-major)         -> TMEM S buffer j & 1
//   softmax warps: S_j TMEM -> 64 registers per thread (thread == query row), buffer released at once; scale,
//                  key-padding mask (-1e4 like the reference), p = exp2(s - m) with a lazily moved reference
//                  point m (O and l are rescaled only when the row maximum grew by more than 2^8),
//                  P (bf16) -> smem in the SWIZZLE_128B K-major layout
//   tcgen05.mma  O += P_j V_j         (M128 N96 K64, V is the MN-major B operand)   -> TMEM cols [128,224)
// S is issued up to two tiles ahead, K/V loads two tiles ahead: no TMA round trip and no
// TMEM -> register copy sits on the per-tile critical path.  O never leaves TMEM until the final O / l.
// Two CTAs fit per SM (101 KB smem, 256 TMEM columns each) so one CTA's softmax overlaps the other's MMAs and
// prologue/epilogue.  Fully padded key tiles are skipped: with the reference's -1e4 fill their probabilities
// underflow to exactly 0 in fp32 whenever the utterance has one valid key.  Key tiles inside the all-valid
// prefix of the mask (kv_info[b][1]) never touch the mask.
#include "common.cuh"
#include "../../include/ttsb200.h"

namespace ttsb {

constexpr int kDh = 96;
constexpr int kBQ = 128;
constexpr int kBKV = 64;                       // keys per tile
// Softmax column halves (threads per query row): 2 = eight softmax warps per CTA.  -DTTSB_ATT_HALVES=1 builds the
// one-thread-per-row kernel.
#ifndef TTSB_ATT_HALVES
#define TTSB_ATT_HALVES 1
#endif
constexpr int kAttHalves = TTSB_ATT_HALVES;
constexpr int kAttThreads = 64 + 128 * kAttHalves;
// A head's 96 dims are fetched as one 64-dim box (128-byte rows, SWIZZLE_128B) and one 32-dim box (64-byte rows,
// SWIZZLE_64B): the TMA engine is paced by 128-byte lines, and three 64-byte boxes per row made the kernel
// TMA-bound (19 B/clk/SM, 176 us with the softmax arithmetic removed).
constexpr int kQBox0 = 128 * kBQ;              // 16 KB
constexpr int kQBox1 = 64 * kBQ;               // 8 KB
constexpr int kQBytes = kQBox0 + kQBox1;       // 24 KB
constexpr int kKVBox0 = 128 * kBKV;            // 8 KB
constexpr int kKVBox1 = 64 * kBKV;             // 4 KB
constexpr int kKVBytes = kKVBox0 + kKVBox1;    // 12 KB
constexpr int kKStages = 2;
constexpr int kVStages = 2;
constexpr int kAttSmem = kQBytes + (kKStages + kVStages) * kKVBytes + 1024 /*align*/ + 256 /*barriers*/ +
                         256 /*flags*/ + 6 * kBQ * 4 /*row max / row sum exchange between the column halves*/;
constexpr int kTmemQ = 224;                    // Q dims [0,64) as packed bf16 pairs: columns [224,256)
constexpr int kAttTmemCols = 256;              // S0 [0,64) S1 [64,128) O [128,224); P_j (bf16 pairs) overwrites the
                                               // first 32 columns of its own S buffer once S_j sits in registers
constexpr float kLog2e = 1.4426950408889634f;

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

#if defined(TTSB_ATT_TRACE)
__device__ long long g_att_trace[8192];
#define ATT_TR(slot) do { if (tr_on) g_att_trace[tr_n++] = ((long long)(slot) << 56) | (clock64() & 0xFFFFFFFFFFFFFFLL); } while (0)
#else
#define ATT_TR(slot) do { } while (0)
#endif

// 2^x for x <= ~9 on the FMA/ALU pipes: x = n + f, n = round(x), f in [-0.5, 0.5]; 2^f by a degree-4 polynomial,
// 2^n by adding n to the exponent field.  x is clamped at -126 (the result then flushes towards 2^-126 ~ 0).
#ifndef TTSB_POLY
#define TTSB_POLY 0
#endif
constexpr int kPolyOf4 = TTSB_POLY;  // elements per group of 4 evaluated by the polynomial instead of MUFU
__device__ __forceinline__ float poly_exp2(float x) {
  x = fmaxf(x, -126.f);
  const float t = x + 12582912.f;          // 1.5 * 2^23: the low mantissa bits of t hold round(x)
  const float f = x - (t - 12582912.f);
  float p = fmaf(f, 0.0096181291f, 0.0555041087f);
  p = fmaf(p, f, 0.2402265070f);
  p = fmaf(p, f, 0.6931471806f);
  p = fmaf(p, f, 1.0f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}

// Packed variant (TTSB_POLY2 = pairs per group of 8 elements that take the FMA pipe instead of MUFU.EX2): the two
// softmax warps that share an SM sub-partition contend for its 4-lane MUFU (a warp-wide EX2 holds it for 8 clk),
// which is what the exponential phase of a tile is bound by; FFMA2/FADD2 evaluate two elements per issue slot.
#ifndef TTSB_POLY2
#define TTSB_POLY2 0
#endif
constexpr int kPoly2 = TTSB_POLY2;
__device__ __forceinline__ void poly_exp2x2(float& x, float& y) {
  x = fmaxf(x, -126.f);
  y = fmaxf(y, -126.f);
  float tx = x, ty = y;
  add2(tx, ty, 12582912.f, 12582912.f);            // t = x + 1.5 * 2^23: low mantissa bits hold round(x)
  float rx = tx, ry = ty;
  add2(rx, ry, -12582912.f, -12582912.f);          // r = round(x)
  fma2(rx, ry, -1.f, -1.f, x, y);                  // f = x - r in [-0.5, 0.5]
  float px = rx, py = ry;
  fma2(px, py, 0.0096181291f, 0.0096181291f, 0.0555041087f, 0.0555041087f);
  fma2(px, py, rx, ry, 0.2402265070f, 0.2402265070f);
  fma2(px, py, rx, ry, 0.6931471806f, 0.6931471806f);
  fma2(px, py, rx, ry, 1.0f, 1.0f);
  x = __int_as_float(__float_as_int(px) + (__float_as_int(tx) << 23));
  y = __int_as_float(__float_as_int(py) + (__float_as_int(ty) << 23));
}

constexpr int kSk0 = 4, kSk1 = 2;   // k-steps of S: dims [0,64) with Q from TMEM, dims [64,96) with Q from smem
constexpr int kPVk = kBKV / 16;     // k-steps of P V

__device__ __forceinline__ void att_tma_store_3d(const CUtensorMap* m, uint32_t src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(m)), "r"(src), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}

enum {
  BAR_Q = 0,
  BAR_KF,                       // [kKStages] K stage filled
  BAR_KE = BAR_KF + kKStages,   // [kKStages] K stage consumed
  BAR_VF = BAR_KE + kKStages,   // [2]
  BAR_VE = BAR_VF + kVStages,   // [2]
  BAR_SF = BAR_VE + kVStages,   // [2] S buffer written by the tensor core
  BAR_PF = BAR_SF + 2,          // [2] P_j written over S buffer j & 1
  BAR_OF = BAR_PF + 2,          // [2] P_j V_j retired (S/P buffer j & 1 reusable, O current)
  BAR_QT = BAR_OF + 2,          // Q dims [0,64) copied into TMEM by the softmax threads
  BAR_COUNT
};

// 64-key tiles, K and V double-buffered in smem, S double-buffered in TMEM, so that neither a TMA
// round trip nor the S -> registers copy sits on the per-tile critical path (v2 had single K/V buffers: its
// per-tile time was 181 us of 226 us with the softmax arithmetic removed).
// DROP: training with dropout on the probabilities (Models/modules.py:18-19): the row sum l is taken BEFORE the mask
// (softmax, then dropout), P * m is what multiplies V.  Its own instantiation: the plain kernel keeps its registers.
template <int NH, bool DROP = false, bool CAUSAL = false>
__global__ void __launch_bounds__(64 + 128 * NH, 2)
attention_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmQ32,
                     const __grid_constant__ CUtensorMap tmKV, const __grid_constant__ CUtensorMap tmKV32,
                     const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmO32,
                     const uint8_t* __restrict__ mask, const int32_t* __restrict__ kv_info,
                     float* __restrict__ lse2, int T, int H, float scale_log2, int skip_padded_q,
                     const DropCfg drop = DropCfg()) {
  extern __shared__ uint8_t smem_raw[];
#if defined(TTSB_ATT_TRACE)
  const long long t_entry = clock64();
#endif
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + kQBytes;                   // [kKStages] x 12 KB (each 1024-aligned: 12 KB = 12 * 1024)
  uint8_t* sV = sK + kKStages * kKVBytes;       // [2] x 12 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + kVStages * kKVBytes);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + BAR_COUNT);
  uint8_t* sflag = reinterpret_cast<uint8_t*>(tmem_slot + 4);  // [2][64] key flags of a tile outside the clean prefix
  float* sxch = reinterpret_cast<float*>(sflag + 240);         // [2 S buffers + final][2 halves][128 rows]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * kBQ;
  const int h = blockIdx.y;
  const int b = blockIdx.z;
  const int D = H * kDh;
  const int klen = kv_info[2 * b];          // 1 + index of the last valid key
  const int kprefix = kv_info[2 * b + 1];   // keys [0, kprefix) are all valid
  const int kv_eff = klen > 0 ? klen : T;   // no valid key: the reference degenerates to uniform over T
  // CAUSAL (its own instantiation: the plain kernel keeps its registers and its 127 us): key j attends only for
  // j <= query index, on top of the key-padding mask (the target mask of Transformer-TTS, train.py:26-58) — key tiles
  // wholly above the tile's last query are never loaded, tiles that cross the diagonal take the flagged path below
  const int nkv = CAUSAL ? min((kv_eff + kBKV - 1) / kBKV, (q0 + kBQ - 1) / kBKV + 1) : (kv_eff + kBKV - 1) / kBKV;
  // ragged batches: a query tile wholly beyond the utterance's last valid row is padding (CTA-uniform exit, before
  // any barrier or TMEM allocation); its ctx rows stay unwritten
  if (skip_padded_q && klen > 0 && q0 >= klen) return;

  auto load_k = [&](int j) {
    const int ks = j % kKStages;
    const uint32_t kf = smem_u32(&bars[BAR_KF + ks]);
    mbar_arrive_expect_tx(kf, kKVBytes);
    tma_load_3d(smem_u32(sK + ks * kKVBytes), &tmKV, kf, D + h * kDh, j * kBKV, b);
    tma_load_3d(smem_u32(sK + ks * kKVBytes + kKVBox0), &tmKV32, kf, D + h * kDh + 64, j * kBKV, b);
  };
  auto load_v = [&](int j) {
    const int vs = j % kVStages;
    const uint32_t vf = smem_u32(&bars[BAR_VF + vs]);
    mbar_arrive_expect_tx(vf, kKVBytes);
    tma_load_3d(smem_u32(sV + vs * kKVBytes), &tmKV, vf, 2 * D + h * kDh, j * kBKV, b);
    tma_load_3d(smem_u32(sV + vs * kKVBytes + kKVBox0), &tmKV32, vf, 2 * D + h * kDh + 64, j * kBKV, b);
  };
  // Prologue.  Thread 0 initialises the barriers and then issues ten TMA loads (~150 clk each for the issuing
  // thread); the other warps only need the barriers and the TMEM base, so warp 0 ARRIVES on the start barrier and
  // the rest SYNC on it: nobody waits for the load issue (this was ~1500 clk of every CTA's 3900-clk prologue).
  if (warp == 0) {
    if (lane == 0) {
      tma_prefetch_desc(&tmQ);
      tma_prefetch_desc(&tmQ32);
      tma_prefetch_desc(&tmKV);
      tma_prefetch_desc(&tmKV32);
      for (int i = 0; i < BAR_COUNT; ++i)
        mbar_init(smem_u32(&bars[i]), (i == BAR_PF || i == BAR_PF + 1 || i == BAR_QT) ? 128 * NH : 1);
      fence_mbar_init();
    }
    __syncwarp();
    asm volatile("bar.arrive 2, %0;" ::"n"(64 + 128 * NH) : "memory");
    if (lane == 0) {
      const uint32_t qb = smem_u32(&bars[BAR_Q]);
      mbar_arrive_expect_tx(qb, kQBytes);
      tma_load_3d(smem_u32(sQ), &tmQ, qb, h * kDh, q0, b);
      tma_load_3d(smem_u32(sQ + kQBox0), &tmQ32, qb, h * kDh + 64, q0, b);
      load_k(0);
      load_v(0);
      if (nkv > 1) {
        load_k(1);
        load_v(1);
      }
      tma_prefetch_desc(&tmO);
      tma_prefetch_desc(&tmO32);
    }
  } else {
    if (warp == 1) {
      tmem_alloc(smem_u32(tmem_slot), kAttTmemCols);
      tmem_relinquish();
    }
    tc_fence_before();
    asm volatile("bar.sync 2, %0;" ::"n"(64 + 128 * NH) : "memory");
    tc_fence_after();
  }
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_O = tmem_base + 128;

  if (warp == 0) {
    if (lane == 0) {
      // V_i reuses the stage of V_{i-2} once P V_{i-2} has retired; K_{i+2} the stage of K_i once S_i has (S_i is
      // issued right behind P V_{i-2}), so both loads run two tiles ahead of their consumers
      for (int i = 0; i < nkv; ++i) {
        if (i >= 2) {
          mbar_wait(smem_u32(&bars[BAR_VE + (i & 1)]), ((i >> 1) - 1) & 1);
          load_v(i);
        }
        if (i + 2 < nkv) {
          mbar_wait(smem_u32(&bars[BAR_KE + (i & 1)]), (i >> 1) & 1);
          load_k(i + 2);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // The WHOLE warp runs this loop converged (waits included) and only the tcgen05 instructions are predicated on
    // one elected lane: every operand is then warp-uniform and lives in uniform registers.  Under `if (lane == 0)`
    // the compiler wraps each UTCHMMA in an ELECT / R2UR.BROADCAST / BRA.U.ANY waterfall loop (~100 clk per MMA),
    // which made this single thread the bottleneck of the kernel (2500 clk per 64-key tile).
    constexpr uint32_t idesc_s = umma_idesc_bf16(kBQ, kBKV, 0, 0);
    constexpr uint32_t idesc_o64 = umma_idesc_bf16(kBQ, 64, 0, 1);  // B = V is MN-major: dims [0,64) ...
    constexpr uint32_t idesc_o32 = umma_idesc_bf16(kBQ, 32, 0, 1);  // ... and dims [64,96)
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t q_base = smem_u32(sQ);
    const uint32_t leader = elect_one();
#if defined(TTSB_ATT_TRACE)
    const bool tr_on = TTSB_ATT_TRACE == 2 && blockIdx.x == 4 && blockIdx.y == 1 && blockIdx.z == 10 && lane == 0;
    int tr_n = 0;
#endif
    // S_j = Q K_j^T goes to S buffer j & 1 and is issued up to two tiles ahead of O += P_j V_j.
    // S buffer j & 1 last held P_{j-2}: P_{j-2} V_{j-2} was issued by this warp before S_j and the tensor pipe
    // executes one thread's MMAs in issue order, so the overwrite needs no barrier.
    auto s_mmas = [&](int j) {  // leader lane only
      const int ks = j % kKStages;
      const uint32_t k_base = smem_u32(sK + ks * kKVBytes);
      const uint32_t d_tmem = tb + (j & 1) * kBKV;
#pragma unroll
      for (int kk = 0; kk < kSk0; ++kk)  // dims [0,64): Q from TMEM (columns [224,256), 8 packed columns per k-step)
        umma_bf16_ts(d_tmem, tb + kTmemQ + kk * 8, umma_smem_desc(k_base + kk * 32, 16, 1024, kLayoutSW128), idesc_s,
                     kk != 0);
#pragma unroll
      for (int kk = 0; kk < kSk1; ++kk)  // dims [64,96): 64-byte rows
        umma_bf16(d_tmem, umma_smem_desc(q_base + kQBox0 + kk * 32, 16, 512, kLayoutSW64),
                  umma_smem_desc(k_base + kKVBox0 + kk * 32, 16, 512, kLayoutSW64), idesc_s, 1u);
      umma_commit(smem_u32(&bars[BAR_KE + ks]));
      umma_commit(smem_u32(&bars[BAR_SF + (j & 1)]));
    };
    auto issue_s = [&](int j) {
      mbar_wait(smem_u32(&bars[BAR_KF + j % kKStages]), (j / kKStages) & 1);
      tc_fence_after();
      if (leader) s_mmas(j);
      __syncwarp();
    };
    mbar_wait(smem_u32(&bars[BAR_Q]), 0);
    mbar_wait(smem_u32(&bars[BAR_QT]), 0);
    tc_fence_after();
    issue_s(0);
    if (nkv > 1) issue_s(1);
    for (int j = 0; j < nkv; ++j) {
      const int vs = j % kVStages;
      const bool more = j + 2 < nkv;
      ATT_TR(10);
      // ONE wait for everything this iteration consumes — P_j (softmax), V_j and K_{j+2} (TMA, both long landed):
      // a poll of a completed barrier still costs the issuer ~150 clk, and this warp's round trip (P_j ready ->
      // S_{j+2} ready) is half of the tile period, so three polls in a row were ~20 % of it.
      const uint32_t vf = smem_u32(&bars[BAR_VF + vs]), vpar = (j / kVStages) & 1;
      mbar_wait3(smem_u32(&bars[BAR_PF + (j & 1)]), (j >> 1) & 1, vf, vpar,
                 more ? smem_u32(&bars[BAR_KF + (j + 2) % kKStages]) : vf, more ? ((j + 2) / kKStages) & 1 : vpar);
      ATT_TR(12);
      tc_fence_after();
      const uint32_t v_base = smem_u32(sV + vs * kKVBytes);
      if (leader) {
#pragma unroll
        for (int kk = 0; kk < kPVk; ++kk) {
          // A = P read from TMEM (16 keys = 8 packed columns per MMA), never through shared memory;
          // B = V: 16 rows of 128 B (box 0) / of 64 B (box 1); SBO = 8 rows
          const uint32_t a_tmem = tb + (j & 1) * kBKV + kk * 8;
          umma_bf16_ts(tb + 128, a_tmem, umma_smem_desc(v_base + kk * 2048, kKVBox0, 1024, kLayoutSW128), idesc_o64,
                       (j | kk) != 0);
          umma_bf16_ts(tb + 128 + 64, a_tmem, umma_smem_desc(v_base + kKVBox0 + kk * 1024, kKVBox1, 512, kLayoutSW64),
                       idesc_o32, (j | kk) != 0);
        }
        umma_commit(smem_u32(&bars[BAR_VE + vs]));
        umma_commit(smem_u32(&bars[BAR_OF + (j & 1)]));
        if (more) s_mmas(j + 2);
      }
      __syncwarp();
      ATT_TR(13);
    }
  } else {
    // ===================== softmax / output warps =====================
    // NH = 1: thread == query row.  NH = 2: two threads per row — warps 2..5 take key columns [0,32) of every tile,
    // warps 6..9 columns [32,64) (a warp reaches the TMEM lane quarter warp % 4, so warps w and w + 4 share rows).
    // The tile period is a round trip (softmax chain + issuer chain) / 2, and with one softmax warp per CTA and
    // sub-partition the chain is latency-bound; four shorter chains per sub-partition keep its MUFU busy.
    constexpr int CW = kBKV / NH;     // S columns per thread
    constexpr int OW = kDh / NH;      // O columns per thread in the rescale / output passes
    const int q = warp & 3;
    const int half = NH == 1 ? 0 : (warp - 2) >> 2;
    const int row = q * 32 + lane;
    const int st = threadIdx.x - 64;  // 0 .. 128 * NH - 1 among softmax threads
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    const float masked = -1e4f * kLog2e;  // reference fill value (modules.py:14), in log2 units
    // m is the reference point of the exponentials, not necessarily the running maximum: it is only
    // moved (and O, l rescaled) when the maximum has grown by more than 2^8, so p <= 256 stays well
    // inside bf16/fp32 range and the final O / l is unchanged.  With NH = 2 both threads of a row hold the same m
    // and a partial l each.
    float m = -INFINITY, l = 0.f;
    // Q dims [0,64) -> TMEM once per CTA: the S MMAs then read their A operand from TMEM and only K from shared
    // memory (re-reading Q for every key tile was 16 KB of the 72 KB of shared-memory traffic per tile).
#if defined(TTSB_ATT_TRACE)
    long long t_sync, t_q;
#endif
    {
#if defined(TTSB_ATT_TRACE)
      t_sync = clock64();
#endif
      mbar_wait(smem_u32(&bars[BAR_Q]), 0);
#if defined(TTSB_ATT_TRACE)
      t_q = clock64();
#endif
      const uint8_t* qrow = sQ + row * 128;  // SWIZZLE_128B box: 16-byte chunk c of row r sits at (c ^ (r & 7)) * 16
      uint32_t qv[32 / NH];
#pragma unroll
      for (int c = 0; c < 8 / NH; ++c) {
        const int cc = c + half * (8 / NH);
        const uint4 u = *reinterpret_cast<const uint4*>(qrow + ((cc ^ (row & 7)) << 4));
        qv[4 * c] = u.x; qv[4 * c + 1] = u.y; qv[4 * c + 2] = u.z; qv[4 * c + 3] = u.w;
      }
      if constexpr (NH == 1) tmem_st32(tmem_base + kTmemQ + lane_off, reinterpret_cast<const float*>(qv));
      else tmem_st16(tmem_base + kTmemQ + lane_off + half * 16, qv);
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(smem_u32(&bars[BAR_QT]));
    }
#if defined(TTSB_ATT_TRACE)
    const bool tr_on = TTSB_ATT_TRACE == 1 && blockIdx.x == 4 && blockIdx.y == 1 && blockIdx.z == 10 && threadIdx.x == 64;
    int tr_n = 0;
    if (tr_on) {
      g_att_trace[tr_n++] = (9LL << 56) | (t_entry & 0xFFFFFFFFFFFFFFLL);
      g_att_trace[tr_n++] = (20LL << 56) | (t_sync & 0xFFFFFFFFFFFFFFLL);
      g_att_trace[tr_n++] = (21LL << 56) | (t_q & 0xFFFFFFFFFFFFFFLL);
    }
#endif
    ATT_TR(0);

    for (int j = 0; j < nkv; ++j) {
      const uint32_t sb = j & 1;
      // a tile inside the all-valid prefix of the mask needs no flags (CTA-uniform decision)
      const bool clean = CAUSAL ? ((j + 1) * kBKV <= kprefix && (j + 1) * kBKV <= q0 + 1) : ((j + 1) * kBKV <= kprefix);
      uint8_t* flg = sflag + sb * kBKV;
      if (!clean) {
        if (st < kBKV) {
          const int key = j * kBKV + st;
          flg[st] = key < T ? (mask[static_cast<int64_t>(b) * T + key] ? 1 : 0) : 2;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(128 * NH) : "memory");
      }
      ATT_TR(1);
      mbar_wait(smem_u32(&bars[BAR_SF + sb]), (j >> 1) & 1);
      tc_fence_after();
      ATT_TR(2);
      float s[CW];
#pragma unroll
      for (int c = 0; c < CW; c += 32) tmem_ld32(tmem_base + sb * kBKV + lane_off + half * CW + c, s + c);
      tmem_ld_wait();
      ATT_TR(3);
      // scores in log2 units, masked like the reference (-1e4 fill; keys beyond T do not exist)
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
      if (clean) {
#pragma unroll
        for (int i = 0; i < CW; i += 8) {  // three-input max: half the issue slots
          mx4[0] = max3(mx4[0], s[i], s[i + 4]); mx4[1] = max3(mx4[1], s[i + 1], s[i + 5]);
          mx4[2] = max3(mx4[2], s[i + 2], s[i + 6]); mx4[3] = max3(mx4[3], s[i + 3], s[i + 7]);
        }
      } else {
#pragma unroll
        for (int i = 0; i < CW; ++i) {
          const uint8_t f = flg[half * CW + i];
          if constexpr (CAUSAL) {
            // a key beyond this row's own index is filled like a padded one, as the reference's combined mask does
            const bool hidden = f == 0 || (f == 1 && j * kBKV + half * CW + i > q0 + row);
            s[i] = hidden ? masked : (f == 1 ? s[i] * scale_log2 : -INFINITY);
          } else {
            s[i] = f == 1 ? s[i] * scale_log2 : (f == 0 ? masked : -INFINITY);
          }
          mx4[i & 3] = fmaxf(mx4[i & 3], s[i]);
        }
      }
      float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
      if constexpr (NH == 2) {
        // the row maximum over both column halves: exchanged through shared memory, one 64-thread named barrier per
        // lane quarter (the two warps that share these rows)
        sxch[(sb * 2 + half) * kBQ + row] = mx;
        asm volatile("bar.sync %0, 64;" ::"r"(3 + q) : "memory");
        mx = fmaxf(mx, sxch[(sb * 2 + (half ^ 1)) * kBQ + row]);
      }
      if (clean) mx *= scale_log2;  // the positive scale commutes with max
      ATT_TR(4);
      ATT_TR(5);
      const bool grow = mx > m + 8.f;  // also true on the first tile (m = -inf); identical in both halves
      if (__any_sync(0xffffffffu, grow)) {
        // the (rare) rescale of O needs P_{j-1} V_{j-1} retired
        if (j > 0) {
          mbar_wait(smem_u32(&bars[BAR_OF + (sb ^ 1)]), ((j - 1) >> 1) & 1);
          tc_fence_after();
        }
        const float m_new = grow ? mx : m;  // rows that did not grow keep their reference point (alpha = 1)
        const float alpha = fast_exp2(m - m_new);  // m = -inf -> 0
        l *= alpha;
        m = m_new;
        if (j > 0) {
#pragma unroll
          for (int c = 0; c < OW; c += 16) {
            float o[16];
            tmem_ld16(tmem_O + lane_off + half * OW + c, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) o[i] *= alpha;
            tmem_st16(tmem_O + lane_off + half * OW + c, reinterpret_cast<const uint32_t*>(o));
          }
          tmem_st_wait();
        }
      }
      // p = exp2(s - m) -> bf16; fp32 row sum (four independent partial sums)
      float ls4[4] = {0.f, 0.f, 0.f, 0.f};
      const float neg_m = -m;
      const float sc = clean ? scale_log2 : 1.f;  // masked tiles were scaled while masking
#if defined(TTSB_ATT_EXP) && TTSB_ATT_EXP >= 2
      ls4[0] = s[0] + s[CW / 2] + s[CW - 1] + neg_m * sc;  // timing experiment only: no exponentials, no P stores
      for (int c = 0; c < 0; c += 32) {
#else
#pragma unroll
      for (int c = 0; c < CW; c += 32) {
#endif
        float* v = s + c;
#if defined(TTSB_ATT_EXP) || TTSB_POLY
#pragma unroll
        for (int i = 0; i < 32; ++i) {
#if defined(TTSB_ATT_EXP) && TTSB_ATT_EXP >= 1
          v[i] = fmaf(v[i], sc, neg_m);  // timing experiment only: no exponential
#else
          // experiment (TTSB_POLY): part of the elements on the FMA pipe (round-to-nearest split + degree-4
          // polynomial) instead of MUFU.EX2 — measured slower
          if ((i & 3) < kPolyOf4) v[i] = poly_exp2(fmaf(v[i], sc, neg_m));
          else v[i] = fast_exp2(fmaf(v[i], sc, neg_m));
#endif
          ls4[i & 3] += v[i];
        }
#else
        // packed f32x2 scale-subtract and row-sum (FFMA2 / FADD2); MUFU.EX2 (16/clk/SM) is the only per-element
        // instruction that cannot be paired
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
#pragma unroll
          for (int p = 0; p < 4; ++p) fma2(v[i + 2 * p], v[i + 2 * p + 1], sc, sc, neg_m, neg_m);
#pragma unroll
          for (int p = 0; p < 4; ++p) {
            if (p < kPoly2) {
              poly_exp2x2(v[i + 2 * p], v[i + 2 * p + 1]);
            } else {
              v[i + 2 * p] = fast_exp2(v[i + 2 * p]);
              v[i + 2 * p + 1] = fast_exp2(v[i + 2 * p + 1]);
            }
          }
          add2(ls4[0], ls4[1], v[i], v[i + 1]);
          add2(ls4[2], ls4[3], v[i + 2], v[i + 3]);
          add2(ls4[0], ls4[1], v[i + 4], v[i + 5]);
          add2(ls4[2], ls4[3], v[i + 6], v[i + 7]);
        }
#endif
      }
      if constexpr (DROP) {
        const uint64_t e0 = ((static_cast<uint64_t>(b) * H + h) * T + static_cast<uint64_t>(q0 + row < T ? q0 + row : 0)) * T +
                            j * kBKV + half * CW;
#pragma unroll
        for (int i = 0; i < CW; ++i) s[i] *= drop_scale(drop, e0 + i);
      }
      // P (bf16 pairs, key 2c in the low half of column c) -> this thread's TMEM lane: the A operand of P V
      {
        uint32_t pk[CW / 2];
#pragma unroll
        for (int i = 0; i < CW / 2; ++i) pk[i] = pack_bf16(s[2 * i], s[2 * i + 1]);
        if constexpr (NH == 1) tmem_st32(tmem_base + sb * kBKV + lane_off, reinterpret_cast<const float*>(pk));
        else tmem_st16(tmem_base + sb * kBKV + lane_off + half * 16, pk);
        tmem_st_wait();
      }
      l += (ls4[0] + ls4[1]) + (ls4[2] + ls4[3]);
      ATT_TR(6);
      tc_fence_before();   // orders the P store and the O rescale (tcgen05.st) before the MMA
      mbar_arrive(smem_u32(&bars[BAR_PF + sb]));
      ATT_TR(7);
    }
    // ---- O / l ----
    if constexpr (NH == 2) {  // the row sum over both column halves (its own exchange slots: no reuse hazard)
      sxch[(4 + half) * kBQ + row] = l;
      asm volatile("bar.sync %0, 64;" ::"r"(3 + q) : "memory");
      l += sxch[(4 + (half ^ 1)) * kBQ + row];
    }
    mbar_wait(smem_u32(&bars[BAR_OF + ((nkv - 1) & 1)]), ((nkv - 1) >> 1) & 1);
    tc_fence_after();
    const int t = q0 + row;
    const float inv = 1.f / l;
    // training: log2-domain log-sum-exp of the row, p = exp2(s*scale_log2 - lse2) in the backward kernels
    if (lse2 != nullptr && t < T && half == 0) lse2[(static_cast<int64_t>(b) * H + h) * T + t] = m + log2f(l);
    // O / l -> bf16 -> the (dead) Q staging buffer in the layout of the two TMA boxes (128-byte rows SWIZZLE_128B for
    // dims [0,64), 64-byte rows SWIZZLE_64B for [64,96)), then two TMA stores per CTA.  Row-per-thread 16-byte global
    // stores cost one LSU wavefront per row and instruction (~2000 clk per CTA, ncu: long_sb on the next TMEM load);
    // the TMA store also clips the rows beyond T.  Every S MMA (the last reader of the Q buffer) retired before
    // BAR_OF of the last tile completed.
    uint8_t* orow0 = sQ + row * 128;
    uint8_t* orow1 = sQ + kQBox0 + row * 64;
#pragma unroll
    for (int c = 0; c < OW; c += 16) {
      float o[16];
      tmem_ld16(tmem_O + lane_off + half * OW + c, o);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 16; i += 8) {
        uint4 u;
        u.x = pack_bf16(o[i] * inv, o[i + 1] * inv);
        u.y = pack_bf16(o[i + 2] * inv, o[i + 3] * inv);
        u.z = pack_bf16(o[i + 4] * inv, o[i + 5] * inv);
        u.w = pack_bf16(o[i + 6] * inv, o[i + 7] * inv);
        const int g = (half * OW + c + i) >> 3;  // 16-byte chunk of the row
        if (g < 8) *reinterpret_cast<uint4*>(orow0 + ((g ^ (row & 7)) << 4)) = u;
        else *reinterpret_cast<uint4*>(orow1 + (((g - 8) ^ ((row >> 1) & 3)) << 4)) = u;
      }
    }
    fence_async_smem();  // generic-proxy writes -> visible to the TMA engine
    asm volatile("bar.sync 1, %0;" ::"n"(128 * NH) : "memory");
    if (st == 0) {
      att_tma_store_3d(&tmO, smem_u32(sQ), h * kDh, q0, b);
      att_tma_store_3d(&tmO32, smem_u32(sQ + kQBox0), h * kDh + 64, q0, b);
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // smem must outlive the reads of the stores
    }
    tc_fence_before();
    ATT_TR(8);
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

static int attention_fwd_launch(const void* qkv, const uint8_t* mask, const int32_t* kv_info, void* ctx, float* lse2,
                                int32_t B, int32_t T, int32_t H, int32_t Dh, int skip_padded_q, ttsb200_stream_t s,
                                const DropCfg drop = DropCfg()) {
  TTSB_CHECK_ARG(qkv && mask && kv_info && ctx, "attention_fwd: null pointer");
  TTSB_CHECK_ARG(B > 0 && T > 0 && H > 0, "attention_fwd: bad dims");
  TTSB_CHECK_ARG(Dh == kDh, "attention_fwd: head dim %d unsupported (this build: %d)", Dh, kDh);
  TTSB_CHECK_ARG((reinterpret_cast<uintptr_t>(qkv) & 15) == 0 && (reinterpret_cast<uintptr_t>(ctx) & 15) == 0,
                 "attention_fwd: qkv/ctx must be 16-byte aligned");
  CUtensorMap tmQ, tmQ32, tmKV, tmKV32;
  const uint64_t ld = 3ull * H * Dh;
  const uint64_t dims[3] = {ld, static_cast<uint64_t>(T), static_cast<uint64_t>(B)};
  const uint64_t strides[2] = {ld * 2, ld * 2 * static_cast<uint64_t>(T)};
  const uint32_t box_q[3] = {64, kBQ, 1}, box_q32[3] = {32, kBQ, 1};
  const uint32_t box_kv[3] = {64, kBKV, 1}, box_kv32[3] = {32, kBKV, 1};
  int rc = make_tmap_bf16(&tmQ, qkv, 3, dims, strides, box_q, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  rc = make_tmap_bf16(&tmQ32, qkv, 3, dims, strides, box_q32, CU_TENSOR_MAP_SWIZZLE_64B);
  if (rc) return rc;
  rc = make_tmap_bf16(&tmKV, qkv, 3, dims, strides, box_kv, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  rc = make_tmap_bf16(&tmKV32, qkv, 3, dims, strides, box_kv32, CU_TENSOR_MAP_SWIZZLE_64B);
  if (rc) return rc;
  CUtensorMap tmO, tmO32;  // ctx (B, T, H*Dh): the store boxes have the shape of the Q boxes
  const uint64_t odims[3] = {static_cast<uint64_t>(H) * Dh, static_cast<uint64_t>(T), static_cast<uint64_t>(B)};
  const uint64_t ostrides[2] = {odims[0] * 2, odims[0] * 2 * static_cast<uint64_t>(T)};
  rc = make_tmap_bf16(&tmO, ctx, 3, odims, ostrides, box_q, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  rc = make_tmap_bf16(&tmO32, ctx, 3, odims, ostrides, box_q32, CU_TENSOR_MAP_SWIZZLE_64B);
  if (rc) return rc;
  dim3 grid((T + kBQ - 1) / kBQ, H, B);
  const float scale_log2 = kLog2e / sqrtf(static_cast<float>(Dh));
  if ((skip_padded_q & 2) && drop.thresh) {   // causal training with dropout on the probabilities (train.py, dropout 0.1)
    auto kern = attention_fwd_kernel<kAttHalves, true, true>;
    TTSB_SMEM_OPT_IN(kern, kAttSmem);
    kern<<<grid, kAttThreads, kAttSmem, static_cast<cudaStream_t>(s)>>>(tmQ, tmQ32, tmKV, tmKV32, tmO, tmO32, mask, kv_info,
                                                                       lse2, T, H, scale_log2, skip_padded_q & 1, drop);
  } else if (skip_padded_q & 2) {   // flag word: bit 0 = skip padded query tiles, bit 1 = causal
    auto kern = attention_fwd_kernel<kAttHalves, false, true>;
    TTSB_SMEM_OPT_IN(kern, kAttSmem);
    kern<<<grid, kAttThreads, kAttSmem, static_cast<cudaStream_t>(s)>>>(tmQ, tmQ32, tmKV, tmKV32, tmO, tmO32, mask, kv_info,
                                                                       lse2, T, H, scale_log2, skip_padded_q & 1, drop);
  } else if (drop.thresh) {
    auto kern = attention_fwd_kernel<kAttHalves, true>;
    TTSB_SMEM_OPT_IN(kern, kAttSmem);
    kern<<<grid, kAttThreads, kAttSmem, static_cast<cudaStream_t>(s)>>>(tmQ, tmQ32, tmKV, tmKV32, tmO, tmO32, mask, kv_info,
                                                                       lse2, T, H, scale_log2, skip_padded_q, drop);
  } else {
    auto kern = attention_fwd_kernel<kAttHalves, false>;
    TTSB_SMEM_OPT_IN(kern, kAttSmem);
    kern<<<grid, kAttThreads, kAttSmem, static_cast<cudaStream_t>(s)>>>(tmQ, tmQ32, tmKV, tmKV32, tmO, tmO32, mask, kv_info,
                                                                       lse2, T, H, scale_log2, skip_padded_q, drop);
  }
  TTSB_CHECK_CUDA(cudaGetLastError());
  return TTSB_OK;
}

#if defined(TTSB_ATT_TRACE)
extern "C" int ttsb200_debug_att_trace(long long* dst, int n) {
  return cudaMemcpyFromSymbol(dst, g_att_trace, n * sizeof(long long)) == cudaSuccess ? 0 : -1;
}
#endif

extern "C" int ttsb200_attention_fwd_lse(const void* qkv, const uint8_t* mask, const int32_t* kv_info, void* ctx,
                                         float* lse2, int32_t B, int32_t T, int32_t H, int32_t Dh,
                                         ttsb200_stream_t s) {
  return attention_fwd_launch(qkv, mask, kv_info, ctx, lse2, B, T, H, Dh, 0, s);
}

extern "C" int ttsb200_attention_fwd_drop(const void* qkv, const uint8_t* mask, const int32_t* kv_info, void* ctx,
                                          float* lse2, int32_t B, int32_t T, int32_t H, int32_t Dh, float drop_p,
                                          uint64_t drop_seed, int64_t drop_site, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f, "attention_fwd: dropout probability must be in [0, 1)");
  return attention_fwd_launch(qkv, mask, kv_info, ctx, lse2, B, T, H, Dh, 0, s, make_drop(drop_p, drop_seed, drop_site));
}

extern "C" int ttsb200_attention_fwd_ragged(const void* qkv, const uint8_t* mask, const int32_t* kv_info, void* ctx,
                                            int32_t B, int32_t T, int32_t H, int32_t Dh, ttsb200_stream_t s) {
  return attention_fwd_launch(qkv, mask, kv_info, ctx, nullptr, B, T, H, Dh, 1, s);
}

extern "C" int ttsb200_attention_fwd_causal(const void* qkv, const uint8_t* mask, const int32_t* kv_info, void* ctx,
                                            float* lse2, int32_t B, int32_t T, int32_t H, int32_t Dh, ttsb200_stream_t s) {
  return attention_fwd_launch(qkv, mask, kv_info, ctx, lse2, B, T, H, Dh, 2, s);
}

extern "C" int ttsb200_attention_fwd_causal_drop(const void* qkv, const uint8_t* mask, const int32_t* kv_info, void* ctx,
                                                 float* lse2, int32_t B, int32_t T, int32_t H, int32_t Dh, float drop_p,
                                                 uint64_t drop_seed, int64_t drop_site, ttsb200_stream_t s) {
  TTSB_CHECK_ARG(drop_p >= 0.f && drop_p < 1.f, "attention_fwd: dropout probability must be in [0, 1)");
  return attention_fwd_launch(qkv, mask, kv_info, ctx, lse2, B, T, H, Dh, 2, s, make_drop(drop_p, drop_seed, drop_site));
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
