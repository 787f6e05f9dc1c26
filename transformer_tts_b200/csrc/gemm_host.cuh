// Host side of conv_gemm: tensor maps, persistent-grid launch, argument validation, C entry point.
#pragma once

namespace ttsb {

static int g_num_sms = 0;

static int num_sms() {
  if (g_num_sms == 0) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
      g_num_sms = n;
    else
      g_num_sms = 148;
  }
  return g_num_sms;
}

// CTA-pair (cta_group::2) kernels: 0 = never, 1 = where they measured faster, 2 = wherever implemented.
// ttsb200_debug_gemm_pair() or the environment variable TTSB200_GEMM_PAIR (read once) override the default.
static int gemm_pair_default() {
  const char* e = getenv("TTSB200_GEMM_PAIR");
  return (e && e[0] >= '0' && e[0] <= '2') ? e[0] - '0' : 1;
}
static int g_gemm_pair = gemm_pair_default();

template <int BLOCK_N, int EPI, bool A_STAT, bool PAIR = false>
static int launch_gemm(const ttsb200_gemm_args& a, GemmDev& d, cudaStream_t stream) {
  using Cfg = GemmCfg<BLOCK_N, EPI, A_STAT, PAIR>;
  static_assert(Cfg::kSmemBytes <= kMaxSmem, "shared memory budget exceeded");
  CUtensorMap tmA, tmW, tmOut, tmOut32, tmRes32, tmResBf;
  const uint64_t uN = static_cast<uint64_t>(a.N), uT = static_cast<uint64_t>(a.T), uB = static_cast<uint64_t>(a.B);
  {
    const uint64_t dims[3] = {static_cast<uint64_t>(a.C), uT, uB};
    const uint64_t strides[2] = {static_cast<uint64_t>(a.ldx) * 2, static_cast<uint64_t>(a.ldx) * 2 * uT};
    const uint32_t box[3] = {kBlockK, kBlockM, 1};
    int rc = make_tmap_bf16(&tmA, a.x, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  }
  {
    const uint64_t dims[2] = {static_cast<uint64_t>(a.C), static_cast<uint64_t>(a.taps) * uN};
    const uint64_t strides[1] = {static_cast<uint64_t>(a.C) * 2};
    const uint32_t box[2] = {kBlockK, static_cast<uint32_t>(Cfg::kWLoadN)};
    int rc = make_tmap_bf16(&tmW, a.w, 2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  }
  tmOut = tmA;  // placeholders keep the descriptor prefetches valid when a map is unused
  if (a.out_bf16 && a.N % 64 == 0) {
    const uint64_t dims[3] = {uN, uT, uB};
    const uint64_t strides[2] = {uN * 2, uN * 2 * uT};
    const uint32_t box[3] = {32, kBlockM, 1};  // staging boxes: 128 rows x 32 columns, SWIZZLE_64B
    int rc = make_tmap_bf16(&tmOut, a.out_bf16, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_64B);
    if (rc) return rc;
  }
  tmOut32 = tmA;
  tmRes32 = tmA;
  tmResBf = tmA;
  if (Cfg::kRing) {
    const uint32_t box32[3] = {32, kBlockM, 1};
    if (a.out_f32) {
      const uint64_t dims[3] = {uN, uT, uB};
      const uint64_t strides[2] = {uN * 4, uN * 4 * uT};
      int rc = make_tmap(&tmOut32, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, a.out_f32, 3, dims, strides, box32,
                         CU_TENSOR_MAP_SWIZZLE_128B);
      if (rc) return rc;
    }
    {
      const uint64_t dims[3] = {uN, uT, d.res_bcast ? 1 : uB};
      const uint64_t strides[2] = {uN * 4, uN * 4 * uT};
      int rc = make_tmap(&tmRes32, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, a.res_f32, 3, dims, strides, box32,
                         CU_TENSOR_MAP_SWIZZLE_128B);
      if (rc) return rc;
    }
    if (EPI == TTSB200_EPI_FFN2) {
      const uint64_t dims[3] = {uN, uT, uB};
      const uint64_t strides[2] = {uN * 2, uN * 2 * uT};
      const uint32_t box[3] = {64, kBlockM, 1};
      int rc = make_tmap_bf16(&tmResBf, a.res_bf16, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
      if (rc) return rc;
    }
  }
  auto kern = conv_gemm_kernel<BLOCK_N, EPI, A_STAT, PAIR>;
  TTSB_SMEM_OPT_IN(kern, Cfg::kSmemBytes);  // per instantiation and device
  d.m_tiles = (a.T + kBlockM - 1) / kBlockM;
  d.n_tiles = a.N / BLOCK_N;
  d.num_tiles = a.B * d.m_tiles * d.n_tiles;
  if (PAIR) {
    const int mt_units = (a.B * d.m_tiles + 1) / 2;
    const int units = A_STAT ? mt_units : mt_units * d.n_tiles;
    const int pairs = units < num_sms() / 2 ? units : num_sms() / 2;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * pairs);
    cfg.blockDim = dim3(kGemmThreads);
    cfg.dynamicSmemBytes = Cfg::kSmemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    TTSB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, tmA, tmW, tmOut, tmOut32, tmRes32, tmResBf, d));
    return TTSB_OK;
  }
  const int units = A_STAT ? a.B * d.m_tiles : d.num_tiles;
  const int grid = units < num_sms() ? units : num_sms();
  kern<<<grid, kGemmThreads, Cfg::kSmemBytes, stream>>>(tmA, tmW, tmOut, tmOut32, tmRes32, tmResBf, d);
  TTSB_CHECK_CUDA(cudaGetLastError());
  return TTSB_OK;
}

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace ttsb

using namespace ttsb;

extern "C" void ttsb200_debug_gemm_pair(int on) { g_gemm_pair = on; }
// The A-stationary sweep gives one unit per 128-row tile; below this many row tiles (short batches: the encoder's
// 64 x 128 phonemes are 64 tiles for 148 SMs) the tile-parallel schedule fills the machine better.
static int g_astat_min_units = 148;  // measured: encoder q|k|v (64 tiles) 18.5 -> 12.1 us tile-parallel
extern "C" void ttsb200_debug_gemm_astat_min(int units) { g_astat_min_units = units; }

#if defined(TTSB_GEMM_TRACE)
extern "C" int ttsb200_debug_gemm_trace(long long* dst, int n) {
  return cudaMemcpyFromSymbol(dst, g_gemm_trace, n * sizeof(long long)) == cudaSuccess ? 0 : -1;
}
#endif

extern "C" int ttsb200_conv_gemm(const ttsb200_gemm_args* a, ttsb200_stream_t stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  TTSB_CHECK_ARG(a != nullptr, "conv_gemm: null args");
  TTSB_CHECK_ARG(a->B > 0 && a->T > 0 && a->N > 0 && a->C > 0 && a->taps > 0, "conv_gemm: bad dims");
  TTSB_CHECK_ARG(a->x && a->w, "conv_gemm: null x/w");
  TTSB_CHECK_ARG(a->C % 8 == 0 && a->ldx % 8 == 0 && a->ldx >= a->C,
                 "conv_gemm: C (%d) and ldx (%lld) must be multiples of 8, ldx >= C", a->C, (long long)a->ldx);
  TTSB_CHECK_ARG(aligned16(a->x) && aligned16(a->w), "conv_gemm: x/w must be 16-byte aligned");
  TTSB_CHECK_ARG(a->block_n > 0 && a->N % a->block_n == 0, "conv_gemm: N (%d) not a multiple of block_n (%d)",
                 a->N, a->block_n);
  TTSB_CHECK_ARG(a->bias != nullptr, "conv_gemm: bias required");
  if (a->out_bf16) TTSB_CHECK_ARG(aligned16(a->out_bf16), "conv_gemm: out_bf16 must be 16-byte aligned");
  if (a->out_f32) TTSB_CHECK_ARG(aligned16(a->out_f32), "conv_gemm: out_f32 must be 16-byte aligned");
  GemmDev d;
  d.B = a->B; d.T = a->T; d.N = a->N; d.C = a->C; d.taps = a->taps; d.pad_left = a->pad_left;
  d.epi = a->epi; d.act = a->act;
  d.m_tiles = d.n_tiles = d.num_tiles = 0;
  d.res_bcast = 0;
  d.bias = a->bias; d.scale = a->scale; d.res_f32 = a->res_f32;
  d.res_bstride = a->res_bstride; d.res_tstride = a->res_tstride; d.res_scale = a->res_scale;
  d.ln1_g = a->ln1_g; d.ln1_b = a->ln1_b; d.ln2_g = a->ln2_g; d.ln2_b = a->ln2_b; d.eps = a->eps;
  d.dot_w = a->dot_w; d.dot_b = a->dot_b; d.row_mask = a->row_mask;
  d.out_f32 = a->out_f32; d.out_bf16 = static_cast<__nv_bfloat16*>(a->out_bf16);
  d.out2_f32 = a->out2_f32; d.out_row = a->out_row;
  d.pre_f32 = a->pre_f32; d.gate_bf16 = static_cast<const __nv_bfloat16*>(a->gate_bf16);
  TTSB_CHECK_ARG((a->tile_map == nullptr) == (a->tile_count == nullptr), "conv_gemm: tile_map and tile_count go together");
  d.mt_map = a->tile_map; d.mt_count = a->tile_count;
  if (a->pre_f32) TTSB_CHECK_ARG(aligned16(a->pre_f32) && a->epi != TTSB200_EPI_AFFINE_ACT,
                                 "conv_gemm: pre_f32 is an output of the LayerNorm epilogues, 16-byte aligned");
  if (a->act == TTSB200_ACT_RELU_GRAD)
    TTSB_CHECK_ARG(a->gate_bf16 && aligned16(a->gate_bf16) && a->epi == TTSB200_EPI_AFFINE_ACT && a->N % 8 == 0,
                   "conv_gemm: ACT_RELU_GRAD needs gate_bf16 (B,T,N), AFFINE epilogue");

  const int e = a->epi, bn = a->block_n;
  if (e == TTSB200_EPI_AFFINE_ACT) {
    TTSB_CHECK_ARG(a->out_f32 || a->out_bf16, "conv_gemm: AFFINE_ACT needs an output");
    if (a->res_f32)
      TTSB_CHECK_ARG(a->res_tstride % 4 == 0 && a->res_bstride % 4 == 0 && aligned16(a->res_f32),
                     "conv_gemm: res strides % 4, 16-byte aligned");
    // A-stationary sweep: plain linears on <= 384 input channels with bf16-only output
    const bool a_stat = a->taps == 1 && a->C <= 384 && a->C % 64 == 0 && a->out_bf16 && !a->out_f32 &&
                        !a->res_f32 && a->N / bn >= 2 &&
                        static_cast<long long>(a->B) * ((a->T + kBlockM - 1) / kBlockM) >= g_astat_min_units;
    if (bn == 80) return launch_gemm<80, TTSB200_EPI_AFFINE_ACT, false>(*a, d, stream);
    if (bn == 128) {
      if (a_stat) return launch_gemm<128, TTSB200_EPI_AFFINE_ACT, true>(*a, d, stream);
      return launch_gemm<128, TTSB200_EPI_AFFINE_ACT, false>(*a, d, stream);
    }
    // CTA pairs halve the weight stream into each SM; worth it once there are at least two m-tiles per pair
    // (ttsb200_debug_gemm_pair: 0 = never, 1 = where it measured faster, 2 = wherever it is implemented)
    const int pair = static_cast<long long>(a->B) * ((a->T + kBlockM - 1) / kBlockM) >= 2 ? g_gemm_pair : 0;
    if (bn == 192) {  // measured: the 192-wide tiles gain nothing from pairing (q|k|v 70 us single, 75 us paired)
      if (a_stat && pair > 1) return launch_gemm<192, TTSB200_EPI_AFFINE_ACT, true, true>(*a, d, stream);
      if (a_stat) return launch_gemm<192, TTSB200_EPI_AFFINE_ACT, true>(*a, d, stream);
      if (pair > 1) return launch_gemm<192, TTSB200_EPI_AFFINE_ACT, false, true>(*a, d, stream);
      return launch_gemm<192, TTSB200_EPI_AFFINE_ACT, false>(*a, d, stream);
    }
    if (bn == 256) {
      if (a_stat && pair) return launch_gemm<256, TTSB200_EPI_AFFINE_ACT, true, true>(*a, d, stream);
      if (a_stat) return launch_gemm<256, TTSB200_EPI_AFFINE_ACT, true>(*a, d, stream);
      if (pair) return launch_gemm<256, TTSB200_EPI_AFFINE_ACT, false, true>(*a, d, stream);
      return launch_gemm<256, TTSB200_EPI_AFFINE_ACT, false>(*a, d, stream);
    }
  } else if (e == TTSB200_EPI_RES_LN || e == TTSB200_EPI_FFN2) {
    TTSB_CHECK_ARG(bn == a->N, "conv_gemm: LN epilogue needs block_n == N");
    TTSB_CHECK_ARG(a->res_f32 && a->ln1_g && a->ln1_b && a->out_bf16, "conv_gemm: RES_LN/FFN2 operands");
    // the fp32 residual is fetched by TMA: dense rows, per-utterance (T*N) or broadcast (0) batch stride
    TTSB_CHECK_ARG(a->res_tstride == a->N &&
                       (a->res_bstride == static_cast<int64_t>(a->T) * a->N || a->res_bstride == 0) &&
                       aligned16(a->res_f32),
                   "conv_gemm: res_f32 must be dense (tstride == N, bstride == T*N or 0) and 16-byte aligned");
    d.res_bcast = a->res_bstride == 0 ? 1 : 0;
    const bool pair_ln = g_gemm_pair && static_cast<long long>(a->B) * ((a->T + kBlockM - 1) / kBlockM) >= 2;
    if (e == TTSB200_EPI_FFN2) {
      TTSB_CHECK_ARG(a->res_bf16 && a->ln2_g && a->ln2_b && aligned16(a->res_bf16), "conv_gemm: FFN2 operands");
      TTSB_CHECK_ARG(!a->res_scale, "conv_gemm: res_scale is a RES_LN operand");
      if (bn == 384 && pair_ln) return launch_gemm<384, TTSB200_EPI_FFN2, false, true>(*a, d, stream);
      if (bn == 384) return launch_gemm<384, TTSB200_EPI_FFN2, false>(*a, d, stream);
    } else {
      if (bn == 384 && pair_ln) return launch_gemm<384, TTSB200_EPI_RES_LN, false, true>(*a, d, stream);
      if (bn == 384) return launch_gemm<384, TTSB200_EPI_RES_LN, false>(*a, d, stream);
    }
  } else if (e == TTSB200_EPI_RELU_LN) {
    TTSB_CHECK_ARG(bn == a->N, "conv_gemm: LN epilogue needs block_n == N");
    TTSB_CHECK_ARG(a->ln1_g && a->ln1_b && a->out_bf16, "conv_gemm: RELU_LN operands");
    if (bn == 256) return launch_gemm<256, TTSB200_EPI_RELU_LN, false>(*a, d, stream);
  } else if (e == TTSB200_EPI_RELU_LN_DOT) {
    TTSB_CHECK_ARG(bn == a->N, "conv_gemm: LN epilogue needs block_n == N");
    TTSB_CHECK_ARG(a->ln1_g && a->ln1_b && a->dot_w && a->dot_b && a->out_row, "conv_gemm: RELU_LN_DOT operands");
    if (bn == 256) return launch_gemm<256, TTSB200_EPI_RELU_LN_DOT, false>(*a, d, stream);
  }
  set_last_error("conv_gemm: unsupported (epi=%d, block_n=%d)", e, bn);
  return TTSB_ERR_ARG;
}
