// TMA load-throughput probe for sm_100a: how many cycles does one SM need per box, as a function of the
// box shape (inner bytes x rows), swizzle mode and CTAs per SM?  One thread per CTA streams boxes from an
// L2-resident tensor into a ring of smem slots; nobody reads the data.  Also probes 1-D bulk copies.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/tma_probe tools/tma_probe.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t c) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(c) : "memory");
}
__device__ __forceinline__ void mbar_expect(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok = 0;
  while (!ok) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  }
}

// mode 0: 2-D tensor boxes; mode 1: 1-D bulk copies of box_bytes
__global__ void probe(const __grid_constant__ CUtensorMap tm, const uint8_t* base, int mode, int box_bytes,
                      int boxes_per_slot, int slots, int iters, int rows_total, int box_rows, int inner_el,
                      int cols_el, long long* cycles) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bars[8];
  if (threadIdx.x == 0) {
    for (int i = 0; i < slots; ++i) mbar_init(smem_u32(&bars[i]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const int slot_bytes = box_bytes * boxes_per_slot;
    long long t0 = clock64();
    int r = (blockIdx.x * 977) % (rows_total - box_rows);
    int c = 0;
    for (int it = 0; it < iters; ++it) {
      const int s = it % slots;
      if (it >= slots) mbar_wait(smem_u32(&bars[s]), ((it / slots) - 1) & 1);
      const uint32_t bar = smem_u32(&bars[s]);
      mbar_expect(bar, slot_bytes);
      for (int k = 0; k < boxes_per_slot; ++k) {
        const uint32_t dst = smem_u32(smem + s * slot_bytes + k * box_bytes);
        if (mode == 0) {
          asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                       ::"r"(dst), "l"((uint64_t)&tm), "r"(bar), "r"(c), "r"(r) : "memory");
          c += inner_el;
          if (c + inner_el > cols_el) { c = 0; r = (r + box_rows) % (rows_total - box_rows); }
        } else {
          const uint8_t* src = base + ((size_t)(blockIdx.x * 131 + it * boxes_per_slot + k) * box_bytes) % ((size_t)rows_total * cols_el * 2 - box_bytes);
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                       ::"r"(dst), "l"((uint64_t)(((uintptr_t)src) & ~(uintptr_t)15)), "r"(box_bytes), "r"(bar) : "memory");
        }
      }
    }
    for (int it = iters; it < iters + slots; ++it) {
      const int s = it % slots;
      mbar_wait(smem_u32(&bars[s]), ((it / slots) - 1) & 1);
    }
    cycles[blockIdx.x] = clock64() - t0;
  }
}

int main() {
  void* fp = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q);
  EncodeTiledFn enc = (EncodeTiledFn)fp;
  const int rows_total = 16384, cols_el = 1152;  // bf16 tensor 16384 x 1152 = 37.7 MB (L2 resident)
  uint8_t* d;
  cudaMalloc(&d, (size_t)rows_total * cols_el * 2);
  cudaMemset(d, 1, (size_t)rows_total * cols_el * 2);
  long long* dc;
  cudaMalloc(&dc, 1024 * sizeof(long long));
  struct Cfg { const char* name; int mode, inner_el, box_rows; CUtensorMapSwizzle sw; int boxes_per_slot, slots; };
  Cfg cfgs[] = {
      {"2D 128Bx128 SW128 (16KB) x2/slot", 0, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B, 2, 4},
      {"2D 128Bx256 SW128 (32KB) x1/slot", 0, 64, 256, CU_TENSOR_MAP_SWIZZLE_128B, 1, 4},
      {"2D  64Bx128 SW64  ( 8KB) x4/slot", 0, 32, 128, CU_TENSOR_MAP_SWIZZLE_64B, 4, 4},
      {"2D  64Bx256 SW64  (16KB) x2/slot", 0, 32, 256, CU_TENSOR_MAP_SWIZZLE_64B, 2, 4},
      {"2D 128Bx64  SW128 ( 8KB) x4/slot", 0, 64, 64, CU_TENSOR_MAP_SWIZZLE_128B, 4, 4},
      {"2D 256Bx128 NONE  (32KB) x1/slot", 0, 128, 128, CU_TENSOR_MAP_SWIZZLE_NONE, 1, 4},
      {"2D 512Bx64  NONE  (32KB) x1/slot", 0, 256, 64, CU_TENSOR_MAP_SWIZZLE_NONE, 1, 4},
      {"1D bulk 32KB x1/slot", 1, 64, 256, CU_TENSOR_MAP_SWIZZLE_NONE, 1, 4},
      {"1D bulk 8KB x4/slot", 1, 64, 64, CU_TENSOR_MAP_SWIZZLE_NONE, 4, 4},
  };
  const int iters = 400;
  for (int per_sm = 1; per_sm <= 2; ++per_sm) {
    for (auto& c : cfgs) {
      CUtensorMap tm;
      cuuint64_t gdim[2] = {(cuuint64_t)cols_el, (cuuint64_t)rows_total};
      cuuint64_t gstr[1] = {(cuuint64_t)cols_el * 2};
      cuuint32_t box[2] = {(cuuint32_t)c.inner_el, (cuuint32_t)c.box_rows};
      cuuint32_t es[2] = {1, 1};
      CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, d, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                       c.sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) { printf("%s: encode failed %d\n", c.name, (int)r); continue; }
      const int box_bytes = c.inner_el * 2 * c.box_rows;
      const int smem = box_bytes * c.boxes_per_slot * c.slots;  // 128 KB -> 1/SM; for 2/SM use 2 slots
      const int slots = per_sm == 1 ? c.slots : 2;
      const int smem_use = box_bytes * c.boxes_per_slot * slots + (per_sm == 1 ? 0 : 0);
      (void)smem;
      cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
      // force occupancy: per_sm==1 -> pad smem to >114 KB
      const int dyn = per_sm == 1 ? 160 * 1024 : smem_use;
      for (int rep = 0; rep < 2; ++rep)
        probe<<<148 * per_sm, 32, dyn>>>(tm, d, c.mode, box_bytes, c.boxes_per_slot, slots, iters, rows_total, c.box_rows,
                                         c.inner_el, cols_el, dc);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("%s: %s\n", c.name, cudaGetErrorString(e)); return 1; }
      long long h[296];
      cudaMemcpy(h, dc, 148 * per_sm * sizeof(long long), cudaMemcpyDeviceToHost);
      double avg = 0;
      for (int i = 0; i < 148 * per_sm; ++i) avg += (double)h[i];
      avg /= 148 * per_sm;
      const double bytes = (double)iters * box_bytes * c.boxes_per_slot;
      const double rows = (double)iters * c.boxes_per_slot * c.box_rows;
      printf("%d CTA/SM  %-36s  %7.1f B/clk/SM  %6.2f clk/row  (%.0f cycles)\n", per_sm, c.name, bytes * per_sm / avg,
             c.mode == 0 ? avg / rows / per_sm : 0.0, avg);
    }
  }
  return 0;
}
