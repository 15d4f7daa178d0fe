// Micro-benchmark: shared-memory fill rate of TMA tile loads (128B swizzle, 64-element-wide boxes)
// from an L2-resident bf16 matrix, per SM, as a function of box height, boxes in flight and the
// number of CTAs pulling at once.  Prints bytes per SM clock.
#include <cstdio>
#include <cstdlib>
#include "common.cuh"
#include "internal.h"
#include <cuda.h>
using namespace dtds;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

__device__ __forceinline__ void mbar_poll(uint64_t* bar, uint32_t parity) {
  uint32_t ok = 0;
  while (!ok) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  }
}
template <bool POLL>
__global__ void fill_kernel(const __grid_constant__ CUtensorMap tm, long long* out, int rows_per_box, int stages,
                            int iters, int n_rowtiles, int n_kb, int loads_per_stage) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t full[16];
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) mbar_init(&full[s], 1);
    fence_mbar_init();
    tma_prefetch_desc(&tm);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t box_bytes = rows_per_box * 128;
    const uint32_t stage_bytes = box_bytes * loads_per_stage;
    long long t0 = clock64();
    uint32_t ph = 0; int s = 0;
    for (int i = 0; i < iters + stages; ++i) {
      if (i >= stages) {  // wait for the load issued `stages` iterations ago, then reuse its slot
        if (POLL) mbar_poll(&full[s], ph); else mbar_wait(&full[s], ph);
      }
      if (i < iters) {
        mbar_expect_tx(&full[s], stage_bytes);
        const int rt = (blockIdx.x * 7 + i / n_kb) % n_rowtiles, kb = i % n_kb;
        for (int l = 0; l < loads_per_stage; ++l)
          tma_load_2d(smem + s * stage_bytes + l * box_bytes, &tm, &full[s], kb * 64,
                      ((rt + l * 3) % n_rowtiles) * rows_per_box, kEvictNormal);
      }
      if (++s == stages) { s = 0; if (i >= stages) ph ^= 1; }
    }
    long long t1 = clock64();
    out[blockIdx.x] = t1 - t0;
  }
}

// N boxes issued back to back, each on its own mbarrier; then wait for all.
__global__ void burst_kernel(const __grid_constant__ CUtensorMap tm, long long* out, int rows_per_box, int nbox, int one_bar) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t full[16];
  if (threadIdx.x == 0) {
    for (int s = 0; s < 16; ++s) mbar_init(&full[s], 1);
    fence_mbar_init();
    tma_prefetch_desc(&tm);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t box_bytes = rows_per_box * 128;
    for (int rep = 0; rep < 3; ++rep) {   // rep 0 warms L2 and the descriptor
      long long t0 = clock64();
      if (one_bar) mbar_expect_tx(&full[0], box_bytes * nbox);
      for (int i = 0; i < nbox; ++i) {
        if (!one_bar) mbar_expect_tx(&full[i], box_bytes);
        tma_load_2d(smem + i * box_bytes, &tm, &full[one_bar ? 0 : i], (i % 12) * 64, (blockIdx.x % 8) * rows_per_box, kEvictNormal);
      }
      long long t1 = clock64();
      for (int i = 0; i < (one_bar ? 1 : nbox); ++i) mbar_wait(&full[i], rep & 1);
      long long t2 = clock64();
      out[blockIdx.x * 2] = t1 - t0;
      out[blockIdx.x * 2 + 1] = t2 - t0;
    }
  }
}

int main() {
  const int R = 8192, C = 768;
  __nv_bfloat16* d; cudaMalloc(&d, (size_t)R * C * 2); cudaMemset(d, 0, (size_t)R * C * 2);
  long long* o; cudaMalloc(&o, 148 * 2 * 8);
  void* sym = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q);
  EncodeTiledFn enc = (EncodeTiledFn)sym;
  cudaFuncSetAttribute(fill_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(fill_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  {
    cudaFuncSetAttribute(burst_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    CUtensorMap tm;
    cuuint64_t dims[2] = {(cuuint64_t)C, (cuuint64_t)R}; cuuint64_t strides[1] = {(cuuint64_t)C * 2};
    cuuint32_t box[2] = {64, 128}; cuuint32_t es[2] = {1, 1};
    enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    for (int one_bar = 0; one_bar < 2; ++one_bar)
      for (int nbox : {1, 2, 4, 8, 12})
        for (int grid : {1, 148}) {
          burst_kernel<<<grid, 32, 12 * 16384 + 2048>>>(tm, o, 128, nbox, one_bar);
          cudaError_t e = cudaDeviceSynchronize();
          long long h[296]; cudaMemcpy(h, o, grid * 16, cudaMemcpyDeviceToHost);
          printf("burst %2d boxes of 16 KB, %s, %3d CTAs: issued after %5lld clk, all landed after %5lld clk [%s]\n", nbox,
                 one_bar ? "one barrier " : "own barriers", grid, h[0], h[1], cudaGetErrorString(e));
        }
  }
  const int cfgs[][3] = {{64, 1, 12}, {128, 1, 6}, {256, 1, 4}, {128, 2, 6}, {128, 2, 3}, {64, 4, 6}, {256, 1, 6}, {128, 3, 4}};
  for (auto& c : cfgs) {
    const int rows = c[0], lps = c[1], stages = c[2];
    CUtensorMap tm;
    cuuint64_t dims[2] = {(cuuint64_t)C, (cuuint64_t)R}; cuuint64_t strides[1] = {(cuuint64_t)C * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)rows}; cuuint32_t es[2] = {1, 1};
    enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    for (int poll = 0; poll < 2; ++poll)
    for (int grid : {1, 148}) {
      const int iters = 512;
      for (int rep = 0; rep < 2; ++rep)
        if (poll) fill_kernel<true><<<grid, 32, stages * lps * rows * 128 + 2048>>>(tm, o, rows, stages, iters, R / rows, C / 64, lps);
        else fill_kernel<false><<<grid, 32, stages * lps * rows * 128 + 2048>>>(tm, o, rows, stages, iters, R / rows, C / 64, lps);
      cudaError_t e = cudaDeviceSynchronize();
      long long h[148]; cudaMemcpy(h, o, grid * 8, cudaMemcpyDeviceToHost);
      double mx = 0; for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
      printf("%s box 64x%-3d x%d per stage, %d stages (%3d KB in flight), %3d CTAs: %6.1f B/clk/SM, %5.0f clk per stage [%s]\n",
             poll ? "test_wait" : "try_wait ", rows, lps, stages, stages * lps * rows * 128 / 1024, grid, (double)iters * lps * rows * 128 / mx, mx / iters,
             cudaGetErrorString(e));
    }
  }
  return 0;
}
