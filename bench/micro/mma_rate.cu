// Micro-benchmark: issue rate of tcgen05.mma (kind::f16, bf16 operands from shared memory) for
// different N, with nothing else running on the SM.  Prints SM clocks per MMA instruction.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../distributed-training-and-deepspeed_b200/csrc \
//        -I../../include mma_rate.cu -o mma_rate
#include <cstdio>
#include "common.cuh"
using namespace dtds;

template <int N>
__global__ void rate_kernel(long long* out, int iters, int same_addr) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  if (threadIdx.x < 32) { tmem_alloc(&slot, 512); tmem_relinquish(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  if (threadIdx.x == 0) {
    constexpr uint32_t idesc = make_idesc_bf16(128, N, 0, 0);
    const uint32_t a0 = smem_u32(smem), b0 = smem_u32(smem + 4 * 16384);
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      const int st = same_addr ? 0 : (i >> 2) & 3, kk = same_addr ? 0 : i & 3;
      const uint64_t ad = make_smem_desc_sw128(a0 + st * 16384 + kk * 32, 16, 1024);
      const uint64_t bd = make_smem_desc_sw128(b0 + st * (N * 128) + kk * 32, 16, 1024);
      umma_ss(tm + ((i >> 2) & 1) * N, ad, bd, idesc, 1u);
    }
    long long t1 = clock64();
    umma_commit(&bar);
    mbar_wait(&bar, 0);
    long long t2 = clock64();
    out[blockIdx.x * 2] = t1 - t0;
    out[blockIdx.x * 2 + 1] = t2 - t0;
  }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  if (threadIdx.x < 32) tmem_dealloc(tm, 512);
}

template <int N>
void run(int iters, int same) {
  long long* d; cudaMalloc(&d, 148 * 2 * sizeof(long long));
  const int smem = 4 * 16384 + 4 * N * 128 + 2048;
  cudaFuncSetAttribute(rate_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int rep = 0; rep < 2; ++rep) rate_kernel<N><<<148, 128, smem>>>(d, iters, same);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[296]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  printf("N=%3d same_addr=%d: issue %.1f clk/mma, complete %.1f clk/mma  (ideal %d)  [%s]\n", N, same,
         (double)h[0] / iters, (double)h[1] / iters, N / 2, cudaGetErrorString(e));
  cudaFree(d);
}

int main() {
  for (int same = 0; same < 2; ++same) {
    run<64>(2048, same); run<128>(2048, same); run<256>(2048, same);
  }
  return 0;
}
