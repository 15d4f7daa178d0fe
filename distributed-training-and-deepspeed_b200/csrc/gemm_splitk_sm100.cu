// Split-K tcgen05 GEMM for sm_100a: D[M,N] = epilogue(alpha * A·Bᵀ) with `splits` CTAs per tile.
//
// Why: at a few thousand tokens the N = hidden projections of a transformer (out-proj, FFN down,
// and the dgrads of QKV / FFN up) have 48-96 output tiles for 148 SMs and a long K.  The main loop
// of gemm_sm100.cu is paced by the per-SM operand fill rate (0.29 us per 32 KB k-block, see
// profiles/gemm_timeline_warm.txt), so the idle SMs are the loss: splitting K over them shortens
// the main loop by the split factor at the price of one fp32 round trip of the partial tiles
// through L2.
//
// One CTA per (tile, split) unit, all units resident at once (host checks tiles*splits <= SMs).
// Blocks [0, tiles) are the split-0 "owners".  A CTA with split > 0 parks its raw fp32 accumulators
// in the workspace and bumps the tile's arrival counter (release); the owner waits for splits-1
// arrivals (acquire), adds the partials in split order -- bit-reproducible -- runs the same fused
// epilogue as the plain kernel (bias, activation or its derivative, residual, bf16/fp32 store,
// accumulate) and clears the counter for the next launch on the stream.
//
// This file is self-contained on purpose (it repeats the tile constants and row helpers of
// gemm_sm100.cu instead of sharing a header): the plain kernel's code generation stays exactly
// what was profiled.  Warp roles as there: warp 0 TMA producer, warp 1 MMA issuer, warps 2..9
// epilogue (two per TMEM lane quadrant, each half of the tile's columns).
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstring>

#include "common.cuh"
#include "dtds_b200.h"
#include "internal.h"

namespace dtds {
namespace {

constexpr int kBM = 128;
constexpr int kBK = 64;
constexpr int kThreads = 320;

template <int BN>
struct SkCfg {
  static constexpr int kAStage = kBM * kBK * 2;
  static constexpr int kBStage = BN * kBK * 2;
  static constexpr int kStage = kAStage + kBStage;
  static constexpr int kStages = (BN == 256) ? 4 : 6;
  static constexpr int kTmemCols = BN;  // one accumulator: a CTA owns exactly one unit
  static constexpr int kSmemBytes = kStages * kStage + 1024 /*align slack*/ + 256 /*barriers*/;
};

struct SkParams {
  void* D;
  const __nv_bfloat16* bias;
  const __nv_bfloat16* residual;
  __nv_bfloat16* pre_out;
  const __nv_bfloat16* dact_in;
  int M, N, K;
  int ldd, ld_res, ld_dact;
  int act, out_fp32, accumulate;
  float alpha;
  int splits;
  float* ws;            // [(splits-1) * tiles][kBM * BN] fp32 partial tiles
  uint32_t* ws_flags;   // [tiles] arrival counters, zero between launches
};

__device__ __forceinline__ float4 ld_cg_f4(const float* p) {  // L2-coherent: written by other SMs
  float4 r;
  asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
  return r;
}

template <class F>
__device__ __forceinline__ void combine32(const __nv_bfloat16* __restrict__ src, int nvalid, bool full,
                                          float (&v)[32], F f) {
  if (full) {
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const uint4 u = *reinterpret_cast<const uint4*>(src + g * 8);
      f(v[g * 8 + 0], bf16_lo(u.x)); f(v[g * 8 + 1], bf16_hi(u.x));
      f(v[g * 8 + 2], bf16_lo(u.y)); f(v[g * 8 + 3], bf16_hi(u.y));
      f(v[g * 8 + 4], bf16_lo(u.z)); f(v[g * 8 + 5], bf16_hi(u.z));
      f(v[g * 8 + 6], bf16_lo(u.w)); f(v[g * 8 + 7], bf16_hi(u.w));
    }
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (j < nvalid) f(v[j], __bfloat162float(src[j]));
  }
}
__device__ __forceinline__ void store32_bf16(__nv_bfloat16* __restrict__ dst, int nvalid, bool full,
                                             float (&v)[32], bool accumulate) {
  if (accumulate) combine32(dst, nvalid, full, v, [](float& a, float x) { a += x; });
  if (full) {
#pragma unroll
    for (int g = 0; g < 4; ++g)
      *reinterpret_cast<uint4*>(dst + g * 8) =
          make_uint4(pack_bf16x2(v[g * 8 + 0], v[g * 8 + 1]), pack_bf16x2(v[g * 8 + 2], v[g * 8 + 3]),
                     pack_bf16x2(v[g * 8 + 4], v[g * 8 + 5]), pack_bf16x2(v[g * 8 + 6], v[g * 8 + 7]));
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (j < nvalid) dst[j] = __float2bfloat16(v[j]);
  }
}
__device__ __forceinline__ void store32_f32(float* __restrict__ dst, int nvalid, bool full, float (&v)[32],
                                            bool accumulate) {
  if (full) {
#pragma unroll
    for (int g = 0; g < 8; ++g) {
      float4 o = make_float4(v[g * 4], v[g * 4 + 1], v[g * 4 + 2], v[g * 4 + 3]);
      if (accumulate) {
        const float4 old = *reinterpret_cast<const float4*>(dst + g * 4);
        o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
      }
      *reinterpret_cast<float4*>(dst + g * 4) = o;
    }
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (j < nvalid) dst[j] = accumulate ? dst[j] + v[j] : v[j];
  }
}

template <int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(kThreads)
gemm_splitk_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                   const SkParams p) {
  using Cfg = SkCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + Cfg::kStages * Cfg::kAStage;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStage);
  uint64_t* full = bars;
  uint64_t* empty = bars + Cfg::kStages;
  uint64_t* acc_full = bars + 2 * Cfg::kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_m = (p.M + kBM - 1) / kBM;
  const int num_n = (p.N + BN - 1) / BN;
  const int num_tiles = num_m * num_n;
  const int num_kb = (p.K + kBK - 1) / kBK;
  const int tile = static_cast<int>(blockIdx.x) % num_tiles;
  const int split = static_cast<int>(blockIdx.x) / num_tiles;
  const int kb_begin = split * num_kb / p.splits;        // host guarantees splits <= num_kb,
  const int kb_end = (split + 1) * num_kb / p.splits;    // so every range is non-empty
  const int m0 = (tile % num_m) * kBM;
  const int n0 = (tile / num_m) * BN;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(acc_full, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, Cfg::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();
  pdl_wait();   // operands, workspace, counters and the output belong to earlier kernels until here

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(&empty[s], ph ^ 1);
        mbar_expect_tx(&full[s], Cfg::kStage);
        uint8_t* a_dst = sA + s * Cfg::kAStage;
        uint8_t* b_dst = sB + s * Cfg::kBStage;
        if constexpr (!A_MN) {
          tma_load_2d(a_dst, &tmA, &full[s], kb * kBK, m0, kEvictNormal);
        } else {
#pragma unroll
          for (int j = 0; j < kBM / 64; ++j)
            tma_load_2d(a_dst + j * 8192, &tmA, &full[s], m0 + j * 64, kb * kBK, kEvictNormal);
        }
        if constexpr (!B_MN) {
          tma_load_2d(b_dst, &tmB, &full[s], kb * kBK, n0, kEvictNormal);
        } else {
#pragma unroll
          for (int j = 0; j < BN / 64; ++j)
            tma_load_2d(b_dst + j * 8192, &tmB, &full[s], n0 + j * 64, kb * kBK, kEvictNormal);
        }
        if (++s == Cfg::kStages) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(kBM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      int s = 0;
      uint32_t ph = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(sA + s * Cfg::kAStage);
        const uint32_t b_addr = smem_u32(sB + s * Cfg::kBStage);
#pragma unroll
        for (int kk = 0; kk < kBK / 16; ++kk) {
          const uint64_t adesc = A_MN ? make_smem_desc_sw128(a_addr + kk * 2048, 8192, 1024)
                                      : make_smem_desc_sw128(a_addr + kk * 32, 16, 1024);
          const uint64_t bdesc = B_MN ? make_smem_desc_sw128(b_addr + kk * 2048, 8192, 1024)
                                      : make_smem_desc_sw128(b_addr + kk * 32, 16, 1024);
          umma_ss(tmem_base, adesc, bdesc, idesc, (kb != kb_begin || kk != 0) ? 1u : 0u);
        }
        umma_commit(&empty[s]);
        if (++s == Cfg::kStages) { s = 0; ph ^= 1; }
      }
      umma_commit(acc_full);
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps (2..9)
    const int q = warp & 3;             // TMEM lane quadrant this warp may access
    const int chalf = (warp - 2) >> 2;  // which half of the tile's 32-column chunks
    constexpr int kChunksPerWarp = BN / 64;
    constexpr size_t kTileElems = static_cast<size_t>(kBM) * BN;
    const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const int row = m0 + q * 32 + lane;
    const bool row_ok = row < p.M;
    mbar_wait(acc_full, 0);
    tc_fence_after();

    if (split != 0) {
      // partial sums of this k-range: raw accumulators into this unit's tile-sized slot
      float* wrow = p.ws + (static_cast<size_t>(split - 1) * num_tiles + tile) * kTileElems +
                    static_cast<size_t>(q * 32 + lane) * BN;
#pragma unroll 1
      for (int g = 0; g < kChunksPerWarp; ++g) {
        const int c = chalf * kChunksPerWarp + g;
        uint32_t r[32];
        tmem_ld_32x32(t_row + c * 32, r);
        tmem_ld_wait();
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
        store32_f32(wrow + c * 32, 32, true, v, false);
      }
      __threadfence();
      named_bar_sync(4, 256);   // all eight epilogue warps have written their part
      if (warp == 2 && lane == 0)
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(p.ws_flags + tile) : "memory");
    } else {
      if (lane == 0) {   // owner: wait for the other splits of this tile
        uint32_t spins = 0, seen;
        const uint32_t want = static_cast<uint32_t>(p.splits - 1);
        do {
          asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(p.ws_flags + tile) : "memory");
          if (++spins > DTDS_SPIN_LIMIT) {
            printf("[dtds] gemm(split-K): tile %d saw %u of %u partial sums\n", tile, seen, want);
            __trap();
          }
        } while (seen < want);
      }
      __syncwarp();
#pragma unroll 1
      for (int g = 0; g < kChunksPerWarp; ++g) {
        const int c = chalf * kChunksPerWarp + g;
        const int col0 = n0 + c * 32;
        if (col0 >= p.N) break;  // warp-uniform
        uint32_t r[32];
        tmem_ld_32x32(t_row + c * 32, r);
        tmem_ld_wait();
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
        for (int sp = 1; sp < p.splits; ++sp) {   // fixed order: bit-reproducible
          const float* wsrc = p.ws + (static_cast<size_t>(sp - 1) * num_tiles + tile) * kTileElems +
                              static_cast<size_t>(q * 32 + lane) * BN + c * 32;
#pragma unroll
          for (int k4 = 0; k4 < 8; ++k4) {
            const float4 x = ld_cg_f4(wsrc + k4 * 4);
            v[k4 * 4] += x.x; v[k4 * 4 + 1] += x.y; v[k4 * 4 + 2] += x.z; v[k4 * 4 + 3] += x.w;
          }
        }
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] *= p.alpha;
        const bool full_chunk = (col0 + 32 <= p.N);
        const int nvalid = full_chunk ? 32 : p.N - col0;
        const size_t out_off = static_cast<size_t>(row_ok ? row : 0) * p.ldd + col0;
        if (p.bias != nullptr)
          combine32(p.bias + col0, nvalid, full_chunk, v, [](float& a, float x) { a += x; });
        if (p.pre_out != nullptr && row_ok) store32_bf16(p.pre_out + out_off, nvalid, full_chunk, v, false);
        if (p.dact_in != nullptr) {
          if (row_ok) {
            const __nv_bfloat16* dsrc = p.dact_in + static_cast<size_t>(row) * p.ld_dact + col0;
            if (p.act == ACT_GELU_ERF)
              combine32(dsrc, nvalid, full_chunk, v, [](float& a, float x) { a *= gelu_erf_grad(x); });
            else if (p.act == ACT_RELU)
              combine32(dsrc, nvalid, full_chunk, v, [](float& a, float x) { a = x > 0.f ? a : 0.f; });
            else if (p.act == ACT_GELU_TANH)
              combine32(dsrc, nvalid, full_chunk, v, [](float& a, float x) { a *= gelu_tanh_grad(x); });
          }
        } else if (p.act == ACT_GELU_ERF) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
        } else if (p.act == ACT_RELU) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
        } else if (p.act == ACT_GELU_TANH) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = gelu_tanh(v[j]);
        }
        if (p.residual != nullptr && row_ok)
          combine32(p.residual + static_cast<size_t>(row) * p.ld_res + col0, nvalid, full_chunk, v,
                    [](float& a, float x) { a += x; });
        if (row_ok) {
          if (p.out_fp32)
            store32_f32(reinterpret_cast<float*>(p.D) + out_off, nvalid, full_chunk, v, p.accumulate != 0);
          else
            store32_bf16(reinterpret_cast<__nv_bfloat16*>(p.D) + out_off, nvalid, full_chunk, v,
                         p.accumulate != 0);
        }
      }
      named_bar_sync(4, 256);   // every owner warp is past its wait and its workspace reads
      if (warp == 2 && lane == 0) p.ws_flags[tile] = 0u;   // ready for the next launch on this stream
    }
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) tmem_dealloc(tmem_base, Cfg::kTmemCols);
}

template <int BN, bool A_MN, bool B_MN>
int launch_splitk(const CUtensorMap& tmA, const CUtensorMap& tmB, const SkParams& p, int tiles,
                  cudaStream_t stream) {
  using Cfg = SkCfg<BN>;
  auto kern = gemm_splitk_kernel<BN, A_MN, B_MN>;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    DTDS_HOST_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    attr_set[dev & 63] = true;
  }
  DTDS_HOST_CHECK(launch_pdl(kern, dim3(tiles * p.splits), dim3(kThreads), Cfg::kSmemBytes, stream, tmA, tmB, p));
  return 0;
}

}  // namespace
}  // namespace dtds

using namespace dtds;

extern "C" int dtds_gemm_splitk_tiles(int M, int N, int block_n) {
  if (block_n != 128 && block_n != 256) return -1;
  return ((M + kBM - 1) / kBM) * ((N + block_n - 1) / block_n);
}

extern "C" size_t dtds_gemm_splitk_workspace_bytes(int M, int N, int block_n, int splits) {
  const int tiles = dtds_gemm_splitk_tiles(M, N, block_n);
  if (tiles <= 0 || splits < 2) return 0;
  return static_cast<size_t>(tiles) * (splits - 1) * kBM * block_n * sizeof(float);
}

extern "C" int dtds_gemm_bf16_splitk(const void* A, const void* B, void* D, const void* bias,
                                     const void* residual, void* pre_out, const void* dact_in, int M,
                                     int N, int K, int lda, int ldb, int ldd, int ld_res, int ld_dact,
                                     int a_mn_major, int b_mn_major, int act, int out_fp32,
                                     int accumulate, float alpha, int block_n, int splits,
                                     float* workspace, unsigned* flags, cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15) ||
      (reinterpret_cast<uintptr_t>(D) & 15) || (reinterpret_cast<uintptr_t>(workspace) & 15) || (lda % 8) ||
      (ldb % 8) || (ldd % (out_fp32 ? 4 : 8))) {
    fprintf(stderr, "[dtds] gemm(split-K): operands and workspace must be 16-byte aligned with leading "
                    "dimensions that are multiples of 8 elements (lda=%d ldb=%d ldd=%d)\n", lda, ldb, ldd);
    return -3;
  }
  const int tiles = dtds_gemm_splitk_tiles(M, N, block_n);
  const int num_kb = (K + kBK - 1) / kBK;
  if (tiles <= 0 || splits < 2 || splits > num_kb || tiles * splits > sm_count() || workspace == nullptr ||
      flags == nullptr) {
    fprintf(stderr,
            "[dtds] gemm(split-K): needs BLOCK_N 128 or 256, 2 <= splits <= k-blocks (%d), tiles*splits <= "
            "SMs (%d*%d vs %d), a workspace and counters\n",
            num_kb, tiles, splits, sm_count());
    return -8;
  }
  CUtensorMap tmA, tmB;
  int rc;
  if (!a_mn_major) rc = make_tmap_2d_bf16(&tmA, A, K, M, lda, kBK, kBM);
  else             rc = make_tmap_2d_bf16(&tmA, A, M, K, lda, 64, kBK);
  if (rc) return rc;
  if (!b_mn_major) rc = make_tmap_2d_bf16(&tmB, B, K, N, ldb, kBK, block_n);
  else             rc = make_tmap_2d_bf16(&tmB, B, N, K, ldb, 64, kBK);
  if (rc) return rc;

  SkParams p;
  memset(&p, 0, sizeof(p));
  p.D = D;
  p.bias = static_cast<const __nv_bfloat16*>(bias);
  p.residual = static_cast<const __nv_bfloat16*>(residual);
  p.pre_out = static_cast<__nv_bfloat16*>(pre_out);
  p.dact_in = static_cast<const __nv_bfloat16*>(dact_in);
  p.M = M; p.N = N; p.K = K;
  p.ldd = ldd; p.ld_res = ld_res > 0 ? ld_res : ldd; p.ld_dact = ld_dact > 0 ? ld_dact : ldd;
  p.act = act; p.out_fp32 = out_fp32; p.accumulate = accumulate; p.alpha = alpha;
  p.splits = splits; p.ws = workspace; p.ws_flags = flags;

#define DTDS_SK(BN_)                                                                                  \
  if (block_n == BN_) {                                                                               \
    if (!a_mn_major && !b_mn_major) return launch_splitk<BN_, false, false>(tmA, tmB, p, tiles, stream); \
    if (!a_mn_major && b_mn_major) return launch_splitk<BN_, false, true>(tmA, tmB, p, tiles, stream);   \
    if (a_mn_major && !b_mn_major) return launch_splitk<BN_, true, false>(tmA, tmB, p, tiles, stream);   \
    return launch_splitk<BN_, true, true>(tmA, tmB, p, tiles, stream);                                \
  }
  DTDS_SK(128)
  DTDS_SK(256)
#undef DTDS_SK
  return -8;
}
