// tcgen05 / TMEM / TMA GEMM family for sm_100a (bf16 operands, fp32 accumulation).
//
//   D[M,N] = epilogue( alpha * sum_k A[m,k] * B[n,k] )
//
// Operands may be K-major (the contraction index is contiguous in memory) or MN-major (the
// output index is contiguous), which covers forward (x·Wᵀ), dgrad (dy·W) and wgrad (dyᵀ·x)
// without transposing anything in memory.  Replaces the cuBLAS SGEMM calls behind every
// nn.Linear in the reference (model/transformer.py:37-40,50-51; HF BertLayer in
// model/bert_mp.py:22-24) — SURVEY.md rows K1/K2.
//
// Structure: persistent kernel, one CTA per SM, 6 warps:
//   warp 0      TMA producer    (cp.async.bulk.tensor -> 128B-swizzled smem ring, mbarrier tx)
//   warp 1      MMA issuer      (one thread issues tcgen05.mma, accumulators in TMEM, 2 stages)
//   warps 2..9  epilogue        (tcgen05.ld -> bias / activation / residual -> global); two warps
//                               per TMEM lane quadrant, each taking half of the tile's columns
// The epilogue of tile i overlaps the main loop of tile i+1 through the double-buffered TMEM
// accumulator.  When `peer_flags` is given (ZeRO-3 all-gather fused GEMM) the producer waits on
// the per-chunk arrival flag of the B rows it is about to load; see fused_comm.cu.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstring>
#include <mutex>

#include "common.cuh"
#include "dtds_b200.h"
#include "internal.h"

namespace dtds {

constexpr int kBM = 128;
constexpr int kBK = 64;
constexpr int kFetchThreads = 128;  // + 4 fetch warps in the all-gather fused variant
constexpr int kMaxFetchSegs = 16;

// CG = 1: one CTA per 128 x BN tile.  CG = 2: a CTA pair (cluster of 2, tcgen05 cta_group::2) per
// 256 x BN tile; each CTA stages its own 128 rows of A and BN/2 rows of B, so the pair moves half
// of the B bytes per output element, which is what bounds the main loop (L2 -> shared memory fill).
template <int BN, int CG = 1>
struct GemmCfg {
  static constexpr int EPI = 8;  // epilogue warps: two per TMEM lane quadrant
  static constexpr int kAStage = kBM * kBK * 2;
  static constexpr int kBStage = (BN / CG) * kBK * 2;
  static constexpr int kStage = kAStage + kBStage;
  static constexpr int kStages = CG == 2 ? (BN == 256 ? 6 : 8) : ((BN == 256) ? 4 : (BN == 128 ? 6 : 8));
  static constexpr int kTmemCols = 2 * BN;  // two accumulator stages
  static constexpr int kStageOut = EPI * 32 * 128;  // epilogue staging: EPI warps x 32 rows x 128 B
  static constexpr int kSmemBytes = kStages * kStage + kStageOut + 1024 /*align slack*/ + 256 /*barriers*/;
  static constexpr int kThreads = 64 + 32 * EPI;  // TMA warp + MMA warp + epilogue warps
};

struct GemmParams {
  void* D;
  const __nv_bfloat16* bias;      // [N] or null
  const __nv_bfloat16* residual;  // [M, ld_res] or null, added after the activation
  __nv_bfloat16* pre_out;         // [M, ldd] or null: pre-activation values kept for backward
  const __nv_bfloat16* dact_in;   // [M, ld_dact] or null: multiply result by act'(dact_in)
  const uint32_t* ready_flags;    // null, or per-B-chunk arrival flags (all-gather fused mode)
  uint32_t ready_epoch;           // value a flag must reach before its chunk may be read
  int rows_per_flag;              // B rows (n index) covered by one flag
  int M, N, K;
  int ldd, ld_res, ld_dact;
  int act;
  int out_fp32;
  int accumulate;  // D += result
  float alpha;
  long long* dbg;  // optional [grid][16] clock64 timeline (null in production)
  // ---- all-gather fused mode (ZeRO-3): operand B is assembled by this kernel's fetch warps out
  // of the owners' parameter shards (peer loads over NVLink) while the tiles that are already
  // complete are being multiplied.
  const char* seg_src[kMaxFetchSegs];             // source of segment i (peer or local address)
  unsigned long long seg_begin[kMaxFetchSegs + 1];  // byte offset in B where segment i starts
  char* fetch_dst;                                // local full copy of B
  uint32_t* fetch_flags;                          // one arrival flag per chunk (== ready_flags)
  unsigned long long fetch_total;                 // bytes of B
  unsigned int fetch_chunk;                       // bytes per chunk (multiple of 16)
  int fetch_nseg;
  int flags_on_k;                                 // flags index rows of B along K (dgrad) not N
  // Local-first order (all-gather fused mode): tiles / k-blocks / chunks are visited cyclically
  // starting at this rank's own shard, which needs no NVLink transfer, so the tensor pipe has work
  // from the first microsecond while the fetch warps pull the remote shards.
  int n_rot, k_rot;
  unsigned chunk_rot;
};

// Spins (bounded) until flags[f0..f1] have reached `epoch`, then fences generic -> async proxy.
__device__ __forceinline__ void wait_ready_flags(const uint32_t* flags, int f0, int f1, uint32_t epoch) {
  for (int f = f0; f <= f1; ++f) {
    uint32_t spins = 0, v;
    do {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + f) : "memory");
      if (++spins > DTDS_SPIN_LIMIT) {
        printf("[dtds] gemm: ready flag %d stuck at %u (want %u)\n", f, v, epoch);
        __trap();
      }
    } while (v < epoch);
  }
  asm volatile("fence.proxy.async;" ::: "memory");
}
__device__ __forceinline__ uint4 ld_sys_v4(const void* p) {
  uint4 r;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}

// ---- epilogue row helpers: 32 consecutive columns of one output row, registers only ----------
template <class F>
__device__ __forceinline__ void combine_row32(const __nv_bfloat16* __restrict__ src, int nvalid,
                                              bool full, float (&v)[32], F f) {
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
__device__ __forceinline__ void store_row32_bf16(__nv_bfloat16* __restrict__ dst, int nvalid, bool full,
                                                 float (&v)[32], bool accumulate) {
  if (accumulate) combine_row32(dst, nvalid, full, v, [](float& a, float x) { a += x; });
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
__device__ __forceinline__ void store_row32_f32(float* __restrict__ dst, int nvalid, bool full,
                                                float (&v)[32], bool accumulate) {
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

template <int BN, bool A_MN, bool B_MN, bool FETCH, int CG = 1>
// 96 registers: 320 threads x 96 leaves room on the SM for a 512-thread, 64-register
// communication block (overlapped gradient-bucket all-reduce) next to this persistent CTA.
__global__ void __maxnreg__(96)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const GemmParams p) {
  using Cfg = GemmCfg<BN, CG>;
  constexpr int EPI = Cfg::EPI;
  static_assert(!(FETCH && CG == 2), "the all-gather fused mode runs single-CTA tiles");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + Cfg::kStages * Cfg::kAStage;
  uint8_t* s_out = smem + Cfg::kStages * Cfg::kStage;  // per-warp output staging (128B-pitch, XOR swizzle)
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_out + Cfg::kStageOut);
  uint64_t* full = bars;
  uint64_t* empty = bars + Cfg::kStages;
  uint64_t* tmem_full = bars + 2 * Cfg::kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // CTA pair: rank 0 (leader) owns the `full` and `tmem_empty` barriers and issues the MMAs.
  const uint32_t crank = CG == 2 ? cluster_ctarank() : 0u;
  const int unit = CG == 2 ? (blockIdx.x >> 1) : blockIdx.x;       // persistent tile walker id
  const int nunits = CG == 2 ? (gridDim.x >> 1) : gridDim.x;
  long long* dbg = p.dbg != nullptr ? p.dbg + static_cast<size_t>(blockIdx.x) * 16 : nullptr;
  if (dbg != nullptr && threadIdx.x == 0) dbg[0] = clock64();

  const int num_m = (p.M + kBM * CG - 1) / (kBM * CG);
  const int num_n = (p.N + BN - 1) / BN;
  const int num_tiles = num_m * num_n;
  const int num_kb = (p.K + kBK - 1) / kBK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], EPI * CG);
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    if constexpr (CG == 2) {
      tmem_alloc_pair(tmem_slot, Cfg::kTmemCols);
      tmem_relinquish_pair();
    } else {
      tmem_alloc(tmem_slot, Cfg::kTmemCols);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  if constexpr (CG == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();
  pdl_wait();   // operands, bias/residual and the output buffer belong to earlier kernels until here
  if (dbg != nullptr && threadIdx.x == 0) dbg[1] = clock64();  // setup done

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      for (int tile = unit; tile < num_tiles; tile += nunits) {
        const int m0 = (tile % num_m) * (kBM * CG) + static_cast<int>(crank) * kBM;
        int nt = tile / num_m + p.n_rot;
        if (nt >= num_n) nt -= num_n;
        const int n0 = nt * BN;
        const int nb0 = n0 + static_cast<int>(crank) * (BN / CG);   // first B row this CTA stages
        if (p.ready_flags != nullptr && !p.flags_on_k) {
          // All-gather fused mode: B rows [n0, n0+BN) are written by fetch warps (of any CTA of
          // this grid, or by an earlier kernel); wait for the arrival flag(s) covering them, then
          // order those generic-proxy writes before our async-proxy (TMA) reads.
          const int n_last = min(n0 + BN, p.N) - 1;
          wait_ready_flags(p.ready_flags, n0 / p.rows_per_flag, n_last / p.rows_per_flag, p.ready_epoch);
        }
        for (int kbi = 0; kbi < num_kb; ++kbi) {
          int kb = kbi + p.k_rot;   // k order is free: the accumulator does not care
          if (kb >= num_kb) kb -= num_kb;
          if (p.ready_flags != nullptr && p.flags_on_k) {  // dgrad: B rows are indexed by k
            const int k_last = min(kb * kBK + kBK, p.K) - 1;
            wait_ready_flags(p.ready_flags, kb * kBK / p.rows_per_flag, k_last / p.rows_per_flag,
                             p.ready_epoch);
          }
          mbar_wait(&empty[s], ph ^ 1);
          uint8_t* a_dst = sA + s * Cfg::kAStage;
          uint8_t* b_dst = sB + s * Cfg::kBStage;
          if constexpr (CG == 2) {
            // Both CTAs' bytes are counted on the leader's barrier: the MMA needs both halves.
            if (crank == 0) mbar_expect_tx(&full[s], 2 * Cfg::kStage);
            const uint32_t fb = mapa_u32(smem_u32(&full[s]), 0);
            if constexpr (!A_MN) {
              tma_load_2d_pair(a_dst, &tmA, fb, kb * kBK, m0, kEvictNormal);
            } else {
#pragma unroll
              for (int j = 0; j < kBM / 64; ++j)
                tma_load_2d_pair(a_dst + j * 8192, &tmA, fb, m0 + j * 64, kb * kBK, kEvictNormal);
            }
            if constexpr (!B_MN) {
              tma_load_2d_pair(b_dst, &tmB, fb, kb * kBK, nb0, kEvictNormal);
            } else {
#pragma unroll
              for (int j = 0; j < BN / CG / 64; ++j)
                tma_load_2d_pair(b_dst + j * 8192, &tmB, fb, nb0 + j * 64, kb * kBK, kEvictNormal);
            }
          } else {
            mbar_expect_tx(&full[s], Cfg::kStage);
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
          }
          if (++s == Cfg::kStages) { s = 0; ph ^= 1; }
        }
      }
      if (dbg != nullptr) dbg[2] = clock64();  // all TMA loads issued
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0 && crank == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(kBM * CG, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      int s = 0;
      uint32_t ph = 0;
      int it = 0;
      for (int tile = unit; tile < num_tiles; tile += nunits, ++it) {
        const int as = it & 1;
        const uint32_t aph = (it >> 1) & 1;
        mbar_wait(&tmem_empty[as], aph ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full[s], ph);
          tc_fence_after();
          if (dbg != nullptr && it == 0 && kb == 0) dbg[3] = clock64();  // first operands landed
          const uint32_t a_addr = smem_u32(sA + s * Cfg::kAStage);
          const uint32_t b_addr = smem_u32(sB + s * Cfg::kBStage);
#pragma unroll
          for (int kk = 0; kk < kBK / 16; ++kk) {
            const uint64_t adesc = A_MN ? make_smem_desc_sw128(a_addr + kk * 2048, 8192, 1024)
                                        : make_smem_desc_sw128(a_addr + kk * 32, 16, 1024);
            const uint64_t bdesc = B_MN ? make_smem_desc_sw128(b_addr + kk * 2048, 8192, 1024)
                                        : make_smem_desc_sw128(b_addr + kk * 32, 16, 1024);
            if constexpr (CG == 2) umma_ss_pair(d_tmem, adesc, bdesc, idesc, (kb | kk) != 0 ? 1u : 0u);
            else                   umma_ss(d_tmem, adesc, bdesc, idesc, (kb | kk) != 0 ? 1u : 0u);
          }
          // frees the smem slot (in both CTAs of a pair) once these MMAs have read it
          if constexpr (CG == 2) umma_commit_pair(&empty[s], 0x3); else umma_commit(&empty[s]);
          if (++s == Cfg::kStages) { s = 0; ph ^= 1; }
        }
        // accumulator complete -> epilogue warps (of both CTAs of a pair)
        if constexpr (CG == 2) umma_commit_pair(&tmem_full[as], 0x3); else umma_commit(&tmem_full[as]);
        if (dbg != nullptr && it == 0) dbg[4] = clock64();  // all MMAs of the first tile issued
      }
    }
  } else if (FETCH && warp >= 2 + EPI) {
    // ------------------------------------------------------------------ fetch warps (all-gather)
    // Chunk c of B (fetch_chunk bytes) is assembled by CTA (c mod grid): 128 threads stream it
    // from the owning ranks' shard buffers with 4 independent 128-bit peer loads in flight per
    // thread, then publish the chunk's arrival flag with release semantics.
    const int ft = threadIdx.x - Cfg::kThreads;
    const unsigned long long nchunks = (p.fetch_total + p.fetch_chunk - 1) / p.fetch_chunk;
    for (unsigned long long ci = blockIdx.x; ci < nchunks; ci += gridDim.x) {
      unsigned long long c = ci + p.chunk_rot;
      if (c >= nchunks) c -= nchunks;
      const unsigned long long b0 = c * p.fetch_chunk;
      const unsigned long long b1 = min(p.fetch_total, b0 + p.fetch_chunk);
      for (int sg = 0; sg < p.fetch_nseg; ++sg) {
        const unsigned long long lo = max(b0, p.seg_begin[sg]);
        const unsigned long long hi = min(b1, p.seg_begin[sg + 1]);
        if (lo >= hi) continue;
        const char* src = p.seg_src[sg] + (lo - p.seg_begin[sg]);
        char* dst = p.fetch_dst + lo;
        const unsigned long long nv = (hi - lo) >> 4;
        unsigned long long v = ft;
        // 8 independent 128-bit peer loads in flight per thread: 128 threads x 8 x 16 B = 16 KiB per
        // CTA, ~2.4 MB chip-wide, which covers the ~2.5 us NVLink round trip at full link rate.
        for (; v + 7 * kFetchThreads < nv; v += 8 * kFetchThreads) {
          uint4 x[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) x[u] = ld_sys_v4(src + ((v + u * kFetchThreads) << 4));
#pragma unroll
          for (int u = 0; u < 8; ++u) *reinterpret_cast<uint4*>(dst + ((v + u * kFetchThreads) << 4)) = x[u];
        }
        for (; v < nv; v += kFetchThreads)
          *reinterpret_cast<uint4*>(dst + (v << 4)) = ld_sys_v4(src + (v << 4));
      }
      __threadfence();
      named_bar_sync(3, kFetchThreads);
      if (ft == 0)
        asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p.fetch_flags + c), "r"(p.ready_epoch)
                     : "memory");
    }
  } else if (warp >= 2 && warp < 2 + EPI) {
    // ------------------------------------------------------------------ epilogue warps
    const int q = warp & 3;          // TMEM lane quadrant this warp may access
    const int chalf = (warp - 2) >> 2;  // which half of the tile's column chunks this warp owns
    constexpr int kChunksPerWarp = BN / (8 * EPI);  // 32-column chunks per warp (EPI/4 warps share a quadrant)
    constexpr int kGroup = kChunksPerWarp >= 2 ? 2 : 1;  // chunks staged together (64 or 32 columns)
    constexpr int kPieces = kGroup * 4;                   // 16-byte pieces per staged row
    uint8_t* stage = s_out + (warp - 2) * (32 * 128);
    // Staged path: bf16 output, plain overwrite.  Rows go through shared memory so that one warp
    // store instruction writes kPieces*16 contiguous bytes of 32/kPieces rows (full cache lines)
    // instead of 16 bytes of 32 different rows.
    const bool can_stage = !p.out_fp32 && !p.accumulate;
    int it = 0;
    for (int tile = unit; tile < num_tiles; tile += nunits, ++it) {
      const int m0 = (tile % num_m) * (kBM * CG) + static_cast<int>(crank) * kBM;
      int nt = tile / num_m + p.n_rot;
      if (nt >= num_n) nt -= num_n;
      const int n0 = nt * BN;
      const int as = it & 1;
      const uint32_t aph = (it >> 1) & 1;
      mbar_wait(&tmem_full[as], aph);
      tc_fence_after();
      if (dbg != nullptr && it == 0 && threadIdx.x == 64) dbg[5] = clock64();  // accumulator ready
      const int row = m0 + q * 32 + lane;
      const bool row_ok = row < p.M;
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN;
#pragma unroll 1
      for (int g0 = 0; g0 < kChunksPerWarp; g0 += kGroup) {
        const int gcol0 = n0 + (chalf * kChunksPerWarp + g0) * 32;
        if (gcol0 >= p.N) break;  // warp-uniform
        const bool staged = can_stage && (gcol0 + kGroup * 32 <= p.N);
#pragma unroll 1   // one copy of the chunk code: every CTA runs it once or twice, so its
                   // instruction-cache footprint (cold fetches) is part of the latency
        for (int h = 0; h < kGroup; ++h) {
          const int c = chalf * kChunksPerWarp + g0 + h;
          const int col0 = n0 + c * 32;
          if (col0 >= p.N) break;  // warp-uniform
          uint32_t r[32];
          tmem_ld_32x32(t_row + c * 32, r);
          tmem_ld_wait();
          if (dbg != nullptr && it == 0 && threadIdx.x == 64 && g0 == 0) dbg[8 + h * 3] = clock64();
          // Everything below indexes v[] with compile-time constants only (fully unrolled, tails
          // handled by predicates) so the 32 accumulator values stay in registers.
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * p.alpha;
          const bool full_chunk = (col0 + 32 <= p.N);
          const int nvalid = full_chunk ? 32 : p.N - col0;
          const size_t out_off = static_cast<size_t>(row_ok ? row : 0) * p.ldd + col0;
          if (p.bias != nullptr)
            combine_row32(p.bias + col0, nvalid, full_chunk, v, [](float& a, float x) { a += x; });
          if (p.pre_out != nullptr && row_ok)
            store_row32_bf16(p.pre_out + out_off, nvalid, full_chunk, v, false);
          if (p.dact_in != nullptr) {
            if (row_ok) {
              const __nv_bfloat16* dsrc = p.dact_in + static_cast<size_t>(row) * p.ld_dact + col0;
              if (p.act == ACT_GELU_ERF)
                combine_row32(dsrc, nvalid, full_chunk, v, [](float& a, float x) { a *= gelu_erf_grad(x); });
              else if (p.act == ACT_RELU)
                combine_row32(dsrc, nvalid, full_chunk, v, [](float& a, float x) { a = x > 0.f ? a : 0.f; });
              else if (p.act == ACT_GELU_TANH)
                combine_row32(dsrc, nvalid, full_chunk, v, [](float& a, float x) { a *= gelu_tanh_grad(x); });
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
            combine_row32(p.residual + static_cast<size_t>(row) * p.ld_res + col0, nvalid, full_chunk, v,
                          [](float& a, float x) { a += x; });
          if (dbg != nullptr && it == 0 && threadIdx.x == 64 && g0 == 0) dbg[9 + h * 3] = clock64();
          if (staged) {
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              const int piece = h * 4 + g;
              *reinterpret_cast<uint4*>(stage + lane * 128 + ((piece ^ (lane & 7)) << 4)) =
                  make_uint4(pack_bf16x2(v[g * 8 + 0], v[g * 8 + 1]), pack_bf16x2(v[g * 8 + 2], v[g * 8 + 3]),
                             pack_bf16x2(v[g * 8 + 4], v[g * 8 + 5]), pack_bf16x2(v[g * 8 + 6], v[g * 8 + 7]));
            }
          } else if (row_ok) {
            if (p.out_fp32)
              store_row32_f32(reinterpret_cast<float*>(p.D) + out_off, nvalid, full_chunk, v, p.accumulate != 0);
            else
              store_row32_bf16(reinterpret_cast<__nv_bfloat16*>(p.D) + out_off, nvalid, full_chunk, v,
                               p.accumulate != 0);
          }
          if (dbg != nullptr && it == 0 && threadIdx.x == 64 && g0 == 0) dbg[10 + h * 3] = clock64();
        }
        if (staged) {
          __syncwarp();
          constexpr int kRowsPerInstr = 32 / kPieces;
          const int piece = lane % kPieces;
          __nv_bfloat16* dbase = reinterpret_cast<__nv_bfloat16*>(p.D) + gcol0 + piece * 8;
#pragma unroll
          for (int i = 0; i < kPieces; ++i) {
            const int rr = i * kRowsPerInstr + lane / kPieces;
            const uint4 x = *reinterpret_cast<const uint4*>(stage + rr * 128 + ((piece ^ (rr & 7)) << 4));
            const int grow = m0 + q * 32 + rr;
            if (grow < p.M) *reinterpret_cast<uint4*>(dbase + static_cast<size_t>(grow) * p.ldd) = x;
          }
          __syncwarp();
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (CG == 2 && crank != 0) mbar_arrive_cluster(mapa_u32(smem_u32(&tmem_empty[as]), 0));
        else mbar_arrive(&tmem_empty[as]);
      }
      if (dbg != nullptr && it == 0 && threadIdx.x == 64) dbg[6] = clock64();  // first tile stored
    }
  }

  tc_fence_before();
  // A pair leaves together: the leader's commits still arrive on the peer's barriers, and the
  // peer's epilogue still arrives on the leader's, until both have drained their last tile.
  if constexpr (CG == 2) cluster_sync_all(); else __syncthreads();
  if (dbg != nullptr && threadIdx.x == 0) dbg[7] = clock64();
  tc_fence_after();
  if (warp == 1) {
    if constexpr (CG == 2) tmem_dealloc_pair(tmem_base, Cfg::kTmemCols);
    else tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

// ---------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  });
  return fn;
}

// Builds a 2-D bf16 tensor map with 128B swizzle.  dim0 is the contiguous dimension.
int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t dim0, uint64_t dim1,
                      uint64_t ld_elems, uint32_t box0, uint32_t box1) {
  // cuTensorMapEncodeTiled is a driver call: it needs a current context on THIS thread.  PyTorch's
  // autograd worker threads may not have one yet (device 0 is "current" by default without its
  // primary context being bound), so bind it through the runtime once per thread.
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaSetDevice(dev);
    ctx_bound = true;
  }
  EncodeTiledFn fn = get_encode_fn();
  if (fn == nullptr) {
    fprintf(stderr, "[dtds] cuTensorMapEncodeTiled entry point not available\n");
    return -1;
  }
  cuuint64_t dims[2] = {dim0, dim1};
  cuuint64_t strides[1] = {ld_elems * 2};
  cuuint32_t box[2] = {box0, box1};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides,
                  box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    fprintf(stderr,
            "[dtds] cuTensorMapEncodeTiled failed (%d): base=%p dims=(%llu,%llu) ld=%llu "
            "box=(%u,%u)\n",
            (int)r, base, (unsigned long long)dim0, (unsigned long long)dim1,
            (unsigned long long)ld_elems, box0, box1);
    return -2;
  }
  return 0;
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (cached[dev] == 0) {
    int n = 0;
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    cached[dev] = n > 0 ? n : 148;
  }
  return cached[dev];
}

template <int BN, bool A_MN, bool B_MN, bool FETCH = false, int CG = 1>
static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p,
                       int max_ctas, cudaStream_t stream) {
  using Cfg = GemmCfg<BN, CG>;
  auto kern = gemm_bf16_kernel<BN, A_MN, B_MN, FETCH, CG>;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_set[dev & 63]) {
    DTDS_HOST_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         Cfg::kSmemBytes));
    attr_set[dev & 63] = true;
  }
  const int num_tiles = ((p.M + kBM * CG - 1) / (kBM * CG)) * ((p.N + BN - 1) / BN);
  int units = sm_count() / CG;
  if (max_ctas > 0 && max_ctas / CG < units) units = max_ctas / CG > 0 ? max_ctas / CG : 1;
  if (num_tiles < units && !FETCH) units = num_tiles;  // fetch mode keeps every SM pulling chunks
  DTDS_HOST_CHECK(launch_pdl_cluster(kern, dim3(units * CG), dim3(Cfg::kThreads + (FETCH ? kFetchThreads : 0)),
                                     Cfg::kSmemBytes, stream, static_cast<unsigned>(CG), tmA, tmB, p));
  DTDS_HOST_CHECK(cudaGetLastError());
  return 0;
}

// Picks BLOCK_N with a small cost model.  At the token counts of the reference's scripts
// (2048 rows per rank) most GEMMs have fewer than two waves of tiles and are bound by the
// L2 -> shared-memory fill rate rather than by the tensor pipe, so "enough tiles for every SM" is
// the wrong criterion: a 128-wide tile moves 2/3 of the bytes of two 64-wide ones.
static int pick_block_n(int M, int N, int K, int requested) {
  if (requested == 64 || requested == 128 || requested == 256) return requested;
  if (requested == DTDS_GEMM_PAIR + 128 || requested == DTDS_GEMM_PAIR + 256) return requested;
  const int sms = sm_count();
  const int num_m = (M + kBM - 1) / kBM;
  const int num_kb = (K + kBK - 1) / kBK;
  const int cands[3] = {256, 128, 64};
  double best = 1e30;
  int best_bn = 64;
  for (int i = 0; i < 3; ++i) {
    const int bn = cands[i];
    if (bn > 64 && N <= bn / 2) continue;
    const int tiles = num_m * ((N + bn - 1) / bn);
    const int rounds = (tiles + sms - 1) / sms;
    const int active = tiles < sms ? tiles : sms;
    double bw = 6300.0 / active;          // chip-wide L2 fill rate shared by the active SMs
    if (bw > 72.0) bw = 72.0;             // per-SM ceiling (bytes / clock)
    const double fill = (16384.0 + bn * 128.0) / bw;
    const double mma = 2.0 * bn;          // 4 MMAs of K=16 per k-block, bn/2 clocks each
    const double per_kb = fill > mma ? fill : mma;
    const double epi = 40.0 * bn / 8 + 600;  // drain of one tile, overlapped except for the last
    const double t = rounds * (num_kb * per_kb) + epi + 1500;
    if (t < best) { best = t; best_bn = bn; }
  }
  return best_bn;
}

}  // namespace dtds

using namespace dtds;

static long long* g_gemm_dbg = nullptr;
// Development aid: when set, every GEMM CTA writes 8 clock64() stamps (start, setup done, loads
// issued, first operands landed, MMAs issued, accumulator ready, first tile stored, end) to
// buf[blockIdx.x * 8 ...].  Pass NULL to switch it off.
extern "C" int dtds_gemm_set_debug_buffer(long long* buf) {
  g_gemm_dbg = buf;
  return 0;
}

struct FetchSpec {
  const unsigned long long* seg_src;
  const unsigned long long* seg_begin;  // nseg + 1 entries
  int nseg;
  unsigned chunk_bytes;
  int flags_on_k;
  unsigned long long local_begin;   // byte offset of this rank's own shard inside B
};

static int gemm_impl(const void* A, const void* B, void* D, const void* bias, const void* residual,
                     void* pre_out, const void* dact_in, int M, int N, int K, int lda, int ldb, int ldd,
                     int ld_res, int ld_dact, int a_mn_major, int b_mn_major, int act, int out_fp32,
                     int accumulate, float alpha, int block_n, int max_ctas, const void* ready_flags,
                     unsigned ready_epoch, int rows_per_flag, const FetchSpec* fetch,
                     cudaStream_t stream) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15) ||
      (reinterpret_cast<uintptr_t>(D) & 15) || (lda % 8) || (ldb % 8) ||
      (ldd % (out_fp32 ? 4 : 8))) {
    fprintf(stderr,
            "[dtds] gemm: operands must be 16-byte aligned with leading dimensions that are "
            "multiples of 8 elements (lda=%d ldb=%d ldd=%d)\n",
            lda, ldb, ldd);
    return -3;
  }
  const int bn = pick_block_n(M, N, K, block_n);
  CUtensorMap tmA, tmB;
  int rc;
  if (!a_mn_major) rc = make_tmap_2d_bf16(&tmA, A, K, M, lda, kBK, kBM);
  else             rc = make_tmap_2d_bf16(&tmA, A, M, K, lda, 64, kBK);
  if (rc) return rc;
  const bool pair = bn >= DTDS_GEMM_PAIR;
  const int bn_tile = pair ? bn - DTDS_GEMM_PAIR : bn;
  if (pair && fetch != nullptr) {
    fprintf(stderr, "[dtds] gemm(all-gather): CTA-pair tiles are not available in this mode\n");
    return -8;
  }
  if (!b_mn_major) rc = make_tmap_2d_bf16(&tmB, B, K, N, ldb, kBK, pair ? bn_tile / 2 : bn_tile);
  else             rc = make_tmap_2d_bf16(&tmB, B, N, K, ldb, 64, kBK);
  if (rc) return rc;

  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.D = D;
  p.bias = static_cast<const __nv_bfloat16*>(bias);
  p.residual = static_cast<const __nv_bfloat16*>(residual);
  p.pre_out = static_cast<__nv_bfloat16*>(pre_out);
  p.dact_in = static_cast<const __nv_bfloat16*>(dact_in);
  p.ready_flags = static_cast<const uint32_t*>(ready_flags);
  p.ready_epoch = ready_epoch;
  p.rows_per_flag = rows_per_flag > 0 ? rows_per_flag : N;
  p.M = M; p.N = N; p.K = K;
  p.ldd = ldd; p.ld_res = ld_res > 0 ? ld_res : ldd; p.ld_dact = ld_dact > 0 ? ld_dact : ldd;
  p.act = act; p.out_fp32 = out_fp32; p.accumulate = accumulate; p.alpha = alpha;
  p.dbg = g_gemm_dbg;
  if (fetch != nullptr) {
    if (fetch->nseg < 1 || fetch->nseg > kMaxFetchSegs || (fetch->chunk_bytes % 16) || ready_flags == nullptr) {
      fprintf(stderr, "[dtds] gemm(all-gather): need 1..%d segments, a 16-byte chunk size and flags\n",
              kMaxFetchSegs);
      return -5;
    }
    for (int i = 0; i < fetch->nseg; ++i) p.seg_src[i] = reinterpret_cast<const char*>(fetch->seg_src[i]);
    for (int i = 0; i <= fetch->nseg; ++i) p.seg_begin[i] = fetch->seg_begin[i];
    p.fetch_nseg = fetch->nseg;
    p.fetch_dst = static_cast<char*>(const_cast<void*>(B));
    p.fetch_flags = static_cast<uint32_t*>(const_cast<void*>(ready_flags));
    p.fetch_total = fetch->seg_begin[fetch->nseg];
    p.fetch_chunk = fetch->chunk_bytes;
    p.flags_on_k = fetch->flags_on_k;
    const unsigned long long row_bytes = static_cast<unsigned long long>(ldb) * 2;
    const unsigned long long nchunks = (p.fetch_total + p.fetch_chunk - 1) / p.fetch_chunk;
    p.chunk_rot = static_cast<unsigned>((fetch->local_begin / p.fetch_chunk) % (nchunks ? nchunks : 1));
    if (fetch->flags_on_k) {
      const int num_kb = (K + kBK - 1) / kBK;
      p.k_rot = static_cast<int>((fetch->local_begin + kBK * row_bytes - 1) / (kBK * row_bytes)) % num_kb;
    } else {
      const int num_n = (N + bn - 1) / bn;
      p.n_rot = static_cast<int>((fetch->local_begin + bn * row_bytes - 1) / (bn * row_bytes)) % num_n;
    }
  }
  const bool f = fetch != nullptr;

  if (pair) {
#define DTDS_DISPATCH_PAIR(BN_)                                                                                 \
  if (bn_tile == BN_) {                                                                                         \
    if (!a_mn_major && !b_mn_major) return launch_gemm<BN_, false, false, false, 2>(tmA, tmB, p, max_ctas, stream); \
    if (!a_mn_major && b_mn_major) return launch_gemm<BN_, false, true, false, 2>(tmA, tmB, p, max_ctas, stream);   \
    if (a_mn_major && !b_mn_major) return launch_gemm<BN_, true, false, false, 2>(tmA, tmB, p, max_ctas, stream);   \
    return launch_gemm<BN_, true, true, false, 2>(tmA, tmB, p, max_ctas, stream);                               \
  }
    DTDS_DISPATCH_PAIR(128)
    DTDS_DISPATCH_PAIR(256)
#undef DTDS_DISPATCH_PAIR
    return -4;
  }

#define DTDS_DISPATCH(BN_)                                                                          \
  if (bn == BN_) {                                                                                  \
    if (!a_mn_major && !b_mn_major)                                                                 \
      return f ? launch_gemm<BN_, false, false, true>(tmA, tmB, p, max_ctas, stream)                \
               : launch_gemm<BN_, false, false, false>(tmA, tmB, p, max_ctas, stream);              \
    if (!a_mn_major && b_mn_major)                                                                  \
      return f ? launch_gemm<BN_, false, true, true>(tmA, tmB, p, max_ctas, stream)                 \
               : launch_gemm<BN_, false, true, false>(tmA, tmB, p, max_ctas, stream);               \
    if (f) { fprintf(stderr, "[dtds] gemm(all-gather): operand A must be K-major\n"); return -6; } \
    if (a_mn_major && !b_mn_major) return launch_gemm<BN_, true, false, false>(tmA, tmB, p, max_ctas, stream); \
    return launch_gemm<BN_, true, true, false>(tmA, tmB, p, max_ctas, stream);                      \
  }
  DTDS_DISPATCH(64)
  DTDS_DISPATCH(128)
  DTDS_DISPATCH(256)
#undef DTDS_DISPATCH
  return -4;
}

extern "C" int dtds_gemm_bf16_ex(const void* A, const void* B, void* D, const void* bias,
                                 const void* residual, void* pre_out, const void* dact_in, int M,
                                 int N, int K, int lda, int ldb, int ldd, int ld_res, int ld_dact,
                                 int a_mn_major, int b_mn_major, int act, int out_fp32,
                                 int accumulate, float alpha, int block_n, int max_ctas,
                                 const void* ready_flags, unsigned ready_epoch, int rows_per_flag,
                                 cudaStream_t stream) {
  return gemm_impl(A, B, D, bias, residual, pre_out, dact_in, M, N, K, lda, ldb, ldd, ld_res, ld_dact,
                   a_mn_major, b_mn_major, act, out_fp32, accumulate, alpha, block_n, max_ctas,
                   ready_flags, ready_epoch, rows_per_flag, nullptr, stream);
}

// ZeRO-3 parameter all-gather fused with the consuming GEMM.  B (the weight, [N,K] K-major for
// forward, consumed MN-major for dgrad) is NOT complete in memory when the kernel starts: it is
// described by `nseg` byte segments (seg_begin[i] .. seg_begin[i+1]) whose sources are the
// owners' shard buffers (peer-mapped addresses).  Fetch warps inside the GEMM kernel copy them
// into `B_local` chunk by chunk (`chunk_bytes`, a whole number of B rows) and raise flags[chunk]
// to `epoch`; the TMA producer of every CTA waits for the flags of the rows it needs, so tiles
// start as soon as their part of the weight has arrived and the NVLink transfer overlaps the
// tensor-core work tile by tile.  `flags` must hold one uint32 per chunk and `epoch` must be
// larger than any value used with these flags before (monotonic; no reset between calls).
// `local_begin` = byte offset inside the weight of the part this rank owns itself: tiles, k-blocks
// and chunks are visited cyclically starting there (no transfer needed for the first ones).
extern "C" int dtds_gemm_bf16_allgather(const void* A, void* B_local, void* D, const void* bias,
                                        const void* residual, int M, int N, int K, int lda, int ldb,
                                        int ldd, int ld_res, int b_mn_major, int act, int block_n,
                                        const unsigned long long* seg_src,
                                        const unsigned long long* seg_begin, int nseg,
                                        unsigned chunk_bytes, void* flags, unsigned epoch,
                                        unsigned long long local_begin, cudaStream_t stream) {
  const size_t row_bytes = static_cast<size_t>(ldb) * 2;
  if (chunk_bytes == 0 || chunk_bytes % row_bytes != 0) {
    fprintf(stderr, "[dtds] gemm(all-gather): chunk_bytes must be a whole number of B rows\n");
    return -7;
  }
  FetchSpec fs{seg_src, seg_begin, nseg, chunk_bytes, b_mn_major ? 1 : 0, local_begin};
  const int rows_per_flag = static_cast<int>(chunk_bytes / row_bytes);
  return gemm_impl(A, B_local, D, bias, residual, nullptr, nullptr, M, N, K, lda, ldb, ldd, ld_res, 0, 0,
                   b_mn_major, act, 0, 0, 1.0f, block_n, 0, flags, epoch, rows_per_flag, &fs, stream);
}

extern "C" int dtds_gemm_bf16(const void* A, const void* B, void* D, const void* bias, int M, int N,
                              int K, int lda, int ldb, int ldd, int a_mn_major, int b_mn_major,
                              int act, int out_fp32, cudaStream_t stream) {
  return dtds_gemm_bf16_ex(A, B, D, bias, nullptr, nullptr, nullptr, M, N, K, lda, ldb, ldd, 0, 0,
                           a_mn_major, b_mn_major, act, out_fp32, 0, 1.0f, 0, 0, nullptr, 0, 0,
                           stream);
}
