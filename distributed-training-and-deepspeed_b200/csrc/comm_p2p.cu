// Symmetric-memory runtime and in-kernel collectives over NVLink peer memory.
//
// Replaces ProcessGroupNCCL / NCCL ring kernels behind dist.all_reduce (pytorch_allreduce.py:31),
// the DDP Reducer's bucketed all-reduce (data_parallel_training.py:44) and the DeepSpeed ZeRO
// reduce-to-owner / all-gather / FusedAdam sequence (zero_dp_training.py:46-50,88-90) —
// SURVEY.md N1-N4, K13-K16.
//
// Model: every rank allocates identically sized buffers ("symmetric"), exports them with CUDA
// IPC and maps every peer's copy; kernels receive a device table of the n peer addresses and of
// the n signal pads.  Through NVSwitch every peer is at full NVLink bandwidth, so there is no
// ring: each rank pulls its 1/n slice from all peers at once (reduce-scatter phase) and pushes
// the result to all peers at once (all-gather phase).  Cross-GPU ordering uses epoch flag words in
// the signal pads with release/acquire at system scope (one remote store per peer per barrier).
//
// Kernels here: barrier, one-shot all-reduce (latency-bound sizes), two-shot all-reduce
// (bandwidth-bound sizes; scale and dtype conversion fused), reduce-scatter, all-gather,
// broadcast, and the ZeRO step kernel: reduce-scatter + Adam on the owned partition +
// all-gather of the updated bf16 parameters in ONE launch.
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstring>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "dtds_b200.h"
#include "internal.h"

namespace dtds {

constexpr int kMaxRanks = 16;
constexpr int kCommThreads = 512;
constexpr int kMaxCommBlocks = 128;  // signal pad holds kMaxCommBlocks * kMaxRanks flags per use

struct PeerTable {
  void* data[kMaxRanks];
  uint32_t* signal[kMaxRanks];
  int rank, world;
};

// ------------------------------------------------------------------------- signalling
// Signal pad layout (per rank, zero-initialised once at allocation, never reset):
//   flags  [kSlots][kMaxCommBlocks][kMaxRanks] u32 : flags[slot][block][src] is written only by rank
//                                                    `src` (block `block` of its kernel)
//   epochs [kSlots][kMaxCommBlocks]            u32 : this rank's own arrival count per (slot, block)
// A barrier is ONE fire-and-forget remote store per peer plus a spin on local memory: the block
// bumps its private epoch e, stores e into its flag word on every peer (st.release.sys: all earlier
// writes of the block are visible first) and waits until every peer's word in its own pad has
// reached e (ld.acquire.sys).  Epochs only grow, so nothing is consumed or reset and no remote
// atomic round trip is paid (the first version CAS-ed every flag 0->1 remotely and 1->0 locally:
// two NVLink round trips per barrier).  All ranks issue the same collectives with the same grid
// sizes in the same order on the streams that share a pad, so block b's epochs advance in lockstep
// on every rank; comparison is wrap-safe (signed difference).
constexpr int kSlots = 4;
constexpr size_t kFlagWords = static_cast<size_t>(kSlots) * kMaxCommBlocks * kMaxRanks;
constexpr size_t kEpochWords = static_cast<size_t>(kSlots) * kMaxCommBlocks;

__device__ __forceinline__ void st_release_sys(uint32_t* addr, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* addr) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
  return v;
}
// Block-level barrier across all ranks: block b of every rank meets block b of every other rank.
// `slot` selects an independent flag set (0 = entry barrier, 1 = exit barrier, ...).
__device__ __forceinline__ void cross_gpu_barrier(const PeerTable& pt, int slot) {
  __syncthreads();   // every thread's earlier global writes happen-before the releasing stores below
  if (threadIdx.x < pt.world) {
    const int peer = threadIdx.x;
    const size_t cell = static_cast<size_t>(slot) * kMaxCommBlocks + blockIdx.x;
    uint32_t* my_epoch = pt.signal[pt.rank] + kFlagWords + cell;
    const uint32_t e = *reinterpret_cast<volatile uint32_t*>(my_epoch) + 1;
    __syncwarp((1u << pt.world) - 1u);         // everyone has read the old epoch
    if (peer == 0) *reinterpret_cast<volatile uint32_t*>(my_epoch) = e;
    __threadfence_system();                    // cumulativity: the block's writes, then the flag
    st_release_sys(pt.signal[peer] + cell * kMaxRanks + pt.rank, e);
    const uint32_t* mine = pt.signal[pt.rank] + cell * kMaxRanks + peer;
    uint32_t spins = 0;
    unsigned long long t_start = 0;
    while (static_cast<int32_t>(ld_acquire_sys(mine) - e) < 0) {
      // bounded by wall time (60 s), not by iteration count: a peer that died must surface as an
      // error on the survivors soon, a peer that is merely late (host jitter) must not
      if ((++spins & 0x3FF) == 0) {
        unsigned long long now;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (t_start == 0) t_start = now;
      if (now - t_start > 60000000000ull) {
        printf("[dtds] comm: barrier slot %d block %d: rank %d never saw epoch %u from rank %d\n", slot,
               blockIdx.x, pt.rank, e, peer);
        __trap();
      }
      }
    }
  }
  __syncthreads();
}

// 128-bit load of (possibly remote) memory that another GPU wrote earlier: a strong load at system
// scope, so it can never be served from a stale line of this SM's L1.
__device__ __forceinline__ uint4 ld_peer_v4(const void* p) {
  uint4 r;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}

// ------------------------------------------------------------------------- typed access
template <typename T> struct Vec;  // 16-byte vector of T with fp32 (or int64) accumulation
template <> struct Vec<__nv_bfloat16> {
  static constexpr int N = 8;
  using Acc = float;
  __device__ static void add(const uint4& u, Acc (&a)[N]) {
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) { a[2 * i] += bf16_lo(w[i]); a[2 * i + 1] += bf16_hi(w[i]); }
  }
  __device__ static uint4 pack(const Acc (&a)[N]) {
    return make_uint4(pack_bf16x2(a[0], a[1]), pack_bf16x2(a[2], a[3]), pack_bf16x2(a[4], a[5]),
                      pack_bf16x2(a[6], a[7]));
  }
};
template <> struct Vec<float> {
  static constexpr int N = 4;
  using Acc = float;
  __device__ static void add(const uint4& u, Acc (&a)[N]) {
    a[0] += __uint_as_float(u.x); a[1] += __uint_as_float(u.y);
    a[2] += __uint_as_float(u.z); a[3] += __uint_as_float(u.w);
  }
  __device__ static uint4 pack(const Acc (&a)[N]) {
    return make_uint4(__float_as_uint(a[0]), __float_as_uint(a[1]), __float_as_uint(a[2]),
                      __float_as_uint(a[3]));
  }
};
template <> struct Vec<long long> {
  static constexpr int N = 2;
  using Acc = long long;
  __device__ static void add(const uint4& u, Acc (&a)[N]) {
    a[0] += static_cast<long long>((static_cast<unsigned long long>(u.y) << 32) | u.x);
    a[1] += static_cast<long long>((static_cast<unsigned long long>(u.w) << 32) | u.z);
  }
  __device__ static uint4 pack(const Acc (&a)[N]) {
    const unsigned long long x = static_cast<unsigned long long>(a[0]);
    const unsigned long long y = static_cast<unsigned long long>(a[1]);
    return make_uint4(static_cast<uint32_t>(x), static_cast<uint32_t>(x >> 32),
                      static_cast<uint32_t>(y), static_cast<uint32_t>(y >> 32));
  }
};

// Sum over all ranks of the 16-byte vector at byte offset `off` of each rank's buffer.  All peer
// loads are issued before any is consumed (up to 8 outstanding 128-bit NVLink reads per thread);
// peers are visited starting after our own rank so the n ranks spread over the n links.
template <typename T>
__device__ __forceinline__ void reduce_vec(const PeerTable& pt, size_t off,
                                           typename Vec<T>::Acc (&acc)[Vec<T>::N]) {
#pragma unroll
  for (int i = 0; i < Vec<T>::N; ++i) acc[i] = 0;
  for (int base = 0; base < pt.world; base += 8) {
    uint4 raw[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (base + k < pt.world) {
        int peer = pt.rank + base + k;
        if (peer >= pt.world) peer -= pt.world;
        raw[k] = ld_peer_v4(static_cast<const char*>(pt.data[peer]) + off);
      }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (base + k < pt.world) Vec<T>::add(raw[k], acc);
  }
}
// Two vectors at once: all 2 x world peer loads are in flight before the first add.  With 128
// blocks x 512 threads this keeps several MB outstanding, enough to cover NVLink latency
// (~2 us) at full link bandwidth.
template <typename T>
__device__ __forceinline__ void reduce_vec2(const PeerTable& pt, size_t off0, size_t off1,
                                            typename Vec<T>::Acc (&a0)[Vec<T>::N],
                                            typename Vec<T>::Acc (&a1)[Vec<T>::N]) {
#pragma unroll
  for (int i = 0; i < Vec<T>::N; ++i) { a0[i] = 0; a1[i] = 0; }
  for (int base = 0; base < pt.world; base += 8) {
    uint4 r0[8], r1[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (base + k < pt.world) {
        int peer = pt.rank + base + k;
        if (peer >= pt.world) peer -= pt.world;
        const char* b = static_cast<const char*>(pt.data[peer]);
        r0[k] = ld_peer_v4(b + off0);
        r1[k] = ld_peer_v4(b + off1);
      }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (base + k < pt.world) { Vec<T>::add(r0[k], a0); Vec<T>::add(r1[k], a1); }
  }
}
template <typename T>
__device__ __forceinline__ void scale_vec(typename Vec<T>::Acc (&acc)[Vec<T>::N], float scale) {
  if (scale != 1.f) {
#pragma unroll
    for (int i = 0; i < Vec<T>::N; ++i) acc[i] = static_cast<typename Vec<T>::Acc>(acc[i] * scale);
  }
}

// ------------------------------------------------------------------------- all-reduce kernels
// One-shot: every rank reads everything from everyone.  n reads of the whole message per rank, one
// barrier pair; best below a few hundred KB where latency dominates.  Result goes to `out`
// (may alias this rank's symmetric buffer only if out == data[rank], handled by the exit barrier).
template <typename T>
__global__ void __launch_bounds__(kCommThreads)
allreduce_oneshot_kernel(PeerTable pt, void* out, size_t off_bytes, size_t nvec, float scale) {
  cross_gpu_barrier(pt, 0);
  for (size_t v = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < nvec;
       v += static_cast<size_t>(gridDim.x) * blockDim.x) {
    typename Vec<T>::Acc acc[Vec<T>::N];
    reduce_vec<T>(pt, off_bytes + v * 16, acc);
    scale_vec<T>(acc, scale);
    *reinterpret_cast<uint4*>(static_cast<char*>(out) + v * 16) = Vec<T>::pack(acc);
  }
  cross_gpu_barrier(pt, 1);  // nobody may overwrite its input while a peer is still reading it
}

// Two-shot, in place on the symmetric buffer: rank r reduces slice r (pulling from all peers) and
// pushes the reduced slice into every peer's buffer.  Per rank: (n-1)/n of the message in over
// NVLink and (n-1)/n out.  `scale` (e.g. 1/world for gradient averaging) is applied once.
// ILP = 2: standalone / bandwidth-critical (128 registers, one block per SM).  ILP = 1: capped at
// 64 registers so a block fits on an SM next to a resident persistent GEMM CTA — the variant the
// DDP engine uses for buckets reduced while backward is still running.
template <typename T, int ILP>
__global__ void __launch_bounds__(kCommThreads, ILP == 1 ? 2 : 1)
allreduce_twoshot_kernel(PeerTable pt, size_t off_bytes, size_t nvec, float scale) {
  cross_gpu_barrier(pt, 0);
  const size_t per = (nvec + pt.world - 1) / pt.world;
  const size_t v0 = per * pt.rank;
  const size_t v1 = min(nvec, v0 + per);
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  auto push = [&](size_t v, typename Vec<T>::Acc (&acc)[Vec<T>::N]) {
    scale_vec<T>(acc, scale);
    const uint4 packed = Vec<T>::pack(acc);
#pragma unroll 1
    for (int k = 0; k < pt.world; ++k) {
      const int peer = (pt.rank + k) % pt.world;
      *reinterpret_cast<uint4*>(static_cast<char*>(pt.data[peer]) + off_bytes + v * 16) = packed;
    }
  };
  size_t v = v0 + static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if constexpr (ILP == 2) {
    for (; v + stride < v1; v += 2 * stride) {
      typename Vec<T>::Acc a0[Vec<T>::N], a1[Vec<T>::N];
      reduce_vec2<T>(pt, off_bytes + v * 16, off_bytes + (v + stride) * 16, a0, a1);
      push(v, a0);
      push(v + stride, a1);
    }
  } else {
    for (; v + stride < v1; v += stride) {
      typename Vec<T>::Acc acc[Vec<T>::N];
      reduce_vec<T>(pt, off_bytes + v * 16, acc);
      push(v, acc);
    }
  }
  if (v < v1) {
    typename Vec<T>::Acc acc[Vec<T>::N];
    reduce_vec<T>(pt, off_bytes + v * 16, acc);
    push(v, acc);
  }
  cross_gpu_barrier(pt, 1);
}

// Reduce-scatter: out[0:per) = scale * sum over ranks of slice `rank` of the symmetric buffer.
template <typename T>
__global__ void __launch_bounds__(kCommThreads)
reduce_scatter_kernel(PeerTable pt, void* out, size_t off_bytes, size_t nvec_per_rank, float scale) {
  cross_gpu_barrier(pt, 0);
  const size_t v0 = nvec_per_rank * pt.rank;
  for (size_t v = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < nvec_per_rank;
       v += static_cast<size_t>(gridDim.x) * blockDim.x) {
    typename Vec<T>::Acc acc[Vec<T>::N];
    reduce_vec<T>(pt, off_bytes + (v0 + v) * 16, acc);
    scale_vec<T>(acc, scale);
    *reinterpret_cast<uint4*>(static_cast<char*>(out) + v * 16) = Vec<T>::pack(acc);
  }
  cross_gpu_barrier(pt, 1);
}

// All-gather (push): rank r copies its `nvec_per_rank` vectors from `in` into slot r of every
// peer's symmetric buffer.
__global__ void __launch_bounds__(kCommThreads)
all_gather_kernel(PeerTable pt, const void* in, size_t off_bytes, size_t nvec_per_rank) {
  cross_gpu_barrier(pt, 0);  // peers are done reading the previous contents
  const size_t v0 = nvec_per_rank * pt.rank;
  for (size_t v = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < nvec_per_rank;
       v += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const uint4 x = *reinterpret_cast<const uint4*>(static_cast<const char*>(in) + v * 16);
#pragma unroll 1
    for (int k = 0; k < pt.world; ++k) {
      const int peer = (pt.rank + k) % pt.world;
      *reinterpret_cast<uint4*>(static_cast<char*>(pt.data[peer]) + off_bytes + (v0 + v) * 16) = x;
    }
  }
  cross_gpu_barrier(pt, 1);
}

// Broadcast (pull): every rank copies [off, off + nvec*16) of root's buffer into its own.
__global__ void __launch_bounds__(kCommThreads)
broadcast_kernel(PeerTable pt, int root, size_t off_bytes, size_t nvec) {
  cross_gpu_barrier(pt, 0);
  if (pt.rank != root) {
    const char* src = static_cast<const char*>(pt.data[root]) + off_bytes;
    char* dst = static_cast<char*>(pt.data[pt.rank]) + off_bytes;
    for (size_t v = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < nvec;
         v += static_cast<size_t>(gridDim.x) * blockDim.x)
      *reinterpret_cast<uint4*>(dst + v * 16) = ld_peer_v4(src + v * 16);
  }
  cross_gpu_barrier(pt, 1);
}

__global__ void __launch_bounds__(kCommThreads) barrier_kernel(PeerTable pt) {
  cross_gpu_barrier(pt, 0);
}

__device__ __forceinline__ uint4 mm_ld_reduce_bf16(const char* addr);
__device__ __forceinline__ void mm_st(char* addr, const uint4& r);

// ------------------------------------------------------------------------- ZeRO step kernel
// One launch = gradient reduce-scatter + Adam on the owned partition + parameter all-gather.
//   grads : symmetric bf16 buffer (every rank holds its full local gradient); rank r owns elements
//           [r*per, (r+1)*per) of the range [off, off+n).
//   master/m/v : this rank's fp32 partition state (length `per`), private.
//   params: symmetric bf16 buffer; the updated owned slice is pushed into every peer's copy
//           (push_params = 1, ZeRO-1/2 where parameters are replicated) or only written locally
//           into `param_shard` (ZeRO-3 where parameters stay sharded).
struct ZeroStepArgs {
  PeerTable grads;   // data[] = gradient buffers, signal[] = pads
  void* params[kMaxRanks];
  size_t goff_elems, poff_elems, per;  // element offsets into grad / param buffers, partition size
  float* master; float* m; float* v;
  __nv_bfloat16* param_shard;          // optional private destination for the updated slice
  const int* step;
  float lr, beta1, beta2, eps, weight_decay, grad_scale;
  int adamw_mode, bias_correction, push_params;
  // NVLS: multicast addresses of the gradient / parameter buffers (0 = peer loads / peer stores).
  // With them the owned gradient slice is reduced inside the switch (one multimem.ld_reduce per
  // 16 bytes instead of world peer loads) and the updated slice is replicated by one multimem.st.
  const char* grads_mc;
  char* params_mc;
};

__global__ void __launch_bounds__(kCommThreads) zero_rs_adam_ag_kernel(ZeroStepArgs a) {
  const PeerTable& pt = a.grads;
  cross_gpu_barrier(pt, 0);  // every rank's local gradients for this range are complete
  const int t = *a.step + 1;
  float c1 = 1.f, c2 = 1.f;
  if (a.bias_correction) {
    c1 = 1.f - __powf(a.beta1, static_cast<float>(t));
    c2 = 1.f - __powf(a.beta2, static_cast<float>(t));
  }
  const float step_size = a.lr / c1, inv_sqrt_c2 = rsqrtf(c2);
  const size_t nvec = a.per / 8;  // partitions are padded to multiples of 8 elements
  const size_t g_base = (a.goff_elems + a.per * pt.rank) * 2;
  const size_t p_base = (a.poff_elems + a.per * pt.rank) * 2;
  for (size_t vi = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; vi < nvec;
       vi += static_cast<size_t>(gridDim.x) * blockDim.x) {
    float g[8];
    if (a.grads_mc != nullptr) {
      const uint4 r = mm_ld_reduce_bf16(a.grads_mc + g_base + vi * 16);
      const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) { g[2 * i] = bf16_lo(w[i]); g[2 * i + 1] = bf16_hi(w[i]); }
    } else {
      reduce_vec<__nv_bfloat16>(pt, g_base + vi * 16, g);
    }
    float p[8], mm[8], vv[8];
    const size_t e = vi * 8;
    *reinterpret_cast<float4*>(p) = *reinterpret_cast<const float4*>(a.master + e);
    *reinterpret_cast<float4*>(p + 4) = *reinterpret_cast<const float4*>(a.master + e + 4);
    *reinterpret_cast<float4*>(mm) = *reinterpret_cast<const float4*>(a.m + e);
    *reinterpret_cast<float4*>(mm + 4) = *reinterpret_cast<const float4*>(a.m + e + 4);
    *reinterpret_cast<float4*>(vv) = *reinterpret_cast<const float4*>(a.v + e);
    *reinterpret_cast<float4*>(vv + 4) = *reinterpret_cast<const float4*>(a.v + e + 4);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float gk = g[k] * a.grad_scale;
      if (!a.adamw_mode && a.weight_decay != 0.f) gk += a.weight_decay * p[k];
      mm[k] = a.beta1 * mm[k] + (1.f - a.beta1) * gk;
      vv[k] = a.beta2 * vv[k] + (1.f - a.beta2) * gk * gk;
      const float denom = sqrtf(vv[k]) * inv_sqrt_c2 + a.eps;
      if (a.adamw_mode && a.weight_decay != 0.f) p[k] *= (1.f - a.lr * a.weight_decay);
      p[k] -= step_size * (mm[k] / denom);
    }
    *reinterpret_cast<float4*>(a.master + e) = *reinterpret_cast<float4*>(p);
    *reinterpret_cast<float4*>(a.master + e + 4) = *reinterpret_cast<float4*>(p + 4);
    *reinterpret_cast<float4*>(a.m + e) = *reinterpret_cast<float4*>(mm);
    *reinterpret_cast<float4*>(a.m + e + 4) = *reinterpret_cast<float4*>(mm + 4);
    *reinterpret_cast<float4*>(a.v + e) = *reinterpret_cast<float4*>(vv);
    *reinterpret_cast<float4*>(a.v + e + 4) = *reinterpret_cast<float4*>(vv + 4);
    const uint4 lp = make_uint4(pack_bf16x2(p[0], p[1]), pack_bf16x2(p[2], p[3]),
                                pack_bf16x2(p[4], p[5]), pack_bf16x2(p[6], p[7]));
    if (a.param_shard != nullptr) *reinterpret_cast<uint4*>(a.param_shard + e) = lp;
    if (a.push_params && a.params_mc != nullptr) {
      mm_st(a.params_mc + p_base + vi * 16, lp);
    } else if (a.push_params) {
#pragma unroll 1
      for (int k = 0; k < pt.world; ++k) {
        const int peer = (pt.rank + k) % pt.world;
        *reinterpret_cast<uint4*>(static_cast<char*>(a.params[peer]) + p_base + vi * 16) = lp;
      }
    }
  }
  cross_gpu_barrier(pt, 1);  // all pushes have landed; gradient buffers may be reused
}

// ------------------------------------------------------------------------- host runtime
struct SymmAlloc { void* ptr; size_t bytes; };
static std::mutex g_mu;
static std::vector<SymmAlloc> g_allocs;

static int comm_blocks(size_t nvec, int requested) {
  int b = requested > 0 ? requested : 96;
  if (b > kMaxCommBlocks) b = kMaxCommBlocks;
  const size_t need = (nvec + kCommThreads - 1) / kCommThreads;
  if (need < static_cast<size_t>(b)) b = need > 0 ? static_cast<int>(need) : 1;
  return b;
}

static int fill_table(PeerTable* pt, const unsigned long long* data_ptrs,
                      const unsigned long long* signal_ptrs, int rank, int world) {
  if (world < 1 || world > kMaxRanks || rank < 0 || rank >= world) {
    fprintf(stderr, "[dtds] comm: invalid rank/world %d/%d (max %d)\n", rank, world, kMaxRanks);
    return -1;
  }
  memset(pt, 0, sizeof(*pt));
  for (int i = 0; i < world; ++i) {
    pt->data[i] = reinterpret_cast<void*>(data_ptrs[i]);
    pt->signal[i] = reinterpret_cast<uint32_t*>(signal_ptrs[i]);
  }
  pt->rank = rank;
  pt->world = world;
  return 0;
}

}  // namespace dtds

using namespace dtds;

// ---- allocation / IPC -----------------------------------------------------------------------
extern "C" int dtds_symm_alloc(size_t bytes, unsigned long long* out_ptr) {
  void* p = nullptr;
  DTDS_HOST_CHECK(cudaMalloc(&p, bytes));
  DTDS_HOST_CHECK(cudaMemset(p, 0, bytes));
  std::lock_guard<std::mutex> lk(g_mu);
  g_allocs.push_back({p, bytes});
  *out_ptr = reinterpret_cast<unsigned long long>(p);
  return 0;
}
extern "C" int dtds_symm_free(unsigned long long ptr) {
  std::lock_guard<std::mutex> lk(g_mu);
  for (size_t i = 0; i < g_allocs.size(); ++i) {
    if (reinterpret_cast<unsigned long long>(g_allocs[i].ptr) == ptr) {
      cudaFree(g_allocs[i].ptr);
      g_allocs.erase(g_allocs.begin() + i);
      return 0;
    }
  }
  return -1;
}
extern "C" int dtds_symm_export(unsigned long long ptr, void* handle_out /* 64 bytes */) {
  cudaIpcMemHandle_t h;
  DTDS_HOST_CHECK(cudaIpcGetMemHandle(&h, reinterpret_cast<void*>(ptr)));
  static_assert(sizeof(h) == 64, "IPC handle size");
  memcpy(handle_out, &h, sizeof(h));
  return 0;
}
extern "C" int dtds_symm_import(const void* handle /* 64 bytes */, unsigned long long* out_ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  void* p = nullptr;
  DTDS_HOST_CHECK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *out_ptr = reinterpret_cast<unsigned long long>(p);
  return 0;
}
extern "C" int dtds_symm_close(unsigned long long ptr) {
  DTDS_HOST_CHECK(cudaIpcCloseMemHandle(reinterpret_cast<void*>(ptr)));
  return 0;
}
extern "C" size_t dtds_signal_pad_bytes() {
  return (kFlagWords + kEpochWords) * sizeof(uint32_t);
}

// ---- collectives ----------------------------------------------------------------------------
// dtype: 0 = bf16, 1 = fp32, 2 = int64.  `nbytes` and `off_bytes` must be multiples of 16.
extern "C" int dtds_allreduce(const unsigned long long* data_ptrs,
                              const unsigned long long* signal_ptrs, int rank, int world,
                              void* out /* one-shot only; NULL = in place */, size_t off_bytes,
                              size_t nbytes, int dtype, float scale, int algo /*0 auto,1 one,2 two*/,
                              int blocks, cudaStream_t stream) {
  if (nbytes == 0) return 0;
  if ((nbytes % 16) || (off_bytes % 16)) {
    fprintf(stderr, "[dtds] allreduce: offset/size must be multiples of 16 bytes\n");
    return -2;
  }
  PeerTable pt;
  if (fill_table(&pt, data_ptrs, signal_ptrs, rank, world)) return -1;
  const size_t nvec = nbytes / 16;
  // The one-shot kernel writes while peers may still be reading, so it needs a destination
  // outside the symmetric buffer; in-place calls always take the two-shot path.
  if (out == nullptr && algo != 3) algo = 2;
  if (algo == 0) algo = (nbytes <= (512u << 10)) ? 1 : 2;
  const int nb = comm_blocks(algo == 1 ? nvec : (nvec + world - 1) / world, blocks > 0 ? blocks : (algo == 3 ? 32 : 0));
  if (algo == 1) {
    void* dst = out;
    if (dtype == 0) allreduce_oneshot_kernel<__nv_bfloat16><<<nb, kCommThreads, 0, stream>>>(pt, dst, off_bytes, nvec, scale);
    else if (dtype == 1) allreduce_oneshot_kernel<float><<<nb, kCommThreads, 0, stream>>>(pt, dst, off_bytes, nvec, scale);
    else allreduce_oneshot_kernel<long long><<<nb, kCommThreads, 0, stream>>>(pt, dst, off_bytes, nvec, scale);
  } else {
    if (out != nullptr) {
      fprintf(stderr, "[dtds] allreduce: the two-shot algorithm works in place\n");
      return -3;
    }
    const bool overlap = algo == 3;  // co-resident with compute: low-register variant
    if (dtype == 0) {
      if (overlap) allreduce_twoshot_kernel<__nv_bfloat16, 1><<<nb, kCommThreads, 0, stream>>>(pt, off_bytes, nvec, scale);
      else allreduce_twoshot_kernel<__nv_bfloat16, 2><<<nb, kCommThreads, 0, stream>>>(pt, off_bytes, nvec, scale);
    } else if (dtype == 1) {
      if (overlap) allreduce_twoshot_kernel<float, 1><<<nb, kCommThreads, 0, stream>>>(pt, off_bytes, nvec, scale);
      else allreduce_twoshot_kernel<float, 2><<<nb, kCommThreads, 0, stream>>>(pt, off_bytes, nvec, scale);
    } else {
      allreduce_twoshot_kernel<long long, 1><<<nb, kCommThreads, 0, stream>>>(pt, off_bytes, nvec, scale);
    }
  }
  DTDS_HOST_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtds_reduce_scatter(const unsigned long long* data_ptrs,
                                   const unsigned long long* signal_ptrs, int rank, int world,
                                   void* out, size_t off_bytes, size_t nbytes_per_rank, int dtype,
                                   float scale, int blocks, cudaStream_t stream) {
  if (nbytes_per_rank == 0) return 0;
  if ((nbytes_per_rank % 16) || (off_bytes % 16)) return -2;
  PeerTable pt;
  if (fill_table(&pt, data_ptrs, signal_ptrs, rank, world)) return -1;
  const size_t nvec = nbytes_per_rank / 16;
  const int nb = comm_blocks(nvec, blocks);
  if (dtype == 0) reduce_scatter_kernel<__nv_bfloat16><<<nb, kCommThreads, 0, stream>>>(pt, out, off_bytes, nvec, scale);
  else if (dtype == 1) reduce_scatter_kernel<float><<<nb, kCommThreads, 0, stream>>>(pt, out, off_bytes, nvec, scale);
  else reduce_scatter_kernel<long long><<<nb, kCommThreads, 0, stream>>>(pt, out, off_bytes, nvec, scale);
  DTDS_HOST_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtds_all_gather(const unsigned long long* data_ptrs,
                               const unsigned long long* signal_ptrs, int rank, int world,
                               const void* in, size_t off_bytes, size_t nbytes_per_rank, int blocks,
                               cudaStream_t stream) {
  if (nbytes_per_rank == 0) return 0;
  if ((nbytes_per_rank % 16) || (off_bytes % 16)) return -2;
  PeerTable pt;
  if (fill_table(&pt, data_ptrs, signal_ptrs, rank, world)) return -1;
  const size_t nvec = nbytes_per_rank / 16;
  all_gather_kernel<<<comm_blocks(nvec, blocks), kCommThreads, 0, stream>>>(pt, in, off_bytes, nvec);
  DTDS_HOST_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtds_broadcast(const unsigned long long* data_ptrs,
                              const unsigned long long* signal_ptrs, int rank, int world, int root,
                              size_t off_bytes, size_t nbytes, int blocks, cudaStream_t stream) {
  if (nbytes == 0) return 0;
  if ((nbytes % 16) || (off_bytes % 16)) return -2;
  PeerTable pt;
  if (fill_table(&pt, data_ptrs, signal_ptrs, rank, world)) return -1;
  const size_t nvec = nbytes / 16;
  broadcast_kernel<<<comm_blocks(nvec, blocks), kCommThreads, 0, stream>>>(pt, root, off_bytes, nvec);
  DTDS_HOST_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtds_device_barrier(const unsigned long long* signal_ptrs, int rank, int world,
                                   cudaStream_t stream) {
  PeerTable pt;
  if (fill_table(&pt, signal_ptrs, signal_ptrs, rank, world)) return -1;
  barrier_kernel<<<1, kCommThreads, 0, stream>>>(pt);
  DTDS_HOST_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtds_zero_step_nvls(const unsigned long long* grad_ptrs,
                                   const unsigned long long* signal_ptrs,
                                   const unsigned long long* param_ptrs, int rank, int world,
                                   size_t goff_elems, size_t poff_elems, size_t per_rank_elems,
                                   float* master, float* m, float* v, void* param_shard, const int* step,
                                   float lr, float beta1, float beta2, float eps, float weight_decay,
                                   float grad_scale, int adamw_mode, int bias_correction,
                                   int push_params, int blocks, unsigned long long grads_mc,
                                   unsigned long long params_mc, cudaStream_t stream) {
  if (per_rank_elems == 0) return 0;
  if ((per_rank_elems % 8) || (goff_elems % 8) || (poff_elems % 8)) {
    fprintf(stderr, "[dtds] zero_step: offsets and partition size must be multiples of 8 elements\n");
    return -2;
  }
  ZeroStepArgs a;
  memset(&a, 0, sizeof(a));
  if (fill_table(&a.grads, grad_ptrs, signal_ptrs, rank, world)) return -1;
  for (int i = 0; i < world; ++i)
    a.params[i] = param_ptrs != nullptr ? reinterpret_cast<void*>(param_ptrs[i]) : nullptr;
  a.goff_elems = goff_elems; a.poff_elems = poff_elems; a.per = per_rank_elems;
  a.master = master; a.m = m; a.v = v;
  a.param_shard = static_cast<__nv_bfloat16*>(param_shard);
  a.step = step;
  a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.weight_decay = weight_decay;
  a.grad_scale = grad_scale; a.adamw_mode = adamw_mode; a.bias_correction = bias_correction;
  a.push_params = push_params && param_ptrs != nullptr;
  a.grads_mc = reinterpret_cast<const char*>(grads_mc);
  a.params_mc = reinterpret_cast<char*>(params_mc);
  zero_rs_adam_ag_kernel<<<comm_blocks(per_rank_elems / 8, blocks), kCommThreads, 0, stream>>>(a);
  DTDS_HOST_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtds_zero_step(const unsigned long long* grad_ptrs,
                              const unsigned long long* signal_ptrs,
                              const unsigned long long* param_ptrs, int rank, int world,
                              size_t goff_elems, size_t poff_elems, size_t per_rank_elems,
                              float* master, float* m, float* v, void* param_shard, const int* step,
                              float lr, float beta1, float beta2, float eps, float weight_decay,
                              float grad_scale, int adamw_mode, int bias_correction,
                              int push_params, int blocks, cudaStream_t stream) {
  return dtds_zero_step_nvls(grad_ptrs, signal_ptrs, param_ptrs, rank, world, goff_elems, poff_elems,
                             per_rank_elems, master, m, v, param_shard, step, lr, beta1, beta2, eps, weight_decay,
                             grad_scale, adamw_mode, bias_correction, push_params, blocks, 0ull, 0ull, stream);
}

// ------------------------------------------------------------------------- segment gather (pull)
// ZeRO-3 parameter all-gather as a PULL: each rank copies the slices it needs straight out of the
// owners' shard buffers (peer loads over NVLink) into a private destination.  No barrier is
// needed: a shard is only rewritten by the owner's ZeRO step kernel, whose entry barrier cannot
// pass before every rank has issued (in stream order) all of its gathers of the current step, and
// whose exit barrier has completed before any rank starts gathering for the next step.
namespace dtds {
constexpr int kMaxSegs = 48;
struct SegArgs {
  const char* src[kMaxSegs];
  char* dst[kMaxSegs];
  unsigned long long nvec[kMaxSegs];  // 16-byte vectors per segment
  int nseg;
};
__global__ void __launch_bounds__(kCommThreads) copy_segments_kernel(SegArgs a) {
  for (int s = 0; s < a.nseg; ++s) {
    const char* src = a.src[s];
    char* dst = a.dst[s];
    const size_t n = a.nvec[s];
    for (size_t v = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < n;
         v += static_cast<size_t>(gridDim.x) * blockDim.x)
      *reinterpret_cast<uint4*>(dst + v * 16) = ld_peer_v4(src + v * 16);
  }
}
}  // namespace dtds

extern "C" int dtds_copy_segments(const unsigned long long* src_ptrs,
                                  const unsigned long long* dst_ptrs,
                                  const unsigned long long* nbytes, int nseg, int blocks,
                                  cudaStream_t stream) {
  int done = 0;
  while (done < nseg) {
    dtds::SegArgs a;
    memset(&a, 0, sizeof(a));
    size_t total_vec = 0;
    int k = 0;
    for (; k < dtds::kMaxSegs && done + k < nseg; ++k) {
      if ((nbytes[done + k] % 16) || (src_ptrs[done + k] % 16) || (dst_ptrs[done + k] % 16)) {
        fprintf(stderr, "[dtds] copy_segments: segments must be 16-byte aligned\n");
        return -2;
      }
      a.src[k] = reinterpret_cast<const char*>(src_ptrs[done + k]);
      a.dst[k] = reinterpret_cast<char*>(dst_ptrs[done + k]);
      a.nvec[k] = nbytes[done + k] / 16;
      total_vec += a.nvec[k];
    }
    a.nseg = k;
    if (total_vec > 0) {
      int nb = blocks > 0 ? blocks : 64;
      const size_t need = (total_vec + dtds::kCommThreads - 1) / dtds::kCommThreads;
      if (need < static_cast<size_t>(nb)) nb = static_cast<int>(need);
      dtds::copy_segments_kernel<<<nb, dtds::kCommThreads, 0, stream>>>(a);
      DTDS_HOST_CHECK(cudaGetLastError());
    }
    done += k;
  }
  return 0;
}

// ------------------------------------------------------------------------- NVLS (multimem)
// Two-shot all-reduce through the NVSwitch multicast object: `mc` is the multicast address of the
// symmetric buffer (every rank's copy bound to one multicast object).  multimem.ld_reduce makes
// the switch read the vector from all ranks and return the sum (one request, 1/n of the P2P
// inbound traffic); multimem.st writes the result into every rank's copy in one store.  Rank r
// handles slice r.  Barriers are the same signal-pad barriers as the P2P kernels.
namespace dtds {
template <bool BF16>
__global__ void __launch_bounds__(kCommThreads)
allreduce_nvls_kernel(PeerTable pt, char* mc, size_t off_bytes, size_t nvec, float scale) {
  cross_gpu_barrier(pt, 0);
  const size_t per = (nvec + pt.world - 1) / pt.world;
  const size_t v0 = per * pt.rank;
  const size_t v1 = min(nvec, v0 + per);
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  auto ld = [&](char* addr) -> uint4 {
    uint4 r;
    if (BF16)
      asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                   : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(addr) : "memory");
    else
      asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                   : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(addr) : "memory");
    return r;
  };
  auto st = [&](char* addr, uint4 r) {
    if (scale != 1.f) {
      if (BF16) {
        uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) w[i] = pack_bf16x2(bf16_lo(w[i]) * scale, bf16_hi(w[i]) * scale);
        r = make_uint4(w[0], w[1], w[2], w[3]);
      } else {
        r.x = __float_as_uint(__uint_as_float(r.x) * scale);
        r.y = __float_as_uint(__uint_as_float(r.y) * scale);
        r.z = __float_as_uint(__uint_as_float(r.z) * scale);
        r.w = __float_as_uint(__uint_as_float(r.w) * scale);
      }
    }
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "r"(r.x),
                 "r"(r.y), "r"(r.z), "r"(r.w)
                 : "memory");
  };
  char* base = mc + off_bytes;
  size_t v = v0 + static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  for (; v + 3 * stride < v1; v += 4 * stride) {  // four switch reductions in flight per thread
    const uint4 r0 = ld(base + v * 16);
    const uint4 r1 = ld(base + (v + stride) * 16);
    const uint4 r2 = ld(base + (v + 2 * stride) * 16);
    const uint4 r3 = ld(base + (v + 3 * stride) * 16);
    st(base + v * 16, r0);
    st(base + (v + stride) * 16, r1);
    st(base + (v + 2 * stride) * 16, r2);
    st(base + (v + 3 * stride) * 16, r3);
  }
  for (; v < v1; v += stride) st(base + v * 16, ld(base + v * 16));
  cross_gpu_barrier(pt, 1);
}
}  // namespace dtds

namespace dtds {
__device__ __forceinline__ uint4 mm_ld_reduce_bf16(const char* addr) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(addr) : "memory");
  return r;
}
__device__ __forceinline__ uint4 mm_ld_reduce_f32(const char* addr) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(addr) : "memory");
  return r;
}
__device__ __forceinline__ void mm_st(char* addr, const uint4& r) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "r"(r.x), "r"(r.y),
               "r"(r.z), "r"(r.w)
               : "memory");
}
__device__ __forceinline__ uint4 scale_raw(uint4 r, float scale, bool bf16) {
  if (scale == 1.f) return r;
  if (bf16) {
    uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) w[i] = pack_bf16x2(bf16_lo(w[i]) * scale, bf16_hi(w[i]) * scale);
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
  r.x = __float_as_uint(__uint_as_float(r.x) * scale);
  r.y = __float_as_uint(__uint_as_float(r.y) * scale);
  r.z = __float_as_uint(__uint_as_float(r.z) * scale);
  r.w = __float_as_uint(__uint_as_float(r.w) * scale);
  return r;
}

// Reduce-scatter through the switch: out[0:per) = scale * sum over ranks of slice `rank` (one
// multimem.ld_reduce per 16 bytes instead of world-1 peer loads).
template <bool BF16>
__global__ void __launch_bounds__(kCommThreads)
reduce_scatter_nvls_kernel(PeerTable pt, const char* mc, void* out, size_t off_bytes, size_t nvec_per_rank,
                           float scale) {
  cross_gpu_barrier(pt, 0);
  const char* base = mc + off_bytes + nvec_per_rank * pt.rank * 16;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  size_t v = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  for (; v + 3 * stride < nvec_per_rank; v += 4 * stride) {
    uint4 r[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      r[u] = BF16 ? mm_ld_reduce_bf16(base + (v + u * stride) * 16) : mm_ld_reduce_f32(base + (v + u * stride) * 16);
#pragma unroll
    for (int u = 0; u < 4; ++u)
      *reinterpret_cast<uint4*>(static_cast<char*>(out) + (v + u * stride) * 16) = scale_raw(r[u], scale, BF16);
  }
  for (; v < nvec_per_rank; v += stride) {
    const uint4 r = BF16 ? mm_ld_reduce_bf16(base + v * 16) : mm_ld_reduce_f32(base + v * 16);
    *reinterpret_cast<uint4*>(static_cast<char*>(out) + v * 16) = scale_raw(r, scale, BF16);
  }
  cross_gpu_barrier(pt, 1);  // inputs may be overwritten only after every rank has read them
}

// All-gather through the switch: rank r writes its slice once (multimem.st) and the switch
// replicates it into slot r of every rank's buffer.
__global__ void __launch_bounds__(kCommThreads)
all_gather_nvls_kernel(PeerTable pt, char* mc, const void* in, size_t off_bytes, size_t nvec_per_rank) {
  cross_gpu_barrier(pt, 0);  // peers are done reading the previous contents
  char* base = mc + off_bytes + nvec_per_rank * pt.rank * 16;
  for (size_t v = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < nvec_per_rank;
       v += static_cast<size_t>(gridDim.x) * blockDim.x)
    mm_st(base + v * 16, *reinterpret_cast<const uint4*>(static_cast<const char*>(in) + v * 16));
  cross_gpu_barrier(pt, 1);
}

// Broadcast through the switch: the root stores its range once to the multicast address.
__global__ void __launch_bounds__(kCommThreads)
broadcast_nvls_kernel(PeerTable pt, char* mc, int root, size_t off_bytes, size_t nvec) {
  cross_gpu_barrier(pt, 0);
  if (pt.rank == root) {
    const char* src = static_cast<const char*>(pt.data[root]) + off_bytes;
    for (size_t v = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < nvec;
         v += static_cast<size_t>(gridDim.x) * blockDim.x)
      mm_st(mc + off_bytes + v * 16, *reinterpret_cast<const uint4*>(src + v * 16));
  }
  cross_gpu_barrier(pt, 1);
}
}  // namespace dtds

// dtype: 0 bf16, 1 fp32.  `mc_ptr` = multicast address of the buffer (cuMulticast* binding, e.g.
// obtained through torch.distributed._symmetric_memory), `signal_ptrs` = per-rank signal pads.
extern "C" int dtds_allreduce_nvls(unsigned long long mc_ptr, const unsigned long long* signal_ptrs,
                                   int rank, int world, size_t off_bytes, size_t nbytes, int dtype,
                                   float scale, int blocks, cudaStream_t stream) {
  if (nbytes == 0) return 0;
  if ((nbytes % 16) || (off_bytes % 16) || mc_ptr == 0 || dtype < 0 || dtype > 1) return -2;
  dtds::PeerTable pt;
  if (dtds::fill_table(&pt, signal_ptrs, signal_ptrs, rank, world)) return -1;
  const size_t nvec = nbytes / 16;
  const int nb = dtds::comm_blocks((nvec + world - 1) / world, blocks > 0 ? blocks : 96);
  char* mc = reinterpret_cast<char*>(mc_ptr);
  if (dtype == 0) dtds::allreduce_nvls_kernel<true><<<nb, dtds::kCommThreads, 0, stream>>>(pt, mc, off_bytes, nvec, scale);
  else dtds::allreduce_nvls_kernel<false><<<nb, dtds::kCommThreads, 0, stream>>>(pt, mc, off_bytes, nvec, scale);
  DTDS_HOST_CHECK(cudaGetLastError());
  return 0;
}

// NVLS variants of reduce-scatter / all-gather / broadcast on a multicast-bound symmetric buffer
// (csrc/symm_mcast.cu).  `data_ptrs` is only needed by broadcast (the root's local address).
extern "C" int dtds_reduce_scatter_nvls(unsigned long long mc_ptr, const unsigned long long* signal_ptrs, int rank,
                                        int world, void* out, size_t off_bytes, size_t nbytes_per_rank, int dtype,
                                        float scale, int blocks, cudaStream_t stream) {
  if (nbytes_per_rank == 0) return 0;
  if ((nbytes_per_rank % 16) || (off_bytes % 16) || mc_ptr == 0 || dtype < 0 || dtype > 1) return -2;
  dtds::PeerTable pt;
  if (dtds::fill_table(&pt, signal_ptrs, signal_ptrs, rank, world)) return -1;
  const size_t nvec = nbytes_per_rank / 16;
  const int nb = dtds::comm_blocks(nvec, blocks);
  const char* mc = reinterpret_cast<const char*>(mc_ptr);
  if (dtype == 0) dtds::reduce_scatter_nvls_kernel<true><<<nb, dtds::kCommThreads, 0, stream>>>(pt, mc, out, off_bytes, nvec, scale);
  else dtds::reduce_scatter_nvls_kernel<false><<<nb, dtds::kCommThreads, 0, stream>>>(pt, mc, out, off_bytes, nvec, scale);
  DTDS_HOST_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtds_all_gather_nvls(unsigned long long mc_ptr, const unsigned long long* signal_ptrs, int rank,
                                    int world, const void* in, size_t off_bytes, size_t nbytes_per_rank, int blocks,
                                    cudaStream_t stream) {
  if (nbytes_per_rank == 0) return 0;
  if ((nbytes_per_rank % 16) || (off_bytes % 16) || mc_ptr == 0) return -2;
  dtds::PeerTable pt;
  if (dtds::fill_table(&pt, signal_ptrs, signal_ptrs, rank, world)) return -1;
  const size_t nvec = nbytes_per_rank / 16;
  dtds::all_gather_nvls_kernel<<<dtds::comm_blocks(nvec, blocks), dtds::kCommThreads, 0, stream>>>(
      pt, reinterpret_cast<char*>(mc_ptr), in, off_bytes, nvec);
  DTDS_HOST_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtds_broadcast_nvls(unsigned long long mc_ptr, const unsigned long long* data_ptrs,
                                   const unsigned long long* signal_ptrs, int rank, int world, int root,
                                   size_t off_bytes, size_t nbytes, int blocks, cudaStream_t stream) {
  if (nbytes == 0) return 0;
  if ((nbytes % 16) || (off_bytes % 16) || mc_ptr == 0) return -2;
  dtds::PeerTable pt;
  if (dtds::fill_table(&pt, data_ptrs, signal_ptrs, rank, world)) return -1;
  const size_t nvec = nbytes / 16;
  dtds::broadcast_nvls_kernel<<<dtds::comm_blocks(nvec, blocks), dtds::kCommThreads, 0, stream>>>(
      pt, reinterpret_cast<char*>(mc_ptr), root, off_bytes, nvec);
  DTDS_HOST_CHECK(cudaGetLastError());
  return 0;
}
