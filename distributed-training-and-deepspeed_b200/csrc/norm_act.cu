// LayerNorm (+dropout +residual) forward/backward, standalone activations, dropout, bias
// column-sums, row softmax and MSE.  Memory-bound kernels: 128-bit accesses, one pass over the
// data, fp32 math, bf16 storage.  SURVEY.md rows K4, K5, K7, K8, K9, K12.
//
// Dropout masks are never stored: keep/drop is a pure function of (seed, element index), so the
// backward pass recomputes it.  (The memory estimators still account one byte per element for
// the unfused nn.Dropout path, estimate_nn_memory.py:31-33 in the reference.)
#include "common.cuh"
#include "dtds_b200.h"
#include "internal.h"

namespace dtds {

// ------------------------------------------------------------------------------------------
// z = dropout(x) + residual ; y = LN(z) * gamma + beta.   One warp per row when cols <= 1024,
// otherwise one block per row.  Saves z (optional), mean and rstd for the backward pass.
// ------------------------------------------------------------------------------------------
template <int VEC_PER_LANE>  // 8-element vectors per lane; cols == VEC_PER_LANE * 32 * 8 max
__global__ void __launch_bounds__(256)
ln_fwd_warp_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ residual,
                   const __nv_bfloat16* __restrict__ gamma, const __nv_bfloat16* __restrict__ beta,
                   __nv_bfloat16* __restrict__ y, __nv_bfloat16* __restrict__ z_out,
                   float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows, int cols,
                   float eps, float p_drop, uint64_t seed, const uint64_t* __restrict__ seed_dev) {
  pdl_trigger();
  pdl_wait();
  if (seed_dev != nullptr) seed += *seed_dev * 0xD1B54A32D192ED03ull;
  const int warps_per_block = blockDim.x >> 5;
  const int row = blockIdx.x * warps_per_block + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const int nvec = cols >> 3;
  const DropRng rng = make_drop_rng(seed, p_drop);
  float v[VEC_PER_LANE][8];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < VEC_PER_LANE; ++i) {
    const int vi = lane + i * 32;
    if (vi < nvec) {
      const size_t off = static_cast<size_t>(row) * cols + vi * 8;
      const uint4 u = *reinterpret_cast<const uint4*>(x + off);
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        v[i][2 * t] = bf16_lo(w[t]);
        v[i][2 * t + 1] = bf16_hi(w[t]);
      }
      if (p_drop > 0.f) {
        float ks[8];
        drop_scale8(rng, off, ks);
#pragma unroll
        for (int t = 0; t < 8; ++t) v[i][t] *= ks[t];
      }
      if (residual != nullptr) {
        const uint4 r = *reinterpret_cast<const uint4*>(residual + off);
        const uint32_t rw[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          v[i][2 * t] += bf16_lo(rw[t]);
          v[i][2 * t + 1] += bf16_hi(rw[t]);
        }
      }
      if (z_out != nullptr) {
        // Round z to bf16 first so that forward statistics and the saved tensor agree.
        uint4 o = make_uint4(pack_bf16x2(v[i][0], v[i][1]), pack_bf16x2(v[i][2], v[i][3]),
                             pack_bf16x2(v[i][4], v[i][5]), pack_bf16x2(v[i][6], v[i][7]));
        *reinterpret_cast<uint4*>(z_out + off) = o;
        const uint32_t ow[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          v[i][2 * t] = bf16_lo(ow[t]);
          v[i][2 * t + 1] = bf16_hi(ow[t]);
        }
      }
#pragma unroll
      for (int t = 0; t < 8; ++t) sum += v[i][t];
    } else {
#pragma unroll
      for (int t = 0; t < 8; ++t) v[i][t] = 0.f;
    }
  }
  const float mean = warp_sum(sum) / cols;
  float var = 0.f;
#pragma unroll
  for (int i = 0; i < VEC_PER_LANE; ++i) {
    if (lane + i * 32 < nvec) {
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        const float d = v[i][t] - mean;
        var += d * d;
      }
    }
  }
  const float rstd = rsqrtf(warp_sum(var) / cols + eps);
  if (lane == 0) {
    if (mean_out) mean_out[row] = mean;
    if (rstd_out) rstd_out[row] = rstd;
  }
#pragma unroll
  for (int i = 0; i < VEC_PER_LANE; ++i) {
    const int vi = lane + i * 32;
    if (vi < nvec) {
      const uint4 g = *reinterpret_cast<const uint4*>(gamma + vi * 8);
      const uint4 b = *reinterpret_cast<const uint4*>(beta + vi * 8);
      const uint32_t gw[4] = {g.x, g.y, g.z, g.w}, bw[4] = {b.x, b.y, b.z, b.w};
      uint32_t o[4];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const float a0 = (v[i][2 * t] - mean) * rstd * bf16_lo(gw[t]) + bf16_lo(bw[t]);
        const float a1 = (v[i][2 * t + 1] - mean) * rstd * bf16_hi(gw[t]) + bf16_hi(bw[t]);
        o[t] = pack_bf16x2(a0, a1);
      }
      *reinterpret_cast<uint4*>(y + static_cast<size_t>(row) * cols + vi * 8) =
          make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
}

// Generic (any cols multiple of 8) block-per-row variant for very wide rows (e.g. h = 9216).
__global__ void __launch_bounds__(256)
ln_fwd_block_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ residual,
                    const __nv_bfloat16* __restrict__ gamma, const __nv_bfloat16* __restrict__ beta,
                    __nv_bfloat16* __restrict__ y, __nv_bfloat16* __restrict__ z_out,
                    float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows, int cols,
                    float eps, float p_drop, uint64_t seed, const uint64_t* __restrict__ seed_dev) {
  pdl_trigger();
  pdl_wait();
  if (seed_dev != nullptr) seed += *seed_dev * 0xD1B54A32D192ED03ull;
  __shared__ float scratch[32];
  const int row = blockIdx.x;
  const DropRng rng = make_drop_rng(seed, p_drop);
  const size_t base = static_cast<size_t>(row) * cols;
  // pass 1: materialise z (into z_out if given, else into y as temporary) and its sum
  __nv_bfloat16* zbuf = z_out != nullptr ? z_out : y;
  float sum = 0.f;
  for (int c = threadIdx.x * 8; c < cols; c += blockDim.x * 8) {
    const uint4 u = *reinterpret_cast<const uint4*>(x + base + c);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
    float v[8];
#pragma unroll
    for (int t = 0; t < 4; ++t) { v[2 * t] = bf16_lo(w[t]); v[2 * t + 1] = bf16_hi(w[t]); }
    if (p_drop > 0.f) {
      float ks[8];
      drop_scale8(rng, base + c, ks);
#pragma unroll
      for (int t = 0; t < 8; ++t) v[t] *= ks[t];
    }
    if (residual != nullptr) {
      const uint4 r = *reinterpret_cast<const uint4*>(residual + base + c);
      const uint32_t rw[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
      for (int t = 0; t < 4; ++t) { v[2 * t] += bf16_lo(rw[t]); v[2 * t + 1] += bf16_hi(rw[t]); }
    }
    uint4 o = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                         pack_bf16x2(v[6], v[7]));
    *reinterpret_cast<uint4*>(zbuf + base + c) = o;
    const uint32_t ow[4] = {o.x, o.y, o.z, o.w};
#pragma unroll
    for (int t = 0; t < 4; ++t) sum += bf16_lo(ow[t]) + bf16_hi(ow[t]);
  }
  const float mean = block_sum(sum, scratch) / cols;
  float var = 0.f;
  for (int c = threadIdx.x * 8; c < cols; c += blockDim.x * 8) {
    const uint4 u = *reinterpret_cast<const uint4*>(zbuf + base + c);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const float d0 = bf16_lo(w[t]) - mean, d1 = bf16_hi(w[t]) - mean;
      var += d0 * d0 + d1 * d1;
    }
  }
  const float rstd = rsqrtf(block_sum(var, scratch) / cols + eps);
  if (threadIdx.x == 0) {
    if (mean_out) mean_out[row] = mean;
    if (rstd_out) rstd_out[row] = rstd;
  }
  for (int c = threadIdx.x * 8; c < cols; c += blockDim.x * 8) {
    const uint4 u = *reinterpret_cast<const uint4*>(zbuf + base + c);
    const uint4 g = *reinterpret_cast<const uint4*>(gamma + c);
    const uint4 b = *reinterpret_cast<const uint4*>(beta + c);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w}, gw[4] = {g.x, g.y, g.z, g.w},
                   bw[4] = {b.x, b.y, b.z, b.w};
    uint32_t o[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const float a0 = (bf16_lo(w[t]) - mean) * rstd * bf16_lo(gw[t]) + bf16_lo(bw[t]);
      const float a1 = (bf16_hi(w[t]) - mean) * rstd * bf16_hi(gw[t]) + bf16_hi(bw[t]);
      o[t] = pack_bf16x2(a0, a1);
    }
    *reinterpret_cast<uint4*>(y + base + c) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// ------------------------------------------------------------------------------------------
// LayerNorm backward.  dz = rstd * (g*dy - mean(g*dy) - xhat * mean(g*dy*xhat)) (+ dz_extra).
// dx_drop (optional) = dz * dropout_mask_scale  (gradient w.r.t. the pre-dropout tensor).
// Per-block partial dgamma/dbeta go to `partial` [gridDim.x, 2, cols] and are reduced by
// ln_bwd_finalize_kernel: no atomics, deterministic.
// ------------------------------------------------------------------------------------------
template <int VPL>  // 8-wide vectors per lane kept in registers; 0 = generic (smem atomics)
__global__ void __launch_bounds__(256)
ln_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ z,
              const float* __restrict__ mean, const float* __restrict__ rstd,
              const __nv_bfloat16* __restrict__ gamma, const __nv_bfloat16* __restrict__ dz_extra,
              __nv_bfloat16* __restrict__ dz, __nv_bfloat16* __restrict__ dx_drop,
              float* __restrict__ partial, int rows, int cols, float p_drop, uint64_t seed,
              const uint64_t* __restrict__ seed_dev) {
  pdl_trigger();
  pdl_wait();
  if (seed_dev != nullptr) seed += *seed_dev * 0xD1B54A32D192ED03ull;
  extern __shared__ float sm[];  // [2][cols] block-level dgamma/dbeta accumulators
  float* s_dg = sm;
  float* s_db = sm + cols;
  for (int c = threadIdx.x; c < 2 * cols; c += blockDim.x) sm[c] = 0.f;
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const DropRng rng = make_drop_rng(seed, p_drop);
  const int nvec = cols >> 3;
  constexpr int R = VPL > 0 ? VPL : 1;
  float acc_dg[R][8], acc_db[R][8];
#pragma unroll
  for (int i = 0; i < R; ++i)
#pragma unroll
    for (int t = 0; t < 8; ++t) { acc_dg[i][t] = 0.f; acc_db[i][t] = 0.f; }

  for (int row = blockIdx.x * nwarps + warp; row < rows; row += gridDim.x * nwarps) {
    const size_t base = static_cast<size_t>(row) * cols;
    const float mu = mean[row], rs = rstd[row];
    float s1 = 0.f, s2 = 0.f;
    const int iters = VPL > 0 ? R : (nvec + 31) / 32;
#pragma unroll
    for (int i = 0; i < (VPL > 0 ? R : iters); ++i) {
      const int vi = lane + i * 32;
      if (vi >= nvec) break;
      const uint4 d = *reinterpret_cast<const uint4*>(dy + base + vi * 8);
      const uint4 zz = *reinterpret_cast<const uint4*>(z + base + vi * 8);
      const uint4 g = *reinterpret_cast<const uint4*>(gamma + vi * 8);
      const uint32_t dw[4] = {d.x, d.y, d.z, d.w}, zw[4] = {zz.x, zz.y, zz.z, zz.w},
                     gw[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const float dy0 = bf16_lo(dw[t]), dy1 = bf16_hi(dw[t]);
        const float xh0 = (bf16_lo(zw[t]) - mu) * rs, xh1 = (bf16_hi(zw[t]) - mu) * rs;
        const float gd0 = dy0 * bf16_lo(gw[t]), gd1 = dy1 * bf16_hi(gw[t]);
        s1 += gd0 + gd1;
        s2 += gd0 * xh0 + gd1 * xh1;
        if constexpr (VPL > 0) {
          acc_dg[i][2 * t] += dy0 * xh0;
          acc_dg[i][2 * t + 1] += dy1 * xh1;
          acc_db[i][2 * t] += dy0;
          acc_db[i][2 * t + 1] += dy1;
        } else {
          atomicAdd(&s_dg[vi * 8 + 2 * t], dy0 * xh0);
          atomicAdd(&s_dg[vi * 8 + 2 * t + 1], dy1 * xh1);
          atomicAdd(&s_db[vi * 8 + 2 * t], dy0);
          atomicAdd(&s_db[vi * 8 + 2 * t + 1], dy1);
        }
      }
    }
    s1 = warp_sum(s1) / cols;
    s2 = warp_sum(s2) / cols;
    for (int vi = lane; vi < nvec; vi += 32) {
      const uint4 d = *reinterpret_cast<const uint4*>(dy + base + vi * 8);
      const uint4 zz = *reinterpret_cast<const uint4*>(z + base + vi * 8);
      const uint4 g = *reinterpret_cast<const uint4*>(gamma + vi * 8);
      const uint32_t dw[4] = {d.x, d.y, d.z, d.w}, zw[4] = {zz.x, zz.y, zz.z, zz.w},
                     gw[4] = {g.x, g.y, g.z, g.w};
      float o[8];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const float xh0 = (bf16_lo(zw[t]) - mu) * rs, xh1 = (bf16_hi(zw[t]) - mu) * rs;
        o[2 * t] = rs * (bf16_lo(dw[t]) * bf16_lo(gw[t]) - s1 - xh0 * s2);
        o[2 * t + 1] = rs * (bf16_hi(dw[t]) * bf16_hi(gw[t]) - s1 - xh1 * s2);
      }
      if (dz_extra != nullptr) {
        const uint4 e = *reinterpret_cast<const uint4*>(dz_extra + base + vi * 8);
        const uint32_t ew[4] = {e.x, e.y, e.z, e.w};
#pragma unroll
        for (int t = 0; t < 4; ++t) { o[2 * t] += bf16_lo(ew[t]); o[2 * t + 1] += bf16_hi(ew[t]); }
      }
      *reinterpret_cast<uint4*>(dz + base + vi * 8) =
          make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]),
                     pack_bf16x2(o[6], o[7]));
      if (dx_drop != nullptr) {
        float ks[8];
        drop_scale8(rng, base + vi * 8, ks);
#pragma unroll
        for (int t = 0; t < 8; ++t) o[t] *= ks[t];
        *reinterpret_cast<uint4*>(dx_drop + base + vi * 8) =
            make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]),
                       pack_bf16x2(o[6], o[7]));
      }
    }
  }
  if constexpr (VPL > 0) {
#pragma unroll
    for (int i = 0; i < R; ++i) {
      const int vi = lane + i * 32;
      if (vi < nvec) {
#pragma unroll
        for (int t = 0; t < 8; ++t) {
          atomicAdd(&s_dg[vi * 8 + t], acc_dg[i][t]);
          atomicAdd(&s_db[vi * 8 + t], acc_db[i][t]);
        }
      }
    }
  }
  __syncthreads();
  float* out = partial + static_cast<size_t>(blockIdx.x) * 2 * cols;
  for (int c = threadIdx.x; c < 2 * cols; c += blockDim.x) out[c] = sm[c];
}

// Row-parallel variant used when there are too few rows per warp for the register-accumulating
// kernel above to pay off (the reference's 2048-token batches): pass A computes dz (and the
// pre-dropout gradient) one warp per row with no cross-row state; pass B is a column reduction of
// dy*xhat and dy over row slices (same access pattern as colsum) into `partial`.
__global__ void __launch_bounds__(256)
ln_bwd_dx_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ z,
                 const float* __restrict__ mean, const float* __restrict__ rstd,
                 const __nv_bfloat16* __restrict__ gamma, const __nv_bfloat16* __restrict__ dz_extra,
                 __nv_bfloat16* __restrict__ dz, __nv_bfloat16* __restrict__ dx_drop, int rows, int cols,
                 float p_drop, uint64_t seed, const uint64_t* __restrict__ seed_dev) {
  pdl_trigger();
  pdl_wait();
  if (seed_dev != nullptr) seed += *seed_dev * 0xD1B54A32D192ED03ull;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const DropRng rng = make_drop_rng(seed, p_drop);
  const int nvec = cols >> 3;
  const size_t base = static_cast<size_t>(row) * cols;
  const float mu = mean[row], rs = rstd[row];
  float s1 = 0.f, s2 = 0.f;
  for (int vi = lane; vi < nvec; vi += 32) {
    const uint4 d = *reinterpret_cast<const uint4*>(dy + base + vi * 8);
    const uint4 zz = *reinterpret_cast<const uint4*>(z + base + vi * 8);
    const uint4 g = *reinterpret_cast<const uint4*>(gamma + vi * 8);
    const uint32_t dw[4] = {d.x, d.y, d.z, d.w}, zw[4] = {zz.x, zz.y, zz.z, zz.w}, gw[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const float gd0 = bf16_lo(dw[t]) * bf16_lo(gw[t]), gd1 = bf16_hi(dw[t]) * bf16_hi(gw[t]);
      s1 += gd0 + gd1;
      s2 += gd0 * ((bf16_lo(zw[t]) - mu) * rs) + gd1 * ((bf16_hi(zw[t]) - mu) * rs);
    }
  }
  s1 = warp_sum(s1) / cols;
  s2 = warp_sum(s2) / cols;
  for (int vi = lane; vi < nvec; vi += 32) {
    const uint4 d = *reinterpret_cast<const uint4*>(dy + base + vi * 8);
    const uint4 zz = *reinterpret_cast<const uint4*>(z + base + vi * 8);
    const uint4 g = *reinterpret_cast<const uint4*>(gamma + vi * 8);
    const uint32_t dw[4] = {d.x, d.y, d.z, d.w}, zw[4] = {zz.x, zz.y, zz.z, zz.w}, gw[4] = {g.x, g.y, g.z, g.w};
    float o[8];
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const float xh0 = (bf16_lo(zw[t]) - mu) * rs, xh1 = (bf16_hi(zw[t]) - mu) * rs;
      o[2 * t] = rs * (bf16_lo(dw[t]) * bf16_lo(gw[t]) - s1 - xh0 * s2);
      o[2 * t + 1] = rs * (bf16_hi(dw[t]) * bf16_hi(gw[t]) - s1 - xh1 * s2);
    }
    if (dz_extra != nullptr) {
      const uint4 e = *reinterpret_cast<const uint4*>(dz_extra + base + vi * 8);
      const uint32_t ew[4] = {e.x, e.y, e.z, e.w};
#pragma unroll
      for (int t = 0; t < 4; ++t) { o[2 * t] += bf16_lo(ew[t]); o[2 * t + 1] += bf16_hi(ew[t]); }
    }
    *reinterpret_cast<uint4*>(dz + base + vi * 8) =
        make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
    if (dx_drop != nullptr) {
      float ks[8];
      drop_scale8(rng, base + vi * 8, ks);
#pragma unroll
      for (int t = 0; t < 8; ++t) o[t] *= ks[t];
      *reinterpret_cast<uint4*>(dx_drop + base + vi * 8) =
          make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
    }
  }
}
// grid = (cols / 64 tiles, row splits); partial[split][0..cols) = sum dy*xhat, [cols..2cols) = sum dy.
__global__ void __launch_bounds__(256)
ln_bwd_dgb_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ z,
                  const float* __restrict__ mean, const float* __restrict__ rstd, float* __restrict__ partial,
                  int rows, int cols) {
  pdl_trigger();
  pdl_wait();
  __shared__ float sg[8][64], sb[8][64];
  const int c0 = blockIdx.x * 64;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = c0 + tx * 2;
  const int rows_per = (rows + gridDim.y - 1) / gridDim.y;
  const int r_begin = blockIdx.y * rows_per, r_end = min(rows, r_begin + rows_per);
  float g0 = 0.f, g1 = 0.f, b0 = 0.f, b1 = 0.f;
  if (c + 1 < cols + 1 && c < cols) {
#pragma unroll 4
    for (int r = r_begin + ty; r < r_end; r += 8) {
      const uint32_t d = *reinterpret_cast<const uint32_t*>(dy + static_cast<size_t>(r) * cols + c);
      const uint32_t zz = *reinterpret_cast<const uint32_t*>(z + static_cast<size_t>(r) * cols + c);
      const float mu = mean[r], rs = rstd[r];
      const float d0 = bf16_lo(d), d1 = bf16_hi(d);
      g0 += d0 * ((bf16_lo(zz) - mu) * rs);
      g1 += d1 * ((bf16_hi(zz) - mu) * rs);
      b0 += d0;
      b1 += d1;
    }
  }
  sg[ty][tx * 2] = g0; sg[ty][tx * 2 + 1] = g1;
  sb[ty][tx * 2] = b0; sb[ty][tx * 2 + 1] = b1;
  __syncthreads();
  if (threadIdx.x < 128) {
    const int j = threadIdx.x & 63;
    const bool is_b = threadIdx.x >= 64;
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += is_b ? sb[k][j] : sg[k][j];
    if (c0 + j < cols)
      partial[static_cast<size_t>(blockIdx.y) * 2 * cols + (is_b ? cols : 0) + c0 + j] = t;
  }
}

// Reduces [nblocks, 2, cols] partials into dgamma/dbeta (bf16 or fp32, optionally accumulating).
// Block = 32 columns x 8 segments of the partial rows; coalesced loads, smem tree at the end.
__global__ void __launch_bounds__(256)
ln_bwd_finalize_kernel(const float* __restrict__ partial, int nblocks, int cols,
                       void* __restrict__ dgamma, void* __restrict__ dbeta, int out_fp32,
                       int accumulate) {
  pdl_trigger();
  pdl_wait();
  __shared__ float s[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  float a0 = 0.f, a1 = 0.f;
  if (c < 2 * cols) {
    int b = ty;
    for (; b + 8 < nblocks; b += 16) {
      a0 += partial[static_cast<size_t>(b) * 2 * cols + c];
      a1 += partial[static_cast<size_t>(b + 8) * 2 * cols + c];
    }
    if (b < nblocks) a0 += partial[static_cast<size_t>(b) * 2 * cols + c];
  }
  s[ty][tx] = a0 + a1;
  __syncthreads();
  if (ty == 0 && c < 2 * cols) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += s[k][tx];
    void* dst = c < cols ? dgamma : dbeta;
    const int i = c < cols ? c : c - cols;
    if (out_fp32) {
      float* d = static_cast<float*>(dst);
      d[i] = accumulate ? d[i] + t : t;
    } else {
      __nv_bfloat16* d = static_cast<__nv_bfloat16*>(dst);
      d[i] = __float2bfloat16(accumulate ? __bfloat162float(d[i]) + t : t);
    }
  }
}

// ------------------------------------------------------------------------------------------
// Column sum of a [rows, cols] bf16 matrix (bias gradients): out[c] = sum_r x[r, c].
// grid = (cols/64 tiles, row splits); split partials are combined with fp32 atomics when
// gridDim.y > 1 (caller zeroes `scratch` once); the last-arriving row-split block of each
// column tile folds the partial sums, writes the result and re-zeroes the scratch: one launch.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
colsum_kernel(const __nv_bfloat16* __restrict__ x, int rows, int cols, int ld,
              float* __restrict__ scratch, unsigned int* __restrict__ counters,
              void* __restrict__ out, int out_fp32, int accumulate) {
  pdl_trigger();
  pdl_wait();
  __shared__ float s[8][64];
  __shared__ bool is_last;
  const int c0 = blockIdx.x * 64;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 column-pairs x 8 row lanes
  const int c = c0 + tx * 2;
  float a0 = 0.f, a1 = 0.f;
  const int rows_per = (rows + gridDim.y - 1) / gridDim.y;
  const int r_begin = blockIdx.y * rows_per;
  const int r_end = min(rows, r_begin + rows_per);
  if (c + 1 < cols) {
#pragma unroll 4
    for (int r = r_begin + ty; r < r_end; r += 8) {
      const uint32_t u = *reinterpret_cast<const uint32_t*>(x + static_cast<size_t>(r) * ld + c);
      a0 += bf16_lo(u);
      a1 += bf16_hi(u);
    }
  } else if (c < cols) {
    for (int r = r_begin + ty; r < r_end; r += 8)
      a0 += __bfloat162float(x[static_cast<size_t>(r) * ld + c]);
  }
  s[ty][tx * 2] = a0;
  s[ty][tx * 2 + 1] = a1;
  __syncthreads();
  float t = 0.f;
  const int cc = c0 + threadIdx.x;
  if (threadIdx.x < 64) {
#pragma unroll
    for (int k = 0; k < 8; ++k) t += s[k][threadIdx.x];
  }
  if (gridDim.y == 1) {  // no row split: write the result directly
    if (threadIdx.x < 64 && cc < cols) {
      if (out_fp32) {
        float* d = static_cast<float*>(out);
        d[cc] = accumulate ? d[cc] + t : t;
      } else {
        __nv_bfloat16* d = static_cast<__nv_bfloat16*>(out);
        d[cc] = __float2bfloat16(accumulate ? __bfloat162float(d[cc]) + t : t);
      }
    }
    return;
  }
  if (threadIdx.x < 64 && cc < cols) atomicAdd(&scratch[cc], t);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = (atomicAdd(&counters[blockIdx.x], 1u) == gridDim.y - 1);
  __syncthreads();
  if (is_last) {  // the last row-split block of this column tile folds the partial sums
    __threadfence();
    if (threadIdx.x < 64 && cc < cols) {
      const float v = __ldcg(&scratch[cc]);
      scratch[cc] = 0.f;  // leave the scratch zeroed for the next call
      if (out_fp32) {
        float* d = static_cast<float*>(out);
        d[cc] = accumulate ? d[cc] + v : v;
      } else {
        __nv_bfloat16* d = static_cast<__nv_bfloat16*>(out);
        d[cc] = __float2bfloat16(accumulate ? __bfloat162float(d[cc]) + v : v);
      }
    }
    if (threadIdx.x == 0) counters[blockIdx.x] = 0;
  }
}

// ------------------------------------------------------------------------------------------
// Standalone elementwise kernels (used by the unfused, hookable module path).
// ------------------------------------------------------------------------------------------
__global__ void act_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                               size_t n, int act) {
  const size_t i = (static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 8;
  if (i + 8 <= n) {
    const uint4 u = *reinterpret_cast<const uint4*>(x + i);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
    uint32_t o[4];
#pragma unroll
    for (int t = 0; t < 4; ++t)
      o[t] = pack_bf16x2(apply_act(bf16_lo(w[t]), act), apply_act(bf16_hi(w[t]), act));
    *reinterpret_cast<uint4*>(y + i) = make_uint4(o[0], o[1], o[2], o[3]);
  } else {
    for (size_t j = i; j < n; ++j) y[j] = __float2bfloat16(apply_act(__bfloat162float(x[j]), act));
  }
}
__global__ void act_bwd_kernel(const __nv_bfloat16* __restrict__ dy,
                               const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ dx,
                               size_t n, int act) {
  const size_t i = (static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 8;
  if (i + 8 <= n) {
    const uint4 u = *reinterpret_cast<const uint4*>(x + i);
    const uint4 d = *reinterpret_cast<const uint4*>(dy + i);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w}, dw[4] = {d.x, d.y, d.z, d.w};
    uint32_t o[4];
#pragma unroll
    for (int t = 0; t < 4; ++t)
      o[t] = pack_bf16x2(bf16_lo(dw[t]) * act_grad(bf16_lo(w[t]), act),
                         bf16_hi(dw[t]) * act_grad(bf16_hi(w[t]), act));
    *reinterpret_cast<uint4*>(dx + i) = make_uint4(o[0], o[1], o[2], o[3]);
  } else {
    for (size_t j = i; j < n; ++j)
      dx[j] = __float2bfloat16(__bfloat162float(dy[j]) * act_grad(__bfloat162float(x[j]), act));
  }
}
// y = x * keep_scale(seed, i) (+ residual).  Backward is the same kernel applied to dy.
__global__ void dropout_kernel(const __nv_bfloat16* __restrict__ x,
                               const __nv_bfloat16* __restrict__ residual,
                               __nv_bfloat16* __restrict__ y, size_t n, float p, uint64_t seed,
                               const uint64_t* __restrict__ seed_dev) {
  if (seed_dev != nullptr) seed += *seed_dev * 0xD1B54A32D192ED03ull;
  const size_t i = (static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 8;
  const DropRng rng = make_drop_rng(seed, p);
  if (i + 8 <= n) {
    const uint4 u = *reinterpret_cast<const uint4*>(x + i);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
    float v[8], ks[8];
#pragma unroll
    for (int t = 0; t < 4; ++t) { v[2 * t] = bf16_lo(w[t]); v[2 * t + 1] = bf16_hi(w[t]); }
    drop_scale8(rng, i, ks);
#pragma unroll
    for (int t = 0; t < 8; ++t) v[t] *= ks[t];
    if (residual != nullptr) {
      const uint4 r = *reinterpret_cast<const uint4*>(residual + i);
      const uint32_t rw[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
      for (int t = 0; t < 4; ++t) { v[2 * t] += bf16_lo(rw[t]); v[2 * t + 1] += bf16_hi(rw[t]); }
    }
    *reinterpret_cast<uint4*>(y + i) =
        make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                   pack_bf16x2(v[6], v[7]));
  } else {
    for (size_t j = i; j < n; ++j) {
      float v = __bfloat162float(x[j]) * drop_scale1(rng, j);
      if (residual != nullptr) v += __bfloat162float(residual[j]);
      y[j] = __float2bfloat16(v);
    }
  }
}

// Row softmax over the last dimension, one warp per row (rows of the [B*heads, S, S] scores in
// the unfused TransformerBlock path).  `dy == nullptr` -> forward; else backward given y.
__global__ void __launch_bounds__(256)
softmax_rows_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                    __nv_bfloat16* __restrict__ out, int rows, int cols) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const __nv_bfloat16* xr = x + static_cast<size_t>(row) * cols;
  __nv_bfloat16* o = out + static_cast<size_t>(row) * cols;
  if (dy == nullptr) {
    float m = -INFINITY;
    for (int c = lane; c < cols; c += 32) m = fmaxf(m, __bfloat162float(xr[c]));
    m = warp_max(m);
    float s = 0.f;
    for (int c = lane; c < cols; c += 32) s += __expf(__bfloat162float(xr[c]) - m);
    s = warp_sum(s);
    const float inv = 1.f / s;
    for (int c = lane; c < cols; c += 32)
      o[c] = __float2bfloat16(__expf(__bfloat162float(xr[c]) - m) * inv);
  } else {
    // x holds y (the softmax output); dx = y * (dy - sum(dy*y))
    const __nv_bfloat16* dr = dy + static_cast<size_t>(row) * cols;
    float dot = 0.f;
    for (int c = lane; c < cols; c += 32) dot += __bfloat162float(xr[c]) * __bfloat162float(dr[c]);
    dot = warp_sum(dot);
    for (int c = lane; c < cols; c += 32)
      o[c] = __float2bfloat16(__bfloat162float(xr[c]) * (__bfloat162float(dr[c]) - dot));
  }
}

// MSE loss (mean) forward+backward in one pass: loss += sum((x-t)^2)/n ; dx = 2(x-t)/n * gscale.
__global__ void __launch_bounds__(256)
mse_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ t,
           __nv_bfloat16* __restrict__ dx, float* __restrict__ loss, size_t n, float gscale) {
  __shared__ float scratch[32];
  float acc = 0.f;
  const float inv_n = 1.f / static_cast<float>(n);
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const float d = __bfloat162float(x[i]) - __bfloat162float(t[i]);
    acc += d * d;
    if (dx != nullptr) dx[i] = __float2bfloat16(2.f * d * inv_n * gscale);
  }
  acc = block_sum(acc, scratch);
  if (threadIdx.x == 0) atomicAdd(loss, acc * inv_n);
}

}  // namespace dtds

using namespace dtds;

extern "C" int dtds_layernorm_fwd(const void* x, const void* residual, const void* gamma,
                                  const void* beta, void* y, void* z_out, float* mean, float* rstd,
                                  int rows, int cols, float eps, float p_drop,
                                  unsigned long long seed, const unsigned long long* seed_dev,
                                  cudaStream_t stream) {
  if (rows <= 0) return 0;
  if (cols % 8 != 0) {
    fprintf(stderr, "[dtds] layernorm: cols (%d) must be a multiple of 8\n", cols);
    return -1;
  }
  auto X = static_cast<const __nv_bfloat16*>(x);
  auto R = static_cast<const __nv_bfloat16*>(residual);
  auto G = static_cast<const __nv_bfloat16*>(gamma);
  auto Bt = static_cast<const __nv_bfloat16*>(beta);
  auto Y = static_cast<__nv_bfloat16*>(y);
  auto Z = static_cast<__nv_bfloat16*>(z_out);
  const int wpb = 8;
  const int grid = (rows + wpb - 1) / wpb;
  if (cols <= 256) {
    DTDS_HOST_CHECK(launch_pdl(ln_fwd_warp_kernel<1>, dim3(grid), dim3(256), 0, stream, X, R, G, Bt, Y, Z, mean, rstd, rows, cols, eps, p_drop, seed, reinterpret_cast<const uint64_t*>(seed_dev)));
  } else if (cols <= 768) {
    DTDS_HOST_CHECK(launch_pdl(ln_fwd_warp_kernel<3>, dim3(grid), dim3(256), 0, stream, X, R, G, Bt, Y, Z, mean, rstd, rows, cols, eps, p_drop, seed, reinterpret_cast<const uint64_t*>(seed_dev)));
  } else if (cols <= 1024) {
    DTDS_HOST_CHECK(launch_pdl(ln_fwd_warp_kernel<4>, dim3(grid), dim3(256), 0, stream, X, R, G, Bt, Y, Z, mean, rstd, rows, cols, eps, p_drop, seed, reinterpret_cast<const uint64_t*>(seed_dev)));
  } else {
    DTDS_HOST_CHECK(launch_pdl(ln_fwd_block_kernel, dim3(rows), dim3(256), 0, stream, X, R, G, Bt, Y, Z, mean, rstd, rows, cols, eps, p_drop, seed, reinterpret_cast<const uint64_t*>(seed_dev)));
  }
  return 0;
}

extern "C" int dtds_layernorm_bwd_blocks(int rows, int cols) {
  (void)cols;
  if (rows <= 64 * dtds::sm_count() && cols % 2 == 0) {  // row-parallel path: `blocks` = row splits
    int splits = (rows + 127) / 128;
    return splits < 1 ? 1 : (splits > 64 ? 64 : splits);
  }
  int blocks = (rows + 31) / 32;  // >= 4 rows per warp so the register accumulators pay off
  const int cap = dtds::sm_count();
  return blocks < cap ? (blocks < 1 ? 1 : blocks) : cap;
}

// 1 if dtds_layernorm_bwd uses separate kernels for dz and for dgamma/dbeta at this shape, i.e.
// the two halves may be issued by two calls (dgamma = dbeta = NULL, then dz = NULL).
extern "C" int dtds_layernorm_bwd_splittable(int rows, int cols) {
  return (rows <= 64 * dtds::sm_count() && cols % 2 == 0) ? 1 : 0;
}

extern "C" int dtds_layernorm_bwd(const void* dy, const void* z, const float* mean,
                                  const float* rstd, const void* gamma, const void* dz_extra,
                                  void* dz, void* dx_drop, void* dgamma, void* dbeta,
                                  float* partial /* [blocks,2,cols] */, int rows, int cols,
                                  float p_drop, unsigned long long seed,
                                  const unsigned long long* seed_dev, int grads_fp32,
                                  int accumulate, cudaStream_t stream) {
  if (rows <= 0) return 0;
  if (cols % 8 != 0) return -1;
  const int blocks = dtds_layernorm_bwd_blocks(rows, cols);
  if (dtds_layernorm_bwd_splittable(rows, cols)) {
    // dz == NULL: parameter gradients only (the caller runs them on another stream)
    if (dz != nullptr)
    DTDS_HOST_CHECK(launch_pdl(
        ln_bwd_dx_kernel, dim3((rows + 7) / 8), dim3(256), 0, stream,
        static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(z), mean, rstd,
        static_cast<const __nv_bfloat16*>(gamma), static_cast<const __nv_bfloat16*>(dz_extra),
        static_cast<__nv_bfloat16*>(dz), static_cast<__nv_bfloat16*>(dx_drop), rows, cols, p_drop, seed,
        reinterpret_cast<const uint64_t*>(seed_dev)));
    if (dgamma != nullptr && dbeta != nullptr) {
      dim3 grid((cols + 63) / 64, blocks);
      DTDS_HOST_CHECK(launch_pdl(ln_bwd_dgb_kernel, grid, dim3(256), 0, stream,
                                 static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(z),
                                 mean, rstd, partial, rows, cols));
      DTDS_HOST_CHECK(launch_pdl(ln_bwd_finalize_kernel, dim3((2 * cols + 31) / 32), dim3(256), 0, stream,
                                 partial, blocks, cols, dgamma, dbeta, grads_fp32, accumulate));
    }
    return 0;
  }
  if (dz == nullptr) {
    fprintf(stderr, "[dtds] layernorm_bwd: dz may only be NULL on the row-parallel path\n");
    return -1;
  }
  const size_t smem = sizeof(float) * 2 * cols;
#define DTDS_LN_BWD(VPL_)                                                                        \
  ln_bwd_kernel<VPL_><<<blocks, 256, smem, stream>>>(                                             \
      static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(z), mean, rstd,    \
      static_cast<const __nv_bfloat16*>(gamma), static_cast<const __nv_bfloat16*>(dz_extra),      \
      static_cast<__nv_bfloat16*>(dz), static_cast<__nv_bfloat16*>(dx_drop), partial, rows, cols, \
      p_drop, seed, reinterpret_cast<const uint64_t*>(seed_dev))
  if (cols <= 256) DTDS_LN_BWD(1);
  else if (cols <= 768) DTDS_LN_BWD(3);
  else if (cols <= 1024) DTDS_LN_BWD(4);
  else {
    if (smem > 48 * 1024)
      DTDS_HOST_CHECK(cudaFuncSetAttribute(ln_bwd_kernel<0>,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           static_cast<int>(smem)));
    DTDS_LN_BWD(0);
  }
#undef DTDS_LN_BWD
  DTDS_HOST_CHECK(cudaGetLastError());
  if (dgamma != nullptr && dbeta != nullptr) {
    ln_bwd_finalize_kernel<<<(2 * cols + 31) / 32, 256, 0, stream>>>(partial, blocks, cols, dgamma,
                                                                     dbeta, grads_fp32, accumulate);
    DTDS_HOST_CHECK(cudaGetLastError());
  }
  return 0;
}

extern "C" int dtds_colsum(const void* x, int rows, int cols, int ld,
                           float* scratch /* [cols + cols/64 + 130] zeroed once by the caller */,
                           void* out, int out_fp32, int accumulate, cudaStream_t stream) {
  if (rows <= 0 || cols <= 0) return 0;
  const int tiles = (cols + 63) / 64;
  int splits = 1;
  if (tiles < 2 * dtds::sm_count()) {
    splits = (2 * dtds::sm_count() + tiles - 1) / tiles;
    const int max_splits = (rows + 63) / 64;
    if (splits > max_splits) splits = max_splits;
    if (splits > 64) splits = 64;
    if (splits < 1) splits = 1;
  }
  dim3 grid(tiles, splits);
  // per-tile arrival counters live right after the partial sums (same zeroed scratch)
  unsigned int* counters = reinterpret_cast<unsigned int*>(scratch + tiles * 64);
  DTDS_HOST_CHECK(launch_pdl(colsum_kernel, grid, dim3(256), 0, stream, static_cast<const __nv_bfloat16*>(x),
                             rows, cols, ld, scratch, counters, out, out_fp32, accumulate));
  return 0;
}

extern "C" int dtds_act_fwd(const void* x, void* y, size_t n, int act, cudaStream_t stream) {
  if (n == 0) return 0;
  const size_t threads = (n + 7) / 8;
  act_fwd_kernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, stream>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), n, act);
  DTDS_HOST_CHECK(cudaGetLastError());
  return 0;
}
extern "C" int dtds_act_bwd(const void* dy, const void* x, void* dx, size_t n, int act,
                            cudaStream_t stream) {
  if (n == 0) return 0;
  const size_t threads = (n + 7) / 8;
  act_bwd_kernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, stream>>>(
      static_cast<const __nv_bfloat16*>(dy), static_cast<const __nv_bfloat16*>(x),
      static_cast<__nv_bfloat16*>(dx), n, act);
  DTDS_HOST_CHECK(cudaGetLastError());
  return 0;
}
extern "C" int dtds_dropout(const void* x, const void* residual, void* y, size_t n, float p,
                            unsigned long long seed, const unsigned long long* seed_dev,
                            cudaStream_t stream) {
  if (n == 0) return 0;
  const size_t threads = (n + 7) / 8;
  dropout_kernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, stream>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(residual),
      static_cast<__nv_bfloat16*>(y), n, p, seed,
      reinterpret_cast<const uint64_t*>(seed_dev));
  DTDS_HOST_CHECK(cudaGetLastError());
  return 0;
}
extern "C" int dtds_softmax_rows(const void* x, const void* dy, void* out, int rows, int cols,
                                 cudaStream_t stream) {
  if (rows <= 0) return 0;
  softmax_rows_kernel<<<(rows + 7) / 8, 256, 0, stream>>>(
      static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(dy),
      static_cast<__nv_bfloat16*>(out), rows, cols);
  DTDS_HOST_CHECK(cudaGetLastError());
  return 0;
}
extern "C" int dtds_mse(const void* x, const void* t, void* dx, float* loss /* zeroed */, size_t n,
                        float gscale, cudaStream_t stream) {
  if (n == 0) return 0;
  int grid = 4 * dtds::sm_count();
  mse_kernel<<<grid, 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(x),
                                       static_cast<const __nv_bfloat16*>(t),
                                       static_cast<__nv_bfloat16*>(dx), loss, n, gscale);
  DTDS_HOST_CHECK(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------
// Column sum without any allocation by caller or callee.
//
// dtds_colsum() needs a small zero-initialised fp32 scratch (per-column partial sums plus one
// arrival counter per 64-column tile) that its caller has to allocate.  Some hosts of this
// library cannot allocate device memory at the point where they need a bias gradient: code
// running while a CUDA graph is being captured (cudaMalloc is illegal during capture), code
// inside a framework whose caching allocator owns all device memory and must not be re-entered
// (an allocator callback, a destructor, an out-of-memory handler), and plain C callers that only
// hold the tensors they were handed.  For them the scratch lives in this library's image: a
// file-scope __device__ array of fixed size that every GPU which loads the library keeps until
// the process exits (kNoAllocMaxCols + kNoAllocMaxTiles 4-byte words = 132 KiB per GPU).
//
// In host code `g_colsum_scratch` is only a handle for the symbol, not a device address, so the
// entry point asks the runtime for the address (cudaGetSymbolAddress) once per device, with that
// device current, and hands the returned pointer to the kernel.  That first call also clears the
// array with a cudaMemsetAsync on ITS stream: other streams must be ordered behind it before they
// use the entry point.  The kernel restores the zeros it consumed before it exits, so a launch
// that dies part-way leaves the array dirty until the process ends (nothing can repair it without
// the set-up path, which runs once): callers should treat such a failure as fatal for this path.
constexpr int kNoAllocMaxCols = 32768;
constexpr int kNoAllocMaxTiles = kNoAllocMaxCols / 64;
static __device__ __align__(256) float g_colsum_scratch[kNoAllocMaxCols + kNoAllocMaxTiles];

namespace {
constexpr int kMaxDevices = 64;
float* g_scratch_addr[kMaxDevices] = {nullptr};
}  // namespace

extern "C" int dtds_colsum_noalloc(const void* x, int rows, int cols, int ld, void* out,
                                   int out_fp32, int accumulate, cudaStream_t stream) {
  if (rows <= 0 || cols <= 0) return 0;
  if (cols > kNoAllocMaxCols) {
    fprintf(stderr,
            "[dtds] colsum_noalloc: cols = %d exceeds the built-in scratch (max %d columns); use "
            "dtds_colsum with a caller-provided scratch instead\n",
            cols, kNoAllocMaxCols);
    return -2;
  }
  int dev = 0;
  DTDS_HOST_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) {
    fprintf(stderr, "[dtds] colsum_noalloc: device ordinal %d out of range\n", dev);
    return -3;
  }
  if (g_scratch_addr[dev] == nullptr) {
    // One-time set-up for this device (it is the current one): resolve the symbol to its address
    // on THIS device and start from a clean scratch.  Not an allocation, but it is not something
    // to record into a CUDA graph either: make the first call on each device before any capture.
    float* addr = nullptr;
    DTDS_HOST_CHECK(cudaGetSymbolAddress(reinterpret_cast<void**>(&addr), g_colsum_scratch));
    DTDS_HOST_CHECK(cudaMemsetAsync(addr, 0, sizeof(float) * (kNoAllocMaxCols + kNoAllocMaxTiles), stream));
    g_scratch_addr[dev] = addr;
  }
  float* scratch = g_scratch_addr[dev];
  const int tiles = (cols + 63) / 64;
  int splits = 1;
  if (tiles < 2 * dtds::sm_count()) {
    splits = (2 * dtds::sm_count() + tiles - 1) / tiles;
    const int max_splits = (rows + 63) / 64;
    if (splits > max_splits) splits = max_splits;
    if (splits > 64) splits = 64;
    if (splits < 1) splits = 1;
  }
  // Partial sums occupy [0, tiles*64); the per-tile arrival counters sit at a fixed offset behind
  // the largest possible sum region so calls with different `cols` never alias the two.  The
  // kernel re-zeroes everything it touched before it exits, so consecutive calls need no memset.
  unsigned int* counters = reinterpret_cast<unsigned int*>(scratch + kNoAllocMaxCols);
  dim3 grid(tiles, splits);
  dtds::colsum_kernel<<<grid, 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(x), rows, cols, ld,
                                                scratch, counters, out, out_fp32, accumulate);
  DTDS_HOST_CHECK(cudaGetLastError());
  return 0;
}
