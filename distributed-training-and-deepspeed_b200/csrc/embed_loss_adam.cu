// Embedding gather (+position +token-type) forward / scatter-add backward, fused cross-entropy
// forward+backward over the vocabulary, and the flat fused Adam/AdamW step.
// SURVEY.md rows K10 (embedding), K11 (cross-entropy), K13 (Adam; DeepSpeed FusedAdam /
// torch.optim.AdamW in the reference: zero_dp_training.py:28-33, data_parallel_training.py:34).
#include "common.cuh"
#include "dtds_b200.h"
#include "internal.h"

namespace dtds {

// out[t, :] = word[ids[t]] + pos[pos_ids ? pos_ids[t] : t % seq] + type[type_ids ? .. : 0]
// One warp per token, 128-bit accesses.  pos / type tables are optional (BLOOM has neither).
__global__ void __launch_bounds__(256)
embedding_fwd_kernel(const int64_t* __restrict__ ids, const int64_t* __restrict__ type_ids,
                     const __nv_bfloat16* __restrict__ word, const __nv_bfloat16* __restrict__ pos,
                     const __nv_bfloat16* __restrict__ type, __nv_bfloat16* __restrict__ out,
                     int tokens, int seq, int hidden, int pos_offset, int vocab) {
  const int t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (t >= tokens) return;
  const int lane = threadIdx.x & 31;
  int64_t id = ids[t];
  if (id < 0 || id >= vocab) id = 0;  // never index outside the table
  const int p = (t % seq) + pos_offset;
  const int ty = type_ids != nullptr ? static_cast<int>(type_ids[t]) : 0;
  const __nv_bfloat16* w = word + static_cast<size_t>(id) * hidden;
  for (int c = lane * 8; c < hidden; c += 256) {
    const uint4 a = *reinterpret_cast<const uint4*>(w + c);
    float v[8];
    const uint32_t aw[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) { v[2 * k] = bf16_lo(aw[k]); v[2 * k + 1] = bf16_hi(aw[k]); }
    if (pos != nullptr) {
      const uint4 b = *reinterpret_cast<const uint4*>(pos + static_cast<size_t>(p) * hidden + c);
      const uint32_t bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) { v[2 * k] += bf16_lo(bw[k]); v[2 * k + 1] += bf16_hi(bw[k]); }
    }
    if (type != nullptr) {
      const uint4 b = *reinterpret_cast<const uint4*>(type + static_cast<size_t>(ty) * hidden + c);
      const uint32_t bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) { v[2 * k] += bf16_lo(bw[k]); v[2 * k + 1] += bf16_hi(bw[k]); }
    }
    *reinterpret_cast<uint4*>(out + static_cast<size_t>(t) * hidden + c) =
        make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                   pack_bf16x2(v[6], v[7]));
  }
}

__device__ __forceinline__ void atomic_add_bf16x2(__nv_bfloat16* addr, float a, float b) {
  atomicAdd(reinterpret_cast<__nv_bfloat162*>(addr), __floats2bfloat162_rn(a, b));
}

// Scatter-add of d_out rows into the word-embedding gradient (bf16x2 or fp32 atomics): repeated
// token ids collide, so this must be atomic.  One warp per token.
__global__ void __launch_bounds__(256)
embedding_bwd_word_kernel(const int64_t* __restrict__ ids, const __nv_bfloat16* __restrict__ dout,
                          void* __restrict__ dword, int tokens, int hidden, int vocab,
                          int grads_fp32) {
  const int t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (t >= tokens) return;
  const int lane = threadIdx.x & 31;
  const int64_t id = ids[t];
  if (id < 0 || id >= vocab) return;
  const __nv_bfloat16* src = dout + static_cast<size_t>(t) * hidden;
  for (int c = lane * 2; c < hidden; c += 64) {
    const uint32_t u = *reinterpret_cast<const uint32_t*>(src + c);
    if (grads_fp32) {
      float* d = static_cast<float*>(dword) + static_cast<size_t>(id) * hidden + c;
      atomicAdd(d, bf16_lo(u));
      atomicAdd(d + 1, bf16_hi(u));
    } else {
      atomic_add_bf16x2(static_cast<__nv_bfloat16*>(dword) + static_cast<size_t>(id) * hidden + c,
                        bf16_lo(u), bf16_hi(u));
    }
  }
}
// Position-embedding gradient: positions are t % seq, so dpos[s] = sum_b dout[b*seq + s].
__global__ void embedding_bwd_pos_kernel(const __nv_bfloat16* __restrict__ dout,
                                         void* __restrict__ dpos, int batch, int seq, int hidden,
                                         int pos_offset, int grads_fp32, int accumulate) {
  const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= static_cast<size_t>(seq) * hidden) return;
  float s = 0.f;
  for (int b = 0; b < batch; ++b) s += __bfloat162float(dout[static_cast<size_t>(b) * seq * hidden + i]);
  const size_t o = i + static_cast<size_t>(pos_offset) * hidden;
  if (grads_fp32) {
    float* d = static_cast<float*>(dpos);
    d[o] = accumulate ? d[o] + s : s;
  } else {
    __nv_bfloat16* d = static_cast<__nv_bfloat16*>(dpos);
    d[o] = __float2bfloat16(accumulate ? __bfloat162float(d[o]) + s : s);
  }
}

// ------------------------------------------------------------------------------------------
// Cross-entropy over [rows, vocab] bf16 logits with ignore_index, mean reduction.
//   pass A (ce_count_kernel): n_valid = #(label != ignore).
//   pass B (ce_fwd_bwd_kernel): one block per row; the row is read once into registers/smem,
//     lse computed, loss/n_valid accumulated, and (if dlogits) the gradient
//     (softmax - onehot) * gscale / n_valid written — in place is allowed.
// `shift` = 1 implements the causal-LM convention (logits at t predict label t+1 inside each
// sequence of length `seq`; the last position of a sequence is ignored).
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int64_t ce_label(const int64_t* labels, int row, int seq, int shift,
                                            int64_t ignore) {
  if (!shift) return labels[row];
  const int s = row % seq;
  return (s == seq - 1) ? ignore : labels[row + 1];
}

__global__ void ce_count_kernel(const int64_t* __restrict__ labels, int rows, int seq, int shift,
                                int64_t ignore, float* __restrict__ stats /* [0]=n_valid [1]=loss */) {
  __shared__ float scratch[32];
  float n = 0.f;
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += gridDim.x * blockDim.x)
    n += (ce_label(labels, r, seq, shift, ignore) != ignore) ? 1.f : 0.f;
  n = block_sum(n, scratch);
  if (threadIdx.x == 0 && n != 0.f) atomicAdd(&stats[0], n);
}

constexpr int kCeThreads = 512;
constexpr int kCeMaxPerThread = 8;  // 8 x uint4 (8 bf16 each) per thread -> vocab <= 32768 in regs

__global__ void __launch_bounds__(kCeThreads)
ce_fwd_bwd_kernel(const __nv_bfloat16* __restrict__ logits, const int64_t* __restrict__ labels,
                  __nv_bfloat16* __restrict__ dlogits, float* __restrict__ stats,
                  float* __restrict__ row_loss, int rows, int vocab, int ld, int seq, int shift,
                  int64_t ignore, float gscale) {
  __shared__ float scratch[32];
  const int row = blockIdx.x;
  const int64_t label = ce_label(labels, row, seq, shift, ignore);
  const bool valid = (label != ignore) && label >= 0 && label < vocab;
  const __nv_bfloat16* x = logits + static_cast<size_t>(row) * ld;
  __nv_bfloat16* dx = dlogits != nullptr ? dlogits + static_cast<size_t>(row) * ld : nullptr;
  const int nvec = vocab >> 3;  // full 8-wide vectors; the tail (< 8) is handled by scalars
  const int tail0 = nvec << 3;

  if (!valid) {
    if (dx != nullptr) {
      for (int vi = threadIdx.x; vi < nvec; vi += kCeThreads)
        *reinterpret_cast<uint4*>(dx + vi * 8) = make_uint4(0, 0, 0, 0);
      for (int c = tail0 + threadIdx.x; c < vocab; c += kCeThreads) dx[c] = __float2bfloat16(0.f);
    }
    if (threadIdx.x == 0 && row_loss != nullptr) row_loss[row] = 0.f;
    return;
  }
  const bool in_regs = nvec <= kCeThreads * kCeMaxPerThread;
  uint4 cache[kCeMaxPerThread];
  float m = -INFINITY;
  if (in_regs) {
#pragma unroll
    for (int i = 0; i < kCeMaxPerThread; ++i) {
      const int vi = threadIdx.x + i * kCeThreads;
      if (vi < nvec) {
        cache[i] = *reinterpret_cast<const uint4*>(x + vi * 8);
        const uint32_t w[4] = {cache[i].x, cache[i].y, cache[i].z, cache[i].w};
#pragma unroll
        for (int k = 0; k < 4; ++k) m = fmaxf(m, fmaxf(bf16_lo(w[k]), bf16_hi(w[k])));
      }
    }
  } else {
    for (int vi = threadIdx.x; vi < nvec; vi += kCeThreads) {
      const uint4 u = *reinterpret_cast<const uint4*>(x + vi * 8);
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) m = fmaxf(m, fmaxf(bf16_lo(w[k]), bf16_hi(w[k])));
    }
  }
  for (int c = tail0 + threadIdx.x; c < vocab; c += kCeThreads) m = fmaxf(m, __bfloat162float(x[c]));
  m = block_max(m, scratch);

  float s = 0.f;
  if (in_regs) {
#pragma unroll
    for (int i = 0; i < kCeMaxPerThread; ++i) {
      if (threadIdx.x + i * kCeThreads < nvec) {
        const uint32_t w[4] = {cache[i].x, cache[i].y, cache[i].z, cache[i].w};
#pragma unroll
        for (int k = 0; k < 4; ++k) s += __expf(bf16_lo(w[k]) - m) + __expf(bf16_hi(w[k]) - m);
      }
    }
  } else {
    for (int vi = threadIdx.x; vi < nvec; vi += kCeThreads) {
      const uint4 u = *reinterpret_cast<const uint4*>(x + vi * 8);
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) s += __expf(bf16_lo(w[k]) - m) + __expf(bf16_hi(w[k]) - m);
    }
  }
  for (int c = tail0 + threadIdx.x; c < vocab; c += kCeThreads) s += __expf(__bfloat162float(x[c]) - m);
  s = block_sum(s, scratch);
  const float lse = m + __logf(s);
  const float n_valid = fmaxf(stats[0], 1.f);
  const float xl = __bfloat162float(x[label]);
  __syncthreads();  // everyone has read x[label] before any in-place gradient store
  if (threadIdx.x == 0) {
    const float l = lse - xl;
    if (row_loss != nullptr) row_loss[row] = l;
    atomicAdd(&stats[1], l / n_valid);
  }
  if (dx == nullptr) return;
  const float g = gscale / n_valid;
  const float inv_s = 1.f / s;
  const int lab = static_cast<int>(label);
  auto grad_pair = [&](uint32_t w, int c) -> uint32_t {
    float a = __expf(bf16_lo(w) - m) * inv_s, b = __expf(bf16_hi(w) - m) * inv_s;
    if (c == lab) a -= 1.f;
    if (c + 1 == lab) b -= 1.f;
    return pack_bf16x2(a * g, b * g);
  };
  if (in_regs) {
#pragma unroll
    for (int i = 0; i < kCeMaxPerThread; ++i) {
      const int vi = threadIdx.x + i * kCeThreads;
      if (vi < nvec) {
        const int c = vi * 8;
        *reinterpret_cast<uint4*>(dx + c) =
            make_uint4(grad_pair(cache[i].x, c), grad_pair(cache[i].y, c + 2),
                       grad_pair(cache[i].z, c + 4), grad_pair(cache[i].w, c + 6));
      }
    }
  } else {
    for (int vi = threadIdx.x; vi < nvec; vi += kCeThreads) {
      const int c = vi * 8;
      const uint4 u = *reinterpret_cast<const uint4*>(x + c);
      *reinterpret_cast<uint4*>(dx + c) = make_uint4(grad_pair(u.x, c), grad_pair(u.y, c + 2),
                                                     grad_pair(u.z, c + 4), grad_pair(u.w, c + 6));
    }
  }
  for (int c = tail0 + threadIdx.x; c < vocab; c += kCeThreads) {
    float pr = __expf(__bfloat162float(x[c]) - m) * inv_s;
    if (c == lab) pr -= 1.f;
    dx[c] = __float2bfloat16(pr * g);
  }
}

// ------------------------------------------------------------------------------------------
// Flat fused Adam / AdamW.  One launch updates a whole flat parameter range:
//   master fp32 p, fp32 m, v; gradient bf16 or fp32; optional bf16 working copy written back.
// `step` lives in device memory so the launch can be replayed from a CUDA graph; bias
// correction uses (*step + 1) and adam_step_inc_kernel advances it afterwards.
// ------------------------------------------------------------------------------------------
struct AdamArgs {
  float lr, beta1, beta2, eps, weight_decay, grad_scale;
  int adamw_mode;       // 1: decoupled weight decay, 0: L2 added to the gradient
  int bias_correction;  // 1: torch/DeepSpeed semantics
};

template <bool GRAD_FP32>
__global__ void __launch_bounds__(256)
adam_flat_kernel(float* __restrict__ p, __nv_bfloat16* __restrict__ p_lp, const void* __restrict__ g,
                 float* __restrict__ m, float* __restrict__ v, size_t n,
                 const int* __restrict__ step, AdamArgs a) {
  const int t = *step + 1;
  float c1 = 1.f, c2 = 1.f;
  if (a.bias_correction) {
    c1 = 1.f - __powf(a.beta1, static_cast<float>(t));
    c2 = 1.f - __powf(a.beta2, static_cast<float>(t));
  }
  const float step_size = a.lr / c1;
  const float inv_sqrt_c2 = rsqrtf(c2);
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x * 4;
  for (size_t i = (static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 4; i < n;
       i += stride) {
    if (i + 4 <= n) {
      float4 pv = *reinterpret_cast<const float4*>(p + i);
      float4 mv = *reinterpret_cast<const float4*>(m + i);
      float4 vv = *reinterpret_cast<const float4*>(v + i);
      float gg[4];
      if (GRAD_FP32) {
        const float4 gv = *reinterpret_cast<const float4*>(static_cast<const float*>(g) + i);
        gg[0] = gv.x; gg[1] = gv.y; gg[2] = gv.z; gg[3] = gv.w;
      } else {
        const uint2 gv = *reinterpret_cast<const uint2*>(static_cast<const __nv_bfloat16*>(g) + i);
        gg[0] = bf16_lo(gv.x); gg[1] = bf16_hi(gv.x); gg[2] = bf16_lo(gv.y); gg[3] = bf16_hi(gv.y);
      }
      float pp[4] = {pv.x, pv.y, pv.z, pv.w}, mm[4] = {mv.x, mv.y, mv.z, mv.w},
            vs[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float gk = gg[k] * a.grad_scale;
        if (!a.adamw_mode && a.weight_decay != 0.f) gk += a.weight_decay * pp[k];
        mm[k] = a.beta1 * mm[k] + (1.f - a.beta1) * gk;
        vs[k] = a.beta2 * vs[k] + (1.f - a.beta2) * gk * gk;
        const float denom = sqrtf(vs[k]) * inv_sqrt_c2 + a.eps;
        if (a.adamw_mode && a.weight_decay != 0.f) pp[k] *= (1.f - a.lr * a.weight_decay);
        pp[k] -= step_size * (mm[k] / denom);
      }
      *reinterpret_cast<float4*>(p + i) = make_float4(pp[0], pp[1], pp[2], pp[3]);
      *reinterpret_cast<float4*>(m + i) = make_float4(mm[0], mm[1], mm[2], mm[3]);
      *reinterpret_cast<float4*>(v + i) = make_float4(vs[0], vs[1], vs[2], vs[3]);
      if (p_lp != nullptr)
        *reinterpret_cast<uint2*>(p_lp + i) =
            make_uint2(pack_bf16x2(pp[0], pp[1]), pack_bf16x2(pp[2], pp[3]));
    } else {
      for (size_t j = i; j < n; ++j) {
        float gk = (GRAD_FP32 ? static_cast<const float*>(g)[j]
                              : __bfloat162float(static_cast<const __nv_bfloat16*>(g)[j])) *
                   a.grad_scale;
        float pj = p[j];
        if (!a.adamw_mode && a.weight_decay != 0.f) gk += a.weight_decay * pj;
        const float mj = a.beta1 * m[j] + (1.f - a.beta1) * gk;
        const float vj = a.beta2 * v[j] + (1.f - a.beta2) * gk * gk;
        const float denom = sqrtf(vj) * inv_sqrt_c2 + a.eps;
        if (a.adamw_mode && a.weight_decay != 0.f) pj *= (1.f - a.lr * a.weight_decay);
        pj -= step_size * (mj / denom);
        p[j] = pj; m[j] = mj; v[j] = vj;
        if (p_lp != nullptr) p_lp[j] = __float2bfloat16(pj);
      }
    }
  }
}
__global__ void adam_step_inc_kernel(int* step) { *step += 1; }

// fp32 <-> bf16 flat casts (initial master-weight copy, checkpoint export).
__global__ void cast_f32_to_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                        size_t n) {
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x)
    dst[i] = __float2bfloat16(src[i]);
}
__global__ void cast_bf16_to_f32_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst,
                                        size_t n) {
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<size_t>(gridDim.x) * blockDim.x)
    dst[i] = __bfloat162float(src[i]);
}

}  // namespace dtds

using namespace dtds;

extern "C" int dtds_embedding_fwd(const long long* ids, const long long* type_ids, const void* word,
                                  const void* pos, const void* type, void* out, int tokens, int seq,
                                  int hidden, int pos_offset, int vocab, cudaStream_t stream) {
  if (tokens <= 0) return 0;
  if (hidden % 8 != 0) return -1;
  embedding_fwd_kernel<<<(tokens + 7) / 8, 256, 0, stream>>>(
      reinterpret_cast<const int64_t*>(ids), reinterpret_cast<const int64_t*>(type_ids),
      static_cast<const __nv_bfloat16*>(word), static_cast<const __nv_bfloat16*>(pos),
      static_cast<const __nv_bfloat16*>(type), static_cast<__nv_bfloat16*>(out), tokens, seq,
      hidden, pos_offset, vocab);
  DTDS_HOST_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtds_embedding_bwd(const long long* ids, const void* dout, void* dword, void* dpos,
                                  int batch, int seq, int hidden, int pos_offset, int vocab,
                                  int grads_fp32, int accumulate_pos, cudaStream_t stream) {
  const int tokens = batch * seq;
  if (tokens <= 0) return 0;
  if (dword != nullptr) {
    embedding_bwd_word_kernel<<<(tokens + 7) / 8, 256, 0, stream>>>(
        reinterpret_cast<const int64_t*>(ids), static_cast<const __nv_bfloat16*>(dout), dword,
        tokens, hidden, vocab, grads_fp32);
    DTDS_HOST_CHECK(cudaGetLastError());
  }
  if (dpos != nullptr) {
    const size_t n = static_cast<size_t>(seq) * hidden;
    embedding_bwd_pos_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, stream>>>(
        static_cast<const __nv_bfloat16*>(dout), dpos, batch, seq, hidden, pos_offset, grads_fp32,
        accumulate_pos);
    DTDS_HOST_CHECK(cudaGetLastError());
  }
  return 0;
}

// `n_valid_dev` (optional, one float on the device): normalise by this count instead of the
// number of non-ignored labels among `rows` -- a micro-batch of a pipelined batch divides by the
// whole batch's count, so the micro-batch losses simply add up and no count travels to the host.
extern "C" int dtds_cross_entropy_ex(const void* logits, const long long* labels, void* dlogits,
                                     float* stats /* [2], written by this call */, float* row_loss,
                                     int rows, int vocab, int ld, int seq, int shift,
                                     long long ignore_index, float gscale, const float* n_valid_dev,
                                     cudaStream_t stream) {
  if (rows <= 0) return 0;
  if (ld % 8 != 0) {
    fprintf(stderr, "[dtds] cross_entropy: ld (%d) must be a multiple of 8\n", ld);
    return -1;
  }
  DTDS_HOST_CHECK(cudaMemsetAsync(stats, 0, 2 * sizeof(float), stream));
  if (n_valid_dev != nullptr)
    DTDS_HOST_CHECK(cudaMemcpyAsync(stats, n_valid_dev, sizeof(float), cudaMemcpyDeviceToDevice, stream));
  else
    ce_count_kernel<<<8, 256, 0, stream>>>(reinterpret_cast<const int64_t*>(labels), rows, seq, shift,
                                           ignore_index, stats);
  DTDS_HOST_CHECK(cudaGetLastError());
  ce_fwd_bwd_kernel<<<rows, kCeThreads, 0, stream>>>(
      static_cast<const __nv_bfloat16*>(logits), reinterpret_cast<const int64_t*>(labels),
      static_cast<__nv_bfloat16*>(dlogits), stats, row_loss, rows, vocab, ld, seq, shift,
      ignore_index, gscale);
  DTDS_HOST_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int dtds_cross_entropy(const void* logits, const long long* labels, void* dlogits,
                                  float* stats /* [2], zeroed by this call */, float* row_loss,
                                  int rows, int vocab, int ld, int seq, int shift,
                                  long long ignore_index, float gscale, cudaStream_t stream) {
  return dtds_cross_entropy_ex(logits, labels, dlogits, stats, row_loss, rows, vocab, ld, seq, shift,
                               ignore_index, gscale, nullptr, stream);
}

extern "C" int dtds_adam_flat(float* p, void* p_lp, const void* g, float* m, float* v, size_t n,
                              int* step, int inc_step, int grad_fp32, float lr, float beta1,
                              float beta2, float eps, float weight_decay, float grad_scale,
                              int adamw_mode, int bias_correction, cudaStream_t stream) {
  if (n > 0) {
    if ((reinterpret_cast<uintptr_t>(p) & 15) || (reinterpret_cast<uintptr_t>(m) & 15) ||
        (reinterpret_cast<uintptr_t>(v) & 15) ||
        (reinterpret_cast<uintptr_t>(g) & (grad_fp32 ? 15 : 7)) ||
        (p_lp != nullptr && (reinterpret_cast<uintptr_t>(p_lp) & 7))) {
      fprintf(stderr, "[dtds] adam_flat: buffers must be 16-byte aligned\n");
      return -1;
    }
    AdamArgs a{lr, beta1, beta2, eps, weight_decay, grad_scale, adamw_mode, bias_correction};
    size_t blocks = (n / 4 + 255) / 256;
    const size_t cap = static_cast<size_t>(dtds::sm_count()) * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    if (grad_fp32)
      adam_flat_kernel<true><<<static_cast<unsigned>(blocks), 256, 0, stream>>>(
          p, static_cast<__nv_bfloat16*>(p_lp), g, m, v, n, step, a);
    else
      adam_flat_kernel<false><<<static_cast<unsigned>(blocks), 256, 0, stream>>>(
          p, static_cast<__nv_bfloat16*>(p_lp), g, m, v, n, step, a);
    DTDS_HOST_CHECK(cudaGetLastError());
  }
  if (inc_step) {
    adam_step_inc_kernel<<<1, 1, 0, stream>>>(step);
    DTDS_HOST_CHECK(cudaGetLastError());
  }
  return 0;
}

extern "C" int dtds_cast(const void* src, void* dst, size_t n, int to_bf16, cudaStream_t stream) {
  if (n == 0) return 0;
  size_t blocks = (n + 255) / 256;
  const size_t cap = static_cast<size_t>(dtds::sm_count()) * 16;
  if (blocks > cap) blocks = cap;
  if (to_bf16)
    cast_f32_to_bf16_kernel<<<static_cast<unsigned>(blocks), 256, 0, stream>>>(
        static_cast<const float*>(src), static_cast<__nv_bfloat16*>(dst), n);
  else
    cast_bf16_to_f32_kernel<<<static_cast<unsigned>(blocks), 256, 0, stream>>>(
        static_cast<const __nv_bfloat16*>(src), static_cast<float*>(dst), n);
  DTDS_HOST_CHECK(cudaGetLastError());
  return 0;
}
