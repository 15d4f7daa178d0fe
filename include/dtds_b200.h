/* C ABI of libdtds_b200.so — the hand-written sm_100a kernels and the symmetric-memory runtime.
 *
 * Every function returns 0 on success and a non-zero code (a cudaError_t value, or a small
 * negative number for argument errors, with a message on stderr) otherwise.  All device pointers
 * are plain `void*`; bf16 tensors are row-major unless noted; `stream` is a cudaStream_t.
 * Python reaches these through ctypes (distributed-training-and-deepspeed_b200/ops/_lib.py).
 *
 * Each block cites what it replaces in the reference (paths relative to /root/reference).
 */
#ifndef DTDS_B200_H_
#define DTDS_B200_H_

#include <cuda_runtime.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Activation codes shared by the GEMM epilogue and the standalone kernels. */
#define DTDS_ACT_NONE 0
#define DTDS_ACT_RELU 1      /* model/transformer.py:52 (OPT) */
#define DTDS_ACT_GELU_ERF 2  /* HF BERT hidden_act="gelu" (model/bert_mp.py:22-24) */
#define DTDS_ACT_GELU_TANH 3 /* BLOOM (zero_dp_training.py:117) */

/* Added to `block_n` (128 or 256) of dtds_gemm_bf16_ex: 256 x block_n tiles computed by a pair of
 * CTAs on neighbouring SMs (tcgen05 cta_group::2), each staging half of the B tile. */
#define DTDS_GEMM_PAIR 1024

/* ---- tcgen05 / TMEM / TMA GEMM (replaces cuBLAS behind nn.Linear: model/transformer.py:37-40,
 *      50-51; BertLayer / BertOnlyMLMHead in model/bert_mp.py:22-24) -------------------------
 * D[M,N] = act(alpha * A·Bᵀ + bias) (+ residual), A logical [M,K], B logical [N,K].
 * x_mn_major = 0: the operand is stored with K contiguous; 1: stored as [K, M|N] (M|N contiguous).
 * Leading dimensions are in elements and must be multiples of 8; pointers 16-byte aligned.   */
int dtds_gemm_bf16(const void* A, const void* B, void* D, const void* bias, int M, int N, int K,
                   int lda, int ldb, int ldd, int a_mn_major, int b_mn_major, int act,
                   int out_fp32, cudaStream_t stream);
/* Full form.  pre_out: also store the pre-activation (bf16, same ld as D).  dact_in: multiply the
 * result by act'(dact_in) (backward through an activation).  accumulate: D += result.
 * block_n: 0 = auto, or 64/128/256.  max_ctas: 0 = all SMs (lower it to leave SMs to a
 * concurrent communication kernel).  ready_flags/ready_epoch/rows_per_flag: all-gather fused
 * mode — rows [f*rows_per_flag, (f+1)*rows_per_flag) of B may only be read once
 * ready_flags[f] >= ready_epoch (see dtds_allgather_rows_async).                              */
int dtds_gemm_bf16_ex(const void* A, const void* B, void* D, const void* bias,
                      const void* residual, void* pre_out, const void* dact_in, int M, int N,
                      int K, int lda, int ldb, int ldd, int ld_res, int ld_dact, int a_mn_major,
                      int b_mn_major, int act, int out_fp32, int accumulate, float alpha,
                      int block_n, int max_ctas, const void* ready_flags, unsigned ready_epoch,
                      int rows_per_flag, cudaStream_t stream);
/* ZeRO-3: parameter all-gather fused with the consuming GEMM (zero_dp_training.py --stage 3;
 * DeepSpeed gathers with all_gather_into_tensor and then calls cuBLAS).  B_local is filled by
 * fetch warps of the GEMM kernel from `nseg` source segments (peer-mapped shard addresses;
 * segment i covers bytes [seg_begin[i], seg_begin[i+1]) of B) in chunks of `chunk_bytes` (whole
 * rows); flags[chunk] is raised to `epoch` when a chunk has landed and the TMA producers wait on
 * it, so transfer and tensor-core work overlap tile by tile.  b_mn_major = 1 is the dgrad form
 * (dX = dY * W, flags indexed along K).  `epoch` must increase from call to call.              */
int dtds_gemm_bf16_allgather(const void* A, void* B_local, void* D, const void* bias,
                             const void* residual, int M, int N, int K, int lda, int ldb, int ldd,
                             int ld_res, int b_mn_major, int act, int block_n,
                             const unsigned long long* seg_src, const unsigned long long* seg_begin,
                             int nseg, unsigned chunk_bytes, void* flags, unsigned epoch, unsigned long long local_begin,
                             cudaStream_t stream);

/* Development aid: per-CTA clock64() timeline of the GEMM kernel into buf[grid][8]; NULL = off. */
int dtds_gemm_set_debug_buffer(long long* buf);

/* ---- LayerNorm (+dropout +residual), activations, dropout, softmax, column sums, MSE
 *      (ATen kernels behind model/transformer.py:36,43-53; BERT post-LN blocks) ------------- */
int dtds_layernorm_fwd(const void* x, const void* residual, const void* gamma, const void* beta,
                       void* y, void* z_out, float* mean, float* rstd, int rows, int cols,
                       float eps, float p_drop, unsigned long long seed,
                       const unsigned long long* seed_dev, cudaStream_t stream);
int dtds_layernorm_bwd_blocks(int rows, int cols); /* rows of `partial` the backward needs */
/* 1 if dz and dgamma/dbeta come from separate kernels at this shape: the caller may then issue
 * two calls (first with dgamma = dbeta = NULL, then with dz = NULL) on two streams. */
int dtds_layernorm_bwd_splittable(int rows, int cols);
int dtds_layernorm_bwd(const void* dy, const void* z, const float* mean, const float* rstd,
                       const void* gamma, const void* dz_extra, void* dz, void* dx_drop,
                       void* dgamma, void* dbeta, float* partial, int rows, int cols, float p_drop,
                       unsigned long long seed, const unsigned long long* seed_dev, int grads_fp32,
                       int accumulate, cudaStream_t stream);
int dtds_colsum(const void* x, int rows, int cols, int ld, float* scratch, void* out, int out_fp32,
                int accumulate, cudaStream_t stream);
int dtds_act_fwd(const void* x, void* y, size_t n, int act, cudaStream_t stream);
int dtds_act_bwd(const void* dy, const void* x, void* dx, size_t n, int act, cudaStream_t stream);
int dtds_dropout(const void* x, const void* residual, void* y, size_t n, float p,
                 unsigned long long seed, const unsigned long long* seed_dev, cudaStream_t stream);
int dtds_softmax_rows(const void* x, const void* dy, void* out, int rows, int cols,
                      cudaStream_t stream);
int dtds_mse(const void* x, const void* t, void* dx, float* loss, size_t n, float gscale,
             cudaStream_t stream); /* estimate_transformer_memory.py:114-120 */

/* ---- Embedding, cross-entropy, Adam (BertEmbeddings model/bert_mp.py:22; CrossEntropyLoss
 *      model_parallel_training.py:51,73; FusedAdam zero_dp_training.py:28-33; AdamW
 *      data_parallel_training.py:34, model_parallel_training.py:50) ------------------------- */
/* Host-side (no GPU): non-empty lines of a text file -> int64 token rows [rows, max_len] with the
 * word-hash vocabulary of utils/data.py (WordHashTokenizer), padded with pad_id; lengths[r] is the
 * unpadded length.  Multi-threaded; returns rows written, -1 unreadable file, -2 bad arguments.
 * Stands in for the Rust tokenizers + Arrow pipeline of the reference (util.py:47-51). */
long long dtds_text_rows(const char* path, long long vocab, long long first_id, int max_len,
                         long long pad_id, long long* ids, int* lengths, long long max_rows,
                         int threads);
int dtds_embedding_fwd(const long long* ids, const long long* type_ids, const void* word,
                       const void* pos, const void* type, void* out, int tokens, int seq,
                       int hidden, int pos_offset, int vocab, cudaStream_t stream);
int dtds_embedding_bwd(const long long* ids, const void* dout, void* dword, void* dpos, int batch,
                       int seq, int hidden, int pos_offset, int vocab, int grads_fp32,
                       int accumulate_pos, cudaStream_t stream);
int dtds_cross_entropy(const void* logits, const long long* labels, void* dlogits, float* stats,
                       float* row_loss, int rows, int vocab, int ld, int seq, int shift,
                       long long ignore_index, float gscale, cudaStream_t stream);
/* Same with the normalising count supplied on the device (NULL = count the non-ignored labels). */
int dtds_cross_entropy_ex(const void* logits, const long long* labels, void* dlogits, float* stats,
                          float* row_loss, int rows, int vocab, int ld, int seq, int shift,
                          long long ignore_index, float gscale, const float* n_valid_dev,
                          cudaStream_t stream);
int dtds_adam_flat(float* p, void* p_lp, const void* g, float* m, float* v, size_t n, int* step,
                   int inc_step, int grad_fp32, float lr, float beta1, float beta2, float eps,
                   float weight_decay, float grad_scale, int adamw_mode, int bias_correction,
                   cudaStream_t stream);
int dtds_cast(const void* src, void* dst, size_t n, int to_bf16, cudaStream_t stream);

/* ---- Fused attention on tcgen05 (replaces bmm -> Softmax -> Dropout -> bmm,
 *      model/transformer.py:80-86, and HF BertSelfAttention / BLOOM / OPT attention) ----------
 * qkv: [B*S, ld_qkv] bf16 rows = tokens, columns = [Q(H) | K(H) | V(H)], H = heads*64.
 * out: [B*S, H].  lse: [B, heads, S] fp32.  key_bias: [B,S] additive fp32 or NULL (padding mask).
 * alibi_slopes: [heads] fp32 or NULL.  Requires head_dim == 64 and S % 128 == 0.
 * Backward writes dqkv [B*S, 3H]; `delta` is [B, heads, S] fp32 scratch supplied by the caller. */
int dtds_attention_fwd(const void* qkv, int ld_qkv, void* out, float* lse, const float* key_bias,
                       const float* alibi_slopes, int B, int S, int heads, int head_dim,
                       float scale, int causal, float p_drop, unsigned long long seed,
                       const unsigned long long* seed_dev, cudaStream_t stream);
int dtds_attention_bwd(const void* qkv, int ld_qkv, const void* out, const void* dout,
                       const float* lse, float* delta, void* dqkv, const float* key_bias,
                       const float* alibi_slopes, int B, int S, int heads, int head_dim,
                       float scale, int causal, float p_drop, unsigned long long seed,
                       const unsigned long long* seed_dev, cudaStream_t stream);
/* Same, launching only the selected parts: 1 = delta, 2 = dK/dV kernel, 4 = dQ kernel (the last two
 * are independent given delta and may run on different streams).                                */
int dtds_attention_bwd_parts(const void* qkv, int ld_qkv, const void* out, const void* dout,
                             const float* lse, float* delta, void* dqkv, const float* key_bias,
                             const float* alibi_slopes, int B, int S, int heads, int head_dim,
                             float scale, int causal, float p_drop, unsigned long long seed,
                             const unsigned long long* seed_dev, int parts, cudaStream_t stream);

/* ---- Symmetric-memory runtime + in-kernel collectives over NVLink peer memory (replaces
 *      ProcessGroupNCCL/NCCL behind pytorch_allreduce.py:31, the DDP Reducer all-reduce behind
 *      data_parallel_training.py:44, and DeepSpeed ZeRO's reduce/all-gather/FusedAdam sequence
 *      behind zero_dp_training.py:46-50,88-90) ------------------------------------------------
 * Every rank allocates the same sizes in the same order; buffers are exported with CUDA IPC and
 * mapped by all peers.  `data_ptrs` / `signal_ptrs` are host arrays of `world` device addresses
 * (entry i = rank i's buffer / signal pad as mapped in THIS process).  Every rank must make the
 * same call with the same sizes; kernels spin on peer flags (bounded; a stuck wait traps).     */
int dtds_symm_alloc(size_t bytes, unsigned long long* out_ptr);          /* zero-filled       */
int dtds_symm_free(unsigned long long ptr);
int dtds_symm_export(unsigned long long ptr, void* handle_out);          /* 64-byte IPC handle */
int dtds_symm_import(const void* handle, unsigned long long* out_ptr);
int dtds_symm_close(unsigned long long ptr);
size_t dtds_signal_pad_bytes(void);
/* dtype: 0 bf16, 1 fp32, 2 int64 (sum).  algo: 0 auto, 1 one-shot (needs `out` outside the
 * symmetric buffer), 2 two-shot (in place, out = NULL), 3 two-shot low-register variant that can
 * share an SM with a resident GEMM CTA (overlapped gradient buckets).  `scale` is applied once.  */
int dtds_allreduce(const unsigned long long* data_ptrs, const unsigned long long* signal_ptrs,
                   int rank, int world, void* out, size_t off_bytes, size_t nbytes, int dtype,
                   float scale, int algo, int blocks, cudaStream_t stream);
int dtds_reduce_scatter(const unsigned long long* data_ptrs, const unsigned long long* signal_ptrs,
                        int rank, int world, void* out, size_t off_bytes, size_t nbytes_per_rank,
                        int dtype, float scale, int blocks, cudaStream_t stream);
int dtds_all_gather(const unsigned long long* data_ptrs, const unsigned long long* signal_ptrs,
                    int rank, int world, const void* in, size_t off_bytes, size_t nbytes_per_rank,
                    int blocks, cudaStream_t stream);
int dtds_broadcast(const unsigned long long* data_ptrs, const unsigned long long* signal_ptrs,
                   int rank, int world, int root, size_t off_bytes, size_t nbytes, int blocks,
                   cudaStream_t stream);
int dtds_device_barrier(const unsigned long long* signal_ptrs, int rank, int world,
                        cudaStream_t stream);
/* NVLS two-shot all-reduce: multimem.ld_reduce (sum inside the NVSwitch) + multimem.st (one store
 * lands in every rank's copy) on the multicast address `mc_ptr` of a symmetric buffer; in place.
 * dtype 0 bf16 (fp32 accumulation), 1 fp32.                                                      */
int dtds_allreduce_nvls(unsigned long long mc_ptr, const unsigned long long* signal_ptrs, int rank,
                        int world, size_t off_bytes, size_t nbytes, int dtype, float scale,
                        int blocks, cudaStream_t stream);
/* ZeRO step in ONE kernel: reduce-scatter of bf16 gradients + Adam on the owned fp32 partition +
 * all-gather (push) of the updated bf16 parameters.  Offsets / sizes in elements, multiples of 8. */
int dtds_zero_step(const unsigned long long* grad_ptrs, const unsigned long long* signal_ptrs,
                   const unsigned long long* param_ptrs, int rank, int world, size_t goff_elems,
                   size_t poff_elems, size_t per_rank_elems, float* master, float* m, float* v,
                   void* param_shard, const int* step, float lr, float beta1, float beta2, float eps,
                   float weight_decay, float grad_scale, int adamw_mode, int bias_correction,
                   int push_params, int blocks, cudaStream_t stream);
/* Same kernel with the switch doing the reduction and the replication: `grads_mc` / `params_mc` are
 * the multicast addresses of the gradient / parameter buffers (0 = peer loads / peer stores).   */
int dtds_zero_step_nvls(const unsigned long long* grad_ptrs, const unsigned long long* signal_ptrs,
                        const unsigned long long* param_ptrs, int rank, int world, size_t goff_elems,
                        size_t poff_elems, size_t per_rank_elems, float* master, float* m, float* v,
                        void* param_shard, const int* step, float lr, float beta1, float beta2,
                        float eps, float weight_decay, float grad_scale, int adamw_mode,
                        int bias_correction, int push_params, int blocks,
                        unsigned long long grads_mc, unsigned long long params_mc,
                        cudaStream_t stream);
/* NVLS reduce-scatter (multimem.ld_reduce of the own slice), all-gather and broadcast (multimem.st:
 * one store, the switch replicates) on a multicast-bound buffer; same contracts as the P2P ones. */
int dtds_reduce_scatter_nvls(unsigned long long mc_ptr, const unsigned long long* signal_ptrs,
                             int rank, int world, void* out, size_t off_bytes,
                             size_t nbytes_per_rank, int dtype, float scale, int blocks,
                             cudaStream_t stream);
int dtds_all_gather_nvls(unsigned long long mc_ptr, const unsigned long long* signal_ptrs, int rank,
                         int world, const void* in, size_t off_bytes, size_t nbytes_per_rank,
                         int blocks, cudaStream_t stream);
int dtds_broadcast_nvls(unsigned long long mc_ptr, const unsigned long long* data_ptrs,
                        const unsigned long long* signal_ptrs, int rank, int world, int root,
                        size_t off_bytes, size_t nbytes, int blocks, cudaStream_t stream);
/* Multicast-bound symmetric buffers (csrc/symm_mcast.cu): cuMemCreate + cuMulticastCreate /
 * AddDevice / BindMem, file descriptors passed between the rank processes over unix sockets.
 * Call order (every rank, bootstrap barriers where noted): create -> [rank 0: root_create] ->
 * send_fds -> recv_add -> barrier -> bind_map -> barrier.  Replaces the buffers NCCL's NVLS
 * transport sets up internally for the reference's collectives (pytorch_allreduce.py:31).       */
int dtds_mc_supported(void);
int dtds_mc_create(size_t bytes, int rank, int world, size_t* bytes_out);   /* returns a buffer id */
int dtds_mc_root_create(int id);
int dtds_mc_send_fds(int id, const int* pids, const int* ids);
int dtds_mc_recv_add(int id);
int dtds_mc_bind_map(int id, unsigned long long* ptrs_out, unsigned long long* mc_ptr_out);
int dtds_mc_free(int id);
/* Pull-style gather: copies `nseg` byte ranges (local or peer sources) into local destinations in
 * one launch; ZeRO-3 uses it to materialise a block's parameters from the owners' shards.      */
int dtds_copy_segments(const unsigned long long* src_ptrs, const unsigned long long* dst_ptrs,
                       const unsigned long long* nbytes, int nseg, int blocks, cudaStream_t stream);

/* ---- Column sum that allocates nothing (no counterpart in the reference: it has no native code
 *      and no allocation-free entry points) ----------------------------------------------------
 * Same result as dtds_colsum, for callers that cannot allocate device memory where they need it:
 * code running under CUDA-graph capture, code inside a caching allocator's callbacks or
 * out-of-memory path, and C hosts that own no allocator.  The fp32 scratch is a fixed
 * __device__ array inside the library (132 KiB on every GPU that loads it, until process exit).
 * Limits and rules: cols <= 32768 (larger -> error code -2 and a message, nothing is launched);
 * the first call on each device resolves the array's address and clears it, so make that call
 * outside any stream capture -- that clear is enqueued on the FIRST call's stream only, so a second
 * stream may use the entry point only after an event or synchronisation orders it behind that
 * first call; every call uses the same array, so two calls whose kernels could overlap on the GPU
 * (different streams without ordering) race -- serialise them.  The kernel re-zeroes what it
 * touched on its way out; if a launch dies part-way (device fault, reset of a sticky error) the
 * partial sums and arrival counters it left behind stay dirty for the life of the process and
 * later results are undefined: treat a failed call as fatal for this entry point.               */
int dtds_colsum_noalloc(const void* x, int rows, int cols, int ld, void* out, int out_fp32,
                        int accumulate, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* DTDS_B200_H_ */
