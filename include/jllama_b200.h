/* C ABI of libjllama_b200.so — hand-written sm_100a kernels for LLaMA inference on NVIDIA B200.
 *
 * Conventions
 *  - every entry point returns 0 on success, non-zero on failure; jl_last_error() describes the failure;
 *  - `stream` is a cudaStream_t passed as void*; launches are asynchronous on it and are legal inside
 *    CUDA-graph stream capture unless stated otherwise;
 *  - activations / weights are bf16, row-major; linear weights are [N, K] (out, in) i.e. K-major;
 *  - kernels are launched with Programmatic Dependent Launch enabled and are safe in PDL chains.
 *
 * Each declaration cites the reference op it replaces (paths under /root/reference/jax_llama/).
 */
#ifndef JLLAMA_B200_H
#define JLLAMA_B200_H
#include <stddef.h>
#ifndef JL_API
#define JL_API __attribute__((visibility("default")))
#endif
#ifdef __cplusplus
extern "C" {
#endif

const char* jl_last_error(void);
int jl_version(void);
/* number of kernel launches issued through this library since load (bench.py reports it) */
unsigned long long jl_launch_count(void);

/* model.py:611-617,638 — embedding lookup. ids int32 [M]; table [V,H]; out [M,H]. */
int jl_embed(const int* ids, const void* table, void* out, int M, int H, int V, void* stream);

/* model.py:42-48 — RMSNorm over rows of x [M,H]; optional fused residual add: if `delta` is given the
 * row normalised is x+delta and (if x_out != NULL) the sum is stored to x_out (may alias x). */
int jl_rmsnorm(const void* x, const void* delta, const void* w, void* out, void* x_out, int M, int H,
               float eps, void* stream);

/* model.py:58-92 + 177-199 — interleaved-pair RoPE on the q|k column blocks of the fused qkv activation
 * [B*S, (nh+2nkv)*hd] and append of k,v to this layer's cache [B][nkv][T][hd].  Token (b,s) goes to slot
 * *cache_len_ptr + s at rotary position slot - pad_left[b] (or positions[b*S+s] if given).  q is rotated
 * in place, or written to q_out [B*S, nh*hd] when q_out != NULL. */
int jl_rope_kv_write(void* qkv, void* q_out, void* k_cache, void* v_cache, const float* cos_t,
                     const float* sin_t, const int* cache_len_ptr, const int* pad_left, const int* positions,
                     int B, int S, int nh, int nkv, int hd, int T, int max_pos, void* stream);

/* model.py:338 — silu(gate)*up.  in [M,2F] (interleaved pairs if `interleaved`, else two halves). */
int jl_swiglu(const void* in, void* out, int M, int F, int interleaved, void* stream);

/* model.py:119-150,311-338,699-706 at decode — y = f(x) W^T for M <= 8 rows, weight-streaming bound.
 * norm_w != NULL fuses RMSNorm(x)*norm_w; mode 0 plain, 1 out = residual + y, 2 SwiGLU over interleaved
 * (w1,w3) rows (out is [M,N/2]); out_f32 selects fp32 output (plain mode). */
int jl_gemv(const void* x, const void* w, const void* norm_w, const void* residual, void* out, int M, int N,
            int K, float eps, int mode, int out_f32, int num_sms, void* stream);

/* jl_gemv for M <= 4 with the weights streamed by TMA bulk copies (cp.async.bulk + mbarrier) through a
 * shared-memory ring that starts filling before the programmatic-dependent-launch wait.  Returns 3
 * (unsupported, nothing launched) when the activations leave no room for the ring: call jl_gemv then. */
int jl_gemv_tma(const void* x, const void* w, const void* norm_w, const void* residual, void* out, int M, int N,
                int K, float eps, int mode, int out_f32, int num_sms, void* stream);

/* Same linear layers for any M on the tcgen05 tensor cores (TMA-fed, TMEM accumulators).  M > 128 runs
 * token-major tiles (prefill); M <= 128 runs swap-AB so the weight rows fill the 128-wide MMA and the
 * batch is the MMA N (decode at batch 9..128).  Modes / out_f32 as in jl_gemv.  N % 8 == 0, K % 8 == 0;
 * SwiGLU needs N % 16 == 0. */
int jl_gemm(const void* x, const void* w, const void* residual, void* out, int M, int N, int K, int mode,
            int out_f32, int num_sms, void* stream);
/* Optional scratch for jl_gemm on the CURRENT device: with it, small-batch (M <= 128) launches that have fewer
 * 128-row weight tiles than half the SMs (N = 4096: 32 tiles) split K over the idle SMs; partial sums meet in `ws`
 * (fp32, ws_bytes); once all splits of a tile have arrived (`tickets`: n_tickets ints, zero on entry and between
 * calls, half of them arrival and half completion counters) each split adds the partials of its share of the tokens
 * in split order and runs the epilogue — results do not depend on arrival order.  The splits of a tile wait for each
 * other, so such a launch needs its CTAs resident together (at most one per SM is launched).  The caller owns both
 * buffers and keeps them alive; ws == NULL unregisters.  Calls that share the scratch must be ordered on one stream. */
int jl_gemm_set_workspace(void* ws, size_t ws_bytes, int* tickets, int n_tickets);
/* Timeline hook of the single-CTA GEMM kernel for scripts/gemm_trace.py: [grid][16] u64 device buffer of
 * %globaltimer stamps, NULL = off (default). */
void jl_gemm_set_trace(void* buf);

/* model.py:194-198,263-288 at S=1 — split-KV GQA decode attention over [pad_left[b], *cache_len+n_new).
 * q [B,nh,hd] with batch pitch q_bstride elements (0 = dense; lets q live inside the fused qkv
 * activation); caches [B][nkv][T][hd]; out [B,nh*hd].  `scratch` must hold
 * jl_decode_attn_scratch_bytes(B,nh,hd,splits) bytes when splits > 1. */
size_t jl_decode_attn_scratch_bytes(int B, int nh, int hd, int splits);
int jl_decode_attn_splits(int B, int nkv, int max_len, int num_sms);
int jl_decode_attn(const void* q, const void* k_cache, const void* v_cache, const int* cache_len_ptr,
                   const int* pad_left, void* scratch, void* out, int B, int nh, int nkv, int hd, int T,
                   int n_new, float scale, int splits, int q_bstride, void* stream);

/* model.py:201-300 at S=1 in ONE launch — RoPE of q and of the new token's k, append of (k,v) at slot
 * *cache_len, split-KV GQA attention over [pad_left[b], *cache_len] and the split merge (last-arriving CTA,
 * atomic ticket).  qkv [B,(nh+2nkv)*hd] is the un-rotated fused projection; out [B,nh*hd].  `scratch` must
 * hold jl_decode_attn_fused_scratch_bytes(...) bytes and be ZERO when first used (ticket counters; the kernel
 * leaves them zero).  *cache_len is not advanced here. */
size_t jl_decode_attn_fused_scratch_bytes(int B, int nh, int nkv, int hd, int splits);
int jl_decode_attn_fused_splits(int B, int nkv, int max_len, int num_sms);
int jl_decode_attn_fused(const void* qkv, void* k_cache, void* v_cache, const float* cos_t, const float* sin_t,
                         const int* cache_len_ptr, const int* pad_left, void* scratch, void* out, int B, int nh,
                         int nkv, int hd, int T, int max_pos, float scale, int splits, void* stream);
/* Two kernels stand behind jl_decode_attn_fused: a SIMT one with many short splits (lowest latency at batch 1) and,
 * for head_dim 128 with batch * nkv >= 16, a tensor-core one (mma.sync, K/V from global memory straight into MMA
 * fragments) that streams the cache ~3x faster at batch 16-64.  variant: 0 = choose by shape, 1 = SIMT, 2 = tensor
 * core (where it applies).  Process-wide; meant for tests and A/B measurements. */
int jl_decode_attn_fused_set_variant(int variant);

/* No counterpart in the reference: an allocation-free form of jl_decode_attn for callers that cannot get
 * device memory themselves (HBM fully owned by another runtime's arena, hosts that ban cudaMalloc/cudaFree
 * inside the token loop, bare C callers without an allocator).  Same arguments minus `scratch`: the split-KV
 * workspace is a fixed 4 MiB array inside the library image, resident on every GPU that loads the library
 * until process exit.  B*nh*splits*(hd+2) floats must not exceed jl_decode_attn_noalloc_max_floats();
 * otherwise the call fails with a message naming the limit.  It calls no allocating routine.  The first call
 * on each device looks the array's address up — do it before any stream capture.  All calls share the array
 * (it keeps the previous call's contents; nothing needs zeroing), so calls that may overlap on the GPU race:
 * keep them on one stream.  Output equals jl_decode_attn's for the same `splits`. */
size_t jl_decode_attn_noalloc_max_floats(void);
int jl_decode_attn_noalloc(const void* q, const void* k_cache, const void* v_cache, const int* cache_len_ptr,
                           const int* pad_left, void* out, int B, int nh, int nkv, int hd, int T, int n_new,
                           float scale, int splits, int q_bstride, void* stream);

/* model.py:236-288 at S>1 — causal flash attention of S new tokens per row (cache slots start..start+S,
 * already appended) on tcgen05: q [B*S, nh*hd] rotated, caches [B][nkv][T][hd], out [B*S, nh*hd].
 * Keys [pad_left[b], slot of the query] are visible.  head_dim 128 only.  cache_len_ptr is reserved. */
int jl_flash_prefill_occupancy(int smem_override /* -1 */); /* resident CTAs per SM (diagnostic; 2 expected) */
int jl_flash_prefill(const void* q, const void* k_cache, const void* v_cache, const int* cache_len_ptr,
                     const int* pad_left, void* out, int B, int S, int nh, int nkv, int hd, int T, int start,
                     float scale, void* stream);
/* Same contract on warp-level tensor cores (mma.sync, K / V from global memory straight into fragments, one warp per
 * 16 query rows): head_dim 64 and 128.  jl_flash_prefill routes head_dim 64 here (model.py:115 allows any
 * hidden / heads; the 1B-class LLaMA-architecture models have 64-wide heads). */
int jl_prefill_attn_mma(const void* q, const void* k_cache, const void* v_cache, const int* pad_left, void* out, int B,
                        int S, int nh, int nkv, int hd, int T, int start, float scale, void* stream);

/* generation.py:28-41 (HF warpers + categorical/argmax) — logits fp32 [B, ld] (V valid columns). */
int jl_argmax(const float* logits, int B, int V, int ld, int idx_offset, float* out_val, int* out_idx,
              void* stream);
int jl_topk(const float* logits, int B, int V, int ld, int idx_offset, int k, float* cand_val, int* cand_idx,
            void* stream);
int jl_sample(const float* cand_val, const int* cand_idx, int B, int n_lists, size_t list_stride, int k,
              int top_k, float top_p, float temperature, unsigned long long seed, const int* step_ptr,
              int* out_tok, void* stream);
/* HF _sample/_greedy_search body: pad finished rows, mark EOS, append to sequences,
 * advance cur_len / cache_len / step on device. */
int jl_decode_advance(const int* tok, int* next_ids, int* finished, int* sequences, int B, int max_len,
                      int eos, int pad, int* cur_len, int* cache_len, int* step, int* n_unfinished,
                      void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Whole-model decode step (model.py:743-790 at S=1, i.e. everything FlaxLLaMAForCausalLMModule.__call__ does
 * for one new token) in ONE persistent cooperative kernel for M <= 8 rows: embedding, L x (RMSNorm+wqkv, RoPE,
 * cache append, GQA attention, wo+residual, RMSNorm+w13+SwiGLU, w2+residual), final norm and LM head with a
 * running arg-max.  Weights are the per-op stacks [L*N, K].  There is no grid barrier: activations cross CTAs as
 * tagged 8-byte packets that consumers poll (csrc/decode_mega.cu).  `scratch` holds the packet buffers
 * (jl_decode_mega_scratch_bytes() bytes, ZERO when first used, never cleared afterwards); `gen` is a device
 * counter of the steps this engine has run (zero at start; the kernel advances it; it is the tag base, so every
 * scratch / peer buffer ever used with it must only be used with it).  `logits` fp32 [M, V] (V = this rank's
 * vocab shard, vocab_full the whole vocabulary) may be NULL: then only the per-CTA arg-max candidates are
 * produced (read by jl_mega_select).  Row pitch `logits_ld` (0 = V), this rank's columns start at `logits_col0`;
 * with `logits_multicast` the pointer is the NVLS multicast address of a symmetric [M, vocab_full] buffer and every
 * rank receives every shard (input of jl_sample_full under tensor parallelism).  Needs a cooperative launch of one CTA per SM (num_sms CTAs co-resident).
 * With tp > 1 the wo / w2 outputs are all-reduced inside the kernel (partition.py:67,70 is what GSPMD
 * all-reduced there): fp32 partial packets go to every rank with one multimem.st on `mc_partials` (the NVLS
 * multicast address of the symmetric buffer whose per-rank mappings are `peer_partials`; NULL = one unicast
 * store per peer), candidates likewise through `peer_cand` / `mc_cand`; sizes from jl_decode_mega_peer_bytes. */
size_t jl_decode_mega_scratch_bytes(int M, int H, int F, int nh, int nkv, int num_sms);
size_t jl_decode_mega_cand_offset(int M, int H, int F, int nh, int nkv, int num_sms);   /* tp == 1: candidates inside scratch */
size_t jl_decode_mega_peer_bytes(int M, int H, int tp, int num_sms, size_t* cand_offset);
int jl_decode_mega_stages(int M, int H, int F, int nh, int nkv, int hd); /* weight-ring depth; 0 = shape does not fit */
void jl_decode_mega_set_trace(void* buf /* device u64 [num_sms][L][32] %globaltimer stamps, NULL = off */);
int jl_decode_mega(const void* wqkv_all, const void* wo_all, const void* w13_all, const void* w2_all,
                   const void* lm_head, const void* attn_norm, const void* ffn_norm, const void* final_norm,
                   const void* embedding, const int* ids, float* logits, int logits_ld, int logits_col0,
                   int logits_multicast, void* k_cache, void* v_cache,
                   const float* cos_t, const float* sin_t, const int* cache_len, const int* pad_left, void* scratch,
                   unsigned int* gen, int M, int L, int H, int F, int V, int vocab_full, int nh, int nkv, int hd,
                   int T, int max_pos, float eps, float scale, int num_sms, int tp, int rank,
                   void* const* peer_partials, void* mc_partials, void* const* peer_cand, void* mc_cand,
                   void* stream);
/* HF _greedy_search body (generation.py:28-41 with temperature 0) for the step jl_decode_mega just ran on this
 * stream: merges the num_sms x tp arg-max candidates per row (ties: lower vocabulary index; waits for ranks whose
 * kernel is still finishing), then the bookkeeping of jl_decode_advance.  `cand` = this rank's candidate packets. */
int jl_mega_select(const void* cand, const unsigned int* gen, int tp, int rank, int num_sms, int M, int L, int* tok,
                   int* next_ids, int* finished, int* sequences, int max_len, int eos, int pad, int* cur_len,
                   int* cache_len, int* step, int* n_unfinished, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Tensor parallelism over NVLink peer memory (partition.py:62-78 names the sharding; XLA/GSPMD inserted NCCL
 * collectives, here they run inside the compute kernels).  Peer buffers are cudaMalloc'd + exported with
 * CUDA IPC; `peer_*` arguments are host arrays of tp device pointers (index = rank, own entry local). */
int jl_ipc_alloc(size_t bytes, void** ptr, void* handle64 /* out: 64-byte IPC handle */);
int jl_ipc_open(const void* handle64, void** ptr);
int jl_ipc_close(void* ptr);
int jl_ipc_free(void* ptr);

/* Row-parallel linear at decode (wo, w2; M <= 8): local GEMV over this rank's K shard, partial sums pushed to
 * every peer's inbox, one-shot all-reduce in rank order (bitwise identical on all ranks) + residual, in one
 * launch.  The epoch that tags the flags is *seq_ptr + call_index (device counter advanced with jl_comm_bump,
 * so the same CUDA graph replays with fresh epochs).  peer_flags must hold jl_gemv_allreduce_flags(N) ints at
 * flag_offset. */
int jl_gemv_allreduce_flags(int N);
int jl_gemv_allreduce(const void* x, const void* w, const void* residual, void* out, int M, int N, int K, int tp,
                      int rank, void* const* peer_inbox, void* const* peer_flags,
                      size_t inbox_parity_stride_floats, int flag_offset, const int* seq_ptr, int call_index,
                      int num_sms, void* stream);
/* All-gather of a small record (<= slot_bytes, multiples of 16) from every rank: push to every inbox + flag.
 * Used for the cross-rank sampler (per-rank top-k candidates of the vocab-sharded logits). */
int jl_allgather_small(const void* src, void* out, int n_bytes, int slot_bytes, int tp, int rank,
                       void* const* peer_inbox, void* const* peer_flags, size_t parity_stride_bytes,
                       int flag_offset, const int* seq_ptr, int call_index, void* stream);
int jl_comm_bump(int* seq_ptr, int n, void* stream);
/* one warp: flags[r][flag_off + rank] = epoch on every rank r, then wait until this rank's flags[flag_off + 0..tp) >= epoch
 * ("every rank has finished the kernel before this one on its stream") */
int jl_flag_exchange(void* const* peer_flags, int flag_off, int epoch, const int* base /* device int added to epoch, or NULL */,
                     int tp, int rank, void* stream);
/* one warp waits (on the stream) until flags[0..tp) >= epoch (+ *seq_ptr if given: device epoch counter, graph safe):
 * later kernels may read what the peers published first */
int jl_wait_flags(const int* flags, int epoch, const int* seq_ptr, int tp, int rank, void* stream);
int jl_memcpy_d2d(void* dst, const void* src, size_t bytes, void* stream);

/* HF Flax warpers + categorical sampling + `_sample` bookkeeping (generation.py:28-41) over the WHOLE vocabulary, exact
 * for any top_k / top_p (no candidate budget), deterministic ties (lower index), identical on every rank: see
 * csrc/sample_full.cu.  logits fp32 [B, ld] with V valid columns; temperature 0 = arg-max; top_k <= 0 or >= V and
 * top_p >= 1 disable the filters.  `cand`/`gen` (may be NULL) = candidate packets / step counter of jl_decode_mega: wait
 * for every rank's logits shard first.  `next_ids` == NULL: selection only (tok[B]).  `ticket` = one zero int. */
int jl_sample_full(const float* logits, int B, int V, int ld, float temperature, int top_k, float top_p,
                   unsigned long long seed, const int* step, int* tok, const void* cand, const unsigned int* gen, int tp,
                   int rank, int num_sms, int L, int* ticket, int* next_ids, int* finished, int* sequences, int max_len,
                   int eos, int pad, int* cur_len, int* cache_len, int* step_rw, int* n_unfinished, void* stream);

/* All-gather of a column-sharded fp32 block over NVLS: this rank's [rows, cols] goes to columns [col0, col0+cols) of
 * every rank's [rows, dst_ld] buffer (`dst_mc` = its multicast address) with multimem.st; flags[r][flag_off + rank] =
 * epoch (+ *seq_ptr if given) on every rank afterwards (pair with jl_wait_flags).  The reference gathers the vocab-sharded logits with the
 * all-gather GSPMD derives from partition.py:77. */
int jl_nvls_bcast_f32(const float* src, float* dst_mc, int rows, int cols, int src_ld, int dst_ld, int col0, int tp,
                      int rank, void* const* peer_flags, int flag_off, int epoch, const int* seq_ptr, int* ticket,
                      void* stream);

/* The same linear layers on CTA PAIRS (2-CTA clusters, tcgen05.mma.cta_group::2, 256 x 256 tiles, TMA-store epilogue):
 * token-major GEMMs with M >= 256, N % 256 == 0, K % 64 == 0 (every prefill GEMM of model.py:119-150, 311-338).
 * Returns 3 (nothing launched) for other shapes; jl_gemm tries this first and falls back to its single-CTA kernel. */
int jl_gemm_2cta(const void* x, const void* w, const void* residual, void* out, int M, int N, int K, int mode,
                 int num_sms, void* stream);
int jl_gemm_2cta_pairs(void);   /* 2-CTA clusters resident at once on this GPU (the persistent grid uses exactly these) */

/* NVLS form of the owner-side step of the sequence-parallel prefill, and the all-reduce of a row-parallel GEMM for any
 * token count (partition.py:67,70 is what GSPMD all-reduced there).  The GEMM before it stored this rank's bf16 partial
 * product into its own copy of a symmetric buffer; `part_mc` is the multicast address of that buffer at the first of
 * the `rows` rows this call owns: x[rows, H] += sum over ranks (multimem.ld_reduce, fp32 in the switch).  With `xn_mc`
 * (multicast address of the gathered-activation buffer at the same first row) the updated rows are RMS-normalised and
 * stored into every rank's copy with multimem.st, then the "rows ready" flag is published; without it the call is a
 * plain all-reduce into a replicated residual stream (rows = all tokens).  With `xn_local` instead (an ordinary
 * [rows, H] buffer) the normalised rows stay on this rank: all-reduce + residual add + the next RMSNorm of a batched
 * decode step in one launch.  has_partials: 0 = no reduction (chain start), 2 = the caller already exchanged the "GEMM
 * done" flags (jl_flag_exchange), 1 = exchange them inside this kernel (rows <= 128 only: a small resident grid). */
int jl_rs_reduce_mm(void* x, const void* part_mc, const void* norm_w, void* xn_mc, void* xn_local, int rows, int H,
                    int has_partials, int tp, int rank, void* const* peer_flags, int gemm_flag_off, int ready_flag_off,
                    int gemm_epoch, int ready_epoch, const int* epoch_base /* device int added to both epochs, or NULL */,
                    int* ticket, float eps, void* stream);

/* NVLS self-test over a multicast mapping of a symmetric buffer (parallel/symm.py): `mc` is the multicast
 * address.  op 0: out[i] = sum over ranks of fp32 buf[i] (multimem.ld_reduce); 1: 8-byte packets {src[i], tag}
 * stored into every rank's copy (multimem.st); 2: every copy += src[i] (multimem.red); 3 / 4: the bf16 forms of
 * 0 / (plain store); 5: u32 arrival counter at buf[0] += 1 with release semantics.  The reference's collectives
 * are NCCL calls emitted by XLA for partition.py:62-78; these instructions are what the fused kernels use instead. */
int jl_nvls_selftest(void* mc, const void* src, void* out, int n, int op, unsigned int tag, void* stream);

#ifdef __cplusplus
}
#endif
#endif
