// Persistent whole-model decode step ("megakernel") for small batches (M <= 8 tokens).
//
// Why: a decode step of an 8B model is ~225 dependent kernels whose useful work is 5-40 µs each.  Measured
// with ncu on B200, every one of them loses 5-10 µs to CTA launch ramp, prologue and tail (SMs active for
// only ~50% of a 19 µs `wo` GEMV), and programmatic dependent launch does not recover it because the
// streaming kernels fill the SM.  Here ONE cooperative kernel with one CTA per SM runs the entire step:
//
//   * warp 0 (of a 4-warp producer group, see MG_PWARPS) is a TMA producer that streams every weight matrix of the model, layer after layer, through a
//     shared-memory ring: one 16 KB request per stage ([16 k-blocks][8 rows][128 B], 128-byte swizzle).  It
//     depends on nothing but free ring slots, so it runs ahead across op boundaries: while the consumers poll
//     for their inputs or do attention, the next op's weights are already landing.
//   * warps 4..11 are consumers: RMSNorm + tensor-core GEMV (ldmatrix + mma.sync, see decode_mega_mma.cuh),
//     residual / SwiGLU epilogues, RoPE + KV append + split-KV GQA attention (one (batch, kv-head, split) item
//     per CTA, merge spread over the grid), final norm + LM head with a running arg-max.
//   * rows of each GEMV are cut into 148 contiguous, near-equal ranges (±1 group of 8 rows).
//
// DATA-FLOW SYNCHRONISATION (round 2).  The first version separated the ops with 5-6 grid-wide barriers per
// layer plus two all-reduce waits; the %globaltimer trace put ≈ 12 of 44 µs per layer into them at TP=4, which
// is why 8 GPUs bought only 2.07x one GPU.  Now there is NO grid barrier.  Every activation that crosses CTAs
// travels as an 8-byte TAGGED PACKET {payload, tag}: the producer's plain store is data and flag at once (an
// aligned 8-byte store is single-copy atomic), and a consumer polls exactly the packets it is about to use until
// their tag equals the tag of this (step, layer, op).  No fence, no counter, no second round trip; a CTA waits
// only for the data it needs.  Tags grow monotonically over the life of the engine (gen * 2L + 2l + {1,2}), so
// buffers are never cleared and a stale packet can never match.
//   - residual stream: each CTA OWNS the rows of x that its `wo` / `w2` row range produces and keeps them in
//     shared memory for the whole step; the epilogue adds, rounds to bf16 and publishes {bf16 pair, tag} packets
//     that every CTA's next staging pass polls (x never exists as a plain global array);
//   - qkv, attention partials, merged attention output and the SwiGLU output travel the same way;
//   - tensor parallelism: the row-parallel epilogues push their fp32 partials {value, tag} to every rank with ONE
//     `multimem.st` on the NVLS multicast address of the packet buffer (the NVSwitch replicates it; without
//     multicast: one unicast store per peer); the owner CTA of a row sums the tp packets in rank order
//     (bit-identical on all ranks), adds its residual slice and publishes the local x packets.  The wait for the
//     peers' packets is the only cross-GPU synchronisation of the all-reduce;
//   - the LM head keeps a running (max, argmax) per CTA and publishes one candidate per CTA and batch row to every
//     rank (same packet mechanism); `jl_mega_select` merges grid x tp candidates and does the sequence
//     bookkeeping: a greedy decode step is 2 launches at any tp, and the fp32 [M, V] logits are written only when
//     the caller asks for them.
// WAR safety of the single-buffered packet arrays follows from the dependency chain (a CTA can only produce the
// packets of op k+1 after it has consumed ALL packets of op k, and every later overwrite of op k's buffer is
// ordered behind such a chain); the cross-GPU buffers and the residual packets alternate by tag parity.  Every
// poll has a watchdog that names block / kind / tag and traps.
//
// History (all measured on B200, LLaMA-3 8B, batch 1, ms per token): SIMT consumers with one weight row per
// warp 3.69 → one 16 KB TMA request per stage 3.64 → activations in registers 3.18-3.39 (box dependent; ncu:
// ~1 instruction issued per cycle per SM, ~1100 warp-instructions per stage — instruction-latency bound, `w2`
// at 35 GB/s per SM whatever the ring depth) → tensor-core consumers 3.04 on the box that gave 3.39 (GEMV
// phases at 6.4-7.3 TB/s).  An L2 prefetch cursor running ahead of the ring was tried twice and was slower
// every time (+9 % at 16 stages ahead, +60 % at 64); 4-row work units to even out the 24/32-row split were
// neutral once the phases were HBM-bound.
//
// Weights of all layers are stacked per op ([L·N, K]) so five tensor maps describe the whole model.
// Math is identical to the unfused kernels (gemv.cu, decode_attn_fused.cu); the tests compare the two.
// Reference: the body of HF Flax `_sample`'s while-loop that /root/reference/jax_llama/generation.py:28-41
// runs as one compiled XLA program — this is the hand-written equivalent; the all-reduces are the ones GSPMD
// inserts for /root/reference/jax_llama/partition.py:67,70.
#include <stdlib.h>

#include "jllama_b200.h"
#include "nvls.cuh"
#include "tcgen05.cuh"

namespace jl {

constexpr int MG_CONSUMERS = 8;
// Warp roles and registers.  The SM allocates registers per scheduler quarter, so a 9-warp CTA (8 consumers + 1
// producer) caps EVERY thread at 168 registers (three warps share one quarter) and the consumers spilled inside the
// GEMV loop.  The CTA is therefore three warpgroups: warps 0-3 = producer group (warp 0 lane 0 issues the TMA
// stream, the others leave at once), warps 4-11 = consumers; `setmaxnreg` hands the producer group's registers to
// the consumers right after start-up (56 vs 224 per thread; 56·128 + 224·256 = 64512 of the SM's 65536).
#ifndef JL_MG_PWARPS
#define JL_MG_PWARPS 4
#endif
constexpr int MG_PWARPS = JL_MG_PWARPS;            // warps of the producer warpgroup (1: no register hand-over)
constexpr int MG_CT0 = MG_PWARPS * 32;             // first consumer thread
constexpr int MG_THREADS = (MG_CONSUMERS + MG_PWARPS) * 32;
constexpr int MG_PRODUCER_REGS = 56, MG_CONSUMER_REGS = 224;
constexpr int MG_CT = MG_CONSUMERS * 32;           // consumer threads
constexpr int MG_GROUP = 8;                        // weight rows per stage (one per consumer warp)
constexpr int MG_KS = 1024;                        // K elements per stage
constexpr int MG_STAGE_BYTES = MG_GROUP * MG_KS * 2;   // 16 KB
constexpr int MG_HD = 128;
constexpr int MG_KPI = 4;
constexpr int MG_PART = MG_HD + 2;                 // packets of one split-KV partial: 128 outputs, max, sum
constexpr int MG_MAX_OWN = 64;                     // residual rows one CTA may own (H / 8 groups over the grid)

struct alignas(64) MegaTmaps {
  CUtensorMap m[5];          // wqkv_all, wo_all, w13_all, w2_all, lm_head
};

// Cross-GPU packet buffers (symmetric heap of parallel/symm.py; tp == 1: cand_peer[0] is local scratch)
struct MegaAr {
  uint2* p_peer[8];                  // rank r's copy of the partial-sum packets: [parity][src][M][H] {fp32, tag}
  uint2* p_mc;                       // NVLS multicast address of the same buffer (null: unicast to every peer)
  uint2* cand_peer[8];               // rank r's copy of the arg-max candidates: [parity][src][grid][M][2] {bits, tag}
  uint2* cand_mc;
  int tp, rank;
};

struct MegaParams {
  MegaAr ar;
  int L, H, F, V, nh, nkv, NQKV, vocab_full;
  float eps, scale_log2;
  const __nv_bfloat16* attn_norm;    // [L, H]
  const __nv_bfloat16* ffn_norm;     // [L, H]
  const __nv_bfloat16* final_norm;   // [H]
  const __nv_bfloat16* embedding;    // [vocab_full, H]
  const int* ids;                    // [M]
  uint2* r_pk;                       // [2][M][H/2]     residual stream after wo / w2 ({bf16 pair, tag})
  uint2* qkv_pk;                     // [M][NQKV/2]
  uint2* part_pk;                    // [M·nh][splits][130]   split-KV partials ({fp32, tag})
  uint2* attn_pk;                    // [M][nh·hd/2]
  uint2* h_pk;                       // [M][F/2]
  float* logits;                     // [M, logits_ld] or null; this rank's V columns start at logits_col0
  int logits_ld, logits_col0, logits_mc;     // logits_mc: `logits` is a multicast address (every rank gets every shard)
  __nv_bfloat16* kc;                 // [L][M][nkv][T][hd]
  __nv_bfloat16* vc;
  const float* cos_t;
  const float* sin_t;
  const int* cache_len;
  const int* pad_left;
  unsigned int* gen;                 // steps run so far by this engine (tag base; bumped by block 0 at the end)
  int T, max_pos, splits, stages, l2_ahead;
  int w2_passes;                     // 2: w2 is walked in two K ranges (batch rows of its full input do not fit in smem)
  unsigned long long* trace;         // optional [grid][L][32] globaltimer stamps
};

enum { MG_PLAIN = 0, MG_RESIDUAL = 1, MG_SWIGLU = 2, MG_LOGITS = 3 };
enum { MG_W_X = 0, MG_W_QKV = 1, MG_W_PART = 2, MG_W_ATTN = 3, MG_W_H = 4, MG_W_PEER = 5, MG_W_X2 = 6 };      // what a poll waits for

__device__ __forceinline__ void mg_consumer_bar() { asm volatile("bar.sync 1, %0;" ::"n"(MG_CT) : "memory"); }

// ---- tagged packets
__device__ __forceinline__ void st_packet(uint2* dst, uint32_t payload, uint32_t tag) {
  asm volatile("st.relaxed.gpu.global.v2.u32 [%0], {%1, %2};" ::"l"(dst), "r"(payload), "r"(tag) : "memory");
}
__device__ __forceinline__ uint4 ld_relaxed_gpu_v4(const uint4* p) {
  uint4 v;
  asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint2 ld_relaxed_gpu_v2(const uint2* p) {
  uint2 v;
  asm volatile("ld.relaxed.gpu.global.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_relaxed_sys_v4(const uint4* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
// cold path of every poll: never returns, so the caller keeps nothing alive across the call
[[noreturn]] __device__ __noinline__ void mg_poll_timeout(int what, uint32_t want, uint32_t got) {
  printf("jllama_b200: decode kernel block %d thread %d timed out polling packets of kind %d (tag %u, last seen %u)\n",
         (int)blockIdx.x, (int)threadIdx.x, what, want, got);
  __trap();
  for (;;) {}
}
[[noreturn]] __device__ __noinline__ void mg_peer_timeout(int rank, int src, uint32_t tag, uint32_t got, int kind) {
  printf("jllama_b200: rank %d timed out waiting for rank %d (%s, tag %u, last seen %u, block %d)\n", rank, src,
         kind == 0 ? "partial sums of the all-reduce" : "arg-max candidates", tag, got, (int)blockIdx.x);
  __trap();
  for (;;) {}
}

__device__ __forceinline__ void mg_stamp(const MegaParams& p, int layer, int ev) {
  if (p.trace != nullptr && threadIdx.x == MG_CT0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    p.trace[((size_t)blockIdx.x * p.L + layer) * 32 + ev] = t;
    if (ev == 0) {
      unsigned int smid;
      asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
      p.trace[((size_t)blockIdx.x * p.L + layer) * 32 + 31] = smid;
    }
  }
}

// contiguous, near-equal split of `n` groups over the grid (n * gridDim.x < 2^31 for every matrix here)
__device__ __forceinline__ void mg_range(int n, int& g0, int& g1) {
  g0 = (int)(((unsigned)n * blockIdx.x) / gridDim.x);
  g1 = (int)(((unsigned)n * (blockIdx.x + 1u)) / gridDim.x);
}

struct MegaCtx {
  uint32_t ring, bars;
  const uint8_t* ring_ptr;
  uint4* xin;                 // [M][K/8] staged activation (bf16)
  float* s_attn;              // attention fold scratch
  int stages;
  int stage; uint32_t phase;  // ring cursor (producer and consumers keep separate copies)
  const MegaAr* ar;           // tensor-parallel state; ar_tag tags (and its parity places) the current all-reduce
  uint32_t ar_tag;
  unsigned long long* trace_row;   // tracing: this block's stamp row of the current layer (null: off)
  bool trace_on;
};
__device__ __forceinline__ void mg_stamp_raw(const MegaCtx& c, int ev) {
  if (threadIdx.x == MG_CT0 && c.trace_row != nullptr) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    c.trace_row[ev] = t;
  }
}

// row-parallel epilogue under tensor parallelism: hand this rank's fp32 partial to every rank (itself included).
// One 8-byte packet {value, tag}; the receiver polls the slot until the tag matches, so no flag, no fence and a
// single NVLink traversal are on the critical path.  With NVLS the switch replicates one store to all ranks.
template <int M>
__device__ __forceinline__ void mg_push_partial(const MegaCtx& c, int m, int row, int N, float v) {
  const MegaAr& a = *c.ar;
  const size_t off = (((size_t)(c.ar_tag & 1) * a.tp + a.rank) * M + m) * N + row;
  if (a.p_mc != nullptr) {
    mm::st_b64(a.p_mc + off, __float_as_uint(v), c.ar_tag);
  } else {
    for (int peer = 0; peer < a.tp; ++peer)
      asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(a.p_peer[peer] + off), "r"(__float_as_uint(v)),
                   "r"(c.ar_tag)
                   : "memory");
  }
}

// ------------------------------------------------------------------------------------ producer side
// Position in the model-wide weight stream of one CTA: (layer, op, row group, k stage).
struct MegaCursor {
  int l, op, g, g0, g1, ks, ks_end, n_ks, row_base;
  bool done, second_pass;
  __device__ __forceinline__ bool step_op() {
    if (op == 4) { done = true; return false; }
    if (++op == 4) { op = 0; ++l; }
    return true;
  }
  // TWO: the kernel instance supports the two-pass w2 walk (batch > 4); otherwise the plain sequence is compiled
  template <bool TWO>
  __device__ __forceinline__ void load_op(const MegaParams& p, int qdim) {
    for (;;) {
      if (l >= p.L && op != 4) op = 4;                   // all layers streamed → LM head
      int N, K;
      switch (op) {
        case 0: N = p.NQKV; K = p.H; row_base = l * p.NQKV; break;
        case 1: N = p.H; K = qdim; row_base = l * p.H; break;
        case 2: N = 2 * p.F; K = p.H; row_base = l * 2 * p.F; break;
        case 3: N = p.H; K = p.F; row_base = l * p.H; break;
        default: N = p.V; K = p.H; row_base = 0; break;
      }
      mg_range(N / MG_GROUP, g, g1);
      g0 = g;
      ks = 0;
      n_ks = (K + MG_KS - 1) / MG_KS;
      // two-pass w2: every row group over the first K range, then every row group over the second
      second_pass = false;
      ks_end = (TWO && op == 3 && p.w2_passes == 2) ? (n_ks + 1) / 2 : n_ks;
      if (g < g1) return;
      if (!step_op()) return;                            // empty range for this CTA: skip the op
    }
  }
  template <bool TWO>
  __device__ __forceinline__ void init(const MegaParams& p, int qdim) { l = 0; op = 0; done = false; load_op<TWO>(p, qdim); }
  template <bool TWO>
  __device__ __forceinline__ void next(const MegaParams& p, int qdim) {
    if constexpr (TWO) {
      if (++ks < ks_end) return;
      ks = second_pass ? (n_ks + 1) / 2 : 0;
      if (++g < g1) return;
      if (ks_end < n_ks) {                               // first pass of a two-pass op done: same rows, second K range
        second_pass = true;
        g = g0;
        ks = ks_end;
        ks_end = n_ks;
        return;
      }
    } else {
      if (++ks < n_ks) return;
      ks = 0;
      if (++g < g1) return;
    }
    if (step_op()) load_op<TWO>(p, qdim);
  }
  __device__ __forceinline__ int map() const { return op; }
  __device__ __forceinline__ int row() const { return row_base + g * MG_GROUP; }
};

}  // namespace jl
#include "decode_mega_mma.cuh"
namespace jl {

// 8 consecutive bf16 of N packet rows (4 packets = two 16-byte loads each): all loads are issued, then all tags are
// checked, and everything is fetched again until every tag matches (one L2 round trip once the data is there)
template <int N>
__device__ __forceinline__ void mg_poll8(const uint2* const (&pk)[N], uint32_t tag, int what, uint4 (&out)[N]) {
  unsigned spins = 0;
  for (;;) {
    uint4 a[N], b[N];
    bool ok = true;
#pragma unroll
    for (int i = 0; i < N; ++i) {
      a[i] = ld_relaxed_gpu_v4(reinterpret_cast<const uint4*>(pk[i]));
      b[i] = ld_relaxed_gpu_v4(reinterpret_cast<const uint4*>(pk[i]) + 1);
    }
#pragma unroll
    for (int i = 0; i < N; ++i) {
      ok = ok && a[i].y == tag && a[i].w == tag && b[i].y == tag && b[i].w == tag;
      out[i] = make_uint4(a[i].x, a[i].z, b[i].x, b[i].z);
    }
    if (ok) return;
    if (++spins > (1u << 22)) mg_poll_timeout(what, tag, a[0].y);
  }
}

// RoPE + KV append + split-KV attention for one (b, kv head, split) item (8 consumer warps)
struct MgPos { int len, first; float cs[4], sn[4]; };      // per-step constants of one batch row (this lane's pairs)

__device__ __forceinline__ MgPos mg_pos(const MegaParams& p, int b, int lig) {
  MgPos r;
  int len = __ldcg(p.cache_len);
  r.len = len > p.T ? p.T : len;
  r.first = p.pad_left ? p.pad_left[b] : 0;
  int pos = r.len - r.first;
  pos = pos < 0 ? 0 : (pos >= p.max_pos ? p.max_pos - 1 : pos);
  const float4 cs4 = __ldg(reinterpret_cast<const float4*>(p.cos_t) + (size_t)pos * (MG_HD / 8) + lig);
  const float4 sn4 = __ldg(reinterpret_cast<const float4*>(p.sin_t) + (size_t)pos * (MG_HD / 8) + lig);
  r.cs[0] = cs4.x; r.cs[1] = cs4.y; r.cs[2] = cs4.z; r.cs[3] = cs4.w;
  r.sn[0] = sn4.x; r.sn[1] = sn4.y; r.sn[2] = sn4.z; r.sn[3] = sn4.w;
  return r;
}

template <int G>
__device__ __forceinline__ void mg_attention_item(const MegaParams& p, const MegaCtx& c, int layer, int M, int b,
                                                  int kvh, int split, uint32_t tag, const uint2* qkv_pk, uint2* part_pk,
                                                  uint2* attn_pk, const MgPos& pp) {
  constexpr int HD = MG_HD, LPK = HD / 8, GPW = 32 / LPK, NG = MG_CONSUMERS * GPW;
  constexpr int KEYS_PER_ITER = NG * MG_KPI;
  const int nh = p.nkv * G;
  const int lane = threadIdx.x & 31, warp = (threadIdx.x >> 5) - MG_PWARPS, ct = threadIdx.x - MG_CT0;
  const int lig = lane % LPK, gid = warp * GPW + lane / LPK, group_base = lane - lig;
  const int kidx = ((lig / (LPK / 2)) & 1) * 2 + ((lig / (LPK / 4)) & 1);
  float (*s_o)[G][HD] = reinterpret_cast<float (*)[G][HD]>(c.s_attn);
  float (*s_ml)[G][2] = reinterpret_cast<float (*)[G][2]>(c.s_attn + NG * G * HD);

  const int len = pp.len, first = pp.first;
  const int valid = len - first > 0 ? len - first : 0;
  int chunk = (valid + p.splits - 1) / p.splits;
  chunk = ((chunk + KEYS_PER_ITER - 1) / KEYS_PER_ITER) * KEYS_PER_ITER;
  const int lo = first + split * chunk;
  const int hi = lo + chunk < len ? lo + chunk : len;

  mg_stamp(p, layer, 16);
  const float cs[4] = {pp.cs[0], pp.cs[1], pp.cs[2], pp.cs[3]}, sn[4] = {pp.sn[0], pp.sn[1], pp.sn[2], pp.sn[3]};
  // q (all heads of the group) and the new token's k, v arrive as packets from the qkv GEMV of other CTAs
  const uint2* row = qkv_pk + (size_t)b * (p.NQKV / 2);
  const bool owns_new = split == 0 && warp == 0;      // consumes the new k, v after the KV loop
  // The K/V cache rows of this item's first load batch do not depend on q: request them BEFORE polling for q, so the
  // HBM round trip of the cache overlaps the wait for the qkv GEMV of the other CTAs (needs the 224-register
  // consumers: at 168 registers the held rows spilled and this lost, round 1).
  constexpr int UNR = G <= 4 ? 2 : 1;
  const size_t head_off = (((size_t)layer * M + b) * p.nkv + kvh) * (size_t)p.T * HD;
  const __nv_bfloat16* kbase = p.kc + head_off + lig * 8;
  const __nv_bfloat16* vbase = p.vc + head_off + lig * 8;
  uint4 kr[UNR][MG_KPI], vr[UNR][MG_KPI];
  auto load_batch = [&](int base) {
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const int key0 = base + u * KEYS_PER_ITER + gid * MG_KPI;
#pragma unroll
      for (int i = 0; i < MG_KPI; ++i) {
        const bool ok = key0 + i < hi;
        kr[u][i] = ok ? ldg_stream(kbase + (size_t)(key0 + i) * HD) : make_uint4(0, 0, 0, 0);
        vr[u][i] = ok ? ldg_stream(vbase + (size_t)(key0 + i) * HD) : make_uint4(0, 0, 0, 0);
      }
    }
  };
  load_batch(lo);
  uint4 qa[G];
  {
    const uint2* src[G];
#pragma unroll
    for (int g = 0; g < G; ++g) src[g] = row + ((size_t)(kvh * G + g) * HD + lig * 8) / 2;
    mg_poll8<G>(src, tag, MG_W_QKV, qa);
  }
  uint4 ka = make_uint4(0, 0, 0, 0), va = ka;
  if (owns_new) {
    const uint2* src[2] = {row + ((size_t)(nh + kvh) * HD + lig * 8) / 2, row + ((size_t)(nh + p.nkv + kvh) * HD + lig * 8) / 2};
    uint4 kv[2];
    mg_poll8<2>(src, tag, MG_W_QKV, kv);
    ka = kv[0]; va = kv[1];
  }
  float qf[G][8];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const uint32_t aw[4] = {qa[g].x, qa[g].y, qa[g].z, qa[g].w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float re = bf16_lo(aw[j]), im = bf16_hi(aw[j]);
      const uint32_t r = pack_bf16(re * cs[j] - im * sn[j], re * sn[j] + im * cs[j]);
      qf[g][2 * j] = bf16_lo(r) * p.scale_log2;
      qf[g][2 * j + 1] = bf16_hi(r) * p.scale_log2;
    }
  }
  float m[G], l[G], o[G][8];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    m[g] = -INFINITY; l[g] = 0.f;
#pragma unroll
    for (int e = 0; e < 8; ++e) o[g][e] = 0.f;
  }
  auto process = [&](const uint4 (&kr)[MG_KPI], const uint4 (&vr)[MG_KPI], const int key0) {
    float part[G][MG_KPI];
#pragma unroll
    for (int i = 0; i < MG_KPI; ++i) {
      const float kf[8] = {bf16_lo(kr[i].x), bf16_hi(kr[i].x), bf16_lo(kr[i].y), bf16_hi(kr[i].y),
                           bf16_lo(kr[i].z), bf16_hi(kr[i].z), bf16_lo(kr[i].w), bf16_hi(kr[i].w)};
#pragma unroll
      for (int g = 0; g < G; ++g) {
        float s = 0.f;
#pragma unroll
        for (int e = 0; e < 8; ++e) s = fmaf(qf[g][e], kf[e], s);
        part[g][i] = s;
      }
    }
    float pr[G][MG_KPI];
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const bool up = (lig & (LPK / 2)) != 0;
      float k0 = up ? part[g][2] : part[g][0], k1 = up ? part[g][3] : part[g][1];
      const float s0 = up ? part[g][0] : part[g][2], s1 = up ? part[g][1] : part[g][3];
      k0 += __shfl_xor_sync(0xffffffffu, s0, LPK / 2);
      k1 += __shfl_xor_sync(0xffffffffu, s1, LPK / 2);
      const bool up2 = (lig & (LPK / 4)) != 0;
      float s = up2 ? k1 : k0;
      const float snd = up2 ? k0 : k1;
      s += __shfl_xor_sync(0xffffffffu, snd, LPK / 4);
#pragma unroll
      for (int off = LPK / 8; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
      if (key0 + kidx >= hi) s = -INFINITY;
      float mx = fmaxf(s, __shfl_xor_sync(0xffffffffu, s, LPK / 2));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, LPK / 4));
      const float m_new = fmaxf(m[g], mx);
      const float m_safe = m_new == -INFINITY ? 0.f : m_new;
      const float corr = exp2f(m[g] - m_safe);
      const float pe = exp2f(s - m_safe);
      m[g] = m_new;
      float psum = 0.f;
#pragma unroll
      for (int i = 0; i < MG_KPI; ++i) {
        pr[g][i] = __shfl_sync(0xffffffffu, pe, group_base + (i >> 1) * (LPK / 2) + (i & 1) * (LPK / 4));
        psum += pr[g][i];
      }
      l[g] = l[g] * corr + psum;
#pragma unroll
      for (int e = 0; e < 8; ++e) o[g][e] *= corr;
    }
#pragma unroll
    for (int i = 0; i < MG_KPI; ++i) {
      const float vf[8] = {bf16_lo(vr[i].x), bf16_hi(vr[i].x), bf16_lo(vr[i].y), bf16_hi(vr[i].y),
                           bf16_lo(vr[i].z), bf16_hi(vr[i].z), bf16_lo(vr[i].w), bf16_hi(vr[i].w)};
#pragma unroll
      for (int g = 0; g < G; ++g)
#pragma unroll
        for (int e = 0; e < 8; ++e) o[g][e] = fmaf(pr[g][i], vf[e], o[g][e]);
    }
  };

  mg_stamp(p, layer, 17);
#pragma unroll 1
  for (int base = lo; base < hi; base += UNR * KEYS_PER_ITER) {
    if (base != lo) load_batch(base);
#pragma unroll
    for (int u = 0; u < UNR; ++u) process(kr[u], vr[u], base + u * KEYS_PER_ITER + gid * MG_KPI);
  }

  mg_stamp(p, layer, 18);
  // the new token: rotate k, append (k, v), fold it in from registers (split 0, consumer warp 0)
  if (owns_new) {
    const uint32_t kw[4] = {ka.x, ka.y, ka.z, ka.w};
    uint32_t kr4[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float re = bf16_lo(kw[j]), im = bf16_hi(kw[j]);
      kr4[j] = pack_bf16(re * cs[j] - im * sn[j], re * sn[j] + im * cs[j]);
    }
    if (gid == 0 && len < p.T) {
      *reinterpret_cast<uint4*>(p.kc + head_off + (size_t)len * HD + lig * 8) = make_uint4(kr4[0], kr4[1], kr4[2], kr4[3]);
      *reinterpret_cast<uint4*>(p.vc + head_off + (size_t)len * HD + lig * 8) = va;
    }
    const float kf[8] = {bf16_lo(kr4[0]), bf16_hi(kr4[0]), bf16_lo(kr4[1]), bf16_hi(kr4[1]),
                         bf16_lo(kr4[2]), bf16_hi(kr4[2]), bf16_lo(kr4[3]), bf16_hi(kr4[3])};
    const float vf[8] = {bf16_lo(va.x), bf16_hi(va.x), bf16_lo(va.y), bf16_hi(va.y),
                         bf16_lo(va.z), bf16_hi(va.z), bf16_lo(va.w), bf16_hi(va.w)};
#pragma unroll
    for (int g = 0; g < G; ++g) {
      float s = 0.f;
#pragma unroll
      for (int e = 0; e < 8; ++e) s = fmaf(qf[g][e], kf[e], s);
#pragma unroll
      for (int off = LPK / 2; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
      if (gid != 0) s = -INFINITY;
      const float m_new = fmaxf(m[g], s);
      const float m_safe = m_new == -INFINITY ? 0.f : m_new;
      const float corr = exp2f(m[g] - m_safe);
      const float pe = exp2f(s - m_safe);
      m[g] = m_new;
      l[g] = l[g] * corr + pe;
#pragma unroll
      for (int e = 0; e < 8; ++e) o[g][e] = o[g][e] * corr + pe * vf[e];
    }
  }

  mg_stamp(p, layer, 19);
  // fold the lane-groups of this CTA
#pragma unroll
  for (int g = 0; g < G; ++g) {
#pragma unroll
    for (int e = 0; e < 8; ++e) s_o[gid][g][lig * 8 + e] = o[g][e];
    if (lig == 0) { s_ml[gid][g][0] = m[g]; s_ml[gid][g][1] = l[g]; }
  }
  mg_consumer_bar();
  // G·HD is a multiple of 128: the trip count is warp-uniform, so the pair shuffle below has all lanes present
  for (int idx = ct; idx < G * HD; idx += MG_CT) {
    const int g = idx / HD, d = idx - g * HD;
    float Mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < NG; ++j) Mx = fmaxf(Mx, s_ml[j][g][0]);
    const float Ms = Mx == -INFINITY ? 0.f : Mx;
    float acc = 0.f, Ls = 0.f;
#pragma unroll
    for (int j = 0; j < NG; ++j) {
      const float w = exp2f(s_ml[j][g][0] - Ms);
      acc += w * s_o[j][g][d];
      Ls += w * s_ml[j][g][1];
    }
    const int h = kvh * G + g;
    if (p.splits == 1) {
      const float y = Ls > 0.f ? acc / Ls : 0.f;
      const float y1 = __shfl_down_sync(0xffffffffu, y, 1);
      if ((d & 1) == 0) st_packet(attn_pk + ((size_t)b * nh + h) * (HD / 2) + (d >> 1), pack_bf16(y, y1), tag);
    } else {
      uint2* dst = part_pk + (((size_t)b * nh + h) * p.splits + split) * MG_PART;
      st_packet(dst + d, __float_as_uint(acc), tag);
      if (d == 0) { st_packet(dst + HD, __float_as_uint(Mx), tag); st_packet(dst + HD + 1, __float_as_uint(Ls), tag); }
    }
  }
  mg_stamp(p, layer, 20);
  mg_consumer_bar();                 // s_attn is reused by this CTA's next item
}

// Merge of the split-KV partials, spread over the grid one warp at a time (32 outputs of one head per warp, every
// CTA's first warp before any CTA's second).  Splits are taken in chunks of 16: lane s fetches the (max, sum) pair
// of split s for the whole warp, every lane fetches its own output element of the 16 splits, all of it is issued
// before anything is checked (one L2 round trip per chunk once the data is there); the softmax weights travel by
// shuffle.  Indices past the last split are clamped to it and get weight zero, so the code has no tail branches.
// The merged attention output is published as {bf16 pair, tag} packets for the `wo` staging.
__device__ __noinline__ void mg_attn_merge(const uint2* part_pk, uint2* attn_pk, int total, int splits, uint32_t tag) {
  const int lane = threadIdx.x & 31, cw = (threadIdx.x >> 5) - MG_PWARPS;
#pragma unroll 1
  for (int idx = (cw * gridDim.x + blockIdx.x) * 32 + lane; idx < total; idx += gridDim.x * MG_CT) {
    const int bh = idx / MG_HD, d = idx - bh * MG_HD;          // bh is warp-uniform (32 | 128)
    const uint2* base = part_pk + (size_t)bh * splits * MG_PART;
    float Mx = -INFINITY, acc = 0.f, Ls = 0.f;
#pragma unroll 1
    for (int s0 = 0; s0 < splits; s0 += 16) {
      const int ns = splits - s0 < 16 ? splits - s0 : 16;
      const uint4* mlp = reinterpret_cast<const uint4*>(base + (size_t)(s0 + (lane < ns ? lane : 0)) * MG_PART + MG_HD);
      uint2 ov[16];
      uint4 ml;
      unsigned spins = 0;
      for (;;) {
        bool ok = true;
#pragma unroll
        for (int u = 0; u < 16; ++u) ov[u] = ld_relaxed_gpu_v2(base + (size_t)(s0 + (u < ns ? u : ns - 1)) * MG_PART + d);
        ml = ld_relaxed_gpu_v4(mlp);
#pragma unroll
        for (int u = 0; u < 16; ++u) ok = ok && ov[u].y == tag;
        if (ok && ml.y == tag && ml.w == tag) break;
        if (++spins > (1u << 22)) mg_poll_timeout(MG_W_PART, tag, ml.y);
      }
      const float my_m = lane < ns ? __uint_as_float(ml.x) : -INFINITY;
      const float cm = fmaxf(Mx, warp_max(my_m));
      const float Ms = cm == -INFINITY ? 0.f : cm;
      const float corr = exp2f(Mx - Ms);                       // first chunk: Mx = -inf → 0
      const float w = exp2f(my_m - Ms);                        // lanes >= ns: 0
      acc *= corr;
      Ls = Ls * corr + warp_sum(w * (lane < ns ? __uint_as_float(ml.z) : 0.f));
#pragma unroll
      for (int u = 0; u < 16; ++u) {
        const float wu = __shfl_sync(0xffffffffu, w, u);
        acc += wu > 0.f ? wu * __uint_as_float(ov[u].x) : 0.f;
      }
      Mx = cm;
    }
    const float y = Ls > 0.f ? acc / Ls : 0.f;
    const float y1 = __shfl_down_sync(0xffffffffu, y, 1);
    if ((d & 1) == 0) st_packet(attn_pk + (size_t)bh * (MG_HD / 2) + (d >> 1), pack_bf16(y, y1), tag);
  }
}

// Second half of the in-kernel one-shot all-reduce (tp > 1): the owner CTA of a row range waits for the packets of
// the same-index CTA of every rank (all ranks cut rows identically), sums them in rank order (bit-identical on
// all ranks), adds its residual slice and publishes the new residual-stream packets to the CTAs of this GPU.
template <int M>
__device__ __noinline__ void mg_owner_reduce(const uint2* mine, int tp, int rank, uint32_t tag, int N,
                                             float (*res)[MG_MAX_OWN], uint2* r_pk) {
  const int ct = threadIdx.x - MG_CT0;
  int g0, g1;
  mg_range(N / MG_GROUP, g0, g1);
  const int row0 = g0 * MG_GROUP, pairs = (g1 - g0) * MG_GROUP / 2;
#pragma unroll 1
  for (int i = ct; i < pairs * M; i += MG_CT) {
    const int m = i / pairs, row = row0 + 2 * (i - m * pairs);
    uint4 v[8];
    unsigned spins = 0;
    for (;;) {
      bool ok = true;
#pragma unroll
      for (int src = 0; src < 8; ++src)       // sources past the group are clamped to the last rank (re-read, ignored below)
        v[src] = ld_relaxed_sys_v4(reinterpret_cast<const uint4*>(mine + ((size_t)(src < tp ? src : tp - 1) * M + m) * N + row));
#pragma unroll
      for (int src = 0; src < 8; ++src) ok = ok && v[src].y == tag && v[src].w == tag;
      if (ok) break;
      if (++spins > (1u << 22)) {
        int late = 0;
        for (int src = 0; src < 8; ++src) if (src < tp && (v[src].y != tag || v[src].w != tag)) late = src;
        mg_peer_timeout(rank, late, tag, v[late].y, 0);
      }
    }
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int src = 0; src < 8; ++src) {
      if (src < tp) { s0 += __uint_as_float(v[src].x); s1 += __uint_as_float(v[src].z); }
    }
    const float y0 = __bfloat162float(__float2bfloat16(__bfloat162float(__float2bfloat16(s0)) + res[m][row - row0]));
    const float y1 = __bfloat162float(__float2bfloat16(__bfloat162float(__float2bfloat16(s1)) + res[m][row - row0 + 1]));
    res[m][row - row0] = y0;
    res[m][row - row0 + 1] = y1;
    st_packet(r_pk + (size_t)m * (N / 2) + (row >> 1), pack_bf16(y0, y1), tag);
  }
}

template <int M, int G>
__global__ void __launch_bounds__(MG_THREADS, 1)
decode_mega_kernel(const __grid_constant__ MegaTmaps tm, const MegaParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ float red[M][MG_CONSUMERS];
  __shared__ __align__(16) float s_acc[2][MG_CONSUMERS][MG_MMA_MAX_M][8];
  __shared__ float s_carry[M > 4 ? M : 1][64];         // two-pass w2: sums of the first K range (this CTA's rows)
  __shared__ float s_res[M][MG_MAX_OWN];               // this CTA's rows of the residual stream (bf16-exact values)
  __shared__ volatile int s_idle_flag;          // consumers → producer: "no GEMV is consuming the ring right now"
  volatile int* s_idle = &s_idle_flag;
  if (threadIdx.x == 0) s_idle_flag = 0;

  MegaCtx c;
  c.stages = p.stages;
  c.ring = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
  c.ring_ptr = smem_raw + (c.ring - tc::smem_u32(smem_raw));
  c.bars = c.ring + p.stages * MG_STAGE_BYTES;
  uint8_t* after = const_cast<uint8_t*>(c.ring_ptr) + p.stages * MG_STAGE_BYTES + 16 * p.stages;
  c.s_attn = reinterpret_cast<float*>(after);
  // the staged GEMV input and the attention scratch are never live together (a GEMV ends before attention
  // starts and is restaged after it), so they share one region sized for the larger of the two
  c.xin = reinterpret_cast<uint4*>(after);
  c.stage = 0; c.phase = 0;
  c.ar = &p.ar; c.ar_tag = 0;
  c.trace_row = nullptr; c.trace_on = p.trace != nullptr;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 5; ++i) tc::prefetch_tensormap(&tm.m[i]);
    for (int s = 0; s < p.stages; ++s) {
      tc::mbar_init(c.bars + 8u * s, 1);
      tc::mbar_init(c.bars + 8u * (p.stages + s), MG_CONSUMERS);
    }
    tc::fence_barrier_init();
  }
  __syncthreads();

  const int qdim = p.nh * MG_HD;
  if (warp < MG_PWARPS) {
    if constexpr (MG_PWARPS == 4) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(MG_PRODUCER_REGS));
    // ===================================================== producer: the whole model's weights, in order
    if (warp == 0 && lane == 0) {
      // `ld` feeds the ring.  While the consumers are NOT consuming (attention, merge and the polls around
      // them: s_idle is set) and the ring is full, `pf` walks up to p.l2_ahead stages further down the same
      // sequence and only asks the L2 to fetch them, so HBM keeps streaming through the part of a layer in
      // which no GEMV runs.  (Prefetching all the time was measured twice and loses: it queues in front of
      // the loads the GEMV is waiting for.)
      MegaCursor ld, pf;
      ld.template init<(M > 4)>(p, qdim);
      pf = ld;
      int ahead = 0;                       // stages between ld and pf
      while (!ld.done) {
        const uint32_t empty = c.bars + 8u * (c.stages + c.stage);
        bool ready;
        if (p.l2_ahead > 0 && *s_idle != 0 && !pf.done && ahead < p.l2_ahead) {
          ready = tc::mbar_test_wait(empty, c.phase ^ 1);
          if (!ready) {
            if (ahead >= c.stages) tc::tma_prefetch_3d(&tm.m[pf.map()], 0, pf.row(), pf.ks * (MG_KS / 64));
            pf.template next<(M > 4)>(p, qdim);
            ++ahead;
          }
        } else {
          // try_wait suspends the thread in hardware until the slot frees or a time-out elapses; a busy
          // test_wait loop here steals issue slots from the two consumer warps that share this scheduler
          // (measured: 122 vs 108 µs per layer)
          ready = tc::mbar_try_wait(empty, c.phase ^ 1);
        }
        if (ready) {
          const uint32_t full = c.bars + 8u * c.stage;
          tc::mbar_expect_tx(full, MG_STAGE_BYTES);
          tc::tma_load_3d(c.ring + c.stage * MG_STAGE_BYTES, &tm.m[ld.map()], full, 0, ld.row(), ld.ks * (MG_KS / 64),
                          tc::EVICT_FIRST);
          if (++c.stage == c.stages) { c.stage = 0; c.phase ^= 1; }
          ld.template next<(M > 4)>(p, qdim);
          if (ahead > 0) --ahead; else pf = ld;
        }
      }
    }
    return;
  }

  // ===================================================== consumers
  if constexpr (MG_PWARPS == 4) asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(MG_CONSUMER_REGS));
  const unsigned int gen = *reinterpret_cast<volatile unsigned int*>(p.gen);
  const uint32_t base = gen * (uint32_t)(2 * p.L);           // tags of this step: base + 2 l + {1, 2}; base is even
  const MgPos pos0 = mg_pos(p, 0, lane % (MG_HD / 8));      // batch row 0: constant for the whole step
  const int ct = threadIdx.x - MG_CT0;
  int own0, own1;
  mg_range(p.H / MG_GROUP, own0, own1);
  own0 *= MG_GROUP; own1 *= MG_GROUP;                        // rows of the residual stream this CTA owns

  // residual slice ← embedding rows of this step's tokens
  for (int i = ct; i < M * (own1 - own0); i += MG_CT) {
    const int m = i / (own1 - own0), j = i - m * (own1 - own0);
    int id = p.ids[m];
    id = id < 0 ? 0 : (id >= p.vocab_full ? p.vocab_full - 1 : id);
    s_res[m][j] = __bfloat162float(p.embedding[(size_t)id * p.H + own0 + j]);
  }
  // (first read of s_res happens after the consumer barriers of the first staging pass)

  const uint2* ar_mine = p.ar.tp > 1 ? p.ar.p_peer[p.ar.rank] : nullptr;      // this rank's copy of the partial-sum packets
  const size_t r_half = (size_t)M * (p.H / 2);               // packets of one parity half of the residual buffer
  MegaOut o;
  o.res = s_res; o.row0 = own0; o.logits = nullptr; o.best_v = nullptr; o.best_i = nullptr; o.idx_offset = 0;
  o.logits_ld = 0; o.logits_col0 = 0; o.logits_mc = false;
  o.pk = nullptr; o.r_pk = nullptr; o.tag = 0; o.red = nullptr; o.inv_k = 0.f; o.eps = p.eps;
  for (int l = 0; l < p.L; ++l) {
    const uint32_t t1 = base + 2u * l + 1u, t2 = t1 + 1u;      // t1 odd, t2 even
    uint2* r1 = p.r_pk + r_half;                               // residual packets tagged t1 (after wo)
    uint2* r2 = p.r_pk;                                        // residual packets tagged t2 (after w2)
    // qkv / partials / attention output / SwiGLU output alternate between two buffers by layer parity: a CTA whose
    // row range of some GEMV is empty (tiny models) may trail the grid by most of a layer, and must never find
    // the packets it is still polling overwritten by the next layer's
    const int lp = l & 1;
    uint2* qkv_pk = p.qkv_pk + (size_t)lp * M * (p.NQKV / 2);
    uint2* part_pk = p.part_pk + (size_t)lp * M * p.nh * p.splits * MG_PART;
    uint2* attn_pk = p.attn_pk + (size_t)lp * M * (qdim / 2);
    uint2* h_pk = p.h_pk + (size_t)lp * M * (p.F / 2);
    // ---- fused qkv projection (RMSNorm in the prologue); x comes from the previous layer's w2 epilogues
    if (c.trace_on) c.trace_row = p.trace + ((size_t)blockIdx.x * p.L + l) * 32;
    mg_stamp(p, l, 0);
    if (l == 0) mg_stage_embedding<M>(c.xin, p.embedding, p.ids, p.vocab_full, p.H, p.attn_norm, red);
    else mg_stage_packets<M>(c.xin, r2, t1 - 1u, p.H, p.H / 2, p.attn_norm + (size_t)l * p.H, red, MG_W_X, c.trace_row);
    mg_stamp(p, l, 1);
    o.pk = qkv_pk; o.tag = t1; o.red = red; o.inv_k = 1.f / (float)p.H;
    mg_gemv_mma<M, MG_PLAIN>(c, p.NQKV, p.H, o, s_acc);
    mg_stamp(p, l, 2);
    if (threadIdx.x == MG_CT0) *s_idle = 1;
    // ---- attention: each item polls the q / k / v packets it needs
    {
      const int items = M * p.nkv * p.splits;
      for (int it = blockIdx.x; it < items; it += gridDim.x) {
        const int split = it % p.splits, bk = it / p.splits;
        const int bb = bk / p.nkv;
        mg_attention_item<G>(p, c, l, M, bb, bk % p.nkv, split, t1, qkv_pk, part_pk, attn_pk, M == 1 ? pos0 : mg_pos(p, bb, lane % (MG_HD / 8)));
      }
    }
    mg_stamp(p, l, 3);
    if (p.splits > 1) mg_attn_merge(part_pk, attn_pk, M * p.nh * MG_HD, p.splits, t1);
    mg_stamp(p, l, 4);
    // ---- wo + residual
    mg_stage_packets<M>(c.xin, attn_pk, t1, qdim, qdim / 2, nullptr, red, MG_W_ATTN, c.trace_row);
    mg_stamp(p, l, 5);
    if (threadIdx.x == MG_CT0) *s_idle = 0;
    c.ar_tag = t1;
    o.r_pk = r1; o.tag = t1; o.red = nullptr;
    mg_gemv_mma<M, MG_RESIDUAL>(c, p.H, qdim, o, s_acc);
    if (p.ar.tp > 1) mg_owner_reduce<M>(ar_mine + (size_t)(t1 & 1) * p.ar.tp * M * p.H, p.ar.tp, p.ar.rank, t1, p.H, s_res, r1);
    mg_stamp(p, l, 6);
    // ---- w1/w3 + SwiGLU (RMSNorm in the prologue)
    mg_stage_packets<M>(c.xin, r1, t1, p.H, p.H / 2, p.ffn_norm + (size_t)l * p.H, red, MG_W_X2, c.trace_row);
    mg_stamp(p, l, 7);
    o.pk = h_pk; o.tag = t2; o.red = red; o.inv_k = 1.f / (float)p.H;
    mg_gemv_mma<M, MG_SWIGLU>(c, 2 * p.F, p.H, o, s_acc);
    mg_stamp(p, l, 8);
    // ---- w2 + residual
    c.ar_tag = t2;
    o.r_pk = r2; o.tag = t2; o.red = nullptr;
    bool w2_done = false;
    if constexpr (M > 4) {
      if (p.w2_passes == 2) {
        const int ka = ((p.F + MG_KS - 1) / MG_KS + 1) / 2 * MG_KS;       // first K range: whole k-stages
        mg_stage_packets<M>(c.xin, h_pk, t2, ka, p.F / 2, nullptr, red, MG_W_H, c.trace_row);
        mg_stamp(p, l, 9);
        mg_gemv_mma<M, MG_RESIDUAL, 1>(c, p.H, ka, o, s_acc, s_carry);
        mg_stage_packets<M>(c.xin, h_pk + ka / 2, t2, p.F - ka, p.F / 2, nullptr, red, MG_W_H, c.trace_row);
        mg_gemv_mma<M, MG_RESIDUAL, 2>(c, p.H, p.F - ka, o, s_acc, s_carry);
        w2_done = true;
      }
    }
    if (!w2_done) {
      mg_stage_packets<M>(c.xin, h_pk, t2, p.F, p.F / 2, nullptr, red, MG_W_H, c.trace_row);
      mg_stamp(p, l, 9);
      mg_gemv_mma<M, MG_RESIDUAL>(c, p.H, p.F, o, s_acc);
    }
    if (p.ar.tp > 1) mg_owner_reduce<M>(ar_mine + (size_t)(t2 & 1) * p.ar.tp * M * p.H, p.ar.tp, p.ar.rank, t2, p.H, s_res, r2);
    mg_stamp(p, l, 10);
  }
  // ---- final norm + LM head with a running arg-max per (batch row, weight-row lane)
  const uint32_t t_last = base + 2u * p.L;
  c.trace_row = nullptr;
  mg_stage_packets<M>(c.xin, p.r_pk, t_last, p.H, p.H / 2, p.final_norm, red, MG_W_X, c.trace_row);
  float best_v = -INFINITY;
  int best_i = 0x7fffffff;
  o.logits = p.logits; o.logits_ld = p.logits_ld; o.logits_col0 = p.logits_col0; o.logits_mc = p.logits_mc != 0;
  o.best_v = &best_v; o.best_i = &best_i; o.idx_offset = p.ar.rank * p.V;
  o.red = red; o.inv_k = 1.f / (float)p.H;
  mg_gemv_mma<M, MG_LOGITS>(c, p.V, p.H, o, s_acc);
  if (ct < 64) {                                             // thread (m, n): fold the 8 weight-row lanes of a batch row
    // logits stored into the other ranks' buffers must be there before the candidate packet that announces them
    if (p.logits != nullptr && p.logits_mc) __threadfence_system();
    __syncwarp();
#pragma unroll
    for (int off = 1; off < 8; off <<= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best_v, off);
      const int oi = __shfl_xor_sync(0xffffffffu, best_i, off);
      if (ov > best_v || (ov == best_v && oi < best_i)) { best_v = ov; best_i = oi; }
    }
    const int m = ct >> 3;
    if ((ct & 7) == 0 && m < M) {
      // one candidate per (CTA, batch row), published to every rank: {value, tag}, {index, tag}
      const MegaAr& a = p.ar;
      const size_t off = ((((size_t)(gen & 1) * a.tp + a.rank) * gridDim.x + blockIdx.x) * M + m) * 2;
      if (a.cand_mc != nullptr) {
        mm::st_b64(a.cand_mc + off, __float_as_uint(best_v), t_last);
        mm::st_b64(a.cand_mc + off + 1, (uint32_t)best_i, t_last);
      } else {
        for (int peer = 0; peer < a.tp; ++peer) {
          asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(a.cand_peer[peer] + off),
                       "r"(__float_as_uint(best_v)), "r"(t_last) : "memory");
          asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(a.cand_peer[peer] + off + 1),
                       "r"((uint32_t)best_i), "r"(t_last) : "memory");
        }
      }
    }
  }
  // every block read `gen` before it produced anything, and block 0 got here only after consuming packets of the
  // last layer from every block: nobody can still be about to read the old value
  if (blockIdx.x == 0 && threadIdx.x == MG_CT0) *p.gen = gen + 1u;
}

// Greedy selection + sequence bookkeeping after the decode kernel: merges the grid x tp arg-max candidates of every
// batch row (polling the packets of ranks whose kernel is still finishing), ties → lower vocabulary index, then
// does what advance_kernel does.  One CTA.
struct SelectParams {
  const uint2* cand;           // this rank's copy: [parity][tp][grid][M][2]
  const unsigned int* gen;     // already bumped by the decode kernel of this step
  int tp, grid, M, two_l, rank;
  int* tok; int* next_ids; int* finished; int* sequences;
  int max_len, eos, pad;
  int* cur_len; int* cache_len; int* step; int* n_unfinished;
};

__global__ void __launch_bounds__(256) mega_select_kernel(const SelectParams p) {
  __shared__ float sv[8];
  __shared__ int si[8];
  __shared__ int alive;
  const unsigned int gen_now = *reinterpret_cast<const volatile unsigned int*>(p.gen);
  const uint32_t tag = gen_now * (uint32_t)p.two_l;          // = (gen_now - 1) * 2L + 2L
  const int parity = (gen_now - 1u) & 1u;
  const int ncand = p.tp * p.grid;
  const uint2* base = p.cand + (size_t)parity * ncand * p.M * 2;
  if (threadIdx.x == 0) alive = 0;
  const int cl = *p.cur_len;
  for (int m = 0; m < p.M; ++m) {
    float bv = -INFINITY;
    int bi = 0x7fffffff;
    for (int cnd = threadIdx.x; cnd < ncand; cnd += blockDim.x) {
      const uint4* q = reinterpret_cast<const uint4*>(base + ((size_t)cnd * p.M + m) * 2);
      uint4 v = ld_relaxed_sys_v4(q);
      unsigned spins = 0;
      while (v.y != tag || v.w != tag) {
        v = ld_relaxed_sys_v4(q);
        if (++spins > (1u << 24)) mg_peer_timeout(p.rank, cnd / p.grid, tag, v.y, 1);
      }
      const float x = __uint_as_float(v.x);
      const int id = (int)v.z;
      if (x > bv || (x == bv && id < bi)) { bv = x; bi = id; }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, off);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, off);
      if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = bv; si[threadIdx.x >> 5] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int w = 1; w < 8; ++w)
        if (sv[w] > bv || (sv[w] == bv && si[w] < bi)) { bv = sv[w]; bi = si[w]; }
      int t = bi == 0x7fffffff ? 0 : bi;
      int f = p.finished[m];
      if (f) t = p.pad;
      f = f | (t == p.eos);
      p.finished[m] = f;
      p.tok[m] = t;
      p.next_ids[m] = t;
      if (cl < p.max_len) p.sequences[(size_t)m * p.max_len + cl] = t;
      if (!f) ++alive;
    }
  }
  if (threadIdx.x == 0) {
    *p.cur_len = cl + 1;
    *p.cache_len = *p.cache_len + 1;
    *p.step = *p.step + 1;
    *p.n_unfinished = alive;
  }
}

}  // namespace jl

using namespace jl;

// optional per-op timestamp buffer ([grid][L][32] u64, device memory; slot 31 = %smid); NULL disables tracing
static unsigned long long* g_mega_trace = nullptr;
extern "C" JL_API void jl_decode_mega_set_trace(void* buf) { g_mega_trace = (unsigned long long*)buf; }

static int mega_splits(int M, int nh, int nkv, int T, int num_sms) {
  const int G = nkv > 0 ? nh / nkv : 1;
  int splits = num_sms / (M * nkv);
  if (splits > 64) splits = 64;
  // a split shorter than one 128-key batch only adds partials to merge (tp = 8 has one KV head per rank:
  // 148 candidate splits for a 2k context): cap by the cache capacity over one load batch
  const int batch_keys = G <= 4 ? 128 : 64;           // keys one item handles per load batch (UNR x 64 in mg_attention_item)
  if (T > 0 && splits > (T + batch_keys - 1) / batch_keys) splits = (T + batch_keys - 1) / batch_keys;
  if (splits < 1) splits = 1;
  return splits;
}

// Packet buffers of one (batch, shape): residual x 2 parities, qkv, split-KV partials, attention output, SwiGLU
// output, local arg-max candidates.  Must be ZERO when first used (tag 0 never matches) and is never cleared again.
struct MegaLayout { size_t r, qkv, part, attn, h, cand, total; };
static MegaLayout mega_layout(int M, int H, int F, int nh, int nkv, int num_sms) {
  MegaLayout o;
  const int splits = mega_splits(M, nh, nkv, 0, num_sms);            // upper bound over all cache capacities
  size_t at = 0;
  auto take = [&](size_t packets) { const size_t off = at; at += (packets * 8 + 255) / 256 * 256; return off; };
  o.r = take((size_t)2 * M * (H / 2));
  o.qkv = take((size_t)2 * M * ((nh + 2 * nkv) * MG_HD / 2));          // x2: buffers alternate by layer parity
  o.part = take((size_t)2 * M * nh * splits * MG_PART);
  o.attn = take((size_t)2 * M * (nh * MG_HD / 2));
  o.h = take((size_t)2 * M * (F / 2));
  o.cand = take((size_t)2 * num_sms * M * 2);
  o.total = at;
  return o;
}
extern "C" JL_API size_t jl_decode_mega_scratch_bytes(int M, int H, int F, int nh, int nkv, int num_sms) {
  if (num_sms <= 0) num_sms = 148;
  return mega_layout(M, H, F, nh, nkv, num_sms).total;
}
// bytes of the two cross-GPU packet buffers (symmetric heap) for batches up to M on tp ranks
extern "C" JL_API size_t jl_decode_mega_peer_bytes(int M, int H, int tp, int num_sms, size_t* cand_offset) {
  if (num_sms <= 0) num_sms = 148;
  const size_t p_bytes = ((size_t)2 * tp * M * H * 8 + 1023) / 1024 * 1024;
  if (cand_offset) *cand_offset = p_bytes;
  return p_bytes + (size_t)2 * tp * num_sms * M * 2 * 8;
}

// Shared-memory plan: the staged activations (M rows of the widest GEMV input, whole k-stages), the attention
// scratch and whatever is left for the weight ring.  Returns the ring depth (< 4: the configuration does not fit).
static int mega_ring_stages(int M, int H, int F, int nh, int nkv, int hd, size_t* fixed_out, int w2_passes = 1) {
  const int G = nkv > 0 ? nh / nkv : 1;
  int fw = F;                                            // widest K range of w2 that is staged at once
  if (w2_passes == 2) fw = ((F + MG_KS - 1) / MG_KS + 1) / 2 * MG_KS;
  const int kmax = (H > fw ? H : fw) > nh * hd ? (H > fw ? H : fw) : nh * hd;
  const size_t xin_bytes = (size_t)M * ((kmax + MG_KS - 1) / MG_KS * MG_KS) * 2;      // whole k-stages (zero padded)
  const size_t attn_bytes = (size_t)(MG_CONSUMERS * 2) * G * (MG_HD + 2) * sizeof(float);
  const size_t fixed = 1024 + (xin_bytes > attn_bytes ? xin_bytes : attn_bytes) + 64;      // shared region (see kernel)
  const size_t budget = 218 * 1024;      // 227 KB per CTA minus ~7 KB of static shared memory (reduction buffers, residual slice)
  if (fixed_out) *fixed_out = fixed;
  if (fixed + 4 * (MG_STAGE_BYTES + 16) > budget) return 0;
  int stages = (int)((budget - fixed) / (MG_STAGE_BYTES + 16));
  return stages > 12 ? 12 : stages;
}
// ring depth jl_decode_mega would use for this shape, 0 when the batch does not fit (use the per-op path then)
// 1 or 2 K passes for w2: two when the batch is > 4, one pass does not fit, and the CTA's rows fit the carry buffer
static int mega_w2_passes(int M, int H, int F, int nh, int nkv, int hd, int num_sms) {
  const char* force = getenv("JLLAMA_MEGA_W2_PASSES");       // tests: force the two-pass walk at small shapes
  const bool rows_ok = F > MG_KS;
  if (M <= 4 || !rows_ok) return 1;
  if (force && atoi(force) == 2) return 2;
  if (mega_ring_stages(M, H, F, nh, nkv, hd, nullptr, 1) >= 4) return 1;
  return mega_ring_stages(M, H, F, nh, nkv, hd, nullptr, 2) >= 4 ? 2 : 1;
}
static bool mega_rows_fit(int H, int num_sms) {
  return ((H / MG_GROUP + num_sms - 1) / num_sms) * MG_GROUP <= MG_MAX_OWN;       // residual slice of one CTA
}
extern "C" JL_API int jl_decode_mega_stages(int M, int H, int F, int nh, int nkv, int hd) {
  if (M < 1 || M > 8 || hd != MG_HD || nkv <= 0 || nh % nkv || !mega_rows_fit(H, 148)) return 0;
  return mega_ring_stages(M, H, F, nh, nkv, hd, nullptr, mega_w2_passes(M, H, F, nh, nkv, hd, 148));
}

// One decode step of the whole model for M <= 8 rows.  Weight pointers are the per-op stacks [L·N, K].
extern "C" JL_API int jl_decode_mega(const void* wqkv_all, const void* wo_all, const void* w13_all, const void* w2_all,
                                     const void* lm_head, const void* attn_norm, const void* ffn_norm,
                                     const void* final_norm, const void* embedding, const int* ids, float* logits,
                                     int logits_ld, int logits_col0, int logits_multicast, void* k_cache, void* v_cache, const float* cos_t, const float* sin_t,
                                     const int* cache_len, const int* pad_left, void* scratch, unsigned int* gen, int M,
                                     int L, int H, int F, int V, int vocab_full, int nh, int nkv, int hd, int T,
                                     int max_pos, float eps, float scale, int num_sms, int tp, int rank,
                                     void* const* peer_partials, void* mc_partials, void* const* peer_cand,
                                     void* mc_cand, void* stream) {
  JL_CHECK_ARG(M >= 1 && M <= 8, "jl_decode_mega: M=%d outside 1..8", M);
  JL_CHECK_ARG(hd == MG_HD, "jl_decode_mega: head_dim=%d (only 128 is built)", hd);
  JL_CHECK_ARG(nkv > 0 && nh % nkv == 0, "jl_decode_mega: nh=%d not a multiple of nkv=%d", nh, nkv);
  const int G = nh / nkv;
  JL_CHECK_ARG(G == 1 || G == 2 || G == 4 || G == 8, "jl_decode_mega: GQA group %d not in {1,2,4,8}", G);
  const int NQKV = (nh + 2 * nkv) * hd;
  JL_CHECK_ARG(H % 256 == 0 && F % 256 == 0 && (nh * hd) % 256 == 0 && V % 8 == 0 && NQKV % 8 == 0,
               "jl_decode_mega: H=%d, F=%d and nh*hd must be multiples of 256, V=%d of 8", H, F, V);
  JL_CHECK_ARG(gen != nullptr && scratch != nullptr, "jl_decode_mega: scratch and the step counter are required");
  if (num_sms <= 0) num_sms = 148;
  JL_CHECK_ARG(mega_rows_fit(H, num_sms), "jl_decode_mega: H=%d gives a CTA more than %d residual rows on %d SMs", H,
               MG_MAX_OWN, num_sms);
  if (tp < 1) tp = 1;
  MegaTmaps tm;
  int rc;
  if ((rc = make_tmap_kblock_bf16(&tm.m[0], wqkv_all, (uint64_t)L * NQKV, H, H, MG_KS / 64, MG_GROUP))) return rc;
  if ((rc = make_tmap_kblock_bf16(&tm.m[1], wo_all, (uint64_t)L * H, nh * hd, nh * hd, MG_KS / 64, MG_GROUP))) return rc;
  if ((rc = make_tmap_kblock_bf16(&tm.m[2], w13_all, (uint64_t)L * 2 * F, H, H, MG_KS / 64, MG_GROUP))) return rc;
  if ((rc = make_tmap_kblock_bf16(&tm.m[3], w2_all, (uint64_t)L * H, F, F, MG_KS / 64, MG_GROUP))) return rc;
  if ((rc = make_tmap_kblock_bf16(&tm.m[4], lm_head, (uint64_t)V, H, H, MG_KS / 64, MG_GROUP))) return rc;
  MegaParams p{};
  p.L = L; p.H = H; p.F = F; p.V = V; p.nh = nh; p.nkv = nkv; p.NQKV = NQKV; p.vocab_full = vocab_full;
  p.eps = eps; p.scale_log2 = scale * 1.4426950408889634f;
  p.attn_norm = (const __nv_bfloat16*)attn_norm; p.ffn_norm = (const __nv_bfloat16*)ffn_norm;
  p.final_norm = (const __nv_bfloat16*)final_norm; p.embedding = (const __nv_bfloat16*)embedding;
  p.ids = ids; p.logits = logits; p.logits_ld = logits_ld > 0 ? logits_ld : V; p.logits_col0 = logits_col0;
  p.logits_mc = logits_multicast; p.kc = (__nv_bfloat16*)k_cache; p.vc = (__nv_bfloat16*)v_cache;
  p.cos_t = cos_t; p.sin_t = sin_t; p.cache_len = cache_len; p.pad_left = pad_left; p.gen = gen;
  p.splits = mega_splits(M, nh, nkv, T, num_sms);
  const MegaLayout lay = mega_layout(M, H, F, nh, nkv, num_sms);
  uint8_t* sb = (uint8_t*)scratch;
  p.r_pk = (uint2*)(sb + lay.r); p.qkv_pk = (uint2*)(sb + lay.qkv); p.part_pk = (uint2*)(sb + lay.part);
  p.attn_pk = (uint2*)(sb + lay.attn); p.h_pk = (uint2*)(sb + lay.h);
  p.T = T; p.max_pos = max_pos;
  p.ar.tp = tp; p.ar.rank = tp > 1 ? rank : 0;
  if (tp > 1) {
    JL_CHECK_ARG(tp <= 8 && rank >= 0 && rank < tp && peer_partials && peer_cand,
                 "jl_decode_mega: tensor-parallel launch needs the peer tables of the packet buffers");
    for (int i = 0; i < tp; ++i) { p.ar.p_peer[i] = (uint2*)peer_partials[i]; p.ar.cand_peer[i] = (uint2*)peer_cand[i]; }
    p.ar.p_mc = (uint2*)mc_partials; p.ar.cand_mc = (uint2*)mc_cand;
  } else {
    p.ar.cand_peer[0] = (uint2*)(sb + lay.cand);
  }
  p.trace = g_mega_trace;

  size_t fixed = 0;
  p.w2_passes = mega_w2_passes(M, H, F, nh, nkv, hd, num_sms);
  int stages = mega_ring_stages(M, H, F, nh, nkv, hd, &fixed, p.w2_passes);
  JL_CHECK_ARG(stages >= 4, "jl_decode_mega: M=%d, F=%d leave no room for the weight ring", M, F);
  {
    static const int cap = [] { const char* e = getenv("JLLAMA_MEGA_STAGES"); return e ? atoi(e) : 0; }();   // A/B: ring depth
    if (cap >= 4 && cap < stages) stages = cap;
  }
  p.stages = stages;
  {
    // measured on one box, ms/token: 0 → 2.92, 8 → 2.875, 16 → 2.871, 32 → 2.905, 64 → 2.906
    static const int ahead = [] { const char* e = getenv("JLLAMA_MEGA_L2_AHEAD"); return e ? atoi(e) : 16; }();
    p.l2_ahead = ahead;
  }
  const size_t smem = fixed + (size_t)stages * (MG_STAGE_BYTES + 16);

  void (*kern)(const MegaTmaps, const MegaParams) = nullptr;
#define JL_MG(M_, G_) if (M == M_ && G == G_) kern = decode_mega_kernel<M_, G_>;
#define JL_MG_ROW(M_) JL_MG(M_, 1) JL_MG(M_, 2) JL_MG(M_, 4) JL_MG(M_, 8)
  JL_MG_ROW(1) JL_MG_ROW(2) JL_MG_ROW(3) JL_MG_ROW(4) JL_MG_ROW(5) JL_MG_ROW(6) JL_MG_ROW(7) JL_MG_ROW(8)
#undef JL_MG_ROW
#undef JL_MG
  JL_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(num_sms);
  cfg.blockDim = dim3(MG_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;       // all CTAs co-resident: the packet polls cannot deadlock
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  count_launch();
  JL_CUDA(cudaLaunchKernelEx(&cfg, kern, tm, p));
  return JL_OK;
}

// Greedy token of every batch row from the candidates of the decode kernel that just ran on this stream, plus the
// sequence bookkeeping of jl_decode_advance.  `cand` = this rank's copy of the candidate packets (tp > 1: symmetric
// heap; tp == 1: inside the scratch of jl_decode_mega, see jl_decode_mega_cand_offset).
extern "C" JL_API size_t jl_decode_mega_cand_offset(int M, int H, int F, int nh, int nkv, int num_sms) {
  if (num_sms <= 0) num_sms = 148;
  return mega_layout(M, H, F, nh, nkv, num_sms).cand;
}
extern "C" JL_API int jl_mega_select(const void* cand, const unsigned int* gen, int tp, int rank, int num_sms, int M, int L,
                                     int* tok, int* next_ids, int* finished, int* sequences, int max_len, int eos,
                                     int pad, int* cur_len, int* cache_len, int* step, int* n_unfinished, void* stream) {
  JL_CHECK_ARG(cand && gen && M >= 1 && M <= 8 && tp >= 1 && tp <= 8, "jl_mega_select: bad arguments");
  if (num_sms <= 0) num_sms = 148;
  SelectParams p{};
  p.cand = (const uint2*)cand; p.gen = gen; p.tp = tp; p.grid = num_sms; p.M = M; p.two_l = 2 * L; p.rank = rank;
  p.tok = tok; p.next_ids = next_ids; p.finished = finished; p.sequences = sequences;
  p.max_len = max_len; p.eos = eos; p.pad = pad;
  p.cur_len = cur_len; p.cache_len = cache_len; p.step = step; p.n_unfinished = n_unfinished;
  count_launch();
  mega_select_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(p);
  JL_CUDA(cudaGetLastError());
  return JL_OK;
}
