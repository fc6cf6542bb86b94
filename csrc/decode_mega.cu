// Persistent whole-model decode step ("megakernel") for small batches (M <= 8 tokens).
//
// Why: a decode step of an 8B model is ~225 dependent kernels whose useful work is 5-40 µs each.  Measured
// with ncu on B200, every one of them loses 5-10 µs to CTA launch ramp, prologue and tail (SMs active for
// only ~50% of a 19 µs `wo` GEMV), and programmatic dependent launch does not recover it because the
// streaming kernels fill the SM.  Here ONE cooperative kernel with one CTA per SM runs the entire step:
//
//   * warp 0 is a TMA producer that streams every weight matrix of the model, layer after layer, through a
//     shared-memory ring: one 16 KB request per stage ([16 k-blocks][8 rows][128 B], 128-byte swizzle).  It
//     depends on nothing but free ring slots, so it runs ahead across op boundaries: while the consumers sit in
//     a grid barrier or do attention, the next op's weights are already landing.
//   * warps 1..8 are consumers: RMSNorm + tensor-core GEMV (ldmatrix + mma.sync, see decode_mega_mma.cuh),
//     residual / SwiGLU epilogues, RoPE + KV append + split-KV GQA attention (one (batch, kv-head, split) item
//     per CTA, merge spread over the grid), final norm + LM head.
//   * ops are separated by a grid-wide barrier (monotonic atomic counter, acquire spin with a watchdog);
//     activations cross CTAs through L2 (ld.global.cg), never through stale L1 lines.
//   * rows of each GEMV are cut into 148 contiguous, near-equal ranges (±1 group of 8 rows).
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
// runs as one compiled XLA program — this is the hand-written equivalent.
#include <stdlib.h>

#include "jllama_b200.h"
#include "tcgen05.cuh"

namespace jl {

constexpr int MG_CONSUMERS = 8;
constexpr int MG_THREADS = (MG_CONSUMERS + 1) * 32;
constexpr int MG_CT = MG_CONSUMERS * 32;           // consumer threads
constexpr int MG_GROUP = 8;                        // weight rows per stage (one per consumer warp)
constexpr int MG_KS = 1024;                        // K elements per stage
constexpr int MG_STAGE_BYTES = MG_GROUP * MG_KS * 2;   // 16 KB
constexpr int MG_HD = 128;
constexpr int MG_KPI = 4;

struct alignas(64) MegaTmaps {
  CUtensorMap m[5];          // wqkv_all, wo_all, w13_all, w2_all, lm_head
};

// Tensor-parallel all-reduce state (tp == 1: unused).  Same inbox / flag protocol as comm.cu.
struct MegaAr {
  float* inbox[8];                   // peer p's inbox, mapped into this process: [parity][src][M][H] fp32
  int* flags[8];                     // peer p's flag array
  const int* seq_ptr;                // device epoch counter (bumped once per step by the host-side sequence)
  size_t parity_stride;              // floats between the two parity halves
  int flag_offset, tp, rank;
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
  __nv_bfloat16* x;                  // [M, H]   residual stream
  __nv_bfloat16* qkv;                // [M, NQKV]
  __nv_bfloat16* attn;               // [M, nh·hd]
  __nv_bfloat16* hmlp;               // [M, F]
  float* logits;                     // [M, V]
  __nv_bfloat16* kc;                 // [L][M][nkv][T][hd]
  __nv_bfloat16* vc;
  const float* cos_t;
  const float* sin_t;
  const int* cache_len;
  const int* pad_left;
  float* o_part;                     // [M·nh][splits][hd]
  float* ml_part;                    // [M·nh][splits][2]
  int* tickets;                      // [M·nkv]
  unsigned int* bar_counter;         // monotonically increasing arrival counter
  unsigned int* bar_gen;             // barriers completed by previous launches
  int T, max_pos, splits, stages, l2_ahead;
  int w2_passes;                     // 2: w2 is walked in two K ranges (batch rows of its full input do not fit in smem)
  unsigned long long* trace;         // optional [2][L][16] globaltimer stamps (block 0 and the last block)
};

enum { MG_PLAIN = 0, MG_RESIDUAL = 1, MG_SWIGLU = 2, MG_LOGITS = 3 };

__device__ __forceinline__ void mg_consumer_bar() { asm volatile("bar.sync 1, %0;" ::"n"(MG_CT) : "memory"); }
__device__ __forceinline__ uint4 ldcg16(const void* p) { return __ldcg(reinterpret_cast<const uint4*>(p)); }

__device__ __forceinline__ void mg_stamp(const MegaParams& p, int layer, int ev) {
  if (p.trace != nullptr && threadIdx.x == 32) {
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

// contiguous, near-equal split of `n` groups over the grid
__device__ __forceinline__ void mg_range(int n, int& g0, int& g1) {
  g0 = (int)(((long long)n * blockIdx.x) / gridDim.x);
  g1 = (int)(((long long)n * (blockIdx.x + 1)) / gridDim.x);
}

struct MegaCtx {
  uint32_t ring, bars;
  const uint8_t* ring_ptr;
  uint4* xin;                 // [M][K/8] staged activation (bf16)
  float* s_attn;              // attention fold scratch
  int stages;
  int stage; uint32_t phase;  // ring cursor (producer and consumers keep separate copies)
  const MegaAr* ar;           // tensor-parallel state; ar_epoch tags (and its parity places) the current all-reduce
  int ar_epoch;
};

// row-parallel epilogue under tensor parallelism: hand this rank's fp32 partial to every peer's inbox.
// Low-latency protocol: every value travels as one 8-byte store {value, epoch}; the receiver polls the slot
// itself until the tag matches, so no flag, no fence and a single NVLink traversal are on the critical path
// (a release-flag version measured ~11 µs per all-reduce at tp = 2).
template <int M>
__device__ __forceinline__ void mg_push_partial(const MegaCtx& c, int m, int row, int N, float v) {
  const MegaAr& a = *c.ar;
  const size_t off = (size_t)(c.ar_epoch & 1) * (a.parity_stride / 2) + ((size_t)a.rank * M + m) * N + row;
  const uint2 pkt = make_uint2(__float_as_uint(v), (unsigned)c.ar_epoch);
  for (int peer = 0; peer < a.tp; ++peer) {
    uint2* dst = reinterpret_cast<uint2*>(a.inbox[peer]) + off;
    asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(dst), "r"(pkt.x), "r"(pkt.y) : "memory");
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

// grid-wide barrier among the consumer warps of all CTAs
__device__ __forceinline__ void mg_grid_barrier(const MegaParams& p, unsigned int gen_base, int& bar_idx) {
  mg_consumer_bar();                                   // all of this CTA's global writes are issued
  if (threadIdx.x == 32) {
    __threadfence();
    const unsigned int target = (gen_base + (unsigned)bar_idx + 1u) * gridDim.x;
    atomicAdd(p.bar_counter, 1u);
    unsigned int spins = 0;
    while ((int)(ld_acquire_gpu_u32(p.bar_counter) - target) < 0) {
      if (++spins > (1u << 26)) {
        printf("jllama_b200: grid barrier %d timed out in block %d\n", bar_idx, (int)blockIdx.x);
        __trap();
      }
    }
  }
  ++bar_idx;
  mg_consumer_bar();
}

// RoPE + KV append + split-KV attention for one (b, kv head, split) item + ticket merge (8 consumer warps)
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
                                                  int kvh, int split, int* s_last, const MgPos& pp) {
  constexpr int HD = MG_HD, LPK = HD / 8, GPW = 32 / LPK, NG = MG_CONSUMERS * GPW;
  constexpr int KEYS_PER_ITER = NG * MG_KPI;
  const int nh = p.nkv * G;
  const int lane = threadIdx.x & 31, warp = (threadIdx.x >> 5) - 1, ct = threadIdx.x - 32;
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
  const __nv_bfloat16* row = p.qkv + (size_t)b * p.NQKV;
  // the new token's k, v are fetched together with q (consumed after the KV loop by split 0 / warp 0)
  const bool owns_new = split == 0 && warp == 0;
  uint4 ka = make_uint4(0, 0, 0, 0), va = ka;
  if (owns_new) {
    ka = ldcg16(row + (size_t)(nh + kvh) * HD + lig * 8);
    va = ldcg16(row + (size_t)(nh + p.nkv + kvh) * HD + lig * 8);
  }

  float qf[G][8];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const uint4 a = ldcg16(row + (size_t)(kvh * G + g) * HD + lig * 8);
    const uint32_t aw[4] = {a.x, a.y, a.z, a.w};
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
  const size_t head_off = (((size_t)layer * M + b) * p.nkv + kvh) * (size_t)p.T * HD;
  const __nv_bfloat16* kbase = p.kc + head_off + lig * 8;
  const __nv_bfloat16* vbase = p.vc + head_off + lig * 8;

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
  constexpr int UNR = G <= 4 ? 2 : 1;
  for (int base = lo; base < hi; base += UNR * KEYS_PER_ITER) {
    uint4 kr[UNR][MG_KPI], vr[UNR][MG_KPI];
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
      p.attn[((size_t)b * nh + h) * HD + d] = __float2bfloat16(Ls > 0.f ? acc / Ls : 0.f);
    } else {
      const size_t pi = ((size_t)b * nh + h) * p.splits + split;
      p.o_part[pi * HD + d] = acc;
      if (d == 0) { p.ml_part[pi * 2] = Mx; p.ml_part[pi * 2 + 1] = Ls; }
    }
  }
  (void)s_last;
  mg_stamp(p, layer, 20);
  mg_consumer_bar();                 // s_attn is reused by this CTA's next item
}

// Merge of the split-KV partials, spread over the whole grid (one output element per thread): runs after a
// grid barrier.  The earlier "last CTA of each kv head merges" ticket scheme put atomic + two dependent L2
// round trips of a single CTA on the critical path of every layer (13 µs per attention phase).
__device__ __forceinline__ void mg_attn_merge(const MegaParams& p, int M) {
  const int ct = threadIdx.x - 32;
  const int total = M * p.nh * MG_HD;
  for (int idx = blockIdx.x * MG_CT + ct; idx < total; idx += gridDim.x * MG_CT) {
    const int bh = idx / MG_HD, d = idx - bh * MG_HD;
    const float2* ml = reinterpret_cast<const float2*>(p.ml_part) + (size_t)bh * p.splits;
    const float* op = p.o_part + (size_t)bh * p.splits * MG_HD + d;
    float acc = 0.f, Ls = 0.f;
    if (p.splits <= 32) {
      // issue every load of this output up front: one L2 round trip instead of two dependent ones
      float2 e[32];
      float ov[32];
#pragma unroll
      for (int sp = 0; sp < 32; ++sp) {
        if (sp < p.splits) { e[sp] = __ldcg(ml + sp); ov[sp] = __ldcg(op + (size_t)sp * MG_HD); }
        else { e[sp] = make_float2(-INFINITY, 0.f); ov[sp] = 0.f; }
      }
      float Mx = -INFINITY;
#pragma unroll
      for (int sp = 0; sp < 32; ++sp) Mx = fmaxf(Mx, e[sp].x);
      const float Ms = Mx == -INFINITY ? 0.f : Mx;
#pragma unroll
      for (int sp = 0; sp < 32; ++sp) {
        const float w = exp2f(e[sp].x - Ms);
        acc += w > 0.f ? w * ov[sp] : 0.f;
        Ls += w * e[sp].y;
      }
    } else {
      float Mx = -INFINITY;
#pragma unroll 8
      for (int sp = 0; sp < p.splits; ++sp) Mx = fmaxf(Mx, __ldcg(ml + sp).x);
      const float Ms = Mx == -INFINITY ? 0.f : Mx;
#pragma unroll 8
      for (int sp = 0; sp < p.splits; ++sp) {
        const float2 e = __ldcg(ml + sp);
        const float w = exp2f(e.x - Ms);
        const float ov = __ldcg(op + (size_t)sp * MG_HD);
        acc += w > 0.f ? w * ov : 0.f;
        Ls += w * e.y;
      }
    }
    p.attn[idx] = __float2bfloat16(Ls > 0.f ? acc / Ls : 0.f);
  }
}

// Second half of the in-kernel one-shot all-reduce (tp > 1): publish this CTA's slice to every peer, wait for
// the same-index CTA of every peer (all ranks cut rows identically), reduce in rank order (bit-identical on
// all ranks), add the residual and store the new residual-stream rows.  The local grid barrier follows.
template <int M>
__device__ __forceinline__ void mg_ar_finish(const MegaParams& p, const MegaCtx& c, int epoch, int N) {
  const MegaAr& a = p.ar;
  const int ct = threadIdx.x - 32;
  int g0, g1;
  mg_range(N / MG_GROUP, g0, g1);
  const int nrows = (g1 - g0) * MG_GROUP;
  const uint2* mine = reinterpret_cast<const uint2*>(a.inbox[a.rank]) + (size_t)(epoch & 1) * (a.parity_stride / 2);
  for (int i = ct; i < nrows * M; i += MG_CT) {
    const int m = i / nrows, row = g0 * MG_GROUP + (i - m * nrows);
    float s = 0.f;
    for (int src = 0; src < a.tp; ++src) {
      const uint2* slot = mine + ((size_t)src * M + m) * N + row;
      uint2 pkt;
      unsigned spins = 0;
      for (;;) {
        asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(pkt.x), "=r"(pkt.y) : "l"(slot) : "memory");
        if ((int)pkt.y == epoch) break;
        if (++spins > (1u << 26)) {
          printf("jllama_b200: rank %d timed out waiting for peer %d (all-reduce epoch %d, block %d)\n", a.rank, src, epoch,
                 (int)blockIdx.x);
          __trap();
        }
      }
      s += __uint_as_float(pkt.x);
    }
    const float y = __bfloat162float(__float2bfloat16(s)) + __bfloat162float(__ldcg(p.x + (size_t)m * N + row));
    p.x[(size_t)m * N + row] = __float2bfloat16(y);
  }
}

template <int M, int G>
__global__ void __launch_bounds__(MG_THREADS, 1)
decode_mega_kernel(const __grid_constant__ MegaTmaps tm, const MegaParams p) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ float red[M][MG_CONSUMERS];
  __shared__ __align__(16) float s_acc[2][MG_CONSUMERS][MG_MMA_MAX_M][8];
  __shared__ float s_carry[M > 4 ? M : 1][64];         // two-pass w2: sums of the first K range (this CTA's rows)
  __shared__ volatile int s_idle_flag;          // consumers → producer: "no GEMV is consuming the ring right now"
  volatile int* s_idle = &s_idle_flag;
  if (threadIdx.x == 0) s_idle_flag = 0;
  __shared__ int s_last;
  constexpr int NG = MG_CONSUMERS * (32 / (MG_HD / 8));

  MegaCtx c;
  c.stages = p.stages;
  c.ring = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
  c.ring_ptr = smem_raw + (c.ring - tc::smem_u32(smem_raw));
  c.bars = c.ring + p.stages * MG_STAGE_BYTES;
  uint8_t* after = const_cast<uint8_t*>(c.ring_ptr) + p.stages * MG_STAGE_BYTES + 16 * p.stages;
  c.s_attn = reinterpret_cast<float*>(after);
  const int kmax = p.H > p.F ? p.H : p.F;
  // the staged GEMV input and the attention scratch are never live together (a GEMV ends before attention
  // starts and is restaged after it), so they share one region sized for the larger of the two
  c.xin = reinterpret_cast<uint4*>(after);
  (void)kmax;
  c.stage = 0; c.phase = 0;
  c.ar = &p.ar; c.ar_epoch = 0;

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
  if (warp == 0) {
    // ===================================================== producer: the whole model's weights, in order
    if (lane == 0) {
      // `ld` feeds the ring.  While the consumers are NOT consuming (attention, merge and the barriers around
      // them: s_idle is set) and the ring is full, `pf` walks up to p.l2_ahead stages further down the same
      // sequence and only asks the L2 to fetch them, so HBM keeps streaming through the ~12 µs per layer in
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
#define MG_STAGE(SRC_, K_, NORM_) mg_stage_input_mma<M>(c, SRC_, K_, NORM_, p.eps, red)
#define MG_GEMV(MODE_, N_, K_, OUT_, RES_) mg_gemv_mma<M, MODE_>(c, N_, K_, OUT_, RES_, s_acc)
  const unsigned int gen_base = *reinterpret_cast<volatile unsigned int*>(p.bar_gen);
  int bar_idx = 0;
  const MgPos pos0 = mg_pos(p, 0, lane % (MG_HD / 8));      // batch row 0: constant for the whole step
  const int ar_base = p.ar.tp > 1 ? __ldcg(p.ar.seq_ptr) : 0;      // epoch of collective j is ar_base + j + 1
  const int ct = threadIdx.x - 32;

  // embedding rows → residual stream (each CTA copies a slice), then everyone sees x
  {
    const int H8 = p.H >> 3;
    for (int i = blockIdx.x * MG_CT + ct; i < M * H8; i += gridDim.x * MG_CT) {
      const int m = i / H8, j = i - m * H8;
      int id = p.ids[m];
      id = id < 0 ? 0 : (id >= p.vocab_full ? p.vocab_full - 1 : id);
      reinterpret_cast<uint4*>(p.x)[i] = __ldg(reinterpret_cast<const uint4*>(p.embedding) + (size_t)id * H8 + j);
    }
  }
  mg_grid_barrier(p, gen_base, bar_idx);

  for (int l = 0; l < p.L; ++l) {
    // ---- fused qkv projection (RMSNorm in the prologue)
    mg_stamp(p, l, 0);
    MG_STAGE(p.x, p.H, p.attn_norm + (size_t)l * p.H);
    mg_stamp(p, l, 1);
    MG_GEMV(MG_PLAIN, p.NQKV, p.H, p.qkv, nullptr);
    mg_stamp(p, l, 2);
    mg_grid_barrier(p, gen_base, bar_idx);
    mg_stamp(p, l, 3);
    if (threadIdx.x == 32) *s_idle = 1;
    // ---- attention
    {
      const int items = M * p.nkv * p.splits;
      for (int it = blockIdx.x; it < items; it += gridDim.x) {
        const int split = it % p.splits, bk = it / p.splits;
        const int bb = bk / p.nkv;
        mg_attention_item<G>(p, c, l, M, bb, bk % p.nkv, split, &s_last, M == 1 ? pos0 : mg_pos(p, bb, lane % (MG_HD / 8)));
      }
    }
    mg_stamp(p, l, 4);
    mg_grid_barrier(p, gen_base, bar_idx);
    if (p.splits > 1) {                      // grid-uniform: every CTA takes the same branch
      mg_attn_merge(p, M);
      mg_grid_barrier(p, gen_base, bar_idx);
    }
    mg_stamp(p, l, 5);
    // ---- wo + residual
    MG_STAGE(p.attn, qdim, nullptr);
    mg_stamp(p, l, 6);
    if (threadIdx.x == 32) *s_idle = 0;
    c.ar_epoch = ar_base + 2 * l + 1;
    MG_GEMV(MG_RESIDUAL, p.H, qdim, p.x, p.x);
    if (p.ar.tp > 1) mg_ar_finish<M>(p, c, ar_base + 2 * l + 1, p.H);
    mg_stamp(p, l, 7);
    mg_grid_barrier(p, gen_base, bar_idx);
    mg_stamp(p, l, 8);
    // ---- w1/w3 + SwiGLU (RMSNorm in the prologue)
    MG_STAGE(p.x, p.H, p.ffn_norm + (size_t)l * p.H);
    mg_stamp(p, l, 9);
    MG_GEMV(MG_SWIGLU, 2 * p.F, p.H, p.hmlp, nullptr);
    mg_stamp(p, l, 10);
    mg_grid_barrier(p, gen_base, bar_idx);
    mg_stamp(p, l, 11);
    // ---- w2 + residual
    bool w2_done = false;
    if constexpr (M > 4) {
      if (p.w2_passes == 2) {
        const int ka = ((p.F + MG_KS - 1) / MG_KS + 1) / 2 * MG_KS;       // first K range: whole k-stages
        mg_stage_input_mma<M>(c, p.hmlp, ka, nullptr, p.eps, red, p.F);
        mg_stamp(p, l, 12);
        c.ar_epoch = ar_base + 2 * l + 2;
        mg_gemv_mma<M, MG_RESIDUAL, 1>(c, p.H, ka, p.x, p.x, s_acc, s_carry);
        mg_stage_input_mma<M>(c, p.hmlp + ka, p.F - ka, nullptr, p.eps, red, p.F);
        mg_gemv_mma<M, MG_RESIDUAL, 2>(c, p.H, p.F - ka, p.x, p.x, s_acc, s_carry);
        w2_done = true;
      }
    }
    if (!w2_done) {
      MG_STAGE(p.hmlp, p.F, nullptr);
      mg_stamp(p, l, 12);
      c.ar_epoch = ar_base + 2 * l + 2;
      MG_GEMV(MG_RESIDUAL, p.H, p.F, p.x, p.x);
    }
    if (p.ar.tp > 1) mg_ar_finish<M>(p, c, ar_base + 2 * l + 2, p.H);
    mg_stamp(p, l, 13);
    mg_grid_barrier(p, gen_base, bar_idx);
    mg_stamp(p, l, 14);
  }
  // ---- final norm + LM head
  MG_STAGE(p.x, p.H, p.final_norm);
  MG_GEMV(MG_LOGITS, p.V, p.H, p.logits, nullptr);
  mg_grid_barrier(p, gen_base, bar_idx);
  if (blockIdx.x == 0 && threadIdx.x == 32) *p.bar_gen = gen_base + (unsigned)bar_idx;
}

}  // namespace jl

using namespace jl;

extern "C" JL_API int jl_decode_mega_barriers(int L) { return 2 + 5 * L; }

// optional per-op timestamp buffer ([grid][L][32] u64, device memory; slot 31 = %smid); NULL disables tracing
static unsigned long long* g_mega_trace = nullptr;
extern "C" JL_API void jl_decode_mega_set_trace(void* buf) { g_mega_trace = (unsigned long long*)buf; }

extern "C" JL_API size_t jl_decode_mega_scratch_bytes(int M, int nh, int nkv, int num_sms) {
  int splits = num_sms / (M * nkv);
  if (splits < 1) splits = 1;
  if (splits > 64) splits = 64;
  const size_t parts = (size_t)M * nh * splits * (MG_HD + 2) * sizeof(float);
  return parts + (size_t)M * nkv * sizeof(int) + 2 * sizeof(unsigned int) + 64;
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
  const size_t budget = 220 * 1024;      // 227 KB per CTA minus ~5 KB of static shared memory (reduction buffers)
  if (fixed_out) *fixed_out = fixed;
  if (fixed + 4 * (MG_STAGE_BYTES + 16) > budget) return 0;
  int stages = (int)((budget - fixed) / (MG_STAGE_BYTES + 16));
  return stages > 12 ? 12 : stages;
}
// ring depth jl_decode_mega would use for this shape, 0 when the batch does not fit (use the per-op path then)
// 1 or 2 K passes for w2: two when the batch is > 4, one pass does not fit, and the CTA's rows fit the carry buffer
static int mega_w2_passes(int M, int H, int F, int nh, int nkv, int hd, int num_sms) {
  const char* force = getenv("JLLAMA_MEGA_W2_PASSES");       // tests: force the two-pass walk at small shapes
  const bool rows_ok = ((H / MG_GROUP + num_sms - 1) / num_sms) * MG_GROUP <= 64 && F > MG_KS;
  if (M <= 4 || !rows_ok) return 1;
  if (force && atoi(force) == 2) return 2;
  if (mega_ring_stages(M, H, F, nh, nkv, hd, nullptr, 1) >= 4) return 1;
  return mega_ring_stages(M, H, F, nh, nkv, hd, nullptr, 2) >= 4 ? 2 : 1;
}
extern "C" JL_API int jl_decode_mega_stages(int M, int H, int F, int nh, int nkv, int hd) {
  if (M < 1 || M > 8 || hd != MG_HD || nkv <= 0 || nh % nkv) return 0;
  return mega_ring_stages(M, H, F, nh, nkv, hd, nullptr, mega_w2_passes(M, H, F, nh, nkv, hd, 148));
}

// One decode step of the whole model for M <= 8 rows.  Weight pointers are the per-op stacks [L·N, K].
extern "C" JL_API int jl_decode_mega(const void* wqkv_all, const void* wo_all, const void* w13_all, const void* w2_all,
                                     const void* lm_head, const void* attn_norm, const void* ffn_norm,
                                     const void* final_norm, const void* embedding, const int* ids, void* x, void* qkv,
                                     void* attn, void* hmlp, float* logits, void* k_cache, void* v_cache,
                                     const float* cos_t, const float* sin_t, const int* cache_len, const int* pad_left,
                                     void* scratch, int M, int L, int H, int F, int V, int vocab_full, int nh, int nkv,
                                     int hd, int T, int max_pos, float eps, float scale, int num_sms, int tp, int rank,
                                     void* const* peer_inbox, void* const* peer_flags, size_t inbox_parity_stride_floats,
                                     int flag_offset, const int* seq_ptr, void* stream) {
  JL_CHECK_ARG(M >= 1 && M <= 8, "jl_decode_mega: M=%d outside 1..8", M);
  JL_CHECK_ARG(hd == MG_HD, "jl_decode_mega: head_dim=%d (only 128 is built)", hd);
  JL_CHECK_ARG(nkv > 0 && nh % nkv == 0, "jl_decode_mega: nh=%d not a multiple of nkv=%d", nh, nkv);
  const int G = nh / nkv;
  JL_CHECK_ARG(G == 1 || G == 2 || G == 4 || G == 8, "jl_decode_mega: GQA group %d not in {1,2,4,8}", G);
  const int NQKV = (nh + 2 * nkv) * hd;
  JL_CHECK_ARG(H % 256 == 0 && F % 256 == 0 && (nh * hd) % 256 == 0 && V % 8 == 0 && NQKV % 8 == 0,
               "jl_decode_mega: H=%d, F=%d and nh*hd must be multiples of 256, V=%d of 8", H, F, V);
  if (num_sms <= 0) num_sms = 148;
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
  p.ids = ids; p.x = (__nv_bfloat16*)x; p.qkv = (__nv_bfloat16*)qkv; p.attn = (__nv_bfloat16*)attn;
  p.hmlp = (__nv_bfloat16*)hmlp; p.logits = logits; p.kc = (__nv_bfloat16*)k_cache; p.vc = (__nv_bfloat16*)v_cache;
  p.cos_t = cos_t; p.sin_t = sin_t; p.cache_len = cache_len; p.pad_left = pad_left;
  int splits = num_sms / (M * nkv);
  if (splits > 64) splits = 64;
  // a split shorter than one 128-key batch only adds partials to merge (tp = 8 has one KV head per rank:
  // 148 candidate splits for a 2k context): cap by the cache capacity over one load batch
  const int batch_keys = G <= 4 ? 128 : 64;           // keys one item handles per load batch (UNR x 64 in mg_attention_item)
  if (splits > (T + batch_keys - 1) / batch_keys) splits = (T + batch_keys - 1) / batch_keys;
  if (splits < 1) splits = 1;
  p.splits = splits;
  const size_t part_floats = (size_t)M * nh * splits * MG_HD;
  p.o_part = (float*)scratch;
  p.ml_part = p.o_part + part_floats;
  p.tickets = (int*)(p.ml_part + (size_t)M * nh * splits * 2);
  uintptr_t bp = (reinterpret_cast<uintptr_t>(p.tickets + M * nkv) + 15) & ~(uintptr_t)15;
  p.bar_counter = reinterpret_cast<unsigned int*>(bp);
  p.bar_gen = p.bar_counter + 1;
  p.T = T; p.max_pos = max_pos;
  p.ar.tp = tp < 1 ? 1 : tp; p.ar.rank = rank; p.ar.seq_ptr = seq_ptr; p.ar.parity_stride = inbox_parity_stride_floats;
  p.ar.flag_offset = flag_offset;
  if (p.ar.tp > 1) {
    JL_CHECK_ARG(p.ar.tp <= 8 && peer_inbox && peer_flags && seq_ptr, "jl_decode_mega: tensor-parallel launch needs peer tables");
    JL_CHECK_ARG((size_t)p.ar.tp * M * H * 2 <= inbox_parity_stride_floats, "jl_decode_mega: inbox too small for the all-reduce");
    for (int i = 0; i < p.ar.tp; ++i) { p.ar.inbox[i] = (float*)peer_inbox[i]; p.ar.flags[i] = (int*)peer_flags[i]; }
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
  attr[0].id = cudaLaunchAttributeCooperative;       // all CTAs co-resident: the grid barrier cannot deadlock
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  count_launch();
  JL_CUDA(cudaLaunchKernelEx(&cfg, kern, tm, p));
  return JL_OK;
}
