// One-launch decode attention: RoPE(q, k_new) + KV-cache append + split-KV GQA attention + split merge.
//
// Replaces three launches of the unfused path (rope_kv_kernel → decode_attn_kernel → attn_combine_kernel)
// per layer; at batch 1 those launches cost more than the 8 MB of K/V they move.  Differences from
// decode_attn.cu:
//   * q and the new token's k, v are read straight from the fused qkv activation; the rotation is applied
//     in registers from the fp32 cos/sin table (position = cache_len − pad_left[b]);
//   * the new token is appended to the cache by split 0 and folded into its online softmax from registers,
//     so the streamed range is [pad_left, cache_len) and nothing written in this launch is re-read;
//   * many more, shorter splits (64 keys each at 2k context) to cover the load latency with parallelism;
//   * the last CTA of each (batch, kv-head) to finish — found with an atomic ticket — merges the partial
//     (m, l, o) triples, so there is no separate combine kernel.  The ticket counter is reset by that CTA.
// Reference semantics: /root/reference/jax_llama/model.py:201-300 at S = 1.
#include "common.cuh"
#include "jllama_b200.h"

namespace jl {

constexpr int DF_WARPS = 4;
constexpr int DF_KPI = 4;

struct FusedAttnParams {
  const __nv_bfloat16* qkv;      // [B, (nh + 2 nkv) · hd], not rotated
  __nv_bfloat16* kc;             // [B][nkv][T][hd]
  __nv_bfloat16* vc;
  const float* cos_t;            // [max_pos, hd/2]
  const float* sin_t;
  const int* cache_len_ptr;
  const int* pad_left;
  float* o_part;                 // [B·nh][splits][hd]
  float* ml_part;                // [B·nh][splits][2]
  int* tickets;                  // [B·nkv], zero on entry, zero on exit
  __nv_bfloat16* out;            // [B, nh·hd]
  int nkv, T, max_pos, splits;
  float scale_log2;
};

// Shared tail of both kernels: s_o / s_ml hold the (m, l, o) triples of the CTA's NG independent streams.  Folds them,
// then (splits > 1) the last CTA of this (batch, kv-head) to finish — found with an atomic ticket — merges the
// partial triples of all splits.  The ticket counter is reset by that CTA.
template <int NG, int G, int HD>
__device__ __forceinline__ void df_fold_and_merge(const FusedAttnParams& p, float (*s_o)[G][HD], float (*s_ml)[G][2],
                                                  int* s_last_p, const int b, const int kvh, const int split) {
  const int nh = p.nkv * G;
  int& s_last = *s_last_p;
  __syncthreads();
  for (int idx = threadIdx.x; idx < G * HD; idx += (int)blockDim.x) {
    const int g = idx / HD, d = idx - g * HD;
    float M = -INFINITY;
#pragma unroll
    for (int j = 0; j < NG; ++j) M = fmaxf(M, s_ml[j][g][0]);
    const float Ms = M == -INFINITY ? 0.f : M;
    float acc = 0.f, L = 0.f;
#pragma unroll
    for (int j = 0; j < NG; ++j) {
      const float w = exp2f(s_ml[j][g][0] - Ms);
      acc += w * s_o[j][g][d];
      L += w * s_ml[j][g][1];
    }
    const int h = kvh * G + g;
    if (p.splits == 1) {
      p.out[((size_t)b * nh + h) * HD + d] = __float2bfloat16(L > 0.f ? acc / L : 0.f);
    } else {
      const size_t pi = ((size_t)b * nh + h) * p.splits + split;
      p.o_part[pi * HD + d] = acc;
      if (d == 0) { p.ml_part[pi * 2] = M; p.ml_part[pi * 2 + 1] = L; }
    }
  }
  if (p.splits == 1) return;

  // ---- ticket: the last CTA of this (b, kv head) merges all splits
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    int* t = p.tickets + b * p.nkv + kvh;
    const int prev = atomicAdd(t, 1);
    s_last = prev == p.splits - 1;
    if (s_last) *t = 0;                                  // everyone has arrived: safe to re-arm for the next launch
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // stage the (m, l) pairs of all splits in shared memory (one parallel round trip), then every thread
  // folds its outputs with independent, unrolled loads
  float* s_m = &s_o[0][0][0];                       // reuse: G*splits floats each, far below NG*G*HD
  float* s_l = s_m + G * 64;
  __syncthreads();
  for (int i = threadIdx.x; i < G * p.splits; i += (int)blockDim.x) {
    const int g = i / p.splits, sp = i - g * p.splits;
    const size_t bh = (size_t)b * nh + kvh * G + g;
    s_m[g * 64 + sp] = __ldcg(p.ml_part + (bh * p.splits + sp) * 2);
    s_l[g * 64 + sp] = __ldcg(p.ml_part + (bh * p.splits + sp) * 2 + 1);
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < G * HD; idx += (int)blockDim.x) {
    const int g = idx / HD, d = idx - g * HD;
    const size_t bh = (size_t)b * nh + kvh * G + g;
    float M = -INFINITY;
    for (int sp = 0; sp < p.splits; ++sp) M = fmaxf(M, s_m[g * 64 + sp]);
    const float Ms = M == -INFINITY ? 0.f : M;
    float acc = 0.f, L = 0.f;
    const float* op = p.o_part + bh * p.splits * HD + d;
#pragma unroll 8
    for (int sp = 0; sp < p.splits; ++sp) {
      const float w = exp2f(s_m[g * 64 + sp] - Ms);
      const float ov = __ldcg(op + (size_t)sp * HD);
      acc += w > 0.f ? w * ov : 0.f;
      L += w * s_l[g * 64 + sp];
    }
    p.out[bh * HD + d] = __float2bfloat16(L > 0.f ? acc / L : 0.f);
  }
}

template <int G, int HD>
__global__ void __launch_bounds__(DF_WARPS * 32) decode_attn_fused_kernel(const FusedAttnParams p) {
  constexpr int LPK = HD / 8;
  constexpr int GPW = 32 / LPK;
  constexpr int NG = DF_WARPS * GPW;
  constexpr int KEYS_PER_ITER = NG * DF_KPI;
  __shared__ float s_o[NG][G][HD];
  __shared__ float s_ml[NG][G][2];
  __shared__ int s_last;

  pdl_wait();
  pdl_launch_dependents();

  const int split = blockIdx.x, kvh = blockIdx.y, b = blockIdx.z;
  const int nh = p.nkv * G;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int lig = lane % LPK;
  const int gid = warp * GPW + lane / LPK;
  const int group_base = lane - lig;
  const int kidx = ((lig / (LPK / 2)) & 1) * 2 + ((lig / (LPK / 4)) & 1);

  int len = *p.cache_len_ptr;                          // slots already filled; the new token takes slot `len`
  len = len > p.T ? p.T : len;
  const int first = p.pad_left ? p.pad_left[b] : 0;
  const int valid = len - first > 0 ? len - first : 0;
  int pos = len - first;
  pos = pos < 0 ? 0 : (pos >= p.max_pos ? p.max_pos - 1 : pos);
  int chunk = (valid + p.splits - 1) / p.splits;
  chunk = ((chunk + KEYS_PER_ITER - 1) / KEYS_PER_ITER) * KEYS_PER_ITER;
  const int lo = first + split * chunk;
  const int hi = lo + chunk < len ? lo + chunk : len;

  // rotation factors of this lane's 4 pairs
  const float4 cs4 = __ldg(reinterpret_cast<const float4*>(p.cos_t) + (size_t)pos * LPK + lig);
  const float4 sn4 = __ldg(reinterpret_cast<const float4*>(p.sin_t) + (size_t)pos * LPK + lig);
  const float cs[4] = {cs4.x, cs4.y, cs4.z, cs4.w}, sn[4] = {sn4.x, sn4.y, sn4.z, sn4.w};
  const __nv_bfloat16* row = p.qkv + (size_t)b * (nh + 2 * p.nkv) * HD;

  float qf[G][8];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const uint4 a = *reinterpret_cast<const uint4*>(row + (size_t)(kvh * G + g) * HD + lig * 8);
    const uint32_t aw[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float re = bf16_lo(aw[j]), im = bf16_hi(aw[j]);
      // round the rotated q to bf16 like the unfused path does, then fold the softmax scale in
      const uint32_t r = pack_bf16(re * cs[j] - im * sn[j], re * sn[j] + im * cs[j]);
      qf[g][2 * j] = bf16_lo(r) * p.scale_log2;
      qf[g][2 * j + 1] = bf16_hi(r) * p.scale_log2;
    }
  }
  float m[G], l[G], o[G][8];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    m[g] = -INFINITY;
    l[g] = 0.f;
#pragma unroll
    for (int e = 0; e < 8; ++e) o[g][e] = 0.f;
  }

  const size_t head_off = ((size_t)b * p.nkv + kvh) * (size_t)p.T * HD;
  const __nv_bfloat16* kbase = p.kc + head_off + lig * 8;
  const __nv_bfloat16* vbase = p.vc + head_off + lig * 8;

  // one online-softmax update with DF_KPI keys per lane-group (K/V rows already in registers)
  auto process = [&](const uint4 (&kr)[DF_KPI], const uint4 (&vr)[DF_KPI], const int key0) {
      float part[G][DF_KPI];
#pragma unroll
      for (int i = 0; i < DF_KPI; ++i) {
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
      float pr[G][DF_KPI];
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
        for (int i = 0; i < DF_KPI; ++i) {
          pr[g][i] = __shfl_sync(0xffffffffu, pe, group_base + (i >> 1) * (LPK / 2) + (i & 1) * (LPK / 4));
          psum += pr[g][i];
        }
        l[g] = l[g] * corr + psum;
#pragma unroll
        for (int e = 0; e < 8; ++e) o[g][e] *= corr;
      }
#pragma unroll
      for (int i = 0; i < DF_KPI; ++i) {
        const float vf[8] = {bf16_lo(vr[i].x), bf16_hi(vr[i].x), bf16_lo(vr[i].y), bf16_hi(vr[i].y),
                             bf16_lo(vr[i].z), bf16_hi(vr[i].z), bf16_lo(vr[i].w), bf16_hi(vr[i].w)};
#pragma unroll
        for (int g = 0; g < G; ++g)
#pragma unroll
          for (int e = 0; e < 8; ++e) o[g][e] = fmaf(pr[g][i], vf[e], o[g][e]);
      }
  };

  // UNR iterations' worth of K/V loads are issued back to back before any of them is consumed, so a
  // 64-key split is a single memory round trip (warp-uniform trip count; tails are predicated)
  constexpr int UNR = G <= 4 ? 2 : 1;
  for (int base = lo; base < hi; base += UNR * KEYS_PER_ITER) {
    uint4 kr[UNR][DF_KPI], vr[UNR][DF_KPI];
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const int key0 = base + u * KEYS_PER_ITER + gid * DF_KPI;
#pragma unroll
      for (int i = 0; i < DF_KPI; ++i) {
        const bool ok = key0 + i < hi;
        kr[u][i] = ok ? ldg_stream(kbase + (size_t)(key0 + i) * HD) : make_uint4(0, 0, 0, 0);
        vr[u][i] = ok ? ldg_stream(vbase + (size_t)(key0 + i) * HD) : make_uint4(0, 0, 0, 0);
      }
    }
#pragma unroll
    for (int u = 0; u < UNR; ++u) process(kr[u], vr[u], base + u * KEYS_PER_ITER + gid * DF_KPI);
  }

  // ---- the new token: rotate k, append k and v to the cache, fold it in from registers (split 0, warp 0)
  if (split == 0 && warp == 0) {
    const uint4 ka = *reinterpret_cast<const uint4*>(row + (size_t)(nh + kvh) * HD + lig * 8);
    const uint4 va = *reinterpret_cast<const uint4*>(row + (size_t)(nh + p.nkv + kvh) * HD + lig * 8);
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
      if (gid != 0) s = -INFINITY;                       // counted once, by lane-group 0
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

  // ---- fold the lane-groups of this CTA
#pragma unroll
  for (int g = 0; g < G; ++g) {
#pragma unroll
    for (int e = 0; e < 8; ++e) s_o[gid][g][lig * 8 + e] = o[g][e];
    if (lig == 0) { s_ml[gid][g][0] = m[g]; s_ml[gid][g][1] = l[g]; }
  }
  df_fold_and_merge<NG, G, HD>(p, s_o, s_ml, &s_last, b, kvh, split);
}

// =====================================================================================================================
// Tensor-core variant for batched decode (batch x kv-heads large enough to fill the GPU; head_dim 128).
//
// Why: at batch 32 / 2k context the kernel above moves 268 MB per layer in 133 µs (2 TB/s) — ncu shows ~1 issued
// instruction per cycle and no memory stall: the bf16 → fp32 unpacking and the G x 256 FMAs per key are the limit,
// not HBM.  Here both products go through mma.sync.m16n8k16 (bf16 in, fp32 accumulate) with the G query heads of
// the kv head as rows 0..G-1 of the 16-row operand, and K / V travel global memory → registers → MMA with no
// shared memory, no ldmatrix and no barrier in the loop; every warp runs an independent online softmax over its
// own 32-key blocks.
//   S = Q Kᵀ: the contraction index d may be permuted freely as long as Q and K use the same permutation, so thread
//     (g = lane / 4, t = lane % 4) loads the 16-byte chunks t, t + 4, t + 8, t + 12 of K[key g] (four LDG.128, each
//     one 64 contiguous bytes per quad = whole sectors) and k-step s = 2u + v uses words (2v, 2v+1) of chunk u as
//     (b0, b1); Q is held the same way.  Blocks two iterations ahead are requested into L2 with
//     cp.async.bulk.prefetch (no registers), so the register loads mostly hit L2.
//   O = P V: the C fragment of S (row g, keys 2t, 2t+1) is already the A fragment of P.  The B operand needs two
//     KEYS of one d packed per register: thread (g, t) loads V[key][8 g .. 8 g + 7] and V[key][64 + 8 g ..] for its
//     four keys (2t, 2t+1, 2t+8, 2t+9) and pairs them with PRMT; output tile (h, e) column n is d = 64 h + 8 n + e.
// The new token is folded in by the same code as a block whose only valid key is slot 0.
constexpr int DM_WARPS = 4;

__device__ __forceinline__ void mma_16816(float (&d)[4], uint32_t a0, uint32_t a2, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(0u), "r"(a2), "r"(0u), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}
__device__ __forceinline__ uint32_t u4w(const uint4& v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : (i == 2 ? v.z : v.w)); }

// rotate the four (re, im) pairs of one 16-byte chunk; pair j uses cos/sin entry j of cs / sn
__device__ __forceinline__ uint4 rope_chunk(const uint4 a, const float4 cs, const float4 sn) {
  const uint32_t w[4] = {a.x, a.y, a.z, a.w};
  const float c[4] = {cs.x, cs.y, cs.z, cs.w}, s[4] = {sn.x, sn.y, sn.z, sn.w};
  uint32_t r[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float re = bf16_lo(w[j]), im = bf16_hi(w[j]);
    r[j] = pack_bf16(re * c[j] - im * s[j], re * s[j] + im * c[j]);
  }
  return make_uint4(r[0], r[1], r[2], r[3]);
}

template <int G, int KPI>
__global__ void __launch_bounds__(DM_WARPS * 32, KPI == 32 ? 2 : 3) decode_attn_mma_kernel(const FusedAttnParams p) {
  constexpr int HD = 128;
  constexpr int NT = KPI / 8;                  // 8-key S tiles per block
  constexpr int NB = KPI / 16;                 // 16-key P·V steps per block
  __shared__ float s_o[DM_WARPS][G][HD];
  __shared__ float s_ml[DM_WARPS][G][2];
  __shared__ int s_last;

  pdl_wait();
  pdl_launch_dependents();

  const int split = blockIdx.x, kvh = blockIdx.y, b = blockIdx.z;
  const int nh = p.nkv * G;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;

  int len = *p.cache_len_ptr;
  len = len > p.T ? p.T : len;
  const int first = p.pad_left ? p.pad_left[b] : 0;
  const int valid = len - first > 0 ? len - first : 0;
  int pos = len - first;
  pos = pos < 0 ? 0 : (pos >= p.max_pos ? p.max_pos - 1 : pos);
  int chunk = (valid + p.splits - 1) / p.splits;
  chunk = ((chunk + DM_WARPS * KPI - 1) / (DM_WARPS * KPI)) * (DM_WARPS * KPI);
  const int lo = first + split * chunk;
  const int hi = lo + chunk < len ? lo + chunk : len;

  // K and V rows of a block are contiguous (KPI x 256 bytes each): one lane asks for them to be brought into L2
  const size_t head_off = ((size_t)b * p.nkv + kvh) * (size_t)p.T * HD;
  auto prefetch_block = [&](const int base) {
    if (lane == 0 && base < hi) {
      const int n = hi - base < KPI ? hi - base : KPI;
      const size_t off = head_off + (size_t)base * HD;
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p.kc + off), "r"(n * HD * 2) : "memory");
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p.vc + off), "r"(n * HD * 2) : "memory");
    }
  };
  prefetch_block(lo + warp * KPI);                       // overlaps the position → cos/sin → q dependency chain below
  prefetch_block(lo + warp * KPI + DM_WARPS * KPI);

  const __nv_bfloat16* row = p.qkv + (size_t)b * (nh + 2 * p.nkv) * HD;
  const float4* cos4 = reinterpret_cast<const float4*>(p.cos_t) + (size_t)pos * (HD / 8) + t;   // chunk t + 4u: entry 4u
  const float4* sin4 = reinterpret_cast<const float4*>(p.sin_t) + (size_t)pos * (HD / 8) + t;

  // Q of head kvh·G + g, rotated and rounded to bf16 (as the unfused path does): qa[s] = (a0, a2) of k-step s
  uint32_t qa[8][2];
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    uint4 r = make_uint4(0, 0, 0, 0);
    if (g < G)
      r = rope_chunk(*reinterpret_cast<const uint4*>(row + (size_t)(kvh * G + g) * HD + 8 * t + 32 * u),
                     __ldg(cos4 + 4 * u), __ldg(sin4 + 4 * u));
    qa[2 * u][0] = r.x; qa[2 * u][1] = r.y; qa[2 * u + 1][0] = r.z; qa[2 * u + 1][1] = r.w;
  }

  float m = -INFINITY, l = 0.f;                // row g: m is quad-uniform, l is this thread's share of the sum
  float o[2][8][4];
#pragma unroll
  for (int h = 0; h < 2; ++h)
#pragma unroll
    for (int e = 0; e < 8; ++e)
#pragma unroll
      for (int i = 0; i < 4; ++i) o[h][e][i] = 0.f;

  const __nv_bfloat16* kbase = p.kc + head_off + 8 * t;
  const __nv_bfloat16* vbase = p.vc + head_off + 8 * g;

  // one online-softmax update with a KPI-key block whose K / V fragments are in registers; keys >= limit are masked
  auto process = [&](const uint4 (&kr)[NT][4], const uint4 (&vr)[NB][4][2], const int base, const int limit) {
    float c[NT][4];
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      c[j][0] = c[j][1] = c[j][2] = c[j][3] = 0.f;
#pragma unroll
      for (int s = 0; s < 8; ++s)
        mma_16816(c[j], qa[s][0], qa[s][1], u4w(kr[j][s >> 1], (s & 1) * 2), u4w(kr[j][s >> 1], (s & 1) * 2 + 1));
    }
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      const int key = base + 8 * j + 2 * t;
      c[j][0] = key < limit ? c[j][0] * p.scale_log2 : -INFINITY;
      c[j][1] = key + 1 < limit ? c[j][1] * p.scale_log2 : -INFINITY;
      mx = fmaxf(mx, fmaxf(c[j][0], c[j][1]));
    }
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
    const float m_new = fmaxf(m, mx);
    const float m_safe = m_new == -INFINITY ? 0.f : m_new;
    const float corr = exp2f(m - m_safe);
    m = m_new;
    float psum = 0.f;
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      c[j][0] = exp2f(c[j][0] - m_safe);
      c[j][1] = exp2f(c[j][1] - m_safe);
      psum += c[j][0] + c[j][1];
    }
    l = l * corr + psum;
#pragma unroll
    for (int h = 0; h < 2; ++h)
#pragma unroll
      for (int e = 0; e < 8; ++e) { o[h][e][0] *= corr; o[h][e][1] *= corr; }
#pragma unroll
    for (int kb = 0; kb < NB; ++kb) {
      const uint32_t a0 = pack_bf16(c[2 * kb][0], c[2 * kb][1]), a2 = pack_bf16(c[2 * kb + 1][0], c[2 * kb + 1][1]);
#pragma unroll
      for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const uint32_t sel = (e & 1) ? 0x7632u : 0x5410u;
          const uint32_t b0 = prmt(u4w(vr[kb][0][h], e >> 1), u4w(vr[kb][1][h], e >> 1), sel);
          const uint32_t b1 = prmt(u4w(vr[kb][2][h], e >> 1), u4w(vr[kb][3][h], e >> 1), sel);
          mma_16816(o[h][e], a0, a2, b0, b1);
        }
    }
  };

  for (int base = lo + warp * KPI; base < hi; base += DM_WARPS * KPI) {
    prefetch_block(base + 2 * DM_WARPS * KPI);
    uint4 kr[NT][4], vr[NB][4][2];
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      const int key = base + 8 * j + g;
      const __nv_bfloat16* src = kbase + (size_t)key * HD;
#pragma unroll
      for (int u = 0; u < 4; ++u) kr[j][u] = key < hi ? ldg_stream(src + 32 * u) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int kb = 0; kb < NB; ++kb)
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int key = base + 16 * kb + 2 * t + (q & 1) + (q >> 1) * 8;
        const __nv_bfloat16* src = vbase + (size_t)key * HD;
        vr[kb][q][0] = key < hi ? ldg_stream(src) : make_uint4(0, 0, 0, 0);
        vr[kb][q][1] = key < hi ? ldg_stream(src + 64) : make_uint4(0, 0, 0, 0);
      }
    process(kr, vr, base, hi);
  }

  // ---- the new token (split 0, warp 0): rotate k, append k and v to the cache, fold it in as key 0 of a block
  if (split == 0 && warp == 0) {
    uint4 kr[NT][4], vr[NB][4][2];
#pragma unroll
    for (int j = 0; j < NT; ++j)
#pragma unroll
      for (int u = 0; u < 4; ++u) kr[j][u] = make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int kb = 0; kb < NB; ++kb)
#pragma unroll
      for (int q = 0; q < 4; ++q) vr[kb][q][0] = vr[kb][q][1] = make_uint4(0, 0, 0, 0);
    if (g == 0) {                                          // K fragment of key 0 lives in the quad g == 0
      const __nv_bfloat16* src = row + (size_t)(nh + kvh) * HD + 8 * t;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        kr[0][u] = rope_chunk(*reinterpret_cast<const uint4*>(src + 32 * u), __ldg(cos4 + 4 * u), __ldg(sin4 + 4 * u));
        if (len < p.T) *reinterpret_cast<uint4*>(p.kc + head_off + (size_t)len * HD + 8 * t + 32 * u) = kr[0][u];
      }
    }
    if (t == 0) {                                          // V fragments of key 0 = key slot 2t, t == 0
      const __nv_bfloat16* src = row + (size_t)(nh + p.nkv + kvh) * HD + 8 * g;
      vr[0][0][0] = *reinterpret_cast<const uint4*>(src);
      vr[0][0][1] = *reinterpret_cast<const uint4*>(src + 64);
      if (len < p.T) {
        *reinterpret_cast<uint4*>(p.vc + head_off + (size_t)len * HD + 8 * g) = vr[0][0][0];
        *reinterpret_cast<uint4*>(p.vc + head_off + (size_t)len * HD + 64 + 8 * g) = vr[0][0][1];
      }
    }
    process(kr, vr, 0, 1);
  }

  // ---- this warp's triple → shared memory; tile (h, e) holds columns d = 64 h + 8 n + e, n = 2t (c0), 2t + 1 (c1)
  l += __shfl_xor_sync(0xffffffffu, l, 1);
  l += __shfl_xor_sync(0xffffffffu, l, 2);
  if (g < G) {
#pragma unroll
    for (int h = 0; h < 2; ++h)
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        s_o[warp][g][64 * h + 16 * t + e] = o[h][e][0];
        s_o[warp][g][64 * h + 16 * t + 8 + e] = o[h][e][1];
      }
    if (t == 0) { s_ml[warp][g][0] = m; s_ml[warp][g][1] = l; }
  }
  df_fold_and_merge<DM_WARPS, G, HD>(p, s_o, s_ml, &s_last, b, kvh, split);
}

}  // namespace jl

using namespace jl;

static int g_attn_variant = 0;      // 0 = by shape, 1 = SIMT kernel, 2 = tensor-core kernel (head_dim 128 only)

// Pins the kernel behind jl_decode_attn_fused (tests and A/B measurements); 0 restores the shape-based choice.
extern "C" JL_API int jl_decode_attn_fused_set_variant(int variant) {
  JL_CHECK_ARG(variant >= 0 && variant <= 2, "jl_decode_attn_fused_set_variant: %d outside 0..2", variant);
  g_attn_variant = variant;
  return JL_OK;
}

extern "C" JL_API size_t jl_decode_attn_fused_scratch_bytes(int B, int nh, int nkv, int hd, int splits) {
  const size_t parts = splits <= 1 ? 0 : (size_t)B * nh * splits * (hd + 2) * sizeof(float);
  return parts + (size_t)B * nkv * sizeof(int);
}

extern "C" JL_API int jl_decode_attn_fused_splits(int B, int nkv, int max_len, int num_sms) {
  if (num_sms <= 0) num_sms = 148;
  int want = (4 * num_sms + B * nkv - 1) / (B * nkv);
  const int cap = (max_len + 127) / 128;                 // at least 128 keys per split
  int s = want < cap ? want : cap;
  return s < 1 ? 1 : (s > 64 ? 64 : s);
}

extern "C" JL_API int jl_decode_attn_fused(const void* qkv, void* k_cache, void* v_cache, const float* cos_t,
                                           const float* sin_t, const int* cache_len_ptr, const int* pad_left,
                                           void* scratch, void* out, int B, int nh, int nkv, int hd, int T,
                                           int max_pos, float scale, int splits, void* stream) {
  JL_CHECK_ARG(nkv > 0 && nh % nkv == 0, "jl_decode_attn_fused: nh=%d not a multiple of nkv=%d", nh, nkv);
  JL_CHECK_ARG(splits >= 1 && splits <= 64, "jl_decode_attn_fused: splits=%d outside 1..64", splits);
  JL_CHECK_ARG(scratch != nullptr && cache_len_ptr != nullptr, "jl_decode_attn_fused: scratch and cache_len are required");
  if (B == 0) return JL_OK;
  const int G = nh / nkv;
  FusedAttnParams p{};
  p.qkv = (const __nv_bfloat16*)qkv; p.kc = (__nv_bfloat16*)k_cache; p.vc = (__nv_bfloat16*)v_cache;
  p.cos_t = cos_t; p.sin_t = sin_t; p.cache_len_ptr = cache_len_ptr; p.pad_left = pad_left;
  const size_t part_floats = splits <= 1 ? 0 : (size_t)B * nh * splits * hd;
  p.o_part = (float*)scratch;
  p.ml_part = p.o_part + part_floats;
  p.tickets = (int*)(p.ml_part + (splits <= 1 ? 0 : (size_t)B * nh * splits * 2));
  p.out = (__nv_bfloat16*)out;
  p.nkv = nkv; p.T = T; p.max_pos = max_pos; p.splits = splits;
  p.scale_log2 = scale * 1.4426950408889634f;
  auto s = (cudaStream_t)stream;
  cudaError_t e = cudaErrorInvalidValue;
  // 16 or more (batch, kv-head) pairs: the tensor-core kernel (see its header; measured faster from 2 x 8 pairs up:
  // 19.5 vs 27.6 us at 2k context, 55 vs 162 us at batch 32); below that the SIMT kernel with many short splits
  const bool use_mma = hd == 128 && (g_attn_variant == 2 || (g_attn_variant == 0 && B * nkv >= 16));
  if (use_mma) {
    // one resident wave (2 CTAs per SM) of long CTAs: every CTA pays the position → q → first-block latency chain and
    // the fold / ticket / merge tail once, and with 2.6 waves of short CTAs that fixed cost was ~25 of 72 us at
    // batch 32.  The scratch layout (ticket offset) still follows the caller's `splits`.
    static int sms = 0;
    if (sms == 0) {
      int dev = 0;
      if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
        sms = 148;
    }
    // fewest splits whose CTA count fills whole waves of the 2 x sms resident slots best (e.g. 192 pairs x 3 = 576 of
    // 592 slots); at equal fill the smaller count wins (less fixed cost, no merge at 1)
    const int slots = 2 * sms, pairs = B * nkv;
    int ms = 1;
    float best = 0.f;
    for (int c = 1; c <= splits && c <= 16; ++c) {
      const int n = pairs * c;
      const float fill = (float)n / (float)(((n + slots - 1) / slots) * slots);
      if (fill > best + 0.02f) { best = fill; ms = c; }
    }
    p.splits = ms;
    splits = ms;
#define JL_DM(G_) \
    if (G == G_) e = launch_pdl(decode_attn_mma_kernel<G_, 32>, dim3(splits, nkv, B), dim3(DM_WARPS * 32), 0, s, p);
    JL_DM(1) else JL_DM(2) else JL_DM(4) else JL_DM(8)
#undef JL_DM
    if (e != cudaErrorInvalidValue) { JL_CUDA(e); return JL_OK; }
  }
#define JL_DF(G_, HD_) \
  if (G == G_ && hd == HD_) e = launch_pdl(decode_attn_fused_kernel<G_, HD_>, dim3(splits, nkv, B), dim3(DF_WARPS * 32), 0, s, p);
  JL_DF(1, 128) else JL_DF(2, 128) else JL_DF(4, 128) else JL_DF(8, 128)
  else JL_DF(1, 64) else JL_DF(2, 64) else JL_DF(4, 64) else JL_DF(8, 64)
  else {
    set_error("jl_decode_attn_fused: unsupported GQA group %d / head_dim %d (groups 1,2,4,8; head_dim 64,128)", G, hd);
    return JL_ERR_UNSUPPORTED;
  }
#undef JL_DF
  JL_CUDA(e);
  return JL_OK;
}
