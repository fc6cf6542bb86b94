// Causal prefill attention on warp-level tensor cores (mma.sync.m16n8k16) for head_dim 64 — and 128 on request.
//
// Why it exists: the tcgen05 flash kernel (flash_prefill.cu) is built for head_dim 128, the width of every LLaMA-1/2/3
// checkpoint the reference names.  LLaMA-architecture models with 64-wide heads (hidden 2048 / 32 heads: the 1B class)
// are loadable by the reference as well (/root/reference/jax_llama/model.py:115 takes any hidden / heads), so they get
// a kernel here instead of a refusal.  Same contract as jl_flash_prefill; reference semantics model.py:201-300 (S > 1).
//
// Shape of the kernel: one warp owns 16 query rows of one head and walks the visible keys in blocks of 32 (16 at head_dim
// 128) with an online softmax; nothing is shared between warps, so there is no shared memory and no barrier.  As in the batched
// decode kernel (decode_attn_fused.cu) K and V go global memory → registers → MMA fragments directly:
//   S = Q Kᵀ  the head dimension may be permuted as long as Q and K agree: thread (g = lane / 4, t = lane % 4) holds
//             the 16-byte chunks t, t + 4, … of its rows, k-step 2u + v uses words (2v, 2v + 1) of chunk u;
//   O = P V   the C fragment of S is the A fragment of P; the B operand pairs two keys of one column with PRMT and
//             output tile (h, e) column n is d = 64 h + 8 n + e.
// The 4 warps of a CTA take neighbouring row blocks of the same head, so their K / V reads hit L1 / L2.  (Two 16-row
// tiles per warp sharing the K / V fragments were measured at head_dim 64 and lost — 173 vs 160 us per 2k-token layer:
// fewer, fatter warps hide the load latency worse.)
#include "common.cuh"
#include "jllama_b200.h"

namespace jl {

constexpr int PM_WARPS = 4;
constexpr int PM_ROWS = 16;            // query rows per warp
// keys per iteration: 32 at head_dim 64, 16 at 128 (the K / V fragments of a block live in registers)

struct PrefillMmaParams {
  const __nv_bfloat16* q;              // [B, S, nh, HD] rotated
  const __nv_bfloat16* kc;             // [B, nkv, T, HD] (new keys already appended at [start, start + S))
  const __nv_bfloat16* vc;
  const int* pad_left;                 // [B] or null
  __nv_bfloat16* out;                  // [B, S, nh * HD]
  int S, nh, nkv, T, start;
  float scale_log2;
};

__device__ __forceinline__ void pm_mma(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                       uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pm_prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}
__device__ __forceinline__ uint32_t pm_w(const uint4& v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : (i == 2 ? v.z : v.w)); }

template <int HD>
__global__ void __launch_bounds__(PM_WARPS * 32) prefill_attn_mma_kernel(const PrefillMmaParams p) {
  constexpr int CH = HD / 32;                  // 16-byte chunks of a row per thread (Q, K)
  constexpr int KS = HD / 16;                  // k-steps of S = Q Kᵀ
  constexpr int VH = HD / 64;                  // 64-column halves of V / O
  constexpr int PM_KEYS = HD == 64 ? 32 : 16;
  constexpr int NT = PM_KEYS / 8, NB = PM_KEYS / 16;

  pdl_wait();
  pdl_launch_dependents();

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int h = blockIdx.y, b = blockIdx.z;
  const int G = p.nh / p.nkv, kvh = h / G;
  // longest row blocks first (the last rows see the most keys): block index descends with blockIdx.x
  const int q0 = ((int)gridDim.x - 1 - (int)blockIdx.x) * (PM_WARPS * PM_ROWS) + warp * PM_ROWS;
  if (q0 >= p.S) return;
  const int first = p.pad_left ? p.pad_left[b] : 0;
  const int r0 = q0 + g, r1 = q0 + g + 8;                                  // this thread's two query rows
  const int kmax = p.start + (q0 + PM_ROWS - 1 < p.S - 1 ? q0 + PM_ROWS - 1 : p.S - 1);   // last key any row of the warp sees

  // Q fragments: qa[s] = (a0, a1, a2, a3) of k-step s = 2u + v: rows r0 / r1, words (2v, 2v + 1) of chunk t + 4u
  uint32_t qa[KS][4];
  {
    const __nv_bfloat16* qb = p.q + (((size_t)b * p.S) * p.nh + h) * HD + 8 * t;
#pragma unroll
    for (int u = 0; u < CH; ++u) {
      const uint4 x0 = r0 < p.S ? *reinterpret_cast<const uint4*>(qb + (size_t)r0 * p.nh * HD + 32 * u) : make_uint4(0, 0, 0, 0);
      const uint4 x1 = r1 < p.S ? *reinterpret_cast<const uint4*>(qb + (size_t)r1 * p.nh * HD + 32 * u) : make_uint4(0, 0, 0, 0);
      qa[2 * u][0] = x0.x; qa[2 * u][1] = x1.x; qa[2 * u][2] = x0.y; qa[2 * u][3] = x1.y;
      qa[2 * u + 1][0] = x0.z; qa[2 * u + 1][1] = x1.z; qa[2 * u + 1][2] = x0.w; qa[2 * u + 1][3] = x1.w;
    }
  }
  float m[2] = {-INFINITY, -INFINITY}, l[2] = {0.f, 0.f};
  float o[VH][8][4];
#pragma unroll
  for (int hh = 0; hh < VH; ++hh)
#pragma unroll
    for (int e = 0; e < 8; ++e)
#pragma unroll
      for (int i = 0; i < 4; ++i) o[hh][e][i] = 0.f;

  const size_t head_off = ((size_t)b * p.nkv + kvh) * (size_t)p.T * HD;
  const __nv_bfloat16* kbase = p.kc + head_off + 8 * t;
  const __nv_bfloat16* vbase = p.vc + head_off + 8 * g;
  const int lim0 = p.start + r0, lim1 = p.start + r1;                      // causal limit of each row (inclusive)

  for (int base = first; base <= kmax; base += PM_KEYS) {
    uint4 kr[NT][CH], vr[NB][4][VH];
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      const int key = base + 8 * j + g;
      const __nv_bfloat16* src = kbase + (size_t)key * HD;
#pragma unroll
      for (int u = 0; u < CH; ++u) kr[j][u] = key <= kmax ? __ldg(reinterpret_cast<const uint4*>(src + 32 * u)) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int kb = 0; kb < NB; ++kb)
#pragma unroll
      for (int qd = 0; qd < 4; ++qd) {
        const int key = base + 16 * kb + 2 * t + (qd & 1) + (qd >> 1) * 8;
        const __nv_bfloat16* src = vbase + (size_t)key * HD;
#pragma unroll
        for (int hh = 0; hh < VH; ++hh)       // keys past the warp's horizon must read as zero: 0 x garbage would be NaN
          vr[kb][qd][hh] = key <= kmax ? __ldg(reinterpret_cast<const uint4*>(src + 64 * hh)) : make_uint4(0, 0, 0, 0);
      }
    float c[NT][4];
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      c[j][0] = c[j][1] = c[j][2] = c[j][3] = 0.f;
#pragma unroll
      for (int s = 0; s < KS; ++s)
        pm_mma(c[j], qa[s][0], qa[s][1], qa[s][2], qa[s][3], pm_w(kr[j][s >> 1], (s & 1) * 2), pm_w(kr[j][s >> 1], (s & 1) * 2 + 1));
    }
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      const int key = base + 8 * j + 2 * t;
      c[j][0] = key <= lim0 ? c[j][0] * p.scale_log2 : -INFINITY;
      c[j][1] = key + 1 <= lim0 ? c[j][1] * p.scale_log2 : -INFINITY;
      c[j][2] = key <= lim1 ? c[j][2] * p.scale_log2 : -INFINITY;
      c[j][3] = key + 1 <= lim1 ? c[j][3] * p.scale_log2 : -INFINITY;
      mx0 = fmaxf(mx0, fmaxf(c[j][0], c[j][1]));
      mx1 = fmaxf(mx1, fmaxf(c[j][2], c[j][3]));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    const float mn0 = fmaxf(m[0], mx0), mn1 = fmaxf(m[1], mx1);
    const float ms0 = mn0 == -INFINITY ? 0.f : mn0, ms1 = mn1 == -INFINITY ? 0.f : mn1;
    const float corr0 = exp2f(m[0] - ms0), corr1 = exp2f(m[1] - ms1);
    m[0] = mn0; m[1] = mn1;
    float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      c[j][0] = exp2f(c[j][0] - ms0); c[j][1] = exp2f(c[j][1] - ms0);
      c[j][2] = exp2f(c[j][2] - ms1); c[j][3] = exp2f(c[j][3] - ms1);
      ps0 += c[j][0] + c[j][1];
      ps1 += c[j][2] + c[j][3];
    }
    l[0] = l[0] * corr0 + ps0;
    l[1] = l[1] * corr1 + ps1;
#pragma unroll
    for (int hh = 0; hh < VH; ++hh)
#pragma unroll
      for (int e = 0; e < 8; ++e) { o[hh][e][0] *= corr0; o[hh][e][1] *= corr0; o[hh][e][2] *= corr1; o[hh][e][3] *= corr1; }
#pragma unroll
    for (int kb = 0; kb < NB; ++kb) {
      const uint32_t a0 = pack_bf16(c[2 * kb][0], c[2 * kb][1]), a1 = pack_bf16(c[2 * kb][2], c[2 * kb][3]);
      const uint32_t a2 = pack_bf16(c[2 * kb + 1][0], c[2 * kb + 1][1]), a3 = pack_bf16(c[2 * kb + 1][2], c[2 * kb + 1][3]);
#pragma unroll
      for (int hh = 0; hh < VH; ++hh)
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const uint32_t sel = (e & 1) ? 0x7632u : 0x5410u;
          const uint32_t b0 = pm_prmt(pm_w(vr[kb][0][hh], e >> 1), pm_w(vr[kb][1][hh], e >> 1), sel);
          const uint32_t b1 = pm_prmt(pm_w(vr[kb][2][hh], e >> 1), pm_w(vr[kb][3][hh], e >> 1), sel);
          pm_mma(o[hh][e], a0, a1, a2, a3, b0, b1);
        }
    }
  }

  // ---- normalise and store: thread (g, t) holds, per row and 64-column half, d = 64 hh + 16 t + [0, 16)
  l[0] += __shfl_xor_sync(0xffffffffu, l[0], 1); l[0] += __shfl_xor_sync(0xffffffffu, l[0], 2);
  l[1] += __shfl_xor_sync(0xffffffffu, l[1], 1); l[1] += __shfl_xor_sync(0xffffffffu, l[1], 2);
  const float inv0 = l[0] > 0.f ? 1.f / l[0] : 0.f, inv1 = l[1] > 0.f ? 1.f / l[1] : 0.f;
#pragma unroll
  for (int hh = 0; hh < VH; ++hh) {
    uint32_t w0[8], w1[8];                   // columns 16 t + [0, 8) come from c0 of tiles e, + [8, 16) from c1
#pragma unroll
    for (int e = 0; e < 8; e += 2) {
      w0[e / 2] = pack_bf16(o[hh][e][0] * inv0, o[hh][e + 1][0] * inv0);
      w0[4 + e / 2] = pack_bf16(o[hh][e][1] * inv0, o[hh][e + 1][1] * inv0);
      w1[e / 2] = pack_bf16(o[hh][e][2] * inv1, o[hh][e + 1][2] * inv1);
      w1[4 + e / 2] = pack_bf16(o[hh][e][3] * inv1, o[hh][e + 1][3] * inv1);
    }
    if (r0 < p.S) {
      uint4* dst = reinterpret_cast<uint4*>(p.out + (((size_t)b * p.S + r0) * p.nh + h) * HD + 64 * hh + 16 * t);
      dst[0] = make_uint4(w0[0], w0[1], w0[2], w0[3]);
      dst[1] = make_uint4(w0[4], w0[5], w0[6], w0[7]);
    }
    if (r1 < p.S) {
      uint4* dst = reinterpret_cast<uint4*>(p.out + (((size_t)b * p.S + r1) * p.nh + h) * HD + 64 * hh + 16 * t);
      dst[0] = make_uint4(w1[0], w1[1], w1[2], w1[3]);
      dst[1] = make_uint4(w1[4], w1[5], w1[6], w1[7]);
    }
  }
}

}  // namespace jl

using namespace jl;

extern "C" JL_API int jl_prefill_attn_mma(const void* q, const void* k_cache, const void* v_cache, const int* pad_left,
                                          void* out, int B, int S, int nh, int nkv, int hd, int T, int start, float scale,
                                          void* stream) {
  JL_CHECK_ARG(hd == 64 || hd == 128, "jl_prefill_attn_mma: head_dim=%d (64 and 128 are built)", hd);
  JL_CHECK_ARG(nkv > 0 && nh % nkv == 0, "jl_prefill_attn_mma: nh=%d not a multiple of nkv=%d", nh, nkv);
  JL_CHECK_ARG(start >= 0 && start + S <= T, "jl_prefill_attn_mma: start=%d + S=%d exceeds cache capacity %d", start, S, T);
  if (B * S == 0) return JL_OK;
  PrefillMmaParams p{(const __nv_bfloat16*)q, (const __nv_bfloat16*)k_cache, (const __nv_bfloat16*)v_cache, pad_left,
                     (__nv_bfloat16*)out, S, nh, nkv, T, start, scale * 1.4426950408889634f};
  const dim3 grid((S + PM_WARPS * PM_ROWS - 1) / (PM_WARPS * PM_ROWS), nh, B);
  auto s = (cudaStream_t)stream;
  if (hd == 64) JL_CUDA(launch_pdl(prefill_attn_mma_kernel<64>, grid, dim3(PM_WARPS * 32), 0, s, p));
  else JL_CUDA(launch_pdl(prefill_attn_mma_kernel<128>, grid, dim3(PM_WARPS * 32), 0, s, p));
  return JL_OK;
}
