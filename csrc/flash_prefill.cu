// Causal flash attention for prefill on tcgen05 tensor cores (head_dim 128, GQA, left padding).
//
// The reference materialises softmax(QKᵀ/√d + bias) as a [B, nh, Q, K] tensor with an explicit mask
// (/root/reference/jax_llama/model.py:236-288).  Here one CTA owns a 128-query tile of one head:
//   S = Q·Kᵀ   tcgen05.mma (A = Q, B = K tile, both K-major in smem via TMA) → TMEM, double-buffered;
//   softmax    four warps, one query row per thread: tcgen05.ld the row, causal / left-pad predicates
//              in registers (no mask tensor), online max/sum in the log2 domain, P → bf16 → smem in
//              the 128-byte-swizzled K-major layout the tensor core reads;
//   O_j = P·V  tcgen05.mma with V as an MN-major B operand (the cache row layout, no transpose);
//   O accumulates in registers with the online-softmax correction; S_{j+1} is issued before P_j is
//   consumed so the tensor core works on the next tile while the softmax warps finish this one.
// K/V tiles stream through a 2-stage TMA ring; only tiles intersecting [pad_left, diagonal] are visited.
#include "jllama_b200.h"
#include "tcgen05.cuh"

namespace jl {

constexpr int FA_BM = 128, FA_BN = 128, FA_HD = 128;
constexpr int FA_THREADS = 256;
constexpr uint32_t FA_TILE_BYTES = 128 * 128 * 2;   // one [128 x 128] bf16 tile = two swizzled 16 KB halves
constexpr uint32_t FA_HALF_BYTES = 128 * 64 * 2;

struct FlashParams {
  __nv_bfloat16* out;
  const int* pad_left;
  int B, S, nh, nkv, T, start;
  float scale_log2;
};

__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__global__ void __launch_bounds__(FA_THREADS, 1)
flash_prefill_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                     const __grid_constant__ CUtensorMap tmap_v, const FlashParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = base, sP = base + FA_TILE_BYTES;
  auto sK = [&](int st) { return base + 2 * FA_TILE_BYTES + st * 2 * FA_TILE_BYTES; };
  auto sV = [&](int st) { return sK(st) + FA_TILE_BYTES; };
  const uint32_t bars = base + 6 * FA_TILE_BYTES;
  const uint32_t q_full = bars, p_full = bars + 8, o_full = bars + 16;
  auto kv_full = [&](int s) { return bars + 24 + 8u * s; };
  auto kv_empty = [&](int s) { return bars + 40 + 8u * s; };
  auto s_full = [&](int s) { return bars + 56 + 8u * s; };
  const uint32_t tmem_slot = bars + 72;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - tc::smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * FA_BM, h = blockIdx.y, b = blockIdx.z;
  const int kvh = h / (p.nh / p.nkv);

  if (warp == 0 && lane == 0) {
    tc::prefetch_tensormap(&tmap_q); tc::prefetch_tensormap(&tmap_k); tc::prefetch_tensormap(&tmap_v);
  }
  if (warp == 1 && lane == 0) {
    tc::mbar_init(q_full, 1); tc::mbar_init(p_full, 4); tc::mbar_init(o_full, 1);
    for (int s = 0; s < 2; ++s) { tc::mbar_init(kv_full(s), 1); tc::mbar_init(kv_empty(s), 1); tc::mbar_init(s_full(s), 1); }
    tc::fence_barrier_init();
  }
  if (warp == 2) tc::tmem_alloc(tmem_slot, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *tmem_slot_ptr;
  const uint32_t tS[2] = {tmem, tmem + 128};
  const uint32_t tO = tmem + 256;

  pdl_wait();
  pdl_launch_dependents();

  // visible key range for this query tile: [pad_left[b], start + q_last]
  const int pad = p.pad_left ? p.pad_left[b] : 0;
  const int q_hi = min(q0 + FA_BM, p.S) - 1;                 // last valid query row of the tile
  const int first_tile = pad / FA_BN;
  const int last_tile = (p.start + q_hi) / FA_BN;
  // a tile whose queries are all left padding sees no key at all: n_tiles <= 0, every role skips its
  // loop and the epilogue writes zeros (those rows are don't-care)
  const int n_tiles = last_tile - first_tile + 1 > 0 ? last_tile - first_tile + 1 : 0;

  if (warp == 0) {
    if (lane == 0) {
      tc::mbar_expect_tx(q_full, FA_TILE_BYTES);
      tc::tma_load_2d(sQ, &tmap_q, q_full, h * FA_HD, b * p.S + q0, tc::EVICT_FIRST);
      tc::tma_load_2d(sQ + FA_HALF_BYTES, &tmap_q, q_full, h * FA_HD + 64, b * p.S + q0, tc::EVICT_FIRST);
      for (int j = 0; j < n_tiles; ++j) {
        const int st = j & 1; const uint32_t ph = (j >> 1) & 1;
        tc::mbar_wait(kv_empty(st), ph ^ 1);
        const int key0 = (first_tile + j) * FA_BN;
        tc::mbar_expect_tx(kv_full(st), 2 * FA_TILE_BYTES);
        tc::tma_load_3d(sK(st), &tmap_k, kv_full(st), 0, key0, b * p.nkv + kvh, tc::EVICT_LAST);
        tc::tma_load_3d(sK(st) + FA_HALF_BYTES, &tmap_k, kv_full(st), 64, key0, b * p.nkv + kvh, tc::EVICT_LAST);
        tc::tma_load_3d(sV(st), &tmap_v, kv_full(st), 0, key0, b * p.nkv + kvh, tc::EVICT_LAST);
        tc::tma_load_3d(sV(st) + FA_HALF_BYTES, &tmap_v, kv_full(st), 64, key0, b * p.nkv + kvh, tc::EVICT_LAST);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t IDESC_S = tc::idesc_bf16_f32(FA_BM, FA_BN, false, false);
      constexpr uint32_t IDESC_O = tc::idesc_bf16_f32(FA_BM, FA_HD, false, true);    // B = V is MN-major
      auto issue_s = [&](int j) {
        const int st = j & 1;
        tc::mbar_wait(kv_full(st), (j >> 1) & 1);
        tc::fence_after_sync();
#pragma unroll
        for (int k = 0; k < FA_HD / 16; ++k) {
          const uint32_t off = (k >> 2) * FA_HALF_BYTES + (k & 3) * 32;     // 64-wide halves, 32 B per K step
          tc::mma_ss(tS[j & 1], tc::smem_desc_kmajor_sw128(sQ + off), tc::smem_desc_kmajor_sw128(sK(st) + off),
                     IDESC_S, k != 0);
        }
        tc::mma_commit(s_full(j & 1));
      };
      tc::mbar_wait(q_full, 0);
      if (n_tiles > 0) issue_s(0);
      for (int j = 0; j < n_tiles; ++j) {
        if (j + 1 < n_tiles) issue_s(j + 1);           // tensor core computes S_{j+1} during softmax_j
        tc::mbar_wait(p_full, j & 1);
        tc::fence_after_sync();
        const int st = j & 1;
#pragma unroll
        for (int k = 0; k < FA_BN / 16; ++k) {
          // A = P (K-major over keys): halves of 64 keys; B = V (MN-major): 16 keys = 2 KB per K step
          const uint32_t aoff = (k >> 2) * FA_HALF_BYTES + (k & 3) * 32;
          tc::mma_ss(tO, tc::smem_desc_kmajor_sw128(sP + aoff),
                     tc::smem_desc_mnmajor_sw128(sV(st) + k * 2048, FA_HALF_BYTES, 1024), IDESC_O, k != 0);
        }
        tc::mma_commit(o_full);
        tc::mma_commit(kv_empty(st));
      }
    }
  } else if (warp >= 4) {
    const int quarter = warp & 3;
    const int r = quarter * 32 + lane;                     // query row inside the tile
    const int q_slot = p.start + q0 + r;                   // cache slot (= causal limit) of this query
    const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
    float o_acc[FA_HD];
#pragma unroll
    for (int i = 0; i < FA_HD; ++i) o_acc[i] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;

    for (int j = 0; j < n_tiles; ++j) {
      const int key0 = (first_tile + j) * FA_BN;
      const bool need_mask = (key0 + FA_BN - 1 > p.start + q0) || (key0 < pad);     // warp-uniform
      tc::mbar_wait(s_full(j & 1), (j >> 1) & 1);
      tc::fence_after_sync();
      const uint32_t ts = tS[j & 1] + lane_off;
      // ---- pass 1: row maximum of the (masked) scores
      float mx = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        tc::tmem_ld32(ts + c * 32, v);
        tc::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float s = __uint_as_float(v[i]);
          if (need_mask) {
            const int key = key0 + c * 32 + i;
            if (key > q_slot || key < pad) s = -INFINITY;
          }
          mx = fmaxf(mx, s);
        }
      }
      const float m_new = fmaxf(m_run, mx * p.scale_log2);
      const float m_safe = m_new == -INFINITY ? 0.f : m_new;
      const float corr = exp2f(m_run - m_safe);
      m_run = m_new;
      // ---- pass 2: p = exp2(s·scale − m), row sum, P → bf16 → swizzled smem (A operand of P·V)
      float psum = 0.f;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        tc::tmem_ld32(ts + c * 32, v);
        tc::tmem_ld_wait();
        uint32_t w[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float s0 = __uint_as_float(v[2 * i]), s1 = __uint_as_float(v[2 * i + 1]);
          if (need_mask) {
            const int key = key0 + c * 32 + 2 * i;
            if (key > q_slot || key < pad) s0 = -INFINITY;
            if (key + 1 > q_slot || key + 1 < pad) s1 = -INFINITY;
          }
          const float p0 = exp2f(s0 * p.scale_log2 - m_safe), p1 = exp2f(s1 * p.scale_log2 - m_safe);
          w[i] = pack_bf16(p0, p1);
          // sum what the tensor core will actually multiply (the bf16-rounded probabilities)
          psum += bf16_lo(w[i]) + bf16_hi(w[i]);
        }
        // 32 keys = 64 B = four 16-byte chunks of this row; half = c>>1, chunk index = (c&1)*4 + q
        const uint32_t row_base = sP + (c >> 1) * FA_HALF_BYTES + r * 128;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint32_t chunk = ((c & 1) * 4 + q) ^ (r & 7);
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(row_base + chunk * 16), "r"(w[4 * q]),
                       "r"(w[4 * q + 1]), "r"(w[4 * q + 2]), "r"(w[4 * q + 3]) : "memory");
        }
      }
      l_run = l_run * corr + psum;
      fence_proxy_async();                 // generic-proxy smem writes → visible to the tensor core
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(p_full);
      // ---- O += P·V (tile result comes back through TMEM; accumulate with the max correction)
      tc::mbar_wait(o_full, j & 1);
      tc::fence_after_sync();
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        tc::tmem_ld32(tO + lane_off + c * 32, v);
        tc::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o_acc[c * 32 + i] = o_acc[c * 32 + i] * corr + __uint_as_float(v[i]);
      }
      tc::fence_before_sync();
    }
    // ---- epilogue: normalise and store this row (256 contiguous bytes)
    if (q0 + r < p.S) {
      const float inv = l_run > 0.f ? 1.0f / l_run : 0.f;
      uint4* dst = reinterpret_cast<uint4*>(p.out + ((size_t)(b * p.S + q0 + r) * p.nh + h) * FA_HD);
#pragma unroll
      for (int i = 0; i < FA_HD / 8; ++i)
        dst[i] = make_uint4(pack_bf16(o_acc[8 * i] * inv, o_acc[8 * i + 1] * inv),
                            pack_bf16(o_acc[8 * i + 2] * inv, o_acc[8 * i + 3] * inv),
                            pack_bf16(o_acc[8 * i + 4] * inv, o_acc[8 * i + 5] * inv),
                            pack_bf16(o_acc[8 * i + 6] * inv, o_acc[8 * i + 7] * inv));
    }
  }

  tc::fence_before_sync();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc(tmem, 512);
}

}  // namespace jl

using namespace jl;

extern "C" JL_API int jl_flash_prefill(const void* q, const void* k_cache, const void* v_cache,
                                       const int* cache_len_ptr, const int* pad_left, void* out, int B, int S,
                                       int nh, int nkv, int hd, int T, int start, float scale, void* stream) {
  (void)cache_len_ptr;
  JL_CHECK_ARG(hd == FA_HD, "jl_flash_prefill: head_dim=%d (only 128 is built)", hd);
  JL_CHECK_ARG(nkv > 0 && nh % nkv == 0, "jl_flash_prefill: nh=%d not a multiple of nkv=%d", nh, nkv);
  JL_CHECK_ARG(start >= 0 && start + S <= T, "jl_flash_prefill: start=%d + S=%d exceeds cache capacity %d", start, S, T);
  if (B * S == 0) return JL_OK;
  CUtensorMap tq, tk, tv;
  int rc = make_tmap_2d_bf16(&tq, q, (uint64_t)B * S, (uint64_t)nh * hd, (uint64_t)nh * hd, FA_BM);
  if (rc) return rc;
  rc = make_tmap_3d_bf16(&tk, k_cache, (uint64_t)B * nkv, T, hd, (uint64_t)T * hd, hd, FA_BN);
  if (rc) return rc;
  rc = make_tmap_3d_bf16(&tv, v_cache, (uint64_t)B * nkv, T, hd, (uint64_t)T * hd, hd, FA_BN);
  if (rc) return rc;
  FlashParams p{(__nv_bfloat16*)out, pad_left, B, S, nh, nkv, T, start, scale * 1.4426950408889634f};
  const size_t smem = 6 * FA_TILE_BYTES + 128 + 1024;
  JL_CUDA(cudaFuncSetAttribute(flash_prefill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((S + FA_BM - 1) / FA_BM, nh, B);
  JL_CUDA(launch_pdl(flash_prefill_kernel, grid, dim3(FA_THREADS), smem, (cudaStream_t)stream, tq, tk, tv, p));
  return JL_OK;
}
