// Causal flash attention for prefill on tcgen05 tensor cores (head_dim 128, GQA, left padding).
//
// The reference materialises softmax(QKᵀ/√d + bias) as a [B, nh, Q, K] tensor with an explicit mask
// (/root/reference/jax_llama/model.py:236-288).  Here one CTA owns a 128-query tile of one head and walks
// the visible keys in tiles of 64; two such CTAs share an SM so that one CTA's softmax runs under the other's
// MMAs (each SM sub-partition then has two softmax warps to switch between):
//   S = Q·Kᵀ   tcgen05.mma (A = Q, B = K tile, both K-major in smem via TMA) → TMEM, double-buffered;
//   softmax    four warps, one query row per thread: ONE tcgen05.ld pass of the 64 scores into registers,
//              causal / left-pad predicates in registers (no mask tensor), exponentials in the log2 domain
//              against a per-row reference maximum, P → bf16 → smem in the 128-byte-swizzled K-major layout;
//   O += P·V   tcgen05.mma with V as an MN-major B operand (the cache row layout, no transpose), accumulating
//              IN TMEM across all key tiles.  The reference maximum is only raised when the running maximum
//              has grown by more than 2^8 (then the warp rescales its 32 rows of O in TMEM with
//              tcgen05.ld/st), so the common iteration never touches O; probabilities stay ≤ 256, exact in
//              bf16's range, and the final 1/l normalisation makes the result independent of the reference.
//   S_{j+1} is issued before P_j is consumed, K and V tiles stream through a 4-slot 16 KB TMA ring, query
//   tiles are scheduled longest-first, and only tiles intersecting [pad_left, diagonal] are visited.
// flash_prefill_v1_kernel is the first version (128-key tiles, O accumulated in registers through a TMEM
// round trip per tile, one CTA per SM); kept selectable (JLLAMA_FLASH_V1=1) as the A/B yardstick.
#include <cstdio>
#include <cstdlib>
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
flash_prefill_v1_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
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


// ------------------------------------------------------------------------------------------------ v2
constexpr int F2_BN = 64, F2_THREADS = 192, F2_SLOTS = 4;
constexpr uint32_t F2_SLOT_BYTES = F2_BN * FA_HD * 2;        // 16 KB: one K or V tile = two swizzled 8 KB halves
constexpr uint32_t F2_KV_HALF = F2_BN * 64 * 2;              // 8 KB
constexpr uint32_t F2_P_BYTES = FA_BM * F2_BN * 2;           // 16 KB
// Q tile + (P tile unless it lives in TMEM) + K/V ring, then barriers and alignment slack: 97 KB with P in TMEM
// (two CTAs per SM), 113 KB otherwise (one CTA per SM)
__host__ __device__ constexpr uint32_t f2_tiles_bytes(bool p_tmem) {
  return FA_TILE_BYTES + (p_tmem ? 0u : F2_P_BYTES) + F2_SLOTS * F2_SLOT_BYTES;
}
__host__ __device__ constexpr uint32_t f2_smem_bytes(bool p_tmem) { return f2_tiles_bytes(p_tmem) + 1024; }
constexpr float F2_RESCALE_THRESHOLD = 8.0f;                 // log2 units

// P_TMEM: the probabilities go back to the tensor core through TMEM (tcgen05.st over the columns of the
// score tile they were computed from; A operand of P·V read from TMEM) instead of swizzled shared memory:
// no generic→async proxy fence, and the softmax warps no longer wait for P·V_{j-1} before producing P_j.
template <bool P_TMEM>
__global__ void __launch_bounds__(256, 2)   // 256: caps registers at 128 so that two 6-warp CTAs (allocated as 8) fit
flash_prefill_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                     const __grid_constant__ CUtensorMap tmap_v, const FlashParams p) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const uint32_t raw = tc::smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t sQ = base, sP = base + FA_TILE_BYTES;
  auto slot_addr = [&](int s) {
    return base + FA_TILE_BYTES + (P_TMEM ? 0u : F2_P_BYTES) + (uint32_t)s * F2_SLOT_BYTES;
  };
  const uint32_t bars = base + f2_tiles_bytes(P_TMEM);
  const uint32_t q_full = bars, p_full = bars + 8, o_full = bars + 16;
  auto slot_full = [&](int s) { return bars + 24 + 8u * s; };
  auto slot_empty = [&](int s) { return bars + 56 + 8u * s; };
  auto s_full = [&](int s) { return bars + 88 + 8u * s; };
  const uint32_t tmem_slot = bars + 104;
  if (bars + 112 > raw + f2_smem_bytes(P_TMEM)) __trap();            // dynamic smem base less aligned than budgeted
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // CTAs are dispatched in linear block order: walk the query tiles from the last (most keys) to the first,
  // all heads of one tile before the next tile — longest-processing-time-first over the whole grid, and the
  // heads of one GQA group (same K/V tiles) run next to each other
  const int lin = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
  const int n_hb = gridDim.y * gridDim.z;
  const int q0 = ((int)gridDim.x - 1 - lin / n_hb) * FA_BM;
  const int h = (lin % n_hb) % (int)gridDim.y, b = (lin % n_hb) / (int)gridDim.y;
  const int kvh = h / (p.nh / p.nkv);

  if (warp == 0 && lane == 0) {
    tc::prefetch_tensormap(&tmap_q); tc::prefetch_tensormap(&tmap_k); tc::prefetch_tensormap(&tmap_v);
  }
  if (warp == 1 && lane == 0) {
    tc::mbar_init(q_full, 1); tc::mbar_init(p_full, 4); tc::mbar_init(o_full, 1);
    for (int s = 0; s < F2_SLOTS; ++s) { tc::mbar_init(slot_full(s), 1); tc::mbar_init(slot_empty(s), 1); }
    for (int s = 0; s < 2; ++s) tc::mbar_init(s_full(s), 1);
    tc::fence_barrier_init();
  }
  if (warp == 0) tc::tmem_alloc(tmem_slot, 256);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *tmem_slot_ptr;
  auto tS = [&](int j) { return tmem + (uint32_t)(j & 1) * 64u; };     // score tile j (also holds P_j afterwards)
  const uint32_t tO = tmem + 128;

  pdl_wait();
  pdl_launch_dependents();

  const int pad = p.pad_left ? p.pad_left[b] : 0;
  const int q_hi = min(q0 + FA_BM, p.S) - 1;
  const int first_tile = pad / F2_BN;
  const int last_tile = (p.start + q_hi) / F2_BN;
  const int n_tiles = last_tile - first_tile + 1 > 0 ? last_tile - first_tile + 1 : 0;

  if (warp == 0) {
    if (lane == 0) {
      tc::mbar_expect_tx(q_full, FA_TILE_BYTES);
      tc::tma_load_2d(sQ, &tmap_q, q_full, h * FA_HD, b * p.S + q0, tc::EVICT_FIRST);
      tc::tma_load_2d(sQ + FA_HALF_BYTES, &tmap_q, q_full, h * FA_HD + 64, b * p.S + q0, tc::EVICT_FIRST);
      for (int idx = 0; idx < 2 * n_tiles; ++idx) {          // K_0, V_0, K_1, V_1, ...
        const int sl = idx % F2_SLOTS, use = idx / F2_SLOTS;
        tc::mbar_wait(slot_empty(sl), (use & 1) ^ 1);
        const int key0 = (first_tile + (idx >> 1)) * F2_BN;
        const CUtensorMap* m = (idx & 1) ? &tmap_v : &tmap_k;
        tc::mbar_expect_tx(slot_full(sl), F2_SLOT_BYTES);
        tc::tma_load_3d(slot_addr(sl), m, slot_full(sl), 0, key0, b * p.nkv + kvh, tc::EVICT_LAST);
        tc::tma_load_3d(slot_addr(sl) + F2_KV_HALF, m, slot_full(sl), 64, key0, b * p.nkv + kvh, tc::EVICT_LAST);
      }
    }
  } else if (warp == 1) {
    // MMA issue.  The whole warp walks the loop (warp-uniform control flow keeps descriptors and addresses
    // in uniform registers) and one elected lane issues; descriptors are built once and advanced by adding
    // the 16-byte-unit offset to their address field, so an MMA costs a couple of instructions to issue —
    // with 64-key tiles an MMA only runs 32-64 cycles, and a slow issue loop starves the tensor pipe.
    constexpr uint32_t IDESC_S = tc::idesc_bf16_f32(FA_BM, F2_BN, false, false);
    constexpr uint32_t IDESC_O = tc::idesc_bf16_f32(FA_BM, FA_HD, false, true);    // B = V is MN-major
    const uint64_t dQ = tc::smem_desc_kmajor_sw128(sQ);
    const uint64_t dP = tc::smem_desc_kmajor_sw128(sP);
    const uint64_t dK0 = tc::smem_desc_kmajor_sw128(slot_addr(0));
    const uint64_t dV0 = tc::smem_desc_mnmajor_sw128(slot_addr(0), F2_KV_HALF, 1024);
    constexpr uint64_t SLOT_STEP = F2_SLOT_BYTES >> 4;
    auto issue_s = [&](int j) {
      const int idx = 2 * j, sl = idx % F2_SLOTS, use = idx / F2_SLOTS;
      tc::mbar_wait(slot_full(sl), use & 1);
      tc::fence_after_sync();
      if (tc::elect_one()) {
        const uint64_t dK = dK0 + (uint64_t)sl * SLOT_STEP;
        const uint32_t d = tS(j);
#pragma unroll
        for (int k = 0; k < FA_HD / 16; ++k)
          tc::mma_ss(d, dQ + (uint64_t)(((k >> 2) * FA_HALF_BYTES + (k & 3) * 32) >> 4),
                     dK + (uint64_t)(((k >> 2) * F2_KV_HALF + (k & 3) * 32) >> 4), IDESC_S, k != 0);
        tc::mma_commit(s_full(j & 1));
        tc::mma_commit(slot_empty(sl));
      }
      __syncwarp();
    };
    tc::mbar_wait(q_full, 0);
    if (n_tiles > 0) issue_s(0);
    for (int j = 0; j < n_tiles; ++j) {
      if (j + 1 < n_tiles) issue_s(j + 1);             // tensor core computes S_{j+1} during softmax_j
      const int idx = 2 * j + 1, sl = idx % F2_SLOTS, use = idx / F2_SLOTS;
      tc::mbar_wait(slot_full(sl), use & 1);
      tc::mbar_wait(p_full, j & 1);
      tc::fence_after_sync();
      if (tc::elect_one()) {
        const uint64_t dV = dV0 + (uint64_t)sl * SLOT_STEP;
#pragma unroll
        for (int k = 0; k < F2_BN / 16; ++k) {
          if (P_TMEM)       // A = P_j in TMEM: 16 keys = 8 packed-bf16x2 columns per K step
            tc::mma_ts(tO, tS(j) + k * 8, dV + (uint64_t)((k * 2048) >> 4), IDESC_O, j != 0 || k != 0);
          else
            tc::mma_ss(tO, dP + (uint64_t)((k * 32) >> 4), dV + (uint64_t)((k * 2048) >> 4), IDESC_O,
                       j != 0 || k != 0);
        }
        tc::mma_commit(o_full);
        tc::mma_commit(slot_empty(sl));
      }
      __syncwarp();
    }
  } else {
    const int quarter = warp & 3;                          // TMEM lane quadrant this warp may touch
    const int r = quarter * 32 + lane;                     // query row inside the tile
    const int q_slot = p.start + q0 + r;                   // cache slot (= causal limit) of this query
    const uint32_t lane_off = (uint32_t)(quarter * 32) << 16;
    float m_ref = -INFINITY, l_run = 0.f;
    const uint32_t p_row = sP + r * 128;

    for (int j = 0; j < n_tiles; ++j) {
      const int key0 = (first_tile + j) * F2_BN;
      const bool need_mask = (key0 + F2_BN - 1 > p.start + q0 + quarter * 32) || (key0 < pad);   // warp-uniform
      tc::mbar_wait(s_full(j & 1), (j >> 1) & 1);
      tc::fence_after_sync();
      uint32_t v0[32], v1[32];
      tc::tmem_ld32(tS(j) + lane_off, v0);
      tc::tmem_ld32(tS(j) + lane_off + 32, v1);
      tc::tmem_ld_wait();
      if (need_mask) {
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          if (key0 + i > q_slot || key0 + i < pad) v0[i] = 0xff800000u;             // -inf
          if (key0 + 32 + i > q_slot || key0 + 32 + i < pad) v1[i] = 0xff800000u;
        }
      }
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int i = 0; i < 32; ++i)
        mx4[i & 3] = fmaxf(mx4[i & 3], fmaxf(__uint_as_float(v0[i]), __uint_as_float(v1[i])));
      const float t = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])) * p.scale_log2;
      // P·V_{j-1} must be complete before the P tile in shared memory is overwritten
      if (!P_TMEM && j > 0) {
        tc::mbar_wait(o_full, (j - 1) & 1);
        tc::fence_after_sync();
      }
      const bool need = t > m_ref + F2_RESCALE_THRESHOLD || (m_ref == -INFINITY && t > -INFINITY);
      if (j == 0) {
        m_ref = t;
      } else if (__any_sync(0xffffffffu, need)) {
        const float factor = need ? exp2f(m_ref - t) : 1.0f;      // m_ref = -inf → 0 (O and l are 0 there)
        if (P_TMEM) {                                             // O must be at rest: P·V_{j-1} complete
          tc::mbar_wait(o_full, (j - 1) & 1);
          tc::fence_after_sync();
        }
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
          uint32_t o[32];
          tc::tmem_ld32(tO + lane_off + c * 32, o);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * factor);
          tc::tmem_st32(tO + lane_off + c * 32, o);
        }
        tc::tmem_st_wait();
        l_run *= factor;
        if (need) m_ref = t;
      }
      const float m_safe = m_ref == -INFINITY ? 0.f : m_ref;
      float psum = 0.f;
      if (P_TMEM) {
        uint32_t w[32];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float a0 = exp2f(fmaf(__uint_as_float(v0[2 * i]), p.scale_log2, -m_safe));
          const float a1 = exp2f(fmaf(__uint_as_float(v0[2 * i + 1]), p.scale_log2, -m_safe));
          const float b0 = exp2f(fmaf(__uint_as_float(v1[2 * i]), p.scale_log2, -m_safe));
          const float b1 = exp2f(fmaf(__uint_as_float(v1[2 * i + 1]), p.scale_log2, -m_safe));
          psum += (a0 + a1) + (b0 + b1);
          w[i] = pack_bf16(a0, a1);
          w[16 + i] = pack_bf16(b0, b1);
        }
        tc::tmem_st32(tS(j) + lane_off, w);          // keys 2c, 2c+1 of this row → 32-bit column c
        tc::tmem_st_wait();
      } else {
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t w[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float s0 = __uint_as_float(c == 0 ? v0[2 * i] : v1[2 * i]);
            const float s1 = __uint_as_float(c == 0 ? v0[2 * i + 1] : v1[2 * i + 1]);
            const float p0 = exp2f(fmaf(s0, p.scale_log2, -m_safe)), p1 = exp2f(fmaf(s1, p.scale_log2, -m_safe));
            psum += p0 + p1;
            w[i] = pack_bf16(p0, p1);
          }
          // 32 keys = 64 B = four 16-byte chunks of this row: chunk index = c*4 + q, XOR-swizzled by the row
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const uint32_t chunk = (uint32_t)(c * 4 + q) ^ (uint32_t)(r & 7);
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(p_row + chunk * 16), "r"(w[4 * q]),
                         "r"(w[4 * q + 1]), "r"(w[4 * q + 2]), "r"(w[4 * q + 3]) : "memory");
          }
        }
        fence_proxy_async();               // generic-proxy smem writes → visible to the tensor core
      }
      l_run += psum;
      // Phase discipline of p_full: the MMA warp may still be parked on the V-tile barrier when this warp is
      // done with tile j; without a guard all four softmax warps could complete phase j AND phase j+1 before
      // the MMA warp looks at phase j, and a parity wait cannot tell "two phases later" from "not yet"
      // (observed once as a hang at batch 8).  P·V_{j-1} having completed proves the MMA warp consumed phase
      // j-1, so at most one phase is ever outstanding; the wait is almost always already satisfied.
      if (P_TMEM && j > 0) tc::mbar_wait(o_full, (j - 1) & 1);
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(p_full);
    }
    // ---- epilogue: O (TMEM) / l → bf16 row (256 contiguous bytes)
    if (n_tiles > 0) {
      tc::mbar_wait(o_full, (n_tiles - 1) & 1);
      tc::fence_after_sync();
    }
    const float inv = l_run > 0.f ? 1.0f / l_run : 0.f;
    uint4* dst = reinterpret_cast<uint4*>(p.out + ((size_t)(b * p.S + q0 + r) * p.nh + h) * FA_HD);
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      uint32_t o[32];
      if (n_tiles > 0) {
        tc::tmem_ld32(tO + lane_off + c * 32, o);
        tc::tmem_ld_wait();
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] = 0u;
      }
      if (q0 + r < p.S) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
          dst[c * 4 + i] = make_uint4(
              pack_bf16(__uint_as_float(o[8 * i]) * inv, __uint_as_float(o[8 * i + 1]) * inv),
              pack_bf16(__uint_as_float(o[8 * i + 2]) * inv, __uint_as_float(o[8 * i + 3]) * inv),
              pack_bf16(__uint_as_float(o[8 * i + 4]) * inv, __uint_as_float(o[8 * i + 5]) * inv),
              pack_bf16(__uint_as_float(o[8 * i + 6]) * inv, __uint_as_float(o[8 * i + 7]) * inv));
      }
    }
    tc::fence_before_sync();
  }

  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 256);
}

}  // namespace jl

using namespace jl;

// resident CTAs per SM of the v2 kernel at its launch configuration (2 is what the design assumes);
// smem_override >= 0 asks the same question for another dynamic shared-memory size (diagnostic)
extern "C" JL_API int jl_flash_prefill_occupancy(int smem_override) {
  int n = 0;
  const size_t smem = smem_override >= 0 ? (size_t)smem_override : (size_t)f2_smem_bytes(true);
  if (cudaFuncSetAttribute(flash_prefill_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)f2_smem_bytes(true)) != cudaSuccess) return -1;
  if (cudaFuncSetAttribute(flash_prefill_kernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared) != cudaSuccess) return -2;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, flash_prefill_kernel<true>, F2_THREADS, smem) != cudaSuccess) return -3;
  if (getenv("JLLAMA_DEBUG")) {
    cudaFuncAttributes fa{};
    cudaFuncGetAttributes(&fa, flash_prefill_kernel<true>);
    int n128 = 0, n256 = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n128, flash_prefill_kernel<true>, 128, smem);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n256, flash_prefill_kernel<true>, 256, smem);
    fprintf(stderr, "flash occupancy: regs %d static smem %zu maxThreads %d maxDyn %d carveout %d local %zu | blocks/SM at 128/192/256 threads: %d %d %d (dyn smem %zu)\n",
            fa.numRegs, fa.sharedSizeBytes, fa.maxThreadsPerBlock, fa.maxDynamicSharedSizeBytes,
            fa.preferredShmemCarveout, fa.localSizeBytes, n128, n, n256, smem);
  }
  return n;
}

extern "C" JL_API int jl_flash_prefill(const void* q, const void* k_cache, const void* v_cache,
                                       const int* cache_len_ptr, const int* pad_left, void* out, int B, int S,
                                       int nh, int nkv, int hd, int T, int start, float scale, void* stream) {
  (void)cache_len_ptr;
  // 64-wide heads (1B-class LLaMA-architecture models) go to the warp-level tensor-core kernel (prefill_attn_mma.cu);
  // JLLAMA_PREFILL_MMA=1 sends head_dim 128 there too (A/B against the tcgen05 kernel below)
  static const bool force_mma = [] { const char* e = getenv("JLLAMA_PREFILL_MMA"); return e && e[0] == '1'; }();
  if (hd == 64 || (force_mma && hd == 128))
    return jl_prefill_attn_mma(q, k_cache, v_cache, pad_left, out, B, S, nh, nkv, hd, T, start, scale, stream);
  JL_CHECK_ARG(hd == FA_HD, "jl_flash_prefill: head_dim=%d (64 and 128 are built)", hd);
  JL_CHECK_ARG(nkv > 0 && nh % nkv == 0, "jl_flash_prefill: nh=%d not a multiple of nkv=%d", nh, nkv);
  JL_CHECK_ARG(start >= 0 && start + S <= T, "jl_flash_prefill: start=%d + S=%d exceeds cache capacity %d", start, S, T);
  if (B * S == 0) return JL_OK;
  static const bool use_v1 = [] { const char* e = getenv("JLLAMA_FLASH_V1"); return e && e[0] == '1'; }();
  const int bn = use_v1 ? FA_BN : F2_BN;
  CUtensorMap tq, tk, tv;
  int rc = make_tmap_2d_bf16(&tq, q, (uint64_t)B * S, (uint64_t)nh * hd, (uint64_t)nh * hd, FA_BM);
  if (rc) return rc;
  rc = make_tmap_3d_bf16(&tk, k_cache, (uint64_t)B * nkv, T, hd, (uint64_t)T * hd, hd, bn);
  if (rc) return rc;
  rc = make_tmap_3d_bf16(&tv, v_cache, (uint64_t)B * nkv, T, hd, (uint64_t)T * hd, hd, bn);
  if (rc) return rc;
  FlashParams p{(__nv_bfloat16*)out, pad_left, B, S, nh, nkv, T, start, scale * 1.4426950408889634f};
  dim3 grid((S + FA_BM - 1) / FA_BM, nh, B);
  if (use_v1) {
    const size_t smem = 6 * FA_TILE_BYTES + 128 + 1024;
    JL_CUDA(cudaFuncSetAttribute(flash_prefill_v1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    JL_CUDA(launch_pdl(flash_prefill_v1_kernel, grid, dim3(FA_THREADS), smem, (cudaStream_t)stream, tq, tk, tv, p));
    return JL_OK;
  }
  static const bool p_smem = [] { const char* e = getenv("JLLAMA_FLASH_PSMEM"); return e && e[0] == '1'; }();
  if (p_smem) {
    JL_CUDA(cudaFuncSetAttribute(flash_prefill_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)f2_smem_bytes(false)));
    JL_CUDA(launch_pdl(flash_prefill_kernel<false>, grid, dim3(F2_THREADS), f2_smem_bytes(false), (cudaStream_t)stream, tq, tk, tv, p));
    return JL_OK;
  }
  JL_CUDA(cudaFuncSetAttribute(flash_prefill_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)f2_smem_bytes(true)));
  // two CTAs per SM need the full shared-memory carve-out (the default heuristic sizes it for one)
  JL_CUDA(cudaFuncSetAttribute(flash_prefill_kernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  JL_CUDA(launch_pdl(flash_prefill_kernel<true>, grid, dim3(F2_THREADS), f2_smem_bytes(true), (cudaStream_t)stream, tq, tk, tv, p));
  return JL_OK;
}
