// Tensor-core GEMV phase of the persistent decode kernel (included by decode_mega.cu after MegaCtx / mg_range).
//
// Why: ncu on the SIMT GEMV showed the SM issuing ~1 instruction per cycle with every warp stalled on
// dependent-latency (9 warps per SM cannot hide it): a 16 KB weight stage cost ~1100 warp-instructions
// (LDS + 2 unpack + 1 FFMA per weight pair) ≈ 900-1000 cycles, so `w2` ran at 35 GB/s per SM regardless of how
// many SMs were streaming — instruction-bound, not HBM-bound.  Here the same stage costs ~35 instructions per
// warp: the weights go through `ldmatrix` + `mma.sync.m16n8k16` (bf16 in, fp32 accumulate) with the WEIGHT
// rows as the 8-wide N operand and the batch rows (M <= 8) as the 16-wide M operand, so unpacking and the
// multiply-adds are done by the tensor core and batch rows beyond the first are free.
// (tcgen05 is the wrong tool here: its minimum tile is 128 x N x 16 with operands described by shared-memory
// descriptors and accumulators in TMEM — a GEMV has one useful row, and the per-stage commit / TMEM read-back
// round trips would cost more than the whole stage does with the warp-level MMA.)
//
// Stage layout (TMA, SWIZZLE_128B, box {64 k, 8 rows, 16 k-blocks}): [k-block 0..15][row 0..7][128 B], the
// canonical K-major swizzled tile, so each ldmatrix 8x8 fetch (8 rows x 16 B) is bank-conflict free.
// Consumer warp w owns k in [128 w, 128 w + 128) of every stage: 8 k16-steps = 4 ldmatrix.x4 + 8 MMAs, two
// independent accumulator chains.  The activation fragments are read from a PERMUTED copy of x in shared
// memory (written once per op by mg_stage_input_mma) in which the 16 words a thread needs for one stage are
// contiguous: 4 LDS.128 per stage.  Partial sums of the 8 warps meet in shared memory once per 8-row group.
#pragma once

namespace jl {

constexpr int MG_MMA_MAX_M = 8;

__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
// D(16x8, fp32) += A(16x16, bf16: batch rows x k) · B(16x8, bf16: k x weight rows); rows 8..15 of A are zero
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], uint32_t a0, uint32_t a2, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(0u), "r"(a2), "r"(0u), "r"(b0), "r"(b1));
}

// 32-bit word index of the activation pair (k, k+1), k even, of batch row m in the permuted staging buffer:
// [m][k-stage][warp w][quad lane t][16 words: k16-step s, half]
__device__ __forceinline__ int mg_xperm_word(int m, int n_ks, int k) {
  const int ks = k >> 10, w = (k >> 7) & 7, s = (k >> 4) & 7, half = (k >> 3) & 1, t = (k >> 1) & 3;
  return ((((m * n_ks + ks) * 8 + w) * 4 + t) << 4) + s * 2 + half;
}

// stage [M, K] activations (optional RMSNorm) into the permuted layout; positions k >= K are zero
template <int M>
__device__ __forceinline__ void mg_stage_input_mma(MegaCtx& c, const __nv_bfloat16* src, int K,
                                                   const __nv_bfloat16* norm_w, float eps, float (*red)[MG_CONSUMERS],
                                                   int ld = 0) {
  // `ld` = row pitch of src in elements when only a K-range [src, src + K) of wider rows is staged (two-pass w2)
  const int ct = threadIdx.x - 32, lane = threadIdx.x & 31, cw = (threadIdx.x >> 5) - 1;
  const int K8 = K >> 3, n_ks = (K + MG_KS - 1) / MG_KS;
  const int LD8 = ld > 0 ? ld >> 3 : K8;
  uint32_t* xw = reinterpret_cast<uint32_t*>(c.xin);
  float ss[M];
#pragma unroll
  for (int m = 0; m < M; ++m) ss[m] = 0.f;
  for (int i = ct; i < n_ks * (MG_KS / 8); i += MG_CT) {
#pragma unroll
    for (int m = 0; m < M; ++m) {
      const uint4 a = i < K8 ? ldcg16(src + ((size_t)m * LD8 + i) * 8) : make_uint4(0, 0, 0, 0);
      const uint32_t aw[4] = {a.x, a.y, a.z, a.w};
      const int base = mg_xperm_word(m, n_ks, i * 8);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        xw[base + j * 16] = aw[j];                       // pair j of the chunk belongs to quad lane t = j
        const float lo = bf16_lo(aw[j]), hi = bf16_hi(aw[j]);
        ss[m] += lo * lo + hi * hi;
      }
    }
  }
  if (norm_w != nullptr) {
#pragma unroll
    for (int m = 0; m < M; ++m) {
      const float t = warp_sum(ss[m]);
      if (lane == 0) red[m][cw] = t;
    }
    mg_consumer_bar();
    float rinv[M];
#pragma unroll
    for (int m = 0; m < M; ++m) {
      float t = lane < MG_CONSUMERS ? red[m][lane] : 0.f;
      t = warp_sum(t);
      rinv[m] = rsqrtf(t / (float)K + eps);
    }
    for (int i = ct; i < K8; i += MG_CT) {
      const uint4 g = __ldg(reinterpret_cast<const uint4*>(norm_w) + i);
      const uint32_t gw[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
      for (int m = 0; m < M; ++m) {
        const int base = mg_xperm_word(m, n_ks, i * 8);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t a = xw[base + j * 16];          // written by this same thread above
          xw[base + j * 16] = pack_bf16(bf16_lo(a) * rinv[m] * bf16_lo(gw[j]), bf16_hi(a) * rinv[m] * bf16_hi(gw[j]));
        }
      }
    }
  }
  mg_consumer_bar();
}

// CARRY (two-pass GEMV over K, used when M rows of the full input do not fit in shared memory): 1 = this call
// covers the first K range and parks its sums in `carry` ([M][64] floats, this CTA's rows); 2 = this call covers
// the rest, adds the parked sums and runs the epilogue.  0 = ordinary single pass.
template <int M, int MODE, int CARRY = 0>
__device__ __forceinline__ void mg_gemv_mma(MegaCtx& c, int N, int K, void* out, const __nv_bfloat16* residual,
                                            float (*s_acc)[MG_CONSUMERS][MG_MMA_MAX_M][8], float (*carry)[64] = nullptr) {
  static_assert(M <= MG_MMA_MAX_M, "batch rows ride in the first 8 rows of the MMA M dimension");
  const int lane = threadIdx.x & 31, cw = (threadIdx.x >> 5) - 1, ct = threadIdx.x - 32;
  const int g = lane >> 2, t = lane & 3;                   // MMA fragment coordinates: batch row g, quad lane t
  const int n_ks = (K + MG_KS - 1) / MG_KS;
  int g0, g1;
  mg_range(N / MG_GROUP, g0, g1);
  // ldmatrix row addresses of this lane, relative to the stage base: matrix q = lane / 8 covers 16-byte chunk
  // c0 + q of k-block kb, row r = lane % 8; pair pr = (k16-steps 2pr, 2pr+1) → kb = 2 cw + pr/2, c0 = 4 (pr & 1)
  uint32_t boff[4];
  {
    const int q = lane >> 3, r = lane & 7;
#pragma unroll
    for (int pr = 0; pr < 4; ++pr) {
      const int kb = cw * 2 + (pr >> 1), chunk = ((pr & 1) * 4 + q) ^ r;
      boff[pr] = (uint32_t)((kb * 8 + r) * 128 + chunk * 16);
    }
  }
  const bool has_row = g < M;
  const uint4* xq = c.xin + ((size_t)((g * n_ks) * 8 + cw) * 4 + t) * 4;      // + ks * 128 uint4 per k-stage

  for (int grp = g0; grp < g1; ++grp) {
    float d0[4] = {0.f, 0.f, 0.f, 0.f}, d1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 1
    for (int ks = 0; ks < n_ks; ++ks) {
      uint4 xa[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) xa[j] = has_row ? xq[ks * 128 + j] : make_uint4(0, 0, 0, 0);
      tc::mbar_wait(c.bars + 8u * c.stage, c.phase);
      const uint32_t sbase = c.ring + c.stage * MG_STAGE_BYTES;
#pragma unroll
      for (int pr = 0; pr < 4; ++pr) {
        uint32_t b[4];
        ldmatrix_x4(sbase + boff[pr], b);
        // xa[pr] = {a0, a2 of step 2pr; a0, a2 of step 2pr + 1}
        mma_bf16_16816(d0, xa[pr].x, xa[pr].y, b[0], b[1]);
        mma_bf16_16816(d1, xa[pr].z, xa[pr].w, b[2], b[3]);
      }
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(c.bars + 8u * (c.stages + c.stage));
      if (++c.stage == c.stages) { c.stage = 0; c.phase ^= 1; }
    }
    // ---- the 8 k-slices of this group meet in shared memory; thread (g, t) holds outputs (row g, cols 2t, 2t+1)
    float (*buf)[MG_MMA_MAX_M][8] = s_acc[grp & 1];
    if (has_row) *reinterpret_cast<float2*>(&buf[cw][g][2 * t]) = make_float2(d0[0] + d1[0], d0[1] + d1[1]);
    mg_consumer_bar();
    if constexpr (MODE == MG_SWIGLU) {
      if (ct < M * 4) {                                  // one thread per (batch row, (gate, up) pair)
        const int m = ct >> 2, pp = ct & 3;
        float gate = 0.f, up = 0.f;
#pragma unroll
        for (int w = 0; w < MG_CONSUMERS; ++w) { gate += buf[w][m][2 * pp]; up += buf[w][m][2 * pp + 1]; }
        gate = __bfloat162float(__float2bfloat16(gate));
        up = __bfloat162float(__float2bfloat16(up));
        reinterpret_cast<__nv_bfloat16*>(out)[(size_t)m * (N / 2) + grp * (MG_GROUP / 2) + pp] =
            __float2bfloat16(silu(gate) * up);
      }
    } else {
      if (ct < M * 8) {
        const int m = ct >> 3, n = ct & 7;
        float v = 0.f;
#pragma unroll
        for (int w = 0; w < MG_CONSUMERS; ++w) v += buf[w][m][n];
        const int row = grp * MG_GROUP + n;
        if constexpr (CARRY == 1) {
          carry[m][(grp - g0) * MG_GROUP + n] = v;         // read back by this same thread in the second pass
        } else if constexpr (MODE == MG_LOGITS) {
          reinterpret_cast<float*>(out)[(size_t)m * N + row] = v;
        } else {
          bool pushed = false;
          if constexpr (CARRY == 2) v += carry[m][(grp - g0) * MG_GROUP + n];
          if constexpr (MODE == MG_RESIDUAL) {
            if (c.ar->tp > 1) { mg_push_partial<M>(c, m, row, N, v); pushed = true; }     // reduced in mg_ar_finish
            else v = __bfloat162float(__float2bfloat16(v)) + __bfloat162float(__ldcg(residual + (size_t)m * N + row));
          }
          if (!pushed) reinterpret_cast<__nv_bfloat16*>(out)[(size_t)m * N + row] = __float2bfloat16(v);
        }
      }
    }
  }
}

}  // namespace jl
