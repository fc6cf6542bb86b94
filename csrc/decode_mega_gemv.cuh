// GEMV phase of the persistent decode kernel (included by decode_mega.cu after MegaCtx / mg_range).
//
// Thread mapping ("column owner"): a stage holds ROWS weight rows x 1024 columns as [row][slab][256].
// Consumer warp w owns column slab (w & 3) — lane = one 16-byte chunk of that slab — and the ROWS/2 rows of
// half (w >> 2).  A thread therefore meets the SAME activation chunk for every row group of the op, once per
// k-stage: those chunks (K/1024 of them) are kept in registers for the whole op, each is unpacked once per
// stage and reused for all of the thread's weight rows.  Shared memory carries only the TMA write and a
// single read of the weights (the first version also read the activations from smem every stage: three
// 16 KB streams per stage ≈ the 128 B/clk/SM limit).
// Partial sums of the four slab-warps of a half are combined through shared memory once per row group
// (named barrier over the 128 threads of the half, double-buffered by group parity).
// ROWS = 4 (8 KB stages) halves the work-unit size: with 4096 output rows over 148 CTAs, 8-row groups give
// some CTAs 4 groups and others 3 (the slow ones set the barrier time: 23 vs 18 µs for w2); 4-row groups
// give 7 vs 6.
#pragma once

namespace jl {

// NKS_MAX: compile-time bound on k-stages (K <= NKS_MAX·1024) so the register array is statically indexed.
// XREG = false falls back to reading the activations from shared memory (large M·K).
template <int M, int MODE, int NKS_MAX, bool XREG, int ROWS>
__device__ __forceinline__ void mg_gemv2(MegaCtx& c, int N, int K, void* out, const __nv_bfloat16* residual,
                                         float (*s_part)[MG_CONSUMERS][M][4]) {
  constexpr int RPT = ROWS / 2;                              // rows per thread
  const int lane = threadIdx.x & 31, cw = (threadIdx.x >> 5) - 1;
  const int half = cw >> 2, slab = cw & 3;
  const int K8 = K >> 3;
  int g0, g1;
  mg_range(N / ROWS, g0, g1);
  const int n_ks = (K + MG_KS - 1) / MG_KS;
  const int ch0 = slab * 32 + lane;                         // this thread's chunk inside a k-stage

  uint4 xr[XREG ? M : 1][XREG ? NKS_MAX : 1];
  if constexpr (XREG) {
#pragma unroll
    for (int ks = 0; ks < NKS_MAX; ++ks) {
      const int ch = ks * (MG_KS / 8) + ch0;
#pragma unroll
      for (int m = 0; m < M; ++m) xr[m][ks] = (ks < n_ks && ch < K8) ? c.xin[m * K8 + ch] : make_uint4(0, 0, 0, 0);
    }
  }

  for (int g = g0; g < g1; ++g) {
    float acc[M][RPT];
#pragma unroll
    for (int m = 0; m < M; ++m)
#pragma unroll
      for (int r = 0; r < RPT; ++r) acc[m][r] = 0.f;

    auto stage_body = [&](const int ks, const uint4 (&xv)[M]) {
      tc::mbar_wait(c.bars + 8u * c.stage, c.phase);
      const uint4* wr = reinterpret_cast<const uint4*>(c.ring_ptr + c.stage * MG_STAGE_BYTES) + (half * RPT) * 128 + ch0;
      if (ks * (MG_KS / 8) + ch0 < K8) {
        float xf[M][8];
#pragma unroll
        for (int m = 0; m < M; ++m) {
          xf[m][0] = bf16_lo(xv[m].x); xf[m][1] = bf16_hi(xv[m].x); xf[m][2] = bf16_lo(xv[m].y); xf[m][3] = bf16_hi(xv[m].y);
          xf[m][4] = bf16_lo(xv[m].z); xf[m][5] = bf16_hi(xv[m].z); xf[m][6] = bf16_lo(xv[m].w); xf[m][7] = bf16_hi(xv[m].w);
        }
        uint4 wv[RPT];
#pragma unroll
        for (int r = 0; r < RPT; ++r) wv[r] = wr[r * 128];
#pragma unroll
        for (int r = 0; r < RPT; ++r) {
          const float wf[8] = {bf16_lo(wv[r].x), bf16_hi(wv[r].x), bf16_lo(wv[r].y), bf16_hi(wv[r].y),
                               bf16_lo(wv[r].z), bf16_hi(wv[r].z), bf16_lo(wv[r].w), bf16_hi(wv[r].w)};
#pragma unroll
          for (int m = 0; m < M; ++m)
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[m][r] = fmaf(wf[e], xf[m][e], acc[m][r]);
        }
      }
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(c.bars + 8u * (c.stages + c.stage));
      if (++c.stage == c.stages) { c.stage = 0; c.phase ^= 1; }
    };

    if constexpr (XREG) {
#pragma unroll
      for (int ks = 0; ks < NKS_MAX; ++ks) {
        if (ks < n_ks) {
          uint4 xv[M];
#pragma unroll
          for (int m = 0; m < M; ++m) xv[m] = xr[m][ks];
          stage_body(ks, xv);
        }
      }
    } else {
      for (int ks = 0; ks < n_ks; ++ks) {
        uint4 xv[M];
        const int ch = ks * (MG_KS / 8) + ch0;
#pragma unroll
        for (int m = 0; m < M; ++m) xv[m] = ch < K8 ? c.xin[m * K8 + ch] : make_uint4(0, 0, 0, 0);
        stage_body(ks, xv);
      }
    }

    // ---- combine the four slab-warps of this half, then the epilogue for the half's rows
#pragma unroll
    for (int m = 0; m < M; ++m)
#pragma unroll
      for (int r = 0; r < RPT; ++r) acc[m][r] = warp_sum(acc[m][r]);
    float (*buf)[M][4] = s_part[g & 1];
    if (lane == 0) {
#pragma unroll
      for (int m = 0; m < M; ++m)
#pragma unroll
        for (int r = 0; r < RPT; ++r) buf[cw][m][r] = acc[m][r];
    }
    asm volatile("bar.sync %0, 128;" ::"r"(2 + half) : "memory");
    if (slab == 0 && lane < RPT * M) {
      const int m = lane / RPT, r = lane - m * RPT;
      const int wb = half * 4;
      const float v = buf[wb][m][r] + buf[wb + 1][m][r] + buf[wb + 2][m][r] + buf[wb + 3][m][r];
      const int row = g * ROWS + half * RPT + r;
      if constexpr (MODE == MG_SWIGLU) {
        // rows (2j, 2j+1) = (gate, up): the even lane finishes the pair
        const float vb = __bfloat162float(__float2bfloat16(v));
        const float up = __shfl_down_sync((1u << (RPT * M)) - 1u, vb, 1);    // exactly the finishing lanes
        if ((r & 1) == 0)
          reinterpret_cast<__nv_bfloat16*>(out)[(size_t)m * (N / 2) + row / 2] = __float2bfloat16(silu(vb) * up);
      } else if constexpr (MODE == MG_LOGITS) {
        reinterpret_cast<float*>(out)[(size_t)m * N + row] = v;
      } else {
        float y = v;
        bool pushed = false;
        if constexpr (MODE == MG_RESIDUAL) {
          if (c.ar->tp > 1) { mg_push_partial<M>(c, m, row, N, v); pushed = true; }     // reduced in mg_ar_finish
          else y = __bfloat162float(__float2bfloat16(v)) + __bfloat162float(__ldcg(residual + (size_t)m * N + row));
        }
        if (!pushed) reinterpret_cast<__nv_bfloat16*>(out)[(size_t)m * N + row] = __float2bfloat16(y);
      }
    }
  }
}

}  // namespace jl
