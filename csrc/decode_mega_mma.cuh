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
// memory (written once per op by the staging functions) in which the 16 words a thread needs for one stage are
// contiguous: 4 LDS.128 per stage.  Partial sums of the 8 warps meet in shared memory once per 8-row group.
//
// Activations travel between CTAs as TAGGED PACKETS (see decode_mega.cu): the epilogues below write
// {bf16 pair, tag} 8-byte words and the staging functions poll them, so no grid-wide barrier separates the ops.
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

// ---- staging of a GEMV input into the permuted shared-memory layout.
// The two staging functions are deliberately NOT inlined: the layer body calls them six times, and the decode
// kernel must stay small — measured on B200: at 129 KB / 178 KB of code per instance the step ran 2.84 / 2.95 ms,
// at 238 KB and 280 KB 5.0 ms (every GEMV phase twice as slow: the instruction stream no longer stays cached).
//
// RMSNorm is split around the GEMV: staging writes x ⊙ g (the norm weights, fetched before the poll) and leaves the
// per-warp sums of squares of x in `red`; the GEMV epilogue multiplies its outputs by rsqrt(mean x² + eps).  The
// scale is a per-row scalar, so W·(x ⊙ g)·r = W·(x·r ⊙ g); this takes a barrier, a second pass over shared memory
// and the rsqrt out of the critical path between two GEMVs.

// end of a staging pass: per-warp sums of squares → red, then the barrier that publishes the staged activations
// (and red) to all consumer warps
template <int M>
__device__ __forceinline__ void mg_stage_finish(bool norm, float (*red)[MG_CONSUMERS], float (&ss)[M]) {
  const int lane = threadIdx.x & 31, cw = (threadIdx.x >> 5) - MG_PWARPS;
  if (norm) {
#pragma unroll
    for (int m = 0; m < M; ++m) {
      const float t = warp_sum(ss[m]);
      if (lane == 0) red[m][cw] = t;
    }
  }
  mg_consumer_bar();
}

// stage the embedding rows of this step's tokens (first layer only; optional RMSNorm); positions k >= K are zero
template <int M>
__device__ __noinline__ void mg_stage_embedding(uint4* xin, const __nv_bfloat16* table, const int* ids, int vocab,
                                                int K, const __nv_bfloat16* norm_w, float (*red)[MG_CONSUMERS]) {
  const int ct = threadIdx.x - MG_CT0;
  const int K8 = K >> 3, n_ks = (K + MG_KS - 1) / MG_KS;
  uint32_t* xw = reinterpret_cast<uint32_t*>(xin);
  float ss[M];
#pragma unroll
  for (int m = 0; m < M; ++m) ss[m] = 0.f;
  for (int i = ct; i < n_ks * (MG_KS / 8); i += MG_CT) {
    const uint4 g = (norm_w != nullptr && i < K8) ? __ldg(reinterpret_cast<const uint4*>(norm_w) + i)
                                                  : make_uint4(0x3f803f80u, 0x3f803f80u, 0x3f803f80u, 0x3f803f80u);
    const uint32_t gw[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
    for (int m = 0; m < M; ++m) {
      int id = ids[m];
      id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
      const uint4 a = i < K8 ? __ldg(reinterpret_cast<const uint4*>(table) + (size_t)id * K8 + i) : make_uint4(0, 0, 0, 0);
      const uint32_t aw[4] = {a.x, a.y, a.z, a.w};
      const int base = mg_xperm_word(m, n_ks, i * 8);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float lo = bf16_lo(aw[j]), hi = bf16_hi(aw[j]);
        xw[base + j * 16] = pack_bf16(lo * bf16_lo(gw[j]), hi * bf16_hi(gw[j]));      // pair j of the chunk belongs to quad lane t = j
        ss[m] += lo * lo + hi * hi;
      }
    }
  }
  mg_stage_finish<M>(norm_w != nullptr, red, ss);
}

// Stage [M, K] activations that arrive as tagged packets {bf16 pair, tag} (row pitch `ldp` packets): every thread
// polls the packets it stages until their tag is `tag` — this poll IS the synchronisation with the producers of
// the previous op (all CTAs of this GPU), there is no grid barrier.  Up to 8 16-byte loads (16 packets) are in
// flight per thread (one L2 round trip per batch once the data is there); the norm weights of a batch are requested
// before its poll.  Optional RMSNorm (see above); positions k >= K are zero.
template <int M>
__device__ __noinline__ void mg_stage_packets(uint4* xin, const uint2* pk, uint32_t tag, int K, int ldp,
                                              const __nv_bfloat16* norm_w, float (*red)[MG_CONSUMERS], int what,
                                              unsigned long long* trace_row) {
  constexpr int NB = 8;                                    // 16-byte loads in flight per thread
  const int ct = threadIdx.x - MG_CT0;
  const int n_ks = (K + MG_KS - 1) / MG_KS, K4 = K >> 2;
  uint32_t* xw = reinterpret_cast<uint32_t*>(xin);
  float ss[M];
#pragma unroll
  for (int m = 0; m < M; ++m) ss[m] = 0.f;
  // item q = it * M + m covers values k = 4 i .. 4 i + 3 of row m, i = ct + it * MG_CT (exactly n_ks items per row)
  const int items = n_ks * M;
#pragma unroll 1
  for (int q0 = 0; q0 < items; q0 += NB) {
    uint4 v[NB];
    uint2 g[NB];
#pragma unroll
    for (int u = 0; u < NB; ++u) {
      const int q = q0 + u, it = q / M, i = ct + it * MG_CT;
      g[u] = (norm_w != nullptr && q < items && i < K4) ? __ldg(reinterpret_cast<const uint2*>(norm_w) + i)
                                                        : make_uint2(0x3f803f80u, 0x3f803f80u);      // bf16 1.0 pairs
    }
    unsigned spins = 0;
    for (;;) {
      bool ok = true;
#pragma unroll
      for (int u = 0; u < NB; ++u) {
        const int q = q0 + u, it = q / M, m = q - it * M, i = ct + it * MG_CT;
        v[u] = (q < items && i < K4) ? ld_relaxed_gpu_v4(reinterpret_cast<const uint4*>(pk + (size_t)m * ldp) + i)
                                     : make_uint4(0, tag, 0, tag);
      }
#pragma unroll
      for (int u = 0; u < NB; ++u) ok = ok && v[u].y == tag && v[u].w == tag;
      if (ok) break;
      if (++spins > (1u << 22)) mg_poll_timeout(what, tag, v[0].y);
    }
#pragma unroll
    for (int u = 0; u < NB; ++u) {
      const int q = q0 + u, it = q / M, m = q - it * M, i = ct + it * MG_CT;
      if (q < items) {
        const float a = bf16_lo(v[u].x), b = bf16_hi(v[u].x), d = bf16_lo(v[u].z), e = bf16_hi(v[u].z);
        xw[mg_xperm_word(m, n_ks, i * 4)] = pack_bf16(a * bf16_lo(g[u].x), b * bf16_hi(g[u].x));
        xw[mg_xperm_word(m, n_ks, i * 4 + 2)] = pack_bf16(d * bf16_lo(g[u].y), e * bf16_hi(g[u].y));
        ss[m] += a * a + b * b + d * d + e * e;
      }
    }
  }
  if (trace_row != nullptr && threadIdx.x == MG_CT0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    trace_row[21 + what] = t;
  }
  mg_stage_finish<M>(norm_w != nullptr, red, ss);
}

// What the epilogue of a GEMV does with a finished 8-row group (see decode_mega.cu for the packet buffers)
struct MegaOut {
  uint2* pk;                // MG_PLAIN / MG_SWIGLU: packet buffer [M][N/2] ({bf16 pair, tag})
  uint2* r_pk;              // MG_RESIDUAL: packets of the new residual stream [M][H/2]
  float* logits;            // MG_LOGITS: optional fp32 logits (may be null: candidates only): row pitch logits_ld, this
  int logits_ld, logits_col0;   // rank's shard starts at column logits_col0; logits_mc: the pointer is an NVLS multicast
  bool logits_mc;               // address (the store lands in every rank's [M, V] buffer)
  float (*res)[64];         // MG_RESIDUAL: this CTA's rows of the residual stream (shared memory, bf16-exact values)
  float* best_v;            // MG_LOGITS: running (max, argmax) of this thread's rows
  int* best_i;
  const float (*red)[MG_CONSUMERS];   // non-null: the input was staged with RMSNorm, scale outputs by rsqrt(mean x² + eps)
  float inv_k, eps;         // 1 / K of the normalised input
  uint32_t tag;
  int row0;                 // first row of this CTA's range (index base of `res`)
  int idx_offset;           // MG_LOGITS: global vocabulary index of local row 0
};

// CARRY (two-pass GEMV over K, used when M rows of the full input do not fit in shared memory): 1 = this call
// covers the first K range and parks its sums in `carry` ([M][64] floats, this CTA's rows); 2 = this call covers
// the rest, adds the parked sums and runs the epilogue.  0 = ordinary single pass.
template <int M, int MODE, int CARRY = 0>
__device__ __forceinline__ void mg_gemv_mma(MegaCtx& c, int N, int K, const MegaOut& o,
                                            float (*s_acc)[MG_CONSUMERS][MG_MMA_MAX_M][8], float (*carry)[64] = nullptr) {
  static_assert(M <= MG_MMA_MAX_M, "batch rows ride in the first 8 rows of the MMA M dimension");
  const int lane = threadIdx.x & 31, cw = (threadIdx.x >> 5) - MG_PWARPS, ct = threadIdx.x - MG_CT0;
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
  // epilogue threads (see below) scale their outputs by the RMSNorm factor of their batch row
  float rs = 1.f;
  if (o.red != nullptr) {
    const int em = MODE == MG_SWIGLU ? (ct >> 2) : (ct >> 3);
    if (em < M) {
      float t = 0.f;
#pragma unroll
      for (int w = 0; w < MG_CONSUMERS; ++w) t += o.red[em][w];
      rs = rsqrtf(t * o.inv_k + o.eps);
    }
  }
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
      if (ct < 32) {                                     // one thread per (batch row, (gate, up) pair); whole warp for the shuffle
        const int m = ct >> 2, pp = ct & 3;
        float h = 0.f;
        if (m < M) {
          float gate = 0.f, up = 0.f;
#pragma unroll
          for (int w = 0; w < MG_CONSUMERS; ++w) { gate += buf[w][m][2 * pp]; up += buf[w][m][2 * pp + 1]; }
          gate = __bfloat162float(__float2bfloat16(gate * rs));
          up = __bfloat162float(__float2bfloat16(up * rs));
          h = silu(gate) * up;
        }
        const float h1 = __shfl_down_sync(0xffffffffu, h, 1);
        if (m < M && (pp & 1) == 0)
          st_packet(o.pk + (size_t)m * (N / 4) + ((grp * (MG_GROUP / 2) + pp) >> 1), pack_bf16(h, h1), o.tag);
      }
    } else {
      if (ct < 64) {                                     // thread (batch row m, weight row n); two whole warps
        const int m = ct >> 3, n = ct & 7;
        const bool valid = m < M;
        float v = 0.f;
        if (valid) {
#pragma unroll
          for (int w = 0; w < MG_CONSUMERS; ++w) v += buf[w][m][n];
          v *= rs;
        }
        const int row = grp * MG_GROUP + n;
        if constexpr (CARRY == 1) {
          if (valid) carry[m][(grp - g0) * MG_GROUP + n] = v;       // read back by this same thread in the second pass
        } else if constexpr (MODE == MG_LOGITS) {
          if (valid) {
            if (o.logits != nullptr) {
              float* dst = o.logits + (size_t)m * o.logits_ld + o.logits_col0 + row;
              if (o.logits_mc) mm::st_f32(dst, v); else *dst = v;
            }
            if (v > *o.best_v) { *o.best_v = v; *o.best_i = row + o.idx_offset; }      // rows ascend: ties keep the lower index
          }
        } else if constexpr (MODE == MG_RESIDUAL) {
          if constexpr (CARRY == 2) { if (valid) v += carry[m][(grp - g0) * MG_GROUP + n]; }
          if (c.ar->tp > 1) {
            if (valid) mg_push_partial<M>(c, m, row, N, v);          // reduced over ranks in mg_owner_reduce
          } else {
            float y = 0.f;
            if (valid) {
              y = __bfloat162float(__float2bfloat16(__bfloat162float(__float2bfloat16(v)) + o.res[m][row - o.row0]));
              o.res[m][row - o.row0] = y;
            }
            const float y1 = __shfl_down_sync(0xffffffffu, y, 1);
            if (valid && (n & 1) == 0) st_packet(o.r_pk + (size_t)m * (N / 2) + (row >> 1), pack_bf16(y, y1), o.tag);
          }
        } else {
          const float v1 = __shfl_down_sync(0xffffffffu, v, 1);
          if (valid && (n & 1) == 0) st_packet(o.pk + (size_t)m * (N / 2) + (row >> 1), pack_bf16(v, v1), o.tag);
        }
      }
    }
  }
}

}  // namespace jl
