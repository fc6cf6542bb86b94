// Owner-side step of the sequence-parallel prefill and the all-reduce of a row-parallel GEMM, over NVLS.
// What XLA/GSPMD emitted here for the reference: an NCCL all-reduce of the whole [M, H] activation after wo / w2
// (/root/reference/jax_llama/partition.py:67,70), then replicated norms (:73-76).  (Round 1 scattered the GEMM's
// partial tiles into the row owners' buffers with per-thread remote stores and pushed the normalised rows with tp
// unicast stores; it lost to NCCL by 8-19 % and is gone.)
#include "common.cuh"
#include "jllama_b200.h"
#include "nvls.cuh"

namespace jl {

// ------------------------------------------------------------------------------------------------------------------
// NVLS form of the owner-side reduction (round 2).  The row-parallel GEMM no longer scatters anything: it stores its
// bf16 partial product [M, H] into ITS OWN copy of a symmetric buffer with ordinary coalesced stores and runs at full
// speed.  The reduce-scatter is one `multimem.ld_reduce` per 16 bytes on the multicast address of that buffer: the
// NVSwitch fetches the tp partials, sums them in fp32 and returns 8 reduced bf16 — 1/tp of the bytes per rank that
// unicast pulls would move, and no per-source loop.  The all-gather of the normalised rows is one `multimem.st` per
// 16 bytes on the multicast address of the gathered-activation buffer (the switch replicates it into every rank's
// copy) instead of tp unicast stores.  Flags as before: "my GEMM is done" exchange before the pull, "my rows are in
// every copy" publication after the push.
struct RsMmParams {
  __nv_bfloat16* x;                 // [rows, H] residual rows this kernel updates (the rank's shard, or all rows)
  const uint4* part_mc;             // multicast address of the partial-product buffer, at this kernel's first row
  const __nv_bfloat16* norm_w;      // [H] or null (no normalised output)
  uint4* xn_mc;                     // multicast address of the gathered activations at this rank's first row (or null)
  uint4* xn_local;                  // or: normalised rows into an ordinary local buffer [rows, H] (batched decode under TP)
  int* flags[8];                    // every rank's flag array (index rank = local)
  int* ticket;
  const int* epoch_base;            // device epoch base added to ready_epoch (null: 0) — CUDA-graph replays stay correct
  int H8, tp, rank, gemm_flag_off, ready_flag_off, gemm_epoch, ready_epoch, has_partials;
  float inv_h, eps;
};

template <int MAXC>
__global__ void __launch_bounds__(512) rs_reduce_mm_kernel(const RsMmParams p) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ float red[16];
  __shared__ int s_last;
  const size_t row = (size_t)blockIdx.x * p.H8;
  if (p.has_partials == 1) {
    // every rank's GEMM (the previous kernel on its stream) has written its partial product
    // (has_partials == 2: the caller already ran jl_flag_exchange — this kernel then never waits for a peer)
    // (in-kernel exchange: only for grids that are resident at once and small — the batched decode, <= 128 rows; a
    // prompt-sized grid waiting here could keep a peer-facing kernel of another stream off the SMs)
    const int ge = p.gemm_epoch + (p.epoch_base ? *p.epoch_base : 0);
    if (blockIdx.x == 0 && threadIdx.x < p.tp) {
      __threadfence_system();
      st_release_sys(p.flags[threadIdx.x] + p.gemm_flag_off + p.rank, ge);
    }
    if (threadIdx.x < p.tp)
      wait_flag(p.flags[p.rank] + p.gemm_flag_off + threadIdx.x, ge, p.rank, threadIdx.x, 4);
    __syncthreads();
  }
  const uint4* xg = reinterpret_cast<const uint4*>(p.x);
  uint4 part[MAXC], xa[MAXC];
#pragma unroll
  for (int c = 0; c < MAXC; ++c) {                     // all loads of the row in flight before anything is used
    const int i = threadIdx.x + c * blockDim.x;
    if (i < p.H8) {
      xa[c] = xg[row + i];
      if (p.has_partials) part[c] = mm::ld_reduce_v4bf16x2(p.part_mc + row + i);
    }
  }
  float v[MAXC][8];
  float ss = 0.f;
#pragma unroll
  for (int c = 0; c < MAXC; ++c) {
    const int i = threadIdx.x + c * blockDim.x;
    if (i < p.H8) {
      const uint32_t aw[4] = {xa[c].x, xa[c].y, xa[c].z, xa[c].w};
      uint32_t o[4] = {aw[0], aw[1], aw[2], aw[3]};
      if (p.has_partials) {
        const uint32_t pw[4] = {part[c].x, part[c].y, part[c].z, part[c].w};
#pragma unroll
        for (int j = 0; j < 4; ++j)      // bf16(Σ partials) as one GEMM would have produced it, then the bf16 residual add
          o[j] = pack_bf16(bf16_lo(pw[j]) + bf16_lo(aw[j]), bf16_hi(pw[j]) + bf16_hi(aw[j]));
        reinterpret_cast<uint4*>(p.x)[row + i] = make_uint4(o[0], o[1], o[2], o[3]);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        v[c][2 * j] = bf16_lo(o[j]); v[c][2 * j + 1] = bf16_hi(o[j]);
        ss += v[c][2 * j] * v[c][2 * j] + v[c][2 * j + 1] * v[c][2 * j + 1];
      }
    }
  }
  if (p.xn_mc != nullptr || p.xn_local != nullptr) {
    ss = warp_sum(ss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    __syncthreads();
    if (threadIdx.x < 32) {
      float t = threadIdx.x < (blockDim.x + 31) / 32 ? red[threadIdx.x] : 0.f;
      t = warp_sum(t);
      if (threadIdx.x == 0) red[0] = rsqrtf(t * p.inv_h + p.eps);
    }
    __syncthreads();
    const float r = red[0];
#pragma unroll
    for (int c = 0; c < MAXC; ++c) {
      const int i = threadIdx.x + c * blockDim.x;
      if (i < p.H8) {
        const uint4 g = __ldg(reinterpret_cast<const uint4*>(p.norm_w) + i);
        const uint32_t gw[4] = {g.x, g.y, g.z, g.w};
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
          o[j] = pack_bf16(v[c][2 * j] * r * bf16_lo(gw[j]), v[c][2 * j + 1] * r * bf16_hi(gw[j]));
        if (p.xn_local != nullptr) p.xn_local[row + i] = make_uint4(o[0], o[1], o[2], o[3]);
        else mm::st_v4bf16x2(p.xn_mc + row + i, make_uint4(o[0], o[1], o[2], o[3]));      // one store, every rank's copy
      }
    }
    if (p.xn_mc == nullptr) return;         // local output: nothing to publish
    __threadfence_system();                 // this CTA's rows have reached every copy before it takes a ticket
    __syncthreads();
    if (threadIdx.x == 0) {
      const int prev = atomicAdd(p.ticket, 1);
      s_last = prev == (int)gridDim.x - 1;
      if (s_last) *p.ticket = 0;
    }
    __syncthreads();
    if (s_last && threadIdx.x < p.tp) {
      __threadfence_system();
      st_release_sys(p.flags[threadIdx.x] + p.ready_flag_off + p.rank, p.ready_epoch + (p.epoch_base ? *p.epoch_base : 0));
    }
  }
}

}  // namespace jl

using namespace jl;

// NVLS form (see rs_reduce_mm_kernel).  `x` = the rows this call updates ([rows, H]); `part_mc` / `xn_mc` = multicast
// addresses of the partial-product buffer and of the gathered-activation buffer AT THE FIRST OF THOSE ROWS.
//   * sequence-parallel owner step: rows = mloc, xn_mc != NULL (normalised rows broadcast, "ready" flag published);
//   * plain all-reduce of a row-parallel GEMM (any token count): rows = M, xn_mc = NULL — every rank pulls the sum of
//     all rows and adds it to its replicated residual stream (what GSPMD's all-reduce did for
//     /root/reference/jax_llama/partition.py:67,70).
// has_partials = 0 skips the reduction (chain start on the embedding rows).
//   * batched decode under TP: rows = batch, has_partials = 1 (the "GEMM done" exchange happens inside: the grid is a
//     few dozen resident CTAs), xn_mc = NULL, xn_local = where the rows normalised with `norm_w` go — all-reduce,
//     residual add and the next RMSNorm in one launch.
extern "C" JL_API int jl_rs_reduce_mm(void* x, const void* part_mc, const void* norm_w, void* xn_mc, void* xn_local,
                                      int rows, int H, int has_partials, int tp, int rank, void* const* peer_flags,
                                      int gemm_flag_off, int ready_flag_off, int gemm_epoch, int ready_epoch,
                                      const int* epoch_base, int* ticket, float eps, void* stream) {
  JL_CHECK_ARG(H % 8 == 0 && H <= 16384, "jl_rs_reduce_mm: H=%d must be a multiple of 8 and <= 16384", H);
  JL_CHECK_ARG(tp >= 1 && tp <= 8 && rank >= 0 && rank < tp, "jl_rs_reduce_mm: bad tp/rank");
  JL_CHECK_ARG(!has_partials || part_mc != nullptr, "jl_rs_reduce_mm: needs the multicast address of the partials");
  JL_CHECK_ARG(!(xn_mc && xn_local), "jl_rs_reduce_mm: one normalised output, multicast or local");
  JL_CHECK_ARG(!(xn_mc || xn_local) || norm_w != nullptr, "jl_rs_reduce_mm: a normalised output needs norm_w");
  JL_CHECK_ARG(has_partials != 1 || rows <= 128, "jl_rs_reduce_mm: the in-kernel flag exchange is for <= 128 rows");
  if (rows == 0) return JL_OK;
  RsMmParams p{};
  p.x = (__nv_bfloat16*)x; p.part_mc = (const uint4*)part_mc; p.norm_w = (const __nv_bfloat16*)norm_w;
  p.xn_mc = (uint4*)xn_mc; p.xn_local = (uint4*)xn_local; p.ticket = ticket; p.epoch_base = epoch_base;
  for (int i = 0; i < tp; ++i) p.flags[i] = (int*)peer_flags[i];
  p.H8 = H / 8; p.tp = tp; p.rank = rank; p.gemm_flag_off = gemm_flag_off; p.ready_flag_off = ready_flag_off;
  p.gemm_epoch = gemm_epoch; p.ready_epoch = ready_epoch; p.has_partials = has_partials;
  p.inv_h = 1.0f / (float)H; p.eps = eps;
  const int chunks = p.H8 <= 512 ? 1 : (p.H8 <= 1024 ? 2 : 4);
  const int threads = (((p.H8 + chunks - 1) / chunks + 31) / 32) * 32;
  auto s = (cudaStream_t)stream;
  cudaError_t e;
  if (chunks == 1) e = launch_pdl(rs_reduce_mm_kernel<1>, dim3(rows), dim3(threads), 0, s, p);
  else if (chunks == 2) e = launch_pdl(rs_reduce_mm_kernel<2>, dim3(rows), dim3(threads), 0, s, p);
  else e = launch_pdl(rs_reduce_mm_kernel<4>, dim3(rows), dim3(threads), 0, s, p);
  JL_CUDA(e);
  return JL_OK;
}
