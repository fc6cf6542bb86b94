// Sequence-parallel prefill glue between the fused GEMM→reduce-scatter and all-gather→GEMM kernels
// (gemm_tcgen05.cu, COMM_SCATTER / COMM_GATHER): the owner-side reduction.
//
// After jl_gemm_rs every rank holds, for ITS rows of the residual stream, one bf16 partial tile from every
// rank.  `rs_reduce_kernel` (one CTA per row)
//   1. makes sure every peer's GEMM has finished writing (flag exchange: st.release.sys / ld.acquire.sys),
//   2. x_row += Σ_src partial[src][row]   (fp32 sum in rank order → identical on every run),
//   3. writes RMSNorm(x_row)·g — the input of the next column-parallel GEMM — into EVERY rank's gathered
//      activation buffer at this rank's row offset (the all-gather is a push of fire-and-forget NVLink stores
//      from the kernel that produces the rows; a first version let the consuming GEMM pull the shards with
//      one warp per CTA and was latency-bound: 17.9 ms vs 14.0 ms for GEMM + NCCL all-reduce at tp = 8), and
//   4. the last CTA to finish (atomic ticket) publishes "shard ready" to every rank, which is what the next
//      kernel's in-kernel all-gather waits for.
// With n_src == 0 it only performs 3-4 (used once, on the embedding rows, to start the chain).
// What XLA/GSPMD emitted here for the reference: an NCCL all-reduce of the whole [M, H] activation after wo / w2
// (/root/reference/jax_llama/partition.py:67,70), then replicated norms (:73-76).
#include "common.cuh"
#include "jllama_b200.h"
#include "nvls.cuh"

namespace jl {

struct RsParams {
  __nv_bfloat16* x;                 // [mloc, H] residual-stream shard (in/out)
  const __nv_bfloat16* partials;    // [n_src][mloc, H] received partial tiles (local memory)
  const __nv_bfloat16* norm_w;      // [H] (may be null: no normalised output)
  __nv_bfloat16* xn[8];             // every rank's [M, H] gathered-activation buffer (xn[0] null: no normalised output)
  int xn_row0;                      // first row of this rank's shard inside those buffers
  int* flags[8];                    // every rank's flag array (index rank = local)
  int* ticket;                      // self-resetting CTA counter
  int H8, n_src, tp, rank, gemm_flag_off, ready_flag_off, gemm_epoch, ready_epoch;
  float inv_h, eps;
};

template <int MAXC>
__global__ void __launch_bounds__(1024) rs_reduce_kernel(const RsParams p) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ float red[32];
  __shared__ int s_last;
  const size_t row = (size_t)blockIdx.x * p.H8;
  if (p.n_src > 0) {
    // ---- 1. all peers' partial tiles have landed
    if (blockIdx.x == 0 && threadIdx.x < p.tp)
      st_release_sys(p.flags[threadIdx.x] + p.gemm_flag_off + p.rank, p.gemm_epoch);
    if (threadIdx.x < p.tp)
      wait_flag(p.flags[p.rank] + p.gemm_flag_off + threadIdx.x, p.gemm_epoch, p.rank, threadIdx.x, 4);
    __syncthreads();
  }
  // ---- 2. reduce + residual
  const uint4* xg = reinterpret_cast<const uint4*>(p.x);
  float v[MAXC][8];
  float ss = 0.f;
#pragma unroll
  for (int c = 0; c < MAXC; ++c) {
    const int i = threadIdx.x + c * blockDim.x;
    if (i < p.H8) {
      const uint4 a = xg[row + i];
      float acc[8] = {bf16_lo(a.x), bf16_hi(a.x), bf16_lo(a.y), bf16_hi(a.y), bf16_lo(a.z), bf16_hi(a.z), bf16_lo(a.w), bf16_hi(a.w)};
      if (p.n_src > 0) {
        float sum[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int s = 0; s < p.n_src; ++s) {
          const uint4* pp = reinterpret_cast<const uint4*>(p.partials) + ((size_t)s * gridDim.x + blockIdx.x) * p.H8 + i;
          uint4 t;
          asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(t.x), "=r"(t.y), "=r"(t.z), "=r"(t.w) : "l"(pp) : "memory");
          sum[0] += bf16_lo(t.x); sum[1] += bf16_hi(t.x); sum[2] += bf16_lo(t.y); sum[3] += bf16_hi(t.y);
          sum[4] += bf16_lo(t.z); sum[5] += bf16_hi(t.z); sum[6] += bf16_lo(t.w); sum[7] += bf16_hi(t.w);
        }
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          // y = bf16(Σ partials) as a single GEMM would have produced it, then the bf16 residual add
          const uint32_t y = pack_bf16(sum[2 * j], sum[2 * j + 1]);
          o[j] = pack_bf16(bf16_lo(y) + acc[2 * j], bf16_hi(y) + acc[2 * j + 1]);
          acc[2 * j] = bf16_lo(o[j]);
          acc[2 * j + 1] = bf16_hi(o[j]);
        }
        reinterpret_cast<uint4*>(p.x)[row + i] = make_uint4(o[0], o[1], o[2], o[3]);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) { v[c][j] = acc[j]; ss += acc[j] * acc[j]; }
    }
  }
  // ---- 3. RMSNorm of the updated row → peer-visible shard
  if (p.xn[0] != nullptr) {
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
        // all-gather by push: the row goes straight into every rank's copy of the gathered activations
        const uint4 ov = make_uint4(o[0], o[1], o[2], o[3]);
        for (int d = 0; d < p.tp; ++d)
          reinterpret_cast<uint4*>(p.xn[d])[(size_t)p.xn_row0 * p.H8 + row + i] = ov;
      }
    }
    // ---- 4. last CTA publishes "shard ready" to every rank (self included)
    __threadfence_system();                 // this CTA's rows have reached the peers before it takes a ticket
    __syncthreads();
    if (threadIdx.x == 0) {
      const int prev = atomicAdd(p.ticket, 1);
      s_last = prev == (int)gridDim.x - 1;
      if (s_last) *p.ticket = 0;
    }
    __syncthreads();
    if (s_last && threadIdx.x < p.tp) {
      __threadfence_system();
      st_release_sys(p.flags[threadIdx.x] + p.ready_flag_off + p.rank, p.ready_epoch);
    }
  }
}


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
    if (blockIdx.x == 0 && threadIdx.x < p.tp)
      st_release_sys(p.flags[threadIdx.x] + p.gemm_flag_off + p.rank, p.gemm_epoch);
    if (threadIdx.x < p.tp)
      wait_flag(p.flags[p.rank] + p.gemm_flag_off + threadIdx.x, p.gemm_epoch, p.rank, threadIdx.x, 4);
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
  if (p.xn_mc != nullptr) {
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
        mm::st_v4bf16x2(p.xn_mc + row + i, make_uint4(o[0], o[1], o[2], o[3]));      // one store, every rank's copy
      }
    }
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

extern "C" JL_API int jl_rs_reduce(void* x, const void* partials, const void* norm_w, void* const* xn_peers,
                                   int xn_row0, int mloc, int H, int n_src, int tp, int rank,
                                   void* const* peer_flags, int gemm_flag_off,
                                   int ready_flag_off, int gemm_epoch, int ready_epoch, int* ticket, float eps,
                                   void* stream) {
  JL_CHECK_ARG(H % 8 == 0 && H <= 32768, "jl_rs_reduce: H=%d must be a multiple of 8 and <= 32768", H);
  JL_CHECK_ARG(tp >= 1 && tp <= 8 && rank >= 0 && rank < tp && n_src >= 0 && n_src <= tp, "jl_rs_reduce: bad tp/rank/n_src");
  if (mloc == 0) return JL_OK;
  RsParams p{};
  p.x = (__nv_bfloat16*)x; p.partials = (const __nv_bfloat16*)partials; p.norm_w = (const __nv_bfloat16*)norm_w;
  for (int i = 0; i < tp; ++i) p.xn[i] = xn_peers ? (__nv_bfloat16*)xn_peers[i] : nullptr;
  p.xn_row0 = xn_row0; p.ticket = ticket;
  for (int i = 0; i < tp; ++i) p.flags[i] = (int*)peer_flags[i];
  p.H8 = H / 8; p.n_src = n_src; p.tp = tp; p.rank = rank; p.gemm_flag_off = gemm_flag_off;
  p.ready_flag_off = ready_flag_off; p.gemm_epoch = gemm_epoch; p.ready_epoch = ready_epoch;
  p.inv_h = 1.0f / (float)H; p.eps = eps;
  const int chunks = p.H8 <= 512 ? 1 : (p.H8 <= 1024 ? 2 : 4);
  const int threads = (((p.H8 + chunks - 1) / chunks + 31) / 32) * 32;
  auto s = (cudaStream_t)stream;
  cudaError_t e;
  if (chunks == 1) e = launch_pdl(rs_reduce_kernel<1>, dim3(mloc), dim3(threads), 0, s, p);
  else if (chunks == 2) e = launch_pdl(rs_reduce_kernel<2>, dim3(mloc), dim3(threads), 0, s, p);
  else e = launch_pdl(rs_reduce_kernel<4>, dim3(mloc), dim3(threads), 0, s, p);
  JL_CUDA(e);
  return JL_OK;
}

// NVLS form (see rs_reduce_mm_kernel).  `x` = the rows this call updates ([rows, H]); `part_mc` / `xn_mc` = multicast
// addresses of the partial-product buffer and of the gathered-activation buffer AT THE FIRST OF THOSE ROWS.
//   * sequence-parallel owner step: rows = mloc, xn_mc != NULL (normalised rows broadcast, "ready" flag published);
//   * plain all-reduce of a row-parallel GEMM (any token count): rows = M, xn_mc = NULL — every rank pulls the sum of
//     all rows and adds it to its replicated residual stream (what GSPMD's all-reduce did for
//     /root/reference/jax_llama/partition.py:67,70).
// has_partials = 0 skips the reduction (chain start on the embedding rows).
extern "C" JL_API int jl_rs_reduce_mm(void* x, const void* part_mc, const void* norm_w, void* xn_mc, int rows, int H,
                                      int has_partials, int tp, int rank, void* const* peer_flags, int gemm_flag_off,
                                      int ready_flag_off, int gemm_epoch, int ready_epoch, const int* epoch_base, int* ticket,
                                      float eps, void* stream) {
  JL_CHECK_ARG(H % 8 == 0 && H <= 16384, "jl_rs_reduce_mm: H=%d must be a multiple of 8 and <= 16384", H);
  JL_CHECK_ARG(tp >= 1 && tp <= 8 && rank >= 0 && rank < tp, "jl_rs_reduce_mm: bad tp/rank");
  JL_CHECK_ARG(!has_partials || part_mc != nullptr, "jl_rs_reduce_mm: needs the multicast address of the partials");
  if (rows == 0) return JL_OK;
  RsMmParams p{};
  p.x = (__nv_bfloat16*)x; p.part_mc = (const uint4*)part_mc; p.norm_w = (const __nv_bfloat16*)norm_w;
  p.xn_mc = (uint4*)xn_mc; p.ticket = ticket; p.epoch_base = epoch_base;
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
