// bf16 GEMM on 5th-generation tensor cores: D = A[Ra,K] · B[Rb,K]^T with fp32 accumulation in TMEM.
//
//   * operands are K-major and arrive by TMA (128-byte swizzle) into a multi-stage mbarrier ring;
//   * one elected thread issues tcgen05.mma (UMMA 128 x BN x 16), completion is signalled with
//     tcgen05.commit → mbarrier, the accumulator (double-buffered, 2·BN TMEM columns) is drained by four
//     epilogue warps with tcgen05.ld while the next tile's MMAs run;
//   * persistent: one CTA per SM walks the tile list, A-tile index fastest so CTAs that run together
//     share the same weight tile through L2.
//
// Two orientations share the kernel:
//   normal  (prefill, M >= 128): A = activations [M,K], B = weights [N,K]; D tile rows are tokens;
//   swap-AB (decode, 8 < M <= 128): A = weights [N,K] (fills the 128-wide MMA M), B = activations
//           [M,K] (MMA N = padded batch); the D tile is transposed on the way out.
// Epilogues: plain bf16 / fp32, residual add, SwiGLU over interleaved (w1,w3) output columns.
//
// Replaces the cuBLAS GEMMs XLA emitted for /root/reference/jax_llama/model.py:119-150 (wq/wk/wv/wo),
// :311-338 (w1/w2/w3) and :699-706 (lm_head).
#include <stdlib.h>

#include "jllama_b200.h"
#include "tcgen05.cuh"

namespace jl {

constexpr int GEMM_BM = 128;
constexpr int GEMM_BK = 64;
constexpr int GEMM_THREADS = 256;   // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warp3 idle, warps 4-7 epilogue

enum { EPI_PLAIN = 0, EPI_RESIDUAL = 1, EPI_SWIGLU = 2 };

struct GemmParams {
  void* out;
  const __nv_bfloat16* residual;
  int Ra, Rb, K;            // A-side rows, B-side rows, reduction length
  int ldo;                  // output row pitch (elements)
  int a_blocks, b_blocks;   // tile counts
  int stages;
  uint64_t hint_a, hint_b;
  // ---- split-K for swap-AB launches with fewer weight-row tiles than SMs (decode at batch 9..128: N = 4096
  // gives 32 tiles, i.e. 22 % of the SMs streaming): each tile's K range is cut into `ksplit` work items, the
  // fp32 partials meet in `ws` ([ksplit][features][tokens]); once all splits of a tile have arrived (ticket) each of
  // them sums its share of the tokens in split order — deterministic — and runs the epilogue.  ksplit == 1: off.
  int ksplit;
  float* ws;
  int* tickets;                      // [2][ticket_half]: arrivals, completions per tile; zero between launches
  int ticket_half;
  unsigned long long* trace;         // optional [grid][16] %globaltimer stamps (jl_gemm_set_trace), NULL = off
};

__host__ __device__ constexpr int tmem_cols_for(int bn) {
  return 2 * bn <= 32 ? 32 : (2 * bn <= 64 ? 64 : (2 * bn <= 128 ? 128 : (2 * bn <= 256 ? 256 : 512)));
}

// one thread records when it reached a point of the kernel (scripts/gemm_trace.py prints the timeline)
__device__ __forceinline__ void gemm_stamp(const GemmParams& p, int ev) {
  if (p.trace != nullptr) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    p.trace[(size_t)blockIdx.x * 16 + ev] = t;
  }
}

template <int BN, int EPI, bool SWAP, bool OUT_F32>
__global__ void __launch_bounds__(GEMM_THREADS, SWAP ? 2 : 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  constexpr uint32_t A_BYTES = GEMM_BM * GEMM_BK * 2;
  constexpr uint32_t B_BYTES = BN * GEMM_BK * 2;
  constexpr uint32_t STAGE_BYTES = A_BYTES + (B_BYTES < 1024 ? 1024 : B_BYTES);
  constexpr uint32_t TMEM_COLS = tmem_cols_for(BN);
  constexpr uint32_t IDESC = tc::idesc_bf16_f32(GEMM_BM, BN);

  const uint32_t smem_base = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
  const int stages = p.stages;
  const uint32_t bars = smem_base + stages * STAGE_BYTES;       // full[stages] | empty[stages] | tfull[2] | tempty[2]
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (stages + s); };
  auto tfull_bar = [&](int s) { return bars + 8u * (2 * stages + s); };
  auto tempty_bar = [&](int s) { return bars + 8u * (2 * stages + 2 + s); };
  const uint32_t tmem_slot = bars + 8u * (2 * stages + 4);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - tc::smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_kb = (p.K + GEMM_BK - 1) / GEMM_BK;
  const int num_tiles = p.a_blocks * p.b_blocks;
  const int ksplit = SWAP ? p.ksplit : 1;
  const int num_work = num_tiles * ksplit;              // (tile, K split) items; split-major so neighbours differ in tile

  // swap-AB (decode): the next kernel of the chain may start right away — it waits for this grid itself before it
  // touches anything this grid writes — so that ITS weight prefetch (below) overlaps our epilogue
  if constexpr (SWAP) pdl_launch_dependents();
  if (threadIdx.x == 0) gemm_stamp(p, 0);
  if (warp == 0 && lane == 0) {
    tc::prefetch_tensormap(&tmap_a);
    tc::prefetch_tensormap(&tmap_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < stages; ++s) { tc::mbar_init(full_bar(s), 1); tc::mbar_init(empty_bar(s), 1); }
    for (int s = 0; s < 2; ++s) { tc::mbar_init(tfull_bar(s), 1); tc::mbar_init(tempty_bar(s), 4); }
    tc::fence_barrier_init();
  }
  if (warp == 2) tc::tmem_alloc(tmem_slot, TMEM_COLS);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *tmem_slot_ptr;
  if (threadIdx.x == 0) gemm_stamp(p, 1);

  if constexpr (!SWAP) {
    pdl_wait();               // everything above overlapped the previous kernel's tail
    pdl_launch_dependents();
  }

  if (warp == 0) {
    // ===================================================== TMA producer
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      int pre = 0;            // leading k-blocks of the first item whose weight tile is already on its way
      if constexpr (SWAP) {
        // the weights do not depend on the previous kernel: fill the ring with A tiles NOW — they stream from HBM
        // while that kernel runs its tail — and only then wait for it; the activation tiles follow after the wait
        if ((int)blockIdx.x < num_work) {
          const int t = blockIdx.x % num_tiles, split = blockIdx.x / num_tiles;
          const int kb0 = split * num_kb / ksplit, kb1 = (split + 1) * num_kb / ksplit;
          const int a_blk = t % p.a_blocks;
          pre = kb1 - kb0 < stages ? kb1 - kb0 : stages;
          for (int i = 0; i < pre; ++i) {
            tc::mbar_expect_tx(full_bar(i), A_BYTES + B_BYTES);
            tc::tma_load_2d(smem_base + i * STAGE_BYTES, &tmap_a, full_bar(i), (kb0 + i) * GEMM_BK, a_blk * GEMM_BM, p.hint_a);
          }
        }
        pdl_wait();
      }
      gemm_stamp(p, 2);
      for (int wi = blockIdx.x; wi < num_work; wi += gridDim.x) {
        const int t = wi % num_tiles, split = wi / num_tiles;
        const int kb0 = split * num_kb / ksplit, kb1 = (split + 1) * num_kb / ksplit;
        const int b_blk = t / p.a_blocks;
        const int a_blk = t % p.a_blocks;
        for (int kb = kb0; kb < kb1; ++kb) {
          const uint32_t sa = smem_base + stage * STAGE_BYTES, sb = sa + A_BYTES;
          if (pre > 0) {
            --pre;
          } else {
            tc::mbar_wait(empty_bar(stage), phase ^ 1);
            tc::mbar_expect_tx(full_bar(stage), A_BYTES + B_BYTES);
            tc::tma_load_2d(sa, &tmap_a, full_bar(stage), kb * GEMM_BK, a_blk * GEMM_BM, p.hint_a);
          }
          tc::tma_load_2d(sb, &tmap_b, full_bar(stage), kb * GEMM_BK, b_blk * BN, p.hint_b);
          if (++stage == stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer (single thread)
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0; int it = 0;
      for (int wi = blockIdx.x; wi < num_work; wi += gridDim.x, ++it) {
        const int split = wi / num_tiles;
        const int kb0 = split * num_kb / ksplit, kb1 = (split + 1) * num_kb / ksplit;
        const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
        tc::mbar_wait(tempty_bar(as), aphase ^ 1);          // epilogue has drained this accumulator
        tc::fence_after_sync();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          tc::mbar_wait(full_bar(stage), phase);
          tc::fence_after_sync();
          if (it == 0 && kb == kb0) gemm_stamp(p, 3);
          const uint32_t sa = smem_base + stage * STAGE_BYTES, sb = sa + A_BYTES;
          const uint64_t da = tc::smem_desc_kmajor_sw128(sa), db = tc::smem_desc_kmajor_sw128(sb);
#pragma unroll
          for (int k = 0; k < GEMM_BK / 16; ++k)            // +32 bytes per UMMA_K inside the swizzle span
            tc::mma_ss(d_tmem, da + 2 * k, db + 2 * k, IDESC, kb != kb0 || k != 0);
          tc::mma_commit(empty_bar(stage));                 // frees the smem slot when these MMAs retire
          if (++stage == stages) { stage = 0; phase ^= 1; }
        }
        tc::mma_commit(tfull_bar(as));                      // accumulator complete → epilogue
        if (it == 0) gemm_stamp(p, 4);
      }
    }
  } else if (warp >= 4) {
    // ===================================================== epilogue: TMEM → registers → global
    const int quarter = warp & 3;                           // TMEM lanes [32·quarter, 32·quarter+32)
    int it = 0;
    if constexpr (SWAP) pdl_wait();                         // residual, output, split-K scratch: after the previous kernel
    for (int wi = blockIdx.x; wi < num_work; wi += gridDim.x, ++it) {
      const int t = wi % num_tiles, split = wi / num_tiles;
      const int b_blk = t / p.a_blocks;
      const int a_blk = t % p.a_blocks;
      const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
      tc::mbar_wait(tfull_bar(as), aphase);
      tc::fence_after_sync();
      if (it == 0 && threadIdx.x == 128) gemm_stamp(p, 5);
      const int ra = a_blk * GEMM_BM + quarter * 32 + lane;  // A-side row of this thread
      const uint32_t t_row = tmem_base + ((uint32_t)(quarter * 32) << 16) + as * BN;
      constexpr int CH = BN >= 32 ? 32 : 16;
      // swap-AB epilogue of one (token, feature = this lane's weight row) value; whole warps call it together
      // rpre: the residual value when the caller already holds it (split-K finish), else it is loaded here
      auto swap_emit = [&](const int tok, float x, const bool pre = false, const float rpre = 0.f) {
        if constexpr (EPI == EPI_SWIGLU) {
          // (gate, up) sit in adjacent lanes: even lane gets its neighbour's value
          const float xb = __bfloat162float(__float2bfloat16(x));
          const float up = __shfl_down_sync(0xffffffffu, xb, 1);
          if ((lane & 1) == 0 && tok < p.Rb && ra + 1 < p.Ra)
            reinterpret_cast<__nv_bfloat16*>(p.out)[(size_t)tok * p.ldo + ra / 2] = __float2bfloat16(silu(xb) * up);
        } else if (tok < p.Rb && ra < p.Ra) {
          if constexpr (OUT_F32) {
            reinterpret_cast<float*>(p.out)[(size_t)tok * p.ldo + ra] = x;
          } else {
            if constexpr (EPI == EPI_RESIDUAL)
              x = __bfloat162float(__float2bfloat16(x)) +
                  (pre ? rpre : __bfloat162float(p.residual[(size_t)tok * p.ldo + ra]));
            reinterpret_cast<__nv_bfloat16*>(p.out)[(size_t)tok * p.ldo + ra] = __float2bfloat16(x);
          }
        }
      };
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += CH) {
        uint32_t v[CH];
        if constexpr (CH == 32) tc::tmem_ld32(t_row + c0, reinterpret_cast<uint32_t(&)[32]>(v));
        else tc::tmem_ld16(t_row + c0, reinterpret_cast<uint32_t(&)[16]>(v));
        tc::tmem_ld_wait();
        const int cb0 = b_blk * BN + c0;                    // first B-side row of this chunk
        if constexpr (!SWAP) {
          // thread owns one token row, CH consecutive output columns
          if (ra < p.Ra) {
            if constexpr (EPI == EPI_SWIGLU) {
              __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + (size_t)ra * p.ldo + cb0 / 2;
              uint32_t w[CH / 4];
#pragma unroll
              for (int j = 0; j < CH / 4; ++j) {
                float g0 = __bfloat162float(__float2bfloat16(__uint_as_float(v[4 * j])));
                float u0 = __bfloat162float(__float2bfloat16(__uint_as_float(v[4 * j + 1])));
                float g1 = __bfloat162float(__float2bfloat16(__uint_as_float(v[4 * j + 2])));
                float u1 = __bfloat162float(__float2bfloat16(__uint_as_float(v[4 * j + 3])));
                w[j] = pack_bf16(silu(g0) * u0, silu(g1) * u1);
              }
              if (cb0 + CH <= p.Rb) {
#pragma unroll
                for (int j = 0; j < CH / 16; ++j)
                  reinterpret_cast<uint4*>(o)[j] = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
              } else {
                for (int j = 0; j < CH / 2; ++j)
                  if (cb0 + 2 * j + 1 < p.Rb)
                    o[j] = reinterpret_cast<__nv_bfloat16*>(w)[j];
              }
            } else if constexpr (OUT_F32) {
              float* o = reinterpret_cast<float*>(p.out) + (size_t)ra * p.ldo + cb0;
              if (cb0 + CH <= p.Rb) {
#pragma unroll
                for (int j = 0; j < CH / 4; ++j)
                  reinterpret_cast<uint4*>(o)[j] = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
              } else {
                for (int j = 0; j < CH; ++j) if (cb0 + j < p.Rb) o[j] = __uint_as_float(v[j]);
              }
            } else {
              __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + (size_t)ra * p.ldo + cb0;
              const bool full = cb0 + CH <= p.Rb;
              uint32_t w[CH / 2];
              if constexpr (EPI == EPI_RESIDUAL) {
                const __nv_bfloat16* r = p.residual + (size_t)ra * p.ldo + cb0;
                if (full) {
#pragma unroll
                  for (int j = 0; j < CH / 8; ++j) {
                    const uint4 rr = reinterpret_cast<const uint4*>(r)[j];
                    const uint32_t rw[4] = {rr.x, rr.y, rr.z, rr.w};
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                      const uint32_t acc = pack_bf16(__uint_as_float(v[8 * j + 2 * q]), __uint_as_float(v[8 * j + 2 * q + 1]));
                      w[4 * j + q] = pack_bf16(bf16_lo(acc) + bf16_lo(rw[q]), bf16_hi(acc) + bf16_hi(rw[q]));
                    }
                  }
                } else {
                  for (int j = 0; j < CH; ++j)
                    if (cb0 + j < p.Rb)
                      o[j] = __float2bfloat16(__bfloat162float(__float2bfloat16(__uint_as_float(v[j]))) +
                                              __bfloat162float(r[j]));
                }
              } else {
#pragma unroll
                for (int j = 0; j < CH / 2; ++j) w[j] = pack_bf16(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
              }
              if (full) {
#pragma unroll
                for (int j = 0; j < CH / 8; ++j)
                  reinterpret_cast<uint4*>(o)[j] = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
              } else if constexpr (EPI != EPI_RESIDUAL) {
                for (int j = 0; j < CH; ++j)
                  if (cb0 + j < p.Rb) o[j] = reinterpret_cast<__nv_bfloat16*>(w)[j];
              }
            }
          }
        } else {
          // swap-AB: lanes are weight rows (output features), columns are tokens
          if (ksplit > 1) {
            // fp32 partial of this K split: ws[split][feature][token] — a thread's tokens are contiguous, so both this
            // store and the finish below move 4 tokens per instruction
            if (ra < p.Ra) {
              float* wrow = p.ws + ((size_t)split * p.Ra + ra) * (size_t)(p.b_blocks * BN) + cb0;
#pragma unroll
              for (int j = 0; j < CH; j += 4)
                *reinterpret_cast<uint4*>(wrow + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
            }
          } else {
            // transposed store: for a fixed token the 32 lanes write 32 consecutive features (coalesced)
#pragma unroll
            for (int j = 0; j < CH; ++j) swap_emit(cb0 + j, __uint_as_float(v[j]));
          }
        }
      }
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(tempty_bar(as));
      if constexpr (SWAP) {
        if (ksplit > 1) {
          // ---- split-K finish, shared by the K splits of the tile: all of them run at the same time (one work item
          // per CTA, see launch_gemm), so each publishes its partial, waits until the tile's ticket shows all ksplit
          // arrivals and then sums + emits ITS share of the tile's tokens (ksplit 16-byte loads per 4 tokens; when the
          // last split to arrive did the whole tile alone, with scalar loads, the finish was 10 us of a 25 us GEMM).
          // Sums run in split order: deterministic.
          // publish: the four warps' partial stores, one barrier, then ONE fence + arrival (cumulative: it orders the
          // stores of all 128 threads; 128 separate fences cost ~1 us more)
          asm volatile("bar.sync 1, 128;" ::: "memory");                // the four epilogue warps
          if (threadIdx.x == 128) {
            gemm_stamp(p, 6);
            __threadfence();
            atomicAdd(p.tickets + t, 1);
            unsigned spins = 0;
            while (ld_acquire_gpu_u32(reinterpret_cast<const unsigned int*>(p.tickets + t)) < (unsigned)ksplit) {
              if (++spins > (1u << 26)) {
                printf("jllama_b200: gemm split-K timed out waiting for tile %d (block %d)\n", t, (int)blockIdx.x);
                __trap();
              }
            }
          }
          asm volatile("bar.sync 1, 128;" ::: "memory");
          if (threadIdx.x == 128) gemm_stamp(p, 7);
          {
            // this split's share of the tile's tokens, 4 (one 16-byte load per split) at a time; the residual values
            // are requested first, the partials (L2, written by other SMs: ld.cg) of all splits right behind them
            const int per = ((BN + ksplit - 1) / ksplit + 3) & ~3;
            const int j_lo = split * per, j_hi = j_lo + per < BN ? j_lo + per : BN;
            const size_t tok_pitch = (size_t)(p.b_blocks * BN), split_stride = (size_t)p.Ra * tok_pitch;
            const bool row_ok = ra < p.Ra;                  // (all lanes stay in the loop: the SwiGLU emit shuffles)
            const float* wrow = p.ws + (size_t)(row_ok ? ra : 0) * tok_pitch + b_blk * BN;
#pragma unroll 2
            for (int j0 = j_lo; j0 < j_hi; j0 += 4) {
              float rr[4] = {0.f, 0.f, 0.f, 0.f};
              if constexpr (EPI == EPI_RESIDUAL) {
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                  const int tok = b_blk * BN + j0 + u;
                  if (tok < p.Rb && row_ok) rr[u] = __bfloat162float(p.residual[(size_t)tok * p.ldo + ra]);
                }
              }
              float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
              for (int sp = 0; sp < ksplit; ++sp) {
                const float4 pv = __ldcg(reinterpret_cast<const float4*>(wrow + sp * split_stride + j0));
                acc.x += pv.x; acc.y += pv.y; acc.z += pv.z; acc.w += pv.w;
              }
              const float x[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
              for (int u = 0; u < 4; ++u) swap_emit(b_blk * BN + j0 + u, x[u], true, rr[u]);
            }
          }
          // the last split to finish re-arms the tile's two counters for the next launch
          asm volatile("bar.sync 1, 128;" ::: "memory");
          if (threadIdx.x == 128) {
            gemm_stamp(p, 8);
            int* done = p.tickets + p.ticket_half + t;
            if (atomicAdd(done, 1) == ksplit - 1) { *done = 0; p.tickets[t] = 0; }
          }
        }
      }
    }
  }

  tc::fence_before_sync();
  __syncthreads();
  if (threadIdx.x == 0) gemm_stamp(p, 9);
  if (warp == 2) tc::tmem_dealloc(tmem_base, TMEM_COLS);
}

static unsigned long long* g_gemm_trace = nullptr;

// Caller-provided scratch for split-K (jl_gemm_set_workspace), one per device; none registered → no split-K
struct SplitKWorkspace { float* ws; size_t ws_bytes; int* tickets; int n_tickets; };
static SplitKWorkspace g_splitk[16] = {};
static SplitKWorkspace splitk_workspace() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) return SplitKWorkspace{};
  return g_splitk[dev];
}

template <int BN, int EPI, bool SWAP, bool OUT_F32>
static int launch_gemm(const void* a, const void* b, const GemmParams& p0, int Ra, int Rb, int K, int num_sms,
                       cudaStream_t stream) {
  GemmParams p = p0;
  p.trace = g_gemm_trace;
  CUtensorMap ta, tb;
  int rc = make_tmap_2d_bf16(&ta, a, Ra, K, K, GEMM_BM);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&tb, b, Rb, K, K, BN);
  if (rc) return rc;
  constexpr uint32_t A_BYTES = GEMM_BM * GEMM_BK * 2;
  constexpr uint32_t B_BYTES = BN * GEMM_BK * 2;
  constexpr uint32_t STAGE_BYTES = A_BYTES + (B_BYTES < 1024 ? 1024 : B_BYTES);
  const int num_kb = (K + GEMM_BK - 1) / GEMM_BK;
  p.a_blocks = (Ra + GEMM_BM - 1) / GEMM_BM;
  p.b_blocks = (Rb + BN - 1) / BN;
  int grid = p.a_blocks * p.b_blocks;
  p.ksplit = 1;
  if constexpr (SWAP) {
    // few weight-row tiles (N <= ~9k at decode batch sizes): cut K so that (nearly) every SM streams weights.  Every
    // (tile, split) item gets its own CTA — the shared finish in the kernel relies on all splits running at once.
    const SplitKWorkspace wsp = splitk_workspace();
    if (wsp.ws != nullptr && Rb <= 64 && grid * 2 <= num_sms && grid <= wsp.n_tickets / 2) {
      int ks = num_sms / grid;
      if (ks > num_kb / 4) ks = num_kb / 4;               // at least 4 k-blocks (256 columns) per item
      if (ks > 16) ks = 16;
      while (ks > 1 && (size_t)ks * Ra * (size_t)(p.b_blocks * BN) * sizeof(float) > wsp.ws_bytes) --ks;
      if (ks > 1) {
        p.ksplit = ks; p.ws = wsp.ws; p.tickets = wsp.tickets; p.ticket_half = wsp.n_tickets / 2;
        grid *= ks;
      }
    }
  }
  if (grid > num_sms) grid = num_sms;
  // Ring depth.  Token-major GEMMs and long weight streams (> 1.25 MB per CTA: w13, lm_head) take the whole shared
  // memory (10 stages).  Short decode GEMMs (wqkv, wo, w2 at batch 9..128) stay under half of it, so that the CTAs of
  // the NEXT kernel of the chain fit on the same SM and their weight prefetch — issued before they wait for this
  // grid — runs under our epilogue and split-K finish; those prologue / tail latencies were most of a 25 us GEMM.
  // A split-K launch always takes the small ring: its CTAs wait for each other, so all of them must become resident
  // even while CTAs of a neighbouring kernel (the previous one draining, the next one prefetching) share the SMs:
  // two ring-sized CTAs per SM always fit.
  int budget = 200 * 1024;
  if (SWAP && (p.ksplit > 1 || (size_t)Ra * K * 2 / grid <= (size_t)1280 * 1024)) budget = 100 * 1024;
  int stages = (int)(budget / STAGE_BYTES);
  if (stages > 10) stages = 10;
  if (stages > num_kb * 2) stages = num_kb * 2 < 2 ? 2 : num_kb * 2;
  p.stages = stages;
  const size_t smem = (size_t)stages * STAGE_BYTES + 8 * (2 * stages + 4) + 16 + 1024;
  auto kern = gemm_kernel<BN, EPI, SWAP, OUT_F32>;
  JL_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  JL_CUDA(launch_pdl(kern, dim3(grid), dim3(GEMM_THREADS), smem, stream, ta, tb, p));
  return JL_OK;
}

template <int EPI, bool OUT_F32>
static int dispatch_gemm(const void* x, const void* w, const GemmParams& p, int M, int N, int K, int num_sms,
                         cudaStream_t s) {
  if (M > 128) {
    // normal orientation: A = activations, B = weights
    GemmParams q = p;
    q.Ra = M; q.Rb = N;
    q.hint_a = tc::EVICT_LAST; q.hint_b = tc::EVICT_NORMAL;
    if (N >= 256) return launch_gemm<256, EPI, false, OUT_F32>(x, w, q, M, N, K, num_sms, s);
    return launch_gemm<128, EPI, false, OUT_F32>(x, w, q, M, N, K, num_sms, s);
  }
  // swap-AB: A = weights (streamed once → evict-first), B = activations (re-read by every CTA)
  GemmParams q = p;
  q.Ra = N; q.Rb = M;
  q.hint_a = tc::EVICT_FIRST; q.hint_b = tc::EVICT_LAST;
  if (M <= 16) return launch_gemm<16, EPI, true, OUT_F32>(w, x, q, N, M, K, num_sms, s);
  if (M <= 32) return launch_gemm<32, EPI, true, OUT_F32>(w, x, q, N, M, K, num_sms, s);
  if (M <= 64) return launch_gemm<64, EPI, true, OUT_F32>(w, x, q, N, M, K, num_sms, s);
  return launch_gemm<128, EPI, true, OUT_F32>(w, x, q, N, M, K, num_sms, s);
}

}  // namespace jl

using namespace jl;

// Registers (or, with ws == NULL, removes) the split-K scratch of the CURRENT device: `ws` fp32 workspace of
// ws_bytes, `tickets` n_tickets ints that must be zero.  The library keeps the pointers; the caller owns the memory.
// Optional timeline of the single-CTA kernel: [grid][16] u64 device buffer (0 entry, 1 set-up done, 2 producer past
// the dependency wait, 3 first stage landed, 4 last MMA issued, 5 accumulator complete, 6 partial published,
// 7 all splits arrived, 8 share emitted, 9 exit); NULL switches it off.
extern "C" JL_API void jl_gemm_set_trace(void* buf) { g_gemm_trace = (unsigned long long*)buf; }

extern "C" JL_API int jl_gemm_set_workspace(void* ws, size_t ws_bytes, int* tickets, int n_tickets) {
  int dev = 0;
  JL_CUDA(cudaGetDevice(&dev));
  JL_CHECK_ARG(dev >= 0 && dev < 16, "jl_gemm_set_workspace: device %d out of range", dev);
  g_splitk[dev] = SplitKWorkspace{(float*)ws, ws ? ws_bytes : 0, tickets, ws ? n_tickets : 0};
  return JL_OK;
}

extern "C" JL_API int jl_gemm(const void* x, const void* w, const void* residual, void* out, int M, int N, int K,
                       int mode, int out_f32, int num_sms, void* stream) {
  JL_CHECK_ARG(M >= 1 && N >= 1 && K >= 8 && K % 8 == 0, "jl_gemm: bad shape M=%d N=%d K=%d (K %% 8 == 0)", M, N, K);
  JL_CHECK_ARG(mode >= 0 && mode <= 2, "jl_gemm: bad mode %d", mode);
  JL_CHECK_ARG(mode != EPI_SWIGLU || N % 2 == 0, "jl_gemm: SwiGLU needs an even N");
  JL_CHECK_ARG(mode != EPI_RESIDUAL || residual != nullptr, "jl_gemm: residual pointer missing");
  JL_CHECK_ARG(!(out_f32 && mode != EPI_PLAIN), "jl_gemm: fp32 output only in plain mode");
  JL_CHECK_ARG(N % 8 == 0, "jl_gemm: N=%d must be a multiple of 8", N);
  if (num_sms <= 0) num_sms = 148;
  GemmParams p{};
  p.out = out;
  p.residual = (const __nv_bfloat16*)residual;
  p.K = K;
  p.ldo = mode == EPI_SWIGLU ? N / 2 : N;
  auto s = (cudaStream_t)stream;
  // token-major tiles of the prefill shapes go to the CTA-pair kernel (cta_group::2, gemm_2cta.cu); JLLAMA_GEMM_2CTA=0
  // keeps everything on the single-CTA kernel below (A/B measurements)
  static const bool pair_on = [] { const char* e = getenv("JLLAMA_GEMM_2CTA"); return !(e && e[0] == '0'); }();
  if (pair_on && !out_f32 && M >= 256) {
    const int rc = jl_gemm_2cta(x, w, residual, out, M, N, K, mode, num_sms, stream);
    if (rc != JL_ERR_UNSUPPORTED) return rc;
  }
  if (mode == EPI_PLAIN && out_f32) return dispatch_gemm<EPI_PLAIN, true>(x, w, p, M, N, K, num_sms, s);
  if (mode == EPI_PLAIN) return dispatch_gemm<EPI_PLAIN, false>(x, w, p, M, N, K, num_sms, s);
  if (mode == EPI_RESIDUAL) return dispatch_gemm<EPI_RESIDUAL, false>(x, w, p, M, N, K, num_sms, s);
  return dispatch_gemm<EPI_SWIGLU, false>(x, w, p, M, N, K, num_sms, s);
}
