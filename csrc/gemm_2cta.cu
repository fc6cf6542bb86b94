// Prefill GEMM on CTA PAIRS: D[256 x 256] tiles computed by `tcgen05.mma.cta_group::2` (UMMA 256 x 256 x 16).
//
// Two CTAs of a 2-CTA cluster sit on the two SMs of one TPC and behave as one MMA engine:
//   * each CTA stages only HALF of the operands — its 128 token rows of A and its 128 weight rows of B (32 KB per
//     k-stage instead of 48 KB for a 128 x 256 single-CTA tile) — and the tensor cores of both SMs read the two B
//     halves from both shared memories: half the shared-memory reads and half the L2 → SM traffic per FLOP;
//   * both producers issue their TMA loads with `.cta_group::2` so that all four loads of a stage complete on the
//     LEADER's (cluster rank 0) mbarrier; the leader's single MMA thread issues the pair-wide MMAs, and
//     `tcgen05.commit ... .multicast::cluster` releases the stage / publishes the accumulator in BOTH CTAs at once;
//   * each CTA holds the fp32 accumulator rows of its own 128 tokens in its own TMEM (double buffered: 2 x 256
//     columns = all 512), drained by its four epilogue warps while the next tile's MMAs run;
//   * the epilogue goes through shared memory and leaves the SM as TMA tensor stores (`cp.async.bulk.tensor` shared →
//     global, 32 x 64 element boxes): full 128-byte rows on the wire instead of one 64-byte row write per thread, and
//     rows beyond M are clipped by the tensor map.
// All TMA loads here read this GPU's own memory through tensor maps over local buffers.
//
// Used for token-major GEMMs with M >= 256 and N a multiple of 256 (every prefill GEMM of the LLaMA shapes); other
// shapes take the single-CTA kernel (gemm_tcgen05.cu).  Replaces the cuBLAS GEMMs XLA emitted for
// /root/reference/jax_llama/model.py:119-150 (wq/wk/wv/wo) and :311-338 (w1/w2/w3).
#include <stdlib.h>

#include "jllama_b200.h"
#include "tcgen05.cuh"

namespace jl {

constexpr int G2_BM = 128;            // token rows per CTA (256 per pair)
constexpr int G2_BN = 256;            // weight rows per pair tile (128 staged per CTA)
constexpr int G2_BK = 64;
constexpr int G2_THREADS = 256;       // warp0 TMA, warp1 MMA (leader), warp2 TMEM alloc, warp3 idle, warps 4-7 epilogue
constexpr uint32_t G2_A_BYTES = G2_BM * G2_BK * 2;                 // 16 KB
constexpr uint32_t G2_B_BYTES = (G2_BN / 2) * G2_BK * 2;           // 16 KB
constexpr uint32_t G2_STAGE_BYTES = G2_A_BYTES + G2_B_BYTES;
constexpr uint32_t G2_EPI_BYTES = 4 * 32 * 64 * 2;                 // per CTA: 4 warps x [32 rows][64 cols] bf16
constexpr uint32_t G2_PEER_MASK = 0xFEFFFFFFu;                     // shared::cluster address of the same offset in CTA 0

enum { G2_PLAIN = 0, G2_RESIDUAL = 1, G2_SWIGLU = 2 };

struct Gemm2Params {
  const __nv_bfloat16* residual;
  int M, N, K, ldo;           // ldo = output row pitch in elements
  int a_blocks, b_blocks, stages;
};

__device__ __forceinline__ uint32_t g2_cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void g2_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local_smem_addr` in CTA `rank` of this cluster
__device__ __forceinline__ uint32_t g2_mapa(uint32_t local_smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(rank));
  return r;
}
// arrive on an mbarrier of another CTA of the cluster.  Default semantics on purpose: `.release.cluster` compiles to
// MEMBAR.ALL.GPU + ERRBAR in front of the arrive, which in the producer loop waited for the previous stage's TMA
// loads every time (pipeline depth 1: the first version of this kernel ran at half the single-CTA kernel's speed).
__device__ __forceinline__ void g2_remote_arrive(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load issued by either CTA of the pair; the bytes are accounted on the LEADER's mbarrier
__device__ __forceinline__ void g2_tma_load_2d(uint32_t smem_dst, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                                               uint64_t hint) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar & G2_PEER_MASK), "r"(c0), "r"(c1), "l"(hint)
      : "memory");
}
__device__ __forceinline__ void g2_tma_store_2d(const CUtensorMap* map, uint32_t smem_src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(map)),
               "r"(smem_src), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void g2_mma(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, bool acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"((uint32_t)acc)
      : "memory");
}
// arrive on the mbarrier at this offset in BOTH CTAs once all MMAs issued so far have completed
__device__ __forceinline__ void g2_commit_both(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
               "h"((uint16_t)3)
               : "memory");
}

template <int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(G2_THREADS, 1)
gemm2_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
             const __grid_constant__ CUtensorMap tmap_out, const Gemm2Params p) {
  extern __shared__ uint8_t smem_raw[];
  constexpr uint32_t IDESC = tc::idesc_bf16_f32(2 * G2_BM, G2_BN);
  const uint32_t smem_base = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
  const int stages = p.stages;
  const uint32_t epi_smem = smem_base + stages * G2_STAGE_BYTES;
  const uint32_t bars = epi_smem + G2_EPI_BYTES;            // full[stages] | empty[stages] | tfull[2] | tempty[2]
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (stages + s); };
  auto tfull_bar = [&](int s) { return bars + 8u * (2 * stages + s); };
  auto tempty_bar = [&](int s) { return bars + 8u * (2 * stages + 2 + s); };
  const uint32_t tmem_slot = bars + 8u * (2 * stages + 4);
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - tc::smem_u32(smem_raw)));
  uint8_t* epi_ptr = smem_raw + (epi_smem - tc::smem_u32(smem_raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = g2_cluster_rank();                  // 0 = leader
  const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  const int num_kb = (p.K + G2_BK - 1) / G2_BK;
  const int num_tiles = p.a_blocks * p.b_blocks;

  if (warp == 0 && lane == 0) {
    tc::prefetch_tensormap(&tmap_a);
    tc::prefetch_tensormap(&tmap_b);
    tc::prefetch_tensormap(&tmap_out);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < stages; ++s) {
      tc::mbar_init(full_bar(s), 2);                        // leader's: its own expect_tx arrive + the peer producer's arrive
      tc::mbar_init(empty_bar(s), 1);                       // one multicast commit per use
    }
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(tfull_bar(s), 1);                       // one multicast commit per tile
      tc::mbar_init(tempty_bar(s), 8);                      // leader's: 4 epilogue warps of each CTA
    }
    tc::fence_barrier_init();
  }
  if (warp == 2) {                                          // the same warp of both CTAs allocates the pair's TMEM
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc::fence_before_sync();
  g2_cluster_sync();                                        // barriers of both CTAs are initialised before any remote use
  tc::fence_after_sync();
  const uint32_t tmem_base = *tmem_slot_ptr;

  pdl_wait();
  pdl_launch_dependents();

  if (warp == 0) {
    // ===================================================== TMA producer (one per CTA: its halves of A and B)
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int t = pair; t < num_tiles; t += n_pairs) {
        const int b_blk = t / p.a_blocks, a_blk = t % p.a_blocks;
        const int a_row = a_blk * 2 * G2_BM + (int)rank * G2_BM;
        const int b_row = b_blk * G2_BN + (int)rank * (G2_BN / 2);
        for (int kb = 0; kb < num_kb; ++kb) {
          tc::mbar_wait(empty_bar(stage), phase ^ 1);
          const uint32_t sa = smem_base + stage * G2_STAGE_BYTES, sb = sa + G2_A_BYTES;
          if (rank == 0) tc::mbar_expect_tx(full_bar(stage), 2 * G2_STAGE_BYTES);     // both CTAs' bytes land here
          else g2_remote_arrive(g2_mapa(full_bar(stage), 0));
          g2_tma_load_2d(sa, &tmap_a, full_bar(stage), kb * G2_BK, a_row, tc::EVICT_NORMAL);
          g2_tma_load_2d(sb, &tmap_b, full_bar(stage), kb * G2_BK, b_row, tc::EVICT_LAST);
          if (++stage == stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer: one thread of the leader CTA, for the pair
    if (rank == 0 && lane == 0) {
      int stage = 0; uint32_t phase = 0; int it = 0;
      for (int t = pair; t < num_tiles; t += n_pairs, ++it) {
        const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
        tc::mbar_wait(tempty_bar(as), aphase ^ 1);          // both CTAs' epilogues have drained this accumulator
        tc::fence_after_sync();
        const uint32_t d_tmem = tmem_base + as * G2_BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          tc::mbar_wait(full_bar(stage), phase);
          tc::fence_after_sync();
          const uint32_t sa = smem_base + stage * G2_STAGE_BYTES, sb = sa + G2_A_BYTES;
          const uint64_t da = tc::smem_desc_kmajor_sw128(sa), db = tc::smem_desc_kmajor_sw128(sb);
#pragma unroll
          for (int k = 0; k < G2_BK / 16; ++k) g2_mma(d_tmem, da + 2 * k, db + 2 * k, IDESC, kb != 0 || k != 0);
          g2_commit_both(empty_bar(stage));                 // frees the slot in both CTAs when these MMAs retire
          if (++stage == stages) { stage = 0; phase ^= 1; }
        }
        g2_commit_both(tfull_bar(as));                      // accumulator complete → both epilogues
      }
    }
  } else if (warp >= 4) {
    // ===================================================== epilogue: TMEM → registers → shared → TMA store
    const int quarter = warp & 3;                           // TMEM lanes [32·quarter, 32·quarter+32)
    const uint32_t my_smem = epi_smem + quarter * (32 * 64 * 2);
    uint8_t* my_ptr = epi_ptr + quarter * (32 * 64 * 2);
    const uint32_t tempty_leader = g2_mapa(tempty_bar(0), 0);
    int it = 0;
    for (int t = pair; t < num_tiles; t += n_pairs, ++it) {
      const int b_blk = t / p.a_blocks, a_blk = t % p.a_blocks;
      const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
      tc::mbar_wait(tfull_bar(as), aphase);
      tc::fence_after_sync();
      const int row0 = a_blk * 2 * G2_BM + (int)rank * G2_BM + quarter * 32;      // first token row of this warp
      const int ra = row0 + lane;
      const uint32_t t_row = tmem_base + ((uint32_t)(quarter * 32) << 16) + as * G2_BN;
#pragma unroll 1
      for (int c0 = 0; c0 < G2_BN; c0 += 64) {
        // 64 accumulator columns of this thread's row → up to 64 bf16 (32 for SwiGLU) in shared memory
        uint32_t w[32];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t v[32];
          tc::tmem_ld32(t_row + c0 + h * 32, v);
          tc::tmem_ld_wait();
          if constexpr (EPI == G2_SWIGLU) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float g0 = __bfloat162float(__float2bfloat16(__uint_as_float(v[4 * j])));
              const float u0 = __bfloat162float(__float2bfloat16(__uint_as_float(v[4 * j + 1])));
              const float g1 = __bfloat162float(__float2bfloat16(__uint_as_float(v[4 * j + 2])));
              const float u1 = __bfloat162float(__float2bfloat16(__uint_as_float(v[4 * j + 3])));
              w[h * 8 + j] = pack_bf16(silu(g0) * u0, silu(g1) * u1);
            }
          } else if constexpr (EPI == G2_RESIDUAL) {
            const int col = b_blk * G2_BN + c0 + h * 32;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              uint4 rr = make_uint4(0, 0, 0, 0);
              if (ra < p.M) rr = *reinterpret_cast<const uint4*>(p.residual + (size_t)ra * p.ldo + col + 8 * j);
              const uint32_t rw[4] = {rr.x, rr.y, rr.z, rr.w};
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const uint32_t acc = pack_bf16(__uint_as_float(v[8 * j + 2 * q]), __uint_as_float(v[8 * j + 2 * q + 1]));
                w[h * 16 + 4 * j + q] = pack_bf16(bf16_lo(acc) + bf16_lo(rw[q]), bf16_hi(acc) + bf16_hi(rw[q]));
              }
            }
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) w[h * 16 + j] = pack_bf16(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
          }
        }
        // the previous TMA store of this warp must have finished READING the staging buffer
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncwarp();
        constexpr int OUT_COLS = EPI == G2_SWIGLU ? 32 : 64;              // bf16 columns this chunk produces
        // staging layout [32 rows][OUT_COLS] bf16, row-major: what a plain (unswizzled) tensor-map box expects
        uint4* dst = reinterpret_cast<uint4*>(my_ptr + lane * (OUT_COLS * 2));
#pragma unroll
        for (int j = 0; j < OUT_COLS / 8; ++j) dst[j] = make_uint4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy writes → async-proxy (TMA) reads
        __syncwarp();
        if (lane == 0) {
          const int out_col = EPI == G2_SWIGLU ? (b_blk * G2_BN + c0) / 2 : b_blk * G2_BN + c0;
          g2_tma_store_2d(&tmap_out, my_smem, out_col, row0);             // rows >= M are clipped by the tensor map
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) g2_remote_arrive(tempty_leader + 8u * as);           // leader's tempty[as]
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");        // stores complete before exit
  }

  tc::fence_before_sync();
  g2_cluster_sync();                                        // neither CTA frees TMEM / exits while the pair still uses it
  if (warp == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
}

int make_tmap_2d_bf16_plain(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld,
                            uint32_t box_cols, uint32_t box_rows);

}  // namespace jl

using namespace jl;

// 0 = launched; JL_ERR_UNSUPPORTED (3) = shape not covered by the pair kernel (caller takes the single-CTA kernel)
// clusters of this kernel that can be resident at once (0 until the first launch of that epilogue mode)
static int g2_pairs_used = 0;
extern "C" JL_API int jl_gemm_2cta_pairs(void) { return g2_pairs_used; }

extern "C" JL_API int jl_gemm_2cta(const void* x, const void* w, const void* residual, void* out, int M, int N, int K,
                                   int mode, int num_sms, void* stream) {
  if (M < 256 || N % G2_BN != 0 || K % G2_BK != 0 || mode < 0 || mode > 2) return JL_ERR_UNSUPPORTED;
  if (mode == G2_RESIDUAL && residual == nullptr) return JL_ERR_UNSUPPORTED;
  if (num_sms <= 0) num_sms = 148;
  const int odim = mode == G2_SWIGLU ? N / 2 : N;
  Gemm2Params p{};
  p.residual = (const __nv_bfloat16*)residual; p.M = M; p.N = N; p.K = K; p.ldo = odim;
  p.a_blocks = (M + 2 * G2_BM - 1) / (2 * G2_BM);
  p.b_blocks = N / G2_BN;
  CUtensorMap ta, tb, to;
  int rc;
  if ((rc = make_tmap_2d_bf16(&ta, x, M, K, K, G2_BM))) return rc;
  if ((rc = make_tmap_2d_bf16(&tb, w, N, K, K, G2_BN / 2))) return rc;
  if ((rc = make_tmap_2d_bf16_plain(&to, out, M, odim, odim, mode == G2_SWIGLU ? 32 : 64, 32))) return rc;
  int stages = (int)((200 * 1024) / G2_STAGE_BYTES);
  if (stages > 6) stages = 6;
  const int num_kb = K / G2_BK;
  if (stages > num_kb * 2) stages = num_kb * 2 < 2 ? 2 : num_kb * 2;
  p.stages = stages;
  const size_t smem = (size_t)stages * G2_STAGE_BYTES + G2_EPI_BYTES + 8 * (2 * stages + 4) + 16 + 1024;
  void (*kern)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const Gemm2Params) =
      mode == G2_SWIGLU ? gemm2_kernel<G2_SWIGLU> : (mode == G2_RESIDUAL ? gemm2_kernel<G2_RESIDUAL> : gemm2_kernel<G2_PLAIN>);
  JL_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(G2_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = (cudaStream_t)stream;
  // The tile loop is persistent, so every cluster of the grid must be resident at once: a pair that only starts when
  // another one has finished doubles the run time.  Not every TPC can host a 2-CTA cluster of this size on every part
  // (measured here: fewer than num_sms / 2), so ask the runtime how many fit.
  static int max_pairs[3] = {0, 0, 0};
  if (max_pairs[mode] == 0) {
    cfg.gridDim = dim3((num_sms / 2) * 2);
    cudaLaunchAttribute cattr[1];
    cattr[0].id = cudaLaunchAttributeClusterDimension;
    cattr[0].val.clusterDim.x = 2; cattr[0].val.clusterDim.y = 1; cattr[0].val.clusterDim.z = 1;
    cfg.attrs = cattr; cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess || n < 1) { cudaGetLastError(); n = num_sms / 2; }
    if (const char* e = getenv("JLLAMA_GEMM_PAIRS")) { const int v = atoi(e); if (v >= 1) n = v; }
    max_pairs[mode] = n < num_sms / 2 ? n : num_sms / 2;
  }
  int grid = p.a_blocks * p.b_blocks * 2;
  if (grid > 2 * max_pairs[mode]) grid = 2 * max_pairs[mode];
  cfg.gridDim = dim3(grid);
  g2_pairs_used = max_pairs[mode];
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  count_launch();
  JL_CUDA(cudaLaunchKernelEx(&cfg, kern, ta, tb, to, p));
  return JL_OK;
}
