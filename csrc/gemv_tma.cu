// Decode linear layer (M <= 4 tokens), second generation: the weight matrix is streamed by the TMA engine
// (cp.async.bulk.tensor, one 16 KB [32 rows x 256 cols] box per request, mbarrier completion) through a
// deep shared-memory ring, so the bytes in flight per SM are bounded by shared memory (≈ 100 KB) instead
// of by registers, and the stream starts BEFORE griddepcontrol.wait: under programmatic dependent launch
// the next layer's weights are already arriving while the previous kernel drains.  The CTA deliberately
// stays under half an SM's shared memory and registers so that this early launch can actually become
// resident next to its predecessor.
//
//   warp 0          producer: one elected lane issues the tensor-map loads
//   warps 1..8      consumers: warp w owns rows 4w..4w+3 of the stage (two SwiGLU pairs), reads weights and
//                   the (RMS-normalised) activations from shared memory, fp32 accumulate, warp-reduce, epilogue
//
// (A first version issued sixteen 1 KB cp.async.bulk copies per stage; it measured 3.3 TB/s on the LM head,
// i.e. ~87 cycles per request per SM — the TMA unit is request-rate bound at that size, hence the 16 KB boxes.)
// Same math / epilogues as gemv.cu (plain, residual add, SwiGLU, fp32 logits); that kernel remains the
// fallback for activations that do not fit next to the ring.
#include "jllama_b200.h"
#include "tcgen05.cuh"

namespace jl {

constexpr int GT_CONSUMERS = 8;
constexpr int GT_THREADS = (GT_CONSUMERS + 1) * 32;
constexpr int GT_RPW = 4;                          // rows per consumer warp per stage
constexpr int GT_ROWS = GT_RPW * GT_CONSUMERS;     // 32 rows per stage
constexpr int GT_KC = 256;                         // columns per stage (one 16-byte chunk per lane per row)
constexpr int GT_STAGE_BYTES = GT_ROWS * GT_KC * 2;   // 16 KB

struct GemvTmaParams {
  const __nv_bfloat16* x;
  const __nv_bfloat16* norm_w;
  const __nv_bfloat16* residual;
  void* out;
  int N, K, stages;
  float eps;
};

enum { GT_PLAIN = 0, GT_RESIDUAL = 1, GT_SWIGLU = 2 };

__device__ __forceinline__ void consumer_bar() { asm volatile("bar.sync 1, %0;" ::"n"(GT_CONSUMERS * 32) : "memory"); }

template <int M, int MODE, bool OUT_F32>
__global__ void __launch_bounds__(GT_THREADS, 2)
gemv_tma_kernel(const __grid_constant__ CUtensorMap tmap_w, const GemvTmaParams p) {
  extern __shared__ uint8_t smem_raw[];
  const int K8 = p.K >> 3;
  // ring first (TMA destinations want 128-byte alignment), activations behind it
  const uint32_t ring = (tc::smem_u32(smem_raw) + 127u) & ~127u;
  const uint8_t* ring_ptr = smem_raw + (ring - tc::smem_u32(smem_raw));
  const uint32_t bars = ring + p.stages * GT_STAGE_BYTES;          // full[stages] | empty[stages]
  uint4* xs = reinterpret_cast<uint4*>(const_cast<uint8_t*>(ring_ptr) + p.stages * GT_STAGE_BYTES + 16 * p.stages);
  __shared__ float red[M][GT_CONSUMERS];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_items = (p.N + GT_ROWS - 1) / GT_ROWS;
  const int n_kc = (p.K + GT_KC - 1) / GT_KC;

  if (threadIdx.x == 0) {
    tc::prefetch_tensormap(&tmap_w);
    for (int s = 0; s < p.stages; ++s) {
      tc::mbar_init(bars + 8u * s, 1);
      tc::mbar_init(bars + 8u * (p.stages + s), GT_CONSUMERS);
    }
    tc::fence_barrier_init();
  }
  __syncthreads();
  pdl_launch_dependents();

  if (warp == 0) {
    // ===================================================== producer: runs ahead of the dependency wait
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        for (int kc = 0; kc < n_kc; ++kc) {
          tc::mbar_wait(bars + 8u * (p.stages + stage), phase ^ 1);
          tc::mbar_expect_tx(bars + 8u * stage, GT_STAGE_BYTES);     // out-of-range rows/cols arrive as zeros
          tc::tma_load_2d(ring + stage * GT_STAGE_BYTES, &tmap_w, bars + 8u * stage, kc * GT_KC, item * GT_ROWS,
                          tc::EVICT_FIRST);
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // ===================================================== consumers
    const int cw = warp - 1;
    const int ct = threadIdx.x - 32;                          // 0 .. 255
    pdl_wait();                                               // activations come from the previous kernel
    {
      const uint4* xg = reinterpret_cast<const uint4*>(p.x);
      if (p.norm_w != nullptr) {
        float ss[M];
#pragma unroll
        for (int m = 0; m < M; ++m) ss[m] = 0.f;
        for (int i = ct; i < K8; i += GT_CONSUMERS * 32) {
#pragma unroll
          for (int m = 0; m < M; ++m) {
            const uint4 a = ldg_cached(xg + (size_t)m * K8 + i);
            const uint32_t aw[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float lo = bf16_lo(aw[j]), hi = bf16_hi(aw[j]);
              ss[m] += lo * lo + hi * hi;
            }
          }
        }
#pragma unroll
        for (int m = 0; m < M; ++m) {
          const float t = warp_sum(ss[m]);
          if (lane == 0) red[m][cw] = t;
        }
        consumer_bar();
        float rinv[M];
#pragma unroll
        for (int m = 0; m < M; ++m) {
          float t = lane < GT_CONSUMERS ? red[m][lane] : 0.f;
          t = warp_sum(t);
          rinv[m] = rsqrtf(t / (float)p.K + p.eps);
        }
        const uint4* gg = reinterpret_cast<const uint4*>(p.norm_w);
        for (int i = ct; i < K8; i += GT_CONSUMERS * 32) {
          const uint4 g = __ldg(gg + i);
          const uint32_t gw[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
          for (int m = 0; m < M; ++m) {
            const uint4 a = ldg_cached(xg + (size_t)m * K8 + i);
            const uint32_t aw[4] = {a.x, a.y, a.z, a.w};
            uint32_t o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j)
              o[j] = pack_bf16(bf16_lo(aw[j]) * rinv[m] * bf16_lo(gw[j]), bf16_hi(aw[j]) * rinv[m] * bf16_hi(gw[j]));
            xs[m * K8 + i] = make_uint4(o[0], o[1], o[2], o[3]);
          }
        }
      } else {
#pragma unroll 4
        for (int i = ct; i < M * K8; i += GT_CONSUMERS * 32) xs[i] = ldg_cached(xg + i);
      }
    }
    consumer_bar();

    int stage = 0; uint32_t phase = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
      const int row0 = item * GT_ROWS + cw * GT_RPW;
      float acc[M][GT_RPW];
#pragma unroll
      for (int m = 0; m < M; ++m)
#pragma unroll
        for (int r = 0; r < GT_RPW; ++r) acc[m][r] = 0.f;
      for (int kc = 0; kc < n_kc; ++kc) {
        tc::mbar_wait(bars + 8u * stage, phase);
        const int c = kc * (GT_KC / 8) + lane;                // this lane's 16-byte chunk along K
        if (row0 < p.N && c < K8) {
          const uint4* wr = reinterpret_cast<const uint4*>(ring_ptr + stage * GT_STAGE_BYTES) +
                            (cw * GT_RPW) * (GT_KC / 8) + lane;
          uint4 wv[GT_RPW];
#pragma unroll
          for (int r = 0; r < GT_RPW; ++r) wv[r] = wr[r * (GT_KC / 8)];
#pragma unroll
          for (int m = 0; m < M; ++m) {
            const uint4 xv = xs[m * K8 + c];
            const float xf[8] = {bf16_lo(xv.x), bf16_hi(xv.x), bf16_lo(xv.y), bf16_hi(xv.y),
                                 bf16_lo(xv.z), bf16_hi(xv.z), bf16_lo(xv.w), bf16_hi(xv.w)};
#pragma unroll
            for (int r = 0; r < GT_RPW; ++r) {
              const float wf[8] = {bf16_lo(wv[r].x), bf16_hi(wv[r].x), bf16_lo(wv[r].y), bf16_hi(wv[r].y),
                                   bf16_lo(wv[r].z), bf16_hi(wv[r].z), bf16_lo(wv[r].w), bf16_hi(wv[r].w)};
#pragma unroll
              for (int e = 0; e < 8; ++e) acc[m][r] = fmaf(wf[e], xf[e], acc[m][r]);
            }
          }
        }
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(bars + 8u * (p.stages + stage));
        if (++stage == p.stages) { stage = 0; phase ^= 1; }
      }
      // ---- rows finished: reduce across the warp and write
#pragma unroll
      for (int m = 0; m < M; ++m)
#pragma unroll
        for (int r = 0; r < GT_RPW; ++r) acc[m][r] = warp_sum(acc[m][r]);
      if (lane == 0 && row0 < p.N) {
#pragma unroll
        for (int m = 0; m < M; ++m) {
          if constexpr (MODE == GT_SWIGLU) {
#pragma unroll
            for (int pr = 0; pr < GT_RPW / 2; ++pr) {
              if (row0 + 2 * pr + 1 < p.N) {
                const float g = __bfloat162float(__float2bfloat16(acc[m][2 * pr]));
                const float u = __bfloat162float(__float2bfloat16(acc[m][2 * pr + 1]));
                reinterpret_cast<__nv_bfloat16*>(p.out)[(size_t)m * (p.N / 2) + row0 / 2 + pr] =
                    __float2bfloat16(silu(g) * u);
              }
            }
          } else {
#pragma unroll
            for (int r = 0; r < GT_RPW; ++r) {
              const int row = row0 + r;
              if (row < p.N) {
                float v = acc[m][r];
                if constexpr (MODE == GT_RESIDUAL)
                  v = __bfloat162float(__float2bfloat16(v)) + __bfloat162float(p.residual[(size_t)m * p.N + row]);
                if constexpr (OUT_F32) reinterpret_cast<float*>(p.out)[(size_t)m * p.N + row] = v;
                else reinterpret_cast<__nv_bfloat16*>(p.out)[(size_t)m * p.N + row] = __float2bfloat16(v);
              }
            }
          }
        }
      }
    }
  }
}

template <int M>
static cudaError_t gemv_tma_launch(const CUtensorMap& tm, const GemvTmaParams& p, int mode, bool out_f32, int grid,
                                   size_t smem, cudaStream_t s) {
#define JL_GT_CASE(MODE_, F32_)                                                                  \
  {                                                                                              \
    auto k = gemv_tma_kernel<M, MODE_, F32_>;                                                    \
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    if (e != cudaSuccess) return e;                                                              \
    return launch_pdl(k, dim3(grid), dim3(GT_THREADS), smem, s, tm, p);                          \
  }
  if (mode == GT_PLAIN && !out_f32) JL_GT_CASE(GT_PLAIN, false)
  if (mode == GT_PLAIN && out_f32) JL_GT_CASE(GT_PLAIN, true)
  if (mode == GT_RESIDUAL) JL_GT_CASE(GT_RESIDUAL, false)
  JL_GT_CASE(GT_SWIGLU, false)
#undef JL_GT_CASE
}

}  // namespace jl

using namespace jl;

// returns JL_ERR_UNSUPPORTED (nothing launched) when the shape should use jl_gemv instead
extern "C" JL_API int jl_gemv_tma(const void* x, const void* w, const void* norm_w, const void* residual, void* out,
                                  int M, int N, int K, float eps, int mode, int out_f32, int num_sms, void* stream) {
  JL_CHECK_ARG(M >= 1 && M <= 4, "jl_gemv_tma: M=%d outside 1..4", M);
  JL_CHECK_ARG(K % 8 == 0, "jl_gemv_tma: K=%d must be a multiple of 8", K);
  JL_CHECK_ARG(mode >= 0 && mode <= 2, "jl_gemv_tma: bad mode %d", mode);
  JL_CHECK_ARG(mode != GT_SWIGLU || N % 2 == 0, "jl_gemv_tma: SwiGLU needs an even N");
  JL_CHECK_ARG(mode != GT_RESIDUAL || residual != nullptr, "jl_gemv_tma: residual pointer missing");
  JL_CHECK_ARG(!(out_f32 && mode != GT_PLAIN), "jl_gemv_tma: fp32 output only in plain mode");
  const size_t x_bytes = (size_t)M * K * 2;
  // stay under ~half an SM so the next kernel's CTA can become resident early; never fewer than 3 stages
  size_t budget = 110 * 1024;
  if (x_bytes + 3 * GT_STAGE_BYTES + 512 > budget) budget = 216 * 1024;
  if (x_bytes + 3 * GT_STAGE_BYTES + 512 > budget) {
    set_error("jl_gemv_tma: activations (%zu bytes) leave no room for the weight ring", x_bytes);
    return JL_ERR_UNSUPPORTED;
  }
  int stages = (int)((budget - x_bytes - 512) / GT_STAGE_BYTES);
  if (stages > 12) stages = 12;
  const size_t smem = 128 + (size_t)stages * GT_STAGE_BYTES + 16 * stages + x_bytes + 16;
  CUtensorMap tm;
  int rc = make_tmap_2d_bf16_plain(&tm, w, (uint64_t)N, (uint64_t)K, (uint64_t)K, GT_KC, GT_ROWS);
  if (rc) return rc;
  GemvTmaParams p{(const __nv_bfloat16*)x, (const __nv_bfloat16*)norm_w, (const __nv_bfloat16*)residual, out, N, K,
                  stages, eps};
  if (num_sms <= 0) num_sms = 148;
  const int items = (N + GT_ROWS - 1) / GT_ROWS;
  int grid = items < num_sms ? items : num_sms;
  if (items > num_sms) {                      // balanced persistent grid: every CTA gets the same item count
    double best = -1.0;
    for (int g = num_sms; g >= num_sms / 2; --g) {
      const int waves = (items + g - 1) / g;
      const double eff = (double)items / ((double)waves * g) * (0.9 + 0.1 * g / num_sms);
      if (eff > best) { best = eff; grid = g; }
    }
  }
  cudaError_t e;
  auto s = (cudaStream_t)stream;
  switch (M) {
    case 1: e = gemv_tma_launch<1>(tm, p, mode, out_f32, grid, smem, s); break;
    case 2: e = gemv_tma_launch<2>(tm, p, mode, out_f32, grid, smem, s); break;
    case 3: e = gemv_tma_launch<3>(tm, p, mode, out_f32, grid, smem, s); break;
    default: e = gemv_tma_launch<4>(tm, p, mode, out_f32, grid, smem, s); break;
  }
  JL_CUDA(e);
  return JL_OK;
}
