// Decode linear layer (M <= 4 tokens), second generation: the weight matrix is streamed by the TMA bulk-copy
// engine (cp.async.bulk global → shared, mbarrier completion) through a deep shared-memory ring, so the
// bytes in flight per SM are bounded by shared memory (≈ 100 KB) instead of by registers, and the stream
// starts BEFORE griddepcontrol.wait: under programmatic dependent launch the next layer's weights are
// already arriving while the previous kernel drains.  The CTA deliberately stays under half an SM's shared
// memory and registers so that this early launch can actually become resident next to its predecessor.
//
//   warp 0          producer: one 1 KB bulk copy per weight-row segment, 16 rows per stage
//   warps 1..8      consumers: warp w owns rows (2w, 2w+1) of the stage (a SwiGLU pair), reads weights and the
//                   (RMS-normalised) activations from shared memory, fp32 accumulate, warp-reduce, epilogue
//
// Same math / epilogues as gemv.cu (plain, residual add, SwiGLU, fp32 logits); that kernel remains the
// fallback for activations that do not fit next to the ring.
#include "jllama_b200.h"
#include "tcgen05.cuh"

namespace jl {

constexpr int GT_CONSUMERS = 8;
constexpr int GT_THREADS = (GT_CONSUMERS + 1) * 32;
constexpr int GT_ROWS = 2 * GT_CONSUMERS;          // rows per stage
constexpr int GT_KC = 512;                         // elements per row segment (1 KB)
constexpr int GT_SEG_BYTES = GT_KC * 2;
constexpr int GT_STAGE_BYTES = GT_ROWS * GT_SEG_BYTES;   // 16 KB

struct GemvTmaParams {
  const __nv_bfloat16* x;
  const __nv_bfloat16* w;
  const __nv_bfloat16* norm_w;
  const __nv_bfloat16* residual;
  void* out;
  int N, K, stages;
  float eps;
};

enum { GT_PLAIN = 0, GT_RESIDUAL = 1, GT_SWIGLU = 2 };

__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar, uint64_t hint) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
          "r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(hint)
      : "memory");
}
__device__ __forceinline__ void consumer_bar() { asm volatile("bar.sync 1, %0;" ::"n"(GT_CONSUMERS * 32) : "memory"); }

template <int M, int MODE, bool OUT_F32>
__global__ void __launch_bounds__(GT_THREADS, 2) gemv_tma_kernel(const GemvTmaParams p) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  const int K8 = p.K >> 3;
  const uint32_t x_bytes = ((uint32_t)M * p.K * 2 + 127u) & ~127u;
  uint4* xs = reinterpret_cast<uint4*>(smem_raw);
  const uint32_t ring = tc::smem_u32(smem_raw) + x_bytes;
  const uint32_t bars = ring + p.stages * GT_STAGE_BYTES;          // full[stages] | empty[stages]
  const uint8_t* ring_ptr = smem_raw + x_bytes;
  __shared__ float red[M][GT_CONSUMERS];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_items = (p.N + GT_ROWS - 1) / GT_ROWS;
  const int n_kc = (p.K + GT_KC - 1) / GT_KC;

  if (threadIdx.x == 0) {
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
    int stage = 0; uint32_t phase = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
      const int row0 = item * GT_ROWS;
      const int rows = p.N - row0 < GT_ROWS ? p.N - row0 : GT_ROWS;
      for (int kc = 0; kc < n_kc; ++kc) {
        const int k0 = kc * GT_KC;
        const uint32_t seg = (uint32_t)((p.K - k0 < GT_KC ? p.K - k0 : GT_KC) * 2);
        if (lane == 0) {
          tc::mbar_wait(bars + 8u * (p.stages + stage), phase ^ 1);
          tc::mbar_expect_tx(bars + 8u * stage, seg * rows);
        }
        __syncwarp();
        if (lane < rows)
          bulk_g2s(ring + stage * GT_STAGE_BYTES + lane * GT_SEG_BYTES,
                   p.w + (size_t)(row0 + lane) * p.K + k0, seg, bars + 8u * stage, tc::EVICT_FIRST);
        if (++stage == p.stages) { stage = 0; phase ^= 1; }
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
      const int row0 = item * GT_ROWS + cw * 2;
      float acc[M][2];
#pragma unroll
      for (int m = 0; m < M; ++m) acc[m][0] = acc[m][1] = 0.f;
      for (int kc = 0; kc < n_kc; ++kc) {
        tc::mbar_wait(bars + 8u * stage, phase);
        const int k0 = kc * GT_KC;
        const int seg16 = (p.K - k0 < GT_KC ? p.K - k0 : GT_KC) >> 3;          // 16-byte chunks in this segment
        if (row0 < p.N) {
          const uint4* w0 = reinterpret_cast<const uint4*>(ring_ptr + stage * GT_STAGE_BYTES + (cw * 2) * GT_SEG_BYTES);
          const uint4* w1 = w0 + GT_SEG_BYTES / 16;
#pragma unroll
          for (int j = 0; j < GT_KC / 8 / 32; ++j) {
            const int c = j * 32 + lane;
            if (c < seg16) {
              const uint4 a0 = w0[c], a1 = w1[c];
              const float f0[8] = {bf16_lo(a0.x), bf16_hi(a0.x), bf16_lo(a0.y), bf16_hi(a0.y),
                                   bf16_lo(a0.z), bf16_hi(a0.z), bf16_lo(a0.w), bf16_hi(a0.w)};
              const float f1[8] = {bf16_lo(a1.x), bf16_hi(a1.x), bf16_lo(a1.y), bf16_hi(a1.y),
                                   bf16_lo(a1.z), bf16_hi(a1.z), bf16_lo(a1.w), bf16_hi(a1.w)};
#pragma unroll
              for (int m = 0; m < M; ++m) {
                const uint4 xv = xs[m * K8 + (k0 >> 3) + c];
                const float xf[8] = {bf16_lo(xv.x), bf16_hi(xv.x), bf16_lo(xv.y), bf16_hi(xv.y),
                                     bf16_lo(xv.z), bf16_hi(xv.z), bf16_lo(xv.w), bf16_hi(xv.w)};
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                  acc[m][0] = fmaf(f0[e], xf[e], acc[m][0]);
                  acc[m][1] = fmaf(f1[e], xf[e], acc[m][1]);
                }
              }
            }
          }
        }
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(bars + 8u * (p.stages + stage));
        if (++stage == p.stages) { stage = 0; phase ^= 1; }
      }
      // ---- rows finished: reduce across the warp and write
#pragma unroll
      for (int m = 0; m < M; ++m) {
        acc[m][0] = warp_sum(acc[m][0]);
        acc[m][1] = warp_sum(acc[m][1]);
      }
      if (lane == 0 && row0 < p.N) {
#pragma unroll
        for (int m = 0; m < M; ++m) {
          if constexpr (MODE == GT_SWIGLU) {
            const float g = __bfloat162float(__float2bfloat16(acc[m][0]));
            const float u = __bfloat162float(__float2bfloat16(acc[m][1]));
            reinterpret_cast<__nv_bfloat16*>(p.out)[(size_t)m * (p.N / 2) + row0 / 2] = __float2bfloat16(silu(g) * u);
          } else {
#pragma unroll
            for (int r = 0; r < 2; ++r) {
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
static cudaError_t gemv_tma_launch(const GemvTmaParams& p, int mode, bool out_f32, int grid, size_t smem, cudaStream_t s) {
#define JL_GT_CASE(MODE_, F32_)                                                                  \
  {                                                                                              \
    auto k = gemv_tma_kernel<M, MODE_, F32_>;                                                    \
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    if (e != cudaSuccess) return e;                                                              \
    return launch_pdl(k, dim3(grid), dim3(GT_THREADS), smem, s, p);                              \
  }
  if (mode == GT_PLAIN && !out_f32) JL_GT_CASE(GT_PLAIN, false)
  if (mode == GT_PLAIN && out_f32) JL_GT_CASE(GT_PLAIN, true)
  if (mode == GT_RESIDUAL) JL_GT_CASE(GT_RESIDUAL, false)
  JL_GT_CASE(GT_SWIGLU, false)
#undef JL_GT_CASE
}

}  // namespace jl

using namespace jl;

// returns JL_ERR_UNSUPPORTED (without setting up anything) when the shape should use jl_gemv instead
extern "C" JL_API int jl_gemv_tma(const void* x, const void* w, const void* norm_w, const void* residual, void* out,
                                  int M, int N, int K, float eps, int mode, int out_f32, int num_sms, void* stream) {
  JL_CHECK_ARG(M >= 1 && M <= 4, "jl_gemv_tma: M=%d outside 1..4", M);
  JL_CHECK_ARG(K % 8 == 0, "jl_gemv_tma: K=%d must be a multiple of 8", K);
  JL_CHECK_ARG(mode >= 0 && mode <= 2, "jl_gemv_tma: bad mode %d", mode);
  JL_CHECK_ARG(mode != GT_SWIGLU || N % 2 == 0, "jl_gemv_tma: SwiGLU needs an even N");
  JL_CHECK_ARG(mode != GT_RESIDUAL || residual != nullptr, "jl_gemv_tma: residual pointer missing");
  JL_CHECK_ARG(!(out_f32 && mode != GT_PLAIN), "jl_gemv_tma: fp32 output only in plain mode");
  JL_CHECK_ARG((reinterpret_cast<uintptr_t>(w) & 15) == 0, "jl_gemv_tma: weight pointer must be 16-byte aligned");
  const size_t x_bytes = ((size_t)M * K * 2 + 127) & ~(size_t)127;
  // stay under ~half an SM so the next kernel's CTA can become resident early; never fewer than 3 stages
  size_t budget = 110 * 1024;
  int stages = x_bytes + 3 * GT_STAGE_BYTES <= budget ? (int)((budget - x_bytes) / GT_STAGE_BYTES) : 0;
  if (stages == 0) {
    budget = 216 * 1024;
    if (x_bytes + 3 * GT_STAGE_BYTES > budget) {
      set_error("jl_gemv_tma: activations (%zu bytes) leave no room for the weight ring", x_bytes);
      return JL_ERR_UNSUPPORTED;
    }
    stages = (int)((budget - x_bytes) / GT_STAGE_BYTES);
  }
  if (stages > 12) stages = 12;
  const size_t smem = x_bytes + (size_t)stages * GT_STAGE_BYTES + 16 * stages + 16;
  GemvTmaParams p{(const __nv_bfloat16*)x, (const __nv_bfloat16*)w, (const __nv_bfloat16*)norm_w,
                  (const __nv_bfloat16*)residual, out, N, K, stages, eps};
  if (num_sms <= 0) num_sms = 148;
  const int items = (N + GT_ROWS - 1) / GT_ROWS;
  int grid = items < num_sms ? items : num_sms;
  if (items > num_sms) {                      // balanced persistent grid: every CTA gets the same item count
    double best = -1.0;
    for (int g = num_sms; g >= num_sms * 3 / 4; --g) {
      const int waves = (items + g - 1) / g;
      const double eff = (double)items / ((double)waves * g) * (0.97 + 0.03 * g / num_sms);
      if (eff > best) { best = eff; grid = g; }
    }
  }
  cudaError_t e;
  auto s = (cudaStream_t)stream;
  switch (M) {
    case 1: e = gemv_tma_launch<1>(p, mode, out_f32, grid, smem, s); break;
    case 2: e = gemv_tma_launch<2>(p, mode, out_f32, grid, smem, s); break;
    case 3: e = gemv_tma_launch<3>(p, mode, out_f32, grid, smem, s); break;
    default: e = gemv_tma_launch<4>(p, mode, out_f32, grid, smem, s); break;
  }
  JL_CUDA(e);
  return JL_OK;
}
