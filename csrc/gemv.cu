// Small-batch linear layer (decode, M <= 8 tokens): y[M,N] = prologue(x)[M,K] · W[N,K]^T with fused
// RMSNorm prologue and residual / SwiGLU epilogues.  Weight-streaming bound: every weight byte is read
// from HBM exactly once per step with 16-byte non-allocating loads, 8 loads in flight per lane, and the
// first batch is issued *before* griddepcontrol.wait so that under Programmatic Dependent Launch the
// HBM stream of layer i+1 starts while layer i drains.  This SIMT form is instruction-latency bound at
// ~5-6 TB/s (profiles/README.md); the whole-step kernel (decode_mega_mma.cuh) does the same GEMV with ldmatrix +
// mma.sync and reaches the HBM roofline.  M > 8 goes to the tcgen05 swap-AB kernel in gemm_tcgen05.cu.
//
// Covers the reference's wq/wk/wv (fused), wo, w1/w3 (fused, interleaved), w2 and lm_head at decode
// (/root/reference/jax_llama/model.py:119-150,311-338,699-706).
#include "common.cuh"
#include "jllama_b200.h"

namespace jl {

constexpr int GEMV_THREADS = 512;   // 16 warps, one persistent CTA per SM
// rows per warp work-unit R (2 = a SwiGLU pair, 1 = finer balance) is a template parameter; each lane keeps
// R·U = 8 16-byte loads in flight per step

struct GemvParams {
  const __nv_bfloat16* x;         // [M, K]
  const __nv_bfloat16* w;         // [N, K]
  const __nv_bfloat16* norm_w;    // [K] or null: fused RMSNorm on x
  const __nv_bfloat16* residual;  // [M, N] or null (mode RESIDUAL)
  void* out;                      // bf16 or fp32; [M, N] or [M, N/2] for SWIGLU
  int N, K;
  float eps;
};

enum { MODE_PLAIN = 0, MODE_RESIDUAL = 1, MODE_SWIGLU = 2 };

template <int M, int MODE, bool OUT_F32, int R>
__global__ void __launch_bounds__(GEMV_THREADS, 1) gemv_kernel(const GemvParams p) {
  constexpr int U = 8 / R;        // 16-byte loads per row per step: R·U = 8 loads in flight per lane
  extern __shared__ __align__(16) uint8_t smem_raw[];
  uint4* xs = reinterpret_cast<uint4*>(smem_raw);             // [M][K/8] normalised activations (bf16)
  __shared__ float red[M][GEMV_THREADS / 32];

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int K8 = p.K >> 3;
  const int steps_per_unit = (K8 + 32 * U - 1) / (32 * U);
  const int n_units = (p.N + R - 1) / R;
  const int total_warps = gridDim.x * (GEMV_THREADS / 32);
  const int gw = blockIdx.x * (GEMV_THREADS / 32) + warp;     // warps of one CTA take adjacent units
  const uint4* wbase = reinterpret_cast<const uint4*>(p.w);

  // flattened (unit, step) iteration for this warp, software-pipelined one step ahead
  int unit = gw, step = 0;
  uint4 cur[R][U], nxt[R][U];
  auto load = [&](uint4 (&dst)[R][U], int u, int s) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int row = u * R + r;
#pragma unroll
      for (int j = 0; j < U; ++j) {
        const int c = (s * U + j) * 32 + lane;
        if (row < p.N && c < K8) dst[r][j] = ldg_stream(wbase + (size_t)row * K8 + c);
        else dst[r][j] = make_uint4(0, 0, 0, 0);
      }
    }
  };
  if (unit < n_units) load(cur, unit, 0);      // weights do not depend on the previous kernel

  pdl_wait();
  pdl_launch_dependents();

  // ---- prologue: stage x (optionally RMS-normalised, rounded to bf16 like the unfused op) in smem
  {
    const uint4* xg = reinterpret_cast<const uint4*>(p.x);
    if (p.norm_w != nullptr) {
      float ss[M];
#pragma unroll
      for (int m = 0; m < M; ++m) ss[m] = 0.f;
      for (int i = threadIdx.x; i < K8; i += GEMV_THREADS) {
#pragma unroll
        for (int m = 0; m < M; ++m) {
          uint4 a = ldg_cached(xg + (size_t)m * K8 + i);
          uint32_t aw[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float lo = bf16_lo(aw[j]), hi = bf16_hi(aw[j]);
            ss[m] += lo * lo + hi * hi;
          }
        }
      }
#pragma unroll
      for (int m = 0; m < M; ++m) {
        float t = warp_sum(ss[m]);
        if (lane == 0) red[m][warp] = t;
      }
      __syncthreads();
      float rinv[M];
#pragma unroll
      for (int m = 0; m < M; ++m) {
        float t = lane < GEMV_THREADS / 32 ? red[m][lane] : 0.f;
        t = warp_sum(t);
        rinv[m] = rsqrtf(t / (float)p.K + p.eps);
      }
      const uint4* gg = reinterpret_cast<const uint4*>(p.norm_w);
      for (int i = threadIdx.x; i < K8; i += GEMV_THREADS) {
        uint4 g = __ldg(gg + i);
        uint32_t gwd[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
        for (int m = 0; m < M; ++m) {
          uint4 a = ldg_cached(xg + (size_t)m * K8 + i);
          uint32_t aw[4] = {a.x, a.y, a.z, a.w}, o[4];
#pragma unroll
          for (int j = 0; j < 4; ++j)
            o[j] = pack_bf16(bf16_lo(aw[j]) * rinv[m] * bf16_lo(gwd[j]),
                             bf16_hi(aw[j]) * rinv[m] * bf16_hi(gwd[j]));
          xs[m * K8 + i] = make_uint4(o[0], o[1], o[2], o[3]);
        }
      }
    } else {
      for (int i = threadIdx.x; i < M * K8; i += GEMV_THREADS) xs[i] = ldg_cached(xg + i);
    }
  }
  __syncthreads();

  float acc[M][R];
#pragma unroll
  for (int m = 0; m < M; ++m)
#pragma unroll
    for (int r = 0; r < R; ++r) acc[m][r] = 0.f;

  while (unit < n_units) {
    // prefetch the following step (possibly of the next unit) before consuming the current one
    int nunit = unit, nstep = step + 1;
    if (nstep == steps_per_unit) { nstep = 0; nunit = unit + total_warps; }
    if (nunit < n_units) load(nxt, nunit, nstep);

#pragma unroll
    for (int j = 0; j < U; ++j) {
      const int c = (step * U + j) * 32 + lane;
      if (c < K8) {
        float wf[R][8];
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const uint4 wv = cur[r][j];
          wf[r][0] = bf16_lo(wv.x); wf[r][1] = bf16_hi(wv.x); wf[r][2] = bf16_lo(wv.y); wf[r][3] = bf16_hi(wv.y);
          wf[r][4] = bf16_lo(wv.z); wf[r][5] = bf16_hi(wv.z); wf[r][6] = bf16_lo(wv.w); wf[r][7] = bf16_hi(wv.w);
        }
#pragma unroll
        for (int m = 0; m < M; ++m) {
          const uint4 a = xs[m * K8 + c];
          const float xf[8] = {bf16_lo(a.x), bf16_hi(a.x), bf16_lo(a.y), bf16_hi(a.y),
                               bf16_lo(a.z), bf16_hi(a.z), bf16_lo(a.w), bf16_hi(a.w)};
#pragma unroll
          for (int r = 0; r < R; ++r)
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[m][r] = fmaf(wf[r][e], xf[e], acc[m][r]);
        }
      }
    }

    if (step == steps_per_unit - 1) {
      // ---- unit finished: reduce across the warp, apply the epilogue, write
#pragma unroll
      for (int m = 0; m < M; ++m)
#pragma unroll
        for (int r = 0; r < R; ++r) acc[m][r] = warp_sum(acc[m][r]);
      if (lane == 0) {
        const int row0 = unit * R;
#pragma unroll
        for (int m = 0; m < M; ++m) {
          if constexpr (MODE == MODE_SWIGLU) {
            // rows (2j, 2j+1) hold (w1_j, w3_j): emit silu(gate)·up; round the two GEMM outputs to bf16
            // first so the result is bit-compatible with the unfused GEMM → SwiGLU sequence
            float g = __bfloat162float(__float2bfloat16(acc[m][0]));
            float u = __bfloat162float(__float2bfloat16(acc[m][1]));
            reinterpret_cast<__nv_bfloat16*>(p.out)[(size_t)m * (p.N / 2) + unit] =
                __float2bfloat16(silu(g) * u);
          } else {
#pragma unroll
            for (int r = 0; r < R; ++r) {
              const int row = row0 + r;
              if (row < p.N) {
                float v = acc[m][r];
                if (MODE == MODE_RESIDUAL)
                  v = __bfloat162float(__float2bfloat16(v)) +
                      __bfloat162float(p.residual[(size_t)m * p.N + row]);
                if (OUT_F32) reinterpret_cast<float*>(p.out)[(size_t)m * p.N + row] = v;
                else reinterpret_cast<__nv_bfloat16*>(p.out)[(size_t)m * p.N + row] = __float2bfloat16(v);
              }
            }
          }
        }
      }
#pragma unroll
      for (int m = 0; m < M; ++m)
#pragma unroll
        for (int r = 0; r < R; ++r) acc[m][r] = 0.f;
    }
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
      for (int j = 0; j < U; ++j) cur[r][j] = nxt[r][j];
    unit = nunit;
    step = nstep;
  }
}

template <int M, int R>
static cudaError_t gemv_launch(const GemvParams& p, int mode, bool out_f32, int grid, size_t smem,
                               cudaStream_t s) {
#define JL_GEMV_CASE(MODE_, F32_)                                                              \
  {                                                                                            \
    auto k = gemv_kernel<M, MODE_, F32_, R>;                                                   \
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    if (e != cudaSuccess) return e;                                                            \
    return launch_pdl(k, dim3(grid), dim3(GEMV_THREADS), smem, s, p);                          \
  }
  if (mode == MODE_PLAIN && !out_f32) JL_GEMV_CASE(MODE_PLAIN, false)
  if (mode == MODE_PLAIN && out_f32) JL_GEMV_CASE(MODE_PLAIN, true)
  if (mode == MODE_RESIDUAL) JL_GEMV_CASE(MODE_RESIDUAL, false)
  if constexpr (R == 2) JL_GEMV_CASE(MODE_SWIGLU, false)
  return cudaErrorInvalidValue;
#undef JL_GEMV_CASE
}

// Work-unit shape and grid size.  The kernel is HBM-bound, so what matters is (a) every resident warp
// gets the same number of units — a warp with one unit more than the others sets the kernel time — and
// (b) enough bytes stay in flight.  Pick rows-per-unit R ∈ {2, 1} and a grid in [3/4·SMs, SMs] that
// maximise  units / (waves · grid · warps_per_cta).
static void gemv_plan(int N, int mode, int num_sms, int* r_out, int* grid_out) {
  const int wpc = GEMV_THREADS / 32;
  double best = -1.0;
  int best_r = 2, best_g = num_sms;
  for (int r = 2; r >= 1; --r) {
    if (mode == MODE_SWIGLU && r != 2) break;
    const int units = (N + r - 1) / r;
    if (units <= (num_sms * 3 / 4) * wpc) {            // small layer: one unit per warp, as many CTAs as needed
      const double eff = 1.0 - 0.01 * (2 - r);         // prefer R = 2 on ties
      if (eff > best) { best = eff; best_r = r; best_g = (units + wpc - 1) / wpc; }
      continue;
    }
    for (int g = num_sms; g >= num_sms * 3 / 4; --g) {
      const int slots = g * wpc;
      const int waves = (units + slots - 1) / slots;
      const double eff = (double)units / ((double)waves * slots) * (0.97 + 0.03 * g / num_sms) - 0.005 * (2 - r);
      if (eff > best) { best = eff; best_r = r; best_g = g; }
    }
  }
  *r_out = best_r;
  *grid_out = best_g < 1 ? 1 : best_g;
}

}  // namespace jl

using namespace jl;

extern "C" JL_API int jl_gemv(const void* x, const void* w, const void* norm_w, const void* residual, void* out,
                       int M, int N, int K, float eps, int mode, int out_f32, int num_sms, void* stream) {
  JL_CHECK_ARG(M >= 1 && M <= 8, "jl_gemv: M=%d outside 1..8 (use the tcgen05 GEMM)", M);
  JL_CHECK_ARG(K % 8 == 0, "jl_gemv: K=%d must be a multiple of 8", K);
  JL_CHECK_ARG(mode >= 0 && mode <= 2, "jl_gemv: bad mode %d", mode);
  JL_CHECK_ARG(mode != MODE_SWIGLU || N % 2 == 0, "jl_gemv: SwiGLU needs an even N");
  JL_CHECK_ARG(mode != MODE_RESIDUAL || residual != nullptr, "jl_gemv: residual pointer missing");
  JL_CHECK_ARG(!(out_f32 && mode != MODE_PLAIN), "jl_gemv: fp32 output only in plain mode");
  const size_t smem = (size_t)M * K * 2;
  JL_CHECK_ARG(smem <= 220 * 1024, "jl_gemv: M*K=%d*%d does not fit in shared memory", M, K);
  GemvParams p{(const __nv_bfloat16*)x, (const __nv_bfloat16*)w, (const __nv_bfloat16*)norm_w,
               (const __nv_bfloat16*)residual, out, N, K, eps};
  if (num_sms <= 0) num_sms = 148;
  int r = 2, grid = num_sms;
  gemv_plan(N, mode, num_sms, &r, &grid);
  auto s = (cudaStream_t)stream;
  cudaError_t e;
#define JL_GEMV_M(M_)                                                                   \
  case M_:                                                                              \
    e = r == 2 ? gemv_launch<M_, 2>(p, mode, out_f32, grid, smem, s)                    \
               : gemv_launch<M_, 1>(p, mode, out_f32, grid, smem, s);                   \
    break;
  switch (M) {
    JL_GEMV_M(1) JL_GEMV_M(2) JL_GEMV_M(3) JL_GEMV_M(4) JL_GEMV_M(5) JL_GEMV_M(6) JL_GEMV_M(7)
    default: e = r == 2 ? gemv_launch<8, 2>(p, mode, out_f32, grid, smem, s)
                        : gemv_launch<8, 1>(p, mode, out_f32, grid, smem, s);
  }
#undef JL_GEMV_M
  JL_CUDA(e);
  return JL_OK;
}
