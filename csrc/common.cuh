// Shared device/host helpers for the sm_100a kernel library (libjllama_b200.so).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define JL_OK 0
#define JL_ERR_ARG 1
#define JL_ERR_CUDA 2
#define JL_ERR_UNSUPPORTED 3

namespace jl {

// ---------------------------------------------------------------- host-side error plumbing
void set_error(const char* fmt, ...);          // error.cu
void count_launch();                           // error.cu: bumps the counter behind jl_launch_count()
bool pdl_enabled();                            // error.cu: false when JLLAMA_NO_PDL is set
int cuda_fail(cudaError_t e, const char* what); // records + returns JL_ERR_CUDA

#define JL_CHECK_ARG(cond, ...)                  \
  do {                                           \
    if (!(cond)) {                               \
      jl::set_error(__VA_ARGS__);                \
      return JL_ERR_ARG;                         \
    }                                            \
  } while (0)

#define JL_CUDA(expr)                                   \
  do {                                                  \
    cudaError_t _e = (expr);                            \
    if (_e != cudaSuccess) return jl::cuda_fail(_e, #expr); \
  } while (0)

// Launch with Programmatic Dependent Launch enabled: the kernel may start while its predecessor on
// the stream drains, and must execute pdl_wait() before touching anything the predecessor wrote.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                              cudaStream_t stream, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;      // JLLAMA_NO_PDL=1 turns the attribute off (A/B measurements)
  count_launch();
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ---------------------------------------------------------------- device helpers
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// bf16 pair packed in a 32-bit word → two fp32 (bf16 is the top half of an fp32)
__device__ __forceinline__ float bf16_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 p = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&p);
}

// streaming 16-byte load that does not pollute L1 (weights are touched exactly once per step)
__device__ __forceinline__ uint4 ldg_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
// ordinary cached 16-byte load (activations: small, re-read by every CTA)
__device__ __forceinline__ uint4 ldg_cached(const void* p) {
  return *reinterpret_cast<const uint4*>(p);
}

// device-scope acquire load (grid barriers, in-kernel ready counters)
__device__ __forceinline__ unsigned int ld_acquire_gpu_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// ---- system-scope flag protocol used by the NVLink peer-memory collectives (comm.cu, decode_mega.cu)
__device__ __forceinline__ void st_release_sys(int* p, int v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float ld_relaxed_sys_f32(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void wait_flag(const int* flag, int epoch, int rank, int peer, int what) {
  unsigned spins = 0;
  while (ld_acquire_sys(flag) < epoch) {
    if (++spins > (1u << 27)) {
      printf("jllama_b200: rank %d timed out waiting for peer %d (epoch %d, kind %d, block %d)\n", rank, peer,
             epoch, what, (int)blockIdx.x);
      __trap();
    }
  }
}

__device__ __forceinline__ float silu(float x) { return x / (1.0f + __expf(-x)); }

}  // namespace jl
