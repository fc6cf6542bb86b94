// Stand-alone entry points over the NVLS primitives of nvls.cuh: used by the tests and by parallel/symm.py to
// prove, on the box, that the multicast mapping of a symmetric buffer works before any fused kernel relies on it.
// (The reference has no counterpart: its collectives are NCCL calls emitted by XLA.)
#include "common.cuh"
#include "jllama_b200.h"
#include "nvls.cuh"

namespace jl {

// op 0: out[i] = Σ_ranks buf[i]           (multimem.ld_reduce, fp32 x4)
// op 1: every rank's buf[i] = {src[i], tag} as 8-byte packets (multimem.st)
// op 2: every rank's buf[i] += src[i]     (multimem.red, fp32)
// op 3: out[i] = Σ_ranks buf[i]           (bf16 x8, fp32 accumulation in the switch)
// op 4: every rank's buf[i] = src[i]      (bf16 x8, multimem.st)
// op 5: every rank's counter buf[0] += 1  (multimem.red.release, u32)
__global__ void nvls_selftest_kernel(void* mc, const void* src, void* out, int n, int op, uint32_t tag) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  switch (op) {
    case 0:
      if (i * 4 < n) reinterpret_cast<float4*>(out)[i] = mm::ld_reduce_v4f32(reinterpret_cast<const float4*>(mc) + i);
      break;
    case 1:
      if (i < n) mm::st_b64(reinterpret_cast<uint2*>(mc) + i, __float_as_uint(reinterpret_cast<const float*>(src)[i]), tag);
      break;
    case 2:
      if (i < n) mm::red_add_f32(reinterpret_cast<float*>(mc) + i, reinterpret_cast<const float*>(src)[i]);
      break;
    case 3:
      if (i * 8 < n) reinterpret_cast<uint4*>(out)[i] = mm::ld_reduce_v4bf16x2(reinterpret_cast<const uint4*>(mc) + i);
      break;
    case 4:
      if (i * 8 < n) mm::st_v4bf16x2(reinterpret_cast<uint4*>(mc) + i, reinterpret_cast<const uint4*>(src)[i]);
      break;
    default:
      if (i == 0) mm::red_release_add_u32(mc, 1u);
      break;
  }
}

}  // namespace jl

using namespace jl;

extern "C" JL_API int jl_nvls_selftest(void* mc, const void* src, void* out, int n, int op, unsigned int tag, void* stream) {
  JL_CHECK_ARG(mc != nullptr && n > 0 && op >= 0 && op <= 5, "jl_nvls_selftest: bad arguments");
  const int threads = 256;
  const int blocks = (n + threads - 1) / threads;
  count_launch();
  nvls_selftest_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(mc, src, out, n, op, tag);
  JL_CUDA(cudaGetLastError());
  return JL_OK;
}
