// NVLS (NVLink SHARP) primitives: `multimem.*` on a multicast address that maps the same offset of every
// rank's symmetric buffer.  One store reaches every GPU of the group (`multimem.st`), one load returns the
// sum over all GPUs reduced inside the switch (`multimem.ld_reduce`), one reduction updates every copy
// (`multimem.red`).  The multicast mapping itself is created by parallel/symm.py.
//
// What these replace: the reference's collectives are whatever NCCL call GSPMD emits for the shardings of
// /root/reference/jax_llama/partition.py:62-78; here the same traffic is issued from inside the compute kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace jl {
namespace mm {

// one 8-byte store replicated by the switch into every rank's copy (tagged packet {value, tag})
__device__ __forceinline__ void st_b64(void* mc, uint32_t lo, uint32_t hi) {
  const uint64_t v = ((uint64_t)hi << 32) | lo;
  asm volatile("multimem.st.relaxed.sys.global.u64 [%0], %1;" ::"l"(mc), "l"(v) : "memory");
}
__device__ __forceinline__ void st_f32(void* mc, float v) {
  asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(mc), "f"(v) : "memory");
}
__device__ __forceinline__ void st_v4f32(void* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}
// 8 bf16 (16 bytes) replicated into every rank's copy
__device__ __forceinline__ void st_v4bf16x2(void* mc, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}
// sum over all ranks of 4 fp32 at the same offset, reduced in the switch
__device__ __forceinline__ float4 ld_reduce_v4f32(const void* mc) {
  float4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(mc)
               : "memory");
  return r;
}
// sum over all ranks of 8 bf16, accumulated in fp32 inside the switch, returned as bf16 pairs
__device__ __forceinline__ uint4 ld_reduce_v4bf16x2(const void* mc) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(mc)
               : "memory");
  return r;
}
__device__ __forceinline__ void red_add_f32(void* mc, float v) {
  asm volatile("multimem.red.relaxed.sys.global.add.f32 [%0], %1;" ::"l"(mc), "f"(v) : "memory");
}
// arrival counter on every rank's copy at once; release: this thread's earlier writes are visible first
__device__ __forceinline__ void red_release_add_u32(void* mc, uint32_t v) {
  asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" ::"l"(mc), "r"(v) : "memory");
}
__device__ __forceinline__ void red_relaxed_add_u32(void* mc, uint32_t v) {
  asm volatile("multimem.red.relaxed.sys.global.add.u32 [%0], %1;" ::"l"(mc), "r"(v) : "memory");
}

}  // namespace mm
}  // namespace jl
