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


// Every rank's [rows, cols] fp32 block → columns [col0, col0 + cols) of every rank's [rows, dst_ld] buffer: one
// multimem.st per 16 bytes (the switch replicates it), then the last CTA publishes this rank's "done" flag to all
// ranks.  Used to assemble the vocabulary-sharded prefill logits on every rank for the exact sampler.
struct BcastParams {
  const float* src; float* dst_mc;
  int rows, cols, src_ld, dst_ld, col0, tp, rank, flag_off, epoch;
  const int* seq_ptr;        // non-null: the flag value is *seq_ptr + epoch (device epoch counter, CUDA-graph safe)
  int* flags[8];
  int* ticket;
};
__global__ void __launch_bounds__(256) nvls_bcast_f32_kernel(const BcastParams p) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ int s_last;
  const int c4 = p.cols / 4, total = p.rows * c4;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int r = i / c4, c = (i - r * c4) * 4;
    const float4 v = *reinterpret_cast<const float4*>(p.src + (size_t)r * p.src_ld + c);
    mm::st_v4f32(p.dst_mc + (size_t)r * p.dst_ld + p.col0 + c, v);
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const int prev = atomicAdd(p.ticket, 1);
    s_last = prev == (int)gridDim.x - 1;
    if (s_last) *p.ticket = 0;
  }
  __syncthreads();
  if (s_last && threadIdx.x < p.tp) {
    __threadfence_system();
    st_release_sys(p.flags[threadIdx.x] + p.flag_off + p.rank, p.epoch + (p.seq_ptr ? *p.seq_ptr : 0));
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

extern "C" JL_API int jl_nvls_bcast_f32(const float* src, float* dst_mc, int rows, int cols, int src_ld, int dst_ld, int col0,
                                        int tp, int rank, void* const* peer_flags, int flag_off, int epoch, const int* seq_ptr,
                                        int* ticket, void* stream) {
  JL_CHECK_ARG(src && dst_mc && ticket && cols % 4 == 0 && col0 % 4 == 0 && src_ld % 4 == 0 && dst_ld % 4 == 0,
               "jl_nvls_bcast_f32: columns, offsets and pitches must be multiples of 4");
  JL_CHECK_ARG(tp >= 1 && tp <= 8 && rank >= 0 && rank < tp, "jl_nvls_bcast_f32: bad tp/rank");
  BcastParams p{};
  p.src = src; p.dst_mc = dst_mc; p.rows = rows; p.cols = cols; p.src_ld = src_ld; p.dst_ld = dst_ld; p.col0 = col0;
  p.tp = tp; p.rank = rank; p.flag_off = flag_off; p.epoch = epoch; p.seq_ptr = seq_ptr; p.ticket = ticket;
  for (int i = 0; i < tp; ++i) p.flags[i] = (int*)peer_flags[i];
  int blocks = (rows * (cols / 4) + 255) / 256;
  if (blocks > 148) blocks = 148;
  if (blocks < 1) blocks = 1;
  JL_CUDA(launch_pdl(nvls_bcast_f32_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, p));
  return JL_OK;
}
