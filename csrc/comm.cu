// Tensor-parallel communication over NVLink 5 / NVSwitch peer memory, fused into the kernels that
// produce the data — no NCCL call on these paths.
//
// The reference has no explicit collective: GSPMD inserts NCCL all-reduces after the row-parallel `wo`
// and `w2` matmuls and an all-gather of the vocab-sharded logits (SURVEY §2.5, K2-K4;
// /root/reference/jax_llama/partition.py:62-78).  Here every rank maps every peer's "inbox" buffer
// (CUDA IPC) and:
//   * `gemv_ar_kernel`  — row-parallel linear at decode batch sizes + ONE-SHOT all-reduce + residual add:
//     each CTA computes a 32-row slice of the partial product, stores it (fp32) straight into every
//     peer's inbox, publishes a per-CTA flag (st.release.sys) and then waits only for the same-index CTA
//     of each peer (ld.acquire.sys) before reducing its slice in a fixed rank order — every rank gets
//     bit-identical sums, which keeps sampling in lock-step without a broadcast.
//   * `allgather_small_kernel` — tiny push all-gather (top-k candidates of the vocab-parallel head).
// Flags carry a monotonically increasing epoch (device counter + call index), never reset; inbox slots
// alternate by epoch parity, which is sufficient because a rank cannot start epoch e+2 before every
// peer has finished reading epoch e (it needs their epoch e+1 contribution first).
// Spin loops have a watchdog: a lost peer becomes a trap naming rank / peer / epoch, not a hang.
#include "common.cuh"
#include "jllama_b200.h"

namespace jl {

constexpr int MAX_TP = 8;
constexpr int AR_THREADS = 512;
constexpr int AR_ROWS = 32;          // rows per CTA slice = 16 warps x 2 rows
constexpr int AR_MAX_M = 8;

struct PeerTable {
  float* inbox[MAX_TP];              // peer p's inbox base, mapped into this process
  int* flags[MAX_TP];                // peer p's flag array base
};

struct GemvArParams {
  const __nv_bfloat16* x;         // [M, K] local activation slice (K = K_total / tp)
  const __nv_bfloat16* w;         // [N, K] local weight slice
  const __nv_bfloat16* residual;  // [M, N] (replicated); may alias out
  __nv_bfloat16* out;             // [M, N]
  const int* seq_ptr;             // device epoch counter (bumped once per step / forward)
  int call_index;                 // position of this collective since the last bump
  int N, K, tp, rank;
  size_t inbox_parity_stride;     // floats between the two parity halves of an inbox
  int flag_offset;                // first flag of this call site's flag block
  PeerTable peers;
};

// inbox layout (floats): [parity][src_rank][m][N]
template <int M>
__global__ void __launch_bounds__(AR_THREADS, 1) gemv_ar_kernel(const GemvArParams p) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  uint4* xs = reinterpret_cast<uint4*>(smem_raw);          // [M][K/8]
  __shared__ float part[M][AR_ROWS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int K8 = p.K >> 3;
  const int n_blocks = (p.N + AR_ROWS - 1) / AR_ROWS;
  const uint4* wbase = reinterpret_cast<const uint4*>(p.w);

  // first weight chunk of the first slice is independent of the previous kernel: issue before the wait
  const int row_first = blockIdx.x * AR_ROWS + warp * 2;
  uint4 pre[2];
#pragma unroll
  for (int r = 0; r < 2; ++r)
    pre[r] = (row_first + r < p.N && lane < K8) ? ldg_stream(wbase + (size_t)(row_first + r) * K8 + lane)
                                                 : make_uint4(0, 0, 0, 0);
  pdl_wait();
  pdl_launch_dependents();

  const int epoch = *p.seq_ptr + p.call_index + 1;
  const int parity = epoch & 1;
  for (int i = threadIdx.x; i < M * K8; i += AR_THREADS) xs[i] = ldg_cached(reinterpret_cast<const uint4*>(p.x) + i);
  __syncthreads();

  for (int blk = blockIdx.x; blk < n_blocks; blk += gridDim.x) {
    const int row0 = blk * AR_ROWS + warp * 2;
    float acc[M][2];
#pragma unroll
    for (int m = 0; m < M; ++m) acc[m][0] = acc[m][1] = 0.f;
    for (int c = lane; c < K8; c += 32) {
      uint4 wv[2];
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        if (blk == blockIdx.x && c == lane) wv[r] = pre[r];
        else wv[r] = row0 + r < p.N ? ldg_stream(wbase + (size_t)(row0 + r) * K8 + c) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int m = 0; m < M; ++m) {
        const uint4 a = xs[m * K8 + c];
        const float xf[8] = {bf16_lo(a.x), bf16_hi(a.x), bf16_lo(a.y), bf16_hi(a.y),
                             bf16_lo(a.z), bf16_hi(a.z), bf16_lo(a.w), bf16_hi(a.w)};
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          const float wf[8] = {bf16_lo(wv[r].x), bf16_hi(wv[r].x), bf16_lo(wv[r].y), bf16_hi(wv[r].y),
                               bf16_lo(wv[r].z), bf16_hi(wv[r].z), bf16_lo(wv[r].w), bf16_hi(wv[r].w)};
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[m][r] = fmaf(wf[e], xf[e], acc[m][r]);
        }
      }
    }
#pragma unroll
    for (int m = 0; m < M; ++m)
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const float s = warp_sum(acc[m][r]);
        if (lane == 0) part[m][warp * 2 + r] = s;
      }
    __syncthreads();
    // ---- push this slice to every peer's inbox: warp w serves (peer, m) pairs, lanes = 32 rows → 128 B stores
    const int col = blk * AR_ROWS + lane;
    for (int pm = warp; pm < p.tp * M; pm += AR_THREADS / 32) {
      const int peer = pm / M, m = pm - peer * M;
      if (col < p.N)
        p.peers.inbox[peer][parity * p.inbox_parity_stride + ((size_t)p.rank * M + m) * p.N + col] = part[m][lane];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < p.tp)
      st_release_sys(p.peers.flags[threadIdx.x] + p.flag_offset + p.rank * n_blocks + blk, epoch);
    // ---- wait for the same slice from every peer, reduce in rank order, add the residual
    if (threadIdx.x < p.tp)
      wait_flag(p.peers.flags[p.rank] + p.flag_offset + threadIdx.x * n_blocks + blk, epoch, p.rank, threadIdx.x, 0);
    __syncthreads();
    const float* mine = p.peers.inbox[p.rank] + parity * p.inbox_parity_stride;
    for (int i = threadIdx.x; i < M * AR_ROWS; i += AR_THREADS) {
      const int m = i / AR_ROWS, c = blk * AR_ROWS + (i - m * AR_ROWS);
      if (c < p.N) {
        float s = 0.f;
        for (int src = 0; src < p.tp; ++src) s += ld_relaxed_sys_f32(mine + ((size_t)src * M + m) * p.N + c);
        const float y = __bfloat162float(__float2bfloat16(s)) + __bfloat162float(p.residual[(size_t)m * p.N + c]);
        p.out[(size_t)m * p.N + c] = __float2bfloat16(y);
      }
    }
    __syncthreads();
  }
}

// Push all-gather of a small buffer: out[src] (n_bytes each) on every rank.  One CTA.
struct GatherParams {
  const uint8_t* src;
  uint8_t* inbox[MAX_TP];      // peer inbox bases (bytes); layout [parity][src_rank][slot_bytes]
  int* flags[MAX_TP];
  const int* seq_ptr;
  int call_index, tp, rank, n_bytes, slot_bytes, flag_offset;
  size_t parity_stride;
  uint8_t* out;                // local [tp][n_bytes] (may be null: consumers read the inbox directly)
};

__global__ void __launch_bounds__(256) allgather_small_kernel(const GatherParams p) {
  pdl_wait();
  pdl_launch_dependents();
  const int epoch = *p.seq_ptr + p.call_index + 1;
  const int parity = epoch & 1;
  const int n16 = p.n_bytes / 16;
  for (int peer = 0; peer < p.tp; ++peer) {
    uint4* dst = reinterpret_cast<uint4*>(p.inbox[peer] + parity * p.parity_stride + (size_t)p.rank * p.slot_bytes);
    for (int i = threadIdx.x; i < n16; i += blockDim.x) dst[i] = reinterpret_cast<const uint4*>(p.src)[i];
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x < p.tp) {
    st_release_sys(p.flags[threadIdx.x] + p.flag_offset + p.rank, epoch);
    wait_flag(p.flags[p.rank] + p.flag_offset + threadIdx.x, epoch, p.rank, threadIdx.x, 1);
  }
  __syncthreads();
  if (p.out) {
    const uint8_t* mine = p.inbox[p.rank] + parity * p.parity_stride;
    for (int s = 0; s < p.tp; ++s)
      for (int i = threadIdx.x; i < n16; i += blockDim.x) {
        uint4 v;
        const uint4* a = reinterpret_cast<const uint4*>(mine + (size_t)s * p.slot_bytes) + i;
        asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(a) : "memory");
        reinterpret_cast<uint4*>(p.out + (size_t)s * p.n_bytes)[i] = v;
      }
  }
}

// stream-side wait: later kernels on this stream may read what the peers published before raising their flags
// (No early `launch_dependents` in the two waiting kernels: a GEMM launched programmatically behind them would take every
// SM's shared memory and then sit in its dependency wait while this kernel spins — and the kernel this one waits for,
// on another stream of a peer whose GPU is in the same state, could never be scheduled.  Measured: deadlock on 8B.)
__global__ void wait_flags_kernel(const int* flags, int epoch, const int* seq_ptr, int tp, int rank) {
  pdl_wait();
  if (threadIdx.x < tp) wait_flag(flags + threadIdx.x, epoch + (seq_ptr ? *seq_ptr : 0), rank, threadIdx.x, 6);
}

// publish this rank's flag to every rank, then wait for everybody's: one warp.  Kernels that move data never spin on
// a peer themselves (a kernel that fills the SMs while it waits could keep the very kernel it waits for — on another
// stream of a peer — from being scheduled); the waiting is done by these one-warp kernels between them.
__global__ void flag_exchange_kernel(PeerTable peers, int flag_off, int epoch, const int* base, int tp, int rank) {
  pdl_wait();
  if (base != nullptr) epoch += *base;         // device epoch base: the same captured launch is right on every graph replay
  if (threadIdx.x < tp) {
    __threadfence_system();
    st_release_sys(peers.flags[threadIdx.x] + flag_off + rank, epoch);
    wait_flag(peers.flags[rank] + flag_off + threadIdx.x, epoch, rank, threadIdx.x, 7);
  }
}

__global__ void bump_kernel(int* seq, int n) {
  pdl_wait();
  pdl_launch_dependents();
  if (threadIdx.x == 0) *seq += n;
}

}  // namespace jl

using namespace jl;

// ------------------------------------------------------------------------------------ IPC plumbing
extern "C" JL_API int jl_ipc_alloc(size_t bytes, void** ptr, void* handle64) {
  JL_CUDA(cudaMalloc(ptr, bytes));
  JL_CUDA(cudaMemset(*ptr, 0, bytes));
  JL_CUDA(cudaDeviceSynchronize());
  cudaIpcMemHandle_t h;
  JL_CUDA(cudaIpcGetMemHandle(&h, *ptr));
  static_assert(sizeof(h) == 64, "unexpected IPC handle size");
  memcpy(handle64, &h, 64);
  return JL_OK;
}

extern "C" JL_API int jl_ipc_open(const void* handle64, void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  JL_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return JL_OK;
}

extern "C" JL_API int jl_ipc_close(void* ptr) {
  JL_CUDA(cudaIpcCloseMemHandle(ptr));
  return JL_OK;
}

extern "C" JL_API int jl_ipc_free(void* ptr) {
  JL_CUDA(cudaFree(ptr));
  return JL_OK;
}

// ------------------------------------------------------------------------------------ fused collectives
extern "C" JL_API int jl_gemv_allreduce(const void* x, const void* w, const void* residual, void* out, int M, int N,
                                        int K, int tp, int rank, void* const* peer_inbox, void* const* peer_flags,
                                        size_t inbox_parity_stride_floats, int flag_offset, const int* seq_ptr,
                                        int call_index, int num_sms, void* stream) {
  JL_CHECK_ARG(M >= 1 && M <= AR_MAX_M, "jl_gemv_allreduce: M=%d outside 1..8", M);
  JL_CHECK_ARG(tp >= 1 && tp <= MAX_TP && rank >= 0 && rank < tp, "jl_gemv_allreduce: bad tp/rank %d/%d", tp, rank);
  JL_CHECK_ARG(K % 8 == 0, "jl_gemv_allreduce: K=%d must be a multiple of 8", K);
  JL_CHECK_ARG((size_t)tp * M * N <= inbox_parity_stride_floats, "jl_gemv_allreduce: inbox too small");
  const size_t smem = (size_t)M * K * 2;
  JL_CHECK_ARG(smem <= 200 * 1024, "jl_gemv_allreduce: M*K does not fit in shared memory");
  GemvArParams p{};
  p.x = (const __nv_bfloat16*)x; p.w = (const __nv_bfloat16*)w; p.residual = (const __nv_bfloat16*)residual;
  p.out = (__nv_bfloat16*)out; p.seq_ptr = seq_ptr; p.call_index = call_index;
  p.N = N; p.K = K; p.tp = tp; p.rank = rank;
  p.inbox_parity_stride = inbox_parity_stride_floats; p.flag_offset = flag_offset;
  for (int i = 0; i < tp; ++i) { p.peers.inbox[i] = (float*)peer_inbox[i]; p.peers.flags[i] = (int*)peer_flags[i]; }
  if (num_sms <= 0) num_sms = 148;
  int grid = (N + AR_ROWS - 1) / AR_ROWS;
  if (grid > num_sms) grid = num_sms;
  cudaError_t e;
#define JL_AR(M_)                                                                                      \
  case M_: {                                                                                           \
    auto k = gemv_ar_kernel<M_>;                                                                       \
    e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);               \
    if (e == cudaSuccess) e = launch_pdl(k, dim3(grid), dim3(AR_THREADS), smem, (cudaStream_t)stream, p); \
  } break;
  switch (M) { JL_AR(1) JL_AR(2) JL_AR(3) JL_AR(4) JL_AR(5) JL_AR(6) JL_AR(7) default: JL_AR(8) }
#undef JL_AR
  JL_CUDA(e);
  return JL_OK;
}

extern "C" JL_API int jl_gemv_allreduce_flags(int N) { return MAX_TP * ((N + AR_ROWS - 1) / AR_ROWS); }

extern "C" JL_API int jl_allgather_small(const void* src, void* out, int n_bytes, int slot_bytes, int tp, int rank,
                                         void* const* peer_inbox, void* const* peer_flags,
                                         size_t parity_stride_bytes, int flag_offset, const int* seq_ptr,
                                         int call_index, void* stream) {
  JL_CHECK_ARG(n_bytes % 16 == 0 && slot_bytes % 16 == 0 && n_bytes <= slot_bytes, "jl_allgather_small: sizes must be multiples of 16");
  JL_CHECK_ARG(tp >= 1 && tp <= MAX_TP && rank >= 0 && rank < tp, "jl_allgather_small: bad tp/rank");
  JL_CHECK_ARG((size_t)tp * slot_bytes <= parity_stride_bytes, "jl_allgather_small: inbox too small");
  GatherParams p{};
  p.src = (const uint8_t*)src; p.out = (uint8_t*)out; p.seq_ptr = seq_ptr; p.call_index = call_index;
  p.tp = tp; p.rank = rank; p.n_bytes = n_bytes; p.slot_bytes = slot_bytes; p.flag_offset = flag_offset;
  p.parity_stride = parity_stride_bytes;
  for (int i = 0; i < tp; ++i) { p.inbox[i] = (uint8_t*)peer_inbox[i]; p.flags[i] = (int*)peer_flags[i]; }
  JL_CUDA(launch_pdl(allgather_small_kernel, dim3(1), dim3(256), 0, (cudaStream_t)stream, p));
  return JL_OK;
}

extern "C" JL_API int jl_wait_flags(const int* flags, int epoch, const int* seq_ptr, int tp, int rank, void* stream) {
  JL_CHECK_ARG(flags && tp >= 1 && tp <= MAX_TP, "jl_wait_flags: bad arguments");
  JL_CUDA(launch_pdl(wait_flags_kernel, dim3(1), dim3(32), 0, (cudaStream_t)stream, flags, epoch, seq_ptr, tp, rank));
  return JL_OK;
}

extern "C" JL_API int jl_flag_exchange(void* const* peer_flags, int flag_off, int epoch, const int* base, int tp, int rank,
                                       void* stream) {
  JL_CHECK_ARG(peer_flags && tp >= 1 && tp <= MAX_TP && rank >= 0 && rank < tp, "jl_flag_exchange: bad arguments");
  PeerTable t{};
  for (int i = 0; i < tp; ++i) t.flags[i] = (int*)peer_flags[i];
  JL_CUDA(launch_pdl(flag_exchange_kernel, dim3(1), dim3(32), 0, (cudaStream_t)stream, t, flag_off, epoch, base, tp, rank));
  return JL_OK;
}

extern "C" JL_API int jl_comm_bump(int* seq_ptr, int n, void* stream) {
  JL_CUDA(launch_pdl(bump_kernel, dim3(1), dim3(32), 0, (cudaStream_t)stream, seq_ptr, n));
  return JL_OK;
}

extern "C" JL_API int jl_memcpy_d2d(void* dst, const void* src, size_t bytes, void* stream) {
  JL_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return JL_OK;
}
