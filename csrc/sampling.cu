// Token selection on device: argmax, exact top-k candidate extraction, temperature / top-k / top-p
// sampling over (possibly cross-rank) candidate lists, and the per-step sequence bookkeeping that keeps
// the decode loop free of host synchronisation.
//
// Reference behaviour: HF Flax logits warpers + `_sample/_greedy_search` as configured by
// /root/reference/jax_llama/generation.py:28-41 (temperature → top-k → top-p(min_keep=1) → categorical;
// finished rows emit pad; EOS marks a row finished).  The candidate representation is what makes the
// vocab-parallel head exact without gathering [B, V] logits: each rank extracts its local top-k
// (value, global index) pairs, ranks exchange tp·k pairs, and every rank runs the identical final
// selection with a counter-based RNG keyed by (seed, step, row).
#include "common.cuh"
#include "jllama_b200.h"

namespace jl {

constexpr int SM_THREADS = 1024;
constexpr int SURV_CAP = 2048;

__device__ __forceinline__ bool better(float va, int ia, float vb, int ib) {
  return va > vb || (va == vb && ia < ib);          // descending value, ties → lower index
}

// in-place bitonic sort (descending by `better`) of n = power of two pairs held in shared memory
__device__ void bitonic_desc(float* v, int* id, int n) {
  for (int size = 2; size <= n; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      __syncthreads();
      for (int t = threadIdx.x; t < n / 2; t += blockDim.x) {
        const int lo = 2 * t - (t & (stride - 1));
        const int hi = lo + stride;
        const bool desc = ((lo & size) == 0);
        const float a = v[lo], b = v[hi];
        const int ia = id[lo], ib = id[hi];
        const bool swap = desc ? better(b, ib, a, ia) : better(a, ia, b, ib);
        if (swap) { v[lo] = b; v[hi] = a; id[lo] = ib; id[hi] = ia; }
      }
    }
  }
  __syncthreads();
}

// ------------------------------------------------------------------------------------ argmax
// One CTA per row.  Writes (value, index + idx_offset) so that vocab shards can be merged.
__global__ void __launch_bounds__(SM_THREADS)
argmax_kernel(const float* __restrict__ logits, int V, int ld, int idx_offset, float* __restrict__ out_val,
              int* __restrict__ out_idx) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ float sv[32];
  __shared__ int si[32];
  const float* row = logits + (size_t)blockIdx.x * ld;
  float bv = -INFINITY;
  int bi = 0x7fffffff;
  for (int i = threadIdx.x; i < V; i += SM_THREADS) {
    const float x = row[i];
    if (better(x, i, bv, bi)) { bv = x; bi = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (better(ov, oi, bv, bi)) { bv = ov; bi = oi; }
  }
  if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = bv; si[threadIdx.x >> 5] = bi; }
  __syncthreads();
  if (threadIdx.x < 32) {
    bv = sv[threadIdx.x]; bi = si[threadIdx.x];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (better(ov, oi, bv, bi)) { bv = ov; bi = oi; }
    }
    if (threadIdx.x == 0) {
      if (out_val) out_val[blockIdx.x] = bv;
      out_idx[blockIdx.x] = (bi == 0x7fffffff ? 0 : bi) + idx_offset;
    }
  }
}

// ------------------------------------------------------------------------------------ top-k candidates
// Exact top-k (k <= 64... up to 256) of one row per CTA in two passes over the (L2-resident) row:
//   A. every thread keeps the max of its strided slice; the k-th largest of the 1024 thread maxima is
//      a lower bound L of the k-th largest element overall (at least k elements are >= L);
//   B. elements >= L (a handful more than k) are appended to a shared list, sorted, and the first k
//      are emitted in descending order as (value, global index).
__global__ void __launch_bounds__(SM_THREADS)
topk_kernel(const float* __restrict__ logits, int V, int ld, int idx_offset, int k,
            float* __restrict__ cand_val, int* __restrict__ cand_idx) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ float v[SURV_CAP];
  __shared__ int id[SURV_CAP];
  __shared__ int count;
  __shared__ float bound;
  const float* row = logits + (size_t)blockIdx.x * ld;
  float mx = -INFINITY;
  for (int i = threadIdx.x; i < V; i += SM_THREADS) mx = fmaxf(mx, row[i]);
  v[threadIdx.x] = mx;
  id[threadIdx.x] = threadIdx.x;
  if (threadIdx.x == 0) count = 0;
  bitonic_desc(v, id, SM_THREADS);
  if (threadIdx.x == 0) bound = v[(k < SM_THREADS ? k : SM_THREADS) - 1];
  __syncthreads();
  const float L = bound;
  __syncthreads();
  for (int i = threadIdx.x; i < V; i += SM_THREADS) {
    const float x = row[i];
    if (x >= L) {
      const int slot = atomicAdd(&count, 1);
      if (slot < SURV_CAP) { v[slot] = x; id[slot] = i; }
    }
  }
  __syncthreads();
  if (count > SURV_CAP) {
    // More survivors than the list holds (thousands of elements tie with or exceed the bound: constant rows, masked
    // vocabularies, quantised heads): which of them got a slot depended on arrival order.  Exact and deterministic
    // instead: k rounds of a block-wide arg-max over the row, each restricted to elements that come AFTER the
    // previously emitted one in the total order (value descending, index ascending).  Rare and slow (k passes).
    __shared__ float sv[32];
    __shared__ int si[32];
    __shared__ float last_v;
    __shared__ int last_i;
    if (threadIdx.x == 0) { last_v = INFINITY; last_i = -1; }
    __syncthreads();
    for (int r = 0; r < k; ++r) {
      const float lv = last_v;
      const int li = last_i;
      float bv = -INFINITY;
      int bi = 0x7fffffff;
      for (int i = threadIdx.x; i < V; i += SM_THREADS) {
        const float x = row[i];
        const bool after = x < lv || (x == lv && i > li);
        if (after && better(x, i, bv, bi)) { bv = x; bi = i; }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (better(ov, oi, bv, bi)) { bv = ov; bi = oi; }
      }
      if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = bv; si[threadIdx.x >> 5] = bi; }
      __syncthreads();
      if (threadIdx.x < 32) {
        bv = sv[threadIdx.x]; bi = si[threadIdx.x];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
          const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
          if (better(ov, oi, bv, bi)) { bv = ov; bi = oi; }
        }
        if (threadIdx.x == 0) {
          const bool ok = bi != 0x7fffffff && bv > -INFINITY;
          cand_val[(size_t)blockIdx.x * k + r] = ok ? bv : -INFINITY;
          cand_idx[(size_t)blockIdx.x * k + r] = ok ? bi + idx_offset : -1;
          last_v = bv; last_i = ok ? bi : 0x7fffffff;
        }
      }
      __syncthreads();
    }
    return;
  }
  int n = count;
  int n2 = 1;
  while (n2 < n) n2 <<= 1;
  if (n2 < 2) n2 = 2;
  for (int i = n + threadIdx.x; i < n2; i += SM_THREADS) { v[i] = -INFINITY; id[i] = 0x7fffffff; }
  bitonic_desc(v, id, n2);
  for (int i = threadIdx.x; i < k; i += SM_THREADS) {
    const bool ok = i < n && v[i] > -INFINITY;
    cand_val[(size_t)blockIdx.x * k + i] = ok ? v[i] : -INFINITY;
    cand_idx[(size_t)blockIdx.x * k + i] = ok ? id[i] + idx_offset : -1;
  }
}

// ------------------------------------------------------------------------------------ RNG
// Philox-4x32-10: counter-based, so every rank derives the same uniform for (seed, step, row).
__device__ __forceinline__ float philox_uniform(unsigned long long seed, unsigned int step, unsigned int row) {
  unsigned int c0 = step, c1 = row, c2 = 0x9E3779B9u, c3 = 0xBB67AE85u;
  unsigned int k0 = (unsigned int)seed, k1 = (unsigned int)(seed >> 32);
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const unsigned int hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const unsigned int hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const unsigned int n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return ((c0 >> 8) + 0.5f) * (1.0f / 16777216.0f);      // (0, 1)
}

// ------------------------------------------------------------------------------------ final selection
// Candidates: n_lists lists of k (value, index) pairs per row (list stride = list_stride elements,
// e.g. one list per TP rank).  One CTA per row: merge (sort) → temperature → top-k → top-p → sample.
__global__ void __launch_bounds__(512)
sample_kernel(const float* __restrict__ cand_val, const int* __restrict__ cand_idx, int n_lists,
              size_t list_stride, int k, int top_k, float top_p, float inv_temp, int greedy,
              unsigned long long seed, const int* __restrict__ step_ptr, int* __restrict__ out_tok) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ float v[512];
  __shared__ int id[512];
  const int row = blockIdx.x;
  const int total = n_lists * k;
  int n2 = 2;
  while (n2 < total) n2 <<= 1;
  for (int i = threadIdx.x; i < n2; i += blockDim.x) {
    if (i < total) {
      const int li = i / k, j = i - li * k;
      v[i] = cand_val[li * list_stride + (size_t)row * k + j];
      id[i] = cand_idx[li * list_stride + (size_t)row * k + j];
      if (id[i] < 0) { v[i] = -INFINITY; id[i] = 0x7fffffff; }
    } else {
      v[i] = -INFINITY; id[i] = 0x7fffffff;
    }
  }
  bitonic_desc(v, id, n2);
  if (threadIdx.x == 0) {
    int keep = top_k > 0 && top_k < total ? top_k : total;
    while (keep > 1 && !(v[keep - 1] > -INFINITY)) --keep;
    int choice = 0;
    if (!greedy && keep > 1) {
      const float mx = v[0] * inv_temp;
      float z = 0.f;
      for (int i = 0; i < keep; ++i) z += __expf(v[i] * inv_temp - mx);
      // nucleus: keep tokens while the mass of the tokens before them is < top_p
      float cum = 0.f, kept = 0.f;
      int nk = 0;
      for (int i = 0; i < keep; ++i) {
        if (i > 0 && cum >= top_p * z) break;
        const float e = __expf(v[i] * inv_temp - mx);
        cum += e; kept += e; ++nk;
      }
      const float u = philox_uniform(seed, (unsigned)(step_ptr ? *step_ptr : 0), (unsigned)row) * kept;
      float run = 0.f;
      choice = nk - 1;
      for (int i = 0; i < nk; ++i) {
        run += __expf(v[i] * inv_temp - mx);
        if (u <= run) { choice = i; break; }
      }
    }
    out_tok[row] = id[choice] == 0x7fffffff ? 0 : id[choice];
  }
}

// ------------------------------------------------------------------------------------ step bookkeeping
// One thread per row: finished rows emit pad, EOS marks a row finished, the token goes into
// sequences[b, cur_len] and becomes the next input; thread 0 then advances the shared counters.
__global__ void advance_kernel(const int* __restrict__ tok, int* __restrict__ next_ids,
                               int* __restrict__ finished, int* __restrict__ sequences, int B, int max_len,
                               int eos, int pad, int* __restrict__ cur_len, int* __restrict__ cache_len,
                               int* __restrict__ step, int* __restrict__ n_unfinished) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ int alive;
  if (threadIdx.x == 0) alive = 0;
  __syncthreads();
  const int cl = *cur_len;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    int t = tok[b];
    int f = finished[b];
    if (f) t = pad;
    f = f | (t == eos);
    finished[b] = f;
    next_ids[b] = t;
    if (cl < max_len) sequences[(size_t)b * max_len + cl] = t;
    if (!f) atomicAdd(&alive, 1);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    *cur_len = cl + 1;
    *cache_len = *cache_len + 1;
    *step = *step + 1;
    *n_unfinished = alive;
  }
}

}  // namespace jl

using namespace jl;

extern "C" JL_API int jl_argmax(const float* logits, int B, int V, int ld, int idx_offset, float* out_val,
                         int* out_idx, void* stream) {
  if (B == 0) return JL_OK;
  JL_CUDA(launch_pdl(argmax_kernel, dim3(B), dim3(SM_THREADS), 0, (cudaStream_t)stream, logits, V, ld,
                     idx_offset, out_val, out_idx));
  return JL_OK;
}

extern "C" JL_API int jl_topk(const float* logits, int B, int V, int ld, int idx_offset, int k, float* cand_val,
                       int* cand_idx, void* stream) {
  JL_CHECK_ARG(k >= 1 && k <= 256, "jl_topk: k=%d outside 1..256", k);
  if (B == 0) return JL_OK;
  JL_CUDA(launch_pdl(topk_kernel, dim3(B), dim3(SM_THREADS), 0, (cudaStream_t)stream, logits, V, ld,
                     idx_offset, k, cand_val, cand_idx));
  return JL_OK;
}

extern "C" JL_API int jl_sample(const float* cand_val, const int* cand_idx, int B, int n_lists, size_t list_stride,
                         int k, int top_k, float top_p, float temperature, unsigned long long seed,
                         const int* step_ptr, int* out_tok, void* stream) {
  JL_CHECK_ARG(n_lists >= 1 && k >= 1 && n_lists * k <= 512, "jl_sample: n_lists*k=%d exceeds 512", n_lists * k);
  if (B == 0) return JL_OK;
  const int greedy = temperature == 0.0f;
  const float inv_temp = greedy ? 1.0f : 1.0f / temperature;
  JL_CUDA(launch_pdl(sample_kernel, dim3(B), dim3(512), 0, (cudaStream_t)stream, cand_val, cand_idx, n_lists,
                     list_stride, k, top_k, top_p, inv_temp, greedy, seed, step_ptr, out_tok));
  return JL_OK;
}

extern "C" JL_API int jl_decode_advance(const int* tok, int* next_ids, int* finished, int* sequences, int B,
                                 int max_len, int eos, int pad, int* cur_len, int* cache_len, int* step,
                                 int* n_unfinished, void* stream) {
  if (B == 0) return JL_OK;
  JL_CUDA(launch_pdl(advance_kernel, dim3(1), dim3(B < 256 ? ((B + 31) / 32) * 32 : 256), 0,
                     (cudaStream_t)stream, tok, next_ids, finished, sequences, B, max_len, eos, pad, cur_len,
                     cache_len, step, n_unfinished));
  return JL_OK;
}
