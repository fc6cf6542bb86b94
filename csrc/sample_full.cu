// Exact temperature / top-k / top-p / categorical selection over the WHOLE vocabulary, one CTA per row, plus the
// sequence bookkeeping of the step — the second (and last) launch of a sampling decode step at any tp.
//
// Reference behaviour: the HF Flax warpers as configured by /root/reference/jax_llama/generation.py:28-41 —
// logits / T → top-k (if set) → top-p with min_tokens_to_keep = 1 (sorted descending, a token is kept while the
// probability mass before it is < top_p) → categorical; temperature 0 = arg-max.  Round 1 ran this on at most 64
// candidates per row, so `top_k=None` silently meant "nucleus over the 64 most likely tokens" on the GPU and a
// plateau of equal logits could overflow the candidate list.  This kernel has no candidate budget:
//
//   * every token gets the 64-bit sort key (order-preserving bits of its scaled logit, ~index): larger key = better,
//     equal logits order by LOWER index — selection is deterministic whatever the thread arrival order;
//   * "the element whose interval of the descending cumulative sum contains `target`" is found by a radix select
//     (11 + 11 + 10 bits of the value, then the index among exact ties) over integer histograms.  With unit weights
//     this is the k-th largest element (top-k); with probability weights it is the nucleus boundary (top-p) and the
//     sampled token (inverse CDF at u·Z).  Probabilities are exp(z − max) in 2^-40 fixed point, so every histogram
//     is an integer sum: associative, hence bit-identical on every rank and every run (float atomics are not);
//   * tensor parallelism: the decode kernel stores every rank's logits shard into every rank's [B, V] buffer with
//     multimem.st, so all ranks run this kernel on identical inputs with the same Philox stream (seed, step, row) and
//     pick the same token without a broadcast.  The kernel first waits for the arg-max candidate packets of all ranks
//     (they are published after a system fence that follows the logits stores).
#include "common.cuh"
#include "jllama_b200.h"

namespace jl {

constexpr int SF_THREADS = 1024;
constexpr int SF_BINS = 2048;
constexpr float SF_FIX = 1099511627776.f;          // 2^40

__device__ __forceinline__ uint32_t sf_key(float z) {
  if (!(z == z)) return 0u;                          // NaN sorts last
  const uint32_t u = __float_as_uint(z);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// same generator as sampling.cu (Philox-4x32-10 keyed by seed, counter = (step, row))
__device__ __forceinline__ float sf_uniform(unsigned long long seed, unsigned int step, unsigned int row) {
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

struct SfRow {
  const float* x;           // logits of this row
  int V;
  float inv_temp, zmax;     // z = x · inv_temp; zmax = max z
  // current kept set: elements with key > kmin, or key == kmin and index <= imax (lower index = better)
  uint32_t kmin;
  int imax;
};

__device__ __forceinline__ bool sf_in_set(const SfRow& r, uint32_t key, int i) {
  return key > r.kmin || (key == r.kmin && i <= r.imax);
}
__device__ __forceinline__ unsigned long long sf_mass(const SfRow& r, float z) {
  return (unsigned long long)(exp2f((z - r.zmax) * 1.4426950408889634f) * SF_FIX);
}

// Descending scan of the histogram (warp 0): the bin in which the running sum (starting at `base`) first exceeds
// `target`, and the sum before that bin.  A target at or beyond the total (rounding of a fraction) selects the last
// non-empty bin.
__device__ void sf_scan(const unsigned long long* hist, int nbins, unsigned long long target, unsigned long long base,
                        int* out_bin, unsigned long long* out_before, unsigned long long* out_total) {
  if (threadIdx.x >= 32) return;
  const int lane = threadIdx.x;
  const int per = nbins / 32;                            // lane 0 owns the highest bins
  unsigned long long s = 0;
  for (int j = 0; j < per; ++j) s += hist[nbins - 1 - (lane * per + j)];
  unsigned long long incl = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  const unsigned long long total = __shfl_sync(0xffffffffu, incl, 31) + base;
  const unsigned long long excl = incl - s + base;
  const unsigned cross = __ballot_sync(0xffffffffu, s > 0 && target < excl + s);
  const unsigned nonempty = __ballot_sync(0xffffffffu, s > 0);
  int owner;
  bool past_end = false;
  if (cross) owner = __ffs(cross) - 1;
  else { owner = nonempty ? 31 - __clz(nonempty) : 0; past_end = true; }
  if (lane == owner) {
    unsigned long long acc = excl, before = excl;
    int bin = nbins - 1 - lane * per;
    for (int j = 0; j < per; ++j) {
      const int b = nbins - 1 - (lane * per + j);
      const unsigned long long h = hist[b];
      if (h > 0) { bin = b; before = acc; if (!past_end && acc + h > target) break; }
      acc += h;
    }
    *out_bin = bin; *out_before = before;
  }
  if (lane == 0) *out_total = total;
}

// The element of the kept set whose interval [before, before + w) of the descending cumulative sum contains `target`
// (w = 1 per element if BY_COUNT, else its fixed-point probability).  Returns (key, index); *before_out = sum strictly
// before that element, *set_total = sum over the whole kept set (valid after the first level).
template <bool BY_COUNT>
__device__ void sf_select(const SfRow& r, unsigned long long target, bool target_is_fraction, double fraction,
                          unsigned long long* hist, int* s_bin, unsigned long long* s_before, unsigned long long* s_total,
                          int* s_idx, uint32_t* key_out, int* idx_out, unsigned long long* before_out,
                          unsigned long long* total_out) {
  uint32_t prefix = 0;            // key bits fixed so far (top `done` bits)
  int done = 0;
  unsigned long long before = 0, set_total = 0;
  const int widths[3] = {11, 11, 10};
#pragma unroll 1
  for (int level = 0; level < 3; ++level) {
    const int w = widths[level], shift = 32 - done - w, nb = 1 << w;
    for (int b = threadIdx.x; b < SF_BINS; b += SF_THREADS) hist[b] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < r.V; i += SF_THREADS) {
      const float z = r.x[i] * r.inv_temp;
      const uint32_t key = sf_key(z);
      if (!sf_in_set(r, key, i)) continue;
      if (done > 0 && (key >> (32 - done)) != (prefix >> (32 - done))) continue;
      const int bin = (key >> shift) & (nb - 1);
      atomicAdd(&hist[bin], BY_COUNT ? 1ull : sf_mass(r, z));
    }
    __syncthreads();
    if (level == 0 && target_is_fraction) {
      // the target is a fraction of the kept set's total, known only now: target = max(fraction·Z, 1) − 1
      // (a token is kept while the mass before it is < fraction·Z)
      if (threadIdx.x < 32) {
        unsigned long long s = 0;
        for (int b = threadIdx.x; b < nb; b += 32) s += hist[b];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (threadIdx.x == 0) {
          unsigned long long t = (unsigned long long)((double)s * fraction);
          if (t > s) t = s;
          *s_total = t > 0 ? t - 1 : 0;
        }
      }
      __syncthreads();
      target = *s_total;
      __syncthreads();
    }
    sf_scan(hist, nb, target, before, s_bin, s_before, s_total);
    __syncthreads();
    prefix |= (uint32_t)(*s_bin) << shift;
    before = *s_before;
    if (level == 0) set_total = *s_total;
    done += w;
    __syncthreads();
  }
  // exact ties: elements of the set with key == prefix all weigh the same; the r-th lowest index crosses
  unsigned long long wt = 1;
  if (!BY_COUNT) {
    // weight of one tied element: recompute from the key (z is not recoverable bit-exactly from the key only for NaN)
    const uint32_t u = (prefix & 0x80000000u) ? (prefix & 0x7fffffffu) : ~prefix;
    wt = sf_mass(r, __uint_as_float(u));
  }
  unsigned long long rank = wt > 0 ? (target >= before ? (target - before) / wt : 0) : 0;      // 0-based among ties
  // count ties and find the (rank)-th lowest index: lowest index by atomicMin when rank == 0 (the common case)
  if (threadIdx.x == 0) { *s_idx = 0x7fffffff; *s_bin = 0; }
  __syncthreads();
  for (int i = threadIdx.x; i < r.V; i += SF_THREADS) {
    const uint32_t key = sf_key(r.x[i] * r.inv_temp);
    if (key == prefix && sf_in_set(r, key, i)) { atomicMin(s_idx, i); atomicAdd(s_bin, 1); }
  }
  __syncthreads();
  int idx = *s_idx;
  const int n_ties = *s_bin;
  __syncthreads();
  if (rank >= (unsigned long long)n_ties) rank = n_ties > 0 ? n_ties - 1 : 0;
  if (rank > 0) {
    // radix select over the index (17 bits: 9 + 8) among the ties, ascending
    int ipre = 0, idone = 0;
    unsigned long long ibefore = 0;
    const int iw[2] = {9, 8};
#pragma unroll 1
    for (int level = 0; level < 2; ++level) {
      const int w = iw[level], shift = 17 - idone - w, nb = 1 << w;
      for (int b = threadIdx.x; b < SF_BINS; b += SF_THREADS) hist[b] = 0;
      __syncthreads();
      for (int i = threadIdx.x; i < r.V; i += SF_THREADS) {
        const uint32_t key = sf_key(r.x[i] * r.inv_temp);
        if (key != prefix || !sf_in_set(r, key, i)) continue;
        if (idone > 0 && (i >> (17 - idone)) != (ipre >> (17 - idone))) continue;
        // ascending index order = descending (nb − 1 − digit)
        atomicAdd(&hist[nb - 1 - ((i >> shift) & (nb - 1))], 1ull);
      }
      __syncthreads();
      sf_scan(hist, nb, rank, ibefore, s_bin, s_before, s_total);
      __syncthreads();
      ipre |= (nb - 1 - *s_bin) << shift;
      ibefore = *s_before;
      idone += w;
      __syncthreads();
    }
    idx = ipre;
  }
  *key_out = prefix; *idx_out = idx;
  *before_out = before + rank * wt;
  *total_out = set_total;
}

struct SampleFullParams {
  const float* logits;      // [B, ld] fp32 (full vocabulary)
  int V, ld, B;
  float inv_temp, top_p;
  int top_k, greedy;
  unsigned long long seed;
  const int* step;
  int* tok;
  // wait for the other ranks' logits shards (tp > 1): candidate packets of the decode kernel, see decode_mega.cu
  const uint2* cand; const unsigned int* gen; int tp, grid, two_l, rank;
  // bookkeeping (advance_kernel), done by the last CTA to finish
  int* ticket; int* next_ids; int* finished; int* sequences; int max_len, eos, pad;
  int* cur_len; int* cache_len; int* step_rw; int* n_unfinished;
};

__global__ void __launch_bounds__(SF_THREADS) sample_full_kernel(const SampleFullParams p) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ unsigned long long hist[SF_BINS];
  __shared__ unsigned long long s_before, s_total, s_kk;
  __shared__ int s_bin, s_idx, s_last;
  const int row = blockIdx.x;
  if (p.cand != nullptr) {
    // every (rank, CTA) of the decode kernel has published its candidate for this row ⇒ its logits stores landed
    const unsigned int gen_now = *reinterpret_cast<const volatile unsigned int*>(p.gen);
    const uint32_t tag = gen_now * (uint32_t)p.two_l;
    const uint2* base = p.cand + (size_t)((gen_now - 1u) & 1u) * p.tp * p.grid * p.B * 2;
    for (int c = threadIdx.x; c < p.tp * p.grid; c += SF_THREADS) {
      const uint2* q = base + ((size_t)c * p.B + row) * 2;
      unsigned spins = 0;
      for (;;) {
        uint2 a, b;
        asm volatile("ld.relaxed.sys.global.v2.u32 {%0,%1}, [%2];" : "=r"(a.x), "=r"(a.y) : "l"(q) : "memory");
        asm volatile("ld.relaxed.sys.global.v2.u32 {%0,%1}, [%2];" : "=r"(b.x), "=r"(b.y) : "l"(q + 1) : "memory");
        if (a.y == tag && b.y == tag) break;
        if (++spins > (1u << 24)) {
          printf("jllama_b200: rank %d timed out waiting for the logits of rank %d (tag %u)\n", p.rank, c / p.grid, tag);
          __trap();
        }
      }
    }
    __threadfence_system();
    __syncthreads();
  }
  SfRow r;
  r.x = p.logits + (size_t)row * p.ld; r.V = p.V; r.inv_temp = p.inv_temp; r.kmin = 0; r.imax = 0x7fffffff;
  // ---- max of the scaled logits and the best (key, ~index) in one pass
  unsigned long long best = 0;
  for (int i = threadIdx.x; i < p.V; i += SF_THREADS) {
    const unsigned long long kk = ((unsigned long long)sf_key(r.x[i] * p.inv_temp) << 32) | (unsigned)(0x7fffffff - i);
    best = kk > best ? kk : best;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long t = __shfl_xor_sync(0xffffffffu, best, o);
    best = t > best ? t : best;
  }
  if (threadIdx.x == 0) s_kk = 0;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) atomicMax(&s_kk, best);
  __syncthreads();
  best = s_kk;
  int token = 0x7fffffff - (int)(best & 0xffffffffu);
  {
    const uint32_t k = (uint32_t)(best >> 32);
    const uint32_t u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
    r.zmax = __uint_as_float(u);
  }
  if (!p.greedy) {
    uint32_t key; int idx; unsigned long long before, total;
    if (p.top_k > 0 && p.top_k < p.V) {
      sf_select<true>(r, (unsigned long long)(p.top_k - 1), false, 0.0, hist, &s_bin, &s_before, &s_total, &s_idx, &key, &idx,
                      &before, &total);
      r.kmin = key; r.imax = idx;
    }
    if (p.top_p < 1.0f) {
      sf_select<false>(r, 0, true, (double)p.top_p, hist, &s_bin, &s_before, &s_total, &s_idx, &key, &idx, &before, &total);
      r.kmin = key; r.imax = idx;
    }
    // inverse CDF over the kept set, descending order
    const float u01 = sf_uniform(p.seed, (unsigned)(p.step ? *p.step : 0), (unsigned)row);
    sf_select<false>(r, 0, true, (double)u01, hist, &s_bin, &s_before, &s_total, &s_idx, &key, &idx, &before, &total);
    token = idx;
  }
  if (threadIdx.x == 0) {
    if (token < 0 || token >= p.V) token = 0;
    p.tok[row] = token;
  }
  // ---- last CTA: sequence bookkeeping of all rows
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const int prev = atomicAdd(p.ticket, 1);
    s_last = prev == (int)gridDim.x - 1;
    if (s_last) *p.ticket = 0;
  }
  __syncthreads();
  if (s_last && p.next_ids != nullptr) {
    __threadfence();
    __shared__ int alive;
    if (threadIdx.x == 0) alive = 0;
    __syncthreads();
    const int cl = *p.cur_len;
    for (int b = threadIdx.x; b < p.B; b += SF_THREADS) {
      int t = __ldcg(p.tok + b);
      int f = p.finished[b];
      if (f) t = p.pad;
      f = f | (t == p.eos);
      p.finished[b] = f;
      p.next_ids[b] = t;
      if (cl < p.max_len) p.sequences[(size_t)b * p.max_len + cl] = t;
      if (!f) atomicAdd(&alive, 1);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      *p.cur_len = cl + 1;
      *p.cache_len = *p.cache_len + 1;
      *p.step_rw = *p.step_rw + 1;
      *p.n_unfinished = alive;
    }
  }
}

}  // namespace jl

using namespace jl;

// temperature == 0 → arg-max (ties: lower index); top_k <= 0 or >= V → no top-k; top_p >= 1 → no nucleus filter.
// `cand` / `gen` (may be NULL): wait for the decode kernel of every rank first (tensor-parallel logits gathered by
// multimem.st).  `next_ids` == NULL skips the bookkeeping (selection only: tok[B]).
extern "C" JL_API int jl_sample_full(const float* logits, int B, int V, int ld, float temperature, int top_k, float top_p,
                                     unsigned long long seed, const int* step, int* tok, const void* cand,
                                     const unsigned int* gen, int tp, int rank, int num_sms, int L, int* ticket,
                                     int* next_ids, int* finished, int* sequences, int max_len, int eos, int pad,
                                     int* cur_len, int* cache_len, int* step_rw, int* n_unfinished, void* stream) {
  JL_CHECK_ARG(logits && tok && ticket && B >= 1 && V >= 1 && V <= (1 << 17) && (ld >= V || ld == 0),
               "jl_sample_full: bad arguments (V=%d must be <= 131072; ld >= V, or 0 = every row reads the same logits)", V);
  if (num_sms <= 0) num_sms = 148;
  SampleFullParams p{};
  p.logits = logits; p.V = V; p.ld = ld; p.B = B;
  p.greedy = temperature == 0.0f;
  p.inv_temp = p.greedy ? 1.0f : 1.0f / temperature;
  p.top_k = top_k; p.top_p = top_p; p.seed = seed; p.step = step; p.tok = tok;
  p.cand = (const uint2*)cand; p.gen = gen; p.tp = tp; p.grid = num_sms; p.two_l = 2 * L; p.rank = rank;
  p.ticket = ticket; p.next_ids = next_ids; p.finished = finished; p.sequences = sequences;
  p.max_len = max_len; p.eos = eos; p.pad = pad;
  p.cur_len = cur_len; p.cache_len = cache_len; p.step_rw = step_rw; p.n_unfinished = n_unfinished;
  JL_CUDA(launch_pdl(sample_full_kernel, dim3(B), dim3(SF_THREADS), 0, (cudaStream_t)stream, p));
  return JL_OK;
}
