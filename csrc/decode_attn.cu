// Single-token (decode) attention over the KV cache with GQA: split-KV "flash decoding".
//
// The reference attends over all `max_length` cache slots with a mask every step
// (/root/reference/jax_llama/model.py:194-198,263-288); here only the filled, non-padding prefix
// [pad_left[b], cache_len + n_new) is streamed, once, at HBM rate, and no mask tensor or repeated KV
// heads are ever materialised: the G query heads of a KV group share every K/V byte loaded.
//
// Work split: grid = (splits, nkv, B).  A CTA streams a contiguous key range; inside it every group of
// HD/8 lanes owns 4 consecutive keys per iteration (one 16-byte load per lane per key, K and V issued
// together → 8 loads in flight per lane).  QKᵀ partials are combined with a reduce-scatter butterfly
// (5 shuffles per head per 4 keys instead of 16), softmax is online in the log2 domain, and the
// per-split (m, l, o) triples go to a scratch buffer that `attn_combine_kernel` folds.
#include <mutex>

#include "common.cuh"
#include "jllama_b200.h"

namespace jl {

constexpr int DA_WARPS = 4;
constexpr int DA_KPI = 4;   // keys per lane-group per iteration

template <int G, int HD>
__global__ void __launch_bounds__(DA_WARPS * 32)
decode_attn_kernel(const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ kc,
                   const __nv_bfloat16* __restrict__ vc, const int* __restrict__ cache_len_ptr,
                   const int* __restrict__ pad_left, float* __restrict__ o_part,
                   float* __restrict__ ml_part, __nv_bfloat16* __restrict__ out, int nkv, int T,
                   int n_new, float scale_log2, int splits, int q_bstride) {
  constexpr int LPK = HD / 8;            // lanes per key row
  constexpr int GPW = 32 / LPK;          // lane-groups per warp
  constexpr int NG = DA_WARPS * GPW;     // lane-groups per CTA
  __shared__ float s_o[NG][G][HD];
  __shared__ float s_ml[NG][G][2];

  pdl_wait();
  pdl_launch_dependents();

  const int split = blockIdx.x, kvh = blockIdx.y, b = blockIdx.z;
  const int nh = nkv * G;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int lig = lane % LPK;                       // lane in group → dims [8·lig, 8·lig+8)
  const int gid = warp * GPW + lane / LPK;
  const int group_base = lane - lig;
  const int kidx = ((lig / (LPK / 2)) & 1) * 2 + ((lig / (LPK / 4)) & 1);   // key this lane owns post-reduce

  int len = (cache_len_ptr ? *cache_len_ptr : 0) + n_new;
  len = len > T ? T : len;
  const int first = pad_left ? pad_left[b] : 0;
  const int valid = len - first > 0 ? len - first : 0;
  constexpr int KEYS_PER_ITER = NG * DA_KPI;
  int chunk = (valid + splits - 1) / splits;
  chunk = ((chunk + KEYS_PER_ITER - 1) / KEYS_PER_ITER) * KEYS_PER_ITER;
  const int lo = first + split * chunk;
  const int hi = lo + chunk < len ? lo + chunk : len;

  // q for the G heads of this KV group, pre-scaled by softmax_scale·log2(e)
  float qf[G][8];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    const uint4 a = *reinterpret_cast<const uint4*>(q + (size_t)b * q_bstride + (size_t)(kvh * G + g) * HD + lig * 8);
    qf[g][0] = bf16_lo(a.x) * scale_log2; qf[g][1] = bf16_hi(a.x) * scale_log2;
    qf[g][2] = bf16_lo(a.y) * scale_log2; qf[g][3] = bf16_hi(a.y) * scale_log2;
    qf[g][4] = bf16_lo(a.z) * scale_log2; qf[g][5] = bf16_hi(a.z) * scale_log2;
    qf[g][6] = bf16_lo(a.w) * scale_log2; qf[g][7] = bf16_hi(a.w) * scale_log2;
  }
  float m[G], l[G], o[G][8];
#pragma unroll
  for (int g = 0; g < G; ++g) {
    m[g] = -INFINITY;
    l[g] = 0.f;
#pragma unroll
    for (int e = 0; e < 8; ++e) o[g][e] = 0.f;
  }

  const size_t head_off = ((size_t)b * nkv + kvh) * (size_t)T * HD;
  const __nv_bfloat16* kbase = kc + head_off + lig * 8;
  const __nv_bfloat16* vbase = vc + head_off + lig * 8;

  // warp-uniform trip count (the body shuffles across the whole warp); tails are predicated
  for (int base = lo; base < hi; base += KEYS_PER_ITER) {
    const int key0 = base + gid * DA_KPI;
    uint4 kr[DA_KPI], vr[DA_KPI];
#pragma unroll
    for (int i = 0; i < DA_KPI; ++i) {
      const bool ok = key0 + i < hi;
      kr[i] = ok ? ldg_stream(kbase + (size_t)(key0 + i) * HD) : make_uint4(0, 0, 0, 0);
      vr[i] = ok ? ldg_stream(vbase + (size_t)(key0 + i) * HD) : make_uint4(0, 0, 0, 0);
    }
    float part[G][DA_KPI];
#pragma unroll
    for (int i = 0; i < DA_KPI; ++i) {
      const float kf[8] = {bf16_lo(kr[i].x), bf16_hi(kr[i].x), bf16_lo(kr[i].y), bf16_hi(kr[i].y),
                           bf16_lo(kr[i].z), bf16_hi(kr[i].z), bf16_lo(kr[i].w), bf16_hi(kr[i].w)};
#pragma unroll
      for (int g = 0; g < G; ++g) {
        float s = 0.f;
#pragma unroll
        for (int e = 0; e < 8; ++e) s = fmaf(qf[g][e], kf[e], s);
        part[g][i] = s;
      }
    }
    float p[G][DA_KPI];
#pragma unroll
    for (int g = 0; g < G; ++g) {
      // reduce-scatter across the lane group: afterwards this lane holds the full score of key `kidx`
      const bool up = (lig & (LPK / 2)) != 0;
      float k0 = up ? part[g][2] : part[g][0], k1 = up ? part[g][3] : part[g][1];
      const float s0 = up ? part[g][0] : part[g][2], s1 = up ? part[g][1] : part[g][3];
      k0 += __shfl_xor_sync(0xffffffffu, s0, LPK / 2);
      k1 += __shfl_xor_sync(0xffffffffu, s1, LPK / 2);
      const bool up2 = (lig & (LPK / 4)) != 0;
      float s = up2 ? k1 : k0;
      const float snd = up2 ? k0 : k1;
      s += __shfl_xor_sync(0xffffffffu, snd, LPK / 4);
#pragma unroll
      for (int off = LPK / 8; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
      if (key0 + kidx >= hi) s = -INFINITY;
      float mx = fmaxf(s, __shfl_xor_sync(0xffffffffu, s, LPK / 2));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, LPK / 4));
      const float m_new = fmaxf(m[g], mx);
      const float m_safe = m_new == -INFINITY ? 0.f : m_new;
      const float corr = exp2f(m[g] - m_safe);
      const float pe = exp2f(s - m_safe);
      m[g] = m_new;
      float psum = 0.f;
#pragma unroll
      for (int i = 0; i < DA_KPI; ++i) {
        p[g][i] = __shfl_sync(0xffffffffu, pe, group_base + (i >> 1) * (LPK / 2) + (i & 1) * (LPK / 4));
        psum += p[g][i];
      }
      l[g] = l[g] * corr + psum;
#pragma unroll
      for (int e = 0; e < 8; ++e) o[g][e] *= corr;
    }
#pragma unroll
    for (int i = 0; i < DA_KPI; ++i) {
      const float vf[8] = {bf16_lo(vr[i].x), bf16_hi(vr[i].x), bf16_lo(vr[i].y), bf16_hi(vr[i].y),
                           bf16_lo(vr[i].z), bf16_hi(vr[i].z), bf16_lo(vr[i].w), bf16_hi(vr[i].w)};
#pragma unroll
      for (int g = 0; g < G; ++g)
#pragma unroll
        for (int e = 0; e < 8; ++e) o[g][e] = fmaf(p[g][i], vf[e], o[g][e]);
    }
  }

  // ---- fold the NG lane-groups of this CTA
#pragma unroll
  for (int g = 0; g < G; ++g) {
#pragma unroll
    for (int e = 0; e < 8; ++e) s_o[gid][g][lig * 8 + e] = o[g][e];
    if (lig == 0) { s_ml[gid][g][0] = m[g]; s_ml[gid][g][1] = l[g]; }
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < G * HD; idx += DA_WARPS * 32) {
    const int g = idx / HD, d = idx - g * HD;
    float M = -INFINITY;
#pragma unroll
    for (int j = 0; j < NG; ++j) M = fmaxf(M, s_ml[j][g][0]);
    const float Ms = M == -INFINITY ? 0.f : M;
    float acc = 0.f, L = 0.f;
#pragma unroll
    for (int j = 0; j < NG; ++j) {
      const float w = exp2f(s_ml[j][g][0] - Ms);
      acc += w * s_o[j][g][d];
      L += w * s_ml[j][g][1];
    }
    const int h = kvh * G + g;
    if (splits == 1) {
      out[((size_t)b * nh + h) * HD + d] = __float2bfloat16(L > 0.f ? acc / L : 0.f);
    } else {
      const size_t pi = ((size_t)b * nh + h) * splits + split;
      o_part[pi * HD + d] = acc;
      if (d == 0) { ml_part[pi * 2] = M; ml_part[pi * 2 + 1] = L; }
    }
  }
}

template <int HD>
__global__ void __launch_bounds__(HD)
attn_combine_kernel(const float* __restrict__ o_part, const float* __restrict__ ml_part,
                    __nv_bfloat16* __restrict__ out, int splits) {
  pdl_wait();
  pdl_launch_dependents();
  const size_t bh = blockIdx.x;
  const int d = threadIdx.x;
  float M = -INFINITY;
  for (int s = 0; s < splits; ++s) M = fmaxf(M, ml_part[(bh * splits + s) * 2]);
  const float Ms = M == -INFINITY ? 0.f : M;
  float acc = 0.f, L = 0.f;
  for (int s = 0; s < splits; ++s) {
    const float w = exp2f(ml_part[(bh * splits + s) * 2] - Ms);
    acc += w * o_part[(bh * splits + s) * HD + d];
    L += w * ml_part[(bh * splits + s) * 2 + 1];
  }
  out[bh * HD + d] = __float2bfloat16(L > 0.f ? acc / L : 0.f);
}

template <int G, int HD>
static cudaError_t launch_decode(const void* q, const void* kc, const void* vc, const int* cache_len_ptr,
                                 const int* pad_left, float* o_part, float* ml_part, void* out, int B,
                                 int nkv, int T, int n_new, float scale, int splits, int q_bstride, cudaStream_t s) {
  cudaError_t e = launch_pdl(decode_attn_kernel<G, HD>, dim3(splits, nkv, B), dim3(DA_WARPS * 32), 0, s,
                             (const __nv_bfloat16*)q, (const __nv_bfloat16*)kc, (const __nv_bfloat16*)vc,
                             cache_len_ptr, pad_left, o_part, ml_part, (__nv_bfloat16*)out, nkv, T, n_new,
                             scale * 1.4426950408889634f, splits, q_bstride);
  if (e != cudaSuccess || splits == 1) return e;
  return launch_pdl(attn_combine_kernel<HD>, dim3(B * nkv * G), dim3(HD), 0, s, (const float*)o_part,
                    (const float*)ml_part, (__nv_bfloat16*)out, splits);
}

}  // namespace jl

using namespace jl;

extern "C" JL_API size_t jl_decode_attn_scratch_bytes(int B, int nh, int hd, int splits) {
  return splits <= 1 ? 0 : (size_t)B * nh * splits * (hd + 2) * sizeof(float);
}

extern "C" JL_API int jl_decode_attn_splits(int B, int nkv, int max_len, int num_sms) {
  // enough CTAs for ~2 per SM, but never less than 256 keys per split
  if (num_sms <= 0) num_sms = 148;
  int want = (2 * num_sms + B * nkv - 1) / (B * nkv);
  int cap = (max_len + 255) / 256;
  int s = want < cap ? want : cap;
  return s < 1 ? 1 : (s > 64 ? 64 : s);
}

extern "C" JL_API int jl_decode_attn(const void* q, const void* k_cache, const void* v_cache,
                              const int* cache_len_ptr, const int* pad_left, void* scratch, void* out,
                              int B, int nh, int nkv, int hd, int T, int n_new, float scale, int splits,
                              int q_bstride, void* stream) {
  JL_CHECK_ARG(nkv > 0 && nh % nkv == 0, "jl_decode_attn: nh=%d not a multiple of nkv=%d", nh, nkv);
  JL_CHECK_ARG(splits >= 1 && splits <= 64, "jl_decode_attn: splits=%d outside 1..64", splits);
  JL_CHECK_ARG(splits == 1 || scratch != nullptr, "jl_decode_attn: scratch buffer required for splits > 1");
  if (B == 0) return JL_OK;
  const int G = nh / nkv;
  float* o_part = (float*)scratch;
  float* ml_part = o_part ? o_part + (size_t)B * nh * splits * hd : nullptr;
  auto s = (cudaStream_t)stream;
  cudaError_t e = cudaErrorInvalidValue;
#define JL_DA(G_, HD_)                                                                              \
  if (G == G_ && hd == HD_)                                                                         \
    e = launch_decode<G_, HD_>(q, k_cache, v_cache, cache_len_ptr, pad_left, o_part, ml_part, out, B, \
                               nkv, T, n_new, scale, splits, q_bstride > 0 ? q_bstride : nh * hd, s);
  JL_DA(1, 128) else JL_DA(2, 128) else JL_DA(4, 128) else JL_DA(8, 128)
  else JL_DA(1, 64) else JL_DA(2, 64) else JL_DA(4, 64) else JL_DA(8, 64)
  else {
    set_error("jl_decode_attn: unsupported GQA group %d / head_dim %d (groups 1,2,4,8; head_dim 64,128)", G, hd);
    return JL_ERR_UNSUPPORTED;
  }
#undef JL_DA
  JL_CUDA(e);
  return JL_OK;
}

// ------------------------------------------------------------------------------------ allocation-free variant
// Split-KV decode attention for callers that cannot allocate device memory.  The per-split partial
// results live in the fixed array below instead of a caller-supplied buffer.  The array is part of
// the library image: every GPU that loads libjllama_b200.so carries these JL_NOALLOC_SCRATCH_FLOATS
// floats (4 MiB) from load until the process exits, whether or not this entry point is ever used.
//
// `g_noalloc_scratch` names the array for DEVICE code.  In host code the same identifier is only a
// host-side stand-in for the symbol, not a device address, so the host never hands it to a kernel: once
// per device (with that device current) it asks the runtime for the real address with
// cudaGetSymbolAddress and passes that pointer instead.
#define JL_NOALLOC_SCRATCH_FLOATS (1u << 20)
static __device__ __align__(256) float g_noalloc_scratch[JL_NOALLOC_SCRATCH_FLOATS];

namespace jl {
constexpr int NOALLOC_MAX_DEVICES = 64;
static float* g_noalloc_dev_ptr[NOALLOC_MAX_DEVICES];     // resolved device address of the array, per device
static bool g_noalloc_ready[NOALLOC_MAX_DEVICES];
static std::mutex g_noalloc_mu;

// Resolve (once per device) where the static array lives on the CURRENT device.
static int noalloc_scratch_for_current_device(float** out) {
  int dev = -1;
  JL_CUDA(cudaGetDevice(&dev));
  JL_CHECK_ARG(dev >= 0 && dev < NOALLOC_MAX_DEVICES, "jl_decode_attn_noalloc: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_ready[dev]) {
    float* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_noalloc_scratch)");
    JL_CHECK_ARG(ptr != nullptr, "jl_decode_attn_noalloc: runtime returned a null address for the scratch array");
    g_noalloc_dev_ptr[dev] = ptr;
    g_noalloc_ready[dev] = true;
  }
  *out = g_noalloc_dev_ptr[dev];
  return JL_OK;
}
}  // namespace jl

extern "C" JL_API size_t jl_decode_attn_noalloc_max_floats(void) { return JL_NOALLOC_SCRATCH_FLOATS; }

// jl_decode_attn without the `scratch` argument.
//
// Who needs it: code that drives these kernels from a place where it may not, or cannot, obtain device
// memory — a serving process whose HBM is already owned end to end by another runtime's arena, a plugin
// running under a host that forbids cudaMalloc/cudaFree in the token loop because those calls take the
// allocator lock and can synchronise the device, or a thin C caller that has no allocator at all.  Such
// callers still own q, the KV cache and `out`; only the split-KV workspace comes from the library.
//
// Contract
//  * Nothing is allocated or freed here and no allocating library routine is called (no cudaMalloc,
//    cudaFree, no torch).  The workspace is the file-scope array above.
//  * Limit: B*nh*splits*(hd+2) floats must be <= jl_decode_attn_noalloc_max_floats() (2^20 floats = 4 MiB);
//    larger requests fail with JL_ERR_ARG and a message that says so.  splits == 1 needs no workspace.
//  * The first call on each device resolves the array's device address (cudaGetSymbolAddress, with that
//    device current).  Make that first call ahead of time, before any stream capture begins.
//  * All calls on a device share the one array and it keeps whatever the previous call left in it.
//    Nothing in it has to start at zero: every (m, l, o) slot a call reads was written earlier in that
//    same call, so stale contents are never consumed and no clearing pass is needed.
//  * Because the array is shared, two calls whose kernels can overlap on the GPU (different streams,
//    or different host threads on different streams) race on it.  Serialise them on one stream.
//  * Results are identical to jl_decode_attn given the same `splits`.
extern "C" JL_API int jl_decode_attn_noalloc(const void* q, const void* k_cache, const void* v_cache,
                                             const int* cache_len_ptr, const int* pad_left, void* out, int B,
                                             int nh, int nkv, int hd, int T, int n_new, float scale, int splits,
                                             int q_bstride, void* stream) {
  JL_CHECK_ARG(nkv > 0 && nh % nkv == 0, "jl_decode_attn_noalloc: nh=%d not a multiple of nkv=%d", nh, nkv);
  JL_CHECK_ARG(splits >= 1 && splits <= 64, "jl_decode_attn_noalloc: splits=%d outside 1..64", splits);
  JL_CHECK_ARG(B >= 0 && hd > 0, "jl_decode_attn_noalloc: bad shape B=%d hd=%d", B, hd);
  const size_t need = splits <= 1 ? 0 : (size_t)B * nh * splits * (hd + 2);
  JL_CHECK_ARG(need <= JL_NOALLOC_SCRATCH_FLOATS,
               "jl_decode_attn_noalloc: B*nh*splits*(hd+2) = %zu floats exceeds the built-in scratch of %u floats; "
               "lower `splits` or use jl_decode_attn with a caller-provided buffer",
               need, (unsigned)JL_NOALLOC_SCRATCH_FLOATS);
  if (B == 0) return JL_OK;
  float* scratch = nullptr;
  int rc = noalloc_scratch_for_current_device(&scratch);     // first call on a device does the set-up
  if (rc != JL_OK) return rc;
  const int G = nh / nkv;
  float* o_part = scratch;
  float* ml_part = scratch + (size_t)B * nh * splits * hd;
  auto s = (cudaStream_t)stream;
  cudaError_t e = cudaErrorInvalidValue;
#define JL_DA(G_, HD_)                                                                              \
  if (G == G_ && hd == HD_)                                                                         \
    e = launch_decode<G_, HD_>(q, k_cache, v_cache, cache_len_ptr, pad_left, o_part, ml_part, out, B, \
                               nkv, T, n_new, scale, splits, q_bstride > 0 ? q_bstride : nh * hd, s);
  JL_DA(1, 128) else JL_DA(2, 128) else JL_DA(4, 128) else JL_DA(8, 128)
  else JL_DA(1, 64) else JL_DA(2, 64) else JL_DA(4, 64) else JL_DA(8, 64)
  else {
    set_error("jl_decode_attn_noalloc: unsupported GQA group %d / head_dim %d (groups 1,2,4,8; head_dim 64,128)", G, hd);
    return JL_ERR_UNSUPPORTED;
  }
#undef JL_DA
  JL_CUDA(e);
  return JL_OK;
}
