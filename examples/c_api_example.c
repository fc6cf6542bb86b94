/* Using libjllama_b200.so from plain C: one transformer-style step on raw device pointers.
 *
 *   RMSNorm → fused qkv projection (jl_gemm, swap-AB path for a batch of 16 tokens) → RoPE + KV append + GQA
 *   decode attention in one launch (jl_decode_attn_fused) → output projection with residual add.
 *
 * Build (the library is self-contained apart from the CUDA driver; the example itself needs the runtime for
 * cudaMalloc / cudaMemcpy):
 *   gcc -std=c99 -Iinclude -I/usr/local/cuda/include examples/c_api_example.c \
 *       -Ljax_llama_b200/ops/_C -ljllama_b200 -L/usr/local/cuda/lib64 -lcudart -lm -o c_api_example
 *   LD_LIBRARY_PATH=jax_llama_b200/ops/_C ./c_api_example
 *
 * Everything is asynchronous on one stream; consecutive calls are chained with programmatic dependent launch by
 * the library, so no host synchronisation is needed between them.  (tests/test_harness_cli.py compiles and links
 * this file on the CPU; running it needs a B200.)
 */
#include <cuda_runtime_api.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "jllama_b200.h"

#define CK(call)                                                                                  \
  do {                                                                                            \
    cudaError_t e_ = (call);                                                                      \
    if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #call, cudaGetErrorString(e_)); return 1; } \
  } while (0)
#define JL(call)                                                                          \
  do {                                                                                    \
    if ((call) != 0) { fprintf(stderr, "%s: %s\n", #call, jl_last_error()); return 1; }  \
  } while (0)

/* float → bf16 (round to nearest even), host side */
static uint16_t bf16(float f) {
  union { float f; uint32_t u; } v;
  v.f = f;
  v.u += 0x7fffu + ((v.u >> 16) & 1u);
  return (uint16_t)(v.u >> 16);
}

static void* device_bf16(size_t n, float scale, unsigned seed) {
  uint16_t* h = (uint16_t*)malloc(n * sizeof(uint16_t));
  void* d = NULL;
  size_t i;
  srand(seed);
  for (i = 0; i < n; ++i) h[i] = bf16(scale * ((float)rand() / (float)RAND_MAX - 0.5f));
  if (cudaMalloc(&d, n * sizeof(uint16_t)) != cudaSuccess) { free(h); return NULL; }
  cudaMemcpy(d, h, n * sizeof(uint16_t), cudaMemcpyHostToDevice);
  free(h);
  return d;
}

int main(void) {
  /* a LLaMA-3-8B-shaped layer: hidden 4096, 32 query heads, 8 kv heads, head_dim 128; 16 sequences, 512 keys each */
  enum { B = 16, H = 4096, NH = 32, NKV = 8, HD = 128, T = 1024, FILLED = 512, MAXPOS = 2048 };
  const int qkv_cols = (NH + 2 * NKV) * HD;
  cudaStream_t stream;
  void *x, *xn, *norm_w, *wqkv, *wo, *qkv, *attn, *kc, *vc, *scratch;
  float *cos_t, *sin_t, *h_cos, *h_sin;
  int *cache_len, *pad_left;
  int filled = FILLED, splits, i, j;
  size_t scratch_bytes;
  uint16_t first[4];

  printf("libjllama_b200 version %d\n", jl_version());
  CK(cudaStreamCreate(&stream));
  x = device_bf16((size_t)B * H, 2.0f, 1);
  norm_w = device_bf16(H, 0.2f, 2);
  wqkv = device_bf16((size_t)qkv_cols * H, 0.04f, 3);
  wo = device_bf16((size_t)H * NH * HD, 0.04f, 4);
  kc = device_bf16((size_t)B * NKV * T * HD, 1.0f, 5);
  vc = device_bf16((size_t)B * NKV * T * HD, 1.0f, 6);
  if (!x || !norm_w || !wqkv || !wo || !kc || !vc) { fprintf(stderr, "cudaMalloc failed\n"); return 1; }
  CK(cudaMalloc(&xn, (size_t)B * H * 2));
  CK(cudaMalloc(&qkv, (size_t)B * qkv_cols * 2));
  CK(cudaMalloc(&attn, (size_t)B * NH * HD * 2));
  CK(cudaMalloc((void**)&cache_len, sizeof(int)));
  CK(cudaMalloc((void**)&pad_left, B * sizeof(int)));
  CK(cudaMemcpy(cache_len, &filled, sizeof(int), cudaMemcpyHostToDevice));
  CK(cudaMemset(pad_left, 0, B * sizeof(int)));

  /* RoPE tables: fp32 [max_pos, head_dim / 2] */
  h_cos = (float*)malloc((size_t)MAXPOS * HD / 2 * sizeof(float));
  h_sin = (float*)malloc((size_t)MAXPOS * HD / 2 * sizeof(float));
  for (i = 0; i < MAXPOS; ++i)
    for (j = 0; j < HD / 2; ++j) {
      const double a = (double)i * pow(500000.0, -2.0 * j / HD);
      h_cos[(size_t)i * HD / 2 + j] = (float)cos(a);
      h_sin[(size_t)i * HD / 2 + j] = (float)sin(a);
    }
  CK(cudaMalloc((void**)&cos_t, (size_t)MAXPOS * HD / 2 * sizeof(float)));
  CK(cudaMalloc((void**)&sin_t, (size_t)MAXPOS * HD / 2 * sizeof(float)));
  CK(cudaMemcpy(cos_t, h_cos, (size_t)MAXPOS * HD / 2 * sizeof(float), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(sin_t, h_sin, (size_t)MAXPOS * HD / 2 * sizeof(float), cudaMemcpyHostToDevice));

  /* split-KV scratch of the fused decode attention: sized by the library, zeroed once (ticket counters) */
  splits = jl_decode_attn_fused_splits(B, NKV, T, 0);
  scratch_bytes = jl_decode_attn_fused_scratch_bytes(B, NH, NKV, HD, splits);
  CK(cudaMalloc(&scratch, scratch_bytes));
  CK(cudaMemset(scratch, 0, scratch_bytes));

  /* the step: four launches on one stream, no host synchronisation in between */
  JL(jl_rmsnorm(x, NULL, norm_w, xn, NULL, B, H, 1e-5f, stream));
  JL(jl_gemm(xn, wqkv, NULL, qkv, B, qkv_cols, H, /*mode: plain*/ 0, /*fp32 out*/ 0, /*SMs: 0 = B200 (148)*/ 0, stream));
  JL(jl_decode_attn_fused(qkv, kc, vc, cos_t, sin_t, cache_len, pad_left, scratch, attn, B, NH, NKV, HD, T, MAXPOS,
                          1.0f / sqrtf((float)HD), splits, stream));
  JL(jl_gemm(attn, wo, /*residual*/ x, /*out*/ x, B, H, NH * HD, /*mode: residual add*/ 1, 0, 0, stream));
  CK(cudaStreamSynchronize(stream));

  CK(cudaMemcpy(first, x, sizeof(first), cudaMemcpyDeviceToHost));
  printf("kernel launches issued by the library: %llu; first residual values (bf16 bits): %04x %04x %04x %04x\n",
         jl_launch_count(), first[0], first[1], first[2], first[3]);
  free(h_cos);
  free(h_sin);
  return 0;
}
