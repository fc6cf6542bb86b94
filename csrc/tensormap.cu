// Host-side TMA descriptor construction.  cuTensorMapEncodeTiled is resolved through the runtime
// (cudaGetDriverEntryPoint) so the library has no link-time dependency on libcuda.
#include <mutex>

#include "tcgen05.cuh"

namespace jl {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

static int encode(CUtensorMap* out, const void* base, int rank, const cuuint64_t* dims,
                  const cuuint64_t* strides_bytes, const cuuint32_t* box,
                  CUtensorMapSwizzle swizzle = CU_TENSOR_MAP_SWIZZLE_128B) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from this driver");
    return JL_ERR_CUDA;
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15) != 0) {
    set_error("TMA base address %p is not 16-byte aligned", base);
    return JL_ERR_ARG;
  }
  for (int i = 0; i < rank - 1; ++i)
    if (strides_bytes[i] % 16 != 0) {
      set_error("TMA stride %llu bytes is not a multiple of 16", (unsigned long long)strides_bytes[i]);
      return JL_ERR_ARG;
    }
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), dims,
                  strides_bytes, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rank %d dims %llu,%llu box %u,%u)", (int)r, rank,
              (unsigned long long)dims[0], (unsigned long long)dims[1], box[0], box[1]);
    return JL_ERR_CUDA;
  }
  return JL_OK;
}

int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld,
                      uint32_t box_rows) {
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {64, box_rows};
  return encode(out, base, 2, dims, strides, box);
}

int make_tmap_2d_bf16_plain(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld,
                            uint32_t box_cols, uint32_t box_rows) {
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  return encode(out, base, 2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_NONE);
}

// [rows, cols] viewed as [rows, cols/256, 256]: one request moves box_rows x box_slabs slabs of 256 columns
int make_tmap_slab_bf16_plain(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld,
                              uint32_t box_slabs, uint32_t box_rows) {
  if (cols % 256 != 0) {
    set_error("slab tensor map needs a column count that is a multiple of 256 (got %llu)", (unsigned long long)cols);
    return JL_ERR_ARG;
  }
  cuuint64_t dims[3] = {256, cols / 256, rows};
  cuuint64_t strides[2] = {256 * 2, ld * 2};
  cuuint32_t box[3] = {256, box_slabs, box_rows};
  return encode(out, base, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_NONE);
}

// [rows, cols] viewed as [cols/64 k-blocks][rows][64]: one request moves box_kblocks x box_rows rows of 128 B,
// stored k-block-major with the 128-byte swizzle — the canonical K-major tile ldmatrix / UMMA read conflict-free
int make_tmap_kblock_bf16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld,
                          uint32_t box_kblocks, uint32_t box_rows) {
  if (cols % 64 != 0) {
    set_error("k-block tensor map needs a column count that is a multiple of 64 (got %llu)", (unsigned long long)cols);
    return JL_ERR_ARG;
  }
  cuuint64_t dims[3] = {64, rows, cols / 64};
  cuuint64_t strides[2] = {ld * 2, 64 * 2};
  cuuint32_t box[3] = {64, box_rows, box_kblocks};
  return encode(out, base, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

int make_tmap_3d_bf16(CUtensorMap* out, const void* base, uint64_t d2, uint64_t rows, uint64_t cols,
                      uint64_t ld2, uint64_t ld1, uint32_t box_rows) {
  cuuint64_t dims[3] = {cols, rows, d2};
  cuuint64_t strides[2] = {ld1 * 2, ld2 * 2};
  cuuint32_t box[3] = {64, box_rows, 1};
  return encode(out, base, 3, dims, strides, box);
}

}  // namespace jl
