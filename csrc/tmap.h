// Host-side helpers: CUDA error checks and TMA tensor-map creation.
// cuTensorMapEncodeTiled is resolved through cudaGetDriverEntryPoint so the library does not
// link against libcuda directly.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>

#define GF_OK 0
#define GF_ERR_CUDA 1
#define GF_ERR_ARG 2
#define GF_ERR_DRIVER 3

#define GF_CUDA_CHECK(expr)                                                              \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess) {                                                             \
      fprintf(stderr, "[gfedntm] CUDA error %s at %s:%d: %s\n", cudaGetErrorName(_e),    \
              __FILE__, __LINE__, cudaGetErrorString(_e));                               \
      return GF_ERR_CUDA;                                                                \
    }                                                                                    \
  } while (0)

// after every kernel launch: surface launch errors and count the launch
#define GF_LAUNCHED()                      \
  do {                                     \
    GF_CUDA_CHECK(cudaGetLastError());     \
    gf::g_launches.fetch_add(1);           \
  } while (0)

#define GF_ARG_CHECK(cond, msg)                                                          \
  do {                                                                                   \
    if (!(cond)) {                                                                       \
      fprintf(stderr, "[gfedntm] bad argument at %s:%d: %s\n", __FILE__, __LINE__, msg); \
      return GF_ERR_ARG;                                                                 \
    }                                                                                    \
  } while (0)

namespace gf {

extern std::atomic<long long> g_launches;

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  }
  return fn;
}

// 2D bf16 tensor, `inner` contiguous elements per row, `outer` rows, `pitch_bytes` between rows.
// Box = box_inner x box_outer elements, 128B swizzle, zero fill out of bounds.
inline int make_tmap_bf16_2d(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer,
                             uint64_t pitch_bytes, uint32_t box_inner, uint32_t box_outer) {
  PFN_encodeTiled fn = get_encode_fn();
  if (!fn) {
    fprintf(stderr, "[gfedntm] cuTensorMapEncodeTiled not available\n");
    return GF_ERR_DRIVER;
  }
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {pitch_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims,
                  strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    fprintf(stderr,
            "[gfedntm] cuTensorMapEncodeTiled failed (%d): inner=%llu outer=%llu pitch=%llu "
            "box=%ux%u base=%p\n",
            (int)r, (unsigned long long)inner, (unsigned long long)outer,
            (unsigned long long)pitch_bytes, box_inner, box_outer, base);
    return GF_ERR_DRIVER;
  }
  return GF_OK;
}

inline int num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

}  // namespace gf
