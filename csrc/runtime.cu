// harp_b200 runtime glue: error strings, TMA descriptor encoding through the
// driver entry point (no link-time libcuda dependency), device queries.
#include "common.cuh"
#include "../include/harp_b200.h"

#include <stdarg.h>
#include <string.h>
#include <mutex>

namespace hb {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess) {
      fn = reinterpret_cast<EncodeTiledFn>(p);
    }
  });
  return fn;
}

int make_tmap_2d_f32(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols,
                     uint64_t row_stride_bytes, uint32_t box_rows, uint32_t box_cols,
                     CUtensorMapSwizzle swizzle) {
  EncodeTiledFn fn = encode_fn();
  HB_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
  HB_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base must be 16-byte aligned");
  HB_REQUIRE((row_stride_bytes & 15) == 0, "TMA row stride must be a multiple of 16 bytes (got %llu)",
             (unsigned long long)row_stride_bytes);
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {row_stride_bytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), gdim, gstride,
                  box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  HB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return 0;
}

int make_tmap_2d_u16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols,
                     uint64_t row_stride_bytes, uint32_t box_rows, uint32_t box_cols,
                     CUtensorMapSwizzle swizzle) {
  EncodeTiledFn fn = encode_fn();
  HB_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled is not available from this driver");
  HB_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base must be 16-byte aligned");
  HB_REQUIRE((row_stride_bytes & 15) == 0, "TMA row stride must be a multiple of 16 bytes (got %llu)",
             (unsigned long long)row_stride_bytes);
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {row_stride_bytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, const_cast<void*>(base), gdim, gstride,
                  box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  HB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled (16-bit) failed with CUresult %d", (int)r);
  return 0;
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

}  // namespace hb

extern "C" {

const char* hb_last_error(void) { return hb::g_err; }

int hb_version(void) { return 100; }

int hb_device_sm_count(void) { return hb::sm_count(); }

// Writes a buffer larger than L2 so that the next timed kernel starts cold.
__global__ void hb_fill_kernel(float4* p, size_t n, float v) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) p[i] = make_float4(v, v, v, v);
}

int hb_l2_flush(void* buf, size_t bytes, void* stream) {
  HB_REQUIRE(buf != nullptr && bytes >= 16, "hb_l2_flush: empty buffer");
  hb_fill_kernel<<<hb::sm_count() * 4, 256, 0, (cudaStream_t)stream>>>((float4*)buf, bytes / 16,
                                                                        0.f);
  HB_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"
