// Library plumbing: error strings, launch counter, TMA descriptor encoding.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>

#include "common.cuh"

namespace dvc {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

bool pdl_enabled() {
  const char* e = getenv("DVC_PDL");          // read per launch: a CUDA graph keeps whatever was set at capture
  return !(e && e[0] == '0');
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn resolve_encode_fn() {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult qres;
  // The driver symbol is resolved through the runtime so libcuda is not a link-time dependency
  // (the build box has no driver).
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
      qres == cudaDriverEntryPointSuccess) {
    return reinterpret_cast<EncodeTiledFn>(p);
  }
  return nullptr;
}

static EncodeTiledFn get_encode_fn() {
  // function-local static: initialised exactly once even when several engine threads build plans at the same time
  static const EncodeTiledFn fn = resolve_encode_fn();
  return fn;
}

int make_tmap_2d_f16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld_elems,
                     uint32_t box_rows, uint32_t box_cols) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is unavailable (no CUDA driver?)");
    return DVC_ERR_CUDA;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld_elems * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed: CUresult %d (rows=%llu cols=%llu ld=%llu box=%ux%u)", (int)r,
              (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld_elems, box_rows, box_cols);
    return DVC_ERR_CUDA;
  }
  return DVC_OK;
}

int make_tmap_nhwc_f16(CUtensorMap* out, const void* base, uint64_t n, uint64_t h, uint64_t w, uint64_t c,
                       uint32_t box_c, uint32_t box_w, uint32_t box_h) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is unavailable (no CUDA driver?)");
    return DVC_ERR_CUDA;
  }
  cuuint64_t dims[4] = {c, w, h, n};
  cuuint64_t strides[3] = {c * 2, w * c * 2, h * w * c * 2};
  cuuint32_t box[4] = {box_c, box_w, box_h, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled(4d) failed: CUresult %d (n=%llu h=%llu w=%llu c=%llu box=%ux%ux%u)", (int)r,
              (unsigned long long)n, (unsigned long long)h, (unsigned long long)w, (unsigned long long)c, box_c, box_w, box_h);
    return DVC_ERR_CUDA;
  }
  return DVC_OK;
}

// Same 4-D NHWC view with 128-byte swizzle: for TMA *stores* of epilogue tiles staged as 128-byte rows
// (box_c must be 64 fp16 = 128 bytes); coordinates outside the tensor are clipped by the hardware.
int make_tmap_nhwc_f16_sw128(CUtensorMap* out, const void* base, uint64_t n, uint64_t h, uint64_t w, uint64_t c,
                             uint32_t box_c, uint32_t box_w, uint32_t box_h) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is unavailable (no CUDA driver?)");
    return DVC_ERR_CUDA;
  }
  if (box_c * 2 != 128) {
    set_error("make_tmap_nhwc_f16_sw128: box_c=%u must be 64 (128-byte swizzle rows)", box_c);
    return DVC_ERR_ARG;
  }
  cuuint64_t dims[4] = {c, w, h, n};
  cuuint64_t strides[3] = {c * 2, w * c * 2, h * w * c * 2};
  cuuint32_t box[4] = {box_c, box_w, box_h, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled(4d, sw128) failed: CUresult %d (n=%llu h=%llu w=%llu c=%llu box=%ux%ux%u)", (int)r,
              (unsigned long long)n, (unsigned long long)h, (unsigned long long)w, (unsigned long long)c, box_c, box_w, box_h);
    return DVC_ERR_CUDA;
  }
  return DVC_OK;
}

typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeIm2colFn resolve_im2col_fn() {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &p, cudaEnableDefault, &qres) == cudaSuccess &&
      qres == cudaDriverEntryPointSuccess) {
    return reinterpret_cast<EncodeIm2colFn>(p);
  }
  return nullptr;
}

// NHWC fp16 activations as an im2col-mode TMA descriptor for a 3x3 / stride-2 / pad-1 convolution.  One load fetches
// `pixels` consecutive OUTPUT pixels (w fastest, then h, then n) x `box_c` channels of one filter tap; the tap is the
// instruction's {kw, kh} offset pair and the start pixel's coordinate is (2*ow - 1, 2*oh - 1, n).  The pixel box
// [-1, W-2] x [-1, H-2] stepped by 2 holds exactly ceil(W/2) x ceil(H/2) base positions; taps that land in the padding
// halo read as zeros.  Rows are 128 bytes (box_c = 64) with the 128-byte swizzle tcgen05.mma operands use.
int make_tmap_im2col3x3s2_f16(CUtensorMap* out, const void* base, uint64_t n, uint64_t h, uint64_t w, uint64_t c,
                              uint32_t box_c, uint32_t pixels) {
  static const EncodeIm2colFn fn = resolve_im2col_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeIm2col is unavailable (no CUDA driver?)");
    return DVC_ERR_CUDA;
  }
  if (box_c * 2 != 128 || pixels == 0 || pixels > 1024 || h < 2 || w < 2) {
    set_error("make_tmap_im2col3x3s2_f16: box_c=%u pixels=%u h=%llu w=%llu unsupported", box_c, pixels,
              (unsigned long long)h, (unsigned long long)w);
    return DVC_ERR_ARG;
  }
  cuuint64_t dims[4] = {c, w, h, n};
  cuuint64_t strides[3] = {c * 2, w * c * 2, h * w * c * 2};
  int lower[2] = {-1, -1};          // {w, h}: -pad
  int upper[2] = {-1, -1};          // {w, h}: pad - (kernel - 1)
  cuuint32_t estr[4] = {1, 2, 2, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, const_cast<void*>(base), dims, strides, lower, upper, box_c, pixels,
                  estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeIm2col failed: CUresult %d (n=%llu h=%llu w=%llu c=%llu box_c=%u pixels=%u)", (int)r,
              (unsigned long long)n, (unsigned long long)h, (unsigned long long)w, (unsigned long long)c, box_c, pixels);
    return DVC_ERR_CUDA;
  }
  return DVC_OK;
}

}  // namespace dvc

extern "C" int dvc_version(void) { return 100; }
/* This library links its own (static) CUDA runtime, whose per-thread current device is separate from the
 * one PyTorch's runtime tracks.  Callers that drive several GPUs from one process select the device here
 * before preparing or launching anything on it. */
extern "C" int dvc_set_device(int device) {
  DVC_CHECK_CUDA(cudaSetDevice(device));
  return DVC_OK;
}
extern "C" const char* dvc_last_error(void) { return dvc::g_err; }
extern "C" long long dvc_launch_count(void) { return dvc::g_launches.load(); }
