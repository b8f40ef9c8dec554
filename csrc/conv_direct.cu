// K3: depthwise 3x3 (+bias +ReLU) and K4: 3x3/s2 patch gather, NHWC fp16.
//
// Depthwise layers have 9 MACs per output and no reuse across channels, so they are pure
// bandwidth: each thread owns 8 channels (one 16-byte vector) of a strip of 4 output
// pixels, slides the 3x3 window along the strip in registers and leans on L1 for the
// row overlap.  Accumulation is fp32.  Reference op: cv2.dnn Convolution with
// group == channels, prototxt:42-75, 107-141, ... 826-859 (engine: CAFFE, prototxt:60).
#include "common.cuh"

namespace dvc {

constexpr int DW_TW = 4;
constexpr int DW_THREADS = 256;

__device__ __forceinline__ void cvt8(const uint4& u, float* f) {
  float2 a = unpack_half2(u.x), b = unpack_half2(u.y), c = unpack_half2(u.z), d = unpack_half2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}

template <int STRIDE>
__global__ void __launch_bounds__(DW_THREADS)
dw3x3_kernel(const __half* __restrict__ in, int B, int H, int W, int C, int Ho, int Wo, const float* __restrict__ w,
             const float* __restrict__ bias, int relu, __half* __restrict__ out, long long total) {
  constexpr int NCOL = (DW_TW - 1) * STRIDE + 3;
  const long long idx = (long long)blockIdx.x * DW_THREADS + threadIdx.x;
  if (idx >= total) return;
  const int CG = C >> 3;
  const int spr = (Wo + DW_TW - 1) / DW_TW;
  const int cg = (int)(idx % CG);
  long long r = idx / CG;
  const int xs = (int)(r % spr); r /= spr;
  const int y = (int)(r % Ho);
  const int b = (int)(r / Ho);
  const int c0 = cg * 8;
  const int x0 = xs * DW_TW;

  float acc[DW_TW][8];
  {
    const float4 b0 = __ldg(reinterpret_cast<const float4*>(bias + c0));
    const float4 b1 = __ldg(reinterpret_cast<const float4*>(bias + c0 + 4));
#pragma unroll
    for (int t = 0; t < DW_TW; ++t) {
      acc[t][0] = b0.x; acc[t][1] = b0.y; acc[t][2] = b0.z; acc[t][3] = b0.w;
      acc[t][4] = b1.x; acc[t][5] = b1.y; acc[t][6] = b1.z; acc[t][7] = b1.w;
    }
  }
  const __half* in_b = in + (size_t)b * H * W * C;
#pragma unroll
  for (int ky = 0; ky < 3; ++ky) {
    const int iy = y * STRIDE - 1 + ky;
    if (iy < 0 || iy >= H) continue;
    float wr[3][8];
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      const float4 w0 = __ldg(reinterpret_cast<const float4*>(w + (size_t)(ky * 3 + kx) * C + c0));
      const float4 w1 = __ldg(reinterpret_cast<const float4*>(w + (size_t)(ky * 3 + kx) * C + c0 + 4));
      wr[kx][0] = w0.x; wr[kx][1] = w0.y; wr[kx][2] = w0.z; wr[kx][3] = w0.w;
      wr[kx][4] = w1.x; wr[kx][5] = w1.y; wr[kx][6] = w1.z; wr[kx][7] = w1.w;
    }
    const __half* row = in_b + (size_t)iy * W * C + c0;
    uint4 raw[NCOL];
#pragma unroll
    for (int col = 0; col < NCOL; ++col) {
      const int ix = x0 * STRIDE - 1 + col;
      raw[col] = (ix >= 0 && ix < W) ? __ldg(reinterpret_cast<const uint4*>(row + (size_t)ix * C)) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int col = 0; col < NCOL; ++col) {
      float v[8];
      cvt8(raw[col], v);
#pragma unroll
      for (int t = 0; t < DW_TW; ++t) {
        const int kx = col - t * STRIDE;
        if (kx >= 0 && kx < 3) {
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[t][j] = fmaf(v[j], wr[kx][j], acc[t][j]);
        }
      }
    }
  }
  __half* orow = out + ((size_t)(b * Ho + y) * Wo) * C + c0;
#pragma unroll
  for (int t = 0; t < DW_TW; ++t) {
    const int x = x0 + t;
    if (x < Wo) {
      float* a = acc[t];
      if (relu) {
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] = fmaxf(a[j], 0.f);
      }
      *reinterpret_cast<uint4*>(orow + (size_t)x * C) =
          make_uint4(pack_half2(a[0], a[1]), pack_half2(a[2], a[3]), pack_half2(a[4], a[5]), pack_half2(a[6], a[7]));
    }
  }
}

int dwconv_launch(const void* in, int B, int H, int W, int C, int stride, const float* w, const float* b, int relu,
                  void* out, cudaStream_t st) {
  DVC_REQUIRE(C % 8 == 0 && (stride == 1 || stride == 2), "dwconv: C=%d must be a multiple of 8, stride=%d in {1,2}", C, stride);
  DVC_REQUIRE(B > 0 && H > 0 && W > 0, "dwconv: bad shape");
  const int Ho = (H + 2 - 3) / stride + 1, Wo = (W + 2 - 3) / stride + 1;
  const long long total = (long long)B * Ho * ((Wo + DW_TW - 1) / DW_TW) * (C / 8);
  const int grid = (int)ceil_div64(total, DW_THREADS);
  if (stride == 1)
    dw3x3_kernel<1><<<grid, DW_THREADS, 0, st>>>((const __half*)in, B, H, W, C, Ho, Wo, w, b, relu, (__half*)out, total);
  else
    dw3x3_kernel<2><<<grid, DW_THREADS, 0, st>>>((const __half*)in, B, H, W, C, Ho, Wo, w, b, relu, (__half*)out, total);
  count_launch();
  DVC_CHECK_CUDA(cudaGetLastError());
  return DVC_OK;
}

// ---- im2col for the dense 3x3 stride-2 pad-1 extras (conv14_2 .. conv17_2, prototxt:922-1146)
__global__ void __launch_bounds__(256)
im2col3x3s2_kernel(const __half* __restrict__ in, int B, int H, int W, int C, int Ho, int Wo, __half* __restrict__ out,
                   long long total) {
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= total) return;
  const int CG = C >> 3;
  const int cg = (int)(idx % CG);
  long long r = idx / CG;
  const int tap = (int)(r % 9); r /= 9;
  const int ox = (int)(r % Wo); r /= Wo;
  const int oy = (int)(r % Ho);
  const int b = (int)(r / Ho);
  const int iy = oy * 2 - 1 + tap / 3, ix = ox * 2 - 1 + tap % 3;
  uint4 v = make_uint4(0, 0, 0, 0);
  if (iy >= 0 && iy < H && ix >= 0 && ix < W)
    v = __ldg(reinterpret_cast<const uint4*>(in + (((size_t)b * H + iy) * W + ix) * C + cg * 8));
  const size_t m = ((size_t)b * Ho + oy) * Wo + ox;
  *reinterpret_cast<uint4*>(out + m * (size_t)(9 * C) + (size_t)tap * C + cg * 8) = v;
}

int im2col_launch(const void* in, int B, int H, int W, int C, void* out, cudaStream_t st) {
  DVC_REQUIRE(C % 8 == 0 && B > 0 && H > 0 && W > 0, "im2col: bad shape C=%d", C);
  const int Ho = (H + 2 - 3) / 2 + 1, Wo = (W + 2 - 3) / 2 + 1;
  const long long total = (long long)B * Ho * Wo * 9 * (C / 8);
  im2col3x3s2_kernel<<<(int)ceil_div64(total, 256), 256, 0, st>>>((const __half*)in, B, H, W, C, Ho, Wo, (__half*)out, total);
  count_launch();
  DVC_CHECK_CUDA(cudaGetLastError());
  return DVC_OK;
}

}  // namespace dvc

extern "C" int dvc_dwconv3x3(const void* in_f16, int B, int H, int W, int C, int stride, const float* w, const float* b,
                             int relu, void* out_f16, void* stream) {
  return dvc::dwconv_launch(in_f16, B, H, W, C, stride, w, b, relu, out_f16, (cudaStream_t)stream);
}
extern "C" int dvc_im2col3x3s2(const void* in_f16, int B, int H, int W, int C, void* out_f16, void* stream) {
  return dvc::im2col_launch(in_f16, B, H, W, C, out_f16, (cudaStream_t)stream);
}
