// K3: depthwise 3x3 (+bias +ReLU) and K4: 3x3/s2 patch gather, NHWC fp16.
//
// Depthwise layers have 9 MACs per output and no reuse across channels, so they are pure
// bandwidth: each thread owns 8 channels (one 16-byte vector) of a strip of 4 output
// pixels, slides the 3x3 window along the strip in registers and leans on L1 for the
// row overlap.  Accumulation is fp32.  Reference op: cv2.dnn Convolution with
// group == channels, prototxt:42-75, 107-141, ... 826-859 (engine: CAFFE, prototxt:60).
#include <atomic>

#include "common.cuh"

namespace dvc {

// ---- TMA-fed depthwise: one CTA per (image, spatial tile, 64-channel chunk).  A single 4-D TMA box
// load brings the input tile plus its halo into shared memory (out-of-range coordinates read as
// zeros, so the pad-1 border costs nothing); the loads of several co-resident CTAs overlap each
// other's math.  Each thread owns 8 channels and walks the tile's pixels: 9 x LDS.128 and 72
// FHFMA (fp16 x fp16 + fp32) per output vector, no convert instructions.
struct DwArgs {
  int B, H, W, C, Ho, Wo;
  int CT, TX, TY, BW, BH, tiles_x, tiles_y, cchunks;
  const float* w;
  const float* bias;
  int relu;
  __half* out;
};

template <int STRIDE, int NPX>
__global__ void __launch_bounds__(256) dw3x3_tma_kernel(const __grid_constant__ CUtensorMap tm, const DwArgs a) {
  extern __shared__ uint8_t dw_smem_raw[];
  uint8_t* tile = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(dw_smem_raw) + 127) & ~uintptr_t(127));
  __shared__ uint64_t bar;
  const int tid = threadIdx.x;
  int bid = blockIdx.x;
  const int cc = bid % a.cchunks; bid /= a.cchunks;
  const int tx = bid % a.tiles_x; bid /= a.tiles_x;
  const int ty = bid % a.tiles_y;
  const int n = bid / a.tiles_y;
  pdl_launch_dependents();
  if (tid == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  __syncthreads();
  pdl_wait();                   // the input activation is the previous kernel's output
  if (tid == 0) {
    mbar_arrive_expect_tx(&bar, (uint32_t)(a.BH * a.BW * a.CT * 2));
    tma_load_4d(tile, &tm, &bar, cc * a.CT, tx * a.TX * STRIDE - 1, ty * a.TY * STRIDE - 1, n);
  }
  const int cgs = a.CT >> 3;
  const int cg = tid % cgs, pt = tid / cgs, pxt = (int)blockDim.x / cgs;
  const int c0 = cc * a.CT + cg * 8;
  // weights of this thread's 8 channels as packed fp16 pairs, bias as fp32 (overlaps the TMA flight)
  uint32_t wq[9][4];
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    const float4 w0 = __ldg(reinterpret_cast<const float4*>(a.w + (size_t)t * a.C + c0));
    const float4 w1 = __ldg(reinterpret_cast<const float4*>(a.w + (size_t)t * a.C + c0 + 4));
    wq[t][0] = pack_half2(w0.x, w0.y); wq[t][1] = pack_half2(w0.z, w0.w);
    wq[t][2] = pack_half2(w1.x, w1.y); wq[t][3] = pack_half2(w1.z, w1.w);
  }
  const float4 b0 = __ldg(reinterpret_cast<const float4*>(a.bias + c0));
  const float4 b1 = __ldg(reinterpret_cast<const float4*>(a.bias + c0 + 4));
  mbar_wait(&bar, 0);
  const uint32_t tile_s = smem_u32(tile) + (uint32_t)(cg * 16);
  // Work item = NPX horizontally adjacent outputs x 8 channels: they share (NPX-1)*STRIDE+3 input columns per kernel row
  // (NPX = 2: 12 / 15 shared-memory loads instead of 18; NPX = 4, stride 1: 18 instead of 36) and the per-item address
  // and loop instructions are paid once per NPX outputs - the kernel is issue-bound (ncu: 55 % issue utilisation, half
  // of it FHFMA), so instructions per output are what counts.  NPX = 4 needs TX % 4 == 0 (the host picks such a tile).
  constexpr int NC = (NPX - 1) * STRIDE + 3;
  const int ppr = NPX == 2 ? ((a.TX + 1) >> 1) : (a.TX / NPX);
  const int nitems = ppr * a.TY;
  // (row, item-in-row) advance incrementally: one runtime division per CTA instead of one per item
  const int step_y = pxt / ppr, step_x = pxt - step_y * ppr;
  int oy = pt / ppr, oxi = pt - oy * ppr;
  for (int p = pt; p < nitems; p += pxt, oxi += step_x, oy += step_y) {
    if (oxi >= ppr) { oxi -= ppr; ++oy; }
    const int ox = oxi * NPX;
    const int gx = tx * a.TX + ox, gy = ty * a.TY + oy;
    if (gx >= a.Wo || gy >= a.Ho) continue;
    float acc[NPX][8];
#pragma unroll
    for (int q = 0; q < NPX; ++q) {
      acc[q][0] = b0.x; acc[q][1] = b0.y; acc[q][2] = b0.z; acc[q][3] = b0.w;
      acc[q][4] = b1.x; acc[q][5] = b1.y; acc[q][6] = b1.z; acc[q][7] = b1.w;
    }
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      uint4 v[NC];
      const int rowbase = (oy * STRIDE + ky) * a.BW;
#pragma unroll
      for (int c = 0; c < NC; ++c) {
        const int col = NPX == 2 ? min(ox * STRIDE + c, a.BW - 1) : ox * STRIDE + c;
        v[c] = lds_v4(tile_s + (uint32_t)((rowbase + col) * a.CT * 2));
      }
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int t = ky * 3 + kx;
#pragma unroll
        for (int q = 0; q < NPX; ++q) {
          fhfma2(acc[q][0], acc[q][1], v[q * STRIDE + kx].x, wq[t][0]);
          fhfma2(acc[q][2], acc[q][3], v[q * STRIDE + kx].y, wq[t][1]);
          fhfma2(acc[q][4], acc[q][5], v[q * STRIDE + kx].z, wq[t][2]);
          fhfma2(acc[q][6], acc[q][7], v[q * STRIDE + kx].w, wq[t][3]);
        }
      }
    }
    __half* o = a.out + (((size_t)n * a.Ho + gy) * a.Wo + gx) * a.C + c0;
#pragma unroll
    for (int q = 0; q < NPX; ++q) {
      if (q == 0 || ((ox + q < a.TX) && (gx + q < a.Wo))) {
        if (a.relu) {
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[q][j] = fmaxf(acc[q][j], 0.f);
        }
        *reinterpret_cast<uint4*>(o + (size_t)q * a.C) = make_uint4(pack_half2(acc[q][0], acc[q][1]), pack_half2(acc[q][2], acc[q][3]),
                                                                   pack_half2(acc[q][4], acc[q][5]), pack_half2(acc[q][6], acc[q][7]));
      }
    }
  }
}

struct DwLaunch {
  CUtensorMap tm;
  DwArgs args;
  int grid, smem, stride, threads, npx;
  bool tma;
  const void* in;
};

constexpr int DW_SMEM_LIMIT = 56 * 1024;

// Tuning overrides for sweeps (scripts/sweep_dw.py): DVC_DW_SMEM_KB caps the input tile, DVC_DW_THREADS sets the CTA size.
static int dw_env_int(const char* name, int dflt, int lo, int hi) {
  const char* e = getenv(name);
  if (!e || !*e) return dflt;
  const int v = atoi(e);
  return v < lo ? lo : (v > hi ? hi : v);
}

int dw_prepare(DwLaunch* L, const void* in, int B, int H, int W, int C, int stride, const float* w, const float* b, int relu,
               void* out) {
  DVC_REQUIRE(C % 8 == 0 && (stride == 1 || stride == 2), "dwconv: C=%d must be a multiple of 8, stride=%d in {1,2}", C, stride);
  DVC_REQUIRE(B > 0 && H > 0 && W > 0, "dwconv: bad shape");
  DVC_REQUIRE((reinterpret_cast<uintptr_t>(in) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0, "dwconv: unaligned tensors");
  DwArgs& a = L->args;
  a.B = B; a.H = H; a.W = W; a.C = C; a.w = w; a.bias = b; a.relu = relu; a.out = reinterpret_cast<__half*>(out);
  a.Ho = (H + 2 - 3) / stride + 1; a.Wo = (W + 2 - 3) / stride + 1;
  L->stride = stride; L->in = in;
  a.CT = C % 64 == 0 ? 64 : (C % 32 == 0 ? 32 : (C % 16 == 0 ? 16 : 8));
  const int ct_env = dw_env_int("DVC_DW_CT", 0, 0, 128);
  if (ct_env >= 8 && (ct_env & (ct_env - 1)) == 0 && C % ct_env == 0) a.CT = ct_env;
  a.cchunks = C / a.CT;
  // tile search: minimise (input pixels fetched incl. halo + output pixels covered incl. edge waste)
  double best = 1e30;
  int bx = 0, by = 0;
  const int smem_limit = dw_env_int("DVC_DW_SMEM_KB", DW_SMEM_LIMIT / 1024, 4, DW_SMEM_LIMIT / 1024) * 1024;
  L->threads = dw_env_int("DVC_DW_THREADS", 128, 64, 256) & ~31;   // 128: 4 CTAs per SM (register-bound), ~200 KB of TMA in flight per SM
  const int min_outs = L->threads / 4;
  // stride 1: four outputs per work item when a tile whose width is a multiple of 4 costs at most 8 % more than the
  // unconstrained best (DVC_DW_NPX=2 forces the pair kernel)
  double best4 = 1e30;
  int bx4 = 0, by4 = 0;
  for (int TY = 1; TY <= (a.Ho < 64 ? a.Ho : 64); ++TY) {
    for (int TX = 1; TX <= 64; ++TX) {
      const int outs = TX * TY;
      if (TX % 4 == 0 && stride == 1 && TX <= ((a.Wo + 3) & ~3) && outs <= 1024 && !(outs < min_outs && (TX < a.Wo || TY < a.Ho))) {
        const int BW4 = TX + 2, BH4 = TY + 2;
        if (BW4 <= 256 && BH4 <= 256 && BW4 * BH4 * a.CT * 2 <= smem_limit) {
          const int ntx = ceil_div(a.Wo, TX), nty = ceil_div(a.Ho, TY);
          const double cost = (double)ntx * nty * BW4 * BH4 / ((double)H * W) + 1.5 * (double)ntx * TX * nty * TY / ((double)a.Wo * a.Ho)
                              + 0.02 * (256.0 / outs);
          if (cost < best4) { best4 = cost; bx4 = TX; by4 = TY; }
        }
      }
      if (TX > (a.Wo < 64 ? a.Wo : 64)) continue;
      if (outs > 1024) continue;
      if (outs < min_outs && (TX < a.Wo || TY < a.Ho)) continue;
      const int BW = (TX - 1) * stride + 3, BH = (TY - 1) * stride + 3;
      if (BW > 256 || BH > 256 || BW * BH * a.CT * 2 > smem_limit) continue;
      const int ntx = ceil_div(a.Wo, TX), nty = ceil_div(a.Ho, TY);
      const double cost = (double)ntx * nty * BW * BH / ((double)H * W) + 1.5 * (double)ntx * TX * nty * TY / ((double)a.Wo * a.Ho)
                          + 0.02 * (256.0 / outs);
      if (cost < best) { best = cost; bx = TX; by = TY; }
    }
  }
  DVC_REQUIRE(bx > 0, "dwconv: no tile fits shared memory (W=%d H=%d C=%d)", W, H, C);
  L->npx = 2;
  const double slack = 1.0 + dw_env_int("DVC_DW_NPX4_SLACK", 20, 0, 100) / 100.0;
  if (stride == 1 && bx4 > 0 && best4 <= slack * best && dw_env_int("DVC_DW_NPX", 4, 2, 4) == 4) {
    L->npx = 4; bx = bx4; by = by4;
  }
  a.TX = bx; a.TY = by;
  a.BW = (bx - 1) * stride + 3; a.BH = (by - 1) * stride + 3;
  a.tiles_x = ceil_div(a.Wo, bx); a.tiles_y = ceil_div(a.Ho, by);
  const long long grid = (long long)B * a.tiles_y * a.tiles_x * a.cchunks;
  DVC_REQUIRE(grid < (1LL << 31), "dwconv: grid too large");
  L->grid = (int)grid;
  L->smem = a.BW * a.BH * a.CT * 2 + 128;
  L->tma = true;
  return make_tmap_nhwc_f16(&L->tm, in, (uint64_t)B, (uint64_t)H, (uint64_t)W, (uint64_t)C, (uint32_t)a.CT, (uint32_t)a.BW, (uint32_t)a.BH);
}

int dw_launch(const DwLaunch* L, cudaStream_t st) {
  static std::atomic<bool> attr_set[64];   // zero-initialised; engine threads of different GPUs may race here
  int dev = 0;
  DVC_CHECK_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63].load(std::memory_order_acquire)) {
    DVC_CHECK_CUDA(cudaFuncSetAttribute(dw3x3_tma_kernel<1, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, DW_SMEM_LIMIT + 1024));
    DVC_CHECK_CUDA(cudaFuncSetAttribute(dw3x3_tma_kernel<1, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, DW_SMEM_LIMIT + 1024));
    DVC_CHECK_CUDA(cudaFuncSetAttribute(dw3x3_tma_kernel<2, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, DW_SMEM_LIMIT + 1024));
    attr_set[dev & 63].store(true, std::memory_order_release);
  }
  if (L->stride == 1 && L->npx == 4) DVC_CHECK_CUDA(launch_pdl(dw3x3_tma_kernel<1, 4>, dim3(L->grid), dim3(L->threads), (size_t)L->smem, st, 1, L->tm, L->args));
  else if (L->stride == 1) DVC_CHECK_CUDA(launch_pdl(dw3x3_tma_kernel<1, 2>, dim3(L->grid), dim3(L->threads), (size_t)L->smem, st, 1, L->tm, L->args));
  else DVC_CHECK_CUDA(launch_pdl(dw3x3_tma_kernel<2, 2>, dim3(L->grid), dim3(L->threads), (size_t)L->smem, st, 1, L->tm, L->args));
  count_launch();
  DVC_CHECK_CUDA(cudaGetLastError());
  return DVC_OK;
}

DwLaunch* dw_alloc() { return new DwLaunch(); }
void dw_free(DwLaunch* L) { delete L; }

int dwconv_launch(const void* in, int B, int H, int W, int C, int stride, const float* w, const float* b, int relu, void* out,
                  cudaStream_t st) {
  DwLaunch L;
  int rc = dw_prepare(&L, in, B, H, W, C, stride, w, b, relu, out);
  if (rc) return rc;
  return dw_launch(&L, st);
}

// ---- im2col for the dense 3x3 stride-2 pad-1 extras (conv14_2 .. conv17_2, prototxt:922-1146)
__global__ void __launch_bounds__(256)
im2col3x3s2_kernel(const __half* in, int B, int H, int W, int C, int Ho, int Wo, __half* __restrict__ out,
                   long long total) {
  pdl_launch_dependents();
  pdl_wait();
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
    v = ldg_v4(in + (((size_t)b * H + iy) * W + ix) * C + cg * 8);
  const size_t m = ((size_t)b * Ho + oy) * Wo + ox;
  *reinterpret_cast<uint4*>(out + m * (size_t)(9 * C) + (size_t)tap * C + cg * 8) = v;
}

int im2col_launch(const void* in, int B, int H, int W, int C, void* out, cudaStream_t st) {
  DVC_REQUIRE(C % 8 == 0 && B > 0 && H > 0 && W > 0, "im2col: bad shape C=%d", C);
  const int Ho = (H + 2 - 3) / 2 + 1, Wo = (W + 2 - 3) / 2 + 1;
  const long long total = (long long)B * Ho * Wo * 9 * (C / 8);
  DVC_CHECK_CUDA(launch_pdl(im2col3x3s2_kernel, dim3((unsigned)ceil_div64(total, 256)), dim3(256), 0, st, 1, (const __half*)in, B, H, W, C,
                            Ho, Wo, (__half*)out, total));
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
