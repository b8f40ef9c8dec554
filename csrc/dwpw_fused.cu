// EXPERIMENTAL - written after this round's GPU budget was all but spent.  Its one hardware run (the last
// seconds of the budget, profiles/dwpw_fused_first_run.log) was bit-identical to the dwconv3x3 -> gemm_f16 path on five
// small shapes (overhanging tiles, C_in = 32, C_out = 192, several K chunks); it has NOT been timed at the flagship shapes
// and nothing in the plan executor calls it yet.  scripts/check_dwpw_fused.py is the numerics / timing harness.
//
// K3 as the producer stage of K2: depthwise 3x3 (stride 1, pad 1) + bias + ReLU feeding the
// pointwise 1x1 + bias + ReLU of the same MobileNet block without the intermediate ever leaving the SM.
//
//   mid[p, k] = relu( sum_{3x3} in[p + tap, k] * wdw[tap, k] + bdw[k] )          CUDA cores (FHFMA), fp16 result
//   out[p, n] = relu( sum_k mid[p, k] * wpw[n, k] + bpw[n] )                      tcgen05.mma, fp32 accumulate
//
// For conv1 / conv3 / conv5 (prototxt:42-106, 172-236, 302-366) the separate kernels move in + 2*mid + out bytes
// through HBM (460 MB for conv1 at batch 64, both kernels at the copy-rate roofline); fused it is in + out (276 MB).
//
// One persistent CTA per SM, 448 threads, tile = 8 x 16 output pixels (M = 128) x all C_out <= 256 columns:
//   warp 0        TMA producer: per K chunk of CT = min(64, C_in) channels one 4-D box [10 x 18 x CT] of the input
//                 (halo included, out-of-range = zero padding) and the matching [C_out x 64] slice of the 1x1 weights
//   warps 6..13   depthwise producers: 8 channels x 2 adjacent pixels per work item from the staged box, ReLU, fp16,
//                 written as the 128B-swizzled K-major A operand of the chunk (the layout TMA would have produced)
//   warp 1        TMEM allocator + MMA issuer: tcgen05.mma M=128 N=C_out over the chunks, two accumulators in TMEM
//   warps 2..5    epilogue: tcgen05.ld, +bias, ReLU, fp16, swizzled staging, 4-D TMA store of [2 rows x 16 px x 64 ch]
//                 boxes (the hardware clips tile pixels that fall outside the image)
#include <atomic>

#include "common.cuh"

namespace dvc {

namespace fused {

constexpr int TY = 8, TX = 16;                     // output tile (pixels); TX * TY = 128 = UMMA M
constexpr int BH = TY + 2, BW = TX + 2;            // input box incl. halo (stride 1)
constexpr int BK = 64;                             // K block of the A / B operands (128 bytes of fp16)
constexpr int THREADS = 448;                       // 14 warps
constexpr int DW_WARP0 = 6, DW_WARPS = 8;
constexpr int A_STAGE_BYTES = 128 * BK * 2;        // 16 KiB
constexpr int S_IN = 3;                            // input-box stages
constexpr int S_AB = 2;                            // A / B operand stages
constexpr int EPI_CHUNK = 64;
constexpr int EPI_BUF_BYTES = 32 * EPI_CHUNK * 2;  // one warp: 32 pixels x 128 B
constexpr int MAX_SMEM = 227 * 1024;

struct Args {
  int B, H, W, C_in, C_out;
  int CT, num_kb;                  // channels per chunk, chunks per tile
  int tiles_x, tiles_y, num_tiles;
  int tmem_cols;
  const float* wdw;                // [9][C_in]
  const float* bdw;                // [C_in]
  const float* bpw;                // [C_out]
  // shared-memory carve-up (bytes from the 1024-aligned base)
  int off_in, in_stage_bytes, off_a, off_b, b_stage_bytes, off_epi, off_wdw, off_bias, off_bar;
};

__device__ __forceinline__ void tma_store_4d(const CUtensorMap* tmap, const void* smem_src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
               ::"l"(tmap), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}

}  // namespace fused

__global__ void __launch_bounds__(fused::THREADS, 1)
dwpw_fused_s1_kernel(const __grid_constant__ CUtensorMap tmIn, const __grid_constant__ CUtensorMap tmB,
                     const __grid_constant__ CUtensorMap tmOut, const fused::Args g) {
  using namespace fused;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = __shfl_sync(0xffffffff, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;

  uint8_t* sEpi = smem + g.off_epi;                        // [4 warps][2][32 px x 128 B]
  float* sBias = reinterpret_cast<float*>(smem + g.off_bias);
  float* sWdw = reinterpret_cast<float*>(smem + g.off_wdw);   // [9][C_in] then [C_in] bias
  float* sBdw = sWdw + 9 * g.C_in;
  uint8_t* sA = smem + g.off_a;
  uint8_t* sB = smem + g.off_b;
  uint8_t* sIn = smem + g.off_in;
  uint64_t* in_full = reinterpret_cast<uint64_t*>(smem + g.off_bar);
  uint64_t* in_empty = in_full + S_IN;
  uint64_t* a_full = in_empty + S_IN;
  uint64_t* b_full = a_full + S_AB;
  uint64_t* ab_empty = b_full + S_AB;
  uint64_t* tmem_full = ab_empty + S_AB;                   // [2]
  uint64_t* tmem_empty = tmem_full + 2;                    // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  pdl_launch_dependents();
  // constants first (they may overlap the previous kernel's tail under programmatic dependent launch)
  for (int i = threadIdx.x; i < g.C_out; i += THREADS) sBias[i] = __ldg(g.bpw + i);
  for (int i = threadIdx.x; i < 9 * g.C_in; i += THREADS) sWdw[i] = __ldg(g.wdw + i);
  for (int i = threadIdx.x; i < g.C_in; i += THREADS) sBdw[i] = __ldg(g.bdw + i);
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmIn);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmOut);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < S_IN; ++s) {
        mbar_init(&in_full[s], 1);
        mbar_init(&in_empty[s], DW_WARPS);
      }
      for (int s = 0; s < S_AB; ++s) {
        mbar_init(&a_full[s], DW_WARPS);
        mbar_init(&b_full[s], 1);
        mbar_init(&ab_empty[s], 1);
      }
      for (int a = 0; a < 2; ++a) {
        mbar_init(&tmem_full[a], 1);
        mbar_init(&tmem_empty[a], 4);
      }
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, (uint32_t)g.tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int tiles_per_img = g.tiles_x * g.tiles_y;
  const int ksteps = g.CT / 16;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer: input boxes + weight slices
    if (elect_one()) {
      pdl_wait();
      int it = 0;
      for (int t = blockIdx.x; t < g.num_tiles; t += gridDim.x) {
        const int n = t / tiles_per_img, r = t - n * tiles_per_img;
        const int ty0 = (r / g.tiles_x) * TY, tx0 = (r % g.tiles_x) * TX;
        for (int kb = 0; kb < g.num_kb; ++kb, ++it) {
          const int si = it % S_IN, sb = it % S_AB;
          mbar_wait(&in_empty[si], ((uint32_t)(it / S_IN) & 1u) ^ 1u);
          mbar_arrive_expect_tx(&in_full[si], (uint32_t)g.in_stage_bytes);
          tma_load_4d(sIn + si * g.in_stage_bytes, &tmIn, &in_full[si], kb * g.CT, tx0 - 1, ty0 - 1, n);
          mbar_wait(&ab_empty[sb], ((uint32_t)(it / S_AB) & 1u) ^ 1u);
          mbar_arrive_expect_tx(&b_full[sb], (uint32_t)g.b_stage_bytes);
          tma_load_2d(sB + sb * g.b_stage_bytes, &tmB, &b_full[sb], kb * g.CT, 0);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    pdl_wait();
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_f16(128, (uint32_t)g.C_out);
      int it = 0, tl = 0;
      for (int t = blockIdx.x; t < g.num_tiles; t += gridDim.x, ++tl) {
        const int acc = tl & 1;
        mbar_wait(&tmem_empty[acc], (((uint32_t)tl >> 1) & 1u) ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * g.C_out);
        for (int kb = 0; kb < g.num_kb; ++kb, ++it) {
          const int sb = it % S_AB;
          const uint32_t ph = (uint32_t)(it / S_AB) & 1u;
          mbar_wait(&a_full[sb], ph);          // depthwise producers have written (and proxy-fenced) the A block
          mbar_wait(&b_full[sb], ph);          // weight slice has landed
          tc_fence_after();
          const uint64_t a_desc = umma_desc_sw128(smem_u32(sA + sb * A_STAGE_BYTES));
          const uint64_t b_desc = umma_desc_sw128(smem_u32(sB + sb * g.b_stage_bytes));
          for (int k = 0; k < ksteps; ++k)
            umma_f16(d_tmem, a_desc + (uint64_t)(k * 2), b_desc + (uint64_t)(k * 2), idesc, (uint32_t)((kb | k) != 0));
          umma_commit(&ab_empty[sb]);          // A / B stage reusable once these MMAs retire
        }
        umma_commit(&tmem_full[acc]);
      }
    }
    __syncwarp();
  } else if (warp < DW_WARP0) {
    // ------------------------------------------------------------------ epilogue (4 warps)
    pdl_wait();
    const int q = warp & 3;                    // TMEM lane quadrant = tile rows 2q, 2q + 1
    const uint32_t bias_s = smem_u32(sBias);
    uint8_t* my_epi = sEpi + q * 2 * EPI_BUF_BYTES;
    int chunk_it = 0, tl = 0;
    for (int t = blockIdx.x; t < g.num_tiles; t += gridDim.x, ++tl) {
      const int n = t / tiles_per_img, r = t - n * tiles_per_img;
      const int ty0 = (r / g.tiles_x) * TY, tx0 = (r % g.tiles_x) * TX;
      const int acc = tl & 1;
      mbar_wait(&tmem_full[acc], ((uint32_t)tl >> 1) & 1u);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * g.C_out);
      for (int c0 = 0; c0 < g.C_out; c0 += EPI_CHUNK, ++chunk_it) {
        uint8_t* buf = my_epi + (chunk_it & 1) * EPI_BUF_BYTES;
        if (lane == 0) tma_store_wait_read<1>();     // the store issued two chunks ago has finished reading this buffer
        __syncwarp();
        float v[EPI_CHUNK];
#pragma unroll
        for (int j = 0; j < EPI_CHUNK / 16; ++j) tmem_ld16(taddr + (uint32_t)(c0 + 16 * j), v + 16 * j);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < EPI_CHUNK / 8; ++j) {
          uint32_t pk[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 bb = lds_f2(bias_s + (uint32_t)((c0 + 8 * j + 2 * e) * 4));
            const float x0 = fmaxf(v[8 * j + 2 * e] + bb.x, 0.f);
            const float x1 = fmaxf(v[8 * j + 2 * e + 1] + bb.y, 0.f);
            pk[e] = pack_half2(x0, x1);
          }
          sts_v4(smem_u32(buf) + (uint32_t)(lane * 128 + ((j ^ (lane & 7)) << 4)), make_uint4(pk[0], pk[1], pk[2], pk[3]));
        }
        fence_proxy_async_smem();
        __syncwarp();
        // box = [2 tile rows][16 px][64 ch]; lane l holds pixel (row 2q + l / 16, column l % 16); pixels outside the image are clipped
        if (lane == 0 && ty0 + 2 * q < g.H) {
          tma_store_4d(&tmOut, buf, c0, tx0, ty0 + 2 * q, n);
          tma_store_commit();
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
    }
    if (lane == 0) tma_store_wait_all<0>();
  } else {
    // ------------------------------------------------------------------ depthwise producers (8 warps, 256 threads)
    pdl_wait();
    const int dtid = threadIdx.x - DW_WARP0 * 32;
    const int cgs = g.CT >> 3;                 // 8-channel groups per chunk (8 or 4)
    const int cg = dtid % cgs, slot = dtid / cgs, nslots = (DW_WARPS * 32) / cgs;
    constexpr int NPAIRS = TY * (TX / 2);      // two horizontally adjacent outputs per work item
    int it = 0;
    for (int t = blockIdx.x; t < g.num_tiles; t += gridDim.x) {
      for (int kb = 0; kb < g.num_kb; ++kb, ++it) {
        const int si = it % S_IN, sb = it % S_AB;
        const int c0 = kb * g.CT + cg * 8;
        // this thread's 8 channels: 3x3 weights as packed fp16 pairs, bias as fp32 (from the smem copy)
        uint32_t wq[9][4];
#pragma unroll
        for (int tp = 0; tp < 9; ++tp) {
          const float4 w0 = *reinterpret_cast<const float4*>(sWdw + tp * g.C_in + c0);
          const float4 w1 = *reinterpret_cast<const float4*>(sWdw + tp * g.C_in + c0 + 4);
          wq[tp][0] = pack_half2(w0.x, w0.y); wq[tp][1] = pack_half2(w0.z, w0.w);
          wq[tp][2] = pack_half2(w1.x, w1.y); wq[tp][3] = pack_half2(w1.z, w1.w);
        }
        const float4 b0 = *reinterpret_cast<const float4*>(sBdw + c0);
        const float4 b1 = *reinterpret_cast<const float4*>(sBdw + c0 + 4);
        mbar_wait(&in_full[si], (uint32_t)(it / S_IN) & 1u);                   // input box has landed
        mbar_wait(&ab_empty[sb], ((uint32_t)(it / S_AB) & 1u) ^ 1u);           // the MMAs that read this A stage are done
        const uint32_t in_s = smem_u32(sIn + si * g.in_stage_bytes) + (uint32_t)(cg * 16);
        const uint32_t a_s = smem_u32(sA + sb * A_STAGE_BYTES);
        for (int p = slot; p < NPAIRS; p += nslots) {
          const int oy = p / (TX / 2), ox = (p % (TX / 2)) * 2;
          float acc0[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
          float acc1[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
          for (int ky = 0; ky < 3; ++ky) {
            uint4 v[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) v[c] = lds_v4(in_s + (uint32_t)((((oy + ky) * BW) + ox + c) * g.CT * 2));
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
              const int tp = ky * 3 + kx;
              fhfma2(acc0[0], acc0[1], v[kx].x, wq[tp][0]);
              fhfma2(acc0[2], acc0[3], v[kx].y, wq[tp][1]);
              fhfma2(acc0[4], acc0[5], v[kx].z, wq[tp][2]);
              fhfma2(acc0[6], acc0[7], v[kx].w, wq[tp][3]);
              fhfma2(acc1[0], acc1[1], v[1 + kx].x, wq[tp][0]);
              fhfma2(acc1[2], acc1[3], v[1 + kx].y, wq[tp][1]);
              fhfma2(acc1[4], acc1[5], v[1 + kx].z, wq[tp][2]);
              fhfma2(acc1[6], acc1[7], v[1 + kx].w, wq[tp][3]);
            }
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) { acc0[j] = fmaxf(acc0[j], 0.f); acc1[j] = fmaxf(acc1[j], 0.f); }
          // A operand, K-major with 128-byte swizzle: row = tile pixel, 16-byte chunk cg lives at position cg ^ (row & 7)
          const int row0 = oy * TX + ox, row1 = row0 + 1;
          sts_v4(a_s + (uint32_t)(row0 * 128 + ((cg ^ (row0 & 7)) << 4)),
                 make_uint4(pack_half2(acc0[0], acc0[1]), pack_half2(acc0[2], acc0[3]), pack_half2(acc0[4], acc0[5]), pack_half2(acc0[6], acc0[7])));
          sts_v4(a_s + (uint32_t)(row1 * 128 + ((cg ^ (row1 & 7)) << 4)),
                 make_uint4(pack_half2(acc1[0], acc1[1]), pack_half2(acc1[2], acc1[3]), pack_half2(acc1[4], acc1[5]), pack_half2(acc1[6], acc1[7])));
        }
        fence_proxy_async_smem();              // generic-proxy writes of A -> visible to the tensor core's async proxy
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&a_full[sb]);
          mbar_arrive(&in_empty[si]);          // this warp no longer reads the input box
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, (uint32_t)g.tmem_cols);
}

// ------------------------------------------------------------------------------------ host

static int fused_num_sms() {
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return 148;
  return n;
}

int dwpw_fused_s1_launch(const void* in, int B, int H, int W, int C_in, const float* wdw, const float* bdw,
                         const void* wpw, const float* bpw, int C_out, void* out, cudaStream_t st) {
  using namespace fused;
  DVC_REQUIRE(B > 0 && H > 0 && W > 0, "dwpw_fused: bad shape");
  DVC_REQUIRE(C_in == 32 || (C_in % 64 == 0 && C_in <= 1024), "dwpw_fused: C_in=%d must be 32 or a multiple of 64", C_in);
  DVC_REQUIRE(C_out % 64 == 0 && C_out >= 64 && C_out <= 256, "dwpw_fused: C_out=%d must be 64, 128, 192 or 256", C_out);
  DVC_REQUIRE(((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(wpw)) & 15) == 0,
              "dwpw_fused: tensors must be 16-byte aligned");
  Args g{};
  g.B = B; g.H = H; g.W = W; g.C_in = C_in; g.C_out = C_out;
  g.CT = C_in < 64 ? C_in : 64;
  g.num_kb = C_in / g.CT;
  g.tiles_x = ceil_div(W, TX); g.tiles_y = ceil_div(H, TY);
  const long long tiles = (long long)B * g.tiles_x * g.tiles_y;
  DVC_REQUIRE(tiles < (1LL << 31), "dwpw_fused: too many tiles");
  g.num_tiles = (int)tiles;
  int cols = 32;
  while (cols < 2 * C_out) cols <<= 1;
  g.tmem_cols = cols;
  g.wdw = wdw; g.bdw = bdw; g.bpw = bpw;
  auto up = [](int v, int a) { return (v + a - 1) / a * a; };
  int off = 0;
  g.off_epi = off; off += 4 * 2 * EPI_BUF_BYTES;
  g.off_bias = off; off += up(C_out * 4, 1024);
  g.off_wdw = off; off += up(10 * C_in * 4, 1024);
  g.off_a = off; off += S_AB * A_STAGE_BYTES;
  g.b_stage_bytes = C_out * BK * 2;
  g.off_b = off; off += S_AB * g.b_stage_bytes;
  g.in_stage_bytes = BH * BW * g.CT * 2;
  g.off_in = off; off += up(S_IN * g.in_stage_bytes, 1024);
  g.off_bar = off; off += 256;
  const int smem = off + 1024;       // slack for the 1024-byte alignment of the dynamic base
  DVC_REQUIRE(smem <= MAX_SMEM, "dwpw_fused: needs %d bytes of shared memory (C_in=%d C_out=%d)", smem, C_in, C_out);
  CUtensorMap tmIn, tmB, tmOut;
  int rc = make_tmap_nhwc_f16(&tmIn, in, (uint64_t)B, (uint64_t)H, (uint64_t)W, (uint64_t)C_in, (uint32_t)g.CT, BW, BH);
  if (rc) return rc;
  rc = make_tmap_2d_f16(&tmB, wpw, (uint64_t)C_out, (uint64_t)C_in, (uint64_t)C_in, (uint32_t)C_out, BK);
  if (rc) return rc;
  rc = make_tmap_nhwc_f16_sw128(&tmOut, out, (uint64_t)B, (uint64_t)H, (uint64_t)W, (uint64_t)C_out, EPI_CHUNK, TX, 2);
  if (rc) return rc;
  static std::atomic<bool> attr_set[64];
  int dev = 0;
  DVC_CHECK_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63].load(std::memory_order_acquire)) {
    DVC_CHECK_CUDA(cudaFuncSetAttribute(dwpw_fused_s1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_SMEM));
    attr_set[dev & 63].store(true, std::memory_order_release);
  }
  const int sms = fused_num_sms();
  const int grid = g.num_tiles < sms ? g.num_tiles : sms;
  DVC_CHECK_CUDA(launch_pdl(dwpw_fused_s1_kernel, dim3(grid), dim3(THREADS), (size_t)smem, st, 1, tmIn, tmB, tmOut, g));
  count_launch();
  DVC_CHECK_CUDA(cudaGetLastError());
  return DVC_OK;
}

}  // namespace dvc

/* EXPERIMENTAL (see the file header): fused depthwise 3x3 (stride 1) + pointwise 1x1, both with bias + ReLU.
 * in [B,H,W,C_in] fp16 NHWC, wdw [9][C_in] fp32, bdw [C_in], wpw [C_out][C_in] fp16, bpw [C_out], out [B,H,W,C_out] fp16. */
extern "C" int dvc_dwpw_fused_s1(const void* in_f16, int B, int H, int W, int C_in, const float* wdw, const float* bdw,
                                 const void* wpw_f16, const float* bpw, int C_out, void* out_f16, void* stream) {
  return dvc::dwpw_fused_s1_launch(in_f16, B, H, W, C_in, wdw, bdw, wpw_f16, bpw, C_out, out_f16, (cudaStream_t)stream);
}
