// K3 as the producer stage of K2: depthwise 3x3 (pad 1, stride 1 or 2) + bias + ReLU feeding the
// pointwise 1x1 + bias + ReLU of the same MobileNet block without the intermediate ever leaving the SM.
//
//   mid[p, k] = relu( sum_{3x3} in[p*s + tap, k] * wdw[tap, k] + bdw[k] )        CUDA cores (FHFMA), fp16 result
//   out[p, n] = relu( sum_k mid[p, k] * wpw[n, k] + bpw[n] )                      tcgen05.mma, fp32 accumulate
//
// (prototxt:42-106 conv1/dw+conv1 ... :302-366 conv5/dw+conv5.)  The two separate kernels move in + 2*mid + out bytes
// and - what matters more on this network - spend a whole GEMM launch whose tensor work could hide behind the
// depthwise arithmetic: measured, the standalone depthwise kernel is issue-bound on the CUDA cores (55% issue
// utilisation, half of it FHFMA) while the GEMM leaves them idle.  Fused, the MMA of chunk k runs while the
// depthwise warps compute chunk k+1.
//
// One persistent CTA per SM, 896 threads = 28 warps, tile = 8 x 16 output pixels (M = 128) x all C_out <= 256 columns:
//   warp 0        TMA producer: per K chunk of CT channels one 4-D box of the input (halo included, out-of-range =
//                 zero padding) and the matching [C_out x 64] slice of the 1x1 weights
//   warp 1        TMEM allocator + MMA issuer: tcgen05.mma M=128 N=C_out over the chunks, two accumulators in TMEM
//   warp 2        TMA producer of the 1x1 weight chunks (own ring; resident when the matrix is <= 64 KB)
//   warps 4..11   epilogue, two per TMEM lane quadrant: tcgen05.ld, +bias, ReLU, fp16, swizzled staging, 4-D TMA stores
//   warps 12..27  depthwise producers, 512 threads: ONE work item per thread and chunk (CT = 64: 8 channels x 2
//                 adjacent pixels; CT = 32: 8 channels x 1 pixel), weights as packed fp16 in shared memory (keeps the
//                 thread at 72 registers so that 28 warps fit), result written as the 128B-swizzled K-major A
//                 operand of the chunk (the layout TMA would have produced)
// v1 of this kernel had 8 depthwise warps and lost to the two-kernel path (profiles/r2/dwpw_fused_pairs.md).
#include <atomic>

#include "common.cuh"

namespace dvc {

namespace fused {

constexpr int TY = 8, TX = 16;                     // output tile (pixels); TX * TY = 128 = UMMA M
constexpr int BK = 64;                             // K block of the A / B operands (128 bytes of fp16)
constexpr int THREADS = 896;                       // 28 warps
constexpr int EPI_WARP0 = 4, EPI_WARPS = 8;        // two warps per TMEM lane quadrant, 32 of a chunk's 64 columns each
constexpr int DW_WARP0 = 12, DW_WARPS = 16;
constexpr int A_STAGE_BYTES = 128 * BK * 2;        // 16 KiB
constexpr int S_A_MAX = 4, S_B_MAX = 8;             // A ring (depthwise -> MMA) and B ring / resident weight chunks
constexpr int S_IN_MAX = 12;                       // input-box stages: as many as fit (the loads are latency-bound: ~80 KB per SM must be in flight)
constexpr int EPI_CHUNK = 64;
constexpr int EPI_BUF_BYTES = 32 * EPI_CHUNK * 2;  // one warp: 32 pixels x 128 B
constexpr int MAX_SMEM = 227 * 1024;

struct Args {
  int B, H, W, Ho, Wo, C_in, C_out;
  int CT, num_kb;                  // channels per chunk, chunks per tile
  int BH, BW, S_IN;                // input box incl. halo, number of input stages
  int S_A, S_B, b_resident;        // A stages; B stages (= num_kb when the whole 1x1 weight matrix stays in shared memory)
  int tiles_x, tiles_y, num_tiles;
  int tmem_cols;
  const float* wdw;                // [9][C_in]
  const float* bdw;                // [C_in]
  const float* bpw;                // [C_out]
  // shared-memory carve-up (bytes from the 1024-aligned base)
  int off_in, in_stage_bytes, in_box_bytes, off_a, off_b, b_stage_bytes, off_epi, off_wdw, off_bdw, off_bias, off_bar;
};

__device__ __forceinline__ void tma_store_4d(const CUtensorMap* tmap, const void* smem_src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
               ::"l"(tmap), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}

// One depthwise work item: NPX horizontally adjacent outputs x 8 channels.  `in_s` = shared address of input pixel
// (row 0, column 0 of the item's window) for this thread's channel group, `px_bytes` / `row_bytes` the box pitches,
// `w_s` the 9 packed-fp16 weight vectors, `b_s` the 8 fp32 biases.  Results (ReLU, fp16) land in out[NPX].
template <int STRIDE, int NPX>
__device__ __forceinline__ void dw_item(uint32_t in_s, uint32_t px_bytes, uint32_t row_bytes, uint32_t w_s, uint32_t b_s, uint4* out) {
  constexpr int NC = (NPX - 1) * STRIDE + 3;
  const uint4 bq0 = lds_v4(b_s), bq1 = lds_v4(b_s + 16);
  float acc[NPX][8];
#pragma unroll
  for (int p = 0; p < NPX; ++p) {
    acc[p][0] = __uint_as_float(bq0.x); acc[p][1] = __uint_as_float(bq0.y); acc[p][2] = __uint_as_float(bq0.z); acc[p][3] = __uint_as_float(bq0.w);
    acc[p][4] = __uint_as_float(bq1.x); acc[p][5] = __uint_as_float(bq1.y); acc[p][6] = __uint_as_float(bq1.z); acc[p][7] = __uint_as_float(bq1.w);
  }
#pragma unroll
  for (int ky = 0; ky < 3; ++ky) {
    uint4 v[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) v[c] = lds_v4(in_s + (uint32_t)ky * row_bytes + (uint32_t)c * px_bytes);
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      const uint4 w = lds_v4(w_s + (uint32_t)((ky * 3 + kx) * 16));
#pragma unroll
      for (int p = 0; p < NPX; ++p) {
        fhfma2(acc[p][0], acc[p][1], v[p * STRIDE + kx].x, w.x);
        fhfma2(acc[p][2], acc[p][3], v[p * STRIDE + kx].y, w.y);
        fhfma2(acc[p][4], acc[p][5], v[p * STRIDE + kx].z, w.z);
        fhfma2(acc[p][6], acc[p][7], v[p * STRIDE + kx].w, w.w);
      }
    }
  }
#pragma unroll
  for (int p = 0; p < NPX; ++p)
    out[p] = make_uint4(pack_half2(fmaxf(acc[p][0], 0.f), fmaxf(acc[p][1], 0.f)), pack_half2(fmaxf(acc[p][2], 0.f), fmaxf(acc[p][3], 0.f)),
                        pack_half2(fmaxf(acc[p][4], 0.f), fmaxf(acc[p][5], 0.f)), pack_half2(fmaxf(acc[p][6], 0.f), fmaxf(acc[p][7], 0.f)));
}

}  // namespace fused

template <int STRIDE>
__global__ void __launch_bounds__(fused::THREADS, 1)
dwpw_fused_kernel(const __grid_constant__ CUtensorMap tmIn, const __grid_constant__ CUtensorMap tmB,
                  const __grid_constant__ CUtensorMap tmOut, const fused::Args g) {
  using namespace fused;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = __shfl_sync(0xffffffff, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;

  uint8_t* sEpi = smem + g.off_epi;                        // [4 warps][2][32 px x 128 B]
  float* sBias = reinterpret_cast<float*>(smem + g.off_bias);
  uint4* sWpk = reinterpret_cast<uint4*>(smem + g.off_wdw);   // [C_in / 8][9] packed fp16 x 8
  float* sBdw = reinterpret_cast<float*>(smem + g.off_bdw);   // [C_in] fp32
  uint8_t* sA = smem + g.off_a;
  uint8_t* sB = smem + g.off_b;
  uint8_t* sIn = smem + g.off_in;
  uint64_t* in_full = reinterpret_cast<uint64_t*>(smem + g.off_bar);
  uint64_t* in_empty = in_full + S_IN_MAX;
  uint64_t* a_full = in_empty + S_IN_MAX;
  uint64_t* a_empty = a_full + S_A_MAX;
  uint64_t* b_full = a_empty + S_A_MAX;
  uint64_t* b_empty = b_full + S_B_MAX;
  uint64_t* tmem_full = b_empty + S_B_MAX;                 // [2]
  uint64_t* tmem_empty = tmem_full + 2;                    // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  pdl_launch_dependents();
  // constants first (they may overlap the previous kernel's tail under programmatic dependent launch)
  for (int i = threadIdx.x; i < g.C_out; i += THREADS) sBias[i] = __ldg(g.bpw + i);
  for (int i = threadIdx.x; i < g.C_in; i += THREADS) sBdw[i] = __ldg(g.bdw + i);
  for (int i = threadIdx.x; i < (g.C_in >> 3) * 9; i += THREADS) {
    const int cgi = i / 9, tp = i - cgi * 9;
    const float4 w0 = __ldg(reinterpret_cast<const float4*>(g.wdw + (size_t)tp * g.C_in + cgi * 8));
    const float4 w1 = __ldg(reinterpret_cast<const float4*>(g.wdw + (size_t)tp * g.C_in + cgi * 8 + 4));
    sWpk[i] = make_uint4(pack_half2(w0.x, w0.y), pack_half2(w0.z, w0.w), pack_half2(w1.x, w1.y), pack_half2(w1.z, w1.w));
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmIn);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmOut);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < g.S_IN; ++s) {
        mbar_init(&in_full[s], 1);
        mbar_init(&in_empty[s], DW_WARPS);
      }
      for (int s = 0; s < S_A_MAX; ++s) {
        mbar_init(&a_full[s], DW_WARPS);
        mbar_init(&a_empty[s], 1);
      }
      for (int s = 0; s < S_B_MAX; ++s) {
        mbar_init(&b_full[s], 1);
        mbar_init(&b_empty[s], 1);
      }
      for (int a = 0; a < 2; ++a) {
        mbar_init(&tmem_full[a], 1);
        mbar_init(&tmem_empty[a], EPI_WARPS);
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
  const int S_IN = g.S_IN, S_A = g.S_A, S_B = g.S_B;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer: input boxes + weight slices
    if (elect_one()) {
      pdl_wait();
      int it = 0;
      for (int t = blockIdx.x; t < g.num_tiles; t += gridDim.x) {
        const int n = t / tiles_per_img, r = t - n * tiles_per_img;
        const int ty0 = (r / g.tiles_x) * TY, tx0 = (r % g.tiles_x) * TX;
        for (int kb = 0; kb < g.num_kb; ++kb, ++it) {
          const int si = it % S_IN;
          mbar_wait(&in_empty[si], ((uint32_t)(it / S_IN) & 1u) ^ 1u);
          mbar_arrive_expect_tx(&in_full[si], (uint32_t)g.in_box_bytes);
          tma_load_4d(sIn + si * g.in_stage_bytes, &tmIn, &in_full[si], kb * g.CT, tx0 * STRIDE - 1, ty0 * STRIDE - 1, n);
        }
      }
    }
    __syncwarp();
  } else if (warp == 2) {
    // ------------------------------------------------------------------ TMA producer: 1x1 weight chunks
    // their own ring and their own thread: a weight load (L2 latency) must never sit between the depthwise warps and
    // their next A stage.  Small weight matrices are loaded once and stay resident.
    if (elect_one()) {
      if (g.b_resident) {
        for (int kb = 0; kb < g.num_kb; ++kb) {
          mbar_arrive_expect_tx(&b_full[kb], (uint32_t)g.b_stage_bytes);
          tma_load_2d(sB + kb * g.b_stage_bytes, &tmB, &b_full[kb], kb * g.CT, 0);
        }
      } else {
        int it = 0;
        for (int t = blockIdx.x; t < g.num_tiles; t += gridDim.x) {
          for (int kb = 0; kb < g.num_kb; ++kb, ++it) {
            const int sb = it % S_B;
            mbar_wait(&b_empty[sb], ((uint32_t)(it / S_B) & 1u) ^ 1u);
            mbar_arrive_expect_tx(&b_full[sb], (uint32_t)g.b_stage_bytes);
            tma_load_2d(sB + sb * g.b_stage_bytes, &tmB, &b_full[sb], kb * g.CT, 0);
          }
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
          const int sa = it % S_A, sb = g.b_resident ? kb : it % S_B;
          mbar_wait(&a_full[sa], (uint32_t)(it / S_A) & 1u);      // depthwise producers have written (and proxy-fenced) the A block
          mbar_wait(&b_full[sb], g.b_resident ? 0u : ((uint32_t)(it / S_B) & 1u));    // weight chunk has landed
          tc_fence_after();
          const uint64_t a_desc = umma_desc_sw128(smem_u32(sA + sa * A_STAGE_BYTES));
          const uint64_t b_desc = umma_desc_sw128(smem_u32(sB + sb * g.b_stage_bytes));
          for (int k = 0; k < ksteps; ++k)
            umma_f16(d_tmem, a_desc + (uint64_t)(k * 2), b_desc + (uint64_t)(k * 2), idesc, (uint32_t)((kb | k) != 0));
          umma_commit(&a_empty[sa]);           // A stage reusable once these MMAs retire
          if (!g.b_resident) umma_commit(&b_empty[sb]);
        }
        umma_commit(&tmem_full[acc]);
      }
    }
    __syncwarp();
  } else if (warp >= EPI_WARP0 && warp < DW_WARP0) {
    // ------------------------------------------------------------------ epilogue (8 warps)
    // A single warp per quadrant was the bottleneck of the whole kernel (its ~330 dependent instructions per tile took
    // longer than the depthwise stage): two warps share a quadrant, each drains 32 of the chunk's 64 columns.
    pdl_wait();
    const int q = warp & 3;                    // TMEM lane quadrant = tile rows 2q, 2q + 1
    const int hf = (warp - EPI_WARP0) >> 2;    // which 32 columns of each 64-column chunk
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
        float v[32];
        tmem_ld16(taddr + (uint32_t)(c0 + 32 * hf), v);
        tmem_ld16(taddr + (uint32_t)(c0 + 32 * hf + 16), v + 16);
        if (hf == 0 && lane == 0) tma_store_wait_read<1>();     // the store issued two chunks ago has finished reading this buffer
        asm volatile("bar.sync %0, 64;" ::"r"(1 + q) : "memory");
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint32_t pk[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 bb = lds_f2(bias_s + (uint32_t)((c0 + 32 * hf + 8 * j + 2 * e) * 4));
            pk[e] = pack_half2(fmaxf(v[8 * j + 2 * e] + bb.x, 0.f), fmaxf(v[8 * j + 2 * e + 1] + bb.y, 0.f));
          }
          sts_v4(smem_u32(buf) + (uint32_t)(lane * 128 + (((4 * hf + j) ^ (lane & 7)) << 4)), make_uint4(pk[0], pk[1], pk[2], pk[3]));
        }
        fence_proxy_async_smem();
        asm volatile("bar.sync %0, 64;" ::"r"(1 + q) : "memory");
        // box = [2 tile rows][16 px][64 ch]; lane l holds pixel (row 2q + l / 16, column l % 16); pixels outside the image are clipped
        if (hf == 0 && lane == 0 && ty0 + 2 * q < g.Ho) {
          tma_store_4d(&tmOut, buf, c0, tx0, ty0 + 2 * q, n);
          tma_store_commit();
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
    }
    if (hf == 0 && lane == 0) tma_store_wait_all<0>();
  } else if (warp >= DW_WARP0) {
    // ------------------------------------------------------------------ depthwise producers (16 warps, 512 threads)
    pdl_wait();
    const int dtid = threadIdx.x - DW_WARP0 * 32;
    const bool wide = (STRIDE == 1) && g.CT == 64;         // 8 channels x 2 pixels per thread; else 8 channels x 1 pixel
    const int cgs = g.CT >> 3;                             // 8 or 4
    const int cg = dtid & (cgs - 1);
    const int pidx = wide ? (dtid >> 3) : (dtid >> 2);     // pair index 0..63 / pixel index 0..127
    const int oy = wide ? (pidx >> 3) : (pidx >> 4);
    const int ox = wide ? ((pidx & 7) << 1) : (pidx & 15);
    const uint32_t px_bytes = (uint32_t)(g.CT * 2), row_bytes = (uint32_t)(g.BW * g.CT * 2);
    const uint32_t item_off = (uint32_t)((oy * STRIDE) * g.BW + ox * STRIDE) * px_bytes + (uint32_t)(cg * 16);
    const int row0 = oy * TX + ox;
    const uint32_t a_off0 = (uint32_t)(row0 * 128 + ((cg ^ (row0 & 7)) << 4));
    const uint32_t a_off1 = (uint32_t)((row0 + 1) * 128 + ((cg ^ ((row0 + 1) & 7)) << 4));
    int it = 0;
    for (int t = blockIdx.x; t < g.num_tiles; t += gridDim.x) {
      for (int kb = 0; kb < g.num_kb; ++kb, ++it) {
        const int si = it % S_IN, sa = it % S_A;
        const uint32_t w_s = smem_u32(sWpk) + (uint32_t)(((kb * cgs + cg) * 9) * 16);
        const uint32_t b_s = smem_u32(sBdw) + (uint32_t)((kb * g.CT + cg * 8) * 4);
        mbar_wait(&in_full[si], (uint32_t)(it / S_IN) & 1u);                   // input box has landed
        mbar_wait(&a_empty[sa], ((uint32_t)(it / S_A) & 1u) ^ 1u);             // the MMAs that read this A stage are done
        const uint32_t in_s = smem_u32(sIn + si * g.in_stage_bytes) + item_off;
        const uint32_t a_s = smem_u32(sA + sa * A_STAGE_BYTES);
        if (wide) {
          uint4 o[2];
          dw_item<1, 2>(in_s, px_bytes, row_bytes, w_s, b_s, o);
          sts_v4(a_s + a_off0, o[0]);
          sts_v4(a_s + a_off1, o[1]);
        } else {
          uint4 o[1];
          dw_item<STRIDE, 1>(in_s, px_bytes, row_bytes, w_s, b_s, o);
          sts_v4(a_s + a_off0, o[0]);
        }
        fence_proxy_async_smem();              // generic-proxy writes of A -> visible to the tensor core's async proxy
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&a_full[sa]);
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

int dwpw_fused_launch(const void* in, int B, int H, int W, int C_in, int stride, const float* wdw, const float* bdw,
                      const void* wpw, const float* bpw, int C_out, void* out, cudaStream_t st) {
  using namespace fused;
  DVC_REQUIRE(B > 0 && H > 0 && W > 0 && (stride == 1 || stride == 2), "dwpw_fused: bad shape / stride %d", stride);
  DVC_REQUIRE(C_in == 32 || (C_in % 64 == 0 && C_in <= 1024), "dwpw_fused: C_in=%d must be 32 or a multiple of 64", C_in);
  DVC_REQUIRE(C_out % 64 == 0 && C_out >= 64 && C_out <= 256, "dwpw_fused: C_out=%d must be 64, 128, 192 or 256", C_out);
  DVC_REQUIRE(((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(wpw)) & 15) == 0,
              "dwpw_fused: tensors must be 16-byte aligned");
  Args g{};
  g.B = B; g.H = H; g.W = W; g.C_in = C_in; g.C_out = C_out;
  g.Ho = (H + 2 - 3) / stride + 1; g.Wo = (W + 2 - 3) / stride + 1;
  g.CT = (stride == 2 || C_in < 64) ? 32 : 64;          // stride 2: a 17 x 33 box of 64 channels would be 72 KB
  g.num_kb = C_in / g.CT;
  g.BH = (TY - 1) * stride + 3; g.BW = (TX - 1) * stride + 3;
  g.tiles_x = ceil_div(g.Wo, TX); g.tiles_y = ceil_div(g.Ho, TY);
  const long long tiles = (long long)B * g.tiles_x * g.tiles_y;
  DVC_REQUIRE(tiles < (1LL << 31), "dwpw_fused: too many tiles");
  g.num_tiles = (int)tiles;
  int cols = 32;
  while (cols < 2 * C_out) cols <<= 1;
  g.tmem_cols = cols;
  g.wdw = wdw; g.bdw = bdw; g.bpw = bpw;
  auto up = [](int v, int a) { return (v + a - 1) / a * a; };
  g.in_box_bytes = g.BH * g.BW * g.CT * 2;       // what one TMA box delivers (the mbarrier's expect-tx count)
  g.in_stage_bytes = g.in_box_bytes;
  g.b_stage_bytes = C_out * BK * 2;
  // the whole 1x1 weight matrix stays in shared memory when it is small (<= 64 KB); else a ring of 2 chunks
  g.b_resident = (g.num_kb <= S_B_MAX && g.num_kb * g.b_stage_bytes <= 64 * 1024) ? 1 : 0;
  g.S_B = g.b_resident ? g.num_kb : 2;
  g.S_A = (g.S_B * g.b_stage_bytes >= 64 * 1024) ? 3 : S_A_MAX;
  int off = 0;
  g.off_a = off; off += g.S_A * A_STAGE_BYTES;                      // 1024-aligned operand regions first
  g.off_b = off; off += g.S_B * g.b_stage_bytes;
  g.off_epi = off; off += 4 * 2 * EPI_BUF_BYTES;
  g.off_bias = off; off += up(C_out * 4, 128);
  g.off_bdw = off; off += up(C_in * 4, 128);
  g.off_wdw = off; off += up((C_in / 8) * 9 * 16, 128);
  g.off_bar = off; off += 512;
  g.off_in = up(off, 128);
  const int fixed = g.off_in + 1024;                                // slack for the 1024-byte alignment of the dynamic base
  int s_in = (MAX_SMEM - fixed) / up(g.in_stage_bytes, 128);
  if (s_in > S_IN_MAX) s_in = S_IN_MAX;
  DVC_REQUIRE(s_in >= 2, "dwpw_fused: needs more shared memory than there is (C_in=%d C_out=%d stride %d)", C_in, C_out, stride);
  g.S_IN = s_in;
  g.in_stage_bytes = up(g.in_stage_bytes, 128);
  const int smem = g.off_in + s_in * g.in_stage_bytes + 1024;
  CUtensorMap tmIn, tmB, tmOut;
  int rc = make_tmap_nhwc_f16(&tmIn, in, (uint64_t)B, (uint64_t)H, (uint64_t)W, (uint64_t)C_in, (uint32_t)g.CT, (uint32_t)g.BW, (uint32_t)g.BH);
  if (rc) return rc;
  rc = make_tmap_2d_f16(&tmB, wpw, (uint64_t)C_out, (uint64_t)C_in, (uint64_t)C_in, (uint32_t)C_out, BK);
  if (rc) return rc;
  rc = make_tmap_nhwc_f16_sw128(&tmOut, out, (uint64_t)B, (uint64_t)g.Ho, (uint64_t)g.Wo, (uint64_t)C_out, EPI_CHUNK, TX, 2);
  if (rc) return rc;
  static std::atomic<bool> attr_set[64];
  int dev = 0;
  DVC_CHECK_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63].load(std::memory_order_acquire)) {
    DVC_CHECK_CUDA(cudaFuncSetAttribute(dwpw_fused_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_SMEM));
    DVC_CHECK_CUDA(cudaFuncSetAttribute(dwpw_fused_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, MAX_SMEM));
    attr_set[dev & 63].store(true, std::memory_order_release);
  }
  const int sms = fused_num_sms();
  const int grid = g.num_tiles < sms ? g.num_tiles : sms;
  const Args& ga = g;
  if (stride == 1) DVC_CHECK_CUDA(launch_pdl(dwpw_fused_kernel<1>, dim3(grid), dim3(THREADS), (size_t)smem, st, 1, tmIn, tmB, tmOut, ga));
  else DVC_CHECK_CUDA(launch_pdl(dwpw_fused_kernel<2>, dim3(grid), dim3(THREADS), (size_t)smem, st, 1, tmIn, tmB, tmOut, ga));
  count_launch();
  DVC_CHECK_CUDA(cudaGetLastError());
  return DVC_OK;
}

int dwpw_fused_s1_launch(const void* in, int B, int H, int W, int C_in, const float* wdw, const float* bdw,
                         const void* wpw, const float* bpw, int C_out, void* out, cudaStream_t st) {
  return dwpw_fused_launch(in, B, H, W, C_in, 1, wdw, bdw, wpw, bpw, C_out, out, st);
}

}  // namespace dvc

/* Fused depthwise 3x3 (stride 1 or 2, pad 1) + pointwise 1x1, both with bias + ReLU.
 * in [B,H,W,C_in] fp16 NHWC, wdw [9][C_in] fp32, bdw [C_in], wpw [C_out][C_in] fp16, bpw [C_out], out [B,Ho,Wo,C_out] fp16. */
extern "C" int dvc_dwpw_fused(const void* in_f16, int B, int H, int W, int C_in, int stride, const float* wdw, const float* bdw,
                              const void* wpw_f16, const float* bpw, int C_out, void* out_f16, void* stream) {
  return dvc::dwpw_fused_launch(in_f16, B, H, W, C_in, stride, wdw, bdw, wpw_f16, bpw, C_out, out_f16, (cudaStream_t)stream);
}
extern "C" int dvc_dwpw_fused_s1(const void* in_f16, int B, int H, int W, int C_in, const float* wdw, const float* bdw,
                                 const void* wpw_f16, const float* bpw, int C_out, void* out_f16, void* stream) {
  return dvc::dwpw_fused_launch(in_f16, B, H, W, C_in, 1, wdw, bdw, wpw_f16, bpw, C_out, out_f16, (cudaStream_t)stream);
}
