// K2 / K5: pointwise (1x1) convolution and SSD head GEMMs on the 5th-gen tensor cores.
//
//   out[M, N] = act( A[M, K] * Wt[N, K]^T + bias[N] )
//
// A is the NHWC fp16 activation matrix (M = batch * H * W pixels, K = Cin), Wt the
// [Cout, Cin] weights; both are K-major, so both operands are TMA-loaded as 128-byte
// swizzled tiles and consumed by tcgen05.mma straight from shared memory.  The fp32
// accumulator tile (128 x block_n) lives in TMEM and is drained by four epilogue warps
// with tcgen05.ld -> bias -> ReLU -> fp16 NHWC store (or the SSD loc/conf split store,
// which makes Permute/Flatten/Concat of prototxt:1172-1192,1817-1844 free).
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer
// (one elected lane), warps 2..5 = epilogue (TMEM lane quadrant = warp % 4).
// Reference op: cv2.dnn Convolution layers, prototxt:76-106 ... 1083-1113, 1147-1775.
#include "common.cuh"

namespace dvc {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int GEMM_THREADS = 192;
constexpr int A_STAGE_BYTES = BM * BK * 2;  // 16 KiB
constexpr int GEMM_MAX_SMEM = 227 * 1024;

struct GemmArgs {
  int M, N, K;
  int block_n, n_tiles;
  int num_kb, stages;
  int tmem_cols;
  const float* bias;
  int out_mode;
  __half* out;
  int ldc;
  int relu;
  dvc_head_out head;
};

__global__ void __launch_bounds__(GEMM_THREADS)
gemm_f16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmArgs g) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = __shfl_sync(0xffffffff, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  const int n_tile = blockIdx.x % g.n_tiles;
  const int m_tile = blockIdx.x / g.n_tiles;
  const int b_stage_bytes = g.block_n * BK * 2;

  uint8_t* sA = smem;
  uint8_t* sB = smem + g.stages * A_STAGE_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(sB + g.stages * b_stage_bytes);
  uint64_t* empty_bar = full_bar + g.stages;
  uint64_t* tmem_full_bar = empty_bar + g.stages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < g.stages; ++s) {
        mbar_init(&full_bar[s], 1);
        mbar_init(&empty_bar[s], 1);
      }
      mbar_init(tmem_full_bar, 1);
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

  if (warp == 0) {
    // ------------------------------------------------------------- TMA producer
    if (elect_one()) {
      const uint32_t stage_bytes = (uint32_t)(A_STAGE_BYTES + b_stage_bytes);
      for (int kb = 0; kb < g.num_kb; ++kb) {
        const int s = kb % g.stages;
        const uint32_t ph = (uint32_t)(kb / g.stages) & 1u;
        mbar_wait(&empty_bar[s], ph ^ 1u);
        mbar_arrive_expect_tx(&full_bar[s], stage_bytes);
        tma_load_2d(sA + s * A_STAGE_BYTES, &tmA, &full_bar[s], kb * BK, m_tile * BM);
        tma_load_2d(sB + s * b_stage_bytes, &tmB, &full_bar[s], kb * BK, n_tile * g.block_n);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------- MMA issuer
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_f16(BM, (uint32_t)g.block_n);
      for (int kb = 0; kb < g.num_kb; ++kb) {
        const int s = kb % g.stages;
        const uint32_t ph = (uint32_t)(kb / g.stages) & 1u;
        mbar_wait(&full_bar[s], ph);
        tc_fence_after();
        const uint64_t a_desc = umma_desc_sw128(smem_u32(sA + s * A_STAGE_BYTES));
        const uint64_t b_desc = umma_desc_sw128(smem_u32(sB + s * b_stage_bytes));
        int ksteps = (g.K - kb * BK + 15) / 16;
        ksteps = ksteps > 4 ? 4 : ksteps;
        for (int k = 0; k < ksteps; ++k) {
          // advance 16 fp16 = 32 bytes along K inside the 128-byte swizzle atom (+2 in 16-byte units)
          umma_f16(tmem_base, a_desc + (uint64_t)(k * 2), b_desc + (uint64_t)(k * 2), idesc, (uint32_t)((kb | k) != 0));
        }
        umma_commit(&empty_bar[s]);   // frees this smem stage once the MMAs above retire
      }
      umma_commit(tmem_full_bar);     // accumulator complete -> epilogue
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------- epilogue (4 warps)
    const int q = warp & 3;
    const int row = m_tile * BM + q * 32 + lane;
    const bool row_ok = row < g.M;
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
    const int n0 = n_tile * g.block_n;
    if (g.out_mode == 0) {
      __half* orow = g.out + (size_t)row * g.ldc + n0;
      for (int c0 = 0; c0 < g.block_n; c0 += 16) {
        float v[16];
        tmem_ld16(taddr + (uint32_t)c0, v);
        tmem_ld_wait();
        const float4* bp = reinterpret_cast<const float4*>(g.bias + n0 + c0);
        uint32_t pk[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float4 b4 = __ldg(bp + j);
          float x0 = v[4 * j + 0] + b4.x, x1 = v[4 * j + 1] + b4.y, x2 = v[4 * j + 2] + b4.z, x3 = v[4 * j + 3] + b4.w;
          if (g.relu) {
            x0 = fmaxf(x0, 0.f); x1 = fmaxf(x1, 0.f); x2 = fmaxf(x2, 0.f); x3 = fmaxf(x3, 0.f);
          }
          pk[2 * j] = pack_half2(x0, x1);
          pk[2 * j + 1] = pack_half2(x2, x3);
        }
        if (row_ok) {
          uint4* dst = reinterpret_cast<uint4*>(orow + c0);
          dst[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
          dst[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
        }
      }
    } else {
      const int b = row_ok ? row / g.head.hw : 0;
      const int p = row_ok ? row - b * g.head.hw : 0;
      float* lrow = g.head.loc + (size_t)b * g.head.loc_pitch + g.head.loc_off + (size_t)p * g.head.n_loc;
      float* crow = g.head.conf + (size_t)b * g.head.conf_pitch + g.head.conf_off + (size_t)p * g.head.n_conf;
      const int n_valid = g.head.n_loc + g.head.n_conf;
      for (int c0 = 0; c0 < g.block_n; c0 += 16) {
        float v[16];
        tmem_ld16(taddr + (uint32_t)c0, v);
        tmem_ld_wait();
        if (row_ok) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int c = n0 + c0 + j;
            if (c < n_valid) {
              float x = v[j] + __ldg(g.bias + c);
              if (g.relu) x = fmaxf(x, 0.f);
              if (c < g.head.n_loc) lrow[c] = x;
              else crow[c - g.head.n_loc] = x;
            }
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, (uint32_t)g.tmem_cols);
}

// ------------------------------------------------------------------------------------ host

static int pick_block_n(int M, int N) {
  // Widest tile dividing N that still yields >= 2 waves of CTAs; otherwise the narrowest
  // tile >= 64 (more CTAs in flight beats tile width for these bandwidth-bound shapes).
  const int cand[] = {256, 160, 128, 80, 64, 48, 32, 16};
  const int m_tiles = ceil_div(M, BM);
  int fallback = 0;
  for (int c : cand) {
    if (N % c) continue;
    if (m_tiles * (N / c) >= 2 * 148) return c;
    if (c >= 64 || !fallback) fallback = c;
  }
  return fallback ? fallback : 16;
}

struct GemmLaunch {
  CUtensorMap tmA, tmB;
  GemmArgs args;
  int grid;
  int smem;
};

int gemm_prepare(GemmLaunch* L, const void* A, const void* Wt, const float* bias, int M, int N, int K, int lda,
                 int out_mode, void* out, int ldc, int relu, const dvc_head_out* head, int block_n) {
  DVC_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: bad shape M=%d N=%d K=%d", M, N, K);
  DVC_REQUIRE(N % 16 == 0, "gemm: N=%d must be a multiple of 16 (pad the weights)", N);
  DVC_REQUIRE(K % 16 == 0 && lda % 8 == 0, "gemm: K=%d must be a multiple of 16 and lda=%d of 8", K, lda);
  DVC_REQUIRE((reinterpret_cast<uintptr_t>(A) & 15) == 0 && (reinterpret_cast<uintptr_t>(Wt) & 15) == 0,
              "gemm: operands must be 16-byte aligned");
  if (block_n <= 0) block_n = pick_block_n(M, N);
  DVC_REQUIRE(block_n % 16 == 0 && block_n <= 256 && N % block_n == 0, "gemm: block_n=%d invalid for N=%d", block_n, N);
  if (out_mode == 0) {
    DVC_REQUIRE(out != nullptr && ldc % 8 == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0, "gemm: bad output/ldc=%d", ldc);
  } else {
    DVC_REQUIRE(head != nullptr && head->loc && head->conf && head->hw > 0, "gemm: head output missing");
    DVC_REQUIRE(head->n_loc + head->n_conf <= N, "gemm: head columns %d > N=%d", head->n_loc + head->n_conf, N);
  }
  GemmArgs& g = L->args;
  g.M = M; g.N = N; g.K = K;
  g.block_n = block_n;
  g.n_tiles = N / block_n;
  g.num_kb = ceil_div(K, BK);
  g.stages = g.num_kb < 4 ? g.num_kb : 4;
  int cols = 32;
  while (cols < block_n) cols <<= 1;
  g.tmem_cols = cols;
  g.bias = bias;
  g.out_mode = out_mode;
  g.out = reinterpret_cast<__half*>(out);
  g.ldc = ldc;
  g.relu = relu;
  if (head) g.head = *head; else g.head = dvc_head_out{};
  L->grid = ceil_div(M, BM) * g.n_tiles;
  L->smem = g.stages * (A_STAGE_BYTES + block_n * BK * 2) + 1024 + 256;
  DVC_REQUIRE(L->smem <= GEMM_MAX_SMEM, "gemm: smem %d too large", L->smem);
  int rc = make_tmap_2d_f16(&L->tmA, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, BM, BK);
  if (rc) return rc;
  rc = make_tmap_2d_f16(&L->tmB, Wt, (uint64_t)N, (uint64_t)K, (uint64_t)K, (uint32_t)block_n, BK);
  return rc;
}

int gemm_launch(const GemmLaunch* L, cudaStream_t st) {
  static bool attr_set[64] = {};
  int dev = 0;
  DVC_CHECK_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    DVC_CHECK_CUDA(cudaFuncSetAttribute(gemm_f16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_MAX_SMEM));
    attr_set[dev & 63] = true;
  }
  gemm_f16_kernel<<<L->grid, GEMM_THREADS, L->smem, st>>>(L->tmA, L->tmB, L->args);
  count_launch();
  DVC_CHECK_CUDA(cudaGetLastError());
  return DVC_OK;
}

GemmLaunch* gemm_alloc() { return new GemmLaunch(); }
void gemm_free(GemmLaunch* L) { delete L; }

}  // namespace dvc

extern "C" int dvc_gemm_f16(const void* A, const void* Wt, const float* bias, int M, int N, int K, int lda, int out_mode,
                            void* out_f16, int ldc, int relu, const dvc_head_out* head, int block_n, void* stream) {
  dvc::GemmLaunch L;
  int rc = dvc::gemm_prepare(&L, A, Wt, bias, M, N, K, lda, out_mode, out_f16, ldc, relu, head, block_n);
  if (rc) return rc;
  return dvc::gemm_launch(&L, (cudaStream_t)stream);
}
