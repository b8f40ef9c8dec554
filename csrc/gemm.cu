// K2 / K5: pointwise (1x1) convolution and SSD head GEMMs on the 5th-gen tensor cores.
//
//   out[M, N] = act( A[M, K] * Wt[N, K]^T + bias[N] )
//
// A is the NHWC fp16 activation matrix (M = batch * H * W pixels, K = Cin), Wt the
// [Cout, Cin] weights; both are K-major, so both operands are TMA-loaded as 128-byte
// swizzled tiles and consumed by tcgen05.mma straight from shared memory.
//
// Persistent, warp-specialised (192 threads, one CTA per SM, tiles strided by gridDim):
//   warp 0      TMA producer: runs ahead across tiles through a ring of smem stages
//   warp 1      TMEM allocator + MMA issuer (one elected lane); two accumulator buffers in
//               TMEM so tile i+1 is multiplied while tile i is drained
//   warps 2..5  epilogue: tcgen05.ld -> +bias -> ReLU -> fp16 -> swizzled smem staging ->
//               TMA store (asynchronous, coalesced NHWC rows, M tail clipped by the tensor map);
//               SSD heads instead split-store fp32 loc/conf rows at their concat offsets, which
//               makes Permute/Flatten/Concat of prototxt:1172-1192,1817-1844 free.
// These layers have K = 32..1024: most are bandwidth-bound, so what matters is keeping loads
// and stores of several tiles in flight per SM, which is what the producer run-ahead and the
// bulk-async stores are for.
// Reference op: cv2.dnn Convolution layers, prototxt:76-106 ... 1083-1113, 1147-1775.
#include <stdlib.h>

#include <atomic>

#include "common.cuh"

namespace dvc {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int GEMM_THREADS = 320;               // 2 control warps + 2 groups of 4 epilogue warps
constexpr int A_STAGE_BYTES = BM * BK * 2;       // 16 KiB
constexpr int GEMM_MAX_SMEM = 227 * 1024;
constexpr int EPI_CHUNK = 64;                    // columns per staging buffer (128 B of fp16 per row)
constexpr int EPI_BUF_BYTES = 32 * EPI_CHUNK * 2;   // one warp, 32 rows x 128 B
constexpr int EPI_SMEM = 2 * 4 * 2 * EPI_BUF_BYTES;   // 2 groups x 4 warps x 2 buffers
constexpr int BIAS_MAX_N = 2048;
constexpr int BIAS_SMEM = BIAS_MAX_N * 4;        // the whole bias vector, loaded once per CTA

struct GemmArgs {
  int M, N, K;
  int block_n, n_tiles, num_tiles;
  int num_kb, stages;
  int tmem_cols;
  const float* bias;
  int out_mode;        // 0: fp16 [M, ldc]   1: SSD head split store (fp32)
  int tma_store;       // out_mode 0 only: epilogue goes through smem + TMA store
  int cluster;         // CTAs per cluster (2 only in CTA-pair mode, else 1)
  int two_cta;         // cta_group::2: a CTA pair multiplies one 256 x block_n tile (cluster == 2, no multicast)
  int m_tiles, num_ctiles;
  int b_resident;      // whole weight matrix stays in smem (one N tile, few K blocks): loaded once per CTA
  __half* out;
  int ldc;
  int relu;
  dvc_head_out head;
  // implicit 3x3 / stride-2 / pad-1 convolution (conv_c > 0): A is never materialised, tmA is an im2col-mode map over the
  // NHWC input and K runs over (kh, kw, c) with c fastest, i.e. the layout im2col3x3s2_kernel would have written
  int conv_c, conv_wo, conv_howo;
};

// The activation operand of an implicit convolution (gemm_prepare's optional last argument).
struct ConvA {
  int B, H, W, C;
};

// PAIR = 1 is the cta_group::2 build: ptxas marks a kernel containing cta_group::2 instructions as needing an even
// cluster, so the single-CTA path is a separate instantiation.
template <int PAIR>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_f16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const __grid_constant__ CUtensorMap tmC, const GemmArgs g) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = __shfl_sync(0xffffffff, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  // pair mode: each CTA holds only its half of the weight tile (block_n / 2 rows)
  const int b_stage_bytes = (PAIR ? g.block_n / 2 : g.block_n) * BK * 2;

  uint8_t* sEpi = smem;                                   // [4 warps][2][32 rows x 128 B], 1024-aligned
  float* sBias = reinterpret_cast<float*>(smem + EPI_SMEM);
  uint8_t* sA = smem + EPI_SMEM + BIAS_SMEM;
  uint8_t* sB = sA + g.stages * A_STAGE_BYTES;
  const int b_tiles = g.b_resident ? g.num_kb : g.stages;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(sB + b_tiles * b_stage_bytes);
  uint64_t* empty_bar = full_bar + g.stages;
  uint64_t* tmem_full_bar = empty_bar + g.stages;         // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;           // [2]
  uint64_t* bres_bar = tmem_empty_bar + 2;                // weights resident
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bres_bar + 1);

  pdl_launch_dependents();
  // everything up to pdl_wait() touches only constants (bias, weights, descriptors) and on-chip state, so it overlaps
  // the tail of the previous kernel when this one was launched programmatically
  for (int i = threadIdx.x; i < g.N; i += GEMM_THREADS) sBias[i] = __ldg(g.bias + i);
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (g.tma_store) tma_prefetch_desc(&tmC);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < g.stages; ++s) {
        mbar_init(&full_bar[s], 1);
        // multicast clusters: every CTA must have drained the stage; pair mode: one multicast commit from the leader
        mbar_init(&empty_bar[s], PAIR ? 1u : (uint32_t)g.cluster);
      }
      for (int a = 0; a < 2; ++a) {
        mbar_init(&tmem_full_bar[a], 1);
        mbar_init(&tmem_empty_bar[a], PAIR ? 8u : 4u);     // one arrive per epilogue warp (of both CTAs in pair mode)
      }
      mbar_init(bres_bar, 1);
      fence_mbar_init();
    }
    __syncwarp();
    if (PAIR) {
      tmem_alloc_pair(tmem_slot, (uint32_t)g.tmem_cols);
      tmem_relinquish_pair();
    } else {
      tmem_alloc(tmem_slot, (uint32_t)g.tmem_cols);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  if (g.cluster > 1) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // cluster-tile schedule: a cluster walks (m-group, n-tile) pairs; CTA rank r owns m-tile group*cluster + r
  const int crank = g.cluster > 1 ? (int)cluster_ctarank() : 0;
  const int cluster_id = blockIdx.x / g.cluster, num_clusters = gridDim.x / g.cluster;
  const uint16_t cmask = (uint16_t)((1u << g.cluster) - 1u);

  if (warp == 0) {
    // ------------------------------------------------------------- TMA producer
    if (elect_one()) {
      const uint32_t stage_bytes = (uint32_t)(A_STAGE_BYTES + (g.b_resident ? 0 : b_stage_bytes));
      if (g.b_resident) {
        // small weight matrices are fetched once: re-reading the same few KB from every SM for every tile
        // hot-spots one L2 slice and starves the pipeline (measured on conv1: 38% of samples in the
        // epilogue's wait for the accumulator)
        mbar_arrive_expect_tx(bres_bar, (uint32_t)(g.num_kb * b_stage_bytes));
        for (int kb = 0; kb < g.num_kb; ++kb) tma_load_2d(sB + kb * b_stage_bytes, &tmB, bres_bar, kb * BK, 0);
      }
      pdl_wait();               // A is the previous kernel's output; the (constant) resident weights are already in flight
      int it = 0;
      const int slice_rows = g.block_n / g.cluster;
      for (int ct = cluster_id; ct < g.num_ctiles; ct += num_clusters) {
        const int n_tile = ct % g.n_tiles, m_tile = (ct / g.n_tiles) * g.cluster + crank;
        for (int kb = 0; kb < g.num_kb; ++kb, ++it) {
          const int s = it % g.stages;
          const uint32_t ph = (uint32_t)(it / g.stages) & 1u;
          mbar_wait(&empty_bar[s], ph ^ 1u);
          if (PAIR) {
            // both CTAs load their own A rows and their half of the weight tile; all bytes are credited to the
            // leader's barrier, which the leader arms for the pair
            if (crank == 0) mbar_arrive_expect_tx(&full_bar[s], 2u * stage_bytes);
            const uint32_t lead_bar = mapa_u32(smem_u32(&full_bar[s]), 0);
            tma_load_2d_pair(sA + s * A_STAGE_BYTES, &tmA, lead_bar, kb * BK, m_tile * BM);
            tma_load_2d_pair(sB + s * b_stage_bytes, &tmB, lead_bar, kb * BK, n_tile * g.block_n + crank * slice_rows);
            continue;
          }
          mbar_arrive_expect_tx(&full_bar[s], stage_bytes);
          if (g.conv_c) {
            // K block kb = 64 channels of one filter tap; the 128 rows are 128 consecutive output pixels, which the
            // TMA unit walks (w, then h, then n) from the tile's first pixel, zero-filling the padding halo
            const int m0 = m_tile * BM;
            const int n_img = m0 / g.conv_howo, r = m0 - n_img * g.conv_howo;
            const int oh = r / g.conv_wo, ow = r - oh * g.conv_wo;
            const int tap = (kb * BK) / g.conv_c, c0 = kb * BK - tap * g.conv_c;
            const int kh = tap / 3, kw = tap - kh * 3;
            tma_load_im2col_4d(sA + s * A_STAGE_BYTES, &tmA, &full_bar[s], c0, 2 * ow - 1, 2 * oh - 1, n_img, kw, kh);
          } else {
            tma_load_2d(sA + s * A_STAGE_BYTES, &tmA, &full_bar[s], kb * BK, m_tile * BM);
          }
          if (!g.b_resident) {
            tma_load_2d(sB + s * b_stage_bytes, &tmB, &full_bar[s], kb * BK, n_tile * g.block_n);
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------- MMA issuer
    pdl_wait();
    if (PAIR && crank != 0) {
      // pair mode: only the leader CTA issues MMAs (they drive both SMs' tensor cores)
    } else if (elect_one()) {
      const uint32_t idesc = umma_idesc_f16(PAIR ? 2 * BM : BM, (uint32_t)g.block_n);
      if (g.b_resident) mbar_wait(bres_bar, 0);
      int it = 0, t = 0;
      for (int ct = cluster_id; ct < g.num_ctiles; ct += num_clusters, ++t) {
        const int acc = t & 1;
        mbar_wait(&tmem_empty_bar[acc], (((uint32_t)t >> 1) & 1u) ^ 1u);   // epilogue has drained this buffer
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * g.block_n);
        for (int kb = 0; kb < g.num_kb; ++kb, ++it) {
          const int s = it % g.stages;
          const uint32_t ph = (uint32_t)(it / g.stages) & 1u;
          mbar_wait(&full_bar[s], ph);
          tc_fence_after();
          const uint64_t a_desc = umma_desc_sw128(smem_u32(sA + s * A_STAGE_BYTES));
          const uint64_t b_desc = umma_desc_sw128(smem_u32(sB + (g.b_resident ? kb : s) * b_stage_bytes));
          int ksteps = (g.K - kb * BK + 15) / 16;
          ksteps = ksteps > 4 ? 4 : ksteps;
          for (int k = 0; k < ksteps; ++k) {
            // advance 16 fp16 = 32 bytes along K inside the 128-byte swizzle atom (+2 in 16-byte units)
            if (PAIR) umma_f16_pair(d_tmem, a_desc + (uint64_t)(k * 2), b_desc + (uint64_t)(k * 2), idesc, (uint32_t)((kb | k) != 0));
            else umma_f16(d_tmem, a_desc + (uint64_t)(k * 2), b_desc + (uint64_t)(k * 2), idesc, (uint32_t)((kb | k) != 0));
          }
          // frees this smem stage (in every CTA of the cluster: peers multicast into it) once the MMAs retire
          if (PAIR) umma_commit_pair(&empty_bar[s], 3);
          else if (g.cluster > 1) umma_commit_mcast(&empty_bar[s], cmask);
          else umma_commit(&empty_bar[s]);
        }
        // accumulator complete -> epilogue (of both CTAs in pair mode)
        if (PAIR) umma_commit_pair(&tmem_full_bar[acc], 3); else umma_commit(&tmem_full_bar[acc]);
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------- epilogue (2 groups x 4 warps)
    // Group e drains accumulator buffer e, i.e. every second tile of this CTA, so two tiles are in
    // the epilogue at once while the MMA warp fills the buffer the other group has just released.
    pdl_wait();                                // the output buffer may still be read by an earlier kernel
    const int q = warp & 3;                    // TMEM lane quadrant this warp may access
    const int grp = (warp - 2) >> 2;
    const uint32_t bias_s = smem_u32(sBias);
    uint8_t* my_epi = sEpi + (grp * 4 + q) * 2 * EPI_BUF_BYTES;
    int chunk_it = 0;
    for (int t = grp; ; t += 2) {
      const int ct = cluster_id + t * num_clusters;
      if (ct >= g.num_ctiles) break;
      const int n_tile = ct % g.n_tiles, m_tile = (ct / g.n_tiles) * g.cluster + crank;
      const int acc = grp;
      const int n0 = n_tile * g.block_n;
      const int row0 = m_tile * BM + q * 32;
      const int row = row0 + lane;
      const bool row_ok = row < g.M;
      mbar_wait(&tmem_full_bar[acc], ((uint32_t)t >> 1) & 1u);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * g.block_n);
      if (g.out_mode == 0 && g.tma_store) {
        for (int c0 = 0; c0 < g.block_n; c0 += EPI_CHUNK, ++chunk_it) {
          uint8_t* buf = my_epi + (chunk_it & 1) * EPI_BUF_BYTES;
          // the TMA store issued two chunks ago from this buffer must have finished reading it
          if (lane == 0) tma_store_wait_read<1>();
          __syncwarp();
          float v[EPI_CHUNK];
#pragma unroll
          for (int j = 0; j < EPI_CHUNK / 16; ++j) tmem_ld16(taddr + (uint32_t)(c0 + 16 * j), v + 16 * j);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < EPI_CHUNK / 8; ++j) {          // 8 x 16-byte chunks of this lane's row
            uint32_t pk[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float2 bb = lds_f2(bias_s + (uint32_t)((n0 + c0 + 8 * j + 2 * e) * 4));
              float x0 = v[8 * j + 2 * e] + bb.x;
              float x1 = v[8 * j + 2 * e + 1] + bb.y;
              if (g.relu) { x0 = fmaxf(x0, 0.f); x1 = fmaxf(x1, 0.f); }
              pk[e] = pack_half2(x0, x1);
            }
            // 128-byte swizzle: 16-byte chunk j of row r lives at chunk position j ^ (r & 7)
            sts_v4(smem_u32(buf) + (uint32_t)(lane * 128 + ((j ^ (lane & 7)) << 4)), make_uint4(pk[0], pk[1], pk[2], pk[3]));
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0 && row0 < g.M) {
            tma_store_2d(&tmC, buf, n0 + c0, row0);
            tma_store_commit();
          }
        }
      } else if (g.out_mode == 0) {
        __half* orow = g.out + (size_t)row * g.ldc + n0;
        for (int c0 = 0; c0 < g.block_n; c0 += 16) {
          float v[16];
          tmem_ld16(taddr + (uint32_t)c0, v);
          tmem_ld_wait();
          uint32_t pk[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float2 bb = lds_f2(bias_s + (uint32_t)((n0 + c0 + 2 * j) * 4));
            float x0 = v[2 * j] + bb.x, x1 = v[2 * j + 1] + bb.y;
            if (g.relu) { x0 = fmaxf(x0, 0.f); x1 = fmaxf(x1, 0.f); }
            pk[j] = pack_half2(x0, x1);
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
                float x = v[j] + lds_f1(bias_s + (uint32_t)(c * 4));
                if (g.relu) x = fmaxf(x, 0.f);
                if (c < g.head.n_loc) lrow[c] = x;
                else crow[c - g.head.n_loc] = x;
              }
            }
          }
        }
      }
      // all tcgen05.ld of this accumulator have completed (wait::ld above): hand it back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (PAIR) mbar_arrive_cluster(mapa_u32(smem_u32(&tmem_empty_bar[acc]), 0));   // the leader's MMA warp waits
        else mbar_arrive(&tmem_empty_bar[acc]);
      }
    }
    if (g.tma_store && lane == 0) tma_store_wait_all<0>();
  }

  tc_fence_before();
  if (g.cluster > 1) cluster_sync_all(); else __syncthreads();   // peers may still multicast into / arrive on this CTA
  if (warp == 1) {
    if (PAIR) tmem_dealloc_pair(tmem_base, (uint32_t)g.tmem_cols);
    else tmem_dealloc(tmem_base, (uint32_t)g.tmem_cols);
  }
}

// ------------------------------------------------------------------------------------ host

static int num_sms() {
  static std::atomic<int> cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (!cached[dev & 63].load(std::memory_order_relaxed)) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev & 63].store(n, std::memory_order_relaxed);
  }
  return cached[dev & 63].load(std::memory_order_relaxed);
}

static int pick_block_n(int M, int N) {
  // Persistent kernel: the widest tile that still gives most SMs a tile (fewer re-reads of A,
  // wider TMA stores); otherwise the narrowest tile that is still >= 64 columns.  The 60% threshold is from
  // scripts/sweep_gemm.py: conv14_1 (50 M tiles, N=256) runs 16.4 us with 100 tiles of 128 vs 20.5 us with 200 of 64.
  const int cand[] = {256, 160, 128, 80, 64};
  const int m_tiles = ceil_div(M, BM);
  const int sms = num_sms();
  int last = 0;
  for (int c : cand) {
    if (N % c) continue;
    last = c;
    if (m_tiles * (N / c) * 10 >= sms * 6) return c;
  }
  if (last) return last;
  return N <= 256 ? N : 16;     // N in {16, 32, 48}
}

struct GemmLaunch {
  CUtensorMap tmA, tmB, tmC;
  GemmArgs args;
  int grid;
  int smem;
};

static int gemm_prepare_impl(GemmLaunch* L, const void* A, const void* Wt, const float* bias, int M, int N, int K, int lda,
                             int out_mode, void* out, int ldc, int relu, const dvc_head_out* head, int block_n, const ConvA* conv) {
  DVC_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: bad shape M=%d N=%d K=%d", M, N, K);
  DVC_REQUIRE(N % 16 == 0 && N <= BIAS_MAX_N, "gemm: N=%d must be a multiple of 16 (pad the weights) and <= %d", N, BIAS_MAX_N);
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
  g.conv_c = g.conv_wo = g.conv_howo = 0;
  if (conv) {
    const int Ho = (conv->H - 1) / 2 + 1, Wo = (conv->W - 1) / 2 + 1;
    DVC_REQUIRE(conv->C % BK == 0 && conv->H >= 2 && conv->W >= 2 && M == conv->B * Ho * Wo && K == 9 * conv->C,
                "gemm: implicit conv needs C %% %d == 0 (C=%d), H, W >= 2 and M, K matching the input shape", BK, conv->C);
    g.conv_c = conv->C; g.conv_wo = Wo; g.conv_howo = Ho * Wo;
  }
  g.M = M; g.N = N; g.K = K;
  g.block_n = block_n;
  g.n_tiles = N / block_n;
  g.num_tiles = ceil_div(M, BM) * g.n_tiles;
  g.num_kb = ceil_div(K, BK);
  const int b_tile_bytes = block_n * BK * 2;
  g.m_tiles = ceil_div(M, BM);
  g.b_resident = (g.n_tiles == 1 && g.num_kb * b_tile_bytes <= 64 * 1024 && g.num_tiles > 1) ? 1 : 0;
  // The weight pointer is caller-supplied, so weight tiles are fetched per CTA: TMA multicast loads here read only
  // this GPU's memory, since on B200 a TMA multicast load from another GPU's memory hung the GPU.  (The multicast
  // variant measured no gain on these shapes anyway: conv7 24.5 us with 1, 2 or 4 CTAs per cluster.)
  int cluster = 1;
  // CTA pairs (tcgen05 cta_group::2, DVC_GEMM_2CTA=<min K blocks>): one 256 x block_n tile per pair, each CTA stages
  // its 128 rows of A and half of the weight tile, so a stage is 16 + block_n/8 KB instead of 16 + block_n/4 KB, more
  // K blocks fit in flight and each SM reads 8 KB instead of 12 KB of operands per MMA.  Measured on B200
  // (scripts/check_2cta.py, bit-identical outputs): M=23104 N=512 K=4096 105.5 -> 95.0 us, conv13 26.6 -> 25.4 us,
  // conv7 24.6 -> 24.6 us, K<=256 shapes 2-3 us slower (cluster launch + cross-CTA barriers).  The network's K is at
  // most 1024, where fill/drain and wave quantisation dominate, so the pair path is opt-in (off by default).
  g.two_cta = 0;
  {
    const char* env = getenv("DVC_GEMM_2CTA");
    const int want = env ? atoi(env) : 0;
    if (want && !conv && cluster == 1 && !g.b_resident && block_n % 32 == 0 && g.m_tiles >= 2 && g.num_kb >= want) {
      g.two_cta = 1;
      cluster = 2;
    }
  }
  g.cluster = cluster;
  g.num_ctiles = ceil_div(g.m_tiles, cluster) * g.n_tiles;
  const int stage_bytes = A_STAGE_BYTES + (g.b_resident ? 0 : (g.two_cta ? b_tile_bytes / 2 : b_tile_bytes));
  const int fixed = EPI_SMEM + BIAS_SMEM + 1024 + 256 + (g.b_resident ? g.num_kb * b_tile_bytes : 0);
  int stages = (GEMM_MAX_SMEM - fixed) / stage_bytes;
  if (stages > 8) stages = 8;
  DVC_REQUIRE(stages >= 2, "gemm: tile does not fit shared memory");
  g.stages = stages;
  int cols = 32;
  while (cols < 2 * block_n) cols <<= 1;
  g.tmem_cols = cols;
  g.bias = bias;
  g.out_mode = out_mode;
  g.tma_store = (out_mode == 0 && block_n % EPI_CHUNK == 0) ? 1 : 0;
  g.out = reinterpret_cast<__half*>(out);
  g.ldc = ldc;
  g.relu = relu;
  if (head) g.head = *head; else g.head = dvc_head_out{};
  const int sms = num_sms();
  int num_clusters = (cluster == 4 ? 132 : sms) / cluster;      // 4-CTA clusters cannot use every SM (GPC shapes)
  if (g.num_ctiles < num_clusters) num_clusters = g.num_ctiles;
  L->grid = num_clusters * cluster;
  L->smem = g.stages * stage_bytes + fixed;
  int rc = conv ? make_tmap_im2col3x3s2_f16(&L->tmA, A, (uint64_t)conv->B, (uint64_t)conv->H, (uint64_t)conv->W, (uint64_t)conv->C, BK, BM)
                : make_tmap_2d_f16(&L->tmA, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, BM, BK);
  if (rc) return rc;
  rc = make_tmap_2d_f16(&L->tmB, Wt, (uint64_t)N, (uint64_t)K, (uint64_t)K, (uint32_t)(block_n / cluster), BK);
  if (rc) return rc;
  if (g.tma_store) {
    rc = make_tmap_2d_f16(&L->tmC, out, (uint64_t)M, (uint64_t)N, (uint64_t)ldc, 32, EPI_CHUNK);
    if (rc) return rc;
  } else {
    L->tmC = L->tmA;
  }
  return DVC_OK;
}

int gemm_prepare(GemmLaunch* L, const void* A, const void* Wt, const float* bias, int M, int N, int K, int lda,
                 int out_mode, void* out, int ldc, int relu, const dvc_head_out* head, int block_n) {
  return gemm_prepare_impl(L, A, Wt, bias, M, N, K, lda, out_mode, out, ldc, relu, head, block_n, nullptr);
}

int gemm_launch(const GemmLaunch* L, cudaStream_t st) {
  static std::atomic<bool> attr_set[64];   // zero-initialised; engine threads of different GPUs may race here
  int dev = 0;
  DVC_CHECK_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63].load(std::memory_order_acquire)) {
    DVC_CHECK_CUDA(cudaFuncSetAttribute(gemm_f16_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_MAX_SMEM));
    DVC_CHECK_CUDA(cudaFuncSetAttribute(gemm_f16_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_MAX_SMEM));
    attr_set[dev & 63].store(true, std::memory_order_release);
  }
  if (L->args.cluster > 1) {
    cudaError_t le = L->args.two_cta ? launch_pdl(gemm_f16_kernel<1>, dim3(L->grid), dim3(GEMM_THREADS), (size_t)L->smem, st, L->args.cluster,
                                                  L->tmA, L->tmB, L->tmC, L->args)
                                     : launch_pdl(gemm_f16_kernel<0>, dim3(L->grid), dim3(GEMM_THREADS), (size_t)L->smem, st, L->args.cluster,
                                                  L->tmA, L->tmB, L->tmC, L->args);
    if (le != cudaSuccess) (void)cudaGetLastError();
    DVC_REQUIRE(le == cudaSuccess, "gemm: cluster launch failed (%s): grid=%d cluster=%d smem=%d pair=%d block_n=%d stages=%d",
                cudaGetErrorString(le), L->grid, L->args.cluster, L->smem, L->args.two_cta, L->args.block_n, L->args.stages);
  } else {
    DVC_CHECK_CUDA(launch_pdl(gemm_f16_kernel<0>, dim3(L->grid), dim3(GEMM_THREADS), (size_t)L->smem, st, 1, L->tmA, L->tmB, L->tmC, L->args));
  }
  count_launch();
  DVC_CHECK_CUDA(cudaGetLastError());
  return DVC_OK;
}

// 3x3 / stride-2 / pad-1 dense convolution (the SSD extras conv14_2 .. conv17_2, prototxt:922-1146) as an implicit
// GEMM: out[B*Ho*Wo, N] = act(patches(in) * Wt[N, 9*C]^T + bias) with the patch rows gathered by the TMA unit.
int gemm_prepare_conv3x3s2(GemmLaunch* L, const void* in, int B, int H, int W, int C, const void* Wt, const float* bias, int N,
                           void* out, int ldc, int relu, int block_n) {
  DVC_REQUIRE(B > 0 && H >= 2 && W >= 2 && C > 0, "conv3x3s2: bad shape B=%d H=%d W=%d C=%d", B, H, W, C);
  const ConvA conv{B, H, W, C};
  const int Ho = (H - 1) / 2 + 1, Wo = (W - 1) / 2 + 1;
  return gemm_prepare_impl(L, in, Wt, bias, B * Ho * Wo, N, 9 * C, 9 * C, 0, out, ldc, relu, nullptr, block_n, &conv);
}

// Test hook for the im2col tensor map: one CTA fetches one [128 pixels x 64 channels] operand tile exactly as the GEMM
// producer does and writes it back un-swizzled, so the gather can be compared against a host-side unfold.  The wait is
// bounded (status 1 = the expected bytes never arrived).
__global__ void __launch_bounds__(128, 1)
im2col_tile_kernel(const __grid_constant__ CUtensorMap tmA, int c0, int w0, int h0, int n0, int kw, int kh, __half* __restrict__ out,
                   int* __restrict__ status) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* tile = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ int ok;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
    ok = 0;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(&bar, (uint32_t)A_STAGE_BYTES);
    tma_load_im2col_4d(tile, &tmA, &bar, c0, w0, h0, n0, kw, kh);
    const unsigned long long t0 = globaltimer_ns();
    while (!mbar_try_wait(&bar, 0)) {
      if (globaltimer_ns() - t0 > 200000000ull) break;
    }
    ok = mbar_try_wait(&bar, 0) ? 1 : 0;
  }
  __syncthreads();
  if (!ok) {
    if (threadIdx.x == 0) *status = 1;
    return;
  }
  const int r = threadIdx.x;                    // row r: 8 x 16-byte chunks, chunk j stored at position j ^ (r & 7)
  for (int j = 0; j < 8; ++j) {
    const uint4 v = *reinterpret_cast<const uint4*>(tile + r * 128 + ((j ^ (r & 7)) << 4));
    *reinterpret_cast<uint4*>(out + r * BK + j * 8) = v;
  }
  if (threadIdx.x == 0) *status = 0;
}

GemmLaunch* gemm_alloc() { return new GemmLaunch(); }
void gemm_free(GemmLaunch* L) { delete L; }

}  // namespace dvc

extern "C" int dvc_conv3x3s2_f16(const void* in_f16, int B, int H, int W, int C, const void* Wt, const float* bias, int N,
                                 void* out_f16, int ldc, int relu, int block_n, void* stream) {
  dvc::GemmLaunch L;
  int rc = dvc::gemm_prepare_conv3x3s2(&L, in_f16, B, H, W, C, Wt, bias, N, out_f16, ldc, relu, block_n);
  if (rc) return rc;
  return dvc::gemm_launch(&L, (cudaStream_t)stream);
}

extern "C" int dvc_debug_im2col_tile(const void* in_f16, int B, int H, int W, int C, int m0, int kb, void* out_tile_f16,
                                     int* status, void* stream) {
  using namespace dvc;
  DVC_REQUIRE(C % BK == 0 && kb >= 0 && kb < 9 * C / BK && m0 >= 0, "im2col tile: bad C=%d kb=%d m0=%d", C, kb, m0);
  CUtensorMap tm;
  int rc = make_tmap_im2col3x3s2_f16(&tm, in_f16, (uint64_t)B, (uint64_t)H, (uint64_t)W, (uint64_t)C, BK, BM);
  if (rc) return rc;
  const int Ho = (H - 1) / 2 + 1, Wo = (W - 1) / 2 + 1;
  const int n_img = m0 / (Ho * Wo), r = m0 - n_img * Ho * Wo, oh = r / Wo, ow = r - oh * Wo;
  const int tap = (kb * BK) / C, c0 = kb * BK - tap * C, kh = tap / 3, kw = tap - kh * 3;
  im2col_tile_kernel<<<1, 128, A_STAGE_BYTES + 1024, (cudaStream_t)stream>>>(tm, c0, 2 * ow - 1, 2 * oh - 1, n_img, kw, kh,
                                                                               (__half*)out_tile_f16, status);
  count_launch();
  DVC_CHECK_CUDA(cudaGetLastError());
  return DVC_OK;
}

extern "C" int dvc_gemm_f16(const void* A, const void* Wt, const float* bias, int M, int N, int K, int lda, int out_mode,
                            void* out_f16, int ldc, int relu, const dvc_head_out* head, int block_n, void* stream) {
  dvc::GemmLaunch L;
  int rc = dvc::gemm_prepare(&L, A, Wt, bias, M, N, K, lda, out_mode, out_f16, ldc, relu, head, block_n);
  if (rc) return rc;
  return dvc::gemm_launch(&L, (cudaStream_t)stream);
}
