// K6 PriorBox, K7+K8+K9 fused softmax / decode / top-k / NMS / keep-top-k / app filter with
// (optionally remote) result delivery, K10 annotation raster.
//
// Reference ops: cv2.dnn PriorBox / Softmax / DetectionOutput layers
// (prototxt:1239-1256, 1872-1880, 1890-1912), the Python post-filter loop of
// worker.py:253-269 and the cv2.rectangle / cv2.putText overlay of worker.py:268-276.
#include <mutex>

#include <atomic>

#include "common.cuh"

namespace dvc {

// ------------------------------------------------------------------------------ K6
__global__ void priorbox_kernel(float* boxes, float* variances, int prior_off, int gh, int gw, float step_h, float step_w,
                                float img_h, float img_w, const float* __restrict__ anchor_wh, int anchors, float offset,
                                float v0, float v1, float v2, float v3, int clip) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= gh * gw * anchors) return;
  const int a = i % anchors;
  const int cell = i / anchors;
  const int x = cell % gw, y = cell / gw;
  const float cx = __fmul_rn((float)x + offset, step_w);
  const float cy = __fmul_rn((float)y + offset, step_h);
  const float bw = __fmul_rn(anchor_wh[2 * a], 0.5f), bh = __fmul_rn(anchor_wh[2 * a + 1], 0.5f);
  float b0 = __fdiv_rn(__fsub_rn(cx, bw), img_w), b1 = __fdiv_rn(__fsub_rn(cy, bh), img_h);
  float b2 = __fdiv_rn(__fadd_rn(cx, bw), img_w), b3 = __fdiv_rn(__fadd_rn(cy, bh), img_h);
  if (clip) {
    b0 = fminf(fmaxf(b0, 0.f), 1.f); b1 = fminf(fmaxf(b1, 0.f), 1.f);
    b2 = fminf(fmaxf(b2, 0.f), 1.f); b3 = fminf(fmaxf(b3, 0.f), 1.f);
  }
  float* bp = boxes + (size_t)(prior_off + i) * 4;
  float* vp = variances + (size_t)(prior_off + i) * 4;
  bp[0] = b0; bp[1] = b1; bp[2] = b2; bp[3] = b3;
  vp[0] = v0; vp[1] = v1; vp[2] = v2; vp[3] = v3;
}

// ------------------------------------------------------------------------------ K7/K8/K9
// Three launches, all sized to the whole GPU instead of one CTA per frame:
//   K7  softmax_kernel        thread per prior: 21-way softmax, probabilities written class-major to scratch
//   K8  class_nms_kernel      CTA per (frame, class): threshold -> compaction -> rank-by-counting top_k ->
//                             CENTER_SIZE decode -> IoU bit-matrix -> greedy pass over the bit rows;
//                             survivors (sorted) go to scratch
//   K9  merge_deliver_kernel  CTA per frame: global keep_top_k by binary search across the per-class sorted
//                             lists, application filter, rows staged in shared memory and written - possibly
//                             to a peer GPU - followed by a release.sys flag (the detection gather)
constexpr int DET_MAXC = 32;      // classes
constexpr int DET_MAXK = 128;     // per-class top_k capacity
constexpr int NMS_THREADS = 128;
constexpr int MERGE_THREADS = 1024;   // one compare-exchange per thread and stage for up to 2048 survivors

struct DetArgs {
  const float *loc, *conf, *priors, *variances;
  int B;
  dvc_det_params p;
  const int* slot_index;
  float* det; int* det_count; int* boxes_i; int* box_count;
  uint32_t* flags; uint32_t flag_value;
  // K14 epoch gate (all optional): deliver only if this GPU's membership epoch word still equals the epoch the
  // batch was dispatched under; otherwise count the frame in *stale_count and leave the sinks untouched
  const uint32_t* epoch_now; const uint32_t* epoch_expect; uint32_t* stale_count;
  unsigned long long* t_done;      // optional [B]: %globaltimer when a frame's rows have been delivered
  // mailbox binding (csrc/mailbox.cu, optional): the flag value comes from the lane's control block, and the last CTA
  // of the batch publishes it in *done_word (a word on the client GPU) once every frame's rows are system-visible
  const uint32_t* flag_value_ptr; uint32_t* done_word; uint32_t* done_counter;
  // global scratch (per frame): prob [C][P], kept_n [C], kept_score [C][K], kept_idx [C][K], kept_box [C][K][4]
  float* prob; int* kept_n; float* kept_score; int* kept_idx; float* kept_box;
};

__device__ __forceinline__ bool before(float sa, int ia, int ca, float sb, int ib, int cb) {
  // strict total order: score desc, prior index asc, class asc
  if (sa != sb) return sa > sb;
  if (ia != ib) return ia < ib;
  return ca < cb;
}

__device__ __forceinline__ float box_area(const float* b) {
  if (b[2] < b[0] || b[3] < b[1]) return 0.f;
  return __fmul_rn(__fsub_rn(b[2], b[0]), __fsub_rn(b[3], b[1]));
}
__device__ __forceinline__ float box_iou(const float* a, const float* b) {
  float ib[4] = {fmaxf(a[0], b[0]), fmaxf(a[1], b[1]), fminf(a[2], b[2]), fminf(a[3], b[3])};
  const float inter = box_area(ib);
  if (!(inter > 0.f)) return 0.f;
  return __fdiv_rn(inter, __fsub_rn(__fadd_rn(box_area(a), box_area(b)), inter));
}

__global__ void __launch_bounds__(256) softmax_kernel(const DetArgs a) {
  pdl_launch_dependents();
  pdl_wait();
  const int C = a.p.num_classes, P = a.p.num_priors;
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  if (idx >= (long long)a.B * P) return;
  const int f = (int)(idx / P), i = (int)(idx - (long long)f * P);
  const float* x = a.conf + ((size_t)f * P + i) * C;
  float m = x[0];
  for (int c = 1; c < C; ++c) m = fmaxf(m, x[c]);
  float e[DET_MAXC];
  float s = 0.f;
#pragma unroll
  for (int c = 0; c < DET_MAXC; ++c) {
    if (c < C) {
      e[c] = expf(__fsub_rn(x[c], m));
      s = __fadd_rn(s, e[c]);
    }
  }
  float* out = a.prob + (size_t)f * C * P + i;
#pragma unroll
  for (int c = 0; c < DET_MAXC; ++c)
    if (c < C) out[(size_t)c * P] = __fdiv_rn(e[c], s);
}

__global__ void __launch_bounds__(NMS_THREADS) class_nms_kernel(const DetArgs a) {
  extern __shared__ __align__(16) uint8_t smem[];
  pdl_launch_dependents();
  pdl_wait();
  const int C = a.p.num_classes, P = a.p.num_priors, K = a.p.top_k;
  const int c = blockIdx.x, f = blockIdx.y;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int P_pad = (P + 3) & ~3;                               // keeps the float4 regions below 16-byte aligned
  float* cs = reinterpret_cast<float*>(smem);                   // [P] candidate scores
  int* ci = reinterpret_cast<int*>(cs + P_pad);                 // [P] candidate prior indices
  float* sel_score = reinterpret_cast<float*>(ci + P_pad);      // [DET_MAXK] sorted
  int* sel_idx = reinterpret_cast<int*>(sel_score + DET_MAXK);
  float* sel_box = reinterpret_cast<float*>(sel_idx + DET_MAXK);       // [DET_MAXK][4]
  uint32_t* mask = reinterpret_cast<uint32_t*>(sel_box + DET_MAXK * 4);  // [DET_MAXK][4]: later boxes this one suppresses
  uint32_t* keepw = mask + DET_MAXK * 4;                        // [4] survivor bits
  __shared__ int s_n;

  int* kept_n = a.kept_n + (size_t)f * C;
  if (c == a.p.background) {
    if (tid == 0) kept_n[c] = 0;
    return;
  }
  const float* prob = a.prob + ((size_t)f * C + c) * P;
  // ---- compaction of priors above the confidence threshold.  The order inside the candidate list does not matter: the
  // ranking below uses the strict total order (score desc, prior index asc), so an unordered shared-memory counter
  // replaces the per-chunk ballot / prefix / two-barrier scheme (15 chunks of 128 priors).
  if (tid == 0) s_n = 0;
  __syncthreads();
  for (int i = tid; i < P; i += NMS_THREADS) {
    const float sc = prob[i];
    if (sc > a.p.conf_thresh) {
      const int pos = atomicAdd(&s_n, 1);
      cs[pos] = sc;
      ci[pos] = i;
    }
  }
  __syncthreads();
  const int n_c = s_n;
  __syncthreads();                       // s_n is re-used for the survivor count further down
  // ---- rank by counting; ranks < top_k land in sorted position
  for (int e = tid; e < n_c; e += NMS_THREADS) {
    const float s = cs[e];
    const int id = ci[e];
    int rank = 0;
    for (int j = 0; j < n_c; ++j) rank += before(cs[j], ci[j], 0, s, id, 0) ? 1 : 0;
    if (rank < K) {
      sel_score[rank] = s;
      sel_idx[rank] = id;
    }
  }
  const int n = min(n_c, K);
  __syncthreads();
  // ---- decode the selected boxes (CENTER_SIZE, variance not encoded, no clip)
  const float* loc = a.loc + (size_t)f * P * 4;
  if (tid < n) {
    const int i = sel_idx[tid];
    const float4 pr = *reinterpret_cast<const float4*>(a.priors + (size_t)i * 4);
    const float4 va = *reinterpret_cast<const float4*>(a.variances + (size_t)i * 4);
    const float4 l = *reinterpret_cast<const float4*>(loc + (size_t)i * 4);
    const float pw = __fsub_rn(pr.z, pr.x), ph = __fsub_rn(pr.w, pr.y);
    const float pcx = __fmul_rn(__fadd_rn(pr.x, pr.z), 0.5f), pcy = __fmul_rn(__fadd_rn(pr.y, pr.w), 0.5f);
    const float cx = __fadd_rn(__fmul_rn(__fmul_rn(va.x, l.x), pw), pcx);
    const float cy = __fadd_rn(__fmul_rn(__fmul_rn(va.y, l.y), ph), pcy);
    const float w = __fmul_rn(expf(__fmul_rn(va.z, l.z)), pw);
    const float h = __fmul_rn(expf(__fmul_rn(va.w, l.w)), ph);
    float* bx = sel_box + tid * 4;
    bx[0] = __fsub_rn(cx, __fmul_rn(w, 0.5f));
    bx[1] = __fsub_rn(cy, __fmul_rn(h, 0.5f));
    bx[2] = __fadd_rn(cx, __fmul_rn(w, 0.5f));
    bx[3] = __fadd_rn(cy, __fmul_rn(h, 0.5f));
  }
  __syncthreads();
  // ---- IoU bit matrix: row i has bit j set when j > i overlaps i by more than the NMS threshold
  if (tid < DET_MAXK) {
    uint32_t m0 = 0, m1 = 0, m2 = 0, m3 = 0;
    if (tid < n) {
      const float* bi = sel_box + tid * 4;
      for (int j = tid + 1; j < n; ++j) {
        if (box_iou(bi, sel_box + j * 4) > a.p.nms_thresh) {
          const uint32_t bit = 1u << (j & 31);
          if (j < 32) m0 |= bit; else if (j < 64) m1 |= bit; else if (j < 96) m2 |= bit; else m3 |= bit;
        }
      }
    }
    mask[tid * 4 + 0] = m0; mask[tid * 4 + 1] = m1; mask[tid * 4 + 2] = m2; mask[tid * 4 + 3] = m3;
  }
  __syncthreads();
  // ---- greedy pass (one thread, pure bit operations): a kept box removes every later box it suppresses
  if (tid == 0) {
    uint32_t alive[4];
    for (int w = 0; w < 4; ++w) {
      const int lo = w * 32;
      alive[w] = n >= lo + 32 ? 0xffffffffu : (n > lo ? ((1u << (n - lo)) - 1u) : 0u);
    }
    int cnt = 0;
    for (int i = 0; i < n; ++i) {
      if ((alive[i >> 5] >> (i & 31)) & 1u) {
        ++cnt;
        alive[0] &= ~mask[i * 4 + 0]; alive[1] &= ~mask[i * 4 + 1];
        alive[2] &= ~mask[i * 4 + 2]; alive[3] &= ~mask[i * 4 + 3];
      }
    }
    for (int w = 0; w < 4; ++w) keepw[w] = alive[w];
    s_n = cnt;
  }
  __syncthreads();
  // ---- survivors, still in sorted order, to scratch
  if (tid < n && ((keepw[tid >> 5] >> (tid & 31)) & 1u)) {
    int pos = 0;
    for (int w = 0; w < (tid >> 5); ++w) pos += __popc(keepw[w]);
    pos += __popc(keepw[tid >> 5] & ((1u << (tid & 31)) - 1u));
    const size_t o = ((size_t)f * C + c) * DET_MAXK + pos;
    a.kept_score[o] = sel_score[tid];
    a.kept_idx[o] = sel_idx[tid];
    *reinterpret_cast<float4*>(a.kept_box + o * 4) = *reinterpret_cast<const float4*>(sel_box + tid * 4);
  }
  if (tid == 0) kept_n[c] = s_n;
}

__global__ void __launch_bounds__(MERGE_THREADS) merge_deliver_kernel(const DetArgs a) {
  __shared__ int s_n[DET_MAXC];
  __shared__ int s_off[DET_MAXC];
  __shared__ int s_tot;
  __shared__ int s_stale;
  __shared__ float s_rows[DVC_MAX_DET * DVC_DET_ROW];
  extern __shared__ __align__(16) uint8_t msm[];          // sort keys of the survivors
  pdl_launch_dependents();
  pdl_wait();
  const int C = a.p.num_classes;
  if (threadIdx.x == 0) {
    // a dispatch made under an older membership epoch must not land in rings that may have been re-assigned
    int stale = 0;
    if (a.epoch_now && a.epoch_expect) {
      const uint32_t expect = ld_acquire_sys_u32(a.epoch_expect);
      stale = ld_acquire_sys_u32(a.epoch_now) != expect;
      // expect == ~0: the lane's wait saw no dispatch at all (csrc/mailbox.cu) - nothing to deliver, nothing to count
      if (stale && a.stale_count && expect != 0xFFFFFFFFu) atomicAdd(a.stale_count, 1u);
    }
    s_stale = stale;
  }
  __syncthreads();
  if (s_stale) {
    // mailbox: a batch with a rejected frame never publishes its done word (the client times out and re-dispatches);
    // the arrival is still counted so the counter is back at zero for the lane's next batch
    if (threadIdx.x == 0 && a.done_counter) {
      const uint32_t prev = atomicAdd(a.done_counter, 0x10001u);
      if ((prev & 0xFFFFu) == (uint32_t)a.B - 1u) atomicExch(a.done_counter, 0u);
    }
    return;
  }
  unsigned long long* keys = reinterpret_cast<unsigned long long*>(msm);   // survivors of all classes, one sort key each
  const int f = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int slot = a.slot_index ? a.slot_index[f] : f;
  const int* kept_n = a.kept_n + (size_t)f * C;
  const float* gks = a.kept_score + (size_t)f * C * DET_MAXK;
  const int* gki = a.kept_idx + (size_t)f * C * DET_MAXK;
  const float* kb = a.kept_box + (size_t)f * C * DET_MAXK * 4;
  if (tid < DET_MAXC) s_n[tid] = tid < C ? kept_n[tid] : 0;
  for (int e = tid; e < DVC_MAX_DET * DVC_DET_ROW; e += MERGE_THREADS) s_rows[e] = 0.f;
  __syncthreads();
  if (tid == 0) {
    int t = 0;
    for (int c = 0; c < C; ++c) { s_off[c] = t; t += s_n[c]; }
    s_tot = t;
  }
  __syncthreads();
  // Global keep-top-k = one bitonic sort of all survivors in shared memory.  Key (ascending = output order):
  // [63:32] ~orderable(score) (score desc), [24:12] prior index asc, [11:7] class asc, [6:0] slot in the class list.
  int ns = 2;
  while (ns < s_tot) ns <<= 1;
  for (int e = tid; e < C * DET_MAXK; e += MERGE_THREADS) {
    const int c = e / DET_MAXK, r = e - c * DET_MAXK;
    if (r < s_n[c]) {
      uint32_t u = __float_as_uint(gks[e]);
      u ^= (u >> 31) ? 0xFFFFFFFFu : 0x80000000u;                  // monotonic in the float order
      keys[s_off[c] + r] = ((unsigned long long)(~u) << 32) | ((unsigned long long)(uint32_t)gki[e] << 12) |
                           ((unsigned long long)c << 7) | (unsigned long long)r;
    }
  }
  for (int i = s_tot + tid; i < ns; i += MERGE_THREADS) keys[i] = ~0ull;
  __syncthreads();
  for (int k = 2; k <= ns; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < ns; i += MERGE_THREADS) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const unsigned long long x = keys[i], y = keys[ixj];
          if ((x > y) == ((i & k) == 0)) { keys[i] = y; keys[ixj] = x; }
        }
      }
      __syncthreads();
    }
  }
  const int n_out = min(s_tot, min(a.p.keep_top_k, DVC_MAX_DET));
  if (tid < n_out) {
    const unsigned long long key = keys[tid];
    const int r = (int)(key & 127u), c = (int)((key >> 7) & 31u);
    uint32_t u = ~(uint32_t)(key >> 32);
    u = (u & 0x80000000u) ? (u ^ 0x80000000u) : ~u;
    float* row = s_rows + tid * DVC_DET_ROW;
    const float* bx = kb + (size_t)(c * DET_MAXK + r) * 4;
    row[0] = 0.f; row[1] = (float)c; row[2] = __uint_as_float(u);
    row[3] = bx[0]; row[4] = bx[1]; row[5] = bx[2]; row[6] = bx[3];
  }
  __syncthreads();
  // rows (zero padded) go out with coalesced stores; `det` may be peer memory (the gather)
  float* det = a.det + (size_t)slot * DVC_MAX_DET * DVC_DET_ROW;
  for (int e = tid; e < DVC_MAX_DET * DVC_DET_ROW; e += MERGE_THREADS) det[e] = s_rows[e];
  // application filter (conf > 0.2, class in CONSIDER) -> integer boxes, in row order
  if (warp == 0) {
    int base = 0;
    int* bi = a.boxes_i + (size_t)slot * DVC_MAX_DET * 4;
    for (int r0 = 0; r0 < n_out; r0 += 32) {
      const int r = r0 + lane;
      bool ok = false;
      const float* row = s_rows + min(r, DVC_MAX_DET - 1) * DVC_DET_ROW;
      if (r < n_out) {
        const int c = (int)row[1];
        ok = row[2] > a.p.app_conf && ((a.p.consider_mask >> c) & 1u);
      }
      const unsigned m = __ballot_sync(0xffffffff, ok);
      if (ok) {
        const int pos = base + __popc(m & ((1u << lane) - 1));
        const double dw = (double)a.p.out_w, dh = (double)a.p.out_h;
        const double v[4] = {(double)row[3] * dw, (double)row[4] * dh, (double)row[5] * dw, (double)row[6] * dh};
        for (int k = 0; k < 4; ++k) bi[pos * 4 + k] = (int)fmin(fmax(v[k], -2147483000.0), 2147483000.0);
      }
      base += __popc(m);
    }
    if (lane == 0) {
      a.box_count[slot] = base;
      a.det_count[slot] = n_out;
    }
  }
  // delivery: make the rows visible system-wide, then publish the slot flag
  __threadfence_system();
  __syncthreads();
  if (tid == 0) {
    const uint32_t fv = a.flag_value_ptr ? *reinterpret_cast<const volatile uint32_t*>(a.flag_value_ptr) : a.flag_value;
    if (a.flags) st_release_sys_u32(a.flags + slot, fv);
    if (a.done_counter) {
      // batch-level completion: every CTA's rows were fenced to system scope above; the device-scope atomic chains
      // them (cumulativity) to the last arriver, whose release.sys store is what the client's ld.acquire.sys observes
      const uint32_t prev = atomicAdd(a.done_counter, 1u);
      if ((prev & 0xFFFFu) == (uint32_t)a.B - 1u) {
        atomicExch(a.done_counter, 0u);
        __threadfence_system();
        if ((prev >> 16) == 0u && a.done_word) st_release_sys_u32(a.done_word, fv);
      }
    }
    if (a.t_done) a.t_done[f] = globaltimer_ns();
  }
}

// one 8-byte sort key per per-class survivor slot, padded to a power of two for the bitonic network
static int merge_smem_bytes(int C) {
  int n = 2;
  while (n < C * DET_MAXK) n <<= 1;
  return n * 8;
}

struct DetLaunch {
  DetArgs args;
  int smem_nms;
};

static size_t det_per_frame_bytes(int P, int C) {
  // prob [C][P] f32 + kept_n [C] + kept_score/idx [C][K] + kept_box [C][K][4], each region 16-byte aligned
  size_t b = 0;
  b += (((size_t)C * P * 4) + 15) & ~size_t(15);
  b += (((size_t)C * 4) + 15) & ~size_t(15);
  b += (size_t)C * DET_MAXK * 4 * 2;
  b += (size_t)C * DET_MAXK * 16;
  return b;
}
size_t det_scratch_bytes(int B, int P, int C) { return (size_t)B * det_per_frame_bytes(P, C); }

int det_prepare(DetLaunch* L, const float* loc, const float* conf, const float* priors, const float* variances, int B,
                const dvc_det_params* p, const int* slot_index, float* det, int* det_count, int* boxes_i, int* box_count,
                uint32_t* flags, uint32_t flag_value, void* scratch, size_t scratch_bytes) {
  DVC_REQUIRE(p && B > 0, "detection: bad arguments");
  DVC_REQUIRE(p->num_classes >= 2 && p->num_classes <= DET_MAXC, "detection: num_classes=%d unsupported (max %d)", p->num_classes, DET_MAXC);
  DVC_REQUIRE(p->top_k > 0 && p->top_k <= DET_MAXK, "detection: top_k=%d unsupported (max %d)", p->top_k, DET_MAXK);
  DVC_REQUIRE(p->keep_top_k > 0, "detection: keep_top_k must be positive");
  DVC_REQUIRE(p->num_priors > 0 && p->num_priors <= 8192, "detection: num_priors=%d unsupported (max 8192)", p->num_priors);
  const size_t need = det_scratch_bytes(B, p->num_priors, p->num_classes);
  DVC_REQUIRE(scratch != nullptr && scratch_bytes >= need, "detection: scratch %zu bytes < required %zu", scratch_bytes, need);
  DVC_REQUIRE((reinterpret_cast<uintptr_t>(scratch) & 15) == 0, "detection: scratch must be 16-byte aligned");
  DVC_REQUIRE((reinterpret_cast<uintptr_t>(loc) & 15) == 0 && (reinterpret_cast<uintptr_t>(priors) & 15) == 0 &&
              (reinterpret_cast<uintptr_t>(variances) & 15) == 0, "detection: loc/priors/variances must be 16-byte aligned");
  DetArgs& a = L->args;
  a.loc = loc; a.conf = conf; a.priors = priors; a.variances = variances; a.B = B; a.p = *p; a.slot_index = slot_index;
  a.det = det; a.det_count = det_count; a.boxes_i = boxes_i; a.box_count = box_count; a.flags = flags; a.flag_value = flag_value;
  a.epoch_now = nullptr; a.epoch_expect = nullptr; a.stale_count = nullptr;      // gate off until det_set_epoch_gate
  a.t_done = nullptr;
  a.flag_value_ptr = nullptr; a.done_word = nullptr; a.done_counter = nullptr;
  const int C = p->num_classes, P = p->num_priors;
  uint8_t* s = reinterpret_cast<uint8_t*>(scratch);
  a.prob = reinterpret_cast<float*>(s);
  s += (size_t)B * ((((size_t)C * P * 4) + 15) & ~size_t(15));
  a.kept_n = reinterpret_cast<int*>(s);
  s += (size_t)B * ((((size_t)C * 4) + 15) & ~size_t(15));
  a.kept_score = reinterpret_cast<float*>(s);
  s += (size_t)B * C * DET_MAXK * 4;
  a.kept_idx = reinterpret_cast<int*>(s);
  s += (size_t)B * C * DET_MAXK * 4;
  a.kept_box = reinterpret_cast<float*>(s);
  L->smem_nms = ((P + 3) & ~3) * 8 + DET_MAXK * (4 + 4 + 16 + 16) + 16 + 64;
  return DVC_OK;
}

int det_launch(const DetLaunch* L, cudaStream_t st) {
  static std::atomic<bool> attr_set[64];   // zero-initialised; engine threads of different GPUs may race here
  int dev = 0;
  DVC_CHECK_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63].load(std::memory_order_acquire)) {
    DVC_CHECK_CUDA(cudaFuncSetAttribute(class_nms_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    attr_set[dev & 63].store(true, std::memory_order_release);
  }
  const DetArgs& a = L->args;
  const long long n = (long long)a.B * a.p.num_priors;
  DVC_CHECK_CUDA(launch_pdl(softmax_kernel, dim3((unsigned)ceil_div64(n, 256)), dim3(256), 0, st, 1, a));
  DVC_CHECK_CUDA(launch_pdl(class_nms_kernel, dim3(a.p.num_classes, a.B), dim3(NMS_THREADS), (size_t)L->smem_nms, st, 1, a));
  DVC_CHECK_CUDA(launch_pdl(merge_deliver_kernel, dim3(a.B), dim3(MERGE_THREADS), (size_t)merge_smem_bytes(a.p.num_classes), st, 1, a));
  count_launch(3);
  DVC_CHECK_CUDA(cudaGetLastError());
  return DVC_OK;
}

DetLaunch* det_alloc() { return new DetLaunch(); }
void det_free(DetLaunch* L) { delete L; }
void det_set_flag_value(DetLaunch* L, uint32_t v) { L->args.flag_value = v; }
void det_set_timestamps(DetLaunch* L, unsigned long long* t_done) { L->args.t_done = t_done; }
void det_set_epoch_gate(DetLaunch* L, const uint32_t* now, const uint32_t* expect, uint32_t* stale_count) {
  L->args.epoch_now = now; L->args.epoch_expect = expect; L->args.stale_count = stale_count;
}
void det_set_mailbox(DetLaunch* L, const uint32_t* flag_value_ptr, uint32_t* done_word, uint32_t* done_counter) {
  L->args.flag_value_ptr = flag_value_ptr; L->args.done_word = done_word; L->args.done_counter = done_counter;
}
void det_set_sinks(DetLaunch* L, float* det, int* det_count, int* boxes_i, int* box_count, uint32_t* flags, const int* slot_index) {
  L->args.det = det; L->args.det_count = det_count; L->args.boxes_i = boxes_i; L->args.box_count = box_count;
  L->args.flags = flags; L->args.slot_index = slot_index;
}

// ------------------------------------------------------------------------------ K10
struct AnnArgs {
  uint8_t* frames; long long stride; const int* frame_index; int B, h, w;
  const int *boxes_i, *box_count, *slot_index;
  const uint8_t* id_mask; int id_x, id_y, id_w, id_h;
  const uint8_t* label_masks; int n_labels, lab_x, lab_y, lab_w, lab_h;
};

// cv2.rectangle(..., thickness=2): a 3-pixel frame [x0-1,x1+1] x [y0-1,y1+1] minus the interior
// [x0+2,x1-2] x [y0+2,y1-2] minus the four outermost corner pixels (verified against cv2 4.13).
__device__ __forceinline__ bool on_rect(int x, int y, int x0, int y0, int x1, int y1) {
  if (x0 > x1) { int t = x0; x0 = x1; x1 = t; }
  if (y0 > y1) { int t = y0; y0 = y1; y1 = t; }
  if (x < x0 - 1 || x > x1 + 1 || y < y0 - 1 || y > y1 + 1) return false;
  if (x >= x0 + 2 && x <= x1 - 2 && y >= y0 + 2 && y <= y1 - 2) return false;
  if ((x == x0 - 1 || x == x1 + 1) && (y == y0 - 1 || y == y1 + 1)) return false;
  return true;
}

constexpr int ANN_THREADS = 512;

// One CTA per frame.  Work is proportional to what is drawn (box perimeters + text masks), not to
// the frame area: for each box the four 3-pixel bands of its outline are rasterised with the exact
// rule above; the two text overlays follow after a barrier so they win over box pixels, as in the
// reference's draw order (worker.py:268-276).
__global__ void __launch_bounds__(ANN_THREADS) annotate_kernel(const AnnArgs a) {
  __shared__ int s_box[DVC_MAX_DET * 4];
  __shared__ int s_n;
  pdl_launch_dependents();
  pdl_wait();
  const int b = blockIdx.x;
  const int tid = threadIdx.x;
  const int slot = a.slot_index ? a.slot_index[b] : b;
  const int fslot = a.frame_index ? a.frame_index[b] : b;
  if (tid == 0) s_n = min(max(a.box_count[slot], 0), DVC_MAX_DET);
  __syncthreads();
  const int n = s_n;
  for (int i = tid; i < n * 4; i += ANN_THREADS) s_box[i] = a.boxes_i[(size_t)slot * DVC_MAX_DET * 4 + i];
  __syncthreads();
  uint8_t* frame = a.frames + (size_t)fslot * a.stride;
  for (int k = 0; k < n; ++k) {
    int x0 = s_box[4 * k], y0 = s_box[4 * k + 1], x1 = s_box[4 * k + 2], y1 = s_box[4 * k + 3];
    if (x0 > x1) { int t = x0; x0 = x1; x1 = t; }
    if (y0 > y1) { int t = y0; y0 = y1; y1 = t; }
    // bands: 0 top, 1 bottom, 2 left, 3 right (each 3 pixels thick, clipped to the frame)
#pragma unroll
    for (int band = 0; band < 4; ++band) {
      long long bx0, bx1, by0, by1;
      if (band < 2) {
        bx0 = (long long)x0 - 1; bx1 = (long long)x1 + 1;
        by0 = band == 0 ? (long long)y0 - 1 : (long long)y1 - 1;
        by1 = by0 + 2;
      } else {
        by0 = (long long)y0 - 1; by1 = (long long)y1 + 1;
        bx0 = band == 2 ? (long long)x0 - 1 : (long long)x1 - 1;
        bx1 = bx0 + 2;
      }
      const int cx0 = (int)max(bx0, 0LL), cx1 = (int)min(bx1, (long long)a.w - 1);
      const int cy0 = (int)max(by0, 0LL), cy1 = (int)min(by1, (long long)a.h - 1);
      if (cx0 > cx1 || cy0 > cy1) continue;
      const int bw = cx1 - cx0 + 1, cnt = bw * (cy1 - cy0 + 1);
      for (int i = tid; i < cnt; i += ANN_THREADS) {
        const int y = cy0 + i / bw, x = cx0 + i % bw;
        if (on_rect(x, y, x0, y0, x1, y1)) {
          uint8_t* p = frame + ((size_t)y * a.w + x) * 3;
          p[0] = 255; p[1] = 0; p[2] = 0;
        }
      }
    }
  }
  __syncthreads();
  if (a.id_mask) {
    for (int i = tid; i < a.id_w * a.id_h; i += ANN_THREADS) {
      const int y = a.id_y + i / a.id_w, x = a.id_x + i % a.id_w;
      if (a.id_mask[i] && x >= 0 && x < a.w && y >= 0 && y < a.h) {
        uint8_t* p = frame + ((size_t)y * a.w + x) * 3;
        p[0] = 0; p[1] = 0; p[2] = 255;
      }
    }
  }
  __syncthreads();
  if (a.label_masks) {
    const uint8_t* lab = a.label_masks + (size_t)min(n, a.n_labels - 1) * a.lab_w * a.lab_h;
    for (int i = tid; i < a.lab_w * a.lab_h; i += ANN_THREADS) {
      const int y = a.lab_y + i / a.lab_w, x = a.lab_x + i % a.lab_w;
      if (lab[i] && x >= 0 && x < a.w && y >= 0 && y < a.h) {
        uint8_t* p = frame + ((size_t)y * a.w + x) * 3;
        p[0] = 0; p[1] = 255; p[2] = 0;
      }
    }
  }
}

}  // namespace dvc

extern "C" int dvc_priorbox(float* boxes, float* variances, int prior_off, int grid_h, int grid_w, int img_h, int img_w,
                            const float* anchor_wh, int anchors, float offset, const float* h_var4, int clip, void* stream) {
  using namespace dvc;
  DVC_REQUIRE(boxes && variances && anchor_wh && h_var4 && grid_h > 0 && grid_w > 0 && anchors > 0, "priorbox: bad arguments");
  const int n = grid_h * grid_w * anchors;
  priorbox_kernel<<<ceil_div(n, 128), 128, 0, (cudaStream_t)stream>>>(
      boxes, variances, prior_off, grid_h, grid_w, (float)img_h / (float)grid_h, (float)img_w / (float)grid_w, (float)img_h,
      (float)img_w, anchor_wh, anchors, offset, h_var4[0], h_var4[1], h_var4[2], h_var4[3], clip);
  count_launch();
  DVC_CHECK_CUDA(cudaGetLastError());
  return DVC_OK;
}

extern "C" size_t dvc_detection_scratch_bytes(int B, int num_priors, int num_classes) {
  return dvc::det_scratch_bytes(B, num_priors, num_classes);
}

extern "C" int dvc_detection_output(const float* loc, const float* conf, const float* priors, const float* variances, int B,
                                    const dvc_det_params* p, const int* slot_index, float* det, int* det_count,
                                    int* boxes_i, int* box_count, uint32_t* flags, uint32_t flag_value, void* scratch,
                                    size_t scratch_bytes, void* stream) {
  dvc::DetLaunch L;
  int rc = dvc::det_prepare(&L, loc, conf, priors, variances, B, p, slot_index, det, det_count, boxes_i, box_count, flags,
                            flag_value, scratch, scratch_bytes);
  if (rc) return rc;
  return dvc::det_launch(&L, (cudaStream_t)stream);
}

// ---- allocation-free variant of dvc_detection_output -------------------------------------
// The detection scratch (class-major probabilities and the per-class survivor lists) lives in this
// file-scope device array instead of caller memory.  Every
// GPU that loads the library carries these 4 MiB until the process exits.  In host code the
// symbol below is only a stand-in: its device address is looked up once per device with
// cudaGetSymbolAddress (with that device current) and that address is what the kernel gets.
static __device__ __align__(256) unsigned char g_det_static_scratch[DVC_DET_STATIC_SCRATCH_BYTES];

static int det_static_scratch(void** out) {
  static std::mutex mu;
  static void* ptrs[64] = {};
  static bool ready[64] = {};
  int dev = 0;
  DVC_CHECK_CUDA(cudaGetDevice(&dev));
  DVC_REQUIRE(dev >= 0 && dev < 64, "detection (static scratch): device index %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!ready[dev]) {
    void* p = nullptr;
    DVC_CHECK_CUDA(cudaGetSymbolAddress(&p, g_det_static_scratch));
    DVC_REQUIRE(p != nullptr, "detection (static scratch): cudaGetSymbolAddress returned a null address");
    ptrs[dev] = p;
    ready[dev] = true;
  }
  *out = ptrs[dev];
  return DVC_OK;
}

extern "C" int dvc_detection_output_static(const float* loc, const float* conf, const float* priors, const float* variances,
                                           int B, const dvc_det_params* p, const int* slot_index, float* det, int* det_count,
                                           int* boxes_i, int* box_count, uint32_t* flags, uint32_t flag_value, void* stream) {
  using namespace dvc;
  DVC_REQUIRE(p && B > 0, "detection (static scratch): bad arguments");
  const size_t need = det_scratch_bytes(B, p->num_priors, p->num_classes);
  DVC_REQUIRE(need <= (size_t)DVC_DET_STATIC_SCRATCH_BYTES,
              "detection (static scratch): B=%d x %d priors x %d classes needs %zu bytes of scratch, the built-in "
              "pool holds %zu; split the batch or use dvc_detection_output with your own scratch",
              B, p->num_priors, p->num_classes, need, (size_t)DVC_DET_STATIC_SCRATCH_BYTES);
  void* scratch = nullptr;
  int rc = det_static_scratch(&scratch);
  if (rc) return rc;
  DetLaunch L;
  rc = det_prepare(&L, loc, conf, priors, variances, B, p, slot_index, det, det_count, boxes_i, box_count, flags, flag_value,
                   scratch, (size_t)DVC_DET_STATIC_SCRATCH_BYTES);
  if (rc) return rc;
  return det_launch(&L, (cudaStream_t)stream);
}

extern "C" int dvc_annotate(uint8_t* frames, long long frame_stride, const int* frame_index, int B, int h, int w,
                            const int* boxes_i, const int* box_count, const int* slot_index, const uint8_t* id_mask, int id_x,
                            int id_y, int id_w, int id_h, const uint8_t* label_masks, int n_labels, int lab_x, int lab_y,
                            int lab_w, int lab_h, void* stream) {
  using namespace dvc;
  DVC_REQUIRE(frames && boxes_i && box_count && B > 0 && h > 0 && w > 0, "annotate: bad arguments");
  DVC_REQUIRE(!label_masks || n_labels > 0, "annotate: n_labels must be positive");
  AnnArgs a{frames, frame_stride, frame_index, B, h, w, boxes_i, box_count, slot_index, id_mask, id_x, id_y, id_w, id_h,
            label_masks, n_labels, lab_x, lab_y, lab_w, lab_h};
  DVC_CHECK_CUDA(launch_pdl(annotate_kernel, dim3(B), dim3(ANN_THREADS), 0, (cudaStream_t)stream, 1, a));
  count_launch();
  DVC_CHECK_CUDA(cudaGetLastError());
  return DVC_OK;
}
