// K6 PriorBox, K7+K8+K9 fused softmax / decode / top-k / NMS / keep-top-k / app filter with
// (optionally remote) result delivery, K10 annotation raster.
//
// Reference ops: cv2.dnn PriorBox / Softmax / DetectionOutput layers
// (prototxt:1239-1256, 1872-1880, 1890-1912), the Python post-filter loop of
// worker.py:253-269 and the cv2.rectangle / cv2.putText overlay of worker.py:268-276.
#include <mutex>

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
constexpr int DET_THREADS = 672;    // 21 warps: one per class for the NMS phase
constexpr int DET_MAXC = 32;      // classes
constexpr int DET_MAXK = 128;     // per-class top_k capacity

struct DetArgs {
  const float *loc, *conf, *priors, *variances;
  int B;
  dvc_det_params p;
  const int* slot_index;
  float* det; int* det_count; int* boxes_i; int* box_count;
  uint32_t* flags; uint32_t flag_value;
  float* pool_score; int* pool_idx; long long pool_stride;   // global scratch, per frame
};

__device__ __forceinline__ void softmax_row(const float* __restrict__ x, int C, float* prob) {
  float m = x[0];
  for (int c = 1; c < C; ++c) m = fmaxf(m, x[c]);
  float s = 0.f;
  for (int c = 0; c < C; ++c) {
    prob[c] = expf(__fsub_rn(x[c], m));
    s = __fadd_rn(s, prob[c]);
  }
  for (int c = 0; c < C; ++c) prob[c] = __fdiv_rn(prob[c], s);
}

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

__global__ void __launch_bounds__(DET_THREADS) detection_kernel(const DetArgs a) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int C = a.p.num_classes, P = a.p.num_priors, K = a.p.top_k;
  int* s_cnt = reinterpret_cast<int*>(smem);           // [DET_MAXC] candidates per class
  int* s_off = s_cnt + DET_MAXC;                       // [DET_MAXC+1]
  int* s_fill = s_off + DET_MAXC + 1;                  // [DET_MAXC]
  int* s_nsel = s_fill + DET_MAXC;                     // [DET_MAXC]
  int* s_misc = s_nsel + DET_MAXC;                     // [0]=n_kept  [1]=n_boxes
  float* sel_score = reinterpret_cast<float*>(s_misc + 7);      // [C][DET_MAXK]
  int* sel_idx = reinterpret_cast<int*>(sel_score + C * DET_MAXK);
  float* sel_box = reinterpret_cast<float*>(sel_idx + C * DET_MAXK);   // [C][DET_MAXK][4]
  uint8_t* sel_keep = reinterpret_cast<uint8_t*>(sel_box + C * DET_MAXK * 4);
  int* kept = reinterpret_cast<int*>(sel_keep + ((C * DET_MAXK + 15) & ~15));   // [C*DET_MAXK] (c<<16 | r)
  float* s_rows = reinterpret_cast<float*>(kept + C * DET_MAXK);                // [100][7] output staging

  const int f = blockIdx.x;
  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const float* loc = a.loc + (size_t)f * P * 4;
  const float* conf = a.conf + (size_t)f * P * C;
  float* pool_score = a.pool_score + (size_t)f * a.pool_stride;
  int* pool_idx = a.pool_idx + (size_t)f * a.pool_stride;
  const int slot = a.slot_index ? a.slot_index[f] : f;

  if (tid < DET_MAXC) { s_cnt[tid] = 0; s_fill[tid] = 0; s_nsel[tid] = 0; }
  if (tid < 2) s_misc[tid] = 0;
  __syncthreads();

  // ---- phase 1: softmax, count candidates per class
  float prob[DET_MAXC];
  for (int i = tid; i < P; i += DET_THREADS) {
    softmax_row(conf + (size_t)i * C, C, prob);
    for (int c = 0; c < C; ++c)
      if (c != a.p.background && prob[c] > a.p.conf_thresh) atomicAdd(&s_cnt[c], 1);
  }
  __syncthreads();
  if (tid == 0) {
    int acc = 0;
    for (int c = 0; c < C; ++c) { s_off[c] = acc; acc += s_cnt[c]; }
    s_off[C] = acc;
  }
  __syncthreads();
  // ---- phase 2: fill the per-class candidate pool (global scratch: worst case P*(C-1) entries)
  for (int i = tid; i < P; i += DET_THREADS) {
    softmax_row(conf + (size_t)i * C, C, prob);
    for (int c = 0; c < C; ++c)
      if (c != a.p.background && prob[c] > a.p.conf_thresh) {
        const int pos = s_off[c] + atomicAdd(&s_fill[c], 1);
        pool_score[pos] = prob[c];
        pool_idx[pos] = i;
      }
  }
  __syncthreads();
  // ---- phase 3: per-class rank by counting; ranks < top_k are selected in sorted position
  const int total = s_off[C];
  for (int e = tid; e < total; e += DET_THREADS) {
    int c = 0;
    while (e >= s_off[c + 1]) ++c;
    const float s = pool_score[e];
    const int id = pool_idx[e];
    int rank = 0;
    for (int j = s_off[c]; j < s_off[c + 1]; ++j) rank += before(pool_score[j], pool_idx[j], 0, s, id, 0) ? 1 : 0;
    if (rank < K) {
      sel_score[c * DET_MAXK + rank] = s;
      sel_idx[c * DET_MAXK + rank] = id;
    }
  }
  if (tid < C) s_nsel[tid] = min(s_cnt[tid], K);
  __syncthreads();
  // ---- phase 4: decode the selected boxes (CENTER_SIZE, variance not encoded, no clip)
  for (int e = tid; e < C * K; e += DET_THREADS) {
    const int c = e / K, r = e - c * K;
    if (r >= s_nsel[c]) continue;
    const int i = sel_idx[c * DET_MAXK + r];
    const float4 pr = *reinterpret_cast<const float4*>(a.priors + (size_t)i * 4);
    const float4 va = *reinterpret_cast<const float4*>(a.variances + (size_t)i * 4);
    const float4 l = *reinterpret_cast<const float4*>(loc + (size_t)i * 4);
    const float pw = __fsub_rn(pr.z, pr.x), ph = __fsub_rn(pr.w, pr.y);
    const float pcx = __fmul_rn(__fadd_rn(pr.x, pr.z), 0.5f), pcy = __fmul_rn(__fadd_rn(pr.y, pr.w), 0.5f);
    const float cx = __fadd_rn(__fmul_rn(__fmul_rn(va.x, l.x), pw), pcx);
    const float cy = __fadd_rn(__fmul_rn(__fmul_rn(va.y, l.y), ph), pcy);
    const float w = __fmul_rn(expf(__fmul_rn(va.z, l.z)), pw);
    const float h = __fmul_rn(expf(__fmul_rn(va.w, l.w)), ph);
    float* bx = sel_box + (size_t)(c * DET_MAXK + r) * 4;
    bx[0] = __fsub_rn(cx, __fmul_rn(w, 0.5f));
    bx[1] = __fsub_rn(cy, __fmul_rn(h, 0.5f));
    bx[2] = __fadd_rn(cx, __fmul_rn(w, 0.5f));
    bx[3] = __fadd_rn(cy, __fmul_rn(h, 0.5f));
    sel_keep[c * DET_MAXK + r] = 1;
  }
  __syncthreads();
  // ---- phase 5: greedy NMS, one warp per class; a kept box suppresses every later box it overlaps
  for (int c = warp; c < C; c += DET_THREADS / 32) {
    const int n = s_nsel[c];
    for (int i = 0; i < n; ++i) {
      __syncwarp();
      if (!sel_keep[c * DET_MAXK + i]) continue;   // warp-uniform (shared memory value)
      const float* bi = sel_box + (size_t)(c * DET_MAXK + i) * 4;
      for (int r = i + 1 + lane; r < n; r += 32) {
        if (sel_keep[c * DET_MAXK + r] && box_iou(bi, sel_box + (size_t)(c * DET_MAXK + r) * 4) > a.p.nms_thresh)
          sel_keep[c * DET_MAXK + r] = 0;
      }
    }
  }
  __syncthreads();
  // ---- phase 6: global keep_top_k.  Survivors of one class are already in rank order, so the global
  // rank of an entry is the sum over classes of "survivors of that class that sort before it": a
  // binary search per class over its sorted list plus an exclusive prefix count of the keep flags.
  int* kprefix = kept;                                   // [C][DET_MAXK] survivors among r' < r
  for (int c = warp; c < C; c += DET_THREADS / 32) {
    const int n = s_nsel[c];
    int cnt = 0;
    int flags4[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int r = lane * 4 + k;
      flags4[k] = (r < n && sel_keep[c * DET_MAXK + r]) ? 1 : 0;
      cnt += flags4[k];
    }
    int incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int up = __shfl_up_sync(0xffffffff, incl, d);
      if (lane >= d) incl += up;
    }
    int run = incl - cnt;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      kprefix[c * DET_MAXK + lane * 4 + k] = run;
      run += flags4[k];
    }
    if (lane == 31) s_fill[c] = incl;                    // survivors of class c (s_fill is free by now)
  }
  __syncthreads();
  if (tid == 0) {
    int tot = 0;
    for (int c = 0; c < C; ++c) tot += s_fill[c];
    s_misc[0] = tot;
  }
  for (int e = tid; e < DVC_MAX_DET * DVC_DET_ROW; e += DET_THREADS) s_rows[e] = 0.f;
  __syncthreads();
  const int n_kept = s_misc[0];
  const int n_out = min(n_kept, min(a.p.keep_top_k, DVC_MAX_DET));
  for (int e = tid; e < C * K; e += DET_THREADS) {
    const int c = e / K, r = e - c * K;
    if (r >= s_nsel[c] || !sel_keep[c * DET_MAXK + r]) continue;
    const float s = sel_score[c * DET_MAXK + r];
    const int id = sel_idx[c * DET_MAXK + r];
    int rank = 0;
    for (int c2 = 0; c2 < C; ++c2) {
      const int n2 = s_nsel[c2];
      if (n2 == 0) continue;
      int lo = 0, hi = n2;                               // first index whose entry does NOT sort before (s,id,c)
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (before(sel_score[c2 * DET_MAXK + mid], sel_idx[c2 * DET_MAXK + mid], c2, s, id, c)) lo = mid + 1;
        else hi = mid;
      }
      rank += lo == n2 ? s_fill[c2] : kprefix[c2 * DET_MAXK + lo];
    }
    if (rank < n_out) {
      float* row = s_rows + rank * DVC_DET_ROW;
      const float* bx = sel_box + (size_t)(c * DET_MAXK + r) * 4;
      row[0] = 0.f; row[1] = (float)c; row[2] = s;
      row[3] = bx[0]; row[4] = bx[1]; row[5] = bx[2]; row[6] = bx[3];
    }
  }
  __syncthreads();
  // rows (zero padded) go out with coalesced stores; `det` may be peer memory (the gather)
  float* det = a.det + (size_t)slot * DVC_MAX_DET * DVC_DET_ROW;
  for (int e = tid; e < DVC_MAX_DET * DVC_DET_ROW; e += DET_THREADS) det[e] = s_rows[e];
  // ---- phase 7: application filter (conf > 0.2, class in CONSIDER) -> integer boxes, in row order
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
  // ---- delivery: make the rows visible system-wide, then publish the slot flag
  __threadfence_system();
  __syncthreads();
  if (tid == 0 && a.flags) st_release_sys_u32(a.flags + slot, a.flag_value);
}

struct DetLaunch {
  DetArgs args;
  int smem;
};

size_t det_scratch_bytes(int B, int P, int C) { return (size_t)B * P * (C > 1 ? C - 1 : 1) * 8; }

int det_prepare(DetLaunch* L, const float* loc, const float* conf, const float* priors, const float* variances, int B,
                const dvc_det_params* p, const int* slot_index, float* det, int* det_count, int* boxes_i, int* box_count,
                uint32_t* flags, uint32_t flag_value, void* scratch, size_t scratch_bytes) {
  DVC_REQUIRE(p && B > 0, "detection: bad arguments");
  DVC_REQUIRE(p->num_classes >= 2 && p->num_classes <= DET_MAXC, "detection: num_classes=%d unsupported (max %d)", p->num_classes, DET_MAXC);
  DVC_REQUIRE(p->top_k > 0 && p->top_k <= DET_MAXK, "detection: top_k=%d unsupported (max %d)", p->top_k, DET_MAXK);
  DVC_REQUIRE(p->keep_top_k > 0, "detection: keep_top_k must be positive");
  DVC_REQUIRE(p->num_priors > 0 && p->num_priors < 65536, "detection: num_priors=%d unsupported", p->num_priors);
  const size_t need = det_scratch_bytes(B, p->num_priors, p->num_classes);
  DVC_REQUIRE(scratch != nullptr && scratch_bytes >= need, "detection: scratch %zu bytes < required %zu", scratch_bytes, need);
  DVC_REQUIRE((reinterpret_cast<uintptr_t>(scratch) & 15) == 0, "detection: scratch must be 16-byte aligned");
  DetArgs& a = L->args;
  a.loc = loc; a.conf = conf; a.priors = priors; a.variances = variances; a.B = B; a.p = *p; a.slot_index = slot_index;
  a.det = det; a.det_count = det_count; a.boxes_i = boxes_i; a.box_count = box_count; a.flags = flags; a.flag_value = flag_value;
  const long long per = (long long)p->num_priors * (p->num_classes - 1);
  a.pool_stride = per;
  a.pool_score = reinterpret_cast<float*>(scratch);
  a.pool_idx = reinterpret_cast<int*>(reinterpret_cast<float*>(scratch) + (size_t)B * per);
  const int C = p->num_classes;
  L->smem = (4 * DET_MAXC + 1 + 7) * 4 + C * DET_MAXK * (4 + 4 + 16) + ((C * DET_MAXK + 15) & ~15) + C * DET_MAXK * 4 + DVC_MAX_DET * DVC_DET_ROW * 4 + 64;
  return DVC_OK;
}

int det_launch(const DetLaunch* L, cudaStream_t st) {
  static bool attr_set[64] = {};
  int dev = 0;
  DVC_CHECK_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    DVC_CHECK_CUDA(cudaFuncSetAttribute(detection_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    attr_set[dev & 63] = true;
  }
  detection_kernel<<<L->args.B, DET_THREADS, L->smem, st>>>(L->args);
  count_launch();
  DVC_CHECK_CUDA(cudaGetLastError());
  return DVC_OK;
}

DetLaunch* det_alloc() { return new DetLaunch(); }
void det_free(DetLaunch* L) { delete L; }
void det_set_flag_value(DetLaunch* L, uint32_t v) { L->args.flag_value = v; }
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
// The candidate pool lives in this file-scope device array instead of caller memory.  Every
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
              "detection (static scratch): B=%d x %d priors x %d classes needs %zu bytes of candidate pool, the built-in "
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
  annotate_kernel<<<B, ANN_THREADS, 0, (cudaStream_t)stream>>>(a);
  count_launch();
  DVC_CHECK_CUDA(cudaGetLastError());
  return DVC_OK;
}
