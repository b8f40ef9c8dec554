// Native plan executor: the whole MobileNet-SSD forward pass as one pre-planned launch
// sequence (TMA descriptors encoded once, launch geometry frozen), replayed on a stream.
// This replaces cv2.dnn's net.forward() graph walk (reference worker.py:248-249).
#include <vector>

#include "common.cuh"

namespace dvc {

// launch descriptors defined in the kernel translation units
struct GemmArgs;
struct GemmLaunch;
struct PreprocLaunch;
struct DetLaunch;

int gemm_prepare(GemmLaunch* L, const void* A, const void* Wt, const float* bias, int M, int N, int K, int lda, int out_mode,
                 void* out, int ldc, int relu, const dvc_head_out* head, int block_n);
int gemm_prepare_conv3x3s2(GemmLaunch* L, const void* in, int B, int H, int W, int C, const void* Wt, const float* bias, int N,
                           void* out, int ldc, int relu, int block_n);
int gemm_launch(const GemmLaunch* L, cudaStream_t st);
GemmLaunch* gemm_alloc();
void gemm_free(GemmLaunch*);

int preproc_prepare(PreprocLaunch* L, const uint8_t* src400, long long src_frame_stride, const int* frame_index, int B, int sh,
                    int sw, const int* xs, const int* xq, const int* ys, const int* yq, const float* w0, const float* b0,
                    const float* h_mean3, float scale, void* out, int band_rows_max, int slot0);
int preproc_launch(const PreprocLaunch* L, cudaStream_t st);
PreprocLaunch* preproc_alloc();
void preproc_free(PreprocLaunch*);
void preproc_set_source(PreprocLaunch*, const uint8_t* src, long long stride, const int* frame_index);

int det_prepare(DetLaunch* L, const float* loc, const float* conf, const float* priors, const float* variances, int B,
                const dvc_det_params* p, const int* slot_index, float* det, int* det_count, int* boxes_i, int* box_count,
                uint32_t* flags, uint32_t flag_value, void* scratch, size_t scratch_bytes);
int det_launch(const DetLaunch* L, cudaStream_t st);
DetLaunch* det_alloc();
void det_free(DetLaunch*);
void det_set_flag_value(DetLaunch*, uint32_t);
void det_set_timestamps(DetLaunch*, unsigned long long* t_done);
void preproc_set_timestamps(PreprocLaunch*, unsigned long long* t_ingest);
void preproc_set_epoch_gate(PreprocLaunch*, const uint32_t* now, const uint32_t* expect);
void det_set_epoch_gate(DetLaunch*, const uint32_t* now, const uint32_t* expect, uint32_t* stale_count);
void det_set_mailbox(DetLaunch*, const uint32_t* flag_value_ptr, uint32_t* done_word, uint32_t* done_counter);
void det_set_sinks(DetLaunch*, float* det, int* det_count, int* boxes_i, int* box_count, uint32_t* flags, const int* slot_index);

struct DwLaunch;
int dw_prepare(DwLaunch* L, const void* in, int B, int H, int W, int C, int stride, const float* w, const float* b, int relu,
               void* out);
int dw_launch(const DwLaunch* L, cudaStream_t st);
DwLaunch* dw_alloc();
void dw_free(DwLaunch*);
int im2col_launch(const void* in, int B, int H, int W, int C, void* out, cudaStream_t st);
int dwpw_fused_s1_launch(const void* in, int B, int H, int W, int C_in, const float* wdw, const float* bdw, const void* wpw,
                         const float* bpw, int C_out, void* out, cudaStream_t st);

enum OpKind { OP_PREPROC, OP_DW, OP_GEMM, OP_IM2COL, OP_DET, OP_DWPW };

struct Op {
  OpKind kind;
  GemmLaunch* gemm = nullptr;
  PreprocLaunch* pre = nullptr;
  DetLaunch* det = nullptr;
  DwLaunch* dw = nullptr;
  // im2col
  const void* in = nullptr; void* out = nullptr; const float* w = nullptr; const float* b = nullptr;
  int B = 0, H = 0, W = 0, C = 0, stride = 1, relu = 0;
  // fused depthwise -> pointwise (experimental): w / b above are the depthwise ones
  const void* wpw = nullptr; const float* bpw = nullptr; int C_out = 0;
  bool is_head = false;            // SSD head GEMM: may run on the plan's side branch (see dvc_plan_set_head_branch)
};

}  // namespace dvc

struct dvc_plan {
  std::vector<dvc::Op> ops;
  // side branch for the SSD heads: they depend only on their tap layer and are needed only by the detection op, so
  // they leave the critical path (one fork event per head, one join before detection; captured into CUDA graphs as a
  // parallel branch)
  cudaStream_t side = nullptr;
  std::vector<cudaEvent_t> forks;
  cudaEvent_t join = nullptr;
};

using namespace dvc;

extern "C" dvc_plan* dvc_plan_create(void) { return new dvc_plan(); }

extern "C" void dvc_plan_destroy(dvc_plan* p) {
  if (!p) return;
  for (auto e : p->forks) cudaEventDestroy(e);
  if (p->join) cudaEventDestroy(p->join);
  if (p->side) cudaStreamDestroy(p->side);
  for (auto& op : p->ops) {
    if (op.gemm) gemm_free(op.gemm);
    if (op.pre) preproc_free(op.pre);
    if (op.det) det_free(op.det);
    if (op.dw) dw_free(op.dw);
  }
  delete p;
}

extern "C" int dvc_plan_num_ops(const dvc_plan* p) { return p ? (int)p->ops.size() : 0; }

extern "C" int dvc_plan_add_preproc_conv0(dvc_plan* p, const uint8_t* src400, long long src_frame_stride, const int* frame_index,
                                          int B, int sh, int sw, const int* xs, const int* xq, const int* ys, const int* yq,
                                          const float* w0, const float* b0, const float* h_mean3, float scale, void* out_f16,
                                          int slot0) {
  DVC_REQUIRE(p, "plan is null");
  Op op; op.kind = OP_PREPROC; op.pre = preproc_alloc();
  int rc = preproc_prepare(op.pre, src400, src_frame_stride, frame_index, B, sh, sw, xs, xq, ys, yq, w0, b0, h_mean3, scale, out_f16, 0, slot0);
  if (rc) { preproc_free(op.pre); return rc; }
  p->ops.push_back(op);
  return DVC_OK;
}

extern "C" int dvc_plan_add_dwconv(dvc_plan* p, const void* in_f16, int B, int H, int W, int C, int stride, const float* w,
                                   const float* b, int relu, void* out_f16) {
  DVC_REQUIRE(p, "plan is null");
  DVC_REQUIRE(C % 8 == 0 && (stride == 1 || stride == 2), "dwconv: C=%d stride=%d unsupported", C, stride);
  Op op; op.kind = OP_DW; op.dw = dw_alloc();
  int rc = dw_prepare(op.dw, in_f16, B, H, W, C, stride, w, b, relu, out_f16);
  if (rc) { dw_free(op.dw); return rc; }
  p->ops.push_back(op);
  return DVC_OK;
}

extern "C" int dvc_plan_add_gemm(dvc_plan* p, const void* A, const void* Wt, const float* bias, int M, int N, int K, int lda,
                                 int out_mode, void* out_f16, int ldc, int relu, const dvc_head_out* head, int block_n) {
  DVC_REQUIRE(p, "plan is null");
  Op op; op.kind = OP_GEMM; op.gemm = gemm_alloc();
  int rc = gemm_prepare(op.gemm, A, Wt, bias, M, N, K, lda, out_mode, out_f16, ldc, relu, head, block_n);
  if (rc) { gemm_free(op.gemm); return rc; }
  op.is_head = (out_mode == 1);
  p->ops.push_back(op);
  return DVC_OK;
}

/* Run the SSD head GEMMs on a side branch (on = 1).  Only valid when no later op of the plan overwrites a head's input
 * before the detection op - i.e. when every tap layer writes into a buffer of its own (engine.py does that). */
extern "C" int dvc_plan_set_head_branch(dvc_plan* p, int on) {
  DVC_REQUIRE(p, "plan is null");
  if (!on) {
    if (p->side) { cudaStreamDestroy(p->side); p->side = nullptr; }
    return DVC_OK;
  }
  if (!p->side) {
    DVC_CHECK_CUDA(cudaStreamCreateWithFlags(&p->side, cudaStreamNonBlocking));
    DVC_CHECK_CUDA(cudaEventCreateWithFlags(&p->join, cudaEventDisableTiming));
  }
  size_t heads = 0;
  for (auto& op : p->ops) heads += op.is_head ? 1 : 0;
  while (p->forks.size() < heads) {
    cudaEvent_t e;
    DVC_CHECK_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    p->forks.push_back(e);
  }
  return DVC_OK;
}

extern "C" int dvc_plan_add_im2col(dvc_plan* p, const void* in_f16, int B, int H, int W, int C, void* out_f16) {
  DVC_REQUIRE(p, "plan is null");
  DVC_REQUIRE(C % 8 == 0, "im2col: C=%d unsupported", C);
  Op op; op.kind = OP_IM2COL; op.in = in_f16; op.out = out_f16; op.B = B; op.H = H; op.W = W; op.C = C;
  p->ops.push_back(op);
  return DVC_OK;
}

/* Dense 3x3 / stride-2 / pad-1 convolution as an implicit GEMM (no im2col buffer): see gemm_prepare_conv3x3s2. */
extern "C" int dvc_plan_add_conv3x3s2(dvc_plan* p, const void* in_f16, int B, int H, int W, int C, const void* Wt, const float* bias,
                                      int N, void* out_f16, int ldc, int relu, int block_n) {
  DVC_REQUIRE(p, "plan is null");
  Op op; op.kind = OP_GEMM; op.gemm = gemm_alloc();
  int rc = gemm_prepare_conv3x3s2(op.gemm, in_f16, B, H, W, C, Wt, bias, N, out_f16, ldc, relu, block_n);
  if (rc) { gemm_free(op.gemm); return rc; }
  p->ops.push_back(op);
  return DVC_OK;
}

/* EXPERIMENTAL (csrc/dwpw_fused.cu): one op for a stride-1 depthwise 3x3 + its pointwise 1x1, both with ReLU. */
extern "C" int dvc_plan_add_dwpw_fused(dvc_plan* p, const void* in_f16, int B, int H, int W, int C_in, const float* wdw,
                                       const float* bdw, const void* wpw_f16, const float* bpw, int C_out, void* out_f16) {
  DVC_REQUIRE(p, "plan is null");
  DVC_REQUIRE((C_in == 32 || C_in % 64 == 0) && C_out % 64 == 0 && C_out >= 64 && C_out <= 256,
              "dwpw_fused: C_in=%d / C_out=%d unsupported", C_in, C_out);
  Op op; op.kind = OP_DWPW; op.in = in_f16; op.out = out_f16; op.B = B; op.H = H; op.W = W; op.C = C_in;
  op.w = wdw; op.b = bdw; op.wpw = wpw_f16; op.bpw = bpw; op.C_out = C_out;
  p->ops.push_back(op);
  return DVC_OK;
}

extern "C" int dvc_plan_add_detection(dvc_plan* p, const float* loc, const float* conf, const float* priors,
                                      const float* variances, int B, const dvc_det_params* prm, const int* slot_index, float* det,
                                      int* det_count, int* boxes_i, int* box_count, uint32_t* flags, uint32_t flag_value,
                                      void* scratch, size_t scratch_bytes) {
  DVC_REQUIRE(p, "plan is null");
  Op op; op.kind = OP_DET; op.det = det_alloc();
  int rc = det_prepare(op.det, loc, conf, priors, variances, B, prm, slot_index, det, det_count, boxes_i, box_count, flags,
                       flag_value, scratch, scratch_bytes);
  if (rc) { det_free(op.det); return rc; }
  p->ops.push_back(op);
  return DVC_OK;
}

extern "C" int dvc_plan_set_source(dvc_plan* p, const uint8_t* src400, long long src_frame_stride, const int* frame_index) {
  DVC_REQUIRE(p, "plan is null");
  DVC_REQUIRE(src_frame_stride % 16 == 0 && (reinterpret_cast<uintptr_t>(src400) & 15) == 0, "set_source: unaligned frames");
  for (auto& op : p->ops)
    if (op.kind == OP_PREPROC) preproc_set_source(op.pre, src400, src_frame_stride, frame_index);
  return DVC_OK;
}

extern "C" int dvc_plan_set_sinks(dvc_plan* p, float* det, int* det_count, int* boxes_i, int* box_count, uint32_t* flags,
                                  const int* slot_index) {
  DVC_REQUIRE(p, "plan is null");
  for (auto& op : p->ops)
    if (op.kind == OP_DET) det_set_sinks(op.det, det, det_count, boxes_i, box_count, flags, slot_index);
  return DVC_OK;
}

extern "C" int dvc_plan_set_flag_value(dvc_plan* p, uint32_t flag_value) {
  DVC_REQUIRE(p, "plan is null");
  for (auto& op : p->ops)
    if (op.kind == OP_DET) det_set_flag_value(op.det, flag_value);
  return DVC_OK;
}

extern "C" int dvc_plan_set_timestamps(dvc_plan* p, unsigned long long* t_ingest, unsigned long long* t_done) {
  DVC_REQUIRE(p, "plan is null");
  for (auto& op : p->ops) {
    if (op.kind == OP_PREPROC) preproc_set_timestamps(op.pre, t_ingest);
    if (op.kind == OP_DET) det_set_timestamps(op.det, t_done);
  }
  return DVC_OK;
}

extern "C" int dvc_plan_set_epoch_gate(dvc_plan* p, const uint32_t* epoch_now, const uint32_t* epoch_expect, uint32_t* stale_count) {
  DVC_REQUIRE(p, "plan is null");
  DVC_REQUIRE((epoch_now == nullptr) == (epoch_expect == nullptr), "epoch gate: give both epoch words or neither");
  for (auto& op : p->ops) {
    if (op.kind == OP_DET) det_set_epoch_gate(op.det, epoch_now, epoch_expect, stale_count);
    if (op.kind == OP_PREPROC) preproc_set_epoch_gate(op.pre, epoch_now, epoch_expect);     // K1b: do not pull a stale batch
  }
  return DVC_OK;
}

extern "C" int dvc_plan_set_mailbox(dvc_plan* p, const dvc_mbx_lane_ctl* ctl, const uint32_t* live_epoch, uint32_t* done_word,
                                    uint32_t* done_counter, uint32_t* stale_count) {
  DVC_REQUIRE(p && ctl && live_epoch && done_counter, "plan_set_mailbox: plan, control block, epoch word and counter are required");
  for (auto& op : p->ops) {
    if (op.kind == OP_DET) {
      det_set_epoch_gate(op.det, live_epoch, &ctl->gate_expect, stale_count);
      det_set_mailbox(op.det, &ctl->flag_value, done_word, done_counter);
    }
    if (op.kind == OP_PREPROC) preproc_set_epoch_gate(op.pre, &ctl->gate_now, &ctl->gate_expect);
  }
  return DVC_OK;
}

extern "C" int dvc_plan_run(dvc_plan* p, void* stream) {
  DVC_REQUIRE(p, "plan is null");
  cudaStream_t st = (cudaStream_t)stream;
  size_t k = 0;
  bool forked = false;
  for (const auto& op : p->ops) {
    int rc = DVC_OK;
    if (p->side && op.is_head && k < p->forks.size()) {
      DVC_CHECK_CUDA(cudaEventRecord(p->forks[k], st));
      DVC_CHECK_CUDA(cudaStreamWaitEvent(p->side, p->forks[k], 0));
      rc = gemm_launch(op.gemm, p->side);
      if (rc) return rc;
      ++k;
      forked = true;
      continue;
    }
    if (forked && op.kind == OP_DET) {
      DVC_CHECK_CUDA(cudaEventRecord(p->join, p->side));
      DVC_CHECK_CUDA(cudaStreamWaitEvent(st, p->join, 0));
      forked = false;
    }
    switch (op.kind) {
      case OP_PREPROC: rc = preproc_launch(op.pre, st); break;
      case OP_DW: rc = dw_launch(op.dw, st); break;
      case OP_GEMM: rc = gemm_launch(op.gemm, st); break;
      case OP_IM2COL: rc = im2col_launch(op.in, op.B, op.H, op.W, op.C, op.out, st); break;
      case OP_DET: rc = det_launch(op.det, st); break;
      case OP_DWPW: rc = dwpw_fused_s1_launch(op.in, op.B, op.H, op.W, op.C, op.w, op.b, op.wpw, op.bpw, op.C_out, op.out, st); break;
    }
    if (rc) return rc;
  }
  if (forked) {                     // a plan without a detection op: join at the end
    DVC_CHECK_CUDA(cudaEventRecord(p->join, p->side));
    DVC_CHECK_CUDA(cudaStreamWaitEvent(st, p->join, 0));
  }
  return DVC_OK;
}
