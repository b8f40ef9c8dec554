/* dvc_b200 — C ABI of the B200-native MobileNet-SSD video-inference kernels.
 *
 * Every function returns 0 (DVC_OK) or a negative DVC_ERR_* code; dvc_last_error() gives
 * the message for the calling thread.  All pointers are device pointers unless the name
 * starts with h_.  `stream` is a cudaStream_t passed as void*.  Nothing in here allocates
 * device memory unless the comment says so (dvc_plan_create, dvc_p2p_*).
 *
 * Reference parity citations are against /root/reference (worker.py, server.py and
 * MobileNetSSD_deploy.prototxt.txt); the reference itself has no native code.
 */
#ifndef DVC_B200_H_
#define DVC_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DVC_OK 0
#define DVC_ERR_ARG (-1)
#define DVC_ERR_CUDA (-2)
#define DVC_ERR_UNSUPPORTED (-3)

#define DVC_MAX_DET 100      /* keep_top_k, prototxt:1909 */
#define DVC_DET_ROW 7        /* image_id,label,conf,xmin,ymin,xmax,ymax */

int dvc_version(void);
const char* dvc_last_error(void);
/* Number of kernel launches issued by this library in this process (bench "gpu_launches"). */
long long dvc_launch_count(void);
/* Select the device for this thread inside the library's own (statically linked) CUDA runtime.  Needed only
 * when one process drives several GPUs; call it whenever you switch devices. */
int dvc_set_device(int device);

/* ---- K1a: u8 HWC resize, the imutils.resize(width=400) step (worker.py:243).
 * mode 0: INTER_AREA decimation by coefficient tables (xs/xc/xa = start index, tap count,
 *         float weights [dst_w, max_taps]; same for y).
 * mode 1: INTER_AREA with integer ratios (ix, iy), exact integer sums.
 * mode 2: bilinear with Q11 fixed-point taps (enlarging INTER_AREA and plain INTER_LINEAR):
 *         xs = source index, xq = [dst_w,2] int weights; same for y.
 * src/dst are [B, h, w, 3]; src_frame_stride in bytes lets frames live in a ring. */
int dvc_resize_u8(const uint8_t* src, long long src_frame_stride, int B, int sh, int sw,
                  uint8_t* dst, long long dst_frame_stride, int dh, int dw, int mode,
                  const int* xs, const int* xc, const float* xa, int x_taps,
                  const int* ys, const int* yc, const float* ya, int y_taps,
                  const int* xq, const int* yq, int ix, int iy, void* stream);
/* Same, gated on the K14 membership epoch: when *epoch_now != *epoch_expect (two device words, read by the kernel)
 * nothing is read from `src` (no NVLink pull of a stale batch) and `dst` is left untouched. */
int dvc_resize_u8_gated(const uint8_t* src, long long src_frame_stride, int B, int sh, int sw,
                        uint8_t* dst, long long dst_frame_stride, int dh, int dw, int mode,
                        const int* xs, const int* xc, const float* xa, int x_taps,
                        const int* ys, const int* yc, const float* ya, int y_taps,
                        const int* xq, const int* yq, int ix, int iy,
                        const uint32_t* epoch_now, const uint32_t* epoch_expect, void* stream);

/* ---- K1b: fused frame fetch + cv2.resize(300x300, INTER_LINEAR) + (x-mean_c)*scale +
 * conv0 3x3/s2 + bias + ReLU (worker.py:245-246, prototxt:9-41).  `src400` may be a peer
 * (NVLink-mapped) pointer: the loads ARE the frame scatter.  Output NHWC fp16
 * [B,150,150,32].  w0 is fp32 [27][32] ((kh,kw,cin) major, cout innermost), b0 fp32 [32].
 * xs/xq/ys/yq are the Q11 bilinear taps from (sh,sw) to 300x300.  h_mean3 is a HOST pointer to
 * the 3 per-channel means.  frame_index (optional, may be NULL) maps batch slot -> frame slot
 * inside the source ring.  Frames must be 16-byte aligned (base and stride). */
int dvc_preproc_conv0(const uint8_t* src400, long long src_frame_stride, const int* frame_index, int B, int sh, int sw,
                      const int* xs, const int* xq, const int* ys, const int* yq,
                      const float* w0, const float* b0, const float* h_mean3, float scale,
                      void* out_f16, void* stream);

/* ---- K3: depthwise 3x3 (+bias, +ReLU), NHWC fp16, pad 1, stride 1|2 (prototxt:42-75 ...).
 * w is fp32 [9][C] (tap major), b fp32 [C].  The input tile + halo arrives by one 4-D TMA box load
 * (zero-filled outside the image = the padding); products are fp16 x fp16 accumulated in fp32
 * (FHFMA), so the weights are rounded to fp16 inside the kernel. */
int dvc_dwconv3x3(const void* in_f16, int B, int H, int W, int C, int stride, const float* w, const float* b,
                  int relu, void* out_f16, void* stream);

   /* non-TMA variant, for A/B measurements */

/* ---- K2/K5: pointwise conv as GEMM on tcgen05 tensor cores with TMA-fed operands and a
 * TMEM accumulator: out[M,N] = act(A[M,K] * Wt[N,K]^T + bias[N]).  A and Wt are K-major
 * fp16 (NHWC activations / [Cout,Cin] weights).  N % 16 == 0, K % 8 == 0.
 * out_mode 0: fp16 [M, ldc] (+ReLU if relu).
 * out_mode 1: SSD head split store (prototxt:1147-1238 + 1817-1844): row m = b*hw + p;
 *             columns [0,n_loc) -> loc[b][prior_off*4 + p*n_loc + c] (fp32, row pitch loc_pitch),
 *             columns [n_loc,n_loc+n_conf) -> conf[b][prior_off*ncls... ] likewise. */
typedef struct {
  float* loc;            /* [B, loc_pitch]  */
  float* conf;           /* [B, conf_pitch] */
  long long loc_pitch;   /* 1917*4  */
  long long conf_pitch;  /* 1917*21 */
  int hw;                /* pixels per frame of this tap */
  int n_loc;             /* anchors*4  */
  int n_conf;            /* anchors*classes */
  long long loc_off;     /* prior_off*4 */
  long long conf_off;    /* prior_off*classes */
} dvc_head_out;

int dvc_gemm_f16(const void* A, const void* Wt, const float* bias, int M, int N, int K, int lda,
                 int out_mode, void* out_f16, int ldc, int relu, const dvc_head_out* head,
                 int block_n, void* stream);

/* ---- K4: 3x3 stride-2 pad-1 patch gather (im2col) NHWC fp16 -> [B*Ho*Wo, 9*C] with
 * (kh,kw,c) column order; the dense extras convs (prototxt:922-954 ...) are then a GEMM. */
int dvc_im2col3x3s2(const void* in_f16, int B, int H, int W, int C, void* out_f16, void* stream);

/* ---- K4 (implicit): the same convolution without the patch buffer.  out[B*Ho*Wo, N] = act(patches(in) * Wt[N, 9*C]^T +
 * bias): the K2 GEMM kernel with its activation operand fetched by im2col-mode TMA loads (128 output pixels x 64
 * channels of one filter tap per load, padding halo zero-filled by the TMA unit).  C must be a multiple of 64;
 * Wt is in (kh, kw, c) column order as for dvc_im2col3x3s2 + dvc_gemm_f16, and the result is bit-identical to that pair. */
int dvc_conv3x3s2_f16(const void* in_f16, int B, int H, int W, int C, const void* Wt_f16, const float* bias, int N,
                      void* out_f16, int ldc, int relu, int block_n, void* stream);
/* Test hook: fetch operand tile (first output pixel m0, K block kb) of that convolution into out_tile_f16[128][64];
 * *status (device int) = 0 on success, 1 if the load did not complete. */
int dvc_debug_im2col_tile(const void* in_f16, int B, int H, int W, int C, int m0, int kb, void* out_tile_f16, int* status,
                          void* stream);

/* ---- K6: PriorBox for one tap (prototxt:1239-1256), written at prior_off in [P,4] arrays. */
int dvc_priorbox(float* boxes, float* variances, int prior_off, int grid_h, int grid_w, int img_h, int img_w,
                 const float* anchor_wh, int anchors, float offset, const float* h_var4, int clip, void* stream);

/* ---- K7+K8+K9: softmax + CENTER_SIZE decode + per-class top-k + NMS + keep_top_k + app
 * filter, one CTA per frame (prototxt:1872-1912, worker.py:253-269).  Results are written
 * through `out` pointers that may live on a peer GPU (the detection gather); when `flags`
 * is non-NULL, flags[slot] is set to `flag_value` with release.sys semantics after the rows.
 *   det      [B][100][7] fp32  rows sorted by (conf desc, prior asc, label asc), zero padded
 *   det_count[B]               number of valid rows
 *   boxes_i  [B][100][4] int32 app-filtered boxes scaled by (out_w,out_h) and truncated
 *   box_count[B]               number of app-filtered boxes ("person: N")
 * slot_index (optional) maps batch entry -> output slot.
 * Three launches: a softmax pass, one CTA per (frame, class) for top-k + NMS, one CTA per frame for
 * the global top-k, the filter and the delivery.
 * `scratch` is caller-provided global memory that carries the class-major probabilities and the
 * per-class survivor lists between those launches: dvc_detection_scratch_bytes(B, priors, classes)
 * bytes, 16-byte aligned.  Every call rewrites what it reads; concurrent calls must not share it. */
typedef struct {
  int num_priors, num_classes, background;
  float conf_thresh, nms_thresh;
  int top_k, keep_top_k;
  float app_conf;            /* 0.2, worker.py:186,257 */
  uint32_t consider_mask;    /* bit c set => class c is drawn/counted (person = 1<<15) */
  int out_w, out_h;          /* 400 x h' */
} dvc_det_params;

size_t dvc_detection_scratch_bytes(int B, int num_priors, int num_classes);
int dvc_detection_output(const float* loc, const float* conf, const float* priors, const float* variances,
                         int B, const dvc_det_params* p, const int* slot_index,
                         float* det, int* det_count, int* boxes_i, int* box_count,
                         uint32_t* flags, uint32_t flag_value, void* scratch, size_t scratch_bytes, void* stream);

/* ---- Allocation-free variant of dvc_detection_output (no counterpart in the reference, which
 * has no native code at all).
 *
 * Who it is for: embedders that cannot hand this library any device memory beyond the inputs
 * and outputs - a volunteer's host application that reserved the whole GPU in its own pool at
 * start-up, a capture loop that cannot afford the device-wide stall of cudaMalloc/cudaFree, or
 * a plugin host written in C with no allocator to lend.  For them the detection scratch
 * sits in a device array built into the library (DVC_DET_STATIC_SCRATCH_BYTES on every GPU that
 * loads the library, held until the process exits).  Nothing is allocated or freed here, by
 * CUDA or by any other library.
 *
 * Limit: dvc_detection_scratch_bytes(B, num_priors, num_classes) must not exceed
 * DVC_DET_STATIC_SCRATCH_BYTES (the 1917-prior, 21-class SSD fits B <= 18); larger requests
 * are rejected with DVC_ERR_ARG and a message that names both sizes.
 *
 * The first call on each device looks the array's address up (that device must be current);
 * make that first call before any stream capture begins.  All calls on a device share the one
 * array: two calls whose kernels can overlap on that GPU (different streams, or concurrent
 * host threads) race on it - serialise them.  The array keeps whatever the previous call left;
 * a call only ever reads scratch entries it wrote itself earlier in the same call, so nothing has
 * to be cleared between calls.  Outputs are identical to dvc_detection_output. */
#define DVC_DET_STATIC_SCRATCH_BYTES (4 * 1024 * 1024)
int dvc_detection_output_static(const float* loc, const float* conf, const float* priors, const float* variances,
                                int B, const dvc_det_params* p, const int* slot_index,
                                float* det, int* det_count, int* boxes_i, int* box_count,
                                uint32_t* flags, uint32_t flag_value, void* stream);

/* ---- K10: annotation raster on the 400-wide BGR frames (worker.py:268-276): blue
 * thickness-2 rectangles, then two pre-rendered text masks (id at (10,25) red; "person: N"
 * at (10,h-20) green).  id_mask is [id_h][id_w] u8 placed at (id_x,id_y); label_masks is
 * [n_labels][lab_h][lab_w], entry min(count, n_labels-1) is placed at (lab_x,lab_y). */
int dvc_annotate(uint8_t* frames, long long frame_stride, const int* frame_index, int B, int h, int w,
                 const int* boxes_i, const int* box_count, const int* slot_index,
                 const uint8_t* id_mask, int id_x, int id_y, int id_w, int id_h,
                 const uint8_t* label_masks, int n_labels, int lab_x, int lab_y, int lab_w, int lab_h,
                 void* stream);

/* ---- plan executor: the whole network as one native launch sequence. --------------------
 * A plan owns pre-encoded TMA descriptors and launch geometry for a fixed max batch; the
 * caller owns all memory (weights, activations, workspace).  Ops are appended once and then
 * dvc_plan_run() replays them on a stream (capturable in a CUDA graph). */
typedef struct dvc_plan dvc_plan;
dvc_plan* dvc_plan_create(void);
void dvc_plan_destroy(dvc_plan*);
int dvc_plan_num_ops(const dvc_plan*);
int dvc_plan_add_preproc_conv0(dvc_plan*, const uint8_t* src400, long long src_frame_stride, const int* frame_index,
                               int B, int sh, int sw, const int* xs, const int* xq, const int* ys, const int* yq,
                               const float* w0, const float* b0, const float* h_mean3, float scale, void* out_f16,
                               int slot0 /* first ring slot read by this op; src400 is the ring base */);
int dvc_plan_add_dwconv(dvc_plan*, const void* in_f16, int B, int H, int W, int C, int stride, const float* w,
                        const float* b, int relu, void* out_f16);
int dvc_plan_add_gemm(dvc_plan*, const void* A, const void* Wt, const float* bias, int M, int N, int K, int lda,
                      int out_mode, void* out_f16, int ldc, int relu, const dvc_head_out* head, int block_n);
int dvc_plan_add_im2col(dvc_plan*, const void* in_f16, int B, int H, int W, int C, void* out_f16);
int dvc_plan_add_conv3x3s2(dvc_plan*, const void* in_f16, int B, int H, int W, int C, const void* Wt_f16, const float* bias,
                           int N, void* out_f16, int ldc, int relu, int block_n);
/* EXPERIMENTAL: dvc_dwpw_fused_s1 as a plan op (the engine adds it only when DVC_FUSE_DWPW=1). */
int dvc_plan_add_dwpw_fused(dvc_plan*, const void* in_f16, int B, int H, int W, int C_in, const float* wdw, const float* bdw,
                            const void* wpw_f16, const float* bpw, int C_out, void* out_f16);
int dvc_plan_add_detection(dvc_plan*, const float* loc, const float* conf, const float* priors, const float* variances,
                           int B, const dvc_det_params* p, const int* slot_index, float* det, int* det_count,
                           int* boxes_i, int* box_count, uint32_t* flags, uint32_t flag_value,
                           void* scratch, size_t scratch_bytes);
/* Re-point the frame source / result sinks of an existing plan (elastic re-dispatch). */
int dvc_plan_set_source(dvc_plan*, const uint8_t* src400, long long src_frame_stride, const int* frame_index);
int dvc_plan_set_sinks(dvc_plan*, float* det, int* det_count, int* boxes_i, int* box_count, uint32_t* flags,
                       const int* slot_index);
int dvc_plan_set_flag_value(dvc_plan*, uint32_t flag_value);
/* Device-side trace points (the reference has one wall-clock line per job, worker.py:105,219): two optional arrays of
 * B nanosecond %globaltimer stamps indexed by the frame's position in the batch - when K1b starts pulling the frame,
 * and when K9 has delivered its rows.  NULL switches a stamp off. */
int dvc_plan_set_timestamps(dvc_plan*, unsigned long long* t_ingest, unsigned long long* t_done);
/* K14 epoch gate (server.py:104-154 has only host-side lists): the deliver kernel compares the membership epoch
 * word of this GPU (written by dvc_epoch_broadcast) with the epoch the batch was dispatched under (a device word
 * the dispatcher writes); on mismatch the frame is not delivered (sinks and flags untouched) and *stale_count is
 * incremented once per frame so the scheduler can re-dispatch.  Pass NULLs to switch the gate off. */
int dvc_plan_set_epoch_gate(dvc_plan*, const uint32_t* epoch_now, const uint32_t* epoch_expect, uint32_t* stale_count);
/* SSD head GEMMs on a side branch of the plan (they depend only on their tap layer and only the detection op needs
 * them): off the critical path of a single batch.  Requires that no later op overwrites a head's input. */
int dvc_plan_set_head_branch(dvc_plan*, int on);
int dvc_plan_run(dvc_plan*, void* stream);

/* Fused depthwise 3x3 (pad 1, stride 1 or 2, +bias+ReLU) -> pointwise 1x1 (+bias+ReLU) of one MobileNet block
 * (prototxt:42-106 ... 302-366); the intermediate stays in shared memory as the A operand of tcgen05.mma and the
 * tensor work hides behind the depthwise arithmetic (csrc/dwpw_fused.cu).  The reference runs these as two cv2.dnn
 * layers.  in [B,H,W,C_in] fp16, wdw [9][C_in] fp32, bdw [C_in], wpw [C_out][C_in] fp16, bpw [C_out],
 * out [B,Ho,Wo,C_out] fp16; C_in = 32 or a multiple of 64, C_out in {64,128,192,256}. */
int dvc_dwpw_fused(const void* in_f16, int B, int H, int W, int C_in, int stride, const float* wdw, const float* bdw,
                   const void* wpw_f16, const float* bpw, int C_out, void* out_f16, void* stream);
int dvc_dwpw_fused_s1(const void* in_f16, int B, int H, int W, int C_in, const float* wdw, const float* bdw,
                      const void* wpw_f16, const float* bpw, int C_out, void* out_f16, void* stream);

/* ---- P2P substrate (csrc/p2p.cu).  These DO allocate / map device memory (set-up time only).
 * A buffer is owned by one GPU process, exported as a 64-byte CUDA IPC handle and mapped by the
 * other per-GPU processes of the box; kernels then load/store it directly over NVLink.  This
 * replaces the reference's 4-hop TCP relay through the server (server.py:56,89,60,73). */
int dvc_p2p_alloc(size_t bytes, void** out);                 /* cudaMalloc + zero */
int dvc_p2p_free(void* p);
int dvc_p2p_get_handle(void* p, void* h_handle64);           /* export (owner) */
int dvc_p2p_open(const void* h_handle64, void** out);        /* map (peer), enables peer access lazily */
int dvc_p2p_close(void* p);
int dvc_p2p_enable_peer(int peer_device);                    /* same-process multi-GPU */
/* K14 membership epoch: write `value` (release.sys) to each of n peer addresses (server.py:104-154). */
int dvc_epoch_broadcast(void** d_dst_ptrs, int n, uint32_t value, void* stream);
/* K13 weight distribution: push `bytes` from src into n peer buffers (worker.py:194 loads per client). */
int dvc_push_multicast(const void* src, void** d_dst_ptrs, int n, size_t bytes, void* stream);
/* NVLink read-bandwidth probe (16-byte loads of a possibly remote buffer). */
int dvc_peer_read(const void* src, size_t bytes, void* d_sink, void* stream);


/* ---- Mailboxes: device-side flow control for the in-box data plane (csrc/mailbox.cu). --------------------------
 * The reference's flow control is one blocking REQ/REP round trip per chunk through the server (`b'OK'`,
 * server.py:62, worker.py:180) and its dispatch is server.py:77-91.  Here the requesting client's GPU owns, per
 * (worker, lane), one 32-byte descriptor and one completion word; everything is written and polled by kernels
 * with system-scope release / acquire, so no collective and no host round trip orders a step:
 *
 *   client  dvc_mbx_post       fills the descriptor, then st.release.sys(desc.seq)              (client-local memory)
 *   worker  dvc_mbx_wait_desc  polls ld.acquire.sys(desc.seq) over NVLink until it equals the lane's next sequence
 *                              number (bounded wait), then copies source pointer / epoch / sequence into the lane's
 *                              local control block, which K1a, K1b and K9 of that lane read
 *   worker  K1a                pulls the frames from the client's arena with peer loads (the scatter)
 *   worker  K9                 pushes rows with peer stores; the last CTA of the batch does
 *                              st.release.sys(done word = seq)                                    (the gather)
 *   client  dvc_mbx_wait_done  polls ld.acquire.sys(done word) (client-local memory, bounded), then the stream goes
 *                              on to the device->host copy of the rows; a time-out is reported, never a hang
 *
 * Reuse of a lane's arena slot and result slot needs no separate credit word: the next descriptor of the lane is
 * posted only after the client has consumed the previous result, in stream order on the client.
 * Every wait is bounded by a time-out measured with %globaltimer, so a dead peer cannot wedge a GPU. */
typedef struct dvc_mbx_desc {
  uint32_t seq;        /* written last with release.sys; 0xFFFFFFFF = stop serving this lane */
  uint32_t epoch;      /* membership epoch of the worker slot the batch was dispatched under (K14) */
  uint32_t n_valid;    /* frames of the batch that carry data (<= B) */
  uint32_t kind;       /* 0: frames are in the client's arena at src_off (NVLink pull); 1: worker-local frames */
  uint64_t src_off;    /* byte offset of the batch's first frame inside the arena */
  uint32_t batch_id;
  uint32_t reserved;
} dvc_mbx_desc;

/* Worker-local, per lane, device memory: written by dvc_mbx_wait_desc, read by that lane's kernels. */
typedef struct dvc_mbx_lane_ctl {
  const uint8_t* src;  /* where K1a reads this batch's raw frames */
  uint32_t gate_now;   /* K1a / K1b run only when gate_now == gate_expect (nothing arrived: they differ) */
  uint32_t gate_expect;/* K9 delivers only when the client's live epoch word still equals this */
  uint32_t flag_value; /* the sequence number K9 publishes in the done word */
  uint32_t status;     /* DVC_MBX_* below */
  uint32_t batch_id;
  uint32_t n_valid;
  uint32_t next_seq;   /* the lane's position: the sequence number its next wait expects (advanced on every GO) */
  uint32_t reserved;
} dvc_mbx_lane_ctl;

/* Host-visible (pinned, mapped) mirror of what a wait saw; `wait_ns` is how long the kernel polled. */
typedef struct dvc_mbx_host_status {
  uint32_t status, seq, batch_id, n_valid;
  unsigned long long wait_ns;
} dvc_mbx_host_status;

#define DVC_MBX_NOTHING 0u   /* the wait timed out */
#define DVC_MBX_GO 1u
#define DVC_MBX_STOP 2u
#define DVC_MBX_SEQ_STOP 0xFFFFFFFFu
#define DVC_MBX_MAX_GROUP 16

/* Client: publish n descriptors (n <= DVC_MBX_MAX_GROUP).  h_desc_ptrs[i] is the device address of descriptor i
 * (client-local); h_descs[i] its new contents (seq included).  Host arrays, copied into the launch. */
int dvc_mbx_post(void* const* h_desc_ptrs, const dvc_mbx_desc* h_descs, int n, void* stream);
/* Worker: wait (at most timeout_ns) until desc->seq == expect_seq (0: use and advance ctl->next_seq, so several
 * waits can be queued ahead of their batches) or == DVC_MBX_SEQ_STOP, then fill *ctl and
 * *h_status.  `desc` usually is a peer address on the client GPU; `arena_base` is the worker's
 * mapping of the arena, `local_src` the worker's own frame buffer (kind 1).  h_status is pinned mapped host memory
 * (device-accessible address) or NULL. */
int dvc_mbx_wait_desc(const dvc_mbx_desc* desc, uint32_t expect_seq, const uint8_t* arena_base, const uint8_t* local_src,
                      dvc_mbx_lane_ctl* ctl, dvc_mbx_host_status* h_status, unsigned long long timeout_ns, void* stream);
/* Client: wait (at most timeout_ns) until *done_ptrs[i] == seqs[i] for every i < n; h_status[i].status = GO or
 * NOTHING (timed out).  h_status is pinned mapped host memory, one entry per waited word. */
int dvc_mbx_wait_done(void* const* h_done_ptrs, const uint32_t* h_seqs, int n, dvc_mbx_host_status* h_status,
                      unsigned long long timeout_ns, void* stream);
/* Client, one call per dispatch to a remote worker, all on `stream`: optional host->device copy of the frames into
 * the arena, dvc_mbx_post of one descriptor, dvc_mbx_wait_done on its done word, optional device->host copy of the
 * result slot, optional cudaEventRecord(event).  Pointers that are NULL switch their step off. */
int dvc_mbx_dispatch(const void* h_src, void* d_dst, size_t src_bytes, void* desc_ptr, const dvc_mbx_desc* h_desc,
                     void* done_ptr, dvc_mbx_host_status* h_status, unsigned long long timeout_ns,
                     const void* d_result, void* h_result, size_t result_bytes, void* event, void* stream);
/* Device->host copy of a result slot into pinned memory, then cudaEventRecord(event) (event may be NULL). */
int dvc_mbx_copy_out(const void* d_result, void* h_result, size_t bytes, void* event, void* stream);
/* Host-side helpers that fold what would be several Python-level CUDA calls into one (the client dispatches ~10^4
 * batches a second on a full box).  dvc_mbx_poll: cudaEventQuery over n events, h_done[i] = 1 where busy and finished,
 * returns how many.  dvc_mbx_worker_enqueue: [wait for a descriptor] -> [launch the lane's CUDA graph] -> [record event];
 * NULL switches a step off.  dvc_mbx_local_dispatch: the client's own GPU as a worker, everything on one stream:
 * post -> wait_desc (immediate) -> graph -> wait_done -> copy-out -> event. */
int dvc_mbx_poll(void* const* h_events, const unsigned char* h_busy, int n, unsigned char* h_done);
int dvc_mbx_worker_enqueue(const dvc_mbx_desc* desc, uint32_t expect_seq, const uint8_t* arena_base, const uint8_t* local_src,
                           dvc_mbx_lane_ctl* ctl, dvc_mbx_host_status* h_status, unsigned long long timeout_ns,
                           void* graph_exec, void* event, void* stream);
int dvc_mbx_local_dispatch(void* desc_ptr, const dvc_mbx_desc* h_desc, const uint8_t* arena_base, const uint8_t* local_src,
                           dvc_mbx_lane_ctl* ctl, void* graph_exec, void* done_ptr, dvc_mbx_host_status* h_status,
                           unsigned long long timeout_ns, const void* d_result, void* h_result, size_t result_bytes,
                           void* event, void* stream);
/* st.release.sys of one word (epoch bump, hand-rolled completion after an extra copy). */
int dvc_mbx_signal(uint32_t* word, uint32_t value, void* stream);
/* Bind a plan to a lane: K1b's gate reads ctl->gate_now / gate_expect, K9 takes its flag value from ctl->flag_value,
 * compares *live_epoch (the client's word for this worker, a peer address) with ctl->gate_expect, counts rejected
 * frames in *stale_count and lets the last CTA of the batch publish *done_word = flag value.  `done_counter` is one
 * zero-initialised device word owned by the lane. */
int dvc_plan_set_mailbox(dvc_plan*, const dvc_mbx_lane_ctl* ctl, const uint32_t* live_epoch, uint32_t* done_word,
                         uint32_t* done_counter, uint32_t* stale_count);
/* K1a whose source base pointer is read from device memory at run time (`*src_indirect`, e.g. &ctl->src) and which is
 * gated like dvc_resize_u8_gated. */
int dvc_resize_u8_indirect(const uint8_t* const* src_indirect, long long src_frame_stride, int B, int sh, int sw,
                           uint8_t* dst, long long dst_frame_stride, int dh, int dw, int mode,
                           const int* xs, const int* xc, const float* xa, int x_taps,
                           const int* ys, const int* yc, const float* ya, int y_taps,
                           const int* xq, const int* yq, int ix, int iy,
                           const uint32_t* epoch_now, const uint32_t* epoch_expect, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DVC_B200_H_ */
