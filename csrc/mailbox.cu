// Mailboxes: the device-side flow control of the in-box data plane.
//
// The reference orders its data plane with one blocking REQ/REP round trip per chunk (send_image waits for the
// hub's b'OK': server.py:62,73,89 and worker.py:157,180) and dispatches from a host loop (server.py:77-91).
// Here the requesting client's GPU owns one descriptor and one completion word per (worker, lane); the words are
// written with st.release.sys and polled with ld.acquire.sys by the kernels below, so neither a collective nor a
// host round trip sits between "frames are in the arena" -> "worker pulls them" -> "rows are back".
//
// Memory placement: descriptors, completion words, epoch words, the frame arena and the result slots all live in
// memory the CLIENT allocated and exported.  Workers only map it.  A worker that dies therefore takes no memory
// with it that anyone else dereferences; the client notices through the bounded wait and re-dispatches.
//
// Scopes and fences (who writes / who polls):
//   desc.{epoch,n_valid,kind,src_off,batch_id}  client kernel, plain stores, then __threadfence_system()
//   desc.seq                                    client kernel, st.release.sys        | worker: ld.acquire.sys (peer)
//   epoch word of a worker                      client kernel, st.release.sys        | worker K9: ld.acquire.sys (peer)
//   result rows / counts / boxes                worker K9, plain peer stores, then __threadfence_system()
//   done word                                   worker K9 (last CTA), st.release.sys | client: ld.acquire.sys (local)
// Every poll loop is bounded by a %globaltimer deadline: a peer that never answers costs a time-out, not a hang.
#include <string.h>

#include "common.cuh"

namespace dvc {

struct PostArgs {
  int n;
  dvc_mbx_desc* ptr[DVC_MBX_MAX_GROUP];
  dvc_mbx_desc d[DVC_MBX_MAX_GROUP];
};

__global__ void mbx_post_kernel(const PostArgs a) {
  const int i = threadIdx.x;
  if (i >= a.n) return;
  dvc_mbx_desc* p = a.ptr[i];
  p->epoch = a.d[i].epoch;
  p->n_valid = a.d[i].n_valid;
  p->kind = a.d[i].kind;
  p->src_off = a.d[i].src_off;
  p->batch_id = a.d[i].batch_id;
  p->reserved = 0;
  __threadfence_system();
  st_release_sys_u32(&p->seq, a.d[i].seq);
}

__device__ __forceinline__ uint32_t ld_relaxed_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// One thread.  The descriptor is usually on another GPU: every poll is one NVLink round trip (~2 us), so the loop
// generates next to no traffic and needs no back-off beyond a short nanosleep.
__global__ void mbx_wait_desc_kernel(const dvc_mbx_desc* desc, uint32_t expect_seq, const uint8_t* arena_base,
                                     const uint8_t* local_src, dvc_mbx_lane_ctl* ctl, dvc_mbx_host_status* hs,
                                     unsigned long long timeout_ns) {
  if (threadIdx.x != 0) return;
  const unsigned long long t0 = globaltimer_ns();
  uint32_t status = DVC_MBX_NOTHING, seq = 0;
  if (expect_seq == 0u) expect_seq = ctl->next_seq;     // the lane keeps its own position: waits can be queued ahead
  for (;;) {
    seq = ld_acquire_sys_u32(&desc->seq);
    if (seq == expect_seq) { status = DVC_MBX_GO; break; }
    if (seq == DVC_MBX_SEQ_STOP) { status = DVC_MBX_STOP; break; }
    if (globaltimer_ns() - t0 >= timeout_ns) break;
    __nanosleep(100);
  }
  uint32_t batch = 0, nv = 0;
  if (status == DVC_MBX_GO) {
    const uint32_t epoch = ld_relaxed_sys_u32(&desc->epoch);
    const uint32_t kind = ld_relaxed_sys_u32(&desc->kind);
    const unsigned long long off = ld_relaxed_sys_u64(reinterpret_cast<const unsigned long long*>(&desc->src_off));
    nv = ld_relaxed_sys_u32(&desc->n_valid);
    batch = ld_relaxed_sys_u32(&desc->batch_id);
    ctl->src = kind == 1 ? local_src : arena_base + off;
    ctl->gate_now = epoch;
    ctl->gate_expect = epoch;
    ctl->flag_value = seq;
    ctl->next_seq = (seq + 1u >= DVC_MBX_SEQ_STOP) ? 1u : seq + 1u;      // 0 and 0xFFFFFFFF are never sequence numbers (the client wraps to 1 too)
  } else {
    ctl->src = local_src;
    ctl->gate_now = 0u;              // K1a / K1b skip their reads, K9 delivers nothing and does not count it as stale
    ctl->gate_expect = 0xFFFFFFFFu;
    ctl->flag_value = 0u;
  }
  ctl->status = status;
  ctl->batch_id = batch;
  ctl->n_valid = nv;
  if (hs) {
    hs->seq = seq; hs->batch_id = batch; hs->n_valid = nv;
    hs->wait_ns = globaltimer_ns() - t0;
    __threadfence_system();
    *reinterpret_cast<volatile uint32_t*>(&hs->status) = status;
  }
}

struct WaitDoneArgs {
  int n;
  const uint32_t* ptr[DVC_MBX_MAX_GROUP];
  uint32_t seq[DVC_MBX_MAX_GROUP];
};

// One thread per waited word (the words are local to the client GPU; the peers' K9 wrote them over NVLink).
__global__ void mbx_wait_done_kernel(const WaitDoneArgs a, dvc_mbx_host_status* hs, unsigned long long timeout_ns) {
  const int i = threadIdx.x;
  if (i >= a.n) return;
  const unsigned long long t0 = globaltimer_ns();
  uint32_t status = DVC_MBX_NOTHING, seq = 0;
  for (;;) {
    seq = ld_acquire_sys_u32(a.ptr[i]);
    if (seq == a.seq[i]) { status = DVC_MBX_GO; break; }
    if (globaltimer_ns() - t0 >= timeout_ns) break;
    __nanosleep(100);
  }
  if (hs) {
    hs[i].seq = seq; hs[i].batch_id = 0; hs[i].n_valid = 0;
    hs[i].wait_ns = globaltimer_ns() - t0;
    __threadfence_system();
    *reinterpret_cast<volatile uint32_t*>(&hs[i].status) = status;
  }
}

__global__ void mbx_signal_kernel(uint32_t* word, uint32_t value) {
  __threadfence_system();
  st_release_sys_u32(word, value);
}

}  // namespace dvc

using namespace dvc;

extern "C" int dvc_mbx_post(void* const* h_desc_ptrs, const dvc_mbx_desc* h_descs, int n, void* stream) {
  DVC_REQUIRE(h_desc_ptrs && h_descs && n > 0 && n <= DVC_MBX_MAX_GROUP, "mbx_post: n=%d out of range (1..%d)", n, DVC_MBX_MAX_GROUP);
  PostArgs a;
  memset(&a, 0, sizeof(a));
  a.n = n;
  for (int i = 0; i < n; ++i) {
    DVC_REQUIRE(h_desc_ptrs[i] && (reinterpret_cast<uintptr_t>(h_desc_ptrs[i]) & 7) == 0, "mbx_post: descriptor %d is null or unaligned", i);
    a.ptr[i] = reinterpret_cast<dvc_mbx_desc*>(h_desc_ptrs[i]);
    a.d[i] = h_descs[i];
  }
  mbx_post_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(a);
  count_launch();
  DVC_CHECK_CUDA(cudaGetLastError());
  return DVC_OK;
}

extern "C" int dvc_mbx_wait_desc(const dvc_mbx_desc* desc, uint32_t expect_seq, const uint8_t* arena_base, const uint8_t* local_src,
                                 dvc_mbx_lane_ctl* ctl, dvc_mbx_host_status* h_status, unsigned long long timeout_ns,
                                 void* stream) {
  DVC_REQUIRE(desc && ctl, "mbx_wait_desc: descriptor and control block are required");
  DVC_REQUIRE(expect_seq != DVC_MBX_SEQ_STOP, "mbx_wait_desc: sequence number 0xFFFFFFFF is reserved for stop");
  mbx_wait_desc_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(desc, expect_seq, arena_base, local_src, ctl, h_status, timeout_ns);
  count_launch();
  DVC_CHECK_CUDA(cudaGetLastError());
  return DVC_OK;
}

extern "C" int dvc_mbx_wait_done(void* const* h_done_ptrs, const uint32_t* h_seqs, int n, dvc_mbx_host_status* h_status,
                                 unsigned long long timeout_ns, void* stream) {
  DVC_REQUIRE(h_done_ptrs && h_seqs && n > 0 && n <= DVC_MBX_MAX_GROUP, "mbx_wait_done: n=%d out of range (1..%d)", n, DVC_MBX_MAX_GROUP);
  WaitDoneArgs a;
  memset(&a, 0, sizeof(a));
  a.n = n;
  for (int i = 0; i < n; ++i) {
    DVC_REQUIRE(h_done_ptrs[i], "mbx_wait_done: word %d is null", i);
    a.ptr[i] = reinterpret_cast<const uint32_t*>(h_done_ptrs[i]);
    a.seq[i] = h_seqs[i];
  }
  mbx_wait_done_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(a, h_status, timeout_ns);
  count_launch();
  DVC_CHECK_CUDA(cudaGetLastError());
  return DVC_OK;
}

extern "C" int dvc_mbx_dispatch(const void* h_src, void* d_dst, size_t src_bytes, void* desc_ptr, const dvc_mbx_desc* h_desc,
                                void* done_ptr, dvc_mbx_host_status* h_status, unsigned long long timeout_ns,
                                const void* d_result, void* h_result, size_t result_bytes, void* event, void* stream) {
  DVC_REQUIRE(desc_ptr && h_desc && done_ptr, "mbx_dispatch: descriptor and done word are required");
  cudaStream_t st = (cudaStream_t)stream;
  if (h_src && src_bytes) DVC_CHECK_CUDA(cudaMemcpyAsync(d_dst, h_src, src_bytes, cudaMemcpyHostToDevice, st));
  void* dp[1] = {desc_ptr};
  int rc = dvc_mbx_post(dp, h_desc, 1, stream);
  if (rc) return rc;
  void* wp[1] = {done_ptr};
  const uint32_t sq[1] = {h_desc->seq};
  rc = dvc_mbx_wait_done(wp, sq, 1, h_status, timeout_ns, stream);
  if (rc) return rc;
  if (h_result && result_bytes) DVC_CHECK_CUDA(cudaMemcpyAsync(h_result, d_result, result_bytes, cudaMemcpyDeviceToHost, st));
  if (event) DVC_CHECK_CUDA(cudaEventRecord((cudaEvent_t)event, st));
  return DVC_OK;
}

extern "C" int dvc_mbx_copy_out(const void* d_result, void* h_result, size_t bytes, void* event, void* stream) {
  DVC_REQUIRE(d_result && h_result && bytes, "mbx_copy_out: bad arguments");
  DVC_CHECK_CUDA(cudaMemcpyAsync(h_result, d_result, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  if (event) DVC_CHECK_CUDA(cudaEventRecord((cudaEvent_t)event, (cudaStream_t)stream));
  return DVC_OK;
}

extern "C" int dvc_mbx_poll(void* const* h_events, const unsigned char* h_busy, int n, unsigned char* h_done) {
  DVC_REQUIRE(h_events && h_busy && h_done && n >= 0, "mbx_poll: bad arguments");
  int found = 0;
  for (int i = 0; i < n; ++i) {
    h_done[i] = 0;
    if (!h_busy[i]) continue;
    cudaError_t e = cudaEventQuery((cudaEvent_t)h_events[i]);
    if (e == cudaSuccess) { h_done[i] = 1; ++found; }
    else if (e != cudaErrorNotReady) { DVC_CHECK_CUDA(e); }
  }
  cudaGetLastError();      // cudaErrorNotReady is not sticky, but keep the thread's last-error slot clean
  return found;
}

extern "C" int dvc_mbx_worker_enqueue(const dvc_mbx_desc* desc, uint32_t expect_seq, const uint8_t* arena_base,
                                      const uint8_t* local_src, dvc_mbx_lane_ctl* ctl, dvc_mbx_host_status* h_status,
                                      unsigned long long timeout_ns, void* graph_exec, void* event, void* stream) {
  if (desc) {
    int rc = dvc_mbx_wait_desc(desc, expect_seq, arena_base, local_src, ctl, h_status, timeout_ns, stream);
    if (rc) return rc;
  }
  if (graph_exec) DVC_CHECK_CUDA(cudaGraphLaunch((cudaGraphExec_t)graph_exec, (cudaStream_t)stream));
  if (event) DVC_CHECK_CUDA(cudaEventRecord((cudaEvent_t)event, (cudaStream_t)stream));
  return DVC_OK;
}

extern "C" int dvc_mbx_local_dispatch(void* desc_ptr, const dvc_mbx_desc* h_desc, const uint8_t* arena_base,
                                      const uint8_t* local_src, dvc_mbx_lane_ctl* ctl, void* graph_exec, void* done_ptr,
                                      dvc_mbx_host_status* h_status, unsigned long long timeout_ns, const void* d_result,
                                      void* h_result, size_t result_bytes, void* event, void* stream) {
  DVC_REQUIRE(desc_ptr && h_desc && ctl && graph_exec && done_ptr, "mbx_local_dispatch: missing argument");
  void* dp[1] = {desc_ptr};
  int rc = dvc_mbx_post(dp, h_desc, 1, stream);
  if (rc) return rc;
  rc = dvc_mbx_wait_desc(reinterpret_cast<const dvc_mbx_desc*>(desc_ptr), h_desc->seq, arena_base, local_src, ctl, nullptr,
                         1000000ull, stream);
  if (rc) return rc;
  DVC_CHECK_CUDA(cudaGraphLaunch((cudaGraphExec_t)graph_exec, (cudaStream_t)stream));
  void* wp[1] = {done_ptr};
  const uint32_t sq[1] = {h_desc->seq};
  rc = dvc_mbx_wait_done(wp, sq, 1, h_status, timeout_ns, stream);
  if (rc) return rc;
  if (h_result && result_bytes) DVC_CHECK_CUDA(cudaMemcpyAsync(h_result, d_result, result_bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  if (event) DVC_CHECK_CUDA(cudaEventRecord((cudaEvent_t)event, (cudaStream_t)stream));
  return DVC_OK;
}

extern "C" int dvc_mbx_signal(uint32_t* word, uint32_t value, void* stream) {
  DVC_REQUIRE(word, "mbx_signal: null word");
  mbx_signal_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(word, value);
  count_launch();
  DVC_CHECK_CUDA(cudaGetLastError());
  return DVC_OK;
}
