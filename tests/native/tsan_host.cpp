// ThreadSanitizer harness for the host side of libdvc_b200 (the reference shares plain Python lists
// across threads without locks, server.py:20-24, worker.py:33,37-38; here the native host runtime is checked with TSAN).
//
// Runs WITHOUT a GPU: every entry point below is pure host code up to the point where it would touch the driver, and
// that point fails cleanly (no driver -> DVC_ERR_CUDA with a per-thread message).  Engine threads of different GPUs call
// these concurrently in one process, so the shared pieces - the lazily resolved driver entry points, the per-device
// attribute / SM-count caches, the launch counter, the thread-local error text - are hammered from several threads.
//
//   build + run: python -m distributedvolunteercomputing_b200.ops.build --tsan     (or tests/test_tsan_host.py)
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "dvc_b200.h"

static std::atomic<int> g_fail{0};
#define EXPECT(c, ...)                                  \
  do {                                                  \
    if (!(c)) {                                         \
      fprintf(stderr, "tsan_host: " __VA_ARGS__);       \
      fprintf(stderr, "\n");                            \
      g_fail.fetch_add(1);                              \
    }                                                   \
  } while (0)

static void worker(int tid, int iters) {
  // aligned fake device addresses: never dereferenced on the host
  uint8_t* base = reinterpret_cast<uint8_t*>(uintptr_t(0x7f0000000000ull) + (uintptr_t)tid * 0x10000000ull);
  float mean[3] = {127.5f, 0.f, 0.f};
  dvc_det_params dp;
  memset(&dp, 0, sizeof(dp));
  for (int it = 0; it < iters; ++it) {
    dvc_plan* p = dvc_plan_create();
    EXPECT(p != nullptr, "plan_create failed");
    int rc = dvc_plan_add_preproc_conv0(p, base, 480000, nullptr, 8, 400, 400, (const int*)base, (const int*)base, (const int*)base,
                                        (const int*)base, (const float*)base, (const float*)base, mean, 0.007843f, base + 4096, 0);
    EXPECT(rc == DVC_OK, "add_preproc_conv0 rc=%d (%s)", rc, dvc_last_error());
    rc = dvc_plan_add_im2col(p, base, 8, 10, 10, 256, base + 8192);
    EXPECT(rc == DVC_OK, "add_im2col rc=%d (%s)", rc, dvc_last_error());
    EXPECT(dvc_plan_num_ops(p) == 2, "num_ops %d", dvc_plan_num_ops(p));
    // ops that need a TMA descriptor: without a driver they must fail cleanly, leave the plan as it was, and leave the
    // reason in THIS thread's error text
    rc = dvc_plan_add_gemm(p, base, base + 65536, (const float*)base, 1000 + tid, 128, 64, 64, 0, base + 131072, 128, 1, nullptr, 0);
    if (rc != DVC_OK) EXPECT(strstr(dvc_last_error(), "cuTensorMapEncode") != nullptr, "gemm error text: %s", dvc_last_error());
    rc = dvc_plan_add_dwconv(p, base, 8, 19, 19, 64, 1, (const float*)base, (const float*)base, 1, base + 65536);
    if (rc != DVC_OK) EXPECT(dvc_last_error()[0] != 0, "dwconv failed without a message");
    rc = dvc_plan_add_conv3x3s2(p, base, 8, 10, 10, 256, base + 65536, (const float*)base, 512, base + 131072, 512, 1, 0);
    if (rc != DVC_OK) EXPECT(strstr(dvc_last_error(), "cuTensorMapEncodeIm2col") != nullptr, "conv3x3s2 error text: %s", dvc_last_error());
    // argument errors carry this thread's own numbers
    rc = dvc_plan_add_gemm(p, base, base + 65536, (const float*)base, 100, 24 + 1, 64, 64, 0, base + 131072, 128, 1, nullptr, 0);
    EXPECT(rc != DVC_OK && strstr(dvc_last_error(), "N=25") != nullptr, "bad-N error text: %s", dvc_last_error());
    char want[32];
    snprintf(want, sizeof(want), "K=%d", 17 + 16 * tid);
    rc = dvc_gemm_f16(base, base + 65536, (const float*)base, 100, 32, 17 + 16 * tid, 64, 0, base + 131072, 32, 1, nullptr, 0, nullptr);
    EXPECT(rc != DVC_OK && strstr(dvc_last_error(), want) != nullptr, "bad-K error text: %s (want %s)", dvc_last_error(), want);
    (void)dvc_plan_set_head_branch(p, it & 1);
    (void)dvc_plan_set_flag_value(p, (uint32_t)it);
    // launching needs a device: must fail cleanly too (and never count a launch that did not happen twice)
    rc = dvc_plan_run(p, nullptr);
    if (rc == DVC_OK) EXPECT(false, "plan_run succeeded without a GPU?");
    (void)dvc_set_device(0);
    (void)dvc_launch_count();
    dvc_plan_destroy(p);
  }
}

static int g_plain_counter;            // --selftest only: written by two threads without synchronisation

int main(int argc, char** argv) {
  if (argc > 1 && strcmp(argv[1], "--selftest") == 0) {
    // proves the sanitizer is live in this build: TSAN must report this race (exit code from TSAN_OPTIONS=exitcode=..)
    std::thread a([] { for (int i = 0; i < 100000; ++i) g_plain_counter++; });
    std::thread b([] { for (int i = 0; i < 100000; ++i) g_plain_counter++; });
    a.join();
    b.join();
    printf("selftest counter %d\n", g_plain_counter);
    return 0;
  }
  const int threads = argc > 1 ? atoi(argv[1]) : 8;
  const int iters = argc > 2 ? atoi(argv[2]) : 200;
  EXPECT(dvc_version() > 0, "version");
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; ++t) pool.emplace_back(worker, t, iters);
  for (auto& t : pool) t.join();
  if (g_fail.load()) {
    fprintf(stderr, "tsan_host: %d expectation(s) failed\n", g_fail.load());
    return 1;
  }
  printf("tsan_host ok: %d threads x %d iterations\n", threads, iters);
  return 0;
}
