// Host-only check of UploadGroup (csrc/scene_store.h): the barrier the device threads of rtb_render_multi meet at, and
// that a member which fails releases the others instead of leaving them waiting.  No CUDA call is made.
#include <atomic>
#include <chrono>
#include <cstdio>
#include <thread>
#include <vector>
#include <cuda_runtime_api.h>
#include "rtb200.h"
#include "rtb_internal.h"
#include "scene_store.h"

using rtb::UploadGroup;

static int run(int n, int failing, int rounds) {
  UploadGroup g;
  g.n = n;
  std::atomic<int> passed{0}, released{0};
  std::vector<std::thread> th;
  for (int m = 0; m < n; ++m)
    th.emplace_back([&, m] {
      for (int r = 0; r < rounds; ++r) {
        if (m == failing && r == rounds - 1) { std::this_thread::sleep_for(std::chrono::milliseconds(20)); g.fail(); return; }
        if (!g.barrier()) { ++released; return; }
        ++passed;
      }
    });
  for (auto &t : th) t.join();
  const int expect_pass = failing < 0 ? n * rounds : n * (rounds - 1);
  const int expect_rel = failing < 0 ? 0 : n - 1;
  if (passed != expect_pass || released != expect_rel) {
    std::printf("n=%d failing=%d rounds=%d: passed %d (want %d), released %d (want %d)\n", n, failing, rounds, passed.load(), expect_pass,
                released.load(), expect_rel);
    return 1;
  }
  return 0;
}

int main() {
  int bad = 0;
  for (int n : {1, 2, 4, 8}) {
    bad += run(n, -1, 50);                 // a reusable barrier: 50 rounds, everybody passes every round
    if (n > 1) bad += run(n, n - 1, 3);    // the last member fails before the third barrier: the others are released, not stuck
  }
  // after a failure every later barrier call returns false at once
  UploadGroup g;
  g.n = 2;
  g.fail();
  if (g.barrier()) { std::printf("barrier after fail() returned true\n"); ++bad; }
  std::printf(bad ? "FAILED\n" : "OK\n");
  return bad ? 1 : 0;
}
