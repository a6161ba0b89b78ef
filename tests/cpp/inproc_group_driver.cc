// Test driver (host only): the rendezvous of the shards of one multi-GPU handle (InProcGroup,
// csrc/multi_gpu.h) under real threads: barriers in lock step, rank-ordered sums identical on every
// shard, and Abort releasing everybody who waits for a shard that failed during construction.
#include <cstdio>
#include <thread>
#include <vector>

#include <chrono>

#include "../../ceres-solver-ceres-solver_b200/csrc/multi_gpu.h"

int main() {
  const int world = 6;
  cb200::InProcGroup g;
  g.world = world;
  std::vector<double> sums(world * 50, 0.0);
  std::vector<int> order_errors(world, 0);
  std::vector<int> stage(world, 0);
  std::vector<std::thread> th;
  for (int r = 0; r < world; ++r)
    th.emplace_back([&, r] {
      for (int round = 0; round < 50; ++round) {
        stage[r] = round;
        if (!g.Barrier()) { ++order_errors[r]; return; }
        for (int q = 0; q < world; ++q)
          if (stage[q] < round) ++order_errors[r];  // nobody passes a barrier before everybody reached it
        sums[r * 50 + round] = g.Sum(r, 0.1 * (r + 1) + round);
      }
    });
  for (auto& t : th) t.join();
  int bad = 0;
  for (int r = 0; r < world; ++r) bad += order_errors[r];
  int differing = 0;
  for (int round = 0; round < 50; ++round)
    for (int r = 1; r < world; ++r)
      if (sums[r * 50 + round] != sums[round]) ++differing;  // bit-identical: same order on every shard
  double want = 0.0;
  for (int r = 0; r < world; ++r) want += 0.1 * (r + 1) + 7;
  const bool value_ok = sums[7] == want;
  // Abort: world - 1 shards wait, the last one fails instead of arriving
  cb200::InProcGroup a;
  a.world = 3;
  int released = 0;
  std::vector<std::thread> waiters;
  std::mutex m;
  for (int r = 0; r < 2; ++r)
    waiters.emplace_back([&] {
      const bool ok = a.Barrier();
      const double s = a.Sum(0, 1.0);
      std::lock_guard<std::mutex> lock(m);
      if (!ok && s < 0.0) ++released;
    });
  std::this_thread::sleep_for(std::chrono::milliseconds(50));
  a.Abort();
  for (auto& t : waiters) t.join();
  std::printf("order_errors=%d differing=%d value_ok=%d released=%d\n", bad, differing, value_ok ? 1 : 0, released);
  return 0;
}
