// Host-side plumbing of multi-GPU solvers (pure C++: the CPU test-suite compiles it without nvcc): the peer
// channel descriptors, the rendezvous of the shards of one multi-GPU handle and its worker threads.
#ifndef CB200_MULTI_GPU_H_
#define CB200_MULTI_GPU_H_

#include <atomic>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

struct cb200_solver;

namespace cb200 {

// One of the two sets of peer-mapped partial-sum buffers of a multi-GPU solver (cg_step.cuh): channel 0
// belongs to the main stream's sequence of sums, channel 1 to the side stream's (the reduced rhs of a
// forked solve start). Layout of a rank's buffer: 2 slots x `slot` doubles, then 2 flags 128 bytes apart.
struct PeerChannel {
  void* d_buf = nullptr;      // this rank's buffer
  size_t slot = 0;            // doubles per slot
  void* base[16] = {};        // every rank's buffer as addressable from this device (own = d_buf)
  bool ipc = false;           // peers' buffers came from cudaIpcOpenMemHandle (one process per GPU)
};

// Host rendezvous of the shard contexts of ONE multi-GPU handle (cb200_options::num_gpus > 1): the
// shards live in one process, so buffer addresses and small agreement values travel through host memory.
struct InProcGroup {
  int world = 0;
  std::mutex m;
  std::condition_variable cv;
  int arrived = 0;
  unsigned generation = 0;
  void* peer_buf[2][16] = {};
  double value[16] = {};
  char nccl_id[128] = {};
  bool aborted = false;       // a shard failed during construction: nobody waits for it any more
  bool Barrier() {
    std::unique_lock<std::mutex> lock(m);
    const unsigned gen = generation;
    if (++arrived == world) { arrived = 0; ++generation; cv.notify_all(); }
    else cv.wait(lock, [&] { return generation != gen || aborted; });
    return !aborted;
  }
  void Abort() {
    std::lock_guard<std::mutex> lock(m);
    aborted = true;
    cv.notify_all();
  }
  double Sum(int rank, double v) {  // every shard calls it; all get the same sum (rank order); -1 after Abort
    value[rank] = v;
    if (!Barrier()) return -1.0;
    double s = 0.0;
    for (int r = 0; r < world; ++r) s += value[r];
    if (!Barrier()) return -1.0;
    return s;
  }
};

// The worker threads of a multi-GPU handle: one per shard, each bound to its device. A call on the
// handle posts one task, every worker runs it on its shard, the caller waits for all of them.
struct MultiGpu {
  std::vector<cb200_solver*> shards;
  InProcGroup group;
  std::vector<std::thread> threads;
  std::mutex m;
  std::condition_variable cv_work, cv_done;
  std::atomic<unsigned> posted{0};   // tasks posted so far (workers spin on it briefly before sleeping)
  std::function<int(cb200_solver*, int)> task;
  int pending = 0;
  bool stop = false;
  std::vector<int> rc;
  std::vector<std::string> error;
};

}  // namespace cb200

#endif  // CB200_MULTI_GPU_H_
