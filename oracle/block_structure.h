// ORACLE — TEST INFRASTRUCTURE ONLY (see small_blas.h header).
//
// Flat restatement of the reference's block-sparse container:
//   internal/ceres/block_structure.h:52-96      Block / Cell / CompressedList
//   internal/ceres/block_sparse_matrix.cc:784-808  CreateTranspose
//   internal/ceres/block_sparse_matrix.cc:58-67    cumulative_nnz
//   internal/ceres/partition_range_for_parallel_for.h  nnz-balanced partitions
//   internal/ceres/parallel_for.h:69-183 / parallel_invoke.h:150-260 ParallelFor
#ifndef ORACLE_BLOCK_STRUCTURE_H_
#define ORACLE_BLOCK_STRUCTURE_H_

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace oracle {

struct Cell { int block_id; int position; };
struct Block { int size; int position; };
struct CompressedRow {
  Block block;
  std::vector<Cell> cells;
  int nnz = 0;
  int cumulative_nnz = 0;
};

struct BlockStructure {
  std::vector<Block> cols;
  std::vector<CompressedRow> rows;
};

// Transpose structure: rows of the result are the column blocks of bs; each
// cell keeps the ORIGINAL value position (block_sparse_matrix.cc:784-808).
inline BlockStructure CreateTranspose(const BlockStructure& bs) {
  BlockStructure t;
  t.rows.resize(bs.cols.size());
  for (size_t i = 0; i < bs.cols.size(); ++i) {
    t.rows[i].block = bs.cols[i];
    t.rows[i].nnz = 0;
  }
  t.cols.resize(bs.rows.size());
  for (size_t i = 0; i < bs.rows.size(); ++i) {
    const CompressedRow& row = bs.rows[i];
    t.cols[i] = row.block;
    for (const Cell& c : row.cells) {
      t.rows[c.block_id].cells.push_back(Cell{static_cast<int>(i), c.position});
      t.rows[c.block_id].nnz += row.block.size * bs.cols[c.block_id].size;
    }
  }
  int cum = 0;
  for (auto& r : t.rows) { cum += r.nnz * 1; r.cumulative_nnz = cum; }
  return t;
}

// Split [start,end) into at most max_partitions contiguous ranges minimising
// the maximal cumulative cost (partition_range_for_parallel_for.h). Binary
// search on the admissible max cost, greedy feasibility check.
template <typename CumCost>
std::vector<int> PartitionRange(int start, int end, int max_partitions,
                                CumCost cum) {
  auto cost_before = [&](int i) { return i == 0 ? 0LL : (long long)cum(i - 1); };
  auto try_cost = [&](long long max_cost, std::vector<int>* out) {
    out->clear();
    out->push_back(start);
    int cur = start;
    while (cur < end) {
      const long long base = cost_before(cur);
      // largest nxt with cum(nxt-1) - base <= max_cost
      int lo = cur, hi = end;
      while (lo < hi) {
        int mid = (lo + hi + 1) / 2;
        if ((long long)cum(mid - 1) - base <= max_cost) lo = mid; else hi = mid - 1;
      }
      if (lo == cur) return false;
      cur = lo;
      out->push_back(cur);
      if ((int)out->size() - 1 > max_partitions) return false;
    }
    return (int)out->size() - 1 <= max_partitions;
  };
  long long lo = 0, hi = (long long)cum(end - 1) - cost_before(start);
  std::vector<int> best, tmp;
  try_cost(hi, &best);
  while (lo < hi) {
    long long mid = lo + (hi - lo) / 2;
    if (try_cost(mid, &tmp)) { hi = mid; best = tmp; } else { lo = mid + 1; }
  }
  try_cost(hi, &best);
  return best;
}

// Persistent worker pool (stands in for ContextImpl::thread_pool,
// context_impl.cc:211-213 / thread_pool.cc) so that a ParallelFor does not pay
// thread creation: workers sleep on a condition variable between jobs.
class Pool {
 public:
  static Pool& Get() { static Pool p; return p; }
  // Runs job(tid) on tids 1..n-1 in workers and tid 0 on the caller.
  void Run(int n, const std::function<void(int)>& job) {
    Ensure(n - 1);
    {
      std::lock_guard<std::mutex> l(mu_);
      job_ = &job; want_ = n - 1; pending_ = n - 1; ++epoch_;
    }
    cv_.notify_all();
    job(0);
    std::unique_lock<std::mutex> l(mu_);
    done_.wait(l, [&] { return pending_ == 0; });
    job_ = nullptr;
  }
  ~Pool() {
    { std::lock_guard<std::mutex> l(mu_); stop_ = true; }
    cv_.notify_all();
    for (auto& t : th_) t.join();
  }
 private:
  void Ensure(int n) {
    while ((int)th_.size() < n) {
      const int id = (int)th_.size() + 1;
      th_.emplace_back([this, id] {
        unsigned long long seen = 0;
        for (;;) {
          const std::function<void(int)>* j = nullptr;
          {
            std::unique_lock<std::mutex> l(mu_);
            cv_.wait(l, [&] { return stop_ || (epoch_ != seen && id <= want_); });
            if (stop_) return;
            seen = epoch_;
            j = job_;
          }
          (*j)(id);
          {
            std::lock_guard<std::mutex> l(mu_);
            if (--pending_ == 0) done_.notify_one();
          }
        }
      });
    }
  }
  std::mutex mu_;
  std::condition_variable cv_, done_;
  std::vector<std::thread> th_;
  const std::function<void(int)>* job_ = nullptr;
  int want_ = 0, pending_ = 0;
  unsigned long long epoch_ = 0;
  bool stop_ = false;
};

// ParallelFor with the reference's work-block scheme (at most 4 blocks per
// thread, sizes differing by at most one, claimed through an atomic counter;
// parallel_invoke.h:150-260). f(thread_id, begin, end).
inline void ParallelForRange(int start, int end, int num_threads,
                             const std::function<void(int, int, int)>& f,
                             int min_block_size = 1) {
  if (start >= end) return;
  if (num_threads <= 1 || end - start < 2 * min_block_size) {
    f(0, start, end);
    return;
  }
  const int num_blocks =
      std::max(1, std::min((end - start) / min_block_size, num_threads * 4));
  const int base = (end - start) / num_blocks;
  const int p1 = (end - start) % num_blocks;
  std::atomic<int> next{0};
  auto worker = [&](int tid) {
    for (;;) {
      int b = next.fetch_add(1);
      if (b >= num_blocks) break;
      int s = start + b * base + std::min(b, p1);
      int e = s + base + (b < p1 ? 1 : 0);
      f(tid, s, e);
    }
  };
  const int nt = std::min(num_threads, num_blocks);
  Pool::Get().Run(nt, worker);
}

// ParallelFor over user partitions (parallel_for.h:104-137).
inline void ParallelForPartitions(int start, int end, int num_threads,
                                  const std::vector<int>& partitions,
                                  const std::function<void(int, int, int)>& f) {
  if (start >= end) return;
  if (num_threads <= 1 || end - start <= num_threads || partitions.size() < 2) {
    ParallelForRange(start, end, num_threads, f);
    return;
  }
  const int np = (int)partitions.size() - 1;
  ParallelForRange(0, np, num_threads, [&](int tid, int ps, int pe) {
    f(tid, partitions[ps], partitions[pe]);
  });
}

}  // namespace oracle

#endif  // ORACLE_BLOCK_STRUCTURE_H_
