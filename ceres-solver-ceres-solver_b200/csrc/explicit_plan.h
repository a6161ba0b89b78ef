// Host-side plan of the explicit reduced camera matrix (use_explicit_schur_complement):
// the block structure of S and, per point shard, where every outer product lands.
// Pure C++ (no CUDA) so the CPU test-suite can exercise it.
//
// Replaces, for the <2,3,9> bundle-adjustment case only:
//   SparseSchurComplementSolver::InitStorage  internal/ceres/schur_complement_solver.cc:224-290
//   BlockRandomAccessSparseMatrix ctor/layout internal/ceres/block_random_access_sparse_matrix.cc:47-103
//   Chunk::buffer_layout (merging of duplicate (point, camera) rows)
//                                             internal/ceres/schur_eliminator_impl.h:140-148
#ifndef CB200_EXPLICIT_PLAN_H_
#define CB200_EXPLICIT_PLAN_H_

#include <algorithm>
#include <cstdint>
#include <thread>
#include <vector>

#include "flatten.h"

namespace cb200 {

// fn(thread, begin, end) over [0, n) cut into contiguous ranges, one host thread each (planning is
// once per solver lifetime, but 2 x 21 M pairs at Venice shape are seconds on one core).
template <typename Fn>
inline void ParallelRanges(int n, Fn fn) {
  const int nt = std::max(1, std::min<int>({(int)std::thread::hardware_concurrency(), 16, n / 1024 + 1}));
  if (nt == 1) { fn(0, 0, n); return; }
  std::vector<std::thread> pool;
  for (int t = 0; t < nt; ++t)
    pool.emplace_back(fn, t, (int)((long long)n * t / nt), (int)((long long)n * (t + 1) / nt));
  for (auto& th : pool) th.join();
}

constexpr int kFormRows = 16;  // rows of one point whose outer products one warp item owns

// Block CRS of the upper triangle of S (diagonal cell first in every row, columns
// ascending: the order std::set<std::pair<int,int>> gives the reference), cell k at
// values + 81 k, plus the strictly-upper cells listed by COLUMN so that S x can gather
// the lower triangle from the same storage without atomics.
struct ExplicitStructure {
  int num_cameras = 0;
  std::vector<int> row_ptr;   // C + 1
  std::vector<int> col;       // num_cells
  std::vector<int> t_ptr;     // C + 1: cells (i, j), i < j, grouped by j
  std::vector<int> t_cell;    // cell index k
  std::vector<int> t_row;     // its row block i
};

// The block pairs of every point's cameras, from the GLOBAL structure (every rank builds
// the same matrix; the values are summed over the ranks).
inline ExplicitStructure PlanExplicitStructure(const FlatStructure& flat) {
  const int C = flat.num_cameras, P = flat.num_points;
  ExplicitStructure es;
  es.num_cameras = C;
  // rows of every camera (camera-major), to visit the points a camera sees
  std::vector<int> cam_start(C + 1, 0);
  for (int r = 0; r < flat.num_obs; ++r) ++cam_start[flat.cam_idx[r] + 1];
  for (int c = 0; c < C; ++c) cam_start[c + 1] += cam_start[c];
  std::vector<int> cam_rows(flat.num_obs), cursor(cam_start.begin(), cam_start.end() - 1);
  std::vector<int> row_point(flat.num_obs);
  for (int p = 0; p < P; ++p)
    for (int r = flat.point_row_start[p]; r < flat.point_row_start[p + 1]; ++r) row_point[r] = p;
  for (int r = 0; r < flat.num_obs; ++r) cam_rows[cursor[flat.cam_idx[r]]++] = r;
  // columns of every block row (independent per row: camera ranges on host threads)
  std::vector<std::vector<int>> cols_of(C);
  ParallelRanges(C, [&](int, int c_begin, int c_end) {
    std::vector<int> mark(C, -1);
    for (int i = c_begin; i < c_end; ++i) {
      std::vector<int>& row_cols = cols_of[i];
      mark[i] = i;
      row_cols.push_back(i);  // the diagonal pair always exists (:233-235)
      for (int k = cam_start[i]; k < cam_start[i + 1]; ++k) {
        const int p = row_point[cam_rows[k]];
        for (int r = flat.point_row_start[p]; r < flat.point_row_start[p + 1]; ++r) {
          const int j = flat.cam_idx[r];
          if (j > i && mark[j] != i) { mark[j] = i; row_cols.push_back(j); }
        }
      }
      std::sort(row_cols.begin(), row_cols.end());
    }
  });
  es.row_ptr.assign(C + 1, 0);
  for (int i = 0; i < C; ++i) {
    es.col.insert(es.col.end(), cols_of[i].begin(), cols_of[i].end());
    es.row_ptr[i + 1] = (int)es.col.size();
  }
  es.t_ptr.assign(C + 1, 0);
  for (int i = 0; i < C; ++i)
    for (int k = es.row_ptr[i] + 1; k < es.row_ptr[i + 1]; ++k) ++es.t_ptr[es.col[k] + 1];
  for (int c = 0; c < C; ++c) es.t_ptr[c + 1] += es.t_ptr[c];
  es.t_cell.resize(es.t_ptr[C]);
  es.t_row.resize(es.t_ptr[C]);
  std::vector<int> tc(es.t_ptr.begin(), es.t_ptr.end() - 1);
  for (int i = 0; i < C; ++i)
    for (int k = es.row_ptr[i] + 1; k < es.row_ptr[i + 1]; ++k) {
      const int slot = tc[es.col[k]]++;
      es.t_cell[slot] = k;
      es.t_row[slot] = i;  // ascending i within a column: the loop visits rows in order
    }
  return es;
}

// Work of one shard's outer products. The rows of a point are visited in camera order
// (srow), so that a pair (u <= v) always lands in the stored (upper) cell
// (cam[u], cam[v]) untransposed; pairs of DIFFERENT rows with the SAME camera
// (duplicate observations) land in the diagonal cell together with their transpose
// (bit 31 of the destination).
struct FormItem { int point, u0; };  // rows [u0, u0 + kFormRows) of the point's sorted rows x all v >= u
constexpr unsigned int kFormDuplicate = 0x80000000u;

struct ExplicitFormPlan {
  std::vector<int> srow;                // N_loc: local row ids, camera-sorted within each point
  std::vector<long long> pair_start;    // P_loc + 1: first destination of each point's pairs
  std::vector<unsigned int> pair_cell;  // per point, u-major over u <= v: cell index | kFormDuplicate
  std::vector<FormItem> items;
};

inline ExplicitFormPlan PlanExplicitForm(const ExplicitStructure& es, const int* row_start /* local */,
                                         const int* cam /* local rows */, int P_loc) {
  ExplicitFormPlan plan;
  const int N_loc = row_start[P_loc];
  plan.srow.resize(N_loc);
  plan.pair_start.assign(P_loc + 1, 0);
  for (int p = 0; p < P_loc; ++p) {
    const long long k = row_start[p + 1] - row_start[p];
    plan.pair_start[p + 1] = plan.pair_start[p] + k * (k + 1) / 2;
  }
  plan.pair_cell.resize((size_t)plan.pair_start[P_loc]);
  ParallelRanges(P_loc, [&](int, int p_begin, int p_end) {
    for (int p = p_begin; p < p_end; ++p) {
      const int r0 = row_start[p], k = row_start[p + 1] - r0;
      for (int u = 0; u < k; ++u) plan.srow[r0 + u] = r0 + u;
      std::stable_sort(plan.srow.begin() + r0, plan.srow.begin() + r0 + k,
                       [&](int a, int b) { return cam[a] < cam[b]; });
      size_t out = (size_t)plan.pair_start[p];
      for (int u = 0; u < k; ++u) {
        const int ci = cam[plan.srow[r0 + u]];
        const int* first = es.col.data() + es.row_ptr[ci];
        const int* last = es.col.data() + es.row_ptr[ci + 1];
        for (int v = u; v < k; ++v) {
          const int cj = cam[plan.srow[r0 + v]];
          first = std::lower_bound(first, last, cj);  // cj ascends with v: resume where the last search ended
          unsigned int cell = (unsigned int)(first - es.col.data());
          if (v > u && cj == ci) cell |= kFormDuplicate;
          plan.pair_cell[out++] = cell;
        }
      }
    }
  });
  for (int p = 0; p < P_loc; ++p)
    for (int u0 = 0; u0 < row_start[p + 1] - row_start[p]; u0 += kFormRows) plan.items.push_back(FormItem{p, u0});
  // Order of execution. The cells a point touches lie in the block rows of its cameras, and S is 2-3x
  // the L2 at Ladybug / Venice shape: in point order (random home cameras) every reduction into it
  // missed (11 GB of DRAM traffic for 1.3 GB of data at Venice). So items are grouped by their point's
  // smallest camera into camera ranges whose rows of S fit a quarter of the L2, and inside a group
  // they are dealt round-robin over the group's cameras: warps that run at the same time hit the
  // same resident rows, but not the same cells (same-address reductions serialise).
  {
    const int C = es.num_cameras;
    auto key = [&](const FormItem& it) { return cam[plan.srow[row_start[it.point]]]; };
    std::vector<int> bucket_start(C + 1, 0);
    for (const FormItem& it : plan.items) ++bucket_start[key(it) + 1];
    for (int c = 0; c < C; ++c) bucket_start[c + 1] += bucket_start[c];
    std::vector<FormItem> by_camera(plan.items.size());
    {
      std::vector<int> cursor(bucket_start.begin(), bucket_start.end() - 1);
      for (const FormItem& it : plan.items) by_camera[cursor[key(it)]++] = it;
    }
    constexpr long long kGroupBytes = 32LL << 20;
    size_t out = 0;
    for (int c0 = 0; c0 < C;) {
      int c1 = c0;
      long long bytes = 0;
      while (c1 < C && (c1 == c0 || bytes < kGroupBytes)) {
        bytes += 648LL * (es.row_ptr[c1 + 1] - es.row_ptr[c1]);
        ++c1;
      }
      bool any = true;
      for (int round = 0; any; ++round) {
        any = false;
        for (int c = c0; c < c1; ++c) {
          const int idx = bucket_start[c] + round;
          if (idx < bucket_start[c + 1]) { plan.items[out++] = by_camera[idx]; any = true; }
        }
      }
      c0 = c1;
    }
  }
  return plan;
}

}  // namespace cb200

#endif  // CB200_EXPLICIT_PLAN_H_
