// Test driver (host only): invariants of csrc/explicit_plan.h on a random <2,3,9> structure with
// unsorted cameras, duplicate (point, camera) rows, empty points and points of more than kFormRows rows.
// Prints key=value pairs; every check is a count of violations (0 = holds).
#include <cstdio>
#include <map>
#include <random>
#include <set>

#include "../../ceres-solver-ceres-solver_b200/csrc/explicit_plan.h"

using namespace cb200;

int main() {
  const int C = 37, P = 400;
  std::mt19937 rng(11);
  FlatStructure flat;
  flat.num_points = P; flat.num_cameras = C;
  flat.point_row_start.push_back(0);
  for (int p = 0; p < P; ++p) {
    int k = (p % 50 == 7) ? 0 : (p % 97 == 3) ? 40 : 1 + (int)(rng() % 6);
    for (int i = 0; i < k; ++i) flat.cam_idx.push_back((int)(rng() % C));  // unsorted, duplicates possible
    flat.point_row_start.push_back((int)flat.cam_idx.size());
  }
  flat.num_obs = (int)flat.cam_idx.size();
  const ExplicitStructure es = PlanExplicitStructure(flat);
  const ExplicitFormPlan fp = PlanExplicitForm(es, flat.point_row_start.data(), flat.cam_idx.data(), P);

  // expected block pairs, the way SparseSchurComplementSolver::InitStorage collects them
  std::set<std::pair<int, int>> want;
  for (int c = 0; c < C; ++c) want.emplace(c, c);
  for (int p = 0; p < P; ++p) {
    std::set<int> cams(flat.cam_idx.begin() + flat.point_row_start[p], flat.cam_idx.begin() + flat.point_row_start[p + 1]);
    for (int a : cams) for (int b : cams) if (a < b) want.emplace(a, b);
  }
  int bad_structure = (int)want.size() != (int)es.col.size();
  {
    auto it = want.begin();
    for (int i = 0; i < C && !bad_structure; ++i)
      for (int k = es.row_ptr[i]; k < es.row_ptr[i + 1]; ++k, ++it)
        if (it->first != i || it->second != es.col[k]) { bad_structure = 1; break; }
  }
  // column lists: every strictly upper cell exactly once, under its column, rows ascending
  int bad_transpose = 0;
  std::vector<int> seen(es.col.size(), 0);
  for (int j = 0; j < C; ++j)
    for (int t = es.t_ptr[j]; t < es.t_ptr[j + 1]; ++t) {
      const int k = es.t_cell[t], i = es.t_row[t];
      if (es.col[k] != j || k < es.row_ptr[i] || k >= es.row_ptr[i + 1] || i >= j) ++bad_transpose;
      if (t > es.t_ptr[j] && es.t_row[t - 1] >= i) ++bad_transpose;
      ++seen[k];
    }
  for (int i = 0; i < C; ++i)
    for (int k = es.row_ptr[i]; k < es.row_ptr[i + 1]; ++k)
      if (seen[k] != (es.col[k] != i ? 1 : 0)) ++bad_transpose;
  // pair destinations
  int bad_pairs = 0, bad_order = 0, dup_pairs = 0;
  long long pairs = 0;
  for (int p = 0; p < P; ++p) {
    const int r0 = flat.point_row_start[p], k = flat.point_row_start[p + 1] - r0;
    std::multiset<int> rows(fp.srow.begin() + r0, fp.srow.begin() + r0 + k);
    for (int u = 0; u < k; ++u) if (rows.count(r0 + u) != 1) ++bad_order;     // a permutation of the point's rows
    for (int u = 0; u + 1 < k; ++u) if (flat.cam_idx[fp.srow[r0 + u]] > flat.cam_idx[fp.srow[r0 + u + 1]]) ++bad_order;
    long long at = fp.pair_start[p];
    for (int u = 0; u < k; ++u)
      for (int v = u; v < k; ++v, ++at, ++pairs) {
        const unsigned int cell = fp.pair_cell[at];
        const int ci = flat.cam_idx[fp.srow[r0 + u]], cj = flat.cam_idx[fp.srow[r0 + v]];
        const int kk = (int)(cell & ~kFormDuplicate);
        if (kk < es.row_ptr[ci] || kk >= es.row_ptr[ci + 1] || es.col[kk] != cj) ++bad_pairs;
        const bool dup = (cell & kFormDuplicate) != 0;
        if (dup != (v > u && ci == cj)) ++bad_pairs;
        dup_pairs += dup;
      }
    if (at != fp.pair_start[p + 1]) ++bad_pairs;
  }
  // items: every (point, u0) chunk exactly once
  std::map<std::pair<int, int>, int> item_count;
  for (const FormItem& it : fp.items) ++item_count[{it.point, it.u0}];
  int bad_items = 0;
  long long expected_items = 0;
  for (int p = 0; p < P; ++p)
    for (int u0 = 0; u0 < flat.point_row_start[p + 1] - flat.point_row_start[p]; u0 += kFormRows, ++expected_items)
      if (item_count[{p, u0}] != 1) ++bad_items;
  if ((long long)fp.items.size() != expected_items) ++bad_items;
  std::printf("cells=%d pairs=%lld dup_pairs=%d items=%d bad_structure=%d bad_transpose=%d bad_pairs=%d bad_order=%d bad_items=%d\n",
              (int)es.col.size(), pairs, dup_pairs, (int)fp.items.size(), bad_structure, bad_transpose, bad_pairs,
              bad_order, bad_items);
  return 0;
}
