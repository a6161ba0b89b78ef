// Host-side flattening of Ceres' CompressedRowBlockStructure into the flat
// arrays the device kernels read, plus the point sharding and the tile plan.
// Pure C++ (no CUDA) so the CPU test-suite can exercise it.
//
// Replaces, for the <2,3,9> bundle-adjustment case only:
//   DetectStructure                       internal/ceres/detect_structure.cc:39-120
//   PartitionedMatrixView ctor            internal/ceres/partitioned_matrix_view_impl.h:47-105
//   SchurEliminator::Init chunk discovery internal/ceres/schur_eliminator_impl.h:80-174
// and the array-of-vectors container of internal/ceres/block_structure.h:52-96 by
//   point_row_start[P+1], cam_idx[N], e_pos[N], f_pos[N]   (all int32).
#ifndef CB200_FLATTEN_H_
#define CB200_FLATTEN_H_

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/ceres_b200_schur.h"

namespace cb200 {

constexpr int kTileObs = 32;  // rows per tile of the fused per-point kernels: one lane each
constexpr int kTilePoints = 16;  // points (= (E'E)^-1 blocks staged) per tile; real tracks have >= 2 rows
// doubles per row of the Schur-Jacobi weights array (w00, w01, w11 [, pad]). 4 would give the camera-major gather of
// camera_block_kernel one aligned 32-byte sector per row instead of 1.5 on average; measured (round 2, Venice shape,
// same box): camera blocks 250.0 -> 250.9 us, E'E / rhs pass 283.5 -> 285.9 us (8 more bytes written per row). The
// gather is not limited by those bytes; 3 stays.
#ifndef CB200_WEIGHT_STRIDE
#define CB200_WEIGHT_STRIDE 3
#endif
constexpr int kWeightStride = CB200_WEIGHT_STRIDE;
constexpr int kChunkedPointRows = 256;  // points of 33..256 rows run as chunk tiles inside the stream kernel
// flags in TileMeta::np above the point count (low 8 bits)
constexpr int kTileKindShift = 8;        // bits 8-9: 0 whole points, 1 first-pass chunk, 2 second-pass chunk
constexpr int kTileLastChunk = 1 << 10;  // last first-pass chunk of its point: the point's sums are complete
constexpr int kTileContinues = 1 << 11;  // not the first tile of its group: warp ranges may not start here

struct FlatStructure {
  int num_points = 0;      // P: eliminated (E) column blocks, size 3
  int num_cameras = 0;     // C: remaining (F) column blocks, size 9
  int num_obs = 0;         // N: row blocks, size 2, one E cell + one F cell
  std::vector<int> point_row_start;  // P+1
  std::vector<int> cam_idx;          // N, camera index in [0, C)
  std::vector<int> e_pos, f_pos;     // N each, value positions
  bool canonical_layout = false;     // e_pos[r] == 6r and f_pos[r] == 6N + 18r
};

// Returns CB200_OK or an error code with a message.
inline int Flatten(const cb200_structure& s, int num_eliminate_blocks,
                   FlatStructure* out, std::string* err) {
  auto fail = [&](int code, const std::string& m) { *err = m; return code; };
  if (s.num_row_blocks < 0 || s.num_col_blocks <= 0 || !s.col_block_size ||
      !s.col_block_pos || (s.num_row_blocks > 0 && (!s.row_block_size || !s.row_block_pos ||
      !s.cell_ptr || !s.cell_block_id || !s.cell_position)))
    return fail(CB200_ERR_INVALID_ARGUMENT, "null or empty block structure");
  const int P = num_eliminate_blocks;
  if (P < 0 || P > s.num_col_blocks)
    return fail(CB200_ERR_INVALID_ARGUMENT, "num_eliminate_blocks out of range");
  const int C = s.num_col_blocks - P;
  // column blocks: P blocks of 3 then C blocks of 9, densely packed
  int pos = 0;
  for (int c = 0; c < s.num_col_blocks; ++c) {
    const int want = c < P ? 3 : 9;
    if (s.col_block_size[c] != want)
      return fail(CB200_ERR_UNSUPPORTED_STRUCTURE,
                  "column block " + std::to_string(c) + " has size " +
                      std::to_string(s.col_block_size[c]) + ", expected " + std::to_string(want) +
                      " (only the <2,3,9> structure is supported)");
    if (s.col_block_pos[c] != pos)
      return fail(CB200_ERR_UNSUPPORTED_STRUCTURE, "column blocks are not densely packed");
    pos += want;
  }
  const int N = s.num_row_blocks;
  out->num_points = P;
  out->num_cameras = C;
  out->num_obs = N;
  out->point_row_start.assign(P + 1, 0);
  out->cam_idx.resize(N);
  out->e_pos.resize(N);
  out->f_pos.resize(N);
  int last_point = -1;
  bool canonical = true;
  for (int r = 0; r < N; ++r) {
    if (s.row_block_size[r] != 2 || s.row_block_pos[r] != 2 * r)
      return fail(CB200_ERR_UNSUPPORTED_STRUCTURE,
                  "row block " + std::to_string(r) + " is not a densely packed block of size 2");
    const int c0 = s.cell_ptr[r], c1 = s.cell_ptr[r + 1];
    if (c1 - c0 != 2)
      return fail(CB200_ERR_UNSUPPORTED_STRUCTURE,
                  "row block " + std::to_string(r) + " has " + std::to_string(c1 - c0) +
                      " cells; expected one E cell and one F cell");
    const int pe = s.cell_block_id[c0], cf = s.cell_block_id[c0 + 1];
    if (pe < 0 || pe >= P || cf < P || cf >= P + C)
      return fail(CB200_ERR_UNSUPPORTED_STRUCTURE,
                  "row block " + std::to_string(r) + ": first cell must be an E cell, second an F cell");
    if (pe < last_point)
      return fail(CB200_ERR_UNSUPPORTED_STRUCTURE, "rows are not grouped by ascending E block");
    // rows of one point are contiguous; points without rows get empty ranges
    for (int p = last_point + 1; p <= pe; ++p) out->point_row_start[p] = r;
    last_point = pe;
    out->cam_idx[r] = cf - P;
    out->e_pos[r] = s.cell_position[c0];
    out->f_pos[r] = s.cell_position[c0 + 1];
    if (out->e_pos[r] < 0 || out->f_pos[r] < 0)
      return fail(CB200_ERR_INVALID_ARGUMENT, "negative cell position");
    canonical = canonical && out->e_pos[r] == 6 * r &&
                (long long)out->f_pos[r] == 6LL * N + 18LL * r;
  }
  for (int p = last_point + 1; p <= P; ++p) out->point_row_start[p] = N;
  out->canonical_layout = canonical;
  return CB200_OK;
}

// Contiguous point ranges with balanced observation counts (same idea as
// PartitionRangeForParallelFor over cumulative_nnz, block_sparse_matrix.cc:58-67):
// rank g owns points [bounds[g], bounds[g+1]).
inline std::vector<int> ShardPoints(const std::vector<int>& point_row_start, int world) {
  const int P = (int)point_row_start.size() - 1;
  const long long N = point_row_start[P];
  std::vector<int> bounds(world + 1, P);
  bounds[0] = 0;
  for (int g = 1; g < world; ++g) {
    const long long target = N * g / world;
    const int p = (int)(std::lower_bound(point_row_start.begin(), point_row_start.end(), target) -
                        point_row_start.begin());
    bounds[g] = std::max(bounds[g - 1], std::min(p, P));
  }
  return bounds;
}

// Tile plan of the fused per-point kernels: a tile is the unit of work of ONE
// WARP: whole points [p0, p0 + np) whose rows [r0, r0 + n) fit the 32 lanes.
// A point of 33..kChunkedPointRows rows becomes a group of chunk tiles that one
// warp runs back to back: first-pass chunks accumulate the point's sums, then
// second-pass chunks (the same rows again, now L2 hits) apply the result.
// Longer points go to long_points (CTA-per-point kernel), points without rows
// to empty_points. Indices are LOCAL to the shard (point 0
// = first point of the shard). One TileMeta record per tile is the static
// side-band the device streams next to the Jacobian blocks (one bulk copy).
struct alignas(16) TileMeta {
  int r0, n, p0, np;
  int cam[kTileObs];   // 10 * camera of each row (offset into a camera vector padded to 10 per camera)
  int seg[kTileObs];   // tile-local point | first lane of the point << 8 | last lane << 16
};
static_assert(sizeof(TileMeta) % 16 == 0, "bulk copies move multiples of 16 bytes");

struct TilePlan {
  std::vector<TileMeta> tiles;
  std::vector<int> long_points;   // local point ids
  std::vector<int> empty_points;  // local point ids
};

inline TilePlan PlanTiles(const int* row_start /* local, P_loc+1 */, const int* cam /* local rows */,
                          int P_loc) {
  TilePlan plan;
  std::vector<int> members;  // non-empty points of the open tile
  int open_rows = 0;
  auto emit = [&]() {
    if (members.empty()) return;
    TileMeta m;
    std::memset(&m, 0, sizeof(m));
    const int p0 = members.front(), p1 = members.back() + 1;
    m.r0 = row_start[p0];
    m.n = row_start[p1] - m.r0;
    m.p0 = p0;
    m.np = p1 - p0;
    for (int l = 0; l < kTileObs; ++l) m.seg[l] = 0 | (l << 8) | (l << 16);  // idle lanes: own segment
    for (int p : members) {
      const int first = row_start[p] - m.r0, last = row_start[p + 1] - 1 - m.r0;
      for (int l = first; l <= last; ++l) {
        m.cam[l] = 10 * cam[m.r0 + l];
        m.seg[l] = (p - p0) | (first << 8) | (last << 16);
      }
    }
    plan.tiles.push_back(m);
    members.clear();
    open_rows = 0;
  };
  auto emit_chunks = [&](int p) {
    const int r0 = row_start[p], len = row_start[p + 1] - r0;
    const int chunks = (len + kTileObs - 1) / kTileObs;
    for (int pass = 1; pass <= 2; ++pass)
      for (int ch = 0; ch < chunks; ++ch) {
        TileMeta m;
        std::memset(&m, 0, sizeof(m));
        m.r0 = r0 + ch * kTileObs;
        m.n = std::min(kTileObs, len - ch * kTileObs);
        m.p0 = p;
        m.np = 1 | (pass << kTileKindShift) | ((pass == 1 && ch == chunks - 1) ? kTileLastChunk : 0) |
               ((pass == 1 && ch == 0) ? 0 : kTileContinues);
        for (int l = 0; l < kTileObs; ++l) m.seg[l] = 0 | (l << 8) | (l << 16);
        for (int l = 0; l < m.n; ++l) {
          m.cam[l] = 10 * cam[m.r0 + l];
          m.seg[l] = 0 | (0 << 8) | ((m.n - 1) << 16);
        }
        plan.tiles.push_back(m);
      }
  };
  for (int p = 0; p < P_loc; ++p) {
    const int len = row_start[p + 1] - row_start[p];
    if (len == 0) { plan.empty_points.push_back(p); continue; }  // (rows stay contiguous)
    if (len > kTileObs) {
      emit();
      if (len <= kChunkedPointRows) emit_chunks(p);
      else plan.long_points.push_back(p);
      continue;
    }
    if (open_rows + len > kTileObs) emit();
    if (!members.empty() && p - members.front() >= kTilePoints) emit();  // np <= kTilePoints (empty points between)
    members.push_back(p);
    open_rows += len;
  }
  emit();
  return plan;
}

}  // namespace cb200

#endif  // CB200_FLATTEN_H_
