// Test driver: compiles include/ceres_b200_adapter.h against a stand-in for
// ceres/block_structure.h (same member names as block_structure.h:52-96), flattens a small
// <2,3,9> structure, checks the tile plan of csrc/flatten.h on it, and calls cb200_create.
// Prints one line of key=value pairs.
#include <cstdio>
#include <vector>

#include "../../include/ceres_b200_adapter.h"
#include "../../ceres-solver-ceres-solver_b200/csrc/flatten.h"

#ifdef CB200_REAL_BLOCK_STRUCTURE
// the reference's own internal/ceres/block_structure.h (Eigen-free; export.h stubbed), where the tree is present
#include "ceres/block_structure.h"
namespace standin = ceres::internal;
#else
namespace standin {
struct Block { int size = -1; int position = -1; };
struct Cell { int block_id = -1; int position = -1; };
struct CompressedList { Block block; std::vector<Cell> cells; int nnz = -1; int cumulative_nnz = -1; };
struct CompressedRowBlockStructure { std::vector<Block> cols; std::vector<CompressedList> rows; };
}  // namespace standin
#endif

int main(int argc, char** argv) {
  const int P = 40, C = 7;
  const int chunked_point = 5, chunked_len = 45;  // 33..256 rows: chunk tiles inside the stream kernel
  const int long_point = 20, long_len = 300;      // > 256 rows: CTA-per-point kernel
  standin::CompressedRowBlockStructure bs;
  for (int p = 0; p < P; ++p) bs.cols.push_back({3, 3 * p});
  for (int c = 0; c < C; ++c) bs.cols.push_back({9, 3 * P + 9 * c});
  std::vector<int> len(P);
  int N = 0;
  for (int p = 0; p < P; ++p) { len[p] = (p == long_point) ? long_len : (p == chunked_point) ? chunked_len : (p == 9 ? 0 : 2 + p % 5); N += len[p]; }
  int r = 0;
  for (int p = 0; p < P; ++p)
    for (int k = 0; k < len[p]; ++k, ++r) {
      standin::CompressedList row;
      row.block = {2, 2 * r};
      row.cells.push_back({p, 6 * r});
      row.cells.push_back({P + (p + k) % C, 6 * N + 18 * r});
      bs.rows.push_back(row);
    }
  const ceres_b200::FlatBlockStructure flat = ceres_b200::FlattenBlockStructure(bs);
  const cb200_structure view = flat.view();
  cb200::FlatStructure fs;
  std::string err;
  const int frc = cb200::Flatten(view, P, &fs, &err);
  const cb200::TilePlan plan = cb200::PlanTiles(fs.point_row_start.data(), fs.cam_idx.data(), P);
  // every row of a tiled point appears exactly once, in order, with a consistent segment
  int rows_in_tiles = 0, bad = 0, chunk_rows[3] = {0, 0, 0}, group_starts = 0;
  for (const auto& t : plan.tiles) {
    const int kind = (t.np >> cb200::kTileKindShift) & 3, np = t.np & 255;
    if (!(t.np & cb200::kTileContinues)) ++group_starts;
    if (kind != 0) {  // chunk of the chunked point: one segment over the active lanes
      chunk_rows[kind] += t.n;
      if (t.p0 != chunked_point || np != 1 || t.n <= 0 || t.n > cb200::kTileObs) ++bad;
      for (int l = 0; l < t.n; ++l) {
        if (t.seg[l] != ((t.n - 1) << 16) || t.cam[l] != 10 * fs.cam_idx[t.r0 + l]) ++bad;
      }
      const bool last = (t.np & cb200::kTileLastChunk) != 0;
      if (last != (kind == 1 && t.r0 + t.n == fs.point_row_start[chunked_point + 1])) ++bad;
      continue;
    }
    rows_in_tiles += t.n;
    if (t.n > cb200::kTileObs || np > cb200::kTileObs || t.n <= 0) ++bad;
    for (int l = 0; l < t.n; ++l) {
      const int pt = t.seg[l] & 255, head = (t.seg[l] >> 8) & 255, tail = (t.seg[l] >> 16) & 255;
      const int p = t.p0 + pt;
      if (fs.point_row_start[p] != t.r0 + head || fs.point_row_start[p + 1] - 1 != t.r0 + tail) ++bad;
      if (t.cam[l] != 10 * fs.cam_idx[t.r0 + l]) ++bad;
    }
    for (int l = t.n; l < cb200::kTileObs; ++l)
      if (t.cam[l] != 0 || ((t.seg[l] >> 8) & 255) != l || ((t.seg[l] >> 16) & 255) != l) ++bad;
  }
  const std::vector<int> shard = cb200::ShardPoints(fs.point_row_start, 3);
  cb200_options o{};
  o.num_eliminate_blocks = P;
  o.preconditioner_type = CB200_SCHUR_JACOBI;
  o.max_num_iterations = 10;
  o.residual_reset_period = 10;
  o.world_size = 1;
  cb200_solver* h = nullptr;
  const int crc = cb200_create(&view, &o, &h);
  std::printf("flatten=%d canonical=%d N=%d tiles=%d rows_in_tiles=%d chunk1=%d chunk2=%d groups=%d long=%d "
              "empty=%d bad=%d shard=%d,%d,%d,%d create=%d\n",
              frc, (int)fs.canonical_layout, fs.num_obs, (int)plan.tiles.size(), rows_in_tiles, chunk_rows[1],
              chunk_rows[2], group_starts,
              (int)plan.long_points.size(), (int)plan.empty_points.size(), bad, shard[0], shard[1], shard[2],
              shard[3], crc);
  if (h) cb200_destroy(h);
  return 0;
}
