// Stand-in (test only) with the member names of internal/ceres/block_structure.h:52-96.
#pragma once
#include <vector>
namespace ceres::internal {
struct Block { int size = -1; int position = -1; };
struct Cell { int block_id = -1; int position = -1; };
struct CompressedList { Block block; std::vector<Cell> cells; int nnz = -1; int cumulative_nnz = -1; };
using CompressedRow = CompressedList;
struct CompressedRowBlockStructure { std::vector<Block> cols; std::vector<CompressedList> rows; };
}  // namespace ceres::internal
