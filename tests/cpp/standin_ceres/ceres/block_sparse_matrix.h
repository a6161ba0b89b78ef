// Stand-in (test only): the accessors of internal/ceres/block_sparse_matrix.h the adapter uses.
#pragma once
#include <vector>
#include "ceres/block_structure.h"
namespace ceres::internal {
class LinearOperator { public: virtual ~LinearOperator() = default; };
class BlockSparseMatrix : public LinearOperator {
 public:
  BlockSparseMatrix(CompressedRowBlockStructure bs, std::vector<double> values)
      : bs_(std::move(bs)), values_(std::move(values)) {}
  const CompressedRowBlockStructure* block_structure() const { return &bs_; }
  const double* values() const { return values_.data(); }
  int num_nonzeros() const { return static_cast<int>(values_.size()); }
 private:
  CompressedRowBlockStructure bs_;
  std::vector<double> values_;
};
}  // namespace ceres::internal
