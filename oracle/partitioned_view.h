// ORACLE — TEST INFRASTRUCTURE ONLY (see small_blas.h header).
//
// CPU restatement of PartitionedMatrixView (dynamic block sizes, which covers
// <2,3,9> and the scalar / mixed blocks of the reference's known-answer tests):
//   internal/ceres/partitioned_matrix_view_impl.h:47-105   ctor, partitions
//   internal/ceres/partitioned_matrix_view_impl.h:112-137  RightMultiplyAndAccumulateE
//   internal/ceres/partitioned_matrix_view_impl.h:139-191  RightMultiplyAndAccumulateF
//   internal/ceres/partitioned_matrix_view_impl.h:193-264  LeftMultiplyAndAccumulateE{Single,Multi}Threaded
//   internal/ceres/partitioned_matrix_view_impl.h:266-375  LeftMultiplyAndAccumulateF{Single,Multi}Threaded
//   internal/ceres/partitioned_matrix_view_impl.h:382-439  CreateBlockDiagonal{MatrixLayout,EtE,FtF}
//   internal/ceres/partitioned_matrix_view_impl.h:446-523  UpdateBlockDiagonalEtE
//   internal/ceres/partitioned_matrix_view_impl.h:530-660  UpdateBlockDiagonalFtF
#ifndef ORACLE_PARTITIONED_VIEW_H_
#define ORACLE_PARTITIONED_VIEW_H_

#include <cstring>
#include <vector>

#include "block_structure.h"
#include "small_blas.h"

namespace oracle {

// Block-diagonal matrix: block i is (size_i x size_i) row-major at
// values[cell_pos[i]], acting on the vector slice starting at block_pos[i].
struct BlockDiagonal {
  std::vector<int> size, block_pos, cell_pos;
  std::vector<double> values;
  int num_rows = 0;
  // y += B x (block_sparse_matrix.cc:239-274 on a block-diagonal matrix).
  void RightMultiplyAndAccumulate(const double* x, double* y,
                                  int num_threads) const {
    ParallelForRange(0, (int)size.size(), num_threads, [&](int, int s, int e) {
      for (int i = s; i < e; ++i)
        MatVec(&values[cell_pos[i]], size[i], size[i], x + block_pos[i],
               y + block_pos[i], 1);
    });
  }
};

class PartitionedMatrixView {
 public:
  PartitionedMatrixView(const BlockStructure* bs, const BlockStructure* tbs,
                        const double* values, int num_col_blocks_e,
                        int num_threads)
      : bs_(bs), tbs_(tbs), values_(values), num_col_blocks_e_(num_col_blocks_e),
        num_threads_(num_threads) {
    num_col_blocks_f_ = (int)bs->cols.size() - num_col_blocks_e_;
    num_row_blocks_e_ = 0;
    for (const auto& row : bs->rows)
      if (row.cells[0].block_id < num_col_blocks_e_) ++num_row_blocks_e_;
    num_cols_e_ = num_cols_f_ = 0;
    for (int c = 0; c < (int)bs->cols.size(); ++c)
      (c < num_col_blocks_e_ ? num_cols_e_ : num_cols_f_) += bs->cols[c].size;
    num_rows_ = bs->rows.empty() ? 0
                                 : bs->rows.back().block.position +
                                       bs->rows.back().block.size;
    if (tbs != nullptr && num_threads > 1) {
      const int kMax = num_threads * 4;
      auto cum = [&](int i) { return tbs->rows[i].cumulative_nnz; };
      if (num_col_blocks_e_ > 0)
        e_cols_partition_ = PartitionRange(0, num_col_blocks_e_, kMax, cum);
      if (num_col_blocks_f_ > 0)
        f_cols_partition_ =
            PartitionRange(num_col_blocks_e_,
                           num_col_blocks_e_ + num_col_blocks_f_, kMax, cum);
    }
  }

  void set_values(const double* v) { values_ = v; }
  int num_rows() const { return num_rows_; }
  int num_cols_e() const { return num_cols_e_; }
  int num_cols_f() const { return num_cols_f_; }
  int num_col_blocks_e() const { return num_col_blocks_e_; }
  int num_col_blocks_f() const { return num_col_blocks_f_; }
  int num_row_blocks_e() const { return num_row_blocks_e_; }

  void RightMultiplyAndAccumulateE(const double* x, double* y) const {
    ParallelForRange(0, num_row_blocks_e_, num_threads_, [&](int, int s, int e) {
      for (int r = s; r < e; ++r) {
        const CompressedRow& row = bs_->rows[r];
        const Cell& cell = row.cells[0];
        const Block& col = bs_->cols[cell.block_id];
        MatVec(values_ + cell.position, row.block.size, col.size,
               x + col.position, y + row.block.position, 1);
      }
    });
  }

  void RightMultiplyAndAccumulateF(const double* x, double* y) const {
    const int nr = (int)bs_->rows.size();
    ParallelForRange(0, num_row_blocks_e_, num_threads_, [&](int, int s, int e) {
      for (int r = s; r < e; ++r) {
        const CompressedRow& row = bs_->rows[r];
        for (size_t c = 1; c < row.cells.size(); ++c) {
          const Block& col = bs_->cols[row.cells[c].block_id];
          MatVec(values_ + row.cells[c].position, row.block.size, col.size,
                 x + col.position - num_cols_e_, y + row.block.position, 1);
        }
      }
    });
    ParallelForRange(num_row_blocks_e_, nr, num_threads_, [&](int, int s, int e) {
      for (int r = s; r < e; ++r) {
        const CompressedRow& row = bs_->rows[r];
        for (const Cell& cell : row.cells) {
          const Block& col = bs_->cols[cell.block_id];
          MatVec(values_ + cell.position, row.block.size, col.size,
                 x + col.position - num_cols_e_, y + row.block.position, 1);
        }
      }
    });
  }

  void LeftMultiplyAndAccumulateE(const double* x, double* y) const {
    if (!num_col_blocks_e_ || !num_row_blocks_e_) return;
    if (num_threads_ == 1) {
      for (int r = 0; r < num_row_blocks_e_; ++r) {
        const CompressedRow& row = bs_->rows[r];
        const Cell& cell = row.cells[0];
        const Block& col = bs_->cols[cell.block_id];
        MatTVec(values_ + cell.position, row.block.size, col.size,
                x + row.block.position, y + col.position, 1);
      }
      return;
    }
    ParallelForPartitions(
        0, num_col_blocks_e_, num_threads_, e_cols_partition_,
        [&](int, int s, int e) {
          for (int cb = s; cb < e; ++cb) {
            const CompressedRow& trow = tbs_->rows[cb];
            for (const Cell& cell : trow.cells) {
              if (cell.block_id >= num_row_blocks_e_) break;
              const Block& rb = tbs_->cols[cell.block_id];
              MatTVec(values_ + cell.position, rb.size, trow.block.size,
                      x + rb.position, y + trow.block.position, 1);
            }
          }
        });
  }

  void LeftMultiplyAndAccumulateF(const double* x, double* y) const {
    if (!num_col_blocks_f_) return;
    if (num_threads_ == 1) {
      for (int r = 0; r < (int)bs_->rows.size(); ++r) {
        const CompressedRow& row = bs_->rows[r];
        for (size_t c = (r < num_row_blocks_e_ ? 1 : 0); c < row.cells.size();
             ++c) {
          const Block& col = bs_->cols[row.cells[c].block_id];
          MatTVec(values_ + row.cells[c].position, row.block.size, col.size,
                  x + row.block.position, y + col.position - num_cols_e_, 1);
        }
      }
      return;
    }
    ParallelForPartitions(
        num_col_blocks_e_, num_col_blocks_e_ + num_col_blocks_f_, num_threads_,
        f_cols_partition_, [&](int, int s, int e) {
          for (int cb = s; cb < e; ++cb) {
            const CompressedRow& trow = tbs_->rows[cb];
            for (const Cell& cell : trow.cells) {
              const Block& rb = tbs_->cols[cell.block_id];
              MatTVec(values_ + cell.position, rb.size, trow.block.size,
                      x + rb.position, y + trow.block.position - num_cols_e_, 1);
            }
          }
        });
  }

  BlockDiagonal CreateBlockDiagonalLayout(int start_cb, int end_cb) const {
    BlockDiagonal d;
    int bp = 0, cp = 0;
    for (int c = start_cb; c < end_cb; ++c) {
      const int s = bs_->cols[c].size;
      d.size.push_back(s);
      d.block_pos.push_back(bp);
      d.cell_pos.push_back(cp);
      bp += s;
      cp += s * s;
    }
    d.num_rows = bp;
    d.values.assign(cp, 0.0);
    return d;
  }
  BlockDiagonal CreateBlockDiagonalEtE() const {
    BlockDiagonal d = CreateBlockDiagonalLayout(0, num_col_blocks_e_);
    UpdateBlockDiagonalEtE(&d);
    return d;
  }
  BlockDiagonal CreateBlockDiagonalFtF() const {
    BlockDiagonal d = CreateBlockDiagonalLayout(
        num_col_blocks_e_, num_col_blocks_e_ + num_col_blocks_f_);
    UpdateBlockDiagonalFtF(&d);
    return d;
  }

  void UpdateBlockDiagonalEtE(BlockDiagonal* d) const {
    if (num_threads_ == 1) {
      std::fill(d->values.begin(), d->values.end(), 0.0);
      for (int r = 0; r < num_row_blocks_e_; ++r) {
        const CompressedRow& row = bs_->rows[r];
        const Cell& cell = row.cells[0];
        const int cs = bs_->cols[cell.block_id].size;
        MatTMat(values_ + cell.position, row.block.size, cs,
                values_ + cell.position, row.block.size, cs,
                &d->values[d->cell_pos[cell.block_id]], 0, 0, cs, cs, 1);
      }
      return;
    }
    ParallelForPartitions(
        0, num_col_blocks_e_, num_threads_, e_cols_partition_,
        [&](int, int s, int e) {
          for (int cb = s; cb < e; ++cb) {
            const CompressedRow& trow = tbs_->rows[cb];
            const int cs = trow.block.size;
            double* out = &d->values[d->cell_pos[cb]];
            std::memset(out, 0, sizeof(double) * cs * cs);
            for (const Cell& c : trow.cells) {
              const int rs = tbs_->cols[c.block_id].size;
              MatTMat(values_ + c.position, rs, cs, values_ + c.position, rs,
                      cs, out, 0, 0, cs, cs, 1);
            }
          }
        });
  }

  void UpdateBlockDiagonalFtF(BlockDiagonal* d) const {
    if (num_threads_ == 1) {
      std::fill(d->values.begin(), d->values.end(), 0.0);
      for (int r = 0; r < (int)bs_->rows.size(); ++r) {
        const CompressedRow& row = bs_->rows[r];
        for (size_t c = (r < num_row_blocks_e_ ? 1 : 0); c < row.cells.size();
             ++c) {
          const int cbid = row.cells[c].block_id;
          const int cs = bs_->cols[cbid].size;
          MatTMat(values_ + row.cells[c].position, row.block.size, cs,
                  values_ + row.cells[c].position, row.block.size, cs,
                  &d->values[d->cell_pos[cbid - num_col_blocks_e_]], 0, 0, cs,
                  cs, 1);
        }
      }
      return;
    }
    ParallelForPartitions(
        num_col_blocks_e_, num_col_blocks_e_ + num_col_blocks_f_, num_threads_,
        f_cols_partition_, [&](int, int s, int e) {
          for (int cb = s; cb < e; ++cb) {
            const CompressedRow& trow = tbs_->rows[cb];
            const int cs = trow.block.size;
            double* out = &d->values[d->cell_pos[cb - num_col_blocks_e_]];
            std::memset(out, 0, sizeof(double) * cs * cs);
            for (const Cell& c : trow.cells) {
              const int rs = tbs_->cols[c.block_id].size;
              MatTMat(values_ + c.position, rs, cs, values_ + c.position, rs,
                      cs, out, 0, 0, cs, cs, 1);
            }
          }
        });
  }

 private:
  const BlockStructure* bs_;
  const BlockStructure* tbs_;
  const double* values_;
  int num_col_blocks_e_, num_col_blocks_f_, num_row_blocks_e_;
  int num_cols_e_, num_cols_f_, num_rows_, num_threads_;
  std::vector<int> e_cols_partition_, f_cols_partition_;
};

}  // namespace oracle

#endif  // ORACLE_PARTITIONED_VIEW_H_
