// ORACLE — TEST INFRASTRUCTURE ONLY (see small_blas.h header).
//
// CPU restatement of SchurJacobiPreconditioner = SchurEliminator::Eliminate into
// a BlockRandomAccessDiagonalMatrix (b = rhs = nullptr) followed by Invert():
//   internal/ceres/schur_jacobi_preconditioner.cc:45-104
//   internal/ceres/schur_eliminator_impl.h:80-174    Init (chunks, buffer_layout)
//   internal/ceres/schur_eliminator_impl.h:176-304   Eliminate
//   internal/ceres/schur_eliminator_impl.h:441-505   ChunkDiagonalBlockAndGradient
//   internal/ceres/schur_eliminator_impl.h:511-561   ChunkOuterProduct
//   internal/ceres/schur_eliminator_impl.h:566-593   NoEBlockRowsUpdate
//   internal/ceres/schur_eliminator_impl.h:664-714   EBlockRowOuterProduct
//   internal/ceres/block_random_access_diagonal_matrix.cc:50-116 GetCell/Invert/apply
//   internal/ceres/invert_psd_matrix.h:48-72         InvertPSDMatrix (full rank)
// Only (c,c) cells exist in the diagonal LHS, so every off-diagonal update the
// eliminator attempts is dropped by GetCell returning nullptr.
#ifndef ORACLE_SCHUR_JACOBI_H_
#define ORACLE_SCHUR_JACOBI_H_

#include <map>
#include <mutex>
#include <vector>

#include "block_structure.h"
#include "small_blas.h"

namespace oracle {

class SchurJacobiPreconditioner {
 public:
  SchurJacobiPreconditioner(const BlockStructure* bs, int num_eliminate_blocks,
                            int num_threads)
      : bs_(bs), neb_(num_eliminate_blocks), num_threads_(num_threads) {
    const int ncb = (int)bs->cols.size();
    int pos = 0, cp = 0;
    for (int i = neb_; i < ncb; ++i) {
      size_.push_back(bs->cols[i].size);
      block_pos_.push_back(pos);
      cell_pos_.push_back(cp);
      pos += bs->cols[i].size;
      cp += bs->cols[i].size * bs->cols[i].size;
    }
    num_rows_ = pos;
    values_.assign(cp, 0.0);
    locks_ = std::vector<std::mutex>(size_.size());
    // Init: detect chunks.
    const int nrb = (int)bs->rows.size();
    buffer_size_ = 1;
    int r = 0;
    while (r < nrb) {
      const int chunk_block_id = bs->rows[r].cells.front().block_id;
      if (chunk_block_id >= neb_) break;
      Chunk chunk;
      chunk.start = r;
      chunk.size = 0;
      int buffer_size = 0;
      const int es = bs->cols[chunk_block_id].size;
      while (r + chunk.size < nrb) {
        const CompressedRow& row = bs->rows[r + chunk.size];
        if (row.cells.front().block_id != chunk_block_id) break;
        for (size_t c = 1; c < row.cells.size(); ++c) {
          const Cell& cell = row.cells[c];
          if (chunk.buffer_layout.find(cell.block_id) ==
              chunk.buffer_layout.end()) {
            chunk.buffer_layout[cell.block_id] = buffer_size;
            buffer_size += es * bs->cols[cell.block_id].size;
          }
        }
        buffer_size_ = std::max(buffer_size, buffer_size_);
        ++chunk.size;
      }
      r += chunk.size;
      chunks_.push_back(std::move(chunk));
    }
    uneliminated_row_begins_ =
        chunks_.empty() ? 0 : chunks_.back().start + chunks_.back().size;
  }

  int num_rows() const { return num_rows_; }
  const std::vector<double>& values() const { return values_; }

  // UpdateImpl: Eliminate(A, nullptr, D, m_, nullptr); m_->Invert().
  bool Update(const double* values, const double* D) {
    Eliminate(values, D);
    Invert();
    return true;
  }

  // Eliminate only (block-diag(S) before inversion) — exposed for parity tests.
  void Eliminate(const double* values, const double* D) {
    std::fill(values_.begin(), values_.end(), 0.0);
    const int ncb = (int)bs_->cols.size();
    if (D != nullptr) {
      for (int i = neb_; i < ncb; ++i) {
        const int n = bs_->cols[i].size;
        double* m = &values_[cell_pos_[i - neb_]];
        const double* d = D + bs_->cols[i].position;
        for (int k = 0; k < n; ++k) m[k * n + k] += d[k] * d[k];
      }
    }
    const int nt = std::max(1, num_threads_);
    std::vector<double> buffer(buffer_size_ * (size_t)nt);
    std::vector<double> outer(buffer_size_ * (size_t)nt);
    ParallelForRange(0, (int)chunks_.size(), nt, [&](int tid, int s, int e) {
      double* buf = buffer.data() + (size_t)tid * buffer_size_;
      double* b1t = outer.data() + (size_t)tid * buffer_size_;
      for (int i = s; i < e; ++i) {
        const Chunk& chunk = chunks_[i];
        const int e_block_id = bs_->rows[chunk.start].cells.front().block_id;
        const int es = bs_->cols[e_block_id].size;
        std::fill(buf, buf + buffer_size_, 0.0);
        double ete[16 * 16];
        for (int k = 0; k < es * es; ++k) ete[k] = 0.0;
        if (D != nullptr) {
          const double* d = D + bs_->cols[e_block_id].position;
          for (int k = 0; k < es; ++k) ete[k * es + k] = d[k] * d[k];
        }
        // ChunkDiagonalBlockAndGradient (b == nullptr)
        for (int j = 0; j < chunk.size; ++j) {
          const CompressedRow& row = bs_->rows[chunk.start + j];
          if (row.cells.size() > 1) EBlockRowOuterProduct(values, row);
          const Cell& ec = row.cells.front();
          MatTMat(values + ec.position, row.block.size, es,
                  values + ec.position, row.block.size, es, ete, 0, 0, es, es, 1);
          for (size_t c = 1; c < row.cells.size(); ++c) {
            const int fid = row.cells[c].block_id;
            const int fs = bs_->cols[fid].size;
            double* bp = buf + chunk.buffer_layout.at(fid);
            MatTMat(values + ec.position, row.block.size, es,
                    values + row.cells[c].position, row.block.size, fs, bp, 0, 0,
                    es, fs, 1);
          }
        }
        double inverse_ete[16 * 16];
        if (es < 5) {
          InverseSmall(ete, es, inverse_ete);
        } else {
          for (int k = 0; k < es * es; ++k) inverse_ete[k] = ete[k];
          LltInverseUpper(inverse_ete, es);
        }
        // ChunkOuterProduct: S(c,c) -= b_c' ete^-1 b_c (diagonal cells only).
        for (auto it1 = chunk.buffer_layout.begin();
             it1 != chunk.buffer_layout.end(); ++it1) {
          const int block1 = it1->first - neb_;
          const int b1s = bs_->cols[it1->first].size;
          MatTMat(buf + it1->second, es, b1s, inverse_ete, es, es, b1t, 0, 0,
                  b1s, es, 0);
          std::lock_guard<std::mutex> l(locks_[block1]);
          MatMat(b1t, b1s, es, buf + it1->second, es, b1s,
                 &values_[cell_pos_[block1]], 0, 0, b1s, b1s, -1);
        }
      }
    });
    // NoEBlockRowsUpdate: S += F'F for rows without an e-block.
    for (int r = uneliminated_row_begins_; r < (int)bs_->rows.size(); ++r) {
      const CompressedRow& row = bs_->rows[r];
      for (const Cell& cell : row.cells) {
        const int block1 = cell.block_id - neb_;
        const int s = bs_->cols[cell.block_id].size;
        MatTMat(values + cell.position, row.block.size, s,
                values + cell.position, row.block.size, s,
                &values_[cell_pos_[block1]], 0, 0, s, s, 1);
      }
    }
  }

  void Invert() {
    ParallelForRange(0, (int)size_.size(), num_threads_, [&](int, int s, int e) {
      for (int i = s; i < e; ++i) LltInverseUpper(&values_[cell_pos_[i]], size_[i]);
    });
  }

  // y += M^-1 x
  void RightMultiplyAndAccumulate(const double* x, double* y) const {
    ParallelForRange(0, (int)size_.size(), num_threads_, [&](int, int s, int e) {
      for (int i = s; i < e; ++i)
        MatVec(&values_[cell_pos_[i]], size_[i], size_[i], x + block_pos_[i],
               y + block_pos_[i], 1);
    });
  }

 private:
  struct Chunk {
    int start, size;
    std::map<int, int> buffer_layout;
  };
  void EBlockRowOuterProduct(const double* values, const CompressedRow& row) {
    for (size_t i = 1; i < row.cells.size(); ++i) {
      const int block1 = row.cells[i].block_id - neb_;
      const int s = bs_->cols[row.cells[i].block_id].size;
      std::lock_guard<std::mutex> l(locks_[block1]);
      MatTMat(values + row.cells[i].position, row.block.size, s,
              values + row.cells[i].position, row.block.size, s,
              &values_[cell_pos_[block1]], 0, 0, s, s, 1);
    }
  }

  const BlockStructure* bs_;
  int neb_, num_threads_, num_rows_ = 0, buffer_size_ = 1;
  int uneliminated_row_begins_ = 0;
  std::vector<int> size_, block_pos_, cell_pos_;
  std::vector<double> values_;
  std::vector<std::mutex> locks_;
  std::vector<Chunk> chunks_;
};

}  // namespace oracle

#endif  // ORACLE_SCHUR_JACOBI_H_
