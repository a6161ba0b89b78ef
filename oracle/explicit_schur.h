// ORACLE — TEST INFRASTRUCTURE ONLY (see small_blas.h header).
//
// CPU restatement of ITERATIVE_SCHUR with use_explicit_schur_complement = true
// (linear_solver.cc:111-116 returns a SparseSchurComplementSolver):
//   internal/ceres/schur_complement_solver.cc:101-163   SchurComplementSolver::SolveImpl
//   internal/ceres/schur_complement_solver.cc:224-290   SparseSchurComplementSolver::InitStorage
//   internal/ceres/schur_complement_solver.cc:337-408   SolveReducedLinearSystemUsingConjugateGradients
//   internal/ceres/block_random_access_sparse_matrix.cc:47-170  ctor / GetCell /
//                                                       SymmetricRightMultiplyAndAccumulate
//   internal/ceres/schur_eliminator_impl.h:80-174       Init (chunks, buffer_layout)
//   internal/ceres/schur_eliminator_impl.h:176-304      Eliminate
//   internal/ceres/schur_eliminator_impl.h:306-374      BackSubstitute
//   internal/ceres/schur_eliminator_impl.h:379-420      UpdateRhs
//   internal/ceres/schur_eliminator_impl.h:441-505      ChunkDiagonalBlockAndGradient
//   internal/ceres/schur_eliminator_impl.h:511-561      ChunkOuterProduct
//   internal/ceres/schur_eliminator_impl.h:566-593      NoEBlockRowsUpdate
//   internal/ceres/schur_eliminator_impl.h:609-714      NoEBlockRowOuterProduct / EBlockRowOuterProduct
//   internal/ceres/block_random_access_diagonal_matrix.cc:90-116  Invert / multiply
#ifndef ORACLE_EXPLICIT_SCHUR_H_
#define ORACLE_EXPLICIT_SCHUR_H_

#include <algorithm>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <unordered_map>
#include <vector>

#include "solver.h"

namespace oracle {

// BlockRandomAccessSparseMatrix: the upper-triangular block pairs as a block CRS
// matrix, each cell a contiguous row-major dense block.
class BlockRandomAccessSparseMatrix {
 public:
  BlockRandomAccessSparseMatrix(const std::vector<Block>& blocks,
                                const std::set<std::pair<int, int>>& block_pairs)
      : blocks_(blocks) {
    const int nb = (int)blocks.size();
    row_ptr_.assign(nb + 1, 0);
    for (const auto& p : block_pairs) ++row_ptr_[p.first + 1];
    for (int i = 0; i < nb; ++i) row_ptr_[i + 1] += row_ptr_[i];
    col_.reserve(block_pairs.size());
    pos_.reserve(block_pairs.size());
    int nnz = 0;
    for (const auto& p : block_pairs) {  // lexicographic: row by row, columns increasing
      col_.push_back(p.second);
      pos_.push_back(nnz);
      layout_[Key(p.first, p.second)] = nnz;
      nnz += blocks[p.first].size * blocks[p.second].size;
    }
    values_.assign(nnz, 0.0);
    locks_ = std::vector<std::mutex>(col_.size());
    num_rows_ = nb == 0 ? 0 : blocks.back().position + blocks.back().size;
  }

  // nullptr when the pair is not present; *cell is the cell index (for its lock).
  double* GetCell(int r, int c, int* cell) {
    auto it = layout_.find(Key(r, c));
    if (it == layout_.end()) return nullptr;
    if (cell) *cell = LowerBoundCell(r, c);
    return &values_[it->second];
  }
  std::mutex& lock(int cell) { return locks_[cell]; }
  void SetZero() { std::fill(values_.begin(), values_.end(), 0.0); }
  int num_rows() const { return num_rows_; }
  const std::vector<double>& values() const { return values_; }
  const std::vector<int>& row_ptr() const { return row_ptr_; }
  const std::vector<int>& cols() const { return col_; }
  const std::vector<int>& positions() const { return pos_; }
  const std::vector<Block>& blocks() const { return blocks_; }

  // y += S x with only the upper triangle stored (:128-170).
  void SymmetricRightMultiplyAndAccumulate(const double* x, double* y) const {
    const int nb = (int)blocks_.size();
    for (int i = 0; i < nb; ++i) {
      const int rs = blocks_[i].size, rp = blocks_[i].position;
      for (int k = row_ptr_[i]; k < row_ptr_[i + 1]; ++k) {
        const int j = col_[k];
        const int cs = blocks_[j].size, cp = blocks_[j].position;
        MatVec(&values_[pos_[k]], rs, cs, x + cp, y + rp, 1);
        if (j == i) continue;
        MatTVec(&values_[pos_[k]], rs, cs, x + rp, y + cp, 1);
      }
    }
  }

 private:
  static long long Key(int r, int c) { return ((long long)r << 32) | (unsigned int)c; }
  int LowerBoundCell(int r, int c) const {
    return (int)(std::lower_bound(col_.begin() + row_ptr_[r], col_.begin() + row_ptr_[r + 1], c) -
                 col_.begin());
  }
  std::vector<Block> blocks_;
  std::vector<int> row_ptr_, col_, pos_;
  std::unordered_map<long long, int> layout_;
  std::vector<double> values_;
  std::vector<std::mutex> locks_;
  int num_rows_ = 0;
};

// SchurEliminator<Dynamic, Dynamic, Dynamic> writing into a
// BlockRandomAccessSparseMatrix, with the reduced right hand side.
class SchurEliminator {
 public:
  SchurEliminator(const BlockStructure* bs, int num_eliminate_blocks, int num_threads)
      : bs_(bs), neb_(num_eliminate_blocks), nt_(std::max(1, num_threads)) {
    const int ncb = (int)bs->cols.size();
    const int nrb = (int)bs->rows.size();
    int lhs_num_rows = 0;
    lhs_row_layout_.resize(ncb - neb_);
    for (int i = neb_; i < ncb; ++i) {
      lhs_row_layout_[i - neb_] = lhs_num_rows;
      lhs_num_rows += bs->cols[i].size;
    }
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
          if (chunk.buffer_layout.find(cell.block_id) == chunk.buffer_layout.end()) {
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
    uneliminated_row_begins_ = chunks_.empty() ? 0 : chunks_.back().start + chunks_.back().size;
    rhs_locks_ = std::vector<std::mutex>(ncb - neb_);
  }

  void Eliminate(const double* values, const double* b, const double* D,
                 BlockRandomAccessSparseMatrix* lhs, double* rhs) {
    if (lhs->num_rows() > 0) {
      lhs->SetZero();
      if (rhs) std::fill(rhs, rhs + lhs->num_rows(), 0.0);
    }
    const int ncb = (int)bs_->cols.size();
    if (D != nullptr) {
      for (int i = neb_; i < ncb; ++i) {
        const int block_id = i - neb_;
        double* m = lhs->GetCell(block_id, block_id, nullptr);
        if (m == nullptr) continue;
        const int n = bs_->cols[i].size;
        const double* d = D + bs_->cols[i].position;
        for (int k = 0; k < n; ++k) m[k * n + k] += d[k] * d[k];
      }
    }
    std::vector<double> buffer(buffer_size_ * (size_t)nt_), outer(buffer_size_ * (size_t)nt_);
    ParallelForRange(0, (int)chunks_.size(), nt_, [&](int tid, int s, int e) {
      double* buf = buffer.data() + (size_t)tid * buffer_size_;
      double* b1t = outer.data() + (size_t)tid * buffer_size_;
      for (int i = s; i < e; ++i) {
        const Chunk& chunk = chunks_[i];
        const int e_block_id = bs_->rows[chunk.start].cells.front().block_id;
        const int es = bs_->cols[e_block_id].size;
        std::fill(buf, buf + buffer_size_, 0.0);
        double ete[16 * 16], g[16];
        for (int k = 0; k < es * es; ++k) ete[k] = 0.0;
        for (int k = 0; k < es; ++k) g[k] = 0.0;
        if (D != nullptr) {
          const double* d = D + bs_->cols[e_block_id].position;
          for (int k = 0; k < es; ++k) ete[k * es + k] = d[k] * d[k];
        }
        // ChunkDiagonalBlockAndGradient
        int b_pos = bs_->rows[chunk.start].block.position;
        for (int j = 0; j < chunk.size; ++j) {
          const CompressedRow& row = bs_->rows[chunk.start + j];
          if (row.cells.size() > 1) RowOuterProduct(values, row, 1, lhs);
          const Cell& ec = row.cells.front();
          MatTMat(values + ec.position, row.block.size, es, values + ec.position, row.block.size, es,
                  ete, 0, 0, es, es, 1);
          if (b) MatTVec(values + ec.position, row.block.size, es, b + b_pos, g, 1);
          for (size_t c = 1; c < row.cells.size(); ++c) {
            const int fid = row.cells[c].block_id;
            const int fs = bs_->cols[fid].size;
            MatTMat(values + ec.position, row.block.size, es, values + row.cells[c].position,
                    row.block.size, fs, buf + chunk.buffer_layout.at(fid), 0, 0, es, fs, 1);
          }
          b_pos += row.block.size;
        }
        double inverse_ete[16 * 16];
        InvertPSD(ete, es, inverse_ete);
        if (rhs) {
          double inverse_ete_g[16];
          MatVec(inverse_ete, es, es, g, inverse_ete_g, 0);
          UpdateRhs(chunk, values, b, inverse_ete_g, rhs);
        }
        // ChunkOuterProduct: S(i,j) -= b_i' ete^-1 b_j for every stored pair i <= j
        for (auto it1 = chunk.buffer_layout.begin(); it1 != chunk.buffer_layout.end(); ++it1) {
          const int block1 = it1->first - neb_;
          const int b1s = bs_->cols[it1->first].size;
          MatTMat(buf + it1->second, es, b1s, inverse_ete, es, es, b1t, 0, 0, b1s, es, 0);
          for (auto it2 = it1; it2 != chunk.buffer_layout.end(); ++it2) {
            const int block2 = it2->first - neb_;
            int cell = 0;
            double* m = lhs->GetCell(block1, block2, &cell);
            if (m == nullptr) continue;
            const int b2s = bs_->cols[it2->first].size;
            std::lock_guard<std::mutex> l(lhs->lock(cell));
            MatMat(b1t, b1s, es, buf + it2->second, es, b2s, m, 0, 0, b1s, b2s, -1);
          }
        }
      }
    });
    // NoEBlockRowsUpdate
    for (int r = uneliminated_row_begins_; r < (int)bs_->rows.size(); ++r) {
      const CompressedRow& row = bs_->rows[r];
      RowOuterProduct(values, row, 0, lhs);
      if (!rhs) continue;
      for (const Cell& cell : row.cells) {
        const int block = cell.block_id - neb_;
        MatTVec(values + cell.position, row.block.size, bs_->cols[cell.block_id].size,
                b + row.block.position, rhs + lhs_row_layout_[block], 1);
      }
    }
  }

  void BackSubstitute(const double* values, const double* b, const double* D, const double* z,
                      double* y) const {
    ParallelForRange(0, (int)chunks_.size(), nt_, [&](int, int s, int e) {
      for (int i = s; i < e; ++i) {
        const Chunk& chunk = chunks_[i];
        const int e_block_id = bs_->rows[chunk.start].cells.front().block_id;
        const int es = bs_->cols[e_block_id].size;
        double* y_ptr = y + bs_->cols[e_block_id].position;
        double ete[16 * 16];
        for (int k = 0; k < es * es; ++k) ete[k] = 0.0;
        if (D != nullptr) {
          const double* d = D + bs_->cols[e_block_id].position;
          for (int k = 0; k < es; ++k) ete[k * es + k] = d[k] * d[k];
        }
        for (int j = 0; j < chunk.size; ++j) {
          const CompressedRow& row = bs_->rows[chunk.start + j];
          const Cell& ec = row.cells.front();
          double sj[16];
          for (int k = 0; k < row.block.size; ++k) sj[k] = b[row.block.position + k];
          for (size_t c = 1; c < row.cells.size(); ++c) {
            const int fid = row.cells[c].block_id;
            MatVec(values + row.cells[c].position, row.block.size, bs_->cols[fid].size,
                   z + lhs_row_layout_[fid - neb_], sj, -1);
          }
          MatTVec(values + ec.position, row.block.size, es, sj, y_ptr, 1);
          MatTMat(values + ec.position, row.block.size, es, values + ec.position, row.block.size, es,
                  ete, 0, 0, es, es, 1);
        }
        double inverse_ete[16 * 16], t[16];
        InvertPSD(ete, es, inverse_ete);
        MatVec(inverse_ete, es, es, y_ptr, t, 0);
        for (int k = 0; k < es; ++k) y_ptr[k] = t[k];
      }
    });
  }

 private:
  struct Chunk {
    int start, size;
    std::map<int, int> buffer_layout;
  };
  // InvertPSDMatrix<kEBlockSize>(assume_full_rank = true): closed form below 5, LLT otherwise
  static void InvertPSD(const double* m, int n, double* out) {
    if (n < 5) {
      InverseSmall(m, n, out);
    } else {
      for (int k = 0; k < n * n; ++k) out[k] = m[k];
      LltInverseUpper(out, n);
    }
  }
  void UpdateRhs(const Chunk& chunk, const double* values, const double* b, const double* inverse_ete_g,
                 double* rhs) {
    const int e_block_id = bs_->rows[chunk.start].cells.front().block_id;
    const int es = bs_->cols[e_block_id].size;
    int b_pos = bs_->rows[chunk.start].block.position;
    for (int j = 0; j < chunk.size; ++j) {
      const CompressedRow& row = bs_->rows[chunk.start + j];
      const Cell& ec = row.cells.front();
      double sj[16];
      for (int k = 0; k < row.block.size; ++k) sj[k] = b[b_pos + k];
      MatVec(values + ec.position, row.block.size, es, inverse_ete_g, sj, -1);
      for (size_t c = 1; c < row.cells.size(); ++c) {
        const int block_id = row.cells[c].block_id;
        const int block = block_id - neb_;
        std::lock_guard<std::mutex> l(rhs_locks_[block]);
        MatTVec(values + row.cells[c].position, row.block.size, bs_->cols[block_id].size, sj,
                rhs + lhs_row_layout_[block], 1);
      }
      b_pos += row.block.size;
    }
  }
  // EBlockRowOuterProduct (first = 1) / NoEBlockRowOuterProduct (first = 0): S += F'F of one row
  void RowOuterProduct(const double* values, const CompressedRow& row, size_t first,
                       BlockRandomAccessSparseMatrix* lhs) {
    for (size_t i = first; i < row.cells.size(); ++i) {
      const int block1 = row.cells[i].block_id - neb_;
      const int b1s = bs_->cols[row.cells[i].block_id].size;
      int cell = 0;
      double* m = lhs->GetCell(block1, block1, &cell);
      if (m != nullptr) {
        std::lock_guard<std::mutex> l(lhs->lock(cell));
        MatTMat(values + row.cells[i].position, row.block.size, b1s, values + row.cells[i].position,
                row.block.size, b1s, m, 0, 0, b1s, b1s, 1);
      }
      for (size_t j = i + 1; j < row.cells.size(); ++j) {
        const int block2 = row.cells[j].block_id - neb_;
        const int b2s = bs_->cols[row.cells[j].block_id].size;
        double* m2 = lhs->GetCell(block1, block2, &cell);
        if (m2 == nullptr) continue;
        std::lock_guard<std::mutex> l(lhs->lock(cell));
        MatTMat(values + row.cells[i].position, row.block.size, b1s, values + row.cells[j].position,
                row.block.size, b2s, m2, 0, 0, b1s, b2s, 1);
      }
    }
  }

  const BlockStructure* bs_;
  int neb_, nt_, buffer_size_ = 1, uneliminated_row_begins_ = 0;
  std::vector<int> lhs_row_layout_;
  std::vector<Chunk> chunks_;
  std::vector<std::mutex> rhs_locks_;
};

// SparseSchurComplementSolver for options.type == ITERATIVE_SCHUR.
class ExplicitSchurComplementSolver {
 public:
  ExplicitSchurComplementSolver(const BlockStructure* bs, const SolverOptions& options)
      : bs_(bs), options_(options) {}

  const BlockRandomAccessSparseMatrix* lhs() const { return lhs_.get(); }
  const Vec& rhs() const { return rhs_; }
  const Vec& reduced_solution() const { return cg_solution_; }

  // InitStorage + Eliminate only (for unit parity of S and the reduced rhs).
  void Eliminate(const double* values, const double* b, const double* D) {
    Init();
    eliminator_->Eliminate(values, b, D, lhs_.get(), rhs_.data());
  }

  Summary Solve(const double* values, const double* b, const double* D, double q_tolerance,
                double r_tolerance, double* x,
                const std::function<void(int, const Vec&)>& trace = nullptr) {
    Init();
    int num_cols = 0;
    for (const Block& c : bs_->cols) num_cols += c.size;
    std::fill(x, x + num_cols, 0.0);
    eliminator_->Eliminate(values, b, D, lhs_.get(), rhs_.data());
    double* reduced_solution = x + num_cols - lhs_->num_rows();
    const Summary summary = SolveReduced(q_tolerance, r_tolerance, reduced_solution, trace);
    if (summary.termination_type == LINEAR_SOLVER_SUCCESS)
      eliminator_->BackSubstitute(values, b, D, reduced_solution, x);
    return summary;
  }

 private:
  void Init() {
    if (eliminator_) return;
    const int neb = options_.num_eliminate_blocks;
    const int ncb = (int)bs_->cols.size();
    const int nrb = (int)bs_->rows.size();
    // blocks_ = Tail(bs->cols, ...) with positions restarting at 0
    int pos = 0;
    for (int i = neb; i < ncb; ++i) {
      blocks_.push_back(Block{bs_->cols[i].size, pos});
      pos += bs_->cols[i].size;
    }
    std::set<std::pair<int, int>> block_pairs;
    for (int i = 0; i < (int)blocks_.size(); ++i) block_pairs.emplace(i, i);
    int r = 0;
    while (r < nrb) {
      const int e_block_id = bs_->rows[r].cells.front().block_id;
      if (e_block_id >= neb) break;
      std::vector<int> f_blocks;
      for (; r < nrb; ++r) {
        const CompressedRow& row = bs_->rows[r];
        if (row.cells.front().block_id != e_block_id) break;
        for (size_t c = 1; c < row.cells.size(); ++c) f_blocks.push_back(row.cells[c].block_id - neb);
      }
      std::sort(f_blocks.begin(), f_blocks.end());
      f_blocks.erase(std::unique(f_blocks.begin(), f_blocks.end()), f_blocks.end());
      for (size_t i = 0; i < f_blocks.size(); ++i)
        for (size_t j = i + 1; j < f_blocks.size(); ++j) block_pairs.emplace(f_blocks[i], f_blocks[j]);
    }
    for (; r < nrb; ++r) {
      const CompressedRow& row = bs_->rows[r];
      for (const Cell& c1 : row.cells)
        for (const Cell& c2 : row.cells)
          if (c1.block_id <= c2.block_id) block_pairs.emplace(c1.block_id - neb, c2.block_id - neb);
    }
    lhs_ = std::make_unique<BlockRandomAccessSparseMatrix>(blocks_, block_pairs);
    rhs_.assign(lhs_->num_rows(), 0.0);
    eliminator_ = std::make_unique<SchurEliminator>(bs_, neb, options_.num_threads);
  }

  Summary SolveReduced(double q_tolerance, double r_tolerance, double* solution,
                       const std::function<void(int, const Vec&)>& trace) {
    const int num_rows = lhs_->num_rows();
    if (num_rows == 0) {
      Summary s;
      s.num_iterations = 0;
      s.termination_type = LINEAR_SOLVER_SUCCESS;
      s.message = "Success.";
      return s;
    }
    // block diagonal of S -> BlockRandomAccessDiagonalMatrix -> Invert()
    const int nb = (int)blocks_.size();
    std::vector<int> cell_pos(nb);
    int cp = 0;
    for (int i = 0; i < nb; ++i) { cell_pos[i] = cp; cp += blocks_[i].size * blocks_[i].size; }
    preconditioner_.assign(cp, 0.0);
    for (int i = 0; i < nb; ++i) {
      const double* m = lhs_->GetCell(i, i, nullptr);
      const int n = blocks_[i].size;
      for (int k = 0; k < n * n; ++k) preconditioner_[cell_pos[i] + k] = m[k];
      LltInverseUpper(&preconditioner_[cell_pos[i]], n);
    }
    CGOptions cg;
    cg.min_num_iterations = options_.min_num_iterations;
    cg.max_num_iterations = options_.max_num_iterations;
    cg.residual_reset_period = options_.residual_reset_period;
    cg.q_tolerance = q_tolerance;
    cg.r_tolerance = r_tolerance;
    Operator lhs = [this](const Vec& a, Vec& y) {
      lhs_->SymmetricRightMultiplyAndAccumulate(a.data(), y.data());
    };
    Operator precond = [this, cell_pos](const Vec& a, Vec& y) {
      for (size_t i = 0; i < blocks_.size(); ++i)
        MatVec(&preconditioner_[cell_pos[i]], blocks_[i].size, blocks_[i].size,
               a.data() + blocks_[i].position, y.data() + blocks_[i].position, 1);
    };
    cg_solution_.assign(num_rows, 0.0);
    Summary summary = ConjugateGradientsSolver(cg, lhs, rhs_, precond, cg_solution_, trace);
    for (int i = 0; i < num_rows; ++i) solution[i] = cg_solution_[i];
    return summary;
  }

  const BlockStructure* bs_;
  SolverOptions options_;
  std::vector<Block> blocks_;
  std::unique_ptr<BlockRandomAccessSparseMatrix> lhs_;
  std::unique_ptr<SchurEliminator> eliminator_;
  Vec rhs_, cg_solution_;
  std::vector<double> preconditioner_;
};

}  // namespace oracle

#endif  // ORACLE_EXPLICIT_SCHUR_H_
