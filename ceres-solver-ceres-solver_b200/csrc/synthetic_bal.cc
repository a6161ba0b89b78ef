// Synthetic BAL (bundle adjustment in the large) problems of a named shape:
// structure (which camera observes which point), scene geometry, and the
// Snavely reprojection residuals / Jacobian blocks written directly in the
// value layout Ceres' BlockJacobianWriter produces for an ITERATIVE_SCHUR
// problem (all 2x3 point blocks "E" first in row order, then all 2x9 camera
// blocks "F"; rows sorted by point; reference:
// internal/ceres/block_jacobian_writer.cc:102-165, reorder_program.cc:446-540).
//
// Host-only measurement / test support (no CUDA, no oracle): the BAL data files
// are not shipped with the reference and there is no network, so problems of the
// BASELINE.json shapes are synthesised. Camera model follows
// examples/snavely_reprojection_error.h:57-92 and the angle-axis rotation of
// include/ceres/rotation.h:792-850; derivatives come from a small forward-mode
// dual number type instead of ceres::Jet.
//
// Build: g++ -O3 -fopenmp -shared -fPIC synthetic_bal.cc -o libsynthetic_bal.so
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <random>
#include <string>
#include <vector>

#include "snavely_model.h"
#include "bal_io.h"

using namespace cb200_model;

extern "C" {

// Observation structure of a BAL-like problem with exactly C cameras, P points
// and N observations. point_row_start[P+1] (rows are grouped by point), cam_idx[N]
// ascending within a point. Track lengths are 2 + geometric, capped at C and
// nudged to sum to N; cameras of a point cluster around a popularity-weighted
// "home" camera (spatial locality + skewed per-camera counts, like real BAL).
// dup_pairs > 0 forces that many points to observe one camera twice (exercises
// the duplicate (point, camera) merge of schur_eliminator_impl.h:140-148).
// Returns 0, or -1 when the shape is infeasible (N < 2P or N > P*C).
int bal_synth_structure(int C, int P, long long N, unsigned long long seed,
                        int dup_pairs, int* point_row_start, int* cam_idx) {
  if (C < 1 || P < 1) return -1;
  const int min_len = C >= 2 ? 2 : 1;
  if (N < (long long)min_len * P || N > (long long)P * C) return -1;
  std::mt19937_64 rng(seed);
  std::vector<int> len(P);
  const double mean_extra = (double)N / P - min_len;
  std::geometric_distribution<int> geo(1.0 / (1.0 + std::max(mean_extra, 1e-9)));
  long long sum = 0;
  for (int p = 0; p < P; ++p) {
    len[p] = std::min(C, min_len + (mean_extra > 0 ? geo(rng) : 0));
    sum += len[p];
  }
  std::uniform_int_distribution<int> pick(0, P - 1);
  while (sum != N) {
    const int p = pick(rng);
    if (sum < N && len[p] < C) { ++len[p]; ++sum; }
    else if (sum > N && len[p] > min_len) { --len[p]; --sum; }
  }
  point_row_start[0] = 0;
  for (int p = 0; p < P; ++p) {
    const long long next = (long long)point_row_start[p] + len[p];
    point_row_start[p + 1] = (int)next;
  }
  // camera popularity: log-normal weights -> cumulative table
  std::normal_distribution<double> gauss(0.0, 1.0);
  std::vector<double> cum(C);
  double acc = 0.0;
  for (int c = 0; c < C; ++c) { acc += std::exp(0.7 * gauss(rng)); cum[c] = acc; }
  std::uniform_real_distribution<double> uni(0.0, 1.0);
  const double spread = std::max(2.0, 0.02 * C);
  std::vector<int> cams;
  for (int p = 0; p < P; ++p) {
    const int L = len[p];
    cams.clear();
    if (L * 2 > C) {  // dense track: sample without replacement
      std::vector<int> all(C);
      std::iota(all.begin(), all.end(), 0);
      std::shuffle(all.begin(), all.end(), rng);
      cams.assign(all.begin(), all.begin() + L);
    } else {
      const int home = (int)(std::lower_bound(cum.begin(), cum.end(), uni(rng) * acc) - cum.begin());
      cams.push_back(std::min(home, C - 1));
      double width = spread;
      int tries = 0;
      while ((int)cams.size() < L) {
        // Laplace-distributed offset around the home camera, on a ring
        const double u = uni(rng) - 0.5;
        const int off = (int)std::lround(-width * (u < 0 ? -1 : 1) * std::log(1 - 2 * std::fabs(u)));
        int c = ((cams[0] + off) % C + C) % C;
        if (std::find(cams.begin(), cams.end(), c) == cams.end()) cams.push_back(c);
        else if (++tries > 8) { width *= 2; tries = 0; }
      }
    }
    std::sort(cams.begin(), cams.end());
    std::copy(cams.begin(), cams.end(), cam_idx + point_row_start[p]);
  }
  // duplicates: the second observation of a point repeats the first's camera
  for (int k = 0, p = 0; k < dup_pairs && p < P; ++p) {
    if (len[p] >= 2) { cam_idx[point_row_start[p] + 1] = cam_idx[point_row_start[p]]; ++k; }
  }
  return 0;
}

// Scene at the scale BALProblem::Normalize() leaves (points with median absolute
// deviation ~100, examples/bal_problem.cc:249-292): cameras on a ring of radius
// 1000 looking at a point cloud around the origin; observations = projection of
// the true scene + pixel noise; the returned parameters are then perturbed
// (rotation / translation / point sigmas as in BALProblem::Perturb :294-333).
void bal_synth_geometry(int C, int P, long long N, const int* point_row_start,
                        const int* cam_idx, unsigned long long seed, double pixel_sigma,
                        double rotation_sigma, double translation_sigma, double point_sigma,
                        double* cameras, double* points, double* observations) {
  std::mt19937_64 rng(seed ^ 0x9e3779b97f4a7c15ULL);
  std::normal_distribution<double> gauss(0.0, 1.0);
  std::uniform_real_distribution<double> uni(0.0, 1.0);
  const double kPi = 3.14159265358979323846;
  const double radius = 1000.0;
  for (int c = 0; c < C; ++c) {
    double* cam = cameras + 9 * c;
    const double th = -kPi + 2.0 * kPi * (c + 0.5) / C;
    // world position on the ring; R = rot_y(-th) maps it onto +z so the scene
    // sits at negative camera z (Bundler convention)
    const double cw[3] = {radius * std::sin(th), 20.0 * gauss(rng), radius * std::cos(th)};
    double aa[3] = {0.02 * gauss(rng), -th + 0.02 * gauss(rng), 0.02 * gauss(rng)};
    double rc[3];
    RotateByAngleAxis<double>(aa, cw, rc);
    cam[0] = aa[0]; cam[1] = aa[1]; cam[2] = aa[2];
    cam[3] = -rc[0]; cam[4] = -rc[1]; cam[5] = -rc[2];
    cam[6] = 500.0 + 1000.0 * uni(rng);
    cam[7] = 1e-2 * gauss(rng);
    cam[8] = 1e-3 * gauss(rng);
  }
  for (int p = 0; p < P; ++p)
    for (int k = 0; k < 3; ++k) points[3 * p + k] = 120.0 * gauss(rng);
  std::vector<int> row_point(N);
  for (int p = 0; p < P; ++p)
    for (int r = point_row_start[p]; r < point_row_start[p + 1]; ++r) row_point[r] = p;
  for (long long r = 0; r < N; ++r) {
    double pred[2];
    Project<double>(cameras + 9 * cam_idx[r], points + 3 * row_point[r], pred);
    observations[2 * r] = pred[0] + pixel_sigma * gauss(rng);
    observations[2 * r + 1] = pred[1] + pixel_sigma * gauss(rng);
  }
  for (int c = 0; c < C; ++c) {
    double* cam = cameras + 9 * c;
    for (int k = 0; k < 3; ++k) cam[k] += rotation_sigma * gauss(rng);
    for (int k = 3; k < 6; ++k) cam[k] += translation_sigma * gauss(rng);
  }
  for (long long i = 0; i < 3LL * P; ++i) points[i] += point_sigma * gauss(rng);
}

// Residuals [2N] and, when E/F are non-null, the Jacobian blocks E [N x 6]
// (2x3 row-major, d residual / d point) and F [N x 18] (2x9 row-major,
// d residual / d camera). Returns the cost 0.5 * |residuals|^2.
double bal_evaluate(int P, long long N, const int* point_row_start, const int* cam_idx,
                    const double* cameras, const double* points, const double* observations,
                    double* residuals, double* E, double* F) {
  double cost = 0.0;
#pragma omp parallel for schedule(static) reduction(+ : cost)
  for (int p = 0; p < P; ++p) {
    for (int r = point_row_start[p]; r < point_row_start[p + 1]; ++r) {
      const double* cam = cameras + 9 * cam_idx[r];
      const double* pt = points + 3 * p;
      double res[2];
      if (E == nullptr) {
        double pred[2];
        Project<double>(cam, pt, pred);
        res[0] = pred[0] - observations[2 * r];
        res[1] = pred[1] - observations[2 * r + 1];
      } else {
        Dual dc[9], dp[3], pred[2];
        for (int k = 0; k < 9; ++k) dc[k] = Var(cam[k], k);
        for (int k = 0; k < 3; ++k) dp[k] = Var(pt[k], 9 + k);
        Project<Dual>(dc, dp, pred);
        for (int i = 0; i < 2; ++i) {
          res[i] = pred[i].v - observations[2 * r + i];
          for (int k = 0; k < 9; ++k) F[18LL * r + 9 * i + k] = pred[i].d[k];
          for (int k = 0; k < 3; ++k) E[6LL * r + 3 * i + k] = pred[i].d[9 + k];
        }
      }
      if (residuals) { residuals[2LL * r] = res[0]; residuals[2LL * r + 1] = res[1]; }
      cost += 0.5 * (res[0] * res[0] + res[1] * res[1]);
    }
  }
  return cost;
}

// ---- BAL problem files (bal_io.h) for the Python side ------------------------------------
// Two-step read: bal_file_open parses the file and returns a handle + the counts, the
// caller allocates, bal_file_fetch copies out and frees the handle. Non-zero = error
// (message via bal_file_error).
static thread_local std::string g_bal_error;
const char* bal_file_error() { return g_bal_error.c_str(); }

int bal_file_open(const char* filename, void** handle, int* num_cameras, int* num_points, int* num_observations) {
  auto* p = new cb200_bal::BalProblem;
  g_bal_error = cb200_bal::Read(filename, p);
  if (!g_bal_error.empty()) { delete p; return 1; }
  *handle = p;
  *num_cameras = p->num_cameras; *num_points = p->num_points; *num_observations = p->num_observations;
  return 0;
}
void bal_file_fetch(void* handle, int* camera_index, int* point_index, double* observations, double* cameras,
                    double* points) {
  auto* p = static_cast<cb200_bal::BalProblem*>(handle);
  std::copy(p->camera_index.begin(), p->camera_index.end(), camera_index);
  std::copy(p->point_index.begin(), p->point_index.end(), point_index);
  std::copy(p->observations.begin(), p->observations.end(), observations);
  std::copy(p->cameras.begin(), p->cameras.end(), cameras);
  std::copy(p->points.begin(), p->points.end(), points);
  delete p;
}
static cb200_bal::BalProblem MakeBal(int C, int P, int N, const int* camera_index, const int* point_index,
                                     const double* observations, const double* cameras, const double* points) {
  cb200_bal::BalProblem p;
  p.num_cameras = C; p.num_points = P; p.num_observations = N;
  if (camera_index) p.camera_index.assign(camera_index, camera_index + N);
  if (point_index) p.point_index.assign(point_index, point_index + N);
  if (observations) p.observations.assign(observations, observations + 2 * (size_t)N);
  p.cameras.assign(cameras, cameras + 9 * (size_t)C);
  p.points.assign(points, points + 3 * (size_t)P);
  return p;
}
int bal_file_write(const char* filename, int C, int P, int N, const int* camera_index, const int* point_index,
                   const double* observations, const double* cameras, const double* points) {
  g_bal_error = cb200_bal::Write(filename, MakeBal(C, P, N, camera_index, point_index, observations, cameras, points));
  return g_bal_error.empty() ? 0 : 1;
}
int bal_file_write_ply(const char* filename, int C, int P, const double* cameras, const double* points) {
  g_bal_error = cb200_bal::WritePly(filename, MakeBal(C, P, 0, nullptr, nullptr, nullptr, cameras, points));
  return g_bal_error.empty() ? 0 : 1;
}
void bal_normalize(int C, int P, double* cameras, double* points) {
  cb200_bal::BalProblem p = MakeBal(C, P, 0, nullptr, nullptr, nullptr, cameras, points);
  cb200_bal::Normalize(&p);
  std::copy(p.cameras.begin(), p.cameras.end(), cameras);
  std::copy(p.points.begin(), p.points.end(), points);
}
void bal_perturb(int C, int P, double rotation_sigma, double translation_sigma, double point_sigma, double* cameras,
                 double* points) {
  cb200_bal::BalProblem p = MakeBal(C, P, 0, nullptr, nullptr, nullptr, cameras, points);
  cb200_bal::Perturb(&p, rotation_sigma, translation_sigma, point_sigma);
  std::copy(p.cameras.begin(), p.cameras.end(), cameras);
  std::copy(p.points.begin(), p.points.end(), points);
}
// order[N], row_start[P + 1]: see cb200_bal::GroupByPoint
void bal_group_by_point(int P, int N, const int* point_index, int* order, int* row_start) {
  cb200_bal::BalProblem p;
  p.num_points = P; p.num_observations = N;
  p.point_index.assign(point_index, point_index + N);
  std::vector<int> o, r;
  cb200_bal::GroupByPoint(p, &o, &r);
  std::copy(o.begin(), o.end(), order);
  std::copy(r.begin(), r.end(), row_start);
}

}  // extern "C"
