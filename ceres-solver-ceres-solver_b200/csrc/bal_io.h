// Bundle Adjustment in the Large (BAL) problem files: read, write, PLY export,
// Normalize and Perturb, with the semantics of the reference's example loader
// (examples/bal_problem.cc:73-333; angle-axis cameras only). Host-only, header-only;
// used by tools/bundle_adjuster.cc and exported to Python through libsynthetic_bal.so.
//
// File format (bal_problem.cc:83-108): "C P N", then N lines "camera point x y",
// then 9 numbers per camera (angle-axis 3, translation 3, focal, k1, k2) and 3 per
// point, one per line.
#ifndef CB200_BAL_IO_H_
#define CB200_BAL_IO_H_

#include <algorithm>
#include <cstdio>
#include <functional>
#include <numeric>
#include <random>
#include <string>
#include <vector>

#include "snavely_model.h"

namespace cb200_bal {

struct BalProblem {
  int num_cameras = 0, num_points = 0, num_observations = 0;
  std::vector<int> camera_index, point_index;  // per observation, file order
  std::vector<double> observations;            // 2 per observation
  std::vector<double> cameras;                 // 9 per camera
  std::vector<double> points;                  // 3 per point
};

// Returns an empty string on success, else what went wrong (the reference aborts:
// "This will die horribly on invalid files", bal_problem.cc:82).
inline std::string Read(const std::string& filename, BalProblem* p) {
  FILE* f = std::fopen(filename.c_str(), "r");
  if (f == nullptr) return "unable to open file " + filename;
  auto fail = [&](const std::string& what) { std::fclose(f); return "invalid BAL file " + filename + ": " + what; };
  if (std::fscanf(f, "%d %d %d", &p->num_cameras, &p->num_points, &p->num_observations) != 3) return fail("header");
  if (p->num_cameras <= 0 || p->num_points <= 0 || p->num_observations <= 0) return fail("non-positive counts");
  const int N = p->num_observations;
  p->camera_index.resize(N);
  p->point_index.resize(N);
  p->observations.resize(2 * (size_t)N);
  for (int i = 0; i < N; ++i) {
    if (std::fscanf(f, "%d %d %lf %lf", &p->camera_index[i], &p->point_index[i], &p->observations[2 * (size_t)i],
                    &p->observations[2 * (size_t)i + 1]) != 4)
      return fail("observation " + std::to_string(i));
    if (p->camera_index[i] < 0 || p->camera_index[i] >= p->num_cameras || p->point_index[i] < 0 ||
        p->point_index[i] >= p->num_points)
      return fail("index out of range in observation " + std::to_string(i));
  }
  p->cameras.resize(9 * (size_t)p->num_cameras);
  p->points.resize(3 * (size_t)p->num_points);
  for (double& v : p->cameras)
    if (std::fscanf(f, "%lf", &v) != 1) return fail("camera parameters");
  for (double& v : p->points)
    if (std::fscanf(f, "%lf", &v) != 1) return fail("point parameters");
  std::fclose(f);
  return "";
}

// Same text format and precision as BALProblem::WriteToFile (bal_problem.cc:146-185).
inline std::string Write(const std::string& filename, const BalProblem& p) {
  FILE* f = std::fopen(filename.c_str(), "w");
  if (f == nullptr) return "unable to open file " + filename;
  std::fprintf(f, "%d %d %d\n", p.num_cameras, p.num_points, p.num_observations);
  for (int i = 0; i < p.num_observations; ++i)
    std::fprintf(f, "%d %d %g %g\n", p.camera_index[i], p.point_index[i], p.observations[2 * (size_t)i],
                 p.observations[2 * (size_t)i + 1]);
  for (double v : p.cameras) std::fprintf(f, "%.16g\n", v);
  for (double v : p.points) std::fprintf(f, "%.16g\n", v);
  std::fclose(f);
  return "";
}

// c = -R' t  (bal_problem.cc:226-240)
inline void CameraToAngleAxisAndCenter(const double* camera, double* angle_axis, double* center) {
  double inverse_rotation[3];
  for (int k = 0; k < 3; ++k) { angle_axis[k] = camera[k]; inverse_rotation[k] = -camera[k]; }
  cb200_model::RotateByAngleAxis(inverse_rotation, camera + 3, center);
  for (int k = 0; k < 3; ++k) center[k] = -center[k];
}
// t = -R c  (bal_problem.cc:242-255)
inline void AngleAxisAndCenterToCamera(const double* angle_axis, const double* center, double* camera) {
  for (int k = 0; k < 3; ++k) camera[k] = angle_axis[k];
  double t[3];
  cb200_model::RotateByAngleAxis(angle_axis, center, t);
  for (int k = 0; k < 3; ++k) camera[3 + k] = -t[k];
}

// Camera centres in green, points in white (bal_problem.cc:188-224).
inline std::string WritePly(const std::string& filename, const BalProblem& p) {
  FILE* f = std::fopen(filename.c_str(), "w");
  if (f == nullptr) return "unable to open file " + filename;
  std::fprintf(f,
               "ply\nformat ascii 1.0\nelement vertex %d\nproperty float x\nproperty float y\nproperty float z\n"
               "property uchar red\nproperty uchar green\nproperty uchar blue\nend_header\n",
               p.num_cameras + p.num_points);
  for (int i = 0; i < p.num_cameras; ++i) {
    double aa[3], c[3];
    CameraToAngleAxisAndCenter(&p.cameras[9 * (size_t)i], aa, c);
    std::fprintf(f, "%g %g %g 0 255 0\n", c[0], c[1], c[2]);
  }
  for (int i = 0; i < p.num_points; ++i)
    std::fprintf(f, "%g %g %g 255 255 255\n", p.points[3 * (size_t)i], p.points[3 * (size_t)i + 1],
                 p.points[3 * (size_t)i + 2]);
  std::fclose(f);
  return "";
}

inline double Median(std::vector<double>* data) {  // bal_problem.cc:65-69 (upper median)
  auto mid = data->begin() + data->size() / 2;
  std::nth_element(data->begin(), mid, data->end());
  return *mid;
}

// Centre the points on their marginal median and scale so that the median absolute
// (L1) deviation is 100; move the camera centres along (bal_problem.cc:257-300).
inline void Normalize(BalProblem* p) {
  const int P = p->num_points;
  std::vector<double> tmp(P);
  double median[3];
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < P; ++j) tmp[j] = p->points[3 * (size_t)j + i];
    median[i] = Median(&tmp);
  }
  for (int j = 0; j < P; ++j) {
    const double* x = &p->points[3 * (size_t)j];
    tmp[j] = std::fabs(x[0] - median[0]) + std::fabs(x[1] - median[1]) + std::fabs(x[2] - median[2]);
  }
  const double scale = 100.0 / Median(&tmp);
  for (int j = 0; j < P; ++j)
    for (int i = 0; i < 3; ++i) p->points[3 * (size_t)j + i] = scale * (p->points[3 * (size_t)j + i] - median[i]);
  for (int c = 0; c < p->num_cameras; ++c) {
    double* camera = &p->cameras[9 * (size_t)c];
    double aa[3], center[3];
    CameraToAngleAxisAndCenter(camera, aa, center);
    for (int i = 0; i < 3; ++i) center[i] = scale * (center[i] - median[i]);
    AngleAxisAndCenterToCamera(aa, center, camera);
  }
}

// Gaussian noise on points, camera rotations (angle-axis) and translations from a
// default-seeded std::mt19937, in the reference's order of draws
// (bal_problem.cc:302-340). As there, the ROTATION noise is drawn with point_sigma
// as its standard deviation (:318-319) and only applied when rotation_sigma > 0.
inline void Perturb(BalProblem* p, double rotation_sigma, double translation_sigma, double point_sigma) {
  std::mt19937 prng;
  // The reference perturbs each 3-vector through std::bind(distribution, std::ref(prng)), which
  // COPIES the distribution: the spare value std::normal_distribution keeps between calls is
  // dropped after every 3-vector. The same here, so the draws are the same.
  auto perturb3 = [&prng](std::normal_distribution<double> dist, double* v) {
    for (int k = 0; k < 3; ++k) v[k] += dist(prng);
  };
  std::normal_distribution<double> point_noise(0.0, point_sigma);
  if (point_sigma > 0)
    for (int i = 0; i < p->num_points; ++i) perturb3(point_noise, &p->points[3 * (size_t)i]);
  std::normal_distribution<double> rotation_noise(0.0, point_sigma);
  std::normal_distribution<double> translation_noise(0.0, translation_sigma);
  for (int c = 0; c < p->num_cameras; ++c) {
    double* camera = &p->cameras[9 * (size_t)c];
    double aa[3], center[3];
    CameraToAngleAxisAndCenter(camera, aa, center);
    if (rotation_sigma > 0.0) perturb3(rotation_noise, aa);
    AngleAxisAndCenterToCamera(aa, center, camera);
    if (translation_sigma > 0.0) perturb3(translation_noise, camera + 3);
  }
}

// Row order of the Schur-ordered Program: observations grouped by point (points in
// file order, the observations of a point in file order), which is what
// reorder_program.cc:536-539 arranges for the eliminator. order[r] = index of the
// observation that becomes row r; row_start[P + 1] = first row of each point.
inline void GroupByPoint(const BalProblem& p, std::vector<int>* order, std::vector<int>* row_start) {
  row_start->assign((size_t)p.num_points + 1, 0);
  for (int i = 0; i < p.num_observations; ++i) ++(*row_start)[p.point_index[i] + 1];
  for (int j = 0; j < p.num_points; ++j) (*row_start)[j + 1] += (*row_start)[j];
  std::vector<int> next(row_start->begin(), row_start->end() - 1);
  order->resize(p.num_observations);
  for (int i = 0; i < p.num_observations; ++i) (*order)[next[p.point_index[i]]++] = i;
}

}  // namespace cb200_bal

#endif  // CB200_BAL_IO_H_
