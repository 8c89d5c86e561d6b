// ref_mvg_driver.cpp -- plain-pointer C entry points around the reference's OWN RANSAC scoring code
// (src/RansacFitter.h:42-98 process_fundamental_matrix, src/DltTriangulator.h, src/Camera.h:32-59
// Essential2Cameras, and the seven-point solver of src/FundamentalMatrixFitter.h as a hypothesis
// generator), compiled where they lie under /root/reference/src.
//
// TEST INFRASTRUCTURE ONLY (oracle/_ref/libspectavi_ref_mvg.so; see Makefile).  Nothing from the
// reference is copied into this repository: the headers are #included from the read-only tree at build
// time; Eigen is replaced by oracle/eigen_shim_mvg, whose JacobiSVD is a restatement of Eigen's
// published two-sided Jacobi algorithm (parity with a real Eigen build: unpinned, see that header).
// One interposition: `private` -> `public` so that process_fundamental_matrix -- the scoring step, a
// private member reached only from fit_essential's random sampling loop -- can be driven with given
// fundamental matrices.  NdArray.h (included by EigenDefinitions.h) comes from the reference tree.
#include <chrono>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <random>
#include <stdexcept>
#include <unordered_set>
#include <vector>
#include <Eigen/Dense>

#define private public
#include "RansacFitter.h"
#undef private

using namespace spectavi;

extern "C" {

// the restated JacobiSVD itself (square, n <= 16), for direct comparison with LAPACK
void ref_mvg_jacobi_svd(const double *a, int n, double *sv, double *u, double *v) {
  std::vector<double> w(a, a + (size_t)n * n);
  Eigen::internal::jacobi_svd_square(w.data(), n, sv, u, v);
}

// src/Spectavi.cpp:14-36
int ref_mvg_seven_point(const double *x, const double *xp, double *dst) {
  FundamentalMatrixFitter<RowMatrixXd> fitter;
  for (int i = 0; i < 7; ++i) fitter.add_putative_match(x[2 * i], x[2 * i + 1], xp[2 * i], xp[2 * i + 1]);
  RowMatrixXd Fs[3];
  const int nroots = fitter.solve(Fs[0], Fs[1], Fs[2]);
  for (int r = 0; r < nroots; ++r) std::copy(Fs[r].data(), Fs[r].data() + 9, dst + 9 * r);
  return nroots;
}

// src/Spectavi.cpp:54-68
void ref_mvg_dlt_reprojection_error(const double *P0, const double *P1, int npt, const double *x, const double *xp,
                                    double *err, double *X, int *infront) {
  DltTriangulator<RowMatrixXd> dlt(P0, P1);
  for (int i = 0; i < npt; ++i) {
    dlt.solve(x + 3 * i, xp + 3 * i, 3);
    err[i] = dlt.reprojection_error();
    if (infront) infront[i] = dlt.is_infront_both_cameras() ? 1 : 0;
    if (X) { RowMatrixXd xx = dlt.X(); for (int k = 0; k < 4; ++k) X[4 * i + k] = xx(k); }
  }
}

// RansacFitter::process_fundamental_matrix for nF matrices (F: nF x 9 row-major).  Per matrix: ok (its return
// value), inlier count, winning camera (12 doubles) and -- when masks != nullptr -- npt bytes, 1 = inlier of
// the winning camera.  nthreads: OpenMP over the matrices, as fit_essential parallelises over its tries
// (RansacFitter.h:172).  Returns the wall seconds of the scoring loop.
double ref_mvg_process(const double *x0, const double *x1, int npt, double required_percent_inliers,
                       double reprojection_error_allowed, int find_best_even_in_failure,
                       double singular_value_ratio_allowed, const double *F, int nF, int nthreads, int *ok,
                       int *ninlier, double *cams, uint8_t *masks) {
  RansacFitter<> fitter(x0, x1, npt, required_percent_inliers, reprojection_error_allowed, 0,
                        find_best_even_in_failure != 0);
  const auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 1)
  for (int f = 0; f < nF; ++f) {
    RowMatrixXd Fm(3, 3);
    for (int k = 0; k < 9; ++k) Fm(k) = F[(size_t)9 * f + k];
    int cnt = 0;
    Camera<RowMatrixXd> cam;
    RowMatrixXi idx;
    const bool s = fitter.process_fundamental_matrix(Fm, singular_value_ratio_allowed, cnt, cam, idx);
    ok[f] = s ? 1 : 0;
    ninlier[f] = s ? cnt : 0;
    for (int k = 0; k < 12; ++k) cams[(size_t)12 * f + k] = s ? cam.get_matrix()(k) : 0.0;
    if (masks) {
      std::memset(masks + (size_t)npt * f, 0, (size_t)npt);
      if (s) for (int k = 0; k < cnt; ++k) masks[(size_t)npt * f + idx(k)] = 1;
    }
  }
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

}  // extern "C"
