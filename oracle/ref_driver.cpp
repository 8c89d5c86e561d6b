// ref_driver.cpp -- plain-pointer C entry points around the reference's OWN
// matcher headers, compiled where they lie under /root/reference/src.
//
// TEST INFRASTRUCTURE ONLY (oracle/_ref/libspectavi_ref.so; see Makefile).
// Nothing from the reference is copied into this repository: the headers are
// #included from the read-only reference tree at build time, Eigen is replaced
// by oracle/eigen_shim.  Each function mirrors the body of the corresponding
// extern "C" function in src/Spectavi.cpp minus the NdArray allocation
// protocol (the caller passes output buffers).
//
// Two interpositions, both on the cascade only:
//  * std::random_device -> a fixed-seed stand-in, because the reference seeds
//    its dictionary from the OS entropy pool (CascadingHashNn.h:87-88) and would
//    otherwise never produce the same output twice;
//  * `private` -> `public` while including CascadingHashNn.h so the generated
//    dictionary can be read back and handed to the oracle / the CUDA path.
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <algorithm>
#include <chrono>
#include <cmath>
#include <iostream>
#include <list>
#include <queue>
#include <random>
#include <stdexcept>
#include <unordered_map>
#include <unordered_set>
#include <vector>
#include <emmintrin.h>
#include <Eigen/Dense>

static unsigned g_ref_seed = 5489u;
namespace std {
struct spv_seeded_device {
  unsigned operator()() { return g_ref_seed; }
};
} // namespace std

#include "BruteForceNn.h"
#include "BruteForceNnL1K2.h"
#define random_device spv_seeded_device
#define private public
#include "CascadingHashNn.h"
#include "KMedians.h"
#undef private
#undef random_device

using namespace spectavi;

extern "C" {

// src/Spectavi.cpp:284-298
int ref_l1k2(const uint8_t *x, const uint8_t *y, int xrows, int yrows, int dim,
             int nthreads, uint64_t *idx, int32_t *dist) {
  try {
    RowMatrixXsMap _outidx(reinterpret_cast<size_t *>(idx), yrows, 2);
    RowMatrixXiMap _outdist(reinterpret_cast<int *>(dist), yrows, 2);
    BruteForceNnL1K2<> nn(x, y, xrows, yrows, dim);
    auto filter = filter::IdentityFilter(xrows);
    nn.find_neighbours<filter::IdentityFilter>(_outidx, _outdist, filter,
                                               nthreads);
  } catch (const std::exception &e) {
    return -1;
  }
  return 0;
}

// src/Spectavi.cpp:258-269
int ref_bruteforce_f32(const float *x, const float *y, int xrows, int yrows,
                       int dim, int k, float p, float mu, uint64_t *idx,
                       float *dist) {
  RowMatrixXsMap _outidx(reinterpret_cast<size_t *>(idx), yrows, k);
  RowMatrixXfMap _outdist(dist, yrows, k);
  BruteForceNn<> nn(x, y, xrows, yrows, dim, p, mu);
  nn.find_neighbours(_outidx, _outdist, k);
  return 0;
}

// src/Spectavi.cpp:271-282
int ref_bruteforce_i32(const int *x, const int *y, int xrows, int yrows,
                       int dim, int k, float p, float mu, uint64_t *idx,
                       int *dist) {
  RowMatrixXsMap _outidx(reinterpret_cast<size_t *>(idx), yrows, k);
  RowMatrixXiMap _outdist(dist, yrows, k);
  BruteForceNn<RowMatrixXi> nn(x, y, xrows, yrows, dim, p, mu);
  nn.find_neighbours(_outidx, _outdist, k);
  return 0;
}

// src/Spectavi.cpp:321-336 with the dictionary seed pinned.
// dict_out (optional): [n][dim][m] float32, the dictionary the run used.
// codes_out (optional): [xrows][n] int32 m_hashcodes_x.
// ncand_out (optional): [yrows] distinct candidate count per query.
int ref_cascade(const float *x, const float *y, int xrows, int yrows, int dim,
                int k, int m, int n, int g, unsigned seed, uint64_t *idx,
                float *dist, float *dict_out, int32_t *codes_out,
                int32_t *ncand_out) {
  try {
    g_ref_seed = seed;
    CascadingHashNn<> nn(x, y, xrows, yrows, dim, m, n, g);
    RowMatrixXsMap _outidx(reinterpret_cast<size_t *>(idx), yrows, k);
    RowMatrixXfMap _outdist(dist, yrows, k);
    nn.find_neighbours(_outidx, _outdist);
    if (dict_out)
      for (int t = 0; t < n; ++t)
        for (int i = 0; i < dim; ++i)
          for (int j = 0; j < m; ++j)
            dict_out[((size_t)t * dim + i) * m + j] = nn.m_hash_dicts[t](i, j);
    if (codes_out)
      for (int i = 0; i < xrows; ++i)
        for (int t = 0; t < n; ++t)
          codes_out[(size_t)i * n + t] = nn.m_hashcodes_x(i, t);
    if (ncand_out)
      for (int r = 0; r < yrows; ++r) {
        std::unordered_set<int> s;
        nn.filter_potential_neighbours(r, s);
        ncand_out[r] = (int32_t)s.size();
      }
  } catch (const std::exception &e) {
    return -1;
  }
  return 0;
}

// candidate id set of one query, sorted ascending (for candidate-rule parity)
int ref_cascade_candidates(const float *x, const float *y, int xrows,
                           int yrows, int dim, int m, int n, int g,
                           unsigned seed, int row, int32_t *cand_out,
                           int cap) {
  try {
    g_ref_seed = seed;
    CascadingHashNn<> nn(x, y, xrows, yrows, dim, m, n, g);
    std::unordered_set<int> s;
    nn.filter_potential_neighbours(row, s);
    std::vector<int> v(s.begin(), s.end());
    std::sort(v.begin(), v.end());
    int c = (int)v.size() < cap ? (int)v.size() : cap;
    for (int i = 0; i < c; ++i) cand_out[i] = v[i];
    return (int)v.size();
  } catch (const std::exception &e) {
    return -1;
  }
}

// BruteForceNn::find_neighbours(out_idx, out_dist, filter_x, filter_y, k),
// src/BruteForceNn.h:37-120 -- the overload KMedians drives.  idx / dist are
// IN-OUT: only the rows named by filter_y are written (all rows when nfy == 0).
int ref_bruteforce_filtered_f32(const float *x, const float *y, int xrows,
                                int yrows, int dim, int k, float p, float mu,
                                const uint64_t *fx, int nfx,
                                const uint64_t *fy, int nfy, uint64_t *idx,
                                float *dist) {
  RowMatrixXsMap _outidx(reinterpret_cast<size_t *>(idx), yrows, k);
  RowMatrixXfMap _outdist(dist, yrows, k);
  std::unordered_set<size_t> sx(fx, fx + nfx), sy(fy, fy + nfy);
  BruteForceNn<> nn(x, y, xrows, yrows, dim, p, mu);
  nn.find_neighbours(_outidx, _outdist, sx, sy, k);
  return 0;
}

// src/Spectavi.cpp:305-319 (nn_kmedians) with the shuffle seed pinned, the
// reference's y-side row-count slip (:309 passes xrows) kept out by requiring
// the caller's yrows, and the clustering read back so that the GPU path and the
// oracle can be given the SAME clusters: labels (m_point_to_median), medians
// (m_medians), and nothing else.  idx / dist are IN-OUT like above.
int ref_kmedians_nn(const float *x, const float *y, int xrows, int yrows,
                    int dim, int nmx, int nmy, int c, int k, unsigned seed,
                    uint64_t *idx, float *dist, int32_t *xlab, int32_t *ylab,
                    float *xmed, float *ymed, double *seconds_fnn) {
  try {
    g_ref_seed = seed;
    KMedians<> kmedx(x, xrows, dim, nmx);
    kmedx.run();
    g_ref_seed = seed + 1;
    KMedians<> kmedy(y, yrows, dim, nmy);
    kmedy.run();
    RowMatrixXsMap _outidx(reinterpret_cast<size_t *>(idx), yrows, k);
    RowMatrixXfMap _outdist(dist, yrows, k);
    auto t0 = std::chrono::steady_clock::now();
    kmedy.find_nearest_neighbours(kmedx, _outidx, _outdist, c, k);
    if (seconds_fnn)   // the matching half alone (KMedians.h:260-295), for the CPU baseline
      *seconds_fnn = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    for (int g = 0; g < nmx; ++g)
      for (auto &pid : kmedx.m_median_idx[g]) xlab[pid] = g;
    for (int g = 0; g < nmy; ++g)
      for (auto &pid : kmedy.m_median_idx[g]) ylab[pid] = g;
    for (int g = 0; g < nmx; ++g)
      for (int j = 0; j < dim; ++j) xmed[(size_t)g * dim + j] = kmedx.m_medians(g, j);
    for (int g = 0; g < nmy; ++g)
      for (int j = 0; j < dim; ++j) ymed[(size_t)g * dim + j] = kmedy.m_medians(g, j);
  } catch (const std::exception &e) {
    return -1;
  }
  return 0;
}
}
