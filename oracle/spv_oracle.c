/*
 * spv_oracle.c -- CPU restatement of spectavi's descriptor-matching hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the parity checker and the timed CPU
 * baseline.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it.  The product (libspectavi.so) never links,
 * imports or falls back to anything in oracle/.
 *
 * Every function cites the reference file:line it restates (paths relative to
 * the spectavi source tree).  Parity pin: this restatement is checked (tests/
 * test_oracle_pin.py) against
 *   - the reference's own headers compiled from /root/reference through the
 *     Eigen shim in oracle/eigen_shim (oracle/_ref/libspectavi_ref.so), and the
 *     golden vectors generated from that build (tests/golden/), and
 *   - an independent numpy stable (dist, idx) sort (oracle/numpy_oracle.py),
 * on the reference's own test shapes (test/test_feature.py:67-81,102-151).
 *
 * Build: gcc -O3 -DNDEBUG -fopenmp -shared -fPIC (no -march: the reference
 * builds with plain Release flags, CMakeLists.txt:12, i.e. SSE2 only).
 */
#include <emmintrin.h>
#include <limits.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define SPV_NONE_IDX UINT64_MAX /* size_t(-1), BruteForceNnL1K2.h:102-103 */

/* ---- src/BruteForceNnL1K2.h:43-48 (implementation::sad_16) ----------------
 * One psadbw over 16 byte lanes, then the two 64-bit lane sums are added.
 * The reference uses the aligned load; numpy buffers happen to satisfy it.  The
 * oracle uses the unaligned form so that sliced inputs work too -- same sums. */
static inline int sad16(const uint8_t *a, const uint8_t *b) {
  __m128i va = _mm_loadu_si128((const __m128i *)a);
  __m128i vb = _mm_loadu_si128((const __m128i *)b);
  __m128i s = _mm_sad_epu8(va, vb);
  return _mm_cvtsi128_si32(s) + _mm_extract_epi16(s, 4);
}

/* State of one query row's running top-2, BruteForceNnL1K2.h:95-105. */
typedef struct {
  int32_t d0, d1;
  uint64_t i0, i1;
  int32_t worst; /* -1 until two neighbours are known (:105,:140-142) */
} top2_t;

static inline void top2_reset(top2_t *t) {
  t->d0 = INT_MAX;
  t->d1 = INT_MAX;
  t->i0 = SPV_NONE_IDX;
  t->i1 = SPV_NONE_IDX;
  t->worst = -1;
}

/* One candidate x row against one y row: BruteForceNnL1K2.h:110-142.
 * prune != 0 keeps the reference's early exit (:118-121, strict '>');
 * prune == 0 computes the full distance (what the GPU does).  Both give the
 * same top-2 because an abandoned candidate has dist > second-best and the
 * updates below use strict '<'. */
static inline void top2_offer(top2_t *t, const uint8_t *xr, const uint8_t *yr,
                              int ngroups, uint64_t label, int prune) {
  int32_t acc = 0;
  for (int g = 0; g < ngroups; ++g) {
    acc += sad16(xr + 16 * g, yr + 16 * g);
    if (prune && t->worst >= 0 && acc > t->worst) return;
  }
  if (acc < t->d0) { /* :129-134 shift first -> second */
    t->d1 = t->d0;
    t->i1 = t->i0;
    t->d0 = acc;
    t->i0 = label;
  } else if (acc < t->d1) { /* :135-139 */
    t->d1 = acc;
    t->i1 = label;
  }
  if (t->i1 != SPV_NONE_IDX) t->worst = t->d1; /* :140-142 */
}

/* ---- src/BruteForceNnL1K2.h:84-145 with filter::IdentityFilter (:19-38) ----
 * and the entry point src/Spectavi.cpp:284-298.
 * Returns 0, or -1 where the reference constructor throws (:74-81). */
int oracle_l1k2(const uint8_t *x, const uint8_t *y, int xrows, int yrows,
                int dim, int nthreads, int prune, uint64_t *out_idx,
                int32_t *out_dist) {
  if (dim % 16 != 0) return -1;
  const int ngroups = dim / 16;
  if (nthreads < 1) nthreads = 1;
#pragma omp parallel for num_threads(nthreads) schedule(static)
  for (int r = 0; r < yrows; ++r) {
    top2_t t;
    top2_reset(&t);
    const uint8_t *yr = y + (size_t)r * dim;
    for (int c = 0; c < xrows; ++c) /* IdentityFilter: 0,1,...,xrows-1 */
      top2_offer(&t, x + (size_t)c * dim, yr, ngroups, (uint64_t)c, prune);
    out_idx[2 * (size_t)r + 0] = t.i0;
    out_idx[2 * (size_t)r + 1] = t.i1;
    out_dist[2 * (size_t)r + 0] = t.d0;
    out_dist[2 * (size_t)r + 1] = t.d1;
  }
  return 0;
}

/* Same scan restricted to an explicit candidate list per y row (CSR):
 * candidates of row r are cand[off[r] .. off[r+1]) and are visited in the
 * order given.  This is BruteForceNnL1K2.h:84-145 driven by a filter whose
 * operator() yields that list (filter::SetFilter, CascadingHashNn.h:22-50). */
int oracle_l1k2_filtered(const uint8_t *x, const uint8_t *y, int xrows,
                         int yrows, int dim, const int64_t *off,
                         const int32_t *cand, int nthreads, uint64_t *out_idx,
                         int32_t *out_dist) {
  (void)xrows;
  if (dim % 16 != 0) return -1;
  const int ngroups = dim / 16;
  if (nthreads < 1) nthreads = 1;
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 64)
  for (int r = 0; r < yrows; ++r) {
    top2_t t;
    top2_reset(&t);
    const uint8_t *yr = y + (size_t)r * dim;
    for (int64_t p = off[r]; p < off[r + 1]; ++p)
      top2_offer(&t, x + (size_t)cand[p] * dim, yr, ngroups, (uint64_t)cand[p],
                 1);
    out_idx[2 * (size_t)r + 0] = t.i0;
    out_idx[2 * (size_t)r + 1] = t.i1;
    out_dist[2 * (size_t)r + 0] = t.d0;
    out_dist[2 * (size_t)r + 1] = t.d1;
  }
  return 0;
}

/* ---- src/BruteForceNn.h:37-120, no-filter overload (:31-35) -----------------
 * Generic p-norm k-NN.  The reference keeps the k best in a
 * std::priority_queue<pair<Scalar,int>> (max-heap on the pair, :54-56); only
 * its top() (the lexicographically largest kept pair) and its size are ever
 * observed, so a sorted array gives identical behaviour.
 *
 * Documented divergence: when an unpruned candidate has
 * distp == worst_dist the reference falls into its 'else' (:102-108), pushes
 * without popping, and later stores to column -1.  That path is undefined
 * behaviour; here such a candidate is rejected, which is what a lexicographic
 * (dist, idx) top-k does because the scan is ascending in idx.
 *
 * kind: 0 = float32 data (nn_bruteforce, Spectavi.cpp:258-269)
 *       1 = int32 data   (nn_bruteforcei, Spectavi.cpp:271-282) -- diff is
 *           float, pval is truncated to int, the sum is int (:68-79).
 * Fewer than k candidates: the heap is written back to front (:111-118) so the
 * first k-count columns keep whatever the caller put there (numpy zeros). */
typedef struct {
  double d; /* holds a float or an int32 exactly */
  int i;
} kpair_t;

static inline int kpair_less(const kpair_t *a, const kpair_t *b) {
  return a->d < b->d || (a->d == b->d && a->i < b->i);
}

/* insert keeping ascending order; returns new count */
static int kset_push(kpair_t *s, int n, kpair_t v) {
  int p = n;
  while (p > 0 && kpair_less(&v, &s[p - 1])) {
    s[p] = s[p - 1];
    --p;
  }
  s[p] = v;
  return n + 1;
}

static inline float pval_f(float diff, double p) { /* :70-78 */
  if (p == 1) return fabsf(diff);
  if (p == 2) return diff * diff;
  if (p == .5) return sqrtf(fabsf(diff));
  return (float)pow((double)fabsf(diff), p);
}

int oracle_bruteforce(const void *xv, const void *yv, int xrows, int yrows,
                      int dim, int k, float p_in, float mu_in, int kind,
                      uint64_t *out_idx, void *out_dist) {
  const double p = (double)p_in; /* BruteForceNn ctor takes double p (:28) */
  if (k < 1) return -1;
  kpair_t *set = (kpair_t *)malloc(sizeof(kpair_t) * (size_t)(k + 2));
  if (!set) return -2;
  for (int r = 0; r < yrows; ++r) {
    int n = 0;
    double worst = -1; /* Scalar worst_dist = -1 (:60) */
    for (int c = 0; c < xrows; ++c) {
      int pruned = 0;
      double distp;
      if (kind == 0) {
        const float *x = (const float *)xv + (size_t)c * dim;
        const float *y = (const float *)yv + (size_t)r * dim;
        const float mu = mu_in;
        float acc = 0.f;
        for (int j = 0; j < dim; ++j) {
          float diff = x[j] - y[j];
          acc += pval_f(diff, p);
          if (worst >= 0 && (acc + mu * (float)(dim - (j + 1)) > (float)worst)) {
            pruned = 1;
            break;
          }
        }
        distp = acc;
      } else {
        const int32_t *x = (const int32_t *)xv + (size_t)c * dim;
        const int32_t *y = (const int32_t *)yv + (size_t)r * dim;
        const int32_t mu = (int32_t)mu_in; /* Scalar m_mu, Scalar = int */
        int32_t acc = 0;
        for (int j = 0; j < dim; ++j) {
          float diff = (float)(x[j] - y[j]);
          /* Scalar pval = int: the float (or, for pow, double) value is
           * truncated on assignment (:69-78) */
          if (p == 1 || p == 2 || p == .5)
            acc += (int32_t)pval_f(diff, p);
          else
            acc += (int32_t)pow((double)fabsf(diff), p);
          if (worst >= 0 && (acc + mu * (dim - (j + 1)) > (int32_t)worst)) {
            pruned = 1;
            break;
          }
        }
        distp = acc;
      }
      if (pruned) continue;
      kpair_t v = {distp, c};
      if (worst >= 0 && distp < worst) { /* :93-101 */
        if (n >= k) --n;                 /* pop the largest */
        n = kset_push(set, n, v);
        worst = set[n - 1].d;
      } else if (worst < 0) { /* :102-108, heap not yet full */
        n = kset_push(set, n, v);
        if (n == k) worst = set[n - 1].d;
      } /* else distp == worst: rejected, see header comment */
    }
    /* :111-118 -- largest goes to column k-1, and so on downwards */
    for (int q = 0; q < n; ++q) {
      int col = k - n + q;
      out_idx[(size_t)r * k + col] = (uint64_t)set[q].i;
      if (kind == 0)
        ((float *)out_dist)[(size_t)r * k + col] = (float)set[q].d;
      else
        ((int32_t *)out_dist)[(size_t)r * k + col] = (int32_t)set[q].d;
    }
  }
  free(set);
  return 0;
}

/* BruteForceNn::find_neighbours(out_idx, out_dist, filter_x, filter_y, k),
 * src/BruteForceNn.h:37-120 -- the overload KMedians::find_nearest_neighbours
 * drives (src/KMedians.h:260-295).  float32 rows.  The reference walks two
 * std::unordered_set<size_t>, whose iteration order is unspecified; here the
 * filters are explicit lists visited in the order given.  The order only
 * matters when two candidates tie in distance (see the tie note above), so on
 * tie-free data the output equals the reference's for any order.
 *   nfx == 0: every train row is a candidate (filter_x.empty(), :46)
 *   nfy == 0: every query row is processed   (filter_y.empty(), :45)
 * out_idx / out_dist are IN-OUT: only processed rows are written, and a row
 * with n < k candidates only gets its last n columns written (:111-118). */
int oracle_bruteforce_filtered(const float *x, const float *y, int xrows,
                               int yrows, int dim, int k, float p_in,
                               float mu, const uint64_t *fx, int nfx,
                               const uint64_t *fy, int nfy, uint64_t *out_idx,
                               float *out_dist) {
  const double p = (double)p_in;
  if (k < 1) return -1;
  kpair_t *set = (kpair_t *)malloc(sizeof(kpair_t) * (size_t)(k + 2));
  if (!set) return -2;
  const int ylimit = nfy ? nfy : yrows, xlimit = nfx ? nfx : xrows;
  for (int ir = 0; ir < ylimit; ++ir) {
    const size_t r = nfy ? (size_t)fy[ir] : (size_t)ir;
    const float *yr = y + r * dim;
    int n = 0;
    double worst = -1;
    for (int ic = 0; ic < xlimit; ++ic) {
      const size_t c = nfx ? (size_t)fx[ic] : (size_t)ic;
      const float *xr = x + c * dim;
      int pruned = 0;
      float acc = 0.f;
      for (int j = 0; j < dim; ++j) {
        float diff = xr[j] - yr[j];
        acc += pval_f(diff, p);
        if (worst >= 0 && (acc + mu * (float)(dim - (j + 1)) > (float)worst)) {
          pruned = 1;
          break;
        }
      }
      if (pruned) continue;
      kpair_t v = {(double)acc, (int)c};
      if (worst >= 0 && acc < worst) {
        if (n >= k) --n;
        n = kset_push(set, n, v);
        worst = set[n - 1].d;
      } else if (worst < 0) {
        n = kset_push(set, n, v);
        if (n == k) worst = set[n - 1].d;
      }
    }
    for (int q = 0; q < n; ++q) {
      int col = k - n + q;
      out_idx[r * k + col] = (uint64_t)set[q].i;
      out_dist[r * k + col] = (float)set[q].d;
    }
  }
  free(set);
  (void)xrows;
  return 0;
}

/* ---- src/CascadingHashNn.h ---------------------------------------------------
 * The dictionary (generate_hash_dict, :86-100) is an INPUT here: the reference
 * draws it from std::random_device, so no two reference runs agree; tests feed
 * the same dictionary to this oracle and to the CUDA path.
 * dict layout: [table][dim][m] float32 row-major, as hash_dict(i, j) (:93-97).
 */

/* R = in * dict for one row and one table (:119, :140), accumulated in the
 * natural column order in float32.  With an integer-valued input and a
 * dictionary quantised to multiples of 2^-5 every partial sum
 * is exactly representable, so the summation order cannot matter. */
static void project_row(const float *row, const float *dict, int dim, int m,
                        float *out) {
  for (int b = 0; b < m; ++b) out[b] = 0.f;
  for (int j = 0; j < dim; ++j) {
    const float v = row[j];
    const float *dr = dict + (size_t)j * m;
    for (int b = 0; b < m; ++b) out[b] += v * dr[b];
  }
}

/* convert_to_hashcode, :102-111: bit i set iff proj(i) >= 0 */
static int32_t code_of(const float *proj, int m) {
  int32_t c = 0;
  for (int b = 0; b < m; ++b)
    if (proj[b] >= 0) c |= (int32_t)(1u << b);
  return c;
}

/* generate_hashcodes_for_data, :125-135.  codes: [rows][n] int32 row-major
 * (m_hashcodes_x). */
int oracle_hashcodes(const float *in, int rows, int dim, const float *dict,
                     int m, int n, int32_t *codes) {
  if (m < 1 || m > 31) return -1;
  float proj[32];
  for (int r = 0; r < rows; ++r)
    for (int t = 0; t < n; ++t) {
      project_row(in + (size_t)r * dim, dict + (size_t)t * dim * m, dim, m,
                  proj);
      codes[(size_t)r * n + t] = code_of(proj, m);
    }
  return 0;
}

/* generate_y_candidate_hashcodes, :150-185.  The reference keeps the `cutoff`
 * smallest (|proj|, bit) pairs with a max-heap of pairs (:153-159); popping it
 * (:166-169) lists those bits from the largest pair down.  Probe ii sets the
 * q-th listed bit to bit q of ii (:171-179).  Emits 1<<cutoff codes. */
static void probe_codes(const float *proj, int m, int cutoff, int32_t *dst) {
  int bits[32];
  float mags[32];
  int nb = 0;
  for (int b = 0; b < m; ++b) {
    /* insert (|proj|, b) into an ascending list, drop the largest if > cutoff */
    float a = fabsf(proj[b]);
    int p = nb;
    while (p > 0 && (a < mags[p - 1] || (a == mags[p - 1] && b < bits[p - 1]))) {
      if (p < 32) {
        mags[p] = mags[p - 1];
        bits[p] = bits[p - 1];
      }
      --p;
    }
    mags[p] = a;
    bits[p] = b;
    ++nb;
    if (nb > cutoff) nb = cutoff;
  }
  int32_t base = code_of(proj, m);
  int limit = 1 << cutoff;
  for (int ii = 0; ii < limit; ++ii) {
    int32_t c = base;
    /* top_idx order = descending pair order = bits[nb-1], ..., bits[0] */
    for (int q = 0; q < nb; ++q) {
      int bit = bits[nb - 1 - q];
      c &= ~(int32_t)(1u << bit);
      if (ii & (1 << q)) c |= (int32_t)(1u << bit);
    }
    dst[ii] = c;
  }
}

static int cmp_i32(const void *a, const void *b) {
  int32_t x = *(const int32_t *)a, y = *(const int32_t *)b;
  return (x > y) - (x < y);
}

typedef struct {
  int32_t code, idx;
} ci_t;

static int cmp_ci(const void *a, const void *b) {
  const ci_t *p = (const ci_t *)a, *q = (const ci_t *)b;
  if (p->code != q->code) return (p->code > q->code) - (p->code < q->code);
  return (p->idx > q->idx) - (p->idx < q->idx);
}
static void spv_sort_ci(ci_t *base, size_t n) {
  qsort(base, n, sizeof(ci_t), cmp_ci);
}

/* Whole matcher: ctor + find_neighbours, CascadingHashNn.h:187-245, called by
 * nn_cascading_hash, Spectavi.cpp:321-336.
 *   x, y      float32 rows holding integers in [-128, 127]
 *   out_idx   [yrows][2] u64, out_dist [yrows][2] float32 (int dist cast, :244)
 *   ncand_out optional [yrows] number of distinct candidates per query
 * Candidate visiting order: the reference iterates a std::unordered_set<int>
 * (:24,:42-49), i.e. libstdc++ hash order, unspecified.  Here candidates are
 * visited in ascending idx, which makes the result the lexicographic top-2 --
 * the rule the CUDA path implements. */
int oracle_cascade(const float *x, const float *y, int xrows, int yrows,
                   int dim, int m, int n, int g, const float *dict,
                   int nthreads, uint64_t *out_idx, float *out_dist,
                   int32_t *ncand_out) {
  if (dim % 16 != 0 || m < 1 || m > 31 || g < 0 || g > m || g > 16) return -1;
  if (nthreads < 1) nthreads = 1;
  /* find_neighbours :236-239: cast<uint8_t>() then += 128 in uint8 arithmetic */
  uint8_t *bx = (uint8_t *)malloc((size_t)xrows * dim + 16);
  uint8_t *by = (uint8_t *)malloc((size_t)yrows * dim + 16);
  for (size_t i = 0; i < (size_t)xrows * dim; ++i)
    bx[i] = (uint8_t)((int)x[i] + 128);
  for (size_t i = 0; i < (size_t)yrows * dim; ++i)
    by[i] = (uint8_t)((int)y[i] + 128);

  /* initialize_hash_tables :187-196 -- buckets as CSR over sorted (code, idx) */
  int32_t *codes = (int32_t *)malloc(sizeof(int32_t) * (size_t)xrows * n + 4);
  oracle_hashcodes(x, xrows, dim, dict, m, n, codes);
  /* per table: order[] = x indices sorted by (code, idx) */
  ci_t **tab = (ci_t **)malloc(sizeof(ci_t *) * (size_t)n);
  for (int t = 0; t < n; ++t) {
    tab[t] = (ci_t *)malloc(sizeof(ci_t) * (size_t)(xrows + 1));
    for (int i = 0; i < xrows; ++i) {
      tab[t][i].code = codes[(size_t)i * n + t];
      tab[t][i].idx = i;
    }
  }
  for (int t = 0; t < n; ++t) spv_sort_ci(tab[t], (size_t)xrows);

  int err = 0;
#pragma omp parallel num_threads(nthreads)
  {
    float proj[32];
    int32_t *probes = (int32_t *)malloc(sizeof(int32_t) * ((size_t)1 << g));
    size_t cap = 1024, cnt = 0;
    int32_t *cands = (int32_t *)malloc(sizeof(int32_t) * cap);
#pragma omp for schedule(dynamic, 64)
    for (int r = 0; r < yrows; ++r) {
      cnt = 0;
      /* filter_potential_neighbours :208-227 */
      for (int t = 0; t < n; ++t) {
        project_row(y + (size_t)r * dim, dict + (size_t)t * dim * m, dim, m,
                    proj); /* m_intermediate_form_y[t].row(r), :137-142 */
        probe_codes(proj, m, g, probes);
        for (int q = 0; q < (1 << g); ++q) {
          /* lower bound of probes[q] in tab[t] */
          int lo = 0, hi = xrows;
          while (lo < hi) {
            int mid = lo + (hi - lo) / 2;
            if (tab[t][mid].code < probes[q])
              lo = mid + 1;
            else
              hi = mid;
          }
          for (int p = lo; p < xrows && tab[t][p].code == probes[q]; ++p) {
            if (cnt == cap) {
              cap *= 2;
              cands = (int32_t *)realloc(cands, sizeof(int32_t) * cap);
            }
            cands[cnt++] = tab[t][p].idx;
          }
        }
      }
      /* unordered_set dedup (:223) -> sort + unique, ascending idx */
      qsort(cands, cnt, sizeof(int32_t), cmp_i32);
      size_t u = 0;
      for (size_t p = 0; p < cnt; ++p)
        if (u == 0 || cands[p] != cands[u - 1]) cands[u++] = cands[p];
      if (ncand_out) ncand_out[r] = (int32_t)u;
      top2_t tt;
      top2_reset(&tt);
      for (size_t p = 0; p < u; ++p)
        top2_offer(&tt, bx + (size_t)cands[p] * dim, by + (size_t)r * dim,
                   dim / 16, (uint64_t)cands[p], 1);
      out_idx[2 * (size_t)r + 0] = tt.i0;
      out_idx[2 * (size_t)r + 1] = tt.i1;
      out_dist[2 * (size_t)r + 0] = (float)tt.d0; /* :244 int -> float */
      out_dist[2 * (size_t)r + 1] = (float)tt.d1;
    }
    free(probes);
    free(cands);
  }
  for (int t = 0; t < n; ++t) free(tab[t]);
  free(tab);
  free(codes);
  free(bx);
  free(by);
  return err;
}

/* Probe codes of one y row / one table, exposed so tests can compare the
 * candidate rule itself (CascadingHashNn.h:150-185) with the CUDA path. */
int oracle_probe_codes(const float *yrow, int dim, const float *dict_t, int m,
                       int g, int32_t *dst) {
  if (m < 1 || m > 31 || g < 0 || g > m) return -1;
  float proj[32];
  project_row(yrow, dict_t, dim, m, proj);
  probe_codes(proj, m, g, dst);
  return 0;
}

/* ---- example/ex01_essential_estimation.py:102-103 ---------------------------
 * ratio = d[:,1] / float64(d[:,0]);  pass = ratio >= min_ratio.
 * x/0 = inf passes, 0/0 = nan fails -- IEEE semantics of the C division. */
void oracle_ratio_mask_i32(const int32_t *dist, int yrows, double min_ratio,
                           uint8_t *mask) {
  for (int r = 0; r < yrows; ++r) {
    double ratio = (double)dist[2 * r + 1] / (double)dist[2 * r + 0];
    mask[r] = ratio >= min_ratio;
  }
}

/* float32 distances (cascade path): numpy computes float32 / float64 in
 * float64, same expression. */
void oracle_ratio_mask_f32(const float *dist, int yrows, double min_ratio,
                           uint8_t *mask) {
  for (int r = 0; r < yrows; ++r) {
    double ratio = (double)dist[2 * r + 1] / (double)dist[2 * r + 0];
    mask[r] = ratio >= min_ratio;
  }
}
