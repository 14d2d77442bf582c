/*
 * convert_oracle.c -- CPU oracle for the format conversions around the hot path (SURVEY.md
 * section 8f, N2/N4).  TEST INFRASTRUCTURE ONLY (same rule as spgemm_oracle.c: tests/, smoke()
 * and bench.py's CPU legs may load it, the product never does).
 *
 * What it restates.  The reference converts through CuPy/cuSPARSE (closed: coosort/coo2csr/
 * csc2csr, modify_src/cupy-src/cupyx/cusparse.py:908-970, 1092-1114) and sums duplicates with
 * floating-point atomics (cupyx/scipy/sparse/_coo.py:437-447), i.e. in no fixed order; its CPU
 * path is SciPy's sparsetools (not vendored, unpinned, 1.18.1 here):
 *   coo_tocsr            scipy/sparse/sparsetools/coo.h  (count rows, cumsum, scatter in input order)
 *   csr_sort_indices     csr.h  (std::sort of (col, val) pairs per row -- NOT stable)
 *   csr_sum_duplicates   csr.h  (adjacent equal columns added left to right)
 *   csr_tocsc            csr.h  (count, cumsum, scatter in row order => sorted, stable)
 * Restated here with the one degree of freedom SciPy leaves open (the order of equal columns
 * after the unstable sort) fixed to INPUT order, which is the product's contract.  With at most
 * two duplicates per position (fp addition commutes) or rows shorter than 16 entries
 * (std::sort is an insertion sort there) this is bit-identical to SciPy; tests pin exactly that.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

typedef int32_t idx_t;

typedef struct { idx_t col; int64_t pos; } ent_t;

static int cmp_ent(const void *a, const void *b) {
  const ent_t *x = (const ent_t *)a, *y = (const ent_t *)b;
  if (x->col != y->col) return x->col < y->col ? -1 : 1;
  return x->pos < y->pos ? -1 : (x->pos > y->pos);
}

/* returns nnz of the canonical CSR, -1 on a coordinate out of range, -2 on OOM */
#define DEF_COO_TO_CSR(NAME, T)                                                                  \
  int64_t NAME(int64_t m, int64_t n, int64_t nnz, const idx_t *row, const idx_t *col,            \
               const T *val, idx_t *Cp, idx_t *Cj, T *Cx) {                                      \
    int64_t *start = (int64_t *)calloc((size_t)m + 1, sizeof(int64_t));                          \
    ent_t *ent = (ent_t *)malloc(((size_t)nnz + 1) * sizeof(ent_t));                             \
    if (!start || !ent) { free(start); free(ent); return -2; }                                   \
    for (int64_t t = 0; t < nnz; ++t) {                                                          \
      if (row[t] < 0 || row[t] >= m || col[t] < 0 || col[t] >= n) { free(start); free(ent); return -1; } \
      start[row[t] + 1]++;                                                                       \
    }                                                                                            \
    for (int64_t i = 0; i < m; ++i) start[i + 1] += start[i];                                    \
    {                                                                                            \
      int64_t *cur = (int64_t *)malloc(((size_t)m + 1) * sizeof(int64_t));                       \
      if (!cur) { free(start); free(ent); return -2; }                                           \
      memcpy(cur, start, ((size_t)m + 1) * sizeof(int64_t));                                     \
      for (int64_t t = 0; t < nnz; ++t) { /* coo_tocsr: input order inside a row */              \
        ent_t e = {col[t], t};                                                                   \
        ent[cur[row[t]]++] = e;                                                                  \
      }                                                                                          \
      free(cur);                                                                                 \
    }                                                                                            \
    int64_t out = 0;                                                                             \
    Cp[0] = 0;                                                                                   \
    for (int64_t i = 0; i < m; ++i) {                                                            \
      const int64_t b = start[i], e = start[i + 1];                                              \
      qsort(ent + b, (size_t)(e - b), sizeof(ent_t), cmp_ent);                                   \
      int64_t t = b;                                                                             \
      while (t < e) { /* csr_sum_duplicates */                                                   \
        const idx_t j = ent[t].col;                                                              \
        T x = val[ent[t].pos];                                                                   \
        ++t;                                                                                     \
        while (t < e && ent[t].col == j) { x = x + val[ent[t].pos]; ++t; }                       \
        Cj[out] = j;                                                                             \
        Cx[out] = x;                                                                             \
        ++out;                                                                                   \
      }                                                                                          \
      Cp[i + 1] = (idx_t)out;                                                                    \
    }                                                                                            \
    free(start);                                                                                 \
    free(ent);                                                                                   \
    return out;                                                                                  \
  }

DEF_COO_TO_CSR(oracle_coo_to_csr_f32, float)
DEF_COO_TO_CSR(oracle_coo_to_csr_f64, double)

/* csr_tocsc restated (csr.h): input = nmajor segments (ptr, idx < nminor); output compressed
 * along the other axis with every segment in ascending old-major order.  0 ok, -1 bad index. */
#define DEF_TRANSPOSE(NAME, T)                                                                   \
  int NAME(int64_t nmajor, int64_t nminor, const idx_t *ptr, const idx_t *idx, const T *val,     \
           idx_t *optr, idx_t *oidx, T *oval) {                                                  \
    const int64_t nnz = ptr[nmajor];                                                             \
    memset(optr, 0, ((size_t)nminor + 1) * sizeof(idx_t));                                       \
    for (int64_t t = 0; t < nnz; ++t) {                                                          \
      if (idx[t] < 0 || idx[t] >= nminor) return -1;                                             \
      optr[idx[t] + 1]++;                                                                        \
    }                                                                                            \
    for (int64_t j = 0; j < nminor; ++j) optr[j + 1] += optr[j];                                 \
    idx_t *cur = (idx_t *)malloc(((size_t)nminor + 1) * sizeof(idx_t));                          \
    if (!cur) return -2;                                                                         \
    memcpy(cur, optr, ((size_t)nminor + 1) * sizeof(idx_t));                                     \
    for (int64_t s = 0; s < nmajor; ++s)                                                         \
      for (idx_t t = ptr[s]; t < ptr[s + 1]; ++t) {                                              \
        const idx_t d = cur[idx[t]]++;                                                           \
        oidx[d] = (idx_t)s;                                                                      \
        oval[d] = val[t];                                                                        \
      }                                                                                          \
    free(cur);                                                                                   \
    return 0;                                                                                    \
  }

DEF_TRANSPOSE(oracle_transpose_f32, float)
DEF_TRANSPOSE(oracle_transpose_f64, double)

/* max |A - B| over the union pattern of two canonical CSR matrices (two-pointer merge per row);
 * the dense statement is numerical_error/error.py:27-28.  NaN if any compared value is NaN. */
#define DEF_MAXDIFF(NAME, T)                                                                     \
  double NAME(int64_t m, const idx_t *Ap, const idx_t *Aj, const T *Ax, const idx_t *Bp,         \
              const idx_t *Bj, const T *Bx) {                                                    \
    double mx = 0.0;                                                                             \
    int nan = 0;                                                                                 \
    for (int64_t i = 0; i < m; ++i) {                                                            \
      idx_t a = Ap[i], b = Bp[i];                                                                \
      const idx_t ae = Ap[i + 1], be = Bp[i + 1];                                                \
      while (a < ae || b < be) {                                                                 \
        T d;                                                                                     \
        if (b >= be || (a < ae && Aj[a] < Bj[b])) d = Ax[a++];                                   \
        else if (a >= ae || Bj[b] < Aj[a]) d = Bx[b++];                                          \
        else d = (T)(Ax[a++] - Bx[b++]);                                                         \
        const double v = fabs((double)d);                                                        \
        if (v != v) nan = 1; else if (v > mx) mx = v;                                            \
      }                                                                                          \
    }                                                                                            \
    return nan ? NAN : mx;                                                                       \
  }

DEF_MAXDIFF(oracle_csr_max_abs_diff_f32, float)
DEF_MAXDIFF(oracle_csr_max_abs_diff_f64, double)
