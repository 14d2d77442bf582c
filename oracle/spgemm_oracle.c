/*
 * spgemm_oracle.c -- CPU oracle for the CSR SpGEMM / SpMV hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this.  The product (libb200sp.so) never links, imports or calls it and has no CPU
 * fallback.
 *
 * What it restates.  The reference (wjustin784/spMM) has no arithmetic of its own on this
 * path: the GPU side is closed-source cuSPARSE (call sites cupy_cusparse/spgemm_from_txt_alg1.cu:171-188,
 * modify_src/cupy-src/cupyx/cusparse.py:2059-2137) and the CPU side is scipy.sparse
 * (`A_fmt @ B_fmt`, `A_fmt @ C` at SpGEMM_vs_SpMV/profiler.py:408,411).  SciPy is a third-party
 * dependency that is NOT vendored under /root/reference and is unpinned there
 * (requirements.txt:2); this container has scipy 1.18.1.  The algorithm restated here is
 * SciPy's published sparsetools one (scipy/sparse/sparsetools/csr.h: csr_matmat_maxnnz,
 * csr_matmat, csr_matvec, csr_matvecs): row-wise Gustavson with a dense per-column
 * accumulator; every output entry is the sequential sum, in A-row storage order, of
 * separately rounded products (no fused multiply-add; build with -ffp-contract=off).
 * SciPy emits each row's columns in reverse first-touch order; the reference generators call
 * sort_indices() before saving (cupy_cusparse/gen_and_save_alg1_txt.py:23-24) and cuSPARSE
 * returns sorted rows, so the oracle sorts each row's columns.  Entries whose sum is exactly 0.0
 * (cancellation, products of explicit zeros) stay stored: that is the structural pattern
 * cuSPARSE returns (the path the product replaces; checked on the GPU against the reference's own
 * drivers, tests/test_gpu_vs_reference.py).  SciPy's csr_matmat drops them in its second pass
 * (`if (sums[head] != 0)`); oracle.py: spgemm(..., drop_zeros=True) applies that filter, and the
 * tests pin both forms.
 *
 * Pinning.  tests/test_oracle.py checks this file against (a) the literal 3x4 / 4x3 fixtures of
 * modify_src/cupy-src/tests/cupyx_tests/scipy_tests/sparse_tests/test_csr.py:25-135 with
 * hand-computed products, (b) golden vectors under tests/golden/ produced by scipy.sparse (the
 * reference's CPU path) with tests/golden/make_golden.py, bit-for-bit, and (c) live scipy on
 * the 24-case size x density grid of cupy_cusparse/run_all_alg1.sh:5-7.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef int32_t idx_t;

int oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* number of intermediate products: what cusparseSpGEMM_getNumProducts returns
 * (spgemm_from_txt_alg3.cu:190-191); FLOPs = 2 * this. */
int64_t oracle_num_products(int64_t m, const idx_t *Ap, const idx_t *Aj, const idx_t *Bp) {
  int64_t total = 0;
  for (int64_t i = 0; i < m; ++i)
    for (idx_t t = Ap[i]; t < Ap[i + 1]; ++t) total += (int64_t)(Bp[Aj[t] + 1] - Bp[Aj[t]]);
  return total;
}

/* per-row product counts (used for the row-panel partition of the multi-GPU path). */
void oracle_row_products(int64_t m, const idx_t *Ap, const idx_t *Aj, const idx_t *Bp,
                         int64_t *prods) {
  for (int64_t i = 0; i < m; ++i) {
    int64_t s = 0;
    for (idx_t t = Ap[i]; t < Ap[i + 1]; ++t) s += (int64_t)(Bp[Aj[t] + 1] - Bp[Aj[t]]);
    prods[i] = s;
  }
}

/* Symbolic pass (the job of csr_matmat_maxnnz, made exact per row): fills Cp[0..m] and returns
 * nnz(C), or -1 when it would overflow int32 indptr. */
int64_t oracle_spgemm_symbolic(int64_t m, int64_t n, const idx_t *Ap, const idx_t *Aj,
                               const idx_t *Bp, const idx_t *Bj, idx_t *Cp) {
  idx_t *stamp = (idx_t *)malloc((size_t)(n > 0 ? n : 1) * sizeof(idx_t));
  if (!stamp) return -2;
  for (int64_t c = 0; c < n; ++c) stamp[c] = -1;
  int64_t nnz = 0;
  Cp[0] = 0;
  for (int64_t i = 0; i < m; ++i) {
    for (idx_t t = Ap[i]; t < Ap[i + 1]; ++t) {
      idx_t kk = Aj[t];
      for (idx_t u = Bp[kk]; u < Bp[kk + 1]; ++u) {
        idx_t c = Bj[u];
        if (stamp[c] != (idx_t)i) {
          stamp[c] = (idx_t)i;
          ++nnz;
        }
      }
    }
    if (nnz > INT32_MAX) {
      free(stamp);
      return -1;
    }
    Cp[i + 1] = (idx_t)nnz;
  }
  free(stamp);
  return nnz;
}

static int cmp_idx(const void *a, const void *b) {
  idx_t x = *(const idx_t *)a, y = *(const idx_t *)b;
  return (x > y) - (x < y);
}

/* One row of the numeric pass.  acc/stamp are n-long scratch owned by the caller (stamp == row
 * id marks "touched in this row"); cols receives the touched columns, then is sorted. */
#define DEFINE_ROW_KERNEL(NAME, T)                                                               \
  static idx_t NAME(int64_t i, const idx_t *Ap, const idx_t *Aj, const T *Ax, const idx_t *Bp,   \
                    const idx_t *Bj, const T *Bx, T *acc, idx_t *stamp, idx_t *cols, T *vals) {  \
    idx_t cnt = 0;                                                                               \
    for (idx_t t = Ap[i]; t < Ap[i + 1]; ++t) {                                                  \
      idx_t kk = Aj[t];                                                                          \
      T a = Ax[t];                                                                               \
      for (idx_t u = Bp[kk]; u < Bp[kk + 1]; ++u) {                                              \
        idx_t c = Bj[u];                                                                         \
        T prod = a * Bx[u]; /* rounded product ... */                                            \
        if (stamp[c] != (idx_t)i) {                                                              \
          stamp[c] = (idx_t)i;                                                                   \
          acc[c] = (T)0;                                                                         \
          cols[cnt++] = c;                                                                       \
        }                                                                                        \
        acc[c] = acc[c] + prod; /* ... then a separately rounded add (0 + p == p exactly) */     \
      }                                                                                          \
    }                                                                                            \
    qsort(cols, (size_t)cnt, sizeof(idx_t), cmp_idx);                                            \
    for (idx_t q = 0; q < cnt; ++q) vals[q] = acc[cols[q]];                                      \
    return cnt;                                                                                  \
  }

DEFINE_ROW_KERNEL(row_f32, float)
DEFINE_ROW_KERNEL(row_f64, double)

/* Numeric pass: Cp must come from oracle_spgemm_symbolic.  alpha scales the finished sums
 * (C = alpha * (A @ B), the form the upstream test checks: test_cusparse.py:395-409).
 * Rows are independent, so the OpenMP build splits them over threads (nthreads <= 0: all);
 * the per-entry arithmetic and its order do not depend on the thread count. */
#define DEFINE_NUMERIC(NAME, ROWK, T)                                                            \
  int NAME(int64_t m, int64_t n, const idx_t *Ap, const idx_t *Aj, const T *Ax, const idx_t *Bp, \
           const idx_t *Bj, const T *Bx, T alpha, const idx_t *Cp, idx_t *Cj, T *Cx,             \
           int nthreads) {                                                                       \
    int bad = 0;                                                                                 \
    (void)nthreads;                                                                              \
    _Pragma("omp parallel num_threads(nthreads > 0 ? nthreads : oracle_num_threads())")          \
    {                                                                                            \
      size_t nn = (size_t)(n > 0 ? n : 1);                                                       \
      T *acc = (T *)malloc(nn * sizeof(T));                                                      \
      idx_t *stamp = (idx_t *)malloc(nn * sizeof(idx_t));                                        \
      if (!acc || !stamp) {                                                                      \
        _Pragma("omp atomic write") bad = 1;                                                     \
      } else {                                                                                   \
        for (size_t c = 0; c < nn; ++c) stamp[c] = -1;                                           \
        _Pragma("omp for schedule(dynamic, 64)")                                                 \
        for (int64_t i = 0; i < m; ++i) {                                                        \
          idx_t cnt = ROWK(i, Ap, Aj, Ax, Bp, Bj, Bx, acc, stamp, Cj + Cp[i], Cx + Cp[i]);       \
          if (cnt != Cp[i + 1] - Cp[i]) {                                                        \
            _Pragma("omp atomic write") bad = 2;                                                 \
          }                                                                                      \
          if (alpha != (T)1)                                                                     \
            for (idx_t q = Cp[i]; q < Cp[i + 1]; ++q) Cx[q] = alpha * Cx[q];                     \
        }                                                                                        \
      }                                                                                          \
      free(acc);                                                                                 \
      free(stamp);                                                                               \
    }                                                                                            \
    return bad;                                                                                  \
  }

DEFINE_NUMERIC(oracle_spgemm_numeric_f32, row_f32, float)
DEFINE_NUMERIC(oracle_spgemm_numeric_f64, row_f64, double)

/* y = alpha * (A @ x) + beta * y ; csr_matvec order: one sequential sum per row.
 * beta == 0 overwrites y without reading it. */
#define DEFINE_SPMV(NAME, T)                                                                     \
  void NAME(int64_t m, const idx_t *Ap, const idx_t *Aj, const T *Ax, const T *x, T *y, T alpha, \
            T beta) {                                                                            \
    for (int64_t i = 0; i < m; ++i) {                                                            \
      T s = (T)0;                                                                                \
      for (idx_t t = Ap[i]; t < Ap[i + 1]; ++t) {                                                \
        T prod = Ax[t] * x[Aj[t]];                                                               \
        s = s + prod;                                                                            \
      }                                                                                          \
      T r = alpha * s;                                                                           \
      y[i] = (beta == (T)0) ? r : r + beta * y[i];                                               \
    }                                                                                            \
  }

DEFINE_SPMV(oracle_spmv_f32, float)
DEFINE_SPMV(oracle_spmv_f64, double)

/* 1 when every row has strictly increasing in-range columns and indptr is monotone
 * (the has_canonical_format predicate, cupyx/scipy/sparse/_compressed.py:177-192). */
int oracle_csr_is_canonical(int64_t m, int64_t ncols, const idx_t *Ap, const idx_t *Aj) {
  if (Ap[0] != 0) return 0;
  for (int64_t i = 0; i < m; ++i) {
    if (Ap[i + 1] < Ap[i]) return 0;
    for (idx_t t = Ap[i]; t < Ap[i + 1]; ++t) {
      if (Aj[t] < 0 || Aj[t] >= ncols) return 0;
      if (t > Ap[i] && Aj[t - 1] >= Aj[t]) return 0;
    }
  }
  return 1;
}
