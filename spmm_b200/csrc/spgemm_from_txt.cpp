// spgemm_from_txt.cpp -- CLI driver: A_prefix B_prefix C_prefix [chunk_fraction]
//
// Drop-in for the reference's three drivers cupy_cusparse/spgemm_from_txt_alg{1,2,3}.cu; plugs
// into run_all_alg{1,2,3}.sh through the CUEXE environment variable (run_all_alg1.sh:14).
// Same argv, same env (CHUNK_FRACTION), same files in and out, same exit codes (0 ok, 2 usage,
// 1 anything else, with the message on stderr), same one-line banner on stdout.
// The algorithm (1/2/3) comes from the executable's name (..._alg1/_alg2/_alg3, installed as
// symlinks), else from the B200SP_ALG environment variable, else 1.
// All arithmetic happens in libb200sp.so on the GPU; there is no CPU fallback.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../../include/b200sp.h"

static int pick_alg(const char *argv0) {
  std::string name(argv0);
  for (int a = 1; a <= 3; ++a) {
    const std::string suf = "_alg" + std::to_string(a);
    if (name.size() >= suf.size() && name.compare(name.size() - suf.size(), suf.size(), suf) == 0) return a;
  }
  if (const char *s = getenv("B200SP_ALG")) {
    const int a = atoi(s);
    if (a >= 1 && a <= 3) return a;
  }
  return 1;
}

// spgemm_from_txt_alg3.cu:101-115
static float get_chunk_fraction(int argc, char **argv) {
  float cf = 0.2f;
  if (argc >= 5) cf = strtof(argv[4], nullptr);
  else if (const char *s = getenv("CHUNK_FRACTION")) cf = strtof(s, nullptr);
  if (cf <= 0.0f || cf > 1.0f) {
    fprintf(stderr, "chunk_fraction must be in (0,1], got %f\n", cf);
    exit(1);
  }
  return cf;
}

struct HostCSR {
  int64_t rows = 0, cols = -1, nnz = 0;
  int32_t *indptr = nullptr, *indices = nullptr;
  float *data = nullptr;
};

static void read_or_die(const std::string &prefix, HostCSR &M) {
  char err[512] = {0};
  const int rc = b200sp_read_csr_txt(prefix.c_str(), &M.rows, &M.nnz, &M.indptr, &M.indices, &M.data, err, sizeof err);
  if (rc != B200SP_OK) {
    fprintf(stderr, "%s\n", err[0] ? err : b200sp_status_string(rc));
    exit(1);
  }
}

// spgemm_from_txt_alg1.cu:78-99
static void validate_or_die(const HostCSR &M, const char *name) {
  if (M.cols < 0) {
    fprintf(stderr, "[%s] cols not set before validation\n", name);
    exit(1);
  }
  if (M.nnz > 0) {
    const int32_t mx = *std::max_element(M.indices, M.indices + M.nnz);
    if (mx >= M.cols) {
      fprintf(stderr, "[%s] index out of bounds: max index %d >= ncols %lld\n", name, mx, (long long)M.cols);
      exit(1);
    }
    if (*std::min_element(M.indices, M.indices + M.nnz) < 0) {
      fprintf(stderr, "[%s] negative column index detected\n", name);
      exit(1);
    }
  }
  // beyond the reference: indptr must describe exactly nnz entries, or the kernels would read
  // out of bounds (cuSPARSE would, too)
  if (M.indptr[0] != 0 || M.indptr[M.rows] != M.nnz) {
    fprintf(stderr, "[%s] indptr[0]=%d, indptr[rows]=%d inconsistent with nnz=%lld\n", name, M.indptr[0],
            M.indptr[M.rows], (long long)M.nnz);
    exit(1);
  }
  for (int64_t i = 0; i < M.rows; ++i)
    if (M.indptr[i + 1] < M.indptr[i]) {
      fprintf(stderr, "[%s] indptr not monotone at row %lld\n", name, (long long)i);
      exit(1);
    }
}

#define CHECK_B200SP(h, x)                                                                  \
  do {                                                                                      \
    int s_ = (x);                                                                           \
    if (s_ != B200SP_OK) {                                                                  \
      fprintf(stderr, "b200sp error %s:%d: %d (%s) %s\n", __FILE__, __LINE__, s_,          \
              b200sp_status_string(s_), (h) ? b200sp_last_error(h) : "");                   \
      exit(1);                                                                              \
    }                                                                                       \
  } while (0)

int main(int argc, char **argv) {
  const int alg = pick_alg(argv[0]);
  if (argc < 4) {
    fprintf(stderr, alg == 1 ? "Usage: %s A_prefix B_prefix C_prefix\n" : "Usage: %s A_prefix B_prefix C_prefix [chunk_fraction]\n",
            argv[0]);
    return 2;
  }
  const std::string Apre = argv[1], Bpre = argv[2], Cpre = argv[3];
  const float chunk_fraction = alg == 1 ? 0.2f : get_chunk_fraction(argc, argv);

  HostCSR A, B;
  read_or_die(Apre, A);
  read_or_die(Bpre, B);
  if (A.rows <= 0 || B.rows <= 0) {
    fprintf(stderr, "invalid rows: A.rows=%lld B.rows=%lld\n", (long long)A.rows, (long long)B.rows);
    return 1;
  }
  // The format stores no column count: the reference infers A.cols = B.rows, B.cols = A.rows
  // (spgemm_from_txt_alg1.cu:121-122).  Kept as is.
  A.cols = B.rows;
  B.cols = A.rows;
  validate_or_die(A, "A");
  validate_or_die(B, "B");

  b200sp_handle_t h = nullptr;
  {
    const int rc = b200sp_create(&h, -1);
    if (rc != B200SP_OK) {
      fprintf(stderr, "b200sp error %s:%d: %d (%s): no usable CUDA device\n", __FILE__, __LINE__, rc, b200sp_status_string(rc));
      return 1;
    }
  }
  int32_t *Cp = nullptr, *Cj = nullptr;
  void *Cx = nullptr;
  int64_t nnzC = 0, prods = 0;
  CHECK_B200SP(h, b200sp_spgemm_host(h, B200SP_R_32F, A.rows, A.cols, B.cols, A.indptr, A.indices, A.data, B.indptr,
                                     B.indices, B.data, alg, chunk_fraction, nullptr, &Cp, &Cj, &Cx, &nnzC, &prods));
  char err[512] = {0};
  const int wrc = b200sp_write_csr_txt(Cpre.c_str(), A.rows, nnzC, Cp, Cj, (const float *)Cx, err, sizeof err);
  if (wrc != B200SP_OK) {
    fprintf(stderr, "%s\n", err[0] ? err : b200sp_status_string(wrc));
    return 1;
  }
  if (alg == 1)
    printf("[C++] ALG1 wrote %s_* (rows=%lld, cols=%lld, nnz=%lld)\n", Cpre.c_str(), (long long)A.rows, (long long)B.cols,
           (long long)nnzC);
  else
    printf("[C++] ALG%d wrote %s_* (rows=%lld, cols=%lld, nnz=%lld, chunk_fraction=%g)\n", alg, Cpre.c_str(),
           (long long)A.rows, (long long)B.cols, (long long)nnzC, chunk_fraction);
  b200sp_destroy(h);
  b200sp_free_host(A.indptr); b200sp_free_host(A.indices); b200sp_free_host(A.data);
  b200sp_free_host(B.indptr); b200sp_free_host(B.indices); b200sp_free_host(B.data);
  return 0;
}
