// txt_csr.cpp -- the reference's txt-CSR on-disk format, read and written at memory speed.
//
// Format (cupy_cusparse/gen_and_save_alg1_txt.py:8-14): <prefix>_indptr.txt, _indices.txt,
// _data.txt; one number per line; ints "%d", values "%.9g".
// Reader follows read_i32_list / read_f32_list (cupy_cusparse/spgemm_from_txt_alg1.cu:19-34):
// whitespace-separated tokens, integers parsed as long long and narrowed to int32, values
// parsed as double and narrowed to float; parsing stops at the first token that is not a
// number (like `while (f >> x)`).  Writer follows write_i32_list / write_f32_list
// (alg1.cu:35-46): default float format with precision 9, i.e. printf("%.9g").
// The reference spends >99 % of its wall clock for small cases in iostream token parsing;
// here the file is read in one piece and parsed with std::from_chars / written with
// std::to_chars into one buffer.
#include <cerrno>
#include <charconv>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/b200sp.h"

namespace {

void set_err(char *errbuf, size_t n, const std::string &msg) {
  if (errbuf && n) snprintf(errbuf, n, "%s", msg.c_str());
}

bool slurp(const std::string &path, std::vector<char> &out, std::string &err) {
  FILE *f = fopen(path.c_str(), "rb");
  if (!f) {
    err = "open " + path + ": " + strerror(errno);
    return false;
  }
  fseek(f, 0, SEEK_END);
  long sz = ftell(f);
  fseek(f, 0, SEEK_SET);
  if (sz < 0) sz = 0;
  out.resize((size_t)sz + 1);
  size_t got = sz ? fread(out.data(), 1, (size_t)sz, f) : 0;
  fclose(f);
  out[got] = '\0';
  out.resize(got + 1);
  return true;
}

inline const char *skip_ws(const char *p, const char *e) {
  while (p < e && (*p == ' ' || *p == '\n' || *p == '\r' || *p == '\t' || *p == '\v' || *p == '\f')) ++p;
  return p;
}

template <typename T, typename Parse>
bool parse_list(const std::string &path, T **out, int64_t *count, Parse parse, std::string &err) {
  std::vector<char> buf;
  if (!slurp(path, buf, err)) return false;
  const char *p = buf.data(), *e = buf.data() + buf.size() - 1;
  size_t cap = buf.size() / 4 + 16, n = 0;
  T *v = (T *)malloc(cap * sizeof(T));
  if (!v) {
    err = "out of host memory reading " + path;
    return false;
  }
  while (true) {
    p = skip_ws(p, e);
    if (p >= e) break;
    if (*p == '+') ++p;  // iostream accepts a leading '+', from_chars does not
    T x;
    const char *q = parse(p, e, x);
    if (!q) break;  // first non-number ends the list, like `while (f >> x)`
    if (n == cap) {
      cap *= 2;
      T *nv = (T *)realloc(v, cap * sizeof(T));
      if (!nv) {
        free(v);
        err = "out of host memory reading " + path;
        return false;
      }
      v = nv;
    }
    v[n++] = x;
    p = q;
  }
  *out = v;
  *count = (int64_t)n;
  return true;
}

const char *parse_i32(const char *p, const char *e, int32_t &x) {
  long long ll = 0;
  auto r = std::from_chars(p, e, ll);
  if (r.ec != std::errc()) return nullptr;
  x = (int32_t)ll;  // narrowed exactly like the reference's (int32_t)x
  return r.ptr;
}

const char *parse_f32(const char *p, const char *e, float &x) {
  double d = 0;
  auto r = std::from_chars(p, e, d);
  if (r.ec == std::errc::result_out_of_range) {  // strtod semantics: +-inf / denormal-flush
    char *end = nullptr;
    d = strtod(p, &end);
    x = (float)d;
    return end;
  }
  if (r.ec != std::errc()) return nullptr;
  x = (float)d;
  return r.ptr;
}

template <typename T, typename Fmt>
bool write_list(const std::string &path, const T *v, int64_t n, Fmt fmt, std::string &err) {
  FILE *f = fopen(path.c_str(), "wb");
  if (!f) {
    err = "open " + path + ": " + strerror(errno);
    return false;
  }
  std::vector<char> buf(1 << 20);
  size_t used = 0;
  for (int64_t i = 0; i < n; ++i) {
    if (used + 64 > buf.size()) {
      if (fwrite(buf.data(), 1, used, f) != used) {
        err = "write " + path + " failed";
        fclose(f);
        return false;
      }
      used = 0;
    }
    char *p = fmt(buf.data() + used, buf.data() + buf.size(), v[i]);
    *p++ = '\n';
    used = (size_t)(p - buf.data());
  }
  bool ok = fwrite(buf.data(), 1, used, f) == used;
  ok = (fclose(f) == 0) && ok;
  if (!ok) err = "write " + path + " failed";
  return ok;
}

}  // namespace

extern "C" int b200sp_read_csr_txt(const char *prefix, int64_t *rows, int64_t *nnz, int32_t **hp,
                                   int32_t **hj, float **hx, char *errbuf, size_t errbuf_len) {
  if (!prefix || !rows || !nnz || !hp || !hj || !hx) return B200SP_ERR_INVALID;
  *hp = *hj = nullptr;
  *hx = nullptr;
  std::string err, pre(prefix);
  int64_t np = 0, nj = 0, nx = 0;
  bool ok = parse_list<int32_t>(pre + "_indptr.txt", hp, &np, parse_i32, err) &&
            parse_list<int32_t>(pre + "_indices.txt", hj, &nj, parse_i32, err) &&
            parse_list<float>(pre + "_data.txt", hx, &nx, parse_f32, err);
  int rc = B200SP_OK;
  if (!ok) rc = B200SP_ERR_IO;
  else if (np == 0) {  // alg1.cu:57-60
    err = "empty indptr: " + pre;
    rc = B200SP_ERR_INVALID;
  } else if (nj != nx) {  // alg1.cu:61-64
    err = "indices/data length mismatch for prefix " + pre;
    rc = B200SP_ERR_INVALID;
  }
  if (rc != B200SP_OK) {
    set_err(errbuf, errbuf_len, err);
    free(*hp);
    free(*hj);
    free(*hx);
    *hp = *hj = nullptr;
    *hx = nullptr;
    return rc;
  }
  *rows = np - 1;
  *nnz = nj;
  return B200SP_OK;
}

extern "C" int b200sp_write_csr_txt(const char *prefix, int64_t rows, int64_t nnz, const int32_t *hp,
                                    const int32_t *hj, const float *hx, char *errbuf, size_t errbuf_len) {
  if (!prefix || rows < 0 || nnz < 0 || !hp || (nnz && (!hj || !hx))) return B200SP_ERR_INVALID;
  std::string err, pre(prefix);
  auto fi = [](char *b, char *e, int32_t v) { return std::to_chars(b, e, v).ptr; };
  auto ff = [](char *b, char *e, float v) {
    // operator<<(float) widens to double and prints %.9g
    return std::to_chars(b, e, (double)v, std::chars_format::general, 9).ptr;
  };
  bool ok = write_list(pre + "_indptr.txt", hp, rows + 1, fi, err) &&
            write_list(pre + "_indices.txt", hj, nnz, fi, err) && write_list(pre + "_data.txt", hx, nnz, ff, err);
  if (!ok) {
    set_err(errbuf, errbuf_len, err);
    return B200SP_ERR_IO;
  }
  return B200SP_OK;
}
