// convert_api.inl -- host orchestration of the canonicalisation / dense / compare entry points
// (included by b200sp_api.cu; kernels in convert.cuh).  Scratch comes from the handle's second
// arena (ws2): 2 key buffers (8 B) + 2 payload buffers (4 B) per entry + digit counters
// (nnz/4 B) = ~24.3 * nnz bytes, released by b200sp_release_workspace.

namespace {

inline int bits_for(int64_t count) {  // bits needed for values 0 .. count-1
  int b = 0;
  while (b < 32 && ((int64_t)1 << b) < count) ++b;
  return b;
}

struct SortScratch {
  u64 *keys[2];
  int *perm[2];
  int *counts;     // 256 * ntiles digit counters, scanned in place
  int *sums;       // tile sums of the generic scan
  long long *total;  // device scalar
  int *flag;         // device scalar
  int *rowtmp;       // extra_rows ints (row cursors of the bucket path), or null
  int ntiles;
};

int carve_sort_scratch(b200sp_context *h, int64_t nnz, SortScratch &s, int64_t extra_rows = 0) {
  const size_t ntiles = (size_t)((nnz + RS_TILE - 1) / RS_TILE);
  const size_t counters = std::max<size_t>(std::max<size_t>(256 * ntiles, (size_t)((nnz + GS_TILE - 1) / GS_TILE) + 1),
                                           (size_t)extra_rows);
  const size_t nsums = (counters + GS_TILE - 1) / GS_TILE + 1;
  const size_t o_k1 = align_up((size_t)nnz * 8, 256);
  const size_t o_p0 = 2 * o_k1;
  const size_t o_p1 = o_p0 + align_up((size_t)nnz * 4, 256);
  const size_t o_c = o_p1 + align_up((size_t)nnz * 4, 256);
  const size_t o_s = o_c + align_up(counters * 4, 256);
  const size_t o_t = o_s + align_up(nsums * 4, 256);
  const size_t o_r = o_t + 256;
  TRY(ensure(h, h->ws2, o_r + align_up((size_t)extra_rows * 4, 256)));
  char *base = (char *)h->ws2.ptr;
  s.keys[0] = (u64 *)base;
  s.keys[1] = (u64 *)(base + o_k1);
  s.perm[0] = (int *)(base + o_p0);
  s.perm[1] = (int *)(base + o_p1);
  s.counts = (int *)(base + o_c);
  s.sums = (int *)(base + o_s);
  s.total = (long long *)(base + o_t);
  s.flag = (int *)(base + o_t + 8);
  s.rowtmp = extra_rows ? (int *)(base + o_r) : nullptr;
  s.ntiles = (int)ntiles;
  return B200SP_OK;
}

// exclusive scan of data[0..n) in place; *total_dev (optional, device) receives the sum
int dev_exscan(b200sp_context *h, int *data, long long n, int *sums, long long *total_dev) {
  if (n <= 0) return B200SP_OK;
  const int nt = (int)((n + GS_TILE - 1) / GS_TILE);
  k_gscan_sums<<<nt, GS_THREADS, 0, h->stream>>>(data, n, sums);
  TRY(check_launch(h, "k_gscan_sums"));
  k_gscan_offsets<<<1, GS_THREADS, 0, h->stream>>>(sums, nt, total_dev);
  TRY(check_launch(h, "k_gscan_offsets"));
  k_gscan_apply<<<nt, GS_THREADS, 0, h->stream>>>(data, data, n, sums);
  return check_launch(h, "k_gscan_apply");
}

// Stable sort of s.keys[0] (payload = input position) over the digit positions covering bits
// [0, lo_bits) and [32, 32 + hi_bits).  *where = index of the buffer pair holding the result.
int radix_sort(b200sp_context *h, SortScratch &s, long long n, int lo_bits, int hi_bits, int *where) {
  int shifts[8], np = 0;
  for (int b = 0; b < lo_bits; b += 8) shifts[np++] = b;
  for (int b = 0; b < hi_bits; b += 8) shifts[np++] = 32 + b;
  if (np == 0) shifts[np++] = 32 + hi_bits;  // one pass on an all-zero digit: it materialises the identity payload
  int cur = 0;
  for (int p = 0; p < np; ++p) {
    k_rs_count<<<s.ntiles, RS_THREADS, 0, h->stream>>>(s.keys[cur], n, shifts[p], s.ntiles, s.counts);
    TRY(check_launch(h, "k_rs_count"));
    TRY(dev_exscan(h, s.counts, 256ll * s.ntiles, s.sums, nullptr));
    k_rs_scatter<<<s.ntiles, RS_THREADS, 0, h->stream>>>(s.keys[cur], p == 0 ? nullptr : s.perm[cur], s.keys[cur ^ 1],
                                                         s.perm[cur ^ 1], n, shifts[p], s.ntiles, s.counts);
    TRY(check_launch(h, "k_rs_scatter"));
    cur ^= 1;
  }
  *where = cur;
  return B200SP_OK;
}

int check_convert_args(b200sp_context *h, int dtype, int64_t m, int64_t n, int64_t nnz) {
  if (m < 0 || n < 0 || nnz < 0 || m >= INT32_MAX || n >= INT32_MAX || nnz > INT32_MAX)
    return fail(h, B200SP_ERR_INVALID, "bad shape m=%lld n=%lld nnz=%lld", (long long)m, (long long)n, (long long)nnz);
  if (dtype != B200SP_R_32F && dtype != B200SP_R_64F) return fail(h, B200SP_ERR_INVALID, "unsupported dtype %d", dtype);
  return B200SP_OK;
}

inline int flat_grid(b200sp_context *h, int64_t items, int per_cta) {
  return (int)std::max<int64_t>(1, std::min<int64_t>((items + per_cta - 1) / per_cta, h->num_sms * 16));
}

// sorted keys + payload -> summed, duplicate-free CSR; one host sync for nnz_out and the input flag
template <typename T>
int finish_unique(b200sp_context *h, SortScratch &s, int cur, int64_t m, int64_t nnz, const T *val, int *Cp, int *Cj,
                  T *Cx, int64_t *nnz_out, const char *what) {
  cudaStream_t st = h->stream;
  const int nt = (int)((nnz + GS_TILE - 1) / GS_TILE);
  int *tile_heads = s.counts;  // the digit counters are free once the sort is done
  k_uniq_count<<<nt, GS_THREADS, 0, st>>>(s.keys[cur], nnz, tile_heads);
  TRY(check_launch(h, "k_uniq_count"));
  TRY(dev_exscan(h, tile_heads, nt, s.sums, s.total));
  k_uniq_emit<T><<<nt, GS_THREADS, 0, st>>>(s.keys[cur], s.perm[cur], val, nnz, tile_heads, s.keys[cur ^ 1], Cj, Cx);
  TRY(check_launch(h, "k_uniq_emit"));
  k_ptr_from_sorted<<<flat_grid(h, m + 1, 256), 256, 0, st>>>((int)m, s.keys[cur ^ 1], nnz, s.total, Cp);
  TRY(check_launch(h, "k_ptr_from_sorted"));
  long long total = 0;
  int bad = 0;
  CUDA_TRY(h, cudaMemcpyAsync(&total, s.total, sizeof(total), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(h, cudaMemcpyAsync(&bad, s.flag, sizeof(bad), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(h, cudaStreamSynchronize(st));
  if (bad) return fail(h, B200SP_ERR_INVALID, "%s: index out of range or malformed pointer array", what);
  *nnz_out = total;
  return B200SP_OK;
}

template <typename T>
int run_coo_to_csr(b200sp_context *h, int64_t m, int64_t n, int64_t nnz, const int *row, const int *col, const T *val,
                   int *Cp, int *Cj, T *Cx, int64_t *nnz_out) {
  SortScratch s;
  TRY(carve_sort_scratch(h, nnz, s, m + 1));
  cudaStream_t st = h->stream;
  CUDA_TRY(h, cudaMemsetAsync(s.flag, 0, 2 * sizeof(int), st));  // [bad coordinate | longest row]
  // ---- row buckets: histogram -> scan -> drop into the row's segment -> in-row sort (convert.cuh) ----
  const char *force = getenv("B200SP_COO_PATH");  // radix | bucket (tests / measurements)
  if (!(force && strcmp(force, "radix") == 0)) {
    CUDA_TRY(h, cudaMemsetAsync(Cp, 0, (size_t)(m + 1) * sizeof(int), st));
    k_coo_row_count<<<flat_grid(h, nnz, 1024), 256, 0, st>>>(nnz, (int)m, (int)n, row, col, Cp, s.flag);
    TRY(check_launch(h, "k_coo_row_count"));
    TRY(dev_exscan(h, Cp, m + 1, s.sums, nullptr));
    k_max_row_len<<<flat_grid(h, m, 256), 256, 0, st>>>((int)m, Cp, s.flag + 1);
    TRY(check_launch(h, "k_max_row_len"));
    int head[2] = {0, 0};
    CUDA_TRY(h, cudaMemcpyAsync(head, s.flag, sizeof(head), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(h, cudaStreamSynchronize(st));
    if (head[0]) return fail(h, B200SP_ERR_INVALID, "coo_to_csr: index out of range or malformed pointer array");
    if (head[1] <= COO_ROW_CAP) {
      int *cursor = s.counts;  // >= m + 1 ints (carve_sort_scratch sizes the counters for it), free in this path
      CUDA_TRY(h, cudaMemcpyAsync(cursor, Cp, (size_t)m * sizeof(int), cudaMemcpyDeviceToDevice, st));
      k_coo_bucket<<<flat_grid(h, nnz, 1024), 256, 0, st>>>(nnz, (int)m, (int)n, row, col, cursor, s.keys[1]);
      TRY(check_launch(h, "k_coo_bucket"));
      // Cp holds the row segments of the un-summed entries until the distinct-column counts are scanned into it
      int *rowptr = s.rowtmp;
      CUDA_TRY(h, cudaMemcpyAsync(rowptr, Cp, (size_t)(m + 1) * sizeof(int), cudaMemcpyDeviceToDevice, st));
      k_coo_sort_rows<<<flat_grid(h, (m + 1) / 2, 8), 256, 0, st>>>((int)m, rowptr, s.keys[1], Cp);
      TRY(check_launch(h, "k_coo_sort_rows"));
      CUDA_TRY(h, cudaMemsetAsync(Cp + m, 0, sizeof(int), st));
      TRY(dev_exscan(h, Cp, m + 1, s.sums, s.total));
      k_coo_emit_rows<T><<<flat_grid(h, m, 8), 256, 0, st>>>((int)m, rowptr, s.keys[1], val, Cp, Cj, Cx);
      TRY(check_launch(h, "k_coo_emit_rows"));
      long long total = 0;
      CUDA_TRY(h, cudaMemcpyAsync(&total, s.total, sizeof(total), cudaMemcpyDeviceToHost, st));
      CUDA_TRY(h, cudaStreamSynchronize(st));
      *nnz_out = total;
      return B200SP_OK;
    }
  }
  k_keys_from_coo<<<flat_grid(h, nnz, 1024), 256, 0, h->stream>>>(nnz, (int)m, (int)n, row, col, s.keys[0], s.flag);
  TRY(check_launch(h, "k_keys_from_coo"));
  int cur = 0;
  TRY(radix_sort(h, s, nnz, bits_for(n), bits_for(m), &cur));
  return finish_unique<T>(h, s, cur, m, nnz, val, Cp, Cj, Cx, nnz_out, "coo_to_csr");
}

template <typename T>
int run_csr_sum_duplicates(b200sp_context *h, int64_t m, int64_t n, int64_t nnz, const int *Ap, const int *Aj,
                           const T *Ax, int *Cp, int *Cj, T *Cx, int64_t *nnz_out) {
  SortScratch s;
  TRY(carve_sort_scratch(h, nnz, s, m + 1));
  cudaStream_t st = h->stream;
  CUDA_TRY(h, cudaMemsetAsync(s.flag, 0, 2 * sizeof(int), st));
  const char *force = getenv("B200SP_COO_PATH");
  if (!(force && strcmp(force, "radix") == 0)) {
    // short rows: sort every row's (column, position) keys where they lie, then the per-row ordered sums
    k_max_row_len<<<flat_grid(h, m, 256), 256, 0, st>>>((int)m, Ap, s.flag + 1);
    TRY(check_launch(h, "k_max_row_len"));
    k_keyc_from_csr<<<flat_grid(h, m, 8), 256, 0, st>>>((int)m, (int)n, nnz, Ap, Aj, s.keys[1], s.flag);
    TRY(check_launch(h, "k_keyc_from_csr"));
    int head[2] = {0, 0};
    CUDA_TRY(h, cudaMemcpyAsync(head, s.flag, sizeof(head), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(h, cudaStreamSynchronize(st));
    if (head[0]) return fail(h, B200SP_ERR_INVALID, "csr_sum_duplicates: index out of range or malformed pointer array");
    if (head[1] <= COO_ROW_CAP) {
      k_coo_sort_rows<<<flat_grid(h, (m + 1) / 2, 8), 256, 0, st>>>((int)m, Ap, s.keys[1], Cp);
      TRY(check_launch(h, "k_coo_sort_rows"));
      CUDA_TRY(h, cudaMemsetAsync(Cp + m, 0, sizeof(int), st));
      TRY(dev_exscan(h, Cp, m + 1, s.sums, s.total));
      k_coo_emit_rows<T><<<flat_grid(h, m, 8), 256, 0, st>>>((int)m, Ap, s.keys[1], Ax, Cp, Cj, Cx);
      TRY(check_launch(h, "k_coo_emit_rows"));
      long long total = 0;
      CUDA_TRY(h, cudaMemcpyAsync(&total, s.total, sizeof(total), cudaMemcpyDeviceToHost, st));
      CUDA_TRY(h, cudaStreamSynchronize(st));
      *nnz_out = total;
      return B200SP_OK;
    }
    CUDA_TRY(h, cudaMemsetAsync(s.flag, 0, 2 * sizeof(int), st));
  }
  k_keys_from_compressed<true><<<flat_grid(h, m, 8), 256, 0, h->stream>>>((int)m, (int)n, nnz, Ap, Aj, s.keys[0], s.flag);
  TRY(check_launch(h, "k_keys_from_compressed"));
  int cur = 0;
  // the input is already grouped by row: sorting on the column digits and then on the row
  // digits is what restores the grouping after the column passes
  TRY(radix_sort(h, s, nnz, bits_for(n), bits_for(m), &cur));
  return finish_unique<T>(h, s, cur, m, nnz, Ax, Cp, Cj, Cx, nnz_out, "csr_sum_duplicates");
}

// compressed (nmajor segments, minor indices < nminor) -> compressed along the other axis
template <typename T>
int run_transpose(b200sp_context *h, int64_t nmajor, int64_t nminor, int64_t nnz, const int *ptr, const int *idx,
                  const T *val, int *out_ptr, int *out_idx, T *out_val) {
  SortScratch s;
  TRY(carve_sort_scratch(h, nnz, s, nminor + 1));
  cudaStream_t st = h->stream;
  CUDA_TRY(h, cudaMemsetAsync(s.flag, 0, 2 * sizeof(int), st));  // [bad input | longest output segment]
  const char *force = getenv("B200SP_COO_PATH");  // radix | bucket: the same switch as COO -> CSR
  if (!(force && strcmp(force, "radix") == 0)) {
    // ---- buckets: entries per minor index -> scan -> (major, position) keys dropped into the segment -> in-segment sort
    CUDA_TRY(h, cudaMemsetAsync(out_ptr, 0, (size_t)(nminor + 1) * sizeof(int), st));
    k_count_minor<<<flat_grid(h, nnz, 1024), 256, 0, st>>>(nnz, (int)nminor, idx, out_ptr, s.flag);
    TRY(check_launch(h, "k_count_minor"));
    TRY(dev_exscan(h, out_ptr, nminor + 1, s.sums, nullptr));
    k_max_row_len<<<flat_grid(h, nminor, 256), 256, 0, st>>>((int)nminor, out_ptr, s.flag + 1);
    TRY(check_launch(h, "k_max_row_len"));
    int head[2] = {0, 0};
    CUDA_TRY(h, cudaMemcpyAsync(head, s.flag, sizeof(head), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(h, cudaStreamSynchronize(st));
    if (head[0]) return fail(h, B200SP_ERR_INVALID, "transpose: index out of range or malformed pointer array");
    if (head[1] <= COO_ROW_CAP) {
      CUDA_TRY(h, cudaMemcpyAsync(s.rowtmp, out_ptr, (size_t)nminor * sizeof(int), cudaMemcpyDeviceToDevice, st));
      k_bucket_compressed<<<flat_grid(h, nmajor, 8), 256, 0, st>>>((int)nmajor, (int)nminor, nnz, ptr, idx, s.rowtmp, s.keys[1], s.flag);
      TRY(check_launch(h, "k_bucket_compressed"));
      k_coo_sort_rows<<<flat_grid(h, (nminor + 1) / 2, 8), 256, 0, st>>>((int)nminor, out_ptr, s.keys[1], s.counts);
      TRY(check_launch(h, "k_coo_sort_rows"));
      k_emit_bucketed<T><<<flat_grid(h, nnz, 1024), 256, 0, st>>>(nnz, s.keys[1], val, out_idx, out_val);
      TRY(check_launch(h, "k_emit_bucketed"));
      int bad2 = 0;
      CUDA_TRY(h, cudaMemcpyAsync(&bad2, s.flag, sizeof(bad2), cudaMemcpyDeviceToHost, st));
      CUDA_TRY(h, cudaStreamSynchronize(st));
      if (bad2) return fail(h, B200SP_ERR_INVALID, "transpose: index out of range or malformed pointer array");
      return B200SP_OK;
    }
    CUDA_TRY(h, cudaMemsetAsync(s.flag, 0, 2 * sizeof(int), st));
  }
  k_keys_from_compressed<false><<<flat_grid(h, nmajor, 8), 256, 0, st>>>((int)nmajor, (int)nminor, nnz, ptr, idx, s.keys[0],
                                                                    s.flag);
  TRY(check_launch(h, "k_keys_from_compressed"));
  int cur = 0;
  // entries arrive in ascending major order, so a stable sort on the minor index alone leaves
  // every output segment sorted by the old major index
  TRY(radix_sort(h, s, nnz, 0, bits_for(nminor), &cur));
  k_emit_sorted<T><<<flat_grid(h, nnz, 1024), 256, 0, st>>>(nnz, s.keys[cur], s.perm[cur], val, out_idx, out_val);
  TRY(check_launch(h, "k_emit_sorted"));
  k_ptr_from_sorted<<<flat_grid(h, nminor + 1, 256), 256, 0, st>>>((int)nminor, s.keys[cur], nnz, nullptr, out_ptr);
  TRY(check_launch(h, "k_ptr_from_sorted"));
  int bad = 0;
  CUDA_TRY(h, cudaMemcpyAsync(&bad, s.flag, sizeof(bad), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(h, cudaStreamSynchronize(st));
  if (bad) return fail(h, B200SP_ERR_INVALID, "transpose: index out of range or malformed pointer array");
  return B200SP_OK;
}

// In-place stable in-row sort for any row length (the fallback of b200sp_csr_sort_indices): keys
// (row << 32 | column), LSD passes over the column digits then the row digits (the input is grouped
// by row already; the row passes restore the grouping), entries written to scratch and copied back.
// Equal columns keep their storage order, as with cusparseXcsrsort + gather (cupyx/cusparse.py:832-866).
template <typename T>
int run_sort_indices_stable(b200sp_context *h, int64_t m, int64_t nnz, const int *Ap, int *Aj, T *Ax) {
  SortScratch s;
  TRY(carve_sort_scratch(h, nnz, s));
  cudaStream_t st = h->stream;
  CUDA_TRY(h, cudaMemsetAsync(s.flag, 0, sizeof(int), st));
  k_keys_from_compressed<true><<<flat_grid(h, m, 8), 256, 0, st>>>((int)m, INT_MAX, nnz, Ap, Aj, s.keys[0], s.flag);
  TRY(check_launch(h, "k_keys_from_compressed"));
  int cur = 0;
  TRY(radix_sort(h, s, nnz, 31, bits_for(m), &cur));  // the column count is not an argument: all 31 index bits
  int *idx_out = s.perm[cur ^ 1];
  T *val_out = reinterpret_cast<T *>(s.keys[cur ^ 1]);
  k_emit_sorted<T><<<flat_grid(h, nnz, 1024), 256, 0, st>>>(nnz, s.keys[cur], s.perm[cur], Ax, idx_out, val_out);
  TRY(check_launch(h, "k_emit_sorted"));
  int bad = 0;
  CUDA_TRY(h, cudaMemcpyAsync(&bad, s.flag, sizeof(bad), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(h, cudaStreamSynchronize(st));
  if (bad) return fail(h, B200SP_ERR_INVALID, "csr_sort_indices: negative index or malformed pointer array");
  CUDA_TRY(h, cudaMemcpyAsync(Aj, idx_out, (size_t)nnz * sizeof(int), cudaMemcpyDeviceToDevice, st));
  CUDA_TRY(h, cudaMemcpyAsync(Ax, val_out, (size_t)nnz * sizeof(T), cudaMemcpyDeviceToDevice, st));
  return B200SP_OK;
}

}  // namespace

extern "C" int b200sp_coo_to_csr(b200sp_handle_t h, int dtype, int64_t m, int64_t n, int64_t nnz,
                                 const int32_t *dRow, const int32_t *dCol, const void *dVal, int32_t *dCp,
                                 int32_t *dCj, void *dCx, int64_t *nnz_out) {
  if (!h) return B200SP_ERR_INVALID;
  TRY(check_convert_args(h, dtype, m, n, nnz));
  if (!dCp || !nnz_out || (nnz && (!dRow || !dCol || !dVal || !dCj || !dCx)))
    return fail(h, B200SP_ERR_INVALID, "null pointer argument");
  if (nnz && (m == 0 || n == 0)) return fail(h, B200SP_ERR_INVALID, "entries in an empty shape");
  DeviceGuard guard(h->device);
  *nnz_out = 0;
  if (nnz == 0) {
    CUDA_TRY(h, cudaMemsetAsync(dCp, 0, (size_t)(m + 1) * sizeof(int), h->stream));
    return B200SP_OK;
  }
  return dtype == B200SP_R_32F
             ? run_coo_to_csr<float>(h, m, n, nnz, dRow, dCol, (const float *)dVal, dCp, dCj, (float *)dCx, nnz_out)
             : run_coo_to_csr<double>(h, m, n, nnz, dRow, dCol, (const double *)dVal, dCp, dCj, (double *)dCx, nnz_out);
}

extern "C" int b200sp_csr_sum_duplicates(b200sp_handle_t h, int dtype, int64_t m, int64_t n, int64_t nnz,
                                         const int32_t *dAp, const int32_t *dAj, const void *dAx, int32_t *dCp,
                                         int32_t *dCj, void *dCx, int64_t *nnz_out) {
  if (!h) return B200SP_ERR_INVALID;
  TRY(check_convert_args(h, dtype, m, n, nnz));
  if (!dAp || !dCp || !nnz_out || (nnz && (!dAj || !dAx || !dCj || !dCx)))
    return fail(h, B200SP_ERR_INVALID, "null pointer argument");
  DeviceGuard guard(h->device);
  *nnz_out = 0;
  if (nnz == 0) {
    CUDA_TRY(h, cudaMemsetAsync(dCp, 0, (size_t)(m + 1) * sizeof(int), h->stream));
    return B200SP_OK;
  }
  return dtype == B200SP_R_32F ? run_csr_sum_duplicates<float>(h, m, n, nnz, dAp, dAj, (const float *)dAx, dCp, dCj,
                                                               (float *)dCx, nnz_out)
                               : run_csr_sum_duplicates<double>(h, m, n, nnz, dAp, dAj, (const double *)dAx, dCp, dCj,
                                                                (double *)dCx, nnz_out);
}

extern "C" int b200sp_csc_to_csr(b200sp_handle_t h, int dtype, int64_t m, int64_t n, int64_t nnz,
                                 const int32_t *dColPtr, const int32_t *dRowIdx, const void *dVal, int32_t *dCp,
                                 int32_t *dCj, void *dCx) {
  if (!h) return B200SP_ERR_INVALID;
  TRY(check_convert_args(h, dtype, m, n, nnz));
  if (!dColPtr || !dCp || (nnz && (!dRowIdx || !dVal || !dCj || !dCx)))
    return fail(h, B200SP_ERR_INVALID, "null pointer argument");
  DeviceGuard guard(h->device);
  if (nnz == 0) {
    CUDA_TRY(h, cudaMemsetAsync(dCp, 0, (size_t)(m + 1) * sizeof(int), h->stream));
    return B200SP_OK;
  }
  return dtype == B200SP_R_32F
             ? run_transpose<float>(h, n, m, nnz, dColPtr, dRowIdx, (const float *)dVal, dCp, dCj, (float *)dCx)
             : run_transpose<double>(h, n, m, nnz, dColPtr, dRowIdx, (const double *)dVal, dCp, dCj, (double *)dCx);
}

extern "C" int b200sp_csr_to_csc(b200sp_handle_t h, int dtype, int64_t m, int64_t n, int64_t nnz, const int32_t *dAp,
                                 const int32_t *dAj, const void *dAx, int32_t *dColPtr, int32_t *dRowIdx,
                                 void *dVal) {
  // the same re-ordering with the roles of rows and columns exchanged
  return b200sp_csc_to_csr(h, dtype, n, m, nnz, dAp, dAj, dAx, dColPtr, dRowIdx, dVal);
}

extern "C" int b200sp_csr_to_dense(b200sp_handle_t h, int dtype, int64_t m, int64_t n, const int32_t *dAp,
                                   const int32_t *dAj, const void *dAx, void *dOut, int order) {
  if (!h) return B200SP_ERR_INVALID;
  TRY(check_convert_args(h, dtype, m, n, 0));
  if (order != B200SP_ORDER_ROW && order != B200SP_ORDER_COL) return fail(h, B200SP_ERR_INVALID, "bad order %d", order);
  if (m == 0 || n == 0) return B200SP_OK;
  if (!dAp || !dOut) return fail(h, B200SP_ERR_INVALID, "null pointer argument");
  DeviceGuard guard(h->device);
  const size_t vs = dtype == B200SP_R_32F ? 4 : 8;
  CUDA_TRY(h, cudaMemsetAsync(dOut, 0, (size_t)m * (size_t)n * vs, h->stream));
  const long long ldr = order == B200SP_ORDER_ROW ? n : 1, ldc = order == B200SP_ORDER_ROW ? 1 : m;
  const int grid = flat_grid(h, m, 8);
  if (dtype == B200SP_R_32F)
    k_csr_to_dense<float><<<grid, 256, 0, h->stream>>>((int)m, (int)n, dAp, dAj, (const float *)dAx, (float *)dOut, ldr, ldc);
  else
    k_csr_to_dense<double><<<grid, 256, 0, h->stream>>>((int)m, (int)n, dAp, dAj, (const double *)dAx, (double *)dOut, ldr, ldc);
  return check_launch(h, "k_csr_to_dense");
}

extern "C" int b200sp_csr_max_abs_diff(b200sp_handle_t h, int dtype, int64_t m, const int32_t *dAp, const int32_t *dAj,
                                       const void *dAx, const int32_t *dBp, const int32_t *dBj, const void *dBx,
                                       double *result) {
  if (!h) return B200SP_ERR_INVALID;
  TRY(check_convert_args(h, dtype, m, 0, 0));
  if (!result) return fail(h, B200SP_ERR_INVALID, "null pointer argument");
  *result = 0.0;
  if (m == 0) return B200SP_OK;
  if (!dAp || !dBp) return fail(h, B200SP_ERR_INVALID, "null pointer argument");
  DeviceGuard guard(h->device);
  TRY(ensure(h, h->ws2, 256));
  u64 *res = (u64 *)h->ws2.ptr;
  CUDA_TRY(h, cudaMemsetAsync(res, 0, 2 * sizeof(u64), h->stream));
  const int grid = flat_grid(h, m, 8);
  if (dtype == B200SP_R_32F)
    k_csr_max_abs_diff<float><<<grid, 256, 0, h->stream>>>((int)m, dAp, dAj, (const float *)dAx, dBp, dBj, (const float *)dBx, res);
  else
    k_csr_max_abs_diff<double><<<grid, 256, 0, h->stream>>>((int)m, dAp, dAj, (const double *)dAx, dBp, dBj, (const double *)dBx, res);
  TRY(check_launch(h, "k_csr_max_abs_diff"));
  u64 out[2] = {0, 0};
  CUDA_TRY(h, cudaMemcpyAsync(out, res, sizeof(out), cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(h, cudaStreamSynchronize(h->stream));
  double v;
  memcpy(&v, &out[0], sizeof(v));
  *result = out[1] ? NAN : v;
  return B200SP_OK;
}
