// spmv.cuh -- CSR SpMV / one-or-few-column SpMM kernels.
//
// Takes the place of cusparseSpMV / cusparseSpMM (reference call sites:
// modify_src/cupy-src/cupyx/cusparse.py:1430-1435, 1505-1511).  The matrix streams once from HBM
// (nnz * (4 + s) bytes); x is gathered through L1/L2.  For uniformly random columns the gathers,
// not HBM, set the ceiling: tools/gather_probe.cu measures ~1.27 SM-cycles per random 4-byte
// gather on B200 (L1 tag stage), i.e. ~50 us for the 1.1e7 nonzeros of BASELINE config C2.
#pragma once
#include "common.cuh"

namespace b200sp {

template <typename T>
__device__ __forceinline__ T fma_acc(T a, T b, T c) { return a * b + c; }

// "Vector" kernel: a group of G lanes per row (G picked on the host from nnz/m), the Aj/Ax
// reads of a warp cover 32/G adjacent rows, i.e. one contiguous span of the arrays.  R rows are
// in flight per group so that R independent indptr -> (index, value) -> x chains overlap.
// Column `col` of dense X / Y is selected with blockIdx.y (ldx, ldy = leading dimensions).
template <int G, typename T>
__global__ void __launch_bounds__(256)
k_spmv_vector(int m, const int *__restrict__ Ap, const int *__restrict__ Aj,
              const T *__restrict__ Ax, const T *__restrict__ X, long long ldx, T *__restrict__ Y,
              long long ldy, T alpha, T beta) {
  constexpr int R = 4;
  const int sub = threadIdx.x & (G - 1);
  const int groups_per_block = 256 / G;
  const T *x = X + (long long)blockIdx.y * ldx;
  T *y = Y + (long long)blockIdx.y * ldy;
  const long long stride = (long long)gridDim.x * groups_per_block;
  for (long long row0 = blockIdx.x * groups_per_block + threadIdx.x / G; row0 < m; row0 += stride * R) {
    int b[R], e[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const long long row = row0 + r * stride;
      b[r] = e[r] = 0;
      if (row < m) {
        b[r] = __ldg(Ap + row);
        e[r] = __ldg(Ap + row + 1);
      }
    }
    T sum[R], a[R];
    int j[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {  // first stride of every row: all loads issued before any is used
      sum[r] = T(0);
      const int t = b[r] + sub;
      j[r] = t < e[r] ? __ldcs(Aj + t) : -1;
      a[r] = t < e[r] ? __ldcs(Ax + t) : T(0);
    }
#pragma unroll
    for (int r = 0; r < R; ++r)
      if (j[r] >= 0) sum[r] = fma_acc(a[r], __ldg(x + j[r]), sum[r]);
#pragma unroll
    for (int r = 0; r < R; ++r)
      for (int t = b[r] + sub + G; t < e[r]; t += G) sum[r] = fma_acc(__ldcs(Ax + t), __ldg(x + __ldcs(Aj + t)), sum[r]);
#pragma unroll
    for (int r = 0; r < R; ++r) {
#pragma unroll
      for (int o = G / 2; o > 0; o >>= 1) sum[r] += __shfl_xor_sync(FULL, sum[r], o, G);
      const long long row = row0 + r * stride;
      if (row < m && sub == 0) y[row] = beta == T(0) ? alpha * sum[r] : alpha * sum[r] + beta * y[row];
    }
  }
}

// ---- 128-bit variant of the vector kernel -------------------------------------------------------
// Each lane reads FOUR consecutive (index, value) pairs with one 128-bit load each (fp64 values:
// two), from the 16-byte aligned position at or below the row start; the entries outside
// [row begin, row end) are masked.  A group of G lanes therefore covers 4 G entries per trip, so
// G is a quarter of what the scalar kernel needs (C2, 10.5 entries per row: 4 lanes per row, 8 rows
// per warp) and the load-instruction count per nonzero drops 4-fold.  Requires Aj / Ax 16-byte
// aligned (the host checks; allocator blocks are); the last three entries of the arrays are read
// with scalar loads so that nothing past the end is touched.
template <typename T> struct Ld4;
template <> struct Ld4<int> {
  static __device__ __forceinline__ void ld(const int *p, int (&v)[4]) {
    const int4 t = __ldcs(reinterpret_cast<const int4 *>(p));
    v[0] = t.x, v[1] = t.y, v[2] = t.z, v[3] = t.w;
  }
};
template <> struct Ld4<float> {
  static __device__ __forceinline__ void ld(const float *p, float (&v)[4]) {
    const float4 t = __ldcs(reinterpret_cast<const float4 *>(p));
    v[0] = t.x, v[1] = t.y, v[2] = t.z, v[3] = t.w;
  }
};
template <> struct Ld4<double> {
  static __device__ __forceinline__ void ld(const double *p, double (&v)[4]) {
    const double2 t = __ldcs(reinterpret_cast<const double2 *>(p)), u = __ldcs(reinterpret_cast<const double2 *>(p + 2));
    v[0] = t.x, v[1] = t.y, v[2] = u.x, v[3] = u.y;
  }
};

template <int G, typename T>
__global__ void __launch_bounds__(256)
k_spmv_vector4(int m, int nnz, const int *__restrict__ Ap, const int *__restrict__ Aj,
               const T *__restrict__ Ax, const T *__restrict__ X, long long ldx, T *__restrict__ Y,
               long long ldy, T alpha, T beta) {
  constexpr int R = 4;
  const int sub = threadIdx.x & (G - 1);
  const int groups_per_block = 256 / G;
  const T *x = X + (long long)blockIdx.y * ldx;
  T *y = Y + (long long)blockIdx.y * ldy;
  const long long stride = (long long)gridDim.x * groups_per_block;
  const int vec_end = nnz & ~3;  // entries [0, vec_end) may be read with whole 16-byte loads
  for (long long row0 = blockIdx.x * groups_per_block + threadIdx.x / G; row0 < m; row0 += stride * R) {
    int b[R], e[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const long long row = row0 + r * stride;
      b[r] = e[r] = 0;
      if (row < m) {
        b[r] = __ldg(Ap + row);
        e[r] = __ldg(Ap + row + 1);
      }
    }
    T sum[R];
    int j[R][4];
    T a[R][4];
    int t[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {  // first trip of every row: all loads issued before any is used
      sum[r] = T(0);
      t[r] = (b[r] & ~3) + 4 * sub;
      if (t[r] < e[r]) {
        if (t[r] < vec_end) {
          Ld4<int>::ld(Aj + t[r], j[r]);
          Ld4<T>::ld(Ax + t[r], a[r]);
        } else {
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const bool in = t[r] + c < nnz;
            j[r][c] = in ? __ldcs(Aj + t[r] + c) : 0;
            a[r][c] = in ? __ldcs(Ax + t[r] + c) : T(0);
          }
        }
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if (t[r] < e[r]) {
        T xv[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int q = t[r] + c;
          xv[c] = (q >= b[r] && q < e[r]) ? __ldg(x + j[r][c]) : T(0);
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int q = t[r] + c;
          if (q >= b[r] && q < e[r]) sum[r] = fma_acc(a[r][c], xv[c], sum[r]);
        }
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      for (int tt = t[r] + 4 * G; tt < e[r]; tt += 4 * G) {  // tt >= b: only the row end needs a mask
        int jj[4];
        T aa[4];
        if (tt < vec_end) {
          Ld4<int>::ld(Aj + tt, jj);
          Ld4<T>::ld(Ax + tt, aa);
        } else {
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const bool in = tt + c < nnz;
            jj[c] = in ? __ldcs(Aj + tt + c) : 0;
            aa[c] = in ? __ldcs(Ax + tt + c) : T(0);
          }
        }
        T xv[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) xv[c] = tt + c < e[r] ? __ldg(x + jj[c]) : T(0);
#pragma unroll
        for (int c = 0; c < 4; ++c)
          if (tt + c < e[r]) sum[r] = fma_acc(aa[c], xv[c], sum[r]);
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
#pragma unroll
      for (int o = G / 2; o > 0; o >>= 1) sum[r] += __shfl_xor_sync(FULL, sum[r], o, G);
      const long long row = row0 + r * stride;
      if (row < m && sub == 0) y[row] = beta == T(0) ? alpha * sum[r] : alpha * sum[r] + beta * y[row];
    }
  }
}

// Longest row of a CSR (decides between the row-group kernel and merge-path; cached per matrix).
__global__ void __launch_bounds__(256) k_max_row_len(int m, const int *__restrict__ Ap, int *__restrict__ out) {
  int mx = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x)
    mx = max(mx, __ldg(Ap + i + 1) - __ldg(Ap + i));
  mx = __reduce_max_sync(FULL, mx);
  if (lane_id() == 0 && mx > 0) atomicMax(out, mx);
}

// ---- merge-path kernel ------------------------------------------------------------------------
// The work list is the merge of the m row-end markers with the nnz nonzeros (m + nnz items);
// every CTA takes SPMV_TILE consecutive items of it, so a tile holds at most SPMV_TILE nonzeros
// AND at most SPMV_TILE rows whatever the row-length distribution: one heavy row cannot
// serialise a warp, a stretch of empty rows cannot either.  k_spmv_partition finds the row
// coordinate of every tile boundary (a 32-ary search per warp: 4 rounds for 2^20 rows); the main
// kernel then knows its row range [i0, i1] and nonzero range [j0, j1) up front.
// Rows completed in a tile are written by it; the piece of a row begun in an earlier tile
// ("head") and the piece of a row left open at the tile's end ("carry") become per-tile partials
// that k_spmv_fixup adds in tile order, so the result is run-to-run deterministic (no atomics).
constexpr int SPMV_THREADS = 256;
constexpr int SPMV_ITEMS = 8;                           // merge items per thread
constexpr int SPMV_TILE = SPMV_THREADS * SPMV_ITEMS;    // 2048 merge items per tile

struct SpmvTileInfo {
  int head_row;    // row begun earlier whose entries open this tile, or -1
  int head_open;   // 1 when that row is not completed in this tile either (the tile lies inside it)
  int carry_row;   // row that starts in this tile and is not completed in it, or -1
  int pad;
};

// Rows completed before merge diagonal d: the first i with Ap[i+1] + i + 1 > d (that key is
// strictly increasing in i).  One warp per diagonal, 32-ary search.
__global__ void __launch_bounds__(256)
k_spmv_partition(int m, long long total, int ntiles, const int *__restrict__ Ap, int *__restrict__ tile_row) {
  const int lane = lane_id();
  const int k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;  // boundary index 0..ntiles
  if (k > ntiles) return;
  const long long d = min((long long)k * SPMV_TILE, total);
  int lo = 0, hi = m;  // answer in [lo, hi]: rows < lo are completed, rows >= hi are not
  while (hi - lo > 32) {
    const int step = (hi - lo + 31) / 32;
    const long long i = (long long)lo + (long long)(lane + 1) * step - 1;
    const bool done = (i < hi) && ((long long)__ldg(Ap + i + 1) + i + 1 <= d);
    const int cnt = __popc(__ballot_sync(FULL, done));  // monotone: the first cnt probes are completed
    const long long nlo = (long long)lo + (long long)cnt * step;
    const long long nhi = nlo + step - 1;
    lo = (int)nlo;
    hi = (int)(nhi < hi ? nhi : hi);
  }
  const int i = lo + lane;
  const bool done = (i < hi) && ((long long)__ldg(Ap + i + 1) + i + 1 <= d);
  const int res = lo + __popc(__ballot_sync(FULL, done));
  if (lane == 0) tile_row[k] = res;
}

template <typename T>
__device__ __forceinline__ void spmv_store(T *__restrict__ y, int r, T sum, T alpha, T beta) {
  y[r] = beta == T(0) ? alpha * sum : alpha * sum + beta * y[r];
}

// VEC: the tile's indices / values are read with 128-bit loads from the first 16-byte aligned entry
// at or after j0 (two loads of four entries per thread cover the 2048-entry tile); the up to three
// entries before it are read by the first threads with scalar loads.  Needs Aj / Ax 16-byte aligned.
template <typename T, bool VEC>
__global__ void __launch_bounds__(SPMV_THREADS)
k_spmv_merge(int m, int nnz, const int *__restrict__ tile_row, const int *__restrict__ Ap,
             const int *__restrict__ Aj, const T *__restrict__ Ax, const T *__restrict__ x,
             T *__restrict__ y, T alpha, T beta, SpmvTileInfo *__restrict__ info,
             T *__restrict__ head_val, T *__restrict__ carry_val) {
  constexpr int NW = SPMV_THREADS / 32;
  __shared__ T s_prod[SPMV_TILE + 4];
  __shared__ int s_ap[SPMV_TILE + 2];
  __shared__ T s_part[NW];
  __shared__ int s_nlong;
  __shared__ int s_long[SPMV_TILE / 32 + 1];  // rows of the tile longer than 32 entries
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int k = blockIdx.x;
  const long long total = (long long)m + nnz;
  const long long d0 = (long long)k * SPMV_TILE, d1 = min(d0 + SPMV_TILE, total);
  const int i0 = __ldg(tile_row + k), i1 = __ldg(tile_row + k + 1);  // rows [i0, i1) are completed here
  const int j0 = (int)(d0 - i0), j1 = (int)(d1 - i1);               // nonzeros [j0, j1) are consumed here
  const int nz = j1 - j0, nrows = i1 - i0;
  // all requests of the tile go out together: indptr of rows i0 .. i1 (+1), indices and values
  int jj[SPMV_ITEMS];
  T aa[SPMV_ITEMS];
  int apw[SPMV_ITEMS];
  int jh = -1;  // VEC: one of the entries in front of the aligned part
  T ah = T(0);
  const int vbase = (j0 + 3) & ~3;
  if (VEC) {
    const int vec_end = nnz & ~3;
#pragma unroll
    for (int q = 0; q < SPMV_ITEMS / 4; ++q) {
      const int el = vbase + 4 * (q * SPMV_THREADS + tid);
      int j4[4] = {-1, -1, -1, -1};
      T a4[4] = {T(0), T(0), T(0), T(0)};
      if (el < j1) {
        if (el < vec_end) {
          Ld4<int>::ld(Aj + el, j4);
          Ld4<T>::ld(Ax + el, a4);
        } else {
#pragma unroll
          for (int c = 0; c < 4; ++c)
            if (el + c < j1) {
              j4[c] = __ldcs(Aj + el + c);
              a4[c] = __ldcs(Ax + el + c);
            }
        }
      }
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        jj[q * 4 + c] = el + c < j1 ? j4[c] : -1;
        aa[q * 4 + c] = a4[c];
      }
    }
    if (j0 + tid < min(vbase, j1)) {
      jh = __ldcs(Aj + j0 + tid);
      ah = __ldcs(Ax + j0 + tid);
    }
  } else {
#pragma unroll
    for (int q = 0; q < SPMV_ITEMS; ++q) {
      const int e = q * SPMV_THREADS + tid;
      jj[q] = e < nz ? __ldcs(Aj + j0 + e) : -1;
      aa[q] = e < nz ? __ldcs(Ax + j0 + e) : T(0);
    }
  }
#pragma unroll
  for (int q = 0; q < SPMV_ITEMS; ++q) apw[q] = __ldg(Ap + min(i0 + q * SPMV_THREADS + tid, m));
  if (tid < 2) s_ap[SPMV_TILE + tid] = __ldg(Ap + min(i0 + SPMV_TILE + tid, m));
  if (tid == 0) s_nlong = 0;
  {
    T xv[SPMV_ITEMS];
#pragma unroll
    for (int q = 0; q < SPMV_ITEMS; ++q) xv[q] = jj[q] >= 0 ? __ldg(x + jj[q]) : T(0);
    const T xh = jh >= 0 ? __ldg(x + jh) : T(0);
#pragma unroll
    for (int q = 0; q < SPMV_ITEMS; ++q) {
      s_ap[q * SPMV_THREADS + tid] = apw[q];
      // tile-relative position of the entry held in slot q
      const int e = VEC ? (vbase - j0) + 4 * ((q >> 2) * SPMV_THREADS + tid) + (q & 3) : q * SPMV_THREADS + tid;
      if (jj[q] >= 0) s_prod[e] = aa[q] * xv[q];
    }
    if (jh >= 0) s_prod[tid] = ah * xh;
  }
  __syncthreads();
  // rows: s_ap[t] = Ap[i0 + t].  Row i0 may have begun in an earlier tile (head); row i1 is the
  // one left incomplete (its entries, if any, are [max(Ap[i1], j0), j1)).
  const int p0 = s_ap[0];                     // Ap[i0] <= j0
  const bool has_head = j0 > p0;              // row i0 started before this tile
  const bool head_closed = has_head && nrows > 0;
  const int open_b = max(s_ap[nrows], j0);    // first entry of the incomplete row inside the tile
  const bool has_open = (i1 < m) && (j1 > open_b);
  const bool open_is_head = has_open && nrows == 0 && has_head;  // the tile lies inside one row
  // Warp-uniform helper: one warp sums s_prod[b, e) in a fixed order.
  auto warp_sum_range = [&](int b, int e) {
    T sum = T(0);
    for (int q = b + lane; q < e; q += 32) sum += s_prod[q];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(FULL, sum, o);
    return sum;
  };
  // completed rows t = 0 .. nrows-1 cover s_prod[max(Ap[i0+t], j0) - j0, Ap[i0+t+1] - j0)
  if (nrows >= 64) {  // many rows: one thread per short row, long ones are queued for the warps
    for (int t = tid; t < nrows; t += SPMV_THREADS) {
      const int b = max(s_ap[t], j0) - j0, e = s_ap[t + 1] - j0;
      if (e - b > 32) {
        s_long[atomicAdd(&s_nlong, 1)] = t;
        continue;
      }
      T sum = T(0);
      for (int q = b; q < e; q += 4) {  // four loads in flight; the additions stay in order
        const T v0 = s_prod[q];
        const T v1 = q + 1 < e ? s_prod[q + 1] : T(0);
        const T v2 = q + 2 < e ? s_prod[q + 2] : T(0);
        const T v3 = q + 3 < e ? s_prod[q + 3] : T(0);
        sum += v0; sum += v1; sum += v2; sum += v3;
      }
      if (t == 0 && has_head) head_val[k] = sum;
      else spmv_store(y, i0 + t, sum, alpha, beta);
    }
    __syncthreads();
    const int nlong = s_nlong;
    for (int q = w; q < nlong; q += NW) {
      const int t = s_long[q];
      const T sum = warp_sum_range(max(s_ap[t], j0) - j0, s_ap[t + 1] - j0);
      if (lane == 0) {
        if (t == 0 && has_head) head_val[k] = sum;
        else spmv_store(y, i0 + t, sum, alpha, beta);
      }
    }
  } else {  // few rows: one warp per row
    for (int t = w; t < nrows; t += NW) {
      const T sum = warp_sum_range(max(s_ap[t], j0) - j0, s_ap[t + 1] - j0);
      if (lane == 0) {
        if (t == 0 && has_head) head_val[k] = sum;
        else spmv_store(y, i0 + t, sum, alpha, beta);
      }
    }
  }
  // the incomplete row: every warp sums a contiguous chunk, thread 0 adds the chunks in order
  if (has_open) {  // CTA-uniform
    const int ob = open_b - j0, ol = nz - ob;
    const int per = (ol + NW - 1) / NW;
    const int cb = ob + min(w * per, ol), ce = ob + min((w + 1) * per, ol);
    const T part = warp_sum_range(cb, ce);
    if (lane == 0) s_part[w] = part;
    __syncthreads();
    if (tid == 0) {
      T sum = T(0);
#pragma unroll
      for (int q = 0; q < NW; ++q) sum += s_part[q];
      if (open_is_head) head_val[k] = sum;
      else carry_val[k] = sum;
    }
  }
  if (tid == 0) {
    SpmvTileInfo ti;
    ti.head_row = (head_closed || open_is_head) ? i0 : -1;
    ti.head_open = open_is_head;
    ti.carry_row = (has_open && !open_is_head) ? i1 : -1;
    ti.pad = 0;
    info[k] = ti;
  }
}

// One thread per tile: the tile in which a multi-tile row is COMPLETED (head present, not open)
// adds the partials of that row in tile order - the carry of the tile where it started, the
// heads of the tiles that lie inside it, its own head - and writes the row.
template <typename T>
__global__ void __launch_bounds__(256)
k_spmv_fixup(int ntiles, const SpmvTileInfo *__restrict__ info, const T *__restrict__ head_val,
             const T *__restrict__ carry_val, T *__restrict__ y, T alpha, T beta) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= ntiles) return;
  const SpmvTileInfo me = info[t];
  if (me.head_row < 0 || me.head_open) return;
  int ta = t - 1;  // walk back over the tiles inside the row, to the tile where it starts
  while (ta > 0 && info[ta].carry_row != me.head_row) --ta;
  T sum = carry_val[ta];
  for (int q = ta + 1; q <= t; ++q) sum += head_val[q];
  spmv_store(y, me.head_row, sum, alpha, beta);
}

}  // namespace b200sp
