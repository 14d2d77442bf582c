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

// ---- nnz-balanced "tile" kernel (merge-path decomposition) ------------------------------------
// Each tile is SPMV_THREADS * ITEMS consecutive nonzeros regardless of the row structure (one
// heavy row cannot serialise a warp); the tile owns every row that STARTS inside it, empty rows
// included.  The part of a tile that belongs to a row begun in an earlier tile ("head") and a
// row left open at the tile's end ("carry") become per-tile partials; k_spmv_fixup adds them
// in tile order, so the result is run-to-run deterministic (no atomics).
constexpr int SPMV_THREADS = 256;
constexpr int SPMV_ITEMS = 8;                           // nonzeros per thread and tile
constexpr int SPMV_TILE = SPMV_THREADS * SPMV_ITEMS;    // 2048
constexpr int SPMV_WINDOW = 1024;                       // indptr entries staged per batch of rows

struct SpmvTileInfo {
  int head_row;    // row begun earlier whose entries open this tile, or -1
  int head_open;   // 1 when that row continues past the end of this tile
  int carry_row;   // row that starts in this tile and continues into the next, or -1
  int pad;
};

// CTA-wide lower bound over the sorted int array a[0..len): first index whose value is >= key.
// Every thread of the CTA must call it; `len` and `key` are CTA-uniform.  256-ary: 3 rounds of
// one load + one barrier for 2^20 rows instead of the 20 dependent loads of a binary search.
__device__ __forceinline__ int cta_lower_bound(const int *__restrict__ a, int len, int key) {
  int lo = 0, hi = len;  // invariant: a[i] < key for i < lo, a[i] >= key for i >= hi
  while (hi - lo > SPMV_THREADS) {
    const int step = (hi - lo + SPMV_THREADS - 1) / SPMV_THREADS;
    const long long idx = (long long)lo + (long long)(threadIdx.x + 1) * step - 1;
    const int less = (idx < hi) && (__ldg(a + idx) < key);
    const int cnt = __syncthreads_count(less);  // probes are monotone: the first cnt are < key
    const long long nlo = (long long)lo + (long long)cnt * step;
    const long long nhi = nlo + step - 1;
    lo = (int)nlo;
    hi = (int)(nhi < hi ? nhi : hi);
  }
  const int idx = lo + threadIdx.x;
  const int less = (idx < hi) && (__ldg(a + idx) < key);
  return lo + __syncthreads_count(less);
}

template <typename T>
__device__ __forceinline__ void spmv_store(T *__restrict__ y, int r, T sum, T alpha, T beta) {
  y[r] = beta == T(0) ? alpha * sum : alpha * sum + beta * y[r];
}

// 128-bit streaming loads of one full tile: entry 4e+u of the tile lands in slot 4q+u,
// e = q*THREADS + tid.
template <typename T>
__device__ __forceinline__ void spmv_load_tile(const int *__restrict__ Aj, const T *__restrict__ Ax,
                                               int t0, int tid, int (&jj)[SPMV_ITEMS], T (&aa)[SPMV_ITEMS]) {
#pragma unroll
  for (int q = 0; q < SPMV_ITEMS / 4; ++q) {
    const int e = q * SPMV_THREADS + tid;
    const int4 j4 = __ldcs(reinterpret_cast<const int4 *>(Aj + t0) + e);
    jj[4 * q + 0] = j4.x; jj[4 * q + 1] = j4.y; jj[4 * q + 2] = j4.z; jj[4 * q + 3] = j4.w;
    if (sizeof(T) == 4) {
      const float4 a4 = __ldcs(reinterpret_cast<const float4 *>(Ax + t0) + e);
      aa[4 * q + 0] = (T)a4.x; aa[4 * q + 1] = (T)a4.y; aa[4 * q + 2] = (T)a4.z; aa[4 * q + 3] = (T)a4.w;
    } else {
      const double2 a0 = __ldcs(reinterpret_cast<const double2 *>(Ax + t0) + 2 * e);
      const double2 a1 = __ldcs(reinterpret_cast<const double2 *>(Ax + t0) + 2 * e + 1);
      aa[4 * q + 0] = (T)a0.x; aa[4 * q + 1] = (T)a0.y; aa[4 * q + 2] = (T)a1.x; aa[4 * q + 3] = (T)a1.y;
    }
  }
}

// Persistent CTAs: CTA c owns the contiguous tiles [c*ntiles/G, (c+1)*ntiles/G).  Only the first
// tile needs the indptr search; after that r_begin(k+1) = r_end(k).  The rows of a tile are
// walked in batches of SPMV_WINDOW indptr entries staged in shared memory (one batch for all but
// hypersparse tiles), which both finds r_end and feeds the row reduction.
// Software pipeline over tiles, so that the gather path keeps working while a tile is reduced:
//   iteration k:  store products(k) to smem | request indptr window(k), x gathers(k+1),
//                 A vectors(k+2) | barrier | reduce tile k from smem | barrier
template <typename T>
__global__ void __launch_bounds__(SPMV_THREADS, sizeof(T) == 4 ? 3 : 2)
k_spmv_tile(int m, int nnz, int ntiles, const int *__restrict__ Ap, const int *__restrict__ Aj,
            const T *__restrict__ Ax, const T *__restrict__ x, T *__restrict__ y, T alpha, T beta,
            SpmvTileInfo *__restrict__ info, T *__restrict__ head_val, T *__restrict__ carry_val,
            int vec_ok) {
  constexpr int ITEMS = SPMV_ITEMS, TILE = SPMV_TILE;
  constexpr int WPT = SPMV_WINDOW / SPMV_THREADS;  // window entries per thread
  constexpr int NW = SPMV_THREADS / 32;
  __shared__ __align__(16) T s_prod[TILE + 4];
  __shared__ int s_ap[SPMV_WINDOW + 1];
  __shared__ T s_part[NW];
  __shared__ int s_cnt[NW];
  __shared__ int s_carry_row;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int k0 = (int)((long long)ntiles * blockIdx.x / gridDim.x);
  const int k1 = (int)((long long)ntiles * (blockIdx.x + 1) / gridDim.x);
  if (k0 >= k1) return;
  auto tile_full = [&](int k) { return vec_ok && k < k1 && (nnz - k * TILE >= TILE); };
  int jj[ITEMS], jn[ITEMS];
  T aa[ITEMS], an[ITEMS], pr[ITEMS], xg[ITEMS];
  // prologue: products of tile k0 in pr, A vectors of tile k0+1 in jj/aa
  if (tile_full(k0)) spmv_load_tile<T>(Aj, Ax, k0 * TILE, tid, jj, aa);
  int r_begin = cta_lower_bound(Ap, m + 1, k0 * TILE);  // overlaps the loads above
  if (tile_full(k0)) {
#pragma unroll
    for (int u = 0; u < ITEMS; ++u) xg[u] = __ldg(x + jj[u]);
#pragma unroll
    for (int u = 0; u < ITEMS; ++u) pr[u] = aa[u] * xg[u];
  }
  if (tile_full(k0 + 1)) spmv_load_tile<T>(Aj, Ax, (k0 + 1) * TILE, tid, jj, aa);

  for (int k = k0; k < k1; ++k) {
    const int t0 = k * TILE;
    const int t1 = (nnz - t0 > TILE) ? t0 + TILE : nnz;
    const bool last_tile = (t1 == nnz);
    // products of this tile -> shared memory
    if (tile_full(k)) {
#pragma unroll
      for (int q = 0; q < ITEMS / 4; ++q) {
        const int e = q * SPMV_THREADS + tid;
#pragma unroll
        for (int u = 0; u < 4; ++u) s_prod[4 * e + u] = pr[4 * q + u];
      }
    } else {
      for (int e = t0 + tid; e < t1; e += SPMV_THREADS) s_prod[e - t0] = __ldcs(Ax + e) * __ldg(x + __ldcs(Aj + e));
    }
    if (tid == 0) s_carry_row = -1;
    // requests: first indptr window of this tile, x gathers of the next, A vectors of the one after
    int apw[WPT];
#pragma unroll
    for (int q = 0; q < WPT; ++q) apw[q] = __ldg(Ap + min(r_begin + q * SPMV_THREADS + tid, m));
    int ap_last = (tid == 0) ? __ldg(Ap + min(r_begin + SPMV_WINDOW, m)) : 0;
    const bool next_full = tile_full(k + 1), next2_full = tile_full(k + 2);
    if (next_full) {
#pragma unroll
      for (int u = 0; u < ITEMS; ++u) xg[u] = __ldg(x + jj[u]);
    }
    if (next2_full) spmv_load_tile<T>(Aj, Ax, t0 + 2 * TILE, tid, jn, an);

    int r = r_begin;  // first row of the current batch
    int head_row = -1, head_open = 0;
    for (bool first = true;; first = false) {
      if (!first) {  // hypersparse tile: more than SPMV_WINDOW rows start here
#pragma unroll
        for (int q = 0; q < WPT; ++q) apw[q] = __ldg(Ap + min(r + q * SPMV_THREADS + tid, m));
        ap_last = (tid == 0) ? __ldg(Ap + min(r + SPMV_WINDOW, m)) : 0;
      }
      int below = 0;  // batch entries (rows r + i) that start before t1; Ap[m] = nnz >= t1 never counts
#pragma unroll
      for (int q = 0; q < WPT; ++q) {
        s_ap[q * SPMV_THREADS + tid] = apw[q];
        below += (apw[q] < t1);
      }
      if (tid == 0) s_ap[SPMV_WINDOW] = ap_last;
      below = warp_sum(below);
      if (lane == 0) s_cnt[w] = below;
      __syncthreads();  // s_prod, s_ap, s_cnt visible
      int nbelow = 0;
#pragma unroll
      for (int q = 0; q < NW; ++q) nbelow += s_cnt[q];
      if (first) {
        // head: entries [t0, h_end) continue row r_begin-1 (Ap[r_begin] >= t0, r_begin <= m)
        const int p_begin = s_ap[0];
        const int h_end = min(p_begin, t1);
        if (h_end > t0) {  // CTA-uniform
          T part = T(0);
          const int hl = h_end - t0;
          const int per = (hl + NW - 1) / NW;  // contiguous chunk per warp
          const int cb = min(w * per, hl), ce = min(cb + per, hl);
          for (int q = cb + lane; q < ce; q += 32) part += s_prod[q];
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(FULL, part, o);
          if (lane == 0) s_part[w] = part;
          __syncthreads();
          if (tid == 0) {
            T sum = T(0);
#pragma unroll
            for (int q = 0; q < NW; ++q) sum += s_part[q];
            head_val[k] = sum;
          }
          head_row = r_begin - 1;
          head_open = p_begin > t1;  // the whole tile lies inside that row
        }
      }
      // rows of this batch that start in the tile (the last tile also owns trailing empty rows)
      const int rows_here = last_tile ? min(SPMV_WINDOW, m - r) : nbelow;
      if (rows_here >= 64) {  // many short rows: one thread per row
        for (int i = tid; i < rows_here; i += SPMV_THREADS) {
          const int nxt = s_ap[i + 1];
          const int b = s_ap[i] - t0, e = min(nxt, t1) - t0;
          T sum = T(0);
          for (int q = b; q < e; q += 4) {  // four loads in flight; the additions stay in order
            const T v0 = s_prod[q];
            const T v1 = q + 1 < e ? s_prod[q + 1] : T(0);
            const T v2 = q + 2 < e ? s_prod[q + 2] : T(0);
            const T v3 = q + 3 < e ? s_prod[q + 3] : T(0);
            sum += v0; sum += v1; sum += v2; sum += v3;
          }
          if (!last_tile && nxt > t1) {  // the one row left open at the end of the tile
            carry_val[k] = sum;
            s_carry_row = r + i;
          } else {
            spmv_store(y, r + i, sum, alpha, beta);
          }
        }
      } else {  // few long rows: one warp per row
        for (int i = w; i < rows_here; i += NW) {
          const int nxt = s_ap[i + 1];
          const int b = s_ap[i] - t0, e = min(nxt, t1) - t0;
          T sum = T(0);
          for (int q = b + lane; q < e; q += 32) sum += s_prod[q];
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(FULL, sum, o);
          if (lane == 0) {
            if (!last_tile && nxt > t1) {
              carry_val[k] = sum;
              s_carry_row = r + i;
            } else {
              spmv_store(y, r + i, sum, alpha, beta);
            }
          }
        }
      }
      r += rows_here;
      const bool done = last_tile ? (r >= m) : (nbelow < SPMV_WINDOW);
      __syncthreads();  // s_ap / s_cnt / s_prod / s_carry_row settled
      if (done) break;
    }
    if (tid == 0) {
      SpmvTileInfo ti;
      ti.head_row = head_row;
      ti.head_open = head_open;
      ti.carry_row = s_carry_row;
      ti.pad = 0;
      info[k] = ti;
    }
    r_begin = r;
    if (next_full) {
#pragma unroll
      for (int u = 0; u < ITEMS; ++u) pr[u] = aa[u] * xg[u];
    }
    if (next2_full) {
#pragma unroll
      for (int u = 0; u < ITEMS; ++u) {
        jj[u] = jn[u];
        aa[u] = an[u];
      }
    }
  }
}

// One thread per tile: the tile in which a multi-tile row ENDS (head present, not open) adds the
// partials of that row in tile order - the carry of the tile where it started, the heads of the
// tiles it covered completely, its own head - and writes the row.
template <typename T>
__global__ void __launch_bounds__(256)
k_spmv_fixup(int ntiles, const SpmvTileInfo *__restrict__ info, const T *__restrict__ head_val,
             const T *__restrict__ carry_val, T *__restrict__ y, T alpha, T beta) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= ntiles) return;
  const SpmvTileInfo me = info[t];
  if (me.head_row < 0 || me.head_open) return;
  int ta = t - 1;  // walk back over the tiles the row covers completely, to the tile where it starts
  while (ta > 0 && info[ta].carry_row != me.head_row) --ta;
  T sum = carry_val[ta];
  for (int q = ta + 1; q <= t; ++q) sum += head_val[q];
  spmv_store(y, me.head_row, sum, alpha, beta);
}

}  // namespace b200sp
