// spmv.cuh -- CSR SpMV / one-or-few-column SpMM kernels.
//
// Takes the place of cusparseSpMV / cusparseSpMM (reference call sites:
// modify_src/cupy-src/cupyx/cusparse.py:1430-1435, 1505-1511).  HBM-bound: the matrix streams
// once (nnz * (4 + s) bytes), x is gathered through L2.
#pragma once
#include "common.cuh"

namespace b200sp {

template <typename T>
__device__ __forceinline__ T fma_acc(T a, T b, T c) { return a * b + c; }

// "Vector" kernel: a group of G lanes per row (G picked on the host from nnz/m), the Aj/Ax
// reads of a warp cover 32/G adjacent rows, i.e. one contiguous span of the arrays.
// Column `col` of dense X / Y is selected with blockIdx.y (ldx, ldy = leading dimensions).
template <int G, typename T>
__global__ void __launch_bounds__(256)
k_spmv_vector(int m, const int *__restrict__ Ap, const int *__restrict__ Aj,
              const T *__restrict__ Ax, const T *__restrict__ X, long long ldx, T *__restrict__ Y,
              long long ldy, T alpha, T beta) {
  const int sub = threadIdx.x & (G - 1);
  const int groups_per_block = 256 / G;
  const T *x = X + (long long)blockIdx.y * ldx;
  T *y = Y + (long long)blockIdx.y * ldy;
  for (int row0 = blockIdx.x * groups_per_block; row0 < m; row0 += gridDim.x * groups_per_block) {
    const int row = row0 + threadIdx.x / G;
    T sum = T(0);
    if (row < m) {
      const int e = __ldg(Ap + row + 1);
      for (int t = __ldg(Ap + row) + sub; t < e; t += G) sum = fma_acc(__ldg(Ax + t), __ldg(x + __ldg(Aj + t)), sum);
    }
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) sum += __shfl_xor_sync(FULL, sum, o, G);
    if (row < m && sub == 0) y[row] = beta == T(0) ? alpha * sum : alpha * sum + beta * y[row];
  }
}

// ---- nnz-balanced "stream" kernel -----------------------------------------------------------
// Each CTA owns a tile of SPMV_TILE consecutive nonzeros regardless of the row structure (the
// merge-path idea: one heavy row cannot serialise a warp), reads them with coalesced 128-bit
// loads, stages the products in shared memory and reduces them per row.  A row that crosses
// tile boundaries leaves one partial per tile (carry = open at the tile's end, tail = closes
// in this tile after starting earlier); k_spmv_fixup adds them in tile order, so the result
// is run-to-run deterministic (no atomics).
constexpr int SPMV_TILE = 2048;
constexpr int SPMV_THREADS = 256;

struct SpmvTileInfo {
  int carry_row;  // row left open at the end of the tile, or -1
  int tail_row;   // row that started in an earlier tile and closes in this one, or -1
};

// the row r with Ap[r] <= pos < Ap[r+1]  (0 <= pos < nnz; empty rows are skipped)
__device__ __forceinline__ int row_of(const int *__restrict__ Ap, int m, int pos) {
  int lo = 0, hi = m - 1;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(Ap + mid + 1) > pos) hi = mid;
    else lo = mid + 1;
  }
  return lo;
}

template <typename T>
__device__ __forceinline__ void spmv_emit(int r, T sum, int r_lo, int r_hi, bool first_cont,
                                          bool last_open, T *__restrict__ y, T alpha, T beta,
                                          T *__restrict__ carry_val, T *__restrict__ tail_val) {
  if (r == r_hi && last_open) carry_val[blockIdx.x] = sum;       // continues in the next tile
  else if (r == r_lo && first_cont) tail_val[blockIdx.x] = sum;  // finished by the fix-up
  else y[r] = beta == T(0) ? alpha * sum : alpha * sum + beta * y[r];
}

template <typename T>
__global__ void __launch_bounds__(SPMV_THREADS)
k_spmv_stream(int m, int nnz, const int *__restrict__ Ap, const int *__restrict__ Aj,
              const T *__restrict__ Ax, const T *__restrict__ x, T *__restrict__ y, T alpha, T beta,
              SpmvTileInfo *__restrict__ info, T *__restrict__ carry_val,
              T *__restrict__ tail_val, int vec_ok) {
  __shared__ T s_prod[SPMV_TILE];
  __shared__ int s_rows[2];
  const int tid = threadIdx.x;
  const int t0 = blockIdx.x * SPMV_TILE;
  const int t1 = min(t0 + SPMV_TILE, nnz);
  if (tid == 0) s_rows[0] = row_of(Ap, m, t0);
  if (tid == 32) s_rows[1] = row_of(Ap, m, t1 - 1);
  if (vec_ok && t1 - t0 == SPMV_TILE) {
    // 16 bytes of indices + 16 bytes of values per lane and load
    constexpr int V = 16 / sizeof(T);  // values per 128-bit load
    const int4 *j4 = reinterpret_cast<const int4 *>(Aj + t0);
#pragma unroll
    for (int q = 0; q < SPMV_TILE / 4 / SPMV_THREADS; ++q) {
      const int e = q * SPMV_THREADS + tid;
      const int4 jj = __ldcs(j4 + e);
      T a[4];
      if (V == 4) {
        const float4 aa = __ldcs(reinterpret_cast<const float4 *>(Ax + t0) + e);
        a[0] = (T)aa.x; a[1] = (T)aa.y; a[2] = (T)aa.z; a[3] = (T)aa.w;
      } else {
        const double2 a0 = __ldcs(reinterpret_cast<const double2 *>(Ax + t0) + 2 * e);
        const double2 a1 = __ldcs(reinterpret_cast<const double2 *>(Ax + t0) + 2 * e + 1);
        a[0] = (T)a0.x; a[1] = (T)a0.y; a[2] = (T)a1.x; a[3] = (T)a1.y;
      }
      s_prod[4 * e + 0] = a[0] * __ldg(x + jj.x);
      s_prod[4 * e + 1] = a[1] * __ldg(x + jj.y);
      s_prod[4 * e + 2] = a[2] * __ldg(x + jj.z);
      s_prod[4 * e + 3] = a[3] * __ldg(x + jj.w);
    }
  } else {
    for (int e = t0 + tid; e < t1; e += SPMV_THREADS) s_prod[e - t0] = __ldcs(Ax + e) * __ldg(x + __ldcs(Aj + e));
  }
  __syncthreads();
  const int r_lo = s_rows[0], r_hi = s_rows[1];
  const bool last_open = __ldg(Ap + r_hi + 1) > t1;
  const bool first_cont = __ldg(Ap + r_lo) < t0;
  if (tid == 0) {
    SpmvTileInfo ti;
    ti.carry_row = last_open ? r_hi : -1;
    ti.tail_row = (first_cont && !(r_lo == r_hi && last_open)) ? r_lo : -1;
    info[blockIdx.x] = ti;
  }
  if (r_hi - r_lo + 1 >= 64) {  // many short rows: one thread per row
    for (int r = r_lo + tid; r <= r_hi; r += SPMV_THREADS) {
      const int b = max(__ldg(Ap + r), t0), e = min(__ldg(Ap + r + 1), t1);
      if (e <= b) continue;  // empty row: k_spmv_empty_rows owns it
      T sum = T(0);
      for (int q = b; q < e; ++q) sum += s_prod[q - t0];
      spmv_emit(r, sum, r_lo, r_hi, first_cont, last_open, y, alpha, beta, carry_val, tail_val);
    }
  } else {  // few long rows: one warp per row
    const int lane = tid & 31, w = tid >> 5;
    for (int r = r_lo + w; r <= r_hi; r += SPMV_THREADS / 32) {
      const int b = max(__ldg(Ap + r), t0), e = min(__ldg(Ap + r + 1), t1);
      if (e <= b) continue;  // empty row (warp-uniform)
      T sum = T(0);
      for (int q = b + lane; q < e; q += 32) sum += s_prod[q - t0];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(FULL, sum, o);
      if (lane == 0) spmv_emit(r, sum, r_lo, r_hi, first_cont, last_open, y, alpha, beta, carry_val, tail_val);
    }
  }
}

// One thread per tile: a tile whose first row closes here after starting earlier collects the
// carries of the preceding tiles (in order) and writes the row.
template <typename T>
__global__ void __launch_bounds__(256)
k_spmv_fixup(int ntiles, const SpmvTileInfo *__restrict__ info, const T *__restrict__ carry_val,
             const T *__restrict__ tail_val, T *__restrict__ y, T alpha, T beta) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= ntiles) return;
  const int r = info[t].tail_row;
  if (r < 0) return;
  int ta = t;
  while (ta > 0 && info[ta - 1].carry_row == r) --ta;
  T sum = T(0);
  for (int q = ta; q < t; ++q) sum += carry_val[q];
  sum += tail_val[t];
  y[r] = beta == T(0) ? alpha * sum : alpha * sum + beta * y[r];
}

// Rows without nonzeros are never visited by the stream kernel: y = beta * y there.
template <typename T>
__global__ void __launch_bounds__(256)
k_spmv_empty_rows(int m, const int *__restrict__ Ap, T *__restrict__ y, T beta) {
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < m; r += gridDim.x * blockDim.x)
    if (__ldg(Ap + r + 1) == __ldg(Ap + r)) y[r] = beta == T(0) ? T(0) : beta * y[r];
}

}  // namespace b200sp
