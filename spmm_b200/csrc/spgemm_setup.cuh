// spgemm_setup.cuh -- symbolic pass 1 (intermediate products per row), row binning, scan.
//
// Takes the place of cusparseSpGEMM_workEstimation + cusparseSpGEMM_getNumProducts
// (reference: cupy_cusparse/spgemm_from_txt_alg3.cu:182-191).
#pragma once
#include "common.cuh"

namespace b200sp {

constexpr int MAX_BINS = 16;

// Bin limits, passed by value to the classify kernels.  A row of size s > 0 goes to the first
// bin b with s <= limit[b]; the last bin has limit INT_MAX.  Size 0 rows are in no bin.
struct BinLimits {
  int nbins;
  int limit[MAX_BINS];
};

__device__ __forceinline__ int classify(const BinLimits &bl, int s) {
  int b = 0;
#pragma unroll 1
  while (b < bl.nbins - 1 && s > bl.limit[b]) ++b;
  return b;
}

// Device-side counters of one SpGEMM plan (one cudaMemsetAsync clears them all).
struct Counters {
  unsigned long long total_products;
  unsigned long long nnz_c;
  int max_row_products;
  int overflow;  // a row's product count did not fit int32
  int b_bad;     // B is not canonical (unsorted row / index out of range): no column-blocked kernel
  int sym_count[MAX_BINS];
  int sym_cursor[MAX_BINS];
  int num_count[MAX_BINS];
  int num_cursor[MAX_BINS];
  unsigned long long total_ub;  // sum_i min(products_i, n): an upper bound of nnz(C) (single-pass mode)
  int ticket;                   // row ticket of the persistent symbolic kernels (zeroed before each launch)
  int pad_;
};

// prods[i] = sum over the nonzeros (i,k) of A of nnz(B_k); ub[i] = min(prods[i], n) is the
// upper bound on nnz(C_i) that sizes the symbolic hash tables.  G lanes per row (G chosen on
// the host from nnz(A)/m) so that the Aj reads coalesce.  Also: total, max, bin histogram,
// and Cp[i] = 0 for rows without products (the bins never visit them).
template <int G>
__global__ void __launch_bounds__(256)
k_row_products(int m, int n, const int *__restrict__ Ap, const int *__restrict__ Aj,
               const int *__restrict__ Bp, int *__restrict__ rowub, int *__restrict__ Cp,
               BinLimits bl, Counters *ctr, int *__restrict__ rowprod = nullptr) {
  __shared__ int s_hist[MAX_BINS];
  __shared__ unsigned long long s_total, s_ub;
  __shared__ int s_max;
  if (threadIdx.x < MAX_BINS) s_hist[threadIdx.x] = 0;
  if (threadIdx.x == 0) {
    s_total = 0;
    s_ub = 0;
    s_max = 0;
  }
  __syncthreads();
  const int sub = threadIdx.x & (G - 1);
  const int groups_per_block = 256 / G;
  long long my_total = 0, my_ub = 0;
  int my_max = 0;
  for (int row0 = blockIdx.x * groups_per_block; row0 < m; row0 += gridDim.x * groups_per_block) {
    const int row = row0 + threadIdx.x / G;
    long long s = 0;
    if (row < m) {
      const int e = __ldg(Ap + row + 1);
      for (int t = __ldg(Ap + row) + sub; t < e; t += G) {
        const int kk = __ldg(Aj + t);
        s += __ldg(Bp + kk + 1) - __ldg(Bp + kk);
      }
    }
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) s += __shfl_xor_sync(FULL, s, o, G);
    if (row < m && sub == 0) {
      if (s > 0x7fffffffLL) {
        atomicExch(&ctr->overflow, 1);
        s = 0x7fffffffLL;
      }
      const int p = (int)s;
      const int ub = min(p, n);
      rowub[row] = ub;
      if (rowprod) rowprod[row] = p;
      if (ub == 0) {
        Cp[row] = 0;
      } else {
        atomicAdd(&s_hist[classify(bl, ub)], 1);
      }
      my_total += s;
      my_ub += ub;
      my_max = max(my_max, p);
    }
  }
  my_total = warp_sum(my_total);
  my_ub = warp_sum(my_ub);
  my_max = __reduce_max_sync(FULL, my_max);
  if (lane_id() == 0) {
    atomicAdd(&s_total, (unsigned long long)my_total);
    atomicAdd(&s_ub, (unsigned long long)my_ub);
    atomicMax(&s_max, my_max);
  }
  __syncthreads();
  if (threadIdx.x < bl.nbins && s_hist[threadIdx.x]) atomicAdd(&ctr->sym_count[threadIdx.x], s_hist[threadIdx.x]);
  if (threadIdx.x == 0) {
    if (s_total) atomicAdd(&ctr->total_products, s_total);
    if (s_ub) atomicAdd(&ctr->total_ub, s_ub);
    if (s_max) atomicMax(&ctr->max_row_products, s_max);
  }
}

// Histogram of the exact row sizes nnz(C_i) = Cp[i+1]-Cp[i] for the numeric bins.
__global__ void __launch_bounds__(256)
k_hist_numeric(int m, const int *__restrict__ Cp, BinLimits bl, Counters *ctr) {
  __shared__ int s_hist[MAX_BINS];
  if (threadIdx.x < MAX_BINS) s_hist[threadIdx.x] = 0;
  __syncthreads();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x) {
    const int s = Cp[i + 1] - Cp[i];
    if (s > 0) atomicAdd(&s_hist[classify(bl, s)], 1);
  }
  __syncthreads();
  if (threadIdx.x < bl.nbins && s_hist[threadIdx.x]) atomicAdd(&ctr->num_count[threadIdx.x], s_hist[threadIdx.x]);
}

// Scatter row ids into per-bin contiguous lists (rows[bin_offset[b] ...]).  NUMERIC selects
// the size source.  Order inside a bin is arbitrary; results do not depend on it.
template <bool NUMERIC>
__global__ void __launch_bounds__(256)
k_scatter_rows(int m, const int *__restrict__ sizes /* rowub or Cp */, BinLimits bl, Counters *ctr,
               int *__restrict__ rows) {
  __shared__ int s_cnt[MAX_BINS];
  __shared__ int s_base[MAX_BINS];
  const int *count = NUMERIC ? ctr->num_count : ctr->sym_count;
  int *cursor = NUMERIC ? ctr->num_cursor : ctr->sym_cursor;
  const int tile = blockDim.x;
  for (int i0 = blockIdx.x * tile; i0 < m; i0 += gridDim.x * tile) {
    if (threadIdx.x < MAX_BINS) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    const int i = i0 + threadIdx.x;
    int s = 0, b = 0, local = 0;
    if (i < m) s = NUMERIC ? (sizes[i + 1] - sizes[i]) : sizes[i];
    if (s > 0) {
      b = classify(bl, s);
      local = atomicAdd(&s_cnt[b], 1);
    }
    __syncthreads();
    if (threadIdx.x < bl.nbins) {
      int off = 0;
      for (int q = 0; q < (int)threadIdx.x; ++q) off += count[q];
      const int c = s_cnt[threadIdx.x];
      s_base[threadIdx.x] = off + (c ? atomicAdd(&cursor[threadIdx.x], c) : 0);
    }
    __syncthreads();
    if (s > 0) rows[s_base[b] + local] = i;
    __syncthreads();
  }
}

// ------------------------------------------------------------------------ scan ----
// In-place exclusive scan of Cp[0..m): Cp[i] <- sum_{r<i} cnt[r], Cp[m] <- total.
// 64-bit running sums; the int32 result overflows only if total > INT32_MAX, which the host
// checks from ctr->nnz_c before anyone uses Cp.

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 16;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

// block-wide exclusive scan of one value per thread; returns exclusive prefix, *total = sum.
__device__ __forceinline__ long long block_exscan(long long v, long long *s_warp, long long *total) {
  const int lane = lane_id(), w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  long long incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const long long t = __shfl_up_sync(FULL, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_warp[w] = incl;
  __syncthreads();
  if (w == 0) {
    long long x = lane < nw ? s_warp[lane] : 0;
    long long xi = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const long long t = __shfl_up_sync(FULL, xi, o);
      if (lane >= o) xi += t;
    }
    if (lane < nw) s_warp[lane] = xi - x;
    if (lane == 31) s_warp[32] = xi;
  }
  __syncthreads();
  const long long res = s_warp[w] + incl - v;
  *total = s_warp[32];
  __syncthreads();
  return res;
}

// One tile: loads SCAN_ITEMS consecutive counts per thread, scans, stores exclusive + carry.
__device__ __forceinline__ long long scan_tile(int *Cp, int m, int tile_base, long long carry,
                                               long long *s_warp) {
  int item[SCAN_ITEMS];
  long long sum = 0;
  const int base = tile_base + threadIdx.x * SCAN_ITEMS;
#pragma unroll
  for (int q = 0; q < SCAN_ITEMS; ++q) {
    item[q] = (base + q < m) ? Cp[base + q] : 0;
    sum += item[q];
  }
  long long tile_total;
  long long ex = block_exscan(sum, s_warp, &tile_total) + carry;
#pragma unroll
  for (int q = 0; q < SCAN_ITEMS; ++q) {
    if (base + q < m) Cp[base + q] = (int)ex;
    ex += item[q];
  }
  return carry + tile_total;
}

// Small m (<= 64 K rows): one 1024-thread block, each thread owns a contiguous run of rows, one
// block-wide scan.  Also builds the numeric-bin histogram of the row sizes (saves a launch).
constexpr int SCAN1_THREADS = 1024;
constexpr int SCAN1_MAX_ITEMS = 64;
__global__ void __launch_bounds__(SCAN1_THREADS, 1)
k_scan_single(int *Cp, int m, BinLimits bl, Counters *ctr) {
  // warp w owns the contiguous rows [w*seg, (w+1)*seg), walked 32 at a time with coalesced
  // accesses: pass 1 sums the segment (and builds the histogram), pass 2 writes the scan.
  __shared__ long long s_wsum[32];
  __shared__ int s_hist[MAX_BINS];
  const int lane = lane_id(), w = threadIdx.x >> 5;
  if (threadIdx.x < MAX_BINS) s_hist[threadIdx.x] = 0;
  __syncthreads();
  const int seg = (((m + 31) / 32) + 31) & ~31;
  const int b = min(w * seg, m), e = min(b + seg, m);
  long long sum = 0;
  int mybin = 0;  // lane q accumulates the count of bin q (warp-aggregated: no same-address atomics)
  for (int i0 = b; i0 < e; i0 += 128) {
    int c[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) c[u] = (i0 + u * 32 + lane < e) ? Cp[i0 + u * 32 + lane] : 0;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      sum += c[u];
      const int bin = c[u] > 0 ? classify(bl, c[u]) : -1;
      for (int q = 0; q < bl.nbins; ++q) {
        const int k = __popc(__ballot_sync(FULL, bin == q));
        if (lane == q) mybin += k;
      }
    }
  }
  if (lane < bl.nbins && mybin) atomicAdd(&s_hist[lane], mybin);
  sum = warp_sum(sum);
  if (lane == 0) s_wsum[w] = sum;
  __syncthreads();
  long long total = 0, base = 0;
#pragma unroll 1
  for (int q = 0; q < 32; ++q) {
    const long long v = s_wsum[q];
    if (q < w) base += v;
    total += v;
  }
  for (int i0 = b; i0 < e; i0 += 128) {
    int c[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) c[u] = (i0 + u * 32 + lane < e) ? Cp[i0 + u * 32 + lane] : 0;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * 32 + lane;
      int incl = c[u];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(FULL, incl, o);
        if (lane >= o) incl += t;
      }
      if (i < e) Cp[i] = (int)(base + incl - c[u]);
      base += __shfl_sync(FULL, incl, 31);
    }
  }
  if (threadIdx.x == 0) {
    Cp[m] = (int)total;
    ctr->nnz_c = (unsigned long long)total;
  }
  if (threadIdx.x < bl.nbins && s_hist[threadIdx.x]) ctr->num_count[threadIdx.x] = s_hist[threadIdx.x];
}

// Large m: (1) per-tile sums, (2) one block scans the tile sums, (3) per-tile scan + offset.
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_tile_sums(const int *Cp, int m, long long *tile_sums) {
  __shared__ long long s_warp[33];
  const int base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
  long long sum = 0;
#pragma unroll
  for (int q = 0; q < SCAN_ITEMS; ++q) sum += (base + q < m) ? Cp[base + q] : 0;
  long long total;
  block_exscan(sum, s_warp, &total);
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(SCAN_THREADS)
k_scan_tile_offsets(long long *tile_sums, int ntiles, int *Cp, int m, Counters *ctr) {
  __shared__ long long s_warp[33];
  long long carry = 0;
  for (int b0 = 0; b0 < ntiles; b0 += SCAN_THREADS) {
    const int b = b0 + threadIdx.x;
    const long long v = b < ntiles ? tile_sums[b] : 0;
    long long total;
    const long long ex = block_exscan(v, s_warp, &total);
    if (b < ntiles) tile_sums[b] = carry + ex;
    carry += total;
  }
  if (threadIdx.x == 0) {
    Cp[m] = (int)carry;
    ctr->nnz_c = (unsigned long long)carry;
  }
}

__global__ void __launch_bounds__(SCAN_THREADS)
k_scan_apply(int *Cp, int m, const long long *tile_sums) {
  __shared__ long long s_warp[33];
  scan_tile(Cp, m, blockIdx.x * SCAN_TILE, tile_sums[blockIdx.x], s_warp);
}

// Rows of a bin ordered by decreasing work: 16 buckets by the power of two of their intermediate products
// relative to the largest row (bucket 15 = within a factor 2 of it, bucket 0 = 2^15 times lighter or more),
// heaviest bucket first, arbitrary order inside a bucket.  The heavy-row kernels hand rows out by ticket in
// list order: a hub row of a power-law matrix is a whole CTA's work for tens of milliseconds (R-MAT scale 20,
// row 0: ~8e7 products) and must start first, not wherever the scatter happened to put it
// (longest-processing-time-first; measured R-MAT scale 18 / 20 numeric pass 46.4 / 577 -> 43.5 / 510 ms with
// two buckets).
__device__ __forceinline__ int work_bucket(int prod, int top_log2) {
  const int l = 31 - __clz(max(prod, 1));
  return max(0, min(15, l - top_log2 + 15));
}
__global__ void __launch_bounds__(256)
k_work_hist(const int *__restrict__ rows, int cnt, const int *__restrict__ rowprod, int top_log2, int *__restrict__ hist) {
  __shared__ int s_h[16];
  if (threadIdx.x < 16) s_h[threadIdx.x] = 0;
  __syncthreads();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < cnt; i += gridDim.x * blockDim.x)
    atomicAdd(&s_h[work_bucket(rowprod[rows[i]], top_log2)], 1);
  __syncthreads();
  if (threadIdx.x < 16 && s_h[threadIdx.x]) atomicAdd(hist + threadIdx.x, s_h[threadIdx.x]);
}
__global__ void __launch_bounds__(256)
k_work_order(const int *__restrict__ rows, int cnt, const int *__restrict__ rowprod, int top_log2,
             const int *__restrict__ hist, int *__restrict__ cursor, int *__restrict__ out) {
  __shared__ int s_base[16];
  if (threadIdx.x < 16) {
    int base = 0;
    for (int q = 15; q > (int)threadIdx.x; --q) base += hist[q];  // the heavier buckets come first
    s_base[threadIdx.x] = base;
  }
  __syncthreads();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < cnt; i += gridDim.x * blockDim.x) {
    const int r = rows[i];
    const int bk = work_bucket(rowprod[r], top_log2);
    out[s_base[bk] + atomicAdd(cursor + bk, 1)] = r;
  }
}

}  // namespace b200sp
