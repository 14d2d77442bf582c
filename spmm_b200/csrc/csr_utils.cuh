// csr_utils.cuh -- canonical-format check and in-row column sort of a CSR.
//
// Take the place of the has_canonical_format elementwise kernel + .all()
// (reference: modify_src/cupy-src/cupyx/scipy/sparse/_compressed.py:177-192) and of
// cusparseXcsrsort + gather behind sort_indices() (cupyx/cusparse.py:832-866).
#pragma once
#include <limits.h>
#include "common.cuh"
#include "spgemm_numeric.cuh"  // bitonic_sort

namespace b200sp {

// flag[0] is set to 1 when any row is not strictly increasing / in range / indptr not monotone.
__global__ void __launch_bounds__(256)
k_csr_check_canonical(int m, int ncols, const int *__restrict__ Ap, const int *__restrict__ Aj,
                      int *__restrict__ flag) {
  const int lane = lane_id();
  const int warps = (gridDim.x * blockDim.x) >> 5;
  int bad = 0;
  for (int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < m; r += warps) {
    const int b = __ldg(Ap + r), e = __ldg(Ap + r + 1);
    if (e < b || (r == 0 && b != 0)) bad = 1;
    for (int t = b + lane; t < e; t += 32) {
      const int j = __ldg(Aj + t);
      if (j < 0 || j >= ncols) bad = 1;
      if (t > b && __ldg(Aj + t - 1) >= j) bad = 1;
    }
  }
  if (__any_sync(FULL, bad) && lane == 0) atomicExch(flag, 1);
}

// CTA per row: rows of up to `cap` entries are sorted in shared memory (bitonic).  Rows already
// sorted are left untouched.  Longer rows, and rows that turn out to hold duplicate columns (the
// bitonic network is not stable: duplicates would come out in another order than csrsort's stable
// sort gives, and a later sum in storage order would round differently), are left as they are and
// set *too_long: the host then runs the stable radix sort over the matrix.
template <typename T>
__global__ void __launch_bounds__(256)
k_csr_sort_rows(int m, int cap, const int *__restrict__ Ap, int *__restrict__ Aj, T *__restrict__ Ax,
                int *__restrict__ too_long) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T *vals = reinterpret_cast<T *>(smem_raw);
  int *keys = reinterpret_cast<int *>(smem_raw + (size_t)cap * sizeof(T));
  const int tid = threadIdx.x;
  for (int r = blockIdx.x; r < m; r += gridDim.x) {
    const int b = Ap[r], len = Ap[r + 1] - b;
    if (len <= 1) continue;
    int unsorted = 0;
    for (int t = tid + 1; t < len; t += 256) unsorted |= (Aj[b + t - 1] > Aj[b + t]);
    if (!__syncthreads_or(unsorted)) continue;
    if (len > cap) {
      if (tid == 0) atomicExch(too_long, 1);
      continue;
    }
    const int P = next_pow2(len);
    for (int t = tid; t < P; t += 256) {
      keys[t] = t < len ? Aj[b + t] : INT_MAX;
      if (t < len) vals[t] = Ax[b + t];
    }
    __syncthreads();
    bitonic_sort(keys, vals, P, tid, 256, [] { __syncthreads(); });
    int dup = 0;
    for (int t = tid + 1; t < len; t += 256) dup |= (keys[t - 1] == keys[t]);
    if (__syncthreads_or(dup)) {  // CTA-uniform
      if (tid == 0) atomicExch(too_long, 1);
      continue;
    }
    for (int t = tid; t < len; t += 256) {
      Aj[b + t] = keys[t];
      Ax[b + t] = vals[t];
    }
    __syncthreads();
  }
}

}  // namespace b200sp
