// spgemm_symbolic.cuh -- exact nnz(C_i) per row (symbolic pass 2).
//
// Takes the place of the structure half of cusparseSpGEMM_compute
// (reference: cupy_cusparse/spgemm_from_txt_alg1.cu:177-181).  No value is touched here, so
// insertion order is free and all kernels use the whole group/CTA in parallel.
// The count of row r is written to Cp[r]; the scan turns the counts into C.indptr.
#pragma once
#include "common.cuh"

namespace b200sp {

// ---- group-of-G-lanes per row, hash set of S slots per group in shared memory ------------
// G = 8,  S = 64   : rows with <= 32 candidate columns  (4 rows per warp)
// G = 32, S = 256+ : rows with <= S/2 candidate columns (warp per row)
template <int G>
__global__ void __launch_bounds__(128)
k_sym_hash_group(const int *__restrict__ rows, int nrows, int S, const int *__restrict__ Ap,
                 const int *__restrict__ Aj, const int *__restrict__ Bp,
                 const int *__restrict__ Bj, int *__restrict__ Cp) {
  extern __shared__ int smem_i[];
  const int groups_per_block = 128 / G;
  const int g = threadIdx.x / G;
  const int sub = threadIdx.x & (G - 1);
  int *keys = smem_i + g * S;
  const unsigned mask = S - 1;
  const int ridx = blockIdx.x * groups_per_block + g;
  const int row = ridx < nrows ? rows[ridx] : -1;
  for (int s = sub; s < S; s += G) keys[s] = EMPTY;
  __syncwarp();
  int a_begin = 0, a_end = 0;
  if (row >= 0) {
    a_begin = __ldg(Ap + row);
    a_end = __ldg(Ap + row + 1);
  }
  int cnt = 0;
  stream_products<G>(
      a_begin, a_end, Aj, Bp, [&](bool act, int bidx, int) { return act ? __ldg(Bj + bidx) : 0; },
      [&](bool act, int j, bool) {
        if (act) {
          bool fresh;
          hash_insert(keys, mask, j, fresh);
          cnt += fresh;
        }
      });
  cnt = group_sum<G>(cnt);
  if (row >= 0 && sub == 0) Cp[row] = cnt;
}

// ---- CTA per row, hash set of S slots in shared memory ------------------------------------
// Each warp takes whole rows of B (A nonzero q = warp, warp + W, ...) and strides them.
__global__ void __launch_bounds__(256)
k_sym_hash_cta(const int *__restrict__ rows, int nrows, int S, const int *__restrict__ Ap,
               const int *__restrict__ Aj, const int *__restrict__ Bp,
               const int *__restrict__ Bj, int *__restrict__ Cp) {
  extern __shared__ int smem_i[];
  __shared__ int s_cnt;
  int *keys = smem_i;
  const unsigned mask = S - 1;
  const int lane = lane_id(), w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int ridx = blockIdx.x; ridx < nrows; ridx += gridDim.x) {
    const int row = rows[ridx];
    for (int s = threadIdx.x; s < S; s += blockDim.x) keys[s] = EMPTY;
    if (threadIdx.x == 0) s_cnt = 0;
    __syncthreads();
    const int a_end = __ldg(Ap + row + 1);
    int cnt = 0;
    for (int t = __ldg(Ap + row) + w; t < a_end; t += nw) {
      const int kk = __ldg(Aj + t);
      const int be = __ldg(Bp + kk + 1);
      for (int u = __ldg(Bp + kk) + lane; u < be; u += 32) {
        bool fresh;
        hash_insert(keys, mask, __ldg(Bj + u), fresh);
        cnt += fresh;
      }
    }
    cnt = warp_sum(cnt);
    if (lane == 0 && cnt) atomicAdd(&s_cnt, cnt);
    __syncthreads();
    if (threadIdx.x == 0) Cp[row] = s_cnt;
    __syncthreads();
  }
}

// The column indices of the B rows named by A's row [a_begin, a_end), warp w of nw taking every
// nw-th A nonzero: f(column) for every product.  The chain A.indices -> B.indptr -> B.indices is
// three dependent memory round trips per B row; the pointers of the next B row and the A index
// after it are requested while the current row is walked, and a row is read 128 entries per
// trip (four independent loads per lane), so a warp waits for about one round trip per row.
template <class F>
__device__ __forceinline__ void walk_b_columns(int a_begin, int a_end, int w, int nw, int lane,
                                               const int *__restrict__ Aj, const int *__restrict__ Bp,
                                               const int *__restrict__ Bj, F &&f) {
  int t = a_begin + w;
  if (t >= a_end) return;
  int kk = __ldg(Aj + t);
  int kk2 = t + nw < a_end ? __ldg(Aj + t + nw) : -1;
  int bs = __ldg(Bp + kk), be = __ldg(Bp + kk + 1);
  while (true) {
    int bs2 = 0, be2 = 0;
    if (kk2 >= 0) {
      bs2 = __ldg(Bp + kk2);
      be2 = __ldg(Bp + kk2 + 1);
    }
    const int kk3 = t + 2 * nw < a_end ? __ldg(Aj + t + 2 * nw) : -1;
    for (int u = bs + lane; u < be; u += 128) {
      const int c0 = __ldg(Bj + u);
      const int c1 = u + 32 < be ? __ldg(Bj + u + 32) : -1;
      const int c2 = u + 64 < be ? __ldg(Bj + u + 64) : -1;
      const int c3 = u + 96 < be ? __ldg(Bj + u + 96) : -1;
      f(c0);
      if (c1 >= 0) f(c1);
      if (c2 >= 0) f(c2);
      if (c3 >= 0) f(c3);
    }
    if (kk2 < 0) break;
    t += nw;
    bs = bs2;
    be = be2;
    kk2 = kk3;
  }
}

// ---- CTA per row, bitmap over the n columns -----------------------------------------------
// SMEM = true : bitmap in dynamic shared memory (n/8 bytes; n <= ~1.7 M columns)
// SMEM = false: bitmap in the CTA's slice of the global workspace (any n); persistent grid.
// The bitmap is zero on entry to every row: the count sweep clears what it counts.
template <bool SMEM>
__global__ void __launch_bounds__(1024)
k_sym_bitmap_cta(const int *__restrict__ rows, int nrows, int n, const int *__restrict__ Ap,
                 const int *__restrict__ Aj, const int *__restrict__ Bp,
                 const int *__restrict__ Bj, int *__restrict__ Cp, unsigned *__restrict__ gws,
                 int *__restrict__ ticket = nullptr) {
  extern __shared__ int smem_i[];
  __shared__ int s_cnt;
  __shared__ int s_row;
  const int nwords = (n + 31) >> 5;
  unsigned *bm = SMEM ? reinterpret_cast<unsigned *>(smem_i) : gws + (size_t)blockIdx.x * nwords;
  const int lane = lane_id(), w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  if (SMEM) {
    for (int s = threadIdx.x; s < nwords; s += blockDim.x) bm[s] = 0u;
  }  // the global workspace is zeroed by the host before launch
  if (threadIdx.x == 0) s_cnt = 0;
  __syncthreads();
  // ticket != nullptr: rows are taken in list order as CTAs come free (the host orders the list by decreasing
  // work): the rows of a power-law matrix differ 1000-fold in products, a fixed stride leaves most CTAs idle
  // while a few finish their hubs
  for (int ridx = blockIdx.x;; ridx += gridDim.x) {
    if (ticket) {
      if (threadIdx.x == 0) s_row = atomicAdd(ticket, 1);
      __syncthreads();
      ridx = s_row;
    }
    if (ridx >= nrows) break;
    const int row = rows[ridx];
    walk_b_columns(__ldg(Ap + row), __ldg(Ap + row + 1), w, nw, lane, Aj, Bp, Bj,
                   [&](int j) { atomicOr(&bm[j >> 5], 1u << (j & 31)); });
    __syncthreads();
    int cnt = 0;
    for (int s = threadIdx.x; s < nwords; s += blockDim.x) {
      const unsigned bits = SMEM ? bm[s] : __ldcg(bm + s);
      if (bits) {
        cnt += __popc(bits);
        bm[s] = 0u;
      }
    }
    cnt = warp_sum(cnt);
    if (lane == 0 && cnt) atomicAdd(&s_cnt, cnt);
    __syncthreads();
    if (threadIdx.x == 0) {
      Cp[row] = s_cnt;
      s_cnt = 0;
    }
    __syncthreads();
  }
}

// ---- CTA per row, one byte flag per column in shared memory --------------------------------
// For n <= ~200 K columns.  Marking a column is a plain byte store (idempotent, so no atomic
// and no ordering between warps); the count is a popcount sweep over the flag words, which
// also clears them for the next row.
__global__ void __launch_bounds__(256, 2)
k_sym_flags_cta(const int *__restrict__ rows, int nrows, int n, const int *__restrict__ Ap,
                const int *__restrict__ Aj, const int *__restrict__ Bp,
                const int *__restrict__ Bj, int *__restrict__ Cp, int *__restrict__ ticket = nullptr) {
  extern __shared__ int smem_i[];
  __shared__ int s_cnt;
  __shared__ int s_row;
  unsigned char *flags = reinterpret_cast<unsigned char *>(smem_i);
  unsigned *fw = reinterpret_cast<unsigned *>(smem_i);
  const int nwords = (n + 3) >> 2;
  const int lane = lane_id(), w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int s = threadIdx.x; s < nwords; s += blockDim.x) fw[s] = 0u;
  if (threadIdx.x == 0) s_cnt = 0;
  __syncthreads();
  for (int ridx = blockIdx.x;; ridx += gridDim.x) {
    if (ticket) {  // see k_sym_bitmap_cta
      if (threadIdx.x == 0) s_row = atomicAdd(ticket, 1);
      __syncthreads();
      ridx = s_row;
    }
    if (ridx >= nrows) break;
    const int row = rows[ridx];
    // plain walk: with 64 resident warps per SM the kernel is bound by the L1 / shared-memory
    // pipe, and the pipelined walk_b_columns only adds instructions here (C3: 0.36 -> 0.39 ms)
    const int a_end = __ldg(Ap + row + 1);
    for (int t = __ldg(Ap + row) + w; t < a_end; t += nw) {
      const int kk = __ldg(Aj + t);
      const int be = __ldg(Bp + kk + 1);
      for (int u = __ldg(Bp + kk) + lane; u < be; u += 32) flags[__ldg(Bj + u)] = 1;
    }
    __syncthreads();
    int cnt = 0;
    for (int s = threadIdx.x; s < nwords; s += blockDim.x) {
      const unsigned f = fw[s];
      if (f) {
        cnt += __popc(f);
        fw[s] = 0u;
      }
    }
    cnt = warp_sum(cnt);
    if (lane == 0 && cnt) atomicAdd(&s_cnt, cnt);
    __syncthreads();
    if (threadIdx.x == 0) {
      Cp[row] = s_cnt;
      s_cnt = 0;
    }
    __syncthreads();
  }
}

}  // namespace b200sp
