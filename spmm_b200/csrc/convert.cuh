// convert.cuh -- device canonicalisation: COO -> CSR, CSC <-> CSR, duplicate summation.
//
// Take the place of the conversions the reference runs before every non-CSR product
// (SpGEMM_vs_SpMV: {csc,coo} x {csr,csc,coo} operands) and inside its generators:
//   coo_matrix.sum_duplicates + tocsr   modify_src/cupy-src/cupyx/scipy/sparse/_coo.py:356-459, 517-540
//                                       (cusparse.coosort :908-955, coo2csr :957-970)
//   csc_matrix.tocsr / csr.tocsc        _csc.py:304-326, _csr.py:455-477 (cusparse.csc2csr :1092-1114,
//                                       csr2csc :1014-1036)
//   csr_matrix.sum_duplicates           _compressed.py:971-991
//
// One primitive does all of them: a stable LSD radix sort (8-bit digits) of 64-bit keys
// (row << 32 | col) with a 32-bit payload (position in the input), run only over the bit
// ranges that can differ.  Stability is the contract: duplicates end up in INPUT order and are
// summed sequentially in that order with separately rounded adds, so the result is
// deterministic (the reference sums duplicates with floating-point atomics, _coo.py:437-447,
// i.e. in no fixed order).
#pragma once
#include "common.cuh"

namespace b200sp {

constexpr int RS_THREADS = 256;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;  // keys per CTA and pass
constexpr int RS_WARPS = RS_THREADS / 32;

typedef unsigned long long u64;

// ------------------------------------------------------------------ generic int32 exclusive scan
constexpr int GS_THREADS = 256;
constexpr int GS_ITEMS = 16;
constexpr int GS_TILE = GS_THREADS * GS_ITEMS;

__device__ __forceinline__ int block_exscan_i32(int v, int *s_warp /*>=9*/, int *total) {
  const int lane = lane_id(), w = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(FULL, inc, o);
    if (lane >= o) inc += t;
  }
  __syncthreads();  // s_warp may still be read from a previous call
  if (lane == 31) s_warp[w] = inc;
  __syncthreads();
  int base = 0, tot = 0;
#pragma unroll
  for (int q = 0; q < GS_THREADS / 32; ++q) {
    const int s = s_warp[q];
    if (q < w) base += s;
    tot += s;
  }
  *total = tot;
  return base + inc - v;
}

__global__ void __launch_bounds__(GS_THREADS) k_gscan_sums(const int *__restrict__ in, long long n,
                                                             int *__restrict__ sums) {
  __shared__ int s_warp[GS_THREADS / 32];
  const long long base = (long long)blockIdx.x * GS_TILE;
  int sum = 0;
#pragma unroll
  for (int q = 0; q < GS_ITEMS; ++q) {
    const long long i = base + q * GS_THREADS + threadIdx.x;
    sum += i < n ? in[i] : 0;
  }
  int total;
  block_exscan_i32(sum, s_warp, &total);
  if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

// one CTA: exclusive scan of the tile sums in place; total -> *out_total (64-bit)
__global__ void __launch_bounds__(GS_THREADS) k_gscan_offsets(int *sums, int ntiles, long long *out_total) {
  __shared__ int s_warp[GS_THREADS / 32];
  long long carry = 0;
  for (int b0 = 0; b0 < ntiles; b0 += GS_THREADS) {
    const int b = b0 + threadIdx.x;
    const int v = b < ntiles ? sums[b] : 0;
    int total;
    const int ex = block_exscan_i32(v, s_warp, &total);
    if (b < ntiles) sums[b] = (int)(carry + ex);
    carry += total;
  }
  if (threadIdx.x == 0 && out_total) *out_total = carry;
}

__global__ void __launch_bounds__(GS_THREADS) k_gscan_apply(const int *in, int *out /* may alias in */,
                                                              long long n, const int *__restrict__ sums) {
  __shared__ int s_warp[GS_THREADS / 32];
  const long long base = (long long)blockIdx.x * GS_TILE + (long long)threadIdx.x * GS_ITEMS;
  int v[GS_ITEMS], sum = 0;
#pragma unroll
  for (int q = 0; q < GS_ITEMS; ++q) {
    v[q] = base + q < n ? in[base + q] : 0;
    sum += v[q];
  }
  int total;
  int run = sums[blockIdx.x] + block_exscan_i32(sum, s_warp, &total);
#pragma unroll
  for (int q = 0; q < GS_ITEMS; ++q) {
    if (base + q < n) out[base + q] = run;
    run += v[q];
  }
}

// ------------------------------------------------------------------ stable LSD radix sort pass
// counts[d * ntiles + tile] = number of keys of `tile` whose digit is d
__global__ void __launch_bounds__(RS_THREADS) k_rs_count(const u64 *__restrict__ keys, long long n, int shift,
                                                           int ntiles, int *__restrict__ counts) {
  __shared__ int s_hist[256];
  s_hist[threadIdx.x] = 0;
  __syncthreads();
  const long long base = (long long)blockIdx.x * RS_TILE;
#pragma unroll 4
  for (int q = 0; q < RS_ITEMS; ++q) {
    const long long i = base + q * RS_THREADS + threadIdx.x;
    if (i < n) atomicAdd(&s_hist[(int)((keys[i] >> shift) & 255)], 1);
  }
  __syncthreads();
  counts[(size_t)threadIdx.x * ntiles + blockIdx.x] = s_hist[threadIdx.x];
}

// offsets = exclusive scan of counts (digit-major).  Warp w owns a contiguous run of the tile
// and walks it 32 keys at a time; equal digits inside a round are ranked in lane order
// (__match_any_sync), rounds in order, warps in order => stable.
__global__ void __launch_bounds__(RS_THREADS) k_rs_scatter(const u64 *__restrict__ keys_in,
                                                             const int *__restrict__ perm_in /* null = identity */,
                                                             u64 *__restrict__ keys_out, int *__restrict__ perm_out,
                                                             long long n, int shift, int ntiles,
                                                             const int *__restrict__ offsets) {
  __shared__ int s_cnt[RS_WARPS][256];
  const int lane = lane_id(), w = threadIdx.x >> 5;
  for (int t = threadIdx.x; t < RS_WARPS * 256; t += RS_THREADS) (&s_cnt[0][0])[t] = 0;
  __syncthreads();
  const long long wbase = (long long)blockIdx.x * RS_TILE + (long long)w * (32 * RS_ITEMS);
  u64 key[RS_ITEMS];
  int rank[RS_ITEMS];
#pragma unroll
  for (int q = 0; q < RS_ITEMS; ++q) {
    const long long i = wbase + q * 32 + lane;
    key[q] = i < n ? keys_in[i] : 0ull;
  }
  const unsigned lt = (1u << lane) - 1u;
#pragma unroll
  for (int q = 0; q < RS_ITEMS; ++q) {
    const long long i = wbase + q * 32 + lane;
    const bool valid = i < n;
    const int d = (int)((key[q] >> shift) & 255);
    const unsigned peers = __match_any_sync(FULL, valid ? d : 256 + lane);
    const int r = __popc(peers & lt);
    int base = 0;
    if (valid && r == 0) {
      base = s_cnt[w][d];
      s_cnt[w][d] = base + __popc(peers);
    }
    base = __shfl_sync(FULL, base, __ffs(peers) - 1);
    rank[q] = base + r;
    __syncwarp();
  }
  __syncthreads();
  {
    const int d = threadIdx.x;
    int run = offsets[(size_t)d * ntiles + blockIdx.x];
#pragma unroll
    for (int q = 0; q < RS_WARPS; ++q) {
      const int c = s_cnt[q][d];
      s_cnt[q][d] = run;
      run += c;
    }
  }
  __syncthreads();
#pragma unroll
  for (int q = 0; q < RS_ITEMS; ++q) {
    const long long i = wbase + q * 32 + lane;
    if (i < n) {
      const int d = (int)((key[q] >> shift) & 255);
      const int dst = s_cnt[w][d] + rank[q];
      keys_out[dst] = key[q];
      perm_out[dst] = perm_in ? perm_in[i] : (int)i;
    }
  }
}

// ------------------------------------------------------------------ key builders
// COO: key = row << 32 | col; out-of-range coordinates raise *bad.
__global__ void __launch_bounds__(256) k_keys_from_coo(long long nnz, int m, int n, const int *__restrict__ row,
                                                        const int *__restrict__ col, u64 *__restrict__ keys,
                                                        int *__restrict__ bad) {
  int b = 0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (long long)gridDim.x * blockDim.x) {
    const int r = row[i], c = col[i];
    if (r < 0 || r >= m || c < 0 || c >= n) b = 1;
    keys[i] = ((u64)(unsigned)r << 32) | (unsigned)c;
  }
  if (__any_sync(FULL, b) && lane_id() == 0) atomicExch(bad, 1);
}

// Compressed (CSR or CSC) input: the major index of entry t is the segment that holds it.
// MAJOR_HIGH: key = major << 32 | minor (CSR -> keys for a full sort);
// else:       key = minor << 32 | major (CSC -> row-major keys; CSR -> column-major keys).
// Warp per segment.  Raises *bad for a non-monotone pointer array or an out-of-range minor.
template <bool MAJOR_HIGH>
__global__ void __launch_bounds__(256) k_keys_from_compressed(int nmajor, int nminor, long long nnz,
                                                               const int *__restrict__ ptr,
                                                               const int *__restrict__ idx, u64 *__restrict__ keys,
                                                               int *__restrict__ bad) {
  const int lane = lane_id();
  const int warps = (gridDim.x * blockDim.x) >> 5;
  int b = 0;
  for (int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; s < nmajor; s += warps) {
    const int t0 = ptr[s], t1 = ptr[s + 1];
    if (t0 < 0 || t1 < t0 || t1 > nnz || (s == 0 && t0 != 0) || (s == nmajor - 1 && t1 != nnz)) {
      b = 1;
      continue;
    }
    for (int t = t0 + lane; t < t1; t += 32) {
      const int j = idx[t];
      if (j < 0 || j >= nminor) b = 1;
      keys[t] = MAJOR_HIGH ? (((u64)(unsigned)s << 32) | (unsigned)j) : (((u64)(unsigned)j << 32) | (unsigned)s);
    }
  }
  if (__any_sync(FULL, b) && lane == 0) atomicExch(bad, 1);
}

// ------------------------------------------------------------------ COO -> CSR by row buckets
// For inputs whose rows are short (the reference's matrices: C2 holds 10.5 entries per row) the five
// LSD passes over (row, column) keys are mostly wasted motion: a row histogram + scan gives every row
// its segment, the entries are dropped into their row's segment in whatever order the atomics decide,
// and each segment is then sorted on (column, input position) - which is a total order, so the result
// does not depend on the order of arrival, and duplicates end up in input order as the summation
// rule requires.  Rows longer than COO_ROW_CAP take the radix path (the host decides from the longest
// row).
constexpr int COO_ROW_CAP = 512;  // 8 warps x 512 keys x 8 bytes = 32 KB of static shared memory

__global__ void __launch_bounds__(256) k_coo_row_count(long long nnz, int m, int n, const int *__restrict__ row,
                                                        const int *__restrict__ col, int *__restrict__ cnt,
                                                        int *__restrict__ bad) {
  int b = 0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (long long)gridDim.x * blockDim.x) {
    const int r = row[i], c = col[i];
    if (r < 0 || r >= m || c < 0 || c >= n) b = 1;
    else atomicAdd(cnt + r, 1);
  }
  if (__any_sync(FULL, b) && lane_id() == 0) atomicExch(bad, 1);
}

// keyc[pos] = column << 32 | input position, pos drawn from the row's cursor (rows with a bad
// coordinate were not counted and are skipped here as well).
__global__ void __launch_bounds__(256) k_coo_bucket(long long nnz, int m, int n, const int *__restrict__ row,
                                                     const int *__restrict__ col, int *__restrict__ cursor,
                                                     u64 *__restrict__ keyc) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (long long)gridDim.x * blockDim.x) {
    const int r = row[i], c = col[i];
    if (r < 0 || r >= m || c < 0 || c >= n) continue;
    const int pos = atomicAdd(cursor + r, 1);
    keyc[pos] = ((u64)(unsigned)c << 32) | (unsigned)i;
  }
}

// Bitonic network over the G lanes of a subgroup (G = 16 or 32), ascending; `sl` = lane inside the subgroup.
template <int G>
__device__ __forceinline__ u64 subgroup_sort(u64 key, int sl) {
#pragma unroll
  for (int k = 2; k <= G; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
      const u64 other = __shfl_xor_sync(FULL, key, j);  // j < G: the partner is in the same subgroup
      const bool up = (sl & k) == 0, lower = (sl & j) == 0;
      const u64 lo_ = key < other ? key : other, hi_ = key < other ? other : key;
      key = (lower == up) ? lo_ : hi_;
    }
  }
  return key;
}

// Sorts every row's (column, position) keys in place (keyc) and counts its distinct columns (ucnt[r];
// ucnt[m] is left alone).  A warp takes two neighbouring rows at a time: both of at most 16 entries -
// the common case on the reference's matrices - are sorted side by side in the two half warps; rows of up
// to 32 entries by the whole warp in registers, longer ones (<= COO_ROW_CAP) in the warp's slice of
// shared memory.
__global__ void __launch_bounds__(256) k_coo_sort_rows(int m, const int *__restrict__ rowptr, u64 *__restrict__ keyc,
                                                        int *__restrict__ ucnt) {
  __shared__ u64 s_keys[8][COO_ROW_CAP];
  const int lane = lane_id(), w = threadIdx.x >> 5;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int r0 = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 2; r0 < m; r0 += warps * 2) {
    const int b0 = __ldg(rowptr + r0), b1 = __ldg(rowptr + r0 + 1);
    const int b2 = r0 + 1 < m ? __ldg(rowptr + r0 + 2) : b1;
    const int len0 = b1 - b0, len1 = b2 - b1;
    if (len0 <= 16 && len1 <= 16) {
      const int sg = lane >> 4, sl = lane & 15;
      const int b = sg ? b1 : b0, len = sg ? len1 : len0;
      u64 key = sl < len ? keyc[b + sl] : ~0ull;
      key = subgroup_sort<16>(key, sl);
      const u64 prev = __shfl_up_sync(FULL, key, 1);
      const bool head = sl < len && (sl == 0 || (prev >> 32) != (key >> 32));
      const unsigned heads = __ballot_sync(FULL, head);
      if (sl < len) keyc[b + sl] = key;
      if (lane == 0) ucnt[r0] = __popc(heads & 0xffffu);
      if (lane == 16 && r0 + 1 < m) ucnt[r0 + 1] = __popc(heads >> 16);
      continue;
    }
    for (int q = 0; q < 2 && r0 + q < m; ++q) {
      const int b = q ? b1 : b0, len = q ? len1 : len0;
      if (len <= 32) {
        u64 key = lane < len ? keyc[b + lane] : ~0ull;
        key = subgroup_sort<32>(key, lane);
        const u64 prev = __shfl_up_sync(FULL, key, 1);
        const bool head = lane < len && (lane == 0 || (prev >> 32) != (key >> 32));
        const unsigned heads = __ballot_sync(FULL, head);
        if (lane < len) keyc[b + lane] = key;
        if (lane == 0) ucnt[r0 + q] = __popc(heads);
      } else {
        u64 *sk = s_keys[w];
        const int P = next_pow2(len);
        for (int t = lane; t < P; t += 32) sk[t] = t < len ? keyc[b + t] : ~0ull;
        __syncwarp();
        for (int k = 2; k <= P; k <<= 1) {
          for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = lane; t < P; t += 32) {
              const int u = t ^ j;
              if (u > t) {
                const u64 x = sk[t], y = sk[u];
                const bool up = (t & k) == 0;
                if ((x > y) == up) sk[t] = y, sk[u] = x;
              }
            }
            __syncwarp();
          }
        }
        int heads = 0;
        for (int t = lane; t < len; t += 32) {
          const u64 key = sk[t];
          keyc[b + t] = key;
          heads += (t == 0 || (sk[t - 1] >> 32) != (key >> 32));
        }
        heads = warp_sum(heads);
        if (lane == 0) ucnt[r0 + q] = heads;
        __syncwarp();
      }
    }
  }
}

// Warp per row: the row's sorted (column, position) keys -> its distinct columns and their sums (the
// duplicates of a column are neighbours, in input order, and are added sequentially with separately
// rounded adds), written at Cp[r] (the scanned distinct-column counts).
template <typename T>
__global__ void __launch_bounds__(256) k_coo_emit_rows(int m, const int *__restrict__ rowptr, const u64 *__restrict__ keyc,
                                                        const T *__restrict__ val, const int *__restrict__ Cp,
                                                        int *__restrict__ Cj, T *__restrict__ Cx) {
  const int lane = lane_id();
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < m; r += warps) {
    const int b = __ldg(rowptr + r), len = __ldg(rowptr + r + 1) - b;
    int out = __ldg(Cp + r);
    for (int t0 = 0; t0 < len; t0 += 32) {
      const int t = t0 + lane;
      const bool in = t < len;
      const u64 key = in ? keyc[b + t] : ~0ull;
      const int col = (int)(key >> 32);
      const bool head = in && (t == 0 || (int)(keyc[b + t - 1] >> 32) != col);
      const unsigned heads = __ballot_sync(FULL, head);
      if (head) {
        T sum = val[(unsigned)key];
        for (int e = t + 1; e < len; ++e) {  // the rest of the run (usually empty)
          const u64 k2 = keyc[b + e];
          if ((int)(k2 >> 32) != col) break;
          sum = add_rn(sum, val[(unsigned)k2]);
        }
        const int o = out + __popc(heads & ((1u << lane) - 1u));
        Cj[o] = col;
        Cx[o] = sum;
      }
      out += __popc(heads);
    }
  }
}

// CSR with unsorted rows / duplicates -> the (column, position) keys of its entries, in place of the
// bucket pass (the entries are grouped by row already).  Raises *bad like k_keys_from_compressed.
__global__ void __launch_bounds__(256) k_keyc_from_csr(int m, int n, long long nnz, const int *__restrict__ Ap,
                                                        const int *__restrict__ Aj, u64 *__restrict__ keyc,
                                                        int *__restrict__ bad) {
  const int lane = lane_id();
  const int warps = (gridDim.x * blockDim.x) >> 5;
  int b = 0;
  for (int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < m; r += warps) {
    const int t0 = Ap[r], t1 = Ap[r + 1];
    if (t0 < 0 || t1 < t0 || t1 > nnz || (r == 0 && t0 != 0) || (r == m - 1 && t1 != nnz)) {
      b = 1;
      continue;
    }
    for (int t = t0 + lane; t < t1; t += 32) {
      const int j = Aj[t];
      if (j < 0 || j >= n) b = 1;
      keyc[t] = ((u64)(unsigned)j << 32) | (unsigned)t;
    }
  }
  if (__any_sync(FULL, b) && lane == 0) atomicExch(bad, 1);
}

// Transposition (CSC <-> CSR) by the same buckets: count the entries of every minor index, scan, drop
// (major, position) keys into the minor's segment, sort the segments (k_coo_sort_rows), write out.
__global__ void __launch_bounds__(256) k_count_minor(long long nnz, int nminor, const int *__restrict__ idx,
                                                      int *__restrict__ cnt, int *__restrict__ bad) {
  int b = 0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (long long)gridDim.x * blockDim.x) {
    const int j = idx[i];
    if (j < 0 || j >= nminor) b = 1;
    else atomicAdd(cnt + j, 1);
  }
  if (__any_sync(FULL, b) && lane_id() == 0) atomicExch(bad, 1);
}

__global__ void __launch_bounds__(256) k_bucket_compressed(int nmajor, int nminor, long long nnz, const int *__restrict__ ptr,
                                                            const int *__restrict__ idx, int *__restrict__ cursor,
                                                            u64 *__restrict__ keyc, int *__restrict__ bad) {
  const int lane = lane_id();
  const int warps = (gridDim.x * blockDim.x) >> 5;
  int b = 0;
  for (int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; s < nmajor; s += warps) {
    const int t0 = ptr[s], t1 = ptr[s + 1];
    if (t0 < 0 || t1 < t0 || t1 > nnz || (s == 0 && t0 != 0) || (s == nmajor - 1 && t1 != nnz)) {
      b = 1;
      continue;
    }
    for (int t = t0 + lane; t < t1; t += 32) {
      const int j = idx[t];
      if (j < 0 || j >= nminor) continue;  // flagged by k_count_minor
      const int pos = atomicAdd(cursor + j, 1);
      keyc[pos] = ((u64)(unsigned)s << 32) | (unsigned)t;
    }
  }
  if (__any_sync(FULL, b) && lane == 0) atomicExch(bad, 1);
}

template <typename T>
__global__ void __launch_bounds__(256) k_emit_bucketed(long long nnz, const u64 *__restrict__ keyc, const T *__restrict__ val,
                                                        int *__restrict__ idx_out, T *__restrict__ val_out) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (long long)gridDim.x * blockDim.x) {
    const u64 key = keyc[i];
    idx_out[i] = (int)(key >> 32);
    val_out[i] = val[(unsigned)key];
  }
}

// ------------------------------------------------------------------ outputs from sorted keys
// ptr_out[s] = first position of a key whose major (high word) is >= s, for s in [0, nmajor];
// *n_dev (when given) overrides n: the number of keys is only known on the device after the
// duplicate summation.
__global__ void __launch_bounds__(256) k_ptr_from_sorted(int nmajor, const u64 *__restrict__ keys, long long n,
                                                          const long long *__restrict__ n_dev,
                                                          int *__restrict__ ptr_out) {
  if (n_dev) n = *n_dev;
  for (int s = blockIdx.x * blockDim.x + threadIdx.x; s <= nmajor; s += gridDim.x * blockDim.x) {
    const u64 want = (u64)(unsigned)s << 32;
    long long lo = 0, hi = n;
    while (lo < hi) {
      const long long mid = (lo + hi) >> 1;
      if (keys[mid] < want) lo = mid + 1; else hi = mid;
    }
    ptr_out[s] = (int)lo;
  }
}

// No duplicate handling (pure re-ordering: csc2csr / csr2csc): minor index + gathered value.
template <typename T>
__global__ void __launch_bounds__(256) k_emit_sorted(long long nnz, const u64 *__restrict__ keys,
                                                      const int *__restrict__ perm, const T *__restrict__ val,
                                                      int *__restrict__ idx_out, T *__restrict__ val_out) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (long long)gridDim.x * blockDim.x) {
    idx_out[i] = (int)(unsigned)keys[i];
    val_out[i] = val[perm[i]];
  }
}

// Duplicate summation, step 1: number of run heads (key differs from its predecessor) per tile.
__global__ void __launch_bounds__(GS_THREADS) k_uniq_count(const u64 *__restrict__ keys, long long n,
                                                             int *__restrict__ tile_heads) {
  __shared__ int s_warp[GS_THREADS / 32];
  const long long base = (long long)blockIdx.x * GS_TILE;
  int c = 0;
#pragma unroll
  for (int q = 0; q < GS_ITEMS; ++q) {
    const long long i = base + q * GS_THREADS + threadIdx.x;
    if (i < n) c += (i == 0 || keys[i] != keys[i - 1]);
  }
  int total;
  block_exscan_i32(c, s_warp, &total);
  if (threadIdx.x == 0) tile_heads[blockIdx.x] = total;
}

// Step 2: every head writes its key (for the row pointers), its column and the sequential sum of
// its run (input order).  tile_off = exclusive scan of tile_heads.
template <typename T>
__global__ void __launch_bounds__(GS_THREADS) k_uniq_emit(const u64 *__restrict__ keys, const int *__restrict__ perm,
                                                            const T *__restrict__ val, long long n,
                                                            const int *__restrict__ tile_off, u64 *__restrict__ ukeys,
                                                            int *__restrict__ Cj, T *__restrict__ Cx) {
  __shared__ int s_warp[GS_THREADS / 32];
  const long long base = (long long)blockIdx.x * GS_TILE + (long long)threadIdx.x * GS_ITEMS;
  unsigned heads = 0;
#pragma unroll
  for (int q = 0; q < GS_ITEMS; ++q) {
    const long long i = base + q;
    if (i < n && (i == 0 || keys[i] != keys[i - 1])) heads |= 1u << q;
  }
  int total;
  int u = tile_off[blockIdx.x] + block_exscan_i32(__popc(heads), s_warp, &total);
  while (heads) {
    const int q = __ffs(heads) - 1;
    heads &= heads - 1;
    const long long i = base + q;
    const u64 k = keys[i];
    T s = val[perm[i]];
    for (long long e = i + 1; e < n && keys[e] == k; ++e) s = add_rn(s, val[perm[e]]);
    ukeys[u] = k;
    Cj[u] = (int)(unsigned)k;
    Cx[u] = s;
    ++u;
  }
}

// ------------------------------------------------------------------ dense / compare (N4)
// CSR -> dense (row- or column-major), duplicates accumulate in storage order (one warp per row
// keeps the order: lanes of one round hit distinct or equal columns; equal columns are rare and
// resolved by ordered_add-like serialisation through match_any).
template <typename T>
__global__ void __launch_bounds__(256) k_csr_to_dense(int m, int n, const int *__restrict__ Ap,
                                                       const int *__restrict__ Aj, const T *__restrict__ Ax,
                                                       T *__restrict__ out, long long ld_row, long long ld_col) {
  const int lane = lane_id();
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < m; r += warps) {
    const int b = Ap[r], e = Ap[r + 1];
    for (int t0 = b; t0 < e; t0 += 32) {
      const int t = t0 + lane;
      const bool act = t < e;
      const int j = act ? Aj[t] : -1 - lane;
      const T v = act ? Ax[t] : T(0);
      const unsigned peers = __match_any_sync(FULL, j);
      T *dst = out + (long long)r * ld_row + (long long)(act ? j : 0) * ld_col;
      if (peers == (1u << lane)) {
        if (act) *dst = add_rn(*dst, v);
      } else {
        // duplicates inside this round: lowest lane first
        unsigned rest = peers;
        while (rest) {
          const int l = __ffs(rest) - 1;
          rest &= rest - 1;
          if (l == lane && act) *dst = add_rn(*dst, v);
          __syncwarp(peers);
        }
      }
      __syncwarp();
    }
  }
}

// max |A - B| over the union pattern of two canonical CSR matrices of the same shape (absent
// entries count as 0).  Warp per row; each lane looks its entries up in the other row by binary
// search.  result[0] = bit pattern of the max (non-negative doubles order like u64),
// result[1] = 1 if a NaN was met.
template <typename T>
__device__ __forceinline__ int row_find(const int *__restrict__ idx, int b, int e, int j) {
  int lo = b, hi = e;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (idx[mid] < j) lo = mid + 1; else hi = mid;
  }
  return (lo < e && idx[lo] == j) ? lo : -1;
}

template <typename T>
__global__ void __launch_bounds__(256) k_csr_max_abs_diff(int m, const int *__restrict__ Ap, const int *__restrict__ Aj,
                                                           const T *__restrict__ Ax, const int *__restrict__ Bp,
                                                           const int *__restrict__ Bj, const T *__restrict__ Bx,
                                                           u64 *__restrict__ result) {
  const int lane = lane_id();
  const int warps = (gridDim.x * blockDim.x) >> 5;
  double mx = 0.0;
  int nan = 0;
  for (int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < m; r += warps) {
    const int ab = Ap[r], ae = Ap[r + 1], bb = Bp[r], be = Bp[r + 1];
    for (int t = ab + lane; t < ae; t += 32) {
      const int p = row_find<T>(Bj, bb, be, Aj[t]);
      const T d = p >= 0 ? Ax[t] - Bx[p] : Ax[t];
      const double a = fabs((double)d);
      if (a != a) nan = 1; else mx = fmax(mx, a);
    }
    for (int t = bb + lane; t < be; t += 32) {
      if (row_find<T>(Aj, ab, ae, Bj[t]) < 0) {
        const double a = fabs((double)Bx[t]);
        if (a != a) nan = 1; else mx = fmax(mx, a);
      }
    }
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    mx = fmax(mx, __shfl_xor_sync(FULL, mx, o));
    nan |= __shfl_xor_sync(FULL, nan, o);
  }
  if (lane == 0) {
    if (mx > 0.0) atomicMax(result, (u64)__double_as_longlong(mx));
    if (nan) atomicExch(result + 1, 1ull);
  }
}

}  // namespace b200sp
